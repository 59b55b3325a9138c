"""Column sharding + flux gather on CPU with world_size 2 and 3 (gloo): the N>1 host logic of
bench.py / rte_rrtmgp_cpp_b200.shard, including ragged shards."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from rte_rrtmgp_cpp_b200 import shard


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _flux(field, c, lev):
    return 1000.0 * field + lev + 1e-3 * c


def _worker(rank, world, port, ncol, nlev, nfield, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    c0, c1 = shard.shard_bounds(ncol, world, rank)
    cols = torch.arange(c0, c1, dtype=torch.float64)
    levs = torch.arange(nlev, dtype=torch.float64)
    fields = [_flux(f, cols[None, :], levs[:, None]) for f in range(nfield)]      # (nlev, ncol_shard)
    packed = shard.pack_fields(fields)
    recv = shard.gather_fields(packed, ncol, nlev, dst=0)
    mean = shard.domain_mean(fields[0].sum(dim=1).clone(), ncol)
    if rank == 0:
        full = shard.assemble(recv, ncol, nlev)
        q.put((full.numpy(), mean.numpy()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,ncol", [(2, 10), (2, 7), (3, 11)])
def test_gather_reassembles_reference_layout(world, ncol):
    nlev, nfield = 5, 3
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, ncol, nlev, nfield, q)) for r in range(world)]
    for p in procs:
        p.start()
    full, mean = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    c = np.arange(ncol)[None, None, :]; l = np.arange(nlev)[None, :, None]; f = np.arange(nfield)[:, None, None]
    want = _flux(f, c, l)
    assert np.array_equal(full, want)
    assert np.allclose(mean, want[0].mean(axis=1))


def test_shard_bounds_cover_domain():
    for ncol in (0, 1, 7, 1024 * 1024):
        for world in (1, 2, 3, 8):
            b = [shard.shard_bounds(ncol, world, r) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == ncol
            assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            sizes = [c1 - c0 for c0, c1 in b]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard.shard_bounds(10, 2, 2)

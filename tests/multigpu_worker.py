"""Worker of tests/test_zz_multigpu.py (run under torch.distributed.run, one process per GPU): every rank computes the fluxes
of its column shard in `nparts` blocks and gathers each block to rank 0 through the C ABI (rrtmgp_b200_gather_fluxes over
NCCL); rank 0 compares with the whole domain computed on its own GPU: bit for bit."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rte_rrtmgp_cpp_b200 import api                                   # noqa: E402
from rte_rrtmgp_cpp_b200.atmos import rce_atmosphere                   # noqa: E402
from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist                  # noqa: E402
from rte_rrtmgp_cpp_b200.shard import ShardPlan, shard_bounds          # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ncol, nlay, nparts, dtype = 203, 24, 3, torch.float32          # ragged shards and ragged blocks
    atm = rce_atmosphere(ncol, nlay)
    kd_lw, kd_sw = synthetic_kdist("lw"), synthetic_kdist("sw")
    lw = api.Radiation_solver_longwave(kd_lw, atm.vmr.keys(), None, dtype)
    sw = api.Radiation_solver_shortwave(kd_sw, atm.vmr.keys(), None, dtype)
    plan = ShardPlan(ncol, world, rank)
    c0, c1 = plan.bounds()
    assert (c0, c1) == shard_bounds(ncol, world, rank)
    nlev = nlay + 1
    names = [("lw", "flux_up"), ("lw", "flux_dn"), ("sw", "flux_up"), ("sw", "flux_dn"), ("sw", "flux_dn_dir")]
    full = [torch.full((nlev * ncol,), float("nan"), device="cuda", dtype=dtype) for _ in names] if rank == 0 else None
    side = torch.cuda.Stream()
    keep = []
    mode = os.environ.get("RB200_TEST_GATHER", "nccl")
    if mode == "p2p":       # peer-memory mode: every rank writes into rank 0's arrays over NVLink
        nreg = plan.register(full, len(names))
        assert nreg == len(names), f"peer-memory registration failed on rank {rank}: {nreg}"
    for part in range(nparts):
        p0, p1 = shard_bounds(c1 - c0, nparts, part)
        o_lw, o_sw = api.solve_pair(lw, sw, api.DeviceColumns(atm.subset(c0 + p0, c0 + p1), dtype))
        fields = [(o_lw if w == "lw" else o_sw)[k].t for w, k in names]
        keep.append(fields)
        side.wait_stream(torch.cuda.current_stream())
        plan.gather(fields, full, nlev, part, nparts, root=0, stream=side)     # overlaps the next block's kernels
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    ok = True
    if rank == 0:
        r_lw, r_sw = api.solve_pair(lw, sw, api.DeviceColumns(atm, dtype))
        torch.cuda.synchronize()
        for (w, k), got in zip(names, full):
            want = (r_lw if w == "lw" else r_sw)[k].t.reshape(-1)
            if not torch.equal(got, want):
                ok = False
                print(f"MISMATCH {w} {k}: max abs diff {(got - want).abs().max().item()}", flush=True)
        print("MULTIGPU_GATHER_OK" if ok else "MULTIGPU_GATHER_FAILED", flush=True)
    dist.barrier()
    plan.close()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()

"""rrtmgp_b200_gather_fluxes on one GPU (world = 1): a shard processed in column blocks lands in the reference layout."""
import numpy as np
import pytest
import torch

from rte_rrtmgp_cpp_b200 import api
from rte_rrtmgp_cpp_b200.atmos import rce_atmosphere
from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist
from rte_rrtmgp_cpp_b200.shard import ShardPlan, shard_bounds

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_gather_of_column_blocks_reassembles_the_domain(dtype):
    ncol, nlev, nparts = 101, 7, 3
    plan = ShardPlan(ncol, 1, 0)
    assert plan.bounds() == (0, ncol)
    g = torch.Generator(device="cuda").manual_seed(5)
    truth = [torch.rand(nlev, ncol, device="cuda", dtype=dtype, generator=g) for _ in range(2)]     # (ncol, nlev) column-major
    full = [torch.full((nlev, ncol), float("nan"), device="cuda", dtype=dtype) for _ in range(2)]
    for part in range(nparts):
        c0, c1 = shard_bounds(ncol, nparts, part)
        blocks = [t[:, c0:c1].contiguous() for t in truth]                                           # (count, nlev) column-major
        plan.gather(blocks, full, nlev, part, nparts)
    torch.cuda.synchronize()
    for t, f in zip(truth, full):
        assert torch.equal(t, f)
    plan.close()


def test_fluxes_computed_in_column_blocks_and_gathered_equal_the_whole_call():
    """A shard split into blocks (what a host model does per column block), each block's fluxes gathered: bit-identical to one call."""
    kd_lw, kd_sw = synthetic_kdist("lw"), synthetic_kdist("sw")
    atm = rce_atmosphere(90, 20)
    dtype = torch.float32
    lw = api.Radiation_solver_longwave(kd_lw, atm.vmr.keys(), None, dtype)
    sw = api.Radiation_solver_shortwave(kd_sw, atm.vmr.keys(), None, dtype)
    ref_lw, ref_sw = api.solve_pair(lw, sw, api.DeviceColumns(atm, dtype))
    ref = [ref_lw["flux_up"].t.clone(), ref_lw["flux_dn"].t.clone(), ref_sw["flux_up"].t.clone(), ref_sw["flux_dn"].t.clone(), ref_sw["flux_dn_dir"].t.clone()]
    nlev, nparts = atm.nlay + 1, 4
    plan = ShardPlan(atm.ncol, 1, 0)
    full = [torch.zeros(nlev * atm.ncol, device="cuda", dtype=dtype) for _ in range(5)]
    for part in range(nparts):
        c0, c1 = shard_bounds(atm.ncol, nparts, part)
        o_lw, o_sw = api.solve_pair(lw, sw, api.DeviceColumns(atm.subset(c0, c1), dtype))
        plan.gather([o_lw["flux_up"].t, o_lw["flux_dn"].t, o_sw["flux_up"].t, o_sw["flux_dn"].t, o_sw["flux_dn_dir"].t], full, nlev, part, nparts)
    torch.cuda.synchronize()
    for a, b in zip(ref, full):
        assert torch.equal(a.reshape(-1), b)

"""The two fused implementations behind rrtmgp_b200_{lw,sw}_fluxes — column-resident (recurrences
in shared memory, default) and per-band kernels with HBM scratch strips — run the same per-cell
arithmetic, so they must agree to rounding with each other and each within tolerance of the oracle.
Odd sizes exercise partial CTAs / partial layer groups; non-zeroed outputs catch unwritten levels."""
import numpy as np
import pytest
import torch

from rte_rrtmgp_cpp_b200 import api
from rte_rrtmgp_cpp_b200.atmos import rce_atmosphere
from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist, synthetic_cloud_lut, REDUCED_LW_128, REDUCED_SW_112
from helpers import oracle_lw, oracle_sw

pytestmark = pytest.mark.gpu
TOL = {torch.float64: 1e-5, torch.float32: 0.05}
KEYS = {"lw": ("flux_up", "flux_dn", "flux_net", "bnd_flux_up", "bnd_flux_dn", "bnd_flux_net"),
        "sw": ("flux_up", "flux_dn", "flux_net", "flux_dn_dir", "bnd_flux_up", "bnd_flux_dn", "bnd_flux_net", "bnd_flux_dn_dir")}


def _poisoned(sol, cols, nb, sw):
    """output arrays pre-filled with garbage: every element must be written by the call"""
    out = {}
    ncol, nlev = cols.ncol, cols.nlay + 1
    names = [("flux_up", 2), ("flux_dn", 2), ("flux_net", 2), ("bnd_flux_up", 3), ("bnd_flux_dn", 3), ("bnd_flux_net", 3)]
    if sw:
        names += [("flux_dn_dir", 2), ("bnd_flux_dn_dir", 3)]
    for k, nd in names:
        a = api.Array_gpu((ncol, nlev) if nd == 2 else (ncol, nlev, nb), sol.dtype)
        a.t.fill_(-7777.0)
        out[k] = a
    return out


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("kind", ["lw", "sw"])
@pytest.mark.parametrize("ncol,nlay,top_at_1,cloudy,reduced", [
    (70, 48, False, False, False), (33, 37, True, True, False), (1, 128, False, True, False),
    (37, 30, False, True, True), (64, 129, True, False, True),
    (5, 300, False, True, False)])          # taller than the 256 layers one tropopause-mask block covers
def test_paths_agree(dtype, kind, ncol, nlay, top_at_1, cloudy, reduced):
    sizes = (REDUCED_LW_128 if kind == "lw" else REDUCED_SW_112) if reduced else None
    kd = synthetic_kdist(kind, sizes) if reduced else synthetic_kdist(kind)
    lut = synthetic_cloud_lut(kind) if cloudy else None
    atm = rce_atmosphere(ncol, nlay, top_at_1=top_at_1, cloudy=cloudy, night_fraction=0.2 if kind == "sw" else 0.0)
    ref = (oracle_lw if kind == "lw" else oracle_sw)(kd, atm, lut)
    Solver = api.Radiation_solver_longwave if kind == "lw" else api.Radiation_solver_shortwave
    cols = api.DeviceColumns(atm, dtype)
    res = {}
    for path in (1, 0):
        sol = Solver(kd, atm.vmr.keys(), lut, dtype)
        sol.set_fused_path(path)
        out = _poisoned(sol, cols, kd.nbnd, kind == "sw")
        sol.solve_gpu(cols, switch_cloud_optics=cloudy, switch_output_bnd_fluxes=True, out=out)
        res[path] = {k: out[k].numpy() for k in KEYS[kind]}
        for k in KEYS[kind]:
            d = np.max(np.abs(res[path][k] - ref[k]))
            assert d <= TOL[dtype], f"path {path} {k}: max abs diff {d}"
    rt = 1e-11 if dtype == torch.float64 else 2e-5
    for k in KEYS[kind]:
        scale = max(1.0, float(np.max(np.abs(res[0][k]))))
        assert np.max(np.abs(res[1][k] - res[0][k])) <= rt * scale, k


def test_path_is_chosen_by_call_size():
    """set_fused_path(2): calls of up to 6144 columns run the column-resident kernels (one launch for all bands), larger ones
    the per-band kernels; both within tolerance of each other.  Counted through the library's launch counter."""
    from rte_rrtmgp_cpp_b200 import capi
    L = capi.lib()
    kd = synthetic_kdist("lw")
    launches = {}
    for ncol in (64, 6144, 6145):
        atm = rce_atmosphere(ncol, 12)
        sol = api.Radiation_solver_longwave(kd, atm.vmr.keys(), None, torch.float32)
        sol.set_fused_path(2)
        cols = api.DeviceColumns(atm, torch.float32)
        n0 = L.rrtmgp_b200_launch_count()
        out = sol.solve_gpu(cols)
        torch.cuda.synchronize()
        launches[ncol] = L.rrtmgp_b200_launch_count() - n0
        sol.set_fused_path(0)
        ref = sol.solve_gpu(cols, out={})
        assert np.max(np.abs(out["flux_up"].numpy() - ref["flux_up"].numpy())) < 1e-2
    # prologue + one kernel for all bands (+ col_dry, net); a call too small to fill the SMs spreads its bands over the grid of
    # that one launch and adds the per-band partial sums in a second, tiny one
    assert launches[64] == launches[6144] + 1 < kd.nbnd
    assert launches[6145] >= kd.nbnd + 1                      # prologue + one launch per band

"""Committed golden vectors (tests/golden/ref_fluxes.npz): fluxes computed HERE by the reference's own,
unmodified C++ classes (oracle/_ref) on small seeded column sets; generator tests/golden/make_golden.py.
CPU: the numpy oracle must reproduce them bit for bit (so the oracle stays pinned where oracle/_ref is not
built).  GPU: the fused CUDA path through the C ABI must match them within BASELINE.json's tolerances."""
import numpy as np
import pytest
import torch

import golden_cases as G
from helpers import oracle_lw, oracle_sw
from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist, synthetic_cloud_lut, synthetic_aerosol_lut

Z, CASES = G.load()
LW_KEYS = ("flux_up", "flux_dn", "flux_net", "bnd_flux_up", "bnd_flux_dn", "bnd_flux_net")
SW_KEYS = LW_KEYS + ("flux_dn_dir", "bnd_flux_dn_dir")


def test_fixture_is_complete_and_physical():
    for name, (kind, ncol, nlay, top1, *_rest) in CASES.items():
        fl, atm = G.fluxes(Z, name), G.atmosphere(Z, name)
        assert (atm.ncol, atm.nlay, atm.top_at_1) == (ncol, nlay, top1)
        for k in (LW_KEYS if kind == "lw" else SW_KEYS):
            assert fl[k].shape[:2] == (ncol, nlay + 1) and np.all(np.isfinite(fl[k])), (name, k)
        assert np.array_equal(fl["flux_net"], fl["flux_dn"] - fl["flux_up"])
        assert fl["flux_dn"].max() > 100.0


@pytest.mark.parametrize("name", list(CASES))
def test_oracle_reproduces_reference_golden_bitwise(name):
    kind, _, _, _, cloudy, dcl, aer, daer, n_ang = CASES[name]
    atm, want = G.atmosphere(Z, name), G.fluxes(Z, name)
    clut = synthetic_cloud_lut(kind) if cloudy else None
    if kind == "lw":
        got = oracle_lw(synthetic_kdist("lw"), atm, clut, n_ang=n_ang)
    else:
        got = oracle_sw(synthetic_kdist("sw"), atm, clut, delta=dcl, aerosol_lut=synthetic_aerosol_lut() if aer else None,
                        delta_aerosol=daer)
    for k in (LW_KEYS if kind == "lw" else SW_KEYS):
        assert np.array_equal(got[k], want[k]), f"{name}/{k}: max abs diff {np.max(np.abs(got[k] - want[k]))}"


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("name", list(CASES))
def test_cuda_path_matches_reference_golden(name, dtype):
    from rte_rrtmgp_cpp_b200 import api
    kind, _, _, _, cloudy, dcl, aer, daer, n_ang = CASES[name]
    atm, want = G.atmosphere(Z, name), G.fluxes(Z, name)
    clut = synthetic_cloud_lut(kind) if cloudy else None
    Solver = api.Radiation_solver_longwave if kind == "lw" else api.Radiation_solver_shortwave
    sol = Solver(synthetic_kdist(kind), atm.vmr.keys(), clut, dtype, aerosol_lut=synthetic_aerosol_lut() if aer else None)
    kw = dict(switch_cloud_optics=cloudy, switch_delta_cloud=dcl, switch_output_bnd_fluxes=True)
    if kind == "lw":
        sol.n_gauss_angles = n_ang
    else:
        kw.update(switch_aerosol_optics=aer, switch_delta_aerosol=daer)
    out = sol.solve_gpu(api.DeviceColumns(atm, dtype), **kw)
    tol = 1e-5 if dtype == torch.float64 else 0.05          # W/m2, BASELINE.json north_star
    for k in (LW_KEYS if kind == "lw" else SW_KEYS):
        d = np.max(np.abs(out[k].numpy() - want[k]))
        assert d <= tol, f"{name}/{k}: max abs diff {d} > {tol}"

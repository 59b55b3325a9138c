"""GPU parity on REAL inputs of rrtmgp-data (tests/golden/rrtmgp_data.npz): the 100 RFMIP clear-sky sites (60 layers,
top first, day and night) and the cloud-optics LUTs the reference ships — the fused CUDA path through the C ABI
against the CPU oracle (itself bit-exact against the reference's compiled classes on these inputs,
tests/test_real_data.py).  Gas tables stay synthetic: rrtmgp-gas-*.nc are absent from the reference mount.
Tolerances are BASELINE.json's: 1e-5 W/m2 fp64, 0.05 W/m2 fp32."""
import numpy as np
import pytest
import torch

import real_data as RD
from helpers import oracle_lw, oracle_sw, as_f32_valued
from rte_rrtmgp_cpp_b200 import api
from rte_rrtmgp_cpp_b200.atmos import allsky_atmosphere
from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist

pytestmark = pytest.mark.gpu
Z = RD.fixture()
TOL = {torch.float64: 1e-5, torch.float32: 0.05}
LW_KEYS = ("flux_up", "flux_dn", "flux_net", "bnd_flux_up", "bnd_flux_dn", "bnd_flux_net")
SW_KEYS = LW_KEYS + ("flux_dn_dir", "bnd_flux_dn_dir")


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("expt", [0, 3])
def test_rfmip_sites_fused(expt, dtype):
    atm = RD.rfmip_atmosphere(Z, expt, tsi_ref=RD.default_tsi(synthetic_kdist("sw")))
    if dtype == torch.float32:
        atm = as_f32_valued(atm)
    for kind, oracle, Solver, keys in (("lw", oracle_lw, api.Radiation_solver_longwave, LW_KEYS),
                                       ("sw", oracle_sw, api.Radiation_solver_shortwave, SW_KEYS)):
        kd = synthetic_kdist(kind)
        ref = oracle(kd, atm)
        out = Solver(kd, atm.vmr.keys(), dtype=dtype).solve_gpu(api.DeviceColumns(atm, dtype), switch_output_bnd_fluxes=True)
        for k in keys:
            d = np.max(np.abs(out[k].numpy() - ref[k]))
            assert d <= TOL[dtype], f"expt {expt} {kind} {k}: max abs diff {d}"
        # the part of the PUBLISHED fluxes that does not depend on the gas tables (tests/test_real_data.py)
        if kind == "sw":
            day = atm.mu0 > 0
            pub = Z[f"rfmip/{expt}/pub_rsd"].astype(np.float64)
            assert np.max(np.abs(out["flux_dn"].numpy()[day, 0] / pub[day, 0] - 1.0)) < (1e-6 if dtype == torch.float64 else 1e-4)
        else:
            pub = Z[f"rfmip/{expt}/pub_rlu"].astype(np.float64)
            assert np.max(np.abs(out["flux_up"].numpy()[:, -1] / pub[:, -1] - 1.0)) < 0.01


@pytest.mark.parametrize("delta", [True, False])
def test_real_cloud_luts_fused(delta):
    """fp64 build; the fp32 build's two-stream rearrangement is exercised on the synthetic LUT."""
    atm = allsky_atmosphere(64, 72, random_clouds=True)
    for kind, oracle, Solver, keys in (("lw", oracle_lw, api.Radiation_solver_longwave, LW_KEYS),
                                       ("sw", oracle_sw, api.Radiation_solver_shortwave, SW_KEYS)):
        kd, lut = synthetic_kdist(kind), RD.cloud_lut(Z, kind)
        ref = oracle(kd, atm, lut) if kind == "lw" else oracle(kd, atm, lut, delta=delta)
        sol = Solver(kd, atm.vmr.keys(), lut, torch.float64)
        out = sol.solve_gpu(api.DeviceColumns(atm, torch.float64), switch_cloud_optics=True, switch_delta_cloud=delta,
                            switch_output_bnd_fluxes=True)
        for k in keys:
            d = np.max(np.abs(out[k].numpy() - ref[k]))
            assert d <= 1e-5, f"{kind} {k}: max abs diff {d}"


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("kind", ["lw", "sw"])
def test_rfmip_sites_interpolation_indices_bit_exact(kind, dtype):
    """SURVEY T1 on the real RFMIP columns: jtemp, jpress, jeta, tropo bit for bit against the oracle."""
    from oracle import oracle as O
    A = api.Array_gpu
    kd = synthetic_kdist(kind)
    atm = RD.rfmip_atmosphere(Z, 0)
    if dtype == torch.float32:
        atm = as_f32_valued(atm)
    go = api.Gas_optics_rrtmgp_gpu(kd, atm.vmr.keys(), dtype)
    cols = api.DeviceColumns(atm, dtype)
    col_dry = A((atm.ncol, atm.nlay), dtype)
    go.get_col_dry(col_dry, cols.gas_concs.get_vmr("h2o"), cols.plev)
    cd = col_dry.numpy().astype(np.float64)
    out = go.interpolation(cols.play, cols.tlay, col_dry, cols.gas_concs)
    st = O.GasOptics(kd, atm.vmr.keys())._gas_taus(O.F(atm.play), O.F(atm.plev), O.F(atm.tlay), atm.vmr, O.F(cd))
    for k in ("jtemp", "jpress", "jeta"):
        assert np.array_equal(out[k].numpy(), st[k]), k
    assert np.array_equal(out["tropo"].numpy() != 0, st["tropo"] != 0)
    assert 0.2 < (st["tropo"] != 0).mean() < 0.8                # both sides of the tropopause are exercised

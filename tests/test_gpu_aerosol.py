"""GPU parity of the aerosol optics: the Aerosol_optics_gpu stage call, and the fused
shortwave solver with aerosols (with and without clouds, with and without delta scaling) against the CPU
oracle — src/Aerosol_optics.cpp restated, itself pinned bit-exact against the reference's compiled class
(tests/test_oracle_vs_ref.py).  Tolerances are BASELINE.json's: 1e-5 W/m2 fp64, 0.05 W/m2 fp32."""
import numpy as np
import pytest
import torch

from oracle import oracle as O
from rte_rrtmgp_cpp_b200 import api
from rte_rrtmgp_cpp_b200.atmos import rce_atmosphere, add_aerosols
from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist, synthetic_cloud_lut, synthetic_aerosol_lut, REDUCED_SW_112
from helpers import oracle_sw, as_f32_valued

pytestmark = pytest.mark.gpu
TOL = {torch.float64: 1e-5, torch.float32: 0.05}
KEYS = ("flux_up", "flux_dn", "flux_net", "flux_dn_dir", "bnd_flux_up", "bnd_flux_dn", "bnd_flux_net", "bnd_flux_dn_dir")


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("top_at_1", [False, True])
def test_aerosol_optics_stage(dtype, top_at_1):
    atm = add_aerosols(rce_atmosphere(37, 29, top_at_1=top_at_1))
    if dtype == torch.float32:
        atm = as_f32_valued(atm)
    lut = synthetic_aerosol_lut()
    cols = api.DeviceColumns(atm, dtype)
    ao = api.Aerosol_optics_gpu(lut, dtype)
    assert ao.get_nband() == 14
    op = api.Optical_props_2str_gpu(atm.ncol, atm.nlay, ao, dtype)
    ao.aerosol_optics(cols.aerosol_concs, cols.rh, cols.plev, op)
    # bottom-up columns: the CPU source as it stands; top-down: its GPU twin's |dp| (the CPU source gives tau < 0)
    want = O.AerosolOptics(lut).aerosol_optics(atm.aermr, atm.rh, atm.plev, abs_dp=top_at_1)
    rtol = 1e-13 if dtype == torch.float64 else 5e-6
    # the reference floors tau and tau*ssa at epsilon(Float) before dividing (src/Aerosol_optics.cpp:207-222): in the
    # fp32 build ssa and g of cells thinner than ~1e-7 are floored values, so they are compared where tau >> eps(fp32)
    thick = want[0] > (0.0 if dtype == torch.float64 else 1e-4)
    assert thick.mean() > 0.2
    for got, w, k in zip((op.tau, op.ssa, op.g), want, ("tau", "ssa", "g")):
        assert np.all(w > 0)
        m = thick if k != "tau" else np.ones_like(thick)
        assert np.max(np.abs(got.numpy() - w)[m] / np.abs(w)[m]) < rtol, k
    op.delta_scale()
    wd = O.delta_scale(*want)
    for got, w, k in zip((op.tau, op.ssa, op.g), wd, ("tau", "ssa", "g")):       # scaled tau depends on ssa: same mask
        assert np.max(np.abs(got.numpy() - w)[thick] / np.abs(w)[thick]) < rtol * 20, k
    # errors: missing species, wrong shape
    bad = api.Aerosol_concs_gpu(dtype)
    for k, v in atm.aermr.items():
        if k != "aermr07":
            bad.set_vmr(k, v)
    with pytest.raises(RuntimeError, match="aermr07"):
        ao.aerosol_optics(bad, cols.rh, cols.plev, op)


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("cloudy,delta_aerosol", [(False, False), (True, False), (True, True), (False, True)])
def test_sw_fused_with_aerosols(dtype, cloudy, delta_aerosol):
    kd, clut, alut = synthetic_kdist("sw"), synthetic_cloud_lut("sw"), synthetic_aerosol_lut()
    atm = add_aerosols(rce_atmosphere(70, 48, night_fraction=0.15, cloudy=cloudy))
    ref = oracle_sw(kd, atm, clut if cloudy else None, aerosol_lut=alut, delta_aerosol=delta_aerosol)
    clear = oracle_sw(kd, atm, clut if cloudy else None)
    assert np.max(np.abs(ref["flux_dn"] - clear["flux_dn"])) > 1.0      # the aerosols matter in this case
    sol = api.Radiation_solver_shortwave(kd, atm.vmr.keys(), clut if cloudy else None, dtype, aerosol_lut=alut)
    out = sol.solve_gpu(api.DeviceColumns(atm, dtype), switch_cloud_optics=cloudy, switch_output_bnd_fluxes=True,
                        switch_aerosol_optics=True, switch_delta_aerosol=delta_aerosol)
    for k in KEYS:
        d = np.max(np.abs(out[k].numpy() - ref[k]))
        assert d <= TOL[dtype], f"{k}: max abs diff {d}"
    # the aerosol switch off gives the aerosol-free fluxes from the same solver object
    out0 = sol.solve_gpu(api.DeviceColumns(atm, dtype), switch_cloud_optics=cloudy)
    assert np.max(np.abs(out0["flux_dn"].numpy() - clear["flux_dn"])) <= TOL[dtype]


def test_sw_aerosols_paths_tiles_host_and_reduced_set():
    """Column-resident path, small tiles, the host-buffer pipeline and the reduced g-point set give the same
    aerosol-laden fluxes (bit-identical for tiles/host, to rounding for the other kernel family)."""
    kd, alut = synthetic_kdist("sw", REDUCED_SW_112), synthetic_aerosol_lut()
    atm = add_aerosols(rce_atmosphere(150, 33, top_at_1=True, cloudy=True))
    clut = synthetic_cloud_lut("sw")
    ref = oracle_sw(kd, atm, clut, aerosol_lut=alut)
    sol = api.Radiation_solver_shortwave(kd, atm.vmr.keys(), clut, aerosol_lut=alut)
    kw = dict(switch_cloud_optics=True, switch_aerosol_optics=True)
    full = sol.solve_gpu(api.DeviceColumns(atm), **kw)["flux_net"].numpy().copy()
    assert np.max(np.abs(full - ref["flux_net"])) <= 1e-5
    sol.set_tile_columns(64)
    assert np.array_equal(full, sol.solve_gpu(api.DeviceColumns(atm), **kw)["flux_net"].numpy())
    sol.set_tile_columns(0)
    _, oh = api.solve_host(None, sol, api.HostColumns(atm), chunk_columns=37, **kw)
    assert np.array_equal(full, oh["flux_net"].numpy())
    sol.set_fused_path(1)
    colres = sol.solve_gpu(api.DeviceColumns(atm), **kw)["flux_net"].numpy()
    assert np.max(np.abs(colres - ref["flux_net"])) <= 1e-5


def test_aerosol_errors_are_loud():
    kd, alut = synthetic_kdist("sw"), synthetic_aerosol_lut()
    atm = add_aerosols(rce_atmosphere(5, 12))
    sol = api.Radiation_solver_shortwave(kd, atm.vmr.keys(), aerosol_lut=alut)
    cols = api.DeviceColumns(atm)
    cols.aerosol_concs.gas_concs_map.pop("aermr11")
    with pytest.raises(RuntimeError, match="aermr11"):
        sol.solve_gpu(cols, switch_aerosol_optics=True)
    lw = api.Radiation_solver_longwave(synthetic_kdist("lw"), atm.vmr.keys(), aerosol_lut=alut)
    with pytest.raises(RuntimeError, match="shortwave only"):
        lw.solve_gpu(api.DeviceColumns(atm), switch_aerosol_optics=True)
    with pytest.raises(RuntimeError, match="incompatible band"):
        bad = api.Radiation_solver_shortwave(kd, atm.vmr.keys(), aerosol_lut=synthetic_aerosol_lut(nband=9))
        bad.solve_gpu(api.DeviceColumns(atm), switch_aerosol_optics=True)

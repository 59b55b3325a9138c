"""The numpy restatement of the class layer (oracle/oracle.py) against the reference's own,
unmodified C++ class layer (oracle/_ref, built from /root/reference/src over the restated
kernels).  Skipped when oracle/_ref was not built (e.g. /root/reference absent)."""
import numpy as np
import pytest

from oracle import oracle as O
from oracle import ref as R
from rte_rrtmgp_cpp_b200.atmos import rce_atmosphere, allsky_atmosphere, add_aerosols
from rte_rrtmgp_cpp_b200.kdist import (synthetic_kdist, synthetic_cloud_lut, synthetic_aerosol_lut, REDUCED_LW_128,
                                       REDUCED_SW_112)

pytestmark = pytest.mark.skipif(not R.available(), reason="oracle/_ref/liboracle_ref.so not built")


def _eq(a, b, name):
    assert a.shape == b.shape, name
    assert np.array_equal(a, b), f"{name}: max abs diff {np.max(np.abs(a - b))}"


@pytest.mark.parametrize("top_at_1", [False, True])
@pytest.mark.parametrize("sizes", [16, REDUCED_LW_128])
def test_lw_chain_matches_reference_classes(sizes, top_at_1):
    kd = synthetic_kdist("lw", sizes)
    atm = rce_atmosphere(7, 24, top_at_1=top_at_1)
    go = O.GasOptics(kd, atm.vmr.keys()); rf = R.RefGasOptics(kd, atm.vmr.keys())
    assert (go.ngpt, go.nbnd, go.nflav) == (rf.ngpt, rf.nbnd, rf.nflav)
    col_dry = go.get_col_dry(np.broadcast_to(atm.vmr["h2o"], atm.play.shape), atm.plev)
    _eq(col_dry, rf.get_col_dry(atm.vmr["h2o"], atm.plev), "col_dry")
    st = go.gas_optics_lw(atm.play, atm.plev, atm.tlay, atm.tsfc, atm.vmr, col_dry, atm.tlev)
    rs = rf.gas_optics_lw(atm.play, atm.plev, atm.tlay, atm.tsfc, atm.vmr, col_dry, atm.tlev)
    for k in ("tau", "lay_src", "lev_src", "sfc_src", "sfc_src_jac"):
        _eq(st[k], rs[k], k)
    for bb in (False, True):
        for n_ang in (1, 3):
            up, dn = O.rte_lw(go, st["tau"], st["lay_src"], st["lev_src"], st["sfc_src"], atm.emis_sfc,
                              atm.top_at_1, n_gauss_angles=n_ang, broadband=bb)
            rup, rdn = rf.rte_lw(st["tau"], st["lay_src"], st["lev_src"], st["sfc_src"], atm.emis_sfc,
                                 atm.top_at_1, n_ang, bb)
            _eq(up, rup, "lw up"); _eq(dn, rdn, "lw dn")
    up, dn = O.rte_lw(go, st["tau"], st["lay_src"], st["lev_src"], st["sfc_src"], atm.emis_sfc, atm.top_at_1)
    fo = O.fluxes_byband(go, up, dn); fr = rf.fluxes_byband(up, dn)
    for k in ("flux_up", "flux_dn", "flux_net", "bnd_flux_up", "bnd_flux_dn", "bnd_flux_net"):
        _eq(fo[k], fr[k], k)


@pytest.mark.parametrize("top_at_1", [False, True])
@pytest.mark.parametrize("sizes", [16, REDUCED_SW_112])
def test_sw_chain_matches_reference_classes(sizes, top_at_1):
    kd = synthetic_kdist("sw", sizes)
    atm = rce_atmosphere(9, 20, top_at_1=top_at_1, night_fraction=0.3)
    go = O.GasOptics(kd, atm.vmr.keys()); rf = R.RefGasOptics(kd, atm.vmr.keys())
    col_dry = go.get_col_dry(np.broadcast_to(atm.vmr["h2o"], atm.play.shape), atm.plev)
    st = go.gas_optics_sw(atm.play, atm.plev, atm.tlay, atm.vmr, col_dry)
    rs = rf.gas_optics_sw(atm.play, atm.plev, atm.tlay, atm.vmr, col_dry)
    for k in ("tau", "ssa", "g", "toa_src"):
        _eq(st[k], rs[k], k)
    for bb in (False, True):
        o = O.rte_sw(go, st["tau"], st["ssa"], st["g"], atm.mu0, st["toa_src"], atm.sfc_alb_dir,
                     atm.sfc_alb_dif, atm.top_at_1, broadband=bb)
        r = rf.rte_sw(st["tau"], st["ssa"], st["g"], atm.mu0, st["toa_src"], atm.sfc_alb_dir,
                      atm.sfc_alb_dif, atm.top_at_1, bb)
        for a, b, n in zip(o, r, ("up", "dn", "dir")):
            _eq(a, b, "sw " + n)
    o = O.rte_sw(go, st["tau"], st["ssa"], st["g"], atm.mu0, st["toa_src"], atm.sfc_alb_dir,
                 atm.sfc_alb_dif, atm.top_at_1)
    fo = O.fluxes_byband(go, *o); fr = rf.fluxes_byband(*o)
    for k in fo:
        _eq(fo[k], fr[k], k)


def test_cloud_optics_delta_add_matches_reference_classes():
    atm = allsky_atmosphere(12, 30, random_clouds=True)
    for kind in ("lw", "sw"):
        kd = synthetic_kdist(kind); lut = synthetic_cloud_lut(kind)
        go = O.GasOptics(kd, atm.vmr.keys()); rf = R.RefGasOptics(kd, atm.vmr.keys())
        co = O.CloudOptics(lut)
        if kind == "lw":
            st = go.gas_optics_lw(atm.play, atm.plev, atm.tlay, atm.tsfc, atm.vmr, tlev=atm.tlev)
            ctau = co.cloud_optics_1scl(atm.lwp, atm.iwp, atm.rel, atm.dei)
            tau = O.add_to_1scl(go, st["tau"], ctau)
            r = rf.cloud_add(lut, 1, False, atm.lwp, atm.iwp, atm.rel, atm.dei, st["tau"])
            _eq(ctau, r["ctau"], "ctau"); _eq(tau, r["tau"], "tau+cloud")
        else:
            st = go.gas_optics_sw(atm.play, atm.plev, atm.tlay, atm.vmr)
            for dl in (False, True):
                c = co.cloud_optics_2str(atm.lwp, atm.iwp, atm.rel, atm.dei)
                if dl:
                    c = O.delta_scale(*c)
                t = O.add_to_2str(go, st["tau"], st["ssa"], st["g"], *c)
                r = rf.cloud_add(lut, 2, dl, atm.lwp, atm.iwp, atm.rel, atm.dei, st["tau"], st["ssa"], st["g"])
                for a, k in zip(c, ("ctau", "cssa", "cg")):
                    _eq(a, r[k], k)
                for a, k in zip(t, ("tau", "ssa", "g")):
                    _eq(a, r[k], k + "+cloud")


@pytest.mark.parametrize("top_at_1", [False, True])
def test_aerosol_optics_delta_add_matches_reference_classes(top_at_1):
    """Aerosol_optics::aerosol_optics (src/Aerosol_optics.cpp, compiled unmodified) -> delta_scale -> add_to.
    top_at_1=True pins the CPU source's signed layer mass (negative optical depths, :53)."""
    atm = add_aerosols(rce_atmosphere(11, 26, top_at_1=top_at_1))
    kd = synthetic_kdist("sw"); lut = synthetic_aerosol_lut()
    go = O.GasOptics(kd, atm.vmr.keys()); rf = R.RefGasOptics(kd, atm.vmr.keys())
    ao = O.AerosolOptics(lut)
    st = go.gas_optics_sw(atm.play, atm.plev, atm.tlay, atm.vmr)
    for dl in (False, True):
        a = ao.aerosol_optics(atm.aermr, atm.rh, atm.plev)
        assert np.all(a[0] < 0) if top_at_1 else np.all(a[0] > 0)
        if dl:
            a = O.delta_scale(*a)
        t = O.add_to_2str(go, st["tau"], st["ssa"], st["g"], *a)
        r = rf.aerosol_add(lut, dl, atm.aermr, atm.rh, atm.plev, st["tau"], st["ssa"], st["g"])
        for x, k in zip(a, ("atau", "assa", "ag")):
            _eq(x, r[k], k)
        for x, k in zip(t, ("tau", "ssa", "g")):
            _eq(x, r[k], k + "+aerosol")


def test_missing_key_gas_and_range_errors_match():
    kd = synthetic_kdist("lw")
    atm = rce_atmosphere(3, 10)
    gases = [g for g in atm.vmr.keys() if g != "co2"]
    with pytest.raises(RuntimeError):
        O.GasOptics(kd, gases)
    with pytest.raises(RuntimeError):
        R.RefGasOptics(kd, gases)
    go = O.GasOptics(kd, atm.vmr.keys()); rf = R.RefGasOptics(kd, atm.vmr.keys())
    bad = atm.tlay.copy(order="F"); bad[0, 0] = 400.0
    col_dry = go.get_col_dry(np.broadcast_to(atm.vmr["h2o"], atm.play.shape), atm.plev)
    with pytest.raises(ValueError):
        go.gas_optics_lw(atm.play, atm.plev, bad, atm.tsfc, atm.vmr, col_dry, atm.tlev)
    with pytest.raises(RuntimeError, match="tlay is out of range"):
        rf.gas_optics_lw(atm.play, atm.plev, bad, atm.tsfc, atm.vmr, col_dry, atm.tlev)

"""Real inputs of rrtmgp-data (committed fixture tests/golden/rrtmgp_data.npz): the RFMIP clear-sky atmospheres and
the cloud-optics LUTs the reference ships, through the numpy oracle and the reference's own compiled classes
(oracle/_ref).  The gas-optics tables stay synthetic (rrtmgp-gas-*.nc are absent from the mount)."""
import numpy as np
import pytest

import real_data as RD
from helpers import oracle_lw, oracle_sw
from oracle import oracle as O
from oracle import ref as R
from rte_rrtmgp_cpp_b200.atmos import allsky_atmosphere
from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist

Z = RD.fixture()
needs_ref = pytest.mark.skipif(not R.available(), reason="oracle/_ref/liboracle_ref.so not built")


@needs_ref
@pytest.mark.parametrize("expt", [0, 3])
def test_rfmip_columns_oracle_matches_reference_classes_bitwise(expt):
    atm = RD.rfmip_atmosphere(Z, expt, slice(0, 100, 9))          # 12 sites across the globe, day and night
    for kind in ("lw", "sw"):
        kd = synthetic_kdist(kind)
        go, rf = O.GasOptics(kd, atm.vmr.keys()), R.RefGasOptics(kd, atm.vmr.keys())
        col_dry = go.get_col_dry(np.broadcast_to(atm.vmr["h2o"], atm.play.shape), atm.plev)
        assert np.array_equal(col_dry, rf.get_col_dry(atm.vmr["h2o"], atm.plev))
        if kind == "lw":
            a = go.gas_optics_lw(atm.play, atm.plev, atm.tlay, atm.tsfc, atm.vmr, col_dry, atm.tlev)
            b = rf.gas_optics_lw(atm.play, atm.plev, atm.tlay, atm.tsfc, atm.vmr, col_dry, atm.tlev)
            keys = ("tau", "lay_src", "lev_src", "sfc_src")
        else:
            a = go.gas_optics_sw(atm.play, atm.plev, atm.tlay, atm.vmr, col_dry)
            b = rf.gas_optics_sw(atm.play, atm.plev, atm.tlay, atm.vmr, col_dry)
            keys = ("tau", "ssa", "g", "toa_src")
        for k in keys:
            assert np.array_equal(a[k], b[k]), (kind, k)
        if kind == "lw":
            o = O.rte_lw(go, a["tau"], a["lay_src"], a["lev_src"], a["sfc_src"], atm.emis_sfc, atm.top_at_1)
            r = rf.rte_lw(a["tau"], a["lay_src"], a["lev_src"], a["sfc_src"], atm.emis_sfc, atm.top_at_1, 1, False)
        else:
            o = O.rte_sw(go, a["tau"], a["ssa"], a["g"], atm.mu0, a["toa_src"], atm.sfc_alb_dir, atm.sfc_alb_dif, atm.top_at_1)
            r = rf.rte_sw(a["tau"], a["ssa"], a["g"], atm.mu0, a["toa_src"], atm.sfc_alb_dir, atm.sfc_alb_dif, atm.top_at_1, False)
        for x, y in zip(o, r):
            assert np.array_equal(x, y), kind


@needs_ref
def test_real_cloud_lut_oracle_matches_reference_classes_bitwise():
    atm = allsky_atmosphere(12, 30, random_clouds=True)
    for kind in ("lw", "sw"):
        kd, lut = synthetic_kdist(kind), RD.cloud_lut(Z, kind)
        go, rf = O.GasOptics(kd, atm.vmr.keys()), R.RefGasOptics(kd, atm.vmr.keys())
        co = O.CloudOptics(lut)
        if kind == "lw":
            st = go.gas_optics_lw(atm.play, atm.plev, atm.tlay, atm.tsfc, atm.vmr, tlev=atm.tlev)
            ctau = co.cloud_optics_1scl(atm.lwp, atm.iwp, atm.rel, atm.dei)
            r = rf.cloud_add(lut, 1, False, atm.lwp, atm.iwp, atm.rel, atm.dei, st["tau"])
            assert ctau.max() > 1.0 and np.array_equal(ctau, r["ctau"])
            assert np.array_equal(O.add_to_1scl(go, st["tau"], ctau), r["tau"])
        else:
            st = go.gas_optics_sw(atm.play, atm.plev, atm.tlay, atm.vmr)
            for dl in (False, True):
                c = co.cloud_optics_2str(atm.lwp, atm.iwp, atm.rel, atm.dei)
                assert c[1].max() > 0.99                      # the real tables are near-conservative in the visible
                if dl:
                    c = O.delta_scale(*c)
                t = O.add_to_2str(go, st["tau"], st["ssa"], st["g"], *c)
                r = rf.cloud_add(lut, 2, dl, atm.lwp, atm.iwp, atm.rel, atm.dei, st["tau"], st["ssa"], st["g"])
                for x, k in zip(c, ("ctau", "cssa", "cg")):
                    assert np.array_equal(x, r[k]), k
                for x, k in zip(t, ("tau", "ssa", "g")):
                    assert np.array_equal(x, r[k]), k


def test_rfmip_fluxes_are_physical():
    """Synthetic tables, real columns: not the published fluxes, but energy bounds must hold for every site."""
    atm = RD.rfmip_atmosphere(Z, 0)
    lw = oracle_lw(synthetic_kdist("lw"), atm)
    sw = oracle_sw(synthetic_kdist("sw"), atm)
    sigma = 5.670374419e-8
    assert np.all(lw["flux_dn"][:, 0] == 0.0)
    assert np.all(lw["flux_up"][:, -1] < 1.02 * sigma * atm.tsfc ** 4) and np.all(lw["flux_up"][:, -1] > 0.5 * sigma * atm.tsfc ** 4)
    night = atm.mu0 == 0
    assert np.all(sw["flux_dn"][night] == 0.0) and np.all(sw["flux_up"][~night, 0] < sw["flux_dn"][~night, 0])


@pytest.mark.parametrize("expt", [0, 3])
def test_rfmip_boundary_fluxes_match_the_published_files(expt):
    """What the published RFMIP fluxes pin WITHOUT the gas tables: the down-welling shortwave at the top is
    tsi * mu0 and the up-welling longwave at the surface is emissivity * Planck(t_sfc) + reflection, both independent of
    the gas optics to first order.  Checks the whole input recipe (site order, level orientation, units, mu0, tsi
    scaling, band-constant surface properties) against the reference's own published output."""
    ksw = synthetic_kdist("sw")
    atm = RD.rfmip_atmosphere(Z, expt, tsi_ref=RD.default_tsi(ksw))        # tsi_scaling = tsi / get_tsi(), as the driver does
    pub = {k: Z[f"rfmip/{expt}/pub_{k}"].astype(np.float64) for k in ("rld", "rlu", "rsd", "rsu")}
    assert np.all(pub["rld"][:, 0] == 0.0)                                   # level 0 is the top, as in the inputs
    sw = oracle_sw(ksw, atm)
    day = atm.mu0 > 0
    assert day.sum() > 40 and np.all(pub["rsd"][~day, 0] == 0.0) and np.all(sw["flux_dn"][~day, 0] == 0.0)
    assert np.max(np.abs(sw["flux_dn"][day, 0] / pub["rsd"][day, 0] - 1.0)) < 5e-7          # float32 storage of the published file
    lw = oracle_lw(synthetic_kdist("lw"), atm)
    assert np.max(np.abs(lw["flux_up"][:, -1] / pub["rlu"][:, -1] - 1.0)) < 0.01
    # and the synthetic tables are not the real ones: the interior differs (so the two checks above are not vacuous)
    assert np.max(np.abs(lw["flux_dn"][:, -1] - pub["rld"][:, -1])) > 10.0


def test_allsky_recipe_and_boundary_fluxes_match_the_published_example():
    """rte_rrtmgp_cpp_b200.atmos.allsky_atmosphere restates allsky/allsky_init.py; the all-sky example of rrtmgp-data
    publishes the same atmosphere (written by the Fortran example) next to its fluxes.  Inputs agree to rounding (the
    cloud fields exactly; only the top level pressure is floored differently by the two recipes), and the published
    fluxes' table-independent parts are reproduced with the REAL cloud LUT over synthetic gas tables."""
    g = lambda n: Z[f"allsky/{n}"].T                                          # (col, lay|lev), bottom first
    atm = allsky_atmosphere(24, 72)
    rel_err = lambda a, b: np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300))
    assert rel_err(atm.play, g("p_lay")) < 2e-6 and rel_err(atm.plev[:, :-1], g("p_lev")[:, :-1]) < 2e-6
    assert rel_err(atm.tlay, g("t_lay")) < 1e-7 and rel_err(atm.tlev, g("t_lev")) < 1e-7
    assert rel_err(np.broadcast_to(atm.vmr["h2o"], (24, 72)), g("h2o")) < 1e-7
    assert rel_err(np.broadcast_to(atm.vmr["o3"], (24, 72)), g("o3")) < 1e-5
    for ours, name in ((atm.lwp, "lwp"), (atm.iwp, "iwp"), (atm.rel, "rel"), (atm.dei, "rei")):
        assert np.array_equal(ours, g(name)), name
    assert (g("lwp") > 0).sum() > 50 and (g("iwp") > 0).sum() > 50
    sw = oracle_sw(synthetic_kdist("sw"), atm, RD.cloud_lut(Z, "sw"))
    assert np.max(np.abs(sw["flux_dn"][:, -1] / g("sw_flux_dn")[:, -1] - 1.0)) < 1e-4          # tsi * mu0 at the top
    assert np.max(np.abs(sw["flux_dn_dir"][:, -1] / g("sw_flux_dir")[:, -1] - 1.0)) < 1e-4
    lw = oracle_lw(synthetic_kdist("lw"), atm, RD.cloud_lut(Z, "lw"))
    assert np.max(np.abs(lw["flux_up"][:, 0] / g("lw_flux_up")[:, 0] - 1.0)) < 0.01            # emission at the surface

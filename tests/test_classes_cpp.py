"""The C++ drop-in class layer (include/rrtmgp_b200/classes.hpp) driven the way the reference's
test_rte_rrtmgp drives its classes: Radiation_solver_{longwave,shortwave}::solve_gpu, fused
(switch_output_optical=false) and staged (optical properties returned), against the CPU oracle.
The CPU-only test checks that the driver links and that, without a GPU, it throws instead of
computing anything on the host."""
import ctypes as C
import os
import sys

import numpy as np
import pytest
import torch

from rte_rrtmgp_cpp_b200 import capi, ctables
from rte_rrtmgp_cpp_b200.atmos import rce_atmosphere, allsky_atmosphere
from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist, synthetic_cloud_lut
from helpers import oracle_lw, oracle_sw, as_f32_valued

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "cpp"))
import build_driver  # noqa: E402

dp = C.POINTER(C.c_double)
TOL = {"f64": 1e-5, "f32": 0.05}


def _drv(tag):
    capi.lib()                       # librrtmgp_b200.so first (fails loudly if it is not built)
    build_driver.build()
    d = C.CDLL(build_driver.path(tag))
    assert d.classes_float_bytes() == (8 if tag == "f64" else 4)
    return d


def _F(a):
    return None if a is None else np.asfortranarray(a, dtype=np.float64)


def _p(a):
    return None if a is None else a.ctypes.data_as(dp)


def _concs(atm):
    names = list(atm.vmr.keys())
    arrs = [_F(np.atleast_2d(atm.vmr[n])) for n in names]
    ptrs = (dp * len(names))(*[_p(a) for a in arrs])
    nc = (C.c_int * len(names))(*[a.shape[0] for a in arrs])
    nl = (C.c_int * len(names))(*[a.shape[1] for a in arrs])
    return len(names), ctables.pack_strings(names, 32), ptrs, nc, nl, arrs


def _out(*shape):
    return np.zeros(shape, dtype=np.float64, order="F")


def run_lw(tag, kd, atm, lut=None, mode=0, bnd=1, n_ang=1):
    d = _drv(tag)
    kc, keep = ctables.pack_kdist(kd)
    cc, keep2 = ctables.pack_cloud(lut) if lut is not None else (None, None)
    ng, names, ptrs, nc, nl, arrs = _concs(atm)
    ncol, nlay, ngpt, nbnd = atm.ncol, atm.nlay, kd.ngpt, kd.nbnd
    ins = [_F(x) for x in (atm.play, atm.plev, atm.tlay, atm.tlev, atm.tsfc, atm.emis_sfc, atm.lwp, atm.iwp, atm.rel, atm.dei)]
    o = dict(tau=_out(ncol, nlay, ngpt), lay_src=_out(ncol, nlay, ngpt), lev_src=_out(ncol, nlay + 1, ngpt), sfc_src=_out(ncol, ngpt),
             flux_up=_out(ncol, nlay + 1), flux_dn=_out(ncol, nlay + 1), flux_net=_out(ncol, nlay + 1),
             bnd_flux_up=_out(ncol, nlay + 1, nbnd), bnd_flux_dn=_out(ncol, nlay + 1, nbnd), bnd_flux_net=_out(ncol, nlay + 1, nbnd))
    err = C.create_string_buffer(512)
    rc = d.classes_run_lw(C.byref(kc), C.byref(cc) if cc is not None else None, mode, bnd, n_ang, ncol, nlay, ng, names, ptrs, nc, nl,
                          *[_p(x) for x in ins], *[_p(x) for x in o.values()], err, 512)
    return rc, err.value.decode(), o


def run_sw(tag, kd, atm, lut=None, mode=0, bnd=1, delta=1):
    d = _drv(tag)
    kc, keep = ctables.pack_kdist(kd)
    cc, keep2 = ctables.pack_cloud(lut) if lut is not None else (None, None)
    ng, names, ptrs, nc, nl, arrs = _concs(atm)
    ncol, nlay, ngpt, nbnd = atm.ncol, atm.nlay, kd.ngpt, kd.nbnd
    ins = [_F(x) for x in (atm.play, atm.plev, atm.tlay, atm.sfc_alb_dir, atm.sfc_alb_dif, atm.tsi_scaling, atm.mu0,
                           atm.lwp, atm.iwp, atm.rel, atm.dei)]
    o = dict(tau=_out(ncol, nlay, ngpt), ssa=_out(ncol, nlay, ngpt), g=_out(ncol, nlay, ngpt), toa=_out(ncol, ngpt),
             flux_up=_out(ncol, nlay + 1), flux_dn=_out(ncol, nlay + 1), flux_dn_dir=_out(ncol, nlay + 1), flux_net=_out(ncol, nlay + 1),
             bnd_flux_up=_out(ncol, nlay + 1, nbnd), bnd_flux_dn=_out(ncol, nlay + 1, nbnd),
             bnd_flux_dn_dir=_out(ncol, nlay + 1, nbnd), bnd_flux_net=_out(ncol, nlay + 1, nbnd))
    err = C.create_string_buffer(512)
    rc = d.classes_run_sw(C.byref(kc), C.byref(cc) if cc is not None else None, mode, bnd, delta, ncol, nlay, ng, names, ptrs, nc, nl,
                          *[_p(x) for x in ins], *[_p(x) for x in o.values()], err, 512)
    return rc, err.value.decode(), o


def _cmp(o, ref, keys, tol):
    for k in keys:
        dmax = np.max(np.abs(o[k] - ref[k]))
        assert dmax <= tol, f"{k}: max abs diff {dmax} > {tol}"


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
@pytest.mark.parametrize("tag", ["f64", "f32"])
def test_driver_links_and_fails_loudly_without_gpu(tag):
    kd, atm = synthetic_kdist("lw"), rce_atmosphere(4, 8)
    rc, msg, o = run_lw(tag, kd, atm)
    assert rc == 1 and msg, "the class layer must throw without a CUDA device (no CPU fallback)"
    assert not o["flux_up"].any()


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["f64", "f32"])
@pytest.mark.parametrize("mode", [0, 1])
def test_lw_solve_gpu(tag, mode):
    kd, lut = synthetic_kdist("lw"), synthetic_cloud_lut("lw")
    atm = allsky_atmosphere(24, 72)
    ref = oracle_lw(kd, atm, lut)
    rc, msg, o = run_lw(tag, kd, atm, lut, mode=mode)
    assert rc == 0, msg
    _cmp(o, ref, ("flux_up", "flux_dn", "flux_net", "bnd_flux_up", "bnd_flux_dn", "bnd_flux_net"), TOL[tag])
    if mode == 1 and tag == "f64":
        np.testing.assert_allclose(o["tau"], ref["tau"], rtol=1e-12, atol=0)
        np.testing.assert_allclose(o["lay_src"], ref["state"]["lay_src"], rtol=1e-12, atol=0)
        np.testing.assert_allclose(o["lev_src"], ref["state"]["lev_src"], rtol=1e-12, atol=0)
        np.testing.assert_allclose(o["sfc_src"], ref["state"]["sfc_src"], rtol=1e-12, atol=0)


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["f64", "f32"])
@pytest.mark.parametrize("mode", [0, 1])
def test_sw_solve_gpu(tag, mode):
    kd, lut = synthetic_kdist("sw"), synthetic_cloud_lut("sw")
    atm = allsky_atmosphere(24, 72)
    atm.tsi_scaling[:] = np.linspace(0.95, 1.05, atm.ncol)
    ref = oracle_sw(kd, atm, lut)
    rc, msg, o = run_sw(tag, kd, atm, lut, mode=mode)
    assert rc == 0, msg
    _cmp(o, ref, ("flux_up", "flux_dn", "flux_dn_dir", "flux_net", "bnd_flux_up", "bnd_flux_dn", "bnd_flux_dn_dir", "bnd_flux_net"), TOL[tag])
    if mode == 1 and tag == "f64":
        np.testing.assert_allclose(o["tau"], ref["tau"], rtol=1e-12, atol=0)
        np.testing.assert_allclose(o["ssa"], ref["ssa"], rtol=1e-11, atol=1e-15)
        np.testing.assert_allclose(o["g"], ref["g"], rtol=1e-11, atol=1e-15)
        np.testing.assert_allclose(o["toa"], ref["toa"], rtol=1e-14, atol=0)


@pytest.mark.gpu
def test_clear_sky_no_bands_and_range_error():
    kd = synthetic_kdist("lw")
    atm = rce_atmosphere(40, 32, top_at_1=True)
    ref = oracle_lw(kd, atm, n_ang=3)
    rc, msg, o = run_lw("f64", kd, atm, mode=1, bnd=0, n_ang=3)
    assert rc == 0, msg
    _cmp(o, ref, ("flux_up", "flux_dn", "flux_net"), 1e-5)
    assert not o["bnd_flux_up"].any()
    # staged path validates ranges like the CPU reference (src/Gas_optics_rrtmgp.cpp:812-827)
    bad = rce_atmosphere(8, 16)
    bad.tlay[3, 5] = 1000.0
    rc, msg, _ = run_lw("f64", kd, bad, mode=1)
    assert rc == 2 and "out of range" in msg


# ---- set-up half of the boundary: host Array / Array_gpu and the reference's positional constructors -----------------
def _ctor(tag):
    capi.lib()
    build_driver.build()
    d = C.CDLL(build_driver.ctor_path(tag))
    d.ctor_last_message.restype = C.c_char_p
    assert d.ctor_float_bytes() == (8 if tag == "f64" else 4)
    return d


def _f32_tables(t):
    """An RTE_USE_SP caller holds the file variables as float Arrays: the same coefficients rounded to fp32 (kept as fp64),
    so the struct constructor and the positional constructor are handed identical values."""
    import copy
    t = copy.copy(t)
    for k, v in vars(t).items():
        if isinstance(v, np.ndarray) and v.dtype == np.float64:
            setattr(t, k, v.astype(np.float32).astype(np.float64))
        elif isinstance(v, float):
            setattr(t, k, float(np.float32(v)))
    return t


@pytest.mark.parametrize("tag", ["f64", "f32"])
def test_host_array_has_the_reference_semantics(tag):
    """include/Array.h:94-297: column-major, 1-based operator(), offsets, subset with broadcast, move leaves the source
    empty, set_dims / expand_dims errors; Gas_concs scalar / profile / field and column subset (no device needed)."""
    d = _ctor(tag)
    assert d.ctor_array_host_selftest() == 0, d.ctor_last_message().decode()


@pytest.mark.parametrize("tag", ["f64", "f32"])
@pytest.mark.parametrize("kind", ["lw", "sw"])
def test_positional_constructor_builds_identical_tables(kind, tag):
    """Gas_optics_rrtmgp_gpu through the reference's 32 / 35 positional Array arguments (include/Gas_optics_rrtmgp.h:
    293-364), fed like load_and_init_gas_optics feeds them, against the struct constructor: every prepared table
    (hash over coefficients in device layout, index tables, grids, retained gases) bit-identical.  Host-only handles."""
    d = _ctor(tag)
    kd = _f32_tables(synthetic_kdist(kind)) if tag == "f32" else synthetic_kdist(kind)
    kc, keep = ctables.pack_kdist(kd)
    for gases in (["h2o", "co2", "o3", "n2o", "ch4", "n2", "o2"], ["h2o", "co2", "o3", "n2o", "ch4", "n2", "o2", "co", "cfc11", "no2"]):
        rc = d.ctor_tables_equal(C.byref(kc), len(gases), ctables.pack_strings(gases, 32), 32, 1)
        assert rc == 0, d.ctor_last_message().decode()
    # a missing key gas is refused with the reference's message through either constructor
    rc = d.ctor_tables_equal(C.byref(kc), 2, ctables.pack_strings(["h2o", "o3"], 32), 32, 1)
    assert rc == 1 and "missing" in d.ctor_last_message().decode()


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["f64", "f32"])
def test_array_gpu_has_the_reference_semantics(tag):
    """include/Array.h:355-650 on the device: Array <-> Array_gpu conversions, views, deep copies, moves, the subset kernel
    against the host subset (with a broadcast dimension and a 1-byte type), fill, insert, operator(), Gas_concs_gpu subset."""
    d = _ctor(tag)
    assert d.ctor_array_gpu_selftest() == 0, d.ctor_last_message().decode()


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["f64", "f32"])
def test_positional_constructors_solve_identically(tag):
    """All-sky LW + SW (clouds, aerosols) with every optics object built through the reference's positional constructors
    (Gas_optics_rrtmgp_gpu, Cloud_optics_gpu, Aerosol_optics_gpu) against the struct-constructed solvers: fluxes bit-identical,
    and within tolerance of the oracle."""
    from rte_rrtmgp_cpp_b200.atmos import add_aerosols
    from rte_rrtmgp_cpp_b200.kdist import synthetic_aerosol_lut
    d = _ctor(tag)
    kd_lw, kd_sw = synthetic_kdist("lw"), synthetic_kdist("sw")
    lut_lw, lut_sw, alut = synthetic_cloud_lut("lw"), synthetic_cloud_lut("sw"), synthetic_aerosol_lut()
    if tag == "f32":
        kd_lw, kd_sw, lut_lw, lut_sw, alut = (_f32_tables(t) for t in (kd_lw, kd_sw, lut_lw, lut_sw, alut))
    atm = add_aerosols(allsky_atmosphere(24, 72))
    klw, k1 = ctables.pack_kdist(kd_lw); ksw, k2 = ctables.pack_kdist(kd_sw)
    clw, k3 = ctables.pack_cloud(lut_lw); csw, k4 = ctables.pack_cloud(lut_sw); aer, k5 = ctables.pack_aerosol(alut)
    ng, names, ptrs, nc, nl, arrs = _concs(atm)
    ncol, nlay = atm.ncol, atm.nlay
    ins = [_F(x) for x in (atm.play, atm.plev, atm.tlay, atm.tlev, atm.tsfc, atm.emis_sfc, atm.sfc_alb_dir, atm.sfc_alb_dif, atm.mu0,
                           atm.lwp, atm.iwp, atm.rel, atm.dei, atm.rh)]
    am = [_F(np.broadcast_to(atm.aermr[f"aermr{i:02d}"], (ncol, nlay))) for i in range(1, 12)]
    amp = (dp * 11)(*[_p(x) for x in am])
    res = {}
    for positional in (0, 1):
        o = [_out(ncol, nlay + 1) for _ in range(5)]
        rc = d.ctor_solve(C.byref(klw), C.byref(ksw), C.byref(clw), C.byref(csw), C.byref(aer), positional, ncol, nlay, ng, names, ptrs, nc, nl,
                          *[_p(x) for x in ins], amp, *[_p(x) for x in o])
        assert rc == 0, d.ctor_last_message().decode()
        res[positional] = o
    for a, b in zip(res[0], res[1]):
        assert a.any() and np.array_equal(a, b)
    ref_lw = oracle_lw(kd_lw, atm, lut_lw)
    ref_sw = oracle_sw(kd_sw, atm, lut_sw, delta=True, aerosol_lut=alut, delta_aerosol=False)
    for got, want in zip(res[1], (ref_lw["flux_up"], ref_lw["flux_dn"], ref_sw["flux_up"], ref_sw["flux_dn"], ref_sw["flux_dn_dir"])):
        assert np.max(np.abs(got - want)) <= TOL[tag]

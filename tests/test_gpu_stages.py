"""GPU parity of the materialised (class-API call-by-call) stages against the CPU oracle.
Index outputs (jtemp, jpress, jeta, tropo) must be bit-exact; fp64 stage outputs agree to
round-off; the fp32 build is compared with the fp64 oracle fed the same fp32-valued inputs."""
import numpy as np
import pytest
import torch

from oracle import oracle as O
from rte_rrtmgp_cpp_b200 import api
from rte_rrtmgp_cpp_b200.api import Array_gpu as A
from rte_rrtmgp_cpp_b200.atmos import rce_atmosphere, allsky_atmosphere
from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist, synthetic_cloud_lut, REDUCED_LW_128
from helpers import oracle_lw, oracle_sw, as_f32_valued

pytestmark = pytest.mark.gpu


def rel(a, b):
    return np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300))


def _adversarial(atm, kd):
    """put some cells exactly on temperature / pressure grid points"""
    n = min(atm.ncol, kd.temp_ref.size)
    atm.tlay[:n, 1] = kd.temp_ref[:n]
    m = min(atm.ncol, 20)
    atm.play[:m, 2] = kd.press_ref[5:5 + m]
    atm.play[0, 3] = kd.press_ref_trop
    return atm


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("kind", ["lw", "sw"])
def test_interpolation_indices_bit_exact(dtype, kind):
    kd = synthetic_kdist(kind)
    atm = _adversarial(rce_atmosphere(96, 40), kd)
    if dtype == torch.float32:
        atm = as_f32_valued(atm)
    go = api.Gas_optics_rrtmgp_gpu(kd, atm.vmr.keys(), dtype)
    cols = api.DeviceColumns(atm, dtype)
    col_dry = A((atm.ncol, atm.nlay), dtype)
    go.get_col_dry(col_dry, cols.gas_concs.get_vmr("h2o"), cols.plev)
    cd = col_dry.numpy().astype(np.float64)
    ref_cd = O.GasOptics.get_col_dry(atm.vmr["h2o"], atm.plev)
    assert rel(cd, ref_cd) < (1e-14 if dtype == torch.float64 else 1e-6)
    out = go.interpolation(cols.play, cols.tlay, col_dry, cols.gas_concs)
    og = O.GasOptics(kd, atm.vmr.keys())
    st = og._gas_taus(O.F(atm.play), O.F(atm.plev), O.F(atm.tlay), atm.vmr, O.F(cd))
    for k in ("jtemp", "jpress", "jeta"):
        assert np.array_equal(out[k].numpy(), st[k]), k
    assert np.array_equal(out["tropo"].numpy() != 0, st["tropo"] != 0)
    tol = 1e-12 if dtype == torch.float64 else 1e-6
    for k in ("col_mix", "col_gas"):
        assert rel(out[k].numpy(), st[k]) < tol, k
    for k in ("fminor", "fmajor"):
        assert np.max(np.abs(out[k].numpy() - st[k])) < tol, k


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("sizes", [16, REDUCED_LW_128])
def test_gas_optics_lw_stage(dtype, sizes):
    kd = synthetic_kdist("lw", sizes)
    atm = rce_atmosphere(50, 33, top_at_1=(sizes == 16))
    if dtype == torch.float32:
        atm = as_f32_valued(atm)
    ref = oracle_lw(kd, atm)["state"]
    go = api.Gas_optics_rrtmgp_gpu(kd, atm.vmr.keys(), dtype)
    cols = api.DeviceColumns(atm, dtype)
    op = api.Optical_props_1scl_gpu(atm.ncol, atm.nlay, go, dtype)
    src = api.Source_func_lw_gpu(atm.ncol, atm.nlay, go, dtype)
    go.gas_optics(cols.play, cols.plev, cols.tlay, cols.tsfc, cols.gas_concs, op, src, None, cols.tlev)
    tol = 1e-11 if dtype == torch.float64 else 2e-4
    assert rel(op.tau.numpy(), ref["tau"]) < tol
    for k, a in (("lay_src", src.lay_source), ("lev_src", src.lev_source), ("sfc_src", src.sfc_source),
                 ("sfc_src_jac", src.sfc_source_jac)):
        scale = np.abs(ref[k]).max()
        assert np.max(np.abs(a.numpy() - ref[k])) < tol * scale * (50 if k == "sfc_src_jac" else 1), k


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_gas_optics_sw_stage(dtype):
    kd = synthetic_kdist("sw")
    atm = rce_atmosphere(50, 33)
    if dtype == torch.float32:
        atm = as_f32_valued(atm)
    ref = oracle_sw(kd, atm)["state"]
    go = api.Gas_optics_rrtmgp_gpu(kd, atm.vmr.keys(), dtype)
    cols = api.DeviceColumns(atm, dtype)
    op = api.Optical_props_2str_gpu(atm.ncol, atm.nlay, go, dtype)
    toa = A((atm.ncol, go.get_ngpt()), dtype)
    go.gas_optics(cols.play, cols.plev, cols.tlay, cols.gas_concs, op, toa, None)
    tol = 1e-11 if dtype == torch.float64 else 2e-4
    assert rel(op.tau.numpy(), ref["tau"]) < tol
    assert np.max(np.abs(op.ssa.numpy() - ref["ssa"])) < tol
    assert np.all(op.g.numpy() == 0.0)
    assert rel(toa.numpy(), ref["toa_src"]) < (1e-15 if dtype == torch.float64 else 1e-7)


def test_range_errors_match_reference_messages():
    kd = synthetic_kdist("lw")
    atm = rce_atmosphere(5, 10)
    atm.tlay[0, 0] = 400.0
    go = api.Gas_optics_rrtmgp_gpu(kd, atm.vmr.keys())
    cols = api.DeviceColumns(atm)
    op = api.Optical_props_1scl_gpu(atm.ncol, atm.nlay, go)
    src = api.Source_func_lw_gpu(atm.ncol, atm.nlay, go)
    with pytest.raises(api.capi.RangeError, match="tlay is out of range"):
        go.gas_optics(cols.play, cols.plev, cols.tlay, cols.tsfc, cols.gas_concs, op, src, None, cols.tlev)
    g = api.Gas_concs_gpu()
    with pytest.raises(api.capi.RangeError):
        g.set_vmr("h2o", 1.5)


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("top_at_1", [False, True])
def test_rte_lw_and_reduce(dtype, top_at_1):
    kd = synthetic_kdist("lw")
    atm = rce_atmosphere(37, 25, top_at_1=top_at_1)
    ref = oracle_lw(kd, atm)
    og, st = ref["go"], ref["state"]
    go = api.Gas_optics_rrtmgp_gpu(kd, atm.vmr.keys(), dtype)
    op = api.Optical_props_1scl_gpu(atm.ncol, atm.nlay, go, dtype)
    src = api.Source_func_lw_gpu(atm.ncol, atm.nlay, go, dtype)
    op.tau = A.from_numpy(st["tau"], dtype)
    src.lay_source, src.lev_source = A.from_numpy(st["lay_src"], dtype), A.from_numpy(st["lev_src"], dtype)
    src.sfc_source, src.sfc_source_jac = A.from_numpy(st["sfc_src"], dtype), A.from_numpy(st["sfc_src_jac"], dtype)
    emis = A.from_numpy(atm.emis_sfc, dtype)
    nlev, ngpt = atm.nlay + 1, go.get_ngpt()
    tol = 1e-10 if dtype == torch.float64 else 5e-4
    for n_ang in (1, 2, 3, 4):
        up, dn = A((atm.ncol, nlev, ngpt), dtype), A((atm.ncol, nlev, ngpt), dtype)
        api.Rte_lw_gpu.rte_lw(op, top_at_1, src, emis, None, up, dn, n_ang)
        rup, rdn = O.rte_lw(og, st["tau"], st["lay_src"], st["lev_src"], st["sfc_src"], atm.emis_sfc, top_at_1,
                            n_gauss_angles=n_ang)
        assert np.max(np.abs(up.numpy() - rup)) < tol and np.max(np.abs(dn.numpy() - rdn)) < tol
        bup, bdn = A((atm.ncol, nlev), dtype), A((atm.ncol, nlev), dtype)
        api.Rte_lw_gpu.rte_lw(op, top_at_1, src, emis, None, bup, bdn, n_ang)
        assert np.max(np.abs(bup.numpy() - O.sum_broadband(rup))) < tol * 100
    # Fluxes::reduce on the oracle's own g-point fluxes: sequential sum -> bit exact in fp64
    up, dn = A.from_numpy(ref["gpt_up"], dtype), A.from_numpy(ref["gpt_dn"], dtype)
    fl = api.Fluxes_byband_gpu(atm.ncol, nlev, go.get_nband(), dtype)
    fl.reduce(up, dn, go, top_at_1)
    for k in ("flux_up", "flux_dn", "flux_net", "bnd_flux_up", "bnd_flux_dn", "bnd_flux_net"):
        got = getattr(fl, k).numpy()
        if dtype == torch.float64:
            assert np.array_equal(got, ref[k]), k
        else:
            assert np.max(np.abs(got - ref[k])) < 5e-3, k
    # Jacobian of the upward flux to surface temperature (do_Jacobians path, F90:196-215)
    rup, rdn, rjac = O.rte_lw(og, st["tau"], st["lay_src"], st["lev_src"], st["sfc_src"], atm.emis_sfc, top_at_1,
                              sfc_src_jac=st["sfc_src_jac"])
    up, dn, jac = A((atm.ncol, nlev, ngpt), dtype), A((atm.ncol, nlev, ngpt), dtype), A((atm.ncol, nlev), dtype)
    api.Rte_lw_gpu.rte_lw(op, top_at_1, src, emis, None, up, dn, 1, src.sfc_source_jac, jac)
    assert np.max(np.abs(jac.numpy() - rjac)) < (1e-10 if dtype == torch.float64 else 1e-4)


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("top_at_1", [False, True])
def test_rte_sw_stage(dtype, top_at_1):
    kd = synthetic_kdist("sw")
    atm = rce_atmosphere(37, 25, top_at_1=top_at_1, night_fraction=0.2, cloudy=True)
    lut = synthetic_cloud_lut("sw")
    ref = oracle_sw(kd, atm, lut)
    go = api.Gas_optics_rrtmgp_gpu(kd, atm.vmr.keys(), dtype)
    op = api.Optical_props_2str_gpu(atm.ncol, atm.nlay, go, dtype)
    op.tau, op.ssa, op.g = (A.from_numpy(ref[k], dtype) for k in ("tau", "ssa", "g"))
    nlev, ngpt = atm.nlay + 1, go.get_ngpt()
    up, dn, dr = (A((atm.ncol, nlev, ngpt), dtype) for _ in range(3))
    api.Rte_sw_gpu.rte_sw(op, top_at_1, A.from_numpy(atm.mu0, dtype), A.from_numpy(ref["toa"], dtype),
                          A.from_numpy(atm.sfc_alb_dir, dtype), A.from_numpy(atm.sfc_alb_dif, dtype), None, up, dn, dr)
    tol = 1e-10 if dtype == torch.float64 else 2e-3
    for got, k in ((up, "gpt_up"), (dn, "gpt_dn"), (dr, "gpt_dir")):
        assert np.max(np.abs(got.numpy() - ref[k])) < tol, k
    bup, bdn, bdr = (A((atm.ncol, nlev), dtype) for _ in range(3))
    api.Rte_sw_gpu.rte_sw(op, top_at_1, A.from_numpy(atm.mu0, dtype), A.from_numpy(ref["toa"], dtype),
                          A.from_numpy(atm.sfc_alb_dir, dtype), A.from_numpy(atm.sfc_alb_dif, dtype), None, bup, bdn, bdr)
    btol = 1e-9 if dtype == torch.float64 else 0.05
    for got, k in ((bup, "flux_up"), (bdn, "flux_dn"), (bdr, "flux_dn_dir")):
        assert np.max(np.abs(got.numpy() - ref[k])) < btol, k


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_cloud_optics_delta_scale_add_to(dtype):
    atm = allsky_atmosphere(12, 30, random_clouds=True)
    if dtype == torch.float32:
        atm = as_f32_valued(atm)
    cols = api.DeviceColumns(atm, dtype)
    tol = 1e-12 if dtype == torch.float64 else 1e-5
    for kind in ("lw", "sw"):
        kd, lut = synthetic_kdist(kind), synthetic_cloud_lut(kind)
        og, oc = O.GasOptics(kd, atm.vmr.keys()), O.CloudOptics(lut)
        go = api.Gas_optics_rrtmgp_gpu(kd, atm.vmr.keys(), dtype)
        co = api.Cloud_optics_gpu(lut, dtype)
        rng = np.random.default_rng(3)
        shape = (atm.ncol, atm.nlay, go.get_ngpt())
        if kind == "lw":
            cl = api.Optical_props_1scl_gpu(atm.ncol, atm.nlay, co, dtype)
            co.cloud_optics(cols.lwp, cols.iwp, cols.rel, cols.dei, cl)
            rc = oc.cloud_optics_1scl(atm.lwp, atm.iwp, atm.rel, atm.dei)
            assert np.max(np.abs(cl.tau.numpy() - rc)) < tol * max(1.0, rc.max())
            gas = rng.uniform(0.0, 2.0, shape)
            op = api.Optical_props_1scl_gpu(atm.ncol, atm.nlay, go, dtype)
            op.tau = A.from_numpy(gas, dtype)
            api.add_to(op, cl)
            gas64 = A.from_numpy(gas, dtype).numpy().astype(np.float64)
            want = O.add_to_1scl(og, gas64, cl.tau.numpy().astype(np.float64))
            assert np.max(np.abs(op.tau.numpy() - want)) < tol * 10
        else:
            cl = api.Optical_props_2str_gpu(atm.ncol, atm.nlay, co, dtype)
            co.cloud_optics(cols.lwp, cols.iwp, cols.rel, cols.dei, cl)
            rc = oc.cloud_optics_2str(atm.lwp, atm.iwp, atm.rel, atm.dei)
            for got, want in zip((cl.tau, cl.ssa, cl.g), rc):
                assert np.max(np.abs(got.numpy() - want)) < tol * max(1.0, want.max())
            cl.delta_scale()
            rd = O.delta_scale(*rc)
            for got, want in zip((cl.tau, cl.ssa, cl.g), rd):
                assert np.max(np.abs(got.numpy() - want)) < tol * 20 * max(1.0, want.max())
            gas = [rng.uniform(0.0, 2.0, shape), rng.uniform(0.0, 1.0, shape), rng.uniform(-0.2, 0.9, shape)]
            op = api.Optical_props_2str_gpu(atm.ncol, atm.nlay, go, dtype)
            op.tau, op.ssa, op.g = (A.from_numpy(x, dtype) for x in gas)
            gas64 = [A.from_numpy(x, dtype).numpy().astype(np.float64) for x in gas]
            cl64 = [x.numpy().astype(np.float64) for x in (cl.tau, cl.ssa, cl.g)]
            api.add_to(op, cl)
            want = O.add_to_2str(og, *gas64, *cl64)
            for got, w in zip((op.tau, op.ssa, op.g), want):
                assert np.max(np.abs(got.numpy() - w)) < tol * 20
            # same-resolution increment
            op2 = api.Optical_props_2str_gpu(atm.ncol, atm.nlay, go, dtype)
            op2.tau, op2.ssa, op2.g = (A.from_numpy(x, dtype) for x in gas)
            api.add_to(op, op2)
            want2 = O.add_to_2str(og, *want, *gas64)
            for got, w in zip((op.tau, op.ssa, op.g), want2):
                assert np.max(np.abs(got.numpy() - w)) < tol * 40
            with pytest.raises(RuntimeError, match="incompatible band"):
                bad = api.Optical_props_2str_gpu(atm.ncol, atm.nlay, api.Optical_props_gpu(lut.bnd_limits_wavenumber.T[:, :5]), dtype)
                api.add_to(op, bad)


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("top_at_1", [False, True])
def test_rte_lw_rescaled_stage(dtype, top_at_1):
    """do_rescaling = true of rte_lw_solver_noscat (mo_rte_solver_kernels.F90:148-178,753-844): 2-stream
    cloud properties in the longwave, 1 and 3 angles, g-point and broadband outputs, with the Jacobian."""
    kd, lut = synthetic_kdist("lw"), synthetic_cloud_lut("lw")
    atm = rce_atmosphere(33, 27, top_at_1=top_at_1, cloudy=True)
    ref = oracle_lw(kd, atm)
    og, st = ref["go"], ref["state"]
    c = O.CloudOptics(lut).cloud_optics_2str(atm.lwp, atm.iwp, atm.rel, atm.dei)
    tau, ssa, g = O.add_to_2str(og, st["tau"], np.zeros_like(st["tau"]), np.zeros_like(st["tau"]), *c)
    assert ssa.max() > 0.3                                   # the clouds do scatter
    go = api.Gas_optics_rrtmgp_gpu(kd, atm.vmr.keys(), dtype)
    op = api.Optical_props_2str_gpu(atm.ncol, atm.nlay, go, dtype)
    op.tau, op.ssa, op.g = A.from_numpy(tau, dtype), A.from_numpy(ssa, dtype), A.from_numpy(g, dtype)
    src = api.Source_func_lw_gpu(atm.ncol, atm.nlay, go, dtype)
    src.lay_source, src.lev_source = A.from_numpy(st["lay_src"], dtype), A.from_numpy(st["lev_src"], dtype)
    src.sfc_source, src.sfc_source_jac = A.from_numpy(st["sfc_src"], dtype), A.from_numpy(st["sfc_src_jac"], dtype)
    emis = A.from_numpy(atm.emis_sfc, dtype)
    nlev, ngpt = atm.nlay + 1, go.get_ngpt()
    tol = 1e-10 if dtype == torch.float64 else 1e-3
    noscat_up, _ = O.rte_lw(og, tau, st["lay_src"], st["lev_src"], st["sfc_src"], atm.emis_sfc, top_at_1)
    for n_ang in (1, 3):
        rup, rdn, rjac = O.rte_lw(og, tau, st["lay_src"], st["lev_src"], st["sfc_src"], atm.emis_sfc, top_at_1,
                                  n_gauss_angles=n_ang, sfc_src_jac=st["sfc_src_jac"], ssa=ssa, g=g)
        up, dn, jac = A((atm.ncol, nlev, ngpt), dtype), A((atm.ncol, nlev, ngpt), dtype), A((atm.ncol, nlev), dtype)
        up.t.fill_(-7.0); dn.t.fill_(-7.0)
        api.Rte_lw_gpu.rte_lw(op, top_at_1, src, emis, None, up, dn, n_ang, src.sfc_source_jac, jac, do_rescaling=True)
        assert np.max(np.abs(up.numpy() - rup)) < tol and np.max(np.abs(dn.numpy() - rdn)) < tol
        assert np.max(np.abs(jac.numpy() - rjac)) < (1e-10 if dtype == torch.float64 else 1e-4)
        bup, bdn = A((atm.ncol, nlev), dtype), A((atm.ncol, nlev), dtype)
        api.Rte_lw_gpu.rte_lw(op, top_at_1, src, emis, None, bup, bdn, n_ang, do_rescaling=True)
        assert np.max(np.abs(bup.numpy() - O.sum_broadband(rup))) < tol * 100
        assert np.max(np.abs(bdn.numpy() - O.sum_broadband(rdn))) < tol * 100
    assert np.max(np.abs(O.rte_lw(og, tau, st["lay_src"], st["lev_src"], st["sfc_src"], atm.emis_sfc, top_at_1,
                                  ssa=ssa, g=g)[0] - noscat_up)) > 1e-3          # and the option is not a no-op
    # without scattering the re-scaled kernel gives the noscat kernel's numbers (check_variants.F90:381-397)
    op.ssa.t.zero_()
    up0, dn0 = A((atm.ncol, nlev, ngpt), dtype), A((atm.ncol, nlev, ngpt), dtype)
    api.Rte_lw_gpu.rte_lw(op, top_at_1, src, emis, None, up0, dn0, 1)
    up1, dn1 = A((atm.ncol, nlev, ngpt), dtype), A((atm.ncol, nlev, ngpt), dtype)
    api.Rte_lw_gpu.rte_lw(op, top_at_1, src, emis, None, up1, dn1, 1, do_rescaling=True)
    assert np.max(np.abs(up0.numpy() - up1.numpy())) < tol and np.max(np.abs(dn0.numpy() - dn1.numpy())) < tol
    # 1-scalar properties cannot be re-scaled
    op1 = api.Optical_props_1scl_gpu(atm.ncol, atm.nlay, go, dtype)
    with pytest.raises(RuntimeError, match="2-stream"):
        api.Rte_lw_gpu.rte_lw(op1, top_at_1, src, emis, None, up0, dn0, 1, do_rescaling=True)


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("top_at_1", [False, True])
def test_rte_sw_height_dependent_mu0(dtype, top_at_1):
    """mu0 (ncol, nlay), the kernel's own shape (mo_rte_solver_kernels.F90:517,1061-1065,1117-1125):
    twilight columns whose lower layers have the sun below the local horizon and, in the fp64 build, a column dark at
    the top only.  The Fortran sends the beam through dark layers with the nominal mu0 = sqrt(epsilon(wp)), which is
    1.5e-8 in a double build and 3.5e-4 in a single build: the direct flux INSIDE dark layers (and a beam entering
    through a dark top, which is negative) depends on the build's precision in the reference itself, so the fp32
    build is compared on what does not: diffuse fluxes everywhere, the direct flux outside the twilight columns."""
    f64 = dtype == torch.float64
    kd = synthetic_kdist("sw")
    atm = rce_atmosphere(29, 23, top_at_1=top_at_1, night_fraction=0.1, cloudy=True)
    ref = oracle_sw(kd, atm, synthetic_cloud_lut("sw"))
    og = ref["go"]
    rng = np.random.default_rng(11)
    mu = np.tile(atm.mu0[:, None], (1, atm.nlay)) * rng.uniform(0.8, 1.2, (atm.ncol, atm.nlay))
    mu = np.minimum(mu, 1.0)
    low = slice(atm.nlay - 6, atm.nlay) if top_at_1 else slice(0, 6)
    top = slice(0, 3) if top_at_1 else slice(atm.nlay - 3, atm.nlay)
    mu[3:9, low] = -0.1
    if f64:
        mu[12, top] = -0.02
    mu = np.asfortranarray(mu)
    want = O.rte_sw(og, ref["tau"], ref["ssa"], ref["g"], mu, ref["toa"], atm.sfc_alb_dir, atm.sfc_alb_dif, top_at_1)
    go = api.Gas_optics_rrtmgp_gpu(kd, atm.vmr.keys(), dtype)
    op = api.Optical_props_2str_gpu(atm.ncol, atm.nlay, go, dtype)
    op.tau, op.ssa, op.g = (A.from_numpy(ref[k], dtype) for k in ("tau", "ssa", "g"))
    nlev, ngpt = atm.nlay + 1, go.get_ngpt()
    up, dn, dr = (A((atm.ncol, nlev, ngpt), dtype) for _ in range(3))
    args = (A.from_numpy(ref["toa"], dtype), A.from_numpy(atm.sfc_alb_dir, dtype), A.from_numpy(atm.sfc_alb_dif, dtype), None)
    api.Rte_sw_gpu.rte_sw(op, top_at_1, A.from_numpy(mu, dtype), *args, up, dn, dr)
    tol = 1e-10 if f64 else 2e-3
    gu, gd, gr = up.numpy(), dn.numpy(), dr.numpy()
    assert np.max(np.abs(gu - want[0])) < tol, "up"
    assert np.max(np.abs((gd - gr) - (want[1] - want[2]))) < tol, "diffuse dn"
    cols = slice(None) if f64 else np.r_[0:3, 9:atm.ncol]
    assert np.max(np.abs(gr[cols] - want[2][cols])) < tol, "dir"
    assert np.max(np.abs(gd[cols] - want[1][cols])) < tol, "dn"
    assert np.max(np.abs(want[0] - ref["gpt_up"])) > 1e-2          # the layer values matter
    # a constant field is the per-column call, bit for bit
    const = np.asfortranarray(np.tile(atm.mu0[:, None], (1, atm.nlay)))
    up2, dn2, dr2 = (A((atm.ncol, nlev, ngpt), dtype) for _ in range(3))
    api.Rte_sw_gpu.rte_sw(op, top_at_1, A.from_numpy(const, dtype), *args, up2, dn2, dr2)
    up1, dn1, dr1 = (A((atm.ncol, nlev, ngpt), dtype) for _ in range(3))
    api.Rte_sw_gpu.rte_sw(op, top_at_1, A.from_numpy(atm.mu0, dtype), *args, up1, dn1, dr1)
    assert torch.equal(up1.t, up2.t) and torch.equal(dn1.t, dn2.t) and torch.equal(dr1.t, dr2.t)
    with pytest.raises(RuntimeError, match="mu0 must be"):
        api.Rte_sw_gpu.rte_sw(op, top_at_1, A.from_numpy(const[:, :-1], dtype), *args, up2, dn2, dr2)


def test_cloud_particle_size_below_the_table_does_not_read_before_it():
    """Liquid radius below radliq_lwr / ice diameter 0 in a cloudy cell: the reference clamps the LUT index from above only
    (src/Cloud_optics.cpp:86) and would read before the table; here the index is clamped to the first interval, i.e. the
    value is the linear extrapolation of the first two table entries (and equals the oracle for every in-range cell)."""
    dtype = torch.float64
    lut = synthetic_cloud_lut("lw")
    atm = allsky_atmosphere(6, 10, random_clouds=True)
    atm.lwp[:] = 20.0; atm.iwp[:] = 0.0
    atm.rel[:] = lut.radliq_lwr + 3.3                      # in range everywhere ...
    step0 = (lut.radliq_upr - lut.radliq_lwr) / (lut.lut_extliq.shape[1] - 1)
    atm.rel[2, 4] = lut.radliq_lwr - 2.5 * step0           # ... except one cell 2.5 table steps below it (index -1 unclamped)
    atm.dei[:] = 0.0                                       # no ice water: the ice branch must not matter
    cols = api.DeviceColumns(atm, dtype)
    co = api.Cloud_optics_gpu(lut, dtype)
    cl = api.Optical_props_1scl_gpu(atm.ncol, atm.nlay, co, dtype)
    co.cloud_optics(cols.lwp, cols.iwp, cols.rel, cols.dei, cl)
    got = cl.tau.numpy()
    assert np.isfinite(got).all()
    want = O.CloudOptics(lut).cloud_optics_1scl(atm.lwp, atm.iwp, atm.rel, atm.dei)
    mask = np.ones(got.shape[:2], bool); mask[2, 4] = False
    assert np.max(np.abs(got[mask] - want[mask])) < 1e-12 * max(1.0, want[mask].max())
    nl = lut.lut_extliq.shape[1]
    step = (lut.radliq_upr - lut.radliq_lwr) / (nl - 1)
    x = (atm.rel[2, 4] - lut.radliq_lwr) / step            # negative
    for b in range(lut.nbnd):
        ext = lut.lut_extliq[b, 0] + x * (lut.lut_extliq[b, 1] - lut.lut_extliq[b, 0])
        ssa = lut.lut_ssaliq[b, 0] + x * (lut.lut_ssaliq[b, 1] - lut.lut_ssaliq[b, 0])
        assert abs(got[2, 4, b] - 20.0 * ext * (1.0 - ssa)) < 1e-10 * max(1.0, abs(got[2, 4, b]))

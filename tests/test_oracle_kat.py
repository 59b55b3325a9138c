"""Analytic known-answer tests that pin the oracle's solver kernels without any coefficient
file, restated from the reference's upstream Fortran unit tests:
  * LW gray radiative equilibrium  — rte-rrtmgp/tests/rte_lw_solver_unit_tests.F90:56-66,238-340
  * SW thin scattering atmosphere  — rte-rrtmgp/tests/rte_sw_solver_unit_tests.F90:56-66,121-133,226-272
Tolerances are the Fortran ones: |x - ref| <= tol * spacing(ref) (mo_testing_utils.F90:44-70).
"""
import numpy as np
import pytest

from oracle import oracle as O

SIGMA = 5.670374419e-8
D = 1.0 / 0.6096748751
NCOL, NLAY = 8, 16
SFC_T = np.array([285.0] * 4 + [310.0] * 4)
TOTAL_TAU = np.array([0.1, 1.0, 10.0, 50.0] * 2)


def allclose(tst, ref, tol):
    return np.all(np.abs(tst - ref) <= tol * np.spacing(np.abs(ref)))


def gray_olr(T, tau):
    return (2.0 * SIGMA * T ** 4) / (2.0 + D * tau)


def gray_rad_equil(top_at_1):
    disc = O.SpectralDisc(np.array([[0.0], [3250.0]]), np.array([[1], [1]]))
    tau = O.zeros((NCOL, NLAY, 1))
    tau[:, :, 0] = (TOTAL_TAU / NLAY)[:, None]
    olr = gray_olr(SFC_T, TOTAL_TAU)
    sfc = O.F((SIGMA / np.pi * SFC_T ** 4)[:, None])
    jac = O.F((4.0 * SIGMA / np.pi * SFC_T ** 3)[:, None])
    lev = O.zeros((NCOL, NLAY + 1, 1)); lay = O.zeros((NCOL, NLAY, 1))
    lev[:, 0, 0] = 0.5 / np.pi * olr
    for ilay in range(1, NLAY + 1):
        lev[:, ilay, 0] = 0.5 / np.pi * olr * (1.0 + D * tau[:, :ilay, 0].sum(axis=1))
        lay[:, ilay - 1, 0] = 0.5 * (lev[:, ilay, 0] + lev[:, ilay - 1, 0])
    if not top_at_1:
        lev = O.F(lev[:, ::-1]); lay = O.F(lay[:, ::-1])
    return disc, tau, lay, lev, sfc, jac


@pytest.mark.parametrize("top_at_1", [True, False])
def test_lw_gray_radiative_equilibrium(top_at_1):
    disc, tau, lay, lev, sfc, jac = gray_rad_equil(top_at_1)
    emis = np.ones((1, NCOL))
    up, dn = O.rte_lw(disc, tau, lay, lev, sfc, emis, top_at_1)
    fl = O.fluxes_broadband(up, dn)
    toa = 0 if top_at_1 else NLAY
    assert allclose(fl["flux_up"][:, toa], gray_olr(SFC_T, TOTAL_TAU), 8.0)
    net = fl["flux_net"]
    # Fortran: compares to spread(net(:,1)); the sign convention is dn - up
    assert np.all(np.abs(net - net[:, :1]) <= 100.0 * np.spacing(np.abs(fl["flux_up"]).max()))
    assert np.array_equal(net, fl["flux_dn"] - fl["flux_up"])
    # broadband accumulation inside the solver agrees with reduce() for a single g-point
    up1, dn1 = O.rte_lw(disc, tau, lay, lev, sfc, emis, top_at_1, broadband=True)
    assert allclose(up1[:, :, 0], fl["flux_up"], 2.0) and allclose(dn1[:, :, 0], fl["flux_dn"], 2.0)
    # computing the Jacobian does not change fluxes (F90:196-215)
    up2, dn2, jflux = O.rte_lw(disc, tau, lay, lev, sfc, emis, top_at_1, sfc_src_jac=jac)
    assert np.array_equal(up2, up) and np.array_equal(dn2, dn)
    assert np.all(jflux > 0)


def test_lw_vertical_flip_invariance():
    a = gray_rad_equil(True); b = gray_rad_equil(False)
    emis = np.ones((1, NCOL))
    up_a, dn_a = O.rte_lw(a[0], a[1], a[2], a[3], a[4], emis, True)
    up_b, dn_b = O.rte_lw(b[0], b[1], b[2], b[3], b[4], emis, False)
    assert np.array_equal(up_a, up_b[:, ::-1]) and np.array_equal(dn_a, dn_b[:, ::-1])


def test_lw_transparent_increment_and_subset_invariance():
    disc, tau, lay, lev, sfc, _ = gray_rad_equil(True)
    emis = np.ones((1, NCOL))
    up, dn = O.rte_lw(disc, tau, lay, lev, sfc, emis, True)
    tau2 = O.add_to_1scl(disc, tau, O.zeros(tau.shape))
    up2, dn2 = O.rte_lw(disc, tau2, lay, lev, sfc, emis, True)
    assert np.array_equal(up, up2) and np.array_equal(dn, dn2)
    s = slice(2, 6)
    up3, dn3 = O.rte_lw(disc, tau[s], lay[s], lev[s], sfc[s], emis[:, s], True)
    assert np.array_equal(up[s], up3) and np.array_equal(dn[s], dn3)


def thin_scattering():
    g = np.array([0.85, 0.65]); tau = np.array([1e-4, 1e-2]); ssa = 1.0 - np.array([1e-4, 1e-2])
    ntau, nssa, ng = 2, 2, 2
    t_col = np.array([tau[i] for j in range(nssa * ng) for i in range(ntau)])
    s_col = np.array([ssa[i] for j in range(ng) for i in range(nssa) for k in range(ntau)])
    g_col = np.array([g[i] for i in range(ng) for k in range(ntau * ng)])
    T = O.F(np.tile((t_col / NLAY)[:, None, None], (1, NLAY, 1)))
    S = O.F(np.tile(s_col[:, None, None], (1, NLAY, 1)))
    G = O.F(np.tile(g_col[:, None, None], (1, NLAY, 1)))
    return O.delta_scale(T, S, G)


@pytest.mark.parametrize("top_at_1", [True, False])
@pytest.mark.parametrize("mu0", [1.0, 0.5])
def test_sw_thin_scattering_direct_beam_and_linearity(mu0, top_at_1):
    disc = O.SpectralDisc(np.array([[3250.0], [1.0e5]]), np.array([[1], [1]]))
    tau, ssa, g = thin_scattering()
    mu = np.full(NCOL, mu0); toa = O.F(np.ones((NCOL, 1))); alb = np.zeros((1, NCOL))
    up, dn, dr = O.rte_sw(disc, tau, ssa, g, mu, toa, alb, alb, top_at_1)
    sfc = dr[:, NLAY if top_at_1 else 0, 0]
    ref = toa[:, 0] * mu * np.exp(-tau[:, :, 0].sum(axis=1) / mu)
    assert allclose(sfc, ref, 20.0)
    fl = O.fluxes_broadband(up, dn, dr)
    assert np.array_equal(fl["flux_net"], fl["flux_dn"] - fl["flux_up"])
    # linear in TOA flux (F90: factor = 2)
    up2, dn2, dr2 = O.rte_sw(disc, tau, ssa, g, mu, 2.0 * toa, alb, alb, top_at_1)
    assert allclose(up2, 2.0 * up, 2.0) and allclose(dn2, 2.0 * dn, 2.0) and allclose(dr2, 2.0 * dr, 2.0)
    # energy: with zero albedo nothing comes up from the surface
    assert np.all(up[:, 0 if not top_at_1 else NLAY, 0] == 0.0)


def test_sw_flip_and_transparent_increment_invariance():
    disc = O.SpectralDisc(np.array([[3250.0], [1.0e5]]), np.array([[1], [1]]))
    tau, ssa, g = thin_scattering()
    mu = np.full(NCOL, 0.5); toa = O.F(np.ones((NCOL, 1))); alb = np.full((1, NCOL), 0.2)
    a = O.rte_sw(disc, tau, ssa, g, mu, toa, alb, alb, True)
    b = O.rte_sw(disc, O.F(tau[:, ::-1]), O.F(ssa[:, ::-1]), O.F(g[:, ::-1]), mu, toa, alb, alb, False)
    for x, y in zip(a, b):
        assert np.array_equal(x, y[:, ::-1])
    z = O.zeros(tau.shape)
    t2, s2, g2 = O.add_to_2str(disc, tau, ssa, g, z, z, z)
    c = O.rte_sw(disc, t2, s2, g2, mu, toa, alb, alb, True)
    for x, y in zip(a, c):   # (tau*ssa)/tau re-rounds ssa and g: compare at 1e-12 relative
        assert np.allclose(y, x, rtol=1e-12, atol=1e-15)


def test_night_columns_are_dark():
    disc = O.SpectralDisc(np.array([[3250.0], [1.0e5]]), np.array([[1], [1]]))
    tau, ssa, g = thin_scattering()
    mu = np.zeros(NCOL); toa = O.F(np.ones((NCOL, 1))); alb = np.full((1, NCOL), 0.3)
    up, dn, dr = O.rte_sw(disc, tau, ssa, g, mu, toa, alb, alb, True)
    assert np.all(up == 0) and np.all(dn == 0) and np.all(dr == 0)


def test_heating_rate_formula():
    up = O.F(np.array([[10.0, 20.0, 40.0]])); dn = O.F(np.array([[100.0, 90.0, 70.0]]))
    p = O.F(np.array([[1000.0, 2000.0, 4000.0]]))
    h = O.heating_rate(up, dn, p)
    ref0 = (20.0 - 10.0 - 90.0 + 100.0) * 9.80665 / (1004.64 * 1000.0)
    assert np.isclose(h[0, 0], ref0, rtol=1e-15)


# ------------------------------------------------------------------------------------------
# the solver options the reference C++ never switches on.  No fixture of the reference pins
# them (src/Rte_lw.cpp:178 passes do_rescaling = false; src/Rte_sw.cpp:157-160 broadcasts mu0), so the
# restated branches are pinned by the properties upstream's own variant check states
# (rte-rrtmgp/tests/check_variants.F90:381-397: "uses rescaling and should agree with _1scl answers")
# and by the limits of the Tang et al. scaling.
# ------------------------------------------------------------------------------------------
def _lw_case(top_at_1, ncol=5, nlay=18):
    from rte_rrtmgp_cpp_b200.atmos import rce_atmosphere
    from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist
    kd = synthetic_kdist("lw", 16)
    atm = rce_atmosphere(ncol, nlay, top_at_1=top_at_1)
    go = O.GasOptics(kd, atm.vmr.keys())
    st = go.gas_optics_lw(atm.play, atm.plev, atm.tlay, atm.tsfc, atm.vmr, tlev=atm.tlev)
    return go, atm, st


@pytest.mark.parametrize("top_at_1", [False, True])
@pytest.mark.parametrize("n_ang", [1, 3])
def test_lw_rescaling_without_scattering_is_the_noscat_solution(top_at_1, n_ang):
    go, atm, st = _lw_case(top_at_1)
    args = (go, st["tau"], st["lay_src"], st["lev_src"], st["sfc_src"], atm.emis_sfc, top_at_1)
    up0, dn0, jac0 = O.rte_lw(*args, n_gauss_angles=n_ang, sfc_src_jac=st["sfc_src_jac"])
    zero = np.zeros_like(st["tau"])
    g = np.full_like(st["tau"], 0.7)
    up1, dn1, jac1 = O.rte_lw(*args, n_gauss_angles=n_ang, sfc_src_jac=st["sfc_src_jac"], ssa=zero, g=g)
    # ssa = 0: scaleTau = 1, Cn = 0 -> the same operations on the same numbers
    assert np.array_equal(up0, up1) and np.array_equal(dn0, dn1) and np.array_equal(jac0, jac1)


@pytest.mark.parametrize("top_at_1", [False, True])
def test_lw_rescaling_limits(top_at_1):
    go, atm, st = _lw_case(top_at_1)
    rng = np.random.default_rng(3)
    ssa = np.asfortranarray(rng.uniform(0.1, 0.9, st["tau"].shape))
    tau = np.asfortranarray(st["tau"] + rng.uniform(0.0, 2.0, st["tau"].shape) * (rng.uniform(size=st["tau"].shape) < 0.3))
    rest = (st["lay_src"], st["lev_src"], st["sfc_src"], atm.emis_sfc, top_at_1)
    # g = 1 (pure forward scattering): b = 0, Cn = 0 -> absorption-only transport of tau * (1 - ssa)
    one = np.ones_like(tau)
    up_f, dn_f = O.rte_lw(go, tau, *rest, ssa=ssa, g=one)
    up_a, dn_a = O.rte_lw(go, np.asfortranarray(tau * (1.0 - ssa)), *rest)
    assert np.max(np.abs(up_f - up_a)) < 1e-10 and np.max(np.abs(dn_f - dn_a)) < 1e-10
    # back-scattering (g < 1) matters
    g = np.asfortranarray(rng.uniform(0.0, 0.9, tau.shape))
    up_s, dn_s = O.rte_lw(go, tau, *rest, ssa=ssa, g=g)
    assert np.max(np.abs(dn_s - dn_a)) > 1e-3
    assert np.all(np.isfinite(up_s)) and np.all(up_s >= 0) and np.all(dn_s >= 0)
    # a back-scattering layer returns part of the up-welling radiation: more down-welling at the surface than
    # with pure forward scattering, summed over g-points
    sfc = 0 if not top_at_1 else -1
    assert dn_s[:, sfc, :].sum() > dn_f[:, sfc, :].sum()


@pytest.mark.parametrize("top_at_1", [False, True])
def test_sw_height_dependent_mu0(top_at_1):
    from rte_rrtmgp_cpp_b200.atmos import rce_atmosphere
    from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist
    kd = synthetic_kdist("sw", 16)
    atm = rce_atmosphere(6, 20, top_at_1=top_at_1, night_fraction=0.0)
    go = O.GasOptics(kd, atm.vmr.keys())
    st = go.gas_optics_sw(atm.play, atm.plev, atm.tlay, atm.vmr)
    args = (st["toa_src"], atm.sfc_alb_dir, atm.sfc_alb_dif, top_at_1)
    base = O.rte_sw(go, st["tau"], st["ssa"], st["g"], atm.mu0, *args)
    # a constant (ncol, nlay) field is the per-column call, bit for bit
    const = O.rte_sw(go, st["tau"], st["ssa"], st["g"], np.tile(atm.mu0[:, None], (1, atm.nlay)), *args)
    for a, b in zip(base, const):
        assert np.array_equal(a, b)
    # sun below the local horizon in the lowest 5 layers of column 0 (twilight on a round earth): no direct
    # source there, the beam keeps its nominal-value attenuation (F90:1061-1065, 1117-1125)
    mu = np.tile(atm.mu0[:, None], (1, atm.nlay))
    low = slice(atm.nlay - 5, atm.nlay) if top_at_1 else slice(0, 5)
    mu[0, low] = -0.05
    up, dn, dr = O.rte_sw(go, st["tau"], st["ssa"], st["g"], mu, *args)
    assert np.array_equal(up[1:], base[0][1:]) and np.array_equal(dn[1:], base[1][1:])   # other columns untouched
    sfc = -1 if top_at_1 else 0
    assert np.all(dr[0, sfc, :] < 1e-6 * base[2][0, sfc, :] + 1e-300)    # exp(-tau/sqrt(eps)) kills the beam
    assert np.all(up[0, sfc, :] <= base[0][0, sfc, :])
    assert np.all(np.isfinite(up)) and np.all(up >= 0)

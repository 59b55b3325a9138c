"""GPU parity of the fused LW / SW column solvers (C ABI rrtmgp_b200_{lw,sw}_fluxes) against the
CPU oracle on identical seeded inputs.  Tolerances are BASELINE.json's: fluxes within 1e-5 W/m2
in the fp64 build, within 0.05 W/m2 in the fp32 build (vs the fp64 oracle)."""
import numpy as np
import pytest
import torch

from rte_rrtmgp_cpp_b200 import api
from rte_rrtmgp_cpp_b200.atmos import rce_atmosphere, allsky_atmosphere
from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist, synthetic_cloud_lut, REDUCED_LW_128, REDUCED_SW_112
from helpers import oracle_lw, oracle_sw

pytestmark = pytest.mark.gpu
TOL = {torch.float64: 1e-5, torch.float32: 0.05}


def _cmp(out, ref, keys, tol):
    for k in keys:
        d = np.max(np.abs(out[k].numpy() - ref[k]))
        assert d <= tol, f"{k}: max abs diff {d} > {tol}"


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("top_at_1", [False, True])
def test_lw_fused_clear(dtype, top_at_1):
    kd = synthetic_kdist("lw")
    atm = rce_atmosphere(70, 48, top_at_1=top_at_1)
    ref = oracle_lw(kd, atm)
    sol = api.Radiation_solver_longwave(kd, atm.vmr.keys(), dtype=dtype)
    out = sol.solve_gpu(api.DeviceColumns(atm, dtype), switch_output_bnd_fluxes=True, heating=True)
    _cmp(out, ref, ("flux_up", "flux_dn", "flux_net", "bnd_flux_up", "bnd_flux_dn", "bnd_flux_net"), TOL[dtype])
    hr_tol = 1e-9 if dtype == torch.float64 else 2e-3   # K/s; thin top layers amplify flux noise in fp32
    assert np.max(np.abs(out["heating_rate"].numpy() - ref["heating_rate"])) <= hr_tol


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("top_at_1", [False, True])
def test_sw_fused_clear(dtype, top_at_1):
    kd = synthetic_kdist("sw")
    atm = rce_atmosphere(70, 48, top_at_1=top_at_1, night_fraction=0.2)
    atm.tsi_scaling[:] = np.linspace(0.9, 1.1, atm.ncol)
    ref = oracle_sw(kd, atm)
    sol = api.Radiation_solver_shortwave(kd, atm.vmr.keys(), dtype=dtype)
    out = sol.solve_gpu(api.DeviceColumns(atm, dtype), switch_output_bnd_fluxes=True)
    _cmp(out, ref, ("flux_up", "flux_dn", "flux_net", "flux_dn_dir", "bnd_flux_up", "bnd_flux_dn",
                    "bnd_flux_net", "bnd_flux_dn_dir"), TOL[dtype])
    night = atm.mu0 <= 0
    assert night.any() and np.all(out["flux_dn"].numpy()[night] == 0.0)


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("delta", [True, False])
def test_allsky_fused(dtype, delta):
    atm = allsky_atmosphere(24, 72)       # the reference's all-sky case: 2 of 3 columns cloudy
    for kind in ("lw", "sw"):
        kd, lut = synthetic_kdist(kind), synthetic_cloud_lut(kind)
        if kind == "lw":
            ref = oracle_lw(kd, atm, lut)
            sol = api.Radiation_solver_longwave(kd, atm.vmr.keys(), lut, dtype)
            keys = ("flux_up", "flux_dn", "flux_net")
        else:
            ref = oracle_sw(kd, atm, lut, delta)
            sol = api.Radiation_solver_shortwave(kd, atm.vmr.keys(), lut, dtype)
            keys = ("flux_up", "flux_dn", "flux_net", "flux_dn_dir")
        out = sol.solve_gpu(api.DeviceColumns(atm, dtype), switch_cloud_optics=True, switch_delta_cloud=delta)
        _cmp(out, ref, keys, TOL[dtype])


@pytest.mark.parametrize("kind,sizes", [("lw", REDUCED_LW_128), ("sw", REDUCED_SW_112)])
def test_reduced_gpoint_sets(kind, sizes):
    kd = synthetic_kdist(kind, sizes)
    atm = rce_atmosphere(33, 30, cloudy=True)
    lut = synthetic_cloud_lut(kind)
    if kind == "lw":
        ref, sol = oracle_lw(kd, atm, lut), api.Radiation_solver_longwave(kd, atm.vmr.keys(), lut)
    else:
        ref, sol = oracle_sw(kd, atm, lut), api.Radiation_solver_shortwave(kd, atm.vmr.keys(), lut)
    out = sol.solve_gpu(api.DeviceColumns(atm), switch_cloud_optics=True, switch_output_bnd_fluxes=True)
    _cmp(out, ref, ("flux_up", "flux_dn", "bnd_flux_up", "bnd_flux_dn"), 1e-5)


def test_tiling_and_column_sharding_are_bit_identical():
    """T5: a column's result does not depend on the tile or shard it is computed in."""
    kd = synthetic_kdist("lw")
    atm = rce_atmosphere(200, 32)
    sol = api.Radiation_solver_longwave(kd, atm.vmr.keys())
    full = sol.solve_gpu(api.DeviceColumns(atm))["flux_up"].numpy().copy()
    sol.set_tile_columns(64)
    tiled = sol.solve_gpu(api.DeviceColumns(atm))["flux_up"].numpy().copy()
    assert np.array_equal(full, tiled)
    sol.set_tile_columns(0)
    parts = [sol.solve_gpu(api.DeviceColumns(atm.subset(a, b)))["flux_up"].numpy().copy()
             for a, b in ((0, 77), (77, 200))]
    assert np.array_equal(full, np.concatenate(parts, axis=0))


def test_vertical_flip_invariance():
    kd = synthetic_kdist("sw")
    atm = rce_atmosphere(40, 36)
    sol = api.Radiation_solver_shortwave(kd, atm.vmr.keys())
    a = sol.solve_gpu(api.DeviceColumns(atm))["flux_net"].numpy().copy()
    b = sol.solve_gpu(api.DeviceColumns(atm.flipped()))["flux_net"].numpy().copy()
    assert np.max(np.abs(a - b[:, ::-1])) < 1e-9


def test_errors_are_loud():
    kd = synthetic_kdist("lw")
    atm = rce_atmosphere(4, 8)
    with pytest.raises(RuntimeError, match="required gas"):
        api.Gas_optics_rrtmgp_gpu(kd, [g for g in atm.vmr if g != "co2"])
    sol = api.Radiation_solver_shortwave(synthetic_kdist("sw"), atm.vmr.keys())
    lw = api.Radiation_solver_longwave(kd, atm.vmr.keys())
    cols = api.DeviceColumns(atm)
    cols.gas_concs.gas_concs_map.pop("o3")
    with pytest.raises(RuntimeError, match="o3"):
        lw.solve_gpu(cols)


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("n_ang,reduced", [(2, False), (3, False), (4, False), (3, True)])
def test_lw_fused_gauss_angles(dtype, n_ang, reduced):
    """Gauss-Jacobi quadrature with 2..4 angles in the fused call (src/Rte_lw.cpp:139-151; the CPU reference
    supports it, its GPU path does not).  Bands of the common shape run ONE launch with the angles inside the layer loop
    (fused_lw_ma.cu: optical depths and Planck sources once per layer); the narrow bands of the reduced g-point set run one
    pass of the single-angle kernel per angle.  Either way fluxes and by-band fluxes add up over the angles."""
    kd = synthetic_kdist("lw", REDUCED_LW_128) if reduced else synthetic_kdist("lw")
    lut = synthetic_cloud_lut("lw")
    atm = rce_atmosphere(45, 40, cloudy=True)
    ref = oracle_lw(kd, atm, lut, n_ang=n_ang)
    sol = api.Radiation_solver_longwave(kd, atm.vmr.keys(), lut, dtype)
    sol.n_gauss_angles = n_ang
    out = {}
    for k, nd in (("flux_up", 2), ("flux_dn", 2), ("flux_net", 2), ("bnd_flux_up", 3), ("bnd_flux_dn", 3), ("bnd_flux_net", 3)):
        out[k] = api.Array_gpu((45, 41) if nd == 2 else (45, 41, kd.nbnd), dtype)
        out[k].t.fill_(-999.0)                       # every element must be (over)written, not accumulated onto garbage
    sol.solve_gpu(api.DeviceColumns(atm, dtype), switch_cloud_optics=True, switch_output_bnd_fluxes=True, out=out)
    _cmp(out, ref, ("flux_up", "flux_dn", "flux_net", "bnd_flux_up", "bnd_flux_dn", "bnd_flux_net"), TOL[dtype])


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_rfmip_block_shape(dtype):
    """BASELINE config 1: one RFMIP block — 100 columns x 60 layers, LW 256 + SW 224 g-points — against the oracle."""
    atm = rce_atmosphere(100, 60, night_fraction=0.1)
    for kind, oracle, Solver, keys in (("lw", oracle_lw, api.Radiation_solver_longwave, ("flux_up", "flux_dn", "flux_net")),
                                       ("sw", oracle_sw, api.Radiation_solver_shortwave, ("flux_up", "flux_dn", "flux_net", "flux_dn_dir"))):
        kd = synthetic_kdist(kind)
        assert kd.ngpt == (256 if kind == "lw" else 224)
        out = Solver(kd, atm.vmr.keys(), dtype=dtype).solve_gpu(api.DeviceColumns(atm, dtype), heating=True)
        ref = oracle(kd, atm)
        _cmp(out, ref, keys, TOL[dtype])


def replicate_columns(a, ndist, rep):
    """Array_gpu with a column dimension of ndist -> the same with every column repeated `rep` times (column c + i*ndist
    = column c).  Column-major: the column index is the fastest (fields) or the second (per-band surface arrays)."""
    if a is None:
        return None
    t = a.t.view(*reversed(a.dims))              # torch row-major view: last index = first (fastest) dimension
    if a.dims[0] == ndist:
        t = t.repeat(*([1] * (t.dim() - 1)), rep)
        dims = (ndist * rep,) + tuple(a.dims[1:])
    elif len(a.dims) == 2 and a.dims[1] == ndist:
        t = t.repeat(rep, 1)
        dims = (a.dims[0], ndist * rep)
    else:
        return a
    return api.Array_gpu(dims, a.t.dtype, a.t.device, t.contiguous().view(-1))


def test_full_size_domain_replicas_are_bit_identical():
    """BASELINE configs 3/4 at a size the oracle cannot reach: a 512 x 512 x 128 domain made of 4096 distinct columns
    repeated 64 times.  Columns are independent, so every replica must carry exactly the bits of its source column
    computed in a separate 4096-column call: tiles, waves and scratch-strip recycling must not leak between columns."""
    ndist, rep, nlay = 4096, 64, 128
    atm = rce_atmosphere(ndist, nlay, cloudy=True)
    dt = torch.float32
    small = api.DeviceColumns(atm, dt)
    big = api.DeviceColumns(atm, dt)
    for name in ("play", "plev", "tlay", "tlev", "tsfc", "mu0", "tsi_scaling", "lwp", "iwp", "rel", "dei", "emis_sfc",
                 "sfc_alb_dir", "sfc_alb_dif"):
        setattr(big, name, replicate_columns(getattr(small, name), ndist, rep))
    for k, v in list(big.gas_concs.gas_concs_map.items()):
        big.gas_concs.gas_concs_map[k] = replicate_columns(v, ndist, rep)
    big.ncol = ndist * rep
    for kind, Solver in (("lw", api.Radiation_solver_longwave), ("sw", api.Radiation_solver_shortwave)):
        sol = Solver(synthetic_kdist(kind), atm.vmr.keys(), synthetic_cloud_lut(kind), dt)
        ref = {k: v.t.clone() for k, v in sol.solve_gpu(small, switch_cloud_optics=True, heating=True).items()}
        out = sol.solve_gpu(big, switch_cloud_optics=True, heating=True)
        torch.cuda.synchronize()
        for k in ("flux_up", "flux_dn", "flux_net", "heating_rate"):
            got = out[k].t.view(-1, ndist * rep)                      # (nlev | nlay, ncol)
            want = ref[k].view(-1, ndist).repeat(1, rep)
            assert torch.equal(got, want), (kind, k)
        assert torch.isfinite(out["flux_net"].t).all()


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("case", ["clear", "allsky-aerosol-3angles", "aerosol-only", "reduced", "ragged-tiles"])
def test_paired_lw_sw_call_is_bit_identical_to_the_two_calls(case, dtype):
    """rrtmgp_b200_lwsw_fluxes: longwave and shortwave CTAs co-resident in one launch per band pair (fused_pair.cu).  Separate
    strips and flux slabs, the same per-CTA code: every output must equal lw.solve_gpu + sw.solve_gpu bit for bit — full-width
    bands (paired kernel), the reduced g-point sets (narrow bands fall back to the single-role kernels mid-sequence), several
    tiles with a ragged last one, clouds + aerosols, and extra longwave Gauss angles."""
    from rte_rrtmgp_cpp_b200.atmos import add_aerosols
    from rte_rrtmgp_cpp_b200.kdist import synthetic_aerosol_lut, REDUCED_LW_128, REDUCED_SW_112
    kw = {}
    if case == "clear":
        atm = rce_atmosphere(150, 40)
        kd_lw, kd_sw, lut_lw, lut_sw, alut = synthetic_kdist("lw"), synthetic_kdist("sw"), None, None, None
    elif case == "allsky-aerosol-3angles":
        atm = add_aerosols(allsky_atmosphere(70, 48, random_clouds=True))
        kd_lw, kd_sw = synthetic_kdist("lw"), synthetic_kdist("sw")
        lut_lw, lut_sw, alut = synthetic_cloud_lut("lw"), synthetic_cloud_lut("sw"), synthetic_aerosol_lut()
        kw = dict(switch_cloud_optics=True, switch_output_bnd_fluxes=True, heating=True)
    elif case == "aerosol-only":        # shortwave with, longwave without a particulate increment: not paired, same results
        atm = add_aerosols(rce_atmosphere(60, 30))
        kd_lw, kd_sw, lut_lw, lut_sw, alut = synthetic_kdist("lw"), synthetic_kdist("sw"), None, None, synthetic_aerosol_lut()
    elif case == "reduced":
        atm = rce_atmosphere(100, 33, top_at_1=True)
        kd_lw, kd_sw, lut_lw, lut_sw, alut = synthetic_kdist("lw", REDUCED_LW_128), synthetic_kdist("sw", REDUCED_SW_112), None, None, None
    else:
        atm = rce_atmosphere(333, 17)
        kd_lw, kd_sw, lut_lw, lut_sw, alut = synthetic_kdist("lw"), synthetic_kdist("sw"), None, None, None
    lw = api.Radiation_solver_longwave(kd_lw, atm.vmr.keys(), lut_lw, dtype)
    sw = api.Radiation_solver_shortwave(kd_sw, atm.vmr.keys(), lut_sw, dtype, aerosol_lut=alut) if alut is not None else \
        api.Radiation_solver_shortwave(kd_sw, atm.vmr.keys(), lut_sw, dtype)
    if case == "ragged-tiles":
        lw.set_tile_columns(128); sw.set_tile_columns(96)
    if case == "allsky-aerosol-3angles":
        lw.n_gauss_angles = 3
    cols = api.DeviceColumns(atm, dtype)
    kw_sw = dict(kw, switch_aerosol_optics=True) if alut is not None else kw
    ref_lw = {k: v.numpy().copy() for k, v in lw.solve_gpu(cols, **kw).items()}
    ref_sw = {k: v.numpy().copy() for k, v in sw.solve_gpu(cols, **kw_sw).items()}
    out_lw, out_sw = api.solve_pair(lw, sw, cols, **kw_sw)
    torch.cuda.synchronize()
    assert set(out_lw) == set(ref_lw) and set(out_sw) == set(ref_sw)
    for k in ref_lw:
        assert np.array_equal(out_lw[k].numpy(), ref_lw[k]), ("lw", k)
    for k in ref_sw:
        assert np.array_equal(out_sw[k].numpy(), ref_sw[k]), ("sw", k)
    assert np.abs(ref_lw["flux_up"]).max() > 10 and np.abs(ref_sw["flux_dn"]).max() > 10


def test_many_waves_recycle_strips_and_keep_columns_independent():
    """A domain of many waves of CTAs (more CTAs per band launch than scratch strips: strips are claimed, released and claimed
    again by later CTAs) built from a small set of distinct columns repeated periodically.  Size-independent properties: every
    period is bit-identical to the first, and the first equals the same columns computed alone — whatever strip a CTA got,
    whichever wave it ran in, whichever tile it fell into.  Both the paired call and the two single calls."""
    dtype = torch.float32
    nbase, reps, nlay = 2048, 49, 20                     # 100 352 columns = 3136 CTAs per role; 1480 strips per role
    base = rce_atmosphere(nbase, nlay, cloudy=True)
    kd_lw, kd_sw = synthetic_kdist("lw"), synthetic_kdist("sw")
    lw = api.Radiation_solver_longwave(kd_lw, base.vmr.keys(), synthetic_cloud_lut("lw"), dtype)
    sw = api.Radiation_solver_shortwave(kd_sw, base.vmr.keys(), synthetic_cloud_lut("sw"), dtype)
    small = api.DeviceColumns(base, dtype)
    ref_lw, ref_sw = api.solve_pair(lw, sw, small, switch_cloud_optics=True)
    ref = {("lw", k): v.t.clone() for k, v in ref_lw.items()} | {("sw", k): v.t.clone() for k, v in ref_sw.items()}

    def rep2(a):          # (nbase, n2) column-major -> (nbase*reps, n2)
        n2 = a.dims[1]
        return api.Array_gpu((nbase * reps, n2), a.t.dtype, data=a.t.view(n2, nbase).repeat(1, reps).contiguous().view(-1))

    def repb(a):          # (nbnd, nbase) -> (nbnd, nbase*reps)
        n1 = a.dims[0]
        return api.Array_gpu((n1, nbase * reps), a.t.dtype, data=a.t.view(nbase, n1).repeat(reps, 1).contiguous().view(-1))

    big = api.DeviceColumns(base, dtype)
    for k in ("play", "plev", "tlay", "tlev", "lwp", "iwp", "rel", "dei"):
        setattr(big, k, rep2(getattr(big, k)))
    for k in ("emis_sfc", "sfc_alb_dir", "sfc_alb_dif"):
        setattr(big, k, repb(getattr(big, k)))
    for k in ("tsfc", "mu0", "tsi_scaling"):
        a = getattr(big, k)
        setattr(big, k, api.Array_gpu((nbase * reps,), a.t.dtype, data=a.t.repeat(reps).contiguous()))
    for name, a in list(big.gas_concs.gas_concs_map.items()):
        if a.dims[0] > 1:
            big.gas_concs.gas_concs_map[name] = rep2(a)
    big.ncol = nbase * reps
    lw.set_tile_columns(40000); sw.set_tile_columns(40000)            # three ragged tiles, not a multiple of the period
    nlev = nlay + 1
    for mode in ("pair", "single"):
        if mode == "pair":
            o_lw, o_sw = api.solve_pair(lw, sw, big, switch_cloud_optics=True)
        else:
            o_lw, o_sw = lw.solve_gpu(big, switch_cloud_optics=True), sw.solve_gpu(big, switch_cloud_optics=True)
        torch.cuda.synchronize()
        for (w, k), want in ref.items():
            got = (o_lw if w == "lw" else o_sw)[k].t.view(nlev, reps, nbase)
            assert torch.equal(got, want.view(nlev, 1, nbase).expand(nlev, reps, nbase)), (mode, w, k)

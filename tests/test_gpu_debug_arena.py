"""Stand-in for compute-sanitizer (closed on this GPU pool): rrtmgp_b200_set_debug(handle, 1) fills the handle's whole workspace
with NaN bytes before every fused call and checks a canary block behind every carved piece after it.  The fused kernels index
a multi-GB arena by hand and claim scratch strips through an atomicCAS spin, so: at sizes where tiles, chunks and CTAs do not
divide the domain, and at the smallest and largest layer counts, results under the debug mode must be finite, bit-identical to
the normal run, and no canary may be touched (the call fails with E_CUDA otherwise)."""
import numpy as np
import pytest
import torch

from rte_rrtmgp_cpp_b200 import api
from rte_rrtmgp_cpp_b200.atmos import rce_atmosphere, add_aerosols
from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist, synthetic_cloud_lut, synthetic_aerosol_lut, REDUCED_LW_128, REDUCED_SW_112

pytestmark = pytest.mark.gpu


def _solvers(dtype, cloudy=False, aerosol=False, reduced=False):
    kd_lw = synthetic_kdist("lw", REDUCED_LW_128) if reduced else synthetic_kdist("lw")
    kd_sw = synthetic_kdist("sw", REDUCED_SW_112) if reduced else synthetic_kdist("sw")
    gases = ["h2o", "co2", "o3", "n2o", "ch4", "n2", "o2"]
    lw = api.Radiation_solver_longwave(kd_lw, gases, synthetic_cloud_lut("lw") if cloudy else None, dtype)
    sw = api.Radiation_solver_shortwave(kd_sw, gases, synthetic_cloud_lut("sw") if cloudy else None, dtype,
                                        aerosol_lut=synthetic_aerosol_lut() if aerosol else None)
    return lw, sw


def _snapshot(out):
    return {k: v.numpy().copy() for k, v in out.items()}


def _same(a, b, what):
    assert set(a) == set(b)
    for k in a:
        assert np.isfinite(b[k]).all(), (what, k, "not finite under the NaN-poisoned workspace")
        assert np.array_equal(a[k], b[k]), (what, k)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("ncol,nlay,tile,path", [
    (333, 17, 96, 0),       # 4 ragged tiles, last CTA partly empty, strips kernels
    (333, 17, 96, 1),       # column-resident kernels
    (45, 1, 0, 0),          # a single layer
    (45, 2, 0, 0),
    (9, 255, 0, 0),         # odd, and the last layer count of one block of tropopause-mask words
    (9, 256, 0, 0),
    (9, 257, 0, 0),         # one layer into the second block of tropopause-mask words
    (6, 400, 0, 2),         # by size: column-resident in fp32; in fp64 400 layers exceed its shared memory, so the strips kernels
    (70, 40, 64, 2),        # path chosen by size
])
def test_poisoned_workspace_and_canaries(dtype, ncol, nlay, tile, path):
    cloudy = nlay == 17
    atm = rce_atmosphere(ncol, nlay, cloudy=cloudy)
    if cloudy:
        atm = add_aerosols(atm)
    lw, sw = _solvers(dtype, cloudy=cloudy, aerosol=cloudy)
    for sol in (lw, sw):
        sol.set_fused_path(path)
        if tile:
            sol.set_tile_columns(tile)
    cols = api.DeviceColumns(atm, dtype)
    kw = dict(switch_cloud_optics=cloudy, switch_output_bnd_fluxes=True, heating=True)
    kw_sw = dict(kw, switch_aerosol_optics=cloudy)
    ref = (_snapshot(lw.solve_gpu(cols, **kw)), _snapshot(sw.solve_gpu(cols, **kw_sw)))
    ref_pair = api.solve_pair(lw, sw, cols, **kw_sw)
    ref_pair = (_snapshot(ref_pair[0]), _snapshot(ref_pair[1]))
    for sol in (lw, sw):
        sol.set_debug(1)
    for rep in range(2):            # twice: the second call finds the first one's data, re-poisoned
        got = (_snapshot(lw.solve_gpu(cols, **kw)), _snapshot(sw.solve_gpu(cols, **kw_sw)))
        _same(ref[0], got[0], "lw"); _same(ref[1], got[1], "sw")
        o_lw, o_sw = api.solve_pair(lw, sw, cols, **kw_sw)
        _same(ref_pair[0], _snapshot(o_lw), "pair lw"); _same(ref_pair[1], _snapshot(o_sw), "pair sw")
    for sol in (lw, sw):
        sol.set_debug(0)


def test_host_pipeline_with_ragged_chunks_under_debug_mode():
    dtype = torch.float32
    atm = rce_atmosphere(301, 23)
    lw, sw = _solvers(dtype, reduced=True)
    hcols = api.HostColumns(atm, dtype)
    ref_lw, ref_sw = api.solve_host(lw, sw, hcols, chunk_columns=128)
    ref = (_snapshot(ref_lw), _snapshot(ref_sw))
    lw.set_debug(1); sw.set_debug(1)
    o_lw, o_sw = api.solve_host(lw, sw, hcols, chunk_columns=128, out_lw={}, out_sw={})
    _same(ref[0], _snapshot(o_lw), "host lw"); _same(ref[1], _snapshot(o_sw), "host sw")


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("n_ang,reduced", [(2, False), (4, False), (3, True)])
def test_gauss_angles_under_debug_mode(dtype, n_ang, reduced):
    """Several Gauss angles: the strips hold {trans, source_up} per angle (fused_lw_ma.cu), so a strip is n_ang times longer;
    the reduced g-point set mixes those launches with one pass per angle for its narrow bands."""
    lw, _ = _solvers(dtype, cloudy=True, reduced=reduced)
    lw.set_fused_path(0)
    lw.set_tile_columns(64)
    lw.n_gauss_angles = n_ang
    cols = api.DeviceColumns(rce_atmosphere(150, 33, cloudy=True), dtype)
    kw = dict(switch_cloud_optics=True, switch_output_bnd_fluxes=True)
    ref = _snapshot(lw.solve_gpu(cols, **kw))
    lw.set_debug(1)
    for rep in range(2):
        _same(ref, _snapshot(lw.solve_gpu(cols, **kw)), f"lw {n_ang} angles")
    lw.set_debug(0)


def test_an_overwritten_canary_fails_the_call():
    """flags = 3: the library corrupts one canary byte after arming them; the call must fail and name the canary."""
    lw, sw = _solvers(torch.float32)
    cols = api.DeviceColumns(rce_atmosphere(40, 12), torch.float32)
    lw.set_debug(3)
    with pytest.raises(RuntimeError, match="canary"):
        lw.solve_gpu(cols)
    lw.set_debug(1)
    assert np.isfinite(lw.solve_gpu(cols)["flux_up"].numpy()).all()

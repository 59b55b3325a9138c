"""CPU-side checks of the C-ABI library: it loads, exports every symbol include/rrtmgp_b200.h
declares, its host-side table preparation (RRTMGP_B200_HOST_ONLY) reproduces the reference
constructor logic as restated by the oracle, argument errors are reported without a device, and
device entry points fail loudly (no CPU fallback) when there is no GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

from oracle import oracle as O
from rte_rrtmgp_cpp_b200 import api, capi
from rte_rrtmgp_cpp_b200.kdist import (synthetic_kdist, synthetic_cloud_lut, synthetic_aerosol_lut, REDUCED_LW_128,
                                       REDUCED_SW_112)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GASES = ["h2o", "co2", "o3", "n2o", "ch4", "n2", "o2"]


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "rrtmgp_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(rrtmgp_b200_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 30
    L = C.CDLL(capi.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(L, name), f"{name} declared in include/rrtmgp_b200.h but not exported"
    assert declared == set(capi.SYMBOLS), declared ^ set(capi.SYMBOLS)
    assert capi.lib().rrtmgp_b200_version() >= 100


@pytest.mark.parametrize("kind,sizes", [("lw", 16), ("sw", 16), ("lw", REDUCED_LW_128), ("sw", REDUCED_SW_112)])
@pytest.mark.parametrize("gases", [GASES, GASES + ["co", "cfc11", "cfc12", "no2"]])
def test_host_tables_match_reference_constructor(kind, sizes, gases):
    kd = synthetic_kdist(kind, sizes)
    go = api.Gas_optics_rrtmgp_gpu(kd, gases, host_only=True)
    og = O.GasOptics(kd, gases)
    assert go.gas_names == og.gas_names
    assert (go.get_ngpt(), go.get_nband(), go.get_nflav()) == (og.ngpt, og.nbnd, og.nflav)
    assert np.array_equal(go.flavor, og.flavor)
    assert np.array_equal(go.gpoint_flavor, og.gpoint_flavor)
    assert np.array_equal(go.get_band_lims_gpoint(), og.band2gpt)
    assert np.array_equal(go.get_gpoint_bands(), og.gpt2band)
    assert np.array_equal(go.get_band_lims_wavenumber(), og.band_lims_wvn)
    for side, o in ((0, og.lower), (1, og.upper)):
        m = go.get_minor(side)
        for k in ("idx_minor", "idx_scal", "kstart", "limits", "dens", "comp"):
            assert np.array_equal(m[k], o[k]), (side, k)
    assert go.get_press_ref_min() == og.press_ref_min and go.get_press_ref_max() == og.press_ref_max
    assert go.get_temp_min() == og.temp_ref_min and go.get_temp_max() == og.temp_ref_max
    if kind == "sw":
        assert np.array_equal(go.solar_source, og.solar_source) and go.get_tsi() == og.get_tsi()
        assert go.source_is_external() and not go.source_is_internal()
    else:
        assert go.source_is_internal()


def test_constructor_errors_match_reference():
    kd = synthetic_kdist("lw")
    with pytest.raises(RuntimeError, match="required gas co2 is missing"):
        api.Gas_optics_rrtmgp_gpu(kd, [g for g in GASES if g != "co2"], host_only=True)
    with pytest.raises(RuntimeError):
        O.GasOptics(kd, [g for g in GASES if g != "co2"])


def test_argument_errors_need_no_device():
    L = capi.lib()
    rc = L.rrtmgp_b200_delta_scale_2str(capi.F32, 0, 4, 4, None, None, None, None)
    assert rc == capi.E_ARG and b"ncol" in L.rrtmgp_b200_last_error()
    rc = L.rrtmgp_b200_lw_solver_noscat(capi.F64, 4, 4, 4, 1, 7, *([None] * 8), 0, None, None, None)
    assert rc == capi.E_ARG
    lims = (C.c_int * 4)(1, 2, 4, 3)
    rc = L.rrtmgp_b200_sum_byband(capi.F64, 4, 4, 4, 2, lims, C.c_void_p(8), C.c_void_p(8), None)
    assert rc == capi.E_ARG and b"band limits" in L.rrtmgp_b200_last_error()


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback_without_a_device():
    kd = synthetic_kdist("lw")
    with pytest.raises(RuntimeError, match="(?i)cuda"):
        api.Gas_optics_rrtmgp_gpu(kd, GASES)                     # device tables need a GPU
    go = api.Gas_optics_rrtmgp_gpu(kd, GASES, host_only=True)
    concs = (C.c_void_p * go.info.ngas)(*([8] * go.info.ngas))
    one = (C.c_int * go.info.ngas)(*([1] * go.info.ngas))
    gc = capi.GasConcsC(go.info.ngas, concs, one, one)
    rc = capi.lib().rrtmgp_b200_interpolation(go.h, 1, 1, C.c_void_p(8), C.c_void_p(8), C.c_void_p(8), C.byref(gc),
                                              *([None] * 8), None)
    assert rc == capi.E_CUDA and b"no CPU fallback" in capi.lib().rrtmgp_b200_last_error()
    with pytest.raises(RuntimeError, match="(?i)cuda"):
        api.Cloud_optics_gpu(synthetic_cloud_lut("lw"))
    with pytest.raises(RuntimeError, match="(?i)cuda"):
        api.Aerosol_optics_gpu(synthetic_aerosol_lut())


def test_aerosol_table_argument_errors_need_no_device():
    """The reference addresses hydrophobic columns up to 11 and hydrophilic types up to 5
    (src/Aerosol_optics.cpp:64-142): smaller tables are refused before anything touches the device."""
    from rte_rrtmgp_cpp_b200.ctables import pack_aerosol
    lut = synthetic_aerosol_lut()
    lut.mass_ext_sw_hydrophobic = lut.mass_ext_sw_hydrophobic[:9]
    a, keep = pack_aerosol(lut)
    h = C.c_void_p()
    rc = capi.lib().rrtmgp_b200_aerosol_optics_create(C.byref(a), capi.F64, C.byref(h))
    assert rc == capi.E_ARG and b"hydrophobic" in capi.lib().rrtmgp_b200_last_error() and not h.value
    assert capi.lib().rrtmgp_b200_aerosol_optics_nband(None) == -1


def test_shard_bounds_is_the_python_partition():
    """rrtmgp_b200_shard_bounds (HOST function; the partition a C/C++ host uses per rank) == shard.shard_bounds."""
    import ctypes as C
    from rte_rrtmgp_cpp_b200 import capi
    from rte_rrtmgp_cpp_b200.shard import shard_bounds
    L = capi.lib()
    a, b = C.c_longlong(), C.c_longlong()
    for ncol in (0, 1, 7, 100, 1048576, 3 * 10 ** 12 + 5):
        for world in (1, 2, 3, 8):
            prev = 0
            for r in range(world):
                assert L.rrtmgp_b200_shard_bounds(ncol, world, r, C.byref(a), C.byref(b)) == 0
                assert (a.value, b.value) == shard_bounds(ncol, world, r) and a.value == prev
                prev = b.value
            assert prev == ncol
    assert L.rrtmgp_b200_shard_bounds(10, 2, 2, C.byref(a), C.byref(b)) == capi.E_ARG
    assert b"shard_bounds" in L.rrtmgp_b200_last_error()


def test_shard_plan_host_logic_needs_no_device_and_no_nccl():
    """rrtmgp_b200_shard_plan_*: a single-process plan is pure host state; its bounds are rrtmgp_b200_shard_bounds; a
    multi-process plan without a communicator or unique id is refused (no NCCL call is made for any of this)."""
    L = capi.lib()
    h = C.c_void_p()
    assert L.rrtmgp_b200_shard_plan_create(1000003, 1, 0, None, None, C.byref(h)) == 0 and h.value
    a, b = C.c_longlong(), C.c_longlong()
    assert L.rrtmgp_b200_shard_plan_bounds(h, -1, C.byref(a), C.byref(b)) == 0 and (a.value, b.value) == (0, 1000003)
    assert L.rrtmgp_b200_shard_plan_bounds(h, 1, C.byref(a), C.byref(b)) == capi.E_ARG
    L.rrtmgp_b200_shard_plan_destroy(h)
    h = C.c_void_p()
    rc = L.rrtmgp_b200_shard_plan_create(100, 4, 1, None, None, C.byref(h))
    assert rc == capi.E_ARG and not h.value and b"unique id" in L.rrtmgp_b200_last_error()
    assert L.rrtmgp_b200_shard_plan_create(100, 4, 4, None, None, C.byref(h)) == capi.E_ARG
    assert L.rrtmgp_b200_gather_fluxes(None, capi.F32, 3, 1, None, None, 0, 1, 0, None) == capi.E_ARG
    L.rrtmgp_b200_shard_plan_destroy(None)


def test_host_pipeline_chunk_schedule():
    """rrtmgp_b200_host_chunk_schedule: the chunks of the host pipeline are contiguous, cover every column once, are whole CTAs
    (multiples of 32 columns) except the last of the equal run, ramp up 1-2-4 waves from both ends and never exceed 8 waves."""
    import ctypes as C
    L = capi.lib()
    W = 23680                                    # one resident wave of the fp32 paired kernel on a B200: 148 SMs x 5 CTAs x 32 columns

    def sched(ncol, wave, chunk=0):
        first, count, n = (C.c_int * 256)(), (C.c_int * 256)(), C.c_int()
        capi.check(L.rrtmgp_b200_host_chunk_schedule(ncol, wave, chunk, 256, first, count, C.byref(n)))
        return [(first[i], count[i]) for i in range(n.value)]

    for ncol in (1, 31, 100, W - 1, W, W + 1, 40000, 65536, 131072, 262144, 1048576, 1048577):
        s = sched(ncol, W)
        assert s[0][0] == 0 and sum(c for _, c in s) == ncol
        assert all(s[i][0] + s[i][1] == s[i + 1][0] for i in range(len(s) - 1))
        assert all(c > 0 and c <= 8 * W for _, c in s)
    assert sched(20000, W) == [(0, 20000)]                               # under 1.5 waves: one chunk
    assert sched(65536, W) == [(0, W), (W, 65536 - W)]                   # 1.5 .. 3 waves: a one-wave first chunk
    s = sched(131072, W)                                                 # a rank's shard of the 1024 x 1024 domain at 8 GPUs
    assert [c for _, c in s] == [W, 131072 - 2 * W, W]
    s = [c for _, c in sched(1048576, W)]                                # the whole domain on one GPU
    assert s[:3] == [W, 2 * W, 4 * W] and s[-3:] == [4 * W, 2 * W, W]
    mid = s[3:-3]
    assert len(set(mid[:-1])) == 1 and mid[0] % 32 == 0 and mid[0] <= 8 * W and sum(s) == 1048576
    # explicit chunk size: equal chunks of whole CTAs, no ramp
    s = sched(301, 0, 128)
    assert [c for _, c in s] == [128, 128, 45]
    assert L.rrtmgp_b200_host_chunk_schedule(0, W, 0, 0, None, None, None) == capi.E_ARG

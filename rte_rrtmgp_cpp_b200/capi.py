"""ctypes binding of the C ABI (include/rrtmgp_b200.h).  Loading fails loudly: there is no
Python or CPU fallback for any compute entry point."""
from __future__ import annotations

import ctypes as C
import os

from .ctables import KDistFileC, CloudFileC, AerosolFileC

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RB200_LIB", os.path.join(_HERE, "librrtmgp_b200.so"))
F32, F64, HOST_ONLY = 0, 1, 0x100
E_ARG, E_RANGE, E_MISSING, E_CUDA = 1, 2, 3, 4

vp = C.c_void_p
ip = C.POINTER(C.c_int)


class GasConcsC(C.Structure):
    _fields_ = [("ngas", C.c_int), ("d_vmr", C.POINTER(vp)), ("vmr_ncol", ip), ("vmr_nlay", ip)]


class GasOpticsInfoC(C.Structure):
    _fields_ = ([(n, C.c_int) for n in ("dtype", "is_longwave", "ngpt", "nbnd", "nflav", "ngas", "ntemp",
                                        "npres", "neta", "n_minor_lower", "n_minor_upper", "idx_h2o", "gpt_width")]
                + [(n, C.c_double) for n in ("press_ref_min", "press_ref_max", "temp_ref_min", "temp_ref_max", "tsi")])


class AerosolConcsC(C.Structure):
    _fields_ = [("d_mmr", vp * 11), ("mmr_ncol", C.c_int * 11), ("mmr_nlay", C.c_int * 11)]


class ColumnsC(C.Structure):
    _fields_ = [("ncol", C.c_int), ("nlay", C.c_int), ("top_at_1", C.c_int),
                ("d_play", vp), ("d_plev", vp), ("d_tlay", vp), ("d_tlev", vp), ("d_tsfc", vp), ("d_col_dry", vp),
                ("concs", GasConcsC),
                ("d_sfc_emis", vp), ("d_sfc_alb_dir", vp), ("d_sfc_alb_dif", vp), ("d_mu0", vp), ("d_tsi_scaling", vp),
                ("cloud", vp), ("delta_cloud", C.c_int),
                ("d_lwp", vp), ("d_iwp", vp), ("d_rel", vp), ("d_dei", vp), ("n_gauss_angles", C.c_int),
                ("aerosol", vp), ("delta_aerosol", C.c_int), ("d_rh", vp), ("aerosol_concs", AerosolConcsC)]


class FluxesC(C.Structure):
    _fields_ = [(n, vp) for n in ("d_flux_up", "d_flux_dn", "d_flux_net", "d_flux_dn_dir", "d_bnd_flux_up",
                                  "d_bnd_flux_dn", "d_bnd_flux_net", "d_bnd_flux_dn_dir", "d_heating_rate")]


# every symbol include/rrtmgp_b200.h declares: name -> (restype, argtypes)
i, sz, d, cp = C.c_int, C.c_size_t, C.c_double, C.c_char_p
SYMBOLS = {
    "rrtmgp_b200_last_error": (cp, []),
    "rrtmgp_b200_version": (i, []),
    "rrtmgp_b200_gas_optics_create": (i, [C.POINTER(KDistFileC), i, cp, i, i, C.POINTER(vp)]),
    "rrtmgp_b200_gas_optics_destroy": (None, [vp]),
    "rrtmgp_b200_gas_optics_get_info": (i, [vp, C.POINTER(GasOpticsInfoC)]),
    "rrtmgp_b200_gas_optics_get_tables": (i, [vp, ip, C.POINTER(d), ip, ip, ip, C.c_char_p, i, C.POINTER(d)]),
    "rrtmgp_b200_gas_optics_table_hash": (i, [vp, C.POINTER(C.c_ulonglong)]),
    "rrtmgp_b200_gas_optics_get_minor": (i, [vp, i, ip, ip, ip, ip, ip, ip]),
    "rrtmgp_b200_get_col_dry": (i, [i, i, i, vp, vp, vp, vp]),
    "rrtmgp_b200_interpolation": (i, [vp, i, i, vp, vp, vp, C.POINTER(GasConcsC), vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "rrtmgp_b200_gas_optics_lw": (i, [vp, i, i, vp, vp, vp, vp, vp, vp, C.POINTER(GasConcsC), vp, vp, vp, vp, vp, vp]),
    "rrtmgp_b200_gas_optics_sw": (i, [vp, i, i, vp, vp, vp, vp, C.POINTER(GasConcsC), vp, vp, vp, vp, vp]),
    "rrtmgp_b200_check_range": (i, [i, sz, vp, d, d, cp, vp]),
    "rrtmgp_b200_cloud_optics_create": (i, [C.POINTER(CloudFileC), i, C.POINTER(vp)]),
    "rrtmgp_b200_cloud_optics_destroy": (None, [vp]),
    "rrtmgp_b200_cloud_optics_nband": (i, [vp]),
    "rrtmgp_b200_cloud_optics_2str": (i, [vp, i, i, vp, vp, vp, vp, vp, vp, vp, vp]),
    "rrtmgp_b200_cloud_optics_1scl": (i, [vp, i, i, vp, vp, vp, vp, vp, vp]),
    "rrtmgp_b200_aerosol_optics_create": (i, [C.POINTER(AerosolFileC), i, C.POINTER(vp)]),
    "rrtmgp_b200_aerosol_optics_destroy": (None, [vp]),
    "rrtmgp_b200_aerosol_optics_nband": (i, [vp]),
    "rrtmgp_b200_aerosol_optics_2str": (i, [vp, i, i, C.POINTER(AerosolConcsC), vp, vp, vp, vp, vp, vp]),
    "rrtmgp_b200_delta_scale_2str": (i, [i, i, i, i, vp, vp, vp, vp]),
    "rrtmgp_b200_increment_1scl": (i, [i, i, i, i, vp, vp, vp]),
    "rrtmgp_b200_increment_2str": (i, [i, i, i, i, vp, vp, vp, vp, vp, vp, vp]),
    "rrtmgp_b200_increment_1scl_bybnd": (i, [i, i, i, i, vp, vp, i, ip, vp]),
    "rrtmgp_b200_increment_2str_bybnd": (i, [i, i, i, i, vp, vp, vp, vp, vp, vp, i, ip, vp]),
    "rrtmgp_b200_expand_and_transpose": (i, [i, i, i, i, ip, vp, vp, vp]),
    "rrtmgp_b200_lw_solver_noscat": (i, [i, i, i, i, i, i, vp, vp, vp, vp, vp, vp, vp, vp, i, vp, vp, vp]),
    "rrtmgp_b200_lw_solver_rescaled": (i, [i, i, i, i, i, i, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i, vp, vp, vp]),
    "rrtmgp_b200_sw_solver_2stream": (i, [i, i, i, i, i, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i, vp]),
    "rrtmgp_b200_sw_solver_2stream_mu0lay": (i, [i, i, i, i, i, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i, vp]),
    "rrtmgp_b200_sw_solver_2stream_noalloc": (i, [i, i, i, i, i, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "rrtmgp_b200_sum_broadband": (i, [i, i, i, i, vp, vp, vp]),
    "rrtmgp_b200_sum_byband": (i, [i, i, i, i, i, ip, vp, vp, vp]),
    "rrtmgp_b200_net_flux": (i, [i, sz, vp, vp, vp, vp]),
    "rrtmgp_b200_scale_columns": (i, [i, i, i, vp, vp, vp]),
    "rrtmgp_b200_heating_rate": (i, [i, i, i, vp, vp, vp, vp, vp]),
    "rrtmgp_b200_array_fill": (i, [i, C.c_longlong, vp, vp, vp]),
    "rrtmgp_b200_array_subset": (i, [i, i, ip, ip, ip, ip, vp, vp, vp]),
    "rrtmgp_b200_lw_fluxes": (i, [vp, C.POINTER(ColumnsC), C.POINTER(FluxesC), vp]),
    "rrtmgp_b200_sw_fluxes": (i, [vp, C.POINTER(ColumnsC), C.POINTER(FluxesC), vp]),
    "rrtmgp_b200_lwsw_fluxes": (i, [vp, C.POINTER(ColumnsC), C.POINTER(FluxesC), vp, C.POINTER(ColumnsC), C.POINTER(FluxesC), vp]),
    "rrtmgp_b200_profile_pair": (i, [C.POINTER(d), C.POINTER(C.c_longlong)]),
    "rrtmgp_b200_host_chunk_schedule": (i, [i, i, i, i, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "rrtmgp_b200_fluxes_host": (i, [vp, C.POINTER(ColumnsC), C.POINTER(FluxesC), vp, C.POINTER(ColumnsC), C.POINTER(FluxesC), i, vp]),
    "rrtmgp_b200_set_tile_columns": (i, [vp, i]),
    "rrtmgp_b200_set_debug": (i, [vp, i]),
    "rrtmgp_b200_set_fused_path": (i, [vp, i]),
    "rrtmgp_b200_launch_count": (C.c_longlong, []),
    "rrtmgp_b200_shard_bounds": (i, [C.c_longlong, i, i, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]),
    "rrtmgp_b200_nccl_unique_id": (i, [vp]),
    "rrtmgp_b200_shard_plan_create": (i, [C.c_longlong, i, i, vp, vp, C.POINTER(vp)]),
    "rrtmgp_b200_shard_plan_destroy": (None, [vp]),
    "rrtmgp_b200_shard_plan_bounds": (i, [vp, i, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]),
    "rrtmgp_b200_shard_plan_register": (i, [vp, i, C.POINTER(vp), i, vp]),
    "rrtmgp_b200_shard_plan_registered": (i, [vp]),
    "rrtmgp_b200_gather_fluxes": (i, [vp, i, i, i, C.POINTER(vp), C.POINTER(vp), i, i, i, vp]),
    "rrtmgp_b200_profile": (i, [i, C.POINTER(d), C.POINTER(C.c_longlong), C.POINTER(d), C.POINTER(C.c_longlong)]),
}

_lib = None


def lib():
    """Load (never build-on-demand on a GPU box: the .so is built by __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`. "
                               "There is no CPU fallback for the radiation kernels.")
        _lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            f = getattr(_lib, name)
            f.restype, f.argtypes = res, args
    return _lib


class RangeError(ValueError):
    """std::range_error of the reference (inputs outside the k-distribution's validity range)."""


def check(rc: int) -> None:
    if rc == 0:
        return
    msg = lib().rrtmgp_b200_last_error().decode()
    if rc == E_RANGE:
        raise RangeError(msg)
    raise RuntimeError(msg)

"""Host-side mirror of the reference's L3 class API over the C ABI (include/rrtmgp_b200.h).

Class and method names follow the reference's GPU twins (paths relative to the reference root):
  Array_gpu                 include/Array.h:355-650
  Gas_concs_gpu             include/Gas_concs.h:73-91
  Gas_optics_rrtmgp_gpu     include/Gas_optics_rrtmgp.h:293-408
  Optical_props_{1scl,2str}_gpu, add_to   include/Optical_props.h:165-293
  Source_func_lw_gpu        include/Source_functions.h:70-93
  Cloud_optics_gpu          include/Cloud_optics.h
  Aerosol_optics_gpu, Aerosol_concs_gpu   include/Aerosol_optics.h:20,48-84
  Rte_lw_gpu / Rte_sw_gpu   include/Rte_lw.h:62-77, include/Rte_sw.h:62-80
  Fluxes_broadband_gpu / Fluxes_byband_gpu  include/Fluxes.h:127-210
  Radiation_solver_{longwave,shortwave}.solve_gpu  src_test/Radiation_solver.cu:419-950 (fused)
PyTorch is used for device memory and streams only; all arithmetic is in librrtmgp_b200.so.
Arrays are column-major with the reference's dimension order; `Array_gpu.dims` holds those
dims and `.t` is the flat device tensor.  Errors: RuntimeError (std::runtime_error),
capi.RangeError (std::range_error).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Sequence

import numpy as np
import torch

from . import capi
from .capi import check, lib, F32, F64
from .ctables import pack_kdist, pack_cloud, pack_aerosol, pack_strings
from .kdist import KDistFile, CloudLUTFile, AerosolLUTFile, STRING_LEN


def _code(dtype: torch.dtype) -> int:
    if dtype == torch.float32:
        return F32
    if dtype == torch.float64:
        return F64
    raise TypeError("Float must be torch.float32 or torch.float64")


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


class Array_gpu:
    """Column-major device array (include/Array.h:355-650): dims as in the reference."""

    def __init__(self, dims: Sequence[int], dtype=torch.float64, device="cuda", data: Optional[torch.Tensor] = None):
        self.dims = tuple(int(d) for d in dims)
        n = int(np.prod(self.dims)) if self.dims else 0
        self.t = data if data is not None else torch.zeros(n, dtype=dtype, device=device)
        assert self.t.numel() == n and self.t.is_contiguous()

    @classmethod
    def from_numpy(cls, a, dtype=torch.float64, device="cuda") -> "Array_gpu":
        a = np.asarray(a)
        np_dtype = {torch.float32: np.float32, torch.float64: np.float64, torch.int32: np.int32}[dtype]
        flat = np.asfortranarray(a, dtype=np_dtype).reshape(-1, order="F")
        return cls(a.shape, dtype, device, torch.from_numpy(flat.copy()).to(device))

    def numpy(self) -> np.ndarray:
        return self.t.detach().cpu().numpy().reshape(self.dims, order="F")

    def dim(self, i: int) -> int:          # 1-based like the reference
        return self.dims[i - 1]

    def size(self) -> int:
        return self.t.numel()

    @property
    def ptr(self) -> int:
        return self.t.data_ptr()

    def fill(self, v) -> None:
        self.t.fill_(v)


def _ptr(a: Optional[Array_gpu]):
    return None if a is None else C.c_void_p(a.ptr)


class Gas_concs_gpu:
    """name -> vmr as (1,1) scalar, (1,nlay) profile or (ncol,nlay) field; range check 0..1
    (src/Gas_concs.cpp:28-117)."""

    def __init__(self, dtype=torch.float64, device="cuda"):
        self.dtype, self.device = dtype, device
        self.gas_concs_map: Dict[str, Array_gpu] = {}

    def set_vmr(self, name: str, vmr) -> None:
        a = np.atleast_2d(np.asarray(vmr.numpy() if isinstance(vmr, Array_gpu) else vmr, dtype=np.float64))
        if np.any(a < 0.0) or np.any(a > 1.0):
            raise capi.RangeError("Gas concentration values should be between 0 and 1")
        self.gas_concs_map[name] = Array_gpu.from_numpy(a, self.dtype, self.device)

    def get_vmr(self, name: str) -> Array_gpu:
        return self.gas_concs_map[name]

    def exists(self, name: str) -> bool:
        return name in self.gas_concs_map

    def names(self):
        return list(self.gas_concs_map.keys())

    def _pack(self, gas_names: Sequence[str]):
        n = len(gas_names)
        ptrs = (C.c_void_p * n)(); nc = (C.c_int * n)(); nl = (C.c_int * n)()
        for i, g in enumerate(gas_names):
            if g not in self.gas_concs_map:
                raise RuntimeError("Gas_concs: gas " + g + " not set")
            a = self.gas_concs_map[g]
            ptrs[i], nc[i], nl[i] = a.ptr, a.dims[0], a.dims[1]
        s = capi.GasConcsC(n, ptrs, nc, nl)
        s._keep = (ptrs, nc, nl)
        return s


class Optical_props_gpu:
    """band <-> g-point bookkeeping (src/Optical_props.cpp:31-72)."""

    def __init__(self, band_lims_wvn: np.ndarray, band_lims_gpt: Optional[np.ndarray] = None):
        self.band_lims_wvn = np.asfortranarray(band_lims_wvn, dtype=np.float64)   # (2, nbnd)
        nb = self.band_lims_wvn.shape[1]
        if band_lims_gpt is None:
            band_lims_gpt = np.stack([np.arange(1, nb + 1), np.arange(1, nb + 1)])
        self.band2gpt = np.asfortranarray(band_lims_gpt, dtype=np.int32)          # (2, nbnd) 1-based
        self.gpt2band = np.zeros(int(self.band2gpt.max()), dtype=np.int32)
        for ib in range(nb):
            self.gpt2band[self.band2gpt[0, ib] - 1:self.band2gpt[1, ib]] = ib + 1

    def get_nband(self) -> int:
        return self.band2gpt.shape[1]

    def get_ngpt(self) -> int:
        return int(self.band2gpt.max())

    def get_band_lims_gpoint(self) -> np.ndarray:
        return self.band2gpt

    def get_band_lims_wavenumber(self) -> np.ndarray:
        return self.band_lims_wvn

    def get_gpoint_bands(self) -> np.ndarray:
        return self.gpt2band

    def _lims_c(self):
        a = np.ascontiguousarray(self.band2gpt.T.reshape(-1), dtype=np.int32)     # (2,nbnd) column-major flat
        return a, a.ctypes.data_as(C.POINTER(C.c_int))


class Optical_props_1scl_gpu(Optical_props_gpu):
    def __init__(self, ncol: int, nlay: int, optical_props: Optical_props_gpu, dtype=torch.float64):
        super().__init__(optical_props.band_lims_wvn, optical_props.band2gpt)
        self.dtype = dtype
        self.tau = Array_gpu((ncol, nlay, self.get_ngpt()), dtype)

    def get_tau(self): return self.tau
    def get_ncol(self): return self.tau.dims[0]
    def get_nlay(self): return self.tau.dims[1]
    def delta_scale(self, forward_frac=None): pass      # no-op for 1scl (src/Optical_props.cpp:130)


class Optical_props_2str_gpu(Optical_props_gpu):
    def __init__(self, ncol: int, nlay: int, optical_props: Optical_props_gpu, dtype=torch.float64):
        super().__init__(optical_props.band_lims_wvn, optical_props.band2gpt)
        self.dtype = dtype
        dims = (ncol, nlay, self.get_ngpt())
        self.tau, self.ssa, self.g = Array_gpu(dims, dtype), Array_gpu(dims, dtype), Array_gpu(dims, dtype)

    def get_tau(self): return self.tau
    def get_ssa(self): return self.ssa
    def get_g(self): return self.g
    def get_ncol(self): return self.tau.dims[0]
    def get_nlay(self): return self.tau.dims[1]

    def delta_scale(self, forward_frac=None) -> None:
        """src/Optical_props.cpp:207-216 (the forward_frac argument is ignored there too)."""
        n = self.tau.dims
        check(lib().rrtmgp_b200_delta_scale_2str(_code(self.dtype), n[0], n[1], n[2], _ptr(self.tau),
                                                 _ptr(self.ssa), _ptr(self.g), _stream()))


def add_to(op_inout, op_in) -> None:
    """Optical_props increment (src/Optical_props.cpp:219-268)."""
    n = op_inout.tau.dims
    code = _code(op_inout.dtype)
    two = isinstance(op_inout, Optical_props_2str_gpu)
    if two != isinstance(op_in, Optical_props_2str_gpu):
        raise RuntimeError("add_to: optical property types differ")
    if op_in.tau.dims[:2] != n[:2]:
        raise RuntimeError("add_to: ncol/nlay mismatch")
    same = op_in.get_ngpt() == op_inout.get_ngpt()
    if not same and op_in.get_ngpt() != op_inout.get_nband():
        raise RuntimeError("Cannot add optical properties with incompatible band - gpoint combination")
    L = lib()
    if two:
        if same:
            check(L.rrtmgp_b200_increment_2str(code, n[0], n[1], n[2], _ptr(op_inout.tau), _ptr(op_inout.ssa),
                                               _ptr(op_inout.g), _ptr(op_in.tau), _ptr(op_in.ssa), _ptr(op_in.g), _stream()))
        else:
            keep, lims = op_inout._lims_c()
            check(L.rrtmgp_b200_increment_2str_bybnd(code, n[0], n[1], n[2], _ptr(op_inout.tau), _ptr(op_inout.ssa),
                                                     _ptr(op_inout.g), _ptr(op_in.tau), _ptr(op_in.ssa), _ptr(op_in.g),
                                                     op_inout.get_nband(), lims, _stream()))
    else:
        if same:
            check(L.rrtmgp_b200_increment_1scl(code, n[0], n[1], n[2], _ptr(op_inout.tau), _ptr(op_in.tau), _stream()))
        else:
            keep, lims = op_inout._lims_c()
            check(L.rrtmgp_b200_increment_1scl_bybnd(code, n[0], n[1], n[2], _ptr(op_inout.tau), _ptr(op_in.tau),
                                                     op_inout.get_nband(), lims, _stream()))


class Source_func_lw_gpu(Optical_props_gpu):
    """include/Source_functions.h:70-93"""

    def __init__(self, n_col: int, n_lay: int, optical_props: Optical_props_gpu, dtype=torch.float64):
        super().__init__(optical_props.band_lims_wvn, optical_props.band2gpt)
        ng = self.get_ngpt()
        self.sfc_source = Array_gpu((n_col, ng), dtype)
        self.sfc_source_jac = Array_gpu((n_col, ng), dtype)
        self.lay_source = Array_gpu((n_col, n_lay, ng), dtype)
        self.lev_source = Array_gpu((n_col, n_lay + 1, ng), dtype)

    def get_sfc_source(self): return self.sfc_source
    def get_sfc_source_jac(self): return self.sfc_source_jac
    def get_lay_source(self): return self.lay_source
    def get_lev_source(self): return self.lev_source


class Gas_optics_rrtmgp_gpu(Optical_props_gpu):
    """k-distribution object (include/Gas_optics_rrtmgp.h:293-408).  Constructed from the file
    variables (KDistFile) and the caller's Gas_concs, like load_and_init_gas_optics
    (src_test/Radiation_solver.cpp:70-284)."""

    def __init__(self, kdist: KDistFile, gas_concs, dtype=torch.float64, host_only: bool = False):
        names = gas_concs.names() if isinstance(gas_concs, Gas_concs_gpu) else list(gas_concs)
        self.dtype = dtype
        f, self._keep = pack_kdist(kdist)
        self.h = C.c_void_p()
        check(lib().rrtmgp_b200_gas_optics_create(C.byref(f), len(names), pack_strings(names), STRING_LEN,
                                                  _code(dtype) | (capi.HOST_ONLY if host_only else 0), C.byref(self.h)))
        info = capi.GasOpticsInfoC()
        check(lib().rrtmgp_b200_gas_optics_get_info(self.h, C.byref(info)))
        self.info = info
        nb, ng, nf, ngas = info.nbnd, info.ngpt, info.nflav, info.ngas
        lims = np.zeros(2 * nb, np.int32); wvn = np.zeros(2 * nb); gb = np.zeros(ng, np.int32)
        flav = np.zeros(2 * nf, np.int32); gf = np.zeros(2 * ng, np.int32); sol = np.zeros(ng)
        names_buf = C.create_string_buffer(ngas * STRING_LEN)
        check(lib().rrtmgp_b200_gas_optics_get_tables(
            self.h, lims.ctypes.data_as(capi.ip), wvn.ctypes.data_as(C.POINTER(C.c_double)),
            gb.ctypes.data_as(capi.ip), flav.ctypes.data_as(capi.ip), gf.ctypes.data_as(capi.ip),
            names_buf, STRING_LEN, sol.ctypes.data_as(C.POINTER(C.c_double))))
        super().__init__(wvn.reshape(2, nb, order="F"), lims.reshape(2, nb, order="F"))
        self.flavor = flav.reshape(2, nf, order="F")
        self.gpoint_flavor = gf.reshape(2, ng, order="F")
        raw = names_buf.raw
        self.gas_names = [raw[k * STRING_LEN:(k + 1) * STRING_LEN].decode().strip() for k in range(ngas)]
        self.solar_source = sol

    def __del__(self):
        if getattr(self, "h", None) and self.h.value:
            try:
                lib().rrtmgp_b200_gas_optics_destroy(self.h)
            except Exception:
                pass
            self.h = None

    def get_minor(self, side: int) -> dict:
        n = self.info.n_minor_lower if side == 0 else self.info.n_minor_upper
        a = {k: np.zeros(n * (2 if k == "limits" else 1), np.int32) for k in
             ("idx_minor", "idx_scal", "kstart", "limits", "dens", "comp")}
        check(lib().rrtmgp_b200_gas_optics_get_minor(self.h, side, *[a[k].ctypes.data_as(capi.ip) for k in
                                                     ("idx_minor", "idx_scal", "kstart", "limits", "dens", "comp")]))
        a["limits"] = a["limits"].reshape(2, n, order="F")
        return a

    # getters: include/Gas_optics_rrtmgp.h:371-386
    def source_is_internal(self): return bool(self.info.is_longwave)
    def source_is_external(self): return not self.info.is_longwave
    def get_press_ref_min(self): return self.info.press_ref_min
    def get_press_ref_max(self): return self.info.press_ref_max
    def get_temp_min(self): return self.info.temp_ref_min
    def get_temp_max(self): return self.info.temp_ref_max
    def get_tsi(self): return self.info.tsi
    def get_nflav(self): return self.info.nflav

    @staticmethod
    def get_col_dry(col_dry: Array_gpu, vmr_h2o: Array_gpu, plev: Array_gpu) -> None:
        ncol, nlay = vmr_h2o.dims
        check(lib().rrtmgp_b200_get_col_dry(_code(vmr_h2o.t.dtype), ncol, nlay, _ptr(vmr_h2o), _ptr(plev),
                                            _ptr(col_dry), _stream()))

    def _validate(self, **arrs) -> None:
        """std::range_error checks of the CPU path (src/Gas_optics_rrtmgp.cpp:812-827)."""
        i = self.info
        for name, a in arrs.items():
            if a is None:
                continue
            lo, hi = (i.press_ref_min, i.press_ref_max) if name.startswith("p") else (i.temp_ref_min, i.temp_ref_max)
            if name == "col_dry":
                lo, hi = 0.0, float("inf")
            check(lib().rrtmgp_b200_check_range(_code(self.dtype), a.size(), _ptr(a), lo, hi, name.encode(), _stream()))

    def interpolation(self, play, tlay, col_dry, gas_desc: Gas_concs_gpu):
        """materialised rrtmgp_interpolation; returns dict of reference-layout arrays"""
        ncol, nlay = play.dims
        nf, ngas = self.info.nflav, self.info.ngas
        o = dict(jtemp=Array_gpu((ncol, nlay), torch.int32), jpress=Array_gpu((ncol, nlay), torch.int32),
                 tropo=Array_gpu((ncol, nlay), torch.int8), jeta=Array_gpu((2, ncol, nlay, nf), torch.int32),
                 col_mix=Array_gpu((2, ncol, nlay, nf), self.dtype), fminor=Array_gpu((2, 2, ncol, nlay, nf), self.dtype),
                 fmajor=Array_gpu((2, 2, 2, ncol, nlay, nf), self.dtype), col_gas=Array_gpu((ncol, nlay, ngas + 1), self.dtype))
        concs = gas_desc._pack(self.gas_names)
        check(lib().rrtmgp_b200_interpolation(self.h, ncol, nlay, _ptr(play), _ptr(tlay), _ptr(col_dry), C.byref(concs),
                                              _ptr(o["jtemp"]), _ptr(o["jpress"]), _ptr(o["tropo"]), _ptr(o["jeta"]),
                                              _ptr(o["col_mix"]), _ptr(o["fminor"]), _ptr(o["fmajor"]), _ptr(o["col_gas"]),
                                              _stream()))
        return o

    def gas_optics(self, play, plev, tlay, *args, validate: bool = True):
        """LW: gas_optics(play, plev, tlay, tsfc, gas_desc, optical_props, sources, col_dry, tlev)
           SW: gas_optics(play, plev, tlay, gas_desc, optical_props, toa_src, col_dry)
        (include/Gas_optics_rrtmgp.h:389-408); col_dry may be None (computed from h2o)."""
        ncol, nlay = play.dims
        if self.info.is_longwave:
            tsfc, gas_desc, optical_props, sources, col_dry, tlev = args
            if validate:
                self._validate(play=play, plev=plev, tlay=tlay, tlev=tlev, tsfc=tsfc, col_dry=col_dry)
            concs = gas_desc._pack(self.gas_names)
            check(lib().rrtmgp_b200_gas_optics_lw(
                self.h, ncol, nlay, _ptr(play), _ptr(plev), _ptr(tlay), _ptr(tsfc), _ptr(tlev), _ptr(col_dry),
                C.byref(concs), _ptr(optical_props.tau), _ptr(sources.lay_source), _ptr(sources.lev_source),
                _ptr(sources.sfc_source), _ptr(sources.sfc_source_jac), _stream()))
        else:
            gas_desc, optical_props, toa_src, col_dry = args
            if validate:
                self._validate(play=play, plev=plev, tlay=tlay, col_dry=col_dry)
            concs = gas_desc._pack(self.gas_names)
            two = isinstance(optical_props, Optical_props_2str_gpu)
            check(lib().rrtmgp_b200_gas_optics_sw(
                self.h, ncol, nlay, _ptr(play), _ptr(plev), _ptr(tlay), _ptr(col_dry), C.byref(concs),
                _ptr(optical_props.tau), _ptr(optical_props.ssa) if two else None,
                _ptr(optical_props.g) if two else None, _ptr(toa_src), _stream()))


class Cloud_optics_gpu(Optical_props_gpu):
    """include/Cloud_optics.h; LUT cloud optics by band (src/Cloud_optics.cpp:29-232)."""

    def __init__(self, lut: CloudLUTFile, dtype=torch.float64):
        super().__init__(lut.bnd_limits_wavenumber.T)
        self.dtype = dtype
        c, self._keep = pack_cloud(lut)
        self.h = C.c_void_p()
        check(lib().rrtmgp_b200_cloud_optics_create(C.byref(c), _code(dtype), C.byref(self.h)))

    def __del__(self):
        if getattr(self, "h", None) and self.h.value:
            try:
                lib().rrtmgp_b200_cloud_optics_destroy(self.h)
            except Exception:
                pass
            self.h = None

    def cloud_optics(self, clwp, ciwp, reliq, deice, optical_props) -> None:
        ncol, nlay = clwp.dims
        if optical_props.get_ngpt() != self.get_nband():
            raise RuntimeError("cloud_optics: optical_props must be by band")
        if isinstance(optical_props, Optical_props_2str_gpu):
            check(lib().rrtmgp_b200_cloud_optics_2str(self.h, ncol, nlay, _ptr(clwp), _ptr(ciwp), _ptr(reliq), _ptr(deice),
                                                      _ptr(optical_props.tau), _ptr(optical_props.ssa), _ptr(optical_props.g), _stream()))
        else:
            check(lib().rrtmgp_b200_cloud_optics_1scl(self.h, ncol, nlay, _ptr(clwp), _ptr(ciwp), _ptr(reliq), _ptr(deice),
                                                      _ptr(optical_props.tau), _stream()))


Aerosol_concs_gpu = Gas_concs_gpu          # include/Aerosol_optics.h:20
AEROSOL_NAMES = tuple(f"aermr{i:02d}" for i in range(1, 12))


def _pack_aerosol_concs(aerosol_concs: Gas_concs_gpu) -> capi.AerosolConcsC:
    ac = capi.AerosolConcsC()
    for i, name in enumerate(AEROSOL_NAMES):
        if not aerosol_concs.exists(name):
            raise RuntimeError(f"Aerosol_concs: {name} not set")
        a = aerosol_concs.get_vmr(name)
        ac.d_mmr[i], ac.mmr_ncol[i], ac.mmr_nlay[i] = a.ptr, a.dims[0], a.dims[1]
    return ac


class Aerosol_optics_gpu(Optical_props_gpu):
    """include/Aerosol_optics.h:48-84; LUT aerosol optics by shortwave band (src_cuda/Aerosol_optics.cu:9-279,
    CPU src/Aerosol_optics.cpp:10-224).  Like the reference's loader the band limits are left zero."""

    def __init__(self, lut: AerosolLUTFile, dtype=torch.float64):
        super().__init__(np.zeros((2, lut.nbnd)))
        self.dtype = dtype
        c, self._keep = pack_aerosol(lut)
        self.h = C.c_void_p()
        check(lib().rrtmgp_b200_aerosol_optics_create(C.byref(c), _code(dtype), C.byref(self.h)))

    def __del__(self):
        if getattr(self, "h", None) and self.h.value:
            try:
                lib().rrtmgp_b200_aerosol_optics_destroy(self.h)
            except Exception:
                pass
            self.h = None

    def aerosol_optics(self, aerosol_concs: Gas_concs_gpu, rh: Array_gpu, plev: Array_gpu, optical_props) -> None:
        ncol, nlay = rh.dims
        if optical_props.get_ngpt() != self.get_nband():
            raise RuntimeError("aerosol_optics: optical_props must be by band")
        ac = _pack_aerosol_concs(aerosol_concs)
        check(lib().rrtmgp_b200_aerosol_optics_2str(self.h, ncol, nlay, C.byref(ac), _ptr(rh), _ptr(plev),
                                                    _ptr(optical_props.tau), _ptr(optical_props.ssa), _ptr(optical_props.g), _stream()))


class Rte_lw_gpu:
    @staticmethod
    def expand_and_transpose(ops: Optical_props_gpu, arr_in: Array_gpu) -> Array_gpu:
        nbnd, ncol = arr_in.dims
        out = Array_gpu((ncol, ops.get_ngpt()), arr_in.t.dtype)
        keep, lims = ops._lims_c()
        check(lib().rrtmgp_b200_expand_and_transpose(_code(arr_in.t.dtype), ncol, nbnd, ops.get_ngpt(), lims,
                                                     _ptr(arr_in), _ptr(out), _stream()))
        return out

    @staticmethod
    def rte_lw(optical_props, top_at_1: bool, sources: Source_func_lw_gpu, sfc_emis: Array_gpu,
               inc_flux: Optional[Array_gpu], gpt_flux_up: Array_gpu, gpt_flux_dn: Array_gpu, n_gauss_angles: int,
               sfc_src_jac: Optional[Array_gpu] = None, flux_up_jac: Optional[Array_gpu] = None,
               do_rescaling: bool = False) -> None:
        """include/Rte_lw.h:62-77; broadband iff gpt_flux_up.dim(3) == 1 (src/Rte_lw.cpp:170).  do_rescaling (false
        in the reference, src/Rte_lw.cpp:178) needs 2-stream optical_props: the kernel's Tang et al. re-scaled transport."""
        ncol, nlay, ngpt = optical_props.tau.dims
        emis = Rte_lw_gpu.expand_and_transpose(optical_props, sfc_emis)
        bb = len(gpt_flux_up.dims) < 3 or gpt_flux_up.dims[2] == 1
        inc = inc_flux if (inc_flux is not None and inc_flux.size() > 0) else None
        if do_rescaling:
            if getattr(optical_props, "ssa", None) is None:
                raise RuntimeError("rte_lw: do_rescaling needs 2-stream optical properties (tau, ssa, g)")
            check(lib().rrtmgp_b200_lw_solver_rescaled(
                _code(optical_props.dtype), ncol, nlay, ngpt, int(top_at_1), n_gauss_angles, _ptr(optical_props.tau),
                _ptr(optical_props.ssa), _ptr(optical_props.g), _ptr(sources.lay_source), _ptr(sources.lev_source),
                _ptr(emis), _ptr(sources.sfc_source), _ptr(inc), _ptr(gpt_flux_up), _ptr(gpt_flux_dn), int(bb),
                _ptr(sfc_src_jac), _ptr(flux_up_jac), _stream()))
            return
        check(lib().rrtmgp_b200_lw_solver_noscat(
            _code(optical_props.dtype), ncol, nlay, ngpt, int(top_at_1), n_gauss_angles, _ptr(optical_props.tau),
            _ptr(sources.lay_source), _ptr(sources.lev_source), _ptr(emis), _ptr(sources.sfc_source), _ptr(inc),
            _ptr(gpt_flux_up), _ptr(gpt_flux_dn), int(bb), _ptr(sfc_src_jac), _ptr(flux_up_jac), _stream()))


class Rte_sw_gpu:
    @staticmethod
    def rte_sw(optical_props, top_at_1: bool, mu0: Array_gpu, inc_flux_dir: Array_gpu, sfc_alb_dir: Array_gpu,
               sfc_alb_dif: Array_gpu, inc_flux_dif: Optional[Array_gpu], gpt_flux_up: Array_gpu,
               gpt_flux_dn: Array_gpu, gpt_flux_dir: Array_gpu) -> None:
        """include/Rte_sw.h:62-80; has_dif_bc iff inc_flux_dif.size() > 0 (src/Rte_sw.cpp:136-184).  mu0 is (ncol) as in
        the reference, or the kernel's own (ncol, nlay): a height-dependent zenith cosine (F90:517,1061-1065)."""
        ncol, nlay, ngpt = optical_props.tau.dims
        a_dir = Rte_lw_gpu.expand_and_transpose(optical_props, sfc_alb_dir)
        a_dif = Rte_lw_gpu.expand_and_transpose(optical_props, sfc_alb_dif)
        bb = len(gpt_flux_up.dims) < 3 or gpt_flux_up.dims[2] == 1
        dif = inc_flux_dif if (inc_flux_dif is not None and inc_flux_dif.size() > 0) else None
        by_layer = len(mu0.dims) == 2
        if tuple(mu0.dims) not in ((ncol,), (ncol, nlay)):
            raise RuntimeError(f"rte_sw: mu0 must be (ncol) or (ncol, nlay), got {tuple(mu0.dims)}")
        solver = lib().rrtmgp_b200_sw_solver_2stream_mu0lay if by_layer else lib().rrtmgp_b200_sw_solver_2stream
        check(solver(
            _code(optical_props.dtype), ncol, nlay, ngpt, int(top_at_1), _ptr(optical_props.tau), _ptr(optical_props.ssa),
            _ptr(optical_props.g), _ptr(mu0), _ptr(a_dir), _ptr(a_dif), _ptr(inc_flux_dir), _ptr(dif),
            _ptr(gpt_flux_up), _ptr(gpt_flux_dn), _ptr(gpt_flux_dir), int(bb), _stream()))


class Fluxes_broadband_gpu:
    """include/Fluxes.h:127-167"""

    def __init__(self, ncol: int, nlev: int, dtype=torch.float64):
        self.dtype = dtype
        self.flux_up, self.flux_dn = Array_gpu((ncol, nlev), dtype), Array_gpu((ncol, nlev), dtype)
        self.flux_dn_dir, self.flux_net = Array_gpu((ncol, nlev), dtype), Array_gpu((ncol, nlev), dtype)

    def net_flux(self) -> None:
        check(lib().rrtmgp_b200_net_flux(_code(self.dtype), self.flux_up.size(), _ptr(self.flux_dn), _ptr(self.flux_up),
                                         _ptr(self.flux_net), _stream()))

    def reduce(self, gpt_flux_up, gpt_flux_dn, spectral_disc, top_at_1: bool, gpt_flux_dn_dir=None) -> None:
        ncol, nlev, ngpt = gpt_flux_up.dims
        code = _code(self.dtype)
        for src, dst in ((gpt_flux_up, self.flux_up), (gpt_flux_dn, self.flux_dn), (gpt_flux_dn_dir, self.flux_dn_dir)):
            if src is not None:
                check(lib().rrtmgp_b200_sum_broadband(code, ncol, nlev, ngpt, _ptr(src), _ptr(dst), _stream()))
        self.net_flux()

    def get_flux_up(self): return self.flux_up
    def get_flux_dn(self): return self.flux_dn
    def get_flux_dn_dir(self): return self.flux_dn_dir
    def get_flux_net(self): return self.flux_net


class Fluxes_byband_gpu(Fluxes_broadband_gpu):
    """include/Fluxes.h:169-210"""

    def __init__(self, ncol: int, nlev: int, nbnd: int, dtype=torch.float64):
        super().__init__(ncol, nlev, dtype)
        d = (ncol, nlev, nbnd)
        self.bnd_flux_up, self.bnd_flux_dn = Array_gpu(d, dtype), Array_gpu(d, dtype)
        self.bnd_flux_dn_dir, self.bnd_flux_net = Array_gpu(d, dtype), Array_gpu(d, dtype)

    def reduce(self, gpt_flux_up, gpt_flux_dn, spectral_disc, top_at_1: bool, gpt_flux_dn_dir=None) -> None:
        super().reduce(gpt_flux_up, gpt_flux_dn, spectral_disc, top_at_1, gpt_flux_dn_dir)
        ncol, nlev, ngpt = gpt_flux_up.dims
        code = _code(self.dtype)
        keep, lims = spectral_disc._lims_c()
        nb = spectral_disc.get_nband()
        for src, dst in ((gpt_flux_up, self.bnd_flux_up), (gpt_flux_dn, self.bnd_flux_dn), (gpt_flux_dn_dir, self.bnd_flux_dn_dir)):
            if src is not None:
                check(lib().rrtmgp_b200_sum_byband(code, ncol, nlev, ngpt, nb, lims, _ptr(src), _ptr(dst), _stream()))
        check(lib().rrtmgp_b200_net_flux(code, self.bnd_flux_up.size(), _ptr(self.bnd_flux_dn), _ptr(self.bnd_flux_up),
                                         _ptr(self.bnd_flux_net), _stream()))

    def get_bnd_flux_up(self): return self.bnd_flux_up
    def get_bnd_flux_dn(self): return self.bnd_flux_dn
    def get_bnd_flux_dn_dir(self): return self.bnd_flux_dn_dir
    def get_bnd_flux_net(self): return self.bnd_flux_net


def heating_rate(flux_up: Array_gpu, flux_dn: Array_gpu, plev: Array_gpu) -> Array_gpu:
    ncol, nlev = flux_up.dims
    hr = Array_gpu((ncol, nlev - 1), flux_up.t.dtype)
    check(lib().rrtmgp_b200_heating_rate(_code(flux_up.t.dtype), ncol, nlev - 1, _ptr(flux_up), _ptr(flux_dn), _ptr(plev),
                                         _ptr(hr), _stream()))
    return hr


class DeviceColumns:
    """Inputs of one radiation call resident on the device (what test_rte_rrtmgp.cu's run_solver
    uploads before the timed region, src_test/test_rte_rrtmgp.cu)."""

    def __init__(self, atm, dtype=torch.float64, device="cuda"):
        A = lambda a: None if a is None else Array_gpu.from_numpy(a, dtype, device)
        self.dtype = dtype
        self.ncol, self.nlay = atm.play.shape
        self.top_at_1 = bool(atm.top_at_1)
        self.play, self.plev, self.tlay, self.tlev = A(atm.play), A(atm.plev), A(atm.tlay), A(atm.tlev)
        self.tsfc = A(atm.tsfc)
        self.gas_concs = Gas_concs_gpu(dtype, device)
        for k, v in atm.vmr.items():
            self.gas_concs.set_vmr(k, v)
        self.emis_sfc, self.sfc_alb_dir, self.sfc_alb_dif = A(atm.emis_sfc), A(atm.sfc_alb_dir), A(atm.sfc_alb_dif)
        self.mu0, self.tsi_scaling = A(atm.mu0), A(atm.tsi_scaling)
        self.lwp, self.iwp, self.rel, self.dei = A(atm.lwp), A(atm.iwp), A(atm.rel), A(atm.dei)
        self.rh = A(getattr(atm, "rh", None))
        self.aerosol_concs = Aerosol_concs_gpu(dtype, device)
        for k, v in (getattr(atm, "aermr", None) or {}).items():
            self.aerosol_concs.set_vmr(k, v)
        self.col_dry = None

    def input_bytes(self) -> int:
        n = 0
        for v in vars(self).values():
            if isinstance(v, Array_gpu):
                n += v.t.numel() * v.t.element_size()
        for a in list(self.gas_concs.gas_concs_map.values()) + list(self.aerosol_concs.gas_concs_map.values()):
            n += a.t.numel() * a.t.element_size()
        return n


class _Radiation_solver:
    """Fused equivalent of Radiation_solver_{longwave,shortwave}::solve_gpu
    (src_test/Radiation_solver.cu:419-657, 690-950): gas optics -> [cloud optics -> delta scale ->
    add_to] -> rte -> Fluxes::reduce in one C-ABI call, per-g-point arrays never leave the chip."""
    sw = False

    def __init__(self, kdist: KDistFile, gas_concs, cloud_lut: Optional[CloudLUTFile] = None, dtype=torch.float64,
                 aerosol_lut: Optional[AerosolLUTFile] = None):
        self.dtype = dtype
        self.kdist = Gas_optics_rrtmgp_gpu(kdist, gas_concs, dtype)
        self.cloud = Cloud_optics_gpu(cloud_lut, dtype) if cloud_lut is not None else None
        self.aerosol = Aerosol_optics_gpu(aerosol_lut, dtype) if aerosol_lut is not None else None

    def set_tile_columns(self, n: int) -> None:
        check(lib().rrtmgp_b200_set_tile_columns(self.kdist.h, int(n)))

    def _pack(self, cols, out, switch_cloud_optics, switch_delta_cloud, switch_output_bnd_fluxes, heating, device,
              switch_aerosol_optics=False, switch_delta_aerosol=False):
        """ColumnsC / FluxesC for `cols` (device or pinned-host resident), allocating outputs in `out`."""
        k = self.kdist
        ncol, nlay, nlev, nb = cols.ncol, cols.nlay, cols.nlay + 1, k.get_nband()
        dt = self.dtype
        def arr(name, dims):
            if name not in out or out[name].dims != tuple(dims):
                n = int(np.prod(dims))
                data = torch.zeros(n, dtype=dt, pin_memory=True) if device == "cpu" else None
                out[name] = Array_gpu(dims, dt, device, data)
            return out[name]
        f = capi.FluxesC()
        f.d_flux_up, f.d_flux_dn, f.d_flux_net = arr("flux_up", (ncol, nlev)).ptr, arr("flux_dn", (ncol, nlev)).ptr, arr("flux_net", (ncol, nlev)).ptr
        if self.sw:
            f.d_flux_dn_dir = arr("flux_dn_dir", (ncol, nlev)).ptr
        if switch_output_bnd_fluxes:
            f.d_bnd_flux_up, f.d_bnd_flux_dn = arr("bnd_flux_up", (ncol, nlev, nb)).ptr, arr("bnd_flux_dn", (ncol, nlev, nb)).ptr
            f.d_bnd_flux_net = arr("bnd_flux_net", (ncol, nlev, nb)).ptr
            if self.sw:
                f.d_bnd_flux_dn_dir = arr("bnd_flux_dn_dir", (ncol, nlev, nb)).ptr
        if heating:
            f.d_heating_rate = arr("heating_rate", (ncol, nlay)).ptr
        c = capi.ColumnsC()
        c.ncol, c.nlay, c.top_at_1 = ncol, nlay, int(cols.top_at_1)
        c.d_play, c.d_plev, c.d_tlay = cols.play.ptr, cols.plev.ptr, cols.tlay.ptr
        c.d_col_dry = cols.col_dry.ptr if cols.col_dry is not None else None
        concs = cols.gas_concs._pack(k.gas_names)
        c.concs = concs
        if self.sw:
            c.d_sfc_alb_dir, c.d_sfc_alb_dif, c.d_mu0 = cols.sfc_alb_dir.ptr, cols.sfc_alb_dif.ptr, cols.mu0.ptr
            c.d_tsi_scaling = cols.tsi_scaling.ptr if cols.tsi_scaling is not None else None
        else:
            c.d_tlev, c.d_tsfc, c.d_sfc_emis = cols.tlev.ptr, cols.tsfc.ptr, cols.emis_sfc.ptr
        if switch_cloud_optics:
            if self.cloud is None or cols.lwp is None:
                raise RuntimeError("cloud optics requested without a cloud LUT / cloud inputs")
            c.cloud = self.cloud.h
            c.delta_cloud = int(switch_delta_cloud)
            c.d_lwp, c.d_iwp, c.d_rel, c.d_dei = cols.lwp.ptr, cols.iwp.ptr, cols.rel.ptr, cols.dei.ptr
        if switch_aerosol_optics:
            if not self.sw:
                raise RuntimeError("aerosol optics are shortwave only")
            if self.aerosol is None or cols.rh is None:
                raise RuntimeError("aerosol optics requested without an aerosol LUT / aerosol inputs")
            c.aerosol = self.aerosol.h
            c.delta_aerosol = int(switch_delta_aerosol)
            c.d_rh = cols.rh.ptr
            c.aerosol_concs = _pack_aerosol_concs(cols.aerosol_concs)
        c.n_gauss_angles = int(getattr(self, "n_gauss_angles", 1))
        c._keep = concs
        return c, f

    def set_debug(self, flags: int) -> None:
        """1: NaN-poison the workspace before and check canary blocks after every fused call (rrtmgp_b200_set_debug)."""
        check(lib().rrtmgp_b200_set_debug(self.kdist.h, int(flags)))

    def set_fused_path(self, path: int) -> None:
        """0 per-band kernels with HBM scratch strips (default), 1 column-resident kernels."""
        check(lib().rrtmgp_b200_set_fused_path(self.kdist.h, int(path)))

    def solve_gpu(self, cols: DeviceColumns, switch_cloud_optics: bool = False, switch_delta_cloud: bool = True,
                  switch_output_bnd_fluxes: bool = False, heating: bool = False, out: Optional[dict] = None,
                  switch_aerosol_optics: bool = False, switch_delta_aerosol: bool = False) -> dict:
        if out is None:
            out = {}
        c, f = self._pack(cols, out, switch_cloud_optics, switch_delta_cloud, switch_output_bnd_fluxes, heating, "cuda",
                          switch_aerosol_optics, switch_delta_aerosol)
        fn = lib().rrtmgp_b200_sw_fluxes if self.sw else lib().rrtmgp_b200_lw_fluxes
        check(fn(self.kdist.h, C.byref(c), C.byref(f), _stream()))
        return out


class Radiation_solver_longwave(_Radiation_solver):
    sw = False


class Radiation_solver_shortwave(_Radiation_solver):
    sw = True


class HostColumns(DeviceColumns):
    """The same inputs resident in pinned HOST memory: what a host model (MicroHH's CPU fields, a
    NetCDF reader) holds before a radiation call.  Consumed by solve_host()."""

    def __init__(self, atm, dtype=torch.float64):
        super().__init__(atm, dtype, device="cpu")
        for obj in (vars(self), self.gas_concs.gas_concs_map, self.aerosol_concs.gas_concs_map):
            for v in obj.values():
                if isinstance(v, Array_gpu):
                    v.t = v.t.pin_memory()


def solve_pair(lw: Radiation_solver_longwave, sw: Radiation_solver_shortwave, cols: DeviceColumns,
               switch_cloud_optics: bool = False, switch_delta_cloud: bool = True, switch_output_bnd_fluxes: bool = False,
               heating: bool = False, out_lw: Optional[dict] = None, out_sw: Optional[dict] = None,
               switch_aerosol_optics: bool = False, switch_delta_aerosol: bool = False):
    """Longwave and shortwave solve_gpu of the same device-resident columns in one call (rrtmgp_b200_lwsw_fluxes): band i of
    both k-distributions in one launch with longwave and shortwave CTAs co-resident on every SM.  Bit-identical to
    lw.solve_gpu(...) + sw.solve_gpu(...).  Returns (out_lw, out_sw)."""
    out_lw = {} if out_lw is None else out_lw
    out_sw = {} if out_sw is None else out_sw
    cl, fl = lw._pack(cols, out_lw, switch_cloud_optics, switch_delta_cloud, switch_output_bnd_fluxes, heating, "cuda")
    cs, fs = sw._pack(cols, out_sw, switch_cloud_optics, switch_delta_cloud, switch_output_bnd_fluxes, heating, "cuda",
                      switch_aerosol_optics, switch_delta_aerosol)
    check(lib().rrtmgp_b200_lwsw_fluxes(lw.kdist.h, C.byref(cl), C.byref(fl), sw.kdist.h, C.byref(cs), C.byref(fs), _stream()))
    return out_lw, out_sw


def solve_host(lw: Optional[Radiation_solver_longwave], sw: Optional[Radiation_solver_shortwave], cols: HostColumns,
               switch_cloud_optics: bool = False, switch_delta_cloud: bool = True, switch_output_bnd_fluxes: bool = False,
               heating: bool = False, out_lw: Optional[dict] = None, out_sw: Optional[dict] = None, chunk_columns: int = 0,
               switch_aerosol_optics: bool = False, switch_delta_aerosol: bool = False):
    """Host buffers in, host buffers out (rrtmgp_b200_fluxes_host): column chunks stream through
    H2D -> fused LW and SW -> D2H with the copies hidden behind the kernels.  Returns
    (out_lw, out_sw): dicts of pinned host Array_gpu objects.  Blocks until the outputs are ready."""
    out_lw = {} if out_lw is None else out_lw
    out_sw = {} if out_sw is None else out_sw
    cl = fl = cs = fs = None
    if lw is not None:
        cl, fl = lw._pack(cols, out_lw, switch_cloud_optics, switch_delta_cloud, switch_output_bnd_fluxes, heating, "cpu")
    if sw is not None:
        cs, fs = sw._pack(cols, out_sw, switch_cloud_optics, switch_delta_cloud, switch_output_bnd_fluxes, heating, "cpu",
                          switch_aerosol_optics, switch_delta_aerosol)
    check(lib().rrtmgp_b200_fluxes_host(lw.kdist.h if lw else None, C.byref(cl) if cl else None, C.byref(fl) if fl else None,
                                        sw.kdist.h if sw else None, C.byref(cs) if cs else None, C.byref(fs) if fs else None,
                                        int(chunk_columns), _stream()))
    return out_lw, out_sw

"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes binding of oracle/_ref/liboracle_ref.so: the reference's own unmodified C++ class layer
(built by `make -C oracle ref` from /root/reference/src, when that tree is present) running over
the restated kernels.  `available()` is False when the library was not built / did not travel.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict

import numpy as np

from rte_rrtmgp_cpp_b200.ctables import pack_kdist, pack_cloud, pack_aerosol, pack_strings, KDistFileC, CloudFileC
from rte_rrtmgp_cpp_b200.kdist import STRING_LEN

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "_ref", "liboracle_ref.so")
# `make -C oracle ref_fast`: the same sources at the reference's release optimisation level (-O3 -march=x86-64-v3 -DNDEBUG).
# TIMING ONLY — bench.py's CPU legs select it with ORACLE_REF_FAST=1; parity tests always use the -O2 -ffp-contract=off build.
_PATH_FAST = os.path.join(_HERE, "_ref", "liboracle_ref_fast.so")
FLAGS = {"parity": "g++ -O2 -ffp-contract=off", "fast": "g++ -O3 -march=x86-64-v3 -DNDEBUG"}
_lib = None
c_dbl_p = C.POINTER(C.c_double)


def fast_selected() -> bool:
    return os.environ.get("ORACLE_REF_FAST") == "1" and os.path.exists(_PATH_FAST)


def available() -> bool:
    return os.path.exists(_PATH_FAST if fast_selected() else _PATH)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(_PATH_FAST if fast_selected() else _PATH)
        _lib.ref_last_error.restype = C.c_char_p
    return _lib


def _chk(rc):
    if rc != 0:
        raise RuntimeError(lib().ref_last_error().decode())


def _F(a):
    return np.asfortranarray(a, dtype=np.float64)


def _p(a):
    return a.ctypes.data_as(c_dbl_p)


def _z(shape):
    return np.zeros(shape, order="F")


class RefGasOptics:
    def __init__(self, kd, available_gases):
        self._f, self._keep = pack_kdist(kd)
        names = list(available_gases)
        self.h = C.c_void_p()
        _chk(lib().ref_gas_optics_create(C.byref(self._f), len(names), pack_strings(names),
                                         STRING_LEN, C.byref(self.h)))
        ngpt, nbnd, nflav, ngas = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        _chk(lib().ref_gas_optics_info(self.h, C.byref(ngpt), C.byref(nbnd), C.byref(nflav), C.byref(ngas)))
        self.ngpt, self.nbnd, self.nflav = ngpt.value, nbnd.value, nflav.value

    def __del__(self):
        if getattr(self, "h", None):
            lib().ref_gas_optics_destroy(self.h)

    @staticmethod
    def _concs(vmr: Dict[str, np.ndarray]):
        names = list(vmr.keys())
        arrs = [_F(np.atleast_2d(vmr[n])) for n in names]
        ptrs = (c_dbl_p * len(names))(*[_p(a) for a in arrs])
        nc = (C.c_int * len(names))(*[a.shape[0] for a in arrs])
        nl = (C.c_int * len(names))(*[a.shape[1] for a in arrs])
        return len(names), pack_strings(names), ptrs, nc, nl, arrs

    @staticmethod
    def get_col_dry(vmr_h2o, plev):
        vmr_h2o, plev = _F(vmr_h2o), _F(plev)
        ncol, nlay = vmr_h2o.shape
        out = _z((ncol, nlay))
        _chk(lib().ref_get_col_dry(ncol, nlay, _p(vmr_h2o), _p(plev), _p(out)))
        return out

    def gas_optics_lw(self, play, plev, tlay, tsfc, vmr, col_dry, tlev):
        play, plev, tlay, tsfc, tlev, col_dry = map(_F, (play, plev, tlay, tsfc, tlev, col_dry))
        ncol, nlay = play.shape
        n, names, ptrs, nc, nl, keep = self._concs(vmr)
        tau = _z((ncol, nlay, self.ngpt)); lay = _z((ncol, nlay, self.ngpt))
        lev = _z((ncol, nlay + 1, self.ngpt)); sfc = _z((ncol, self.ngpt)); jac = _z((ncol, self.ngpt))
        _chk(lib().ref_gas_optics_lw(self.h, ncol, nlay, _p(play), _p(plev), _p(tlay), _p(tsfc), _p(tlev),
                                     _p(col_dry), n, names, STRING_LEN, ptrs, nc, nl,
                                     _p(tau), _p(lay), _p(lev), _p(sfc), _p(jac)))
        return dict(tau=tau, lay_src=lay, lev_src=lev, sfc_src=sfc, sfc_src_jac=jac)

    def gas_optics_sw(self, play, plev, tlay, vmr, col_dry):
        play, plev, tlay, col_dry = map(_F, (play, plev, tlay, col_dry))
        ncol, nlay = play.shape
        n, names, ptrs, nc, nl, keep = self._concs(vmr)
        tau = _z((ncol, nlay, self.ngpt)); ssa = _z((ncol, nlay, self.ngpt)); g = _z((ncol, nlay, self.ngpt))
        toa = _z((ncol, self.ngpt))
        _chk(lib().ref_gas_optics_sw(self.h, ncol, nlay, _p(play), _p(plev), _p(tlay), _p(col_dry),
                                     n, names, STRING_LEN, ptrs, nc, nl, _p(tau), _p(ssa), _p(g), _p(toa)))
        return dict(tau=tau, ssa=ssa, g=g, toa_src=toa)

    def rte_lw(self, tau, lay_src, lev_src, sfc_src, sfc_emis, top_at_1, n_ang=1, broadband=False):
        tau, lay_src, lev_src, sfc_src, sfc_emis = map(_F, (tau, lay_src, lev_src, sfc_src, sfc_emis))
        ncol, nlay, _ = tau.shape
        ng = 1 if broadband else self.ngpt
        up = _z((ncol, nlay + 1, ng)); dn = _z((ncol, nlay + 1, ng))
        _chk(lib().ref_rte_lw(self.h, ncol, nlay, int(top_at_1), n_ang, ng, _p(tau), _p(lay_src),
                              _p(lev_src), _p(sfc_src), _p(sfc_emis), _p(up), _p(dn)))
        return up, dn

    def rte_sw(self, tau, ssa, g, mu0, inc_flux_dir, sfc_alb_dir, sfc_alb_dif, top_at_1, broadband=False):
        tau, ssa, g, mu0, inc, ad, af = map(_F, (tau, ssa, g, mu0, inc_flux_dir, sfc_alb_dir, sfc_alb_dif))
        ncol, nlay, _ = tau.shape
        ng = 1 if broadband else self.ngpt
        up = _z((ncol, nlay + 1, ng)); dn = _z((ncol, nlay + 1, ng)); dr = _z((ncol, nlay + 1, ng))
        _chk(lib().ref_rte_sw(self.h, ncol, nlay, int(top_at_1), ng, _p(tau), _p(ssa), _p(g), _p(mu0),
                              _p(inc), _p(ad), _p(af), _p(up), _p(dn), _p(dr)))
        return up, dn, dr

    def cloud_add(self, lut, mode, do_delta, lwp, iwp, rel, dei, tau, ssa=None, g=None):
        """Cloud_optics + (delta_scale) + add_to on copies of the gas props; returns dict."""
        c, keep = pack_cloud(lut)
        lwp, iwp, rel, dei = map(_F, (lwp, iwp, rel, dei))
        ncol, nlay = lwp.shape
        tau = _F(tau).copy(order="F")
        ssa = _F(ssa).copy(order="F") if ssa is not None else _z((1,))
        g = _F(g).copy(order="F") if g is not None else _z((1,))
        ctau = _z((ncol, nlay, lut.nbnd)); cssa = _z((ncol, nlay, lut.nbnd)); cg = _z((ncol, nlay, lut.nbnd))
        _chk(lib().ref_cloud_add(self.h, C.byref(c), ncol, nlay, mode, int(do_delta), _p(lwp), _p(iwp),
                                 _p(rel), _p(dei), _p(tau), _p(ssa), _p(g), _p(ctau), _p(cssa), _p(cg)))
        return dict(tau=tau, ssa=ssa, g=g, ctau=ctau, cssa=cssa, cg=cg)

    def aerosol_add(self, lut, do_delta, aermr, rh, plev, tau, ssa, g):
        """Aerosol_optics + (delta_scale) + add_to on copies of the 2-stream gas props; returns dict."""
        a, keep = pack_aerosol(lut)
        rh, plev = _F(rh), _F(plev)
        ncol, nlay = rh.shape
        arrs = [_F(np.atleast_2d(aermr[f"aermr{i:02d}"])) for i in range(1, 12)]
        ptrs = (c_dbl_p * 11)(*[_p(x) for x in arrs])
        nc = (C.c_int * 11)(*[x.shape[0] for x in arrs]); nl = (C.c_int * 11)(*[x.shape[1] for x in arrs])
        tau, ssa, g = (_F(x).copy(order="F") for x in (tau, ssa, g))
        nb = lut.nbnd
        atau = _z((ncol, nlay, nb)); assa = _z((ncol, nlay, nb)); ag = _z((ncol, nlay, nb))
        _chk(lib().ref_aerosol_add(self.h, C.byref(a), ncol, nlay, int(do_delta), ptrs, nc, nl, _p(rh), _p(plev),
                                   _p(tau), _p(ssa), _p(g), _p(atau), _p(assa), _p(ag)))
        return dict(tau=tau, ssa=ssa, g=g, atau=atau, assa=assa, ag=ag)

    def fluxes_byband(self, up, dn, dr=None):
        up, dn = _F(up), _F(dn)
        ncol, nlev, _ = up.shape
        o = {k: _z((ncol, nlev)) for k in ("flux_up", "flux_dn", "flux_net", "flux_dn_dir")}
        o.update({k: _z((ncol, nlev, self.nbnd)) for k in ("bnd_flux_up", "bnd_flux_dn", "bnd_flux_net", "bnd_flux_dn_dir")})
        drp = _p(_F(dr)) if dr is not None else c_dbl_p()
        _chk(lib().ref_fluxes_byband(self.h, ncol, nlev, _p(up), _p(dn), drp,
                                     _p(o["flux_up"]), _p(o["flux_dn"]), _p(o["flux_net"]), _p(o["flux_dn_dir"]),
                                     _p(o["bnd_flux_up"]), _p(o["bnd_flux_dn"]), _p(o["bnd_flux_net"]),
                                     _p(o["bnd_flux_dn_dir"])))
        return o

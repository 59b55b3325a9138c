"""BASELINE — TEST AND BENCH INFRASTRUCTURE ONLY.

ctypes binding of oracle/_ref/librefcuda_{sp,dp}.so: the reference's own unmodified CUDA
implementation of the path (built by `make -C oracle refcuda` from /root/reference when that tree
is present; the .so travels to the GPU box).  Run as a module it times the reference's
solve_gpu call sequence on seeded S-RCE columns and prints one JSON line:

    python -m oracle.refcuda --ncol 65536 --nlay 128 --dtype f32 --reps 3
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(_HERE))
c_dbl_p = C.POINTER(C.c_double)


def path(dtype: str) -> str:
    return os.path.join(_HERE, "_ref", f"librefcuda_{'sp' if dtype == 'f32' else 'dp'}.so")


def available(dtype: str = "f32") -> bool:
    return os.path.exists(path(dtype))


def solve(atm, kd_lw, kd_sw, gases, dtype="f32", reps=3):
    """-> dict(lw_ms, sw_ms, lw_up, lw_dn, sw_up, sw_dn, sw_dir) from the reference CUDA classes."""
    from rte_rrtmgp_cpp_b200.ctables import pack_kdist, pack_strings
    from rte_rrtmgp_cpp_b200.kdist import STRING_LEN
    L = C.CDLL(path(dtype))
    L.refcuda_last_error.restype = C.c_char_p
    assert L.refcuda_float_bytes() == (4 if dtype == "f32" else 8)
    flw, k1 = pack_kdist(kd_lw)
    fsw, k2 = pack_kdist(kd_sw)
    names = list(gases)
    h = C.c_void_p()
    if L.refcuda_create(C.byref(flw), C.byref(fsw), len(names), pack_strings(names), STRING_LEN, C.byref(h)):
        raise RuntimeError(L.refcuda_last_error().decode())
    F = lambda a: np.asfortranarray(a, dtype=np.float64)
    P = lambda a: a.ctypes.data_as(c_dbl_p)
    vn = list(atm.vmr.keys())
    va = [F(np.atleast_2d(atm.vmr[n])) for n in vn]
    ptrs = (c_dbl_p * len(vn))(*[P(a) for a in va])
    nc = (C.c_int * len(vn))(*[a.shape[0] for a in va])
    nl = (C.c_int * len(vn))(*[a.shape[1] for a in va])
    ncol, nlay = atm.play.shape
    ins = [F(x) for x in (atm.play, atm.plev, atm.tlay, atm.tlev, atm.tsfc, atm.emis_sfc, atm.sfc_alb_dir, atm.sfc_alb_dif, atm.mu0)]
    outs = [np.zeros((ncol, nlay + 1), order="F") for _ in range(5)]
    ms_lw, ms_sw = C.c_double(), C.c_double()
    rc = L.refcuda_solve(h, ncol, nlay, len(vn), pack_strings(vn), STRING_LEN, ptrs, nc, nl, *[P(x) for x in ins], int(reps),
                         *[P(x) for x in outs], C.byref(ms_lw), C.byref(ms_sw))
    if rc:
        raise RuntimeError(L.refcuda_last_error().decode())
    L.refcuda_destroy(h)
    return dict(lw_ms=ms_lw.value, sw_ms=ms_sw.value, lw_up=outs[0], lw_dn=outs[1], sw_up=outs[2], sw_dn=outs[3], sw_dir=outs[4])


class RefCuda:
    """Stage-level access to the reference's own CUDA classes and kernels (refcuda_shim.cu `refcuda_stage_*`):
    the reference-authored implementation of rows a5-a9/a15/a17 that can run on the GPU box.  Used by
    tests/test_oracle_vs_refcuda.py to pin the C restatement of the Fortran kernels."""

    def __init__(self, kd_lw, kd_sw, gases, dtype="f64"):
        from rte_rrtmgp_cpp_b200.ctables import pack_kdist, pack_strings
        from rte_rrtmgp_cpp_b200.kdist import STRING_LEN
        self._pack_strings, self._len = pack_strings, STRING_LEN
        self.L = C.CDLL(path(dtype))
        self.L.refcuda_last_error.restype = C.c_char_p
        assert self.L.refcuda_float_bytes() == (4 if dtype == "f32" else 8)
        self.kd = {0: kd_lw, 1: kd_sw}
        self._keep = []
        f = []
        for kd in (kd_lw, kd_sw):
            if kd is None:
                f.append(None)
            else:
                fk, keep = pack_kdist(kd)
                self._keep.append((fk, keep))
                f.append(C.byref(fk))
        names = list(gases)
        self.h = C.c_void_p()
        self._chk(self.L.refcuda_create(f[0], f[1], len(names), pack_strings(names), STRING_LEN, C.byref(self.h)))

    def _chk(self, rc):
        if rc:
            raise RuntimeError(self.L.refcuda_last_error().decode())

    def close(self):
        if getattr(self, "h", None):
            self.L.refcuda_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def gas_optics(self, which, atm, nflav, ngas):
        """which: 'lw' | 'sw'.  -> dict with the same keys and layouts as oracle.GasOptics.gas_optics_{lw,sw}.
        nflav, ngas: flavors and retained gases of the k-distribution (oracle.GasOptics.nflav / .ngas; checked)."""
        w = 0 if which == "lw" else 1
        kd = self.kd[w]
        F = lambda a: np.asfortranarray(a, dtype=np.float64)
        P = lambda a: a.ctypes.data_as(c_dbl_p)
        vn = list(atm.vmr.keys())
        va = [F(np.atleast_2d(atm.vmr[n])) for n in vn]
        ptrs = (c_dbl_p * len(vn))(*[P(a) for a in va])
        nc = (C.c_int * len(vn))(*[a.shape[0] for a in va])
        nl = (C.c_int * len(vn))(*[a.shape[1] for a in va])
        ncol, nlay = atm.play.shape
        ngpt = kd.ngpt
        Z = lambda shape, dt=np.float64: np.zeros(shape, dtype=dt, order="F")
        o = dict(jtemp=Z((ncol, nlay), np.int32), jpress=Z((ncol, nlay), np.int32), tropo=Z((ncol, nlay), np.int8),
                 jeta=Z((2, ncol, nlay, nflav), np.int32), fmajor=Z((2, 2, 2, ncol, nlay, nflav)),
                 fminor=Z((2, 2, ncol, nlay, nflav)), col_mix=Z((2, ncol, nlay, nflav)),
                 col_gas=Z((ncol, nlay, ngas + 1)), tau=Z((ncol, nlay, ngpt)), col_dry=Z((ncol, nlay)))
        if w == 0:
            o.update(lay_src=Z((ncol, nlay, ngpt)), lev_src=Z((ncol, nlay + 1, ngpt)), sfc_src=Z((ncol, ngpt)),
                     sfc_src_jac=Z((ncol, ngpt)))
        else:
            o.update(ssa=Z((ncol, nlay, ngpt)), g=Z((ncol, nlay, ngpt)), toa_src=Z((ncol, ngpt)))
        ins = [F(x) for x in (atm.play, atm.plev, atm.tlay, atm.tlev, atm.tsfc)]
        q = lambda k: P(o[k]) if k in o else c_dbl_p()
        ip = lambda k: o[k].ctypes.data_as(C.POINTER(C.c_int))
        nf, ng = C.c_int(nflav), C.c_int(ngas)          # in: what the buffers are sized for; the shim refuses a mismatch
        self._chk(self.L.refcuda_stage_gas_optics(
            self.h, w, ncol, nlay, len(vn), self._pack_strings(vn), self._len, ptrs, nc, nl, *[P(x) for x in ins],
            ip("jtemp"), ip("jpress"), o["tropo"].ctypes.data_as(C.POINTER(C.c_byte)), ip("jeta"),
            q("fmajor"), q("fminor"), q("col_mix"), q("col_gas"), q("tau"), q("ssa"), q("g"),
            q("lay_src"), q("lev_src"), q("sfc_src"), q("sfc_src_jac"), q("toa_src"), q("col_dry"), C.byref(nf), C.byref(ng)))
        assert (nf.value, ng.value) == (nflav, ngas), ("flavors / gases of the reference class differ from the caller's", nf.value, ng.value, nflav, ngas)
        return o

    def rte_lw(self, tau, lay_src, lev_src, sfc_src, emis, top_at_1):
        F = lambda a: np.asfortranarray(a, dtype=np.float64)
        P = lambda a: a.ctypes.data_as(c_dbl_p)
        tau, lay_src, lev_src, sfc_src, emis = map(F, (tau, lay_src, lev_src, sfc_src, emis))
        ncol, nlay, ngpt = tau.shape
        up, dn = (np.zeros((ncol, nlay + 1, ngpt), order="F") for _ in range(2))
        self._chk(self.L.refcuda_stage_rte_lw(self.h, ncol, nlay, int(top_at_1), P(tau), P(lay_src), P(lev_src), P(sfc_src),
                                              P(emis), P(up), P(dn)))
        return up, dn

    def rte_sw(self, tau, ssa, g, mu0, toa, alb_dir, alb_dif, top_at_1):
        F = lambda a: np.asfortranarray(a, dtype=np.float64)
        P = lambda a: a.ctypes.data_as(c_dbl_p)
        tau, ssa, g, mu0, toa, alb_dir, alb_dif = map(F, (tau, ssa, g, mu0, toa, alb_dir, alb_dif))
        ncol, nlay, ngpt = tau.shape
        up, dn, dr = (np.zeros((ncol, nlay + 1, ngpt), order="F") for _ in range(3))
        self._chk(self.L.refcuda_stage_rte_sw(self.h, ncol, nlay, int(top_at_1), P(tau), P(ssa), P(g), P(mu0), P(toa),
                                              P(alb_dir), P(alb_dif), P(up), P(dn), P(dr)))
        return up, dn, dr


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ncol", type=int, default=65536)
    ap.add_argument("--nlay", type=int, default=128)
    ap.add_argument("--dtype", default="f32", choices=["f32", "f64"])
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--check", action="store_true", help="also compare its fluxes with the CPU oracle on 256 columns")
    a = ap.parse_args()
    from rte_rrtmgp_cpp_b200.atmos import rce_atmosphere
    from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist
    gases = ["h2o", "co2", "o3", "n2o", "ch4", "n2", "o2"]      # bench.py:GASES
    kd_lw, kd_sw = synthetic_kdist("lw"), synthetic_kdist("sw")
    base = rce_atmosphere(min(a.ncol, 16384), a.nlay)
    reps_col = -(-a.ncol // base.ncol)
    atm = base
    if reps_col > 1:
        import copy
        atm = copy.copy(base)
        t2 = lambda x: np.tile(x, (reps_col, 1))[:a.ncol]
        for k in ("play", "plev", "tlay", "tlev"):
            setattr(atm, k, t2(getattr(base, k)))
        for k in ("tsfc", "mu0", "tsi_scaling"):
            setattr(atm, k, np.tile(getattr(base, k), reps_col)[:a.ncol])
        for k in ("emis_sfc", "sfc_alb_dir", "sfc_alb_dif"):
            setattr(atm, k, np.tile(getattr(base, k), (1, reps_col))[:, :a.ncol])
        atm.vmr = {k: (t2(v) if np.ndim(v) == 2 and np.shape(v)[0] == base.ncol else v) for k, v in base.vmr.items()}
    r = solve(atm, kd_lw, kd_sw, gases, a.dtype, a.reps)
    units = a.ncol * (kd_lw.ngpt + kd_sw.ngpt)
    out = {"impl": "reference-cuda", "value": units / ((r["lw_ms"] + r["sw_ms"]) * 1e-3), "unit": "column-gpts/s",
           "lw_ms": r["lw_ms"], "sw_ms": r["sw_ms"], "dtype": a.dtype,
           "sample": f"{a.ncol} seeded S-RCE columns x {a.nlay} layers, LW g256 + SW g224, clear-sky, the reference's "
                     f"solve_gpu call sequence in 1024-column blocks, inputs resident, mean of {a.reps} passes after one "
                     f"untimed (autotuning) pass"}
    if a.check:
        sys.path.insert(0, os.path.join(os.path.dirname(_HERE), "tests"))
        from helpers import oracle_lw, oracle_sw
        sub = base.subset(0, 256)
        o_lw, o_sw = oracle_lw(kd_lw, sub), oracle_sw(kd_sw, sub)
        out["max_abs_diff_vs_cpu_oracle_Wm2"] = {
            "lw_up": float(np.max(np.abs(r["lw_up"][:256] - o_lw["flux_up"]))), "lw_dn": float(np.max(np.abs(r["lw_dn"][:256] - o_lw["flux_dn"]))),
            "sw_up": float(np.max(np.abs(r["sw_up"][:256] - o_sw["flux_up"]))), "sw_dn": float(np.max(np.abs(r["sw_dn"][:256] - o_sw["flux_dn"])))}
    print(json.dumps(out))


if __name__ == "__main__":
    main()

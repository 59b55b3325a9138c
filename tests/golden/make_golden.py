"""Generates tests/golden/ref_fluxes.npz: fluxes of the reference's own, unmodified C++ class layer
(oracle/_ref = /root/reference/src/*.cpp compiled where they lie, over the restated Fortran kernels) on
small seeded column sets.  Run HERE (needs oracle/_ref; `make -C oracle ref`):

    python tests/golden/make_golden.py

The fixture holds the inputs too (it does not depend on the atmosphere generator staying put) and a
SHA-256 of every coefficient table the run used (the synthetic tables are too large to commit: a stale
generator fails the hash check instead of silently comparing different physics).

Chain per case = Radiation_solver_{longwave,shortwave}::solve of src_test/Radiation_solver.cpp:347-560,
572-800: gas optics -> [cloud optics (-> delta_scale) -> add_to] -> [aerosol optics (-> delta_scale)
-> add_to] -> rte_{lw,sw} -> Fluxes_byband::reduce.
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import ref as R                                                      # noqa: E402
from rte_rrtmgp_cpp_b200.atmos import rce_atmosphere, add_aerosols              # noqa: E402
from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist, synthetic_cloud_lut, synthetic_aerosol_lut  # noqa: E402

FIELDS = ("play", "plev", "tlay", "tlev", "tsfc", "emis_sfc", "sfc_alb_dir", "sfc_alb_dif", "mu0", "tsi_scaling",
          "lwp", "iwp", "rel", "dei", "rh")
# name: (kind, ncol, nlay, top_at_1, cloudy, delta_cloud, aerosols, delta_aerosol, n_gauss_angles)
CASES = {
    "lw_clear":        ("lw", 10, 32, False, False, False, False, False, 1),
    "lw_clear_top1":   ("lw", 6, 24, True, False, False, False, False, 1),
    "lw_cloudy_3ang":  ("lw", 8, 28, False, True, False, False, False, 3),
    "sw_clear":        ("sw", 10, 32, False, False, False, False, False, 1),
    "sw_cloudy_delta": ("sw", 8, 28, True, True, True, False, False, 1),
    "sw_cloud_aer":    ("sw", 8, 28, False, True, True, True, True, 1),
}


def table_hash(obj) -> str:
    h = hashlib.sha256()
    for k in sorted(vars(obj)):
        v = getattr(obj, k)
        if isinstance(v, np.ndarray):
            h.update(k.encode()); h.update(np.ascontiguousarray(v).tobytes())
    return h.hexdigest()


def case_atmosphere(ncol, nlay, top_at_1, cloudy, aerosols):
    atm = rce_atmosphere(ncol, nlay, top_at_1=top_at_1, night_fraction=0.2, cloudy=cloudy)
    return add_aerosols(atm) if aerosols else atm


def reference_fluxes(kind, atm, cloudy, delta_cloud, aerosols, delta_aerosol, n_ang):
    kd = synthetic_kdist(kind)
    rf = R.RefGasOptics(kd, atm.vmr.keys())
    col_dry = rf.get_col_dry(atm.vmr["h2o"], atm.plev)
    if kind == "lw":
        st = rf.gas_optics_lw(atm.play, atm.plev, atm.tlay, atm.tsfc, atm.vmr, col_dry, atm.tlev)
        tau = st["tau"]
        if cloudy:
            tau = rf.cloud_add(synthetic_cloud_lut("lw"), 1, False, atm.lwp, atm.iwp, atm.rel, atm.dei, tau)["tau"]
        up, dn = rf.rte_lw(tau, st["lay_src"], st["lev_src"], st["sfc_src"], atm.emis_sfc, atm.top_at_1, n_ang, False)
        return rf.fluxes_byband(up, dn)
    st = rf.gas_optics_sw(atm.play, atm.plev, atm.tlay, atm.vmr, col_dry)
    tau, ssa, g = st["tau"], st["ssa"], st["g"]
    if cloudy:
        r = rf.cloud_add(synthetic_cloud_lut("sw"), 2, delta_cloud, atm.lwp, atm.iwp, atm.rel, atm.dei, tau, ssa, g)
        tau, ssa, g = r["tau"], r["ssa"], r["g"]
    if aerosols:
        r = rf.aerosol_add(synthetic_aerosol_lut(), delta_aerosol, atm.aermr, atm.rh, atm.plev, tau, ssa, g)
        tau, ssa, g = r["tau"], r["ssa"], r["g"]
    toa = st["toa_src"] * np.asarray(atm.tsi_scaling)[:, None]
    up, dn, dr = rf.rte_sw(tau, ssa, g, atm.mu0, np.asfortranarray(toa), atm.sfc_alb_dir, atm.sfc_alb_dif, atm.top_at_1, False)
    return rf.fluxes_byband(up, dn, dr)


def main():
    assert R.available(), "oracle/_ref is not built: make -C oracle ref"
    out = {}
    for kind in ("lw", "sw"):
        out[f"sha/kdist_{kind}"] = np.array(table_hash(synthetic_kdist(kind)))
        out[f"sha/cloud_{kind}"] = np.array(table_hash(synthetic_cloud_lut(kind)))
    out["sha/aerosol"] = np.array(table_hash(synthetic_aerosol_lut()))
    for name, (kind, ncol, nlay, top1, cloudy, dcl, aer, daer, n_ang) in CASES.items():
        atm = case_atmosphere(ncol, nlay, top1, cloudy, aer)
        for f in FIELDS:
            v = getattr(atm, f, None)
            if v is not None:
                out[f"{name}/in/{f}"] = np.asarray(v)
        for k, v in atm.vmr.items():
            out[f"{name}/vmr/{k}"] = np.asarray(v)
        for k, v in (atm.aermr or {}).items():
            out[f"{name}/aermr/{k}"] = np.asarray(v)
        fl = reference_fluxes(kind, atm, cloudy, dcl, aer, daer, n_ang)
        for k, v in fl.items():
            out[f"{name}/out/{k}"] = np.asarray(v)
        print(name, {k: tuple(v.shape) for k, v in fl.items() if not k.startswith("bnd")},
              "max flux_dn", float(np.max(fl["flux_dn"])))
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_fluxes.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()

"""Generates tests/golden/rrtmgp_data.npz from the rrtmgp-data files of the reference mount (run HERE:
/root/reference does not exist on the GPU box):

    python tests/golden/make_real_fixtures.py

  cloud_{lw,sw}/...   the cloud-optics LUT variables of rrtmgp-clouds-{lw,sw}.nc the reference reads
                      (src_test/Radiation_solver.cpp:286-326), as stored (float32 / float64)
  rfmip/...           experiment 0 ("Present day (PD)") and 3 ("future": other gas amounts) of the RFMIP clear-sky input
                      (examples/rfmip-clear-sky/inputs/multiple_input4MIPs_...nc) as rfmip/rfmip_init.py builds them

  rfmip/<e>/pub_*     the PUBLISHED fluxes of those experiments (examples/rfmip-clear-sky/reference/r{ld,lu,sd,su}_Efx_...nc,
                      float32 (site, level), level 0 = top): what the reference's own rfmip test compares with

  allsky/...          inputs and PUBLISHED fluxes of the all-sky example (examples/all-sky/reference/rrtmgp-allsky-{lw,sw}-no-aerosols.nc,
                      float64 (lay|lev, col), bottom first): what allsky/allsky_init.py restates and allsky_run compares with

The gas-optics coefficient files (rrtmgp-gas-*.nc) are absent from the mount (.MISSING_LARGE_BLOBS), so the
published RFMIP / all-sky reference fluxes stay out of reach; these fixtures put REAL cloud tables and REAL
atmospheres through the oracle and the CUDA path instead of synthetic ones.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from rte_rrtmgp_cpp_b200 import ncfiles                        # noqa: E402

DATA = "/root/reference/rrtmgp-data"
RFMIP = DATA + "/examples/rfmip-clear-sky/inputs/multiple_input4MIPs_radiation_RFMIP_UColorado-RFMIP-1-2_none.nc"
PUBLISHED = DATA + "/examples/rfmip-clear-sky/reference/{}_Efx_RTE-RRTMGP-181204_rad-irf_r1i1p1f1_gn.nc"
LUT_VARS = ("bnd_limits_wavenumber", "radliq_lwr", "radliq_upr", "radliq_fac", "diamice_lwr", "diamice_upr", "diamice_fac",
            "lut_extliq", "lut_ssaliq", "lut_asyliq", "lut_extice", "lut_ssaice", "lut_asyice")
EXPTS = (0, 3)


def collect():
    out = {}
    for kind in ("lw", "sw"):
        r = ncfiles._Reader(f"{DATA}/rrtmgp-clouds-{kind}.nc")
        for v in LUT_VARS:
            out[f"cloud_{kind}/{v}"] = r.get(v)
        r.close()
    for e in EXPTS:
        d = ncfiles.read_rfmip_input(RFMIP, e)
        for k, v in d.items():
            if k != "vmr":
                out[f"rfmip/{e}/{k}"] = np.asarray(v, dtype=np.float32 if k not in ("plev",) else np.float64)
        for g, v in d["vmr"].items():
            out[f"rfmip/{e}/vmr/{g}"] = np.asarray(v, dtype=np.float32 if np.ndim(v) else np.float64)
    for k in ("rld", "rlu", "rsd", "rsu"):
        r = ncfiles._Reader(PUBLISHED.format(k))
        a = r.get(k)
        for e in EXPTS:
            out[f"rfmip/{e}/pub_{k}"] = a[e]
        r.close()
    for kind in ("lw", "sw"):
        r = ncfiles._Reader(f"{DATA}/examples/all-sky/reference/rrtmgp-allsky-{kind}-no-aerosols.nc")
        for v in ("p_lay", "p_lev", "t_lay", "t_lev", "h2o", "o3", "lwp", "iwp", "rel", "rei"):
            a = r.get(v)
            if kind == "lw":
                out[f"allsky/{v}"] = a
            else:
                assert np.array_equal(a, out[f"allsky/{v}"]), v             # the two files hold the same atmosphere
        for v in (("lw_flux_up", "lw_flux_dn") if kind == "lw" else ("sw_flux_up", "sw_flux_dn", "sw_flux_dir")):
            out[f"allsky/{v}"] = r.get(v)
        r.close()
    return out


if __name__ == "__main__":
    out = collect()
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "rrtmgp_data.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")

"""Loader of tests/golden/rrtmgp_data.npz (real cloud LUTs and RFMIP atmospheres of rrtmgp-data; made by
tests/golden/make_real_fixtures.py from the reference mount's netCDF-4 files through h5lite)."""
import os

import numpy as np

from rte_rrtmgp_cpp_b200.atmos import Atmosphere
from rte_rrtmgp_cpp_b200.kdist import CloudLUTFile

HERE = os.path.dirname(os.path.abspath(__file__))
PATH = os.path.join(HERE, "golden", "rrtmgp_data.npz")
_F = lambda a: np.asfortranarray(np.asarray(a, dtype=np.float64))


def fixture():
    return np.load(PATH)


def cloud_lut(z, kind: str) -> CloudLUTFile:
    g = lambda n: z[f"cloud_{kind}/{n}"]
    return CloudLUTFile(
        bnd_limits_wavenumber=np.ascontiguousarray(g("bnd_limits_wavenumber"), dtype=np.float64),
        **{n: float(g(n)) for n in ("radliq_lwr", "radliq_upr", "radliq_fac", "diamice_lwr", "diamice_upr", "diamice_fac")},
        **{n: np.ascontiguousarray(g(n), dtype=np.float64)
           for n in ("lut_extliq", "lut_ssaliq", "lut_asyliq", "lut_extice", "lut_ssaice", "lut_asyice")},
        label=f"rrtmgp-clouds-{kind}.nc")


def default_tsi(kd) -> float:
    """Gas_optics_rrtmgp::get_tsi of the default solar source (src/Gas_optics_rrtmgp.cpp:745-756): what the driver
    divides the input file's `tsi` by (test_rte_rrtmgp.cpp)."""
    s = (kd.solar_source_quiet + (kd.mg_default - 0.1495954) * kd.solar_source_facular
         + (kd.sb_default - 0.00066696) * kd.solar_source_sunspot)
    return float(np.sum(s))


def rfmip_atmosphere(z, expt: int, sites=slice(None), tsi_ref: float = 1360.85, nbnd_lw: int = 16, nbnd_sw: int = 14) -> Atmosphere:
    """The sites of one RFMIP experiment as the reference's driver sees them after rfmip_init.py: band-constant surface
    emissivity / albedo (rfmip_init.py:42-50), tsi_scaling = tsi / the k-distribution's tsi (test_rte_rrtmgp.cpp)."""
    g = lambda n: z[f"rfmip/{expt}/{n}"]
    ncol = g("play")[sites].shape[0]
    vmr = {}
    for k in z.files:
        if k.startswith(f"rfmip/{expt}/vmr/"):
            v = z[k]
            vmr[k.split("/")[-1]] = _F(v[sites]) if v.ndim == 2 else _F(np.full((1, 1), float(v)))
    return Atmosphere(
        play=_F(g("play")[sites]), plev=_F(g("plev")[sites]), tlay=_F(g("tlay")[sites]), tlev=_F(g("tlev")[sites]),
        tsfc=_F(g("tsfc")[sites]), vmr=vmr,
        emis_sfc=_F(np.tile(g("emis")[sites][None, :], (nbnd_lw, 1))),
        sfc_alb_dir=_F(np.tile(g("albedo")[sites][None, :], (nbnd_sw, 1))),
        sfc_alb_dif=_F(np.tile(g("albedo")[sites][None, :], (nbnd_sw, 1))),
        mu0=_F(g("mu0")[sites]), tsi_scaling=_F(g("tsi")[sites] / tsi_ref))

"""ctypes mirrors of include/rrtmgp_b200_tables.h and packers from KDistFile / CloudLUTFile."""
from __future__ import annotations

import ctypes as C
from typing import List, Sequence

import numpy as np

from .kdist import KDistFile, CloudLUTFile, STRING_LEN

c_dbl_p = C.POINTER(C.c_double)
c_int_p = C.POINTER(C.c_int)


class KDistFileC(C.Structure):
    _fields_ = (
        [(n, C.c_int) for n in (
            "n_temps", "n_press", "n_absorbers", "n_minorabsorbers", "n_extabsorbers",
            "n_mixingfracs", "n_layers", "n_bnds", "n_gpts", "n_pairs", "n_minor_lower",
            "n_minor_upper", "n_contrib_lower", "n_contrib_upper", "n_planck_temps", "string_len")]
        + [("gas_names", C.c_char_p), ("key_species", c_int_p),
           ("bnd_limits_wavenumber", c_dbl_p), ("bnd_limits_gpt", c_int_p),
           ("press_ref", c_dbl_p), ("temp_ref", c_dbl_p),
           ("absorption_coefficient_ref_P", C.c_double), ("absorption_coefficient_ref_T", C.c_double),
           ("press_ref_trop", C.c_double),
           ("vmr_ref", c_dbl_p), ("kmajor", c_dbl_p), ("kminor_lower", c_dbl_p), ("kminor_upper", c_dbl_p),
           ("gas_minor", C.c_char_p), ("identifier_minor", C.c_char_p),
           ("minor_gases_lower", C.c_char_p), ("minor_gases_upper", C.c_char_p),
           ("minor_limits_gpt_lower", c_int_p), ("minor_limits_gpt_upper", c_int_p),
           ("minor_scales_with_density_lower", c_int_p), ("minor_scales_with_density_upper", c_int_p),
           ("scale_by_complement_lower", c_int_p), ("scale_by_complement_upper", c_int_p),
           ("scaling_gas_lower", C.c_char_p), ("scaling_gas_upper", C.c_char_p),
           ("kminor_start_lower", c_int_p), ("kminor_start_upper", c_int_p),
           ("totplnk", c_dbl_p), ("plank_fraction", c_dbl_p),
           ("rayl_lower", c_dbl_p), ("rayl_upper", c_dbl_p),
           ("solar_source_quiet", c_dbl_p), ("solar_source_facular", c_dbl_p),
           ("solar_source_sunspot", c_dbl_p),
           ("tsi_default", C.c_double), ("mg_default", C.c_double), ("sb_default", C.c_double)])


class CloudFileC(C.Structure):
    _fields_ = [("n_band", C.c_int), ("n_rghice", C.c_int), ("n_size_liq", C.c_int), ("n_size_ice", C.c_int),
                ("bnd_limits_wavenumber", c_dbl_p),
                ("radliq_lwr", C.c_double), ("radliq_upr", C.c_double), ("radliq_fac", C.c_double),
                ("diamice_lwr", C.c_double), ("diamice_upr", C.c_double), ("diamice_fac", C.c_double),
                ("lut_extliq", c_dbl_p), ("lut_ssaliq", c_dbl_p), ("lut_asyliq", c_dbl_p),
                ("lut_extice", c_dbl_p), ("lut_ssaice", c_dbl_p), ("lut_asyice", c_dbl_p)]


class AerosolFileC(C.Structure):
    _fields_ = [("n_band", C.c_int), ("n_hum", C.c_int), ("n_philic", C.c_int), ("n_phobic", C.c_int),
                ("mass_ext_sw_hydrophobic", c_dbl_p), ("ssa_sw_hydrophobic", c_dbl_p), ("asymmetry_sw_hydrophobic", c_dbl_p),
                ("mass_ext_sw_hydrophilic", c_dbl_p), ("ssa_sw_hydrophilic", c_dbl_p), ("asymmetry_sw_hydrophilic", c_dbl_p),
                ("relative_humidity2", c_dbl_p)]


def pack_strings(names: Sequence[str], length: int = STRING_LEN) -> bytes:
    """Blank-padded fixed-length block, as netCDF char arrays store strings."""
    return b"".join(s.encode("ascii")[:length].ljust(length, b" ") for s in names)


def _d(a, keep: List):
    if a is None:
        return c_dbl_p()
    b = np.ascontiguousarray(a, dtype=np.float64)
    keep.append(b)
    return b.ctypes.data_as(c_dbl_p)


def _i(a, keep: List):
    b = np.ascontiguousarray(a, dtype=np.int32)
    keep.append(b)
    return b.ctypes.data_as(c_int_p)


def _s(names, keep: List):
    b = pack_strings(names) + b"\0"
    keep.append(b)
    return C.c_char_p(b)


def pack_kdist(kd: KDistFile):
    """-> (KDistFileC, keepalive list). The struct borrows the numpy buffers in keepalive."""
    keep: List = []
    f = KDistFileC()
    f.n_temps = kd.temp_ref.size
    f.n_press = kd.press_ref.size
    f.n_absorbers = len(kd.gas_names)
    f.n_minorabsorbers = len(kd.gas_minor)
    f.n_extabsorbers = kd.vmr_ref.shape[1]
    f.n_mixingfracs = kd.kmajor.shape[2]
    f.n_layers = kd.vmr_ref.shape[2]
    f.n_bnds = kd.nbnd
    f.n_gpts = kd.ngpt
    f.n_pairs = 2
    f.n_minor_lower = len(kd.minor_gases_lower)
    f.n_minor_upper = len(kd.minor_gases_upper)
    f.n_contrib_lower = kd.kminor_lower.shape[2]
    f.n_contrib_upper = kd.kminor_upper.shape[2]
    f.n_planck_temps = kd.totplnk.shape[1] if kd.totplnk is not None else 0
    f.string_len = STRING_LEN
    f.gas_names = _s(kd.gas_names, keep)
    f.key_species = _i(kd.key_species, keep)
    f.bnd_limits_wavenumber = _d(kd.bnd_limits_wavenumber, keep)
    f.bnd_limits_gpt = _i(kd.bnd_limits_gpt, keep)
    f.press_ref = _d(kd.press_ref, keep)
    f.temp_ref = _d(kd.temp_ref, keep)
    f.absorption_coefficient_ref_P = kd.absorption_coefficient_ref_P
    f.absorption_coefficient_ref_T = kd.absorption_coefficient_ref_T
    f.press_ref_trop = kd.press_ref_trop
    f.vmr_ref = _d(kd.vmr_ref, keep)
    f.kmajor = _d(kd.kmajor, keep)
    f.kminor_lower = _d(kd.kminor_lower, keep)
    f.kminor_upper = _d(kd.kminor_upper, keep)
    f.gas_minor = _s(kd.gas_minor, keep)
    f.identifier_minor = _s(kd.identifier_minor, keep)
    f.minor_gases_lower = _s(kd.minor_gases_lower, keep)
    f.minor_gases_upper = _s(kd.minor_gases_upper, keep)
    f.minor_limits_gpt_lower = _i(kd.minor_limits_gpt_lower, keep)
    f.minor_limits_gpt_upper = _i(kd.minor_limits_gpt_upper, keep)
    f.minor_scales_with_density_lower = _i(kd.minor_scales_with_density_lower, keep)
    f.minor_scales_with_density_upper = _i(kd.minor_scales_with_density_upper, keep)
    f.scale_by_complement_lower = _i(kd.scale_by_complement_lower, keep)
    f.scale_by_complement_upper = _i(kd.scale_by_complement_upper, keep)
    f.scaling_gas_lower = _s(kd.scaling_gas_lower, keep)
    f.scaling_gas_upper = _s(kd.scaling_gas_upper, keep)
    f.kminor_start_lower = _i(kd.kminor_start_lower, keep)
    f.kminor_start_upper = _i(kd.kminor_start_upper, keep)
    f.totplnk = _d(kd.totplnk, keep)
    f.plank_fraction = _d(kd.plank_fraction, keep)
    f.rayl_lower = _d(kd.rayl_lower, keep)
    f.rayl_upper = _d(kd.rayl_upper, keep)
    f.solar_source_quiet = _d(kd.solar_source_quiet, keep)
    f.solar_source_facular = _d(kd.solar_source_facular, keep)
    f.solar_source_sunspot = _d(kd.solar_source_sunspot, keep)
    f.tsi_default, f.mg_default, f.sb_default = kd.tsi_default, kd.mg_default, kd.sb_default
    return f, keep


def pack_cloud(lut: CloudLUTFile):
    keep: List = []
    c = CloudFileC()
    c.n_band = lut.nbnd
    c.n_rghice = lut.lut_extice.shape[0]
    c.n_size_liq = lut.lut_extliq.shape[1]
    c.n_size_ice = lut.lut_extice.shape[2]
    c.bnd_limits_wavenumber = _d(lut.bnd_limits_wavenumber, keep)
    c.radliq_lwr, c.radliq_upr, c.radliq_fac = lut.radliq_lwr, lut.radliq_upr, lut.radliq_fac
    c.diamice_lwr, c.diamice_upr, c.diamice_fac = lut.diamice_lwr, lut.diamice_upr, lut.diamice_fac
    c.lut_extliq = _d(lut.lut_extliq, keep); c.lut_ssaliq = _d(lut.lut_ssaliq, keep)
    c.lut_asyliq = _d(lut.lut_asyliq, keep); c.lut_extice = _d(lut.lut_extice, keep)
    c.lut_ssaice = _d(lut.lut_ssaice, keep); c.lut_asyice = _d(lut.lut_asyice, keep)
    return c, keep


def pack_aerosol(lut):
    keep: List = []
    a = AerosolFileC()
    a.n_phobic, a.n_band = lut.mass_ext_sw_hydrophobic.shape
    a.n_philic, a.n_hum, _ = lut.mass_ext_sw_hydrophilic.shape
    for k in ("mass_ext_sw_hydrophobic", "ssa_sw_hydrophobic", "asymmetry_sw_hydrophobic", "mass_ext_sw_hydrophilic",
              "ssa_sw_hydrophilic", "asymmetry_sw_hydrophilic", "relative_humidity2"):
        setattr(a, k, _d(getattr(lut, k), keep))
    return a, keep

"""netCDF (classic, 64-bit offset) files with the reference's variable names and dimension orders.

Writers for the three kinds of file the reference's executable reads and a reader for the one it
writes, so that the command-line driver (cli/test_rte_rrtmgp.cpp) can be exercised end to end:
  coefficients_{lw,sw}.nc        <- KDistFile      (schema: src_test/Radiation_solver.cpp:78-243)
  cloud_coefficients_{lw,sw}.nc  <- CloudLUTFile   (schema: src_test/Radiation_solver.cpp:286-326)
  aerosol_optics.nc              <- AerosolLUTFile (schema: src_test/Radiation_solver.cpp:331-367)
  rte_rrtmgp_input.nc            <- Atmosphere     (schema: src_test/test_rte_rrtmgp.cpp:205-330,455-478)
  rte_rrtmgp_output.nc           -> dict of numpy arrays (schema: :300-311,394-439,559-630)
scipy.io.netcdf_file writes CDF-1/CDF-2 only.  The readers (read_kdist, read_cloud_lut, read_aerosol_lut,
read_rfmip_input) take either format: netCDF classic through scipy, the netCDF-4/HDF5 files rrtmgp-data
ships through h5lite (no HDF5 library in this image).  The C++ driver reads netCDF classic only and needs
rrtmgp-data's files converted once with `nccopy -k cdf5`."""
from __future__ import annotations

from typing import Dict, Sequence

import numpy as np
from scipy.io import netcdf_file

from . import h5lite
from .atmos import Atmosphere
from .kdist import AerosolLUTFile, CloudLUTFile, KDistFile, STRING_LEN


def _dims(f, **sizes):
    for k, v in sizes.items():
        f.createDimension(k, int(v))


def _var(f, name: str, data, dims: Sequence[str], kind: str = "f8"):
    v = f.createVariable(name, kind, tuple(dims))
    a = np.asarray(data)
    if dims:
        v[:] = a.reshape([f.dimensions[d] for d in dims])
    else:
        v[...] = a.reshape(())
    return v


def _strings(f, name: str, names: Sequence[str], dim: str):
    v = f.createVariable(name, "c", (dim, "string_len"))
    block = np.array([list(s.encode("ascii")[:STRING_LEN].ljust(STRING_LEN, b" ")) for s in names], dtype=np.uint8)
    v[:] = block.view("S1").reshape(len(names), STRING_LEN)


def write_kdist(kd: KDistFile, path: str) -> None:
    """rrtmgp-gas-*.nc layout: variables in file dimension order, strings as blank-padded char blocks."""
    f = netcdf_file(path, "w", version=2)
    nT, nPi, nE, nG = kd.kmajor.shape
    _dims(f, temperature=nT, pressure=kd.press_ref.size, pressure_interp=nPi, absorber=len(kd.gas_names),
          minor_absorber=len(kd.gas_minor), absorber_ext=kd.vmr_ref.shape[1], mixing_fraction=nE,
          atmos_layer=kd.vmr_ref.shape[2], bnd=kd.nbnd, gpt=nG, pair=2,
          minor_absorber_intervals_lower=len(kd.minor_gases_lower), minor_absorber_intervals_upper=len(kd.minor_gases_upper),
          contributors_lower=kd.kminor_lower.shape[2], contributors_upper=kd.kminor_upper.shape[2], string_len=STRING_LEN)
    _strings(f, "gas_names", kd.gas_names, "absorber")
    _var(f, "key_species", kd.key_species, ("bnd", "atmos_layer", "pair"), "i4")
    _var(f, "bnd_limits_wavenumber", kd.bnd_limits_wavenumber, ("bnd", "pair"))
    _var(f, "bnd_limits_gpt", kd.bnd_limits_gpt, ("bnd", "pair"), "i4")
    _var(f, "press_ref", kd.press_ref, ("pressure",))
    _var(f, "temp_ref", kd.temp_ref, ("temperature",))
    _var(f, "absorption_coefficient_ref_P", kd.absorption_coefficient_ref_P, ())
    _var(f, "absorption_coefficient_ref_T", kd.absorption_coefficient_ref_T, ())
    _var(f, "press_ref_trop", kd.press_ref_trop, ())
    _var(f, "vmr_ref", kd.vmr_ref, ("temperature", "absorber_ext", "atmos_layer"))
    _var(f, "kmajor", kd.kmajor, ("temperature", "pressure_interp", "mixing_fraction", "gpt"))
    _var(f, "kminor_lower", kd.kminor_lower, ("temperature", "mixing_fraction", "contributors_lower"))
    _var(f, "kminor_upper", kd.kminor_upper, ("temperature", "mixing_fraction", "contributors_upper"))
    _strings(f, "gas_minor", kd.gas_minor, "minor_absorber")
    _strings(f, "identifier_minor", kd.identifier_minor, "minor_absorber")
    for side in ("lower", "upper"):
        d = "minor_absorber_intervals_" + side
        _strings(f, "minor_gases_" + side, getattr(kd, "minor_gases_" + side), d)
        _var(f, "minor_limits_gpt_" + side, getattr(kd, "minor_limits_gpt_" + side), (d, "pair"), "i4")
        _var(f, "minor_scales_with_density_" + side, getattr(kd, "minor_scales_with_density_" + side), (d,), "i4")
        _var(f, "scale_by_complement_" + side, getattr(kd, "scale_by_complement_" + side), (d,), "i4")
        _strings(f, "scaling_gas_" + side, getattr(kd, "scaling_gas_" + side), d)
        _var(f, "kminor_start_" + side, getattr(kd, "kminor_start_" + side), (d,), "i4")
    if kd.is_lw:
        _dims(f, temperature_Planck=kd.totplnk.shape[1])
        _var(f, "totplnk", kd.totplnk, ("bnd", "temperature_Planck"))
        _var(f, "plank_fraction", kd.plank_fraction, ("temperature", "pressure_interp", "mixing_fraction", "gpt"))
    else:
        _var(f, "rayl_lower", kd.rayl_lower, ("temperature", "mixing_fraction", "gpt"))
        _var(f, "rayl_upper", kd.rayl_upper, ("temperature", "mixing_fraction", "gpt"))
        for n in ("quiet", "facular", "sunspot"):
            _var(f, "solar_source_" + n, getattr(kd, "solar_source_" + n), ("gpt",))
        for n in ("tsi_default", "mg_default", "sb_default"):
            _var(f, n, getattr(kd, n), ())
    f.close()


def write_cloud_lut(lut: CloudLUTFile, path: str) -> None:
    f = netcdf_file(path, "w", version=2)
    _dims(f, nband=lut.nbnd, nrghice=lut.lut_extice.shape[0], nsize_liq=lut.lut_extliq.shape[1],
          nsize_ice=lut.lut_extice.shape[2], pair=2)
    _var(f, "bnd_limits_wavenumber", lut.bnd_limits_wavenumber, ("nband", "pair"))
    for n in ("radliq_lwr", "radliq_upr", "radliq_fac", "diamice_lwr", "diamice_upr", "diamice_fac"):
        _var(f, n, getattr(lut, n), ())
    for n in ("lut_extliq", "lut_ssaliq", "lut_asyliq"):
        _var(f, n, getattr(lut, n), ("nband", "nsize_liq"))
    for n in ("lut_extice", "lut_ssaice", "lut_asyice"):
        _var(f, n, getattr(lut, n), ("nrghice", "nband", "nsize_ice"))
    f.close()


def write_aerosol_lut(lut: AerosolLUTFile, path: str) -> None:
    f = netcdf_file(path, "w", version=2)
    nphob, nb = lut.mass_ext_sw_hydrophobic.shape
    nphil, nh, _ = lut.mass_ext_sw_hydrophilic.shape
    _dims(f, band_sw=nb, relative_humidity=nh, hydrophilic=nphil, hydrophobic=nphob)
    for n in ("mass_ext_sw_hydrophobic", "ssa_sw_hydrophobic", "asymmetry_sw_hydrophobic"):
        _var(f, n, getattr(lut, n), ("hydrophobic", "band_sw"))
    for n in ("mass_ext_sw_hydrophilic", "ssa_sw_hydrophilic", "asymmetry_sw_hydrophilic"):
        _var(f, n, getattr(lut, n), ("hydrophilic", "relative_humidity", "band_sw"))
    _var(f, "relative_humidity2", lut.relative_humidity2, ("relative_humidity",))
    f.close()


def write_input(atm: Atmosphere, path: str, nx: int, ny: int, kind: str = "f8", tsi_ref: float | None = None) -> None:
    """rte_rrtmgp_input.nc: fields as (lay|lev, y, x), surface fields as (y, x[, band]); x runs fastest
    over the columns.  With `tsi_ref` the per-column `tsi` variable is written (tsi_scaling * tsi_ref)."""
    assert nx * ny == atm.ncol
    f = netcdf_file(path, "w", version=2)
    nlay = atm.nlay
    _dims(f, x=nx, y=ny, lay=nlay, lev=nlay + 1, band_lw=atm.emis_sfc.shape[0], band_sw=atm.sfc_alb_dir.shape[0])

    def field(name, a, dim):      # (ncol, n) column-major-by-column -> (n, y, x)
        _var(f, name, np.asarray(a).T, (dim, "y", "x"), kind)

    field("p_lay", atm.play, "lay"); field("t_lay", atm.tlay, "lay")
    field("p_lev", atm.plev, "lev"); field("t_lev", atm.tlev, "lev")
    for gas, v in atm.vmr.items():
        v = np.atleast_2d(np.asarray(v))
        if v.shape == (1, 1):
            _var(f, "vmr_" + gas, v, (), kind)
        elif v.shape[0] == 1:
            _var(f, "vmr_" + gas, v[0], ("lay",), kind)
        else:
            field("vmr_" + gas, v, "lay")
    _var(f, "emis_sfc", atm.emis_sfc.T, ("y", "x", "band_lw"), kind)
    _var(f, "t_sfc", atm.tsfc, ("y", "x"), kind)
    _var(f, "mu0", atm.mu0, ("y", "x"), kind)
    _var(f, "sfc_alb_dir", atm.sfc_alb_dir.T, ("y", "x", "band_sw"), kind)
    _var(f, "sfc_alb_dif", atm.sfc_alb_dif.T, ("y", "x", "band_sw"), kind)
    if tsi_ref is not None:
        _var(f, "tsi", np.asarray(atm.tsi_scaling) * tsi_ref, ("y", "x"), kind)
    if atm.lwp is not None:
        for n in ("lwp", "iwp", "rel", "dei"):
            field(n, getattr(atm, n), "lay")
    if atm.rh is not None and atm.aermr is not None:      # src_test/test_rte_rrtmgp.cpp:274-293: (lay) or (lay, y, x)
        field("rh", atm.rh, "lay")
        for name, v in atm.aermr.items():
            v = np.atleast_2d(np.asarray(v))
            if v.shape[0] == 1:
                _var(f, name, np.broadcast_to(v[0], (nlay,)), ("lay",), kind)
            else:
                field(name, v, "lay")
    f.close()


GAS_NAMES = ("h2o", "co2", "o3", "n2o", "co", "ch4", "o2", "n2", "ccl4", "cfc11", "cfc12", "cfc22", "hfc143a", "hfc125",
             "hfc23", "hfc32", "hfc134a", "cf4", "no2")      # src_test/test_rte_rrtmgp.cpp:232-250


def read_input(path: str, tsi_ref: float | None = None):
    """rte_rrtmgp_input.nc (either format; written by the reference's rfmip_init.py / allsky_init.py / rcemip scripts or by
    write_input) -> (Atmosphere, nx, ny), the way test_rte_rrtmgp.cpp:205-330,455-478 reads it: fields (lay|lev, y, x) ->
    (ncol, n) with x fastest; vmr_<gas> scalar, (lay) or (lay, y, x), absent gases skipped; `tsi` (needs `tsi_ref`, the
    k-distribution's own total solar irradiance) or `tsi_scaling` -> tsi_scaling, else 1."""
    r = _Reader(path)
    dims3 = r.get("p_lay").shape
    if len(dims3) != 3:
        raise RuntimeError(f"{path}: p_lay must be (lay, y, x)")
    nlay, ny, nx = dims3
    ncol = nx * ny
    F = np.asfortranarray

    def field(name):
        a = r.f64(name)
        return F(a.reshape(a.shape[0], ncol).T)

    def surface(name):                           # (y, x, band) -> (band, ncol)
        a = r.f64(name)
        return F(a.reshape(ncol, -1).T)

    def concentration(name):
        a = r.f64(name)
        if a.ndim == 0 or a.size == 1 and a.ndim <= 1 and a.shape in ((), (1,)):
            return F(a.reshape(1, 1))
        if a.ndim == 1 and a.shape[0] == nlay:
            return F(a.reshape(1, nlay))
        if a.shape == (nlay, ny, nx):
            return F(a.reshape(nlay, ncol).T)
        raise RuntimeError(f"Illegal dimensions of \"{name}\" in input")

    vmr = {g: concentration("vmr_" + g) for g in GAS_NAMES if ("vmr_" + g) in r}
    if "tsi" in r:
        if tsi_ref is None:
            raise RuntimeError(f"{path} holds `tsi`: pass tsi_ref (the k-distribution's total solar irradiance)")
        tsi_scaling = r.f64("tsi").reshape(ncol) / tsi_ref
    elif "tsi_scaling" in r:
        tsi_scaling = r.f64("tsi_scaling").reshape(ncol)
    else:
        tsi_scaling = np.ones(ncol)
    cloud = {n: (field(n) if n in r else None) for n in ("lwp", "iwp", "rel", "dei")}
    aer = {n: concentration(n) for n in (f"aermr{i:02d}" for i in range(1, 12)) if n in r}
    atm = Atmosphere(play=field("p_lay"), plev=field("p_lev"), tlay=field("t_lay"), tlev=field("t_lev"),
                     tsfc=F(r.f64("t_sfc").reshape(ncol)), vmr=vmr, emis_sfc=surface("emis_sfc"),
                     sfc_alb_dir=surface("sfc_alb_dir"), sfc_alb_dif=surface("sfc_alb_dif"),
                     mu0=F(r.f64("mu0").reshape(ncol)), tsi_scaling=F(tsi_scaling),
                     rh=field("rh") if "rh" in r else None, aermr=aer or None, **cloud)
    r.close()
    return atm, nx, ny


def read_output(path: str) -> Dict[str, np.ndarray]:
    """rte_rrtmgp_output.nc -> arrays in the class-API layout: (lev|lay, y, x) -> (ncol, nlev|nlay),
    (band|gpt, lev|lay, y, x) -> (ncol, n, nband|ngpt), (gpt, y, x) -> (ncol, ngpt); band limits as stored."""
    f = netcdf_file(path, "r", mmap=False)
    out: Dict[str, np.ndarray] = {}
    for name, v in f.variables.items():
        a = np.array(v[...])
        d = v.dimensions
        if len(d) >= 2 and d[-2:] == ("y", "x"):
            a = a.reshape(a.shape[:-2] + (-1,))
            a = np.transpose(a, tuple(range(a.ndim))[::-1])
        out[name] = a
    f.close()
    return out


# ------------------------------------------------------------------------------------------
# Readers, either on-disk format
# ------------------------------------------------------------------------------------------
class _Reader:
    """get(name) -> array in file (C) dimension order; strings(name) -> stripped python strings."""

    def __init__(self, path: str):
        self.path = path
        with open(path, "rb") as f:
            magic = f.read(8)
        if magic[:3] == b"CDF":
            self.nc, self.h5 = netcdf_file(path, "r", mmap=False), None
        elif magic == b"\x89HDF\r\n\x1a\n":
            self.nc, self.h5 = None, h5lite.File(path)
        else:
            raise RuntimeError(f"{path}: neither netCDF classic nor netCDF-4/HDF5")

    def __contains__(self, name: str) -> bool:
        return name in (self.nc.variables if self.nc is not None else self.h5)

    def get(self, name: str) -> np.ndarray:
        if name not in self:
            raise RuntimeError(f"{self.path}: variable '{name}' is missing")      # Netcdf_interface.h throws likewise
        if self.nc is not None:
            v = self.nc.variables[name]
            return np.array(v.data if v.shape else v.getValue()).reshape(v.shape)
        return self.h5.get(name)

    def f64(self, name: str) -> np.ndarray:
        return np.ascontiguousarray(self.get(name), dtype=np.float64)

    def i32(self, name: str) -> np.ndarray:
        return np.ascontiguousarray(self.get(name), dtype=np.int32)

    def scalar(self, name: str) -> float:
        return float(np.asarray(self.get(name)).reshape(-1)[0])

    def strings(self, name: str):
        a = self.get(name)
        if a.dtype.kind == "S" and a.dtype.itemsize == 1:       # (n, string_len) char block
            a = a.reshape(a.shape[0], -1)
            rows = [b"".join(r.tolist()) for r in a]
        elif a.dtype.kind == "S":
            rows = a.reshape(-1).tolist()
        else:
            raise RuntimeError(f"{self.path}: '{name}' is not a character variable")
        return [r.decode("ascii", "replace").replace("\0", " ").strip() for r in rows]

    def attr(self, name: str, key: str):
        if self.nc is not None:
            v = getattr(self.nc.variables[name], key, None)
            return v.decode() if isinstance(v, bytes) else v
        return self.h5.attrs(name).get(key)

    def close(self) -> None:
        if self.nc is not None:
            self.nc.close()


def read_kdist(path: str) -> KDistFile:
    """A gas-optics coefficient file (rrtmgp-gas-{lw,sw}-g*.nc or coefficients_{lw,sw}.nc) -> KDistFile; the variable
    list is load_and_init_gas_optics's (src_test/Radiation_solver.cpp:78-243)."""
    r = _Reader(path)
    kw = dict(
        gas_names=r.strings("gas_names"), key_species=r.i32("key_species"),
        bnd_limits_wavenumber=r.f64("bnd_limits_wavenumber"), bnd_limits_gpt=r.i32("bnd_limits_gpt"),
        press_ref=r.f64("press_ref"), temp_ref=r.f64("temp_ref"),
        absorption_coefficient_ref_P=r.scalar("absorption_coefficient_ref_P"),
        absorption_coefficient_ref_T=r.scalar("absorption_coefficient_ref_T"), press_ref_trop=r.scalar("press_ref_trop"),
        vmr_ref=r.f64("vmr_ref"), kmajor=r.f64("kmajor"), kminor_lower=r.f64("kminor_lower"), kminor_upper=r.f64("kminor_upper"),
        gas_minor=r.strings("gas_minor"), identifier_minor=r.strings("identifier_minor"))
    for side in ("lower", "upper"):
        kw["minor_gases_" + side] = r.strings("minor_gases_" + side)
        kw["minor_limits_gpt_" + side] = r.i32("minor_limits_gpt_" + side)
        kw["minor_scales_with_density_" + side] = r.i32("minor_scales_with_density_" + side)
        kw["scale_by_complement_" + side] = r.i32("scale_by_complement_" + side)
        kw["scaling_gas_" + side] = r.strings("scaling_gas_" + side)
        kw["kminor_start_" + side] = r.i32("kminor_start_" + side)
    if "totplnk" in r:                                                          # :189 the reference's own test
        kw.update(totplnk=r.f64("totplnk"), plank_fraction=r.f64("plank_fraction"))
    else:
        kw.update(rayl_lower=r.f64("rayl_lower"), rayl_upper=r.f64("rayl_upper"),
                  solar_source_quiet=r.f64("solar_source_quiet"), solar_source_facular=r.f64("solar_source_facular"),
                  solar_source_sunspot=r.f64("solar_source_sunspot"), tsi_default=r.scalar("tsi_default"),
                  mg_default=r.scalar("mg_default"), sb_default=r.scalar("sb_default"))
    r.close()
    return KDistFile(label=path, **kw)


def read_cloud_lut(path: str) -> CloudLUTFile:
    """rrtmgp-clouds-{lw,sw}.nc / cloud_coefficients_{lw,sw}.nc (src_test/Radiation_solver.cpp:286-326)."""
    r = _Reader(path)
    lut = CloudLUTFile(
        bnd_limits_wavenumber=r.f64("bnd_limits_wavenumber"),
        **{n: r.scalar(n) for n in ("radliq_lwr", "radliq_upr", "radliq_fac", "diamice_lwr", "diamice_upr", "diamice_fac")},
        **{n: r.f64(n) for n in ("lut_extliq", "lut_ssaliq", "lut_asyliq", "lut_extice", "lut_ssaice", "lut_asyice")},
        label=path)
    r.close()
    return lut


def read_aerosol_lut(path: str) -> AerosolLUTFile:
    """aerosol_optics.nc (src_test/Radiation_solver.cpp:331-367).  NOT rrtmgp-data's rrtmgp-aerosols-merra-*.nc: that is the
    Fortran MERRA scheme's schema (aero_*_tbl), which the C++ reference does not read."""
    r = _Reader(path)
    lut = AerosolLUTFile(**{n: r.f64(n) for n in (
        "mass_ext_sw_hydrophobic", "ssa_sw_hydrophobic", "asymmetry_sw_hydrophobic",
        "mass_ext_sw_hydrophilic", "ssa_sw_hydrophilic", "asymmetry_sw_hydrophilic", "relative_humidity2")})
    r.close()
    return lut


def read_rfmip_input(path: str, expt: int = 0):
    """One experiment of the RFMIP clear-sky input (rrtmgp-data/examples/rfmip-clear-sky/inputs/multiple_input4MIPs_...nc)
    as the arrays rfmip/rfmip_init.py:14-112 builds from it (site-major (100, 60|61), top first as in the file): the
    pressure floor at the top (:34), mu0 = max(0, cos sza) (:52), tsi per site (:55), and the mole fractions of the
    gases RFMIP_GASES lists, each multiplied by the factor in its `units` attribute as the recipe does (:82-103)."""
    r = _Reader(path)

    def unit(var):
        return float(r.attr(var, "units"))

    out = dict(play=r.f64("pres_layer"), plev=r.f64("pres_level"), tlay=r.f64("temp_layer")[expt], tlev=r.f64("temp_level")[expt],
               tsfc=r.f64("surface_temperature")[expt], emis=r.f64("surface_emissivity"), albedo=r.f64("surface_albedo"),
               mu0=np.maximum(0.0, np.cos(np.deg2rad(r.f64("solar_zenith_angle")))), tsi=r.f64("total_solar_irradiance"),
               vmr={"h2o": r.f64("water_vapor")[expt] * unit("water_vapor"), "o3": r.f64("ozone")[expt] * unit("ozone")})
    for gas, var in RFMIP_GASES:
        out["vmr"][gas] = float(r.f64(var)[expt]) * unit(var)
    out["plev"] = np.maximum(out["plev"], np.nextafter(1.005183574463, 1e8))
    r.close()
    return out


# gas name in the k-distribution <- global-mean variable of the RFMIP file (rfmip/rfmip_init.py:82-103)
RFMIP_GASES = (("co2", "carbon_dioxide_GM"), ("n2o", "nitrous_oxide_GM"), ("co", "carbon_monoxide_GM"), ("ch4", "methane_GM"),
               ("o2", "oxygen_GM"), ("n2", "nitrogen_GM"), ("ccl4", "carbon_tetrachloride_GM"), ("cfc11", "cfc11_GM"),
               ("cfc12", "cfc12_GM"), ("cfc22", "hcfc22_GM"), ("hfc143a", "hfc143a_GM"), ("hfc125", "hfc125_GM"),
               ("hfc23", "hfc23_GM"), ("hfc32", "hfc32_GM"), ("hfc134a", "hfc134a_GM"), ("cf4", "cf4_GM"))

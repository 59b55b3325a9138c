"""k-distribution "file" container and a seeded synthetic generator.

`KDistFile` holds the variables of an rrtmgp-gas-{lw,sw}-g*.nc coefficient file under the
netCDF variable names and dimension orders the reference reads them with
(/root/reference/src_test/Radiation_solver.cpp:78-243).  Arrays are C-ordered numpy arrays in
FILE dimension order, e.g. kmajor(temperature, pressure_interp, mixing_fraction, gpt).

The real coefficient files are absent from the reference mount
(/root/reference/.MISSING_LARGE_BLOBS), so `synthetic_kdist` builds tables with the same
dimensions, dtypes, monotonic grids and descriptor structure (flavors, minor contributors
with density/complement scaling), with smooth positive coefficients and a true band-integrated
Planck table.  Runs that use it must be labelled "synthetic-tables".
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional

import numpy as np

STRING_LEN = 32

LW_BAND_LIMS = np.array(
    [[10, 250], [250, 500], [500, 630], [630, 700], [700, 820], [820, 980], [980, 1080],
     [1080, 1180], [1180, 1390], [1390, 1480], [1480, 1800], [1800, 2080], [2080, 2250],
     [2250, 2390], [2390, 2680], [2680, 3250]], dtype=np.float64)
SW_BAND_LIMS = np.array(
    [[820, 2680], [2680, 3250], [3250, 4000], [4000, 4650], [4650, 5150], [5150, 6150],
     [6150, 7700], [7700, 8050], [8050, 12850], [12850, 16000], [16000, 22650],
     [22650, 29000], [29000, 38000], [38000, 50000]], dtype=np.float64)

GAS_NAMES = ["h2o", "co2", "o3", "n2o", "co", "ch4", "o2", "n2", "ccl4", "cfc11", "cfc12",
             "cfc22", "hfc143a", "hfc125", "hfc23", "hfc32", "hfc134a", "cf4", "no2"]


@dataclass
class KDistFile:
    """Variables of a gas-optics coefficient file, file dimension order, C-contiguous."""
    gas_names: List[str]
    key_species: np.ndarray            # int32 (bnd, atmos_layer, pair)
    bnd_limits_wavenumber: np.ndarray  # f64 (bnd, 2)
    bnd_limits_gpt: np.ndarray         # int32 (bnd, 2), 1-based inclusive
    press_ref: np.ndarray              # f64 (pressure), descending
    temp_ref: np.ndarray               # f64 (temperature), ascending
    absorption_coefficient_ref_P: float
    absorption_coefficient_ref_T: float
    press_ref_trop: float
    vmr_ref: np.ndarray                # f64 (temperature, absorber_ext, atmos_layer)
    kmajor: np.ndarray                 # f64 (temperature, pressure_interp, mixing_fraction, gpt)
    kminor_lower: np.ndarray           # f64 (temperature, mixing_fraction, contributors_lower)
    kminor_upper: np.ndarray
    gas_minor: List[str]
    identifier_minor: List[str]
    minor_gases_lower: List[str]
    minor_gases_upper: List[str]
    minor_limits_gpt_lower: np.ndarray  # int32 (minor_absorber_intervals_lower, pair)
    minor_limits_gpt_upper: np.ndarray
    minor_scales_with_density_lower: np.ndarray  # int32 0/1
    minor_scales_with_density_upper: np.ndarray
    scale_by_complement_lower: np.ndarray
    scale_by_complement_upper: np.ndarray
    scaling_gas_lower: List[str]
    scaling_gas_upper: List[str]
    kminor_start_lower: np.ndarray     # int32, 1-based
    kminor_start_upper: np.ndarray
    # longwave only
    totplnk: Optional[np.ndarray] = None         # f64 (bnd, temperature_Planck)
    plank_fraction: Optional[np.ndarray] = None  # f64 like kmajor
    # shortwave only
    rayl_lower: Optional[np.ndarray] = None      # f64 (temperature, mixing_fraction, gpt)
    rayl_upper: Optional[np.ndarray] = None
    solar_source_quiet: Optional[np.ndarray] = None
    solar_source_facular: Optional[np.ndarray] = None
    solar_source_sunspot: Optional[np.ndarray] = None
    tsi_default: float = 0.0
    mg_default: float = 0.0
    sb_default: float = 0.0
    label: str = "synthetic-tables"

    @property
    def is_lw(self) -> bool:
        return self.totplnk is not None

    @property
    def ngpt(self) -> int:
        return int(self.kmajor.shape[3])

    @property
    def nbnd(self) -> int:
        return int(self.bnd_limits_gpt.shape[0])


def _planck_band_table(band_lims_cm: np.ndarray, temps: np.ndarray) -> np.ndarray:
    """Band-integrated Planck radiance B [W/m2/sr], shape (bnd, ntemp) (the solver applies pi)."""
    h, c, kb = 6.62607015e-34, 2.99792458e8, 1.380649e-23
    out = np.zeros((band_lims_cm.shape[0], temps.size))
    for ib, (w0, w1) in enumerate(band_lims_cm):
        nu = np.linspace(w0, w1, 400) * 100.0  # 1/m
        x = h * c * nu[None, :] / (kb * temps[:, None])
        b = 2.0 * h * c * c * nu[None, :] ** 3 / np.expm1(x)  # W/m2/sr/(1/m)
        out[ib, :] = np.trapezoid(b, nu, axis=1)
    return out


def _smooth_field(rng, shape, amp):
    """Smooth positive-ish random field: low-order separable cosine series."""
    f = np.zeros(shape)
    grids = np.meshgrid(*[np.linspace(0.0, 1.0, n) for n in shape], indexing="ij")
    for _ in range(4):
        term = np.ones(shape)
        for gax in grids:
            term = term * np.cos(np.pi * rng.uniform(0.2, 1.5) * gax + rng.uniform(0, 2 * np.pi))
        f += rng.normal(0.0, amp) * term
    return f


# key species (1-based indices into GAS_NAMES; 0 = none), (lower, upper) per band
_LW_KEY = [((1, 0), (1, 0)), ((1, 0), (1, 0)), ((1, 2), (1, 2)), ((1, 2), (2, 3)),
           ((1, 2), (2, 3)), ((1, 0), (0, 0)), ((1, 3), (3, 0)), ((1, 0), (3, 0)),
           ((1, 6), (6, 0)), ((1, 0), (1, 0)), ((1, 0), (1, 0)), ((1, 2), (0, 0)),
           ((1, 4), (3, 0)), ((2, 0), (2, 0)), ((4, 2), (0, 0)), ((1, 6), (6, 0))]
_SW_KEY = [((1, 6), (6, 0)), ((1, 2), (1, 2)), ((1, 6), (6, 0)), ((1, 2), (2, 0)),
           ((1, 0), (1, 0)), ((1, 2), (2, 0)), ((1, 0), (1, 0)), ((1, 7), (7, 0)),
           ((1, 0), (0, 0)), ((1, 7), (7, 0)), ((1, 0), (0, 0)), ((0, 0), (0, 0)),
           ((3, 0), (3, 0)), ((3, 7), (3, 7))]

_TYPICAL_VMR = {"h2o": 5e-3, "co2": 4e-4, "o3": 1e-6, "n2o": 3e-7, "co": 1e-7, "ch4": 1.7e-6,
                "o2": 0.209, "n2": 0.781, "ccl4": 1e-10, "cfc11": 2.5e-10, "cfc12": 5e-10,
                "cfc22": 2e-10, "hfc143a": 1e-11, "hfc125": 1e-11, "hfc23": 2e-11,
                "hfc32": 5e-12, "hfc134a": 8e-11, "cf4": 8e-11, "no2": 1e-10}

_W16 = np.array([0.1527534276, 0.1491729617, 0.1420961469, 0.1316886544, 0.1181945205,
                 0.1019300893, 0.0832767040, 0.0626720116, 0.0424925, 0.0046269894,
                 0.0038279891, 0.0030260086, 0.0022199750, 0.0014140010, 0.000533, 0.000075])


def _gweights(n: int) -> np.ndarray:
    if n == 16:
        w = _W16.copy()
    else:
        x = (np.arange(n) + 0.5) / n
        w = np.interp(x, (np.arange(16) + 0.5) / 16, _W16)
    return w / w.sum()


REDUCED_LW_128 = [10, 12, 10, 8, 8, 8, 6, 6, 8, 6, 8, 8, 6, 8, 8, 8]
REDUCED_SW_112 = [8] * 14


def synthetic_kdist(kind: str = "lw", gpt_per_band=16, seed: int = 20240607) -> KDistFile:
    """Seeded synthetic coefficient file with the structure of rrtmgp-gas-{lw,sw}-g*.nc."""
    assert kind in ("lw", "sw")
    rng = np.random.default_rng(seed + (0 if kind == "lw" else 1))
    band_lims = LW_BAND_LIMS if kind == "lw" else SW_BAND_LIMS
    keys = _LW_KEY if kind == "lw" else _SW_KEY
    nbnd = band_lims.shape[0]
    sizes = [gpt_per_band] * nbnd if np.isscalar(gpt_per_band) else list(gpt_per_band)
    assert len(sizes) == nbnd
    ngpt = int(sum(sizes))
    ntemp, npres, neta, nplanck = 14, 59, 9, 196
    ngas = len(GAS_NAMES)

    lims = np.zeros((nbnd, 2), dtype=np.int32)
    s = 1
    for ib, n in enumerate(sizes):
        lims[ib] = (s, s + n - 1)
        s += n

    temp_ref = 160.0 + 15.0 * np.arange(ntemp)
    press_ref = np.exp(np.log(109663.31) - 0.2 * np.arange(npres))
    press_ref_trop = 9948.431564193395

    key_species = np.zeros((nbnd, 2, 2), dtype=np.int32)
    for ib in range(nbnd):
        key_species[ib, 0, :] = keys[ib][0]
        key_species[ib, 1, :] = keys[ib][1]

    # vmr_ref(temperature, absorber_ext, atmos_layer): slot 0 is dry air (1.0)
    vmr_ref = np.ones((ntemp, ngas + 1, 2))
    for ig, name in enumerate(GAS_NAMES):
        base = _TYPICAL_VMR[name]
        tdep = np.exp(0.4 * (temp_ref - 250.0) / 100.0 * (1.0 if name == "h2o" else 0.1))
        vmr_ref[:, ig + 1, 0] = base * tdep
        vmr_ref[:, ig + 1, 1] = base * (0.02 if name == "h2o" else 1.0) * tdep ** 0.5

    col_typ = 1.7e23
    kmajor = np.zeros((ntemp, npres + 1, neta, ngpt))
    for ib in range(nbnd):
        g0, g1 = lims[ib, 0] - 1, lims[ib, 1]
        n = g1 - g0
        k1 = keys[ib][0][0] or 2
        mix_typ = 2.0 * _TYPICAL_VMR[GAS_NAMES[k1 - 1]] * col_typ
        base = np.log10(0.02 / mix_typ) + rng.uniform(-0.5, 0.5)
        gdep = 5.0 * (np.arange(n) / max(n - 1, 1)) ** 1.5 - 1.5
        fld = _smooth_field(rng, (ntemp, npres + 1, neta), 0.25)
        ptrend = np.linspace(0.3, -0.6, npres + 1)[None, :, None]
        for ig in range(n):
            wig = _smooth_field(rng, (ntemp, npres + 1, neta), 0.05)
            kmajor[:, :, :, g0 + ig] = 10.0 ** (base + gdep[ig] + fld + ptrend + wig)

    # ---- minor contributors ----
    identifier_minor = ["h2o_self", "h2o_frgn"] + GAS_NAMES[1:]
    gas_minor = ["h2o", "h2o"] + GAS_NAMES[1:]

    def make_minor(side: int):
        names, lim, dens, comp, scal, start, ks = [], [], [], [], [], [], []
        pos = 1
        for ib in range(nbnd):
            n = sizes[ib]
            ncontrib = int(rng.integers(2, 7)) if side == 0 else int(rng.integers(1, 5))
            pool = list(range(len(identifier_minor)))
            chosen = []
            if side == 0 and rng.uniform() < 0.8:
                chosen += [0, 1]
            while len(chosen) < ncontrib:
                c = int(rng.choice(pool[2:12] if rng.uniform() < 0.8 else pool[12:]))
                if c not in chosen:
                    chosen.append(c)
            for c in chosen:
                ident = identifier_minor[c]
                names.append(ident)
                lim.append((lims[ib, 0], lims[ib, 1]))
                if ident == "h2o_self":
                    d, cm, sg = 1, 0, "h2o"
                elif ident == "h2o_frgn":
                    d, cm, sg = 1, 1, "h2o"
                else:
                    r = rng.uniform()
                    if r < 0.4:
                        d, cm, sg = 0, 0, ""
                    elif r < 0.6:
                        d, cm, sg = 1, 0, ""
                    elif r < 0.8:
                        d, cm, sg = 1, 0, str(rng.choice(["n2", "o2", "h2o"]))
                    else:
                        d, cm, sg = 1, 1, str(rng.choice(["n2", "o2", "h2o", "cfc11"]))
                dens.append(d); comp.append(cm); scal.append(sg); start.append(pos)
                pos += n
                gname = gas_minor[c]
                sc_typ = _TYPICAL_VMR[gname] * col_typ * (3.0 if d else 1.0)
                if sg:
                    v = _TYPICAL_VMR[sg]
                    sc_typ *= (1.0 - v) if cm else v
                base = np.log10(0.003 / max(sc_typ, 1e-300)) + rng.uniform(-1.0, 0.5)
                gdep = 3.0 * (np.arange(n) / max(n - 1, 1)) - 1.0
                fld = _smooth_field(rng, (ntemp, neta), 0.2)
                ks.append(10.0 ** (base + gdep[None, None, :] + fld[:, :, None]))
        k = np.concatenate(ks, axis=2) if ks else np.zeros((ntemp, neta, 0))
        return (names, np.array(lim, dtype=np.int32).reshape(-1, 2), np.array(dens, dtype=np.int32),
                np.array(comp, dtype=np.int32), scal, np.array(start, dtype=np.int32),
                np.ascontiguousarray(k))

    mlo = make_minor(0)
    mup = make_minor(1)

    kd = KDistFile(
        gas_names=list(GAS_NAMES), key_species=key_species,
        bnd_limits_wavenumber=band_lims.copy(), bnd_limits_gpt=lims,
        press_ref=press_ref, temp_ref=temp_ref,
        absorption_coefficient_ref_P=101325.0, absorption_coefficient_ref_T=296.0,
        press_ref_trop=press_ref_trop, vmr_ref=vmr_ref, kmajor=np.ascontiguousarray(kmajor),
        kminor_lower=mlo[6], kminor_upper=mup[6], gas_minor=gas_minor,
        identifier_minor=identifier_minor, minor_gases_lower=mlo[0], minor_gases_upper=mup[0],
        minor_limits_gpt_lower=mlo[1], minor_limits_gpt_upper=mup[1],
        minor_scales_with_density_lower=mlo[2], minor_scales_with_density_upper=mup[2],
        scale_by_complement_lower=mlo[3], scale_by_complement_upper=mup[3],
        scaling_gas_lower=mlo[4], scaling_gas_upper=mup[4],
        kminor_start_lower=mlo[5], kminor_start_upper=mup[5],
        label=f"synthetic-tables-{kind}-g{ngpt}-seed{seed}")

    if kind == "lw":
        tpl = np.linspace(temp_ref[0], temp_ref[-1], nplanck)
        kd.totplnk = np.ascontiguousarray(_planck_band_table(band_lims, tpl))
        pf = np.zeros((ntemp, npres + 1, neta, ngpt))
        for ib in range(nbnd):
            g0, g1 = lims[ib, 0] - 1, lims[ib, 1]
            n = g1 - g0
            w = _gweights(n)
            raw = np.stack([w[ig] * np.exp(_smooth_field(rng, (ntemp, npres + 1, neta), 0.15))
                            for ig in range(n)], axis=3)
            pf[:, :, :, g0:g1] = raw / raw.sum(axis=3, keepdims=True)
        kd.plank_fraction = np.ascontiguousarray(pf)
    else:
        rl = np.zeros((ntemp, neta, ngpt)); ru = np.zeros((ntemp, neta, ngpt))
        sol = np.zeros(ngpt)
        sun = _planck_band_table(band_lims, np.array([5772.0]))[:, 0]
        sun = sun / sun.sum() * 1360.85
        for ib in range(nbnd):
            g0, g1 = lims[ib, 0] - 1, lims[ib, 1]
            n = g1 - g0
            numid = 0.5 * (band_lims[ib, 0] + band_lims[ib, 1])
            kr = 4.4e-27 * (numid / 18000.0) ** 4
            f1 = 1.0 + _smooth_field(rng, (ntemp, neta), 0.02)
            f2 = 1.0 + _smooth_field(rng, (ntemp, neta), 0.02)
            gv = 1.0 + 0.1 * np.linspace(-1, 1, n)
            rl[:, :, g0:g1] = kr * f1[:, :, None] * gv[None, None, :]
            ru[:, :, g0:g1] = kr * f2[:, :, None] * gv[None, None, :]
            sol[g0:g1] = sun[ib] * _gweights(n)
        kd.rayl_lower, kd.rayl_upper = rl, ru
        kd.solar_source_quiet = sol
        kd.solar_source_facular = sol * 0.02 * rng.uniform(0.5, 1.5, ngpt)
        kd.solar_source_sunspot = -sol * 1e-7 * rng.uniform(0.5, 1.5, ngpt)
        kd.tsi_default, kd.mg_default, kd.sb_default = 1360.85, 0.1567652, 909.21
    return kd


@dataclass
class CloudLUTFile:
    """Variables of rrtmgp-clouds-{lw,sw}.nc in file dimension order
    (/root/reference/src_test/Radiation_solver.cpp:286-326)."""
    bnd_limits_wavenumber: np.ndarray  # (nband, 2)
    radliq_lwr: float
    radliq_upr: float
    radliq_fac: float
    diamice_lwr: float
    diamice_upr: float
    diamice_fac: float
    lut_extliq: np.ndarray  # (nband, nsize_liq)
    lut_ssaliq: np.ndarray
    lut_asyliq: np.ndarray
    lut_extice: np.ndarray  # (nrghice, nband, nsize_ice)
    lut_ssaice: np.ndarray
    lut_asyice: np.ndarray
    label: str = "synthetic-cloud-lut"

    @property
    def nbnd(self) -> int:
        return int(self.bnd_limits_wavenumber.shape[0])


def synthetic_cloud_lut(kind: str = "lw", seed: int = 20240607) -> CloudLUTFile:
    """Cloud optics LUT with the dimensions of rrtmgp-clouds-{lw,sw}.nc and Mie-like trends."""
    rng = np.random.default_rng(seed + 17 + (0 if kind == "lw" else 1))
    band_lims = LW_BAND_LIMS if kind == "lw" else SW_BAND_LIMS
    nb, nliq, nice, nrgh = band_lims.shape[0], 20, 18, 3
    rliq = np.linspace(2.5, 21.5, nliq)
    dice = np.linspace(10.0, 180.0, nice)
    bvar = 1.0 + 0.15 * rng.uniform(-1, 1, nb)
    extliq = 1.5 / rliq[None, :] * bvar[:, None]              # m2/g ~ 3/(2 rho r)
    extice = 3.3 / dice[None, None, :] * bvar[None, :, None] * (1.0 + 0.1 * np.arange(nrgh)[:, None, None])
    if kind == "lw":
        ssaliq = np.clip(0.3 + 0.02 * rliq[None, :] * bvar[:, None], 0.05, 0.8)
        ssaice = np.clip(0.4 + 0.001 * dice[None, None, :] * bvar[None, :, None], 0.05, 0.8) * np.ones((nrgh, 1, 1))
    else:
        absb = np.linspace(0.25, 1e-6, nb) ** 2
        ssaliq = 1.0 - absb[:, None] * (rliq[None, :] / 10.0)
        ssaice = (1.0 - absb[None, :, None] * (dice[None, None, :] / 60.0)) * np.ones((nrgh, 1, 1))
        ssaliq, ssaice = np.clip(ssaliq, 0.3, 0.999999), np.clip(ssaice, 0.3, 0.999999)
    asyliq = np.clip(0.78 + 0.004 * rliq[None, :] + 0.03 * (bvar[:, None] - 1), 0.5, 0.95) * np.ones((nb, 1))
    asyice = np.clip(0.74 + 0.0008 * dice[None, None, :] - 0.02 * np.arange(nrgh)[:, None, None]
                     + 0.03 * (bvar[None, :, None] - 1), 0.5, 0.95)
    c = np.ascontiguousarray
    return CloudLUTFile(band_lims.copy(), 2.5, 21.5, 1.0, 10.0, 180.0, 1.0,
                        c(extliq), c(ssaliq), c(asyliq), c(extice), c(ssaice), c(asyice),
                        label=f"synthetic-cloud-lut-{kind}")


@dataclass
class AerosolLUTFile:
    """Variables of aerosol_optics.nc in file dimension order
    (/root/reference/src_test/Radiation_solver.cpp:331-367)."""
    mass_ext_sw_hydrophobic: np.ndarray     # (hydrophobic, band_sw)
    ssa_sw_hydrophobic: np.ndarray
    asymmetry_sw_hydrophobic: np.ndarray
    mass_ext_sw_hydrophilic: np.ndarray     # (hydrophilic, relative_humidity, band_sw)
    ssa_sw_hydrophilic: np.ndarray
    asymmetry_sw_hydrophilic: np.ndarray
    relative_humidity2: np.ndarray          # (relative_humidity) upper bound of each class
    label: str = "synthetic-aerosol-lut"

    @property
    def nbnd(self) -> int:
        return int(self.mass_ext_sw_hydrophobic.shape[1])


def synthetic_aerosol_lut(seed: int = 20240607, nband: int = 14) -> AerosolLUTFile:
    """Aerosol LUT with the dimensions of the ECMWF/CAMS-style aerosol_optics.nc the reference reads
    (14 shortwave bands, 12 humidity classes, 13 hydrophilic / 14 hydrophobic types): mass extinction in
    m2/kg decreasing with wavelength, hydrophilic extinction growing with humidity, black carbon absorbing."""
    rng = np.random.default_rng(seed + 29)
    nhum, nphil, nphob = 12, 13, 14
    rh_upper = np.array([0.1, 0.2, 0.3, 0.4, 0.5, 0.6, 0.7, 0.8, 0.85, 0.9, 0.95, 1.0])
    wl = np.linspace(0.3, 1.0, nband)                              # relative spectral weight, band 1 = near-IR
    mext_phob = 200.0 + 3000.0 * rng.uniform(0.2, 1.0, (nphob, 1)) * wl[None, :]
    ssa_phob = np.clip(0.95 - 0.25 * rng.uniform(0, 1, (nphob, 1)) * (1.0 - wl[None, :]), 0.2, 0.999)
    ssa_phob[10] = np.clip(0.15 + 0.2 * wl, 0.1, 0.5)              # black carbon (column 11)
    asy_phob = np.clip(0.55 + 0.25 * rng.uniform(0, 1, (nphob, 1)) * wl[None, :], 0.1, 0.9)
    growth = 1.0 + 4.0 * rh_upper ** 3
    mext_phil = (300.0 + 4000.0 * rng.uniform(0.2, 1.0, (nphil, 1, 1)) * wl[None, None, :]) * growth[None, :, None]
    ssa_phil = np.clip(0.90 + 0.099 * rh_upper[None, :, None] * np.ones((nphil, 1, nband))
                       - 0.1 * rng.uniform(0, 1, (nphil, 1, 1)) * (1.0 - wl[None, None, :]), 0.5, 0.99999)
    asy_phil = np.clip(0.6 + 0.18 * rh_upper[None, :, None] + 0.05 * rng.uniform(-1, 1, (nphil, 1, nband)), 0.1, 0.9)
    c = np.ascontiguousarray
    return AerosolLUTFile(c(mext_phob), c(ssa_phob), c(asy_phob), c(mext_phil), c(ssa_phil), c(asy_phil), rh_upper)

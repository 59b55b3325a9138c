"""Seeded synthetic atmospheric columns of the shapes BASELINE.json names.

Profiles follow the reference's own analytic test-case generators (recipes only, re-derived):
  * RCEMIP-style:  /root/reference/rcemip/test_rcemip_input.py:20-75
  * all-sky:       /root/reference/allsky/allsky_init.py:14-90,157-176
with the per-column perturbations SURVEY.md §8(d) specifies (S-RCE / S-ALLSKY).  All arrays are
fp64, Fortran-ordered with the column index fastest: (ncol, nlay) / (ncol, nlay+1) /
(nbnd, ncol), exactly as the reference class API takes them.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, Optional

import numpy as np

SEED = 20240607


@dataclass
class Atmosphere:
    play: np.ndarray   # (ncol, nlay) Pa
    plev: np.ndarray   # (ncol, nlay+1)
    tlay: np.ndarray
    tlev: np.ndarray
    tsfc: np.ndarray   # (ncol,)
    vmr: Dict[str, np.ndarray]  # each (1,1), (1,nlay) or (ncol,nlay)
    emis_sfc: np.ndarray        # (nbnd_lw, ncol)
    sfc_alb_dir: np.ndarray     # (nbnd_sw, ncol)
    sfc_alb_dif: np.ndarray
    mu0: np.ndarray             # (ncol,)
    tsi_scaling: np.ndarray     # (ncol,)
    lwp: Optional[np.ndarray] = None  # (ncol, nlay) g/m2
    iwp: Optional[np.ndarray] = None
    rel: Optional[np.ndarray] = None  # micron
    dei: Optional[np.ndarray] = None
    rh: Optional[np.ndarray] = None   # (ncol, nlay) relative humidity 0..1 (aerosol optics)
    aermr: Optional[Dict[str, np.ndarray]] = None   # "aermr01".."aermr11": (1,1), (1,nlay) or (ncol,nlay) kg/kg

    @property
    def ncol(self) -> int:
        return self.play.shape[0]

    @property
    def nlay(self) -> int:
        return self.play.shape[1]

    @property
    def top_at_1(self) -> bool:
        return bool(self.play[0, 0] < self.play[0, -1])

    def subset(self, c0: int, c1: int) -> "Atmosphere":
        """Columns [c0, c1) — the column shard of one rank / one block."""
        def cut(a, axis=0):
            if a is None:
                return None
            sl = [slice(None)] * a.ndim
            sl[axis] = slice(c0, c1)
            return np.asfortranarray(a[tuple(sl)])
        vmr = {k: (cut(v) if v.shape[0] > 1 else v) for k, v in self.vmr.items()}
        aer = None if self.aermr is None else {k: (cut(v) if v.shape[0] > 1 else v) for k, v in self.aermr.items()}
        return Atmosphere(cut(self.play), cut(self.plev), cut(self.tlay), cut(self.tlev),
                          cut(self.tsfc), vmr, cut(self.emis_sfc, 1), cut(self.sfc_alb_dir, 1),
                          cut(self.sfc_alb_dif, 1), cut(self.mu0), cut(self.tsi_scaling),
                          cut(self.lwp), cut(self.iwp), cut(self.rel), cut(self.dei), cut(self.rh), aer)

    def flipped(self) -> "Atmosphere":
        """Same columns with the vertical axis reversed (top_at_1 toggled)."""
        def fl(a):
            return None if a is None else np.asfortranarray(a[:, ::-1])
        vmr = {k: (np.asfortranarray(v[:, ::-1]) if v.shape[1] > 1 else v) for k, v in self.vmr.items()}
        aer = None if self.aermr is None else {k: (np.asfortranarray(v[:, ::-1]) if v.shape[1] > 1 else v)
                                               for k, v in self.aermr.items()}
        return Atmosphere(fl(self.play), fl(self.plev), fl(self.tlay), fl(self.tlev), self.tsfc,
                          vmr, self.emis_sfc, self.sfc_alb_dir, self.sfc_alb_dif, self.mu0,
                          self.tsi_scaling, fl(self.lwp), fl(self.iwp), fl(self.rel), fl(self.dei), fl(self.rh), aer)


def _rce_profile(z):
    q_0, z_q1, z_q2, z_t = 0.01864, 4.0e3, 7.5e3, 15.0e3
    q = q_0 * np.exp(-z / z_q1) * np.exp(-(z / z_q2) ** 2)
    q_t = q_0 * np.exp(-z_t / z_q1) * np.exp(-(z_t / z_q2) ** 2)
    above = z > z_t
    q = np.where(above, q_t, q)
    T_0, gamma = 300.0, 6.7e-3
    Tv_0 = (1.0 + 0.608 * q_0) * T_0
    Tv_t = Tv_0 - gamma * z_t
    Tv = np.where(above, Tv_t, Tv_0 - gamma * z)
    T = Tv / (1.0 + 0.608 * q)
    g, Rd, p0 = 9.79764, 287.04, 101480.0
    p = p0 * (Tv / Tv_0) ** (g / (Rd * gamma))
    p_tmp = p0 * (Tv_t / Tv_0) ** (g / (Rd * gamma)) * np.exp(-((g * (z - z_t)) / (Rd * Tv_t)))
    p = np.where(above, p_tmp, p)
    return p, q, T


def _ozone(p_lay):
    g1, g2, g3 = 3.6478, 0.83209, 11.3515
    p_hpa = p_lay / 100.0
    return np.maximum(1e-13, g1 * p_hpa ** g2 * np.exp(-p_hpa / g3) * 1e-6)


def _F(a):
    return np.asfortranarray(a, dtype=np.float64)


def rce_atmosphere(ncol: int, nlay: int = 128, seed: int = SEED, nbnd_lw: int = 16,
                   nbnd_sw: int = 14, p_min: float = 1.1, p_max: float = 109000.0,
                   top_at_1: bool = False, night_fraction: float = 0.1,
                   cloudy: bool = False, z_top: float = 70.0e3) -> Atmosphere:
    """S-RCE columns: RCEMIP analytic profile + per-column perturbations."""
    rng = np.random.default_rng(seed)
    dz = z_top / nlay
    z = np.arange(nlay) * dz + 0.5 * dz
    zh = np.arange(nlay + 1) * dz
    p_lay, q, T_lay = _rce_profile(z)
    p_lev, _, T_lev = _rce_profile(zh)

    # smooth-in-z temperature perturbation, N(0, 2 K) amplitude
    amp = rng.normal(0.0, 2.0, size=(ncol, 3))
    ph = rng.uniform(0, 2 * np.pi, size=(ncol, 3))
    def tpert(zz):
        s = np.zeros((ncol, zz.size))
        for k in range(3):
            s += amp[:, k:k + 1] / np.sqrt(3.0) * np.sin((k + 1) * np.pi * zz[None, :] / z_top + ph[:, k:k + 1])
        return s
    tlay = np.clip(T_lay[None, :] + tpert(z), 161.0, 354.0)
    tlev = np.clip(T_lev[None, :] + tpert(zh), 161.0, 354.0)
    qfac = np.exp(rng.normal(0.0, 0.3, size=(ncol, 1)))
    Rd_Rv = 287.04 / 461.5
    h2o = np.clip(q[None, :] * qfac, 0.0, 0.04)
    h2o = h2o / (Rd_Rv * (1.0 - h2o))
    play = np.clip(np.tile(p_lay[None, :], (ncol, 1)), p_min, p_max)
    plev = np.clip(np.tile(p_lev[None, :], (ncol, 1)), p_min, p_max)
    o3 = np.tile(_ozone(p_lay)[None, :], (ncol, 1))

    vmr = {"h2o": _F(h2o), "o3": _F(o3),
           "co2": _F([[348e-6]]), "ch4": _F([[1650e-9]]), "n2o": _F([[306e-9]]),
           "n2": _F([[0.7808]]), "o2": _F([[0.2095]])}
    tsfc = np.clip(tlev[:, 0] + rng.uniform(0.0, 3.0, ncol), 161.0, 354.0)
    emis = rng.uniform(0.95, 1.0, size=(nbnd_lw, ncol))
    alb = rng.uniform(0.05, 0.3, size=(nbnd_sw, ncol))
    mu0 = rng.uniform(0.05, 1.0, ncol)
    mu0[rng.uniform(size=ncol) < night_fraction] = 0.0
    atm = Atmosphere(_F(play), _F(plev), _F(tlay), _F(tlev), _F(tsfc), vmr, _F(emis), _F(alb),
                     _F(alb.copy()), _F(mu0), _F(np.ones(ncol)))
    if cloudy:
        _add_clouds(atm, rng)
    return atm.flipped() if top_at_1 else atm


def add_aerosols(atm: Atmosphere, seed: int = SEED) -> Atmosphere:
    """Relative humidity and the 11 aerosol mixing ratios (aermr01..11) for the columns of `atm`:
    boundary-layer-weighted profiles of CAMS-like magnitude (1e-11..1e-8 kg/kg).  aermr02 is a (1,nlay)
    profile and aermr10 a scalar so the broadcast of fill_aerosols_3d (src/Aerosol_optics.cpp:152-165) is
    exercised; the rest are (ncol,nlay) fields."""
    rng = np.random.default_rng(seed + 41)
    ncol, nlay = atm.play.shape
    sig = atm.play / np.max(atm.plev)                               # ~1 at the surface
    shape = np.exp(-(1.0 - sig) * 6.0)
    atm.rh = _F(np.clip(0.85 * sig ** 0.5 * rng.uniform(0.3, 1.15, (ncol, nlay)), 0.0, 0.99))
    aer = {}
    for i in range(1, 12):
        amp = 10.0 ** rng.uniform(-10.5, -8.0)
        f = amp * shape * rng.uniform(0.5, 1.5, (ncol, nlay))
        if i == 2:
            f = f[:1, :]
        elif i == 10:
            f = np.array([[amp * 0.3]])
        aer[f"aermr{i:02d}"] = _F(f)
    atm.aermr = aer
    return atm


def _add_clouds(atm: Atmosphere, rng) -> None:
    ncol, nlay = atm.play.shape
    flag = rng.uniform(size=(ncol, 1)) < (2.0 / 3.0)
    mask = (atm.play > 1.0e4) & (atm.play < 9.0e4) & flag
    lwp = np.where(mask & (atm.tlay > 263.0), rng.uniform(1.0, 50.0, (ncol, nlay)), 0.0)
    iwp = np.where(mask & (atm.tlay < 273.0), rng.uniform(1.0, 50.0, (ncol, nlay)), 0.0)
    atm.lwp, atm.iwp = _F(lwp), _F(iwp)
    atm.rel = _F(np.where(lwp > 0, rng.uniform(2.5, 21.5, (ncol, nlay)), 0.0))
    atm.dei = _F(np.where(iwp > 0, rng.uniform(10.0, 180.0, (ncol, nlay)), 0.0))


def allsky_atmosphere(ncol: int = 24, nlay: int = 72, seed: int = SEED, nbnd_lw: int = 16,
                      nbnd_sw: int = 14, random_clouds: bool = False, p_min: float = 1.1) -> Atmosphere:
    """All-sky test columns (allsky_init.py): identical columns, 2 of 3 cloudy; or S-ALLSKY."""
    rng = np.random.default_rng(seed)
    z_top, z_trop = 70.0e3, 15.0e3
    zh = np.zeros(nlay + 1)
    for i in range(1, nlay // 2 + 1):
        zh[i] = 2.0 * i * z_trop / nlay
        zh[i + nlay // 2] = z_trop + 2.0 * i * (z_top - z_trop) / nlay
    z = 0.5 * (zh[1:] + zh[:-1])

    def prof(zz):
        q_0, z_q1, z_q2 = 0.01864, 4.0e3, 7.5e3
        q = q_0 * np.exp(-zz / z_q1) * np.exp(-(zz / z_q2) ** 2)
        above = zz > z_trop
        q = np.where(above, 1.0e-8, q)
        T_0, gamma = 300.0, 6.7e-3
        Tv_0 = (1.0 + 0.608 * q_0) * T_0
        T = T_0 - gamma * zz / (1.0 + 0.608 * q)
        T = np.where(above, T_0 - gamma * z_trop / (1.0 + 0.608 * q_0), T)
        Tv = T * (1.0 + 0.608 * q)
        g, Rd, p0 = 9.79764, 287.04, 101480.0
        p = p0 * (Tv / Tv_0) ** (g / (Rd * gamma))
        p_tmp = p * np.exp(-((g * (zz - z_trop)) / (Rd * Tv)))
        return np.where(above, p_tmp, p), q, T

    p_lay, q, T_lay = prof(z)
    p_lev, _, T_lev = prof(zh)
    p_lay = np.maximum(p_lay, p_min); p_lev = np.maximum(p_lev, p_min)
    tile = lambda a: _F(np.tile(a[None, :], (ncol, 1)))
    play, plev, tlay, tlev = tile(p_lay), tile(p_lev), tile(T_lay), tile(T_lev)
    vmr = {"h2o": tile(q), "o3": tile(_ozone(p_lay)),
           "co2": _F([[348e-6]]), "ch4": _F([[1650e-9]]), "n2o": _F([[306e-9]]),
           "n2": _F([[0.7808]]), "o2": _F([[0.2095]])}
    atm = Atmosphere(play, plev, tlay, tlev, _F(np.full(ncol, 300.0)), vmr,
                     _F(np.full((nbnd_lw, ncol), 0.98)), _F(np.full((nbnd_sw, ncol), 0.06)),
                     _F(np.full((nbnd_sw, ncol), 0.06)), _F(np.full(ncol, 0.86)), _F(np.ones(ncol)))
    if random_clouds:
        _add_clouds(atm, rng)
    else:
        flag = (np.arange(1, ncol + 1) % 3 > 0)[:, None]
        mask = (play > 1.0e4) & (play < 9.0e4) & flag
        atm.lwp = _F(np.where(mask & (tlay > 263.0), 10.0, 0.0))
        atm.iwp = _F(np.where(mask & (tlay < 273.0), 10.0, 0.0))
        atm.rel = _F(np.where(atm.lwp > 0, 0.5 * (2.5 + 21.5), 0.0))
        atm.dei = _F(np.where(atm.iwp > 0, 0.5 * (10.0 + 180.0), 0.0))
    return atm

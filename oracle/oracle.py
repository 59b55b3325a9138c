"""ORACLE — TEST INFRASTRUCTURE ONLY (never imported by the product package).

CPU restatement of the reference's column radiation path:
  * the numeric kernels are oracle/rrtmgp_kernels_restated.c (C port of the Fortran bind(C)
    kernels), loaded through ctypes;
  * this module restates, in numpy, the thin C++ class layer above them
    (/root/reference/src/{Gas_optics_rrtmgp,Optical_props,Rte_lw,Rte_sw,Fluxes,Cloud_optics}.cpp),
    each function citing the lines it follows.
The restatement of the class layer is itself validated against the reference's unmodified
class layer in tests/test_oracle_vs_ref.py whenever oracle/_ref/liboracle_ref.so was built.

Arrays: numpy fp64, Fortran order, column index fastest, 1-based integer tables as in Fortran.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Dict, List, Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "liboracle.so")
_lib = None

c_int_p = C.POINTER(C.c_int)
c_dbl_p = C.POINTER(C.c_double)
c_bool_p = C.POINTER(C.c_byte)


def build(force: bool = False) -> str:
    """Compile the C restatement (gcc only)."""
    src = os.path.join(_HERE, "rrtmgp_kernels_restated.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        os.makedirs(os.path.dirname(_LIB_PATH), exist_ok=True)
        subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared",
                               "-o", _LIB_PATH, src, "-lm"])
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
    return _lib


def _i(v):
    return C.byref(C.c_int(int(v)))


def _d(v):
    return C.byref(C.c_double(float(v)))


def _b(v):
    return C.byref(C.c_byte(1 if v else 0))


def _p(a: np.ndarray):
    assert a.flags.f_contiguous or a.ndim <= 1, "oracle arrays must be Fortran-ordered"
    if a.dtype == np.float64:
        return a.ctypes.data_as(c_dbl_p)
    if a.dtype == np.int32:
        return a.ctypes.data_as(c_int_p)
    if a.dtype == np.int8:
        return a.ctypes.data_as(c_bool_p)
    raise TypeError(a.dtype)


def F(a, dtype=np.float64):
    return np.asfortranarray(a, dtype=dtype)


def zeros(shape, dtype=np.float64):
    return np.zeros(shape, dtype=dtype, order="F")


# ------------------------------------------------------------------------------------------
# Optical_props band/g-point bookkeeping: src/Optical_props.cpp:31-72
# ------------------------------------------------------------------------------------------
class SpectralDisc:
    def __init__(self, band_lims_wvn: np.ndarray, band_lims_gpt: Optional[np.ndarray] = None):
        self.band_lims_wvn = F(band_lims_wvn)              # (2, nbnd)
        nb = self.band_lims_wvn.shape[1]
        if band_lims_gpt is None:                          # by-band: one "g-point" per band
            band_lims_gpt = np.stack([np.arange(1, nb + 1), np.arange(1, nb + 1)])
        self.band2gpt = F(band_lims_gpt, np.int32)         # (2, nbnd), 1-based
        self.ngpt = int(self.band2gpt.max())
        self.nbnd = nb
        self.gpt2band = np.zeros(self.ngpt, dtype=np.int32)
        for ib in range(nb):
            self.gpt2band[self.band2gpt[0, ib] - 1:self.band2gpt[1, ib]] = ib + 1


# ------------------------------------------------------------------------------------------
# Gas_optics_rrtmgp: src/Gas_optics_rrtmgp.cpp
# ------------------------------------------------------------------------------------------
class GasOptics(SpectralDisc):
    def __init__(self, kd, available_gases: Sequence[str]):
        """Constructor + init_abs_coeffs (src/Gas_optics_rrtmgp.cpp:391-742)."""
        super().__init__(kd.bnd_limits_wavenumber.T, kd.bnd_limits_gpt.T)
        avail = set(available_gases)
        file_names = [s.strip() for s in kd.gas_names]
        # :574-586 gases known to the k-distribution AND present in the host model
        self.gas_names = [s for s in file_names if s in avail]
        ngas = len(self.gas_names)
        self.ngas = ngas

        # :589-605 vmr_ref(2, 0:ngas, ntemp); file order (temperature, absorber_ext, atmos_layer)
        vfile = np.transpose(kd.vmr_ref, (2, 1, 0))          # (atmos_layer, absorber_ext, temp)
        vmr_ref = np.zeros((2, ngas + 1, vfile.shape[2]))
        vmr_ref[:, 0, :] = vfile[:, 0, :]
        for i, name in enumerate(self.gas_names):
            vmr_ref[:, i + 1, :] = vfile[:, file_names.index(name) + 1, :]
        self.vmr_ref = F(vmr_ref)

        # :54-162 reduce_minor_arrays for lower / upper
        def reduce(kminor, minor_gases, limits, dens, scal_gas, comp, kstart):
            ident = [s.strip() for s in kd.identifier_minor]
            gminor = [s.strip() for s in kd.gas_minor]
            keep = [gminor[ident.index(m.strip())] in avail for m in minor_gases]
            kfile = np.transpose(kminor, (2, 1, 0))          # (contributors, eta, temp)
            lim_r, dens_r, scal_r, comp_r, start_r, names_r, blocks = [], [], [], [], [], [], []
            n_elim = 0
            for i, m in enumerate(minor_gases):
                ng = int(limits[i, 1] - limits[i, 0] + 1)
                if keep[i]:
                    names_r.append(m.strip()); lim_r.append(limits[i]); dens_r.append(dens[i])
                    scal_r.append(scal_gas[i].strip()); comp_r.append(comp[i])
                    start_r.append(int(kstart[i]) - n_elim)
                    blocks.append(kfile[kstart[i] - 1:kstart[i] - 1 + ng])
                else:
                    n_elim += ng
            kred = np.concatenate(blocks, axis=0) if blocks else np.zeros((0,) + kfile.shape[1:])
            # :156-161 reshape to (ntemp, neta, nk)
            kred = F(np.transpose(kred, (2, 1, 0)))
            lim = F(np.array(lim_r, dtype=np.int32).reshape(-1, 2).T, np.int32)  # (2, nminor)
            # :164-196 create_idx_minor / create_idx_minor_scaling (find_index returns -1 if absent)
            def find(name):
                return self.gas_names.index(name) + 1 if name in self.gas_names else -1
            idx_minor = np.array([find(gminor[ident.index(n)]) for n in names_r], dtype=np.int32)
            idx_scal = np.array([find(s) for s in scal_r], dtype=np.int32)
            return dict(kminor=kred, limits=lim, dens=np.array(dens_r, dtype=np.int8),
                        comp=np.array(comp_r, dtype=np.int8), kstart=np.array(start_r, dtype=np.int32),
                        idx_minor=idx_minor, idx_scal=idx_scal)

        self.lower = reduce(kd.kminor_lower, kd.minor_gases_lower, kd.minor_limits_gpt_lower,
                            kd.minor_scales_with_density_lower, kd.scaling_gas_lower,
                            kd.scale_by_complement_lower, kd.kminor_start_lower)
        self.upper = reduce(kd.kminor_upper, kd.minor_gases_upper, kd.minor_limits_gpt_upper,
                            kd.minor_scales_with_density_upper, kd.scaling_gas_upper,
                            kd.scale_by_complement_upper, kd.kminor_start_upper)

        self.press_ref = F(kd.press_ref)
        self.temp_ref = F(kd.temp_ref)
        # :657-663 kmajor(ntemp, neta, npres+1, ngpt); file (temperature, pressure_interp, eta, gpt)
        self.kmajor = F(np.transpose(kd.kmajor, (0, 2, 1, 3)))
        self.krayl = None
        if kd.rayl_lower is not None:
            # :667-677 krayl(ntemp, neta, ngpt, 2)
            self.krayl = F(np.stack([kd.rayl_lower, kd.rayl_upper], axis=3))
        self.press_ref_log = F(np.log(self.press_ref))
        self.press_ref_trop_log = float(np.log(kd.press_ref_trop))

        # :198-236 create_key_species_reduce; key_species(2, n_layers, n_bnds)
        ks = np.transpose(kd.key_species, (2, 1, 0)).astype(np.int32)
        ks_red = np.zeros_like(ks)
        for idx in np.ndindex(ks.shape):
            k = int(ks[idx])
            if k != 0:
                name = file_names[k - 1]
                if name not in self.gas_names:
                    raise RuntimeError("Gas optics: required gas " + name + " is missing")
                ks_red[idx] = self.gas_names.index(name) + 1
        # :247-311 create_flavor
        pairs = []
        for ibnd in range(ks_red.shape[2]):
            for iatm in range(ks_red.shape[1]):
                p = (int(ks_red[0, iatm, ibnd]), int(ks_red[1, iatm, ibnd]))
                pairs.append((2, 2) if p == (0, 0) else p)
        flav: List[tuple] = []
        for p in pairs:
            if p not in flav:
                flav.append(p)
        self.flavor = F(np.array(flav, dtype=np.int32).T, np.int32)       # (2, nflav)
        self.nflav = len(flav)
        # :331-362 create_gpoint_flavor -> (2, ngpt)
        gf = np.zeros((2, self.ngpt), dtype=np.int32)
        for ig in range(self.ngpt):
            for iatm in range(2):
                p = (int(ks_red[0, iatm, self.gpt2band[ig] - 1]), int(ks_red[1, iatm, self.gpt2band[ig] - 1]))
                p = (2, 2) if p == (0, 0) else p
                gf[iatm, ig] = flav.index(p) + 1
        self.gpoint_flavor = F(gf, np.int32)

        # :712-723
        self.temp_ref_min = float(self.temp_ref[0]); self.temp_ref_max = float(self.temp_ref[-1])
        self.press_ref_min = float(self.press_ref[-1]); self.press_ref_max = float(self.press_ref[0])
        self.press_ref_log_delta = (np.log(self.press_ref_min) - np.log(self.press_ref_max)) / (self.press_ref.size - 1)
        self.temp_ref_delta = (self.temp_ref_max - self.temp_ref_min) / (self.temp_ref.size - 1)
        self.ntemp, self.neta, self.npres = self.kmajor.shape[0], self.kmajor.shape[1], self.kmajor.shape[2] - 1

        self.totplnk = None
        if kd.totplnk is not None:
            # :426-433 planck_frac reorder; totplnk(nPlanckTemp, nbnd); :458 totplnk_delta
            self.planck_frac = F(np.transpose(kd.plank_fraction, (0, 2, 1, 3)))
            self.totplnk = F(kd.totplnk.T)
            self.totplnk_delta = (self.temp_ref_max - self.temp_ref_min) / (self.totplnk.shape[0] - 1)
        else:
            # :745-756 set_solar_variability
            self.solar_source = (kd.solar_source_quiet
                                 + (kd.mg_default - 0.1495954) * kd.solar_source_facular
                                 + (kd.sb_default - 0.00066696) * kd.solar_source_sunspot)
        if "h2o" not in self.gas_names:
            raise RuntimeError("idx_h2o cannot be found")
        self.idx_h2o = self.gas_names.index("h2o") + 1

    # -- src/Gas_optics_rrtmgp.cpp:764-792
    @staticmethod
    def get_col_dry(vmr_h2o: np.ndarray, plev: np.ndarray) -> np.ndarray:
        g0, avogad, m_dry, m_h2o = 9.80665, 6.02214076e23, 0.028964, 0.018016
        delta_plev = np.abs(plev[:, :-1] - plev[:, 1:])
        m_air = (m_dry + m_h2o * vmr_h2o) / (1.0 + vmr_h2o)
        col_dry = 10.0 * delta_plev * avogad / (1000.0 * m_air * 100.0 * g0)
        return F(col_dry / (1.0 + vmr_h2o))

    def _check(self, name, a, lo, hi):
        if np.any(a < lo) or np.any(a > hi):
            raise ValueError(name + " is out of range")   # std::range_error, :812-826

    # -- compute_gas_taus: src/Gas_optics_rrtmgp.cpp:1090-1275
    def _gas_taus(self, play, plev, tlay, vmr: Dict[str, np.ndarray], col_dry):
        ncol, nlay = play.shape
        ngas, nflav, ngpt = self.ngas, self.nflav, self.ngpt
        col_gas = zeros((ncol, nlay, ngas + 1))
        col_gas[:, :, 0] = col_dry
        for ig, name in enumerate(self.gas_names):            # :1121-1160
            v = np.broadcast_to(np.asarray(vmr[name], dtype=np.float64), (ncol, nlay))
            col_gas[:, :, ig + 1] = v * col_dry
        jtemp = zeros((ncol, nlay), np.int32); jpress = zeros((ncol, nlay), np.int32)
        tropo = zeros((ncol, nlay), np.int8)
        jeta = zeros((2, ncol, nlay, nflav), np.int32)
        col_mix = zeros((2, ncol, nlay, nflav)); fminor = zeros((2, 2, ncol, nlay, nflav))
        fmajor = zeros((2, 2, 2, ncol, nlay, nflav))
        play, tlay = F(play), F(tlay)
        lib().rrtmgp_interpolation(
            _i(ncol), _i(nlay), _i(ngas), _i(nflav), _i(self.neta), _i(self.npres), _i(self.ntemp),
            _p(self.flavor), _p(self.press_ref_log), _p(self.temp_ref),
            _d(self.press_ref_log_delta), _d(self.temp_ref_min), _d(self.temp_ref_delta),
            _d(self.press_ref_trop_log), _p(self.vmr_ref), _p(play), _p(tlay), _p(col_gas),
            _p(jtemp), _p(fmajor), _p(fminor), _p(col_mix), _p(tropo), _p(jeta), _p(jpress))
        tau = zeros((ncol, nlay, ngpt))
        lo, up = self.lower, self.upper
        lib().rrtmgp_compute_tau_absorption(
            _i(ncol), _i(nlay), _i(self.nbnd), _i(ngpt), _i(ngas), _i(nflav), _i(self.neta),
            _i(self.npres), _i(self.ntemp),
            _i(lo["dens"].size), _i(lo["kminor"].shape[2]), _i(up["dens"].size), _i(up["kminor"].shape[2]),
            _i(self.idx_h2o), _p(self.gpoint_flavor), _p(self.band2gpt), _p(self.kmajor),
            _p(lo["kminor"]), _p(up["kminor"]), _p(lo["limits"]), _p(up["limits"]),
            _p(lo["dens"]), _p(up["dens"]), _p(lo["comp"]), _p(up["comp"]),
            _p(lo["idx_minor"]), _p(up["idx_minor"]), _p(lo["idx_scal"]), _p(up["idx_scal"]),
            _p(lo["kstart"]), _p(up["kstart"]), _p(tropo), _p(col_mix), _p(fmajor), _p(fminor),
            _p(play), _p(tlay), _p(col_gas), _p(jeta), _p(jtemp), _p(jpress), _p(tau))
        st = dict(jtemp=jtemp, jpress=jpress, tropo=tropo, jeta=jeta, col_mix=col_mix,
                  fminor=fminor, fmajor=fmajor, col_gas=col_gas)
        if self.krayl is not None:                             # :1196-1241
            tau_r = zeros((ncol, nlay, ngpt))
            lib().rrtmgp_compute_tau_rayleigh(
                _i(ncol), _i(nlay), _i(self.nbnd), _i(ngpt), _i(ngas), _i(nflav), _i(self.neta),
                _i(self.npres), _i(self.ntemp), _p(self.gpoint_flavor), _p(self.band2gpt),
                _p(self.krayl), _i(self.idx_h2o), _p(F(col_dry)), _p(col_gas), _p(fminor), _p(jeta),
                _p(tropo), _p(jtemp), _p(tau_r))
            st["tau_abs"], st["tau_rayleigh"] = tau, tau_r
            # combine_abs_and_rayleigh_kernel :365-385 (threshold 2*epsilon), g zeroed :1308
            t = tau + tau_r
            ssa = np.where(t > 2.0 * np.finfo(np.float64).eps, tau_r / np.where(t > 0, t, 1.0), 0.0)
            st["tau"], st["ssa"], st["g"] = F(t), F(ssa), zeros((ncol, nlay, ngpt))
        else:
            st["tau"] = tau
        return st

    # -- gas_optics, longwave: src/Gas_optics_rrtmgp.cpp:796-849 + source() :1312-1344
    def gas_optics_lw(self, play, plev, tlay, tsfc, vmr, col_dry=None, tlev=None, check=True):
        play, plev, tlay, tlev, tsfc = F(play), F(plev), F(tlay), F(tlev), F(tsfc)
        if col_dry is None:
            col_dry = self.get_col_dry(np.broadcast_to(vmr["h2o"], play.shape), plev)
        if check:
            self._check("play", play, self.press_ref_min, self.press_ref_max)
            self._check("plev", plev, self.press_ref_min, self.press_ref_max)
            self._check("tlay", tlay, self.temp_ref_min, self.temp_ref_max)
            self._check("tlev", tlev, self.temp_ref_min, self.temp_ref_max)
            self._check("tsfc", tsfc, self.temp_ref_min, self.temp_ref_max)
            if np.any(col_dry < 0):
                raise ValueError("col_dry is out of range")
        st = self._gas_taus(play, plev, tlay, vmr, F(col_dry))
        ncol, nlay = play.shape
        ngpt = self.ngpt
        sfc_lay = 1 if play[0, 0] > play[0, nlay - 1] else nlay      # :1333
        sfc_src = zeros((ncol, ngpt)); sfc_jac = zeros((ncol, ngpt))
        lay_src = zeros((ncol, nlay, ngpt)); lev_src = zeros((ncol, nlay + 1, ngpt))
        lib().rrtmgp_compute_Planck_source(
            _i(ncol), _i(nlay), _i(self.nbnd), _i(ngpt), _i(self.nflav), _i(self.neta), _i(self.npres),
            _i(self.ntemp), _i(self.totplnk.shape[0]), _p(tlay), _p(tlev), _p(tsfc), _i(sfc_lay),
            _p(st["fmajor"]), _p(st["jeta"]), _p(st["tropo"]), _p(st["jtemp"]), _p(st["jpress"]),
            _p(self.gpt2band), _p(self.band2gpt), _p(self.planck_frac), _d(self.temp_ref_min),
            _d(self.totplnk_delta), _p(self.totplnk), _p(self.gpoint_flavor),
            _p(sfc_src), _p(lay_src), _p(lev_src), _p(sfc_jac))
        st.update(sfc_src=sfc_src, sfc_src_jac=sfc_jac, lay_src=lay_src, lev_src=lev_src,
                  col_dry=F(col_dry))
        return st

    # -- gas_optics, shortwave: src/Gas_optics_rrtmgp.cpp:853-898
    def gas_optics_sw(self, play, plev, tlay, vmr, col_dry=None, check=True):
        play, plev, tlay = F(play), F(plev), F(tlay)
        if col_dry is None:
            col_dry = self.get_col_dry(np.broadcast_to(vmr["h2o"], play.shape), plev)
        if check:
            self._check("play", play, self.press_ref_min, self.press_ref_max)
            self._check("plev", plev, self.press_ref_min, self.press_ref_max)
            self._check("tlay", tlay, self.temp_ref_min, self.temp_ref_max)
            if np.any(col_dry < 0):
                raise ValueError("col_dry is out of range")
        st = self._gas_taus(play, plev, tlay, vmr, F(col_dry))
        ncol = play.shape[0]
        st["toa_src"] = F(np.tile(self.solar_source[None, :], (ncol, 1)))   # :895-897
        st["col_dry"] = F(col_dry)
        return st

    def get_tsi(self) -> float:                                             # :1347-1356
        return float(np.cumsum(self.solar_source)[-1])   # sequential igpt = 1..ngpt, like the reference loop


# ------------------------------------------------------------------------------------------
# Cloud_optics: src/Cloud_optics.cpp (pure C++ in the reference; restated in numpy)
# ------------------------------------------------------------------------------------------
class CloudOptics(SpectralDisc):
    def __init__(self, lut):
        super().__init__(lut.bnd_limits_wavenumber.T)
        icergh = 2                                                          # :59
        self.liq_nsteps = lut.lut_extliq.shape[1]; self.ice_nsteps = lut.lut_extice.shape[2]
        self.liq_step = (lut.radliq_upr - lut.radliq_lwr) / (self.liq_nsteps - 1.0)
        self.ice_step = (lut.diamice_upr - lut.diamice_lwr) / (self.ice_nsteps - 1.0)
        self.radliq_lwr, self.diamice_lwr = lut.radliq_lwr, lut.diamice_lwr
        self.liq = [F(a.T) for a in (lut.lut_extliq, lut.lut_ssaliq, lut.lut_asyliq)]      # (nsize, nbnd)
        self.ice = [F(a[icergh - 1].T) for a in (lut.lut_extice, lut.lut_ssaice, lut.lut_asyice)]

    @staticmethod
    def _from_table(cwp, re, nsteps, step, offset, tabs):
        """compute_all_from_table, src/Cloud_optics.cpp:72-107 -> (ncol,nlay,nbnd) x3"""
        mask = cwp > 0.0
        x = np.where(mask, (re - offset) / step, 0.0)
        index = np.minimum(x.astype(np.int64) + 1, nsteps - 1)
        fint = x - (index - 1)
        i0 = index - 1
        def interp(tab):
            return tab[i0, :] + fint[..., None] * (tab[i0 + 1, :] - tab[i0, :])
        tau = cwp[..., None] * interp(tabs[0])
        taussa = tau * interp(tabs[1])
        taussag = taussa * interp(tabs[2])
        m = mask[..., None]
        return np.where(m, tau, 0.0), np.where(m, taussa, 0.0), np.where(m, taussag, 0.0)

    def cloud_optics_2str(self, lwp, iwp, rel, dei):                        # :111-172
        lt, lts, ltsg = self._from_table(lwp, rel, self.liq_nsteps, self.liq_step, self.radliq_lwr, self.liq)
        it, its, itsg = self._from_table(iwp, dei, self.ice_nsteps, self.ice_step, self.diamice_lwr, self.ice)
        eps = np.finfo(np.float64).eps
        tau = lt + it; taussa = lts + its; taussag = ltsg + itsg
        return F(tau), F(taussa / np.maximum(tau, eps)), F(taussag / np.maximum(taussa, eps))

    def cloud_optics_1scl(self, lwp, iwp, rel, dei):                        # :176-232
        lt, lts, _ = self._from_table(lwp, rel, self.liq_nsteps, self.liq_step, self.radliq_lwr, self.liq)
        it, its, _ = self._from_table(iwp, dei, self.ice_nsteps, self.ice_step, self.diamice_lwr, self.ice)
        return F((lt - lts) + (it - its))


# ------------------------------------------------------------------------------------------
# Aerosol_optics: src/Aerosol_optics.cpp (pure C++ in the reference; restated in numpy)
# ------------------------------------------------------------------------------------------
class AerosolOptics(SpectralDisc):
    # (species name, "phobic" column | "philic" type (1-based), aermr index) in the summation order of
    # list_aerosol_types, src/Aerosol_optics.cpp:51,64-142
    SPECIES = (("SS1", -1, 1), ("SS2", -2, 2), ("SS3", -3, 3), ("DU1", 1, 4), ("DU2", 8, 5), ("DU3", 6, 6),
               ("OM1", 10, 8), ("OM2", -4, 7), ("BC1", 11, 9), ("BC2", 11, 10), ("SU", -5, 11))

    def __init__(self, lut):
        nb = lut.mass_ext_sw_hydrophobic.shape[1]
        super().__init__(np.zeros((2, nb)))          # the reference's loader leaves band_lims_wvn zero (:342)
        self.phobic = [F(a) for a in (lut.mass_ext_sw_hydrophobic, lut.ssa_sw_hydrophobic, lut.asymmetry_sw_hydrophobic)]
        self.philic = [F(a) for a in (lut.mass_ext_sw_hydrophilic, lut.ssa_sw_hydrophilic, lut.asymmetry_sw_hydrophilic)]
        self.rh_upper = F(lut.relative_humidity2)

    def aerosol_optics(self, aermr: Dict[str, np.ndarray], rh, plev, abs_dp: bool = False):
        """Aerosol_optics::aerosol_optics, src/Aerosol_optics.cpp:167-224 -> tau, ssa, g (ncol,nlay,nbnd).
        abs_dp=False is the CPU source (dpg signed, :53); True the GPU twin (src_cuda/Aerosol_optics.cu:63)."""
        rh = F(rh); plev = F(plev)
        ncol, nlay = rh.shape
        # rh_class (:27-38): first class with rh_upper >= rh
        ihum = np.searchsorted(self.rh_upper, rh, side="left")
        if np.any(ihum >= self.rh_upper.size):
            raise ValueError("relative humidity above the last class: the reference reads past the table")
        dpg = (plev[:, :-1] - plev[:, 1:]) / 9.81
        if abs_dp:
            dpg = np.abs(dpg)
        nb = self.nbnd
        tau = np.zeros((ncol, nlay, nb)); taussa = np.zeros_like(tau); taussag = np.zeros_like(tau)
        for _, col, iaer in self.SPECIES:
            mmr = np.broadcast_to(F(aermr[f"aermr{iaer:02d}"]), (ncol, nlay))            # fill_aerosols_3d (:152-165)
            if col > 0:
                mext, ssa, g = (np.broadcast_to(t[col - 1][None, None, :], (ncol, nlay, nb)) for t in self.phobic)
            else:
                mext, ssa, g = (t[-col - 1][ihum, :] for t in self.philic)
            local_od = (mmr * dpg)[..., None] * mext
            tau = tau + local_od
            taussa = taussa + local_od * ssa
            taussag = taussag + local_od * ssa * g
        eps = np.finfo(np.float64).eps
        return F(tau), F(taussa / np.maximum(tau, eps)), F(taussag / np.maximum(taussa, eps))


# ------------------------------------------------------------------------------------------
# Optical_props: delta_scale / add_to — src/Optical_props.cpp:207-268
# ------------------------------------------------------------------------------------------
def delta_scale(tau, ssa, g):
    tau, ssa, g = F(tau).copy(order="F"), F(ssa).copy(order="F"), F(g).copy(order="F")
    n = tau.shape
    lib().rte_delta_scale_2str_k(_i(n[0]), _i(n[1]), _i(n[2]), _p(tau), _p(ssa), _p(g))
    return tau, ssa, g


def add_to_1scl(disc: SpectralDisc, tau, tau_in):
    tau = F(tau).copy(order="F"); tau_in = F(tau_in)
    n = tau.shape
    if tau_in.shape[2] == n[2]:
        lib().rte_increment_1scalar_by_1scalar(_i(n[0]), _i(n[1]), _i(n[2]), _p(tau), _p(tau_in))
    else:
        if tau_in.shape[2] != disc.nbnd:
            raise RuntimeError("Cannot add optical properties with incompatible band - gpoint combination")
        lib().rte_inc_1scalar_by_1scalar_bybnd(_i(n[0]), _i(n[1]), _i(n[2]), _p(tau), _p(tau_in),
                                               _i(disc.nbnd), _p(disc.band2gpt))
    return tau


def add_to_2str(disc: SpectralDisc, tau, ssa, g, tau_in, ssa_in, g_in):
    tau, ssa, g = F(tau).copy(order="F"), F(ssa).copy(order="F"), F(g).copy(order="F")
    tau_in, ssa_in, g_in = F(tau_in), F(ssa_in), F(g_in)
    n = tau.shape
    if tau_in.shape[2] == n[2]:
        lib().rte_increment_2stream_by_2stream(_i(n[0]), _i(n[1]), _i(n[2]), _p(tau), _p(ssa), _p(g),
                                               _p(tau_in), _p(ssa_in), _p(g_in))
    else:
        if tau_in.shape[2] != disc.nbnd:
            raise RuntimeError("Cannot add optical properties with incompatible band - gpoint combination")
        lib().rte_inc_2stream_by_2stream_bybnd(_i(n[0]), _i(n[1]), _i(n[2]), _p(tau), _p(ssa), _p(g),
                                               _p(tau_in), _p(ssa_in), _p(g_in),
                                               _i(disc.nbnd), _p(disc.band2gpt))
    return tau, ssa, g


# ------------------------------------------------------------------------------------------
# Rte_lw / Rte_sw: src/Rte_lw.cpp:126-217, src/Rte_sw.cpp:136-201
# ------------------------------------------------------------------------------------------
GAUSS_DS = np.array([[1. / 0.6096748751, 0., 0., 0.],
                     [1. / 0.2509907356, 1 / 0.7908473988, 0., 0.],
                     [1. / 0.1024922169, 1 / 0.4417960320, 1. / 0.8633751621, 0.],
                     [1. / 0.0454586727, 1 / 0.2322334416, 1. / 0.5740198775, 1. / 0.903077597]])
GAUSS_WTS = np.array([[1., 0., 0., 0.],
                      [0.2300253764, 0.7699746236, 0., 0.],
                      [0.0437820218, 0.3875796738, 0.5686383044, 0.],
                      [0.0092068785, 0.1285704278, 0.4323381850, 0.4298845087]])


def expand_and_transpose(disc: SpectralDisc, arr_in):
    """(nbnd, ncol) -> (ncol, ngpt); src/Rte_lw.cpp:204-217"""
    arr_in = np.asarray(arr_in)
    out = zeros((arr_in.shape[1], disc.ngpt))
    for ib in range(disc.nbnd):
        out[:, disc.band2gpt[0, ib] - 1:disc.band2gpt[1, ib]] = arr_in[ib, :, None]
    return out


def rte_lw(disc, tau, lay_src, lev_src, sfc_src, sfc_emis, top_at_1, inc_flux=None,
           n_gauss_angles=1, broadband=False, sfc_src_jac=None, ssa=None, g=None):
    """ssa and g given: the re-scaled transport (do_rescaling, mo_rte_solver_kernels.F90:148-178,753-844) that the
    Fortran front end uses for 2-stream optical properties; the reference C++ always passes false (src/Rte_lw.cpp:178)."""
    tau, lay_src, lev_src, sfc_src = F(tau), F(lay_src), F(lev_src), F(sfc_src)
    rescale = ssa is not None
    ssa_, g_ = (F(ssa), F(g)) if rescale else (tau, tau)
    ncol, nlay, ngpt = tau.shape
    sfc_emis_gpt = expand_and_transpose(disc, sfc_emis)
    n = n_gauss_angles
    secants = zeros((ncol, ngpt, n))
    for imu in range(n):
        secants[:, :, imu] = GAUSS_DS[n - 1, imu]
    wts = F(GAUSS_WTS[n - 1, :n].copy())
    inc = zeros((ncol, ngpt)) if inc_flux is None else F(inc_flux)
    ng_out = 1 if broadband else ngpt
    up = zeros((ncol, nlay + 1, ng_out)); dn = zeros((ncol, nlay + 1, ng_out))
    do_jac = sfc_src_jac is not None
    jac_in = F(sfc_src_jac) if do_jac else zeros((ncol, ngpt))
    jac = zeros((ncol, nlay + 1))
    lib().rte_lw_solver_noscat(
        _i(ncol), _i(nlay), _i(ngpt), _b(top_at_1), _i(n), _p(secants), _p(wts), _p(tau),
        _p(lay_src), _p(lev_src), _p(sfc_emis_gpt), _p(sfc_src), _p(inc), _p(up), _p(dn),
        _b(broadband), _p(up), _p(dn), _b(do_jac), _p(jac_in), _p(jac), _b(rescale), _p(ssa_), _p(g_))
    return (up, dn, jac) if do_jac else (up, dn)


def rte_sw(disc, tau, ssa, g, mu0, inc_flux_dir, sfc_alb_dir, sfc_alb_dif, top_at_1,
           inc_flux_dif=None, broadband=False):
    tau, ssa, g = F(tau), F(ssa), F(g)
    ncol, nlay, ngpt = tau.shape
    a_dir = expand_and_transpose(disc, sfc_alb_dir)
    a_dif = expand_and_transpose(disc, sfc_alb_dif)
    mu0 = np.asarray(mu0)                        # (ncol) as src/Rte_sw.cpp:157-160 expands it, or (ncol, nlay)
    mu0_2d = F(mu0) if mu0.ndim == 2 else F(np.tile(mu0[:, None], (1, nlay)))
    ng_out = 1 if broadband else ngpt
    up = zeros((ncol, nlay + 1, ng_out)); dn = zeros((ncol, nlay + 1, ng_out))
    dr = zeros((ncol, nlay + 1, ng_out))
    has_dif = inc_flux_dif is not None
    dif = F(inc_flux_dif) if has_dif else zeros((ncol, ngpt))
    lib().rte_sw_solver_2stream(
        _i(ncol), _i(nlay), _i(ngpt), _b(top_at_1), _p(tau), _p(ssa), _p(g), _p(mu0_2d),
        _p(a_dir), _p(a_dif), _p(F(inc_flux_dir)), _p(up), _p(dn), _p(dr), _b(has_dif), _p(dif),
        _b(broadband), _p(up), _p(dn), _p(dr))
    return up, dn, dr


# ------------------------------------------------------------------------------------------
# Fluxes: src/Fluxes.cpp:95-196
# ------------------------------------------------------------------------------------------
def sum_broadband(gpt_flux):
    gpt_flux = F(gpt_flux)
    ncol, nlev, ngpt = gpt_flux.shape
    out = zeros((ncol, nlev))
    lib().rte_sum_broadband(_i(ncol), _i(nlev), _i(ngpt), _p(gpt_flux), _p(out))
    return out


def sum_byband(disc, gpt_flux):
    gpt_flux = F(gpt_flux)
    ncol, nlev, ngpt = gpt_flux.shape
    out = zeros((ncol, nlev, disc.nbnd))
    lib().sum_byband(_i(ncol), _i(nlev), _i(ngpt), _i(disc.nbnd), _p(disc.band2gpt), _p(gpt_flux), _p(out))
    return out


def fluxes_broadband(up, dn, dn_dir=None):
    r = dict(flux_up=sum_broadband(up), flux_dn=sum_broadband(dn))
    r["flux_net"] = F(r["flux_dn"] - r["flux_up"])
    if dn_dir is not None:
        r["flux_dn_dir"] = sum_broadband(dn_dir)
    return r


def fluxes_byband(disc, up, dn, dn_dir=None):
    r = fluxes_broadband(up, dn, dn_dir)
    r["bnd_flux_up"] = sum_byband(disc, up); r["bnd_flux_dn"] = sum_byband(disc, dn)
    r["bnd_flux_net"] = F(r["bnd_flux_dn"] - r["bnd_flux_up"])
    if dn_dir is not None:
        r["bnd_flux_dn_dir"] = sum_byband(disc, dn_dir)
    return r


def heating_rate(flux_up, flux_dn, plev):
    flux_up, flux_dn, plev = F(flux_up), F(flux_dn), F(plev)
    ncol, nlev = flux_up.shape
    out = zeros((ncol, nlev - 1))
    lib().oracle_heating_rate.argtypes = [C.c_int, C.c_int, c_dbl_p, c_dbl_p, c_dbl_p, c_dbl_p]
    lib().oracle_heating_rate(ncol, nlev - 1, _p(flux_up), _p(flux_dn), _p(plev), _p(out))
    return out

"""Shared test helpers: the oracle chain for a set of columns (CPU, fp64) — checker only."""
import numpy as np

from oracle import oracle as O


def oracle_lw(kd, atm, cloud_lut=None, n_ang=1, col_dry=None):
    go = O.GasOptics(kd, atm.vmr.keys())
    st = go.gas_optics_lw(atm.play, atm.plev, atm.tlay, atm.tsfc, atm.vmr, col_dry, atm.tlev)
    tau = st["tau"]
    if cloud_lut is not None:
        co = O.CloudOptics(cloud_lut)
        ctau = co.cloud_optics_1scl(atm.lwp, atm.iwp, atm.rel, atm.dei)
        tau = O.add_to_1scl(go, tau, ctau)
    up, dn = O.rte_lw(go, tau, st["lay_src"], st["lev_src"], st["sfc_src"], atm.emis_sfc, atm.top_at_1,
                      n_gauss_angles=n_ang)
    fl = O.fluxes_byband(go, up, dn)
    fl["heating_rate"] = O.heating_rate(fl["flux_up"], fl["flux_dn"], atm.plev)
    fl["state"] = st
    fl["gpt_up"], fl["gpt_dn"], fl["tau"] = up, dn, tau
    fl["go"] = go
    return fl


def oracle_sw(kd, atm, cloud_lut=None, delta=True, col_dry=None, aerosol_lut=None, delta_aerosol=False):
    go = O.GasOptics(kd, atm.vmr.keys())
    st = go.gas_optics_sw(atm.play, atm.plev, atm.tlay, atm.vmr, col_dry)
    tau, ssa, g = st["tau"], st["ssa"], st["g"]
    if cloud_lut is not None:
        co = O.CloudOptics(cloud_lut)
        c = co.cloud_optics_2str(atm.lwp, atm.iwp, atm.rel, atm.dei)
        if delta:
            c = O.delta_scale(*c)
        tau, ssa, g = O.add_to_2str(go, tau, ssa, g, *c)
    if aerosol_lut is not None:
        # after the clouds, src_test/Radiation_solver.cpp:770-786; layer mass as the GPU twin takes it (|dp|),
        # identical to the CPU source for bottom-up columns
        a = O.AerosolOptics(aerosol_lut).aerosol_optics(atm.aermr, atm.rh, atm.plev, abs_dp=True)
        if delta_aerosol:
            a = O.delta_scale(*a)
        tau, ssa, g = O.add_to_2str(go, tau, ssa, g, *a)
    toa = st["toa_src"] * np.asarray(atm.tsi_scaling)[:, None]
    up, dn, dr = O.rte_sw(go, tau, ssa, g, atm.mu0, toa, atm.sfc_alb_dir, atm.sfc_alb_dif, atm.top_at_1)
    fl = O.fluxes_byband(go, up, dn, dr)
    fl["heating_rate"] = O.heating_rate(fl["flux_up"], fl["flux_dn"], atm.plev)
    fl["state"] = st
    fl["gpt_up"], fl["gpt_dn"], fl["gpt_dir"] = up, dn, dr
    fl["tau"], fl["ssa"], fl["g"], fl["toa"] = tau, ssa, g, toa
    fl["go"] = go
    return fl


def as_f32_valued(atm):
    """The same columns with every input rounded to fp32 (what the fp32 build is handed), kept
    as fp64 arrays so the fp64 oracle sees identical input values."""
    import copy
    a = copy.copy(atm)
    r = lambda x: None if x is None else np.asfortranarray(np.asarray(x, dtype=np.float32).astype(np.float64))
    for k in ("play", "plev", "tlay", "tlev", "tsfc", "emis_sfc", "sfc_alb_dir", "sfc_alb_dif", "mu0",
              "tsi_scaling", "lwp", "iwp", "rel", "dei", "rh"):
        setattr(a, k, r(getattr(atm, k)))
    a.vmr = {k: r(v) for k, v in atm.vmr.items()}
    a.aermr = None if atm.aermr is None else {k: r(v) for k, v in atm.aermr.items()}
    return a

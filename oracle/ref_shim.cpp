/*
 * ORACLE — TEST INFRASTRUCTURE ONLY.
 *
 * Flat C entry points over the reference's OWN, UNMODIFIED C++ class layer
 * (/root/reference/src/{Gas_optics_rrtmgp,Optical_props,Rte_lw,Rte_sw,Fluxes,Source_functions,
 * Gas_concs,Cloud_optics,Aerosol_optics}.cpp, compiled where they lie by oracle/Makefile into
 * oracle/_ref/liboracle_ref.so), linked over oracle/rrtmgp_kernels_restated.c in place of the
 * Fortran kernels that cannot be compiled here.  Construction follows what the reference's
 * driver does with the netCDF variables (src_test/Radiation_solver.cpp:70-284), using the same
 * Array dims.  Used to validate oracle/oracle.py's numpy restatement of the class layer.
 */
#include <cstring>
#include <cstdio>
#include <memory>
#include <string>
#include <vector>
#include <stdexcept>
#include <algorithm>
#include <cmath>

#include "Array.h"
#include "Gas_concs.h"
#include "Gas_optics_rrtmgp.h"
#include "Optical_props.h"
#include "Source_functions.h"
#include "Rte_lw.h"
#include "Rte_sw.h"
#include "Fluxes.h"
#include "Cloud_optics.h"
#include "Aerosol_optics.h"

#include "rrtmgp_b200_tables.h"
#include "ref_kdist.hpp"

namespace
{
    thread_local std::string last_error;
    using refbuild::strings; using refbuild::barr;
    template<typename T, int N> Array<T, N> arr(const T* p, std::array<int, N> dims) { return refbuild::arr<T, N, T>(p, dims); }

    Gas_concs make_concs(int ncol, int nlay, int ngas, const char* names, int len,
                         const double* const* vmr, const int* vmr_ncol, const int* vmr_nlay)
    {
        Gas_concs gc;
        auto nm = strings(names, ngas, len, true);
        for (int i = 0; i < ngas; ++i)
        {
            (void)ncol; (void)nlay;
            gc.set_vmr(nm[i], arr<Float, 2>(vmr[i], {vmr_ncol[i], vmr_nlay[i]}));
        }
        return gc;
    }

    struct Ref_gas_optics
    {
        std::unique_ptr<Gas_optics_rrtmgp> kdist;
    };
}

#define SHIM_TRY try {
#define SHIM_CATCH } catch (const std::exception& e) { last_error = e.what(); return 1; } \
                     catch (...) { last_error = "unknown exception"; return 1; } return 0;

extern "C"
{
const char* ref_last_error() { return last_error.c_str(); }

/* Build Gas_optics_rrtmgp as load_and_init_gas_optics does (Radiation_solver.cpp:70-284). */
int ref_gas_optics_create(const rrtmgp_b200_kdist_file* f, int n_avail, const char* avail_names,
                          int avail_len, void** out)
{
    SHIM_TRY
    Gas_concs avail;
    for (auto& s : strings(avail_names, n_avail, avail_len, true)) avail.set_vmr(s, Float(0.));

    auto* h = new Ref_gas_optics;
    h->kdist = refbuild::build_kdist<Gas_optics_rrtmgp, Gas_concs>(f, avail);
    *out = h;
    SHIM_CATCH
}

void ref_gas_optics_destroy(void* h) { delete static_cast<Ref_gas_optics*>(h); }

int ref_gas_optics_info(void* hv, int* ngpt, int* nbnd, int* nflav, int* ngas)
{
    SHIM_TRY
    auto* h = static_cast<Ref_gas_optics*>(hv);
    *ngpt = h->kdist->get_ngpt(); *nbnd = h->kdist->get_nband();
    *nflav = h->kdist->get_nflav(); *ngas = -1; /* get_ngas() is private in the reference */
    SHIM_CATCH
}

/* Gas_optics_rrtmgp::get_col_dry (src/Gas_optics_rrtmgp.cpp:764-792) */
int ref_get_col_dry(int ncol, int nlay, const double* vmr_h2o, const double* plev, double* col_dry)
{
    SHIM_TRY
    Array<Float,2> cd({ncol, nlay});
    Gas_optics_rrtmgp::get_col_dry(cd, arr<Float,2>(vmr_h2o, {ncol, nlay}), arr<Float,2>(plev, {ncol, nlay + 1}));
    std::copy(cd.v().begin(), cd.v().end(), col_dry);
    SHIM_CATCH
}

/* LW gas_optics (src/Gas_optics_rrtmgp.cpp:796-849) */
int ref_gas_optics_lw(void* hv, int ncol, int nlay,
                      const double* play, const double* plev, const double* tlay, const double* tsfc,
                      const double* tlev, const double* col_dry,
                      int ngas, const char* names, int len, const double* const* vmr,
                      const int* vmr_ncol, const int* vmr_nlay,
                      double* tau, double* lay_src, double* lev_src, double* sfc_src, double* sfc_src_jac)
{
    SHIM_TRY
    auto* h = static_cast<Ref_gas_optics*>(hv);
    Gas_concs gc = make_concs(ncol, nlay, ngas, names, len, vmr, vmr_ncol, vmr_nlay);
    std::unique_ptr<Optical_props_arry> op = std::make_unique<Optical_props_1scl>(ncol, nlay, *h->kdist);
    Source_func_lw src(ncol, nlay, *h->kdist);
    h->kdist->gas_optics(arr<Float,2>(play, {ncol, nlay}), arr<Float,2>(plev, {ncol, nlay + 1}),
                         arr<Float,2>(tlay, {ncol, nlay}), arr<Float,1>(tsfc, {ncol}), gc, op, src,
                         arr<Float,2>(col_dry, {ncol, nlay}), arr<Float,2>(tlev, {ncol, nlay + 1}));
    std::copy(op->get_tau().v().begin(), op->get_tau().v().end(), tau);
    std::copy(src.get_lay_source().v().begin(), src.get_lay_source().v().end(), lay_src);
    std::copy(src.get_lev_source().v().begin(), src.get_lev_source().v().end(), lev_src);
    std::copy(src.get_sfc_source().v().begin(), src.get_sfc_source().v().end(), sfc_src);
    std::copy(src.get_sfc_source_jac().v().begin(), src.get_sfc_source_jac().v().end(), sfc_src_jac);
    SHIM_CATCH
}

/* SW gas_optics (src/Gas_optics_rrtmgp.cpp:853-898) */
int ref_gas_optics_sw(void* hv, int ncol, int nlay,
                      const double* play, const double* plev, const double* tlay, const double* col_dry,
                      int ngas, const char* names, int len, const double* const* vmr,
                      const int* vmr_ncol, const int* vmr_nlay,
                      double* tau, double* ssa, double* g, double* toa_src)
{
    SHIM_TRY
    auto* h = static_cast<Ref_gas_optics*>(hv);
    Gas_concs gc = make_concs(ncol, nlay, ngas, names, len, vmr, vmr_ncol, vmr_nlay);
    std::unique_ptr<Optical_props_arry> op = std::make_unique<Optical_props_2str>(ncol, nlay, *h->kdist);
    Array<Float,2> toa({ncol, h->kdist->get_ngpt()});
    h->kdist->gas_optics(arr<Float,2>(play, {ncol, nlay}), arr<Float,2>(plev, {ncol, nlay + 1}),
                         arr<Float,2>(tlay, {ncol, nlay}), gc, op, toa, arr<Float,2>(col_dry, {ncol, nlay}));
    std::copy(op->get_tau().v().begin(), op->get_tau().v().end(), tau);
    std::copy(op->get_ssa().v().begin(), op->get_ssa().v().end(), ssa);
    std::copy(op->get_g().v().begin(), op->get_g().v().end(), g);
    std::copy(toa.v().begin(), toa.v().end(), toa_src);
    SHIM_CATCH
}

/* Rte_lw::rte_lw (src/Rte_lw.cpp:126-201); ng_out = ngpt or 1 (broadband) */
int ref_rte_lw(void* hv, int ncol, int nlay, int top_at_1, int n_ang, int ng_out,
               const double* tau, const double* lay_src, const double* lev_src, const double* sfc_src,
               const double* sfc_emis /* (nbnd,ncol) */, double* flux_up, double* flux_dn)
{
    SHIM_TRY
    auto* h = static_cast<Ref_gas_optics*>(hv);
    const int ngpt = h->kdist->get_ngpt(), nbnd = h->kdist->get_nband();
    std::unique_ptr<Optical_props_arry> op = std::make_unique<Optical_props_1scl>(ncol, nlay, *h->kdist);
    op->get_tau() = std::vector<Float>(tau, tau + (size_t)ncol * nlay * ngpt);
    Source_func_lw src(ncol, nlay, *h->kdist);
    src.get_lay_source() = std::vector<Float>(lay_src, lay_src + (size_t)ncol * nlay * ngpt);
    src.get_lev_source() = std::vector<Float>(lev_src, lev_src + (size_t)ncol * (nlay + 1) * ngpt);
    src.get_sfc_source() = std::vector<Float>(sfc_src, sfc_src + (size_t)ncol * ngpt);
    Array<Float,3> up({ncol, nlay + 1, ng_out}), dn({ncol, nlay + 1, ng_out});
    Rte_lw::rte_lw(op, (Bool)top_at_1, src, arr<Float,2>(sfc_emis, {nbnd, ncol}),
                   Array<Float,2>({ncol, ngpt}), up, dn, n_ang);
    std::copy(up.v().begin(), up.v().end(), flux_up);
    std::copy(dn.v().begin(), dn.v().end(), flux_dn);
    SHIM_CATCH
}

/* Rte_sw::rte_sw (src/Rte_sw.cpp:136-184) */
int ref_rte_sw(void* hv, int ncol, int nlay, int top_at_1, int ng_out,
               const double* tau, const double* ssa, const double* g, const double* mu0,
               const double* inc_flux_dir, const double* sfc_alb_dir, const double* sfc_alb_dif,
               double* flux_up, double* flux_dn, double* flux_dir)
{
    SHIM_TRY
    auto* h = static_cast<Ref_gas_optics*>(hv);
    const int ngpt = h->kdist->get_ngpt(), nbnd = h->kdist->get_nband();
    const size_t n = (size_t)ncol * nlay * ngpt;
    std::unique_ptr<Optical_props_arry> op = std::make_unique<Optical_props_2str>(ncol, nlay, *h->kdist);
    op->get_tau() = std::vector<Float>(tau, tau + n);
    op->get_ssa() = std::vector<Float>(ssa, ssa + n);
    op->get_g() = std::vector<Float>(g, g + n);
    Array<Float,3> up({ncol, nlay + 1, ng_out}), dn({ncol, nlay + 1, ng_out}), dr({ncol, nlay + 1, ng_out});
    Rte_sw::rte_sw(op, (Bool)top_at_1, arr<Float,1>(mu0, {ncol}), arr<Float,2>(inc_flux_dir, {ncol, ngpt}),
                   arr<Float,2>(sfc_alb_dir, {nbnd, ncol}), arr<Float,2>(sfc_alb_dif, {nbnd, ncol}),
                   Array<Float,2>(), up, dn, dr);
    std::copy(up.v().begin(), up.v().end(), flux_up);
    std::copy(dn.v().begin(), dn.v().end(), flux_dn);
    std::copy(dr.v().begin(), dr.v().end(), flux_dir);
    SHIM_CATCH
}

/* Cloud_optics (src/Cloud_optics.cpp) + optional delta_scale + add_to onto gas props.
 * mode 1: 1scl (tau_gas in/out); mode 2: 2str (tau,ssa,g in/out). Also returns cloud props. */
int ref_cloud_add(void* hv, const rrtmgp_b200_cloud_file* c, int ncol, int nlay, int mode, int do_delta,
                  const double* lwp, const double* iwp, const double* rel, const double* dei,
                  double* tau, double* ssa, double* g,
                  double* ctau, double* cssa, double* cg)
{
    SHIM_TRY
    auto* h = static_cast<Ref_gas_optics*>(hv);
    const int ngpt = h->kdist->get_ngpt();
    const size_t n = (size_t)ncol * nlay * ngpt, nb = (size_t)ncol * nlay * c->n_band;
    Cloud_optics co(arr<Float,2>(c->bnd_limits_wavenumber, {2, c->n_band}),
                    c->radliq_lwr, c->radliq_upr, c->radliq_fac, c->diamice_lwr, c->diamice_upr, c->diamice_fac,
                    arr<Float,2>(c->lut_extliq, {c->n_size_liq, c->n_band}),
                    arr<Float,2>(c->lut_ssaliq, {c->n_size_liq, c->n_band}),
                    arr<Float,2>(c->lut_asyliq, {c->n_size_liq, c->n_band}),
                    arr<Float,3>(c->lut_extice, {c->n_size_ice, c->n_band, c->n_rghice}),
                    arr<Float,3>(c->lut_ssaice, {c->n_size_ice, c->n_band, c->n_rghice}),
                    arr<Float,3>(c->lut_asyice, {c->n_size_ice, c->n_band, c->n_rghice}));
    auto a_lwp = arr<Float,2>(lwp, {ncol, nlay}); auto a_iwp = arr<Float,2>(iwp, {ncol, nlay});
    auto a_rel = arr<Float,2>(rel, {ncol, nlay}); auto a_dei = arr<Float,2>(dei, {ncol, nlay});
    if (mode == 1)
    {
        Optical_props_1scl cl(ncol, nlay, co);
        co.cloud_optics(a_lwp, a_iwp, a_rel, a_dei, cl);
        std::copy(cl.get_tau().v().begin(), cl.get_tau().v().end(), ctau);
        Optical_props_1scl gas(ncol, nlay, *h->kdist);
        gas.get_tau() = std::vector<Float>(tau, tau + n);
        add_to(gas, cl);
        std::copy(gas.get_tau().v().begin(), gas.get_tau().v().end(), tau);
    }
    else
    {
        Optical_props_2str cl(ncol, nlay, co);
        co.cloud_optics(a_lwp, a_iwp, a_rel, a_dei, cl);
        if (do_delta) cl.delta_scale();
        std::copy(cl.get_tau().v().begin(), cl.get_tau().v().begin() + nb, ctau);
        std::copy(cl.get_ssa().v().begin(), cl.get_ssa().v().begin() + nb, cssa);
        std::copy(cl.get_g().v().begin(), cl.get_g().v().begin() + nb, cg);
        Optical_props_2str gas(ncol, nlay, *h->kdist);
        gas.get_tau() = std::vector<Float>(tau, tau + n);
        gas.get_ssa() = std::vector<Float>(ssa, ssa + n);
        gas.get_g() = std::vector<Float>(g, g + n);
        add_to(gas, cl);
        std::copy(gas.get_tau().v().begin(), gas.get_tau().v().end(), tau);
        std::copy(gas.get_ssa().v().begin(), gas.get_ssa().v().end(), ssa);
        std::copy(gas.get_g().v().begin(), gas.get_g().v().end(), g);
    }
    SHIM_CATCH
}

/* Aerosol_optics (src/Aerosol_optics.cpp) built like load_and_init_aerosol_optics
 * (src_test/Radiation_solver.cpp:331-367) + optional delta_scale + add_to onto 2-stream gas props
 * (src_test/Radiation_solver.cpp:770-786).  mmr[i] is aermr<i+1>, shape (mmr_ncol[i], mmr_nlay[i]).
 * Returns the aerosol props (atau,assa,ag) and the incremented gas props in place. */
int ref_aerosol_add(void* hv, const rrtmgp_b200_aerosol_file* a, int ncol, int nlay, int do_delta,
                    const double* const* mmr, const int* mmr_ncol, const int* mmr_nlay,
                    const double* rh, const double* plev,
                    double* tau, double* ssa, double* g, double* atau, double* assa, double* ag)
{
    SHIM_TRY
    auto* h = static_cast<Ref_gas_optics*>(hv);
    const int ngpt = h->kdist->get_ngpt();
    const int nb = a->n_band, nh = a->n_hum, nphil = a->n_philic, nphob = a->n_phobic;
    const size_t n = (size_t)ncol * nlay * ngpt, nbn = (size_t)ncol * nlay * nb;
    Array<Float,2> band_lims_wvn({2, nb});
    Aerosol_optics ao(band_lims_wvn, arr<Float,1>(a->relative_humidity2, {nh}),
                      arr<Float,2>(a->mass_ext_sw_hydrophobic, {nb, nphob}),
                      arr<Float,2>(a->ssa_sw_hydrophobic, {nb, nphob}),
                      arr<Float,2>(a->asymmetry_sw_hydrophobic, {nb, nphob}),
                      arr<Float,3>(a->mass_ext_sw_hydrophilic, {nb, nh, nphil}),
                      arr<Float,3>(a->ssa_sw_hydrophilic, {nb, nh, nphil}),
                      arr<Float,3>(a->asymmetry_sw_hydrophilic, {nb, nh, nphil}));
    Aerosol_concs ac;
    for (int i = 0; i < 11; ++i)
    {
        char name[16]; snprintf(name, sizeof name, "aermr%02d", i + 1);
        ac.set_vmr(name, arr<Float,2>(mmr[i], {mmr_ncol[i], mmr_nlay[i]}));
    }
    Optical_props_2str ae(ncol, nlay, ao);
    ao.aerosol_optics(ac, arr<Float,2>(rh, {ncol, nlay}), arr<Float,2>(plev, {ncol, nlay + 1}), ae);
    if (do_delta) ae.delta_scale();
    std::copy(ae.get_tau().v().begin(), ae.get_tau().v().begin() + nbn, atau);
    std::copy(ae.get_ssa().v().begin(), ae.get_ssa().v().begin() + nbn, assa);
    std::copy(ae.get_g().v().begin(), ae.get_g().v().begin() + nbn, ag);
    Optical_props_2str gas(ncol, nlay, *h->kdist);
    gas.get_tau() = std::vector<Float>(tau, tau + n);
    gas.get_ssa() = std::vector<Float>(ssa, ssa + n);
    gas.get_g() = std::vector<Float>(g, g + n);
    add_to(gas, ae);
    std::copy(gas.get_tau().v().begin(), gas.get_tau().v().end(), tau);
    std::copy(gas.get_ssa().v().begin(), gas.get_ssa().v().end(), ssa);
    std::copy(gas.get_g().v().begin(), gas.get_g().v().end(), g);
    SHIM_CATCH
}

/* Fluxes_byband::reduce (src/Fluxes.cpp:142-196); dir may be NULL */
int ref_fluxes_byband(void* hv, int ncol, int nlev, const double* up, const double* dn, const double* dir,
                      double* f_up, double* f_dn, double* f_net, double* f_dir,
                      double* b_up, double* b_dn, double* b_net, double* b_dir)
{
    SHIM_TRY
    auto* h = static_cast<Ref_gas_optics*>(hv);
    const int ngpt = h->kdist->get_ngpt(), nbnd = h->kdist->get_nband();
    std::unique_ptr<Optical_props_arry> op = std::make_unique<Optical_props_1scl>(1, 1, *h->kdist);
    Fluxes_byband fl(ncol, nlev, nbnd);
    auto a_up = arr<Float,3>(up, {ncol, nlev, ngpt}); auto a_dn = arr<Float,3>(dn, {ncol, nlev, ngpt});
    if (dir)
        fl.reduce(a_up, a_dn, arr<Float,3>(dir, {ncol, nlev, ngpt}), op, true);
    else
        fl.reduce(a_up, a_dn, op, true);
    auto cp = [](const auto& a, double* o) { std::copy(a.v().begin(), a.v().end(), o); };
    cp(fl.get_flux_up(), f_up); cp(fl.get_flux_dn(), f_dn); cp(fl.get_flux_net(), f_net);
    cp(fl.get_bnd_flux_up(), b_up); cp(fl.get_bnd_flux_dn(), b_dn); cp(fl.get_bnd_flux_net(), b_net);
    if (dir) { cp(fl.get_flux_dn_dir(), f_dir); cp(fl.get_bnd_flux_dn_dir(), b_dir); }
    SHIM_CATCH
}
}

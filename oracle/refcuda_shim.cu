/*
 * ORACLE / BASELINE — TEST AND BENCH INFRASTRUCTURE ONLY (never on the product path).
 *
 * Flat C entry points over the reference's OWN, UNMODIFIED CUDA implementation of the path
 * (/root/reference/src_cuda/*.cu + src_kernels_cuda/*_launchers.cu, compiled where they lie by
 * oracle/Makefile `make refcuda` for sm_100a into oracle/_ref/librefcuda.so).  It exists so the
 * B200 kernels can be timed beside the reference's GPU path on the same device and inputs
 * (SURVEY.md 8d "Reference CUDA build (secondary)").
 *
 * The driver below issues the same class calls, in the same order and with the same 1024-column
 * blocking, as Radiation_solver_{longwave,shortwave}::solve_gpu (src_test/Radiation_solver.cu:
 * 419-657, 690-950) with switch_fluxes=true, no clouds/aerosols, broadband output only.  That
 * file itself cannot be built here (it needs netCDF), so the call sequence is restated.
 */
#include <cstring>
#include <memory>
#include <string>
#include <vector>
#include <stdexcept>
#include <cuda_runtime.h>

#include <algorithm>
#include <array>
#include <iostream>
#include <fstream>
#include <map>
#include <utility>

#include "Array.h"
#include "Gas_concs.h"
#include "Optical_props.h"
#include "Source_functions.h"
#include "Gas_optics.h"
// The stage entry points below (refcuda_stage_*) read the interpolation state the reference class keeps in
// locals and private members.  Only the access specifiers of this one header are lifted, in this one test
// translation unit; the class itself is compiled unmodified (oracle/Makefile) and its layout does not change.
#define private public
#include "Gas_optics_rrtmgp.h"
#undef private
#include "gas_optics_rrtmgp_kernels_cuda.h"
#include "Rte_lw.h"
#include "Rte_sw.h"
#include "Fluxes.h"
#include "subset_kernels_cuda.h"

#include "rrtmgp_b200_tables.h"
#include "ref_kdist.hpp"

namespace
{
    thread_local std::string last_error;

    struct Ref_cuda
    {
        std::unique_ptr<Gas_optics_rrtmgp_gpu> lw, sw;
        Rte_lw_gpu rte_lw;
        Rte_sw_gpu rte_sw;
    };

    template<int N> Array_gpu<Float, N> up(const double* p, std::array<int, N> dims)
    {
        return Array_gpu<Float, N>(refbuild::arr<Float, N, double>(p, dims));
    }
    template<int N> void down(const Array_gpu<Float, N>& a, double* out)
    {
        if (!out) return;
        Array<Float, N> h(a);
        for (size_t i = 0; i < h.v().size(); ++i) out[i] = (double)h.v()[i];
    }

    template<typename T, int N> void down_raw(const Array_gpu<T, N>& a, T* out)
    {
        if (!out) return;
        Array<T, N> h(a);
        std::copy(h.v().begin(), h.v().end(), out);
    }

    Gas_concs host_concs(int ngas, const char* names, int len, const double* const* vmr, const int* vnc, const int* vnl)
    {
        Gas_concs gc;
        auto nm = refbuild::strings(names, ngas, len, true);
        for (int i = 0; i < ngas; ++i) gc.set_vmr(nm[i], refbuild::arr<Float, 2, double>(vmr[i], {vnc[i], vnl[i]}));
        return gc;
    }

    // one longwave pass over all columns in blocks of 1024 (Radiation_solver.cu:444, 606-657)
    void lw_pass(Ref_cuda& R, const Gas_concs_gpu& gas, const Array_gpu<Float,2>& p_lay, const Array_gpu<Float,2>& p_lev,
                 const Array_gpu<Float,2>& t_lay, const Array_gpu<Float,2>& t_lev, const Array_gpu<Float,2>& col_dry,
                 const Array_gpu<Float,1>& t_sfc, const Array_gpu<Float,2>& emis, Bool top_at_1,
                 Array_gpu<Float,2>& f_up, Array_gpu<Float,2>& f_dn, Array_gpu<Float,2>& f_net)
    {
        const int n_col = p_lay.dim(1), n_lay = p_lay.dim(2), n_lev = p_lev.dim(2);
        const int n_gpt = R.lw->get_ngpt(), n_bnd = R.lw->get_nband();
        constexpr int blk = 1024;
        for (int s = 1; s <= n_col; s += blk)
        {
            const int e = std::min(s + blk - 1, n_col), n = e - s + 1;
            std::unique_ptr<Optical_props_arry_gpu> op = std::make_unique<Optical_props_1scl_gpu>(n, n_lay, *R.lw);
            Source_func_lw_gpu src(n, n_lay, *R.lw);
            Gas_concs_gpu gsub(gas, s, n);
            auto p_lev_s = p_lev.subset({{ {s, e}, {1, n_lev} }});
            Array_gpu<Float,2> cd = col_dry.subset({{ {s, e}, {1, n_lay} }});
            R.lw->gas_optics(p_lay.subset({{ {s, e}, {1, n_lay} }}), p_lev_s, t_lay.subset({{ {s, e}, {1, n_lay} }}),
                             t_sfc.subset({{ {s, e} }}), gsub, op, src, cd, t_lev.subset({{ {s, e}, {1, n_lev} }}));
            Array_gpu<Float,3> g_up({n, n_lev, n_gpt}), g_dn({n, n_lev, n_gpt});
            R.rte_lw.rte_lw(op, top_at_1, src, emis.subset({{ {1, n_bnd}, {s, e} }}), Array_gpu<Float,2>(), g_up, g_dn, 1);
            Fluxes_broadband_gpu fl(n, n_lev);
            fl.reduce(g_up, g_dn, op, top_at_1);
            Subset_kernels_cuda::get_from_subset(n_col, n_lev, n, s, f_up.ptr(), f_dn.ptr(), f_net.ptr(),
                                                 fl.get_flux_up().ptr(), fl.get_flux_dn().ptr(), fl.get_flux_net().ptr());
        }
    }

    // one shortwave pass (Radiation_solver.cu:713, 752-950), tsi scaling = 1
    void sw_pass(Ref_cuda& R, const Gas_concs_gpu& gas, const Array_gpu<Float,2>& p_lay, const Array_gpu<Float,2>& p_lev,
                 const Array_gpu<Float,2>& t_lay, const Array_gpu<Float,2>& col_dry, const Array_gpu<Float,2>& alb_dir,
                 const Array_gpu<Float,2>& alb_dif, const Array_gpu<Float,1>& mu0, Bool top_at_1,
                 Array_gpu<Float,2>& f_up, Array_gpu<Float,2>& f_dn, Array_gpu<Float,2>& f_dir, Array_gpu<Float,2>& f_net)
    {
        const int n_col = p_lay.dim(1), n_lay = p_lay.dim(2), n_lev = p_lev.dim(2);
        const int n_gpt = R.sw->get_ngpt(), n_bnd = R.sw->get_nband();
        constexpr int blk = 1024;
        for (int s = 1; s <= n_col; s += blk)
        {
            const int e = std::min(s + blk - 1, n_col), n = e - s + 1;
            std::unique_ptr<Optical_props_arry_gpu> op = std::make_unique<Optical_props_2str_gpu>(n, n_lay, *R.sw);
            Gas_concs_gpu gsub(gas, s, n);
            auto p_lev_s = p_lev.subset({{ {s, e}, {1, n_lev} }});
            Array_gpu<Float,2> cd = col_dry.subset({{ {s, e}, {1, n_lay} }});
            Array_gpu<Float,2> toa({n, n_gpt});
            R.sw->gas_optics(p_lay.subset({{ {s, e}, {1, n_lay} }}), p_lev_s, t_lay.subset({{ {s, e}, {1, n_lay} }}), gsub, op, toa, cd);
            Array_gpu<Float,3> g_up({n, n_lev, n_gpt}), g_dn({n, n_lev, n_gpt}), g_dir({n, n_lev, n_gpt});
            R.rte_sw.rte_sw(op, top_at_1, mu0.subset({{ {s, e} }}), toa, alb_dir.subset({{ {1, n_bnd}, {s, e} }}),
                            alb_dif.subset({{ {1, n_bnd}, {s, e} }}), Array_gpu<Float,2>(), g_up, g_dn, g_dir);
            Fluxes_broadband_gpu fl(n, n_lev);
            fl.reduce(g_up, g_dn, g_dir, op, top_at_1);
            Subset_kernels_cuda::get_from_subset(n_col, n_lev, n, s, f_up.ptr(), f_dn.ptr(), f_dir.ptr(), f_net.ptr(),
                                                 fl.get_flux_up().ptr(), fl.get_flux_dn().ptr(), fl.get_flux_dn_dir().ptr(),
                                                 fl.get_flux_net().ptr());
        }
    }
}

#define SHIM_TRY try {
#define SHIM_CATCH } catch (const std::exception& e) { last_error = e.what(); return 1; } \
                     catch (...) { last_error = "reference CUDA path threw (tools_gpu.h: printf + throw 1)"; return 1; } return 0;

extern "C"
{
const char* refcuda_last_error() { return last_error.c_str(); }
int refcuda_float_bytes() { return (int)sizeof(Float); }

int refcuda_create(const rrtmgp_b200_kdist_file* lw, const rrtmgp_b200_kdist_file* sw, int n_avail,
                   const char* avail_names, int avail_len, void** out)
{
    SHIM_TRY
    Gas_concs avail;
    for (auto& s : refbuild::strings(avail_names, n_avail, avail_len, true)) avail.set_vmr(s, Float(0.));
    Gas_concs_gpu avail_gpu(avail);
    auto* h = new Ref_cuda;
    if (lw) h->lw = refbuild::build_kdist<Gas_optics_rrtmgp_gpu, Gas_concs_gpu>(lw, avail_gpu);
    if (sw) h->sw = refbuild::build_kdist<Gas_optics_rrtmgp_gpu, Gas_concs_gpu>(sw, avail_gpu);
    *out = h;
    SHIM_CATCH
}
void refcuda_destroy(void* h) { delete static_cast<Ref_cuda*>(h); }

/* Host fp64 arrays in (converted to the build's Float and uploaded before the timed region, like
 * test_rte_rrtmgp.cu does), `reps` timed passes after one untimed pass (kernel autotuning happens
 * there), CUDA-event milliseconds per pass out, last pass's fluxes out as fp64. */
int refcuda_solve(void* hv, int ncol, int nlay, int ngas, const char* names, int len, const double* const* vmr,
                  const int* vnc, const int* vnl, const double* play, const double* plev, const double* tlay,
                  const double* tlev, const double* tsfc, const double* emis, const double* alb_dir,
                  const double* alb_dif, const double* mu0, int reps,
                  double* lw_up, double* lw_dn, double* sw_up, double* sw_dn, double* sw_dir,
                  double* ms_lw, double* ms_sw)
{
    SHIM_TRY
    auto* R = static_cast<Ref_cuda*>(hv);
    const int nlev = nlay + 1;
    Gas_concs gc = host_concs(ngas, names, len, vmr, vnc, vnl);
    Gas_concs_gpu gas(gc);
    auto p_lay = up<2>(play, {ncol, nlay}); auto p_lev = up<2>(plev, {ncol, nlev}); auto t_lay = up<2>(tlay, {ncol, nlay});
    const Bool top_at_1 = play[0] < play[(size_t)ncol * (nlay - 1)];
    Array_gpu<Float,2> col_dry({ncol, nlay});
    Gas_optics_rrtmgp_gpu::get_col_dry(col_dry, gas.get_vmr("h2o"), p_lev);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    if (R->lw)
    {
        auto t_lev = up<2>(tlev, {ncol, nlev}); auto t_sfc = up<1>(tsfc, {ncol});
        auto em = up<2>(emis, {R->lw->get_nband(), ncol});
        Array_gpu<Float,2> f_up({ncol, nlev}), f_dn({ncol, nlev}), f_net({ncol, nlev});
        lw_pass(*R, gas, p_lay, p_lev, t_lay, t_lev, col_dry, t_sfc, em, top_at_1, f_up, f_dn, f_net);
        cudaDeviceSynchronize();
        cudaEventRecord(e0);
        for (int r = 0; r < reps; ++r) lw_pass(*R, gas, p_lay, p_lev, t_lay, t_lev, col_dry, t_sfc, em, top_at_1, f_up, f_dn, f_net);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
        if (ms_lw) *ms_lw = ms / std::max(reps, 1);
        down(f_up, lw_up); down(f_dn, lw_dn);
    }
    if (R->sw)
    {
        auto ad = up<2>(alb_dir, {R->sw->get_nband(), ncol}); auto af = up<2>(alb_dif, {R->sw->get_nband(), ncol});
        auto m0 = up<1>(mu0, {ncol});
        Array_gpu<Float,2> f_up({ncol, nlev}), f_dn({ncol, nlev}), f_dir({ncol, nlev}), f_net({ncol, nlev});
        sw_pass(*R, gas, p_lay, p_lev, t_lay, col_dry, ad, af, m0, top_at_1, f_up, f_dn, f_dir, f_net);
        cudaDeviceSynchronize();
        cudaEventRecord(e0);
        for (int r = 0; r < reps; ++r) sw_pass(*R, gas, p_lay, p_lev, t_lay, col_dry, ad, af, m0, top_at_1, f_up, f_dn, f_dir, f_net);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
        if (ms_sw) *ms_sw = ms / std::max(reps, 1);
        down(f_up, sw_up); down(f_dn, sw_dn); down(f_dir, sw_dir);
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    const cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) throw std::runtime_error(std::string("CUDA: ") + cudaGetErrorString(err));
    SHIM_CATCH
}
/* ---- stage entry points: what the reference's own CUDA kernels compute at every class-API boundary of the
 * path, on host fp64 inputs (converted to the build's Float), returned as fp64 / int32 / int8 host arrays in
 * the reference layouts (column fastest).  tests/test_oracle_vs_refcuda.py compares them with the C
 * restatement of the Fortran kernels (oracle/rrtmgp_kernels_restated.c) — the pin of rows a5-a9. ---- */

/* which: 0 longwave, 1 shortwave.  Outputs (any may be null):
 *   jtemp, jpress (ncol,nlay) int32; tropo (ncol,nlay) int8; jeta (2,ncol,nlay,nflav) int32;
 *   fmajor (2,2,2,ncol,nlay,nflav); fminor (2,2,ncol,nlay,nflav); col_mix (2,ncol,nlay,nflav); col_gas (ncol,nlay,ngas+1);
 *   tau, ssa, g (ncol,nlay,ngpt) [ssa, g shortwave only]; lay_src (ncol,nlay,ngpt), lev_src (ncol,nlay+1,ngpt),
 *   sfc_src, sfc_src_jac (ncol,ngpt) [longwave]; toa_src (ncol,ngpt) [shortwave]; col_dry (ncol,nlay).
 * Call sequence: get_col_dry -> public gas_optics() (src_cuda/Gas_optics_rrtmgp.cu:907-977) for the optical
 * outputs; compute_gas_taus() (:980-1148) once more for the interpolation state it returns by reference; and the
 * interpolation launcher (gas_optics_rrtmgp_kernels_launchers.cu:91-125) on the class's own tables for fminor /
 * col_mix, which compute_gas_taus keeps local.  The launcher's indices must equal compute_gas_taus's (checked). */
int refcuda_stage_gas_optics(void* hv, int which, int ncol, int nlay, int ngas, const char* names, int len,
                             const double* const* vmr, const int* vnc, const int* vnl,
                             const double* play, const double* plev, const double* tlay, const double* tlev, const double* tsfc,
                             int* jtemp_o, int* jpress_o, signed char* tropo_o, int* jeta_o,
                             double* fmajor_o, double* fminor_o, double* col_mix_o, double* col_gas_o,
                             double* tau_o, double* ssa_o, double* g_o,
                             double* lay_src_o, double* lev_src_o, double* sfc_src_o, double* sfc_jac_o, double* toa_o,
                             double* col_dry_o, int* nflav_o, int* ngas_o)
{
    SHIM_TRY
    auto* R = static_cast<Ref_cuda*>(hv);
    Gas_optics_rrtmgp_gpu* K = which == 0 ? R->lw.get() : R->sw.get();
    if (!K) throw std::runtime_error("refcuda_stage_gas_optics: k-distribution not loaded");
    const int nlev = nlay + 1, ngpt = K->get_ngpt(), nband = K->get_nband(), nflav = K->get_nflav();
    // in: the sizes the caller allocated for (negative = unknown); out: the class's own.  A mismatch stops before any write.
    const int nflav_in = nflav_o ? *nflav_o : -1, ngas_in = ngas_o ? *ngas_o : -1;
    if (nflav_o) *nflav_o = nflav;
    if (ngas_o) *ngas_o = K->get_ngas();
    if ((nflav_in >= 0 && nflav_in != nflav) || (ngas_in >= 0 && ngas_in != K->get_ngas()))
        throw std::runtime_error("refcuda_stage_gas_optics: caller's nflav / ngas differ from the reference class's (different available gases?)");
    Gas_concs gc = host_concs(ngas, names, len, vmr, vnc, vnl);
    Gas_concs_gpu gas(gc);
    auto p_lay = up<2>(play, {ncol, nlay}); auto p_lev = up<2>(plev, {ncol, nlev}); auto t_lay = up<2>(tlay, {ncol, nlay});
    Array_gpu<Float,2> col_dry({ncol, nlay});
    Gas_optics_rrtmgp_gpu::get_col_dry(col_dry, gas.get_vmr("h2o"), p_lev);
    down(col_dry, col_dry_o);

    Array_gpu<int,2> jtemp({ncol, nlay}), jpress({ncol, nlay});
    Array_gpu<Bool,2> tropo({ncol, nlay});
    Array_gpu<Float,6> fmajor({2, 2, 2, ncol, nlay, nflav});
    Array_gpu<int,4> jeta({2, ncol, nlay, nflav});
    std::unique_ptr<Optical_props_arry_gpu> op;
    if (which == 0)
    {
        op = std::make_unique<Optical_props_1scl_gpu>(ncol, nlay, *K);
        Source_func_lw_gpu src(ncol, nlay, *K);
        auto t_lev = up<2>(tlev, {ncol, nlev}); auto t_sfc = up<1>(tsfc, {ncol});
        K->gas_optics(p_lay, p_lev, t_lay, t_sfc, gas, op, src, col_dry, t_lev);
        down(op->get_tau(), tau_o);
        down(src.get_lay_source(), lay_src_o); down(src.get_lev_source(), lev_src_o);
        down(src.get_sfc_source(), sfc_src_o); down(src.get_sfc_source_jac(), sfc_jac_o);
    }
    else
    {
        op = std::make_unique<Optical_props_2str_gpu>(ncol, nlay, *K);
        Array_gpu<Float,2> toa({ncol, ngpt});
        K->gas_optics(p_lay, p_lev, t_lay, gas, op, toa, col_dry);
        down(op->get_tau(), tau_o); down(op->get_ssa(), ssa_o); down(op->get_g(), g_o);
        down(toa, toa_o);
    }
    // the interpolation state, as the class computes it for itself
    K->compute_gas_taus(ncol, nlay, ngpt, nband, p_lay, p_lev, t_lay, gas, op, jtemp, jpress, jeta, tropo, fmajor, col_dry);
    down_raw(jtemp, jtemp_o); down_raw(jpress, jpress_o); down_raw(jeta, jeta_o);
    if (tropo_o) { Array<Bool,2> h(tropo); for (size_t i = 0; i < h.v().size(); ++i) tropo_o[i] = (signed char)(h.v()[i] ? 1 : 0); }
    down(fmajor, fmajor_o);

    // fminor / col_mix: the reference's interpolation launcher on the class's own device tables; col_gas as
    // fill_gases_kernel builds it (src_cuda/Gas_optics_rrtmgp.cu:392-422: slot 0 = col_dry, slot i = vmr_i * col_dry,
    // one exactly-rounded product per element, so the host product in Float is the same number)
    {
        const int ng = K->get_ngas();
        Array<Float,2> cd_h(col_dry);
        Array<Float,3> cg_h({ncol, nlay, ng + 1});
        for (int il = 0; il < nlay; ++il) for (int ic = 0; ic < ncol; ++ic) cg_h.v()[ic + (size_t)ncol * il] = cd_h.v()[ic + (size_t)ncol * il];
        for (int ig = 1; ig <= ng; ++ig)
        {
            const Array<Float,2>& v = gc.get_vmr(K->gas_names({ig}));
            const int vc = v.dim(1), vl = v.dim(2);
            for (int il = 0; il < nlay; ++il) for (int ic = 0; ic < ncol; ++ic)
            {
                const Float x = v.v()[(vc > 1 ? ic : 0) + (size_t)vc * (vl > 1 ? il : 0)];
                cg_h.v()[ic + (size_t)ncol * (il + (size_t)nlay * ig)] = x * cd_h.v()[ic + (size_t)ncol * il];
            }
        }
        Array_gpu<Float,3> col_gas(cg_h);
        Array_gpu<Float,4> col_mix({2, ncol, nlay, nflav});
        Array_gpu<Float,5> fminor({2, 2, ncol, nlay, nflav});
        Array_gpu<int,2> jtemp2({ncol, nlay}), jpress2({ncol, nlay});
        Array_gpu<Bool,2> tropo2({ncol, nlay});
        Array_gpu<Float,6> fmajor2({2, 2, 2, ncol, nlay, nflav});
        Array_gpu<int,4> jeta2({2, ncol, nlay, nflav});
        Gas_optics_rrtmgp_kernels_cuda::interpolation(
                ncol, nlay, ng, nflav, K->get_neta(), K->get_npres(), K->get_ntemp(),
                K->flavor_gpu.ptr(), K->press_ref_log_gpu.ptr(), K->temp_ref_gpu.ptr(),
                K->press_ref_log_delta, K->temp_ref_min, K->temp_ref_delta, K->press_ref_trop_log,
                K->vmr_ref_gpu.ptr(), p_lay.ptr(), t_lay.ptr(), col_gas.ptr(),
                jtemp2.ptr(), fmajor2.ptr(), fminor.ptr(), col_mix.ptr(), tropo2.ptr(), jeta2.ptr(), jpress2.ptr());
        cudaDeviceSynchronize();
        Array<int,2> a(jtemp), b(jtemp2); Array<int,4> c(jeta), d(jeta2); Array<Float,6> e(fmajor), f(fmajor2);
        if (a.v() != b.v() || c.v() != d.v() || e.v() != f.v())
            throw std::runtime_error("refcuda_stage_gas_optics: the launcher call and compute_gas_taus disagree (col_gas restatement?)");
        down(fminor, fminor_o); down(col_mix, col_mix_o); down(col_gas, col_gas_o);
    }
    const cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) throw std::runtime_error(std::string("CUDA: ") + cudaGetErrorString(err));
    SHIM_CATCH
}

/* Rte_lw_gpu::rte_lw (src_cuda/Rte_lw.cu:60-159; kernels rte_solver_kernels.cu:35-193), one Gauss angle (the
 * reference GPU launcher throws for more), no incident flux: per-g-point fluxes (ncol,nlay+1,ngpt). */
int refcuda_stage_rte_lw(void* hv, int ncol, int nlay, int top_at_1, const double* tau, const double* lay_src,
                         const double* lev_src, const double* sfc_src, const double* emis, double* gpt_up, double* gpt_dn)
{
    SHIM_TRY
    auto* R = static_cast<Ref_cuda*>(hv);
    if (!R->lw) throw std::runtime_error("refcuda_stage_rte_lw: longwave k-distribution not loaded");
    const int nlev = nlay + 1, ngpt = R->lw->get_ngpt(), nbnd = R->lw->get_nband();
    std::unique_ptr<Optical_props_arry_gpu> op = std::make_unique<Optical_props_1scl_gpu>(ncol, nlay, *R->lw);
    op->get_tau() = up<3>(tau, {ncol, nlay, ngpt});
    Source_func_lw_gpu src(ncol, nlay, *R->lw);
    src.get_lay_source() = up<3>(lay_src, {ncol, nlay, ngpt});
    src.get_lev_source() = up<3>(lev_src, {ncol, nlev, ngpt});
    src.get_sfc_source() = up<2>(sfc_src, {ncol, ngpt});
    src.get_sfc_source_jac().fill(Float(0.));
    auto em = up<2>(emis, {nbnd, ncol});
    Array_gpu<Float,3> g_up({ncol, nlev, ngpt}), g_dn({ncol, nlev, ngpt});
    R->rte_lw.rte_lw(op, (Bool)top_at_1, src, em, Array_gpu<Float,2>(), g_up, g_dn, 1);
    cudaDeviceSynchronize();
    down(g_up, gpt_up); down(g_dn, gpt_dn);
    const cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) throw std::runtime_error(std::string("CUDA: ") + cudaGetErrorString(err));
    SHIM_CATCH
}

/* Rte_sw_gpu::rte_sw (src_cuda/Rte_sw.cu:116-185; kernels rte_solver_kernels.cu:196-286,543-655), no incident
 * diffuse flux: per-g-point up / down / direct fluxes (ncol,nlay+1,ngpt). */
int refcuda_stage_rte_sw(void* hv, int ncol, int nlay, int top_at_1, const double* tau, const double* ssa, const double* g,
                         const double* mu0, const double* toa, const double* alb_dir, const double* alb_dif,
                         double* gpt_up, double* gpt_dn, double* gpt_dir)
{
    SHIM_TRY
    auto* R = static_cast<Ref_cuda*>(hv);
    if (!R->sw) throw std::runtime_error("refcuda_stage_rte_sw: shortwave k-distribution not loaded");
    const int nlev = nlay + 1, ngpt = R->sw->get_ngpt(), nbnd = R->sw->get_nband();
    std::unique_ptr<Optical_props_arry_gpu> op = std::make_unique<Optical_props_2str_gpu>(ncol, nlay, *R->sw);
    op->get_tau() = up<3>(tau, {ncol, nlay, ngpt});
    op->get_ssa() = up<3>(ssa, {ncol, nlay, ngpt});
    op->get_g() = up<3>(g, {ncol, nlay, ngpt});
    auto m0 = up<1>(mu0, {ncol}); auto t = up<2>(toa, {ncol, ngpt});
    auto ad = up<2>(alb_dir, {nbnd, ncol}); auto af = up<2>(alb_dif, {nbnd, ncol});
    Array_gpu<Float,3> g_up({ncol, nlev, ngpt}), g_dn({ncol, nlev, ngpt}), g_dir({ncol, nlev, ngpt});
    R->rte_sw.rte_sw(op, (Bool)top_at_1, m0, t, ad, af, Array_gpu<Float,2>(), g_up, g_dn, g_dir);
    cudaDeviceSynchronize();
    down(g_up, gpt_up); down(g_dn, gpt_dn); down(g_dir, gpt_dir);
    const cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) throw std::runtime_error(std::string("CUDA: ") + cudaGetErrorString(err));
    SHIM_CATCH
}
}

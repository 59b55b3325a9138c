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

#include "Array.h"
#include "Gas_concs.h"
#include "Gas_optics_rrtmgp.h"
#include "Optical_props.h"
#include "Source_functions.h"
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
}

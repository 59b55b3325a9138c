// stages.hpp — launchers of the materialised stage kernels (stages.cu, gasopt.cu).
#pragma once
#include <cuda_runtime.h>
#include "common.cuh"

namespace rb200 {

constexpr int MAX_BANDS = 32;
struct BandLims { int nbnd; int lims[2 * MAX_BANDS]; };   // (2, nbnd) 1-based; nbnd = 0: same resolution

template<typename Real> struct CloudDev
{
    int nbnd, liq_nsteps, ice_nsteps;
    Real liq_step, ice_step, radliq_lwr, diamice_lwr;
    const Real *extliq, *ssaliq, *asyliq, *extice, *ssaice, *asyice;   // [b][nsize]
};

template<typename Real>
void launch_cloud_optics(const CloudDev<Real>& L, bool twostr, int ncol, int c0, int nc, int nlay,
                         const void* lwp, const void* iwp, const void* rel, const void* dei,
                         int nc_out, int c_out0, int delta, void* tau, void* ssa, void* g, cudaStream_t s);
// Aerosol look-up tables, species-ordered [species][humidity class][band]; species in the summation order of
// src/Aerosol_optics.cpp:51 (SS1 SS2 SS3 DU1 DU2 DU3 OM1 OM2 BC1 BC2 SU)
constexpr int AEROSOL_SPECIES = 11;
template<typename Real> struct AerosolDev
{
    int nbnd, nhum;
    const Real *rh_upper, *mext, *ssa, *asy;
};
struct AerosolMmr { const void* p[AEROSOL_SPECIES]; int ncol[AEROSOL_SPECIES]; int nlay[AEROSOL_SPECIES]; };   // in species order
template<typename Real>
void launch_aerosol_optics(const AerosolDev<Real>& A, const AerosolMmr& m, bool increment, int ncol, int c0, int nc, int nlay,
                           const void* rh, const void* plev, int nc_out, int c_out0, int delta,
                           void* tau, void* ssa, void* g, cudaStream_t s);
template<typename Real> void launch_delta_scale(size_t n, void* tau, void* ssa, void* g, cudaStream_t s);
template<typename Real>
void launch_increment(bool twostr, size_t ncl, int ngpt, const BandLims& bl, void* tau1, void* ssa1, void* g1,
                      const void* tau2, const void* ssa2, const void* g2, cudaStream_t s);
template<typename Real> void launch_expand(int ncol, const BandLims& bl, const void* in, void* out, cudaStream_t s);

template<typename Real>
void launch_lw_solver(int ncol, int nlay, int ngpt, int top_at_1, int nmus, const void* tau, const void* lay_src,
                      const void* lev_src, const void* sfc_emis, const void* sfc_src, const void* inc_flux,
                      void* flux_up, void* flux_dn, const void* sfc_src_jac, void* flux_up_jac, cudaStream_t s);
template<typename Real>
void launch_lw_solver_rescl(int ncol, int nlay, int ngpt, int top_at_1, int nmus, const void* tau, const void* ssa,
                            const void* g, const void* lay_src, const void* lev_src, const void* sfc_emis,
                            const void* sfc_src, const void* inc_flux, void* flux_up, void* flux_dn,
                            const void* sfc_src_jac, void* flux_up_jac, void* work2, cudaStream_t s);
template<typename Real>
void launch_sw_solver(int ncol, int nlay, int ngpt, int top_at_1, int mu0_by_layer, const void* tau, const void* ssa,
                      const void* g, const void* mu0, const void* alb_dir, const void* alb_dif, const void* inc_dir,
                      const void* inc_dif, void* flux_up, void* flux_dn, void* flux_dir, void* work4, cudaStream_t s);
template<typename Real>
void launch_sum_broadband(int ncol, int nlev, int ngpt, const void* gpt_flux, void* flux, cudaStream_t s);
template<typename Real>
void launch_sum_byband(int ncol, int nlev, const BandLims& bl, const void* gpt_flux, void* bnd_flux, cudaStream_t s);
template<typename Real> void launch_net(size_t n, const void* dn, const void* up, void* net, cudaStream_t s);
template<typename Real> void launch_scale_columns(int ncol, int n2, void* x, const void* scale, cudaStream_t s);
template<typename Real>
void launch_heating_rate(int ncol, int nlay, const void* up, const void* dn, const void* plev, void* hr, cudaStream_t s);
template<typename Real>
void launch_range_flag(size_t n, const void* x, double lo, double hi, int* d_flag, cudaStream_t s);

// arrays.cu : Array_gpu::fill / Array_gpu::subset, type-erased by element size (false = unsupported size)
struct SubsetGeom
{
    static constexpr int MAXD = 8;
    int ndim;
    int extent[MAXD];          // extents of the window (= the output array)
    int start0[MAXD];          // 0-based first source index of the window per dimension
    size_t stride[MAXD];       // column-major strides of the source
    bool spread[MAXD];         // source extent 1: broadcast
};
bool launch_array_fill(int elem_bytes, void* a, const void* value, size_t n, cudaStream_t s);
bool launch_array_subset(int elem_bytes, void* out, const void* in, const SubsetGeom& g, size_t n, cudaStream_t s);

// gasopt.cu : materialised gas optics of one band from a TileState (c0 = 0, nc = ncol)
template<typename Real> struct GasOut
{
    Real* tau; Real* ssa; Real* g;                       // (ncol, nlay, ngpt)
    Real* lay_src; Real* lev_src; Real* sfc_src; Real* sfc_src_jac;
    Real* toa_src;                                       // (ncol, ngpt)
    const Real* tlay; const Real* tlev; const Real* tsfc;
    int sfc_lay;                                         // 0-based surface layer
};
template<typename Real>
void launch_gas_band(const DevPlan<Real>& P, const BandArgs& B, const TileState<Real>& st, int ncol, int nlay,
                     const GasOut<Real>& o, cudaStream_t s);

}  // namespace rb200

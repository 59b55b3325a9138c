// launch.hpp — host-callable launchers implemented in the .cu translation units.
#pragma once
#include <cuda_runtime.h>
#include "common.cuh"

namespace rb200 {

void count_launch();          // bumps the process-wide kernel-launch counter (cabi.cu)

// ---- k1_interp.cu
template<typename Real>
void launch_col_dry(int ncol, int nlay, const void* vmr_h2o, const void* plev, void* col_dry, cudaStream_t s);
template<typename Real>
void launch_interp(const DevPlan<Real>& P, const VmrArgs& vmr, int ncol, int c0, int nc, int nlay,
                   const void* play, const void* tlay, const void* col_dry, const void* tlev,
                   const TileState<Real>& st, int* jtemp, int* jpress, signed char* tropo, int* jeta, void* col_mix,
                   void* fminor, void* fmajor, void* col_gas, cudaStream_t s);

// ---- fused_lw.cu / fused_sw.cu : one launch = one band of one column tile
template<typename Real> struct FusedArgs
{
    int ncol, c0, nc, nlay, top_at_1;      // caller stride, tile offset, tile columns
    const Real* tlay; const Real* tlev; const Real* tsfc;         // LW
    const Real* sfc_emis;                                           // LW (nbnd, ncol)
    const Real* mu0; const Real* alb_dir; const Real* alb_dif;     // SW (ncol), (nbnd, ncol)
    const Real* tsi_scaling;                                        // SW (ncol) or null
    const Real* cld_tau; const Real* cld_ssa; const Real* cld_g;   // tile-local (nc,nlay,nbnd) or null
    V4<Real>* scratch;                                              // recurrence state, one strip per slot
    int* slot_busy; int nslots;                                     // strips are claimed by resident CTAs
    Real* flux_up; Real* flux_dn; Real* flux_dir;                   // (ncol, nlay+1), accumulated
    Real* bnd_up; Real* bnd_dn; Real* bnd_dir;                      // (ncol, nlay+1, nbnd) or null
    Real secant, weight;                                            // LW quadrature angle of this launch
    int bnd_add;                                                    // LW: add this angle onto the by-band outputs
};
constexpr int FUSED_THREADS = 128;
template<typename Real> int lw_blocks_per_wave(int gw);
template<typename Real> int sw_blocks_per_wave(int gw);
template<typename Real> size_t fused_scratch_bytes(int nslots, int nlay, bool sw);
template<typename Real> void launch_lw_band(const DevPlan<Real>& P, const BandArgs& B, const TileState<Real>& st,
                                            const FusedArgs<Real>& a, cudaStream_t s);
template<typename Real> void launch_sw_band(const DevPlan<Real>& P, const BandArgs& B, const TileState<Real>& st,
                                            const FusedArgs<Real>& a, cudaStream_t s);

// ---- colres.cu : one launch = all bands of one column tile, recurrences resident in shared memory
template<typename Real> size_t colres_smem_bytes(int nlay, int gw, bool sw);
template<typename Real> void launch_cols(bool sw, const DevPlan<Real>& P, const BandArgs* bands_dev, const TileState<Real>& st,
                                         const FusedArgs<Real>& a, cudaStream_t s);

}  // namespace rb200

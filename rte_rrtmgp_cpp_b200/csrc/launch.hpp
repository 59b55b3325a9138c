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
    Real* part; int nsplit;                                         // column-resident kernels, bands split over CTAs: partial sums
    Real secants[4], weights[4];                                    // LW, all angles in one launch (fused_lw_ma.cu)
};
constexpr int FUSED_THREADS = 128;

// Loop-invariant addresses and strides of one band launch, computed on the host and handed over in the constant bank:
// at 96 registers ptxas rematerialises such values inside the layer loop, and a value that is ONE constant load costs one
// instruction there instead of the 64-bit multiply chain it was derived with (~10 % of the longwave kernel's instructions).
template<typename Real> struct BandLaunch
{
    BandArgs B;
    const Real* kmaj;            // the band's major (+ interleaved Planck-fraction) block
    const Real* kmin[2];         // the band's first minor contributor block, per side
    const Real* kray[2];         // SW: the band's Rayleigh block, per side (null without Rayleigh)
    const Real* plk;             // LW: the band's row of totplnk
    size_t mstride;              // elements per minor contributor block
    size_t tcl;                  // tile columns * layers
    size_t flv[2];               // V4 offset of the band's flavor plane, per side
    size_t scal0[2], scal1[2];   // V4 offset of the band's first / second scaling group plane, per side
    size_t strip;                // scratch vectors per strip (nlay * words * threads)
    size_t nlev_stride;          // ncol * (nlay + 1)
    size_t o_top;                // ncol * (level index of the top of the atmosphere)
    long lstep;                  // +-ncol: one level down in the caller's arrays
};
template<typename Real>
inline BandLaunch<Real> make_band_launch(const DevPlan<Real>& P, const BandArgs& B, const FusedArgs<Real>& a, int flav0, int flav1,
                                         bool lw, int words)
{
    BandLaunch<Real> L{};
    L.B = B;
    const size_t rs = (size_t)(lw ? 2 : 1) * P.gw;
    L.kmaj = P.kmajor + (size_t)P.ntemp * (P.npres + 1) * P.neta * rs * B.band;
    L.mstride = (size_t)P.ntemp * P.neta * P.gw;
    for (int s = 0; s < 2; ++s) L.kmin[s] = P.kminor[s] + L.mstride * B.first_minor[s];
    const size_t ray_elems = (size_t)P.ntemp * P.neta * P.gw;
    for (int s = 0; s < 2; ++s) L.kray[s] = P.krayl ? P.krayl + ray_elems * ((size_t)s * P.nbnd + B.band) : nullptr;
    L.plk = lw ? P.totplnk + (size_t)P.nplanck * B.band : nullptr;
    L.tcl = (size_t)a.nc * a.nlay;
    L.flv[0] = L.tcl * flav0; L.flv[1] = L.tcl * flav1;
    for (int s = 0; s < 2; ++s) { L.scal0[s] = L.tcl * B.scal_off[s]; L.scal1[s] = L.tcl * (B.scal_off[s] + (B.nminor[s] > 4 ? 1 : 0)); }
    L.strip = (size_t)a.nlay * words * FUSED_THREADS;
    L.nlev_stride = (size_t)a.ncol * (a.nlay + 1);
    L.o_top = a.top_at_1 ? 0 : (size_t)a.ncol * a.nlay;
    L.lstep = a.top_at_1 ? (long)a.ncol : -(long)a.ncol;
    return L;
}
template<typename Real> int lw_blocks_per_wave(int gw);
template<typename Real> int sw_blocks_per_wave(int gw);
template<typename Real> size_t fused_scratch_bytes(int nslots, int nlay, int words);     // words per layer: 6 SW, 2 per angle LW
template<typename Real> void launch_lw_band(const DevPlan<Real>& P, const BandArgs& B, const TileState<Real>& st,
                                            const FusedArgs<Real>& a, cudaStream_t s);
template<typename Real> void launch_sw_band(const DevPlan<Real>& P, const BandArgs& B, const TileState<Real>& st,
                                            const FusedArgs<Real>& a, cudaStream_t s);

// ---- fused_lw_ma.cu : longwave band launch with the 2..4 Gauss angles inside the layer loop (16-wide bands, 9 mixing fractions)
template<typename Real> bool lw_multi_angle_supported(const DevPlan<Real>& P, const BandArgs& B);
template<typename Real> void launch_lw_band_angles(const DevPlan<Real>& P, const BandArgs& B, const TileState<Real>& st,
                                                   const FusedArgs<Real>& a, int nang, cudaStream_t s);

// ---- fused_pair.cu : one launch = one longwave band + one shortwave band of the same tile, CTAs of both kinds co-resident
template<typename Real> bool pair_supported(const DevPlan<Real>& Plw, const BandArgs& Blw, const DevPlan<Real>& Psw, const BandArgs& Bsw);
template<typename Real> void launch_lwsw_band(const DevPlan<Real>& Plw, const BandArgs& Blw, const TileState<Real>& stlw, const FusedArgs<Real>& alw,
                                              const DevPlan<Real>& Psw, const BandArgs& Bsw, const TileState<Real>& stsw, const FusedArgs<Real>& asw,
                                              cudaStream_t s);
template<typename Real> void launch_role_band(bool shortwave, const DevPlan<Real>& P, const BandArgs& B, const TileState<Real>& st,
                                              const FusedArgs<Real>& a, cudaStream_t s);
template<typename Real> int pair_blocks_per_wave();

// ---- colres.cu : one launch = all bands of one column tile, recurrences resident in shared memory
template<typename Real> size_t colres_smem_bytes(int nlay, int gw, bool sw);
int colres_band_split(int ncol, int gw, int nbnd);       // nbnd when a call of ncol columns should spread its bands over CTAs, else 1
template<typename Real> void launch_cols(bool sw, const DevPlan<Real>& P, const BandArgs* bands_dev, const TileState<Real>& st,
                                         const FusedArgs<Real>& a, cudaStream_t s);

}  // namespace rb200

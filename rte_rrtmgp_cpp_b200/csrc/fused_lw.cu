// fused_lw.cu — launchers and occupancy queries of the longwave band kernel (kernel: fused_lw.cuh).
#include "fused_lw.cuh"

namespace rb200 {

template<typename Real> size_t fused_scratch_bytes(int nslots, int nlay, int words)
{
    return (size_t)nslots * nlay * words * FUSED_THREADS * sizeof(V4<Real>);
}

template<typename Real>
void launch_lw_band(const DevPlan<Real>& P, const BandArgs& B, const TileState<Real>& st,
                    const FusedArgs<Real>& a, cudaStream_t s)
{
    // lanes per column follow the BAND's width (a 16-wide table layout still runs its <= 8-g-point bands on
    // two lanes: the reduced LW set mixes 6..12 g-points per band)
    const int lpc = (P.gw == 16 && B.ngb <= 8) ? 2 : P.gw / 4, cpb = FUSED_THREADS / lpc;
    const unsigned grid = (unsigned)((a.nc + cpb - 1) / cpb);
    const bool cloud = a.cld_tau != nullptr;
    // fp32: the common variant (16-wide bands, 9 mixing fractions) runs in the kernel it shares with the paired longwave +
    // shortwave launches (fused_pair.cu): same machine code whether the two are launched together or apart
    if (sizeof(Real) == 4 && P.gw == 16 && lpc == 4 && P.neta == 9) { launch_role_band<Real>(false, P, B, st, a, s); return; }
    const BandLaunch<Real> L = make_band_launch<Real>(P, B, a, B.flav[0], B.flav[1], true, 2);
#define RB_LAUNCH(LPC_, GW_, CL_, NE_) lw_band_kernel<Real, LPC_, GW_, CL_, NE_><<<grid, FUSED_THREADS, 0, s>>>(P, L, st, a)
#define RB_NE(LPC_, GW_, CL_) { if (P.neta == 9) RB_LAUNCH(LPC_, GW_, CL_, 9); else RB_LAUNCH(LPC_, GW_, CL_, 0); }
#define RB_CL(LPC_, GW_) { if (cloud) RB_NE(LPC_, GW_, true) else RB_NE(LPC_, GW_, false) }
    if (P.gw == 16) { if (lpc == 4) RB_CL(4, 16) else RB_CL(2, 16) }
    else RB_CL(2, 8)
#undef RB_CL
#undef RB_NE
#undef RB_LAUNCH
    count_launch();
}

template size_t fused_scratch_bytes<float>(int, int, int);
template size_t fused_scratch_bytes<double>(int, int, int);
template void launch_lw_band<float>(const DevPlan<float>&, const BandArgs&, const TileState<float>&, const FusedArgs<float>&, cudaStream_t);
template void launch_lw_band<double>(const DevPlan<double>&, const BandArgs&, const TileState<double>&, const FusedArgs<double>&, cudaStream_t);

template<typename Real> int lw_blocks_per_wave(int gw)
{
    int dev = 0, sms = 0, occ = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    static int cache[64][2] = {};                       // per device, per table width: asked on every call of the C ABI
    int* slot = dev >= 0 && dev < 64 ? &cache[dev][gw == 16] : nullptr;
    if (slot && *slot > 0) return *slot;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (gw == 16 && sizeof(Real) == 4) return pair_blocks_per_wave<Real>();
    if (gw == 16) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, lw_band_kernel<Real, 4, 16, false, 9>, FUSED_THREADS, 0);
    else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, lw_band_kernel<Real, 2, 8, false, 9>, FUSED_THREADS, 0);
    if (slot) *slot = sms * occ;
    return sms * occ;
}
template int lw_blocks_per_wave<float>(int);
template int lw_blocks_per_wave<double>(int);

}  // namespace rb200

// fused_pair.cu — one launch = one LONGWAVE band and one SHORTWAVE band of the same column tile, their CTAs interleaved so
// both kinds are resident on every SM at the same time.
//
// Why: the two band kernels saturate different units (ncu, fp32, 1024^2 x 128, profiles/ncu_{lw,sw}_r02d.txt + L1 metrics in
// DESIGN.md): longwave = L1TEX LSU data pipe 75 % / DRAM 43 %; shortwave = L1 53 % / DRAM 67 %.  Run back to back each leaves
// the other's unit half idle; launched on two streams they do not mix either (the block scheduler drains the first grid
// before it starts the second).  Here the role of a CTA follows from its index: even = longwave CTA bid/2, odd = shortwave
// CTA bid/2 (both grids have the same number of CTAs: same tile, same 32 columns per CTA).  The bodies are the single-role
// kernels' (fused_lw.cuh / fused_sw.cuh), unchanged: separate scratch strips, separate flux slabs, so every flux element is
// still touched by one thread per launch and the results are bit-identical to the single-role launches.
#include "fused_lw.cuh"
#include "fused_sw.cuh"

namespace rb200 {

#ifndef RB_PAIR_MINB
#define RB_PAIR_MINB 5
#endif
#ifndef RB_PAIR_MINB64
#define RB_PAIR_MINB64 3
#endif
template<typename Real> struct PairMinBlocks { static constexpr int v = sizeof(Real) == 4 ? RB_PAIR_MINB : RB_PAIR_MINB64; };

template<typename Real> struct PairSide
{
    DevPlan<Real> P; BandLaunch<Real> L; TileState<Real> st; FusedArgs<Real> a;
};

// CLOUD: both roles with (true) or without (false) the by-band particulate increment — one instantiation serves the paired
// launch and both single-role launches of a call, so all three execute the same machine code per CTA.  (A run-time flag
// instead measured 4-6 % slower.)
template<typename Real, bool CLOUD>
__global__ void __launch_bounds__(FUSED_THREADS, PairMinBlocks<Real>::v)
lwsw_band_kernel(const __grid_constant__ PairSide<Real> lw, const __grid_constant__ PairSide<Real> sw, const int nblk_lw, const int nblk_sw)
{
    // interleave while both roles have CTAs left, then the rest of the longer grid
    const int b = (int)blockIdx.x, both = 2 * min(nblk_lw, nblk_sw);
    int bid; bool is_lw;
    if (b < both) { is_lw = (b & 1) == 0; bid = b >> 1; }
    else { is_lw = nblk_lw > nblk_sw; bid = b - both + min(nblk_lw, nblk_sw); }
    if (is_lw) lw_band_body<Real, 4, 16, CLOUD, 9>(lw.P, lw.L, lw.st, lw.a, bid);
    else sw_band_body<Real, 4, 16, CLOUD, 9>(sw.P, sw.L, sw.st, sw.a, bid);
}

// true when the paired kernel covers this pair of bands: fp32 (the fp64 build measured slower paired: 3 resident CTAs per
// SM), full-width bands of 16-wide tables, 9 mixing fractions
template<typename Real>
bool pair_supported(const DevPlan<Real>& Plw, const BandArgs& Blw, const DevPlan<Real>& Psw, const BandArgs& Bsw)
{
    return sizeof(Real) == 4 && Plw.gw == 16 && Psw.gw == 16 && Blw.ngb > 8 && Bsw.ngb > 8 && Plw.neta == 9 && Psw.neta == 9;
}

template<typename Real>
void launch_lwsw_band(const DevPlan<Real>& Plw, const BandArgs& Blw, const TileState<Real>& stlw, const FusedArgs<Real>& alw,
                      const DevPlan<Real>& Psw, const BandArgs& Bsw, const TileState<Real>& stsw, const FusedArgs<Real>& asw, cudaStream_t s)
{
    constexpr int cpb = FUSED_THREADS / 4;
    const int nl = (alw.nc + cpb - 1) / cpb, ns = (asw.nc + cpb - 1) / cpb;
    PairSide<Real> lw{Plw, make_band_launch<Real>(Plw, Blw, alw, Blw.flav[0], Blw.flav[1], true, 2), stlw, alw};
    PairSide<Real> sw{Psw, make_band_launch<Real>(Psw, Bsw, asw, Bsw.flav[0], Bsw.flav[1], false, 6), stsw, asw};
    const unsigned grid = (unsigned)(nl + ns);
    // the caller pairs only calls with the same particulate flag (cabi.cu: fused_pair)
    if (alw.cld_tau != nullptr) lwsw_band_kernel<Real, true><<<grid, FUSED_THREADS, 0, s>>>(lw, sw, nl, ns);
    else lwsw_band_kernel<Real, false><<<grid, FUSED_THREADS, 0, s>>>(lw, sw, nl, ns);
    count_launch();
}

// One role alone through the SAME compiled kernel (the other role's grid is empty): the single-role calls and the paired
// call then execute identical machine code per CTA, so their results are bit-identical by construction (two separately
// compiled kernels of the same source differed in the last bit in fp64: the compiler contracts multiply-adds per kernel).
template<typename Real>
void launch_role_band(bool shortwave, const DevPlan<Real>& P, const BandArgs& B, const TileState<Real>& st, const FusedArgs<Real>& a, cudaStream_t s)
{
    constexpr int cpb = FUSED_THREADS / 4;
    const int n = (a.nc + cpb - 1) / cpb;
    PairSide<Real> lw{}, sw{};
    if (shortwave) sw = PairSide<Real>{P, make_band_launch<Real>(P, B, a, B.flav[0], B.flav[1], false, 6), st, a};
    else lw = PairSide<Real>{P, make_band_launch<Real>(P, B, a, B.flav[0], B.flav[1], true, 2), st, a};
    const int nl = shortwave ? 0 : n, ns = shortwave ? n : 0;
    if (a.cld_tau != nullptr) lwsw_band_kernel<Real, true><<<(unsigned)n, FUSED_THREADS, 0, s>>>(lw, sw, nl, ns);
    else lwsw_band_kernel<Real, false><<<(unsigned)n, FUSED_THREADS, 0, s>>>(lw, sw, nl, ns);
    count_launch();
}

template<typename Real> int pair_blocks_per_wave()
{
    int dev = 0, sms = 0, occ = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    static int cache[64] = {};                            // per device
    if (dev >= 0 && dev < 64 && cache[dev] > 0) return cache[dev];
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, lwsw_band_kernel<Real, false>, FUSED_THREADS, 0);
    if (dev >= 0 && dev < 64) cache[dev] = sms * occ;
    return sms * occ;
}

#define INST(Real) \
    template bool pair_supported<Real>(const DevPlan<Real>&, const BandArgs&, const DevPlan<Real>&, const BandArgs&); \
    template void launch_lwsw_band<Real>(const DevPlan<Real>&, const BandArgs&, const TileState<Real>&, const FusedArgs<Real>&, \
                                         const DevPlan<Real>&, const BandArgs&, const TileState<Real>&, const FusedArgs<Real>&, cudaStream_t); \
    template void launch_role_band<Real>(bool, const DevPlan<Real>&, const BandArgs&, const TileState<Real>&, const FusedArgs<Real>&, cudaStream_t); \
    template int pair_blocks_per_wave<Real>();
INST(float)
INST(double)
#undef INST

}  // namespace rb200

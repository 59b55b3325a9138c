// fused_lw_ma.cu — the longwave band kernel with all Gauss angles of the quadrature in one launch (kernel body: fused_lw.cuh
// with NANG > 1): gas optics and Planck sources once per layer, transmissivity, layer sources and transport per angle.
// The reference sums the angles in its CPU solver only (src/Rte_lw.cpp:139-151, mo_rte_solver_kernels.F90:250-321); its
// CUDA path has one angle.  Instantiated for the common shape (16-wide bands, 9 mixing fractions); other bands run one
// launch of the single-angle kernel per angle (cabi.cu).
#include "fused_lw.cuh"

namespace rb200 {

// 4 resident CTAs (128 registers) in fp32, 2 (255) in fp64: the per-angle radiances are live through the layer loop
template<typename Real> struct LWAnglesMinBlocks { static constexpr int v = sizeof(Real) == 4 ? 4 : 2; };

template<typename Real, bool CLOUD, int NANG>
__global__ void __launch_bounds__(FUSED_THREADS, LWAnglesMinBlocks<Real>::v)
lw_band_angles_kernel(const DevPlan<Real> P, const __grid_constant__ BandLaunch<Real> L, const TileState<Real> st,
                      const FusedArgs<Real> a)
{
    lw_band_body<Real, 4, 16, CLOUD, 9, NANG>(P, L, st, a, (int)blockIdx.x);
}

template<typename Real> bool lw_multi_angle_supported(const DevPlan<Real>& P, const BandArgs& B)
{
    return P.gw == 16 && P.neta == 9 && B.ngb > 8;
}

template<typename Real>
void launch_lw_band_angles(const DevPlan<Real>& P, const BandArgs& B, const TileState<Real>& st, const FusedArgs<Real>& a, int nang,
                           cudaStream_t s)
{
    constexpr int cpb = FUSED_THREADS / 4;
    const unsigned grid = (unsigned)((a.nc + cpb - 1) / cpb);
    const bool cloud = a.cld_tau != nullptr;
    const BandLaunch<Real> L = make_band_launch<Real>(P, B, a, B.flav[0], B.flav[1], true, 2 * nang);
#define RB_GO(CL_, N_) lw_band_angles_kernel<Real, CL_, N_><<<grid, FUSED_THREADS, 0, s>>>(P, L, st, a)
#define RB_N(CL_) { if (nang == 2) RB_GO(CL_, 2); else if (nang == 3) RB_GO(CL_, 3); else RB_GO(CL_, 4); }
    if (cloud) RB_N(true) else RB_N(false)
#undef RB_N
#undef RB_GO
    count_launch();
}

template bool lw_multi_angle_supported<float>(const DevPlan<float>&, const BandArgs&);
template bool lw_multi_angle_supported<double>(const DevPlan<double>&, const BandArgs&);
template void launch_lw_band_angles<float>(const DevPlan<float>&, const BandArgs&, const TileState<float>&, const FusedArgs<float>&, int, cudaStream_t);
template void launch_lw_band_angles<double>(const DevPlan<double>&, const BandArgs&, const TileState<double>&, const FusedArgs<double>&, int, cudaStream_t);

}  // namespace rb200

// gasopt.cu — materialised gas optics: Gas_optics_rrtmgp::gas_optics LW / SW filling the
// reference-layout arrays tau[,ssa,g], lay_src, lev_src, sfc_src, sfc_src_jac, toa_src
// (src/Gas_optics_rrtmgp.cpp:796-898, 1090-1344; kernels mo_gas_optics_rrtmgp_kernels.F90:178-710).
// Same (column, 4 g-points) thread mapping and the same device functions as the fused kernels;
// a warp's stores cover 32/LPC consecutive columns per g-point (full 32-byte sectors in fp32).
#include "physics.cuh"
#include "launch.hpp"
#include "stages.hpp"

namespace rb200 {

template<typename Real, int LPC, bool LW>
__global__ void __launch_bounds__(FUSED_THREADS)
gas_band_kernel(const DevPlan<Real> P, const __grid_constant__ BandArgs B, const TileState<Real> st,
                int ncol, int nlay, const GasOut<Real> o)
{
    using G = Geo<4 * LPC, 0, LW>;
    constexpr int CPB = FUSED_THREADS / LPC;
    const int tid = threadIdx.x;
    const int q = tid % LPC, cl = tid / LPC;
    const int c = blockIdx.x * CPB + cl;
    if (c >= ncol) return;                       // no warp-collective operations in this kernel
    const size_t ncl = (size_t)ncol * nlay, ncv = (size_t)ncol * (nlay + 1);
    const int q4 = q * 4;
    const int band = B.band;
    const size_t band_elems = (size_t)P.ntemp * (P.npres + 1) * P.neta * G::RS;
    const size_t ray_elems = (size_t)P.ntemp * P.neta * G::GW;
    const Real* kmaj = P.kmajor + band_elems * band + q4;
    const int flav[2] = {P.band_flavor[2 * band], P.band_flavor[2 * band + 1]};
    constexpr bool lw = LW;
    const Real* plk = lw ? P.totplnk + (size_t)P.nplanck * band : nullptr;
    int nvalid = B.ngb - q4; nvalid = nvalid < 0 ? 0 : (nvalid > 4 ? 4 : nvalid);
    const int g0 = B.gpt0 + q4;

    Real pf_prev[4] = {0, 0, 0, 0};
    for (int l = 0; l < nlay; ++l)
    {
        const size_t icl = (size_t)c + (size_t)ncol * l;
        const V4<Real> cell = st.cell[icl];
        const int f = flav[cell_side(cell)];
        Cell<Real> ce;
        decode_cell(cell, st.flvA[icl + ncl * f], st.flvB[icl + ncl * f], ce);
        Real tau[4], s0[4], s1[4];
        V4<Real> sc0, sc1;
        load_scal4(st.scal + icl, ncl, B, ce.side, 0, sc0);
        load_scal4(st.scal + icl, ncl, B, ce.side, 4, sc1);
        gas_tau<G>(P, B, ce, kmaj, q4, sc0, sc1, st.scal + icl, ncl, tau);
        if (lw)
        {
            Real pf[4];
            interp_major<G>(P.pfrac + band_elems * band + q4, P, ce, s0, s1);
            const Real blay = planck_band(plk, P.nplanck, P.planck_tmin, P.planck_delta, o.tlay[icl]);
            const Real blev = planck_band(plk, P.nplanck, P.planck_tmin, P.planck_delta, o.tlev[icl]);
            for (int j = 0; j < nvalid; ++j)
            {
                pf[j] = s0[j] + s1[j];
                const size_t og = ncl * (g0 + j), ov = ncv * (g0 + j);
                o.tau[icl + og] = tau[j];
                o.lay_src[icl + og] = pf[j] * blay;
                o.lev_src[icl + ov] = (l == 0 ? pf[j] : M<Real>::sqrt(pf_prev[j] * pf[j])) * blev;
                if (l == nlay - 1)
                {
                    const Real btop = planck_band(plk, P.nplanck, P.planck_tmin, P.planck_delta, o.tlev[icl + ncol]);
                    o.lev_src[icl + ncol + ov] = pf[j] * btop;
                }
                if (l == o.sfc_lay)
                {
                    const Real b0 = planck_band(plk, P.nplanck, P.planck_tmin, P.planck_delta, o.tsfc[c]);
                    const Real b1 = planck_band(plk, P.nplanck, P.planck_tmin, P.planck_delta, o.tsfc[c] + (Real)1);
                    o.sfc_src[c + (size_t)ncol * (g0 + j)] = pf[j] * b0;
                    o.sfc_src_jac[c + (size_t)ncol * (g0 + j)] = pf[j] * (b1 - b0);
                }
                pf_prev[j] = pf[j];
            }
        }
        else
        {
            Real ray[4] = {0, 0, 0, 0};
            if (P.krayl)
            {
                interp_minor<G>(P.krayl + ray_elems * ((size_t)ce.side * P.nbnd + band) + q4, ce, ray);
#pragma unroll
                for (int j = 0; j < 4; ++j) ray[j] *= ce.wet;
            }
            for (int j = 0; j < nvalid; ++j)
            {
                const size_t og = ncl * (g0 + j);
                if (o.ssa)
                {
                    Real t, w;
                    combine_rayleigh(tau[j], ray[j], t, w);
                    o.tau[icl + og] = t; o.ssa[icl + og] = w; o.g[icl + og] = (Real)0;
                }
                else o.tau[icl + og] = tau[j];
                if (l == 0 && o.toa_src) o.toa_src[c + (size_t)ncol * (g0 + j)] = P.solar_source[g0 + j];
            }
        }
    }
}

template<typename Real>
void launch_gas_band(const DevPlan<Real>& P, const BandArgs& B, const TileState<Real>& st, int ncol, int nlay,
                     const GasOut<Real>& o, cudaStream_t s)
{
    const int lpc = P.gw / 4, cpb = FUSED_THREADS / lpc;
    const unsigned grid = (unsigned)((ncol + cpb - 1) / cpb);
    const bool lw = o.lay_src != nullptr;
    if (lpc == 4) { if (lw) gas_band_kernel<Real, 4, true><<<grid, FUSED_THREADS, 0, s>>>(P, B, st, ncol, nlay, o);
                    else gas_band_kernel<Real, 4, false><<<grid, FUSED_THREADS, 0, s>>>(P, B, st, ncol, nlay, o); }
    else { if (lw) gas_band_kernel<Real, 2, true><<<grid, FUSED_THREADS, 0, s>>>(P, B, st, ncol, nlay, o);
           else gas_band_kernel<Real, 2, false><<<grid, FUSED_THREADS, 0, s>>>(P, B, st, ncol, nlay, o); }
    count_launch();
}

template void launch_gas_band<float>(const DevPlan<float>&, const BandArgs&, const TileState<float>&, int, int, const GasOut<float>&, cudaStream_t);
template void launch_gas_band<double>(const DevPlan<double>&, const BandArgs&, const TileState<double>&, int, int, const GasOut<double>&, cudaStream_t);

}  // namespace rb200

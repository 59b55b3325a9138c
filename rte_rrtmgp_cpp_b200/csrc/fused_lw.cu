// fused_lw.cu — longwave: gas optics + Planck sources + no-scattering solver + g-point reduction
// in one kernel per (column tile, band).  tau / lay_src / lev_src / gpt_flux never exist in HBM.
//
// Restates (fp order aside) the chain the oracle pins:
//   compute_tau_absorption   mo_gas_optics_rrtmgp_kernels.F90:178-502
//   compute_Planck_source    mo_gas_optics_rrtmgp_kernels.F90:567-710
//   inc_1scalar_by_1scalar_bybnd (clouds) mo_optical_props_kernels.F90:366-382
//   lw_solver_noscat_oneangle mo_rte_solver_kernels.F90:51-240 (+ :620-745)
//   sum_broadband / sum_byband mo_fluxes_broadband_kernels.F90:32-61, mo_fluxes_byband_kernels.F90:22-41
//
// Mapping: a thread owns (column, 4 consecutive g-points); LPC lanes cover a band of gw = 4*LPC
// g-points, a warp covers 32/LPC columns, so table rows are read as 64-byte (fp32) contiguous
// segments and per-column state as 128-byte coalesced vectors.  The thread walks its column top
// to bottom (down-welling recurrence in registers), parks {trans, source_up} of each layer in a
// per-CTA scratch strip (coalesced 16/32-byte vectors, consumed LIFO by the up-sweep, so the
// hot end stays in L2), then walks back up.  g-point sums are 4-wide in-thread adds followed by
// log2(LPC) warp shuffles; one lane per column adds the band total into the broadband flux.
#include "physics.cuh"
#include "launch.hpp"

namespace rb200 {

template<typename Real, int LPC, bool CLOUD>
__global__ void __launch_bounds__(FUSED_THREADS)
lw_band_kernel(const DevPlan<Real> P, const __grid_constant__ BandArgs B, const TileState<Real> st,
               const FusedArgs<Real> a)
{
    constexpr int CPB = FUSED_THREADS / LPC;
    const int tid = threadIdx.x;
    const int q = tid % LPC, cl = tid / LPC;
    const int craw = blockIdx.x * CPB + cl;
    const bool active = craw < a.nc;
    const int c = active ? craw : a.nc - 1;
    const int nlay = a.nlay, nc = a.nc;
    const size_t tcl = (size_t)nc * nlay;
    const size_t cg = (size_t)a.c0 + c;                 // column in the caller's arrays
    const int q4 = q * 4;
    const int band = B.band;
    const size_t band_elems = (size_t)P.ntemp * (P.npres + 1) * P.neta * P.gw;
    const Real* kmaj = P.kmajor + band_elems * band;
    const Real* pfr = P.pfrac + band_elems * band + q4;
    const Real* plk = P.totplnk + (size_t)P.nplanck * band;
    const int flav[2] = {P.band_flavor[2 * band], P.band_flavor[2 * band + 1]};
    V4<Real>* scr = a.scratch + (size_t)blockIdx.x * nlay * 2 * FUSED_THREADS + tid;
    const Real piw = (Real)3.14159265358979323846 * a.weight;
    const bool lead = active && q == 0;
    const size_t nlev_stride = (size_t)a.ncol * (nlay + 1);

    Real radn[4] = {0, 0, 0, 0};        // incident diffuse flux = 0 (Radiation_solver.cu passes zeros)
    Real pf_prev[4] = {0, 0, 0, 0}, tau_prev[4] = {0, 0, 0, 0}, lev_top[4] = {0, 0, 0, 0};
    Real blay_prev = 0;

    auto level = [&](int j) { return a.top_at_1 ? j : nlay - j; };

    // ---- down-sweep, top of atmosphere -> surface; layer k-1 is finalised at iteration k because
    // its lower level source needs the Planck fraction of layer k (F90:681-708).
    for (int k = 0; k <= nlay; ++k)
    {
        Real pf[4] = {0, 0, 0, 0}, tau[4] = {0, 0, 0, 0};
        Real blay = 0;
        if (k < nlay)
        {
            const int l = a.top_at_1 ? k : nlay - 1 - k;
            const size_t icl = (size_t)c + (size_t)nc * l;
            const V4<Real> cell = st.cell[icl];
            const int f = flav[cell_side(cell)];
            Cell<Real> ce;
            decode_cell(P, cell, st.flvA[icl + tcl * f], st.flvB[icl + tcl * f], ce);
            Real s0[4], s1[4];
            gas_tau(P, B, ce, kmaj, q4, st.col_gas, icl, tcl, tau, s0, s1);
            if (CLOUD)
            {
                const Real ct = a.cld_tau[icl + tcl * band];
#pragma unroll
                for (int j = 0; j < 4; ++j) tau[j] += ct;
            }
            interp_major(pfr, P, ce, s0, s1);
#pragma unroll
            for (int j = 0; j < 4; ++j) pf[j] = s0[j] + s1[j];
            blay = planck_band(plk, P.nplanck, P.planck_tmin, P.planck_delta, a.tlay[cg + (size_t)a.ncol * l]);
        }
        const Real blev = planck_band(plk, P.nplanck, P.planck_tmin, P.planck_delta,
                                      a.tlev[cg + (size_t)a.ncol * level(k)]);
        if (k == 0)
        {
#pragma unroll
            for (int j = 0; j < 4; ++j) lev_top[j] = pf[j] * blev;
        }
        else
        {
            V4<Real> tr, su;
            Real* trp = &tr.x; Real* sup = &su.x;
#pragma unroll
            for (int j = 0; j < 4; ++j)
            {
                const Real lev_bot = (k < nlay) ? M<Real>::sqrt(pf_prev[j] * pf[j]) * blev : pf_prev[j] * blev;
                Real t, s_up, s_dn;
                lw_layer(tau_prev[j] * a.secant, pf_prev[j] * blay_prev, lev_top[j], lev_bot, t, s_up, s_dn);
                radn[j] = fma(t, radn[j], s_dn);
                trp[j] = t; sup[j] = s_up;
                lev_top[j] = lev_bot;
            }
            scr[(size_t)((k - 1) * 2 + 0) * FUSED_THREADS] = tr;
            scr[(size_t)((k - 1) * 2 + 1) * FUSED_THREADS] = su;
            const Real s = quad_sum<Real, LPC>(radn) * piw;
            if (lead)
            {
                const size_t o = cg + (size_t)a.ncol * level(k);
                a.flux_dn[o] += s;
                if (a.bnd_dn) a.bnd_dn[o + nlev_stride * band] = s;
            }
        }
        if (k < nlay)
        {
#pragma unroll
            for (int j = 0; j < 4; ++j) { pf_prev[j] = pf[j]; tau_prev[j] = tau[j]; }
            blay_prev = blay;
        }
    }

    // ---- surface: sfc_src = pfrac(sfc layer) * B(tsfc) (F90:640-656); reflection + emission
    {
        const Real bsfc = planck_band(plk, P.nplanck, P.planck_tmin, P.planck_delta, a.tsfc[cg]);
        const Real emis = a.sfc_emis[band + (size_t)P.nbnd * cg];
#pragma unroll
        for (int j = 0; j < 4; ++j) radn[j] = radn[j] * ((Real)1 - emis) + emis * (pf_prev[j] * bsfc);
        const Real s = quad_sum<Real, LPC>(radn) * piw;
        if (lead)
        {
            const size_t o = cg + (size_t)a.ncol * level(nlay);
            a.flux_up[o] += s;
            if (a.bnd_up) a.bnd_up[o + nlev_stride * band] = s;
        }
    }
    // ---- up-sweep, surface -> top (F90:710-745), state read back LIFO
    for (int k = nlay - 1; k >= 0; --k)
    {
        const V4<Real> tr = scr[(size_t)(k * 2 + 0) * FUSED_THREADS];
        const V4<Real> su = scr[(size_t)(k * 2 + 1) * FUSED_THREADS];
        radn[0] = fma(tr.x, radn[0], su.x); radn[1] = fma(tr.y, radn[1], su.y);
        radn[2] = fma(tr.z, radn[2], su.z); radn[3] = fma(tr.w, radn[3], su.w);
        const Real s = quad_sum<Real, LPC>(radn) * piw;
        if (lead)
        {
            const size_t o = cg + (size_t)a.ncol * level(k);
            a.flux_up[o] += s;
            if (a.bnd_up) a.bnd_up[o + nlev_stride * band] = s;
        }
    }
}

template<typename Real> size_t fused_scratch_bytes(int nc, int nlay, int gw, bool sw)
{
    const int cpb = FUSED_THREADS / (gw / 4);
    const size_t nblk = (size_t)(nc + cpb - 1) / cpb;
    return nblk * nlay * (sw ? 4 : 2) * FUSED_THREADS * sizeof(V4<Real>);
}

template<typename Real>
void launch_lw_band(const DevPlan<Real>& P, const BandArgs& B, const TileState<Real>& st,
                    const FusedArgs<Real>& a, cudaStream_t s)
{
    const int lpc = P.gw / 4, cpb = FUSED_THREADS / lpc;
    const unsigned grid = (unsigned)((a.nc + cpb - 1) / cpb);
    const bool cloud = a.cld_tau != nullptr;
    if (lpc == 4)
    {
        if (cloud) lw_band_kernel<Real, 4, true><<<grid, FUSED_THREADS, 0, s>>>(P, B, st, a);
        else lw_band_kernel<Real, 4, false><<<grid, FUSED_THREADS, 0, s>>>(P, B, st, a);
    }
    else
    {
        if (cloud) lw_band_kernel<Real, 2, true><<<grid, FUSED_THREADS, 0, s>>>(P, B, st, a);
        else lw_band_kernel<Real, 2, false><<<grid, FUSED_THREADS, 0, s>>>(P, B, st, a);
    }
    count_launch();
}

template size_t fused_scratch_bytes<float>(int, int, int, bool);
template size_t fused_scratch_bytes<double>(int, int, int, bool);
template void launch_lw_band<float>(const DevPlan<float>&, const BandArgs&, const TileState<float>&, const FusedArgs<float>&, cudaStream_t);
template void launch_lw_band<double>(const DevPlan<double>&, const BandArgs&, const TileState<double>&, const FusedArgs<double>&, cudaStream_t);

template<typename Real> int lw_blocks_per_wave(int gw)
{
    int dev = 0, sms = 0, occ = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (gw == 16) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, lw_band_kernel<Real, 4, false>, FUSED_THREADS, 0);
    else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, lw_band_kernel<Real, 2, false>, FUSED_THREADS, 0);
    return sms * occ;
}
template int lw_blocks_per_wave<float>(int);
template int lw_blocks_per_wave<double>(int);

}  // namespace rb200

// fused_lw.cuh — longwave: gas optics + Planck sources + no-scattering solver + g-point reduction
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
#pragma once
#include "physics.cuh"
#include "launch.hpp"

namespace rb200 {

#ifndef RB_LW_MINB
#define RB_LW_MINB 5
#endif
#ifndef RB_LW_MINB64
#define RB_LW_MINB64 3
#endif
#ifndef RB_LW_PFS
#define RB_LW_PFS 2             // down-sweep: layers beyond the register prefetch whose state lines are requested into L2 (0 = off)
#endif
#ifndef RB_LW_PF2
#define RB_LW_PF2 12            // up-sweep: layers ahead of the consumed ones whose strip lines are requested into L2 (0 = off)
#endif
// resident CTAs per SM requested from ptxas: 5 (96 registers) for fp32, 3 (168 registers) for fp64 — measured best
template<typename Real> struct LWMinBlocks { static constexpr int v = sizeof(Real) == 4 ? RB_LW_MINB : RB_LW_MINB64; };

// The whole band computation of one CTA; `bid` is the CTA's index among the launch's LW CTAs (blockIdx.x in the
// single-role kernel below, a role-local index in the paired longwave + shortwave kernel of fused_pair.cu).
// NANG > 1: the Gauss angles of the quadrature (src/Rte_lw.cpp:139-151) inside the layer loop — optical depths and Planck
// sources are computed once per layer, only exp(-tau * secant), the layer sources and the transport are per angle; the strip
// holds {trans, source_up} per (layer, angle).
template<typename Real, int LPC, int GWT, bool CLOUD, int NE, int NANG = 1>
__device__ __forceinline__ void lw_band_body(const DevPlan<Real>& P, const BandLaunch<Real>& L, const TileState<Real>& st,
                                             const FusedArgs<Real>& a, const int bid)
{
    const BandArgs& B = L.B;
    using G = Geo<GWT, NE, true>;      // table rows are GWT wide; LPC lanes cover this band's (<= 4*LPC) g-points
    constexpr int CPB = FUSED_THREADS / LPC;
    constexpr int T = FUSED_THREADS;
    const int tid = threadIdx.x;
    const int q = tid % LPC, cl = tid / LPC;
    const int craw = bid * CPB + cl;
    const bool active = craw < a.nc;
    const int c = active ? craw : a.nc - 1;
    const int nlay = a.nlay, nc = a.nc;
    const size_t tcl = L.tcl;
    const size_t cg = (size_t)a.c0 + c;                 // column in the caller's arrays
    const int q4 = q * 4;
    const int band = B.band;
    const Real* kmaj = L.kmaj + q4;
    const Real* plk = L.plk;
    const int flav0 = B.flav[0], flav1 = B.flav[1];
    const unsigned* tm = st.tmask + (size_t)c * st.tmw;
    const int slot = claim_slot(a.slot_busy, a.nslots, bid);
    V4<Real>* const scr = a.scratch + (size_t)slot * L.strip + tid;
    const Real pi = (Real)3.14159265358979323846;
    const Real piw = pi * a.weight;
    auto secant = [&](int ia) -> Real { return NANG == 1 ? a.secant : a.secants[ia]; };
    auto weight = [&](int ia) -> Real { return NANG == 1 ? a.weight : a.weights[ia]; };
    // pi * sum over angles of weight * radiance, summed over the band's g-points (one angle: the launch's weight)
    auto band_flux = [&](const Real (&r)[NANG][4]) -> Real
    {
        if (NANG == 1) return quad_sum<Real, LPC>(r[0]) * piw;
        Real ws[4] = {0, 0, 0, 0};
#pragma unroll
        for (int ia = 0; ia < NANG; ++ia)
#pragma unroll
            for (int j = 0; j < 4; ++j) ws[j] += weight(ia) * r[ia][j];
        return quad_sum<Real, LPC>(ws) * pi;
    };
    const bool lead = active && q == 0;
    const size_t nlev_stride = L.nlev_stride;
    const int top = a.top_at_1;
    const long lstep = L.lstep;
    const int tstep = top ? nc : -nc;                                          // same, tile arrays
    const size_t o_top = cg + L.o_top;                                         // level index * ncol

    Real radn[NANG][4] = {};            // incident diffuse flux = 0 (Radiation_solver.cu passes zeros)
    Real pf_prev[4] = {0, 0, 0, 0}, tau_prev[4] = {0, 0, 0, 0}, lev_top[4] = {0, 0, 0, 0};
    Real blay_prev = 0;

    // ---- down-sweep, top of atmosphere -> surface; layer k-1 is finalised at iteration k because
    // its lower level source needs the Planck fraction of layer k (F90:681-708).  State and
    // Planck-table positions of the next layer are fetched one iteration ahead, INTO THE SAME VARIABLES and
    // after their last use (the minor-gas loop): ptxas gives all these loads one scoreboard and drains it at
    // that loop's header, and a separate current/next pair costs a register copy that waits for the loads at once.
    int l = top ? 0 : nlay - 1;                         // current layer, array order
    const V2<Real>* plev_p = st.pl_lev + (size_t)c + (size_t)nc * (top ? 0 : nlay);
    RawState<Real> cur = load_state(st, L, tm, c, l);
    V2<Real> play_c = st.pl_lay[(size_t)c + (size_t)nc * l], plev_c = *plev_p;
    V2<Real> rwn{};                                     // table rows of the NEXT layer (see prefetch_tables)
    if (RB_TAB_PF && nlay > 1) rwn = load_rows(st, L, tm, c, top ? 1 : nlay - 2);
    size_t o = o_top;
    StatePrefetch<Real, LPC, true> pfs;
    if (RB_LW_PFS > 0) pfs.init(st, B, flav0, flav1, c, tcl);
    for (int k = 0; k <= nlay; ++k)
    {
        Real pf[4] = {0, 0, 0, 0}, tau[4] = {0, 0, 0, 0};
        Real blay = 0;
        const Real blev = planck_at(plk, plev_c);
        if (k < nlay)
        {
            const size_t icl = (size_t)c + (size_t)nc * l;
            if (RB_TAB_PF && k + 1 < nlay) prefetch_tables<G, LPC, false>(L, P, rwn, layer_side(tm, top ? l + 1 : l - 1), q);
            blay = planck_at(plk, play_c);
            Cell<Real> ce;
            decode_cell(cur.cell, cur.fa, cur.fb, ce);
            Real s0[4], s1[4], p0[4], p1[4];
            interp_major_pair<G>(kmaj, P, ce, s0, s1, p0, p1);
#pragma unroll
            for (int h = 0; h < 2; ++h)
            {
                pput(tau, h, pfma(ce.cm1, pget(s1, h), ce.cm0 * pget(s0, h)));
                pput(pf, h, pget(p0, h) + pget(p1, h));
            }
            add_minor_tau<G>(L, ce, q4, cur.sc0, cur.sc1, st.scal + icl, tau);
            if (CLOUD)          // before the prefetch below: a load consumed after it would wait for the prefetch too
            {
                const Real ct = a.cld_tau[icl + tcl * band];
#pragma unroll
                for (int h = 0; h < 2; ++h) pput(tau, h, pget(tau, h) + ct);
            }
            plev_p += tstep;
            plev_c = *plev_p;
            if (k + 1 < nlay)
            {
                const int ln = top ? l + 1 : l - 1;
                cur = load_state(st, L, tm, c, ln);
                play_c = st.pl_lay[(size_t)c + (size_t)nc * ln];
                if (RB_TAB_PF && k + 2 < nlay) rwn = load_rows(st, L, tm, c, top ? l + 2 : l - 2);
            }
            if (RB_LW_PFS > 0 && k + 1 + RB_LW_PFS < nlay)
            {
                // state of a layer further down goes to L2 now, so that the register prefetch of the next iteration
                // is an L2 hit (the state arrays stream from DRAM once per band, ~1 us under load)
                const int lp = top ? l + 1 + RB_LW_PFS : l - 1 - RB_LW_PFS;
                pfs.layer(lp, layer_side(tm, lp));
            }
            l += top ? 1 : -1;
        }
        if (k == 0)
        {
#pragma unroll
            for (int h = 0; h < 2; ++h) pput(lev_top, h, pget(pf, h) * blev);
            if (lead && a.bnd_dn && !a.bnd_add) a.bnd_dn[o_top + nlev_stride * band] = 0;      // no incident diffuse flux
        }
        else
        {
            Pr<Real> lev_bot[2];
#pragma unroll
            for (int h = 0; h < 2; ++h)
            {
                const Pr<Real> pfp = pget(pf_prev, h);
                lev_bot[h] = ((k < nlay) ? psqrt(pfp * pget(pf, h)) : pfp) * blev;
            }
#pragma unroll
            for (int ia = 0; ia < NANG; ++ia)
            {
                Pr<Real> t2[2], su2[2];
#pragma unroll
                for (int h = 0; h < 2; ++h)
                {
                    Pr<Real> s_dn;
                    lw_layer2(pget(tau_prev, h) * secant(ia), pget(pf_prev, h) * blay_prev, pget(lev_top, h), lev_bot[h], t2[h], su2[h], s_dn);
                    pput(radn[ia], h, pfma(t2[h], pget(radn[ia], h), s_dn));
                }
                V4<Real>* w = scr + ((size_t)(k - 1) * NANG + ia) * 2 * T;
                st_stream(w, v4_of(t2[0], t2[1])); st_stream(w + T, v4_of(su2[0], su2[1]));
            }
#pragma unroll
            for (int h = 0; h < 2; ++h) pput(lev_top, h, lev_bot[h]);
            const Real s = band_flux(radn);
            if (lead)
            {
                atomicAdd(a.flux_dn + o, s);
                if (a.bnd_dn) { Real* b_ = a.bnd_dn + o + nlev_stride * band; *b_ = a.bnd_add ? *b_ + s : s; }
            }
        }
        if (k < nlay)
        {
#pragma unroll
            for (int j = 0; j < 4; ++j) { pf_prev[j] = pf[j]; tau_prev[j] = tau[j]; }
            blay_prev = blay;
            o += lstep;
        }
    }

    // ---- surface: sfc_src = pfrac(sfc layer) * B(tsfc) (F90:640-656); reflection + emission
    {
        const Real bsfc = planck_band(plk, P.nplanck, P.planck_tmin, P.planck_delta, a.tsfc[cg]);
        const Real emis = a.sfc_emis[band + (size_t)P.nbnd * cg];
        const Real refl = (Real)1 - emis, eb = emis * bsfc;
#pragma unroll
        for (int ia = 0; ia < NANG; ++ia)
#pragma unroll
            for (int h = 0; h < 2; ++h) pput(radn[ia], h, pfma(eb, pget(pf_prev, h), pget(radn[ia], h) * refl));
        const Real s = band_flux(radn);
        if (lead)
        {
            atomicAdd(a.flux_up + o, s);
            if (a.bnd_up) { Real* b_ = a.bnd_up + o + nlev_stride * band; *b_ = a.bnd_add ? *b_ + s : s; }
        }
    }
    // ---- up-sweep, surface -> top (F90:710-745): the parked {trans, source_up} come back LIFO through
    // two register buffers of 4 layers; the loads of the next 4 layers are issued before the
    // arithmetic of the current 4.
    const size_t o_sfc = o;
#pragma unroll
    for (int ia = 0; ia < NANG; ++ia)
    {
        Real (&rad)[4] = radn[ia];
        const Real pw = NANG == 1 ? piw : pi * weight(ia);
        const bool add = a.bnd_add || ia > 0;           // by-band outputs: the first angle stores, the others add
        o = o_sfc;
        auto ld = [&](V4<Real> (&tr)[4], V4<Real> (&su)[4], int k0)
        {
#pragma unroll
            for (int u = 0; u < 4; ++u)
            {
                const V4<Real>* w = scr + ((size_t)max(k0 - u, 0) * NANG + ia) * 2 * T;
                tr[u] = ld_stream(w); su[u] = ld_stream(w + T);
            }
        };
        auto sweep = [&](const V4<Real> (&tr)[4], const V4<Real> (&su)[4], int k0)
        {
#pragma unroll
            for (int u = 0; u < 4; ++u)
            {
                if (k0 - u < 0) break;
                pput(rad, 0, pfma(plo(tr[u]), pget(rad, 0), plo(su[u])));
                pput(rad, 1, pfma(phi(tr[u]), pget(rad, 1), phi(su[u])));
                const Real s = quad_sum<Real, LPC>(rad) * pw;
                o -= lstep;
                if (lead)
                {
                    atomicAdd(a.flux_up + o, s);
                    if (a.bnd_up) { Real* b_ = a.bnd_up + o + nlev_stride * band; *b_ = add ? *b_ + s : s; }
                }
            }
        };
        // strip lines of the layers RB_LW_PF2 further up go to L2 while the current ones are consumed
        auto pf4 = [&](int k0)
        {
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (RB_LW_PF2 > 0 && k0 - u - RB_LW_PF2 >= 0) prefetch_strip_layer<Real, 2>(scr - tid, (k0 - u - RB_LW_PF2) * NANG + ia);
        };
        V4<Real> trA[4], suA[4], trB[4], suB[4];
        ld(trA, suA, nlay - 1);
        for (int k0 = nlay - 1; k0 >= 0; k0 -= 8)
        {
            ld(trB, suB, k0 - 4);
            pf4(k0);
            sweep(trA, suA, k0);
            if (k0 - 4 < 0) break;
            ld(trA, suA, k0 - 8);
            pf4(k0 - 4);
            sweep(trB, suB, k0 - 4);
        }
    }
    release_slot(a.slot_busy, slot);
}

template<typename Real, int LPC, int GWT, bool CLOUD, int NE>
__global__ void __launch_bounds__(FUSED_THREADS, LWMinBlocks<Real>::v)
lw_band_kernel(const DevPlan<Real> P, const __grid_constant__ BandLaunch<Real> L, const TileState<Real> st,
               const FusedArgs<Real> a)
{
    lw_band_body<Real, LPC, GWT, CLOUD, NE>(P, L, st, a, (int)blockIdx.x);
}

}  // namespace rb200

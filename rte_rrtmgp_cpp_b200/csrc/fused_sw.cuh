// fused_sw.cuh — shortwave: gas optics (+Rayleigh, +cloud increment) + two-stream coefficients +
// direct beam + adding + g-point reduction in one kernel per (column tile, band).
//
// Restates (fp order aside) the chain the oracle pins:
//   compute_tau_absorption / compute_tau_rayleigh  mo_gas_optics_rrtmgp_kernels.F90:178-564
//   combine_abs_and_rayleigh                       src/Gas_optics_rrtmgp.cpp:365-385
//   inc_2stream_by_2stream_bybnd (clouds)          mo_optical_props_kernels.F90:451-482
//   sw_solver_2stream = sw_dif_and_source + adding mo_rte_solver_kernels.F90:503-609,985-1245
//   sum_broadband / sum_byband
// Same thread mapping as fused_lw.cu.  Two vertical passes per (column, g-point):
//   1 top->bottom : optics, two-stream, direct beam, and the adding recurrences for the stack above
//                   each level (reflectance Ra seen from below, emerging downward diffuse flux Sa);
//                   parks {Rdif,Tdif,src_up,src_dn,Ra,Sa} (6 words per layer and g-point)
//   2 bottom->top : the reference's adding recurrences for the stack below each level (albedo A,
//                   source Sb); where the two meet the fluxes follow in closed form,
//                   flux_dn = (Sa + Ra*Sb)/(1 - Ra*A), flux_up = A*flux_dn + Sb, + reduction.
// The reference (and the oracle) sweeps a third time, top->bottom, for the fluxes; combining the
// two half-stacks instead is algebraically the same adding method (same interaction principle,
// same denominators 1 - R*A > 0) and saves re-parking 4 words per cell: 12 instead of 16 words of
// HBM scratch traffic per (cell, g-point).
#pragma once
#include "physics.cuh"
#include "launch.hpp"

namespace rb200 {

#ifndef RB_SW_MINB
#define RB_SW_MINB 5
#endif
#ifndef RB_SW_MINB64
#define RB_SW_MINB64 3
#endif
// L2 prefetch hints (see physics.cuh): the state two layers ahead helps a little (-1.5 %) since the prefetch was moved
// behind the iteration's last load; strip lines ahead of pass 2 do not (this kernel is DRAM-bound: +1 % at 4 layers, +11 % at 8)
#ifndef RB_SW_PFS
#define RB_SW_PFS 2
#endif
#ifndef RB_SW_PF2
#define RB_SW_PF2 0
#endif
// resident CTAs per SM requested from ptxas: 5 (96 registers) for fp32, 3 (168 registers) for fp64 — measured best
template<typename Real> struct SWMinBlocks { static constexpr int v = sizeof(Real) == 4 ? RB_SW_MINB : RB_SW_MINB64; };

// The whole band computation of one CTA; `bid` is the CTA's index among the launch's SW CTAs (blockIdx.x in the
// single-role kernel below, a role-local index in the paired longwave + shortwave kernel of fused_pair.cu).
template<typename Real, int LPC, int GWT, bool CLOUD, int NE>
__device__ __forceinline__ void sw_band_body(const DevPlan<Real>& P, const BandLaunch<Real>& L, const TileState<Real>& st,
                                             const FusedArgs<Real>& a, const int bid)
{
    const BandArgs& B = L.B;
    using G = Geo<GWT, NE, false>;      // table rows are GWT wide; LPC lanes cover this band's (<= 4*LPC) g-points
    constexpr int CPB = FUSED_THREADS / LPC;
    constexpr int T = FUSED_THREADS;
    const int tid = threadIdx.x;
    const int q = tid % LPC, cl = tid / LPC;
    const int craw = bid * CPB + cl;
    const bool active = craw < a.nc;
    const int c = active ? craw : a.nc - 1;
    const int nlay = a.nlay, nc = a.nc;
    const size_t tcl = L.tcl;
    const size_t cg = (size_t)a.c0 + c;
    const int q4 = q * 4;
    const int band = B.band;
    const Real* kmaj = L.kmaj + q4;
    const int flav0 = B.flav[0], flav1 = B.flav[1];
    const unsigned* tm = st.tmask + (size_t)c * st.tmw;
    const int slot = claim_slot(a.slot_busy, a.nslots, bid);
    V4<Real>* const scr = a.scratch + (size_t)slot * L.strip + tid;
    const bool lead = active && q == 0;
    const size_t nlev_stride = L.nlev_stride;
    const int top = a.top_at_1;
    // level j / layer k counted from the top of the atmosphere -> array index, times ncol
    const long lstep = L.lstep;
    const size_t o_top = cg + L.o_top;

    const Real mu0 = a.mu0[cg];
    const Real mu0s = M<Real>::max(M<Real>::sqrt(M<Real>::eps), mu0);
    const bool sun = mu0 > (Real)0;
    const Real tsi = a.tsi_scaling ? a.tsi_scaling[cg] : (Real)1;

    // top-of-domain boundary condition (F90:575-583): direct = toa_src * mu0, diffuse = 0
    Real dir[4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
    {
        const int gl = q4 + j;
        dir[j] = gl < B.ngb ? P.solar_source[B.gpt0 + gl] * tsi * mu0 : (Real)0;
    }
    {
        const Real s = quad_sum<Real, LPC>(dir);
        if (lead)
        {
            atomicAdd(a.flux_dir + o_top, s); atomicAdd(a.flux_dn + o_top, s);
            if (a.bnd_dir) a.bnd_dir[o_top + nlev_stride * band] = s;
        }
    }

    // ---- pass 1: optics + two-stream + direct beam + adding of the stack ABOVE each level, top -> surface.
    // Ra(j) = diffuse reflectance of everything above level j seen from below, Sa(j) = downward diffuse
    // flux leaving that stack (driven by the direct-beam sources), Ra(0) = Sa(0) = 0.  The next layer's
    // state is requested one iteration ahead (its address does not depend on loaded data: the flavor
    // comes from the column's tropopause bit mask), into the same variable and after its last use (the
    // minor-gas loop, see fused_lw.cu): the Rayleigh gather and the two-stream arithmetic cover the latency.
    Real Ra[4] = {0, 0, 0, 0}, Sa[4] = {0, 0, 0, 0};
    RawState<Real> cur = load_state(st, L, tm, c, top ? 0 : nlay - 1);
    V2<Real> rwn{};                                     // table rows of the NEXT layer (see prefetch_tables)
    if (RB_TAB_PF && nlay > 1) rwn = load_rows(st, L, tm, c, top ? 1 : nlay - 2);
    size_t o = o_top;
    StatePrefetch<Real, LPC, false> pfs;
    if (RB_SW_PFS > 0) pfs.init(st, B, flav0, flav1, c, tcl);
    for (int k = 0; k < nlay; ++k)
    {
        const int l = top ? k : nlay - 1 - k;
        const size_t icl = (size_t)c + (size_t)nc * l;
        if (RB_TAB_PF && k + 1 < nlay) prefetch_tables<G, LPC, true>(L, P, rwn, layer_side(tm, top ? l + 1 : l - 1), q);
        Cell<Real> ce;
        decode_cell(cur.cell, cur.fa, cur.fb, ce);
        Real tau[4];
        {
            Real s0[4], s1[4];
            interp_major<G>(kmaj, P, ce, s0, s1);
#pragma unroll
            for (int h = 0; h < 2; ++h) pput(tau, h, pfma(ce.cm1, pget(s1, h), ce.cm0 * pget(s0, h)));
            add_minor_tau<G>(L, ce, q4, cur.sc0, cur.sc1, st.scal + icl, tau);
        }
        Real ray[4] = {0, 0, 0, 0};
        if (L.kray[0])
        {
            interp_minor<G>(L.kray[ce.side] + q4, ce, ray);
#pragma unroll
            for (int h = 0; h < 2; ++h) pput(ray, h, pget(ray, h) * ce.wet);
        }
        // optical properties of the 4 g-points: gas + Rayleigh (+ by-band cloud/aerosol increment)
        Pr<Real> tt[2], ww[2], gg[2];
        {
            Real ct = 0, cs = 0, cgg = 0;
            if (CLOUD) { ct = a.cld_tau[icl + tcl * band]; cs = a.cld_ssa[icl + tcl * band]; cgg = a.cld_g[icl + tcl * band]; }
#pragma unroll
            for (int h = 0; h < 2; ++h)
            {
                gg[h] = Pr<Real>((Real)0);
                combine_rayleigh2(pget(tau, h), pget(ray, h), tt[h], ww[h]);
                if (CLOUD) increment_2str2(tt[h], ww[h], gg[h], ct, cs, cgg);
            }
        }
        // The next layer's state is requested HERE: after the last load of this iteration has been consumed (every wait on
        // the scoreboard these loads share drains all of them, so a later table or cloud load would wait for the prefetch
        // at once) and before the two-stream arithmetic, which covers the latency.
        if (k + 1 < nlay) cur = load_state(st, L, tm, c, top ? k + 1 : nlay - 2 - k);
        if (RB_TAB_PF && k + 2 < nlay) rwn = load_rows(st, L, tm, c, top ? k + 2 : nlay - 3 - k);
        if (RB_SW_PFS > 0 && k + 1 + RB_SW_PFS < nlay)
        {
            const int lp = top ? k + 1 + RB_SW_PFS : nlay - 2 - k - RB_SW_PFS;      // to L2, see fused_lw.cu
            pfs.layer(lp, layer_side(tm, lp));
        }
        // two g-points at a time: packed fp32 arithmetic (FFMA2 / FMUL2 / FADD2), exp / rcp / sqrt per lane
        Pr<Real> Rdif[2], Tdif[2], su[2], sd[2], Ra0[2], Sa0[2];
#pragma unroll
        for (int h = 0; h < 2; ++h)
        {
            Pr<Real> Rdir, Tdir, Tnoscat;
            two_stream2<Real>(tt[h], ww[h], gg[h], mu0s, Rdif[h], Tdif[h], Rdir, Tdir, Tnoscat);
            const Pr<Real> dirh = pget(dir, h), Rah = pget(Ra, h), Sah = pget(Sa, h);
            su[h] = sun ? Rdir * dirh : Pr<Real>((Real)0);     // source_up
            sd[h] = sun ? Tdir * dirh : Pr<Real>((Real)0);     // source_dn
            pput(dir, h, Tnoscat * dirh);
            Ra0[h] = Rah; Sa0[h] = Sah;
            // add layer k below the stack: multiple reflections between Ra (above) and Rdif (the layer)
            const Pr<Real> td = Tdif[h] * prcp(pfma(-Rah, Rdif[h], (Real)1));
            pput(Sa, h, pfma(td, pfma(Rah, su[h], Sah), sd[h]));
            pput(Ra, h, pfma(Tdif[h] * td, Rah, Rdif[h]));
        }
        V4<Real>* w = scr + (size_t)k * 6 * T;
        st_stream(w, v4_of(Rdif[0], Rdif[1])); st_stream(w + T, v4_of(Tdif[0], Tdif[1])); st_stream(w + 2 * T, v4_of(su[0], su[1]));
        st_stream(w + 3 * T, v4_of(sd[0], sd[1])); st_stream(w + 4 * T, v4_of(Ra0[0], Ra0[1])); st_stream(w + 5 * T, v4_of(Sa0[0], Sa0[1]));
        const Real s = quad_sum<Real, LPC>(dir);
        o += lstep;
        if (lead)
        {
            atomicAdd(a.flux_dir + o, s); atomicAdd(a.flux_dn + o, s);
            if (a.bnd_dir) a.bnd_dir[o + nlev_stride * band] = s;
        }
    }

    // ---- pass 2: adding of the stack BELOW each level, surface -> top (F90:1135-1245), and the fluxes
    // where the two stacks meet:  dn(j) = (Sa + Ra*Sb) / (1 - Ra*A),  up(j) = A*dn(j) + Sb.
    // `o` is at the surface level; scratch is read two layers ahead.
    // Scratch comes back LIFO through two register buffers: the loads of layer k-1 are issued before
    // the arithmetic of layer k, so a full layer step sits between a load and its first use.
    auto ld6 = [&](V4<Real> (&d)[6], int k)
    {
        const V4<Real>* w = scr + (size_t)max(k, 0) * 6 * T;
#pragma unroll
        for (int m = 0; m < 6; ++m) d[m] = ld_stream(w + m * T);
    };
    V4<Real> bufA[6], bufB[6], bufC[6];
    ld6(bufA, nlay - 1);
    ld6(bufB, nlay - 2);
    Real alb[4], src[4];
    {
        const Real adir = a.alb_dir[band + (size_t)P.nbnd * cg];
        const Real adif = a.alb_dif[band + (size_t)P.nbnd * cg];
        Real fdn[4], fup[4];
#pragma unroll
        for (int h = 0; h < 2; ++h)
        {
            const Pr<Real> albh(adif), srch = sun ? pget(dir, h) * adir : Pr<Real>((Real)0), Rah = pget(Ra, h);
            const Pr<Real> fd = pfma(Rah, srch, pget(Sa, h)) * prcp(pfma(-Rah, albh, (Real)1));
            pput(alb, h, albh); pput(src, h, srch);
            pput(fdn, h, fd); pput(fup, h, pfma(albh, fd, srch));
        }
        const Real su = quad_sum<Real, LPC>(fup), sd = quad_sum<Real, LPC>(fdn);
        if (lead)
        {
            atomicAdd(a.flux_up + o, su); atomicAdd(a.flux_dn + o, sd);
            if (a.bnd_up) { a.bnd_up[o + nlev_stride * band] = su; a.bnd_dn[o + nlev_stride * band] = sd + a.bnd_dir[o + nlev_stride * band]; }
        }
    }
    int kpf = nlay - 1 - RB_SW_PF2;     // strip layer whose lines go to L2 while layer kpf + RB_SW_PF2 is consumed
    auto step = [&](const V4<Real> (&in)[6])
    {
        if (RB_SW_PF2 > 0 && kpf >= 0) prefetch_strip_layer<Real, 6>(scr - tid, kpf);
        --kpf;
        Real fdn[4], fup[4];
#pragma unroll
        for (int h = 0; h < 2; ++h)
        {
            const Pr<Real> rdif = h ? phi(in[0]) : plo(in[0]), tdif = h ? phi(in[1]) : plo(in[1]), sup = h ? phi(in[2]) : plo(in[2]);
            const Pr<Real> sdn = h ? phi(in[3]) : plo(in[3]), ra = h ? phi(in[4]) : plo(in[4]), sa = h ? phi(in[5]) : plo(in[5]);
            const Pr<Real> albh = pget(alb, h), srch = pget(src, h);
            const Pr<Real> td = tdif * prcp(pfma(-rdif, albh, (Real)1));
            const Pr<Real> nsrc = pfma(td, pfma(albh, sdn, srch), sup);
            const Pr<Real> nalb = pfma(tdif * td, albh, rdif);
            const Pr<Real> fd = pfma(ra, nsrc, sa) * prcp(pfma(-ra, nalb, (Real)1));
            pput(alb, h, nalb); pput(src, h, nsrc);
            pput(fdn, h, fd); pput(fup, h, pfma(nalb, fd, nsrc));
        }
        const Real su = quad_sum<Real, LPC>(fup), sd = quad_sum<Real, LPC>(fdn);
        o -= lstep;
        if (lead)
        {
            atomicAdd(a.flux_up + o, su); atomicAdd(a.flux_dn + o, sd);
            if (a.bnd_up) { a.bnd_up[o + nlev_stride * band] = su; a.bnd_dn[o + nlev_stride * band] = sd + a.bnd_dir[o + nlev_stride * band]; }
        }
    };
    // three rotating buffers: two layers (192 B per thread) in flight while the third is consumed
    for (int k0 = nlay - 1; k0 >= 0; k0 -= 3)
    {
        ld6(bufC, k0 - 2);
        step(bufA);
        if (k0 - 1 < 0) break;
        ld6(bufA, k0 - 3);
        step(bufB);
        if (k0 - 2 < 0) break;
        ld6(bufB, k0 - 4);
        step(bufC);
    }
    release_slot(a.slot_busy, slot);
}

template<typename Real, int LPC, int GWT, bool CLOUD, int NE>
__global__ void __launch_bounds__(FUSED_THREADS, SWMinBlocks<Real>::v)
sw_band_kernel(const DevPlan<Real> P, const __grid_constant__ BandLaunch<Real> L, const TileState<Real> st,
               const FusedArgs<Real> a)
{
    sw_band_body<Real, LPC, GWT, CLOUD, NE>(P, L, st, a, (int)blockIdx.x);
}

}  // namespace rb200

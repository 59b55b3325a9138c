// colres.cu — column-resident fused kernels: the per-(column, g-point) vertical recurrences of the
// RTE solvers live in shared memory, never in HBM.
//
// One CTA owns 16 recurrence chains = (16 / gw) columns x the gw (padded) g-points of a band and
// loops over all bands of the k-distribution:
//   phase 1 (all 128 threads, layer-parallel): thread = (layer group, 4 g-points).  Gas optics
//            (+ Rayleigh / cloud increment) and the layer's transfer coefficients (two-stream
//            R/T or exp(-tau) + Planck sources) -> shared memory rows [word][level][chain].
//   phase 2 (16 lanes of warp 0, one chain each): the sequential sweeps (direct beam, adding,
//            fluxes / down- and up-welling radiances) read and overwrite the rows in place.
//   phase 3 (all threads, level-parallel): g-point sums by quad shuffle into per-CTA broadband
//            accumulators (and the optional by-band outputs).
// After the last band the accumulators are stored once to (ncol, nlay+1): no atomics, no memset.
// DRAM traffic is the prologue kernel's interpolation state (~100 B per (column, layer, band))
// and the k-tables (L2-resident); 5 CTAs x (41 KB + 1.5 KB) fit an SM's 227 KB at fp32, nlay 128.
//
// Restates the same reference chain as fused_lw.cu / fused_sw.cu (same per-cell arithmetic from
// physics.cuh); see those files for the Fortran line references.
#include "physics.cuh"
#include "launch.hpp"
#include <algorithm>
#include <atomic>

namespace rb200 {

constexpr int CR_T = FUSED_THREADS;      // 128
constexpr int CR_CH = 16;                // chains per CTA = floats per shared-memory word row
constexpr int CR_NL = CR_T / 4;          // layers in flight in phase 1 (4 quads per row)
constexpr int CR_RS = 80;                // floats per level row: up to 5 words x 16 chains (80 = 16 mod 32 keeps
                                         // the phase-1 vector stores of consecutive levels on disjoint banks)

template<typename Real> struct CRMinBlocks { static constexpr int v = sizeof(Real) == 4 ? 5 : 2; };

template<typename Real> __device__ __forceinline__ V4<Real> lds4(const Real* p) { return *reinterpret_cast<const V4<Real>*>(p); }
template<typename Real> __device__ __forceinline__ void sts4(Real* p, const V4<Real>& v) { *reinterpret_cast<V4<Real>*>(p) = v; }

// Shared memory: row j (level j; layer k lives in row k+1 unless noted) holds the words of the 16
// chains back to back, sm[j*CR_RS + w*16 + chain], so a sweep lane walks one pointer with
// compile-time word offsets.  Sweeps load a group of rows into registers before touching any of
// them (the compiler cannot move shared loads above the stores of the previous row itself).

// ------------------------------------------------------------------------------------ shortwave
// words: 0 Rdif -> a, 1 Tdif -> b -> flux_dn(diffuse), 2 Rdir -> source_up -> albedo, 3 Tdir -> source_dn -> src -> flux_up,
//        4 Tnoscat -> direct flux
template<typename Real, int LPC, bool CLOUD, int NE>
__global__ void __launch_bounds__(CR_T, CRMinBlocks<Real>::v)
sw_col_kernel(const DevPlan<Real> P, const BandArgs* __restrict__ bands, const TileState<Real> st,
              const FusedArgs<Real> a)
{
    using G = Geo<4 * LPC, NE, false>;
    constexpr int GW = 4 * LPC, NCB = CR_CH / GW;
    constexpr int W0 = 0, W1 = 16, W2 = 32, W3 = 48, W4 = 64;
    extern __shared__ __align__(16) unsigned char cr_smem[];
    __shared__ BandArgs sB;
    const int nlay = a.nlay, nlev = nlay + 1, nc = a.nc;
    Real* const sm = reinterpret_cast<Real*>(cr_smem);
    Real* const acc = sm + (size_t)nlev * CR_RS;            // [NCB][3][nlev] broadband up, dn, dir

    const int tid = threadIdx.x;
    const int quad = tid & 3, lsub = tid >> 2;
    const int cb = quad / LPC, q = quad % LPC, q4 = q * 4;
    const int craw = blockIdx.x * NCB + cb;
    const bool active = craw < nc;
    const int c = active ? craw : nc - 1;
    const size_t tcl = (size_t)nc * nlay;
    const size_t cg = (size_t)a.c0 + c;
    const unsigned* tm = st.tmask + (size_t)c * st.tmw;
    const int top = a.top_at_1;
    const size_t nlev_stride = (size_t)a.ncol * nlev;
    const Real mu0 = a.mu0[cg];
    const Real mu0s = M<Real>::max(M<Real>::sqrt(M<Real>::eps), mu0);

    // sweep lanes: chain ch = tid (< 16) -> (column cbs, g-point gs)
    const int cbs = (tid & 15) / GW, gs = (tid & 15) % GW;
    const int cs = min(blockIdx.x * NCB + cbs, nc - 1);
    const size_t cgs = (size_t)a.c0 + cs;

    // one CTA per column group: the broadband sums are stored once.  Bands split over gridDim.y: every y stores its own
    // partial sums (part[y][field][ncol * nlev]) and sum_parts_kernel adds them in band order (no atomics: reproducible)
    const bool split = gridDim.y > 1;
    Real* const o_up = split ? a.part + ((size_t)blockIdx.y * 3 + 0) * nlev_stride : a.flux_up;
    Real* const o_dn = split ? a.part + ((size_t)blockIdx.y * 3 + 1) * nlev_stride : a.flux_dn;
    Real* const o_dir = split ? a.part + ((size_t)blockIdx.y * 3 + 2) * nlev_stride : a.flux_dir;
    for (int i = tid; i < NCB * 3 * nlev; i += CR_T) acc[i] = (Real)0;

    for (int ib = blockIdx.y; ib < P.nbnd; ib += gridDim.y)     // gridDim.y > 1: the bands of a small call spread over CTAs
    {
        if (tid == 0) sB = bands[ib];
        __syncthreads();                 // also orders the previous band's phase 3 before this phase 1
        const BandArgs& B = sB;
        const int band = B.band;
        const size_t band_elems = (size_t)P.ntemp * (P.npres + 1) * G::neta(P) * G::RS;
        const size_t ray_elems = (size_t)P.ntemp * G::neta(P) * G::GW;
        const Real* kmaj = P.kmajor + band_elems * band + q4;
        const Real* kray = P.krayl ? P.krayl + ray_elems * band + q4 : nullptr;
        const size_t ray_side = ray_elems * P.nbnd;
        const int flav0 = P.band_flavor[2 * band], flav1 = P.band_flavor[2 * band + 1];

        // ---- phase 1: optics + two-stream coefficients of every layer, 32 layers at a time
        {
            int k = lsub;
            RawState<Real> nxt;
            if (k < nlay) nxt = load_state(st, P, B, tm, flav0, flav1, c, top ? k : nlay - 1 - k, tcl);
            for (; k < nlay; k += CR_NL)
            {
                const int l = top ? k : nlay - 1 - k;
                const size_t icl = (size_t)c + (size_t)nc * l;
                const RawState<Real> cur = nxt;
                if (k + CR_NL < nlay) nxt = load_state(st, P, B, tm, flav0, flav1, c, top ? k + CR_NL : nlay - 1 - k - CR_NL, tcl);
                Cell<Real> ce;
                decode_cell(cur.cell, cur.fa, cur.fb, ce);
                Real tau[4];
                gas_tau<G>(P, B, ce, kmaj, q4, cur.sc0, cur.sc1, st.scal + icl, tcl, tau);
                Real ray[4] = {0, 0, 0, 0};
                if (kray)
                {
                    interp_minor<G>(kray + (ce.side ? ray_side : 0), ce, ray);
#pragma unroll
                    for (int j = 0; j < 4; ++j) ray[j] *= ce.wet;
                }
                Real ct = 0, cs_ = 0, cgg = 0;
                if (CLOUD) { ct = a.cld_tau[icl + tcl * band]; cs_ = a.cld_ssa[icl + tcl * band]; cgg = a.cld_g[icl + tcl * band]; }
                V4<Real> v0, v1, v2, v3, v4;
                Real* p0 = &v0.x; Real* p1 = &v1.x; Real* p2 = &v2.x; Real* p3 = &v3.x; Real* p4 = &v4.x;
#pragma unroll
                for (int j = 0; j < 4; ++j)
                {
                    Real t, w0, g = 0;
                    combine_rayleigh(tau[j], ray[j], t, w0);
                    if (CLOUD) increment_2str(t, w0, g, ct, cs_, cgg);
                    two_stream(t, w0, g, mu0s, p0[j], p1[j], p2[j], p3[j], p4[j]);
                }
                Real* r = sm + (k + 1) * CR_RS + quad * 4;
                sts4(r + W0, v0); sts4(r + W1, v1); sts4(r + W2, v2); sts4(r + W3, v3); sts4(r + W4, v4);
            }
        }
        __syncthreads();

        // ---- phase 2: the three sweeps of one chain per lane (mo_rte_solver_kernels.F90:575-609, 1135-1245)
        if (tid < CR_CH)
        {
            Real* const p0 = sm + tid;
            const Real mu0c = a.mu0[cgs];
            const bool sun = mu0c > (Real)0;
            const Real tsi = a.tsi_scaling ? a.tsi_scaling[cgs] : (Real)1;
            Real dir = gs < B.ngb ? P.solar_source[B.gpt0 + gs] * tsi * mu0c : (Real)0;
            p0[W4] = dir;
            // direct beam, top -> surface: source_up/dn = Rdir/Tdir * incident direct flux
            {
                Real* p = p0 + CR_RS;
                int k = 0;
                for (; k + 8 <= nlay; k += 8, p += 8 * CR_RS)
                {
                    Real rd[8], td[8], tn[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) { rd[u] = p[u * CR_RS + W2]; td[u] = p[u * CR_RS + W3]; tn[u] = p[u * CR_RS + W4]; }
#pragma unroll
                    for (int u = 0; u < 8; ++u)
                    {
                        p[u * CR_RS + W2] = sun ? rd[u] * dir : (Real)0;
                        p[u * CR_RS + W3] = sun ? td[u] * dir : (Real)0;
                        dir = tn[u] * dir;
                        p[u * CR_RS + W4] = dir;
                    }
                }
                for (; k < nlay; ++k, p += CR_RS)
                {
                    const Real rd = p[W2], td = p[W3], tn = p[W4];
                    p[W2] = sun ? rd * dir : (Real)0; p[W3] = sun ? td * dir : (Real)0;
                    dir = tn * dir; p[W4] = dir;
                }
            }
            // adding, surface -> top: leaves {a, b, albedo(k+1), src(k+1)} with
            //   flux_dn(k+1) = a*flux_dn(k) + b,  flux_up(k+1) = flux_dn(k+1)*albedo(k+1) + src(k+1)
            Real alb = a.alb_dif[band + (size_t)P.nbnd * cgs];
            Real src = sun ? dir * a.alb_dir[band + (size_t)P.nbnd * cgs] : (Real)0;
            auto add_row = [&](Real* pr, Real rdif, Real tdif, Real sup, Real sdn)
            {
                const Real denom = M<Real>::rcp((Real)1 - rdif * alb);
                const Real td = tdif * denom;
                pr[W0] = td;
                pr[W1] = (rdif * src + sdn) * denom;
                pr[W2] = alb; pr[W3] = src;
                const Real nsrc = sup + td * (src + alb * sdn);
                alb = rdif + tdif * td * alb;
                src = nsrc;
            };
            {
                Real* p = p0 + nlay * CR_RS;           // row of layer nlay-1
                int k = nlay - 1;
                for (; k >= 3; k -= 4, p -= 4 * CR_RS)
                {
                    Real r0[4], r1[4], r2[4], r3[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) { const Real* pr = p - u * CR_RS; r0[u] = pr[W0]; r1[u] = pr[W1]; r2[u] = pr[W2]; r3[u] = pr[W3]; }
#pragma unroll
                    for (int u = 0; u < 4; ++u) add_row(p - u * CR_RS, r0[u], r1[u], r2[u], r3[u]);
                }
                for (; k >= 0; --k, p -= CR_RS) add_row(p, p[W0], p[W1], p[W2], p[W3]);
            }
            // fluxes, top -> surface; diffuse flux_dn(top) = 0
            Real fdn = 0, fup = fdn * alb + src;
            p0[W1] = fdn; p0[W3] = fup;
            {
                Real* p = p0 + CR_RS;
                int k = 0;
                for (; k + 8 <= nlay; k += 8, p += 8 * CR_RS)
                {
                    Real ca[8], cb_[8], cc[8], cd[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) { const Real* pr = p + u * CR_RS; ca[u] = pr[W0]; cb_[u] = pr[W1]; cc[u] = pr[W2]; cd[u] = pr[W3]; }
#pragma unroll
                    for (int u = 0; u < 8; ++u)
                    {
                        fdn = fma(ca[u], fdn, cb_[u]);
                        fup = fma(fdn, cc[u], cd[u]);
                        p[u * CR_RS + W1] = fdn; p[u * CR_RS + W3] = fup;
                    }
                }
                for (; k < nlay; ++k, p += CR_RS)
                {
                    fdn = fma(p[W0], fdn, p[W1]);
                    fup = fma(fdn, p[W2], p[W3]);
                    p[W1] = fdn; p[W3] = fup;
                }
            }
        }
        __syncthreads();

        // ---- phase 3: g-point sums per level -> broadband accumulators (+ by-band outputs)
        for (int j0 = 0; j0 < nlev; j0 += CR_NL)
        {
            const int jr = j0 + lsub;
            const bool valid = jr < nlev;
            const int j = valid ? jr : nlev - 1;
            const Real* r = sm + j * CR_RS + quad * 4;
            const V4<Real> vu = lds4(r + W3), vd = lds4(r + W1), vr = lds4(r + W4);
            const Real au[4] = {vu.x, vu.y, vu.z, vu.w}, ad[4] = {vd.x, vd.y, vd.z, vd.w}, ar[4] = {vr.x, vr.y, vr.z, vr.w};
            const Real su = quad_sum<Real, LPC>(au), sd = quad_sum<Real, LPC>(ad), sr = quad_sum<Real, LPC>(ar);
            if (valid && q == 0 && active)
            {
                Real* ac = acc + (size_t)cb * 3 * nlev;
                ac[j] += su;
                ac[nlev + j] = (ac[nlev + j] + sr) + sd;
                ac[2 * nlev + j] += sr;
                if (a.bnd_up)
                {
                    const size_t o = cg + (size_t)a.ncol * (top ? j : nlay - j) + nlev_stride * band;
                    a.bnd_up[o] = su; a.bnd_dn[o] = sd + sr;
                    if (a.bnd_dir) a.bnd_dir[o] = sr;
                }
            }
        }
    }
    __syncthreads();
    for (int i = tid; i < NCB * nlev; i += CR_T)
    {
        const int cbi = i / nlev, j = i - cbi * nlev;
        const int cc = blockIdx.x * NCB + cbi;
        if (cc < nc)
        {
            const size_t o = (size_t)a.c0 + cc + (size_t)a.ncol * (top ? j : nlay - j);
            const Real* ac = acc + (size_t)cbi * 3 * nlev;
            o_up[o] = ac[j]; o_dn[o] = ac[nlev + j];
            if (o_dir) o_dir[o] = ac[2 * nlev + j];
        }
    }
}

// ------------------------------------------------------------------------------------- longwave
// words (rows are layers k / levels j, no shift except word 3):
//   0 tau(k) -> source_up(k) -> radiance up at level k; row nlay: radiance up at the surface
//   1 Planck fraction of layer k
//   2 trans(k); row nlay: surface source
//   3 row k+1: source_dn(k) -> radiance down at level k+1; row 0: incident (0)
template<typename Real, int LPC, bool CLOUD, int NE>
__global__ void __launch_bounds__(CR_T, CRMinBlocks<Real>::v)
lw_col_kernel(const DevPlan<Real> P, const BandArgs* __restrict__ bands, const TileState<Real> st,
              const FusedArgs<Real> a)
{
    using G = Geo<4 * LPC, NE, true>;
    constexpr int GW = 4 * LPC, NCB = CR_CH / GW;
    constexpr int W0 = 0, W1 = 16, W2 = 32, W3 = 48;
    extern __shared__ __align__(16) unsigned char cr_smem[];
    __shared__ BandArgs sB;
    const int nlay = a.nlay, nlev = nlay + 1, nc = a.nc;
    Real* const sm = reinterpret_cast<Real*>(cr_smem);
    Real* const acc = sm + (size_t)nlev * CR_RS;         // [NCB][2][nlev] broadband up, dn

    const int tid = threadIdx.x;
    const int quad = tid & 3, lsub = tid >> 2;
    const int cb = quad / LPC, q = quad % LPC, q4 = q * 4;
    const int craw = blockIdx.x * NCB + cb;
    const bool active = craw < nc;
    const int c = active ? craw : nc - 1;
    const size_t tcl = (size_t)nc * nlay;
    const size_t cg = (size_t)a.c0 + c;
    const unsigned* tm = st.tmask + (size_t)c * st.tmw;
    const int top = a.top_at_1;
    const size_t nlev_stride = (size_t)a.ncol * nlev;
    const Real piw = (Real)3.14159265358979323846 * a.weight;
    const Real tsfc = a.tsfc[cg];

    const int cbs = (tid & 15) / GW;
    const int cs = min(blockIdx.x * NCB + cbs, nc - 1);
    const size_t cgs = (size_t)a.c0 + cs;

    const bool split = gridDim.y > 1;
    Real* const o_up = split ? a.part + ((size_t)blockIdx.y * 3 + 0) * nlev_stride : a.flux_up;
    Real* const o_dn = split ? a.part + ((size_t)blockIdx.y * 3 + 1) * nlev_stride : a.flux_dn;
    for (int i = tid; i < NCB * 2 * nlev; i += CR_T) acc[i] = (Real)0;

    for (int ib = blockIdx.y; ib < P.nbnd; ib += gridDim.y)     // gridDim.y > 1: the bands of a small call spread over CTAs
    {
        if (tid == 0) sB = bands[ib];
        __syncthreads();
        const BandArgs& B = sB;
        const int band = B.band;
        const size_t band_elems = (size_t)P.ntemp * (P.npres + 1) * G::neta(P) * G::RS;
        const Real* kmaj = P.kmajor + band_elems * band + q4;
        const Real* plk = P.totplnk + (size_t)P.nplanck * band;
        const int flav0 = P.band_flavor[2 * band], flav1 = P.band_flavor[2 * band + 1];

        // ---- phase 1a: absorption optical depth and Planck fraction of every layer
        {
            int k = lsub;
            RawState<Real> nxt;
            if (k < nlay) nxt = load_state(st, P, B, tm, flav0, flav1, c, top ? k : nlay - 1 - k, tcl);
            for (; k < nlay; k += CR_NL)
            {
                const int l = top ? k : nlay - 1 - k;
                const size_t icl = (size_t)c + (size_t)nc * l;
                const RawState<Real> cur = nxt;
                if (k + CR_NL < nlay) nxt = load_state(st, P, B, tm, flav0, flav1, c, top ? k + CR_NL : nlay - 1 - k - CR_NL, tcl);
                Cell<Real> ce;
                decode_cell(cur.cell, cur.fa, cur.fb, ce);
                Real s0[4], s1[4], p0[4], p1[4];
                interp_major_pair<G>(kmaj, P, ce, s0, s1, p0, p1);
                V4<Real> vt, vp;
                Real* tau = &vt.x; Real* pf = &vp.x;
                Real tl[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) { tl[j] = ce.cm0 * s0[j] + ce.cm1 * s1[j]; pf[j] = p0[j] + p1[j]; }
                add_minor_tau<G>(P, B, ce, q4, cur.sc0, cur.sc1, st.scal + icl, tcl, tl);
                if (CLOUD)
                {
                    const Real ct = a.cld_tau[icl + tcl * band];
#pragma unroll
                    for (int j = 0; j < 4; ++j) tl[j] += ct;
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) tau[j] = tl[j];
                Real* r = sm + k * CR_RS + quad * 4;
                sts4(r + W0, vt); sts4(r + W1, vp);
            }
        }
        __syncthreads();
        // ---- phase 1b: level sources need the Planck fraction of both neighbours (F90:681-708);
        // then lw_source_noscat + transmissivity per layer (F90:620-680)
        for (int k = lsub; k < nlay; k += CR_NL)
        {
            const int l = top ? k : nlay - 1 - k;
            Real* r = sm + k * CR_RS + quad * 4;
            const V4<Real> vt = lds4(r + W0), vp = lds4(r + W1);
            const V4<Real> vm = lds4(r + W1 - (k > 0 ? CR_RS : 0)), vn = lds4(r + W1 + (k + 1 < nlay ? CR_RS : 0));
            const Real blay = planck_at(plk, st.pl_lay[(size_t)c + (size_t)nc * l]);
            const Real bla = planck_at(plk, st.pl_lev[(size_t)c + (size_t)nc * (top ? k : nlay - k)]);
            const Real blb = planck_at(plk, st.pl_lev[(size_t)c + (size_t)nc * (top ? k + 1 : nlay - k - 1)]);
            const Real* tau = &vt.x; const Real* pf = &vp.x; const Real* pm = &vm.x; const Real* pn = &vn.x;
            V4<Real> ot, ou, od;
            Real* trp = &ot.x; Real* sup = &ou.x; Real* sdp = &od.x;
#pragma unroll
            for (int j = 0; j < 4; ++j)
            {
                const Real lev_a = k > 0 ? M<Real>::sqrt(pm[j] * pf[j]) * bla : pf[j] * bla;
                const Real lev_b = k + 1 < nlay ? M<Real>::sqrt(pf[j] * pn[j]) * blb : pf[j] * blb;
                lw_layer(tau[j] * a.secant, pf[j] * blay, lev_a, lev_b, trp[j], sup[j], sdp[j]);
            }
            sts4(r + W2, ot); sts4(r + W0, ou); sts4(r + CR_RS + W3, od);
            if (k == nlay - 1)
            {
                // sfc_src = pfrac(surface layer) * B(tsfc)  (F90:640-656)
                const Real bsfc = planck_band(plk, P.nplanck, P.planck_tmin, P.planck_delta, tsfc);
                V4<Real> os; os.x = pf[0] * bsfc; os.y = pf[1] * bsfc; os.z = pf[2] * bsfc; os.w = pf[3] * bsfc;
                sts4(r + CR_RS + W2, os);
            }
        }
        __syncthreads();

        // ---- phase 2: down-sweep, surface, up-sweep of one chain per lane (F90:51-240, 710-745)
        if (tid < CR_CH)
        {
            Real* const p0 = sm + tid;
            const Real emis = a.sfc_emis[band + (size_t)P.nbnd * cgs];
            Real rad = 0;                       // incident diffuse flux = 0
            p0[W3] = rad;
            {
                Real* p = p0;
                int k = 0;
                for (; k + 8 <= nlay; k += 8, p += 8 * CR_RS)
                {
                    Real tr[8], sd[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) { tr[u] = p[u * CR_RS + W2]; sd[u] = p[(u + 1) * CR_RS + W3]; }
#pragma unroll
                    for (int u = 0; u < 8; ++u) { rad = fma(tr[u], rad, sd[u]); p[(u + 1) * CR_RS + W3] = rad; }
                }
                for (; k < nlay; ++k, p += CR_RS) { rad = fma(p[W2], rad, p[CR_RS + W3]); p[CR_RS + W3] = rad; }
            }
            rad = rad * ((Real)1 - emis) + emis * p0[nlay * CR_RS + W2];
            p0[nlay * CR_RS + W0] = rad;
            {
                Real* p = p0 + (nlay - 1) * CR_RS;
                int k = nlay - 1;
                for (; k >= 7; k -= 8, p -= 8 * CR_RS)
                {
                    Real tr[8], su[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) { tr[u] = p[-u * CR_RS + W2]; su[u] = p[-u * CR_RS + W0]; }
#pragma unroll
                    for (int u = 0; u < 8; ++u) { rad = fma(tr[u], rad, su[u]); p[-u * CR_RS + W0] = rad; }
                }
                for (; k >= 0; --k, p -= CR_RS) { rad = fma(p[W2], rad, p[W0]); p[W0] = rad; }
            }
        }
        __syncthreads();

        // ---- phase 3: g-point sums per level
        for (int j0 = 0; j0 < nlev; j0 += CR_NL)
        {
            const int jr = j0 + lsub;
            const bool valid = jr < nlev;
            const int j = valid ? jr : nlev - 1;
            const Real* r = sm + j * CR_RS + quad * 4;
            const V4<Real> vu = lds4(r + W0), vd = lds4(r + W3);
            const Real au[4] = {vu.x, vu.y, vu.z, vu.w}, ad[4] = {vd.x, vd.y, vd.z, vd.w};
            const Real su = quad_sum<Real, LPC>(au) * piw, sd = quad_sum<Real, LPC>(ad) * piw;
            if (valid && q == 0 && active)
            {
                Real* ac = acc + (size_t)cb * 2 * nlev;
                ac[j] += su; ac[nlev + j] += sd;
                if (a.bnd_up)
                {
                    const size_t o = cg + (size_t)a.ncol * (top ? j : nlay - j) + nlev_stride * band;
                    a.bnd_up[o] = su; a.bnd_dn[o] = sd;
                }
            }
        }
    }
    __syncthreads();
    for (int i = tid; i < NCB * nlev; i += CR_T)
    {
        const int cbi = i / nlev, j = i - cbi * nlev;
        const int cc = blockIdx.x * NCB + cbi;
        if (cc < nc)
        {
            const size_t o = (size_t)a.c0 + cc + (size_t)a.ncol * (top ? j : nlay - j);
            const Real* ac = acc + (size_t)cbi * 2 * nlev;
            o_up[o] = ac[j]; o_dn[o] = ac[nlev + j];
        }
    }
}

// ------------------------------------------------------------------------------------ launchers
template<typename Real> size_t colres_smem_bytes(int nlay, int gw, bool sw)
{
    const int nlev = nlay + 1, ncb = CR_CH / gw;
    return (size_t)(nlev * CR_RS + (sw ? 3 : 2) * ncb * nlev) * sizeof(Real);
}

// flux(c, j) = sum over y of part[y][field](c, j), y ascending (= band order), for the nc columns of the tile
template<typename Real>
__global__ void __launch_bounds__(256)
sum_parts_kernel(const Real* __restrict__ part, const int nsplit, const int nfield, const int ncol, const int c0, const int nc,
                 const int nlev, Real* __restrict__ up, Real* __restrict__ dn, Real* __restrict__ dir)
{
    const size_t n = (size_t)nc * nlev, plane = (size_t)ncol * nlev;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n * nfield; i += (size_t)gridDim.x * blockDim.x)
    {
        const int f = (int)(i / n);
        const size_t r = i - (size_t)f * n;
        const size_t o = (size_t)c0 + r % nc + (size_t)ncol * (r / nc);
        Real acc = 0;
        for (int y = 0; y < nsplit; ++y) acc += part[((size_t)y * 3 + f) * plane + o];
        Real* dst = f == 0 ? up : (f == 1 ? dn : dir);
        if (dst) dst[o] = acc;
    }
}

template<typename K> static void cr_launch(K kern, dim3 grid, size_t smem, cudaStream_t s,
                                           const void* P, const void* bands, const void* st, const void* a)
{
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    void* args[4] = {const_cast<void*>(P), const_cast<void*>(bands), const_cast<void*>(st), const_cast<void*>(a)};
    cudaLaunchKernel((const void*)kern, grid, dim3(CR_T), args, smem, s);
}

// A call of fewer column groups than this leaves SMs idle with one CTA per group looping over every band (100 RFMIP columns:
// 100 CTAs, 16 bands in series each); its bands are spread over the grid's y dimension instead.
int colres_band_split(int ncol, int gw, int nbnd)
{
    static std::atomic<int> sms{0};
    int n = sms.load();
    if (!n)
    {
        int dev = 0; cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        sms.store(n);
    }
    const int ncb = CR_CH / gw, groups = (ncol + ncb - 1) / ncb;
    return groups < 4 * n ? nbnd : 1;
}

template<typename Real>
void launch_cols(bool sw, const DevPlan<Real>& P, const BandArgs* bands, const TileState<Real>& st,
                 const FusedArgs<Real>& a, cudaStream_t s)
{
    const int lpc = P.gw / 4, ncb = CR_CH / P.gw;
    const int nsplit = a.part ? a.nsplit : 1;
    const dim3 grid((unsigned)((a.nc + ncb - 1) / ncb), (unsigned)nsplit);
    const size_t smem = colres_smem_bytes<Real>(a.nlay, P.gw, sw);
    const bool cloud = a.cld_tau != nullptr;
    const void* pb = &bands;
#define RB_GO(K_) cr_launch(K_, grid, smem, s, &P, pb, &st, &a)
#define RB_NE(KN_, LPC_, CL_) { if (P.neta == 9) RB_GO((KN_<Real, LPC_, CL_, 9>)); else RB_GO((KN_<Real, LPC_, CL_, 0>)); }
#define RB_ALL(KN_) { if (lpc == 4) { if (cloud) RB_NE(KN_, 4, true) else RB_NE(KN_, 4, false) } \
                      else { if (cloud) RB_NE(KN_, 2, true) else RB_NE(KN_, 2, false) } }
    if (sw) RB_ALL(sw_col_kernel) else RB_ALL(lw_col_kernel)
#undef RB_ALL
#undef RB_NE
#undef RB_GO
    count_launch();
    if (nsplit > 1)
    {
        const int nlev = a.nlay + 1, nfield = sw ? 3 : 2;
        const size_t n = (size_t)a.nc * nlev * nfield;
        const unsigned blocks = (unsigned)std::min<size_t>((n + 255) / 256, 4096);
        sum_parts_kernel<Real><<<blocks, 256, 0, s>>>(a.part, nsplit, nfield, a.ncol, a.c0, a.nc, nlev, a.flux_up, a.flux_dn,
                                                      sw ? a.flux_dir : nullptr);
        count_launch();
    }
}

template size_t colres_smem_bytes<float>(int, int, bool);
template size_t colres_smem_bytes<double>(int, int, bool);
template void launch_cols<float>(bool, const DevPlan<float>&, const BandArgs*, const TileState<float>&, const FusedArgs<float>&, cudaStream_t);
template void launch_cols<double>(bool, const DevPlan<double>&, const BandArgs*, const TileState<double>&, const FusedArgs<double>&, cudaStream_t);

}  // namespace rb200

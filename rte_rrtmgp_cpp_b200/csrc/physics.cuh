// physics.cuh — per-cell device functions shared by the fused and the materialised kernels.
// A thread owns one column and NG=4 consecutive g-points of one band ("quad"); the tables are
// band-blocked with the g-point fastest (plan.hpp), so every interpolation corner is one
// aligned 4-wide vector load.
#pragma once
#include "common.cuh"
#include "launch.hpp"

namespace rb200 {

// Compile-time table geometry of a kernel instantiation: GW padded g-points per band,
// NE mixing-fraction points (0 = take it from the plan at run time), RS row stride in elements
// (2*GW when planck_frac rows are interleaved with kmajor rows, GW otherwise).
template<int GW_, int NE_, bool LW_> struct Geo
{
    static constexpr int GW = GW_, NE = NE_, RS = LW_ ? 2 * GW_ : GW_;
    template<typename Real> static __device__ __forceinline__ int neta(const DevPlan<Real>& P) { return NE_ > 0 ? NE_ : P.neta; }
};

// decoded (column, layer, band) interpolation state
template<typename Real> struct Cell
{
    Real w[8];        // fmajor: [it*4 + ip*2 + ie]
    Real fm[4];       // fminor: [it*2 + ie]
    Real cm0, cm1;    // col_mix
    Real wet;         // col_gas(h2o) + col_dry (Rayleigh)
    int maj0, maj1;   // row numbers of the (it, ip=0, ie=0) corners in a band's major block
    int min0, min1;   // row numbers of the (it, ie=0) corners in a contributor / Rayleigh block
    int side;         // 0 lower (tropo), 1 upper
};

template<typename Real>
__device__ __forceinline__ int cell_side(const V4<Real>& cell) { return ((M<Real>::as_int(cell.w) >> 16) & 1) ? 0 : 1; }

template<typename Real>
__device__ __forceinline__ void decode_cell(const V4<Real>& cell, const V4<Real>& fa, const V4<Real>& fb, Cell<Real>& o)
{
    o.side = cell_side(cell);
    const int r0 = M<Real>::as_int(fb.z), r1 = M<Real>::as_int(fb.w);
    o.maj0 = r0 & 0xfffff; o.min0 = (r0 >> 20) & 0xfff;
    o.maj1 = r1 & 0xfffff; o.min1 = (r1 >> 20) & 0xfff;
    const Real fpress = cell.x, omf = (Real)1 - fpress;
    // cell.z is 0: the addition is exact, and it makes every lane of the 128-bit load READ here.  A lane nothing reads is
    // handed to another value by ptxas while the load is still in flight, and that write waits for the whole prefetch.
    o.wet = cell.y + cell.z;
    o.cm0 = fa.x; o.cm1 = fa.y;
    o.fm[0] = fa.z; o.fm[1] = fa.w; o.fm[2] = fb.x; o.fm[3] = fb.y;
    const Pr<Real> f01(o.fm[0], o.fm[1]), f23(o.fm[2], o.fm[3]);
    const Pr<Real> w01 = omf * f01, w23 = fpress * f01, w45 = omf * f23, w67 = fpress * f23;
    o.w[0] = w01.lo(); o.w[1] = w01.hi(); o.w[2] = w23.lo(); o.w[3] = w23.hi();
    o.w[4] = w45.lo(); o.w[5] = w45.hi(); o.w[6] = w67.lo(); o.w[7] = w67.hi();
}

// raw per-(column, layer) state of one band: what the prologue kernel left in HBM
template<typename Real> struct RawState { V4<Real> cell, fa, fb, sc0, sc1; };

// flavor of layer l from the column's tropopause bit mask (no dependent load on `cell`)
__device__ __forceinline__ int layer_side(const unsigned* __restrict__ tm, int l)
{
    return ((__ldg(tm + (l >> 5)) >> (l & 31)) & 1u) ? 0 : 1;
}

// scalings of contributors m0..m0+3 of the band on this side: one vector from the group's plane
// (planes are [group][l][c], so a warp's columns read one contiguous 128-byte line)
template<typename Real>
__device__ __forceinline__ void load_scal4(const V4<Real>* __restrict__ scal_icl, size_t tcl, const BandArgs& B, int side, int m0, V4<Real>& out)
{
    out = scal_icl[(size_t)(B.scal_off[side] + (m0 >> 2)) * tcl];
}

template<typename Real>
__device__ __forceinline__ RawState<Real> load_state(const TileState<Real>& st, const DevPlan<Real>& P, const BandArgs& B,
                                                    const unsigned* tm, int flav0, int flav1, int c, int l, size_t tcl)
{
    const size_t icl = (size_t)c + (size_t)st.nc * l;
    const int side = layer_side(tm, l);
    const int f = side == 0 ? flav0 : flav1;
    RawState<Real> r;
    load_scal4(st.scal + icl, tcl, B, side, 0, r.sc0);
    load_scal4(st.scal + icl, tcl, B, side, B.nminor[side] > 4 ? 4 : 0, r.sc1);   // a second load, never a copy of sc0
    r.cell = st.cell[icl];
    r.fa = st.flvA[icl + tcl * f];
    r.fb = st.flvB[icl + tcl * f];
    return r;
}

// ---- L2 prefetch ("prefetch.global.L2": no register, no scoreboard) ------------------------------------------
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" :: "l"(p)); }

// The per-layer state of a warp's columns is one run of 32/LPC * sizeof(V4) bytes per array (a few 128-byte lines): lane i
// of a warp requests ONE line of ONE array, so the whole warp prefetches a layer's state with a single instruction.
// Arrays: cell, flvA, flvB, the band's first two scaling groups, and (longwave) the two Planck-position arrays.
// Measured on B200 (1024^2 x 128): longwave (issue/latency-bound) -3 %, shortwave (DRAM-bound) -1.5 % for the state lines and
// slower for the strip lines (RB_LW_PFS / RB_SW_PFS = layers ahead of the register prefetch, RB_*_PF2 = strip layers, 0 = off).
template<typename Real, int LPC, bool LW> struct StatePrefetch
{
    const char* base; unsigned lstep, off0, off1; bool on;           // layer step in units of 8 bytes, plane offsets in units of 16
    __device__ __forceinline__ void init(const TileState<Real>& st, const BandArgs& B, int flav0, int flav1, int c, size_t tcl)
    {
        constexpr int LINES = (32 / LPC) * (int)sizeof(V4<Real>) / 128;       // lines per array per warp (V4 arrays)
        static_assert(LINES >= 1, "a warp's state run must be at least one line");
        constexpr int NARR = LW ? 7 : 5;
        constexpr unsigned U4 = sizeof(V4<Real>) / 8, U2 = sizeof(V2<Real>) / 8;
        const int lane = threadIdx.x & 31, arr = lane / LINES, sub = lane % LINES;
        on = arr < NARR;
        const size_t csh = (size_t)(c - (c % (32 / LPC)));                    // the warp's first column
        lstep = (unsigned)st.nc * U4; off0 = off1 = 0;
        const unsigned t = (unsigned)tcl * (U4 / 2);
        if ((size_t)tcl * (U4 / 2) * (size_t)(arr < 3 ? max(flav0, flav1) : max(B.scal_off[0], B.scal_off[1]) + 1) > 0xffffffffull) on = false;
        switch (arr)
        {
        case 0: base = (const char*)(st.cell + csh); break;
        case 1: base = (const char*)(st.flvA + csh); off0 = t * flav0; off1 = t * flav1; break;
        case 2: base = (const char*)(st.flvB + csh); off0 = t * flav0; off1 = t * flav1; break;
        case 3: base = (const char*)(st.scal + csh); off0 = t * B.scal_off[0]; off1 = t * B.scal_off[1]; break;
        case 4: base = (const char*)(st.scal + csh);
                off0 = t * (B.scal_off[0] + (B.nminor[0] > 4 ? 1 : 0)); off1 = t * (B.scal_off[1] + (B.nminor[1] > 4 ? 1 : 0)); break;
        case 5: base = (const char*)(st.pl_lay + csh); lstep = (unsigned)st.nc * U2; on = on && 2 * sub < LINES + 1; break;
        case 6: base = (const char*)(st.pl_lev + csh); lstep = (unsigned)st.nc * U2; on = on && 2 * sub < LINES + 1; break;
        default: base = (const char*)st.cell; break;
        }
        base += (size_t)sub * 128;
    }
    // layer l of the tile (array order); `side` = 0 lower / 1 upper atmosphere of that layer
    __device__ __forceinline__ void layer(int l, int side) const
    {
        if (on) prefetch_l2(base + ((size_t)((unsigned)l * lstep) << 3) + ((size_t)(side ? off1 : off0) << 4));
    }
};

// scratch strip of a CTA: layer k is W*T vectors = W*T*sizeof(V4)/128 lines; thread t requests lines t, t+T, ...
template<typename Real, int W>
__device__ __forceinline__ void prefetch_strip_layer(const V4<Real>* strip_thread0, int k)
{
    constexpr int T = FUSED_THREADS;
    constexpr int LINES = W * T * (int)sizeof(V4<Real>) / 128;
    const char* p = (const char*)(strip_thread0 + (size_t)k * W * T);
#pragma unroll
    for (int i = 0; i < (LINES + T - 1) / T; ++i)
    {
        const int line = (int)threadIdx.x + i * T;
        if (line < LINES) prefetch_l2(p + (size_t)line * 128);
    }
}

// the same with the plane offsets of the band launch precomputed on the host (BandLaunch)
template<typename Real>
__device__ __forceinline__ RawState<Real> load_state(const TileState<Real>& st, const BandLaunch<Real>& L, const unsigned* tm, int c, int l)
{
    const size_t icl = (size_t)c + (size_t)st.nc * l;
    const int side = layer_side(tm, l);
    RawState<Real> r;
    r.sc0 = st.scal[icl + L.scal0[side]];
    r.sc1 = st.scal[icl + L.scal1[side]];
    r.cell = st.cell[icl];
    r.fa = st.flvA[icl + L.flv[side]];
    r.fb = st.flvB[icl + L.flv[side]];
    return r;
}

// acc (+)= s * v on a thread's quad: two packed instructions in fp32 (the scalar weight is a broadcast operand)
template<typename Real> __device__ __forceinline__ void q_fma(Real (&acc)[4], Real s, const V4<Real>& v)
{
    pput(acc, 0, pfma(s, plo(v), pget(acc, 0))); pput(acc, 1, pfma(s, phi(v), pget(acc, 1)));
}
template<typename Real> __device__ __forceinline__ void q_mul(Real (&acc)[4], Real s, const V4<Real>& v)
{
    pput(acc, 0, s * plo(v)); pput(acc, 1, s * phi(v));
}
template<typename Real> __device__ __forceinline__ void q_fma(Real (&acc)[4], Real s, const Real (&v)[4])
{
    pput(acc, 0, pfma(s, pget(v, 0), pget(acc, 0))); pput(acc, 1, pfma(s, pget(v, 1), pget(acc, 1)));
}
// ---- L1 prefetch of the NEXT layer's table rows ("prefetch.global.L1": no register, no scoreboard) ---------------------
// MEASURED SLOWER, off by default (-DRB_TAB_PF=1 enables it): step 569 -> 934 ms on B200.  The band kernels are bound by
// the L1TEX LSU data pipe (ncu: l1tex__data_pipe_lsu_wavefronts 75 % of peak in the longwave kernel), not by the latency
// of the gathers, and these uncoalesced requests (one line per lane) add as many L1 wavefronts as the gathers themselves.
// Mechanism kept for the record: the packed row numbers of the layer AFTER next arrive together with the next layer's
// state (same scoreboard wait); at the top of a layer step the lanes of a column request the lines the NEXT step gathers.
#ifndef RB_TAB_PF
#define RB_TAB_PF 0
#endif
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" :: "l"(p)); }

// {rows(itemp 1), rows(itemp 2)} of layer l for the band's flavor on that layer's side: the .z/.w half of flvB
template<typename Real>
__device__ __forceinline__ V2<Real> load_rows(const TileState<Real>& st, const BandLaunch<Real>& L, const unsigned* tm, int c, int l)
{
    const size_t icl = (size_t)c + (size_t)st.nc * l;
    return reinterpret_cast<const V2<Real>*>(st.flvB + icl + L.flv[layer_side(tm, l)])[1];
}

// lane q of the LPC lanes of a column requests its share of the lines; `side` is the side of the layer the rows belong to
template<typename G, int LPC, bool RAYLEIGH, typename Real>
__device__ __forceinline__ void prefetch_tables(const BandLaunch<Real>& L, const DevPlan<Real>& P, const V2<Real>& rows, int side, int q)
{
    const int r0 = M<Real>::as_int(rows.x), r1 = M<Real>::as_int(rows.y);
    const int dp = G::neta(P);
    // major block: 8 corner rows (itemp, ipress, ieta); a row is RS elements (k, and the Planck fraction behind it in LW)
#pragma unroll
    for (int j = 0; j < 8 / LPC; ++j)
    {
        const int i = q + j * LPC;
        const int row = ((i & 4) ? (r1 & 0xfffff) : (r0 & 0xfffff)) + ((i >> 1) & 1) * dp + (i & 1);
        const char* p = (const char*)(L.kmaj + (size_t)((unsigned)row * (unsigned)G::RS));
        prefetch_l1(p);
        if (G::RS * sizeof(Real) > 128) prefetch_l1(p + 128);
    }
    // minor contributors of this side: per contributor two row pairs (itemp 1 / 2), the two eta rows of a pair adjacent
    const int m0 = (r0 >> 20) & 0xfff, m1 = (r1 >> 20) & 0xfff;
    const int npairs = 2 * L.B.nminor[side];
    const Real* km = L.kmin[side];
    for (int p2 = q; p2 < npairs; p2 += LPC)
    {
        const Real* t = km + (size_t)(p2 >> 1) * L.mstride + ((p2 & 1) ? m1 : m0) * G::GW;
        prefetch_l1(t); prefetch_l1(t + G::GW);
    }
    if (RAYLEIGH && L.kray[0] && q < 2)
    {
        const Real* t = L.kray[side] + (q ? m1 : m0) * G::GW;
        prefetch_l1(t); prefetch_l1(t + G::GW);
    }
}

#define RB_FMA4(acc, s, v) q_fma(acc, s, v);
#define RB_MUL4(acc, s, v) q_mul(acc, s, v);

// interpolate3D_byflav (mo_gas_optics_rrtmgp_kernels.F90:760-789): partial sums per temperature.
// `tab` points at the band block (+ slot and quad offset); with G::NE > 0 the 4 corners of a
// temperature are at compile-time byte offsets from one computed address.
template<typename G, typename Real>
__device__ __forceinline__ void interp_major(const Real* __restrict__ tab, const DevPlan<Real>& P,
                                             const Cell<Real>& c, Real (&s0)[4], Real (&s1)[4])
{
    const int de = G::RS, dp = G::neta(P) * G::RS;
    const Real* t0 = tab + (unsigned)(c.maj0 * G::RS);
    const Real* t1 = tab + (unsigned)(c.maj1 * G::RS);
    const V4<Real> a0 = ldg4(t0), a1 = ldg4(t0 + de), a2 = ldg4(t0 + dp), a3 = ldg4(t0 + dp + de);
    const V4<Real> b0 = ldg4(t1), b1 = ldg4(t1 + de), b2 = ldg4(t1 + dp), b3 = ldg4(t1 + dp + de);
    RB_MUL4(s0, c.w[0], a0) RB_FMA4(s0, c.w[1], a1) RB_FMA4(s0, c.w[2], a2) RB_FMA4(s0, c.w[3], a3)
    RB_MUL4(s1, c.w[4], b0) RB_FMA4(s1, c.w[5], b1) RB_FMA4(s1, c.w[6], b2) RB_FMA4(s1, c.w[7], b3)
}

// kmajor and planck_frac of the same 8 corners (interleaved rows, longwave)
template<typename G, typename Real>
__device__ __forceinline__ void interp_major_pair(const Real* __restrict__ tab, const DevPlan<Real>& P, const Cell<Real>& c,
                                                  Real (&s0)[4], Real (&s1)[4], Real (&p0)[4], Real (&p1)[4])
{
    const int de = G::RS, dp = G::neta(P) * G::RS, pf = G::GW;
    const Real* t0 = tab + (unsigned)(c.maj0 * G::RS);
    const Real* t1 = tab + (unsigned)(c.maj1 * G::RS);
    {
        const V4<Real> a0 = ldg4(t0), a1 = ldg4(t0 + de), a2 = ldg4(t0 + dp), a3 = ldg4(t0 + dp + de);
        const V4<Real> f0 = ldg4(t0 + pf), f1 = ldg4(t0 + de + pf), f2 = ldg4(t0 + dp + pf), f3 = ldg4(t0 + dp + de + pf);
        RB_MUL4(s0, c.w[0], a0) RB_FMA4(s0, c.w[1], a1) RB_FMA4(s0, c.w[2], a2) RB_FMA4(s0, c.w[3], a3)
        RB_MUL4(p0, c.w[0], f0) RB_FMA4(p0, c.w[1], f1) RB_FMA4(p0, c.w[2], f2) RB_FMA4(p0, c.w[3], f3)
    }
    {
        const V4<Real> b0 = ldg4(t1), b1 = ldg4(t1 + de), b2 = ldg4(t1 + dp), b3 = ldg4(t1 + dp + de);
        const V4<Real> f0 = ldg4(t1 + pf), f1 = ldg4(t1 + de + pf), f2 = ldg4(t1 + dp + pf), f3 = ldg4(t1 + dp + de + pf);
        RB_MUL4(s1, c.w[4], b0) RB_FMA4(s1, c.w[5], b1) RB_FMA4(s1, c.w[6], b2) RB_FMA4(s1, c.w[7], b3)
        RB_MUL4(p1, c.w[4], f0) RB_FMA4(p1, c.w[5], f1) RB_FMA4(p1, c.w[6], f2) RB_FMA4(p1, c.w[7], f3)
    }
}

// interpolate2D_byflav (F90:738-758); minor / Rayleigh blocks have row stride GW
template<typename G, typename Real>
__device__ __forceinline__ void interp_minor(const Real* __restrict__ tab, const Cell<Real>& c, Real (&s)[4])
{
    const Real* t0 = tab + c.min0 * G::GW;
    const Real* t1 = tab + c.min1 * G::GW;
    const V4<Real> a0 = ldg4(t0), a1 = ldg4(t0 + G::GW), b0 = ldg4(t1), b1 = ldg4(t1 + G::GW);
    RB_MUL4(s, c.fm[0], a0) RB_FMA4(s, c.fm[1], a1) RB_FMA4(s, c.fm[2], b0) RB_FMA4(s, c.fm[3], b1)
}

// minor-gas part of compute_tau_absorption (F90:404-502) for the layer's side, added onto tau.
// The band's contributors are consecutive table blocks and consecutive entries of the scaling
// scalings the prologue kernel prepared; `pre0/pre1` hold those of the first 8 contributors (fetched a layer ahead).
template<typename G, typename Real>
__device__ __forceinline__ void add_minor_tau(const DevPlan<Real>& P, const BandArgs& B, const Cell<Real>& c, int q4,
                                              const V4<Real>& pre0, const V4<Real>& pre1,
                                              const V4<Real>* __restrict__ scal_icl, size_t tcl, Real (&tau)[4])
{
    const int side = c.side;
    const int nm = B.nminor[side];
    const size_t mstride = (size_t)P.ntemp * G::neta(P) * G::GW;
    const Real* km = P.kminor[side] + mstride * B.first_minor[side] + q4;
    for (int m0 = 0; m0 < nm; m0 += 4)
    {
        V4<Real> sc;
        if (m0 == 0) sc = pre0; else if (m0 == 4) sc = pre1; else load_scal4(scal_icl, tcl, B, side, m0, sc);
        const Real* scp = &sc.x;
#pragma unroll
        for (int u = 0; u < 4; ++u)
        {
            if (m0 + u < nm)
            {
                Real s[4];
                interp_minor<G>(km, c, s);
                q_fma(tau, scp[u], s);
                km += mstride;
            }
        }
    }
}

template<typename G, typename Real>
__device__ __forceinline__ void add_minor_tau(const BandLaunch<Real>& L, const Cell<Real>& c, int q4,
                                              const V4<Real>& pre0, const V4<Real>& pre1,
                                              const V4<Real>* __restrict__ scal_icl, Real (&tau)[4])
{
    const int side = c.side;
    const int nm = L.B.nminor[side];
    const Real* km = L.kmin[side] + q4;
    for (int m0 = 0; m0 < nm; m0 += 4)
    {
        V4<Real> sc;
        if (m0 == 0) sc = pre0; else if (m0 == 4) sc = pre1; else load_scal4(scal_icl, L.tcl, L.B, side, m0, sc);
        const Real* scp = &sc.x;
#pragma unroll
        for (int u = 0; u < 4; ++u)
        {
            if (m0 + u < nm)
            {
                Real s[4];
                interp_minor<G>(km, c, s);
                q_fma(tau, scp[u], s);
                km += L.mstride;
            }
        }
    }
}

// gas absorption optical depth of 4 g-points: major + minor (compute_tau_absorption, F90:178-502)
template<typename G, typename Real>
__device__ __forceinline__ void gas_tau(const DevPlan<Real>& P, const BandArgs& B, const Cell<Real>& c,
                                        const Real* __restrict__ kmaj_quad, int q4,
                                        const V4<Real>& pre0, const V4<Real>& pre1,
                                        const V4<Real>* __restrict__ scal_icl, size_t tcl, Real (&tau)[4])
{
    Real s0[4], s1[4];
    interp_major<G>(kmaj_quad, P, c, s0, s1);
#pragma unroll
    for (int h = 0; h < 2; ++h) pput(tau, h, pfma(c.cm1, pget(s1, h), c.cm0 * pget(s0, h)));
    add_minor_tau<G>(P, B, c, q4, pre0, pre1, scal_icl, tcl, tau);
}

// interpolate1D (F90:715-734) for one band
template<typename Real>
__device__ __forceinline__ Real planck_band(const Real* __restrict__ tab, int n, Real tmin, Real delta, Real T)
{
    const Real val0 = M<Real>::div(T - tmin, delta);
    const int iv = (int)val0;
    const Real frac = val0 - (Real)iv;
    const int index = min(n - 1, max(1, iv + 1));
    const Real t0 = tab[index - 1], t1 = tab[index];
    return t0 + frac * (t1 - t0);
}
// the same from a position {frac, lower index} precomputed by the prologue kernel
template<typename Real>
__device__ __forceinline__ Real planck_at(const Real* __restrict__ tab, const V2<Real>& pos)
{
    const int i = M<Real>::as_int(pos.y);
    const Real t0 = tab[i], t1 = tab[i + 1];
    return t0 + pos.x * (t1 - t0);
}

// A resident CTA claims one scratch strip for its lifetime (grids are larger than one wave, so
// strips are recycled as CTAs retire; which strip a CTA gets does not affect any result).
__device__ __forceinline__ int claim_slot(int* busy, int nslots, int bid)
{
    __shared__ int s_slot;
    if (threadIdx.x == 0)
    {
        int s = (int)(((unsigned)bid * 2654435761u) % (unsigned)nslots);
        while (atomicCAS(&busy[s], 0, 1) != 0) s = (s + 1 == nslots) ? 0 : s + 1;
        s_slot = s;
    }
    __syncthreads();
    return s_slot;
}
__device__ __forceinline__ void release_slot(int* busy, int slot)
{
    __syncthreads();
    if (threadIdx.x == 0) atomicExch(&busy[slot], 0);
}

// sum of a quad's 4 values across the LPC lanes of a column; result valid in every lane
template<typename Real, int LPC>
__device__ __forceinline__ Real quad_sum(const Real (&v)[4])
{
    const Pr<Real> h = pget(v, 0) + pget(v, 1);          // (v0 + v2, v1 + v3): one packed add
    Real s = h.lo() + h.hi();
    if (LPC >= 2) s += __shfl_xor_sync(0xffffffffu, s, 1);
    if (LPC >= 4) s += __shfl_xor_sync(0xffffffffu, s, 2);
    return s;
}

// lw_source_noscat + one transport step (mo_rte_solver_kernels.F90:620-708) for one g-point
template<typename Real>
__device__ __forceinline__ void lw_layer(Real tau_loc, Real lay_src, Real lev_above, Real lev_below,
                                         Real& trans, Real& src_up, Real& src_dn)
{
    // tau_thresh = sqrt(sqrt(epsilon))
    const Real tau_thresh = sizeof(Real) == 4 ? (Real)1.857e-2 : (Real)1.220703125e-4;
    trans = M<Real>::exp(-tau_loc);
    const Real omt = (Real)1 - trans;
    const Real f_big = M<Real>::div(omt, M<Real>::max(tau_loc, tau_thresh)) - trans;
    const Real f_small = tau_loc * ((Real)0.5 + tau_loc * ((Real)(-1.0 / 3.0) + tau_loc * (Real)0.125));
    const Real fact = tau_loc > tau_thresh ? f_big : f_small;
    src_dn = omt * lev_below + (Real)2 * fact * (lay_src - lev_below);
    src_up = omt * lev_above + (Real)2 * fact * (lay_src - lev_above);
}

// the same for a pair of g-points (packed arithmetic in fp32; exp / rcp / select per lane)
template<typename Real>
__device__ __forceinline__ void lw_layer2(Pr<Real> tau_loc, Pr<Real> lay_src, Pr<Real> lev_above, Pr<Real> lev_below,
                                          Pr<Real>& trans, Pr<Real>& src_up, Pr<Real>& src_dn)
{
    const Real tau_thresh = sizeof(Real) == 4 ? (Real)1.857e-2 : (Real)1.220703125e-4;
    trans = pexp(-tau_loc);
    const Pr<Real> omt = (Real)1 - trans;
    const Pr<Real> f_big = omt * prcp(pmax(tau_loc, tau_thresh)) - trans;
    const Pr<Real> f_small = tau_loc * pfma(tau_loc, pfma(tau_loc, (Real)0.125, (Real)(-1.0 / 3.0)), (Real)0.5);
    const Pr<Real> fact(tau_loc.lo() > tau_thresh ? f_big.lo() : f_small.lo(), tau_loc.hi() > tau_thresh ? f_big.hi() : f_small.hi());
    const Pr<Real> fact2 = fact + fact;
    src_dn = pfma(fact2, lay_src - lev_below, omt * lev_below);
    src_up = pfma(fact2, lay_src - lev_above, omt * lev_above);
}

}  // namespace rb200

namespace rb200 {

// combine_abs_and_rayleigh (src/Gas_optics_rrtmgp.cpp:365-385): ssa threshold 2*epsilon
template<typename Real>
__device__ __forceinline__ void combine_rayleigh(Real tau_abs, Real tau_ray, Real& tau, Real& ssa)
{
    tau = tau_abs + tau_ray;
    // CPU reference: 2*epsilon (fp64).  In the fp32 build 2*epsilon(float) would zero the
    // single-scattering albedo of every cell thinner than 2.4e-7; use the reference CUDA path's
    // 2*tiny there (src_kernels_cuda/gas_optics_rrtmgp_kernels.cu:721-746).
    const Real thr = sizeof(Real) == 4 ? (Real)2 * M<Real>::tiny : (Real)2 * M<Real>::eps;
    ssa = (tau > thr) ? M<Real>::div(tau_ray, tau) : (Real)0;
}

// inc_2stream_by_2stream_bybnd for one cell (mo_optical_props_kernels.F90:451-482), eps = 3*tiny
template<typename Real>
__device__ __forceinline__ void increment_2str(Real& tau, Real& ssa, Real& g, Real t2, Real s2, Real g2)
{
    const Real eps3 = (Real)3 * M<Real>::tiny;
    const Real tau12 = tau + t2;
    const Real ts1 = tau * ssa, ts2 = t2 * s2;
    const Real tauscat12 = ts1 + ts2;
    g = M<Real>::div(ts1 * g + ts2 * g2, M<Real>::max(eps3, tauscat12));
    ssa = M<Real>::div(tauscat12, M<Real>::max(eps3, tau12));
    tau = tau12;
}

// delta_scale_2str_k for one cell (mo_optical_props_kernels.F90:76-98)
template<typename Real>
__device__ __forceinline__ void delta_scale_cell(Real& tau, Real& ssa, Real& g)
{
    const Real eps3 = (Real)3 * M<Real>::tiny;
    const Real f = g * g, wf = ssa * f;
    tau = ((Real)1 - wf) * tau;
    ssa = M<Real>::div(ssa - wf, M<Real>::max(eps3, (Real)1 - wf));
    g = M<Real>::div(g - f, M<Real>::max(eps3, (Real)1 - f));
}

// sw_dif_and_source for one cell (mo_rte_solver_kernels.F90:985-1127): two-stream coefficients
// (Zdunkowski PIFM; Meador & Weaver eqs. 14, 15, 25, 26) — mu0s is max(min_mu0, mu0).
template<typename Real>
__device__ __forceinline__ void two_stream_ref(Real tau, Real w0, Real g, Real mu0s,
                                           Real& Rdif, Real& Tdif, Real& Rdir, Real& Tdir, Real& Tnoscat)
{
    const Real min_k = (Real)1.e4 * M<Real>::eps;
    const Real gamma1 = ((Real)8 - w0 * ((Real)5 + (Real)3 * g)) * (Real).25;
    const Real gamma2 = (Real)3 * (w0 * ((Real)1 - g)) * (Real).25;
    const Real k = M<Real>::sqrt(M<Real>::max((gamma1 - gamma2) * (gamma1 + gamma2), min_k));
    const Real exp_minusktau = M<Real>::exp(-tau * k);
    const Real exp_minus2ktau = exp_minusktau * exp_minusktau;
    Real RT_term = M<Real>::rcp(k * ((Real)1 + exp_minus2ktau) + gamma1 * ((Real)1 - exp_minus2ktau));
    Rdif = RT_term * gamma2 * ((Real)1 - exp_minus2ktau);
    Tdif = RT_term * (Real)2 * k * exp_minusktau;

    const Real k_mu = k * mu0s;
    const Real d = (Real)1 - k_mu * k_mu;
    RT_term = M<Real>::div(w0 * RT_term, (M<Real>::fabs(d) >= M<Real>::eps) ? d : M<Real>::eps);
    const Real gamma3 = ((Real)2 - (Real)3 * mu0s * g) * (Real).25;
    const Real gamma4 = (Real)1 - gamma3;
    const Real alpha1 = gamma1 * gamma4 + gamma2 * gamma3;
    const Real alpha2 = gamma1 * gamma3 + gamma2 * gamma4;
    const Real k_gamma3 = k * gamma3, k_gamma4 = k * gamma4;
    Tnoscat = M<Real>::exp(-M<Real>::div(tau, mu0s));
    Rdir = RT_term * (((Real)1 - k_mu) * (alpha2 + k_gamma3)
                      - ((Real)1 + k_mu) * (alpha2 - k_gamma3) * exp_minus2ktau
                      - (Real)2 * (k_gamma3 - alpha2 * k_mu) * exp_minusktau * Tnoscat);
    Tdir = -RT_term * (((Real)1 + k_mu) * (alpha1 + k_gamma4) * Tnoscat
                       - ((Real)1 - k_mu) * (alpha1 - k_gamma4) * exp_minus2ktau * Tnoscat
                       - (Real)2 * (k_gamma4 + alpha1 * k_mu) * exp_minusktau);
    Rdir = M<Real>::max((Real)0, M<Real>::min(Rdir, (Real)1 - Tnoscat));
    Tdir = M<Real>::max((Real)0, M<Real>::min(Tdir, (Real)1 - Tnoscat - Rdir));
}

// (1 - exp(-x)) / x for x >= 0 given e = exp(-x): series below 0.03 (no cancellation)
__device__ __forceinline__ float omexp_over_x(float x, float e)
{
    const float ser = fmaf(-0.5f * x, fmaf(-(1.f / 3.f) * x, fmaf(-0.25f, x, 1.f), 1.f), 1.f);
    const float dir = M<float>::div(1.f - e, fmaxf(x, 1e-30f));
    return x < 0.03f ? ser : dir;
}

// fp32 two-stream: the same Meador-Weaver / PIFM expressions as two_stream_ref, algebraically
// rearranged so nothing is divided by k and no O(1) terms cancel to O(tau) or O(k):
//   P = (1-e2)/k = tau*g(a)*(1+e1),  g(a) = (1-e^-a)/a,  a = k*tau
//   D = (1+e2) + gamma1*P            (RT_term = 1/(k*D))
//   Rdif = gamma2*P/D, Tdif = 2*e1/D
//   Rdir =  w0*[alpha2*P + (gamma3 - mu*alpha2)*Q  - k^2*mu*gamma3*P   ] / (D*(1-k^2 mu^2))
//   Tdir = -w0*[T0*alpha1*P + (gamma4 + mu*alpha1)*Q' + k^2*mu*gamma4*P*T0] / (D*(1-k^2 mu^2))
//   Q = (1-e1)^2 + 2*e1*(1-T0),  Q' = T0*(1-e1)^2 - 2*e1*(1-T0)
// so k may go to zero (conservative scattering) without the reference's min_k = 1e4*eps floor,
// which in fp32 (k >= 0.035) biases thick near-conservative clouds by O(0.3 W/m2).  Within
// |1 - k^2 mu^2| < 1e-3 of the removable singularity mu is nudged to the edge of that band for the
// scattering terms only (the direct-beam transmission keeps the true mu).
__device__ __forceinline__ void two_stream_f32(float tau, float w0, float g, float mu0s,
                                               float& Rdif, float& Tdif, float& Rdir, float& Tdir, float& Tnoscat)
{
    const float gamma1 = (8.f - w0 * (5.f + 3.f * g)) * .25f;
    const float gamma2 = 3.f * (w0 * (1.f - g)) * .25f;
    const float k2 = fmaxf(2.f * (1.f - w0) * (2.f - 0.5f * w0 * (1.f + 3.f * g)), 0.f);   // (g1-g2)(g1+g2)
    const float k = M<float>::sqrt(k2);
    const float a = k * tau;
    const float e1 = M<float>::exp(-a), e2 = e1 * e1;
    const float ga = omexp_over_x(a, e1);
    const float om1 = a * ga;
    const float P = tau * ga * (1.f + e1);
    const float rD = M<float>::rcp((1.f + e2) + gamma1 * P);
    Rdif = gamma2 * P * rD;
    Tdif = 2.f * e1 * rD;

    const float x = M<float>::div(tau, mu0s);
    Tnoscat = M<float>::exp(-x);
    const float omT_true = x * omexp_over_x(x, Tnoscat);
    float mu = mu0s, T0 = Tnoscat, omT = omT_true;
    float kmu = k * mu, d = 1.f - kmu * kmu;
    const float band = 1.e-3f;
    if (fabsf(d) < band)
    {
        d = d >= 0.f ? band : -band;
        mu = M<float>::div(M<float>::sqrt(1.f - d), k);
        const float xe = M<float>::div(tau, mu);
        T0 = M<float>::exp(-xe);
        omT = xe * omexp_over_x(xe, T0);
    }
    const float gamma3 = (2.f - 3.f * mu * g) * .25f;
    const float gamma4 = 1.f - gamma3;
    const float alpha1 = gamma1 * gamma4 + gamma2 * gamma3;
    const float alpha2 = gamma1 * gamma3 + gamma2 * gamma4;
    const float Q = om1 * om1 + 2.f * e1 * omT;
    const float Qp = T0 * om1 * om1 - 2.f * e1 * omT;
    const float k2mu = k2 * mu;
    const float NR = alpha2 * P + (gamma3 - mu * alpha2) * Q - k2mu * gamma3 * P;
    const float NT = T0 * (alpha1 * P + k2mu * gamma4 * P) + (gamma4 + mu * alpha1) * Qp;
    const float f = M<float>::div(w0 * rD, d);
    Rdir = f * NR;
    Tdir = -f * NT;
    Rdir = fmaxf(0.f, fminf(Rdir, omT_true));
    Tdir = fmaxf(0.f, fminf(Tdir, omT_true - Rdir));
}

template<typename Real>
__device__ __forceinline__ void two_stream(Real tau, Real w0, Real g, Real mu0s,
                                           Real& Rdif, Real& Tdif, Real& Rdir, Real& Tdir, Real& Tnoscat)
{
    two_stream_ref<Real>(tau, w0, g, mu0s, Rdif, Tdif, Rdir, Tdir, Tnoscat);
}
#ifndef RB_F32_REFERENCE_TWOSTREAM
template<>
__device__ __forceinline__ void two_stream<float>(float tau, float w0, float g, float mu0s,
                                                  float& Rdif, float& Tdif, float& Rdir, float& Tdir, float& Tnoscat)
{
    two_stream_f32(tau, w0, g, mu0s, Rdif, Tdif, Rdir, Tdir, Tnoscat);
}
#endif

// ---- the same per-cell functions for a pair of g-points ---------------------------------------------------------------
template<typename Real>
__device__ __forceinline__ void combine_rayleigh2(Pr<Real> tau_abs, Pr<Real> tau_ray, Pr<Real>& tau, Pr<Real>& ssa)
{
    tau = tau_abs + tau_ray;
    const Real thr = sizeof(Real) == 4 ? (Real)2 * M<Real>::tiny : (Real)2 * M<Real>::eps;
    const Pr<Real> q = sizeof(Real) == 4 ? tau_ray * prcp(tau) : Pr<Real>(M<Real>::div(tau_ray.lo(), tau.lo()), M<Real>::div(tau_ray.hi(), tau.hi()));
    ssa = Pr<Real>(tau.lo() > thr ? q.lo() : (Real)0, tau.hi() > thr ? q.hi() : (Real)0);
}

template<typename Real>
__device__ __forceinline__ void increment_2str2(Pr<Real>& tau, Pr<Real>& ssa, Pr<Real>& g, Real t2, Real s2, Real g2)
{
    const Real eps3 = (Real)3 * M<Real>::tiny;
    const Real ts2 = t2 * s2;
    const Pr<Real> tau12 = tau + t2;
    const Pr<Real> ts1 = tau * ssa;
    const Pr<Real> tauscat12 = ts1 + ts2;
    const Pr<Real> num = pfma(ts1, g, ts2 * g2);
    const Pr<Real> d1 = pmax(tauscat12, eps3), d2 = pmax(tau12, eps3);
    g = Pr<Real>(M<Real>::div(num.lo(), d1.lo()), M<Real>::div(num.hi(), d1.hi()));
    ssa = Pr<Real>(M<Real>::div(tauscat12.lo(), d2.lo()), M<Real>::div(tauscat12.hi(), d2.hi()));
    tau = tau12;
}

// generic (fp64): the reference expressions per lane
template<typename Real>
__device__ __forceinline__ void two_stream2(Pr<Real> tau, Pr<Real> w0, Pr<Real> g, Real mu0s,
                                            Pr<Real>& Rdif, Pr<Real>& Tdif, Pr<Real>& Rdir, Pr<Real>& Tdir, Pr<Real>& Tnoscat)
{
    Real a0, a1, a2, a3, a4, b0, b1, b2, b3, b4;
    two_stream<Real>(tau.lo(), w0.lo(), g.lo(), mu0s, a0, a1, a2, a3, a4);
    two_stream<Real>(tau.hi(), w0.hi(), g.hi(), mu0s, b0, b1, b2, b3, b4);
    Rdif = Pr<Real>(a0, b0); Tdif = Pr<Real>(a1, b1); Rdir = Pr<Real>(a2, b2); Tdir = Pr<Real>(a3, b3); Tnoscat = Pr<Real>(a4, b4);
}

#ifndef RB_F32_REFERENCE_TWOSTREAM
// (1 - exp(-x)) / x for a pair, e = exp(-x)
__device__ __forceinline__ PrF omexp_over_x2(PrF x, PrF e)
{
    const PrF ser = pfma(x * -0.5f, pfma(x * -(1.f / 3.f), pfma(x, -0.25f, 1.f), 1.f), 1.f);
    const PrF dir = (1.f - e) * prcp(pmax(x, 1e-30f));
    return PrF(x.lo() < 0.03f ? ser.lo() : dir.lo(), x.hi() < 0.03f ? ser.hi() : dir.hi());
}
// one lane of the removable singularity |1 - k^2 mu^2| < band (two_stream_f32): rare, scalar
__device__ __forceinline__ void two_stream_nudge(float k, float tau, float& d, float& mu, float& T0, float& omT)
{
    const float band = 1.e-3f;
    d = d >= 0.f ? band : -band;
    mu = M<float>::div(M<float>::sqrt(1.f - d), k);
    const float xe = M<float>::div(tau, mu);
    T0 = M<float>::exp(-xe);
    omT = xe * omexp_over_x(xe, T0);
}
// fp32: two_stream_f32 on two g-points with packed arithmetic (same expressions, same order)
template<>
__device__ __forceinline__ void two_stream2<float>(PrF tau, PrF w0, PrF g, float mu0s,
                                                   PrF& Rdif, PrF& Tdif, PrF& Rdir, PrF& Tdir, PrF& Tnoscat)
{
    const PrF g3 = g * 3.f;
    const PrF gamma1 = pfma(w0 * -0.25f, g3 + 5.f, 2.f);                       // (8 - w0 (5 + 3g)) / 4
    const PrF gamma2 = (w0 * 0.75f) * (1.f - g);
    const PrF k2 = pmax((2.f * (1.f - w0)) * pfma(w0 * -0.5f, g3 + 1.f, 2.f), 0.f);
    const PrF k = psqrt(k2);
    const PrF a = k * tau;
    const PrF e1 = pexp(-a), e2 = e1 * e1;
    const PrF ga = omexp_over_x2(a, e1);
    const PrF om1 = a * ga;
    const PrF P = (tau * ga) * (1.f + e1);
    const PrF rD = prcp(pfma(gamma1, P, 1.f + e2));
    Rdif = (gamma2 * P) * rD;
    Tdif = (e1 + e1) * rD;

    const PrF x = tau * M<float>::rcp(mu0s);
    Tnoscat = pexp(-x);
    const PrF omT_true = x * omexp_over_x2(x, Tnoscat);
    const PrF kmu = k * mu0s;
    const PrF d0 = pfma(-kmu, kmu, 1.f);
    float dl = d0.lo(), dh = d0.hi(), mul = mu0s, muh = mu0s, T0l = Tnoscat.lo(), T0h = Tnoscat.hi(), oml = omT_true.lo(), omh = omT_true.hi();
    if (fabsf(dl) < 1.e-3f) two_stream_nudge(k.lo(), tau.lo(), dl, mul, T0l, oml);
    if (fabsf(dh) < 1.e-3f) two_stream_nudge(k.hi(), tau.hi(), dh, muh, T0h, omh);
    const PrF d(dl, dh), mu(mul, muh), T0(T0l, T0h), omT(oml, omh);
    const PrF gamma3 = pfma(mu * -0.75f, g, 0.5f);                              // (2 - 3 mu g) / 4
    const PrF gamma4 = 1.f - gamma3;
    const PrF alpha1 = pfma(gamma1, gamma4, gamma2 * gamma3);
    const PrF alpha2 = pfma(gamma1, gamma3, gamma2 * gamma4);
    const PrF om1sq = om1 * om1, e1omT2 = (e1 + e1) * omT;
    const PrF Q = om1sq + e1omT2;
    const PrF Qp = pfma(T0, om1sq, -e1omT2);
    const PrF k2muP = (k2 * mu) * P;
    const PrF NR = pfma(pfma(-mu, alpha2, gamma3), Q, pfma(alpha2, P, -(k2muP * gamma3)));
    const PrF NT = pfma(pfma(mu, alpha1, gamma4), Qp, T0 * pfma(alpha1, P, k2muP * gamma4));
    const PrF f = (w0 * rD) * prcp(d);
    Rdir = pmax(pmin(f * NR, omT_true), 0.f);
    Tdir = pmax(pmin(-(f * NT), omT_true - Rdir), 0.f);
}
#endif

}  // namespace rb200

// common.cuh — shared device-side types of the B200 column radiation kernels.
#pragma once
#include <cuda_runtime.h>
#include <cfloat>
#include <cstdint>

namespace rb200 {

constexpr int MAX_GAS = 32;
constexpr int MAX_MINOR_DEV = 16;
constexpr int MAX_UNIQ = 63;       // distinct minor-gas scalings per side the prologue kernel keeps per thread

// ---- 4-wide vectors: one 16-byte (fp32) or 32-byte (fp64) aligned access -------------------
template<typename Real> struct V4;
template<> struct alignas(16) V4<float>  { float x, y, z, w; };
template<> struct alignas(32) V4<double> { double x, y, z, w; };

template<typename Real> __device__ __forceinline__ V4<Real> ldg4(const Real* p)
{
    return *reinterpret_cast<const V4<Real>*>(p);
}
template<> __device__ __forceinline__ V4<float> ldg4<float>(const float* p)
{
    const float4 v = __ldg(reinterpret_cast<const float4*>(p));
    return V4<float>{v.x, v.y, v.z, v.w};
}
template<> __device__ __forceinline__ V4<double> ldg4<double>(const double* p)
{
    const double2 a = __ldg(reinterpret_cast<const double2*>(p));
    const double2 b = __ldg(reinterpret_cast<const double2*>(p) + 1);
    return V4<double>{a.x, a.y, b.x, b.y};
}

// Vector access for the recurrence scratch strips.  -DRB_STREAM_HINT makes them streaming (evict-first):
// measured neutral to slightly slower on B200 (the LIFO re-reads of the deepest layers do hit L2), so
// plain accesses are the default.
__device__ __forceinline__ void st_stream(V4<float>* p, const V4<float>& v)
{
#ifndef RB_STREAM_HINT
    *p = v;
#else
    __stcs(reinterpret_cast<float4*>(p), make_float4(v.x, v.y, v.z, v.w));
#endif
}
__device__ __forceinline__ void st_stream(V4<double>* p, const V4<double>& v)
{
#ifndef RB_STREAM_HINT
    *p = v;
#else
    __stcs(reinterpret_cast<double2*>(p), make_double2(v.x, v.y));
    __stcs(reinterpret_cast<double2*>(p) + 1, make_double2(v.z, v.w));
#endif
}
__device__ __forceinline__ V4<float> ld_stream(const V4<float>* p)
{
#ifndef RB_STREAM_HINT
    return *p;
#else
    const float4 v = __ldcs(reinterpret_cast<const float4*>(p));
    return V4<float>{v.x, v.y, v.z, v.w};
#endif
}
__device__ __forceinline__ V4<double> ld_stream(const V4<double>* p)
{
#ifndef RB_STREAM_HINT
    return *p;
#else
    const double2 a = __ldcs(reinterpret_cast<const double2*>(p));
    const double2 b = __ldcs(reinterpret_cast<const double2*>(p) + 1);
    return V4<double>{a.x, a.y, b.x, b.y};
#endif
}

// ---- per-precision math: fp32 uses the fast SFU paths, fp64 the accurate library ------------
template<typename Real> struct M;
template<> struct M<float>
{
    static constexpr float eps = FLT_EPSILON;
    static constexpr float tiny = FLT_MIN;
#ifdef RB_F32_PRECISE
    static __device__ __forceinline__ float exp(float x) { return expf(x); }
    static __device__ __forceinline__ float sqrt(float x) { return sqrtf(x); }
    static __device__ __forceinline__ float div(float a, float b) { return a / b; }
#else
    // bare SFU forms: ex2/rcp.approx.ftz without the range fix-ups of __expf / __fdividef (2 instructions
    // instead of 5; arguments here are never denormal or beyond 2^126)
    static __device__ __forceinline__ float exp(float x) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x * 1.4426950408889634f)); return r; }
    static __device__ __forceinline__ float sqrt(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
    static __device__ __forceinline__ float div(float a, float b) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(b)); return a * r; }
#endif
    static __device__ __forceinline__ float rcp(float a) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a)); return r; }
    static __device__ __forceinline__ float fabs(float a) { return fabsf(a); }
    static __device__ __forceinline__ float max(float a, float b) { return fmaxf(a, b); }
    static __device__ __forceinline__ float min(float a, float b) { return fminf(a, b); }
    static __device__ __forceinline__ int as_int(float a) { return __float_as_int(a); }
    static __device__ __forceinline__ float from_int(int a) { return __int_as_float(a); }
};
template<> struct M<double>
{
    static constexpr double eps = DBL_EPSILON;
    static constexpr double tiny = DBL_MIN;
    static __device__ __forceinline__ double exp(double x) { return ::exp(x); }
    static __device__ __forceinline__ double sqrt(double x) { return ::sqrt(x); }
    static __device__ __forceinline__ double div(double a, double b) { return a / b; }
    static __device__ __forceinline__ double rcp(double a) { return 1.0 / a; }
    static __device__ __forceinline__ double fabs(double a) { return ::fabs(a); }
    static __device__ __forceinline__ double max(double a, double b) { return ::fmax(a, b); }
    static __device__ __forceinline__ double min(double a, double b) { return ::fmin(a, b); }
    static __device__ __forceinline__ int as_int(double a) { return (int)__double_as_longlong(a); }
    static __device__ __forceinline__ double from_int(int a) { return __longlong_as_double((long long)a); }
};

// ---- Pr<Real>: two adjacent g-points of a thread's quad --------------------------------------------------------------
// fp32: one 64-bit register pair driven by sm_100's packed fp32 instructions (FFMA2 / FMUL2 / FADD2: two lanes per issue
// slot, scalar operands broadcast for free as `R.F32`); the band kernels are issue-bound on the FMA pipe, so every
// multiply-add written on a pair costs half an instruction per g-point.  fp64: two scalars, same interface, so the
// kernels are written once.  SFU functions (ex2, rcp, sqrt) and min/max/select stay per lane.
template<typename Real> struct Pr;
template<> struct Pr<float>
{
    float2 v;
    __device__ __forceinline__ Pr() {}
    __device__ __forceinline__ Pr(float a, float b) : v(make_float2(a, b)) {}
    __device__ __forceinline__ explicit Pr(float s) : v(make_float2(s, s)) {}
    __device__ __forceinline__ explicit Pr(float2 f) : v(f) {}
    __device__ __forceinline__ float lo() const { return v.x; }
    __device__ __forceinline__ float hi() const { return v.y; }
};
template<> struct Pr<double>
{
    double x, y;
    __device__ __forceinline__ Pr() {}
    __device__ __forceinline__ Pr(double a, double b) : x(a), y(b) {}
    __device__ __forceinline__ explicit Pr(double s) : x(s), y(s) {}
    __device__ __forceinline__ double lo() const { return x; }
    __device__ __forceinline__ double hi() const { return y; }
};
using PrF = Pr<float>;
using PrD = Pr<double>;
__device__ __forceinline__ PrF operator-(PrF a) { return PrF(-a.v.x, -a.v.y); }                 // folded into operand modifiers
__device__ __forceinline__ PrF operator+(PrF a, PrF b) { return PrF(__fadd2_rn(a.v, b.v)); }
__device__ __forceinline__ PrF operator-(PrF a, PrF b) { return PrF(__fadd2_rn(a.v, (-b).v)); }
__device__ __forceinline__ PrF operator*(PrF a, PrF b) { return PrF(__fmul2_rn(a.v, b.v)); }
__device__ __forceinline__ PrF operator+(PrF a, float s) { return a + PrF(s); }
__device__ __forceinline__ PrF operator+(float s, PrF a) { return a + PrF(s); }
__device__ __forceinline__ PrF operator-(PrF a, float s) { return a + PrF(-s); }
__device__ __forceinline__ PrF operator-(float s, PrF a) { return PrF(s) - a; }
__device__ __forceinline__ PrF operator*(PrF a, float s) { return a * PrF(s); }
__device__ __forceinline__ PrF operator*(float s, PrF a) { return a * PrF(s); }
__device__ __forceinline__ PrF pfma(PrF a, PrF b, PrF c) { return PrF(__ffma2_rn(a.v, b.v, c.v)); }   // a*b + c
__device__ __forceinline__ PrF pfma(float a, PrF b, PrF c) { return pfma(PrF(a), b, c); }
__device__ __forceinline__ PrF pfma(PrF a, float b, PrF c) { return pfma(a, PrF(b), c); }
__device__ __forceinline__ PrF pfma(PrF a, PrF b, float c) { return pfma(a, b, PrF(c)); }
__device__ __forceinline__ PrF pfma(float a, PrF b, float c) { return pfma(PrF(a), b, PrF(c)); }
__device__ __forceinline__ PrF pfma(PrF a, float b, float c) { return pfma(a, PrF(b), PrF(c)); }
__device__ __forceinline__ PrD operator-(PrD a) { return PrD(-a.x, -a.y); }
__device__ __forceinline__ PrD operator+(PrD a, PrD b) { return PrD(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ PrD operator-(PrD a, PrD b) { return PrD(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ PrD operator*(PrD a, PrD b) { return PrD(a.x * b.x, a.y * b.y); }
__device__ __forceinline__ PrD operator+(PrD a, double s) { return PrD(a.x + s, a.y + s); }
__device__ __forceinline__ PrD operator+(double s, PrD a) { return PrD(a.x + s, a.y + s); }
__device__ __forceinline__ PrD operator-(PrD a, double s) { return PrD(a.x - s, a.y - s); }
__device__ __forceinline__ PrD operator-(double s, PrD a) { return PrD(s - a.x, s - a.y); }
__device__ __forceinline__ PrD operator*(PrD a, double s) { return PrD(a.x * s, a.y * s); }
__device__ __forceinline__ PrD operator*(double s, PrD a) { return PrD(a.x * s, a.y * s); }
__device__ __forceinline__ PrD pfma(PrD a, PrD b, PrD c) { return PrD(fma(a.x, b.x, c.x), fma(a.y, b.y, c.y)); }
__device__ __forceinline__ PrD pfma(double a, PrD b, PrD c) { return PrD(fma(a, b.x, c.x), fma(a, b.y, c.y)); }
__device__ __forceinline__ PrD pfma(PrD a, double b, PrD c) { return PrD(fma(a.x, b, c.x), fma(a.y, b, c.y)); }
__device__ __forceinline__ PrD pfma(PrD a, PrD b, double c) { return PrD(fma(a.x, b.x, c), fma(a.y, b.y, c)); }
__device__ __forceinline__ PrD pfma(double a, PrD b, double c) { return PrD(fma(a, b.x, c), fma(a, b.y, c)); }
__device__ __forceinline__ PrD pfma(PrD a, double b, double c) { return PrD(fma(a.x, b, c), fma(a.y, b, c)); }
// per-lane functions
template<typename Real> __device__ __forceinline__ Pr<Real> prcp(Pr<Real> a) { return Pr<Real>(M<Real>::rcp(a.lo()), M<Real>::rcp(a.hi())); }
template<typename Real> __device__ __forceinline__ Pr<Real> psqrt(Pr<Real> a) { return Pr<Real>(M<Real>::sqrt(a.lo()), M<Real>::sqrt(a.hi())); }
template<typename Real> __device__ __forceinline__ Pr<Real> pmax(Pr<Real> a, Real s) { return Pr<Real>(M<Real>::max(a.lo(), s), M<Real>::max(a.hi(), s)); }
template<typename Real> __device__ __forceinline__ Pr<Real> pmin(Pr<Real> a, Pr<Real> b) { return Pr<Real>(M<Real>::min(a.lo(), b.lo()), M<Real>::min(a.hi(), b.hi())); }
template<typename Real> __device__ __forceinline__ Pr<Real> pmax(Pr<Real> a, Pr<Real> b) { return Pr<Real>(M<Real>::max(a.lo(), b.lo()), M<Real>::max(a.hi(), b.hi())); }
// exp(x): fp32 = one packed multiply by log2(e) + two MUFU.EX2; fp64 = the library exp
__device__ __forceinline__ PrF pexp(PrF x)
{
#ifdef RB_F32_PRECISE
    return PrF(expf(x.lo()), expf(x.hi()));
#else
    const PrF t = x * 1.4426950408889634f;
    float a, b;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(a) : "f"(t.lo()));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(b) : "f"(t.hi()));
    return PrF(a, b);
#endif
}
__device__ __forceinline__ PrD pexp(PrD x) { return PrD(::exp(x.x), ::exp(x.y)); }
// halves of a 4-vector / of a thread's 4-element register array
template<typename Real> __device__ __forceinline__ Pr<Real> plo(const V4<Real>& v) { return Pr<Real>(v.x, v.y); }
template<typename Real> __device__ __forceinline__ Pr<Real> phi(const V4<Real>& v) { return Pr<Real>(v.z, v.w); }
template<typename Real> __device__ __forceinline__ Pr<Real> pget(const Real (&a)[4], int h) { return Pr<Real>(a[2 * h], a[2 * h + 1]); }
template<typename Real> __device__ __forceinline__ void pput(Real (&a)[4], int h, Pr<Real> p) { a[2 * h] = p.lo(); a[2 * h + 1] = p.hi(); }
template<typename Real> __device__ __forceinline__ V4<Real> v4_of(Pr<Real> lo, Pr<Real> hi) { return V4<Real>{lo.lo(), lo.hi(), hi.lo(), hi.hi()}; }

// ---- device view of the k-distribution (tables in Real, index grids in double) -------------
template<typename Real> struct DevPlan
{
    int is_lw, ntemp, npres, neta, nbnd, ngpt, nflav, ngas, gw, nplanck, idx_h2o, row_stride;
    const Real* kmajor;        // [b][it][ip][ie][slot][gw], row_stride = (LW ? 2 : 1) * gw
    const Real* pfrac;         // LW: kmajor + gw (slot 1 of the same rows)
    const Real* krayl;         // [side][b][it][ie][gw] (SW) or null
    const Real* kminor[2];     // [m][it][ie][gw]
    const Real* totplnk;       // [b][nplanck] (LW)
    const Real* solar_source;  // [ngpt] (SW)
    const double* temp_ref;    // [ntemp]
    const double* vmr_ref;     // (2, ngas+1, ntemp)
    const double* eta_half;    // (2, nflav, ntemp): vmr_ref(key species 1) / vmr_ref(key species 2) of a flavor
    const int* flavor;         // (2, nflav) 1-based gas index, 0 = dry air
    const int* band_flavor;    // (2, nbnd) 0-based flavor
    const int* band_lims;      // (2, nbnd) 1-based
    const int4* minor_slots[2]; // per side, nuniq entries {idx_minor (-1 = unused), idx_scaling, flags, 0}: the DISTINCT scalings
    const int4* slot_uid[2];    // per side, nscal/4 groups: distinct-scaling number of each slot (nuniq = padding, value 0)
    int nscal, nuniq;          // slots of the band-blocked scaling vector (multiple of 4); distinct scalings
    double press_ref_log0, press_ref_log_delta, press_ref_trop_log;
    double temp_ref_min, temp_ref_delta;
    Real planck_tmin, planck_delta;
};

// vmr inputs of one call: scalar / profile / field per gas (Gas_concs semantics)
struct VmrArgs
{
    const void* p[MAX_GAS];
    int ncol[MAX_GAS];
    int nlay[MAX_GAS];
};

// ---- per-tile interpolation state produced by the prologue kernel, g-point independent ------
// All arrays are column-fastest over the tile: index (c + nc*l) [+ nc*nlay*f].
template<typename Real> struct V2;
template<> struct alignas(8) V2<float> { float x, y; };
template<> struct alignas(16) V2<double> { double x, y; };

template<typename Real> struct TileState
{
    int nc, nlay;            // columns in this tile, layers
    V4<Real>* cell;          // {fpress, col_gas(h2o) + col_dry, 0, bits}
                             //   bits = jtemp | jpress<<8 | tropo<<16
    V4<Real>* scal;          // [nscal/4][l][c]: minor-gas scalings of the contributors of the layer's side
                             //   (col_gas * density * scaling-gas factor, F90:453-475), band-blocked groups of 4
    V4<Real>* flvA;          // [f]: {col_mix0, col_mix1, fminor(1,1), fminor(2,1)}
    V4<Real>* flvB;          // [f]: {fminor(1,2), fminor(2,2), rows(itemp 1), rows(itemp 2)}
                             //   rows = major row | minor row << 20 (row = offset / row stride)
    V2<Real>* pl_lay;        // LW: Planck table position of tlay {frac, lower index}  (nc, nlay)
    V2<Real>* pl_lev;        // LW: idem for tlev                                      (nc, nlay+1)
    unsigned* tmask;         // [c][tmw]: bit l of a column = tropo(c, l)  (fused path only)
    int tmw;                 // mask words per column: tmask_words(nlay)
};
constexpr int TMASK_WORDS = 8;   // at least 8 words (256 layers) per column, more for taller columns
inline int tmask_words(int nlay) { const int w = (nlay + 31) / 32; return w > TMASK_WORDS ? w : TMASK_WORDS; }

struct MinorDev { int idx_minor, idx_scaling, flags, contributor; };   // flags: 1 density, 2 complement

struct BandArgs
{
    int band, gpt0, ngb;     // 0-based band, first g-point (0-based), g-points in band
    int nminor[2];
    int first_minor[2];      // device block index of the band's first contributor (blocks are consecutive)
    int scal_off[2];         // first group (V4 plane) of the band's contributors in TileState::scal
    int flav[2];             // 0-based flavor of the band per side (= DevPlan::band_flavor)
};

}  // namespace rb200

// k1_interp.cu — per-(column, layer) prologue: col_dry, vmr fill and the RRTMGP interpolation.
//
// Follows the CPU path the oracle restates:
//   get_col_dry        src/Gas_optics_rrtmgp.cpp:764-792
//   vmr fill / col_gas src/Gas_optics_rrtmgp.cpp:1103-1160
//   interpolation      rte-rrtmgp/rrtmgp-kernels/mo_gas_optics_rrtmgp_kernels.F90:37-172
// All of it is O(ncol*nlay*nflav) — g-point independent — so it is done ONCE per column here and
// never per g-point.  The index arithmetic (jtemp, jpress, tropo, jeta) runs in fp64 in both
// builds and this translation unit is compiled with --fmad=false, so the integer outputs are
// bit-identical to the fp64 CPU path on the same input values.
#include "common.cuh"
#include "launch.hpp"
#include <algorithm>

namespace rb200 {

template<typename Real>
__global__ void col_dry_kernel(int ncol, int nlay, const Real* __restrict__ vmr_h2o,
                               const Real* __restrict__ plev, Real* __restrict__ col_dry)
{
    const size_t n = (size_t)ncol * nlay;
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double g0 = 9.80665, avogad = 6.02214076e23, m_dry = 0.028964, m_h2o = 0.018016;
    const double q = (double)vmr_h2o[i];
    const double dp = fabs((double)plev[i] - (double)plev[i + ncol]);
    const double m_air = (m_dry + m_h2o * q) / (1.0 + q);
    const double cd = 10.0 * dp * avogad / (1000.0 * m_air * 100.0 * g0);
    col_dry[i] = (Real)(cd / (1.0 + q));
}

template<typename Real> struct RefInterpOut
{
    int* jtemp; int* jpress; signed char* tropo; int* jeta;
    Real* col_mix; Real* fminor; Real* fmajor; Real* col_gas;
};

template<typename Real>
__device__ __forceinline__ double vmr_at(const VmrArgs& v, int g, size_t i_full, int l)
{
    const Real* p = static_cast<const Real*>(v.p[g]);
    if (v.ncol[g] > 1) return (double)p[i_full];
    if (v.nlay[g] > 1) return (double)p[l];
    return (double)p[0];
}

template<typename Real>
__global__ void __launch_bounds__(128)
interp_kernel(const DevPlan<Real> P, const VmrArgs vmr, int ncol, int c0, int nc, int nlay,
              const Real* __restrict__ play, const Real* __restrict__ tlay,
              const Real* __restrict__ col_dry, const Real* __restrict__ tlev,
              TileState<Real> st, RefInterpOut<Real> ro)
{
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)nc * nlay) return;
    const int c = (int)(idx % nc), l = (int)(idx / nc);
    const size_t i = (size_t)(c0 + c) + (size_t)ncol * l;      // index in the caller's arrays
    const size_t ncl = (size_t)ncol * nlay;
    const size_t it = (size_t)c + (size_t)nc * l;              // index in the tile
    const size_t tcl = (size_t)nc * nlay;

    const double t = (double)tlay[i], pr = (double)play[i], cd = (double)col_dry[i];
    // F90:109-125
    int jt = (int)((t - (P.temp_ref_min - P.temp_ref_delta)) / P.temp_ref_delta);
    jt = min(P.ntemp - 1, max(1, jt));
    const double ftemp = (t - P.temp_ref[jt - 1]) / P.temp_ref_delta;
    const double lp = log(pr);
    const double locpress = 1.0 + (lp - P.press_ref_log0) / P.press_ref_log_delta;
    const int jp = min(P.npres - 1, max(1, (int)locpress));
    const double fpress = locpress - (double)jp;
    const int tropo = lp > P.press_ref_trop_log;
    const int side = tropo ? 0 : 1;

    auto cgas = [&](int ig) -> double { return ig == 0 ? cd : vmr_at<Real>(vmr, ig - 1, i, l) * cd; };

    // gridDim.y > 1 (small calls, too few cells to fill the SMs with one thread per cell): the work of a cell is spread over
    // y — y = 0 writes the cell record, the minor-gas scalings, the Planck positions and the tropopause bit, y = 1.. take the
    // flavors round-robin.  Every thread repeats the few operations above; what each stores is computed exactly as before.
    const int ny = (int)gridDim.y, ry = (int)blockIdx.y;
    const bool do_cell = ry == 0;
    const int f_first = ny == 1 ? 0 : ry - 1, f_step = ny == 1 ? 1 : ny - 1;

    if (do_cell)
    {
        if (ro.jtemp) ro.jtemp[i] = jt;
        if (ro.jpress) ro.jpress[i] = jp;
        if (ro.tropo) ro.tropo[i] = (signed char)tropo;
        if (ro.col_gas)
            for (int ig = 0; ig <= P.ngas; ++ig) ro.col_gas[i + ncl * ig] = (Real)cgas(ig);
    }
    if (st.cell && do_cell)
    {
        V4<Real> rec;
        rec.x = (Real)fpress;
        rec.y = (Real)(cgas(P.idx_h2o) + cd);
        rec.z = (Real)0;
        rec.w = M<Real>::from_int(jt | (jp << 8) | (tropo << 16));
        st.cell[it] = rec;
        // minor-gas scaling of every contributor of this layer's side (F90:453-475)
        const double vmr_fact = 1.0 / cd;
        const double dry_fact = 1.0 / (1.0 + cgas(P.idx_h2o) * vmr_fact);
        const double dens = 0.01 * pr / t;
        // each DISTINCT scaling once (many contributors share one: same gas, scaling gas and flags) ...
        const int4* slots = P.minor_slots[side];
        Real uval[MAX_UNIQ + 1];
        for (int u = 0; u < P.nuniq; ++u)
        {
            const int4 dsc = slots[u];
            double scaling = 0.0;
            if (dsc.x >= 0)
            {
                scaling = cgas(dsc.x);
                if (dsc.z & 1)
                {
                    scaling = scaling * dens;
                    if (dsc.y > 0)
                    {
                        const double x = cgas(dsc.y) * vmr_fact * dry_fact;
                        scaling = scaling * ((dsc.z & 2) ? (1.0 - x) : x);
                    }
                }
            }
            uval[u] = (Real)scaling;
        }
        uval[P.nuniq] = (Real)0;
        // ... then scattered to the band-blocked groups of 4 the band kernels read as one vector
        const int4* suid = P.slot_uid[side];
        for (int g4 = 0; g4 < (P.nscal >> 2); ++g4)
        {
            const int4 id = suid[g4];
            st.scal[(size_t)g4 * tcl + it] = V4<Real>{uval[id.x], uval[id.y], uval[id.z], uval[id.w]};
        }
        if (st.pl_lay && tlev)
        {
            // interpolate1D position in totplnk (F90:715-734), band independent
            auto pos = [&](Real T) {
                const Real val0 = (T - P.planck_tmin) / P.planck_delta;
                const int iv = (int)val0;
                const int index = min(P.nplanck - 1, max(1, iv + 1));
                return V2<Real>{val0 - (Real)iv, M<Real>::from_int(index - 1)};
            };
            st.pl_lay[it] = pos(tlay[i]);
            st.pl_lev[it] = pos(tlev[i]);
            if (l == nlay - 1) st.pl_lev[it + nc] = pos(tlev[i + ncol]);
        }
        if (st.tmask && tropo) atomicOr(&st.tmask[(size_t)c * st.tmw + (l >> 5)], 1u << (l & 31));
    }

    // F90:127-169
    if (ny > 1 && ry == 0) return;
    for (int f = f_first; f < P.nflav; f += f_step)
    {
        const int ig1 = P.flavor[2 * f], ig2 = P.flavor[2 * f + 1];
        const double cg1 = cgas(ig1), cg2 = cgas(ig2);
        double cm[2], fm[2][2]; int je[2];
#pragma unroll
        for (int itemp = 0; itemp < 2; ++itemp)
        {
            const int jtt = jt + itemp;   // 1-based
            const double ratio_eta_half = P.eta_half[side + 2 * (f + (size_t)P.nflav * (jtt - 1))];
            cm[itemp] = cg1 + ratio_eta_half * cg2;
            const double eta = (cm[itemp] > 2.0 * DBL_MIN) ? cg1 / cm[itemp] : 0.5;
            const double loceta = eta * (double)(P.neta - 1);
            je[itemp] = min((int)loceta + 1, P.neta - 1);
            const double feta = loceta - trunc(loceta);      // == fmod(loceta, 1.0) exactly for loceta >= 0
            const double ftt = itemp == 0 ? (1.0 - ftemp) : ftemp;
            fm[itemp][0] = (1.0 - feta) * ftt;
            fm[itemp][1] = feta * ftt;
        }
        const size_t ifl = i + ncl * f;
        if (ro.jeta) { ro.jeta[2 * ifl] = je[0]; ro.jeta[2 * ifl + 1] = je[1]; }
        if (ro.col_mix) { ro.col_mix[2 * ifl] = (Real)cm[0]; ro.col_mix[2 * ifl + 1] = (Real)cm[1]; }
        if (ro.fminor)
        {
            ro.fminor[4 * ifl + 0] = (Real)fm[0][0]; ro.fminor[4 * ifl + 1] = (Real)fm[0][1];
            ro.fminor[4 * ifl + 2] = (Real)fm[1][0]; ro.fminor[4 * ifl + 3] = (Real)fm[1][1];
        }
        if (ro.fmajor)
#pragma unroll
            for (int itemp = 0; itemp < 2; ++itemp)
            {
                ro.fmajor[8 * ifl + 4 * itemp + 0] = (Real)((1.0 - fpress) * fm[itemp][0]);
                ro.fmajor[8 * ifl + 4 * itemp + 1] = (Real)((1.0 - fpress) * fm[itemp][1]);
                ro.fmajor[8 * ifl + 4 * itemp + 2] = (Real)(fpress * fm[itemp][0]);
                ro.fmajor[8 * ifl + 4 * itemp + 3] = (Real)(fpress * fm[itemp][1]);
            }
        if (st.flvA)
        {
            const size_t k = it + tcl * f;
            st.flvA[k] = V4<Real>{(Real)cm[0], (Real)cm[1], (Real)fm[0][0], (Real)fm[0][1]};
            const int ip0 = jp + side - 1;     // 0-based lower pressure row (Fortran jpress+itropo-1)
            const int maj0 = ((jt - 1) * (P.npres + 1) + ip0) * P.neta + (je[0] - 1);
            const int maj1 = (jt * (P.npres + 1) + ip0) * P.neta + (je[1] - 1);
            const int min0 = (jt - 1) * P.neta + (je[0] - 1), min1 = jt * P.neta + (je[1] - 1);
            st.flvB[k] = V4<Real>{(Real)fm[1][0], (Real)fm[1][1], M<Real>::from_int(maj0 | (min0 << 20)),
                                  M<Real>::from_int(maj1 | (min1 << 20))};
        }
    }
}

template<typename Real>
void launch_col_dry(int ncol, int nlay, const void* vmr_h2o, const void* plev, void* col_dry, cudaStream_t s)
{
    const size_t n = (size_t)ncol * nlay;
    col_dry_kernel<Real><<<(unsigned)((n + 255) / 256), 256, 0, s>>>(
            ncol, nlay, (const Real*)vmr_h2o, (const Real*)plev, (Real*)col_dry);
    count_launch();
}

template<typename Real>
void launch_interp(const DevPlan<Real>& P, const VmrArgs& vmr, int ncol, int c0, int nc, int nlay,
                   const void* play, const void* tlay, const void* col_dry, const void* tlev,
                   const TileState<Real>& st, int* jtemp, int* jpress, signed char* tropo, int* jeta, void* col_mix,
                   void* fminor, void* fmajor, void* col_gas, cudaStream_t s)
{
    RefInterpOut<Real> ro{jtemp, jpress, tropo, jeta, (Real*)col_mix, (Real*)fminor, (Real*)fmajor, (Real*)col_gas};
    const size_t n = (size_t)nc * nlay;
    // a small call (under 65 536 cells; 100 RFMIP sites x 60 layers = 47 CTAs of one thread per cell): spread each cell's flavors
    // over the grid's y dimension, the kernel is a 20 us latency chain per thread otherwise
    const unsigned ny = n < 65536 ? (unsigned)std::min(P.nflav + 1, 16) : 1u;
    interp_kernel<Real><<<dim3((unsigned)((n + 127) / 128), ny), 128, 0, s>>>(
            P, vmr, ncol, c0, nc, nlay, (const Real*)play, (const Real*)tlay, (const Real*)col_dry,
            (const Real*)tlev, st, ro);
    count_launch();
}

#define INST(Real) \
    template void launch_col_dry<Real>(int, int, const void*, const void*, void*, cudaStream_t); \
    template void launch_interp<Real>(const DevPlan<Real>&, const VmrArgs&, int, int, int, int, const void*, \
        const void*, const void*, const void*, const TileState<Real>&, int*, int*, signed char*, int*, void*, void*, void*, void*, cudaStream_t);
INST(float)
INST(double)
#undef INST

}  // namespace rb200

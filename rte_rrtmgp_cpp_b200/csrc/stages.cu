// stages.cu — materialised (API call-by-call) kernels: every class method of the reference fills
// the same reference-layout arrays (ncol, nlay|nlev, ngpt), column index fastest, so each stage is
// parity-checkable against the oracle and callers that want tau/ssa/g get them.
#include "physics.cuh"
#include "launch.hpp"
#include "stages.hpp"
#include <algorithm>

namespace rb200 {

// ------------------------------------------------------------------------------------------
// Cloud optics: compute_all_from_table + combine (src/Cloud_optics.cpp:72-232)
// ------------------------------------------------------------------------------------------
template<typename Real>
__device__ __forceinline__ void cloud_from_table(Real cwp, Real re, int nsteps, Real step, Real offset,
                                                 const Real* ext, const Real* ssa, const Real* asy,
                                                 Real& tau, Real& taussa, Real& taussag)
{
    if (cwp > (Real)0)
    {
        const Real x = (re - offset) / step;
        // the reference clamps from above only (src/Cloud_optics.cpp:86): a particle size below the table's lower bound
        // (x < 0: rel < radliq_lwr, or dei = 0 in partly specified cells) would read ext[-1...]: clamp from below too.
        // In-range sizes (x >= 0) give index >= 1 anyway, so results inside the table are unchanged.
        const int index = max(1, min((int)x + 1, nsteps - 1));
        const Real fint = x - (Real)(index - 1);
        tau = cwp * (ext[index - 1] + fint * (ext[index] - ext[index - 1]));
        taussa = tau * (ssa[index - 1] + fint * (ssa[index] - ssa[index - 1]));
        taussag = taussa * (asy[index - 1] + fint * (asy[index] - asy[index - 1]));
    }
    else { tau = 0; taussa = 0; taussag = 0; }
}

template<typename Real, bool TWOSTR>
__global__ void cloud_optics_kernel(const CloudDev<Real> L, int ncol, int c0, int nc, int nlay,
                                    const Real* __restrict__ lwp, const Real* __restrict__ iwp,
                                    const Real* __restrict__ rel, const Real* __restrict__ dei,
                                    int nc_out, int c_out0, int delta,
                                    Real* __restrict__ otau, Real* __restrict__ ossa, Real* __restrict__ og)
{
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t n = (size_t)nc * nlay * L.nbnd;
    if (idx >= n) return;
    const int c = (int)(idx % nc);
    const int l = (int)((idx / nc) % nlay);
    const int b = (int)(idx / ((size_t)nc * nlay));
    const size_t i = (size_t)(c0 + c) + (size_t)ncol * l;
    Real lt, lts, ltsg, it, its, itsg;
    cloud_from_table(lwp[i], rel[i], L.liq_nsteps, L.liq_step, L.radliq_lwr,
                     L.extliq + (size_t)b * L.liq_nsteps, L.ssaliq + (size_t)b * L.liq_nsteps,
                     L.asyliq + (size_t)b * L.liq_nsteps, lt, lts, ltsg);
    cloud_from_table(iwp[i], dei[i], L.ice_nsteps, L.ice_step, L.diamice_lwr,
                     L.extice + (size_t)b * L.ice_nsteps, L.ssaice + (size_t)b * L.ice_nsteps,
                     L.asyice + (size_t)b * L.ice_nsteps, it, its, itsg);
    const size_t o = (size_t)(c_out0 + c) + (size_t)nc_out * (l + (size_t)nlay * b);
    if (TWOSTR)
    {
        const Real tau = lt + it, taussa = lts + its, taussag = ltsg + itsg;
        Real t = tau;
        Real w = taussa / M<Real>::max(tau, M<Real>::eps);
        Real g = taussag / M<Real>::max(taussa, M<Real>::eps);
        if (delta) delta_scale_cell(t, w, g);
        otau[o] = t; ossa[o] = w; og[o] = g;
    }
    else
        otau[o] = (lt - lts) + (it - its);
}

template<typename Real>
void launch_cloud_optics(const CloudDev<Real>& L, bool twostr, int ncol, int c0, int nc, int nlay,
                         const void* lwp, const void* iwp, const void* rel, const void* dei,
                         int nc_out, int c_out0, int delta, void* tau, void* ssa, void* g, cudaStream_t s)
{
    const size_t n = (size_t)nc * nlay * L.nbnd;
    const unsigned grid = (unsigned)((n + 255) / 256);
    if (twostr)
        cloud_optics_kernel<Real, true><<<grid, 256, 0, s>>>(L, ncol, c0, nc, nlay, (const Real*)lwp, (const Real*)iwp,
                (const Real*)rel, (const Real*)dei, nc_out, c_out0, delta, (Real*)tau, (Real*)ssa, (Real*)g);
    else
        cloud_optics_kernel<Real, false><<<grid, 256, 0, s>>>(L, ncol, c0, nc, nlay, (const Real*)lwp, (const Real*)iwp,
                (const Real*)rel, (const Real*)dei, nc_out, c_out0, delta, (Real*)tau, nullptr, nullptr);
    count_launch();
}

// ------------------------------------------------------------------------------------------
// Aerosol optics: compute_all_from_table + combine (src/Aerosol_optics.cpp:40-224; CUDA twin
// src_cuda/Aerosol_optics.cu:9-279).  One thread per (column, layer): the 11 mixing ratios, the
// humidity class and the layer mass are fetched once and reused for every band (the reference's GPU
// kernel re-reads them per band), the per-band stores are coalesced.  The look-up tables arrive as
// one species-ordered block [species][humidity class][band] (hydrophobic rows replicated over the
// classes), so every species is the same three loads.  MODE 0: write the aerosol properties;
// MODE 1: increment the by-band properties already in (tau,ssa,g) (the tile's cloud properties) —
// gas (+) cloud (+) aerosol is then one increment of the gas inside the band kernels.
// ------------------------------------------------------------------------------------------
template<typename Real, int MODE>
__global__ void aerosol_optics_kernel(const AerosolDev<Real> A, const AerosolMmr mm, int ncol, int c0, int nc, int nlay,
                                      const Real* __restrict__ rh, const Real* __restrict__ plev,
                                      int nc_out, int c_out0, int delta,
                                      Real* __restrict__ otau, Real* __restrict__ ossa, Real* __restrict__ og)
{
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)nc * nlay) return;
    const int c = (int)(idx % nc), l = (int)(idx / nc);
    const size_t cgl = (size_t)(c0 + c);
    const size_t i = cgl + (size_t)ncol * l;
    // humidity class: first class whose upper bound is not below rh (src/Aerosol_optics.cpp:27-38)
    const Real h = rh[i];
    int ihum = 0;
    while (ihum < A.nhum - 1 && A.rh_upper[ihum] < h) ++ihum;
    const Real dpg = M<Real>::fabs(plev[i] - plev[i + ncol]) / (Real)9.81;
    Real od[AEROSOL_SPECIES];
#pragma unroll
    for (int k = 0; k < AEROSOL_SPECIES; ++k)
    {
        const Real* m = (const Real*)mm.p[k];
        const Real mmr = m[(mm.ncol[k] > 1 ? cgl : (size_t)0) + (mm.nlay[k] > 1 ? (size_t)mm.ncol[k] * l : (size_t)0)];
        od[k] = mmr * dpg;
    }
    const size_t sstride = (size_t)A.nhum * A.nbnd;
    const size_t cl_out = (size_t)nc_out * nlay;
    size_t o = (size_t)(c_out0 + c) + (size_t)nc_out * l;
    const Real* te = A.mext + (size_t)ihum * A.nbnd;
    const Real* ts = A.ssa + (size_t)ihum * A.nbnd;
    const Real* tg = A.asy + (size_t)ihum * A.nbnd;
    for (int b = 0; b < A.nbnd; ++b, o += cl_out)
    {
        Real tau = 0, taussa = 0, taussag = 0;
#pragma unroll
        for (int k = 0; k < AEROSOL_SPECIES; ++k)
        {
            const Real local_od = od[k] * te[k * sstride + b];
            const Real s = local_od * ts[k * sstride + b];
            tau += local_od; taussa += s; taussag += s * tg[k * sstride + b];
        }
        Real t = tau;
        Real w = taussa / M<Real>::max(tau, M<Real>::eps);
        Real g = taussag / M<Real>::max(taussa, M<Real>::eps);
        if (delta) delta_scale_cell(t, w, g);
        if (MODE == 1)
        {
            Real t1 = otau[o], w1 = ossa[o], g1 = og[o];
            increment_2str(t1, w1, g1, t, w, g);
            t = t1; w = w1; g = g1;
        }
        otau[o] = t; ossa[o] = w; og[o] = g;
    }
}

template<typename Real>
void launch_aerosol_optics(const AerosolDev<Real>& A, const AerosolMmr& m, bool increment, int ncol, int c0, int nc, int nlay,
                           const void* rh, const void* plev, int nc_out, int c_out0, int delta,
                           void* tau, void* ssa, void* g, cudaStream_t s)
{
    const size_t n = (size_t)nc * nlay;
    const unsigned grid = (unsigned)((n + 127) / 128);
    if (increment)
        aerosol_optics_kernel<Real, 1><<<grid, 128, 0, s>>>(A, m, ncol, c0, nc, nlay, (const Real*)rh, (const Real*)plev,
                                                            nc_out, c_out0, delta, (Real*)tau, (Real*)ssa, (Real*)g);
    else
        aerosol_optics_kernel<Real, 0><<<grid, 128, 0, s>>>(A, m, ncol, c0, nc, nlay, (const Real*)rh, (const Real*)plev,
                                                            nc_out, c_out0, delta, (Real*)tau, (Real*)ssa, (Real*)g);
    count_launch();
}

// ------------------------------------------------------------------------------------------
// Optical_props arithmetic (mo_optical_props_kernels.F90)
// ------------------------------------------------------------------------------------------
template<typename Real>
__global__ void delta_scale_kernel(size_t n, Real* tau, Real* ssa, Real* g)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Real t = tau[i], w = ssa[i], gg = g[i];
    delta_scale_cell(t, w, gg);
    tau[i] = t; ssa[i] = w; g[i] = gg;
}

// op2 index: same resolution (nb2 == 0) or by band (g-point range [g1,g2) handled by blockIdx.y)
template<typename Real, bool TWOSTR>
__global__ void increment_kernel(size_t ncl, int ngpt, BandLims bl, Real* tau1, Real* ssa1, Real* g1,
                                 const Real* __restrict__ tau2, const Real* __restrict__ ssa2,
                                 const Real* __restrict__ g2)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ncl) return;
    const int b = blockIdx.y;
    int gs, ge; size_t i2;
    if (bl.nbnd > 0) { gs = bl.lims[2 * b] - 1; ge = bl.lims[2 * b + 1]; i2 = i + ncl * b; }
    else { gs = b; ge = b + 1; i2 = i + ncl * b; }
    const Real t2 = tau2[i2];
    Real s2 = 0, gg2 = 0;
    if (TWOSTR) { s2 = ssa2[i2]; gg2 = g2[i2]; }
    for (int g = gs; g < ge; ++g)
    {
        const size_t a = i + ncl * g;
        if (TWOSTR)
        {
            Real t = tau1[a], w = ssa1[a], gg = g1[a];
            increment_2str(t, w, gg, t2, s2, gg2);
            tau1[a] = t; ssa1[a] = w; g1[a] = gg;
        }
        else tau1[a] = tau1[a] + t2;
    }
}

template<typename Real>
void launch_delta_scale(size_t n, void* tau, void* ssa, void* g, cudaStream_t s)
{
    delta_scale_kernel<Real><<<(unsigned)((n + 255) / 256), 256, 0, s>>>(n, (Real*)tau, (Real*)ssa, (Real*)g);
    count_launch();
}

template<typename Real>
void launch_increment(bool twostr, size_t ncl, int ngpt, const BandLims& bl, void* tau1, void* ssa1, void* g1,
                      const void* tau2, const void* ssa2, const void* g2, cudaStream_t s)
{
    dim3 grid((unsigned)((ncl + 255) / 256), bl.nbnd > 0 ? bl.nbnd : ngpt);
    if (twostr)
        increment_kernel<Real, true><<<grid, 256, 0, s>>>(ncl, ngpt, bl, (Real*)tau1, (Real*)ssa1, (Real*)g1,
                                                          (const Real*)tau2, (const Real*)ssa2, (const Real*)g2);
    else
        increment_kernel<Real, false><<<grid, 256, 0, s>>>(ncl, ngpt, bl, (Real*)tau1, nullptr, nullptr,
                                                           (const Real*)tau2, nullptr, nullptr);
    count_launch();
}

// expand_and_transpose (src/Rte_lw.cpp:204-217): (nbnd, ncol) -> (ncol, ngpt)
template<typename Real>
__global__ void expand_kernel(int ncol, BandLims bl, const Real* __restrict__ in, Real* __restrict__ out)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    const int b = blockIdx.y;
    if (c >= ncol) return;
    const Real v = in[b + (size_t)bl.nbnd * c];
    for (int g = bl.lims[2 * b] - 1; g < bl.lims[2 * b + 1]; ++g) out[c + (size_t)ncol * g] = v;
}

template<typename Real>
void launch_expand(int ncol, const BandLims& bl, const void* in, void* out, cudaStream_t s)
{
    dim3 grid((unsigned)((ncol + 127) / 128), bl.nbnd);
    expand_kernel<Real><<<grid, 128, 0, s>>>(ncol, bl, (const Real*)in, (Real*)out);
    count_launch();
}

// ------------------------------------------------------------------------------------------
// rte_lw_solver_noscat (mo_rte_solver_kernels.F90:51-367): thread = (column, g-point), column
// fastest (the reference layout is coalesced along columns).  The up-sweep recomputes trans and
// source_up from the inputs instead of round-tripping 4 scratch arrays through HBM.
// ------------------------------------------------------------------------------------------
__constant__ double c_gauss_Ds[4][4] = {
    {1. / 0.6096748751, 0., 0., 0.},
    {1. / 0.2509907356, 1. / 0.7908473988, 0., 0.},
    {1. / 0.1024922169, 1. / 0.4417960320, 1. / 0.8633751621, 0.},
    {1. / 0.0454586727, 1. / 0.2322334416, 1. / 0.5740198775, 1. / 0.903077597}};
__constant__ double c_gauss_wts[4][4] = {
    {1., 0., 0., 0.},
    {0.2300253764, 0.7699746236, 0., 0.},
    {0.0437820218, 0.3875796738, 0.5686383044, 0.},
    {0.0092068785, 0.1285704278, 0.4323381850, 0.4298845087}};

template<typename Real>
__global__ void __launch_bounds__(128)
lw_solver_kernel(int ncol, int nlay, int ngpt, int top_at_1, int nmus,
                 const Real* __restrict__ tau, const Real* __restrict__ lay_src,
                 const Real* __restrict__ lev_src, const Real* __restrict__ sfc_emis,
                 const Real* __restrict__ sfc_src, const Real* __restrict__ inc_flux,
                 Real* __restrict__ flux_up, Real* __restrict__ flux_dn,
                 const Real* __restrict__ sfc_src_jac, Real* __restrict__ flux_up_jac)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    const int g = blockIdx.y;
    if (c >= ncol) return;
    const size_t ncl = (size_t)ncol * nlay, ncv = (size_t)ncol * (nlay + 1);
    const Real* tg = tau + ncl * g + c;
    const Real* lay = lay_src + ncl * g + c;
    const Real* lev = lev_src + ncv * g + c;
    Real* up = flux_up + ncv * g + c;
    Real* dn = flux_dn + ncv * g + c;
    const size_t cg = c + (size_t)ncol * g;
    const Real emis = sfc_emis[cg], ssrc = sfc_src[cg];
    const Real inc = inc_flux ? inc_flux[cg] : (Real)0;
    const Real pi = (Real)3.14159265358979323846;
    auto level = [&](int j) { return (size_t)ncol * (top_at_1 ? j : nlay - j); };
    auto layer = [&](int k) { return (size_t)ncol * (top_at_1 ? k : nlay - 1 - k); };

    for (int imu = 0; imu < nmus; ++imu)
    {
        const Real D = (Real)c_gauss_Ds[nmus - 1][imu], w = (Real)c_gauss_wts[nmus - 1][imu];
        const Real piw = pi * w;
        Real radn = inc / piw;
        { const Real v = piw * radn; Real* o = dn + level(0); *o = imu == 0 ? v : *o + v; }
        for (int k = 0; k < nlay; ++k)
        {
            const size_t il = layer(k);
            Real t, s_up, s_dn;
            lw_layer<Real>(tg[il] * D, lay[il], lev[level(k)], lev[level(k + 1)], t, s_up, s_dn);
            radn = t * radn + s_dn;
            const Real v = piw * radn; Real* o = dn + level(k + 1); *o = imu == 0 ? v : *o + v;
        }
        radn = radn * ((Real)1 - emis) + emis * ssrc;
        Real jac = sfc_src_jac ? emis * sfc_src_jac[cg] : (Real)0;
        { const Real v = piw * radn; Real* o = up + level(nlay); *o = imu == 0 ? v : *o + v; }
        if (flux_up_jac) { const Real v = piw * jac; Real* o = flux_up_jac + ncv * g + c + level(nlay); *o = imu == 0 ? v : *o + v; }
        for (int k = nlay - 1; k >= 0; --k)
        {
            const size_t il = layer(k);
            Real t, s_up, s_dn;
            lw_layer<Real>(tg[il] * D, lay[il], lev[level(k)], lev[level(k + 1)], t, s_up, s_dn);
            radn = t * radn + s_up;
            const Real v = piw * radn; Real* o = up + level(k); *o = imu == 0 ? v : *o + v;
            if (flux_up_jac) { jac = t * jac; const Real v = piw * jac; Real* o = flux_up_jac + ncv * g + c + level(k); *o = imu == 0 ? v : *o + v; }
        }
    }
}

template<typename Real>
void launch_lw_solver(int ncol, int nlay, int ngpt, int top_at_1, int nmus, const void* tau, const void* lay_src,
                      const void* lev_src, const void* sfc_emis, const void* sfc_src, const void* inc_flux,
                      void* flux_up, void* flux_dn, const void* sfc_src_jac, void* flux_up_jac, cudaStream_t s)
{
    dim3 grid((unsigned)((ncol + 127) / 128), ngpt);
    lw_solver_kernel<Real><<<grid, 128, 0, s>>>(ncol, nlay, ngpt, top_at_1, nmus, (const Real*)tau, (const Real*)lay_src,
            (const Real*)lev_src, (const Real*)sfc_emis, (const Real*)sfc_src, (const Real*)inc_flux,
            (Real*)flux_up, (Real*)flux_dn, (const Real*)sfc_src_jac, (Real*)flux_up_jac);
    count_launch();
}

// ------------------------------------------------------------------------------------------
// rte_lw_solver_noscat with do_rescaling (mo_rte_solver_kernels.F90:148-178, 206-214, 753-844; Tang et al. 2018):
// the optical depth is scaled by 1 - w(1 - b), b = (1 - g)/2, and the up sweep and a SECOND down sweep carry an
// adjustment term Cn * (An * I_opposite - ...).  The first down sweep and the up sweep must both be complete
// before the second down sweep, so one angle's radiances live in work2 = 2 x (ncol, nlay+1, ngpt) (thread-private
// lines: thread = (column, g-point)) and are added onto the outputs at the end of the angle.  The level whose
// opposite-direction radiance enters each adjustment is the Fortran's, which differs between the orientations in
// the second down sweep (top of the layer when top_at_1, bottom of the layer otherwise).
// ------------------------------------------------------------------------------------------
template<typename Real>
__global__ void __launch_bounds__(128)
lw_solver_rescl_kernel(int ncol, int nlay, int ngpt, int top_at_1, int nmus,
                       const Real* __restrict__ tau, const Real* __restrict__ ssa, const Real* __restrict__ asy,
                       const Real* __restrict__ lay_src, const Real* __restrict__ lev_src,
                       const Real* __restrict__ sfc_emis, const Real* __restrict__ sfc_src,
                       const Real* __restrict__ inc_flux, Real* __restrict__ flux_up, Real* __restrict__ flux_dn,
                       const Real* __restrict__ sfc_src_jac, Real* __restrict__ flux_up_jac, Real* __restrict__ work2)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    const int g = blockIdx.y;
    if (c >= ncol) return;
    const size_t ncl = (size_t)ncol * nlay, ncv = (size_t)ncol * (nlay + 1);
    const Real* tg = tau + ncl * g + c; const Real* wg = ssa + ncl * g + c; const Real* ag = asy + ncl * g + c;
    const Real* lay = lay_src + ncl * g + c;
    const Real* lev = lev_src + ncv * g + c;
    Real* up = flux_up + ncv * g + c; Real* dn = flux_dn + ncv * g + c;
    Real* ru = work2 + ncv * g + c;   Real* rd = ru + ncv * ngpt;
    const size_t cg = c + (size_t)ncol * g;
    const Real emis = sfc_emis[cg], ssrc = sfc_src[cg];
    const Real inc = inc_flux ? inc_flux[cg] : (Real)0;
    const Real pi = (Real)3.14159265358979323846;
    auto level = [&](int j) { return (size_t)ncol * (top_at_1 ? j : nlay - j); };
    auto layer = [&](int k) { return (size_t)ncol * (top_at_1 ? k : nlay - 1 - k); };
    // scaled path, transmissivity, sources, An, Cn of top-down layer k at secant D
    auto cell = [&](int k, Real D, Real& t, Real& s_up, Real& s_dn, Real& An, Real& Cn)
    {
        const size_t il = layer(k);
        const Real ssal = wg[il];
        const Real wb = ssal * ((Real)1 - ag[il]) * (Real)0.5;
        const Real scaleTau = (Real)1 - ssal + wb;
        Cn = M<Real>::div((Real)0.4 * wb, scaleTau);
        lw_layer<Real>(tg[il] * D * scaleTau, lay[il], lev[level(k)], lev[level(k + 1)], t, s_up, s_dn);
        An = (Real)1 - t * t;
    };

    for (int imu = 0; imu < nmus; ++imu)
    {
        const Real D = (Real)c_gauss_Ds[nmus - 1][imu], w = (Real)c_gauss_wts[nmus - 1][imu];
        const Real piw = pi * w;
        Real t, s_up, s_dn, An, Cn;
        Real radn = inc / piw;
        rd[level(0)] = radn;
        for (int k = 0; k < nlay; ++k)
        {
            cell(k, D, t, s_up, s_dn, An, Cn);
            radn = t * radn + s_dn;
            rd[level(k + 1)] = radn;
        }
        radn = radn * ((Real)1 - emis) + emis * ssrc;
        ru[level(nlay)] = radn;
        Real jac = sfc_src_jac ? emis * sfc_src_jac[cg] : (Real)0;
        if (flux_up_jac) { const Real v = piw * jac; Real* o = flux_up_jac + ncv * g + c + level(nlay); *o = imu == 0 ? v : *o + v; }
        for (int k = nlay - 1; k >= 0; --k)
        {
            cell(k, D, t, s_up, s_dn, An, Cn);
            const Real adj = Cn * (An * rd[level(k)] - t * s_dn - s_up);          // down-welling at the layer top
            radn = t * radn + s_up + adj;
            ru[level(k)] = radn;
            if (flux_up_jac) { jac = t * jac; const Real v = piw * jac; Real* o = flux_up_jac + ncv * g + c + level(k); *o = imu == 0 ? v : *o + v; }
        }
        radn = rd[level(0)];
        for (int k = 0; k < nlay; ++k)
        {
            cell(k, D, t, s_up, s_dn, An, Cn);
            const Real adj = Cn * (An * ru[level(top_at_1 ? k : k + 1)] - t * s_up - s_dn);
            radn = t * radn + s_dn + adj;
            rd[level(k + 1)] = radn;
        }
        for (int j = 0; j <= nlay; ++j)
        {
            const size_t lv = level(j);
            const Real vu = piw * ru[lv], vd = piw * rd[lv];
            up[lv] = imu == 0 ? vu : up[lv] + vu;
            dn[lv] = imu == 0 ? vd : dn[lv] + vd;
        }
    }
}

template<typename Real>
void launch_lw_solver_rescl(int ncol, int nlay, int ngpt, int top_at_1, int nmus, const void* tau, const void* ssa,
                            const void* g, const void* lay_src, const void* lev_src, const void* sfc_emis,
                            const void* sfc_src, const void* inc_flux, void* flux_up, void* flux_dn,
                            const void* sfc_src_jac, void* flux_up_jac, void* work2, cudaStream_t s)
{
    dim3 grid((unsigned)((ncol + 127) / 128), ngpt);
    lw_solver_rescl_kernel<Real><<<grid, 128, 0, s>>>(ncol, nlay, ngpt, top_at_1, nmus, (const Real*)tau, (const Real*)ssa,
            (const Real*)g, (const Real*)lay_src, (const Real*)lev_src, (const Real*)sfc_emis, (const Real*)sfc_src,
            (const Real*)inc_flux, (Real*)flux_up, (Real*)flux_dn, (const Real*)sfc_src_jac, (Real*)flux_up_jac,
            (Real*)work2);
    count_launch();
}

// ------------------------------------------------------------------------------------------
// rte_sw_solver_2stream (mo_rte_solver_kernels.F90:503-609, 985-1245); work4 = 4 arrays
// (ncol, nlay, ngpt) holding the per-layer recurrence coefficients between the passes.
// ------------------------------------------------------------------------------------------
template<typename Real>
__global__ void __launch_bounds__(128)
sw_solver_kernel(int ncol, int nlay, int ngpt, int top_at_1, int mu0_by_layer,
                 const Real* __restrict__ tau, const Real* __restrict__ ssa, const Real* __restrict__ gg,
                 const Real* __restrict__ mu0_, const Real* __restrict__ alb_dir, const Real* __restrict__ alb_dif,
                 const Real* __restrict__ inc_dir, const Real* __restrict__ inc_dif,
                 Real* __restrict__ flux_up, Real* __restrict__ flux_dn, Real* __restrict__ flux_dir,
                 Real* __restrict__ work)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    const int g = blockIdx.y;
    if (c >= ncol) return;
    const size_t ncl = (size_t)ncol * nlay, ncv = (size_t)ncol * (nlay + 1);
    const size_t ngl = ncl * ngpt;
    const Real* tg = tau + ncl * g + c; const Real* wg = ssa + ncl * g + c; const Real* ag = gg + ncl * g + c;
    Real* w0 = work + ncl * g + c; Real* w1 = w0 + ngl; Real* w2 = w1 + ngl; Real* w3 = w2 + ngl;
    Real* up = flux_up + ncv * g + c; Real* dn = flux_dn + ncv * g + c; Real* dr = flux_dir + ncv * g + c;
    const size_t cg = c + (size_t)ncol * g;
    auto level = [&](int j) { return (size_t)ncol * (top_at_1 ? j : nlay - j); };
    auto layer = [&](int k) { return (size_t)ncol * (top_at_1 ? k : nlay - 1 - k); };
    // mu0 (ncol), or (ncol, nlay) on a round earth (F90:517, 1061-1065, 1117-1125): each layer's own value, the
    // top layer's for the boundary condition, the surface layer's for the surface source
    Real mu0 = mu0_[c + (mu0_by_layer ? layer(0) : 0)];
    Real mu0s = M<Real>::max(M<Real>::sqrt(M<Real>::eps), mu0);
    bool sun = mu0 > (Real)0;

    Real dir = inc_dir[cg] * mu0;
    dr[level(0)] = dir;
    for (int k = 0; k < nlay; ++k)
    {
        const size_t il = layer(k);
        if (mu0_by_layer)
        {
            mu0 = mu0_[c + il];
            mu0s = M<Real>::max(M<Real>::sqrt(M<Real>::eps), mu0);
            sun = mu0 > (Real)0;
        }
        Real Rdif, Tdif, Rdir, Tdir, Tnoscat;
        two_stream<Real>(tg[il], wg[il], ag[il], mu0s, Rdif, Tdif, Rdir, Tdir, Tnoscat);
        w0[il] = Rdif; w1[il] = Tdif;
        w2[il] = sun ? Rdir * dir : (Real)0;
        w3[il] = sun ? Tdir * dir : (Real)0;
        dir = Tnoscat * dir;
        dr[level(k + 1)] = dir;
    }
    Real alb = alb_dif[cg], src = sun ? dir * alb_dir[cg] : (Real)0;
    for (int k = nlay - 1; k >= 0; --k)
    {
        const size_t il = layer(k);
        const Real rdif = w0[il], tdif = w1[il], sup = w2[il], sdn = w3[il];
        const Real denom = (Real)1 / ((Real)1 - rdif * alb);
        w0[il] = tdif; w1[il] = denom; w2[il] = alb; w3[il] = src;
        // keep rdif, src_dn for pass 3 in the (not yet final) flux arrays
        up[level(k + 1)] = rdif; dn[level(k + 1)] = sdn;
        const Real nsrc = sup + tdif * denom * (src + alb * sdn);
        alb = rdif + tdif * tdif * alb * denom;
        src = nsrc;
    }
    Real fdn = inc_dif ? inc_dif[cg] : (Real)0;
    Real fup = fdn * alb + src;
    up[level(0)] = fup; dn[level(0)] = fdn + dr[level(0)];
    for (int k = 0; k < nlay; ++k)
    {
        const size_t il = layer(k), lv = level(k + 1);
        const Real tdif = w0[il], denom = w1[il], albl = w2[il], srcl = w3[il];
        const Real rdif = up[lv], sdn = dn[lv];
        fdn = (tdif * fdn + rdif * srcl + sdn) * denom;      // F90:1196-1200, same association
        fup = fdn * albl + srcl;
        up[lv] = fup; dn[lv] = fdn + dr[lv];
    }
}

template<typename Real>
void launch_sw_solver(int ncol, int nlay, int ngpt, int top_at_1, int mu0_by_layer, const void* tau, const void* ssa,
                      const void* g, const void* mu0, const void* alb_dir, const void* alb_dif, const void* inc_dir,
                      const void* inc_dif, void* flux_up, void* flux_dn, void* flux_dir, void* work4, cudaStream_t s)
{
    dim3 grid((unsigned)((ncol + 127) / 128), ngpt);
    sw_solver_kernel<Real><<<grid, 128, 0, s>>>(ncol, nlay, ngpt, top_at_1, mu0_by_layer, (const Real*)tau, (const Real*)ssa,
            (const Real*)g, (const Real*)mu0, (const Real*)alb_dir, (const Real*)alb_dif, (const Real*)inc_dir,
            (const Real*)inc_dif, (Real*)flux_up, (Real*)flux_dn, (Real*)flux_dir, (Real*)work4);
    count_launch();
}

// ------------------------------------------------------------------------------------------
// Fluxes: sum_broadband (sequential over g-points, the oracle's order), sum_byband, net
// ------------------------------------------------------------------------------------------
template<typename Real>
__global__ void sum_broadband_kernel(size_t ncv, int ngpt, const Real* __restrict__ gpt_flux, Real* __restrict__ flux)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ncv) return;
    Real s = 0;
    for (int g = 0; g < ngpt; ++g) s = s + gpt_flux[i + ncv * g];
    flux[i] = s;
}

template<typename Real>
__global__ void sum_byband_kernel(size_t ncv, BandLims bl, const Real* __restrict__ gpt_flux, Real* __restrict__ bnd_flux)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int b = blockIdx.y;
    if (i >= ncv) return;
    const int g1 = bl.lims[2 * b] - 1, g2 = bl.lims[2 * b + 1];
    Real s = gpt_flux[i + ncv * g1];
    for (int g = g1 + 1; g < g2; ++g) s = s + gpt_flux[i + ncv * g];
    bnd_flux[i + ncv * b] = s;
}

template<typename Real>
__global__ void net_kernel(size_t n, const Real* __restrict__ dn, const Real* __restrict__ up, Real* __restrict__ net)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) net[i] = dn[i] - up[i];
}

// mo_heating_rates.F90:44-73
template<typename Real>
__global__ void heating_rate_kernel(int ncol, int nlay, const Real* __restrict__ up, const Real* __restrict__ dn,
                                    const Real* __restrict__ plev, Real* __restrict__ hr)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)ncol * nlay) return;
    const size_t b = i + ncol;
    const Real grav = (Real)9.80665, cp_dry = (Real)1004.64;
    hr[i] = (up[b] - up[i] - dn[b] + dn[i]) * grav / (cp_dry * (plev[b] - plev[i]));
}

template<typename Real>
__global__ void range_flag_kernel(size_t n, const Real* __restrict__ x, Real lo, Real hi, int* flag)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    int bad = 0;
    for (; i < n; i += (size_t)gridDim.x * blockDim.x) bad |= !(x[i] >= lo && x[i] <= hi);
    if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicOr(flag, 1);
}

template<typename Real>
void launch_sum_broadband(int ncol, int nlev, int ngpt, const void* gpt_flux, void* flux, cudaStream_t s)
{
    const size_t ncv = (size_t)ncol * nlev;
    sum_broadband_kernel<Real><<<(unsigned)((ncv + 127) / 128), 128, 0, s>>>(ncv, ngpt, (const Real*)gpt_flux, (Real*)flux);
    count_launch();
}
template<typename Real>
void launch_sum_byband(int ncol, int nlev, const BandLims& bl, const void* gpt_flux, void* bnd_flux, cudaStream_t s)
{
    const size_t ncv = (size_t)ncol * nlev;
    dim3 grid((unsigned)((ncv + 127) / 128), bl.nbnd);
    sum_byband_kernel<Real><<<grid, 128, 0, s>>>(ncv, bl, (const Real*)gpt_flux, (Real*)bnd_flux);
    count_launch();
}
template<typename Real> void launch_net(size_t n, const void* dn, const void* up, void* net, cudaStream_t s)
{
    net_kernel<Real><<<(unsigned)((n + 255) / 256), 256, 0, s>>>(n, (const Real*)dn, (const Real*)up, (Real*)net);
    count_launch();
}
// toa_src(icol,igpt) *= tsi_scaling(icol)  (src_test/Radiation_solver.cu:42-70)
template<typename Real> __global__ void scale_columns_kernel(size_t n, int ncol, Real* x, const Real* __restrict__ scale)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) x[i] *= scale[i % ncol];
}
template<typename Real> void launch_scale_columns(int ncol, int n2, void* x, const void* scale, cudaStream_t s)
{
    const size_t n = (size_t)ncol * n2;
    scale_columns_kernel<Real><<<(unsigned)((n + 255) / 256), 256, 0, s>>>(n, ncol, (Real*)x, (const Real*)scale);
    count_launch();
}
template<typename Real>
void launch_heating_rate(int ncol, int nlay, const void* up, const void* dn, const void* plev, void* hr, cudaStream_t s)
{
    const size_t n = (size_t)ncol * nlay;
    heating_rate_kernel<Real><<<(unsigned)((n + 255) / 256), 256, 0, s>>>(ncol, nlay, (const Real*)up, (const Real*)dn,
                                                                          (const Real*)plev, (Real*)hr);
    count_launch();
}
template<typename Real>
void launch_range_flag(size_t n, const void* x, double lo, double hi, int* d_flag, cudaStream_t s)
{
    const unsigned grid = (unsigned)std::min<size_t>((n + 255) / 256, 148 * 8);
    range_flag_kernel<Real><<<grid ? grid : 1, 256, 0, s>>>(n, (const Real*)x, (Real)lo, (Real)hi, d_flag);
    count_launch();
}

#define INST(Real) \
    template void launch_cloud_optics<Real>(const CloudDev<Real>&, bool, int, int, int, int, const void*, const void*, const void*, const void*, int, int, int, void*, void*, void*, cudaStream_t); \
    template void launch_aerosol_optics<Real>(const AerosolDev<Real>&, const AerosolMmr&, bool, int, int, int, int, const void*, const void*, int, int, int, void*, void*, void*, cudaStream_t); \
    template void launch_delta_scale<Real>(size_t, void*, void*, void*, cudaStream_t); \
    template void launch_increment<Real>(bool, size_t, int, const BandLims&, void*, void*, void*, const void*, const void*, const void*, cudaStream_t); \
    template void launch_expand<Real>(int, const BandLims&, const void*, void*, cudaStream_t); \
    template void launch_lw_solver<Real>(int, int, int, int, int, const void*, const void*, const void*, const void*, const void*, const void*, void*, void*, const void*, void*, cudaStream_t); \
    template void launch_lw_solver_rescl<Real>(int, int, int, int, int, const void*, const void*, const void*, const void*, const void*, const void*, const void*, const void*, void*, void*, const void*, void*, void*, cudaStream_t); \
    template void launch_sw_solver<Real>(int, int, int, int, int, const void*, const void*, const void*, const void*, const void*, const void*, const void*, const void*, void*, void*, void*, void*, cudaStream_t); \
    template void launch_sum_broadband<Real>(int, int, int, const void*, void*, cudaStream_t); \
    template void launch_sum_byband<Real>(int, int, const BandLims&, const void*, void*, cudaStream_t); \
    template void launch_net<Real>(size_t, const void*, const void*, void*, cudaStream_t); \
    template void launch_scale_columns<Real>(int, int, void*, const void*, cudaStream_t); \
    template void launch_heating_rate<Real>(int, int, const void*, const void*, const void*, void*, cudaStream_t); \
    template void launch_range_flag<Real>(size_t, const void*, double, double, int*, cudaStream_t);
INST(float)
INST(double)
#undef INST

}  // namespace rb200

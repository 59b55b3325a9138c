// cabi.cu — the extern "C" boundary (include/rrtmgp_b200.h): handles, table upload, argument
// checks, tile loop of the fused path.  No compute happens on the host and there is no CPU
// fallback: every entry that needs the device fails with RRTMGP_B200_E_CUDA when CUDA is absent.
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <mutex>
#include <cstring>
#include <string>
#include <vector>
#include "../../include/rrtmgp_b200.h"
#include "plan.hpp"
#include "launch.hpp"
#include "stages.hpp"

namespace rb200 {

static thread_local std::string g_err;
static std::atomic<long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

// optional device-side timing of the fused band kernels (bench.py's roofline leg)
struct Profile { bool on = false; std::vector<cudaEvent_t> ev[3]; double ms[3] = {0, 0, 0}; long long launches[3] = {0, 0, 0}; };   // 0 LW, 1 SW, 2 paired LW+SW
static Profile g_prof;
// host-buffer entry: run the LW and the SW band kernels of a chunk on two streams (RB200_HOST_CONCURRENT=0 disables)
// RB200_PATH=colres selects the column-resident kernels (recurrences in shared memory) instead of the
// per-band kernels with recurrence scratch strips in HBM, which measure faster (kernels)
// unset: chosen per call by size (small calls are column-resident); "strips" / "colres" force one
// RB200_ANGLES=launches: longwave calls with several Gauss angles run one pass of the single-angle kernels per angle (the
// round-1 path, kept for A/B timing) instead of the kernels with the angles inside the layer loop
static const bool g_angle_launches = [] { const char* e = getenv("RB200_ANGLES"); return e && e[0] == 'l'; }();
static const int g_colres = [] { const char* e = getenv("RB200_PATH"); return !e || !*e ? -1 : (e[0] == 'c' ? 1 : 0); }();
// Calls of at most this many columns take the column-resident kernels when the path is not forced.  Measured on B200 (fp32):
// 100 columns x 60 layers 1.6 ms (strips, paired) vs 0.42 ms; 16384 x 72 8.0 ms vs 10.0 ms; the strips launches cost a
// fixed ~0.23 ms per band pair until they fill a wave (23680 columns), the column-resident kernel ~0.56 us per column.
constexpr int kColresAutoMaxCols = 6144;
static const bool g_host_concurrent = [] { const char* e = getenv("RB200_HOST_CONCURRENT"); return !(e && e[0] == '0'); }();

struct CudaError : std::runtime_error { using std::runtime_error::runtime_error; };
struct RangeError : std::runtime_error { using std::runtime_error::runtime_error; };

#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
    throw CudaError(std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)

static int fail(int code, const std::string& msg) { g_err = msg; return code; }

template<typename F> static int guarded(F&& f, bool check_cuda = true)
{
    try { f(); cudaError_t e = check_cuda ? cudaPeekAtLastError() : cudaSuccess; if (e != cudaSuccess) { cudaGetLastError(); return fail(RRTMGP_B200_E_CUDA, std::string("kernel launch: ") + cudaGetErrorString(e)); } return 0; }
    catch (const ArgError& e) { return fail(RRTMGP_B200_E_ARG, e.what()); }
    catch (const MissingError& e) { return fail(RRTMGP_B200_E_MISSING, e.what()); }
    catch (const RangeError& e) { return fail(RRTMGP_B200_E_RANGE, e.what()); }
    catch (const CudaError& e) { return fail(RRTMGP_B200_E_CUDA, e.what()); }
    catch (const std::exception& e) { return fail(RRTMGP_B200_E_ARG, e.what()); }
}

// device buffer that converts host doubles to Real
struct DevBuf
{
    void* p = nullptr; size_t bytes = 0;
    ~DevBuf() { if (p) cudaFree(p); }
    DevBuf() = default; DevBuf(const DevBuf&) = delete; DevBuf& operator=(const DevBuf&) = delete;
    void alloc(size_t n) { if (p) { cudaFree(p); p = nullptr; } bytes = n; if (n) CK(cudaMalloc(&p, n)); }
    template<typename T> void upload(const std::vector<T>& v)
    {
        alloc(v.size() * sizeof(T));
        if (!v.empty()) CK(cudaMemcpy(p, v.data(), bytes, cudaMemcpyHostToDevice));
    }
    template<typename Real> void upload_as(const std::vector<double>& v)
    {
        std::vector<Real> t(v.begin(), v.end());
        upload(t);
    }
};

// grow-only workspace carved into aligned pieces per call
struct Arena
{
    char* base = nullptr; size_t cap = 0, off = 0;
    // debug mode (rrtmgp_b200_set_debug; the pool's compute-sanitizer is closed): reserve() fills the whole arena with 0xFF
    // bytes (NaN as float / double, -1 as int) so that anything read before it is written shows up in the fluxes, and every
    // piece is followed by a 256-byte canary block that check() expects untouched after the call.
    bool debug = false, selftest = false;      // selftest: arm() corrupts the first canary on purpose (the check must then fail)
    static constexpr size_t kCanary = 256;
    static constexpr unsigned char kCanaryByte = 0xA5;
    std::vector<size_t> canaries;       // offsets of the canary blocks carved since the last reserve()
    ~Arena() { if (base) cudaFree(base); }
    void reserve(size_t n, cudaStream_t s)
    {
        if (debug) n += 64 * (kCanary + 256);
        if (n > cap)
        {
            if (base) { CK(cudaStreamSynchronize(s)); cudaFree(base); base = nullptr; }
            CK(cudaMalloc((void**)&base, n)); cap = n;
        }
        off = 0;
        canaries.clear();
        if (debug && cap) CK(cudaMemsetAsync(base, 0xFF, cap, s));
    }
    template<typename T> T* take(size_t count)
    {
        off = (off + 255) & ~(size_t)255;
        T* r = reinterpret_cast<T*>(base + off);
        off += count * sizeof(T);
        if (debug)
        {
            off = (off + 255) & ~(size_t)255;
            canaries.push_back(off);
            off += kCanary;
        }
        if (off > cap) throw ArgError("internal: workspace overflow");
        return r;
    }
    // after the last take(): write the canary blocks (stream-ordered, before the kernels of the call)
    void arm(cudaStream_t s)
    {
        if (!debug) return;
        for (size_t o : canaries) CK(cudaMemsetAsync(base + o, kCanaryByte, kCanary, s));
        if (selftest && !canaries.empty()) CK(cudaMemsetAsync(base + canaries[0] + 7, 0, 1, s));
    }
    // after the call: synchronises; throws when a kernel wrote past the end of a piece
    void check(cudaStream_t s, const char* what)
    {
        if (!debug) return;
        CK(cudaStreamSynchronize(s));
        std::vector<unsigned char> h(kCanary);
        for (size_t i = 0; i < canaries.size(); ++i)
        {
            CK(cudaMemcpy(h.data(), base + canaries[i], kCanary, cudaMemcpyDeviceToHost));
            for (unsigned char b : h)
                if (b != kCanaryByte)
                    throw CudaError(std::string(what) + ": workspace canary " + std::to_string(i) + " of " + std::to_string(canaries.size()) +
                                    " overwritten (a kernel wrote past the end of a workspace piece)");
        }
    }
};
static size_t pad256(size_t n) { return (n + 255) & ~(size_t)255; }

// stream-ordered temporary of a stage call, released on every exit path (an exception included)
struct AsyncTemp
{
    char* p = nullptr; cudaStream_t s = nullptr;
    explicit AsyncTemp(cudaStream_t s_) : s(s_) {}
    AsyncTemp(const AsyncTemp&) = delete; AsyncTemp& operator=(const AsyncTemp&) = delete;
    void alloc(size_t n) { if (n) CK(cudaMallocAsync((void**)&p, n, s)); }
    ~AsyncTemp() { if (p) cudaFreeAsync(p, s); }
};

}  // namespace rb200

using namespace rb200;

struct rrtmgp_b200_gas_optics
{
    int dtype = 0, device = 0;
    bool host_only = false;
    HostPlan h;
    DevBuf kmajor, pfrac, krayl, kminor[2], totplnk, solar, temp_ref, vmr_ref, eta_half, flavor, band_flavor, band_lims, minor_slots[2], slot_uid[2];
    DevPlan<float> pf{};
    DevPlan<double> pd{};
    std::vector<BandArgs> bands;
    DevBuf bands_dev;               // the same on the device (column-resident kernels loop over bands)
    BandLims lims{};
    Arena ws;
    int tile_cols = 0;
    int colres = -1;                // fused path: 1 column-resident, 0 strips, -1 default (RB200_PATH)
    std::shared_ptr<void> pipe;     // HostPipe of the host-buffer entry, created on first use
    std::recursive_mutex mu;        // HandleCall: one call at a time carves the workspace ...
    cudaEvent_t last_done = nullptr; // ... and its device work is ordered behind the previous call's, whatever their streams
    ~rrtmgp_b200_gas_optics() { if (last_done) cudaEventDestroy(last_done); }
};

// Every call that carves a handle's workspace (fused fluxes, materialised gas optics, the host pipeline) holds one of these: the
// arena is re-carved from offset 0 by each call, so two calls on one handle from different host threads or CUDA streams would
// otherwise overwrite each other's tile state and scratch strips.  Host side: the handle's mutex for the duration of the call
// (recursive: the host pipeline calls the fused entries).  Device side: the call's stream first waits for the event the
// previous call on this handle recorded after its last launch.  Calls on one handle therefore serialise; calls on different
// handles stay independent.
struct HandleCall
{
    rrtmgp_b200_gas_optics* g; cudaStream_t s;
    HandleCall(rrtmgp_b200_gas_optics* g_, cudaStream_t s_) : g(g_), s(s_)
    {
        g->mu.lock();
        if (!g->last_done) { if (cudaEventCreateWithFlags(&g->last_done, cudaEventDisableTiming) != cudaSuccess) { g->last_done = nullptr; cudaGetLastError(); } }
        else cudaStreamWaitEvent(s, g->last_done, 0);
    }
    ~HandleCall()
    {
        if (g->last_done) cudaEventRecord(g->last_done, s);
        g->mu.unlock();
    }
    HandleCall(const HandleCall&) = delete;
    HandleCall& operator=(const HandleCall&) = delete;
};

struct rrtmgp_b200_cloud_optics
{
    int dtype = 0;
    DevBuf t[6];
    CloudDev<float> cf{};
    CloudDev<double> cd{};
};

struct rrtmgp_b200_aerosol_optics
{
    int dtype = 0;
    DevBuf t[4];                    // rh_upper, mext, ssa, asy as [species][humidity class][band]
    AerosolDev<float> af{};
    AerosolDev<double> ad{};
};

template<typename Real> static const DevPlan<Real>& dplan(const rrtmgp_b200_gas_optics* g);
template<> const DevPlan<float>& dplan<float>(const rrtmgp_b200_gas_optics* g) { return g->pf; }
template<> const DevPlan<double>& dplan<double>(const rrtmgp_b200_gas_optics* g) { return g->pd; }
template<typename Real> static const CloudDev<Real>& dcloud(const rrtmgp_b200_cloud_optics* c);
template<> const CloudDev<float>& dcloud<float>(const rrtmgp_b200_cloud_optics* c) { return c->cf; }
template<> const CloudDev<double>& dcloud<double>(const rrtmgp_b200_cloud_optics* c) { return c->cd; }
template<typename Real> static const AerosolDev<Real>& daerosol(const rrtmgp_b200_aerosol_optics* a);
template<> const AerosolDev<float>& daerosol<float>(const rrtmgp_b200_aerosol_optics* a) { return a->af; }
template<> const AerosolDev<double>& daerosol<double>(const rrtmgp_b200_aerosol_optics* a) { return a->ad; }

// aermr01..aermr11 of the caller -> the kernel's species order (summation order of src/Aerosol_optics.cpp:51:
// SS1 SS2 SS3 DU1 DU2 DU3 OM1 OM2 BC1 BC2 SU; OM1 is aermr08 and OM2 aermr07, :131-142)
static const int kAerosolMmrOfSpecies[AEROSOL_SPECIES] = {0, 1, 2, 3, 4, 5, 7, 6, 8, 9, 10};
static AerosolMmr make_mmr(const rrtmgp_b200_aerosol_concs* ac, int ncol, int nlay)
{
    if (!ac) throw ArgError("aerosol optics: aerosol concentrations missing");
    AerosolMmr m{};
    for (int k = 0; k < AEROSOL_SPECIES; ++k)
    {
        const int i = kAerosolMmrOfSpecies[k];
        char name[16]; snprintf(name, sizeof name, "aermr%02d", i + 1);
        if (!ac->d_mmr[i]) throw MissingError(std::string("Aerosol_concs: ") + name + " not set");
        if ((ac->mmr_ncol[i] != 1 && ac->mmr_ncol[i] != ncol) || (ac->mmr_nlay[i] != 1 && ac->mmr_nlay[i] != nlay))
            throw ArgError(std::string("Aerosol_concs: ") + name + " must be a scalar, a (1,nlay) profile or an (ncol,nlay) field");
        m.p[k] = ac->d_mmr[i]; m.ncol[k] = ac->mmr_ncol[i]; m.nlay[k] = ac->mmr_nlay[i];
    }
    return m;
}

template<typename Real>
static void fill_devplan(rrtmgp_b200_gas_optics* g, DevPlan<Real>& P)
{
    const HostPlan& h = g->h;
    g->kmajor.upload_as<Real>(h.kmajor);
    if (h.is_lw) g->totplnk.upload_as<Real>(h.totplnk);
    else { g->solar.upload_as<Real>(h.solar_source); if (!h.krayl.empty()) g->krayl.upload_as<Real>(h.krayl); }
    for (int s = 0; s < 2; ++s) g->kminor[s].upload_as<Real>(h.minor[s].table);
    g->temp_ref.upload(h.temp_ref); g->vmr_ref.upload(h.vmr_ref);
    {
        // ratio_eta_half of interpolation (F90:137-139) depends on (side, flavor, reference temperature)
        // only: the same IEEE double division, done once here instead of per (column, layer) on the device
        std::vector<double> eh((size_t)2 * h.nflav * h.ntemp);
        for (int it = 0; it < h.ntemp; ++it)
            for (int f = 0; f < h.nflav; ++f)
                for (int sd = 0; sd < 2; ++sd)
                {
                    const int ig1 = h.flavor[2 * f], ig2 = h.flavor[2 * f + 1];
                    eh[sd + 2 * (f + (size_t)h.nflav * it)] = h.vmr_ref[sd + 2 * (ig1 + (size_t)(h.ngas + 1) * it)] /
                                                              h.vmr_ref[sd + 2 * (ig2 + (size_t)(h.ngas + 1) * it)];
                }
        g->eta_half.upload(eh);
    }
    g->flavor.upload(h.flavor); g->band_flavor.upload(h.band_flavor); g->band_lims.upload(h.band_lims_gpt);
    P.is_lw = h.is_lw; P.ntemp = h.ntemp; P.npres = h.npres; P.neta = h.neta; P.nbnd = h.nbnd; P.ngpt = h.ngpt;
    P.nflav = h.nflav; P.ngas = h.ngas; P.gw = h.gw; P.nplanck = h.nplanck; P.idx_h2o = h.idx_h2o;
    P.kmajor = (const Real*)g->kmajor.p; P.pfrac = h.is_lw ? (const Real*)g->kmajor.p + h.gw : nullptr; P.krayl = (const Real*)g->krayl.p;
    P.row_stride = h.row_stride;
    P.kminor[0] = (const Real*)g->kminor[0].p; P.kminor[1] = (const Real*)g->kminor[1].p;
    P.totplnk = (const Real*)g->totplnk.p; P.solar_source = (const Real*)g->solar.p;
    P.temp_ref = (const double*)g->temp_ref.p; P.vmr_ref = (const double*)g->vmr_ref.p; P.eta_half = (const double*)g->eta_half.p;
    P.flavor = (const int*)g->flavor.p; P.band_flavor = (const int*)g->band_flavor.p; P.band_lims = (const int*)g->band_lims.p;
    for (int sd = 0; sd < 2; ++sd)
    {
        std::vector<int> slots((size_t)4 * h.nuniq, 0);
        for (int k = 0; k < h.nuniq; ++k)
        {
            slots[4 * k] = -1;
            if (k < (int)h.minor[sd].uniq_rep.size())
            {
                const MinorDesc& dd = h.minor[sd].desc[h.minor[sd].uniq_rep[k]];
                slots[4 * k] = dd.idx_minor; slots[4 * k + 1] = dd.idx_scaling;
                slots[4 * k + 2] = (dd.scales_with_density ? 1 : 0) | (dd.scale_by_complement ? 2 : 0);
            }
        }
        std::vector<int> suid((size_t)h.nscal, h.nuniq);       // nuniq = the zero entry
        for (size_t k = 0; k < h.minor[sd].slot_desc.size(); ++k)
            if (h.minor[sd].slot_desc[k] >= 0) suid[k] = h.minor[sd].uid[h.minor[sd].slot_desc[k]];
        g->slot_uid[sd].upload(suid);
        P.slot_uid[sd] = (const int4*)g->slot_uid[sd].p;
        g->minor_slots[sd].upload(slots);
        P.minor_slots[sd] = (const int4*)g->minor_slots[sd].p;
    }
    P.nscal = h.nscal; P.nuniq = h.nuniq;
    P.press_ref_log0 = h.press_ref_log[0]; P.press_ref_log_delta = h.press_ref_log_delta;
    P.press_ref_trop_log = h.press_ref_trop_log; P.temp_ref_min = h.temp_ref_min; P.temp_ref_delta = h.temp_ref_delta;
    P.planck_tmin = (Real)h.temp_ref_min; P.planck_delta = (Real)h.totplnk_delta;
}

extern "C" {

const char* rrtmgp_b200_last_error(void) { return g_err.c_str(); }
int rrtmgp_b200_version(void) { return 100; }
int rrtmgp_b200_set_fused_path(rrtmgp_b200_gas_optics* g, int path)   /* -1 default (RB200_PATH, else by size), 0, 1, 2 = by size */
{
    return guarded([&] {
        if (!g || path < -1 || path > 2) throw ArgError("set_fused_path: handle and path in {-1, 0, 1, 2}");
        g->colres = path;
    }, false);
}
long long rrtmgp_b200_launch_count(void) { return g_launches.load(); }

int rrtmgp_b200_shard_bounds(long long ncol, int world, int rank, long long* c0, long long* c1)
{
    return guarded([&] {
        if (ncol < 0 || world < 1 || rank < 0 || rank >= world || !c0 || !c1) throw ArgError("shard_bounds: bad ncol / world / rank");
        *c0 = (long long)((__int128)rank * ncol / world);
        *c1 = (long long)((__int128)(rank + 1) * ncol / world);
    }, false);
}

int rrtmgp_b200_gas_optics_create(const rrtmgp_b200_kdist_file* kd, int n_avail, const char* names,
                                  int string_len, int dtype, rrtmgp_b200_gas_optics** out)
{
    if (!kd || !out || !names || n_avail <= 0) return fail(RRTMGP_B200_E_ARG, "gas_optics_create: null argument");
    const bool host_only = (dtype & RRTMGP_B200_HOST_ONLY) != 0;
    dtype &= ~RRTMGP_B200_HOST_ONLY;
    if (dtype != RRTMGP_B200_F32 && dtype != RRTMGP_B200_F64) return fail(RRTMGP_B200_E_ARG, "gas_optics_create: bad dtype");
    *out = nullptr;
    std::unique_ptr<rrtmgp_b200_gas_optics> g(new rrtmgp_b200_gas_optics);
    const int rc = guarded([&] {
        g->dtype = dtype;
        g->h = build_host_plan(*kd, n_avail, names, string_len);
        if (g->h.ngas > MAX_GAS) throw ArgError("too many gases");
        if (g->h.nbnd > MAX_BANDS) throw ArgError("too many bands");
        g->host_only = host_only;
        if (!host_only)
        {
            CK(cudaGetDevice(&g->device));
            if (dtype == RRTMGP_B200_F32) fill_devplan(g.get(), g->pf); else fill_devplan(g.get(), g->pd);
        }
        const HostPlan& h = g->h;
        g->lims.nbnd = h.nbnd;
        for (int i = 0; i < 2 * h.nbnd; ++i) g->lims.lims[i] = h.band_lims_gpt[i];
        g->bands.resize(h.nbnd);
        for (int b = 0; b < h.nbnd; ++b)
        {
            BandArgs& B = g->bands[b];
            B.band = b; B.gpt0 = h.band_lims_gpt[2 * b] - 1; B.ngb = h.band_lims_gpt[2 * b + 1] - B.gpt0;
            for (int s = 0; s < 2; ++s)
            {
                const auto& lst = h.minor[s].by_band[b];
                B.nminor[s] = (int)lst.size();
                B.first_minor[s] = lst.empty() ? 0 : h.minor[s].desc[lst[0]].contributor;
                B.scal_off[s] = h.minor[s].scal_off[b] / 4;
                B.flav[s] = h.band_flavor[s + 2 * b];
                for (size_t m = 0; m < lst.size(); ++m)
                    if (h.minor[s].desc[lst[m]].idx_minor < 0) throw ArgError("internal: minor gas index");
            }
        }
        if (!host_only) g->bands_dev.upload(g->bands);
    }, !host_only);
    if (rc) return rc;
    *out = g.release();
    return 0;
}

void rrtmgp_b200_gas_optics_destroy(rrtmgp_b200_gas_optics* g) { if (g) { if (!g->host_only) cudaDeviceSynchronize(); delete g; } }

int rrtmgp_b200_gas_optics_get_minor(const rrtmgp_b200_gas_optics* g, int side, int* idx_minor, int* idx_scaling,
                                     int* kstart, int* limits, int* dens, int* comp)
{
    if (!g || side < 0 || side > 1) return fail(RRTMGP_B200_E_ARG, "get_minor: bad argument");
    const MinorSide& m = g->h.minor[side];
    for (int i = 0; i < m.n; ++i)
    {
        if (idx_minor) idx_minor[i] = m.desc[i].idx_minor;
        if (idx_scaling) idx_scaling[i] = m.desc[i].idx_scaling;
        if (kstart) kstart[i] = m.kstart[i];
        if (limits) { limits[2 * i] = m.limits[2 * i]; limits[2 * i + 1] = m.limits[2 * i + 1]; }
        if (dens) dens[i] = m.desc[i].scales_with_density;
        if (comp) comp[i] = m.desc[i].scale_by_complement;
    }
    return 0;
}

int rrtmgp_b200_gas_optics_get_info(const rrtmgp_b200_gas_optics* g, rrtmgp_b200_gas_optics_info* i)
{
    if (!g || !i) return fail(RRTMGP_B200_E_ARG, "get_info: null argument");
    const HostPlan& h = g->h;
    i->dtype = g->dtype; i->is_longwave = h.is_lw; i->ngpt = h.ngpt; i->nbnd = h.nbnd; i->nflav = h.nflav;
    i->ngas = h.ngas; i->ntemp = h.ntemp; i->npres = h.npres; i->neta = h.neta;
    i->n_minor_lower = h.minor[0].n; i->n_minor_upper = h.minor[1].n; i->idx_h2o = h.idx_h2o; i->gpt_width = h.gw;
    i->press_ref_min = h.press_ref_min; i->press_ref_max = h.press_ref_max;
    i->temp_ref_min = h.temp_ref_min; i->temp_ref_max = h.temp_ref_max; i->tsi = h.tsi;
    return 0;
}

int rrtmgp_b200_gas_optics_get_tables(const rrtmgp_b200_gas_optics* g, int* band_lims_gpt, double* band_lims_wvn,
                                      int* gpoint_bands, int* flavor, int* gpoint_flavor, char* gas_names,
                                      int string_len, double* solar_source)
{
    if (!g) return fail(RRTMGP_B200_E_ARG, "get_tables: null handle");
    const HostPlan& h = g->h;
    if (band_lims_gpt) std::copy(h.band_lims_gpt.begin(), h.band_lims_gpt.end(), band_lims_gpt);
    if (band_lims_wvn) std::copy(h.band_lims_wvn.begin(), h.band_lims_wvn.end(), band_lims_wvn);
    if (gpoint_bands) std::copy(h.gpoint_bands.begin(), h.gpoint_bands.end(), gpoint_bands);
    if (flavor) std::copy(h.flavor.begin(), h.flavor.end(), flavor);
    if (gpoint_flavor) std::copy(h.gpoint_flavor.begin(), h.gpoint_flavor.end(), gpoint_flavor);
    if (gas_names)
        for (int i = 0; i < h.ngas; ++i)
        {
            std::memset(gas_names + (size_t)i * string_len, ' ', string_len);
            std::memcpy(gas_names + (size_t)i * string_len, h.gas_names[i].data(), std::min<size_t>(string_len, h.gas_names[i].size()));
        }
    if (solar_source && !h.solar_source.empty()) std::copy(h.solar_source.begin(), h.solar_source.end(), solar_source);
    return 0;
}

/* 64-bit FNV-1a over every prepared table of the handle (device-layout kmajor / planck_frac, Rayleigh, both minor
 * sides, Planck, solar source, reference grids and all index tables): two handles with equal hashes were built from
 * bit-identical coefficients and available-gas lists.  HOST function. */
int rrtmgp_b200_gas_optics_table_hash(const rrtmgp_b200_gas_optics* g, unsigned long long* out)
{
    if (!g || !out) return fail(RRTMGP_B200_E_ARG, "table_hash: null argument");
    unsigned long long hsh = 1469598103934665603ull;
    auto mix = [&](const void* p, size_t n) { const unsigned char* b = (const unsigned char*)p; for (size_t i = 0; i < n; ++i) { hsh ^= b[i]; hsh *= 1099511628211ull; } };
    auto vd = [&](const std::vector<double>& v) { const size_t n = v.size(); mix(&n, sizeof n); if (n) mix(v.data(), n * sizeof(double)); };
    auto vi = [&](const std::vector<int>& v) { const size_t n = v.size(); mix(&n, sizeof n); if (n) mix(v.data(), n * sizeof(int)); };
    const HostPlan& h = g->h;
    const int dims[] = {h.is_lw, h.ntemp, h.npres, h.neta, h.nbnd, h.ngpt, h.nflav, h.ngas, h.gw, h.nplanck, h.idx_h2o, h.row_stride, h.nscal, h.nuniq};
    mix(dims, sizeof dims);
    for (auto& s : h.gas_names) mix(s.data(), s.size() + 1);
    vi(h.band_lims_gpt); vd(h.band_lims_wvn); vi(h.gpoint_bands); vi(h.flavor); vi(h.gpoint_flavor); vi(h.band_flavor);
    vd(h.vmr_ref); vd(h.press_ref); vd(h.press_ref_log); vd(h.temp_ref);
    const double sc[] = {h.press_ref_log_delta, h.temp_ref_min, h.temp_ref_max, h.temp_ref_delta, h.press_ref_min, h.press_ref_max,
                         h.press_ref_trop_log, h.totplnk_delta, h.tsi};
    mix(sc, sizeof sc);
    vd(h.kmajor); vd(h.krayl); vd(h.totplnk); vd(h.solar_source);
    for (int s = 0; s < 2; ++s)
    {
        const MinorSide& m = h.minor[s];
        mix(&m.n, sizeof m.n); mix(&m.nk, sizeof m.nk);
        vi(m.limits); vi(m.kstart); vd(m.kminor_ref); vd(m.table); vi(m.uid); vi(m.uniq_rep); vi(m.scal_off); vi(m.slot_desc);
        if (!m.desc.empty()) mix(m.desc.data(), m.desc.size() * sizeof(m.desc[0]));
    }
    *out = hsh;
    return 0;
}

static double g_pair_ms = 0; static long long g_pair_launches = 0;
int rrtmgp_b200_profile(int enable, double* lw_ms, long long* lw_launches, double* sw_ms, long long* sw_launches)
{
    // collect: synchronises the device, folds the recorded event pairs into milliseconds
    cudaDeviceSynchronize();
    for (int k = 0; k < 3; ++k)
    {
        for (size_t i = 0; i + 1 < g_prof.ev[k].size(); i += 2)
        {
            float ms = 0; cudaEventElapsedTime(&ms, g_prof.ev[k][i], g_prof.ev[k][i + 1]); g_prof.ms[k] += ms;
        }
        for (cudaEvent_t e : g_prof.ev[k]) cudaEventDestroy(e);
        g_prof.ev[k].clear();
    }
    if (lw_ms) *lw_ms = g_prof.ms[0];
    if (lw_launches) *lw_launches = g_prof.launches[0];
    if (sw_ms) *sw_ms = g_prof.ms[1];
    if (sw_launches) *sw_launches = g_prof.launches[1];
    g_pair_ms = g_prof.ms[2]; g_pair_launches = g_prof.launches[2];
    for (int k = 0; k < 3; ++k) { g_prof.ms[k] = 0; g_prof.launches[k] = 0; }
    g_prof.on = enable != 0;
    return 0;
}
int rrtmgp_b200_profile_pair(double* pair_ms, long long* pair_launches)
{
    if (pair_ms) *pair_ms = g_pair_ms;
    if (pair_launches) *pair_launches = g_pair_launches;
    return 0;
}

int rrtmgp_b200_set_debug(rrtmgp_b200_gas_optics* g, int flags)
{
    if (!g || flags < 0 || flags > 3) return fail(RRTMGP_B200_E_ARG, "set_debug: handle and flags in 0..3");
    g->ws.debug = (flags & 1) != 0; g->ws.selftest = (flags & 3) == 3;
    return 0;
}

int rrtmgp_b200_set_tile_columns(rrtmgp_b200_gas_optics* g, int n)
{
    if (!g || n < 0) return fail(RRTMGP_B200_E_ARG, "set_tile_columns: bad argument");
    g->tile_cols = n; return 0;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
namespace {

VmrArgs make_vmr(const rrtmgp_b200_gas_optics* g, const rrtmgp_b200_gas_concs* c, int ncol, int nlay)
{
    if (g->host_only) throw CudaError("this handle was created HOST_ONLY: no device tables, and there is no CPU fallback");
    if (!c || c->ngas != g->h.ngas || !c->d_vmr || !c->vmr_ncol || !c->vmr_nlay)
        throw ArgError("gas concentrations: need one entry per gas of the k-distribution object");
    VmrArgs v{};
    for (int i = 0; i < c->ngas; ++i)
    {
        if (!c->d_vmr[i]) throw MissingError("Gas_concs: gas " + g->h.gas_names[i] + " not set");
        const int nc = c->vmr_ncol[i], nl = c->vmr_nlay[i];
        if (!((nc == 1 && nl == 1) || (nc == 1 && nl == nlay) || (nc == ncol && nl == nlay)))
            throw ArgError("Gas_concs: vmr of " + g->h.gas_names[i] + " must be (1,1), (1,nlay) or (ncol,nlay)");
        v.p[i] = c->d_vmr[i]; v.ncol[i] = (nc == 1 && ncol > 1) ? 1 : nc; v.nlay[i] = nl;
        if (ncol == 1 && nc == 1 && nl == nlay) { v.ncol[i] = 1; }
    }
    return v;
}

template<typename Real>
TileState<Real> carve_state(Arena& ws, const HostPlan& h, int nc, int nlay)
{
    TileState<Real> st{};
    const size_t tcl = (size_t)nc * nlay;
    st.nc = nc; st.nlay = nlay;
    st.cell = ws.take<V4<Real>>(tcl);
    st.flvA = ws.take<V4<Real>>(tcl * h.nflav);
    st.flvB = ws.take<V4<Real>>(tcl * h.nflav);
    st.scal = ws.take<V4<Real>>(tcl * (h.nscal / 4));
    st.pl_lay = nullptr; st.pl_lev = nullptr;
    if (h.is_lw) { st.pl_lay = ws.take<V2<Real>>(tcl); st.pl_lev = ws.take<V2<Real>>(tcl + nc); }
    st.tmw = tmask_words(nlay);
    st.tmask = ws.take<unsigned>((size_t)nc * st.tmw);
    return st;
}
template<typename Real> size_t state_bytes(const HostPlan& h, int nc, int nlay)
{
    const size_t tcl = (size_t)nc * nlay;
    return pad256(tcl * sizeof(V4<Real>)) + 2 * pad256(tcl * h.nflav * sizeof(V4<Real>))
         + pad256(tcl * h.nscal * sizeof(Real)) + pad256((size_t)nc * tmask_words(nlay) * sizeof(unsigned))
         + 2 * pad256((tcl + nc) * sizeof(V2<Real>)) + 1024;
}

// Pin the k-major table (the bulk of the table bytes) in L2 for the band kernels of one fused call: an
// access-policy window on the launching stream, persisting on hit.  The recurrence scratch streams tens
// of GB through L2 per launch and would otherwise keep evicting table lines.  RB200_L2_PIN=0 disables.
struct L2Pin
{
    cudaStream_t s; bool on = false;
    L2Pin(cudaStream_t s_, const void* base, size_t bytes) : s(s_)
    {
        static const bool enabled = [] { const char* e = getenv("RB200_L2_PIN"); return !(e && e[0] == '0'); }();
        static const size_t max_window = [] {
            int dev = 0, persist = 0, window = 0;
            if (cudaGetDevice(&dev) != cudaSuccess) return (size_t)0;
            cudaDeviceGetAttribute(&persist, cudaDevAttrMaxPersistingL2CacheSize, dev);
            cudaDeviceGetAttribute(&window, cudaDevAttrMaxAccessPolicyWindowSize, dev);
            if (persist <= 0 || window <= 0) return (size_t)0;
            const size_t set_aside = std::min<size_t>((size_t)persist, (size_t)32 << 20);
            if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, set_aside) != cudaSuccess) { cudaGetLastError(); return (size_t)0; }
            return std::min<size_t>((size_t)window, set_aside);
        }();
        if (!enabled || !max_window || !base || !bytes) return;
        cudaStreamAttrValue v{};
        v.accessPolicyWindow.base_ptr = const_cast<void*>(base);
        v.accessPolicyWindow.num_bytes = std::min(bytes, max_window);
        v.accessPolicyWindow.hitRatio = 1.0f;
        v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
        on = cudaStreamSetAttribute(s, cudaStreamAttributeAccessPolicyWindow, &v) == cudaSuccess;
        if (!on) cudaGetLastError();
    }
    ~L2Pin()
    {
        if (!on) return;
        cudaStreamAttrValue v{};
        v.accessPolicyWindow.num_bytes = 0;
        cudaStreamSetAttribute(s, cudaStreamAttributeAccessPolicyWindow, &v);
    }
};

void check_dims(int ncol, int nlay)
{
    if (ncol < 1 || nlay < 1) throw ArgError("ncol and nlay must be positive");
}

template<typename Real>
void gas_optics_materialised(rrtmgp_b200_gas_optics* g, int ncol, int nlay, const void* play, const void* plev,
                             const void* tlay, const void* tsfc, const void* tlev, const void* col_dry,
                             const rrtmgp_b200_gas_concs* concs, GasOut<Real> o, cudaStream_t s)
{
    check_dims(ncol, nlay);
    const HostPlan& h = g->h;
    if (!play || !tlay || !o.tau) throw ArgError("gas_optics: null array");
    HandleCall hc(g, s);
    const VmrArgs vmr = make_vmr(g, concs, ncol, nlay);
    const size_t tcl = (size_t)ncol * nlay;
    g->ws.reserve(state_bytes<Real>(h, ncol, nlay) + pad256(tcl * sizeof(Real)) + 2 * sizeof(Real) + 512, s);
    TileState<Real> st = carve_state<Real>(g->ws, h, ncol, nlay);
    st.tmask = nullptr;
    const void* cd = col_dry;
    if (!cd)
    {
        if (!plev) throw ArgError("gas_optics: plev is needed to compute col_dry");
        Real* t = g->ws.take<Real>(tcl);
        const int ih = h.idx_h2o - 1;
        if (!(concs->vmr_ncol[ih] == ncol && concs->vmr_nlay[ih] == nlay))
            throw ArgError("gas_optics: computing col_dry needs an (ncol,nlay) h2o field");
        launch_col_dry<Real>(ncol, nlay, concs->d_vmr[ih], plev, t, s);
        cd = t;
    }
    launch_interp<Real>(dplan<Real>(g), vmr, ncol, 0, ncol, nlay, play, tlay, cd, tlev, st,
                        nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, s);
    if (h.is_lw)
    {
        // sfc_lay from play(1,1) > play(1,nlay) (src/Gas_optics_rrtmgp.cpp:1333); one blocking read
        Real p2[2];
        CK(cudaMemcpyAsync(&p2[0], play, sizeof(Real), cudaMemcpyDeviceToHost, s));
        CK(cudaMemcpyAsync(&p2[1], (const Real*)play + (size_t)ncol * (nlay - 1), sizeof(Real), cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
        o.sfc_lay = p2[0] > p2[1] ? 0 : nlay - 1;
    }
    o.tlay = (const Real*)tlay; o.tlev = (const Real*)tlev; o.tsfc = (const Real*)tsfc;
    for (int b = 0; b < h.nbnd; ++b) launch_gas_band<Real>(dplan<Real>(g), g->bands[b], st, ncol, nlay, o, s);
}

BandLims make_lims(int nbnd, const int* lims, int ngpt)
{
    if (nbnd < 1 || nbnd > MAX_BANDS || !lims) throw ArgError("band limits: bad argument");
    BandLims bl{}; bl.nbnd = nbnd;
    for (int i = 0; i < 2 * nbnd; ++i) bl.lims[i] = lims[i];
    for (int b = 0; b < nbnd; ++b)
        if (lims[2 * b] < 1 || lims[2 * b + 1] > ngpt || lims[2 * b + 1] < lims[2 * b]) throw ArgError("band limits out of range");
    return bl;
}

}  // namespace

#define DISPATCH(dtype, CALL_F32, CALL_F64) \
    if ((dtype) == RRTMGP_B200_F32) { CALL_F32; } else if ((dtype) == RRTMGP_B200_F64) { CALL_F64; } \
    else throw ArgError("bad dtype")

extern "C" {

int rrtmgp_b200_get_col_dry(int dtype, int ncol, int nlay, const void* vmr_h2o, const void* plev, void* col_dry, void* stream)
{
    return guarded([&] {
        check_dims(ncol, nlay);
        if (!vmr_h2o || !plev || !col_dry) throw ArgError("get_col_dry: null array");
        DISPATCH(dtype, launch_col_dry<float>(ncol, nlay, vmr_h2o, plev, col_dry, (cudaStream_t)stream),
                        launch_col_dry<double>(ncol, nlay, vmr_h2o, plev, col_dry, (cudaStream_t)stream));
    });
}

int rrtmgp_b200_interpolation(const rrtmgp_b200_gas_optics* g, int ncol, int nlay, const void* play, const void* tlay,
                              const void* col_dry, const rrtmgp_b200_gas_concs* concs, int* jtemp, int* jpress,
                              signed char* tropo, int* jeta, void* col_mix, void* fminor, void* fmajor,
                              void* col_gas, void* stream)
{
    return guarded([&] {
        if (!g) throw ArgError("interpolation: null handle");
        check_dims(ncol, nlay);
        if (!play || !tlay || !col_dry) throw ArgError("interpolation: null array");
        const VmrArgs vmr = make_vmr(g, concs, ncol, nlay);
        DISPATCH(g->dtype,
            launch_interp<float>(g->pf, vmr, ncol, 0, ncol, nlay, play, tlay, col_dry, nullptr, TileState<float>{}, jtemp, jpress, tropo, jeta, col_mix, fminor, fmajor, col_gas, (cudaStream_t)stream),
            launch_interp<double>(g->pd, vmr, ncol, 0, ncol, nlay, play, tlay, col_dry, nullptr, TileState<double>{}, jtemp, jpress, tropo, jeta, col_mix, fminor, fmajor, col_gas, (cudaStream_t)stream));
    });
}

int rrtmgp_b200_gas_optics_lw(const rrtmgp_b200_gas_optics* g_, int ncol, int nlay, const void* play, const void* plev,
                              const void* tlay, const void* tsfc, const void* tlev, const void* col_dry,
                              const rrtmgp_b200_gas_concs* concs, void* tau, void* lay_src, void* lev_src,
                              void* sfc_src, void* sfc_src_jac, void* stream)
{
    return guarded([&] {
        auto* g = const_cast<rrtmgp_b200_gas_optics*>(g_);
        if (!g || !g->h.is_lw) throw ArgError("gas_optics_lw: needs a longwave k-distribution");
        if (!tsfc || !tlev || !lay_src || !lev_src || !sfc_src || !sfc_src_jac) throw ArgError("gas_optics_lw: null array");
        if (g->dtype == RRTMGP_B200_F32)
        {
            GasOut<float> o{}; o.tau = (float*)tau; o.lay_src = (float*)lay_src; o.lev_src = (float*)lev_src;
            o.sfc_src = (float*)sfc_src; o.sfc_src_jac = (float*)sfc_src_jac;
            gas_optics_materialised<float>(g, ncol, nlay, play, plev, tlay, tsfc, tlev, col_dry, concs, o, (cudaStream_t)stream);
        }
        else
        {
            GasOut<double> o{}; o.tau = (double*)tau; o.lay_src = (double*)lay_src; o.lev_src = (double*)lev_src;
            o.sfc_src = (double*)sfc_src; o.sfc_src_jac = (double*)sfc_src_jac;
            gas_optics_materialised<double>(g, ncol, nlay, play, plev, tlay, tsfc, tlev, col_dry, concs, o, (cudaStream_t)stream);
        }
    });
}

int rrtmgp_b200_gas_optics_sw(const rrtmgp_b200_gas_optics* g_, int ncol, int nlay, const void* play, const void* plev,
                              const void* tlay, const void* col_dry, const rrtmgp_b200_gas_concs* concs,
                              void* tau, void* ssa, void* gg, void* toa_src, void* stream)
{
    return guarded([&] {
        auto* g = const_cast<rrtmgp_b200_gas_optics*>(g_);
        if (!g || g->h.is_lw) throw ArgError("gas_optics_sw: needs a shortwave k-distribution");
        if ((ssa == nullptr) != (gg == nullptr)) throw ArgError("gas_optics_sw: ssa and g go together");
        if (g->dtype == RRTMGP_B200_F32)
        {
            GasOut<float> o{}; o.tau = (float*)tau; o.ssa = (float*)ssa; o.g = (float*)gg; o.toa_src = (float*)toa_src;
            gas_optics_materialised<float>(g, ncol, nlay, play, plev, tlay, nullptr, nullptr, col_dry, concs, o, (cudaStream_t)stream);
        }
        else
        {
            GasOut<double> o{}; o.tau = (double*)tau; o.ssa = (double*)ssa; o.g = (double*)gg; o.toa_src = (double*)toa_src;
            gas_optics_materialised<double>(g, ncol, nlay, play, plev, tlay, nullptr, nullptr, col_dry, concs, o, (cudaStream_t)stream);
        }
    });
}

int rrtmgp_b200_check_range(int dtype, size_t n, const void* x, double lo, double hi, const char* name, void* stream)
{
    return guarded([&] {
        if (!x || !name) throw ArgError("check_range: null argument");
        cudaStream_t s = (cudaStream_t)stream;
        int* d_flag = nullptr; int flag = 0;
        CK(cudaMallocAsync((void**)&d_flag, sizeof(int), s));
        CK(cudaMemsetAsync(d_flag, 0, sizeof(int), s));
        DISPATCH(dtype, launch_range_flag<float>(n, x, lo, hi, d_flag, s), launch_range_flag<double>(n, x, lo, hi, d_flag, s));
        CK(cudaMemcpyAsync(&flag, d_flag, sizeof(int), cudaMemcpyDeviceToHost, s));
        CK(cudaFreeAsync(d_flag, s));
        CK(cudaStreamSynchronize(s));
        if (flag) throw RangeError(std::string(name) + " is out of range");
    });
}

// ---- cloud optics
int rrtmgp_b200_cloud_optics_create(const rrtmgp_b200_cloud_file* lut, int dtype, rrtmgp_b200_cloud_optics** out)
{
    if (!lut || !out) return fail(RRTMGP_B200_E_ARG, "cloud_optics_create: null argument");
    *out = nullptr;
    std::unique_ptr<rrtmgp_b200_cloud_optics> c(new rrtmgp_b200_cloud_optics);
    const int rc = guarded([&] {
        if (dtype != RRTMGP_B200_F32 && dtype != RRTMGP_B200_F64) throw ArgError("bad dtype");
        if (lut->n_rghice < 2 || lut->n_size_liq < 2 || lut->n_size_ice < 2 || lut->n_band < 1) throw ArgError("cloud LUT: bad dimensions");
        c->dtype = dtype;
        const int nb = lut->n_band, nl = lut->n_size_liq, ni = lut->n_size_ice;
        const int icergh = 2;                                       // src/Cloud_optics.cpp:59
        const double* src[6] = {lut->lut_extliq, lut->lut_ssaliq, lut->lut_asyliq,
                                lut->lut_extice + (size_t)(icergh - 1) * nb * ni,
                                lut->lut_ssaice + (size_t)(icergh - 1) * nb * ni,
                                lut->lut_asyice + (size_t)(icergh - 1) * nb * ni};
        for (int k = 0; k < 6; ++k)
        {
            std::vector<double> v(src[k], src[k] + (size_t)nb * (k < 3 ? nl : ni));
            if (dtype == RRTMGP_B200_F32) c->t[k].upload_as<float>(v); else c->t[k].upload_as<double>(v);
        }
        const double liq_step = (lut->radliq_upr - lut->radliq_lwr) / (nl - 1.0);
        const double ice_step = (lut->diamice_upr - lut->diamice_lwr) / (ni - 1.0);
        c->cf = CloudDev<float>{nb, nl, ni, (float)liq_step, (float)ice_step, (float)lut->radliq_lwr, (float)lut->diamice_lwr,
                                (const float*)c->t[0].p, (const float*)c->t[1].p, (const float*)c->t[2].p,
                                (const float*)c->t[3].p, (const float*)c->t[4].p, (const float*)c->t[5].p};
        c->cd = CloudDev<double>{nb, nl, ni, liq_step, ice_step, lut->radliq_lwr, lut->diamice_lwr,
                                 (const double*)c->t[0].p, (const double*)c->t[1].p, (const double*)c->t[2].p,
                                 (const double*)c->t[3].p, (const double*)c->t[4].p, (const double*)c->t[5].p};
    });
    if (rc) return rc;
    *out = c.release();
    return 0;
}
void rrtmgp_b200_cloud_optics_destroy(rrtmgp_b200_cloud_optics* c) { if (c) { cudaDeviceSynchronize(); delete c; } }
int rrtmgp_b200_cloud_optics_nband(const rrtmgp_b200_cloud_optics* c) { return c ? c->cf.nbnd : -1; }

int rrtmgp_b200_cloud_optics_2str(const rrtmgp_b200_cloud_optics* c, int ncol, int nlay, const void* lwp, const void* iwp,
                                  const void* rel, const void* dei, void* tau, void* ssa, void* g, void* stream)
{
    return guarded([&] {
        if (!c || !lwp || !iwp || !rel || !dei || !tau || !ssa || !g) throw ArgError("cloud_optics: null argument");
        check_dims(ncol, nlay);
        DISPATCH(c->dtype,
            launch_cloud_optics<float>(c->cf, true, ncol, 0, ncol, nlay, lwp, iwp, rel, dei, ncol, 0, 0, tau, ssa, g, (cudaStream_t)stream),
            launch_cloud_optics<double>(c->cd, true, ncol, 0, ncol, nlay, lwp, iwp, rel, dei, ncol, 0, 0, tau, ssa, g, (cudaStream_t)stream));
    });
}
int rrtmgp_b200_cloud_optics_1scl(const rrtmgp_b200_cloud_optics* c, int ncol, int nlay, const void* lwp, const void* iwp,
                                  const void* rel, const void* dei, void* tau, void* stream)
{
    return guarded([&] {
        if (!c || !lwp || !iwp || !rel || !dei || !tau) throw ArgError("cloud_optics: null argument");
        check_dims(ncol, nlay);
        DISPATCH(c->dtype,
            launch_cloud_optics<float>(c->cf, false, ncol, 0, ncol, nlay, lwp, iwp, rel, dei, ncol, 0, 0, tau, nullptr, nullptr, (cudaStream_t)stream),
            launch_cloud_optics<double>(c->cd, false, ncol, 0, ncol, nlay, lwp, iwp, rel, dei, ncol, 0, 0, tau, nullptr, nullptr, (cudaStream_t)stream));
    });
}

// ---- aerosol optics
int rrtmgp_b200_aerosol_optics_create(const rrtmgp_b200_aerosol_file* lut, int dtype, rrtmgp_b200_aerosol_optics** out)
{
    if (!lut || !out) return fail(RRTMGP_B200_E_ARG, "aerosol_optics_create: null argument");
    *out = nullptr;
    std::unique_ptr<rrtmgp_b200_aerosol_optics> a(new rrtmgp_b200_aerosol_optics);
    const int rc = guarded([&] {
        if (dtype != RRTMGP_B200_F32 && dtype != RRTMGP_B200_F64) throw ArgError("bad dtype");
        const int nb = lut->n_band, nh = lut->n_hum, nphil = lut->n_philic, nphob = lut->n_phobic;
        // the reference addresses hydrophobic columns 1, 6, 8, 10, 11 and hydrophilic types 1..5 (src/Aerosol_optics.cpp:64-142)
        if (nb < 1 || nh < 1 || nphil < 5 || nphob < 11) throw ArgError("aerosol LUT: bad dimensions (need >= 5 hydrophilic, >= 11 hydrophobic types)");
        if (!lut->mass_ext_sw_hydrophobic || !lut->ssa_sw_hydrophobic || !lut->asymmetry_sw_hydrophobic || !lut->mass_ext_sw_hydrophilic ||
            !lut->ssa_sw_hydrophilic || !lut->asymmetry_sw_hydrophilic || !lut->relative_humidity2) throw ArgError("aerosol LUT: null table");
        a->dtype = dtype;
        // species in kernel order; > 0: hydrophobic column (1-based), < 0: hydrophilic type (1-based)
        static const int column[AEROSOL_SPECIES] = {-1, -2, -3, 1, 8, 6, 10, -4, 11, 11, -5};
        const double* phob[3] = {lut->mass_ext_sw_hydrophobic, lut->ssa_sw_hydrophobic, lut->asymmetry_sw_hydrophobic};
        const double* phil[3] = {lut->mass_ext_sw_hydrophilic, lut->ssa_sw_hydrophilic, lut->asymmetry_sw_hydrophilic};
        std::vector<double> rh(lut->relative_humidity2, lut->relative_humidity2 + nh);
        if (dtype == RRTMGP_B200_F32) a->t[0].upload_as<float>(rh); else a->t[0].upload_as<double>(rh);
        for (int p = 0; p < 3; ++p)
        {
            std::vector<double> v((size_t)AEROSOL_SPECIES * nh * nb);
            for (int k = 0; k < AEROSOL_SPECIES; ++k)
                for (int ih = 0; ih < nh; ++ih)
                    for (int b = 0; b < nb; ++b)
                        v[((size_t)k * nh + ih) * nb + b] = column[k] > 0 ? phob[p][(size_t)(column[k] - 1) * nb + b]
                                                                          : phil[p][((size_t)(-column[k] - 1) * nh + ih) * nb + b];
            if (dtype == RRTMGP_B200_F32) a->t[1 + p].upload_as<float>(v); else a->t[1 + p].upload_as<double>(v);
        }
        a->af = AerosolDev<float>{nb, nh, (const float*)a->t[0].p, (const float*)a->t[1].p, (const float*)a->t[2].p, (const float*)a->t[3].p};
        a->ad = AerosolDev<double>{nb, nh, (const double*)a->t[0].p, (const double*)a->t[1].p, (const double*)a->t[2].p, (const double*)a->t[3].p};
    });
    if (rc) return rc;
    *out = a.release();
    return 0;
}
void rrtmgp_b200_aerosol_optics_destroy(rrtmgp_b200_aerosol_optics* a) { if (a) { cudaDeviceSynchronize(); delete a; } }
int rrtmgp_b200_aerosol_optics_nband(const rrtmgp_b200_aerosol_optics* a) { return a ? a->af.nbnd : -1; }

int rrtmgp_b200_aerosol_optics_2str(const rrtmgp_b200_aerosol_optics* a, int ncol, int nlay, const rrtmgp_b200_aerosol_concs* concs,
                                    const void* rh, const void* plev, void* tau, void* ssa, void* g, void* stream)
{
    return guarded([&] {
        if (!a || !rh || !plev || !tau || !ssa || !g) throw ArgError("aerosol_optics: null argument");
        check_dims(ncol, nlay);
        const AerosolMmr m = make_mmr(concs, ncol, nlay);
        DISPATCH(a->dtype,
            launch_aerosol_optics<float>(a->af, m, false, ncol, 0, ncol, nlay, rh, plev, ncol, 0, 0, tau, ssa, g, (cudaStream_t)stream),
            launch_aerosol_optics<double>(a->ad, m, false, ncol, 0, ncol, nlay, rh, plev, ncol, 0, 0, tau, ssa, g, (cudaStream_t)stream));
    });
}

}  // extern "C"

extern "C" {

int rrtmgp_b200_delta_scale_2str(int dtype, int ncol, int nlay, int ngpt, void* tau, void* ssa, void* g, void* stream)
{
    return guarded([&] {
        check_dims(ncol, nlay);
        if (ngpt < 1 || !tau || !ssa || !g) throw ArgError("delta_scale: bad argument");
        const size_t n = (size_t)ncol * nlay * ngpt;
        DISPATCH(dtype, launch_delta_scale<float>(n, tau, ssa, g, (cudaStream_t)stream),
                        launch_delta_scale<double>(n, tau, ssa, g, (cudaStream_t)stream));
    });
}

static int increment_impl(int dtype, bool twostr, int ncol, int nlay, int ngpt, void* tau1, void* ssa1, void* g1,
                          const void* tau2, const void* ssa2, const void* g2, int nbnd, const int* lims, void* stream)
{
    return guarded([&] {
        check_dims(ncol, nlay);
        if (ngpt < 1 || !tau1 || !tau2 || (twostr && (!ssa1 || !g1 || !ssa2 || !g2))) throw ArgError("increment: bad argument");
        BandLims bl{};
        if (nbnd > 0) bl = make_lims(nbnd, lims, ngpt);
        const size_t ncl = (size_t)ncol * nlay;
        DISPATCH(dtype, launch_increment<float>(twostr, ncl, ngpt, bl, tau1, ssa1, g1, tau2, ssa2, g2, (cudaStream_t)stream),
                        launch_increment<double>(twostr, ncl, ngpt, bl, tau1, ssa1, g1, tau2, ssa2, g2, (cudaStream_t)stream));
    });
}
int rrtmgp_b200_increment_1scl(int dtype, int ncol, int nlay, int ngpt, void* tau1, const void* tau2, void* stream)
{ return increment_impl(dtype, false, ncol, nlay, ngpt, tau1, nullptr, nullptr, tau2, nullptr, nullptr, 0, nullptr, stream); }
int rrtmgp_b200_increment_2str(int dtype, int ncol, int nlay, int ngpt, void* tau1, void* ssa1, void* g1,
                               const void* tau2, const void* ssa2, const void* g2, void* stream)
{ return increment_impl(dtype, true, ncol, nlay, ngpt, tau1, ssa1, g1, tau2, ssa2, g2, 0, nullptr, stream); }
int rrtmgp_b200_increment_1scl_bybnd(int dtype, int ncol, int nlay, int ngpt, void* tau1, const void* tau2,
                                     int nbnd, const int* lims, void* stream)
{
    if (nbnd < 1) return fail(RRTMGP_B200_E_ARG, "increment_bybnd: nbnd must be positive");
    return increment_impl(dtype, false, ncol, nlay, ngpt, tau1, nullptr, nullptr, tau2, nullptr, nullptr, nbnd, lims, stream);
}
int rrtmgp_b200_increment_2str_bybnd(int dtype, int ncol, int nlay, int ngpt, void* tau1, void* ssa1, void* g1,
                                     const void* tau2, const void* ssa2, const void* g2, int nbnd, const int* lims, void* stream)
{
    if (nbnd < 1) return fail(RRTMGP_B200_E_ARG, "increment_bybnd: nbnd must be positive");
    return increment_impl(dtype, true, ncol, nlay, ngpt, tau1, ssa1, g1, tau2, ssa2, g2, nbnd, lims, stream);
}

int rrtmgp_b200_expand_and_transpose(int dtype, int ncol, int nbnd, int ngpt, const int* lims, const void* in, void* out, void* stream)
{
    return guarded([&] {
        if (ncol < 1 || !in || !out) throw ArgError("expand_and_transpose: bad argument");
        const BandLims bl = make_lims(nbnd, lims, ngpt);
        DISPATCH(dtype, launch_expand<float>(ncol, bl, in, out, (cudaStream_t)stream),
                        launch_expand<double>(ncol, bl, in, out, (cudaStream_t)stream));
    });
}

// rte_lw_solver_noscat; ssa and g non-null select its do_rescaling branch
static int lw_solver_any(int dtype, int ncol, int nlay, int ngpt, int top_at_1, int nmus, const void* tau,
                         const void* ssa, const void* g,
                         const void* lay_src, const void* lev_src, const void* sfc_emis, const void* sfc_src,
                         const void* inc_flux, void* flux_up, void* flux_dn, int do_broadband,
                         const void* sfc_src_jac, void* flux_up_jac, void* stream)
{
    return guarded([&] {
        check_dims(ncol, nlay);
        if ((ssa == nullptr) != (g == nullptr)) throw ArgError("lw_solver: ssa and g go together");
        if (ngpt < 1 || !tau || !lay_src || !lev_src || !sfc_emis || !sfc_src || !flux_up || !flux_dn) throw ArgError("lw_solver: null array");
        if (nmus < 1 || nmus > 4) throw ArgError("lw_solver: n_gauss_angles must be 1..4");   // src/Rte_lw.cpp:139-151
        if ((sfc_src_jac == nullptr) != (flux_up_jac == nullptr)) throw ArgError("lw_solver: Jacobian input and output go together");
        if (dtype != RRTMGP_B200_F32 && dtype != RRTMGP_B200_F64) throw ArgError("bad dtype");      // before anything is allocated
        cudaStream_t s = (cudaStream_t)stream;
        const size_t esz = dtype == RRTMGP_B200_F32 ? 4 : 8;
        const size_t ncv = (size_t)ncol * (nlay + 1);
        void *up = flux_up, *dn = flux_dn;
        AsyncTemp temp(s);
        const size_t fb = ncv * ngpt * esz;                       // one (ncol, nlay+1, ngpt) array
        // the surface-temperature Jacobian is computed per g-point and summed in g-point order afterwards (deterministic; it
        // used to be accumulated with atomicAdd across g-point blocks)
        const size_t nfl = do_broadband ? 2 : 0, nwk = ssa ? 2 : 0, njac = flux_up_jac ? 1 : 0;
        temp.alloc((nfl + nwk + njac) * fb);
        char* const tmp = temp.p;
        void* const jac_gpt = njac ? tmp + (nfl + nwk) * fb : nullptr;
        if (do_broadband) { up = tmp; dn = tmp + fb; }
        void* work2 = tmp + nfl * fb;
        if (ssa)
        {
            DISPATCH(dtype,
                launch_lw_solver_rescl<float>(ncol, nlay, ngpt, top_at_1, nmus, tau, ssa, g, lay_src, lev_src, sfc_emis, sfc_src, inc_flux, up, dn, sfc_src_jac, jac_gpt, work2, s),
                launch_lw_solver_rescl<double>(ncol, nlay, ngpt, top_at_1, nmus, tau, ssa, g, lay_src, lev_src, sfc_emis, sfc_src, inc_flux, up, dn, sfc_src_jac, jac_gpt, work2, s));
        }
        else
        {
            DISPATCH(dtype,
                launch_lw_solver<float>(ncol, nlay, ngpt, top_at_1, nmus, tau, lay_src, lev_src, sfc_emis, sfc_src, inc_flux, up, dn, sfc_src_jac, jac_gpt, s),
                launch_lw_solver<double>(ncol, nlay, ngpt, top_at_1, nmus, tau, lay_src, lev_src, sfc_emis, sfc_src, inc_flux, up, dn, sfc_src_jac, jac_gpt, s));
        }
        if (flux_up_jac)
        { DISPATCH(dtype, launch_sum_broadband<float>(ncol, nlay + 1, ngpt, jac_gpt, flux_up_jac, s), launch_sum_broadband<double>(ncol, nlay + 1, ngpt, jac_gpt, flux_up_jac, s)); }
        if (do_broadband)
        {
            DISPATCH(dtype, (launch_sum_broadband<float>(ncol, nlay + 1, ngpt, up, flux_up, s), launch_sum_broadband<float>(ncol, nlay + 1, ngpt, dn, flux_dn, s)),
                            (launch_sum_broadband<double>(ncol, nlay + 1, ngpt, up, flux_up, s), launch_sum_broadband<double>(ncol, nlay + 1, ngpt, dn, flux_dn, s)));
        }
    });
}

int rrtmgp_b200_lw_solver_noscat(int dtype, int ncol, int nlay, int ngpt, int top_at_1, int nmus, const void* tau,
                                 const void* lay_src, const void* lev_src, const void* sfc_emis, const void* sfc_src,
                                 const void* inc_flux, void* flux_up, void* flux_dn, int do_broadband,
                                 const void* sfc_src_jac, void* flux_up_jac, void* stream)
{
    return lw_solver_any(dtype, ncol, nlay, ngpt, top_at_1, nmus, tau, nullptr, nullptr, lay_src, lev_src, sfc_emis, sfc_src,
                         inc_flux, flux_up, flux_dn, do_broadband, sfc_src_jac, flux_up_jac, stream);
}

int rrtmgp_b200_lw_solver_rescaled(int dtype, int ncol, int nlay, int ngpt, int top_at_1, int nmus, const void* tau,
                                   const void* ssa, const void* g, const void* lay_src, const void* lev_src,
                                   const void* sfc_emis, const void* sfc_src, const void* inc_flux, void* flux_up,
                                   void* flux_dn, int do_broadband, const void* sfc_src_jac, void* flux_up_jac, void* stream)
{
    if (!ssa || !g) return guarded([&] { throw ArgError("lw_solver_rescaled: ssa and g are required"); });
    return lw_solver_any(dtype, ncol, nlay, ngpt, top_at_1, nmus, tau, ssa, g, lay_src, lev_src, sfc_emis, sfc_src,
                         inc_flux, flux_up, flux_dn, do_broadband, sfc_src_jac, flux_up_jac, stream);
}

static int sw_solver_any(int dtype, int ncol, int nlay, int ngpt, int top_at_1, int mu0_by_layer, const void* tau, const void* ssa,
                         const void* g, const void* mu0, const void* alb_dir, const void* alb_dif,
                         const void* inc_dir, const void* inc_dif, void* flux_up, void* flux_dn, void* flux_dir,
                         int do_broadband, void* stream)
{
    return guarded([&] {
        check_dims(ncol, nlay);
        if (ngpt < 1 || !tau || !ssa || !g || !mu0 || !alb_dir || !alb_dif || !inc_dir || !flux_up || !flux_dn || !flux_dir)
            throw ArgError("sw_solver: null array");
        if (dtype != RRTMGP_B200_F32 && dtype != RRTMGP_B200_F64) throw ArgError("bad dtype");      // before anything is allocated
        cudaStream_t s = (cudaStream_t)stream;
        const size_t esz = dtype == RRTMGP_B200_F32 ? 4 : 8;
        const size_t ncv = (size_t)ncol * (nlay + 1), ncl = (size_t)ncol * nlay;
        AsyncTemp temp(s);
        const size_t wbytes = 4 * ncl * ngpt * esz, fbytes = ncv * ngpt * esz;
        temp.alloc(wbytes + (do_broadband ? 3 * fbytes : 0));
        char* const tmp = temp.p;
        void *up = flux_up, *dn = flux_dn, *dr = flux_dir;
        if (do_broadband) { up = tmp + wbytes; dn = tmp + wbytes + fbytes; dr = tmp + wbytes + 2 * fbytes; }
        DISPATCH(dtype,
            launch_sw_solver<float>(ncol, nlay, ngpt, top_at_1, mu0_by_layer, tau, ssa, g, mu0, alb_dir, alb_dif, inc_dir, inc_dif, up, dn, dr, tmp, s),
            launch_sw_solver<double>(ncol, nlay, ngpt, top_at_1, mu0_by_layer, tau, ssa, g, mu0, alb_dir, alb_dif, inc_dir, inc_dif, up, dn, dr, tmp, s));
        if (do_broadband)
        {
            void* src[3] = {up, dn, dr}; void* dst[3] = {flux_up, flux_dn, flux_dir};
            for (int k = 0; k < 3; ++k)
            { DISPATCH(dtype, launch_sum_broadband<float>(ncol, nlay + 1, ngpt, src[k], dst[k], s), launch_sum_broadband<double>(ncol, nlay + 1, ngpt, src[k], dst[k], s)); }
        }
    });
}

int rrtmgp_b200_sw_solver_2stream(int dtype, int ncol, int nlay, int ngpt, int top_at_1, const void* tau, const void* ssa,
                                  const void* g, const void* mu0, const void* alb_dir, const void* alb_dif,
                                  const void* inc_dir, const void* inc_dif, void* flux_up, void* flux_dn, void* flux_dir,
                                  int do_broadband, void* stream)
{
    return sw_solver_any(dtype, ncol, nlay, ngpt, top_at_1, 0, tau, ssa, g, mu0, alb_dir, alb_dif, inc_dir, inc_dif,
                         flux_up, flux_dn, flux_dir, do_broadband, stream);
}

int rrtmgp_b200_sw_solver_2stream_mu0lay(int dtype, int ncol, int nlay, int ngpt, int top_at_1, const void* tau, const void* ssa,
                                         const void* g, const void* mu0_lay, const void* alb_dir, const void* alb_dif,
                                         const void* inc_dir, const void* inc_dif, void* flux_up, void* flux_dn,
                                         void* flux_dir, int do_broadband, void* stream)
{
    return sw_solver_any(dtype, ncol, nlay, ngpt, top_at_1, 1, tau, ssa, g, mu0_lay, alb_dir, alb_dif, inc_dir, inc_dif,
                         flux_up, flux_dn, flux_dir, do_broadband, stream);
}

int rrtmgp_b200_sum_broadband(int dtype, int ncol, int nlev, int ngpt, const void* gpt_flux, void* flux, void* stream)
{
    return guarded([&] {
        if (ncol < 1 || nlev < 1 || ngpt < 1 || !gpt_flux || !flux) throw ArgError("sum_broadband: bad argument");
        DISPATCH(dtype, launch_sum_broadband<float>(ncol, nlev, ngpt, gpt_flux, flux, (cudaStream_t)stream),
                        launch_sum_broadband<double>(ncol, nlev, ngpt, gpt_flux, flux, (cudaStream_t)stream));
    });
}
int rrtmgp_b200_sum_byband(int dtype, int ncol, int nlev, int ngpt, int nbnd, const int* lims, const void* gpt_flux,
                           void* bnd_flux, void* stream)
{
    return guarded([&] {
        if (ncol < 1 || nlev < 1 || !gpt_flux || !bnd_flux) throw ArgError("sum_byband: bad argument");
        const BandLims bl = make_lims(nbnd, lims, ngpt);
        DISPATCH(dtype, launch_sum_byband<float>(ncol, nlev, bl, gpt_flux, bnd_flux, (cudaStream_t)stream),
                        launch_sum_byband<double>(ncol, nlev, bl, gpt_flux, bnd_flux, (cudaStream_t)stream));
    });
}
int rrtmgp_b200_net_flux(int dtype, size_t n, const void* dn, const void* up, void* net, void* stream)
{
    return guarded([&] {
        if (!dn || !up || !net) throw ArgError("net_flux: null array");
        DISPATCH(dtype, launch_net<float>(n, dn, up, net, (cudaStream_t)stream), launch_net<double>(n, dn, up, net, (cudaStream_t)stream));
    });
}
int rrtmgp_b200_scale_columns(int dtype, int ncol, int n2, void* x, const void* scale, void* stream)
{
    return guarded([&] {
        if (ncol < 1 || n2 < 1 || !x || !scale) throw ArgError("scale_columns: bad argument");
        DISPATCH(dtype, launch_scale_columns<float>(ncol, n2, x, scale, (cudaStream_t)stream),
                        launch_scale_columns<double>(ncol, n2, x, scale, (cudaStream_t)stream));
    });
}
/* Array_gpu::fill / Array_gpu::subset (include/Array.h:311-350, 515-524, 579-622) */
int rrtmgp_b200_array_fill(int elem_bytes, long long n, void* d_a, const void* value, void* stream)
{
    return guarded([&] {
        if (n < 0 || (n > 0 && !d_a) || !value) throw ArgError("array_fill: bad argument");
        if (n == 0) return;
        if (!launch_array_fill(elem_bytes, d_a, value, (size_t)n, (cudaStream_t)stream)) throw ArgError("array_fill: element size must be 1, 2, 4, 8 or 16 bytes");
    });
}
int rrtmgp_b200_array_subset(int elem_bytes, int ndim, const int* src_dims, const int* src_offsets, const int* first,
                             const int* sub_dims, const void* d_src, void* d_sub, void* stream)
{
    return guarded([&] {
        if (ndim < 1 || ndim > SubsetGeom::MAXD || !src_dims || !first || !sub_dims || !d_src || !d_sub) throw ArgError("array_subset: bad argument");
        SubsetGeom g{};
        g.ndim = ndim;
        size_t n = 1, stride = 1;
        for (int k = 0; k < ndim; ++k)
        {
            const int off = src_offsets ? src_offsets[k] : 0;
            g.extent[k] = sub_dims[k]; g.spread[k] = src_dims[k] == 1; g.stride[k] = stride;
            g.start0[k] = first[k] - off - 1;
            if (sub_dims[k] < 1 || src_dims[k] < 1) throw ArgError("array_subset: empty dimension");
            if (!g.spread[k] && (g.start0[k] < 0 || g.start0[k] + sub_dims[k] > src_dims[k])) throw ArgError("array_subset: range outside the array");
            if (g.spread[k]) g.start0[k] = -off;          // index 1 of a broadcast dimension (include/Array.h:329)
            n *= (size_t)sub_dims[k]; stride *= (size_t)src_dims[k];
        }
        if (!launch_array_subset(elem_bytes, d_sub, d_src, g, n, (cudaStream_t)stream)) throw ArgError("array_subset: element size must be 1, 2, 4, 8 or 16 bytes");
    });
}
int rrtmgp_b200_heating_rate(int dtype, int ncol, int nlay, const void* up, const void* dn, const void* plev, void* hr, void* stream)
{
    return guarded([&] {
        check_dims(ncol, nlay);
        if (!up || !dn || !plev || !hr) throw ArgError("heating_rate: null array");
        DISPATCH(dtype, launch_heating_rate<float>(ncol, nlay, up, dn, plev, hr, (cudaStream_t)stream),
                        launch_heating_rate<double>(ncol, nlay, up, dn, plev, hr, (cudaStream_t)stream));
    });
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// Fused column solvers
// ---------------------------------------------------------------------------------------------
namespace {

// One fused call (= one solve_gpu of the reference driver) as an object, so that the longwave and the shortwave call of a
// step can also be advanced tile by tile TOGETHER (fused_pair below).  setup() = validation, workspace, zeroing;
// prologue(c0) = the tile's interpolation state and by-band particulate optics; bands() = the band launches of the tile;
// finish() = net flux and heating rate.
template<typename Real>
struct FusedJob
{
    rrtmgp_b200_gas_optics* g = nullptr; const rrtmgp_b200_columns* cols = nullptr; const rrtmgp_b200_fluxes* out = nullptr;
    bool sw = false, cloud = false, aerosol = false, bnd = false, colres = false;
    int ncol = 0, nlay = 0, nlev = 0, nang = 1, top_at_1 = 0, cpb = 32, wave = 1, nslots = 0, tile = 0;
    AerosolMmr mmr{}; VmrArgs vmr{};
    TileState<Real> st{}; FusedArgs<Real> a{};
    const Real* col_dry = nullptr; Real *ctau = nullptr, *cssa = nullptr, *cg = nullptr;
    size_t ncv = 0;

    // everything before the tile loop.  forced_tile > 0: use that tile size (the paired driver gives both calls the same)
    void setup(rrtmgp_b200_gas_optics* g_, const rrtmgp_b200_columns* cols_, const rrtmgp_b200_fluxes* out_, cudaStream_t s, bool sw_,
               int forced_tile = 0, int forced_wave = 0)
    {
        g = g_; cols = cols_; out = out_; sw = sw_;
        if (!g || !cols || !out) throw ArgError("fused fluxes: null argument");
        const HostPlan& h = g->h;
        if (h.is_lw == sw) throw ArgError(sw ? "sw_fluxes needs a shortwave k-distribution" : "lw_fluxes needs a longwave k-distribution");
        ncol = cols->ncol; nlay = cols->nlay; nlev = nlay + 1;
        check_dims(ncol, nlay);
        if (!cols->d_play || !cols->d_plev || !cols->d_tlay || !out->d_flux_up || !out->d_flux_dn) throw ArgError("fused fluxes: null array");
        if (!sw && (!cols->d_tlev || !cols->d_tsfc || !cols->d_sfc_emis)) throw ArgError("lw_fluxes: tlev, tsfc and sfc_emis are required");
        if (sw && (!cols->d_mu0 || !cols->d_sfc_alb_dir || !cols->d_sfc_alb_dif)) throw ArgError("sw_fluxes: mu0 and surface albedos are required");
        nang = sw ? 1 : std::max(cols->n_gauss_angles, 1);
        if (nang > 4) throw ArgError("lw_fluxes: n_gauss_angles must be 1..4");   // src/Rte_lw.cpp:139-151
        cloud = cols->cloud != nullptr;
        if (cloud)
        {
            if (cols->cloud->dtype != g->dtype) throw ArgError("cloud optics object has a different precision");
            if (cols->cloud->cf.nbnd != h.nbnd) throw ArgError("Cannot add optical properties with incompatible band - gpoint combination");
            if (!cols->d_lwp || !cols->d_iwp || !cols->d_rel || !cols->d_dei) throw ArgError("cloud inputs missing");
        }
        aerosol = cols->aerosol != nullptr;
        if (aerosol)
        {
            if (!sw) throw ArgError("lw_fluxes: aerosol optics are shortwave only (Radiation_solver_longwave has none)");
            if (cols->aerosol->dtype != g->dtype) throw ArgError("aerosol optics object has a different precision");
            if (cols->aerosol->af.nbnd != h.nbnd) throw ArgError("Cannot add optical properties with incompatible band - gpoint combination");
            if (!cols->d_rh) throw ArgError("aerosol inputs missing: rh");
            mmr = make_mmr(&cols->aerosol_concs, ncol, nlay);
        }
        bnd = out->d_bnd_flux_up != nullptr;
        if (bnd && !out->d_bnd_flux_dn) throw ArgError("by-band output needs both bnd_flux_up and bnd_flux_dn");
        vmr = make_vmr(g, &cols->concs, ncol, nlay);

        top_at_1 = cols->top_at_1;
        if (top_at_1 < 0)
        {
            Real p2[2];
            CK(cudaMemcpyAsync(&p2[0], cols->d_play, sizeof(Real), cudaMemcpyDeviceToHost, s));
            CK(cudaMemcpyAsync(&p2[1], (const Real*)cols->d_play + (size_t)ncol * (nlay - 1), sizeof(Real), cudaMemcpyDeviceToHost, s));
            CK(cudaStreamSynchronize(s));
            top_at_1 = p2[0] < p2[1];
        }

        colres = wants_colres(g, cols, sw);
        cpb = FUSED_THREADS / (h.gw / 4);
        // strips path: a tile is several waves of CTAs (CTAs retire and start continuously, so compute-heavy
        // and memory-heavy passes of different CTAs overlap); scratch strips are per resident CTA.
        // column-resident path: a tile only bounds the interpolation state (one CTA per 16/gw columns).
        wave = forced_wave > 0 ? forced_wave : std::max(sw ? sw_blocks_per_wave<Real>(h.gw) : lw_blocks_per_wave<Real>(h.gw), 1);
        nslots = colres ? 0 : 2 * wave;
        tile = forced_tile > 0 ? forced_tile : auto_tile(g, ncol, nlay, wave, cpb);
        const size_t tcl = (size_t)tile * nlay, ncl = (size_t)ncol * nlay;
        ncv = (size_t)ncol * nlev;
        // longwave with several Gauss angles: the strip holds {trans, source_up} per angle (fused_lw_ma.cu)
        const int words = sw ? 6 : 2 * (nang > 1 && !g_angle_launches && h.gw == 16 && h.neta == 9 ? nang : 1);
        size_t need = state_bytes<Real>(h, tile, nlay) + pad256(fused_scratch_bytes<Real>(nslots, nlay, words)) + pad256(nslots * sizeof(int)) + 4096;
        if (!cols->d_col_dry) need += pad256(ncl * sizeof(Real));
        if (cloud || aerosol) need += 3 * pad256(tcl * h.nbnd * sizeof(Real));
        if (sw && !out->d_flux_dn_dir) need += pad256(ncv * sizeof(Real));
        if (sw && bnd && !out->d_bnd_flux_dn_dir) need += pad256(ncv * h.nbnd * sizeof(Real));
        // a small column-resident call spreads its bands over CTAs: per-band partial sums, added in band order afterwards
        const int nsplit = colres ? colres_band_split(tile, h.gw, h.nbnd) : 1;
        if (nsplit > 1) need += pad256((size_t)nsplit * 3 * ncv * sizeof(Real));
        g->ws.reserve(need, s);
        st = carve_state<Real>(g->ws, h, tile, nlay);

        a = FusedArgs<Real>{};
        a.ncol = ncol; a.nlay = nlay; a.top_at_1 = top_at_1;
        a.tlay = (const Real*)cols->d_tlay; a.tlev = (const Real*)cols->d_tlev; a.tsfc = (const Real*)cols->d_tsfc;
        a.sfc_emis = (const Real*)cols->d_sfc_emis; a.mu0 = (const Real*)cols->d_mu0;
        a.alb_dir = (const Real*)cols->d_sfc_alb_dir; a.alb_dif = (const Real*)cols->d_sfc_alb_dif;
        a.tsi_scaling = (const Real*)cols->d_tsi_scaling;
        a.scratch = (V4<Real>*)g->ws.take<char>(fused_scratch_bytes<Real>(nslots, nlay, words));
        a.slot_busy = g->ws.take<int>(nslots); a.nslots = nslots;
        if (nslots) CK(cudaMemsetAsync(a.slot_busy, 0, nslots * sizeof(int), s));
        a.flux_up = (Real*)out->d_flux_up; a.flux_dn = (Real*)out->d_flux_dn;
        a.flux_dir = (Real*)out->d_flux_dn_dir;
        a.bnd_up = (Real*)out->d_bnd_flux_up; a.bnd_dn = (Real*)out->d_bnd_flux_dn; a.bnd_dir = (Real*)out->d_bnd_flux_dn_dir;
        a.secant = (Real)(1. / 0.6096748751); a.weight = (Real)1;      // src/Rte_lw.cpp:139-151, n = 1
        col_dry = (const Real*)cols->d_col_dry;
        if (!col_dry)
        {
            const int ih = h.idx_h2o - 1;
            if (!(cols->concs.vmr_ncol[ih] == ncol && cols->concs.vmr_nlay[ih] == nlay))
                throw ArgError("fused fluxes: computing col_dry needs an (ncol,nlay) h2o field");
            Real* t = g->ws.take<Real>(ncl);
            launch_col_dry<Real>(ncol, nlay, cols->concs.d_vmr[ih], cols->d_plev, t, s);
            col_dry = t;
        }
        if (cloud || aerosol) { ctau = g->ws.take<Real>(tcl * h.nbnd); cssa = g->ws.take<Real>(tcl * h.nbnd); cg = g->ws.take<Real>(tcl * h.nbnd); }
        if (sw && !a.flux_dir) a.flux_dir = g->ws.take<Real>(ncv);
        if (sw && bnd && !a.bnd_dir) a.bnd_dir = g->ws.take<Real>(ncv * h.nbnd);
        if (nsplit > 1) { a.part = g->ws.take<Real>((size_t)nsplit * 3 * ncv); a.nsplit = nsplit; }

        if (!colres)        // the strips kernels accumulate band by band; the column-resident ones store once
        {
            CK(cudaMemsetAsync(a.flux_up, 0, ncv * sizeof(Real), s));
            CK(cudaMemsetAsync(a.flux_dn, 0, ncv * sizeof(Real), s));
            if (sw) CK(cudaMemsetAsync(a.flux_dir, 0, ncv * sizeof(Real), s));
        }
        g->ws.arm(s);
    }

    // path: the handle's setting (0 strips, 1 column-resident, 2 by size), else RB200_PATH, else by size
    static bool wants_colres(const rrtmgp_b200_gas_optics* g, const rrtmgp_b200_columns* cols, bool sw)
    {
        const int want = g->colres >= 0 ? (g->colres == 2 ? -1 : g->colres) : g_colres;
        const bool one_angle = sw || cols->n_gauss_angles <= 1;          // column-resident LW: one angle
        const bool colres_ok = one_angle && colres_smem_bytes<Real>(cols->nlay, g->h.gw, sw) <= 227 * 1024;
        if (want == 1 && one_angle && !colres_ok) throw ArgError("fused fluxes: column too tall for the shared-memory resident path");
        return want == 1 ? colres_ok : (want < 0 && colres_ok && cols->ncol <= kColresAutoMaxCols);
    }

    static int auto_tile(rrtmgp_b200_gas_optics* g, int ncol, int nlay, int wave, int cpb)
    {
        int tile = g->tile_cols;
        if (tile <= 0)
        {
            // 8 waves per tile; 16 when the interpolation state of such a tile takes under a quarter of the
            // free device memory (longer launches: fewer tails, fewer prologue launches)
            tile = 8 * wave * cpb;
            size_t free_b = 0, total_b = 0;
            if (ncol > tile &&          // a call that fits one tile needs no memory query (it costs ~50 us per call)
                cudaMemGetInfo(&free_b, &total_b) == cudaSuccess &&
                state_bytes<Real>(g->h, 2 * tile, nlay) < (free_b + g->ws.cap) / 4) tile *= 2;
        }
        tile = std::min(tile, ncol);
        // equal tiles (rounded up to whole CTAs) instead of full tiles plus a remainder
        const int ntiles = (ncol + tile - 1) / tile;
        tile = (((ncol + ntiles - 1) / ntiles + cpb - 1) / cpb) * cpb;
        return std::min(tile, ncol);
    }

    void prologue(int c0, cudaStream_t s)
    {
        const int nc = std::min(tile, ncol - c0);
        st.nc = nc; a.c0 = c0; a.nc = nc;
        CK(cudaMemsetAsync(st.tmask, 0, (size_t)nc * st.tmw * sizeof(unsigned), s));
        launch_interp<Real>(dplan<Real>(g), vmr, ncol, c0, nc, nlay, cols->d_play, cols->d_tlay, col_dry, cols->d_tlev, st,
                            nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, s);
        if (cloud)
        {
            launch_cloud_optics<Real>(dcloud<Real>(cols->cloud), sw, ncol, c0, nc, nlay, cols->d_lwp, cols->d_iwp,
                                      cols->d_rel, cols->d_dei, nc, 0, sw && cols->delta_cloud, ctau, cssa, cg, s);
            a.cld_tau = ctau; a.cld_ssa = cssa; a.cld_g = cg;
        }
        if (aerosol)
        {
            // the tile's by-band particulate properties = cloud (+) aerosol (increment is associative: tau, tau*ssa and
            // tau*ssa*g add), so the band kernels still apply ONE by-band increment to the gas
            launch_aerosol_optics<Real>(daerosol<Real>(cols->aerosol), mmr, cloud, ncol, c0, nc, nlay, cols->d_rh, cols->d_plev,
                                        nc, 0, cols->delta_aerosol, ctau, cssa, cg, s);
            a.cld_tau = ctau; a.cld_ssa = cssa; a.cld_g = cg;
        }
    }

    // Gauss-Jacobi-5 quadrature (src/Rte_lw.cpp:139-151): secant and weight of angle ia of nang
    void set_angle(int ia)
    {
        static const double gauss_Ds[4][4] = {{1. / 0.6096748751, 0., 0., 0.}, {1. / 0.2509907356, 1. / 0.7908473988, 0., 0.},
            {1. / 0.1024922169, 1. / 0.4417960320, 1. / 0.8633751621, 0.}, {1. / 0.0454586727, 1. / 0.2322334416, 1. / 0.5740198775, 1. / 0.903077597}};
        static const double gauss_wts[4][4] = {{1., 0., 0., 0.}, {0.2300253764, 0.7699746236, 0., 0.},
            {0.0437820218, 0.3875796738, 0.5686383044, 0.}, {0.0092068785, 0.1285704278, 0.4323381850, 0.4298845087}};
        a.secant = (Real)gauss_Ds[nang - 1][ia]; a.weight = (Real)gauss_wts[nang - 1][ia]; a.bnd_add = ia > 0;
        for (int k = 0; k < 4; ++k) { a.secants[k] = (Real)gauss_Ds[nang - 1][k]; a.weights[k] = (Real)gauss_wts[nang - 1][k]; }
    }

    void bands(cudaStream_t s)
    {
        const HostPlan& h = g->h;
        int nlaunch = 0;
        if (g_prof.on) { cudaEvent_t e; CK(cudaEventCreate(&e)); CK(cudaEventRecord(e, s)); g_prof.ev[sw].push_back(e); }
        if (colres) launch_cols<Real>(sw, dplan<Real>(g), (const BandArgs*)g->bands_dev.p, st, a, s);
        else if (sw)
            for (int b = 0; b < h.nbnd; ++b) launch_sw_band<Real>(dplan<Real>(g), g->bands[b], st, a, s);
        else
            // several Gauss angles (the reference's GPU path has a single one; its CPU path sums them, src/Rte_lw.cpp:139-151):
            // one launch per band with the angles inside the layer loop where that kernel exists, else one pass of the
            // single-angle kernel per angle; fluxes accumulate either way
            for (int b = 0; b < h.nbnd; ++b)
            {
                if (nang > 1 && !g_angle_launches && lw_multi_angle_supported<Real>(dplan<Real>(g), g->bands[b]))
                {
                    set_angle(0);
                    launch_lw_band_angles<Real>(dplan<Real>(g), g->bands[b], st, a, nang, s);
                    ++nlaunch;
                    continue;
                }
                for (int ia = 0; ia < nang; ++ia)
                {
                    set_angle(ia);
                    launch_lw_band<Real>(dplan<Real>(g), g->bands[b], st, a, s);
                    ++nlaunch;
                }
            }
        if (g_prof.on)
        {
            cudaEvent_t e; CK(cudaEventCreate(&e)); CK(cudaEventRecord(e, s)); g_prof.ev[sw].push_back(e);
            g_prof.launches[sw] += colres ? 1 : (sw ? h.nbnd : nlaunch);
        }
    }

    void finish(cudaStream_t s)
    {
        const HostPlan& h = g->h;
        if (out->d_flux_net) launch_net<Real>(ncv, a.flux_dn, a.flux_up, out->d_flux_net, s);
        if (bnd && out->d_bnd_flux_net) launch_net<Real>(ncv * h.nbnd, a.bnd_dn, a.bnd_up, out->d_bnd_flux_net, s);
        if (out->d_heating_rate) launch_heating_rate<Real>(ncol, nlay, a.flux_up, a.flux_dn, cols->d_plev, out->d_heating_rate, s);
        g->ws.check(s, sw ? "sw_fluxes" : "lw_fluxes");
    }
};

template<typename Real>
void fused_fluxes(rrtmgp_b200_gas_optics* g, const rrtmgp_b200_columns* cols, const rrtmgp_b200_fluxes* out,
                  cudaStream_t s, bool sw)
{
    if (!g) throw ArgError("fused fluxes: null handle");
    HandleCall hc(g, s);
    FusedJob<Real> j;
    j.setup(g, cols, out, s, sw);
    L2Pin pin(s, g->kmajor.p, g->kmajor.bytes);
    for (int c0 = 0; c0 < j.ncol; c0 += j.tile)
    {
        j.prologue(c0, s);
        j.bands(s);
    }
    j.finish(s);
}

// ---- host-buffer entry: chunks of columns stream H2D -> fused LW/SW -> D2H, double-buffered on
// three streams so the PCIe copies hide behind the band kernels of the neighbouring chunks.
struct HostPipe
{
    Arena in[2], out[2];
    cudaStream_t h2d = nullptr, d2h = nullptr, side = nullptr;
    cudaEvent_t e_h2d[2]{}, e_comp[2]{}, e_d2h[2]{}, e_side[2]{};
    HostPipe()
    {
        CK(cudaStreamCreateWithFlags(&h2d, cudaStreamNonBlocking)); CK(cudaStreamCreateWithFlags(&d2h, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&side, cudaStreamNonBlocking));
        for (int i = 0; i < 2; ++i)
        {
            CK(cudaEventCreateWithFlags(&e_h2d[i], cudaEventDisableTiming)); CK(cudaEventCreateWithFlags(&e_comp[i], cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&e_d2h[i], cudaEventDisableTiming)); CK(cudaEventCreateWithFlags(&e_side[i], cudaEventDisableTiming));
        }
    }
    ~HostPipe()
    {
        for (int i = 0; i < 2; ++i) { cudaEventDestroy(e_h2d[i]); cudaEventDestroy(e_comp[i]); cudaEventDestroy(e_d2h[i]); cudaEventDestroy(e_side[i]); }
        cudaStreamDestroy(h2d); cudaStreamDestroy(d2h); cudaStreamDestroy(side);
    }
};
HostPipe& host_pipe(rrtmgp_b200_gas_optics* g)
{
    if (!g->pipe) g->pipe = std::make_shared<HostPipe>();
    return *static_cast<HostPipe*>(g->pipe.get());
}

// Longwave and shortwave of the same columns in ONE call: per tile both prologues, then band i of the longwave and band i of
// the shortwave k-distribution in one launch whose CTAs alternate between the two roles (fused_pair.cu), so the L1-bound
// longwave CTAs and the DRAM-bound shortwave CTAs share every SM; bands without a partner, and every band the paired
// kernel does not cover (narrow bands of the reduced g-point sets, the column-resident path), run through
// the single-role launches.  Results are bit-identical to rrtmgp_b200_lw_fluxes + rrtmgp_b200_sw_fluxes.
template<typename Real>
void fused_pair(rrtmgp_b200_gas_optics* glw, const rrtmgp_b200_columns* clw, const rrtmgp_b200_fluxes* olw,
                rrtmgp_b200_gas_optics* gsw, const rrtmgp_b200_columns* csw, const rrtmgp_b200_fluxes* osw, cudaStream_t s)
{
    if (!glw || !gsw || !clw || !csw || !olw || !osw) throw ArgError("lwsw_fluxes: null argument");
    if (glw == gsw) throw ArgError("lwsw_fluxes: the longwave and the shortwave call need their own k-distribution objects");
    if (clw->ncol != csw->ncol || clw->nlay != csw->nlay) throw ArgError("lwsw_fluxes: the two calls must cover the same columns");
    HandleCall hl(glw, s), hs(gsw, s);
    // the two calls side by side: the shortwave one on a side stream, forked from and joined back into the caller's stream
    auto two_streams = [&](bool prof)
    {
        HostPipe& pp = host_pipe(gsw);
        // profile (large calls only: two more events would show in a 40 us call): the band times of the two calls overlap, so
        // the wall time from fork to join goes to the third counter
        prof = prof && g_prof.on;
        if (prof) { cudaEvent_t e; CK(cudaEventCreate(&e)); CK(cudaEventRecord(e, s)); g_prof.ev[2].push_back(e); }
        CK(cudaEventRecord(pp.e_side[0], s)); CK(cudaStreamWaitEvent(pp.side, pp.e_side[0], 0));
        fused_fluxes<Real>(gsw, csw, osw, pp.side, true);
        fused_fluxes<Real>(glw, clw, olw, s, false);
        CK(cudaEventRecord(pp.e_side[1], pp.side)); CK(cudaStreamWaitEvent(s, pp.e_side[1], 0));
        if (prof) { cudaEvent_t e; CK(cudaEventCreate(&e)); CK(cudaEventRecord(e, s)); g_prof.ev[2].push_back(e); g_prof.launches[2] += 1; }
    };
    if (FusedJob<Real>::wants_colres(glw, clw, false) && FusedJob<Real>::wants_colres(gsw, csw, true))
    {
        two_streams(false);     // two small calls (column-resident kernels): neither fills the SMs
        return;
    }
    // paired launches need both roles with the same particulate flag (one kernel instantiation per flag), one longwave angle
    // (several run in their own kernels, fused_lw_ma.cu) and the per-band path on both sides; the fp64 build measured slower
    // paired (pair_supported refuses it).  Decided before any set-up work.
    const bool pairable = sizeof(Real) == 4 && !FusedJob<Real>::wants_colres(glw, clw, false) && !FusedJob<Real>::wants_colres(gsw, csw, true) &&
                          glw->h.gw == 16 && gsw->h.gw == 16 && clw->n_gauss_angles <= 1 &&
                          ((clw->cloud != nullptr) == (csw->cloud != nullptr || csw->aerosol != nullptr));
    if (!pairable)
    {
        // nothing to pair (each call with its own tiling): on two streams the launches of one fill the tails of the other's
        // (what the host pipeline has always done for the fp64 build: +5 %); RB200_HOST_CONCURRENT=0: one after the other
        if (g_host_concurrent) two_streams(true);
        else
        {
            fused_fluxes<Real>(glw, clw, olw, s, false);
            fused_fluxes<Real>(gsw, csw, osw, s, true);
        }
        return;
    }
    const int wave = std::max(pair_blocks_per_wave<Real>(), 1);
    // one tile size for both: the smaller of the two automatic choices
    const int cpb = FUSED_THREADS / 4;
    const int tile = std::min(FusedJob<Real>::auto_tile(glw, clw->ncol, clw->nlay, wave, cpb), FusedJob<Real>::auto_tile(gsw, csw->ncol, csw->nlay, wave, cpb));
    FusedJob<Real> L, S;
    L.setup(glw, clw, olw, s, false, tile, wave);
    S.setup(gsw, csw, osw, s, true, tile, wave);
    L2Pin pin(s, gsw->kmajor.p, gsw->kmajor.bytes);
    const int nbl = glw->h.nbnd, nbs = gsw->h.nbnd;
    for (int c0 = 0; c0 < L.ncol; c0 += tile)
    {
        L.prologue(c0, s);
        S.prologue(c0, s);
        if (g_prof.on) { cudaEvent_t e; CK(cudaEventCreate(&e)); CK(cudaEventRecord(e, s)); g_prof.ev[2].push_back(e); }
        L.set_angle(0);
        int il = 0, is = 0, npair = 0, nsingle = 0;
        while (il < nbl || is < nbs)
        {
            if (il < nbl && is < nbs && pair_supported<Real>(dplan<Real>(glw), glw->bands[il], dplan<Real>(gsw), gsw->bands[is]))
            {
                launch_lwsw_band<Real>(dplan<Real>(glw), glw->bands[il], L.st, L.a, dplan<Real>(gsw), gsw->bands[is], S.st, S.a, s);
                ++il; ++is; ++npair;
            }
            // no pair: the band the paired kernel does not cover (or the only kind left) goes through its own kernel
            else if (il < nbl && (is >= nbs || glw->bands[il].ngb <= 8))
            { launch_lw_band<Real>(dplan<Real>(glw), glw->bands[il], L.st, L.a, s); ++il; ++nsingle; }
            else
            { launch_sw_band<Real>(dplan<Real>(gsw), gsw->bands[is], S.st, S.a, s); ++is; ++nsingle; }
        }
        if (g_prof.on)
        {
            cudaEvent_t e; CK(cudaEventCreate(&e)); CK(cudaEventRecord(e, s)); g_prof.ev[2].push_back(e);
            g_prof.launches[2] += npair + nsingle;
        }
    }
    L.finish(s);
    S.finish(s);
}


// The schedule of column chunks of the host pipeline: (first column, columns), contiguous, covering [0, ncol).
// wave_cols == 0: equal chunks of `chunk` columns (an explicit chunk or tile size).  wave_cols = W > 0: a ramp.  The pipeline's
// exposed time is the upload of the FIRST chunk and the download of the LAST one (the rest hides behind the kernels as long as a
// chunk's copies are shorter than its neighbour's kernels), and every band launch of a chunk ends in a partly empty wave, so
// chunks should be large.  Hence 1, 2, 4 waves of CTAs per role at both ends and equal chunks of up to 8 waves between (copies of
// a chunk take ~0.3 of its kernels' time, so a chunk twice its neighbour still hides).  Measured: two equal chunks exposed 18 ms
// of an 87 ms step at 8 GPUs (131 072 columns per rank); equal 4-wave chunks cost 6 % in tails at 1 GPU.
static std::vector<std::pair<int, int>> chunk_schedule(int ncol, int wave_cols, int chunk)
{
    std::vector<std::pair<int, int>> sched;
    std::vector<int> left, right;
    int remaining = ncol;
    if (wave_cols > 0)
    {
        const int W = wave_cols;
        chunk = 8 * W;
        for (int size = W; size < chunk && remaining > 3 * size; size *= 2) { left.push_back(size); right.push_back(size); remaining -= 2 * size; }
        // a call of 1.5 to 3 waves still gets a one-wave first chunk: one chunk would expose its whole upload and download, two
        // hide the larger one's upload and the smaller one's download
        if (left.empty() && 2 * remaining > 3 * W) { left.push_back(W); remaining -= W; }
    }
    chunk = std::max(std::min(chunk, remaining), 1);
    const int nmid = std::max((remaining + chunk - 1) / chunk, 1);
    const int mid = std::min((((remaining + nmid - 1) / nmid + 31) / 32) * 32, remaining);   // equal chunks, whole CTAs
    int c0 = 0;
    for (int n : left) { sched.push_back({c0, n}); c0 += n; }
    const int end = c0 + remaining;
    for (; c0 < end; c0 += mid) sched.push_back({c0, std::min(mid, end - c0)});
    c0 = end;                      // the last equal chunk may be short: the loop variable overshoots
    for (auto it = right.rbegin(); it != right.rend(); ++it) { sched.push_back({c0, *it}); c0 += *it; }
    return sched;
}

// one host array of the column set. kind 0: (ncol, n2) column-major, chunked by rows of the first index;
// kind 1: (n2, ncol) band-major, a chunk is contiguous; kind 2: not column-dependent, uploaded whole
struct HostItem { const void* h; int kind; size_t n2; void* d[2]; };

template<typename Real>
void fluxes_host(rrtmgp_b200_gas_optics* glw, const rrtmgp_b200_columns* clw, const rrtmgp_b200_fluxes* olw,
                 rrtmgp_b200_gas_optics* gsw, const rrtmgp_b200_columns* csw, const rrtmgp_b200_fluxes* osw,
                 int chunk_cols, cudaStream_t s)
{
    rrtmgp_b200_gas_optics* g0 = glw ? glw : gsw;
    const rrtmgp_b200_columns* c0p = glw ? clw : csw;
    if (!g0 || !c0p) throw ArgError("fluxes_host: need a longwave and/or a shortwave solver with its columns");
    if (glw && (!clw || !olw)) throw ArgError("fluxes_host: longwave columns/outputs missing");
    if (gsw && (!csw || !osw)) throw ArgError("fluxes_host: shortwave columns/outputs missing");
    if (glw && gsw && (clw->ncol != csw->ncol || clw->nlay != csw->nlay)) throw ArgError("fluxes_host: LW and SW column sets differ in shape");
    if (g0->host_only || (gsw && gsw->host_only)) throw CudaError("this handle was created HOST_ONLY: no device tables, and there is no CPU fallback");
    const int ncol = c0p->ncol, nlay = c0p->nlay, nlev = nlay + 1;
    check_dims(ncol, nlay);
    if (!c0p->d_play) throw ArgError("fluxes_host: play missing");
    HandleCall hp(g0, s);            // the pipeline's buffers live in g0's HostPipe; the fused calls below guard their own handles
    HostPipe& P = host_pipe(g0);
    int top_at_1 = c0p->top_at_1;
    if (top_at_1 < 0) { const Real* pl = (const Real*)c0p->d_play; top_at_1 = pl[0] < pl[(size_t)ncol * (nlay - 1)]; }

    // chunk = one default tile of the wider solver, so every chunk fills whole waves
    int chunk = chunk_cols;
    if (chunk <= 0)
        for (auto* g : {glw, gsw})
            if (g)
            {
                const int cpb = FUSED_THREADS / (g->h.gw / 4);
                const int wave = std::max(g->h.is_lw ? lw_blocks_per_wave<Real>(g->h.gw) : sw_blocks_per_wave<Real>(g->h.gw), 1);
                chunk = std::max(chunk, g->tile_cols > 0 ? g->tile_cols : 4 * wave * cpb);   // finer than a tile: shorter pipeline ramps
            }
    chunk = std::min(chunk, ncol);
    const bool ramp = chunk_cols <= 0 && !(glw && glw->tile_cols > 0) && !(gsw && gsw->tile_cols > 0);   // a set tile size is the chunk size
    int W = 0;
    if (ramp)
    {
        rrtmgp_b200_gas_optics* g = glw ? glw : gsw;
        const int cpb = FUSED_THREADS / (g->h.gw / 4);
        W = std::max(g->h.is_lw ? lw_blocks_per_wave<Real>(g->h.gw) : sw_blocks_per_wave<Real>(g->h.gw), 1) * cpb;
    }
    const std::vector<std::pair<int, int>> sched = chunk_schedule(ncol, ramp ? W : 0, chunk);      // (first column, columns)
    chunk = 0;
    for (auto& sc : sched) chunk = std::max(chunk, sc.second);

    // inputs, de-duplicated by host address (LW and SW normally share play/plev/tlay/vmr/col_dry)
    std::vector<HostItem> items;
    auto item = [&](const void* h, int kind, size_t n2) -> int {
        if (!h) return -1;
        for (size_t i = 0; i < items.size(); ++i) if (items[i].h == h) return (int)i;
        items.push_back({h, kind, n2, {nullptr, nullptr}});
        return (int)items.size() - 1;
    };
    struct Side
    {
        rrtmgp_b200_gas_optics* g; const rrtmgp_b200_columns* c; const rrtmgp_b200_fluxes* o; bool sw;
        int play, plev, tlay, tlev, tsfc, coldry, emis, adir, adif, mu0, tsi, lwp, iwp, rel, dei, rh, aer[11]; std::vector<int> vmr;
        std::vector<const void*> dv; std::vector<int> vnc, vnl;
    };
    std::vector<Side> sides;
    for (int k = 0; k < 2; ++k)
    {
        rrtmgp_b200_gas_optics* g = k ? gsw : glw; const rrtmgp_b200_columns* c = k ? csw : clw;
        if (!g) continue;
        if (g->dtype != g0->dtype) throw ArgError("fluxes_host: LW and SW solvers differ in precision");
        if (g->h.is_lw == (k == 1)) throw ArgError("fluxes_host: solver passed on the wrong side (LW first, SW second)");
        Side sd{}; sd.g = g; sd.c = c; sd.o = k ? osw : olw; sd.sw = k == 1;
        sd.play = item(c->d_play, 0, nlay); sd.plev = item(c->d_plev, 0, nlev); sd.tlay = item(c->d_tlay, 0, nlay);
        sd.tlev = k ? -1 : item(c->d_tlev, 0, nlev); sd.tsfc = k ? -1 : item(c->d_tsfc, 0, 1);
        sd.coldry = item(c->d_col_dry, 0, nlay);
        sd.emis = k ? -1 : item(c->d_sfc_emis, 1, g->h.nbnd);
        sd.adir = k ? item(c->d_sfc_alb_dir, 1, g->h.nbnd) : -1; sd.adif = k ? item(c->d_sfc_alb_dif, 1, g->h.nbnd) : -1;
        sd.mu0 = k ? item(c->d_mu0, 0, 1) : -1; sd.tsi = k ? item(c->d_tsi_scaling, 0, 1) : -1;
        const bool cl = c->cloud != nullptr;
        sd.lwp = cl ? item(c->d_lwp, 0, nlay) : -1; sd.iwp = cl ? item(c->d_iwp, 0, nlay) : -1;
        sd.rel = cl ? item(c->d_rel, 0, nlay) : -1; sd.dei = cl ? item(c->d_dei, 0, nlay) : -1;
        const bool ae = k == 1 && c->aerosol != nullptr;
        sd.rh = ae ? item(c->d_rh, 0, nlay) : -1;
        for (int i = 0; i < 11; ++i)
        {
            sd.aer[i] = -1;
            if (!ae) continue;
            const rrtmgp_b200_aerosol_concs& ac = c->aerosol_concs;
            if (!ac.d_mmr[i]) throw MissingError("Aerosol_concs: aermr" + std::string(i < 9 ? "0" : "") + std::to_string(i + 1) + " not set");
            const bool field = ac.mmr_ncol[i] == ncol && ncol > 1;
            sd.aer[i] = item(ac.d_mmr[i], field ? 0 : 2, field ? (size_t)ac.mmr_nlay[i] : (size_t)ac.mmr_ncol[i] * ac.mmr_nlay[i]);
        }
        const rrtmgp_b200_gas_concs& gc = c->concs;
        if (gc.ngas != g->h.ngas || !gc.d_vmr || !gc.vmr_ncol || !gc.vmr_nlay)
            throw ArgError("gas concentrations: need one entry per gas of the k-distribution object");
        for (int i = 0; i < gc.ngas; ++i)
        {
            if (!gc.d_vmr[i]) throw MissingError("Gas_concs: gas " + g->h.gas_names[i] + " not set");
            const bool field = gc.vmr_ncol[i] == ncol && ncol > 1;
            sd.vmr.push_back(item(gc.d_vmr[i], field ? 0 : 2, field ? (size_t)gc.vmr_nlay[i] : (size_t)gc.vmr_ncol[i] * gc.vmr_nlay[i]));
        }
        sd.dv.resize(gc.ngas); sd.vnc.assign(gc.vmr_ncol, gc.vmr_ncol + gc.ngas); sd.vnl.assign(gc.vmr_nlay, gc.vmr_nlay + gc.ngas);
        sides.push_back(std::move(sd));
    }
    // outputs: (ncol, rows) host arrays; device chunk arrays (chunk, rows)
    struct OutItem { void* h; size_t rows; void* d[2]; };
    std::vector<OutItem> outs;
    auto oitem = [&](void* h, size_t rows) -> int { if (!h) return -1; outs.push_back({h, rows, {nullptr, nullptr}}); return (int)outs.size() - 1; };
    struct SideOut { int up, dn, net, dir, bup, bdn, bnet, bdir, hr; };
    std::vector<SideOut> so;
    for (auto& sd : sides)
    {
        const rrtmgp_b200_fluxes* o = sd.o; const size_t nb = sd.g->h.nbnd;
        if (!o->d_flux_up || !o->d_flux_dn) throw ArgError("fluxes_host: flux_up and flux_dn are required");
        so.push_back({oitem(o->d_flux_up, nlev), oitem(o->d_flux_dn, nlev), oitem(o->d_flux_net, nlev), sd.sw ? oitem(o->d_flux_dn_dir, nlev) : -1,
                      oitem(o->d_bnd_flux_up, nlev * nb), oitem(o->d_bnd_flux_dn, nlev * nb), oitem(o->d_bnd_flux_net, nlev * nb),
                      sd.sw ? oitem(o->d_bnd_flux_dn_dir, nlev * nb) : -1, oitem(o->d_heating_rate, nlay)});
    }
    size_t in_bytes = 4096, out_bytes = 4096;
    for (auto& it : items) in_bytes += pad256((it.kind == 2 ? it.n2 : (size_t)chunk * it.n2) * sizeof(Real)) + 256;
    for (auto& ot : outs) out_bytes += pad256((size_t)chunk * ot.rows * sizeof(Real)) + 256;
    for (int b = 0; b < 2; ++b)
    {
        P.in[b].reserve(in_bytes, s); P.out[b].reserve(out_bytes, s);
        for (auto& it : items) it.d[b] = P.in[b].take<Real>(it.kind == 2 ? it.n2 : (size_t)chunk * it.n2);
        for (auto& ot : outs) ot.d[b] = P.out[b].take<Real>((size_t)chunk * ot.rows);
    }
    for (auto& it : items)          // column-independent inputs once, into both buffers
        if (it.kind == 2)
            for (int b = 0; b < 2; ++b) CK(cudaMemcpyAsync(it.d[b], it.h, it.n2 * sizeof(Real), cudaMemcpyHostToDevice, P.h2d));

    // both sides present, fp32: each chunk runs through the paired driver (longwave and shortwave CTAs co-resident, one
    // stream); otherwise longwave on the caller's stream and shortwave on a side stream
    const bool paired = sides.size() == 2 && sizeof(Real) == 4 && g_host_concurrent;
    const bool concurrent = sides.size() == 2 && g_host_concurrent && !paired;
    for (int i = 0; i < (int)sched.size(); ++i)
    {
        const int c0 = sched[i].first, nc = sched[i].second, b = i & 1;
        // H2D of chunk i once the kernels of chunk i-2 have drained this buffer
        if (i >= 2) CK(cudaStreamWaitEvent(P.h2d, P.e_comp[b], 0));
        for (auto& it : items)
        {
            if (it.kind == 0)
                CK(cudaMemcpy2DAsync(it.d[b], (size_t)nc * sizeof(Real), (const Real*)it.h + c0, (size_t)ncol * sizeof(Real),
                                     (size_t)nc * sizeof(Real), it.n2, cudaMemcpyHostToDevice, P.h2d));
            else if (it.kind == 1)
                CK(cudaMemcpyAsync(it.d[b], (const Real*)it.h + (size_t)c0 * it.n2, (size_t)nc * it.n2 * sizeof(Real), cudaMemcpyHostToDevice, P.h2d));
        }
        CK(cudaEventRecord(P.e_h2d[b], P.h2d));
        CK(cudaStreamWaitEvent(s, P.e_h2d[b], 0));
        if (i >= 2) CK(cudaStreamWaitEvent(s, P.e_d2h[b], 0));
        if (concurrent) { CK(cudaEventRecord(P.e_side[0], s)); CK(cudaStreamWaitEvent(P.side, P.e_side[0], 0)); }
        rrtmgp_b200_columns dcs[2]; rrtmgp_b200_fluxes dfs[2];
        for (size_t k = 0; k < sides.size(); ++k)
        {
            Side& sd = sides[k];
            auto D = [&](int idx) -> const void* { return idx < 0 ? nullptr : items[idx].d[b]; };
            auto O = [&](int idx) -> void* { return idx < 0 ? nullptr : outs[idx].d[b]; };
            rrtmgp_b200_columns dc = *sd.c;
            dc.ncol = nc; dc.top_at_1 = top_at_1;
            dc.d_play = D(sd.play); dc.d_plev = D(sd.plev); dc.d_tlay = D(sd.tlay); dc.d_tlev = D(sd.tlev); dc.d_tsfc = D(sd.tsfc);
            dc.d_col_dry = D(sd.coldry); dc.d_sfc_emis = D(sd.emis); dc.d_sfc_alb_dir = D(sd.adir); dc.d_sfc_alb_dif = D(sd.adif);
            dc.d_mu0 = D(sd.mu0); dc.d_tsi_scaling = D(sd.tsi);
            dc.d_lwp = D(sd.lwp); dc.d_iwp = D(sd.iwp); dc.d_rel = D(sd.rel); dc.d_dei = D(sd.dei);
            dc.d_rh = D(sd.rh);
            for (int ai = 0; ai < 11; ++ai)
                if (sd.aer[ai] >= 0)
                {
                    dc.aerosol_concs.d_mmr[ai] = D(sd.aer[ai]);
                    dc.aerosol_concs.mmr_ncol[ai] = items[sd.aer[ai]].kind == 0 ? nc : sd.c->aerosol_concs.mmr_ncol[ai];
                }
            for (size_t gi = 0; gi < sd.vmr.size(); ++gi)
            {
                sd.dv[gi] = D(sd.vmr[gi]);
                sd.vnc[gi] = items[sd.vmr[gi]].kind == 0 ? nc : sd.c->concs.vmr_ncol[gi];
            }
            dc.concs.d_vmr = sd.dv.data(); dc.concs.vmr_ncol = sd.vnc.data(); dc.concs.vmr_nlay = sd.vnl.data();
            rrtmgp_b200_fluxes df{};
            df.d_flux_up = O(so[k].up); df.d_flux_dn = O(so[k].dn); df.d_flux_net = O(so[k].net); df.d_flux_dn_dir = O(so[k].dir);
            df.d_bnd_flux_up = O(so[k].bup); df.d_bnd_flux_dn = O(so[k].bdn); df.d_bnd_flux_net = O(so[k].bnet);
            df.d_bnd_flux_dn_dir = O(so[k].bdir); df.d_heating_rate = O(so[k].hr);
            dcs[k] = dc; dfs[k] = df;
            if (!paired) fused_fluxes<Real>(sd.g, &dc, &df, (concurrent && k == 1) ? P.side : s, sd.sw);
        }
        if (paired) fused_pair<Real>(sides[0].g, &dcs[0], &dfs[0], sides[1].g, &dcs[1], &dfs[1], s);
        if (concurrent) { CK(cudaEventRecord(P.e_side[1], P.side)); CK(cudaStreamWaitEvent(s, P.e_side[1], 0)); }
        CK(cudaEventRecord(P.e_comp[b], s));
        CK(cudaStreamWaitEvent(P.d2h, P.e_comp[b], 0));
        for (auto& ot : outs)
            CK(cudaMemcpy2DAsync((Real*)ot.h + c0, (size_t)ncol * sizeof(Real), ot.d[b], (size_t)nc * sizeof(Real),
                                 (size_t)nc * sizeof(Real), ot.rows, cudaMemcpyDeviceToHost, P.d2h));
        CK(cudaEventRecord(P.e_d2h[b], P.d2h));
    }
    CK(cudaStreamSynchronize(P.d2h));
    CK(cudaStreamSynchronize(s));
}
}  // namespace

extern "C" {

int rrtmgp_b200_lw_fluxes(rrtmgp_b200_gas_optics* g, const rrtmgp_b200_columns* cols, const rrtmgp_b200_fluxes* out, void* stream)
{
    return guarded([&] {
        if (!g) throw ArgError("lw_fluxes: null handle");
        if (g->dtype == RRTMGP_B200_F32) fused_fluxes<float>(g, cols, out, (cudaStream_t)stream, false);
        else fused_fluxes<double>(g, cols, out, (cudaStream_t)stream, false);
    });
}
int rrtmgp_b200_fluxes_host(rrtmgp_b200_gas_optics* go_lw, const rrtmgp_b200_columns* cols_lw, const rrtmgp_b200_fluxes* out_lw,
                            rrtmgp_b200_gas_optics* go_sw, const rrtmgp_b200_columns* cols_sw, const rrtmgp_b200_fluxes* out_sw,
                            int chunk_cols, void* stream)
{
    return guarded([&] {
        rrtmgp_b200_gas_optics* g0 = go_lw ? go_lw : go_sw;
        if (!g0) throw ArgError("fluxes_host: null handles");
        if (g0->dtype == RRTMGP_B200_F32) fluxes_host<float>(go_lw, cols_lw, out_lw, go_sw, cols_sw, out_sw, chunk_cols, (cudaStream_t)stream);
        else fluxes_host<double>(go_lw, cols_lw, out_lw, go_sw, cols_sw, out_sw, chunk_cols, (cudaStream_t)stream);
    });
}
int rrtmgp_b200_host_chunk_schedule(int ncol, int wave_columns, int chunk_columns, int max_chunks, int* first, int* count, int* nchunks)
{
    return guarded([&] {
        if (ncol < 1 || wave_columns < 0 || (wave_columns == 0 && chunk_columns < 1) || !nchunks) throw ArgError("host_chunk_schedule: bad argument");
        const auto sched = chunk_schedule(ncol, wave_columns, std::min(chunk_columns, ncol));
        *nchunks = (int)sched.size();
        for (int i = 0; i < (int)sched.size() && i < max_chunks; ++i)
        {
            if (first) first[i] = sched[i].first;
            if (count) count[i] = sched[i].second;
        }
    }, false);
}
int rrtmgp_b200_lwsw_fluxes(rrtmgp_b200_gas_optics* go_lw, const rrtmgp_b200_columns* cols_lw, const rrtmgp_b200_fluxes* out_lw,
                            rrtmgp_b200_gas_optics* go_sw, const rrtmgp_b200_columns* cols_sw, const rrtmgp_b200_fluxes* out_sw, void* stream)
{
    return guarded([&] {
        if (!go_lw || !go_sw) throw ArgError("lwsw_fluxes: null handle");
        if (go_lw->dtype != go_sw->dtype) throw ArgError("lwsw_fluxes: the two k-distribution objects have different precisions");
        if (go_lw->dtype == RRTMGP_B200_F32) fused_pair<float>(go_lw, cols_lw, out_lw, go_sw, cols_sw, out_sw, (cudaStream_t)stream);
        else fused_pair<double>(go_lw, cols_lw, out_lw, go_sw, cols_sw, out_sw, (cudaStream_t)stream);
    });
}
int rrtmgp_b200_sw_fluxes(rrtmgp_b200_gas_optics* g, const rrtmgp_b200_columns* cols, const rrtmgp_b200_fluxes* out, void* stream)
{
    return guarded([&] {
        if (!g) throw ArgError("sw_fluxes: null handle");
        if (g->dtype == RRTMGP_B200_F32) fused_fluxes<float>(g, cols, out, (cudaStream_t)stream, true);
        else fused_fluxes<double>(g, cols, out, (cudaStream_t)stream, true);
    });
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// Allocation-free shortwave solver (no counterpart in the reference).
// For callers that may not allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc.  Same kernel as rrtmgp_b200_sw_solver_2stream (sw_solver_kernel, stages.cu); its four recurrence
// arrays (ncol, nlay, ngpt) live in this file-scope device array instead of a cudaMallocAsync block.
//   * 32 MiB per GPU that loads the library, held until the process exits;
//   * every call on a device shares it: calls that overlap on the GPU (two streams, two host threads) RACE;
//   * it keeps the previous call's contents.  Nothing has to start at zero: each thread writes every scratch cell
//     of its (column, g-point) before it reads it back, so no clearing is needed;
//   * in host code `g_noalloc_scratch` is a host-side stand-in, not a device address: the address the kernel gets
//     comes from cudaGetSymbolAddress, looked up once per device with that device current, below.
// ---------------------------------------------------------------------------------------------
namespace {
constexpr size_t kNoAllocScratchBytes = RRTMGP_B200_NOALLOC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

constexpr int kNoAllocMaxDevices = 64;
std::atomic<void*> g_noalloc_addr[kNoAllocMaxDevices];      // zero-initialised: not looked up yet

// device address of the scratch array on the CURRENT device (first call per device: one cudaGetSymbolAddress)
void* noalloc_scratch_address()
{
    int dev = -1;
    CK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kNoAllocMaxDevices) throw ArgError("sw_solver_2stream_noalloc: device ordinal beyond 63");
    void* p = g_noalloc_addr[dev].load(std::memory_order_acquire);
    if (!p)
    {
        CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p) throw CudaError("cudaGetSymbolAddress returned a null address for the static scratch array");
        g_noalloc_addr[dev].store(p, std::memory_order_release);      // racing first calls store the same address
    }
    return p;
}
}  // namespace

extern "C" int rrtmgp_b200_sw_solver_2stream_noalloc(int dtype, int ncol, int nlay, int ngpt, int top_at_1, const void* tau,
                                                     const void* ssa, const void* g, const void* mu0, const void* alb_dir,
                                                     const void* alb_dif, const void* inc_dir, const void* inc_dif,
                                                     void* gpt_flux_up, void* gpt_flux_dn, void* gpt_flux_dir, void* stream)
{
    return guarded([&] {
        check_dims(ncol, nlay);
        if (dtype != RRTMGP_B200_F32 && dtype != RRTMGP_B200_F64) throw ArgError("bad dtype");
        if (ngpt < 1 || !tau || !ssa || !g || !mu0 || !alb_dir || !alb_dif || !inc_dir || !gpt_flux_up || !gpt_flux_dn || !gpt_flux_dir)
            throw ArgError("sw_solver_2stream_noalloc: null array");
        const size_t esz = dtype == RRTMGP_B200_F32 ? 4 : 8;
        const size_t cells = (size_t)ncol * (size_t)nlay * (size_t)ngpt;
        const size_t max_cells = kNoAllocScratchBytes / (4 * esz);
        if (cells > max_cells)
            throw ArgError("sw_solver_2stream_noalloc: ncol*nlay*ngpt = " + std::to_string(cells) + " exceeds the static scratch ("
                           + std::to_string(max_cells) + " cells for this dtype, " + std::to_string(kNoAllocScratchBytes)
                           + " bytes); split the columns or use rrtmgp_b200_sw_solver_2stream");
        void* work4 = noalloc_scratch_address();
        DISPATCH(dtype,
            launch_sw_solver<float>(ncol, nlay, ngpt, top_at_1, 0, tau, ssa, g, mu0, alb_dir, alb_dif, inc_dir, inc_dif,
                                    gpt_flux_up, gpt_flux_dn, gpt_flux_dir, work4, (cudaStream_t)stream),
            launch_sw_solver<double>(ncol, nlay, ngpt, top_at_1, 0, tau, ssa, g, mu0, alb_dir, alb_dif, inc_dir, inc_dif,
                                     gpt_flux_up, gpt_flux_dn, gpt_flux_dir, work4, (cudaStream_t)stream));
    });
}

// ---------------------------------------------------------------------------------------------
// Multi-GPU: column-shard plan + flux gather over NCCL.  The library has no link-time dependency on NCCL:
// libnccl.so.2 is resolved with dlopen on first use (the copy already in the process when the host model or PyTorch loaded
// one; RB200_NCCL_LIB overrides the name), and only the eight entry points below are used.
// ---------------------------------------------------------------------------------------------
#include <dlfcn.h>
namespace {
struct NcclUniqueId { char internal[128]; };                    // nccl.h: NCCL_UNIQUE_ID_BYTES
struct NcclApi
{
    void* lib = nullptr;
    int (*GetUniqueId)(NcclUniqueId*) = nullptr;
    int (*CommInitRank)(void**, int, NcclUniqueId, int) = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    int (*CommCount)(void*, int*) = nullptr;
    int (*CommUserRank)(void*, int*) = nullptr;
    int (*Send)(const void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
NcclApi& nccl()
{
    static NcclApi api;
    static std::once_flag once;
    static std::string err;
    std::call_once(once, [] {
        const char* names[] = {getenv("RB200_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
        for (const char* n : names)
        {
            if (!n || !*n) continue;
            api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
            if (api.lib) break;
            err = dlerror();
        }
        if (!api.lib) return;
        auto sym = [&](const char* s) { void* p = dlsym(api.lib, s); if (!p) err = std::string("missing symbol ") + s; return p; };
        api.GetUniqueId = (int (*)(NcclUniqueId*))sym("ncclGetUniqueId");
        api.CommInitRank = (int (*)(void**, int, NcclUniqueId, int))sym("ncclCommInitRank");
        api.CommDestroy = (int (*)(void*))sym("ncclCommDestroy");
        api.CommCount = (int (*)(void*, int*))sym("ncclCommCount");
        api.CommUserRank = (int (*)(void*, int*))sym("ncclCommUserRank");
        api.Send = (int (*)(const void*, size_t, int, int, void*, cudaStream_t))sym("ncclSend");
        api.Recv = (int (*)(void*, size_t, int, int, void*, cudaStream_t))sym("ncclRecv");
        api.GroupStart = (int (*)())sym("ncclGroupStart");
        api.GroupEnd = (int (*)())sym("ncclGroupEnd");
        api.GetErrorString = (const char* (*)(int))sym("ncclGetErrorString");
    });
    if (!api.lib || !api.GetUniqueId || !api.CommInitRank || !api.CommDestroy || !api.Send || !api.Recv || !api.GroupStart || !api.GroupEnd)
        throw CudaError("NCCL is not available (dlopen libnccl.so.2: " + (err.empty() ? std::string("unknown error") : err) + "); set RB200_NCCL_LIB");
    return api;
}
void nccl_ck(int rc, const char* what)
{
    if (rc == 0) return;
    NcclApi& n = nccl();
    throw CudaError(std::string(what) + ": " + (n.GetErrorString ? n.GetErrorString(rc) : "NCCL error") + " (" + std::to_string(rc) + ")");
}
// the partition rule of rrtmgp_b200_shard_bounds
void shard_range(long long n, int parts, int i, long long& a, long long& b)
{
    a = (long long)((__int128)i * n / parts); b = (long long)((__int128)(i + 1) * n / parts);
}
}  // namespace

struct rrtmgp_b200_shard_plan
{
    long long ncol = 0; int world = 1, rank = 0;
    void* comm = nullptr; bool owns_comm = false;
    Arena stage;                    // root: landing area of the peers' slabs (grow-only, allocated at the first gather)
    // peer-memory mode (rrtmgp_b200_shard_plan_register): the root's whole-domain arrays mapped into this process
    int reg_root = -1; int reg_nfields = 0;
    std::vector<void*> reg_full;    // root: the registered arrays; peers: their addresses in THIS process (CUDA IPC mappings)
    std::vector<void*> ipc_opened;  // peers: mappings to close
    void* signal = nullptr;         // one device word for the completion messages
};
// RB200_GATHER=direct: every level of a slab received straight into its place (no staging copy, but nlev x nfields small
// messages per peer: measured 11.9 ms for 1.35 GB at N=2, 113 GB/s).  Default "staged": one message per field and peer into a
// staging buffer on the root, then one strided device copy per (peer, field) into the reference layout.
static const bool g_gather_direct = [] { const char* e = getenv("RB200_GATHER"); return e && e[0] == 'd'; }();

extern "C" {

int rrtmgp_b200_nccl_unique_id(void* id128)
{
    return guarded([&] {
        if (!id128) throw ArgError("nccl_unique_id: null buffer");
        NcclUniqueId id;
        nccl_ck(nccl().GetUniqueId(&id), "ncclGetUniqueId");
        std::memcpy(id128, id.internal, sizeof id.internal);
    }, false);
}

int rrtmgp_b200_shard_plan_create(long long ncol, int world, int rank, const void* id128, void* nccl_comm, rrtmgp_b200_shard_plan** out)
{
    return guarded([&] {
        if (!out) throw ArgError("shard_plan_create: null output");
        *out = nullptr;
        if (ncol < 0 || world < 1 || rank < 0 || rank >= world) throw ArgError("shard_plan_create: need ncol >= 0 and 0 <= rank < world");
        std::unique_ptr<rrtmgp_b200_shard_plan> p(new rrtmgp_b200_shard_plan);
        p->ncol = ncol; p->world = world; p->rank = rank;
        if (nccl_comm)
        {
            // the host model's communicator (ncclComm_t): checked against world / rank, never destroyed here
            NcclApi& n = nccl();
            int cw = world, cr = rank;
            if (n.CommCount && n.CommUserRank) { nccl_ck(n.CommCount(nccl_comm, &cw), "ncclCommCount"); nccl_ck(n.CommUserRank(nccl_comm, &cr), "ncclCommUserRank"); }
            if (cw != world || cr != rank) throw ArgError("shard_plan_create: communicator size / rank differ from world / rank");
            p->comm = nccl_comm;
        }
        else if (world > 1)
        {
            if (!id128) throw ArgError("shard_plan_create: world > 1 needs a communicator or the 128-byte unique id of rank 0");
            NcclUniqueId id;
            std::memcpy(id.internal, id128, sizeof id.internal);
            nccl_ck(nccl().CommInitRank(&p->comm, world, id, rank), "ncclCommInitRank");
            p->owns_comm = true;
        }
        *out = p.release();
    }, false);
}

void rrtmgp_b200_shard_plan_destroy(rrtmgp_b200_shard_plan* p)
{
    if (!p) return;
    for (void* m : p->ipc_opened) cudaIpcCloseMemHandle(m);
    if (p->signal) cudaFree(p->signal);
    if (p->owns_comm && p->comm) { try { nccl().CommDestroy(p->comm); } catch (...) {} }
    delete p;
}

int rrtmgp_b200_shard_plan_bounds(const rrtmgp_b200_shard_plan* p, int rank, long long* c0, long long* c1)
{
    if (!p || !c0 || !c1) return fail(RRTMGP_B200_E_ARG, "shard_plan_bounds: null argument");
    if (rank < 0) rank = p->rank;
    if (rank >= p->world) return fail(RRTMGP_B200_E_ARG, "shard_plan_bounds: rank outside the plan");
    shard_range(p->ncol, p->world, rank, *c0, *c1);
    return 0;
}

// base address and size of the allocation a device pointer lies in (driver entry point resolved at run time: the library
// does not link libcuda)
static bool allocation_base(const void* ptr, char** base)
{
    typedef int (*Fn)(unsigned long long*, size_t*, unsigned long long);
    static Fn fn = [] { void* h = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL); return h ? (Fn)dlsym(h, "cuMemGetAddressRange_v2") : (Fn) nullptr; }();
    unsigned long long b = 0; size_t n = 0;
    if (!fn || fn(&b, &n, (unsigned long long)(uintptr_t)ptr) != 0) return false;
    *base = (char*)(uintptr_t)b;
    return true;
}

int rrtmgp_b200_shard_plan_register(rrtmgp_b200_shard_plan* p, int nfields, void* const* d_full_fields, int root, void* stream)
{
    return guarded([&] {
        if (!p || nfields < 1 || nfields > 64 || root < 0 || root >= p->world) throw ArgError("shard_plan_register: bad argument");
        const bool is_root = p->rank == root;
        if (is_root && !d_full_fields) throw ArgError("shard_plan_register: the root needs the whole-domain arrays");
        for (void* m : p->ipc_opened) cudaIpcCloseMemHandle(m);
        p->ipc_opened.clear(); p->reg_full.clear(); p->reg_root = -1; p->reg_nfields = 0;
        if (p->world == 1) return;
        if (!p->comm) throw ArgError("shard_plan_register: the plan has no communicator");
        NcclApi& n = nccl();
        cudaStream_t s = (cudaStream_t)stream;
        struct Entry { cudaIpcMemHandle_t h; unsigned long long offset; int ok; int pad; };
        std::vector<Entry> host((size_t)nfields);
        Entry* dev = nullptr;
        CK(cudaMalloc((void**)&dev, sizeof(Entry) * nfields));
        int* dflag = nullptr;
        CK(cudaMalloc((void**)&dflag, sizeof(int) * (size_t)(p->world + 1)));
        if (!p->signal) CK(cudaMalloc(&p->signal, 256));
        auto cleanup = [&] { cudaFree(dev); cudaFree(dflag); };
        try
        {
            int ok = 1;
            if (is_root)
            {
                for (int f = 0; f < nfields; ++f)
                {
                    char* base = nullptr;
                    host[f].ok = d_full_fields[f] && allocation_base(d_full_fields[f], &base) &&
                                 cudaIpcGetMemHandle(&host[f].h, base) == cudaSuccess;
                    host[f].offset = host[f].ok ? (unsigned long long)((char*)d_full_fields[f] - base) : 0;
                    if (!host[f].ok) { ok = 0; cudaGetLastError(); }
                }
                CK(cudaMemcpyAsync(dev, host.data(), sizeof(Entry) * nfields, cudaMemcpyHostToDevice, s));
            }
            // 1. handles root -> peers
            nccl_ck(n.GroupStart(), "ncclGroupStart");
            if (is_root) { for (int r = 0; r < p->world; ++r) if (r != root) nccl_ck(n.Send(dev, sizeof(Entry) * nfields, 1 /*ncclUint8*/, r, p->comm, s), "ncclSend"); }
            else nccl_ck(n.Recv(dev, sizeof(Entry) * nfields, 1, root, p->comm, s), "ncclRecv");
            nccl_ck(n.GroupEnd(), "ncclGroupEnd");
            CK(cudaStreamSynchronize(s));
            std::vector<void*> mapped((size_t)nfields, nullptr);
            if (!is_root)
            {
                CK(cudaMemcpy(host.data(), dev, sizeof(Entry) * nfields, cudaMemcpyDeviceToHost));
                std::vector<std::pair<cudaIpcMemHandle_t, void*>> seen;        // arrays may share an allocation: open it once
                for (int f = 0; f < nfields && ok; ++f)
                {
                    if (!host[f].ok) { ok = 0; break; }
                    void* m = nullptr;
                    for (auto& sv : seen) if (std::memcmp(&sv.first, &host[f].h, sizeof(cudaIpcMemHandle_t)) == 0) m = sv.second;
                    if (!m)
                    {
                        if (cudaIpcOpenMemHandle(&m, host[f].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = 0; break; }
                        seen.push_back({host[f].h, m}); p->ipc_opened.push_back(m);
                    }
                    mapped[f] = (char*)m + host[f].offset;
                }
            }
            // 2. every rank's verdict -> root; 3. the decision root -> peers (all or nothing)
            CK(cudaMemcpyAsync(dflag + p->rank, &ok, sizeof(int), cudaMemcpyHostToDevice, s));
            nccl_ck(n.GroupStart(), "ncclGroupStart");
            if (is_root) { for (int r = 0; r < p->world; ++r) if (r != root) nccl_ck(n.Recv(dflag + r, 1, 2 /*ncclInt32*/, r, p->comm, s), "ncclRecv"); }
            else nccl_ck(n.Send(dflag + p->rank, 1, 2, root, p->comm, s), "ncclSend");
            nccl_ck(n.GroupEnd(), "ncclGroupEnd");
            CK(cudaStreamSynchronize(s));
            int all = ok;
            if (is_root)
            {
                std::vector<int> flags((size_t)p->world);
                CK(cudaMemcpy(flags.data(), dflag, sizeof(int) * p->world, cudaMemcpyDeviceToHost));
                for (int r = 0; r < p->world; ++r) if (r != root && !flags[r]) all = 0;
                CK(cudaMemcpyAsync(dflag + p->world, &all, sizeof(int), cudaMemcpyHostToDevice, s));
            }
            nccl_ck(n.GroupStart(), "ncclGroupStart");
            if (is_root) { for (int r = 0; r < p->world; ++r) if (r != root) nccl_ck(n.Send(dflag + p->world, 1, 2, r, p->comm, s), "ncclSend"); }
            else nccl_ck(n.Recv(dflag + p->world, 1, 2, root, p->comm, s), "ncclRecv");
            nccl_ck(n.GroupEnd(), "ncclGroupEnd");
            CK(cudaStreamSynchronize(s));
            if (!is_root) CK(cudaMemcpy(&all, dflag + p->world, sizeof(int), cudaMemcpyDeviceToHost));
            if (all)
            {
                p->reg_root = root; p->reg_nfields = nfields;
                if (is_root) p->reg_full.assign(d_full_fields, d_full_fields + nfields); else p->reg_full = mapped;
            }
            else
            {
                for (void* m : p->ipc_opened) cudaIpcCloseMemHandle(m);
                p->ipc_opened.clear();
            }
        }
        catch (...) { cleanup(); throw; }
        cleanup();
    });
}

int rrtmgp_b200_shard_plan_registered(const rrtmgp_b200_shard_plan* p) { return p && p->reg_root >= 0 ? p->reg_nfields : 0; }

int rrtmgp_b200_gather_fluxes(rrtmgp_b200_shard_plan* p, int dtype, int nlev, int nfields, const void* const* d_part_fields,
                              void* const* d_full_fields, int part, int nparts, int root, void* stream)
{
    return guarded([&] {
        if (!p) throw ArgError("gather_fluxes: null plan");
        if (dtype != RRTMGP_B200_F32 && dtype != RRTMGP_B200_F64) throw ArgError("bad dtype");
        if (nlev < 1 || nfields < 1 || nparts < 1 || part < 0 || part >= nparts || root < 0 || root >= p->world) throw ArgError("gather_fluxes: bad argument");
        if (!d_part_fields) throw ArgError("gather_fluxes: null field list");
        const bool is_root = p->rank == root;
        if (is_root && !d_full_fields) throw ArgError("gather_fluxes: the root needs the full-domain arrays");
        const size_t esz = dtype == RRTMGP_B200_F32 ? 4 : 8;
        const int dt = dtype == RRTMGP_B200_F32 ? 7 : 8;             // ncclFloat32 / ncclFloat64
        cudaStream_t s = (cudaStream_t)stream;
        // columns of part `part` of rank r's shard, in domain coordinates
        auto span = [&](int r, long long& first, long long& count) {
            long long a, b, pa, pb;
            shard_range(p->ncol, p->world, r, a, b);
            shard_range(b - a, nparts, part, pa, pb);
            first = a + pa; count = pb - pa;
        };
        long long my_first, my_count;
        span(p->rank, my_first, my_count);
        for (int f = 0; f < nfields; ++f)
            if ((my_count > 0 && !d_part_fields[f]) || (is_root && !d_full_fields[f])) throw ArgError("gather_fluxes: null field");
        if (is_root && my_count > 0)        // the root's own columns: one strided device copy per field
            for (int f = 0; f < nfields; ++f)
                CK(cudaMemcpy2DAsync((char*)d_full_fields[f] + (size_t)my_first * esz, (size_t)p->ncol * esz, d_part_fields[f],
                                     (size_t)my_count * esz, (size_t)my_count * esz, (size_t)nlev, cudaMemcpyDeviceToDevice, s));
        if (p->world == 1) return;
        if (!p->comm) throw ArgError("gather_fluxes: the plan has no communicator");
        NcclApi& n = nccl();
        if (p->reg_root == root && p->reg_nfields == nfields)
        {
            // peer-memory mode: every rank WRITES its slab into the root's arrays over NVLink (one strided peer copy per
            // field, straight into the reference layout: no staging buffer, no layout copy on the root, all peers at once),
            // then tells the root with a one-element message that its columns have landed
            if (is_root)
                for (int f = 0; f < nfields; ++f)
                    if (d_full_fields[f] != p->reg_full[f]) throw ArgError("gather_fluxes: the whole-domain arrays differ from the registered ones");
            // "clear to write": the root's stream has reached this gather, i.e. everything it enqueued before (the consumers
            // of the previous result) is ordered ahead; a peer's copies wait for it
            nccl_ck(n.GroupStart(), "ncclGroupStart");
            if (is_root) { for (int r = 0; r < p->world; ++r) if (r != root) nccl_ck(n.Send(p->signal, 1, 2, r, p->comm, s), "ncclSend"); }
            else nccl_ck(n.Recv((int*)p->signal + 40, 1, 2, root, p->comm, s), "ncclRecv");
            nccl_ck(n.GroupEnd(), "ncclGroupEnd");
            if (!is_root && my_count > 0)
                for (int f = 0; f < nfields; ++f)
                    CK(cudaMemcpy2DAsync((char*)p->reg_full[f] + (size_t)my_first * esz, (size_t)p->ncol * esz, d_part_fields[f],
                                         (size_t)my_count * esz, (size_t)my_count * esz, (size_t)nlev, cudaMemcpyDeviceToDevice, s));
            nccl_ck(n.GroupStart(), "ncclGroupStart");
            if (is_root) { for (int r = 0; r < p->world; ++r) if (r != root) nccl_ck(n.Recv((int*)p->signal + 1 + (r % 32), 1, 2, r, p->comm, s), "ncclRecv"); }
            else nccl_ck(n.Send(p->signal, 1, 2 /*ncclInt32*/, root, p->comm, s), "ncclSend");
            nccl_ck(n.GroupEnd(), "ncclGroupEnd");
            return;
        }
        if (!g_gather_direct)
        {
            // staged: a peer's (count, nlev) slab of a field is ONE message; the root lays it out with a strided copy
            char* base = nullptr;
            std::vector<size_t> off((size_t)p->world * nfields, 0);
            if (is_root)
            {
                size_t total = 0;
                for (int r = 0; r < p->world; ++r)
                {
                    if (r == root) continue;
                    long long first, count; span(r, first, count);
                    for (int f = 0; f < nfields; ++f) { off[(size_t)r * nfields + f] = total; total += pad256((size_t)count * nlev * esz); }
                }
                p->stage.reserve(total + 256, s);       // grows (synchronising) only while the sizes still increase
                base = p->stage.take<char>(total);
            }
            nccl_ck(n.GroupStart(), "ncclGroupStart");
            if (is_root)
            {
                for (int r = 0; r < p->world; ++r)
                {
                    if (r == root) continue;
                    long long first, count; span(r, first, count);
                    if (count == 0) continue;
                    for (int f = 0; f < nfields; ++f)
                        nccl_ck(n.Recv(base + off[(size_t)r * nfields + f], (size_t)count * nlev, dt, r, p->comm, s), "ncclRecv");
                }
            }
            else if (my_count > 0)
                for (int f = 0; f < nfields; ++f)
                    nccl_ck(n.Send(d_part_fields[f], (size_t)my_count * nlev, dt, root, p->comm, s), "ncclSend");
            nccl_ck(n.GroupEnd(), "ncclGroupEnd");
            if (is_root)
                for (int r = 0; r < p->world; ++r)
                {
                    if (r == root) continue;
                    long long first, count; span(r, first, count);
                    if (count == 0) continue;
                    for (int f = 0; f < nfields; ++f)
                        CK(cudaMemcpy2DAsync((char*)d_full_fields[f] + (size_t)first * esz, (size_t)p->ncol * esz, base + off[(size_t)r * nfields + f],
                                             (size_t)count * esz, (size_t)count * esz, (size_t)nlev, cudaMemcpyDeviceToDevice, s));
                }
            return;
        }
        // direct: every level of a part is one contiguous run of `count` columns on both sides: it goes straight into the
        // reference-layout (ncol, nlev) array of the root, no staging buffer and no reordering pass
        nccl_ck(n.GroupStart(), "ncclGroupStart");
        if (is_root)
        {
            for (int r = 0; r < p->world; ++r)
            {
                if (r == root) continue;
                long long first, count;
                span(r, first, count);
                if (count == 0) continue;
                for (int f = 0; f < nfields; ++f)
                    for (int l = 0; l < nlev; ++l)
                        nccl_ck(n.Recv((char*)d_full_fields[f] + ((size_t)l * p->ncol + (size_t)first) * esz, (size_t)count, dt, r, p->comm, s), "ncclRecv");
            }
        }
        else if (my_count > 0)
        {
            for (int f = 0; f < nfields; ++f)
                for (int l = 0; l < nlev; ++l)
                    nccl_ck(n.Send((const char*)d_part_fields[f] + (size_t)l * my_count * esz, (size_t)my_count, dt, root, p->comm, s), "ncclSend");
        }
        nccl_ck(n.GroupEnd(), "ncclGroupEnd");
    });
}

}  // extern "C"

// plan.hpp — host-side table preparation of the k-distribution object.
//
// Re-derives, from the raw coefficient-file variables, the member tables the reference builds
// in Gas_optics_rrtmgp's constructor + init_abs_coeffs (src/Gas_optics_rrtmgp.cpp:391-742:
// reduce_minor_arrays :54-162, create_idx_minor :164, create_idx_minor_scaling :185,
// create_key_species_reduce :198, create_flavor :247, create_gpoint_flavor :331,
// set_solar_variability :745-756), and lays the big tables out for the B200 kernels:
// band-blocked with the g-point index fastest and every band padded to `gw` g-points, so the
// 8 (major) / 4 (minor, Rayleigh) interpolation corners of a (column, layer, band) are
// contiguous, 16-byte aligned rows.  Pure host C++ (no CUDA) — unit-testable on a CPU box.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>
#include "../../include/rrtmgp_b200_tables.h"

namespace rb200 {

struct ArgError : std::runtime_error { using std::runtime_error::runtime_error; };
struct MissingError : std::runtime_error { using std::runtime_error::runtime_error; };

inline std::string trimmed(const char* block, int i, int len)
{
    std::string s(block + (size_t)i * len, len);
    size_t b = s.find_first_not_of(" \0", 0, 2), e = s.find_last_not_of(" \0", std::string::npos, 2);
    if (b == std::string::npos) return std::string();
    return s.substr(b, e - b + 1);
}

constexpr int MAX_MINOR = 16;   // per (band, side)

struct MinorDesc
{
    int idx_minor;     // index into col_gas (0 = dry air slot, gases 1..ngas); -1 if absent
    int idx_scaling;   // idem for the scaling gas, <=0: none
    int scales_with_density;
    int scale_by_complement;
    int contributor;   // position in the reduced contributor list (table block index)
};

struct MinorSide
{
    int n = 0;                               // reduced contributors
    std::vector<int> limits;                 // (2,n) 1-based g-point limits
    std::vector<int> kstart;                 // 1-based start in the reduced (unpadded) kminor
    std::vector<MinorDesc> desc;             // n
    std::vector<std::vector<int>> by_band;   // nbnd lists of contributor positions
    std::vector<double> kminor_ref;          // reduced, reference layout (ntemp, neta, nk)
    int nk = 0;
    std::vector<double> table;               // device layout [m][it][ie][gw]
    // many contributors share one scaling (same gas, same scaling gas, same flags; e.g. the h2o continua of
    // every band): the prologue kernel computes each distinct one once per (column, layer)
    std::vector<int> uid;                    // per contributor: its distinct-scaling number
    std::vector<int> uniq_rep;               // per distinct scaling: a contributor that has it
    std::vector<int> scal_off;               // per band: first slot of the band's contributors (slots per band padded to 4)
    std::vector<int> slot_desc;              // per slot: index into desc, -1 = padding
};

struct HostPlan
{
    bool is_lw = false;
    int ntemp = 0, npres = 0, neta = 0, nbnd = 0, ngpt = 0, nflav = 0, ngas = 0, gw = 0;
    int nplanck = 0, idx_h2o = 0, row_stride = 0;
    std::vector<std::string> gas_names;          // reduced list, slot i+1 of col_gas
    std::vector<int> band_lims_gpt;              // (2,nbnd) 1-based
    std::vector<double> band_lims_wvn;           // (2,nbnd)
    std::vector<int> gpoint_bands;               // (ngpt) 1-based
    std::vector<int> flavor;                     // (2,nflav) 1-based gas indices (0 = dry air)
    std::vector<int> gpoint_flavor;              // (2,ngpt) 1-based
    std::vector<int> band_flavor;                // (2,nbnd) 0-based flavor of each band
    std::vector<double> vmr_ref;                 // (2, ngas+1, ntemp)
    std::vector<double> press_ref, press_ref_log, temp_ref;
    double press_ref_log_delta = 0, temp_ref_min = 0, temp_ref_max = 0, temp_ref_delta = 0;
    double press_ref_min = 0, press_ref_max = 0, press_ref_trop_log = 0, totplnk_delta = 0;
    std::vector<double> kmajor;                  // device layout [b][it][ip][ie][slot][gw]; slot 1 = planck_frac (LW)
    std::vector<double> krayl;                   // device layout [side][b][it][ie][gw]
    std::vector<double> totplnk;                 // [b][nplanck]
    std::vector<double> solar_source;            // (ngpt)
    double tsi = 0;
    MinorSide minor[2];                          // 0 = lower (tropo true), 1 = upper
    int nscal = 0;                               // slots of the band-blocked scaling vector (max over sides, multiple of 4)
    int nuniq = 0;                               // distinct scalings per (column, layer): max over the two sides, >= 1
};

inline HostPlan build_host_plan(const rrtmgp_b200_kdist_file& kd, int n_avail,
                                const char* avail_block, int avail_len)
{
    HostPlan p;
    if (!kd.kmajor || !kd.key_species || !kd.bnd_limits_gpt || !kd.vmr_ref || !kd.press_ref || !kd.temp_ref)
        throw ArgError("k-distribution: missing mandatory table");
    const int sl = kd.string_len;
    std::vector<std::string> avail;
    for (int i = 0; i < n_avail; ++i) avail.push_back(trimmed(avail_block, i, avail_len));
    auto is_avail = [&](const std::string& s) { return std::find(avail.begin(), avail.end(), s) != avail.end(); };

    std::vector<std::string> file_names;
    for (int i = 0; i < kd.n_absorbers; ++i) file_names.push_back(trimmed(kd.gas_names, i, sl));
    // src/Gas_optics_rrtmgp.cpp:574-586
    for (auto& s : file_names) if (is_avail(s)) p.gas_names.push_back(s);
    p.ngas = (int)p.gas_names.size();
    auto find_gas = [&](const std::string& s) {   // 1-based, -1 if absent (find_index)
        auto it = std::find(p.gas_names.begin(), p.gas_names.end(), s);
        return it == p.gas_names.end() ? -1 : (int)(it - p.gas_names.begin()) + 1;
    };
    auto file_index = [&](const std::string& s) {
        auto it = std::find(file_names.begin(), file_names.end(), s);
        return it == file_names.end() ? -1 : (int)(it - file_names.begin());
    };

    p.ntemp = kd.n_temps; p.npres = kd.n_press; p.neta = kd.n_mixingfracs;
    p.nbnd = kd.n_bnds; p.ngpt = kd.n_gpts;
    p.is_lw = kd.totplnk != nullptr;
    const int ntemp = p.ntemp, npres = p.npres, neta = p.neta, nbnd = p.nbnd, ngpt = p.ngpt;
    if (ntemp < 2 || npres < 2 || neta < 2 || nbnd < 1 || ngpt < 1) throw ArgError("k-distribution: bad dimensions");

    // band bookkeeping (src/Optical_props.cpp:31-72); file (bnd, 2) -> (2, nbnd)
    p.band_lims_gpt.resize(2 * nbnd); p.band_lims_wvn.resize(2 * nbnd); p.gpoint_bands.assign(ngpt, 0);
    int maxw = 0;
    for (int b = 0; b < nbnd; ++b)
    {
        p.band_lims_gpt[2 * b] = kd.bnd_limits_gpt[2 * b]; p.band_lims_gpt[2 * b + 1] = kd.bnd_limits_gpt[2 * b + 1];
        p.band_lims_wvn[2 * b] = kd.bnd_limits_wavenumber[2 * b]; p.band_lims_wvn[2 * b + 1] = kd.bnd_limits_wavenumber[2 * b + 1];
        const int g1 = p.band_lims_gpt[2 * b], g2 = p.band_lims_gpt[2 * b + 1];
        if (g1 < 1 || g2 > ngpt || g2 < g1) throw ArgError("k-distribution: bad band limits");
        for (int g = g1; g <= g2; ++g) p.gpoint_bands[g - 1] = b + 1;
        maxw = std::max(maxw, g2 - g1 + 1);
    }
    if (maxw > 16) throw ArgError("k-distribution: bands wider than 16 g-points are not supported");
    p.gw = maxw <= 8 ? 8 : 16;
    const int gw = p.gw;

    // :589-605 vmr_ref(2, 0:ngas, ntemp); file order (temperature, absorber_ext, atmos_layer)
    const int next = kd.n_extabsorbers;
    p.vmr_ref.assign((size_t)2 * (p.ngas + 1) * ntemp, 0.0);
    for (int it = 0; it < ntemp; ++it)
        for (int ia = 0; ia < 2; ++ia)
        {
            p.vmr_ref[ia + 2 * (0 + (size_t)(p.ngas + 1) * it)] = kd.vmr_ref[((size_t)it * next + 0) * 2 + ia];
            for (int i = 0; i < p.ngas; ++i)
            {
                const int fi = file_index(p.gas_names[i]) + 1;
                p.vmr_ref[ia + 2 * ((i + 1) + (size_t)(p.ngas + 1) * it)] = kd.vmr_ref[((size_t)it * next + fi) * 2 + ia];
            }
        }

    p.press_ref.assign(kd.press_ref, kd.press_ref + npres);
    p.temp_ref.assign(kd.temp_ref, kd.temp_ref + ntemp);
    p.press_ref_log.resize(npres);
    for (int i = 0; i < npres; ++i) p.press_ref_log[i] = std::log(p.press_ref[i]);
    p.press_ref_trop_log = std::log(kd.press_ref_trop);
    // :712-723
    p.temp_ref_min = p.temp_ref[0]; p.temp_ref_max = p.temp_ref[ntemp - 1];
    p.press_ref_min = p.press_ref[npres - 1]; p.press_ref_max = p.press_ref[0];
    p.press_ref_log_delta = (std::log(p.press_ref_min) - std::log(p.press_ref_max)) / (npres - 1);
    p.temp_ref_delta = (p.temp_ref_max - p.temp_ref_min) / (ntemp - 1);

    // :198-236 create_key_species_reduce; file key_species(bnd, atmos_layer, pair)
    std::vector<int> ks((size_t)2 * 2 * nbnd, 0);   // (pair, atmos_layer, bnd)
    for (int b = 0; b < nbnd; ++b)
        for (int ia = 0; ia < 2; ++ia)
            for (int ip = 0; ip < 2; ++ip)
            {
                const int k = kd.key_species[((size_t)b * 2 + ia) * 2 + ip];
                int red = 0;
                if (k != 0)
                {
                    if (k < 1 || k > kd.n_absorbers) throw ArgError("k-distribution: bad key species");
                    red = find_gas(file_names[k - 1]);
                    if (red < 0) throw MissingError("Gas optics: required gas " + file_names[k - 1] + " is missing");
                }
                ks[ip + 2 * (ia + 2 * b)] = red;
            }
    // :247-311 create_flavor — unique pairs in (band, atmos_layer) order; (0,0) -> (2,2)
    std::vector<std::pair<int, int>> flav;
    auto pair_of = [&](int ia, int b) {
        std::pair<int, int> q(ks[0 + 2 * (ia + 2 * b)], ks[1 + 2 * (ia + 2 * b)]);
        if (q.first == 0 && q.second == 0) q = {2, 2};
        return q;
    };
    for (int b = 0; b < nbnd; ++b)
        for (int ia = 0; ia < 2; ++ia)
        {
            auto q = pair_of(ia, b);
            if (std::find(flav.begin(), flav.end(), q) == flav.end()) flav.push_back(q);
        }
    p.nflav = (int)flav.size();
    p.flavor.resize(2 * p.nflav);
    for (int f = 0; f < p.nflav; ++f) { p.flavor[2 * f] = flav[f].first; p.flavor[2 * f + 1] = flav[f].second; }
    // :331-362 create_gpoint_flavor
    p.gpoint_flavor.resize(2 * ngpt); p.band_flavor.resize(2 * nbnd);
    for (int b = 0; b < nbnd; ++b)
        for (int ia = 0; ia < 2; ++ia)
        {
            auto q = pair_of(ia, b);
            p.band_flavor[ia + 2 * b] = (int)(std::find(flav.begin(), flav.end(), q) - flav.begin());
        }
    for (int g = 0; g < ngpt; ++g)
        for (int ia = 0; ia < 2; ++ia)
            p.gpoint_flavor[ia + 2 * g] = p.band_flavor[ia + 2 * (p.gpoint_bands[g] - 1)] + 1;

    p.idx_h2o = find_gas("h2o");
    if (p.idx_h2o < 0) throw MissingError("idx_h2o cannot be found");

    // Big tables.  File kmajor(temperature, pressure_interp, mixing_fraction, gpt) (C order);
    // reference member kmajor(ntemp, neta, npres+1, ngpt) (:657-663).  Device: [b][it][ip][ie][gw].
    // Longwave: kmajor and planck_frac rows are interleaved ([...][ie][2][gw]) so the two tables'
    // corners of a cell share one 128-byte line (fp32) and one address computation.
    const int rs = p.is_lw ? 2 * gw : gw;          // row stride in elements
    p.row_stride = rs;
    if ((long long)ntemp * (npres + 1) * neta >= (1 << 20) || ntemp * neta >= (1 << 12))
        throw ArgError("k-distribution: table too large for the packed row indices");
    auto block4 = [&](const double* src, int slot, std::vector<double>& out) {
        out.resize((size_t)nbnd * ntemp * (npres + 1) * neta * rs, 0.0);
        for (int b = 0; b < nbnd; ++b)
        {
            const int g1 = p.band_lims_gpt[2 * b] - 1, n = p.band_lims_gpt[2 * b + 1] - g1;
            for (int it = 0; it < ntemp; ++it)
                for (int ip = 0; ip <= npres; ++ip)
                    for (int ie = 0; ie < neta; ++ie)
                    {
                        const double* s = src + (((size_t)it * (npres + 1) + ip) * neta + ie) * ngpt + g1;
                        double* d = &out[((((size_t)b * ntemp + it) * (npres + 1) + ip) * neta + ie) * rs + (size_t)slot * gw];
                        for (int g = 0; g < n; ++g) d[g] = s[g];
                    }
        }
    };
    block4(kd.kmajor, 0, p.kmajor);
    if (p.is_lw)
    {
        if (!kd.plank_fraction || kd.n_planck_temps < 2) throw ArgError("longwave k-distribution without Planck tables");
        block4(kd.plank_fraction, 1, p.kmajor);
        p.nplanck = kd.n_planck_temps;
        p.totplnk.assign(kd.totplnk, kd.totplnk + (size_t)nbnd * p.nplanck);   // file (bnd, temperature_Planck)
        p.totplnk_delta = (p.temp_ref_max - p.temp_ref_min) / (p.nplanck - 1);  // :458
    }
    else
    {
        if (!kd.solar_source_quiet) throw ArgError("shortwave k-distribution without solar source");
        // :745-756 set_solar_variability
        p.solar_source.resize(ngpt);
        const double a_offset = 0.1495954, b_offset = 0.00066696;
        p.tsi = 0;
        for (int g = 0; g < ngpt; ++g)
        {
            p.solar_source[g] = kd.solar_source_quiet[g] + (kd.mg_default - a_offset) * kd.solar_source_facular[g]
                              + (kd.sb_default - b_offset) * kd.solar_source_sunspot[g];
            p.tsi += p.solar_source[g];
        }
        if (kd.rayl_lower && kd.rayl_upper)
        {
            // :667-677 krayl(ntemp, neta, ngpt, 2); file rayl_*(temperature, mixing_fraction, gpt)
            p.krayl.assign((size_t)2 * nbnd * ntemp * neta * gw, 0.0);
            for (int s = 0; s < 2; ++s)
            {
                const double* src = s == 0 ? kd.rayl_lower : kd.rayl_upper;
                for (int b = 0; b < nbnd; ++b)
                {
                    const int g1 = p.band_lims_gpt[2 * b] - 1, n = p.band_lims_gpt[2 * b + 1] - g1;
                    for (int it = 0; it < ntemp; ++it)
                        for (int ie = 0; ie < neta; ++ie)
                            for (int g = 0; g < n; ++g)
                                p.krayl[((((size_t)s * nbnd + b) * ntemp + it) * neta + ie) * gw + g] =
                                        src[((size_t)it * neta + ie) * ngpt + g1 + g];
                }
            }
        }
    }

    // :54-162 reduce_minor_arrays + :164-196 idx_minor / idx_minor_scaling, per side
    std::vector<std::string> ident, gminor;
    for (int i = 0; i < kd.n_minorabsorbers; ++i)
    {
        ident.push_back(trimmed(kd.identifier_minor, i, sl));
        gminor.push_back(trimmed(kd.gas_minor, i, sl));
    }
    auto gas_of_ident = [&](const std::string& id) -> std::string {
        auto it = std::find(ident.begin(), ident.end(), id);
        if (it == ident.end()) throw ArgError("k-distribution: unknown minor identifier " + id);
        return gminor[it - ident.begin()];
    };
    for (int side = 0; side < 2; ++side)
    {
        MinorSide& ms = p.minor[side];
        const int nfile = side == 0 ? kd.n_minor_lower : kd.n_minor_upper;
        const int nkfile = side == 0 ? kd.n_contrib_lower : kd.n_contrib_upper;
        const char* names = side == 0 ? kd.minor_gases_lower : kd.minor_gases_upper;
        const char* scal = side == 0 ? kd.scaling_gas_lower : kd.scaling_gas_upper;
        const int* lim = side == 0 ? kd.minor_limits_gpt_lower : kd.minor_limits_gpt_upper;
        const int* dens = side == 0 ? kd.minor_scales_with_density_lower : kd.minor_scales_with_density_upper;
        const int* comp = side == 0 ? kd.scale_by_complement_lower : kd.scale_by_complement_upper;
        const int* kst = side == 0 ? kd.kminor_start_lower : kd.kminor_start_upper;
        const double* kmin = side == 0 ? kd.kminor_lower : kd.kminor_upper;   // (temperature, eta, contributors)
        ms.by_band.assign(nbnd, {});
        std::vector<int> src_start;
        int n_elim = 0;
        for (int i = 0; i < nfile; ++i)
        {
            const std::string id = trimmed(names, i, sl);
            const std::string gas = gas_of_ident(id);
            const int g1 = lim[2 * i], g2 = lim[2 * i + 1], ng = g2 - g1 + 1;
            if (!is_avail(gas)) { n_elim += ng; continue; }
            MinorDesc d;
            d.idx_minor = find_gas(gas);
            d.idx_scaling = find_gas(trimmed(scal, i, sl));
            d.scales_with_density = dens[i] != 0; d.scale_by_complement = comp[i] != 0;
            d.contributor = ms.n;
            const int b = p.gpoint_bands[g1 - 1] - 1;
            if (p.band_lims_gpt[2 * b] != g1 || p.band_lims_gpt[2 * b + 1] != g2)
                throw ArgError("k-distribution: minor contributor g-point range is not a whole band (unsupported)");
            ms.by_band[b].push_back(ms.n);
            if ((int)ms.by_band[b].size() > MAX_MINOR) throw ArgError("k-distribution: too many minor contributors in one band");
            ms.desc.push_back(d);
            ms.limits.push_back(g1); ms.limits.push_back(g2);
            ms.kstart.push_back(kst[i] - n_elim);
            src_start.push_back(kst[i] - 1);
            ms.n++;
        }
        ms.nk = 0;
        for (int m = 0; m < ms.n; ++m) ms.nk += ms.limits[2 * m + 1] - ms.limits[2 * m] + 1;
        ms.kminor_ref.assign((size_t)ntemp * neta * std::max(ms.nk, 1), 0.0);
        ms.table.assign((size_t)std::max(ms.n, 1) * ntemp * neta * gw, 0.0);
        // device blocks in band order, so a band's contributors are consecutive blocks
        {
            int pos = 0;
            for (int b = 0; b < nbnd; ++b)
                for (int m : ms.by_band[b]) ms.desc[m].contributor = pos++;
        }
        ms.uid.assign(ms.n, 0);
        for (int m = 0; m < ms.n; ++m)
        {
            const MinorDesc& d = ms.desc[m];
            int u = -1;
            for (size_t k = 0; k < ms.uniq_rep.size() && u < 0; ++k)
            {
                const MinorDesc& e = ms.desc[ms.uniq_rep[k]];
                if (e.idx_minor == d.idx_minor && e.idx_scaling == d.idx_scaling && e.scales_with_density == d.scales_with_density &&
                    e.scale_by_complement == d.scale_by_complement) u = (int)k;
            }
            if (u < 0) { u = (int)ms.uniq_rep.size(); ms.uniq_rep.push_back(m); }
            ms.uid[m] = u;
        }
        if (ms.uniq_rep.size() > 63) throw ArgError("k-distribution: more than 63 distinct minor-gas scalings on one side (unsupported)");
        p.nuniq = std::max<int>(p.nuniq, std::max<int>((int)ms.uniq_rep.size(), 1));
        ms.scal_off.assign(nbnd, 0);
        for (int b = 0; b < nbnd; ++b)
        {
            ms.scal_off[b] = (int)ms.slot_desc.size();
            for (int m : ms.by_band[b]) ms.slot_desc.push_back(m);
            while (ms.slot_desc.size() % 4) ms.slot_desc.push_back(-1);
        }
        p.nscal = std::max<int>(p.nscal, std::max<int>((int)ms.slot_desc.size(), 4));
        for (int m = 0; m < ms.n; ++m)
        {
            const int ng = ms.limits[2 * m + 1] - ms.limits[2 * m] + 1;
            for (int it = 0; it < ntemp; ++it)
                for (int ie = 0; ie < neta; ++ie)
                    for (int g = 0; g < ng; ++g)
                    {
                        const double v = kmin[((size_t)it * neta + ie) * nkfile + src_start[m] + g];
                        ms.kminor_ref[it + (size_t)ntemp * (ie + (size_t)neta * (ms.kstart[m] - 1 + g))] = v;
                        ms.table[(((size_t)ms.desc[m].contributor * ntemp + it) * neta + ie) * gw + g] = v;
                    }
        }
    }
    return p;
}

}  // namespace rb200

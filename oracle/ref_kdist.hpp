/*
 * ORACLE — TEST INFRASTRUCTURE ONLY.
 * Builds the reference's own k-distribution object (CPU class Gas_optics_rrtmgp or its CUDA twin
 * Gas_optics_rrtmgp_gpu, both unmodified) from the file-variable struct, the way the reference's
 * driver does with the netCDF variables (src_test/Radiation_solver.cpp:70-284), same Array dims.
 * Shared by ref_shim.cpp (CPU) and refcuda_shim.cu (CUDA).
 */
#pragma once
#include <array>
#include <memory>
#include <string>
#include <vector>
#include "Array.h"
#include "Gas_concs.h"
#include "Gas_optics_rrtmgp.h"
#include "rrtmgp_b200_tables.h"

namespace refbuild
{
    inline std::vector<std::string> strings(const char* block, int n, int len, bool trim)
    {
        std::vector<std::string> out;
        for (int i = 0; i < n; ++i)
        {
            std::string s(block + (size_t)i * len, len);
            if (trim)
            {
                size_t e = s.find_last_not_of(" \0", std::string::npos, 2);
                s = (e == std::string::npos) ? "" : s.substr(0, e + 1);
                size_t b = s.find_first_not_of(' ');
                s = (b == std::string::npos) ? "" : s.substr(b);
            }
            out.push_back(s);
        }
        return out;
    }

    // host doubles -> Array<T,N> (T = Float of the build: double, or float with RTE_USE_SP)
    template<typename T, int N, typename S>
    Array<T, N> arr(const S* p, std::array<int, N> dims)
    {
        size_t n = 1;
        for (int d : dims) n *= d;
        std::vector<T> v(n);
        for (size_t i = 0; i < n; ++i) v[i] = (T)p[i];
        return Array<T, N>(std::move(v), dims);
    }

    inline Array<Bool, 1> barr(const int* p, int n)
    {
        std::vector<Bool> v(n);
        for (int i = 0; i < n; ++i) v[i] = (Bool)(p[i] != 0);
        return Array<Bool, 1>(std::move(v), {n});
    }

    // KD = Gas_optics_rrtmgp with GC = Gas_concs, or KD = Gas_optics_rrtmgp_gpu with GC = Gas_concs_gpu
    template<typename KD, typename GC>
    std::unique_ptr<KD> build_kdist(const rrtmgp_b200_kdist_file* f, const GC& avail)
    {
        std::unique_ptr<KD> kdist;
    const int L = f->string_len;
    Array<std::string,1> gas_names(strings(f->gas_names, f->n_absorbers, L, true), {f->n_absorbers});
    auto key_species = arr<int,3>(f->key_species, {2, f->n_layers, f->n_bnds});
    auto band_lims = arr<Float,2>(f->bnd_limits_wavenumber, {2, f->n_bnds});
    auto band2gpt = arr<int,2>(f->bnd_limits_gpt, {2, f->n_bnds});
    auto press_ref = arr<Float,1>(f->press_ref, {f->n_press});
    auto temp_ref = arr<Float,1>(f->temp_ref, {f->n_temps});
    auto kminor_lower = arr<Float,3>(f->kminor_lower, {f->n_contrib_lower, f->n_mixingfracs, f->n_temps});
    auto kminor_upper = arr<Float,3>(f->kminor_upper, {f->n_contrib_upper, f->n_mixingfracs, f->n_temps});
    Array<std::string,1> gas_minor(strings(f->gas_minor, f->n_minorabsorbers, L, true), {f->n_minorabsorbers});
    Array<std::string,1> identifier_minor(strings(f->identifier_minor, f->n_minorabsorbers, L, true), {f->n_minorabsorbers});
    Array<std::string,1> minor_gases_lower(strings(f->minor_gases_lower, f->n_minor_lower, L, true), {f->n_minor_lower});
    Array<std::string,1> minor_gases_upper(strings(f->minor_gases_upper, f->n_minor_upper, L, true), {f->n_minor_upper});
    auto minor_limits_gpt_lower = arr<int,2>(f->minor_limits_gpt_lower, {f->n_pairs, f->n_minor_lower});
    auto minor_limits_gpt_upper = arr<int,2>(f->minor_limits_gpt_upper, {f->n_pairs, f->n_minor_upper});
    auto dens_lower = barr(f->minor_scales_with_density_lower, f->n_minor_lower);
    auto dens_upper = barr(f->minor_scales_with_density_upper, f->n_minor_upper);
    auto comp_lower = barr(f->scale_by_complement_lower, f->n_minor_lower);
    auto comp_upper = barr(f->scale_by_complement_upper, f->n_minor_upper);
    Array<std::string,1> scaling_gas_lower(strings(f->scaling_gas_lower, f->n_minor_lower, L, true), {f->n_minor_lower});
    Array<std::string,1> scaling_gas_upper(strings(f->scaling_gas_upper, f->n_minor_upper, L, true), {f->n_minor_upper});
    auto kminor_start_lower = arr<int,1>(f->kminor_start_lower, {f->n_minor_lower});
    auto kminor_start_upper = arr<int,1>(f->kminor_start_upper, {f->n_minor_upper});
    auto vmr_ref = arr<Float,3>(f->vmr_ref, {f->n_layers, f->n_extabsorbers, f->n_temps});
    auto kmajor = arr<Float,4>(f->kmajor, {f->n_gpts, f->n_mixingfracs, f->n_press + 1, f->n_temps});

    Array<Float,3> rayl_lower, rayl_upper;
    if (f->rayl_lower)
    {
        rayl_lower = arr<Float,3>(f->rayl_lower, {f->n_gpts, f->n_mixingfracs, f->n_temps});
        rayl_upper = arr<Float,3>(f->rayl_upper, {f->n_gpts, f->n_mixingfracs, f->n_temps});
    }

    if (f->totplnk)
    {
        auto totplnk = arr<Float,2>(f->totplnk, {f->n_planck_temps, f->n_bnds});
        auto planck_frac = arr<Float,4>(f->plank_fraction, {f->n_gpts, f->n_mixingfracs, f->n_press + 1, f->n_temps});
        kdist = std::make_unique<KD>(
                avail, gas_names, key_species, band2gpt, band_lims, press_ref, (Float)f->press_ref_trop,
                temp_ref, (Float)f->absorption_coefficient_ref_P, (Float)f->absorption_coefficient_ref_T, vmr_ref,
                kmajor, kminor_lower, kminor_upper, gas_minor, identifier_minor,
                minor_gases_lower, minor_gases_upper, minor_limits_gpt_lower, minor_limits_gpt_upper,
                dens_lower, dens_upper, scaling_gas_lower, scaling_gas_upper, comp_lower, comp_upper,
                kminor_start_lower, kminor_start_upper, totplnk, planck_frac, rayl_lower, rayl_upper);
    }
    else
    {
        auto sq = arr<Float,1>(f->solar_source_quiet, {f->n_gpts});
        auto sf = arr<Float,1>(f->solar_source_facular, {f->n_gpts});
        auto ss = arr<Float,1>(f->solar_source_sunspot, {f->n_gpts});
        kdist = std::make_unique<KD>(
                avail, gas_names, key_species, band2gpt, band_lims, press_ref, (Float)f->press_ref_trop,
                temp_ref, (Float)f->absorption_coefficient_ref_P, (Float)f->absorption_coefficient_ref_T, vmr_ref,
                kmajor, kminor_lower, kminor_upper, gas_minor, identifier_minor,
                minor_gases_lower, minor_gases_upper, minor_limits_gpt_lower, minor_limits_gpt_upper,
                dens_lower, dens_upper, scaling_gas_lower, scaling_gas_upper, comp_lower, comp_upper,
                kminor_start_lower, kminor_start_upper, sq, sf, ss,
                (Float)f->tsi_default, (Float)f->mg_default, (Float)f->sb_default, rayl_lower, rayl_upper);
    }
        return kdist;
    }
}

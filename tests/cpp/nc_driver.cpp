// nc_driver.cpp — checks include/rrtmgp_b200/coefficients.hpp (+ netcdf_classic.hpp) from pytest:
// load a coefficient file and compare every field with the struct the Python side packed from the
// same tables (rte_rrtmgp_cpp_b200/ctables.py).  Host-only; returns 0 when identical, else the
// 1-based index of the first differing field (-1: exception, message in nc_last_error()).
#include <cstring>
#include <string>
#include "rrtmgp_b200/coefficients.hpp"

static std::string g_err;
extern "C" const char* nc_last_error() { return g_err.c_str(); }

namespace {
template<typename T> bool same(const T* a, const T* b, size_t n)
{
    if ((a == nullptr) != (b == nullptr)) return false;
    return a == nullptr || std::memcmp(a, b, n * sizeof(T)) == 0;
}
}

extern "C" int nc_kdist_equals(const char* path, const rrtmgp_b200_kdist_file* r)
{
    try
    {
        rrtmgp_b200::Kdist_tables t(path);
        const rrtmgp_b200_kdist_file& f = t.file();
        int k = 0;
#define CHECK(cond) do { ++k; if (!(cond)) return k; } while (0)
        CHECK(f.n_temps == r->n_temps); CHECK(f.n_press == r->n_press); CHECK(f.n_absorbers == r->n_absorbers);
        CHECK(f.n_minorabsorbers == r->n_minorabsorbers); CHECK(f.n_extabsorbers == r->n_extabsorbers);
        CHECK(f.n_mixingfracs == r->n_mixingfracs); CHECK(f.n_layers == r->n_layers); CHECK(f.n_bnds == r->n_bnds);
        CHECK(f.n_gpts == r->n_gpts); CHECK(f.n_pairs == r->n_pairs); CHECK(f.n_minor_lower == r->n_minor_lower);
        CHECK(f.n_minor_upper == r->n_minor_upper); CHECK(f.n_contrib_lower == r->n_contrib_lower);
        CHECK(f.n_contrib_upper == r->n_contrib_upper); CHECK(f.n_planck_temps == r->n_planck_temps);
        CHECK(f.string_len == r->string_len);
        const size_t sl = f.string_len, nT = f.n_temps, nP = f.n_press, nE = f.n_mixingfracs, nG = f.n_gpts, nB = f.n_bnds;
        CHECK(same(f.gas_names, r->gas_names, f.n_absorbers * sl));
        CHECK(same(f.key_species, r->key_species, nB * f.n_layers * 2));
        CHECK(same(f.bnd_limits_wavenumber, r->bnd_limits_wavenumber, nB * 2));
        CHECK(same(f.bnd_limits_gpt, r->bnd_limits_gpt, nB * 2));
        CHECK(same(f.press_ref, r->press_ref, nP)); CHECK(same(f.temp_ref, r->temp_ref, nT));
        CHECK(f.absorption_coefficient_ref_P == r->absorption_coefficient_ref_P);
        CHECK(f.absorption_coefficient_ref_T == r->absorption_coefficient_ref_T);
        CHECK(f.press_ref_trop == r->press_ref_trop);
        CHECK(same(f.vmr_ref, r->vmr_ref, nT * f.n_extabsorbers * f.n_layers));
        CHECK(same(f.kmajor, r->kmajor, nT * (nP + 1) * nE * nG));
        CHECK(same(f.kminor_lower, r->kminor_lower, nT * nE * f.n_contrib_lower));
        CHECK(same(f.kminor_upper, r->kminor_upper, nT * nE * f.n_contrib_upper));
        CHECK(same(f.gas_minor, r->gas_minor, f.n_minorabsorbers * sl));
        CHECK(same(f.identifier_minor, r->identifier_minor, f.n_minorabsorbers * sl));
        CHECK(same(f.minor_gases_lower, r->minor_gases_lower, f.n_minor_lower * sl));
        CHECK(same(f.minor_gases_upper, r->minor_gases_upper, f.n_minor_upper * sl));
        CHECK(same(f.minor_limits_gpt_lower, r->minor_limits_gpt_lower, (size_t)f.n_minor_lower * 2));
        CHECK(same(f.minor_limits_gpt_upper, r->minor_limits_gpt_upper, (size_t)f.n_minor_upper * 2));
        CHECK(same(f.minor_scales_with_density_lower, r->minor_scales_with_density_lower, (size_t)f.n_minor_lower));
        CHECK(same(f.minor_scales_with_density_upper, r->minor_scales_with_density_upper, (size_t)f.n_minor_upper));
        CHECK(same(f.scale_by_complement_lower, r->scale_by_complement_lower, (size_t)f.n_minor_lower));
        CHECK(same(f.scale_by_complement_upper, r->scale_by_complement_upper, (size_t)f.n_minor_upper));
        CHECK(same(f.scaling_gas_lower, r->scaling_gas_lower, f.n_minor_lower * sl));
        CHECK(same(f.scaling_gas_upper, r->scaling_gas_upper, f.n_minor_upper * sl));
        CHECK(same(f.kminor_start_lower, r->kminor_start_lower, (size_t)f.n_minor_lower));
        CHECK(same(f.kminor_start_upper, r->kminor_start_upper, (size_t)f.n_minor_upper));
        CHECK(same(f.totplnk, r->totplnk, nB * f.n_planck_temps));
        CHECK(same(f.plank_fraction, r->plank_fraction, nT * (nP + 1) * nE * nG));
        CHECK(same(f.rayl_lower, r->rayl_lower, nT * nE * nG)); CHECK(same(f.rayl_upper, r->rayl_upper, nT * nE * nG));
        CHECK(same(f.solar_source_quiet, r->solar_source_quiet, nG));
        CHECK(same(f.solar_source_facular, r->solar_source_facular, nG));
        CHECK(same(f.solar_source_sunspot, r->solar_source_sunspot, nG));
        CHECK(f.tsi_default == r->tsi_default); CHECK(f.mg_default == r->mg_default); CHECK(f.sb_default == r->sb_default);
        return 0;
    }
    catch (const std::exception& e) { g_err = e.what(); return -1; }
}

extern "C" int nc_cloud_equals(const char* path, const rrtmgp_b200_cloud_file* r)
{
    try
    {
        rrtmgp_b200::Cloud_tables t(path);
        const rrtmgp_b200_cloud_file& f = t.file();
        int k = 0;
        CHECK(f.n_band == r->n_band); CHECK(f.n_rghice == r->n_rghice); CHECK(f.n_size_liq == r->n_size_liq);
        CHECK(f.n_size_ice == r->n_size_ice);
        CHECK(same(f.bnd_limits_wavenumber, r->bnd_limits_wavenumber, (size_t)f.n_band * 2));
        CHECK(f.radliq_lwr == r->radliq_lwr); CHECK(f.radliq_upr == r->radliq_upr); CHECK(f.radliq_fac == r->radliq_fac);
        CHECK(f.diamice_lwr == r->diamice_lwr); CHECK(f.diamice_upr == r->diamice_upr); CHECK(f.diamice_fac == r->diamice_fac);
        const size_t nl = (size_t)f.n_band * f.n_size_liq, ni = (size_t)f.n_rghice * f.n_band * f.n_size_ice;
        CHECK(same(f.lut_extliq, r->lut_extliq, nl)); CHECK(same(f.lut_ssaliq, r->lut_ssaliq, nl)); CHECK(same(f.lut_asyliq, r->lut_asyliq, nl));
        CHECK(same(f.lut_extice, r->lut_extice, ni)); CHECK(same(f.lut_ssaice, r->lut_ssaice, ni)); CHECK(same(f.lut_asyice, r->lut_asyice, ni));
        return 0;
    }
    catch (const std::exception& e) { g_err = e.what(); return -1; }
}

extern "C" int nc_aerosol_equals(const char* path, const rrtmgp_b200_aerosol_file* r)
{
    try
    {
        rrtmgp_b200::Aerosol_tables t(path);
        const rrtmgp_b200_aerosol_file& f = t.file();
        int k = 0;
        CHECK(f.n_band == r->n_band); CHECK(f.n_hum == r->n_hum); CHECK(f.n_philic == r->n_philic); CHECK(f.n_phobic == r->n_phobic);
        const size_t nb = (size_t)f.n_phobic * f.n_band, nl = (size_t)f.n_philic * f.n_hum * f.n_band;
        CHECK(same(f.mass_ext_sw_hydrophobic, r->mass_ext_sw_hydrophobic, nb)); CHECK(same(f.ssa_sw_hydrophobic, r->ssa_sw_hydrophobic, nb));
        CHECK(same(f.asymmetry_sw_hydrophobic, r->asymmetry_sw_hydrophobic, nb));
        CHECK(same(f.mass_ext_sw_hydrophilic, r->mass_ext_sw_hydrophilic, nl)); CHECK(same(f.ssa_sw_hydrophilic, r->ssa_sw_hydrophilic, nl));
        CHECK(same(f.asymmetry_sw_hydrophilic, r->asymmetry_sw_hydrophilic, nl));
        CHECK(same(f.relative_humidity2, r->relative_humidity2, (size_t)f.n_hum));
        return 0;
    }
    catch (const std::exception& e) { g_err = e.what(); return -1; }
}

// writer check: a file with one variable of each supported type, read back by scipy in the test
extern "C" int nc_write_sample(const char* path, int nx, int nlev)
{
    try
    {
        rrtmgp_b200::Netcdf_file dst(path, rrtmgp_b200::Netcdf_mode::Create);
        dst.add_dimension("x", nx); dst.add_dimension("lev", nlev); dst.add_dimension("pair", 2);
        std::vector<double> d((size_t)nx * nlev); std::vector<float> f(d.size()); std::vector<int> i2((size_t)nx * 2);
        for (size_t k = 0; k < d.size(); ++k) { d[k] = 0.125 * (double)k - 3.; f[k] = (float)(1.5 * (double)k); }
        for (size_t k = 0; k < i2.size(); ++k) i2[k] = (int)k * 7 - 5;
        dst.add_variable<double>("flux_d", {"lev", "x"}).insert(d, {0, 0});
        dst.add_variable<float>("flux_f", {"lev", "x"}).insert(f, {0, 0});
        dst.add_variable<int>("lims", {"x", "pair"}).insert(i2, {0, 0});
        dst.add_variable<double>("scalar").insert(42.5);
        dst.add_variable<float>("never_inserted", {"x"});
        dst.sync();
        // and read it back through the reader, converting types on the way
        rrtmgp_b200::Netcdf_file src(path, rrtmgp_b200::Netcdf_mode::Read);
        const std::vector<float> back = src.get_variable<float>("flux_d", {nlev, nx});
        for (size_t k = 0; k < d.size(); ++k) if (back[k] != (float)d[k]) return 1;
        if (src.get_variable<int>("scalar") != 42) return 2;
        if (src.get_dimension_size("lev") != nlev || src.get_variable_dimensions("lims").at("pair") != 2) return 3;
        try { src.get_variable<double>("flux_d", {nx, nlev}); if (nx != nlev) return 4; } catch (const std::runtime_error&) {}
        try { src.get_variable<double>("absent", {nx}); return 5; } catch (const std::runtime_error&) {}
        return 0;
    }
    catch (const std::exception& e) { g_err = e.what(); return -1; }
}

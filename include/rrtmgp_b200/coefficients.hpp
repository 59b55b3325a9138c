// coefficients.hpp — read rrtmgp-gas-{lw,sw}-g*.nc / rrtmgp-clouds-{lw,sw}.nc (netCDF classic, see
// netcdf_classic.hpp) into the plain-C table structs of include/rrtmgp_b200_tables.h.
//
// Same variables and dimension names as the reference's loaders read
// (src_test/Radiation_solver.cpp:78-243 gas optics, :286-326 cloud optics, :331-367 aerosol optics); the arrays stay in FILE
// order — the reordering the reference does on the host is part of rrtmgp_b200_gas_optics_create.
#pragma once
#include <deque>
#include <string>
#include <vector>
#include "../rrtmgp_b200_tables.h"
#include "netcdf_classic.hpp"

namespace rrtmgp_b200 {

// Owns the arrays a rrtmgp_b200_kdist_file points into.
class Kdist_tables
{
    public:
        explicit Kdist_tables(const std::string& file_name)
        {
            Netcdf_file nc(file_name, Netcdf_mode::Read);
            f.n_temps = nc.get_dimension_size("temperature");
            f.n_press = nc.get_dimension_size("pressure");
            f.n_absorbers = nc.get_dimension_size("absorber");
            f.n_minorabsorbers = nc.get_dimension_size("minor_absorber");
            f.n_extabsorbers = nc.get_dimension_size("absorber_ext");
            f.n_mixingfracs = nc.get_dimension_size("mixing_fraction");
            f.n_layers = nc.get_dimension_size("atmos_layer");
            f.n_bnds = nc.get_dimension_size("bnd");
            f.n_gpts = nc.get_dimension_size("gpt");
            f.n_pairs = nc.get_dimension_size("pair");
            f.n_minor_lower = nc.get_dimension_size("minor_absorber_intervals_lower");
            f.n_minor_upper = nc.get_dimension_size("minor_absorber_intervals_upper");
            f.n_contrib_lower = nc.get_dimension_size("contributors_lower");
            f.n_contrib_upper = nc.get_dimension_size("contributors_upper");
            f.string_len = nc.get_dimension_size("string_len");
            const int nT = f.n_temps, nP = f.n_press, nE = f.n_mixingfracs, nG = f.n_gpts, nB = f.n_bnds, sl = f.string_len;

            f.gas_names = chars(nc, "gas_names", {f.n_absorbers, sl});
            f.key_species = ints(nc, "key_species", {nB, f.n_layers, 2});
            f.bnd_limits_wavenumber = dbls(nc, "bnd_limits_wavenumber", {nB, 2});
            f.bnd_limits_gpt = ints(nc, "bnd_limits_gpt", {nB, 2});
            f.press_ref = dbls(nc, "press_ref", {nP});
            f.temp_ref = dbls(nc, "temp_ref", {nT});
            f.absorption_coefficient_ref_P = nc.get_variable<double>("absorption_coefficient_ref_P");
            f.absorption_coefficient_ref_T = nc.get_variable<double>("absorption_coefficient_ref_T");
            f.press_ref_trop = nc.get_variable<double>("press_ref_trop");
            f.vmr_ref = dbls(nc, "vmr_ref", {nT, f.n_extabsorbers, f.n_layers});
            f.kmajor = dbls(nc, "kmajor", {nT, nP + 1, nE, nG});
            f.kminor_lower = dbls(nc, "kminor_lower", {nT, nE, f.n_contrib_lower});
            f.kminor_upper = dbls(nc, "kminor_upper", {nT, nE, f.n_contrib_upper});
            f.gas_minor = chars(nc, "gas_minor", {f.n_minorabsorbers, sl});
            f.identifier_minor = chars(nc, "identifier_minor", {f.n_minorabsorbers, sl});
            f.minor_gases_lower = chars(nc, "minor_gases_lower", {f.n_minor_lower, sl});
            f.minor_gases_upper = chars(nc, "minor_gases_upper", {f.n_minor_upper, sl});
            f.minor_limits_gpt_lower = ints(nc, "minor_limits_gpt_lower", {f.n_minor_lower, 2});
            f.minor_limits_gpt_upper = ints(nc, "minor_limits_gpt_upper", {f.n_minor_upper, 2});
            f.minor_scales_with_density_lower = ints(nc, "minor_scales_with_density_lower", {f.n_minor_lower});
            f.minor_scales_with_density_upper = ints(nc, "minor_scales_with_density_upper", {f.n_minor_upper});
            f.scale_by_complement_lower = ints(nc, "scale_by_complement_lower", {f.n_minor_lower});
            f.scale_by_complement_upper = ints(nc, "scale_by_complement_upper", {f.n_minor_upper});
            f.scaling_gas_lower = chars(nc, "scaling_gas_lower", {f.n_minor_lower, sl});
            f.scaling_gas_upper = chars(nc, "scaling_gas_upper", {f.n_minor_upper, sl});
            f.kminor_start_lower = ints(nc, "kminor_start_lower", {f.n_minor_lower});
            f.kminor_start_upper = ints(nc, "kminor_start_upper", {f.n_minor_upper});

            if (nc.variable_exists("totplnk"))                 // longwave file (src_test/Radiation_solver.cpp:193)
            {
                f.n_planck_temps = nc.get_dimension_size("temperature_Planck");
                f.totplnk = dbls(nc, "totplnk", {nB, f.n_planck_temps});
                f.plank_fraction = dbls(nc, "plank_fraction", {nT, nP + 1, nE, nG});
            }
            else if (nc.variable_exists("rayl_lower"))         // shortwave file (:145, :220-243)
            {
                f.rayl_lower = dbls(nc, "rayl_lower", {nT, nE, nG});
                f.rayl_upper = dbls(nc, "rayl_upper", {nT, nE, nG});
                f.solar_source_quiet = dbls(nc, "solar_source_quiet", {nG});
                f.solar_source_facular = dbls(nc, "solar_source_facular", {nG});
                f.solar_source_sunspot = dbls(nc, "solar_source_sunspot", {nG});
                f.tsi_default = nc.get_variable<double>("tsi_default");
                f.mg_default = nc.get_variable<double>("mg_default");
                f.sb_default = nc.get_variable<double>("sb_default");
            }
            else
                throw std::runtime_error(file_name + " holds neither totplnk (longwave) nor rayl_lower (shortwave)");
        }
        Kdist_tables(const Kdist_tables&) = delete;
        const rrtmgp_b200_kdist_file& file() const { return f; }

    private:
        const double* dbls(const Netcdf_file& nc, const char* n, const std::vector<int>& s) { d.push_back(nc.get_variable<double>(n, s)); return d.back().data(); }
        const int* ints(const Netcdf_file& nc, const char* n, const std::vector<int>& s) { i.push_back(nc.get_variable<int>(n, s)); return i.back().data(); }
        const char* chars(const Netcdf_file& nc, const char* n, const std::vector<int>& s) { c.push_back(nc.get_variable<char>(n, s)); return c.back().data(); }
        rrtmgp_b200_kdist_file f{};
        std::deque<std::vector<double>> d; std::deque<std::vector<int>> i; std::deque<std::vector<char>> c;   // deque: stable addresses
};

class Cloud_tables
{
    public:
        explicit Cloud_tables(const std::string& file_name)
        {
            Netcdf_file nc(file_name, Netcdf_mode::Read);
            f.n_band = nc.get_dimension_size("nband");
            f.n_rghice = nc.get_dimension_size("nrghice");
            f.n_size_liq = nc.get_dimension_size("nsize_liq");
            f.n_size_ice = nc.get_dimension_size("nsize_ice");
            f.bnd_limits_wavenumber = dbls(nc, "bnd_limits_wavenumber", {f.n_band, 2});
            f.radliq_lwr = nc.get_variable<double>("radliq_lwr"); f.radliq_upr = nc.get_variable<double>("radliq_upr");
            f.radliq_fac = nc.get_variable<double>("radliq_fac");
            f.diamice_lwr = nc.get_variable<double>("diamice_lwr"); f.diamice_upr = nc.get_variable<double>("diamice_upr");
            f.diamice_fac = nc.get_variable<double>("diamice_fac");
            f.lut_extliq = dbls(nc, "lut_extliq", {f.n_band, f.n_size_liq});
            f.lut_ssaliq = dbls(nc, "lut_ssaliq", {f.n_band, f.n_size_liq});
            f.lut_asyliq = dbls(nc, "lut_asyliq", {f.n_band, f.n_size_liq});
            f.lut_extice = dbls(nc, "lut_extice", {f.n_rghice, f.n_band, f.n_size_ice});
            f.lut_ssaice = dbls(nc, "lut_ssaice", {f.n_rghice, f.n_band, f.n_size_ice});
            f.lut_asyice = dbls(nc, "lut_asyice", {f.n_rghice, f.n_band, f.n_size_ice});
        }
        Cloud_tables(const Cloud_tables&) = delete;
        const rrtmgp_b200_cloud_file& file() const { return f; }
    private:
        const double* dbls(const Netcdf_file& nc, const char* n, const std::vector<int>& s) { d.push_back(nc.get_variable<double>(n, s)); return d.back().data(); }
        rrtmgp_b200_cloud_file f{};
        std::deque<std::vector<double>> d;
};

// aerosol_optics.nc (src_test/Radiation_solver.cpp:331-367)
class Aerosol_tables
{
    public:
        explicit Aerosol_tables(const std::string& file_name)
        {
            Netcdf_file nc(file_name, Netcdf_mode::Read);
            f.n_band = nc.get_dimension_size("band_sw");
            f.n_hum = nc.get_dimension_size("relative_humidity");
            f.n_philic = nc.get_dimension_size("hydrophilic");
            f.n_phobic = nc.get_dimension_size("hydrophobic");
            f.mass_ext_sw_hydrophobic = dbls(nc, "mass_ext_sw_hydrophobic", {f.n_phobic, f.n_band});
            f.ssa_sw_hydrophobic = dbls(nc, "ssa_sw_hydrophobic", {f.n_phobic, f.n_band});
            f.asymmetry_sw_hydrophobic = dbls(nc, "asymmetry_sw_hydrophobic", {f.n_phobic, f.n_band});
            f.mass_ext_sw_hydrophilic = dbls(nc, "mass_ext_sw_hydrophilic", {f.n_philic, f.n_hum, f.n_band});
            f.ssa_sw_hydrophilic = dbls(nc, "ssa_sw_hydrophilic", {f.n_philic, f.n_hum, f.n_band});
            f.asymmetry_sw_hydrophilic = dbls(nc, "asymmetry_sw_hydrophilic", {f.n_philic, f.n_hum, f.n_band});
            f.relative_humidity2 = dbls(nc, "relative_humidity2", {f.n_hum});
        }
        Aerosol_tables(const Aerosol_tables&) = delete;
        const rrtmgp_b200_aerosol_file& file() const { return f; }
    private:
        const double* dbls(const Netcdf_file& nc, const char* n, const std::vector<int>& s) { d.push_back(nc.get_variable<double>(n, s)); return d.back().data(); }
        rrtmgp_b200_aerosol_file f{};
        std::deque<std::vector<double>> d;
};

}  // namespace rrtmgp_b200

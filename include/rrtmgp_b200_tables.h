/*
 * rrtmgp_b200_tables.h — plain-C description of the coefficient files handed across the C ABI.
 *
 * The structs mirror, variable by variable and in FILE dimension order (C order, last index
 * fastest), what the reference reads from rrtmgp-gas-{lw,sw}-g*.nc and rrtmgp-clouds-{lw,sw}.nc
 * before constructing Gas_optics_rrtmgp / Cloud_optics
 * (reference: src_test/Radiation_solver.cpp:70-284 and :286-326).
 * Strings are fixed-length, blank padded char blocks [n][string_len] as netCDF stores them.
 * All floating-point tables are double (the files' precision); a plan converts to its own Real.
 */
#ifndef RRTMGP_B200_TABLES_H
#define RRTMGP_B200_TABLES_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rrtmgp_b200_kdist_file
{
    /* dimensions */
    int n_temps;            /* "temperature" */
    int n_press;            /* "pressure" (kmajor has n_press+1 = "pressure_interp") */
    int n_absorbers;        /* "absorber" */
    int n_minorabsorbers;   /* "minor_absorber" */
    int n_extabsorbers;     /* "absorber_ext" = n_absorbers+1 */
    int n_mixingfracs;      /* "mixing_fraction" (eta) */
    int n_layers;           /* "atmos_layer" = 2 (lower, upper) */
    int n_bnds;             /* "bnd" */
    int n_gpts;             /* "gpt" */
    int n_pairs;            /* "pair" = 2 */
    int n_minor_lower;      /* "minor_absorber_intervals_lower" */
    int n_minor_upper;      /* "minor_absorber_intervals_upper" */
    int n_contrib_lower;    /* "contributors_lower" */
    int n_contrib_upper;    /* "contributors_upper" */
    int n_planck_temps;     /* "temperature_Planck"; 0 for a shortwave file */
    int string_len;         /* "string_len" */

    const char*   gas_names;                 /* [n_absorbers][string_len] */
    const int*    key_species;               /* (bnd, atmos_layer, pair) */
    const double* bnd_limits_wavenumber;     /* (bnd, 2) */
    const int*    bnd_limits_gpt;            /* (bnd, 2), 1-based inclusive */
    const double* press_ref;                 /* (pressure) */
    const double* temp_ref;                  /* (temperature) */
    double        absorption_coefficient_ref_P;
    double        absorption_coefficient_ref_T;
    double        press_ref_trop;
    const double* vmr_ref;                   /* (temperature, absorber_ext, atmos_layer) */
    const double* kmajor;                    /* (temperature, pressure_interp, mixing_fraction, gpt) */
    const double* kminor_lower;              /* (temperature, mixing_fraction, contributors_lower) */
    const double* kminor_upper;
    const char*   gas_minor;                 /* [n_minorabsorbers][string_len] */
    const char*   identifier_minor;          /* [n_minorabsorbers][string_len] */
    const char*   minor_gases_lower;         /* [n_minor_lower][string_len] */
    const char*   minor_gases_upper;
    const int*    minor_limits_gpt_lower;    /* (n_minor_lower, pair) */
    const int*    minor_limits_gpt_upper;
    const int*    minor_scales_with_density_lower;  /* 0/1 */
    const int*    minor_scales_with_density_upper;
    const int*    scale_by_complement_lower;
    const int*    scale_by_complement_upper;
    const char*   scaling_gas_lower;         /* [n_minor_lower][string_len] */
    const char*   scaling_gas_upper;
    const int*    kminor_start_lower;        /* 1-based */
    const int*    kminor_start_upper;

    /* longwave files only (NULL otherwise) */
    const double* totplnk;                   /* (bnd, temperature_Planck) */
    const double* plank_fraction;            /* like kmajor */

    /* shortwave files only (NULL otherwise) */
    const double* rayl_lower;                /* (temperature, mixing_fraction, gpt) */
    const double* rayl_upper;
    const double* solar_source_quiet;        /* (gpt) */
    const double* solar_source_facular;
    const double* solar_source_sunspot;
    double        tsi_default;
    double        mg_default;
    double        sb_default;
} rrtmgp_b200_kdist_file;

typedef struct rrtmgp_b200_cloud_file
{
    int n_band;        /* "nband" */
    int n_rghice;      /* "nrghice" */
    int n_size_liq;    /* "nsize_liq" */
    int n_size_ice;    /* "nsize_ice" */
    const double* bnd_limits_wavenumber;   /* (nband, 2) */
    double radliq_lwr, radliq_upr, radliq_fac;
    double diamice_lwr, diamice_upr, diamice_fac;
    const double* lut_extliq;   /* (nband, nsize_liq) */
    const double* lut_ssaliq;
    const double* lut_asyliq;
    const double* lut_extice;   /* (nrghice, nband, nsize_ice) */
    const double* lut_ssaice;
    const double* lut_asyice;
} rrtmgp_b200_cloud_file;

/* aerosol_optics.nc as read by load_and_init_aerosol_optics (src_test/Radiation_solver.cpp:331-367,
 * src_test/Radiation_solver.cu:366-402): shortwave MERRA/IFS-style look-up tables.  Arrays are in
 * FILE order (C order, last dimension fastest), i.e. band fastest. */
typedef struct rrtmgp_b200_aerosol_file
{
    int n_band;        /* "band_sw" */
    int n_hum;         /* "relative_humidity" */
    int n_philic;      /* "hydrophilic" */
    int n_phobic;      /* "hydrophobic" */
    const double* mass_ext_sw_hydrophobic;    /* (hydrophobic, band_sw) */
    const double* ssa_sw_hydrophobic;
    const double* asymmetry_sw_hydrophobic;
    const double* mass_ext_sw_hydrophilic;    /* (hydrophilic, relative_humidity, band_sw) */
    const double* ssa_sw_hydrophilic;
    const double* asymmetry_sw_hydrophilic;
    const double* relative_humidity2;         /* (relative_humidity): upper bound of each class */
} rrtmgp_b200_aerosol_file;

#ifdef __cplusplus
}
#endif
#endif

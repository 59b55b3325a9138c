/*
 * ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path.
 *
 * Plain-C (fp64) restatement of the Fortran `bind(C)` numeric kernels that carry all the
 * arithmetic of the reference CPU path (microhh/rte-rrtmgp-cpp links them from the
 * rte-rrtmgp v1.7 submodule; there is no Fortran compiler in this environment).
 * Symbol names and argument lists are exactly those declared in the reference's
 * include/rrtmgp_kernels.h:34-288, so this file can be linked underneath the reference's own,
 * unmodified C++ class layer (see oracle/Makefile target _ref) as well as driven directly
 * through ctypes (oracle/oracle.py).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may
 * load this.  Every function cites the Fortran it follows (paths relative to /root/reference).
 *
 * Layout: all arrays are column-major (first index fastest), integer tables hold 1-based
 * Fortran indices, Bool is `signed char` (RTE_USE_CBOOL), Float is double.
 *
 * Parity pinning: the solver kernels are pinned by the analytic known-answer tests restated
 * from rte-rrtmgp/tests/rte_{lw,sw}_solver_unit_tests.F90 (tests/test_oracle_kat.py); the gas
 * optics kernels can only be pinned through the reference's own class layer (oracle/_ref)
 * because the reference's golden flux files need the rrtmgp-gas-*.nc coefficient files, which
 * are absent from the mount (/root/reference/.MISSING_LARGE_BLOBS).
 */
#include <math.h>
#include <float.h>
#include <stdlib.h>
#include <string.h>

typedef double Float;
typedef signed char Bool;

#define MIN(a, b) ((a) < (b) ? (a) : (b))
#define MAX(a, b) ((a) > (b) ? (a) : (b))

/* ---------------------------------------------------------------------------------------------
 * rte-rrtmgp/rte-kernels/mo_rte_util_array.F90:56-90 (zero_array_3D / _4D)
 * ------------------------------------------------------------------------------------------- */
void zero_array_3D(int* ni, int* nj, int* nk, Float* array)
{
    const size_t n = (size_t)(*ni) * (*nj) * (*nk);
    for (size_t i = 0; i < n; ++i) array[i] = 0.0;
}

void zero_array_4D(int* ni, int* nj, int* nk, int* nl, Float* array)
{
    const size_t n = (size_t)(*ni) * (*nj) * (*nk) * (*nl);
    for (size_t i = 0; i < n; ++i) array[i] = 0.0;
}

/* ---------------------------------------------------------------------------------------------
 * rte-rrtmgp/rte-kernels/mo_fluxes_broadband_kernels.F90:32-61 (sum_broadband): the g-point sum
 * is sequential igpt = 1..ngpt for each (icol, ilev).
 * ------------------------------------------------------------------------------------------- */
void rte_sum_broadband(int* ncol_, int* nlev_, int* ngpt_, Float* spectral_flux, Float* broadband_flux)
{
    const int ncol = *ncol_, nlev = *nlev_, ngpt = *ngpt_;
    for (int ilev = 0; ilev < nlev; ++ilev)
        for (int icol = 0; icol < ncol; ++icol)
        {
            Float s = 0.0;
            for (int igpt = 0; igpt < ngpt; ++igpt)
                s = s + spectral_flux[icol + (size_t)ncol * (ilev + (size_t)nlev * igpt)];
            broadband_flux[icol + (size_t)ncol * ilev] = s;
        }
}

/* mo_fluxes_broadband_kernels.F90:109-128 (net_broadband_precalc) */
void rte_net_broadband_precalc(int* ncol_, int* nlev_, Float* flux_dn, Float* flux_up, Float* flux_net)
{
    const size_t n = (size_t)(*ncol_) * (*nlev_);
    for (size_t i = 0; i < n; ++i) flux_net[i] = flux_dn[i] - flux_up[i];
}

/* src_kernels/mo_fluxes_byband_kernels.F90:22-41 (sum_byband); band_lims(2,nbnd) is 1-based. */
void sum_byband(int* ncol_, int* nlev_, int* ngpt_, int* nbnd_, int* band_lims,
                Float* spectral_flux, Float* byband_flux)
{
    const int ncol = *ncol_, nlev = *nlev_, nbnd = *nbnd_;
    (void)ngpt_;
    for (int ibnd = 0; ibnd < nbnd; ++ibnd)
        for (int ilev = 0; ilev < nlev; ++ilev)
            for (int icol = 0; icol < ncol; ++icol)
            {
                const int g1 = band_lims[0 + 2 * ibnd], g2 = band_lims[1 + 2 * ibnd];
                Float s = spectral_flux[icol + (size_t)ncol * (ilev + (size_t)nlev * (g1 - 1))];
                for (int igpt = g1 + 1; igpt <= g2; ++igpt)
                    s = s + spectral_flux[icol + (size_t)ncol * (ilev + (size_t)nlev * (igpt - 1))];
                byband_flux[icol + (size_t)ncol * (ilev + (size_t)nlev * ibnd)] = s;
            }
}

/* src_kernels/mo_fluxes_byband_kernels.F90:71-78 (net_byband_precalc) */
void net_byband_precalc(int* ncol_, int* nlev_, int* nbnd_, Float* dn, Float* up, Float* net)
{
    const size_t n = (size_t)(*ncol_) * (*nlev_) * (*nbnd_);
    for (size_t i = 0; i < n; ++i) net[i] = dn[i] - up[i];
}

/* ---------------------------------------------------------------------------------------------
 * rte-rrtmgp/rrtmgp-kernels/mo_gas_optics_rrtmgp_kernels.F90:37-172 (interpolation)
 *
 * flavor(2,nflav), vmr_ref(2,0:ngas,ntemp), col_gas(ncol,nlay,0:ngas),
 * jeta(2,ncol,nlay,nflav), col_mix(2,ncol,nlay,nflav), fminor(2,2,ncol,nlay,nflav),
 * fmajor(2,2,2,ncol,nlay,nflav).
 * ------------------------------------------------------------------------------------------- */
void rrtmgp_interpolation(
        int* ncol_, int* nlay_,
        int* ngas_, int* nflav_, int* neta_, int* npres_, int* ntemp_,
        int* flavor,
        Float* press_ref_log,
        Float* temp_ref,
        Float* press_ref_log_delta_,
        Float* temp_ref_min_,
        Float* temp_ref_delta_,
        Float* press_ref_trop_log_,
        Float* vmr_ref,
        Float* play,
        Float* tlay,
        Float* col_gas,
        int* jtemp,
        Float* fmajor, Float* fminor,
        Float* col_mix,
        Bool* tropo,
        int* jeta,
        int* jpress)
{
    const int ncol = *ncol_, nlay = *nlay_, ngas = *ngas_, nflav = *nflav_;
    const int neta = *neta_, npres = *npres_, ntemp = *ntemp_;
    const Float press_ref_log_delta = *press_ref_log_delta_;
    const Float temp_ref_min = *temp_ref_min_;
    const Float temp_ref_delta = *temp_ref_delta_;
    const Float press_ref_trop_log = *press_ref_trop_log_;
    const size_t ncl = (size_t)ncol * nlay;

    Float* ftemp = (Float*)malloc(sizeof(Float) * ncl);
    Float* fpress = (Float*)malloc(sizeof(Float) * ncl);

    /* F90:109-125 */
    for (int ilay = 0; ilay < nlay; ++ilay)
        for (int icol = 0; icol < ncol; ++icol)
        {
            const size_t i = icol + (size_t)ncol * ilay;
            int jt = (int)((tlay[i] - (temp_ref_min - temp_ref_delta)) / temp_ref_delta);
            jt = MIN(ntemp - 1, MAX(1, jt));
            jtemp[i] = jt;
            ftemp[i] = (tlay[i] - temp_ref[jt - 1]) / temp_ref_delta;

            const Float locpress = 1.0 + (log(play[i]) - press_ref_log[0]) / press_ref_log_delta;
            const int jp = MIN(npres - 1, MAX(1, (int)locpress));
            jpress[i] = jp;
            fpress[i] = locpress - (Float)jp;

            tropo[i] = (Bool)(log(play[i]) > press_ref_trop_log);
        }

    /* F90:127-169 */
    for (int iflav = 0; iflav < nflav; ++iflav)
    {
        const int ig1 = flavor[0 + 2 * iflav];
        const int ig2 = flavor[1 + 2 * iflav];
        for (int ilay = 0; ilay < nlay; ++ilay)
            for (int icol = 0; icol < ncol; ++icol)
            {
                const size_t i = icol + (size_t)ncol * ilay;
                const int itropo = tropo[i] ? 1 : 2;
                for (int itemp = 1; itemp <= 2; ++itemp)
                {
                    const int jt = jtemp[i] + itemp - 1;  /* 1-based temperature index */
                    /* vmr_ref(itropo, igas, jt): dims (2, ngas+1, ntemp), igas 0-based */
                    const Float v1 = vmr_ref[(itropo - 1) + 2 * (ig1 + (size_t)(ngas + 1) * (jt - 1))];
                    const Float v2 = vmr_ref[(itropo - 1) + 2 * (ig2 + (size_t)(ngas + 1) * (jt - 1))];
                    const Float ratio_eta_half = v1 / v2;
                    const Float cg1 = col_gas[i + ncl * ig1];
                    const Float cg2 = col_gas[i + ncl * ig2];
                    const size_t i2 = (itemp - 1) + 2 * (i + ncl * iflav);
                    const Float cm = cg1 + ratio_eta_half * cg2;
                    col_mix[i2] = cm;
                    Float eta;
                    if (cm > 2.0 * DBL_MIN)
                        eta = cg1 / cm;
                    else
                        eta = 0.5;
                    const Float loceta = eta * (Float)(neta - 1);
                    jeta[i2] = MIN((int)loceta + 1, neta - 1);
                    const Float feta = fmod(loceta, 1.0);
                    const Float ftemp_term = ((Float)(2 - itemp) + (Float)(2 * itemp - 3) * ftemp[i]);
                    /* fminor(2,2,ncol,nlay,nflav): index (ieta, itemp, icol, ilay, iflav) */
                    const size_t im = 2 * (size_t)(itemp - 1) + 4 * (i + ncl * iflav);
                    fminor[0 + im] = (1.0 - feta) * ftemp_term;
                    fminor[1 + im] = feta * ftemp_term;
                    /* fmajor(2,2,2,...): index (ieta, ipress, itemp, icol, ilay, iflav) */
                    const size_t iM = 4 * (size_t)(itemp - 1) + 8 * (i + ncl * iflav);
                    fmajor[0 + 2 * 0 + iM] = (1.0 - fpress[i]) * fminor[0 + im];
                    fmajor[1 + 2 * 0 + iM] = (1.0 - fpress[i]) * fminor[1 + im];
                    fmajor[0 + 2 * 1 + iM] = fpress[i] * fminor[0 + im];
                    fmajor[1 + 2 * 1 + iM] = fpress[i] * fminor[1 + im];
                }
            }
    }
    free(ftemp);
    free(fpress);
}

/* mo_gas_optics_rrtmgp_kernels.F90:760-789 (interpolate3D_byflav) for one g-point.
 * k(ntemp, neta, npres+1, ngpt); jtemp, jeta, jpress 1-based; igpt 0-based here. */
static inline Float interpolate3D(const Float* scaling, const Float* fmajor, const Float* k,
                                  int igpt, const int* jeta, int jtemp, int jpress,
                                  int ntemp, int neta, int npres1)
{
#define K4(it, ie, ip) k[((it) - 1) + (size_t)ntemp * (((ie) - 1) + (size_t)neta * (((ip) - 1) + (size_t)npres1 * igpt))]
#define FM(ie, ip, it) fmajor[((ie) - 1) + 2 * ((ip) - 1) + 4 * ((it) - 1)]
    return scaling[0] *
               (FM(1, 1, 1) * K4(jtemp, jeta[0], jpress - 1) +
                FM(2, 1, 1) * K4(jtemp, jeta[0] + 1, jpress - 1) +
                FM(1, 2, 1) * K4(jtemp, jeta[0], jpress) +
                FM(2, 2, 1) * K4(jtemp, jeta[0] + 1, jpress)) +
           scaling[1] *
               (FM(1, 1, 2) * K4(jtemp + 1, jeta[1], jpress - 1) +
                FM(2, 1, 2) * K4(jtemp + 1, jeta[1] + 1, jpress - 1) +
                FM(1, 2, 2) * K4(jtemp + 1, jeta[1], jpress) +
                FM(2, 2, 2) * K4(jtemp + 1, jeta[1] + 1, jpress));
#undef K4
#undef FM
}

/* mo_gas_optics_rrtmgp_kernels.F90:738-758 (interpolate2D_byflav) for one g-point.
 * k(ntemp, neta, nk); ik 0-based index into the third dimension. */
static inline Float interpolate2D(const Float* fminor, const Float* k, int ik,
                                  const int* jeta, int jtemp, int ntemp, int neta)
{
#define K3(it, ie) k[((it) - 1) + (size_t)ntemp * (((ie) - 1) + (size_t)neta * ik)]
    return fminor[0 + 2 * 0] * K3(jtemp, jeta[0]) +
           fminor[1 + 2 * 0] * K3(jtemp, jeta[0] + 1) +
           fminor[0 + 2 * 1] * K3(jtemp + 1, jeta[1]) +
           fminor[1 + 2 * 1] * K3(jtemp + 1, jeta[1] + 1);
#undef K3
}

/* mo_gas_optics_rrtmgp_kernels.F90:347-398 (gas_optical_depths_major) */
static void gas_optical_depths_major(
        int ncol, int nlay, int nbnd, int ngpt, int nflav, int neta, int npres, int ntemp,
        const int* gpoint_flavor, const int* band_lims_gpt, const Float* kmajor,
        const Float* col_mix, const Float* fmajor, const int* jeta, const Bool* tropo,
        const int* jtemp, const int* jpress, Float* tau)
{
    const size_t ncl = (size_t)ncol * nlay;
    (void)ngpt; (void)nflav;
    for (int ibnd = 0; ibnd < nbnd; ++ibnd)
    {
        const int gptS = band_lims_gpt[0 + 2 * ibnd];
        const int gptE = band_lims_gpt[1 + 2 * ibnd];
        for (int ilay = 0; ilay < nlay; ++ilay)
            for (int icol = 0; icol < ncol; ++icol)
            {
                const size_t i = icol + (size_t)ncol * ilay;
                const int itropo = tropo[i] ? 1 : 2;
                const int iflav = gpoint_flavor[(itropo - 1) + 2 * (gptS - 1)];  /* 1-based */
                const size_t ifl = i + ncl * (iflav - 1);
                for (int igpt = gptS; igpt <= gptE; ++igpt)
                {
                    const Float tau_major = interpolate3D(
                            &col_mix[2 * ifl], &fmajor[8 * ifl], kmajor, igpt - 1,
                            &jeta[2 * ifl], jtemp[i], jpress[i] + itropo, ntemp, neta, npres + 1);
                    tau[i + ncl * (igpt - 1)] = tau[i + ncl * (igpt - 1)] + tau_major;
                }
            }
    }
}

/* mo_gas_optics_rrtmgp_kernels.F90:404-502 (gas_optical_depths_minor).
 * gpt_flv is gpoint_flavor(idx_tropo,:) i.e. stride-2 into gpoint_flavor. layer_limits(ncol,2). */
static void gas_optical_depths_minor(
        int ncol, int nlay, int ngpt, int ngas, int nflav, int ntemp, int neta,
        int nminor, int nminork, int idx_h2o,
        const int* gpoint_flavor, int flv_row,
        const Float* kminor, const int* minor_limits_gpt,
        const Bool* minor_scales_with_density, const Bool* scale_by_complement,
        const int* idx_minor, const int* idx_minor_scaling, const int* kminor_start,
        const Float* play, const Float* tlay, const Float* col_gas, const Float* fminor,
        const int* jeta, const int* layer_limits, const int* jtemp, Float* tau)
{
    const Float PaTohPa = 0.01;
    const size_t ncl = (size_t)ncol * nlay;
    (void)ngpt; (void)ngas; (void)nflav; (void)nminork;

    int any = 0;
    for (int icol = 0; icol < ncol; ++icol)
        if (layer_limits[icol] > 0) any = 1;
    if (!any) return;

    for (int imnr = 0; imnr < nminor; ++imnr)
        for (int icol = 0; icol < ncol; ++icol)
        {
            if (layer_limits[icol] <= 0) continue;
            for (int ilay1 = layer_limits[icol]; ilay1 <= layer_limits[icol + ncol]; ++ilay1)
            {
                const size_t i = icol + (size_t)ncol * (ilay1 - 1);
                Float scaling = col_gas[i + ncl * idx_minor[imnr]];
                if (minor_scales_with_density[imnr])
                {
                    scaling = scaling * (PaTohPa * play[i] / tlay[i]);
                    if (idx_minor_scaling[imnr] > 0)
                    {
                        const Float vmr_fact = 1.0 / col_gas[i];
                        const Float dry_fact = 1.0 / (1.0 + col_gas[i + ncl * idx_h2o] * vmr_fact);
                        if (scale_by_complement[imnr])
                            scaling = scaling * (1.0 - col_gas[i + ncl * idx_minor_scaling[imnr]] * vmr_fact * dry_fact);
                        else
                            scaling = scaling * (col_gas[i + ncl * idx_minor_scaling[imnr]] * vmr_fact * dry_fact);
                    }
                }
                const int gptS = minor_limits_gpt[0 + 2 * imnr];
                const int gptE = minor_limits_gpt[1 + 2 * imnr];
                const int iflav = gpoint_flavor[flv_row + 2 * (gptS - 1)];
                const size_t ifl = i + ncl * (iflav - 1);
                for (int igpt = gptS; igpt <= gptE; ++igpt)
                {
                    const int ik = kminor_start[imnr] - 1 + (igpt - gptS);
                    const Float tau_minor = scaling * interpolate2D(
                            &fminor[4 * ifl], kminor, ik, &jeta[2 * ifl], jtemp[i], ntemp, neta);
                    tau[i + ncl * (igpt - 1)] = tau[i + ncl * (igpt - 1)] + tau_minor;
                }
            }
        }
}

/* Fortran minloc/maxloc(play, dim=2, mask=...) for one column: 1-based location of the first
 * extremum among masked elements, 0 if the mask is all false. */
static int minloc_masked(const Float* play, const Bool* tropo, int want, int ncol, int nlay, int icol)
{
    int loc = 0; Float best = 0.0;
    for (int ilay = 0; ilay < nlay; ++ilay)
    {
        const size_t i = icol + (size_t)ncol * ilay;
        if ((tropo[i] != 0) != (want != 0)) continue;
        if (loc == 0 || play[i] < best) { best = play[i]; loc = ilay + 1; }
    }
    return loc;
}

static int maxloc_masked(const Float* play, const Bool* tropo, int want, int ncol, int nlay, int icol)
{
    int loc = 0; Float best = 0.0;
    for (int ilay = 0; ilay < nlay; ++ilay)
    {
        const size_t i = icol + (size_t)ncol * ilay;
        if ((tropo[i] != 0) != (want != 0)) continue;
        if (loc == 0 || play[i] > best) { best = play[i]; loc = ilay + 1; }
    }
    return loc;
}

/* mo_gas_optics_rrtmgp_kernels.F90:178-340 (compute_tau_absorption) */
void rrtmgp_compute_tau_absorption(
        int* ncol_, int* nlay_, int* nband_, int* ngpt_,
        int* ngas_, int* nflav_, int* neta_, int* npres_, int* ntemp_,
        int* nminorlower_, int* nminorklower_,
        int* nminorupper_, int* nminorkupper_,
        int* idx_h2o_,
        int* gpoint_flavor,
        int* band_lims_gpt,
        Float* kmajor,
        Float* kminor_lower,
        Float* kminor_upper,
        int* minor_limits_gpt_lower,
        int* minor_limits_gpt_upper,
        Bool* minor_scales_with_density_lower,
        Bool* minor_scales_with_density_upper,
        Bool* scale_by_complement_lower,
        Bool* scale_by_complement_upper,
        int* idx_minor_lower,
        int* idx_minor_upper,
        int* idx_minor_scaling_lower,
        int* idx_minor_scaling_upper,
        int* kminor_start_lower,
        int* kminor_start_upper,
        Bool* tropo,
        Float* col_mix, Float* fmajor, Float* fminor,
        Float* play, Float* tlay, Float* col_gas,
        int* jeta, int* jtemp, int* jpress,
        Float* tau)
{
    const int ncol = *ncol_, nlay = *nlay_, nbnd = *nband_, ngpt = *ngpt_;
    const int ngas = *ngas_, nflav = *nflav_, neta = *neta_, npres = *npres_, ntemp = *ntemp_;

    int* itropo_lower = (int*)malloc(sizeof(int) * 2 * ncol);
    int* itropo_upper = (int*)malloc(sizeof(int) * 2 * ncol);

    /* F90:276-287 */
    const int top_at_1 = play[0] < play[(size_t)ncol * (nlay - 1)];
    for (int icol = 0; icol < ncol; ++icol)
    {
        if (top_at_1)
        {
            itropo_lower[icol] = minloc_masked(play, tropo, 1, ncol, nlay, icol);
            itropo_lower[icol + ncol] = nlay;
            itropo_upper[icol] = 1;
            itropo_upper[icol + ncol] = maxloc_masked(play, tropo, 0, ncol, nlay, icol);
        }
        else
        {
            itropo_lower[icol] = 1;
            itropo_lower[icol + ncol] = minloc_masked(play, tropo, 1, ncol, nlay, icol);
            itropo_upper[icol] = maxloc_masked(play, tropo, 0, ncol, nlay, icol);
            itropo_upper[icol + ncol] = nlay;
        }
    }

    gas_optical_depths_major(ncol, nlay, nbnd, ngpt, nflav, neta, npres, ntemp,
                             gpoint_flavor, band_lims_gpt, kmajor, col_mix, fmajor,
                             jeta, tropo, jtemp, jpress, tau);

    gas_optical_depths_minor(ncol, nlay, ngpt, ngas, nflav, ntemp, neta,
                             *nminorlower_, *nminorklower_, *idx_h2o_,
                             gpoint_flavor, 0,
                             kminor_lower, minor_limits_gpt_lower,
                             minor_scales_with_density_lower, scale_by_complement_lower,
                             idx_minor_lower, idx_minor_scaling_lower, kminor_start_lower,
                             play, tlay, col_gas, fminor, jeta, itropo_lower, jtemp, tau);

    gas_optical_depths_minor(ncol, nlay, ngpt, ngas, nflav, ntemp, neta,
                             *nminorupper_, *nminorkupper_, *idx_h2o_,
                             gpoint_flavor, 1,
                             kminor_upper, minor_limits_gpt_upper,
                             minor_scales_with_density_upper, scale_by_complement_upper,
                             idx_minor_upper, idx_minor_scaling_upper, kminor_start_upper,
                             play, tlay, col_gas, fminor, jeta, itropo_upper, jtemp, tau);

    free(itropo_lower);
    free(itropo_upper);
}

/* mo_gas_optics_rrtmgp_kernels.F90:507-564 (compute_tau_rayleigh); krayl(ntemp,neta,ngpt,2) */
void rrtmgp_compute_tau_rayleigh(
        int* ncol_, int* nlay_, int* nband_, int* ngpt_,
        int* ngas_, int* nflav_, int* neta_, int* npres_, int* ntemp_,
        int* gpoint_flavor,
        int* band_lims_gpt,
        Float* krayl,
        int* idx_h2o_, Float* col_dry, Float* col_gas,
        Float* fminor, int* jeta,
        Bool* tropo, int* jtemp,
        Float* tau_rayleigh)
{
    const int ncol = *ncol_, nlay = *nlay_, nbnd = *nband_, ngpt = *ngpt_;
    const int neta = *neta_, ntemp = *ntemp_, idx_h2o = *idx_h2o_;
    const size_t ncl = (size_t)ncol * nlay;
    (void)ngas_; (void)nflav_; (void)npres_;
    for (int ibnd = 0; ibnd < nbnd; ++ibnd)
    {
        const int gptS = band_lims_gpt[0 + 2 * ibnd];
        const int gptE = band_lims_gpt[1 + 2 * ibnd];
        for (int ilay = 0; ilay < nlay; ++ilay)
            for (int icol = 0; icol < ncol; ++icol)
            {
                const size_t i = icol + (size_t)ncol * ilay;
                const int itropo = tropo[i] ? 1 : 2;
                const int iflav = gpoint_flavor[(itropo - 1) + 2 * (gptS - 1)];
                const size_t ifl = i + ncl * (iflav - 1);
                const Float* kr = krayl + (size_t)ntemp * neta * ngpt * (itropo - 1);
                for (int igpt = gptS; igpt <= gptE; ++igpt)
                {
                    const Float k = interpolate2D(&fminor[4 * ifl], kr, igpt - 1,
                                                  &jeta[2 * ifl], jtemp[i], ntemp, neta);
                    tau_rayleigh[i + ncl * (igpt - 1)] = k * (col_gas[i + ncl * idx_h2o] + col_dry[i]);
                }
            }
    }
}

/* mo_gas_optics_rrtmgp_kernels.F90:715-734 (interpolate1D): all bands; table(nPlanckTemp,nbnd) */
static void interpolate1D(Float val, Float offset, Float delta, const Float* table,
                          int n1, int nbnd, Float* res)
{
    const Float val0 = (val - offset) / delta;
    const Float frac = val0 - (Float)(int)val0;
    const int index = MIN(n1 - 1, MAX(1, (int)val0 + 1));
    for (int ibnd = 0; ibnd < nbnd; ++ibnd)
    {
        const Float t0 = table[(index - 1) + (size_t)n1 * ibnd];
        const Float t1 = table[index + (size_t)n1 * ibnd];
        res[ibnd] = t0 + frac * (t1 - t0);
    }
}

/* mo_gas_optics_rrtmgp_kernels.F90:567-710 (compute_Planck_source) */
void rrtmgp_compute_Planck_source(
        int* ncol_, int* nlay_, int* nbnd_, int* ngpt_,
        int* nflav_, int* neta_, int* npres_, int* ntemp_, int* nPlanckTemp_,
        Float* tlay, Float* tlev, Float* tsfc, int* sfc_lay_,
        Float* fmajor, int* jeta, Bool* tropo, int* jtemp, int* jpress,
        int* gpoint_bands, int* band_lims_gpt, Float* pfracin, Float* temp_ref_min_,
        Float* totplnk_delta_, Float* totplnk, int* gpoint_flavor,
        Float* sfc_src, Float* lay_src, Float* lev_src,
        Float* sfc_src_jac)
{
    const int ncol = *ncol_, nlay = *nlay_, nbnd = *nbnd_, ngpt = *ngpt_;
    const int neta = *neta_, npres = *npres_, ntemp = *ntemp_, nPlanckTemp = *nPlanckTemp_;
    const int sfc_lay = *sfc_lay_;
    const Float temp_ref_min = *temp_ref_min_, totplnk_delta = *totplnk_delta_;
    const Float delta_Tsurf = 1.0;
    const Float one[2] = {1.0, 1.0};
    const size_t ncl = (size_t)ncol * nlay;
    const size_t ncv = (size_t)ncol * (nlay + 1);
    (void)nflav_; (void)gpoint_bands;

    Float* pfrac = (Float*)malloc(sizeof(Float) * ncl * ngpt);
    Float* planck_function = (Float*)malloc(sizeof(Float) * ncv * nbnd);  /* (ncol,nlay+1,nbnd) */
    Float* tmp = (Float*)malloc(sizeof(Float) * nbnd);
    Float* tmp2 = (Float*)malloc(sizeof(Float) * nbnd);

    /* F90:619-634 */
    for (int ibnd = 0; ibnd < nbnd; ++ibnd)
    {
        const int gptS = band_lims_gpt[0 + 2 * ibnd];
        const int gptE = band_lims_gpt[1 + 2 * ibnd];
        for (int ilay = 0; ilay < nlay; ++ilay)
            for (int icol = 0; icol < ncol; ++icol)
            {
                const size_t i = icol + (size_t)ncol * ilay;
                const int itropo = tropo[i] ? 1 : 2;
                const int iflav = gpoint_flavor[(itropo - 1) + 2 * (gptS - 1)];
                const size_t ifl = i + ncl * (iflav - 1);
                for (int igpt = gptS; igpt <= gptE; ++igpt)
                    pfrac[i + ncl * (igpt - 1)] = interpolate3D(
                            one, &fmajor[8 * ifl], pfracin, igpt - 1, &jeta[2 * ifl],
                            jtemp[i], jpress[i] + itropo, ntemp, neta, npres + 1);
            }
    }

    /* F90:640-656: surface */
    for (int icol = 0; icol < ncol; ++icol)
    {
        interpolate1D(tsfc[icol], temp_ref_min, totplnk_delta, totplnk, nPlanckTemp, nbnd, tmp);
        interpolate1D(tsfc[icol] + delta_Tsurf, temp_ref_min, totplnk_delta, totplnk, nPlanckTemp, nbnd, tmp2);
        for (int ibnd = 0; ibnd < nbnd; ++ibnd)
        {
            const int gptS = band_lims_gpt[0 + 2 * ibnd];
            const int gptE = band_lims_gpt[1 + 2 * ibnd];
            for (int igpt = gptS; igpt <= gptE; ++igpt)
            {
                const Float pf = pfrac[icol + (size_t)ncol * (sfc_lay - 1) + ncl * (igpt - 1)];
                sfc_src[icol + (size_t)ncol * (igpt - 1)] = pf * tmp[ibnd];
                sfc_src_jac[icol + (size_t)ncol * (igpt - 1)] = pf * (tmp2[ibnd] - tmp[ibnd]);
            }
        }
    }

    /* F90:658-679: layer sources */
    for (int ilay = 0; ilay < nlay; ++ilay)
        for (int icol = 0; icol < ncol; ++icol)
        {
            interpolate1D(tlay[icol + (size_t)ncol * ilay], temp_ref_min, totplnk_delta, totplnk,
                          nPlanckTemp, nbnd, tmp);
            for (int ibnd = 0; ibnd < nbnd; ++ibnd)
                planck_function[icol + (size_t)ncol * ilay + ncv * ibnd] = tmp[ibnd];
        }
    for (int ibnd = 0; ibnd < nbnd; ++ibnd)
    {
        const int gptS = band_lims_gpt[0 + 2 * ibnd];
        const int gptE = band_lims_gpt[1 + 2 * ibnd];
        for (int igpt = gptS; igpt <= gptE; ++igpt)
            for (int ilay = 0; ilay < nlay; ++ilay)
                for (int icol = 0; icol < ncol; ++icol)
                {
                    const size_t i = icol + (size_t)ncol * ilay;
                    lay_src[i + ncl * (igpt - 1)] =
                            pfrac[i + ncl * (igpt - 1)] * planck_function[i + ncv * ibnd];
                }
    }

    /* F90:681-708: level sources */
    for (int ilev = 0; ilev < nlay + 1; ++ilev)
        for (int icol = 0; icol < ncol; ++icol)
        {
            interpolate1D(tlev[icol + (size_t)ncol * ilev], temp_ref_min, totplnk_delta, totplnk,
                          nPlanckTemp, nbnd, tmp);
            for (int ibnd = 0; ibnd < nbnd; ++ibnd)
                planck_function[icol + (size_t)ncol * ilev + ncv * ibnd] = tmp[ibnd];
        }
    for (int ibnd = 0; ibnd < nbnd; ++ibnd)
    {
        const int gptS = band_lims_gpt[0 + 2 * ibnd];
        const int gptE = band_lims_gpt[1 + 2 * ibnd];
        for (int igpt = gptS; igpt <= gptE; ++igpt)
        {
            const size_t og = ncl * (igpt - 1), ov = ncv * (igpt - 1);
            for (int icol = 0; icol < ncol; ++icol)
                lev_src[icol + ov] = pfrac[icol + og] * planck_function[icol + ncv * ibnd];
            for (int ilay = 1; ilay < nlay; ++ilay)
                for (int icol = 0; icol < ncol; ++icol)
                {
                    const size_t i = icol + (size_t)ncol * ilay;
                    lev_src[i + ov] = sqrt(pfrac[i - ncol + og] * pfrac[i + og]) *
                                      planck_function[i + ncv * ibnd];
                }
            for (int icol = 0; icol < ncol; ++icol)
            {
                const size_t i = icol + (size_t)ncol * nlay;
                lev_src[i + ov] = pfrac[icol + (size_t)ncol * (nlay - 1) + og] *
                                  planck_function[i + ncv * ibnd];
            }
        }
    }

    free(pfrac); free(planck_function); free(tmp); free(tmp2);
}

/* Declared in rrtmgp_kernels.h:128-135 but unused by the reference's src/; provided so the
 * symbols resolve.  Follows mo_gas_optics_rrtmgp_kernels reorder semantics: out(k,j,i)=in(i,j,k). */
void reorder_123x321_kernel(int* d1_, int* d2_, int* d3_, Float* array, Float* array_out)
{
    const int d1 = *d1_, d2 = *d2_, d3 = *d3_;
    for (int i3 = 0; i3 < d3; ++i3)
        for (int i2 = 0; i2 < d2; ++i2)
            for (int i1 = 0; i1 < d1; ++i1)
                array_out[i3 + (size_t)d3 * (i2 + (size_t)d2 * i1)] =
                        array[i1 + (size_t)d1 * (i2 + (size_t)d2 * i3)];
}

/* ---------------------------------------------------------------------------------------------
 * rte-rrtmgp/rte-kernels/mo_optical_props_kernels.F90
 * ------------------------------------------------------------------------------------------- */
#define OP_EPS (3.0 * DBL_MIN)  /* F90:38 eps = 3*tiny */

/* F90:76-98 (delta_scale_2str_k) */
void rte_delta_scale_2str_k(int* ncol, int* nlay, int* ngpt, Float* tau, Float* ssa, Float* g)
{
    const size_t n = (size_t)(*ncol) * (*nlay) * (*ngpt);
    for (size_t i = 0; i < n; ++i)
    {
        const Float f = g[i] * g[i];
        const Float wf = ssa[i] * f;
        tau[i] = (1.0 - wf) * tau[i];
        ssa[i] = (ssa[i] - wf) / MAX(OP_EPS, (1.0 - wf));
        g[i] = (g[i] - f) / MAX(OP_EPS, (1.0 - f));
    }
}

/* F90:116-132 */
void rte_increment_1scalar_by_1scalar(int* ncol, int* nlay, int* ngpt, Float* tau1, Float* tau2)
{
    const size_t n = (size_t)(*ncol) * (*nlay) * (*ngpt);
    for (size_t i = 0; i < n; ++i) tau1[i] = tau1[i] + tau2[i];
}

/* F90:199-226 */
void rte_increment_2stream_by_2stream(int* ncol, int* nlay, int* ngpt,
                                      Float* tau1, Float* ssa1, Float* g1,
                                      Float* tau2, Float* ssa2, Float* g2)
{
    const size_t n = (size_t)(*ncol) * (*nlay) * (*ngpt);
    for (size_t i = 0; i < n; ++i)
    {
        const Float tau12 = tau1[i] + tau2[i];
        const Float tauscat12 = tau1[i] * ssa1[i] + tau2[i] * ssa2[i];
        g1[i] = (tau1[i] * ssa1[i] * g1[i] + tau2[i] * ssa2[i] * g2[i]) / MAX(OP_EPS, tauscat12);
        ssa1[i] = tauscat12 / MAX(OP_EPS, tau12);
        tau1[i] = tau12;
    }
}

/* F90:366-382 */
void rte_inc_1scalar_by_1scalar_bybnd(int* ncol, int* nlay, int* ngpt, Float* tau1, Float* tau2,
                                      int* nbnd, int* gpt_lims)
{
    const size_t ncl = (size_t)(*ncol) * (*nlay);
    (void)ngpt;
    for (int ibnd = 0; ibnd < *nbnd; ++ibnd)
        for (int igpt = gpt_lims[0 + 2 * ibnd]; igpt <= gpt_lims[1 + 2 * ibnd]; ++igpt)
            for (size_t i = 0; i < ncl; ++i)
                tau1[i + ncl * (igpt - 1)] = tau1[i + ncl * (igpt - 1)] + tau2[i + ncl * ibnd];
}

/* F90:451-482 */
void rte_inc_2stream_by_2stream_bybnd(int* ncol, int* nlay, int* ngpt,
                                      Float* tau1, Float* ssa1, Float* g1,
                                      Float* tau2, Float* ssa2, Float* g2,
                                      int* nbnd, int* gpt_lims)
{
    const size_t ncl = (size_t)(*ncol) * (*nlay);
    (void)ngpt;
    for (int ibnd = 0; ibnd < *nbnd; ++ibnd)
        for (int igpt = gpt_lims[0 + 2 * ibnd]; igpt <= gpt_lims[1 + 2 * ibnd]; ++igpt)
            for (size_t i = 0; i < ncl; ++i)
            {
                const size_t a = i + ncl * (igpt - 1), b = i + ncl * ibnd;
                const Float tau12 = tau1[a] + tau2[b];
                const Float tauscat12 = tau1[a] * ssa1[a] + tau2[b] * ssa2[b];
                g1[a] = (tau1[a] * ssa1[a] * g1[a] + tau2[b] * ssa2[b] * g2[b]) / MAX(OP_EPS, tauscat12);
                ssa1[a] = tauscat12 / MAX(OP_EPS, tau12);
                tau1[a] = tau12;
            }
}

/* ---------------------------------------------------------------------------------------------
 * rte-rrtmgp/rte-kernels/mo_rte_solver_kernels.F90 — longwave
 * ------------------------------------------------------------------------------------------- */

/* F90:620-675 (lw_source_noscat) for one g-point; arrays (ncol,nlay) / (ncol,nlay+1) */
static void lw_source_noscat(int ncol, int nlay, int top_at_1,
                             const Float* lay_source, const Float* lev_source,
                             const Float* tau, const Float* trans,
                             Float* source_dn, Float* source_up)
{
    const Float tau_thresh = sqrt(sqrt(DBL_EPSILON));
    Float* source_inc = top_at_1 ? source_dn : source_up;
    Float* source_dec = top_at_1 ? source_up : source_dn;
    for (int ilay = 0; ilay < nlay; ++ilay)
        for (int icol = 0; icol < ncol; ++icol)
        {
            const size_t i = icol + (size_t)ncol * ilay;
            Float fact;
            if (tau[i] > tau_thresh)
                fact = (1.0 - trans[i]) / tau[i] - trans[i];
            else
                fact = tau[i] * (0.5 + tau[i] * (-1.0 / 3.0 + tau[i] * 1.0 / 8.0));
            source_inc[i] = (1.0 - trans[i]) * lev_source[i + ncol] +
                            2.0 * fact * (lay_source[i] - lev_source[i + ncol]);
            source_dec[i] = (1.0 - trans[i]) * lev_source[i] +
                            2.0 * fact * (lay_source[i] - lev_source[i]);
        }
}

/* F90:681-708 */
static void lw_transport_noscat_dn(int ncol, int nlay, int top_at_1,
                                   const Float* trans, const Float* source_dn, Float* radn_dn)
{
    if (top_at_1)
    {
        for (int ilev = 1; ilev < nlay + 1; ++ilev)
            for (int icol = 0; icol < ncol; ++icol)
                radn_dn[icol + (size_t)ncol * ilev] =
                        trans[icol + (size_t)ncol * (ilev - 1)] * radn_dn[icol + (size_t)ncol * (ilev - 1)] +
                        source_dn[icol + (size_t)ncol * (ilev - 1)];
    }
    else
    {
        for (int ilev = nlay - 1; ilev >= 0; --ilev)
            for (int icol = 0; icol < ncol; ++icol)
                radn_dn[icol + (size_t)ncol * ilev] =
                        trans[icol + (size_t)ncol * ilev] * radn_dn[icol + (size_t)ncol * (ilev + 1)] +
                        source_dn[icol + (size_t)ncol * ilev];
    }
}

/* F90:710-745 */
static void lw_transport_noscat_up(int ncol, int nlay, int top_at_1,
                                   const Float* trans, const Float* source_up, Float* radn_up,
                                   int do_Jacobians, Float* radn_upJac)
{
    if (top_at_1)
    {
        for (int ilev = nlay - 1; ilev >= 0; --ilev)
            for (int icol = 0; icol < ncol; ++icol)
            {
                const size_t i = icol + (size_t)ncol * ilev;
                radn_up[i] = trans[i] * radn_up[i + ncol] + source_up[i];
                if (do_Jacobians) radn_upJac[i] = trans[i] * radn_upJac[i + ncol];
            }
    }
    else
    {
        for (int ilev = 1; ilev < nlay + 1; ++ilev)
            for (int icol = 0; icol < ncol; ++icol)
            {
                const size_t i = icol + (size_t)ncol * ilev;
                radn_up[i] = trans[i - ncol] * radn_up[i - ncol] + source_up[i - ncol];
                if (do_Jacobians) radn_upJac[i] = trans[i - ncol] * radn_upJac[i - ncol];
            }
    }
}

/* F90:753-844 (lw_transport_1rescl): upward and second downward transport of the re-scaled solution
 * (Tang et al. 2018).  The level whose radiance enters each adjustment is the Fortran's, per orientation. */
static void lw_transport_1rescl(int ncol, int nlay, int top_at_1,
                                const Float* trans, const Float* source_dn, const Float* source_up,
                                Float* radn_up, Float* radn_dn, const Float* An, const Float* Cn,
                                int do_Jacobians, Float* radn_up_Jac)
{
    if (top_at_1)
    {
        for (int ilev = nlay - 1; ilev >= 0; --ilev)
            for (int icol = 0; icol < ncol; ++icol)
            {
                const size_t i = icol + (size_t)ncol * ilev;
                const Float adjustmentFactor = Cn[i] * (An[i] * radn_dn[i] - trans[i] * source_dn[i] - source_up[i]);
                radn_up[i] = trans[i] * radn_up[i + ncol] + source_up[i] + adjustmentFactor;
                if (do_Jacobians) radn_up_Jac[i] = trans[i] * radn_up_Jac[i + ncol];
            }
        for (int ilev = 0; ilev < nlay; ++ilev)
            for (int icol = 0; icol < ncol; ++icol)
            {
                const size_t i = icol + (size_t)ncol * ilev;
                const Float adjustmentFactor = Cn[i] * (An[i] * radn_up[i] - trans[i] * source_up[i] - source_dn[i]);
                radn_dn[i + ncol] = trans[i] * radn_dn[i] + source_dn[i] + adjustmentFactor;
            }
    }
    else
    {
        for (int ilev = 0; ilev < nlay; ++ilev)
            for (int icol = 0; icol < ncol; ++icol)
            {
                const size_t i = icol + (size_t)ncol * ilev;
                const Float adjustmentFactor = Cn[i] * (An[i] * radn_dn[i + ncol] - trans[i] * source_dn[i] - source_up[i]);
                radn_up[i + ncol] = trans[i] * radn_up[i] + source_up[i] + adjustmentFactor;
                if (do_Jacobians) radn_up_Jac[i + ncol] = trans[i] * radn_up_Jac[i];
            }
        for (int ilev = nlay - 1; ilev >= 0; --ilev)
            for (int icol = 0; icol < ncol; ++icol)
            {
                const size_t i = icol + (size_t)ncol * ilev;
                const Float adjustmentFactor = Cn[i] * (An[i] * radn_up[i] - trans[i] * source_up[i] - source_dn[i]);
                radn_dn[i] = trans[i] * radn_dn[i + ncol] + source_dn[i] + adjustmentFactor;
            }
    }
}

/* F90:51-240 (lw_solver_noscat_oneangle), with the re-scaling branch :148-178,206-214 (the reference C++
 * passes do_rescaling = false, src/Rte_lw.cpp:178; the branch is restated for SURVEY 8f-4). D(ncol,ngpt). */
static void lw_solver_noscat_oneangle(
        int ncol, int nlay, int ngpt, int top_at_1, const Float* D, Float weight,
        const Float* tau, const Float* lay_source, const Float* lev_source,
        const Float* sfc_emis, const Float* sfc_src, const Float* incident_flux,
        Float* flux_up, Float* flux_dn,
        int do_broadband, Float* broadband_up, Float* broadband_dn,
        int do_Jacobians, const Float* sfc_srcJac, Float* flux_upJac,
        int do_rescaling, const Float* ssa, const Float* g)
{
    const Float pi = acos(-1.0);
    const size_t ncl = (size_t)ncol * nlay, ncv = (size_t)ncol * (nlay + 1);
    const int top_level = top_at_1 ? 0 : nlay;
    const int sfc_level = top_at_1 ? nlay : 0;

    Float* tau_loc = (Float*)malloc(sizeof(Float) * ncl);
    Float* trans = (Float*)malloc(sizeof(Float) * ncl);
    Float* source_dn = (Float*)malloc(sizeof(Float) * ncl);
    Float* source_up = (Float*)malloc(sizeof(Float) * ncl);
    Float* loc_flux_up = (Float*)malloc(sizeof(Float) * ncv);
    Float* loc_flux_dn = (Float*)malloc(sizeof(Float) * ncv);
    Float* gpt_flux_Jac = (Float*)malloc(sizeof(Float) * ncv);
    Float* An = do_rescaling ? (Float*)malloc(sizeof(Float) * ncl) : NULL;
    Float* Cn = do_rescaling ? (Float*)malloc(sizeof(Float) * ncl) : NULL;

    if (do_broadband)
        for (size_t i = 0; i < ncv; ++i) { broadband_up[i] = 0.0; broadband_dn[i] = 0.0; }
    if (do_Jacobians)
        for (size_t i = 0; i < ncv; ++i) flux_upJac[i] = 0.0;

    for (int igpt = 0; igpt < ngpt; ++igpt)
    {
        Float* gpt_flux_up = do_broadband ? loc_flux_up : flux_up + ncv * igpt;
        Float* gpt_flux_dn = do_broadband ? loc_flux_dn : flux_dn + ncv * igpt;

        for (int icol = 0; icol < ncol; ++icol)
            gpt_flux_dn[icol + (size_t)ncol * top_level] =
                    incident_flux[icol + (size_t)ncol * igpt] / (pi * weight);

        if (do_rescaling)
            for (int ilay = 0; ilay < nlay; ++ilay)
                for (int icol = 0; icol < ncol; ++icol)
                {
                    const size_t i = icol + (size_t)ncol * ilay;
                    const Float ssal = ssa[i + ncl * igpt];
                    const Float wb = ssal * (1.0 - g[i + ncl * igpt]) * 0.5;
                    const Float scaleTau = (1.0 - ssal + wb);
                    Cn[i] = 0.4 * wb / scaleTau;
                    tau_loc[i] = tau[i + ncl * igpt] * D[icol + (size_t)ncol * igpt] * scaleTau;
                    trans[i] = exp(-tau_loc[i]);
                    An[i] = (1.0 - trans[i] * trans[i]);
                }
        else
            for (int ilay = 0; ilay < nlay; ++ilay)
                for (int icol = 0; icol < ncol; ++icol)
                {
                    const size_t i = icol + (size_t)ncol * ilay;
                    tau_loc[i] = tau[i + ncl * igpt] * D[icol + (size_t)ncol * igpt];
                    trans[i] = exp(-tau_loc[i]);
                }

        lw_source_noscat(ncol, nlay, top_at_1, lay_source + ncl * igpt, lev_source + ncv * igpt,
                         tau_loc, trans, source_dn, source_up);
        lw_transport_noscat_dn(ncol, nlay, top_at_1, trans, source_dn, gpt_flux_dn);

        for (int icol = 0; icol < ncol; ++icol)
        {
            const Float emis = sfc_emis[icol + (size_t)ncol * igpt];
            const Float sfc_albedo = 1.0 - emis;
            gpt_flux_up[icol + (size_t)ncol * sfc_level] =
                    gpt_flux_dn[icol + (size_t)ncol * sfc_level] * sfc_albedo +
                    emis * sfc_src[icol + (size_t)ncol * igpt];
            if (do_Jacobians)
                gpt_flux_Jac[icol + (size_t)ncol * sfc_level] = emis * sfc_srcJac[icol + (size_t)ncol * igpt];
        }

        if (do_rescaling)
            lw_transport_1rescl(ncol, nlay, top_at_1, trans, source_dn, source_up, gpt_flux_up, gpt_flux_dn,
                                An, Cn, do_Jacobians, gpt_flux_Jac);
        else
            lw_transport_noscat_up(ncol, nlay, top_at_1, trans, source_up, gpt_flux_up,
                                   do_Jacobians, gpt_flux_Jac);

        if (do_broadband)
            for (size_t i = 0; i < ncv; ++i)
            {
                broadband_up[i] = broadband_up[i] + gpt_flux_up[i];
                broadband_dn[i] = broadband_dn[i] + gpt_flux_dn[i];
            }
        else
            for (size_t i = 0; i < ncv; ++i)
            {
                gpt_flux_dn[i] = pi * weight * gpt_flux_dn[i];
                gpt_flux_up[i] = pi * weight * gpt_flux_up[i];
            }
        if (do_Jacobians)
            for (size_t i = 0; i < ncv; ++i) flux_upJac[i] = flux_upJac[i] + gpt_flux_Jac[i];
    }

    if (do_broadband)
        for (size_t i = 0; i < ncv; ++i)
        {
            broadband_up[i] = pi * weight * broadband_up[i];
            broadband_dn[i] = pi * weight * broadband_dn[i];
        }
    if (do_Jacobians)
        for (size_t i = 0; i < ncv; ++i) flux_upJac[i] = pi * weight * flux_upJac[i];

    free(tau_loc); free(trans); free(source_dn); free(source_up);
    free(loc_flux_up); free(loc_flux_dn); free(gpt_flux_Jac);
    free(An); free(Cn);
}

/* F90:248-367 (lw_solver_noscat).  Scalars arrive by reference (rrtmgp_kernels.h:207-232). */
void rte_lw_solver_noscat(
        const int* ncol_, const int* nlay_, const int* ngpt_, const Bool* top_at_1_,
        const int* nmus_, const Float* Ds, const Float* weights,
        const Float* tau, const Float* lay_source, const Float* lev_source,
        const Float* sfc_emis, const Float* sfc_src, const Float* inc_flux,
        Float* flux_up, Float* flux_dn,
        const Bool* do_broadband_, Float* broadband_up, Float* broadband_dn,
        const Bool* do_Jacobians_, const Float* sfc_srcJac, Float* flux_upJac,
        const Bool* do_rescaling_, const Float* ssa, const Float* g)
{
    const int ncol = *ncol_, nlay = *nlay_, ngpt = *ngpt_, nmus = *nmus_;
    const int top_at_1 = *top_at_1_ != 0, do_broadband = *do_broadband_ != 0;
    const int do_Jacobians = *do_Jacobians_ != 0;
    const size_t ncv = (size_t)ncol * (nlay + 1);
    const int do_rescaling = *do_rescaling_ != 0;

    lw_solver_noscat_oneangle(ncol, nlay, ngpt, top_at_1, Ds, weights[0],
                              tau, lay_source, lev_source, sfc_emis, sfc_src, inc_flux,
                              flux_up, flux_dn, do_broadband, broadband_up, broadband_dn,
                              do_Jacobians, sfc_srcJac, flux_upJac, do_rescaling, ssa, g);
    if (nmus > 1)
    {
        Float *this_bb_up = NULL, *this_bb_dn = NULL, *this_up = NULL, *this_dn = NULL, *this_jac = NULL;
        if (do_broadband)
        {
            this_bb_up = (Float*)malloc(sizeof(Float) * ncv);
            this_bb_dn = (Float*)malloc(sizeof(Float) * ncv);
            this_up = flux_up; this_dn = flux_dn;
        }
        else
        {
            this_up = (Float*)malloc(sizeof(Float) * ncv * ngpt);
            this_dn = (Float*)malloc(sizeof(Float) * ncv * ngpt);
            this_bb_up = broadband_up; this_bb_dn = broadband_dn;
        }
        this_jac = do_Jacobians ? (Float*)malloc(sizeof(Float) * ncv) : flux_upJac;

        for (int imu = 1; imu < nmus; ++imu)
        {
            lw_solver_noscat_oneangle(ncol, nlay, ngpt, top_at_1,
                                      Ds + (size_t)ncol * ngpt * imu, weights[imu],
                                      tau, lay_source, lev_source, sfc_emis, sfc_src, inc_flux,
                                      this_up, this_dn, do_broadband, this_bb_up, this_bb_dn,
                                      do_Jacobians, sfc_srcJac, this_jac, do_rescaling, ssa, g);
            if (do_broadband)
                for (size_t i = 0; i < ncv; ++i)
                {
                    broadband_up[i] = broadband_up[i] + this_bb_up[i];
                    broadband_dn[i] = broadband_dn[i] + this_bb_dn[i];
                }
            else
                for (size_t i = 0; i < ncv * ngpt; ++i)
                {
                    flux_up[i] = flux_up[i] + this_up[i];
                    flux_dn[i] = flux_dn[i] + this_dn[i];
                }
            if (do_Jacobians)
                for (size_t i = 0; i < ncv; ++i) flux_upJac[i] = flux_upJac[i] + this_jac[i];
        }
        if (do_broadband) { free(this_bb_up); free(this_bb_dn); }
        else { free(this_up); free(this_dn); }
        if (do_Jacobians) free(this_jac);
    }
}

/* ---------------------------------------------------------------------------------------------
 * mo_rte_solver_kernels.F90 — shortwave
 * ------------------------------------------------------------------------------------------- */

/* F90:985-1127 (sw_dif_and_source) for one g-point. mu0(ncol,nlay). */
static void sw_dif_and_source(int ncol, int nlay, int top_at_1, const Float* mu0,
                              const Float* sfc_albedo, const Float* tau, const Float* w0,
                              const Float* g, Float* Rdif, Float* Tdif, Float* source_dn,
                              Float* source_up, Float* source_sfc, Float* flux_dn_dir)
{
    const Float min_k = 1.e4 * DBL_EPSILON;
    const Float min_mu0 = sqrt(DBL_EPSILON);
    int lay_index = 0;
    Float* dir_flux_trans = NULL;

    for (int j = 0; j < nlay; ++j)
    {
        Float* dir_flux_inc;
        if (top_at_1)
        {
            lay_index = j;
            dir_flux_inc = flux_dn_dir + (size_t)ncol * lay_index;
            dir_flux_trans = flux_dn_dir + (size_t)ncol * (lay_index + 1);
        }
        else
        {
            lay_index = nlay - 1 - j;
            dir_flux_inc = flux_dn_dir + (size_t)ncol * (lay_index + 1);
            dir_flux_trans = flux_dn_dir + (size_t)ncol * lay_index;
        }
        for (int i = 0; i < ncol; ++i)
        {
            const size_t ii = i + (size_t)ncol * lay_index;
            const Float tau_s = tau[ii], w0_s = w0[ii], g_s = g[ii];

            const Float gamma1 = (8.0 - w0_s * (5.0 + 3.0 * g_s)) * .25;
            const Float gamma2 = 3.0 * (w0_s * (1.0 - g_s)) * .25;

            const Float k = sqrt(MAX((gamma1 - gamma2) * (gamma1 + gamma2), min_k));
            const Float exp_minusktau = exp(-tau_s * k);
            const Float exp_minus2ktau = exp_minusktau * exp_minusktau;

            Float RT_term = 1.0 / (k * (1.0 + exp_minus2ktau) + gamma1 * (1.0 - exp_minus2ktau));
            Rdif[ii] = RT_term * gamma2 * (1.0 - exp_minus2ktau);
            Tdif[ii] = RT_term * 2.0 * k * exp_minusktau;

            const Float mu0_s = MAX(min_mu0, mu0[ii]);
            const Float k_mu = k * mu0_s;
            const Float d = 1.0 - k_mu * k_mu;
            RT_term = w0_s * RT_term / ((fabs(d) >= DBL_EPSILON) ? d : DBL_EPSILON);

            const Float gamma3 = (2.0 - 3.0 * mu0_s * g_s) * .25;
            const Float gamma4 = 1.0 - gamma3;
            const Float alpha1 = gamma1 * gamma4 + gamma2 * gamma3;
            const Float alpha2 = gamma1 * gamma3 + gamma2 * gamma4;

            const Float k_gamma3 = k * gamma3;
            const Float k_gamma4 = k * gamma4;
            const Float Tnoscat = exp(-tau_s / mu0_s);
            Float Rdir = RT_term *
                    ((1.0 - k_mu) * (alpha2 + k_gamma3) -
                     (1.0 + k_mu) * (alpha2 - k_gamma3) * exp_minus2ktau -
                     2.0 * (k_gamma3 - alpha2 * k_mu) * exp_minusktau * Tnoscat);
            Float Tdir = -RT_term *
                    ((1.0 + k_mu) * (alpha1 + k_gamma4) * Tnoscat -
                     (1.0 - k_mu) * (alpha1 - k_gamma4) * exp_minus2ktau * Tnoscat -
                     2.0 * (k_gamma4 + alpha1 * k_mu) * exp_minusktau);
            Rdir = MAX(0.0, MIN(Rdir, (1.0 - Tnoscat)));
            Tdir = MAX(0.0, MIN(Tdir, (1.0 - Tnoscat - Rdir)));

            source_up[ii] = Rdir * dir_flux_inc[i];
            source_dn[ii] = Tdir * dir_flux_inc[i];
            dir_flux_trans[i] = Tnoscat * dir_flux_inc[i];
        }
    }
    /* F90:1120-1125: lay_index and dir_flux_trans keep their last values (surface layer) */
    for (int i = 0; i < ncol; ++i)
        source_sfc[i] = (mu0[i + (size_t)ncol * lay_index] > 0.0) ? dir_flux_trans[i] * sfc_albedo[i] : 0.0;
    for (size_t ii = 0; ii < (size_t)ncol * nlay; ++ii)
        if (mu0[ii] <= 0.0) { source_up[ii] = 0.0; source_dn[ii] = 0.0; }
}

/* F90:1135-1245 (adding) for one g-point */
static void adding(int ncol, int nlay, int top_at_1, const Float* albedo_sfc,
                   const Float* rdif, const Float* tdif, const Float* src_dn, const Float* src_up,
                   const Float* src_sfc, Float* flux_up, Float* flux_dn)
{
    const size_t ncl = (size_t)ncol * nlay, ncv = (size_t)ncol * (nlay + 1);
    Float* albedo = (Float*)malloc(sizeof(Float) * ncv);
    Float* src = (Float*)malloc(sizeof(Float) * ncv);
    Float* denom = (Float*)malloc(sizeof(Float) * ncl);
#define A2(a, i, l) a[(i) + (size_t)ncol * (l)]
    if (top_at_1)
    {
        for (int i = 0; i < ncol; ++i) { A2(albedo, i, nlay) = albedo_sfc[i]; A2(src, i, nlay) = src_sfc[i]; }
        for (int l = nlay - 1; l >= 0; --l)
            for (int i = 0; i < ncol; ++i)
            {
                A2(denom, i, l) = 1.0 / (1.0 - A2(rdif, i, l) * A2(albedo, i, l + 1));
                A2(albedo, i, l) = A2(rdif, i, l) +
                        A2(tdif, i, l) * A2(tdif, i, l) * A2(albedo, i, l + 1) * A2(denom, i, l);
                A2(src, i, l) = A2(src_up, i, l) +
                        A2(tdif, i, l) * A2(denom, i, l) *
                        (A2(src, i, l + 1) + A2(albedo, i, l + 1) * A2(src_dn, i, l));
            }
        for (int i = 0; i < ncol; ++i)
            A2(flux_up, i, 0) = A2(flux_dn, i, 0) * A2(albedo, i, 0) + A2(src, i, 0);
        for (int l = 1; l < nlay + 1; ++l)
            for (int i = 0; i < ncol; ++i)
            {
                A2(flux_dn, i, l) = (A2(tdif, i, l - 1) * A2(flux_dn, i, l - 1) +
                                     A2(rdif, i, l - 1) * A2(src, i, l) +
                                     A2(src_dn, i, l - 1)) * A2(denom, i, l - 1);
                A2(flux_up, i, l) = A2(flux_dn, i, l) * A2(albedo, i, l) + A2(src, i, l);
            }
    }
    else
    {
        for (int i = 0; i < ncol; ++i) { A2(albedo, i, 0) = albedo_sfc[i]; A2(src, i, 0) = src_sfc[i]; }
        for (int l = 0; l < nlay; ++l)
            for (int i = 0; i < ncol; ++i)
            {
                A2(denom, i, l) = 1.0 / (1.0 - A2(rdif, i, l) * A2(albedo, i, l));
                A2(albedo, i, l + 1) = A2(rdif, i, l) +
                        A2(tdif, i, l) * A2(tdif, i, l) * A2(albedo, i, l) * A2(denom, i, l);
                A2(src, i, l + 1) = A2(src_up, i, l) +
                        A2(tdif, i, l) * A2(denom, i, l) *
                        (A2(src, i, l) + A2(albedo, i, l) * A2(src_dn, i, l));
            }
        for (int i = 0; i < ncol; ++i)
            A2(flux_up, i, nlay) = A2(flux_dn, i, nlay) * A2(albedo, i, nlay) + A2(src, i, nlay);
        for (int l = nlay - 1; l >= 0; --l)
            for (int i = 0; i < ncol; ++i)
            {
                A2(flux_dn, i, l) = (A2(tdif, i, l) * A2(flux_dn, i, l + 1) +
                                     A2(rdif, i, l) * A2(src, i, l) +
                                     A2(src_dn, i, l)) * A2(denom, i, l);
                A2(flux_up, i, l) = A2(flux_dn, i, l) * A2(albedo, i, l) + A2(src, i, l);
            }
    }
#undef A2
    free(albedo); free(src); free(denom);
}

/* F90:503-609 (sw_solver_2stream) */
void rte_sw_solver_2stream(
        const int* ncol_, const int* nlay_, const int* ngpt_, const Bool* top_at_1_,
        const Float* tau, const Float* ssa, const Float* g, const Float* mu0,
        const Float* sfc_alb_dir, const Float* sfc_alb_dif, const Float* inc_flux_dir,
        Float* flux_up, Float* flux_dn, Float* flux_dir,
        const Bool* has_dif_bc_, const Float* inc_flux_dif,
        const Bool* do_broadband_, Float* broadband_up, Float* broadband_dn, Float* broadband_dir)
{
    const int ncol = *ncol_, nlay = *nlay_, ngpt = *ngpt_;
    const int top_at_1 = *top_at_1_ != 0, has_dif_bc = *has_dif_bc_ != 0;
    const int do_broadband = *do_broadband_ != 0;
    const size_t ncl = (size_t)ncol * nlay, ncv = (size_t)ncol * (nlay + 1);
    const int top_level = top_at_1 ? 0 : nlay;
    const int top_layer = top_at_1 ? 0 : nlay - 1;

    Float* Rdif = (Float*)malloc(sizeof(Float) * ncl);
    Float* Tdif = (Float*)malloc(sizeof(Float) * ncl);
    Float* source_up = (Float*)malloc(sizeof(Float) * ncl);
    Float* source_dn = (Float*)malloc(sizeof(Float) * ncl);
    Float* source_srf = (Float*)malloc(sizeof(Float) * ncol);
    Float* loc_up = (Float*)malloc(sizeof(Float) * ncv);
    Float* loc_dn = (Float*)malloc(sizeof(Float) * ncv);
    Float* loc_dir = (Float*)malloc(sizeof(Float) * ncv);

    if (do_broadband)
        for (size_t i = 0; i < ncv; ++i) { broadband_up[i] = 0.0; broadband_dn[i] = 0.0; broadband_dir[i] = 0.0; }

    for (int igpt = 0; igpt < ngpt; ++igpt)
    {
        Float* gpt_flux_up = do_broadband ? loc_up : flux_up + ncv * igpt;
        Float* gpt_flux_dn = do_broadband ? loc_dn : flux_dn + ncv * igpt;
        Float* gpt_flux_dir = do_broadband ? loc_dir : flux_dir + ncv * igpt;

        for (int i = 0; i < ncol; ++i)
        {
            gpt_flux_dir[i + (size_t)ncol * top_level] =
                    inc_flux_dir[i + (size_t)ncol * igpt] * mu0[i + (size_t)ncol * top_layer];
            gpt_flux_dn[i + (size_t)ncol * top_level] =
                    has_dif_bc ? inc_flux_dif[i + (size_t)ncol * igpt] : 0.0;
        }

        sw_dif_and_source(ncol, nlay, top_at_1, mu0, sfc_alb_dir + (size_t)ncol * igpt,
                          tau + ncl * igpt, ssa + ncl * igpt, g + ncl * igpt,
                          Rdif, Tdif, source_dn, source_up, source_srf, gpt_flux_dir);

        adding(ncol, nlay, top_at_1, sfc_alb_dif + (size_t)ncol * igpt, Rdif, Tdif,
               source_dn, source_up, source_srf, gpt_flux_up, gpt_flux_dn);

        if (do_broadband)
            for (size_t i = 0; i < ncv; ++i)
            {
                broadband_up[i] = broadband_up[i] + gpt_flux_up[i];
                broadband_dn[i] = broadband_dn[i] + gpt_flux_dn[i] + gpt_flux_dir[i];
                broadband_dir[i] = broadband_dir[i] + gpt_flux_dir[i];
            }
        else
            for (size_t i = 0; i < ncv; ++i) gpt_flux_dn[i] = gpt_flux_dn[i] + gpt_flux_dir[i];
    }

    free(Rdif); free(Tdif); free(source_up); free(source_dn); free(source_srf);
    free(loc_up); free(loc_dn); free(loc_dir);
}

/* ---------------------------------------------------------------------------------------------
 * Heating rate: rte-rrtmgp/extensions/mo_heating_rates.F90:44-62, constants
 * rte-rrtmgp/gas-optics/mo_gas_optics_constants.F90:51-54 (grav 9.80665, cp_dry 1004.64).
 * Not a symbol of the reference C++; exported for the a19 row of SURVEY.md §8.
 * ------------------------------------------------------------------------------------------- */
void oracle_heating_rate(int ncol, int nlay, const Float* flux_up, const Float* flux_dn,
                         const Float* p_lev, Float* heating_rate)
{
    const Float grav = 9.80665, cp_dry = 1004.64;
    for (int ilay = 0; ilay < nlay; ++ilay)
        for (int icol = 0; icol < ncol; ++icol)
        {
            const size_t a = icol + (size_t)ncol * ilay, b = a + ncol;
            heating_rate[a] = (flux_up[b] - flux_up[a] - flux_dn[b] + flux_dn[a]) * grav /
                              (cp_dry * (p_lev[b] - p_lev[a]));
        }
}

/*
 * rrtmgp_b200.h — C ABI of the B200-native (sm_100a) column radiation hot path.
 *
 * This is the drop-in boundary.  The reference has no FFI: callers link its
 * static libs and call the L3 C++ classes.  Each entry point below names the reference
 * interface it replaces (paths relative to the reference root).  The thin C++ class layer
 * in include/rrtmgp_b200/ (same class and method names as the reference) and the Python mirror
 * in rte_rrtmgp_cpp_b200/api.py call only these symbols.
 *
 * Conventions
 *  - every pointer named d_* is a DEVICE pointer; arrays are column-major (first index
 *    fastest) in exactly the reference's shapes; integer tables hold 1-based Fortran indices;
 *  - `Real` is float for a plan created with RRTMGP_B200_F32, double for RRTMGP_B200_F64
 *    (the reference's compile-time `Float`, include/types.h:19-25); Bool is signed char
 *    (RTE_USE_CBOOL, include/types.h:13-17);
 *  - every call returns 0 on success or an RRTMGP_B200_E* code; the message is available from
 *    rrtmgp_b200_last_error() (thread-local).  The C++ class layer maps codes back to the
 *    reference's exception types (E_RANGE -> std::range_error, others -> std::runtime_error);
 *  - every call takes the CUDA stream to launch on (as void*, a cudaStream_t); no call
 *    synchronises the device unless documented; no global state besides the error string;
 *  - there is NO CPU fallback: without a CUDA device every compute call fails with E_CUDA.
 */
#ifndef RRTMGP_B200_H
#define RRTMGP_B200_H

#include <stddef.h>
#include "rrtmgp_b200_tables.h"

#ifdef __cplusplus
extern "C" {
#endif

enum { RRTMGP_B200_F32 = 0, RRTMGP_B200_F64 = 1,
       /* OR-ed into dtype at gas_optics_create: build the host tables only (no device upload), so
        * the constructor logic can be checked on a machine without a GPU.  Compute calls on such a
        * handle fail with E_CUDA. */
       RRTMGP_B200_HOST_ONLY = 0x100 };

enum {
    RRTMGP_B200_OK = 0,
    RRTMGP_B200_E_ARG = 1,      /* bad dimensions / null pointer / unsupported combination  */
    RRTMGP_B200_E_RANGE = 2,    /* input outside the k-distribution's validity range        */
    RRTMGP_B200_E_MISSING = 3,  /* required gas missing (std::runtime_error in the reference) */
    RRTMGP_B200_E_CUDA = 4      /* CUDA runtime failure (reference: printf + throw 1)        */
};

/* Concurrency: a handle owns ONE workspace that every compute call re-carves, so the calls on one handle SERIALISE: the library
 * holds a per-handle mutex for the duration of a call (host threads) and makes the call's stream wait for an event recorded
 * after the previous call's last launch (different streams).  That is safe, not concurrent: for overlapping calls use one handle
 * per stream (handles are independent, tables are a few tens of MB).  Calls on different handles overlap freely;
 * rrtmgp_b200_profile's counters and rrtmgp_b200_last_error's message are process-wide and not thread-safe. */
typedef struct rrtmgp_b200_gas_optics rrtmgp_b200_gas_optics;     /* Gas_optics_rrtmgp_gpu  */
typedef struct rrtmgp_b200_cloud_optics rrtmgp_b200_cloud_optics; /* Cloud_optics_gpu       */
typedef struct rrtmgp_b200_aerosol_optics rrtmgp_b200_aerosol_optics; /* Aerosol_optics_gpu */

const char* rrtmgp_b200_last_error(void);
int rrtmgp_b200_version(void);

/* ---- Gas_optics_rrtmgp_gpu constructor / destructor / getters -------------------------------
 * replaces Gas_optics_rrtmgp_gpu::Gas_optics_rrtmgp_gpu (include/Gas_optics_rrtmgp.h:293-364,
 * src_cuda/Gas_optics_rrtmgp.cu:427-803; CPU twin src/Gas_optics_rrtmgp.cpp:391-742) fed with
 * the file variables read at src_test/Radiation_solver.cpp:78-243.  `gas_names_avail` is the
 * blank-padded [n_avail][string_len] list of gases present in the caller's Gas_concs. */
int rrtmgp_b200_gas_optics_create(const rrtmgp_b200_kdist_file* kd, int n_avail,
                                  const char* gas_names_avail, int string_len,
                                  int dtype, rrtmgp_b200_gas_optics** out);
void rrtmgp_b200_gas_optics_destroy(rrtmgp_b200_gas_optics* go);

/* include/Optical_props.h:191-200 and include/Gas_optics_rrtmgp.h:371-386 getters */
typedef struct rrtmgp_b200_gas_optics_info
{
    int dtype, is_longwave, ngpt, nbnd, nflav, ngas, ntemp, npres, neta, n_minor_lower,
        n_minor_upper, idx_h2o, gpt_width; /* gpt_width: padded g-points per band on device */
    double press_ref_min, press_ref_max, temp_ref_min, temp_ref_max, tsi;
} rrtmgp_b200_gas_optics_info;
int rrtmgp_b200_gas_optics_get_info(const rrtmgp_b200_gas_optics* go, rrtmgp_b200_gas_optics_info* info);
/* host copies; band_lims_gpt(2,nbnd) 1-based; band_lims_wvn(2,nbnd); gpoint_bands(ngpt) 1-based;
 * flavor(2,nflav); gpoint_flavor(2,ngpt); gas_names [ngas][string_len]; solar_source(ngpt) */
int rrtmgp_b200_gas_optics_get_tables(const rrtmgp_b200_gas_optics* go, int* band_lims_gpt,
                                      double* band_lims_wvn, int* gpoint_bands, int* flavor,
                                      int* gpoint_flavor, char* gas_names, int string_len,
                                      double* solar_source);

/* 64-bit hash over every prepared table of the handle (coefficients in device layout, index tables, reference grids,
 * retained gas names): equal hashes = built from bit-identical coefficient arrays and available-gas lists, whichever
 * constructor packed them.  HOST function. */
int rrtmgp_b200_gas_optics_table_hash(const rrtmgp_b200_gas_optics* go, unsigned long long* hash);

/* reduced minor-contributor descriptors of side 0 (lower) / 1 (upper): arrays of n_minor_* ints;
 * idx_minor / idx_minor_scaling as create_idx_minor[_scaling] (src/Gas_optics_rrtmgp.cpp:164-196),
 * kminor_start and limits(2,n) as reduce_minor_arrays (:54-162).  Any pointer may be NULL. */
int rrtmgp_b200_gas_optics_get_minor(const rrtmgp_b200_gas_optics* go, int side, int* idx_minor,
                                     int* idx_minor_scaling, int* kminor_start, int* limits_gpt,
                                     int* scales_with_density, int* scale_by_complement);

/* Gas concentrations handed to gas_optics: replaces Gas_concs_gpu (include/Gas_concs.h:73-91).
 * For gas i of the plan (order of get_tables' gas_names) d_vmr[i] points at a device array that
 * is a scalar (ncol_i=1,nlay_i=1), a profile (1,nlay) or a field (ncol,nlay).
 * The three arrays themselves are HOST arrays of length ngas. */
typedef struct rrtmgp_b200_gas_concs
{
    int ngas;
    const void* const* d_vmr;
    const int* vmr_ncol;
    const int* vmr_nlay;
} rrtmgp_b200_gas_concs;

/* Gas_optics_rrtmgp_gpu::get_col_dry (include/Gas_optics_rrtmgp.h:366, src_cuda/…cu:806-903) */
int rrtmgp_b200_get_col_dry(int dtype, int ncol, int nlay, const void* d_vmr_h2o /*(ncol,nlay)*/,
                            const void* d_plev /*(ncol,nlay+1)*/, void* d_col_dry, void* stream);

/* rrtmgp_interpolation + the vmr fill of compute_gas_taus (include/rrtmgp_kernels.h:34-58,
 * src/Gas_optics_rrtmgp.cpp:1103-1194; CUDA src_kernels_cuda/gas_optics_rrtmgp_kernels.cu:317-395).
 * Outputs in reference layouts; any output pointer may be NULL.  Index outputs are bit-exact
 * with the fp64 CPU path for the same input values in both precisions (index math is fp64). */
int rrtmgp_b200_interpolation(const rrtmgp_b200_gas_optics* go, int ncol, int nlay,
                              const void* d_play, const void* d_tlay, const void* d_col_dry,
                              const rrtmgp_b200_gas_concs* concs,
                              int* d_jtemp, int* d_jpress, signed char* d_tropo, int* d_jeta,
                              void* d_col_mix, void* d_fminor, void* d_fmajor, void* d_col_gas,
                              void* stream);

/* Gas_optics_rrtmgp_gpu::gas_optics, longwave (include/Gas_optics_rrtmgp.h:389-398,
 * src_cuda/Gas_optics_rrtmgp.cu:907-977 -> compute_gas_taus :980 + source :1167).
 * tau,lay_src (ncol,nlay,ngpt); lev_src (ncol,nlay+1,ngpt); sfc_src,sfc_src_jac (ncol,ngpt). */
int rrtmgp_b200_gas_optics_lw(const rrtmgp_b200_gas_optics* go, int ncol, int nlay,
                              const void* d_play, const void* d_plev, const void* d_tlay,
                              const void* d_tsfc, const void* d_tlev, const void* d_col_dry,
                              const rrtmgp_b200_gas_concs* concs,
                              void* d_tau, void* d_lay_src, void* d_lev_src, void* d_sfc_src,
                              void* d_sfc_src_jac, void* stream);

/* Gas_optics_rrtmgp_gpu::gas_optics, shortwave (include/Gas_optics_rrtmgp.h:401-408):
 * tau,ssa,g (ncol,nlay,ngpt) incl. Rayleigh + combine (src/Gas_optics_rrtmgp.cpp:365-385);
 * toa_src (ncol,ngpt) = solar_source (src/Gas_optics_rrtmgp.cpp:895-897). */
int rrtmgp_b200_gas_optics_sw(const rrtmgp_b200_gas_optics* go, int ncol, int nlay,
                              const void* d_play, const void* d_plev, const void* d_tlay,
                              const void* d_col_dry, const rrtmgp_b200_gas_concs* concs,
                              void* d_tau, void* d_ssa, void* d_g, void* d_toa_src, void* stream);

/* range validation of src/Gas_optics_rrtmgp.cpp:812-827 (the CPU path throws std::range_error;
 * the reference GPU path skips it).  Synchronises `stream`.  Returns E_RANGE with the reference's
 * message ("<name> is out of range") when any value is outside [lo, hi]. */
int rrtmgp_b200_check_range(int dtype, size_t n, const void* d_x, double lo, double hi,
                            const char* name, void* stream);

/* ---- Cloud_optics_gpu (include/Cloud_optics.h, src_cuda/Cloud_optics.cu:31-329; CPU
 * src/Cloud_optics.cpp:29-232).  Outputs are by band: (ncol,nlay,nband). */
int rrtmgp_b200_cloud_optics_create(const rrtmgp_b200_cloud_file* lut, int dtype,
                                    rrtmgp_b200_cloud_optics** out);
void rrtmgp_b200_cloud_optics_destroy(rrtmgp_b200_cloud_optics* co);
int rrtmgp_b200_cloud_optics_nband(const rrtmgp_b200_cloud_optics* co);
int rrtmgp_b200_cloud_optics_2str(const rrtmgp_b200_cloud_optics* co, int ncol, int nlay,
                                  const void* d_lwp, const void* d_iwp, const void* d_rel,
                                  const void* d_dei, void* d_tau, void* d_ssa, void* d_g, void* stream);
int rrtmgp_b200_cloud_optics_1scl(const rrtmgp_b200_cloud_optics* co, int ncol, int nlay,
                                  const void* d_lwp, const void* d_iwp, const void* d_rel,
                                  const void* d_dei, void* d_tau, void* stream);

/* ---- Aerosol_optics_gpu (include/Aerosol_optics.h:48-84, src_cuda/Aerosol_optics.cu:9-279; CPU
 * src/Aerosol_optics.cpp:10-224).  Eleven species in the reference's fixed order aermr01..aermr11
 * (SS1,SS2,SS3,DU1,DU2,DU3,OM2(philic),OM1(phobic),BC1,BC2,SU) with its fixed table columns
 * (src/Aerosol_optics.cpp:64-142).  Mixing ratios replace Aerosol_concs_gpu (= Gas_concs_gpu,
 * include/Aerosol_optics.h:20): species i is a device scalar (1,1), profile (1,nlay) or field
 * (ncol,nlay); fill_aerosols_3d's broadcast (src/Aerosol_optics.cpp:152-165) happens in the kernel.
 * Layer mass dpg = |plev(l) - plev(l+1)| / 9.81 as in the reference's GPU twin
 * (src_cuda/Aerosol_optics.cu:63); its CPU twin omits the abs (src/Aerosol_optics.cpp:53) and agrees
 * for the bottom-up columns its driver uses.  The humidity class is the first i with rh <= rh_upper(i)
 * (src/Aerosol_optics.cpp:27-38), clamped to the last class (the reference reads past the table).
 * Output by band: tau, ssa = taussa/max(tau,eps), g = taussag/max(taussa,eps), (ncol,nlay,nband). */
typedef struct rrtmgp_b200_aerosol_concs
{
    const void* d_mmr[11];     /* aermr01 .. aermr11 */
    int mmr_ncol[11];          /* 1 or ncol */
    int mmr_nlay[11];          /* 1 or nlay */
} rrtmgp_b200_aerosol_concs;
int rrtmgp_b200_aerosol_optics_create(const rrtmgp_b200_aerosol_file* lut, int dtype,
                                      rrtmgp_b200_aerosol_optics** out);
void rrtmgp_b200_aerosol_optics_destroy(rrtmgp_b200_aerosol_optics* ao);
int rrtmgp_b200_aerosol_optics_nband(const rrtmgp_b200_aerosol_optics* ao);
int rrtmgp_b200_aerosol_optics_2str(const rrtmgp_b200_aerosol_optics* ao, int ncol, int nlay,
                                    const rrtmgp_b200_aerosol_concs* concs, const void* d_rh,
                                    const void* d_plev, void* d_tau, void* d_ssa, void* d_g,
                                    void* stream);

/* ---- Optical_props: delta_scale / add_to (include/Optical_props.h:165-176,214-293;
 * kernels rte_delta_scale_2str_k, rte_increment_*, rte_inc_*_bybnd, include/rrtmgp_kernels.h:
 * 137-205; CUDA src_kernels_cuda/optical_props_kernels.cu:30-161).  In-place on op1. */
int rrtmgp_b200_delta_scale_2str(int dtype, int ncol, int nlay, int ngpt,
                                 void* d_tau, void* d_ssa, void* d_g, void* stream);
int rrtmgp_b200_increment_1scl(int dtype, int ncol, int nlay, int ngpt, void* d_tau1,
                               const void* d_tau2, void* stream);
int rrtmgp_b200_increment_2str(int dtype, int ncol, int nlay, int ngpt, void* d_tau1, void* d_ssa1,
                               void* d_g1, const void* d_tau2, const void* d_ssa2,
                               const void* d_g2, void* stream);
/* band_lims_gpt: HOST (2,nbnd) 1-based */
int rrtmgp_b200_increment_1scl_bybnd(int dtype, int ncol, int nlay, int ngpt, void* d_tau1,
                                     const void* d_tau2, int nbnd, const int* band_lims_gpt,
                                     void* stream);
int rrtmgp_b200_increment_2str_bybnd(int dtype, int ncol, int nlay, int ngpt, void* d_tau1,
                                     void* d_ssa1, void* d_g1, const void* d_tau2,
                                     const void* d_ssa2, const void* d_g2, int nbnd,
                                     const int* band_lims_gpt, void* stream);

/* ---- Rte_lw_gpu::rte_lw / Rte_sw_gpu::rte_sw (include/Rte_lw.h:62-77, include/Rte_sw.h:62-80).
 * expand_and_transpose (src/Rte_lw.cpp:204-217): (nbnd,ncol) -> (ncol,ngpt). */
int rrtmgp_b200_expand_and_transpose(int dtype, int ncol, int nbnd, int ngpt,
                                     const int* band_lims_gpt /*HOST (2,nbnd)*/,
                                     const void* d_in, void* d_out, void* stream);
/* rte_lw_solver_noscat (include/rrtmgp_kernels.h:207-232; Fortran mo_rte_solver_kernels.F90:
 * 248-367; CUDA src_kernels_cuda/rte_solver_kernels_launchers.cu:61-286).  n_gauss_angles 1..4
 * with the Gauss-Jacobi-5 secants/weights of src/Rte_lw.cpp:139-151 applied internally.
 * sfc_emis_gpt,sfc_src,inc_flux (ncol,ngpt) [inc_flux may be NULL = 0].
 * do_broadband=0: flux_up/dn are (ncol,nlay+1,ngpt); =1: (ncol,nlay+1) sums over g-points.
 * d_sfc_src_jac/d_flux_up_jac may be NULL (no Jacobian). */
int rrtmgp_b200_lw_solver_noscat(int dtype, int ncol, int nlay, int ngpt, int top_at_1,
                                 int n_gauss_angles, const void* d_tau, const void* d_lay_src,
                                 const void* d_lev_src, const void* d_sfc_emis_gpt,
                                 const void* d_sfc_src, const void* d_inc_flux,
                                 void* d_flux_up, void* d_flux_dn, int do_broadband,
                                 const void* d_sfc_src_jac, void* d_flux_up_jac, void* stream);
/* The same kernel with do_rescaling = true (include/rrtmgp_kernels.h:231-232; Fortran
 * mo_rte_solver_kernels.F90:148-178,206-214, lw_transport_1rescl :753-844; Tang et al. 2018): d_ssa, d_g
 * (ncol,nlay,ngpt) are the 2-stream properties; tau is scaled by 1 - ssa(1 - b), b = (1 - g)/2, and the up
 * sweep and a second down sweep carry the scattering adjustment.  The reference C++ passes false
 * (src/Rte_lw.cpp:178); this is the solver the Fortran front end uses for 2-stream LW properties. */
int rrtmgp_b200_lw_solver_rescaled(int dtype, int ncol, int nlay, int ngpt, int top_at_1,
                                   int n_gauss_angles, const void* d_tau, const void* d_ssa,
                                   const void* d_g, const void* d_lay_src, const void* d_lev_src,
                                   const void* d_sfc_emis_gpt, const void* d_sfc_src,
                                   const void* d_inc_flux, void* d_flux_up, void* d_flux_dn,
                                   int do_broadband, const void* d_sfc_src_jac, void* d_flux_up_jac,
                                   void* stream);
/* rte_sw_solver_2stream (include/rrtmgp_kernels.h:234-264; Fortran :503-609,985-1245).
 * mu0 (ncol); albedos, inc_flux_dir (ncol,ngpt); inc_flux_dif may be NULL. flux_dn includes
 * the direct beam on return, as in the reference. */
int rrtmgp_b200_sw_solver_2stream(int dtype, int ncol, int nlay, int ngpt, int top_at_1,
                                  const void* d_tau, const void* d_ssa, const void* d_g,
                                  const void* d_mu0, const void* d_sfc_alb_dir_gpt,
                                  const void* d_sfc_alb_dif_gpt, const void* d_inc_flux_dir,
                                  const void* d_inc_flux_dif, void* d_flux_up, void* d_flux_dn,
                                  void* d_flux_dir, int do_broadband, void* stream);
/* The kernel's own mu0 shape (ncol,nlay) (Fortran :517,1061-1065,1117-1125; src/Rte_sw.cpp:157-160 only
 * ever broadcasts a per-column value): a height-dependent solar zenith cosine on a round earth.  The top
 * layer's value scales the boundary condition, layers with mu0 <= 0 have no direct source, the surface
 * layer's sign gates the surface source.  As in the Fortran, the beam crosses dark layers with the nominal
 * mu0 = sqrt(epsilon(Real)), so the direct flux INSIDE dark layers depends on the build's precision. */
int rrtmgp_b200_sw_solver_2stream_mu0lay(int dtype, int ncol, int nlay, int ngpt, int top_at_1,
                                         const void* d_tau, const void* d_ssa, const void* d_g,
                                         const void* d_mu0_lay, const void* d_sfc_alb_dir_gpt,
                                         const void* d_sfc_alb_dif_gpt, const void* d_inc_flux_dir,
                                         const void* d_inc_flux_dif, void* d_flux_up, void* d_flux_dn,
                                         void* d_flux_dir, int do_broadband, void* stream);

/* ---- Fluxes_broadband_gpu / Fluxes_byband_gpu ::reduce (include/Fluxes.h:127-210; kernels
 * rte_sum_broadband, rte_net_broadband_precalc, sum_byband, net_byband_precalc,
 * include/rrtmgp_kernels.h:266-288; CUDA src_kernels_cuda/fluxes_kernels.cu:26-114). */
int rrtmgp_b200_sum_broadband(int dtype, int ncol, int nlev, int ngpt, const void* d_gpt_flux,
                              void* d_flux, void* stream);
int rrtmgp_b200_sum_byband(int dtype, int ncol, int nlev, int ngpt, int nbnd,
                           const int* band_lims_gpt /*HOST*/, const void* d_gpt_flux,
                           void* d_bnd_flux, void* stream);
int rrtmgp_b200_net_flux(int dtype, size_t n, const void* d_dn, const void* d_up, void* d_net,
                         void* stream);
/* x(icol, j) *= scale(icol), j < n2: the TSI scaling of the top-of-atmosphere source
 * (scaling_to_subset, src_test/Radiation_solver.cu:42-70). */
int rrtmgp_b200_scale_columns(int dtype, int ncol, int n2, void* d_x, const void* d_scale, void* stream);
/* Array_gpu<T,N>::fill and ::subset (include/Array.h:311-350 kernels, :515-524 and :579-622 members), type-erased:
 * elem_bytes = sizeof(T) in {1, 2, 4, 8, 16}.  fill: a[i] = *value (HOST pointer to one element), i < n.
 * subset: both arrays column-major; src has extents src_dims and index offsets src_offsets (NULL = zeros; element
 * (i1..iN) is at sum (i_k - offset_k - 1) * stride_k like the reference's calc_index); sub(j1..jN) = src(first + j - 1);
 * a source dimension of extent 1 is broadcast.  Ranges outside the source return E_ARG (the reference reads out of
 * bounds).  One stream-ordered kernel each, no synchronisation. */
int rrtmgp_b200_array_fill(int elem_bytes, long long n, void* d_a, const void* value /*HOST*/, void* stream);
int rrtmgp_b200_array_subset(int elem_bytes, int ndim, const int* src_dims /*HOST*/, const int* src_offsets /*HOST*/,
                             const int* first /*HOST, 1-based + offset*/, const int* sub_dims /*HOST*/,
                             const void* d_src, void* d_sub, void* stream);
/* heating rate (not in the C++ reference; rte-rrtmgp/extensions/mo_heating_rates.F90:44-73) */
int rrtmgp_b200_heating_rate(int dtype, int ncol, int nlay, const void* d_flux_up,
                             const void* d_flux_dn, const void* d_plev, void* d_hr, void* stream);

/* ---- Fused column solvers: one call = Radiation_solver_{longwave,shortwave}::solve_gpu for a
 * set of columns (src_test/Radiation_solver.cu:419-657, 690-950), i.e. gas_optics ->
 * [cloud_optics -> delta_scale -> add_to] -> [aerosol_optics -> delta_scale -> add_to] ->
 * rte_lw / rte_sw -> Fluxes::reduce, with the
 * per-g-point arrays kept on chip.  All inputs as the class API takes them. */
typedef struct rrtmgp_b200_columns
{
    int ncol, nlay;
    int top_at_1;                  /* -1: derive from play like Radiation_solver.cu:442 (syncs) */
    const void* d_play;            /* (ncol,nlay)   */
    const void* d_plev;            /* (ncol,nlay+1) */
    const void* d_tlay;            /* (ncol,nlay)   */
    const void* d_tlev;            /* (ncol,nlay+1) longwave only */
    const void* d_tsfc;            /* (ncol)        longwave only */
    const void* d_col_dry;         /* (ncol,nlay) or NULL -> get_col_dry from h2o */
    rrtmgp_b200_gas_concs concs;
    const void* d_sfc_emis;        /* LW: (nbnd,ncol) */
    const void* d_sfc_alb_dir;     /* SW: (nbnd,ncol) */
    const void* d_sfc_alb_dif;     /* SW: (nbnd,ncol) */
    const void* d_mu0;             /* SW: (ncol) */
    const void* d_tsi_scaling;     /* SW: (ncol) or NULL (=1) ; Radiation_solver.cu:744-746 */
    /* clouds (NULL cloud handle = clear sky) */
    const rrtmgp_b200_cloud_optics* cloud;
    int delta_cloud;               /* --delta-cloud, default on in the reference (SW only) */
    const void* d_lwp; const void* d_iwp; const void* d_rel; const void* d_dei; /* (ncol,nlay) */
    int n_gauss_angles;            /* LW, 1..4 (0 -> 1): Gauss-Jacobi-5 angles of src/Rte_lw.cpp:139-151 */
    /* aerosols, SW only (NULL handle = none): aerosol_optics -> [delta_scale] -> add_to after the
     * clouds, src_test/Radiation_solver.cu:794-810 */
    const rrtmgp_b200_aerosol_optics* aerosol;
    int delta_aerosol;             /* --delta-aerosol, default off in the reference */
    const void* d_rh;              /* (ncol,nlay) relative humidity */
    rrtmgp_b200_aerosol_concs aerosol_concs;
} rrtmgp_b200_columns;

typedef struct rrtmgp_b200_fluxes
{
    void* d_flux_up;       /* (ncol,nlay+1) */
    void* d_flux_dn;       /* (ncol,nlay+1) */
    void* d_flux_net;      /* (ncol,nlay+1) or NULL */
    void* d_flux_dn_dir;   /* SW (ncol,nlay+1) or NULL */
    void* d_bnd_flux_up;   /* (ncol,nlay+1,nbnd) or NULL */
    void* d_bnd_flux_dn;
    void* d_bnd_flux_net;
    void* d_bnd_flux_dn_dir;
    void* d_heating_rate;  /* (ncol,nlay) or NULL */
} rrtmgp_b200_fluxes;

int rrtmgp_b200_lw_fluxes(rrtmgp_b200_gas_optics* go, const rrtmgp_b200_columns* cols,
                          const rrtmgp_b200_fluxes* out, void* stream);
int rrtmgp_b200_sw_fluxes(rrtmgp_b200_gas_optics* go, const rrtmgp_b200_columns* cols,
                          const rrtmgp_b200_fluxes* out, void* stream);
/* Host-buffer form of the two calls above: every array pointer inside cols_* and out_* is a HOST
 * pointer here (same shapes and layouts; pinned memory gives full PCIe speed, pageable works).
 * The columns are processed in chunks (chunk_cols, 0 = one default tile); chunk i+1 is uploaded
 * and chunk i-1 downloaded while chunk i computes, so the copies hide behind the kernels.  Arrays
 * that LW and SW share (same host address: play, plev, tlay, vmr, col_dry) are uploaded once.
 * Either solver may be NULL.  Returns when the host outputs are complete (synchronises). */
int rrtmgp_b200_fluxes_host(rrtmgp_b200_gas_optics* go_lw, const rrtmgp_b200_columns* cols_lw,
                            const rrtmgp_b200_fluxes* out_lw, rrtmgp_b200_gas_optics* go_sw,
                            const rrtmgp_b200_columns* cols_sw, const rrtmgp_b200_fluxes* out_sw,
                            int chunk_cols, void* stream);
/* The column chunks rrtmgp_b200_fluxes_host cuts a call of ncol columns into (no GPU needed; for tests and for hosts that size
 * their pinned buffers): wave_columns > 0 = the automatic ramp for a GPU whose resident wave of CTAs covers that many columns
 * (1, 2, 4 waves at both ends, equal chunks of at most 8 waves between); wave_columns = 0 = equal chunks of chunk_columns.
 * Writes up to max_chunks (first column, columns) pairs and the number of chunks. */
int rrtmgp_b200_host_chunk_schedule(int ncol, int wave_columns, int chunk_columns, int max_chunks, int* first, int* count,
                                    int* nchunks);
/* tuning / introspection of the fused path: columns per tile (0 = default), and the number of
 * kernels the last fused call launched (for bench.py's gpu_launches). */
int rrtmgp_b200_set_tile_columns(rrtmgp_b200_gas_optics* go, int ncol_tile);
/* Debug aid standing in for compute-sanitizer (closed on this GPU pool).  flags = 1: before every fused call of this handle
 * the whole workspace (tile state, scratch strips, strip flags, temporaries) is filled with 0xFF bytes — NaN as float or
 * double — so a value read before it is written poisons the fluxes, and every carved piece is followed by a 256-byte canary
 * block; after the call the stream is synchronised and an overwritten canary fails the call with E_CUDA naming the piece.
 * Results must be bit-identical with and without it (tests/test_gpu_debug_arena.py).  flags = 3 additionally corrupts one
 * canary byte on purpose (self-test of the check: the call must fail).  flags = 0 switches it off. */
int rrtmgp_b200_set_debug(rrtmgp_b200_gas_optics* go, int flags);
/* which fused kernels a handle uses: 0 = per-band kernels with recurrence scratch strips in HBM (measured faster
 * on large domains), 1 = column-resident (recurrences in shared memory, every band of a tile in one launch, no atomics,
 * ~30x less DRAM traffic; measured 3.8x faster on 100-column calls), 2 = chosen per call by size (column-resident up to
 * 6144 columns), -1 = default: environment RB200_PATH=strips|colres when set, else by size.  All give the same fluxes to
 * rounding (each is compared with the oracle). */
int rrtmgp_b200_set_fused_path(rrtmgp_b200_gas_optics* go, int path);
long long rrtmgp_b200_launch_count(void);
/* Device-side timing of the fused band kernels: each call first returns (and resets) the time
 * and launch count accumulated since the previous call — CUDA events recorded on the launching
 * stream around every tile's band-kernel loop — then switches recording on/off.  Synchronises. */
int rrtmgp_b200_profile(int enable, double* lw_ms, long long* lw_launches, double* sw_ms,
                        long long* sw_launches);

/* Longwave AND shortwave fluxes of the same columns in one call: Radiation_solver_longwave::solve_gpu followed by
 * Radiation_solver_shortwave::solve_gpu (src_test/Radiation_solver.cu:419-657, 690-950; the reference driver calls them back to
 * back, src_test/test_rte_rrtmgp.cu:560-900) with the two advanced together tile by tile: band i of both k-distributions
 * runs in ONE launch whose CTAs alternate between the longwave and the shortwave role, so the L1-bound longwave work and the
 * DRAM-bound shortwave work share every SM.  Arguments as rrtmgp_b200_lw_fluxes / rrtmgp_b200_sw_fluxes; ncol and nlay must
 * agree; results are bit-identical to the two separate calls.  Single bands the paired kernel does not cover run through the
 * single-role launches; whole calls it does not cover (fp64, reduced g-point sets, several Gauss angles, different particulate
 * flags, small calls on the column-resident path) run as the two calls side by side: the shortwave one on an internal side stream
 * forked from and joined back into `stream`, so the caller still sees one stream-ordered call. */
int rrtmgp_b200_lwsw_fluxes(rrtmgp_b200_gas_optics* go_lw, const rrtmgp_b200_columns* cols_lw, const rrtmgp_b200_fluxes* out_lw,
                            rrtmgp_b200_gas_optics* go_sw, const rrtmgp_b200_columns* cols_sw, const rrtmgp_b200_fluxes* out_sw,
                            void* stream);
/* time and launches of the paired band loops collected by the last rrtmgp_b200_profile call (large calls run as two calls on two
 * streams: the wall time from fork to join, one count per call pair) */
int rrtmgp_b200_profile_pair(double* pair_ms, long long* pair_launches);

/* ---- Allocation-free variant.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that may not allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  The same computation and arguments as rrtmgp_b200_sw_solver_2stream with per-g-point
 * outputs (ncol,nlay+1,ngpt) and mu0 (ncol); the solver's four recurrence arrays live in a static device array of
 * RRTMGP_B200_NOALLOC_SCRATCH_BYTES inside the library (held by every GPU that loads it until the process exits)
 * instead of a stream-ordered allocation.  Calls no cudaMalloc / cudaFree and no allocating library routine.
 * Limit: ncol*nlay*ngpt <= RRTMGP_B200_NOALLOC_SCRATCH_BYTES / (4 * sizeof(Real)) = 2 097 152 cells (fp32) /
 * 1 048 576 (fp64); larger calls return E_ARG with the limit in the message.  Every call on a device shares the
 * array: calls that overlap on the GPU (different streams or threads) race — serialise them.  The array keeps the
 * previous call's contents; the kernel writes each scratch cell before reading it, so nothing needs clearing.  The
 * first call on a device resolves the array's device address (cudaGetSymbolAddress, not capturable): make one call
 * per device outside graph capture first. */
#define RRTMGP_B200_NOALLOC_SCRATCH_BYTES (32u * 1024u * 1024u)
int rrtmgp_b200_sw_solver_2stream_noalloc(int dtype, int ncol, int nlay, int ngpt, int top_at_1,
                                          const void* d_tau, const void* d_ssa, const void* d_g, const void* d_mu0,
                                          const void* d_sfc_alb_dir_gpt, const void* d_sfc_alb_dif_gpt,
                                          const void* d_inc_flux_dir, const void* d_inc_flux_dif /*may be NULL*/,
                                          void* d_gpt_flux_up, void* d_gpt_flux_dn, void* d_gpt_flux_dir, void* stream);

/* ---- Multi-GPU: columns are independent, so N processes (one per GPU) each run the calls
 * above on a contiguous shard of the x-fastest flattened domain (src_test/test_rte_rrtmgp.cpp:217);
 * the library holds no communicator and makes no collective call.  This is the partition rule
 * (shards differ by at most one column; HOST function, no device needed): rank r of `world` owns
 * columns [*c0, *c1).  The host model gathers the flux slabs with its own MPI / NCCL;
 * rte_rrtmgp_cpp_b200/shard.py does it with torch.distributed for bench.py. */
int rrtmgp_b200_shard_bounds(long long ncol, int world, int rank, long long* c0, long long* c1);

/* Shard plan + flux gather over NCCL (SURVEY 8b "shard_plan_create(ncol, ngpu, ncclComm_t)", "gather_fluxes"; the reference
 * itself is single-GPU).  One plan per process (= per GPU).  The communicator is either the host model's own ncclComm_t
 * (nccl_comm, borrowed) or created here from the 128-byte NCCL unique id that rank 0 obtains with
 * rrtmgp_b200_nccl_unique_id and hands to every rank by whatever means it has (MPI_Bcast, torch.distributed, a file).
 * NCCL is resolved with dlopen("libnccl.so.2") on first use (RB200_NCCL_LIB overrides); world == 1 needs no NCCL at all. */
typedef struct rrtmgp_b200_shard_plan rrtmgp_b200_shard_plan;
int rrtmgp_b200_nccl_unique_id(void* id128 /*HOST, 128 bytes out*/);
int rrtmgp_b200_shard_plan_create(long long ncol, int world, int rank, const void* id128 /*HOST or NULL*/,
                                  void* nccl_comm /*ncclComm_t or NULL*/, rrtmgp_b200_shard_plan** out);
void rrtmgp_b200_shard_plan_destroy(rrtmgp_b200_shard_plan* plan);
/* columns [*c0, *c1) of `rank` (rank < 0: this process) — the rule of rrtmgp_b200_shard_bounds */
int rrtmgp_b200_shard_plan_bounds(const rrtmgp_b200_shard_plan* plan, int rank, long long* c0, long long* c1);
/* Gather flux slabs to `root` in the reference layout.  Every rank calls it with the same (nlev, nfields, part, nparts, root).
 * A rank's shard may be processed in `nparts` column blocks (the rule of rrtmgp_b200_shard_bounds applied to the shard):
 * d_part_fields[f] is this rank's (count_of_part, nlev) column-major slab of field f for block `part`; on the root
 * d_full_fields[f] is the whole-domain (ncol, nlev) array (ignored elsewhere, may be NULL).  Default: one ncclSend / ncclRecv
 * per field and peer (grouped) into a staging buffer the plan keeps on the root, then one strided device copy per (peer, field)
 * into the reference layout; RB200_GATHER=direct receives every level straight into its place instead (no staging copy, but
 * nlev x nfields small messages per peer: measured 3x slower).  The root's own columns are one strided device copy per field.
 * Stream-ordered, returns without synchronising (the staging buffer grows, synchronising, only while sizes still increase).
 * Calling it per block on a side stream overlaps the gather with the next block's kernels. */
/* Peer-memory mode of the gather (optional, collective, blocking; call once at set-up from every rank with the same nfields and
 * root).  The root hands the CUDA-IPC handles of its whole-domain arrays to the other processes, which map them; from then
 * on rrtmgp_b200_gather_fluxes with the same root / nfields / arrays makes every rank WRITE its slab straight into the root's
 * reference-layout arrays over NVLink (one strided peer-to-peer copy per field, all ranks at once) and signal completion with a
 * one-element message: no staging buffer and no layout copy on the root.  All or nothing: when any rank cannot map the arrays
 * (no peer access, memory not from cudaMalloc) nothing is registered and the gather keeps the NCCL path; returns OK either
 * way — rrtmgp_b200_shard_plan_registered tells (number of registered fields, 0 = none).  d_full_fields is read on the root
 * only; the arrays must stay allocated while registered; registering again replaces the registration. */
int rrtmgp_b200_shard_plan_register(rrtmgp_b200_shard_plan* plan, int nfields, void* const* d_full_fields /*HOST array, root only*/,
                                    int root, void* stream);
int rrtmgp_b200_shard_plan_registered(const rrtmgp_b200_shard_plan* plan);
int rrtmgp_b200_gather_fluxes(rrtmgp_b200_shard_plan* plan, int dtype, int nlev, int nfields,
                              const void* const* d_part_fields /*HOST array of device pointers*/,
                              void* const* d_full_fields /*HOST array of device pointers, root only*/,
                              int part, int nparts, int root, void* stream);

#ifdef __cplusplus
}
#endif
#endif

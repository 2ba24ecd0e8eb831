/* dartray_engine.h — C ABI of the B200 ray-tracing engine that replaces the
 * DART-Ray hot path.
 *
 * The reference has no FFI: the path sits behind four argument-less Fortran
 * procedures whose state is module variables.  Each entry point below names the
 * reference procedure (file:line under /root/reference) it replaces.  All
 * arrays are caller-owned, column-major exactly as the Fortran arrays
 * (first index fastest), 0-based cell/wavelength/pixel ids, fp64 reals,
 * int32 integers except `cindex` (int64).
 *
 *   dre_run_pass   <- rt_loop            rt_routines.f90:985-1058 (+ main_dir_loop
 *                                        :1343-1467, ray_tracing :1618-1857,
 *                                        deposit :2362-2557, ray_list.f90)
 *   dre_run_iobs   <- rt_loop_iobs       rt_routines.f90:1470-1603
 *   dre_allreduce  <- rt_mpi + reduce_*  io_routines.f90:3667-3881
 *   dre_prepare_scatt <- rt_prepare(rtt_scatt) + calc_total_luminosity_sca
 *                                        dartray_hub.f90:432-457, rt_routines.f90:663-758
 *   dre_dust_emission <- set_dust_emission ('eff')  sed_routines.f90:249-276, 377-430
 *   dre_run_pass_2d   <- rt_loop_2D + fix_symmetry       rt_routines.f90:1061-1140, 3605-4025
 *
 * Return value: 0 on success, non-zero on error (message via dre_last_error);
 * the Fortran wrapper calls stop_prc (smooth_grid_routines.f90:869) on non-zero.
 * Not re-entrant; one engine per GPU / MPI rank; call from the master thread
 * outside any OpenMP region.
 */
#ifndef DARTRAY_ENGINE_H
#define DARTRAY_ENGINE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* rt_type ids, rt_routines.f90:193-231 */
enum {
  DRE_RTT_PRECALC_CELL   = 1,
  DRE_RTT_PRECALC_SRC    = 2,
  DRE_RTT_DIR_CELL       = 5,
  DRE_RTT_I_OBS_DIR_CELL = 6,
  DRE_RTT_DIR_SRC        = 7,
  DRE_RTT_I_OBS_DIR_SRC  = 8,
  DRE_RTT_SCATT          = 10,
  DRE_RTT_I_OBS          = 11
};

#define DRE_NCCL_ID_BYTES 128

typedef struct dre_engine dre_engine;

/* Scalars of the path (module variables of rt_routines.f90). */
typedef struct dre_params {
  int32_t bm_par;        /* rt_routines.f90:16 (caller passes bm_par_sca during scattering, dartray_hub.f90:442) */
  int32_t bm_par_max;    /* :20 */
  int32_t pabs_max;      /* :22, = int(10*modelsize) io_routines.f90:678 */
  int32_t print_psel_av; /* :268 */
  double  rad_lim;       /* :24 */
  double  en_lim;        /* :26, set_en_lim :3137 */
  double  cs;            /* :121, set_units :5192 */
  double  modelsize;     /* smooth_grid_routines.f90 */
  int32_t wall_on[3];    /* x_wall_on.. rt_routines.f90:499-515 */
  int32_t reserved0;
  double  wall_coord[6]; /* x(1:2), y(1:2), z(1:2) in model units (after set_walls :2600) */
} dre_params;

/* Event counters of one pass (the roofline is computed from these). */
typedef struct dre_counters {
  int64_t sources;        /* sources processed (leaf, luminous) */
  int64_t rays;           /* ray_tracing calls */
  int64_t intersections;  /* iterations of the main loop rt_routines.f90:1689 */
  int64_t deposits;       /* intersections reaching :2445 / :2499 */
  int64_t beam_tests;     /* intersections entering the too-coarse branch :2473 */
  int64_t splits;         /* rays sent to the high list */
  int64_t merges;         /* rays sent to the low list */
  int64_t sca_elems;      /* scaspe elements updated (sum of P(lambda) over deposits) */
  int64_t active_lambda;  /* sum over deposits+beam tests of active wavelengths */
  int64_t steps;          /* resolution steps executed (wavefront launches) */
  double  kernel_ms;      /* device time of the ray-march kernel(s), CUDA events */
  double  pass_ms;        /* device time of the whole pass */
  uint64_t isect_hash;    /* order-independent hash of (unit, cell, isel, psel bits) */
  int64_t active_lambda_dep; /* sum over deposits only of active wavelengths */
  int64_t descent_loads;  /* cchild loads of ccindd_to_cc (sum of descent depths) */
  int64_t launches;       /* CUDA kernels launched by the pass */
  int64_t trace_launches; /* of which ray-march kernels */
} dre_counters;

/* One recorded intersection (debug/parity mode, dre_set_record). */
typedef struct dre_isect {
  int32_t unit;    /* source_ordinal*48 + sector (i_obs passes: source_ordinal*ndir + j) */
  int32_t step;    /* resolution step within the unit's sector loop */
  int32_t ray;     /* ordinal of the ray within (unit, step), reference processing order */
  int32_t k;       /* intersection ordinal along the ray */
  int32_t cell;    /* cc */
  int32_t isel;    /* 1,2,3 */
  double  psel;
} dre_isect;

/* ---- lifetime ---------------------------------------------------------- */
int  dre_get_unique_id(void* id128);  /* ncclGetUniqueId, rank 0 only; initialize_mpi io_routines.f90:3610 */
int  dre_create(dre_engine** out, int device, int rank, int nranks, const void* nccl_id128);
void dre_destroy(dre_engine*);
const char* dre_last_error(const dre_engine*);

/* ---- static inputs (copied to HBM) -------------------------------------- */
/* grid arrays of smooth_grid_routines.f90:41-70, csize_arr/carea_arr/cvol_arr io_routines.f90:669-696 */
int dre_set_grid(dre_engine*, int32_t tot_ncell, int32_t max_lvl, const int32_t base[2],
                 const int32_t* cchild, const int32_t* lvl, const int64_t* cindex,
                 const double* ccoord /*(3,N)*/, const double* csize /*(N)*/,
                 const double* csize_arr /*(0:max_lvl)*/, const double* carea_arr, const double* cvol_arr);
/* dens_arr smooth_grid_routines.f90:62; ksca_arr_norm/gsca_arr slices at i0 (set_i_opacity_arrays
 * rt_routines.f90:5074); kp_sca_arr rt_routines.f90:185 (-1 = isotropic); dir_i_out (ndir,2) */
int dre_set_opacity(dre_engine*, int32_t lnum, const double* dens_arr /*(L,N)*/,
                    const double* ksca_arr_norm /*(L)*/, const double* gsca_arr /*(L)*/,
                    const int32_t* kp_sca_arr /*(L)*/, int32_t tot_ndir_scaspe,
                    const double* dir_i_out /*(ndir,2): theta col, phi col*/);
/* lumcell :188, point sources :94,:167,:182,:185 (prepare_p_src rt_routines.f90:535) */
int dre_set_sources(dre_engine*, const double* lumcell /*(L,N) or NULL*/, int32_t tot_p_src,
                    const double* ccoord_p_src /*(3,S)*/, const double* lum_p_src_arr /*(L,S)*/,
                    const int32_t* src_cell /*(N)*/, const int32_t* cell_src /*(S)*/);
int dre_set_params(dre_engine*, const dre_params*);

/* ---- device-resident state (mirrors of u_fest_arr, u_final_arr, scaspe_arr,
 *      scaspe_prev_arr, scaspe_tot_arr, lum_lost, psel_av_arr) ----------------- */
enum { DRE_ARR_U_FEST = 0, DRE_ARR_U_FINAL = 1, DRE_ARR_SCASPE = 2, DRE_ARR_SCASPE_PREV = 3,
       DRE_ARR_SCASPE_TOT = 4, DRE_ARR_LUM_LOST = 5, DRE_ARR_PSEL_AV = 6, DRE_ARR_LUMCELL = 7,
       DRE_ARR_I_OBS = 8, DRE_ARR_I_OBS_IN = 9 };
/* (L,N) arrays: `planes` = pointer to the array.  scaspe arrays: `planes` = array of L
 * pointers, plane l is (P(l),N) pixel-fastest, as scaspe_arr(l)%a. */
int dre_upload  (dre_engine*, int which, const void* planes);
int dre_download(dre_engine*, int which, void* planes);
int dre_zero    (dre_engine*, int which);

/* ---- passes -------------------------------------------------------------- */
/* rt_loop on device-resident state; accumulates into u_fest (precalc) or u_final+scaspe+lum_lost. */
int dre_run_pass_resident(dre_engine*, int rt_type, int iterations, dre_counters* out);
/* rt_loop_2D (rt_routines.f90:1061-1140), the 2D mode for models with the three mirror symmetries (dartray_main_2D
 * dartray_hub.f90:69-126; rt_type = rtt_precalc_cell, rtt_dir_cell or rtt_scatt): the sources of the positive octant
 * (skip_cell_2D_rt_loop :1174-1187), fix_symmetry (:3605-4025: every non-negative leaf cell receives what its seven mirror images
 * got -- find_linked_cells :3364-3578, scaspe rows mirrored pixel by pixel, calc_scaspe_indices :4083-4217 -- and lum_lost is
 * scaled by 8), then the sources on the Cartesian planes.  Device-resident state; needs tot_ndir_scaspe = 0 (the reference's
 * symmetrisation covers the HEALPix part of scaspe only; dartray_main_2D computes no i_obs).  Counters: both loops added. */
int dre_run_pass_2d(dre_engine*, int rt_type, int iterations, dre_counters* out);
/* the mirrored-pixel table of calc_scaspe_indices for nside 2^kp; which = 0..6 = X, Y, Z, XY, XZ, YZ, XYZ; out (12*4^kp) */
int dre_scaspe_symmetry_index(dre_engine*, int32_t kp, int32_t which, int32_t* out);
/* rt_loop with HOST arrays: upload -> pass -> download (the drop-in call). */
int dre_run_pass(dre_engine*, int rt_type, int iterations,
                 const double* u_fest /*(L,N) in*/, double* u_accum /*(L,N) inout*/,
                 double* const* scaspe /*L ptrs inout or NULL*/, const double* const* scaspe_prev /*in or NULL*/,
                 double* lum_lost /*(L) inout or NULL*/, double* psel_av /*(N+S) out or NULL*/,
                 dre_counters* out);
/* rt_loop_iobs; i_obs (L, N+S, ndir) += , i_obs_in (L, N+S, ndir_in) += */
int dre_run_iobs(dre_engine*, int rt_type, const double* const* scaspe_tot /*L ptrs or NULL*/,
                 int32_t tot_ndir_in, const double* ccoord_obs /*(3,ndir_in)*/,
                 double* i_obs, double* i_obs_in, dre_counters* out);
/* rt_mpi: combine the per-rank accumulators (NCCL all-reduce, fp64 sum). */
int dre_allreduce(dre_engine*, int rt_type);
/* rt_prepare(rtt_scatt): u_fest<-u_final, lum_lost_prev<-lum_lost, scaspe_prev<-scaspe,
 * lumcell(l,c)=sum_{p<npix_hp} scaspe, tot_rad_en at iteration 1, convergence flag,
 * sequential: scaspe_tot += scaspe (iterations>1), scaspe = 0. */
int dre_prepare_scatt(dre_engine*, int iterations, double conv_en_lim,
                      double* tot_rad_en /*(L) inout*/, int32_t* cnflag /*out*/);

/* calc_sed / sum_i_obs (sed_routines.f90:90-146, 212-246) on the i_obs array left on the device by dre_run_iobs:
 * sed(l, j) = sum over cells and point sources of i_obs(l, c, j) * conv(l) * (csize / dist_obs)^2  [Jy].
 * units: 0 'erg/s/Hz/pc^2/sr', 1 'W/Hz/pc^2/sr', 2 'W/m/pc^2/sr' (lambda_si (L) required).  sed is (lnum, ndir). */
int dre_calc_sed(dre_engine*, double dist_obs, int units, const double* lambda_si /*(L) or NULL*/,
                 double* sed /*(L,ndir) out*/);

/* make_maps / map_projection (visual_routines.f90:59-126, 570-772) + convert_maps_to_MJy_sr (:180-246) on the i_obs array
 * left on the device by dre_run_iobs: external-observer surface-brightness maps of every wavelength and line of sight
 * (dir_i_out of dre_set_opacity), npixel_maps x npixel_maps pixels over map_size_factor * modelsize, in MJy/sr.
 * maps is map_arr_out(0:npixel_maps-1, 0:npixel_maps-1, 0:lnum-1, 0:tot_ndir-1), column-major (X' pixel fastest).
 * units as for dre_calc_sed. */
int dre_make_maps(dre_engine*, int npixel_maps, double map_size_factor, int units, const double* lambda_si /*(L) or NULL*/,
                  double* maps /*(npix,npix,L,ndir) out*/);

/* map_in_projection (visual_routines.f90:917-1076, calc_cube_texture_sphere :1080-1182) + convert_maps_to_MJy_sr on the i_obs_in
 * array left on the device by dre_run_iobs (same observers): every leaf cell is projected on the observer's HEALPix sphere of
 * nside 2**kp_maps.  maps_in is map_in_arr_out(0:12*4**kp_maps-1, 0:lnum-1, 0:tot_ndir_in-1), column-major, MJy/sr. */
int dre_make_maps_in(dre_engine*, int32_t ndir_in, const double* ccoord_obs /*(3,ndir_in)*/, int kp_maps, int units,
                     const double* lambda_si /*(L) or NULL*/, double* maps_in /*(npix_maps,L,ndir_in) out*/);

/* set_dust_emission, dust_heating_type 'eff' (sed_routines.f90:249-276: calc_dens_dustem :377-430, calc_t_dust_equil :846-874,
 * zbrent_tdust :993-1078, bplanck :939-948, set_units_dens_stars_arr :904-914): the dust-emission luminosity density of every
 * dusty leaf cell from the radiation field it absorbs.  The engine must be configured for the dust-emission wavelengths
 * (dre_set_opacity with lnum = lnum_dust).
 *   kabs_stars (lnum_stars) = kabs_arr(0:lnum_stars-1), kabs_dust (lnum_dust) = kabs_arr(i0:), lambda_si_dust = lambda_arr_SI(i0:),
 *   dl_stars / dl_dust = delta_lambda_bin_stars / _dust, rf_stars (lnum_stars, N) = u_final_uv_opt after convert_ufield_ifield,
 *   rf_dust (lnum_dust, N) = u_final_arr/parsec^3*cspeed, or NULL to read the device-resident u_final of the last dust pass
 *   (ignored when iterations_dustem == 0, sed_routines.f90:857).
 * Out: dens_stars (lnum_dust, N) [W/m/pc^3], t_dust (N) or NULL; to_lumcell != 0 also leaves lumcell = dens_stars * csize^3 on the
 * device (calc_total_luminosity rt_routines.f90:605-660), ready for the next rtt_precalc_cell / rtt_dir_cell pass. */
int dre_dust_emission(dre_engine*, int32_t lnum_stars, int32_t lnum_dust, const double* kabs_stars, const double* kabs_dust,
                      const double* lambda_si_dust, const double* dl_stars, const double* dl_dust, double kext_ref,
                      const double* dens_ref /*(N)*/, const double* rf_stars /*(lnum_stars,N)*/, const double* rf_dust /*(lnum_dust,N) or NULL*/,
                      int32_t iterations_dustem, double* dens_stars /*(lnum_dust,N) out*/, double* t_dust /*(N) out or NULL*/,
                      int32_t to_lumcell);

/* ---- separate small kernels (north star: cell-list and bounding-box kernels) */
/* leaf && sum(lum)>0 gate of rt_loop rt_routines.f90:1016; returns the list length */
int dre_build_source_list(dre_engine*, int rt_type, int32_t* n_sources, int32_t* list_out /*(N) or NULL*/);
/* find_cc_new2 rt_routines.f90:1962: locate points; cell_out=-1 when outside the model */
int dre_locate_points(dre_engine*, int32_t npts, const double* xyz /*(3,npts)*/,
                      const int32_t* cc_old /*(npts)*/, int32_t* cell_out /*(npts)*/);
/* dre_locate_points for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).
 * No counterpart in the reference.  Same arguments and results as
 * dre_locate_points; the device staging lives in a static 2-MiB device array of the library, so npts <= 65536
 * (DRE_NOALLOC_MAX_POINTS; larger calls are rejected with an error) and no cudaMalloc / cudaFree is called.  Every call on a
 * GPU shares that array: calls that overlap on one GPU race. */
int dre_locate_points_noalloc(dre_engine*, int32_t npts, const double* xyz /*(3,npts)*/,
                              const int32_t* cc_old /*(npts)*/, int32_t* cell_out /*(npts)*/);
/* pix2ang_nest + calc_cproj (healpix_routines.f90:30, rt_routines.f90:1945) */
int dre_pix2vec(dre_engine*, int32_t nside, int32_t npix, const int32_t* ipix,
                double* theta, double* phi, double* vec /*(3,npix)*/);

/* ---- parity / debug ------------------------------------------------------- */
int dre_set_record(dre_engine*, int64_t max_records);   /* 0 disables */
int dre_get_records(dre_engine*, dre_isect* out, int64_t cap, int64_t* n);
/* Tuning options (none changes a result beyond the summation order) and one mode switch:
 * "deterministic" = 1: order-independent accumulation.  The north star asks for deposits "written in a fixed order so the
 * result is deterministic"; the reference itself is not (OpenMP dynamic schedule + atomic updates, rt_routines.f90:994,
 * 2667-2684).  Instead of fixing the order, every contribution to u_fest / u_final / scaspe / lum_lost / psel_av is first
 * rounded to a per-wavelength grid q(lambda) = 2^-52 x (a-priori bound of the sum), which makes every fp64 partial sum
 * exact: the accumulators are bit-identical for any schedule, unit batching, reduction path and number of GPUs.  Cost:
 * absolute accuracy ~sqrt(n) q / 2 per element after n contributions (dre_get_quanta returns q).  The grids are running
 * maxima over the passes made since dre_set_grid / dre_set_opacity, so the bits are reproducible between runs that make the
 * same sequence of calls (any schedule, any number of ranks), not between the first and a repeated solve on one engine. */
int dre_set_option(dre_engine*, const char* name, double value);
/* quanta of the last deterministic pass: q_out[0..L) scaspe, [L..2L) u_final / u_fest, [2L..3L) lum_lost (0 = no grid) */
int dre_get_quanta(dre_engine*, double* q_out /*(3L)*/);
/* evaluates the ray-march kernel's exp() (attenuation, deposit rt_routines.f90:2380-2381) on n host values */
int dre_eval_exp(dre_engine*, int32_t n, const double* x, double* y);
/* timing hooks: CUDA events on the engine's stream (print_time io_routines.f90:4268 is wall clock) */
int dre_mark(dre_engine*, int slot /*0..7*/);
int dre_elapsed_ms(dre_engine*, int slot_a, int slot_b, double* ms /*out; synchronises*/);
int64_t dre_launch_count(const dre_engine*);  /* kernels launched since dre_create */

#ifdef __cplusplus
}
#endif
#endif /* DARTRAY_ENGINE_H */

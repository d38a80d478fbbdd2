/* gmcore_gpu.h — C ABI of the B200 dynamics hot path (drop-in for GMCORE's
 * time_integrate and the pieces of gmcore_run around it).
 *
 * Every entry point cites the reference routine it replaces (paths relative to
 * the GMCORE source tree).  All arrays are plain double pointers to the
 * reference's halo'd Fortran storage, so a Fortran host passes
 * c_loc(state%u_lon) unchanged:
 *   lon  : 1-hx : nlon+hx                       (hx = mesh%lon_halo_width)
 *   lat  : lat_ibeg-hy : lat_iend+hy  (full)    (hy = mesh%lat_halo_width = 2)
 *          lat_ibeg-hy : half_iend+hy (half; half_iend = lat_iend, or
 *                                      lat_iend-1 on the band owning the north pole)
 *   lev  : -1 : nlev+2 (full levels), 1 : nlev+1 (half levels)
 * Halo values supplied by the host are ignored; the library regenerates them.
 * Return value: 0 on success, non-zero on error (text via gmc_last_error),
 * mirroring log_error/process_stop in the reference.
 */
#ifndef GMCORE_GPU_H
#define GMCORE_GPU_H

#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct gmc_handle gmc_handle;

enum { GMC_EULER = 0, GMC_PC2 = 1, GMC_WRFRK3 = 2 };           /* time_schemes_mod.F90:67-76 */
enum { GMC_PV_MIDPOINT = 0, GMC_PV_UPWIND = 1, GMC_PV_TVD = 2 }; /* operators_mod.F90:58-69 ('ffsl' is not functional upstream) */
enum { GMC_TVD_NONE = 0, GMC_TVD_VAN_LEER = 1, GMC_TVD_MC = 2 }; /* tvd_limiter_type, adv/tvd_mod.F90:27-40 */
enum { GMC_BIG_FILTER = 0, GMC_SMALL_FILTER1 = 1, GMC_SMALL_FILTER2 = 2 }; /* block_mod.F90:62-64 */
enum { GMC_FORCING_NONE = 0, GMC_FORCING_HELD_SUAREZ = 1 };    /* test_forcing_mod.F90:14-28 */
enum { GMC_LIMITER_NONE = 0, GMC_LIMITER_MONO = 1, GMC_LIMITER_PD = 2 }; /* limiter_type, ffsl_mod.F90:131-140 */
enum { GMC_FLUX_PPM = 0, GMC_FLUX_VAN_LEER = 1 };              /* ffsl_flux_type, ffsl_mod.F90:116-129 */

/* The namelist keys and constants the hot path reads
 * (namelist_mod.F90:10-122, const_mod.F90:19-33). */
typedef struct gmc_config {
  int nlon, nlat, nlev;          /* num_lon, num_lat, num_lev (global)                 */
  int hx, hy;                    /* halo widths of the HOST arrays                      */
  int lat_ibeg, lat_iend;        /* full-lat rows owned by this band, 1-based inclusive */
  int device;                    /* CUDA device ordinal                                 */
  int rank, nranks;              /* position in the latitude-band chain                 */
  int baroclinic;                /* num_lev > 1 (hydrostatic core) or 0 (shallow water) */
  int time_scheme;               /* GMC_EULER / GMC_PC2 / GMC_WRFRK3                    */
  int ke_scheme;                 /* 1 or 2                                              */
  int pv_scheme;                 /* GMC_PV_*                                            */
  int upwind_order_pv;           /* 1 or 3                                              */
  int pv_pole_stokes;
  int coriolis_scheme;           /* 1 or 2                                              */
  int use_div_damp, div_damp_order, div_damp_k0;
  int use_smag_damp;
  int test_forcing;              /* GMC_FORCING_*                                       */
  int ntracers;                  /* tracers in the single 'moist'-like batch (0 = none) */
  int limiter_type;              /* GMC_LIMITER_* : slope of the sub-grid profile       */
  int ffsl_flux_type;            /* GMC_FLUX_* : sub-grid profile of the FFSL fluxes    */
  int tvd_limiter_type;          /* GMC_TVD_* : flux limiter of pv_scheme = GMC_PV_TVD  */
  double ke_cell_wgt, upwind_wgt_pv;
  double div_damp_coef2, div_damp_coef4, div_damp_top, div_damp_pole, div_damp_lat0;
  double smag_damp_coef;
  double dt_dyn, dt_adv;
  double omega, radius, g, Rd, cpd, p0, Rd_o_cpd;
  double ptop, hybrid_p0;        /* hybrid_coord_mod.F90:146-171                        */
} gmc_config;

/* Per-latitude coefficient vectors of mesh_type (mesh_mod.F90:84-110), GLOBAL,
 * with the reference's bounds: full-lat vectors 1-hy : nlat+hy (nlat+2hy values),
 * half-lat vectors 1-hy : nlat-1+hy (nlat-1+2hy values); the (2,:) arrays are
 * Fortran column-major (2, lb:ub).  half_lon is 1-hx : nlon+hx.
 * Level vectors: hyai/hybi/half_lev/half_dlev 1 : nlev+1; full_lev/full_dlev
 * -1 : nlev+2 (NULL for shallow water). */
typedef struct gmc_mesh_vectors {
  const double *full_lat, *half_lat, *full_cos_lat, *half_cos_lat, *full_sin_lat, *half_sin_lat;
  const double *area_cell, *area_lon, *area_lon_west, *area_lon_east, *area_lon_north, *area_lon_south;
  const double *area_lat, *area_lat_west, *area_lat_east, *area_lat_north, *area_lat_south, *area_vtx;
  const double *area_subcell;                  /* (2, full) */
  const double *de_lon, *le_lon, *de_lat, *le_lat;
  const double *full_f, *half_f;
  const double *full_tangent_wgt, *half_tangent_wgt; /* (2, full), (2, half) */
  const double *half_lon;
  const double *hyai, *hybi, *full_lev, *half_lev, *full_dlev, *half_dlev;
} gmc_mesh_vectors;

/* gmcore_init (gmcore_mod.F90:45-95): allocate device state for one band. */
int gmc_create(const gmc_config* cfg, gmc_handle** h);
void gmc_destroy(gmc_handle* h);

/* mesh_type vectors (mesh_mod.F90:134-447) */
int gmc_set_mesh(gmc_handle* h, const gmc_mesh_vectors* mv);

/* filter_type (filter_types_mod.F90:49-147): ngrid_lon/wgt_lon over full-lat
 * rows 1-hy:nlat+hy, ngrid_lat/wgt_lat over half-lat rows; wgt_* are (ld, rows). */
int gmc_set_filters(gmc_handle* h, int which, const int* ngrid_lon, const double* wgt_lon, int ld_lon,
                    const int* ngrid_lat, const double* wgt_lat, int ld_lat);

/* static_type%gzs (static_mod.F90:14-26), (lon,lat) halo'd */
int gmc_set_static(gmc_handle* h, const double* gzs);

/* state(itime)%u_lon, v_lat, pt|gz, phs (restart_mod.F90:82-95 defines this set).
 * itime is 1..3 as in block%state(:).  phs is NULL for shallow water. */
int gmc_upload_state(gmc_handle* h, int itime, const double* u_lon, const double* v_lat,
                     const double* pt_or_gz, const double* phs);
int gmc_download_state(gmc_handle* h, int itime, double* u_lon, double* v_lat,
                       double* pt_or_gz, double* phs);

/* adv_batches(1)%q(:,:,:,l,old) (adv_mod.F90:390), (lon,lat,lev,ntracers) */
int gmc_set_tracers(gmc_handle* h, const double* q);
int gmc_get_tracers(gmc_handle* h, double* q);

/* More advection batches (block%adv_batches(:)): adv_add_tracer + adv_allocate_tracers (adv_mod.F90:314-418).
 * Batch 0 is the one the config struct describes (ntracers, dt_adv); each call adds a 'cell' batch of `ntracers`
 * tracers advected every `dt_adv` seconds (nstep = dt_adv / dt_dyn wind samples, adv_batch_mod.F90:73-149,239-252)
 * and returns its index.  Call after gmc_create and before gmc_prepare_run, on every band.  adv_run
 * (adv_mod.F90:165-312) then loops over the batches in this order.  q is (lon,lat,lev,ntracers) of that batch. */
int gmc_add_tracer_batch(gmc_handle* h, int ntracers, double dt_adv, int* batch);
int gmc_set_batch_tracers(gmc_handle* h, int batch, const double* q);
int gmc_get_batch_tracers(gmc_handle* h, int batch, double* q);

/* The pre-loop block of gmcore_run (gmcore_mod.F90:101-136): filters -> u_f,
 * v_f, phs_f|gz_f; copy_old_m; adv_accum_wind; operators_prepare(all). */
int gmc_prepare_run(gmc_handle* h, int old_idx);

/* time_integrate (gmcore_mod.F90:577-591): time_integrator (pc2|wrfrk3|euler,
 * time_schemes_mod.F90:244-288) + test_forcing_run + damp_run + adv_run.
 * Asynchronous on the handle's stream.  time_step is the host's step counter
 * before this step (adv_run skips while it is 0, adv_mod.F90:166). */
int gmc_time_integrate(gmc_handle* h, int old_idx, int new_idx, double dt, int time_step);

/* Convenience for drivers that do not need per-step host control: n steps of
 * { time_integrate; time_advance (swap old/new); [diagnose] } as in the loop of
 * gmcore_run (gmcore_mod.F90:138-144).  On return *old_idx is the index holding
 * the newest state.  If diag is non-NULL it receives 8 doubles per step. */
int gmc_run_steps(gmc_handle* h, int* old_idx, int* new_idx, int* time_step, int nsteps, double* diag);

/* diagnose (gmcore_mod.F90:213-341): out = tm, te, tav, tpe, te_ke, te_ie, te_pe, tpt */
int gmc_diagnose(gmc_handle* h, int itime, double out[8]);
/* The same sums without a host round trip per step: gmc_diagnose_enqueue computes the sums of state(itime) into a
 * device ring (asynchronous, no all-reduce yet); gmc_diagnose_collect reduces every queued set over the bands,
 * returns them oldest first in out[nsets][8] and reports a raised guard like gmc_diagnose.  For hosts that look at
 * the sums only when the 'print' alert rings (gmcore_mod.F90:140-142). */
int gmc_diagnose_enqueue(gmc_handle* h, int itime);
int gmc_diagnose_collect(gmc_handle* h, double* out, int max_sets, int* nsets);
/* gmc_diagnose without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same sums as gmc_diagnose of a single-band handle, same out[8] order and guard behaviour; the block sums and
 * the eight results live in ONE file-scope `static __device__` array of GMC_STATIC_SCRATCH_DOUBLES doubles (4 MiB)
 * that every GPU loading the library holds until exit, instead of the handle's cudaMalloc'd buffers.  The call
 * allocates nothing (no cudaMalloc / cudaFree / allocating library routine, no host heap).  It needs
 * 32 + 7 * lat_rows * nlev doubles and returns non-zero with a text in gmc_last_error for larger grids (up to
 * 0.25 deg x 64 levels fit) and for band handles (nranks > 1, linked groups, NCCL).  Every call on a GPU shares the
 * array, which keeps the last call's contents: the used part is cleared first, and calls that overlap on one GPU
 * (several handles / streams / host threads) race -- serialise them. */
#define GMC_STATIC_SCRATCH_DOUBLES (512 * 1024)
int gmc_diagnose_static(gmc_handle* h, int itime, double out[8]);

/* Read back one named derived field of state(itime) for parity tests/history
 * (history_mod.F90 h0 variable set and the per-state work arrays).  buf uses the
 * halo'd host layout of that field's grid location.  Returns non-zero for an
 * unknown name. */
int gmc_get_field(gmc_handle* h, int itime, const char* name, double* buf);

/* Pressure-level diagnostics of the h1 / diag_state output (utils/diag_state_mod.F90:241-258): one variable of
 * state(itime) -- "u", "v" (averaged to the cell centre), "pt", "t", "z" (gz_lev / g) -- interpolated to the pressure
 * surface po, linearly in p or in log p, exactly as interp_*_to_pressure_level (dynamics/interp_mod.F90:727-897).
 * y2d is a (lon, lat) array in the halo'd host layout; columns that do not reach po get `fill`.  z is bracketed
 * with ph_lev (the reference passes the full-level pressure into the half-level dummy there -- an index slip that
 * is not reproduced). */
int gmc_interp_pressure_level(gmc_handle* h, int itime, const char* var, double po, int logp, double fill, double* y2d);

/* Latitude bands (the 1-D latitude decomposition of process_mod.F90:93-139; halo exchange =
 * fill_halo S/N of parallel_mod.F90:501-527).  Each band is one handle created with its own
 * lat_ibeg/lat_iend, rank, nranks (hydrostatic core).  Either link all bands of a single process
 * (same device) with gmc_group_link and drive them through hs[0], or run one process per GPU:
 * rank 0 calls gmc_nccl_unique_id, the host broadcasts the 128 bytes (MPI_Bcast in GMCORE),
 * every rank calls gmc_nccl_init; diagnose then returns globally reduced sums on every rank. */
int gmc_group_link(gmc_handle** hs, int n);
int gmc_nccl_unique_id(char id[128]);
int gmc_nccl_init(gmc_handle* h, const char id[128]);

/* Advection-only driver (drivers/gmcore_adv_driver.F90:84-88): gmc_adv_prescribe is the set_uv of
 * tests/adv/*_test_mod.F90 (state(itime)%m = 1, u_f / v_f given, mfx_lon = u_f, mfy_lat = v_f); gmc_adv_step is
 * adv_run(itime) (adv_mod.F90:152-312).  time_step as in gmc_time_integrate. */
int gmc_adv_prescribe(gmc_handle* h, int itime, const double* u_f, const double* v_f);
int gmc_adv_prescribe_we(gmc_handle* h, int itime, const double* we_lev /* (lon, lat, half lev) halo'd: state%we_lev of dcmip12_test_set_uv */);
int gmc_adv_step(gmc_handle* h, int itime, int time_step);

/* time_init (time_mod.F90:61-98): clear the step / tracer-batch sample counters before a fresh upload. */
int gmc_reset_clock(gmc_handle* h);

int gmc_synchronize(gmc_handle* h);

/* Page-lock (cudaHostRegister) / release a host array that is uploaded or downloaded repeatedly -- the arrays of
 * state_type allocated once in state_mod.F90:103-199.  Optional: pageable arrays work, at a fraction of the PCIe
 * rate (the transfers of gmc_upload_state / gmc_download_state then go through the driver's bounce buffer). */
int gmc_pin_host(gmc_handle* h, void* ptr, size_t bytes);
int gmc_unpin_host(gmc_handle* h, void* ptr);
/* Device milliseconds spent in, and kernels launched by, calls on this handle
 * since the last reset (CUDA events on the handle's stream). */
/* Verification hook for the positive-argument pow / log the Exner function and the hydrostatic integral use on the
 * device (gmcore_b200/csrc/fastmath.cuh, replacing the ** and log of calc_ph / calc_t / calc_gz_lev,
 * operators_mod.F90:136-190,433-469): out[i] = pow_pos(x[i], y) (what = 0) or log_pos(x[i]) (what = 1), evaluated on
 * the current device from host arrays.  The same source compiles for the host, where it is held to < 1 ulp against
 * libquadmath; the GPU test checks that the device returns the host's bits. */
int gmc_fastmath_eval(int what, const double* x, double y, double* out, int n);
int gmc_launch_count(gmc_handle* h, long long* launches);
int gmc_last_error(gmc_handle* h, char* buf, int len);

#ifdef __cplusplus
}
#endif
#endif /* GMCORE_GPU_H */

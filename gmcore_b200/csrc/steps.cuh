// Stage sequencing: step_forward_backward + update_state
// (src/dynamics/time_schemes_mod.F90:102-242) and space_operators
// (src/dynamics/gmcore_mod.F90:467-571) as a fixed list of fused kernels.
#pragma once
#include "handle.cuh"
#include "kernels_dyn.cuh"
#include "kernels_update.cuh"
#include "kernels_hydro.cuh"
#include "halo.cuh"

static inline dim3 grid3(const gmc_handle* h, int nk) {
  return dim3((h->g.nlon + 127) / 128, h->g.je - h->g.jb + 1, nk);
}
static inline size_t filt_smem(const gmc_handle* h) { return (size_t)(2 * (h->g.nlon + 2 * h->max_hn) + 2 * (2 * h->max_hn + 1)) * sizeof(double); }

// operators_prepare(forward_pass) for the horizontal part: calc_mf, calc_ke, calc_div,
// calc_pv, interp_pv (operators_mod.F90:117-122)
static void run_forward_prepare(gmc_handle* h, DState& S, double dt /* interp_pv_tvd only */) {
  const Geo& g = h->g;
  Fwd1Args f{S.u_f, S.v_f, S.m, S.mfx_lon, S.mfy_lat, S.u_lat, S.v_lon, S.ke, S.pv, S.div, h->cfg.ke_scheme, h->cfg.ke_cell_wgt};
  LAUNCH(h, k_fwd1, tile_grid(g, 1), tile_block(), 0, g, h->M, f);
  Pole1Args p{S.v_f, S.m, S.u_lat, S.mfy_lat, S.ke, S.div, S.pv, h->pole_dmf, h->cfg.pv_pole_stokes};
  LAUNCH(h, k_pole1, dim3(g.nlev, 2), 256, 0, g, h->M, p);
  PvEdgeArgs e{S.pv, S.u_f, S.v_f, S.u_lat, S.v_lon, S.pv_lon, S.pv_lat, h->cfg.pv_scheme, h->cfg.upwind_order_pv, h->cfg.upwind_wgt_pv, h->M.rows, dt, h->cfg.tvd_limiter_type};
  LAUNCH(h, k_pv_edge, tile_grid(g, g.nlev), tile_block(), 0, g, e);
}

static void run_filter(gmc_handle* h, const double* in, int half_lat, int nlev, int f1, double* out1, int f2, double* out2,
                       int post = POST_NONE, const double* mdiv = nullptr, double* out_m = nullptr, const Geo* gp = nullptr) {
  const Geo& g = gp ? *gp : h->g;
  FilterArgs a;
  a.in = in; a.half_lat = half_lat; a.nlev = nlev;
  a.f1 = h->filt[f1 >= 0 ? f1 : 0]; a.out1 = out1; a.use_f1 = f1 >= 0;
  a.f2 = h->filt[f2 >= 0 ? f2 : 0]; a.out2 = out2; a.use_f2 = f2 >= 0;
  a.post = post; a.gzs = h->gzs; a.grav = h->cfg.g; a.mdiv = mdiv; a.out_m = out_m;
  a.chain = 0; a.row2_off = 0; a.skip_a = a.skip_b = -1000; a.mh = h->max_hn;
  const int nrows = half_lat ? (g.hje - g.hjb + 1) : (g.je - g.jb + 1);
  LAUNCH(h, k_filter_rows, dim3(nlev, nrows), 256, filt_smem(h), g, a);
}

// out2 = f2(in), out1 = f1(out2) in one launch (one read of the raw row instead of two passes)
static void run_filter_chain(gmc_handle* h, const double* in, int half_lat, int nlev, int f2, double* out2, int f1, double* out1,
                             int skip_a = -1000, int skip_b = -1000, const Geo* gp = nullptr) {
  const Geo& g = gp ? *gp : h->g;
  FilterArgs a;
  a.in = in; a.half_lat = half_lat; a.nlev = nlev;
  a.f1 = h->filt[f1]; a.out1 = out1; a.use_f1 = 1;
  a.f2 = h->filt[f2]; a.out2 = out2; a.use_f2 = 1;
  a.post = POST_NONE; a.gzs = h->gzs; a.grav = h->cfg.g; a.mdiv = nullptr; a.out_m = nullptr;
  a.chain = 1; a.row2_off = g.nlon + 2 * h->max_hn; a.skip_a = skip_a; a.skip_b = skip_b; a.mh = h->max_hn;
  const int nrows = half_lat ? (g.hje - g.hjb + 1) : (g.je - g.jb + 1);
  if (nrows <= 0) return;
  LAUNCH(h, k_filter_rows, dim3(nlev, nrows), 256, filt_smem(h), g, a);
}

#include "hydro.cuh"

// One shallow-water step(old, star, new, dt) (time_schemes_mod.F90:102-118).
static void swm_step(gmc_handle* h, DState& old_, DState& star, DState& nw, double dt) {
  const Geo& g = h->g;
  run_forward_prepare(h, star, dt);
  TendArgs t{};
  t.mfx_lon = star.mfx_lon; t.mfy_lat = star.mfy_lat; t.pv_lon = star.pv_lon; t.pv_lat = star.pv_lat; t.ke = star.ke;
  t.we_lev = nullptr; t.du = h->du; t.dv = h->dv;
  t.gz_old = old_.gz; t.pole_dmf = h->pole_dmf; t.gz_new = nw.gz; t.dt = dt; t.grav = h->cfg.g;
  t.coriolis_scheme = h->cfg.coriolis_scheme; t.swm = 1;
  LAUNCH(h, k_tend, tile_grid(g, 1), tile_block(), 0, g, h->M, t);
  // update_state(tend1): gz_f = big(gz); gz = small1(gz); backward prepare: m = (gz - gzs)/g
  run_filter(h, nw.gz, 0, g.nlev, GMC_BIG_FILTER, nw.gz_f, GMC_SMALL_FILTER1, nw.gz, POST_SWM_M, nullptr, nw.m);
  // backward pass + update_state(tend2)
  UvArgs u{};
  u.gz_f = nw.gz_f; u.du_fwd = h->du; u.dv_fwd = h->dv; u.u_old = old_.u; u.v_old = old_.v; u.u_new = nw.u; u.v_new = nw.v;
  u.dt = dt; u.swm = 1; u.blend = 1;
  LAUNCH(h, k_uv_update, grid3(h, g.nlev), 128, 0, g, h->M, u);
  run_filter(h, nw.u, 0, g.nlev, GMC_BIG_FILTER, nw.u_f, -1, nullptr);
  run_filter(h, nw.v, 1, g.nlev, GMC_BIG_FILTER, nw.v_f, -1, nullptr);
}

// One step(old, star, new, dt) on every band of this process, then the stage exchange.
static int do_step(gmc_handle* h, int old_, int star, int nw, double dt) {
  if (!h->cfg.baroclinic) { swm_step(h, h->st[old_], h->st[star], h->st[nw], dt); return 0; }
  for_bands(h, [&](gmc_handle* b) { hydro_stage(b, b->st[old_], b->st[star], b->st[nw], dt); });
  if (h->comm_stream) {
    // the edge rows of new%{u_f, v_f, pt, phs_f} are final at ev_edge: the exchange runs on comm_stream next to
    // the interior rows of the last filter launches; the compute stream rejoins before it touches halo rows
    // (band group: every band's edge event; all bands share the leader's compute stream)
    cudaStream_t s0 = h->stream;
    for_bands(h, [&](gmc_handle* b) { cudaStreamWaitEvent(h->comm_stream, b->ev_edge, 0); });
    for_bands(h, [&](gmc_handle* b) { b->stream = h->comm_stream; });
    const int rc = halo_exchange(h, [&](gmc_handle* b) { return stage_halo_set(b, nw); });
    cudaEventRecord(h->ev_xdone, h->comm_stream);
    for_bands(h, [&](gmc_handle* b) { b->stream = s0; });
    cudaStreamWaitEvent(s0, h->ev_xdone, 0);
    if (rc) return 1;
  } else if (halo_exchange(h, [&](gmc_handle* b) { return stage_halo_set(b, nw); })) return 1;
  for_bands(h, [&](gmc_handle* b) { hydro_after_exchange(b, b->st[nw]); });
  return 0;
}

// pc2 / wrfrk3 / euler (time_schemes_mod.F90:244-288)
static int time_integrator(gmc_handle* h, int old_, int nw, double dt) {
  switch (h->cfg.time_scheme) {
    case GMC_EULER: return do_step(h, old_, old_, nw, dt);
    case GMC_PC2:
      return do_step(h, old_, old_, nw, dt / 2.0) || do_step(h, old_, nw, 3, dt / 2.0) || do_step(h, old_, 3, nw, dt);
    default:
      return do_step(h, old_, old_, nw, dt / 3.0) || do_step(h, old_, nw, 3, dt / 2.0) || do_step(h, old_, 3, nw, dt);
  }
}

// C ABI, run-time half: gmc_prepare_run, gmc_time_integrate, gmc_run_steps,
// gmc_diagnose (src/dynamics/gmcore_mod.F90:97-149,213-341,577-591).
#pragma once
#include "steps.cuh"

static int check_ready(gmc_handle* h) {
  if (!h->mesh_set || !h->filters_set[0] || !h->filters_set[1] || !h->filters_set[2]) {
    snprintf(h->err, sizeof(h->err), "gmc_set_mesh / gmc_set_filters must be called first");
    return 1;
  }
  return 0;
}

static int launch_diag(gmc_handle* h, int itime, double* dev_out) {
  const Geo& g = h->g; DState& S = h->st[itime];
  DiagArgs a{S.m, S.mfx_lon, S.mfy_lat, S.u, S.v, S.pv, S.pv_lon, S.pv_lat, S.t, S.pt, S.phs, h->gzs,
             h->diag_partial, h->cfg.g, h->cfg.cpd, h->cfg.baroclinic};
  const int nrows = g.je - g.jb + 1;
  LAUNCH(h, k_diag_partial, dim3(nrows, g.nlev), 256, 0, g, h->M, a);
  LAUNCH(h, k_diag_final, 8, 1024, 0, h->diag_partial, nrows * g.nlev, dev_out, h->dflags);
  return 0;
}
// device sums (tm, te_ke, te_ie, te_pe, tav, tpe, tpt) -> ABI order
static void diag_to_abi(const double* s, double* out) {
  out[0] = s[0]; out[1] = s[1] + s[2] + s[3]; out[2] = s[4]; out[3] = s[5];
  out[4] = s[1]; out[5] = s[2]; out[6] = s[3]; out[7] = s[6];
}

// time_integrate (gmcore_mod.F90:577-591) for every band driven by this handle
static int time_integrate_impl(gmc_handle* h, int old_, int nw, double dt, int time_step) {
  if (time_integrator(h, old_, nw, dt)) return 1;
  if (!h->cfg.baroclinic) {      // damp_run (damp_mod.F90:49-63): div damping, then Smagorinsky
    if (h->cfg.use_div_damp) run_div_damp(h, h->st[nw]);
    if (h->cfg.use_smag_damp) run_smag_damp(h, h->st[nw], dt);
    return 0;
  }
  bool adv = false;
  for_bands(h, [&](gmc_handle* b) { hydro_post_compute_1(b, nw, dt); });
  if (h->cfg.nranks > 1 && h->cfg.use_smag_damp &&      // strain rates read u, v one row beyond the band
      halo_exchange(h, [&](gmc_handle* b) { HaloSet s = halo_set(); halo_add(s, b->st[nw].u, b->g.nlev); halo_add(s, b->st[nw].v, b->g.nlev); return s; })) return 1;
  for_bands(h, [&](gmc_handle* b) { adv = hydro_post_compute_2(b, nw, dt, time_step); });
  if (h->cfg.nranks > 1) {
    // one packed exchange per advected batch (the alert clocks are the same on every band), pt with the first
    bool hs = h->cfg.test_forcing == GMC_FORCING_HELD_SUAREZ;
    for (size_t m = 0; m < h->batches.size(); ++m) {
      if (!h->batches[m].advected) continue;
      const bool with_pt = hs; hs = false;
      if (halo_exchange(h, [&](gmc_handle* b) { HaloSet s = tracer_halo_set(b, (int)m, 1); if (with_pt) halo_add(s, b->st[nw].pt, b->g.nlev); return s; })) return 1;
    }
    if (hs && halo_exchange(h, [&](gmc_handle* b) { HaloSet s = halo_set(); halo_add(s, b->st[nw].pt, b->g.nlev); return s; })) return 1;
  }
  (void)adv;
  for_bands(h, [&](gmc_handle* b) { hydro_adv_finish(b, nw); });
  return 0;
}
// Shallow-water time_integrate through a captured graph (no host-side state changes inside a shallow-water step:
// fixed pointers, fixed dt).  Falls back to plain launches while profiling, for non-standard buffer pairings, or when
// capture is unavailable.
static int swm_time_integrate(gmc_handle* h, int old_, int nw, double dt, int time_step) {
  static const bool off = getenv("GMCORE_NO_GRAPH") && getenv("GMCORE_NO_GRAPH")[0] == '1';
  const bool pair = (old_ == 1 && nw == 2) || (old_ == 2 && nw == 1);
  if (off || h->prof.on || !pair || dt != h->cfg.dt_dyn) return time_integrate_impl(h, old_, nw, dt, time_step);
  const int gi = old_ - 1;
  if (!h->swm_graph[gi]) {
    cudaGraph_t graph = nullptr;
    const long long l0 = h->launches;
    if (cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) { cudaGetLastError(); return time_integrate_impl(h, old_, nw, dt, time_step); }
    const int rc = time_integrate_impl(h, old_, nw, dt, time_step);
    const cudaError_t e = cudaStreamEndCapture(h->stream, &graph);
    h->swm_graph_launches[gi] = h->launches - l0; h->launches = l0;
    if (rc || e != cudaSuccess || !graph || cudaGraphInstantiate(&h->swm_graph[gi], graph, 0) != cudaSuccess) {
      cudaGetLastError(); if (graph) cudaGraphDestroy(graph); h->swm_graph[gi] = nullptr;
      snprintf(h->err, sizeof(h->err), "shallow-water step: graph capture failed"); return 1;
    }
    cudaGraphDestroy(graph);
  }
  CUDA_CHECK(cudaGraphLaunch(h->swm_graph[gi], h->stream));
  h->launches += h->swm_graph_launches[gi];
  return 0;
}
// time_advance bookkeeping of the tracer alerts (time_mod.F90:128-159), every band, every batch
static void alert_advance(gmc_handle* h, int time_step) {
  for_bands(h, [&](gmc_handle* b) {
    for (TrBatch& T : b->batches) if (T.alert_ring) { T.alert_last = time_step * b->cfg.dt_dyn; T.alert_ring = false; }
  });
}

extern "C" {

int gmc_prepare_run(gmc_handle* h, int old_) {
  if (check_ready(h)) return 1;
  if (old_ < 1 || old_ > 3) { snprintf(h->err, sizeof(h->err), "bad time index %d", old_); return 1; }
  if (!h->cfg.baroclinic) {
    const Geo& g = h->g; DState& S = h->st[old_];
    run_filter(h, S.u, 0, g.nlev, GMC_BIG_FILTER, S.u_f, -1, nullptr);
    run_filter(h, S.v, 1, g.nlev, GMC_BIG_FILTER, S.v_f, -1, nullptr);
    // gz_f = big(gz); calc_m: m = (gz - gzs) / g
    run_filter(h, S.gz, 0, g.nlev, GMC_BIG_FILTER, S.gz_f, -1, nullptr, POST_SWM_M, nullptr, S.m);
    run_forward_prepare(h, S, h->cfg.dt_dyn);
  } else {
    for_bands(h, [&](gmc_handle* b) {
      DState& S = b->st[old_];
      run_filter(b, S.u, 0, b->g.nlev, GMC_BIG_FILTER, S.u_f, -1, nullptr);
      run_filter(b, S.v, 1, b->g.nlev, GMC_BIG_FILTER, S.v_f, -1, nullptr);
      hydro_prepare_run_1(b, old_);
    });
    if (halo_exchange(h, [&](gmc_handle* b) { return prepare_halo_set(b, old_); })) return 1;
    for (size_t m = 0; m < h->batches.size(); ++m)
      if (halo_exchange(h, [&](gmc_handle* b) { return tracer_halo_set(b, (int)m, 0); })) return 1;
    for_bands(h, [&](gmc_handle* b) { hydro_prepare_run_2(b, old_); });
  }
  CUDA_CHECK(cudaGetLastError());
  return 0;
}

int gmc_time_integrate(gmc_handle* h, int old_, int nw, double dt, int time_step) {
  if (check_ready(h)) return 1;
  alert_advance(h, time_step - 1);     // the host's time_advance ran since the previous call
  if (h->cfg.baroclinic ? time_integrate_impl(h, old_, nw, dt, time_step) : swm_time_integrate(h, old_, nw, dt, time_step)) return 1;
  CUDA_CHECK(cudaGetLastError());
  return 0;
}

// Global sums over all bands: single-process group -> host sum in band order;
// one process per GPU -> ncclAllReduce of the partial sums (global_sum, parallel_mod.F90:805-816).
// Slot 7 of every set is the guard word (k_diag_final): a non-zero value anywhere means one of the reference's
// log_error guards fired on some band -> the call fails with the guard text, as the reference stops in
// log_error / process_stop, and the device words are cleared once reported.
static int reduce_diag(gmc_handle* h, std::vector<double*>& dbufs, int nsets, double* out_abi) {
  std::vector<double> acc((size_t)nsets * 8, 0.0), tmp((size_t)nsets * 8);
  if (h->nccl_comm) {
    NcclApi& n = nccl_api();
    if (n.AllReduce(dbufs[0], dbufs[0], (size_t)nsets * 8, GMC_NCCL_DOUBLE, GMC_NCCL_SUM, h->nccl_comm, h->stream)) {
      snprintf(h->err, sizeof(h->err), "ncclAllReduce failed"); return 1; }
  }
  for (double* d : dbufs) {
    CUDA_CHECK(cudaMemcpyAsync(tmp.data(), d, tmp.size() * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CUDA_CHECK(cudaStreamSynchronize(h->stream));
    for (size_t q = 0; q < acc.size(); ++q) acc[q] += tmp[q];
  }
  int guards = 0;
  for (int s = 0; s < nsets; ++s) {
    if (out_abi) diag_to_abi(&acc[(size_t)s * 8], out_abi + (size_t)s * 8);
    double v = acc[(size_t)s * 8 + 7];
    for (int b = 0; b < 5 && v > 0; ++b) { const double r = fmod(v, 256.0); if (r > 0) guards |= 1 << b; v = (v - r) / 256.0; }
  }
  if (guards) {
    hydro_flags_text(h, guards);
    for_bands(h, [&](gmc_handle* b) { if (b->dflags) cudaMemsetAsync(b->dflags, 0, sizeof(int), b->stream); });
    return 1;
  }
  return 0;
}

// grow-only per-band device buffer for the per-step sums of gmc_run_steps (no cudaMalloc inside a timed loop)
static double* diag_steps_buf(gmc_handle* b, int nsteps) {
  if (b->diag_steps_cap < nsteps) {
    const int cap = nsteps < 64 ? 64 : nsteps;
    double* nb = nullptr;
    if (cudaMalloc(&nb, (size_t)cap * 8 * sizeof(double)) != cudaSuccess) return nullptr;
    if (b->diag_steps) {      // keep the sets already queued by gmc_diagnose_enqueue
      cudaMemcpyAsync(nb, b->diag_steps, (size_t)b->diag_steps_cap * 8 * sizeof(double), cudaMemcpyDeviceToDevice, b->stream);
      cudaStreamSynchronize(b->stream);
      cudaFree(b->diag_steps);
    }
    b->diag_steps = nb; b->diag_steps_cap = cap;
  }
  return b->diag_steps;
}

int gmc_diagnose(gmc_handle* h, int itime, double out[8]) {
  if (check_ready(h)) return 1;
  std::vector<double*> dbufs;
  for_bands(h, [&](gmc_handle* b) { launch_diag(b, itime, b->diag_out); dbufs.push_back(b->diag_out); });
  return reduce_diag(h, dbufs, 1, out);
}

// Deferred diagnostics: the reference computes `diagnose` every step but only looks at the sums when the 'print'
// alert rings (gmcore_mod.F90:140-142, log_print_diag).  gmc_diagnose_enqueue launches the sums of state(itime) into
// the next slot of a device ring without touching the host; gmc_diagnose_collect reduces all queued sets over the
// bands (one ncclAllReduce), copies them back in one transfer and checks the guard word.  Same numbers as
// gmc_diagnose per step, without a host round trip and an all-reduce latency inside every time step.
int gmc_diagnose_enqueue(gmc_handle* h, int itime) {
  if (check_ready(h)) return 1;
  int rc = 0;
  const int slot = h->diag_queued;
  for_bands(h, [&](gmc_handle* b) {
    double* d = diag_steps_buf(b, slot + 1 > 64 ? 2 * (slot + 1) : 64);      // grow-only; existing sets are kept
    if (!d) { rc = 1; return; }
    launch_diag(b, itime, d + (size_t)slot * 8);
  });
  if (rc) { snprintf(h->err, sizeof(h->err), "cudaMalloc of the diagnostics ring failed"); return 1; }
  h->diag_queued = slot + 1;
  return 0;
}
int gmc_diagnose_collect(gmc_handle* h, double* out /* [max_sets][8] */, int max_sets, int* nsets) {
  if (check_ready(h)) return 1;
  const int n = h->diag_queued;
  if (nsets) *nsets = n;
  if (n == 0) return 0;
  if (n > max_sets) { snprintf(h->err, sizeof(h->err), "gmc_diagnose_collect: %d sets queued, room for %d", n, max_sets); return 1; }
  std::vector<double*> dbufs;
  for_bands(h, [&](gmc_handle* b) { dbufs.push_back(b->diag_steps); });
  h->diag_queued = 0;
  return reduce_diag(h, dbufs, n, out);
}

int gmc_run_steps(gmc_handle* h, int* old_idx, int* new_idx, int* time_step, int nsteps, double* diag) {
  if (check_ready(h)) return 1;
  if (nsteps <= 0) return 0;
  if (h->diag_queued) { snprintf(h->err, sizeof(h->err), "gmc_run_steps: %d deferred diagnostics are queued, call gmc_diagnose_collect first", h->diag_queued); return 1; }
  std::vector<double*> dbufs;
  int rc = 0;
  // with diag == NULL one set is still taken after the last step: it carries the guard word back
  const int nsets = diag ? nsteps : 1;
  for_bands(h, [&](gmc_handle* b) {
    double* d = diag_steps_buf(b, nsets);
    if (!d) { rc = 1; snprintf(h->err, sizeof(h->err), "cudaMalloc of the diagnostics buffer failed"); }
    dbufs.push_back(d);
  });
  for (int s = 0; s < nsteps && !rc; ++s) {
    rc = h->cfg.baroclinic ? time_integrate_impl(h, *old_idx, *new_idx, h->cfg.dt_dyn, *time_step)
                           : swm_time_integrate(h, *old_idx, *new_idx, h->cfg.dt_dyn, *time_step);
    alert_advance(h, *time_step);   // time_advance (time_mod.F90:128-159)
    int t = *old_idx; *old_idx = *new_idx; *new_idx = t;
    *time_step += 1;
    if (diag || s == nsteps - 1) { int bi = 0; for_bands(h, [&](gmc_handle* b) { launch_diag(b, *old_idx, dbufs[bi++] + (size_t)(diag ? s : 0) * 8); }); }
  }
  if (!rc && cudaGetLastError() != cudaSuccess) { snprintf(h->err, sizeof(h->err), "run_steps: CUDA launch error"); rc = 1; }
  if (!rc) rc = reduce_diag(h, dbufs, nsets, diag);
  return rc;
}

// Back to the state right after gmc_create + set_*: step counter bookkeeping of the tracer batch and the
// alert clock (time_init, time_mod.F90:61-98).  The host then uploads a state and calls gmc_prepare_run again.
int gmc_reset_clock(gmc_handle* h) {
  for_bands(h, [&](gmc_handle* b) {
    for (TrBatch& T : b->batches) {
      T.uv_step = T.mf_step = T.we_step = 0; T.q_old = 0; T.q_new = 1;
      T.alert_last = -T.dt; T.alert_ring = false; T.advected = false;
    }
  });
  return 0;
}
// ---- advection-only driver (gmcore_adv_driver.F90:84-88) -------------------------------------------------
// gmc_adv_prescribe = set_uv of src/tests/adv/*_test_mod.F90: state(itime)%m = 1, u_f / v_f given, mfx_lon = u_f,
// mfy_lat = v_f (no vertical motion).  gmc_adv_step = adv_run(itime) for the tracer batch.
__global__ void k_fill(double* p, size_t n, double v) {
  for (size_t x = blockIdx.x * (size_t)blockDim.x + threadIdx.x; x < n; x += (size_t)gridDim.x * blockDim.x) p[x] = v;
}
int gmc_adv_prescribe(gmc_handle* h, int itime, const double* u_f, const double* v_f) {
  if (!h->cfg.baroclinic || h->batches.empty() || h->cfg.nranks > 1) {
    snprintf(h->err, sizeof(h->err), "the advection-only driver needs the hydrostatic core with a tracer batch on one band"); return 1; }
  if (itime < 1 || itime > 3) { snprintf(h->err, sizeof(h->err), "bad time index %d", itime); return 1; }
  const Geo& g = h->g; DState& S = h->st[itime];
  const size_t n3 = (size_t)g.plane * g.nlev, n3h = (size_t)g.plane * (g.nlev + 1);
  if (xfer(h, S.u_f, const_cast<double*>(u_f), 0, 0) || xfer(h, S.v_f, const_cast<double*>(v_f), 1, 0)) return 1;
  LAUNCH(h, k_fill, 1184, 256, 0, S.m, n3, 1.0);
  LAUNCH(h, k_fill, 1184, 256, 0, S.m_lev, n3h, 1.0);
  CUDA_CHECK(cudaMemsetAsync(S.we_lev, 0, n3h * sizeof(double), h->stream));
  CUDA_CHECK(cudaMemcpyAsync(S.mfx_lon, S.u_f, n3 * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
  CUDA_CHECK(cudaMemcpyAsync(S.mfy_lat, S.v_f, n3 * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
  return 0;
}
// 3-D advection cases (tests/adv/dcmip12_test_mod.F90:75-161): the prescribed vertical mass flux we_lev on the half
// levels (interior-plus-halo Fortran storage), after gmc_adv_prescribe has set m = m_lev = 1
int gmc_adv_prescribe_we(gmc_handle* h, int itime, const double* we_lev) {
  if (!h->cfg.baroclinic || h->batches.empty() || h->cfg.nranks > 1) {
    snprintf(h->err, sizeof(h->err), "the advection-only driver needs the hydrostatic core with a tracer batch on one band"); return 1; }
  if (itime < 1 || itime > 3) { snprintf(h->err, sizeof(h->err), "bad time index %d", itime); return 1; }
  return xfer(h, h->st[itime].we_lev, const_cast<double*>(we_lev), 2, 0);
}
int gmc_adv_step(gmc_handle* h, int itime, int time_step) {
  if (check_ready(h)) return 1;
  if (!h->cfg.baroclinic || h->batches.empty() || h->cfg.nranks > 1) {
    snprintf(h->err, sizeof(h->err), "the advection-only driver needs the hydrostatic core with a tracer batch on one band"); return 1; }
  alert_advance(h, time_step - 1);
  hydro_adv_run(h, itime, time_step);
  hydro_adv_finish(h, itime);
  CUDA_CHECK(cudaGetLastError());
  return hydro_check_flags(h);
}
// ---- pressure-level diagnostics (utils/diag_state_mod.F90:241-258 -> dynamics/interp_mod.F90:727-897) -------
// One variable on one pressure surface, as the reference's interp_{cell,lon_edge,lat_edge,lev_edge}_to_pressure_level:
// search the column from the surface up for the layer that brackets po, interpolate linearly in p (or log p).
// u / v are first averaged to the cell centre (:795-796, :843-844).  Columns that do not reach po keep `fill`.
struct PlevArgs { const double *x, *phs_f; double* y; double po, p0x, fill, rg; int var, logp; VC vc; };
__global__ void __launch_bounds__(128) k_plev(Geo g, DMesh M, PlevArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = g.jb + blockIdx.y;
  if (i >= g.nlon) return;
  const double phs = __ldg(a.phs_f + gidx(g, i, j, 0));
  auto pfull = [&](int k1) { return 0.5 * (ph_lev_at(M, a.vc, k1, phs) + ph_lev_at(M, a.vc, k1 + 1, phs)); };
  const bool half = a.var == 4;             // z lives on the half levels (gz_lev), bracketed by ph_lev
  double y = a.fill;
  for (int k = half ? g.nlev + 1 : g.nlev; k >= 2; --k) {
    const double pm = half ? ph_lev_at(M, a.vc, k - 1, phs) : pfull(k - 1), pk = half ? ph_lev_at(M, a.vc, k, phs) : pfull(k);
    if (pm <= a.po && a.po <= pk) {
      const double dp1 = a.logp ? a.p0x - log(pm) : a.p0x - pm, dp2 = a.logp ? log(pk) - a.p0x : pk - a.p0x;
      double x1, x2;
      if (a.var == 0) {          // u_lon -> cell centre
        x1 = 0.5 * (ldW(a.x, g, i - 1, j, k - 2) + ldW(a.x, g, i, j, k - 2)); x2 = 0.5 * (ldW(a.x, g, i - 1, j, k - 1) + ldW(a.x, g, i, j, k - 1));
      } else if (a.var == 1) {   // v_lat -> cell centre (pole rows read the mirrored half row, fill_halo semantics)
        x1 = 0.5 * (ldH(a.x, g, i, j - 1, k - 2) + ldH(a.x, g, i, j, k - 2)); x2 = 0.5 * (ldH(a.x, g, i, j - 1, k - 1) + ldH(a.x, g, i, j, k - 1));
      } else { x1 = __ldg(a.x + gidx(g, i, j, k - 2)); x2 = __ldg(a.x + gidx(g, i, j, k - 1)); if (half) { x1 = x1 * a.rg; x2 = x2 * a.rg; } }
      y = (dp2 * x1 + dp1 * x2) / (dp1 + dp2);
      break;
    }
  }
  a.y[gidx(g, i, j, 0)] = y;
}
int gmc_interp_pressure_level(gmc_handle* h, int itime, const char* var, double po, int logp, double fill, double* y2d) {
  if (!h->cfg.baroclinic) { snprintf(h->err, sizeof(h->err), "pressure-level diagnostics need the hydrostatic core"); return 1; }
  if (itime < 1 || itime > 3) { snprintf(h->err, sizeof(h->err), "bad time index %d", itime); return 1; }
  DState& S = h->st[itime];
  const std::string v(var);
  PlevArgs a{nullptr, S.phs_f, h->dphs, po, logp ? std::log(po) : po, fill, 1.0 / h->cfg.g, 0, logp, make_vc(h)};
  if (v == "u") { a.x = S.u; a.var = 0; } else if (v == "v") { a.x = S.v; a.var = 1; } else if (v == "pt") { a.x = S.pt; a.var = 2; }
  else if (v == "t") { a.x = S.t; a.var = 3; } else if (v == "z") { a.x = S.gz_lev; a.var = 4; }
  else { snprintf(h->err, sizeof(h->err), "unknown pressure-level variable %s (u, v, pt, t, z)", var); return 1; }
  LAUNCH(h, k_plev, gridg(h->g, 1), 128, 0, h->g, h->M, a);
  return xfer(h, h->dphs, y2d, 3, 1);
}
int gmc_set_tracers(gmc_handle* h, const double* q) { return gmc_set_batch_tracers(h, 0, q); }
int gmc_get_tracers(gmc_handle* h, double* q) { return gmc_get_batch_tracers(h, 0, q); }
int gmc_add_tracer_batch(gmc_handle* h, int ntracers, double dt_adv, int* batch) {
  if (!h->cfg.baroclinic) { snprintf(h->err, sizeof(h->err), "tracer batches need the hydrostatic core"); return 1; }
  return hydro_add_batch(h, ntracers, dt_adv, batch);
}
int gmc_set_batch_tracers(gmc_handle* h, int batch, const double* q) {
  if (!h->cfg.baroclinic) { snprintf(h->err, sizeof(h->err), "no tracer batch configured"); return 1; }
  return hydro_set_tracers(h, batch, q);
}
int gmc_get_batch_tracers(gmc_handle* h, int batch, double* q) {
  if (!h->cfg.baroclinic) { snprintf(h->err, sizeof(h->err), "no tracer batch configured"); return 1; }
  return hydro_get_tracers(h, batch, q);
}

}  // extern "C"

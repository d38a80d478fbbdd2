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
  LAUNCH(h, k_diag_final, 1, 1024, 0, h->diag_partial, nrows * g.nlev, dev_out);
  return 0;
}
// device sums (tm, te_ke, te_ie, te_pe, tav, tpe, tpt) -> ABI order
static void diag_to_abi(const double* s, double* out) {
  out[0] = s[0]; out[1] = s[1] + s[2] + s[3]; out[2] = s[4]; out[3] = s[5];
  out[4] = s[1]; out[5] = s[2]; out[6] = s[3]; out[7] = s[6];
}

static int time_integrate_impl(gmc_handle* h, int old_, int nw, double dt, int time_step) {
  time_integrator(h, old_, nw, dt);
  if (h->cfg.baroclinic) hydro_post_step(h, nw, dt, time_step);
  return 0;
}

extern "C" {

int gmc_prepare_run(gmc_handle* h, int old_) {
  if (check_ready(h)) return 1;
  if (old_ < 1 || old_ > 3) { snprintf(h->err, sizeof(h->err), "bad time index %d", old_); return 1; }
  const Geo& g = h->g; DState& S = h->st[old_];
  run_filter(h, S.u, 0, g.nlev, GMC_BIG_FILTER, S.u_f, -1, nullptr);
  run_filter(h, S.v, 1, g.nlev, GMC_BIG_FILTER, S.v_f, -1, nullptr);
  if (h->cfg.baroclinic) {
    hydro_prepare_run(h, old_);
  } else {
    // gz_f = big(gz); calc_m: m = (gz - gzs) / g
    run_filter(h, S.gz, 0, g.nlev, GMC_BIG_FILTER, S.gz_f, -1, nullptr, POST_SWM_M, nullptr, S.m);
    run_forward_prepare(h, S);
  }
  CUDA_CHECK(cudaGetLastError());
  return 0;
}

int gmc_time_integrate(gmc_handle* h, int old_, int nw, double dt, int time_step) {
  if (check_ready(h)) return 1;
  time_integrate_impl(h, old_, nw, dt, time_step);
  CUDA_CHECK(cudaGetLastError());
  return 0;
}

int gmc_diagnose(gmc_handle* h, int itime, double out[8]) {
  if (check_ready(h)) return 1;
  launch_diag(h, itime, h->diag_out);
  double s[8];
  CUDA_CHECK(cudaMemcpyAsync(s, h->diag_out, 7 * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  CUDA_CHECK(cudaStreamSynchronize(h->stream));
  diag_to_abi(s, out);
  return 0;
}

int gmc_run_steps(gmc_handle* h, int* old_idx, int* new_idx, int* time_step, int nsteps, double* diag) {
  if (check_ready(h)) return 1;
  double* dbuf = nullptr;
  if (diag) {
    cudaError_t e = cudaMalloc(&dbuf, (size_t)nsteps * 8 * sizeof(double));
    if (e != cudaSuccess) { snprintf(h->err, sizeof(h->err), "cudaMalloc diag: %s", cudaGetErrorString(e)); return 1; }
  }
  for (int s = 0; s < nsteps; ++s) {
    time_integrate_impl(h, *old_idx, *new_idx, h->cfg.dt_dyn, *time_step);
    // time_advance (time_mod.F90:128-159)
    if (h->alert_ring) { h->alert_last = (*time_step) * h->cfg.dt_dyn; h->alert_ring = false; }
    int t = *old_idx; *old_idx = *new_idx; *new_idx = t;
    *time_step += 1;
    if (diag) launch_diag(h, *old_idx, dbuf + (size_t)s * 8);
  }
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess && diag) {
    std::vector<double> tmp((size_t)nsteps * 8);
    e = cudaMemcpyAsync(tmp.data(), dbuf, tmp.size() * sizeof(double), cudaMemcpyDeviceToHost, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    for (int s = 0; s < nsteps; ++s) diag_to_abi(&tmp[(size_t)s * 8], diag + (size_t)s * 8);
  }
  if (dbuf) cudaFree(dbuf);
  if (e != cudaSuccess) { snprintf(h->err, sizeof(h->err), "run_steps: %s", cudaGetErrorString(e)); return 1; }
  return hydro_check_flags(h);
}

int gmc_set_tracers(gmc_handle* h, const double* q) { return hydro_set_tracers(h, q); }
int gmc_get_tracers(gmc_handle* h, double* q) { return hydro_get_tracers(h, q); }

}  // extern "C"

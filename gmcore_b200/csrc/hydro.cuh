// Host-side sequencing of the hydrostatic core: allocation, one step(old, star, new, dt)
// (time_schemes_mod.F90:102-242 with gmcore_mod.F90:467-571), the post-step operators
// of time_integrate (gmcore_mod.F90:577-591) and the tracer batch (adv_mod.F90:152-455).
#pragma once
#include <cmath>
#include "kernels_hydro.cuh"

static inline VC make_vc(const gmc_handle* h) {
  const gmc_config& c = h->cfg;
  return VC{c.ptop, c.hybrid_p0, c.p0, c.Rd_o_cpd, c.Rd};
}

static int hydro_alloc(gmc_handle* h) {
  const Geo& g = h->g;
  const size_t n3 = (size_t)g.plane * g.nlev, n3h = (size_t)g.plane * (g.nlev + 1), n2 = (size_t)g.plane;
  int rc = 0;
  auto A = [&](double** p, size_t n) { if (!rc) rc = dev_alloc(h, p, n); };
  A(&h->dphs, n2);
  A(&h->cflx, n3); A(&h->cfly, n3); A(&h->cflz, n3h); A(&h->divx, n3); A(&h->divy, n3);
  A(&h->qx, n3); A(&h->qy, n3); A(&h->ptf_lon, n3); A(&h->ptf_lat, n3); A(&h->qmfx, n3); A(&h->qmfy, n3);
  A(&h->dpt, n3);
  if (h->cfg.ntracers > 0) {
    A(&h->q[0], n3 * h->cfg.ntracers); A(&h->q[1], n3 * h->cfg.ntracers);
    A(&h->old_m, n3); A(&h->bu, n3); A(&h->bv, n3); A(&h->bmfx, n3); A(&h->bmfy, n3);
    A(&h->bwe, n3h); A(&h->bm_lev, n3h);
    A(&h->bcflx, n3); A(&h->bcfly, n3); A(&h->bcflz, n3h); A(&h->bdivx, n3); A(&h->bdivy, n3);
  }
  if (!rc) { int* f = nullptr; rc = dev_alloc(h, &f, 4); h->dflags = f; }
  return rc;
}

// damp_init pieces that are pure functions of the mesh: div_damp coefficients
// (div_damp_mod.F90:27-87) and the lon_damp latitude profile (lon_damp_mod.F90:43-68).
static int hydro_set_mesh(gmc_handle* h, const gmc_mesh_vectors* mv) {
  const gmc_config& c = h->cfg;
  const int nlat = c.nlat, nhalf = c.nlat - 1, nlev = c.nlev, hy = c.hy;
  auto F = [&](const double* v, int j) { return v[j - (1 - hy)]; };
  std::vector<double> cl((size_t)nlat * nlev, 0.0), ca((size_t)nhalf * nlev, 0.0);
  for (int k = 1; k <= nlev; ++k) {
    double fac = 1.0;
    if (c.div_damp_order == 2) {
      const double hyam = 0.5 * (mv->hyai[k - 1] + mv->hyai[k]), hybm = 0.5 * (mv->hybi[k - 1] + mv->hybi[k]);
      const double ph = hyam * (c.hybrid_p0 - c.ptop) + hybm * (c.p0 - c.ptop) + c.ptop;
      fac = std::fmax(1.0, 8 * (1 + std::tanh(std::log(c.ptop / ph))));
    }
    for (int j = 2; j <= nlat - 1; ++j) {
      const double le = F(mv->le_lon, j), de = F(mv->de_lon, j);
      cl[(size_t)(k - 1) * nlat + (j - 1)] = c.div_damp_order == 2 ? c.div_damp_coef2 * fac * le * de : c.div_damp_coef4 * (le * le) * (de * de);
    }
    for (int j = 1; j <= nhalf; ++j) {
      const double le = F(mv->le_lat, j), de = F(mv->de_lat, j);
      ca[(size_t)(k - 1) * nhalf + (j - 1)] = c.div_damp_order == 2 ? c.div_damp_coef2 * fac * le * de : c.div_damp_coef4 * (le * le) * (de * de);
    }
  }
  const double* p = nullptr;
  int rc = dev_upload(h, &p, cl.data(), cl.size()); h->c_lon = const_cast<double*>(p);
  if (!rc) { rc = dev_upload(h, &p, ca.data(), ca.size()); h->c_lat = const_cast<double*>(p); }
  // lon_damp weights; full_lat_deg(j) = full_lat(j) * deg (mesh_mod.F90:212)
  const double pi = std::atan(1.0) * 4.0, deg = 180.0 / pi;
  std::vector<double> w(nlat + 1, 0.0);
  const double lat0 = -(F(mv->full_lat, 2) * deg), l0 = 80;
  for (int j = 2; j <= nlat - 1; ++j) {
    const double lat = std::fabs(F(mv->full_lat, j) * deg);
    const double x = std::exp((lat0 - lat) * (lat0 - lat) * std::log(1.0e-3) / ((lat0 - l0) * (lat0 - l0)));
    w[j] = x < 1.0e-6 ? 0.0 : x;
  }
  if (!rc) { rc = dev_upload(h, &p, w.data(), w.size()); h->lon_damp_wgt = const_cast<double*>(p); }
  return rc;
}

static bool hydro_work_field(gmc_handle* h, const std::string& n, double** p, int* kind) {
  struct E { const char* n; double* p; int kind; };
  const E tab[] = {{"cflx", h->cflx, 0}, {"cfly", h->cfly, 1}, {"cflz", h->cflz, 2}, {"divx", h->divx, 0}, {"divy", h->divy, 0},
                   {"qx", h->qx, 0}, {"qy", h->qy, 0}, {"ptf_lon", h->ptf_lon, 0}, {"ptf_lat", h->ptf_lat, 1},
                   {"dpt", h->dpt, 0}, {"dphs", h->dphs, 3}};
  for (const E& e : tab) if (n == e.n && e.p) { *p = e.p; *kind = e.kind; return true; }
  return false;
}

__global__ void k_axpy2d(Geo g, const double* x, const double* d, double* y, double dt) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = g.jb + blockIdx.y;
  if (i >= g.nlon) return;
  const long long c = gidx(g, i, j, 0);
  y[c] = x[c] + dt * d[c];
}
__global__ void k_copy3(Geo g, const double* x, double* y) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = g.jb + blockIdx.y, k = blockIdx.z;
  if (i >= g.nlon) return;
  const long long c = gidx(g, i, j, k);
  y[c] = x[c];
}

static inline dim3 grid2(const gmc_handle* h, int bx = 128) { return dim3((h->g.nlon + bx - 1) / bx, h->g.je - h->g.jb + 1, 1); }

// calc_ph / calc_m / calc_t / calc_gz_lev of one state from its phs_f and pt
// (operators_prepare backward_pass, operators_mod.F90:123-130)
static void run_ph(gmc_handle* h, DState& S) {
  const Geo& g = h->g;
  PhArgs p{S.phs_f, S.ph_exn_lev, S.m, S.m_lev, make_vc(h), h->dflags};
  LAUNCH(h, k_ph, grid3(h, g.nlev + 1), 128, 0, g, h->M, p);
}
static void run_gz(gmc_handle* h, DState& S) {
  const Geo& g = h->g;
  GzArgs a{S.pt, S.phs_f, h->gzs, S.t, S.gz_lev, make_vc(h)};
  LAUNCH(h, k_gz_col, grid2(h, 64), 64, (size_t)g.nlev * 64 * sizeof(double), g, h->M, a);
}

// FFSL horizontal flux of field q with the winds/CFL of one batch (ffsl_mod.F90:318-413)
struct BatchRef { const double *u, *v, *mfx, *mfy, *cflx, *cfly, *divx, *divy; };
static void run_ffsl_h(gmc_handle* h, const BatchRef& B, const double* q, double* flx, double* fly, double dt) {
  const Geo& g = h->g;
  HflxArgs in{B.u, B.v, q, q, B.cflx, B.cfly, h->qmfx, h->qmfy};
  LAUNCH(h, k_hflx, grid3(h, g.nlev), 128, 0, g, in);
  PoleSumArgs ps{h->qmfy, h->pole_work, nullptr, 0};
  LAUNCH(h, k_pole_sum, dim3(g.nlev, 2), 256, 0, g, h->M, ps);
  FfslQArgs qa{q, h->qmfx, h->qmfy, B.divx, B.divy, h->pole_work, h->qx, h->qy, dt};
  LAUNCH(h, k_ffsl_q, grid3(h, g.nlev), 128, 0, g, h->M, qa);
  HflxArgs out{B.mfx, B.mfy, h->qy, h->qx, B.cflx, B.cfly, flx, fly};
  LAUNCH(h, k_hflx, grid3(h, g.nlev), 128, 0, g, out);
}

static void run_cfl(gmc_handle* h, const double* u, const double* v, double* cflx, double* cfly, double* divx, double* divy, double dt) {
  const Geo& g = h->g;
  CflArgs c{u, v, cflx, cfly, divx, divy, dt, h->cfg.radius, h->dflags};
  LAUNCH(h, k_cfl, grid3(h, g.nlev), 128, 0, g, h->M, c);
  PoleSumArgs ps{v, nullptr, divy, 1};
  LAUNCH(h, k_pole_sum, dim3(g.nlev, 2), 256, 0, g, h->M, ps);
}

static void hydro_accum_wind(gmc_handle* h, int itime);
extern "C" { static int xfer(gmc_handle* h, double* dev, double* host, int kind, int dir); }

// One hydrostatic step(old, star, new, dt).
static void hydro_step(gmc_handle* h, DState& old_, DState& star, DState& nw, double dt) {
  const Geo& g = h->g;
  // ---- forward pass on star -------------------------------------------------------
  run_forward_prepare(h, star);
  run_cfl(h, star.u_f, star.v_f, h->cflx, h->cfly, h->divx, h->divy, dt);
  WeArgs w{star.mfx_lon, star.mfy_lat, h->pole_dmf, star.m_lev, star.we_lev, h->dphs, h->cflz, dt, h->dflags};
  LAUNCH(h, k_we_col, grid2(h), 128, 0, g, h->M, w);
  TendArgs t{};
  t.mfx_lon = star.mfx_lon; t.mfy_lat = star.mfy_lat; t.pv_lon = star.pv_lon; t.pv_lat = star.pv_lat; t.ke = star.ke;
  t.we_lev = star.we_lev; t.u_f = star.u_f; t.v_f = star.v_f; t.m = star.m; t.du = h->du; t.dv = h->dv;
  t.coriolis_scheme = h->cfg.coriolis_scheme; t.swm = 0;
  LAUNCH(h, k_tend, grid3(h, g.nlev), 128, 0, g, h->M, t);
  BatchRef B{star.u_f, star.v_f, star.mfx_lon, star.mfy_lat, h->cflx, h->cfly, h->divx, h->divy};
  run_ffsl_h(h, B, star.pt, h->ptf_lon, h->ptf_lat, dt);
  PoleSumArgs ps{h->ptf_lat, h->pole_work + 2 * g.nlev, nullptr, 1};
  LAUNCH(h, k_pole_sum, dim3(g.nlev, 2), 256, 0, g, h->M, ps);
  PtColArgs pc{star.pt, h->ptf_lon, h->ptf_lat, h->pole_work + 2 * g.nlev, star.we_lev, h->cflz, old_.pt, old_.m, nw.pt, h->dpt, dt};
  LAUNCH(h, k_pt_col, grid2(h), 128, 0, g, h->M, pc);
  // ---- update_state(tend1) ----------------------------------------------------------
  LAUNCH(h, k_axpy2d, grid2(h), 128, 0, g, old_.phs, h->dphs, nw.phs, dt);
  run_filter(h, nw.phs, 0, 1, GMC_BIG_FILTER, nw.phs_f, -1, nullptr);
  run_ph(h, nw);
  run_filter(h, nw.pt, 0, g.nlev, -1, nullptr, GMC_SMALL_FILTER1, nw.pt, POST_DIV_M, nw.m, nullptr);
  // ---- backward pass + update_state(tend2) --------------------------------------------
  run_gz(h, nw);
  UvArgs u{};
  u.ph_exn_lev = nw.ph_exn_lev; u.gz_lev = nw.gz_lev; u.du_fwd = h->du; u.dv_fwd = h->dv;
  u.u_old = old_.u; u.v_old = old_.v; u.u_new = nw.u; u.v_new = nw.v; u.dt = dt; u.swm = 0; u.blend = 0;
  LAUNCH(h, k_uv_update, grid3(h, g.nlev), 128, 0, g, h->M, u);
  run_filter(h, nw.u, 0, g.nlev, -1, nullptr, GMC_SMALL_FILTER2, nw.u);
  run_filter(h, nw.v, 1, g.nlev, -1, nullptr, GMC_SMALL_FILTER2, nw.v);
  LAUNCH(h, k_v_pole_blend, dim3((g.nlon + 127) / 128, g.nlev, 2), 128, 0, g, nw.v);
  run_filter(h, nw.u, 0, g.nlev, GMC_BIG_FILTER, nw.u_f, -1, nullptr);
  run_filter(h, nw.v, 1, g.nlev, GMC_BIG_FILTER, nw.v_f, -1, nullptr);
}

// gmcore_run pre-loop, hydrostatic part (gmcore_mod.F90:104-136) + operators_prepare_1
static void hydro_prepare_run(gmc_handle* h, int old_) {
  DState& S = h->st[old_];
  run_filter(h, S.phs, 0, 1, GMC_BIG_FILTER, S.phs_f, -1, nullptr);
  hydro_accum_wind(h, old_);
  run_ph(h, S);
  run_gz(h, S);
  run_forward_prepare(h, S);
  run_cfl(h, S.u_f, S.v_f, h->cflx, h->cfly, h->divx, h->divy, h->cfg.dt_dyn);
}

// adv_accum_wind for the single tracer batch (adv_mod.F90:420-455, adv_batch_mod.F90:
// 239-252,379-403,512-525): copy / add / average the winds of state(itime), and
// recompute the batch CFL numbers when the sample counter wraps (T8).
static void accum_pair(gmc_handle* h, double* d0, const double* s0, int n0, double* d1, const double* s1, int n1, int mode) {
  const Geo& g = h->g;
  AccumArgs a; a.dst[0] = d0; a.src[0] = s0; a.nlev_[0] = n0; a.dst[1] = d1; a.src[1] = s1; a.nlev_[1] = n1;
  a.mode = mode; a.nstep = h->nstep;
  LAUNCH(h, k_accum, grid3(h, n0 > n1 ? n0 : n1), 128, 0, g, a);
}
static int accum_mode(int& step, int nstep) {
  // returns the mode for this call and advances the counter (non-dynamic batch)
  if (step == 0) { step = 1; return 0; }
  if (step == nstep) { step = 0; return 2; }
  step += 1; return 1;
}
static void hydro_accum_wind(gmc_handle* h, int itime) {
  if (h->cfg.ntracers <= 0) return;
  const Geo& g = h->g; DState& S = h->st[itime];
  int mode = accum_mode(h->uv_step, h->nstep);
  accum_pair(h, h->bu, S.u_f, g.nlev, h->bv, S.v_f, g.nlev, mode);
  if (h->uv_step == 0) run_cfl(h, h->bu, h->bv, h->bcflx, h->bcfly, h->bdivx, h->bdivy, h->cfg.dt_adv);
  mode = accum_mode(h->mf_step, h->nstep);
  accum_pair(h, h->bmfx, S.mfx_lon, g.nlev, h->bmfy, S.mfy_lat, g.nlev, mode);
  mode = accum_mode(h->we_step, h->nstep);
  accum_pair(h, h->bwe, S.we_lev, g.nlev + 1, h->bm_lev, S.m_lev, g.nlev + 1, mode);
  if (h->we_step == 0) {
    CflzArgs c{h->bwe, h->bm_lev, h->bcflz, h->cfg.dt_adv, h->dflags};
    LAUNCH(h, k_cflz, grid2(h), 128, 0, g, h->M, c);
  }
}

// adv_run (adv_mod.F90:152-312)
static void hydro_adv_run(gmc_handle* h, int itime, int time_step) {
  hydro_accum_wind(h, itime);
  if (h->cfg.ntracers <= 0) return;
  const Geo& g = h->g; DState& S = h->st[itime];
  const size_t n3 = (size_t)g.plane * g.nlev;
  const double dt_adv = h->cfg.dt_adv;
  bool alerted = false;
  if (h->alert_last + h->alert_period <= time_step * h->cfg.dt_dyn) { h->alert_ring = true; alerted = true; }
  if (alerted && time_step > 0) {
    BatchRef B{h->bu, h->bv, h->bmfx, h->bmfy, h->bcflx, h->bcfly, h->bdivx, h->bdivy};
    for (int l = 0; l < h->cfg.ntracers; ++l) {
      const double* qo = h->q[h->q_old] + l * n3; double* qn = h->q[h->q_new] + l * n3;
      run_ffsl_h(h, B, qo, h->ptf_lon, h->ptf_lat, dt_adv);   // ptf_* double as qmf_lon / qmf_lat
      PoleSumArgs ps{h->ptf_lat, h->pole_work + 2 * g.nlev, nullptr, 0};
      LAUNCH(h, k_pole_sum, dim3(g.nlev, 2), 256, 0, g, h->M, ps);
      TrHArgs th{qo, h->old_m, S.m, h->ptf_lon, h->ptf_lat, h->pole_work + 2 * g.nlev, qn, dt_adv};
      LAUNCH(h, k_tracer_h, grid3(h, g.nlev), 128, 0, g, h->M, th);
      NegXArgs nx{qn, S.m, h->dflags};
      LAUNCH(h, k_negfill_x, dim3(g.je - g.jb + 1, g.nlev), 128, 0, g, nx);
      TrVArgs tv{qn, S.m, h->bwe, h->bcflz, dt_adv, h->dflags};
      LAUNCH(h, k_tracer_v, grid2(h, 64), 64, (size_t)2 * g.nlev * 64 * sizeof(double), g, tv);
      LonDampArgs ld{qn, h->lon_damp_wgt, 3};
      LAUNCH(h, k_lon_damp, dim3(g.je - g.jb + 1, g.nlev), 256, (size_t)2 * g.nlon * sizeof(double), g, ld);
    }
    int t = h->q_old; h->q_old = h->q_new; h->q_new = t;
  }
  LAUNCH(h, k_copy3, grid3(h, g.nlev), 128, 0, g, S.m, h->old_m);   // copy_old_m
}

// time_integrate after the time integrator (gmcore_mod.F90:586-590)
static void hydro_post_step(gmc_handle* h, int nw, double dt, int time_step) {
  const Geo& g = h->g; DState& S = h->st[nw];
  if (h->cfg.test_forcing == GMC_FORCING_HELD_SUAREZ) {
    HsArgs a{S.u, S.v, S.pt, S.t, S.phs_f, dt, make_vc(h)};
    LAUNCH(h, k_held_suarez, grid3(h, g.nlev), 128, 0, g, h->M, a);
  }
  if (h->cfg.use_div_damp) {
    DivArgs d{S.u_f, S.v_f, S.div};
    LAUNCH(h, k_div, grid3(h, g.nlev), 128, 0, g, h->M, d);
    PoleSumArgs ps{S.v_f, nullptr, S.div, 1};
    LAUNCH(h, k_pole_sum, dim3(g.nlev, 2), 256, 0, g, h->M, ps);
    DivDampArgs dd{S.div, h->c_lon, h->c_lat, S.u, S.v};
    LAUNCH(h, k_div_damp, grid3(h, g.nlev), 128, 0, g, h->M, dd);
  }
  hydro_adv_run(h, nw, time_step);
}

// Maps the reference's log_error guards onto the ABI error text.
static int hydro_check_flags(gmc_handle* h) {
  if (!h->dflags) return 0;
  int f = 0;
  if (cudaMemcpyAsync(&f, h->dflags, sizeof(int), cudaMemcpyDeviceToHost, h->stream) != cudaSuccess) return 0;
  cudaStreamSynchronize(h->stream);
  if (!f) return 0;
  snprintf(h->err, sizeof(h->err), "%s%s%s%s%s",
           f & 1 ? "cflx exceeds mesh%%lon_halo_width; " : "", f & 2 ? "cfly exceeds 1; " : "", f & 4 ? "cflz exceeds range; " : "",
           f & 8 ? "The pressure levels are not monotonic!; " : "", f & 16 ? "Negative tracer!; " : "");
  return 1;
}

static int hydro_set_tracers(gmc_handle* h, const double* q) {
  if (h->cfg.ntracers <= 0) { snprintf(h->err, sizeof(h->err), "no tracer batch configured"); return 1; }
  const size_t n3 = (size_t)h->g.plane * h->g.nlev;
  const size_t host_n = (size_t)(h->cfg.nlon + 2 * h->cfg.hx) * ((h->g.je - h->g.jb + 1) + 2 * h->cfg.hy) * (h->cfg.nlev + 4);
  for (int l = 0; l < h->cfg.ntracers; ++l)
    if (xfer(h, h->q[h->q_old] + l * n3, const_cast<double*>(q) + l * host_n, 0, 0)) return 1;
  return 0;
}
static int hydro_get_tracers(gmc_handle* h, double* q) {
  if (h->cfg.ntracers <= 0) { snprintf(h->err, sizeof(h->err), "no tracer batch configured"); return 1; }
  const size_t n3 = (size_t)h->g.plane * h->g.nlev;
  const size_t host_n = (size_t)(h->cfg.nlon + 2 * h->cfg.hx) * ((h->g.je - h->g.jb + 1) + 2 * h->cfg.hy) * (h->cfg.nlev + 4);
  for (int l = 0; l < h->cfg.ntracers; ++l)
    if (xfer(h, h->q[h->q_old] + l * n3, q + l * host_n, 0, 1)) return 1;
  return 0;
}

// Host-side sequencing of the hydrostatic core: allocation, one step(old, star, new, dt)
// (time_schemes_mod.F90:102-242 with gmcore_mod.F90:467-571), the post-step operators
// of time_integrate (gmcore_mod.F90:577-591) and the tracer batch (adv_mod.F90:152-455).
#pragma once
#include <cmath>
#include "kernels_hydro.cuh"
#include "kernels_ffsl.cuh"

static inline VC make_vc(const gmc_handle* h) {
  const gmc_config& c = h->cfg;
  return VC{c.ptop, c.hybrid_p0, c.p0, c.Rd_o_cpd, c.Rd};
}

// adv_add_tracer + adv_allocate_tracers_1 (adv_mod.F90:314-418) for one more 'cell' batch of ntr tracers advected
// every dt seconds; adv_batch_type%init (adv_batch_mod.F90:73-149): nstep = dt / dt_dyn.
static int halo_ensure(gmc_handle* h, int ktot);
static int hydro_add_batch(gmc_handle* h, int ntr, double dt, int* id) {
  const Geo& g = h->g;
  const size_t n3 = (size_t)g.plane * g.nlev, n3h = (size_t)g.plane * (g.nlev + 1);
  if (ntr < 1) { snprintf(h->err, sizeof(h->err), "a tracer batch needs at least one tracer"); return 1; }
  if (h->cfg.nranks > 1 && ntr + 1 > GMC_HALO_MAX) { snprintf(h->err, sizeof(h->err), "at most %d tracers per batch in band mode", GMC_HALO_MAX - 1); return 1; }
  if (!(dt >= h->cfg.dt_dyn)) { snprintf(h->err, sizeof(h->err), "batch dt %g is shorter than dt_dyn %g", dt, h->cfg.dt_dyn); return 1; }
  TrBatch B; B.ntr = ntr; B.dt = dt; B.nstep = (int)(dt / h->cfg.dt_dyn);
  B.alert_period = dt; B.alert_last = -dt;
  int rc = 0;
  auto A = [&](double** p, size_t n) { if (!rc) rc = dev_alloc(h, p, n); };
  A(&B.q[0], n3 * ntr); A(&B.q[1], n3 * ntr);
  if (!h->old_m) A(&h->old_m, n3);
  B.wind_of = (int)h->batches.size();
  for (size_t m = 0; m < h->batches.size(); ++m) if (h->batches[m].wind_of == (int)m && h->batches[m].dt == dt) { B.wind_of = (int)m; break; }
  if (B.wind_of == (int)h->batches.size()) {
    A(&B.bu, n3); A(&B.bv, n3); A(&B.bmfx, n3); A(&B.bmfy, n3); A(&B.bwe, n3h); A(&B.bm_lev, n3h);
    A(&B.bcflx, n3); A(&B.bcfly, n3); A(&B.bcflz, n3h); A(&B.bdivx, n3); A(&B.bdivy, n3);
  }
  if (!rc) rc = halo_ensure(h, g.nlev * (ntr + 1));
  if (rc) return 1;
  if (id) *id = (int)h->batches.size();
  h->batches.push_back(B);
  return 0;
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
  if (!rc) { int* f = nullptr; rc = dev_alloc(h, &f, 4); h->dflags = f; }
  if (!rc && h->cfg.ntracers > 0) rc = hydro_add_batch(h, h->cfg.ntracers, h->cfg.dt_adv, nullptr);
  return rc;
}

// damp_init pieces that are pure functions of the mesh: div_damp coefficients
// (div_damp_mod.F90:27-87) and the lon_damp latitude profile (lon_damp_mod.F90:43-68).
// div_damp_init (div_damp_mod.F90:27-87), both cores: the level profile applies to the baroclinic core only (:46-51)
static int damp_set_mesh(gmc_handle* h, const gmc_mesh_vectors* mv) {
  const gmc_config& c = h->cfg;
  if (!c.use_div_damp) return 0;
  const int nlat = c.nlat, nhalf = c.nlat - 1, nlev = c.nlev, hy = c.hy;
  auto F = [&](const double* v, int j) { return v[j - (1 - hy)]; };
  std::vector<double> cl((size_t)nlat * nlev, 0.0), ca((size_t)nhalf * nlev, 0.0);
  for (int k = 1; k <= nlev; ++k) {
    double fac = 1.0;
    if (c.div_damp_order == 2 && c.baroclinic) {
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
  if (!rc && c.div_damp_order == 4 && !h->div2) rc = dev_alloc(h, &h->div2, (size_t)h->g.plane * h->g.nlev);
  return rc;
}

static int hydro_set_mesh(gmc_handle* h, const gmc_mesh_vectors* mv) {
  const gmc_config& c = h->cfg;
  {  // column kernels stage whole columns in dynamic shared memory: opt in above 48 KB (deep level counts)
    const int big = 200 * 1024;
    if (we_col_smem(c.nlev) > (size_t)big || (size_t)c.nlev * 128 * sizeof(double) > (size_t)big) {
      snprintf(h->err, sizeof(h->err), "num_lev %d: a column does not fit the 200 KB shared-memory staging", c.nlev); return 1;
    }
    cudaFuncSetAttribute(k_we_col, cudaFuncAttributeMaxDynamicSharedMemorySize, big);
    cudaFuncSetAttribute(k_cflz, cudaFuncAttributeMaxDynamicSharedMemorySize, big);
#define SMEM_OPT_IN(kern) cudaFuncSetAttribute(kern<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, big); \
  cudaFuncSetAttribute(kern<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, big); cudaFuncSetAttribute(kern<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, big); \
  cudaFuncSetAttribute(kern<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, big); cudaFuncSetAttribute(kern<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, big); \
  cudaFuncSetAttribute(kern<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, big);
    SMEM_OPT_IN(k_pt_col) SMEM_OPT_IN(k_tracer_v)
#undef SMEM_OPT_IN
  }
  {  // pure-pressure levels (hybi == 0 from the top down): per-level pow / log table, see DMesh.lev_tab
    int nconst = 0;
    while (nconst <= c.nlev && mv->hybi[nconst] == 0.0) ++nconst;
    if (getenv("GMCORE_NO_LEVTAB")) nconst = 0;
    double* tab = nullptr;
    if (dev_alloc(h, &tab, (size_t)3 * (c.nlev + 1))) return 1;
    cudaMemset(tab, 0, (size_t)3 * (c.nlev + 1) * sizeof(double));
    h->M.lev_tab = tab; h->M.nconst = 0;
    k_lev_tab<<<(c.nlev + 128) / 128, 128, 0, h->stream>>>(h->M, make_vc(h), c.nlev, tab);
    if (cudaStreamSynchronize(h->stream) != cudaSuccess) { snprintf(h->err, sizeof(h->err), "k_lev_tab failed"); return 1; }
    h->M.nconst = nconst;
  }
  const int nlat = c.nlat, nlev = c.nlev, hy = c.hy;
  auto F = [&](const double* v, int j) { return v[j - (1 - hy)]; };
  const double* p = nullptr;
  int rc = 0;
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

// ph = (ph_lev(k) + ph_lev(k+1)) / 2 or gz = (gz_lev(k) + gz_lev(k+1)) / 2 for gmc_get_field
struct FullLevArgs { const double *phs_f, *gz_lev; double* out; int what; VC vc; };
__global__ void __launch_bounds__(128) k_full_lev(Geo g, DMesh M, FullLevArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = g.jb + blockIdx.y, k = blockIdx.z;
  if (i >= g.nlon) return;
  const int c = gidx(g, i, j, k);
  if (a.what == 0) { const double phs = __ldg(a.phs_f + gidx(g, i, j, 0)); a.out[c] = 0.5 * (ph_lev_at(M, a.vc, k + 1, phs) + ph_lev_at(M, a.vc, k + 2, phs)); }
  else a.out[c] = 0.5 * (__ldg(a.gz_lev + c) + __ldg(a.gz_lev + c + (int)g.plane));
}
__global__ void k_axpy2d(Geo g, const double* x, const double* d, double* y, double dt) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = g.jb + blockIdx.y;
  if (i >= g.nlon) return;
  const long long c = gidx(g, i, j, 0);
  y[c] = x[c] + dt * d[c];
}
__global__ void k_copy3(Geo g, const double* __restrict__ x, double* __restrict__ y) {   // one block per (row, level)
  const int j = g.jb + blockIdx.x, k = blockIdx.y;
  const long long rb = gidx(g, 0, j, k);
  for (int i0 = threadIdx.x; i0 < g.nlon; i0 += 4 * blockDim.x) {      // four loads in flight per thread
    double v[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) { const int i = i0 + r * blockDim.x; v[r] = i < g.nlon ? __ldg(x + rb + i) : 0.0; }
#pragma unroll
    for (int r = 0; r < 4; ++r) { const int i = i0 + r * blockDim.x; if (i < g.nlon) y[rb + i] = v[r]; }
  }
}

// Row ranges.  In band mode (cfg.nranks > 1) most kernels run on a few rows beyond the owned
// band (redundant compute on the halo rows) so that only ONE packed exchange per stage is needed
// (design B).  xg() widens the owned range by (es, en) rows, clipped to the sphere; on a
// single device it is the identity.
static inline Geo xg(const gmc_handle* h, int es, int en) {
  Geo g = h->g;
  if (h->cfg.nranks > 1) {
    g.jb = g.jb - es < 1 ? 1 : g.jb - es;
    g.je = g.je + en > g.nlat ? g.nlat : g.je + en;
    g.hjb = g.jb; g.hje = g.je > g.nhalf ? g.nhalf : g.je;
  }
  return g;
}
static inline dim3 gridg(const Geo& g, int nk, int bx = 128) { return dim3((g.nlon + bx - 1) / bx, g.je - g.jb + 1, nk); }

// calc_ph / calc_m / calc_t / calc_gz_lev of one state from its phs_f and pt
// (operators_prepare backward_pass, operators_mod.F90:123-130)
static void run_ph(gmc_handle* h, DState& S, const Geo& g) {
  PhArgs p{S.phs_f, S.ph_exn_lev, S.m, S.m_lev, make_vc(h), h->dflags};
  LAUNCH(h, k_ph, gridg(g, PH_COL ? 1 : g.nlev + 1), 128, 0, g, h->M, p);
}
static void run_gz(gmc_handle* h, DState& S, const Geo& g) {
  GzArgs a{S.pt, S.phs_f, h->gzs, S.t, S.gz_lev, make_vc(h)};
  LAUNCH(h, k_gz_col, gridg(g, 1), 128, 0, g, h->M, a);
}
static void run_pole_sum(gmc_handle* h, const double* f, double* out, double* row_out, int signed_north) {
  const Geo& g = h->g;
  if (!g.has_south && !g.has_north) return;
  PoleSumArgs ps{f, out, row_out, signed_north};
  LAUNCH(h, k_pole_sum, dim3(g.nlev, 2), 256, 0, g, h->M, ps);
}

// FFSL horizontal flux of field q with the winds/CFL of one batch (ffsl_mod.F90:318-413).
// Row needs (owned rows jb..je): outer y flux on half rows jb-1..je(+1) reads qx on jb-3..je+4;
// qx needs the inner x flux on its own row; qy (rows jb..je+1) needs the inner y flux on jb-1..je+1.
struct BatchRef { const double *u, *v, *mfx, *mfy, *cflx, *cfly, *divx, *divy; };
static void run_ffsl_h(gmc_handle* h, const BatchRef& B, const double* q, double* flx, double* fly, double dt) {
  const Geo gw = xg(h, 3, 4), gn = xg(h, 1, 1), g0 = h->g;
  static const bool old_path = getenv("GMCORE_OLD_HFLX") && getenv("GMCORE_OLD_HFLX")[0] == '1';   // round-1 kernels (A/B runs)
  // x and y sweeps in one marching kernel or as two kernels.  Measured at 0.25 deg L32 (tools/kprof.py): inner call
  // fused 0.741 ms, split 0.407 + 0.295 ms; outer call fused 0.466 ms, split 0.249 + 0.250 ms -> default: inner split,
  // outer fused.  GMCORE_FFSL_FUSED = 1 / 0 forces both calls fused / split (A/B runs).
  static const char* fenv = getenv("GMCORE_FFSL_FUSED");
  static const bool fused_in = fenv && fenv[0] == '1', fused_out = !(fenv && fenv[0] == '0');
  static const bool xblock = getenv("GMCORE_FFSL_XBLOCK") && getenv("GMCORE_FFSL_XBLOCK")[0] == '1';   // block-staged x sweep (A/B)
  if (!old_path) {
    // inner advective half step: the fluxes stay on chip, the kernels write qx / qy (and the pole-adjacent y fluxes)
    const Geo g01 = xg(h, 0, 1);
    FfslArgs in{B.u, B.v, q, q, B.cflx, B.cfly, h->qx, h->qy, B.divx, B.divy, h->qmfy, dt, gw.jb, gw.je, gn.jb, gn.je, g01.jb, g01.je};
    const int w = ffsl_block_width(gw.nlon);
    if (fused_in) { const int rj = ffsl_strip_rows(gw); LAUNCH_LIM2(h, k_ffsl_march, true, ffsl_march_grid(gw, gw.nlev, rj, w), w, 0, gw, h->M, in, rj); }
    else {
      if (xblock) LAUNCH_LIM2(h, k_ffsl_x, true, ffsl_x_grid(gw, gw.nlev, w), dim3(w, FX_ROWS), 0, gw, h->M, in);
      else LAUNCH_LIM2(h, k_ffsl_xw, true, ffsl_xw_grid(gw, gw.nlev, true), 32 * XW_WARPS, 0, gw, h->M, in);
      const int rj = ffsl_strip_rows(gw); LAUNCH_LIM2(h, k_ffsl_y, true, ffsl_y_grid(gw, gw.nlev, rj), 128, 0, gw, h->M, in, rj);
    }
    if (g0.has_south || g0.has_north) {
      run_pole_sum(h, h->qmfy, h->pole_work, nullptr, 0);
      FfslQArgs qa{q, h->qmfx, h->qmfy, B.divx, B.divy, h->pole_work, h->qx, h->qy, dt, g01.jb, g01.je};
      for (int north = 0; north < 2; ++north) {            // the two pole rows (ffsl_mod.F90:375-406)
        if (north ? !g0.has_north : !g0.has_south) continue;
        Geo gp = gw; gp.jb = gp.je = north ? gw.nlat : 1;
        LAUNCH(h, k_ffsl_q, tile_grid(gp, gp.nlev), tile_block(), 0, gp, h->M, qa);
      }
    }
    FfslArgs out{B.mfx, B.mfy, h->qy, h->qx, B.cflx, B.cfly, flx, fly, nullptr, nullptr, nullptr, 0.0, g0.jb, g01.je, gn.jb, gn.je, 0, 0};
    if (fused_out) { const int rj = ffsl_strip_rows(gn); LAUNCH_LIM2(h, k_ffsl_march, false, ffsl_march_grid(gn, gn.nlev, rj, w), w, 0, gn, h->M, out, rj); }
    else {
      if (xblock) LAUNCH_LIM2(h, k_ffsl_x, false, ffsl_x_grid(gn, gn.nlev, w), dim3(w, FX_ROWS), 0, gn, h->M, out);
      else LAUNCH_LIM2(h, k_ffsl_xw, false, ffsl_xw_grid(gn, gn.nlev, false), 32 * XW_WARPS, 0, gn, h->M, out);
      const int rj = ffsl_strip_rows(gn); LAUNCH_LIM2(h, k_ffsl_y, false, ffsl_y_grid(gn, gn.nlev, rj), 128, 0, gn, h->M, out, rj);
    }
    return;
  }
  HflxArgs in{B.u, B.v, q, q, B.cflx, B.cfly, h->qmfx, h->qmfy, gw.jb, gw.je, gn.jb, gn.je};
  LAUNCH_LIM(h, k_hflx, tile_grid(gw, gw.nlev), tile_block(), 0, gw, in);
  run_pole_sum(h, h->qmfy, h->pole_work, nullptr, 0);
  FfslQArgs qa{q, h->qmfx, h->qmfy, B.divx, B.divy, h->pole_work, h->qx, h->qy, dt, xg(h, 0, 1).jb, xg(h, 0, 1).je};
  LAUNCH(h, k_ffsl_q, tile_grid(gw, gw.nlev), tile_block(), 0, gw, h->M, qa);
  HflxArgs out{B.mfx, B.mfy, h->qy, h->qx, B.cflx, B.cfly, flx, fly, g0.jb, xg(h, 0, 1).je, gn.jb, gn.je};
  LAUNCH_LIM(h, k_hflx, tile_grid(gn, gn.nlev), tile_block(), 0, gn, out);
}

// RK4 CFL numbers + divx/divy of a batch: x part on rows jb-3..je+4, y part on half rows jb-1..je+1
static void run_cfl(gmc_handle* h, const double* u, const double* v, double* cflx, double* cfly, double* divx, double* divy, double dt) {
  const Geo gw = xg(h, 3, 4), gn = xg(h, 1, 1);
  CflArgs c{u, v, cflx, cfly, divx, divy, dt, h->cfg.radius, h->dflags, gw.jb, gw.je, gn.jb, gn.je};
  LAUNCH(h, k_cfl, tile_grid(gw, gw.nlev), tile_block(), 0, gw, h->M, c);
  run_pole_sum(h, v, nullptr, divy, 1);
}

static void hydro_accum_wind(gmc_handle* h, int itime);
extern "C" { static int xfer(gmc_handle* h, double* dev, double* host, int kind, int dir); }

// forward-pass diagnostics of a state on rows jb-2..je+2 (pv needs two rows either side)
static void run_forward_prepare_x(gmc_handle* h, DState& S, double dt /* interp_pv_tvd only */) {
  const Geo g2 = xg(h, 2, 2), g1 = xg(h, 1, 1);
  Fwd1Args f{S.u_f, S.v_f, S.m, S.mfx_lon, S.mfy_lat, S.u_lat, S.v_lon, S.ke, S.pv, S.div, h->cfg.ke_scheme, h->cfg.ke_cell_wgt};
  LAUNCH(h, k_fwd1, tile_grid(g2, 1), tile_block(), 0, g2, h->M, f);
  if (h->g.has_south || h->g.has_north) {
    Pole1Args p{S.v_f, S.m, S.u_lat, S.mfy_lat, S.ke, S.div, S.pv, h->pole_dmf, h->cfg.pv_pole_stokes};
    LAUNCH(h, k_pole1, dim3(h->g.nlev, 2), 256, 0, h->g, h->M, p);
  }
  PvEdgeArgs e{S.pv, S.u_f, S.v_f, S.u_lat, S.v_lon, S.pv_lon, S.pv_lat, h->cfg.pv_scheme, h->cfg.upwind_order_pv, h->cfg.upwind_wgt_pv, h->M.rows, dt, h->cfg.tvd_limiter_type};
  LAUNCH(h, k_pv_edge, tile_grid(g1, g1.nlev), tile_block(), 0, g1, e);
}

// One hydrostatic step(old, star, new, dt), compute part.  In band mode the caller then
// exchanges new%{u_f, v_f, pt, phs_f} and calls hydro_after_exchange.
static void hydro_stage(gmc_handle* h, DState& old_, DState& star, DState& nw, double dt) {
  const Geo g = h->g, g01 = xg(h, 0, 1);
  // ---- forward pass on star -------------------------------------------------------
  run_forward_prepare_x(h, star, dt);
  run_cfl(h, star.u_f, star.v_f, h->cflx, h->cfly, h->divx, h->divy, dt);
  WeArgs w{star.mfx_lon, star.mfy_lat, h->pole_dmf, star.m_lev, star.we_lev, h->dphs, h->cflz, dt, h->dflags};
  LAUNCH(h, k_we_col, gridg(g01, 1, WE_TX), dim3(WE_TX, WE_TY), we_col_smem(g.nlev), g01, h->M, w);
  TendArgs t{};
  t.mfx_lon = star.mfx_lon; t.mfy_lat = star.mfy_lat; t.pv_lon = star.pv_lon; t.pv_lat = star.pv_lat; t.ke = star.ke;
  t.we_lev = star.we_lev; t.u_f = star.u_f; t.v_f = star.v_f; t.m = star.m; t.du = h->du; t.dv = h->dv;
  t.coriolis_scheme = h->cfg.coriolis_scheme; t.swm = 0;
  LAUNCH(h, k_tend, tile_grid(g, 1), tile_block(), 0, g, h->M, t);
  BatchRef B{star.u_f, star.v_f, star.mfx_lon, star.mfy_lat, h->cflx, h->cfly, h->divx, h->divy};
  run_ffsl_h(h, B, star.pt, h->ptf_lon, h->ptf_lat, dt);
  run_pole_sum(h, h->ptf_lat, h->pole_work + 2 * g.nlev, nullptr, 1);
  PtColArgs pc{star.pt, h->ptf_lon, h->ptf_lat, h->pole_work + 2 * g.nlev, star.we_lev, h->cflz, old_.pt, old_.m, nw.pt, h->keep_dpt ? h->dpt : nullptr, dt};
  LAUNCH_LIM(h, k_pt_col, gridg(g01, 1), 128, (size_t)g.nlev * 128 * sizeof(double), g01, h->M, pc);
  // ---- update_state(tend1): rows jb..je+1 (the Lin97 PGF of half row je reads row je+1) ----
  LAUNCH(h, k_axpy2d, gridg(g01, 1), 128, 0, g01, old_.phs, h->dphs, nw.phs, dt);
  run_filter(h, nw.phs, 0, 1, GMC_BIG_FILTER, nw.phs_f, -1, nullptr, POST_NONE, nullptr, nullptr, &g01);
  run_ph(h, nw, g01);
  run_filter(h, nw.pt, 0, g.nlev, -1, nullptr, GMC_SMALL_FILTER1, nw.pt, POST_DIV_M, nw.m, nullptr, &g01);
  // ---- backward pass + update_state(tend2) --------------------------------------------
  run_gz(h, nw, g01);
  UvArgs u{};
  u.ph_exn_lev = nw.ph_exn_lev; u.gz_lev = nw.gz_lev; u.du_fwd = h->du; u.dv_fwd = h->dv;
  u.u_old = old_.u; u.v_old = old_.v; u.u_new = nw.u; u.v_new = nw.v; u.dt = dt; u.swm = 0; u.blend = 0;
  LAUNCH(h, k_uv_update_h, tile_grid(g, g.nlev), tile_block(), 0, g, h->M, u);
  // small_filter2 in place, then big_filter -> u_f / v_f from the same staged row.  The two pole-adjacent
  // v rows are blended with their small-filtered neighbours in between (time_schemes_mod.F90:225-233), inside
  // the same kernel; a pole-adjacent block re-reads its neighbour's RAW row, so v is filtered out of place
  // (into the free dv work array) and the two pointers are swapped.
  const int sa = g.has_south ? 1 : -1000, sb = g.has_north ? g.nhalf : -1000;
  auto chain_rows = [&](int j0, int j1) {       // full rows j0..j1 of u, half rows j0..min(j1, hje) of v
    if (j1 < j0) return;
    Geo gr = g; gr.jb = j0; gr.je = j1; gr.hjb = j0; gr.hje = j1 < g.hje ? j1 : g.hje;
    run_filter_chain(h, nw.u, 0, g.nlev, GMC_SMALL_FILTER2, nw.u, GMC_BIG_FILTER, nw.u_f, -1000, -1000, &gr);
    run_filter_chain(h, nw.v, 1, g.nlev, GMC_SMALL_FILTER2, h->dv, GMC_BIG_FILTER, nw.v_f, sa, sb, &gr);
  };
  if (h->comm_stream && g.je - g.jb + 1 > 2 * h->H) {
    // rows that travel in the stage exchange first, then the interior (overlapped with the exchange, do_step)
    const int s1 = g.has_south ? g.jb - 1 : g.jb + h->H - 1, n0 = g.has_north ? g.je + 1 : g.je - h->H + 1;
    chain_rows(g.jb, s1); chain_rows(n0, g.je);
    cudaEventRecord(h->ev_edge, h->stream);
    chain_rows(s1 + 1, n0 - 1);
  } else {
    chain_rows(g.jb, g.je);
    if (h->comm_stream) cudaEventRecord(h->ev_edge, h->stream);
  }
  { double* t = nw.v; nw.v = h->dv; h->dv = t; nw.names["v_lat"] = {nw.v, 1}; }
}
// arrays of a state that travel at the per-stage exchange point
static HaloSet stage_halo_set(gmc_handle* b, int it) {
  DState& S = b->st[it]; HaloSet s = halo_set();
  halo_add(s, S.u_f, b->g.nlev); halo_add(s, S.v_f, b->g.nlev); halo_add(s, S.pt, b->g.nlev); halo_add(s, S.phs_f, 1);
  return s;
}
// after the exchange: m, m_lev, ph_exn_lev of the received halo rows (pure functions of phs_f)
static void hydro_after_exchange(gmc_handle* h, DState& S) {
  if (h->cfg.nranks <= 1) return;
  const Geo& g = h->g;
  if (!g.has_south) { Geo gs = g; gs.jb = g.jb - h->H; gs.je = g.jb - 1; gs.hjb = gs.jb; gs.hje = gs.je; run_ph(h, S, gs); }
  if (!g.has_north) { Geo gn = g; gn.jb = g.je + 1; gn.je = g.je + h->H; gn.hjb = gn.jb; gn.hje = gn.je; run_ph(h, S, gn); }
}

// gmcore_run pre-loop, hydrostatic part (gmcore_mod.F90:104-136) + operators_prepare_1.
// part 1 (before the exchange): filtered fields on the owned rows; part 2: everything else.
static void hydro_prepare_run_1(gmc_handle* h, int old_) {
  DState& S = h->st[old_];
  run_filter(h, S.phs, 0, 1, GMC_BIG_FILTER, S.phs_f, -1, nullptr);
}
static void hydro_prepare_run_2(gmc_handle* h, int old_) {
  DState& S = h->st[old_];
  hydro_accum_wind(h, old_);
  const Geo g44 = xg(h, 4, 4), g01 = xg(h, 0, 1);
  run_ph(h, S, g44);
  run_gz(h, S, g01);
  run_forward_prepare_x(h, S, h->cfg.dt_dyn);
  run_cfl(h, S.u_f, S.v_f, h->cflx, h->cfly, h->divx, h->divy, h->cfg.dt_dyn);
}
// extra arrays of the start-up exchange: phs (old%phs is read on row je+1) and the tracers
static HaloSet prepare_halo_set(gmc_handle* b, int it) {
  HaloSet s = stage_halo_set(b, it);
  halo_add(s, b->st[it].phs, 1);
  halo_add(s, b->gzs, 1);          // static%gzs: the hydrostatic integral runs on row je+1 too
  return s;
}

// adv_accum_wind for the single tracer batch (adv_mod.F90:420-455, adv_batch_mod.F90:
// 239-252,379-403,512-525): copy / add / average the winds of state(itime), and
// recompute the batch CFL numbers when the sample counter wraps (T8).
static void accum_pair(gmc_handle* h, const Geo& g, double* d0, const double* s0, int n0, double* d1, const double* s1, int n1, int mode) {
  AccumArgs a; a.dst[0] = d0; a.src[0] = s0; a.nlev_[0] = n0; a.dst[1] = d1; a.src[1] = s1; a.nlev_[1] = n1;
  a.mode = mode; a.nstep = h->nstep;
  LAUNCH(h, k_accum, gridg(g, n0 > n1 ? n0 : n1), 128, 0, g, a);
}
static int accum_mode(int& step, int nstep) {
  // returns the mode for this call and advances the counter (non-dynamic batch)
  if (step == 0) { step = 1; return 0; }
  if (step == nstep) { step = 0; return 2; }
  step += 1; return 1;
}
static void hydro_accum_wind(gmc_handle* h, int itime) {
  const Geo& g = h->g; DState& S = h->st[itime];
  for (size_t m = 0; m < h->batches.size(); ++m) {
    TrBatch& B = h->batches[m];
    if (B.wind_of != (int)m) continue;      // shares the samples of an earlier batch with the same dt
    h->nstep = B.nstep;
    int mode = accum_mode(B.uv_step, B.nstep);
    accum_pair(h, xg(h, 4, 4), B.bu, S.u_f, g.nlev, B.bv, S.v_f, g.nlev, mode);
    if (B.uv_step == 0) run_cfl(h, B.bu, B.bv, B.bcflx, B.bcfly, B.bdivx, B.bdivy, B.dt);
    mode = accum_mode(B.mf_step, B.nstep);
    accum_pair(h, xg(h, 1, 1), B.bmfx, S.mfx_lon, g.nlev, B.bmfy, S.mfy_lat, g.nlev, mode);
    mode = accum_mode(B.we_step, B.nstep);
    accum_pair(h, g, B.bwe, S.we_lev, g.nlev + 1, B.bm_lev, S.m_lev, g.nlev + 1, mode);
    if (B.we_step == 0) {
      CflzArgs c{B.bwe, B.bm_lev, B.bcflz, B.dt, h->dflags};
      LAUNCH(h, k_cflz, gridg(g, 1, WE_TX), dim3(WE_TX, WE_TY), cflz_smem(g.nlev), g, h->M, c);
    }
  }
}

// adv_run (adv_mod.F90:152-312), compute part; returns true when some batch was advected
// (band mode: the caller then exchanges q(new) of those batches and calls hydro_adv_finish on every band).
static bool hydro_adv_run(gmc_handle* h, int itime, int time_step) {
  hydro_accum_wind(h, itime);
  const Geo& g = h->g; DState& S = h->st[itime];
  const size_t n3 = (size_t)g.plane * g.nlev;
  const double dt_adv = h->cfg.dt_adv;      // the update uses the namelist dt_adv, not batch%dt (adv_mod.F90:189,203,213,264)
  bool any = false;
  for (TrBatch& T : h->batches) {
    const TrBatch& W = h->batches[T.wind_of];
    bool alerted = false;
    if (T.alert_last + T.alert_period <= time_step * h->cfg.dt_dyn) { T.alert_ring = true; alerted = true; }
    T.advected = alerted && time_step > 0;
    if (!T.advected) continue;
    any = true;
    BatchRef B{W.bu, W.bv, W.bmfx, W.bmfy, W.bcflx, W.bcfly, W.bdivx, W.bdivy};
    for (int l = 0; l < T.ntr; ++l) {
      const double* qo = T.q[T.q_old] + l * n3; double* qn = T.q[T.q_new] + l * n3;
      run_ffsl_h(h, B, qo, h->ptf_lon, h->ptf_lat, T.dt);   // ptf_* double as qmf_lon / qmf_lat; inner step with batch%dt (ffsl_mod.F90:333)
      run_pole_sum(h, h->ptf_lat, h->pole_work + 2 * g.nlev, nullptr, 0);
      TrHArgs th{qo, h->old_m, S.m, h->ptf_lon, h->ptf_lat, h->pole_work + 2 * g.nlev, qn, dt_adv};
      LAUNCH(h, k_tracer_h, gridg(g, g.nlev), 128, 0, g, h->M, th);
      NegXArgs nx{qn, S.m, h->dflags};
      LAUNCH(h, k_negfill_x, dim3(g.je - g.jb + 1, g.nlev), 128, 0, g, nx);
      TrVArgs tv{qn, S.m, W.bwe, W.bcflz, dt_adv, h->dflags};
      LAUNCH_LIM(h, k_tracer_v, gridg(g, 1), 128, (size_t)g.nlev * 128 * sizeof(double), g, tv);
      LonDampArgs ld{qn, h->lon_damp_wgt, 3};
      LAUNCH(h, k_lon_damp, dim3(g.je - g.jb + 1, g.nlev), 256, (size_t)2 * g.nlon * sizeof(double), g, ld);
    }
  }
  return any;
}
static void hydro_adv_finish(gmc_handle* h, int itime) {
  if (h->batches.empty()) return;
  const Geo& g = h->g;
  for (TrBatch& T : h->batches) if (T.advected) { int t = T.q_old; T.q_old = T.q_new; T.q_new = t; T.advected = false; }
  LAUNCH(h, k_copy3, dim3(g.je - g.jb + 1, g.nlev), 128, 0, g, h->st[itime].m, h->old_m);   // copy_old_m, every batch
}
// tracer arrays of batch m that travel at an exchange point (which: 0 q_old, 1 q_new)
static HaloSet tracer_halo_set(gmc_handle* b, int m, int which) {
  HaloSet s = halo_set();
  const size_t n3 = (size_t)b->g.plane * b->g.nlev;
  const TrBatch& T = b->batches[m];
  for (int l = 0; l < T.ntr; ++l) halo_add(s, T.q[which ? T.q_new : T.q_old] + l * n3, b->g.nlev);
  return s;
}

// smag_damp_run (damp_mod.F90:60-62) on state(new)%u_lon, v_lat.  In band mode the caller has exchanged the halo
// rows of u, v first (steps: strain on rows jb-1 .. je+1, tendencies and update on the owned rows).
static void run_smag_damp(gmc_handle* h, DState& S, double dt) {
  const Geo& g = h->g; const Geo g11 = xg(h, 1, 1);
  SmagArgs a{S.u, S.v, h->smag_t, h->smag_s, h->smag_du, h->smag_dv, h->cfg.smag_damp_coef, dt};
  LAUNCH(h, k_smag_strain, gridg(g11, g.nlev), 128, 0, g11, h->M, a);
  LAUNCH(h, k_smag_tend, gridg(g, g.nlev), 128, 0, g, h->M, a);
  LAUNCH(h, k_smag_apply, gridg(g, g.nlev), 128, 0, g, a, S.u, S.v);
}

// div_damp_run (div_damp_mod.F90:89-139) on state S, both cores.  div is recomputed from the filtered winds
// (calc_div); order 4 needs div one row beyond the rows of div2 (bands: u_f, v_f halo rows are there, H = 4).
static void run_div_damp(gmc_handle* h, DState& S) {
  const Geo& g = h->g;
  const bool o4 = h->cfg.div_damp_order == 4;
  const Geo gd = o4 ? xg(h, 1, 2) : xg(h, 0, 1);
  DivArgs d{S.u_f, S.v_f, S.div};
  LAUNCH(h, k_div, gridg(gd, g.nlev), 128, 0, gd, h->M, d);
  run_pole_sum(h, S.v_f, nullptr, S.div, 1);
  const double* dsrc = S.div; double sgn = 1.0;
  if (o4) {     // u -= c d(div2)/dx with div2 = lap(div) (div_damp_mod.F90:120-134); pole rows of div2 stay 0
    const Geo g01 = xg(h, 0, 1);
    Div2Args d2{S.div, h->div2};
    LAUNCH(h, k_div2, gridg(g01, g.nlev), 128, 0, g01, h->M, d2);
    dsrc = h->div2; sgn = -1.0;
  }
  DivDampArgs dd{dsrc, h->c_lon, h->c_lat, S.u, S.v, sgn};
  LAUNCH(h, k_div_damp, gridg(g, g.nlev), 128, 0, g, h->M, dd);
}

// time_integrate after the time integrator (gmcore_mod.F90:586-590): forcing and divergence damping (part 1), then --
// after the u, v halo exchange that Smagorinsky damping needs across bands -- Smagorinsky damping and tracers (part 2)
static void hydro_post_compute_1(gmc_handle* h, int nw, double dt) {
  const Geo& g = h->g; DState& S = h->st[nw];
  if (h->cfg.test_forcing == GMC_FORCING_HELD_SUAREZ) {
    HsArgs a{S.u, S.v, S.pt, S.t, S.phs_f, dt, make_vc(h)};
    LAUNCH(h, k_held_suarez, gridg(g, g.nlev), 128, 0, g, h->M, a);
  }
  if (h->cfg.use_div_damp) run_div_damp(h, S);
}
static bool hydro_post_compute_2(gmc_handle* h, int nw, double dt, int time_step) {
  if (h->cfg.use_smag_damp) run_smag_damp(h, h->st[nw], dt);
  return hydro_adv_run(h, nw, time_step);
}

// Maps the reference's log_error guards onto the ABI error text.
static void hydro_flags_text(gmc_handle* h, int f) {
  snprintf(h->err, sizeof(h->err), "%s%s%s%s%s",
           f & 1 ? "cflx exceeds mesh%lon_halo_width; " : "", f & 2 ? "cfly exceeds 1; " : "", f & 4 ? "cflz exceeds range; " : "",
           f & 8 ? "The pressure levels are not monotonic!; " : "", f & 16 ? "Negative tracer!; " : "");
}
// Reads the device guard word; non-zero -> error text, word cleared (reported once).
static int hydro_check_flags(gmc_handle* h) {
  if (!h->dflags) return 0;
  int f = 0;
  if (cudaMemcpyAsync(&f, h->dflags, sizeof(int), cudaMemcpyDeviceToHost, h->stream) != cudaSuccess) return 0;
  cudaStreamSynchronize(h->stream);
  if (!f) return 0;
  hydro_flags_text(h, f);
  cudaMemsetAsync(h->dflags, 0, sizeof(int), h->stream);
  return 1;
}

static size_t host_cell_elems(const gmc_handle* h) {
  return (size_t)(h->cfg.nlon + 2 * h->cfg.hx) * ((h->g.je - h->g.jb + 1) + 2 * h->cfg.hy) * (h->cfg.nlev + 4);
}
static int hydro_set_tracers(gmc_handle* h, int m, const double* q) {
  if (m < 0 || m >= (int)h->batches.size()) { snprintf(h->err, sizeof(h->err), "no tracer batch %d configured", m); return 1; }
  const size_t n3 = (size_t)h->g.plane * h->g.nlev, host_n = host_cell_elems(h);
  TrBatch& T = h->batches[m];
  for (int l = 0; l < T.ntr; ++l)
    if (xfer(h, T.q[T.q_old] + l * n3, const_cast<double*>(q) + l * host_n, 0, 0)) return 1;
  return 0;
}
static int hydro_get_tracers(gmc_handle* h, int m, double* q) {
  if (m < 0 || m >= (int)h->batches.size()) { snprintf(h->err, sizeof(h->err), "no tracer batch %d configured", m); return 1; }
  const size_t n3 = (size_t)h->g.plane * h->g.nlev, host_n = host_cell_elems(h);
  TrBatch& T = h->batches[m];
  for (int l = 0; l < T.ntr; ++l)
    if (xfer(h, T.q[T.q_old] + l * n3, q + l * host_n, 0, 1)) return 1;
  return 0;
}

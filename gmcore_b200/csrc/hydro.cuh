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
  LAUNCH(h, k_ph, gridg(g, g.nlev + 1), 128, 0, g, h->M, p);
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
  if (h->cfg.ntracers <= 0) return;
  const Geo& g = h->g; DState& S = h->st[itime];
  int mode = accum_mode(h->uv_step, h->nstep);
  accum_pair(h, xg(h, 4, 4), h->bu, S.u_f, g.nlev, h->bv, S.v_f, g.nlev, mode);
  if (h->uv_step == 0) run_cfl(h, h->bu, h->bv, h->bcflx, h->bcfly, h->bdivx, h->bdivy, h->cfg.dt_adv);
  mode = accum_mode(h->mf_step, h->nstep);
  accum_pair(h, xg(h, 1, 1), h->bmfx, S.mfx_lon, g.nlev, h->bmfy, S.mfy_lat, g.nlev, mode);
  mode = accum_mode(h->we_step, h->nstep);
  accum_pair(h, g, h->bwe, S.we_lev, g.nlev + 1, h->bm_lev, S.m_lev, g.nlev + 1, mode);
  if (h->we_step == 0) {
    CflzArgs c{h->bwe, h->bm_lev, h->bcflz, h->cfg.dt_adv, h->dflags};
    LAUNCH(h, k_cflz, gridg(g, 1, WE_TX), dim3(WE_TX, WE_TY), cflz_smem(g.nlev), g, h->M, c);
  }
}

// adv_run (adv_mod.F90:152-312), compute part; returns true when the tracers were advected
// (band mode: the caller then exchanges q(new) and calls hydro_adv_finish on every band).
static bool hydro_adv_run(gmc_handle* h, int itime, int time_step) {
  hydro_accum_wind(h, itime);
  if (h->cfg.ntracers <= 0) return false;
  const Geo& g = h->g; DState& S = h->st[itime];
  const size_t n3 = (size_t)g.plane * g.nlev;
  const double dt_adv = h->cfg.dt_adv;
  bool alerted = false;
  if (h->alert_last + h->alert_period <= time_step * h->cfg.dt_dyn) { h->alert_ring = true; alerted = true; }
  const bool adv = alerted && time_step > 0;
  if (adv) {
    BatchRef B{h->bu, h->bv, h->bmfx, h->bmfy, h->bcflx, h->bcfly, h->bdivx, h->bdivy};
    for (int l = 0; l < h->cfg.ntracers; ++l) {
      const double* qo = h->q[h->q_old] + l * n3; double* qn = h->q[h->q_new] + l * n3;
      run_ffsl_h(h, B, qo, h->ptf_lon, h->ptf_lat, dt_adv);   // ptf_* double as qmf_lon / qmf_lat
      run_pole_sum(h, h->ptf_lat, h->pole_work + 2 * g.nlev, nullptr, 0);
      TrHArgs th{qo, h->old_m, S.m, h->ptf_lon, h->ptf_lat, h->pole_work + 2 * g.nlev, qn, dt_adv};
      LAUNCH(h, k_tracer_h, gridg(g, g.nlev), 128, 0, g, h->M, th);
      NegXArgs nx{qn, S.m, h->dflags};
      LAUNCH(h, k_negfill_x, dim3(g.je - g.jb + 1, g.nlev), 128, 0, g, nx);
      TrVArgs tv{qn, S.m, h->bwe, h->bcflz, dt_adv, h->dflags};
      LAUNCH_LIM(h, k_tracer_v, gridg(g, 1), 128, (size_t)g.nlev * 128 * sizeof(double), g, tv);
      LonDampArgs ld{qn, h->lon_damp_wgt, 3};
      LAUNCH(h, k_lon_damp, dim3(g.je - g.jb + 1, g.nlev), 256, (size_t)2 * g.nlon * sizeof(double), g, ld);
    }
  }
  return adv;
}
static void hydro_adv_finish(gmc_handle* h, int itime, bool advected) {
  if (h->cfg.ntracers <= 0) return;
  const Geo& g = h->g;
  if (advected) { int t = h->q_old; h->q_old = h->q_new; h->q_new = t; }
  LAUNCH(h, k_copy3, dim3(g.je - g.jb + 1, g.nlev), 128, 0, g, h->st[itime].m, h->old_m);   // copy_old_m
}
static HaloSet tracer_halo_set(gmc_handle* b, int which /*0: q_old, 1: q_new*/, int with_pt_of = 0) {
  HaloSet s = halo_set();
  const size_t n3 = (size_t)b->g.plane * b->g.nlev;
  if (b->cfg.ntracers > 0)
    for (int l = 0; l < b->cfg.ntracers; ++l) halo_add(s, b->q[which ? b->q_new : b->q_old] + l * n3, b->g.nlev);
  if (with_pt_of) halo_add(s, b->st[with_pt_of].pt, b->g.nlev);
  return s;
}

// smag_damp_run (damp_mod.F90:60-62) on state(new)%u_lon, v_lat
static void run_smag_damp(gmc_handle* h, DState& S, double dt) {
  const Geo& g = h->g;
  SmagArgs a{S.u, S.v, h->smag_t, h->smag_s, h->smag_du, h->smag_dv, h->cfg.smag_damp_coef, dt};
  LAUNCH(h, k_smag_strain, gridg(g, g.nlev), 128, 0, g, h->M, a);
  LAUNCH(h, k_smag_tend, gridg(g, g.nlev), 128, 0, g, h->M, a);
  LAUNCH(h, k_smag_apply, gridg(g, g.nlev), 128, 0, g, a, S.u, S.v);
}

// time_integrate after the time integrator (gmcore_mod.F90:586-590): forcing, damping, tracers
static bool hydro_post_compute(gmc_handle* h, int nw, double dt, int time_step) {
  const Geo& g = h->g; DState& S = h->st[nw];
  if (h->cfg.test_forcing == GMC_FORCING_HELD_SUAREZ) {
    HsArgs a{S.u, S.v, S.pt, S.t, S.phs_f, dt, make_vc(h)};
    LAUNCH(h, k_held_suarez, gridg(g, g.nlev), 128, 0, g, h->M, a);
  }
  if (h->cfg.use_div_damp) {
    const Geo g01 = xg(h, 0, 1);
    DivArgs d{S.u_f, S.v_f, S.div};
    LAUNCH(h, k_div, gridg(g01, g.nlev), 128, 0, g01, h->M, d);
    run_pole_sum(h, S.v_f, nullptr, S.div, 1);
    const double* dsrc = S.div; double sgn = 1.0;
    if (h->cfg.div_damp_order == 4) {     // u -= c d(div2)/dx with div2 = lap(div) (div_damp_mod.F90:120-134)
      Div2Args d2{S.div, h->qx};          // qx is free here; its pole rows hold FFSL data, so clear them first
      cudaMemsetAsync(h->qx, 0, (size_t)g.plane * g.nlev * sizeof(double), h->stream);
      LAUNCH(h, k_div2, gridg(g, g.nlev), 128, 0, g, h->M, d2);
      dsrc = h->qx; sgn = -1.0;
    }
    DivDampArgs dd{dsrc, h->c_lon, h->c_lat, S.u, S.v, sgn};
    LAUNCH(h, k_div_damp, gridg(g, g.nlev), 128, 0, g, h->M, dd);
  }
  if (h->cfg.use_smag_damp) run_smag_damp(h, S, dt);
  return hydro_adv_run(h, nw, time_step);
}

// Maps the reference's log_error guards onto the ABI error text.
static int hydro_check_flags(gmc_handle* h) {
  if (!h->dflags) return 0;
  int f = 0;
  if (cudaMemcpyAsync(&f, h->dflags, sizeof(int), cudaMemcpyDeviceToHost, h->stream) != cudaSuccess) return 0;
  cudaStreamSynchronize(h->stream);
  if (!f) return 0;
  snprintf(h->err, sizeof(h->err), "%s%s%s%s%s",
           f & 1 ? "cflx exceeds mesh%lon_halo_width; " : "", f & 2 ? "cfly exceeds 1; " : "", f & 4 ? "cflz exceeds range; " : "",
           f & 8 ? "The pressure levels are not monotonic!; " : "", f & 16 ? "Negative tracer!; " : "");
  return 1;
}

static size_t host_cell_elems(const gmc_handle* h) {
  return (size_t)(h->cfg.nlon + 2 * h->cfg.hx) * ((h->g.je - h->g.jb + 1) + 2 * h->cfg.hy) * (h->cfg.nlev + 4);
}
static int hydro_set_tracers(gmc_handle* h, const double* q) {
  if (h->cfg.ntracers <= 0 || !h->cfg.baroclinic) { snprintf(h->err, sizeof(h->err), "no tracer batch configured"); return 1; }
  const size_t n3 = (size_t)h->g.plane * h->g.nlev, host_n = host_cell_elems(h);
  for (int l = 0; l < h->cfg.ntracers; ++l)
    if (xfer(h, h->q[h->q_old] + l * n3, const_cast<double*>(q) + l * host_n, 0, 0)) return 1;
  return 0;
}
static int hydro_get_tracers(gmc_handle* h, double* q) {
  if (h->cfg.ntracers <= 0 || !h->cfg.baroclinic) { snprintf(h->err, sizeof(h->err), "no tracer batch configured"); return 1; }
  const size_t n3 = (size_t)h->g.plane * h->g.nlev, host_n = host_cell_elems(h);
  for (int l = 0; l < h->cfg.ntracers; ++l)
    if (xfer(h, h->q[h->q_old] + l * n3, q + l * host_n, 0, 1)) return 1;
  return 0;
}

// libgmcore_b200.so — C ABI of include/gmcore_gpu.h: the GMCORE dynamics hot path
// (time_integrate and the pre-loop of gmcore_run) as sm_100a CUDA kernels.
// There is no CPU fallback: every entry point fails if no CUDA device is usable.
#include <algorithm>
#include <cstdlib>
#include <mutex>
#include "handle.cuh"
#include "kernels_dyn.cuh"
#include "kernels_update.cuh"
#include "kernels_hydro.cuh"
extern "C" { static int xfer(gmc_handle* h, double* dev, double* host, int kind, int dir); }
#include "steps.cuh"

extern "C" {

int gmc_last_error(gmc_handle* h, char* buf, int len) {
  if (!h) { snprintf(buf, len, "null handle"); return 1; }
  snprintf(buf, len, "%s", h->err);
  return h->err[0] ? 1 : 0;
}

static char g_create_err[512];
const char* gmc_create_error() { return g_create_err; }

int gmc_create(const gmc_config* cfg, gmc_handle** out) {
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    snprintf(g_create_err, sizeof(g_create_err), "no CUDA device (%s); the GMCORE GPU core has no CPU fallback",
             cudaGetErrorString(e));
    return 2;
  }
  gmc_handle* h = new gmc_handle();
  h->cfg = *cfg;
  if (cudaSetDevice(cfg->device) != cudaSuccess) { snprintf(g_create_err, sizeof(g_create_err), "cudaSetDevice(%d) failed", cfg->device); delete h; return 2; }
  if (cfg->limiter_type < 0 || cfg->limiter_type > 2) { snprintf(g_create_err, sizeof(g_create_err), "Invalid limiter_type %d!", cfg->limiter_type); delete h; return 3; }
  if (cfg->pv_scheme < 0 || cfg->pv_scheme > 2) { snprintf(g_create_err, sizeof(g_create_err), "Invalid pv_scheme %d! (midpoint, upwind, tvd; 'ffsl' is not functional in the reference either)", cfg->pv_scheme); delete h; return 3; }
  if (cfg->tvd_limiter_type < 0 || cfg->tvd_limiter_type > 2) { snprintf(g_create_err, sizeof(g_create_err), "Invalid tvd_limiter_type %d!", cfg->tvd_limiter_type); delete h; return 3; }
  if (cfg->ffsl_flux_type < 0 || cfg->ffsl_flux_type > 1) { snprintf(g_create_err, sizeof(g_create_err), "Invalid ffsl_flux_type %d!", cfg->ffsl_flux_type); delete h; return 3; }
  if (cfg->nlon % 2 != 0) { snprintf(g_create_err, sizeof(g_create_err), "num_lon must be even"); delete h; return 3; }
  if (cfg->coriolis_scheme != 1 && cfg->coriolis_scheme != 2) { snprintf(g_create_err, sizeof(g_create_err), "coriolis_scheme=%d is not a GMCORE scheme (1 or 2)", cfg->coriolis_scheme); delete h; return 3; }
  if (cfg->use_div_damp && cfg->div_damp_order != 2 && cfg->div_damp_order != 4) { snprintf(g_create_err, sizeof(g_create_err), "Unsupported div_damp_order %d!", cfg->div_damp_order); delete h; return 3; }
  if (cfg->nranks > 1 && !cfg->baroclinic) { snprintf(g_create_err, sizeof(g_create_err), "latitude bands (nranks>1) are built for the hydrostatic core only"); delete h; return 3; }
  if (cfg->nranks > 1 && cfg->lat_iend - cfg->lat_ibeg + 1 < 6) { snprintf(g_create_err, sizeof(g_create_err), "a latitude band needs at least 6 rows"); delete h; return 3; }
  if ((cfg->rank == 0) != (cfg->lat_ibeg == 1) || (cfg->rank == cfg->nranks - 1) != (cfg->lat_iend == cfg->nlat)) {
    snprintf(g_create_err, sizeof(g_create_err), "rank %d of %d does not match rows %d..%d", cfg->rank, cfg->nranks, cfg->lat_ibeg, cfg->lat_iend); delete h; return 3; }
  Geo& g = h->g;
  g.nlon = cfg->nlon; g.nlat = cfg->nlat; g.nlev = cfg->nlev; g.nhalf = cfg->nlat - 1;
  g.jb = cfg->lat_ibeg; g.je = cfg->lat_iend;
  g.has_south = g.jb == 1; g.has_north = g.je == g.nlat;
  g.hjb = g.jb; g.hje = g.has_north ? g.je - 1 : g.je;
  const int HL = cfg->nranks > 1 ? 6 : GMC_HL;   // band mode: 4 exchanged rows + 2 rows of stencil over-reach
  g.jlo = g.jb - HL; g.NY = (g.je - g.jb + 1) + 2 * HL;
  g.plane = (long long)g.NY * g.nlon;
  if (g.plane * (g.nlev + 2) >= (1LL << 31)) { snprintf(g_create_err, sizeof(g_create_err), "band too large for 32-bit field indices"); delete h; return 3; }
  cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
  const size_t n3 = (size_t)g.plane * g.nlev, n3h = (size_t)g.plane * (g.nlev + 1), n2 = (size_t)g.plane;
  const bool baro = cfg->baroclinic != 0;
  int rc = 0;
  for (int s = 1; s <= 3 && !rc; ++s) {
    DState& S = h->st[s];
    auto A = [&](const char* name, double** p, size_t n, int kind) { if (!rc) { rc = dev_alloc(h, p, n); S.names[name] = {*p, kind}; } };
    A("u_lon", &S.u, n3, 0); A("v_lat", &S.v, n3, 1); A("u_f", &S.u_f, n3, 0); A("v_f", &S.v_f, n3, 1);
    A("m", &S.m, n3, 0); A("mfx_lon", &S.mfx_lon, n3, 0); A("mfy_lat", &S.mfy_lat, n3, 1);
    A("u_lat", &S.u_lat, n3, 1); A("v_lon", &S.v_lon, n3, 0); A("ke", &S.ke, n3, 0); A("pv", &S.pv, n3, 1);
    A("pv_lon", &S.pv_lon, n3, 0); A("pv_lat", &S.pv_lat, n3, 1); A("div", &S.div, n3, 0);
    if (!baro) { A("gz", &S.gz, n3, 0); A("gz_f", &S.gz_f, n3, 0); }
    else {
      A("pt", &S.pt, n3, 0); A("phs", &S.phs, n2, 3); A("phs_f", &S.phs_f, n2, 3);
      A("ph_exn_lev", &S.ph_exn_lev, n3h, 2); A("t", &S.t, n3, 0);
      A("gz_lev", &S.gz_lev, n3h, 2); A("we_lev", &S.we_lev, n3h, 2); A("m_lev", &S.m_lev, n3h, 2);
    }
  }
  if (!rc) rc = dev_alloc(h, &h->du, n3);
  if (!rc) rc = dev_alloc(h, &h->dv, n3);
  if (!rc) rc = dev_alloc(h, &h->gzs, n2);
  if (!rc) rc = dev_alloc(h, &h->pole_dmf, (size_t)2 * g.nlev);
  if (!rc) rc = dev_alloc(h, &h->pole_work, (size_t)16 * g.nlev);
  if (!rc) rc = dev_alloc(h, &h->diag_partial, (size_t)7 * (g.je - g.jb + 1) * g.nlev);
  if (!rc) rc = dev_alloc(h, &h->diag_out, 8);
  if (!rc && cfg->use_smag_damp) { rc = dev_alloc(h, &h->smag_t, n3); if (!rc) rc = dev_alloc(h, &h->smag_s, n3); if (!rc) rc = dev_alloc(h, &h->smag_du, n3); if (!rc) rc = dev_alloc(h, &h->smag_dv, n3); }
  h->H = cfg->nranks > 1 ? 4 : 0;      // halo rows exchanged per side and stage (band mode)
  if (!rc && baro) rc = hydro_alloc(h);
  if (!rc) rc = halo_alloc(h);
  h->stage_bytes = (size_t)(g.nlon + 2 * cfg->hx) * ((g.je - g.jb + 1) + 2 * cfg->hy) * (g.nlev + 4) * sizeof(double);
  if (!rc) { double* p; rc = dev_alloc(h, &p, h->stage_bytes / sizeof(double)); h->stage_img = p; }
  if (rc) { snprintf(g_create_err, sizeof(g_create_err), "%s", h->err); gmc_destroy(h); return 4; }
  { const char* e = getenv("GMCORE_DEBUG_FIELDS"); h->keep_dpt = e && e[0] == '1'; }
  *out = h;
  return 0;
}

void gmc_destroy(gmc_handle* h) {
  if (!h) return;
  cudaSetDevice(h->cfg.device);
  cudaDeviceSynchronize();
  for (void* p : h->allocs) cudaFree(p);
  if (h->diag_steps) cudaFree(h->diag_steps);
  for (int q = 0; q < 2; ++q) if (h->swm_graph[q]) cudaGraphExecDestroy(h->swm_graph[q]);
  if (h->nccl_comm && nccl_api().CommDestroy) nccl_api().CommDestroy(h->nccl_comm);
  if (h->group) {   // the group vector is shared: the last member to go frees it
    auto* grp = h->group;
    for (auto& m : *grp) if (m == h) m = nullptr;
    bool any = false; for (auto m : *grp) any = any || m;
    if (!any) delete grp;
  }
  if (h->comm_stream) { cudaStreamDestroy(h->comm_stream); cudaEventDestroy(h->ev_edge); cudaEventDestroy(h->ev_xdone); }
  if (h->stream && h->owns_stream) cudaStreamDestroy(h->stream);
  delete h;
}

int gmc_set_mesh(gmc_handle* h, const gmc_mesh_vectors* mv) {
  const gmc_config& c = h->cfg;
  const size_t nf = c.nlat + 2 * c.hy, nh = c.nlat - 1 + 2 * c.hy, nl = c.nlon + 2 * c.hx;
  if (c.hy != 2) { snprintf(h->err, sizeof(h->err), "lat_halo_width must be 2"); return 1; }
  DMesh& M = h->M;
  int rc = 0;
#define UPF(x) if (!rc) rc = dev_upload(h, &M.x, mv->x, nf)
#define UPH(x) if (!rc) rc = dev_upload(h, &M.x, mv->x, nh)
  UPF(full_lat); UPH(half_lat); UPF(full_cos_lat); UPH(half_cos_lat); UPF(full_sin_lat); UPH(half_sin_lat);
  UPF(area_cell); UPF(area_lon); UPF(area_lon_west); UPF(area_lon_east); UPF(area_lon_north); UPF(area_lon_south);
  UPH(area_lat); UPH(area_lat_west); UPH(area_lat_east); UPH(area_lat_north); UPH(area_lat_south); UPH(area_vtx);
  UPF(de_lon); UPF(le_lon); UPH(de_lat); UPH(le_lat); UPF(full_f); UPH(half_f);
  if (!rc) rc = dev_upload(h, &M.area_subcell, mv->area_subcell, 2 * nf);
  if (!rc) rc = dev_upload(h, &M.full_tangent_wgt, mv->full_tangent_wgt, 2 * nf);
  if (!rc) rc = dev_upload(h, &M.half_tangent_wgt, mv->half_tangent_wgt, 2 * nh);
  if (!rc) rc = dev_upload(h, &M.half_lon, mv->half_lon, nl);
  {
    auto recip = [&](const double** dst, const double* src, size_t n) {
      std::vector<double> r(n);
      for (size_t q = 0; q < n; ++q) r[q] = src[q] != 0.0 ? 1.0 / src[q] : 0.0;
      if (!rc) rc = dev_upload(h, dst, r.data(), n);
    };
    recip(&M.r_area_cell, mv->area_cell, nf); recip(&M.r_area_vtx, mv->area_vtx, nh);
    recip(&M.r_de_lon, mv->de_lon, nf); recip(&M.r_de_lat, mv->de_lat, nh);
    recip(&M.r_area_lon, mv->area_lon, nf); recip(&M.r_area_lat, mv->area_lat, nh);
    std::vector<RowCoef> rows(nf);
    auto rc1 = [](double x) { return x != 0.0 ? 1.0 / x : 0.0; };
    for (size_t q = 0; q < nf; ++q) {
      RowCoef& r = rows[q]; memset(&r, 0, sizeof(r));
      r.area_cell = mv->area_cell[q]; r.r_area_cell = rc1(r.area_cell);
      r.area_lon_west = mv->area_lon_west[q]; r.area_lon_east = mv->area_lon_east[q];
      r.area_lon_north = mv->area_lon_north[q]; r.area_lon_south = mv->area_lon_south[q];
      r.area_lon = mv->area_lon[q]; r.r_area_lon = rc1(r.area_lon);
      r.de_lon = mv->de_lon[q]; r.r_de_lon = rc1(r.de_lon); r.le_lon = mv->le_lon[q];
      r.full_cos_lat = mv->full_cos_lat[q]; r.full_sin_lat = mv->full_sin_lat[q];
      r.ftw1 = mv->full_tangent_wgt[2 * q]; r.ftw2 = mv->full_tangent_wgt[2 * q + 1];
      r.sub1 = mv->area_subcell[2 * q]; r.sub2 = mv->area_subcell[2 * q + 1];
      if (q < nh) {
        r.area_lat_west = mv->area_lat_west[q]; r.area_lat_east = mv->area_lat_east[q];
        r.area_lat_north = mv->area_lat_north[q]; r.area_lat_south = mv->area_lat_south[q];
        r.area_lat = mv->area_lat[q]; r.r_area_lat = rc1(r.area_lat); r.r_area_vtx = rc1(mv->area_vtx[q]);
        r.de_lat = mv->de_lat[q]; r.r_de_lat = rc1(r.de_lat); r.le_lat = mv->le_lat[q];
        r.half_f = mv->half_f[q]; r.htw1 = mv->half_tangent_wgt[2 * q]; r.htw2 = mv->half_tangent_wgt[2 * q + 1];
        r.half_lat = mv->half_lat[q];
      }
    }
    if (!rc) rc = dev_upload(h, &M.rows, rows.data(), nf);
  }
  if (c.baroclinic) {
    if (!mv->hyai) { snprintf(h->err, sizeof(h->err), "hybrid coefficients missing"); return 1; }
    if (!rc) rc = dev_upload(h, &M.hyai, mv->hyai, (size_t)c.nlev + 1);
    if (!rc) rc = dev_upload(h, &M.hybi, mv->hybi, (size_t)c.nlev + 1);
    if (!rc) rc = dev_upload(h, &M.half_lev, mv->half_lev, (size_t)c.nlev + 1);
    if (!rc) rc = dev_upload(h, &M.half_dlev, mv->half_dlev, (size_t)c.nlev + 1);
    if (!rc) rc = dev_upload(h, &M.full_lev, mv->full_lev, (size_t)c.nlev + 4);
    if (!rc) rc = dev_upload(h, &M.full_dlev, mv->full_dlev, (size_t)c.nlev + 4);
    {
      std::vector<double> r((size_t)c.nlev + 4);
      for (size_t q = 0; q < r.size(); ++q) r[q] = mv->full_dlev[q] != 0.0 ? 1.0 / mv->full_dlev[q] : 0.0;
      if (!rc) rc = dev_upload(h, &M.r_full_dlev, r.data(), r.size());
    }
    if (!rc) rc = hydro_set_mesh(h, mv);
  }
  if (!rc) rc = damp_set_mesh(h, mv);
#undef UPF
#undef UPH
  M.hx = c.hx;
  h->mesh_set = !rc;
  return rc;
}

int gmc_set_filters(gmc_handle* h, int which, const int* ngrid_lon, const double* wgt_lon, int ld_lon,
                    const int* ngrid_lat, const double* wgt_lat, int ld_lat) {
  if (which < 0 || which > 2) { snprintf(h->err, sizeof(h->err), "bad filter id %d", which); return 1; }
  const gmc_config& c = h->cfg;
  const size_t nf = c.nlat + 2 * c.hy, nh = c.nlat - 1 + 2 * c.hy;
  DFilter& F = h->filt[which];
  int rc = dev_upload(h, &F.ngrid_lon, ngrid_lon, nf);
  if (!rc) rc = dev_upload(h, &F.ngrid_lat, ngrid_lat, nh);
  if (!rc) rc = dev_upload(h, &F.wgt_lon, wgt_lon, nf * ld_lon);
  if (!rc) rc = dev_upload(h, &F.wgt_lat, wgt_lat, nh * ld_lat);
  F.ld_lon = ld_lon; F.ld_lat = ld_lat;
  int mx = 0;
  for (size_t j = 0; j < nf; ++j) mx = ngrid_lon[j] > mx ? ngrid_lon[j] : mx;
  for (size_t j = 0; j < nh; ++j) mx = ngrid_lat[j] > mx ? ngrid_lat[j] : mx;
  F.max_hn = mx > 1 ? (mx - 1) / 2 : 0;
  if (F.max_hn > h->max_hn) h->max_hn = F.max_hn;
  if (F.max_hn >= c.nlon) { snprintf(h->err, sizeof(h->err), "filter wider than the zonal circle"); return 1; }
  // the row filter stages two rows and two weight vectors in shared memory (up to 227 KB per block on sm_100a)
  cudaFuncSetAttribute(k_filter_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  h->filters_set[which] = !rc;
  return rc;
}

// kind: 0 full-lat full-lev, 1 half-lat full-lev, 2 full-lat half-lev, 3 2-D full-lat
// Host (halo'd Fortran storage of one field) <-> device (dense rows, no longitude halo): ONE strided 3-D copy of the
// interior, straight between the caller's array and the field -- no staging image, no pack kernel, and the halo
// columns / rows / ghost levels (a quarter of the bytes at 0.25 deg: 246 of 1686 columns, 4 of 36 levels) never cross
// PCIe.  Host halo cells are left untouched on download (the reference refills them with fill_halo before use).
// kind: 0 full-lat 3-D, 1 half-lat 3-D, 2 full-lat on half levels, 3 2-D.   GMCORE_XFER_STAGED=1: round-1 path.
static int xfer_staged(gmc_handle* h, double* dev, double* host, int kind, int dir);
static int xfer(gmc_handle* h, double* dev, double* host, int kind, int dir) {
  static const bool staged = getenv("GMCORE_XFER_STAGED") && getenv("GMCORE_XFER_STAGED")[0] == '1';
  if (staged) return xfer_staged(h, dev, host, kind, dir);
  const gmc_config& c = h->cfg; const Geo& g = h->g;
  const int nrows = kind == 1 ? g.hje - g.hjb + 1 : g.je - g.jb + 1;
  const int nx = c.nlon + 2 * c.hx, ny = nrows + 2 * c.hy;
  const int klev0 = (kind == 0 || kind == 1) ? 2 : 0, nk = kind == 3 ? 1 : (kind == 2 ? c.nlev + 1 : c.nlev);
  if (nrows <= 0) return 0;
  cudaMemcpy3DParms p = {};
  const cudaPitchedPtr hp = make_cudaPitchedPtr(host, (size_t)nx * sizeof(double), (size_t)nx, (size_t)ny);
  const cudaPitchedPtr dp = make_cudaPitchedPtr(dev, (size_t)g.nlon * sizeof(double), (size_t)g.nlon, (size_t)g.NY);
  const cudaPos hpos = make_cudaPos((size_t)c.hx * sizeof(double), (size_t)c.hy, (size_t)klev0);
  const cudaPos dpos = make_cudaPos(0, (size_t)(g.jb - g.jlo), 0);
  p.extent = make_cudaExtent((size_t)g.nlon * sizeof(double), (size_t)nrows, (size_t)nk);
  if (dir == 0) { p.srcPtr = hp; p.srcPos = hpos; p.dstPtr = dp; p.dstPos = dpos; p.kind = cudaMemcpyHostToDevice; }
  else { p.srcPtr = dp; p.srcPos = dpos; p.dstPtr = hp; p.dstPos = hpos; p.kind = cudaMemcpyDeviceToHost; }
  CUDA_CHECK(cudaMemcpy3DAsync(&p, h->stream));
  if (dir != 0) CUDA_CHECK(cudaStreamSynchronize(h->stream));
  return 0;
}
static int xfer_staged(gmc_handle* h, double* dev, double* host, int kind, int dir) {
  const gmc_config& c = h->cfg; const Geo& g = h->g;
  PackArgs a;
  a.dev = dev; a.img = h->stage_img; a.hx = c.hx; a.hy = c.hy; a.nx = c.nlon + 2 * c.hx;
  const int nrows_full = g.je - g.jb + 1, nrows_half = g.hje - g.hjb + 1;
  a.nrows = kind == 1 ? nrows_half : nrows_full;
  a.ny = a.nrows + 2 * c.hy;
  a.jfirst = g.jb;
  a.klev0 = (kind == 0 || kind == 1) ? 2 : 0;
  a.nk = kind == 3 ? 1 : (kind == 2 ? c.nlev + 1 : c.nlev);
  const int nz_img = kind == 3 ? 1 : (kind == 2 ? c.nlev + 1 : c.nlev + 4);
  const size_t bytes = (size_t)a.nx * a.ny * nz_img * sizeof(double);
  dim3 grid((g.nlon + 127) / 128, a.nrows, a.nk);
  if (dir == 0) {
    CUDA_CHECK(cudaMemcpyAsync(h->stage_img, host, bytes, cudaMemcpyHostToDevice, h->stream));
    LAUNCH(h, k_pack, grid, 128, 0, g, a, 0);
  } else {
    CUDA_CHECK(cudaMemsetAsync(h->stage_img, 0, bytes, h->stream));
    LAUNCH(h, k_pack, grid, 128, 0, g, a, 1);
    CUDA_CHECK(cudaMemcpyAsync(host, h->stage_img, bytes, cudaMemcpyDeviceToHost, h->stream));
    CUDA_CHECK(cudaStreamSynchronize(h->stream));
  }
  CUDA_CHECK(cudaGetLastError());
  return 0;
}

int gmc_set_static(gmc_handle* h, const double* gzs) { return xfer(h, h->gzs, const_cast<double*>(gzs), 3, 0); }

int gmc_upload_state(gmc_handle* h, int itime, const double* u, const double* v, const double* s, const double* phs) {
  if (itime < 1 || itime > 3) { snprintf(h->err, sizeof(h->err), "bad time index %d", itime); return 1; }
  DState& S = h->st[itime];
  if (xfer(h, S.u, const_cast<double*>(u), 0, 0)) return 1;
  if (xfer(h, S.v, const_cast<double*>(v), 1, 0)) return 1;
  if (h->cfg.baroclinic) {
    if (!phs) { snprintf(h->err, sizeof(h->err), "phs is required for the hydrostatic core"); return 1; }
    if (xfer(h, S.pt, const_cast<double*>(s), 0, 0)) return 1;
    if (xfer(h, S.phs, const_cast<double*>(phs), 3, 0)) return 1;
  } else {
    if (xfer(h, S.gz, const_cast<double*>(s), 0, 0)) return 1;
  }
  return 0;
}

int gmc_download_state(gmc_handle* h, int itime, double* u, double* v, double* s, double* phs) {
  if (itime < 1 || itime > 3) { snprintf(h->err, sizeof(h->err), "bad time index %d", itime); return 1; }
  DState& S = h->st[itime];
  if (u && xfer(h, S.u, u, 0, 1)) return 1;
  if (v && xfer(h, S.v, v, 1, 1)) return 1;
  if (h->cfg.baroclinic) {
    if (s && xfer(h, S.pt, s, 0, 1)) return 1;
    if (phs && xfer(h, S.phs, phs, 3, 1)) return 1;
  } else {
    if (s && xfer(h, S.gz, s, 0, 1)) return 1;
  }
  return 0;
}

int gmc_get_field(gmc_handle* h, int itime, const char* name, double* buf) {
  if (itime < 1 || itime > 3) { snprintf(h->err, sizeof(h->err), "bad time index %d", itime); return 1; }
  std::string n(name);
  double* p = nullptr; int kind = 0;
  auto it = h->st[itime].names.find(n);
  if (it != h->st[itime].names.end()) { p = it->second.first; kind = it->second.second; }
  else if (hydro_work_field(h, n, &p, &kind)) {}
  else if (n == "du") { p = h->du; kind = 0; }
  else if (n == "dv") { p = h->dv; kind = 1; }
  else if (n == "gzs") { p = h->gzs; kind = 3; }
  else if (h->cfg.baroclinic && (n == "ph" || n == "gz")) {
    // h0 variables that are never stored on the device (history_mod.F90:519-597): full-level pressure
    // (calc_ph, operators_mod.F90:157-164) and geopotential (calc_gz_lev tail, :459-466) are re-formed on demand
    FullLevArgs fa{h->st[itime].phs_f, h->st[itime].gz_lev, h->du, n == "ph" ? 0 : 1, make_vc(h)};
    LAUNCH(h, k_full_lev, gridg(h->g, h->g.nlev), 128, 0, h->g, h->M, fa);
    p = h->du; kind = 0;
  }
  if (!p) { snprintf(h->err, sizeof(h->err), "unknown field %s", name); return 1; }
  return xfer(h, p, buf, kind, 1);
}

// ---- latitude-band set-up (process_mod.F90:93-139 decompose_domains, 1-D in latitude) --------
// (a) single process, all bands on one device: link the handles (rank order) into a group;
//     gmc_prepare_run / gmc_time_integrate / gmc_run_steps / gmc_diagnose on hs[0] then drive all bands.
int gmc_group_link(gmc_handle** hs, int n) {
  if (n < 1 || !hs[0]) return 1;
  auto* grp = new std::vector<gmc_handle*>(hs, hs + n);
  for (int r = 0; r < n; ++r) {
    gmc_handle* b = hs[r];
    if (b->cfg.nranks != n || b->cfg.rank != r || b->cfg.device != hs[0]->cfg.device) {
      snprintf(hs[0]->err, sizeof(hs[0]->err), "gmc_group_link: handle %d is rank %d of %d on device %d", r, b->cfg.rank, b->cfg.nranks, b->cfg.device);
      delete grp; return 1; }
  }
  for (int r = 0; r < n; ++r) {
    gmc_handle* b = hs[r];
    if (r > 0) { if (b->owns_stream && b->stream) cudaStreamDestroy(b->stream); b->stream = hs[0]->stream; b->owns_stream = false; }
    b->group = grp;
  }
  // The production ordering (edge rows -> ev_edge -> exchange on a second stream next to the interior rows ->
  // ev_xdone) on ONE device too, so that the single-GPU band tests exercise the same row split and event
  // dependencies as the NCCL path (GMCORE_NO_OVERLAP=1 restores the single-stream order in both modes).
  { const char* e = getenv("GMCORE_NO_OVERLAP"); if (!(e && e[0] == '1')) for (int r = 0; r < n; ++r) {
    gmc_handle* b = hs[r];
    if (b->comm_stream) continue;
    cudaStreamCreateWithFlags(&b->comm_stream, cudaStreamNonBlocking);
    cudaEventCreateWithFlags(&b->ev_edge, cudaEventDisableTiming); cudaEventCreateWithFlags(&b->ev_xdone, cudaEventDisableTiming);
  } }
  return 0;
}
// (b) one process per GPU: rank 0 creates the id, the host broadcasts its 128 bytes (MPI_Bcast /
//     torch.distributed), every rank calls gmc_nccl_init.
int gmc_nccl_unique_id(char id[128]) {
  NcclApi& n = nccl_api();
  if (!n.ok) return 1;
  gmc_ncclUniqueId u; if (n.GetUniqueId(&u)) return 1;
  memcpy(id, u.internal, 128); return 0;
}
int gmc_nccl_init(gmc_handle* h, const char id[128]) {
  NcclApi& n = nccl_api();
  if (!n.ok) { snprintf(h->err, sizeof(h->err), "libnccl.so.2 not found in the process"); return 1; }
  gmc_ncclUniqueId u; memcpy(u.internal, id, 128);
  CUDA_CHECK(cudaSetDevice(h->cfg.device));
  int rc = n.CommInitRank(&h->nccl_comm, h->cfg.nranks, u, h->cfg.rank);
  { const char* e = getenv("GMCORE_NO_OVERLAP"); if (!rc && !(e && e[0] == '1')) {
    cudaStreamCreateWithFlags(&h->comm_stream, cudaStreamNonBlocking);
    cudaEventCreateWithFlags(&h->ev_edge, cudaEventDisableTiming); cudaEventCreateWithFlags(&h->ev_xdone, cudaEventDisableTiming);
  } }
  if (rc) { snprintf(h->err, sizeof(h->err), "ncclCommInitRank: %s", n.GetErrorString ? n.GetErrorString(rc) : "error"); h->nccl_comm = nullptr; return 1; }
  return 0;
}

int gmc_synchronize(gmc_handle* h) {
  CUDA_CHECK(cudaStreamSynchronize(h->stream));
  CUDA_CHECK(cudaGetLastError());
  return 0;
}
int gmc_pin_host(gmc_handle* h, void* ptr, size_t bytes) {
  if (!ptr || !bytes) { snprintf(h->err, sizeof(h->err), "gmc_pin_host: empty range"); return 1; }
  const cudaError_t e = cudaHostRegister(ptr, bytes, cudaHostRegisterPortable);
  if (e != cudaSuccess) { cudaGetLastError(); snprintf(h->err, sizeof(h->err), "gmc_pin_host: %s", cudaGetErrorString(e)); return 1; }
  return 0;
}
int gmc_unpin_host(gmc_handle* h, void* ptr) {
  const cudaError_t e = cudaHostUnregister(ptr);   // a failure here must not poison later cudaGetLastError checks
  if (e != cudaSuccess) { cudaGetLastError(); snprintf(h->err, sizeof(h->err), "gmc_unpin_host: %s", cudaGetErrorString(e)); return 1; }
  return 0;
}
int gmc_launch_count(gmc_handle* h, long long* launches) { *launches = h->launches; return 0; }
// Timing helpers (CUDA events on the handle's stream; torch.cuda.Event only sees
// torch's current stream).  slot 0 = start, 1 = stop.
int gmc_timer_record(gmc_handle* h, int slot) {
  if (!h->tev[slot]) CUDA_CHECK(cudaEventCreate(&h->tev[slot]));
  CUDA_CHECK(cudaEventRecord(h->tev[slot], h->stream));
  return 0;
}
int gmc_timer_elapsed_ms(gmc_handle* h, double* ms) {
  CUDA_CHECK(cudaEventSynchronize(h->tev[1]));
  float f = 0; CUDA_CHECK(cudaEventElapsedTime(&f, h->tev[0], h->tev[1]));
  *ms = f; return 0;
}
int gmc_profile(gmc_handle* h, int on) {
  h->prof.on = on != 0;
  if (on) { h->prof.entries.clear(); h->prof.used = 0; }
  return 0;
}
// Text report "name count total_ms\n" per kernel, sorted by total time.
int gmc_profile_report(gmc_handle* h, char* buf, int len) {
  CUDA_CHECK(cudaStreamSynchronize(h->stream));
  std::map<std::string, std::pair<long long, double>> agg;
  for (auto& e : h->prof.entries) {
    float ms = 0; cudaEventElapsedTime(&ms, e.e0, e.e1);
    std::string nm(e.name); nm = nm.substr(0, nm.find('<'));     // k_hflx<1> -> k_hflx
    auto& a = agg[nm]; a.first++; a.second += ms;
  }
  std::vector<std::pair<double, std::string>> order;
  for (auto& kv : agg) order.push_back({-kv.second.second, kv.first});
  std::sort(order.begin(), order.end());
  int off = 0;
  for (auto& o : order) {
    auto& a = agg[o.second];
    off += snprintf(buf + off, off < len ? len - off : 0, "%s %lld %.6f\n", o.second.c_str(), a.first, a.second);
    if (off >= len) break;
  }
  return 0;
}
void* gmc_stream(gmc_handle* h) { return (void*)h->stream; }
int gmc_struct_sizes(int* cfg, int* mv) { *cfg = (int)sizeof(gmc_config); *mv = (int)sizeof(gmc_mesh_vectors); return 0; }

}  // extern "C"

#include "api_run.cuh"

// ---- gmc_diagnose_static: `diagnose` without any allocation -----------------------------------------------------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  gmc_diagnose keeps the block sums
// of k_diag_partial and the eight results in buffers that gmc_create took from cudaMalloc; this variant keeps both
// in ONE file-scope device array of a fixed size, which every GPU that loads the library holds until exit:
//   g_static_scratch[0 .. 7]                      the eight sums (k_diag_final)
//   g_static_scratch[32 .. 32 + 7 * rows * nlev)  block sums of k_diag_partial, [7][rows * nlev]
// In host code `g_static_scratch` is only a stand-in for the symbol, not a device address: the address is taken once
// per device with cudaGetSymbolAddress (that device current) and kept in g_static_ptr[].
// Every call shares the array (it keeps the last call's contents): the used part is cleared on the handle's stream
// before the kernels, and calls that overlap on one GPU -- two handles on different streams, two host threads --
// race.  Single-band handles only (the band all-reduce of gmc_diagnose is not part of it).
#define GMC_STATIC_HEAD 32        // doubles in front of the block sums (keeps them 256-byte aligned)
static __device__ __align__(256) double g_static_scratch[GMC_STATIC_SCRATCH_DOUBLES];
#define GMC_STATIC_MAXDEV 64
static double* g_static_ptr[GMC_STATIC_MAXDEV];
static std::mutex g_static_mu;

static int static_scratch_address(gmc_handle* h, double** out) {
  const int dev = h->cfg.device;
  if (dev < 0 || dev >= GMC_STATIC_MAXDEV) { snprintf(h->err, sizeof(h->err), "gmc_diagnose_static: device %d outside 0..%d", dev, GMC_STATIC_MAXDEV - 1); return 1; }
  std::lock_guard<std::mutex> lk(g_static_mu);
  if (!g_static_ptr[dev]) {
    CUDA_CHECK(cudaSetDevice(dev));
    double* p = nullptr;
    CUDA_CHECK(cudaGetSymbolAddress((void**)&p, g_static_scratch));
    if (!p) { snprintf(h->err, sizeof(h->err), "gmc_diagnose_static: cudaGetSymbolAddress returned a null address"); return 1; }
    g_static_ptr[dev] = p;
  }
  *out = g_static_ptr[dev];
  return 0;
}

extern "C" int gmc_diagnose_static(gmc_handle* h, int itime, double out[8]) {
  if (check_ready(h)) return 1;
  if (itime < 0 || itime >= 4 || !out) { snprintf(h->err, sizeof(h->err), "gmc_diagnose_static: bad argument (itime %d)", itime); return 1; }
  if (h->cfg.nranks > 1 || h->group || h->nccl_comm) {
    snprintf(h->err, sizeof(h->err), "gmc_diagnose_static: single-band handles only (this one is rank %d of %d); use gmc_diagnose", h->cfg.rank, h->cfg.nranks);
    return 1;
  }
  const Geo& g = h->g;
  const int nrows = g.je - g.jb + 1;
  const long long need = GMC_STATIC_HEAD + 7LL * nrows * g.nlev;
  if (need > (long long)GMC_STATIC_SCRATCH_DOUBLES) {
    snprintf(h->err, sizeof(h->err), "gmc_diagnose_static: %d rows x %d levels need %lld doubles of scratch, the static array holds %lld (GMC_STATIC_SCRATCH_DOUBLES); use gmc_diagnose",
             nrows, g.nlev, need, (long long)GMC_STATIC_SCRATCH_DOUBLES);
    return 1;
  }
  double* sc = nullptr;
  if (static_scratch_address(h, &sc)) return 1;
  CUDA_CHECK(cudaSetDevice(h->cfg.device));
  CUDA_CHECK(cudaMemsetAsync(sc, 0, (size_t)need * sizeof(double), h->stream));
  DState& S = h->st[itime];
  DiagArgs a{S.m, S.mfx_lon, S.mfy_lat, S.u, S.v, S.pv, S.pv_lon, S.pv_lat, S.t, S.pt, S.phs, h->gzs,
             sc + GMC_STATIC_HEAD, h->cfg.g, h->cfg.cpd, h->cfg.baroclinic};
  // launched directly: the LAUNCH macro's profiler keeps its events in a growing host vector
  k_diag_partial<<<dim3(nrows, g.nlev), 256, 0, h->stream>>>(g, h->M, a);
  k_diag_final<<<8, 1024, 0, h->stream>>>(sc + GMC_STATIC_HEAD, nrows * g.nlev, sc, h->dflags);
  h->launches += 2;
  CUDA_CHECK(cudaGetLastError());
  double s[8];
  CUDA_CHECK(cudaMemcpyAsync(s, sc, sizeof(s), cudaMemcpyDeviceToHost, h->stream));
  CUDA_CHECK(cudaStreamSynchronize(h->stream));
  diag_to_abi(s, out);
  int guards = 0;
  { double v = s[7]; for (int b = 0; b < 5 && v > 0; ++b) { const double r = fmod(v, 256.0); if (r > 0) guards |= 1 << b; v = (v - r) / 256.0; } }
  if (guards) {      // as reduce_diag: report the reference's log_error guard and clear the device word
    hydro_flags_text(h, guards);
    if (h->dflags) cudaMemsetAsync(h->dflags, 0, sizeof(int), h->stream);
    return 1;
  }
  return 0;
}

// ---- gmc_fastmath_eval: the device results of pow_pos / log_pos for the bit-identity test ---------------------------
// out[i] = pow_pos(x[i], y) (what = 0) or log_pos(x[i]) (what = 1), evaluated on the current device.  Host arrays in and
// out; a verification hook (tests/test_fastmath_gpu.py compares with the host build of the same source), not part of
// the time-stepping path.
__global__ void k_fastmath_eval(int what, const double* x, double y, double* out, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = what == 0 ? pow_pos(x[i], y) : log_pos(x[i]);
}
extern "C" int gmc_fastmath_eval(int what, const double* x, double y, double* out, int n) {
  if (n <= 0 || !x || !out || what < 0 || what > 1) return 1;
  double *dx = nullptr, *dout = nullptr;
  int rc = 1;
  if (cudaMalloc(&dx, (size_t)n * 8) == cudaSuccess && cudaMalloc(&dout, (size_t)n * 8) == cudaSuccess &&
      cudaMemcpy(dx, x, (size_t)n * 8, cudaMemcpyHostToDevice) == cudaSuccess) {
    k_fastmath_eval<<<(n + 255) / 256, 256>>>(what, dx, y, dout, n);
    if (cudaMemcpy(out, dout, (size_t)n * 8, cudaMemcpyDeviceToHost) == cudaSuccess && cudaGetLastError() == cudaSuccess) rc = 0;
  }
  cudaFree(dx); cudaFree(dout);
  return rc;
}

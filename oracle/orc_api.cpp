// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_core.hpp).  C entry points for
// ctypes; same inputs as the product's C ABI (include/gmcore_gpu.h) so tests feed
// both sides identical mesh vectors, filter weights and states.
#include "orc_model.hpp"
#include "orc_forcing.hpp"
#include <map>
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace orc;

namespace {
A3* field_of(Model& m, int itime, const std::string& n) {
  State& s = m.b.state[itime]; Tend& t = m.b.tend[itime]; AdvBatch& B = m.b.adv_batch_mass;
  std::map<std::string, A3*> f = {
    {"u_lon", &s.u_lon}, {"v_lat", &s.v_lat}, {"u_f", &s.u_f}, {"v_f", &s.v_f}, {"gz", &s.gz}, {"gz_f", &s.gz_f},
    {"m", &s.m}, {"m_lon", &s.m_lon}, {"m_lat", &s.m_lat}, {"m_vtx", &s.m_vtx}, {"m_lev", &s.m_lev},
    {"mfx_lon", &s.mfx_lon}, {"mfy_lat", &s.mfy_lat}, {"mfx_lat", &s.mfx_lat}, {"mfy_lon", &s.mfy_lon},
    {"u_lat", &s.u_lat}, {"v_lon", &s.v_lon}, {"pv", &s.pv}, {"pv_lon", &s.pv_lon}, {"pv_lat", &s.pv_lat},
    {"ke", &s.ke}, {"vor", &s.vor}, {"div", &s.div}, {"pt", &s.pt}, {"t", &s.t}, {"ph", &s.ph},
    {"ph_lev", &s.ph_lev}, {"ph_exn_lev", &s.ph_exn_lev}, {"phs", &s.phs}, {"phs_f", &s.phs_f},
    {"gz_lev", &s.gz_lev}, {"we_lev", &s.we_lev}, {"we_lev_lon", &s.we_lev_lon}, {"we_lev_lat", &s.we_lev_lat},
    {"ptf_lon", &s.ptf_lon}, {"ptf_lat", &s.ptf_lat}, {"ptf_lev", &s.ptf_lev},
    {"du", &t.du}, {"dv", &t.dv}, {"dgz", &t.dgz}, {"dpt", &t.dpt}, {"dphs", &t.dphs}, {"qhu", &t.qhu}, {"qhv", &t.qhv},
    {"dkedlon", &t.dkedlon}, {"dkedlat", &t.dkedlat}, {"dmfdlon", &t.dmfdlon}, {"dmfdlat", &t.dmfdlat},
    {"pgf_lon", &t.pgf_lon}, {"pgf_lat", &t.pgf_lat}, {"wedudlev", &t.wedudlev}, {"wedvdlev", &t.wedvdlev},
    {"dptfdlon", &t.dptfdlon}, {"dptfdlat", &t.dptfdlat}, {"dptfdlev", &t.dptfdlev},
    {"cflx", &B.cflx}, {"cfly", &B.cfly}, {"cflz", &B.cflz}, {"divx", &B.divx}, {"divy", &B.divy},
    {"qx", &B.qx}, {"qy", &B.qy}, {"gzs", &m.b.st.gzs},
  };
  auto it = f.find(n);
  return it == f.end() ? nullptr : it->second;
}
void copy_interior_full3(const Mesh& mesh, A3& dst, const double* src, bool full_lat) {
  // src has the same halo'd layout as dst
  (void)mesh; (void)full_lat;
  std::memcpy(dst.d.data(), src, dst.d.size() * sizeof(double));
}
}  // namespace

extern "C" {

// adv_add_tracer + adv_allocate_tracers_1 (adv_mod.F90:314-418); returns the batch index
int orc_add_tracer_batch(void* h, int ntracers, double dt) {
  Model* m = (Model*)h; const Mesh& mesh = m->b.mesh;
  m->b.adv_batches.emplace_back();
  AdvBatch& B = m->b.adv_batches.back();
  B.init(mesh, dt, m->b.cfg.dt_dyn, false);
  B.alloc_tracers(mesh, ntracers);
  B.alert_period = dt; B.alert_last = -dt;
  return (int)m->b.adv_batches.size() - 1;
}
void* orc_create(const gmc_config* cfg, const gmc_mesh_vectors* mv, const double* full_lat_deg) {
  Model* m = new Model();
  m->b.cfg = *cfg;
  m->b.mesh.init(*cfg, *mv);
  const Mesh& mesh = m->b.mesh;
  for (int i = 1; i <= 3; ++i) { m->b.state[i].init(mesh, *cfg); m->b.tend[i].init(mesh); }
  mesh.alloc2(m->b.st.gzs, true);
  m->b.adv_batch_mass.init(mesh, cfg->dt_dyn, cfg->dt_dyn, true);
  if (cfg->ntracers > 0) orc_add_tracer_batch(m, cfg->ntracers, cfg->dt_adv);
  m->div_damp_init();
  m->lon_damp_init(full_lat_deg);
  return m;
}
void orc_destroy(void* h) { delete (Model*)h; }

int orc_set_filters(void* h, int which, const int* nl, const double* wl, int ldl, const int* na, const double* wa, int lda) {
  Model* m = (Model*)h;
  Filter& f = which == GMC_BIG_FILTER ? m->b.big_filter : which == GMC_SMALL_FILTER1 ? m->b.small_filter1 : m->b.small_filter2;
  f.set(m->b.mesh, nl, wl, ldl, na, wa, lda);
  return 0;
}
int orc_set_static(void* h, const double* gzs) {
  Model* m = (Model*)h;
  std::memcpy(m->b.st.gzs.d.data(), gzs, m->b.st.gzs.d.size() * sizeof(double));
  fill_halo(m->b.mesh, m->b.st.gzs, true);
  return 0;
}
// Host arrays are interior-valid; halos are regenerated here the way the
// reference's set_ic routines do (fill_halo after each field).
int orc_upload_state(void* h, int itime, const double* u, const double* v, const double* pt_or_gz, const double* phs) {
  Model* m = (Model*)h; State& s = m->b.state[itime]; const Mesh& mesh = m->b.mesh;
  std::memcpy(s.u_lon.d.data(), u, s.u_lon.d.size() * sizeof(double)); fill_halo(mesh, s.u_lon, true);
  std::memcpy(s.v_lat.d.data(), v, s.v_lat.d.size() * sizeof(double)); fill_halo(mesh, s.v_lat, false);
  if (m->b.cfg.baroclinic) {
    std::memcpy(s.pt.d.data(), pt_or_gz, s.pt.d.size() * sizeof(double)); fill_halo(mesh, s.pt, true);
    std::memcpy(s.phs.d.data(), phs, s.phs.d.size() * sizeof(double)); fill_halo(mesh, s.phs, true);
  } else {
    std::memcpy(s.gz.d.data(), pt_or_gz, s.gz.d.size() * sizeof(double)); fill_halo(mesh, s.gz, true);
  }
  return 0;
}
int orc_set_batch_tracers(void* h, int batch, const double* q) {
  Model* m = (Model*)h; AdvBatch& B = m->b.adv_batches.at(batch);
  for (int l = 0; l < B.ntracers; ++l) {
    A3& a = B.q[B.old - 1][l];
    std::memcpy(a.d.data(), q + (size_t)l * a.d.size(), a.d.size() * sizeof(double));
    fill_halo(m->b.mesh, a, true);
  }
  return 0;
}
// Advection-only driver (gmcore_adv_driver.F90:84-88): set_uv(state(new)) then adv_run(new), time_advance.
// orc_adv_prescribe = the set_uv of tests/adv/*_test_mod.F90: m = 1, u_f/v_f given, mfx = u, mfy = v.
int orc_adv_prescribe(void* h, int itime, const double* u_f, const double* v_f) {
  Model* m = (Model*)h; State& s = m->b.state[itime]; const Mesh& mesh = m->b.mesh;
  std::fill(s.m.d.begin(), s.m.d.end(), 1.0);
  std::memcpy(s.u_f.d.data(), u_f, s.u_f.d.size() * sizeof(double)); fill_halo(mesh, s.u_f, true);
  s.mfx_lon.d = s.u_f.d;
  std::memcpy(s.v_f.d.data(), v_f, s.v_f.d.size() * sizeof(double)); fill_halo(mesh, s.v_f, false);
  s.mfy_lat.d = s.v_f.d;
  if (m->b.cfg.baroclinic) { std::fill(s.we_lev.d.begin(), s.we_lev.d.end(), 0.0); std::fill(s.m_lev.d.begin(), s.m_lev.d.end(), 1.0); }
  return 0;
}
// dcmip12_test_set_uv (tests/adv/dcmip12_test_mod.F90:146-158): prescribed we_lev
int orc_adv_prescribe_we(void* h, int itime, const double* we_lev) {
  Model* m = (Model*)h; State& s = m->b.state[itime];
  std::memcpy(s.we_lev.d.data(), we_lev, s.we_lev.d.size() * sizeof(double));
  return 0;
}
// interp_{cell,lon_edge,lat_edge,lev_edge}_to_pressure_level (interp_mod.F90:727-897) as diag_state_run calls them
// (utils/diag_state_mod.F90:241-258); z is bracketed with ph_lev (see include/gmcore_gpu.h)
int orc_interp_pressure_level(void* h, int itime, const char* var, double po, int logp, double fill, double* y2d) {
  Model* m = (Model*)h; State& s = m->b.state[itime]; const Mesh& mesh = m->b.mesh;
  const std::string v(var);
  A3 y; mesh.alloc2(y, true);
  const double p0 = logp ? std::log(po) : po;
  const bool half = v == "z";
  const A3& p = half ? s.ph_lev : s.ph;
  for (int j = 1; j <= mesh.nlat; ++j) for (int i = 1; i <= mesh.nlon; ++i) {
    y(i, j) = fill;
    for (int k = half ? mesh.nlev + 1 : mesh.nlev; k >= 2; --k)
      if (p(i, j, k - 1) <= po && po <= p(i, j, k)) {
        const double dp1 = logp ? p0 - std::log(p(i, j, k - 1)) : p0 - p(i, j, k - 1);
        const double dp2 = logp ? std::log(p(i, j, k)) - p0 : p(i, j, k) - p0;
        double x1, x2;
        if (v == "u") { x1 = 0.5 * (s.u_lon(i - 1, j, k - 1) + s.u_lon(i, j, k - 1)); x2 = 0.5 * (s.u_lon(i - 1, j, k) + s.u_lon(i, j, k)); }
        else if (v == "v") { x1 = 0.5 * (s.v_lat(i, j - 1, k - 1) + s.v_lat(i, j, k - 1)); x2 = 0.5 * (s.v_lat(i, j - 1, k) + s.v_lat(i, j, k)); }
        else if (v == "pt") { x1 = s.pt(i, j, k - 1); x2 = s.pt(i, j, k); }
        else if (v == "t") { x1 = s.t(i, j, k - 1); x2 = s.t(i, j, k); }
        else { x1 = s.gz_lev(i, j, k - 1) * (1.0 / m->b.cfg.g); x2 = s.gz_lev(i, j, k) * (1.0 / m->b.cfg.g); }
        y(i, j) = (dp2 * x1 + dp1 * x2) / (dp1 + dp2);
        break;
      }
  }
  std::memcpy(y2d, y.d.data(), y.d.size() * sizeof(double));
  return 0;
}
int orc_adv_step(void* h) {
  Model* m = (Model*)h;
  m->adv_run(m->new_idx);
  m->time_advance();
  return m->b.error.empty() ? 0 : 1;
}
int orc_get_batch_tracers(void* h, int batch, double* q) {
  Model* m = (Model*)h; AdvBatch& B = m->b.adv_batches.at(batch);
  for (int l = 0; l < B.ntracers; ++l) {
    A3& a = B.q[B.old - 1][l];
    std::memcpy(q + (size_t)l * a.d.size(), a.d.data(), a.d.size() * sizeof(double));
  }
  return 0;
}
int orc_set_tracers(void* h, const double* q) { return orc_set_batch_tracers(h, 0, q); }
int orc_get_tracers(void* h, double* q) { return orc_get_batch_tracers(h, 0, q); }
int orc_prepare_run(void* h) { Model* m = (Model*)h; m->prepare_run(m->old_idx); return m->b.error.empty() ? 0 : 1; }

// n x { time_integrate; time_advance; diagnose(old) }  (gmcore_mod.F90:138-144)
int orc_run_steps(void* h, int n, double* diag) {
  Model* m = (Model*)h;
  for (int s = 0; s < n; ++s) {
    m->time_integrate(m->old_idx, m->new_idx, m->b.cfg.dt_dyn);
    m->time_advance();
    double d[8];
    m->diagnose(m->old_idx, d);
    if (diag) std::memcpy(diag + (size_t)s * 8, d, sizeof(d));
    if (!m->b.error.empty()) return 1;
  }
  return 0;
}
int orc_diagnose(void* h, int itime, double* out) { ((Model*)h)->diagnose(itime, out); return 0; }
int orc_old_index(void* h) { return ((Model*)h)->old_idx; }
int orc_new_index(void* h) { return ((Model*)h)->new_idx; }
int orc_time_step(void* h) { return ((Model*)h)->time_step; }

// Copies the whole halo'd array; returns its element count (or -1).
long long orc_get_field(void* h, int itime, const char* name, double* buf) {
  Model* m = (Model*)h;
  A3* a = field_of(*m, itime, name);
  if (!a || !a->allocated()) return -1;
  if (buf) std::memcpy(buf, a->d.data(), a->d.size() * sizeof(double));
  return (long long)a->d.size();
}
int orc_field_shape(void* h, int itime, const char* name, int* shape /*ni,nj,nk,ilb,jlb,klb*/) {
  Model* m = (Model*)h;
  A3* a = field_of(*m, itime, name);
  if (!a || !a->allocated()) return -1;
  shape[0] = a->ni; shape[1] = a->nj; shape[2] = a->nk; shape[3] = a->ilb; shape[4] = a->jlb; shape[5] = a->klb;
  return 0;
}
// Run one named operator on state(itime) for operator-level parity tests.
int orc_num_threads() {
#if defined(_OPENMP) && defined(ORC_OPENMP_LOOPS)
  return omp_get_max_threads();
#else
  return 1;   // the restatement's loops are scalar (no OpenMP pragmas yet)
#endif
}
int orc_last_error(void* h, char* buf, int len) {
  Model* m = (Model*)h;
  std::snprintf(buf, len, "%s", m->b.error.c_str());
  return m->b.error.empty() ? 0 : 1;
}

}  // extern "C"

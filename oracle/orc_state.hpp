// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_core.hpp).
// Containers: state_type (src/dynamics/state_mod.F90:105-155), tend_type
// (src/dynamics/tend_mod.F90:67-107), static_type, filter_type
// (src/dynamics/filter_types_mod.F90:14-27), adv_batch_type
// (src/dynamics/adv/adv_batch_mod.F90:19-69), and the zonal filters
// (src/dynamics/filter_mod.F90:25-163).
#pragma once
#include "orc_core.hpp"

namespace orc {

struct State {
  A3 u_lon, v_lon, u_f, u_lat, v_lat, v_f, we_lev, we_lev_lon, we_lev_lat, gz, gz_lev, m, m_vtx, m_lon, m_lat,
     m_lev, mfx_lon, mfy_lon, mfy_lat, mfx_lat, pv, pv_lon, pv_lat, ke, pt, ptf_lon, ptf_lat, ptf_lev, t, ph,
     ph_lev, ph_exn_lev, phs, phs_f, div, vor, gz_f, div2, smag_t, smag_s, kmh_lon, kmh_lat;
  double tm = 0, te = 0, tav = 0, tpe = 0, te_ke = 0, te_ie = 0, te_pe = 0, tpt = 0;
  void init(const Mesh& m_, const gmc_config& c) {
    const bool F = true, H = false;
    m_.alloc(u_lon, F); m_.alloc(v_lon, F); m_.alloc(u_f, F); m_.alloc(u_lat, H); m_.alloc(v_lat, H);
    m_.alloc(v_f, H); m_.alloc(we_lev, F, false); m_.alloc(we_lev_lon, F, false); m_.alloc(we_lev_lat, H, false);
    m_.alloc(gz, F); m_.alloc(gz_lev, F, false); m_.alloc(m, F); m_.alloc(m_vtx, H); m_.alloc(m_lon, F);
    m_.alloc(m_lat, H); m_.alloc(m_lev, F, false); m_.alloc(mfx_lon, F); m_.alloc(mfy_lon, F);
    m_.alloc(mfy_lat, H); m_.alloc(mfx_lat, H); m_.alloc(pv, H); m_.alloc(pv_lon, F); m_.alloc(pv_lat, H);
    m_.alloc(ke, F); m_.alloc(pt, F); m_.alloc(ptf_lon, F); m_.alloc(ptf_lat, H); m_.alloc(ptf_lev, F, false);
    m_.alloc(t, F); m_.alloc(ph, F); m_.alloc(ph_lev, F, false); m_.alloc(ph_exn_lev, F, false);
    m_.alloc2(phs, F); m_.alloc2(phs_f, F); m_.alloc(div, F); m_.alloc(vor, H);
    if (!c.baroclinic) m_.alloc(gz_f, F);
    if (c.div_damp_order == 4) m_.alloc(div2, F);
    if (c.use_smag_damp) { m_.alloc(smag_t, F); m_.alloc(smag_s, H); m_.alloc(kmh_lon, F); m_.alloc(kmh_lat, H); }
  }
};

struct Tend {
  A3 du, dv, dgz, dpt, dphs, qhv, qhu, dkedlon, dkedlat, dmfdlon, dmfdlat, dptfdlon, dptfdlat, dptfdlev,
     pgf_lon, pgf_lat, wedudlev, wedvdlev;
  bool update_u = false, update_v = false, update_gz = false, update_pt = false, update_phs = false;
  bool copy_gz = false, copy_pt = false, copy_phs = false;
  void init(const Mesh& m_) {
    const bool F = true, H = false;
    m_.alloc(du, F); m_.alloc(dv, H); m_.alloc(dgz, F); m_.alloc(dpt, F); m_.alloc2(dphs, F);
    m_.alloc(qhv, F); m_.alloc(qhu, H); m_.alloc(dkedlon, F); m_.alloc(dkedlat, H); m_.alloc(dmfdlon, F);
    m_.alloc(dmfdlat, F); m_.alloc(dptfdlon, F); m_.alloc(dptfdlat, F); m_.alloc(dptfdlev, F);
    m_.alloc(pgf_lon, F); m_.alloc(pgf_lat, H); m_.alloc(wedudlev, F); m_.alloc(wedvdlev, H);
  }
  void reset_flags() {  // tend_mod.F90:109-124
    update_u = update_v = update_gz = update_pt = update_phs = false;
    copy_gz = copy_pt = copy_phs = false;
  }
};

struct Static { A3 gzs; };

struct Filter {
  std::vector<int> ngrid_lon, ngrid_lat; int lon_lb = 0, lat_lb = 0, ld_lon = 1, ld_lat = 1;
  std::vector<double> wgt_lon, wgt_lat;
  int nlon_at(int j) const { return ngrid_lon[j - lon_lb]; }
  int nlat_at(int j) const { return ngrid_lat[j - lat_lb]; }
  const double* wlon(int j) const { return &wgt_lon[(size_t)(j - lon_lb) * ld_lon]; }
  const double* wlat(int j) const { return &wgt_lat[(size_t)(j - lat_lb) * ld_lat]; }
  void set(const Mesh& m, const int* nl, const double* wl, int ldl, const int* na, const double* wa, int lda) {
    lon_lb = m.full_lat_lb; lat_lb = m.half_lat_lb; ld_lon = ldl; ld_lat = lda;
    int nf = m.full_lat_ub - m.full_lat_lb + 1, nh = m.half_lat_ub - m.half_lat_lb + 1;
    ngrid_lon.assign(nl, nl + nf); ngrid_lat.assign(na, na + nh);
    wgt_lon.assign(wl, wl + (size_t)ldl * nf); wgt_lat.assign(wa, wa + (size_t)lda * nh);
  }
};

// filter_on_cell_3d / filter_on_lon_edge (same code; filter_mod.F90:57-127):
// rows full_lat_ibeg..iend, weights wgt_lon.  y==nullptr -> in place.
// T3: the whole tmp row incl. (stale) halo entries is copied back.
inline void filter_full_lat(const Mesh& m, const Filter& f, A3& x, A3* y, int k0, int k1) {
  std::vector<double> tmp(x.ni, 0.0);
  ORC_PAR_FP(tmp) for (int k = k0; k <= k1; ++k)
    for (int j = m.full_lat_ibeg; j <= m.full_lat_iend; ++j) {
      if (f.nlon_at(j) > 1) {
        int n = f.nlon_at(j), hn = (n - 1) / 2; const double* w = f.wlon(j);
        for (int i = 1; i <= m.nlon; ++i) {
          double s = 0;
          for (int t = 0; t < n; ++t) s += w[t] * x(i - hn + t, j, k);
          tmp[i - x.ilb] = s;
        }
        A3& dst = y ? *y : x;
        for (int i = x.ilb; i < x.ilb + x.ni; ++i) dst(i, j, k) = tmp[i - x.ilb];
      } else if (y) {
        for (int i = x.ilb; i < x.ilb + x.ni; ++i) (*y)(i, j, k) = x(i, j, k);
      }
    }
}
// filter_on_lat_edge (filter_mod.F90:129-163)
inline void filter_half_lat(const Mesh& m, const Filter& f, A3& x, A3* y, int k0, int k1) {
  std::vector<double> tmp(x.ni, 0.0);
  ORC_PAR_FP(tmp) for (int k = k0; k <= k1; ++k)
    for (int j = m.half_lat_ibeg; j <= m.half_lat_iend; ++j) {
      if (f.nlat_at(j) > 1) {
        int n = f.nlat_at(j), hn = (n - 1) / 2; const double* w = f.wlat(j);
        for (int i = 1; i <= m.nlon; ++i) {
          double s = 0;
          for (int t = 0; t < n; ++t) s += w[t] * x(i - hn + t, j, k);
          tmp[i - x.ilb] = s;
        }
        A3& dst = y ? *y : x;
        for (int i = x.ilb; i < x.ilb + x.ni; ++i) dst(i, j, k) = tmp[i - x.ilb];
      } else if (y) {
        for (int i = x.ilb; i < x.ilb + x.ni; ++i) (*y)(i, j, k) = x(i, j, k);
      }
    }
}

// adv_batch_type, loc='cell' (adv_batch_mod.F90:73-149)
struct AdvBatch {
  double alert_last = 0, alert_period = 0; bool alert_ring = false;   // time_mod.F90:253-291 alert named after the batch
  bool dynamic = false; int nstep = 0, uv_step = 0, we_step = 0, mf_step = 0, old = 1, nw = 2; double dt = 0;
  int ntracers = 0;
  A3 old_m, mfx, mfy, m_lev, u, v, we, cflx, cfly, cflz, divx, divy, qx, qy, qlx, qly, dqx, dqy, q6x, q6y;
  std::vector<A3> q[2];   // q(:,:,:,l,old/new)
  A3 qmf_lon, qmf_lat, qmf_lev;
  void init(const Mesh& m_, double dt_, double dt_dyn, bool dynamic_) {
    const bool F = true, H = false;
    dt = dt_; dynamic = dynamic_; nstep = (int)(dt_ / dt_dyn); uv_step = we_step = mf_step = 0;
    m_.alloc(old_m, F); m_.alloc(mfx, F); m_.alloc(mfy, H); m_.alloc(m_lev, F, false); m_.alloc(u, F);
    m_.alloc(v, H); m_.alloc(we, F, false); m_.alloc(cflx, F); m_.alloc(cfly, H); m_.alloc(cflz, F, false);
    m_.alloc(divx, F); m_.alloc(divy, F); m_.alloc(qx, F); m_.alloc(qy, F); m_.alloc(qlx, F); m_.alloc(qly, F);
    m_.alloc(dqx, F); m_.alloc(dqy, F); m_.alloc(q6x, F); m_.alloc(q6y, F);
  }
  void alloc_tracers(const Mesh& m_, int n) {
    ntracers = n;
    for (int s = 0; s < 2; ++s) { q[s].resize(n); for (auto& a : q[s]) m_.alloc(a, true); }
    m_.alloc(qmf_lon, true); m_.alloc(qmf_lat, false); m_.alloc(qmf_lev, true, false);
  }
};

// block_type (src/utils/block_mod.F90:27-45) for the one global block.
struct Block {
  gmc_config cfg;
  Mesh mesh;
  State state[4];   // 1..3 used (Fortran indices)
  Tend tend[4];
  Static st;
  Filter big_filter, small_filter1, small_filter2;
  AdvBatch adv_batch_mass;
  std::vector<AdvBatch> adv_batches;   // block%adv_batches(:) (adv_mod.F90:314-418): non-dynamic 'cell' batches, each with its own samples
  std::string error;           // first log_error raised (reference aborts)
  void fail(const std::string& s) {
#if defined(_OPENMP) && defined(ORC_OPENMP_LOOPS)
#pragma omp critical(orc_fail)
#endif
    if (error.empty()) error = s;
  }
};

}  // namespace orc

// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_core.hpp).
// C-grid operators, same loop order and summation order as
//   src/dynamics/operators_mod.F90 (line ranges cited per function)
//   src/dynamics/interp_mod.F90:233-275,361-369,426-463
//   src/dynamics/pgf/pgf_swm_mod.F90:18-46, pgf_lin97_mod.F90:20-105
//   src/dynamics/adv/upwind_mod.F90:14-39
#pragma once
#include "orc_adv.hpp"

namespace orc {

inline double upwind1(double dir, double wgt, double f1, double f2) {
  return 0.5 * (f2 + f1) + (-0.5) * (f2 - f1) * wgt * dir;
}
inline double upwind3(double dir, double wgt, double f1, double f2, double f3, double f4) {
  const double c31 = 7.0 / 12.0, c32 = -1.0 / 12.0, c33 = 1.0 / 12.0;
  return c31 * (f3 + f2) + c32 * (f4 + f1) + c33 * (f4 - f1 - 3 * (f3 - f2)) * wgt * dir;
}

// hybrid_coord_mod.F90:154-171 (local_ptop = ptop for the wrf_* / equal_eta templates)
inline double calc_ph_lev_k(const Block& b, int k, double phs) {
  const double ptop = b.cfg.ptop, p0 = b.cfg.hybrid_p0;
  return b.mesh.hyai(k) * (p0 - ptop) + b.mesh.hybi(k) * (phs - ptop) + ptop;
}
inline double calc_dphdt_lev_k(const Block& b, int k, double dphsdt) { return b.mesh.hybi(k) * dphsdt; }

// operators_mod.F90:136-168
inline void calc_ph(Block& b, State& s) {
  const Mesh& mesh = b.mesh;
  ORC_PAR for (int k = mesh.half_lev_ibeg; k <= mesh.half_lev_iend; ++k)
    for (int j = 1; j <= mesh.nlat; ++j) for (int i = 1; i <= mesh.nlon; ++i) {
      s.ph_lev(i, j, k) = calc_ph_lev_k(b, k, s.phs_f(i, j));
      s.ph_exn_lev(i, j, k) = std::pow(s.ph_lev(i, j, k), b.cfg.Rd_o_cpd);
    }
  fill_halo(mesh, s.ph_exn_lev, true, false, true, false, true);
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = 1; j <= mesh.nlat; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      s.ph(i, j, k) = 0.5 * (s.ph_lev(i, j, k) + s.ph_lev(i, j, k + 1));
  fill_halo(mesh, s.ph, true);
}

// formula_mod.F90:15-31: temperature = pt * (p / p0)**Rd_o_cpd
inline double temperature(const Block& b, double pt, double p) { return pt * std::pow(p / b.cfg.p0, b.cfg.Rd_o_cpd); }

// operators_mod.F90:170-190
inline void calc_t(Block& b, State& s) {
  const Mesh& mesh = b.mesh;
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = 1; j <= mesh.nlat; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      s.t(i, j, k) = temperature(b, s.pt(i, j, k), s.ph(i, j, k));
}

// operators_mod.F90:471-543 (+ interp_mod.F90:233-275,438-463)
inline void calc_m(Block& b, State& s) {
  const Mesh& mesh = b.mesh;
  if (b.cfg.baroclinic) {
    ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
      for (int j = 1; j <= mesh.nlat; ++j) for (int i = 1; i <= mesh.nlon; ++i) {
        s.m(i, j, k) = s.ph_lev(i, j, k + 1) - s.ph_lev(i, j, k);
        if (s.m(i, j, k) <= 0) b.fail("The pressure levels are not monotonic!");
      }
    ORC_PAR for (int k = mesh.half_lev_ibeg + 1; k <= mesh.half_lev_iend - 1; ++k)
      for (int j = 1; j <= mesh.nlat; ++j) for (int i = 1; i <= mesh.nlon; ++i)
        s.m_lev(i, j, k) = s.ph(i, j, k) - s.ph(i, j, k - 1);
    int k = mesh.half_lev_ibeg;
    for (int j = 1; j <= mesh.nlat; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      s.m_lev(i, j, k) = s.ph(i, j, k) - s.ph_lev(i, j, k);
    k = mesh.half_lev_iend;
    for (int j = 1; j <= mesh.nlat; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      s.m_lev(i, j, k) = s.ph_lev(i, j, k) - s.ph(i, j, k - 1);
    fill_halo(mesh, s.m_lev, true);
  } else {
    for (int j = 1; j <= mesh.nlat; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      s.m(i, j, 1) = (s.gz(i, j, 1) - b.st.gzs(i, j)) / b.cfg.g;
  }
  fill_halo(mesh, s.m, true);
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = mesh.full_lat_ibeg_no_pole; j <= mesh.full_lat_iend_no_pole; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      s.m_lon(i, j, k) = (s.m(i, j, k) + s.m(i + 1, j, k)) * 0.5;
  fill_halo(mesh, s.m_lon, true);
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = 1; j <= mesh.nhalf; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      s.m_lat(i, j, k) = (s.m(i, j, k) + s.m(i, j + 1, k)) * 0.5;
  fill_halo(mesh, s.m_lat, false);
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = 1; j <= mesh.nhalf; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      s.m_vtx(i, j, k) = ((s.m(i, j, k) + s.m(i + 1, j, k)) * mesh.area_subcell(2, j) +
                          (s.m(i, j + 1, k) + s.m(i + 1, j + 1, k)) * mesh.area_subcell(1, j + 1)) / mesh.area_vtx(j);
}

// operators_mod.F90:545-610
inline void calc_mf(Block& b, State& s, double dt) {
  const Mesh& mesh = b.mesh;
  accum_uv_cell(b, b.adv_batch_mass, s.u_f, s.v_f, dt, true);
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = mesh.full_lat_ibeg_no_pole; j <= mesh.full_lat_iend_no_pole; ++j) for (int i = 0; i <= mesh.nlon; ++i)
      s.mfx_lon(i, j, k) = s.m_lon(i, j, k) * s.u_f(i, j, k);
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = 1; j <= mesh.nhalf; ++j) for (int i = 1; i <= mesh.nlon + 1; ++i)
      s.mfy_lat(i, j, k) = s.m_lat(i, j, k) * s.v_f(i, j, k);
  accum_mf_cell(b.adv_batch_mass, s.mfx_lon, s.mfy_lat);
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = 1; j <= mesh.nhalf; ++j) for (int i = 1; i <= mesh.nlon; ++i) {
      s.mfx_lat(i, j, k) = mesh.half_tangent_wgt(1, j) * (s.mfx_lon(i - 1, j, k) + s.mfx_lon(i, j, k)) +
                           mesh.half_tangent_wgt(2, j) * (s.mfx_lon(i - 1, j + 1, k) + s.mfx_lon(i, j + 1, k));
      s.u_lat(i, j, k) = s.mfx_lat(i, j, k) / s.m_lat(i, j, k);
    }
  fill_halo(mesh, s.u_lat, false);
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = mesh.full_lat_ibeg_no_pole; j <= mesh.full_lat_iend_no_pole; ++j) for (int i = 1; i <= mesh.nlon; ++i) {
      s.mfy_lon(i, j, k) = mesh.full_tangent_wgt(1, j) * (s.mfy_lat(i, j - 1, k) + s.mfy_lat(i + 1, j - 1, k)) +
                           mesh.full_tangent_wgt(2, j) * (s.mfy_lat(i, j, k) + s.mfy_lat(i + 1, j, k));
      s.v_lon(i, j, k) = s.mfy_lon(i, j, k) / s.m_lon(i, j, k);
    }
  fill_halo(mesh, s.v_lon, true);
}

// operators_mod.F90:234-352
inline void calc_ke(Block& b, State& s) {
  const Mesh& mesh = b.mesh; const A3& u = s.u_f; const A3& v = s.v_f; A3& ke = s.ke;
  auto sq = [](double x) { return x * x; };
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = mesh.full_lat_ibeg_no_pole; j <= mesh.full_lat_iend_no_pole; ++j) for (int i = 1; i <= mesh.nlon + 1; ++i)
      ke(i, j, k) = (mesh.area_lon_west(j) * sq(u(i - 1, j, k)) + mesh.area_lon_east(j) * sq(u(i, j, k)) +
                     mesh.area_lat_north(j - 1) * sq(v(i, j - 1, k)) + mesh.area_lat_south(j) * sq(v(i, j, k))) / mesh.area_cell(j);
  if (b.cfg.ke_scheme == 2) {
    const double w = b.cfg.ke_cell_wgt;
    ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
      for (int j = mesh.full_lat_ibeg_no_pole; j <= mesh.full_lat_iend_no_pole; ++j) for (int i = 1; i <= mesh.nlon + 1; ++i) {
        double kv1 = (mesh.area_lat_east(j) * sq(v(i - 1, j, k)) + mesh.area_lat_west(j) * sq(v(i, j, k)) +
                      mesh.area_lon_north(j) * sq(u(i - 1, j, k)) + mesh.area_lon_south(j + 1) * sq(u(i - 1, j + 1, k))) / mesh.area_vtx(j);
        double kv2 = (mesh.area_lat_east(j - 1) * sq(v(i - 1, j - 1, k)) + mesh.area_lat_west(j - 1) * sq(v(i, j - 1, k)) +
                      mesh.area_lon_north(j - 1) * sq(u(i - 1, j - 1, k)) + mesh.area_lon_south(j) * sq(u(i - 1, j, k))) / mesh.area_vtx(j - 1);
        double kv3 = (mesh.area_lat_east(j - 1) * sq(v(i, j - 1, k)) + mesh.area_lat_west(j - 1) * sq(v(i + 1, j - 1, k)) +
                      mesh.area_lon_north(j - 1) * sq(u(i, j - 1, k)) + mesh.area_lon_south(j) * sq(u(i, j, k))) / mesh.area_vtx(j - 1);
        double kv4 = (mesh.area_lat_east(j) * sq(v(i, j, k)) + mesh.area_lat_west(j) * sq(v(i + 1, j, k)) +
                      mesh.area_lon_north(j) * sq(u(i, j, k)) + mesh.area_lon_south(j + 1) * sq(u(i, j + 1, k))) / mesh.area_vtx(j);
        ke(i, j, k) = (1.0 - w) * ((kv1 + kv4) * mesh.area_subcell(2, j) + (kv2 + kv3) * mesh.area_subcell(1, j)) / mesh.area_cell(j) +
                      w * ke(i, j, k);
      }
  }
  std::vector<double> work(mesh.nlon);
  for (int pole_i = 0; pole_i < 2; ++pole_i) {
    int j = pole_i == 0 ? mesh.full_lat_ibeg : mesh.full_lat_iend;
    int jv = pole_i == 0 ? j : j - 1;
    for (int k = 1; k <= mesh.nlev; ++k) {
      for (int i = 1; i <= mesh.nlon; ++i) work[i - 1] = sq(v(i, jv, k));
      double pole = zonal_sum(work) / mesh.nlon;
      for (int i = 1; i <= mesh.nlon; ++i) ke(i, j, k) = pole;
    }
  }
}

// operators_mod.F90:354-431
inline void calc_div(Block& b, State& s) {
  const Mesh& mesh = b.mesh; const A3& u = s.u_f; const A3& v = s.v_f; A3& div = s.div;
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = mesh.full_lat_ibeg_no_pole; j <= mesh.full_lat_iend_no_pole; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      div(i, j, k) = ((u(i, j, k) * mesh.le_lon(j) - u(i - 1, j, k) * mesh.le_lon(j)) +
                      (v(i, j, k) * mesh.le_lat(j) - v(i, j - 1, k) * mesh.le_lat(j - 1))) / mesh.area_cell(j);
  std::vector<double> work(mesh.nlon);
  {
    int j = mesh.full_lat_ibeg;
    for (int k = 1; k <= mesh.nlev; ++k) {
      for (int i = 1; i <= mesh.nlon; ++i) work[i - 1] = v(i, j, k);
      double pole = zonal_sum(work) * mesh.le_lat(j) / mesh.nlon / mesh.area_cell(j);
      for (int i = 1; i <= mesh.nlon; ++i) div(i, j, k) = pole;
    }
  }
  {
    int j = mesh.full_lat_iend;
    for (int k = 1; k <= mesh.nlev; ++k) {
      for (int i = 1; i <= mesh.nlon; ++i) work[i - 1] = -v(i, j - 1, k);
      double pole = zonal_sum(work) * mesh.le_lat(j - 1) / mesh.nlon / mesh.area_cell(j);
      for (int i = 1; i <= mesh.nlon; ++i) div(i, j, k) = pole;
    }
  }
  if (b.cfg.div_damp_order == 4) fill_halo(mesh, div, true);
  else fill_halo(mesh, div, true, false, true, false, true);
  if (b.cfg.div_damp_order == 4) {
    A3& div2 = s.div2;
    ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
      for (int j = mesh.full_lat_ibeg_no_pole; j <= mesh.full_lat_iend_no_pole; ++j) for (int i = 1; i <= mesh.nlon; ++i)
        div2(i, j, k) = (div(i + 1, j, k) - 2 * div(i, j, k) + div(i - 1, j, k)) / (mesh.de_lon(j) * mesh.de_lon(j)) +
                        ((div(i, j + 1, k) - div(i, j, k)) * mesh.half_cos_lat(j) / mesh.de_lat(j) -
                         (div(i, j, k) - div(i, j - 1, k)) * mesh.half_cos_lat(j - 1) / mesh.de_lat(j - 1)) /
                            mesh.le_lon(j) / mesh.full_cos_lat(j);
    fill_halo(mesh, div2, true, false, true, false, true);
  }
}

// operators_mod.F90:612-677
inline void calc_vor(Block& b, State& s, const A3& u, const A3& v) {
  const Mesh& mesh = b.mesh; A3& vor = s.vor;
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = 1; j <= mesh.nhalf; ++j) for (int i = 0; i <= mesh.nlon; ++i)
      vor(i, j, k) = (u(i, j, k) * mesh.de_lon(j) - u(i, j + 1, k) * mesh.de_lon(j + 1) +
                      v(i + 1, j, k) * mesh.de_lat(j) - v(i, j, k) * mesh.de_lat(j)) / mesh.area_vtx(j);
  if (b.cfg.pv_pole_stokes) {
    std::vector<double> work(mesh.nlon);
    {
      int j = mesh.half_lat_ibeg;
      for (int k = 1; k <= mesh.nlev; ++k) {
        for (int i = 1; i <= mesh.nlon; ++i) work[i - 1] = -s.u_lat(i, j, k) * mesh.le_lat(j);
        double pole = zonal_sum(work) / mesh.nlon / mesh.area_cell(j);
        for (int i = 1; i <= mesh.nlon; ++i) vor(i, j, k) = pole;
      }
    }
    {
      int j = mesh.half_lat_iend;
      for (int k = 1; k <= mesh.nlev; ++k) {
        for (int i = 1; i <= mesh.nlon; ++i) work[i - 1] = s.u_lat(i, j, k) * mesh.le_lat(j);
        double pole = zonal_sum(work) / mesh.nlon / mesh.area_cell(j + 1);
        for (int i = 1; i <= mesh.nlon; ++i) vor(i, j, k) = pole;
      }
    }
  }
}

// operators_mod.F90:679-703
inline void calc_pv(Block& b, State& s) {
  const Mesh& mesh = b.mesh;
  calc_vor(b, s, s.u_f, s.v_f);
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = 1; j <= mesh.nhalf; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      s.pv(i, j, k) = (s.vor(i, j, k) + mesh.half_f(j)) / s.m_vtx(i, j, k);
  fill_halo(mesh, s.pv, false);
}


// operators_mod.F90:705-736
inline void interp_pv_midpoint(Block& b, State& s) {
  const Mesh& mesh = b.mesh;
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = 1; j <= mesh.nhalf; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      s.pv_lat(i, j, k) = 0.5 * (s.pv(i - 1, j, k) + s.pv(i, j, k));
  fill_halo(mesh, s.pv_lat, false, true, true, true, false);
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = mesh.full_lat_ibeg_no_pole; j <= mesh.full_lat_iend_no_pole; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      s.pv_lon(i, j, k) = 0.5 * (s.pv(i, j, k) + s.pv(i, j - 1, k));
  fill_halo(mesh, s.pv_lon, true, true, true, false, true);
}

// operators_mod.F90:738-795
inline void interp_pv_upwind(Block& b, State& s) {
  const Mesh& mesh = b.mesh;
  const A3 &un = s.u_f, &vn = s.v_f, &ut = s.u_lat, &vt = s.v_lon, &pv = s.pv;
  const double eps = 2.220446049250313e-16, wgt = b.cfg.upwind_wgt_pv;
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k) {
    for (int j = mesh.full_lat_ibeg_no_pole; j <= mesh.full_lat_iend_no_pole; ++j) for (int i = 1; i <= mesh.nlon; ++i) {
      double bb = std::fabs(vt(i, j, k)) / (std::sqrt(un(i, j, k) * un(i, j, k) + vt(i, j, k) * vt(i, j, k)) + eps);
      double up = b.cfg.upwind_order_pv == 1
        ? upwind1(sign1(vt(i, j, k)), wgt, pv(i, j - 1, k), pv(i, j, k))
        : upwind3(sign1(vt(i, j, k)), wgt, pv(i, j - 2, k), pv(i, j - 1, k), pv(i, j, k), pv(i, j + 1, k));
      s.pv_lon(i, j, k) = bb * up + (1 - bb) * 0.5 * (pv(i, j - 1, k) + pv(i, j, k));
    }
    for (int j = 1; j <= mesh.nhalf; ++j) for (int i = 1; i <= mesh.nlon; ++i) {
      double bb = std::fabs(ut(i, j, k)) / (std::sqrt(ut(i, j, k) * ut(i, j, k) + vn(i, j, k) * vn(i, j, k)) + eps);
      double up = b.cfg.upwind_order_pv == 1
        ? upwind1(sign1(ut(i, j, k)), wgt, pv(i - 1, j, k), pv(i, j, k))
        : upwind3(sign1(ut(i, j, k)), wgt, pv(i - 2, j, k), pv(i - 1, j, k), pv(i, j, k), pv(i + 1, j, k));
      s.pv_lat(i, j, k) = bb * up + (1 - bb) * 0.5 * (pv(i - 1, j, k) + pv(i, j, k));
    }
  }
  fill_halo(mesh, s.pv_lon, true, true, true, false, true);
  fill_halo(mesh, s.pv_lat, false, true, true, true, false);
}

// adv/tvd_mod.F90:42-85.  eps is the default-real literal 1.0e-6 widened to r8 (tvd_mod.F90:50).
inline double tvd_flux_limiter(int type, double r) {
  if (type == GMC_TVD_NONE) return 1.0;
  if (type == GMC_TVD_MC) return std::fmax(0.0, std::fmin(std::fmin(2.0 * r, 0.5 * (1 + r)), 2.0));
  return (r + std::fabs(r)) / (1 + std::fabs(r));
}
inline double tvd(int type, double cfl, double fm1, double f, double fp1, double fp2) {
  const double eps = (double)1.0e-6f;
  double r = cfl > 0 ? (f - fm1 + eps) / (fp1 - f + eps) : (fp2 - fp1 + eps) / (fp1 - f + eps);
  double C = tvd_flux_limiter(type, r);
  return 0.5 * (fp1 + f - ((1 - C) * sign1(cfl) + cfl * C) * (fp1 - f));
}

// operators_mod.F90:797-832
inline void interp_pv_tvd(Block& b, State& s, double dt) {
  const Mesh& mesh = b.mesh;
  const A3 &ut = s.u_lat, &vt = s.v_lon, &pv = s.pv;
  const int type = b.cfg.tvd_limiter_type;
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k) {
    for (int j = mesh.full_lat_ibeg_no_pole; j <= mesh.full_lat_iend_no_pole; ++j) for (int i = 1; i <= mesh.nlon; ++i) {
      double cfl = vt(i, j, k) * dt / mesh.le_lon(j);
      s.pv_lon(i, j, k) = tvd(type, cfl, pv(i, j - 2, k), pv(i, j - 1, k), pv(i, j, k), pv(i, j + 1, k));
    }
    for (int j = 1; j <= mesh.nhalf; ++j) for (int i = 1; i <= mesh.nlon; ++i) {
      double cfl = ut(i, j, k) * dt / mesh.le_lat(j);
      s.pv_lat(i, j, k) = tvd(type, cfl, pv(i - 2, j, k), pv(i - 1, j, k), pv(i, j, k), pv(i + 1, j, k));
    }
  }
  fill_halo(mesh, s.pv_lon, true, true, true, false, true);
  fill_halo(mesh, s.pv_lat, false, true, true, true, false);
}

// procedure pointer bound by operators_init (operators_mod.F90:56-70)
inline void interp_pv(Block& b, State& s, double dt) {
  if (b.cfg.pv_scheme == GMC_PV_MIDPOINT) interp_pv_midpoint(b, s);
  else if (b.cfg.pv_scheme == GMC_PV_TVD) interp_pv_tvd(b, s, dt);
  else interp_pv_upwind(b, s);
}

// operators_mod.F90:846-906
inline void calc_coriolis(Block& b, State& s, Tend& t) {
  const Mesh& mesh = b.mesh;
  const A3 &mfx_lon = s.mfx_lon, &mfy_lat = s.mfy_lat, &pv_lon = s.pv_lon, &pv_lat = s.pv_lat;
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = 1; j <= mesh.nhalf; ++j) for (int i = 1; i <= mesh.nlon; ++i) {
      if (b.cfg.coriolis_scheme == 1)
        t.qhu(i, j, k) = (mesh.half_tangent_wgt(1, j) * (mfx_lon(i - 1, j, k) * (pv_lat(i, j, k) + pv_lon(i - 1, j, k)) +
                                                          mfx_lon(i, j, k) * (pv_lat(i, j, k) + pv_lon(i, j, k))) +
                          mesh.half_tangent_wgt(2, j) * (mfx_lon(i - 1, j + 1, k) * (pv_lat(i, j, k) + pv_lon(i - 1, j + 1, k)) +
                                                          mfx_lon(i, j + 1, k) * (pv_lat(i, j, k) + pv_lon(i, j + 1, k)))) * 0.5;
      else
        t.qhu(i, j, k) = s.mfx_lat(i, j, k) * pv_lat(i, j, k);
    }
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = mesh.full_lat_ibeg_no_pole; j <= mesh.full_lat_iend_no_pole; ++j) for (int i = 1; i <= mesh.nlon; ++i) {
      if (b.cfg.coriolis_scheme == 1)
        t.qhv(i, j, k) = (mesh.full_tangent_wgt(1, j) * (mfy_lat(i, j - 1, k) * (pv_lon(i, j, k) + pv_lat(i, j - 1, k)) +
                                                          mfy_lat(i + 1, j - 1, k) * (pv_lon(i, j, k) + pv_lat(i + 1, j - 1, k))) +
                          mesh.full_tangent_wgt(2, j) * (mfy_lat(i, j, k) * (pv_lon(i, j, k) + pv_lat(i, j, k)) +
                                                          mfy_lat(i + 1, j, k) * (pv_lon(i, j, k) + pv_lat(i + 1, j, k)))) * 0.5;
      else
        t.qhv(i, j, k) = s.mfy_lon(i, j, k) * pv_lon(i, j, k);
    }
}

// operators_mod.F90:908-937
inline void calc_grad_ke(Block& b, State& s, Tend& t) {
  const Mesh& mesh = b.mesh;
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = mesh.full_lat_ibeg_no_pole; j <= mesh.full_lat_iend_no_pole; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      t.dkedlon(i, j, k) = (s.ke(i + 1, j, k) - s.ke(i, j, k)) / mesh.de_lon(j);
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = 1; j <= mesh.nhalf; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      t.dkedlat(i, j, k) = (s.ke(i, j + 1, k) - s.ke(i, j, k)) / mesh.de_lat(j);
}

// Shared pole treatment of a flux divergence (operators_mod.F90:974-1003 and the
// identical blocks at :1053-1082): d(:,pole,k) = +-sum_i f(i,j*,k) * le_lat / N / area_cell
inline void pole_flux_div(const Mesh& mesh, const A3& f_lat, A3& d) {
  std::vector<double> work(mesh.nlon);
  {
    int j = mesh.full_lat_ibeg;
    for (int k = 1; k <= mesh.nlev; ++k) {
      for (int i = 1; i <= mesh.nlon; ++i) work[i - 1] = f_lat(i, j, k);
      double pole = zonal_sum(work) * mesh.le_lat(j) / mesh.nlon / mesh.area_cell(j);
      for (int i = 1; i <= mesh.nlon; ++i) d(i, j, k) = pole;
    }
  }
  {
    int j = mesh.full_lat_iend;
    for (int k = 1; k <= mesh.nlev; ++k) {
      for (int i = 1; i <= mesh.nlon; ++i) work[i - 1] = -f_lat(i, j - 1, k);
      double pole = zonal_sum(work) * mesh.le_lat(j - 1) / mesh.nlon / mesh.area_cell(j);
      for (int i = 1; i <= mesh.nlon; ++i) d(i, j, k) = pole;
    }
  }
}

// operators_mod.F90:939-1006
inline void calc_grad_mf(Block& b, State& s, Tend& t) {
  const Mesh& mesh = b.mesh;
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = mesh.full_lat_ibeg_no_pole; j <= mesh.full_lat_iend_no_pole; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      t.dmfdlon(i, j, k) = (s.mfx_lon(i, j, k) - s.mfx_lon(i - 1, j, k)) * mesh.le_lon(j) / mesh.area_cell(j);
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = mesh.full_lat_ibeg_no_pole; j <= mesh.full_lat_iend_no_pole; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      t.dmfdlat(i, j, k) = (s.mfy_lat(i, j, k) * mesh.le_lat(j) - s.mfy_lat(i, j - 1, k) * mesh.le_lat(j - 1)) / mesh.area_cell(j);
  pole_flux_div(mesh, s.mfy_lat, t.dmfdlat);
}

// operators_mod.F90:1120-1143
inline void calc_dphsdt(Block& b, Tend& t) {
  const Mesh& mesh = b.mesh;
  std::fill(t.dphs.d.begin(), t.dphs.d.end(), 0.0);
  ORC_PAR for (int j = 1; j <= mesh.nlat; ++j)       // per-point sums over k in the reference's order (k outermost there)
    for (int k = 1; k <= mesh.nlev; ++k) for (int i = 1; i <= mesh.nlon; ++i)
      t.dphs(i, j) = t.dphs(i, j) - t.dmfdlon(i, j, k) - t.dmfdlat(i, j, k);
}

// operators_mod.F90:192-232 (+ interp_mod.F90:361-369,426-434)
inline void calc_we_lev(Block& b, State& s, Tend& t, double dt) {
  const Mesh& mesh = b.mesh;
  ORC_PAR for (int k = mesh.half_lev_ibeg + 1; k <= mesh.half_lev_iend - 1; ++k)
    for (int j = 1; j <= mesh.nlat; ++j) for (int i = 1; i <= mesh.nlon; ++i) {
      double mf = 0.0;
      for (int l = 1; l <= k - 1; ++l) mf = mf + t.dmfdlon(i, j, l) + t.dmfdlat(i, j, l);
      s.we_lev(i, j, k) = -calc_dphdt_lev_k(b, k, t.dphs(i, j)) - mf;
    }
  for (int j = s.we_lev.jlb; j < s.we_lev.jlb + s.we_lev.nj; ++j) for (int i = s.we_lev.ilb; i < s.we_lev.ilb + s.we_lev.ni; ++i) {
    s.we_lev(i, j, mesh.half_lev_ibeg) = 0.0; s.we_lev(i, j, mesh.half_lev_iend) = 0.0;
  }
  fill_halo(mesh, s.we_lev, true);
  accum_we_lev(b, b.adv_batch_mass, s.we_lev, s.m_lev, dt, true);
  ORC_PAR for (int k = mesh.half_lev_ibeg; k <= mesh.half_lev_iend; ++k)
    for (int j = mesh.full_lat_ibeg_no_pole; j <= mesh.full_lat_iend_no_pole; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      s.we_lev_lon(i, j, k) = (mesh.area_lon_west(j) * s.we_lev(i, j, k) + mesh.area_lon_east(j) * s.we_lev(i + 1, j, k)) / mesh.area_lon(j);
  ORC_PAR for (int k = mesh.half_lev_ibeg; k <= mesh.half_lev_iend; ++k)
    for (int j = 1; j <= mesh.nhalf; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      s.we_lev_lat(i, j, k) = (mesh.area_lat_north(j) * s.we_lev(i, j + 1, k) + mesh.area_lat_south(j) * s.we_lev(i, j, k)) / mesh.area_lat(j);
}

// operators_mod.F90:1145-1216
inline void calc_wedudlev_wedvdlev(Block& b, State& s, Tend& t) {
  const Mesh& mesh = b.mesh; const A3 &u = s.u_f, &v = s.v_f;
  const int kb = mesh.full_lev_ibeg, ke = mesh.full_lev_iend;
  const int jb = mesh.full_lat_ibeg_no_pole, je = mesh.full_lat_iend_no_pole;
  ORC_PAR for (int k = kb + 1; k <= ke - 1; ++k)
    for (int j = jb; j <= je; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      t.wedudlev(i, j, k) = (s.we_lev_lon(i, j, k + 1) * (u(i, j, k + 1) - u(i, j, k)) +
                             s.we_lev_lon(i, j, k) * (u(i, j, k) - u(i, j, k - 1))) / s.m_lon(i, j, k) / 2.0;
  for (int j = jb; j <= je; ++j) for (int i = 1; i <= mesh.nlon; ++i)
    t.wedudlev(i, j, kb) = (s.we_lev_lon(i, j, kb + 1) * (u(i, j, kb + 1) - u(i, j, kb))) / s.m_lon(i, j, kb) / 2.0;
  for (int j = jb; j <= je; ++j) for (int i = 1; i <= mesh.nlon; ++i)
    t.wedudlev(i, j, ke) = (s.we_lev_lon(i, j, ke) * (u(i, j, ke) - u(i, j, ke - 1))) / s.m_lon(i, j, ke) / 2.0;
  ORC_PAR for (int k = kb + 1; k <= ke - 1; ++k)
    for (int j = 1; j <= mesh.nhalf; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      t.wedvdlev(i, j, k) = (s.we_lev_lat(i, j, k + 1) * (v(i, j, k + 1) - v(i, j, k)) +
                             s.we_lev_lat(i, j, k) * (v(i, j, k) - v(i, j, k - 1))) / s.m_lat(i, j, k) / 2.0;
  for (int j = 1; j <= mesh.nhalf; ++j) for (int i = 1; i <= mesh.nlon; ++i)
    t.wedvdlev(i, j, kb) = (s.we_lev_lat(i, j, kb + 1) * (v(i, j, kb + 1) - v(i, j, kb))) / s.m_lat(i, j, kb) / 2.0;
  for (int j = 1; j <= mesh.nhalf; ++j) for (int i = 1; i <= mesh.nlon; ++i)
    t.wedvdlev(i, j, ke) = (s.we_lev_lat(i, j, ke) * (v(i, j, ke) - v(i, j, ke - 1))) / s.m_lat(i, j, ke) / 2.0;
}

// operators_mod.F90:433-469
inline void calc_gz_lev(Block& b, State& s) {
  const Mesh& mesh = b.mesh;
  ORC_PAR for (int k = mesh.half_lev_ibeg; k <= mesh.half_lev_iend - 1; ++k)
    for (int j = 1; j <= mesh.nlat; ++j) for (int i = 1; i <= mesh.nlon; ++i) {
      double dgz = 0.0;
      for (int l = k; l <= mesh.nlev; ++l)
        dgz = dgz + b.cfg.Rd * s.t(i, j, l) * std::log(s.ph_lev(i, j, l + 1) / s.ph_lev(i, j, l));
      s.gz_lev(i, j, k) = b.st.gzs(i, j) + dgz;
    }
  fill_halo(mesh, s.gz_lev, true);
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = 1; j <= mesh.nlat; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      s.gz(i, j, k) = 0.5 * (s.gz_lev(i, j, k) + s.gz_lev(i, j, k + 1));
}

// pgf_swm_mod.F90:18-46 / pgf_lin97_mod.F90:20-105 (hydrostatic branch)
inline void pgf_run(Block& b, State& s, Tend& t) {
  const Mesh& mesh = b.mesh;
  const int jb = mesh.full_lat_ibeg_no_pole, je = mesh.full_lat_iend_no_pole;
  if (!b.cfg.baroclinic) {
    const A3& gz = s.gz_f;
    ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
      for (int j = jb; j <= je; ++j) for (int i = 1; i <= mesh.nlon; ++i)
        t.pgf_lon(i, j, k) = (gz(i + 1, j, k) - gz(i, j, k)) / mesh.de_lon(j);
    ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
      for (int j = 1; j <= mesh.nhalf; ++j) for (int i = 1; i <= mesh.nlon; ++i)
        t.pgf_lat(i, j, k) = (gz(i, j + 1, k) - gz(i, j, k)) / mesh.de_lat(j);
    return;
  }
  const A3 &pe = s.ph_exn_lev, &gz = s.gz_lev;
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k) {
    for (int j = jb; j <= je; ++j) for (int i = 1; i <= mesh.nlon; ++i) {
      double dph1 = pe(i + 1, j, k + 1) - pe(i, j, k);
      double dph2 = pe(i, j, k + 1) - pe(i + 1, j, k);
      double dgz1 = gz(i, j, k + 1) - gz(i + 1, j, k);
      double dgz2 = gz(i, j, k) - gz(i + 1, j, k + 1);
      t.pgf_lon(i, j, k) = -(dph1 * dgz1 + dph2 * dgz2) / mesh.de_lon(j) / (dph1 + dph2);
    }
    for (int j = 1; j <= mesh.nhalf; ++j) for (int i = 1; i <= mesh.nlon; ++i) {
      double dph1 = pe(i, j + 1, k + 1) - pe(i, j, k);
      double dph2 = pe(i, j, k + 1) - pe(i, j + 1, k);
      double dgz1 = gz(i, j, k + 1) - gz(i, j + 1, k);
      double dgz2 = gz(i, j, k) - gz(i, j + 1, k + 1);
      t.pgf_lat(i, j, k) = -(dph1 * dgz1 + dph2 * dgz2) / mesh.de_lat(j) / (dph1 + dph2);
    }
  }
}

}  // namespace orc

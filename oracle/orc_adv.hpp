// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_core.hpp).
// Advection batches and FFSL, same loop/summation order as
//   src/dynamics/adv/adv_batch_mod.F90:207-632 (copy_old_m, accum_uv_cell RK4 CFL,
//                                               accum_mf_cell, accum_we_lev)
//   src/dynamics/adv/ffsl_mod.F90:318-432,778-941,1147-1194 (tracer h/v flux, PPM)
#pragma once
#include "orc_state.hpp"

namespace orc {

inline double sign1(double x) { return std::signbit(x) ? -1.0 : 1.0; }  // sign(1.0_r8, x)
inline double fsign(double a, double b) { return std::signbit(b) ? -std::fabs(a) : std::fabs(a); }
inline double min3(double a, double b, double c) { return std::fmin(std::fmin(a, b), c); }
inline double max3(double a, double b, double c) { return std::fmax(std::fmax(a, b), c); }

inline void copy_arr(A3& dst, const A3& src) {
#if defined(_OPENMP) && defined(ORC_OPENMP_LOOPS)
  if (dst.d.size() == src.d.size()) {
    const long long N = (long long)src.d.size(); double* d = dst.d.data(); const double* s = src.d.data();
    ORC_PAR for (long long n = 0; n < N; ++n) d[n] = s[n];
    return;
  }
#endif
  dst.d = src.d;
}
inline void axpby_arr(A3& x, const A3& y, bool average, int nstep) {
  // this%u = this%u + u   |  this%u = (this%u + u) / (nstep + 1)
  const long long N = (long long)x.d.size();
  ORC_PAR for (long long n = 0; n < N; ++n) {
    double s = x.d[n] + y.d[n];
    x.d[n] = average ? s / (nstep + 1) : s;
  }
}

// adv_batch_mod.F90:218-377
inline void accum_uv_cell(Block& b, AdvBatch& B, const A3& u, const A3& v, double dt, bool has_dt) {
  const Mesh& mesh = b.mesh;
  const double radius = b.cfg.radius;
  double dt_ = has_dt ? dt : B.dt;
  if (B.uv_step == 0) {
    copy_arr(B.u, u); copy_arr(B.v, v);
    if (!B.dynamic) B.uv_step = B.uv_step + 1;
  } else if (B.uv_step == B.nstep) {
    axpby_arr(B.u, u, true, B.nstep); axpby_arr(B.v, v, true, B.nstep);
    B.uv_step = 0;
  } else {
    axpby_arr(B.u, u, false, B.nstep); axpby_arr(B.v, v, false, B.nstep);
    B.uv_step = B.uv_step + 1;
  }
  if (B.uv_step != 0) return;
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k) {
    for (int j = mesh.full_lat_ibeg_no_pole; j <= mesh.full_lat_iend_no_pole; ++j) {
      double dx = mesh.de_lon(j);
      for (int i = 1; i <= mesh.nlon; ++i) {
        double x0 = radius * mesh.half_lon(i) * mesh.full_cos_lat(j);
        double u1 = B.u(i, j, k);
        double x1 = x0 + dt_ / 2 * u1;
        int l = (int)std::floor(i + (x1 - x0) / dx);
        double u2 = ((radius * mesh.half_lon(l + 1) * mesh.full_cos_lat(j) - x1) * B.u(l, j, k) +
                     (x1 - radius * mesh.half_lon(l) * mesh.full_cos_lat(j)) * B.u(l + 1, j, k)) / dx;
        double x2 = x0 + dt_ / 2 * u2;
        l = (int)std::floor(i + (x2 - x0) / dx);
        double u3 = ((radius * mesh.half_lon(l + 1) * mesh.full_cos_lat(j) - x2) * B.u(l, j, k) +
                     (x2 - radius * mesh.half_lon(l) * mesh.full_cos_lat(j)) * B.u(l + 1, j, k)) / dx;
        double x3 = x0 + dt_ * u3;
        l = (int)std::floor(i + (x3 - x0) / dx);
        double u4 = ((radius * mesh.half_lon(l + 1) * mesh.full_cos_lat(j) - x3) * B.u(l, j, k) +
                     (x3 - radius * mesh.half_lon(l) * mesh.full_cos_lat(j)) * B.u(l + 1, j, k)) / dx;
        B.cflx(i, j, k) = (u1 + 2 * u2 + 2 * u3 + u4) / 6 * dt_ / dx;
        if (std::fabs(B.cflx(i, j, k)) > mesh.hx) b.fail("cflx exceeds mesh%lon_halo_width");
      }
    }
    for (int j = 1; j <= mesh.nhalf; ++j) {
      double dy = mesh.de_lat(j);
      for (int i = 1; i <= mesh.nlon; ++i) {
        double y0 = radius * mesh.half_lat(j);
        double v1 = B.v(i, j, k);
        double y1 = y0 + dt_ / 2 * v1;
        int l = (int)std::floor(j + (y1 - y0) / dy);
        double v2, v3, v4;
        if (l < 1 || l > mesh.nhalf - 1) v2 = v1;
        else v2 = ((radius * mesh.half_lat(l + 1) - y1) * B.v(i, l, k) + (y1 - radius * mesh.half_lat(l)) * B.v(i, l + 1, k)) / dy;
        double y2 = y0 + dt_ / 2 * v2;
        l = (int)std::floor(j + (y2 - y0) / dy);
        if (l < 1 || l > mesh.nhalf - 1) v3 = v1;
        else v3 = ((radius * mesh.half_lat(l + 1) - y2) * B.v(i, l, k) + (y2 - radius * mesh.half_lat(l)) * B.v(i, l + 1, k)) / dy;
        double y3 = y0 + dt_ * v3;
        l = (int)std::floor(j + (y3 - y0) / dy);
        if (l < 1 || l > mesh.nhalf - 1) v4 = v1;
        else v4 = ((radius * mesh.half_lat(l + 1) - y3) * B.v(i, l, k) + (y3 - radius * mesh.half_lat(l)) * B.v(i, l + 1, k)) / dy;
        B.cfly(i, j, k) = (v1 + 2 * v2 + 2 * v3 + v4) / 6 * dt_ / dy;
        if (std::fabs(B.cfly(i, j, k)) > 1) b.fail("cfly exceeds 1");
      }
    }
    for (int j = mesh.full_lat_ibeg_no_pole; j <= mesh.full_lat_iend_no_pole; ++j)
      for (int i = 1; i <= mesh.nlon; ++i) {
        B.divx(i, j, k) = (B.u(i, j, k) - B.u(i - 1, j, k)) * mesh.le_lon(j) / mesh.area_cell(j);
        B.divy(i, j, k) = (B.v(i, j, k) * mesh.le_lat(j) - B.v(i, j - 1, k) * mesh.le_lat(j - 1)) / mesh.area_cell(j);
      }
  }
  std::vector<double> work(mesh.nlon);
  {
    int j = mesh.full_lat_ibeg;
    for (int k = 1; k <= mesh.nlev; ++k) {
      for (int i = 1; i <= mesh.nlon; ++i) work[i - 1] = B.v(i, j, k);
      double pole = zonal_sum(work);
      pole = pole * mesh.le_lat(j) / mesh.nlon / mesh.area_cell(j);
      for (int i = 1; i <= mesh.nlon; ++i) B.divy(i, j, k) = pole;
    }
  }
  {
    int j = mesh.full_lat_iend;
    for (int k = 1; k <= mesh.nlev; ++k) {
      for (int i = 1; i <= mesh.nlon; ++i) work[i - 1] = -B.v(i, j - 1, k);
      double pole = zonal_sum(work);
      pole = pole * mesh.le_lat(j - 1) / mesh.nlon / mesh.area_cell(j);
      for (int i = 1; i <= mesh.nlon; ++i) B.divy(i, j, k) = pole;
    }
  }
}

// adv_batch_mod.F90:379-403
inline void accum_mf_cell(AdvBatch& B, const A3& mfx, const A3& mfy) {
  if (B.mf_step == 0) {
    copy_arr(B.mfx, mfx); copy_arr(B.mfy, mfy);
    if (!B.dynamic) B.mf_step = B.mf_step + 1;
  } else if (B.mf_step == B.nstep) {
    axpby_arr(B.mfx, mfx, true, B.nstep); axpby_arr(B.mfy, mfy, true, B.nstep);
    B.mf_step = 0;
  } else {
    axpby_arr(B.mfx, mfx, false, B.nstep); axpby_arr(B.mfy, mfy, false, B.nstep);
    B.mf_step = B.mf_step + 1;
  }
}

// adv_batch_mod.F90:493-632
inline void accum_we_lev(Block& b, AdvBatch& B, const A3& we, const A3& m_lev, double dt, bool has_dt) {
  const Mesh& mesh = b.mesh;
  double dt_ = has_dt ? dt : B.dt;
  if (B.we_step == 0) {
    copy_arr(B.we, we); copy_arr(B.m_lev, m_lev);
    if (!B.dynamic) B.we_step = B.we_step + 1;
  } else if (B.we_step == B.nstep) {
    axpby_arr(B.we, we, true, B.nstep); axpby_arr(B.m_lev, m_lev, true, B.nstep);
    B.we_step = 0;
  } else {
    axpby_arr(B.we, we, false, B.nstep); axpby_arr(B.m_lev, m_lev, false, B.nstep);
    B.we_step = B.we_step + 1;
  }
  if (B.we_step != 0) return;
  const int kfb = mesh.full_lev_ibeg, kfe = mesh.full_lev_iend;
  auto search = [&](double w, double z, int k, int i, int j) -> int {
    int l;
    if (w > 0) {
      for (l = k; l <= kfe; ++l) if (mesh.half_lev(l) <= z && z <= mesh.half_lev(l + 1)) break;
      if (l == kfe + 1) { b.fail("cflz exceeds range"); l = kfe; }
    } else {
      for (l = k - 1; l >= kfb; --l) if (mesh.half_lev(l) <= z && z <= mesh.half_lev(l + 1)) break;
      if (l == kfb - 1) { b.fail("cflz exceeds range"); l = kfb; }
    }
    (void)i; (void)j;
    return l;
  };
  auto interp = [&](int i, int j, int l, double z) -> double {
    return ((mesh.half_lev(l + 1) - z) * B.we(i, j, l) / B.m_lev(i, j, l) * mesh.half_dlev(l) +
            (z - mesh.half_lev(l)) * B.we(i, j, l + 1) / B.m_lev(i, j, l + 1) * mesh.half_dlev(l + 1)) / mesh.full_dlev(l);
  };
  ORC_PAR for (int k = mesh.half_lev_ibeg + 1; k <= mesh.half_lev_iend - 1; ++k)
    for (int j = mesh.full_lat_ibeg; j <= mesh.full_lat_iend; ++j)
      for (int i = 1; i <= mesh.nlon; ++i) {
        double z0 = mesh.half_lev(k);
        double w1 = B.we(i, j, k) / B.m_lev(i, j, k) * mesh.half_dlev(k);
        double z1 = z0 + dt_ / 2 * w1;
        int l = search(w1, z1, k, i, j);
        double w2 = interp(i, j, l, z1);
        double z2 = z0 + dt_ / 2 * w2;
        l = search(w2, z2, k, i, j);
        double w3 = interp(i, j, l, z2);
        double z3 = z0 + dt_ * w3;
        l = search(w3, z3, k, i, j);
        double w4 = interp(i, j, l, z3);
        double deta = (w1 + 2 * w2 + 2 * w3 + w4) / 6 * dt_;
        int ks, ke, s;
        if (deta < 0) { ks = k - 1; ke = kfb; s = -1; } else { ks = k; ke = kfe; s = 1; }
        deta = std::fabs(deta);
        for (l = ks; s > 0 ? l <= ke : l >= ke; l += s) {
          deta = deta - mesh.full_dlev(l);
          if (deta < 0) {
            B.cflz(i, j, k) = s * ((1 + deta / mesh.full_dlev(l)) + std::abs(l - ks));
            break;
          }
        }
      }
}


// ffsl_mod.F90:1183-1194 (limiter_type='mono')
inline double slope_mono(double fm1, double f, double fp1) {
  double df = (fp1 - fm1) * 0.5;
  double df_min = 2 * (f - min3(fm1, f, fp1));
  double df_max = 2 * (max3(fm1, f, fp1) - f);
  return fsign(min3(std::fabs(df), df_min, df_max), df);
}

// ffsl_mod.F90:1175-1181, 1196-1205 (limiter_type = 'none' | 'pd'); selection :131-140
inline double slope_simple(double fm1, double, double fp1) { return (fp1 - fm1) * 0.5; }
inline double slope_pd(double fm1, double f, double fp1) {
  double df = (fp1 - fm1) * 0.5;
  return fsign(std::fmin(std::fabs(df), 2 * f), df);
}
inline double slope(int lim, double fm1, double f, double fp1) {
  return lim == GMC_LIMITER_NONE ? slope_simple(fm1, f, fp1) : lim == GMC_LIMITER_PD ? slope_pd(fm1, f, fp1) : slope_mono(fm1, f, fp1);
}

// ffsl_mod.F90:1147-1173
inline void ppm(int lim, double fm2, double fm1, double f, double fp1, double fp2, double& fl, double& df, double& f6) {
  double dfl = slope(lim, fm2, fm1, f);
  df = slope(lim, fm1, f, fp1);
  double dfr = slope(lim, f, fp1, fp2);
  fl = 0.5 * (fm1 + f) + (dfl - df) / 6.0;
  double fr = 0.5 * (fp1 + f) + (df - dfr) / 6.0;
  fl = f - fsign(std::fmin(std::fabs(df), std::fabs(fl - f)), df);
  fr = f + fsign(std::fmin(std::fabs(df), std::fabs(fr - f)), df);
  f6 = 6 * f - 3 * (fl + fr);
  df = fr - fl;
}

// ffsl_mod.F90:778-881
inline void hflx_ppm_cell(Block& b, AdvBatch& B, const A3& u, const A3& v, const A3& mx, const A3& my, A3& mfx, A3& mfy) {
  const Mesh& mesh = b.mesh;
  A3 &mlx = B.qlx, &mly = B.qly, &dmx = B.dqx, &dmy = B.dqy, &m6x = B.q6x, &m6y = B.q6y;
  const A3 &cflx = B.cflx, &cfly = B.cfly;
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = 1; j <= mesh.nlat; ++j) for (int i = 1; i <= mesh.nlon; ++i) {
      ppm(b.cfg.limiter_type, mx(i - 2, j, k), mx(i - 1, j, k), mx(i, j, k), mx(i + 1, j, k), mx(i + 2, j, k), mlx(i, j, k), dmx(i, j, k), m6x(i, j, k));
      ppm(b.cfg.limiter_type, my(i, j - 2, k), my(i, j - 1, k), my(i, j, k), my(i, j + 1, k), my(i, j + 2, k), mly(i, j, k), dmy(i, j, k), m6y(i, j, k));
    }
  fill_halo(mesh, mlx, true, true, true, false, false);
  fill_halo(mesh, dmx, true, true, true, false, false);
  fill_halo(mesh, m6x, true, true, true, false, false);
  fill_halo(mesh, mly, true, false, false, true, true);
  fill_halo(mesh, dmy, true, false, false, true, true);
  fill_halo(mesh, m6y, true, false, false, true, true);
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k) {
    for (int j = mesh.full_lat_ibeg_no_pole; j <= mesh.full_lat_iend_no_pole; ++j) for (int i = 1; i <= mesh.nlon; ++i) {
      int ci = (int)cflx(i, j, k);
      double cf = cflx(i, j, k) - ci;
      if (cflx(i, j, k) > 0) {
        int iu = i - ci;
        double s1 = 1 - cf, s2 = 1;
        double ds1 = s2 - s1, ds2 = s2 * s2 - s1 * s1, ds3 = s2 * s2 * s2 - s1 * s1 * s1;
        double sm = 0; for (int l = i + 1 - ci; l <= i; ++l) sm += mx(l, j, k);
        mfx(i, j, k) = u(i, j, k) * (sm + mlx(iu, j, k) * ds1 + 0.5 * dmx(iu, j, k) * ds2 + m6x(iu, j, k) * (ds2 / 2.0 - ds3 / 3.0)) / cflx(i, j, k);
      } else if (cflx(i, j, k) < 0) {
        int iu = i - ci + 1;
        double s1 = 0, s2 = -cf;
        double ds1 = s2 - s1, ds2 = s2 * s2 - s1 * s1, ds3 = s2 * s2 * s2 - s1 * s1 * s1;
        double sm = 0; for (int l = i + 1; l <= i - ci; ++l) sm += mx(l, j, k);
        mfx(i, j, k) = -u(i, j, k) * (sm + mlx(iu, j, k) * ds1 + 0.5 * dmx(iu, j, k) * ds2 + m6x(iu, j, k) * (ds2 / 2.0 - ds3 / 3.0)) / cflx(i, j, k);
      } else {
        mfx(i, j, k) = 0;
      }
    }
    for (int j = 1; j <= mesh.nhalf; ++j) for (int i = 1; i <= mesh.nlon; ++i) {
      if (cfly(i, j, k) > 0) {
        int ju = j;
        double s1 = 1 - cfly(i, j, k), s2 = 1;
        double ds1 = s2 - s1, ds2 = s2 * s2 - s1 * s1, ds3 = s2 * s2 * s2 - s1 * s1 * s1;
        mfy(i, j, k) = v(i, j, k) * (mly(i, ju, k) * ds1 + 0.5 * dmy(i, ju, k) * ds2 + m6y(i, ju, k) * (ds2 / 2.0 - ds3 / 3.0)) / cfly(i, j, k);
      } else if (cfly(i, j, k) < 0) {
        int ju = j + 1;
        double s1 = 0, s2 = -cfly(i, j, k);
        double ds1 = s2 - s1, ds2 = s2 * s2 - s1 * s1, ds3 = s2 * s2 * s2 - s1 * s1 * s1;
        mfy(i, j, k) = -v(i, j, k) * (mly(i, ju, k) * ds1 + 0.5 * dmy(i, ju, k) * ds2 + m6y(i, ju, k) * (ds2 / 2.0 - ds3 / 3.0)) / cfly(i, j, k);
      } else {
        mfy(i, j, k) = 0;
      }
    }
  }
}

// ffsl_mod.F90:883-941
inline void vflx_ppm_cell(Block& b, AdvBatch& B, const A3& w, const A3& m, A3& mfz) {
  const Mesh& mesh = b.mesh;
  A3 &mlz = B.qlx, &dmz = B.dqx, &m6z = B.q6x; const A3& cflz = B.cflz;
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = 1; j <= mesh.nlat; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      ppm(b.cfg.limiter_type, m(i, j, k - 2), m(i, j, k - 1), m(i, j, k), m(i, j, k + 1), m(i, j, k + 2), mlz(i, j, k), dmz(i, j, k), m6z(i, j, k));
  ORC_PAR for (int k = mesh.half_lev_ibeg + 1; k <= mesh.half_lev_iend - 1; ++k)
    for (int j = 1; j <= mesh.nlat; ++j) for (int i = 1; i <= mesh.nlon; ++i) {
      int ci = (int)cflz(i, j, k);
      double cf = cflz(i, j, k) - ci;
      if (cflz(i, j, k) > 0) {
        int ku = k - ci - 1;
        double s1 = 1 - cf, s2 = 1;
        double ds1 = s2 - s1, ds2 = s2 * s2 - s1 * s1, ds3 = s2 * s2 * s2 - s1 * s1 * s1;
        double sm = 0; for (int l = k - ci; l <= k - 1; ++l) sm += m(i, j, l);
        mfz(i, j, k) = w(i, j, k) * (sm + mlz(i, j, ku) * ds1 + 0.5 * dmz(i, j, ku) * ds2 + m6z(i, j, ku) * (ds2 / 2.0 - ds3 / 3.0)) / cflz(i, j, k);
      } else if (cflz(i, j, k) < 0) {
        int ku = k - ci;
        double s1 = 0, s2 = -cf;
        double ds1 = s2 - s1, ds2 = s2 * s2 - s1 * s1, ds3 = s2 * s2 * s2 - s1 * s1 * s1;
        double sm = 0; for (int l = k; l <= k - ci - 1; ++l) sm += m(i, j, l);
        mfz(i, j, k) = -w(i, j, k) * (sm + mlz(i, j, ku) * ds1 + 0.5 * dmz(i, j, ku) * ds2 + m6z(i, j, ku) * (ds2 / 2.0 - ds3 / 3.0)) / cflz(i, j, k);
      } else {
        mfz(i, j, k) = 0;
      }
    }
}

// ffsl_mod.F90:556-616 (ffsl_flux_type = 'van_leer')
inline void hflx_van_leer_cell(Block& b, AdvBatch& B, const A3& u, const A3& v, const A3& mx, const A3& my, A3& mfx, A3& mfy) {
  const Mesh& mesh = b.mesh;
  const int lim = b.cfg.limiter_type;
  const A3 &cflx = B.cflx, &cfly = B.cfly;
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k) {
    for (int j = mesh.full_lat_ibeg_no_pole; j <= mesh.full_lat_iend_no_pole; ++j) for (int i = 1; i <= mesh.nlon; ++i) {
      int ci = (int)cflx(i, j, k);
      double cf = cflx(i, j, k) - ci;
      if (cflx(i, j, k) > 0) {
        int iu = i - ci;
        double dm = slope(lim, mx(iu - 1, j, k), mx(iu, j, k), mx(iu + 1, j, k));
        double sm = 0; for (int l = i + 1 - ci; l <= i; ++l) sm += mx(l, j, k);
        mfx(i, j, k) = u(i, j, k) * (cf * (mx(iu, j, k) + dm * 0.5 * (1 - cf)) + sm) / cflx(i, j, k);
      } else if (cflx(i, j, k) < 0) {
        int iu = i - ci + 1;
        double dm = slope(lim, mx(iu - 1, j, k), mx(iu, j, k), mx(iu + 1, j, k));
        double sm = 0; for (int l = i + 1; l <= i - ci; ++l) sm += mx(l, j, k);
        mfx(i, j, k) = u(i, j, k) * (cf * (mx(iu, j, k) - dm * 0.5 * (1 + cf)) - sm) / cflx(i, j, k);
      } else {
        mfx(i, j, k) = 0;
      }
    }
    for (int j = 1; j <= mesh.nhalf; ++j) for (int i = 1; i <= mesh.nlon; ++i) {
      double cf = cfly(i, j, k);
      int ju = cf > 0 ? j : j + 1;
      double dm = slope(lim, my(i, ju - 1, k), my(i, ju, k), my(i, ju + 1, k));
      mfy(i, j, k) = v(i, j, k) * (my(i, ju, k) + dm * 0.5 * (std::copysign(1.0, cf) - cf));
    }
  }
}

// ffsl_mod.F90:618-658
inline void vflx_van_leer_cell(Block& b, AdvBatch& B, const A3& w, const A3& m, A3& mfz) {
  const Mesh& mesh = b.mesh;
  const int lim = b.cfg.limiter_type;
  const A3& cflz = B.cflz;
  ORC_PAR for (int k = mesh.half_lev_ibeg + 1; k <= mesh.half_lev_iend - 1; ++k)
    for (int j = 1; j <= mesh.nlat; ++j) for (int i = 1; i <= mesh.nlon; ++i) {
      int ci = (int)cflz(i, j, k);
      double cf = cflz(i, j, k) - ci;
      if (cflz(i, j, k) > 0) {
        int ku = k - ci - 1;
        double dm = slope(lim, m(i, j, ku - 1), m(i, j, ku), m(i, j, ku + 1));
        double sm = 0; for (int l = k - ci; l <= k - 1; ++l) sm += m(i, j, l);
        mfz(i, j, k) = w(i, j, k) * (cf * (m(i, j, ku) + dm * 0.5 * (1 - cf)) + sm) / cflz(i, j, k);
      } else if (cflz(i, j, k) < 0) {
        int ku = k - ci;
        double dm = slope(lim, m(i, j, ku - 1), m(i, j, ku), m(i, j, ku + 1));
        double sm = 0; for (int l = k; l <= k - ci - 1; ++l) sm += m(i, j, l);
        mfz(i, j, k) = w(i, j, k) * (cf * (m(i, j, ku) - dm * 0.5 * (1 + cf)) - sm) / cflz(i, j, k);
      } else {
        mfz(i, j, k) = 0;
      }
    }
}

// the hflx / vflx procedure pointers bound by ffsl_init (ffsl_mod.F90:116-129)
inline void hflx_cell(Block& b, AdvBatch& B, const A3& u, const A3& v, const A3& mx, const A3& my, A3& mfx, A3& mfy) {
  if (b.cfg.ffsl_flux_type == GMC_FLUX_VAN_LEER) hflx_van_leer_cell(b, B, u, v, mx, my, mfx, mfy);
  else hflx_ppm_cell(b, B, u, v, mx, my, mfx, mfy);
}
inline void vflx_cell(Block& b, AdvBatch& B, const A3& w, const A3& m, A3& mfz) {
  if (b.cfg.ffsl_flux_type == GMC_FLUX_VAN_LEER) vflx_van_leer_cell(b, B, w, m, mfz);
  else vflx_ppm_cell(b, B, w, m, mfz);
}

// ffsl_mod.F90:318-413
inline void ffsl_calc_tracer_hflx_cell(Block& b, AdvBatch& B, const A3& q, A3& qmfx, A3& qmfy, double dt, bool has_dt) {
  const Mesh& mesh = b.mesh;
  double dt_ = has_dt ? dt : B.dt;
  A3 &qx = B.qx, &qy = B.qy; const A3 &divx = B.divx, &divy = B.divy;
  hflx_cell(b, B, B.u, B.v, q, q, qmfx, qmfy);
  fill_halo(mesh, qmfx, true, true, true, false, false);
  fill_halo(mesh, qmfy, false, false, false, true, true);
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = mesh.full_lat_ibeg_no_pole; j <= mesh.full_lat_iend_no_pole; ++j) for (int i = 1; i <= mesh.nlon; ++i) {
      qx(i, j, k) = q(i, j, k) - 0.5 * ((qmfx(i, j, k) - qmfx(i - 1, j, k)) * mesh.le_lon(j) / mesh.area_cell(j) - divx(i, j, k) * q(i, j, k)) * dt_;
      qy(i, j, k) = q(i, j, k) - 0.5 * ((qmfy(i, j, k) * mesh.le_lat(j) - qmfy(i, j - 1, k) * mesh.le_lat(j - 1)) / mesh.area_cell(j) - divy(i, j, k) * q(i, j, k)) * dt_;
    }
  std::vector<double> work(mesh.nlon);
  {
    int j = mesh.full_lat_ibeg;
    for (int k = 1; k <= mesh.nlev; ++k) {
      for (int i = 1; i <= mesh.nlon; ++i) work[i - 1] = qmfy(i, j, k);
      double pole = zonal_sum(work) * mesh.le_lat(j) / mesh.nlon / mesh.area_cell(j);
      for (int i = 1; i <= mesh.nlon; ++i) {
        qx(i, j, k) = q(i, j, k);
        qy(i, j, k) = q(i, j, k) - 0.5 * (pole - divy(i, j, k) * q(i, j, k)) * dt_;
      }
    }
  }
  {
    int j = mesh.full_lat_iend;
    for (int k = 1; k <= mesh.nlev; ++k) {
      for (int i = 1; i <= mesh.nlon; ++i) work[i - 1] = qmfy(i, j - 1, k);
      double pole = zonal_sum(work) * mesh.le_lat(j - 1) / mesh.nlon / mesh.area_cell(j);
      for (int i = 1; i <= mesh.nlon; ++i) {
        qx(i, j, k) = q(i, j, k);
        qy(i, j, k) = q(i, j, k) + 0.5 * (pole - divy(i, j, k) * q(i, j, k)) * dt_;
      }
    }
  }
  fill_halo(mesh, qx, true, false, false, true, true);
  fill_halo(mesh, qy, true, true, true, false, false);
  hflx_cell(b, B, B.mfx, B.mfy, qy, qx, qmfx, qmfy);
}

// ffsl_mod.F90:415-432
inline void ffsl_calc_tracer_vflx_cell(Block& b, AdvBatch& B, const A3& q, A3& qmfz) {
  vflx_cell(b, B, B.we, q, qmfz);
}

}  // namespace orc

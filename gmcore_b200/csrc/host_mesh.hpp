// Host-side stand-in for the Fortran host's one-time set-up (NOT the hot path).
//
// GMCORE's Fortran drivers own the mesh, the filter weights and the initial
// conditions and would hand them to the GPU library through the C ABI in
// include/gmcore_gpu.h.  No Fortran compiler exists on the build or GPU
// machines, so this header restates that set-up in C++ for the stand-alone
// driver (bench.py, tests).  References (read-only tree /root/reference):
//   src/dynamics/mesh_mod.F90:134-447      (mesh_init_global)
//   src/utils/sphere_geometry_mod.F90:110-122,302-427 (real(16) areas)
//   src/utils/const_mod.F90:51-85          (planet constants)
#pragma once
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>
#include <quadmath.h>

namespace gmh {

typedef __float128 q16;

struct Consts {
  double pi, pi2, pi05, deg, rad, eps;
  double omega, radius, g, Rd, Rv, cpd, cvd, lapse_rate, Rd_o_Rv, p0;
  double Rd_o_g, Rd_o_cpd, cpd_o_cvd, cvd_o_cpd;
  Consts() {
    pi = std::atan(1.0) * 4.0; pi2 = pi * 2; pi05 = pi * 0.5;
    deg = 180.0 / pi; rad = pi / 180.0; eps = 2.220446049250313e-16;
    // const_mod.F90:56-66 'earth'
    omega = 2 * pi / 86400.0; radius = 6.37122e6; g = 9.80616; Rd = 287.04;
    Rv = 461.497; cpd = 1004.0; cvd = 717.0; lapse_rate = 0.006;
    Rd_o_Rv = Rd / Rv; p0 = 1.0e5;
    Rd_o_g = Rd / g; Rd_o_cpd = Rd / cpd; cpd_o_cvd = cpd / cvd; cvd_o_cpd = cvd / cpd;
  }
};

// 1-D vector with a Fortran lower bound.
struct Vec {
  std::vector<double> d; int lb = 0;
  void alloc(int lb_, int ub_, double fill) { lb = lb_; d.assign(ub_ - lb_ + 1, fill); }
  double& operator()(int j) { return d[j - lb]; }
  double operator()(int j) const { return d[j - lb]; }
  int ub() const { return lb + (int)d.size() - 1; }
};
struct Vec2 {  // (2, lb:ub)
  std::vector<double> d; int lb = 0;
  void alloc(int lb_, int ub_, double fill) { lb = lb_; d.assign(2 * (ub_ - lb_ + 1), fill); }
  double& operator()(int c, int j) { return d[(j - lb) * 2 + (c - 1)]; }
  double operator()(int c, int j) const { return d[(j - lb) * 2 + (c - 1)]; }
};

// ---- real(16) spherical geometry (sphere_geometry_mod.F90) -------------------
struct P3 { q16 x, y, z; };

inline P3 cart16(const Consts& c, double lon, double lat) {
  // cartesian_transform_1_r16: cos/sin are evaluated in r8, then widened (T16)
  q16 cos_lat = (q16)std::cos(lat);
  P3 p;
  p.x = (q16)c.radius * cos_lat * (q16)std::cos(lon);
  p.y = (q16)c.radius * cos_lat * (q16)std::sin(lon);
  p.z = (q16)c.radius * (q16)std::sin(lat);
  return p;
}
inline void cross(const q16 a[3], const q16 b[3], q16 r[3]) {
  r[0] = a[1] * b[2] - a[2] * b[1];
  r[1] = a[2] * b[0] - a[0] * b[2];
  r[2] = a[0] * b[1] - a[1] * b[0];
}
inline void normv(q16 x[3]) {
  q16 n = sqrtq(x[0] * x[0] + x[1] * x[1] + x[2] * x[2]);
  if (n != 0) { x[0] /= n; x[1] /= n; x[2] /= n; }
}
inline q16 sphere_angle(const Consts& c, const P3& A, const P3& B, const P3& C) {
  q16 a[3] = {A.x, A.y, A.z}, b[3] = {B.x, B.y, B.z}, cc[3] = {C.x, C.y, C.z};
  q16 nab[3], nbc[3], t[3];
  cross(a, b, nab); normv(nab);
  cross(b, cc, nbc); normv(nbc);
  q16 dp = nab[0] * nbc[0] + nab[1] * nbc[1] + nab[2] * nbc[2];
  if (dp > (q16)1.0) dp = (q16)1.0;
  if (dp < -(q16)1.0) dp = -(q16)1.0;
  q16 res = acosq(-dp);
  cross(nab, nbc, t);
  if (t[0] * a[0] + t[1] * a[1] + t[2] * a[2] < (q16)0.0) res = (q16)c.pi2 - res;
  return res;
}
inline q16 calc_area(const Consts& c, const P3 p[3]) {
  q16 res = 0;
  for (int i = 0; i < 3; ++i) {
    int im1 = (i == 0) ? 2 : i - 1, ip1 = (i == 2) ? 0 : i + 1;
    res += sphere_angle(c, p[im1], p[i], p[ip1]);
  }
  q16 r2 = (q16)c.radius * (q16)c.radius;
  return r2 * (res - (q16)(3 - 2) * (q16)c.pi);
}
inline void inv_cart16(const Consts& c, const P3& p, q16& lon, q16& lat) {
  lon = atan2q(p.y, p.x);
  lat = asinq(p.z / (q16)c.radius);
  if (lon < (q16)0.0) lon += (q16)c.pi2;
}
inline q16 calc_area_small_arc(const Consts& c, const P3 p[3]) {
  q16 res = calc_area(c, p);
  q16 lon0, lat0, lon1, lat1, lon2, lat2;
  inv_cart16(c, p[0], lon0, lat0);
  inv_cart16(c, p[1], lon1, lat1);
  inv_cart16(c, p[2], lon2, lat2);
  if (lat1 != lat2) { std::fprintf(stderr, "gmh: small arc is not valid\n"); std::abort(); }
  if (lat1 == (q16)0.0) return res;
  q16 dlon = (lat0 > lat1) ? (lon2 - lon1) : (lon1 - lon2);
  if (dlon < (q16)0.0) dlon += (q16)c.pi2;
  P3 v[3];
  v[0].x = 0; v[0].y = 0;
  bool pole_side = (lat0 * lat1 >= 0) && (fabsq(lat0) > fabsq(lat1));
  if (pole_side) { v[1] = p[1]; v[2] = p[2]; } else { v[1] = p[2]; v[2] = p[1]; }
  q16 r2 = (q16)c.radius * (q16)c.radius, area1;
  if (lat1 > (q16)0.0) { v[0].z = (q16)c.radius;  area1 = r2 * dlon * ((q16)1.0 - sinq(lat1)); }
  else             { v[0].z = -(q16)c.radius; area1 = r2 * dlon * (sinq(lat1) + (q16)1.0); }
  q16 area2 = calc_area(c, v);
  q16 area3 = area1 - area2;
  if (area3 < (q16)0.0 && fabsq(area3) > (q16)1.0e-10f) area3 = 0;
  if (pole_side) res += area3; else res -= area3;
  if (res < (q16)0.0) { std::fprintf(stderr, "gmh: failed small-arc area\n"); std::abort(); }
  return res;
}

// ---- mesh (mesh_mod.F90:134-447) -------------------------------------------
struct Mesh {
  int nlon, nlat, nlev, hx, hy;
  int full_lon_lb, full_lon_ub, full_lat_lb, full_lat_ub, half_lat_lb, half_lat_ub;
  int num_half_lat;
  double dlon;
  Vec dlat, full_lon, half_lon, full_lat, half_lat;
  Vec full_cos_lon, half_cos_lon, full_sin_lon, half_sin_lon;
  Vec full_cos_lat, half_cos_lat, full_sin_lat, half_sin_lat;
  Vec full_lon_deg, half_lon_deg, full_lat_deg, half_lat_deg;
  Vec area_cell, area_lon, area_lon_west, area_lon_east, area_lon_north, area_lon_south;
  Vec area_lat, area_lat_west, area_lat_east, area_lat_north, area_lat_south, area_vtx;
  Vec2 area_subcell;
  Vec de_lon, de_lat, le_lat, le_lon, full_f, half_f;
  Vec2 full_tangent_wgt, half_tangent_wgt;
  // vertical (filled by vert coord set-up; bounds full -1:nlev+2, half 1:nlev+1)
  Vec full_lev, half_lev, full_dlev, half_dlev;
  // namelist_mod.F90:85-86 and tangent_wgt_scheme ('classic' | 'thuburn09'); set before init()
  double coarse_pole_mul = 0, coarse_pole_decay = 100.0;
  int thuburn09 = 0;

  void init(const Consts& c, int nlon_, int nlat_, int nlev_, int hx_, int hy_ = 2) {
    const double inf = 1.79769313486231570815e+308;  // huge(1.0_r8)
    nlon = nlon_; nlat = nlat_; nlev = nlev_; hx = hx_; hy = hy_;
    num_half_lat = nlat - 1;
    full_lon_lb = 1 - hx; full_lon_ub = nlon + hx;
    full_lat_lb = 1 - hy; full_lat_ub = nlat + hy;
    half_lat_lb = 1 - hy; half_lat_ub = num_half_lat + hy;
    auto FL = [&](Vec& v, double f) { v.alloc(full_lat_lb, full_lat_ub, f); };
    auto HL = [&](Vec& v, double f) { v.alloc(half_lat_lb, half_lat_ub, f); };
    auto LO = [&](Vec& v, double f) { v.alloc(full_lon_lb, full_lon_ub, f); };
    HL(dlat, 0); LO(full_lon, inf); LO(half_lon, inf); FL(full_lat, inf); HL(half_lat, inf);
    LO(full_cos_lon, inf); LO(half_cos_lon, inf); LO(full_sin_lon, inf); LO(half_sin_lon, inf);
    FL(full_cos_lat, inf); HL(half_cos_lat, inf); FL(full_sin_lat, inf); HL(half_sin_lat, inf);
    LO(full_lon_deg, inf); LO(half_lon_deg, inf); FL(full_lat_deg, inf); HL(half_lat_deg, inf);
    FL(area_cell, 0); FL(area_lon, 0); FL(area_lon_west, 0); FL(area_lon_east, 0);
    FL(area_lon_north, 0); FL(area_lon_south, 0);
    HL(area_lat, 0); HL(area_lat_west, 0); HL(area_lat_east, 0); HL(area_lat_north, 0);
    HL(area_lat_south, 0); HL(area_vtx, 0);
    area_subcell.alloc(full_lat_lb, full_lat_ub, 0);
    FL(de_lon, 0); HL(de_lat, 0); HL(le_lat, 0); FL(le_lon, 0); FL(full_f, inf); HL(half_f, inf);
    full_tangent_wgt.alloc(full_lat_lb, full_lat_ub, inf);
    half_tangent_wgt.alloc(half_lat_lb, half_lat_ub, inf);
    full_lev.alloc(-1, nlev + 2, inf); half_lev.alloc(1, nlev + 1, inf);
    full_dlev.alloc(-1, nlev + 2, 0); half_dlev.alloc(1, nlev + 1, 0);

    const double start_lon = 0.0, end_lon = c.pi2, start_lat = -c.pi05, end_lat = c.pi05;
    const double radius = c.radius;
    dlon = (end_lon - start_lon) / nlon;
    for (int i = full_lon_lb; i <= full_lon_ub; ++i) {
      full_lon(i) = start_lon + (i - 1) * dlon;
      half_lon(i) = full_lon(i) + 0.5 * dlon;
      full_lon_deg(i) = full_lon(i) * c.deg;
      half_lon_deg(i) = half_lon(i) * c.deg;
    }
    double dlat0 = (end_lat - start_lat) / num_half_lat;
    for (int j = 1; j <= num_half_lat; ++j) {
      half_lat(j) = start_lat + (j - 0.5) * dlat0;
      if (std::fabs(half_lat(j)) < (double)1.0e-12f) half_lat(j) = 0.0;
    }
    if (coarse_pole_mul != 0) {          // mesh_mod.F90:195-202: dlat stretched towards the poles, renormalised to pi
      dlat0 = dlon;
      double sum = 0;
      for (int j = 1; j <= num_half_lat; ++j) {
        const double d = std::fabs(half_lat(j)) - c.pi05;
        dlat(j) = dlat0 * (1 + (coarse_pole_mul - 1) * std::exp(-coarse_pole_decay * (d * d)));
        sum += dlat(j);
      }
      for (int j = 1; j <= num_half_lat; ++j) dlat(j) = dlat(j) / sum * c.pi;
    } else {
      for (int j = 1; j <= num_half_lat; ++j) dlat(j) = dlat0;
    }
    full_lat(1) = start_lat; full_lat_deg(1) = start_lat * c.deg;
    for (int j = 2; j <= nlat - 1; ++j) {
      full_lat(j) = full_lat(j - 1) + dlat(j - 1);
      if (std::fabs(full_lat(j)) < (double)1.0e-12f) full_lat(j) = 0.0;
      full_lat_deg(j) = full_lat(j) * c.deg;
    }
    full_lat(nlat) = end_lat; full_lat_deg(nlat) = end_lat * c.deg;
    for (int j = 1; j <= num_half_lat; ++j) {
      if (full_lat(j) == c.pi05) continue;
      half_lat(j) = full_lat(j) + 0.5 * dlat(j);
      if (std::fabs(half_lat(j)) < (double)1.0e-12f) half_lat(j) = 0.0;
      half_lat_deg(j) = half_lat(j) * c.deg;
    }
    for (int j = 1; j <= nlat; ++j)
      if (full_lat(j) > 0) {
        full_lat(j) = -full_lat(nlat - j + 1);
        full_lat_deg(j) = -full_lat_deg(nlat - j + 1);
      }
    for (int j = 1; j <= num_half_lat; ++j)
      if (half_lat(j) > 0) {
        half_lat(j) = -half_lat(num_half_lat - j + 1);
        half_lat_deg(j) = -half_lat_deg(num_half_lat - j + 1);
      }
    for (int i = full_lon_lb; i <= full_lon_ub; ++i) {
      full_cos_lon(i) = std::cos(full_lon(i)); full_sin_lon(i) = std::sin(full_lon(i));
      half_cos_lon(i) = std::cos(half_lon(i)); half_sin_lon(i) = std::sin(half_lon(i));
    }
    for (int j = half_lat_lb; j <= half_lat_ub; ++j)
      if (half_lat(j) >= -c.pi05 && half_lat(j) <= c.pi05) {
        half_cos_lat(j) = std::cos(half_lat(j)); half_sin_lat(j) = std::sin(half_lat(j));
      }
    for (int j = full_lat_lb; j <= full_lat_ub; ++j)
      if (full_lat(j) >= -c.pi05 && full_lat(j) <= c.pi05) {
        full_cos_lat(j) = std::cos(full_lat(j)); full_sin_lat(j) = std::sin(full_lat(j));
      }
    full_cos_lat(1) = 0.0; full_sin_lat(1) = -1.0;
    full_cos_lat(nlat) = 0.0; full_sin_lat(nlat) = 1.0;

    P3 p[3];
    for (int j = 1; j <= nlat; ++j) {
      if (j == 1) {
        area_cell(j) = radius * radius * dlon * (half_sin_lat(j) + 1.0);
        area_subcell(2, j) = radius * radius * 0.5 * dlon * (half_sin_lat(j) + 1.0);
      } else if (j == nlat) {
        area_cell(j) = radius * radius * dlon * (1.0 - half_sin_lat(j - 1));
        area_subcell(1, j) = radius * radius * 0.5 * dlon * (1.0 - half_sin_lat(j - 1));
      } else {
        area_cell(j) = radius * radius * dlon * (half_sin_lat(j) - half_sin_lat(j - 1));
        area_subcell(1, j) = radius * radius * 0.5 * dlon * (full_sin_lat(j) - half_sin_lat(j - 1));
        area_subcell(2, j) = radius * radius * 0.5 * dlon * (half_sin_lat(j) - full_sin_lat(j));
        p[0] = cart16(c, full_lon(1), full_lat(j));
        p[1] = cart16(c, half_lon(1), half_lat(j - 1));
        p[2] = cart16(c, half_lon(1), half_lat(j));
        area_lon_west(j) = (double)calc_area(c, p);
        area_lon_east(j) = area_lon_west(j);
        area_lon(j) = area_lon_west(j) + area_lon_east(j);
        p[0] = cart16(c, half_lon(1), half_lat(j));
        p[1] = cart16(c, full_lon(1), full_lat(j));
        p[2] = cart16(c, full_lon(2), full_lat(j));
        area_lon_north(j) = (double)calc_area_small_arc(c, p);
        p[0] = cart16(c, half_lon(1), half_lat(j - 1));
        p[1] = cart16(c, full_lon(2), full_lat(j));
        p[2] = cart16(c, full_lon(1), full_lat(j));
        area_lon_south(j) = (double)calc_area_small_arc(c, p);
      }
    }
    for (int j = 1; j <= num_half_lat; ++j) {
      area_vtx(j) = radius * radius * dlon * (full_sin_lat(j + 1) - full_sin_lat(j));
      p[0] = cart16(c, half_lon(1), half_lat(j));
      p[1] = cart16(c, full_lon(2), full_lat(j));
      p[2] = cart16(c, full_lon(2), full_lat(j + 1));
      area_lat_west(j) = (double)calc_area(c, p);
      area_lat_east(j) = area_lat_west(j);
      p[0] = cart16(c, full_lon(2), full_lat(j + 1));
      p[1] = cart16(c, half_lon(1), half_lat(j));
      p[2] = cart16(c, half_lon(2), half_lat(j));
      area_lat_north(j) = (double)calc_area_small_arc(c, p);
      p[0] = cart16(c, full_lon(2), full_lat(j));
      p[1] = cart16(c, half_lon(2), half_lat(j));
      p[2] = cart16(c, half_lon(1), half_lat(j));
      area_lat_south(j) = (double)calc_area_small_arc(c, p);
      if (j == 1) area_lat_south(j) = area_cell(j);
      else if (j + 1 == nlat) area_lat_north(j) = area_cell(j + 1);
      area_lat(j) = area_lat_north(j) + area_lat_south(j);
    }
    for (int j = 2; j <= nlat - 1; ++j) {
      de_lon(j) = radius * full_cos_lat(j) * dlon;
      le_lon(j) = 2.0 * area_lon(j) / de_lon(j);
    }
    le_lon(1) = 0; de_lon(1) = 0; le_lon(nlat) = 0; de_lon(nlat) = 0;
    for (int j = 1; j <= num_half_lat; ++j) {
      le_lat(j) = radius * half_cos_lat(j) * dlon;
      de_lat(j) = 2.0 * area_lat(j) / le_lat(j);
    }
    // tangent_wgt_scheme = 'classic' | 'thuburn09' (mesh_mod.F90:417-438)
    for (int j = 2; j <= nlat - 1; ++j) {
      full_tangent_wgt(1, j) = le_lat(j - 1) / de_lon(j) * (thuburn09 ? area_subcell(2, j) / area_cell(j) : 0.25);
      full_tangent_wgt(2, j) = le_lat(j) / de_lon(j) * (thuburn09 ? area_subcell(1, j) / area_cell(j) : 0.25);
    }
    for (int j = 1; j <= num_half_lat; ++j) {
      half_tangent_wgt(1, j) = le_lon(j) / de_lat(j) * (thuburn09 ? area_subcell(1, j) / area_cell(j) : 0.25);
      half_tangent_wgt(2, j) = le_lon(j + 1) / de_lat(j) * (thuburn09 ? area_subcell(2, j + 1) / area_cell(j + 1) : 0.25);
    }
    for (int j = 1; j <= nlat; ++j) full_f(j) = 2.0 * c.omega * full_sin_lat(j);
    for (int j = 1; j <= num_half_lat; ++j) half_f(j) = 2.0 * c.omega * half_sin_lat(j);
  }
};

}  // namespace gmh

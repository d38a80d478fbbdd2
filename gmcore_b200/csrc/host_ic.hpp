// Host-side stand-in: analytic initial conditions of the BASELINE configs (NOT
// the hot path).  Arrays use the reference's halo'd Fortran storage
// (lon 1-hx:nlon+hx, lat 1-2:nlat(+/-)+2, full lev -1:nlev+2) so that the same
// buffers a Fortran host would pass by c_loc() go through the C ABI unchanged.
// Restates
//   src/tests/swm/rossby_haurwitz_wave_test_mod.F90:31-87   (Williamson 6)
//   src/tests/swm/mountain_zonal_flow_test_mod.F90:24-81    (Williamson 5)
#pragma once
#include "host_mesh.hpp"

namespace gmh {

// Index helper for the halo'd Fortran layout.
struct Lay {
  int hx, hy, nlon, nlat, nlev;
  int nx() const { return nlon + 2 * hx; }
  int ny_full() const { return nlat + 2 * hy; }
  int ny_half() const { return nlat - 1 + 2 * hy; }
  int nz_full() const { return nlev + 4; }
  int nz_half() const { return nlev + 1; }
  size_t n2_full() const { return (size_t)nx() * ny_full(); }
  size_t n2_half() const { return (size_t)nx() * ny_half(); }
  // full-lat array (3-D, full lev)
  size_t f(int i, int j, int k) const { return (size_t)(i - 1 + hx) + (size_t)nx() * ((size_t)(j - 1 + hy) + (size_t)ny_full() * (k + 1)); }
  // half-lat array (3-D, full lev)
  size_t h(int i, int j, int k) const { return (size_t)(i - 1 + hx) + (size_t)nx() * ((size_t)(j - 1 + hy) + (size_t)ny_half() * (k + 1)); }
  size_t f2(int i, int j) const { return (size_t)(i - 1 + hx) + (size_t)nx() * (size_t)(j - 1 + hy); }
};

inline void ic_rossby_haurwitz(const Consts& c, const Mesh& m, const Lay& L,
                               double* u, double* v, double* gz, double* gzs) {
  const double R = 4.0, omg = 7.848e-6;
  const double gz0 = 8.0e3 * c.g;
  const double radius = c.radius, omega = c.omega;
  for (size_t n = 0; n < L.n2_full(); ++n) gzs[n] = 0.0;
  for (int j = 1; j <= m.nlat; ++j) {
    double cos_lat = m.full_cos_lat(j), sin_lat = m.full_sin_lat(j);
    for (int i = 1; i <= m.nlon; ++i) {
      double lon = m.half_lon(i);
      double a = cos_lat;
      double b = R * std::pow(cos_lat, R - 1) * (sin_lat * sin_lat) * std::cos(R * lon);
      double cc = std::pow(cos_lat, R + 1) * std::cos(R * lon);
      u[L.f(i, j, 1)] = radius * omg * (a + b - cc);
    }
  }
  for (int j = 1; j <= m.num_half_lat; ++j) {
    double cos_lat = m.half_cos_lat(j), sin_lat = m.half_sin_lat(j);
    for (int i = 1; i <= m.nlon; ++i) {
      double lon = m.full_lon(i);
      double a = R * std::pow(cos_lat, R - 1) * sin_lat * std::sin(R * lon);
      v[L.h(i, j, 1)] = -radius * omg * a;
    }
  }
  for (int j = 1; j <= m.nlat; ++j) {
    double cos_lat = m.full_cos_lat(j);
    double a = 0.5 * omg * (2 * omega + omg) * (cos_lat * cos_lat) +
               0.25 * (omg * omg) * ((R + 1) * std::pow(cos_lat, 2 * R + 2) +
                                     (2 * (R * R) - R - 2) * std::pow(cos_lat, 2 * R) -
                                     2 * (R * R) * std::pow(cos_lat, 2 * R - 2));
    double b = 2 * (omega + omg) * omg * std::pow(cos_lat, R) *
               (R * R + 2 * R + 2 - (R + 1) * (R + 1) * (cos_lat * cos_lat)) / (R + 1) / (R + 2);
    double cc = 0.25 * (omg * omg) * std::pow(cos_lat, 2 * R) * ((R + 1) * (cos_lat * cos_lat) - R - 2);
    for (int i = 1; i <= m.nlon; ++i) {
      double lon = m.full_lon(i);
      gz[L.f(i, j, 1)] = gz0 + radius * radius * (a + b * std::cos(R * lon) + cc * std::cos(2 * R * lon));
    }
  }
}

inline void ic_mountain_zonal_flow(const Consts& c, const Mesh& m, const Lay& L,
                                   double* u, double* v, double* gz, double* gzs) {
  const double alpha = 0.0, u0 = 20.0;
  const double lon0 = c.pi * 1.5, lat0 = c.pi / 6.0, R = c.pi / 9.0;
  const double gz0 = 5960.0 * c.g, gzs0 = 2000.0 * c.g;
  const double cos_alpha = std::cos(alpha), sin_alpha = std::sin(alpha);
  for (int j = 1; j <= m.nlat; ++j)
    for (int i = 1; i <= m.nlon; ++i) {
      double dlon = std::fabs(m.full_lon(i) - lon0);
      dlon = std::fmin(dlon, 2 * c.pi - dlon);
      double dl = m.full_lat(j) - lat0;
      double d = std::fmin(R, std::sqrt(dlon * dlon + dl * dl));
      gzs[L.f2(i, j)] = gzs0 * (1.0 - d / R);
    }
  for (int j = 1; j <= m.nlat; ++j) {
    double cos_lat = m.full_cos_lat(j), sin_lat = m.full_sin_lat(j);
    for (int i = 1; i <= m.nlon; ++i) {
      double cos_lon = m.half_cos_lon(i);
      u[L.f(i, j, 1)] = u0 * (cos_lat * cos_alpha + cos_lon * sin_lat * sin_alpha);
    }
  }
  for (int j = 1; j <= m.num_half_lat; ++j)
    for (int i = 1; i <= m.nlon; ++i) {
      double sin_lon = m.full_sin_lon(i);
      v[L.h(i, j, 1)] = -u0 * sin_lon * sin_alpha;
    }
  for (int j = 1; j <= m.nlat; ++j) {
    double cos_lat = m.full_cos_lat(j), sin_lat = m.full_sin_lat(j);
    for (int i = 1; i <= m.nlon; ++i) {
      double cos_lon = m.full_cos_lon(i);
      double t = sin_lat * cos_alpha - cos_lon * cos_lat * sin_alpha;
      gz[L.f(i, j, 1)] = gz0 - (c.radius * c.omega * u0 + u0 * u0 * 0.5) * (t * t);
    }
  }
}

}  // namespace gmh

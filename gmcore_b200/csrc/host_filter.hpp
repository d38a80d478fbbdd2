// Host-side stand-in: zonal polar-filter weights (NOT the hot path; the
// convolution itself is the CUDA kernel).  Restates
//   src/dynamics/filter_types_mod.F90:31-147 (gaussian_weight, filter_init)
//   src/utils/math_mod.F90:146-159          (exp_two_values)
// Namelist defaults are single-precision literals widened to r8.
#pragma once
#include "host_mesh.hpp"

namespace gmh {

struct FilterCfg {
  double dt_dyn = 0;
  double max_wave_speed = 300;
  double max_cfl = (double)0.5f;
  double coef_a = (double)1.0f;
  double coef_b = (double)0.4f;
  double coef_c = (double)0.2f;
  double coef_d = (double)0.375f;
  double coef_e = (double)0.16f;
};

struct Filter {
  Vec width_lon, width_lat;
  std::vector<int> ngrid_lon, ngrid_lat;   // index j - lb
  int lon_lb = 0, lat_lb = 0, ld_lon = 0, ld_lat = 0;
  std::vector<double> wgt_lon, wgt_lat;    // (ld, lb:ub) column-major like Fortran
  int& nlon_at(int j) { return ngrid_lon[j - lon_lb]; }
  int& nlat_at(int j) { return ngrid_lat[j - lat_lb]; }
  double* wlon(int j) { return &wgt_lon[(size_t)(j - lon_lb) * ld_lon]; }
  double* wlat(int j) { return &wgt_lat[(size_t)(j - lat_lb) * ld_lat]; }
};

inline void gaussian_weight(const Consts& c, const FilterCfg& f, double width, int ngrid, double* w) {
  double s = f.coef_b * width / 2.0;
  for (int i = 1; i <= ngrid; ++i) {
    int x = i - (ngrid + 1) / 2;
    w[i - 1] = std::exp(-(x * x) / (2 * (s * s))) / (s * std::sqrt(c.pi2));
  }
  double sum = 0;
  for (int i = 0; i < ngrid; ++i) sum += w[i];
  for (int i = 0; i < ngrid; ++i) w[i] = w[i] / sum;
}

inline double exp_two_values(double val1, double val2, double x0, double x1, double x) {
  double w = std::exp((x - x0) * (x - x0) * std::log(1.0e-3) / ((x1 - x0) * (x1 - x0)));
  return w * val1 + (1 - w) * val2;
}

// type: 0 big_filter, 1 small_filter1, 2 small_filter2
inline void filter_init(const Consts& c, const Mesh& m, const FilterCfg& f, int type, Filter& F) {
  const int nlat = m.nlat, nhalf = m.num_half_lat;
  F.lon_lb = m.full_lat_lb; F.lat_lb = m.half_lat_lb;
  F.width_lon.alloc(m.full_lat_lb, m.full_lat_ub, 0);
  F.width_lat.alloc(m.half_lat_lb, m.half_lat_ub, 0);
  F.ngrid_lon.assign(m.full_lat_ub - m.full_lat_lb + 1, 0);
  F.ngrid_lat.assign(m.half_lat_ub - m.half_lat_lb + 1, 0);
  const double dt = f.dt_dyn;
  if (f.max_wave_speed > 0) {
    for (int j = 2; j <= nlat - 1; ++j) {
      double dx = m.de_lon(j);
      if (dx > 0) {
        double cfl = f.max_wave_speed * dt / dx;
        double w = f.coef_a * cfl / f.max_cfl * (f.coef_c * (std::tanh(90 - std::fabs(m.full_lat_deg(j))) - 1) + 1);
        int n = (int)std::ceil(w) + 2; if (n % 2 == 0) n = n + 1;
        F.width_lon(j) = w; F.nlon_at(j) = n;
      }
    }
    for (int j = 1; j <= nhalf; ++j) {
      double dx = m.le_lat(j);
      if (dx > 0) {
        double cfl = f.max_wave_speed * dt / dx;
        double w = f.coef_a * cfl / f.max_cfl * (f.coef_c * (std::tanh(90 - std::fabs(m.half_lat_deg(j))) - 1) + 1);
        int n = (int)std::ceil(w) + 2; if (n % 2 == 0) n = n + 1;
        F.width_lat(j) = w; F.nlat_at(j) = n;
      }
    }
  }
  int mx = 0; for (int n : F.ngrid_lon) mx = n > mx ? n : mx; F.ld_lon = mx > 0 ? mx : 1;
  mx = 0;     for (int n : F.ngrid_lat) mx = n > mx ? n : mx; F.ld_lat = mx > 0 ? mx : 1;
  F.wgt_lon.assign((size_t)F.ld_lon * F.ngrid_lon.size(), 0.0);
  F.wgt_lat.assign((size_t)F.ld_lat * F.ngrid_lat.size(), 0.0);
  if (type == 0) {
    for (int j = 2; j <= nlat - 1; ++j)
      if (F.nlon_at(j) > 1) gaussian_weight(c, f, F.width_lon(j), F.nlon_at(j), F.wlon(j));
    for (int j = 1; j <= nhalf; ++j)
      if (F.nlat_at(j) > 1) gaussian_weight(c, f, F.width_lat(j), F.nlat_at(j), F.wlat(j));
  } else {
    for (int pass = 0; pass < 2; ++pass) {
      // pass 0: full-lat rows (lon/cell filter), pass 1: half-lat rows
      int jb = pass == 0 ? 2 : 1, je = pass == 0 ? nlat - 1 : nhalf;
      double lat0 = pass == 0 ? -m.full_lat_deg(2) : -m.half_lat_deg(1);
      for (int j = jb; j <= je; ++j) {
        int& ng = pass == 0 ? F.nlon_at(j) : F.nlat_at(j);
        double& wd = pass == 0 ? F.width_lon(j) : F.width_lat(j);
        if (ng > 1) {
          double w;
          if (type == 1) w = f.coef_d;
          else w = exp_two_values(f.coef_e, f.coef_d, lat0, 80.0,
                                  std::fabs(pass == 0 ? m.full_lat_deg(j) : m.half_lat_deg(j)));
          int n = (int)(w * ng); if (n % 2 == 0) n = n + 1; ng = n;
          wd = w * wd;
          gaussian_weight(c, f, wd, ng, pass == 0 ? F.wlon(j) : F.wlat(j));
        }
      }
    }
  }
}

}  // namespace gmh

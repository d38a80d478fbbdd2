// Host-side stand-in: hybrid sigma-pressure coordinate set-up and the
// Jablonowski–Williamson initial condition (NOT the hot path).  Restates
//   src/dynamics/vert_coord/hybrid_coord_mod.F90:34-171
//   src/dynamics/vert_coord/vert_coord_mod.F90:47-102
//   src/dynamics/vert_coord/hybrid_coord_test_mod.F90:565-647 (wrf_l32), :1330-1352 (equal_eta)
//   src/tests/baroclinic_wave_test_mod.F90:28-170
// The coefficient tables are single-precision literals in the reference and are
// therefore rounded through float here as well.
#pragma once
#include "host_ic.hpp"

namespace gmh {

struct VertCoord {
  std::vector<double> hyai, hybi, hyam, hybm;  // 1-based via [k-1]
};

inline void vert_coord_init(Mesh& m, VertCoord& vc) {
  const int nlev = m.nlev;
  vc.hyai.assign(nlev + 1, 0.0); vc.hybi.assign(nlev + 1, 0.0);
  vc.hyam.assign(nlev, 0.0); vc.hybm.assign(nlev, 0.0);
  if (nlev == 32) {
    static const float a[33] = {
      0.0000000000000000f, 0.0027008056640625f, 0.0076868347823620f, 0.0158526937798342f, 0.0276367470939261f,
      0.0414515552099706f, 0.0562277844155789f, 0.0724737972365225f, 0.0893795253309864f, 0.1063548465048298f,
      0.1254725257849640f, 0.1479642011314251f, 0.1744233770857367f, 0.2054476619883455f, 0.2362396334209435f,
      0.2617261684535087f, 0.2783057827853432f, 0.2816243248017372f, 0.2683280168286641f, 0.2429473354807992f,
      0.2108099294331058f, 0.1759998110016148f, 0.1415334037742616f, 0.1095123294032484f, 0.0812761592574764f,
      0.0575423961814687f, 0.0385285881906738f, 0.0240694984015316f, 0.0137226170696702f, 0.0068720184783044f,
      0.0027973066609278f, 0.0007580797514921f, 0.0000000000000000f};
    static const float bq[33] = {
      0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f,
      0.0001056069805219f, 0.0059397906834944f, 0.0235374270991938f, 0.0576439176162314f, 0.1139578493851603f,
      0.1932313936022966f, 0.2808764820852269f, 0.3715629284436812f, 0.4612091687348547f, 0.5467954414732804f,
      0.6262219272258057f, 0.6981497395977654f, 0.7618575996945027f, 0.8171268253424389f, 0.8641240158900223f,
      0.9032994538662682f, 0.9352560448161740f, 0.9607227915792738f, 0.9804356127137464f, 1.0000000000000000f};
    for (int k = 0; k < 33; ++k) { vc.hyai[k] = (double)a[k]; vc.hybi[k] = (double)bq[k]; }
  } else if (nlev > 1) {
    // equal_eta
    double deta = 1.0 / nlev;
    double etac = (double)0.2f;
    double d3 = (1 - etac) * (1 - etac) * (1 - etac);
    double c1 = 2 * (etac * etac) / d3;
    double c2 = -etac * (4 + etac + etac * etac) / d3;
    double c3 = 2 * (1 + etac + etac * etac) / d3;
    double c4 = -(1 + etac) / d3;
    for (int k = 1; k <= nlev + 1; ++k) {
      double eta = (k - 1) * deta;
      vc.hybi[k - 1] = eta < etac ? 0.0 : c1 + c2 * eta + c3 * (eta * eta) + c4 * (eta * eta * eta);
      vc.hyai[k - 1] = eta - vc.hybi[k - 1];
    }
  }
  for (int k = 1; k <= nlev; ++k) {
    vc.hyam[k - 1] = 0.5 * (vc.hyai[k - 1] + vc.hyai[k]);
    vc.hybm[k - 1] = 0.5 * (vc.hybi[k - 1] + vc.hybi[k]);
  }
  for (int k = 1; k <= nlev; ++k) m.full_lev(k) = vc.hyam[k - 1] + vc.hybm[k - 1];
  for (int k = 1; k <= nlev + 1; ++k) m.half_lev(k) = vc.hyai[k - 1] + vc.hybi[k - 1];
  if (nlev > 1) for (int k = 1; k <= nlev; ++k) m.full_dlev(k) = m.half_lev(k + 1) - m.half_lev(k);
  else m.full_dlev(1) = 1;
  for (int k = 2; k <= nlev; ++k) m.half_dlev(k) = m.full_lev(k) - m.full_lev(k - 1);
  m.half_dlev(1) = m.full_lev(1) - m.half_lev(1);
  m.half_dlev(nlev + 1) = m.half_lev(nlev + 1) - m.full_lev(nlev);
}

inline double ipow(double x, int n) {  // integer power by repeated multiplication
  double r = 1; for (int i = 0; i < n; ++i) r *= x; return r;
}

// baroclinic_wave_test_mod.F90:28-170 — u, v, pt, phs, gzs only (the prognostic set)
inline void ic_baroclinic_wave(const Consts& c, const Mesh& m, const VertCoord& vc, const Lay& L, double ptop, double hyb_p0,
                               double* u, double* v, double* pt, double* phs, double* gzs) {
  const double u0 = 35.0, t0 = 288.0, gamma = 0.005, dtk = 4.8e5;
  const double eta0 = (double)0.252f, etat = (double)0.2f;
  const double lonc = c.pi / 9.0, latc = c.pi2 / 9.0, up = 1.0;
  const int nlev = m.nlev;
  (void)v;
  for (int j = 1; j <= m.nlat; ++j) for (int i = 1; i <= m.nlon; ++i) phs[L.f2(i, j)] = 1.0e5;
  for (int k = 1; k <= nlev; ++k) {
    double eta = m.full_lev(k);
    double etav = (eta - eta0) * c.pi05;
    for (int j = 1; j <= m.nlat; ++j) {
      double sin_lat = m.full_sin_lat(j), cos_lat = m.full_cos_lat(j);
      for (int i = 1; i <= m.nlon; ++i) {
        double half_lon = m.half_lon(i);
        double r = 10.0 * std::acos(std::sin(latc) * sin_lat + std::cos(latc) * cos_lat * std::cos(half_lon - lonc));
        double s2 = std::sin(2 * m.full_lat(j));
        u[L.f(i, j, k)] = u0 * std::pow(std::cos(etav), 1.5) * (s2 * s2) + up * ((r * r <= 50) ? std::exp(-(r * r)) : 0.0);
      }
    }
  }
  for (int k = 1; k <= nlev; ++k) {
    double eta = m.full_lev(k);
    double etav = (eta - eta0) * c.pi / 2.0;
    double tbar;
    if (etat <= eta && eta <= 1) tbar = t0 * std::pow(eta, c.Rd * gamma / c.g);
    else tbar = t0 * std::pow(eta, c.Rd * gamma / c.g) + dtk * ipow(etat - eta, 5);
    for (int j = 1; j <= m.nlat; ++j) {
      double sin_lat = m.full_sin_lat(j), cos_lat = m.full_cos_lat(j);
      for (int i = 1; i <= m.nlon; ++i) {
        double phs_ij = phs[L.f2(i, j)];
        double pl0 = vc.hyai[k - 1] * (hyb_p0 - ptop) + vc.hybi[k - 1] * (phs_ij - ptop) + ptop;
        double pl1 = vc.hyai[k] * (hyb_p0 - ptop) + vc.hybi[k] * (phs_ij - ptop) + ptop;
        double ph = 0.5 * (pl0 + pl1);
        double t = tbar + 3.0 / 4.0 * eta * c.pi * u0 / c.Rd * std::sin(etav) * std::sqrt(std::cos(etav)) * (
            (-2 * ipow(sin_lat, 6) * (cos_lat * cos_lat + 1.0 / 3.0) + 10.0 / 63.0) * 2 * u0 * std::pow(std::cos(etav), 1.5) +
            (8.0 / 5.0 * ipow(cos_lat, 3) * (sin_lat * sin_lat + 2.0 / 3.0) - c.pi / 4.0) * c.radius * c.omega);
        pt[L.f(i, j, k)] = t * std::pow(c.p0 / ph, c.Rd_o_cpd);
      }
    }
  }
  double etav = (1 - eta0) * c.pi / 2;
  for (int j = 1; j <= m.nlat; ++j) {
    double sin_lat = m.full_sin_lat(j), cos_lat = m.full_cos_lat(j);
    for (int i = 1; i <= m.nlon; ++i)
      gzs[L.f2(i, j)] = u0 * std::pow(std::cos(etav), 1.5) * (
          (-2 * ipow(sin_lat, 6) * (cos_lat * cos_lat + 1.0 / 3.0) + 10.0 / 63.0) * u0 * std::pow(std::cos(etav), 1.5) +
          (8.0 / 5.0 * ipow(cos_lat, 3) * (sin_lat * sin_lat + 2.0 / 3.0) - c.pi / 4.0) * c.radius * c.omega);
  }
}

// SplitMix64 -> U[0,1): fixed-seed stand-in for random_number (the reference seeds from
// the OS, held_suarez_test_mod.F90:37, so it is not reproducible run to run).
struct SplitMix64 {
  unsigned long long s;
  explicit SplitMix64(unsigned long long seed) : s(seed) {}
  double next() {
    unsigned long long z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; z ^= z >> 31;
    return (double)(z >> 11) * (1.0 / 9007199254740992.0);
  }
};

// src/tests/rossby_haurwitz_wave_3d_test_mod.F90:26-136 followed by the perturbation of
// src/tests/held_suarez_test_mod.F90:28-63 (perturb = false gives the plain 3-D RH wave).
inline void ic_held_suarez(const Consts& c, const Mesh& m, const VertCoord& vc, const Lay& L, double ptop, double hyb_p0,
                           double* u, double* v, double* pt, double* phs, double* gzs, unsigned long long seed, bool perturb) {
  const int n = 4, nlev = m.nlev;
  const double u0 = 50.0, t0 = 288.0, gamma = 0.0065, pref = 955.0e2;
  const double a = c.radius, M = u0 / (n * c.radius), g = c.g, omega = c.omega;
  for (size_t q = 0; q < L.n2_full(); ++q) gzs[q] = 0.0;
  for (int k = 1; k <= nlev; ++k) {
    for (int j = 1; j <= m.nlat; ++j) {
      const double cl = m.full_cos_lat(j), sl = m.full_sin_lat(j);
      for (int i = 1; i <= m.nlon; ++i) {
        const double lon = m.half_lon(i);
        u[L.f(i, j, k)] = a * M * cl + a * M * ipow(cl, n - 1) * std::cos(n * lon) * (n * (sl * sl) - cl * cl);
      }
    }
    for (int j = 1; j <= m.num_half_lat; ++j) {
      const double cl = m.half_cos_lat(j), sl = m.half_sin_lat(j);
      for (int i = 1; i <= m.nlon; ++i) v[L.h(i, j, k)] = -a * M * n * ipow(cl, n - 1) * sl * std::sin(n * m.full_lon(i));
    }
  }
  for (int j = 1; j <= m.nlat; ++j) {
    const double cl = m.full_cos_lat(j);
    for (int i = 1; i <= m.nlon; ++i) {
      const double lon = m.full_lon(i);
      const double a_lat = M * (2 * omega + M) / 2 * (cl * cl) + (M * M) / 4.0 * ipow(cl, 2 * n) * ((n + 1) * (cl * cl) + (2 * n * n - n - 2)) -
                           n * n * (M * M) / 2 * ipow(cl, 2 * (n - 1));
      const double b_lat = 2 * (omega + M) * M / (n + 1) / (n + 2) * ipow(cl, n) * ((n * n + 2 * n + 2) - (n + 1) * (n + 1) * (cl * cl));
      const double c_lat = (M * M) / 4.0 * ipow(cl, 2 * n) * ((n + 1) * (cl * cl) - (n + 2));
      const double phi_p = (a * a) * (a_lat + b_lat * std::cos(n * lon) + c_lat * std::cos(2 * n * lon));
      phs[L.f2(i, j)] = pref * std::pow(1 + gamma / g / t0 * phi_p, g / gamma / c.Rd);
    }
  }
  auto ph_lev = [&](int k, double ps) { return vc.hyai[k - 1] * (hyb_p0 - ptop) + vc.hybi[k - 1] * (ps - ptop) + ptop; };
  // t and pt from the UNPERTURBED phs (the reference perturbs phs first but keeps t, ph)
  for (int k = 1; k <= nlev; ++k)
    for (int j = 1; j <= m.nlat; ++j) for (int i = 1; i <= m.nlon; ++i) {
      const double ps = phs[L.f2(i, j)];
      const double ph = 0.5 * (ph_lev(k, ps) + ph_lev(k + 1, ps));
      const double t = t0 * std::pow(ph / pref, gamma * c.Rd / g);
      pt[L.f(i, j, k)] = t * std::pow(c.p0 / ph, c.Rd_o_cpd);
    }
  if (!perturb) return;
  SplitMix64 rng(seed);
  for (int j = 1; j <= m.nlat; ++j) for (int i = 1; i <= m.nlon; ++i) {
    const double cl = m.full_cos_lat(j);
    phs[L.f2(i, j)] = phs[L.f2(i, j)] - (0.5 + rng.next()) * (cl * cl);
  }
  for (int k = 1; k <= nlev; ++k)
    for (int j = 1; j <= m.nlat; ++j) for (int i = 1; i <= m.nlon; ++i) {
      const double cl = m.full_cos_lat(j);
      pt[L.f(i, j, k)] = pt[L.f(i, j, k)] - (0.5 + rng.next()) * (cl * cl);
    }
}

}  // namespace gmh

// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_core.hpp).
// Held–Suarez Rayleigh drag + Newtonian relaxation:
//   src/tests/held_suarez_test_mod.F90:65-113
#pragma once
#include "orc_model.hpp"

namespace orc {

inline void Model::held_suarez_forcing(State& s, double dt) {
  const Mesh& mesh = b.mesh;
  const double sig_b = 0.7, kf = 1.0 / 86400.0, ka = 0.025 / 86400.0, ks = 0.25 / 86400.0;
  const double dt_lat = 60.0, dpt_lev = 10.0, p0 = 1.0e5;
  A3 &u = s.u_lon, &v = s.v_lat, &pt = s.pt; const A3 &ph = s.ph, &t = s.t;
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k) {
    double kv = kf * std::fmax(0.0, (mesh.full_lev(k) - sig_b) / (1.0 - sig_b));
    for (int j = 2; j <= mesh.nlat - 1; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      u(i, j, k) = u(i, j, k) - dt * kv * u(i, j, k);
  }
  fill_halo(mesh, u, true);
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k) {
    double kv = kf * std::fmax(0.0, (mesh.full_lev(k) - sig_b) / (1.0 - sig_b));
    for (int j = 1; j <= mesh.nhalf; ++j) for (int i = 1; i <= mesh.nlon; ++i)
      v(i, j, k) = v(i, j, k) - dt * kv * v(i, j, k);
  }
  fill_halo(mesh, v, false);
  ORC_PAR for (int k = 1; k <= mesh.nlev; ++k)
    for (int j = 1; j <= mesh.nlat; ++j) {
      double c = mesh.full_cos_lat(j), sn = mesh.full_sin_lat(j);
      double c4 = (c * c) * (c * c);   // cos_lat**4 (gfortran: (x*x)*(x*x))
      double kt = ka + (ks - ka) * std::fmax(0.0, (mesh.full_lev(k) - sig_b) / (1.0 - sig_b)) * c4;
      for (int i = 1; i <= mesh.nlon; ++i) {
        double p_p0 = ph(i, j, k) / p0;
        double teq = std::fmax(200.0, (315.0 - dt_lat * (sn * sn) - dpt_lev * std::log(p_p0) * (c * c)) * std::pow(p_p0, b.cfg.Rd_o_cpd));
        pt(i, j, k) = pt(i, j, k) - dt * kt * pt(i, j, k) * (1.0 - teq / t(i, j, k));
      }
    }
  fill_halo(mesh, pt, true);
}

}  // namespace orc

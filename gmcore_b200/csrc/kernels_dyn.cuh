// Horizontal C-grid operator kernels shared by the shallow-water and hydrostatic
// cores.  One thread per (lon i, lat row j, level k); longitude is the contiguous,
// coalesced axis.  Each kernel cites the reference loops it fuses
// (src/dynamics/operators_mod.F90 unless noted).
#pragma once
#include "common.cuh"

// ---- loads with the single-process halo semantics of fill_halo -------------------
// Full-lat field beyond a pole: j<1 -> 2-j, j>nlat -> 2*nlat-j, rotated by nlon/2
// (src/utils/parallel_mod.F90:529-563, halo_mod.F90:183-238).  Rows owned by a
// neighbouring band are materialised halo rows and are read directly.
__device__ __forceinline__ double ldF(const double* __restrict__ p, const Geo& g, int i, int j, int k) {
  if (j < 1) { j = 2 - j; i += g.nlon >> 1; }
  else if (j > g.nlat) { j = 2 * g.nlat - j; i += g.nlon >> 1; }
  i = wrapi(i, g.nlon);
  return __ldg(p + gidx(g, i, j, k));
}
__device__ __forceinline__ double ldH(const double* __restrict__ p, const Geo& g, int i, int j, int k) {
  if (j < 1) { j = 1 - j; i += g.nlon >> 1; }
  else if (j > g.nhalf) { j = 2 * g.nhalf + 1 - j; i += g.nlon >> 1; }
  i = wrapi(i, g.nlon);
  return __ldg(p + gidx(g, i, j, k));
}
// plain periodic load (row known to be inside the band + halo)
__device__ __forceinline__ double ldW(const double* __restrict__ p, const Geo& g, int i, int j, int k) {
  return __ldg(p + gidx(g, wrapi(i, g.nlon), j, k));
}

// m_lon / m_lat / m_vtx are pure functions of m (interp_mod.F90:233-275,438-463);
// they are recomputed where needed instead of being stored (bit-identical).
__device__ __forceinline__ double m_lon_at(const double* __restrict__ m, const Geo& g, int i, int j, int k) {
  return (ldW(m, g, i, j, k) + ldW(m, g, i + 1, j, k)) * 0.5;
}
__device__ __forceinline__ double m_lat_at(const double* __restrict__ m, const Geo& g, int i, int j, int k) {
  return (ldW(m, g, i, j, k) + ldW(m, g, i, j + 1, k)) * 0.5;
}
__device__ __forceinline__ double m_vtx_at(const double* __restrict__ m, const Geo& g, const DMesh& M, int i, int j, int k) {
  return ((ldW(m, g, i, j, k) + ldW(m, g, i + 1, j, k)) * MJ2(M.area_subcell, 2, j) +
          (ldW(m, g, i, j + 1, k) + ldW(m, g, i + 1, j + 1, k)) * MJ2(M.area_subcell, 1, j + 1)) / MJ(M.area_vtx, j);
}
// calc_mf :566-579.  mfx_lon is never computed on pole rows (stays 0 from allocation).
__device__ __forceinline__ double mfx_at(const double* __restrict__ m, const double* __restrict__ u, const Geo& g, int i, int j, int k) {
  if (j <= 1 || j >= g.nlat) return 0.0;
  return m_lon_at(m, g, i, j, k) * ldW(u, g, i, j, k);
}
__device__ __forceinline__ double mfy_at(const double* __restrict__ m, const double* __restrict__ v, const Geo& g, int i, int j, int k) {
  return m_lat_at(m, g, i, j, k) * ldW(v, g, i, j, k);
}

struct Fwd1Args {
  const double *u, *v, *m;                       // u_f, v_f, m of the star state
  double *mfx_lon, *mfy_lat, *u_lat, *v_lon, *ke, *pv, *div;
  int ke_scheme; double ke_cell_wgt;
};

// K1: calc_mf (:545-610) + calc_ke (:234-317) + calc_div (:368-377) + calc_vor/calc_pv
// (:630-639,693-699) on all non-pole-special points.
__global__ void __launch_bounds__(128) k_fwd1(Geo g, DMesh M, Fwd1Args a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y, k = blockIdx.z;
  if (i >= g.nlon) return;
  const long long c = gidx(g, i, j, k);
  const bool pole_row = (j == 1 || j == g.nlat);
  if (!pole_row) {
    // full-lat, non-pole row: mfx_lon, v_lon, ke, div
    const double u0 = ldW(a.u, g, i, j, k), um = ldW(a.u, g, i - 1, j, k);
    const double v0 = ldW(a.v, g, i, j, k), vs = ldW(a.v, g, i, j - 1, k);
    const double mlon = m_lon_at(a.m, g, i, j, k);
    a.mfx_lon[c] = mlon * u0;
    const double mfy_lon = MJ2(M.full_tangent_wgt, 1, j) * (mfy_at(a.m, a.v, g, i, j - 1, k) + mfy_at(a.m, a.v, g, i + 1, j - 1, k)) +
                           MJ2(M.full_tangent_wgt, 2, j) * (mfy_at(a.m, a.v, g, i, j, k) + mfy_at(a.m, a.v, g, i + 1, j, k));
    a.v_lon[c] = mfy_lon / mlon;
    double ke = (MJ(M.area_lon_west, j) * (um * um) + MJ(M.area_lon_east, j) * (u0 * u0) +
                 MJ(M.area_lat_north, j - 1) * (vs * vs) + MJ(M.area_lat_south, j) * (v0 * v0)) / MJ(M.area_cell, j);
    if (a.ke_scheme == 2) {
      const double vm = ldW(a.v, g, i - 1, j, k), vp = ldW(a.v, g, i + 1, j, k);
      const double vsm = ldW(a.v, g, i - 1, j - 1, k), vsp = ldW(a.v, g, i + 1, j - 1, k);
      const double umn = ldW(a.u, g, i - 1, j + 1, k), u0n = ldW(a.u, g, i, j + 1, k);
      const double ums = ldW(a.u, g, i - 1, j - 1, k), u0s = ldW(a.u, g, i, j - 1, k);
      const double kv1 = (MJ(M.area_lat_east, j) * (vm * vm) + MJ(M.area_lat_west, j) * (v0 * v0) +
                          MJ(M.area_lon_north, j) * (um * um) + MJ(M.area_lon_south, j + 1) * (umn * umn)) / MJ(M.area_vtx, j);
      const double kv2 = (MJ(M.area_lat_east, j - 1) * (vsm * vsm) + MJ(M.area_lat_west, j - 1) * (vs * vs) +
                          MJ(M.area_lon_north, j - 1) * (ums * ums) + MJ(M.area_lon_south, j) * (um * um)) / MJ(M.area_vtx, j - 1);
      const double kv3 = (MJ(M.area_lat_east, j - 1) * (vs * vs) + MJ(M.area_lat_west, j - 1) * (vsp * vsp) +
                          MJ(M.area_lon_north, j - 1) * (u0s * u0s) + MJ(M.area_lon_south, j) * (u0 * u0)) / MJ(M.area_vtx, j - 1);
      const double kv4 = (MJ(M.area_lat_east, j) * (v0 * v0) + MJ(M.area_lat_west, j) * (vp * vp) +
                          MJ(M.area_lon_north, j) * (u0 * u0) + MJ(M.area_lon_south, j + 1) * (u0n * u0n)) / MJ(M.area_vtx, j);
      ke = (1.0 - a.ke_cell_wgt) * ((kv1 + kv4) * MJ2(M.area_subcell, 2, j) + (kv2 + kv3) * MJ2(M.area_subcell, 1, j)) / MJ(M.area_cell, j) +
           a.ke_cell_wgt * ke;
    }
    a.ke[c] = ke;
    a.div[c] = ((u0 * MJ(M.le_lon, j) - um * MJ(M.le_lon, j)) + (v0 * MJ(M.le_lat, j) - vs * MJ(M.le_lat, j - 1))) / MJ(M.area_cell, j);
  }
  if (j <= g.nhalf) {
    // half-lat row: mfy_lat, u_lat, pv
    const double mlat = m_lat_at(a.m, g, i, j, k);
    const double v0 = ldW(a.v, g, i, j, k);
    a.mfy_lat[c] = mlat * v0;
    const double mfx_lat = MJ2(M.half_tangent_wgt, 1, j) * (mfx_at(a.m, a.u, g, i - 1, j, k) + mfx_at(a.m, a.u, g, i, j, k)) +
                           MJ2(M.half_tangent_wgt, 2, j) * (mfx_at(a.m, a.u, g, i - 1, j + 1, k) + mfx_at(a.m, a.u, g, i, j + 1, k));
    a.u_lat[c] = mfx_lat / mlat;
    const double vor = (ldW(a.u, g, i, j, k) * MJ(M.de_lon, j) - ldW(a.u, g, i, j + 1, k) * MJ(M.de_lon, j + 1) +
                        ldW(a.v, g, i + 1, j, k) * MJ(M.de_lat, j) - v0 * MJ(M.de_lat, j)) / MJ(M.area_vtx, j);
    a.pv[c] = (vor + MJ(M.half_f, j)) / m_vtx_at(a.m, g, M, i, j, k);
  }
}

// Pole rows: zonal means replace the stencils (calc_ke :320-349, calc_div :378-407,
// calc_vor :640-674 with pv_pole_stokes, calc_grad_mf :974-1003).  One block per
// (pole, level); the zonal circle is band-local so no collective is needed
// (zonal_sum, src/utils/parallel_mod.F90:741-764).
struct Pole1Args {
  const double *v, *m, *u_lat, *mfy_lat;
  double *ke, *div, *pv;
  double *pole_dmf;        // [2][nlev]: dmfdlat on the south / north pole row
  int pv_pole_stokes;
};
__global__ void __launch_bounds__(256) k_pole1(Geo g, DMesh M, Pole1Args a) {
  __shared__ double sm[32];
  const int k = blockIdx.x, north = blockIdx.y;
  if ((north && !g.has_north) || (!north && !g.has_south)) return;
  const int j = north ? g.nlat : 1;          // pole full row
  const int jh = north ? g.nhalf : 1;        // adjacent half row
  double s_v2 = 0, s_v = 0, s_ul = 0, s_mf = 0;
  for (int i = threadIdx.x; i < g.nlon; i += blockDim.x) {
    const double v = __ldg(a.v + gidx(g, i, jh, k));
    s_v2 += v * v;
    s_v += north ? -v : v;
    const double ul = __ldg(a.u_lat + gidx(g, i, jh, k)) * MJ(M.le_lat, jh);
    s_ul += north ? ul : -ul;
    const double mf = __ldg(a.mfy_lat + gidx(g, i, jh, k));
    s_mf += north ? -mf : mf;
  }
  s_v2 = block_sum(s_v2, sm); s_v = block_sum(s_v, sm); s_ul = block_sum(s_ul, sm); s_mf = block_sum(s_mf, sm);
  const double ke = s_v2 / g.nlon;
  const double dv = s_v * MJ(M.le_lat, jh) / g.nlon / MJ(M.area_cell, j);
  const double vor = s_ul / g.nlon / MJ(M.area_cell, j);
  if (threadIdx.x == 0) a.pole_dmf[north * g.nlev + k] = s_mf * MJ(M.le_lat, jh) / g.nlon / MJ(M.area_cell, j);
  for (int i = threadIdx.x; i < g.nlon; i += blockDim.x) {
    a.ke[gidx(g, i, j, k)] = ke;
    a.div[gidx(g, i, j, k)] = dv;
    if (a.pv_pole_stokes) a.pv[gidx(g, i, jh, k)] = (vor + MJ(M.half_f, jh)) / m_vtx_at(a.m, g, M, i, jh, k);
  }
}

__device__ __forceinline__ double upwind1(double dir, double wgt, double f1, double f2) {
  return 0.5 * (f2 + f1) + (-0.5) * (f2 - f1) * wgt * dir;
}
__device__ __forceinline__ double upwind3(double dir, double wgt, double f1, double f2, double f3, double f4) {
  const double c31 = 7.0 / 12.0, c32 = -1.0 / 12.0, c33 = 1.0 / 12.0;
  return c31 * (f3 + f2) + c32 * (f4 + f1) + c33 * (f4 - f1 - 3 * (f3 - f2)) * wgt * dir;
}

// K2: interp_pv_upwind / interp_pv_midpoint (:705-795)
struct PvEdgeArgs {
  const double *pv, *u, *v, *u_lat, *v_lon;
  double *pv_lon, *pv_lat;
  int scheme, order; double wgt;
};
__global__ void __launch_bounds__(128) k_pv_edge(Geo g, PvEdgeArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y, k = blockIdx.z;
  if (i >= g.nlon) return;
  const long long c = gidx(g, i, j, k);
  const double eps = 2.220446049250313e-16;
  if (j > 1 && j < g.nlat) {
    const double pm1 = ldH(a.pv, g, i, j - 1, k), p0 = ldH(a.pv, g, i, j, k);
    if (a.scheme == 0) {
      a.pv_lon[c] = 0.5 * (p0 + pm1);
    } else {
      const double vt = __ldg(a.v_lon + c), un = __ldg(a.u + c);
      const double b = fabs(vt) / (sqrt(un * un + vt * vt) + eps);
      const double up = a.order == 1 ? upwind1(sign1(vt), a.wgt, pm1, p0)
                                     : upwind3(sign1(vt), a.wgt, ldH(a.pv, g, i, j - 2, k), pm1, p0, ldH(a.pv, g, i, j + 1, k));
      a.pv_lon[c] = b * up + (1 - b) * 0.5 * (pm1 + p0);
    }
  }
  if (j <= g.nhalf) {
    const double pm1 = ldW(a.pv, g, i - 1, j, k), p0 = __ldg(a.pv + c);
    if (a.scheme == 0) {
      a.pv_lat[c] = 0.5 * (pm1 + p0);
    } else {
      const double ut = __ldg(a.u_lat + c), vn = __ldg(a.v + c);
      const double b = fabs(ut) / (sqrt(ut * ut + vn * vn) + eps);
      const double up = a.order == 1 ? upwind1(sign1(ut), a.wgt, pm1, p0)
                                     : upwind3(sign1(ut), a.wgt, ldW(a.pv, g, i - 2, j, k), pm1, p0, ldW(a.pv, g, i + 1, j, k));
      a.pv_lat[c] = b * up + (1 - b) * 0.5 * (pm1 + p0);
    }
  }
}

// K3: calc_coriolis (:846-906) + calc_grad_ke (:908-937) [+ vertical advection terms
// passed in for the hydrostatic core] -> forward-pass du, dv (gmcore_mod.F90:478-489,
// 507-518); shallow water also forms dgz (:520-524) and the raw gz update
// (time_schemes_mod.F90:179-185).
struct TendArgs {
  const double *mfx_lon, *mfy_lat, *pv_lon, *pv_lat, *ke, *u_lat_unused;
  const double *we_lev, *u_f, *v_f, *m;   // hydrostatic: vertical advection of u, v formed inline; NULL for shallow water
  double *du, *dv;
  // shallow water only
  const double *gz_old, *pole_dmf; double *gz_new; double dt, grav;
  int coriolis_scheme, swm;
};
__global__ void __launch_bounds__(128) k_tend(Geo g, DMesh M, TendArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y, k = blockIdx.z;
  if (i >= g.nlon) return;
  const long long c = gidx(g, i, j, k);
  const bool pole_row = (j == 1 || j == g.nlat);
  if (!pole_row) {
    const double pl = __ldg(a.pv_lon + c);
    const double qhv = (MJ2(M.full_tangent_wgt, 1, j) * (ldW(a.mfy_lat, g, i, j - 1, k) * (pl + ldW(a.pv_lat, g, i, j - 1, k)) +
                                                          ldW(a.mfy_lat, g, i + 1, j - 1, k) * (pl + ldW(a.pv_lat, g, i + 1, j - 1, k))) +
                        MJ2(M.full_tangent_wgt, 2, j) * (ldW(a.mfy_lat, g, i, j, k) * (pl + ldW(a.pv_lat, g, i, j, k)) +
                                                          ldW(a.mfy_lat, g, i + 1, j, k) * (pl + ldW(a.pv_lat, g, i + 1, j, k)))) * 0.5;
    const double ke0 = __ldg(a.ke + c);
    const double dkedlon = (ldW(a.ke, g, i + 1, j, k) - ke0) / MJ(M.de_lon, j);
    double du = qhv - dkedlon;
    if (a.we_lev) {
      // calc_wedudlev_wedvdlev (:1145-1180) with we_lev_lon from interp_mod.F90:361-369
      const int kb = 0, ke = g.nlev - 1;
      const double aw = MJ(M.area_lon_west, j), ae = MJ(M.area_lon_east, j), al = MJ(M.area_lon, j);
      const double mlon = m_lon_at(a.m, g, i, j, k);
      const double u0 = __ldg(a.u_f + c);
      double w;
      if (k > kb && k < ke) {
        const double wl1 = (aw * ldW(a.we_lev, g, i, j, k + 1) + ae * ldW(a.we_lev, g, i + 1, j, k + 1)) / al;
        const double wl0 = (aw * ldW(a.we_lev, g, i, j, k) + ae * ldW(a.we_lev, g, i + 1, j, k)) / al;
        w = (wl1 * (ldW(a.u_f, g, i, j, k + 1) - u0) + wl0 * (u0 - ldW(a.u_f, g, i, j, k - 1))) / mlon / 2.0;
      } else if (k == kb) {
        const double wl1 = (aw * ldW(a.we_lev, g, i, j, k + 1) + ae * ldW(a.we_lev, g, i + 1, j, k + 1)) / al;
        w = (wl1 * (ldW(a.u_f, g, i, j, k + 1) - u0)) / mlon / 2.0;
      } else {
        const double wl0 = (aw * ldW(a.we_lev, g, i, j, k) + ae * ldW(a.we_lev, g, i + 1, j, k)) / al;
        w = (wl0 * (u0 - ldW(a.u_f, g, i, j, k - 1))) / mlon / 2.0;
      }
      du = du - w;
    }
    a.du[c] = du;
    if (a.swm) {
      const double dmfdlon = (__ldg(a.mfx_lon + c) - ldW(a.mfx_lon, g, i - 1, j, k)) * MJ(M.le_lon, j) / MJ(M.area_cell, j);
      const double dmfdlat = (__ldg(a.mfy_lat + c) * MJ(M.le_lat, j) - ldW(a.mfy_lat, g, i, j - 1, k) * MJ(M.le_lat, j - 1)) / MJ(M.area_cell, j);
      const double dgz = -(dmfdlon + dmfdlat) * a.grav;
      a.gz_new[c] = __ldg(a.gz_old + c) + a.dt * dgz;
    }
  } else if (a.swm) {
    const double dgz = -(0.0 + __ldg(a.pole_dmf + (j == 1 ? 0 : g.nlev) + k)) * a.grav;
    a.gz_new[c] = __ldg(a.gz_old + c) + a.dt * dgz;
  }
  if (j <= g.nhalf) {
    const double pt = __ldg(a.pv_lat + c);
    // mfx_lon on pole rows is 0 (never computed by the reference)
    const bool s0 = j > 1, n0 = j + 1 < g.nlat;
    const double mx_a = s0 ? ldW(a.mfx_lon, g, i - 1, j, k) : 0.0, mx_b = s0 ? __ldg(a.mfx_lon + c) : 0.0;
    const double mx_c = n0 ? ldW(a.mfx_lon, g, i - 1, j + 1, k) : 0.0, mx_d = n0 ? ldW(a.mfx_lon, g, i, j + 1, k) : 0.0;
    const double pl_a = s0 ? ldW(a.pv_lon, g, i - 1, j, k) : 0.0, pl_b = s0 ? __ldg(a.pv_lon + c) : 0.0;
    const double pl_c = n0 ? ldW(a.pv_lon, g, i - 1, j + 1, k) : 0.0, pl_d = n0 ? ldW(a.pv_lon, g, i, j + 1, k) : 0.0;
    const double qhu = (MJ2(M.half_tangent_wgt, 1, j) * (mx_a * (pt + pl_a) + mx_b * (pt + pl_b)) +
                        MJ2(M.half_tangent_wgt, 2, j) * (mx_c * (pt + pl_c) + mx_d * (pt + pl_d))) * 0.5;
    const double dkedlat = (ldW(a.ke, g, i, j + 1, k) - __ldg(a.ke + c)) / MJ(M.de_lat, j);
    double dv = -qhu - dkedlat;
    if (a.we_lev) {
      // :1181-1214 with we_lev_lat from interp_mod.F90:426-434
      const int kb = 0, ke = g.nlev - 1;
      const double an = MJ(M.area_lat_north, j), as = MJ(M.area_lat_south, j), al = MJ(M.area_lat, j);
      const double mlat = m_lat_at(a.m, g, i, j, k);
      const double v0 = __ldg(a.v_f + c);
      double w;
      if (k > kb && k < ke) {
        const double wl1 = (an * ldW(a.we_lev, g, i, j + 1, k + 1) + as * ldW(a.we_lev, g, i, j, k + 1)) / al;
        const double wl0 = (an * ldW(a.we_lev, g, i, j + 1, k) + as * ldW(a.we_lev, g, i, j, k)) / al;
        w = (wl1 * (ldW(a.v_f, g, i, j, k + 1) - v0) + wl0 * (v0 - ldW(a.v_f, g, i, j, k - 1))) / mlat / 2.0;
      } else if (k == kb) {
        const double wl1 = (an * ldW(a.we_lev, g, i, j + 1, k + 1) + as * ldW(a.we_lev, g, i, j, k + 1)) / al;
        w = (wl1 * (ldW(a.v_f, g, i, j, k + 1) - v0)) / mlat / 2.0;
      } else {
        const double wl0 = (an * ldW(a.we_lev, g, i, j + 1, k) + as * ldW(a.we_lev, g, i, j, k)) / al;
        w = (wl0 * (v0 - ldW(a.v_f, g, i, j, k - 1))) / mlat / 2.0;
      }
      dv = dv - w;
    }
    a.dv[c] = dv;
  }
}

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

// Blocks whose i-1 .. i+1 (or +-2) neighbourhood stays inside the row skip the periodic wrap: the
// neighbour offsets become compile-time immediates off one base address per (array, row).
template <bool EDGE> __device__ __forceinline__ int wrapE(int i, int n) { return EDGE ? wrapi(i, n) : i; }
__device__ __forceinline__ bool edge_block(int n, int reach) {
  return blockIdx.x == 0 || (int)((blockIdx.x + 1) * blockDim.x) + reach > n;
}
// L2 prefetch of the line holding *p (no register is held while the line travels): the level-marching
// kernels ask for a few levels ahead of the one they work on.  Measured at 0.25 deg L32: k_fwd1 (3 input
// streams) 0.66 -> 0.55 ms with 2-3 levels ahead; k_tend (9 input streams) gets slower with it (the
// prefetch instructions cost more issue slots than they hide).
#ifndef FWD1_PF
#define FWD1_PF 3
#endif
#ifndef TEND_PF
#define TEND_PF 0
#endif
__device__ __forceinline__ void pf_l2(const double* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
// a / b for a field divisor: correctly rounded reciprocal then one multiply (<= 1 ulp off the IEEE
// divide, about half its instruction count).
__device__ __forceinline__ double qdiv(double a, double b) { return a * __drcp_rn(b); }

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
#ifndef FWD1_MINB
#define FWD1_MINB 4
#endif
// One thread per (i, j) marching down the levels: the ~30 row coefficients and the neighbour
// offsets are formed once per column instead of once per point (no vertical coupling here -- the
// loop only amortises the per-row work and keeps 32 levels of independent loads behind one thread).
template <bool EDGE> __device__ __forceinline__ void fwd1_col(const Geo& g, const DMesh& M, const Fwd1Args& a, int i, int j) {
  const int n = g.nlon, pl = (int)g.plane;
  const int dm = wrapE<EDGE>(i - 1, n) - i, dp = wrapE<EDGE>(i + 1, n) - i;   // -1 / +1 away from the row ends
  const bool full = j > 1 && j < g.nlat;     // non-pole full-lat row (warp-uniform)
  const bool half = j <= g.nhalf, n0 = j + 1 < g.nlat;
  const bool ke2 = a.ke_scheme == 2;
  const RowCoef* __restrict__ R = M.rows + (j + 1);
  const double rac = R->r_area_cell, ftw1 = R->ftw1, ftw2 = R->ftw2, htw1 = R->htw1, htw2 = R->htw2;
  const double alw = R->area_lon_west, ale = R->area_lon_east, atn_s = R[-1].area_lat_north, ats = R->area_lat_south;
  const double rv = R->r_area_vtx, rvs = R[-1].r_area_vtx;
  const double ae = R->area_lat_east, aw = R->area_lat_west, aes = R[-1].area_lat_east, aws = R[-1].area_lat_west;
  const double an = R->area_lon_north, asn = R[1].area_lon_south, ans = R[-1].area_lon_north, as0 = R->area_lon_south;
  const double sub1 = R->sub1, sub2 = R->sub2, sub1n = R[1].sub1;
  const double lel = R->le_lon, lea = R->le_lat, leas = R[-1].le_lat;
  const double del = R->de_lon, deln = R[1].de_lon, dea = R->de_lat, hf = R->half_f;
  const double kw = a.ke_cell_wgt;
  int c = gidx(g, i, j, 0);
#pragma unroll 1
  for (int k = 0; k < g.nlev; ++k, c += pl) {
    // one memory phase: the 20 distinct neighbours of u, v, m (rows j-1 and j+1 always exist on the
    // device: >= 2 halo rows); everything below is register arithmetic in the reference's order
    const double* __restrict__ ur = a.u + c; const double* __restrict__ vr = a.v + c; const double* __restrict__ mr = a.m + c;
    if (FWD1_PF && (i & 15) == 0 && k + FWD1_PF < g.nlev) { pf_l2(ur + FWD1_PF * pl); pf_l2(vr + FWD1_PF * pl); pf_l2(mr + FWD1_PF * pl); }
    const double u_m = __ldg(ur + dm), u_0 = __ldg(ur);
    const double u_mn = __ldg(ur + n + dm), u_0n = __ldg(ur + n);
    const double v_0 = __ldg(vr), v_p = __ldg(vr + dp);
    const double m_m = __ldg(mr + dm), m_0 = __ldg(mr), m_p = __ldg(mr + dp);
    const double m_mn = __ldg(mr + n + dm), m_0n = __ldg(mr + n), m_pn = __ldg(mr + n + dp);
    double u_ms = 0, u_0s = 0, v_m = 0, v_ms = 0, v_0s = 0, v_ps = 0, m_0s = 0, m_ps = 0;
    if (full) {
      u_ms = __ldg(ur - n + dm); u_0s = __ldg(ur - n);
      v_m = __ldg(vr + dm); v_ms = __ldg(vr - n + dm); v_0s = __ldg(vr - n); v_ps = __ldg(vr - n + dp);
      m_0s = __ldg(mr - n); m_ps = __ldg(mr - n + dp);
    }
    const double mlon = (m_0 + m_p) * 0.5;            // m_lon(i,j)
    const double mlat = (m_0 + m_0n) * 0.5;           // m_lat(i,j)
    if (full) {
      // full-lat, non-pole row: mfx_lon, v_lon, ke, div
      a.mfx_lon[c] = mlon * u_0;
      const double mfy_lon = ftw1 * (((m_0s + m_0) * 0.5) * v_0s + ((m_ps + m_p) * 0.5) * v_ps) +
                             ftw2 * (mlat * v_0 + ((m_p + m_pn) * 0.5) * v_p);
      a.v_lon[c] = qdiv(mfy_lon, mlon);
      double ke = (alw * (u_m * u_m) + ale * (u_0 * u_0) + atn_s * (v_0s * v_0s) + ats * (v_0 * v_0)) * rac;
      if (ke2) {
        const double kv1 = (ae * (v_m * v_m) + aw * (v_0 * v_0) + an * (u_m * u_m) + asn * (u_mn * u_mn)) * rv;
        const double kv2 = (aes * (v_ms * v_ms) + aws * (v_0s * v_0s) + ans * (u_ms * u_ms) + as0 * (u_m * u_m)) * rvs;
        const double kv3 = (aes * (v_0s * v_0s) + aws * (v_ps * v_ps) + ans * (u_0s * u_0s) + as0 * (u_0 * u_0)) * rvs;
        const double kv4 = (ae * (v_0 * v_0) + aw * (v_p * v_p) + an * (u_0 * u_0) + asn * (u_0n * u_0n)) * rv;
        ke = (1.0 - kw) * ((kv1 + kv4) * sub2 + (kv2 + kv3) * sub1) * rac + kw * ke;
      }
      a.ke[c] = ke;
      a.div[c] = ((u_0 * lel - u_m * lel) + (v_0 * lea - v_0s * leas)) * rac;
    }
    if (half) {
      // half-lat row: mfy_lat, u_lat, pv.  mfx_lon is 0 on pole rows (never computed by the reference).
      a.mfy_lat[c] = mlat * v_0;
      const double fx_m = full ? ((m_m + m_0) * 0.5) * u_m : 0.0, fx_0 = full ? mlon * u_0 : 0.0;
      const double fx_mn = n0 ? ((m_mn + m_0n) * 0.5) * u_mn : 0.0, fx_0n = n0 ? ((m_0n + m_pn) * 0.5) * u_0n : 0.0;
      const double mfx_lat = htw1 * (fx_m + fx_0) + htw2 * (fx_mn + fx_0n);
      a.u_lat[c] = qdiv(mfx_lat, mlat);
      const double vor = (u_0 * del - u_0n * deln + v_p * dea - v_0 * dea) * rv;
      const double mvtx = ((m_0 + m_p) * sub2 + (m_0n + m_pn) * sub1n) * rv;
      a.pv[c] = qdiv(vor + hf, mvtx);
    }
  }
}
__global__ void __launch_bounds__(128, FWD1_MINB) k_fwd1(Geo g, DMesh M, Fwd1Args a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y * blockDim.y + threadIdx.y;
  if (i >= g.nlon || j > g.je) return;
  if (edge_block(g.nlon, 1)) fwd1_col<true>(g, M, a, i, j); else fwd1_col<false>(g, M, a, i, j);
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

// adv/tvd_mod.F90:42-85 (eps = default-real 1.0e-6 widened, :50); lim = tvd_limiter_type
__device__ __forceinline__ double tvd(int lim, double cfl, double fm1, double f, double fp1, double fp2) {
  const double eps = (double)1.0e-6f;
  const double r = qdiv((cfl > 0 ? f - fm1 : fp2 - fp1) + eps, fp1 - f + eps);
  double C = 1.0;
  if (lim == GMC_TVD_MC) C = fmax(0.0, fmin(fmin(2.0 * r, 0.5 * (1 + r)), 2.0));
  else if (lim == GMC_TVD_VAN_LEER) C = qdiv(r + fabs(r), 1 + fabs(r));
  return 0.5 * (fp1 + f - ((1 - C) * sign1(cfl) + cfl * C) * (fp1 - f));
}

// K2: interp_pv_upwind / interp_pv_midpoint / interp_pv_tvd (:705-832)
struct PvEdgeArgs {
  const double *pv, *u, *v, *u_lat, *v_lon;
  double *pv_lon, *pv_lat;
  int scheme, order; double wgt;
  const RowCoef* rows; double dt; int tvd_lim;     // pv_scheme = tvd only
};
template <bool EDGE> __device__ __forceinline__ void pv_edge_body(const Geo& g, const PvEdgeArgs& a, int i, int j, int k) {
  const int n = g.nlon, c = gidx(g, i, j, k);
  const double eps = 2.220446049250313e-16;
  const bool full = j > 1 && j < g.nlat, half = j <= g.nhalf;      // warp-uniform
  const bool tv = a.scheme == GMC_PV_TVD, up = a.scheme != 0, o3 = a.order != 1 || tv;
  const double* __restrict__ pv = a.pv + c;
  // one memory phase: the pv cross (4 rows x 4 columns of it) and the four wind components
  double pm1 = 0, p0 = 0, pm2 = 0, pp1 = 0, vt = 0, un = 0;
  if (full) {
    if (j >= 3 && j + 1 <= g.nhalf) { pm1 = __ldg(pv - n); p0 = __ldg(pv); if (up && o3) { pm2 = __ldg(pv - 2 * n); pp1 = __ldg(pv + n); } }
    else { pm1 = ldH(a.pv, g, i, j - 1, k); p0 = ldH(a.pv, g, i, j, k); if (up && o3) { pm2 = ldH(a.pv, g, i, j - 2, k); pp1 = ldH(a.pv, g, i, j + 1, k); } }
    if (up) { vt = __ldg(a.v_lon + c); un = __ldg(a.u + c); }
  }
  double qm1 = 0, q0 = 0, qm2 = 0, qp1 = 0, ut = 0, vn = 0;
  if (half) {
    qm1 = __ldg(pv + (wrapE<EDGE>(i - 1, n) - i)); q0 = full && j >= 3 && j + 1 <= g.nhalf ? p0 : __ldg(pv);
    if (up && o3) { qm2 = __ldg(pv + (wrapE<EDGE>(i - 2, n) - i)); qp1 = __ldg(pv + (wrapE<EDGE>(i + 1, n) - i)); }
    if (up) { ut = __ldg(a.u_lat + c); vn = __ldg(a.v + c); }
  }
  if (full) {
    if (!up) a.pv_lon[c] = 0.5 * (p0 + pm1);
    else if (tv) a.pv_lon[c] = tvd(a.tvd_lim, qdiv(vt * a.dt, a.rows[j + 1].le_lon), pm2, pm1, p0, pp1);
    else {
      const double b = qdiv(fabs(vt), sqrt(un * un + vt * vt) + eps);
      const double upw = o3 ? upwind3(sign1(vt), a.wgt, pm2, pm1, p0, pp1) : upwind1(sign1(vt), a.wgt, pm1, p0);
      a.pv_lon[c] = b * upw + (1 - b) * 0.5 * (pm1 + p0);
    }
  }
  if (half) {
    if (!up) a.pv_lat[c] = 0.5 * (qm1 + q0);
    else if (tv) a.pv_lat[c] = tvd(a.tvd_lim, qdiv(ut * a.dt, a.rows[j + 1].le_lat), qm2, qm1, q0, qp1);
    else {
      const double b = qdiv(fabs(ut), sqrt(ut * ut + vn * vn) + eps);
      const double upw = o3 ? upwind3(sign1(ut), a.wgt, qm2, qm1, q0, qp1) : upwind1(sign1(ut), a.wgt, qm1, q0);
      a.pv_lat[c] = b * upw + (1 - b) * 0.5 * (qm1 + q0);
    }
  }
}
__global__ void __launch_bounds__(128) k_pv_edge(Geo g, PvEdgeArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y * blockDim.y + threadIdx.y, k = blockIdx.z;
  if (i >= g.nlon || j > g.je) return;
  if (edge_block(g.nlon, 2) || blockIdx.x * blockDim.x < 2) pv_edge_body<true>(g, a, i, j, k); else pv_edge_body<false>(g, a, i, j, k);
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
#ifndef TEND_MINB
#define TEND_MINB 4
#endif
// One thread per (i, j) column edge pair, marching down the levels: the row coefficients and the
// neighbour offsets are formed once per column, and the vertical-advection operands (u_f, v_f at
// k-1, k, k+1 and we_lev at k, k+1 on three columns) roll through registers, so each is read from
// memory once instead of two or three times by different level-blocks.
template <bool EDGE> __device__ __forceinline__ void tend_col(const Geo& g, const DMesh& M, const TendArgs& a, int i, int j) {
  const int n = g.nlon, pl = (int)g.plane, nlev = g.nlev;
  const int dm = wrapE<EDGE>(i - 1, n) - i, dp = wrapE<EDGE>(i + 1, n) - i;
  const bool full = j > 1 && j < g.nlat, half = j <= g.nhalf;     // warp-uniform
  const bool hyd = a.we_lev != nullptr;
  const bool s0 = j > 1, n0 = j + 1 < g.nlat;
  const RowCoef* __restrict__ R = M.rows + (j + 1);
  const double ftw1 = R->ftw1, ftw2 = R->ftw2, htw1 = R->htw1, htw2 = R->htw2, rdl = R->r_de_lon, rda = R->r_de_lat;
  const double aw = R->area_lon_west, ae = R->area_lon_east, ralo = R->r_area_lon;
  const double an = R->area_lat_north, as = R->area_lat_south, rala = R->r_area_lat;
  int c = gidx(g, i, j, 0);
  double u0 = 0, u_up = 0, u_dn = 0, v0 = 0, v_up = 0, v_dn = 0;
  double w_0 = 0, w_p = 0, w_n = 0, w1_0 = 0, w1_p = 0, w1_n = 0;      // we_lev(k) is 0 on the top interface
  if (hyd) {
    if (full) u_dn = __ldg(a.u_f + c);
    if (half) v_dn = __ldg(a.v_f + c);
  }
#pragma unroll 1
  for (int k = 0; k < nlev; ++k, c += pl) {
    // ---- this level's loads, all up front ---------------------------------------------------
    const double* __restrict__ fy = a.mfy_lat + c; const double* __restrict__ fx = a.mfx_lon + c;
    const double* __restrict__ pa = a.pv_lat + c;  const double* __restrict__ po = a.pv_lon + c;
    const double* __restrict__ ker = a.ke + c;
    if (TEND_PF && (i & 15) == 0 && k + TEND_PF < nlev) {
      pf_l2(fy + TEND_PF * pl); pf_l2(fx + TEND_PF * pl); pf_l2(pa + TEND_PF * pl); pf_l2(po + TEND_PF * pl); pf_l2(ker + TEND_PF * pl);
      if (hyd) { pf_l2(a.m + c + TEND_PF * pl); pf_l2(a.we_lev + c + (TEND_PF + 1) * pl); if (k + TEND_PF + 1 < nlev) { pf_l2(a.u_f + c + (TEND_PF + 1) * pl); pf_l2(a.v_f + c + (TEND_PF + 1) * pl); } }
    }
    const double ke0 = __ldg(ker);
    const double plo_0 = __ldg(po), pla_0 = half ? __ldg(pa) : 0.0;
    double fy_0 = 0, fy_p = 0, fy_0s = 0, fy_ps = 0, pla_p = 0, pla_0s = 0, pla_ps = 0, ke_p = 0;
    if (full) {
      fy_0 = __ldg(fy); fy_p = __ldg(fy + dp); fy_0s = __ldg(fy - n); fy_ps = __ldg(fy - n + dp);
      pla_p = __ldg(pa + dp); pla_0s = __ldg(pa - n); pla_ps = __ldg(pa - n + dp);
      ke_p = __ldg(ker + dp);
    }
    double fx_m = 0, fx_0 = 0, fx_mn = 0, fx_0n = 0, plo_m = 0, plo_mn = 0, plo_0n = 0, ke_n = 0;
    if (half) {
      if (s0) { fx_m = __ldg(fx + dm); fx_0 = __ldg(fx); plo_m = __ldg(po + dm); }
      if (n0) { fx_mn = __ldg(fx + n + dm); fx_0n = __ldg(fx + n); plo_mn = __ldg(po + n + dm); plo_0n = __ldg(po + n); }
      ke_n = __ldg(ker + n);
    }
    double m_0 = 0, m_p = 0, m_n = 0;
    if (hyd) {
      const double* __restrict__ mr = a.m + c; const double* __restrict__ wr = a.we_lev + c + pl;   // interface k+1
      const bool kdn = k < nlev - 1;
      m_0 = __ldg(mr);
      u_up = u0; u0 = u_dn; v_up = v0; v0 = v_dn;
      w_0 = w1_0; w_p = w1_p; w_n = w1_n;
      w1_0 = __ldg(wr);
      if (full) { m_p = __ldg(mr + dp); w1_p = __ldg(wr + dp); if (kdn) u_dn = __ldg(a.u_f + c + pl); }
      if (half) { m_n = __ldg(mr + n); w1_n = __ldg(wr + n); if (kdn) v_dn = __ldg(a.v_f + c + pl); }
    }
    // ---- du (lon edges of full-lat rows) -------------------------------------------------------
    if (full) {
      // coriolis_scheme 1: energy-conserving 4-point form; 2: tangential mass flux (calc_mf :597-604) x edge pv
      const double qhv = a.coriolis_scheme == 1
          ? (ftw1 * (fy_0s * (plo_0 + pla_0s) + fy_ps * (plo_0 + pla_ps)) + ftw2 * (fy_0 * (plo_0 + pla_0) + fy_p * (plo_0 + pla_p))) * 0.5
          : (ftw1 * (fy_0s + fy_ps) + ftw2 * (fy_0 + fy_p)) * plo_0;
      const double dkedlon = (ke_p - ke0) * rdl;
      double du = qhv - dkedlon;
      if (hyd) {
        // calc_wedudlev_wedvdlev (:1145-1180) with we_lev_lon from interp_mod.F90:361-369
        const double mlon = (m_0 + m_p) * 0.5;
        const double wl1 = (aw * w1_0 + ae * w1_p) * ralo, wl0 = (aw * w_0 + ae * w_p) * ralo;
        double w;
        if (k > 0 && k < nlev - 1) w = wl1 * (u_dn - u0) + wl0 * (u0 - u_up);
        else if (k == 0) w = wl1 * (u_dn - u0);
        else w = wl0 * (u0 - u_up);
        du = du - qdiv(w, mlon) * 0.5;
      }
      a.du[c] = du;
      if (a.swm) {
        const double rac = R->r_area_cell;
        const double dmfdlon = (__ldg(fx) - __ldg(fx + dm)) * R->le_lon * rac;
        const double dmfdlat = (fy_0 * R->le_lat - fy_0s * R[-1].le_lat) * rac;
        const double dgz = -(dmfdlon + dmfdlat) * a.grav;
        a.gz_new[c] = __ldg(a.gz_old + c) + a.dt * dgz;
      }
    } else if (a.swm) {
      const double dgz = -(0.0 + __ldg(a.pole_dmf + (j == 1 ? 0 : nlev) + k)) * a.grav;
      a.gz_new[c] = __ldg(a.gz_old + c) + a.dt * dgz;
    }
    // ---- dv (lat edges) -------------------------------------------------------------------------
    if (half) {
      // mfx_lon / pv_lon terms on pole rows are 0 (never computed by the reference)
      const double pl_b = s0 ? plo_0 : 0.0;
      const double qhu = a.coriolis_scheme == 1
          ? (htw1 * (fx_m * (pla_0 + plo_m) + fx_0 * (pla_0 + pl_b)) + htw2 * (fx_mn * (pla_0 + plo_mn) + fx_0n * (pla_0 + plo_0n))) * 0.5
          : (htw1 * (fx_m + fx_0) + htw2 * (fx_mn + fx_0n)) * pla_0;
      const double dkedlat = (ke_n - ke0) * rda;
      double dv = -qhu - dkedlat;
      if (hyd) {
        // :1181-1214 with we_lev_lat from interp_mod.F90:426-434
        const double mlat = (m_0 + m_n) * 0.5;
        const double wl1 = (an * w1_n + as * w1_0) * rala, wl0 = (an * w_n + as * w_0) * rala;
        double w;
        if (k > 0 && k < nlev - 1) w = wl1 * (v_dn - v0) + wl0 * (v0 - v_up);
        else if (k == 0) w = wl1 * (v_dn - v0);
        else w = wl0 * (v0 - v_up);
        dv = dv - qdiv(w, mlat) * 0.5;
      }
      a.dv[c] = dv;
    }
  }
}
__global__ void __launch_bounds__(128, TEND_MINB) k_tend(Geo g, DMesh M, TendArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y * blockDim.y + threadIdx.y;
  if (i >= g.nlon || j > g.je) return;
  if (edge_block(g.nlon, 1)) tend_col<true>(g, M, a, i, j); else tend_col<false>(g, M, a, i, j);
}

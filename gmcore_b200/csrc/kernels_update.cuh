// State-update kernels: zonal polar filters, pressure-gradient force + u,v update,
// global diagnostics, host<->device layout conversion.
#pragma once
#include "kernels_dyn.cuh"

// ---- zonal Gaussian filter (src/dynamics/filter_mod.F90:25-163) -------------------
// One block per (latitude row, level).  The row is staged once in shared memory with
// its periodic wrap, then up to two filters are applied to it (update_state applies
// big_filter and a small filter back to back to the same raw row,
// time_schemes_mod.F90:186-190,216-236).  Taps are accumulated left to right exactly
// as sum(wgt(:n) * x(i-hn:i+hn)) so results are bit-identical to the CPU order.
enum { POST_NONE = 0, POST_SWM_M = 1, POST_DIV_M = 2 };
struct FilterArgs {
  const double* in;
  int half_lat;                 // rows are half-lat rows (filter_on_lat_edge) or full-lat
  int nlev;
  DFilter f1; double* out1;     // out1 = f1(in)          (may be NULL)
  DFilter f2; double* out2;     // out2 = post(f2(in))    (may be NULL, may alias in)
  int use_f1, use_f2;           // 0: plain copy for that output
  int post; const double* gzs; double grav; const double* mdiv; double* out_m;
  // chain: out2 = f2(in) then out1 = f1(out2) from a second staged row (small_filter2 -> big_filter on u, v,
  // time_schemes_mod.F90:216-236); f1 is skipped on rows skip_a / skip_b (the pole-adjacent v rows are blended first)
  int chain, row2_off, skip_a, skip_b;
  int mh;                       // largest filter half-width (shared-memory layout)
};
// Row convolution, 4 outputs per thread a quarter row apart (i, i+S, i+2S, i+3S): one shared-memory read of
// the weight serves four multiply-adds, the four x streams stay bank-conflict free and the stores coalesced.
// Every output accumulates its taps left to right (sum(wgt(1:n) * x(i-hn:i+hn))) as before.
__device__ __forceinline__ double conv1(const double* __restrict__ x, const double* __restrict__ w, int n) {
  double s = 0;
  for (int t = 0; t < n; ++t) s += w[t] * x[t];
  return s;
}
// out[i] = post(f(x))[i] for the whole row; xs = staged row with halo hn_s, filter half-width hnf (<= hn_s)
template <class Store> __device__ __forceinline__ void conv_row(const double* __restrict__ xs, int hn_s, const double* __restrict__ w, int n, int nlon, Store store) {
  const int hnf = n > 1 ? (n - 1) / 2 : 0;
  const double* __restrict__ x0 = xs + (hn_s - hnf);
  if (n <= 1) { for (int i = threadIdx.x; i < nlon; i += blockDim.x) store(i, xs[i + hn_s]); return; }
  const int S = (nlon + 3) >> 2;
  for (int r = threadIdx.x; r < S; r += blockDim.x) {
    if (r + 3 * S < nlon) {
      const double* __restrict__ xa = x0 + r; const double* __restrict__ xb = xa + S; const double* __restrict__ xc = xb + S; const double* __restrict__ xd = xc + S;
      double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
      for (int t = 0; t < n; ++t) { const double wt = w[t]; s0 += wt * xa[t]; s1 += wt * xb[t]; s2 += wt * xc[t]; s3 += wt * xd[t]; }
      store(r, s0); store(r + S, s1); store(r + 2 * S, s2); store(r + 3 * S, s3);
    } else {
      for (int q = 0; q < 4; ++q) if (r + q * S < nlon) store(r + q * S, conv1(x0 + r + q * S, w, n));
    }
  }
}
// shared memory: [row: nlon+2*mh][row2: nlon+2*mh][w1: 2*mh+1][w2: 2*mh+1], mh = a.mh (largest half-width of any filter)
#ifndef FILT_DIRECT3
#define FILT_DIRECT3 1
#endif
#ifndef FILT_STAGE4
#define FILT_STAGE4 1
#endif
__global__ void __launch_bounds__(256) k_filter_rows(Geo g, FilterArgs a) {
  extern __shared__ double srow[];
  // grid = (level, row rank): rows are issued from the two ends of the band inwards, i.e. widest filters (most
  // taps, longest blocks) first and all levels of a row together, so the launch does not end on a polar tail
  const int rr = blockIdx.y, nr = gridDim.y, k = blockIdx.x;
  const int j = (a.half_lat ? g.hjb : g.jb) + ((rr & 1) ? nr - 1 - (rr >> 1) : (rr >> 1));
  if (a.half_lat ? (j > g.hje) : (j > g.je)) return;
  const int n1 = a.use_f1 ? (a.half_lat ? a.f1.ngrid_lat[j + 1] : a.f1.ngrid_lon[j + 1]) : 0;
  const int n2 = a.use_f2 ? (a.half_lat ? a.f2.ngrid_lat[j + 1] : a.f2.ngrid_lon[j + 1]) : 0;
  const int hn1 = n1 > 1 ? (n1 - 1) / 2 : 0, hn2 = n2 > 1 ? (n2 - 1) / 2 : 0;
  // chain rows the first filter leaves alone (most rows: small_filter2 is polar only) stage the raw row once,
  // with the second filter's halo, and skip the intermediate row
  const bool chain1 = a.chain && n2 <= 1 && j != a.skip_a && j != a.skip_b;
  const int hn = a.chain ? (chain1 ? hn1 : hn2) : (hn1 > hn2 ? hn1 : hn2);
  const int nlon = g.nlon, rs = nlon + 2 * a.mh;
  if (!a.chain && n1 <= 1 && n2 <= 1) {      // unfiltered row (block-uniform): straight from global memory
    const int base = gidx(g, 0, j, k);
#if FILT_DIRECT3
    if (a.post == POST_DIV_M && !a.out1 && !a.out_m) {     // pt: three points per thread in flight
      const int B = blockDim.x;
      for (int i = threadIdx.x; i < nlon; i += 3 * B) {
        const bool v1 = i + B < nlon, v2 = i + 2 * B < nlon;
        const double s0 = __ldg(a.in + base + i), m0 = __ldg(a.mdiv + base + i);
        const double s1 = v1 ? __ldg(a.in + base + i + B) : 0.0, m1 = v1 ? __ldg(a.mdiv + base + i + B) : 1.0;
        const double s2 = v2 ? __ldg(a.in + base + i + 2 * B) : 0.0, m2 = v2 ? __ldg(a.mdiv + base + i + 2 * B) : 1.0;
        a.out2[base + i] = s0 / m0;
        if (v1) a.out2[base + i + B] = s1 / m1;
        if (v2) a.out2[base + i + 2 * B] = s2 / m2;
      }
      return;
    }
#endif
    for (int i = threadIdx.x; i < nlon; i += blockDim.x) {
      double s = __ldg(a.in + base + i);
      if (a.out1) a.out1[base + i] = s;
      if (a.post == POST_DIV_M) s = s / __ldg(a.mdiv + base + i);
      if (a.out2) a.out2[base + i] = s;
      if (a.post == POST_SWM_M) a.out_m[base + i] = (s - __ldg(a.gzs + gidx(g, i, j, 0))) / a.grav;
    }
    return;
  }
  double* __restrict__ srow2 = srow + rs;
  double* __restrict__ ws1 = srow + 2 * rs; double* __restrict__ ws2 = ws1 + (2 * a.mh + 1);
  const double* __restrict__ row = a.in + gidx(g, 0, j, k);
#if FILT_STAGE4
  {  // four loads per thread in flight before the first shared-memory store (the plain loop waits on every load)
    const int tot = nlon + 2 * hn, B = blockDim.x;
    for (int x0 = threadIdx.x; x0 < tot; x0 += 4 * B) {
      double v[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int x = x0 + q * B;
        int i = x - hn; i = i < 0 ? i + nlon : (i >= nlon ? i - nlon : i);
        v[q] = x < tot ? row[i] : 0.0;
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) if (x0 + q * B < tot) srow[x0 + q * B] = v[q];
    }
  }
#else
  for (int x = threadIdx.x; x < nlon + 2 * hn; x += blockDim.x) {
    int i = x - hn; i = i < 0 ? i + nlon : (i >= nlon ? i - nlon : i);
    srow[x] = row[i];
  }
#endif
  const double* __restrict__ w1 = a.half_lat ? a.f1.wgt_lat + (size_t)(j + 1) * a.f1.ld_lat : a.f1.wgt_lon + (size_t)(j + 1) * a.f1.ld_lon;
  const double* __restrict__ w2 = a.half_lat ? a.f2.wgt_lat + (size_t)(j + 1) * a.f2.ld_lat : a.f2.wgt_lon + (size_t)(j + 1) * a.f2.ld_lon;
  for (int t = threadIdx.x; t < n1; t += blockDim.x) ws1[t] = __ldg(w1 + t);
  for (int t = threadIdx.x; t < n2; t += blockDim.x) ws2[t] = __ldg(w2 + t);
  __syncthreads();
  const int base = gidx(g, 0, j, k);
  if (chain1) {
    if (a.out2 != a.in) for (int i = threadIdx.x; i < nlon; i += blockDim.x) a.out2[base + i] = srow[i + hn];
    conv_row(srow, hn, ws1, n1, nlon, [&](int i, double s) { a.out1[base + i] = s; });
    return;
  }
  if (a.chain) {
    conv_row(srow, hn, ws2, n2, nlon, [&](int i, double s) { a.out2[base + i] = s; srow2[i + hn1] = s; });
    __syncthreads();
    if (j == a.skip_a || j == a.skip_b) {     // block-uniform
      // pole-adjacent v row: blend with the small-filtered neighbour row first (time_schemes_mod.F90:225-233:
      // v(j) = 0.4 v(j) + 0.6 v(jn)); the neighbour's filtered row is recomputed here from its raw row.
      const int jn = j == a.skip_a ? j + 1 : j - 1;
      const int n2n = a.f2.ngrid_lat[jn + 1], hn2n = n2n > 1 ? (n2n - 1) / 2 : 0;
      const double* __restrict__ rown = a.in + gidx(g, 0, jn, k);
      const double* __restrict__ w2n = a.f2.wgt_lat + (size_t)(jn + 1) * a.f2.ld_lat;
      for (int x = threadIdx.x; x < nlon + 2 * hn2n; x += blockDim.x) {
        int i = x - hn2n; i = i < 0 ? i + nlon : (i >= nlon ? i - nlon : i);
        srow[x] = rown[i];
      }
      for (int t = threadIdx.x; t < n2n; t += blockDim.x) ws2[t] = __ldg(w2n + t);
      __syncthreads();
      const double wgt = 0.6;
      conv_row(srow, hn2n, ws2, n2n, nlon, [&](int i, double sn) {
        const double v = (1 - wgt) * srow2[i + hn1] + wgt * sn;
        a.out2[base + i] = v; srow2[i + hn1] = v;
      });
      __syncthreads();
    }
    if (n1 > 1) {
      for (int x = threadIdx.x; x < hn1; x += blockDim.x) { srow2[x] = srow2[x + nlon]; srow2[hn1 + nlon + x] = srow2[hn1 + x]; }
      __syncthreads();
    }
    conv_row(srow2, hn1, ws1, n1, nlon, [&](int i, double s) { a.out1[base + i] = s; });
    return;
  }
  if (a.out1) conv_row(srow, hn, ws1, n1, nlon, [&](int i, double s) { a.out1[base + i] = s; });
  if (a.out2 || a.out_m)
    conv_row(srow, hn, ws2, n2, nlon, [&](int i, double s) {
      if (a.post == POST_DIV_M) s = s / __ldg(a.mdiv + base + i);
      if (a.out2) a.out2[base + i] = s;
      if (a.post == POST_SWM_M) a.out_m[base + i] = (s - __ldg(a.gzs + gidx(g, i, j, 0))) / a.grav;
    });
}

// ---- backward pass: PGF + du,dv + u,v update (gmcore_mod.F90:529-571,
// pgf_swm_mod.F90:18-46, pgf_lin97_mod.F90:57-102, time_schemes_mod.F90:198-233) ------
struct UvArgs {
  const double *gz_f;                 // shallow water
  const double *ph_exn_lev, *gz_lev;  // hydrostatic (half-level arrays)
  const double *du_fwd, *dv_fwd, *u_old, *v_old;
  double *u_new, *v_new;
  double dt; int swm; int blend;      // blend: apply the pole-adjacent v blend here (no small_filter2 in between)
};
__device__ __forceinline__ double pgf_lat_at(const Geo& g, const DMesh& M, const UvArgs& a, int i, int j, int k) {
  if (a.swm) return (ldW(a.gz_f, g, i, j + 1, k) - ldW(a.gz_f, g, i, j, k)) / MJ(M.de_lat, j);
  const double dph1 = ldW(a.ph_exn_lev, g, i, j + 1, k + 1) - ldW(a.ph_exn_lev, g, i, j, k);
  const double dph2 = ldW(a.ph_exn_lev, g, i, j, k + 1) - ldW(a.ph_exn_lev, g, i, j + 1, k);
  const double dgz1 = ldW(a.gz_lev, g, i, j, k + 1) - ldW(a.gz_lev, g, i, j + 1, k);
  const double dgz2 = ldW(a.gz_lev, g, i, j, k) - ldW(a.gz_lev, g, i, j + 1, k + 1);
  return -(dph1 * dgz1 + dph2 * dgz2) / MJ(M.de_lat, j) / (dph1 + dph2);
}
__device__ __forceinline__ double v_raw_at(const Geo& g, const DMesh& M, const UvArgs& a, int i, int j, int k) {
  const long long c = gidx(g, i, j, k);
  const double dv = __ldg(a.dv_fwd + c) - pgf_lat_at(g, M, a, i, j, k);
  return __ldg(a.v_old + c) + a.dt * dv;
}
__global__ void __launch_bounds__(128) k_uv_update(Geo g, DMesh M, UvArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y, k = blockIdx.z;
  if (i >= g.nlon) return;
  const long long c = gidx(g, i, j, k);
  if (j > 1 && j < g.nlat) {
    double pgf;
    if (a.swm) pgf = (ldW(a.gz_f, g, i + 1, j, k) - __ldg(a.gz_f + c)) / MJ(M.de_lon, j);
    else {
      const double dph1 = ldW(a.ph_exn_lev, g, i + 1, j, k + 1) - ldW(a.ph_exn_lev, g, i, j, k);
      const double dph2 = ldW(a.ph_exn_lev, g, i, j, k + 1) - ldW(a.ph_exn_lev, g, i + 1, j, k);
      const double dgz1 = ldW(a.gz_lev, g, i, j, k + 1) - ldW(a.gz_lev, g, i + 1, j, k);
      const double dgz2 = ldW(a.gz_lev, g, i, j, k) - ldW(a.gz_lev, g, i + 1, j, k + 1);
      pgf = -(dph1 * dgz1 + dph2 * dgz2) / MJ(M.de_lon, j) / (dph1 + dph2);
    }
    const double du = __ldg(a.du_fwd + c) - pgf;
    a.u_new[c] = __ldg(a.u_old + c) + a.dt * du;
  }
  if (j <= g.nhalf) {
    double v = v_raw_at(g, M, a, i, j, k);
    if (a.blend) {
      const double wgt = 0.6;
      if (j == 1) v = (1 - wgt) * v + wgt * v_raw_at(g, M, a, i, j + 1, k);
      else if (j == g.nhalf) v = (1 - wgt) * v + wgt * v_raw_at(g, M, a, i, j - 1, k);
    }
    a.v_new[c] = v;
  }
}
// Hydrostatic specialisation: Lin97 PGF on both edges from one up-front load of the 6 distinct
// (ph_exn_lev, gz_lev) corner pairs; /de_lon, /de_lat are reciprocal multiplies.
template <bool EDGE> __device__ __forceinline__ void uv_h_body(const Geo& g, const DMesh& M, const UvArgs& a, int i, int j, int k) {
  const int n = g.nlon, c = gidx(g, i, j, k), pl = (int)g.plane;
  const int dp = wrapE<EDGE>(i + 1, n) - i;
  const bool full = j > 1 && j < g.nlat, half = j <= g.nhalf;
  const RowCoef* __restrict__ R = M.rows + (j + 1);
  const double* __restrict__ pe = a.ph_exn_lev + c; const double* __restrict__ gz = a.gz_lev + c;
  const double p00 = __ldg(pe), p01 = __ldg(pe + pl), g00 = __ldg(gz), g01 = __ldg(gz + pl);
  double pp0 = 0, pp1 = 0, gp0 = 0, gp1 = 0, pn0 = 0, pn1 = 0, gn0 = 0, gn1 = 0, duf = 0, dvf = 0, uo = 0, vo = 0;
  if (full) { pp0 = __ldg(pe + dp); pp1 = __ldg(pe + pl + dp); gp0 = __ldg(gz + dp); gp1 = __ldg(gz + pl + dp); duf = __ldg(a.du_fwd + c); uo = __ldg(a.u_old + c); }
  if (half) { pn0 = __ldg(pe + n); pn1 = __ldg(pe + pl + n); gn0 = __ldg(gz + n); gn1 = __ldg(gz + pl + n); dvf = __ldg(a.dv_fwd + c); vo = __ldg(a.v_old + c); }
  if (full) {
    const double dph1 = pp1 - p00, dph2 = p01 - pp0, dgz1 = g01 - gp0, dgz2 = g00 - gp1;
    const double pgf = qdiv(-(dph1 * dgz1 + dph2 * dgz2) * R->r_de_lon, dph1 + dph2);
    a.u_new[c] = uo + a.dt * (duf - pgf);
  }
  if (half) {
    const double dph1 = pn1 - p00, dph2 = p01 - pn0, dgz1 = g01 - gn0, dgz2 = g00 - gn1;
    const double pgf = qdiv(-(dph1 * dgz1 + dph2 * dgz2) * R->r_de_lat, dph1 + dph2);
    a.v_new[c] = vo + a.dt * (dvf - pgf);
  }
}
__global__ void __launch_bounds__(128) k_uv_update_h(Geo g, DMesh M, UvArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y * blockDim.y + threadIdx.y, k = blockIdx.z;
  if (i >= g.nlon || j > g.je) return;
  if (edge_block(g.nlon, 1)) uv_h_body<true>(g, M, a, i, j, k); else uv_h_body<false>(g, M, a, i, j, k);
}
// Pole-adjacent v blend as a separate pass (hydrostatic core: it follows small_filter2;
// time_schemes_mod.F90:225-233).  Rows 1 and nhalf only; north uses the final j-1 row.
__global__ void k_v_pole_blend(Geo g, double* v) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, k = blockIdx.y, north = blockIdx.z;
  if (i >= g.nlon) return;
  if ((north && !g.has_north) || (!north && !g.has_south)) return;
  const double wgt = 0.6;
  const int j = north ? g.nhalf : 1, jn = north ? j - 1 : j + 1;
  v[gidx(g, i, j, k)] = (1 - wgt) * v[gidx(g, i, j, k)] + wgt * v[gidx(g, i, jn, k)];
}

// ---- diagnose (gmcore_mod.F90:213-341) ----------------------------------------------
// Stage 1: one block per (row, level) -> 7 partial sums; stage 2: one block folds
// them in a fixed order (deterministic run to run).
struct DiagArgs {
  const double *m, *mfx_lon, *mfy_lat, *u_lon, *v_lat, *pv, *pv_lon, *pv_lat, *t, *pt, *phs, *gzs;
  double* partial;   // [7][nrows*nlev]
  double grav, cpd; int baroclinic;
};
__global__ void __launch_bounds__(256) k_diag_partial(Geo g, DMesh M, DiagArgs a) {
  __shared__ double sm7[7][32];
  const int j = g.jb + blockIdx.x, k = blockIdx.y;
  const int nrows = g.je - g.jb + 1;
  double tm = 0, ke = 0, ie = 0, pe = 0, av = 0, tpe = 0, tpt = 0;
  const bool nopole = j > 1 && j < g.nlat, half = j <= g.nhalf;
  // Interior rows of the hydrostatic core (all but the two pole rows): one straight-line body, so that the twelve
  // loads of a point are issued together; the general loop below has one basic block (= one memory round trip)
  // per term.  Same terms in the same order.
  const bool common = nopole && half && a.baroclinic && j + 1 <= g.nlat;
  if (common) {
    const double ac = MJ(M.area_cell, j), alon = MJ(M.area_lon, j), alat = MJ(M.area_lat, j), avtx = MJ(M.area_vtx, j);
    const int n = g.nlon;
    const long long rb = gidx(g, 0, j, k);
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const long long c = rb + i;
      const int ip = i + 1 < n ? i + 1 : 0;
      const double m = __ldg(a.m + c), me = __ldg(a.m + rb + ip), mn = __ldg(a.m + c + n);
      const double fx = __ldg(a.mfx_lon + c), ul = __ldg(a.u_lon + c), pl = __ldg(a.pv_lon + c);
      const double fy = __ldg(a.mfy_lat + c), vl = __ldg(a.v_lat + c), pv = __ldg(a.pv + c), pa = __ldg(a.pv_lat + c);
      const double t = __ldg(a.t + c), pt = __ldg(a.pt + c);
      tm += m * ac;
      ke += fx * 0.5 * ul * alon * 2;
      tpe += (m + me) * 0.5 * (pl * pl) * 0.5 * alon;
      ke += fy * 0.5 * vl * alat * 2;
      av += pv * avtx;
      tpe += (m + mn) * 0.5 * (pa * pa) * 0.5 * alat;
      ie += m * a.cpd * t * ac;
      tpt += m * pt * ac;
      if (k == 0) pe += __ldg(a.gzs + gidx(g, i, j, 0)) * __ldg(a.phs + gidx(g, i, j, 0)) * ac;
    }
  } else
  for (int i = threadIdx.x; i < g.nlon; i += blockDim.x) {
    const long long c = gidx(g, i, j, k);
    const double m = __ldg(a.m + c);
    tm += m * MJ(M.area_cell, j);
    if (nopole) {
      ke += __ldg(a.mfx_lon + c) * 0.5 * __ldg(a.u_lon + c) * MJ(M.area_lon, j) * 2;
      const double pl = __ldg(a.pv_lon + c);
      tpe += m_lon_at(a.m, g, i, j, k) * (pl * pl) * 0.5 * MJ(M.area_lon, j);
    }
    if (half) {
      ke += __ldg(a.mfy_lat + c) * 0.5 * __ldg(a.v_lat + c) * MJ(M.area_lat, j) * 2;
      av += __ldg(a.pv + c) * MJ(M.area_vtx, j);
      const double pl = __ldg(a.pv_lat + c);
      tpe += m_lat_at(a.m, g, i, j, k) * (pl * pl) * 0.5 * MJ(M.area_lat, j);
    }
    if (a.baroclinic) {
      ie += m * a.cpd * __ldg(a.t + c) * MJ(M.area_cell, j);
      tpt += m * __ldg(a.pt + c) * MJ(M.area_cell, j);
      if (k == 0) pe += __ldg(a.gzs + gidx(g, i, j, 0)) * __ldg(a.phs + gidx(g, i, j, 0)) * MJ(M.area_cell, j);
    } else {
      pe += (m * m * a.grav * 0.5 + m * __ldg(a.gzs + gidx(g, i, j, 0))) * MJ(M.area_cell, j);
    }
  }
  const int slot = blockIdx.y * nrows + blockIdx.x, ns = nrows * g.nlev;
  // seven block sums with the summation tree of block_sum (warp shuffles, then one warp per diagnostic folds the
  // per-warp values) behind one barrier instead of 21
  double v[7] = {tm, ke, ie, pe, av, tpe, tpt};
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int q = 0; q < 7; ++q) {
    double x = v[q];
    for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
    if (lane == 0) sm7[q][w] = x;
  }
  __syncthreads();
  if (w < 7) {
    double r = lane < nw ? sm7[w][lane] : 0.0;
    for (int o = 16; o > 0; o >>= 1) r += __shfl_down_sync(0xffffffffu, r, o);
    if (lane == 0) a.partial[(size_t)w * ns + slot] = r;
  }
}
// Block 7 carries the device guard word (the reference's log_error guards, hydro_check_flags) out with the sums:
// out[7] = sum_b [bit b set] * 256^b, so that an all-reduce over <= 255 bands still tells which guards fired.
__global__ void __launch_bounds__(1024) k_diag_final(const double* partial, int ns, double* out /*[8]*/, const int* flags) {
  __shared__ double sm[32];
  const int q = blockIdx.x;            // one block per diagnostic; the summation order within one is fixed
  if (q == 7) {
    if (threadIdx.x == 0) {
      const int f = flags ? *flags : 0; double v = 0, w = 1;
      for (int b = 0; b < 5; ++b, w *= 256.0) if (f & (1 << b)) v += w;
      out[7] = v;
    }
    return;
  }
  double s = 0;
  for (int x = threadIdx.x; x < ns; x += blockDim.x) s += partial[(size_t)q * ns + x];
  s = block_sum(s, sm);
  if (threadIdx.x == 0) out[q] = s;
}

// ---- host (halo'd Fortran storage) <-> device (compact) layout conversion --------------
// dir 0: staged host image -> device field; dir 1: device field -> staged image.
struct PackArgs { double* dev; double* img; int hx, hy, nx, ny, klev0, nk, nrows, jfirst; };
__global__ void k_pack(Geo g, PackArgs a, int dir) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y, k = blockIdx.z;
  if (i >= g.nlon) return;
  const int j = a.jfirst + r;   // global row
  const long long d = gidx(g, i, j, k);
  const long long s = (long long)(i + a.hx) + (long long)a.nx * ((long long)(j - g.jb + a.hy) + (long long)a.ny * (k + a.klev0));
  if (dir == 0) a.dev[d] = a.img[s]; else a.img[s] = a.dev[d];
}

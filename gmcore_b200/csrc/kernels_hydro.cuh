// Hydrostatic-core kernels: RK4 departure-point CFL, vertical mass flux, FFSL/PPM
// advection (horizontal inner/outer + vertical), hydrostatic integration, damping,
// Held-Suarez forcing, tracer batch.  References (GMCORE source tree):
//   src/dynamics/adv/adv_batch_mod.F90:218-403,493-632
//   src/dynamics/adv/ffsl_mod.F90:318-432,778-941,1147-1194
//   src/dynamics/operators_mod.F90:136-232,433-543,1008-1216
//   src/dynamics/damp/div_damp_mod.F90:27-139, lon_damp_mod.F90:43-68,158-241
//   src/dynamics/adv/adv_mod.F90:152-312,420-455
#pragma once
#include "handle.cuh"
#include "kernels_dyn.cuh"
#include "fastmath.cuh"

// Vertical-coordinate constants (hybrid_coord_mod.F90:146-171), by value.
struct VC { double ptop, p0h, p0, kappa, Rd; };
// Exner function and layer thickness (k_ph, k_gz_col, k_lev_tab): positive-argument pow / log of fastmath.cuh
// (GMC_FASTMATH=0: the CUDA library routines, for A/B runs)
#ifndef GMC_FASTMATH
#define GMC_FASTMATH 1
#endif
#if GMC_FASTMATH
#define GM_POW(x, y) pow_pos(x, y)
#define GM_LOG(x) log_pos(x)
#else
#define GM_POW(x, y) pow(x, y)
#define GM_LOG(x) log(x)
#endif

__device__ __forceinline__ double ph_lev_at(const DMesh& M, const VC& vc, int k1 /*1-based half level*/, double phs) {
  return M.hyai[k1 - 1] * (vc.p0h - vc.ptop) + M.hybi[k1 - 1] * (phs - vc.ptop) + vc.ptop;
}

__device__ __forceinline__ double fsign(double a, double b) { return copysign(fabs(a), b); }
// min/max as compare+select: the inputs are never NaN and the sign of a zero result never
// matters below (it is either subtracted or its magnitude is re-signed), so the IEEE fmin/fmax
// NaN plumbing (6 instructions each on sm_100a, no native DMNMX) is not needed.
__device__ __forceinline__ double dmin(double a, double b) { return a < b ? a : b; }
__device__ __forceinline__ double dmax(double a, double b) { return a > b ? a : b; }
__device__ __forceinline__ double min3(double a, double b, double c) { return dmin(dmin(a, b), c); }
__device__ __forceinline__ double max3(double a, double b, double c) { return dmax(dmax(a, b), c); }

// ffsl_mod.F90:1183-1194 (limiter_type = 'mono'): sign(df) * min(|df|, 2 (f - min3), 2 (max3 - f)).
// With da = f - fm1, db = fp1 - f:  f - min3 = max(da, 0, -db) and max3 - f = max(-da, 0, db) exactly,
// so the limiter is min(|df|, 2|da|, 2|db|) when da, db have the same sign and 0 at a local extremum --
// two min-selects instead of six min/max-selects, identical values (up to the sign of a zero).
// LIM = limiter_type (ffsl_mod.F90:131-140): 0 'none' (slope_simple :1175), 1 'mono', 2 'pd' (slope_pd :1196, f = centre value)
// Kernels carry MODE = limiter_type + 4 * ffsl_flux_type; only the two low bits matter here.
template <int MODE> __device__ __forceinline__ double slope_mono_d(double da, double db, double df, double f) {
  constexpr int LIM = MODE & 3;
  if (LIM == 0) return df;
  if (LIM == 2) return copysign(dmin(fabs(df), 2 * f), df);
  const double lim = dmin(fabs(df), 2 * dmin(fabs(da), fabs(db)));
  return da * db > 0 ? copysign(lim, df) : 0.0;
}
// ffsl_mod.F90:1147-1173; the four first differences are shared by the three slopes
template <int LIM> __device__ __forceinline__ void ppm(double fm2, double fm1, double f, double fp1, double fp2, double& fl, double& df, double& f6) {
  const double d1 = fm1 - fm2, d2 = f - fm1, d3 = fp1 - f, d4 = fp2 - fp1;
  const double dfl = slope_mono_d<LIM>(d1, d2, (f - fm2) * 0.5, fm1);
  df = slope_mono_d<LIM>(d2, d3, (fp1 - fm1) * 0.5, f);
  const double dfr = slope_mono_d<LIM>(d3, d4, (fp2 - f) * 0.5, fp1);
  fl = 0.5 * (fm1 + f) + (dfl - df) * (1.0 / 6.0);
  double fr = 0.5 * (fp1 + f) + (df - dfr) * (1.0 / 6.0);
  fl = f - fsign(dmin(fabs(df), fabs(fl - f)), df);
  fr = f + fsign(dmin(fabs(df), fabs(fr - f)), df);
  f6 = 6 * f - 3 * (fl + fr);
  df = fr - fl;
}
// ffsl_flux_type = 'van_leer' (ffsl_mod.F90:556-658): piecewise-linear profile, slope of the upwind cell
template <int MODE> __device__ __forceinline__ double vl_slope(double fm1, double f, double fp1) {
  return slope_mono_d<MODE>(f - fm1, fp1 - f, (fp1 - fm1) * 0.5, f);
}
// fractional-cell part + whole-cell sum along x / z (ffsl_mod.F90:589-601, 640-652); vel = u | w, c = CFL != 0
template <int MODE> __device__ __forceinline__ double vl_flux(double fm1, double f, double fp1, double sm, double cf, double c, double vel) {
  const double dm = vl_slope<MODE>(fm1, f, fp1);
  const double t = c > 0 ? cf * (f + dm * 0.5 * (1 - cf)) + sm : cf * (f - dm * 0.5 * (1 + cf)) - sm;
  return qdiv(vel * t, c);
}
// The PPM flux integral shared by the three directions (ffsl_mod.F90:831-851).
__device__ __forceinline__ double ppm_integral(double sm, double ml, double dm, double m6, double s1, double s2) {
  const double ds1 = s2 - s1, ds2 = s2 * s2 - s1 * s1, ds3 = s2 * s2 * s2 - s1 * s1 * s1;
  return sm + ml * ds1 + 0.5 * dm * ds2 + m6 * (ds2 * 0.5 - ds3 * (1.0 / 3.0));
}

// ---- RK4 departure-point CFL + divx/divy (adv_batch_mod.F90:253-373) ---------------
struct CflArgs {
  const double *u, *v;                 // batch%u, batch%v
  double *cflx, *cfly, *divx, *divy;
  double dt, radius; int* flags;
  int jx0, jx1, jy0, jy1;              // rows of the x part (cflx, divx, divy) / the y part (cfly); band mode
};
// The reference's RK4 loop for one edge with gathers at any distance (adv_batch_mod.F90:256-288 / 290-335): the slow
// path of k_cfl, taken by the lanes whose departure point leaves the three-point window (polar rows in x).
__device__ __noinline__ double cfl_chain_x(const double* __restrict__ ur, const double* __restrict__ hl, int i1, int n, int hx,
                                           double rcl, double rdx, double x0, double u1, double dt, int* flags) {
  double x = x0 + dt / 2 * u1, us[3];
#pragma unroll
  for (int s = 0; s < 3; ++s) {
    const double h = s == 1 ? dt : dt / 2;
    int lx = __double2int_rd(i1 + (x - x0) * rdx), d = lx - i1;
    // a departure point outside the longitude halo is the reference's "cflx exceeds mesh%lon_halo_width" abort
    // (adv_batch_mod.F90:284-287): raise the guard and keep the gather inside the coefficient table
    if (d < -hx || d > hx - 2) { atomicOr(flags, 1); d = d < 0 ? -hx : hx - 2; lx = i1 + d; }
    const double ua = __ldg(ur + wrapi(lx - 1, n)), ub = __ldg(ur + wrapi(lx, n)), xl = rcl * hl[lx], xr = rcl * hl[lx + 1];
    us[s] = ((xr - x) * ua + (x - xl) * ub) * rdx; x = x0 + h * us[s];
  }
  return (u1 + 2 * us[0] + 2 * us[1] + us[2]) * (1.0 / 6.0) * dt * rdx;
}
// vc = &v(i, j, k); half_lat = mesh vector (entry j at [j + 1])
__device__ __noinline__ double cfl_chain_y(const double* __restrict__ vc, const double* __restrict__ half_lat, int j, int n, int nhalf,
                                           double radius, double rdy, double y0, double v1, double dt) {
  double y = y0 + dt / 2 * v1, vs[3];
#pragma unroll
  for (int s = 0; s < 3; ++s) {
    const double h = s == 1 ? dt : dt / 2;
    const int ly = __double2int_rd(j + (y - y0) * rdy);
    if (ly < 1 || ly > nhalf - 1) vs[s] = v1;
    else {
      const double va = __ldg(vc + (ly - j) * n), vb = __ldg(vc + (ly + 1 - j) * n), yl = radius * half_lat[ly + 1], yr = radius * half_lat[ly + 2];
      vs[s] = ((yr - y) * va + (y - yl) * vb) * rdy;
    }
    y = y0 + h * vs[s];
  }
  return (v1 + 2 * vs[0] + 2 * vs[1] + vs[2]) * (1.0 / 6.0) * dt * rdy;
}
// One memory phase and straight-line code in the common case: a departure point less than one cell away (|CFL| < 1,
// everywhere except the polar rows in x) interpolates between the edge's own value and one of its two neighbours, so the
// three-point windows u(i-1..i+1), v(j-1..j+1) and the edge coordinates are loaded up front and the three RK4 stages are
// branch-free register arithmetic (the window side is a select); a lane whose point leaves the window at any stage redoes
// its chain with the reference's gathers (cfl_chain_x / _y) -- same expressions, hence the same bits where both apply.
// x and y chains interleave; divides by dx, dy, area_cell are multiplies by host-formed reciprocals.
template <bool EDGE> __device__ __forceinline__ void cfl_body(const Geo& g, const DMesh& M, const CflArgs& a, int i, int j, int k) {
  const int n = g.nlon, c = gidx(g, i, j, k), rb = c - i;
  const int i1 = i + 1;                       // Fortran index
  const double* __restrict__ hl = M.half_lon + (M.hx - 1);   // hl[l] = half_lon(l), l Fortran index
  const bool dox = j > 1 && j < g.nlat && j >= a.jx0 && j <= a.jx1;   // warp-uniform
  const bool doy = j <= g.nhalf && j >= a.jy0 && j <= a.jy1;
  const double* __restrict__ ur = a.u + rb;
  const RowCoef* __restrict__ R = M.rows + (j + 1);
  const double rdx = R->r_de_lon, rcl = a.radius * R->full_cos_lat, rdy = R->r_de_lat;
  double u_m = 0, u1 = 0, u_p = 0, v_s = 0, v1 = 0, v_n = 0;
  if (dox) { u_m = __ldg(ur + wrapE<EDGE>(i - 1, n)); u1 = __ldg(ur + i); u_p = __ldg(ur + wrapE<EDGE>(i + 1, n)); }
  if (doy || dox) { v_s = __ldg(a.v + c - n); v1 = __ldg(a.v + c); }
  if (doy) v_n = __ldg(a.v + c + n);
  const double xm = rcl * hl[i1 - 1], x0 = rcl * hl[i1], xp = rcl * hl[i1 + 1];
  const double ym = a.radius * R[-1].half_lat, y0 = a.radius * R->half_lat, yp = a.radius * R[1].half_lat;
  const double fi1 = (double)i1, fj = (double)j;
  double x = x0 + a.dt / 2 * u1, y = y0 + a.dt / 2 * v1;
  double us[3], vs[3];
  bool slowx = false, slowy = false;
#pragma unroll
  for (int s = 0; s < 3; ++s) {
    const double h = s == 1 ? a.dt : a.dt / 2;
    {
      const int d = __double2int_rd(fi1 + (x - x0) * rdx) - i1;
      const bool e = d == 0;
      slowx |= (unsigned)(d + 1) > 1u;
      const double ua = e ? u1 : u_m, ub = e ? u_p : u1, xl = e ? x0 : xm, xr = e ? xp : x0;
      us[s] = ((xr - x) * ua + (x - xl) * ub) * rdx; x = x0 + h * us[s];
    }
    {
      const int ly = __double2int_rd(fj + (y - y0) * rdy), d = ly - j;
      const bool inr = !(ly < 1 || ly > g.nhalf - 1), e = d == 0;
      slowy |= inr && (unsigned)(d + 1) > 1u;
      const double va = e ? v1 : v_s, vb = e ? v_n : v1, yl = e ? y0 : ym, yr = e ? yp : y0;
      const double r = ((yr - y) * va + (y - yl) * vb) * rdy;
      vs[s] = inr ? r : v1;                   // adv_batch_mod.F90:301-303: outside the half rows -> v1
      y = y0 + h * vs[s];
    }
  }
  if (dox) {
    const double rac = R->r_area_cell;
    double cf = (u1 + 2 * us[0] + 2 * us[1] + us[2]) * (1.0 / 6.0) * a.dt * rdx;
    if (slowx) cf = cfl_chain_x(ur, hl, i1, n, M.hx, rcl, rdx, x0, u1, a.dt, a.flags);
    a.cflx[c] = cf;
    if (fabs(cf) > M.hx) atomicOr(a.flags, 1);
    a.divx[c] = (u1 - u_m) * R->le_lon * rac;
    a.divy[c] = (v1 * R->le_lat - v_s * R[-1].le_lat) * rac;
  }
  if (doy) {
    double cf = (v1 + 2 * vs[0] + 2 * vs[1] + vs[2]) * (1.0 / 6.0) * a.dt * rdy;
    if (slowy) cf = cfl_chain_y(a.v + c, M.half_lat, j, n, g.nhalf, a.radius, rdy, y0, v1, a.dt);
    a.cfly[c] = cf;
    if (fabs(cf) > 1) atomicOr(a.flags, 2);
  }
}
// 10 blocks per SM (48 registers, a 136-byte spill on the gather path) measured 2.5 % faster than 64 registers / 8 blocks
#ifndef CFL_MINB
#define CFL_MINB 10
#endif
__global__ void __launch_bounds__(128, CFL_MINB) k_cfl(Geo g, DMesh M, CflArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y * blockDim.y + threadIdx.y, k = blockIdx.z;
  if (i >= g.nlon || j > g.je) return;
  if (edge_block(g.nlon, 1)) cfl_body<true>(g, M, a, i, j, k); else cfl_body<false>(g, M, a, i, j, k);
}

// ---- zonal sums on the pole-adjacent half rows (zonal_sum, parallel_mod.F90:741-764) ---
// out[north*nlev + k] = (+-) sum_i f(i, jh, k) * le_lat(jh) / nlon / area_cell(pole) * scale
// sign: south +f, north -f when signed != 0 (flux divergences); north +f otherwise.
// If row_out != NULL the value is also broadcast along the pole row of row_out.
struct PoleSumArgs { const double* f; double* out; double* row_out; int signed_north; };
__global__ void __launch_bounds__(256) k_pole_sum(Geo g, DMesh M, PoleSumArgs a) {
  __shared__ double sm[32];
  const int k = blockIdx.x, north = blockIdx.y;
  if ((north && !g.has_north) || (!north && !g.has_south)) return;
  const int j = north ? g.nlat : 1, jh = north ? g.nhalf : 1;
  double s = 0;
  const double* __restrict__ fr = a.f + gidx(g, 0, jh, k);
  for (int i0 = threadIdx.x; i0 < g.nlon; i0 += 4 * blockDim.x) {      // four loads in flight, added in the same order
    double v[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) { const int i = i0 + r * blockDim.x; v[r] = i < g.nlon ? __ldg(fr + i) : 0.0; }
#pragma unroll
    for (int r = 0; r < 4; ++r) if (i0 + r * blockDim.x < g.nlon) s += (north && a.signed_north) ? -v[r] : v[r];
  }
  s = block_sum(s, sm);
  const double pole = s * MJ(M.le_lat, jh) / g.nlon / MJ(M.area_cell, j);
  if (a.out && threadIdx.x == 0) a.out[north * g.nlev + k] = pole;
  if (a.row_out) for (int i = threadIdx.x; i < g.nlon; i += blockDim.x) a.row_out[gidx(g, i, j, k)] = pole;
}

// ---- column kernel: calc_grad_mf + calc_dphsdt + calc_we_lev + accum_we_lev ------------
// (operators_mod.F90:939-1006,1120-1143,192-232; adv_batch_mod.F90:526-627).
// One thread per (i, j) column; level loops keep the reference's summation order.
// m_lev is a pure function of phs_f (operators_mod.F90:498-521) and is formed on the fly.
struct WeArgs {
  const double *mfx_lon, *mfy_lat, *pole_dmf, *m_lev;
  double *we_lev;       // (nlev+1 levels) of the star state
  double *dphs;         // 2-D
  double *cflz;         // (nlev+1 levels)
  double dt; int* flags;
};
__device__ __forceinline__ double m_lev_at(const DMesh& M, const VC& vc, int nlev, int k1, double phs) {
  // k1 = 1..nlev+1
  if (k1 == 1) return 0.5 * (ph_lev_at(M, vc, 1, phs) + ph_lev_at(M, vc, 2, phs)) - ph_lev_at(M, vc, 1, phs);
  if (k1 == nlev + 1) return ph_lev_at(M, vc, nlev + 1, phs) - 0.5 * (ph_lev_at(M, vc, nlev, phs) + ph_lev_at(M, vc, nlev + 1, phs));
  return 0.5 * (ph_lev_at(M, vc, k1, phs) + ph_lev_at(M, vc, k1 + 1, phs)) - 0.5 * (ph_lev_at(M, vc, k1 - 1, phs) + ph_lev_at(M, vc, k1, phs));
}

// RK4 in eta for one half level (adv_batch_mod.F90:529-627).  wm is this column's eta velocity
// we_lev / m_lev * half_dlev per half level (index 0 = half level 1), staged in shared memory with
// stride `st`: one divide per level instead of the reference's seven per RK4 evaluation.
// lev = the block's shared-memory copy of the level tables (cflz_tables): the departure searches read half_lev in a
// dependent loop, which through global memory was the hottest line of k_we_col (ncu r02: 32 % of its instructions)
__device__ __forceinline__ void cflz_tables(const DMesh& M, int nlev, double* lev, int tid, int nthreads) {
  for (int l = tid; l <= nlev + 1; l += nthreads) {        // l = 0 .. nlev+1
    lev[l] = l >= 1 ? M.half_lev[l - 1] : 0.0;             // lev[l] = half_lev(l), l = 1 .. nlev+1
    lev[(nlev + 2) + l] = M.full_dlev[l + 1];              // full_dlev(l), l = 0 .. nlev+1 (stored from -1)
    lev[2 * (nlev + 2) + l] = M.r_full_dlev[l + 1];
  }
}
static inline size_t cflz_tables_bytes(int nlev) { return (size_t)3 * (nlev + 2) * sizeof(double); }
__device__ __forceinline__ bool cflz_one(const double* __restrict__ lev, int nlev, int k1, const double* wm, int st,
                                         double dt, double* out, int* flags) {
  const double* hl = lev;                      // hl[l] = half_lev(l)
  const double* fd = lev + (nlev + 2);         // fd[l] = full_dlev(l)
  const double* rfd = lev + 2 * (nlev + 2);
  auto search = [&](double w, double z) {
    int l;
    if (w > 0) { for (l = k1; l <= nlev; ++l) if (hl[l] <= z && z <= hl[l + 1]) break; if (l == nlev + 1) { atomicOr(flags, 4); l = nlev; } }
    else { for (l = k1 - 1; l >= 1; --l) if (hl[l] <= z && z <= hl[l + 1]) break; if (l == 0) { atomicOr(flags, 4); l = 1; } }
    return l;
  };
  auto interp = [&](int l, double z) { return ((hl[l + 1] - z) * wm[(l - 1) * st] + (z - hl[l]) * wm[l * st]) * rfd[l]; };
  const double z0 = hl[k1];
  const double w1 = wm[(k1 - 1) * st];
  double z = z0 + dt / 2 * w1;
  int l = search(w1, z);
  const double w2 = interp(l, z);
  z = z0 + dt / 2 * w2; l = search(w2, z);
  const double w3 = interp(l, z);
  z = z0 + dt * w3; l = search(w3, z);
  const double w4 = interp(l, z);
  double deta = (w1 + 2 * w2 + 2 * w3 + w4) * (1.0 / 6.0) * dt;
  int ks, ke, s;
  if (deta < 0) { ks = k1 - 1; ke = 1; s = -1; } else { ks = k1; ke = nlev; s = 1; }
  deta = fabs(deta);
  for (l = ks; s > 0 ? l <= ke : l >= ke; l += s) {
    deta = deta - fd[l];
    if (deta < 0) { *out = s * ((1 + deta * rfd[l]) + abs(l - ks)); return true; }
  }
  return false;
}

// Column kernels with a serial part (running sums in k) and a parallel part (one RK4 departure search per
// half level): a block is WE_TX columns x WE_TY level groups, so the loads, the eta-velocity divides and the
// RK4 searches of one column are spread over WE_TY threads (levels interleaved, a warp = 32 adjacent columns
// at one level -> coalesced), and only the two running sums are walked by one thread per column, in the
// reference's k order.  4x the threads in flight of the one-thread-per-column version for the same shared memory.
#ifndef WE_TX
#define WE_TX 32
#endif
#ifndef WE_TY
#define WE_TY 8
#endif
#ifndef WECOL_PFM
#define WECOL_PFM 1
#endif
static inline size_t we_col_smem(int nlev) { return (size_t)((nlev + 1) + 2 * nlev + 1) * WE_TX * sizeof(double) + cflz_tables_bytes(nlev); }
static inline size_t cflz_smem(int nlev) { return (size_t)(nlev + 1) * WE_TX * sizeof(double) + cflz_tables_bytes(nlev); }
__global__ void __launch_bounds__(WE_TX * WE_TY) k_we_col(Geo g, DMesh M, WeArgs a) {
  extern __shared__ double wm_s[];       // wm[nlev+1][TX] | dlon[nlev][TX] | dlat[nlev][TX] | dphs[TX]
  const int tx = threadIdx.x, ty = threadIdx.y, nlev = g.nlev;
  const int i = blockIdx.x * WE_TX + tx;
  const int j = g.jb + blockIdx.y;
  const bool act = i < g.nlon;
  const bool pole = (j == 1 || j == g.nlat);
  const int st = (int)g.plane, n = g.nlon;
  const int c0 = act ? gidx(g, i, j, 0) : 0, im = act ? wrapi(i - 1, n) - i : 0;
  const RowCoef* __restrict__ R = M.rows + (j + 1);
  const double rac = R->r_area_cell, lel = R->le_lon, lea = R->le_lat, leas = R[-1].le_lat;
  double* wm = wm_s + tx;
  double* dl = wm_s + (nlev + 1) * WE_TX + tx;
  double* da = dl + nlev * WE_TX;
  double* dps = wm_s + ((nlev + 1) + 2 * nlev) * WE_TX;
  double* lev = dps + WE_TX;
  cflz_tables(M, nlev, lev, ty * WE_TX + tx, WE_TX * WE_TY);      // visible after the first barrier, used by the RK4 phase
  // calc_grad_mf (operators_mod.F90:1114-1166): the two divergence parts of every level
  if (act) {
#pragma unroll 4
    for (int k = ty; k < nlev; k += WE_TY) {
      double dlon, dlat;
      if (WECOL_PFM && (tx & 15) == 0) pf_l2(a.m_lev + c0 + (k + 1) * st);    // read two barriers later
      if (pole) { dlon = 0.0; dlat = __ldg(a.pole_dmf + (j == 1 ? 0 : nlev) + k); }
      else {
        const double* fx = a.mfx_lon + c0 + k * st; const double* fy = a.mfy_lat + c0 + k * st;
        dlon = (__ldg(fx) - __ldg(fx + im)) * lel * rac;
        dlat = (__ldg(fy) * lea - __ldg(fy - n) * leas) * rac;
      }
      dl[k * WE_TX] = dlon; da[k * WE_TX] = dlat;
    }
  }
  __syncthreads();
  // calc_dphsdt + the column prefix of calc_we_lev: running sums in the reference's k order
  if (act && ty == 0) {
    double dphs = 0, mf = 0;
    for (int k = 0; k < nlev; ++k) {
      const double dlon = dl[k * WE_TX], dlat = da[k * WE_TX];
      dphs = dphs - dlon - dlat;
      mf = mf + dlon + dlat;
      wm[(k + 1) * WE_TX] = mf;
    }
    a.dphs[c0] = dphs; dps[tx] = dphs;
  }
  __syncthreads();
  // calc_we_lev, and the eta velocity of accum_we_lev
  if (act) {
    const double dphs = dps[tx];
    double* we = a.we_lev + c0;
    const double* ml = a.m_lev + c0;
    const double* hd = M.half_dlev - 1;
#pragma unroll 4
    for (int k1 = 2 + ty; k1 <= nlev; k1 += WE_TY) {
      const double w = -(M.hybi[k1 - 1] * dphs) - wm[(k1 - 1) * WE_TX];
      we[(k1 - 1) * st] = w;
      wm[(k1 - 1) * WE_TX] = qdiv(w, __ldg(ml + (k1 - 1) * st)) * hd[k1];
    }
    if (ty == 0) { we[0] = 0.0; wm[0] = 0.0; }
    if (ty == 1) { we[nlev * st] = 0.0; wm[nlev * WE_TX] = 0.0; }
  }
  __syncthreads();
  if (act) {
    for (int k1 = 2 + ty; k1 <= nlev; k1 += WE_TY) {
      double cz;
      if (cflz_one(lev, nlev, k1, wm, WE_TX, a.dt, &cz, a.flags)) a.cflz[c0 + (k1 - 1) * st] = cz;
    }
  }
}

// cflz for a tracer batch from its accumulated we / m_lev (adv_batch_mod.F90:526-627)
struct CflzArgs { const double *we, *m_lev; double* cflz; double dt; int* flags; };
__global__ void __launch_bounds__(WE_TX * WE_TY) k_cflz(Geo g, DMesh M, CflzArgs a) {
  extern __shared__ double wm_s[];
  const int tx = threadIdx.x, ty = threadIdx.y, nlev = g.nlev;
  const int i = blockIdx.x * WE_TX + tx;
  const int j = g.jb + blockIdx.y;
  const bool act = i < g.nlon;
  const int st = (int)g.plane, c0 = act ? gidx(g, i, j, 0) : 0;
  double* wm = wm_s + tx;
  double* lev = wm_s + (nlev + 1) * WE_TX;
  cflz_tables(M, nlev, lev, ty * WE_TX + tx, WE_TX * WE_TY);
  const double* hd = M.half_dlev - 1;
  if (act)
#pragma unroll 4
    for (int k1 = 1 + ty; k1 <= nlev + 1; k1 += WE_TY)
      wm[(k1 - 1) * WE_TX] = qdiv(__ldg(a.we + c0 + (k1 - 1) * st), __ldg(a.m_lev + c0 + (k1 - 1) * st)) * hd[k1];
  __syncthreads();
  if (act)
    for (int k1 = 2 + ty; k1 <= nlev; k1 += WE_TY) {
      double cz;
      if (cflz_one(lev, nlev, k1, wm, WE_TX, a.dt, &cz, a.flags)) a.cflz[c0 + (k1 - 1) * st] = cz;
    }
}

// ---- FFSL horizontal PPM flux (hflx_ppm_cell, ffsl_mod.F90:778-881) ----------------
// The PPM coefficients of the upwind cell are formed on the fly from the 5-point
// stencil instead of being stored in six work arrays and halo-filled (:813-826).
struct HflxArgs {
  const double *u, *v;          // advecting wind (inner) or mass flux (outer)
  const double *mx, *my;        // field swept in x / in y
  const double *cflx, *cfly;
  double *mfx, *mfy;
  int jx0, jx1, jy0, jy1;       // rows of the x flux / half rows of the y flux (band mode)
};
// One memory phase in the common case: |cfly| <= 1 always and |cflx| < 1 away from the poles, so the
// upwind cell is the edge's west/east (south/north) neighbour and its 5-point stencil lies inside the
// 6-point window i-2..i+3 (j-2..j+3), whose addresses do not depend on the CFL numbers -- the window
// loads go out together with the CFL / velocity loads.  Lanes with |cflx| >= 1 (polar rows) re-gather
// around their own upwind cell and add the whole-cell segment sum (ffsl_mod.F90:831-848).
template <bool EDGE, int LIM> __device__ __forceinline__ void hflx_body(const Geo& g, const HflxArgs& a, int i, int j, int k) {
  const int n = g.nlon, c = gidx(g, i, j, k), rb = c - i;
  const bool dox = j > 1 && j < g.nlat && j >= a.jx0 && j <= a.jx1;     // warp-uniform
  const bool doy = j <= g.nhalf && j >= a.jy0 && j <= a.jy1;
  double cx = 0, uu = 0, cy = 0, vv = 0;
  double xw0 = 0, xw1 = 0, xw2 = 0, xw3 = 0, xw4 = 0, xw5 = 0, yw0 = 0, yw1 = 0, yw2 = 0, yw3 = 0, yw4 = 0, yw5 = 0;
  const double* __restrict__ mxr = a.mx + rb;
  if (dox) {
    cx = __ldg(a.cflx + c); uu = __ldg(a.u + c);
    xw0 = __ldg(mxr + wrapE<EDGE>(i - 2, n)); xw1 = __ldg(mxr + wrapE<EDGE>(i - 1, n)); xw2 = __ldg(mxr + i);
    xw3 = __ldg(mxr + wrapE<EDGE>(i + 1, n)); xw4 = __ldg(mxr + wrapE<EDGE>(i + 2, n)); xw5 = __ldg(mxr + wrapE<EDGE>(i + 3, n));
  }
  if (doy) {
    cy = __ldg(a.cfly + c); vv = __ldg(a.v + c);
    if (j >= 3 && j + 3 <= g.nlat) {   // away from the poles: no mirror, rows at fixed strides from one address
      const double* __restrict__ myc = a.my + c;
      yw0 = __ldg(myc - 2 * n); yw1 = __ldg(myc - n); yw2 = __ldg(myc); yw3 = __ldg(myc + n); yw4 = __ldg(myc + 2 * n); yw5 = __ldg(myc + 3 * n);
    } else {
      yw0 = ldF(a.my, g, i, j - 2, k); yw1 = ldF(a.my, g, i, j - 1, k); yw2 = ldF(a.my, g, i, j, k);
      yw3 = ldF(a.my, g, i, j + 1, k); yw4 = ldF(a.my, g, i, j + 2, k); yw5 = ldF(a.my, g, i, j + 3, k);
    }
  }
  if (dox) {
    double flux = 0;
    if (cx != 0) {
      int ci = (int)cx;
      if (ci >= n || ci <= -n) ci = 0;      // |cflx| > lon_halo_width: the guard word is already up (k_cfl); stay in the row
      const double cf = cx - ci;
      const bool pos = cx > 0;
      double x0, x1, x2, x3, x4, sm = 0;
      if (ci == 0) {
        x0 = pos ? xw0 : xw1; x1 = pos ? xw1 : xw2; x2 = pos ? xw2 : xw3; x3 = pos ? xw3 : xw4; x4 = pos ? xw4 : xw5;
      } else {
        const int iu = wrapi(pos ? i - ci : i - ci + 1, n);
        x0 = __ldg(mxr + wrapi(iu - 2, n)); x1 = __ldg(mxr + wrapi(iu - 1, n)); x2 = __ldg(mxr + iu);
        x3 = __ldg(mxr + wrapi(iu + 1, n)); x4 = __ldg(mxr + wrapi(iu + 2, n));
        if (pos) { for (int l = i + 1 - ci; l <= i; ++l) sm += __ldg(mxr + wrapi(l, n)); }
        else { for (int l = i + 1; l <= i - ci; ++l) sm += __ldg(mxr + wrapi(l, n)); }
      }
      if (LIM >= 4) {
        flux = vl_flux<LIM>(x1, x2, x3, sm, cf, cx, uu);
      } else {
        const double s1 = pos ? 1 - cf : 0.0, s2 = pos ? 1.0 : -cf;
        double ml, dm, m6;
        ppm<LIM>(x0, x1, x2, x3, x4, ml, dm, m6);
        flux = qdiv((pos ? uu : -uu) * ppm_integral(sm, ml, dm, m6, s1, s2), cx);
      }
    }
    a.mfx[c] = flux;
  }
  if (doy) {
    double flux = 0;
    if (LIM >= 4) {                       // ffsl_mod.F90:605-611: no zero-CFL branch, ju = j + 1 unless cfly > 0
      const bool pos = cy > 0;
      const double y1 = pos ? yw1 : yw2, y2 = pos ? yw2 : yw3, y3 = pos ? yw3 : yw4;
      flux = vv * (y2 + vl_slope<LIM>(y1, y2, y3) * 0.5 * (copysign(1.0, cy) - cy));
    } else if (cy != 0) {
      const bool pos = cy > 0;
      const double y0 = pos ? yw0 : yw1, y1 = pos ? yw1 : yw2, y2 = pos ? yw2 : yw3, y3 = pos ? yw3 : yw4, y4 = pos ? yw4 : yw5;
      const double s1 = pos ? 1 - cy : 0.0, s2 = pos ? 1.0 : -cy;
      double ml, dm, m6;
      ppm<LIM>(y0, y1, y2, y3, y4, ml, dm, m6);
      flux = qdiv((pos ? vv : -vv) * ppm_integral(0.0, ml, dm, m6, s1, s2), cy);
    }
    a.mfy[c] = flux;
  }
}
template <int LIM> __global__ void __launch_bounds__(128) k_hflx(Geo g, HflxArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y * blockDim.y + threadIdx.y, k = blockIdx.z;
  if (i >= g.nlon || j > g.je) return;
  if (edge_block(g.nlon, 3) || blockIdx.x * blockDim.x < 2) hflx_body<true, LIM>(g, a, i, j, k); else hflx_body<false, LIM>(g, a, i, j, k);
}

// Intermediate fields of the inner advective step (ffsl_mod.F90:354-406)
struct FfslQArgs { const double *q, *qmfx, *qmfy, *divx, *divy, *pole; double *qx, *qy; double dt; int jy0, jy1; /* rows of qy */ };
template <bool EDGE> __device__ __forceinline__ void ffsl_q_body(const Geo& g, const DMesh& M, const FfslQArgs& a, int i, int j, int k) {
  const int n = g.nlon, c = gidx(g, i, j, k);
  const RowCoef* __restrict__ R = M.rows + (j + 1);
  const double q = __ldg(a.q + c);
  if (j > 1 && j < g.nlat) {
    const bool doy = j >= a.jy0 && j <= a.jy1;
    const double fx0 = __ldg(a.qmfx + c), fxm = __ldg(a.qmfx + c + (wrapE<EDGE>(i - 1, n) - i)), dx = __ldg(a.divx + c);
    double fy0 = 0, fys = 0, dy = 0;
    if (doy) { fy0 = __ldg(a.qmfy + c); fys = __ldg(a.qmfy + c - n); dy = __ldg(a.divy + c); }
    const double rac = R->r_area_cell;
    a.qx[c] = q - 0.5 * ((fx0 - fxm) * R->le_lon * rac - dx * q) * a.dt;
    if (doy) a.qy[c] = q - 0.5 * ((fy0 * R->le_lat - fys * R[-1].le_lat) * rac - dy * q) * a.dt;
  } else {
    const double pole = __ldg(a.pole + (j == 1 ? 0 : g.nlev) + k);
    const double t = 0.5 * (pole - __ldg(a.divy + c) * q) * a.dt;
    a.qx[c] = q;
    a.qy[c] = j == 1 ? q - t : q + t;
  }
}
__global__ void __launch_bounds__(128) k_ffsl_q(Geo g, DMesh M, FfslQArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y * blockDim.y + threadIdx.y, k = blockIdx.z;
  if (i >= g.nlon || j > g.je) return;
  if (edge_block(g.nlon, 1)) ffsl_q_body<true>(g, M, a, i, j, k); else ffsl_q_body<false>(g, M, a, i, j, k);
}

// ---- FFSL vertical PPM flux (vflx_ppm_cell, ffsl_mod.F90:883-941) --------------------
// Column accessor with the two ghost-level rules: GH_LINEAR = linear extrapolation
// (pt, operators_mod.F90:1085-1090), GH_COPY = copies of the end levels (tracers,
// adv_mod.F90:260-265).  l is the 1-based full level, valid for -1..nlev+2.
enum { GH_LINEAR = 0, GH_COPY = 1 };
struct ColQ {
  const double* p; long long st; int nlev, mode;
  __device__ __forceinline__ double in(int l) const { return p[(long long)(l - 1) * st]; }   // plain load: p may be shared memory
  __device__ double operator()(int l) const {
    if (l >= 1 && l <= nlev) return in(l);
    if (mode == GH_COPY) return l < 1 ? in(1) : in(nlev);
    if (l == 0) return 2 * in(1) - in(2);
    if (l == -1) return 2 * (2 * in(1) - in(2)) - in(1);
    if (l == nlev + 1) return 2 * in(nlev) - in(nlev - 1);
    return 2 * (2 * in(nlev) - in(nlev - 1)) - in(nlev);
  }
};
// flux through half level k1 (2..nlev)
template <int LIM> __device__ __forceinline__ double vflx_one(const ColQ& q, int k1, double cz, double w) {
  if (cz == 0) return 0.0;
  int ci = (int)cz;
  if (ci > q.nlev || ci < -q.nlev) ci = 0;      // only reachable after a raised cflz guard (garbage in, bounded loop out)
  const double cf = cz - ci;
  int ku; double s1, s2, sm = 0;
  if (cz > 0) { ku = k1 - ci - 1; s1 = 1 - cf; s2 = 1; for (int l = k1 - ci; l <= k1 - 1; ++l) sm += q(l); }
  else { ku = k1 - ci; s1 = 0; s2 = -cf; for (int l = k1; l <= k1 - ci - 1; ++l) sm += q(l); }
  if (LIM >= 4) return vl_flux<LIM>(q(ku - 1), q(ku), q(ku + 1), sm, cf, cz, w);
  double ml = 0, dm = 0, m6 = 0;
  if (ku >= 1 && ku <= q.nlev) ppm<LIM>(q(ku - 2), q(ku - 1), q(ku), q(ku + 1), q(ku + 2), ml, dm, m6);
  const double integ = ppm_integral(sm, ml, dm, m6, s1, s2);
  return qdiv((cz > 0 ? w : -w) * integ, cz);
}

// calc_grad_ptf tail (operators_mod.F90:1040-1112) + dpt (gmcore_mod.F90:490-495) + the
// raw pt*m update of update_state (time_schemes_mod.F90:150-156), one thread per column.
struct PtColArgs {
  const double *pt, *ptf_lon, *ptf_lat, *pole_ptf, *we_lev, *cflz;   // star state / mass batch
  const double *pt_old, *m_old;
  double *ptm_new;      // new%pt (raw pt*m, filtered and divided by m afterwards)
  double *dpt;          // optional (parity tests), may be NULL
  double dt;
};
template <int LIM> __global__ void __launch_bounds__(128) k_pt_col(Geo g, DMesh M, PtColArgs a) {
  extern __shared__ double pt_s[];       // [nlev][blockDim.x]: the pt column, loaded once
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y;
  if (i >= g.nlon) return;
  const bool pole = (j == 1 || j == g.nlat);
  const int st = (int)g.plane, B = blockDim.x, n = g.nlon, c0 = gidx(g, i, j, 0), im = wrapi(i - 1, n) - i;
  const RowCoef* __restrict__ R = M.rows + (j + 1);
  const double rac = R->r_area_cell, lel = R->le_lon, lea = R->le_lat, leas = R[-1].le_lat;
  double* ps = pt_s + threadIdx.x;
#pragma unroll 8
  for (int k = 0; k < g.nlev; ++k) ps[k * B] = __ldg(a.pt + c0 + k * st);
  ColQ q{ps, (long long)B, g.nlev, GH_LINEAR};
  double f_up = 0.0;   // ptf_lev(k), k = 1 -> 0
  for (int k = 0; k < g.nlev; ++k) {
    const int c = c0 + k * st;
    // all of this level's loads first, then the PPM arithmetic
    const int k1 = k + 2;   // lower half level of full level k+1
#ifndef PTCOL_PF
#define PTCOL_PF 3
#endif
    if (PTCOL_PF && (i & 15) == 0 && k + PTCOL_PF < g.nlev) {
      const int cp = c + PTCOL_PF * st;
      pf_l2(a.cflz + cp + st); pf_l2(a.we_lev + cp + st); pf_l2(a.ptf_lon + cp); pf_l2(a.ptf_lat + cp); pf_l2(a.pt_old + cp); pf_l2(a.m_old + cp);
    }
    double cz = 0, w = 0, fx0 = 0, fxm = 0, fy0 = 0, fys = 0, pdm = 0;
    if (k1 <= g.nlev) { cz = __ldg(a.cflz + c + st); w = __ldg(a.we_lev + c + st); }
    if (pole) pdm = __ldg(a.pole_ptf + (j == 1 ? 0 : g.nlev) + k);
    else { fx0 = __ldg(a.ptf_lon + c); fxm = __ldg(a.ptf_lon + c + im); fy0 = __ldg(a.ptf_lat + c); fys = __ldg(a.ptf_lat + c - n); }
    const double po = __ldg(a.pt_old + c), mo = __ldg(a.m_old + c);
    const double dlon = pole ? 0.0 : (fx0 - fxm) * lel * rac;
    const double dlat = pole ? pdm : (fy0 * lea - fys * leas) * rac;
    const double f_dn = k1 <= g.nlev ? vflx_one<LIM>(q, k1, cz, w) : 0.0;
    const double dlev = f_dn - f_up;
    f_up = f_dn;
    const double dpt = -dlon - dlat - dlev;
    if (a.dpt) a.dpt[c] = dpt;
    a.ptm_new[c] = po * mo + a.dt * dpt;
  }
}

// ---- calc_ph + calc_m (operators_mod.F90:136-168,471-543) from the filtered phs ------
// ph_lev / ph are never stored: they are affine in phs_f and re-formed where needed.
struct PhArgs { const double* phs_f; double *ph_exn_lev, *m, *m_lev; VC vc; int* flags; };
// PH_COL 1 (default): one thread per column walking down the half levels with the three interface pressures a level
// needs in registers, so that every ph_lev is formed once per column instead of up to six times per point and the
// index arithmetic is paid once (the per-point version below spent 110 of its ~235 instructions outside pow).
// Same expressions per point: bit-identical to the per-point kernel (grid z = nlev + 1), kept for A/B runs.
#ifndef PH_COL
#define PH_COL 1
#endif
#if PH_COL
__global__ void __launch_bounds__(128) k_ph(Geo g, DMesh M, PhArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y;
  if (i >= g.nlon) return;
  const int st = (int)g.plane, nlev = g.nlev, c0 = gidx(g, i, j, 0);
  const double phs = __ldg(a.phs_f + c0);
  double plm = 0.0, pl = ph_lev_at(M, a.vc, 1, phs), plp = ph_lev_at(M, a.vc, 2, phs);      // ph_lev(k), (k+1), (k+2)
  bool bad = false;
  for (int k = 0; k <= nlev; ++k) {                    // half level k+1
    const int c = c0 + k * st;
    a.ph_exn_lev[c] = k + 1 <= M.nconst ? __ldg(M.lev_tab + k) : GM_POW(pl, a.vc.kappa);
    double ml;                                         // m_lev_at(k + 1)
    if (k == 0) ml = 0.5 * (pl + plp) - pl;
    else if (k == nlev) ml = pl - 0.5 * (plm + pl);
    else ml = 0.5 * (pl + plp) - 0.5 * (plm + pl);
    a.m_lev[c] = ml;
    if (k < nlev) { const double m = plp - pl; a.m[c] = m; bad |= m <= 0; }
    plm = pl; pl = plp;
    plp = k + 3 <= nlev + 1 ? ph_lev_at(M, a.vc, k + 3, phs) : 0.0;
  }
  if (bad) atomicOr(a.flags, 8);
}
#else
__global__ void __launch_bounds__(128) k_ph(Geo g, DMesh M, PhArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y, k = blockIdx.z;     // k = 0..nlev (half level k+1)
  if (i >= g.nlon) return;
  const double phs = __ldg(a.phs_f + gidx(g, i, j, 0));
  const long long c = gidx(g, i, j, k);
  const double pl = ph_lev_at(M, a.vc, k + 1, phs);
  a.ph_exn_lev[c] = k + 1 <= M.nconst ? __ldg(M.lev_tab + k) : GM_POW(pl, a.vc.kappa);
  a.m_lev[c] = m_lev_at(M, a.vc, g.nlev, k + 1, phs);
  if (k < g.nlev) {
    const double m = ph_lev_at(M, a.vc, k + 2, phs) - pl;
    a.m[c] = m;
    if (m <= 0) atomicOr(a.flags, 8);
  }
}
#endif

// Per-level table of the pure-pressure levels (DMesh.lev_tab); phs is irrelevant there (hybi == 0).
__global__ void k_lev_tab(DMesh M, VC vc, int nlev, double* tab) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k > nlev) return;
  const double rp0 = 1.0 / vc.p0;
  const double pl0 = ph_lev_at(M, vc, k + 1, vc.p0);
  tab[k] = GM_POW(pl0, vc.kappa);
  if (k < nlev) {
    const double pl1 = ph_lev_at(M, vc, k + 2, vc.p0);
    const double ph = 0.5 * (pl0 + pl1);
    tab[nlev + 1 + k] = GM_POW(ph * rp0, vc.kappa);
    tab[2 * (nlev + 1) + k] = GM_LOG(qdiv(pl1, pl0));
  }
}

// ---- calc_t + calc_gz_lev (operators_mod.F90:170-190,433-469) ------------------------
// One thread per column, walking UP from the surface with a running sum: gz_lev(k) = gzs +
// sum_{l=k}^{nlev} Rd t(l) log(ph_lev(l+1)/ph_lev(l)).  The reference adds each level's terms in
// ascending l (an O(nlev^2) loop); the running suffix re-associates that sum (<= nlev ulps of the
// column thickness, the same size as the libm-vs-CUDA pow/log differences) and needs no scratch.
struct GzArgs { const double *pt, *phs_f, *gzs; double *t, *gz_lev; VC vc; };
__global__ void __launch_bounds__(128) k_gz_col(Geo g, DMesh M, GzArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y;
  if (i >= g.nlon) return;
  const int st = (int)g.plane, c0 = gidx(g, i, j, 0);
  const double phs = __ldg(a.phs_f + c0);
  const double gzs = __ldg(a.gzs + c0);
  const double rp0 = 1.0 / a.vc.p0;
  a.gz_lev[c0 + g.nlev * st] = gzs;    // T14: bottom interface = surface geopotential
  double pl1 = ph_lev_at(M, a.vc, g.nlev + 1, phs), dgz = 0.0;
  for (int k = g.nlev - 1; k >= 0; --k) {
    const double pl0 = ph_lev_at(M, a.vc, k + 1, phs);
    const double ph = 0.5 * (pl0 + pl1);
    const bool cst = k + 2 <= M.nconst;       // both interfaces of this layer are pure-pressure levels
    const double t = __ldg(a.pt + c0 + k * st) * (cst ? __ldg(M.lev_tab + (g.nlev + 1) + k) : GM_POW(ph * rp0, a.vc.kappa));
    a.t[c0 + k * st] = t;
    dgz = dgz + a.vc.Rd * t * (cst ? __ldg(M.lev_tab + 2 * (g.nlev + 1) + k) : GM_LOG(qdiv(pl1, pl0)));
    a.gz_lev[c0 + k * st] = gzs + dgz;
    pl1 = pl0;
  }
}

// ---- div_damp_run, order 2 (div_damp_mod.F90:89-139); div is recomputed from the
// filtered winds by k_div first (calc_div, operators_mod.F90:354-412). ---------------
struct DivArgs { const double *u, *v; double* div; };
__global__ void __launch_bounds__(128) k_div(Geo g, DMesh M, DivArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y, k = blockIdx.z;
  if (i >= g.nlon || j == 1 || j == g.nlat) return;
  const long long c = gidx(g, i, j, k);
  a.div[c] = ((__ldg(a.u + c) * MJ(M.le_lon, j) - ldW(a.u, g, i - 1, j, k) * MJ(M.le_lon, j)) +
              (__ldg(a.v + c) * MJ(M.le_lat, j) - ldW(a.v, g, i, j - 1, k) * MJ(M.le_lat, j - 1))) / MJ(M.area_cell, j);
}
// div2 = Laplacian of div for 4th-order damping (calc_div, operators_mod.F90:414-428); pole rows stay 0
struct Div2Args { const double* div; double* div2; };
__global__ void __launch_bounds__(128) k_div2(Geo g, DMesh M, Div2Args a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y, k = blockIdx.z;
  if (i >= g.nlon || j == 1 || j == g.nlat) return;
  const int c = gidx(g, i, j, k);
  const double d0 = __ldg(a.div + c), de = MJ(M.de_lon, j);
  a.div2[c] = (ldW(a.div, g, i + 1, j, k) - 2 * d0 + ldW(a.div, g, i - 1, j, k)) / (de * de) +
              ((ldW(a.div, g, i, j + 1, k) - d0) * MJ(M.half_cos_lat, j) / MJ(M.de_lat, j) -
               (d0 - ldW(a.div, g, i, j - 1, k)) * MJ(M.half_cos_lat, j - 1) / MJ(M.de_lat, j - 1)) / MJ(M.le_lon, j) / MJ(M.full_cos_lat, j);
}
struct DivDampArgs { const double *div, *c_lon, *c_lat; double *u, *v; double sgn; };
__global__ void __launch_bounds__(128) k_div_damp(Geo g, DMesh M, DivDampArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y, k = blockIdx.z;
  if (i >= g.nlon) return;
  const long long c = gidx(g, i, j, k);
  const double d0 = __ldg(a.div + c);
  if (j > 1 && j < g.nhalf) {      // interior rows: every load before the first store (one memory round trip, same arithmetic)
    const double de = ldW(a.div, g, i + 1, j, k), dn = __ldg(a.div + c + g.nlon), uo = a.u[c], vo = a.v[c];
    const double cl = __ldg(a.c_lon + (size_t)k * g.nlat + (j - 1)), ca = __ldg(a.c_lat + (size_t)k * g.nhalf + (j - 1));
    a.u[c] = uo + a.sgn * (cl * (de - d0) / MJ(M.de_lon, j));
    a.v[c] = vo + a.sgn * (ca * (dn - d0) / MJ(M.de_lat, j));
    return;
  }
  if (j > 1 && j < g.nlat)
    a.u[c] = a.u[c] + a.sgn * (__ldg(a.c_lon + (size_t)k * g.nlat + (j - 1)) * (ldW(a.div, g, i + 1, j, k) - d0) / MJ(M.de_lon, j));
  if (j <= g.nhalf)
    a.v[c] = a.v[c] + a.sgn * (__ldg(a.c_lat + (size_t)k * g.nhalf + (j - 1)) * (ldW(a.div, g, i, j + 1, k) - d0) / MJ(M.de_lat, j));
}

// ---- smag_damp_run (smag_damp_mod.F90:45-205): tension / shear strain -> eddy viscosity on the
// edges -> Laplacian damping of u, v.  kmh_lon / kmh_lat are formed on the fly (never stored); the
// tendencies are formed from the un-updated winds and applied by a second pass, as in the reference.
struct SmagArgs { const double *u, *v; double *st, *ss, *du, *dv; double coef, dt; };
__global__ void __launch_bounds__(128) k_smag_strain(Geo g, DMesh M, SmagArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y, k = blockIdx.z;
  if (i >= g.nlon) return;
  const int c = gidx(g, i, j, k);
  if (j > 1 && j < g.nlat)
    a.st[c] = (__ldg(a.u + c) - ldW(a.u, g, i - 1, j, k)) / MJ(M.de_lon, j) -
              (__ldg(a.v + c) * MJ(M.half_cos_lat, j) - ldW(a.v, g, i, j - 1, k) * MJ(M.half_cos_lat, j - 1)) / MJ(M.le_lon, j) / MJ(M.full_cos_lat, j);
  if (j <= g.nhalf)
    a.ss[c] = (ldW(a.v, g, i + 1, j, k) - __ldg(a.v + c)) / MJ(M.le_lat, j) +
              (ldW(a.u, g, i, j + 1, k) * MJ(M.full_cos_lat, j + 1) - __ldg(a.u + c) * MJ(M.full_cos_lat, j)) / MJ(M.de_lat, j) / MJ(M.half_cos_lat, j);
}
__global__ void __launch_bounds__(128) k_smag_tend(Geo g, DMesh M, SmagArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y, k = blockIdx.z;
  if (i >= g.nlon) return;
  const int c = gidx(g, i, j, k);
  const double t0 = __ldg(a.st + c), s0 = j <= g.nhalf ? __ldg(a.ss + c) : 0.0;
  if (j > 1 && j < g.nlat) {
    const double de = MJ(M.de_lon, j), le = MJ(M.le_lon, j);
    const double ls2 = a.coef / (1 / (de * de) + 1 / (le * le)) * 1.0;      // decay_from_top == 1 (:35-36)
    const double tp = ldW(a.st, g, i + 1, j, k), sm = ldW(a.ss, g, i, j - 1, k);
    const double kmh = ls2 * sqrt(0.5 * (t0 * t0 + tp * tp) + 0.5 * (s0 * s0 + sm * sm));
    const double u0 = __ldg(a.u + c);
    a.du[c] = kmh * ((ldW(a.u, g, i - 1, j, k) - 2 * u0 + ldW(a.u, g, i + 1, j, k)) / (de * de) +
                     ((ldW(a.u, g, i, j + 1, k) - u0) / MJ(M.de_lat, j) * MJ(M.half_cos_lat, j) -
                      (u0 - ldW(a.u, g, i, j - 1, k)) / MJ(M.de_lat, j - 1) * MJ(M.half_cos_lat, j - 1)) / le / MJ(M.full_cos_lat, j));
  }
  if (j <= g.nhalf) {
    const double le = MJ(M.le_lat, j), de = MJ(M.de_lat, j);
    const double ls2 = a.coef / (1 / (le * le) + 1 / (de * de)) * 1.0;
    const double tn = ldW(a.st, g, i, j + 1, k), sw = ldW(a.ss, g, i - 1, j, k);
    const double kmh = ls2 * sqrt(0.5 * (t0 * t0 + tn * tn) + 0.5 * (s0 * s0 + sw * sw));
    const double v0 = __ldg(a.v + c);
    double lap = (ldW(a.v, g, i - 1, j, k) - 2 * v0 + ldW(a.v, g, i + 1, j, k)) / (le * le);
    if (!(j == 1 || j == g.nhalf))
      lap = lap + ((ldW(a.v, g, i, j + 1, k) - v0) / MJ(M.le_lon, j + 1) * MJ(M.full_cos_lat, j + 1) -
                   (v0 - ldW(a.v, g, i, j - 1, k)) / MJ(M.le_lon, j) * MJ(M.full_cos_lat, j)) / de / MJ(M.half_cos_lat, j);
    a.dv[c] = kmh * lap;
  }
}
__global__ void __launch_bounds__(128) k_smag_apply(Geo g, SmagArgs a, double* u, double* v) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y, k = blockIdx.z;
  if (i >= g.nlon) return;
  const int c = gidx(g, i, j, k);
  if (j > 1 && j < g.nlat) u[c] = u[c] + a.dt * a.du[c];
  if (j <= g.nhalf) v[c] = v[c] + a.dt * a.dv[c];
}

// ---- tracer batch (adv_batch_mod.F90:207-252,379-403; adv_mod.F90:152-312) -----------
// wind / mass-flux accumulation: mode 0 copy, 1 add, 2 add and divide by nstep+1
struct AccumArgs { double* dst[2]; const double* src[2]; int nlev_[2]; int mode, nstep; };
__global__ void __launch_bounds__(128) k_accum(Geo g, AccumArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y, k = blockIdx.z;
  if (i >= g.nlon) return;
  const long long c = gidx(g, i, j, k);
  // both pairs' loads before the first store (the second pair's accumulator read would otherwise wait for it)
  bool on[2]; double s[2], d[2];
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    on[q] = a.dst[q] && k < a.nlev_[q];
    s[q] = on[q] ? __ldg(a.src[q] + c) : 0.0;
    d[q] = on[q] && a.mode != 0 ? a.dst[q][c] : 0.0;
  }
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    if (!on[q]) continue;
    if (a.mode == 0) a.dst[q][c] = s[q];
    else { const double t = d[q] + s[q]; a.dst[q][c] = a.mode == 2 ? t / (a.nstep + 1) : t; }
  }
}

// horizontal update q* = (old_m q - dt div F) / m, pole rows by zonal means (:183-233)
struct TrHArgs { const double *qo, *old_m, *m, *qmf_lon, *qmf_lat, *pole; double* qn; double dt; };
__global__ void __launch_bounds__(128) k_tracer_h(Geo g, DMesh M, TrHArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y, k = blockIdx.z;
  if (i >= g.nlon) return;
  const long long c = gidx(g, i, j, k);
  const double mq = __ldg(a.old_m + c) * __ldg(a.qo + c);
  const double mnew = __ldg(a.m + c);       // loaded with the others, not after the flux divergence
  double r;
  if (j > 1 && j < g.nlat)
    r = mq - ((__ldg(a.qmf_lon + c) - ldW(a.qmf_lon, g, i - 1, j, k)) * MJ(M.le_lon, j) +
              (__ldg(a.qmf_lat + c) * MJ(M.le_lat, j) - ldW(a.qmf_lat, g, i, j - 1, k) * MJ(M.le_lat, j - 1))) / MJ(M.area_cell, j) * a.dt;
  else {
    const double pole = __ldg(a.pole + (j == 1 ? 0 : g.nlev) + k) * a.dt;
    r = j == 1 ? mq - pole : mq + pole;
  }
  a.qn[c] = r / mnew;
}

// sequential negative filling along x (:236-258); one block per (row, level), the sweep
// itself only runs on rows that contain a negative value.  The reference reads/writes
// the (stale) lon-halo cells at i=1 / i=nlon; here the row is periodic.
struct NegXArgs { double* q; const double* m; int* flags; };
__global__ void __launch_bounds__(128) k_negfill_x(Geo g, NegXArgs a) {
  const int j = g.jb + blockIdx.x, k = blockIdx.y;
  double* q = a.q + gidx(g, 0, j, k); const double* m = a.m + gidx(g, 0, j, k);
  int neg = 0;
  for (int i0 = threadIdx.x; i0 < g.nlon; i0 += 4 * blockDim.x) {      // four loads in flight per thread
    double v[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) { const int i = i0 + r * blockDim.x; v[r] = i < g.nlon ? q[i] : 0.0; }
#pragma unroll
    for (int r = 0; r < 4; ++r) neg |= v[r] < 0;
  }
  neg = __syncthreads_or(neg);
  if (!neg || threadIdx.x != 0) return;
  const int n = g.nlon;
  for (int i = 0; i < n; ++i) {
    if (q[i] < 0) {
      const int im = i == 0 ? n - 1 : i - 1, ip = i == n - 1 ? 0 : i + 1;
      const double qm0 = q[i] * m[i], qm1 = q[im] * m[im], qm2 = q[ip] * m[ip], h = 0.5 * qm0;
      if (qm1 >= h && qm2 >= h) { q[im] = (qm1 - h) / m[im]; q[ip] = (qm2 - h) / m[ip]; }
      else if (qm1 > qm0) q[im] = (qm1 - qm0) / m[im];
      else if (qm2 > qm0) q[ip] = (qm2 - qm0) / m[ip];
      else atomicOr(a.flags, 16);
      q[i] = 0;
    }
  }
}

// vertical flux + update + negative filling in z (:259-297), one thread per column,
// the column staged in shared memory (the update is in place).
struct TrVArgs { double* q; const double *m, *we, *cflz; double dt; int* flags; };
template <int LIM> __global__ void __launch_bounds__(128) k_tracer_v(Geo g, TrVArgs a) {
  extern __shared__ double col[];     // [nlev][blockDim.x]: the old q column
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y;
  if (i >= g.nlon) return;
  const int st = (int)g.plane, c0 = gidx(g, i, j, 0);
  const int B = blockDim.x, nlev = g.nlev;
  double* qs = col + threadIdx.x;
#pragma unroll 8
  for (int k = 0; k < nlev; ++k) qs[k * B] = a.q[c0 + k * st];
  ColQ q{qs, (long long)B, nlev, GH_COPY};
  double f_up = 0;
  bool neg = false;
  for (int k = 0; k < nlev; ++k) {
    const int k1 = k + 2, c = c0 + k * st;
    double cz = 0, w = 0;
    if ((i & 15) == 0 && k + 3 < nlev) {     // one lane per 128-byte line: level k+3 into L2 (as k_pt_col)
      const int cp = c + 3 * st;
      pf_l2(a.cflz + cp + st); pf_l2(a.we + cp + st); pf_l2(a.m + cp);
    }
    if (k1 <= nlev) { cz = __ldg(a.cflz + c + st); w = __ldg(a.we + c + st); }
    const double m = __ldg(a.m + c);
    const double f_dn = k1 <= nlev ? vflx_one<LIM>(q, k1, cz, w) : 0.0;
    const double qn = qs[k * B] - qdiv((f_dn - f_up) * a.dt, m);
    a.q[c] = qn;                       // the update is in place: the stencil reads the staged column
    neg = neg || qn < 0;
    f_up = f_dn;
  }
  if (!neg) return;
  // negative filling in z (adv_mod.F90:275-297): rare, sequential, on this thread's own column
  double* qn = a.q + c0; const double* mm_ = a.m + c0;
  for (int k = 0; k < nlev; ++k) {
    if (qn[k * st] < 0) {
      const double m0 = mm_[k * st];
      const double mm = k > 0 ? mm_[(k - 1) * st] : 0.0, mp = k < nlev - 1 ? mm_[(k + 1) * st] : 0.0;
      const double qm0 = qn[k * st] * m0, qm1 = k > 0 ? qn[(k - 1) * st] * mm : 0.0, qm2 = k < nlev - 1 ? qn[(k + 1) * st] * mp : 0.0;
      const double h = 0.5 * qm0;
      if (qm1 >= h && qm2 >= h) { if (qm1 > 0) qn[(k - 1) * st] = (qm1 - h) / mm; if (qm2 > 0) qn[(k + 1) * st] = (qm2 - h) / mp; }
      else if (qm1 > qm0) { if (qm1 > 0) qn[(k - 1) * st] = (qm1 - qm0) / mm; }
      else if (qm2 > qm0) { if (qm2 > 0) qn[(k + 1) * st] = (qm2 - qm0) / mp; }
      else atomicOr(a.flags, 16);
      qn[k * st] = 0;
    }
  }
}

// 3 x lon_damp_on_cell(order 2) (lon_damp_mod.F90:158-177,231-236), one block per row
struct LonDampArgs { double* q; const double* wgt; /* [nlat+1], entry j */ int npass; };
__global__ void __launch_bounds__(256) k_lon_damp(Geo g, LonDampArgs a) {
  extern __shared__ double row[];     // [2][nlon]
  const int j = g.jb + blockIdx.x, k = blockIdx.y;
  if (j == 1 || j == g.nlat) return;
  const double wgt = __ldg(a.wgt + j);
  if (wgt == 0) return;
  double* q = a.q + gidx(g, 0, j, k);
  const int n = g.nlon;
  double *f = row, *gn = row + n;
  for (int i = threadIdx.x; i < n; i += blockDim.x) f[i] = q[i];
  __syncthreads();
  const double c = 0.25;
  for (int p = 0; p < a.npass; ++p) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const double fm = f[i == 0 ? n - 1 : i - 1], f0 = f[i], fp = f[i == n - 1 ? 0 : i + 1];
      gn[i] = wgt * (f0 + c * (fm - 2.0 * f0 + fp)) + (1 - wgt) * f0;
    }
    __syncthreads();
    double* t = f; f = gn; gn = t;
  }
  for (int i = threadIdx.x; i < n; i += blockDim.x) q[i] = f[i];
}

// ---- Held-Suarez forcing (held_suarez_test_mod.F90:65-113) ---------------------------
struct HsArgs { double *u, *v, *pt; const double *t, *phs_f; double dt; VC vc; };
__global__ void __launch_bounds__(128) k_held_suarez(Geo g, DMesh M, HsArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int j = g.jb + blockIdx.y, k = blockIdx.z;
  if (i >= g.nlon) return;
  const long long c = gidx(g, i, j, k);
  const double sig_b = 0.7, kf = 1.0 / 86400.0, ka = 0.025 / 86400.0, ks = 0.25 / 86400.0;
  const double fl = M.full_lev[k + 2];      // full_lev(k+1), stored from -1
  const double prof = fmax(0.0, (fl - sig_b) / (1.0 - sig_b));
  const double kv = kf * prof;
  if (j > 1 && j < g.nlat) a.u[c] = a.u[c] - a.dt * kv * a.u[c];
  if (j <= g.nhalf) a.v[c] = a.v[c] - a.dt * kv * a.v[c];
  const double cl = MJ(M.full_cos_lat, j), sn = MJ(M.full_sin_lat, j);
  const double c4 = (cl * cl) * (cl * cl);
  const double kt = ka + (ks - ka) * prof * c4;
  const double phs = __ldg(a.phs_f + gidx(g, i, j, 0));
  const double ph = 0.5 * (ph_lev_at(M, a.vc, k + 1, phs) + ph_lev_at(M, a.vc, k + 2, phs));
  const double p_p0 = ph / a.vc.p0;
  const double teq = fmax(200.0, (315.0 - 60.0 * (sn * sn) - 10.0 * GM_LOG(p_p0) * (cl * cl)) * GM_POW(p_p0, a.vc.kappa));
  a.pt[c] = a.pt[c] - a.dt * kt * a.pt[c] * (1.0 - teq / __ldg(a.t + c));
}

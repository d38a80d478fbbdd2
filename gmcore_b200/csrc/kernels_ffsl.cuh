// FFSL horizontal sweeps with shared limited slopes (round 2).
//
// hflx_ppm_cell (ffsl_mod.F90:778-881) builds the PPM coefficients of EVERY cell once (three mono-limited slopes
// per cell and direction, :813-826) and then lets each edge pick the coefficients of its upwind cell.  The round-1
// kernel (k_hflx, kernels_hydro.cuh) re-formed those coefficients per EDGE from a 5-point window -- three limited
// slopes per edge and direction, every cell's slope evaluated three times by three neighbouring threads -- and was
// bound by FP64 / issue slots (432 instructions per point, FP64 pipe 64 %), not by bytes.  Here every cell's slope is
// formed ONCE and shared, and the two directions are separate kernels, because they share nothing but the inner
// call's tracer field (x: wind u, cflx, divx; y: wind v, cfly, divy; the outer call sweeps qy in x and qx in y):
//   k_ffsl_y   marches over a strip of rows with the y window of its column in REGISTERS (5 values, 4 slopes, rolled
//              by one row per iteration: one new load and one new slope per row); no shared memory, no barrier;
//   k_ffsl_x   one block per row segment: the row is staged in shared memory with 3 apron cells either side, each
//              thread forms the slope of its own cell, each edge reads the 3 slopes / values of its upwind cell.
// The inner call of ffsl_calc_tracer_hflx_cell (:354-372) writes the half-step fields qx, qy directly: the inner
// fluxes never go to HBM and the separate k_ffsl_q pass is gone (only the two pole rows, whose qy needs a zonal mean
// of the pole-adjacent y fluxes, :375-406, are finished by k_ffsl_q).  Lanes with |cflx| >= 1 (polar rows) take the
// round-1 gather path (own 5-point stencil + whole-cell segment sum, :831-848).
// Arithmetic is operation-for-operation the round-1 one (same slope, edge-value and integral expressions): results
// are bit-identical to k_hflx + k_ffsl_q.
//   algorithmic passes per call: inner x 4+1 (u, q, cflx, divx -> qx), inner y 4+1 (v, q, cfly, divy -> qy),
//                                outer x 3+1 (mfx, qy, cflx -> flux), outer y 3+1 (mfy, qx, cfly -> flux)   = 18
//   (round 1: inner 5+2, k_ffsl_q 5+2, outer 6+2 = 22;  a fused x+y marching kernel, 9 + 8 = 17 passes, is kept below
//   as k_ffsl_march_* for A/B runs: it needs ~125 registers for both windows and measured slower)
#pragma once
#include "kernels_hydro.cuh"

struct FfslArgs {
  const double *u, *v;          // advecting wind (inner) or mass flux (outer)
  const double *mx, *my;        // field swept in x / in y
  const double *cflx, *cfly;
  double *ox, *oy;              // outer: mfx, mfy      inner: qx, qy
  const double *divx, *divy;    // inner only
  double *qmfy;                 // inner only: y flux on the two pole-adjacent half rows (for the zonal means)
  double dt;
  int jx0, jx1, jy0, jy1;       // rows of the x flux / half rows of the y flux (band mode)
  int jq0, jq1;                 // inner only: rows of qy
};

// limited slope of cell (f) from its two neighbours: exactly the slope expressions inside ppm() / vl_slope()
template <int MODE> __device__ __forceinline__ double cell_slope(double fm1, double f, double fp1) {
  return slope_mono_d<MODE>(f - fm1, fp1 - f, (fp1 - fm1) * 0.5, f);
}
// PPM edge values + parabola of one cell from its three slopes (ffsl_mod.F90:1156-1172; same expressions as ppm())
__device__ __forceinline__ void ppm_from_slopes(double fm1, double f, double fp1, double dfl, double dfc, double dfr,
                                                double& fl, double& df, double& f6) {
  fl = 0.5 * (fm1 + f) + (dfl - dfc) * (1.0 / 6.0);
  double fr = 0.5 * (fp1 + f) + (dfc - dfr) * (1.0 / 6.0);
  fl = f - fsign(dmin(fabs(dfc), fabs(fl - f)), dfc);
  fr = f + fsign(dmin(fabs(dfc), fabs(fr - f)), dfc);
  f6 = 6 * f - 3 * (fl + fr);
  df = fr - fl;
}


// ---- row-marching variant -------------------------------------------------------------------------------
// A block owns blockDim.x columns of one level and marches over a strip of rows.  The y sweep lives entirely in
// registers: every thread keeps the 5 values / 4 limited slopes of its own column that the two candidate upwind
// cells of the current half row need, and rolls them by one row per iteration (one new load and ONE new slope per
// row -- no y apron at all except the 5 start-up rows of a strip).  The x sweep shares the row through shared
// memory: the row of mx (+3 apron cells either side) is staged, every thread forms the slope of its own cell, and
// each edge reads the three slopes and values of its upwind cell.  Within a strip every load address is valid for
// every row (rows are clamped to the strip / to the y window), so the loop body carries no load predicates; which
// results are stored is decided by block-uniform row ranges.
#define FM_WMAX 256
#define FM_RJMAX 64        // rows per strip (ffsl_strip_rows stays below this)
// measured at 0.25 deg L32 (tools/kprof.py): outer 0.538 / 0.466 / 0.487 ms at 2 / 3 / 4 blocks of 256 per SM
// (109 / 72 / 64 registers), inner 0.741 / 0.788 / 0.932 ms (128 / 80 / 64 registers, the last two spill)
#ifndef FM_MINB_IN
#define FM_MINB_IN 2
#endif
#ifndef FM_MINB_OUT
#define FM_MINB_OUT 3
#endif
// |cflx| >= 1 (polar rows): own 5-point stencil around the upwind cell + whole-cell segment sum (ffsl_mod.F90:831-848)
template <int LIM> __device__ __noinline__ double ffsl_flux_x_gather(const double* __restrict__ mxr, int ii, int n, double cx, double uu) {
  int ci = (int)cx;
  const double cf = cx - ci; const bool pos = cx > 0;
  const int iu = wrapi(pos ? ii - ci : ii - ci + 1, n);
  const double x0 = __ldg(mxr + wrapi(iu - 2, n)), x1 = __ldg(mxr + wrapi(iu - 1, n)), x2 = __ldg(mxr + iu);
  const double x3 = __ldg(mxr + wrapi(iu + 1, n)), x4 = __ldg(mxr + wrapi(iu + 2, n));
  double sm = 0;
  if (pos) { for (int l = ii + 1 - ci; l <= ii; ++l) sm += __ldg(mxr + wrapi(l, n)); }
  else { for (int l = ii + 1; l <= ii - ci; ++l) sm += __ldg(mxr + wrapi(l, n)); }
  if (LIM >= 4) return vl_flux<LIM>(x1, x2, x3, sm, cf, cx, uu);
  const double s1 = pos ? 1 - cf : 0.0, s2 = pos ? 1.0 : -cf;
  double ml, dm, m6;
  ppm<LIM>(x0, x1, x2, x3, x4, ml, dm, m6);
  return qdiv((pos ? uu : -uu) * ppm_integral(sm, ml, dm, m6, s1, s2), cx);
}
// flux through the east edge of tile-local cell lc (global column ii) from the staged row xs / slopes ss
template <int LIM> __device__ __forceinline__ double ffsl_flux_x_smem(const double* __restrict__ xs, const double* __restrict__ ss,
                                                                      const double* __restrict__ mxr, int lc, int ii, int n, double cx, double uu) {
  if (cx == 0) return 0.0;
  const int ci = (int)cx;
  if (ci != 0 && ci < n && ci > -n) return ffsl_flux_x_gather<LIM>(mxr, ii, n, cx, uu);
  // (|cflx| >= nlon: the guard word is already up (k_cfl); treated as a fractional CFL to stay in the row)
  const double cf = cx;
  const bool pos = cx > 0;
  const int o = lc + (pos ? 0 : 1);                       // tile-local upwind cell
  const double f = xs[o + 3], dfc = ss[o + 2];
  if (LIM >= 4) { const double t = pos ? cf * (f + dfc * 0.5 * (1 - cf)) : cf * (f - dfc * 0.5 * (1 + cf)); return qdiv(uu * t, cx); }
  const double s1 = pos ? 1 - cf : 0.0, s2 = pos ? 1.0 : -cf;
  double ml, dm, m6;
  ppm_from_slopes(xs[o + 2], f, xs[o + 4], ss[o + 1], dfc, ss[o + 3], ml, dm, m6);
  return qdiv((pos ? uu : -uu) * ppm_integral(0.0, ml, dm, m6, s1, s2), cx);
}
template <int LIM, bool INNER> __device__ __forceinline__ void ffsl_march_body(const Geo& g, const DMesh& M, const FfslArgs& a, int rj) {
  __shared__ double xr[2][FM_WMAX + 6];     // x-swept field of the current row, cells x0-3 .. x0+W+2
  __shared__ double sxr[2][FM_WMAX + 4];    // its limited slopes, cells x0-2 .. x0+W+1
  __shared__ double fxr[INNER ? 2 : 1][INNER ? FM_WMAX + 1 : 1];    // INNER: x fluxes of the edges x0-1 .. x0+W-1
  __shared__ double rcf[INNER ? FM_RJMAX : 1][4];                   // INNER: row coefficients of the strip
  const int n = g.nlon, W = blockDim.x, tx = threadIdx.x, x0 = blockIdx.x * W, i = x0 + tx;
  const int rs = g.jb + blockIdx.y * rj;
  if (rs > g.je) return;
  const int re = rs + rj - 1 > g.je ? g.je : rs + rj - 1;
  const int r0 = INNER ? rs - 1 : rs;       // INNER: qy of the strip's first row needs the y flux below it
  const bool act = i < n;
  // idle lanes of the last block: the first three stand in for the wrapped cells n, n+1, n+2 of the x sweep (their
  // row values feed the slopes of the row's last cells), the rest read a valid column; none of them stores anything
  const int ic = act ? i : (i <= n + 2 ? i - n : n - 1);
  // block-uniform row ranges: x flux (and qx), y flux, qy; y windows may only read rows ylo .. yhi
  const int xlo = max(max(a.jx0, rs), 2), xhi = min(min(a.jx1, re), g.nlat - 1);
  const int flo = max(a.jy0, 1), fhi = min(a.jy1, g.nhalf);
  const int ylo = a.jy0 - 2, yhi = a.jy1 + 3;
  const bool direct = r0 - 2 >= 1 && re + 3 <= g.nlat && r0 - 2 >= ylo && re + 3 <= yhi;   // plain strided y windows
  const bool same = a.mx == a.my && rs >= ylo && re <= yhi;       // inner call: the row value is already in the y window
  const double* __restrict__ pmy = a.my; const double* __restrict__ pmx = a.mx;
  const double* __restrict__ pcx = a.cflx; const double* __restrict__ pcy = a.cfly;
  const double* __restrict__ pu = a.u; const double* __restrict__ pv = a.v;
  auto ldy = [&](int r) -> double {         // generic y-window load (strips touching a pole or the band's y range)
    if (r < ylo || r > yhi) return 0.0;
    return ldF(pmy, g, ic, r, blockIdx.z);
  };
  int c = gidx(g, ic, r0, blockIdx.z);      // element index of (ic, r, k); advances by n per row
  // apron cell of this thread (threads 0..5): x0-3 .. x0-1 and x0+W .. x0+W+2
  const int alc = tx < 3 ? tx - 3 : W + tx - 3;
  const bool hasap = tx < 6 && x0 + alc <= n + 2;
  const int dca = wrapi(x0 + alc, n) - ic;  // column offset of the apron cell from this thread's own column
  const int dcw = wrapi(x0 - 1, n) - ic;    // ... of the west edge x0-1 (thread 0, INNER)
  // ---- start-up: the y window of row r0 ----------------------------------------------------------------
  double fm1, f0, fp1, fp2, fp3, sm1, s0, sp1;
  {
    double fm2;
    if (direct) { const double* __restrict__ p = pmy + c; fm2 = __ldg(p - 2 * n); fm1 = __ldg(p - n); f0 = __ldg(p); fp1 = __ldg(p + n); fp2 = __ldg(p + 2 * n); fp3 = __ldg(p + 3 * n); }
    else { fm2 = ldy(r0 - 2); fm1 = ldy(r0 - 1); f0 = ldy(r0); fp1 = ldy(r0 + 1); fp2 = ldy(r0 + 2); fp3 = ldy(r0 + 3); }
    sm1 = cell_slope<LIM>(fm2, fm1, f0); s0 = cell_slope<LIM>(fm1, f0, fp1); sp1 = cell_slope<LIM>(f0, fp1, fp2);
  }
  // every operand of a row is loaded one iteration ahead (the loop body never waits on HBM)
  double xv = same ? 0.0 : __ldg(pmx + c), xa = hasap ? __ldg(pmx + c + dca) : 0.0;
  double cy = __ldg(pcy + c), vv = __ldg(pv + c), cx = __ldg(pcx + c), uu = __ldg(pu + c);
  double dvx = 0, dvy = 0, cxw = 0, uuw = 0;
  if (INNER) {
    dvx = __ldg(a.divx + c); dvy = __ldg(a.divy + c);
    if (tx == 0) { cxw = __ldg(pcx + c + dcw); uuw = __ldg(pu + c + dcw); }
    // this strip's row coefficients (le_lon, 1/area_cell, le_lat of rows r and r-1) once, in shared memory
    for (int t = tx; t <= re - rs; t += W) {
      const RowCoef* __restrict__ R = M.rows + (rs + t + 1);
      rcf[t][0] = R->le_lon; rcf[t][1] = R->r_area_cell; rcf[t][2] = R->le_lat; rcf[t][3] = R[-1].le_lat;
    }
    __syncthreads();
  }
  double fy_prev = 0;                        // INNER: y flux of half row r-1
  for (int r = r0; r <= re; ++r, c += n) {
    const int b = r & 1;
    const int cn = r < re ? c + n : c;      // next row, clamped to the strip: always a valid address
    // ---- loads for the NEXT row, and the y window's new row -----------------------------------------------
    const double ncy = __ldg(pcy + cn), nvv = __ldg(pv + cn), ncx = __ldg(pcx + cn), nuu = __ldg(pu + cn);
    double ndvx = 0, ndvy = 0, ncxw = 0, nuuw = 0;
    if (INNER) { ndvx = __ldg(a.divx + cn); ndvy = __ldg(a.divy + cn); if (tx == 0) { ncxw = __ldg(pcx + cn + dcw); nuuw = __ldg(pu + cn + dcw); } }
    const double nxv = same ? 0.0 : __ldg(pmx + cn), nxa = hasap ? __ldg(pmx + cn + dca) : 0.0;
    double fp4;
    if (direct) fp4 = __ldg(pmy + (r < re ? c + 4 * n : c + 3 * n));          // the last iteration re-reads row re+3 (unused)
    else fp4 = ldy(r + 4);
    const bool dox = r >= xlo && r <= xhi, doy = r >= flo && r <= fhi;          // block-uniform
    // ---- x sweep, part 1: stage the row ------------------------------------------------------------------
    const double qv = same ? f0 : xv;
    if (dox) { xr[b][tx + 3] = qv; if (hasap) xr[b][alc + 3] = xa; }
    // ---- y sweep: registers only --------------------------------------------------------------------------
    const double sp2 = cell_slope<LIM>(fp1, fp2, fp3);
    double fy = 0;
    if (doy) {
      const bool pos = cy > 0;
      const double f = pos ? f0 : fp1, dfc = pos ? s0 : sp1;
      if (LIM >= 4) fy = vv * (f + dfc * 0.5 * (copysign(1.0, cy) - cy));       // ffsl_mod.F90:605-611
      else if (cy != 0) {
        const double fa = pos ? fm1 : f0, fb = pos ? fp1 : fp2, dfl = pos ? sm1 : s0, dfr = pos ? sp1 : sp2;
        const double s1 = pos ? 1 - cy : 0.0, s2 = pos ? 1.0 : -cy;
        double ml, dm, m6;
        ppm_from_slopes(fa, f, fb, dfl, dfc, dfr, ml, dm, m6);
        fy = qdiv((pos ? vv : -vv) * ppm_integral(0.0, ml, dm, m6, s1, s2), cy);
      }
      if (!INNER) { if (act) a.oy[c] = fy; }
      else if (act && (r == 1 || r == g.nhalf)) a.qmfy[c] = fy;     // pole-adjacent half rows: zonal means of the pole rows
    }
    // ---- x sweep, part 2 -----------------------------------------------------------------------------------
    if (dox) {
      __syncthreads();
      sxr[b][tx + 2] = cell_slope<LIM>(xr[b][tx + 2], qv, xr[b][tx + 4]);
      if (tx < 4) { const int lc = tx < 2 ? tx - 2 : W + tx - 2; sxr[b][lc + 2] = cell_slope<LIM>(xr[b][lc + 2], xr[b][lc + 3], xr[b][lc + 4]); }
      __syncthreads();
      const double* __restrict__ mxr = pmx + (c - ic);
      const double fx = ffsl_flux_x_smem<LIM>(xr[b], sxr[b], mxr, tx, ic, n, cx, uu);
      if (!INNER) { if (act) a.ox[c] = fx; }
      else {
        fxr[b][tx + 1] = fx;
        if (tx == 0) fxr[b][0] = ffsl_flux_x_smem<LIM>(xr[b], sxr[b], mxr, -1, ic + dcw, n, cxw, uuw);
        __syncthreads();
        if (act) {
          const double* __restrict__ rc = rcf[r - rs];
          const double rac = rc[1];
          a.ox[c] = qv - 0.5 * ((fx - fxr[b][tx]) * rc[0] * rac - dvx * qv) * a.dt;
          if (r >= a.jq0 && r <= a.jq1) a.oy[c] = qv - 0.5 * ((fy * rc[2] - fy_prev * rc[3]) * rac - dvy * qv) * a.dt;
        }
      }
    }
    // ---- roll the window ----------------------------------------------------------------------------------
    fy_prev = fy;
    fm1 = f0; f0 = fp1; fp1 = fp2; fp2 = fp3; fp3 = fp4;
    sm1 = s0; s0 = sp1; sp1 = sp2;
    xv = nxv; xa = nxa; cy = ncy; vv = nvv; cx = ncx; uu = nuu;
    if (INNER) { dvx = ndvx; dvy = ndvy; cxw = ncxw; uuw = nuuw; }
  }
}
// named entry points (the profiler keys on the kernel name): inner = half-step fields, outer = fluxes
template <int LIM> __global__ void __launch_bounds__(FM_WMAX, FM_MINB_IN) k_ffsl_march_in(Geo g, DMesh M, FfslArgs a, int rj) { ffsl_march_body<LIM, true>(g, M, a, rj); }
template <int LIM> __global__ void __launch_bounds__(FM_WMAX, FM_MINB_OUT) k_ffsl_march_out(Geo g, DMesh M, FfslArgs a, int rj) { ffsl_march_body<LIM, false>(g, M, a, rj); }
// strips of about 40 rows: the 5 start-up rows of a strip cost ~12 % more loads of `my`, the grid keeps >= 4 waves
static inline int ffsl_strip_rows(const Geo& g) {
  const int rows = g.je - g.jb + 1;
  const int strips = (rows + 39) / 40;
  return (rows + strips - 1) / strips;
}
// block width: the multiple of 32 in 96..256 that wastes the fewest lanes on the last block of a row
static inline int ffsl_block_width(int nlon) {
  int best = 128, waste = 1 << 30;
  for (int w = 96; w <= FM_WMAX; w += 32) { const int x = (nlon + w - 1) / w * w - nlon; if (x < waste || (x == waste && w == 128)) { waste = x; best = w; } }
  return best;
}
static inline dim3 ffsl_march_grid(const Geo& g, int nk, int rj, int w) { return dim3((g.nlon + w - 1) / w, (g.je - g.jb + 1 + rj - 1) / rj, nk); }

// ---- y sweep: register marching, no shared memory ----------------------------------------------------------
#ifndef FY_MINB
#define FY_MINB 8          // 64 registers: measured 0.301 / 0.253 ms (inner / outer) against 0.330 / 0.289 at 6 blocks, 80 registers
#endif
template <int LIM, bool INNER> __device__ __forceinline__ void ffsl_y_body(const Geo& g, const DMesh& M, const FfslArgs& a, int rj) {
  const int n = g.nlon, i = blockIdx.x * blockDim.x + threadIdx.x, k = blockIdx.z;
  if (i >= n) return;
  const int rs = g.jb + blockIdx.y * rj;
  if (rs > g.je) return;
  const int re = rs + rj - 1 > g.je ? g.je : rs + rj - 1;
  const int r0 = INNER ? rs - 1 : rs;       // INNER: qy of the strip's first row needs the y flux below it
  const int flo = max(a.jy0, 1), fhi = min(a.jy1, g.nhalf);       // half rows of the y flux
  const int ylo = a.jy0 - 2, yhi = a.jy1 + 3;                     // rows the y windows may read (band mode)
  const int qlo = max(max(a.jq0, 2), rs), qhi = min(a.jq1, g.nlat - 1);   // INNER: rows of qy (never the pole rows)
  const bool direct = r0 - 2 >= 1 && re + 3 <= g.nlat && r0 - 2 >= ylo && re + 3 <= yhi;   // plain strided windows
  const double* __restrict__ pmy = a.my; const double* __restrict__ pcy = a.cfly; const double* __restrict__ pv = a.v;
  auto ldy = [&](int r) -> double {         // generic window load (strips touching a pole or the band's y range)
    if (r < ylo || r > yhi) return 0.0;
    return ldF(pmy, g, i, r, k);
  };
  int c = gidx(g, i, r0, k);                // element index of (i, r, k); advances by n per row
  double fm1, f0, fp1, fp2, fp3, sm1, s0, sp1;
  {
    double fm2;
    if (direct) { const double* __restrict__ p = pmy + c; fm2 = __ldg(p - 2 * n); fm1 = __ldg(p - n); f0 = __ldg(p); fp1 = __ldg(p + n); fp2 = __ldg(p + 2 * n); fp3 = __ldg(p + 3 * n); }
    else { fm2 = ldy(r0 - 2); fm1 = ldy(r0 - 1); f0 = ldy(r0); fp1 = ldy(r0 + 1); fp2 = ldy(r0 + 2); fp3 = ldy(r0 + 3); }
    sm1 = cell_slope<LIM>(fm2, fm1, f0); s0 = cell_slope<LIM>(fm1, f0, fp1); sp1 = cell_slope<LIM>(f0, fp1, fp2);
  }
  // the operands of a row are loaded one iteration ahead; addresses are clamped to the strip, so no load is predicated
  double cy = __ldg(pcy + c), vv = __ldg(pv + c), dvy = INNER ? __ldg(a.divy + c) : 0.0;
  double fy_prev = 0;
  // INNER: row coefficients le_lat(r), le_lat(r-1), 1 / area_cell(r), also one row ahead
  const RowCoef* __restrict__ Rw = M.rows + (r0 + 1);
  double lea = INNER ? Rw->le_lat : 0.0, leas = INNER ? Rw[-1].le_lat : 0.0, rac = INNER ? Rw->r_area_cell : 0.0;
  for (int r = r0; r <= re; ++r, c += n, ++Rw) {
    const int cn = r < re ? c + n : c;
    const double ncy = __ldg(pcy + cn), nvv = __ldg(pv + cn), ndvy = INNER ? __ldg(a.divy + cn) : 0.0;
    const double nlea = INNER ? Rw[1].le_lat : 0.0, nrac = INNER ? Rw[1].r_area_cell : 0.0;
    double fp4;
    if (direct) fp4 = __ldg(pmy + (r < re ? c + 4 * n : c + 3 * n));          // the last iteration re-reads row re+3 (unused)
    else fp4 = ldy(r + 4);
    const double sp2 = cell_slope<LIM>(fp1, fp2, fp3);
    double fy = 0;
    if (r >= flo && r <= fhi) {
      const bool pos = cy > 0;
      const double f = pos ? f0 : fp1, dfc = pos ? s0 : sp1;
      if (LIM >= 4) fy = vv * (f + dfc * 0.5 * (copysign(1.0, cy) - cy));       // ffsl_mod.F90:605-611
      else if (cy != 0) {
        const double fa = pos ? fm1 : f0, fb = pos ? fp1 : fp2, dfl = pos ? sm1 : s0, dfr = pos ? sp1 : sp2;
        const double s1 = pos ? 1 - cy : 0.0, s2 = pos ? 1.0 : -cy;
        double ml, dm, m6;
        ppm_from_slopes(fa, f, fb, dfl, dfc, dfr, ml, dm, m6);
        fy = qdiv((pos ? vv : -vv) * ppm_integral(0.0, ml, dm, m6, s1, s2), cy);
      }
      if (!INNER) a.oy[c] = fy;
      else if (r == 1 || r == g.nhalf) a.qmfy[c] = fy;              // pole-adjacent half rows: zonal means of the pole rows
    }
    if (INNER && r >= qlo && r <= qhi)                              // ffsl_mod.F90:363-367
      a.oy[c] = f0 - 0.5 * ((fy * lea - fy_prev * leas) * rac - dvy * f0) * a.dt;
    fy_prev = fy;
    fm1 = f0; f0 = fp1; fp1 = fp2; fp2 = fp3; fp3 = fp4;
    sm1 = s0; s0 = sp1; sp1 = sp2;
    cy = ncy; vv = nvv; dvy = ndvy;
    if (INNER) { leas = lea; lea = nlea; rac = nrac; }
  }
}

// ---- x sweep: one row segment per (block row), shared slopes --------------------------------------------------
#ifndef FX_ROWS
#define FX_ROWS 1          // rows per block (blockDim = (W, FX_ROWS)): independent rows, common barriers (measured: 1 best)
#endif
#ifndef FX_MINB
#define FX_MINB 4
#endif
template <int LIM, bool INNER> __device__ __forceinline__ void ffsl_x_body(const Geo& g, const DMesh& M, const FfslArgs& a) {
  __shared__ double xr[FX_ROWS][FM_WMAX + 6];     // x-swept field of the row, cells x0-3 .. x0+W+2
  __shared__ double sxr[FX_ROWS][FM_WMAX + 4];    // its limited slopes, cells x0-2 .. x0+W+1
  __shared__ double fxr[INNER ? FX_ROWS : 1][INNER ? FM_WMAX + 1 : 1];    // INNER: x fluxes of the edges x0-1 .. x0+W-1
  const int n = g.nlon, W = blockDim.x, tx = threadIdx.x, ty = threadIdx.y, x0 = blockIdx.x * W, i = x0 + tx, k = blockIdx.z;
  const int j = g.jb + blockIdx.y * FX_ROWS + ty;
  const bool act = i < n;
  const int ic = act ? i : (i <= n + 2 ? i - n : n - 1);            // idle lanes: wrapped cells n..n+2, then any valid column
  const bool dox = j <= g.je && j > 1 && j < g.nlat && j >= a.jx0 && j <= a.jx1;      // warp-uniform (a warp = part of one row)
  const int jc = j <= g.je ? j : g.je;                                // rows past the band: clamped loads, no stores
  const int c = gidx(g, ic, jc, k);
  const int alc = tx < 3 ? tx - 3 : W + tx - 3;                     // apron cell of threads 0..5
  const bool hasap = tx < 6 && x0 + alc <= n + 2;
  // ---- one memory phase ---------------------------------------------------------------------------------------
  const double qv = __ldg(a.mx + c), cx = __ldg(a.cflx + c), uu = __ldg(a.u + c);
  const double xa = hasap ? __ldg(a.mx + c + (wrapi(x0 + alc, n) - ic)) : 0.0;
  double dvx = 0, cxw = 0, uuw = 0;
  const int dcw = wrapi(x0 - 1, n) - ic;
  double lel = 0, rac = 0;
  if (INNER) {
    dvx = __ldg(a.divx + c); if (tx == 0) { cxw = __ldg(a.cflx + c + dcw); uuw = __ldg(a.u + c + dcw); }
    const RowCoef* __restrict__ R = M.rows + (jc + 1); lel = R->le_lon; rac = R->r_area_cell;
  }
  xr[ty][tx + 3] = qv;
  if (hasap) xr[ty][alc + 3] = xa;
  __syncthreads();
  sxr[ty][tx + 2] = cell_slope<LIM>(xr[ty][tx + 2], qv, xr[ty][tx + 4]);
  if (tx < 4) { const int lc = tx < 2 ? tx - 2 : W + tx - 2; sxr[ty][lc + 2] = cell_slope<LIM>(xr[ty][lc + 2], xr[ty][lc + 3], xr[ty][lc + 4]); }
  __syncthreads();
  const double* __restrict__ mxr = a.mx + (c - ic);
  double fx = 0;
  if (dox) fx = ffsl_flux_x_smem<LIM>(xr[ty], sxr[ty], mxr, tx, ic, n, cx, uu);
  if (!INNER) { if (dox && act) a.ox[c] = fx; return; }
  fxr[ty][tx + 1] = fx;
  __syncthreads();
  if (dox && act) {                                                  // ffsl_mod.F90:358-362
    // the flux through the edge west of the block is needed by thread 0 alone: it forms it AFTER the barrier, so that
    // only its own warp, not the whole block, waits for the extra flux evaluation
    const double fxm = tx == 0 ? ffsl_flux_x_smem<LIM>(xr[ty], sxr[ty], mxr, -1, ic + dcw, n, cxw, uuw) : fxr[ty][tx];
    a.ox[c] = qv - 0.5 * ((fx - fxm) * lel * rac - dvx * qv) * a.dt;
  }
}
// ---- x sweep, warp-private: no block barrier at all ------------------------------------------------------------
// Each WARP owns a segment of one row and stages the 37 values / forms the 35 slopes its 32 edges need in its own
// slice of shared memory (__syncwarp only).  Outer call: 32 fluxes per warp, segments aligned to 32 cells.  Inner
// call: qx of a cell needs the fluxes through both of its edges, so a warp evaluates 32 consecutive edges and
// writes the 31 cells between them (the west neighbour's flux comes from a shuffle): segments of 31 cells,
// one flux in 32 evaluated twice, no barrier and no special thread.
#ifndef XW_WARPS
#define XW_WARPS 4
#endif
template <int LIM, bool INNER> __device__ __forceinline__ void ffsl_xw_body(const Geo& g, const DMesh& M, const FfslArgs& a) {
  __shared__ double xs_[XW_WARPS][40];      // values of cells e0-2 .. e0+34, index = local cell + 2
  __shared__ double ss_[XW_WARPS][36];      // slopes of cells e0-1 .. e0+33, index = local cell + 1
  const int n = g.nlon, lane = threadIdx.x & 31, w = threadIdx.x >> 5, k = blockIdx.z;
  const int j = g.jb + blockIdx.y;
  if (!(j > 1 && j < g.nlat && j >= a.jx0 && j <= a.jx1)) return;             // block-uniform
  constexpr int STR = INNER ? 31 : 32;      // cells written per warp
  const int o0 = (blockIdx.x * XW_WARPS + w) * STR;
  if (o0 >= n) return;                      // warp-uniform
  const int e0 = INNER ? o0 - 1 : o0;       // lane l evaluates the flux through the east edge of cell e0 + l
  double* __restrict__ xs = xs_[w]; double* __restrict__ ss = ss_[w];
  const int rb = gidx(g, 0, j, k);
  const double* __restrict__ mxr = a.mx + rb;
  const int ce = e0 + lane, cw = wrapi(ce, n);
  // ---- one memory phase ---------------------------------------------------------------------------------------
  const double v = __ldg(mxr + cw), cx = __ldg(a.cflx + rb + cw), uu = __ldg(a.u + rb + cw);
  const int alc = lane < 2 ? lane - 2 : 30 + lane;                            // apron cells -2, -1, 32, 33, 34 (lanes 0..4)
  const double xa = lane < 5 ? __ldg(mxr + wrapi(e0 + alc, n)) : 0.0;
  double dvx = 0, lel = 0, rac = 0;
  if (INNER) { dvx = __ldg(a.divx + rb + cw); const RowCoef* __restrict__ R = M.rows + (j + 1); lel = R->le_lon; rac = R->r_area_cell; }
  xs[lane + 2] = v;
  if (lane < 5) xs[alc + 2] = xa;
  __syncwarp();
  ss[lane + 1] = cell_slope<LIM>(xs[lane + 1], v, xs[lane + 3]);
  if (lane < 3) { const int sl = lane == 0 ? -1 : 31 + lane; ss[sl + 1] = cell_slope<LIM>(xs[sl + 1], xs[sl + 2], xs[sl + 3]); }
  __syncwarp();
  const double fx = ffsl_flux_x_smem<LIM>(xs - 1, ss - 1, mxr, lane, cw, n, cx, uu);
  if (!INNER) { if (ce < n) a.ox[rb + ce] = fx; return; }
  const double fxm = __shfl_up_sync(0xffffffffu, fx, 1);
  if (lane >= 1 && ce < n) a.ox[rb + ce] = v - 0.5 * ((fx - fxm) * lel * rac - dvx * v) * a.dt;      // ffsl_mod.F90:358-362
}
template <int LIM> __global__ void __launch_bounds__(32 * XW_WARPS) k_ffsl_xw_in(Geo g, DMesh M, FfslArgs a) { ffsl_xw_body<LIM, true>(g, M, a); }
template <int LIM> __global__ void __launch_bounds__(32 * XW_WARPS) k_ffsl_xw_out(Geo g, DMesh M, FfslArgs a) { ffsl_xw_body<LIM, false>(g, M, a); }
static inline dim3 ffsl_xw_grid(const Geo& g, int nk, bool inner) {
  const int str = inner ? 31 : 32, segs = (g.nlon + str - 1) / str;
  return dim3((segs + XW_WARPS - 1) / XW_WARPS, g.je - g.jb + 1, nk);
}
template <int LIM> __global__ void __launch_bounds__(128, FY_MINB) k_ffsl_y_in(Geo g, DMesh M, FfslArgs a, int rj) { ffsl_y_body<LIM, true>(g, M, a, rj); }
template <int LIM> __global__ void __launch_bounds__(128, FY_MINB) k_ffsl_y_out(Geo g, DMesh M, FfslArgs a, int rj) { ffsl_y_body<LIM, false>(g, M, a, rj); }
template <int LIM> __global__ void __launch_bounds__(FM_WMAX * FX_ROWS, FX_MINB) k_ffsl_x_in(Geo g, DMesh M, FfslArgs a) { ffsl_x_body<LIM, true>(g, M, a); }
template <int LIM> __global__ void __launch_bounds__(FM_WMAX * FX_ROWS, FX_MINB) k_ffsl_x_out(Geo g, DMesh M, FfslArgs a) { ffsl_x_body<LIM, false>(g, M, a); }
static inline dim3 ffsl_x_grid(const Geo& g, int nk, int w) { return dim3((g.nlon + w - 1) / w, (g.je - g.jb + 1 + FX_ROWS - 1) / FX_ROWS, nk); }
static inline dim3 ffsl_y_grid(const Geo& g, int nk, int rj) { return dim3((g.nlon + 127) / 128, (g.je - g.jb + 1 + rj - 1) / rj, nk); }

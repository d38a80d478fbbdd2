// ORACLE — TEST INFRASTRUCTURE ONLY.  CPU restatement of GMCORE's dynamics hot
// path (reference tree /root/reference, commit 73c59ed), used by tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline leg as the checker.  It is
// never linked into or called by the product library.
//
// PARITY UNPINNED: the reference ships no golden vectors for this path and cannot
// be compiled here (no Fortran compiler, MPI or netCDF; five submodules missing),
// so this restatement is anchored on the reference's own invariants instead
//  — see tests/.
//
// This header: halo'd Fortran-bounds arrays, mesh coefficient vectors, and the
// single-process semantics of fill_halo / zonal_sum:
//   src/utils/allocator_mod.F90:168-201   (zero-initialised halo'd allocation)
//   src/utils/parallel_mod.F90:464-565    (fill_halo_3d_r8, incl. pole mirror)
//   src/utils/halo_mod.F90:46-267         (which rows/columns travel)
//   src/utils/parallel_mod.F90:741-764    (zonal_sum)
#pragma once
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include "../include/gmcore_gpu.h"

// Loop-level OpenMP for the TIMED build only (liboracle_fast.so, -DORC_OPENMP_LOOPS): the level loops of the
// restatement are independent, which is how the reference's MPI ranks split the work too.  The parity build
// (liboracle.so) compiles these macros away and stays scalar, left-to-right.
#if defined(_OPENMP) && defined(ORC_OPENMP_LOOPS)
#define ORC_PRAGMA(x) _Pragma(#x)
#define ORC_PAR ORC_PRAGMA(omp parallel for schedule(static))
#define ORC_PAR_FP(v) ORC_PRAGMA(omp parallel for schedule(static) firstprivate(v))
#else
#define ORC_PAR
#define ORC_PAR_FP(v)
#endif

namespace orc {

struct A3 {
  std::vector<double> d;
  int ilb = 1, jlb = 1, klb = 1, ni = 0, nj = 0, nk = 0;
  void alloc(int ilb_, int iub, int jlb_, int jub, int klb_, int kub) {
    ilb = ilb_; jlb = jlb_; klb = klb_;
    ni = iub - ilb + 1; nj = jub - jlb + 1; nk = kub - klb + 1;
    d.assign((size_t)ni * nj * nk, 0.0);
  }
  inline double& operator()(int i, int j, int k = 1) {
    return d[(size_t)(i - ilb) + (size_t)ni * ((size_t)(j - jlb) + (size_t)nj * (size_t)(k - klb))];
  }
  inline double operator()(int i, int j, int k = 1) const {
    return d[(size_t)(i - ilb) + (size_t)ni * ((size_t)(j - jlb) + (size_t)nj * (size_t)(k - klb))];
  }
  int kub() const { return klb + nk - 1; }
  bool allocated() const { return !d.empty(); }
};

struct V1 {
  std::vector<double> d; int lb = 1;
  void set(const double* p, int lb_, int ub_) { lb = lb_; d.assign(p, p + (ub_ - lb_ + 1)); }
  inline double operator()(int j) const { return d[j - lb]; }
};
struct V2 {  // (2, lb:ub)
  std::vector<double> d; int lb = 1;
  void set(const double* p, int lb_, int ub_) { lb = lb_; d.assign(p, p + 2 * (ub_ - lb_ + 1)); }
  inline double operator()(int c, int j) const { return d[(size_t)(j - lb) * 2 + (c - 1)]; }
};

struct Mesh {
  int nlon, nlat, nlev, hx, hy, nhalf;
  // index ranges of the (single, global) block; mesh_mod.F90:148-166,586-609
  int full_lon_ibeg, full_lon_iend, half_lon_ibeg, half_lon_iend;
  int full_lat_ibeg, full_lat_iend, half_lat_ibeg, half_lat_iend;
  int full_lat_ibeg_no_pole, full_lat_iend_no_pole;
  int full_lev_ibeg, full_lev_iend, half_lev_ibeg, half_lev_iend;
  int full_lon_lb, full_lon_ub, full_lat_lb, full_lat_ub, half_lat_lb, half_lat_ub;
  int full_lev_lb, full_lev_ub, half_lev_lb, half_lev_ub;
  V1 full_lat, half_lat, full_cos_lat, half_cos_lat, full_sin_lat, half_sin_lat;
  V1 area_cell, area_lon, area_lon_west, area_lon_east, area_lon_north, area_lon_south;
  V1 area_lat, area_lat_west, area_lat_east, area_lat_north, area_lat_south, area_vtx;
  V2 area_subcell, full_tangent_wgt, half_tangent_wgt;
  V1 de_lon, le_lon, de_lat, le_lat, full_f, half_f, half_lon;
  V1 hyai, hybi, full_lev, half_lev, full_dlev, half_dlev;

  void init(const gmc_config& c, const gmc_mesh_vectors& mv) {
    nlon = c.nlon; nlat = c.nlat; nlev = c.nlev; hx = c.hx; hy = c.hy; nhalf = nlat - 1;
    full_lon_ibeg = half_lon_ibeg = 1; full_lon_iend = half_lon_iend = nlon;
    full_lat_ibeg = 1; full_lat_iend = nlat; half_lat_ibeg = 1; half_lat_iend = nhalf;
    full_lat_ibeg_no_pole = 2; full_lat_iend_no_pole = nlat - 1;
    full_lev_ibeg = 1; full_lev_iend = nlev; half_lev_ibeg = 1; half_lev_iend = nlev + 1;
    full_lon_lb = 1 - hx; full_lon_ub = nlon + hx;
    full_lat_lb = 1 - hy; full_lat_ub = nlat + hy;
    half_lat_lb = 1 - hy; half_lat_ub = nhalf + hy;
    full_lev_lb = -1; full_lev_ub = nlev + 2; half_lev_lb = 1; half_lev_ub = nlev + 1;
    auto F = [&](V1& v, const double* p) { v.set(p, full_lat_lb, full_lat_ub); };
    auto H = [&](V1& v, const double* p) { v.set(p, half_lat_lb, half_lat_ub); };
    F(full_lat, mv.full_lat); H(half_lat, mv.half_lat);
    F(full_cos_lat, mv.full_cos_lat); H(half_cos_lat, mv.half_cos_lat);
    F(full_sin_lat, mv.full_sin_lat); H(half_sin_lat, mv.half_sin_lat);
    F(area_cell, mv.area_cell); F(area_lon, mv.area_lon); F(area_lon_west, mv.area_lon_west);
    F(area_lon_east, mv.area_lon_east); F(area_lon_north, mv.area_lon_north); F(area_lon_south, mv.area_lon_south);
    H(area_lat, mv.area_lat); H(area_lat_west, mv.area_lat_west); H(area_lat_east, mv.area_lat_east);
    H(area_lat_north, mv.area_lat_north); H(area_lat_south, mv.area_lat_south); H(area_vtx, mv.area_vtx);
    area_subcell.set(mv.area_subcell, full_lat_lb, full_lat_ub);
    full_tangent_wgt.set(mv.full_tangent_wgt, full_lat_lb, full_lat_ub);
    half_tangent_wgt.set(mv.half_tangent_wgt, half_lat_lb, half_lat_ub);
    F(de_lon, mv.de_lon); F(le_lon, mv.le_lon); H(de_lat, mv.de_lat); H(le_lat, mv.le_lat);
    F(full_f, mv.full_f); H(half_f, mv.half_f);
    half_lon.set(mv.half_lon, full_lon_lb, full_lon_ub);
    if (mv.hyai) {
      hyai.set(mv.hyai, 1, nlev + 1); hybi.set(mv.hybi, 1, nlev + 1);
      half_lev.set(mv.half_lev, 1, nlev + 1); half_dlev.set(mv.half_dlev, 1, nlev + 1);
      full_lev.set(mv.full_lev, -1, nlev + 2); full_dlev.set(mv.full_dlev, -1, nlev + 2);
    }
  }
  // allocate_array (allocator_mod.F90): zero-initialised, halo'd
  void alloc(A3& a, bool full_lat, bool full_lev = true) const {
    a.alloc(full_lon_lb, full_lon_ub, full_lat ? full_lat_lb : half_lat_lb, full_lat ? full_lat_ub : half_lat_ub,
            full_lev ? full_lev_lb : half_lev_lb, full_lev ? full_lev_ub : half_lev_ub);
  }
  void alloc2(A3& a, bool full_lat) const {
    a.alloc(full_lon_lb, full_lon_ub, full_lat ? full_lat_lb : half_lat_lb, full_lat ? full_lat_ub : half_lat_ub, 1, 1);
  }
};

// fill_halo for one process owning the whole sphere (both poles): W/E exchange is
// a periodic self-copy over rows lat_ibeg..lat_iend; S/N are the pole mirrors.
// full_lon is irrelevant for a single process (same bounds).
inline void fill_halo(const Mesh& m, A3& a, bool full_lat, bool west = true, bool east = true,
                      bool south = true, bool north = true) {
  const int hx = m.hx, hy = m.hy, nlon = m.nlon, nlat = m.nlat;
  const int N = full_lat ? nlat : m.nhalf;  // last interior row
  const int k0 = a.klb, k1 = a.kub();
  // W/E types cover local rows of the *full* index range lat_ibeg..lat_iend
  // (halo_mod.F90:90-96: half_lat_ibeg/iend = full ones unless orient==north).
  const int jb = 1, je = nlat;
  if (west) {
    ORC_PAR for (int k = k0; k <= k1; ++k) for (int j = jb; j <= je; ++j) for (int i = 1 - hx; i <= 0; ++i)
      a(i, j, k) = a(i + nlon, j, k);
  }
  if (east) {
    ORC_PAR for (int k = k0; k <= k1; ++k) for (int j = jb; j <= je; ++j) for (int i = nlon + 1; i <= nlon + hx; ++i)
      a(i, j, k) = a(i - nlon, j, k);
  }
  const int half = nlon / 2;
  std::vector<double> tmp((size_t)a.ni * hy);
  if (south) {
    ORC_PAR_FP(tmp) for (int k = k0; k <= k1; ++k) {
      // Sendrecv: send rows s=1..hy -> halo rows (s-hy), all columns
      for (int s = 1; s <= hy; ++s) {
        int src = full_lat ? 1 + s : s;
        for (int i = a.ilb; i < a.ilb + a.ni; ++i) a(i, s - hy, k) = a(i, src, k);
      }
      for (int s = 1; s <= hy; ++s)
        for (int i = a.ilb; i < a.ilb + a.ni; ++i) tmp[(size_t)(s - 1) * a.ni + (i - a.ilb)] = a(i, s - hy, k);
      for (int s = 1; s <= hy; ++s) {       // reverse rows, rotate interior columns by nlon/2
        const double* t = &tmp[(size_t)(s - 1) * a.ni];
        for (int i = 1; i <= half; ++i) a(i, 1 - s, k) = t[i + half - a.ilb];
        for (int i = half + 1; i <= nlon; ++i) a(i, 1 - s, k) = t[i - half - a.ilb];
      }
    }
  }
  if (north) {
    ORC_PAR_FP(tmp) for (int k = k0; k <= k1; ++k) {
      for (int s = 1; s <= hy; ++s) {
        int src = N - hy - (full_lat ? 1 : 0) + s;
        for (int i = a.ilb; i < a.ilb + a.ni; ++i) a(i, N + s, k) = a(i, src, k);
      }
      for (int s = 1; s <= hy; ++s)
        for (int i = a.ilb; i < a.ilb + a.ni; ++i) tmp[(size_t)(s - 1) * a.ni + (i - a.ilb)] = a(i, N + s, k);
      for (int s = 1; s <= hy; ++s) {
        const double* t = &tmp[(size_t)(hy - s) * a.ni];   // tmp(:, hy-s+1)
        for (int i = 1; i <= half; ++i) a(i, N + s, k) = t[i + half - a.ilb];
        for (int i = half + 1; i <= nlon; ++i) a(i, N + s, k) = t[i - half - a.ilb];
      }
    }
  }
}

// zonal_sum over the full circle, left to right (ENSURE_ORDER semantics).
inline double zonal_sum(const std::vector<double>& work) {
  double s = 0; for (double x : work) s += x; return s;
}

}  // namespace orc

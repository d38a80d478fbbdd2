// Common device-side definitions for the GMCORE dynamics kernels (sm_100a).
//
// Data layout in HBM: every field is a dense FP64 array
// (lev, lat_row, lon) with longitude contiguous and NO longitude halo — the
// periodic wrap is an index map inside the kernels (fill_halo W/E in
// src/utils/parallel_mod.F90:489-499 becomes `wrapi`).  Latitude rows carry HL
// materialised halo rows on each side (pole mirror or neighbour band rows,
// parallel_mod.F90:501-563).  Rows are 128-byte aligned when nlon % 16 == 0.
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>

#define GMC_HL 2   // device latitude halo rows (== reference lat_halo_width)

struct Geo {
  int nlon, nlat, nlev;   // global sizes
  int nhalf;              // nlat - 1
  int jlo;                // global full-lat index of device row 0 (= lat_ibeg - HL)
  int NY;                 // device rows per level (owned + 2*HL)
  int jb, je;             // owned full-lat rows (global, inclusive)
  int hjb, hje;           // owned half-lat rows (global, inclusive)
  int has_south, has_north;
  long long plane;        // NY * nlon
};

// 32-bit element index: every field has fewer than 2^31 elements (checked in gmc_create),
// which keeps the per-load address arithmetic off the 64-bit IMAD path.
__host__ __device__ __forceinline__ int gidx(const Geo& g, int i, int j, int k) {
  return (k * g.NY + (j - g.jlo)) * g.nlon + i;
}
__device__ __forceinline__ int wrapi(int i, int nlon) {
  return i < 0 ? i + nlon : (i >= nlon ? i - nlon : i);
}

// Per-latitude coefficient vectors (mesh_type, src/dynamics/mesh_mod.F90:84-110),
// stored with the reference's lower bound 1-hy = -1, so entry j lives at [j+1].
// All per-latitude constants of one row packed together (array of structs over j, same lower
// bound as the vectors: entry j at rows[j+1]).  A kernel forms ONE address (&rows[j+1]) and reaches
// every coefficient of rows j-1, j, j+1 with immediate offsets, instead of one 64-bit address
// computation per coefficient vector (the integer pipe was the busiest one in the stencil kernels).
struct RowCoef {
  // full-lat row j
  double area_cell, r_area_cell, area_lon_west, area_lon_east, area_lon_north, area_lon_south, r_area_lon;
  double de_lon, r_de_lon, le_lon, full_cos_lat, full_sin_lat, ftw1, ftw2, sub1, sub2;
  // half-lat row j
  double area_lat_west, area_lat_east, area_lat_north, area_lat_south, r_area_lat, r_area_vtx;
  double de_lat, r_de_lat, le_lat, half_f, htw1, htw2, half_lat, area_lon, area_lat, pad_;
};
static_assert(sizeof(RowCoef) == 256, "RowCoef is one 256-byte line pair");

struct DMesh {
  const RowCoef* rows;
  const double *full_lat, *half_lat, *full_cos_lat, *half_cos_lat, *full_sin_lat, *half_sin_lat;
  const double *area_cell, *area_lon, *area_lon_west, *area_lon_east, *area_lon_north, *area_lon_south;
  const double *area_lat, *area_lat_west, *area_lat_east, *area_lat_north, *area_lat_south, *area_vtx;
  const double *area_subcell;   // (2, j)
  const double *de_lon, *le_lon, *de_lat, *le_lat, *full_f, *half_f;
  const double *full_tangent_wgt, *half_tangent_wgt;   // (2, j)
  const double *half_lon;       // lower bound 1-hx
  const double *hyai, *hybi, *full_lev, *half_lev, *full_dlev, *half_dlev;  // 1-based / -1-based as in the header
  // reciprocals of row constants (formed once on the host, 0 where the constant is 0): a divide by
  // a per-latitude coefficient becomes a multiply (<= 1 ulp from the reference's divide, far inside
  // the 1e-12 parity bar) -- FP64 divides are ~30 issue slots each and dominated several kernels.
  const double *r_area_cell, *r_area_vtx, *r_de_lon, *r_de_lat, *r_area_lon, *r_area_lat;
  const double *r_full_dlev;   // 1 / full_dlev, same bounds (-1 : nlev+2)
  // Pure-pressure levels: the leading nconst interfaces have hybi == 0, so ph_lev there is the same
  // number in every column and its pow / log are formed once per level ON THE DEVICE (k_lev_tab, the
  // same expressions as k_ph / k_gz_col, hence the same bits) instead of once per grid point:
  // lev_tab[k] = ph_lev(k+1)^kappa, lev_tab[nlev+1+k] = (ph(k+1)/p0)^kappa, lev_tab[2(nlev+1)+k] = log(ph_lev(k+2)/ph_lev(k+1)).
  const double *lev_tab; int nconst;
  int hx;
};
#define MJ(v, j) (v[(j) + 1])                  // vector entry at latitude index j
#define MJ2(v, c, j) (v[((j) + 1) * 2 + ((c) - 1)])   // (c, j) of a (2,:) array, c = 1|2

struct DFilter {
  const int* ngrid_lon; const double* wgt_lon; int ld_lon;
  const int* ngrid_lat; const double* wgt_lat; int ld_lat;
  int max_hn;
};

// 2-D thread tiles for the horizontal stencil kernels: TBX threads along longitude x (128/TBX) latitude
// rows per block, so that the j-1 / j+1 neighbours of a warp's row are fetched by warps of the
// same block (L1 hits instead of L2 round trips).
#ifndef TBX
#define TBX 32
#endif
static inline dim3 tile_block() { return dim3(TBX, 128 / TBX, 1); }
static inline dim3 tile_grid(const Geo& g, int nk) {
  const dim3 b = tile_block();
  return dim3((g.nlon + b.x - 1) / b.x, (g.je - g.jb + 1 + b.y - 1) / b.y, nk);
}

#define CUDA_CHECK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  snprintf(h->err, sizeof(h->err), "%s:%d %s: %s", __FILE__, __LINE__, #x, cudaGetErrorString(e_)); return 1; } } while (0)

__device__ __forceinline__ double sign1(double x) { return copysign(1.0, x); }

// Block-wide sum in a fixed (deterministic) order: warp shuffles then warp 0.
__device__ __forceinline__ double block_sum(double v, double* sm /* >= 32 doubles */) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  __syncthreads();
  if (lane == 0) sm[w] = v;
  __syncthreads();
  double r = 0;
  if (w == 0) {
    r = lane < nw ? sm[lane] : 0.0;
    for (int o = 16; o > 0; o >>= 1) r += __shfl_down_sync(0xffffffffu, r, o);
    if (lane == 0) sm[0] = r;
  }
  __syncthreads();
  r = sm[0];
  return r;
}

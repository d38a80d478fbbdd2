// Latitude-band halo exchange (fill_halo S/N, src/utils/parallel_mod.F90:501-527, for the
// 1-D latitude decomposition of process_mod.F90:122-125).  All arrays of one exchange point
// are packed into ONE message per neighbour, sent with ncclSend/ncclRecv on the
// compute stream (one process per GPU) or copied device-to-device (single-process band group).
#pragma once
#include <dlfcn.h>
#include "handle.cuh"

#define GMC_HALO_MAX 24     // arrays per exchange point; gmc_create rejects tracer counts that would not fit
struct HaloSet { double* p[GMC_HALO_MAX]; int nk[GMC_HALO_MAX]; int n; int ktot; int overflow; };
static inline HaloSet halo_set() { HaloSet s; s.n = 0; s.ktot = 0; s.overflow = 0; return s; }
static inline void halo_add(HaloSet& s, double* p, int nk) {
  if (s.n >= GMC_HALO_MAX) { s.overflow = 1; return; }      // reported by halo_exchange, never written past the arrays
  s.p[s.n] = p; s.nk[s.n] = nk; s.n++; s.ktot += nk;
}

// grid (ceil(nlon/128), 2*H, ktot).  Buffer layout [kk][r][i], kk running over all levels of all arrays.
__global__ void __launch_bounds__(128) k_halo(Geo g, HaloSet s, double* buf_s, double* buf_n, int H, int unpack) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.nlon) return;
  const int side = blockIdx.y / H, r = blockIdx.y % H;
  if ((side == 0 && g.has_south) || (side == 1 && g.has_north)) return;
  int kk = blockIdx.z, a = 0;
  while (kk >= s.nk[a]) { kk -= s.nk[a]; ++a; }
  double* buf = side == 0 ? buf_s : buf_n;
  const size_t b = ((size_t)blockIdx.z * H + r) * g.nlon + i;
  int j;
  if (!unpack) j = side == 0 ? g.jb + r : g.je - H + 1 + r;
  else j = side == 0 ? g.jb - H + r : g.je + 1 + r;
  double* f = s.p[a] + gidx(g, i, j, kk);
  if (unpack) *f = buf[b]; else buf[b] = *f;
}

// ---- NCCL through dlsym: the symbols come from the libnccl.so.2 the host process has
// already loaded (torch bundles it), so the library has no link-time NCCL dependency.
typedef struct { char internal[128]; } gmc_ncclUniqueId;
struct NcclApi {
  int (*GetUniqueId)(gmc_ncclUniqueId*) = nullptr;
  int (*CommInitRank)(void**, int, gmc_ncclUniqueId, int) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  int (*Send)(const void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*Recv)(void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  bool ok = false;
};
static NcclApi& nccl_api() {
  static NcclApi api;
  if (api.ok) return api;
  void* lib = RTLD_DEFAULT;
  if (!dlsym(lib, "ncclSend")) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!lib) return api;
#define SYM(f) *(void**)(&api.f) = dlsym(lib, "nccl" #f)
  SYM(GetUniqueId); SYM(CommInitRank); SYM(CommDestroy); SYM(Send); SYM(Recv); SYM(AllReduce); SYM(GroupStart); SYM(GroupEnd); SYM(GetErrorString);
#undef SYM
  api.ok = api.GetUniqueId && api.CommInitRank && api.Send && api.Recv && api.AllReduce && api.GroupStart && api.GroupEnd;
  return api;
}
enum { GMC_NCCL_DOUBLE = 8, GMC_NCCL_SUM = 0 };   // ncclFloat64, ncclSum (nccl.h)

// Exchange buffers (grow-only): room for ktot levels x H rows per neighbour and direction.
static int halo_ensure(gmc_handle* h, int ktot) {
  if (h->cfg.nranks <= 1) return 0;
  const size_t need = (size_t)ktot * h->H * h->g.nlon;
  if (need <= h->xbuf_elems) return 0;
  int rc = dev_alloc(h, &h->send_s, need);
  if (!rc) rc = dev_alloc(h, &h->send_n, need);
  if (!rc) rc = dev_alloc(h, &h->recv_s, need);
  if (!rc) rc = dev_alloc(h, &h->recv_n, need);
  if (!rc) h->xbuf_elems = need;
  return rc;
}
static int halo_alloc(gmc_handle* h) { return halo_ensure(h, 3 * h->g.nlev + 3); }

template <class F> static void for_bands(gmc_handle* h, F f) {
  if (h->group) { for (gmc_handle* b : *h->group) f(b); } else f(h);
}

// One exchange point.  make_set(band) lists the arrays of that band to exchange.
template <class MakeSet> static int halo_exchange(gmc_handle* h, MakeSet make_set) {
  if (h->cfg.nranks <= 1) return 0;
  int rc = 0;
  for_bands(h, [&](gmc_handle* b) {
    HaloSet s = make_set(b);
    if (s.overflow || (size_t)s.ktot * b->H * b->g.nlon > b->xbuf_elems) { rc = 1; return; }
    LAUNCH(b, k_halo, dim3((b->g.nlon + 127) / 128, 2 * b->H, s.ktot), 128, 0, b->g, s, b->send_s, b->send_n, b->H, 0);
  });
  if (rc) { snprintf(h->err, sizeof(h->err), "halo exchange: more than %d arrays or a payload beyond the exchange buffer", GMC_HALO_MAX); return 1; }
  if (h->group) {
    // same device, same stream: plain device-to-device copies between neighbouring bands
    for (gmc_handle* b : *h->group) {
      HaloSet s = make_set(b);
      const size_t bytes = (size_t)s.ktot * b->H * b->g.nlon * sizeof(double);
      const int r = b->cfg.rank;
      if (r > 0) cudaMemcpyAsync((*h->group)[r - 1]->recv_n, b->send_s, bytes, cudaMemcpyDeviceToDevice, b->stream);
      if (r + 1 < b->cfg.nranks) cudaMemcpyAsync((*h->group)[r + 1]->recv_s, b->send_n, bytes, cudaMemcpyDeviceToDevice, b->stream);
    }
  } else {
    NcclApi& n = nccl_api();
    if (!h->nccl_comm || !n.ok) { snprintf(h->err, sizeof(h->err), "band mode needs gmc_nccl_init or gmc_group_link first"); return 1; }
    HaloSet s = make_set(h);
    const size_t cnt = (size_t)s.ktot * h->H * h->g.nlon;
    const int r = h->cfg.rank;
    n.GroupStart();
    if (r > 0) { rc |= n.Send(h->send_s, cnt, GMC_NCCL_DOUBLE, r - 1, h->nccl_comm, h->stream); rc |= n.Recv(h->recv_s, cnt, GMC_NCCL_DOUBLE, r - 1, h->nccl_comm, h->stream); }
    if (r + 1 < h->cfg.nranks) { rc |= n.Send(h->send_n, cnt, GMC_NCCL_DOUBLE, r + 1, h->nccl_comm, h->stream); rc |= n.Recv(h->recv_n, cnt, GMC_NCCL_DOUBLE, r + 1, h->nccl_comm, h->stream); }
    rc |= n.GroupEnd();
    if (rc) { snprintf(h->err, sizeof(h->err), "NCCL halo exchange failed (%d)", rc); return 1; }
  }
  for_bands(h, [&](gmc_handle* b) {
    HaloSet s = make_set(b);
    LAUNCH(b, k_halo, dim3((b->g.nlon + 127) / 128, 2 * b->H, s.ktot), 128, 0, b->g, s, b->recv_s, b->recv_n, b->H, 1);
  });
  return 0;
}

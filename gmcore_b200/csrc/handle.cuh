// Host-side handle of the GPU dynamics core: device memory pool, per-state field
// sets (mirroring state_type, src/dynamics/state_mod.F90:17-101), launch helpers.
#pragma once
#include <cstring>
#include <map>
#include <string>
#include <vector>
#include "../../include/gmcore_gpu.h"
#include "common.cuh"

struct DState {
  // prognostic + filtered copies
  double *u = 0, *v = 0, *gz = 0, *u_f = 0, *v_f = 0, *gz_f = 0, *pt = 0, *phs = 0, *phs_f = 0;
  // derived, kept per time level because stale copies are consumed
  double *m = 0, *mfx_lon = 0, *mfy_lat = 0, *u_lat = 0, *v_lon = 0, *ke = 0, *pv = 0, *pv_lon = 0, *pv_lat = 0, *div = 0;
  // hydrostatic core
  double *ph_lev = 0, *ph_exn_lev = 0, *t = 0, *gz_lev = 0, *we_lev = 0, *m_lev = 0;
  std::map<std::string, std::pair<double*, int>> names;   // name -> (ptr, kind) kind: 0 full-lat 3D, 1 half-lat 3D, 2 full-lat half-lev, 3 2-D
};

// Every kernel goes through LAUNCH: it counts launches and, in profiling mode
// (gmc_profile), brackets the launch with CUDA events on the launching stream.
struct ProfEntry { const char* name; cudaEvent_t e0, e1; };
struct Profiler {
  bool on = false;
  std::vector<ProfEntry> entries;
  std::vector<cudaEvent_t> pool; size_t used = 0;
  cudaEvent_t get() {
    if (used == pool.size()) { cudaEvent_t e; cudaEventCreate(&e); pool.push_back(e); }
    return pool[used++];
  }
};
#define LAUNCH(h, kern, grid, block, smem, ...) do { \
  ProfEntry pe_{#kern, 0, 0}; \
  if ((h)->prof.on) { pe_.e0 = (h)->prof.get(); cudaEventRecord(pe_.e0, (h)->stream); } \
  kern<<<grid, block, smem, (h)->stream>>>(__VA_ARGS__); (h)->launches++; \
  if ((h)->prof.on) { pe_.e1 = (h)->prof.get(); cudaEventRecord(pe_.e1, (h)->stream); (h)->prof.entries.push_back(pe_); } \
} while (0)

// kernels templated on MODE = limiter_type + 4 * ffsl_flux_type (ffsl_mod.F90:116-140)
#define LAUNCH_LIM(h, kern, ...) do { switch ((h)->cfg.limiter_type + 4 * (h)->cfg.ffsl_flux_type) { \
  case 0: LAUNCH(h, kern<0>, __VA_ARGS__); break; case 2: LAUNCH(h, kern<2>, __VA_ARGS__); break; \
  case 4: LAUNCH(h, kern<4>, __VA_ARGS__); break; case 5: LAUNCH(h, kern<5>, __VA_ARGS__); break; \
  case 6: LAUNCH(h, kern<6>, __VA_ARGS__); break; \
  default: LAUNCH(h, kern<1>, __VA_ARGS__); } } while (0)

// the same for kernels with a second (bool) template parameter; profiled under the name kern_<flag>
#define LAUNCH2_(h, kern, M_, flag, ...) do { if (flag) LAUNCH(h, kern##_in<M_>, __VA_ARGS__); else LAUNCH(h, kern##_out<M_>, __VA_ARGS__); } while (0)
#define LAUNCH_LIM2(h, kern, flag, ...) do { switch ((h)->cfg.limiter_type + 4 * (h)->cfg.ffsl_flux_type) { \
  case 0: LAUNCH2_(h, kern, 0, flag, __VA_ARGS__); break; case 2: LAUNCH2_(h, kern, 2, flag, __VA_ARGS__); break; \
  case 4: LAUNCH2_(h, kern, 4, flag, __VA_ARGS__); break; case 5: LAUNCH2_(h, kern, 5, flag, __VA_ARGS__); break; \
  case 6: LAUNCH2_(h, kern, 6, flag, __VA_ARGS__); break; \
  default: LAUNCH2_(h, kern, 1, flag, __VA_ARGS__); } } while (0)

// adv_batch_type (adv_batch_mod.F90:19-69) of a non-dynamic 'cell' batch
struct TrBatch {
  int ntr = 0; double dt = 0; int nstep = 1;
  double* q[2] = {0, 0}; int q_old = 0, q_new = 1;        // q(:,:,:,l,old|new), [l][lev][lat][lon]
  double alert_last = 0, alert_period = 0; bool alert_ring = false;   // time_mod.F90:253-291 alert of this batch
  bool advected = false;                                    // set by adv_run for the exchange / swap that follows
  int wind_of = 0;                                          // batch owning the accumulators below
  int uv_step = 0, mf_step = 0, we_step = 0;               // sample counters (owner only)
  double *bu = 0, *bv = 0, *bmfx = 0, *bmfy = 0, *bwe = 0, *bm_lev = 0;
  double *bcflx = 0, *bcfly = 0, *bcflz = 0, *bdivx = 0, *bdivy = 0;
};

struct gmc_handle {
  gmc_config cfg;
  Geo g;
  DMesh M;
  DFilter filt[3];
  cudaStream_t stream = 0;
  char err[512] = {0};
  long long launches = 0;
  Profiler prof;
  cudaEvent_t tev[2] = {0, 0};
  std::vector<void*> allocs;
  DState st[4];
  double *du = 0, *dv = 0;            // forward-pass tendencies (tend_type%du/dv)
  double *gzs = 0;                    // static%gzs (2-D)
  double *pole_dmf = 0;               // [2][nlev]
  double *diag_partial = 0, *diag_out = 0;
  double *diag_steps = 0; int diag_steps_cap = 0;    // per-step sums of gmc_run_steps (grow-only)
  double *stage_img = 0; size_t stage_bytes = 0;
  // hydrostatic work arrays (allocated only when baroclinic)
  double *dphs = 0; int* dflags = 0;   // dflags: device word of the reference's log_error guards
  double *cflx = 0, *cfly = 0, *cflz = 0, *divx = 0, *divy = 0;   // adv_batch_mass
  double *qx = 0, *qy = 0, *ptf_lon = 0, *ptf_lat = 0, *qmfx = 0, *qmfy = 0, *dpt = 0;
  double *c_lon = 0, *c_lat = 0;      // div_damp coefficients (j,k)
  double *div2 = 0;                   // Laplacian of div (div_damp_order = 4)
  int diag_queued = 0;                // sets waiting in the diagnostics ring (gmc_diagnose_enqueue)
  // shallow-water time step as a CUDA graph, one per (old, new) pairing: the ~20 kernels of a 0.5 degree step last
  // 3-15 us each, shorter than a launch, so the step is replayed with one cudaGraphLaunch (GMCORE_NO_GRAPH=1: off)
  cudaGraphExec_t swm_graph[2] = {nullptr, nullptr}; long long swm_graph_launches[2] = {0, 0};
  double *lon_damp_wgt = 0;
  double *pole_work = 0;
  double *smag_t = 0, *smag_s = 0, *smag_du = 0, *smag_dv = 0;   // smag_damp work arrays (use_smag_damp)
  // tracer batches (block%adv_batches(:), adv_mod.F90:314-418): batch 0 is cfg.ntracers / cfg.dt_adv, more come from
  // gmc_add_tracer_batch.  Batches with the same dt see the same wind samples, so they share one accumulator set
  // (wind_of = the first batch with that dt); old_m is the same array for every batch (copy_old_m runs for all of
  // them at every adv_run, adv_mod.F90:309).
  std::vector<TrBatch> batches;
  double *old_m = 0; int nstep = 1;   // nstep: divisor of the accumulation launch in flight (accum_pair)
  // latitude-band mode (cfg.nranks > 1): packed halo exchange with the two neighbours
  int H = 0;                          // halo rows exchanged per side
  double *send_s = 0, *send_n = 0, *recv_s = 0, *recv_n = 0; size_t xbuf_elems = 0;
  std::vector<gmc_handle*>* group = nullptr;   // single-process band group (all bands on this device/stream)
  bool owns_stream = true;
  void* nccl_comm = nullptr;          // one-process-per-GPU mode
  // overlap of the per-stage exchange with interior work (one-process-per-GPU mode): edge rows are finished
  // first, ev_edge releases pack + ncclSend/Recv + unpack on comm_stream while the interior rows are filtered
  cudaStream_t comm_stream = 0; cudaEvent_t ev_edge = 0, ev_xdone = 0;
  bool mesh_set = false, filters_set[3] = {false, false, false};
  int max_hn = 0;
  bool keep_dpt = false;              // GMCORE_DEBUG_FIELDS=1: also store the dpt tendency (parity tests read it back)
};


template <typename T>
static int dev_alloc(gmc_handle* h, T** p, size_t n) {
  void* q = nullptr;
  cudaError_t e = cudaMalloc(&q, n * sizeof(T));
  if (e != cudaSuccess) { snprintf(h->err, sizeof(h->err), "cudaMalloc(%zu B): %s", n * sizeof(T), cudaGetErrorString(e)); return 1; }
  e = cudaMemset(q, 0, n * sizeof(T));
  if (e != cudaSuccess) { snprintf(h->err, sizeof(h->err), "cudaMemset: %s", cudaGetErrorString(e)); return 1; }
  h->allocs.push_back(q);
  *p = (T*)q;
  return 0;
}
template <typename T>
static int dev_upload(gmc_handle* h, const T** dst, const T* src, size_t n) {
  T* p = nullptr;
  if (dev_alloc(h, &p, n)) return 1;
  cudaError_t e = cudaMemcpy(p, src, n * sizeof(T), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) { snprintf(h->err, sizeof(h->err), "cudaMemcpy H2D: %s", cudaGetErrorString(e)); return 1; }
  *dst = p;
  return 0;
}

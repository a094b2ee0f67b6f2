// Shared declarations for libmxrcnn_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "mxrcnn_b200.h"

namespace mxr {

typedef unsigned long long u64;
typedef unsigned int u32;

void set_error(const char* fmt, ...);

#define MXR_CUDA(expr)                                                                                       \
  do {                                                                                                       \
    cudaError_t e_ = (expr);                                                                                 \
    if (e_ != cudaSuccess) {                                                                                 \
      mxr::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__);            \
      return MXR_ERR_CUDA;                                                                                   \
    }                                                                                                        \
  } while (0)

#define MXR_REQUIRE(cond, code, ...)  \
  do {                                \
    if (!(cond)) {                    \
      mxr::set_error(__VA_ARGS__);    \
      return (code);                  \
    }                                 \
  } while (0)

#define MXR_TRY(expr)          \
  do {                         \
    int s_ = (expr);           \
    if (s_ != MXR_OK) return s_; \
  } while (0)

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
static inline int round_up_int(int v, int a) { return (v + a - 1) / a * a; }
static inline int pow2_ceil(int v) {
  int p = 1;
  while (p < v) p <<= 1;
  return p;
}

// Per-problem (image / class) control block, zeroed at the start of every call.
struct Ctrl {
  int n_valid;   // decode: boxes that pass the min-size filter
  int nan_flag;  // decode / prepare: NaN seen in the inputs
  int n;         // boxes entering NMS (after top-k)
  int nkept;     // scan: boxes kept so far
  int done;      // scan: problem finished (post-NMS quota reached or list exhausted)
  int pad[3];
};

// Kernel stages, for the optional per-kernel CUDA-event profile (mxr_profile_*).
enum Stage {
  ST_MEMSET = 0,
  ST_DECODE = 1,
  ST_TOPK = 2,
  ST_NMS_STRIPE = 3,
  ST_NMS_SCAN = 4,
  ST_NMS_PREP = 5,
  ST_BOXES = 6,
  ST_OVERLAPS = 7,
  ST_ANCHOR_TARGET = 8,
  ST_PROPOSAL_TARGET = 9,
  ST_NMS_COOP = 10,
  ST_CLUSTER = 11,
  ST_NMS_CLUSTER = 12,
  ST_BOX_HEAD = 13,
  ST_FUSED = 14,
  ST_NMS_FUSED = 15,
  ST_ANCHOR_PHASE2 = 16,
  ST_COUNT = 17
};

struct ProfRec {
  cudaEvent_t a, b;
  int stage;
};

struct Handle {
  int device;
  cudaStream_t stream;
  char* d_ws;
  size_t d_bytes;
  char* h_pin;
  size_t h_bytes;
  cudaEvent_t ev;
  // profiling (off by default: no events are recorded in the hot path)
  int prof_on;
  ProfRec* prof;
  int prof_cap, prof_used;
  int last_launches;  // kernels launched by the most recent call
  // Both point into the process-wide table of this handle's DEVICE (api.cu): kernel attributes belong to
  // (kernel, device), not to a handle, so two handles must not cache them separately.
  int* smem_attr;     // [12] largest dynamic-smem opt-in already set per kernel (1 prep f32, 2 prep f64, 3 coop,
                      // 4/5 first-generation fused proposal, 6/7 cluster NMS, 8/9 fused proposal, 10/11 fused NMS); raise it through
                      // raise_dyn_smem() only
  int* coop_grid;     // [4] [0] co-resident grid of the cooperative NMS kernel (0: not queried)
  int* cluster_cap;   // [4][8] fused proposal kernels (first generation 4 / 12 keys per thread, second generation 4 / 12)
                      // x cluster size (16, 14, ... 2): clusters resident at once; 0 not queried, -1 none
  void* at_state;     // anchor-target state carried from phase 1 to phase 2 (malloc'd, targets.cu)
  void* pt_state;     // proposal-target state carried from phase 1 to phase 2 (malloc'd, targets.cu)
  int no_cluster;     // MXRCNN_B200_NO_CLUSTER=1: use the multi-kernel proposal path instead of the fused one
  int cluster_override; // MXRCNN_B200_CLUSTER=2|4|8|16 pins the cluster size of the fused proposal kernel (tests cover
                        // them all); MXRCNN_B200_CLUSTER8=1 is the older spelling of 8.  0: chosen per call
  int no_pdl;           // MXRCNN_B200_NO_PDL=1: launch the fused kernel without programmatic dependent launch
  int use_v1;           // MXRCNN_B200_V1=1: first-generation fused kernel (proposal_cluster.cuh) instead of proposal_fused.cuh
  int count_pairs;      // mxr_proposal_count_pairs(h, 1): the next fused calls count the IoU pairs they evaluate (un-timed runs)
  unsigned int* pair_ptr;  // device: per-image pair counters of the last such call
  int pair_images;
  void* ctrl_ptr;       // device: control blocks of the last proposal call (mxr_proposal_counts)
  int ctrl_images;
  // stream that last enqueued work on the handle's workspace / pinned block (ws_order)
  cudaStream_t ws_last;
  int ws_used, ws_last_private;
  cudaEvent_t ev2;
  unsigned long long* dbg_ptr;  // device: phase time stamps of the last fused proposal call ([dbg_images][16])
  int dbg_images;
  cudaStream_t stream2;  // second private stream: slot 1 of the pipelined host path
  void* prop_state;      // per-slot state of mxr_proposal_submit_host / _wait_host (malloc'd, proposal.cu)
};

// Bracket one kernel launch.  No-ops unless mxr_profile_enable(h, 1) was called.
static inline void prof_begin(Handle* h, cudaStream_t s, int stage) {
  h->last_launches++;
  if (!h->prof_on || h->prof_used >= h->prof_cap) return;
  ProfRec& r = h->prof[h->prof_used];
  r.stage = stage;
  cudaEventRecord(r.a, s);
}
static inline void prof_end(Handle* h, cudaStream_t s) {
  if (!h->prof_on || h->prof_used >= h->prof_cap) return;
  cudaEventRecord(h->prof[h->prof_used].b, s);
  h->prof_used++;
}

// Makes the calling thread's current device the handle's for the duration of an entry point and puts the caller's
// back afterwards (a framework that drives several GPUs from one thread must not find its device changed).
struct DeviceGuard {
  int prev;
  bool changed, good;
  explicit DeviceGuard(int dev) : prev(-1), changed(false), good(true) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != dev) {
      good = cudaSetDevice(dev) == cudaSuccess;
      changed = good && prev >= 0;
    }
  }
  ~DeviceGuard() {
    if (changed) cudaSetDevice(prev);
  }
};
#define MXR_DEVICE(h_)                                                                          \
  mxr::DeviceGuard dev_guard_((h_)->device);                                                    \
  if (!dev_guard_.good) {                                                                       \
    mxr::set_error("cudaSetDevice(%d) failed: %s", (h_)->device, cudaGetErrorString(cudaGetLastError())); \
    return MXR_ERR_CUDA;                                                                        \
  }

// The handle's workspace and pinned block are shared by every entry point.  Work enqueued on the handle's own streams
// orders itself; work on a caller's stream does not: before stream `s` touches the workspace, make it wait for what the
// previous user stream still has queued (an event recorded there now -- conservative, no cost while the stream stays
// the same).  Call it in every entry point once the stream is known, before anything is enqueued.
int ws_order(Handle* h, cudaStream_t s);

// Grow-only workspaces.  Growth synchronises the device (never in steady state).
int ensure_device_ws(Handle* h, size_t bytes);
int ensure_pinned_ws(Handle* h, size_t bytes);
// Makes sure kernel `func` may be launched with `bytes` of dynamic shared memory on the handle's device.  The
// opt-in is cached per (device, slot) for the whole process and only ever raised.  *raised (optional) tells the
// caller that the limit moved, i.e. that occupancy figures derived from it must be queried again.
int raise_dyn_smem(Handle* h, int slot, const void* func, size_t bytes, int* raised);

// Bump allocator over the handle's device workspace.
struct Carver {
  char* base;
  size_t off;
  explicit Carver(char* b) : base(b), off(0) {}
  template <typename T>
  T* take(size_t count) {
    off = align_up(off, 256);
    T* p = reinterpret_cast<T*>(base ? base + off : nullptr);
    off += count * sizeof(T);
    return p;
  }
};

struct AnchorTable {
  double a[64][4];
};

}  // namespace mxr

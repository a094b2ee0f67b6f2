// nms(dets, thresh) (helper/processing/nms.py:4-38) as ONE launch: a thread-block cluster per problem
// (image, or class in the per-class loop of rcnn/tester.py:50-56).  The leader CTA orders the scores
// (nms.py:19; ties: higher index first), all CTAs gather boxes and areas (nms.py:18) into processing order,
// then the cluster walks the list greedily in 256-box stripes (cluster_nms.cuh).
#pragma once
#include "cluster_nms.cuh"
#include "nms_coop.cuh"
#include "sort.cuh"

namespace mxr {

template <typename T>
struct NmsClusterArgs {
  const T* dets;        // (n_total, 5)
  const int* offsets;   // device, (P+1)
  NmsArrays<T> a;       // box / area / orig / kept_slot scratch (cap per problem); M, R unused
  NmsOutputs o;
  Ctrl* ctrl;
  T thr;
  int max_keep;
  int sort_cap;         // sort slots in the leader's shared memory (power of two >= the largest problem, >= 1024)
};

template <typename T>
struct NmsKey;
template <>
struct NmsKey<float> {
  static const bool kPayload = false;  // 32-bit score key and row index share one 64-bit word
  static __device__ __forceinline__ u64 make(float s, u32 i) { return ((u64)score_key32(s) << 32) | i; }
};
template <>
struct NmsKey<double> {
  static const bool kPayload = true;   // 64-bit score key, row index (+1) rides as payload
  static __device__ __forceinline__ u64 make(double s, u32) { return score_key64(s); }
};

// grid P * csize CTAs, cluster (csize,1,1), block 1024.
template <typename T>
__global__ void __launch_bounds__(1024, 1) nms_cluster_kernel(NmsClusterArgs<T> g) {
  namespace cgx = cooperative_groups;
  typedef typename Ar<T>::Box Box;
  extern __shared__ __align__(16) unsigned char dyn_smem[];
  __shared__ __align__(8) int s_status[2];
  __shared__ int s_local[4];
  cgx::cluster_group cluster = cgx::this_cluster();
  const int rank = (int)cluster.block_rank();
  const int csize = (int)cluster.num_blocks();
  const int b = blockIdx.x / csize;
  const int tid = threadIdx.x, warp = tid >> 5;
  const int row0 = g.offsets[b];
  const int n = g.offsets[b + 1] - row0;
  const T* d = g.dets + (size_t)row0 * 5;
  u64* sortbuf = reinterpret_cast<u64*>(dyn_smem);                                             // leader: [sort_cap]
  u32* sortpl = reinterpret_cast<u32*>(sortbuf + g.sort_cap);                                  // leader: [sort_cap] (f64 scores only)
  u64* sM = reinterpret_cast<u64*>(sortpl + (NmsKey<T>::kPayload ? g.sort_cap : 0));           // leader: [256][5]
  u64* s_removed = sM + kCW * kCPitch;
  Box* t_box = reinterpret_cast<Box*>(s_removed + 8) + warp * 16;
  T* t_area = reinterpret_cast<T*>(reinterpret_cast<Box*>(s_removed + 8) + 32 * 16) + warp * 16;
  u64* L_sort = cluster.map_shared_rank(sortbuf, 0);
  u32* L_sortpl = cluster.map_shared_rank(sortpl, 0);
  if (rank == 0) {
    if (tid < kCSW) s_removed[tid] = 0ull;
    if (tid < 2) s_status[tid] = 0;
  }
  if (tid < 4) s_local[tid] = 0;
  // ---- order = scores.argsort()[::-1] (leader) ----
  bool nan = false;
  if (rank == 0 && n > 0) {
    const int n_pow2 = max(2, (int)(1u << (32 - __clz(max(n, 2) - 1))));
    const int fill = (!NmsKey<T>::kPayload && n_pow2 <= 1024) ? 1024 : n_pow2;
    for (int i = tid; i < fill; i += blockDim.x) {
      if (i < n) {
        const T s = d[(size_t)i * 5 + 4];
        nan = nan || (s != s);
        sortbuf[i] = NmsKey<T>::make(s, (u32)i);
        if (NmsKey<T>::kPayload) sortpl[i] = (u32)i + 1u;  // +1 keeps every real pair above the (0,0) padding
      } else {
        sortbuf[i] = 0ull;
        if (NmsKey<T>::kPayload) sortpl[i] = 0u;
      }
    }
    __syncthreads();
    if (NmsKey<T>::kPayload) bitonic_sort_desc(sortbuf, sortpl, n_pow2);
    else if (n_pow2 <= 1024) bitonic_sort_1024_x4(sortbuf, n_pow2);
    else bitonic_sort_desc_u64(sortbuf, n_pow2);
  }
  cluster.sync();
  // ---- gather boxes and areas into processing order (all CTAs) ----
  Box* box = const_cast<Box*>(g.a.box) + (size_t)b * g.a.cap;
  T* area = const_cast<T*>(g.a.area) + (size_t)b * g.a.cap;
  int* orig = const_cast<int*>(g.a.orig) + (size_t)b * g.a.cap;
  {
    const int per = (n + csize - 1) / csize;
    for (int i = rank * per + tid; i < min((rank + 1) * per, n); i += blockDim.x) {
      const int src = NmsKey<T>::kPayload ? (int)L_sortpl[i] - 1 : (int)(u32)L_sort[i];
      const T x1 = d[(size_t)src * 5], y1 = d[(size_t)src * 5 + 1], x2 = d[(size_t)src * 5 + 2], y2 = d[(size_t)src * 5 + 3];
      nan = nan || (x1 != x1) || (y1 != y1) || (x2 != x2) || (y2 != y2);
      Box bx;
      bx.x = x1;
      bx.y = y1;
      bx.z = x2;
      bx.w = y2;
      box[i] = bx;
      area[i] = box_area<T>(x1, y1, x2, y2);
      orig[i] = src;
    }
  }
  if (__syncthreads_or(nan) && tid == 0) atomicOr(&g.ctrl[b].nan_flag, 1);
  cluster.sync();
  // ---- greedy walk ----
  StripeWalker<T> wk;
  wk.box = box;
  wk.area = area;
  wk.kept_slot = g.a.kept_slot + (size_t)b * g.a.keep_cap;
  wk.sM = sM;
  wk.s_removed = s_removed;
  wk.s_status = s_status;
  wk.L_sM = cluster.map_shared_rank(sM, 0);
  wk.L_removed = cluster.map_shared_rank(s_removed, 0);
  wk.t_box = t_box;
  wk.t_area = t_area;
  wk.thr = g.thr;
  wk.max_keep = g.max_keep;
  wk.nk = 0;
  wk.fin = 0;
  wk.stripes_done = 0;
  wk.walk(cluster, 0, n);
  if (rank == 0) {
    Ctrl* cb = g.ctrl + b;
    if (tid == 0) {
      cb->n = n;
      cb->nkept = wk.nk;
      cb->done = 1;
    }
    __syncthreads();
    coop_write_outputs<T>(g.a, g.o, cb, b, wk.nk, n);
  }
  cluster.sync();
}

template <typename T>
static inline size_t nms_cluster_smem_bytes(int sort_cap) {
  return (size_t)sort_cap * (NmsKey<T>::kPayload ? 12 : 8) + cluster_nms_smem_bytes<T>();
}

}  // namespace mxr

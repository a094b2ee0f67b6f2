// nms(dets, thresh) (helper/processing/nms.py:4-38) as ONE launch, second generation: a thread-block cluster per problem
// (image, or class in the per-class loop of rcnn/tester.py:50-56) with the stripe engine of the fused proposal kernel
// (proposal_fused.cuh):
//   * the leader CTA orders the scores (nms.py:19; ties: higher index first), all CTAs gather boxes and areas
//     (nms.py:18) into processing order -- as in nms_cluster.cuh;
//   * stripes of up to 512 boxes: every CTA stages the float32 image of the stripe's boxes with a per-box tolerance,
//     32 x 32 IoU bit blocks are spread over all warps of the cluster and tested in float32 (pair_d); a pair inside the
//     tolerance is decided by the reference's arithmetic in the dtype of `dets` (iou_exceeds<T>), so the keep list is
//     the one numpy computes, bit for bit, for float32 and float64 dets alike;
//   * every CTA resolves the stripe itself (block-wide fixed point), one cluster barrier per stripe.
#pragma once
#include <cooperative_groups.h>

#include "nms_cluster.cuh"
#include "pair_screen.cuh"

namespace mxr {
namespace fz {

template <typename T>
struct NmsFusedArgs {
  const T* dets;        // (n_total, 5)
  const int* offsets;   // device, (P+1)
  typename Ar<T>::Box* box;  // [P][cap] processing order
  T* area;
  int* orig;
  int cap;
  float4* kbox;         // [P][keep_cap] kept boxes as the float32 screen sees them
  float4* kaux;         // [P][keep_cap] scaled area, tolerance, position (bits), -
  int keep_cap;
  NmsOutputs o;
  Ctrl* ctrl;
  T thr;                // in the dtype of dets, like numpy
  int max_keep;
  int sort_cap;         // sort slots in the leader's shared memory (power of two >= the largest problem, >= 1024)
};

// float32 image of one box given in T, with the tolerance of its pairs (see decode32 / DESIGN.md section 3).  The
// coordinates are either converted from float64 (error 2^-24 relative each) or exact float32 whose differences the
// reference itself rounds in float32: both are covered by e = 2^-22 (largest |coordinate| + 1).
template <typename T>
__device__ __forceinline__ Box32 box32_of(const typename Ar<T>::Box bx, T area_t, float c32, bool screen_ok) {
  const float u2 = 1.1920929e-07f;  // 2^-23
  Box32 r;
  r.x1 = (float)bx.x;
  r.y1 = (float)bx.y;
  r.x2p = (float)bx.z + 1.0f;
  r.y2p = (float)bx.w + 1.0f;
  const float L = fmaxf(fmaxf(fabsf(r.x1), fabsf(r.y1)), fmaxf(fabsf(r.x2p), fabsf(r.y2p))) + 1.0f;
  const float e = 2.0f * u2 * L;
  const float bw = r.x2p - r.x1, bh = r.y2p - r.y1;
  const float area = (float)area_t;  // the reference's own area (nms.py:18), rounded once more
  r.a = c32 * area;
  const float t = 2.02f * ((bw + bh + 2.0f) * (4.0f * e + u2 * (L + 1.0f)) + 16.0f * (0.5f * u2) * fabsf(area) + 4.0f * e * e);
  const bool fine = screen_ok && e < 0.0625f && bw >= 1.0f && bh >= 1.0f && t < 1e30f && area > 0.0f;
  r.t = fine ? t : __int_as_float(0x7f800000);
  return r;
}

// grid (csize, P), cluster (csize,1,1), block 1024.
template <typename T>
__global__ void __launch_bounds__(1024, 1) nms_fused_kernel(const __grid_constant__ NmsFusedArgs<T> g) {
  namespace cgx = cooperative_groups;
  typedef typename Ar<T>::Box Box;
  extern __shared__ __align__(16) unsigned char dyn_smem[];
  __shared__ __align__(16) u32 s_keep[32];
  __shared__ __align__(16) u32 s_kfin[16];
  cgx::cluster_group cluster = cgx::this_cluster();
  const int rank = (int)cluster.block_rank();
  const int csize = (int)cluster.num_blocks();
  const int b = blockIdx.y;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int row0 = g.offsets[b];
  const int n = g.offsets[b + 1] - row0;
  const T* d = g.dets + (size_t)row0 * 5;
  // phase 1 (sort) and phase 2 (stripes) use the dynamic shared memory one after the other
  u64* sortbuf = reinterpret_cast<u64*>(dyn_smem);                                             // leader: [sort_cap]
  u32* sortpl = reinterpret_cast<u32*>(sortbuf + g.sort_cap);                                  // leader: [sort_cap] (f64 scores only)
  float4* s_cbox = reinterpret_cast<float4*>(dyn_smem);                 // [kWMax]
  float2* s_caux = reinterpret_cast<float2*>(s_cbox + kWMax);           // [kWMax]
  u32* s_rem = reinterpret_cast<u32*>(s_caux + kWMax);                  // [16]
  u32* s_removed = s_rem + 16;                                          // [2][16]
  u32* s_M = s_removed + 32;                                            // [2][kMWords][kWMax]
  u64* L_sort = cluster.map_shared_rank(sortbuf, 0);
  u32* L_sortpl = cluster.map_shared_rank(sortpl, 0);
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
  Box* box = g.box + (size_t)b * g.cap;
  T* area = g.area + (size_t)b * g.cap;
  int* orig = g.orig + (size_t)b * g.cap;
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
  cluster.sync();  // boxes visible to the cluster; the sort buffer is dead: its memory becomes the stripe staging
  // ---- greedy walk in stripes ----
  if (tid < 32) s_removed[tid] = 0u;
  if (tid < 16) s_rem[tid] = 0u;
  __syncthreads();
  const bool screen_ok = g.thr > T(0) && g.thr < T(1);
  const float c32 = screen_ok ? (float)((double)g.thr / (1.0 + (double)g.thr)) : 0.0f;
  float4* kbox = g.kbox + (size_t)b * g.keep_cap;
  float4* kaux = g.kaux + (size_t)b * g.keep_cap;
  const NmsOutputs& o = g.o;
  const size_t orow = o.row_offsets ? (size_t)o.row_offsets[b] : (size_t)b * o.rows;
  const int rows_out = o.row_offsets ? (o.row_offsets[b + 1] - o.row_offsets[b]) : o.rows;
  int nk = 0, fin = 0, stripes = 0;
  for (int c0 = 0; c0 < n && !fin; c0 += kWMax) {
    const int n_s = min(kWMax, n - c0);
    const int par = stripes & 1;
    u32* Mcur = s_M + par * (kWMax * kMWords);
    u32* Rcur = s_removed + par * 16;
    // stage: float32 images of the stripe's boxes
    if (tid < n_s) {
      const Box32 bx = box32_of<T>(ldcg_box<T>(box + c0 + tid), __ldcg(area + c0 + tid), c32, screen_ok);
      s_cbox[tid] = make_float4(bx.x1, bx.y1, bx.x2p, bx.y2p);
      s_caux[tid] = make_float2(bx.a, bx.t);
    }
    if (tid < 16) s_rem[tid] = 0u;
    __syncthreads();
    // IoU bit blocks of 32 x 32 pairs: in-stripe lower triangle + (kept so far) x stripe
    {
      const int n_cg = (n_s + 31) >> 5, n_rg = (nk + 31) >> 5;
      const int n_tri = n_cg * (n_cg + 1) / 2;
      const int n_tasks = n_tri + n_rg * n_cg;
#pragma unroll 1
      for (int u = warp * csize + rank; u < n_tasks; u += 32 * csize) {
        const bool tri = u < n_tri;
        int cg, ri;
        u32 mask = 0xffffffffu;
        Box32 r;
        if (tri) {
          int rb = (int)((sqrtf(8.0f * (float)u + 1.0f) - 1.0f) * 0.5f);
          while (rb * (rb + 1) / 2 > u) --rb;
          while ((rb + 1) * (rb + 2) / 2 <= u) ++rb;
          cg = u - rb * (rb + 1) / 2;
          ri = rb * 32 + lane;
          if (cg == rb) mask = (1u << lane) - 1u;
          if (ri >= n_s) mask = 0u;
          const float4 v = s_cbox[ri < n_s ? ri : 0];
          const float2 a = s_caux[ri < n_s ? ri : 0];
          r.x1 = v.x;
          r.y1 = v.y;
          r.x2p = v.z;
          r.y2p = v.w;
          r.a = a.x;
          r.t = a.y;
        } else {
          const int v = u - n_tri;
          const int rg = v / n_cg;
          cg = v - rg * n_cg;
          ri = rg * 32 + lane;
          if (ri >= nk) mask = 0u;
          const float4 v4 = __ldcg(kbox + (ri < nk ? ri : 0));
          const float4 a4 = __ldcg(kaux + (ri < nk ? ri : 0));
          r.x1 = v4.x;
          r.y1 = v4.y;
          r.x2p = v4.z;
          r.y2p = v4.w;
          r.a = a4.x;
          r.t = a4.y;
        }
        const int jn = min(32, n_s - cg * 32);
        if (jn < 32) mask &= (1u << jn) - 1u;
        u32 supp, und;
        block_bits(r, s_cbox + cg * 32, s_caux + cg * 32, jn, supp, und);
        supp &= mask;
        und &= mask;
        if (und) {  // rare: the reference's arithmetic in the dtype of dets decides
          const int prow = tri ? c0 + ri : __float_as_int(__ldcg(kaux + ri).z);
          const Box rb_ = ldcg_box<T>(box + prow);
          const T ra_ = __ldcg(area + prow);
          do {
            const int j = __ffs(und) - 1;
            und &= und - 1u;
            const Box cb_ = ldcg_box<T>(box + c0 + cg * 32 + j);
            if (iou_exceeds<T>(rb_.x, rb_.y, rb_.z, rb_.w, ra_, cb_.x, cb_.y, cb_.z, cb_.w, __ldcg(area + c0 + cg * 32 + j), g.thr))
              supp |= 1u << j;
            else
              supp &= ~(1u << j);
          } while (und);
        }
        if (tri) {
          if (ri < n_s) {
            u32* wp = Mcur + cg * kWMax + ri;
#pragma unroll 1
            for (int t = 0; t < csize; ++t) *cluster.map_shared_rank(wp, t) = supp;
          }
        } else {
          supp = __reduce_or_sync(0xffffffffu, supp);
          if (lane == 0 && supp) atomicOr(&s_rem[cg], supp);
        }
      }
      __syncthreads();
      if (tid < 16 * csize) {
        const u32 v = s_rem[tid & 15];
        if (v) atomicOr(cluster.map_shared_rank(Rcur + (tid & 15), tid >> 4), v);
      }
    }
    cluster.sync();  // ONE barrier per stripe (bit rows and removed words double-buffered by stripe parity)
    // every CTA resolves the stripe: fixed point of  keep_i = alive_i and no kept j < i suppresses i
    {
      if (tid < 16) {
        const int left = n_s - tid * 32;
        const u32 valid = left >= 32 ? 0xffffffffu : (left > 0 ? (1u << left) - 1u : 0u);
        s_keep[tid] = ~Rcur[tid] & valid;
        Rcur[tid] = 0u;
      }
      __syncthreads();
      if (warp < 8) {
        u32 ra[8], rb[16];
#pragma unroll
        for (int w = 0; w < 8; ++w) ra[w] = (tid < n_s && w <= warp) ? Mcur[w * kWMax + tid] : 0u;
#pragma unroll
        for (int w = 0; w < 16; ++w) rb[w] = (tid + 256 < n_s && w <= warp + 8) ? Mcur[w * kWMax + tid + 256] : 0u;
        const bool alive_a = tid < n_s && ((s_keep[warp] >> lane) & 1u);
        const bool alive_b = tid + 256 < n_s && ((s_keep[8 + warp] >> lane) & 1u);
        int cur = 0;
        u32 bal_a = 0, bal_b = 0;
        for (int it = 0; it <= n_s; ++it) {
          const uint4* kp4 = reinterpret_cast<const uint4*>(s_keep + cur * 16);
          const uint4 k0 = kp4[0], k1 = kp4[1], k2 = kp4[2], k3 = kp4[3];
          const u32 bad_a = (ra[0] & k0.x) | (ra[1] & k0.y) | (ra[2] & k0.z) | (ra[3] & k0.w) | (ra[4] & k1.x) | (ra[5] & k1.y) |
                            (ra[6] & k1.z) | (ra[7] & k1.w);
          const u32 bad_b = (rb[0] & k0.x) | (rb[1] & k0.y) | (rb[2] & k0.z) | (rb[3] & k0.w) | (rb[4] & k1.x) | (rb[5] & k1.y) |
                            (rb[6] & k1.z) | (rb[7] & k1.w) | (rb[8] & k2.x) | (rb[9] & k2.y) | (rb[10] & k2.z) |
                            (rb[11] & k2.w) | (rb[12] & k3.x) | (rb[13] & k3.y) | (rb[14] & k3.z) | (rb[15] & k3.w);
          bal_a = __ballot_sync(0xffffffffu, alive_a && bad_a == 0u);
          bal_b = __ballot_sync(0xffffffffu, alive_b && bad_b == 0u);
          int changed = 0;
          if (lane == 0) {
            changed = (bal_a != s_keep[cur * 16 + warp] || bal_b != s_keep[cur * 16 + 8 + warp]) ? 1 : 0;
            s_keep[(cur ^ 1) * 16 + warp] = bal_a;
            s_keep[(cur ^ 1) * 16 + 8 + warp] = bal_b;
          }
          cur ^= 1;
          int any;
          asm volatile(
              "{\n .reg .pred p, q;\n setp.ne.s32 p, %1, 0;\n bar.red.or.pred q, 1, 256, p;\n selp.s32 %0, 1, 0, q;\n}"
              : "=r"(any)
              : "r"(changed)
              : "memory");
          if (!any) break;
        }
        if (lane == 0) {
          s_kfin[warp] = bal_a;
          s_kfin[8 + warp] = bal_b;
        }
      }
      __syncthreads();
      int before = 0, total = 0;
      {
        const uint4* kp4 = reinterpret_cast<const uint4*>(s_kfin);
        const uint4 k0 = kp4[0], k1 = kp4[1], k2 = kp4[2], k3 = kp4[3];
        const u32 kw[16] = {k0.x, k0.y, k0.z, k0.w, k1.x, k1.y, k1.z, k1.w, k2.x, k2.y, k2.z, k2.w, k3.x, k3.y, k3.z, k3.w};
#pragma unroll
        for (int w = 0; w < 16; ++w) {
          const int c = __popc(kw[w]);
          before += w < warp ? c : 0;
          total += c;
        }
      }
      const u32 myw = s_kfin[warp & 15];
      const bool mine = tid < n_s && ((myw >> lane) & 1u);
      const int idx = nk + before + __popc(myw & ((1u << lane) - 1u));
      int nk2 = nk + total, f = 0;
      if (g.max_keep > 0 && nk2 >= g.max_keep) {
        nk2 = g.max_keep;
        f = 1;
      }
      if (mine && (g.max_keep <= 0 || idx < g.max_keep)) {
        if (!f && c0 + kWMax < n) {  // rows of the later stripes' blocks (same values from every CTA)
          kbox[idx] = s_cbox[tid];
          const float2 a = s_caux[tid];
          kaux[idx] = make_float4(a.x, a.y, __int_as_float(c0 + tid), 0.0f);
        }
        const int lg = warp / csize;
        if (warp - lg * csize == rank && o.keep_orig && idx < rows_out) o.keep_orig[orow + idx] = __ldcg(orig + c0 + tid);
      }
      nk = nk2;
      fin = f;
      __syncthreads();  // the staging arrays are rewritten by the next stripe
    }
    ++stripes;
  }
  if (rank == 0) {
    Ctrl* cb = g.ctrl + b;
    if (tid == 0) {
      cb->n = n;
      cb->nkept = nk;
      cb->done = 1;
    }
    for (int j = nk + tid; j < rows_out; j += blockDim.x)
      if (o.keep_orig) o.keep_orig[orow + j] = -1;
    __syncthreads();
    if (tid == 0 && o.nkeep) {
      o.nkeep[b * 2] = nk;
      o.nkeep[b * 2 + 1] = __ldcg(&cb->nan_flag) ? 1 : 0;
    }
  }
}

template <typename T>
static inline size_t nms_fused_smem_bytes(int sort_cap) {
  const size_t sort_bytes = (size_t)sort_cap * (NmsKey<T>::kPayload ? 12 : 8);
  const size_t stripe_bytes = (size_t)kWMax * (sizeof(float4) + sizeof(float2)) + (16 + 32) * sizeof(u32) +
                              (size_t)2 * kWMax * kMWords * sizeof(u32);
  return (sort_bytes > stripe_bytes ? sort_bytes : stripe_bytes) + 64;
}

}  // namespace fz
}  // namespace mxr

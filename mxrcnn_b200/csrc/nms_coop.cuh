// Greedy NMS as ONE persistent cooperative kernel (148 CTAs x 1024 threads, one per SM).
//
// The sorted list is walked in stripes of 1024 boxes.  Per stripe:
//   phase A (all CTAs, one warp per task): 32x64 IoU bit tiles of the stripe's upper triangle -> M (global),
//            and (boxes kept in earlier stripes) x (stripe columns) OR-reduced into R.
//   grid.sync
//   phase B (one CTA per problem): the stripe's bit rows are pulled into shared memory (17-word pitch)
//            and walked greedily, 64 boxes per step: warp 0 resolves the 64x64 diagonal tile with a
//            shuffle-fed chain, then all warps OR the kept rows into the later words.
//   grid.sync
// The loop ends as soon as every problem is done (post-NMS quota reached or list exhausted), so the
// common early-exit case costs one stripe and no further launches.  An optional deferred phase
// ("tier 2") lets the proposal op sort only the head of its candidate list up front and sort the
// rest only if some image actually needs it.
#pragma once
#include <cooperative_groups.h>

#include "nms_kernels.cuh"

namespace mxr {

namespace cg = cooperative_groups;

static const int kCoopW = 1024;      // stripe width
static const int kCoopSW = 16;       // words per stripe row
static const int kCoopPitch = 17;    // shared-memory row pitch in words (bank spread)
static const int kCoopTri = 136;     // 64x64 tiles in the stripe's upper triangle (16*17/2)

// Data that other CTAs rewrite while this persistent kernel runs is read with ld.global.cg
// (L2 only): L1 is not coherent across SMs.
__device__ __forceinline__ Ctrl ld_ctrl(const Ctrl* p) {
  Ctrl c;
  const int4 v = __ldcg(reinterpret_cast<const int4*>(p));
  const int4 w = __ldcg(reinterpret_cast<const int4*>(p) + 1);
  c.n_valid = v.x;
  c.nan_flag = v.y;
  c.n = v.z;
  c.nkept = v.w;
  c.done = w.x;
  c.pad[0] = w.y;
  c.pad[1] = w.z;
  c.pad[2] = w.w;
  return c;
}

struct NoTier2 {
  __device__ __forceinline__ bool wanted(int) const { return false; }
  __device__ __forceinline__ void run(int, Ctrl*, unsigned char*, u32*, int*) const {}
};

template <typename T>
struct CoopArgs {
  Ctrl* ctrl;
  NmsArrays<T> a;      // M: [b][1024][16] words; R, kept_slot as in the striped kernels
  NmsOutputs o;
  int B;
  int n_max;           // upper bound of ctrl[b].n
  int max_keep;        // > 0: stop a problem once this many boxes are kept
  int tier2_at;        // stripe index before which the deferred phase runs (-1: never)
  T thr;
};

template <typename T>
__device__ __forceinline__ void coop_write_outputs(const NmsArrays<T>& a, const NmsOutputs& o, const Ctrl* cb, int b, int nk,
                                                   int n) {
  const typename Ar<T>::Box* box = a.box + (size_t)b * a.cap;
  const int* orig = a.orig + (size_t)b * a.cap;
  const float* score = a.score ? a.score + (size_t)b * a.cap : nullptr;
  const int* kept_slot = a.kept_slot + (size_t)b * a.keep_cap;
  const size_t orow = o.row_offsets ? (size_t)o.row_offsets[b] : (size_t)b * o.rows;
  const int rows = o.row_offsets ? (o.row_offsets[b + 1] - o.row_offsets[b]) : o.rows;
  for (int j = threadIdx.x; j < rows; j += blockDim.x) {
    if (j < nk) {
      const int s = kept_slot[j];
      if (o.rois) {
        const typename Ar<T>::Box bx = box[s];
        float* r = o.rois + (orow + j) * 5;
        r[0] = (float)b;
        r[1] = (float)bx.x;
        r[2] = (float)bx.y;
        r[3] = (float)bx.z;
        r[4] = (float)bx.w;
      }
      if (o.scores) o.scores[orow + j] = score[s];
      if (o.keep_orig) o.keep_orig[orow + j] = orig[s];
      if (o.keep_pos) {
        int rs_ = s, re_ = s + 1;  // undo the tie-run reversal of nms.py:19 (see nms_scan_kernel)
        if (score) {
          const float sc = score[s];
          while (rs_ > 0 && score[rs_ - 1] == sc) --rs_;
          while (re_ < n && score[re_] == sc) ++re_;
        }
        o.keep_pos[orow + j] = rs_ + (re_ - 1 - s);
      }
    } else {
      if (o.rois) {
        float* r = o.rois + (orow + j) * 5;
        r[0] = r[1] = r[2] = r[3] = r[4] = 0.f;
      }
      if (o.scores) o.scores[orow + j] = 0.f;
      if (o.keep_orig) o.keep_orig[orow + j] = -1;
      if (o.keep_pos) o.keep_pos[orow + j] = -1;
    }
  }
  if (threadIdx.x == 0 && o.nkeep) {
    o.nkeep[b * 2] = nk;
    o.nkeep[b * 2 + 1] = cb->nan_flag ? 1 : 0;
  }
}

// Phase A, one warp task.  u < 2*kCoopTri: half (32 rows) of in-stripe tile u>>1; else a 32-row chunk of
// the kept list against column word cb.
template <typename T>
__device__ __forceinline__ void coop_tile_task(const CoopArgs<T>& g, int b, int u, int c0, typename Ar<T>::Box* s_box,
                                               T* s_area) {
  const Ctrl c = ld_ctrl(g.ctrl + b);
  if (c.done || c0 >= c.n || c0 < c.pad[0] * kCoopW) return;  // pad[0]: first stripe still to walk
  const int lane = threadIdx.x & 31;
  const int c1 = min(c0 + kCoopW, c.n);
  const typename Ar<T>::Box* box = g.a.box + (size_t)b * g.a.cap;
  const T* area = g.a.area + (size_t)b * g.a.cap;
  int cb, rb = 0, half = 0, rc = 0;
  const bool tri = u < 2 * kCoopTri;
  if (tri) {
    int t = u >> 1;
    half = u & 1;
    while (t >= kCoopSW - rb) {  // row-major walk of the upper triangle
      t -= kCoopSW - rb;
      ++rb;
    }
    cb = rb + t;
  } else {
    const int v = u - 2 * kCoopTri;
    cb = v % kCoopSW;
    rc = v / kCoopSW;
    if (rc * 32 >= c.nkept) return;
  }
  const int col0 = c0 + cb * 64;
  if (col0 >= c1) return;
  const int ncols = min(64, c1 - col0);
  for (int j = lane; j < ncols; j += 32) {
    s_box[j] = box[col0 + j];
    s_area[j] = area[col0 + j];
  }
  __syncwarp();
  if (tri) {
    const int rl = rb * 64 + half * 32 + lane;  // row within the stripe
    const int row = c0 + rl;
    if (row < c1) {
      const u64 w = tile_word<T>(box[row], area[row], s_box, s_area, rb == cb ? (half * 32 + lane + 1) : 0, ncols, g.thr);
      g.a.M[(size_t)b * g.a.m_stride + (size_t)rl * kCoopSW + cb] = w;
    }
  } else {
    u64 w = 0;
    const int kidx = rc * 32 + lane;
    if (kidx < c.nkept) {
      const int slot = __ldcg(g.a.kept_slot + (size_t)b * g.a.keep_cap + kidx);
      w = tile_word<T>(box[slot], area[slot], s_box, s_area, 0, ncols, g.thr);
    }
    const u32 lo = __reduce_or_sync(0xffffffffu, (u32)w);
    const u32 hi = __reduce_or_sync(0xffffffffu, (u32)(w >> 32));
    if (lane == 0 && (lo | hi)) atomicOr(&g.a.R[(size_t)b * g.a.r_words + (col0 >> 6)], ((u64)hi << 32) | lo);
  }
  __syncwarp();
}

// Phase B for one problem (whole CTA).  smem: sM[1024*17] u64, removed[16] u64.
template <typename T>
__device__ __forceinline__ void coop_scan(const CoopArgs<T>& g, int b, int c0, u64* sM, u64* s_removed, u64* s_keptbits,
                                          int* s_finish) {
  Ctrl* cb = g.ctrl + b;
  const Ctrl cv = ld_ctrl(cb);
  if (cv.done || c0 < cv.pad[0] * kCoopW) return;
  const int n = cv.n;
  const int tid = threadIdx.x;
  int nk = cv.nkept;
  bool finished = (n <= c0);
  const int c1 = min(c0 + kCoopW, n);
  int* kept_slot = g.a.kept_slot + (size_t)b * g.a.keep_cap;
  if (!finished) {
    const int rows = c1 - c0;
    const int nblk = (rows + 63) >> 6;
    const u64* M = g.a.M + (size_t)b * g.a.m_stride;
    // pull the stripe's bit rows into shared memory (only words at or right of the diagonal are defined)
    for (int i = tid; i < rows * kCoopSW; i += blockDim.x) {
      const int r = i >> 4, w = i & 15;
      if (w >= (r >> 6) && w < nblk) sM[r * kCoopPitch + w] = __ldcg(M + i);
    }
    if (tid < kCoopSW) s_removed[tid] = (tid < nblk) ? __ldcg(g.a.R + (size_t)b * g.a.r_words + (c0 >> 6) + tid) : 0ull;
    for (int blk = 0; blk < nblk; ++blk) {
      __syncthreads();
      if (tid < 32) {
        const u64 word0 = s_removed[blk];
        const int nvalid = min(64, rows - blk * 64);
        const u64 valid = nvalid == 64 ? ~0ull : ((1ull << nvalid) - 1ull);
        const u64 d0 = (blk * 64 + tid < rows) ? sM[(blk * 64 + tid) * kCoopPitch + blk] : 0ull;
        const u64 d1 = (blk * 64 + 32 + tid < rows) ? sM[(blk * 64 + 32 + tid) * kCoopPitch + blk] : 0ull;
        u32 lo = (u32)(word0 | ~valid), hi = (u32)((word0 | ~valid) >> 32);
        const u32 d0lo = (u32)d0, d0hi = (u32)(d0 >> 32), d1hi = (u32)(d1 >> 32);
#pragma unroll
        for (int t = 0; t < 32; ++t) {
          const u32 rlo = __shfl_sync(0xffffffffu, d0lo, t);
          const u32 rhi = __shfl_sync(0xffffffffu, d0hi, t);
          if (!((lo >> t) & 1u)) {
            lo |= rlo;
            hi |= rhi;
          }
        }
#pragma unroll
        for (int t = 0; t < 32; ++t) {
          const u32 rhi = __shfl_sync(0xffffffffu, d1hi, t);  // rows 32..63 only reach columns > 32
          if (!((hi >> t) & 1u)) hi |= rhi;
        }
        u64 kept = ~(((u64)hi << 32) | lo) & valid;
        int fin = 0;
        if (g.max_keep > 0 && nk + __popcll(kept) >= g.max_keep) {
          int want = g.max_keep - nk;
          u64 k2 = 0, rest = kept;
          while (want > 0 && rest) {
            const u64 low = rest & (~rest + 1ull);
            k2 |= low;
            rest ^= low;
            --want;
          }
          kept = k2;
          fin = 1;
        }
        if (tid == 0) {
          *s_keptbits = kept;
          *s_finish = fin;
        }
      }
      __syncthreads();
      const u64 kept = *s_keptbits;
      const int fin = *s_finish;
      if (tid < 64 && ((kept >> tid) & 1ull)) kept_slot[nk + __popcll(kept & ((1ull << tid) - 1ull))] = c0 + blk * 64 + tid;
      nk += __popcll(kept);
      if (fin) {
        finished = true;
        break;
      }
      // OR the rows kept in this block into the removed words of the later blocks: thread (w, row)
      const int nw = nblk - blk - 1;
      const int w = tid >> 6, row = tid & 63;
      if (w < nw) {
        const u64 v = ((kept >> row) & 1ull) ? sM[(blk * 64 + row) * kCoopPitch + blk + 1 + w] : 0ull;
        const u32 vlo = __reduce_or_sync(0xffffffffu, (u32)v);
        const u32 vhi = __reduce_or_sync(0xffffffffu, (u32)(v >> 32));
        if ((tid & 31) == 0 && (vlo | vhi)) atomicOr(&s_removed[blk + 1 + w], ((u64)vhi << 32) | vlo);
      }
    }
    if (c1 >= n) finished = true;
  }
  __syncthreads();
  if (tid == 0) {
    cb->nkept = nk;
    if (finished) cb->done = 1;
  }
  if (finished) coop_write_outputs<T>(g.a, g.o, cb, b, nk, n);
  __syncthreads();
}

// Dynamic smem: max(1024*17*8 + 128 (scan), 32 warps * 64 * (sizeof(Box)+sizeof(T)) (tiles), tier-2 needs).
template <typename T, typename Tier2>
__global__ void __launch_bounds__(1024, 1) nms_coop_kernel(CoopArgs<T> g, Tier2 t2) {
  extern __shared__ __align__(16) unsigned char coop_smem[];
  __shared__ u64 s_keptbits;
  __shared__ int s_finish;
  __shared__ u32 s_st[8];
  __shared__ int s_cnt;
  cg::grid_group grid = cg::this_grid();
  const int warp = threadIdx.x >> 5;
  const int n64 = (max(g.n_max, 1) + 63) & ~63;
  const int nstripes = (n64 + kCoopW - 1) / kCoopW;
  for (int s = 0; s < nstripes; ++s) {
    const int c0 = s * kCoopW;
    // anything left to do?  (ctrl is stable here: written before the previous grid.sync / kernel boundary)
    int active = 0, remaining = 0;
    for (int b = threadIdx.x; b < g.B; b += blockDim.x) {
      const Ctrl c = ld_ctrl(g.ctrl + b);
      remaining |= c.done ? 0 : 1;
      active |= (!c.done && s >= c.pad[0] && (c0 < c.n || s == 0)) ? 1 : 0;
    }
    remaining = __syncthreads_or(remaining);
    active = __syncthreads_or(active);
    if (!remaining) break;
    if (!active) continue;  // e.g. every remaining problem resumes at a later stripe
    if (s == g.tier2_at) {
      for (int b = blockIdx.x; b < g.B; b += gridDim.x) {
        if (!ld_ctrl(g.ctrl + b).done && t2.wanted(b)) t2.run(b, g.ctrl, coop_smem, s_st, &s_cnt);
        __syncthreads();
      }
      grid.sync();
    }
    // ---- phase A ----
    {
      int prev_rows = c0;
      if (g.max_keep > 0 && prev_rows > g.max_keep - 1) prev_rows = g.max_keep - 1;
      const int tp = 2 * kCoopTri + kCoopSW * ((prev_rows + 31) / 32);
      typename Ar<T>::Box* s_box = reinterpret_cast<typename Ar<T>::Box*>(coop_smem) + warp * 64;
      T* s_area = reinterpret_cast<T*>(coop_smem + 32 * 64 * sizeof(typename Ar<T>::Box)) + warp * 64;
      const int total = g.B * tp;
      for (int t = blockIdx.x * 32 + warp; t < total; t += gridDim.x * 32) coop_tile_task<T>(g, t / tp, t % tp, c0, s_box, s_area);
    }
    grid.sync();
    // ---- phase B ----
    {
      u64* sM = reinterpret_cast<u64*>(coop_smem);
      u64* s_removed = sM + kCoopW * kCoopPitch;
      for (int b = blockIdx.x; b < g.B; b += gridDim.x) coop_scan<T>(g, b, c0, sM, s_removed, &s_keptbits, &s_finish);
    }
    grid.sync();
  }
}

static inline size_t coop_smem_bytes(size_t box_bytes, size_t elem_bytes, size_t tier2_bytes) {
  size_t a = (size_t)kCoopW * kCoopPitch * 8 + 256;
  size_t b = 32 * 64 * (box_bytes + elem_bytes);
  size_t m = a > b ? a : b;
  return m > tier2_bytes ? m : tier2_bytes;
}

}  // namespace mxr

// Greedy NMS (helper/processing/nms.py:4-38) on boxes that are already in processing
// order, as "stripes": the sorted list is cut into column stripes [c0, c1).  For each
// stripe one launch evaluates
//   (a) the in-stripe 64x64 IoU bit tiles (upper triangle)            -> M
//   (b) kept boxes of earlier stripes x this stripe's columns          -> R (OR-reduced)
// and a second launch (one CTA per problem) walks the stripe greedily, 64 boxes per step.
// Work is (kept so far) x (stripe width) + (stripe triangle), never the full n^2/2
// triangle, scratch is bounded by the stripe width, and a problem that reaches its
// post-NMS quota sets `done`, which turns every later launch into a no-op for it.
#pragma once
#include "device_math.cuh"

namespace mxr {

template <typename T>
struct NmsArrays {
  // per-problem arrays, problem b at [b * cap]
  const typename Ar<T>::Box* box;  // processing order
  const T* area;
  const float* score;   // may be null (generic nms)
  const int* orig;      // original row (generic nms) / (h,w,a) anchor row (proposal)
  int cap;
  // stripe scratch
  u64* M;               // [b][stripe_rows][m_pitch]
  size_t m_stride;      // words per problem
  int m_pitch;          // words per row
  u64* R;               // [b][r_words] removed-by-earlier-stripes, absolute word index
  int r_words;
  int* kept_slot;       // [b][keep_cap]
  int keep_cap;
};

template <typename T>
__device__ __forceinline__ u64 tile_word(const typename Ar<T>::Box rb, T rarea, const typename Ar<T>::Box* cols,
                                         const T* careas, int j_begin, int j_end, T thr) {
  u64 word = 0;
#pragma unroll 4
  for (int j = j_begin; j < j_end; ++j) {
    const typename Ar<T>::Box cb = cols[j];
    if (iou_exceeds<T>(rb.x, rb.y, rb.z, rb.w, rarea, cb.x, cb.y, cb.z, cb.w, careas[j], thr)) word |= 1ull << j;
  }
  return word;
}

// grid (sw*sw + sw*row_chunks, 1, B), block 64.
template <typename T>
__global__ void __launch_bounds__(64) nms_stripe_kernel(const Ctrl* __restrict__ ctrl, NmsArrays<T> a, int c0, int width,
                                                        int sw, T thr) {
  const int b = blockIdx.z;
  const Ctrl c = ctrl[b];
  if (c.done) return;
  const int n = c.n;
  if (c0 >= n) return;
  const int c1 = min(c0 + width, n);
  __shared__ typename Ar<T>::Box s_box[64];
  __shared__ T s_area[64];
  __shared__ u32 s_red[4];
  const typename Ar<T>::Box* box = a.box + (size_t)b * a.cap;
  const T* area = a.area + (size_t)b * a.cap;
  const int i = threadIdx.x;
  const int t = blockIdx.x;
  if (t < sw * sw) {
    const int rb = t / sw, cb = t % sw;
    if (rb > cb) return;
    const int row0 = c0 + rb * 64, col0 = c0 + cb * 64;
    if (col0 >= c1) return;
    const int ncols = min(64, c1 - col0);
    if (i < ncols) {
      s_box[i] = box[col0 + i];
      s_area[i] = area[col0 + i];
    }
    __syncthreads();
    const int row = row0 + i;
    if (row < c1) {
      const typename Ar<T>::Box r = box[row];
      const u64 w = tile_word<T>(r, area[row], s_box, s_area, rb == cb ? i + 1 : 0, ncols, thr);
      a.M[(size_t)b * a.m_stride + (size_t)(rb * 64 + i) * a.m_pitch + cb] = w;
    }
  } else {
    const int u = t - sw * sw;
    const int cb = u % sw, rc = u / sw;
    const int nk = c.nkept;
    if (rc * 64 >= nk) return;
    const int col0 = c0 + cb * 64;
    if (col0 >= c1) return;
    const int ncols = min(64, c1 - col0);
    if (i < ncols) {
      s_box[i] = box[col0 + i];
      s_area[i] = area[col0 + i];
    }
    __syncthreads();
    u64 w = 0;
    const int kidx = rc * 64 + i;
    if (kidx < nk) {
      const int slot = a.kept_slot[(size_t)b * a.keep_cap + kidx];
      w = tile_word<T>(box[slot], area[slot], s_box, s_area, 0, ncols, thr);
    }
    u32 lo = __reduce_or_sync(0xffffffffu, (u32)w);
    u32 hi = __reduce_or_sync(0xffffffffu, (u32)(w >> 32));
    if ((i & 31) == 0) {
      s_red[(i >> 5) * 2] = lo;
      s_red[(i >> 5) * 2 + 1] = hi;
    }
    __syncthreads();
    if (i == 0) {
      const u64 all = ((u64)(s_red[1] | s_red[3]) << 32) | (u64)(s_red[0] | s_red[2]);
      if (all) atomicOr(&a.R[(size_t)b * a.r_words + (col0 >> 6)], all);
    }
  }
}

struct NmsOutputs {
  float* rois;        // [b][rows][5] or null
  float* scores;      // [b][rows] or null
  int* keep_pos;      // [b][rows] or null: index into the descending pre-NMS list (reference `keep`)
  int* keep_orig;     // [b][rows] or null: orig[] of the kept boxes
  int* nkeep;         // [b][2]: nkeep, flags
  int rows;           // rows per problem in the outputs above
  const int* row_offsets;  // optional: problem b writes keep_orig at row_offsets[b] (batched per-class nms)
};

// One CTA per problem.  Walks stripe [c0, c0+width) greedily in blocks of 64 boxes.
// max_keep > 0: stop once that many boxes are kept (keep[:post] is a prefix of the full
// greedy result, rcnn/rpn/proposal.py:132-133).
template <typename T>
__global__ void __launch_bounds__(256) nms_scan_kernel(Ctrl* __restrict__ ctrl, NmsArrays<T> a, NmsOutputs o, int c0,
                                                       int width, int max_keep, int last_stripe) {
  const int b = blockIdx.x;
  Ctrl* cb = ctrl + b;
  if (cb->done) return;
  const int n = cb->n;
  const int tid = threadIdx.x;
  extern __shared__ u64 s_removed[];  // width/64 words
  __shared__ u64 s_keptbits;
  __shared__ int s_finish;
  int nk = cb->nkept;
  bool finished = (n <= c0);
  const int c1 = min(c0 + width, n);
  int* kept_slot = a.kept_slot + (size_t)b * a.keep_cap;
  if (!finished) {
    const int nblk = (c1 - c0 + 63) >> 6;
    const u64* M = a.M + (size_t)b * a.m_stride;
    for (int w = tid; w < nblk; w += blockDim.x) s_removed[w] = a.R[(size_t)b * a.r_words + (c0 >> 6) + w];
    for (int blk = 0; blk < nblk; ++blk) {
      __syncthreads();
      if (tid < 32) {
        u64 word = s_removed[blk];
        const int nvalid = min(64, c1 - c0 - blk * 64);
        const u64 valid = nvalid == 64 ? ~0ull : ((1ull << nvalid) - 1ull);
        word |= ~valid;
        const u64 d0 = M[(size_t)(blk * 64 + tid) * a.m_pitch + blk];
        const u64 d1 = M[(size_t)(blk * 64 + 32 + tid) * a.m_pitch + blk];
#pragma unroll 8
        for (int t = 0; t < 32; ++t) {
          const u64 row = __shfl_sync(0xffffffffu, d0, t);
          if (!((word >> t) & 1ull)) word |= row;
        }
#pragma unroll 8
        for (int t = 0; t < 32; ++t) {
          const u64 row = __shfl_sync(0xffffffffu, d1, t);
          if (!((word >> (t + 32)) & 1ull)) word |= row;
        }
        u64 kept = ~word & valid;
        int fin = 0;
        if (max_keep > 0 && nk + __popcll(kept) >= max_keep) {
          // keep only the first (max_keep - nk) set bits
          int want = max_keep - nk;
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
          s_keptbits = kept;
          s_finish = fin;
        }
      }
      __syncthreads();
      const u64 kept = s_keptbits;
      const int fin = s_finish;
      if (tid < 64 && ((kept >> tid) & 1ull)) {
        const int j = nk + __popcll(kept & ((1ull << tid) - 1ull));
        kept_slot[j] = c0 + blk * 64 + tid;
      }
      nk += __popcll(kept);
      if (fin) {
        finished = true;
        break;
      }
      // OR the rows of the boxes kept in this block into the removed words of later blocks
      const int nw = nblk - blk - 1;
      if (nw > 0 && kept) {
        const int w = tid % nw;
        const int nslots = max(1, (int)blockDim.x / nw);
        const int rs = tid / nw;
        if (rs < nslots) {
          u64 acc = 0;
          u64 rest = kept;
          int idx = 0;
          while (rest) {
            const int tbit = __ffsll((long long)rest) - 1;
            rest &= rest - 1ull;
            if ((idx % nslots) == rs) acc |= M[(size_t)(blk * 64 + tbit) * a.m_pitch + blk + 1 + w];
            ++idx;
          }
          if (acc) atomicOr(&s_removed[blk + 1 + w], acc);
        }
      }
    }
    if (c1 >= n) finished = true;
  }
  __syncthreads();
  if (tid == 0) {
    cb->nkept = nk;
    if (finished) cb->done = 1;
  }
  if (!finished && !last_stripe) return;
  // ---- outputs (rcnn/rpn/proposal.py:140-149) ----
  const typename Ar<T>::Box* box = a.box + (size_t)b * a.cap;
  const int* orig = a.orig + (size_t)b * a.cap;
  const float* score = a.score ? a.score + (size_t)b * a.cap : nullptr;
  const size_t orow = o.row_offsets ? (size_t)o.row_offsets[b] : (size_t)b * o.rows;
  const int rows = o.row_offsets ? (o.row_offsets[b + 1] - o.row_offsets[b]) : o.rows;
  for (int j = tid; j < rows; j += blockDim.x) {
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
        // the processing order reverses runs of equal scores (nms.py:19 re-sorts the
        // already sorted list); map back to the position in the top-k list
        int rs_ = s, re_ = s + 1;
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
  if (tid == 0 && o.nkeep) {
    o.nkeep[b * 2] = nk;
    o.nkeep[b * 2 + 1] = cb->nan_flag ? 1 : 0;
  }
}

}  // namespace mxr

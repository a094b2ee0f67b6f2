// Greedy NMS inside one thread-block cluster (shared by the fused proposal kernel and the generic
// nms kernel): boxes already in processing order in global memory, bit rows and suppression words in
// the leader CTA's shared memory, reached from the other CTAs through distributed shared memory.
#pragma once
#include <cooperative_groups.h>

#include "device_math.cuh"

namespace mxr {

static const int kCW = 256;        // stripe width (cluster barriers are cheap, so stripes are narrow)
static const int kCSW = 4;         // 64-bit words per stripe row
static const int kCPitch = 5;      // shared-memory row pitch in words
static const int kCTriTasks = 80;  // 10 lower-triangle tiles x 2 row halves x 4 column quarters

template <typename T>
__device__ __forceinline__ typename Ar<T>::Box ldcg_box(const typename Ar<T>::Box* p);
template <>
__device__ __forceinline__ double4 ldcg_box<double>(const double4* p) {
  const double2 a = __ldcg(reinterpret_cast<const double2*>(p));
  const double2 b = __ldcg(reinterpret_cast<const double2*>(p) + 1);
  return make_double4(a.x, a.y, b.x, b.y);
}
template <>
__device__ __forceinline__ float4 ldcg_box<float>(const float4* p) {
  return __ldcg(p);
}

// Descending bitonic sort of n_pow2 <= 1024 distinct 64-bit keys (entries beyond the real keys are zero-padded by the
// caller up to 1024), FOUR keys per thread on the first 256 threads (blockDim.x == 1024, all threads call it; the other
// 768 only wait at the end): strides 1-2 are register compare-exchanges, 4..64 warp shuffles, and only strides
// >= 128 -- 6 of the 55 stages -- go through shared memory, synchronised by a 256-thread named barrier
// instead of a block-wide one.
__device__ __forceinline__ void order_pair(u64& a, u64& b, bool desc) {
  const bool sw = desc ? (b > a) : (a > b);
  const u64 x = sw ? b : a, y = sw ? a : b;
  a = x;
  b = y;
}
__device__ __forceinline__ void bitonic_sort_1024_x4(u64* buf, int n_pow2) {
  const int t = threadIdx.x;
  if (t < 256) {
    u64 v[4];
    {
      const ulonglong2 lo = reinterpret_cast<const ulonglong2*>(buf)[2 * t], hi = reinterpret_cast<const ulonglong2*>(buf)[2 * t + 1];
      v[0] = lo.x;
      v[1] = lo.y;
      v[2] = hi.x;
      v[3] = hi.y;
    }
    // size 2: pairs (0,1) descending, (2,3) ascending
    order_pair(v[0], v[1], true);
    order_pair(v[2], v[3], false);
    for (int size = 4; size <= n_pow2; size <<= 1) {
      const bool desc = ((4 * t) & size) == 0;  // the same for the thread's four elements once size >= 4
      for (int stride = size >> 1; stride >= 4; stride >>= 1) {
        const int ts = stride >> 2;  // partner thread = t ^ ts
        u64 o[4];
        if (stride >= 128) {
          asm volatile("bar.sync 1, 256;" ::: "memory");
          reinterpret_cast<ulonglong2*>(buf)[2 * t] = make_ulonglong2(v[0], v[1]);
          reinterpret_cast<ulonglong2*>(buf)[2 * t + 1] = make_ulonglong2(v[2], v[3]);
          asm volatile("bar.sync 1, 256;" ::: "memory");
          const ulonglong2 lo = reinterpret_cast<const ulonglong2*>(buf)[2 * (t ^ ts)];
          const ulonglong2 hi = reinterpret_cast<const ulonglong2*>(buf)[2 * (t ^ ts) + 1];
          o[0] = lo.x;
          o[1] = lo.y;
          o[2] = hi.x;
          o[3] = hi.y;
        } else {
#pragma unroll
          for (int q = 0; q < 4; ++q) o[q] = __shfl_xor_sync(0xffffffffu, v[q], ts);
        }
        const bool take_max = (((t & ts) == 0) == desc);
#pragma unroll
        for (int q = 0; q < 4; ++q) v[q] = (take_max == (o[q] > v[q])) ? o[q] : v[q];
      }
      order_pair(v[0], v[2], desc);  // stride 2
      order_pair(v[1], v[3], desc);
      order_pair(v[0], v[1], desc);  // stride 1
      order_pair(v[2], v[3], desc);
    }
    asm volatile("bar.sync 1, 256;" ::: "memory");
    reinterpret_cast<ulonglong2*>(buf)[2 * t] = make_ulonglong2(v[0], v[1]);
    reinterpret_cast<ulonglong2*>(buf)[2 * t + 1] = make_ulonglong2(v[2], v[3]);
  }
  __syncthreads();
}

// 16 columns of one row: bit j set when column j is suppressed by the row box (nms.py:25-36).
template <typename T>
__device__ __forceinline__ u32 tile_bits16(const typename Ar<T>::Box rb, T rarea, const typename Ar<T>::Box* cols,
                                           const T* careas, int j_begin, int j_end, T thr, T thr_lo, T thr_hi) {
  u32 bits = 0;
  if (thr > T(0)) {
    // branch-free pass over the columns; the rare undecided pairs (inside the band, degenerate) are settled afterwards
    u32 undecided = 0;
#pragma unroll 4
    for (int j = j_begin; j < j_end; ++j) {
      const typename Ar<T>::Box cb = cols[j];
      const int r = iou_band<T>(rb.x, rb.y, rb.z, rb.w, rarea, cb.x, cb.y, cb.z, cb.w, careas[j], thr_lo, thr_hi);
      bits |= (u32)(r & 1) << j;
      undecided |= (u32)(r >> 1) << j;
    }
    while (undecided) {
      const int j = __ffs(undecided) - 1;
      undecided &= undecided - 1u;
      const typename Ar<T>::Box cb = cols[j];
      if (iou_exceeds<T>(rb.x, rb.y, rb.z, rb.w, rarea, cb.x, cb.y, cb.z, cb.w, careas[j], thr)) bits |= 1u << j;
    }
  } else {
    for (int j = j_begin; j < j_end; ++j) {
      const typename Ar<T>::Box cb = cols[j];
      if (iou_exceeds<T>(rb.x, rb.y, rb.z, rb.w, rarea, cb.x, cb.y, cb.z, cb.w, careas[j], thr)) bits |= 1u << j;
    }
  }
  return bits;
}

// Greedy walk of one stripe whose bit rows already sit in shared memory.  ONE WARP, no block barriers: the
// suppression words of the stripe's (at most four) 64-box blocks live in registers, each lane owns rows
// `lane` and `lane + 32` of the current block.
// Diagonal tiles hold their LOWER triangle (bit j of row i, j < i: "j suppresses i"; the IoU predicate is
// symmetric bit for bit), off-diagonal words hold "row suppresses column".
// Returns the new kept count; *fin_out != 0 when the quota was reached.
__device__ __forceinline__ int scan_rows_warp(const u64* sM, int pitch, const u64* s_removed, int rows, int c0, int nk,
                                              int max_keep, int* kept_slot, int* fin_out) {
  const int lane = threadIdx.x & 31;
  const int nblk = (rows + 63) >> 6;
  u64 rem[kCSW];
#pragma unroll
  for (int w = 0; w < kCSW; ++w) rem[w] = s_removed[w];
  int finished = 0;
#pragma unroll
  for (int blk = 0; blk < kCSW; ++blk) {
    if (blk < nblk && !finished) {
      const int nvalid = min(64, rows - blk * 64);
      const u64 valid = nvalid == 64 ? ~0ull : ((1ull << nvalid) - 1ull);
      const u64 alive = ~rem[blk] & valid;
      const int r0 = blk * 64 + lane, r1 = r0 + 32;
      // Greedy keep is the unique fixed point of  keep_i = alive_i and no kept j < i suppresses i ;
      // iterating from keep = alive fixes at least one more leading position per round, and in practice
      // converges in a handful of rounds instead of a 64-step dependent chain.
      const u64 m0 = (r0 < rows ? sM[r0 * pitch + blk] : 0ull) & ((1ull << lane) - 1ull);
      const u64 m1 = (r1 < rows ? sM[r1 * pitch + blk] : 0ull) & ((1ull << (lane + 32)) - 1ull);
      const bool a0 = (alive >> lane) & 1ull, a1 = (alive >> (lane + 32)) & 1ull;
      u64 keep = alive;
      for (int it = 0; it < 64; ++it) {
        const u32 k0 = __ballot_sync(0xffffffffu, a0 && (m0 & keep) == 0ull);
        const u32 k1 = __ballot_sync(0xffffffffu, a1 && (m1 & keep) == 0ull);
        const u64 nxt = ((u64)k1 << 32) | k0;
        if (nxt == keep) break;
        keep = nxt;
      }
      u64 kept = keep;
      if (max_keep > 0 && nk + __popcll(kept) >= max_keep) {
        int want = max_keep - nk;
        u64 k2 = 0, rest = kept;
        while (want > 0 && rest) {
          const u64 low = rest & (~rest + 1ull);
          k2 |= low;
          rest ^= low;
          --want;
        }
        kept = k2;
        finished = 1;
      }
      const bool k0 = (kept >> lane) & 1ull, k1 = (kept >> (lane + 32)) & 1ull;
      if (k0) kept_slot[nk + __popcll(kept & ((1ull << lane) - 1ull))] = c0 + r0;
      if (k1) kept_slot[nk + __popcll(kept & ((1ull << (lane + 32)) - 1ull))] = c0 + r1;
      nk += __popcll(kept);
      if (!finished) {
#pragma unroll
        for (int w = 0; w < kCSW; ++w) {
          if (w > blk && w < nblk) {  // boxes kept in this block suppress columns of the later blocks
            const u64 v = (k0 ? sM[r0 * pitch + w] : 0ull) | (k1 ? sM[r1 * pitch + w] : 0ull);
            const u32 vlo = __reduce_or_sync(0xffffffffu, (u32)v);
            const u32 vhi = __reduce_or_sync(0xffffffffu, (u32)(v >> 32));
            rem[w] |= ((u64)vhi << 32) | vlo;
          }
        }
      }
    }
  }
  *fin_out = finished;
  return nk;
}

// Everything the stripe walk needs; all shared-memory pointers are this CTA's own copies, the L_* ones
// address the leader (rank 0) through DSMEM.
template <typename T>
struct StripeWalker {
  const typename Ar<T>::Box* box;   // [n] processing order, global
  const T* area;
  int* kept_slot;                   // global, kept boxes so far
  u64* sM;                          // leader smem: [kCW][kCPitch]
  u64* s_removed;                   // leader smem: [kCSW]
  int* s_status;                    // every CTA: [2] kept so far, quota reached (written by the leader, 8-byte aligned)
  u64* L_sM;
  u64* L_removed;
  typename Ar<T>::Box* t_box;       // this warp's staging: [16]
  T* t_area;
  T thr;
  int max_keep;
  // walk state (identical in every thread of the cluster)
  int nk, fin, stripes_done;

  template <typename Cluster>
  __device__ __forceinline__ void walk(Cluster& cluster, int s_begin, int n_sorted) {
    const int rank = (int)cluster.block_rank();
    const int csize = (int)cluster.num_blocks();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const T thr_lo = Ar<T>::mul(thr, Ar<T>::band_lo()), thr_hi = Ar<T>::mul(thr, Ar<T>::band_hi());
    const int nstripes = (n_sorted + kCW - 1) / kCW;
    for (int s = s_begin; s < nstripes && !fin; ++s) {
      const int c0 = s * kCW;
      const int c1 = min(c0 + kCW, n_sorted);
      {
        // one warp per task (32 rows x 16 columns); task u runs on CTA u % csize, warp u / csize, so that a
        // stripe's few dozen tasks spread over all the cluster's SMs
        const int tp = kCTriTasks + 16 * ((nk + 31) / 32);
        for (int u = warp * csize + rank; u < tp; u += csize * 32) {
          int cb, qd, rb = 0, half = 0, rc = 0;
          const bool tri = u < kCTriTasks;
          if (tri) {
            int t = u >> 3;
            half = (u >> 2) & 1;
            qd = u & 3;
            while (t >= kCSW - rb) {
              t -= kCSW - rb;
              ++rb;
            }
            cb = rb + t;
          } else {
            const int v = u - kCTriTasks;
            qd = v & 3;
            cb = (v >> 2) & 3;
            rc = v >> 4;
          }
          const int col0 = c0 + cb * 64 + qd * 16;
          if (col0 >= c1) continue;
          const int ncols = min(16, c1 - col0);
          // this lane's row box is requested BEFORE the column boxes are staged, so that the two L2 round trips overlap
          const int rt = half * 32 + lane;           // tri: row inside the 64x64 tile
          const int rl = rb * 64 + rt;               //      row inside the stripe
          const int kidx = rc * 32 + lane;           // cross: index into the kept list
          int row = -1;
          if (tri) {
            if (c0 + rl < c1) row = c0 + rl;
          } else if (kidx < nk) {
            row = __ldcg(kept_slot + kidx);
          }
          typename Ar<T>::Box rbox = typename Ar<T>::Box();
          T rarea = T(0);
          if (row >= 0) {
            rbox = ldcg_box<T>(box + row);
            rarea = __ldcg(area + row);
          }
          __syncwarp();
          if (lane < ncols) {
            t_box[lane] = ldcg_box<T>(box + col0 + lane);
            t_area[lane] = __ldcg(area + col0 + lane);
          }
          __syncwarp();
          if (tri) {
            if (row >= 0) {
              int je = ncols;
              if (rb == cb) je = min(ncols, max(0, rt - qd * 16));   // diagonal tile: strictly BELOW the diagonal
              const u32 bits = tile_bits16<T>(rbox, rarea, t_box, t_area, 0, je, thr, thr_lo, thr_hi);
              reinterpret_cast<unsigned short*>(L_sM + rl * kCPitch + cb)[qd] = (unsigned short)bits;
            }
          } else {
            u32 bits = 0;
            if (row >= 0) bits = tile_bits16<T>(rbox, rarea, t_box, t_area, 0, ncols, thr, thr_lo, thr_hi);
            bits = __reduce_or_sync(0xffffffffu, bits);
            if (lane == 0 && bits) atomicOr(reinterpret_cast<u32*>(L_removed + cb) + (qd >> 1), bits << (16 * (qd & 1)));
          }
        }
      }
      cluster.sync();
      if (rank == 0 && warp == 0) {  // the leader's first warp walks the stripe; everybody else waits at the barrier
        int f = 0;
        const int nk2 = scan_rows_warp(sM, kCPitch, s_removed, c1 - c0, c0, nk, max_keep, kept_slot, &f);
        __syncwarp();
        if (lane < kCSW) s_removed[lane] = 0ull;
        // the result goes INTO every CTA's shared memory (one 8-byte remote store per CTA): after the barrier nobody
        // reads anybody else's shared memory any more, so the kernel needs no closing cluster barrier after a walk
        if (lane < csize) *reinterpret_cast<int2*>(cluster.map_shared_rank(s_status, lane)) = make_int2(nk2, f);
      }
      cluster.sync();
      nk = s_status[0];  // rewritten only after the next stripe's first barrier, which every thread passes after this read
      fin = s_status[1];
      stripes_done = s + 1;
    }
  }
};

// bytes of dynamic shared memory after the sort buffer: bit rows + suppression words + per-warp staging
template <typename T>
static inline size_t cluster_nms_smem_bytes() {
  return (size_t)kCW * kCPitch * 8 + 8 * 8 + 32 * 16 * (sizeof(typename Ar<T>::Box) + sizeof(T)) + 256;
}

}  // namespace mxr

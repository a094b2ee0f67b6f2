// Block-wide selection and sorting primitives (one CTA per problem, hand-written: no CUB).
#pragma once
#include "common.cuh"

namespace mxr {

// (key, payload) pairs ordered lexicographically, DESCENDING.
__device__ __forceinline__ bool pair_greater(u64 ka, u32 pa, u64 kb, u32 pb) {
  return ka > kb || (ka == kb && pa > pb);
}

// In-place bitonic sort of n_pow2 pairs into descending order.  `keys`/`pl` may live in
// shared or global memory (single CTA, so __syncthreads orders either).
static __device__ void bitonic_sort_desc(u64* keys, u32* pl, int n_pow2) {
  const int half = n_pow2 >> 1;
  for (int size = 2; size <= n_pow2; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int t = threadIdx.x; t < half; t += blockDim.x) {
        const int i = 2 * t - (t & (stride - 1));
        const int l = i + stride;
        const u64 ki = keys[i], kl = keys[l];
        const u32 pi = pl[i], pq = pl[l];
        const bool desc_block = (i & size) == 0;
        const bool l_greater = pair_greater(kl, pq, ki, pi);
        if (l_greater == desc_block) {
          keys[i] = kl;
          keys[l] = ki;
          pl[i] = pq;
          pl[l] = pi;
        }
      }
      __syncthreads();
    }
  }
}

// Same for bare 64-bit keys that are already distinct (proposal path: score key << 32 | ~row).
static __device__ void bitonic_sort_desc_u64(u64* keys, int n_pow2) {
  const int half = n_pow2 >> 1;
  for (int size = 2; size <= n_pow2; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int t = threadIdx.x; t < half; t += blockDim.x) {
        const int i = 2 * t - (t & (stride - 1));
        const int l = i + stride;
        const u64 ki = keys[i], kl = keys[l];
        const bool desc_block = (i & size) == 0;
        if ((kl > ki) == desc_block) {
          keys[i] = kl;
          keys[l] = ki;
        }
      }
      __syncthreads();
    }
  }
}

// Per-warp private 256-bin histograms (no atomics: lanes with the same digit are grouped
// with match.any and the group's leader adds the group's size).
__device__ __forceinline__ void warp_hist_add(u32* warp_hist, bool valid, u32 digit) {
  const u32 active = __ballot_sync(0xffffffffu, valid);
  if (valid) {
    const u32 peers = __match_any_sync(active, digit);
    const int leader = __ffs(peers) - 1;
    if ((int)(threadIdx.x & 31) == leader) warp_hist[digit] += __popc(peers);
  }
}

// Radix select on 64-bit composites: finds the threshold T such that exactly `k` valid
// composites are >= T (composites are distinct).  The items of the CTA are enumerated as
// (i, q = threadIdx.x + i*blockDim.x), i < n_iter; `compose(i, q, &c)` returns false for invalid
// / out-of-range entries.  ITEMS > 0 makes the loop a compile-time unrolled one (keys held in
// registers by the caller), ITEMS == 0 iterates n_iter times.  Needs blockDim.x/32 * 256 u32 of
// shared scratch (`hist`) and 8 u32 of shared state (`st`).  All threads must call; returns T
// to every thread.
template <int ITEMS, typename Compose>
__device__ __forceinline__ u64 block_radix_select(int n_iter, int k, u32* hist, u32* st, Compose compose) {
  const int nwarps = blockDim.x >> 5;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  u64 prefix = 0;
  int k_rem = k;
  for (int pass = 0; pass < 8; ++pass) {
    const int shift = 56 - 8 * pass;
    for (int i = threadIdx.x; i < nwarps * 256; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    u32* wh = hist + warp * 256;
    if (ITEMS > 0) {
#pragma unroll
      for (int i = 0; i < ITEMS; ++i) {
        u64 c = 0;
        bool valid = compose(i, (int)(threadIdx.x + i * blockDim.x), &c);
        if (valid && pass > 0) valid = ((c ^ prefix) >> (shift + 8)) == 0;
        warp_hist_add(wh, valid, (u32)(c >> shift) & 255u);
      }
    } else {
      for (int i = 0; i < n_iter; ++i) {
        u64 c = 0;
        bool valid = compose(i, (int)(threadIdx.x + i * blockDim.x), &c);
        if (valid && pass > 0) valid = ((c ^ prefix) >> (shift + 8)) == 0;
        warp_hist_add(wh, valid, (u32)(c >> shift) & 255u);
      }
    }
    __syncthreads();
    // reduce the per-warp histograms into hist[0..255]
    for (int d = threadIdx.x; d < 256; d += blockDim.x) {
      u32 s = 0;
      for (int w = 0; w < nwarps; ++w) s += hist[w * 256 + d];
      hist[d] = s;
    }
    __syncthreads();
    if (warp == 0) {
      // lane l owns digits 255-8l .. 248-8l (descending); inclusive scan over lanes
      u32 mine = 0;
      for (int j = 0; j < 8; ++j) mine += hist[255 - 8 * lane - j];
      u32 incl = mine;
      for (int o = 1; o < 32; o <<= 1) {
        u32 v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      const u32 excl = incl - mine;
      const bool here = (excl < (u32)k_rem) && ((u32)k_rem <= incl);
      if (here) {
        u32 cum = excl;
        for (int j = 0; j < 8; ++j) {
          const int d = 255 - 8 * lane - j;
          const u32 cnt = hist[d];
          if (cum + cnt >= (u32)k_rem) {
            st[0] = (u32)d;
            st[1] = (u32)k_rem - cum;  // still needed from inside bucket d
            st[2] = cnt;
            break;
          }
          cum += cnt;
        }
      }
    }
    __syncthreads();
    const u32 d = st[0];
    k_rem = (int)st[1];
    const u32 cnt = st[2];
    prefix |= (u64)d << shift;
    __syncthreads();
    if ((int)cnt == k_rem) break;  // the whole bucket is selected: lower bits of T stay 0
  }
  return prefix;
}

}  // namespace mxr

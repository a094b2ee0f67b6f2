// Fused RPN proposal kernel: ONE launch, one thread-block cluster (8 CTAs x 1024 threads) per image.
//
// rcnn/rpn/proposal.py:39-149 end to end inside the cluster, phases separated by cluster barriers
// (~0.2 us) instead of kernel boundaries, with distributed shared memory as the exchange medium:
//   P1 decode + clip + min-size filter (each CTA an eighth of the anchors, keys stay in registers)
//   P2 radix select of the head of the candidate order: per-CTA digit histograms are summed into
//      the leader CTA's shared memory with DSMEM atomics, every CTA reads the sum back
//   P3 the selected composites are written straight into the leader's sort buffer (DSMEM stores)
//   P4 the leader sorts the head (<= 4096 boxes)          P5 all CTAs re-decode the head boxes
//   P6 per 1024-box stripe: the other CTAs evaluate the IoU bit tiles and store them directly into
//      the leader's shared bit matrix (DSMEM), the leader walks the stripe greedily
//   P7 the leader writes rois / scores / keep lists.
// Greedy NMS walks the candidate list in order and usually fills its post-NMS quota long before the
// tail of the list matters, so only the HEAD of the order (<= 4096 boxes) is selected and sorted up
// front.  If the head does not settle the image (quota not reached, or a run of equal scores is cut
// by the head boundary) the same cluster selects and sorts the full pre-NMS list and carries on;
// the kept list and every stripe already walked stay valid because the head is a prefix of it.
#pragma once
#include <cooperative_groups.h>

#include "cluster_nms.cuh"
#include "nms_coop.cuh"
#include "proposal_kernels.cuh"

namespace mxr {

static const int kClusterSize = 8;      // portable cluster size: the default
static const int kClusterSizeMax = 16;  // opt-in size, used when every image's cluster can be resident at once
static const int kClusterSortMax = 16384;  // largest list the leader sorts in shared memory (128 KiB)

struct ClusterArgs {
  ProposalGeom pg;
  AnchorTable at;
  const float* cls_prob;
  const float* bbox_pred;
  const float* im_info;
  float min_size;
  int pre_nms;
  int head;          // sorted prefix handled here (multiple of 64, <= 4096)
  int max_keep;      // > 0: post_nms_top_n
  double thr;
  int sort_cap;      // u64 slots of the leader's sort buffer (power of two >= the largest list sorted)
  NmsArrays<double> a;
  NmsOutputs o;
  Ctrl* ctrl;
  u64* dbg;          // optional [b][16] globaltimer stamps at the phase boundaries (leader CTA), or null
};

__device__ __forceinline__ u64 global_ns() {
  u64 t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define MXR_STAMP(i)                                                              \
  do {                                                                            \
    if (g.dbg && rank == 0 && tid == 0) g.dbg[(size_t)b * 16 + (i)] = global_ns(); \
  } while (0)

__device__ __forceinline__ void remote_or64(u64* p, u64 v) {
  u32* q = reinterpret_cast<u32*>(p);
  const u32 lo = (u32)v, hi = (u32)(v >> 32);
  if (lo) atomicOr(q, lo);
  if (hi) atomicOr(q + 1, hi);
}

// Head selection bins: bin(s) = clamp(int(s * kHeadBins), 0, kHeadBins - 1) is monotone in the score, so "bin >= t" is
// exactly "score >= t / kHeadBins" -- a cut between two distinct scores, never inside a run of equal ones.
static const int kHeadBins = 1024;
__device__ __forceinline__ int head_bin_of_key(u32 key) {
  const u32 bits = (key & 0x80000000u) ? (key & 0x7fffffffu) : ~key;  // inverse of score_key32
  const int v = (int)(__uint_as_float(bits) * (float)kHeadBins);        // saturating conversion
  return v < 0 ? 0 : (v > kHeadBins - 1 ? kHeadBins - 1 : v);
}

// grid B * csize CTAs, cluster (csize,1,1) with csize = 8 or 16, block 1024.
// ITEMS: keys per thread; the image's A*hc*wc anchors must fit csize*ITEMS*1024.
template <int ITEMS>
__global__ void __launch_bounds__(1024, 1) proposal_cluster_kernel(ClusterArgs g) {
  namespace cgx = cooperative_groups;
  extern __shared__ __align__(16) unsigned char dyn_smem[];
  __shared__ u32 s_hist[256];
  __shared__ u32 s_bins[kHeadBins];      // this CTA's counts of its valid scores per linear score bin (head selection)
  __shared__ u32 s_pack[kHeadBins / 2];  // the same as pairs of 16-bit counts: what the other CTAs read
  __shared__ u32 s_cursor[kHeadBins];    // fill level of each bin's bucket while the head is ordered
  __shared__ u32 s_ghist[12][256];  // leader: cluster-wide digit histograms: head select 0..3, full select 4..10
  __shared__ int s_sel[kClusterSizeMax];  // leader: selected count of each CTA
  __shared__ __align__(8) int s_status[2];  // kept so far, quota reached: written by the leader into every CTA
  __shared__ u32 s_st[4];
  __shared__ int s_local[4];
  __shared__ int s_pub[2];  // this CTA's valid / NaN counts of the decode phase, read by the other CTAs
  cgx::cluster_group cluster = cgx::this_cluster();
  const int rank = (int)cluster.block_rank();
  const int csize = (int)cluster.num_blocks();
  const int b = blockIdx.x / csize;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const ProposalGeom& pg = g.pg;
  const DecodeGeom& dg = pg.g;
  // leader-resident structures, addressed through distributed shared memory
  u64* sortbuf = reinterpret_cast<u64*>(dyn_smem);                           // leader: [sort_cap] sort buffer
  u64* sM = sortbuf + g.sort_cap;                                            // leader: [256][5] bit rows of one stripe
  u64* s_removed = sM + kCW * kCPitch;                                       // leader: [4]
  double4* t_box = reinterpret_cast<double4*>(s_removed + 8) + warp * 16;    // every CTA: per-warp column staging
  double* t_area = reinterpret_cast<double*>(reinterpret_cast<double4*>(s_removed + 8) + 32 * 16) + warp * 16;
  u64* L_sort = cluster.map_shared_rank(sortbuf, 0);
  u64* L_sM = cluster.map_shared_rank(sM, 0);
  u64* L_removed = cluster.map_shared_rank(s_removed, 0);
  u32* L_ghist = cluster.map_shared_rank(&s_ghist[0][0], 0);
  int* L_sel = cluster.map_shared_rank(s_sel, 0);
  if (rank == 0) {
    for (int i = tid; i < 12 * 256; i += blockDim.x) (&s_ghist[0][0])[i] = 0;
    if (tid < kCSW) s_removed[tid] = 0ull;
    if (tid < 2) s_status[tid] = 0;
  }
  if (tid < 4) s_local[tid] = 0;
  // programmatic dependent launch (proposal.cu): let the next kernel of the stream be scheduled, then wait until
  // everything launched before this kernel has completed and its writes are visible (no-ops for a plain launch)
  asm volatile("griddepcontrol.launch_dependents;");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  // No start-up barrier: nothing touches another CTA's shared memory before the barrier that ends the decode phase.
  for (int i = tid; i < kHeadBins; i += blockDim.x) s_bins[i] = 0;
  __syncthreads();
  MXR_STAMP(0);
  MXR_STAMP(1);

  // ---- P1: decode + clip + min-size filter (proposal.py:77-120); keys stay in registers ----
  const int hwc = dg.hc * dg.wc;
  const size_t hw = (size_t)dg.H * dg.W;
  const float* fg = g.cls_prob + (size_t)b * pg.cls_img_stride + pg.cls_fg_offset;
  const float* deltas = g.bbox_pred + (size_t)b * 4 * dg.A * hw;
  const float im_h = __ldg(g.im_info + b * 3), im_w = __ldg(g.im_info + b * 3 + 1), im_s = __ldg(g.im_info + b * 3 + 2);
  const int chunk = (pg.n_crop + csize - 1) / csize;
  const int q_begin = rank * chunk;
  const int q_end = min(q_begin + chunk, pg.n_crop);
  const float rcp_hwc = 1.0f / (float)hwc, rcp_wc = 1.0f / (float)dg.wc, rcp_A = 1.0f / (float)dg.A;
  u32 rk[ITEMS];
  {
    int nvalid = 0, nnan = 0;
    // Groups of four anchors per thread: first ALL the loads of the group (five floats per anchor, independent of each
    // other), then the arithmetic -- otherwise the compiler walks the anchors one at a time and a thread pays one DRAM
    // round trip per anchor.
    constexpr int kGroup = ITEMS < 4 ? ITEMS : 4;
#pragma unroll
    for (int i0 = 0; i0 < ITEMS; i0 += kGroup) {
      int ia[kGroup], ih[kGroup], iw[kGroup];
      float dx[kGroup], dy[kGroup], dw[kGroup], dh[kGroup], sc[kGroup];
#pragma unroll
      for (int j = 0; j < kGroup; ++j) {
        const int q = q_begin + tid + (i0 + j) * 1024;
        ia[j] = -1;
        ih[j] = iw[j] = 0;
        dx[j] = dy[j] = dw[j] = dh[j] = sc[j] = 0.0f;
        if (q < q_end) {
          const int a = fast_div(q, hwc, rcp_hwc), r = q - a * hwc;
          const int h = fast_div(r, dg.wc, rcp_wc), w = r - h * dg.wc;
          ia[j] = a;
          ih[j] = h;
          iw[j] = w;
          const float* d = deltas + ((size_t)a * 4) * hw + (size_t)h * dg.W + w;
          dx[j] = __ldg(d);
          dy[j] = __ldg(d + hw);
          dw[j] = __ldg(d + 2 * hw);
          dh[j] = __ldg(d + 3 * hw);
          if (!pg.cropped) {
            sc[j] = __ldg(fg + (size_t)a * hw + (size_t)h * dg.W + w);
          } else {  // quirk Q1: scores are read at the flat (H,W,A) index of the UNCROPPED map
            const int p = r * dg.A + a;
            const int a2 = p % dg.A, w2 = (p / dg.A) % dg.W, h2 = p / (dg.A * dg.W);
            sc[j] = __ldg(fg + (size_t)a2 * hw + (size_t)h2 * dg.W + w2);
          }
        }
      }
#pragma unroll
      for (int j = 0; j < kGroup; ++j) {
        u32 key = 0;
        bool nan = false;
        if (ia[j] >= 0) {
          double bx[4];
          decode_clip_values(g.at, dg, dx[j], dy[j], dw[j], dh[j], ih[j], iw[j], ia[j], im_h, im_w, bx);
          const bool valid = passes_min_size(bx, g.min_size, im_s);
          // the clip drops NaNs, so the reference's NaN check (proposal.py:58-62) looks at the raw values
          nan = (sc[j] != sc[j]) || (dx[j] != dx[j]) || (dy[j] != dy[j]) || (dw[j] != dw[j]) || (dh[j] != dh[j]);
          key = valid ? score_key32(sc[j]) : 0u;
        }
        rk[i0 + j] = key;
        if (key != 0u) atomicAdd(&s_bins[head_bin_of_key(key)], 1u);  // head selection table, see cluster_bin_select
        nvalid += __popc(__ballot_sync(0xffffffffu, key != 0u));
        nnan += __popc(__ballot_sync(0xffffffffu, nan));
      }
    }
    if (pg.cropped) {  // proposal.py:58-62 checks the whole arrays, cropped or not
      const int fchunk = (pg.n_full + csize - 1) / csize;
      for (int q = rank * fchunk + tid; q < min((rank + 1) * fchunk, pg.n_full); q += 1024) {
        const float v0 = __ldg(fg + q), v1 = __ldg(deltas + q), v2 = __ldg(deltas + q + pg.n_full);
        const float v3 = __ldg(deltas + q + 2 * (size_t)pg.n_full), v4 = __ldg(deltas + q + 3 * (size_t)pg.n_full);
        if ((v0 != v0) || (v1 != v1) || (v2 != v2) || (v3 != v3) || (v4 != v4)) nnan = 1;
      }
      nnan = __popc(__ballot_sync(0xffffffffu, nnan != 0));
    }
    if (lane == 0) {
      if (nvalid) atomicAdd(&s_local[0], nvalid);
      if (nnan) atomicAdd(&s_local[1], nnan);
    }
    __syncthreads();
    if (tid < 2) s_pub[tid] = s_local[tid];
    if (tid < kHeadBins / 2) s_pack[tid] = s_bins[2 * tid] | (s_bins[2 * tid + 1] << 16);  // bins 2t (low half), 2t+1
  }
  cluster.sync();  // ONE barrier ends the decode phase: counts and score-bin tables of every CTA are published
  if (warp == 0) {  // every CTA sums the published counts itself (one DSMEM round trip, one lane per CTA)
    const int* pub = cluster.map_shared_rank(s_pub, lane < csize ? lane : 0);
    const int nv = __reduce_add_sync(0xffffffffu, lane < csize ? pub[0] : 0);
    const int nn = __reduce_add_sync(0xffffffffu, lane < csize ? pub[1] : 0);
    if (lane == 0) {
      s_local[2] = nv;
      s_local[3] = nn;
    }
  }
  __syncthreads();
  const int n_valid = s_local[2];
  const int nan_flag = s_local[3] ? 1 : 0;
  const int k = (g.pre_nms > 0) ? min(g.pre_nms, n_valid) : n_valid;  // proposal.py:124-125
  const int k1 = min(g.head, k);
  MXR_STAMP(2);

  auto p_of_item = [&](int i) -> u32 { return (u32)q_to_p(q_begin + tid + i * 1024, hwc, dg.A); };

  // ---- radix select over the cluster: threshold T such that exactly k_sel valid composites
  //      (score key << 32 | row index) are >= T.  The four key bytes come first; the row-index bytes are
  //      only looked at when a run of equal scores straddles the cut and use_p is set.  Per-CTA digit
  //      histograms are summed in the leader's shared memory with DSMEM atomics and read back by all.
  //      Returns false when the cut is not exact (only possible without use_p). ----
  auto cluster_select = [&](int k_sel, int hist_base, bool use_p, u64* T_out) -> bool {
    u64 prefix = 0;
    int k_rem = k_sel;
    u32 cnt = 0;
    const int npass = use_p ? 7 : 4;
    for (int pass = 0; pass < npass; ++pass) {
      const int shift = pass < 4 ? 56 - 8 * pass : 16 - 8 * (pass - 4);
      if (tid < 256) s_hist[tid] = 0;
      __syncthreads();
#pragma unroll
      for (int i = 0; i < ITEMS; ++i) {
        const u32 key = rk[i];
        bool valid = key != 0u;
        u64 c = (u64)key << 32;
        if (pass >= 4) c |= p_of_item(i);
        if (valid && pass > 0) valid = ((c ^ prefix) >> (shift + 8)) == 0ull;
        const u32 digit = (u32)(c >> shift) & 255u;
        const u32 active = __ballot_sync(0xffffffffu, valid);
        if (valid) {
          const u32 peers = __match_any_sync(active, digit);
          if (lane == __ffs(peers) - 1) atomicAdd(&s_hist[digit], (u32)__popc(peers));
        }
      }
      __syncthreads();
      if (tid < 256 && s_hist[tid]) atomicAdd(L_ghist + (hist_base + pass) * 256 + tid, s_hist[tid]);
      cluster.sync();
      if (tid < 256) s_hist[tid] = L_ghist[(hist_base + pass) * 256 + tid];
      __syncthreads();
      if (warp == 0) {
        u32 mine = 0;
        for (int j = 0; j < 8; ++j) mine += s_hist[255 - 8 * lane - j];
        u32 incl = mine;
        for (int o = 1; o < 32; o <<= 1) {
          const u32 v = __shfl_up_sync(0xffffffffu, incl, o);
          if (lane >= o) incl += v;
        }
        const u32 excl = incl - mine;
        if ((excl < (u32)k_rem) && ((u32)k_rem <= incl)) {
          u32 cum = excl;
          for (int j = 0; j < 8; ++j) {
            const int d = 255 - 8 * lane - j;
            const u32 c = s_hist[d];
            if (cum + c >= (u32)k_rem) {
              s_st[0] = (u32)d;
              s_st[1] = (u32)k_rem - cum;
              s_st[2] = c;
              break;
            }
            cum += c;
          }
        }
      }
      __syncthreads();
      prefix |= (u64)s_st[0] << shift;
      k_rem = (int)s_st[1];
      cnt = s_st[2];
      __syncthreads();
      if ((int)cnt == k_rem) break;
    }
    *T_out = prefix;
    return (int)cnt == k_rem;
  };

  // ---- one-pass head selection: every CTA has counted its valid scores per linear bin in its own shared memory
  //      (decode loop); every CTA sums all the cluster's tables (DSMEM reads, no remote atomics) and scans them from the top: the
  //      head is everything in the highest bins that together hold at most `cap` candidates.  Returns false when
  //      that is fewer than 3/4 of cap (scores crowded into one bin): the caller falls back to the exact radix select. ----
  auto cluster_bin_select = [&](int cap, u64* T_out, int* m_out) -> bool {
    // the per-CTA tables were filled during the decode loop and published (16-bit pairs) before its closing barrier
    // this thread's two bins, descending: hi = kHeadBins-1 - 2*tid, lo = hi - 1   (threads below kHeadBins / 2).
    // DSMEM moves only ~10-20 B per clock per SM, so the tables travel as 16-bit counts (a CTA holds < 65536 keys).
    const bool owner = tid < kHeadBins / 2;
    u32 c_hi = 0, c_lo = 0;
    if (owner) {
      for (int r0 = 0; r0 < csize; r0 += 8) {  // eight remote loads in flight; start at the own table and go round,
        u32 v[8];                              // so that no CTA is read by all the others at once
#pragma unroll
        for (int r = 0; r < 8; ++r) {
          const int o = rank + r0 + r, src = o < csize ? o : o - csize;
          v[r] = 0u;
          if (r0 + r < csize) v[r] = cluster.map_shared_rank(s_pack, src)[kHeadBins / 2 - 1 - tid];
        }
#pragma unroll
        for (int r = 0; r < 8; ++r) {
          c_lo += v[r] & 0xffffu;
          c_hi += v[r] >> 16;
        }
      }
    }
    u32 incl = c_hi + c_lo;
    for (int o = 1; o < 32; o <<= 1) {
      const u32 v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    if (lane == 31) s_hist[warp] = incl;
    __syncthreads();
    u32 wincl = s_hist[lane];
    for (int o = 1; o < 32; o <<= 1) {
      const u32 v = __shfl_up_sync(0xffffffffu, wincl, o);
      if (lane >= o) wincl += v;
    }
    const u32 before = __shfl_sync(0xffffffffu, wincl, (warp + 31) & 31);
    const u32 S_lo = (warp > 0 ? before : 0u) + incl, S_hi = S_lo - c_lo;  // candidates in this bin and above
    if (owner) {  // the local counts are spent: the table now holds, per bin, how many candidates sit in HIGHER bins
      s_bins[kHeadBins - 1 - 2 * tid] = S_hi - c_hi;
      s_bins[kHeadBins - 2 - 2 * tid] = S_lo - c_lo;
    }
    const int q = __syncthreads_count(owner && S_hi <= (u32)cap) + __syncthreads_count(owner && S_lo <= (u32)cap);
    if (q > 0) {
      if (2 * tid == q - 1) {
        s_st[0] = (u32)(kHeadBins - 1 - 2 * tid);
        s_st[1] = S_hi;
      }
      if (2 * tid + 1 == q - 1) {
        s_st[0] = (u32)(kHeadBins - 2 - 2 * tid);
        s_st[1] = S_lo;
      }
    }
    __syncthreads();
    const int bin_t = (int)s_st[0], m = (int)s_st[1];
    __syncthreads();
    if (q == 0 || 4 * m < 3 * cap) return false;
    *T_out = bin_t == 0 ? (1ull << 32) : ((u64)score_key32((float)bin_t * (1.0f / (float)kHeadBins)) << 32);
    *m_out = m;
    return true;
  };

  // ---- head of at most 1024 candidates: order it WITHOUT a sort.  Every CTA compacts its own selected composites
  //      locally, pulls everybody's into its shared memory (one DSMEM load per thread), and counts for each of its
  //      own composites how many are greater: that count is the composite's position in (score desc, row index ASC)
  //      order -- nms.py:19 re-sorts the already descending list, which reverses runs of equal scores.  The owner
  //      then decodes the box straight into that position.  ~1M comparisons spread over the cluster's threads
  //      instead of a 55-stage bitonic network on one SM followed by a separate re-decode pass. ----
  auto rank_and_redecode = [&](int k_sel, u64 T, int stamp, bool binned) {
    const bool need_p = (u32)T != 0u;  // the threshold cuts inside a run of equal scores
    auto selected = [&](int i) -> bool {
      const u32 key = rk[i];
      if (key == 0u) return false;
      const u64 c = ((u64)key << 32) | ((need_p && key == (u32)(T >> 32)) ? p_of_item(i) : 0xffffffffu);
      return c >= T;
    };
    u64* loc = sM;       // this CTA's selected composites, unordered (the bit-row area is free until the NMS walk)
    u64* all = sortbuf;  // this CTA's copy of all k_sel composites
    __syncthreads();
    if (tid == 0) s_local[0] = 0;
    __syncthreads();
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const bool sel = selected(i);
      const u32 m = __ballot_sync(0xffffffffu, sel);
      int base = 0;
      if (lane == 0 && m) base = atomicAdd(&s_local[0], __popc(m));
      base = __shfl_sync(0xffffffffu, base, 0);
      if (sel) loc[base + __popc(m & ((1u << lane) - 1u))] = ((u64)rk[i] << 32) | (u32)(~p_of_item(i));
    }
    cluster.sync();  // every CTA's list and count are in place
    MXR_STAMP(stamp);
    if (warp == 0) {  // offsets of the CTAs' lists inside `all`
      const int cnt = lane < csize ? cluster.map_shared_rank(s_local, lane)[0] : 0;
      int incl = cnt;
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      if (lane < kClusterSizeMax) s_sel[lane] = incl - cnt;
    }
    if (binned)
      for (int i = tid; i < kHeadBins; i += blockDim.x) s_cursor[i] = 0;
    __syncthreads();
    for (int j = tid; j < k_sel; j += blockDim.x) {  // one remote load per thread
      int r = 0;
      while (r + 1 < csize && s_sel[r + 1] <= j) ++r;
      const u64 c = cluster.map_shared_rank(loc, r)[j - s_sel[r]];
      int slot = j;
      if (binned) {  // bucket by score bin: s_bins[bin] candidates sit in higher bins (cluster_bin_select)
        const int bin = head_bin_of_key((u32)(c >> 32));
        slot = (int)(s_bins[bin] + atomicAdd(&s_cursor[bin], 1u));
      }
      all[slot] = c;
    }
    __syncthreads();
    const int n_loc = s_local[0];
    int G = 32;  // threads per own composite
    while (G > 1 && n_loc * G > (int)blockDim.x) G >>= 1;
    for (int i0 = 0; i0 < n_loc; i0 += blockDim.x / G) {
      const int i = i0 + tid / G, sub = tid & (G - 1);
      const u64 mine = i < n_loc ? loc[i] : ~0ull;
      int pos = 0;
      if (i < n_loc) {
        int lo = 0, n_cmp = k_sel;  // only the own bucket has to be looked at when the copy is bucketed
        if (binned) {
          const int bin = head_bin_of_key((u32)(mine >> 32));
          lo = (int)s_bins[bin];
          n_cmp = (int)s_cursor[bin];
        }
        for (int j = sub; j < n_cmp; j += G) pos += all[lo + j] > mine ? 1 : 0;
        if (sub == 0) pos += lo;
      }
      for (int o = 1; o < G; o <<= 1) pos += __shfl_xor_sync(0xffffffffu, pos, o);
      if (sub == 0 && i < n_loc) {
        const int p = (int)(~(u32)mine);
        const int r = fast_div(p, dg.A, rcp_A), a = p - r * dg.A;
        const int h = fast_div(r, dg.wc, rcp_wc), w = r - h * dg.wc;
        double bx[4];
        decode_clip(g.at, dg, deltas, h, w, a, im_h, im_w, bx);
        const int a2 = a, w2 = r % dg.W, h2 = r / dg.W;  // the score sits at the flat (H,W,A) index of the FULL map (quirk Q1)
        const float s = __ldg(fg + (size_t)a2 * hw + (size_t)h2 * dg.W + w2);
        const size_t o = (size_t)b * g.a.cap + pos;
        const_cast<double4*>(g.a.box)[o] = make_double4(bx[0], bx[1], bx[2], bx[3]);
        const_cast<double*>(g.a.area)[o] = box_area<double>(bx[0], bx[1], bx[2], bx[3]);
        const_cast<float*>(g.a.score)[o] = s;
        const_cast<int*>(g.a.orig)[o] = p;
      }
    }
    MXR_STAMP(stamp + 1);
    cluster.sync();  // the boxes are visible to every CTA of the cluster; nobody reads `loc` any more
  };

  // ---- compact the k_sel selected composites into the leader's sort buffer (DSMEM stores), let the
  //      leader order them by (score desc, row index ASC) -- nms.py:19 re-sorts the already descending list,
  //      which reverses runs of equal scores -- and re-decode them, spread over all CTAs, into processing order ----
  auto sort_and_redecode = [&](int k_sel, u64 T, int stamp) {
    const int n_pow2 = max(2, (int)(1u << (32 - __clz(k_sel - 1))));
    const bool need_p = (u32)T != 0u;  // the threshold cuts inside a run of equal scores
    auto selected = [&](int i) -> bool {
      const u32 key = rk[i];
      if (key == 0u) return false;
      const u64 c = ((u64)key << 32) | ((need_p && key == (u32)(T >> 32)) ? p_of_item(i) : 0xffffffffu);
      return c >= T;
    };
    {
      int mine = 0;
#pragma unroll
      for (int i = 0; i < ITEMS; ++i) mine += __popc(__ballot_sync(0xffffffffu, selected(i)));
      __syncthreads();
      if (tid < 4) s_local[tid] = 0;
      __syncthreads();
      if (lane == 0 && mine) atomicAdd(&s_local[0], mine);
      __syncthreads();
      if (tid == 0) L_sel[rank] = s_local[0];
    }
    cluster.sync();
    if (warp == 0) {  // one DSMEM round trip: the lower ranks' counts are fetched by separate lanes
      const int off = __reduce_add_sync(0xffffffffu, lane < rank ? L_sel[lane] : 0);
      if (lane == 0) {
        s_local[1] = off;   // base of this CTA inside the sort buffer
        s_local[0] = 0;     // running position inside this CTA
      }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const bool sel = selected(i);
      const u32 m = __ballot_sync(0xffffffffu, sel);
      int base = 0;
      if (lane == 0 && m) base = atomicAdd(&s_local[0], __popc(m));
      base = __shfl_sync(0xffffffffu, base, 0);
      if (sel) L_sort[s_local[1] + base + __popc(m & ((1u << lane) - 1u))] = ((u64)rk[i] << 32) | (u32)(~p_of_item(i));
    }
    if (rank == 0)
      for (int i = k_sel + tid; i < max(n_pow2, 1024); i += blockDim.x)
        if (i < g.sort_cap) sortbuf[i] = 0ull;
    cluster.sync();
    MXR_STAMP(stamp);
    if (rank == 0) {
      if (n_pow2 <= 1024) bitonic_sort_1024_x4(sortbuf, n_pow2);
      else bitonic_sort_desc_u64(sortbuf, n_pow2);
    }
    cluster.sync();
    MXR_STAMP(stamp + 1);
    const int per = (k_sel + csize - 1) / csize;
    for (int i = rank * per + tid; i < min((rank + 1) * per, k_sel); i += blockDim.x) {
      const int p = (int)(~(u32)L_sort[i]);
      const int a = p % dg.A, r = p / dg.A;
      const int h = r / dg.wc, w = r - h * dg.wc;
      double bx[4];
      decode_clip(g.at, dg, deltas, h, w, a, im_h, im_w, bx);
      const int a2 = p % dg.A, w2 = (p / dg.A) % dg.W, h2 = p / (dg.A * dg.W);
      const float s = __ldg(fg + (size_t)a2 * hw + (size_t)h2 * dg.W + w2);
      const size_t o = (size_t)b * g.a.cap + i;
      const_cast<double4*>(g.a.box)[o] = make_double4(bx[0], bx[1], bx[2], bx[3]);
      const_cast<double*>(g.a.area)[o] = box_area<double>(bx[0], bx[1], bx[2], bx[3]);
      const_cast<float*>(g.a.score)[o] = s;
      const_cast<int*>(g.a.orig)[o] = p;
    }
    cluster.sync();  // release/acquire at cluster scope: the boxes are visible to every CTA of the cluster
  };

  // ---- greedy NMS over the sorted positions, in stripes of 256 boxes (cluster_nms.cuh) ----
  StripeWalker<double> wk;
  wk.box = g.a.box + (size_t)b * g.a.cap;
  wk.area = g.a.area + (size_t)b * g.a.cap;
  wk.kept_slot = g.a.kept_slot + (size_t)b * g.a.keep_cap;
  wk.sM = sM;
  wk.s_removed = s_removed;
  wk.s_status = s_status;
  wk.L_sM = L_sM;
  wk.L_removed = L_removed;
  wk.t_box = t_box;
  wk.t_area = t_area;
  wk.thr = g.thr;
  wk.max_keep = g.max_keep;
  wk.nk = 0;
  wk.fin = 0;
  wk.stripes_done = 0;

  // ---- head: select + sort + walk about the first k1 candidates ----
  bool head_ok = false;
  int m_head = 0;  // candidates in the sorted head
  bool head_binned = false;  // the head was cut at a score-bin boundary
  if (k1 > 0) {
    u64 T = 1ull << 32;  // every valid composite
    if (k1 == n_valid) {
      head_ok = true;
      m_head = k1;
    } else if (cluster_bin_select(k1, &T, &m_head)) {
      head_ok = true;
      head_binned = true;
    } else {
      head_ok = cluster_select(k1, 0, false, &T);
      m_head = k1;
    }
    MXR_STAMP(3);
    if (head_ok) {
      if (m_head <= 1024) rank_and_redecode(m_head, T, 4, head_binned);
      else sort_and_redecode(m_head, T, 4);
      MXR_STAMP(6);
      // while more candidates follow the head only whole stripes of it are walked, so that the full list (of which
      // the head is a prefix) can take over at a stripe boundary
      wk.walk(cluster, 0, m_head < k ? (m_head / kCW) * kCW : m_head);
    }
  }
  // ---- full list, only if the head did not settle the image ----
  int full_path = 0;
  if (k > 0 && !wk.fin && (!head_ok || m_head < k)) {
    full_path = 1;
    u64 T = 1ull << 32;
    if (k < n_valid) cluster_select(k, 4, true, &T);
    sort_and_redecode(k, T, 10);
    if (!head_ok) {
      wk.nk = 0;            // nothing was walked (a run of equal scores is cut by the head boundary)
      wk.stripes_done = 0;
    }
    wk.walk(cluster, wk.stripes_done, k);
  }
  MXR_STAMP(7);
  if (!head_ok && g.dbg && rank == 0 && tid == 0)
    for (int i = 3; i <= 6; ++i) g.dbg[(size_t)b * 16 + i] = g.dbg[(size_t)b * 16 + 2];  // head phases did not run
  // ---- results (proposal.py:140-149) ----
  if (rank == 0) {
    Ctrl* cb = g.ctrl + b;
    if (tid == 0) {
      cb->n_valid = n_valid;
      cb->nan_flag = nan_flag;
      cb->n = k;
      cb->nkept = wk.nk;
      cb->done = 1;
      cb->pad[0] = 0;
    }
    __syncthreads();
    coop_write_outputs<double>(g.a, g.o, cb, b, wk.nk, k);
  }
  // Nobody may leave while its shared memory can still be addressed by another CTA.  Every stripe of the walk and every
  // (re)ordering step ends with a cluster barrier AFTER its last remote access, so a closing barrier is only needed
  // when neither ran (no valid candidate at all: the last remote access was the count pull after the decode phase).
  if (wk.stripes_done == 0 && !head_ok && !full_path) cluster.sync();
  MXR_STAMP(8);
  if (g.dbg && rank == 0 && tid == 0) {
    g.dbg[(size_t)b * 16 + 9] = (u64)wk.stripes_done;
    g.dbg[(size_t)b * 16 + 14] = (u64)full_path;
  }
}

static inline size_t cluster_smem_bytes(int sort_cap) { return (size_t)sort_cap * 8 + cluster_nms_smem_bytes<double>(); }

}  // namespace mxr

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

#include "nms_coop.cuh"
#include "proposal_kernels.cuh"

namespace mxr {

static const int kClusterSize = 8;
static const int kCW = 256;        // stripe width inside the cluster kernel (cluster barriers are cheap)
static const int kCSW = 4;         // words per stripe row
static const int kCPitch = 5;      // shared-memory row pitch in words
static const int kCTriTasks = 80;  // 10 upper-triangle tiles x 2 row halves x 4 column quarters
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

__device__ __forceinline__ double4 ldcg_box(const double4* p) {
  const double2 a = __ldcg(reinterpret_cast<const double2*>(p));
  const double2 b = __ldcg(reinterpret_cast<const double2*>(p) + 1);
  return make_double4(a.x, a.y, b.x, b.y);
}

__device__ __forceinline__ void remote_or64(u64* p, u64 v) {
  u32* q = reinterpret_cast<u32*>(p);
  const u32 lo = (u32)v, hi = (u32)(v >> 32);
  if (lo) atomicOr(q, lo);
  if (hi) atomicOr(q + 1, hi);
}

// Descending bitonic sort of n_pow2 <= 1024 distinct 64-bit keys, one key per thread (blockDim.x == 1024):
// strides below 32 are warp shuffles, only strides >= 32 go through shared memory (15 of the 55 stages).
__device__ __forceinline__ void bitonic_sort_1024_regs(u64* buf, int n_pow2) {
  const int t = threadIdx.x;
  u64 v = buf[t];  // entries beyond the real keys are zero-padded by the caller up to n_pow2; beyond that unused
  for (int size = 2; size <= n_pow2; size <<= 1) {
    const bool desc = (t & size) == 0;
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      u64 other;
      if (stride < 32) {
        other = __shfl_xor_sync(0xffffffffu, v, stride);
      } else {
        __syncthreads();
        buf[t] = v;
        __syncthreads();
        other = buf[t ^ stride];
      }
      const bool lower = (t & stride) == 0;
      const bool take_max = (lower == desc);
      const bool other_gt = other > v;
      v = (take_max == other_gt) ? other : v;
    }
  }
  __syncthreads();
  buf[t] = v;
  __syncthreads();
}

// 16 columns of one row: bit j set when column j is suppressed by the row box (nms.py:25-36).
__device__ __forceinline__ u32 tile_bits16(const double4 rb, double rarea, const double4* cols, const double* careas,
                                           int j_begin, int j_end, double thr, double thr_lo, double thr_hi) {
  u32 bits = 0;
  if (thr > 0.0) {
#pragma unroll 4
    for (int j = j_begin; j < j_end; ++j) {
      const double4 cb = cols[j];
      int r = iou_band_f64(rb.x, rb.y, rb.z, rb.w, rarea, cb.x, cb.y, cb.z, cb.w, careas[j], thr_lo, thr_hi);
      if (r == 2) r = iou_exceeds<double>(rb.x, rb.y, rb.z, rb.w, rarea, cb.x, cb.y, cb.z, cb.w, careas[j], thr) ? 1 : 0;
      bits |= (u32)r << j;
    }
  } else {
    for (int j = j_begin; j < j_end; ++j) {
      const double4 cb = cols[j];
      if (iou_exceeds<double>(rb.x, rb.y, rb.z, rb.w, rarea, cb.x, cb.y, cb.z, cb.w, careas[j], thr)) bits |= 1u << j;
    }
  }
  return bits;
}

// Greedy walk of one stripe whose bit rows already sit in shared memory (pitch kCoopPitch).
// Whole CTA.  Returns the new kept count; *fin_out != 0 when the quota was reached.
__device__ __forceinline__ int scan_rows_smem(const u64* sM, int pitch, u64* s_removed, int rows, int c0, int nk, int max_keep,
                                              int* kept_slot, u64* s_keptbits, int* s_finish, int* fin_out) {
  const int tid = threadIdx.x;
  const int nblk = (rows + 63) >> 6;
  int finished = 0;
  for (int blk = 0; blk < nblk; ++blk) {
    __syncthreads();
    if (tid < 32) {
      const u64 word0 = s_removed[blk];
      const int nvalid = min(64, rows - blk * 64);
      const u64 valid = nvalid == 64 ? ~0ull : ((1ull << nvalid) - 1ull);
      const u64 alive = ~word0 & valid;
      // Diagonal tile, stored as its LOWER triangle: bit j of row i (j < i) says "j suppresses i" (the IoU
      // predicate is symmetric bit for bit).  Greedy keep is the unique fixed point of
      //   keep_i = alive_i and no kept j < i suppresses i ;
      // iterating from keep = alive fixes at least one more leading position per round, and in practice
      // converges in a handful of rounds instead of a 64-step dependent chain.
      const u64 m0 = ((blk * 64 + tid < rows) ? sM[(blk * 64 + tid) * pitch + blk] : 0ull) & ((1ull << tid) - 1ull);
      const u64 m1 = ((blk * 64 + 32 + tid < rows) ? sM[(blk * 64 + 32 + tid) * pitch + blk] : 0ull) & ((1ull << (tid + 32)) - 1ull);
      const bool a0 = (alive >> tid) & 1ull, a1 = (alive >> (tid + 32)) & 1ull;
      u64 keep = alive;
      for (int it = 0; it < 64; ++it) {
        const u32 k0 = __ballot_sync(0xffffffffu, a0 && (m0 & keep) == 0ull);
        const u32 k1 = __ballot_sync(0xffffffffu, a1 && (m1 & keep) == 0ull);
        const u64 nxt = ((u64)k1 << 32) | k0;
        if (nxt == keep) break;
        keep = nxt;
      }
      u64 kept = keep;
      int fin = 0;
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
      finished = 1;
      break;
    }
    const int nw = nblk - blk - 1;
    const int w = tid >> 6, row = tid & 63;
    if (w < nw) {
      const u64 v = ((kept >> row) & 1ull) ? sM[(blk * 64 + row) * pitch + blk + 1 + w] : 0ull;
      const u32 vlo = __reduce_or_sync(0xffffffffu, (u32)v);
      const u32 vhi = __reduce_or_sync(0xffffffffu, (u32)(v >> 32));
      if ((tid & 31) == 0 && (vlo | vhi)) atomicOr(&s_removed[blk + 1 + w], ((u64)vhi << 32) | vlo);
    }
  }
  __syncthreads();
  *fin_out = finished;
  return nk;
}

// grid B * kClusterSize CTAs, cluster (kClusterSize,1,1), block 1024.
// ITEMS: keys per thread; the image's A*hc*wc anchors must fit kClusterSize*ITEMS*1024.
template <int ITEMS>
__global__ void __launch_bounds__(1024, 1) proposal_cluster_kernel(ClusterArgs g) {
  namespace cgx = cooperative_groups;
  extern __shared__ __align__(16) unsigned char dyn_smem[];
  __shared__ u32 s_hist[256];
  __shared__ u32 s_ghist[12][256];  // leader: cluster-wide digit histograms: head select 0..3, full select 4..10
  __shared__ int s_cnt[4];          // leader: n_valid, nan
  __shared__ int s_sel[kClusterSize];  // leader: selected count of each CTA
  __shared__ int s_status[2];       // leader: kept so far, quota reached
  __shared__ u32 s_st[4];
  __shared__ int s_local[4];
  __shared__ u64 s_keptbits;
  __shared__ int s_finish;
  cgx::cluster_group cluster = cgx::this_cluster();
  const int rank = (int)cluster.block_rank();
  const int b = blockIdx.x / kClusterSize;
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
  int* L_cnt = cluster.map_shared_rank(s_cnt, 0);
  int* L_sel = cluster.map_shared_rank(s_sel, 0);
  int* L_status = cluster.map_shared_rank(s_status, 0);
  if (rank == 0) {
    for (int i = tid; i < 12 * 256; i += blockDim.x) (&s_ghist[0][0])[i] = 0;
    if (tid < 4) s_cnt[tid] = 0;
    if (tid < kCSW) s_removed[tid] = 0ull;
    if (tid < 2) s_status[tid] = 0;
  }
  if (tid < 4) s_local[tid] = 0;
  MXR_STAMP(0);
  cluster.sync();
  MXR_STAMP(1);

  // ---- P1: decode + clip + min-size filter (proposal.py:77-120); keys stay in registers ----
  const int hwc = dg.hc * dg.wc;
  const size_t hw = (size_t)dg.H * dg.W;
  const float* fg = g.cls_prob + (size_t)b * pg.cls_img_stride + pg.cls_fg_offset;
  const float* deltas = g.bbox_pred + (size_t)b * 4 * dg.A * hw;
  const float im_h = __ldg(g.im_info + b * 3), im_w = __ldg(g.im_info + b * 3 + 1), im_s = __ldg(g.im_info + b * 3 + 2);
  const int chunk = (pg.n_crop + kClusterSize - 1) / kClusterSize;
  const int q_begin = rank * chunk;
  const int q_end = min(q_begin + chunk, pg.n_crop);
  u32 rk[ITEMS];
  {
    int nvalid = 0, nnan = 0;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const int q = q_begin + tid + i * 1024;
      u32 key = 0;
      bool nan = false;
      if (q < q_end) {
        const int a = q / hwc, r = q - a * hwc;
        const int h = r / dg.wc, w = r - h * dg.wc;
        double bx[4];
        bool dnan;
        decode_clip(g.at, dg, deltas, h, w, a, im_h, im_w, bx, &dnan);
        float s;
        if (!pg.cropped) {
          s = __ldg(fg + (size_t)a * hw + (size_t)h * dg.W + w);
        } else {  // quirk Q1: scores are read at the flat (H,W,A) index of the UNCROPPED map
          const int p = r * dg.A + a;
          const int a2 = p % dg.A, w2 = (p / dg.A) % dg.W, h2 = p / (dg.A * dg.W);
          s = __ldg(fg + (size_t)a2 * hw + (size_t)h2 * dg.W + w2);
        }
        const bool valid = passes_min_size(bx, g.min_size, im_s);
        nan = (s != s) || dnan;
        key = valid ? score_key32(s) : 0u;
      }
      rk[i] = key;
      nvalid += __popc(__ballot_sync(0xffffffffu, key != 0u));
      nnan += __popc(__ballot_sync(0xffffffffu, nan));
    }
    if (pg.cropped) {  // proposal.py:58-62 checks the whole arrays, cropped or not
      const int fchunk = (pg.n_full + kClusterSize - 1) / kClusterSize;
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
    if (tid == 0) {
      if (s_local[0]) atomicAdd(L_cnt, s_local[0]);
      if (s_local[1]) atomicAdd(L_cnt + 1, s_local[1]);
    }
  }
  cluster.sync();
  if (tid == 0) {
    s_local[2] = L_cnt[0];
    s_local[3] = L_cnt[1];
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
    if (tid == 0) {
      int off = 0;
      for (int r = 0; r < rank; ++r) off += L_sel[r];
      s_local[1] = off;   // base of this CTA inside the sort buffer
      s_local[0] = 0;     // running position inside this CTA
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
      if (n_pow2 <= 1024) bitonic_sort_1024_regs(sortbuf, n_pow2);
      else bitonic_sort_desc_u64(sortbuf, n_pow2);
    }
    cluster.sync();
    MXR_STAMP(stamp + 1);
    const int per = (k_sel + kClusterSize - 1) / kClusterSize;
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

  // ---- greedy NMS over sorted positions [s_begin*256, n_sorted) in stripes of 256 boxes: cluster barriers
  //      cost ~0.2 us, so narrow stripes keep the evaluated IoU pairs close to what the walk really needs ----
  const double4* box = g.a.box + (size_t)b * g.a.cap;
  const double* area = g.a.area + (size_t)b * g.a.cap;
  int* kept_slot = g.a.kept_slot + (size_t)b * g.a.keep_cap;
  const double thr_lo = Ar<double>::mul(g.thr, Ar<double>::band_lo()), thr_hi = Ar<double>::mul(g.thr, Ar<double>::band_hi());
  int nk = 0, fin = 0, stripes_done = 0;
  auto walk_stripes = [&](int s_begin, int n_sorted) {
    const int nstripes = (n_sorted + kCW - 1) / kCW;
    for (int s = s_begin; s < nstripes && !fin; ++s) {
      const int c0 = s * kCW;
      const int c1 = min(c0 + kCW, n_sorted);
      {
        // one warp per task (32 rows x 16 columns); task u runs on CTA u % 8, warp u / 8, so that a stripe's
        // few dozen tasks spread over all 8 SMs' FP64 pipes
        const int tp = kCTriTasks + 16 * ((nk + 31) / 32);
        for (int u = warp * kClusterSize + rank; u < tp; u += kClusterSize * 32) {
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
          __syncwarp();
          if (lane < ncols) {
            t_box[lane] = ldcg_box(box + col0 + lane);
            t_area[lane] = __ldcg(area + col0 + lane);
          }
          __syncwarp();
          if (tri) {
            const int rt = half * 32 + lane;           // row inside the 64x64 tile
            const int rl = rb * 64 + rt;               // row inside the stripe
            const int row = c0 + rl;
            if (row < c1) {
              int je = ncols;
              if (rb == cb) je = min(ncols, max(0, rt - qd * 16));   // diagonal tile: strictly BELOW the diagonal
              const u32 bits = tile_bits16(ldcg_box(box + row), __ldcg(area + row), t_box, t_area, 0, je, g.thr, thr_lo, thr_hi);
              reinterpret_cast<unsigned short*>(L_sM + rl * kCPitch + cb)[qd] = (unsigned short)bits;
            }
          } else {
            u32 bits = 0;
            const int kidx = rc * 32 + lane;
            if (kidx < nk) {
              const int slot = __ldcg(kept_slot + kidx);
              bits = tile_bits16(ldcg_box(box + slot), __ldcg(area + slot), t_box, t_area, 0, ncols, g.thr, thr_lo, thr_hi);
            }
            bits = __reduce_or_sync(0xffffffffu, bits);
            if (lane == 0 && bits) atomicOr(reinterpret_cast<u32*>(L_removed + cb) + (qd >> 1), bits << (16 * (qd & 1)));
          }
        }
      }
      cluster.sync();
      if (rank == 0) {
        int f = 0;
        nk = scan_rows_smem(sM, kCPitch, s_removed, c1 - c0, c0, nk, g.max_keep, kept_slot, &s_keptbits, &s_finish, &f);
        if (tid < kCSW) s_removed[tid] = 0ull;
        if (tid == 0) {
          s_status[0] = nk;
          s_status[1] = f;
        }
      }
      cluster.sync();
      if (tid == 0) {
        s_local[2] = L_status[0];
        s_local[3] = L_status[1];
      }
      __syncthreads();
      nk = s_local[2];
      fin = s_local[3];
      stripes_done = s + 1;
      __syncthreads();
    }
  };

  // ---- head: select + sort + walk the first k1 candidates ----
  bool head_ok = false;
  if (k1 > 0) {
    u64 T = 1ull << 32;  // every valid composite
    head_ok = (k1 == n_valid) ? true : cluster_select(k1, 0, false, &T);
    MXR_STAMP(3);
    if (head_ok) {
      sort_and_redecode(k1, T, 4);
      MXR_STAMP(6);
      walk_stripes(0, k1);
    }
  }
  // ---- full list, only if the head did not settle the image ----
  int full_path = 0;
  if (k > 0 && !fin && (!head_ok || k1 < k)) {
    full_path = 1;
    u64 T = 1ull << 32;
    if (k < n_valid) cluster_select(k, 4, true, &T);
    sort_and_redecode(k, T, 10);
    if (!head_ok) {
      nk = 0;            // nothing was walked (a run of equal scores is cut by the head boundary)
      stripes_done = 0;
    }
    walk_stripes(stripes_done, k);
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
      cb->nkept = nk;
      cb->done = 1;
      cb->pad[0] = 0;
    }
    __syncthreads();
    coop_write_outputs<double>(g.a, g.o, cb, b, nk, k);
  }
  cluster.sync();  // nobody leaves while its shared memory may still be addressed
  MXR_STAMP(8);
  if (g.dbg && rank == 0 && tid == 0) {
    g.dbg[(size_t)b * 16 + 9] = (u64)stripes_done;
    g.dbg[(size_t)b * 16 + 14] = (u64)full_path;
  }
}

static inline size_t cluster_smem_bytes(int sort_cap) {
  return (size_t)sort_cap * 8 + (size_t)kCW * kCPitch * 8 + 8 * 8 + 32 * 16 * (sizeof(double4) + sizeof(double)) + 256;
}

}  // namespace mxr

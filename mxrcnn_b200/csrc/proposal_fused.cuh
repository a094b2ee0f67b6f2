// Fused RPN proposal kernel, second generation: ONE launch, one thread-block cluster per image, any list length.
//
// rcnn/rpn/proposal.py:39-149 end to end.  What changed against proposal_cluster.cuh (kept as the A/B path):
//   * no sort and no sort buffer.  Every valid anchor's composite (score key << 32 | ~row) is scattered into the
//     slot range of its linear score bin ("bucket") in global memory; bucket starts come from the cluster-wide bin
//     table, which travels through distributed shared memory in two levels (32 coarse counts per CTA, then only the
//     fine counts of the coarse bins that are actually materialised).  Buckets are small (N / 1024 candidates), so
//     the exact order inside a bucket is established by counting, per stripe, while the stripe is staged.
//   * the candidate list is materialised in windows of coarse bins, top scores first: the first window holds about
//     twice the post-NMS quota, the second one (only if the walk gets that far) everything down to the pre-NMS cut.
//     pre_nms_top_n = -1 (train_alternate.py:30-31) and the TRAIN defaults (config.py:47-50) take the same path.
//   * greedy NMS in stripes of up to 512 candidates.  Every CTA stages the stripe (composites -> exact order ->
//     float32 boxes) in its own shared memory; IoU bit blocks of 32 x 32 pairs are spread over all warps of the
//     cluster.  A pair is first tested in float32 against a rigorous error bound (16 instructions); only pairs
//     inside the bound are decided by the float64 arithmetic of the reference (nms.py:25-36), so the keep list
//     is the same bit for bit as with float64 everywhere.
//   * the stripe is resolved by the leader CTA as ONE block-wide fixed point (one thread per candidate, ballots into a
//     keep word array) instead of a one-warp walk over 64-box blocks.
#pragma once
#include <cooperative_groups.h>

#include "nms_kernels.cuh"
#include "pair_screen.cuh"
#include "proposal_kernels.cuh"

namespace mxr {
namespace fz {

#ifndef FZ_P1_GROUP
#define FZ_P1_GROUP 4  // anchors per decode round (loads first, then arithmetic)
#endif
static const int kBins = 1024;              // fine linear score bins: bin(s) = clamp(int(s * 1024))
static const int kCoarse = 32;              // coarse bins of 32 fine bins each

struct FusedArgs {
  ProposalGeom pg;
  AnchorTable at;
  float4 at32[64];   // per anchor: width, height, centre x, centre y of the base anchor (float32)
  float at_err[64];  // per anchor: largest conversion error of those four values
  float fs32, fs_err;
  const float* cls_prob;
  const float* bbox_pred;
  const float* im_info;
  float min_size;
  int pre_nms;
  int max_keep;      // > 0: post_nms_top_n
  int items;         // score keys per thread: ceil(n_crop / (cluster size * 1024))
  int null_run;      // MXRCNN_B200_NULL=1 (measurement only): leave right after the dependency wait
  // launch constants, computed once on the host instead of by every thread (integer divisions, reciprocals)
  int chunk;         // anchors per CTA: ceil(n_crop / cluster size)
  int hwc, hw32;     // hc * wc, H * W
  float rcp_hwc, rcp_wc, rcp_A;
  int a_step, r_step, h_step, w_step;  // a thread's next anchor is 1024 further: 1024 = a_step * hwc + r_step,
                                       // r_step = h_step * wc + w_step
  int first_stripe;  // width of the first stripe (multiple of 32, <= kWMax)
  int next_stripe;   // width of the later stripes
  int window0;       // candidates the first window should hold
  double thr;
  u64* comp;         // [B][n_crop] bucketed composites
  float4* kbox;      // [B][keep_cap] kept boxes: x1, y1, x2 + 1, y2 + 1 (float32)
  float4* kaux;      // [B][keep_cap] kept boxes: scaled area, tolerance, row index (bits), keep position (bits)
  int keep_cap;
  u64* sortbuf;      // [B][sort_cap] scratch for buckets too large to be ranked by counting (rare)
  int sort_cap;
  NmsOutputs o;
  Ctrl* ctrl;
  u64* dbg;          // optional [b][16] globaltimer stamps (leader CTA), or null
  u64* clk;          // optional [b][32] clock64 stamps of the leader's first thread (first stripe), or null
  u32* pair_count;   // optional [b][2] IoU pairs evaluated / sent to the float64 test (instrumented runs only), or null
};

__device__ __forceinline__ u64 global_ns() {
  u64 t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define FZ_STAMP(i)                                                               \
  do {                                                                            \
    if (g.dbg && rank == 0 && tid == 0) g.dbg[(size_t)b * 16 + (i)] = global_ns(); \
  } while (0)

#define FZ_CLK(i)                                                                             \
  do {                                                                                        \
    if (g.clk && rank == 0 && tid == 0 && stripes == 0) g.clk[(size_t)b * 32 + (i)] = (u64)clock64(); \
  } while (0)

__device__ __forceinline__ int bin_of_key(u32 key) {
  const u32 bits = (key & 0x80000000u) ? (key & 0x7fffffffu) : ~key;  // inverse of score_key32
  const int v = (int)(__uint_as_float(bits) * (float)kBins);           // saturating conversion
  return v < 0 ? 0 : (v > kBins - 1 ? kBins - 1 : v);
}

// The same bin from the score itself (bin_of_key(score_key32(s)) == bin_of_score(s): the key is a bijection of s + 0).
__device__ __forceinline__ int bin_of_score(float s) {
  const int v = (int)(s * (float)kBins);
  return v < 0 ? 0 : (v > kBins - 1 ? kBins - 1 : v);
}

// float32 image of bbox_pred + clip_boxes for one anchor, with a bound `e` on the distance of every coordinate
// from the float64 value decode_clip_values() computes for the same inputs (same expf value in both).
//   t1 = dx * bw, pcx = t1 + cx, pw = exp(dw) * bw, x1 = pcx - 0.5 * (pw - 1): each operation is within 2^-24 relative,
//   so |x1 - X1| <= u (|t1| + 2 |pcx| + 1.5 pw + 1) (+ the conversion errors of the anchor and the shift); twice that
//   is used.  The clip is 1-Lipschitz and its limits are the same float32 scalars in both computations.
struct Dec32 {
  float x1, y1, x2, y2, e;
};
__device__ __forceinline__ Dec32 decode32_core(const FusedArgs& g, float dx, float dy, float dw, float dh, int h, int w, int a,
                                               float lim_x, float lim_y) {
  const float u2 = 1.1920929e-07f;  // 2^-23
  const float4 an = g.at32[a];
  const float sx = (float)w * g.fs32, sy = (float)h * g.fs32;
  const float cx = an.z + sx, cy = an.w + sy;
  const float t1 = dx * an.x, t2 = dy * an.y;
  const float pcx = t1 + cx, pcy = t2 + cy;
  const float pw = expf(dw) * an.x, ph = expf(dh) * an.y;
  const float hx = 0.5f * (pw - 1.0f), hy = 0.5f * (ph - 1.0f);
  Dec32 r;
  r.x1 = fmaxf(fminf(pcx - hx, lim_x), 0.0f);  // same order as clip_0_lim: min with the limit first
  r.y1 = fmaxf(fminf(pcy - hy, lim_y), 0.0f);
  r.x2 = fmaxf(fminf(pcx + hx, lim_x), 0.0f);
  r.y2 = fmaxf(fminf(pcy + hy, lim_y), 0.0f);
  const float conv = g.at_err[a] + (float)(w > h ? w : h) * g.fs_err;
  const float mx = fabsf(t1) + 3.0f * fabsf(pcx) + 2.0f * pw + sx;
  const float my = fabsf(t2) + 3.0f * fabsf(pcy) + 2.0f * ph + sy;
  r.e = __fmaf_rn(u2, fmaxf(mx, my) + 2.0f, conv);
  return r;
}

__device__ __forceinline__ Box32 decode32(const FusedArgs& g, float dx, float dy, float dw, float dh, int h, int w, int a,
                                          float lim_x, float lim_y, float c32) {
  const float u2 = 1.1920929e-07f;
  const Dec32 d = decode32_core(g, dx, dy, dw, dh, h, w, a, lim_x, lim_y);
  const float L = fmaxf(lim_x, lim_y);
  const float e = d.e + u2 * (L + 2.0f);
  Box32 r;
  r.x1 = d.x1;
  r.y1 = d.y1;
  r.x2p = d.x2 + 1.0f;
  r.y2p = d.y2 + 1.0f;
  const float bw = r.x2p - d.x1, bh = r.y2p - d.y1;
  const float area = bw * bh;
  r.a = c32 * area;
  // |D^ - D| for a pair is at most t_i + t_j (derivation: DESIGN.md section 3, "float32 screen"); doubled again here
  const float t = 2.02f * ((bw + bh + 2.0f) * (4.0f * e + u2 * (L + 1.0f)) + 10.0f * (0.5f * u2) * area + 4.0f * e * e);
  // not a finite small bound, or a degenerate box: every pair with this box goes to the exact test
  r.t = (e < 0.0625f && bw >= 1.0f && bh >= 1.0f && t < 1e30f) ? t : __int_as_float(0x7f800000);
  return r;
}

// proposal.py:113-120 for one anchor in the reference's float64 arithmetic (the float32 screen could not decide).
__device__ __noinline__ bool exact_min_size(const FusedArgs& g, float dx, float dy, float dw, float dh, int h, int w, int a,
                                            float im_h, float im_w, float im_s) {
  double bx[4];
  decode_clip_values(g.at, g.pg.g, dx, dy, dw, dh, h, w, a, im_h, im_w, bx);
  return passes_min_size(bx, g.min_size, im_s);
}

// bbox_pred + clip_boxes for one row index in the reference's float64 arithmetic (proposal.py:110-113).
__device__ __noinline__ void exact_box(const FusedArgs& g, const float* __restrict__ deltas, float im_h, float im_w, int p,
                                       double out[4]) {
  const DecodeGeom& dg = g.pg.g;
  const int a = p % dg.A, r = p / dg.A, h = r / dg.wc, w = r - h * dg.wc;
  decode_clip(g.at, dg, deltas, h, w, a, im_h, im_w, out);
}

// The reference's own arithmetic for one pair (float64, nms.py:25-36), both boxes decoded from their row indices.
__device__ __noinline__ bool exact_pair(const FusedArgs& g, const float* __restrict__ deltas, float im_h, float im_w,
                                        int p_row, int p_col) {
  double br[4], bc[4];
  exact_box(g, deltas, im_h, im_w, p_row, br);
  exact_box(g, deltas, im_h, im_w, p_col, bc);
  return iou_exceeds<double>(br[0], br[1], br[2], br[3], box_area<double>(br[0], br[1], br[2], br[3]), bc[0], bc[1], bc[2],
                             bc[3], box_area<double>(bc[0], bc[1], bc[2], bc[3]), g.thr);
}

// Min-size filter (proposal.py:113-120, 156-161) on the float32 image of the box; the reference's float64 arithmetic
// decides only when a side is within the error bound of the threshold (or nothing is finite).  Same decision as
// passes_min_size(decode_clip_values(...)) for every input.
__device__ __forceinline__ bool min_size_screened(const FusedArgs& g, float dx, float dy, float dw, float dh, int h, int w,
                                                  int a, float lim_x, float lim_y, float ms32, float im_h, float im_w,
                                                  float im_s) {
  const Dec32 d = decode32_core(g, dx, dy, dw, dh, h, w, a, lim_x, lim_y);
  const float ws = (d.x2 - d.x1) + 1.0f, hs = (d.y2 - d.y1) + 1.0f;
  const float m = 2.0f * d.e + 2.3841858e-07f * (fmaxf(ws, hs) + 1.0f);  // 2e + 2^-22 (side + 1)
  const float lo = fminf(ws, hs) - ms32;
  bool valid = lo > m;
  if (!valid && !(lo < -m)) valid = exact_min_size(g, dx, dy, dw, dh, h, w, a, im_h, im_w, im_s);
  return valid;
}

// Stage entry (mxr_decode_clip): what the fused kernel decides and decodes, for EVERY anchor of the cropped map, in the
// reference's row order p = (h * wc + w) * A + a (proposal.py:88-120).  boxes: the float64 decode the kernel's outputs
// are cast from; valid: the min-size decision of the screened filter above.  grid (ceil(n_crop / 256), B).
__global__ void __launch_bounds__(256) decode_stage_kernel(const __grid_constant__ FusedArgs g, double* __restrict__ boxes,
                                                           int* __restrict__ valid_out) {
  const ProposalGeom& pg = g.pg;
  const DecodeGeom& dg = pg.g;
  const int b = blockIdx.y;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= pg.n_crop) return;
  const size_t hw = (size_t)dg.H * dg.W;
  const float* deltas = g.bbox_pred + (size_t)b * 4 * dg.A * hw;
  const float im_h = __ldg(g.im_info + b * 3), im_w = __ldg(g.im_info + b * 3 + 1), im_s = __ldg(g.im_info + b * 3 + 2);
  const float lim_x = __fsub_rn(im_w, 1.0f), lim_y = __fsub_rn(im_h, 1.0f);
  const float ms32 = __fmul_rn(g.min_size, im_s);
  const int a = p % dg.A, r = p / dg.A, h = r / dg.wc, w = r - h * dg.wc;
  const float* d = deltas + ((size_t)a * 4) * hw + (size_t)h * dg.W + w;
  const float dx = __ldg(d), dy = __ldg(d + hw), dw = __ldg(d + 2 * hw), dh = __ldg(d + 3 * hw);
  double bx[4];
  decode_clip_values(g.at, dg, dx, dy, dw, dh, h, w, a, im_h, im_w, bx);
  double* o = boxes + ((size_t)b * pg.n_crop + p) * 4;
  o[0] = bx[0];
  o[1] = bx[1];
  o[2] = bx[2];
  o[3] = bx[3];
  valid_out[(size_t)b * pg.n_crop + p] = min_size_screened(g, dx, dy, dw, dh, h, w, a, lim_x, lim_y, ms32, im_h, im_w, im_s) ? 1 : 0;
}

// A bucket with more candidates than a stripe is wide (scores crowded into one bin) cannot be ranked by counting while
// it is staged: the leader CTA orders it in place with a sort in global memory -- first by the selection order (score
// desc, row desc: proposal.py:123-125 with the documented tie rule) to apply the pre-NMS cut to the first r, then by the
// processing order (score desc, row asc: nms.py:19 re-sorts the already descending list).  All threads of ONE CTA call it.
__device__ __noinline__ void presort_bucket(u64* __restrict__ bucket, int m, int r, u64* __restrict__ sb) {
  const int tid = threadIdx.x;
  const int n_pow2 = max(2, (int)(1u << (32 - __clz(max(m, 2) - 1))));
  for (int i = tid; i < n_pow2; i += blockDim.x) sb[i] = i < m ? (__ldcg(bucket + i) ^ 0xffffffffull) : 0ull;
  __syncthreads();
  bitonic_sort_desc_u64(sb, n_pow2);
  for (int i = tid; i < n_pow2; i += blockDim.x) sb[i] = i < r ? (sb[i] ^ 0xffffffffull) : 0ull;
  __syncthreads();
  bitonic_sort_desc_u64(sb, n_pow2);
  for (int i = tid; i < r; i += blockDim.x) bucket[i] = sb[i];
}

// First bin >= lo_mat whose bucket ends at or before `lim` (s_end is non-increasing in the bin index), kBins if none:
// two ballots over 32 groups of 32 bins instead of a ten-step dependent binary search.  Whole warps call it.
__device__ __forceinline__ int first_bin_ending_by(const u32* s_end, int lo_mat, u32 lim, int lane) {
  const int gb = 32 * lane + 31;
  const u32 m1 = __ballot_sync(0xffffffffu, gb >= lo_mat && s_end[gb] <= lim);
  if (!m1) return kBins;
  const int g1 = __ffs(m1) - 1;
  const int bin = 32 * g1 + lane;
  const u32 m2 = __ballot_sync(0xffffffffu, bin >= lo_mat && s_end[bin] <= lim);
  return 32 * g1 + __ffs(m2) - 1;
}

// grid (csize, B), cluster (csize,1,1), block 1024.  g.items keys per thread (n_crop <= csize * items * 1024).
__global__ void __launch_bounds__(1024, 1) proposal_fused_kernel(const __grid_constant__ FusedArgs g) {
  namespace cgx = cooperative_groups;
  extern __shared__ __align__(16) unsigned char dyn_smem[];
  __shared__ u32 s_bins[kBins];       // decode: this CTA's valid scores per bin; later: scatter cursors
  __shared__ u32 s_pack[kBins / 2];   // the same as pairs of 16-bit counts: what the other CTAs read
  __shared__ u32 s_cpub[kCoarse];     // this CTA's counts per coarse bin (read by the other CTAs)
  __shared__ u32 s_base[kBins];       // cluster-wide: candidates in higher bins = first slot of the bucket
  __shared__ u32 s_cnt[kBins];        // cluster-wide: size of the bucket
  __shared__ u32 s_own[kBins];        // first slot of this CTA's candidates inside the bucket
  __shared__ u32 s_end[kBins];        // cluster-wide: first slot after the bucket (s_base + s_cnt)
  __shared__ u32 s_ctot[kCoarse], s_csuf[kCoarse];
  __shared__ u32 s_scan[34];
  __shared__ int s_plan[8];           // 0 cut coarse bin, 1 first window's lowest coarse bin, 2 cut bin, 3 r_cut, 4 NaN count
  __shared__ int s_pub[2];            // this CTA's NaN flag (read by the other CTAs)
  __shared__ __align__(16) u32 s_keep[32];  // [2][16] keep words of the stripe being resolved (ping-pong)
  __shared__ __align__(16) u32 s_kfin[16];  // keep words of the resolved stripe
  cgx::cluster_group cluster = cgx::this_cluster();
  const int rank = (int)cluster.block_rank();
  const int csize = (int)cluster.num_blocks();
  const int b = blockIdx.y;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const ProposalGeom& pg = g.pg;
  const DecodeGeom& dg = pg.g;
  // dynamic shared memory: stripe staging (every CTA), bit rows (leader), score keys of this CTA's anchors
  float4* s_cbox = reinterpret_cast<float4*>(dyn_smem);                 // [kWMax] x1, y1, x2+1, y2+1
  float2* s_caux = reinterpret_cast<float2*>(s_cbox + kWMax);           // [kWMax] scaled area, tolerance
  u64* s_comp = reinterpret_cast<u64*>(s_caux + kWMax);                 // [kWMax] raw composites of the stripe
  int* s_p = reinterpret_cast<int*>(s_comp + kWMax);                    // [kWMax] row index
  int* s_kpos = s_p + kWMax;                                            // [kWMax] position in the pre-NMS order
  u32* s_skey = reinterpret_cast<u32*>(s_kpos + kWMax);                 // [kWMax] score key
  float4* s_roi = reinterpret_cast<float4*>(s_skey + kWMax);            // [kWMax / 2] output rows of this CTA's share of the stripe
  u32* s_rem = reinterpret_cast<u32*>(s_roi + kWMax / 2);               // [16] this CTA's share of the removed bits
  u32* s_removed = s_rem + 16;                                          // [2][16] removed by earlier stripes (stripe parity)
  u32* s_M = s_removed + 32;                                            // [2][kMWords][kWMax] bit rows (stripe parity)
  u32* s_keys = s_M + 2 * kWMax * kMWords;                              // [items rounded up to 4][1024] score keys, 0 = filtered out
  int stripes = 0;
  if (tid < 32) s_removed[tid] = 0u;
  if (tid < 16) s_rem[tid] = 0u;
  asm volatile("griddepcontrol.launch_dependents;");
#ifndef FZ_NO_PREFETCH
  // While this CTA waits for the previous kernel in the stream (programmatic dependent launch), it asks L2 for the slab
  // of each input plane it is about to decode: the scores of the anchors rank * chunk ... and their four delta planes,
  // at most two anchors' worth, i.e. ten contiguous pieces.  A prefetch has no architectural effect (L2 is the point
  // of coherence: a producer kernel that is still writing these planes simply overwrites the prefetched lines), so it
  // may run ahead of griddepcontrol.wait; the decode loop then finds its five loads per anchor in L2.
  if (tid < 10) {
    const DecodeGeom& dg0 = g.pg.g;
    const int q_lo = rank * g.chunk, q_hi = min(q_lo + g.chunk, g.pg.n_crop);
    if (q_lo < q_hi) {
      const int a_lo = q_lo / g.hwc, a_hi = (q_hi - 1) / g.hwc;
      const int piece = tid / 5, plane = tid - piece * 5;  // plane 0: score, 1..4: dx, dy, dw, dh
      const int a = a_lo + piece;
      if (a <= a_hi) {
        const int r_lo = a == a_lo ? q_lo - a * g.hwc : 0, r_hi = a == a_hi ? q_hi - 1 - a * g.hwc : g.hwc - 1;
        const int h_lo = r_lo / dg0.wc, h_hi = r_hi / dg0.wc;
        const size_t first = (size_t)h_lo * dg0.W + (r_lo - h_lo * dg0.wc), last = (size_t)h_hi * dg0.W + (r_hi - h_hi * dg0.wc);
        const float* base = plane == 0 ? g.cls_prob + (size_t)b * g.pg.cls_img_stride + g.pg.cls_fg_offset + (size_t)a * g.hw32
                                       : g.bbox_pred + ((size_t)b * 4 * dg0.A + (size_t)a * 4 + (plane - 1)) * g.hw32;
        // whole 16-byte units inside [first, last]: never a byte outside the planes this CTA reads anyway
        const unsigned long long lo = ((unsigned long long)(base + first) + 15ull) & ~15ull;
        const unsigned long long hi = (unsigned long long)(base + last + 1) & ~15ull;
        if (hi > lo) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(lo), "r"((unsigned)(hi - lo)) : "memory");
      }
    }
  }
#endif
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (g.null_run) return;  // launch + dependency floor of this configuration (tools/cluster_sweep2.py)
  s_bins[tid] = 0;  // kBins == blockDim.x
  __syncthreads();
  FZ_STAMP(0);
  FZ_STAMP(1);

  // ---- P1: decode + clip + min-size filter (proposal.py:77-120); score keys go to shared memory ----
  const int hwc = g.hwc;
  const int hw32 = g.hw32;
  const size_t hw = (size_t)hw32;
  const float* fg = g.cls_prob + (size_t)b * pg.cls_img_stride + pg.cls_fg_offset;
  const float* deltas = g.bbox_pred + (size_t)b * 4 * dg.A * hw;
  const float im_h = __ldg(g.im_info + b * 3), im_w = __ldg(g.im_info + b * 3 + 1), im_s = __ldg(g.im_info + b * 3 + 2);
  const int chunk = g.chunk;
  const int q_begin = rank * chunk;
  const int q_end = min(q_begin + chunk, pg.n_crop);
  const float rcp_hwc = g.rcp_hwc, rcp_wc = g.rcp_wc, rcp_A = g.rcp_A;
  const float lim_x = __fsub_rn(im_w, 1.0f), lim_y = __fsub_rn(im_h, 1.0f);
  const float ms32 = __fmul_rn(g.min_size, im_s);  // proposal.py:117: an f32 product under numpy >= 2 (NEP 50)
  const int items = g.items;
  {
    bool nanflag = false;
    // this thread's anchors are q_begin + tid + 1024 i: (a, h, w) of the first by division, of the others by stepping
    int qa, qr, qh, qw;
    {
      const int q0 = q_begin + tid;
      qa = fast_div(q0, hwc, rcp_hwc);
      qr = q0 - qa * hwc;
      qh = fast_div(qr, dg.wc, rcp_wc);
      qw = qr - qh * dg.wc;
    }
    // Four anchors per round: first ALL twenty loads (independent of each other: one DRAM round trip for an image of up
    // to four keys per thread), then the arithmetic.
#pragma unroll 1
    for (int i0 = 0; i0 < items; i0 += FZ_P1_GROUP) {
      int ia[FZ_P1_GROUP], ih[FZ_P1_GROUP], iw[FZ_P1_GROUP];
      float dx[FZ_P1_GROUP], dy[FZ_P1_GROUP], dw[FZ_P1_GROUP], dh[FZ_P1_GROUP], sc[FZ_P1_GROUP];
#pragma unroll
      for (int j = 0; j < FZ_P1_GROUP; ++j) {
        const int q = q_begin + tid + (i0 + j) * 1024;
        ia[j] = -1;
        ih[j] = qh;
        iw[j] = qw;
        dx[j] = dy[j] = dw[j] = dh[j] = sc[j] = 0.0f;
        if (q < q_end) {
          ia[j] = qa;
          const int od = qa * 4 * hw32 + qh * dg.W + qw;
          dx[j] = __ldg(deltas + od);
          dy[j] = __ldg(deltas + od + hw32);
          dw[j] = __ldg(deltas + od + 2 * hw32);
          dh[j] = __ldg(deltas + od + 3 * hw32);
          if (!pg.cropped) {
            sc[j] = __ldg(fg + q);  // q = a * H * W + h * W + w on an uncropped map
          } else {  // quirk Q1: scores are read at the flat (H,W,A) index of the UNCROPPED map
            const int p = qr * dg.A + qa;
            const int a2 = p % dg.A, w2 = (p / dg.A) % dg.W, h2 = p / (dg.A * dg.W);
            sc[j] = __ldg(fg + (size_t)a2 * hw + (size_t)h2 * dg.W + w2);
          }
        }
        // step to this thread's next anchor (1024 further)
        qa += g.a_step;
        qr += g.r_step;
        qh += g.h_step;
        qw += g.w_step;
        if (qw >= dg.wc) {
          qw -= dg.wc;
          ++qh;
        }
        if (qr >= hwc) {
          qr -= hwc;
          qh -= dg.hc;
          ++qa;
        }
      }
#pragma unroll
      for (int j = 0; j < FZ_P1_GROUP; ++j) {
        u32 key = 0;
        if (ia[j] >= 0) {
          const bool valid = min_size_screened(g, dx[j], dy[j], dw[j], dh[j], ih[j], iw[j], ia[j], lim_x, lim_y, ms32, im_h,
                                               im_w, im_s);
          // the clip drops NaNs, so the reference's NaN check (proposal.py:58-62) looks at the raw values
          nanflag |= (sc[j] != sc[j]) || (dx[j] != dx[j]) || (dy[j] != dy[j]) || (dw[j] != dw[j]) || (dh[j] != dh[j]);
          if (valid) {
            key = score_key32(sc[j]);
            atomicAdd(&s_bins[bin_of_score(sc[j])], 1u);
          }
        }
        s_keys[(i0 + j) * 1024 + tid] = key;
      }
    }
    if (pg.cropped) {  // proposal.py:58-62 checks the whole arrays, cropped or not
      const int fchunk = (pg.n_full + csize - 1) / csize;  // (cropped maps only)
      for (int q = rank * fchunk + tid; q < min((rank + 1) * fchunk, pg.n_full); q += 1024) {
        const float v0 = __ldg(fg + q), v1 = __ldg(deltas + q), v2 = __ldg(deltas + q + pg.n_full);
        const float v3 = __ldg(deltas + q + 2 * (size_t)pg.n_full), v4 = __ldg(deltas + q + 3 * (size_t)pg.n_full);
        nanflag |= (v0 != v0) || (v1 != v1) || (v2 != v2) || (v3 != v3) || (v4 != v4);
      }
    }
    const int any_nan = __syncthreads_or(nanflag ? 1 : 0);
    if (tid == 0) s_pub[0] = any_nan;
    if (tid < kBins / 2) s_pack[tid] = s_bins[2 * tid] | (s_bins[2 * tid + 1] << 16);
    {  // coarse counts: warp w sums fine bins 32w .. 32w+31
      const u32 v = __reduce_add_sync(0xffffffffu, s_bins[warp * 32 + lane]);
      if (lane == 0) s_cpub[warp] = v;
    }
  }
  cluster.sync();  // ONE barrier ends the decode phase: NaN flags and bin tables of every CTA are published
  FZ_STAMP(2);

  // ---- P2: cluster-wide coarse table, pre-NMS cut, window plan ----
  if (warp == 0) {
    const int c = kCoarse - 1 - lane;  // descending
    u32 tot = 0;
#pragma unroll 4
    for (int r = 0; r < csize; ++r) {
      const int src = rank + r < csize ? rank + r : rank + r - csize;
      tot += cluster.map_shared_rank(s_cpub, src)[c];
    }
    u32 incl = tot;
    for (int o = 1; o < 32; o <<= 1) {
      const u32 v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    s_ctot[c] = tot;
    s_csuf[c] = incl - tot;
  } else if (warp == 1) {
    const int nn = __reduce_add_sync(0xffffffffu, lane < csize ? cluster.map_shared_rank(s_pub, lane)[0] : 0);
    if (lane == 0) s_plan[4] = nn;
  }
  __syncthreads();
  const int n_valid = (int)(s_csuf[0] + s_ctot[0]);
  const int nan_flag = s_plan[4] ? 1 : 0;
  const int k = (g.pre_nms > 0) ? min(g.pre_nms, n_valid) : n_valid;  // proposal.py:124-125
  if (warp == 0) {
    const u32 suf = s_csuf[lane], tot = s_ctot[lane];
    const u32 t0 = (u32)min(k, g.window0 > 0 ? g.window0 : k);
    if (suf < (u32)k && (u32)k <= suf + tot) s_plan[0] = lane;
    if (suf < t0 && t0 <= suf + tot) s_plan[1] = lane;
  }
  __syncthreads();
  const int cc = s_plan[0];     // coarse bin of the pre-NMS cut (valid when k > 0)
  const int c_w0 = s_plan[1];   // lowest coarse bin of the first window
  FZ_STAMP(3);

  // ---- greedy NMS over the bucketed list ----
  u64* comp = g.comp + (size_t)b * pg.n_crop;
  const bool screen_ok = g.thr > 0.0 && g.thr < 1.0;
  const float c32 = screen_ok ? (float)(g.thr / (1.0 + g.thr)) : 0.0f;
  float4* kbox = g.kbox + (size_t)b * g.keep_cap;
  float4* kaux = g.kaux + (size_t)b * g.keep_cap;
  int nk = 0, fin = 0;
  int c0 = 0;          // next raw slot
  int windows = 0;     // windows materialised
  int lo_mat = kBins;  // lowest fine bin materialised so far
  int ps_bin = -1;     // bucket currently ordered in place (too large to be ranked by counting)
  int ps_end = 0;      // end of its selected part
  u32 pairs = 0, n_und = 0;
  while (k > 0 && !fin && stripes < (1 << 20)) {
    // ---- the candidate list is materialised in (at most) two windows of coarse bins, top scores first ----
    if (windows == 0 || (c0 >= (int)s_end[lo_mat] && windows == 1 && c_w0 > cc)) {
      const int c_lo = windows == 0 ? c_w0 : cc, c_hi = windows == 0 ? kCoarse - 1 : c_w0 - 1;
      // fine table of the window's bins (DSMEM pull), bucket starts, scatter of this CTA's composites
      const int top = 32 * c_hi + 31, bot = 32 * c_lo;
      const int npair = 16 * (c_hi - c_lo + 1);  // thread t < npair owns bins hi = top - 2t and lo = hi - 1
      const bool owner = tid < npair;
      const int bhi = top - 2 * tid, blo = bhi - 1;
      u32 n_hi = 0, n_lo = 0, pre_hi = 0, pre_lo = 0;
      if (owner) {  // all the cluster's tables in one round trip (a cluster has at most 16 CTAs)
        u32 v[16];
#pragma unroll
        for (int r = 0; r < 16; ++r) {
          const int o = rank + r, src = o < csize ? o : o - csize;
          v[r] = r < csize ? cluster.map_shared_rank(s_pack, src)[blo >> 1] : 0u;
        }
#pragma unroll
        for (int r = 0; r < 16; ++r) {
          const int o = rank + r, src = o < csize ? o : o - csize;
          n_lo += v[r] & 0xffffu;
          n_hi += v[r] >> 16;
          if (r < csize && src < rank) {
            pre_lo += v[r] & 0xffffu;
            pre_hi += v[r] >> 16;
          }
        }
      }
      // block-wide exclusive scan of (n_hi + n_lo) in thread order = descending bins
      u32 incl = n_hi + n_lo;
      for (int o = 1; o < 32; o <<= 1) {
        const u32 v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      if (lane == 31) s_scan[warp] = incl;
      __syncthreads();
      if (warp == 0) {
        u32 w = s_scan[lane];
        for (int o = 1; o < 32; o <<= 1) {
          const u32 v = __shfl_up_sync(0xffffffffu, w, o);
          if (lane >= o) w += v;
        }
        s_scan[lane] = w;  // inclusive over warps
      }
      __syncthreads();
      const u32 before = (warp > 0 ? s_scan[warp - 1] : 0u) + incl - (n_hi + n_lo);
      if (owner) {
        const u32 base_hi = s_csuf[c_hi] + before, base_lo = base_hi + n_hi;
        const bool in_hi = base_hi < (u32)k, in_lo = base_lo < (u32)k;  // bins below the pre-NMS cut stay empty
        s_base[bhi] = base_hi;
        s_base[blo] = base_lo;
        s_cnt[bhi] = in_hi ? n_hi : 0u;
        s_cnt[blo] = in_lo ? n_lo : 0u;
        s_end[bhi] = base_hi + (in_hi ? n_hi : 0u);
        s_end[blo] = base_lo + (in_lo ? n_lo : 0u);
        s_own[bhi] = base_hi + pre_hi;
        s_own[blo] = base_lo + pre_lo;
        s_bins[bhi] = 0u;
        s_bins[blo] = 0u;
        if (in_hi && (u32)k <= base_hi + n_hi) {
          s_plan[2] = bhi;
          s_plan[3] = k - (int)base_hi;
        }
        if (in_lo && (u32)k <= base_lo + n_lo) {
          s_plan[2] = blo;
          s_plan[3] = k - (int)base_lo;
        }
      }
      __syncthreads();
      // bin(s) >= bot  <=>  s >= bot / 1024 (the bin function is monotone and bot / 1024 is exact), so most keys are
      // turned away by one comparison.  The few that pass are first compacted per warp (a warp-private queue in the
      // staging area, free at this point), so that the slot arithmetic runs on full warps of candidates instead of on
      // every warp that happens to hold one.
      const u32 key_bot = bot == 0 ? 1u : score_key32((float)bot * (1.0f / (float)kBins));
      u64* wq = reinterpret_cast<u64*>(dyn_smem) + warp * 64;  // [64] (key << 32 | item index), 16 KB for 32 warps
      int cnt = 0;
      auto place = [&](int n) {  // lanes < n take queue entries 0 .. n-1
        if (lane < n) {
          const u64 e = wq[lane];
          const u32 key = (u32)(e >> 32);
          const int bin = bin_of_key(key);
          if (bin <= top && s_cnt[bin] != 0u) {
            const u32 slot = s_own[bin] + atomicAdd(&s_bins[bin], 1u);
            const int q = q_begin + tid - lane + (int)(u32)e;  // (u32)e = item * 1024 + lane of the key's owner
            const int a = fast_div(q, hwc, rcp_hwc);
            const u32 p = (u32)((q - a * hwc) * dg.A + a);  // row index of the reference: (h * wc + w) * A + a
            comp[slot] = ((u64)key << 32) | (u32)(~p);
          }
        }
      };
#pragma unroll 1
      for (int i = 0; i < items; ++i) {
        const u32 key = s_keys[i * 1024 + tid];
        const bool sel = key >= key_bot;
        const u32 m = __ballot_sync(0xffffffffu, sel);
        if (m) {
          if (sel) wq[cnt + __popc(m & ((1u << lane) - 1u))] = ((u64)key << 32) | (u32)(i * 1024 + lane);
          cnt += __popc(m);
          __syncwarp();
          if (cnt >= 32) {
            place(32);
            __syncwarp();
            const u64 e = lane < cnt - 32 ? wq[32 + lane] : 0ull;  // move the rest to the front
            __syncwarp();
            if (lane < cnt - 32) wq[lane] = e;
            cnt -= 32;
            __syncwarp();
          }
        }
      }
      place(cnt);
      __syncthreads();  // the queues alias the stripe staging area
      lo_mat = (c_lo == cc) ? s_plan[2] : bot;
      ++windows;
      cluster.sync();  // the window's composites are visible to every CTA of the cluster
      if (windows == 1) {
        FZ_STAMP(4);
        FZ_STAMP(5);
        FZ_STAMP(6);
      }
    }
    const int mat_end = (int)s_end[lo_mat];
    if (c0 >= mat_end) break;
    FZ_CLK(0);
    const int W = stripes == 0 ? g.first_stripe : g.next_stripe;
    const int cut_bin = (windows == 2 || c_w0 == cc) ? s_plan[2] : -1;
    const int r_cut = s_plan[3];
    // ---- plan the stripe [c0, c1): whole buckets, at most W raw slots ----
    int c1, n_s;
    bool presorted = ps_bin >= 0 && c0 < ps_end;
    if (!presorted) {
      const int lo = first_bin_ending_by(s_end, lo_mat, (u32)(c0 + W), lane);
      c1 = lo < kBins ? (int)s_end[lo] : 0;
      if (c1 <= c0) {  // the bucket that starts at c0 holds more than W candidates (scores crowded into one bin)
        const int x = lo - 1;
        const int m = (int)s_cnt[x], base = (int)s_base[x];
        const int r = (x == cut_bin) ? r_cut : m;
        if (rank == 0) presort_bucket(comp + base, m, r, g.sortbuf + (size_t)b * g.sort_cap);
        cluster.sync();
        ps_bin = x;
        ps_end = base + r;
        presorted = true;
      }
    }
    if (presorted) {
      c1 = min(c0 + W, ps_end);
      n_s = c1 - c0;
    } else {
      n_s = c1 - c0;
      if (cut_bin >= 0 && (int)s_end[cut_bin] <= c1 && (int)s_base[cut_bin] >= c0) n_s -= (int)s_cnt[cut_bin] - r_cut;
    }
    const int m_raw = c1 - c0;
    FZ_CLK(1);
    // ---- stage the stripe: composites -> exact order inside each bucket -> float32 boxes ----
    {
      u64 c = 0ull;
      int p = 0, ah = 0, aw = 0, aa = 0;
      float dx = 0.f, dy = 0.f, dw = 0.f, dh = 0.f;
      if (tid < m_raw) {
        c = __ldcg(comp + c0 + tid);
        s_comp[tid] = c;
        p = (int)(~(u32)c);
        const int r = fast_div(p, dg.A, rcp_A);
        aa = p - r * dg.A;
        ah = fast_div(r, dg.wc, rcp_wc);
        aw = r - ah * dg.wc;
        const float* d = deltas + ((size_t)aa * 4) * hw + (size_t)ah * dg.W + aw;
        dx = __ldg(d);
        dy = __ldg(d + hw);
        dw = __ldg(d + 2 * hw);
        dh = __ldg(d + 3 * hw);
      }
      if (tid < 16) s_rem[tid] = 0u;
      FZ_CLK(2);
      __syncthreads();
      FZ_CLK(3);
      if (tid < m_raw) {
        int pos = tid, kpos = -1 - (c0 + tid);  // ordered in place: the position in the pre-NMS order is found on output
        bool selected = true;
        if (!presorted) {
          const u32 key = (u32)(c >> 32);
          const int bin = bin_of_key(key);
          const int bs = (int)s_base[bin] - c0, bn = (int)s_cnt[bin];
          // count, inside the bucket, the greater scores and the equal ones; equal scores are rare and take the second
          // loop, which counts the ones that come first among them (smaller row: composites hold ~row)
          int n_gt = 0, n_eq_tot = 0;
          const u64* bk = s_comp + bs;
          const u32* bk32 = reinterpret_cast<const u32*>(bk) + 1;  // score keys: the high words
          int j = 0;
          for (; j + 4 <= bn; j += 4) {
            const u32 o0 = bk32[2 * j], o1 = bk32[2 * j + 2], o2 = bk32[2 * j + 4], o3 = bk32[2 * j + 6];
            n_gt += (o0 > key ? 1 : 0) + (o1 > key ? 1 : 0) + (o2 > key ? 1 : 0) + (o3 > key ? 1 : 0);
            n_eq_tot += (o0 == key ? 1 : 0) + (o1 == key ? 1 : 0) + (o2 == key ? 1 : 0) + (o3 == key ? 1 : 0);
          }
          for (; j < bn; ++j) {
            const u32 o = bk32[2 * j];
            n_gt += o > key ? 1 : 0;
            n_eq_tot += o == key ? 1 : 0;
          }
          int n_eq_lt = 0;
          if (n_eq_tot > 1) {
            for (j = 0; j < bn; ++j) {
              const u64 o = bk[j];
              n_eq_lt += ((u32)(o >> 32) == key && (u32)o > (u32)c) ? 1 : 0;
            }
          }
          const int n_eq_gt = n_eq_tot - 1 - n_eq_lt;
          const int r_b = (bin == cut_bin) ? r_cut : bn;
          const int sel_rank = n_gt + n_eq_gt;  // (score desc, row desc): proposal.py:123-125 with the documented tie rule
          selected = sel_rank < r_b;
          const int n_sel_run = min(n_eq_tot, r_b - n_gt);
          pos = bs + n_gt + n_eq_lt - (n_eq_tot - n_sel_run);  // (score desc, row asc): nms.py:19 re-sorts the list
          kpos = (int)s_base[bin] + sel_rank;
        }
        FZ_CLK(4);
        if (selected) {
          const Box32 bx = decode32(g, dx, dy, dw, dh, ah, aw, aa, lim_x, lim_y, c32);
          s_cbox[pos] = make_float4(bx.x1, bx.y1, bx.x2p, bx.y2p);
          s_caux[pos] = make_float2(bx.a, screen_ok ? bx.t : __int_as_float(0x7f800000));
          s_p[pos] = p;
          s_kpos[pos] = kpos;
          s_skey[pos] = (u32)(c >> 32);
        }
      }
      FZ_CLK(5);
      __syncthreads();
      FZ_CLK(6);
    }
    if (stripes == 0) FZ_STAMP(10);
    // ---- IoU bit blocks of 32 x 32 pairs: in-stripe lower triangle + (kept so far) x stripe ----
    const int par = stripes & 1;
    u32* Mcur = s_M + par * (kWMax * kMWords);
    u32* Rcur = s_removed + par * 16;
    {
      // The output row of a stripe position (proposal.py:140-147: the reference's float64 box, cast to float32) is
      // computed by the CTA that owns the position's group of 32 (group % cluster size) while the bit blocks are
      // evaluated: the highest warps of a CTA have the fewest blocks.  After the stripe is resolved the owner only
      // copies the rows of the kept positions out.
      {
        const int lg = 31 - warp;               // this warp's local group: positions (lg * csize + rank) * 32 + lane
        const int pos = (lg * csize + rank) * 32 + lane;
        if (lg < kWMax / 64 && pos < n_s && g.o.rois) {
          double bx[4];
          exact_box(g, deltas, im_h, im_w, s_p[pos], bx);
          s_roi[lg * 32 + lane] = make_float4((float)bx[0], (float)bx[1], (float)bx[2], (float)bx[3]);
        }
      }
      const int n_cg = (n_s + 31) >> 5, n_rg = (nk + 31) >> 5;
      const int n_tri = n_cg * (n_cg + 1) / 2;
      const int n_tasks = n_tri + n_rg * n_cg;
#pragma unroll 1
      for (int u = warp * csize + rank; u < n_tasks; u += 32 * csize) {
        const bool tri = u < n_tri;
        int cg, ri;      // column group; row = stripe position (tri) or index into the kept list (cross)
        u32 mask = 0xffffffffu;
        Box32 r;
        if (tri) {
          int rb = (int)((sqrtf(8.0f * (float)u + 1.0f) - 1.0f) * 0.5f);
          while (rb * (rb + 1) / 2 > u) --rb;
          while ((rb + 1) * (rb + 2) / 2 <= u) ++rb;
          cg = u - rb * (rb + 1) / 2;
          ri = rb * 32 + lane;
          if (cg == rb) mask = (1u << lane) - 1u;  // diagonal block: strictly below the diagonal
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
        if (g.pair_count) {
          pairs += __popc(mask);
          n_und += __popc(und);
        }
        if (und) {  // rare: the reference's float64 arithmetic decides
          const int p_row = tri ? s_p[ri] : __float_as_int(__ldcg(kaux + ri).z);
          do {
            const int j = __ffs(und) - 1;
            und &= und - 1u;
            if (exact_pair(g, deltas, im_h, im_w, p_row, s_p[cg * 32 + j])) supp |= 1u << j;
            else supp &= ~(1u << j);
          } while (und);
        }
        if (tri) {
          if (ri < n_s) {  // every CTA resolves the stripe itself: the word goes into everybody's bit rows
            u32* wp = Mcur + cg * kWMax + ri;
#pragma unroll 1
            for (int t = 0; t < csize; ++t) *cluster.map_shared_rank(wp, t) = supp;
          }
        } else {
          supp = __reduce_or_sync(0xffffffffu, supp);
          if (lane == 0 && supp) atomicOr(&s_rem[cg], supp);
        }
      }
      FZ_CLK(7);
      __syncthreads();
      FZ_CLK(8);
      if (tid < 16 * csize) {
        const u32 v = s_rem[tid & 15];
        if (v) atomicOr(cluster.map_shared_rank(Rcur + (tid & 15), tid >> 4), v);
      }
    }
    if (stripes == 0) FZ_STAMP(11);
    // ONE barrier per stripe: the bit rows and removed words are complete in every CTA's shared memory.  The next
    // stripe's words go to the other parity's buffers, so nobody overwrites rows that a slower CTA is still reading.
    cluster.sync();
    if (stripes == 0) FZ_STAMP(12);
    FZ_CLK(9);
    // ---- every CTA resolves the stripe: one fixed point of  keep_i = alive_i and no kept j < i suppresses i ----
    {
      if (tid < 16) {
        const int left = n_s - tid * 32;
        const u32 valid = left >= 32 ? 0xffffffffu : (left > 0 ? (1u << left) - 1u : 0u);
        s_keep[tid] = ~Rcur[tid] & valid;
        Rcur[tid] = 0u;  // written again two stripes from now, after two more barriers
      }
      __syncthreads();
      FZ_CLK(10);
      // Eight warps resolve the stripe, two bit rows per thread in registers (rows tid and tid + 256; words above the
      // diagonal are never written and read as zero).  A round costs every participating warp ~50 instructions, so
      // fewer warps with more rows each make a shorter round than one row per thread on sixteen warps.
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
          if (it < 8) FZ_CLK(12 + it);
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
          if (g.dbg && rank == 0 && tid == 0 && stripes == 0) g.dbg[(size_t)b * 16 + 15] = (u64)(it + 1);
        }
        if (lane == 0) {  // the last round changed nothing: its ballots are the result
          s_kfin[warp] = bal_a;
          s_kfin[8 + warp] = bal_b;
        }
      }
      __syncthreads();
      FZ_CLK(20);
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
        if (!f) {
          // rows of the later stripes' (kept so far) x stripe blocks.  Every CTA writes the SAME values to the same
          // addresses and reads them back only after its own __syncthreads: no barrier between CTAs is needed.
          kbox[idx] = s_cbox[tid];
          const float2 a = s_caux[tid];
          kaux[idx] = make_float4(a.x, a.y, __int_as_float(s_p[tid]), 0.0f);
        }
        const int lg = warp / csize;
        if (warp - lg * csize == rank) {  // this CTA owns the position's group: results (proposal.py:140-149)
          const NmsOutputs& o = g.o;
          const size_t orow = (size_t)b * o.rows + idx;
          if (o.rois) {
            const float4 bx = s_roi[lg * 32 + lane];
            float* ro = o.rois + orow * 5;
            ro[0] = (float)b;
            ro[1] = bx.x;
            ro[2] = bx.y;
            ro[3] = bx.z;
            ro[4] = bx.w;
          }
          if (o.scores) {  // the staged key is an exact image of the float32 score (-0.0 reads back as +0.0)
            const u32 key = s_skey[tid];
            o.scores[orow] = __uint_as_float((key & 0x80000000u) ? (key & 0x7fffffffu) : ~key);
          }
          if (o.keep_orig) o.keep_orig[orow] = s_p[tid];
          if (o.keep_pos) {
            int kpos = s_kpos[tid];
            if (kpos < 0) {  // bucket ordered in place: undo the tie-run reversal of nms.py:19 by looking at the neighbours
              const int sl = -1 - kpos;
              const u32 key = s_skey[tid];
              const int bin = bin_of_key(key);
              const int lo = (int)s_base[bin], hi = lo + ((bin == cut_bin) ? r_cut : (int)s_cnt[bin]);
              int rs_ = sl, re_ = sl + 1;
              while (rs_ > lo && (u32)(__ldcg(comp + rs_ - 1) >> 32) == key) --rs_;
              while (re_ < hi && (u32)(__ldcg(comp + re_) >> 32) == key) ++re_;
              kpos = rs_ + (re_ - 1 - sl);
            }
            o.keep_pos[orow] = kpos;
          }
        }
      }
      nk = nk2;
      fin = f;
    }
    if (stripes == 0) FZ_STAMP(13);
    FZ_CLK(22);
    ++stripes;
    c0 = presorted ? (c1 >= ps_end ? (int)s_end[ps_bin] : c1) : c1;
    if (presorted && ps_bin == cut_bin && c1 >= ps_end) break;
  }
  FZ_STAMP(7);
  // ---- the kept rows were written stripe by stripe; what is left: empty rows, counters ----
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
    const NmsOutputs& o = g.o;
    const size_t orow = (size_t)b * o.rows;
    for (int j = nk + tid; j < o.rows; j += blockDim.x) {
      if (o.rois) {
        float* ro = o.rois + (orow + j) * 5;
        ro[0] = ro[1] = ro[2] = ro[3] = ro[4] = 0.f;
      }
      if (o.scores) o.scores[orow + j] = 0.f;
      if (o.keep_orig) o.keep_orig[orow + j] = -1;
      if (o.keep_pos) o.keep_pos[orow + j] = -1;
    }
    if (tid == 0 && o.nkeep) {
      o.nkeep[b * 2] = nk;
      o.nkeep[b * 2 + 1] = nan_flag;
    }
  }
  if (g.pair_count) {
    pairs = __reduce_add_sync(0xffffffffu, pairs);
    n_und = __reduce_add_sync(0xffffffffu, n_und);
    if (lane == 0 && pairs) atomicAdd(g.pair_count + 2 * b, pairs);
    if (lane == 0 && n_und) atomicAdd(g.pair_count + 2 * b + 1, n_und);
  }
  // Nobody may leave while its shared memory can still be read by another CTA: the coarse tables are pulled right
  // after the decode barrier, and every later remote access is followed by a barrier of the walk.
  if (windows == 0) cluster.sync();
  FZ_STAMP(8);
  if (g.dbg && rank == 0 && tid == 0) {
    g.dbg[(size_t)b * 16 + 9] = (u64)stripes;
    g.dbg[(size_t)b * 16 + 14] = (u64)windows;
  }
}

static inline size_t fused_smem_bytes(int items) {
  const int items2 = (items + 3) / 4 * 4;
  return (size_t)kWMax * (sizeof(float4) + sizeof(float2) + sizeof(u64) + 3 * sizeof(int)) + (size_t)(kWMax / 2) * sizeof(float4) +
         (16 + 32) * sizeof(u32) + (size_t)2 * kWMax * kMWords * sizeof(u32) + (size_t)items2 * 1024 * sizeof(u32) + 64;
}

}  // namespace fz
}  // namespace mxr

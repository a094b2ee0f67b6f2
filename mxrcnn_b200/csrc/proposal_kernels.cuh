// RPN proposal kernels (rcnn/rpn/proposal.py:39-149).
//   K1 decode_filter_kernel : anchors x deltas -> clip -> min-size filter -> sortable score key
//   K2 topk_sort_kernel     : radix select of the top pre_nms boxes, ordered, boxes re-decoded
//                             into NMS processing order (one CTA per image)
// NMS itself is in nms_kernels.cuh.
#pragma once
#include "device_math.cuh"
#include "sort.cuh"

namespace mxr {

struct ProposalGeom {
  DecodeGeom g;
  int n_crop;    // A*hc*wc  proposals enumerated
  int n_full;    // A*H*W
  int cropped;   // hc<H or wc<W
  size_t cls_img_stride;  // floats between consecutive images of cls_prob
  size_t cls_fg_offset;   // floats from an image's start to its foreground planes (proposal.py:57)
};

// Key layout: keys[b][q], q = a*(hc*wc) + h*wc + w (anchor-major, the NCHW order of the
// inputs, so loads AND stores are coalesced).  The reference's row index of the same
// proposal is p = (h*wc + w)*A + a  (proposal.py:88-107).
__device__ __forceinline__ int q_to_p(int q, int hwc, int A) { return (q % hwc) * A + q / hwc; }

// grid (ceil(max(n_crop,n_full)/256), B)
__global__ void __launch_bounds__(256) decode_filter_kernel(ProposalGeom pg, AnchorTable at,
                                                            const float* __restrict__ cls_prob,
                                                            const float* __restrict__ bbox_pred,
                                                            const float* __restrict__ im_info, float min_size,
                                                            u32* __restrict__ keys, Ctrl* __restrict__ ctrl) {
  const int b = blockIdx.y;
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  const DecodeGeom& g = pg.g;
  const size_t hw = (size_t)g.H * g.W;
  const float* fg = cls_prob + (size_t)b * pg.cls_img_stride + pg.cls_fg_offset;  // proposal.py:57
  const float* deltas = bbox_pred + (size_t)b * 4 * g.A * hw;
  const float im_h = __ldg(im_info + b * 3), im_w = __ldg(im_info + b * 3 + 1), im_s = __ldg(im_info + b * 3 + 2);
  bool nan = false;
  bool valid = false;
  if (q < pg.n_crop) {
    const int hwc = g.hc * g.wc;
    const int a = q / hwc, r = q - a * hwc;
    const int h = r / g.wc, w = r - h * g.wc;
    double bx[4];
    bool dnan;
    decode_clip(at, g, deltas, h, w, a, im_h, im_w, bx, &dnan);
    // score of proposal p is read at flat index p of the UNCROPPED (H,W,A) flattening (quirk Q1,
    // proposal.py:107 does not crop the scores); without a crop this is (a,h,w) itself.
    float s;
    if (!pg.cropped) {
      s = __ldg(fg + (size_t)a * hw + (size_t)h * g.W + w);
    } else {
      const int p = r * g.A + a;
      const int a2 = p % g.A, w2 = (p / g.A) % g.W, h2 = p / (g.A * g.W);
      s = __ldg(fg + (size_t)a2 * hw + (size_t)h2 * g.W + w2);
    }
    valid = passes_min_size(bx, min_size, im_s);
    nan = (s != s) || dnan;
    keys[(size_t)b * pg.n_crop + q] = valid ? score_key32(s) : 0u;
  }
  if (pg.cropped && q < pg.n_full) {
    // proposal.py:58-62 checks the WHOLE fg-score and delta arrays, cropped or not
    const float v0 = __ldg(fg + q);
    const float v1 = __ldg(deltas + q), v2 = __ldg(deltas + q + pg.n_full);
    const float v3 = __ldg(deltas + q + 2 * (size_t)pg.n_full), v4 = __ldg(deltas + q + 3 * (size_t)pg.n_full);
    nan = nan || (v0 != v0) || (v1 != v1) || (v2 != v2) || (v3 != v3) || (v4 != v4);
  }
  const u32 vmask = __ballot_sync(0xffffffffu, valid);
  const u32 nmask = __ballot_sync(0xffffffffu, nan);
  if ((threadIdx.x & 31) == 0) {
    if (vmask) atomicAdd(&ctrl[b].n_valid, __popc(vmask));
    if (nmask) atomicOr(&ctrl[b].nan_flag, 1);
  }
}

struct TopkArrays {
  const u32* keys;      // [b][n_crop]
  double4* box;         // [b][cap]   processing order
  double* area;
  float* score;
  int* orig;            // p index
  int cap;
  u64* g_sort_keys;     // global fallback for the sort buffer (n_pow2 > smem capacity), [b][g_sort_cap]
  int g_sort_cap;
};

struct TopkCall {
  ProposalGeom pg;
  TopkArrays ta;
  const float* cls_prob;
  const float* bbox_pred;
  const float* im_info;
  int pre_nms;        // proposal.py:124-125 (<= 0: all)
  int smem_pairs;     // u64 slots available in shared memory for the sort buffer (0: use the global one)
};

// Body shared by the stand-alone kernel and the cooperative NMS kernel's deferred ("tier 2") phase.
// One CTA (blockDim.x threads, all of them) per image:
//   1. radix-select the threshold of the top-k_sel boxes under (score desc, row index desc)  -- proposal.py:123-125
//   2. compact them, sort by (score desc, row index ASC)  -- nms.py:19 re-sorts the already
//      descending list, which reverses runs of equal scores
//   3. re-decode the selected boxes into processing order.
// k_limit > 0 selects only the first min(k_limit, k) boxes of the order ("head"): greedy NMS walks the list
// in order and usually fills its post-NMS quota long before the tail is needed.
// ITEMS > 0: the image's keys live in registers (ITEMS per thread, n_crop <= ITEMS*blockDim.x).
// smem: max(8 * n_pow2, 32 KiB) bytes; the select's histograms alias the sort buffer.
template <int ITEMS>
__device__ __forceinline__ void topk_body(const TopkCall& tc, const AnchorTable& at, int k_limit, Ctrl* __restrict__ ctrl, int b,
                                          unsigned char* smem_raw, u32* st, int* s_cnt) {
  const int tid = threadIdx.x;
  const ProposalGeom& pg = tc.pg;
  const DecodeGeom& g = pg.g;
  const TopkArrays& ta = tc.ta;
  const int hwc = g.hc * g.wc;
  const u32* keys = ta.keys + (size_t)b * pg.n_crop;
  const int n_valid = ctrl[b].n_valid;
  const int k = (tc.pre_nms > 0) ? min(tc.pre_nms, n_valid) : n_valid;  // proposal.py:124-125
  const int k_sel = (k_limit > 0 && k_limit < k) ? k_limit : k;
  if (tid == 0) {
    ctrl[b].n = k;
    *s_cnt = 0;
  }
  if (k_sel == 0) return;
  const int n_pow2 = max(2, (int)(1u << (32 - __clz(k_sel - 1))));
  u64* skeys = (n_pow2 <= tc.smem_pairs) ? reinterpret_cast<u64*>(smem_raw) : ta.g_sort_keys + (size_t)b * ta.g_sort_cap;
  u32* hist = reinterpret_cast<u32*>(smem_raw);
  const int n_iter = (pg.n_crop + blockDim.x - 1) / blockDim.x;
  u32 rk[ITEMS > 0 ? ITEMS : 1];
  if (ITEMS > 0) {
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
      const int q = tid + i * blockDim.x;
      rk[i] = (q < pg.n_crop) ? __ldg(keys + q) : 0u;
    }
  }
  auto key_of = [&](int i, int q) -> u32 {
    if (ITEMS > 0) return rk[i];
    return (q < pg.n_crop) ? __ldg(keys + q) : 0u;
  };
  u64 T = 1;  // every valid composite is >= 2^32
  if (k_sel < n_valid) {
    T = block_radix_select<ITEMS>(n_iter, k_sel, hist, st, [&](int i, int q, u64* c) {
      const u32 key = key_of(i, q);
      *c = ((u64)key << 32) | (u32)q_to_p(q, hwc, g.A);
      return key != 0u;
    });
  }
  __syncthreads();
  auto compact = [&](int i, int q) {
    const u32 key = key_of(i, q);
    const u32 p = (u32)q_to_p(q, hwc, g.A);
    const bool sel = key != 0u && (((u64)key << 32) | p) >= T;
    const u32 m = __ballot_sync(0xffffffffu, sel);
    int base = 0;
    if ((tid & 31) == 0 && m) base = atomicAdd(s_cnt, __popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (sel) skeys[base + __popc(m & ((1u << (tid & 31)) - 1u))] = ((u64)key << 32) | (u32)(~p);
  };
  if (ITEMS > 0) {
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) compact(i, tid + i * blockDim.x);
  } else {
    for (int i = 0; i < n_iter; ++i) compact(i, tid + i * blockDim.x);
  }
  for (int i = k_sel + tid; i < n_pow2; i += blockDim.x) skeys[i] = 0;
  __syncthreads();
  bitonic_sort_desc_u64(skeys, n_pow2);
  // ---- gather: re-decode the selected boxes into processing order ----
  const size_t hw = (size_t)g.H * g.W;
  const float* fg = tc.cls_prob + (size_t)b * pg.cls_img_stride + pg.cls_fg_offset;
  const float* deltas = tc.bbox_pred + (size_t)b * 4 * g.A * hw;
  const float im_h = __ldg(tc.im_info + b * 3), im_w = __ldg(tc.im_info + b * 3 + 1);
  for (int i = tid; i < k_sel; i += blockDim.x) {
    const int p = (int)(~(u32)skeys[i]);
    const int a = p % g.A, r = p / g.A;
    const int h = r / g.wc, w = r - h * g.wc;
    double bx[4];
    decode_clip(at, g, deltas, h, w, a, im_h, im_w, bx);
    const int a2 = p % g.A, w2 = (p / g.A) % g.W, h2 = p / (g.A * g.W);
    const float s = __ldg(fg + (size_t)a2 * hw + (size_t)h2 * g.W + w2);
    const size_t o = (size_t)b * ta.cap + i;
    ta.box[o] = make_double4(bx[0], bx[1], bx[2], bx[3]);
    ta.area[o] = box_area<double>(bx[0], bx[1], bx[2], bx[3]);
    ta.score[o] = s;
    ta.orig[o] = p;
  }
}

// grid B, block 1024.  gated != 0: skip images that already finished NMS.
template <int ITEMS>
__global__ void __launch_bounds__(1024) topk_kernel(TopkCall tc, AnchorTable at, int k_limit, int gated,
                                                    Ctrl* __restrict__ ctrl) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ u32 st[8];
  __shared__ int s_cnt;
  const int b = blockIdx.x;
  if (gated && ctrl[b].done) return;
  topk_body<ITEMS>(tc, at, k_limit, ctrl, b, smem_raw, st, &s_cnt);
}

}  // namespace mxr

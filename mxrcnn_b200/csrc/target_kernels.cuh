// Kernels of the training-target generators:
//   assign_anchor         rcnn/minibatch.py:204-395          (anchor_* kernels: phase 1 of maps too large for
//                                                             anchor_fused.cuh, which holds the batched kernels)
//   _sample_rois          rcnn/rpn/proposal_target.py:135-196 (fg_bg_lists_kernel, proposal_target_gather_kernel)
//   sample_rois           rcnn/minibatch.py:151-201           (fg_bg_lists_kernel)
// Random draws never happen here: the host layer draws from numpy's RNG in the reference's call
// order and hands over ranks / indices.
#pragma once
#include "box_kernels.cuh"

namespace mxr {

// Block-wide ORDERED compaction step: every thread passes a flag, gets its exclusive rank within
// the CTA (in thread order) and the CTA total.  s_warp: >= 33 ints of shared memory.
__device__ __forceinline__ int block_exclusive_rank(bool flag, int* s_warp, int* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const unsigned m = __ballot_sync(0xffffffffu, flag);
  const int in_warp = __popc(m & ((1u << lane) - 1u));
  __syncthreads();
  if (lane == 0) s_warp[warp] = __popc(m);
  __syncthreads();
  if (warp == 0) {
    int v = lane < nwarps ? s_warp[lane] : 0;
    int incl = v;
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += u;
    }
    s_warp[lane] = incl - v;
    if (lane == 31) s_warp[32] = incl;
  }
  __syncthreads();
  *total = s_warp[32];
  return s_warp[warp] + in_warp;
}

struct AnchorTargetGeom {
  int fh, fw, A, N;      // feature map, anchors per cell, A*fh*fw
  int feat_stride;       // python int (minibatch.py:260-261)
  int allowed_border;
  float im_h, im_w;      // im_info[0][0], im_info[0][1]
  int G;
};

struct AnchorTargetWs {
  int* inside_rank;      // [N]  rank among inside anchors or -1
  double* anchors_in;    // [N][4] inside anchors, rank order
  double* gt64;          // [G][4]  gt boxes cast to float64 (minibatch.py:295)
  int* argmax;           // [N]
  double* maxov;         // [N]
  double* colmax;        // [G]
  float* labels;         // [N] rank-indexed, float32 like the reference (:289)
  int* fg_list;          // [N] ranks with label 1, ascending
  int* bg_list;          // [N] ranks with label 0, ascending
  int* counts;           // [4] n_inside, n_fg, n_bg, -
};

// minibatch.py:260-278.  One CTA walks all N anchors in (h,w,a) order and compacts the inside ones.
static __global__ void __launch_bounds__(1024) anchor_inside_kernel(AnchorTargetGeom g, AnchorTable at, const float* __restrict__ gt,
                                                             AnchorTargetWs w) {
  __shared__ int s_warp[33];
  typedef Ar<double> A;
  for (int t = threadIdx.x; t < g.G * 4; t += blockDim.x) w.gt64[t] = (double)gt[(t >> 2) * 5 + (t & 3)];
  // im_info[k] + allowed_border is a float32 sum (NEP 50), compared against the float64 anchors
  const double lim_w = (double)__fadd_rn(g.im_w, (float)g.allowed_border);
  const double lim_h = (double)__fadd_rn(g.im_h, (float)g.allowed_border);
  const double lo = -(double)g.allowed_border;
  int base = 0;
  for (int p0 = 0; p0 < g.N; p0 += blockDim.x) {
    const int p = p0 + threadIdx.x;
    bool inside = false;
    double b[4] = {0, 0, 0, 0};
    if (p < g.N) {
      const int a = p % g.A, cell = p / g.A;
      const int wx = cell % g.fw, hy = cell / g.fw;
      const double sx = (double)((long long)wx * g.feat_stride), sy = (double)((long long)hy * g.feat_stride);
      b[0] = A::add(at.a[a][0], sx);
      b[1] = A::add(at.a[a][1], sy);
      b[2] = A::add(at.a[a][2], sx);
      b[3] = A::add(at.a[a][3], sy);
      inside = (b[0] >= lo) && (b[1] >= lo) && (b[2] < lim_w) && (b[3] < lim_h);
    }
    int total;
    const int r = block_exclusive_rank(inside, s_warp, &total);
    if (p < g.N) {
      w.inside_rank[p] = inside ? base + r : -1;
      if (inside) {
        double* o = w.anchors_in + (size_t)(base + r) * 4;
        o[0] = b[0];
        o[1] = b[1];
        o[2] = b[2];
        o[3] = b[3];
      }
    }
    base += total;
  }
  if (threadIdx.x == 0) w.counts[0] = base;
}

// bbox_overlaps + reductions over the inside anchors, row count read from the device.
static __global__ void __launch_bounds__(128) anchor_overlaps_kernel(AnchorTargetGeom g, AnchorTargetWs w) {
  __shared__ double sq[kOvChunk][5];
  const int n = w.counts[0];
  if ((int)(blockIdx.x * blockDim.x) >= n) return;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = i < n;
  double b[4] = {0, 0, 0, 0};
  if (live)
    for (int k = 0; k < 4; ++k) b[k] = w.anchors_in[(size_t)i * 4 + k];
  double best = 0.0;
  int arg = 0;
  for (int q0 = 0; q0 < g.G; q0 += kOvChunk) {
    const int qn = min(kOvChunk, g.G - q0);
    __syncthreads();
    for (int t = threadIdx.x; t < qn; t += blockDim.x) {
      const double* q = w.gt64 + (size_t)(q0 + t) * 4;
      sq[t][0] = q[0];
      sq[t][1] = q[1];
      sq[t][2] = q[2];
      sq[t][3] = q[3];
      sq[t][4] = Ar<double>::mul(Ar<double>::add(Ar<double>::sub(q[2], q[0]), 1.0), Ar<double>::add(Ar<double>::sub(q[3], q[1]), 1.0));
    }
    __syncthreads();
    for (int j = 0; j < qn; ++j) {
      const double ov = live ? overlap_one(b[0], b[1], b[2], b[3], sq[j][0], sq[j][1], sq[j][2], sq[j][3], sq[j][4]) : 0.0;
      if (live && ((q0 + j == 0) || ov > best)) {
        best = ov;
        arg = q0 + j;
      }
      unsigned long long bits = (unsigned long long)__double_as_longlong(ov);
      for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, bits, o);
        bits = other > bits ? other : bits;
      }
      if ((threadIdx.x & 31) == 0 && bits != 0ull) atomicMax(reinterpret_cast<unsigned long long*>(w.colmax) + q0 + j, bits);
    }
  }
  if (live) {
    w.argmax[i] = arg;
    w.maxov[i] = best;
  }
}

// minibatch.py:289-316.  Recomputes each row of overlaps to evaluate `overlaps == gt_max_overlaps`
// (quirk Q4: every anchor tying a column maximum -- including a maximum of 0 -- becomes positive).
static __global__ void __launch_bounds__(128) anchor_label_kernel(AnchorTargetGeom g, AnchorTargetWs w, double neg_ov, double pos_ov,
                                                           int clobber) {
  __shared__ double sq[kOvChunk][6];
  const int n = w.counts[0];
  if ((int)(blockIdx.x * blockDim.x) >= n) return;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = i < n;
  double b[4] = {0, 0, 0, 0};
  if (live)
    for (int k = 0; k < 4; ++k) b[k] = w.anchors_in[(size_t)i * 4 + k];
  bool is_gt_argmax = false;
  for (int q0 = 0; q0 < g.G; q0 += kOvChunk) {
    const int qn = min(kOvChunk, g.G - q0);
    __syncthreads();
    for (int t = threadIdx.x; t < qn; t += blockDim.x) {
      const double* q = w.gt64 + (size_t)(q0 + t) * 4;
      sq[t][0] = q[0];
      sq[t][1] = q[1];
      sq[t][2] = q[2];
      sq[t][3] = q[3];
      sq[t][4] = Ar<double>::mul(Ar<double>::add(Ar<double>::sub(q[2], q[0]), 1.0), Ar<double>::add(Ar<double>::sub(q[3], q[1]), 1.0));
      sq[t][5] = w.colmax[q0 + t];
    }
    __syncthreads();
    if (live)
      for (int j = 0; j < qn; ++j)
        is_gt_argmax = is_gt_argmax ||
                       (overlap_one(b[0], b[1], b[2], b[3], sq[j][0], sq[j][1], sq[j][2], sq[j][3], sq[j][4]) == sq[j][5]);
  }
  if (!live) return;
  float label = -1.f;
  if (g.G > 0) {
    const double mo = w.maxov[i];
    if (!clobber && mo < neg_ov) label = 0.f;   // :302-304
    if (is_gt_argmax) label = 1.f;              // :307
    if (mo >= pos_ov) label = 1.f;              // :310
    if (clobber && mo < neg_ov) label = 0.f;    // :312-314
  } else {
    label = 0.f;                                // :316
  }
  w.labels[i] = label;
}

// Ordered lists of the fg (label 1) and bg (label 0) ranks: np.where(labels == 1)[0] (:320, :329).
static __global__ void __launch_bounds__(1024) anchor_lists_kernel(AnchorTargetWs w) {
  __shared__ int s_warp[33];
  const int n = w.counts[0];
  int nf = 0, nb = 0;
  for (int i0 = 0; i0 < n; i0 += blockDim.x) {
    const int i = i0 + threadIdx.x;
    const float l = i < n ? w.labels[i] : -1.f;
    int total;
    int r = block_exclusive_rank(l == 1.f, s_warp, &total);
    if (l == 1.f) w.fg_list[nf + r] = i;
    nf += total;
    r = block_exclusive_rank(l == 0.f, s_warp, &total);
    if (l == 0.f) w.bg_list[nb + r] = i;
    nb += total;
  }
  if (threadIdx.x == 0) {
    w.counts[1] = nf;
    w.counts[2] = nb;
  }
}

// np.where(ov >= fg_thresh)[0] and np.where((ov < bg_hi) & (ov >= bg_lo))[0], both ascending
// (proposal_target.py:148,160-161; minibatch.py:166,174), comparisons in the dtype of `ov`.
// counts: [n_fg, n_bg].
template <typename T>
static __global__ void __launch_bounds__(1024) fg_bg_lists_kernel(const T* __restrict__ ov, int n, T fg_thresh, T bg_hi, T bg_lo,
                                                           int* __restrict__ fg_list, int* __restrict__ bg_list,
                                                           int* __restrict__ counts) {
  __shared__ int s_warp[33];
  int nf = 0, nb = 0;
  for (int i0 = 0; i0 < n; i0 += blockDim.x) {
    const int i = i0 + threadIdx.x;
    const bool live = i < n;
    const T v = live ? ov[i] : T(0);
    const bool fg = live && (v >= fg_thresh);
    const bool bg = live && (v < bg_hi) && (v >= bg_lo);
    int total;
    int r = block_exclusive_rank(fg, s_warp, &total);
    if (fg) fg_list[nf + r] = i;
    nf += total;
    r = block_exclusive_rank(bg, s_warp, &total);
    if (bg) bg_list[nb + r] = i;
    nb += total;
  }
  if (threadIdx.x == 0) {
    counts[0] = nf;
    counts[1] = nb;
  }
}

// proposal_target.py:140-166 in ONE launch (one CTA): IoU of every roi with every gt box in float64, first-maximum
// argmax and max per roi (:143-144), and the three ordered index lists the sampler draws from: fg (max >= FG_THRESH,
// :148), bg (BG_THRESH_LO <= max < BG_THRESH_HI, :160-161) and the TRAIN-mode fallback bg (0 <= max < HI + 0.2, :162-166).
// rois (R,5) f32 [batch index, x1, y1, x2, y2], gt (G,5) f32.  lists: [3][R] i32; counts: [3].
static const int kPtChunk = 256;
static __global__ void __launch_bounds__(1024) proposal_target_phase1_kernel(const float* __restrict__ rois, int R,
                                                                      const float* __restrict__ gt, int G, double fg_thresh,
                                                                      double bg_hi, double bg_lo, double fb_hi, double fb_lo,
                                                                      int* __restrict__ assignment, double* __restrict__ maxov,
                                                                      int* __restrict__ lists, int* __restrict__ counts) {
  __shared__ double sq[kPtChunk][5];
  __shared__ int s_warp[33];
  int nf = 0, nb = 0, nfb = 0;
  for (int i0 = 0; i0 < R; i0 += blockDim.x) {
    const int i = i0 + threadIdx.x;
    const bool live = i < R;
    double bx1 = 0, by1 = 0, bx2 = 0, by2 = 0;
    if (live) {  // all_rois[:, 1:5].astype(np.float), :141
      bx1 = (double)rois[(size_t)i * 5 + 1];
      by1 = (double)rois[(size_t)i * 5 + 2];
      bx2 = (double)rois[(size_t)i * 5 + 3];
      by2 = (double)rois[(size_t)i * 5 + 4];
    }
    double best = 0.0;
    int arg = 0;
    for (int q0 = 0; q0 < G; q0 += kPtChunk) {
      const int qn = min(kPtChunk, G - q0);
      __syncthreads();
      for (int t = threadIdx.x; t < qn; t += blockDim.x) {
        const float* q = gt + (size_t)(q0 + t) * 5;
        const double x1 = (double)q[0], y1 = (double)q[1], x2 = (double)q[2], y2 = (double)q[3];
        sq[t][0] = x1;
        sq[t][1] = y1;
        sq[t][2] = x2;
        sq[t][3] = y2;
        sq[t][4] = Ar<double>::mul(Ar<double>::add(Ar<double>::sub(x2, x1), 1.0), Ar<double>::add(Ar<double>::sub(y2, y1), 1.0));
      }
      __syncthreads();
      if (live) {
        for (int j = 0; j < qn; ++j) {
          const double ov = overlap_one(bx1, by1, bx2, by2, sq[j][0], sq[j][1], sq[j][2], sq[j][3], sq[j][4]);
          if ((q0 + j == 0) || ov > best) {
            best = ov;
            arg = q0 + j;
          }
        }
      }
    }
    if (live) {
      assignment[i] = arg;
      maxov[i] = best;
    }
    const bool fg = live && (best >= fg_thresh);
    const bool bg = live && (best < bg_hi) && (best >= bg_lo);
    const bool fb = live && (best < fb_hi) && (best >= fb_lo);
    int total;
    int r = block_exclusive_rank(fg, s_warp, &total);
    if (fg) lists[nf + r] = i;
    nf += total;
    r = block_exclusive_rank(bg, s_warp, &total);
    if (bg) lists[R + nb + r] = i;
    nb += total;
    r = block_exclusive_rank(fb, s_warp, &total);
    if (fb) lists[2 * R + nfb + r] = i;
    nfb += total;
  }
  if (threadIdx.x == 0) {
    counts[0] = nf;
    counts[1] = nb;
    counts[2] = nfb;
  }
}

struct PtGather {
  const float* rois;       // (R,5)
  const float* gt;         // (G,5)
  const int* assignment;   // (R) argmax gt per roi
  const int* keep;         // (K) sampled roi indices
  int K, fg_this, num_classes, normalize;
  double mean[4], stdv[4];
  float inside_w[4];
  float* out_rois;         // (K,5)
  float* labels;           // (K)
  float* targets;          // (K,4C)
  float* inside;           // (K,4C)
  float* outside;          // (K,4C)
};

// proposal_target.py:183-194 + _compute_targets :120-133 + expand_bbox_regression_targets.
// One thread per (sampled roi, class slot).
static __global__ void __launch_bounds__(256) proposal_target_gather_kernel(PtGather a) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long long)a.K * a.num_classes) return;
  const int k = (int)(t / a.num_classes), c = (int)(t - (long long)k * a.num_classes);
  const int idx = a.keep[k];
  const float* roi = a.rois + (size_t)idx * 5;
  const float* g = a.gt + (size_t)a.assignment[idx] * 5;
  const float lab = (k < a.fg_this) ? g[4] : 0.f;   // labels[fg_rois_per_this_image:] = 0 (:187, quirk Q3)
  if (c == 0) {
    for (int j = 0; j < 5; ++j) a.out_rois[(size_t)k * 5 + j] = roi[j];
    a.labels[k] = lab;
  }
  float tv[4];
  {
    float o[4];
    bbox_transform_one<float, float, float>(roi[1], roi[2], roi[3], roi[4], g[0], g[1], g[2], g[3], o);  // f32 rois, f32 gt
    for (int j = 0; j < 4; ++j)
      tv[j] = a.normalize ? (float)Ar<double>::div(Ar<double>::sub((double)o[j], a.mean[j]), a.stdv[j]) : o[j];  // :129-131
  }
  const int start = lab > 0.f ? (int)(4.f * lab) : -1000;
  for (int e = 0; e < 4; ++e) {
    const int col = 4 * c + e, off = col - start;
    const bool hit = off >= 0 && off < 4;
    const size_t o = (size_t)k * 4 * a.num_classes + col;
    const float iw = hit ? a.inside_w[off] : 0.f;
    a.targets[o] = hit ? tv[off] : 0.f;
    a.inside[o] = iw;
    a.outside[o] = iw > 0.f ? 1.f : 0.f;   // np.array(bbox_inside_weights > 0).astype(float32), :76
  }
}

}  // namespace mxr

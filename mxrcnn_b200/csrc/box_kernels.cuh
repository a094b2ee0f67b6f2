// Kernels of the box helpers (helper/processing/bbox_transform.py, bbox_regression.py).
#pragma once
#include "device_math.cuh"

namespace mxr {

template <typename T>
__device__ __forceinline__ T exp_as_numpy(T v);
template <>
__device__ __forceinline__ float exp_as_numpy<float>(float v) { return expf(v); }
template <>
__device__ __forceinline__ double exp_as_numpy<double>(double v) { return exp(v); }

template <typename T>
__device__ __forceinline__ T log_as_numpy(T v);
template <>
__device__ __forceinline__ float log_as_numpy<float>(float v) { return logf(v); }
template <>
__device__ __forceinline__ double log_as_numpy<double>(double v) { return log(v); }

// helper/processing/bbox_transform.py:35-78.  One thread per (row, class).  Boxes are cast to
// float64 (:46); exp is evaluated in the dtype of the deltas (:67-68), everything else in f64.
template <typename BT, typename DT>
__global__ void __launch_bounds__(256) bbox_pred_kernel(const BT* __restrict__ boxes, const DT* __restrict__ deltas, int n,
                                                        int c4, double* __restrict__ out) {
  typedef Ar<double> A;
  const int nc = c4 >> 2;
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long long)n * nc) return;
  const int i = (int)(t / nc), c = (int)(t - (long long)i * nc);
  const double x1 = (double)boxes[(size_t)i * 4], y1 = (double)boxes[(size_t)i * 4 + 1];
  const double x2 = (double)boxes[(size_t)i * 4 + 2], y2 = (double)boxes[(size_t)i * 4 + 3];
  const double w = A::add(A::sub(x2, x1), 1.0), h = A::add(A::sub(y2, y1), 1.0);
  const double cx = A::add(x1, A::mul(0.5, A::sub(w, 1.0))), cy = A::add(y1, A::mul(0.5, A::sub(h, 1.0)));
  const DT* d = deltas + (size_t)i * c4 + 4 * c;
  const DT dx = d[0], dy = d[1], dw = d[2], dh = d[3];
  const double pcx = A::add(A::mul((double)dx, w), cx), pcy = A::add(A::mul((double)dy, h), cy);
  const double pw = A::mul((double)exp_as_numpy<DT>(dw), w), ph = A::mul((double)exp_as_numpy<DT>(dh), h);
  const double hx = A::mul(0.5, A::sub(pw, 1.0)), hy = A::mul(0.5, A::sub(ph, 1.0));
  double* o = out + (size_t)i * c4 + 4 * c;
  o[0] = A::sub(pcx, hx);
  o[1] = A::sub(pcy, hy);
  o[2] = A::add(pcx, hx);
  o[3] = A::add(pcy, hy);
}

// helper/processing/bbox_transform.py:81-96, in place: max(min(v, lim), 0); x columns use lim_x.
template <typename T>
__global__ void __launch_bounds__(256) clip_boxes_kernel(T* __restrict__ boxes, long long total, T lim_x, T lim_y) {
  const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= total) return;
  const T lim = (e & 1) ? lim_y : lim_x;  // row length is a multiple of 4, so column parity == element parity
  boxes[e] = Ar<T>::vmax(Ar<T>::vmin(boxes[e], lim), T(0));
}

// helper/processing/bbox_transform.py:8-32 for one (ex, gt) pair.  numpy evaluates the ex-side
// quantities (:14-17) in the dtype of ex_rois, the gt-side ones (:19-22) in the dtype of gt_rois,
// and the four targets (:24-27) in the promoted dtype TO -- e.g. rcnn/minibatch.py:338 passes
// float64 anchors with float32 gt boxes.
template <typename TE, typename TG, typename TO>
__device__ __forceinline__ void bbox_transform_one(TE e0, TE e1, TE e2, TE e3, TG g0, TG g1, TG g2, TG g3, TO out[4]) {
  typedef Ar<TE> AE;
  typedef Ar<TG> AG;
  typedef Ar<TO> AO;
  const TE ew = AE::add(AE::sub(e2, e0), TE(1)), eh = AE::add(AE::sub(e3, e1), TE(1));
  const TE ecx = AE::add(e0, AE::mul(TE(0.5), AE::sub(ew, TE(1)))), ecy = AE::add(e1, AE::mul(TE(0.5), AE::sub(eh, TE(1))));
  const TG gw = AG::add(AG::sub(g2, g0), TG(1)), gh = AG::add(AG::sub(g3, g1), TG(1));
  const TG gcx = AG::add(g0, AG::mul(TG(0.5), AG::sub(gw, TG(1)))), gcy = AG::add(g1, AG::mul(TG(0.5), AG::sub(gh, TG(1))));
  // ex_widths + 1e-14 stays in the ex dtype (python float is a weak scalar)
  out[0] = AO::div(AO::sub((TO)gcx, (TO)ecx), (TO)AE::add(ew, TE(1e-14)));
  out[1] = AO::div(AO::sub((TO)gcy, (TO)ecy), (TO)AE::add(eh, TE(1e-14)));
  out[2] = log_as_numpy<TO>(AO::div((TO)gw, (TO)ew));
  out[3] = log_as_numpy<TO>(AO::div((TO)gh, (TO)eh));
}

// gt row of pair i is gt_index ? gt_index[i] : i.
template <typename TE, typename TG, typename TO>
__global__ void __launch_bounds__(256) bbox_transform_kernel(const TE* __restrict__ ex, int ex_stride, const TG* __restrict__ gt,
                                                             int gt_stride, const int* __restrict__ gt_index, int n,
                                                             TO* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const TE* e = ex + (size_t)i * ex_stride;
  const TG* g = gt + (size_t)(gt_index ? gt_index[i] : i) * gt_stride;
  TO o[4];
  bbox_transform_one<TE, TG, TO>(e[0], e[1], e[2], e[3], g[0], g[1], g[2], g[3], o);
  for (int k = 0; k < 4; ++k) out[(size_t)i * 4 + k] = o[k];
}

// helper/processing/bbox_regression.py:22-30 for one pair (f64, the reference's operation order).
__device__ __forceinline__ double overlap_one(double bx1, double by1, double bx2, double by2, double qx1, double qy1, double qx2,
                                              double qy2, double qarea) {
  typedef Ar<double> A;
  const double iw = A::add(A::sub(fmin(bx2, qx2), fmax(bx1, qx1)), 1.0);
  if (!(iw > 0.0)) return 0.0;
  const double ih = A::add(A::sub(fmin(by2, qy2), fmax(by1, qy1)), 1.0);
  if (!(ih > 0.0)) return 0.0;
  const double barea = A::mul(A::add(A::sub(bx2, bx1), 1.0), A::add(A::sub(by2, by1), 1.0));
  const double inter = A::mul(iw, ih);
  return A::div(inter, A::sub(A::add(barea, qarea), inter));
}

struct OverlapArgs {
  const double* boxes;   // (n, box_stride), columns 0..3 read
  int box_stride, n;
  const double* query;   // (k, query_stride)
  int query_stride, k;
  double* out;           // (n,k) or null
  int* row_argmax;       // (n) or null   -- overlaps.argmax(axis=1): first maximum
  double* row_max;       // (n) or null
  double* col_max;       // (k) or null, zero-initialised by the caller (overlaps are >= 0)
};

static const int kOvChunk = 256;

// bbox_overlaps fused with its consumers' reductions (rcnn/minibatch.py:295-300,
// rcnn/rpn/proposal_target.py:140-144).  One thread per box row; query boxes staged through
// shared memory; the column max is a warp shuffle-reduce on the ordered bit pattern of the
// (non-negative) doubles followed by one atomicMax per warp and column.
static __global__ void __launch_bounds__(128) overlaps_kernel(OverlapArgs a) {
  __shared__ double sq[kOvChunk][5];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = i < a.n;
  double bx1 = 0, by1 = 0, bx2 = 0, by2 = 0;
  if (live) {
    const double* b = a.boxes + (size_t)i * a.box_stride;
    bx1 = b[0];
    by1 = b[1];
    bx2 = b[2];
    by2 = b[3];
  }
  double best = 0.0;
  int arg = 0;
  for (int q0 = 0; q0 < a.k; q0 += kOvChunk) {
    const int qn = min(kOvChunk, a.k - q0);
    __syncthreads();
    for (int t = threadIdx.x; t < qn; t += blockDim.x) {
      const double* q = a.query + (size_t)(q0 + t) * a.query_stride;
      const double x1 = q[0], y1 = q[1], x2 = q[2], y2 = q[3];
      sq[t][0] = x1;
      sq[t][1] = y1;
      sq[t][2] = x2;
      sq[t][3] = y2;
      sq[t][4] = Ar<double>::mul(Ar<double>::add(Ar<double>::sub(x2, x1), 1.0), Ar<double>::add(Ar<double>::sub(y2, y1), 1.0));
    }
    __syncthreads();
    for (int j = 0; j < qn; ++j) {
      const double ov = live ? overlap_one(bx1, by1, bx2, by2, sq[j][0], sq[j][1], sq[j][2], sq[j][3], sq[j][4]) : 0.0;
      if (live) {
        if (a.out) a.out[(size_t)i * a.k + q0 + j] = ov;
        if ((q0 + j == 0) || ov > best) {
          best = ov;
          arg = q0 + j;
        }
      }
      if (a.col_max) {
        unsigned long long bits = (unsigned long long)__double_as_longlong(ov);
        for (int o = 16; o > 0; o >>= 1) {
          const unsigned long long other = __shfl_xor_sync(0xffffffffu, bits, o);
          bits = other > bits ? other : bits;
        }
        if ((threadIdx.x & 31) == 0 && bits != 0ull) atomicMax(reinterpret_cast<unsigned long long*>(a.col_max) + q0 + j, bits);
      }
    }
  }
  if (live) {
    if (a.row_argmax) a.row_argmax[i] = arg;
    if (a.row_max) a.row_max[i] = best;
  }
}

// helper/processing/bbox_regression.py:67-85.  One thread per (row, class slot).
static __global__ void __launch_bounds__(256) expand_targets_kernel(const float* __restrict__ data, int k, int num_classes, float4 iw,
                                                             float* __restrict__ targets, float* __restrict__ weights) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long long)k * num_classes) return;
  const int row = (int)(t / num_classes), c = (int)(t - (long long)row * num_classes);
  const float cls = data[(size_t)row * 5];
  const int start = cls > 0.f ? (int)(4.f * cls) : -1000;  // int(4 * cls), :81
  const float w4[4] = {iw.x, iw.y, iw.z, iw.w};
  for (int e = 0; e < 4; ++e) {
    const int col = 4 * c + e;
    const int off = col - start;
    const bool hit = off >= 0 && off < 4;
    targets[(size_t)row * 4 * num_classes + col] = hit ? data[(size_t)row * 5 + 1 + off] : 0.f;
    weights[(size_t)row * 4 * num_classes + col] = hit ? w4[off] : 0.f;
  }
}


// ---- one-to-one matching of proposals and gt boxes for the recall report (helper/dataset/imdb.py:153-170) ----
// k rounds over the (n,k) IoU matrix of bbox_overlaps: take the gt with the best remaining cover
// (overlaps.max(axis=0).argmax(): first maximum), the proposal that covers it best (overlaps[:, g].argmax(): first
// maximum), record that IoU, retire both (their row / column become -1).  The matrix is never stored: each column keeps
// its maximum over the live rows and the first row that attains it; after a round only the columns whose best row was
// just retired are evaluated again.  One CTA; warp w owns the columns w, w + 32, ...
struct RecallCoverArgs {
  const double* boxes;  // (n,4)
  int n;
  const double* gt;     // (k,4)
  int k;
  double* col_max;      // (k) scratch
  int* col_arg;         // (k) scratch
  unsigned* row_dead;   // (ceil(n/32)) scratch, zeroed by the kernel
  unsigned char* col_dead;  // (k) scratch
  double* covered;      // (k) out
};

__device__ __forceinline__ void recall_column(const RecallCoverArgs& a, int g, int lane) {
  const double qx1 = a.gt[(size_t)g * 4], qy1 = a.gt[(size_t)g * 4 + 1], qx2 = a.gt[(size_t)g * 4 + 2], qy2 = a.gt[(size_t)g * 4 + 3];
  const double qarea = Ar<double>::mul(Ar<double>::add(Ar<double>::sub(qx2, qx1), 1.0), Ar<double>::add(Ar<double>::sub(qy2, qy1), 1.0));
  double best = -1.0;  // a retired row reads -1 in the reference's matrix
  int arg = 0x7fffffff;
  for (int i = lane; i < a.n; i += 32) {
    if ((a.row_dead[i >> 5] >> (i & 31)) & 1u) continue;
    const double* b = a.boxes + (size_t)i * 4;
    const double ov = overlap_one(b[0], b[1], b[2], b[3], qx1, qy1, qx2, qy2, qarea);
    if (ov > best) {  // ascending i per lane: the first maximum of the lane
      best = ov;
      arg = i;
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    const double ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
    if (ob > best || (ob == best && oa < arg)) {
      best = ob;
      arg = oa;
    }
  }
  if (lane == 0) {
    a.col_max[g] = best;
    a.col_arg[g] = best < 0.0 ? 0 : arg;  // no live row left: argmax of a column of -1 is row 0
  }
}

static __global__ void __launch_bounds__(1024) recall_cover_kernel(RecallCoverArgs a) {
  __shared__ double s_val[32];
  __shared__ int s_idx[32];
  __shared__ int s_pick[2];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < (a.n + 31) / 32; i += blockDim.x) a.row_dead[i] = 0u;
  for (int g = tid; g < a.k; g += blockDim.x) a.col_dead[g] = 0;
  __syncthreads();
  for (int g = warp; g < a.k; g += 32) recall_column(a, g, lane);
  __syncthreads();
  for (int j = 0; j < a.k; ++j) {
    // g* = first column with the largest maximum (retired columns hold -1 and take part, like the reference's matrix)
    double v = -2.0;
    int gi = 0x7fffffff;
    for (int g = tid; g < a.k; g += blockDim.x) {
      const double c = a.col_max[g];
      if (c > v) {
        v = c;
        gi = g;
      }
    }
    for (int o = 16; o > 0; o >>= 1) {
      const double ov = __shfl_xor_sync(0xffffffffu, v, o);
      const int oi = __shfl_xor_sync(0xffffffffu, gi, o);
      if (ov > v || (ov == v && oi < gi)) {
        v = ov;
        gi = oi;
      }
    }
    if (lane == 0) {
      s_val[warp] = v;
      s_idx[warp] = gi;
    }
    __syncthreads();
    if (warp == 0) {
      v = s_val[lane];
      gi = s_idx[lane];
      for (int o = 16; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, v, o);
        const int oi = __shfl_xor_sync(0xffffffffu, gi, o);
        if (ov > v || (ov == v && oi < gi)) {
          v = ov;
          gi = oi;
        }
      }
      if (lane == 0) {
        const int b = a.col_arg[gi];
        a.covered[j] = v;
        s_pick[0] = gi;
        s_pick[1] = b;
        a.row_dead[b >> 5] |= 1u << (b & 31);  // overlaps[b, :] = -1
        a.col_dead[gi] = 1;                    // overlaps[:, g] = -1
        a.col_max[gi] = -1.0;
        a.col_arg[gi] = 0;
      }
    }
    __syncthreads();
    const int b_star = s_pick[1];
    for (int g = warp; g < a.k; g += 32)
      if (!a.col_dead[g] && a.col_arg[g] == b_star) recall_column(a, g, lane);
    __syncthreads();
  }
}

}  // namespace mxr

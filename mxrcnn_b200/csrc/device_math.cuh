// Parity-critical device arithmetic.  Every operation that the reference performs in
// numpy is spelled with an explicit round-to-nearest intrinsic, so that neither ptxas nor
// nvcc can contract a multiply-add (the library is also compiled with -fmad=false).
#pragma once
#include "common.cuh"

namespace mxr {

template <typename T>
struct Ar;

template <>
struct Ar<double> {
  typedef double4 Box;
  static __device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
  static __device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
  static __device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
  static __device__ __forceinline__ double div(double a, double b) { return __ddiv_rn(a, b); }
  // compare-select (DSETP + 2 SEL) instead of fmax/fmin, whose NaN handling costs ~10 instructions
  // per call on sm_100; identical for the non-NaN values that reach these kernels
  static __device__ __forceinline__ double vmax(double a, double b) { return a > b ? a : b; }
  static __device__ __forceinline__ double vmin(double a, double b) { return a < b ? a : b; }
  // guard band of the division-free IoU test: 1 -/+ 2^-48 (rounding error of the test is < 2^-51)
  static __device__ __forceinline__ double band_lo() { return 1.0 - 3.552713678800501e-15; }
  static __device__ __forceinline__ double band_hi() { return 1.0 + 3.552713678800501e-15; }
};

template <>
struct Ar<float> {
  typedef float4 Box;
  static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
  static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
  static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
  static __device__ __forceinline__ float div(float a, float b) { return __fdiv_rn(a, b); }
  static __device__ __forceinline__ float vmax(float a, float b) { return fmaxf(a, b); }
  static __device__ __forceinline__ float vmin(float a, float b) { return fminf(a, b); }
  // 1 -/+ 2^-19 (rounding error of the test is < 2^-22)
  static __device__ __forceinline__ float band_lo() { return 1.0f - 1.9073486328125e-06f; }
  static __device__ __forceinline__ float band_hi() { return 1.0f + 1.9073486328125e-06f; }
};

// area exactly as helper/processing/nms.py:18:  (x2 - x1 + 1) * (y2 - y1 + 1)
template <typename T>
__device__ __forceinline__ T box_area(T x1, T y1, T x2, T y2) {
  typedef Ar<T> A;
  return A::mul(A::add(A::sub(x2, x1), T(1)), A::add(A::sub(y2, y1), T(1)));
}

// helper/processing/nms.py:25-36.  Box a is the kept box i, box b the candidate j.
// Returns true when j is SUPPRESSED, i.e. when `ovr <= thresh` is false for
//   ovr = inter / (area_a + area_b - inter)
// with every operation rounded as numpy rounds it.  The IEEE division is evaluated only
// inside a 2^-48 (f64) / 2^-19 (f32) relative band around the threshold; outside it
// `inter <= thr*union*(1-band)` (resp. >=) decides the same predicate exactly, because
// round-to-nearest is monotone and thresh is representable.
template <typename T>
__device__ __forceinline__ bool iou_exceeds(T ax1, T ay1, T ax2, T ay2, T aarea, T bx1, T by1, T bx2, T by2, T barea,
                                            T thr) {
  typedef Ar<T> A;
  const T xx1 = A::vmax(ax1, bx1);
  const T yy1 = A::vmax(ay1, by1);
  const T xx2 = A::vmin(ax2, bx2);
  const T yy2 = A::vmin(ay2, by2);
  const T w = A::vmax(T(0), A::add(A::sub(xx2, xx1), T(1)));
  const T h = A::vmax(T(0), A::add(A::sub(yy2, yy1), T(1)));
  const T inter = A::mul(w, h);
  const T uni = A::sub(A::add(aarea, barea), inter);
  if (uni > T(0) && thr > T(0)) {
    const T t = A::mul(thr, uni);
    if (inter <= A::mul(t, A::band_lo())) return false;
    if (inter >= A::mul(t, A::band_hi())) return true;
  }
  return !(A::div(inter, uni) <= thr);
}

// Same predicate with the loop-invariant parts hoisted: thr_lo = thr*(1-band), thr_hi = thr*(1+band)
// (thr > 0 checked by the caller; the extra rounding of the hoisted products is < 2^-52, far inside
// the 2^-48 band).  Returns 0 = keep, 1 = suppressed, 2 = inside the band / degenerate: the caller
// re-evaluates with iou_exceeds().  Branch-free so that a warp does not diverge on it.
template <typename T>
__device__ __forceinline__ int iou_band(T ax1, T ay1, T ax2, T ay2, T aarea, T bx1, T by1, T bx2, T by2, T barea, T thr_lo,
                                        T thr_hi) {
  typedef Ar<T> A;
  const T xx1 = A::vmax(ax1, bx1), yy1 = A::vmax(ay1, by1);
  const T xx2 = A::vmin(ax2, bx2), yy2 = A::vmin(ay2, by2);
  const T w0 = A::add(A::sub(xx2, xx1), T(1)), h0 = A::add(A::sub(yy2, yy1), T(1));
  // max(0, w0) * max(0, h0) without the two clamps (the compiler expands `x > 0 ? x : 0` on doubles into a NaN-quieting
  // max of ~8 instructions): the product is +0 as soon as one extent is <= 0.  That differs from the clamped
  // form only when the other extent is infinite or NaN (0 * inf), and then an area, hence `uni`, is infinite or NaN
  // too: such pairs are sent to the exact evaluation by the `finite` test below.
  const T inter = (w0 > T(0) && h0 > T(0)) ? A::mul(w0, h0) : T(0);
  const T uni = A::sub(A::add(aarea, barea), inter);
  const bool pos = uni > T(0) && uni <= (sizeof(T) == 8 ? T(1.7976931348623157e308) : T(3.4028234663852886e38f));
  const bool keep = inter <= A::mul(thr_lo, uni);
  const bool supp = inter >= A::mul(thr_hi, uni);
  return !pos ? 2 : (keep ? 0 : (supp ? 1 : 2));
}

// Order-preserving map f32 -> u32 (larger float => larger key).  -0.0 is folded onto
// +0.0 so that the two compare equal, as they do for numpy's argsort.  A non-NaN score
// never maps to 0, which marks "filtered out".
__device__ __forceinline__ u32 score_key32(float s) {
  u32 b = __float_as_uint(s + 0.0f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ u64 score_key64(double s) {
  u64 b = (u64)__double_as_longlong(s + 0.0);
  return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
}

struct DecodeGeom {
  int A, H, W;     // full maps
  int hc, wc;      // cropped extent actually enumerated (proposal.py:70-71, 99)
  double fs;       // feat stride
};

// n / d for 0 <= n < 2^24, 1 <= d < 2^24 (anchor indices: A*H*W <= 2^24) through a float reciprocal plus one
// correction loop; `rcp` = 1.0f / d computed once.  An integer division by a run-time value costs ~25 instructions.
// (float)n is exact; rcp and the product carry a relative error of 2^-24 each.
__device__ __forceinline__ int fast_div(int n, int d, float rcp) {
  int q = (int)((float)n * rcp);
  int r = n - q * d;
  // the estimate is off by less than (n / d) 2^-22 + 1: one step for quotients below 2^22 (every use but a
  // single-anchor map of more than 4M cells), at most three in general
  while (r >= d) {
    ++q;
    r -= d;
  }
  while (r < 0) {
    --q;
    r += d;
  }
  return q;
}

// max(min(v, lim), 0) of bbox_transform.py:86-94 as two compare-selects.  The library fmin / fmax cost ~8
// instructions each on sm_100 (NaN quieting), and the compiler turns `x > 0.0 ? x : 0.0` into the same sequence, so
// the selects are spelled on the 32-bit halves.  Same values as fmax(fmin(v, lim), 0.0), including NaN -> lim and
// -0.0 -> +0.0 (the NaN check of proposal.py:61 looks at the raw deltas, not at the clipped boxes).
__device__ __forceinline__ double select_f64(bool p, double a, double b) {
  return __hiloint2double(p ? __double2hiint(a) : __double2hiint(b), p ? __double2loint(a) : __double2loint(b));
}
__device__ __forceinline__ double clip_0_lim(double v, double lim) {
  const double t = select_f64(v < lim, v, lim);
  return select_f64(t > 0.0, t, 0.0);
}

// rcnn/rpn/proposal.py:77-113 + helper/processing/bbox_transform.py:35-96 for ONE anchor.
// (h, w, a) index the cropped map; deltas are read from the full (4A,H,W) map of the image.
// f64 arithmetic, exp in f32 (numpy evaluates np.exp on the f32 deltas), clip limits are
// the f32 scalars im_info[k]-1.
// The arithmetic, for deltas (dx, dy, dw, dh) already in registers.
__device__ __forceinline__ void decode_clip_values(const AnchorTable& at, const DecodeGeom& g, float dx, float dy, float dw,
                                                   float dh, int h, int w, int a, float im_h, float im_w, double out[4]) {
  typedef Ar<double> A;
  const double sx = A::mul((double)w, g.fs);
  const double sy = A::mul((double)h, g.fs);
  const double ax1 = A::add(at.a[a][0], sx), ay1 = A::add(at.a[a][1], sy);
  const double ax2 = A::add(at.a[a][2], sx), ay2 = A::add(at.a[a][3], sy);
  const double bw = A::add(A::sub(ax2, ax1), 1.0);
  const double bh = A::add(A::sub(ay2, ay1), 1.0);
  const double cx = A::add(ax1, A::mul(0.5, A::sub(bw, 1.0)));
  const double cy = A::add(ay1, A::mul(0.5, A::sub(bh, 1.0)));
  const double pcx = A::add(A::mul((double)dx, bw), cx);
  const double pcy = A::add(A::mul((double)dy, bh), cy);
  const double pw = A::mul((double)expf(dw), bw);
  const double ph = A::mul((double)expf(dh), bh);
  const double hx = A::mul(0.5, A::sub(pw, 1.0));
  const double hy = A::mul(0.5, A::sub(ph, 1.0));
  const double lim_x = (double)__fsub_rn(im_w, 1.0f);
  const double lim_y = (double)__fsub_rn(im_h, 1.0f);
  out[0] = clip_0_lim(A::sub(pcx, hx), lim_x);
  out[1] = clip_0_lim(A::sub(pcy, hy), lim_y);
  out[2] = clip_0_lim(A::add(pcx, hx), lim_x);
  out[3] = clip_0_lim(A::add(pcy, hy), lim_y);
}

// Loads + arithmetic for one anchor.
__device__ __forceinline__ void decode_clip(const AnchorTable& at, const DecodeGeom& g, const float* __restrict__ deltas,
                                            int h, int w, int a, float im_h, float im_w, double out[4],
                                            bool* delta_nan = nullptr) {
  const size_t hw = (size_t)g.H * g.W;
  const size_t base = ((size_t)a * 4) * hw + (size_t)h * g.W + w;
  const float dx = __ldg(deltas + base);
  const float dy = __ldg(deltas + base + hw);
  const float dw = __ldg(deltas + base + 2 * hw);
  const float dh = __ldg(deltas + base + 3 * hw);
  // the clip drops NaNs, so the reference's NaN check (proposal.py:61) looks at the raw deltas
  if (delta_nan) *delta_nan = (dx != dx) || (dy != dy) || (dw != dw) || (dh != dh);
  decode_clip_values(at, g, dx, dy, dw, dh, h, w, a, im_h, im_w, out);
}

// proposal.py:156-161 with min_size*im_info[2] evaluated in f32 (NEP 50) and promoted.
__device__ __forceinline__ bool passes_min_size(const double b[4], float min_size, float scale) {
  typedef Ar<double> A;
  const double ms = (double)__fmul_rn(min_size, scale);
  const double ws = A::add(A::sub(b[2], b[0]), 1.0);
  const double hs = A::add(A::sub(b[3], b[1]), 1.0);
  return (ws >= ms) && (hs >= ms);
}

}  // namespace mxr

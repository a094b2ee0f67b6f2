// Host-side driver of the striped NMS (shared by the proposal op and nms()).
#pragma once
#include "nms_kernels.cuh"

namespace mxr {

static const int kMaxStripe = 4096;  // widest stripe: in-stripe mask = 4096 x 64 words = 2 MiB / problem

struct StripePlan {
  int first;      // width of stripe 0
  int max_width;  // widest stripe that will be launched
};

static inline StripePlan plan_stripes(int n_max, int max_keep, int first_override) {
  StripePlan p;
  int w0;
  if (first_override > 0) {
    w0 = round_up_int(first_override, 64);
  } else if (max_keep > 0) {
    w0 = pow2_ceil(2 * max_keep);
    if (w0 < 1024) w0 = 1024;
  } else {
    w0 = kMaxStripe;
  }
  if (w0 > kMaxStripe) w0 = kMaxStripe;
  const int n64 = round_up_int(n_max < 1 ? 1 : n_max, 64);
  if (w0 > n64) w0 = n64;
  p.first = w0;
  int c0 = 0, w = w0, mw = 0;
  do {
    int width = w < n64 - c0 ? w : n64 - c0;
    if (width > mw) mw = width;
    c0 += width;
    w = w * 2 > kMaxStripe ? kMaxStripe : w * 2;
  } while (c0 < n64);
  p.max_width = mw;
  return p;
}

// Enqueue all stripes.  `a` must have M sized for plan.max_width; R and ctrl zeroed
// (ctrl.n set by the preceding kernel).
template <typename T>
int run_nms_stripes(Handle* h, cudaStream_t s, Ctrl* ctrl, NmsArrays<T> a, NmsOutputs o, int B, int n_max, int max_keep,
                    const StripePlan& plan, T thr) {
  const int n64 = round_up_int(n_max < 1 ? 1 : n_max, 64);
  int c0 = 0, w = plan.first;
  do {
    const int width = w < n64 - c0 ? w : n64 - c0;
    const int sw = width / 64;
    int prev_rows = c0;
    if (max_keep > 0 && prev_rows > max_keep - 1) prev_rows = max_keep - 1;
    const int row_chunks = (prev_rows + 63) / 64;
    dim3 grid(sw * sw + sw * row_chunks, 1, B);
    prof_begin(h, s, ST_NMS_STRIPE);
    nms_stripe_kernel<T><<<grid, 64, 0, s>>>(ctrl, a, c0, width, sw, thr);
    prof_end(h, s);
    const int last = (c0 + width >= n64) ? 1 : 0;
    prof_begin(h, s, ST_NMS_SCAN);
    nms_scan_kernel<T><<<B, 256, sw * sizeof(u64), s>>>(ctrl, a, o, c0, width, max_keep, last);
    prof_end(h, s);
    c0 += width;
    w = w * 2 > kMaxStripe ? kMaxStripe : w * 2;
  } while (c0 < n64);
  MXR_CUDA(cudaGetLastError());
  return MXR_OK;
}

}  // namespace mxr

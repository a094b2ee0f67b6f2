// Box helpers of helper/processing/bbox_transform.py and bbox_regression.py:
//   bbox_pred (A2), clip_boxes (A3), bbox_transform (A10), bbox_overlaps fused with the row
//   argmax / row max / column max (A7), expand_bbox_regression_targets (A9).
// All are streaming kernels (HBM-bound at scale, launch-bound at the reference's sizes).
#include <string.h>

#include "box_kernels.cuh"

using namespace mxr;

namespace {

// Copies host arrays into the workspace, runs `fn` on the device copies, copies results back.
struct Staged {
  Handle* h;
  cudaStream_t s;
  size_t off;
  explicit Staged(Handle* h_) : h(h_), s(h_->stream), off(0) {}
  size_t reserve(size_t bytes) {
    off = align_up(off, 256);
    const size_t o = off;
    off += bytes;
    return o;
  }
};

}  // namespace

extern "C" {

// ---- bbox_pred ---------------------------------------------------------------------------
int mxr_bbox_pred(mxr_handle_t hh, void* stream, const void* boxes, int boxes_dtype, const void* deltas,
                  int deltas_dtype, int n, int c4, double* out) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_bbox_pred: NULL handle");
  MXR_REQUIRE(n >= 0 && c4 >= 0 && c4 % 4 == 0, MXR_ERR_INVALID, "mxr_bbox_pred: bad shape n=%d cols=%d", n, c4);
  if (n == 0 || c4 == 0) return MXR_OK;  // bbox_transform.py:43-44: empty in, empty out
  MXR_REQUIRE(boxes && deltas && out, MXR_ERR_INVALID, "mxr_bbox_pred: NULL array");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  cudaStream_t s = (cudaStream_t)stream;
  MXR_TRY(ws_order(h, s));
  const long long total = (long long)n * (c4 / 4);
  const int grid = (int)((total + 255) / 256);
  h->last_launches = 0;
  prof_begin(h, s, ST_BOXES);
  if (boxes_dtype == MXR_DTYPE_F64 && deltas_dtype == MXR_DTYPE_F32)
    bbox_pred_kernel<double, float><<<grid, 256, 0, s>>>((const double*)boxes, (const float*)deltas, n, c4, out);
  else if (boxes_dtype == MXR_DTYPE_F32 && deltas_dtype == MXR_DTYPE_F32)
    bbox_pred_kernel<float, float><<<grid, 256, 0, s>>>((const float*)boxes, (const float*)deltas, n, c4, out);
  else if (boxes_dtype == MXR_DTYPE_F64 && deltas_dtype == MXR_DTYPE_F64)
    bbox_pred_kernel<double, double><<<grid, 256, 0, s>>>((const double*)boxes, (const double*)deltas, n, c4, out);
  else if (boxes_dtype == MXR_DTYPE_F32 && deltas_dtype == MXR_DTYPE_F64)
    bbox_pred_kernel<float, double><<<grid, 256, 0, s>>>((const float*)boxes, (const double*)deltas, n, c4, out);
  else {
    set_error("mxr_bbox_pred: dtypes must be f32 or f64");
    return MXR_ERR_INVALID;
  }
  prof_end(h, s);
  MXR_CUDA(cudaGetLastError());
  return MXR_OK;
}

int mxr_bbox_pred_host(mxr_handle_t hh, const void* boxes, int boxes_dtype, const void* deltas, int deltas_dtype, int n,
                       int c4, double* out) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_bbox_pred_host: NULL handle");
  if (n <= 0 || c4 <= 0) return n < 0 || c4 < 0 ? MXR_ERR_INVALID : MXR_OK;
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  const size_t bs = (size_t)n * 4 * (boxes_dtype == MXR_DTYPE_F64 ? 8 : 4);
  const size_t ds = (size_t)n * c4 * (deltas_dtype == MXR_DTYPE_F64 ? 8 : 4);
  const size_t os = (size_t)n * c4 * 8;
  Staged st(h);
  const size_t ob = st.reserve(bs), od = st.reserve(ds), oo = st.reserve(os);
  MXR_TRY(ensure_device_ws(h, st.off));
  char* w = h->d_ws;
  MXR_CUDA(cudaMemcpyAsync(w + ob, boxes, bs, cudaMemcpyHostToDevice, st.s));
  MXR_CUDA(cudaMemcpyAsync(w + od, deltas, ds, cudaMemcpyHostToDevice, st.s));
  MXR_TRY(mxr_bbox_pred(hh, st.s, w + ob, boxes_dtype, w + od, deltas_dtype, n, c4, (double*)(w + oo)));
  MXR_CUDA(cudaMemcpyAsync(out, w + oo, os, cudaMemcpyDeviceToHost, st.s));
  MXR_CUDA(cudaStreamSynchronize(st.s));
  return MXR_OK;
}

// ---- clip_boxes --------------------------------------------------------------------------
int mxr_clip_boxes(mxr_handle_t hh, void* stream, void* boxes, int dtype, int n, int c4, double lim_x, double lim_y) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_clip_boxes: NULL handle");
  MXR_REQUIRE(n >= 0 && c4 >= 0 && c4 % 4 == 0, MXR_ERR_INVALID, "mxr_clip_boxes: bad shape n=%d cols=%d", n, c4);
  if (n == 0 || c4 == 0) return MXR_OK;
  MXR_REQUIRE(boxes != nullptr, MXR_ERR_INVALID, "mxr_clip_boxes: NULL array");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  cudaStream_t s = (cudaStream_t)stream;
  MXR_TRY(ws_order(h, s));
  const long long total = (long long)n * c4;
  const int grid = (int)((total + 255) / 256);
  h->last_launches = 0;
  prof_begin(h, s, ST_BOXES);
  if (dtype == MXR_DTYPE_F64)
    clip_boxes_kernel<double><<<grid, 256, 0, s>>>((double*)boxes, total, lim_x, lim_y);
  else if (dtype == MXR_DTYPE_F32)
    clip_boxes_kernel<float><<<grid, 256, 0, s>>>((float*)boxes, total, (float)lim_x, (float)lim_y);
  else {
    set_error("mxr_clip_boxes: dtype must be f32 or f64");
    return MXR_ERR_INVALID;
  }
  prof_end(h, s);
  MXR_CUDA(cudaGetLastError());
  return MXR_OK;
}

int mxr_clip_boxes_host(mxr_handle_t hh, void* boxes, int dtype, int n, int c4, double lim_x, double lim_y) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_clip_boxes_host: NULL handle");
  if (n <= 0 || c4 <= 0) return n < 0 || c4 < 0 ? MXR_ERR_INVALID : MXR_OK;
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  const size_t bs = (size_t)n * c4 * (dtype == MXR_DTYPE_F64 ? 8 : 4);
  MXR_TRY(ensure_device_ws(h, bs));
  MXR_CUDA(cudaMemcpyAsync(h->d_ws, boxes, bs, cudaMemcpyHostToDevice, h->stream));
  MXR_TRY(mxr_clip_boxes(hh, h->stream, h->d_ws, dtype, n, c4, lim_x, lim_y));
  MXR_CUDA(cudaMemcpyAsync(boxes, h->d_ws, bs, cudaMemcpyDeviceToHost, h->stream));
  MXR_CUDA(cudaStreamSynchronize(h->stream));
  return MXR_OK;
}

// ---- bbox_transform ----------------------------------------------------------------------
int mxr_bbox_transform(mxr_handle_t hh, void* stream, const void* ex, int ex_dtype, const void* gt, int gt_dtype, int n,
                       void* out) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_bbox_transform: NULL handle");
  MXR_REQUIRE(n >= 0, MXR_ERR_INVALID, "mxr_bbox_transform: n < 0");
  if (n == 0) return MXR_OK;
  MXR_REQUIRE(ex && gt && out, MXR_ERR_INVALID, "mxr_bbox_transform: NULL array");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  cudaStream_t s = (cudaStream_t)stream;
  MXR_TRY(ws_order(h, s));
  h->last_launches = 0;
  const int g = (n + 255) / 256;
  prof_begin(h, s, ST_BOXES);
  if (ex_dtype == MXR_DTYPE_F64 && gt_dtype == MXR_DTYPE_F64)
    bbox_transform_kernel<double, double, double><<<g, 256, 0, s>>>((const double*)ex, 4, (const double*)gt, 4, nullptr, n, (double*)out);
  else if (ex_dtype == MXR_DTYPE_F32 && gt_dtype == MXR_DTYPE_F32)
    bbox_transform_kernel<float, float, float><<<g, 256, 0, s>>>((const float*)ex, 4, (const float*)gt, 4, nullptr, n, (float*)out);
  else if (ex_dtype == MXR_DTYPE_F64 && gt_dtype == MXR_DTYPE_F32)
    bbox_transform_kernel<double, float, double><<<g, 256, 0, s>>>((const double*)ex, 4, (const float*)gt, 4, nullptr, n, (double*)out);
  else if (ex_dtype == MXR_DTYPE_F32 && gt_dtype == MXR_DTYPE_F64)
    bbox_transform_kernel<float, double, double><<<g, 256, 0, s>>>((const float*)ex, 4, (const double*)gt, 4, nullptr, n, (double*)out);
  else {
    set_error("mxr_bbox_transform: dtypes must be f32 or f64");
    return MXR_ERR_INVALID;
  }
  prof_end(h, s);
  MXR_CUDA(cudaGetLastError());
  return MXR_OK;
}

int mxr_bbox_transform_host(mxr_handle_t hh, const void* ex, int ex_dtype, const void* gt, int gt_dtype, int n, void* out) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_bbox_transform_host: NULL handle");
  if (n <= 0) return n < 0 ? MXR_ERR_INVALID : MXR_OK;
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  const size_t es = (size_t)n * 4 * (ex_dtype == MXR_DTYPE_F64 ? 8 : 4);
  const size_t gs = (size_t)n * 4 * (gt_dtype == MXR_DTYPE_F64 ? 8 : 4);
  const size_t os = (size_t)n * 4 * ((ex_dtype == MXR_DTYPE_F32 && gt_dtype == MXR_DTYPE_F32) ? 4 : 8);
  Staged st(h);
  const size_t oe = st.reserve(es), og = st.reserve(gs), oo = st.reserve(os);
  MXR_TRY(ensure_device_ws(h, st.off));
  char* w = h->d_ws;
  MXR_CUDA(cudaMemcpyAsync(w + oe, ex, es, cudaMemcpyHostToDevice, st.s));
  MXR_CUDA(cudaMemcpyAsync(w + og, gt, gs, cudaMemcpyHostToDevice, st.s));
  MXR_TRY(mxr_bbox_transform(hh, st.s, w + oe, ex_dtype, w + og, gt_dtype, n, w + oo));
  MXR_CUDA(cudaMemcpyAsync(out, w + oo, os, cudaMemcpyDeviceToHost, st.s));
  MXR_CUDA(cudaStreamSynchronize(st.s));
  return MXR_OK;
}

// ---- bbox_overlaps -----------------------------------------------------------------------
int mxr_bbox_overlaps(mxr_handle_t hh, void* stream, const double* boxes, int n, int box_stride, const double* query,
                      int k, int query_stride, double* out, int32_t* row_argmax, double* row_max, double* col_max) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_bbox_overlaps: NULL handle");
  MXR_REQUIRE(n >= 0 && k >= 0 && box_stride >= 4 && query_stride >= 4, MXR_ERR_INVALID,
              "mxr_bbox_overlaps: bad shape n=%d k=%d strides %d/%d", n, k, box_stride, query_stride);
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  cudaStream_t s = (cudaStream_t)stream;
  MXR_TRY(ws_order(h, s));
  h->last_launches = 0;
  if (col_max && k > 0) MXR_CUDA(cudaMemsetAsync(col_max, 0, (size_t)k * sizeof(double), s));  // overlaps are >= 0
  if (n == 0 || k == 0) return MXR_OK;
  MXR_REQUIRE(boxes && query, MXR_ERR_INVALID, "mxr_bbox_overlaps: NULL array");
  OverlapArgs oa;
  oa.boxes = boxes;
  oa.box_stride = box_stride;
  oa.n = n;
  oa.query = query;
  oa.query_stride = query_stride;
  oa.k = k;
  oa.out = out;
  oa.row_argmax = row_argmax;
  oa.row_max = row_max;
  oa.col_max = col_max;
  prof_begin(h, s, ST_OVERLAPS);
  overlaps_kernel<<<(n + 127) / 128, 128, 0, s>>>(oa);
  prof_end(h, s);
  MXR_CUDA(cudaGetLastError());
  return MXR_OK;
}

int mxr_bbox_overlaps_host(mxr_handle_t hh, const double* boxes, int n, int box_stride, const double* query, int k,
                           int query_stride, double* out, int32_t* row_argmax, double* row_max, double* col_max) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_bbox_overlaps_host: NULL handle");
  MXR_REQUIRE(n >= 0 && k >= 0, MXR_ERR_INVALID, "mxr_bbox_overlaps_host: bad shape");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  Staged st(h);
  const size_t bb = (size_t)n * box_stride * 8, qb = (size_t)k * query_stride * 8;
  const size_t ob = st.reserve(bb ? bb : 8), oq = st.reserve(qb ? qb : 8);
  const size_t oo = st.reserve(out ? (size_t)n * k * 8 + 8 : 8);
  const size_t oa = st.reserve((size_t)n * 4 + 8), om = st.reserve((size_t)n * 8 + 8), oc = st.reserve((size_t)k * 8 + 8);
  MXR_TRY(ensure_device_ws(h, st.off));
  char* w = h->d_ws;
  if (bb) MXR_CUDA(cudaMemcpyAsync(w + ob, boxes, bb, cudaMemcpyHostToDevice, st.s));
  if (qb) MXR_CUDA(cudaMemcpyAsync(w + oq, query, qb, cudaMemcpyHostToDevice, st.s));
  MXR_TRY(mxr_bbox_overlaps(hh, st.s, (const double*)(w + ob), n, box_stride, (const double*)(w + oq), k, query_stride,
                            out ? (double*)(w + oo) : nullptr, row_argmax ? (int32_t*)(w + oa) : nullptr,
                            row_max ? (double*)(w + om) : nullptr, col_max ? (double*)(w + oc) : nullptr));
  if (out && n && k) MXR_CUDA(cudaMemcpyAsync(out, w + oo, (size_t)n * k * 8, cudaMemcpyDeviceToHost, st.s));
  if (row_argmax && n) MXR_CUDA(cudaMemcpyAsync(row_argmax, w + oa, (size_t)n * 4, cudaMemcpyDeviceToHost, st.s));
  if (row_max && n) MXR_CUDA(cudaMemcpyAsync(row_max, w + om, (size_t)n * 8, cudaMemcpyDeviceToHost, st.s));
  if (col_max && k) MXR_CUDA(cudaMemcpyAsync(col_max, w + oc, (size_t)k * 8, cudaMemcpyDeviceToHost, st.s));
  MXR_CUDA(cudaStreamSynchronize(st.s));
  return MXR_OK;
}

// ---- recall matching ---------------------------------------------------------------------
int mxr_recall_cover_host(mxr_handle_t hh, const double* boxes, int n, const double* gt_boxes, int k, double* covered_out) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_recall_cover_host: NULL handle");
  MXR_REQUIRE(n >= 0 && k >= 0, MXR_ERR_INVALID, "mxr_recall_cover_host: bad shape");
  if (k == 0) return MXR_OK;
  MXR_REQUIRE(gt_boxes && covered_out && (n == 0 || boxes), MXR_ERR_INVALID, "mxr_recall_cover_host: NULL array");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  Staged st(h);
  MXR_TRY(ws_order(h, st.s));
  const size_t ob = st.reserve((size_t)(n ? n : 1) * 32), og = st.reserve((size_t)k * 32);
  const size_t om = st.reserve((size_t)k * 8), oa = st.reserve((size_t)k * 4), ord = st.reserve((size_t)(n / 32 + 1) * 4);
  const size_t ocd = st.reserve((size_t)k), oc = st.reserve((size_t)k * 8);
  MXR_TRY(ensure_device_ws(h, st.off));
  char* w = h->d_ws;
  if (n) MXR_CUDA(cudaMemcpyAsync(w + ob, boxes, (size_t)n * 32, cudaMemcpyHostToDevice, st.s));
  MXR_CUDA(cudaMemcpyAsync(w + og, gt_boxes, (size_t)k * 32, cudaMemcpyHostToDevice, st.s));
  RecallCoverArgs a;
  a.boxes = reinterpret_cast<const double*>(w + ob);
  a.n = n;
  a.gt = reinterpret_cast<const double*>(w + og);
  a.k = k;
  a.col_max = reinterpret_cast<double*>(w + om);
  a.col_arg = reinterpret_cast<int*>(w + oa);
  a.row_dead = reinterpret_cast<unsigned*>(w + ord);
  a.col_dead = reinterpret_cast<unsigned char*>(w + ocd);
  a.covered = reinterpret_cast<double*>(w + oc);
  h->last_launches = 0;
  prof_begin(h, st.s, ST_OVERLAPS);
  recall_cover_kernel<<<1, 1024, 0, st.s>>>(a);
  prof_end(h, st.s);
  MXR_CUDA(cudaGetLastError());
  MXR_CUDA(cudaMemcpyAsync(covered_out, w + oc, (size_t)k * 8, cudaMemcpyDeviceToHost, st.s));
  MXR_CUDA(cudaStreamSynchronize(st.s));
  return MXR_OK;
}

// ---- expand_bbox_regression_targets ------------------------------------------------------
int mxr_expand_bbox_targets(mxr_handle_t hh, void* stream, const float* data, int k, int num_classes,
                            const float* inside_weights4, float* targets, float* weights) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_expand_bbox_targets: NULL handle");
  MXR_REQUIRE(k >= 0 && num_classes >= 1 && inside_weights4 != nullptr, MXR_ERR_INVALID, "mxr_expand_bbox_targets: bad argument");
  if (k == 0) return MXR_OK;
  MXR_REQUIRE(data && targets && weights, MXR_ERR_INVALID, "mxr_expand_bbox_targets: NULL array");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  cudaStream_t s = (cudaStream_t)stream;
  MXR_TRY(ws_order(h, s));
  float4 iw = make_float4(inside_weights4[0], inside_weights4[1], inside_weights4[2], inside_weights4[3]);
  const long long total = (long long)k * num_classes;
  h->last_launches = 0;
  prof_begin(h, s, ST_BOXES);
  expand_targets_kernel<<<(int)((total + 255) / 256), 256, 0, s>>>(data, k, num_classes, iw, targets, weights);
  prof_end(h, s);
  MXR_CUDA(cudaGetLastError());
  return MXR_OK;
}

int mxr_expand_bbox_targets_host(mxr_handle_t hh, const float* data, int k, int num_classes, const float* inside_weights4,
                                 float* targets, float* weights) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_expand_bbox_targets_host: NULL handle");
  if (k <= 0) return k < 0 ? MXR_ERR_INVALID : MXR_OK;
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  Staged st(h);
  const size_t db = (size_t)k * 5 * 4, tb = (size_t)k * 4 * num_classes * 4;
  const size_t od = st.reserve(db), ot = st.reserve(tb), ow = st.reserve(tb);
  MXR_TRY(ensure_device_ws(h, st.off));
  char* w = h->d_ws;
  MXR_CUDA(cudaMemcpyAsync(w + od, data, db, cudaMemcpyHostToDevice, st.s));
  MXR_TRY(mxr_expand_bbox_targets(hh, st.s, (const float*)(w + od), k, num_classes, inside_weights4, (float*)(w + ot),
                                  (float*)(w + ow)));
  MXR_CUDA(cudaMemcpyAsync(targets, w + ot, tb, cudaMemcpyDeviceToHost, st.s));
  MXR_CUDA(cudaMemcpyAsync(weights, w + ow, tb, cudaMemcpyDeviceToHost, st.s));
  MXR_CUDA(cudaStreamSynchronize(st.s));
  return MXR_OK;
}

}  // extern "C"

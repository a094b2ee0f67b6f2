// helper/processing/nms.py:40-70 -- nest(dets, thresh): drop box i when, for some other box j,
// inter(i,j) / area(i) > thresh (box i lies almost entirely inside box j).  O(n^2), one thread per box,
// the other boxes staged through shared memory; arithmetic in the dtype of dets with IEEE division.
#include <string.h>

#include "device_math.cuh"

using namespace mxr;

namespace {

template <typename T>
__global__ void __launch_bounds__(128) nest_kernel(const T* __restrict__ dets, int n, T thr, int* __restrict__ keep_flag) {
  typedef Ar<T> A;
  __shared__ T sb[128][4];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = i < n;
  T x1 = 0, y1 = 0, x2 = 0, y2 = 0, area = 1;
  if (live) {
    x1 = dets[(size_t)i * 5];
    y1 = dets[(size_t)i * 5 + 1];
    x2 = dets[(size_t)i * 5 + 2];
    y2 = dets[(size_t)i * 5 + 3];
    area = box_area<T>(x1, y1, x2, y2);  // :49
  }
  bool keep = true;
  for (int j0 = 0; j0 < n; j0 += 128) {
    const int jn = min(128, n - j0);
    __syncthreads();
    if ((int)threadIdx.x < jn)
      for (int k = 0; k < 4; ++k) sb[threadIdx.x][k] = dets[(size_t)(j0 + threadIdx.x) * 5 + k];
    __syncthreads();
    if (live) {
      for (int j = 0; j < jn; ++j) {
        if (j0 + j == i) continue;  // :56
        const T xx1 = A::vmax(x1, sb[j][0]), yy1 = A::vmax(y1, sb[j][1]);
        const T xx2 = A::vmin(x2, sb[j][2]), yy2 = A::vmin(y2, sb[j][3]);
        const T w = A::vmax(T(0), A::add(A::sub(xx2, xx1), T(1)));
        const T h = A::vmax(T(0), A::add(A::sub(yy2, yy1), T(1)));
        if (A::div(A::mul(w, h), area) > thr) keep = false;  // :63-64
      }
    }
  }
  if (live) keep_flag[i] = keep ? 1 : 0;
}

}  // namespace

extern "C" {

int mxr_nest(mxr_handle_t hh, void* stream, const void* dets, int n, int dtype, double thresh, int32_t* keep_flag) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_nest: NULL handle");
  MXR_REQUIRE(n >= 0, MXR_ERR_INVALID, "mxr_nest: n < 0");
  if (n == 0) return MXR_OK;
  MXR_REQUIRE(dets && keep_flag, MXR_ERR_INVALID, "mxr_nest: NULL array");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  cudaStream_t s = (cudaStream_t)stream;
  MXR_TRY(ws_order(h, s));
  h->last_launches = 0;
  prof_begin(h, s, ST_BOXES);
  if (dtype == MXR_DTYPE_F64)
    nest_kernel<double><<<(n + 127) / 128, 128, 0, s>>>((const double*)dets, n, thresh, keep_flag);
  else if (dtype == MXR_DTYPE_F32)
    nest_kernel<float><<<(n + 127) / 128, 128, 0, s>>>((const float*)dets, n, (float)thresh, keep_flag);
  else {
    set_error("mxr_nest: dtype must be f32 or f64");
    return MXR_ERR_INVALID;
  }
  prof_end(h, s);
  MXR_CUDA(cudaGetLastError());
  return MXR_OK;
}

int mxr_nest_host(mxr_handle_t hh, const void* dets, int n, int dtype, double thresh, int32_t* keep_flag) {
  MXR_REQUIRE(hh != nullptr, MXR_ERR_INVALID, "mxr_nest_host: NULL handle");
  if (n <= 0) return n < 0 ? MXR_ERR_INVALID : MXR_OK;
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  const size_t db = (size_t)n * 5 * (dtype == MXR_DTYPE_F64 ? 8 : 4);
  const size_t off = align_up(db, 256);
  MXR_TRY(ensure_device_ws(h, off + (size_t)n * 4));
  MXR_CUDA(cudaMemcpyAsync(h->d_ws, dets, db, cudaMemcpyHostToDevice, h->stream));
  MXR_TRY(mxr_nest(hh, h->stream, h->d_ws, n, dtype, thresh, reinterpret_cast<int32_t*>(h->d_ws + off)));
  MXR_CUDA(cudaMemcpyAsync(keep_flag, h->d_ws + off, (size_t)n * 4, cudaMemcpyDeviceToHost, h->stream));
  MXR_CUDA(cudaStreamSynchronize(h->stream));
  return MXR_OK;
}

}  // extern "C"

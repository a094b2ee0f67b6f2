// helper/processing/nms.py:40-70 -- nest(dets, thresh): drop box i when, for some other box j,
// inter(i,j) / area(i) > thresh (box i lies almost entirely inside box j).  O(n^2), eight threads per box,
// the other boxes staged through shared memory; arithmetic in the dtype of dets with IEEE division.
#include <string.h>

#include "device_math.cuh"

using namespace mxr;

namespace {

// 256 threads: lane = one of 32 boxes i of the CTA, warp = one of 8 interleaved shares of the other boxes j.
template <typename T>
__global__ void __launch_bounds__(256) nest_kernel(const T* __restrict__ dets, int n, T thr, int* __restrict__ keep_flag) {
  typedef Ar<T> A;
  __shared__ T sb[256][4];
  __shared__ int s_drop[32];
  const int lane = threadIdx.x & 31, share = threadIdx.x >> 5;
  const int i = blockIdx.x * 32 + lane;
  const bool live = i < n;
  T x1 = 0, y1 = 0, x2 = 0, y2 = 0, area = 1;
  if (live) {
    x1 = dets[(size_t)i * 5];
    y1 = dets[(size_t)i * 5 + 1];
    x2 = dets[(size_t)i * 5 + 2];
    y2 = dets[(size_t)i * 5 + 3];
    area = box_area<T>(x1, y1, x2, y2);  // :49
  }
  if (threadIdx.x < 32) s_drop[threadIdx.x] = 0;
  // a pair without intersection gives 0 / area, which is "> thresh" only for odd inputs (negative threshold, area <= 0,
  // NaN): with those excluded the division is done for intersecting pairs only
  const bool skip_empty = area > T(0) && thr >= T(0);
  bool drop = false;
  for (int j0 = 0; j0 < n; j0 += 256) {
    const int jn = min(256, n - j0);
    __syncthreads();
    if ((int)threadIdx.x < jn)
      for (int k = 0; k < 4; ++k) sb[threadIdx.x][k] = dets[(size_t)(j0 + threadIdx.x) * 5 + k];
    __syncthreads();
    if (live) {
      for (int j = share; j < jn; j += 8) {
        if (j0 + j == i) continue;  // :56
        const T xx1 = A::vmax(x1, sb[j][0]), yy1 = A::vmax(y1, sb[j][1]);
        const T xx2 = A::vmin(x2, sb[j][2]), yy2 = A::vmin(y2, sb[j][3]);
        const T w = A::vmax(T(0), A::add(A::sub(xx2, xx1), T(1)));
        const T h = A::vmax(T(0), A::add(A::sub(yy2, yy1), T(1)));
        const T inter = A::mul(w, h);
        if (skip_empty && inter == T(0)) continue;
        if (A::div(inter, area) > thr) drop = true;  // :63-64
      }
    }
  }
  if (drop) s_drop[lane] = 1;
  __syncthreads();
  if (share == 0 && live) keep_flag[i] = s_drop[lane] ? 0 : 1;
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
    nest_kernel<double><<<(n + 31) / 32, 256, 0, s>>>((const double*)dets, n, thresh, keep_flag);
  else if (dtype == MXR_DTYPE_F32)
    nest_kernel<float><<<(n + 31) / 32, 256, 0, s>>>((const float*)dets, n, (float)thresh, keep_flag);
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

// Measured non-tensor ALU peaks for the roofline of the NMS kernels ("ALU peak measured by an
// in-repo micro-benchmark at -fmad=false"): streams of independent adds / multiplies, no FMA contraction.
#include "common.cuh"

using namespace mxr;

namespace {

template <typename T>
__global__ void __launch_bounds__(256) alu_stream_kernel(T* out, int iters, T seed) {
  T a[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = seed + (T)(threadIdx.x + i);
  const T m = (T)1.0000001, c = (T)0.5;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      a[i] = a[i] * m;   // -fmad=false: stays a separate multiply ...
      a[i] = a[i] + c;   // ... and a separate add
    }
  }
  T s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += a[i];
  if (s == (T)12345.678) out[0] = s;  // keep the chains alive
}

template <typename T>
int run_stream(Handle* h, int iters, double* gops) {
  T* d = reinterpret_cast<T*>(h->d_ws);
  int sms = 0;
  MXR_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device));
  const int grid = sms * 8;
  cudaEvent_t e0, e1;
  MXR_CUDA(cudaEventCreate(&e0));
  MXR_CUDA(cudaEventCreate(&e1));
  alu_stream_kernel<T><<<grid, 256, 0, h->stream>>>(d, iters, (T)1.0);
  MXR_CUDA(cudaEventRecord(e0, h->stream));
  alu_stream_kernel<T><<<grid, 256, 0, h->stream>>>(d, iters, (T)1.0);
  MXR_CUDA(cudaEventRecord(e1, h->stream));
  MXR_CUDA(cudaEventSynchronize(e1));
  float ms = 0.f;
  MXR_CUDA(cudaEventElapsedTime(&ms, e0, e1));
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  const double ops = (double)grid * 256.0 * 16.0 * (double)iters;  // 8 chains x (mul + add)
  *gops = ops / (ms * 1e-3) * 1e-9;
  return MXR_OK;
}

}  // namespace

extern "C" int mxr_bench_alu_peak(mxr_handle_t hh, int dtype, double* gops_out) {
  MXR_REQUIRE(hh != nullptr && gops_out != nullptr, MXR_ERR_INVALID, "mxr_bench_alu_peak: NULL argument");
  Handle* h = reinterpret_cast<Handle*>(hh);
  MXR_DEVICE(h);
  MXR_TRY(ensure_device_ws(h, 4096));
  if (dtype == MXR_DTYPE_F64) return run_stream<double>(h, 4096, gops_out);
  if (dtype == MXR_DTYPE_F32) return run_stream<float>(h, 16384, gops_out);
  set_error("mxr_bench_alu_peak: dtype must be f32 or f64");
  return MXR_ERR_INVALID;
}

// Standalone micro-benchmark: per-SM throughput of the instruction kinds in the IoU predicate (FP64 add / mul /
// compare-select, FP32 counterparts).  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o dp_rate tools/dp_rate.cu
// Throughput of the instruction kinds in the IoU predicate, per SM: independent chains, 1024 threads x 148 CTAs.
#include <cstdio>
#include <cuda_runtime.h>
template <int KIND>
__global__ void k(double* out, int iters, double seed) {
  double a[8];
  for (int i = 0; i < 8; ++i) a[i] = seed + threadIdx.x * 1e-3 + i;
  double b = seed * 0.5;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (KIND == 0) a[i] = __dadd_rn(a[i], b);
      if (KIND == 1) a[i] = __dmul_rn(a[i], b);
      if (KIND == 2) a[i] = a[i] > b ? a[i] : b + i;          // DSETP + 64-bit select (compare-select max)
      if (KIND == 3) a[i] = fmax(a[i], b + i);                 // library fmax
      if (KIND == 4) { float f = (float)a[i]; f = fmaxf(f, (float)b); a[i] = f; }
      if (KIND == 5) a[i] = (a[i] <= b) ? 1.0 : a[i];
    }
    b += 1e-9;
  }
  double s = 0;
  for (int i = 0; i < 8; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int KIND>
__global__ void kf(float* out, int iters, float seed) {
  float a[8];
  for (int i = 0; i < 8; ++i) a[i] = seed + threadIdx.x * 1e-3f + i;
  float b = seed * 0.5f;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (KIND == 0) a[i] = __fadd_rn(a[i], b);
      if (KIND == 1) a[i] = fmaxf(a[i], b + i);
      if (KIND == 2) a[i] = a[i] > b ? a[i] : b + i;
    }
    b += 1e-6f;
  }
  float s = 0;
  for (int i = 0; i < 8; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <typename F>
double timeit(F f) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}
int main() {
  double* o; cudaMalloc(&o, 148 * 1024 * 8);
  const int iters = 20000; const double ops = 148.0 * 1024 * 8 * iters;
  const char* names[] = {"DADD", "DMUL", "DSETP+SEL max", "fmax(double)", "cvt+fmaxf+cvt", "DSETP+SEL const"};
  double ms;
  ms = timeit([&] { k<0><<<148, 1024>>>(o, iters, 1.5); }); printf("%-18s %7.1f Gop/s  %5.1f lane-ops/clk/SM\n", names[0], ops / ms / 1e6, ops / ms / 1e6 / 148 / 1.965);
  ms = timeit([&] { k<1><<<148, 1024>>>(o, iters, 1.0000001); }); printf("%-18s %7.1f Gop/s  %5.1f\n", names[1], ops / ms / 1e6, ops / ms / 1e6 / 148 / 1.965);
  ms = timeit([&] { k<2><<<148, 1024>>>(o, iters, 1.5); }); printf("%-18s %7.1f Gop/s  %5.1f\n", names[2], ops / ms / 1e6, ops / ms / 1e6 / 148 / 1.965);
  ms = timeit([&] { k<3><<<148, 1024>>>(o, iters, 1.5); }); printf("%-18s %7.1f Gop/s  %5.1f\n", names[3], ops / ms / 1e6, ops / ms / 1e6 / 148 / 1.965);
  ms = timeit([&] { k<4><<<148, 1024>>>(o, iters, 1.5); }); printf("%-18s %7.1f Gop/s  %5.1f\n", names[4], ops / ms / 1e6, ops / ms / 1e6 / 148 / 1.965);
  ms = timeit([&] { k<5><<<148, 1024>>>(o, iters, 1.5); }); printf("%-18s %7.1f Gop/s  %5.1f\n", names[5], ops / ms / 1e6, ops / ms / 1e6 / 148 / 1.965);
  float* of = (float*)o;
  ms = timeit([&] { kf<0><<<148, 1024>>>(of, iters, 1.5f); }); printf("%-18s %7.1f Gop/s  %5.1f\n", "FADD", ops / ms / 1e6, ops / ms / 1e6 / 148 / 1.965);
  ms = timeit([&] { kf<1><<<148, 1024>>>(of, iters, 1.5f); }); printf("%-18s %7.1f Gop/s  %5.1f\n", "fmaxf", ops / ms / 1e6, ops / ms / 1e6 / 148 / 1.965);
  ms = timeit([&] { kf<2><<<148, 1024>>>(of, iters, 1.5f); }); printf("%-18s %7.1f Gop/s  %5.1f\n", "FSETP+SEL max", ops / ms / 1e6, ops / ms / 1e6 / 148 / 1.965);
  return 0;
}

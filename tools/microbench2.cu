// Do the fp64 vector pipe (DFMA) and the fp64 tensor pipe (DMMA) share execution resources on B200?
// Even warps issue DMMA, odd warps issue DFMA; compare with each alone.
#include <cstdio>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "CUDA error %s\n", cudaGetErrorString(e_)); return 1; } } while (0)
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__global__ void __launch_bounds__(512) k_mix(double* out, int iters, double a, double b, int mode) {
  double r[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) r[i] = threadIdx.x * 1e-3 + i;
  const int warp = threadIdx.x >> 5;
  const bool do_mma = mode == 0 || (mode == 2 && (warp & 1) == 0);
  const bool do_fma = mode == 1 || (mode == 2 && (warp & 1) == 1);
  if (do_mma) for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) dmma884(r[2 * i], r[2 * i + 1], a, b);
  }
  if (do_fma) for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) r[i] = fma(r[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += r[i];
  if (s == 123.456) out[0] = s;
}
int main() {
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
  double* out; CK(cudaMalloc(&out, 64));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const int iters = 20000; dim3 g(prop.multiProcessorCount * 2), b(512);
  for (int mode = 0; mode < 3; ++mode) {
    k_mix<<<g, b>>>(out, 100, 1.0000001, 1e-9, mode); CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0)); k_mix<<<g, b>>>(out, iters, 1.0000001, 1e-9, mode); CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    const double warps = (double)g.x * 16;
    const double fl_mma = 2.0 * 256 * 8 * iters * (mode == 0 ? warps : (mode == 2 ? warps / 2 : 0));
    const double fl_fma = 2.0 * 16 * 32 * iters * (mode == 1 ? warps : (mode == 2 ? warps / 2 : 0));
    printf("{\"mode\": \"%s\", \"ms\": %.3f, \"dmma_tflops\": %.2f, \"dfma_tflops\": %.2f, \"total_tflops\": %.2f}\n",
           mode == 0 ? "dmma only" : (mode == 1 ? "dfma only" : "half warps dmma, half dfma"), ms, fl_mma / ms * 1e-9, fl_fma / ms * 1e-9, (fl_mma + fl_fma) / ms * 1e-9);
  }
  return 0;
}

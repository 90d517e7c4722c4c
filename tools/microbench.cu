// Standalone B200 microbenchmarks that size the KMeans hot-path design:
//   fp64 DFMA peak, fp64 DMMA (mma.sync m8n8k4 / m16n8k8) peak,
//   streaming-read bandwidth with plain LDG.128 and with a cp.async.bulk + mbarrier ring.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tools/microbench tools/microbench.cu
// Not part of the product library; results are recorded under profiles/.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1);} } while (0)

// ---------------------------------------------------------------- DFMA peak
__global__ void __launch_bounds__(256) k_dfma(double* out, int iters, double a, double b) {
  double r[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) r[i] = threadIdx.x * 1e-3 + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) r[i] = fma(r[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += r[i];
  if (s == 123.456) out[0] = s;
}

// ---------------------------------------------------------------- DMMA peak
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__global__ void __launch_bounds__(256) k_dmma884(double* out, int iters, double a, double b) {
  double c[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) c[i] = threadIdx.x * 1e-3 + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) dmma884(c[2 * i], c[2 * i + 1], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += c[i];
  if (s == 123.456) out[0] = s;
}
__device__ __forceinline__ void dmma1688(double* c, const double* a, const double* b) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
               : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
}
__global__ void __launch_bounds__(256) k_dmma1688(double* out, int iters, double a, double b) {
  double c[16], av[4] = {a, a + 1, a + 2, a + 3}, bv[2] = {b, b + 1};
#pragma unroll
  for (int i = 0; i < 16; ++i) c[i] = threadIdx.x * 1e-3 + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 4; ++i) dmma1688(c + 4 * i, av, bv);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += c[i];
  if (s == 123.456) out[0] = s;
}
__device__ __forceinline__ void dmma1684(double* c, const double* a, double b) {
  asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
               : "d"(a[0]), "d"(a[1]), "d"(b));
}
__global__ void __launch_bounds__(256) k_dmma1684(double* out, int iters, double a, double b) {
  double c[16], av[2] = {a, a + 1};
#pragma unroll
  for (int i = 0; i < 16; ++i) c[i] = threadIdx.x * 1e-3 + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 4; ++i) dmma1684(c + 4 * i, av, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += c[i];
  if (s == 123.456) out[0] = s;
}

// ---------------------------------------------------------------- LDG.128 streaming read
__global__ void __launch_bounds__(256) k_ldg(const double2* __restrict__ x, size_t n2, double* out) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  double s = 0;
  for (; i + 3 * stride < n2; i += 4 * stride) {
    double2 a = __ldcs(x + i), b = __ldcs(x + i + stride), c = __ldcs(x + i + 2 * stride), d = __ldcs(x + i + 3 * stride);
    s += a.x + a.y + b.x + b.y + c.x + c.y + d.x + d.y;
  }
  for (; i < n2; i += stride) { double2 a = __ldcs(x + i); s += a.x + a.y; }
  if (s == 123.456) out[0] = s;
}

// ---------------------------------------------------------------- cp.async.bulk ring
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t cnt) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(cnt));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* b) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
  asm volatile(
      "{\n.reg .pred p;\nWAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

template <int STAGES>
__global__ void __launch_bounds__(288) k_bulk(const char* __restrict__ x, size_t ntiles, int tile_bytes, double* out) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ uint64_t full[STAGES], empty[STAGES];
  const int tid = threadIdx.x;
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 8); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int warp = tid >> 5, lane = tid & 31;
  if (warp == 8) {
    if (lane == 0) {
      int s = 0; uint32_t ph = 0;
      for (size_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
        mbar_wait(&empty[s], ph ^ 1);
        mbar_expect_tx(&full[s], tile_bytes);
        bulk_g2s(smem + (size_t)s * tile_bytes, x + t * (size_t)tile_bytes, tile_bytes, &full[s]);
        if (++s == STAGES) { s = 0; ph ^= 1; }
      }
    }
  } else {
    int s = 0; uint32_t ph = 0; double acc = 0;
    for (size_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
      mbar_wait(&full[s], ph);
      const double2* p = reinterpret_cast<const double2*>(smem + (size_t)s * tile_bytes);
      for (int i = tid; i < tile_bytes / 16; i += 256) { double2 v = p[i]; acc += v.x + v.y; }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[s]);
      if (++s == STAGES) { s = 0; ph ^= 1; }
    }
    if (acc == 123.456) out[0] = acc;
  }
}

static float time_ms(cudaEvent_t a, cudaEvent_t b) { float ms; CK(cudaEventElapsedTime(&ms, a, b)); return ms; }

int main(int argc, char** argv) {
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
  int sms = prop.multiProcessorCount;
  printf("{\"device\": \"%s\", \"sms\": %d, \"cc\": \"%d.%d\"}\n", prop.name, sms, prop.major, prop.minor);
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  double* out; CK(cudaMalloc(&out, 64));

  // ---- fp64 peaks
  {
    const int iters = 20000;
    for (int bps = 1; bps <= 8; bps *= 2) {
      dim3 g(sms * bps), b(256);
      k_dfma<<<g, b>>>(out, 100, 1.0000001, 1e-9); CK(cudaDeviceSynchronize());
      CK(cudaEventRecord(e0)); k_dfma<<<g, b>>>(out, iters, 1.0000001, 1e-9); CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
      float ms = time_ms(e0, e1);
      double fl = 2.0 * 16 * iters * (double)g.x * b.x;
      printf("{\"bench\": \"dfma\", \"blocks_per_sm\": %d, \"ms\": %.3f, \"tflops\": %.2f}\n", bps, ms, fl / ms * 1e-9);
      k_dmma884<<<g, b>>>(out, 100, 1.0000001, 1e-9); CK(cudaDeviceSynchronize());
      CK(cudaEventRecord(e0)); k_dmma884<<<g, b>>>(out, iters, 1.0000001, 1e-9); CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
      ms = time_ms(e0, e1);
      fl = 2.0 * 256 * 8 * iters * (double)g.x * (b.x / 32);
      printf("{\"bench\": \"dmma_m8n8k4\", \"blocks_per_sm\": %d, \"ms\": %.3f, \"tflops\": %.2f}\n", bps, ms, fl / ms * 1e-9);
      k_dmma1688<<<g, b>>>(out, 100, 1.0000001, 1e-9); CK(cudaDeviceSynchronize());
      CK(cudaEventRecord(e0)); k_dmma1688<<<g, b>>>(out, iters, 1.0000001, 1e-9); CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
      ms = time_ms(e0, e1);
      fl = 2.0 * 1024 * 4 * iters * (double)g.x * (b.x / 32);
      printf("{\"bench\": \"dmma_m16n8k8\", \"blocks_per_sm\": %d, \"ms\": %.3f, \"tflops\": %.2f}\n", bps, ms, fl / ms * 1e-9);
      k_dmma1684<<<g, b>>>(out, 100, 1.0000001, 1e-9); CK(cudaDeviceSynchronize());
      CK(cudaEventRecord(e0)); k_dmma1684<<<g, b>>>(out, iters, 1.0000001, 1e-9); CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
      ms = time_ms(e0, e1);
      fl = 2.0 * 512 * 4 * iters * (double)g.x * (b.x / 32);
      printf("{\"bench\": \"dmma_m16n8k4\", \"blocks_per_sm\": %d, \"ms\": %.3f, \"tflops\": %.2f}\n", bps, ms, fl / ms * 1e-9);
    }
  }

  // ---- streaming read bandwidth over 4 GiB (> L2)
  {
    size_t bytes = (size_t)4 << 30;
    char* x; CK(cudaMalloc(&x, bytes)); CK(cudaMemset(x, 0, bytes));
    for (int bps = 2; bps <= 8; bps *= 2) {
      dim3 g(sms * bps), b(256);
      k_ldg<<<g, b>>>((const double2*)x, bytes / 16, out); CK(cudaDeviceSynchronize());
      float best = 1e30f;
      for (int r = 0; r < 5; ++r) {
        CK(cudaEventRecord(e0)); k_ldg<<<g, b>>>((const double2*)x, bytes / 16, out); CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
        best = fminf(best, time_ms(e0, e1));
      }
      printf("{\"bench\": \"ldg128_read\", \"blocks_per_sm\": %d, \"ms\": %.3f, \"GBps\": %.1f}\n", bps, best, bytes / best * 1e-6);
    }
    const int tile_sizes[] = {8192, 16384, 32768};
    for (int ti = 0; ti < 3; ++ti) {
      int tb = tile_sizes[ti];
      size_t ntiles = bytes / tb;
      for (int cfg = 0; cfg < 3; ++cfg) {
        // cfg 0: 1 CTA/SM deep ring; cfg 1: 2 CTA/SM; cfg 2: 3 CTA/SM
        int bps = cfg + 1;
        int smem_budget = 200 * 1024 / bps;
        int stages = smem_budget / tb; if (stages > 8) stages = 8; if (stages < 2) continue;
        size_t smem = (size_t)stages * tb;
        float best = 1e30f;
        for (int r = 0; r < 6; ++r) {
          CK(cudaEventRecord(e0));
          dim3 g(sms * bps), b(288);
#define LAUNCH(S) case S: CK(cudaFuncSetAttribute(k_bulk<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
                  k_bulk<S><<<g, b, smem>>>(x, ntiles, tb, out); break;
          switch (stages) { LAUNCH(2) LAUNCH(3) LAUNCH(4) LAUNCH(5) LAUNCH(6) LAUNCH(7) LAUNCH(8) }
          CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
          if (r > 0) best = fminf(best, time_ms(e0, e1));
        }
        printf("{\"bench\": \"bulk_ring_read\", \"tile_bytes\": %d, \"stages\": %d, \"ctas_per_sm\": %d, \"ms\": %.3f, \"GBps\": %.1f}\n",
               tb, stages, bps, best, bytes / best * 1e-6);
      }
    }
    // plain device copy for the same denominator the driver used
    char* y; CK(cudaMalloc(&y, bytes / 2));
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
      CK(cudaEventRecord(e0)); CK(cudaMemcpyAsync(y, x, bytes / 2, cudaMemcpyDeviceToDevice)); CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
      best = fminf(best, time_ms(e0, e1));
    }
    printf("{\"bench\": \"memcpy_d2d\", \"ms\": %.3f, \"GBps_rw\": %.1f}\n", best, (double)bytes / best * 1e-6);
    CK(cudaFree(x)); CK(cudaFree(y));
  }
  return 0;
}

// Probe: how does tcgen05 kind::tf32 treat the 13 low mantissa bits of its fp32 operands (truncate or round)?
// Bring-up kernel of tools/tcgen05_test.cu: D[128x64] = A[128x64] * B[64x64]^T with operands written
// into shared memory by ordinary threads in the canonical K-major no-swizzle layout, accumulator in TMEM.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/tcgen05_test tools/tcgen05_test.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cmath>
#include <vector>
#include <cstring>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1);} } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// K-major, no swizzle: element (row r, 16-byte chunk c) at (r%8)*16 + (r/8)*SBO + c*LBO bytes
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;   // descriptor version (Blackwell)
  return d;                 // layout_type = 0 (no swizzle), base_offset = 0
}

constexpr int M = 128, N = 64, K = 64;

__global__ void __launch_bounds__(128) k_test(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ D, int variant) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ uint64_t mbar;
  __shared__ uint32_t tmem_base_s;
  float* sA = reinterpret_cast<float*>(smem);                 // 128 x 64 tf32 = 32 KB
  float* sB = reinterpret_cast<float*>(smem + M * K * 4);     // 64 x 64 = 16 KB
  const int tid = threadIdx.x, warp = tid >> 5;
  const uint32_t LBO_A = (M / 8) * 128, LBO_B = (N / 8) * 128, SBO = 128;

  // fill operands in the canonical layout: thread handles (row, chunk) pairs
  for (int i = tid; i < M * (K / 4); i += 128) {
    const int r = i % M, c = i / M;
    const float4 v = *reinterpret_cast<const float4*>(A + r * K + 4 * c);
    *reinterpret_cast<float4*>(reinterpret_cast<unsigned char*>(sA) + c * LBO_A + (r / 8) * SBO + (r % 8) * 16) = v;
  }
  for (int i = tid; i < N * (K / 4); i += 128) {
    const int r = i % N, c = i / N;
    const float4 v = *reinterpret_cast<const float4*>(B + r * K + 4 * c);
    *reinterpret_cast<float4*>(reinterpret_cast<unsigned char*>(sB) + c * LBO_B + (r / 8) * SBO + (r % 8) * 16) = v;
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mbar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(smem_u32(&tmem_base_s)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy smem writes -> visible to the tensor core (async proxy)
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_base_s;

  if (tid == 0) {
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
    const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
    for (int ks = 0; ks < K / 8; ++ks) {        // one UMMA consumes K = 8 tf32 = two 16-byte chunks
      const uint64_t da = make_desc(a0 + ks * 2 * LBO_A, LBO_A, SBO);
      const uint64_t db = make_desc(b0 + ks * 2 * LBO_B, LBO_B, SBO);
      const uint32_t acc = ks > 0 ? 1u : 0u;
      asm volatile(
          "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
          "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem), "l"(da), "l"(db), "r"(idesc), "r"(acc)
          : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&mbar)) : "memory");
  }
  // everyone waits for the MMAs
  asm volatile(
      "{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(smem_u32(&mbar))
      : "memory");
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

  // each warp reads its 32 TMEM lanes (= rows), 64 columns
  uint32_t v[64];
  const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16);
#define LD16(off) asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];" \
    : "=r"(v[off+0]), "=r"(v[off+1]), "=r"(v[off+2]), "=r"(v[off+3]), "=r"(v[off+4]), "=r"(v[off+5]), "=r"(v[off+6]), "=r"(v[off+7]), \
      "=r"(v[off+8]), "=r"(v[off+9]), "=r"(v[off+10]), "=r"(v[off+11]), "=r"(v[off+12]), "=r"(v[off+13]), "=r"(v[off+14]), "=r"(v[off+15]) \
    : "r"(taddr + off))
  LD16(0); LD16(16); LD16(32); LD16(48);
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  const int row = warp * 32 + (tid & 31);
  for (int j = 0; j < N; ++j) D[row * N + j] = __uint_as_float(v[j]);

  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(tmem) : "memory");
  (void)variant;
}


int main() {
  std::vector<float> A(M * K, 0.f), B(N * K, 0.f), D(M * N);
  auto bits = [](uint32_t u) { float f; memcpy(&f, &u, 4); return f; };
  // 1.0f = 0x3f800000; one tf32 ulp at 1.0 is 2^-10 = 0x2000 in the fp32 mantissa
  const uint32_t probes[8] = {0x3f800000u + 0x1000u + 0x0200u,   // 1 + half ulp + a bit      (RN -> 1 + ulp, truncation -> 1)
                              0x3f800000u + 0x1000u - 0x0200u,   // 1 + half ulp - a bit      (both -> 1)
                              0x3f800000u + 0x1fffu,             // just under one ulp         (RN -> 1 + ulp, truncation -> 1)
                              0x3f800000u + 0x0800u,             // quarter ulp                (both -> 1)
                              0x3f800000u + 0x2000u + 0x1800u,   // 1 + 1.75 ulp               (RN -> 1 + 2 ulp, truncation -> 1 + ulp)
                              0xbf800000u + 0x1fffu,             // -(just under one ulp more) (sign-magnitude truncation -> -1)
                              0x3f800000u + 0x1000u,             // exactly half               (ties)
                              0x3f800000u};
  for (int r = 0; r < 8; ++r) { A[r * K + 0] = bits(probes[r]); }          // A-side probes against B = 1
  for (int j = 0; j < N; ++j) B[j * K + 0] = 0.f;
  B[0 * K + 0] = 1.0f;
  for (int j = 1; j < 9; ++j) { B[j * K + 1] = bits(probes[j - 1]); }      // B-side probes against A[., 1] = 1 in row 8
  A[8 * K + 1] = 1.0f;
  float *dA, *dB, *dD;
  CK(cudaMalloc(&dA, A.size() * 4)); CK(cudaMalloc(&dB, B.size() * 4)); CK(cudaMalloc(&dD, D.size() * 4));
  CK(cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice));
  const int smem = M * K * 4 + N * K * 4;
  CK(cudaFuncSetAttribute(k_test, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  k_test<<<1, 128, smem>>>(dA, dB, dD, 0);
  CK(cudaGetLastError()); CK(cudaDeviceSynchronize());
  CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
  for (int r = 0; r < 8; ++r) { uint32_t u; memcpy(&u, &D[r * N + 0], 4); printf("A probe %d: in 0x%08x -> out 0x%08x (%.9g)\n", r, probes[r], u, D[r * N + 0]); }
  for (int j = 1; j < 9; ++j) { uint32_t u; memcpy(&u, &D[8 * N + j], 4); printf("B probe %d: in 0x%08x -> out 0x%08x (%.9g)\n", j - 1, probes[j - 1], u, D[8 * N + j]); }
  return 0;
}

// Bring-up test for the integer "limb" reduction on tcgen05: D[128 x N] (s32, TMEM) += A[128 x K]^(MN-major) * B[N x K]^T with
// A = the bytes of a tile of rows (K = 128 rows, M = 128 bytes per row, stored row by row with the 128-byte swizzle: exactly what a
// SWIZZLE_128B tensor-map load of 128-byte rows leaves in shared memory) and B = one-hot bytes (cluster x row), K-major without
// swizzle.  Variants probe the descriptor fields (K advance, swizzle on/off).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/i8_limb_test tools/i8_limb_test.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1);} } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

constexpr int M = 128, K = 128;   // M = bytes per row, K = rows

// layout_type: 0 none, 2 = SWIZZLE_128B
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout_type) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)layout_type << 61;
  return d;
}

template <int N>
__global__ void __launch_bounds__(128) k_test(const uint8_t* __restrict__ A, const uint8_t* __restrict__ B, int32_t* __restrict__ D, int variant) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ uint64_t mbar;
  __shared__ uint32_t tmem_base_s;
  unsigned char* sA = smem;                       // 128 rows x 128 bytes, swizzled
  unsigned char* sB = smem + K * M;               // one-hot: chunk c (16 rows) at c * LBO_B, cluster n at n * 16, row r%16
  const int tid = threadIdx.x, warp = tid >> 5;
  const uint32_t LBO_B = (N / 8) * 128 + 16, SBO_B = 128;
  // A: logical byte (row k, byte m) -> k*128 + ((m/16) ^ (k%8))*16 + m%16
  for (int i = tid; i < K * M; i += 128) {
    const int k = i / M, m = i % M;
    const int chunk = variant == 2 ? (m / 16) : ((m / 16) ^ (k % 8));
    sA[k * 128 + chunk * 16 + (m % 16)] = A[k * M + m];
  }
  for (int i = tid; i < N * K; i += 128) {
    const int n = i / K, k = i % K;
    sB[(k / 16) * LBO_B + n * 16 + (k % 16)] = B[n * K + k];
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mbar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(smem_u32(&tmem_base_s)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_base_s;
  if (tid == 0) {
    // c_format S32 (2), a/b format u8 (0), A MN-major (bit 15), B K-major, N >> 3, M >> 4
    const uint32_t idesc = (2u << 4) | (1u << 15) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    for (int ks = 0; ks < K / 32; ++ks) {
      // A: 32 rows per instruction = four 8-row atoms of 1024 bytes; B: two 16-byte chunks per instruction
      const uint64_t da = make_desc(smem_u32(sA) + ks * 4096, variant == 1 ? 128 : 16, 1024, variant == 2 ? 0 : 2);
      const uint64_t db = make_desc(smem_u32(sB) + 2 * ks * LBO_B, LBO_B, SBO_B, 0);
      const uint32_t acc = ks > 0 ? 1u : 0u;
      asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem), "l"(da), "l"(db),
                   "r"(idesc), "r"(acc)
                   : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&mbar)) : "memory");
  }
  // wait for the commit
  {
    uint32_t ok = 0;
    while (!ok) asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(ok) : "r"(smem_u32(&mbar)), "r"(0u) : "memory");
  }
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  // lane = M index (byte of the row), columns = clusters
  for (int c0 = 0; c0 < N; c0 += 16) {
    uint32_t v[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]),
                   "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(tmem + ((uint32_t)(32 * warp) << 16) + (uint32_t)c0));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int c = 0; c < 16; ++c) D[tid * N + c0 + c] = (int32_t)v[c];
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(tmem) : "memory");
}

template <int N>
static int run(int variant) {
  std::vector<uint8_t> hA(K * M), hB(N * K);
  srand(7 + variant);
  for (auto& v : hA) v = (uint8_t)(rand() & 255);
  for (int n = 0; n < N; ++n) for (int k = 0; k < K; ++k) hB[n * K + k] = (uint8_t)((rand() % N) == n ? 1 : ((rand() & 31) == 0 ? 3 : 0));
  uint8_t *dA, *dB; int32_t* dD;
  CK(cudaMalloc(&dA, hA.size())); CK(cudaMalloc(&dB, hB.size())); CK(cudaMalloc(&dD, M * N * 4));
  CK(cudaMemcpy(dA, hA.data(), hA.size(), cudaMemcpyHostToDevice)); CK(cudaMemcpy(dB, hB.data(), hB.size(), cudaMemcpyHostToDevice));
  CK(cudaMemset(dD, 0xff, M * N * 4));
  const int smem = K * M + 8 * ((N / 8) * 128 + 16) + 1024;
  CK(cudaFuncSetAttribute(k_test<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  k_test<N><<<1, 128, smem>>>(dA, dB, dD, variant);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("N=%d variant %d: launch error %s\n", N, variant, cudaGetErrorString(e)); return 1; }
  std::vector<int32_t> hD(M * N);
  CK(cudaMemcpy(hD.data(), dD, M * N * 4, cudaMemcpyDeviceToHost));
  long bad = 0; int first = -1;
  for (int m = 0; m < M; ++m) for (int n = 0; n < N; ++n) {
    long ref = 0;
    for (int k = 0; k < K; ++k) ref += (long)hA[k * M + m] * hB[n * K + k];
    if (ref != hD[m * N + n]) { if (first < 0) first = m * N + n; ++bad; }
  }
  printf("N=%d variant %d: %ld of %d wrong", N, variant, bad, M * N);
  if (first >= 0) printf(" (first at m=%d n=%d: got %d)", first / N, first % N, hD[first]);
  printf("\n");
  cudaFree(dA); cudaFree(dB); cudaFree(dD);
  return bad != 0;
}

// Four issuing threads, one per warp, each adding its own K slice (32 rows) into the SAME accumulator, `rounds` times without any
// ordering between them: integer accumulation is order free, the question is whether concurrent accumulating MMAs from different
// threads are applied atomically by the tensor pipe.
template <int N>
__global__ void __launch_bounds__(128) k_test4(const uint8_t* __restrict__ A, const uint8_t* __restrict__ B, int32_t* __restrict__ D, int rounds,
                                                 const int32_t* __restrict__ ref, unsigned int* __restrict__ errs) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ uint64_t mbar[4];
  __shared__ uint32_t tmem_base_s;
  unsigned char* sA = smem;
  unsigned char* sB = smem + K * M;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t LBO_B = (N / 8) * 128 + 16, SBO_B = 128;
  for (int i = tid; i < K * M; i += 128) { const int k = i / M, m = i % M; sA[k * 128 + (((m / 16) ^ (k % 8)) * 16) + (m % 16)] = A[k * M + m]; }
  for (int i = tid; i < N * K; i += 128) { const int n = i / K, k = i % K; sB[(k / 16) * LBO_B + n * 16 + (k % 16)] = B[n * K + k]; }
  if (tid == 0) {
    for (int w = 0; w < 4; ++w) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mbar[w])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(smem_u32(&tmem_base_s)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_base_s;
  for (int c0 = 0; c0 < N; c0 += 8)     // zero the accumulator: every warp its lane quarter
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1};" ::"r"(tmem + ((uint32_t)(32 * warp) << 16) + (uint32_t)c0), "r"(0u) : "memory");
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t idesc = (2u << 4) | (1u << 15) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  if (lane == 0) {
    const uint64_t da = make_desc(smem_u32(sA) + warp * 4096, 16, 1024, 2);
    const uint64_t db = make_desc(smem_u32(sB) + 2 * warp * LBO_B, LBO_B, SBO_B, 0);
    for (int r = 0; r < rounds; ++r) {
      asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem), "l"(da), "l"(db), "r"(idesc), "r"(1u) : "memory");
      if ((r & 3) == warp) {   // commits and waits at different times in the four threads: the streams drift apart
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&mbar[warp])) : "memory");
        uint32_t ok = 0; const uint32_t par = (uint32_t)((r >> 2) & 1);
        while (!ok) asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(ok) : "r"(smem_u32(&mbar[warp])), "r"(par) : "memory");
      }
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&mbar[warp])) : "memory");
    const int commits = (rounds + 3 - warp) / 4;   // rounds r with r % 4 == warp
    uint32_t ok = 0; const uint32_t par = (uint32_t)(commits & 1);
    while (!ok) asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(ok) : "r"(smem_u32(&mbar[warp])), "r"(par) : "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  for (int c0 = 0; c0 < N; c0 += 16) {
    uint32_t v[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]),
                   "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(tmem + ((uint32_t)(32 * warp) << 16) + (uint32_t)c0));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int c = 0; c < 16; ++c) { D[tid * N + c0 + c] = (int32_t)v[c]; if ((int32_t)v[c] != ref[tid * N + c0 + c] * rounds) atomicAdd(errs, 1u); }   // every CTA checks its own result
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(tmem) : "memory");
}

template <int N>
static int run4(int rounds) {
  std::vector<uint8_t> hA(K * M), hB(N * K);
  srand(11);
  for (auto& v : hA) v = (uint8_t)(rand() & 255);
  for (int n = 0; n < N; ++n) for (int k = 0; k < K; ++k) hB[n * K + k] = (uint8_t)((rand() % N) == n ? 1 : 0);
  uint8_t *dA, *dB; int32_t* dD;
  CK(cudaMalloc(&dA, hA.size())); CK(cudaMalloc(&dB, hB.size())); CK(cudaMalloc(&dD, M * N * 4));
  CK(cudaMemcpy(dA, hA.data(), hA.size(), cudaMemcpyHostToDevice)); CK(cudaMemcpy(dB, hB.data(), hB.size(), cudaMemcpyHostToDevice));
  std::vector<int32_t> href(M * N);
  for (int m = 0; m < M; ++m) for (int n = 0; n < N; ++n) { long r = 0; for (int k = 0; k < K; ++k) r += (long)hA[k * M + m] * hB[n * K + k]; href[m * N + n] = (int32_t)r; }
  int32_t* dref; unsigned int* derr; CK(cudaMalloc(&dref, M * N * 4)); CK(cudaMalloc(&derr, 4)); CK(cudaMemset(derr, 0, 4));
  CK(cudaMemcpy(dref, href.data(), M * N * 4, cudaMemcpyHostToDevice));
  const int smem = K * M + 8 * ((N / 8) * 128 + 16) + 1024;
  CK(cudaFuncSetAttribute(k_test4<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  long bad = 0;
  for (int rep = 0; rep < 20; ++rep) {
    k_test4<N><<<148, 128, smem>>>(dA, dB, dD, rounds, dref, derr);      // every SM busy; all CTAs write the same (identical) result
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("concurrent N=%d: launch error %s\n", N, cudaGetErrorString(e)); return 1; }
    std::vector<int32_t> hD(M * N);
    CK(cudaMemcpy(hD.data(), dD, M * N * 4, cudaMemcpyDeviceToHost));
    for (int m = 0; m < M; ++m) for (int n = 0; n < N; ++n) {
      long ref = 0;
      for (int k = 0; k < K; ++k) ref += (long)hA[k * M + m] * hB[n * K + k];
      if (ref * rounds != hD[m * N + n]) ++bad;
    }
  }
  unsigned int herr = 0; CK(cudaMemcpy(&herr, derr, 4, cudaMemcpyDeviceToHost)); bad += herr;
  printf("concurrent issue, N=%d, %d rounds x 4 threads, 20 launches x 148 CTAs: %ld wrong (%u found on device, every CTA checked)\n", N, rounds, bad, herr);
  cudaFree(dA); cudaFree(dB); cudaFree(dD);
  return bad != 0;
}

// M = 64 (one 8-feature slab: 64 bytes per row), SWIZZLE_64B: byte (row k, byte m) at k*64 + ((m/16) ^ ((k>>1)&3))*16 + m%16.
// Prints where each of the 64 result rows lands in tensor memory (lane of which warp).
template <int N>
__global__ void __launch_bounds__(128) k_test64(const uint8_t* __restrict__ A, const uint8_t* __restrict__ B, int32_t* __restrict__ D) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ uint64_t mbar;
  __shared__ uint32_t tmem_base_s;
  unsigned char* sA = smem;                       // 128 rows x 64 bytes
  unsigned char* sB = smem + K * 64;
  const int tid = threadIdx.x, warp = tid >> 5;
  const uint32_t LBO_B = (N / 8) * 128 + 16, SBO_B = 128;
  for (int i = tid; i < K * 64; i += 128) { const int k = i / 64, m = i % 64; sA[k * 64 + (((m / 16) ^ ((k >> 1) & 3)) * 16) + (m % 16)] = A[k * 64 + m]; }
  for (int i = tid; i < N * K; i += 128) { const int n = i / K, k = i % K; sB[(k / 16) * LBO_B + n * 16 + (k % 16)] = B[n * K + k]; }
  if (tid == 0) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mbar))); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(smem_u32(&tmem_base_s)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_base_s;
  for (int c0 = 0; c0 < N; c0 += 8)
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1};" ::"r"(tmem + ((uint32_t)(32 * warp) << 16) + (uint32_t)c0), "r"(0x7fffffffu) : "memory");
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (tid == 0) {
    const uint32_t idesc = (2u << 4) | (1u << 15) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(64 >> 4) << 24);
    for (int ks = 0; ks < K / 32; ++ks) {
      const uint64_t da = make_desc(smem_u32(sA) + ks * 2048, 16, 512, 4);
      const uint64_t db = make_desc(smem_u32(sB) + 2 * ks * LBO_B, LBO_B, SBO_B, 0);
      asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem), "l"(da), "l"(db), "r"(idesc), "r"(ks > 0 ? 1u : 0u) : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&mbar)) : "memory");
  }
  { uint32_t ok = 0; while (!ok) asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(ok) : "r"(smem_u32(&mbar)), "r"(0u) : "memory"); }
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  for (int c0 = 0; c0 < N; c0 += 16) {
    uint32_t v[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]),
                   "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(tmem + ((uint32_t)(32 * warp) << 16) + (uint32_t)c0));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int c = 0; c < 16; ++c) D[tid * N + c0 + c] = (int32_t)v[c];
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(tmem) : "memory");
}

template <int N>
static int run64() {
  std::vector<uint8_t> hA(K * 64), hB(N * K);
  srand(21);
  for (auto& v : hA) v = (uint8_t)(rand() & 255);
  for (int n = 0; n < N; ++n) for (int k = 0; k < K; ++k) hB[n * K + k] = (uint8_t)((rand() % N) == n ? 1 : 0);
  uint8_t *dA, *dB; int32_t* dD;
  CK(cudaMalloc(&dA, hA.size())); CK(cudaMalloc(&dB, hB.size())); CK(cudaMalloc(&dD, 128 * N * 4));
  CK(cudaMemcpy(dA, hA.data(), hA.size(), cudaMemcpyHostToDevice)); CK(cudaMemcpy(dB, hB.data(), hB.size(), cudaMemcpyHostToDevice));
  const int smem = K * 64 + 8 * ((N / 8) * 128 + 16) + 1024;
  CK(cudaFuncSetAttribute(k_test64<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  k_test64<N><<<1, 128, smem>>>(dA, dB, dD);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("M=64 N=%d: launch error %s\n", N, cudaGetErrorString(e)); return 1; }
  std::vector<int32_t> hD(128 * N);
  CK(cudaMemcpy(hD.data(), dD, 128 * N * 4, cudaMemcpyDeviceToHost));
  std::vector<long> ref(64 * N);
  for (int m = 0; m < 64; ++m) for (int n = 0; n < N; ++n) { long r = 0; for (int k = 0; k < K; ++k) r += (long)hA[k * 64 + m] * hB[n * K + k]; ref[m * N + n] = r; }
  // which lane holds result row m?
  int found = 0; printf("M=64 N=%d: row -> lane:", N);
  for (int m = 0; m < 64; ++m) {
    int where = -1;
    for (int l = 0; l < 128 && where < 0; ++l) { bool eq = true; for (int n = 0; n < N; ++n) if (hD[l * N + n] != ref[m * N + n]) { eq = false; break; } if (eq) where = l; }
    if (where >= 0) ++found;
    if (m < 20 || m % 16 == 0) printf(" %d->%d", m, where);
  }
  printf("  (%d of 64 rows found)\n", found);
  cudaFree(dA); cudaFree(dB); cudaFree(dD);
  return found != 64;
}

int main() {
  int rc = 0;
  rc |= run<16>(0);        // swizzled A, LBO field 16
  rc |= run<16>(1);        // swizzled A, LBO field 128 (must not matter: one atom along M)
  rc |= run<64>(0);
  rc |= run<32>(0);
  rc |= run64<16>();
  rc |= run64<32>();
  rc |= run4<16>(2000);
  rc |= run4<64>(500);
  return rc;
}

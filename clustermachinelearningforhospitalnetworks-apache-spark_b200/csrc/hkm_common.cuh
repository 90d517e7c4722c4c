// Shared device helpers for libhkm (sm_100a only): mbarrier / bulk-async (TMA 1-D) copies,
// fp64 tensor MMA, error plumbing for the C ABI.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdarg>

namespace hkm {

// ---- thread-local error text behind hkm_last_error()
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);
#define HKM_CUDA_TRY(expr)                                                   \
  do {                                                                       \
    cudaError_t e__ = (expr);                                                \
    if (e__ != cudaSuccess) return ::hkm::cuda_fail(e__, #expr);             \
  } while (0)

int sm_count();            // SMs of the current device (cached per device)

// ---- PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
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
      "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n" ::"r"(smem_u32(b)), "r"(parity)
      : "memory");
}
// 1-D bulk async copy global -> shared (TMA engine; SASS UBLKCP), completion on an mbarrier.
// dst/src 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// ---- explicit shared-space accessors on 32-bit shared addresses.  They keep every hot-loop address in one
// register and spare the generic<->shared conversions the compiler otherwise re-derives per tile.
__device__ __forceinline__ double lds_f64(uint32_t a) { double v; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds_u32(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint4 lds_v4(uint32_t a) {
  uint4 v; asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a)); return v;
}
__device__ __forceinline__ void sts_f64(uint32_t a, double v) { asm volatile("st.shared.f64 [%0], %1;" ::"r"(a), "d"(v) : "memory"); }
__device__ __forceinline__ void sts_u32(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
// store only where `on` is set, without a divergent branch
__device__ __forceinline__ void sts_u32_if(uint32_t a, uint32_t v, bool on) {
  asm volatile("{\n.reg .pred p;\nsetp.ne.u32 p, %2, 0;\n@p st.shared.u32 [%0], %1;\n}\n" ::"r"(a), "r"(v), "r"((uint32_t)on) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx_a(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait_a(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n.reg .pred p;\nWAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n" ::"r"(bar), "r"(parity)
      : "memory");
}
__device__ __forceinline__ bool mbar_test_a(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}
// non-blocking probe (test_wait never suspends the thread)
__device__ __forceinline__ bool mbar_probe_a(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n.reg .pred p;\nmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void bulk_g2s_a(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
               "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
// fp64 tensor MMA, D(8x8) += A(8x4, row) * B(4x8, col).  Lane l: a = A[l>>2][l&3],
// b = B[l&3][l>>2], c0/c1 = C[l>>2][2*(l&3) + {0,1}].   SASS: DMMA.8x8x4
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
// 16x8x4 variant: rows 0-7 (c0,c1 / a_lo) and rows 8-15 (c2,c3 / a_hi) share B.  ptxas lowers it to two
// adjacent DMMA.8x8x4 on independent accumulators, which keeps the two row tiles interleaved in the
// issue stream (separate m8n8k4 calls get serialised into one long dependent chain by the scheduler).
__device__ __forceinline__ void dmma1684(double& c0, double& c1, double& c2, double& c3, double a_lo, double a_hi, double b) {
  asm("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
      : "+d"(c0), "+d"(c1), "+d"(c2), "+d"(c3) : "d"(a_lo), "d"(a_hi), "d"(b));
}
// Sum of n doubles base[0], base[stride], base[2 stride], ... in index order (the last-CTA fold of per-CTA partials).  The loads
// of a batch are issued together and then added in order: same result as the plain loop, without a memory round trip per term.
__device__ __forceinline__ double ordered_sum_ldcg(const double* base, size_t stride, unsigned n) {
  double s = 0;
  unsigned b = 0;
#pragma unroll 1
  for (; b + 16 <= n; b += 16) {
    double v[16];
#pragma unroll
    for (int u = 0; u < 16; ++u) v[u] = __ldcg(base + (size_t)(b + u) * stride);
#pragma unroll
    for (int u = 0; u < 16; ++u) s += v[u];
  }
  if (b < n) {
    double v[16];
#pragma unroll
    for (int u = 0; u < 16; ++u) v[u] = (b + u < n) ? __ldcg(base + (size_t)(b + u) * stride) : 0.0;
#pragma unroll
    for (int u = 0; u < 16; ++u) if (b + u < n) s += v[u];
  }
  return s;
}
__device__ __forceinline__ double shfl_xor_d(double v, int m) { return __shfl_xor_sync(0xffffffffu, v, m); }

}  // namespace hkm

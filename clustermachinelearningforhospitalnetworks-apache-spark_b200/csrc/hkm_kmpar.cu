// k-means|| seeding kernels (mllib KMeans.initKMeansParallel) and candidate
// weighting.  The distance work reuses the Lloyd pass in its min-merge output mode; what is
// new here is the partition-exact Bernoulli selection: Spark draws ONE nextDouble() per row
// from a per-partition XORShiftRandom, sequentially.  XORShift is linear over GF(2), so row r's
// state is T^(2r) s0; a precomputed table of T^(2^j) lets every thread jump straight to its
// strip of rows and then step sequentially inside it.
#include "../../include/hkm.h"
#include "hkm_lloyd_launch.cuh"
#include <cmath>
#include <mutex>

namespace hkm {

constexpr int kJumpBits = 42;
constexpr int kStripRows = 256;
__device__ uint64_t g_jump[kJumpBits][64];  // column form: g_jump[j][i] = T^(2^j) e_i

__device__ __forceinline__ uint64_t xs_step(uint64_t x) {
  x ^= x << 21; x ^= x >> 35; x ^= x << 4; return x;
}
__device__ __forceinline__ uint64_t apply_cols(const uint64_t* cols, uint64_t x) {
  uint64_t y = 0;
#pragma unroll 8
  for (int i = 0; i < 64; ++i) y ^= ((x >> i) & 1ull) ? cols[i] : 0ull;
  return y;
}
// one block of 64 threads: repeated squaring of the step matrix
__global__ void jump_table_init_kernel() {
  __shared__ uint64_t cur[64], nxt[64];
  const int i = threadIdx.x;
  cur[i] = xs_step(1ull << i);
  __syncthreads();
  for (int j = 0; j < kJumpBits; ++j) {
    g_jump[j][i] = cur[i];
    nxt[i] = apply_cols(cur, cur[i]);
    __syncthreads();
    cur[i] = nxt[i];
    __syncthreads();
  }
}

// Built once per device.  The first caller waits for the build, so a table marked ready is ready for every stream.
static int ensure_jump_table(cudaStream_t st) {
  static bool ready[64];
  static std::mutex mu;
  int dev = 0; HKM_CUDA_TRY(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lock(mu);
  if (dev < 64 && ready[dev]) return HKM_OK;
  jump_table_init_kernel<<<1, 64, 0, st>>>();
  HKM_CUDA_TRY(cudaGetLastError());
  HKM_CUDA_TRY(cudaStreamSynchronize(st));
  if (dev < 64) ready[dev] = true;
  return HKM_OK;
}

// MurmurHash3.bytesHash over the 8 big-endian bytes of a long (XORShiftRandom.hashSeed)
__device__ __forceinline__ uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }
__device__ uint32_t mm3_hash8(uint64_t v, uint32_t seed) {
  uint32_t h = seed;
#pragma unroll
  for (int w = 0; w < 2; ++w) {
    // bytes b[4w..4w+3] of the big-endian encoding, packed little-endian into k
    uint32_t be = (uint32_t)(v >> (32 * (1 - w)));
    uint32_t k = __byte_perm(be, 0, 0x0123);
    k *= 0xcc9e2d51u; k = rotl32(k, 15); k *= 0x1b873593u;
    h ^= k; h = rotl32(h, 13); h = h * 5u + 0xe6546b64u;
  }
  h ^= 8u;
  h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
  return h;
}
__device__ uint64_t hash_seed_dev(int64_t seed) {
  const uint32_t lo = mm3_hash8((uint64_t)seed, 0x3c074a61u);
  const uint32_t hi = mm3_hash8((uint64_t)seed, lo);
  return ((uint64_t)hi << 32) | lo;
}

// part_rows[p] .. part_rows[p+1]: local row range of partition p; strip_start[p]: first strip id
__global__ void __launch_bounds__(128) kmpar_sample_kernel(const double* __restrict__ costs, const int64_t* __restrict__ part_rows,
                                                          const int64_t* __restrict__ strip_start, int P, int part_index0, int32_t seed,
                                                          int step, int k, const double* __restrict__ sum_costs,
                                                          int64_t* __restrict__ idx_out, int cap, int32_t* __restrict__ count_out) {
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= strip_start[P]) return;
  int lo = 0, hi = P;  // partition p with strip_start[p] <= g < strip_start[p+1]
  while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (strip_start[mid] <= g) lo = mid; else hi = mid; }
  const int p = lo;
  const int64_t r0 = part_rows[p] + (g - strip_start[p]) * kStripRows;
  const int64_t r1 = min(r0 + (int64_t)kStripRows, part_rows[p + 1]);
  // new XORShiftRandom(seed ^ (step << 16) ^ index): Int arithmetic, widened to Long
  const int32_t s32 = seed ^ (int32_t)((uint32_t)step << 16) ^ (int32_t)(part_index0 + p);
  uint64_t x = hash_seed_dev((int64_t)s32);
  uint64_t steps = 2ull * (uint64_t)(r0 - part_rows[p]);  // two next() calls per nextDouble()
  for (int j = 0; steps != 0 && j < kJumpBits; ++j, steps >>= 1)
    if (steps & 1ull) x = apply_cols(g_jump[j], x);
  const double sum = *sum_costs;
  const double kk = (double)k;
  for (int64_t r = r0; r < r1; ++r) {
    x = xs_step(x); const uint64_t a = x & ((1ull << 26) - 1);
    x = xs_step(x); const uint64_t b = x & ((1ull << 27) - 1);
    const double u = (double)(int64_t)((a << 27) + b) * 0x1.0p-53;
    const double c = costs[r];
    if (u < 2.0 * c * kk / sum) {  // rand.nextDouble() < 2.0 * c * k / sumCosts
      const int pos = atomicAdd(count_out, 1);
      if (pos < cap) idx_out[pos] = r;
    }
  }
}

__global__ void __launch_bounds__(256) label_counts_kernel(const int32_t* __restrict__ labels, int64_t n, int m, unsigned long long* __restrict__ counts) {
  extern __shared__ unsigned int hist[];
  for (int i = threadIdx.x; i < m; i += blockDim.x) hist[i] = 0;
  __syncthreads();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int l = labels[i];
    if (l >= 0 && l < m) atomicAdd(&hist[l], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < m; i += blockDim.x) if (hist[i]) atomicAdd(&counts[i], (unsigned long long)hist[i]);
}

__global__ void fill_double_kernel(double* p, int64_t n, double v) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}

}  // namespace hkm

using namespace hkm;

extern "C" {

// A k-means|| step scores the rows against chunks of 1..64 new centres; the pass's grid (CTAs per SM) depends on the
// chunk's k, so the scratch must cover the largest requirement over all chunk sizes, not only k = 64.
size_t hkm_kmpar_scratch_bytes(int d) {
  size_t mx = 0;
  for (int k = 1; k <= 64; ++k) { const size_t b = hkm_assign_reduce_scratch_bytes(d, k); if (b > mx) mx = b; }
  return mx;
}

int hkm_kmpar_cost_update(const double* X, int64_t n, int d, const double* newC, int m, double* costs_inout, double* sum_out,
                          void* scratch, size_t scratch_bytes, void* stream) {
  if (m < 1 || costs_inout == nullptr || sum_out == nullptr) { set_error("hkm_kmpar_cost_update: bad arguments"); return HKM_ERR_INVALID; }
  const int chunk = 64;
  for (int c0 = 0; c0 < m; c0 += chunk) {
    AssignArgs a{};
    a.X = X; a.n = n; a.d = d; a.C = newC + (size_t)c0 * d; a.k = (m - c0 < chunk) ? (m - c0) : chunk;
    a.partial_out = sum_out; a.mind2_out = costs_inout; a.do_reduce = 2; a.out_mode = 1;
    a.scratch = scratch; a.scratch_bytes = scratch_bytes; a.stream = stream;
    const int rc = assign_pass(a, "hkm_kmpar_cost_update");
    if (rc != HKM_OK) return rc;
  }
  return HKM_OK;
}

int hkm_kmpar_sample(const double* costs, int64_t n, const int64_t* part_rows, const int64_t* strip_start, int64_t total_strips, int P,
                     int part_index0, int32_t seed, int step, int k, const double* sum_costs, int64_t* idx_out, int cap,
                     int32_t* count_out, void* stream) {
  if (costs == nullptr || part_rows == nullptr || strip_start == nullptr || P < 1 || sum_costs == nullptr || idx_out == nullptr || count_out == nullptr || cap < 1) {
    set_error("hkm_kmpar_sample: bad arguments"); return HKM_ERR_INVALID;
  }
  (void)n;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  int rc = ensure_jump_table(st);
  if (rc != HKM_OK) return rc;
  HKM_CUDA_TRY(cudaMemsetAsync(count_out, 0, sizeof(int32_t), st));
  if (total_strips > 0) {
    const int64_t blocks = (total_strips + 127) / 128;
    kmpar_sample_kernel<<<(unsigned)blocks, 128, 0, st>>>(costs, part_rows, strip_start, P, part_index0, seed, step, k, sum_costs, idx_out, cap, count_out);
    HKM_CUDA_TRY(cudaGetLastError());
  }
  return HKM_OK;
}

int hkm_kmpar_strip_rows(void) { return kStripRows; }

int hkm_fill_f64(double* p, int64_t n, double value, void* stream) {
  if (n < 0 || (n > 0 && p == nullptr)) { set_error("hkm_fill_f64: bad arguments"); return HKM_ERR_INVALID; }
  if (n > 0) { fill_double_kernel<<<sm_count() * 4, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(p, n, value); HKM_CUDA_TRY(cudaGetLastError()); }
  return HKM_OK;
}

int hkm_closest_counts(const double* X, int64_t n, int d, const double* cand, int m, int32_t* labels_work, double* mind2_work,
                       unsigned long long* counts_out, void* stream) {
  if (m < 1 || labels_work == nullptr || mind2_work == nullptr || counts_out == nullptr) { set_error("hkm_closest_counts: bad arguments"); return HKM_ERR_INVALID; }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (n > 0) {
    fill_double_kernel<<<sm_count() * 4, 256, 0, st>>>(mind2_work, n, HUGE_VAL);
    HKM_CUDA_TRY(cudaGetLastError());
  }
  const int chunk = 64;
  for (int c0 = 0; c0 < m; c0 += chunk) {
    AssignArgs a{};
    a.X = X; a.n = n; a.d = d; a.C = cand + (size_t)c0 * d; a.k = (m - c0 < chunk) ? (m - c0) : chunk;
    a.labels_out = labels_work; a.mind2_out = mind2_work; a.do_reduce = 0; a.out_mode = 1; a.label_base = c0; a.stream = stream;
    const int rc = assign_pass(a, "hkm_closest_counts");
    if (rc != HKM_OK) return rc;
  }
  HKM_CUDA_TRY(cudaMemsetAsync(counts_out, 0, sizeof(unsigned long long) * (size_t)m, st));
  if (n > 0) {
    if ((size_t)m * 4 > 200 * 1024) { set_error("hkm_closest_counts: too many candidates (%d)", m); return HKM_ERR_UNSUPPORTED; }
    if ((size_t)m * 4 > 48 * 1024) HKM_CUDA_TRY(cudaFuncSetAttribute(label_counts_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, m * 4));
    label_counts_kernel<<<sm_count() * 2, 256, (size_t)m * 4, st>>>(labels_work, n, m, counts_out);
    HKM_CUDA_TRY(cudaGetLastError());
  }
  return HKM_OK;
}

int hkm_label_counts(const int32_t* labels, int64_t n, int m, unsigned long long* counts_out, void* stream) {
  if (m < 1 || n < 0 || counts_out == nullptr || (n > 0 && labels == nullptr)) { set_error("hkm_label_counts: bad arguments"); return HKM_ERR_INVALID; }
  if ((size_t)m * 4 > 200 * 1024) { set_error("hkm_label_counts: too many labels (%d)", m); return HKM_ERR_UNSUPPORTED; }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  HKM_CUDA_TRY(cudaMemsetAsync(counts_out, 0, sizeof(unsigned long long) * (size_t)m, st));
  if (n > 0) {
    if ((size_t)m * 4 > 48 * 1024) HKM_CUDA_TRY(cudaFuncSetAttribute(label_counts_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, m * 4));
    label_counts_kernel<<<sm_count() * 2, 256, (size_t)m * 4, st>>>(labels, n, m, counts_out);
    HKM_CUDA_TRY(cudaGetLastError());
  }
  return HKM_OK;
}

}  // extern "C"

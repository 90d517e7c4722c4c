// General-shape kernels for the Lloyd data pass: any d up to 4096 and any k, for the (d, k) pairs the tiled fast kernels
// refuse (d > 96, or a tile ring + centre tile that does not fit in shared memory).  Plain global loads, no TMA, no tensor
// cores: a warp scores one row against one chunk of centres (lane = feature, shuffle reduction per centre), and the
// per-cluster sums are built one row at a time per CTA with thread = feature, so every (cluster, feature) accumulator has
// exactly one writer and the summation order is fixed.  Same arithmetic as the fast kernels (score x.c - 0.5|c|^2, strict '>'
// in ascending centre index, distance max(|x|^2 - 2 score, 0)); several times below the HBM rate -- a correct general path, not a
// tuned one.  hkm_api.cu runs these under the same chunk loop it uses for k > 64.
#include "../../include/hkm.h"
#include "hkm_lloyd_launch.cuh"

namespace hkm {

constexpr int kGenMaxD = 4096;
constexpr int kGenRegs = kGenMaxD / 32;   // features per lane

// labels/mind2: out_mode 0 writes, out_mode 1 keeps the running minimum (labels + label_base follow the winner); cost_out
// (nullable) receives the sum of the per-row values after the update: per-CTA sums in cost_parts, folded in CTA order by the last
// CTA (ticket counter, zero on entry, left zero).  Grid-stride over rows, warp per row; NR = features per lane held in registers.
template <int NR>
__global__ void __launch_bounds__(256) generic_assign_kernel(const double* __restrict__ X, int64_t n, int d, const double* __restrict__ C, int k,
                                                             int32_t* __restrict__ labels, double* __restrict__ mind2, int out_mode, int label_base,
                                                             const int32_t* __restrict__ skip, double* __restrict__ cost_out,
                                                             double* __restrict__ cost_parts, unsigned int* counter) {
  if (skip != nullptr && *skip != 0) return;
  extern __shared__ double cn_s[];                 // [k] -0.5 |c_j|^2
  __shared__ double s_cost[8];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int j = warp; j < k; j += 8) {
    double s = 0;
    for (int f = lane; f < d; f += 32) { const double c = C[(size_t)j * d + f]; s = fma(c, c, s); }
#pragma unroll
    for (int m = 16; m >= 1; m >>= 1) s += shfl_xor_d(s, m);
    if (lane == 0) cn_s[j] = -0.5 * s;
  }
  __syncthreads();
  double cost = 0;
  const int nreg = (d + 31) / 32;
  for (int64_t r = (int64_t)blockIdx.x * 8 + warp; r < n; r += (int64_t)gridDim.x * 8) {
    const double* x = X + (size_t)r * d;
    double xr[NR];
    double xx = 0;
#pragma unroll
    for (int i = 0; i < NR; ++i) { const int f = lane + 32 * i; const double v = (i < nreg && f < d) ? x[f] : 0.0; xr[i] = v; xx = fma(v, v, xx); }
#pragma unroll
    for (int m = 16; m >= 1; m >>= 1) xx += shfl_xor_d(xx, m);
    double best = -__longlong_as_double(0x7ff0000000000000LL); int bj = 0;
    for (int j = 0; j < k; ++j) {
      const double* c = C + (size_t)j * d;
      double s = 0;
#pragma unroll
      for (int i = 0; i < NR; ++i) { const int f = lane + 32 * i; if (i < nreg && f < d) s = fma(xr[i], __ldg(c + f), s); }
#pragma unroll
      for (int m = 16; m >= 1; m >>= 1) s += shfl_xor_d(s, m);
      s += cn_s[j];
      if (s > best) { best = s; bj = j; }
    }
    if (lane == 0) {
      const double dist = fmax(fma(-2.0, best, xx), 0.0);
      if (out_mode == 0) {
        if (labels != nullptr) labels[r] = bj;
        if (mind2 != nullptr) mind2[r] = dist;
        cost += dist;
      } else {
        const double old = mind2[r];
        if (dist < old) { mind2[r] = dist; if (labels != nullptr) labels[r] = bj + label_base; cost += dist; } else cost += old;
      }
    }
  }
  if (cost_out != nullptr) {
    __shared__ unsigned int s_ticket;
    if (lane == 0) s_cost[warp] = cost;
    __syncthreads();
    if (tid == 0) {
      double t = 0; for (int w = 0; w < 8; ++w) t += s_cost[w];
      cost_parts[blockIdx.x] = t;
      __threadfence();
      s_ticket = atomicAdd(counter, 1u);
      if (s_ticket == gridDim.x - 1) {
        __threadfence();
        double tot = 0;
        tot = ordered_sum_ldcg(cost_parts, 1, gridDim.x);
        *cost_out = tot; *counter = 0u;
      }
    }
  }
}

// Per-cluster sums with labels given: CTA b takes the rows [b * rows_per_cta, ...) one at a time, thread = feature; shared
// accumulators [k][d] + [k] weights + cost, then the CTA's slot of `partials`; the last CTA folds the slots in CTA order.
// stats layout (as the fast kernels' statistics pass): [k*d sums][k weights][cost = sum w * mind2 over the chunk's rows].
__global__ void __launch_bounds__(256) generic_reduce_kernel(const double* __restrict__ X, int64_t n, int d, int k, const int32_t* __restrict__ labels,
                                                             int label_base, const double* __restrict__ w, const double* __restrict__ mind2,
                                                             const int32_t* __restrict__ skip, double* __restrict__ partials, unsigned int* counter,
                                                             double* __restrict__ stats) {
  if (skip != nullptr && *skip != 0) return;
  extern __shared__ double acc[];                  // [k*d] sums, [k] weights, [1] cost
  const int tid = threadIdx.x;
  const int S = k * d + k + 1;
  for (int i = tid; i < S; i += 256) acc[i] = 0.0;
  __syncthreads();
  const int64_t per = (n + gridDim.x - 1) / gridDim.x;
  const int64_t lo = min(n, (int64_t)blockIdx.x * per), hi = min(n, lo + per);
  for (int64_t r = lo; r < hi; ++r) {
    const int l = labels[r] - label_base;          // same value in every thread
    if (l < 0 || l >= k) continue;
    const double wr = w != nullptr ? w[r] : 1.0;
    const double* x = X + (size_t)r * d;
    double* a = acc + (size_t)l * d;
    for (int f = tid; f < d; f += 256) a[f] = fma(wr, x[f], a[f]);
    if (tid == 0) { acc[k * d + l] += wr; acc[k * d + k] = fma(wr, mind2 != nullptr ? mind2[r] : 0.0, acc[k * d + k]); }
  }
  __syncthreads();
  double* out = partials + (size_t)blockIdx.x * S;
  for (int i = tid; i < S; i += 256) out[i] = acc[i];
  __threadfence();
  __syncthreads();
  __shared__ unsigned int s_ticket;
  if (tid == 0) s_ticket = atomicAdd(counter, 1u);
  __syncthreads();
  if (s_ticket == gridDim.x - 1) {
    __threadfence();
    for (int i = tid; i < S; i += 256) {
      double s = 0;
      s = ordered_sum_ldcg(partials + i, (size_t)S, gridDim.x);
      stats[i] = s;
    }
    if (tid == 0) *counter = 0u;
  }
}

int generic_grid() { return sm_count() * 2; }
// centres per chunk so that the shared accumulators [kc][d] stay within ~190 KB
int generic_chunk_k(int d) { int kc = 24000 / d; return kc > 64 ? 64 : (kc < 1 ? 1 : kc); }
size_t generic_scratch_bytes(int d) {
  const int kc = generic_chunk_k(d);
  const size_t red = (size_t)(generic_grid() + 1) * ((size_t)kc * d + kc + 1) * sizeof(double), asg = (size_t)sm_count() * 8 * sizeof(double);
  return 256 + (red > asg ? red : asg);
}

// scratch (needed only with cost_out): 256-byte header with the ticket counter, then sm_count() * 8 doubles
int generic_assign(const double* X, int64_t n, int d, const double* C, int k, int32_t* labels, double* mind2, int out_mode, int label_base,
                   const int32_t* skip, double* cost_out, void* scratch, cudaStream_t st) {
  if (d > kGenMaxD) { set_error("d=%d above the general kernels' limit (%d)", d, kGenMaxD); return HKM_ERR_UNSUPPORTED; }
  if (out_mode != 0 && mind2 == nullptr) { set_error("generic_assign: min-merge needs mind2"); return HKM_ERR_INVALID; }
  if (cost_out != nullptr && scratch == nullptr) { set_error("generic_assign: the cost sum needs scratch"); return HKM_ERR_INVALID; }
  if (n == 0) { if (cost_out != nullptr) HKM_CUDA_TRY(cudaMemsetAsync(cost_out, 0, sizeof(double), st)); return HKM_OK; }
  double* parts = scratch ? reinterpret_cast<double*>(reinterpret_cast<unsigned char*>(scratch) + 256) : nullptr;
  unsigned int* counter = reinterpret_cast<unsigned int*>(scratch);
  const int grid = sm_count() * 8;
  const size_t smem = (size_t)k * sizeof(double);
  if (d <= 128) generic_assign_kernel<4><<<grid, 256, smem, st>>>(X, n, d, C, k, labels, mind2, out_mode, label_base, skip, cost_out, parts, counter);
  else if (d <= 512) generic_assign_kernel<16><<<grid, 256, smem, st>>>(X, n, d, C, k, labels, mind2, out_mode, label_base, skip, cost_out, parts, counter);
  else generic_assign_kernel<kGenRegs><<<grid, 256, smem, st>>>(X, n, d, C, k, labels, mind2, out_mode, label_base, skip, cost_out, parts, counter);
  HKM_CUDA_TRY(cudaGetLastError());
  return HKM_OK;
}

// stats: device [kc*d + kc + 1]; scratch: generic_scratch_bytes(d) (its first word is the ticket counter, zero on entry, left zero)
int generic_reduce(const double* X, int64_t n, int d, int kc, const int32_t* labels, int label_base, const double* w, const double* mind2,
                   const int32_t* skip, void* scratch, double* stats, cudaStream_t st) {
  const size_t smem = ((size_t)kc * d + kc + 1) * sizeof(double);
  if (smem > 200 * 1024) { set_error("generic_reduce: chunk of %d centres x %d features does not fit", kc, d); return HKM_ERR_UNSUPPORTED; }
  static bool attr_done[64];
  int dev = 0; HKM_CUDA_TRY(cudaGetDevice(&dev));
  if (dev >= 64 || !attr_done[dev]) {
    HKM_CUDA_TRY(cudaFuncSetAttribute(generic_reduce_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    if (dev < 64) attr_done[dev] = true;
  }
  generic_reduce_kernel<<<generic_grid(), 256, smem, st>>>(X, n, d, kc, labels, label_base, w, mind2, skip,
                                                           reinterpret_cast<double*>(reinterpret_cast<unsigned char*>(scratch) + 256),
                                                           reinterpret_cast<unsigned int*>(scratch), stats);
  HKM_CUDA_TRY(cudaGetLastError());
  return HKM_OK;
}

}  // namespace hkm

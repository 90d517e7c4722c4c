// ClusteringEvaluator silhouette, squared-Euclidean closed form (ml/evaluation/ClusteringMetrics.scala
// SquaredEuclideanSilhouette): two streaming passes over X, O(n*k*d), no pairwise work.
//   pass 1  per cluster G: Y_G = sum w x, N_G = sum w, Psi_G = sum w |x|^2      (Lloyd kernel, MODE 1)
//   pass 2  per row: D(x,G) = |x|^2 + Psi_G/N_G - 2 x.Y_G/N_G for every G, silhouette s(x), sum s*w  (MODE 2)
#include "../../include/hkm.h"
#include "hkm_lloyd_launch.cuh"

namespace hkm {

// work[0 .. k*d) = Y_G / N_G (the "centre" of the score pass), then [k] N_G, then [k] Psi_G / N_G
__global__ void silhouette_prep_kernel(const double* __restrict__ stats, int k, int d, double* __restrict__ work) {
  const int P = k * d + k + 1;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < k * d + 2 * k; i += gridDim.x * blockDim.x) {
    if (i < k * d) {
      const double N = stats[k * d + i / d];
      work[i] = N > 0 ? stats[i] / N : 0.0;
    } else if (i < k * d + k) {
      work[i] = stats[i];
    } else {
      const int j = i - k * d - k;
      const double N = stats[k * d + j];
      work[i] = N > 0 ? stats[P + j] / N : 0.0;
    }
  }
}

// smallest and largest cluster id of the prediction column in one pass (the evaluator needs the id range before it can size the
// per-cluster statistics); out2 starts at {INT_MAX, INT_MIN}
__global__ void __launch_bounds__(256) label_range_kernel(const int32_t* __restrict__ labels, int64_t n, int32_t* __restrict__ out2) {
  int lo = 0x7fffffff, hi = (int)0x80000000;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) { const int l = labels[i]; lo = min(lo, l); hi = max(hi, l); }
  lo = __reduce_min_sync(0xffffffffu, lo); hi = __reduce_max_sync(0xffffffffu, hi);
  if ((threadIdx.x & 31) == 0) { atomicMin(out2, lo); atomicMax(out2 + 1, hi); }
}
__global__ void label_range_init_kernel(int32_t* out2) { out2[0] = 0x7fffffff; out2[1] = (int)0x80000000; }

}  // namespace hkm

using namespace hkm;

extern "C" {

int hkm_label_range(const int32_t* labels, int64_t n, int32_t* out2, void* stream) {
  if (out2 == nullptr || (n > 0 && labels == nullptr)) { set_error("hkm_label_range: bad arguments"); return HKM_ERR_INVALID; }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  label_range_init_kernel<<<1, 1, 0, st>>>(out2);
  if (n > 0) label_range_kernel<<<sm_count() * 4, 256, 0, st>>>(labels, n, out2);
  HKM_CUDA_TRY(cudaGetLastError());
  return HKM_OK;
}

int hkm_silhouette_stats(const double* X, const int32_t* labels, const double* weights, int64_t n, int d, int k, double* stats_out,
                         void* scratch, size_t scratch_bytes, void* stream) {
  if (labels == nullptr || stats_out == nullptr) { set_error("hkm_silhouette_stats: bad arguments"); return HKM_ERR_INVALID; }
  AssignArgs a{};
  a.X = X; a.n = n; a.d = d; a.C = X;  /* centres unused in this mode; any valid pointer */
  a.k = k; a.weights = weights; a.partial_out = stats_out; a.do_reduce = 1; a.mode = 1; a.labels_in = labels;
  a.scratch = scratch; a.scratch_bytes = scratch_bytes; a.stream = stream;
  if (n == 0) {  // nothing to read: all statistics are zero
    HKM_CUDA_TRY(cudaMemsetAsync(stats_out, 0, sizeof(double) * ((size_t)k * d + 2 * (size_t)k + 1), reinterpret_cast<cudaStream_t>(stream)));
    return HKM_OK;
  }
  return assign_pass(a, "hkm_silhouette_stats");
}

int hkm_silhouette_score(const double* X, const int32_t* labels, const double* weights, int64_t n, int d, int k, const double* stats,
                         double* work, double* out2, void* scratch, size_t scratch_bytes, void* stream) {
  if (labels == nullptr || stats == nullptr || work == nullptr || out2 == nullptr) { set_error("hkm_silhouette_score: bad arguments"); return HKM_ERR_INVALID; }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (n == 0) { HKM_CUDA_TRY(cudaMemsetAsync(out2, 0, 2 * sizeof(double), st)); return HKM_OK; }
  silhouette_prep_kernel<<<(k * d + 2 * k + 255) / 256, 256, 0, st>>>(stats, k, d, work);
  HKM_CUDA_TRY(cudaGetLastError());
  AssignArgs a{};
  a.X = X; a.n = n; a.d = d; a.C = work; a.k = k; a.weights = weights; a.partial_out = out2; a.do_reduce = 1; a.mode = 2;
  a.labels_in = labels; a.aux = work + (size_t)k * d;
  a.scratch = scratch; a.scratch_bytes = scratch_bytes; a.stream = stream;
  return assign_pass(a, "hkm_silhouette_score");
}

}  // extern "C"

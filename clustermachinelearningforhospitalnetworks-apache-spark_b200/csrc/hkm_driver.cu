// Driver-side (host) pieces of the KMeans path that Spark also runs on the driver JVM, O(candidates) work:
// LocalKMeans.kMeansPlusPlus (mllib/clustering/LocalKMeans.scala) on the k-means|| candidates, with java.util.Random.
// No device code here; the pointers are HOST pointers.
#include <cstdint>
#include <cstring>
#include <vector>

#include "../../include/hkm.h"
#include "hkm_common.cuh"

namespace {

// java.util.Random: 48-bit LCG
struct JavaRandom {
  uint64_t s;
  explicit JavaRandom(int64_t seed) : s(((uint64_t)seed ^ 0x5DEECE66DULL) & ((1ULL << 48) - 1)) {}
  int32_t next(int bits) {
    s = (s * 0x5DEECE66DULL + 0xBULL) & ((1ULL << 48) - 1);
    return (int32_t)((int64_t)s >> (48 - bits));
  }
  double next_double() { return (double)(((int64_t)next(26) << 27) + next(27)) * (1.0 / (double)(1LL << 53)); }
  int32_t next_int(int32_t bound) {
    int32_t r = next(31);
    const int32_t m = bound - 1;
    if ((bound & m) == 0) return (int32_t)(((int64_t)bound * (int64_t)r) >> 31);
    for (int32_t u = r; (int32_t)((uint32_t)u - (uint32_t)(r = u % bound) + (uint32_t)m) < 0; u = next(31)) {}
    return r;
  }
};

// Vectors.sqdist, dense x dense: sum of squared differences in index order
inline double sqdist(const double* a, const double* b, int d) {
  double s = 0;
  for (int i = 0; i < d; ++i) { const double t = a[i] - b[i]; s += t * t; }
  return s;
}

}  // namespace

extern "C" int hkm_local_kmeanspp(int32_t seed, const double* points, const double* weights, int m, int d, int k, int max_iter,
                                  double* centers_out) {
  if (points == nullptr || weights == nullptr || centers_out == nullptr || m < 1 || d < 1 || k < 1 || max_iter < 0) {
    hkm::set_error("hkm_local_kmeanspp: bad arguments");
    return HKM_ERR_INVALID;
  }
  JavaRandom rand(seed);
  auto pt = [&](int i) { return points + (size_t)i * d; };
  auto ctr = [&](int j) { return centers_out + (size_t)j * d; };
  // centers(0) = pickWeighted(rand, points, weights)
  {
    double wsum = 0;
    for (int i = 0; i < m; ++i) wsum += weights[i];
    const double r = rand.next_double() * wsum;
    int i = 0; double cur = 0;
    while (i < m && cur < r) { cur += weights[i]; ++i; }
    std::memcpy(ctr(0), pt(i > 0 ? i - 1 : 0), sizeof(double) * d);
  }
  std::vector<double> cost(m);
  for (int i = 0; i < m; ++i) cost[i] = sqdist(pt(i), ctr(0), d);
  for (int c = 1; c < k; ++c) {
    double sum = 0;
    for (int i = 0; i < m; ++i) sum += cost[i] * weights[i];
    const double r = rand.next_double() * sum;
    double cum = 0; int j = 0;
    while (j < m && cum < r) { cum += weights[j] * cost[j]; ++j; }
    std::memcpy(ctr(c), pt(j == 0 ? 0 : j - 1), sizeof(double) * d);
    for (int i = 0; i < m; ++i) { const double v = sqdist(pt(i), ctr(c), d); if (v < cost[i]) cost[i] = v; }
  }
  // Lloyd on the candidates until no assignment moves
  std::vector<int> old(m, -1);
  std::vector<double> counts(k), sums((size_t)k * d);
  bool moved = true;
  for (int it = 0; moved && it < max_iter; ++it) {
    moved = false;
    std::fill(counts.begin(), counts.end(), 0.0);
    std::fill(sums.begin(), sums.end(), 0.0);
    for (int i = 0; i < m; ++i) {
      int best = 0; double bd = sqdist(ctr(0), pt(i), d);
      for (int j = 1; j < k; ++j) { const double v = sqdist(ctr(j), pt(i), d); if (v < bd) { bd = v; best = j; } }
      double* s = sums.data() + (size_t)best * d;
      for (int f = 0; f < d; ++f) s[f] += weights[i] * pt(i)[f];     // axpy(weights(i), p, sums(index))
      counts[best] += weights[i];
      if (best != old[i]) { moved = true; old[i] = best; }
    }
    for (int j = 0; j < k; ++j) {
      if (counts[j] == 0.0) std::memcpy(ctr(j), pt(rand.next_int(m)), sizeof(double) * d);
      else { const double inv = 1.0 / counts[j]; for (int f = 0; f < d; ++f) ctr(j)[f] = sums[(size_t)j * d + f] * inv; }
    }
  }
  return HKM_OK;
}

/*
 * hkm_oracle.c -- CPU restatement of the Spark MLlib algorithms on the KMeans hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product package may import, link or call this
 * file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs use it, and only as the checker / the timed CPU arm.
 *
 * PARITY UNPINNED BY THE REFERENCE: /root/reference/mllearnforhospitalnetwork.py holds no
 * KMeans code and no tests (it does not even parse, ref:L87-91).  The arithmetic of the path
 * lives in an un-vendored, un-pinned dependency -- Apache Spark MLlib (pyspark imported at
 * ref:L26-33; estimator idiom ref:L146-148; evaluator idiom ref:L162-163).  What follows
 * restates the published Spark 3.5.x algorithms ("row" solver, euclidean) file by file; each
 * function names the upstream file it follows.  It is pinned by the known answers that CAN be
 * reproduced offline (tests/test_oracle_kat.py): java.util.Random outputs, the pyspark
 * rand(seed=42) doc values (XORShiftRandom + hashSeed), the ClusteringEvaluatorSuite iris
 * silhouette constants, the KMeansSuite single-cluster centre, the pyspark KMeans doctest cost,
 * and scikit-learn cross-checks.
 *
 * Conventions: row-major fp64 X[n*d]; optional weights w (NULL = all 1.0); partitions are the
 * ParallelCollectionRDD slices [floor(i*n/P), floor((i+1)*n/P)); per-partition work is
 * sequential in row order and partitions are merged in index order.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define HKMO_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------------------------
 * scala.util.hashing.MurmurHash3.bytesHash  (scala-library, used by XORShiftRandom.hashSeed)
 * ---------------------------------------------------------------------------------------- */
static inline uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }
static uint32_t mm3_mix_last(uint32_t h, uint32_t k) {
  k *= 0xcc9e2d51u; k = rotl32(k, 15); k *= 0x1b873593u; return h ^ k;
}
static uint32_t mm3_mix(uint32_t h, uint32_t k) {
  h = mm3_mix_last(h, k); h = rotl32(h, 13); return h * 5u + 0xe6546b64u;
}
static uint32_t mm3_avalanche(uint32_t h) {
  h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16; return h;
}
static uint32_t mm3_bytes_hash(const uint8_t* data, int len0, uint32_t seed) {
  int len = len0, i = 0; uint32_t h = seed;
  while (len >= 4) {
    uint32_t k = data[i] | ((uint32_t)data[i + 1] << 8) | ((uint32_t)data[i + 2] << 16) | ((uint32_t)data[i + 3] << 24);
    h = mm3_mix(h, k); i += 4; len -= 4;
  }
  uint32_t k = 0;
  if (len == 3) k ^= (uint32_t)data[i + 2] << 16;
  if (len >= 2) k ^= (uint32_t)data[i + 1] << 8;
  if (len >= 1) { k ^= data[i]; h = mm3_mix_last(h, k); }
  return mm3_avalanche(h ^ (uint32_t)len0);
}

/* core/src/main/scala/org/apache/spark/util/random/XORShiftRandom.scala: hashSeed */
HKMO_API uint64_t hkmo_hash_seed(int64_t seed) {
  uint8_t b[8]; uint64_t u = (uint64_t)seed;
  for (int i = 0; i < 8; ++i) b[i] = (uint8_t)(u >> (56 - 8 * i)); /* ByteBuffer.putLong: big endian */
  uint32_t lo = mm3_bytes_hash(b, 8, 0x3c074a61u);                 /* MurmurHash3.arraySeed */
  uint32_t hi = mm3_bytes_hash(b, 8, lo);
  return ((uint64_t)hi << 32) | (uint64_t)lo;
}

/* XORShiftRandom extends java.util.Random, overriding next(bits) only. */
typedef struct { uint64_t s; } hkmo_xorshift;
HKMO_API void hkmo_xorshift_init(hkmo_xorshift* r, int64_t seed) { r->s = hkmo_hash_seed(seed); }
HKMO_API int32_t hkmo_xorshift_next(hkmo_xorshift* r, int bits) {
  uint64_t x = r->s;
  x ^= x << 21; x ^= x >> 35; x ^= x << 4;
  r->s = x;
  return (int32_t)(uint32_t)(x & ((1ULL << bits) - 1ULL));
}
HKMO_API int32_t hkmo_xorshift_next_int(hkmo_xorshift* r) { return hkmo_xorshift_next(r, 32); }
HKMO_API double hkmo_xorshift_next_double(hkmo_xorshift* r) {
  int64_t a = hkmo_xorshift_next(r, 26), b = hkmo_xorshift_next(r, 27);
  return (double)((a << 27) + b) * 0x1.0p-53;
}
HKMO_API int64_t hkmo_xorshift_next_long(hkmo_xorshift* r) {
  int64_t a = hkmo_xorshift_next(r, 32), b = hkmo_xorshift_next(r, 32);
  return (int64_t)((uint64_t)a << 32) + b;
}
/* java.util.Random.nextGaussian (polar method) driven by XORShift next(); state kept by caller */
HKMO_API double hkmo_xorshift_next_gaussian(hkmo_xorshift* r, int* have, double* saved) {
  if (*have) { *have = 0; return *saved; }
  double v1, v2, s;
  do { v1 = 2 * hkmo_xorshift_next_double(r) - 1; v2 = 2 * hkmo_xorshift_next_double(r) - 1; s = v1 * v1 + v2 * v2; } while (s >= 1 || s == 0);
  double m = sqrt(-2 * log(s) / s);
  *saved = v2 * m; *have = 1; return v1 * m;
}

/* java.util.Random (48-bit LCG) -- used by RDD.takeSample, PartitionwiseSampledRDD, LocalKMeans */
typedef struct { uint64_t s; } hkmo_jrandom;
#define JR_MASK ((1ULL << 48) - 1)
HKMO_API void hkmo_jrandom_init(hkmo_jrandom* r, int64_t seed) { r->s = ((uint64_t)seed ^ 0x5DEECE66DULL) & JR_MASK; }
HKMO_API int32_t hkmo_jrandom_next(hkmo_jrandom* r, int bits) {
  r->s = (r->s * 0x5DEECE66DULL + 0xBULL) & JR_MASK;
  return (int32_t)((int64_t)r->s >> (48 - bits));
}
HKMO_API int32_t hkmo_jrandom_next_int(hkmo_jrandom* r) { return hkmo_jrandom_next(r, 32); }
HKMO_API int64_t hkmo_jrandom_next_long(hkmo_jrandom* r) {
  int64_t a = hkmo_jrandom_next(r, 32), b = hkmo_jrandom_next(r, 32);
  return (int64_t)((uint64_t)a << 32) + b;
}
HKMO_API double hkmo_jrandom_next_double(hkmo_jrandom* r) {
  int64_t a = hkmo_jrandom_next(r, 26), b = hkmo_jrandom_next(r, 27);
  return (double)((a << 27) + b) * 0x1.0p-53;
}
HKMO_API int32_t hkmo_jrandom_next_int_bound(hkmo_jrandom* r, int32_t bound) {
  int32_t v = hkmo_jrandom_next(r, 31), m = bound - 1;
  if ((bound & m) == 0) return (int32_t)(((int64_t)bound * (int64_t)v) >> 31);
  for (int32_t u = v; (int32_t)((uint32_t)u - (uint32_t)(v = u % bound) + (uint32_t)m) < 0; u = hkmo_jrandom_next(r, 31)) {}
  return v;
}

/* ------------------------------------------------------------------------------------------
 * mllib/linalg/Vectors.scala: norm(v, 2.0), sqdist (dense x dense: sum (a_i-b_i)^2 in index order)
 * mllib/util/MLUtils.scala: fastSquaredDistance -- dense x dense goes straight to Vectors.sqdist
 * (Spark >= 3.1); form=1 selects the older norm-expansion branch for the switch SURVEY 8a-4 names.
 * ---------------------------------------------------------------------------------------- */
static double vnorm2(const double* v, int d) { double s = 0; for (int i = 0; i < d; ++i) s += v[i] * v[i]; return sqrt(s); }
static double sqdist(const double* a, const double* b, int d) {
  double s = 0; for (int i = 0; i < d; ++i) { double t = a[i] - b[i]; s += t * t; } return s;
}
static const double EPSILON_ML = 2.220446049250313e-16; /* MLUtils.EPSILON: smallest eps with 1+eps != 1 */
/* form 2 = CosineDistanceMeasure.distance(v1, v2) = 1 - dot(v1, v2) / v1.norm / v2.norm (DistanceMeasure.scala);
 * callers pass (centre, point) in that order, as DistanceMeasure.findClosest does. */
#define HKMO_FORM_COSINE 2
static double fast_sqdist(const double* a, double na, const double* b, double nb, int d, int form) {
  if (form == HKMO_FORM_COSINE) { double dot = 0; for (int i = 0; i < d; ++i) dot += a[i] * b[i]; return 1 - dot / na / nb; }
  if (form == 0) return sqdist(a, b, d);
  double ssn = na * na + nb * nb, nd = na - nb;
  double bound = 2.0 * EPSILON_ML * ssn / (nd * nd + EPSILON_ML);
  if (bound < 1e-6) { double dot = 0; for (int i = 0; i < d; ++i) dot += a[i] * b[i]; return ssn - 2.0 * dot; }
  return sqdist(a, b, d);
}

/* mllib/clustering/DistanceMeasure.scala: indexUpperTriangular / computeStatistics (euclidean) */
static inline int idx_ut(int n, int i, int j) { if (i > j) { int t = i; i = j; j = t; } return i * (2 * n - i - 1) / 2 + j; }
HKMO_API void hkmo_compute_statistics(const double* C, int k, int d, int form, double* packed /* k(k+1)/2 */) {
  if (k == 1) { packed[0] = NAN; return; }
  double* diag = (double*)malloc(sizeof(double) * k);
  double* cn = (double*)malloc(sizeof(double) * k);
  for (int i = 0; i < k; ++i) { diag[i] = INFINITY; cn[i] = vnorm2(C + (size_t)i * d, d); }
  for (int i = 0; i < k; ++i)
    for (int j = i + 1; j < k; ++j) {
      double dd = sqrt(fast_sqdist(C + (size_t)i * d, cn[i], C + (size_t)j * d, cn[j], d, form));
      double s = 0.25 * dd * dd;
      packed[idx_ut(k, i, j)] = s;
      if (s < diag[i]) diag[i] = s;
      if (s < diag[j]) diag[j] = s;
    }
  for (int i = 0; i < k; ++i) packed[idx_ut(k, i, i)] = diag[i];
  free(diag); free(cn);
}

/* EuclideanDistanceMeasure.findClosest(centers, point): strict '<', lowest index wins ties,
 * norm lower-bound skip. */
static int find_closest(const double* C, const double* cn, int k, int d, const double* x, double xn, int form, double* best_out) {
  double best = INFINITY; int bi = 0;
  for (int i = 0; i < k; ++i) {
    double lb = cn[i] - xn; lb *= lb;
    if (form == HKMO_FORM_COSINE) lb = -1;                /* base DistanceMeasure.findClosest: no norm bound */
    if (lb < best) {
      double dist = fast_sqdist(C + (size_t)i * d, cn[i], x, xn, d, form);
      if (dist < best) { best = dist; bi = i; }
    }
  }
  *best_out = best; return bi;
}
/* EuclideanDistanceMeasure.findClosest(centers, statistics, point): triangle-inequality pruning */
static int find_closest_stats(const double* C, const double* cn, const double* st, int k, int d, const double* x, double xn, int form, double* best_out) {
  double best = fast_sqdist(C, cn[0], x, xn, d, form);
  if (best < st[0]) { *best_out = best; return 0; }
  int bi = 0;
  for (int i = 1; i < k; ++i) {
    double nd = cn[i] - xn, lb = nd * nd;
    if (lb < best) {
      if (st[idx_ut(k, i, bi)] < best) {
        double dist = fast_sqdist(C + (size_t)i * d, cn[i], x, xn, d, form);
        if (dist < st[idx_ut(k, i, i)]) { *best_out = dist; return i; }
        if (dist < best) { best = dist; bi = i; }
      }
    }
  }
  *best_out = best; return bi;
}

HKMO_API int hkmo_find_closest(const double* C, int k, int d, const double* x, int use_stats, int form, double* dist_out) {
  double* cn = (double*)malloc(sizeof(double) * k);
  for (int i = 0; i < k; ++i) cn[i] = vnorm2(C + (size_t)i * d, d);
  int r;
  if (use_stats && k > 1) {
    double* st = (double*)malloc(sizeof(double) * (k * (k + 1) / 2));
    hkmo_compute_statistics(C, k, d, form, st);
    r = find_closest_stats(C, cn, st, k, d, x, vnorm2(x, d), form, dist_out);
    free(st);
  } else r = find_closest(C, cn, k, d, x, vnorm2(x, d), form, dist_out);
  free(cn); return r;
}

static inline int64_t part_lo(int64_t n, int P, int i) { return (int64_t)(((__int128)i * n) / P); }

/* ------------------------------------------------------------------------------------------
 * One Lloyd pass: mllib/clustering/KMeans.scala runAlgorithmWithWeight, the mapPartitions body
 * plus reduceByKey merge.  sums[k*d], wsum[k], cost.  labels optional.  Partitions in parallel
 * (OpenMP) when threads>1 -- each partition is sequential, merge is in index order, so the
 * result does not depend on the thread count.
 * ---------------------------------------------------------------------------------------- */
static void lloyd_pass_impl(const double* X, const double* w, int64_t n, int d, const double* C, const double* cn_in, int k,
                            int P, int use_stats, int form, int threads,
                            double* sums, double* wsum, double* cost_out, int32_t* labels) {
  double* cn = (double*)malloc(sizeof(double) * k);
  for (int i = 0; i < k; ++i) cn[i] = cn_in ? cn_in[i] : vnorm2(C + (size_t)i * d, d);
  double* st = NULL;
  if (use_stats && k > 1 && form != HKMO_FORM_COSINE) { st = (double*)malloc(sizeof(double) * (k * (k + 1) / 2)); hkmo_compute_statistics(C, k, d, form, st); }
  size_t per = (size_t)k * d + k + 1;
  double* part = (double*)calloc((size_t)P * per, sizeof(double));
#ifdef _OPENMP
  if (threads <= 0) threads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads)
#endif
  for (int p = 0; p < P; ++p) {
    double* ps = part + (size_t)p * per; double* pw = ps + (size_t)k * d; double* pc = pw + k;
    int64_t lo = part_lo(n, P, p), hi = part_lo(n, P, p + 1);
    for (int64_t r = lo; r < hi; ++r) {
      const double* x = X + (size_t)r * d;
      double xn = vnorm2(x, d), dist;
      int b = st ? find_closest_stats(C, cn, st, k, d, x, xn, form, &dist) : find_closest(C, cn, k, d, x, xn, form, &dist);
      double wt = w ? w[r] : 1.0;
      *pc += dist * wt;                                   /* costAccum.add(cost * point.weight) */
      double* s = ps + (size_t)b * d;
      if (form == HKMO_FORM_COSINE) {                     /* CosineDistanceMeasure.updateClusterSum: axpy(weight / norm, v, sum) */
        double a = wt / xn;                               /* xn == 0: Spark asserts; here the sums turn NaN/inf and the test sees it */
        for (int j = 0; j < d; ++j) s[j] += a * x[j];
      } else
      for (int j = 0; j < d; ++j) s[j] += wt * x[j];      /* axpy(point.weight, point.vector, sum) */
      pw[b] += wt;
      if (labels) labels[r] = b;
    }
  }
  memset(sums, 0, sizeof(double) * (size_t)k * d); memset(wsum, 0, sizeof(double) * k);
  double cost = 0;
  for (int p = 0; p < P; ++p) {
    double* ps = part + (size_t)p * per; double* pw = ps + (size_t)k * d;
    for (int j = 0; j < k; ++j) {
      if (pw[j] > 0) {                                    /* .filter(_._2._2 > 0) then reduceByKey */
        for (int f = 0; f < d; ++f) sums[(size_t)j * d + f] += ps[(size_t)j * d + f];
        wsum[j] += pw[j];
      }
    }
    cost += pw[k];
  }
  *cost_out = cost;
  free(part); free(cn); free(st);
}

HKMO_API void hkmo_lloyd_pass(const double* X, const double* w, int64_t n, int d, const double* C, int k,
                              int P, int use_stats, int form, int threads,
                              double* sums, double* wsum, double* cost_out, int32_t* labels) {
  lloyd_pass_impl(X, w, n, d, C, NULL, k, P, use_stats, form, threads, sums, wsum, cost_out, labels);
}

/* Cosine driver side: CosineDistanceMeasure.centroid = scal(1/w, sum); scal(1/norm(sum), sum); VectorWithNorm(sum, 1)
 * -- the stored norm is exactly 1.0 from then on (cn[j]); isCenterConverged = distance(old, new) <= epsilon (NOT squared). */
HKMO_API int hkmo_update_centers_cosine(const double* sums, const double* wsum, int k, int d, double tol, double* C, double* cn) {
  int converged = 1;
  double* nc = (double*)malloc(sizeof(double) * d);
  for (int j = 0; j < k; ++j) {
    if (!(wsum[j] > 0)) continue;
    double inv = 1.0 / wsum[j];
    for (int f = 0; f < d; ++f) nc[f] = sums[(size_t)j * d + f] * inv;
    double inv2 = 1.0 / vnorm2(nc, d);
    for (int f = 0; f < d; ++f) nc[f] *= inv2;
    double* c = C + (size_t)j * d;
    if (converged && !(fast_sqdist(c, cn[j], nc, 1.0, d, HKMO_FORM_COSINE) <= tol)) converged = 0;
    memcpy(c, nc, sizeof(double) * d); cn[j] = 1.0;
  }
  free(nc); return converged;
}

/* Driver side of the iteration: centroid = scal(1/w, sum); converged iff every updated centre
 * moved by sqdist <= eps^2; clusters with no points keep their centre. */
HKMO_API int hkmo_update_centers(const double* sums, const double* wsum, int k, int d, double tol, int form, double* C) {
  int converged = 1;
  double* nc = (double*)malloc(sizeof(double) * d);
  for (int j = 0; j < k; ++j) {
    if (!(wsum[j] > 0)) continue;
    double inv = 1.0 / wsum[j];
    for (int f = 0; f < d; ++f) nc[f] = sums[(size_t)j * d + f] * inv;
    double* c = C + (size_t)j * d;
    if (converged && !(fast_sqdist(nc, vnorm2(nc, d), c, vnorm2(c, d), d, form) <= tol * tol)) converged = 0;
    memcpy(c, nc, sizeof(double) * d);
  }
  free(nc); return converged;
}

/* runAlgorithmWithWeight loop: while (iteration < maxIterations && !converged).  cost is the
 * cost accumulated in the LAST executed iteration (w.r.t. the centres before its update). */
HKMO_API void hkmo_lloyd(const double* X, const double* w, int64_t n, int d, double* C, int k, int max_iter, double tol,
                         int P, int use_stats, int form, int threads, int* num_iter, double* cost_out) {
  double* sums = (double*)malloc(sizeof(double) * (size_t)k * d);
  double* wsum = (double*)malloc(sizeof(double) * k);
  int it = 0, converged = 0; double cost = 0.0;
  double* cn = (double*)malloc(sizeof(double) * k);
  for (int j = 0; j < k; ++j) cn[j] = vnorm2(C + (size_t)j * d, d);   /* new VectorWithNorm(initial centre) */
  while (it < max_iter && !converged) {
    if (form == HKMO_FORM_COSINE) {
      lloyd_pass_impl(X, w, n, d, C, cn, k, P, 0, form, threads, sums, wsum, &cost, NULL);
      converged = hkmo_update_centers_cosine(sums, wsum, k, d, tol, C, cn);
    } else {
      hkmo_lloyd_pass(X, w, n, d, C, k, P, use_stats, form, threads, sums, wsum, &cost, NULL);
      converged = hkmo_update_centers(sums, wsum, k, d, tol, form, C);
    }
    ++it;
  }
  free(cn);
  *num_iter = it; *cost_out = cost; free(sums); free(wsum);
}

/* KMeansModel.predict / transform: findClosest index (statistics used when k < 1000 upstream). */
HKMO_API void hkmo_predict(const double* X, int64_t n, int d, const double* C, int k, int use_stats, int form, int threads,
                           int32_t* labels, double* dist /* nullable */) {
  double* cn = (double*)malloc(sizeof(double) * k);
  for (int i = 0; i < k; ++i) cn[i] = vnorm2(C + (size_t)i * d, d);
  double* st = NULL;
  if (use_stats && k > 1 && form != HKMO_FORM_COSINE) { st = (double*)malloc(sizeof(double) * (k * (k + 1) / 2)); hkmo_compute_statistics(C, k, d, form, st); }
#ifdef _OPENMP
  if (threads <= 0) threads = omp_get_max_threads();
#pragma omp parallel for schedule(static) num_threads(threads)
#endif
  for (int64_t r = 0; r < n; ++r) {
    const double* x = X + (size_t)r * d; double dd;
    int b = st ? find_closest_stats(C, cn, st, k, d, x, vnorm2(x, d), form, &dd) : find_closest(C, cn, k, d, x, vnorm2(x, d), form, &dd);
    labels[r] = b; if (dist) dist[r] = dd;
  }
  free(cn); free(st);
}

/* ------------------------------------------------------------------------------------------
 * RDD.takeSample(withReplacement=false, num, seed)  (core/rdd/RDD.scala) with
 * SamplingUtils.computeFractionForSampleSize, PartitionwiseSampledRDD, BernoulliSampler /
 * GapSampling (util/random/RandomSampler.scala), Utils.randomizeInPlace.
 * Returns the number of row indices written to idx_out (<= num).
 * ---------------------------------------------------------------------------------------- */
static int64_t bernoulli_sample_partition(int64_t lo, int64_t hi, double fraction, int64_t pseed, int64_t* out) {
  int64_t m = 0;
  if (fraction <= 0.0) return 0;
  if (fraction >= 1.0) { for (int64_t r = lo; r < hi; ++r) out[m++] = r; return m; }
  hkmo_xorshift rng; hkmo_xorshift_init(&rng, pseed);     /* rng.setSeed(split.seed) */
  if (fraction <= 0.4) {                                  /* defaultMaxGapSamplingFraction */
    double lnq = log1p(-fraction);
    int32_t drop; int started = 0;
    for (int64_t r = lo; r < hi; ++r) {
      if (!started) {                                     /* lazy val gapSampling: advance() at construction */
        double u = fmax(hkmo_xorshift_next_double(&rng), 5e-11);
        double q = log(u) / lnq; drop = q >= 2147483647.0 ? 2147483647 : (int32_t)q; started = 1;
      }
      if (drop > 0) { --drop; }
      else {
        double u = fmax(hkmo_xorshift_next_double(&rng), 5e-11);
        double q = log(u) / lnq; drop = q >= 2147483647.0 ? 2147483647 : (int32_t)q;
        out[m++] = r;
      }
    }
  } else {
    for (int64_t r = lo; r < hi; ++r) if (hkmo_xorshift_next_double(&rng) <= fraction) out[m++] = r;
  }
  return m;
}

HKMO_API int64_t hkmo_take_sample(int64_t n, int P, int num, int64_t seed, int64_t* idx_out) {
  if (num == 0 || n == 0) return 0;
  hkmo_jrandom rand; hkmo_jrandom_init(&rand, seed);
  int64_t* samples = (int64_t*)malloc(sizeof(int64_t) * (size_t)n);
  int64_t m = 0;
  if (num >= n) {
    for (int64_t r = 0; r < n; ++r) samples[r] = r;
    m = n;
    for (int64_t i = m - 1; i >= 1; --i) {
      int32_t j = hkmo_jrandom_next_int_bound(&rand, (int32_t)(i + 1));
      int64_t t = samples[j]; samples[j] = samples[i]; samples[i] = t;
    }
    memcpy(idx_out, samples, sizeof(int64_t) * (size_t)m); free(samples); return m;
  }
  double fr0 = (double)num / (double)n, gamma = -log(1e-4) / (double)n;  /* BinomialBounds.getUpperBound */
  double fraction = fmin(1.0, fmax(1e-10, fr0 + gamma + sqrt(gamma * gamma + 2 * gamma * fr0)));
  do {
    int64_t sseed = (int64_t)hkmo_jrandom_next_int(&rand);                /* this.sample(false, fraction, rand.nextInt()) */
    hkmo_jrandom prng; hkmo_jrandom_init(&prng, sseed);                   /* PartitionwiseSampledRDD.getPartitions */
    m = 0;
    for (int p = 0; p < P; ++p) {
      int64_t ps = hkmo_jrandom_next_long(&prng);
      m += bernoulli_sample_partition(part_lo(n, P, p), part_lo(n, P, p + 1), fraction, ps, samples + m);
    }
  } while (m < num);
  for (int64_t i = m - 1; i >= 1; --i) {                                  /* Utils.randomizeInPlace(samples, rand) */
    int32_t j = hkmo_jrandom_next_int_bound(&rand, (int32_t)(i + 1));
    int64_t t = samples[j]; samples[j] = samples[i]; samples[i] = t;
  }
  memcpy(idx_out, samples, sizeof(int64_t) * (size_t)num);
  free(samples); return num;
}

/* distinct on dense vectors, first occurrence kept (Scala Seq.distinct). Returns new count. */
static int distinct_rows(double* V, int m, int d) {
  int out = 0;
  for (int i = 0; i < m; ++i) {
    int dup = 0;
    for (int j = 0; j < out && !dup; ++j) {
      int eq = 1;
      for (int f = 0; f < d; ++f) if (!(V[(size_t)i * d + f] == V[(size_t)j * d + f])) { eq = 0; break; }
      dup = eq;
    }
    if (!dup) { if (out != i) memmove(V + (size_t)out * d, V + (size_t)i * d, sizeof(double) * d); ++out; }
  }
  return out;
}

/* KMeans.initRandom: takeSample(false, k, new XORShiftRandom(seed).nextInt()), distinct. */
HKMO_API int hkmo_init_random(const double* X, int64_t n, int d, int k, int64_t seed, int P, double* C_out) {
  hkmo_xorshift r; hkmo_xorshift_init(&r, seed);
  int32_t s2 = hkmo_xorshift_next_int(&r);
  int64_t* idx = (int64_t*)malloc(sizeof(int64_t) * (size_t)(k > n ? n : k) + 8);
  int64_t m = hkmo_take_sample(n, P, k, (int64_t)s2, idx);
  for (int64_t i = 0; i < m; ++i) memcpy(C_out + (size_t)i * d, X + (size_t)idx[i] * d, sizeof(double) * d);
  free(idx);
  return distinct_rows(C_out, (int)m, d);
}

/* ------------------------------------------------------------------------------------------
 * mllib/clustering/LocalKMeans.scala: kMeansPlusPlus(seed, points, weights, k, maxIterations)
 * ---------------------------------------------------------------------------------------- */
HKMO_API void hkmo_local_kmeanspp(int32_t seed, const double* pts, const double* wts, int m, int d, int k, int max_iter, int form, double* centers) {
  hkmo_jrandom rand; hkmo_jrandom_init(&rand, (int64_t)seed);
  double* pn = (double*)malloc(sizeof(double) * m);
  for (int i = 0; i < m; ++i) pn[i] = vnorm2(pts + (size_t)i * d, d);
  /* pickWeighted */
  double wsum = 0; for (int i = 0; i < m; ++i) wsum += wts[i];
  {
    double r = hkmo_jrandom_next_double(&rand) * wsum, cur = 0; int i = 0;
    while (i < m && cur < r) { cur += wts[i]; ++i; }
    memcpy(centers, pts + (size_t)(i - 1 < 0 ? 0 : i - 1) * d, sizeof(double) * d);
  }
  double* cost = (double*)malloc(sizeof(double) * m);
  for (int i = 0; i < m; ++i) cost[i] = fast_sqdist(pts + (size_t)i * d, pn[i], centers, vnorm2(centers, d), d, form);
  for (int c = 1; c < k; ++c) {
    double sum = 0; for (int i = 0; i < m; ++i) sum += cost[i] * wts[i];
    double r = hkmo_jrandom_next_double(&rand) * sum, cum = 0; int j = 0;
    while (j < m && cum < r) { cum += wts[j] * cost[j]; ++j; }
    const double* pick = pts + (size_t)(j == 0 ? 0 : j - 1) * d;
    double* cc = centers + (size_t)c * d; memcpy(cc, pick, sizeof(double) * d);
    double ccn = vnorm2(cc, d);
    for (int p = 0; p < m; ++p) { double v = fast_sqdist(pts + (size_t)p * d, pn[p], cc, ccn, d, form); if (v < cost[p]) cost[p] = v; }
  }
  int* old = (int*)malloc(sizeof(int) * m); for (int i = 0; i < m; ++i) old[i] = -1;
  double* counts = (double*)malloc(sizeof(double) * k);
  double* sums = (double*)malloc(sizeof(double) * (size_t)k * d);
  double* cn = (double*)malloc(sizeof(double) * k);
  int iteration = 0, moved = 1;
  while (moved && iteration < max_iter) {
    moved = 0;
    memset(counts, 0, sizeof(double) * k); memset(sums, 0, sizeof(double) * (size_t)k * d);
    for (int j = 0; j < k; ++j) cn[j] = vnorm2(centers + (size_t)j * d, d);
    for (int i = 0; i < m; ++i) {
      double dd; int idx = find_closest(centers, cn, k, d, pts + (size_t)i * d, pn[i], form, &dd);
      for (int f = 0; f < d; ++f) sums[(size_t)idx * d + f] += wts[i] * pts[(size_t)i * d + f];
      counts[idx] += wts[i];
      if (idx != old[i]) { moved = 1; old[i] = idx; }
    }
    for (int j = 0; j < k; ++j) {
      if (counts[j] == 0.0) memcpy(centers + (size_t)j * d, pts + (size_t)hkmo_jrandom_next_int_bound(&rand, m) * d, sizeof(double) * d);
      else { double inv = 1.0 / counts[j]; for (int f = 0; f < d; ++f) centers[(size_t)j * d + f] = sums[(size_t)j * d + f] * inv; }
    }
    ++iteration;
  }
  free(pn); free(cost); free(old); free(counts); free(sums); free(cn);
}

/* ------------------------------------------------------------------------------------------
 * KMeans.initKMeansParallel (mllib/clustering/KMeans.scala).  first_idx >= 0 overrides the
 * takeSample draw (keeps downstream tests independent of the unverifiable sampler).
 * Outputs: C_out (capacity cap rows), returns the number of centres (<= k).
 * Optional debug: cand_out/cand_count (all distinct candidates), step_counts[init_steps].
 * ---------------------------------------------------------------------------------------- */
HKMO_API int hkmo_init_kmeans_parallel(const double* X, int64_t n, int d, int k, int init_steps, int64_t seed, int P,
                                       int64_t first_idx, int form, int threads,
                                       double* C_out, double* cand_out, int cand_cap, int* cand_count, int64_t* chosen_idx_out, int* step_counts) {
  hkmo_xorshift r0; hkmo_xorshift_init(&r0, seed);
  int32_t seed2 = hkmo_xorshift_next_int(&r0);
  int64_t first = first_idx;
  if (first < 0) { int64_t t; if (hkmo_take_sample(n, P, 1, (int64_t)seed2, &t) < 1) return 0; first = t; }
  double* costs = (double*)malloc(sizeof(double) * (size_t)n);
  for (int64_t i = 0; i < n; ++i) costs[i] = INFINITY;
  int cap = cand_cap > 0 ? cand_cap : 1024, m = 0;
  double* cand = (double*)malloc(sizeof(double) * (size_t)cap * d);
  memcpy(cand, X + (size_t)first * d, sizeof(double) * d); m = 1;
  if (chosen_idx_out) chosen_idx_out[0] = first;
  int new_lo = 0, new_hi = 1;
  double* psum = (double*)malloc(sizeof(double) * P);
  int64_t** pch = (int64_t**)calloc(P, sizeof(int64_t*)); int64_t* pcn = (int64_t*)calloc(P, sizeof(int64_t));
#ifdef _OPENMP
  if (threads <= 0) threads = omp_get_max_threads();
#endif
  for (int step = 0; step < init_steps; ++step) {
    int nn = new_hi - new_lo;
    const double* NC = cand + (size_t)new_lo * d;
    double* ncn = (double*)malloc(sizeof(double) * (nn > 0 ? nn : 1));
    for (int i = 0; i < nn; ++i) ncn[i] = vnorm2(NC + (size_t)i * d, d);
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads)
    for (int p = 0; p < P; ++p) {
      int64_t lo = part_lo(n, P, p), hi = part_lo(n, P, p + 1); double s = 0;
      for (int64_t r = lo; r < hi; ++r) {
        if (nn > 0) {
          double dd; find_closest(NC, ncn, nn, d, X + (size_t)r * d, vnorm2(X + (size_t)r * d, d), form, &dd);
          if (dd < costs[r]) costs[r] = dd;               /* math.min(pointCost(newCenters, point), cost) */
        }
        s += costs[r];
      }
      psum[p] = s;
    }
    free(ncn);
    double sum_costs = 0; for (int p = 0; p < P; ++p) sum_costs += psum[p];
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads)
    for (int p = 0; p < P; ++p) {
      int64_t lo = part_lo(n, P, p), hi = part_lo(n, P, p + 1);
      int32_t s32 = seed2 ^ (int32_t)((uint32_t)step << 16) ^ (int32_t)p;   /* Int arithmetic, then widened */
      hkmo_xorshift rng; hkmo_xorshift_init(&rng, (int64_t)s32);
      int64_t cnt = 0, capp = 16; int64_t* buf = (int64_t*)malloc(sizeof(int64_t) * capp);
      for (int64_t r = lo; r < hi; ++r) {
        if (hkmo_xorshift_next_double(&rng) < 2.0 * costs[r] * k / sum_costs) {
          if (cnt == capp) { capp *= 2; buf = (int64_t*)realloc(buf, sizeof(int64_t) * capp); }
          buf[cnt++] = r;
        }
      }
      pch[p] = buf; pcn[p] = cnt;
    }
    new_lo = m;
    for (int p = 0; p < P; ++p) {
      for (int64_t i = 0; i < pcn[p]; ++i) {
        if (m == cap) { cap *= 2; cand = (double*)realloc(cand, sizeof(double) * (size_t)cap * d); }
        memcpy(cand + (size_t)m * d, X + (size_t)pch[p][i] * d, sizeof(double) * d);
        if (chosen_idx_out && m < cand_cap) chosen_idx_out[m] = pch[p][i];
        ++m;
      }
      free(pch[p]); pch[p] = NULL;
    }
    new_hi = m;
    if (step_counts) step_counts[step] = new_hi - new_lo;
  }
  free(costs); free(psum); free(pch); free(pcn);
  m = distinct_rows(cand, m, d);
  if (cand_count) *cand_count = m;
  if (cand_out) memcpy(cand_out, cand, sizeof(double) * (size_t)(m < cand_cap ? m : cand_cap) * d);
  int kout;
  if (m <= k) { memcpy(C_out, cand, sizeof(double) * (size_t)m * d); kout = m; }
  else {
    int32_t* lab = (int32_t*)malloc(sizeof(int32_t) * (size_t)n);
    hkmo_predict(X, n, d, cand, m, 0, form, threads, lab, NULL);          /* findClosest(candidates, x)._1, countByValue */
    double* wts = (double*)calloc(m, sizeof(double));
    for (int64_t i = 0; i < n; ++i) wts[lab[i]] += 1.0;
    /* LocalKMeans works with EuclideanDistanceMeasure.fastSquaredDistance whatever the distance measure of the run */
    hkmo_local_kmeanspp(0, cand, wts, m, d, k, 30, form == HKMO_FORM_COSINE ? 0 : form, C_out);
    free(lab); free(wts); kout = k;
  }
  free(cand);
  return kout;
}

/* ------------------------------------------------------------------------------------------
 * ml/evaluation/ClusteringMetrics.scala: SquaredEuclideanSilhouette (closed form, two passes)
 * labels may be any int32 ids in [0, kmax); only ids that occur take part.
 * ---------------------------------------------------------------------------------------- */
HKMO_API double hkmo_silhouette_sqeuclid(const double* X, const int32_t* labels, const double* w, int64_t n, int d, int kmax, int P, int threads) {
  size_t per = (size_t)kmax * (d + 2);
  double* part = (double*)calloc((size_t)P * per, sizeof(double));
#ifdef _OPENMP
  if (threads <= 0) threads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads)
#endif
  for (int p = 0; p < P; ++p) {
    double* st = part + (size_t)p * per;
    for (int64_t r = part_lo(n, P, p); r < part_lo(n, P, p + 1); ++r) {
      const double* x = X + (size_t)r * d; double wt = w ? w[r] : 1.0;
      double sn = pow(vnorm2(x, d), 2.0);                 /* math.pow(Vectors.norm(features, 2.0), 2.0) */
      double* s = st + (size_t)labels[r] * (d + 2);
      for (int f = 0; f < d; ++f) s[f] += wt * x[f];
      s[d] += sn * wt; s[d + 1] += wt;
    }
  }
  double* st = (double*)calloc(per, sizeof(double));
  for (int p = 0; p < P; ++p) for (size_t i = 0; i < per; ++i) st[i] += part[(size_t)p * per + i];
  free(part);
  int present = 0; for (int c = 0; c < kmax; ++c) if (st[(size_t)c * (d + 2) + d + 1] > 0) ++present;
  if (present < 2) { free(st); return NAN; }              /* require(clustersStatsMap.size > 1) */
  double* ps = (double*)calloc((size_t)P * 2, sizeof(double));
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads)
#endif
  for (int p = 0; p < P; ++p) {
    double a0 = 0, a1 = 0;
    for (int64_t r = part_lo(n, P, p); r < part_lo(n, P, p + 1); ++r) {
      const double* x = X + (size_t)r * d; double wt = w ? w[r] : 1.0;
      double sn = pow(vnorm2(x, d), 2.0);
      int own = labels[r];
      double ws_own = st[(size_t)own * (d + 2) + d + 1], s;
      if (ws_own == wt) s = 0.0;
      else {
        double nb = INFINITY, cur = 0;
        for (int c = 0; c < kmax; ++c) {
          const double* cs = st + (size_t)c * (d + 2);
          if (!(cs[d + 1] > 0)) continue;
          double dot = 0; for (int f = 0; f < d; ++f) dot += x[f] * cs[f];
          double v = sn + cs[d] / cs[d + 1] - 2 * dot / cs[d + 1];
          if (c == own) cur = v * ws_own / (ws_own - wt); else if (v < nb) nb = v;
        }
        if (cur < nb) s = 1 - (cur / nb); else if (cur > nb) s = (nb / cur) - 1; else s = 0.0;
      }
      a0 += s * wt; a1 += wt;
    }
    ps[2 * p] = a0; ps[2 * p + 1] = a1;
  }
  double num = 0, den = 0; for (int p = 0; p < P; ++p) { num += ps[2 * p]; den += ps[2 * p + 1]; }
  free(ps); free(st);
  return num / den;
}

/* CosineSilhouette: normalised features x/||x||; D(x,G) = 1 - x_hat . Omega_G / N_G */
HKMO_API double hkmo_silhouette_cosine(const double* X, const int32_t* labels, const double* w, int64_t n, int d, int kmax) {
  size_t per = (size_t)kmax * (d + 1);
  double* st = (double*)calloc(per, sizeof(double));
  for (int64_t r = 0; r < n; ++r) {
    const double* x = X + (size_t)r * d; double wt = w ? w[r] : 1.0, nm = vnorm2(x, d);
    double* s = st + (size_t)labels[r] * (d + 1);
    for (int f = 0; f < d; ++f) s[f] += wt * (x[f] / nm);
    s[d] += wt;
  }
  int present = 0; for (int c = 0; c < kmax; ++c) if (st[(size_t)c * (d + 1) + d] > 0) ++present;
  if (present < 2) { free(st); return NAN; }
  double num = 0, den = 0;
  double* xh = (double*)malloc(sizeof(double) * d);
  for (int64_t r = 0; r < n; ++r) {
    const double* x = X + (size_t)r * d; double wt = w ? w[r] : 1.0, nm = vnorm2(x, d);
    for (int f = 0; f < d; ++f) xh[f] = x[f] / nm;
    int own = labels[r]; double ws_own = st[(size_t)own * (d + 1) + d], s;
    if (ws_own == wt) s = 0.0;
    else {
      double nb = INFINITY, cur = 0;
      for (int c = 0; c < kmax; ++c) {
        const double* cs = st + (size_t)c * (d + 1);
        if (!(cs[d] > 0)) continue;
        double dot = 0; for (int f = 0; f < d; ++f) dot += xh[f] * cs[f];
        double v = 1 - dot / cs[d];
        if (c == own) cur = v * ws_own / (ws_own - wt); else if (v < nb) nb = v;
      }
      if (cur < nb) s = 1 - (cur / nb); else if (cur > nb) s = (nb / cur) - 1; else s = 0.0;
    }
    num += s * wt; den += wt;
  }
  free(xh); free(st);
  return num / den;
}

/* ------------------------------------------------------------------------------------------
 * ml/stat/Summarizer.scala SummarizerBuffer (mean + std, unit weights): per-partition online
 * update that skips zeros, pairwise merge in partition order, unbiased variance clamped >= 0.
 * ml/feature/StandardScaler.scala: scale = std==0 ? 0 : 1/std;  y = (x - shift) * scale.
 * ---------------------------------------------------------------------------------------- */
HKMO_API void hkmo_scaler_fit(const double* X, int64_t n, int d, int P, double* mean_out, double* std_out) {
  double* mean = (double*)calloc(d, sizeof(double)); double* m2n = (double*)calloc(d, sizeof(double));
  double* nnz = (double*)calloc(d, sizeof(double)); double tot = 0;
  double* pm = (double*)malloc(sizeof(double) * d); double* p2 = (double*)malloc(sizeof(double) * d); double* pz = (double*)malloc(sizeof(double) * d);
  for (int p = 0; p < P; ++p) {
    memset(pm, 0, sizeof(double) * d); memset(p2, 0, sizeof(double) * d); memset(pz, 0, sizeof(double) * d);
    double ptot = 0;
    for (int64_t r = part_lo(n, P, p); r < part_lo(n, P, p + 1); ++r) {
      const double* x = X + (size_t)r * d;
      for (int f = 0; f < d; ++f) {
        double v = x[f];
        if (v != 0.0) {                                   /* add(): foreachNonZero */
          double prev = pm[f], diff = v - prev;
          pm[f] = prev + 1.0 * diff / (pz[f] + 1.0);
          p2[f] += 1.0 * (v - pm[f]) * diff;
          pz[f] += 1.0;
        }
      }
      ptot += 1.0;
    }
    if (ptot == 0) continue;
    if (tot == 0) { memcpy(mean, pm, sizeof(double) * d); memcpy(m2n, p2, sizeof(double) * d); memcpy(nnz, pz, sizeof(double) * d); tot = ptot; continue; }
    for (int f = 0; f < d; ++f) {                         /* merge() */
      double tn = nnz[f], on = pz[f], tt = tn + on;
      if (tt != 0.0) {
        double delta = pm[f] - mean[f];
        mean[f] += delta * on / tt;
        m2n[f] += p2[f] + delta * delta * tn * on / tt;
      }
      nnz[f] = tt;
    }
    tot += ptot;
  }
  double denom = tot - (tot / tot);                       /* totalWeightSum - weightSquareSum/totalWeightSum, unit weights */
  for (int f = 0; f < d; ++f) {
    mean_out[f] = mean[f] * (nnz[f] / tot);
    double var = 0;
    if (denom > 0.0) var = fmax((m2n[f] + mean[f] * mean[f] * nnz[f] * (tot - nnz[f]) / tot) / denom, 0.0);
    std_out[f] = sqrt(var);
  }
  free(mean); free(m2n); free(nnz); free(pm); free(p2); free(pz);
}

HKMO_API void hkmo_scaler_transform(const double* X, int64_t n, int d, const double* mean, const double* std, int with_mean, int with_std, double* Y) {
  for (int64_t r = 0; r < n; ++r)
    for (int f = 0; f < d; ++f) {
      double v = X[(size_t)r * d + f];
      double sc = std[f] == 0 ? 0.0 : 1.0 / std[f];
      if (with_mean && with_std) v = (v - mean[f]) * sc;
      else if (with_std) v = v * sc;
      else if (with_mean) v = v - mean[f];
      Y[(size_t)r * d + f] = v;
    }
}

HKMO_API int hkmo_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

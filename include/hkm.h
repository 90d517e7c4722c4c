/*
 * hkm.h -- C ABI of the B200-native KMeans hot path (libhkm.so).
 *
 * Drop-in boundary: the reference (/root/reference/mllearnforhospitalnetwork.py) has no FFI of
 * its own; it drives pyspark.ml Estimators/Transformers/Evaluators (idiom: ref:L135-136
 * VectorAssembler.transform, ref:L146-148 Estimator.fit / Model.transform, ref:L162-163
 * Evaluator.evaluate), which in Spark cross Py4J into the JVM.  These entry points are what a
 * Py4J/JNI replacement of that seam binds for the KMeans path; each comment names the upstream
 * Spark routine whose per-row work it replaces.
 *
 * Conventions
 *  - All pointers are DEVICE pointers unless the name ends in _host.  X is row-major fp64
 *    [n, d], 16-byte aligned.  Labels are int32.  Sizes are int64/int.
 *  - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *  - The library allocates nothing: callers pass `scratch` buffers sized by the matching
 *    *_scratch_bytes() query and zero-filled once before first use (the kernels leave them
 *    zeroed where that matters).
 *  - No entry point synchronises the host.  Return value 0 = ok, < 0 = error; the text of the
 *    last error on the calling thread is hkm_last_error().
 *  - Calls are asynchronous on `stream`; two calls sharing one scratch buffer must be ordered
 *    on the same stream.
 */
#ifndef HKM_H_
#define HKM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define HKM_API __attribute__((visibility("default")))
#else
#define HKM_API
#endif

#define HKM_OK 0
#define HKM_ERR_INVALID (-1)     /* bad argument (null pointer, size, alignment) */
#define HKM_ERR_UNSUPPORTED (-2) /* shape outside what the built kernels cover */
#define HKM_ERR_CUDA (-3)        /* a CUDA runtime call failed */

#define HKM_XNORM_PER_ROW 0
#define HKM_XNORM_STORE 1
#define HKM_XNORM_REUSE 2

/* Device-side iteration state shared by the Lloyd entry points (32 bytes). */
typedef struct hkm_lloyd_state {
  double cost;        /* cost accumulated by the last executed iteration (KMeans.runAlgorithmWithWeight `cost`) */
  int32_t iter;       /* iterations executed so far */
  int32_t converged;  /* 1 once every updated centre moved by <= tol^2 */
  double max_shift2;  /* largest squared centre movement of the last executed iteration */
  int64_t reserved;   /* hkm_lloyd_step_peer: accumulated exchange timing of this rank in ns -- low 32 bits: storing the vector into the peers,
                         high 32 bits: waiting for the peers' flags (both wrap after ~4 s of accumulated time) */
} hkm_lloyd_state;

HKM_API const char* hkm_last_error(void);
HKM_API int hkm_version(void);

/* Largest k and d the library accepts (a (d, k) pair whose tile ring does not fit in shared memory is refused with
 * HKM_ERR_UNSUPPORTED).  One launch scores up to hkm_max_k_single_pass() = 64 centres; hkm_assign_reduce, hkm_update_centers,
 * hkm_kmpar_cost_update and hkm_closest_counts take larger k in chunks of 64 centres (a running-minimum assignment pass per chunk,
 * then a labels-given statistics pass per chunk: 2 * ceil(k / 64) passes over X, with labels_out and mind2_out as workspace);
 * the fused single-launch steps (hkm_lloyd_step, hkm_lloyd_step_peer) and the silhouette passes stay at k <= 64. */
HKM_API int hkm_max_k(void);
HKM_API int hkm_max_k_single_pass(void);
HKM_API int hkm_max_d(void);
/* 1 when hkm_assign_reduce runs (d, k) in chunks -- k > 64, d > 96 (general-shape kernels, hkm_generic.cu: plain loads, one warp per
 * row, several times below the HBM rate) or a tile ring that does not fit in shared memory -- and therefore needs labels_out and
 * mind2_out as per-row workspace; such shapes have no fused step (use hkm_assign_reduce + hkm_update_centers) and no silhouette pass. */
HKM_API int hkm_needs_workspace(int d, int k);

/* Tuning / test hook, process wide: pins the kernel family the Lloyd data passes use wherever that family is built for the
 * shape.  0 = pick per shape (default), 1 = gather, 2 = fp64 tensor pipe, 3 = warp-specialised fp64, 4 = pipelined tcgen05
 * screening, 5 = the selection without the tcgen05 variants of round 2, 6 = warp-group tcgen05 screening, 7 = as 0.  Every family
 * gives the same labels and counts, and sums / cost equal to fp64 rounding (tests/test_gpu_lloyd.py, tests/test_gpu_limb.py compare
 * them), so this never changes results beyond the last bits.  No counterpart in Spark. */
HKM_API int hkm_set_lloyd_variant(int variant);

/* Bytes of scratch hkm_assign_reduce needs for (d, k) on the current device: the per-CTA partials, 256 + CTAs * (k*d + 2k + 2) * 8,
 * and behind them a 2 KB record of the table's column ranges.  The record is written by a HKM_XNORM_STORE pass and lets the
 * HKM_XNORM_REUSE passes of a fit on the same scratch form the per-cluster sums on the integer tensor cores (8 < k <= 64, even
 * d <= 64).  A scratch of only the per-CTA size is accepted too; those passes then run the other kernels. */
HKM_API size_t hkm_assign_reduce_scratch_bytes(int d, int k);

/*
 * One data pass of Lloyd's algorithm over this rank's rows.
 * Replaces the mapPartitions body + reduceByKey of mllib KMeans.runAlgorithmWithWeight
 * (findClosest / updateClusterSum / costAccum).
 *   partial_out [k*d + k + 1]: per-cluster sums (row-major k x d), weight sums, then cost.
 *   labels_out  [n] or NULL: argmin index per row (KMeansModel.transform / predict, 8a-9).
 *   mind2_out   [n] or NULL: squared distance to the closest centre (pointCost, 8a-6).
 *   weights     [n] or NULL: per-row weights (weightCol); NULL = 1.0.
 *   xnorm2_mode HKM_XNORM_PER_ROW (0): |x|^2 is recomputed per row inside the pass (always
 *               correct; required for mind2_out).  HKM_XNORM_STORE (1): same, and the pass
 *               also stores sum_i w_i |x_i|^2 to *xnorm2_sum.  HKM_XNORM_REUSE (2): the pass
 *               skips the per-row norms and forms cost = *xnorm2_sum - 2 sum_i w_i score_i
 *               (X and weights must be those of the STORE pass; pass the same scratch as well, see
 *               hkm_assign_reduce_scratch_bytes).  The sum is exact to fp64 rounding of ~1e-16 * sum|x|^2,
 *               so REUSE is meant for cost / sum|x|^2 > 1e-9.
 *   state       NULL, or a device hkm_lloyd_state: when state->converged != 0 the pass is a
 *               no-op (lets a host enqueue maxIter passes without reading back in between).
 *   do_reduce   0 = only labels/mind2 are produced (partial_out may be NULL).
 */
HKM_API int hkm_assign_reduce(const double* X, int64_t n, int d, const double* C, int k, const double* weights,
                      double* partial_out, int32_t* labels_out, double* mind2_out,
                      double* xnorm2_sum, int xnorm2_mode,
                      const hkm_lloyd_state* state, int do_reduce,
                      void* scratch, size_t scratch_bytes, void* stream);

/*
 * One whole Lloyd iteration on a single rank in ONE launch: hkm_assign_reduce followed by the centre update of
 * hkm_update_centers, the latter done by the last CTA to finish (no second launch, no gap).  C_inout is read at
 * the start of the pass and overwritten at its end; state as in hkm_update_centers.  Ranks > 1 need the
 * all-reduce in between and use the two separate entry points.
 */
HKM_API int hkm_lloyd_step(const double* X, int64_t n, int d, double* C_inout, int k, const double* weights,
                           double* partial_out, double* xnorm2_sum, int xnorm2_mode, double tol,
                           hkm_lloyd_state* state, void* scratch, size_t scratch_bytes, void* stream);

/*
 * hkm_lloyd_step for ranks > 1: the k*(d+1)+1 partial sums are all-reduced INSIDE the same launch, over NVLink
 * peer memory, instead of by a separate NCCL call (Spark: reduceByKey + collectAsMap + broadcast per iteration).
 * Every rank owns an exchange buffer of hkm_peer_exchange_bytes(world) bytes that is mapped into all peers
 * (CUDA IPC / symmetric memory; zero it once).  peer_bufs_dev is a DEVICE array of the `world` buffer addresses
 * as seen from this rank.  The last CTA of the pass stores this rank's vector into every peer's buffer, raises a
 * flag, waits for the other ranks' flags and sums the slots in rank order, so all ranks get the bit-identical
 * total and update their (replicated) centres identically -- one launch per Lloyd iteration, no collective.
 * `epoch` must be 1, 2, 3, ... on successive calls and equal on all ranks; all ranks must make the call (a rank
 * that never arrives makes the others give up after a few seconds and set *err_flag = 1).
 */
HKM_API size_t hkm_peer_exchange_bytes(int world);
/* Largest vector (doubles) one exchange slot holds: hkm_lloyd_step_peer needs k*d + k + 1 <= this, larger shapes use
 * hkm_assign_reduce + an all-reduce + hkm_update_centers. */
HKM_API int hkm_peer_exchange_slot_doubles(void);
HKM_API int hkm_lloyd_step_peer(const double* X, int64_t n, int d, double* C_inout, int k, const double* weights,
                                double* partial_out, double* xnorm2_sum, int xnorm2_mode, double tol,
                                hkm_lloyd_state* state, void* scratch, size_t scratch_bytes,
                                void* const* peer_bufs_dev, int rank, int world, unsigned long long epoch,
                                int* err_flag, void* stream);

/*
 * hkm_assign_reduce for callers that cannot allocate device memory.  (No counterpart in the
 * reference or in Spark: it exists for the way this library gets embedded.)
 *
 * Some hosts of this code are handed their device buffers and may not call cudaMalloc themselves:
 * a plugin inside a JVM/Spark executor whose memory pool owns the GPU, a serving process with a
 * pre-sized arena, a loop that must stay off the allocator lock.  They can pass X, C and the output
 * vectors, but not the per-CTA scratch of hkm_assign_reduce -- so here the scratch is a fixed
 * HKM_NOALLOC_SCRATCH_BYTES array that lives inside libhkm.so on every GPU that loads it, until
 * the process exits.  Arguments and results are those of hkm_assign_reduce, minus scratch.
 *
 *  - Limit: the per-CTA partials, 256 + CTAs * (k*d + 2k + 2) * 8 (hkm_assign_reduce_scratch_bytes(d, k) without the
 *    column range record, which this entry point does not use), must fit in
 *    HKM_NOALLOC_SCRATCH_BYTES (4 MiB).  On a 148-SM B200 the pass launches 296 CTAs, so k*d + 2k + 2 <= 1771
 *    (e.g. k = 32, d = 32 is the largest square shape), and 444 CTAs for thin rows (d <= 8 and k <= 8), so
 *    k*d + 2k + 2 <= 1180 there (always met).  Larger shapes -- and every shape hkm_needs_workspace() reports --
 *    are refused with HKM_ERR_UNSUPPORTED and a message naming both sizes.
 *  - The call itself performs no cudaMalloc / cudaFree and calls no allocating library routine.
 *  - The first call on each device looks up the array's device address (one cudaGetSymbolAddress
 *    with that device current).  Make that first call before starting any stream capture.
 *  - All calls on a device share the one array, which still holds the previous call's contents;
 *    the entry point clears the 256-byte header it needs zeroed (on `stream`) before launching.
 *    Two calls that overlap on the GPU -- different streams, no ordering between them -- race
 *    on the array and produce garbage: keep calls to this entry point ordered on one stream.
 */
#define HKM_NOALLOC_SCRATCH_BYTES (4u << 20)
HKM_API int hkm_assign_reduce_noalloc(const double* X, int64_t n, int d, const double* C, int k, const double* weights,
                                      double* partial_out, int32_t* labels_out, double* mind2_out,
                                      double* xnorm2_sum, int xnorm2_mode,
                                      const hkm_lloyd_state* state, int do_reduce, void* stream);

/*
 * Driver side of the iteration (KMeans.runAlgorithmWithWeight after collectAsMap, 8a-5):
 * new_j = sums_j * (1/w_j) for clusters with w_j > 0 (others keep their centre), converged iff
 * every updated centre moved by sqdist <= tol^2; state->cost/iter/converged updated.
 * `partial` is the (all-reduced) vector hkm_assign_reduce produced.  No-op once converged.
 * tol < 0 never converges (fixed iteration count; Spark itself requires tol >= 0).
 */
HKM_API int hkm_update_centers(const double* partial, int k, int d, double* C_inout, double tol,
                       hkm_lloyd_state* state, void* stream);

/* distanceMeasure="cosine" (CosineDistanceMeasure, mllib/clustering/DistanceMeasure.scala): the rows are normalised once
 * (hkm_normalize_rows), the centres are kept at unit length, and the passes above run unchanged on the unit rows --
 * argmin of 1 - x.c equals argmin of |x - c|^2 there, the sums are the sums of unit rows (updateClusterSum) and the
 * cost is half the Euclidean one.  This is the driver step for that mode: centre = normalise(sum / w) (centroid),
 * converged iff every updated centre has cosine distance <= tol to its predecessor (isCenterConverged; tol is NOT
 * squared here), state->cost = partial cost / 2.  tol < 0 never converges. */
HKM_API int hkm_update_centers_cosine(const double* partial, int k, int d, double* C_inout, double tol,
                                      hkm_lloyd_state* state, void* stream);

/* ---------------------------------------------------------------------------------------------
 * k-means|| seeding (mllib KMeans.initKMeansParallel)
 * ------------------------------------------------------------------------------------------- */

/* Scratch bytes for hkm_kmpar_cost_update on the current device. */
HKM_API size_t hkm_kmpar_scratch_bytes(int d);

/*
 * costs[i] = min(costs[i], min_j |x_i - newC_j|^2) and *sum_out = sum_i costs[i]
 * (`costs = data.zip(preCosts).map(min(pointCost(newCenters, point), cost))` + `costs.sum()`).
 * m new centres; any m >= 1 (processed in chunks of hkm_max_k()).
 */
HKM_API int hkm_kmpar_cost_update(const double* X, int64_t n, int d, const double* newC, int m, double* costs_inout,
                                  double* sum_out, void* scratch, size_t scratch_bytes, void* stream);

/*
 * Bernoulli selection of one k-means|| step, partition-exact: row i of partition p is kept iff
 * the i-th nextDouble() of XORShiftRandom(seed ^ (step << 16) ^ (part_index0 + p)) is
 * < 2.0 * costs[i] * k / *sum_costs.  part_rows[P+1] are this rank's partition boundaries (local
 * row offsets); strip_start[P+1] is the prefix sum of ceil(len_p / hkm_kmpar_strip_rows()).
 * Selected local row indices are appended to idx_out (unordered; sort on the host), their number
 * to *count_out (may exceed cap: then re-run with a larger buffer).  The first call on a device
 * also builds the XORShift jump table on `stream`.
 */
HKM_API int hkm_kmpar_sample(const double* costs, int64_t n, const int64_t* part_rows, const int64_t* strip_start,
                             int64_t total_strips, int P, int part_index0, int32_t seed, int step, int k,
                             const double* sum_costs, int64_t* idx_out, int cap, int32_t* count_out, void* stream);
HKM_API int hkm_kmpar_strip_rows(void);
/* p[0 .. n) = value: the +infinity start of the k-means|| cost column (initKMeansParallel: costs = Double.PositiveInfinity). */
HKM_API int hkm_fill_f64(double* p, int64_t n, double value, void* stream);

/*
 * counts_out[j] = number of rows whose closest candidate (strict '<', lowest index on ties) is j
 * (`data.map(findClosest(candidates, _)._1).countByValue()`); m candidates, any m >= 1.
 * labels_work [n] int32 and mind2_work [n] double are caller-provided work arrays.
 */
HKM_API int hkm_closest_counts(const double* X, int64_t n, int d, const double* cand, int m, int32_t* labels_work,
                               double* mind2_work, unsigned long long* counts_out, void* stream);

/*
 * counts_out[j] = number of rows with labels[i] == j, j in [0, m) (labels outside the range are not counted): the cluster sizes of a
 * KMeansSummary (`KMeansModel.summary.clusterSizes`, pyspark.ml.clustering) from the labels of a transform pass
 * (hkm_assign_reduce with do_reduce = 0), so that the summary needs no second reducing pass over the rows.
 */
HKM_API int hkm_label_counts(const int32_t* labels, int64_t n, int m, unsigned long long* counts_out, void* stream);

/* ---------------------------------------------------------------------------------------------
 * ClusteringEvaluator silhouette, squared Euclidean (ml/evaluation/ClusteringMetrics.scala,
 * SURVEY 8a-10 / 8a-11).  Scratch: hkm_assign_reduce_scratch_bytes(d, k).
 * ------------------------------------------------------------------------------------------- */

/* Smallest and largest id of a prediction column in one pass: out2 = {min, max} (device int32[2]; {INT_MAX, INT_MIN} when
 * n = 0).  ClusteringEvaluator needs the id range to size the per-cluster statistics (Spark gets it from groupBy(prediction)). */
HKM_API int hkm_label_range(const int32_t* labels, int64_t n, int32_t* out2, void* stream);

/*
 * computeClusterStats: for every cluster id G in [0, k): Y_G = sum w x, N_G = sum w, Psi_G = sum w |x|^2
 * (|x|^2 as pow(norm(x, 2), 2), like the Scala UDF).  labels outside [0, k) are ignored.
 * stats_out [k*d + 2k + 1]: Y (k x d), N (k), one unused slot, Psi (k).  All-reduce it across ranks
 * before the score pass.
 */
HKM_API int hkm_silhouette_stats(const double* X, const int32_t* labels, const double* weights, int64_t n, int d, int k,
                                 double* stats_out, void* scratch, size_t scratch_bytes, void* stream);

/*
 * computeSilhouetteCoefficient + overallScore: out2[0] = sum_i s_i w_i, out2[1] = sum_i w_i over this
 * rank's rows (score = out2[0] / out2[1] after an all-reduce).  Only clusters with N_G > 0 take part;
 * s_i = 0 when the point is alone in its cluster (N_A == w_i).  work: k*d + 2k doubles.
 */
HKM_API int hkm_silhouette_score(const double* X, const int32_t* labels, const double* weights, int64_t n, int d, int k,
                                 const double* stats, double* work, double* out2, void* scratch, size_t scratch_bytes,
                                 void* stream);

/* ---------------------------------------------------------------------------------------------
 * Loader: VectorAssembler + StandardScaler (ml/feature/{VectorAssembler,StandardScaler}.scala,
 * SURVEY 8a-1 / 8a-2).  Column pointer tables (`*_host`) are HOST arrays of DEVICE pointers.
 * ------------------------------------------------------------------------------------------- */
#define HKM_DTYPE_F64 0
#define HKM_DTYPE_F32 1
#define HKM_DTYPE_I32 2
#define HKM_DTYPE_I64 3
#define HKM_DTYPE_I16 4
#define HKM_DTYPE_I8 5
#define HKM_DTYPE_U8 6 /* also bool */
#define HKM_VALID_NONE 0
#define HKM_VALID_BYTES 1 /* one byte per row, non-zero = present */
#define HKM_VALID_BITS 2  /* Arrow validity bitmap, LSB first */

HKM_API size_t hkm_assemble_scratch_bytes(int64_t n);

/*
 * VectorAssembler.transform (ref:L134-136) with StandardScalerModel.transform fused in:
 * X_out[r, c] = (col_c[r] - shift[c]) * scale[c]  (shift / scale nullable), row-major fp64, written once.
 * A row is invalid when any column is null or NaN.  skip_invalid = 0: every row is written (nulls as
 * NaN) and *invalid_count tells the caller whether handleInvalid='error' must raise;
 * skip_invalid = 1 (handleInvalid='skip' / DataFrame.na.drop, ref:L128): valid rows are compacted in
 * order; *n_out = rows written.  scratch (hkm_assemble_scratch_bytes(n)) is needed for skip only.
 * X_out = NULL with skip_invalid = 0 only counts the invalid rows (nothing is written).
 */
HKM_API int hkm_assemble(const void* const* cols_host, const int32_t* dtypes_host, const void* const* valid_host,
                         const int32_t* vkind_host, int d, int64_t n, const double* shift, const double* scale,
                         int skip_invalid, double* X_out, int64_t* n_out, unsigned long long* invalid_count,
                         void* scratch, size_t scratch_bytes, void* stream);

/*
 * hkm_assemble with the row selection of the reference's batch extraction fused into the compaction pass:
 *   SELECT * FROM t WHERE event_time BETWEEN start AND end   (ref:L123-127; window_col = the timestamps as int64, inclusive
 *                                                              bounds, a null timestamp -- window_valid byte 0 -- never matches)
 *   .na.drop()                                                (ref:L128; besides the assembled columns' own nulls / NaNs, up to 8
 *                                                              more validity byte masks: columns that are not features)
 * followed by VectorAssembler (+ the fused StandardScalerModel).  Kept rows are written in order, *n_out of them;
 * *outside_count = rows the window rejects, *invalid_count = rows inside the window that hold a null; keep_out (nullable, n
 * bytes) marks the kept rows so the caller can filter the table's other columns.  window_col = NULL: no window.
 */
HKM_API int hkm_assemble_filtered(const void* const* cols_host, const int32_t* dtypes_host, const void* const* valid_host,
                                  const int32_t* vkind_host, int d, int64_t n, const double* shift, const double* scale,
                                  const int64_t* window_col, const uint8_t* window_valid, int64_t window_lo, int64_t window_hi,
                                  const uint8_t* const* extra_valid_host, int n_extra, double* X_out, int64_t* n_out,
                                  unsigned long long* invalid_count, unsigned long long* outside_count, uint8_t* keep_out,
                                  void* scratch, size_t scratch_bytes, void* stream);

/*
 * Moments for StandardScaler.fit (Summarizer mean/std): out2d[2c] = sum_r (x_rc - pivot_c),
 * out2d[2c+1] = sum_r (x_rc - pivot_c)^2.  The pivot (any row of the data) keeps the sums free of
 * cancellation; mean = pivot + s1/n, var = (s2 - s1^2/n)/(n-1).  All-reduce out2d across ranks.
 */
HKM_API size_t hkm_moments_scratch_bytes(int d);
HKM_API int hkm_moments(const double* X, int64_t n, int d, const double* pivot, double* out2d, void* scratch,
                        size_t scratch_bytes, void* stream);
HKM_API int hkm_column_moments(const void* const* cols_host, const int32_t* dtypes_host, const void* const* valid_host,
                               const int32_t* vkind_host, int d, int64_t n, const double* pivot, double* out2d,
                               void* scratch, size_t scratch_bytes, void* stream);

/* Driver side of k-means|| (KMeans.initKMeansParallel's last step): LocalKMeans.kMeansPlusPlus(seed, points, weights,
 * k, maxIterations) -- weighted k-means++ with java.util.Random(seed), then Lloyd on the m candidates until no
 * assignment moves or max_iter (Spark passes seed 0, 30).  HOST pointers (m is a few hundred rows: Spark runs this on
 * the driver JVM as well); points [m, d], weights [m], centers_out [k, d].  Always Euclidean, whatever the measure. */
HKM_API int hkm_local_kmeanspp(int32_t seed, const double* points, const double* weights, int m, int d, int k, int max_iter,
                               double* centers_out);

/* Cosine distance measure (ClusteringEvaluator distanceMeasure="cosine"): out = x / |x|_2 per row (multiply by the
 * reciprocal norm); *zero_rows = rows of zero length, for which Spark refuses the cosine distance.  In place allowed. */
HKM_API int hkm_normalize_rows(const double* X_in, double* X_out, int64_t n, int d, unsigned long long* zero_rows, void* stream);

/* StandardScalerModel.transform on an existing matrix: out = (in - shift) * scale (in place allowed). */
HKM_API int hkm_scale_rows(const double* X_in, double* X_out, int64_t n, int d, const double* shift, const double* scale,
                           void* stream);

#ifdef __cplusplus
}
#endif
#endif /* HKM_H_ */

// Columnar -> row-major fp64 loader with fused standardisation (VectorAssembler + StandardScaler,
// SURVEY 8a-1 / 8a-2), plus the moment kernels StandardScaler.fit needs.
//
// Spark materialises one boxed Vector per row on the JVM heap; here d typed columns (int32 / int64 /
// float32 / float64 / int16 / int8 / bool, optional validity bytes or Arrow bitmaps) are transposed
// through shared memory into X[n, d] in HBM in one pass, optionally applying
// y = (x - shift) * scale on the way so the standardised matrix is written exactly once.
// Traffic per row: sum of the input column widths + 8*d bytes written.
#include "../../include/hkm.h"
#include "hkm_common.cuh"

namespace hkm {

constexpr int kMaxCols = 64;
constexpr int kRowsPerBlockSmall = 128, kRowsPerBlockBig = 256;
inline int assemble_rows_per_block(int d) { (void)d; return kRowsPerBlockSmall; }   // measured (also with the warp-per-column loop): 256-row blocks are slower, three resident blocks per SM

struct ColDesc { const void* data; const void* valid; int dtype; int vkind; };
struct AssembleParams {
  ColDesc cols[kMaxCols];
  int d; int64_t n;
  double* X;
  const double* shift; const double* scale;    // nullable, [d]
  const int64_t* block_base;                    // skip mode: first output row of each block (exclusive scan), else null
  unsigned long long* invalid_count;            // rows holding a null or NaN
  int64_t* block_valid;                         // count mode: valid rows per block
  int mode;                                     // 0: write rows as they are (NaN for nulls); 1: compact valid rows; 2: only count
  // row predicate fused into the compaction (hkm_assemble_filtered): `win_col BETWEEN win_lo AND win_hi` (null timestamps never
  // match) and up to kMaxExtraValid further validity masks (na.drop over columns that are not features)
  const int64_t* win_col; const uint8_t* win_valid; int64_t win_lo, win_hi;
  const uint8_t* extra_valid[8]; int n_extra;
  unsigned long long* outside_count;            // rows the window rejects
  uint8_t* keep_out;                            // nullable: 1 for every row that is written
};
constexpr int kMaxExtraValid = 8;

__device__ __forceinline__ double load_col(const ColDesc& c, int64_t i, bool& ok) {
  if (c.vkind == 1) ok = reinterpret_cast<const uint8_t*>(c.valid)[i] != 0;
  else if (c.vkind == 2) ok = (reinterpret_cast<const uint8_t*>(c.valid)[i >> 3] >> (i & 7)) & 1;
  else ok = true;
  double v;
  switch (c.dtype) {
    case HKM_DTYPE_F64: v = reinterpret_cast<const double*>(c.data)[i]; break;
    case HKM_DTYPE_F32: v = (double)reinterpret_cast<const float*>(c.data)[i]; break;
    case HKM_DTYPE_I32: v = (double)reinterpret_cast<const int32_t*>(c.data)[i]; break;
    case HKM_DTYPE_I64: v = (double)reinterpret_cast<const int64_t*>(c.data)[i]; break;
    case HKM_DTYPE_I16: v = (double)reinterpret_cast<const int16_t*>(c.data)[i]; break;
    case HKM_DTYPE_I8: v = (double)reinterpret_cast<const int8_t*>(c.data)[i]; break;
    default: v = (double)reinterpret_cast<const uint8_t*>(c.data)[i]; break;
  }
  if (v != v) ok = false;  // VectorAssembler treats NaN like null
  return ok ? v : __longlong_as_double(0x7ff8000000000000LL);
}

// ROWS rows per block (256 threads): 256 -> one row per thread, every column read is a 1-2 KB contiguous burst;
// 128 -> two columns in flight (used when 256 rows of d doubles would not leave room for two blocks per SM).
template <int ROWS>
__global__ void __launch_bounds__(256) assemble_kernel(const AssembleParams p) {
  extern __shared__ double tile[];            // [ROWS][ld], ld odd -> conflict-free transposed stores
  __shared__ int s_rank[ROWS];
  __shared__ int s_bad[ROWS];
  __shared__ int s_wsum[ROWS / 32];
  constexpr int kRowsPerBlock = ROWS;
  const int d = p.d, ld = d | 1;
  const int64_t row0 = (int64_t)blockIdx.x * kRowsPerBlock;
  const int rows = (int)min((int64_t)kRowsPerBlock, p.n - row0);
  if (threadIdx.x < kRowsPerBlock) {
    // bit 0: null / NaN somewhere in the row; bit 1: outside the time window
    int b0 = 0;
    if (threadIdx.x < rows) {
      const int64_t r = row0 + threadIdx.x;
      if (p.win_col != nullptr) {
        const int64_t ts = p.win_col[r];
        if ((p.win_valid != nullptr && p.win_valid[r] == 0) || ts < p.win_lo || ts > p.win_hi) b0 |= 2;
      }
      for (int e = 0; e < p.n_extra; ++e) if (p.extra_valid[e][r] == 0) b0 |= 1;
    }
    s_bad[threadIdx.x] = b0;
  }
  __syncthreads();
  // a warp takes a column at a time: the element type is decided once per warp and column (not per value), every lane reads
  // ROWS / 32 rows of it back to back, and the values go transposed into the tile (row stride ld is odd: conflict free)
  {
    constexpr int RPL = ROWS / 32;
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const double qnan = __longlong_as_double(0x7ff8000000000000LL);
    for (int c = w; c < d; c += 8) {
      const ColDesc col = p.cols[c];
      const double sh = p.shift != nullptr ? p.shift[c] : 0.0, sc = p.scale != nullptr ? p.scale[c] : 1.0;
      double v[RPL];
      const int64_t i0 = row0 + lane;
      auto fetch = [&](auto* ptr) {
#pragma unroll
        for (int j = 0; j < RPL; ++j) v[j] = (lane + 32 * j < rows) ? (double)ptr[i0 + 32 * j] : 0.0;
      };
      switch (col.dtype) {
        case HKM_DTYPE_F64: fetch(reinterpret_cast<const double*>(col.data)); break;
        case HKM_DTYPE_F32: fetch(reinterpret_cast<const float*>(col.data)); break;
        case HKM_DTYPE_I32: fetch(reinterpret_cast<const int32_t*>(col.data)); break;
        case HKM_DTYPE_I64: fetch(reinterpret_cast<const int64_t*>(col.data)); break;
        case HKM_DTYPE_I16: fetch(reinterpret_cast<const int16_t*>(col.data)); break;
        case HKM_DTYPE_I8: fetch(reinterpret_cast<const int8_t*>(col.data)); break;
        default: fetch(reinterpret_cast<const uint8_t*>(col.data)); break;
      }
#pragma unroll
      for (int j = 0; j < RPL; ++j) {
        const int r = lane + 32 * j;
        if (r < rows) {
          bool ok = v[j] == v[j];                      // VectorAssembler treats NaN like null
          if (col.vkind == 1) ok = ok && reinterpret_cast<const uint8_t*>(col.valid)[i0 + 32 * j] != 0;
          else if (col.vkind == 2) ok = ok && ((reinterpret_cast<const uint8_t*>(col.valid)[(i0 + 32 * j) >> 3] >> ((i0 + 32 * j) & 7)) & 1);
          if (!ok) atomicOr(&s_bad[r], 1);
          if (p.mode != 2) tile[r * ld + c] = ok ? (v[j] - sh) * sc : qnan;
        }
      }
    }
  }
  __syncthreads();
  // rank of every valid row inside the block (exclusive), and the block's valid count
  int valid = 0, rank = 0;
  if (threadIdx.x < kRowsPerBlock) {
    valid = (threadIdx.x < rows && !s_bad[threadIdx.x]) ? 1 : 0;
    const unsigned m = __ballot_sync(0xffffffffu, valid);
    rank = __popc(m & ((1u << (threadIdx.x & 31)) - 1u));
    if ((threadIdx.x & 31) == 0) s_wsum[threadIdx.x >> 5] = __popc(m);
  }
  __syncthreads();
  int nvalid = 0;
#pragma unroll
  for (int w = 0; w < ROWS / 32; ++w) nvalid += s_wsum[w];
  if (threadIdx.x < kRowsPerBlock) {
    for (int w = 0; w < (threadIdx.x >> 5); ++w) rank += s_wsum[w];
    s_rank[threadIdx.x] = valid ? rank : -1;
  }
  if (threadIdx.x == 0) {
    if (p.win_col == nullptr) {
      if (rows - nvalid > 0 && p.invalid_count != nullptr) atomicAdd(p.invalid_count, (unsigned long long)(rows - nvalid));
    } else if (p.invalid_count != nullptr || p.outside_count != nullptr) {
      int nout = 0, nnull = 0;                       // WHERE first, na.drop second: nulls are counted among the rows inside the window
      for (int r = 0; r < rows; ++r) { if (s_bad[r] & 2) ++nout; else if (s_bad[r] & 1) ++nnull; }
      if (nnull && p.invalid_count != nullptr) atomicAdd(p.invalid_count, (unsigned long long)nnull);
      if (nout && p.outside_count != nullptr) atomicAdd(p.outside_count, (unsigned long long)nout);
    }
    if (p.block_valid != nullptr) p.block_valid[blockIdx.x] = nvalid;
  }
  if (p.keep_out != nullptr && p.mode != 2 && threadIdx.x < rows) p.keep_out[row0 + threadIdx.x] = valid ? 1 : 0;
  if (p.mode == 2) return;
  __syncthreads();
  if (p.mode == 0) {
    // a warp writes a row at a time: rows * d doubles, contiguous and coalesced
    double* out = p.X + row0 * d;
    for (int rr = threadIdx.x >> 5; rr < rows; rr += 8)
      for (int c = threadIdx.x & 31; c < d; c += 32) out[(int64_t)rr * d + c] = tile[rr * ld + c];
  } else {
    // compaction: scatter inside shared memory is avoided by reading through the rank table
    double* out = p.X + p.block_base[blockIdx.x] * d;
    for (int rr = threadIdx.x / 32; rr < rows; rr += blockDim.x / 32) {
      const int dst = s_rank[rr];
      if (dst < 0) continue;
      for (int c = threadIdx.x & 31; c < d; c += 32) out[(int64_t)dst * d + c] = tile[rr * ld + c];
    }
  }
}

// exclusive scan of per-block valid counts (one block; n_blocks up to tens of millions is still ~ms)
__global__ void __launch_bounds__(1024) scan_blocks_kernel(const int64_t* __restrict__ counts, int64_t nb, int64_t* __restrict__ base, int64_t* __restrict__ total) {
  __shared__ int64_t s_part[1024];
  const int64_t per = (nb + blockDim.x - 1) / blockDim.x;
  const int64_t lo = min(nb, (int64_t)threadIdx.x * per), hi = min(nb, lo + per);
  int64_t s = 0;
  for (int64_t i = lo; i < hi; ++i) s += counts[i];
  s_part[threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.x == 0) { int64_t run = 0; for (int i = 0; i < (int)blockDim.x; ++i) { const int64_t v = s_part[i]; s_part[i] = run; run += v; } *total = run; }
  __syncthreads();
  int64_t run = s_part[threadIdx.x];
  for (int64_t i = lo; i < hi; ++i) { base[i] = run; run += counts[i]; }
}

// ---- moments: per feature (count, sum (x - K), sum (x - K)^2) with a caller-chosen pivot K ---------------
// Row-major input: thread handles feature t % dp of every (t / dp)-th row of its block's slab.
__global__ void __launch_bounds__(256) moments_rows_kernel(const double* __restrict__ X, int64_t n, int d, const double* __restrict__ pivot,
                                                           double* __restrict__ part /* [grid][2*d] */) {
  extern __shared__ double sm[];   // [rows_par][2*d]
  int dp = 1; while (dp < d) dp <<= 1; if (dp > 256) dp = 256;
  const int f = threadIdx.x % dp, rl = threadIdx.x / dp, rows_par = blockDim.x / dp;
  double s1 = 0, s2 = 0;
  for (int f0 = f; f0 < d; f0 += dp) {   // d > 256 is not a case here, loop kept for generality
    const double K = pivot[f0];
    double a1 = 0, a2 = 0;
    // eight independent loads in flight per thread, four accumulator pairs (fixed order -> deterministic)
    double b1[4] = {0, 0, 0, 0}, b2[4] = {0, 0, 0, 0};
    const int64_t step = (int64_t)gridDim.x * rows_par;
    int64_t r = (int64_t)blockIdx.x * rows_par + rl;
    for (; r + 7 * step < n; r += 8 * step) {
      double v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = __ldcs(X + (r + u * step) * d + f0) - K;
#pragma unroll
      for (int u = 0; u < 8; ++u) { b1[u & 3] += v[u]; b2[u & 3] = fma(v[u], v[u], b2[u & 3]); }
    }
    for (; r < n; r += step) { const double v = X[r * d + f0] - K; b1[0] += v; b2[0] = fma(v, v, b2[0]); }
    a1 = (b1[0] + b1[1]) + (b1[2] + b1[3]); a2 = (b2[0] + b2[1]) + (b2[2] + b2[3]);
    if (f0 == f) { s1 = a1; s2 = a2; }
  }
  if (f < d) { sm[(rl * d + f) * 2] = s1; sm[(rl * d + f) * 2 + 1] = s2; }
  __syncthreads();
  for (int i = threadIdx.x; i < d; i += blockDim.x) {
    double t1 = 0, t2 = 0;
    for (int j = 0; j < rows_par; ++j) { t1 += sm[(j * d + i) * 2]; t2 += sm[(j * d + i) * 2 + 1]; }
    part[((size_t)blockIdx.x * d + i) * 2] = t1; part[((size_t)blockIdx.x * d + i) * 2 + 1] = t2;
  }
}

// one typed column: eight independent loads in flight per thread, four accumulator pairs (fixed order -> deterministic)
template <typename T>
__device__ __forceinline__ void col_moments(const T* __restrict__ x, int64_t n, int64_t r, int64_t step, double K, double& a1, double& a2) {
  double b1[4] = {0, 0, 0, 0}, b2[4] = {0, 0, 0, 0};
  for (; r + 7 * step < n; r += 8 * step) {
    T raw[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) raw[u] = __ldcs(x + r + u * step);
#pragma unroll
    for (int u = 0; u < 8; ++u) { const double v = (double)raw[u] - K; b1[u & 3] += v; b2[u & 3] = fma(v, v, b2[u & 3]); }
  }
  for (; r < n; r += step) { const double v = (double)x[r] - K; b1[0] += v; b2[0] = fma(v, v, b2[0]); }
  a1 = (b1[0] + b1[1]) + (b1[2] + b1[3]); a2 = (b2[0] + b2[1]) + (b2[2] + b2[3]);
}

// Columnar input: blockIdx.y = column, blocks stride over rows.
__global__ void __launch_bounds__(256) moments_cols_kernel(const AssembleParams p, const double* __restrict__ pivot, double* __restrict__ part /* [gridDim.x][2*d] */) {
  __shared__ double s_a[8], s_b[8];
  const int c = blockIdx.y;
  const double K = pivot[c];
  double a1 = 0, a2 = 0;
  {
    const ColDesc& col = p.cols[c];
    const int64_t step = (int64_t)gridDim.x * blockDim.x, r0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (col.vkind != 0) {           // columns with a validity mask: element-wise path (nulls turn the moments NaN, as Summarizer does)
      for (int64_t r = r0; r < p.n; r += step) { bool ok; const double v = load_col(col, r, ok) - K; a1 += v; a2 = fma(v, v, a2); }
    } else {
      switch (col.dtype) {           // the type switch sits outside the streaming loop
        case HKM_DTYPE_F64: col_moments(reinterpret_cast<const double*>(col.data), p.n, r0, step, K, a1, a2); break;
        case HKM_DTYPE_F32: col_moments(reinterpret_cast<const float*>(col.data), p.n, r0, step, K, a1, a2); break;
        case HKM_DTYPE_I32: col_moments(reinterpret_cast<const int32_t*>(col.data), p.n, r0, step, K, a1, a2); break;
        case HKM_DTYPE_I64: col_moments(reinterpret_cast<const int64_t*>(col.data), p.n, r0, step, K, a1, a2); break;
        case HKM_DTYPE_I16: col_moments(reinterpret_cast<const int16_t*>(col.data), p.n, r0, step, K, a1, a2); break;
        case HKM_DTYPE_I8: col_moments(reinterpret_cast<const int8_t*>(col.data), p.n, r0, step, K, a1, a2); break;
        default: col_moments(reinterpret_cast<const uint8_t*>(col.data), p.n, r0, step, K, a1, a2); break;
      }
    }
  }
#pragma unroll
  for (int m = 16; m >= 1; m >>= 1) { a1 += shfl_xor_d(a1, m); a2 += shfl_xor_d(a2, m); }
  if ((threadIdx.x & 31) == 0) { s_a[threadIdx.x >> 5] = a1; s_b[threadIdx.x >> 5] = a2; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double t1 = 0, t2 = 0;
    for (int w = 0; w < 8; ++w) { t1 += s_a[w]; t2 += s_b[w]; }
    part[((size_t)blockIdx.x * p.d + c) * 2] = t1; part[((size_t)blockIdx.x * p.d + c) * 2 + 1] = t2;
  }
}

__global__ void store_i64_kernel(int64_t* p, int64_t v) { *p = v; }

__global__ void moments_fold_kernel(const double* __restrict__ part, int nblocks, int d, double* __restrict__ out /* [2*d] */) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < 2 * d; i += gridDim.x * blockDim.x) {
    double s = 0;
    for (int b = 0; b < nblocks; ++b) s += part[(size_t)b * 2 * d + i];
    out[i] = s;
  }
}

__global__ void __launch_bounds__(256) scale_rows_kernel(const double* __restrict__ in, double* __restrict__ out, int64_t total, int d,
                                                        const double* __restrict__ shift, const double* __restrict__ scale) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % d);
    double v = in[i];
    if (shift != nullptr) v -= shift[c];
    if (scale != nullptr) v *= scale[c];
    out[i] = v;
  }
}

// x_hat = x * (1 / |x|_2) per row (cosine distance measure: Spark normalises every vector before use).
// One warp per row; rows of zero length are counted (Spark: "Cosine distance is not defined for zero-length vectors").
__global__ void __launch_bounds__(256) normalize_rows_kernel(const double* __restrict__ in, double* __restrict__ out, int64_t n, int d,
                                                            unsigned long long* __restrict__ zero_rows) {
  const int lane = threadIdx.x & 31;
  for (int64_t r = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); r < n; r += (int64_t)gridDim.x * 8) {
    double v[3], s = 0;
#pragma unroll
    for (int e = 0; e < 3; ++e) { const int f = lane + 32 * e; v[e] = f < d ? in[r * d + f] : 0.0; s = fma(v[e], v[e], s); }
#pragma unroll
    for (int m = 16; m >= 1; m >>= 1) s += shfl_xor_d(s, m);
    const double nrm = sqrt(s);
    if (nrm == 0.0 && lane == 0) atomicAdd(zero_rows, 1ull);
    const double inv = nrm > 0.0 ? 1.0 / nrm : 0.0;
#pragma unroll
    for (int e = 0; e < 3; ++e) { const int f = lane + 32 * e; if (f < d) out[r * d + f] = v[e] * inv; }
  }
}

static int fill_params(AssembleParams& p, const void* const* cols_host, const int32_t* dtypes_host, const void* const* valid_host,
                       const int32_t* vkind_host, int d, int64_t n, const char* who) {
  if (d < 1 || d > kMaxCols || n < 0 || cols_host == nullptr || dtypes_host == nullptr) { set_error("%s: bad d/n/columns (d <= %d)", who, kMaxCols); return HKM_ERR_INVALID; }
  p.d = d; p.n = n;
  for (int c = 0; c < d; ++c) {
    if (dtypes_host[c] < 0 || dtypes_host[c] > HKM_DTYPE_U8 || (n > 0 && cols_host[c] == nullptr)) { set_error("%s: column %d has a bad dtype/pointer", who, c); return HKM_ERR_INVALID; }
    p.cols[c].data = cols_host[c];
    p.cols[c].dtype = dtypes_host[c];
    p.cols[c].valid = valid_host ? valid_host[c] : nullptr;
    p.cols[c].vkind = (valid_host && vkind_host && valid_host[c]) ? vkind_host[c] : 0;
  }
  return HKM_OK;
}

}  // namespace hkm

using namespace hkm;

extern "C" {

size_t hkm_assemble_scratch_bytes(int64_t n) {
  const int64_t nb = (n + kRowsPerBlockSmall - 1) / kRowsPerBlockSmall;   // sized for the smaller block
  return (size_t)(2 * nb + 4) * sizeof(int64_t);
}

int hkm_assemble(const void* const* cols_host, const int32_t* dtypes_host, const void* const* valid_host, const int32_t* vkind_host,
                 int d, int64_t n, const double* shift, const double* scale, int skip_invalid, double* X_out, int64_t* n_out,
                 unsigned long long* invalid_count, void* scratch, size_t scratch_bytes, void* stream) {
  AssembleParams p{};
  int rc = fill_params(p, cols_host, dtypes_host, valid_host, vkind_host, d, n, "hkm_assemble");
  if (rc != HKM_OK) return rc;
  if (invalid_count == nullptr || n_out == nullptr || (X_out == nullptr && skip_invalid)) { set_error("hkm_assemble: null output"); return HKM_ERR_INVALID; }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  HKM_CUDA_TRY(cudaMemsetAsync(invalid_count, 0, sizeof(unsigned long long), st));
  const int rpb = assemble_rows_per_block(d);
  const int64_t nb = (n + rpb - 1) / rpb;
  p.X = X_out; p.shift = shift; p.scale = scale; p.invalid_count = invalid_count;
  const size_t smem = (size_t)rpb * (d | 1) * sizeof(double);
  static bool attr_done[64];
  int dev = 0; HKM_CUDA_TRY(cudaGetDevice(&dev));
  if (dev >= 64 || !attr_done[dev]) {
    HKM_CUDA_TRY(cudaFuncSetAttribute(assemble_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * (kMaxCols | 1) * (int)sizeof(double)));
    HKM_CUDA_TRY(cudaFuncSetAttribute(assemble_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, 256 * 41 * (int)sizeof(double)));
    if (dev < 64) attr_done[dev] = true;
  }
  auto launch = [&](const AssembleParams& q) {
    if (rpb == 256) assemble_kernel<256><<<(unsigned)nb, 256, smem, st>>>(q); else assemble_kernel<128><<<(unsigned)nb, 256, smem, st>>>(q);
  };
  if (n == 0) { HKM_CUDA_TRY(cudaMemsetAsync(n_out, 0, sizeof(int64_t), st)); return HKM_OK; }
  if (nb > 0x7fffffff) { set_error("hkm_assemble: too many rows for one launch"); return HKM_ERR_UNSUPPORTED; }
  if (!skip_invalid) {
    p.mode = X_out == nullptr ? 2 : 0;   // no matrix asked for: only count the invalid rows
    launch(p);
    HKM_CUDA_TRY(cudaGetLastError());
    store_i64_kernel<<<1, 1, 0, st>>>(n_out, n);   // every row is kept
    HKM_CUDA_TRY(cudaGetLastError());
    return HKM_OK;
  }
  if (scratch == nullptr || scratch_bytes < hkm_assemble_scratch_bytes(n)) { set_error("hkm_assemble: scratch too small"); return HKM_ERR_INVALID; }
  int64_t* counts = reinterpret_cast<int64_t*>(scratch);
  int64_t* base = counts + nb;
  p.mode = 2; p.block_valid = counts; p.invalid_count = invalid_count;
  launch(p);
  HKM_CUDA_TRY(cudaGetLastError());
  scan_blocks_kernel<<<1, 1024, 0, st>>>(counts, nb, base, n_out);
  HKM_CUDA_TRY(cudaGetLastError());
  p.mode = 1; p.block_valid = nullptr; p.block_base = base; p.invalid_count = nullptr;
  launch(p);
  HKM_CUDA_TRY(cudaGetLastError());
  return HKM_OK;
}

int hkm_assemble_filtered(const void* const* cols_host, const int32_t* dtypes_host, const void* const* valid_host, const int32_t* vkind_host,
                          int d, int64_t n, const double* shift, const double* scale, const int64_t* window_col, const uint8_t* window_valid,
                          int64_t window_lo, int64_t window_hi, const uint8_t* const* extra_valid_host, int n_extra, double* X_out,
                          int64_t* n_out, unsigned long long* invalid_count, unsigned long long* outside_count, uint8_t* keep_out,
                          void* scratch, size_t scratch_bytes, void* stream) {
  AssembleParams p{};
  int rc = fill_params(p, cols_host, dtypes_host, valid_host, vkind_host, d, n, "hkm_assemble_filtered");
  if (rc != HKM_OK) return rc;
  if (invalid_count == nullptr || outside_count == nullptr || n_out == nullptr || X_out == nullptr || n_extra < 0 || n_extra > kMaxExtraValid ||
      (n_extra > 0 && extra_valid_host == nullptr)) { set_error("hkm_assemble_filtered: bad arguments (at most %d extra validity masks)", kMaxExtraValid); return HKM_ERR_INVALID; }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  HKM_CUDA_TRY(cudaMemsetAsync(invalid_count, 0, sizeof(unsigned long long), st));
  HKM_CUDA_TRY(cudaMemsetAsync(outside_count, 0, sizeof(unsigned long long), st));
  if (n == 0) { HKM_CUDA_TRY(cudaMemsetAsync(n_out, 0, sizeof(int64_t), st)); return HKM_OK; }
  const int rpb = assemble_rows_per_block(d);
  const int64_t nb = (n + rpb - 1) / rpb;
  if (nb > 0x7fffffff) { set_error("hkm_assemble_filtered: too many rows for one launch"); return HKM_ERR_UNSUPPORTED; }
  if (scratch == nullptr || scratch_bytes < hkm_assemble_scratch_bytes(n)) { set_error("hkm_assemble_filtered: scratch too small"); return HKM_ERR_INVALID; }
  p.X = X_out; p.shift = shift; p.scale = scale;
  p.win_col = window_col; p.win_valid = window_valid; p.win_lo = window_lo; p.win_hi = window_hi;
  p.n_extra = n_extra;
  for (int e = 0; e < n_extra; ++e) { if (extra_valid_host[e] == nullptr) { set_error("hkm_assemble_filtered: null validity mask %d", e); return HKM_ERR_INVALID; } p.extra_valid[e] = extra_valid_host[e]; }
  if (window_col == nullptr) { p.win_col = nullptr; }
  const size_t smem = (size_t)rpb * (d | 1) * sizeof(double);
  HKM_CUDA_TRY(cudaFuncSetAttribute(assemble_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * (kMaxCols | 1) * (int)sizeof(double)));
  int64_t* counts = reinterpret_cast<int64_t*>(scratch);
  int64_t* base = counts + nb;
  // pass 1 counts (per block: rows kept; totals: nulls inside the window, rows outside it), pass 2 writes the kept rows in order
  p.mode = 2; p.block_valid = counts; p.invalid_count = invalid_count; p.outside_count = outside_count;
  if (window_col == nullptr) {            // no window: the plain counting rule (every dropped row is an invalid row)
    p.outside_count = nullptr;
  }
  assemble_kernel<128><<<(unsigned)nb, 256, smem, st>>>(p);
  HKM_CUDA_TRY(cudaGetLastError());
  scan_blocks_kernel<<<1, 1024, 0, st>>>(counts, nb, base, n_out);
  HKM_CUDA_TRY(cudaGetLastError());
  p.mode = 1; p.block_valid = nullptr; p.block_base = base; p.invalid_count = nullptr; p.outside_count = nullptr; p.keep_out = keep_out;
  assemble_kernel<128><<<(unsigned)nb, 256, smem, st>>>(p);
  HKM_CUDA_TRY(cudaGetLastError());
  return HKM_OK;
}

size_t hkm_moments_scratch_bytes(int d) { return (size_t)sm_count() * 4 * 2 * (size_t)d * sizeof(double); }

int hkm_moments(const double* X, int64_t n, int d, const double* pivot, double* out2d, void* scratch, size_t scratch_bytes, void* stream) {
  if (d < 1 || d > 256 || n < 0 || pivot == nullptr || out2d == nullptr || scratch == nullptr || scratch_bytes < hkm_moments_scratch_bytes(d)) {
    set_error("hkm_moments: bad arguments"); return HKM_ERR_INVALID;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int grid = sm_count() * 4;
  int dp = 1; while (dp < d) dp <<= 1;
  const int rows_par = 256 / dp;
  moments_rows_kernel<<<grid, 256, (size_t)rows_par * 2 * d * sizeof(double), st>>>(X, n, d, pivot, reinterpret_cast<double*>(scratch));
  HKM_CUDA_TRY(cudaGetLastError());
  moments_fold_kernel<<<1, 256, 0, st>>>(reinterpret_cast<double*>(scratch), grid, d, out2d);
  HKM_CUDA_TRY(cudaGetLastError());
  return HKM_OK;
}

int hkm_column_moments(const void* const* cols_host, const int32_t* dtypes_host, const void* const* valid_host, const int32_t* vkind_host,
                       int d, int64_t n, const double* pivot, double* out2d, void* scratch, size_t scratch_bytes, void* stream) {
  AssembleParams p{};
  int rc = fill_params(p, cols_host, dtypes_host, valid_host, vkind_host, d, n, "hkm_column_moments");
  if (rc != HKM_OK) return rc;
  if (pivot == nullptr || out2d == nullptr || scratch == nullptr || scratch_bytes < hkm_moments_scratch_bytes(d)) { set_error("hkm_column_moments: bad arguments"); return HKM_ERR_INVALID; }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int gx = sm_count() * 4 / (d < 4 ? 1 : (d < 16 ? 2 : 4));
  moments_cols_kernel<<<dim3(gx, d), 256, 0, st>>>(p, pivot, reinterpret_cast<double*>(scratch));
  HKM_CUDA_TRY(cudaGetLastError());
  moments_fold_kernel<<<1, 256, 0, st>>>(reinterpret_cast<double*>(scratch), gx, d, out2d);
  HKM_CUDA_TRY(cudaGetLastError());
  return HKM_OK;
}

int hkm_normalize_rows(const double* X_in, double* X_out, int64_t n, int d, unsigned long long* zero_rows, void* stream) {
  if (d < 1 || d > 96 || n < 0 || zero_rows == nullptr || (n > 0 && (X_in == nullptr || X_out == nullptr))) { set_error("hkm_normalize_rows: bad arguments (d <= 96)"); return HKM_ERR_INVALID; }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  HKM_CUDA_TRY(cudaMemsetAsync(zero_rows, 0, sizeof(unsigned long long), st));
  if (n == 0) return HKM_OK;
  normalize_rows_kernel<<<sm_count() * 8, 256, 0, st>>>(X_in, X_out, n, d, zero_rows);
  HKM_CUDA_TRY(cudaGetLastError());
  return HKM_OK;
}

int hkm_scale_rows(const double* X_in, double* X_out, int64_t n, int d, const double* shift, const double* scale, void* stream) {
  if (d < 1 || n < 0 || (n > 0 && (X_in == nullptr || X_out == nullptr))) { set_error("hkm_scale_rows: bad arguments"); return HKM_ERR_INVALID; }
  if (n == 0) return HKM_OK;
  scale_rows_kernel<<<sm_count() * 8, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(X_in, X_out, n * d, d, shift, scale);
  HKM_CUDA_TRY(cudaGetLastError());
  return HKM_OK;
}

}  // extern "C"

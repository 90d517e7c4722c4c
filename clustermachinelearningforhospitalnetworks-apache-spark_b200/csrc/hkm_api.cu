// C ABI of libhkm (see include/hkm.h): argument checks, kernel dispatch, small driver kernels.
#include "../../include/hkm.h"
#include "hkm_lloyd_launch.cuh"
#include <cstring>
#include <cstdlib>
#include <mutex>

namespace hkm {

static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof(g_err), fmt, ap); va_end(ap);
}
int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
  return HKM_ERR_CUDA;
}
static int g_lloyd_variant = 0;   // hkm_set_lloyd_variant: 0 pick per shape; 1..5 pin one kernel family (tuning / tests)
int lloyd_variant_override() { return g_lloyd_variant; }
int sm_count() {
  static int cached[64]; static bool have[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (!have[dev]) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
    cached[dev] = v; have[dev] = true;
  }
  return cached[dev];
}

// ------------------------------------------------------------------ Lloyd pass dispatch (per-NT translation units)
static int dispatch_lloyd(const LloydParams& p, const LaunchPlan& pl, int grid, int mode, cudaStream_t st) {
  switch ((p.k + 7) / 8) {
    case 1: return dispatch_lloyd_nt1(p, pl, grid, mode, st);
    case 2: return dispatch_lloyd_nt2(p, pl, grid, mode, st);
    case 3: return dispatch_lloyd_nt3(p, pl, grid, mode, st);
    case 4: return dispatch_lloyd_nt4(p, pl, grid, mode, st);
    case 5: return dispatch_lloyd_nt5(p, pl, grid, mode, st);
    case 6: return dispatch_lloyd_nt6(p, pl, grid, mode, st);
    case 7: return dispatch_lloyd_nt7(p, pl, grid, mode, st);
    case 8: return dispatch_lloyd_nt8(p, pl, grid, mode, st);
  }
  set_error("k=%d outside the fp64 kernel range (max %d)", p.k, 64);
  return HKM_ERR_UNSUPPORTED;
}

// ------------------------------------------------------------------ centre update
__global__ void update_centers_kernel(const double* __restrict__ partial, int k, int d, double* __restrict__ C, double tol2,
                                      hkm_lloyd_state* __restrict__ state) {
  if (state->converged != 0) return;
  __shared__ double s_shift[64];
  update_centers_body(partial, k, d, C, tol2, state, s_shift, threadIdx.x);
}


// CosineDistanceMeasure.centroid / isCenterConverged on partial sums of UNIT rows: new_j = normalise(sums_j * (1/w_j)),
// converged iff every updated centre has 1 - old.new / |old| <= tol (not squared).  The pass ran the Euclidean
// machinery on unit vectors, where |x - c|^2 = 2 (1 - x.c): cost = partial cost / 2.
__global__ void update_centers_cosine_kernel(const double* __restrict__ partial, int k, int d, double* __restrict__ C, double tol,
                                             hkm_lloyd_state* __restrict__ state) {
  if (state->converged != 0) return;
  __shared__ double s_shift[64];
  const int tid = threadIdx.x;
  double shift = -1.0;
  if (tid < 64) {
    for (int j = tid; j < k; j += 64) {
      const double wsum = partial[(size_t)k * d + j];
      if (wsum > 0) {
        const double inv = 1.0 / wsum;
        double nn = 0, oo = 0;
        for (int f = 0; f < d; ++f) { const double v = partial[(size_t)j * d + f] * inv; nn = fma(v, v, nn); const double o = C[(size_t)j * d + f]; oo = fma(o, o, oo); }
        const double inv2 = 1.0 / sqrt(nn), on = sqrt(oo);
        double dot = 0;
        for (int f = 0; f < d; ++f) {
          const double nc = partial[(size_t)j * d + f] * inv * inv2;
          dot = fma(C[(size_t)j * d + f], nc, dot);
          C[(size_t)j * d + f] = nc;
        }
        shift = fmax(shift, fmax(1.0 - dot / on, 0.0));
      }
    }
    s_shift[tid] = shift;
  }
  __syncthreads();
  if (tid == 0) {
    double mx = -1.0;
    for (int i = 0; i < 64; ++i) mx = fmax(mx, s_shift[i]);
    state->cost = 0.5 * partial[(size_t)k * d + k];
    state->iter += 1;
    state->max_shift2 = mx < 0 ? 0.0 : mx;
    state->converged = (mx <= tol) ? 1 : 0;
  }
}

}  // namespace hkm

namespace hkm {

constexpr int kMaxFastK = 64;     // centres one launch handles; larger k runs in chunks (assign_pass_large)
constexpr int kMaxFastD = 96;     // widest row of the tiled kernels; wider rows take the general kernels (hkm_generic.cu)

// general-shape kernels (hkm_generic.cu)
int generic_chunk_k(int d);
size_t generic_scratch_bytes(int d);
int generic_assign(const double* X, int64_t n, int d, const double* C, int k, int32_t* labels, double* mind2, int out_mode, int label_base,
                   const int32_t* skip, double* cost_out, void* scratch, cudaStream_t st);
int generic_reduce(const double* X, int64_t n, int d, int kc, const int32_t* labels, int label_base, const double* w, const double* mind2,
                   const int32_t* skip, void* scratch, double* stats, cudaStream_t st);

size_t assign_reduce_scratch_core(int d, int k);
size_t wl_tail_offset(int d, int k);
// the tiled (TMA ring) kernels take (d, k) when the row fits and at least two ring stages do
static bool fits_fast(int d, int k) {
  if (d > kMaxFastD || k > kMaxFastK) return false;
  return plan_lloyd(d, ((k + 7) / 8) * 8).stages >= 2;
}

int assign_pass(const AssignArgs& a, const char* who) {
  const int d = a.d, k = a.k;
  if (a.n < 0 || d < 1 || k < 1 || (a.n > 0 && a.X == nullptr) || a.C == nullptr) { set_error("%s: bad X/C/n/d/k", who); return HKM_ERR_INVALID; }
  if (d > hkm_max_d() || k > kMaxFastK) { set_error("%s: d=%d k=%d outside fp64 kernel range (d<=%d, k<=%d per launch)", who, d, k, hkm_max_d(), kMaxFastK); return HKM_ERR_UNSUPPORTED; }
  if (!fits_fast(d, k)) {
    // general kernels: label / distance passes (plain or running-minimum) and the labels-given statistics of one chunk
    cudaStream_t st = reinterpret_cast<cudaStream_t>(a.stream);
    const int32_t* skip = a.state ? &a.state->converged : nullptr;
    if (a.mode == 0 && a.weights == nullptr && a.do_reduce != 1 && a.C_update == nullptr && a.peer_bufs == nullptr) {
      if (a.do_reduce == 2 && (a.partial_out == nullptr || a.scratch == nullptr || a.scratch_bytes < generic_scratch_bytes(d))) { set_error("%s: cost sum needs partial_out and scratch", who); return HKM_ERR_INVALID; }
      return generic_assign(a.X, a.n, d, a.C, k, a.labels_out, a.mind2_out, a.out_mode, a.label_base, skip, a.do_reduce == 2 ? a.partial_out : nullptr, a.scratch, st);
    }
    set_error("%s: d=%d k=%d does not fit the tiled kernels (%s); this entry point has no general-shape path", who, d, k, d > kMaxFastD ? "d > 96" : "shared-memory ring");
    return HKM_ERR_UNSUPPORTED;
  }
  if ((reinterpret_cast<uintptr_t>(a.X) & 15) != 0) { set_error("%s: X must be 16-byte aligned", who); return HKM_ERR_INVALID; }
  if (a.do_reduce && (a.partial_out == nullptr || a.scratch == nullptr)) { set_error("%s: reduce needs partial_out and scratch", who); return HKM_ERR_INVALID; }
  if (a.do_reduce && a.scratch_bytes < assign_reduce_scratch_core(d, k)) { set_error("%s: scratch too small (%zu < %zu)", who, a.scratch_bytes, assign_reduce_scratch_core(d, k)); return HKM_ERR_INVALID; }
  if (a.xnorm2_mode < 0 || a.xnorm2_mode > 2 || (a.xnorm2_mode != 0 && a.xnorm2_sum == nullptr)) { set_error("%s: bad xnorm2_mode/xnorm2_sum", who); return HKM_ERR_INVALID; }
  if (a.xnorm2_mode == 2 && a.mind2_out != nullptr) { set_error("%s: mind2_out needs per-row norms (xnorm2_mode 0 or 1)", who); return HKM_ERR_INVALID; }
  if (a.out_mode != 0 && a.mind2_out == nullptr) { set_error("%s: min-merge needs mind2", who); return HKM_ERR_INVALID; }
  const int k_pad = ((k + 7) / 8) * 8;
  LaunchPlan pl = plan_lloyd(d, k_pad);
  if (pl.stages < 2) { set_error("%s: d=%d does not fit the shared-memory ring", who, d); return HKM_ERR_UNSUPPORTED; }
  LloydParams p{};
  p.X = a.X; p.n = a.n; p.d = d; p.k = k; p.C = a.C; p.w = a.weights; p.partial_out = a.partial_out; p.labels_out = a.labels_out;
  p.mind2_out = a.mind2_out; p.skip = a.state ? &a.state->converged : nullptr; p.do_reduce = a.do_reduce;
  p.out_mode = a.out_mode; p.label_base = a.label_base; p.labels_in = a.labels_in; p.aux = a.aux;
  p.C_update = a.C_update; p.tol2 = a.tol < 0 ? -1.0 : a.tol * a.tol; p.state_rw = a.state_rw;
  p.peer_bufs = a.peer_bufs; p.xrank = a.xrank; p.xworld = a.xworld; p.xepoch = a.xepoch; p.xerr = a.xerr;
  p.xnorm2_sum = a.xnorm2_sum; p.xx_mode = a.do_reduce == 1 ? a.xnorm2_mode : 0;
  p.counter = reinterpret_cast<unsigned int*>(a.scratch);
  p.cta_partials = a.scratch ? reinterpret_cast<double*>(reinterpret_cast<unsigned char*>(a.scratch) + 256) : nullptr;
  p.stages = pl.stages; p.ks = (d + 3) / 4;
  const int grid = sm_count() * pl.ctas_per_sm;
  const int mode = (a.mode == 0 && a.weights != nullptr) ? 3 : a.mode;   // weights are a compile-time property of the Lloyd kernels
  if (a.n > 0 && lloyd_wd_shape(p, mode) && (p.do_reduce == 0 || a.scratch != nullptr))   // distance-returning passes with 8 < k <= 64, even d <= 48
    return launch_lloyd_wd(p, reinterpret_cast<cudaStream_t>(a.stream));
  if (lloyd_wl_shape(p, mode) && a.scratch != nullptr && a.scratch_bytes >= wl_tail_offset(d, k) + wl_tail_bytes() && a.n > 0) {
    // limb-reduction variant: the STORE pass also measures the column ranges (one more read of the table, once per fit); REUSE
    // passes run the variant, which declines on its own when the record does not admit the table -- the launch below then works
    void* tail = reinterpret_cast<unsigned char*>(a.scratch) + wl_tail_offset(d, k);
    if (p.xx_mode == 1) { const int rc = launch_column_range(a.X, a.n, d, tail, reinterpret_cast<cudaStream_t>(a.stream)); if (rc != HKM_OK) return rc; }
    if (p.xx_mode == 2) {
      const int rc = launch_lloyd_wl(p, tail, reinterpret_cast<cudaStream_t>(a.stream));
      if (rc != HKM_OK) return rc;
      p.skip2 = reinterpret_cast<const int32_t*>(reinterpret_cast<unsigned char*>(tail) + wl_done_offset());
    }
  }
  if (lloyd_wg_eligible(p, mode))                                        // 8 < k <= 32, d <= 32 plain passes: warp-group tcgen05 screening
    return launch_lloyd_wg(p, reinterpret_cast<cudaStream_t>(a.stream));
  if (lloyd_tp_eligible(p, mode))                                        // k > 8 plain passes: pipelined tcgen05 screening + fp64 refinement
    return launch_lloyd_tp(p, reinterpret_cast<cudaStream_t>(a.stream));
  const int variant = lloyd_variant_override();
  if (lloyd_tc_eligible(p, mode) && variant != 1 && variant != 4 && variant < 40)   // its unpipelined predecessor (weights / per-row distances)
    return launch_lloyd_tc(p, reinterpret_cast<cudaStream_t>(a.stream));
  return dispatch_lloyd(p, pl, grid, mode, reinterpret_cast<cudaStream_t>(a.stream));
}

// ------------------------------------------------------------------ k > 64: the pass in chunks of 64 centres
constexpr int kChunkK = 64;
__global__ void fill_inf_kernel(double* p, int64_t n) {
  const double inf = __longlong_as_double(0x7ff0000000000000LL);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = inf;
}
// chunk statistics [kc*d sums][kc N][-][kc Psi] -> rows c0 .. c0 + kc of partial_out's layout [k*d][k][cost]; the chunk's share of
// cost = sum_j (Psi_j - 2 S_j.c_j + N_j |c_j|^2) = sum w |x - c_label|^2 is added to partial_out[k*d + k]
// (cost_given: the general kernels hand the chunk's cost over directly, in the slot after the weights)
__global__ void place_chunk_kernel(const double* __restrict__ st, int kc, int c0, int k, int d, const double* __restrict__ C, double* __restrict__ out, int cost_given) {
  __shared__ double s_cost[256];
  const int Pc = kc * d + kc + 1;
  double cost = 0;
  for (int i = threadIdx.x; i < kc * d; i += blockDim.x) out[(size_t)c0 * d + i] = st[i];
  for (int j = threadIdx.x; j < kc; j += blockDim.x) {
    const double N = st[kc * d + j];
    out[(size_t)k * d + c0 + j] = N;
    if (!cost_given) {
      double sc = 0, cc = 0;
      for (int f = 0; f < d; ++f) { const double c = C[(size_t)(c0 + j) * d + f]; sc = fma(st[j * d + f], c, sc); cc = fma(c, c, cc); }
      cost += st[Pc + j] - 2.0 * sc + N * cc;
    }
  }
  s_cost[threadIdx.x] = cost;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0;
    for (int i = 0; i < (int)blockDim.x; ++i) t += s_cost[i];
    if (cost_given) t = st[kc * d + kc];
    out[(size_t)k * d + k] = (c0 == 0 ? 0.0 : out[(size_t)k * d + k]) + fmax(t, 0.0);
  }
}

// One data pass for k > 64: running-minimum assignment over chunks of 64 centres (strict '<': the lower index keeps a tie, as
// findClosest), then -- when sums are wanted -- the labels-given statistics pass per chunk.  Needs labels_out and mind2_out as
// per-row workspace.  2 (or 1) passes over X per chunk: a general path, not a fast one.
static int assign_pass_large(const AssignArgs& a, const char* who) {
  const int d = a.d, k = a.k;
  if (d < 1 || k < 1 || d > hkm_max_d() || k > hkm_max_k() || a.n < 0 || (a.n > 0 && a.X == nullptr) || a.C == nullptr) { set_error("%s: bad X/C/n/d/k (d <= %d, k <= %d)", who, hkm_max_d(), hkm_max_k()); return HKM_ERR_INVALID; }
  const bool general = !fits_fast(d, k < kChunkK ? k : kChunkK);       // the chunks themselves do not fit the tiled kernels
  const int chunk = general ? generic_chunk_k(d) : kChunkK;
  if (a.labels_out == nullptr || a.mind2_out == nullptr) { set_error("%s: d=%d k=%d runs in centre chunks (k > %d, or a shape outside the tiled kernels) and needs labels_out and mind2_out (n entries each) as workspace", who, d, k, kChunkK); return HKM_ERR_INVALID; }
  if (a.mode != 0 || a.out_mode != 0 || a.C_update != nullptr || a.peer_bufs != nullptr) { set_error("%s: k=%d > %d supports plain passes only", who, k, kChunkK); return HKM_ERR_UNSUPPORTED; }
  if (a.do_reduce && (a.partial_out == nullptr || a.scratch == nullptr || a.scratch_bytes < assign_reduce_scratch_core(d, k))) { set_error("%s: reduce needs partial_out and scratch of hkm_assign_reduce_scratch_bytes(d, k)", who); return HKM_ERR_INVALID; }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(a.stream);
  if (a.n > 0) { fill_inf_kernel<<<sm_count() * 4, 256, 0, st>>>(a.mind2_out, a.n); HKM_CUDA_TRY(cudaGetLastError()); }
  for (int c0 = 0; c0 < k; c0 += chunk) {
    AssignArgs b{};
    b.X = a.X; b.n = a.n; b.d = d; b.C = a.C + (size_t)c0 * d; b.k = (k - c0 < chunk) ? (k - c0) : chunk;
    b.labels_out = a.labels_out; b.mind2_out = a.mind2_out; b.do_reduce = 0; b.out_mode = 1; b.label_base = c0; b.state = a.state; b.stream = a.stream;
    const int rc = assign_pass(b, who);
    if (rc != HKM_OK) return rc;
  }
  if (!a.do_reduce) return HKM_OK;
  const int32_t* skip = a.state ? &a.state->converged : nullptr;
  // chunk statistics land behind the per-CTA partials of one chunk pass
  const size_t chunk_scratch = general ? generic_scratch_bytes(d) : assign_reduce_scratch_core(d, kChunkK);
  double* stats = reinterpret_cast<double*>(reinterpret_cast<unsigned char*>(a.scratch) + chunk_scratch);
  for (int c0 = 0; c0 < k; c0 += chunk) {
    const int kc = (k - c0 < chunk) ? (k - c0) : chunk;
    if (a.n == 0) HKM_CUDA_TRY(cudaMemsetAsync(stats, 0, sizeof(double) * ((size_t)kc * d + 2 * (size_t)kc + 1), st));
    else if (general) {
      const int rc = generic_reduce(a.X, a.n, d, kc, a.labels_out, c0, a.weights, a.mind2_out, skip, a.scratch, stats, st);
      if (rc != HKM_OK) return rc;
    } else {
      AssignArgs b{};
      b.X = a.X; b.n = a.n; b.d = d; b.C = a.X; b.k = kc; b.weights = a.weights; b.partial_out = stats; b.do_reduce = 1; b.mode = 1;
      b.labels_in = a.labels_out; b.label_base = c0; b.state = a.state; b.scratch = a.scratch; b.scratch_bytes = chunk_scratch; b.stream = a.stream;
      const int rc = assign_pass(b, who); if (rc != HKM_OK) return rc;
    }
    place_chunk_kernel<<<1, 256, 0, st>>>(stats, kc, c0, k, d, a.C, a.partial_out, general && a.n > 0 ? 1 : 0);
    HKM_CUDA_TRY(cudaGetLastError());
  }
  return HKM_OK;
}

size_t assign_reduce_scratch_core(int d, int k) {
  if (d < 1 || k < 1) return 0;
  if (!fits_fast(d, k < 64 ? k : 64)) return generic_scratch_bytes(d) + ((size_t)64 * d + 2 * 64 + 2) * sizeof(double);   // general kernels: CTA partials + chunk statistics
  if (k > 64) return assign_reduce_scratch_core(d, 64) + ((size_t)64 * d + 2 * 64 + 2) * sizeof(double);   // chunk pass + chunk statistics
  const int k_pad = ((k + 7) / 8) * 8;
  LaunchPlan pl = plan_lloyd(d, k_pad);
  size_t grid = (size_t)sm_count() * (pl.ctas_per_sm > 2 ? pl.ctas_per_sm : 2);   // every variant launches at most this many CTAs
  return 256 + grid * ((size_t)k * d + 2 * (size_t)k + 2) * sizeof(double);
}
size_t wl_tail_offset(int d, int k) { return (assign_reduce_scratch_core(d, k) + 255) & ~(size_t)255; }

}  // namespace hkm

using namespace hkm;

extern "C" {

const char* hkm_last_error(void) { return g_err; }
int hkm_version(void) { return 100; }
int hkm_max_k(void) { return 4096; }
int hkm_max_k_single_pass(void) { return 64; }
int hkm_max_d(void) { return 4096; }
int hkm_needs_workspace(int d, int k) { return (k > kMaxFastK || !fits_fast(d, k)) ? 1 : 0; }
int hkm_set_lloyd_variant(int variant) {
  if (!((variant >= 0 && variant <= 7) || (variant == 40 || variant == 42 || variant == 43) || (variant >= 50 && variant <= 54))) { set_error("hkm_set_lloyd_variant: 0..7 (40..43 / 50..54: variant 4 with a pinned reduce scheme / warp budget)"); return HKM_ERR_INVALID; }
  g_lloyd_variant = variant;
  return HKM_OK;
}

// what the kernels' per-CTA partials need; hkm_assign_reduce_scratch_bytes adds the column range record of the limb-reduction
// variant behind it (a scratch of only the core size still works: that variant is then not used)
size_t hkm_assign_reduce_scratch_bytes(int d, int k) {
  if (d < 1 || k < 1) return 0;
  return wl_tail_offset(d, k) + wl_tail_bytes();
}
int hkm_assign_reduce(const double* X, int64_t n, int d, const double* C, int k, const double* weights, double* partial_out,
                      int32_t* labels_out, double* mind2_out, double* xnorm2_sum, int xnorm2_mode, const hkm_lloyd_state* state,
                      int do_reduce, void* scratch, size_t scratch_bytes, void* stream) {
  AssignArgs a{};
  a.X = X; a.n = n; a.d = d; a.C = C; a.k = k; a.weights = weights; a.partial_out = partial_out; a.labels_out = labels_out;
  a.mind2_out = mind2_out; a.xnorm2_sum = xnorm2_sum; a.xnorm2_mode = xnorm2_mode; a.state = state; a.do_reduce = do_reduce ? 1 : 0;
  a.scratch = scratch; a.scratch_bytes = scratch_bytes; a.stream = stream;
  if (k > 64 || (d >= 1 && k >= 1 && !fits_fast(d, k))) return assign_pass_large(a, "hkm_assign_reduce");
  return assign_pass(a, "hkm_assign_reduce");
}

int hkm_lloyd_step(const double* X, int64_t n, int d, double* C_inout, int k, const double* weights, double* partial_out,
                   double* xnorm2_sum, int xnorm2_mode, double tol, hkm_lloyd_state* state, void* scratch, size_t scratch_bytes,
                   void* stream) {
  if (state == nullptr || C_inout == nullptr) { set_error("hkm_lloyd_step: bad arguments"); return HKM_ERR_INVALID; }
  if (k > 64) { set_error("hkm_lloyd_step: k=%d > 64 has no fused step; use hkm_assign_reduce + hkm_update_centers", k); return HKM_ERR_UNSUPPORTED; }
  AssignArgs a{};
  a.X = X; a.n = n; a.d = d; a.C = C_inout; a.k = k; a.weights = weights; a.partial_out = partial_out; a.xnorm2_sum = xnorm2_sum;
  a.xnorm2_mode = xnorm2_mode; a.state = state; a.do_reduce = 1; a.scratch = scratch; a.scratch_bytes = scratch_bytes; a.stream = stream;
  a.C_update = C_inout; a.tol = tol; a.state_rw = state;
  return assign_pass(a, "hkm_lloyd_step");
}

size_t hkm_peer_exchange_bytes(int world) { return world < 1 ? 0 : xchg_buffer_bytes(world); }
int hkm_peer_exchange_slot_doubles(void) { return kXchgSlot; }

int hkm_lloyd_step_peer(const double* X, int64_t n, int d, double* C_inout, int k, const double* weights, double* partial_out,
                        double* xnorm2_sum, int xnorm2_mode, double tol, hkm_lloyd_state* state, void* scratch, size_t scratch_bytes,
                        void* const* peer_bufs_dev, int rank, int world, unsigned long long epoch, int* err_flag, void* stream) {
  if (state == nullptr || C_inout == nullptr || k > 64 || peer_bufs_dev == nullptr || world < 2 || world > 64 || rank < 0 || rank >= world || epoch == 0) {
    set_error("hkm_lloyd_step_peer: bad arguments"); return HKM_ERR_INVALID;
  }
  if ((size_t)k * d + k + 1 > (size_t)kXchgSlot) { set_error("hkm_lloyd_step_peer: k*d + k + 1 exceeds the exchange slot (%d doubles)", kXchgSlot); return HKM_ERR_UNSUPPORTED; }
  AssignArgs a{};
  a.X = X; a.n = n; a.d = d; a.C = C_inout; a.k = k; a.weights = weights; a.partial_out = partial_out; a.xnorm2_sum = xnorm2_sum;
  a.xnorm2_mode = xnorm2_mode; a.state = state; a.do_reduce = 1; a.scratch = scratch; a.scratch_bytes = scratch_bytes; a.stream = stream;
  a.C_update = C_inout; a.tol = tol; a.state_rw = state;
  a.peer_bufs = peer_bufs_dev; a.xrank = rank; a.xworld = world; a.xepoch = epoch; a.xerr = err_flag;
  return assign_pass(a, "hkm_lloyd_step_peer");
}

int hkm_update_centers(const double* partial, int k, int d, double* C_inout, double tol, hkm_lloyd_state* state, void* stream) {
  if (partial == nullptr || C_inout == nullptr || state == nullptr || k < 1 || k > hkm_max_k() || d < 1) { set_error("hkm_update_centers: bad arguments"); return HKM_ERR_INVALID; }
  update_centers_kernel<<<1, 64, 0, reinterpret_cast<cudaStream_t>(stream)>>>(partial, k, d, C_inout, tol < 0 ? -1.0 : tol * tol, state);
  HKM_CUDA_TRY(cudaGetLastError());
  return HKM_OK;
}

int hkm_update_centers_cosine(const double* partial, int k, int d, double* C_inout, double tol, hkm_lloyd_state* state, void* stream) {
  if (partial == nullptr || C_inout == nullptr || state == nullptr || k < 1 || k > hkm_max_k() || d < 1) { set_error("hkm_update_centers_cosine: bad arguments"); return HKM_ERR_INVALID; }
  update_centers_cosine_kernel<<<1, 64, 0, reinterpret_cast<cudaStream_t>(stream)>>>(partial, k, d, C_inout, tol < 0 ? -1.0 : tol, state);
  HKM_CUDA_TRY(cudaGetLastError());
  return HKM_OK;
}

}  // extern "C"

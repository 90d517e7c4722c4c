// Host side of the limb-reduction Lloyd pass (hkm_lloyd_wl.cuh): the column range pass, tensor-map construction, dispatch.
#include "../../include/hkm.h"
#include "hkm_lloyd_launch.cuh"
#include "hkm_lloyd_wl.cuh"

namespace hkm {

static unsigned long long* g_wl_prof = nullptr;
static int g_wl_off = 0;       // developer switch: 1 = never use this variant
static int g_wl_flags = -1;    // developer switch: bit 0 = close the previous tile at the top of the iteration when its product is done
static int g_wl_nohalf = 0;    // developer switch: 1 = pad rows to whole 16-feature slabs (no half slab)
static int g_wl_pw = 1;        // developer switch: 0 = no loading warp anywhere
static int g_wl_groups = 0;    // developer switch: group count for d = 16, k <= 16 (4, 5 or 6)

// ---- column range pass: per column the largest magnitude (its exponent) and the mean exponent of the nonzero values; one read of
// the table, run in the STORE pass of a fit.  Thread = (row of a group, feature pair); integer statistics only, so the result does
// not depend on the order in which CTAs finish.
__global__ void __launch_bounds__(256) column_range_kernel(const double* __restrict__ X, int64_t n, int d, RangeRec* rec) {
  __shared__ unsigned int s_ticket;
  __shared__ int s_ok;
  const int tid = threadIdx.x, hp = d / 2, rpb = blockDim.x / hp;          // blockDim.x is a multiple of d / 2
  const int pf = tid % hp, r = tid / hp;
  unsigned int mx0 = 0, mx1 = 0, es0 = 0, es1 = 0, nz0 = 0, nz1 = 0;
  unsigned long long ES0 = 0, ES1 = 0, NZ0 = 0, NZ1 = 0;
  const int64_t step = (int64_t)gridDim.x * rpb;
  int flush = 0;
  for (int64_t row = (int64_t)blockIdx.x * rpb + r; row < n; row += 4 * step) {
    double2 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t rr = row + u * step;
      v[u] = rr < n ? __ldcs(reinterpret_cast<const double2*>(X + (size_t)rr * d) + pf) : make_double2(0.0, 0.0);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const unsigned int h0 = (unsigned int)__double2hiint(v[u].x) & 0x7fffffffu, h1 = (unsigned int)__double2hiint(v[u].y) & 0x7fffffffu;
      mx0 = max(mx0, h0); mx1 = max(mx1, h1);
      const unsigned int e0 = h0 >> 20, e1 = h1 >> 20;
      es0 += e0; es1 += e1; nz0 += e0 != 0u; nz1 += e1 != 0u;
    }
    if (++flush == 256) { ES0 += es0; ES1 += es1; NZ0 += nz0; NZ1 += nz1; es0 = es1 = nz0 = nz1 = 0u; flush = 0; }   // 32-bit partials: 4 * 256 * 2047 fits
  }
  ES0 += es0; ES1 += es1; NZ0 += nz0; NZ1 += nz1;
  // combine the CTA's rows per feature pair in shared memory, then one set of integer atomics per CTA and feature
  __shared__ unsigned int s_mx[2][256];
  __shared__ unsigned long long s_es[2][256], s_nz[2][256];
  s_mx[0][tid] = mx0; s_mx[1][tid] = mx1; s_es[0][tid] = ES0; s_es[1][tid] = ES1; s_nz[0][tid] = NZ0; s_nz[1][tid] = NZ1;
  __syncthreads();
  if (tid < d) {
    const int pp = tid >> 1, e = tid & 1;
    unsigned int m = 0; unsigned long long es = 0, nz = 0;
    for (int q = 0; q < rpb; ++q) { m = max(m, s_mx[e][q * hp + pp]); es += s_es[e][q * hp + pp]; nz += s_nz[e][q * hp + pp]; }
    atomicMax(&rec->emax_hi[tid], m); atomicAdd(&rec->esum[tid], es); atomicAdd(&rec->nnz[tid], nz);
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) s_ticket = atomicAdd(&rec->ticket, 1u);
  __syncthreads();
  if (s_ticket != gridDim.x - 1) return;
  __threadfence();
  if (tid == 0) s_ok = 1;
  __syncthreads();
  if (tid < d) {
    const unsigned int e = __ldcg(&rec->emax_hi[tid]) >> 20;                 // biased exponent of the largest magnitude
    const unsigned long long es = __ldcg(&rec->esum[tid]), nz = __ldcg(&rec->nnz[tid]);
    bool ok = e + 2u < 0x7ffu;                                               // finite, and room for M_f
    if (ok && nz > 0) ok = (unsigned long long)e * nz <= es + 20ull * nz;    // e - mean exponent <= 20
    if (!ok) s_ok = 0;
  }
  __syncthreads();
  if (tid == 0) {
    rec->tag_x = (unsigned long long)(uintptr_t)X; rec->n = (long long)n; rec->d = d; rec->done = 0; rec->ticket = 0u;
    __threadfence();
    rec->ok = s_ok;
  }
}

size_t wl_done_offset() { return offsetof(RangeRec, done); }
size_t wl_tail_bytes() { return 2048; }

// Shapes and passes this variant takes (given a valid record): plain Lloyd passes with sums, 8 < k <= 64, even d <= 64
bool lloyd_wl_shape(const LloydParams& p, int mode) {
  const int variant = lloyd_variant_override();
  if (g_wl_off || (variant != 0 && variant != 7)) return false;
  if (mode != 0 || p.w != nullptr || p.out_mode != 0 || p.mind2_out != nullptr || p.do_reduce != 1) return false;
  if (p.k <= 8 || p.k > 64 || p.d > 64 || p.d < 2 || (p.d & 1)) return false;          // even d: rows stay 16-byte aligned (tensor map, fixed point in place)
  if (p.n >= (int64_t)1 << 31) return false;                                   // tensor-map row coordinates are 32-bit
  if (p.n / kTileRows / sm_count() > 60000) return false;                      // s32 limb columns: 65 000 tiles per group
  const int kp = ((p.k + 15) / 16) * 16;
  // measured against the warp-group variant on the same box (tools/wl_check.py bench): with k <= 16 and rows of 18..24 features
  // the padding to two slabs costs more than the cheaper reduction saves
  if (variant == 0 && kp == 16 && p.d > 16 && p.d <= 24 && g_wl_nohalf) return false;
  if (variant == 0 && kp == 16 && p.d > 32) return false;                     // one group of four warps: the fp64 tensor-pipe kernel is ahead there
  return wl_groups(16 * ((p.d + 15) / 16), kp) > 0;
}

// Distance-returning passes the same kernel takes without the limb sums (k-means|| cost update and candidate weights: running
// minimum of the squared distance, labels, cost sum) and labels-only passes (transform): 8 < k <= 64 centres per launch, even d <= 48
bool lloyd_wd_shape(const LloydParams& p, int mode) {
  const int variant = lloyd_variant_override();
  if (g_wl_off || (variant != 0 && variant != 7)) return false;
  if (mode != 0 || p.w != nullptr || (p.do_reduce != 0 && p.do_reduce != 2)) return false;
  // with distances: any output mode; labels only (transform / predict): a plain pass that writes every row's label
  // (k <= 16: the warp-group kernel's labels-only pass is already at the HBM rate -- 1.01 against 0.96 here)
  if (p.mind2_out == nullptr && (p.labels_out == nullptr || p.do_reduce != 0 || p.out_mode != 0 || p.k <= 16)) return false;
  if (p.k <= 8 || p.k > 64 || p.d > 48 || p.d < 2 || (p.d & 1)) return false;
  if (p.n >= (int64_t)1 << 31) return false;
  const int kp = ((p.k + 15) / 16) * 16;
  return wl_groups_dist(16 * ((p.d + 15) / 16), kp) > 0;
}

int launch_column_range(const double* X, int64_t n, int d, void* tail, cudaStream_t st) {
  static_assert(sizeof(RangeRec) <= 2048, "range record");
  RangeRec* rec = reinterpret_cast<RangeRec*>(tail);
  HKM_CUDA_TRY(cudaMemsetAsync(rec, 0, sizeof(RangeRec), st));
  const int tpb = (256 / (d / 2)) * (d / 2);
  column_range_kernel<<<4 * sm_count(), tpb, 0, st>>>(X, n, d, rec);
  HKM_CUDA_TRY(cudaGetLastError());
  return HKM_OK;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess && qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  }
  return fn;
}

template <int KP, int NS, int GO = 0, int HS = 0, int PW = 0>
static int launch_wl(const LloydParams& p, const CUtensorMap& tmap, const CUtensorMap& tmap_h, RangeRec* rec, cudaStream_t st) {
  constexpr int DPAD = 16 * NS + 8 * HS;
  constexpr int H = (GO > 0 || HS) ? 1 : wl_halves(DPAD, KP);
  constexpr int G = GO > 0 ? GO : (HS ? wl_groups1(DPAD, KP) : wl_groups(DPAD, KP));
  if constexpr (G == 0) { set_error("lloyd_wl: d=%d k=%d does not fit", p.d, p.k); return HKM_ERR_UNSUPPORTED; }
  else {
    constexpr WlPlan L = wl_plan(DPAD, KP, G, wl_stages(DPAD, KP, G, H), H);
    auto kern = lloyd_wl_kernel<KP, NS, G, H, 0, HS, PW>;
    static bool attr_done[64];
    int dev = 0; HKM_CUDA_TRY(cudaGetDevice(&dev));
    if (dev >= 64 || !attr_done[dev]) {
      HKM_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L.total));
      if (dev < 64) attr_done[dev] = true;
    }
    kern<<<sm_count(), G * 128 * H + 32 * PW, L.total, st>>>(p, tmap, tmap_h, rec, g_wl_prof, g_wl_flags >= 0 ? g_wl_flags : (((NS == 2 && KP == 16) || (NS == 1 && KP == 32)) ? 1 : 0));
    HKM_CUDA_TRY(cudaGetLastError());
    return HKM_OK;
  }
}

static int make_tmap(const LloydParams& p, CUtensorMap& tmap, int box_cols = 16) {
  EncodeTiledFn enc = encode_tiled();
  if (enc == nullptr) { set_error("lloyd_wl: cuTensorMapEncodeTiled is not available"); return HKM_ERR_UNSUPPORTED; }
  const cuuint64_t gdim[2] = {(cuuint64_t)p.d, (cuuint64_t)p.n};
  const cuuint64_t gstride[1] = {(cuuint64_t)p.d * 8};
  const cuuint32_t box[2] = {(cuuint32_t)box_cols, 32}, estr[2] = {1, 1};
  const CUresult cr = enc(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<double*>(p.X), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          box_cols == 16 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (cr != CUDA_SUCCESS) { set_error("lloyd_wl: cuTensorMapEncodeTiled failed (%d)", (int)cr); return HKM_ERR_CUDA; }
  return HKM_OK;
}

template <int KP, int NS, int HS = 0>
static int launch_wd(const LloydParams& p, const CUtensorMap& tmap, const CUtensorMap& tmap_h, cudaStream_t st) {
  constexpr int DPAD = 16 * NS + 8 * HS;
  constexpr int G = wl_groups_dist(DPAD, KP);
  static_assert(HS == 0 || G > 0, "half-slab shapes must fit wherever the padded shape does");
  if constexpr (G == 0) { set_error("lloyd_wd: d=%d k=%d does not fit", p.d, p.k); return HKM_ERR_UNSUPPORTED; }
  else {
    constexpr WlPlan L = wl_plan(DPAD, KP, G, wl_stages(DPAD, KP, G, 1, 1), 1, 1);
    auto kern = lloyd_wl_kernel<KP, NS, G, 1, 1, HS>;
    static bool attr_done[64];
    int dev = 0; HKM_CUDA_TRY(cudaGetDevice(&dev));
    if (dev >= 64 || !attr_done[dev]) {
      HKM_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L.total));
      if (dev < 64) attr_done[dev] = true;
    }
    kern<<<sm_count(), G * 128, L.total, st>>>(p, tmap, tmap_h, nullptr, g_wl_prof, 0);
    HKM_CUDA_TRY(cudaGetLastError());
    return HKM_OK;
  }
}

int launch_lloyd_wd(const LloydParams& p, cudaStream_t st) {
  CUtensorMap tmap;
  const int rc = make_tmap(p, tmap);
  if (rc != HKM_OK) return rc;
  const int kp = ((p.k + 15) / 16) * 16, ns = (p.d + 15) / 16;
  // rows of 16 ns + 1 .. 16 ns + 8 features: ns slabs and a half slab, as in the reducing variant
  const int r16 = p.d % 16;
  if (r16 >= 1 && r16 <= 8 && p.d > 16 && p.d <= 40 && !g_wl_nohalf) {
    CUtensorMap tmap_h;
    { const int rc2 = make_tmap(p, tmap_h, 8); if (rc2 != HKM_OK) return rc2; }
    const int nsf = p.d / 16;
#define HKM_WDH_CASE(KPV, NSV) if (kp == KPV && nsf == NSV) return launch_wd<KPV, NSV, 1>(p, tmap, tmap_h, st);
    HKM_WDH_CASE(16, 1) HKM_WDH_CASE(32, 1) HKM_WDH_CASE(48, 1) HKM_WDH_CASE(64, 1)
    HKM_WDH_CASE(16, 2) HKM_WDH_CASE(32, 2) HKM_WDH_CASE(48, 2) HKM_WDH_CASE(64, 2)
#undef HKM_WDH_CASE
  }
#define HKM_WD_CASE(KPV, NSV) if (kp == KPV && ns == NSV) return launch_wd<KPV, NSV>(p, tmap, tmap, st);
  HKM_WD_CASE(16, 1) HKM_WD_CASE(32, 1) HKM_WD_CASE(48, 1) HKM_WD_CASE(64, 1)
  HKM_WD_CASE(16, 2) HKM_WD_CASE(32, 2) HKM_WD_CASE(48, 2) HKM_WD_CASE(64, 2)
  HKM_WD_CASE(16, 3) HKM_WD_CASE(32, 3) HKM_WD_CASE(48, 3) HKM_WD_CASE(64, 3)
#undef HKM_WD_CASE
  set_error("lloyd_wd: no instantiation for d=%d k=%d", p.d, p.k);
  return HKM_ERR_UNSUPPORTED;
}

int launch_lloyd_wl(const LloydParams& p, void* tail, cudaStream_t st) {
  CUtensorMap tmap;
  { const int rc = make_tmap(p, tmap); if (rc != HKM_OK) return rc; }
  RangeRec* rec = reinterpret_cast<RangeRec*>(tail);
  const int kp = ((p.k + 15) / 16) * 16;
  // rows of 16 ns + 1 .. 16 ns + 8 features (ns = 1, 2) take ns slabs of 16 and a half slab of 8 instead of padding to ns + 1 slabs
  const int r16 = p.d % 16;
  if (r16 >= 1 && r16 <= 8 && p.d > 16 && p.d <= 40 && !g_wl_nohalf) {
    CUtensorMap tmap_h;
    { const int rc = make_tmap(p, tmap_h, 8); if (rc != HKM_OK) return rc; }
    const int nsf = p.d / 16;
#define HKM_WLH_CASE(KPV, NSV) if (kp == KPV && nsf == NSV) return launch_wl<KPV, NSV, 0, 1>(p, tmap, tmap_h, rec, st);
    if (kp == 16 && nsf == 1) return launch_wl<16, 1, 5, 1>(p, tmap, tmap_h, rec, st);     // five groups: 96 registers (six spill at 80)
    HKM_WLH_CASE(32, 1) HKM_WLH_CASE(48, 1) HKM_WLH_CASE(64, 1)
    HKM_WLH_CASE(16, 2) HKM_WLH_CASE(32, 2) HKM_WLH_CASE(48, 2) HKM_WLH_CASE(64, 2)
#undef HKM_WLH_CASE
  }
  const int ns = (p.d + 15) / 16;
#define HKM_WL_CASE(KPV, NSV) if (kp == KPV && ns == NSV) return launch_wl<KPV, NSV>(p, tmap, tmap, rec, st);
  if (kp == 16 && ns == 1 && g_wl_groups == 4) return launch_wl<16, 1, 4>(p, tmap, tmap, rec, st);
  if (kp == 16 && ns == 1 && g_wl_groups == 5) return launch_wl<16, 1, 5>(p, tmap, tmap, rec, st);
#define HKM_WLP_CASE(KPV, NSV) if (kp == KPV && ns == NSV) return launch_wl<KPV, NSV, 0, 0, 1>(p, tmap, tmap, rec, st);
  // the loading warp where it measured faster: 64-feature rows (0.657 -> 0.677 at k = 64, 0.68 -> 0.69 at k = 32) and d = 32 with k <= 16
  // (0.84 -> 0.91).  With many groups one warp cannot issue the boxes fast enough (d = 16, k <= 16: 0.86 -> 0.61; d = 48: 0.73 -> 0.68).
  if (g_wl_pw >= 1) { HKM_WLP_CASE(16, 4) HKM_WLP_CASE(32, 4) HKM_WLP_CASE(48, 4) HKM_WLP_CASE(64, 4) HKM_WLP_CASE(16, 2) }
#undef HKM_WLP_CASE
  HKM_WL_CASE(16, 1) HKM_WL_CASE(32, 1) HKM_WL_CASE(48, 1) HKM_WL_CASE(64, 1)
  HKM_WL_CASE(16, 2) HKM_WL_CASE(32, 2) HKM_WL_CASE(48, 2) HKM_WL_CASE(64, 2)
  HKM_WL_CASE(16, 3) HKM_WL_CASE(32, 3) HKM_WL_CASE(48, 3) HKM_WL_CASE(64, 3)
  HKM_WL_CASE(16, 4) HKM_WL_CASE(32, 4) HKM_WL_CASE(48, 4) HKM_WL_CASE(64, 4)
#undef HKM_WL_CASE
  set_error("lloyd_wl: no instantiation for d=%d k=%d", p.d, p.k);
  return HKM_ERR_UNSUPPORTED;
}

}  // namespace hkm

extern "C" __attribute__((visibility("default"))) void hkm_debug_wl_profile(void* buf) { hkm::g_wl_prof = reinterpret_cast<unsigned long long*>(buf); }
extern "C" __attribute__((visibility("default"))) void hkm_debug_wl_off(int off) { hkm::g_wl_off = off; }
extern "C" __attribute__((visibility("default"))) void hkm_debug_wl_pw(int f) { hkm::g_wl_pw = f; }
extern "C" __attribute__((visibility("default"))) void hkm_debug_wl_groups(int g) { hkm::g_wl_groups = g; }
extern "C" __attribute__((visibility("default"))) void hkm_debug_wl_flags(int f) { hkm::g_wl_flags = f; }
extern "C" __attribute__((visibility("default"))) void hkm_debug_wl_nohalf(int f) { hkm::g_wl_nohalf = f; }

// Lloyd data pass, warp-group variant with the per-cluster sums on the integer tensor cores (8 < k <= 64, d a multiple of 16,
// later passes of a fit: xnorm2 mode REUSE).  Screening and refinement are those of hkm_lloyd_wg.cuh; what changes is how a tile's
// rows reach the per-cluster sums, which in the other variants costs k d / 64 fp64 tensor-pipe clocks per row (one-hot DMMA) or
// 24 d bytes of shared-memory traffic per row (read-modify-write):
//
//   * a row x is turned, IN PLACE in its ring stage, into fixed point: y = x + M_f with M_f = 1.5 * 2^(E_f + 2), E_f the exponent
//     of the column's largest magnitude (measured once per fit, see hkm_range.cu).  y stays in one binade, so its 64 bits are the
//     bits of M_f plus round(x / 2^(E_f - 50)) as an integer -- one DADD per value, and the eight bytes of y are eight base-256
//     digits ("limbs") of that integer.
//   * the stage is loaded by tensor-map TMA with the 128-byte swizzle, so after the conversion it is, as it lies, the canonical
//     MN-major SWIZZLE_128B operand of tcgen05.mma kind::i8: A[m = byte of the row (128 per 16 features)][K = row].
//   * the labels become one byte each in a K-major one-hot operand B[n = cluster][K = row] (one STS.U8 per row, cleared after use).
//   * D[byte][cluster] += A B^T in s32, accumulator in TMEM for the whole pass: four tcgen05.mma per 128 rows and 16 features.
//     Integer sums are exact and order independent; a limb column grows by at most 128 * 255 per tile, so s32 holds 65 000 tiles
//     of one group (the host refuses more rows per CTA than that).
//   * at the end S[j][f] = (sum_l D[8f + l][j] 256^l - n_j bits(M_f)) * 2^(E_f - 50), in 128-bit integers, rounded once.
//
// Error: each value is rounded once to a multiple of 2^(E_f - 50) (|error| <= 2^(E_f - 51), unbiased), the sums are then exact; the
// range pass only admits columns whose largest exponent is at most 20 above the mean exponent of their nonzero values (a measure
// one outlier cannot move), so the quantum is below 2^-30 of a typical value.  Columns outside that (or non-finite data) leave the record invalid and this kernel returns at once,
// leaving the pass to the variant launched behind it (LloydParams::skip2).
#pragma once
#include <cuda.h>
#include "hkm_lloyd_wg.cuh"

namespace hkm {

constexpr int kWlMaxGroups = 6;
constexpr int kWlMaxWarps = 4 * kWlMaxGroups;
constexpr int kWlMaxStages = 12;
constexpr int kRangeMaxD = 64;

// Column range record of one table (device memory, tail of the Lloyd scratch): written by column_range_kernel in the STORE pass
struct RangeRec {
  unsigned long long tag_x;            // address of the table the record describes
  long long n;
  int d;
  int ok;                              // 1: every column is finite and within the dynamic range the fixed-point sums accept
  int done;                            // set by lloyd_wl_kernel when it has produced the pass (the fallback launch then returns)
  unsigned int ticket;
  unsigned int emax_hi[kRangeMaxD];    // max over rows of the high word of |x_f|
  unsigned long long esum[kRangeMaxD]; // sum of the biased exponents of the column's nonzero values
  unsigned long long nnz[kRangeMaxD];  // how many there are
};

struct WlPlan {
  int groups, stages, kchunks, lbo_b, lbo_h, h_bytes, tile_bytes, tcols, acols, tstride, gstride, rcol0;
  int x_off, b_off, h_off, ex_off, ct_off, ct_pitch, cn_off, mu_off, mg_off, cnt_off, red_off, bar_off, seq_off, rel_off, tmem_off, cs_off, total;
};

// Tensor memory, per group: [tf32 operand: d data columns + 8 constant columns][scores: k_pad], then behind all groups the s32 limb
// sums [slab][k_pad], one set for the whole CTA (every warp adds its products into it).  Shared memory: ring stages (1024-byte aligned), the centres' tf32 operand, the one-hot operands.
// halves = warps per 32 rows (1, or 2: the features and the score columns of a row split over two threads)
// dist = 1: the distance-returning variant (no limb sums: no one-hot operands, no limb columns; a padded fp64 centre table instead)
__host__ __device__ constexpr inline WlPlan wl_plan(int d, int kp, int groups, int stages, int halves = 1, int dist = 0) {
  WlPlan L{};
  L.groups = groups; L.stages = stages;
  const int ns = d / 16, hs = (d % 16) ? 1 : 0;     // full 16-feature slabs, and a trailing 8-feature half slab (d = 16 ns + 8)
  L.kchunks = (d + 8) / 4;
  L.lbo_b = (kp / 8) * 128 + 16;
  L.lbo_h = (kp / 8) * 128 + 16;                   // one-hot operand: chunk of 16 rows, then cluster * 16, then row % 16
  L.h_bytes = (8 * L.lbo_h + 127) & ~127;
  L.tile_bytes = ns * kTileRows * 128 + hs * kTileRows * 64;
  L.acols = (d + 8 + 15) & ~15;
  L.tstride = kp <= 32 ? 32 : 64;
  L.gstride = L.acols + L.tstride;
  L.rcol0 = groups * L.gstride;
  int tc = 32; while (tc < L.rcol0 + (dist ? 0 : (ns + hs) * kp)) tc *= 2;
  L.tcols = tc;
  int o = 0;
  L.x_off = o; o += stages * L.tile_bytes;         // 1024-byte aligned (swizzle atoms)
  L.b_off = o; o += (L.kchunks * L.lbo_b + 127) & ~127;
  L.h_off = o; o += dist ? 0 : groups * L.h_bytes;
  L.ct_pitch = d * 8 + 16;                         // centre rows 16 bytes apart modulo 128: gathers by label spread over the banks
  L.ct_off = o; o += dist ? kp * L.ct_pitch : 0;
  L.ex_off = o; o += halves > 1 ? groups * kTileRows * 2 * 16 : 0;     // exchange slots of the two threads of a row
  L.cn_off = o; o += kp * 8;
  L.mu_off = o; o += (d + 8) * 8;
  L.mg_off = o; o += d * 8;                        // M_f per feature
  L.cnt_off = o; o += kp * 4;
  L.red_off = o; o += (2 * 4 * groups * halves + 64) * 8;
  const int nst = stages > 0 ? stages : 1;
  L.bar_off = o; o += (4 * halves * nst + kWlMaxGroups + kWlMaxWarps) * 8;     // full[stage][quarter][half], scores[group], sums[group][quarter]
  L.seq_off = o; o += 4 * halves * nst * 4; o = (o + 15) & ~15;
  L.rel_off = o; o += kWlMaxWarps * 4;                // tiles closed per warp (read by the loading warp)
  L.tmem_off = o; o += 16;
  if (kp * d * 8 <= 8 * 1024) { L.cs_off = o; o += kp * d * 8; }
  L.total = o;
  return L;
}

__host__ __device__ constexpr inline int wl_stages(int d, int kp, int g, int halves = 1, int dist = 0) {
  if (g < 1 || g * halves > kWlMaxGroups) return 0;
  const WlPlan z = wl_plan(d, kp, g, 0, halves, dist);
  if (z.tcols > 512) return 0;
  const int budget = 227 * 1024 - 1024;             // the static words are padded to 1 KB (alignment of the dynamic part)
  int st = kWlMaxStages;
  while (st > 0 && wl_plan(d, kp, g, st, halves, dist).total > budget) --st;
  const int comb = dist ? 0 : ((d + 15) / 16) * kp * 128 * 4;   // s32 accumulators parked in the idle ring for the final combine
  if (st < g + 2 || (st - g) * z.tile_bytes < 40 * 1024 || st * z.tile_bytes < comb) return 0;
  if (g == 1 && halves == 1 && d <= 32) return 0;   // a single group only where nothing else fits
  return st;
}
// group count per shape (0: shape not taken), and threads per row: where only one group of four warps fits (64 KB tiles), the
// same tile is worked by eight warps instead, two threads per row
__host__ __device__ constexpr inline int wl_groups1(int d, int kp) {
  for (int g = 6; g >= 1; --g) if (wl_stages(d, kp, g, 1) > 0) return g;
  return 0;
}
__host__ __device__ constexpr inline int wl_halves(int d, int kp) { return (wl_groups1(d, kp) == 1 && wl_stages(d, kp, 1, 2) > 0) ? 2 : 1; }
__host__ __device__ constexpr inline int wl_groups(int d, int kp) { return wl_halves(d, kp) == 2 ? 1 : wl_groups1(d, kp); }
// the distance-returning variant: one thread per row only
__host__ __device__ constexpr inline int wl_groups_dist(int d, int kp) {
  for (int g = 6; g >= 2; --g) if (wl_stages(d, kp, g, 1, 1) > 0) return g;
  return 0;
}

__device__ __forceinline__ void sts_u8(uint32_t a, uint32_t v) { asm volatile("st.shared.u8 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ uint64_t wl_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout_type) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32) |
         ((uint64_t)1 << 46) | ((uint64_t)layout_type << 61);
}

__device__ __forceinline__ void wl_tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr), "r"(v[0]), "r"(v[1]),
               "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]),
               "r"(v[15])
               : "memory");
}
__device__ __forceinline__ void wl_tmem_st8(uint32_t taddr, const uint32_t (&v)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]),
               "r"(v[5]), "r"(v[6]), "r"(v[7])
               : "memory");
}
__device__ __forceinline__ void wl_tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// KP = padded cluster count (16..64); NS = ceil(d / 16) (slabs of 128 bytes per row; d even); GFIX = groups of four warps.
// Thread = row throughout: a thread reads its row through the swizzle (eight consecutive rows hit eight different 16-byte columns),
// writes the tf32 values into its lane of the group's tensor-memory operand (the tensor core takes A from TMEM: no shared-memory
// copy of the operand at all), keeps |x - mu|^2 in a register, and overwrites the row in place with its fixed-point form.
template <int KP, int NS, int GFIX, int H = 1, int DIST = 0, int HS = 0, int PW = 0>
__global__ void __launch_bounds__(GFIX * 128 * H + 32 * PW, 1) lloyd_wl_kernel(const LloydParams p, const __grid_constant__ CUtensorMap tmap,
                                                                 const __grid_constant__ CUtensorMap tmap_h, RangeRec* rec,
                                                                 unsigned long long* prof_buf, int flags) {
  if (p.skip != nullptr && *p.skip != 0) return;
  // DP: features padded to whole slabs -- NS slabs of 16 and, with HS, a trailing half slab of 8 (64-byte rows, SWIZZLE_64B boxes from
  // the second tensor map; its limb product is an M = 64 instruction whose result row m sits in TMEM lane m % 16 + 32 (m / 16))
  constexpr int DP = 16 * NS + 8 * HS, ND = DP / 8, G = GFIX, S = wl_stages(DP, KP, GFIX, H, DIST);
  static_assert(HS == 0 || (H == 1 && NS <= 2), "half slab: one thread per row, d <= 40");
  constexpr uint32_t HSLAB_OFF = NS * kTileRows * 128u;   // the half slab sits behind the full slabs of a stage
  constexpr WlPlan L = wl_plan(DP, KP, GFIX, S, H, DIST);
  static_assert(H == 1 || H == 2, "one or two threads per row");
  static_assert(DIST == 0 || H == 1, "the distance-returning variant keeps a row in one thread");
  const int d = p.d;                                  // even, DP - 16 < d <= DP; the padding columns arrive as zeros (tensor-map bounds)
  static_assert(S >= GFIX + 2, "shape does not fit");
  constexpr int NW = (KP + 31) / 32;
  constexpr int IDXBITS = 6;
  constexpr uint32_t IDXMASK = (1u << IDXBITS) - 1u;
  // PW = 1: one more warp that issues every refill of the ring.  A bulk tensor copy holds the issuing thread for 40-70 clocks per
  // 4 KB box while the copy engine is busy; with the sixteen boxes of a 64-feature tile requested at the same moment, the computing
  // warps lost 500-1000 clocks per tile there.
  static_assert(PW == 0 || (DIST == 0 && HS == 0), "the loading warp is for the reducing variant with whole slabs");
  constexpr int nthreads = G * 128 * H + 32 * PW, nwarps = G * 4 * H;
  constexpr int NCH = KP / 16;                          // score columns in chunks of 16
  constexpr bool HOIST = nwarps <= 8;
  constexpr bool XPROF = nwarps <= 8;                  // the finer clock counters of the profile hook: only where registers are plentiful
  constexpr uint32_t LBO_B = L.lbo_b, LBO_H = L.lbo_h, tile_bytes = L.tile_bytes;
  constexpr uint32_t SLAB = kTileRows * 128;          // bytes of one 16-feature slab of a tile
  // the record must describe this table and admit it; otherwise the launch behind this one does the pass
  if constexpr (DIST == 0) {
    if (rec->ok != 1 || rec->tag_x != (unsigned long long)(uintptr_t)p.X || rec->n != (long long)p.n || rec->d != p.d) {
      if (blockIdx.x == 0 && threadIdx.x == 0) rec->done = 0;
      return;
    }
  }

  extern __shared__ __align__(1024) unsigned char smem_raw[];
  // the swizzle is a function of the shared address: stages must start on 1024-byte boundaries (the declared alignment gives that)
  unsigned char* smem = smem_raw;
  if ((smem_u32(smem_raw) & 1023u) != 0u) __trap();
  const int k = p.k;
  const uint32_t sb = smem_u32(smem);
  const uint32_t a_x = sb + L.x_off, a_B = sb + L.b_off, a_cnt = sb + L.cnt_off;
  const uint32_t a_bar = sb + L.bar_off, a_seq = sb + L.seq_off;
  const uint32_t b_full = a_bar, b_tfull = a_bar + 8u * 4u * H * S, b_rfull = b_tfull + 8u * kWlMaxGroups;
  double* cn_s = reinterpret_cast<double*>(smem + L.cn_off);
  double* mu_s = reinterpret_cast<double*>(smem + L.mu_off);
  double* mg_s = reinterpret_cast<double*>(smem + L.mg_off);
  double* red_s = reinterpret_cast<double*>(smem + L.red_off);
  unsigned int* cnt_s = reinterpret_cast<unsigned int*>(smem + L.cnt_off);

  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  const int grp = warp / (4 * H), q = warp & 3, hh = H == 1 ? 0 : (warp >> 2) & 1;   // group, lane quarter of the tile, half of the row
  // this thread's share of its row: slabs [sl_lo, sl_hi) and score chunks [ch_lo, ch_hi)
  const int sl_lo = hh ? (NS + 1) / 2 : 0, sl_hi = (H == 1 || hh) ? NS : (NS + 1) / 2;
  const int ch_lo = hh ? (NCH + 1) / 2 : 0, ch_hi = (H == 1 || hh) ? NCH : (NCH + 1) / 2;
  const int64_t full_tiles = p.n / kTileRows;
  const int tail_rows = (int)(p.n - full_tiles * kTileRows);
  const int my_tiles = full_tiles > (int64_t)blockIdx.x ? (int)((full_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;
  const bool tail_mine = tail_rows > 0 && (int)(full_tiles % gridDim.x) == (int)blockIdx.x;
  const int T = my_tiles + (tail_mine ? 1 : 0);    // the ragged tail is this CTA's last tile (rows past the end arrive as zeros)
  const bool prof = prof_buf != nullptr && blockIdx.x == 0;
  const bool compute = warp < nwarps;                 // the loading warp (PW) takes part in the set-up and in the combine only
  long long pw0 = 0, pw1 = 0, pw2 = 0, pcnt = 0, pw3 = 0, pw4 = 0, pw5 = 0, pw6 = 0, pw7 = 0, pw8 = 0;

  // ------------------------------------------------------------------ one-time set-up
  if (tid == 0) {
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bar_off);
    for (int s = 0; s < 4 * H * S; ++s) mbar_init(bars + s, 1);
    for (int g = 0; g < kWlMaxGroups + kWlMaxWarps; ++g) mbar_init(bars + 4 * H * S + g, 1);
    mbar_fence_init();
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap) : "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(sb + L.tmem_off), "n"(L.tcols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // zero: B and the one-hot operands (one contiguous range), counters
  for (int i = tid; i < ((DIST ? L.ct_off : L.cn_off) - L.b_off) / 16; i += nthreads) reinterpret_cast<uint4*>(smem + L.b_off)[i] = make_uint4(0, 0, 0, 0);
  for (int i = tid; i < KP; i += nthreads) cnt_s[i] = 0u;
  for (int i = tid; i < kWlMaxWarps; i += nthreads) reinterpret_cast<uint32_t*>(smem + L.rel_off)[i] = 0u;
  for (int i = tid; i < 4 * H * S; i += nthreads) reinterpret_cast<uint32_t*>(smem + L.seq_off)[i] = (uint32_t)(i / (4 * H));
  if (L.cs_off != 0) for (int i = tid; i < k * d; i += nthreads) reinterpret_cast<double*>(smem + L.cs_off)[i] = p.C[i];
  if constexpr (DIST != 0) {
    // fp64 centres for the exact distance of every row's winner: row j at j * ct_pitch, padding columns zero
    for (int i = tid; i < KP * DP; i += nthreads) {
      const int j = i / DP, f = i - j * DP;
      *reinterpret_cast<double*>(smem + L.ct_off + j * L.ct_pitch + f * 8) = (j < k && f < d) ? p.C[(size_t)j * d + f] : 0.0;
    }
  }
  if (tid < DP + 8) { double s = 0; if (tid < d) for (int j = 0; j < k; ++j) s += p.C[(size_t)j * d + tid]; mu_s[tid] = tid < d ? s / k : 0.0; }
  if (tid < DP) {
    // M_f = 1.5 * 2^(E_f + 2): exponent field of the column's largest magnitude + 2, top mantissa bit set (padding columns: 1.5)
    const unsigned eb = (DIST == 0 && tid < d) ? (rec->emax_hi[tid] >> 20) & 0x7ffu : 1021u;
    mg_s[tid] = __longlong_as_double(((long long)(eb + 2u) << 52) | (1ll << 51));
  }
  __syncthreads();
  double* h_s = red_s + 2 * nwarps;
  if (tid < KP) {
    double s = 0, sc = 0;
    if (tid < k) for (int f = 0; f < d; ++f) { const double c = p.C[(size_t)tid * d + f], cc = c - mu_s[f]; s = fma(c, c, s); sc = fma(cc, cc, sc); }
    cn_s[tid] = tid < k ? -0.5 * s : -__longlong_as_double(0x7ff0000000000000LL);
    h_s[tid] = tid < k ? sc : 0.0;
  }
  for (int i = tid; i < KP * (DP / 4); i += nthreads) {
    const int j = i % KP, c = i / KP;
    uint32_t vv[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) { const int f = 4 * c + e; vv[e] = (j < k && f < d) ? f32_to_tf32_rn((float)(p.C[(size_t)j * d + f] - mu_s[f])) : 0u; }
    *reinterpret_cast<uint4*>(smem + L.b_off + c * LBO_B + (j >> 3) * 128 + (j & 7) * 16) = make_uint4(vv[0], vv[1], vv[2], vv[3]);
  }
  __syncthreads();
  if (tid < KP) {
    const double h = tid < k ? -0.5 * h_s[tid] : -1e30;
    const float hi = __uint_as_float(f32_to_tf32_rn((float)h));
    const double r1 = h - (double)hi;
    const float mid = __uint_as_float(f32_to_tf32_rn((float)r1));
    const double r2 = r1 - (double)mid;
    const float lo = __uint_as_float(f32_to_tf32_rn((float)r2));
    *reinterpret_cast<uint4*>(smem + L.b_off + (DP / 4) * LBO_B + (tid >> 3) * 128 + (tid & 7) * 16) =
        make_uint4(__float_as_uint(hi), __float_as_uint(mid), __float_as_uint(lo), 0u);
  }
  float vmax2 = 0.f;
  for (int j = 0; j < k; ++j) vmax2 = fmaxf(vmax2, (float)h_s[j]);
  const float vmax = sqrtf(vmax2) * 1.0001f;
  float mun2 = 0.f;
  for (int f = 0; f < d; ++f) mun2 = fmaf((float)mu_s[f], (float)mu_s[f], mun2);
  const float munorm = sqrtf(mun2) * 1.0001f;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *reinterpret_cast<const uint32_t*>(smem + L.tmem_off);
  const uint32_t tlane = ((uint32_t)(32 * q)) << 16;
  const uint32_t t_A = tmem + tlane + (uint32_t)grp * (uint32_t)L.gstride;       // this lane's row of the group's tf32 operand
  const uint32_t t_D = t_A + (uint32_t)L.acols;                                  // ... and of its score accumulator
  if (compute) {
    // the eight constant columns behind the data columns: (1, 1, 1, 0, ...) against the hi/mid/lo split of -0.5|c - mu|^2
    const uint32_t one = __float_as_uint(1.0f);
    const uint32_t cv[8] = {one, one, one, 0u, 0u, 0u, 0u, 0u};
    wl_tmem_st8(t_A + (uint32_t)DP, cv);
    // the limb sums start at zero (every product below accumulates; the four warps of a group issue theirs independently)
    const uint32_t zv[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
#pragma unroll 1
    if (DIST == 0 && grp == 0 && hh == 0) for (int c0 = 0; c0 < (NS + HS) * KP; c0 += 8) wl_tmem_st8(tmem + tlane + (uint32_t)L.rcol0 + (uint32_t)c0, zv);
    wl_tmem_wait_st();
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  }
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

  double cost = 0;
  const int P = k * d + k + 1;
  const int SL = P + 1 + k;
  double* out = p.cta_partials + (size_t)blockIdx.x * SL;
  const long long pt0 = prof ? clock64() : 0;

  // ------------------------------------------------------------------ per-warp constants
  // loads: this warp's quarter (rows 32q .. 32q+31) of tile i, one box of 16 features x 32 rows per slab; rows past the end of
  // the table arrive as zeros, so the ragged tail needs no other path
  auto tma_load = [&](int i, uint32_t xs_q, uint32_t bar, uint32_t seqa) {
    const int row0 = (int)(((int64_t)blockIdx.x + (int64_t)i * gridDim.x) * kTileRows) + 32 * q;
    // one elected lane, inside the asm: the C++ control flow stays convergent and the operands stay in uniform registers
    if constexpr (NS == 1 && HS == 0) {
      asm volatile(
          "{\n.reg .pred pe;\nelect.sync _|pe, 0xffffffff;\n"
          "@pe st.shared.u32 [%0], %1;\n"
          "@pe mbarrier.arrive.expect_tx.shared::cta.b64 _, [%2], 4096;\n"
          "@pe cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%3], [%4, {%5, %6}], [%2];\n}\n"
          ::"r"(seqa), "r"((uint32_t)i), "r"(bar), "r"(xs_q), "l"(&tmap), "r"(0), "r"(row0)
          : "memory");
    } else {
      asm volatile(
          "{\n.reg .pred pe;\nelect.sync _|pe, 0xffffffff;\n"
          "@pe st.shared.u32 [%0], %1;\n"
          "@pe mbarrier.arrive.expect_tx.shared::cta.b64 _, [%2], %3;\n}\n"
          ::"r"(seqa), "r"((uint32_t)i), "r"(bar), "r"((uint32_t)(sl_hi - sl_lo) * 4096u + (uint32_t)HS * 2048u)
          : "memory");
      if constexpr (HS != 0)
        asm volatile(
            "{\n.reg .pred pe;\nelect.sync _|pe, 0xffffffff;\n"
            "@pe cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];\n}\n"
            ::"r"(xs_q - (uint32_t)q * 4096u + HSLAB_OFF + (uint32_t)q * 2048u), "l"(&tmap_h), "r"(16 * NS), "r"(row0), "r"(bar)
            : "memory");
#pragma unroll
      for (int sl = 0; sl < NS; ++sl)
        if (sl >= sl_lo && sl < sl_hi)
        asm volatile(
            "{\n.reg .pred pe;\nelect.sync _|pe, 0xffffffff;\n"
            "@pe cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];\n}\n"
            ::"r"(xs_q + (uint32_t)sl * SLAB), "l"(&tmap), "r"(16 * sl), "r"(row0), "r"(bar)
            : "memory");
    }
  };
  if (compute) for (int i = grp; i < S && i < T; i += G)
    tma_load(i, a_x + (uint32_t)i * tile_bytes + (uint32_t)q * 4096u, b_full + 8u * (uint32_t)((4 * i + q) * H + hh), a_seq + 4u * (uint32_t)((4 * i + q) * H + hh));

  const int row = 32 * q + lane;                                                  // tile row of this thread = TMEM lane
  const uint32_t xrow = (uint32_t)lane * 128u, xsw = (uint32_t)(lane & 7) << 4;   // row offset inside the quarter, swizzle term
  uint32_t idxclr;
  asm volatile("mov.u32 %0, 0xffffffc0;" : "=r"(idxclr));
  const float kapv = 1.04f * 0.0009765625f * vmax;
  const float etab = 0.5f * 1.52587890625e-05f * vmax * vmax + 2.5e-7f * munorm * vmax;
  const int rlane = 2 * lane;
  const bool ract = rlane < d;
  constexpr int RLANES = tp_dclass(ND) / 2 > 32 ? 32 : tp_dclass(ND) / 2;
  const double* Cr = L.cs_off != 0 ? reinterpret_cast<const double*>(smem + L.cs_off) : p.C;
  const uint32_t a_H = sb + L.h_off + (uint32_t)grp * (uint32_t)L.h_bytes;
  const uint32_t bar_t = b_tfull + 8u * (uint32_t)grp, bar_r = b_rfull + 8u * (uint32_t)(4 * grp + q);
  const uint32_t a_ex = sb + L.ex_off + (uint32_t)grp * (kTileRows * 32u) + (uint32_t)row * 32u;   // this row's two exchange slots
  const int pair_id = 1 + G + 4 * grp + q;                          // named barrier of the two warps that share 32 rows
  const uint32_t hrow = a_H + (uint32_t)(row >> 4) * LBO_H + (uint32_t)(row & 15);       // + 16 * label: this row's one-hot byte
  const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(KP >> 3) << 17) | ((uint32_t)(kTileRows >> 4) << 24);
  const uint32_t idesc2 = (2u << 4) | (1u << 15) | ((uint32_t)(KP >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);   // u8 x u8 -> s32, A MN-major
  const uint64_t db0 = tp_desc_kmajor(a_B, LBO_B, 128u);
  const uint32_t db_step = (2u * LBO_B) >> 4;
  const uint32_t acol = tmem + (uint32_t)grp * (uint32_t)L.gstride, dcol = acol + (uint32_t)L.acols;
  const uint32_t rcol = tmem + (uint32_t)L.rcol0;
  const uint64_t dh0 = wl_desc(a_H, LBO_H, 128u, 0u);
  const uint32_t dh_step = (2u * LBO_H) >> 4;
  // mu as packed float pairs: in registers for one slab, re-read (broadcast) from shared memory for wider rows
  float2* mu2_s = reinterpret_cast<float2*>(smem + L.mu_off);                     // overlays mu_s once every thread has read it
  uint64_t mu_r[NS == 1 ? 8 : 1];
  {
    float2 t[NS == 1 ? 8 : 1];
    if (NS == 1) {
#pragma unroll
      for (int c = 0; c < 8; ++c) t[c] = make_float2((float)mu_s[2 * c], (float)mu_s[2 * c + 1]);
#pragma unroll
      for (int c = 0; c < 8; ++c) asm("mov.b64 %0, {%1,%2};" : "=l"(mu_r[c]) : "f"(t[c].x), "f"(t[c].y));
    } else mu_r[0] = 0;
    float2 mine = make_float2(0.f, 0.f);
    if (tid < DP / 2) mine = make_float2((float)mu_s[2 * tid], (float)mu_s[2 * tid + 1]);
    __syncthreads();
    if (tid < DP / 2) mu2_s[tid] = mine;                     // the table (used beyond the first slab) overlays mu_s
    __syncthreads();
  }
  const uint32_t a_mu2 = sb + L.mu_off, a_mg = sb + L.mg_off;

  // ------------------------------------------------------------------ tiles grp, grp + G, ...
  int s = grp; uint32_t ph = 0, tph = 0;
  uint32_t hb_prev = 0u, xs_prev = 0u, bar_f_prev = 0u, seqa_prev = 0u;
  bool pend = false;                                  // the previous tile's limb product has not been waited for yet
  const bool early = (flags & 1) != 0;
  // a closed tile's quarter of the ring is refilled with tile `nxt`: by this warp, or (PW) by the loading warp, which is told through
  // this warp's counter of closed tiles
  uint32_t nclosed = 0;
  const uint32_t a_rel = sb + L.rel_off;
  auto refill = [&](int nxt, uint32_t xs_q, uint32_t bar, uint32_t seqa) {
    if constexpr (PW != 0) {
      ++nclosed;
      if (lane == 0) asm volatile("st.release.cta.shared.u32 [%0], %1;" ::"r"(a_rel + 4u * (uint32_t)warp), "r"(nclosed) : "memory");
    } else {
      if (nxt < T) tma_load(nxt, xs_q, bar, seqa);
    }
  };
  if constexpr (PW != 0) {
    if (!compute) {
      // ---- the loading warp: lane w serves computing warp w.  That warp's m-th closed tile j = group + m G frees its quarter of
      // stage j % S, which takes tile j + S.  The counters only grow, so a lane that falls behind misses nothing.
      const int w = lane;
      const int g_w = w / (4 * H), q_w = w & 3, hh_w = (w >> 2) & (H - 1);
      const int lo_w = hh_w ? (NS + 1) / 2 : 0, hi_w = (H == 1 || hh_w) ? NS : (NS + 1) / 2;
      int m = 0;
      bool fin = w >= nwarps;
      while (true) {
        bool issued = false;
        if (!fin) {
          const int j = g_w + m * G;
          if (j + S >= T) fin = true;
          else {
            uint32_t have;
            asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(have) : "r"(a_rel + 4u * (uint32_t)w) : "memory");
            if (have > (uint32_t)m) {
              const int t = j + S, st = j % S;
              const int row0 = (int)(((int64_t)blockIdx.x + (int64_t)t * gridDim.x) * kTileRows) + 32 * q_w;
              const uint32_t xs_q = a_x + (uint32_t)st * tile_bytes + (uint32_t)q_w * 4096u;
              const uint32_t bar = b_full + 8u * (uint32_t)((4 * st + q_w) * H + hh_w), seqa = a_seq + 4u * (uint32_t)((4 * st + q_w) * H + hh_w);
              asm volatile("st.shared.u32 [%0], %1;\n"
                           "mbarrier.arrive.expect_tx.shared::cta.b64 _, [%2], %3;\n"
                           ::"r"(seqa), "r"((uint32_t)t), "r"(bar), "r"((uint32_t)(hi_w - lo_w) * 4096u) : "memory");
              for (int sl = lo_w; sl < hi_w; ++sl)
                asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                             ::"r"(xs_q + (uint32_t)sl * SLAB), "l"(&tmap), "r"(16 * sl), "r"(row0), "r"(bar) : "memory");
              ++m; issued = true;
            }
          }
        }
        if (__all_sync(0xffffffffu, fin)) break;
        if (!__any_sync(0xffffffffu, issued)) __nanosleep(32);
      }
    }
  }
  const int Tc = compute ? T : 0;                      // tiles this warp computes on
  for (int i = grp; i < Tc; i += G) {
    const int rows_valid = i < my_tiles ? kTileRows : tail_rows;
    const long long tt0 = (XPROF && prof) ? clock64() : 0;
    const uint32_t xs = a_x + (uint32_t)s * tile_bytes + (uint32_t)q * 4096u;       // this warp's 32 rows of slab 0
    const uint32_t bar_f = b_full + 8u * (uint32_t)((4 * s + q) * H + hh), seqa = a_seq + 4u * (uint32_t)((4 * s + q) * H + hh);
    if (pend && early && mbar_probe_a(bar_r, tph ^ 1u)) {
      // the previous tile's limb product is already complete: close it now, so its quarter of the ring refills at once
      // (otherwise after the next product has been issued, which hides the wait)
      if (hb_prev != 0u) sts_u8(hb_prev, 0u);
      __syncwarp();
      refill(i - G + S, xs_prev, bar_f_prev, seqa_prev);
      pend = false;
    }
    {
      // the quarter must have been requested for THIS tile before its parity means anything (hkm_lloyd_wg.cuh)
      uint32_t have;
      asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(have) : "r"(seqa) : "memory");
      while (have != (uint32_t)i) {
        __nanosleep(64);
        asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(have) : "r"(seqa) : "memory");
      }
    }
    if (XPROF && prof) pw6 += clock64() - tt0;
    tp_wait_p(bar_f, ph, 3, prof, pw0);
    const long long tc0 = (XPROF && prof) ? clock64() : 0;
    // ---- this thread's row: tf32 operand -> TMEM, |x - mu|^2, fixed point in place
    float xa = 0.f, xb = 0.f;
#pragma unroll
    for (int sl = 0; sl < NS; ++sl) {
      if (H == 2 && (sl < sl_lo || sl >= sl_hi)) continue;       // the other thread of the row converts these
      const uint32_t xr = xs + (uint32_t)sl * SLAB + xrow;
      double v[16];
#pragma unroll
      for (int c = 0; c < 8; ++c) lds_v2f64(xr + (((uint32_t)c << 4) ^ xsw), v[2 * c], v[2 * c + 1]);
      // with registers to spare (at most 16 warps) the per-feature constants of the slab are fetched up front, so the arithmetic
      // below does not wait for one broadcast load after the other
      uint64_t mu8[HOIST ? 8 : 1];
      double mg8[HOIST ? 16 : 1];
      if constexpr (HOIST) {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          if (NS > 1) asm volatile("ld.shared.b64 %0, [%1];" : "=l"(mu8[c]) : "r"(a_mu2 + (uint32_t)(8 * sl + c) * 8u)); else mu8[c] = mu_r[c];
          lds_v2f64(a_mg + (uint32_t)(16 * sl + 2 * c) * 8u, mg8[2 * c], mg8[2 * c + 1]);
        }
      }
      uint32_t tv[16];
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        uint64_t mu01 = HOIST ? mu8[HOIST ? c : 0] : mu_r[NS == 1 ? c : 0], vv, rr;
        if (!HOIST && NS > 1) asm volatile("ld.shared.b64 %0, [%1];" : "=l"(mu01) : "r"(a_mu2 + (uint32_t)(8 * sl + c) * 8u));
        asm("mov.b64 %0, {%1,%2};" : "=l"(vv) : "f"((float)v[2 * c]), "f"((float)v[2 * c + 1]));
        asm("sub.f32x2 %0, %1, %2;" : "=l"(rr) : "l"(vv), "l"(mu01));
        tv[2 * c] = (uint32_t)rr + 0x1000u; tv[2 * c + 1] = (uint32_t)(rr >> 32) + 0x1000u;
        const float a0 = __uint_as_float(tv[2 * c]), a1 = __uint_as_float(tv[2 * c + 1]);
        xa = fmaf(a0, a0, xa); xb = fmaf(a1, a1, xb);
      }
      wl_tmem_st16(t_A + (uint32_t)(16 * sl), tv);
      if constexpr (DIST == 0)
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        double m0, m1;
        if constexpr (HOIST) { m0 = mg8[HOIST ? 2 * c : 0]; m1 = mg8[HOIST ? 2 * c + 1 : 0]; }
        else lds_v2f64(a_mg + (uint32_t)(16 * sl + 2 * c) * 8u, m0, m1);
        sts_v2f64(xr + (((uint32_t)c << 4) ^ xsw), v[2 * c] + m0, v[2 * c + 1] + m1);
      }
    }
    if constexpr (HS != 0) {
      // the trailing 8 features: 64-byte rows, 16-byte column c at c ^ ((row >> 1) & 3)
      const uint32_t xr = xs - (uint32_t)q * 4096u + HSLAB_OFF + (uint32_t)q * 2048u + (uint32_t)lane * 64u;
      const uint32_t hsw = (uint32_t)((lane >> 1) & 3) << 4;
      double v[8];
#pragma unroll
      for (int c = 0; c < 4; ++c) lds_v2f64(xr + (((uint32_t)c << 4) ^ hsw), v[2 * c], v[2 * c + 1]);
      uint32_t tv[8];
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint64_t mu01, vv, rr;
        asm volatile("ld.shared.b64 %0, [%1];" : "=l"(mu01) : "r"(a_mu2 + (uint32_t)(8 * NS + c) * 8u));
        asm("mov.b64 %0, {%1,%2};" : "=l"(vv) : "f"((float)v[2 * c]), "f"((float)v[2 * c + 1]));
        asm("sub.f32x2 %0, %1, %2;" : "=l"(rr) : "l"(vv), "l"(mu01));
        tv[2 * c] = (uint32_t)rr + 0x1000u; tv[2 * c + 1] = (uint32_t)(rr >> 32) + 0x1000u;
        const float a0 = __uint_as_float(tv[2 * c]), a1 = __uint_as_float(tv[2 * c + 1]);
        xa = fmaf(a0, a0, xa); xb = fmaf(a1, a1, xb);
      }
      wl_tmem_st8(t_A + (uint32_t)(16 * NS), tv);
      if constexpr (DIST == 0)
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        double m0, m1;
        lds_v2f64(a_mg + (uint32_t)(16 * NS + 2 * c) * 8u, m0, m1);
        sts_v2f64(xr + (((uint32_t)c << 4) ^ hsw), v[2 * c] + m0, v[2 * c + 1] + m1);
      }
    }
    float xn2 = xa + xb;
    wl_tmem_wait_st();
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");       // fixed-point rows -> visible to the tensor core (second product)
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    const long long tc1 = (XPROF && prof) ? clock64() : 0;
    named_bar_sync(1 + grp, 128 * H);
    const long long tm0 = (XPROF && prof) ? clock64() : 0;
    if (XPROF && prof) { pw3 += tc1 - tc0; pw4 += tm0 - tc1; }
    if (q == 0 && hh == 0 && lane == 0) {
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
      for (int ks = 0; ks < ND + 1; ++ks) {                          // K = 8: eight operand columns, two 16-byte chunks of B
        const uint64_t db = db0 + (uint64_t)((uint32_t)ks * db_step);
        const uint32_t accf = ks > 0 ? 1u : 0u;
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n}\n" ::"r"(dcol),
                     "r"(acol + (uint32_t)(8 * ks)), "l"(db), "r"(idesc), "r"(accf)
                     : "memory");
      }
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar_t) : "memory");
    }
    __syncwarp();
    // ---- while the scores are computed: close this warp's previous tile.  Its limb product (issued at the end of the last
    // iteration) has long finished: clear the one-hot byte and fetch the tile that uses that quarter of the ring next
    if (pend) {
      tp_wait_p(bar_r, tph ^ 1u, 8, prof, pw1);
      if (hb_prev != 0u) sts_u8(hb_prev, 0u);
      __syncwarp();
      refill(i - G + S, xs_prev, bar_f_prev, seqa_prev);
      pend = false;
    }
    if (XPROF && prof) pw7 += clock64() - tm0;
    tp_wait_p(bar_t, tph, 5, prof, pw1);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const long long ts0 = (XPROF && prof) ? clock64() : 0;
    // ---- sweep 1: top two of the packed scores (tournament per 16 columns)
    float m1 = 0.f, m2 = 0.f;
#pragma unroll
    for (int ch = 0; ch < NCH; ++ch) {
      if (H == 2 && (ch < ch_lo || ch >= ch_hi)) continue;       // the other thread of the row sweeps these columns
      uint32_t v[16];
      tp_tmem_ld16(t_D + (uint32_t)(ch * 16), v);
      tp_tmem_wait_ld();
      float hi[8], lo[8];
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        uint32_t pa, pb;
        asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(pa) : "r"(v[2 * c]), "r"(idxclr), "r"((uint32_t)(ch * 16 + 2 * c)));
        asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(pb) : "r"(v[2 * c + 1]), "r"(idxclr), "r"((uint32_t)(ch * 16 + 2 * c + 1)));
        hi[c] = fmaxf(__uint_as_float(pa), __uint_as_float(pb)); lo[c] = fminf(__uint_as_float(pa), __uint_as_float(pb));
      }
#pragma unroll
      for (int w2 = 4; w2 >= 1; w2 >>= 1)
#pragma unroll
        for (int c = 0; c < w2; ++c) {
          const float t = fminf(hi[c], hi[c + w2]);
          asm("max.f32 %0, %1, %2, %3;" : "=f"(lo[c]) : "f"(t), "f"(lo[c]), "f"(lo[c + w2]));
          hi[c] = fmaxf(hi[c], hi[c + w2]);
        }
      if (ch == ch_lo) { m1 = hi[0]; m2 = lo[0]; }
      else {
        const float t = fminf(m1, hi[0]);
        asm("max.f32 %0, %1, %2, %3;" : "=f"(m2) : "f"(t), "f"(m2), "f"(lo[0]));
        m1 = fmaxf(m1, hi[0]);
      }
    }
    if constexpr (H == 2) {
      // the two threads of the row swap their top two and their share of |x - mu|^2, and both form the same decision
      asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(a_ex + 16u * (uint32_t)hh), "f"(m1), "f"(m2), "f"(xn2), "f"(0.f) : "memory");
      named_bar_sync(pair_id, 64);
      float o1, o2, o3, o4;
      asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(o1), "=f"(o2), "=f"(o3), "=f"(o4) : "r"(a_ex + 16u * (uint32_t)(hh ^ 1)) : "memory");
      const float t = fminf(m1, o1);
      asm("max.f32 %0, %1, %2, %3;" : "=f"(m2) : "f"(t), "f"(m2), "f"(o2));
      m1 = fmaxf(m1, o1);
      xn2 += o3;
    }
    const long long te0 = (XPROF && prof) ? clock64() : 0;
    if (XPROF && prof) pw5 += te0 - ts0;
    const bool valid = row < rows_valid;
    const float eps = fmaf(kapv, sqrtf(xn2) * 1.001f, etab) + 6e-5f * fabsf(m1);
    const float thr = m1 - 2.0f * eps;
    const bool amb = valid && !(m2 < thr);
    int lab = (int)(__float_as_uint(m1) & IDXMASK);
    unsigned ambmask = __ballot_sync(0xffffffffu, amb);
    if (ambmask) {
      // sweep 2 + exact re-scoring of the undecided rows (hkm_lloyd_wg.cuh).  The stage holds the fixed-point form by now, so the
      // row's fp64 values come from global memory again (the table was just streamed: an L2 hit more often than not)
      const long long tr0 = prof ? clock64() : 0;
      // with registers to spare, the first undecided row's values are requested before the second sweep, each further row's while
      // its predecessor is scored
      const double* xg = p.X + (((int64_t)blockIdx.x + (int64_t)i * gridDim.x) * kTileRows + 32 * q) * d + rlane;
      int srcn = __ffs(ambmask) - 1;
      double nx0 = 0.0, nx1 = 0.0;
      if (HOIST && ract) { const double2 xv = __ldg(reinterpret_cast<const double2*>(xg + (size_t)srcn * d)); nx0 = xv.x; nx1 = xv.y; }
      if constexpr (H == 2) named_bar_sync(pair_id, 64);         // the other thread has read this row's sweep slot: the slots are reused below
      uint32_t w[NW];
#pragma unroll
      for (int qq = 0; qq < NW; ++qq) w[qq] = 0u;
#pragma unroll
      for (int ch = 0; ch < NCH; ++ch) {
        if (H == 2 && (ch < ch_lo || ch >= ch_hi)) continue;
        uint32_t v[16];
        tp_tmem_ld16(t_D + (uint32_t)(ch * 16), v);
        tp_tmem_wait_ld();
#pragma unroll
        for (int c = 0; c < 16; ++c) w[ch >> 1] |= (__uint_as_float(v[c]) >= thr) ? (1u << ((ch & 1) * 16 + c)) : 0u;
      }
      while (ambmask) {
        const int src = HOIST ? srcn : __ffs(ambmask) - 1; ambmask &= ambmask - 1;
        uint32_t cm[NW];
#pragma unroll
        for (int qq = 0; qq < NW; ++qq) cm[qq] = __shfl_sync(0xffffffffu, w[qq], src);
        double x0 = nx0, x1 = nx1;
        if constexpr (HOIST) {
          if (ambmask) {
            srcn = __ffs(ambmask) - 1;
            if (ract) { const double2 xv = __ldg(reinterpret_cast<const double2*>(xg + (size_t)srcn * d)); nx0 = xv.x; nx1 = xv.y; }
          }
        } else if (ract) { const double2 xv = __ldg(reinterpret_cast<const double2*>(xg + (size_t)src * d)); x0 = xv.x; x1 = xv.y; }
        double best = -__longlong_as_double(0x7ff0000000000000LL); int bj = -1;
        constexpr int CB = 2;                                   // candidates scored side by side: an undecided row mostly has two
#pragma unroll
        for (int qq = 0; qq < NW; ++qq) {
          uint32_t msk = cm[qq];
          while (msk) {
            int jj[CB]; double pp[CB];
#pragma unroll
            for (int u = 0; u < CB; ++u) { jj[u] = msk ? 32 * qq + __ffs(msk) - 1 : -1; msk &= msk - 1; }
#pragma unroll
            for (int u = 0; u < CB; ++u) {
              double c0 = 0.0, c1 = 0.0;
              if (ract && jj[u] >= 0) { const double* cr = Cr + (size_t)jj[u] * d + rlane; c0 = cr[0]; c1 = cr[1]; }
              pp[u] = fma(x0, c0, x1 * c1);
            }
#pragma unroll
            for (int mk = RLANES / 2; mk >= 1; mk >>= 1) {
#pragma unroll
              for (int u = 0; u < CB; ++u) pp[u] += shfl_xor_d(pp[u], mk);
            }
#pragma unroll
            for (int u = 0; u < CB; ++u)
              if (jj[u] >= 0) { const double sc = pp[u] + cn_s[jj[u]]; if (sc > best) { best = sc; bj = jj[u]; } }
          }
        }
        bj = __shfl_sync(0xffffffffu, bj, 0);
        if constexpr (H == 1) { if (lane == src) lab = bj; }
        else if (lane == 0) {
          // best of this thread's columns; the row's two halves meet below
          asm volatile("st.shared.v2.u32 [%0], {%1,%2};" ::"r"(a_ex + (uint32_t)(src - lane) * 32u + 16u * (uint32_t)hh),
                       "r"((uint32_t)__double2loint(best)), "r"((uint32_t)__double2hiint(best)) : "memory");
          sts_u32(a_ex + (uint32_t)(src - lane) * 32u + 16u * (uint32_t)hh + 8u, (uint32_t)bj);
        }
        ++pcnt;
      }
      if constexpr (H == 2) {
        named_bar_sync(pair_id, 64);
        if (amb) {
          double b0 = lds_f64(a_ex), b1 = lds_f64(a_ex + 16u);
          const int j0 = (int)lds_u32(a_ex + 8u), j1 = (int)lds_u32(a_ex + 24u);
          lab = (j1 >= 0 && (j0 < 0 || b1 > b0)) ? j1 : j0;    // equal scores: the lower index (first half) wins
        }
      }
      if (prof) pw2 += clock64() - tr0;
    }
    if constexpr (DIST != 0) {
      if (p.mind2_out == nullptr) {
        // labels only (transform / predict): nothing else to compute for the row
        if (valid) p.labels_out[((int64_t)blockIdx.x + (int64_t)i * gridDim.x) * kTileRows + row] = lab;
        __syncwarp();
        if (i + S < T) tma_load(i + S, xs, bar_f, seqa);
        s += G; if (s >= S) { s -= S; ph ^= 1u; }
        tph ^= 1u;
        continue;
      }
      // ---- exact distance to the winner from the row's fp64 values (the stage still holds them) and the padded centre table:
      // score = x.c - 0.5|c|^2 and |x|^2 in ascending feature order, dist = max(|x|^2 - 2 score, 0), as every other variant
      double sc = 0.0, xx = 0.0;
      const uint32_t crow = sb + L.ct_off + (uint32_t)lab * (uint32_t)L.ct_pitch;
#pragma unroll
      for (int sl = 0; sl < NS; ++sl) {
        const uint32_t xr = xs + (uint32_t)sl * SLAB + xrow;
        double xv[16], cv[16];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          lds_v2f64(xr + (((uint32_t)c << 4) ^ xsw), xv[2 * c], xv[2 * c + 1]);
          lds_v2f64(crow + (uint32_t)(16 * sl + 2 * c) * 8u, cv[2 * c], cv[2 * c + 1]);
        }
#pragma unroll
        for (int f = 0; f < 16; ++f) { sc = fma(xv[f], cv[f], sc); xx = fma(xv[f], xv[f], xx); }
      }
      if constexpr (HS != 0) {
        const uint32_t xr = xs - (uint32_t)q * 4096u + HSLAB_OFF + (uint32_t)q * 2048u + (uint32_t)lane * 64u;
        const uint32_t hsw = (uint32_t)((lane >> 1) & 3) << 4;
        double xv[8], cv[8];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          lds_v2f64(xr + (((uint32_t)c << 4) ^ hsw), xv[2 * c], xv[2 * c + 1]);
          lds_v2f64(crow + (uint32_t)(16 * NS + 2 * c) * 8u, cv[2 * c], cv[2 * c + 1]);
        }
#pragma unroll
        for (int f = 0; f < 8; ++f) { sc = fma(xv[f], cv[f], sc); xx = fma(xv[f], xv[f], xx); }
      }
      const double dist = fmax(fma(-2.0, sc + cn_s[lab], xx), 0.0);
      if (valid) {
        const int64_t r = ((int64_t)blockIdx.x + (int64_t)i * gridDim.x) * kTileRows + row;
        if (p.out_mode == 0) {
          if (p.labels_out != nullptr) p.labels_out[r] = lab;
          p.mind2_out[r] = dist;
          cost += dist;
        } else {
          // running minimum over successive centre chunks: strict '<' keeps the earlier (lower-index) centre on ties
          const double old = p.mind2_out[r];
          if (dist < old) {
            p.mind2_out[r] = dist;
            if (p.labels_out != nullptr) p.labels_out[r] = lab + p.label_base;
            cost += dist;
          } else cost += old;
        }
      }
      __syncwarp();
      if (i + S < T) tma_load(i + S, xs, bar_f, seqa);
      s += G; if (s >= S) { s -= S; ph ^= 1u; }
      tph ^= 1u;
      continue;
    }
    // ---- this row's one-hot byte, label, count (rows past the end take no part)
    const uint32_t hb = hrow + 16u * (uint32_t)lab;
    if (valid && hh == 0) {
      sts_u8(hb, 1u);
      asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(a_cnt + 4u * (uint32_t)lab) : "memory");
      if (p.labels_out != nullptr) {
        const int64_t r0 = ((int64_t)blockIdx.x + (int64_t)i * gridDim.x) * kTileRows;
        p.labels_out[r0 + row] = lab;
      }
    }
    // ---- the limb product of this warp's 32 rows (one K step of the tile), issued by the warp itself: A = its quarter of the
    // stage, B = its two chunks of the one-hot operand.  Integer accumulation is order free, and accumulating products issued by
    // different threads into one accumulator are applied one after the other by the tensor pipe (tools/i8_limb_test.cu checks that)
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncwarp();
    if (lane == 0 && hh == 0) {
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint64_t dh = dh0 + (uint64_t)((uint32_t)q * dh_step);
#pragma unroll
      for (int sl = 0; sl < NS; ++sl) {
        const uint64_t da = wl_desc(xs + (uint32_t)sl * SLAB, 16u, 1024u, 2u);
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n" ::"r"(rcol + (uint32_t)(sl * KP)),
                     "l"(da), "l"(dh), "r"(idesc2), "r"(1u)
                     : "memory");
      }
      if constexpr (HS != 0) {
        const uint64_t da = wl_desc(xs - (uint32_t)q * 4096u + HSLAB_OFF + (uint32_t)q * 2048u, 16u, 512u, 4u);   // SWIZZLE_64B, 8 rows x 64 bytes per atom
        const uint32_t idesc2h = (2u << 4) | (1u << 15) | ((uint32_t)(KP >> 3) << 17) | ((uint32_t)(64 >> 4) << 24);
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n" ::"r"(rcol + (uint32_t)(NS * KP)),
                     "l"(da), "l"(dh), "r"(idesc2h), "r"(1u)
                     : "memory");
      }
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar_r) : "memory");
    }
    __syncwarp();
    if (XPROF && prof) pw8 += clock64() - te0;
    hb_prev = (valid && hh == 0) ? hb : 0u; xs_prev = xs; bar_f_prev = bar_f; seqa_prev = seqa; pend = true;
    s += G; if (s >= S) { s -= S; ph ^= 1u; }
    tph ^= 1u;
  }
  // close the last tile: its product must be complete before the accumulators are read
  if (pend) { tp_wait_p(bar_r, tph ^ 1u, 8, prof, pw1); }

  if (prof && lane == 0) {
    unsigned long long* o = prof_buf + warp * 8;
    o[0] = (unsigned long long)(clock64() - pt0); o[1] = (unsigned long long)pw0; o[2] = (unsigned long long)pw1; o[3] = (unsigned long long)pw2;
    o[4] = (unsigned long long)pw3; o[5] = (unsigned long long)pcnt; o[6] = (unsigned long long)T; o[7] = (unsigned long long)pw4;
    // the first eight warps also report: sweep + exchange; top of the iteration; product issue + close of the previous tile; rest of the iteration
    if (XPROF && warp < 8) { prof_buf[192 + warp] = (unsigned long long)pw5; prof_buf[200 + 4 * warp] = (unsigned long long)pw6; prof_buf[201 + 4 * warp] = (unsigned long long)pw7; prof_buf[202 + 4 * warp] = (unsigned long long)pw8; }
  }
  if constexpr (DIST != 0) {
    // ===================================================================== distance variant: only the cost is folded
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(L.tcols) : "memory");
    if (p.do_reduce != 2) return;
#pragma unroll
    for (int msk = 16; msk >= 1; msk >>= 1) cost += shfl_xor_d(cost, msk);
    if (lane == 0) red_s[warp] = cost;
    __syncthreads();
    __shared__ unsigned int s_tk;
    if (tid == 0) {
      double c = 0;
      for (int wv = 0; wv < nwarps + PW; ++wv) c += red_s[wv];
      p.cta_partials[blockIdx.x] = c;
      __threadfence();
      s_tk = atomicAdd(p.counter, 1u);
      if (s_tk == gridDim.x - 1) {
        __threadfence();
        p.partial_out[0] = ordered_sum_ldcg(p.cta_partials, 1, gridDim.x);   // CTA order: bit-stable
        *p.counter = 0u;
      }
    }
    return;
  }
  // ===================================================================== CTA-level combine (all warps)
  // every tile is reduced and no copy is in flight: the s32 limb sums go from TMEM to the idle ring, acc[sl KP + c][byte]
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  int32_t* acc_s = reinterpret_cast<int32_t*>(smem + L.x_off);
  if (grp == 0 && hh == 0) {
#pragma unroll 1
    for (int sl = 0; sl < NS + HS; ++sl)
#pragma unroll 1
      for (int c0 = 0; c0 < KP; c0 += 16) {
        uint32_t v[16];
        tp_tmem_ld16(tmem + tlane + (uint32_t)L.rcol0 + (uint32_t)(sl * KP + c0), v);
        tp_tmem_wait_ld();
#pragma unroll
        for (int c = 0; c < 16; ++c) acc_s[(size_t)(sl * KP + c0 + c) * 128 + row] = (int32_t)v[c];
      }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(L.tcols) : "memory");
  for (int idx = tid; idx < k * d; idx += nthreads) {
    const int j = idx / d, f = idx - j * d, sl = f >> 4;
    int b0 = 8 * (f & 15);
    if (HS != 0 && sl == NS) b0 = (b0 & 15) + 32 * (b0 >> 4);   // M = 64 product: result row m in lane m % 16 + 32 (m / 16)
    __int128 tot = 0;
    const int32_t* a = acc_s + (size_t)(sl * KP + j) * 128 + b0;
#pragma unroll
    for (int l = 0; l < 8; ++l) tot += (__int128)(long long)a[l] << (8 * l);
    const unsigned long long mbits = (unsigned long long)__double_as_longlong(mg_s[f]);
    tot -= (__int128)(unsigned long long)cnt_s[j] * (__int128)mbits;            // every row of the cluster carried the bits of M_f
    // tot = sum of round(x / 2^(E_f - 50)); |tot| < 2^83: two halves, one rounding each
    const bool neg = tot < 0;
    const unsigned __int128 mag = neg ? (unsigned __int128)(-tot) : (unsigned __int128)tot;
    const double hi = (double)(unsigned long long)(mag >> 64), lo = (double)(unsigned long long)mag;
    const int pexp = (int)((mbits >> 52) & 0x7ffull) - 1023 - 52;                 // value of one unit
    double sacc = scalbn(fma(hi, 18446744073709551616.0, lo), pexp);
    if (neg) sacc = -sacc;
    out[idx] = sacc;
    cost = fma(sacc, p.C[idx], cost);                           // sum_j S_j . c_j
  }
  if (tid < k) { const double nj = (double)cnt_s[tid]; out[(size_t)k * d + tid] = nj; cost = fma(nj, cn_s[tid], cost); }
#pragma unroll
  for (int msk = 16; msk >= 1; msk >>= 1) cost += shfl_xor_d(cost, msk);
  if (lane == 0) red_s[warp] = cost;
  __syncthreads();
  if (tid == 0) {
    double c = 0;
    for (int wv = 0; wv < nwarps + PW; ++wv) c += red_s[wv];
    out[P - 1] = c; out[P] = 0.0;
  }
  // ===================================================================== two-level fold of the per-CTA partials
  // The CTAs form groups of 16; the last CTA of a group to finish folds the group's partials into a group slot (behind the
  // per-CTA slots: this kernel launches one CTA per SM, the scratch is sized for two), and the last group folds the group slots.
  // Fixed shape, so the result is the same whatever the order of arrival; a single CTA summing gridDim.x values per entry would
  // pull CTAs x P doubles through one SM.
  __threadfence();
  __syncthreads();
  __shared__ unsigned int s_ticket;
  constexpr unsigned GS = 16u;
  const unsigned ngroups = (gridDim.x + GS - 1u) / GS, mygroup = blockIdx.x / GS;
  const unsigned gsize = min(GS, gridDim.x - mygroup * GS);
  if (tid == 0) s_ticket = atomicAdd(p.counter + 2 + mygroup, 1u);
  __syncthreads();
  if (s_ticket != gsize - 1u) return;
  __threadfence();
  {
    double* gslot = p.cta_partials + (size_t)(gridDim.x + mygroup) * SL;
    const double* src = p.cta_partials + (size_t)(mygroup * GS) * SL;
    for (int i = tid; i < P; i += nthreads) gslot[i] = ordered_sum_ldcg(src + i, (size_t)SL, gsize);
    if (tid == 0) p.counter[2 + mygroup] = 0u;
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) s_ticket = atomicAdd(p.counter, 1u);
  __syncthreads();
  if (s_ticket == ngroups - 1u) {
    __threadfence();
    const double* gs = p.cta_partials + (size_t)gridDim.x * SL;
    for (int i = tid; i < P - 1; i += nthreads) p.partial_out[i] = ordered_sum_ldcg(gs + i, (size_t)SL, ngroups);
    __syncthreads();
    if (tid == 0) {
      const double ssc = ordered_sum_ldcg(gs + (P - 1), (size_t)SL, ngroups);
      p.partial_out[P - 1] = fmax(fma(-2.0, ssc, *p.xnorm2_sum), 0.0);      // cost = sum |x|^2 - 2 sum score (REUSE pass)
      *p.counter = 0u;
      rec->done = 1;                                                         // the launch behind this one has nothing to do
    }
    bool peer_ok = true;
    if (p.peer_bufs != nullptr) {
      __threadfence_block();
      __syncthreads();
      peer_ok = peer_allreduce(p, p.partial_out, P, tid, nthreads, [] { __syncthreads(); });
    }
    if (p.C_update != nullptr && peer_ok) {
      __threadfence_block();
      __syncthreads();
      update_centers_body(p.partial_out, k, d, p.C_update, p.tol2, p.state_rw, red_s + 2 * nwarps, tid);
    }
  }
}

}  // namespace hkm

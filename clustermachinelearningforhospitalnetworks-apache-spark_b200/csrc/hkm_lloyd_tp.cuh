// Lloyd data pass, pipelined tensor-core variant (8 < k <= 64, d <= 64): tcgen05 screening + fp64 refinement with the
// roles of one CTA per SM split over specialised warps, so conversion, MMA, screening and the per-cluster reduction of
// consecutive 128-row tiles overlap.
//
// Why: for k > 8 the fp64 tensor pipe cannot score AND reduce at the HBM rate (k = 16, d = 16 needs 8 fp64 FMA per
// byte streamed).  tcgen05 has no fp64 kind, so the k scores of a row are first computed approximately in tf32
// (operands centred by mu = mean of the centres, rounded to nearest tf32 by the CUDA cores, accumulator in TMEM) with a
// rigorous error bound
//     |approx_j - exact_j| <= eps = kappa |x - mu| max_j |c_j - mu| + eta max_j |c_j - mu|^2,   kappa = 1.04 * 2^-10
// (two operands rounded to 11 bits each, fp32 accumulation over K <= 72 terms, 6 index bits packed into the score).
// A row whose best approximate score beats the second best by more than 2 eps is decided; any other row is re-scored
// exactly in fp64 against every centre within 2 eps of the best, with the same score, tie rule (lowest index) and
// arithmetic as the fp64 kernels.  The tensor cores only prune; labels are the fp64 argmax.
//
// Warp roles (18 warps, 576 threads, one CTA per SM):
//   screen   4 or 8 warps (one or two groups of four; group g takes tiles i = g mod groups): thread = tile row; reads
//            the row's tf32 operand back for |x - mu|, sweeps the TMEM accumulator once (top-2 with the centre index
//            packed into the low mantissa bits) and decides the row; undecided rows of the warp are then re-scored
//            by that warp cooperatively (lane = feature pair).  Publishes the tile's labels.
//   reduce   8 or 4 warps: per-cluster fp64 sums, three schemes picked per shape:
//            k <= 16: S += onehot(labels)^T X on the fp64 tensor pipe (DMMA.8x8x4), accumulators in registers, every
//                     warp a fixed share of the rows -- k*d/64 pipe clocks per row, 0.7 of the pipe at the HBM rate;
//            small k*d: private accumulator sets in shared memory, a fixed share of the rows per warp, several rows
//                     per instruction, each in its own set (no ballots, no imbalance, but 24 d bytes of shared-memory
//                     traffic per row);
//            otherwise: warp r owns the clusters j = r mod R, accumulators in registers, rows found by ballot.
//   convert  4 warps: fp64 stage -> (x - mu) rounded to tf32 -> canonical K-major UMMA operand.
//   producer 1 warp: 1-D bulk-async copies into the ring;   mma 1 warp: one thread issues tcgen05.mma + commit.
// Everything meets through mbarriers (full/empty per ring stage, per operand buffer, per TMEM buffer, labels-ready
// per stage); the only CTA barrier is at the end.  Constant -0.5|c - mu|^2 rides inside the MMA as three extra K
// columns (a tf32 hi/mid/lo split against a column of ones), so the sweep is ~4.5 instructions per centre.
// Rows per cluster are counted with shared-memory integer atomics; the cost needs no per-row work:
//     sum_rows (x.c_label - 0.5|c_label|^2) = sum_j S_j.c_j - 0.5 sum_j n_j |c_j|^2.
// Summation order is fixed (row order within a reduce warp, warp / set order in the combine, CTA order in the fold),
// so results are bit-stable.  Plain passes only (unit weights, no per-row distances); the dispatcher sends the rest
// elsewhere.
#pragma once
#include "hkm_lloyd.cuh"

namespace hkm {

// warp budgets (WCFG): 0 = 1 screen group + 8 reduce + 4 convert (18 warps), 1 = 2 + 8 + 8 (26 warps, thin rows: every role is
// latency bound, so more warps per role), 2 = 2 + 4 + 4 (18 warps, k > 32 with private sets); + producer + MMA issuer
// 3 = 1 + 16 + 8 (30 warps: large k*d, where one tile's convert -> score -> reduce chain is what the two-stage ring waits for);
// 4 = budget 0 with two CTAs per SM (thin rows: twice the warps per role, half the shared memory each)
constexpr int kTpMaxWarps = 30;
__host__ __device__ constexpr int tp_sg(int wcfg) { return (wcfg == 1 || wcfg == 2) ? 2 : 1; }
__host__ __device__ constexpr int tp_reduce_warps(int wcfg) { return wcfg == 2 ? 4 : (wcfg == 3 ? 16 : 8); }
__host__ __device__ constexpr int tp_conv_warps(int wcfg) { return (wcfg == 1 || wcfg == 3) ? 8 : 4; }
__host__ __device__ constexpr int tp_ctas_per_sm(int wcfg) { return wcfg == 4 ? 2 : 1; }
__host__ __device__ constexpr int tp_warps(int wcfg) { return 4 * tp_sg(wcfg) + tp_reduce_warps(wcfg) + tp_conv_warps(wcfg) + 2; }
__host__ __device__ constexpr int tp_tmem_stride(int kp) { return kp <= 32 ? 32 : (kp <= 64 ? 64 : (kp <= 128 ? 128 : 256)); }
// feature-count class of ND = ceil(d / 8): rows are handled as if they had 8, 16, 32 or 64 features
__host__ __device__ constexpr int tp_dclass(int nd) { return nd <= 1 ? 8 : (nd <= 2 ? 16 : (nd <= 4 ? 32 : 64)); }
__host__ __device__ constexpr int tp_lanes_per_row(int nd, int vec) { return tp_dclass(nd) / vec < 4 ? 4 : (tp_dclass(nd) / vec > 32 ? 32 : tp_dclass(nd) / vec); }

// reduce modes: 0 private sets in shared memory, lane = feature (d <= 32); 2 one shared set, clusters owned per warp, lane =
//               feature pair (even d <= 64); 3 one-hot DMMA (k <= 16), accumulators in registers
struct TpPlan {
  int stages, rmode, na, wcfg, nd;   // ring depth, reduce mode, tf32 operand buffers (1 or 2), warp budget, ceil(d/8)
  int ald, nsets;                  // rmode 0/1: accumulator row (doubles), sets
  int kchunks, lbo_a, lbo_b, a_bytes;
  int x_off, a_off, b_off, acc_off, cn_off, mu_off, cnt_off, lab_off, red_off, bar_off, tmem_off, cs_off, total;   // cs_off: fp64 centres for the refine (0: read from global)
  unsigned long long* prof;   // developer profile (tools/tp_check.py prof): [warp][8] cycle counters of CTA 0, or null
};

__host__ __device__ inline TpPlan tp_plan(int d, int kp, int stages, int rmode, int na, int wcfg) {
  TpPlan L{};
  L.stages = stages; L.rmode = rmode; L.na = na; L.wcfg = wcfg; L.nd = (d + 7) / 8;
  const int cvec = (d & 1) ? 1 : 2;
  const int c_g = 32 / tp_lanes_per_row(L.nd, cvec);
  const int R = tp_reduce_warps(wcfg);
  if (rmode == 0) {
    const int rl = tp_lanes_per_row(L.nd, 1);
    L.ald = rl; L.nsets = R * (32 / rl);
  } else if (rmode == 2) { L.ald = 64; L.nsets = 1; }
  const int d8 = (d + 7) & ~7;
  L.kchunks = (d8 + 8) / 4;
  const int pad_a = 8 * c_g > 16 ? 8 * c_g : 16;
  L.lbo_a = (kTileRows / 8) * 128 + pad_a;
  L.lbo_b = (kp / 8) * 128 + 16;
  L.a_bytes = (L.kchunks * L.lbo_a + 127) & ~127;
  int o = 0;
  L.x_off = o; o += stages * stage_bytes(d); o = (o + 127) & ~127;
  L.a_off = o; o += na * L.a_bytes;
  L.b_off = o; o += (L.kchunks * L.lbo_b + 127) & ~127;
  L.acc_off = o; o += L.nsets * kp * L.ald * 8;
  L.cn_off = o; o += kp * 8;
  L.mu_off = o; o += (d8 + 8) * 8;
  L.cnt_off = o; o += kp * 4;
  L.lab_off = o; o += (stages > 0 ? stages : 2) * kTileRows * 4;
  L.red_off = o; o += (2 * kTpMaxWarps + 64) * 8;
  L.bar_off = o; o += (3 * kMaxStages + 8) * 8;
  L.tmem_off = o; o += 16;
  // the exact re-scoring reads whole centres; a shared-memory copy keeps that off the L2 round trip when it is small
  if (kp * d * 8 <= 8 * 1024) { L.cs_off = o; o += kp * d * 8; }
  L.total = o;
  return L;
}

// Picks reduce mode, warp budget, operand buffering and ring depth for a shape; stages < 2 => the shape does not fit.
// force_rmode / force_wcfg >= 0 pin the reduce mode / warp budget (tuning).
__host__ inline TpPlan tp_choose(int d, int kp, int force_rmode = -1, int force_wcfg = -1) {
  const int nd = (d + 7) / 8;
  // candidates in order of preference: (reduce mode, warp budget)
  int cand[10][2]; int nc = 0;
  auto add = [&](int rm, int w) { if (force_wcfg < 0 || force_wcfg == w) { cand[nc][0] = rm; cand[nc][1] = w; ++nc; } };
  if (kp == 16 && (force_rmode < 0 || force_rmode == 3)) { if (force_wcfg == 4) add(3, 4); add(3, nd <= 3 ? 1 : 0); if (force_wcfg == 0 || force_wcfg == 1) add(3, 1 - (nd <= 3 ? 1 : 0)); }   // measured: 26 warps up to d = 24, 18 above
  if (d <= 32 && (force_rmode < 0 || force_rmode == 0)) { if (force_wcfg == 4) add(0, 4); add(0, 1); if (kp >= 48) add(0, 2); add(0, 0); }
  if (!(d & 1) && (force_rmode < 0 || force_rmode == 2)) { add(2, 3); add(2, 0); }
  TpPlan best{}; best.stages = 0;
  for (int ci = 0; ci < nc; ++ci)
    for (int na = 2; na >= 1; --na) {
      const int rm = cand[ci][0], w = cand[ci][1];
      const int budget = (227 * 1024) / tp_ctas_per_sm(w) - 1024 - (tp_ctas_per_sm(w) > 1 ? 1024 : 0);
      const int fixed = tp_plan(d, kp, 2, rm, na, w).total - 2 * stage_bytes(d) - 2 * kTileRows * 4;
      int st = (budget - fixed) / (stage_bytes(d) + kTileRows * 4);
      if (st > kMaxStages) st = kMaxStages;
      if (st < 2) continue;
      if (st >= 3) return tp_plan(d, kp, st, rm, na, w);
      if (st > best.stages) best = tp_plan(d, kp, st, rm, na, w);
    }
  return best;
}

__device__ __forceinline__ uint32_t f32_to_tf32_rn(float v) { uint32_t r; asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v)); return r; }
__device__ __forceinline__ uint64_t tp_desc_kmajor(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  // K-major, no swizzle: element (row r, 16-byte chunk c) sits at (r % 8) * 16 + (r / 8) * SBO + c * LBO
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32) |
         ((uint64_t)1 << 46);
}
__device__ __forceinline__ void tp_tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
                 "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
               : "r"(taddr));
}
// mbarrier wait: the thread sleeps in hardware until the phase flips (suspend-time hint 1 ms per try, so a waiting warp
// costs no issue slots); a protocol bug ends the kernel with a message after ~10 s instead of hanging the GPU
__device__ __forceinline__ bool tp_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\nselp.u32 %0, 1, 0, p;\n}\n"
               : "=r"(ok) : "r"(bar), "r"(parity), "r"(1000000u) : "memory");
  return ok != 0;
}
static __device__ __noinline__ void tp_wait_slow(uint32_t bar, uint32_t parity, int tag) {
  // the sleep ends at every barrier event of the CTA, so a wait wakes a few dozen times per tile: keep the wake-up path short
  // (eight tries per loop trip; the watchdog reads the wall clock every 64 trips and gives up after 10 s)
  unsigned long long t0 = 0;
#pragma unroll 1
  for (unsigned it = 0;; ++it) {
#pragma unroll
    for (int u = 0; u < 8; ++u) if (tp_try_wait(bar, parity)) return;
    if ((it & 63u) == 63u) {
      unsigned long long now;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      if (t0 == 0) t0 = now; else if (now - t0 > 10000000000ull) break;
    }
  }
  printf("lloyd_tp: block %d warp %d stuck on barrier tag %d (smem 0x%x, parity %u)\n", (int)blockIdx.x, (int)(threadIdx.x >> 5), tag, bar, parity);
  __trap();
}
__device__ __forceinline__ void tp_wait(uint32_t bar, uint32_t parity, int tag) {
  if (tp_try_wait(bar, parity)) return;
  tp_wait_slow(bar, parity, tag);
}
// wait that adds the cycles it spent to *acc when profiling
__device__ __forceinline__ void tp_wait_p(uint32_t bar, uint32_t parity, int tag, bool prof, long long& acc) {
  if (!prof) { tp_wait(bar, parity, tag); return; }
  const long long t0 = clock64();
  tp_wait(bar, parity, tag);
  acc += clock64() - t0;
}
__device__ __forceinline__ void tp_tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive_a(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void lds_v2f64(uint32_t a, double& x, double& y) { asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(x), "=d"(y) : "r"(a)); }
__device__ __forceinline__ void sts_v2f64(uint32_t a, double x, double y) { asm volatile("st.shared.v2.f64 [%0], {%1,%2};" ::"r"(a), "d"(x), "d"(y) : "memory"); }

// KP = padded cluster count (16, 32, 48, 64); CVEC = 2 (even d) or 1; RMODE as in TpPlan; WCFG = warp budget; ND = ceil(d/8).
template <int KP, int CVEC, int RMODE, int WCFG, int ND>
__global__ void __launch_bounds__(tp_warps(WCFG) * 32, tp_ctas_per_sm(WCFG)) lloyd_tp_kernel(const LloydParams p, const TpPlan L) {
  if (p.skip != nullptr && *p.skip != 0) return;
  if (p.skip2 != nullptr && *p.skip2 != 0) return;
  constexpr int SG = tp_sg(WCFG), R = tp_reduce_warps(WCFG), CW = tp_conv_warps(WCFG);
  constexpr int kTpWarps = tp_warps(WCFG), kTpThreads = kTpWarps * 32;
  constexpr int kTpConvWarp0 = 4 * SG + R, kTpProdWarp = kTpConvWarp0 + CW, kTpMmaWarp = kTpProdWarp + 1;
  constexpr int NW = (KP + 31) / 32;              // candidate mask words
  constexpr int TSTRIDE = tp_tmem_stride(KP);
  constexpr int TCOLS = 2 * TSTRIDE;
  constexpr int IDXBITS = 6;
  constexpr uint32_t IDXMASK = (1u << IDXBITS) - 1u;
  static_assert(KP % 16 == 0 && KP >= 16 && KP <= 64, "k_pad: 16..64 in steps of 16");
  static_assert(RMODE != 3 || (KP == 16 && R == 8), "one-hot DMMA reduce: k <= 16, eight reduce warps");
  static_assert(RMODE == 0 || RMODE == 2 || RMODE == 3, "reduce mode");

  extern __shared__ __align__(1024) unsigned char smem[];
  const int d = p.d, k = p.k;
  const int S = L.stages;
  const uint32_t sb = smem_u32(smem);
  const uint32_t a_x = sb + L.x_off, a_A = sb + L.a_off, a_B = sb + L.b_off, a_acc = sb + L.acc_off;
  const uint32_t a_lab = sb + L.lab_off, a_cnt = sb + L.cnt_off;
  const uint32_t a_bar = sb + L.bar_off;
  const uint32_t b_full = a_bar, b_sempty = a_bar + 8u * kMaxStages, b_lfull = a_bar + 16u * kMaxStages;
  const uint32_t b_afull = a_bar + 24u * kMaxStages, b_aempty = b_afull + 16u, b_tfull = b_afull + 32u, b_tempty = b_afull + 48u;
  double* cn_s = reinterpret_cast<double*>(smem + L.cn_off);
  double* mu_s = reinterpret_cast<double*>(smem + L.mu_off);
  double* red_s = reinterpret_cast<double*>(smem + L.red_off);
  unsigned int* cnt_s = reinterpret_cast<unsigned int*>(smem + L.cnt_off);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t cstride = (uint32_t)chunk_stride_bytes(d);
  const uint32_t sbytes = (uint32_t)stage_bytes(d);
  const uint32_t row_bytes = (uint32_t)d * 8u;
  const int d8 = (d + 7) & ~7;
  const uint32_t LBO_A = (uint32_t)L.lbo_a, LBO_B = (uint32_t)L.lbo_b;
  const int64_t full_tiles = p.n / kTileRows;
  const int tail_rows = (int)(p.n - full_tiles * kTileRows);
  const int my_tiles = full_tiles > (int64_t)blockIdx.x ? (int)((full_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;
  const bool tail_mine = tail_rows > 0 && (int)(full_tiles % gridDim.x) == (int)blockIdx.x;
  const int T = my_tiles + (tail_mine ? 1 : 0);    // the ragged tail is this CTA's last tile
  const bool prof = L.prof != nullptr && blockIdx.x == 0;
  long long pw0 = 0, pw1 = 0, pw2 = 0, pcnt = 0;     // profile: cycles in wait sites 0/1, refine cycles, refined rows
  const bool do_reduce = p.do_reduce == 1;
  const int xx_mode = p.xx_mode;
  const bool na2 = L.na == 2;

  // ------------------------------------------------------------------ one-time set-up
  if (tid == 0) {
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bar_off);
    for (int s = 0; s < S; ++s) {
      mbar_init(bars + s, 1);
      mbar_init(bars + kMaxStages + s, do_reduce ? R : 4);
      mbar_init(bars + 2 * kMaxStages + s, 4);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(bars + 3 * kMaxStages + b, CW);                // afull
      mbar_init(bars + 3 * kMaxStages + 2 + b, 4);             // aempty
      mbar_init(bars + 3 * kMaxStages + 4 + b, 1);             // tfull
      mbar_init(bars + 3 * kMaxStages + 6 + b, 4);             // tempty
    }
    mbar_fence_init();
  }
  if (warp == kTpProdWarp) {   // the producer warp owns the TMEM allocation
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(sb + L.tmem_off), "n"(TCOLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // zero: the A operands, B, the accumulators, counters
  for (int i = tid; i < (L.acc_off - L.a_off) / 16; i += kTpThreads) reinterpret_cast<uint4*>(smem + L.a_off)[i] = make_uint4(0, 0, 0, 0);
  for (int i = tid; i < (L.nsets * KP * L.ald) / 2; i += kTpThreads) reinterpret_cast<uint4*>(smem + L.acc_off)[i] = make_uint4(0, 0, 0, 0);
  for (int i = tid; i < KP; i += kTpThreads) cnt_s[i] = 0u;
  if (L.cs_off != 0) for (int i = tid; i < k * d; i += kTpThreads) reinterpret_cast<double*>(smem + L.cs_off)[i] = p.C[i];
  if (tid < d8 + 8) { double s = 0; if (tid < d) for (int j = 0; j < k; ++j) s += p.C[(size_t)j * d + tid]; mu_s[tid] = tid < d ? s / k : 0.0; }   // mu = mean of the centres
  __syncthreads();
  double* h_s = red_s + 2 * kTpMaxWarps;   // [KP] |c - mu|^2, later the centre-update scratch
  if (tid < KP) {
    double s = 0, sc = 0;
    if (tid < k) for (int f = 0; f < d; ++f) { const double c = p.C[(size_t)tid * d + f], cc = c - mu_s[f]; s = fma(c, c, s); sc = fma(cc, cc, sc); }
    cn_s[tid] = tid < k ? -0.5 * s : -__longlong_as_double(0x7ff0000000000000LL);
    h_s[tid] = tid < k ? sc : 0.0;
  }
  // B operand: tf32(c - mu) in the canonical K-major layout, then the constant -0.5|c - mu|^2 as a hi/mid/lo tf32 split in
  // K columns d8, d8+1, d8+2 (padding centres get -1e30 there: never the best, never a candidate)
  for (int i = tid; i < KP * (d8 / 4); i += kTpThreads) {
    const int j = i % KP, c = i / KP;
    uint32_t vv[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) { const int f = 4 * c + e; vv[e] = (j < k && f < d) ? f32_to_tf32_rn((float)(p.C[(size_t)j * d + f] - mu_s[f])) : 0u; }
    *reinterpret_cast<uint4*>(smem + L.b_off + c * LBO_B + (j >> 3) * 128 + (j & 7) * 16) = make_uint4(vv[0], vv[1], vv[2], vv[3]);
  }
  // A operands: the column of ones that multiplies the constant (written once; the converters never touch it)
  for (int i = tid; i < L.na * kTileRows; i += kTpThreads) {
    const int b = i / kTileRows, r = i % kTileRows;
    const uint32_t one = __float_as_uint(1.0f);
    *reinterpret_cast<uint4*>(smem + L.a_off + b * L.a_bytes + (d8 / 4) * LBO_A + r * 16) = make_uint4(one, one, one, 0u);
  }
  __syncthreads();
  if (tid < KP) {
    const double h = tid < k ? -0.5 * h_s[tid] : -1e30;
    const float hi = __uint_as_float(f32_to_tf32_rn((float)h));
    const double r1 = h - (double)hi;
    const float mid = __uint_as_float(f32_to_tf32_rn((float)r1));
    const double r2 = r1 - (double)mid;
    const float lo = __uint_as_float(f32_to_tf32_rn((float)r2));
    *reinterpret_cast<uint4*>(smem + L.b_off + (d8 / 4) * LBO_B + (tid >> 3) * 128 + (tid & 7) * 16) =
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

  double cost = 0, xxs = 0;
  const int P = k * d + k + 1;
  const int SL = P + 1 + k;
  double* out = p.cta_partials + (size_t)blockIdx.x * SL;       // this CTA's slot of the scratch
  const long long pt0 = prof ? clock64() : 0;

  if (warp == kTpProdWarp) {
    // ===================================================================== producer
    const uint32_t cbytes = (uint32_t)kChunkRows * row_bytes;
    const size_t tile_gbytes = (size_t)kTileRows * row_bytes;
    const unsigned char* src = reinterpret_cast<const unsigned char*>(p.X) + (size_t)blockIdx.x * tile_gbytes;
    const size_t src_step = (size_t)gridDim.x * tile_gbytes;
    int s = 0; uint32_t ph = 0;
    for (int i = 0; i < T; ++i) {
      if (lane == 0) tp_wait_p(b_sempty + 8u * (uint32_t)s, ph ^ 1u, 7, prof, pw0);
      __syncwarp();
      const uint32_t bar = b_full + 8u * (uint32_t)s, dst = a_x + (uint32_t)s * sbytes;
      if (i < my_tiles) {
        if (lane == 0) {
          mbar_expect_tx_a(bar, cbytes * kChunks);
#pragma unroll
          for (int c = 0; c < kChunks; ++c) bulk_g2s_a(dst + c * cstride, src + (size_t)c * cbytes, cbytes, bar);
        }
        src += src_step;
      } else {
        // ragged tail (< 128 rows): plain loads by the whole warp, zero filled; rows_valid masks the rest downstream
        const double* tsrc = p.X + (size_t)full_tiles * kTileRows * d;
        for (int e = lane; e < kTileRows * d; e += 32) {
          const int tr = e / d, f = e - tr * d;
          sts_f64(dst + (uint32_t)(tr >> 4) * cstride + (uint32_t)((tr & 15) * d + f) * 8u, tr < tail_rows ? tsrc[(size_t)tr * d + f] : 0.0);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive_a(bar);
      }
      if (++s == S) { s = 0; ph ^= 1u; }
    }
  } else if (warp == kTpMmaWarp) {
    // ===================================================================== MMA issuer
    if (lane == 0) {
      const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(KP >> 3) << 17) | ((uint32_t)(kTileRows >> 4) << 24);
      const int nmma = L.kchunks / 2;                            // one UMMA consumes K = 8 tf32 = two 16-byte chunks
      for (int i = 0; i < T; ++i) {
        const uint32_t b = (uint32_t)(i & 1), ph = (uint32_t)((i >> 1) & 1);          // TMEM buffer / phase
        const uint32_t ab = na2 ? b : 0u, aph = na2 ? ph : (uint32_t)(i & 1);          // operand buffer / phase
        tp_wait_p(b_afull + 8u * ab, aph, 1, prof, pw0);
        tp_wait_p(b_tempty + 8u * b, ph ^ 1u, 2, prof, pw1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t abase = a_A + ab * (uint32_t)L.a_bytes, dcol = tmem + b * (uint32_t)TSTRIDE;
        for (int ks = 0; ks < nmma; ++ks) {
          const uint64_t da = tp_desc_kmajor(abase + (uint32_t)(2 * ks) * LBO_A, LBO_A, 128u);
          const uint64_t db = tp_desc_kmajor(a_B + (uint32_t)(2 * ks) * LBO_B, LBO_B, 128u);
          const uint32_t accf = ks > 0 ? 1u : 0u;
          asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n" ::"r"(dcol),
                       "l"(da), "l"(db), "r"(idesc), "r"(accf)
                       : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(b_tfull + 8u * b) : "memory");
      }
    }
  } else if (warp >= kTpConvWarp0) {
    // ===================================================================== converters
    // lane = (row of the group, feature pair); a warp converts chunks cw, cw + CW, ..., C_G rows per instruction.  The tensor
    // core TRUNCATES its fp32 operands to tf32 (it ignores the 13 low mantissa bits: tools/tf32_rounding_probe.cu), so half a
    // tf32 ulp is added to the bit pattern first -- one integer add per value makes the truncation a round-to-nearest
    constexpr int C_LPR = tp_lanes_per_row(ND, CVEC), C_G = 32 / C_LPR, NIT = kChunkRows / C_G;
    const int cw = warp - kTpConvWarp0;
    const int subc = lane / C_LPR, fc = (lane % C_LPR) * CVEC;
    const bool act = fc < d;
    const float mu0 = act ? (float)mu_s[fc] : 0.f, mu1 = (CVEC == 2 && act) ? (float)mu_s[fc + 1] : 0.f;
    const uint32_t dst_lane = (uint32_t)(fc >> 2) * LBO_A + (uint32_t)(fc & 3) * 4u + (uint32_t)subc * 16u;
    const uint32_t src_lane = (uint32_t)subc * row_bytes + (uint32_t)fc * 8u;
    const uint32_t src_it = (uint32_t)C_G * row_bytes;
    const bool need_xx = do_reduce && xx_mode != 2;
    int s = 0; uint32_t sph = 0;
    for (int i = 0; i < T; ++i) {
      const uint32_t b = na2 ? (uint32_t)(i & 1) : 0u, ph = na2 ? (uint32_t)((i >> 1) & 1) : (uint32_t)(i & 1);
      tp_wait_p(b_full + 8u * (uint32_t)s, sph, 3, prof, pw0);
      tp_wait_p(b_aempty + 8u * b, ph ^ 1u, 4, prof, pw1);
      const uint32_t xs = a_x + (uint32_t)s * sbytes, ab = a_A + b * (uint32_t)L.a_bytes;
      if (act) {
#pragma unroll
        for (int cq = 0; cq < kChunks / CW; ++cq) {
          const int cc = cw + cq * CW;
          const uint32_t src0 = xs + (uint32_t)cc * cstride + src_lane;
          const uint32_t dst0 = ab + (uint32_t)cc * 256u + dst_lane;
          constexpr int NB = NIT > 4 ? 4 : NIT;              // rows in flight per lane (register budget)
#pragma unroll
          for (int it0 = 0; it0 < NIT; it0 += NB) {
            double v0[NB], v1[NB];
#pragma unroll
            for (int it = 0; it < NB; ++it) {
              v1[it] = 0.0;
              if (CVEC == 2) lds_v2f64(src0 + (uint32_t)(it0 + it) * src_it, v0[it], v1[it]); else v0[it] = lds_f64(src0 + (uint32_t)(it0 + it) * src_it);
            }
            if (need_xx) {
#pragma unroll
              for (int it = 0; it < NB; ++it) { xxs = fma(v0[it], v0[it], xxs); if (CVEC == 2) xxs = fma(v1[it], v1[it], xxs); }
            }
#pragma unroll
            for (int it = 0; it < NB; ++it) {
              const uint32_t t0 = __float_as_uint((float)v0[it] - mu0) + 0x1000u;
              if (CVEC == 2) {
                const uint32_t t1 = __float_as_uint((float)v1[it] - mu1) + 0x1000u;
                asm volatile("st.shared.v2.u32 [%0], {%1,%2};" ::"r"(dst0 + (uint32_t)((it0 + it) * C_G * 16)), "r"(t0), "r"(t1) : "memory");
              } else sts_u32(dst0 + (uint32_t)((it0 + it) * C_G * 16), t0);
            }
          }
        }
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");       // conversion stores -> visible to the tensor core
      __syncwarp();
      if (lane == 0) mbar_arrive_a(b_afull + 8u * b);
      if (++s == S) { s = 0; sph ^= 1u; }
    }
  } else if (warp < 4 * SG) {
    // ===================================================================== screen warps
    const int grp = warp >> 2, wsub = warp & 3;
    const int row = 32 * wsub + lane;                              // tile row of this thread = TMEM lane
    const uint32_t tlane = ((uint32_t)(32 * wsub)) << 16;
    const float kapv = 1.04f * 0.0009765625f * vmax;               // kappa * max|c - mu|
    // eta * max 0.5|c - mu|^2 (eta = 2^-16), plus the fp32 rounding of x and mu ahead of the subtraction: 2^-24 (|x| + |mu|) per
    // feature, |x| <= |x - mu| + |mu| (the |x - mu| part is inside kappa's slack)
    const float etab = 0.5f * 1.52587890625e-05f * vmax * vmax + 2.5e-7f * munorm * vmax;
    constexpr int DCH = 2 * ND;                                    // 16-byte chunks of the converted row
    const bool conv2 = (d & 1) == 0;
    const int rlane = conv2 ? 2 * lane : lane;                     // refine: first feature of this lane
    const bool ract = rlane < d;
    constexpr int RLANES = tp_dclass(ND) / CVEC > 32 ? 32 : tp_dclass(ND) / CVEC;   // lanes that hold features in the refine (power of two)
    const double* Cr = L.cs_off != 0 ? reinterpret_cast<const double*>(smem + L.cs_off) : p.C;   // centres for the exact re-scoring
    int s = grp % S;
    for (int i = grp; i < T; i += SG) {
      const uint32_t b = (uint32_t)(i & 1), ph = (uint32_t)((i >> 1) & 1);
      const int rows_valid = i < my_tiles ? kTileRows : tail_rows;
      const int64_t r0 = i < my_tiles ? ((int64_t)blockIdx.x + (int64_t)i * gridDim.x) * kTileRows : full_tiles * kTileRows;
      const uint32_t lab_b = a_lab + (uint32_t)s * (kTileRows * 4);
      tp_wait_p(b_tfull + 8u * b, ph, 5, prof, pw0);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      // |x - mu|^2 of this row from the tf32 operand the tensor core just consumed
      const uint32_t abuf = na2 ? b : 0u;
      const uint32_t arow = a_A + abuf * (uint32_t)L.a_bytes + (uint32_t)row * 16u;
      float xa = 0.f, xb = 0.f;
#pragma unroll
      for (int c = 0; c < DCH; ++c) {
        float4 a;
        asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w) : "r"(arow + (uint32_t)c * LBO_A));
        xa = fmaf(a.x, a.x, xa); xb = fmaf(a.y, a.y, xb); xa = fmaf(a.z, a.z, xa); xb = fmaf(a.w, a.w, xb);
      }
      const float xn2 = xa + xb;
      __syncwarp();
      if (lane == 0) mbar_arrive_a(b_aempty + 8u * abuf);        // the operand buffer can take the next conversion
      // sweep 1: best and second best approximate score, centre index in the low mantissa bits
      const uint32_t tbase = tmem + tlane + b * (uint32_t)TSTRIDE;
      const float ninf = -__int_as_float(0x7f800000);
      float m1 = ninf, m2 = ninf;
#pragma unroll
      for (int ch = 0; ch < KP / 16; ++ch) {
        uint32_t v[16];
        tp_tmem_ld16(tbase + (uint32_t)(ch * 16), v);
        tp_tmem_wait_ld();
#pragma unroll
        for (int c = 0; c < 16; ++c) {
          const float pv = __uint_as_float((v[c] & ~IDXMASK) | (uint32_t)(ch * 16 + c));
          m2 = fmaxf(m2, fminf(m1, pv));
          m1 = fmaxf(m1, pv);
        }
      }
      const bool valid = row < rows_valid;
      const float eps = fmaf(kapv, sqrtf(xn2) * 1.001f, etab) + 6e-5f * fabsf(m1);
      const float thr = m1 - 2.0f * eps;
      const bool amb = valid && !(m2 < thr);
      const int j1 = (int)(__float_as_uint(m1) & IDXMASK);
      unsigned ambmask = __ballot_sync(0xffffffffu, amb);
      uint32_t w[NW];
#pragma unroll
      for (int q = 0; q < NW; ++q) w[q] = 0u;
      if (ambmask) {
        // sweep 2 (only when some row of the warp is undecided): every centre within 2 eps of the best is a candidate
        const long long ts0 = prof ? clock64() : 0;
#pragma unroll
        for (int ch = 0; ch < KP / 16; ++ch) {
          uint32_t v[16];
          tp_tmem_ld16(tbase + (uint32_t)(ch * 16), v);
          tp_tmem_wait_ld();
#pragma unroll
          for (int c = 0; c < 16; ++c) w[ch >> 1] |= (__uint_as_float(v[c]) >= thr) ? (1u << ((ch & 1) * 16 + c)) : 0u;
        }
        if (prof) pw1 += clock64() - ts0;
      }
      // the TMEM buffer is free again
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive_a(b_tempty + 8u * b);
      if (!amb) {
        // rows past the end of the table are zero filled: label 0 adds nothing to the sums and is never counted or stored
        sts_u32(lab_b + 4u * (uint32_t)row, valid ? (uint32_t)j1 : 0u);
        if (valid) {
          if (p.labels_out != nullptr) p.labels_out[r0 + row] = j1;
          if (do_reduce) asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(a_cnt + 4u * (uint32_t)j1) : "memory");
        }
      }
      if (ambmask) {
        // exact re-scoring of this warp's undecided rows, one at a time, lane = feature (pair), ascending centre index
        const long long tr0 = prof ? clock64() : 0;
        const uint32_t xs = a_x + (uint32_t)s * sbytes;
        while (ambmask) {
          const int src = __ffs(ambmask) - 1; ambmask &= ambmask - 1;
          const int rr = 32 * wsub + src;
          uint32_t cm[NW];
#pragma unroll
          for (int q = 0; q < NW; ++q) cm[q] = __shfl_sync(0xffffffffu, w[q], src);
          const uint32_t ra = xs + (uint32_t)(rr >> 4) * cstride + (uint32_t)(rr & 15) * row_bytes + (uint32_t)rlane * 8u;
          double x0 = 0.0, x1 = 0.0;
          if (ract) { if (conv2) lds_v2f64(ra, x0, x1); else x0 = lds_f64(ra); }
          double best = -__longlong_as_double(0x7ff0000000000000LL); int bj = -1;
          // candidates in ascending index, four at a time: all centre loads first, then the four shuffle reductions interleaved
#pragma unroll
          for (int q = 0; q < NW; ++q) {
            uint32_t msk = cm[q];
            while (msk) {
              int jj[4]; double pp[4];
#pragma unroll
              for (int u = 0; u < 4; ++u) { jj[u] = msk ? 32 * q + __ffs(msk) - 1 : -1; msk &= msk - 1; }
#pragma unroll
              for (int u = 0; u < 4; ++u) {
                double c0 = 0.0, c1 = 0.0;
                if (ract && jj[u] >= 0) {
                  const double* cr = Cr + (size_t)jj[u] * d + rlane;
                  c0 = cr[0]; if (conv2) c1 = cr[1];
                }
                pp[u] = fma(x0, c0, x1 * c1);
              }
              // only the lanes below RLANES hold features: the butterfly starts at RLANES / 2 (lane 0 ends with the full sum)
#pragma unroll
              for (int mk = RLANES / 2; mk >= 1; mk >>= 1) {
#pragma unroll
                for (int u = 0; u < 4; ++u) pp[u] += shfl_xor_d(pp[u], mk);
              }
#pragma unroll
              for (int u = 0; u < 4; ++u)
                if (jj[u] >= 0) { const double sc = pp[u] + cn_s[jj[u]]; if (sc > best) { best = sc; bj = jj[u]; } }
            }
          }
          if (lane == 0) {
            sts_u32(lab_b + 4u * (uint32_t)rr, (uint32_t)bj);
            if (p.labels_out != nullptr) p.labels_out[r0 + rr] = bj;
            if (do_reduce) asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(a_cnt + 4u * (uint32_t)bj) : "memory");
          }
          ++pcnt;
        }
        if (prof) pw2 += clock64() - tr0;
      }
      __syncwarp();
      if (lane == 0) {
        mbar_arrive_a(b_lfull + 8u * (uint32_t)s);
        if (!do_reduce) mbar_arrive_a(b_sempty + 8u * (uint32_t)s);
      }
      s += SG; if (s >= S) s -= S;
    }
  } else {
    // ===================================================================== reduce warps
    const int ridx = warp - 4 * SG;
    if (do_reduce) {
      int s = 0; uint32_t sph = 0;
      if constexpr (RMODE == 3) {
        // S[k, d] += onehot(labels)^T X on the fp64 tensor pipe.  The warp takes sub-tiles i0 = ridx and i1 = ridx + 8: row i of
        // every 16-row chunk, so the four rows of a k-step sit in four different chunks (32 bytes apart modulo 128: the
        // fragment loads are conflict free).  racc[(mt * ND + nd) * 2 + e] = S[8 mt + g][8 nd + 2 t + e].
        const int g = lane >> 2, t = lane & 3;
        double racc[2 * ND * 2];
#pragma unroll
        for (int q = 0; q < 2 * ND * 2; ++q) racc[q] = 0.0;
        const uint32_t b_frag = (uint32_t)t * cstride + (uint32_t)g * 8u;
        const uint32_t l_frag = (uint32_t)(16 * t) * 4u;
        for (int i = 0; i < T; ++i) {
          tp_wait_p(b_lfull + 8u * (uint32_t)s, sph, 6, prof, pw0);
          const uint32_t xs = a_x + (uint32_t)s * sbytes + b_frag, lab_b = a_lab + (uint32_t)s * (kTileRows * 4) + l_frag;
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {
            const int isub = (ks < 2) ? ridx : ridx + 8;           // sub-tile: row isub of chunks 4 (ks & 1) + t
            const int labt = (int)lds_u32(lab_b + (uint32_t)(64 * (ks & 1) + isub) * 4u);
            const double a0 = labt == g ? 1.0 : 0.0, a1 = labt == g + 8 ? 1.0 : 0.0;
            const uint32_t xb = xs + (uint32_t)(4 * (ks & 1)) * cstride + (uint32_t)isub * row_bytes;
#pragma unroll
            for (int nd = 0; nd < ND; ++nd) {
              const bool ok = (nd < ND - 1) || (8 * nd + g < d);
              const double b2 = ok ? lds_f64(xb + 64u * nd) : 0.0;
              dmma884(racc[(0 * ND + nd) * 2], racc[(0 * ND + nd) * 2 + 1], a0, b2);
              dmma884(racc[(1 * ND + nd) * 2], racc[(1 * ND + nd) * 2 + 1], a1, b2);
            }
          }
          __syncwarp();
          if (lane == 0) mbar_arrive_a(b_sempty + 8u * (uint32_t)s);
          if (++s == S) { s = 0; sph ^= 1u; }
        }
        // every tile is reduced, so the ring is idle: the reduce warps park their fragments there, slab r = [KP][8 ND] doubles
        named_bar_sync(1, R * 32);
        {
          double* slab = reinterpret_cast<double*>(smem + L.x_off) + (size_t)ridx * KP * (8 * ND);
#pragma unroll
          for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nd = 0; nd < ND; ++nd) {
              double* c2 = slab + (8 * mt + g) * (8 * ND) + 8 * nd + 2 * t;
              c2[0] = racc[(mt * ND + nd) * 2]; c2[1] = racc[(mt * ND + nd) * 2 + 1];
            }
        }
      } else if constexpr (RMODE == 2) {
        // one shared set (row j at j * 512 bytes): warp r owns the clusters j = r mod R, lane = feature pair; its rows come from
        // one ballot per 32 rows, two rows in flight -- when both carry the same label the second one stores the sum of both
        const bool act = 2 * lane < d;
        const uint32_t my_acc = a_acc + (uint32_t)lane * 16u;
        for (int i = 0; i < T; ++i) {
          tp_wait_p(b_lfull + 8u * (uint32_t)s, sph, 6, prof, pw0);
          const uint32_t xs = a_x + (uint32_t)s * sbytes + (uint32_t)lane * 16u, lab_b = a_lab + (uint32_t)s * (kTileRows * 4);
#pragma unroll 1
          for (int gq = 0; gq < kTileRows / 32; ++gq) {
            const int lab = (int)lds_u32(lab_b + 4u * (uint32_t)(gq * 32 + lane));
            unsigned mask = __ballot_sync(0xffffffffu, (lab & (R - 1)) == ridx);
            const uint32_t xg = xs + (uint32_t)(2 * gq) * cstride;       // rows 32 gq .. 32 gq + 31 = chunks 2 gq, 2 gq + 1
            while (mask) {
              const int q0 = __ffs(mask) - 1; mask &= mask - 1;
              const bool two = mask != 0;
              const int q1 = two ? __ffs(mask) - 1 : q0; mask &= mask - 1;
              const int l0 = __shfl_sync(0xffffffffu, lab, q0), l1 = __shfl_sync(0xffffffffu, lab, q1);
              if (act) {
                const uint32_t aa0 = my_acc + (uint32_t)l0 * 512u, aa1 = my_acc + (uint32_t)l1 * 512u;
                double v00, v01, v10, v11, s00, s01, s10, s11;
                lds_v2f64(xg + (uint32_t)(q0 >> 4) * cstride + (uint32_t)(q0 & 15) * row_bytes, v00, v01);
                lds_v2f64(xg + (uint32_t)(q1 >> 4) * cstride + (uint32_t)(q1 & 15) * row_bytes, v10, v11);
                lds_v2f64(aa0, s00, s01); lds_v2f64(aa1, s10, s11);
                const bool same = l0 == l1;                      // also when the second row is the first one again (odd count)
                if (!same) sts_v2f64(aa0, s00 + v00, s01 + v01);
                const double w0 = two ? v10 : 0.0, w1 = two ? v11 : 0.0;
                sts_v2f64(aa1, s10 + (same ? v00 + w0 : w0), s11 + (same ? v01 + w1 : w1));
              }
            }
          }
          __syncwarp();
          if (lane == 0) mbar_arrive_a(b_sempty + 8u * (uint32_t)s);
          if (++s == S) { s = 0; sph ^= 1u; }
        }
      } else {
        // private sets in shared memory: this warp's rows are the groups gi = ridx, ridx + R, ... of R_G consecutive rows;
        // sub-row `sub` of a group accumulates into set ridx * R_G + sub, so no two lanes ever meet on an address
        constexpr int R_LPR = tp_lanes_per_row(ND, 1), R_G = 32 / R_LPR, NG = kTileRows / R_G;
        const int sub = lane / R_LPR, f = lane % R_LPR;
        const bool act = f < d;
        const uint32_t acc_row = (uint32_t)L.ald * 8u;
        const uint32_t my_acc = a_acc + (uint32_t)(ridx * R_G + sub) * (uint32_t)KP * acc_row + (uint32_t)f * 8u;
        for (int i = 0; i < T; ++i) {
          tp_wait_p(b_lfull + 8u * (uint32_t)s, sph, 6, prof, pw0);
          const uint32_t xs = a_x + (uint32_t)s * sbytes + (uint32_t)f * 8u, lab_b = a_lab + (uint32_t)s * (kTileRows * 4);
          if (act) {
#pragma unroll
            for (int gi0 = 0; gi0 < NG; gi0 += 2 * R) {
              // two groups in flight; if a sub-row meets the same accumulator row in both, the second carries the sum
              const int rA = (gi0 + ridx) * R_G + sub, rB = (gi0 + R + ridx) * R_G + sub;
              const int labA = (int)lds_u32(lab_b + 4u * (uint32_t)rA), labB = (int)lds_u32(lab_b + 4u * (uint32_t)rB);
              const uint32_t xA = xs + (uint32_t)(rA >> 4) * cstride + (uint32_t)(rA & 15) * row_bytes;
              const uint32_t xB = xs + (uint32_t)(rB >> 4) * cstride + (uint32_t)(rB & 15) * row_bytes;
              const uint32_t aA = my_acc + (uint32_t)labA * acc_row, aB = my_acc + (uint32_t)labB * acc_row;
              const bool same = labA == labB;
              const double vA = lds_f64(xA), vB = lds_f64(xB);
              const double sA = lds_f64(aA), sB = lds_f64(aB);
              if (!same) sts_f64(aA, sA + vA);
              sts_f64(aB, sB + (same ? vA + vB : vB));
            }
          }
          __syncwarp();
          if (lane == 0) mbar_arrive_a(b_sempty + 8u * (uint32_t)s);
          if (++s == S) { s = 0; sph ^= 1u; }
        }
      }
    }
  }

  if (prof && lane == 0) {
    unsigned long long* o = L.prof + warp * 8;
    o[0] = (unsigned long long)(clock64() - pt0); o[1] = (unsigned long long)pw0; o[2] = (unsigned long long)pw1; o[3] = (unsigned long long)pw2;
    o[4] = 0ull; o[5] = (unsigned long long)pcnt; o[6] = (unsigned long long)T; o[7] = 0ull;
  }
  // ===================================================================== CTA-level combine (all warps)
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == kTpProdWarp) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(TCOLS) : "memory");
  if (!p.do_reduce) return;

  if constexpr (RMODE == 3) {
    const double* slab = reinterpret_cast<const double*>(smem + L.x_off);
    constexpr int LDS_ = 8 * ND;
    for (int i = tid; i < k * d; i += kTpThreads) {
      const int j = i / d, f = i - j * d;
      double sacc = 0;
#pragma unroll
      for (int w = 0; w < R; ++w) sacc += slab[(size_t)w * KP * LDS_ + j * LDS_ + f];
      out[i] = sacc;
      cost = fma(sacc, p.C[i], cost);                           // sum_j S_j . c_j
    }
  } else {
    const double* acc0 = reinterpret_cast<const double*>(smem + L.acc_off);
    const int ald = L.ald, setn = KP * L.ald, nsets = L.nsets;
    for (int i = tid; i < k * d; i += kTpThreads) {
      const int j = i / d, f = i - j * d;
      double sacc = 0;
      for (int w = 0; w < nsets; ++w) sacc += acc0[w * setn + j * ald + f];
      out[i] = sacc;
      cost = fma(sacc, p.C[i], cost);
    }
  }
  if (tid < k) { const double nj = (double)cnt_s[tid]; out[(size_t)k * d + tid] = nj; cost = fma(nj, cn_s[tid], cost); }
#pragma unroll
  for (int msk = 16; msk >= 1; msk >>= 1) { cost += shfl_xor_d(cost, msk); xxs += shfl_xor_d(xxs, msk); }
  if (lane == 0) { red_s[warp] = cost; red_s[kTpMaxWarps + warp] = xxs; }
  __syncthreads();
  if (tid == 0) {
    double c = 0, x2 = 0;
#pragma unroll
    for (int wv = 0; wv < kTpWarps; ++wv) { c += red_s[wv]; x2 += red_s[kTpMaxWarps + wv]; }
    out[P - 1] = c; out[P] = x2;                                 // [sum of winner scores][sum |x|^2 of this CTA's rows]
  }
  // ===================================================================== last CTA folds all partials
  __threadfence();
  __syncthreads();
  __shared__ unsigned int s_ticket;
  if (tid == 0) s_ticket = atomicAdd(p.counter, 1u);
  __syncthreads();
  if (s_ticket == gridDim.x - 1) {
    __threadfence();
    for (int i = tid; i < P + 1; i += kTpThreads) {
      if (i == P - 1) continue;
      double sum = 0;
      sum = ordered_sum_ldcg(p.cta_partials + i, (size_t)SL, gridDim.x);
      if (i < P) p.partial_out[i] = sum; else red_s[0] = sum;    // i == P: sum |x|^2 over the whole pass
    }
    __syncthreads();
    if (tid == 0) {
      double ssc = 0;
      ssc = ordered_sum_ldcg(p.cta_partials + (P - 1), (size_t)SL, gridDim.x);
      const double xx = xx_mode == 2 ? *p.xnorm2_sum : red_s[0];
      if (xx_mode == 1) *p.xnorm2_sum = xx;
      p.partial_out[P - 1] = fmax(fma(-2.0, ssc, xx), 0.0);      // cost = sum |x|^2 - 2 sum score
      *p.counter = 0u;
    }
    bool peer_ok = true;
    if (p.peer_bufs != nullptr) {                                // ranks > 1: sum the vector over NVLink peer memory right here
      __threadfence_block();
      __syncthreads();
      peer_ok = peer_allreduce(p, p.partial_out, P, tid, kTpThreads, [] { __syncthreads(); });
    }
    if (p.C_update != nullptr && peer_ok) {                      // fused driver step
      __threadfence_block();
      __syncthreads();
      update_centers_body(p.partial_out, k, d, p.C_update, p.tol2, p.state_rw, red_s + 2 * kTpMaxWarps, tid);
    }
  }
}

}  // namespace hkm

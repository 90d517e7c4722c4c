// Lloyd data pass, warp-group variant (8 < k <= 32, d <= 32): tcgen05 screening + fp64 refinement as in hkm_lloyd_tp.cuh, but
// without specialised roles.  The CTA (one per SM) is G independent groups of four warps; a group takes every G-th tile of the
// CTA and carries it through all steps itself, each warp working on its own 32 rows (= its TMEM lane quarter):
//
//   wait for the warp's 32 rows (bulk-async copies it issued itself)           -- mbarrier, one waiter
//   convert them: fp64 -> (x - mu) rounded to tf32 -> rows 32q.. of the group's K-major UMMA operand
//   group barrier (bar.sync on 128 threads); one thread issues the tcgen05.mma chain + commit
//   |x - mu|^2 of the warp's rows from the operand (while the tensor core runs)
//   wait for the commit; sweep the 32 x k_pad accumulator quarter once (lane = row), decide the row or re-score it exactly
//   per-cluster sums of the warp's 32 rows: S += onehot(labels)^T X on the fp64 tensor pipe (DMMA.8x8x4), labels passed
//   between lanes by shuffle, accumulators in registers for the whole pass
//   re-issue the bulk copies for the tile that next uses this quarter of the ring stage
//
// So per tile a warp meets one group barrier and two mbarriers and nothing else: no producer, converter, issuer or reduce warps
// to hand tiles to, no labels through shared memory, no empty barriers (a quarter stage is refilled by the warp that just
// finished with it).  Latency is hidden by the other groups, which sit at other steps of other tiles.  Ring stages are shared by
// the groups (tile i lives in stage i mod S), so all of shared memory beyond the operands holds tiles in flight.
//
// A stage quarter is waited for by a different group on every lap, so the parity of its mbarrier alone cannot tell "my tile has
// landed" from "the tile of the lap before has not even been fetched yet" (a group that runs ahead would pass the wait on the
// stale phase).  Each quarter therefore carries the index of the tile last requested into it (seq, written by the requesting
// lane ahead of the expect_tx); a warp first sees its own tile index there, then waits on the parity.
//
// Arithmetic (screening bound, tie rule, refinement, cost identity) is that of hkm_lloyd_tp.cuh; summation order is fixed
// (row order within a warp, warp order in the combine, CTA order in the fold).
#pragma once
#include "hkm_lloyd_tp.cuh"

namespace hkm {

constexpr int kWgMaxGroups = 6;
constexpr int kWgMaxWarps = 4 * kWgMaxGroups;
constexpr int kWgMaxStages = 16;
constexpr int kWgSubRows = 8;                  // rows per bulk copy; sub-chunks are padded so DMMA fragment loads spread over the banks
constexpr int kWgSubPad = 32;

struct WgPlan {
  int groups, stages, nd, kchunks, lbo_a, lbo_b, a_bytes, sub_stride, tile_bytes, tcols;
  int x_off, a_off, one_off, b_off, cn_off, mu_off, cnt_off, red_off, bar_off, seq_off, tmem_off, cs_off, total;
  unsigned long long* prof;   // developer profile: [warp][8] cycle counters of CTA 0, or null
};

__host__ __device__ constexpr inline WgPlan wg_plan(int d, int kp, int groups, int stages) {
  WgPlan L{};
  L.groups = groups; L.stages = stages; L.nd = (d + 7) / 8;
  const int cvec = (d & 1) ? 1 : 2;
  const int c_g = 32 / tp_lanes_per_row(L.nd, cvec);
  const int d8 = (d + 7) & ~7;
  L.kchunks = (d8 + 8) / 4;
  const int pad_a = 8 * c_g > 16 ? 8 * c_g : 16;
  L.lbo_a = (kTileRows / 8) * 128 + pad_a;
  L.lbo_b = (kp / 8) * 128 + 16;
  L.a_bytes = ((L.kchunks - 2) * L.lbo_a + 127) & ~127;     // data chunks; the constant pair (ones, zeros) is shared by the groups
  L.sub_stride = kWgSubRows * d * 8 + kWgSubPad;
  L.tile_bytes = (kTileRows / kWgSubRows) * L.sub_stride;
  int tc = 32; while (tc < groups * 32) tc *= 2;
  L.tcols = tc;
  int o = 0;
  L.x_off = o; o += stages * L.tile_bytes; o = (o + 127) & ~127;
  L.a_off = o; o += groups * L.a_bytes;
  L.one_off = o; o += (2 * L.lbo_a + 127) & ~127;
  L.b_off = o; o += (L.kchunks * L.lbo_b + 127) & ~127;
  L.cn_off = o; o += kp * 8;
  L.mu_off = o; o += (d8 + 8) * 8;
  L.cnt_off = o; o += kp * 4;
  L.red_off = o; o += (2 * kWgMaxWarps + 64) * 8;
  L.bar_off = o; o += (4 * kWgMaxStages + kWgMaxGroups) * 8;
  L.seq_off = o; o += 4 * kWgMaxStages * 4;
  L.tmem_off = o; o += 16;
  if (kp * d * 8 <= 8 * 1024) { L.cs_off = o; o += kp * d * 8; }
  L.total = o;
  return L;
}

// Groups and ring depth for a shape; stages == 0 => the shape does not fit.  At least two tiles (and 40 KB) beyond the ones the
// groups work on must be in flight, and the idle ring must hold the per-warp accumulator slabs of the final combine.
// ring depth for a given group count (0: does not fit)
__host__ __device__ constexpr inline int wg_stages(int d, int kp, int g) {
  const int nd = (d + 7) / 8;
  if (g < 1 || g > kWgMaxGroups) return 0;
  const WgPlan z = wg_plan(d, kp, g, 0);
  int st = (227 * 1024 - 1024 - z.total) / z.tile_bytes;
  if (st > kWgMaxStages) st = kWgMaxStages;
  const int slab = 4 * g * kp * 8 * nd * 8;
  if (st < g + 2 || (st - g) * z.tile_bytes < 40 * 1024 || st * z.tile_bytes < slab) return 0;
  return st;
}
__host__ inline WgPlan wg_choose(int d, int kp, int force_groups = -1) {
  const int nd = (d + 7) / 8;
  int cand[4]; int nc = 0;
  if (force_groups > 0) cand[nc++] = force_groups;
  else if (d == 16 && kp == 16) { cand[nc++] = 6; cand[nc++] = 5; }                       // built with the shape folded in, 768 threads
  else if (nd <= 2 && kp == 16) { cand[nc++] = 5; cand[nc++] = 4; cand[nc++] = 3; }
  else if (nd <= 2) { cand[nc++] = 4; cand[nc++] = 3; }
  else if (nd == 3) { cand[nc++] = 4; cand[nc++] = 3; cand[nc++] = 2; }
  else { cand[nc++] = 3; cand[nc++] = 2; }
  for (int ci = 0; ci < nc; ++ci) {
    const int st = wg_stages(d, kp, cand[ci]);
    if (st > 0) return wg_plan(d, kp, cand[ci], st);
  }
  WgPlan none{}; return none;
}

// KP = padded cluster count (16, 32); CVEC = 2 (even d) or 1; ND = ceil(d/8) (1..4); MAXT = thread bound (register budget).
// DFIX / GFIX > 0: d and the group count are compile-time constants (the plan then folds into immediates: the hot shapes spend a
// third of their issue slots on address arithmetic otherwise); the host launches these only with the matching plan.
template <int KP, int CVEC, int ND, int MAXT, int DFIX = 0, int GFIX = 0>
__global__ void __launch_bounds__(MAXT, 1) lloyd_wg_kernel(const LloydParams p, const WgPlan Lrt) {
  if (p.skip != nullptr && *p.skip != 0) return;
  if (p.skip2 != nullptr && *p.skip2 != 0) return;
  constexpr int NW = (KP + 31) / 32;              // candidate mask words
  constexpr int MT = KP / 16;                     // 16-cluster row tiles of the one-hot product
  constexpr int TSTRIDE = 32;
  constexpr int IDXBITS = 6;
  constexpr uint32_t IDXMASK = (1u << IDXBITS) - 1u;
  static_assert(KP == 16 || KP == 32, "k_pad: 16 or 32");
  static_assert(ND >= 1 && ND <= 4, "d <= 32");

  static_assert(DFIX == 0 || ((DFIX + 7) / 8 == ND && ((DFIX & 1) ? 1 : 2) == CVEC && GFIX > 0 && GFIX * 128 <= MAXT), "fixed shape");
  WgPlan L = Lrt;
  if constexpr (DFIX > 0) { L = wg_plan(DFIX, KP, GFIX, wg_stages(DFIX, KP, GFIX)); L.prof = Lrt.prof; }
  extern __shared__ __align__(1024) unsigned char smem[];
  const int d = DFIX > 0 ? DFIX : p.d, k = p.k;
  const int S = L.stages, G = L.groups;
  const int nthreads = G * 128, nwarps = G * 4;
  const uint32_t sb = smem_u32(smem);
  const uint32_t a_x = sb + L.x_off, a_B = sb + L.b_off, a_cnt = sb + L.cnt_off;
  const uint32_t a_bar = sb + L.bar_off, a_seq = sb + L.seq_off;
  const uint32_t b_full = a_bar, b_tfull = a_bar + 8u * 4u * kWgMaxStages;
  double* cn_s = reinterpret_cast<double*>(smem + L.cn_off);
  double* mu_s = reinterpret_cast<double*>(smem + L.mu_off);
  double* red_s = reinterpret_cast<double*>(smem + L.red_off);
  unsigned int* cnt_s = reinterpret_cast<unsigned int*>(smem + L.cnt_off);

  // the warp index through a shuffle, so the compiler keeps everything derived from it (stage, barrier and copy addresses) in
  // uniform registers and the bulk copies are issued without a per-operand election loop
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
  const int grp = warp >> 2, q = warp & 3;
  const uint32_t row_bytes = (uint32_t)d * 8u;
  const uint32_t sub_stride = (uint32_t)L.sub_stride, tile_bytes = (uint32_t)L.tile_bytes;
  const int d8 = (d + 7) & ~7;
  const uint32_t LBO_A = (uint32_t)L.lbo_a, LBO_B = (uint32_t)L.lbo_b;
  const int64_t full_tiles = p.n / kTileRows;
  const int tail_rows = (int)(p.n - full_tiles * kTileRows);
  const int my_tiles = full_tiles > (int64_t)blockIdx.x ? (int)((full_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;
  const bool tail_mine = tail_rows > 0 && (int)(full_tiles % gridDim.x) == (int)blockIdx.x;
  const int T = my_tiles + (tail_mine ? 1 : 0);    // the ragged tail is this CTA's last tile
  const bool prof = L.prof != nullptr && blockIdx.x == 0;
  long long pw0 = 0, pw1 = 0, pw2 = 0, pcnt = 0;
  const bool do_reduce = p.do_reduce == 1;
  const int xx_mode = p.xx_mode;

  // ------------------------------------------------------------------ one-time set-up
  if (tid == 0) {
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bar_off);
    for (int s = 0; s < 4 * S; ++s) mbar_init(bars + s, 1);
    for (int g = 0; g < G; ++g) mbar_init(bars + 4 * kWgMaxStages + g, 1);
    mbar_fence_init();
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(sb + L.tmem_off), "r"((uint32_t)L.tcols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // zero: the A operands and B (one contiguous range), counters
  for (int i = tid; i < (L.cn_off - L.a_off) / 16; i += nthreads) reinterpret_cast<uint4*>(smem + L.a_off)[i] = make_uint4(0, 0, 0, 0);
  for (int i = tid; i < KP; i += nthreads) cnt_s[i] = 0u;
  for (int i = tid; i < 4 * S; i += nthreads) reinterpret_cast<uint32_t*>(smem + L.seq_off)[i] = (uint32_t)(i >> 2);   // stage st first holds tile st
  if (L.cs_off != 0) for (int i = tid; i < k * d; i += nthreads) reinterpret_cast<double*>(smem + L.cs_off)[i] = p.C[i];
  if (tid < d8 + 8) { double s = 0; if (tid < d) for (int j = 0; j < k; ++j) s += p.C[(size_t)j * d + tid]; mu_s[tid] = tid < d ? s / k : 0.0; }   // mu = mean of the centres
  __syncthreads();
  double* h_s = red_s + 2 * kWgMaxWarps;   // [KP] |c - mu|^2, later the centre-update scratch
  if (tid < KP) {
    double s = 0, sc = 0;
    if (tid < k) for (int f = 0; f < d; ++f) { const double c = p.C[(size_t)tid * d + f], cc = c - mu_s[f]; s = fma(c, c, s); sc = fma(cc, cc, sc); }
    cn_s[tid] = tid < k ? -0.5 * s : -__longlong_as_double(0x7ff0000000000000LL);
    h_s[tid] = tid < k ? sc : 0.0;
  }
  // B operand: tf32(c - mu), canonical K-major; the constant -0.5|c - mu|^2 as a hi/mid/lo tf32 split in K columns d8 .. d8+2
  for (int i = tid; i < KP * (d8 / 4); i += nthreads) {
    const int j = i % KP, c = i / KP;
    uint32_t vv[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) { const int f = 4 * c + e; vv[e] = (j < k && f < d) ? f32_to_tf32_rn((float)(p.C[(size_t)j * d + f] - mu_s[f])) : 0u; }
    *reinterpret_cast<uint4*>(smem + L.b_off + c * LBO_B + (j >> 3) * 128 + (j & 7) * 16) = make_uint4(vv[0], vv[1], vv[2], vv[3]);
  }
  // the column of ones that multiplies the constant: one K step (two chunks: ones, zeros) shared by every group and tile
  for (int r = tid; r < kTileRows; r += nthreads) {
    const uint32_t one = __float_as_uint(1.0f);
    *reinterpret_cast<uint4*>(smem + L.one_off + r * 16) = make_uint4(one, one, one, 0u);
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

  // ------------------------------------------------------------------ per-warp constants
  // loads: this warp's quarter (rows 32q .. 32q+31) of tile i into stage st; four bulk copies of eight rows on the quarter's
  // own mbarrier.  The ragged tail tile (< 128 rows) is fetched by plain loads, zero filled.
  const unsigned char* Xb = reinterpret_cast<const unsigned char*>(p.X);
  const size_t tile_gbytes = (size_t)kTileRows * row_bytes;
  auto issue_load = [&](int i, int st) {
    const uint32_t bar = b_full + 8u * (uint32_t)(4 * st + q), dst = a_x + (uint32_t)st * tile_bytes + (uint32_t)(4 * q) * sub_stride;
    if (lane == 0) sts_u32(a_seq + 4u * (uint32_t)(4 * st + q), (uint32_t)i);
    if (i < my_tiles) {
      if (lane == 0) {
        const unsigned char* src = Xb + ((size_t)blockIdx.x + (size_t)i * gridDim.x) * tile_gbytes + (size_t)(32 * q) * row_bytes;
        const uint32_t cb = (uint32_t)kWgSubRows * row_bytes;
        mbar_expect_tx_a(bar, 4u * cb);
#pragma unroll
        for (int c = 0; c < 4; ++c) bulk_g2s_a(dst + (uint32_t)c * sub_stride, src + (size_t)c * cb, cb, bar);
      }
    } else {
      const double* tsrc = p.X + (size_t)full_tiles * kTileRows * d;
      for (int e = lane; e < 32 * d; e += 32) {
        const int tr = e / d, f = e - tr * d, row = 32 * q + tr;
        sts_f64(dst + (uint32_t)(tr >> 3) * sub_stride + (uint32_t)((tr & 7) * d + f) * 8u, row < tail_rows ? tsrc[(size_t)row * d + f] : 0.0);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive_a(bar);
    }
  };
  // prologue: tile i < S is fetched by group i mod G (later tiles by whoever frees their stage)
  for (int i = grp; i < S && i < T; i += G) issue_load(i, i);
  // steady state: the warp that finishes tile i fetches tile i + S into the same quarter stage; source pointer kept incrementally
  const unsigned char* nsrc = Xb + ((size_t)blockIdx.x + (size_t)(grp + S) * gridDim.x) * tile_gbytes + (size_t)(32 * q) * row_bytes;
  const size_t nstep = (size_t)G * gridDim.x * tile_gbytes;
  const uint32_t sub_bytes = (uint32_t)kWgSubRows * row_bytes;

  // conversion: lane = (row of a group of C_G rows, feature pair)
  constexpr int C_LPR = tp_lanes_per_row(ND, CVEC), C_G = 32 / C_LPR, NIT = 32 / C_G;
  constexpr int NB = NIT > 8 ? 8 : NIT;                            // rows in flight per lane
  const int subc = lane / C_LPR, fc = (lane % C_LPR) * CVEC;
  const bool cact = fc < d;
  const float mu0 = cact ? (float)mu_s[fc] : 0.f, mu1 = (CVEC == 2 && cact) ? (float)mu_s[fc + 1] : 0.f;
  uint64_t mu01;
  asm("mov.b64 %0, {%1,%2};" : "=l"(mu01) : "f"(mu0), "f"(mu1));
  const uint32_t a_A = sb + L.a_off + (uint32_t)grp * (uint32_t)L.a_bytes;
  const uint32_t cdst_lane = a_A + (uint32_t)(32 * q) * 16u + (uint32_t)(fc >> 2) * LBO_A + (uint32_t)(fc & 3) * 4u + (uint32_t)subc * 16u;
  const uint32_t csrc_lane = (uint32_t)subc * row_bytes + (uint32_t)fc * 8u;
  const bool need_xx = do_reduce && xx_mode != 2;
  // screening constants (hkm_lloyd_tp.cuh)
  const uint32_t tlane = ((uint32_t)(32 * q)) << 16;
  uint32_t idxclr;                                                  // ~IDXMASK in a register: (v & idxclr) | index is then one LOP3
  asm volatile("mov.u32 %0, 0xffffffc0;" : "=r"(idxclr));
  const float kapv = 1.04f * 0.0009765625f * vmax;
  const float etab = 0.5f * 1.52587890625e-05f * vmax * vmax + 2.5e-7f * munorm * vmax;
  constexpr int DCH = 2 * ND;
  const uint32_t arow = a_A + (uint32_t)(32 * q + lane) * 16u;
  const bool conv2 = CVEC == 2;
  const int rlane = conv2 ? 2 * lane : lane;
  const bool ract = rlane < d;
  constexpr int RLANES = tp_dclass(ND) / CVEC > 32 ? 32 : tp_dclass(ND) / CVEC;
  const double* Cr = L.cs_off != 0 ? reinterpret_cast<const double*>(smem + L.cs_off) : p.C;
  const uint32_t tbase = tmem + tlane + (uint32_t)grp * (uint32_t)TSTRIDE;
  const uint32_t bar_t = b_tfull + 8u * (uint32_t)grp;
  // one-hot product: lane (g, t) of the fragment layout; k-step ks takes rows 8 t + ks (one from each eight-row sub-chunk, so the
  // four rows sit 32 bytes apart modulo 128 and the fragment loads are conflict free).  N tiles come in pairs that share one
  // 16-byte load: tile 2m holds features 16m + 2g, tile 2m+1 features 16m + 2g + 1 (g = fragment column); an odd last tile holds
  // features 8 (ND-1) + g.  Columns past d read the next row's values (finite, or padding) into accumulator columns nobody reads.
  const int fg = lane >> 2, ft = lane & 3;
  const uint32_t frag_lane = (uint32_t)ft * sub_stride + (uint32_t)fg * 16u;
  const uint32_t frag_last = (uint32_t)ft * sub_stride + (uint32_t)fg * 8u + 64u * (ND - 1);
  double racc[MT * ND * 4];
#pragma unroll
  for (int i = 0; i < MT * ND * 4; ++i) racc[i] = 0.0;
  const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(KP >> 3) << 17) | ((uint32_t)(kTileRows >> 4) << 24);
  // descriptors of the first K step; the next ones advance the 14-bit address field (shared addresses stay below 256 KB)
  const uint64_t da0 = tp_desc_kmajor(a_A, LBO_A, 128u), db0 = tp_desc_kmajor(a_B, LBO_B, 128u);
  const uint64_t da1 = tp_desc_kmajor(sb + L.one_off, LBO_A, 128u);
  const uint32_t da_step = (2u * LBO_A) >> 4, db_step = (2u * LBO_B) >> 4;
  const uint32_t dcol = tmem + (uint32_t)grp * (uint32_t)TSTRIDE;

  // ------------------------------------------------------------------ tiles grp, grp + G, ...
  int s = grp; uint32_t ph = 0, tph = 0;
  for (int i = grp; i < T; i += G) {
    const int rows_valid = i < my_tiles ? kTileRows : tail_rows;
    const uint32_t xs = a_x + (uint32_t)s * tile_bytes + (uint32_t)(4 * q) * sub_stride;     // this warp's 32 rows
    {
      // the quarter must have been requested for THIS tile before its parity means anything (see the header)
      const uint32_t sa = a_seq + 4u * (uint32_t)(4 * s + q);
      uint32_t have;
      asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(have) : "r"(sa) : "memory");
      while (have != (uint32_t)i) {
        __nanosleep(64);
        asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(have) : "r"(sa) : "memory");
      }
    }
    tp_wait_p(b_full + 8u * (uint32_t)(4 * s + q), ph, 3, prof, pw0);
    // ---- conversion of the warp's rows
    if (cact) {
#pragma unroll
      for (int it0 = 0; it0 < NIT; it0 += NB) {
        double v0[NB], v1[NB];
#pragma unroll
        for (int it = 0; it < NB; ++it) {
          const int rb = (it0 + it) * C_G;                       // first row of this instruction's group
          const uint32_t src = xs + (uint32_t)(rb >> 3) * sub_stride + (uint32_t)(rb & 7) * row_bytes + csrc_lane;
          v1[it] = 0.0;
          if (CVEC == 2) lds_v2f64(src, v0[it], v1[it]); else v0[it] = lds_f64(src);
        }
        if (need_xx) {
#pragma unroll
          for (int it = 0; it < NB; ++it) { xxs = fma(v0[it], v0[it], xxs); if (CVEC == 2) xxs = fma(v1[it], v1[it], xxs); }
        }
#pragma unroll
        for (int it = 0; it < NB; ++it) {
          const uint32_t dst = cdst_lane + (uint32_t)((it0 + it) * C_G * 16);
          if (CVEC == 2) {
            // both features of the pair in one packed subtract (FADD2)
            uint64_t vv, rr;
            asm("mov.b64 %0, {%1,%2};" : "=l"(vv) : "f"((float)v0[it]), "f"((float)v1[it]));
            asm("sub.f32x2 %0, %1, %2;" : "=l"(rr) : "l"(vv), "l"(mu01));
            const uint32_t t0 = (uint32_t)rr + 0x1000u, t1 = (uint32_t)(rr >> 32) + 0x1000u;
            asm volatile("st.shared.v2.u32 [%0], {%1,%2};" ::"r"(dst), "r"(t0), "r"(t1) : "memory");
          } else sts_u32(dst, __float_as_uint((float)v0[it] - mu0) + 0x1000u);
        }
      }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");       // conversion stores -> visible to the tensor core
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");   // (and last tile's accumulator reads are done)
    named_bar_sync(1 + grp, 128);
    if (q == 0 && lane == 0) {
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
      for (int ks = 0; ks < ND + 1; ++ks) {                        // K = 8 per instruction: two 16-byte chunks
        const uint64_t da = ks < ND ? da0 + (uint64_t)((uint32_t)ks * da_step) : da1, db = db0 + (uint64_t)((uint32_t)ks * db_step);
        const uint32_t accf = ks > 0 ? 1u : 0u;
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n" ::"r"(dcol),
                     "l"(da), "l"(db), "r"(idesc), "r"(accf)
                     : "memory");
      }
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar_t) : "memory");
    }
    __syncwarp();
    // ---- |x - mu|^2 of this lane's row from the tf32 operand (the tensor core reads the same bytes meanwhile)
    float xa = 0.f, xb = 0.f;
#pragma unroll
    for (int c = 0; c < DCH; ++c) {
      float4 a;
      asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w) : "r"(arow + (uint32_t)c * LBO_A));
      xa = fmaf(a.x, a.x, xa); xb = fmaf(a.y, a.y, xb); xa = fmaf(a.z, a.z, xa); xb = fmaf(a.w, a.w, xb);
    }
    const float xn2 = xa + xb;
    tp_wait_p(bar_t, tph, 5, prof, pw1);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    // ---- sweep 1: best and second best approximate score, centre index in the low mantissa bits
    const float ninf = -__int_as_float(0x7f800000);
    float m1 = ninf, m2 = ninf;
#pragma unroll
    for (int ch = 0; ch < KP / 16; ++ch) {
      uint32_t v[16];
      tp_tmem_ld16(tbase + (uint32_t)(ch * 16), v);
      tp_tmem_wait_ld();
      // top two of the 16 packed values by a tournament: 16 min/max for the pairs, then 7 merges of 3 (max, min, 3-input max)
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
      if (ch == 0) { m1 = hi[0]; m2 = lo[0]; }
      else {
        const float t = fminf(m1, hi[0]);
        asm("max.f32 %0, %1, %2, %3;" : "=f"(m2) : "f"(t), "f"(m2), "f"(lo[0]));
        m1 = fmaxf(m1, hi[0]);
      }
    }
    const int row = 32 * q + lane;
    const bool valid = row < rows_valid;
    const float eps = fmaf(kapv, sqrtf(xn2) * 1.001f, etab) + 6e-5f * fabsf(m1);
    const float thr = m1 - 2.0f * eps;
    const bool amb = valid && !(m2 < thr);
    int lab = valid ? (int)(__float_as_uint(m1) & IDXMASK) : 0;   // rows past the end are zero filled: label 0 adds nothing
    unsigned ambmask = __ballot_sync(0xffffffffu, amb);
    if (ambmask) {
      // sweep 2: every centre within 2 eps of the best is a candidate; then the exact re-scoring, one row at a time, lane =
      // feature (pair), candidates in ascending index (lowest index wins ties, strict '>')
      const long long tr0 = prof ? clock64() : 0;
      uint32_t w[NW];
#pragma unroll
      for (int qq = 0; qq < NW; ++qq) w[qq] = 0u;
#pragma unroll
      for (int ch = 0; ch < KP / 16; ++ch) {
        uint32_t v[16];
        tp_tmem_ld16(tbase + (uint32_t)(ch * 16), v);
        tp_tmem_wait_ld();
#pragma unroll
        for (int c = 0; c < 16; ++c) w[ch >> 1] |= (__uint_as_float(v[c]) >= thr) ? (1u << ((ch & 1) * 16 + c)) : 0u;
      }
      while (ambmask) {
        const int src = __ffs(ambmask) - 1; ambmask &= ambmask - 1;
        uint32_t cm[NW];
#pragma unroll
        for (int qq = 0; qq < NW; ++qq) cm[qq] = __shfl_sync(0xffffffffu, w[qq], src);
        const uint32_t ra = xs + (uint32_t)(src >> 3) * sub_stride + (uint32_t)(src & 7) * row_bytes + (uint32_t)rlane * 8u;
        double x0 = 0.0, x1 = 0.0;
        if (ract) { if (conv2) lds_v2f64(ra, x0, x1); else x0 = lds_f64(ra); }
        double best = -__longlong_as_double(0x7ff0000000000000LL); int bj = -1;
#pragma unroll
        for (int qq = 0; qq < NW; ++qq) {
          uint32_t msk = cm[qq];
          while (msk) {
            int jj[4]; double pp[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) { jj[u] = msk ? 32 * qq + __ffs(msk) - 1 : -1; msk &= msk - 1; }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              double c0 = 0.0, c1 = 0.0;
              if (ract && jj[u] >= 0) {
                const double* cr = Cr + (size_t)jj[u] * d + rlane;
                c0 = cr[0]; if (conv2) c1 = cr[1];
              }
              pp[u] = fma(x0, c0, x1 * c1);
            }
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
        bj = __shfl_sync(0xffffffffu, bj, 0);                    // lane 0 holds the full sums
        if (lane == src) lab = bj;
        ++pcnt;
      }
      if (prof) pw2 += clock64() - tr0;
    }
    if (valid) {
      if (p.labels_out != nullptr) {
        const int64_t r0 = i < my_tiles ? ((int64_t)blockIdx.x + (int64_t)i * gridDim.x) * kTileRows : full_tiles * kTileRows;
        p.labels_out[r0 + row] = lab;
      }
      if (do_reduce) asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(a_cnt + 4u * (uint32_t)lab) : "memory");
    }
    // ---- per-cluster sums of the warp's rows on the fp64 tensor pipe
    if (do_reduce) {
      uint32_t xf = xs;
#pragma unroll
      for (int ks = 0; ks < 8; ++ks, xf += row_bytes) {
        const int labt = __shfl_sync(0xffffffffu, lab, 8 * ft + ks);
        double a[2 * MT];
#pragma unroll
        for (int m = 0; m < 2 * MT; ++m) a[m] = __hiloint2double(labt == fg + 8 * m ? 0x3ff00000 : 0, 0);
#pragma unroll
        for (int m = 0; m < ND / 2; ++m) {
          double b0, b1;
          if (CVEC == 2) lds_v2f64(xf + frag_lane + 128u * m, b0, b1);             // rows are 16-byte aligned only for even d
          else { b0 = lds_f64(xf + frag_lane + 128u * m); b1 = lds_f64(xf + frag_lane + 128u * m + 8u); }
#pragma unroll
          for (int mt = 0; mt < MT; ++mt) {
            double* c = racc + (mt * ND + 2 * m) * 4;
            dmma1684(c[0], c[1], c[2], c[3], a[2 * mt], a[2 * mt + 1], b0);
            dmma1684(c[4], c[5], c[6], c[7], a[2 * mt], a[2 * mt + 1], b1);
          }
        }
        if (ND & 1) {
          const double b2 = lds_f64(xf + frag_last);
#pragma unroll
          for (int mt = 0; mt < MT; ++mt) {
            double* c = racc + (mt * ND + ND - 1) * 4;
            dmma1684(c[0], c[1], c[2], c[3], a[2 * mt], a[2 * mt + 1], b2);
          }
        }
      }
    }
    // ---- this quarter of the stage is free: fetch the tile that uses it next
    __syncwarp();
    if (i + S < my_tiles) {
      // one elected lane, inside the asm: the C++ control flow stays convergent, so the operands stay in uniform registers
      const uint32_t bar = b_full + 8u * (uint32_t)(4 * s + q);
      asm volatile(
          "{\n.reg .pred pe;\nelect.sync _|pe, 0xffffffff;\n"
          "@pe st.shared.u32 [%11], %12;\n"
          "@pe mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n"
          "@pe cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%2], [%6], %10, [%0];\n"
          "@pe cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%3], [%7], %10, [%0];\n"
          "@pe cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%4], [%8], %10, [%0];\n"
          "@pe cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%5], [%9], %10, [%0];\n}\n"
          ::"r"(bar), "r"(4u * sub_bytes), "r"(xs), "r"(xs + sub_stride), "r"(xs + 2u * sub_stride), "r"(xs + 3u * sub_stride),
            "l"(nsrc), "l"(nsrc + sub_bytes), "l"(nsrc + 2u * (size_t)sub_bytes), "l"(nsrc + 3u * (size_t)sub_bytes), "r"(sub_bytes),
            "r"(a_seq + 4u * (uint32_t)(4 * s + q)), "r"((uint32_t)(i + S))
          : "memory");
    } else if (i + S < T) issue_load(i + S, s);
    nsrc += nstep;
    s += G; if (s >= S) { s -= S; ph ^= 1u; }
    tph ^= 1u;
  }

  if (prof && lane == 0) {
    unsigned long long* o = L.prof + warp * 8;
    o[0] = (unsigned long long)(clock64() - pt0); o[1] = (unsigned long long)pw0; o[2] = (unsigned long long)pw1; o[3] = (unsigned long long)pw2;
    o[4] = 0ull; o[5] = (unsigned long long)pcnt; o[6] = (unsigned long long)T; o[7] = 0ull;
  }
  // ===================================================================== CTA-level combine (all warps)
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)L.tcols) : "memory");
  if (!p.do_reduce) return;

  // every tile is reduced and no copy is in flight, so the ring is idle: the warps park their fragments there,
  // slab w = [KP][8 ND] doubles
  constexpr int LDS_ = 8 * ND;
  {
    double* slab = reinterpret_cast<double*>(smem + L.x_off) + (size_t)warp * KP * LDS_;
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int nd = 0; nd < ND; ++nd) {
        const double* c = racc + (mt * ND + nd) * 4;
        const bool paired = nd < 2 * (ND / 2);
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int f = paired ? 16 * (nd >> 1) + 2 * (2 * ft + e) + (nd & 1) : 8 * nd + 2 * ft + e;
          slab[(16 * mt + fg) * LDS_ + f] = c[e];
          slab[(16 * mt + fg + 8) * LDS_ + f] = c[2 + e];
        }
      }
  }
  __syncthreads();
  {
    const double* slab = reinterpret_cast<const double*>(smem + L.x_off);
    for (int i = tid; i < k * d; i += nthreads) {
      const int j = i / d, f = i - j * d;
      double sacc = 0;
      for (int w = 0; w < nwarps; ++w) sacc += slab[(size_t)w * KP * LDS_ + j * LDS_ + f];
      out[i] = sacc;
      cost = fma(sacc, p.C[i], cost);                           // sum_j S_j . c_j
    }
  }
  if (tid < k) { const double nj = (double)cnt_s[tid]; out[(size_t)k * d + tid] = nj; cost = fma(nj, cn_s[tid], cost); }
#pragma unroll
  for (int msk = 16; msk >= 1; msk >>= 1) { cost += shfl_xor_d(cost, msk); xxs += shfl_xor_d(xxs, msk); }
  if (lane == 0) { red_s[warp] = cost; red_s[kWgMaxWarps + warp] = xxs; }
  __syncthreads();
  if (tid == 0) {
    double c = 0, x2 = 0;
    for (int wv = 0; wv < nwarps; ++wv) { c += red_s[wv]; x2 += red_s[kWgMaxWarps + wv]; }
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
    for (int i = tid; i < P + 1; i += nthreads) {
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
      peer_ok = peer_allreduce(p, p.partial_out, P, tid, nthreads, [] { __syncthreads(); });
    }
    if (p.C_update != nullptr && peer_ok) {                      // fused driver step
      __threadfence_block();
      __syncthreads();
      update_centers_body(p.partial_out, k, d, p.C_update, p.tol2, p.state_rw, red_s + 2 * kWgMaxWarps, tid);
    }
  }
}

}  // namespace hkm

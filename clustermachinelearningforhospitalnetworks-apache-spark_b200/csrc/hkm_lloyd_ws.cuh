// Lloyd data pass, warp-specialised variant for 8 < k <= 64, d <= 32: "score" warps keep the fp64 tensor pipe busy,
// "reduce" warps build the per-cluster sums one tile behind them.
//
// Why: for k > 8 the DMMA score tiles (k*d/32 tensor-pipe clocks per row) are the scarce resource, and the variants in
// hkm_lloyd.cuh / hkm_lloyd_mma.cuh leave that pipe half idle -- every warp alternates between a DMMA phase and a
// DMMA-free phase (argmax, label exchange, gather), and DSETP/DADD share the fp64 pipe with DMMA.  Here the roles are
// split.  One CTA of 16 warps per SM (4 per scheduler partition, 128 registers):
//   * SW score warps take (tile, 16-row sub-tile pair) items round robin: wait for the tile's
//     "full" mbarrier, run the score DMMAs + the shuffle argmax, add the cost, store the 16 labels into the stage's
//     label slots, arrive on the stage's "scored" mbarrier (8 arrivals per tile).
//   * RW = 16 - SW reduce warps take (tile, 16-row chunk) items round robin as well and own NSET private [k_pad, row]
//     fp64 accumulator blocks in shared memory: wait for "scored", then four rows at a time (lane = feature): one LDS per row, and one
//     LDS / DADD / STS on the accumulator row of its label in block (row % NSET).  The four read-modify-writes are
//     issued together unless two rows of the group share a block AND a label (one ballot per tile finds those groups;
//     they take the row-by-row path).  For d < 32 lane d adds 1.0, so the rows per cluster are column d of the
//     accumulators; for d = 32 they are counted in registers (lane c counts clusters c and c + 32).  The reduce warp
//     that finishes the last chunk of a tile (an acq_rel counter per stage) issues the bulk copies of the tile that
//     reuses the stage.
// No CTA barrier in the steady state; summation order is fixed (row order within a reduce warp, warp order in the
// combine), so results are bit-stable.  Same tile geometry, TMA ring and last-CTA fold as hkm_lloyd.cuh.
// Plain Lloyd passes only (unit weights, do_reduce 1, out_mode 0); the dispatcher sends everything else elsewhere.
#pragma once
#include "hkm_lloyd.cuh"

namespace hkm {

constexpr int kWsWarps = 16;
constexpr int kWsThreads = kWsWarps * 32;
constexpr int kWsPairs = kTileRows / 16;   // 16-row score items per tile

struct WsSmemLayout { int x_off, bar_off, lab_off, cs_off, cn_off, red_off, acc_off, cnt_off, total; };
// rw reduce warps, each with nset accumulator blocks (row r of its slice goes to block r % nset: nset rows in flight
// without two of them ever touching the same address)
// accumulator row (doubles): d sums, plus the row count in column d when the kernel keeps it there (d <= 28)
__host__ __device__ inline int ws_acc_ld(int d) { return d <= 28 ? d + 1 : d; }
__host__ __device__ inline WsSmemLayout ws_smem_layout(int d, int k_pad, int stages, int rw, int nset) {
  WsSmemLayout L; int o = 0;
  L.x_off = o; o += stages * stage_bytes(d); o = (o + 127) & ~127;
  L.bar_off = o; o += 3 * kMaxStages * 8;                 // full | scored | "reduce warps done" counters
  L.lab_off = o; o += kMaxStages * kTileRows * 4;
  L.cs_off = o; o += k_pad * cs_ld(d) * 8;
  L.cn_off = o; o += k_pad * 8;
  L.red_off = o; o += 2 * kWsWarps * 8;
  L.acc_off = o; o += rw * nset * k_pad * ws_acc_ld(d) * 8;
  L.cnt_off = o; o += rw * k_pad * 4;
  L.total = o;
  return L;
}
__host__ __device__ inline int ws_stages(int d, int k_pad, int rw, int nset) {
  const int fixed = ws_smem_layout(d, k_pad, 0, rw, nset).total + 128;
  const int st = (227 * 1024 - fixed) / stage_bytes(d);
  return st > kMaxStages ? kMaxStages : st;
}
// accumulator blocks per reduce warp: two when four stages still fit next to them, else one
__host__ __device__ inline int ws_nset(int d, int k_pad, int rw) { return ws_stages(d, k_pad, rw, 2) >= 4 ? 2 : 1; }
// score warps: always 8 (+ 8 reduce warps).  A 12 + 4 split was faster for k > 48, but a score warp then skips tiles and can
// come back to a ring stage two wraps later, which a parity wait cannot tell from the current wrap; with 8 score warps every
// warp visits every tile, so every wait sees consecutive phases.  (The pipelined tcgen05 variant serves these shapes now.)
__host__ __device__ inline int ws_score_warps(int nt) { (void)nt; return 8; }

// KS_T > 0: centre B-fragments in registers (d_pad = 4*KS_T); KS_T == 0: re-read from shared memory, p.ks k-steps.
template <int KS_T, int NT, int SW, int NSET, int MAXREG>
__global__ void __launch_bounds__(kWsThreads) __maxnreg__(MAXREG) lloyd_ws_kernel(const LloydParams p) {
  if (p.skip != nullptr && *p.skip != 0) return;
  if (p.skip2 != nullptr && *p.skip2 != 0) return;
  constexpr int k_pad = NT * 8;
  constexpr int RW = kWsWarps - SW;
  constexpr int RROWS = 16;                 // reduce item: one 16-row chunk ("slice") of a tile, 8 per tile
  constexpr bool CNTLANE = KS_T >= 1 && KS_T <= 7;   // d <= 28: lane d adds 1.0 per row, the count is column d of the accumulators
  static_assert(SW >= 8 && SW <= 12, "score warps: 8..12");

  extern __shared__ __align__(128) unsigned char smem[];
  const int d = p.d, k = p.k;
  const int KS = KS_T > 0 ? KS_T : p.ks;
  const WsSmemLayout L = ws_smem_layout(d, k_pad, p.stages, RW, NSET);
  double* Cs = reinterpret_cast<double*>(smem + L.cs_off);
  double* cn_s = reinterpret_cast<double*>(smem + L.cn_off);
  double* red_s = reinterpret_cast<double*>(smem + L.red_off);
  const uint32_t sb = smem_u32(smem);
  const uint32_t a_x = sb + L.x_off, a_full = sb + L.bar_off, a_scored = a_full + 8u * kMaxStages, a_done = a_scored + 8u * kMaxStages;
  const uint32_t a_lab = sb + L.lab_off;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int stages = p.stages;
  const int csld = cs_ld(d);
  const uint32_t cstride = (uint32_t)chunk_stride_bytes(d);
  const uint32_t sbytes = (uint32_t)stage_bytes(d);
  const uint32_t row_bytes = (uint32_t)d * 8u;
  const int64_t full_tiles = p.n / kTileRows;
  const int tail_rows = (int)(p.n - full_tiles * kTileRows);
  const int my_tiles = full_tiles > (int64_t)blockIdx.x ? (int)((full_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;
  const bool tail_mine = tail_rows > 0 && (int)(full_tiles % gridDim.x) == (int)blockIdx.x;

  if (tid == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(reinterpret_cast<uint64_t*>(smem + L.bar_off) + s, 1);
      mbar_init(reinterpret_cast<uint64_t*>(smem + L.bar_off) + kMaxStages + s, kWsPairs);
      reinterpret_cast<unsigned long long*>(smem + L.bar_off)[2 * kMaxStages + s] = 0ull;
    }
    mbar_fence_init();
  }
  for (int i = tid; i < k_pad * csld; i += kWsThreads) {
    int j = i / csld, f = i - j * csld;
    Cs[i] = (j < k && f < d) ? p.C[(size_t)j * d + f] : 0.0;
  }
  {
    double* acc0 = reinterpret_cast<double*>(smem + L.acc_off);
    for (int i = tid; i < RW * NSET * k_pad * ws_acc_ld(d); i += kWsThreads) acc0[i] = 0.0;
  }
  __syncthreads();
  for (int j = tid; j < k_pad; j += kWsThreads) {
    double s = 0;
    for (int f = 0; f < d; ++f) s = fma(Cs[j * csld + f], Cs[j * csld + f], s);
    cn_s[j] = j < k ? -0.5 * s : -__longlong_as_double(0x7ff0000000000000LL);
  }
  __syncthreads();

  // ---- TMA refills: local tile j of this CTA is global tile blockIdx.x + j*gridDim.x and lands in stage j % stages
  const uint32_t cbytes = (uint32_t)kChunkRows * row_bytes;
  const size_t tile_gbytes = (size_t)kTileRows * row_bytes;
  const unsigned char* src0 = reinterpret_cast<const unsigned char*>(p.X) + (size_t)blockIdx.x * tile_gbytes;
  const size_t src_step = (size_t)gridDim.x * tile_gbytes;
  auto issue_tile = [&](int j, int st) {
    const uint32_t bar = a_full + 8u * (uint32_t)st, dst = a_x + (uint32_t)st * sbytes;
    const unsigned char* src = src0 + (size_t)j * src_step;
    mbar_expect_tx_a(bar, cbytes * kChunks);
#pragma unroll
    for (int c = 0; c < kChunks; ++c) bulk_g2s_a(dst + c * cstride, src + (size_t)c * cbytes, cbytes, bar);
  };
  if (tid == 0) {
    for (int s = 0; s < stages && s < my_tiles; ++s) issue_tile(s, s);
  }

  double cost = 0, xxs = 0;          // score warps
  unsigned cnt0 = 0, cnt1 = 0;       // reduce warps: rows seen for clusters `lane` and `lane + 32`
  const int xx_mode = p.xx_mode;

  // warps 0..SW-1 score, the rest reduce (spread over the four scheduler partitions; putting the reduce warps on one
  // partition to keep their DADDs away from the DMMA queue was measured slower: that partition becomes the bottleneck)
  const bool is_score = warp < SW;
  const int sidx = warp, ridx = warp - SW;
  if (is_score) {
    // ===================================================================== score warps
    const int g = lane >> 2, t = lane & 3;       // MMA fragment coordinates
    double breg[KS_T > 0 ? KS_T * NT : 1];        // centre B-fragments: C[8*nt + g][4*s + t]
    if (KS_T > 0) {
#pragma unroll
      for (int s = 0; s < (KS_T > 0 ? KS_T : 1); ++s)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) breg[s * NT + nt] = Cs[(8 * nt + g) * csld + 4 * s + t];
    }
    double cinit[NT][2];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) { cinit[nt][0] = cn_s[8 * nt + 2 * t]; cinit[nt][1] = cn_s[8 * nt + 2 * t + 1]; }
    const uint32_t a_frag = (uint32_t)g * cstride + (uint32_t)t * 8u;        // score A-fragment: chunk g, feature t (+4s)
    const uint32_t a_cb = sb + L.cs_off + (uint32_t)((g * csld + t) * 8);   // KS_T == 0: B fragments re-read from smem

    // one item: rows (chunk g, pr) and (chunk g, pr + 8) of the tile in stage memory xs; labels -> lab_b
    auto score_item = [&](auto need_xx_tag, const uint32_t xs, const uint32_t lab_b, const int pr, const int rows_valid, const int64_t row0) {
      constexpr bool NEED_XX = decltype(need_xx_tag)::value;
      const int t0 = g * kChunkRows + pr, t1 = t0 + 8;
      const uint32_t xa0 = xs + a_frag + (uint32_t)pr * row_bytes, xa1 = xa0 + 8u * row_bytes;
      double s0[NT][2], s1[NT][2];
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) { s0[nt][0] = cinit[nt][0]; s0[nt][1] = cinit[nt][1]; s1[nt][0] = cinit[nt][0]; s1[nt][1] = cinit[nt][1]; }
      double xx0 = 0, xx1 = 0;
      if (KS_T > 0) {
#pragma unroll
        for (int s = 0; s < (KS_T > 0 ? KS_T : 1); ++s) {
          const bool ok = (s < KS_T - 1) || (4 * s + t < d);
          const double a0 = ok ? lds_f64(xa0 + 32u * s) : 0.0, a1 = ok ? lds_f64(xa1 + 32u * s) : 0.0;
          if (NEED_XX) { xx0 = fma(a0, a0, xx0); xx1 = fma(a1, a1, xx1); }
#pragma unroll
          for (int nt = 0; nt < NT; ++nt) dmma1684(s0[nt][0], s0[nt][1], s1[nt][0], s1[nt][1], a0, a1, breg[s * NT + nt]);
        }
      } else {
        for (int s = 0; s < KS; ++s) {
          const bool ok = 4 * s + t < d;
          const double a0 = ok ? lds_f64(xa0 + 32u * s) : 0.0, a1 = ok ? lds_f64(xa1 + 32u * s) : 0.0;
          if (NEED_XX) { xx0 = fma(a0, a0, xx0); xx1 = fma(a1, a1, xx1); }
#pragma unroll
          for (int nt = 0; nt < NT; ++nt) {
            const double b = lds_f64(a_cb + (uint32_t)((8 * nt) * csld * 8) + 32u * s);
            dmma1684(s0[nt][0], s0[nt][1], s1[nt][0], s1[nt][1], a0, a1, b);
          }
        }
      }
      // argmax (lowest index wins ties)
      double b0 = s0[0][0], b1 = s1[0][0]; int j0 = 2 * t, j1 = 2 * t;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          if (nt == 0 && e == 0) continue;
          const int j = 8 * nt + 2 * t + e;
          if (s0[nt][e] > b0) { b0 = s0[nt][e]; j0 = j; }
          if (s1[nt][e] > b1) { b1 = s1[nt][e]; j1 = j; }
        }
#pragma unroll
      for (int msk = 1; msk <= 2; msk <<= 1) {
        const double o0 = shfl_xor_d(b0, msk), o1 = shfl_xor_d(b1, msk);
        const int oj0 = __shfl_xor_sync(0xffffffffu, j0, msk), oj1 = __shfl_xor_sync(0xffffffffu, j1, msk);
        if (o0 > b0 || (o0 == b0 && oj0 < j0)) { b0 = o0; j0 = oj0; }
        if (o1 > b1 || (o1 == b1 && oj1 < j1)) { b1 = o1; j1 = oj1; }
        if (NEED_XX) { xx0 += shfl_xor_d(xx0, msk); xx1 += shfl_xor_d(xx1, msk); }
      }
      if (t == 0) {
        const bool v0 = t0 < rows_valid, v1 = t1 < rows_valid;
        double d0 = b0, d1 = b1;
        if (NEED_XX) { d0 = fmax(fma(-2.0, b0, xx0), 0.0); d1 = fmax(fma(-2.0, b1, xx1), 0.0); }
        const double c0 = (NEED_XX && xx_mode == 2) ? b0 : d0, c1 = (NEED_XX && xx_mode == 2) ? b1 : d1;
        if (v0) cost += c0;
        if (v1) cost += c1;
        if (NEED_XX && xx_mode == 1) { if (v0) xxs += xx0; if (v1) xxs += xx1; }
        sts_u32(lab_b + 4u * (uint32_t)t0, (uint32_t)(v0 ? j0 : -1));
        sts_u32(lab_b + 4u * (uint32_t)t1, (uint32_t)(v1 ? j1 : -1));
        if (p.labels_out != nullptr) { if (v0) p.labels_out[row0 + t0] = j0; if (v1) p.labels_out[row0 + t1] = j1; }
        if (NEED_XX && p.mind2_out != nullptr) { if (v0) p.mind2_out[row0 + t0] = d0; if (v1) p.mind2_out[row0 + t1] = d1; }
      }
    };

    // items q = tile * 8 + pair, taken round robin by the SW score warps
    int tile = sidx >> 3, pr = sidx & 7;                 // q = sidx
    int s = tile, wraps = 0;                             // stage of `tile` (stages >= 2 > tile) and how often the ring wrapped
    constexpr int DT = SW / kWsPairs, DP = SW % kWsPairs;   // q += SW  ->  tile += DT (+1 on pair overflow), pair += DP
    while (tile < my_tiles) {
      mbar_wait_a(a_full + 8u * (uint32_t)s, (uint32_t)(wraps & 1));
      const uint32_t xs = a_x + (uint32_t)s * sbytes, lab_b = a_lab + (uint32_t)s * (kTileRows * 4);
      const int64_t row0 = ((int64_t)blockIdx.x + (int64_t)tile * gridDim.x) * kTileRows;
      if (xx_mode != 2) score_item(std::true_type{}, xs, lab_b, pr, kTileRows, row0);
      else score_item(std::false_type{}, xs, lab_b, pr, kTileRows, row0);
      __syncwarp();
      if (lane == 0) mbar_arrive(reinterpret_cast<uint64_t*>(smem + L.bar_off) + kMaxStages + s);
      int dt = DT; pr += DP;
      if (pr >= kWsPairs) { pr -= kWsPairs; ++dt; }
      tile += dt; s += dt;
      while (s >= stages) { s -= stages; ++wraps; }
    }
    if (tail_mine) {
      // ragged tail (< 128 rows): loaded by all threads with plain loads into stage 0 after everything else is consumed
      named_bar_sync(1, kWsThreads);
      unsigned char* xsp = smem + L.x_off;
      const double* src = p.X + (size_t)full_tiles * kTileRows * d;
      for (int i = tid; i < kTileRows * d; i += kWsThreads) {
        int tr = i / d, f = i - tr * d;
        double v = tr < tail_rows ? src[(size_t)tr * d + f] : 0.0;
        reinterpret_cast<double*>(xsp + (tr >> 4) * cstride)[(tr & 15) * d + f] = v;
      }
      named_bar_sync(1, kWsThreads);
      if (sidx < kWsPairs) score_item(std::true_type{}, a_x, a_lab, sidx, tail_rows, full_tiles * kTileRows);
      named_bar_sync(1, kWsThreads);       // labels of the tail tile are in place
    }
  } else {
    // ===================================================================== reduce warps
    const int rw = ridx;
    const bool one_lane = CNTLANE && lane == d;                               // this lane adds 1.0 per row (the count column)
    const uint32_t acc_row = (uint32_t)ws_acc_ld(d) * 8u;
    const bool fl_ok = lane < ws_acc_ld(d);                                    // lane = feature (or the count column)
    const uint32_t set_bytes = (uint32_t)k_pad * acc_row;
    const uint32_t a_acc = sb + L.acc_off + (uint32_t)(rw * NSET) * set_bytes + (uint32_t)lane * 8u;
    const int q4 = lane & 3;

    // one item: the 16 rows of chunk `slice` of the tile in stage memory xs
    auto reduce_rows = [&](const uint32_t xs, const uint32_t lab_b, const int slice) {
      const int lab = (int)lds_u32(lab_b + 4u * (uint32_t)(slice * RROWS + (lane & (RROWS - 1))));   // lane r holds the label of row r of the slice
      // groups of four rows whose read-modify-writes may not be issued together: two rows in the same block with the
      // same label, or an invalid row (tail tile)
      bool clash = lab < 0;
      if (NSET <= 2) { const int o = __shfl_up_sync(0xffffffffu, lab, 2); clash |= q4 >= 2 && o == lab; }
      if (NSET == 1) {
        const int o1 = __shfl_up_sync(0xffffffffu, lab, 1), o3 = __shfl_up_sync(0xffffffffu, lab, 3);
        clash |= (q4 >= 1 && o1 == lab) || (q4 == 3 && o3 == lab);
      }
      const unsigned cm = __ballot_sync(0xffffffffu, clash);
      uint32_t xr = xs + (uint32_t)slice * cstride + (uint32_t)lane * 8u;
#pragma unroll
      for (int r = 0; r < RROWS; r += 4) {
        int lb[4]; double v[4], a[4]; uint32_t aa[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          lb[u] = __shfl_sync(0xffffffffu, lab, r + u);
          v[u] = lds_f64(xr);                              // lanes >= d read past the row: replaced or never stored
          xr += row_bytes;
          if (CNTLANE) v[u] = one_lane ? 1.0 : v[u];
          else { cnt0 += (lb[u] == lane ? 1u : 0u); if (NT > 4) cnt1 += (lb[u] == lane + 32 ? 1u : 0u); }
          aa[u] = a_acc + (uint32_t)(u % NSET) * set_bytes + (uint32_t)lb[u] * acc_row;
        }
        if (((cm >> r) & 0xFu) == 0u) {                    // warp-uniform
#pragma unroll
          for (int u = 0; u < 4; ++u) a[u] = fl_ok ? lds_f64(aa[u]) : 0.0;
#pragma unroll
          for (int u = 0; u < 4; ++u) a[u] += v[u];
#pragma unroll
          for (int u = 0; u < 4; ++u) if (fl_ok) sts_f64(aa[u], a[u]);
        } else {
#pragma unroll
          for (int u = 0; u < 4; ++u)
            if (lb[u] >= 0 && fl_ok) { double t2 = lds_f64(aa[u]); t2 += v[u]; sts_f64(aa[u], t2); }
        }
      }
    };

    // items u = tile * 8 + slice, taken round robin by the RW reduce warps
    int tile = rw >> 3, slice = rw & 7;
    int s = tile, wraps = 0;
    constexpr int DT = RW / kWsPairs, DP = RW % kWsPairs;
    while (tile < my_tiles) {
      mbar_wait_a(a_scored + 8u * (uint32_t)s, (uint32_t)(wraps & 1));
      reduce_rows(a_x + (uint32_t)s * sbytes, a_lab + (uint32_t)s * (kTileRows * 4), slice);
      __syncwarp();
      if (lane == 0) {
        // stage hand-back: whoever finishes the last of the tile's 8 slices refills the stage with tile + stages
        unsigned old;
        asm volatile("atom.acq_rel.cta.shared.add.u32 %0, [%1], 1;" : "=r"(old) : "r"(a_done + 8u * (uint32_t)s) : "memory");
        if (old == kWsPairs - 1) {
          sts_u32(a_done + 8u * (uint32_t)s, 0u);
          if (tile + stages < my_tiles) issue_tile(tile + stages, s);
        }
      }
      int dt = DT; slice += DP;
      if (slice >= kWsPairs) { slice -= kWsPairs; ++dt; }
      tile += dt; s += dt;
      while (s >= stages) { s -= stages; ++wraps; }
    }
    if (tail_mine) {
      named_bar_sync(1, kWsThreads);
      unsigned char* xsp = smem + L.x_off;
      const double* src = p.X + (size_t)full_tiles * kTileRows * d;
      for (int i = tid; i < kTileRows * d; i += kWsThreads) {
        int tr = i / d, f = i - tr * d;
        double v = tr < tail_rows ? src[(size_t)tr * d + f] : 0.0;
        reinterpret_cast<double*>(xsp + (tr >> 4) * cstride)[(tr & 15) * d + f] = v;
      }
      named_bar_sync(1, kWsThreads);
      named_bar_sync(1, kWsThreads);       // score warps have written the tail tile's labels
      for (int sl = rw; sl < kWsPairs; sl += RW) reduce_rows(a_x, a_lab, sl);
    }
  }

  // ===================================================================== CTA-level combine
  const int P = k * d + k + 1;
  const int S = P + 1 + k;
  double* out = p.cta_partials + (size_t)blockIdx.x * S;
#pragma unroll
  for (int msk = 16; msk >= 1; msk >>= 1) { cost += shfl_xor_d(cost, msk); xxs += shfl_xor_d(xxs, msk); }
  if (lane == 0) { red_s[warp] = cost; red_s[kWsWarps + warp] = xxs; }
  unsigned* cnt_s = reinterpret_cast<unsigned*>(smem + L.cnt_off);
  if (!is_score) {
    if (lane < k_pad) cnt_s[ridx * k_pad + lane] = cnt0;
    if (NT > 4 && lane + 32 < k_pad) cnt_s[ridx * k_pad + lane + 32] = cnt1;
  }
  __syncthreads();
  {
    const double* acc0 = reinterpret_cast<const double*>(smem + L.acc_off);
    const int kd = k * d, ald = ws_acc_ld(d), wstride = k_pad * ald;
    {
      for (int i = tid; i < kd; i += kWsThreads) {
        const int j = i / d, f = i - j * d;
        double s = 0;
#pragma unroll
        for (int w = 0; w < RW * NSET; ++w) s += acc0[w * wstride + j * ald + f];
        out[i] = s;
      }
      for (int j = tid; j < k; j += kWsThreads) {
        double c = 0;
        if (CNTLANE) {
#pragma unroll
          for (int w = 0; w < RW * NSET; ++w) c += acc0[w * wstride + j * ald + d];
        } else {
          unsigned cu = 0;
#pragma unroll
          for (int w = 0; w < RW; ++w) cu += cnt_s[w * k_pad + j];
          c = (double)cu;
        }
        out[(size_t)kd + j] = c;
      }
    }
    if (tid == 0) {
      double c = 0, x2 = 0;
#pragma unroll
      for (int wv = 0; wv < kWsWarps; ++wv) { c += red_s[wv]; x2 += red_s[kWsWarps + wv]; }
      out[P - 1] = c; out[P] = x2;
    }
  }
  // ===================================================================== last CTA folds all partials
  __threadfence();
  __syncthreads();
  __shared__ unsigned int s_ticket;
  if (tid == 0) s_ticket = atomicAdd(p.counter, 1u);
  __syncthreads();
  if (s_ticket == gridDim.x - 1) {
    __threadfence();
    for (int i = tid; i < P + 1; i += kWsThreads) {
      double s = 0;
      s = ordered_sum_ldcg(p.cta_partials + i, (size_t)S, gridDim.x);
      if (i == P - 1 && xx_mode == 2) s = fmax(fma(-2.0, s, *p.xnorm2_sum), 0.0);  // cost = sum |x|^2 - 2 sum score
      if (i < P) p.partial_out[i] = s;
      else if (xx_mode == 1) *p.xnorm2_sum = s;
    }
    if (tid == 0) *p.counter = 0u;
    bool peer_ok = true;
    if (p.peer_bufs != nullptr) {                            // ranks > 1: sum the vector over NVLink peer memory right here
      __threadfence_block();
      __syncthreads();
      peer_ok = peer_allreduce(p, p.partial_out, P, tid, kWsThreads, [] { __syncthreads(); });
    }
    if (p.C_update != nullptr && peer_ok) {                             // fused driver step
      __threadfence_block();
      __syncthreads();
      update_centers_body(p.partial_out, k, d, p.C_update, p.tol2, p.state_rw, reinterpret_cast<double*>(smem + L.x_off), tid);
    }
  }
}

}  // namespace hkm

// Lloyd data pass, small-k variant: assign AND reduce on the fp64 tensor pipe, no CTA barrier.
//
// For k <= 16 the per-cluster sums are themselves a skinny matrix product,
//     S[k, d] += onehot(labels)[rows, k]^T * X[rows, d],
// so after the score tiles (x.c_j - 0.5|c_j|^2, DMMA.8x8x4) and the shuffle argmax, each warp turns its 16 labels
// into 8x4 one-hot A-fragments (four shuffles) and feeds the very rows it just scored through a second set of
// DMMAs whose accumulators -- the warp's private [k, d] partial sums -- never leave registers.  Compared with
// the gather variant (hkm_lloyd.cuh) this doubles the DMMA count but removes the label exchange through shared
// memory, the per-tile CTA barrier, the row lists and the per-warp load imbalance: warps are independent and
// only meet the producer through full/empty mbarriers.  At k = 8, d = 32 the tensor pipe needs ~75 % of its
// 64 FMA/clk/SM at the HBM roof; k >= 32 would exceed it, which is where the gather variant takes over.
//
// Same tile geometry, TMA ring and last-CTA fold as hkm_lloyd.cuh.
#pragma once
#include "hkm_lloyd.cuh"

namespace hkm {

constexpr int kMmaThreads = kComputeThreads + 32;   // 8 compute warps + 1 producer warp

// dynamic smem: stages | full/empty barriers | centres (staging at start) | combine buffer (aliases the stages)
struct MmaSmemLayout { int x_off, bar_off, cs_off, cn_off, red_off, total; };
__host__ __device__ inline MmaSmemLayout mma_smem_layout(int d, int k_pad, int stages) {
  MmaSmemLayout L; int o = 0;
  L.x_off = o; o += stages * stage_bytes(d); o = (o + 127) & ~127;
  L.bar_off = o; o += 2 * kMaxStages * 8;
  L.cs_off = o; o += k_pad * cs_ld(d) * 8;
  L.cn_off = o; o += 2 * k_pad * 8;                       // second half: cluster weight sums (silhouette score pass)
  L.red_off = o; o += 2 * kComputeWarps * 8 + 16;
  L.total = o;
  return L;
}

// KS_T = ceil(d/4) (score k-steps), NT = ceil(k/8) (1 or 2), ND = ceil(d/8) feature tiles of the reduce product.
// MODE 0: Lloyd / predict pass.  MODE 1: silhouette cluster statistics (labels given; the one-hot product also
// carries the weights, and one extra feature tile accumulates Psi = sum w|x|^2).  MODE 2: silhouette score pass.
template <int KS_T, int NT, int MAXREG, int MODE = 0>
__global__ void __launch_bounds__(kMmaThreads) __maxnreg__(MAXREG) lloyd_mma_kernel(const LloydParams p) {
  if (p.skip != nullptr && *p.skip != 0) return;
  if (p.skip2 != nullptr && *p.skip2 != 0) return;
  constexpr int ND = (KS_T + 1) / 2;
  constexpr int NDX = MODE == 1 ? ND + 1 : ND;   // accumulator feature tiles (one more for Psi)
  constexpr int k_pad = NT * 8;

  extern __shared__ __align__(128) unsigned char smem[];
  const int d = p.d, k = p.k;
  const MmaSmemLayout L = mma_smem_layout(d, k_pad, p.stages);
  double* Cs = reinterpret_cast<double*>(smem + L.cs_off);
  double* cn_s = reinterpret_cast<double*>(smem + L.cn_off);
  double* red_s = reinterpret_cast<double*>(smem + L.red_off);
  const uint32_t sb = smem_u32(smem);
  const uint32_t a_x = sb + L.x_off, a_full = sb + L.bar_off, a_empty = a_full + 8u * kMaxStages;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int stages = p.stages;
  const int csld = cs_ld(d);
  const uint32_t cstride = (uint32_t)chunk_stride_bytes(d);
  const uint32_t sbytes = (uint32_t)stage_bytes(d);
  const uint32_t row_bytes = (uint32_t)d * 8u;
  const int64_t full_tiles = p.n / kTileRows;
  const int tail_rows = (int)(p.n - full_tiles * kTileRows);
  const int my_tiles = full_tiles > (int64_t)blockIdx.x ? (int)((full_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;

  if (tid == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(reinterpret_cast<uint64_t*>(smem + L.bar_off) + s, 1);
      mbar_init(reinterpret_cast<uint64_t*>(smem + L.bar_off) + kMaxStages + s, kComputeWarps);
    }
    mbar_fence_init();
  }
  for (int i = tid; i < k_pad * csld; i += kMmaThreads) {
    int j = i / csld, f = i - j * csld;
    Cs[i] = (MODE != 1 && j < k && f < d) ? p.C[(size_t)j * d + f] : 0.0;   // the statistics pass has no centres
  }
  __syncthreads();
  for (int j = tid; j < k_pad; j += kMmaThreads) {
    double s = 0;
    for (int f = 0; f < d; ++f) s = fma(Cs[j * csld + f], Cs[j * csld + f], s);
    const double ninf = -__longlong_as_double(0x7ff0000000000000LL);
    if (MODE == 2) {   // D(x, G) = |x|^2 + Psi_G/N_G - 2 x.(Y_G/N_G): the Lloyd score with -0.5*Psi_G/N_G as the constant
      const double Nj = j < k ? p.aux[j] : 0.0;
      cn_s[j] = (j < k && Nj > 0) ? -0.5 * p.aux[k + j] : ninf;
      cn_s[k_pad + j] = Nj;
    } else cn_s[j] = j < k ? -0.5 * s : ninf;
  }
  __syncthreads();

  // ===================================================================== producer warp
  if (warp == kComputeWarps) {
    if (lane == 0) {
      const uint32_t cbytes = (uint32_t)kChunkRows * row_bytes;
      const size_t tile_gbytes = (size_t)kTileRows * row_bytes;
      const unsigned char* src = reinterpret_cast<const unsigned char*>(p.X) + (size_t)blockIdx.x * tile_gbytes;
      const size_t src_step = (size_t)gridDim.x * tile_gbytes;
      int s = 0; uint32_t ph = 0;
      for (int i = 0; i < my_tiles; ++i) {
        mbar_wait_a(a_empty + 8u * (uint32_t)s, ph ^ 1u);
        const uint32_t bar = a_full + 8u * (uint32_t)s, dst = a_x + (uint32_t)s * sbytes;
        mbar_expect_tx_a(bar, cbytes * kChunks);
#pragma unroll
        for (int c = 0; c < kChunks; ++c) bulk_g2s_a(dst + c * cstride, src + (size_t)c * cbytes, cbytes, bar);
        src += src_step;
        if (++s == stages) { s = 0; ph ^= 1u; }
      }
    }
    // the producer warp takes no part in the compute-warp barriers below (named barrier 1, 256 threads)
  } else {
    // ===================================================================== compute warps
    const int g = lane >> 2, t = lane & 3;       // MMA fragment coordinates (row group / k index)

    double breg[KS_T * NT];                       // centre B-fragments: C[8*nt + g][4*s + t]
#pragma unroll
    for (int s = 0; s < KS_T; ++s)
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) breg[s * NT + nt] = Cs[(8 * nt + g) * csld + 4 * s + t];
    double cinit[NT][2];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) { cinit[nt][0] = cn_s[8 * nt + 2 * t]; cinit[nt][1] = cn_s[8 * nt + 2 * t + 1]; }

    // private partial sums: acc[mt][nd][e] = S[8*mt + g][8*nd + 2*t + e]
    double acc[NT][NDX][2];
    unsigned cnt[NT];
    double wcnt[NT];                              // MODE 1: weighted counts
#pragma unroll
    for (int mt = 0; mt < NT; ++mt) {
      cnt[mt] = 0; wcnt[mt] = 0;
#pragma unroll
      for (int nd = 0; nd < NDX; ++nd) { acc[mt][nd][0] = 0; acc[mt][nd][1] = 0; }
    }
    const bool has_w = p.w != nullptr;
    double cost = 0, xxs = 0;
    const int xx_mode = p.xx_mode;
    const uint32_t a_frag = (uint32_t)g * cstride + (uint32_t)t * 8u;        // score A-fragment: chunk g, feature t (+4s)
    const uint32_t b_frag = (uint32_t)t * cstride + (uint32_t)g * 8u;        // reduce B-fragment: chunk t (+4 per k-step), feature g (+8nd)
    const int i0 = warp, i1 = warp + 8;
    const int t0 = g * kChunkRows + i0, t1 = g * kChunkRows + i1;            // tile rows of this lane's score rows

    auto process_tile = [&](auto need_xx_tag, const uint32_t xs, const int rows_valid, const int64_t row0) {
      constexpr bool NEED_XX = decltype(need_xx_tag)::value;
      // silhouette passes: the given labels / weights of this lane's two rows are fetched first, so their global-memory
      // latency runs under the score tiles
      int lin0 = -1, lin1 = -1; double win0 = 1.0, win1 = 1.0;
      if constexpr (MODE == 1 || MODE == 2) {
        if (t0 < rows_valid) { lin0 = __ldg(p.labels_in + row0 + t0) - (MODE == 1 ? p.label_base : 0); if (has_w) win0 = __ldg(p.w + row0 + t0); }
        if (t1 < rows_valid) { lin1 = __ldg(p.labels_in + row0 + t1) - (MODE == 1 ? p.label_base : 0); if (has_w) win1 = __ldg(p.w + row0 + t1); }
        if (lin0 >= k) lin0 = -1;
        if (lin1 >= k) lin1 = -1;
      }
      // ---------------- score tiles
      const uint32_t xa0 = xs + a_frag + (uint32_t)i0 * row_bytes, xa1 = xs + a_frag + (uint32_t)i1 * row_bytes;
      double s0[NT][2], s1[NT][2];
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) { s0[nt][0] = cinit[nt][0]; s0[nt][1] = cinit[nt][1]; s1[nt][0] = cinit[nt][0]; s1[nt][1] = cinit[nt][1]; }
      double xx0 = 0, xx1 = 0;
#pragma unroll
      for (int s = 0; s < KS_T; ++s) {
        const bool ok = (s < KS_T - 1) || (4 * s + t < d);
        const double a0 = ok ? lds_f64(xa0 + 32u * s) : 0.0, a1 = ok ? lds_f64(xa1 + 32u * s) : 0.0;
        if (NEED_XX) { xx0 = fma(a0, a0, xx0); xx1 = fma(a1, a1, xx1); }
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) { if (MODE != 1) dmma1684(s0[nt][0], s0[nt][1], s1[nt][0], s1[nt][1], a0, a1, breg[s * NT + nt]); }
      }
      if constexpr (MODE == 2) {
        // ---------------- silhouette score (ClusteringMetrics.pointSilhouetteCoefficient), one value per row
        const bool v0 = t0 < rows_valid, v1 = t1 < rows_valid;
        const int a0l = lin0, a1l = lin1;
        const double ninf = -__longlong_as_double(0x7ff0000000000000LL);
        double own0 = ninf, own1 = ninf, oth0 = ninf, oth1 = ninf;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int j = 8 * nt + 2 * t + e;
            if (j == a0l) own0 = s0[nt][e]; else oth0 = fmax(oth0, s0[nt][e]);
            if (j == a1l) own1 = s1[nt][e]; else oth1 = fmax(oth1, s1[nt][e]);
          }
#pragma unroll
        for (int msk = 1; msk <= 2; msk <<= 1) {
          own0 = fmax(own0, shfl_xor_d(own0, msk)); own1 = fmax(own1, shfl_xor_d(own1, msk));
          oth0 = fmax(oth0, shfl_xor_d(oth0, msk)); oth1 = fmax(oth1, shfl_xor_d(oth1, msk));
          xx0 += shfl_xor_d(xx0, msk); xx1 += shfl_xor_d(xx1, msk);
        }
        if (t < 2) {   // lane t = 0 finishes row t0, lane t = 1 row t1: one pass through the fp64 divisions instead of two
          auto point = [&](bool valid, int a, double xx, double own, double oth, double wv) {
            if (!valid || a < 0) return;
            const double nrm = sqrt(xx), sn = nrm * nrm;
            const double NA = cn_s[k_pad + a];
            double sil = 0.0;
            if (NA != wv) {                                  // single-element clusters have silhouette 0
              const double cur = fma(-2.0, own, sn) * NA / (NA - wv);
              const double nb = fma(-2.0, oth, sn);
              if (cur < nb) sil = 1.0 - (cur / nb); else if (cur > nb) sil = (nb / cur) - 1.0;
            }
            cost = fma(sil, wv, cost); xxs += wv;
          };
          point(t ? v1 : v0, t ? a1l : a0l, t ? xx1 : xx0, t ? own1 : own0, t ? oth1 : oth0, t ? win1 : win0);
        }
        return;
      }
      if constexpr (MODE == 1) {
        // ---------------- silhouette statistics: labels and weights are given, phase 1 only supplied |x|^2
#pragma unroll
        for (int msk = 1; msk <= 2; msk <<= 1) { xx0 += shfl_xor_d(xx0, msk); xx1 += shfl_xor_d(xx1, msk); }
        const int l0 = lin0, l1 = lin1;
        const double w0 = win0, w1 = win1;
        const double n0 = sqrt(xx0), n1 = sqrt(xx1);           // squaredNorm = math.pow(Vectors.norm(features, 2.0), 2.0)
        xx0 = n0 * n0; xx1 = n1 * n1;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const int src = (t + 4 * (ks & 1)) * 4;
          const int labt = __shfl_sync(0xffffffffu, (ks < 2) ? l0 : l1, src);
          const double wt = __shfl_sync(0xffffffffu, (ks < 2) ? w0 : w1, src);
          const double xxt = __shfl_sync(0xffffffffu, (ks < 2) ? xx0 : xx1, src);
          double a2[NT];
#pragma unroll
          for (int mt = 0; mt < NT; ++mt) { a2[mt] = (labt == g + 8 * mt) ? wt : 0.0; wcnt[mt] += a2[mt]; }
          const uint32_t xb = xs + b_frag + (uint32_t)(4 * (ks & 1)) * cstride + (uint32_t)((ks < 2) ? i0 : i1) * row_bytes;
#pragma unroll
          for (int nd = 0; nd < ND; ++nd) {
            const bool ok = (nd < ND - 1) || (8 * nd + g < d);
            const double b2 = ok ? lds_f64(xb + 64u * nd) : 0.0;
#pragma unroll
            for (int mt = 0; mt < NT; ++mt) dmma884(acc[mt][nd][0], acc[mt][nd][1], a2[mt], b2);
          }
          const double bx = g == 0 ? xxt : 0.0;                // extra feature tile: column 0 carries |x|^2 -> Psi
#pragma unroll
          for (int mt = 0; mt < NT; ++mt) dmma884(acc[mt][NDX - 1][0], acc[mt][NDX - 1][1], a2[mt], bx);
        }
        return;
      }
      // ---------------- argmax (lowest index wins ties)
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
      const bool v0 = t0 < rows_valid, v1 = t1 < rows_valid;
      if (!v0) j0 = -1;
      if (!v1) j1 = -1;
      if (t == 0) {
        double d0 = b0, d1 = b1;
        if (NEED_XX) { d0 = fmax(fma(-2.0, b0, xx0), 0.0); d1 = fmax(fma(-2.0, b1, xx1), 0.0); }
        const double c0 = (NEED_XX && xx_mode == 2) ? b0 : d0, c1 = (NEED_XX && xx_mode == 2) ? b1 : d1;
        if (v0) cost += c0;
        if (v1) cost += c1;
        if (NEED_XX && xx_mode == 1) { if (v0) xxs += xx0; if (v1) xxs += xx1; }
        if (p.labels_out != nullptr) { if (v0) p.labels_out[row0 + t0] = j0; if (v1) p.labels_out[row0 + t1] = j1; }
        if (NEED_XX && p.mind2_out != nullptr) { if (v0) p.mind2_out[row0 + t0] = d0; if (v1) p.mind2_out[row0 + t1] = d1; }
      }
      // ---------------- reduce: S += onehot^T * X, 4 k-steps of 4 rows (rows t, t+4 of sub-tile i0, then of i1)
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        const int jsrc = (ks < 2) ? j0 : j1;
        const int labt = __shfl_sync(0xffffffffu, jsrc, (t + 4 * (ks & 1)) * 4);   // label of MMA row t + 4*(ks&1)
        double a2[NT];
#pragma unroll
        for (int mt = 0; mt < NT; ++mt) { const bool hit = labt == g + 8 * mt; a2[mt] = hit ? 1.0 : 0.0; cnt[mt] += hit ? 1u : 0u; }
        const uint32_t xb = xs + b_frag + (uint32_t)(4 * (ks & 1)) * cstride + (uint32_t)((ks < 2) ? i0 : i1) * row_bytes;
#pragma unroll
        for (int nd = 0; nd < ND; ++nd) {
          const bool ok = (nd < ND - 1) || (8 * nd + g < d);
          const double b2 = ok ? lds_f64(xb + 64u * nd) : 0.0;
#pragma unroll
          for (int mt = 0; mt < NT; ++mt) dmma884(acc[mt][nd][0], acc[mt][nd][1], a2[mt], b2);
        }
      }
    };

    {
      int s = 0; uint32_t ph = 0;
      int64_t row0 = (int64_t)blockIdx.x * kTileRows;
      const int64_t row_step = (int64_t)gridDim.x * kTileRows;
      for (int i = 0; i < my_tiles; ++i) {
        mbar_wait_a(a_full + 8u * (uint32_t)s, ph);
        const uint32_t xs = a_x + (uint32_t)s * sbytes;
        if (xx_mode != 2) process_tile(std::true_type{}, xs, kTileRows, row0);
        else process_tile(std::false_type{}, xs, kTileRows, row0);
        __syncwarp();
        if (lane == 0) mbar_arrive(reinterpret_cast<uint64_t*>(smem + L.bar_off) + kMaxStages + s);
        if (++s == stages) { s = 0; ph ^= 1u; }
        row0 += row_step;
      }
      // ragged tail (< 128 rows): the CTA next in round-robin order loads it with plain loads
      if (tail_rows > 0 && (int)(full_tiles % gridDim.x) == (int)blockIdx.x) {
        named_bar_sync(1, kComputeThreads);  // every compute warp is done with all stages (the producer has no work left)
        unsigned char* xsp = smem + L.x_off;
        const double* src = p.X + (size_t)full_tiles * kTileRows * d;
        for (int i = tid; i < kTileRows * d; i += kComputeThreads) {
          int tr = i / d, f = i - tr * d;
          double v = tr < tail_rows ? src[(size_t)tr * d + f] : 0.0;
          reinterpret_cast<double*>(xsp + (tr >> 4) * cstride)[(tr & 15) * d + f] = v;
        }
        named_bar_sync(1, kComputeThreads);
        process_tile(std::true_type{}, a_x, tail_rows, full_tiles * kTileRows);
      }
    }

    // ===================================================================== CTA-level combine (compute warps only)
    const int P = k * d + k + 1;
    const int S = P + 1 + k;
    double* out = p.cta_partials + (size_t)blockIdx.x * S;
#pragma unroll
    for (int msk = 16; msk >= 1; msk >>= 1) { cost += shfl_xor_d(cost, msk); xxs += shfl_xor_d(xxs, msk); }
    if (lane == 0) { red_s[warp] = cost; red_s[kComputeWarps + warp] = xxs; }
    // counts: every lane of a (g, *) group saw the same hits per k-step? no -- lane (g, t) counted row t only: fold over t
#pragma unroll
    for (int mt = 0; mt < NT; ++mt) {
      cnt[mt] += __shfl_xor_sync(0xffffffffu, cnt[mt], 1); cnt[mt] += __shfl_xor_sync(0xffffffffu, cnt[mt], 2);
      if (MODE == 1) { wcnt[mt] += shfl_xor_d(wcnt[mt], 1); wcnt[mt] += shfl_xor_d(wcnt[mt], 2); } else wcnt[mt] = (double)cnt[mt];
    }
    // sums: warps take turns adding their fragments into a [k_pad][d_pad8] buffer that aliases stage 0
    named_bar_sync(1, kComputeThreads);
    double* comb = reinterpret_cast<double*>(smem + L.x_off);       // [k_pad][8*ND] sums, then [k_pad] counts
    const int ldc = 8 * NDX;
    double* ccnt = comb + k_pad * ldc;
    for (int w = 0; w < kComputeWarps; ++w) {
      if (warp == w) {
#pragma unroll
        for (int mt = 0; mt < NT; ++mt) {
#pragma unroll
          for (int nd = 0; nd < NDX; ++nd) {
            double* c2 = comb + (8 * mt + g) * ldc + 8 * nd + 2 * t;
            if (w == 0) { c2[0] = acc[mt][nd][0]; c2[1] = acc[mt][nd][1]; }
            else { c2[0] += acc[mt][nd][0]; c2[1] += acc[mt][nd][1]; }
          }
          if (t == 0) { if (w == 0) ccnt[8 * mt + g] = wcnt[mt]; else ccnt[8 * mt + g] += wcnt[mt]; }
        }
      }
      named_bar_sync(1, kComputeThreads);
    }
    for (int i = tid; i < k * d; i += kComputeThreads) { const int j = i / d, f = i - j * d; out[i] = comb[j * ldc + f]; }
    for (int j = tid; j < k; j += kComputeThreads) { out[(size_t)k * d + j] = ccnt[j]; if (MODE == 1) out[P + 1 + j] = comb[j * ldc + 8 * ND]; }
    if (tid == 0) {
      double c = 0, x2 = 0;
#pragma unroll
      for (int wv = 0; wv < kComputeWarps; ++wv) { c += red_s[wv]; x2 += red_s[kComputeWarps + wv]; }
      out[P - 1] = c; out[P] = x2;
    }
    // ===================================================================== last CTA folds all partials
    __threadfence();
    named_bar_sync(1, kComputeThreads);
    __shared__ unsigned int s_ticket;
    if (tid == 0) s_ticket = atomicAdd(p.counter, 1u);
    named_bar_sync(1, kComputeThreads);
    if (s_ticket == gridDim.x - 1) {
      __threadfence();
      const int i_lo = MODE == 2 ? P - 1 : 0, i_hi = MODE == 1 ? S : P + 1;
      for (int i = i_lo + tid; i < i_hi; i += kComputeThreads) {
        double s = 0;
        s = ordered_sum_ldcg(p.cta_partials + i, (size_t)S, gridDim.x);
        if (i == P - 1 && xx_mode == 2) s = fmax(fma(-2.0, s, *p.xnorm2_sum), 0.0);  // cost = sum |x|^2 - 2 sum score
        if (MODE == 2) p.partial_out[i - (P - 1)] = s;           // [sum s*w, sum w]
        else if (i < P) p.partial_out[i] = s;
        else if (i == P) { if (xx_mode == 1) *p.xnorm2_sum = s; }
        else p.partial_out[i - 1] = s;                           // MODE 1: psi follows [sums][N][cost]
      }
      if (tid == 0) *p.counter = 0u;
      bool peer_ok = true;
      if (p.peer_bufs != nullptr) {                            // ranks > 1: sum the vector over NVLink peer memory right here
        __threadfence_block();
        named_bar_sync(1, kComputeThreads);
        peer_ok = peer_allreduce(p, p.partial_out, P, tid, kComputeThreads, [] { named_bar_sync(1, kComputeThreads); });
      }
      if (p.C_update != nullptr && peer_ok) {                             // fused driver step: compute warps only
        __threadfence_block();
        named_bar_sync(1, kComputeThreads);
        double shift = -1.0;
        if (tid < k) {
          const double wsum = p.partial_out[(size_t)k * d + tid];
          if (wsum > 0) {
            const double inv = 1.0 / wsum;
            double sft = 0;
            for (int f = 0; f < d; ++f) {
              const double nc = p.partial_out[(size_t)tid * d + f] * inv;
              const double df = nc - p.C_update[(size_t)tid * d + f];
              sft = fma(df, df, sft);
              p.C_update[(size_t)tid * d + f] = nc;
            }
            shift = sft;
          }
        }
        if (tid < 16) red_s[tid] = shift;
        named_bar_sync(1, kComputeThreads);
        if (tid == 0) {
          double mx = -1.0;
          for (int i = 0; i < k; ++i) mx = fmax(mx, red_s[i]);
          p.state_rw->cost = p.partial_out[(size_t)k * d + k];
          p.state_rw->iter += 1;
          p.state_rw->max_shift2 = mx < 0 ? 0.0 : mx;
          p.state_rw->converged = (mx <= p.tol2) ? 1 : 0;
        }
      }
    }
  }
}

}  // namespace hkm

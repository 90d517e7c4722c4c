// Lloyd data pass, large-k variant (24 < k <= 64, even 32 < d <= 64): tcgen05 tensor-core screening + fp64 refinement.
//
// For k*d this large the fp64 pipe cannot keep up with HBM (k = 64, d = 64 needs 4096 fp64 FMA per 512-byte row), and
// tcgen05 has no fp64 kind.  So the k distances of a row are first computed APPROXIMATELY on the 5th-gen tensor cores
// (kind::tf32, operands staged by the CUDA cores from the fp64 tile into the canonical UMMA shared-memory layout,
// accumulator in TMEM, read back with tcgen05.ld), together with a rigorous per-row error bound
//     |approx_j - exact_j| <= eps = 2^-8 * |x - mu| * max_j |c_j - mu|          (tf32 keeps 10 mantissa bits)
// where mu is the mean of the centres (both operands are centred by mu: the argmax is translation invariant and the
// bound shrinks with it).  Only centres with approx_j >= max_j approx_j - 2 eps can be the true argmax; those (usually
// one, sometimes two or three) are re-scored exactly in fp64 from the fp64 row still sitting in shared memory, with the
// same score x.c - 0.5|c|^2, the same lowest-index tie rule and the same cost as the fp64 kernels.  The per-cluster
// sums use the gather scheme of hkm_lloyd.cuh (register accumulators owned per warp, no atomics).
//
// One CTA per SM: warp 8 is the TMA producer, warps 0-7 convert / gather / refine, thread 0 issues the UMMAs.
#pragma once
#include "hkm_lloyd.cuh"

namespace hkm {

constexpr int kTcThreads = kComputeThreads + 32;
constexpr int kTcLboPad = 16;   // K-chunk stride = groups*128 + 16 bytes: spreads the conversion stores over the banks

struct TcSmemLayout { int x_off, a_off, b_off, cs_off, mu_off, h_off, cn_off, xn_off, labels_off, pair_off, list_off, zero_off, red_off, bar_off, tmem_off, total; };
__host__ __device__ inline int tc_lbo_a() { return (kTileRows / 8) * 128 + kTcLboPad; }
__host__ __device__ inline int tc_lbo_b(int k_pad) { return (k_pad / 8) * 128 + kTcLboPad; }
__host__ __device__ inline TcSmemLayout tc_smem_layout(int d, int k_pad, int stages) {
  TcSmemLayout L; int o = 0;
  const int kchunks = ((d + 7) / 8) * 2;                       // 16-byte tf32 chunks along K (K padded to 8)
  L.x_off = o; o += stages * stage_bytes(d); o = (o + 1023) & ~1023;
  L.a_off = o; o += kchunks * tc_lbo_a(); o = (o + 127) & ~127;
  L.b_off = o; o += kchunks * tc_lbo_b(k_pad); o = (o + 127) & ~127;
  L.cs_off = o; o += k_pad * 64 * 8;                           // fp64 per-cluster sums of this CTA, dense 512-byte rows (+ counts below)
  L.mu_off = o; o += 64 * 8;
  L.h_off = o; o += k_pad * 4;                                 // float -0.5|c-mu|^2
  L.cn_off = o; o += k_pad * 8;                                // double -0.5|c|^2
  L.xn_off = o; o += kTileRows * 4;                            // float |x-mu|^2 per tile row
  L.labels_off = o; o += 2 * kTileRows * 4;
  L.pair_off = o; o += kTileRows * 2 * 16;                     // per (row, half): {double score, int idx}
  L.list_off = o; o += kComputeWarps * list_entries(k_pad) * 4; o = (o + 15) & ~15;
  L.zero_off = o; o += 64 * 8;                                 // per-cluster row counts (unsigned, k_pad <= 64 entries)
  L.red_off = o; o += 2 * kComputeWarps * 8 + 64 * 8;
  L.bar_off = o; o += (2 * kMaxStages + 2) * 8;
  L.tmem_off = o; o += 16;
  L.total = o;
  return L;
}

__device__ __forceinline__ uint64_t umma_desc_kmajor(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  // K-major, no swizzle: element (row r, 16-byte chunk c) sits at (r % 8) * 16 + (r / 8) * SBO + c * LBO
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32) |
         ((uint64_t)1 << 46);
}

// NT = k_pad / 8 (4 or 8), FPL = 2 (32 < d <= 64) or 1 (d <= 32)
template <int NT, int FPL>
__global__ void __launch_bounds__(kTcThreads, 1) lloyd_tc_kernel(const LloydParams p) {
  if (p.skip != nullptr && *p.skip != 0) return;
  if (p.skip2 != nullptr && *p.skip2 != 0) return;
  constexpr int k_pad = NT * 8;
  constexpr int HALF = k_pad / 2;               // centres scanned per thread in the epilogue

  extern __shared__ __align__(1024) unsigned char smem[];
  const int d = p.d, k = p.k;
  const TcSmemLayout L = tc_smem_layout(d, k_pad, p.stages);
  double* Cs = reinterpret_cast<double*>(smem + L.cs_off);
  double* mu_s = reinterpret_cast<double*>(smem + L.mu_off);
  float* h_s = reinterpret_cast<float*>(smem + L.h_off);
  double* cn_s = reinterpret_cast<double*>(smem + L.cn_off);
  double* red_s = reinterpret_cast<double*>(smem + L.red_off);
  const uint32_t sb = smem_u32(smem);
  const uint32_t a_x = sb + L.x_off, a_A = sb + L.a_off, a_B = sb + L.b_off, a_cs = sb + L.cs_off, a_xn = sb + L.xn_off;
  const uint32_t a_labels = sb + L.labels_off, a_pair = sb + L.pair_off;
  const uint32_t a_full = sb + L.bar_off, a_empty = a_full + 8u * kMaxStages, a_mma = a_full + 16u * kMaxStages;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int stages = p.stages;
  const uint32_t cstride = (uint32_t)chunk_stride_bytes(d);
  const uint32_t sbytes = (uint32_t)stage_bytes(d);
  const uint32_t row_bytes = (uint32_t)d * 8u;
  const int kchunks = ((d + 7) / 8) * 2;
  const uint32_t LBO_A = (uint32_t)tc_lbo_a(), LBO_B = (uint32_t)tc_lbo_b(k_pad);
  const int64_t full_tiles = p.n / kTileRows;
  const int tail_rows = (int)(p.n - full_tiles * kTileRows);
  const int my_tiles = full_tiles > (int64_t)blockIdx.x ? (int)((full_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;

  // ------------------------------------------------------------------ one-time set-up
  if (tid == 0) {
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.bar_off);
    for (int s = 0; s < stages; ++s) { mbar_init(bars + s, 1); mbar_init(bars + kMaxStages + s, kComputeWarps); }
    mbar_init(bars + 2 * kMaxStages, 1);
    mbar_fence_init();
  }
  if (warp == kComputeWarps) {   // the producer warp owns the TMEM allocation
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(sb + L.tmem_off) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  for (int i = tid; i < k_pad * 64; i += kTcThreads) Cs[i] = 0.0;             // Cs doubles as the per-cluster accumulator block
  for (int i = tid; i < 64; i += kTcThreads) reinterpret_cast<unsigned int*>(smem + L.zero_off)[i] = 0u;
  if (tid == 0) *reinterpret_cast<unsigned int*>(smem + L.tmem_off + 8) = 0u;   // ambiguous-row counter
  if (tid < 64) { double s = 0; if (tid < d) for (int j = 0; j < k; ++j) s += p.C[(size_t)j * d + tid]; mu_s[tid] = tid < d ? s / k : 0.0; }   // mu = mean of the centres
  __syncthreads();
  if (tid < k_pad) {
    double s = 0, sc = 0;
    if (tid < k) for (int f = 0; f < d; ++f) { const double c = p.C[(size_t)tid * d + f], cc = c - mu_s[f]; s = fma(c, c, s); sc = fma(cc, cc, sc); }
    cn_s[tid] = tid < k ? -0.5 * s : -__longlong_as_double(0x7ff0000000000000LL);
    h_s[tid] = tid < k ? (float)(-0.5 * sc) : -__int_as_float(0x7f800000);
    red_s[2 * kComputeWarps + tid] = tid < k ? sc : 0.0;
  }
  // B operand: tf32(c - mu), canonical K-major layout; padded centres / features are zero
  for (int i = tid; i < k_pad * kchunks; i += kTcThreads) {
    const int j = i % k_pad, c = i / k_pad;
    float vv[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) { const int f = 4 * c + e; vv[e] = (j < k && f < d) ? (float)(p.C[(size_t)j * d + f] - mu_s[f]) : 0.f; }
    *reinterpret_cast<float4*>(smem + L.b_off + c * LBO_B + (j >> 3) * 128 + (j & 7) * 16) = make_float4(vv[0], vv[1], vv[2], vv[3]);
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *reinterpret_cast<const uint32_t*>(smem + L.tmem_off);
  float cmax2 = 0.f;   // max_j |c_j - mu|^2
  for (int j = 0; j < k; ++j) cmax2 = fmaxf(cmax2, (float)red_s[2 * kComputeWarps + j]);
  const float cmax = sqrtf(cmax2) * 1.0001f;

  // ===================================================================== producer warp
  if (warp == kComputeWarps) {
    if (lane == 0) {
      const uint32_t cbytes = (uint32_t)kChunkRows * row_bytes;
      const size_t tile_gbytes = (size_t)kTileRows * row_bytes;
      const unsigned char* src = reinterpret_cast<const unsigned char*>(p.X) + (size_t)blockIdx.x * tile_gbytes;
      const size_t src_step = (size_t)gridDim.x * tile_gbytes;
      int s = 0; uint32_t ph = 0;
      for (int i = 0; i < my_tiles; ++i) {
        while (!mbar_test_a(a_empty + 8u * (uint32_t)s, ph ^ 1u)) __nanosleep(128);   // do not spin on the issue port the compute warps need
        const uint32_t bar = a_full + 8u * (uint32_t)s, dst = a_x + (uint32_t)s * sbytes;
        mbar_expect_tx_a(bar, cbytes * kChunks);
#pragma unroll
        for (int c = 0; c < kChunks; ++c) bulk_g2s_a(dst + c * cstride, src + (size_t)c * cbytes, cbytes, bar);
        src += src_step;
        if (++s == stages) { s = 0; ph ^= 1u; }
      }
    }
  } else {
    // ===================================================================== compute warps
    const int fl = lane;
    const uint32_t fl8 = (uint32_t)fl * 8u;
    // gather: warp w owns clusters w, w+8, ... (slot m = label >> 3); lane = feature (and feature + 32)
    double cost = 0, xxs = 0;
    const uint32_t a_acc = a_cs + fl8;                       // per-cluster sums in shared memory: row j at j*512, this lane's feature
    const int xx_mode = p.xx_mode;
    // per-row distances wanted (first Lloyd pass, mind2_out): every row is re-scored exactly; label-only passes
    // (predict) and steady-state Lloyd passes only re-score the ambiguous rows
    const bool refine_all = p.mind2_out != nullptr || (p.do_reduce == 1 && xx_mode != 2);
    const uint32_t group_stride = 32u * row_bytes + 2u * kChunkPad;
    // conversion / cooperative refine: lane l owns features 2l, 2l+1 of the row the warp is working on
    const bool conv_ok = 2 * lane < d;
    const double mu0 = conv_ok ? mu_s[2 * lane] : 0.0, mu1 = conv_ok ? mu_s[2 * lane + 1] : 0.0;
    const uint32_t a_conv = a_A + (uint32_t)(lane >> 1) * LBO_A + (uint32_t)(lane & 1) * 8u;   // + group*128 + (row&7)*16
    // screening: row = 32*(warp&3) + lane, centre half = warp>>2
    const int erow = 32 * (warp & 3) + lane, hf = warp >> 2;
    const uint32_t taddr = tmem + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)(hf * HALF);
    const uint32_t a_amb = sb + L.list_off;                 // ambiguous-row list: {row, mask_lo, mask_hi, -} per entry
    unsigned int* amb_count = reinterpret_cast<unsigned int*>(smem + L.tmem_off + 8);

    auto convert_tile = [&](const uint32_t xs) {
      // rows 8g .. 8g+7 form row group g of the A operand; warp w converts groups w and w+8
#pragma unroll
      for (int gg = 0; gg < 2; ++gg) {
        const int g = warp + 8 * gg;
        float part[8];
        double2 vin[8];
        // rows 8g .. 8g+7 sit in one 16-row chunk, row_bytes apart: all eight loads go out before the first use
        const uint32_t ra0 = xs + (uint32_t)((8 * g) >> 4) * cstride + (uint32_t)((8 * g) & 15) * row_bytes + 16u * lane;
#pragma unroll
        for (int rr = 0; rr < 8; ++rr) {
          vin[rr] = make_double2(mu0, mu1);
          if (conv_ok) asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(vin[rr].x), "=d"(vin[rr].y) : "r"(ra0 + (uint32_t)rr * row_bytes));
        }
#pragma unroll
        for (int rr = 0; rr < 8; ++rr) {
          float2 o = make_float2(0.f, 0.f);
          if (conv_ok) { o.x = (float)(vin[rr].x - mu0); o.y = (float)(vin[rr].y - mu1); }
          part[rr] = fmaf(o.x, o.x, o.y * o.y);
          if (2 * lane < kchunks * 4)
            asm volatile("st.shared.v2.f32 [%0], {%1,%2};" ::"r"(a_conv + (uint32_t)g * 128u + (uint32_t)rr * 16u), "f"(o.x), "f"(o.y) : "memory");
        }
        // |x - mu|^2 of the 8 rows: reduce-scatter over the 32 lanes (8 -> 4 -> 2 -> 1 values, then 2 plain rounds)
#pragma unroll
        for (int half = 4, msk = 16; half >= 1; half >>= 1, msk >>= 1) {
          const bool up = (lane & msk) != 0;
#pragma unroll
          for (int i = 0; i < half; ++i) {
            const float send = up ? part[i] : part[i + half], keep = up ? part[i + half] : part[i];
            part[i] = keep + __shfl_xor_sync(0xffffffffu, send, msk);
          }
        }
        part[0] += __shfl_xor_sync(0xffffffffu, part[0], 2);
        part[0] += __shfl_xor_sync(0xffffffffu, part[0], 1);
        if ((lane & 3) == 0) {
          const int rr = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
          asm volatile("st.shared.f32 [%0], %1;" ::"r"(a_xn + 4u * (uint32_t)(8 * g + rr)), "f"(part[0]) : "memory");
        }
      }
    };

    // per-cluster sums: one ballot per 32 rows finds this warp's rows (label & 7 == warp, so no two warps ever touch
    // the same cluster); each row is added to its cluster's shared-memory accumulator row, one feature (pair) per lane.
    // Two rows are in flight per step; equal labels are merged first so the read-modify-write stays race free.
    unsigned gcnt0 = 0, gcnt1 = 0;
    auto gather_tile = [&](const uint32_t xs, const uint32_t lab_b) {
#pragma unroll 1
      for (int g = 0; g < kTileRows / 32; ++g) {
        const int lab = (int)lds_u32(lab_b + 4u * (g * 32 + lane));
        unsigned mask = __ballot_sync(0xffffffffu, lab >= 0 && (lab & 7) == warp);
        const uint32_t gbase = xs + (uint32_t)g * group_stride + fl8;
        while (mask) {
          const int r0 = __ffs(mask) - 1; mask &= mask - 1;
          bool two = mask != 0;
          const int r1 = two ? __ffs(mask) - 1 : r0; mask &= mask - 1;
          const int l0 = __shfl_sync(0xffffffffu, lab, r0), l1 = __shfl_sync(0xffffffffu, lab, r1);
          const uint32_t ra0 = gbase + (uint32_t)r0 * row_bytes + (uint32_t)(r0 >> 4) * kChunkPad;
          const uint32_t ra1 = gbase + (uint32_t)r1 * row_bytes + (uint32_t)(r1 >> 4) * kChunkPad;
          double v0[FPL], v1[FPL];
#pragma unroll
          for (int e = 0; e < FPL; ++e) { v0[e] = lds_f64(ra0 + 256u * e); v1[e] = two ? lds_f64(ra1 + 256u * e) : 0.0; }   // lanes past d touch padding columns that are never stored
          unsigned n0 = 1;
          if (two && l1 == l0) {
#pragma unroll
            for (int e = 0; e < FPL; ++e) v0[e] += v1[e];
            two = false; n0 = 2;
          }
          const uint32_t aa0 = a_acc + (uint32_t)l0 * 512u, aa1 = a_acc + (uint32_t)l1 * 512u;
          double s0[FPL], s1[FPL];
#pragma unroll
          for (int e = 0; e < FPL; ++e) { s0[e] = lds_f64(aa0 + 256u * e); if (two) s1[e] = lds_f64(aa1 + 256u * e); }
#pragma unroll
          for (int e = 0; e < FPL; ++e) { sts_f64(aa0 + 256u * e, s0[e] + v0[e]); if (two) sts_f64(aa1 + 256u * e, s1[e] + v1[e]); }
          // rows per cluster, in registers: lane c counts clusters c and c + 32 (this warp only ever sees its own clusters)
          gcnt0 += (l0 == lane ? n0 : 0u) + ((two && l1 == lane) ? 1u : 0u);
          gcnt1 += (l0 == lane + 32 ? n0 : 0u) + ((two && l1 == lane + 32) ? 1u : 0u);
        }
      }
    };

    // exact fp64 scores of one row against a set of candidate centres, cooperatively: lane l holds x[2l], x[2l+1]
    auto refine_row = [&](const uint32_t xs, const int row, uint32_t mlo, uint32_t mhi, const bool valid, const int64_t r0, const uint32_t lab_b,
                          const bool need_xx) {
      const uint32_t ra = xs + (uint32_t)(row >> 4) * cstride + (uint32_t)(row & 15) * row_bytes + 16u * lane;
      double2 x = make_double2(0.0, 0.0);
      if (conv_ok) asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(x.x), "=d"(x.y) : "r"(ra));
      double best = -__longlong_as_double(0x7ff0000000000000LL); int bj = -1;
      for (int h = 0; h < 2; ++h) {
        uint32_t msk = h == 0 ? mlo : mhi;
        while (msk) {                                            // ascending index: strict '>' keeps the lowest index on ties
          const int j = h * HALF + __ffs(msk) - 1; msk &= msk - 1;
          double2 c = make_double2(0.0, 0.0);
          if (conv_ok) c = __ldg(reinterpret_cast<const double2*>(p.C + (size_t)j * d) + lane);   // centres stay in L1/L2; only ambiguous rows get here
          double sdot = fma(x.x, c.x, x.y * c.y);
#pragma unroll
          for (int mk = 16; mk >= 1; mk >>= 1) sdot += shfl_xor_d(sdot, mk);
          const double e = sdot + cn_s[j];
          if (e > best) { best = e; bj = j; }
        }
      }
      double xx = 0;
      if (need_xx) {
        xx = fma(x.x, x.x, x.y * x.y);
#pragma unroll
        for (int mk = 16; mk >= 1; mk >>= 1) xx += shfl_xor_d(xx, mk);
      }
      if (lane == 0) {
        sts_u32(lab_b + 4u * (uint32_t)row, (uint32_t)(valid ? bj : -1));
        if (valid && refine_all) {
          const double dist = fmax(fma(-2.0, best, xx), 0.0);
          cost += dist;                                          // only xx_mode 0 / 1 reach here with refine_all set by xx_mode
          if (xx_mode == 1) xxs += xx;
          if (p.labels_out != nullptr) p.labels_out[r0 + row] = bj;
          if (p.mind2_out != nullptr) p.mind2_out[r0 + row] = dist;
        } else if (valid && p.labels_out != nullptr) p.labels_out[r0 + row] = bj;
      }
    };

    auto epilogue_tile = [&](const uint32_t xs, const int rows_valid, const int64_t r0, const uint32_t lab_b) {
      uint32_t v[HALF];
      if constexpr (HALF == 32) {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
                       "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
                       "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
                       "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                     : "r"(taddr));
      } else {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
                       "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                     : "r"(taddr));
      }
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      // approximate scores of this thread's half, their maximum
      float sc[HALF];
      float mx = -__int_as_float(0x7f800000);
#pragma unroll
      for (int c = 0; c < HALF; ++c) { sc[c] = __uint_as_float(v[c]) + h_s[hf * HALF + c]; mx = fmaxf(mx, sc[c]); }
      const uint32_t pair_me = a_pair + (uint32_t)(erow * 2 + hf) * 16u, pair_ot = a_pair + (uint32_t)(erow * 2 + (hf ^ 1)) * 16u;
      asm volatile("st.shared.f32 [%0], %1;" ::"r"(pair_me), "f"(mx) : "memory");
      named_bar_sync(1, kComputeThreads);
      float omx;
      asm volatile("ld.shared.f32 %0, [%1];" : "=f"(omx) : "r"(pair_ot));
      const float gmax = fmaxf(mx, omx);
      float xn2;
      asm volatile("ld.shared.f32 %0, [%1];" : "=f"(xn2) : "r"(a_xn + 4u * (uint32_t)erow));
      // tf32 keeps 10 mantissa bits: |approx - exact| <= 2^-8 |x-mu| max|c-mu| (factor 2 of slack) + fp32 rounding of the sums
      const float eps = 0.00390625f * sqrtf(xn2) * 1.0001f * cmax + 4e-6f * (fabsf(gmax) + 1.0f);
      const float thr = gmax - 2.0f * eps;
      uint32_t cand = 0;
#pragma unroll
      for (int c = 0; c < HALF; ++c) cand |= (sc[c] >= thr && hf * HALF + c < k) ? (1u << c) : 0u;
      sts_u32(pair_me + 4u, cand);
      named_bar_sync(1, kComputeThreads);
      const bool valid = erow < rows_valid;
      if (!refine_all && hf == 1) {
        // steady state: a row with a single candidate is decided here; its exact score reaches the cost through the
        // gather (sum_j S_j . c_j).  Rows with two or more candidates go to the cooperative exact re-scoring below.
        const uint32_t olo = lds_u32(pair_ot + 4u);
        const int total = __popc(olo) + __popc(cand);
        if (!valid) sts_u32(lab_b + 4u * (uint32_t)erow, 0xffffffffu);
        else if (total == 1) {
          const int j = olo ? (__ffs(olo) - 1) : (HALF + __ffs(cand) - 1);
          sts_u32(lab_b + 4u * (uint32_t)erow, (uint32_t)j);
          if (p.labels_out != nullptr) p.labels_out[r0 + erow] = j;
        } else {
          const unsigned slot = atomicAdd(amb_count, 1u);
          asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(a_amb + 16u * slot), "r"((uint32_t)erow), "r"(olo), "r"(cand), "r"(0u) : "memory");
        }
      }
      named_bar_sync(1, kComputeThreads);
      if (refine_all) {
        // every row of the tile, in row order (bit-stable cost): warp w takes rows w, w+8, ...
        for (int row = warp; row < kTileRows; row += kComputeWarps) {
          const uint32_t mlo = lds_u32(a_pair + (uint32_t)(row * 2) * 16u + 4u), mhi = lds_u32(a_pair + (uint32_t)(row * 2 + 1) * 16u + 4u);
          refine_row(xs, row, mlo, mhi, row < rows_valid, r0, lab_b, true);
        }
      } else {
        const unsigned namb = *reinterpret_cast<volatile unsigned int*>(amb_count);
        for (unsigned i = warp; i < namb; i += kComputeWarps) {
          const uint4 e = lds_v4(a_amb + 16u * i);
          refine_row(xs, (int)e.x, e.y, e.z, true, r0, lab_b, false);
        }
      }
      named_bar_sync(1, kComputeThreads);                                 // labels of this tile are complete
      if (tid == 0) *amb_count = 0u;
    };

    // ------------------------------------------------------------------ main loop
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(k_pad >> 3) << 17) | ((uint32_t)(kTileRows >> 4) << 24);
    auto issue_mma = [&]() {
      for (int ks = 0; ks < kchunks / 2; ++ks) {                 // one UMMA consumes K = 8 tf32 = two 16-byte chunks
        const uint64_t da = umma_desc_kmajor(a_A + (uint32_t)(2 * ks) * LBO_A, LBO_A, 128u);
        const uint64_t db = umma_desc_kmajor(a_B + (uint32_t)(2 * ks) * LBO_B, LBO_B, 128u);
        const uint32_t accf = ks > 0 ? 1u : 0u;
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem),
                     "l"(da), "l"(db), "r"(idesc), "r"(accf)
                     : "memory");
      }
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(a_mma) : "memory");
    };

    int s = 0; uint32_t ph = 0, mph = 0; int buf = 0;
    int64_t row0 = (int64_t)blockIdx.x * kTileRows;
    const int64_t row_step = (int64_t)gridDim.x * kTileRows;
    auto run_tile = [&](const uint32_t xs, const int rows_valid, const int64_t r0) {
      const uint32_t lab_b = a_labels + (uint32_t)buf * (kTileRows * 4);
      convert_tile(xs);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");       // conversion stores -> visible to the tensor core
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");   // previous tile's tcgen05.ld are done before TMEM is overwritten
      named_bar_sync(1, kComputeThreads);
      if (tid == 0) { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); issue_mma(); }
      mbar_wait_a(a_mma, mph); mph ^= 1u;
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      epilogue_tile(xs, rows_valid, r0, lab_b);
      if (p.do_reduce == 1) gather_tile(xs, lab_b);                       // per-cluster sums, then the stage can be refilled
    };
    for (int i = 0; i < my_tiles; ++i) {
      mbar_wait_a(a_full + 8u * (uint32_t)s, ph);
      const uint32_t xs = a_x + (uint32_t)s * sbytes;
      run_tile(xs, kTileRows, row0);
      __syncwarp();
      if (lane == 0) mbar_arrive(reinterpret_cast<uint64_t*>(smem + L.bar_off) + kMaxStages + s);
      if (++s == stages) { s = 0; ph ^= 1u; }
      buf ^= 1; row0 += row_step;
    }
    if (tail_rows > 0 && (int)(full_tiles % gridDim.x) == (int)blockIdx.x) {
      named_bar_sync(1, kComputeThreads);
      unsigned char* xsp = smem + L.x_off;
      const double* src = p.X + (size_t)full_tiles * kTileRows * d;
      for (int i = tid; i < kTileRows * d; i += kComputeThreads) {
        const int tr = i / d, f = i - tr * d;
        reinterpret_cast<double*>(xsp + (tr >> 4) * cstride)[(tr & 15) * d + f] = tr < tail_rows ? src[(size_t)tr * d + f] : 0.0;
      }
      named_bar_sync(1, kComputeThreads);
      run_tile(a_x, tail_rows, full_tiles * kTileRows);
    }
    if (p.do_reduce) {
      // ===================================================================== CTA-level combine (as hkm_lloyd.cuh)
      const int P = k * d + k + 1;
      const int S = P + 1 + k;
      double* out = p.cta_partials + (size_t)blockIdx.x * S;
      if ((lane & 7) == warp) {                               // this warp's clusters: w, w+8, ... (counts kept in registers by the gather)
        unsigned int* cnt_s = reinterpret_cast<unsigned int*>(smem + L.zero_off);
        cnt_s[lane] = gcnt0; cnt_s[lane + 32] = gcnt1;
      }
      named_bar_sync(1, kComputeThreads);                     // every warp's gather has landed in shared memory
      if (!refine_all) {
        // steady state: the sum of the exact winner scores needs no per-row work at all:
        //   sum_rows (x.c_label - 0.5|c_label|^2) = sum_j S_j . c_j - 0.5 sum_j n_j |c_j|^2      (the fold turns it into the cost)
        cost = 0;
        for (int i = tid; i < k * d; i += kComputeThreads) { const int j = i / d, f = i - j * d; cost = fma(Cs[j * 64 + f], p.C[i], cost); }
        if (tid < k) cost = fma((double)reinterpret_cast<const unsigned int*>(smem + L.zero_off)[tid], cn_s[tid], cost);
      }
#pragma unroll
      for (int msk = 16; msk >= 1; msk >>= 1) { cost += shfl_xor_d(cost, msk); xxs += shfl_xor_d(xxs, msk); }
      if (lane == 0) { red_s[warp] = cost; red_s[kComputeWarps + warp] = xxs; }
      if (p.do_reduce == 1) {
        for (int i = tid; i < k * d; i += kComputeThreads) { const int j = i / d, f = i - j * d; out[i] = Cs[j * 64 + f]; }
        if (tid < k) out[(size_t)k * d + tid] = (double)reinterpret_cast<const unsigned int*>(smem + L.zero_off)[tid];
      }
      named_bar_sync(1, kComputeThreads);
      if (tid == 0) {
        double c = 0, x2 = 0;
#pragma unroll
        for (int wv = 0; wv < kComputeWarps; ++wv) { c += red_s[wv]; x2 += red_s[kComputeWarps + wv]; }
        out[P - 1] = c; out[P] = x2;
      }
      __threadfence();
      named_bar_sync(1, kComputeThreads);
      __shared__ unsigned int s_ticket;
      if (tid == 0) s_ticket = atomicAdd(p.counter, 1u);
      named_bar_sync(1, kComputeThreads);
      if (s_ticket == gridDim.x - 1) {
        __threadfence();
        const int i_lo = p.do_reduce == 2 ? P - 1 : 0, i_hi = p.do_reduce == 2 ? P : P + 1;
        for (int i = i_lo + tid; i < i_hi; i += kComputeThreads) {
          double sum = 0;
          sum = ordered_sum_ldcg(p.cta_partials + i, (size_t)S, gridDim.x);
          if (i == P - 1 && xx_mode == 2) sum = fmax(fma(-2.0, sum, *p.xnorm2_sum), 0.0);
          if (p.do_reduce == 2) p.partial_out[0] = sum;
          else if (i < P) p.partial_out[i] = sum;
          else if (xx_mode == 1) *p.xnorm2_sum = sum;
        }
        if (tid == 0) *p.counter = 0u;
        bool peer_ok = true;
        if (p.peer_bufs != nullptr && p.do_reduce == 1) {
          __threadfence_block();
          named_bar_sync(1, kComputeThreads);
          peer_ok = peer_allreduce(p, p.partial_out, P, tid, kComputeThreads, [] { named_bar_sync(1, kComputeThreads); });
        }
        if (p.C_update != nullptr && peer_ok) {
          __threadfence_block();
          named_bar_sync(1, kComputeThreads);
          double shift = -1.0;
          if (tid < k) {
            const double wsum = p.partial_out[(size_t)k * d + tid];
            if (wsum > 0) {
              const double inv = 1.0 / wsum; double sft = 0;
              for (int f = 0; f < d; ++f) {
                const double nc = p.partial_out[(size_t)tid * d + f] * inv, df = nc - p.C_update[(size_t)tid * d + f];
                sft = fma(df, df, sft); p.C_update[(size_t)tid * d + f] = nc;
              }
              shift = sft;
            }
          }
          if (tid < 64) red_s[2 * kComputeWarps + tid] = shift;
          named_bar_sync(1, kComputeThreads);
          if (tid == 0) {
            double mxs = -1.0;
            for (int i = 0; i < k; ++i) mxs = fmax(mxs, red_s[2 * kComputeWarps + i]);
            p.state_rw->cost = p.partial_out[(size_t)k * d + k];
            p.state_rw->iter += 1;
            p.state_rw->max_shift2 = mxs < 0 ? 0.0 : mxs;
            p.state_rw->converged = (mxs <= p.tol2) ? 1 : 0;
          }
        }
      }
    }
  }
  // ------------------------------------------------------------------ release TMEM (all warps reach this point)
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == kComputeWarps) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(tmem) : "memory");
}

}  // namespace hkm

// Lloyd data pass: fused assign (fp64 DMMA score tiles + shuffle argmin) and per-cluster
// sum/count/cost reduction, X streamed once from HBM through a bulk-async (TMA 1-D) ring.
//
// Replaces, per row, mllib EuclideanDistanceMeasure.findClosest + updateClusterSum + costAccum
// (SURVEY 8a-4/8a-5); Spark runs these as scalar JVM loops inside mapPartitions.
//
// Tile geometry (kTileRows = 128 rows per stage): the tile is fetched as 8 chunks of 16
// consecutive rows; chunk c lands at smem offset c * (16*d*8 + 32).  MMA sub-tile i (0..15)
// takes row i of every chunk, so the 8 rows of one DMMA A-fragment sit 32 bytes apart modulo
// 128 and the LDS.64 fragment loads are bank-conflict free for every d.
//
// Phase 1 (per warp, two sub-tiles): score_j = x.c_j - 0.5|c_j|^2 by DMMA.8x8x4, argmax over j
//   (lowest index wins ties, as findClosest's strict '<'), dist = max(|x|^2 - 2 score, 0).
// Phase 2 (after one CTA barrier): warp w owns clusters w, w+8, ...; it ballots the tile's
//   labels and adds the rows of its clusters, one feature per lane, into REGISTER accumulators
//   -- no atomics, no contention, fixed summation order.
// Tail: per-CTA partials -> global scratch; the last CTA to finish sums them in CTA order, so
// results are bit-stable run to run.
#pragma once
#include "hkm_common.cuh"
#include <type_traits>

namespace hkm {

constexpr int kComputeWarps = 8;
constexpr int kComputeThreads = kComputeWarps * 32;
constexpr int kThreads = kComputeThreads;  // warp 0 lane 0 doubles as the TMA producer
constexpr int kTileRows = 128;
constexpr int kChunks = 8;
constexpr int kChunkRows = kTileRows / kChunks;  // 16
constexpr int kChunkPad = 32;                    // bytes
constexpr int kMaxStages = 8;

struct LloydParams {
  const double* X;
  int64_t n;
  int d;
  int k;
  const double* C;
  const double* w;         // nullable
  double* partial_out;     // [k*d + k + 1]
  int32_t* labels_out;     // nullable
  double* mind2_out;       // nullable
  const int32_t* skip;     // nullable: device flag, nonzero => no-op
  const int32_t* skip2;    // nullable: second flag of the same kind (set by a variant launched ahead that already produced the pass)
  int do_reduce;           // 0: labels/mind2 only; 1: sums + counts + cost; 2: cost only (partial_out[0])
  const int32_t* labels_in;  // MODE 1/2 (silhouette): given cluster ids; ids outside [0, k) are ignored
  const double* aux;         // MODE 2: [k] cluster weight sums N, then [k] Psi/N
  int out_mode;            // 0: labels/mind2 are written; 1: mind2_out = min(mind2_out, new), labels follow the winner
  int label_base;          // added to labels written in out_mode 1, subtracted from labels_in in MODE 1 (centre chunk offset)
  double* xnorm2_sum;      // see xx_mode
  int xx_mode;             // 0: per-row |x|^2 in-pass; 1: same, and store sum_w|x|^2; 2: reuse stored sum, skip per-row norms
  double* C_update;        // nullable: fused driver step -- the last CTA also updates these centres in place
  double tol2;             //   convergence threshold (squared); < 0 never converges
  hkm_lloyd_state* state_rw;  // state written by the fused update
  // fused cross-GPU reduction (hkm_lloyd_step_peer): peer-mapped exchange buffers, one per rank
  void* const* peer_bufs;  // nullable: device array [xworld] of buffer base addresses (this rank's own at [xrank])
  int xrank, xworld;
  unsigned long long xepoch;  // strictly increasing call counter, identical on every rank
  int* xerr;               // set to 1 if a peer never showed up (spin time-out)
  double* cta_partials;    // scratch: [gridDim.x][k*d + 2k + 2]
  unsigned int* counter;   // scratch: zero before launch; left zero
  int stages;
  int ks;                  // ceil(d/4)
};

// Driver side of one Lloyd iteration (KMeans.runAlgorithmWithWeight after collectAsMap): new_j = sums_j * (1/w_j)
// for clusters with w_j > 0, others keep their centre; converged iff every updated centre moved <= tol^2.
// Called by one CTA (any blockDim >= 64 threads participate as j = tid < k); s_shift: >= 64 doubles of smem.
__device__ __forceinline__ void update_centers_body(const double* partial, int k, int d, double* C, double tol2,
                                                    hkm_lloyd_state* state, double* s_shift, int tid) {
  double shift = -1.0;
  if (tid < 64) {
    for (int j = tid; j < k; j += 64) {                    // k <= 64 in the fused kernels; the stand-alone launch takes any k
      const double wsum = partial[(size_t)k * d + j];
      if (wsum > 0) {
        const double inv = 1.0 / wsum;  // scal(1.0 / weightSum, sum)
        double s = 0;
        for (int f = 0; f < d; ++f) {
          const double nc = partial[(size_t)j * d + f] * inv;
          const double df = nc - C[(size_t)j * d + f];
          s = fma(df, df, s);
          C[(size_t)j * d + f] = nc;
        }
        shift = fmax(shift, s);
      }
    }
    s_shift[tid] = shift;
  }
  __syncthreads();
  if (tid == 0) {
    double mx = -1.0;
    for (int i = 0; i < 64; ++i) mx = fmax(mx, s_shift[i]);
    state->cost = partial[(size_t)k * d + k];
    state->iter += 1;
    state->max_shift2 = mx < 0 ? 0.0 : mx;
    state->converged = (mx <= tol2) ? 1 : 0;  // every updated centre within tol (none updated => converged)
  }
}

// ---- one-shot all-reduce of a small fp64 vector over NVLink peer memory, run by the last CTA of every rank's pass.
// Each rank owns an exchange buffer mapped into every peer: slots[2][world][kXchgSlot] doubles, then flags[2][world].
// A rank stores its vector into slot [epoch&1][rank] of EVERY peer (plain stores through the peer mapping), fences,
// raises flag [epoch&1][rank] = epoch on every peer, waits until all `world` flags of its own buffer carry `epoch`,
// and sums the `world` slots in rank order -- so every rank obtains the bit-identical total without a collective
// launch.  Two slot sets (epoch parity) keep a fast rank's next step from overwriting what a slow rank still reads.
constexpr int kXchgSlot = 4352;   // doubles per slot: >= 64*65 + 64 + 2
__host__ __device__ inline size_t xchg_flags_offset(int world) { return (size_t)2 * world * kXchgSlot * sizeof(double); }
__host__ __device__ inline size_t xchg_buffer_bytes(int world) { return xchg_flags_offset(world) + (size_t)2 * world * sizeof(unsigned long long) + 64; }

// Returns false (on every thread) if a peer never showed up: the caller must then leave the centres alone.
template <class Sync>
__device__ __forceinline__ bool peer_allreduce(const LloydParams& p, double* vec, int len, int tid, int nthreads, Sync sync) {
  __shared__ int s_peer_ok;
  if (tid == 0) s_peer_ok = 1;
  // timing of the exchange (nanoseconds, %globaltimer), accumulated per rank in hkm_lloyd_state.reserved: low word = time to
  // store this rank's vector into every peer, high word = time spent waiting for the peers' flags (arrival skew + flight)
  unsigned long long t_begin = 0, t_stored = 0;
  if (tid == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_begin));
  const int world = p.xworld, rank = p.xrank, set = (int)(p.xepoch & 1ull);
  const size_t slot_off = ((size_t)set * world + rank) * kXchgSlot;
  for (int q = 0; q < world; ++q) {
    double* dst = reinterpret_cast<double*>(p.peer_bufs[q]) + slot_off;
    for (int i = tid; i < len; i += nthreads) dst[i] = vec[i];
  }
  __threadfence_system();
  sync();
  if (tid == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_stored));
  if (tid < world) {
    unsigned long long* f = reinterpret_cast<unsigned long long*>(reinterpret_cast<unsigned char*>(p.peer_bufs[tid]) + xchg_flags_offset(world)) + set * world + rank;
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(f), "l"(p.xepoch) : "memory");
    const unsigned long long* mine = reinterpret_cast<const unsigned long long*>(reinterpret_cast<const unsigned char*>(p.peer_bufs[rank]) + xchg_flags_offset(world)) + set * world + tid;
    unsigned long long v = 0;
    const long long t0 = clock64();
    do {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(mine) : "memory");
      if (v != p.xepoch && clock64() - t0 > 8000000000LL) { if (p.xerr != nullptr) *p.xerr = 1; s_peer_ok = 0; break; }   // ~4 s: a peer is missing, do not hang the GPU
    } while (v != p.xepoch);
  }
  sync();
  if (tid == 0 && p.state_rw != nullptr) {
    unsigned long long t_end;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end));
    const unsigned long long prev = (unsigned long long)p.state_rw->reserved;
    const unsigned long long lo = (prev & 0xffffffffull) + (t_stored - t_begin), hi = (prev >> 32) + (t_end - t_stored);
    p.state_rw->reserved = (long long)((hi << 32) | (lo & 0xffffffffull));
  }
  if (!s_peer_ok) {
    // incomplete exchange: stop this rank's loop (later passes become no-ops) instead of updating from partial sums;
    // the host sees err_flag and raises on every rank
    if (tid == 0 && p.state_rw != nullptr) p.state_rw->converged = 1;
    return false;
  }
  const double* base = reinterpret_cast<const double*>(p.peer_bufs[rank]) + (size_t)set * world * kXchgSlot;
  for (int i = tid; i < len; i += nthreads) {
    double s = 0;
    for (int q = 0; q < world; ++q) s += __ldcv(base + (size_t)q * kXchgSlot + i);
    vec[i] = s;
  }
  return true;
}

__host__ __device__ inline int chunk_stride_bytes(int d) { return kChunkRows * d * 8 + kChunkPad; }
__host__ __device__ inline int stage_bytes(int d) { return kChunks * chunk_stride_bytes(d); }
__host__ __device__ inline int cs_ld(int d) { return ((d + 3) & ~3) + 4; }  // padded centre row (doubles), == 4 mod 8 when d%8==0

// dynamic smem carve-up (bytes), all 16-byte aligned
struct SmemLayout {
  int x_off, labels_off, mind2_off, cs_off, cn_off, red_off, bar_off, list_off, zero_off, w_off, total;
};
__host__ __device__ inline int list_entries(int k_pad) { return kTileRows + 4 * (k_pad / 8) + 4; }  // per warp
__host__ __device__ inline SmemLayout smem_layout(int d, int k_pad, int stages) {
  SmemLayout L;
  int o = 0;
  L.x_off = o; o += stages * stage_bytes(d); o = (o + 127) & ~127;
  L.labels_off = o; o += 2 * kTileRows * 4;
  L.mind2_off = o; o += 2 * kTileRows * 8;
  L.cs_off = o; o += k_pad * cs_ld(d) * 8;
  L.cn_off = o; o += 2 * k_pad * 8;
  L.red_off = o; o += 2 * kComputeWarps * 8 + 16;
  L.bar_off = o; o += kMaxStages * 8;
  L.list_off = o; o += kComputeWarps * list_entries(k_pad) * 4; o = (o + 15) & ~15;
  L.zero_off = o; o += ((d + 3) & ~3) * 8;
  L.w_off = o; o += 2 * kTileRows * 8;
  L.total = o;
  return L;
}

// KS_T > 0: centre B-fragments live in registers (d_pad = 4*KS_T, fully unrolled);
// KS_T == 0: B-fragments are re-read from shared memory, runtime k-step count.
// NT = ceil(k/8) score column tiles == clusters owned per warp in phase 2.
// LPR = lanes per row in phase 2 (8/16/32), FPL = features per lane (d > 32).
// MODE 0: Lloyd / predict pass, unit weights.  MODE 3: the same with per-row weights.
// MODE 1: silhouette cluster statistics (labels given: per-cluster sum w*x, sum w, sum w*|x|^2).
// MODE 2: silhouette score pass (labels + statistics given).
//
// All hot-loop shared-memory traffic goes through 32-bit shared addresses (lds_*/sts_* helpers).
template <int KS_T, int NT, int LPR, int FPL, int MAXREG, int MODE = 0>
__global__ void __launch_bounds__(kThreads) __maxnreg__(MAXREG) lloyd_pass_kernel(const LloydParams p) {
  if (p.skip != nullptr && *p.skip != 0) return;
  if (p.skip2 != nullptr && *p.skip2 != 0) return;

  extern __shared__ __align__(128) unsigned char smem[];
  const int d = p.d, k = p.k;
  constexpr int k_pad = NT * 8;
  const int KS = KS_T > 0 ? KS_T : p.ks;
  const SmemLayout L = smem_layout(d, k_pad, p.stages);
  double* Cs = reinterpret_cast<double*>(smem + L.cs_off);
  double* cn_s = reinterpret_cast<double*>(smem + L.cn_off);
  double* red_s = reinterpret_cast<double*>(smem + L.red_off);

  const uint32_t sb = smem_u32(smem);                 // shared-window address of the dynamic smem base
  const uint32_t a_x = sb + L.x_off, a_labels = sb + L.labels_off, a_mind2 = sb + L.mind2_off, a_w = sb + L.w_off;
  const uint32_t a_full = sb + L.bar_off, a_zero = sb + L.zero_off;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int stages = p.stages;
  const int csld = cs_ld(d);
  const uint32_t cstride = (uint32_t)chunk_stride_bytes(d);
  const uint32_t sbytes = (uint32_t)stage_bytes(d);
  const uint32_t row_bytes = (uint32_t)d * 8u;

  const int64_t full_tiles = p.n / kTileRows;
  const int tail_rows = (int)(p.n - full_tiles * kTileRows);

  constexpr bool WEIGHTED = MODE == 3;
  // weights are a compile-time fact on the Lloyd paths; the silhouette passes take them at run time
  const bool has_w = WEIGHTED || ((MODE == 1 || MODE == 2) && p.w != nullptr);
  constexpr bool EXTRAS = MODE == 1 || MODE == 3;   // phase 2 needs per-row scalars (weight, |x|^2) next to the row

  if (tid == 0) {
    for (int s = 0; s < stages; ++s) mbar_init(reinterpret_cast<uint64_t*>(smem + L.bar_off) + s, 1);
    mbar_fence_init();
  }
  // stage centres: Cs[j][f] (zero padded), cn_s[j] = -0.5|c_j|^2 (-inf for padding)
  for (int i = tid; i < k_pad * csld; i += kThreads) {
    int j = i / csld, f = i - j * csld;
    Cs[i] = (MODE != 1 && j < k && f < d) ? p.C[(size_t)j * d + f] : 0.0;   // the statistics pass has no centres
  }
  for (int i = tid; i < ((d + 3) & ~3); i += kThreads) reinterpret_cast<double*>(smem + L.zero_off)[i] = 0.0;
  __syncthreads();
  for (int j = tid; j < k_pad; j += kThreads) {
    double s = 0;
    for (int f = 0; f < d; ++f) s = fma(Cs[j * csld + f], Cs[j * csld + f], s);
    const double ninf = -__longlong_as_double(0x7ff0000000000000LL);
    if (MODE == 2) {
      // D(x, G) = |x|^2 + Psi_G/N_G - 2 x.(Y_G/N_G): same shape as the Lloyd score with -0.5*Psi_G/N_G as the constant
      const double Nj = j < k ? p.aux[j] : 0.0;
      cn_s[j] = (j < k && Nj > 0) ? -0.5 * p.aux[k + j] : ninf;
      cn_s[k_pad + j] = Nj;
    } else {
      cn_s[j] = j < k ? -0.5 * s : ninf;
    }
  }
  __syncthreads();

  // ---- TMA producer (thread 0): local tile i of this CTA is global tile blockIdx.x + i*gridDim.x and lands in
  // stage i % stages.  A stage is refilled right after the CTA barrier of the NEXT tile, which every warp reaches
  // only after it has finished reading the stage -- so no "empty" mbarriers are needed.
  const int my_tiles = full_tiles > (int64_t)blockIdx.x ? (int)((full_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;
  const uint32_t cbytes = (uint32_t)kChunkRows * row_bytes;
  const size_t tile_gbytes = (size_t)kTileRows * row_bytes;
  const unsigned char* src_next = reinterpret_cast<const unsigned char*>(p.X) + (size_t)blockIdx.x * tile_gbytes;  // next tile to fetch
  const size_t src_step = (size_t)gridDim.x * tile_gbytes;
  int next_fill = 0;                                   // local index of the next tile to fetch
  auto issue_tile = [&](int s) {                       // fetch tile `next_fill` into stage s
    const uint32_t bar = a_full + 8u * (uint32_t)s, dst = a_x + (uint32_t)s * sbytes;
    mbar_expect_tx_a(bar, cbytes * kChunks);
#pragma unroll
    for (int c = 0; c < kChunks; ++c) bulk_g2s_a(dst + c * cstride, src_next + (size_t)c * cbytes, cbytes, bar);
    src_next += src_step; ++next_fill;
  };
  if (tid == 0) {
    for (int s = 0; s < stages && next_fill < my_tiles; ++s) issue_tile(s);
  }

  const int r = lane >> 2, q = lane & 3;

  // centre B-fragments: lane (n = r, q) holds C[8*nt + n][4*s + q]
  double breg[KS_T > 0 ? KS_T * NT : 1];
  if (KS_T > 0) {
#pragma unroll
    for (int s = 0; s < (KS_T > 0 ? KS_T : 1); ++s)
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) breg[s * NT + nt] = Cs[(8 * nt + r) * csld + 4 * s + q];
  }
  double cinit[NT][2];
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) { cinit[nt][0] = cn_s[8 * nt + 2 * q]; cinit[nt][1] = cn_s[8 * nt + 2 * q + 1]; }
  const uint32_t a_cb = sb + L.cs_off + (uint32_t)((r * csld + q) * 8);   // KS_T == 0: B fragments re-read from smem
  const uint32_t a_nn = sb + L.cn_off + (uint32_t)k_pad * 8u;             // MODE 2: cluster weight sums

  // phase-2 register accumulators: cluster warp + 8*m, features (lane % LPR) + 32*e
  double acc_a[NT][FPL], acc_b[NT][FPL];
  double wacc[NT];        // weighted runs: per-lane partial weight sums
  double psi[NT];         // MODE 1: sum of w*|x|^2 per owned cluster (lane fl == 0 only)
  unsigned cnt[NT];       // unweighted runs: rows per owned cluster (warp-uniform)
#pragma unroll
  for (int m = 0; m < NT; ++m) {
    cnt[m] = 0; wacc[m] = 0; psi[m] = 0;
#pragma unroll
    for (int e = 0; e < FPL; ++e) { acc_a[m][e] = 0; acc_b[m][e] = 0; }
  }
  double cost = 0;
  double xxs = 0;  // xx_mode 1: sum of w*|x|^2 over this thread's rows;  MODE 2: sum of w
  const int xx_mode = p.xx_mode;
  const int sub = lane / LPR, fl = lane % LPR;
  constexpr int RPS = 32 / LPR;  // rows handled per warp load step
  const uint32_t a_list = sb + L.list_off + (uint32_t)(warp * list_entries(k_pad)) * 4u;
  const unsigned lt_mask = (1u << lane) - 1u;
  // byte offset (from the stage base) of tile row `lane`; row lane+32g adds g*group_stride
  const uint32_t off_lane = (uint32_t)lane * row_bytes + (uint32_t)(lane >> 4) * kChunkPad;
  const uint32_t group_stride = 32u * row_bytes + 2u * kChunkPad;
  const uint32_t fl8 = (uint32_t)fl * 8u;
  // A-fragment base of this lane inside a stage: chunk r, feature q; sub-tile i adds i*row_bytes
  const uint32_t a_frag = (uint32_t)r * cstride + (uint32_t)q * 8u;
  const int i0 = warp, i1 = warp + 8;
  const int t0 = r * kChunkRows + i0, t1 = r * kChunkRows + i1;   // tile rows of this lane's two MMA rows

  auto process_tile = [&](auto need_xx_tag, const uint32_t xs, const int rows_valid, const int64_t row0, const int buf, const int refill_stage) {
    constexpr bool NEED_XX = decltype(need_xx_tag)::value;
    const uint32_t lab_b = a_labels + (uint32_t)buf * (kTileRows * 4), m2_b = a_mind2 + (uint32_t)buf * (kTileRows * 8),
                   w_b = a_w + (uint32_t)buf * (kTileRows * 8);
    // ---------------- phase 1: two MMA sub-tiles per warp (rows i0 / i1 of every chunk)
    const uint32_t xa0 = xs + a_frag + (uint32_t)i0 * row_bytes, xa1 = xs + a_frag + (uint32_t)i1 * row_bytes;
    double s0[NT][2], s1[NT][2];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) { s0[nt][0] = cinit[nt][0]; s0[nt][1] = cinit[nt][1]; s1[nt][0] = cinit[nt][0]; s1[nt][1] = cinit[nt][1]; }
    double xx0 = 0, xx1 = 0;
    if (KS_T > 0) {
#pragma unroll
      for (int s = 0; s < (KS_T > 0 ? KS_T : 1); ++s) {
        const bool ok = (s < KS_T - 1) || (4 * s + q < d);
        const double a0 = ok ? lds_f64(xa0 + 32u * s) : 0.0, a1 = ok ? lds_f64(xa1 + 32u * s) : 0.0;
        if (NEED_XX) { xx0 = fma(a0, a0, xx0); xx1 = fma(a1, a1, xx1); }
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) { if (MODE != 1) dmma1684(s0[nt][0], s0[nt][1], s1[nt][0], s1[nt][1], a0, a1, breg[s * NT + nt]); }
      }
    } else {
      for (int s = 0; s < KS; ++s) {
        const bool ok = 4 * s + q < d;
        const double a0 = ok ? lds_f64(xa0 + 32u * s) : 0.0, a1 = ok ? lds_f64(xa1 + 32u * s) : 0.0;
        if (NEED_XX) { xx0 = fma(a0, a0, xx0); xx1 = fma(a1, a1, xx1); }
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
          if (MODE != 1) {
            const double b = lds_f64(a_cb + (uint32_t)((8 * nt) * csld * 8) + 32u * s);
            dmma1684(s0[nt][0], s0[nt][1], s1[nt][0], s1[nt][1], a0, a1, b);
          }
        }
      }
    }
    if constexpr (MODE == 1) {
      // ---- silhouette statistics: labels are given; phase 1 only supplies |x|^2 per row
#pragma unroll
      for (int msk = 1; msk <= 2; msk <<= 1) { xx0 += shfl_xor_d(xx0, msk); xx1 += shfl_xor_d(xx1, msk); }
      if (q == 0) {
        const bool v0 = t0 < rows_valid, v1 = t1 < rows_valid;
        int l0 = v0 ? p.labels_in[row0 + t0] - p.label_base : -1, l1 = v1 ? p.labels_in[row0 + t1] - p.label_base : -1;   // label_base: centre chunk offset (k > 64)
        if (l0 >= k) l0 = -1;
        if (l1 >= k) l1 = -1;
        sts_u32(lab_b + 4u * t0, (uint32_t)l0); sts_u32(lab_b + 4u * t1, (uint32_t)l1);
        sts_f64(w_b + 8u * t0, (v0 && has_w) ? p.w[row0 + t0] : 1.0);
        sts_f64(w_b + 8u * t1, (v1 && has_w) ? p.w[row0 + t1] : 1.0);
        // ClusteringMetrics: squaredNorm = math.pow(Vectors.norm(features, 2.0), 2.0)
        const double n0 = sqrt(xx0), n1 = sqrt(xx1);
        sts_f64(m2_b + 8u * t0, n0 * n0); sts_f64(m2_b + 8u * t1, n1 * n1);
      }
    } else if constexpr (MODE == 2) {
      // ---- silhouette score (ClusteringMetrics.pointSilhouetteCoefficient), one value per row
      const bool v0 = t0 < rows_valid, v1 = t1 < rows_valid;
      int a0l = v0 ? p.labels_in[row0 + t0] : -1, a1l = v1 ? p.labels_in[row0 + t1] : -1;   // same address on the 4 lanes of a row
      if (a0l >= k) a0l = -1;
      if (a1l >= k) a1l = -1;
      const double ninf = -__longlong_as_double(0x7ff0000000000000LL);
      double own0 = ninf, own1 = ninf, oth0 = ninf, oth1 = ninf;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int j = 8 * nt + 2 * q + e;
          if (j == a0l) own0 = s0[nt][e]; else oth0 = fmax(oth0, s0[nt][e]);
          if (j == a1l) own1 = s1[nt][e]; else oth1 = fmax(oth1, s1[nt][e]);
        }
#pragma unroll
      for (int msk = 1; msk <= 2; msk <<= 1) {
        own0 = fmax(own0, shfl_xor_d(own0, msk)); own1 = fmax(own1, shfl_xor_d(own1, msk));
        oth0 = fmax(oth0, shfl_xor_d(oth0, msk)); oth1 = fmax(oth1, shfl_xor_d(oth1, msk));
        xx0 += shfl_xor_d(xx0, msk); xx1 += shfl_xor_d(xx1, msk);
      }
      if (q < 2) {   // lane q = 0 finishes row t0, lane q = 1 row t1: one pass through the fp64 divisions instead of two
        auto point = [&](bool valid, int a, double xx, double own, double oth, int64_t row) {
          if (!valid || a < 0) return;
          const double wv = has_w ? p.w[row] : 1.0;
          const double nrm = sqrt(xx), sn = nrm * nrm;
          const double NA = lds_f64(a_nn + 8u * a);
          double sil = 0.0;
          if (NA != wv) {                                  // single-element clusters have silhouette 0
            const double cur = fma(-2.0, own, sn) * NA / (NA - wv);
            const double nb = fma(-2.0, oth, sn);
            if (cur < nb) sil = 1.0 - (cur / nb); else if (cur > nb) sil = (nb / cur) - 1.0;
          }
          cost = fma(sil, wv, cost); xxs += wv;
        };
        point(q ? v1 : v0, q ? a1l : a0l, q ? xx1 : xx0, q ? own1 : own0, q ? oth1 : oth0, row0 + (q ? t1 : t0));
      }
    } else {
      // argmax over this lane's columns (ascending index), then across the 4 lanes of the row
      double b0 = s0[0][0], b1 = s1[0][0]; int j0 = 2 * q, j1 = 2 * q;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          if (nt == 0 && e == 0) continue;
          const int j = 8 * nt + 2 * q + e;
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
      if (q == 0) {
        const bool v0 = t0 < rows_valid, v1 = t1 < rows_valid;
        // per-row term of the cost: the clamped distance, or (xx_mode 2) the score itself, folded
        // against the stored sum of |x|^2 by the last CTA
        double d0 = b0, d1 = b1;
        if (NEED_XX) { d0 = fmax(fma(-2.0, b0, xx0), 0.0); d1 = fmax(fma(-2.0, b1, xx1), 0.0); }
        const double c0 = (NEED_XX && xx_mode == 2) ? b0 : d0, c1 = (NEED_XX && xx_mode == 2) ? b1 : d1;
        if (p.out_mode != 0) {
          // min-merge mode: the output threads below own the cost
        } else if (WEIGHTED) {
          const double w0 = v0 ? p.w[row0 + t0] : 0.0, w1 = v1 ? p.w[row0 + t1] : 0.0;
          cost = fma(c0, w0, cost); cost = fma(c1, w1, cost);
          if (NEED_XX && xx_mode == 1) { xxs = fma(xx0, w0, xxs); xxs = fma(xx1, w1, xxs); }
          sts_f64(w_b + 8u * t0, w0); sts_f64(w_b + 8u * t1, w1);
        } else {
          if (v0) cost += c0;
          if (v1) cost += c1;
          if (NEED_XX && xx_mode == 1) { if (v0) xxs += xx0; if (v1) xxs += xx1; }
        }
        sts_u32(lab_b + 4u * t0, (uint32_t)(v0 ? j0 : -1));
        sts_u32(lab_b + 4u * t1, (uint32_t)(v1 ? j1 : -1));
        if (NEED_XX && p.mind2_out != nullptr) { sts_f64(m2_b + 8u * t0, d0); sts_f64(m2_b + 8u * t1, d1); }
      }
    }
    __syncthreads();
    if (tid == 0 && refill_stage >= 0 && next_fill < my_tiles) issue_tile(refill_stage);  // the previous tile's stage is free now

    // ---------------- outputs (coalesced from smem)
    if (MODE != 0 && MODE != 3) {
    } else if (p.out_mode == 0) {
      if (p.labels_out != nullptr && tid < kTileRows && tid < rows_valid) p.labels_out[row0 + tid] = (int)lds_u32(lab_b + 4u * tid);
      if (p.mind2_out != nullptr && tid >= kTileRows && (tid - kTileRows) < rows_valid)
        p.mind2_out[row0 + tid - kTileRows] = lds_f64(m2_b + 8u * (tid - kTileRows));
    } else if (tid < kTileRows && tid < rows_valid) {
      // running minimum over successive centre chunks (k-means|| cost update, candidate weights):
      // strict '<' keeps the earlier (lower-index) centre on ties, as findClosest does
      const double nv = lds_f64(m2_b + 8u * tid), old = p.mind2_out[row0 + tid];
      if (nv < old) {
        p.mind2_out[row0 + tid] = nv;
        if (p.labels_out != nullptr) p.labels_out[row0 + tid] = (int)lds_u32(lab_b + 4u * tid) + p.label_base;
        cost += nv;
      } else cost += old;
    }

    // ---------------- phase 2: owner warps gather their clusters' rows
    // Each owner warp compacts the shared addresses of its clusters' rows into a private list (ballot + prefix
    // popcount, one predicated store per 32 rows), pads every segment to a multiple of 4 with the address of an
    // all-zero row, then streams the list: one broadcast LDS.128 of four addresses, four row loads, four adds --
    // no per-row bit twiddling, no predication, fixed order.
    if (p.do_reduce == 1 && MODE != 2) {
      int lab[kTileRows / 32];
#pragma unroll
      for (int g = 0; g < kTileRows / 32; ++g) lab[g] = (int)lds_u32(lab_b + 4u * (g * 32 + lane));
      const uint32_t e_lane = xs + off_lane;   // shared address of tile row `lane`
      unsigned seg_begin[NT], seg_cnt[NT];
      unsigned pos = 0;
#pragma unroll
      for (int m = 0; m < NT; ++m) {
        const int j = warp + 8 * m;
        seg_begin[m] = pos;
#pragma unroll
        for (int g = 0; g < kTileRows / 32; ++g) {
          const bool mine = lab[g] == j;
          const unsigned mask = __ballot_sync(0xffffffffu, mine);
          uint32_t e = e_lane + g * group_stride;
          if (EXTRAS) e |= (uint32_t)(g * 32 + lane) << 20;   // shared addresses stay below 2^20
          sts_u32_if(a_list + 4u * (pos + __popc(mask & lt_mask)), e, mine);
          pos += __popc(mask);
        }
        seg_cnt[m] = pos - seg_begin[m];
        sts_u32_if(a_list + 4u * (pos + lane), a_zero, lane < (int)((4u - (pos & 3u)) & 3u));  // padding -> zero row
        pos = (pos + 3) & ~3u;
      }
      __syncwarp();
#pragma unroll
      for (int m = 0; m < NT; ++m) {
        cnt[m] += seg_cnt[m];
        const uint32_t lp = a_list + 4u * seg_begin[m];
        const unsigned nquad = (seg_cnt[m] + 3) >> 2;
        if (!EXTRAS) {
#pragma unroll 2
          for (unsigned i = 0; i < nquad; ++i) {
            const uint4 o = lds_v4(lp + 16u * i);
            const uint32_t oo[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
            for (int u = 0; u < 4 / RPS; ++u) {
              uint32_t osel = oo[u * RPS];
#pragma unroll
              for (int z = 1; z < RPS; ++z) osel = (sub == z) ? oo[u * RPS + z] : osel;
#pragma unroll
              for (int e = 0; e < FPL; ++e) {
                const double v = lds_f64(osel + fl8 + 256u * e);  // lanes past d read neighbouring smem into accumulators that are never stored
                if (u & 1) acc_b[m][e] += v; else acc_a[m][e] += v;
              }
            }
          }
        } else {
          for (unsigned i = 0; i < nquad; ++i) {
            const uint4 o = lds_v4(lp + 16u * i);
            const uint32_t oo[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
            for (int u = 0; u < 4 / RPS; ++u) {
              uint32_t osel = oo[u * RPS];
#pragma unroll
              for (int z = 1; z < RPS; ++z) osel = (sub == z) ? oo[u * RPS + z] : osel;
              const uint32_t addr = osel & 0xFFFFFu, trow = osel >> 20;
              const bool pad = addr == a_zero;
              const double wv = pad ? 0.0 : lds_f64(w_b + 8u * trow);
              if (fl == 0) {
                wacc[m] += wv;
                if (MODE == 1) psi[m] = fma(wv, pad ? 0.0 : lds_f64(m2_b + 8u * trow), psi[m]);
              }
#pragma unroll
              for (int e = 0; e < FPL; ++e) {
                const double v = (fl + 32 * e) < d ? lds_f64(addr + fl8 + 256u * e) : 0.0;  // keep 0*garbage out of the fma
                if (u & 1) acc_b[m][e] = fma(v, wv, acc_b[m][e]); else acc_a[m][e] = fma(v, wv, acc_a[m][e]);
              }
            }
          }
        }
      }
      __syncwarp();
    }
  };

  // main loop over this CTA's full tiles
  {
    int s = 0, buf = 0; uint32_t ph = 0;
    int64_t row0 = (int64_t)blockIdx.x * kTileRows;
    const int64_t row_step = (int64_t)gridDim.x * kTileRows;
    for (int i = 0; i < my_tiles; ++i) {
      const int refill_stage = i == 0 ? -1 : (s == 0 ? stages - 1 : s - 1);   // stage of tile i-1
      mbar_wait_a(a_full + 8u * (uint32_t)s, ph);
      const uint32_t xs = a_x + (uint32_t)s * sbytes;
      if (xx_mode != 2) process_tile(std::true_type{}, xs, kTileRows, row0, buf, refill_stage);
      else process_tile(std::false_type{}, xs, kTileRows, row0, buf, refill_stage);
      if (++s == stages) { s = 0; ph ^= 1; }
      buf ^= 1;
      row0 += row_step;
    }
    // ragged tail (< 128 rows): the CTA next in round-robin order loads it with plain loads
    if (tail_rows > 0 && (int)(full_tiles % gridDim.x) == (int)blockIdx.x) {
      __syncthreads();  // every warp is done with all stages
      unsigned char* xsp = smem + L.x_off;
      const double* src = p.X + (size_t)full_tiles * kTileRows * d;
      for (int i = tid; i < kTileRows * d; i += kComputeThreads) {
        int t = i / d, f = i - t * d;
        double v = t < tail_rows ? src[(size_t)t * d + f] : 0.0;
        reinterpret_cast<double*>(xsp + (t >> 4) * cstride)[(t & 15) * d + f] = v;
      }
      __syncthreads();
      process_tile(std::true_type{}, a_x, tail_rows, full_tiles * kTileRows, buf, -1);
    }
  }
  if (!p.do_reduce) return;

  // ===================================================================== CTA-level combine
  const int P = k * d + k + 1;
  // cost: warp shuffle, then across warps through smem
#pragma unroll
  for (int msk = 16; msk >= 1; msk >>= 1) { cost += shfl_xor_d(cost, msk); xxs += shfl_xor_d(xxs, msk); }
  if (lane == 0) { red_s[warp] = cost; red_s[kComputeWarps + warp] = xxs; }
  // fold lane sub-groups and the two interleaved accumulators
#pragma unroll
  for (int m = 0; m < NT; ++m) {
#pragma unroll
    for (int e = 0; e < FPL; ++e) {
      double v = acc_a[m][e] + acc_b[m][e];
#pragma unroll
      for (int msk = LPR; msk < 32; msk <<= 1) v += shfl_xor_d(v, msk);
      acc_a[m][e] = v;
    }
    if (EXTRAS) {
      double c = wacc[m], ps = psi[m];
#pragma unroll
      for (int msk = 16; msk >= 1; msk >>= 1) { c += shfl_xor_d(c, msk); ps += shfl_xor_d(ps, msk); }
      wacc[m] = c; psi[m] = ps;
    } else wacc[m] = (double)cnt[m];
  }
  // per-CTA slot: [k*d sums][k weights][cost][sum w|x|^2][k psi (MODE 1)]
  const int S = P + 1 + k;
  double* out = p.cta_partials + (size_t)blockIdx.x * S;
#pragma unroll
  for (int m = 0; m < NT; ++m) {
    const int j = warp + 8 * m;
    if (j < k && p.do_reduce == 1) {
      if (sub == 0) {
#pragma unroll
        for (int e = 0; e < FPL; ++e) { const int f = fl + 32 * e; if (f < d) out[(size_t)j * d + f] = acc_a[m][e]; }
      }
      if (lane == 0) { out[(size_t)k * d + j] = wacc[m]; if (MODE == 1) out[P + 1 + j] = psi[m]; }
    }
  }
  __syncthreads();
  if (tid == 0) {
    double c = 0, x2 = 0;
#pragma unroll
    for (int wv = 0; wv < kComputeWarps; ++wv) { c += red_s[wv]; x2 += red_s[kComputeWarps + wv]; }
    out[P - 1] = c; out[P] = x2;
  }
  // ===================================================================== last CTA folds all partials
  __threadfence();
  __syncthreads();
  __shared__ unsigned int s_ticket;
  if (tid == 0) s_ticket = atomicAdd(p.counter, 1u);
  __syncthreads();
  if (s_ticket == gridDim.x - 1) {
    __threadfence();
    const int i_lo = (MODE == 2 || p.do_reduce == 2) ? P - 1 : 0;
    const int i_hi = MODE == 2 ? P + 1 : (p.do_reduce == 2 ? P : (MODE == 1 ? S : P + 1));
    for (int i = i_lo + tid; i < i_hi; i += kComputeThreads) {
      double s = 0;
      s = ordered_sum_ldcg(p.cta_partials + i, (size_t)S, gridDim.x);
      if (i == P - 1 && xx_mode == 2) s = fmax(fma(-2.0, s, *p.xnorm2_sum), 0.0);  // cost = sum w|x|^2 - 2 sum w*score
      if (MODE == 2) p.partial_out[i - (P - 1)] = s;           // [sum s*w, sum w]
      else if (p.do_reduce == 2) p.partial_out[0] = s;
      else if (i < P) p.partial_out[i] = s;
      else if (i == P) { if (xx_mode == 1) *p.xnorm2_sum = s; }
      else p.partial_out[i - 1] = s;                           // MODE 1: psi follows [sums][N][cost]
    }
    if (tid == 0) *p.counter = 0u;
    bool peer_ok = true;
    if (p.peer_bufs != nullptr) {                              // ranks > 1: sum the vector over NVLink peer memory right here
      __threadfence_block();
      __syncthreads();
      peer_ok = peer_allreduce(p, p.partial_out, P, tid, kComputeThreads, [] { __syncthreads(); });
    }
    if (p.C_update != nullptr && peer_ok) {                               // fused driver step: uniform branch
      __threadfence_block();
      __syncthreads();
      update_centers_body(p.partial_out, k, d, p.C_update, p.tol2, p.state_rw, red_s, tid);
    }
  }
}

}  // namespace hkm

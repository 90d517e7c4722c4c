// Instantiations of lloyd_pass_kernel for one NT = ceil(k/8) (compile with -DHKM_NT=<1..8>).
#include <cstdlib>
#include "../../include/hkm.h"
#include "hkm_lloyd_launch.cuh"

#ifndef HKM_NT
#error "compile with -DHKM_NT=<1..8>"
#endif

namespace hkm {

template <int KS_T, int NT, int LPR, int FPL, int MAXREG, int MODE = 0>
static int launch_lloyd(const LloydParams& p, const LaunchPlan& pl, int grid, cudaStream_t st) {
  auto kern = lloyd_pass_kernel<KS_T, NT, LPR, FPL, MAXREG, MODE>;
  static int attr_smem[64];  // per device: largest dynamic smem opted into so far
  int dev = 0; HKM_CUDA_TRY(cudaGetDevice(&dev));
  if (dev >= 64 || attr_smem[dev] < pl.smem) {
    HKM_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, pl.smem));
    if (dev < 64) attr_smem[dev] = pl.smem;
  }
  kern<<<grid, kThreads, pl.smem, st>>>(p);
  HKM_CUDA_TRY(cudaGetLastError());
  return HKM_OK;
}

// register-resident centre fragments: LPR/FPL follow from KS
template <int KS_T, int NT>
static int launch_regb(const LloydParams& p, const LaunchPlan& pl, int grid, cudaStream_t st) {
  constexpr int LPR = KS_T <= 2 ? 8 : (KS_T <= 4 ? 16 : 32);
  constexpr int FPL = KS_T <= 8 ? 1 : 2;
  // 16 warps/SM = 4 per scheduler partition (16K registers each) -> 128 regs when two CTAs share an SM
  constexpr int MAXREG = ((KS_T * NT <= 16 && NT <= 2) || (NT >= 3 && NT <= 6)) ? 128 : 255;
  return launch_lloyd<KS_T, NT, LPR, FPL, MAXREG>(p, pl, grid, st);
}
template <int NT, int MODE = 0>
static int launch_smemb(const LloydParams& p, const LaunchPlan& pl, int grid, cudaStream_t st) {
  const int d = p.d;
  constexpr int MINB = NT <= 6 ? 128 : 255;   // two CTAs per SM (d <= 32 rings fit twice): 16 warps hide the phase latencies
  if (d <= 8) return launch_lloyd<0, NT, 8, 1, MINB, MODE>(p, pl, grid, st);
  if (d <= 16) return launch_lloyd<0, NT, 16, 1, MINB, MODE>(p, pl, grid, st);
  if (d <= 32) return launch_lloyd<0, NT, 32, 1, MINB, MODE>(p, pl, grid, st);
  if (d <= 64) return launch_lloyd<0, NT, 32, 2, (NT <= 2 ? 128 : 255), MODE>(p, pl, grid, st);
  if (d <= 96) return launch_lloyd<0, NT, 32, 3, 255, MODE>(p, pl, grid, st);
  return launch_lloyd<0, NT, 32, 4, 255, MODE>(p, pl, grid, st);
}

// small-k tensor-pipe variant (assign + reduce on DMMA, no CTA barrier)
template <int KS_T, int NT, int MODE = 0>
static int launch_mma(LloydParams p, int grid, cudaStream_t st) {
  const LaunchPlan pl = plan_lloyd_mma(p.d, NT * 8);
  p.stages = pl.stages;
  constexpr int MAXREG = (KS_T <= 2 && NT == 1 && MODE == 0) ? 64 : 96;   // thin rows: three CTAs per SM
  auto kern = lloyd_mma_kernel<KS_T, NT, MAXREG, MODE>;
  static int attr_smem[64];
  int dev = 0; HKM_CUDA_TRY(cudaGetDevice(&dev));
  if (dev >= 64 || attr_smem[dev] < pl.smem) {
    HKM_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, pl.smem));
    if (dev < 64) attr_smem[dev] = pl.smem;
  }
  kern<<<grid, kMmaThreads, pl.smem, st>>>(p);
  HKM_CUDA_TRY(cudaGetLastError());
  return HKM_OK;
}

// mid-k variant: score warps + reduce warps, one CTA of 16 warps per SM (hkm_lloyd_ws.cuh)
template <int KS_T, int NT, int SW, int NSET>
static int launch_ws_n(LloydParams p, int grid, cudaStream_t st) {
  p.stages = ws_stages(p.d, NT * 8, kWsWarps - SW, NSET);
  if (p.stages < 2) { set_error("lloyd_ws: %d stages", p.stages); return HKM_ERR_UNSUPPORTED; }
  const int smem = ws_smem_layout(p.d, NT * 8, p.stages, kWsWarps - SW, NSET).total;
  auto kern = lloyd_ws_kernel<KS_T, NT, SW, NSET, 128>;
  static int attr_smem[64];
  int dev = 0; HKM_CUDA_TRY(cudaGetDevice(&dev));
  if (dev >= 64 || attr_smem[dev] < smem) {
    HKM_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    if (dev < 64) attr_smem[dev] = smem;
  }
  kern<<<grid, kWsThreads, smem, st>>>(p);
  HKM_CUDA_TRY(cudaGetLastError());
  return HKM_OK;
}
template <int KS_T, int NT>
static int launch_ws(const LloydParams& p, int grid, cudaStream_t st) {
  constexpr int SW = 8;   // see ws_score_warps
#ifdef HKM_DEV_FAST
  if (const char* e = getenv("HKM_WS_SW")) {     // developer sweep of the score/reduce split
    const int sw = atoi(e), ns = getenv("HKM_WS_NSET") ? atoi(getenv("HKM_WS_NSET")) : 2;
#define HKM_WS_TRY(S, N2) if (sw == S && ns == N2) return launch_ws_n<KS_T, NT, S, N2>(p, grid, st);
    HKM_WS_TRY(8, 1) HKM_WS_TRY(8, 2) HKM_WS_TRY(9, 1) HKM_WS_TRY(9, 2) HKM_WS_TRY(10, 1) HKM_WS_TRY(10, 2) HKM_WS_TRY(11, 1) HKM_WS_TRY(11, 2)
    HKM_WS_TRY(12, 1) HKM_WS_TRY(12, 2)
#undef HKM_WS_TRY
  }
#endif
  if (ws_nset(p.d, NT * 8, kWsWarps - SW) == 2) return launch_ws_n<KS_T, NT, SW, 2>(p, grid, st);
  return launch_ws_n<KS_T, NT, SW, 1>(p, grid, st);
}

#define HKM_CAT2(a, b) a##b
#define HKM_CAT(a, b) HKM_CAT2(a, b)

int HKM_CAT(dispatch_lloyd_nt, HKM_NT)(const LloydParams& p, const LaunchPlan& pl, int grid, int mode, cudaStream_t st) {
  constexpr int NT = HKM_NT;
  const int ks = p.ks;
  // silhouette passes: k <= 16 on the barrier-free tensor-pipe kernel, larger k on the gather kernel
  if constexpr (NT <= 2) {
    if ((mode == 1 || mode == 2) && ks <= 16 && ks * NT <= 32 && pl.ctas_per_sm >= 2 && plan_lloyd_mma(p.d, NT * 8).stages >= 3 &&
        lloyd_variant_override() != 1 && (mode == 2 || NT * ((ks + 1) / 2 + 1) <= 14)) {
#define HKM_SIL(KS) if (ks == KS) return mode == 1 ? launch_mma<KS, NT, 1>(p, grid, st) : launch_mma<KS, NT, 2>(p, grid, st);
#ifndef HKM_DEV_FAST
      HKM_SIL(1) HKM_SIL(2) HKM_SIL(3) HKM_SIL(5) HKM_SIL(7) HKM_SIL(10) HKM_SIL(12) HKM_SIL(16)
#endif
      HKM_SIL(4) HKM_SIL(6) HKM_SIL(8)
#undef HKM_SIL
    }
  }
  if (mode == 1) return launch_smemb<NT, 1>(p, pl, grid, st);
  if (mode == 2) return launch_smemb<NT, 2>(p, pl, grid, st);
  if (mode == 3) return launch_smemb<NT, 3>(p, pl, grid, st);   // weighted Lloyd pass (weightCol)
  if constexpr (NT >= 2) {
    // 8 < k, d <= 32: warp-specialised variant (score warps on the fp64 tensor pipe, reduce warps with shared-memory
    // accumulators) whenever at least three stages fit next to the accumulators
    const int variant = lloyd_variant_override();
    const bool fits = p.d <= 32 && ws_stages(p.d, NT * 8, kWsWarps - ws_score_warps(NT), 1) >= 3;
    // measured on the same box against the gather / tensor-pipe variants (tools/ws_check.py): k <= 24 wins here only
    // at d = 29..32, k = 25..32 from d = 17 up except d = 25..28 (centre fragments no longer fit in registers there),
    // k > 32 everywhere
    const bool wins = NT >= 5 ? true : (NT == 4 ? (p.d > 16 && ks != 7) : ks == 8);
    const bool wanted = variant == 3 || ((variant == 0 || variant == 5) && wins);
    if (mode == 0 && p.do_reduce == 1 && p.out_mode == 0 && fits && wanted) {
      const int grid1 = grid / (pl.ctas_per_sm > 0 ? pl.ctas_per_sm : 1);
#define HKM_WS(KS) if constexpr (KS * NT <= 24) { if (ks == KS) return launch_ws<KS, NT>(p, grid1, st); }
#ifndef HKM_DEV_FAST
      HKM_WS(1) HKM_WS(2) HKM_WS(3) HKM_WS(5) HKM_WS(7)
#endif
      HKM_WS(4) HKM_WS(6) HKM_WS(8)
#undef HKM_WS
      return launch_ws<0, NT>(p, grid1, st);
    }
  }
  if constexpr (NT <= 2) {
    // k <= 16: reduce on the tensor pipe as well (see hkm_lloyd_mma.cuh) when the pass is a plain Lloyd step
    const int variant = lloyd_variant_override();
    const bool eligible = p.do_reduce == 1 && p.out_mode == 0 && ks <= 16 && ks * NT <= 32 && plan_lloyd_mma(p.d, NT * 8).stages >= 3;
    if (eligible && variant != 1 && pl.ctas_per_sm >= 2) {
#define HKM_MMA(KS) if (ks == KS) return launch_mma<KS, NT>(p, grid, st);
#ifndef HKM_DEV_FAST
      HKM_MMA(1) HKM_MMA(2) HKM_MMA(3) HKM_MMA(5) HKM_MMA(7) HKM_MMA(10) HKM_MMA(12) HKM_MMA(16)
#endif
      HKM_MMA(4) HKM_MMA(6) HKM_MMA(8)
#undef HKM_MMA
    }
  }
  // register-resident B fragments while KS*NT doubles per lane stay affordable (24 under the 128-register cap of NT 3..6)
#define HKM_REGB(KS) if constexpr (KS * NT <= ((NT >= 3 && NT <= 6) ? 24 : 32)) { if (ks == KS) return launch_regb<KS, NT>(p, pl, grid, st); }
#ifndef HKM_DEV_FAST
  HKM_REGB(1) HKM_REGB(2) HKM_REGB(3) HKM_REGB(4) HKM_REGB(5) HKM_REGB(6) HKM_REGB(7)
#endif
  HKM_REGB(8)
#ifndef HKM_DEV_FAST
  HKM_REGB(10) HKM_REGB(12) HKM_REGB(16)
#else
  HKM_REGB(4) HKM_REGB(6)
#endif
#undef HKM_REGB
  return launch_smemb<NT>(p, pl, grid, st);
}

}  // namespace hkm

// Instantiations of lloyd_tp_kernel for one padded cluster count (compile with -DHKM_TP_KP=16|32|48|64).
#include "../../include/hkm.h"
#include "hkm_lloyd_launch.cuh"
#include "hkm_lloyd_tp.cuh"

#ifndef HKM_TP_KP
#error "compile with -DHKM_TP_KP=<16|32|48|64>"
#endif

namespace hkm {

template <int KP, int CVEC, int RMODE, int WCFG, int ND>
static int launch_tp(const LloydParams& p, const TpPlan& L, cudaStream_t st) {
  auto kern = lloyd_tp_kernel<KP, CVEC, RMODE, WCFG, ND>;
  static int attr_smem[64];
  int dev = 0; HKM_CUDA_TRY(cudaGetDevice(&dev));
  if (dev >= 64 || attr_smem[dev] < L.total) {
    HKM_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L.total));
    if (dev < 64) attr_smem[dev] = L.total;
  }
  kern<<<sm_count() * tp_ctas_per_sm(WCFG), tp_warps(WCFG) * 32, L.total, st>>>(p, L);
  HKM_CUDA_TRY(cudaGetLastError());
  return HKM_OK;
}

// ND = ceil(d / 8) is a compile-time loop bound of the converter, the norm read and the DMMA reduce
template <int KP, int CVEC, int RMODE, int WCFG, int LO, int HI>
static int launch_tp_nd(const LloydParams& p, const TpPlan& L, cudaStream_t st) {
  switch (L.nd) {
#define HKM_TP_ND(N) case N: if constexpr (N >= LO && N <= HI) return launch_tp<KP, CVEC, RMODE, WCFG, N>(p, L, st); break;
    HKM_TP_ND(1) HKM_TP_ND(2) HKM_TP_ND(3) HKM_TP_ND(4) HKM_TP_ND(5) HKM_TP_ND(6) HKM_TP_ND(7) HKM_TP_ND(8)
#undef HKM_TP_ND
  }
  set_error("lloyd_tp: no instantiation for d=%d (reduce mode %d, warp budget %d)", p.d, RMODE, WCFG);
  return HKM_ERR_UNSUPPORTED;
}

#define HKM_CAT2(a, b) a##b
#define HKM_CAT(a, b) HKM_CAT2(a, b)

int HKM_CAT(dispatch_lloyd_tp, HKM_TP_KP)(const LloydParams& p, const TpPlan& L, cudaStream_t st) {
  constexpr int KP = HKM_TP_KP;
  const int cvec = (p.d & 1) ? 1 : 2;
#define HKM_TP_TRY(CV, RM, W, LO, HI) if (cvec == CV && L.rmode == RM && L.wcfg == W) return launch_tp_nd<KP, CV, RM, W, LO, HI>(p, L, st);
  // private sets (d <= 32), every warp budget the chooser can pick
  HKM_TP_TRY(1, 0, 1, 1, 4) HKM_TP_TRY(2, 0, 1, 1, 4) HKM_TP_TRY(1, 0, 0, 1, 4) HKM_TP_TRY(2, 0, 0, 1, 4)
  if constexpr (KP >= 48) { HKM_TP_TRY(1, 0, 2, 1, 4) HKM_TP_TRY(2, 0, 2, 1, 4) }
  // one shared set, clusters owned per warp (even d)
  HKM_TP_TRY(2, 2, 0, 1, 8) HKM_TP_TRY(2, 2, 3, 1, 8)
  // one-hot DMMA (k <= 16)
  if constexpr (KP == 16) {
    HKM_TP_TRY(1, 3, 1, 1, 4) HKM_TP_TRY(2, 3, 1, 1, 4) HKM_TP_TRY(2, 3, 0, 1, 8) HKM_TP_TRY(1, 3, 0, 1, 4)
    HKM_TP_TRY(2, 3, 4, 1, 2) HKM_TP_TRY(2, 0, 4, 1, 2)     // two CTAs per SM (d <= 16)
  }
#undef HKM_TP_TRY
  set_error("lloyd_tp: no instantiation for d=%d k=%d (reduce mode %d, warp budget %d)", p.d, p.k, L.rmode, L.wcfg);
  return HKM_ERR_UNSUPPORTED;
}

}  // namespace hkm

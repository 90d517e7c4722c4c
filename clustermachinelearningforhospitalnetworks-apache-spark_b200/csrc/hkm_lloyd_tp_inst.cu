// Instantiations of lloyd_tp_kernel for one padded cluster count (compile with -DHKM_TP_KP=16|32|48|64).
#include "../../include/hkm.h"
#include "hkm_lloyd_launch.cuh"
#include "hkm_lloyd_tp.cuh"

#ifndef HKM_TP_KP
#error "compile with -DHKM_TP_KP=<16|32|48|64>"
#endif

namespace hkm {

template <int KP, int CVEC, int RMODE, int SG>
static int launch_tp(const LloydParams& p, const TpPlan& L, cudaStream_t st) {
  auto kern = lloyd_tp_kernel<KP, CVEC, RMODE, SG>;
  static int attr_smem[64];
  int dev = 0; HKM_CUDA_TRY(cudaGetDevice(&dev));
  if (dev >= 64 || attr_smem[dev] < L.total) {
    HKM_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L.total));
    if (dev < 64) attr_smem[dev] = L.total;
  }
  kern<<<sm_count(), kTpThreads, L.total, st>>>(p, L);
  HKM_CUDA_TRY(cudaGetLastError());
  return HKM_OK;
}

#define HKM_CAT2(a, b) a##b
#define HKM_CAT(a, b) HKM_CAT2(a, b)

int HKM_CAT(dispatch_lloyd_tp, HKM_TP_KP)(const LloydParams& p, const TpPlan& L, cudaStream_t st) {
  constexpr int KP = HKM_TP_KP;
  const bool odd = (p.d & 1) != 0;
#define HKM_TP_TRY(CV, RM, SGV) if ((odd ? 1 : 2) == CV && L.rmode == RM && L.sg == SGV) return launch_tp<KP, CV, RM, SGV>(p, L, st);
  HKM_TP_TRY(1, 0, 1) HKM_TP_TRY(2, 0, 1) HKM_TP_TRY(2, 1, 1)
  HKM_TP_TRY(2, 2, 1)
  if constexpr (KP >= 48) { HKM_TP_TRY(1, 0, 2) HKM_TP_TRY(2, 0, 2) HKM_TP_TRY(2, 1, 2) HKM_TP_TRY(2, 2, 2) }
#undef HKM_TP_TRY
  set_error("lloyd_tp: no instantiation for d=%d k=%d (rmode %d, groups %d)", p.d, p.k, L.rmode, L.sg);
  return HKM_ERR_UNSUPPORTED;
}

}  // namespace hkm

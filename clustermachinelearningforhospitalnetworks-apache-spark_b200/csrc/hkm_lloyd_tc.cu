// Launch glue for the tcgen05 screening variant of the Lloyd pass (hkm_lloyd_tc.cuh).
#include "../../include/hkm.h"
#include "hkm_lloyd_launch.cuh"
#include "hkm_lloyd_tc.cuh"

namespace hkm {

bool lloyd_tc_eligible(const LloydParams& p, int mode) {
  return mode == 0 && p.out_mode == 0 && p.k > 32 && p.k <= 64 && p.d > 32 && p.d <= 64 && (p.d % 2) == 0;   // measured: beats the fp64 gather variant from k > 32
}

template <int NT>
static int launch_tc(LloydParams p, cudaStream_t st) {
  const int k_pad = NT * 8;
  p.stages = 2;
  const int smem = tc_smem_layout(p.d, k_pad, p.stages).total;
  auto kern = lloyd_tc_kernel<NT, 2>;
  static int attr_smem[64];
  int dev = 0; HKM_CUDA_TRY(cudaGetDevice(&dev));
  if (dev >= 64 || attr_smem[dev] < smem) {
    HKM_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    if (dev < 64) attr_smem[dev] = smem;
  }
  kern<<<sm_count(), kTcThreads, smem, st>>>(p);
  HKM_CUDA_TRY(cudaGetLastError());
  return HKM_OK;
}

int launch_lloyd_tc(const LloydParams& p, cudaStream_t st) {
  return p.k <= 32 ? launch_tc<4>(p, st) : launch_tc<8>(p, st);
}

}  // namespace hkm

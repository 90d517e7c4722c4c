// Launch glue for the tcgen05 screening variant of the Lloyd pass (hkm_lloyd_tc.cuh).
#include "../../include/hkm.h"
#include "hkm_lloyd_launch.cuh"
#include "hkm_lloyd_tc.cuh"

namespace hkm {

bool lloyd_tc_eligible(const LloydParams& p, int mode) {
  // measured against the fp64 variants on the same box (5 M rows): wins from k > 24 at 32 < d <= 64 (d = 64, k = 32:
  // 0.97 vs 1.20 ms; d = 48, k = 32: 0.94 vs 1.01); loses at k <= 24 and, to the warp-specialised variant, at d <= 32
  return mode == 0 && p.out_mode == 0 && p.k > 24 && p.k <= 64 && p.d > 32 && p.d <= 64 && (p.d % 2) == 0;
}

template <int NT, int FPL>
static int launch_tc(LloydParams p, cudaStream_t st) {
  const int k_pad = NT * 8;
  p.stages = 2;
  while (p.stages < 4 && tc_smem_layout(p.d, k_pad, p.stages + 1).total <= 226 * 1024) ++p.stages;
  const int smem = tc_smem_layout(p.d, k_pad, p.stages).total;
  auto kern = lloyd_tc_kernel<NT, FPL>;
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
  if (p.d <= 32) return p.k <= 32 ? launch_tc<4, 1>(p, st) : launch_tc<8, 1>(p, st);
  return p.k <= 32 ? launch_tc<4, 2>(p, st) : launch_tc<8, 2>(p, st);
}

}  // namespace hkm

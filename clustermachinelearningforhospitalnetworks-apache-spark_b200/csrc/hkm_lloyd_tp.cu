// Host side of the pipelined tensor-core Lloyd pass (hkm_lloyd_tp.cuh): shape rules and per-k_pad dispatch.
#include "../../include/hkm.h"
#include "hkm_lloyd_launch.cuh"
#include "hkm_lloyd_tp.cuh"

namespace hkm {

int dispatch_lloyd_tp16(const LloydParams&, const TpPlan&, cudaStream_t);
int dispatch_lloyd_tp32(const LloydParams&, const TpPlan&, cudaStream_t);
int dispatch_lloyd_tp48(const LloydParams&, const TpPlan&, cudaStream_t);
int dispatch_lloyd_tp64(const LloydParams&, const TpPlan&, cudaStream_t);

// Plain passes (unit weights, labels and/or sums, no per-row distances) with 8 < k <= 64 and d <= 64 (odd d only up to 32).
bool lloyd_tp_eligible(const LloydParams& p, int mode) {
  const int variant = lloyd_variant_override();
  if (variant != 0 && variant != 4 && !(variant == 40 || variant == 42 || variant == 43 || (variant >= 50 && variant <= 54))) return false;
  if (mode != 0 || p.w != nullptr || p.out_mode != 0 || p.mind2_out != nullptr || p.do_reduce > 1) return false;
  if (p.k > 64 || p.d > 64 || ((p.d & 1) && p.d > 32)) return false;
  if (p.k <= 8 && variant == 0) return false;               // k <= 8 stays on the fp64 tensor-pipe kernel (HBM-bound already)
  const int kp = ((p.k + 15) / 16) * 16;
  // measured against the fp64 kernels on the same box (tools/tp_check.py bench wide, profiles/r02_tp_vs_round1.jsonl): with k <= 16
  // the pipelined variant wins for 8 < d <= 32; thin rows (d <= 8) and d > 32 stay on the fp64 tensor-pipe kernel
  if (variant == 0 && kp == 16 && (p.d <= 8 || p.d > 32)) return false;
  return tp_choose(p.d, kp, (variant >= 40 && variant < 50) ? variant - 40 : -1, variant >= 50 ? variant - 50 : -1).stages >= 2;
}

static unsigned long long* g_tp_prof = nullptr;

int launch_lloyd_tp(const LloydParams& p, cudaStream_t st) {
  const int kp = ((p.k + 15) / 16) * 16;
  const int variant = lloyd_variant_override();
  TpPlan L = tp_choose(p.d, kp, (variant >= 40 && variant < 50) ? variant - 40 : -1, variant >= 50 ? variant - 50 : -1);
  if (L.stages < 2) { set_error("lloyd_tp: d=%d k=%d does not fit the pinned scheme", p.d, p.k); return HKM_ERR_UNSUPPORTED; }
  L.prof = g_tp_prof;
  switch (kp) {
    case 16: return dispatch_lloyd_tp16(p, L, st);
    case 32: return dispatch_lloyd_tp32(p, L, st);
    case 48: return dispatch_lloyd_tp48(p, L, st);
    case 64: return dispatch_lloyd_tp64(p, L, st);
  }
  set_error("lloyd_tp: k=%d outside 1..64", p.k);
  return HKM_ERR_UNSUPPORTED;
}

}  // namespace hkm

// Developer hook (not part of include/hkm.h): device buffer of 18 x 8 u64 that CTA 0 of the pipelined pass fills with
// per-warp cycle counters (total, wait sites, refine, listed rows, tiles); null switches the profile off.
extern "C" __attribute__((visibility("default"))) void hkm_debug_tp_profile(void* buf) { hkm::g_tp_prof = reinterpret_cast<unsigned long long*>(buf); }

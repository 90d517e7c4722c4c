// Host side of the warp-group tensor-core Lloyd pass (hkm_lloyd_wg.cuh): shape rule, plan and dispatch.
#include "../../include/hkm.h"
#include "hkm_lloyd_launch.cuh"
#include "hkm_lloyd_wg.cuh"

namespace hkm {

static unsigned long long* g_wg_prof = nullptr;
static int g_wg_groups = -1;    // developer override of the group count

// Plain passes (unit weights, labels and/or sums, no per-row distances) with 8 < k <= 32 and d <= 32.
bool lloyd_wg_eligible(const LloydParams& p, int mode) {
  const int variant = lloyd_variant_override();
  if (variant != 0 && variant != 6) return false;
  if (mode != 0 || p.w != nullptr || p.out_mode != 0 || p.mind2_out != nullptr || p.do_reduce > 1) return false;
  if (p.k > 32 || p.d > 32) return false;
  if (p.k <= 8 && variant == 0) return false;
  const int kp = ((p.k + 15) / 16) * 16;
  // measured against the pipelined variant on the same box (tools/tp_check.py wgbench): ahead for k <= 16 with 8 < d <= 24
  if (variant == 0 && !(kp == 16 && p.d > 8 && p.d <= 24)) return false;
  return wg_choose(p.d, kp, g_wg_groups).stages > 0;
}

template <int KP, int CVEC, int ND, int DFIX = 0, int GFIX = 0>
static int launch_wg(const LloydParams& p, const WgPlan& L, cudaStream_t st) {
  constexpr int MAXT = GFIX > 0 ? GFIX * 128 : ((ND <= 2 && KP == 16) ? 640 : 512);
  if (L.groups * 128 > MAXT) { set_error("lloyd_wg: %d groups exceed the thread bound %d of this shape", L.groups, MAXT); return HKM_ERR_UNSUPPORTED; }
  auto kern = lloyd_wg_kernel<KP, CVEC, ND, MAXT, DFIX, GFIX>;
  static int attr_smem[64];
  int dev = 0; HKM_CUDA_TRY(cudaGetDevice(&dev));
  if (dev >= 64 || attr_smem[dev] < L.total) {
    HKM_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L.total));
    if (dev < 64) attr_smem[dev] = L.total;
  }
  kern<<<sm_count(), L.groups * 128, L.total, st>>>(p, L);
  HKM_CUDA_TRY(cudaGetLastError());
  return HKM_OK;
}

template <int KP, int CVEC>
static int launch_wg_nd(const LloydParams& p, const WgPlan& L, cudaStream_t st) {
  switch (L.nd) {
    case 1: return launch_wg<KP, CVEC, 1>(p, L, st);
    case 2: return launch_wg<KP, CVEC, 2>(p, L, st);
    case 3: return launch_wg<KP, CVEC, 3>(p, L, st);
    case 4: return launch_wg<KP, CVEC, 4>(p, L, st);
  }
  set_error("lloyd_wg: d=%d outside 1..32", p.d);
  return HKM_ERR_UNSUPPORTED;
}

int launch_lloyd_wg(const LloydParams& p, cudaStream_t st) {
  const int kp = ((p.k + 15) / 16) * 16;
  WgPlan L = wg_choose(p.d, kp, g_wg_groups);
  if (L.stages == 0) { set_error("lloyd_wg: d=%d k=%d does not fit", p.d, p.k); return HKM_ERR_UNSUPPORTED; }
  L.prof = g_wg_prof;
  const bool odd = (p.d & 1) != 0;
  // the hot shapes with d and the group count folded into the code (same plan by construction: checked)
#define HKM_WG_FIXED(KPV, DV, GV) \
  if (kp == KPV && p.d == DV && L.groups == GV && L.stages == wg_stages(DV, KPV, GV)) return launch_wg<KPV, 2, (DV + 7) / 8, DV, GV>(p, L, st);
  HKM_WG_FIXED(16, 16, 6) HKM_WG_FIXED(16, 16, 5) HKM_WG_FIXED(16, 16, 4)
#undef HKM_WG_FIXED
  if (kp == 16) return odd ? launch_wg_nd<16, 1>(p, L, st) : launch_wg_nd<16, 2>(p, L, st);
  if (kp == 32) return odd ? launch_wg_nd<32, 1>(p, L, st) : launch_wg_nd<32, 2>(p, L, st);
  set_error("lloyd_wg: k=%d outside 1..32", p.k);
  return HKM_ERR_UNSUPPORTED;
}

}  // namespace hkm

// Developer hooks (not part of include/hkm.h): per-warp cycle counters of CTA 0 ([24][8] u64, null = off) and the group count
// (<= 0: pick per shape).
extern "C" __attribute__((visibility("default"))) void hkm_debug_wg_profile(void* buf) { hkm::g_wg_prof = reinterpret_cast<unsigned long long*>(buf); }
extern "C" __attribute__((visibility("default"))) void hkm_debug_wg_groups(int g) { hkm::g_wg_groups = g > 0 ? g : -1; }

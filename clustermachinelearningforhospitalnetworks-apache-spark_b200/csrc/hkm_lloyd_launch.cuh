// Host-side launch plan and per-NT dispatch of lloyd_pass_kernel.  The kernel template is
// instantiated in hkm_lloyd_inst.cu, compiled once per NT (-DHKM_NT=1..8) so the build
// parallelises across translation units.
#pragma once
#include "../../include/hkm.h"
#include "hkm_lloyd.cuh"
#include "hkm_lloyd_mma.cuh"
#include "hkm_lloyd_ws.cuh"

namespace hkm {

struct LaunchPlan { int stages; int ctas_per_sm; int smem; };

// thin rows (d <= 8, k <= 8: the hospital table's own shape): the pass is bound by per-row latency, not by bytes or
// the tensor pipe, so three CTAs share an SM (64 registers)
inline bool plan_thin(int d, int k_pad) { return d <= 8 && k_pad <= 8; }

inline LaunchPlan plan_lloyd(int d, int k_pad) {
  // two CTAs per SM when at least 3 stages fit in ~112 KB each, else one CTA with a deep ring
  LaunchPlan pl{};
  const int budget2 = 112 * 1024, budget1 = 224 * 1024;
  const int fixed = smem_layout(d, k_pad, 0).total;
  if (plan_thin(d, k_pad)) {
    const int st = (74 * 1024 - fixed) / stage_bytes(d);
    pl.stages = st > kMaxStages ? kMaxStages : st; pl.ctas_per_sm = 3;
    pl.smem = smem_layout(d, k_pad, pl.stages).total;
    return pl;
  }
  int st2 = (budget2 - fixed) / stage_bytes(d);
  if (st2 >= 3) { pl.stages = st2 > kMaxStages ? kMaxStages : st2; pl.ctas_per_sm = 2; }
  else {
    int st1 = (budget1 - fixed) / stage_bytes(d);
    pl.stages = st1 > kMaxStages ? kMaxStages : st1; pl.ctas_per_sm = 1;
  }
  pl.smem = smem_layout(d, k_pad, pl.stages).total;
  return pl;
}

// Plan for the small-k tensor-pipe variant (hkm_lloyd_mma.cuh): two CTAs of 9 warps per SM (three for thin rows).
inline LaunchPlan plan_lloyd_mma(int d, int k_pad) {
  LaunchPlan pl{};
  const int budget2 = plan_thin(d, k_pad) ? 74 * 1024 : 112 * 1024;
  const int fixed = mma_smem_layout(d, k_pad, 0).total;
  int st = (budget2 - fixed) / stage_bytes(d);
  pl.stages = st > kMaxStages ? kMaxStages : st;
  pl.ctas_per_sm = plan_thin(d, k_pad) ? 3 : 2;
  pl.smem = mma_smem_layout(d, k_pad, pl.stages > 0 ? pl.stages : 0).total;
  return pl;
}
// tcgen05 screening variant (hkm_lloyd_tc.cuh): 16 < k <= 64, even 32 < d <= 64, plain Lloyd / predict passes
bool lloyd_tc_eligible(const LloydParams& p, int mode);
int launch_lloyd_tc(const LloydParams& p, cudaStream_t st);
// pipelined tcgen05 variant (hkm_lloyd_tp.cuh): 8 < k <= 64, d <= 64, plain passes without per-row distances
bool lloyd_tp_eligible(const LloydParams& p, int mode);
int launch_lloyd_tp(const LloydParams& p, cudaStream_t st);
// warp-group tcgen05 variant (hkm_lloyd_wg.cuh): 8 < k <= 32, d <= 32, plain passes without per-row distances
bool lloyd_wg_eligible(const LloydParams& p, int mode);
int launch_lloyd_wg(const LloydParams& p, cudaStream_t st);
// limb-reduction variant (hkm_lloyd_wl.cuh): REUSE passes with 8 < k <= 64, d in {16, 32}, given the column range record that
// launch_column_range leaves in the tail of the scratch during the STORE pass
bool lloyd_wl_shape(const LloydParams& p, int mode);
size_t wl_tail_bytes();
size_t wl_done_offset();      // offset of the "pass produced" flag inside the record
size_t assign_reduce_scratch_core(int d, int k);
int launch_column_range(const double* X, int64_t n, int d, void* tail, cudaStream_t st);
int launch_lloyd_wl(const LloydParams& p, void* tail, cudaStream_t st);
// the same kernel as a distance-returning pass (k-means|| cost update, candidate weights): no record needed
bool lloyd_wd_shape(const LloydParams& p, int mode);
int launch_lloyd_wd(const LloydParams& p, cudaStream_t st);
// 0: pick per shape (default); 1: always the gather variant; 2: the tensor-pipe variant wherever it is built;
// 3: the warp-specialised variant wherever it is built; 4: the pipelined tcgen05 variant wherever it is built;
// 5: the round-1 choice (no pipelined tcgen05 variant); 6: the warp-group tcgen05 variant wherever it is built; 7: as 0 (the limb-reduction variant is a REUSE-pass refinement of 0).  Set through hkm_set_lloyd_variant (tuning and tests).
int lloyd_variant_override();

// Validated, dispatched Lloyd data pass shared by the C entry points (hkm_api.cu).
struct AssignArgs {
  const double* X; int64_t n; int d; const double* C; int k; const double* weights;
  double* partial_out; int32_t* labels_out; double* mind2_out; double* xnorm2_sum; int xnorm2_mode;
  const hkm_lloyd_state* state; int do_reduce; int out_mode; int label_base;
  double* C_update; double tol; hkm_lloyd_state* state_rw;   // fused single-rank driver step (hkm_lloyd_step)
  void* const* peer_bufs; int xrank, xworld; unsigned long long xepoch; int* xerr;   // hkm_lloyd_step_peer
  int mode; const int32_t* labels_in; const double* aux;   // mode 1/2: silhouette statistics / score pass
  void* scratch; size_t scratch_bytes; void* stream;
};
int assign_pass(const AssignArgs& a, const char* who);

// one function per NT, defined in hkm_lloyd_inst.cu
int dispatch_lloyd_nt1(const LloydParams&, const LaunchPlan&, int grid, int mode, cudaStream_t);
int dispatch_lloyd_nt2(const LloydParams&, const LaunchPlan&, int grid, int mode, cudaStream_t);
int dispatch_lloyd_nt3(const LloydParams&, const LaunchPlan&, int grid, int mode, cudaStream_t);
int dispatch_lloyd_nt4(const LloydParams&, const LaunchPlan&, int grid, int mode, cudaStream_t);
int dispatch_lloyd_nt5(const LloydParams&, const LaunchPlan&, int grid, int mode, cudaStream_t);
int dispatch_lloyd_nt6(const LloydParams&, const LaunchPlan&, int grid, int mode, cudaStream_t);
int dispatch_lloyd_nt7(const LloydParams&, const LaunchPlan&, int grid, int mode, cudaStream_t);
int dispatch_lloyd_nt8(const LloydParams&, const LaunchPlan&, int grid, int mode, cudaStream_t);

}  // namespace hkm

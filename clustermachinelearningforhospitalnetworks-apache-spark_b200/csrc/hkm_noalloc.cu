// hkm_assign_reduce_noalloc: the Lloyd data pass for callers that are not able to allocate.
//
// Who that is: code running where cudaMalloc is off the table -- a plugin handed a fixed set of
// device buffers by its host process (a JVM/Spark executor that owns the GPU memory pool, a serving
// runtime with a pre-sized arena), or a latency-critical loop that must not take the allocator lock.
// Such a caller can still pass its inputs and outputs, but has nowhere to get the per-CTA partial
// sums the kernel needs.  This variant keeps that scratch inside the library instead, in a
// file-scope device array of a fixed, modest size.  Every GPU that loads libhkm.so holds the array
// until the process exits.
#include "../../include/hkm.h"
#include "hkm_lloyd_launch.cuh"
#include <mutex>

namespace {

// The scratch itself.  On the host this name is only a stand-in for the symbol, not a device address:
// it must never be handed to a kernel directly -- cudaGetSymbolAddress below asks the runtime for the
// real address on the current device.
static __device__ __align__(256) unsigned char g_noalloc_scratch[HKM_NOALLOC_SCRATCH_BYTES];

// Device address of g_noalloc_scratch per device ordinal, resolved on first use.
void* g_scratch_addr[64];

int scratch_for_current_device(void** out) {
  int dev = 0;
  HKM_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) { hkm::set_error("hkm_assign_reduce_noalloc: device ordinal %d not supported", dev); return HKM_ERR_UNSUPPORTED; }
  static std::mutex mu;                              // two host threads may make their first call on a device together
  std::lock_guard<std::mutex> lock(mu);
  if (g_scratch_addr[dev] == nullptr) {
    void* ptr = nullptr;
    HKM_CUDA_TRY(cudaGetSymbolAddress(&ptr, g_noalloc_scratch));  // with `dev` current: its copy of the array
    if (ptr == nullptr) { hkm::set_error("hkm_assign_reduce_noalloc: cudaGetSymbolAddress returned null"); return HKM_ERR_CUDA; }
    g_scratch_addr[dev] = ptr;
  }
  *out = g_scratch_addr[dev];
  return HKM_OK;
}

}  // namespace

extern "C" int hkm_assign_reduce_noalloc(const double* X, int64_t n, int d, const double* C, int k, const double* weights,
                                         double* partial_out, int32_t* labels_out, double* mind2_out, double* xnorm2_sum,
                                         int xnorm2_mode, const hkm_lloyd_state* state, int do_reduce, void* stream) {
  if (d < 1 || k < 1) { hkm::set_error("hkm_assign_reduce_noalloc: bad d/k"); return HKM_ERR_INVALID; }
  if (hkm_needs_workspace(d, k)) {
    hkm::set_error("hkm_assign_reduce_noalloc: k=%d, d=%d runs in centre chunks with per-row workspace; use hkm_assign_reduce for this shape", k, d);
    return HKM_ERR_UNSUPPORTED;
  }
  const size_t need = hkm::assign_reduce_scratch_core(d, k);   // without the optional column range record
  if (need > (size_t)HKM_NOALLOC_SCRATCH_BYTES) {
    hkm::set_error("hkm_assign_reduce_noalloc: k=%d, d=%d needs %zu bytes of scratch but the built-in array holds %zu; "
                   "use hkm_assign_reduce with caller-allocated scratch for this shape",
                   k, d, need, (size_t)HKM_NOALLOC_SCRATCH_BYTES);
    return HKM_ERR_UNSUPPORTED;
  }
  void* scratch = nullptr;
  int rc = scratch_for_current_device(&scratch);
  if (rc != HKM_OK) return rc;
  // The array keeps whatever the previous call left in it.  Only the 256-byte header (the "last CTA
  // done" ticket) has to start at zero; the per-CTA slots are fully rewritten before they are read.
  HKM_CUDA_TRY(cudaMemsetAsync(scratch, 0, 256, reinterpret_cast<cudaStream_t>(stream)));
  hkm::AssignArgs a{};
  a.X = X; a.n = n; a.d = d; a.C = C; a.k = k; a.weights = weights; a.partial_out = partial_out; a.labels_out = labels_out;
  a.mind2_out = mind2_out; a.xnorm2_sum = xnorm2_sum; a.xnorm2_mode = xnorm2_mode; a.state = state; a.do_reduce = do_reduce ? 1 : 0;
  a.scratch = scratch; a.scratch_bytes = (size_t)HKM_NOALLOC_SCRATCH_BYTES; a.stream = stream;
  return hkm::assign_pass(a, "hkm_assign_reduce_noalloc");
}

// sf_dcn_forward_static_f32: the tensor-core DCN forward (dcn_tc.cu) for callers that cannot allocate.
//
// No counterpart in the reference; added for code under CUDA-graph capture,
// tight serving loops and harnesses that forbid cudaMalloc.  The tcgen05 path needs scratch memory for the bf16
// hi/mid copy of the weights (sf_dcn_workspace_bytes); sf_dcn_forward_ws_f32 takes it from the caller, this entry
// point takes it from a file-scope device array of SF_DCN_STATIC_SCRATCH_BYTES that every GPU loading the library
// holds until exit.  Nothing here (or below it) calls cudaMalloc / cudaFree or an allocating library routine.
//
// The array's NAME is only a host-side stand-in in host code: the device address comes from cudaGetSymbolAddress,
// asked once per device with that device current, and that address is what the kernels receive.
//
// Shared state: every call on a device uses the same array.  The weight-split kernel rewrites the whole region a
// call uses (including the zero padding of the reduction dimension), so nothing has to start at zero and a call
// never sees a previous call's contents; calls that overlap on the GPU (different streams, or graphs replayed
// concurrently) RACE on it -- serialise them or use sf_dcn_forward_ws_f32 with one buffer per stream.
#include <atomic>
#include <mutex>

#include "common.cuh"
#include "../../include/stereoformer_b200.h"

static __device__ __align__(256) unsigned char g_dcn_static_scratch[SF_DCN_STATIC_SCRATCH_BYTES];

namespace {

constexpr int kMaxDevices = 64;
std::atomic<void*> g_scratch_ptr[kMaxDevices];
std::mutex g_scratch_mutex;

// Device address of the scratch array on the CURRENT device (looked up once per device).
int static_scratch(void** out) {
  int dev = -1;
  SF_CUDA(cudaGetDevice(&dev));
  SF_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "dcn_forward_static: device ordinal %d out of range", dev);
  void* ptr = g_scratch_ptr[dev].load(std::memory_order_acquire);
  if (ptr == nullptr) {
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    ptr = g_scratch_ptr[dev].load(std::memory_order_relaxed);
    if (ptr == nullptr) {
      SF_CUDA(cudaGetSymbolAddress(&ptr, g_dcn_static_scratch));
      SF_CHECK_ARG(ptr != nullptr && (reinterpret_cast<uintptr_t>(ptr) & 255) == 0,
                   "dcn_forward_static: bad scratch address %p", ptr);
      g_scratch_ptr[dev].store(ptr, std::memory_order_release);
    }
  }
  *out = ptr;
  return SF_OK;
}

}  // namespace

extern "C" size_t sf_dcn_static_scratch_bytes(void) { return SF_DCN_STATIC_SCRATCH_BYTES; }

extern "C" int sf_dcn_forward_static_f32(const float* input, const float* weight, const float* bias,
                                         const float* offset, const float* mask, float* out, int B, int Cin, int H,
                                         int W, int Cout, int kh, int kw, int stride_h, int stride_w, int pad_h,
                                         int pad_w, int dil_h, int dil_w, int group, int deformable_group, int algo,
                                         void* stream) {
  const size_t need = sf_dcn_workspace_bytes(Cin, Cout, kh, kw, group);
  SF_CHECK_ARG(need <= static_cast<size_t>(SF_DCN_STATIC_SCRATCH_BYTES),
               "dcn_forward_static: this shape needs %zu scratch bytes, the built-in scratch holds %zu "
               "(SF_DCN_STATIC_SCRATCH_BYTES); use sf_dcn_forward_ws_f32 with a caller-provided buffer",
               need, static_cast<size_t>(SF_DCN_STATIC_SCRATCH_BYTES));
  void* scratch = nullptr;
  const int rc = static_scratch(&scratch);
  if (rc) return rc;
  return sf_dcn_forward_ws_f32(input, weight, bias, offset, mask, out, B, Cin, H, W, Cout, kh, kw, stride_h, stride_w,
                               pad_h, pad_w, dil_h, dil_w, group, deformable_group, scratch,
                               static_cast<size_t>(SF_DCN_STATIC_SCRATCH_BYTES), algo, stream);
}

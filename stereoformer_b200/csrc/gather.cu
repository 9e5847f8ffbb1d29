// Consumer side of the fused multi-GPU disparity gather: see the protocol in
// include/stereoformer_b200.h (sf_local_window_gather_signal_f32 / sf_gather_wait_f32).
//
// The reference gathers its outputs on device 0 with nn.DataParallel (train_file/trainer_DLC_GRU.py:102-106).  Here the
// producers' window kernels store straight into the consumer's slots over NVLink; this kernel is what makes step s
// *complete* on the consumer: it waits for every source's arrival word, moves the slot into the caller's gathered
// tensor and tells the sources the slot is free again.
#include "common.cuh"
#include "../../include/stereoformer_b200.h"

namespace sf {

struct GatherWaitParams {
  unsigned long long* sig_local;
  unsigned long long* src_sig[16];
  int src_ranks[16];
  int n_src, my_rank;
  const float* slots;
  long long slot_stride;
  float* out;
  long long n_elems;
};

__global__ void __launch_bounds__(256) gather_wait_kernel(const GatherWaitParams p) {
  const unsigned long long step = p.sig_local[SF_SIG_WAIT_EPOCH];
  // every CTA polls for itself: the acquire also drops this SM's stale L1 lines of the slot
  if (threadIdx.x < p.n_src)
    spin_until_ge_sys(p.sig_local + SF_SIG_ARRIVED + p.src_ranks[threadIdx.x], step + 1, p.sig_local + SF_SIG_ERROR);
  __syncthreads();
  if (p.out) {
    const float* src = p.slots + static_cast<long long>(step & 1ULL) * p.slot_stride;
    const long long n4 = p.n_elems >> 2;
    const float4* s4 = reinterpret_cast<const float4*>(src);
    float4* d4 = reinterpret_cast<float4*>(p.out);
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4;
         i += static_cast<long long>(gridDim.x) * blockDim.x)
      d4[i] = __ldcv(s4 + i);
    if (blockIdx.x == 0)
      for (long long i = (n4 << 2) + threadIdx.x; i < p.n_elems; i += blockDim.x) p.out[i] = __ldcv(src + i);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned long long prev = atomicAdd(p.sig_local + SF_SIG_WAIT_CTA, 1ULL);
    if (prev == gridDim.x - 1) {
      p.sig_local[SF_SIG_WAIT_CTA] = 0ULL;
      __threadfence_system();
      for (int r = 0; r < p.n_src; ++r) st_release_sys_u64(p.src_sig[r] + SF_SIG_CONSUMED + p.my_rank, step + 1);
      p.sig_local[SF_SIG_WAIT_EPOCH] = step + 1;
    }
  }
}

}  // namespace sf

extern "C" int sf_gather_wait_f32(unsigned long long* sig_local, unsigned long long* const* h_src_sig,
                                  const int* h_src_ranks, int n_src, int my_rank, const float* slots,
                                  long long slot_stride, float* out, long long n_elems, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(sig_local && h_src_sig && h_src_ranks && slots, "gather_wait: null pointer");
  SF_CHECK_ARG(n_src > 0 && n_src <= 16 && my_rank >= 0 && my_rank < 16, "gather_wait: bad rank count");
  SF_CHECK_ARG(n_elems > 0 && slot_stride >= n_elems, "gather_wait: slot stride smaller than the slot");
  SF_CHECK_ARG(!out || ((reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(slots)) & 15) == 0 &&
                           slot_stride % 4 == 0,
               "gather_wait: slots / out must be 16-byte aligned");
  GatherWaitParams p;
  p.sig_local = sig_local; p.n_src = n_src; p.my_rank = my_rank; p.slots = slots; p.slot_stride = slot_stride;
  p.out = out; p.n_elems = n_elems;
  for (int r = 0; r < 16; ++r) { p.src_sig[r] = nullptr; p.src_ranks[r] = 0; }
  for (int r = 0; r < n_src; ++r) {
    SF_CHECK_ARG(h_src_sig[r] && h_src_ranks[r] >= 0 && h_src_ranks[r] < 16, "gather_wait: bad source %d", r);
    p.src_sig[r] = h_src_sig[r]; p.src_ranks[r] = h_src_ranks[r];
  }
  const long long n4 = n_elems >> 2;
  long long want = out ? (n4 + 255) / 256 / 8 : 1;
  int grid = static_cast<int>(want < 1 ? 1 : (want > 32 ? 32 : want));
  gather_wait_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

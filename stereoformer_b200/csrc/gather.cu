// Consumer side of the fused multi-GPU disparity gather: see the protocol in
// include/stereoformer_b200.h (sf_local_window_gather_signal_f32 / sf_gather_wait_f32).
//
// The reference gathers its outputs on device 0 with nn.DataParallel (train_file/trainer_DLC_GRU.py:102-106).  Here the
// producers' window kernels store {value, step flag} words straight into the consumer's slots over NVLink; this kernel
// is what makes step s *complete* on the consumer: it polls every word until its flag says "step s", unpacks the
// values into the caller's gathered tensor and tells the sources the slot is free again.
#include "common.cuh"
#include "../../include/stereoformer_b200.h"

namespace sf {

struct GatherWaitParams {
  unsigned long long* sig_local;
  unsigned long long* src_sig[16];
  int n_src, my_rank;
  const unsigned long long* slots;
  long long slot_stride;
  float* out;
  long long n_elems;
};

__global__ void __launch_bounds__(256) gather_wait_kernel(const GatherWaitParams p) {
  __shared__ unsigned long long s_step;
  if (threadIdx.x == 0) s_step = atomicAdd(p.sig_local + SF_SIG_WAIT_TICKET, 1ULL) / gridDim.x;
  __syncthreads();
  const unsigned long long step = s_step;
  const uint32_t want = static_cast<uint32_t>(step + 1);
  const unsigned long long* src = p.slots + static_cast<long long>(step % SF_GATHER_SLOTS) * p.slot_stride;
  bool timed_out = false;
  // 4 independent polls in flight per thread (a system-scope load is a long round trip; one at a time made this
  // kernel 17 us for 1 MB); words that have not arrived yet are re-polled individually
  constexpr int U = 4;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i0 = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i0 < p.n_elems; i0 += U * stride) {
    uint2 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * stride;
      v[u] = i < p.n_elems ? ld_ll_sys(src + i) : make_uint2(0u, want);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * stride;
      if (i >= p.n_elems) continue;
      if (v[u].y != want && !timed_out) {
        const long long t0 = clock64();
        do {
          __nanosleep(32);
          v[u] = ld_ll_sys(src + i);
          if (clock64() - t0 > 4000000000LL) { timed_out = true; break; }
        } while (v[u].y != want);
      }
      p.out[i] = __uint_as_float(v[u].x);
    }
  }
  if (timed_out) p.sig_local[SF_SIG_ERROR] = 1ULL;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned long long prev = atomicAdd(p.sig_local + SF_SIG_WAIT_CTA, 1ULL);
    if (prev == gridDim.x - 1) {
      p.sig_local[SF_SIG_WAIT_CTA] = 0ULL;
      __threadfence_system();
      for (int r = 0; r < p.n_src; ++r) st_relaxed_sys_u64(p.src_sig[r] + SF_SIG_CONSUMED + p.my_rank, step + 1);
    }
  }
}

}  // namespace sf

extern "C" int sf_gather_wait_f32(unsigned long long* sig_local, unsigned long long* const* h_src_sig, int n_src,
                                  int my_rank, const unsigned long long* slots, long long slot_stride, float* out,
                                  long long n_elems, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(sig_local && h_src_sig && slots && out, "gather_wait: null pointer");
  SF_CHECK_ARG(n_src > 0 && n_src <= 16 && my_rank >= 0 && my_rank < 16, "gather_wait: bad rank count");
  SF_CHECK_ARG(n_elems > 0 && slot_stride >= n_elems, "gather_wait: slot stride smaller than the slot");
  SF_CHECK_ARG((reinterpret_cast<uintptr_t>(slots) & 7) == 0, "gather_wait: slots must be 8-byte aligned");
  GatherWaitParams p;
  p.sig_local = sig_local; p.n_src = n_src; p.my_rank = my_rank; p.slots = slots; p.slot_stride = slot_stride;
  p.out = out; p.n_elems = n_elems;
  for (int r = 0; r < 16; ++r) p.src_sig[r] = nullptr;
  for (int r = 0; r < n_src; ++r) {
    SF_CHECK_ARG(h_src_sig[r] != nullptr, "gather_wait: null signal block of source %d", r);
    p.src_sig[r] = h_src_sig[r];
  }
  // a fixed grid per gather size (the ticket counter divides by it): 4 words per thread in one round of polls, at most
  // 148 CTAs.  (The CTAs cannot share an SM with the consumer's persistent cost-volume CTA -- registers -- so what counts
  // is their total SM time: 32 CTAs with 4-5 poll rounds each measured slower than 148 with one, 35.6 vs 33.3 us/step.)
  long long want = (n_elems + 256 * 4 - 1) / (256 * 4);
  int grid = static_cast<int>(want < 1 ? 1 : (want > 148 ? 148 : want));
  gather_wait_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

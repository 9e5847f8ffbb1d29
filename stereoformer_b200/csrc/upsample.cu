// Convex upsampling  (the step right after the hot path).
//
// Reference: models/utils/upsample.py:22-38 (upsample_convex8); RAFT-Stereo's upsample_flow uses factor 4.
//   out[n,0,F*y+i,F*x+j] = sum_k softmax_k(mask[n,k*F*F+i*F+j,y,x]) * F * disp[n,0,y+ky-1,x+kx-1]   (zero padded)
// One thread per low-resolution pixel: mask reads are coalesced along x, each thread writes F rows of F
// consecutive floats (full 32-byte sectors for F = 8).
#include "common.cuh"
#include "../../include/stereoformer_b200.h"

namespace sf {

template <int F>
__global__ void __launch_bounds__(128)
upsample_convex_kernel(const float* __restrict__ disp, const float* __restrict__ mask, float* __restrict__ out, int N,
                       int H, int W) {
  const size_t HW = static_cast<size_t>(H) * W, n = static_cast<size_t>(N) * HW;
  for (size_t idx = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; idx < n;
       idx += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int x = idx % W, y = (idx / W) % H;
    const size_t b = idx / HW;
    float nb[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      const int yy = y + k / 3 - 1, xx = x + k % 3 - 1;
      nb[k] = (yy >= 0 && yy < H && xx >= 0 && xx < W) ? __fmul_rn((float)F, __ldg(disp + b * HW + yy * W + xx)) : 0.f;
    }
    const float* mp = mask + b * 9 * F * F * HW + (idx % HW);
    float* op = out + b * F * F * HW + (static_cast<size_t>(y) * F) * (W * F) + x * F;
    for (int i = 0; i < F; ++i) {
      float res[F];
#pragma unroll
      for (int j = 0; j < F; ++j) {
        float mv[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) mv[k] = __ldg(mp + static_cast<size_t>(k * F * F + i * F + j) * HW);
        float m = mv[0];
#pragma unroll
        for (int k = 1; k < 9; ++k) m = fmaxf(m, mv[k]);
        float s = 0.f, a = 0.f;
#pragma unroll
        for (int k = 0; k < 9; ++k) {
          const float e = __expf(mv[k] - m);
          s += e; a = fmaf(e, nb[k], a);
        }
        res[j] = a / s;
      }
      float* row = op + static_cast<size_t>(i) * (W * F);
      if (F % 4 == 0) {
#pragma unroll
        for (int j = 0; j < F; j += 4)
          *reinterpret_cast<float4*>(row + j) = make_float4(res[j], res[j + 1], res[j + 2], res[j + 3]);
      } else {
#pragma unroll
        for (int j = 0; j < F; ++j) row[j] = res[j];
      }
    }
  }
}

}  // namespace sf

extern "C" int sf_upsample_convex_f32(const float* disp, const float* mask, float* out, int N, int H, int W,
                                      int factor, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(disp && mask && out, "upsample_convex: null pointer");
  SF_CHECK_ARG(N > 0 && H > 0 && W > 0, "upsample_convex: non-positive dimension");
  SF_CHECK_ARG(factor == 8 || factor == 4, "upsample_convex: factor must be 8 or 4, got %d", factor);
  SF_CHECK_ARG((reinterpret_cast<uintptr_t>(out) & 15) == 0, "upsample_convex: out must be 16-byte aligned");
  size_t n = static_cast<size_t>(N) * H * W;
  int grid = static_cast<int>(min(static_cast<size_t>(sm_count()) * 16, (n + 127) / 128));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (factor == 8) upsample_convex_kernel<8><<<grid, 128, 0, s>>>(disp, mask, out, N, H, W);
  else upsample_convex_kernel<4><<<grid, 128, 0, s>>>(disp, mask, out, N, H, W);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

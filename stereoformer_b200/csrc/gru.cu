// ConvGRU gate arithmetic around the three cuDNN convolutions  (the caller of the DLC window).
//
// Reference: models/LocalCostVolume/Attempts/update.py:36-49 (ConvGRU.forward)
//   x1 = cat(input, prev_h);  z = sigmoid(conv_z(x1));  b = sigmoid(conv_b(x1))
//   s  = cat(b * prev_h, input);  g = tanh(conv_g(s))  (Softsign variant: g / (1 + |g|))
//   h  = (1 - z) * prev_h + z * g
// Eager PyTorch spends 9 elementwise launches and 7 temporaries on this per iteration; here it is two streaming
// kernels: `reset` writes s (gate, product and concatenation in one pass) and `blend` writes h.  The z and b
// pre-activations may live in one [B, 2C, H, W] tensor (one convolution with stacked weights), hence the batch strides.
// Algorithmic bytes per pixel: reset (2C + Cx) read + (C + Cx) written, blend 3C read + C written, 4 B each.
#include <algorithm>
#include "common.cuh"
#include "../../include/stereoformer_b200.h"

namespace sf {

__device__ __forceinline__ float sigmoid_f(float x) { return 1.0f / (1.0f + expf(-x)); }
__device__ __forceinline__ float gate_act(float x, int softsign) { return softsign ? x / (1.0f + fabsf(x)) : tanhf(x); }

// one thread per VEC consecutive elements of s_out; per batch the first C*HW elements are gated, the rest copied
template <int VEC>
__global__ void __launch_bounds__(256)
gru_reset_kernel(const float* __restrict__ b_pre, long long b_bstride, const float* __restrict__ prev_h,
                 const float* __restrict__ input, float* __restrict__ s_out, long long CHW, long long XHW, long long n) {
  const long long per = CHW + XHW;
  for (long long i = (blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) * VEC; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x * VEC) {
    const long long b = i / per, r = i - b * per;
    float v[VEC];
    if (r < CHW) {
      float g[VEC], h[VEC];
      if (VEC == 4) {
        *reinterpret_cast<float4*>(g) = *reinterpret_cast<const float4*>(b_pre + b * b_bstride + r);
        if (prev_h) *reinterpret_cast<float4*>(h) = *reinterpret_cast<const float4*>(prev_h + b * CHW + r);
      } else {
        g[0] = b_pre[b * b_bstride + r];
        if (prev_h) h[0] = prev_h[b * CHW + r];
      }
#pragma unroll
      for (int k = 0; k < VEC; ++k) v[k] = prev_h ? __fmul_rn(sigmoid_f(g[k]), h[k]) : 0.f;
    } else {
      if (VEC == 4) *reinterpret_cast<float4*>(v) = *reinterpret_cast<const float4*>(input + b * XHW + (r - CHW));
      else v[0] = input[b * XHW + (r - CHW)];
    }
    if (VEC == 4) *reinterpret_cast<float4*>(s_out + i) = *reinterpret_cast<float4*>(v);
    else s_out[i] = v[0];
  }
}

template <int VEC>
__global__ void __launch_bounds__(256)
gru_blend_kernel(const float* __restrict__ z_pre, long long z_bstride, const float* __restrict__ g_pre,
                 const float* __restrict__ prev_h, float* __restrict__ h_out, long long CHW, long long n, int softsign) {
  for (long long i = (blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) * VEC; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x * VEC) {
    const long long b = i / CHW, r = i - b * CHW;
    float z[VEC], g[VEC], h[VEC], o[VEC];
    if (VEC == 4) {
      *reinterpret_cast<float4*>(z) = *reinterpret_cast<const float4*>(z_pre + b * z_bstride + r);
      *reinterpret_cast<float4*>(g) = *reinterpret_cast<const float4*>(g_pre + i);
      if (prev_h) *reinterpret_cast<float4*>(h) = *reinterpret_cast<const float4*>(prev_h + i);
    } else {
      z[0] = z_pre[b * z_bstride + r];
      g[0] = g_pre[i];
      if (prev_h) h[0] = prev_h[i];
    }
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
      const float zz = sigmoid_f(z[k]);
      const float keep = __fmul_rn(__fsub_rn(1.0f, zz), prev_h ? h[k] : 0.f);      // (1 - z) * prev_h
      o[k] = __fadd_rn(keep, __fmul_rn(zz, gate_act(g[k], softsign)));              // + z * g, no contraction
    }
    if (VEC == 4) *reinterpret_cast<float4*>(h_out + i) = *reinterpret_cast<float4*>(o);
    else h_out[i] = o[0];
  }
}

// backward of the two kernels above (scalar; training only).  reset: gs [B,C+Cx,H,W] -> g_b_pre, g_prev_h, g_input
__global__ void __launch_bounds__(256)
gru_reset_bwd_kernel(const float* __restrict__ b_pre, long long b_bstride, const float* __restrict__ prev_h,
                     const float* __restrict__ gs, float* __restrict__ g_b_pre, float* __restrict__ g_prev_h,
                     float* __restrict__ g_input, long long CHW, long long XHW, long long n) {
  const long long per = CHW + XHW;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long b = i / per, r = i - b * per;
    const float g = gs[i];
    if (r < CHW) {
      const float sg = sigmoid_f(b_pre[b * b_bstride + r]);
      const float h = prev_h ? prev_h[b * CHW + r] : 0.f;
      if (g_b_pre) g_b_pre[b * CHW + r] = g * h * sg * (1.0f - sg);
      if (g_prev_h) g_prev_h[b * CHW + r] = g * sg;
    } else if (g_input) {
      g_input[b * XHW + (r - CHW)] = g;
    }
  }
}

__global__ void __launch_bounds__(256)
gru_blend_bwd_kernel(const float* __restrict__ z_pre, long long z_bstride, const float* __restrict__ g_pre,
                     const float* __restrict__ prev_h, const float* __restrict__ go, float* __restrict__ g_z_pre,
                     float* __restrict__ g_g_pre, float* __restrict__ g_prev_h, long long CHW, long long n, int softsign) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long b = i / CHW, r = i - b * CHW;
    const float z = sigmoid_f(z_pre[b * z_bstride + r]);
    const float x = g_pre[i], a = gate_act(x, softsign), h = prev_h ? prev_h[i] : 0.f, g = go[i];
    const float da = softsign ? 1.0f / ((1.0f + fabsf(x)) * (1.0f + fabsf(x))) : 1.0f - a * a;
    if (g_z_pre) g_z_pre[i] = g * (a - h) * z * (1.0f - z);
    if (g_g_pre) g_g_pre[i] = g * z * da;
    if (g_prev_h) g_prev_h[i] = g * (1.0f - z);
  }
}

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace sf

extern "C" int sf_gru_reset_f32(const float* b_pre, long long b_bstride, const float* prev_h, const float* input,
                                float* s_out, int B, int C, int Cx, int H, int W, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(b_pre && input && s_out, "gru_reset: null pointer");
  SF_CHECK_ARG(B > 0 && C > 0 && Cx > 0 && H > 0 && W > 0, "gru_reset: non-positive dimension");
  const long long HW = 1LL * H * W, CHW = C * HW, XHW = Cx * HW, n = B * (CHW + XHW);
  SF_CHECK_ARG(b_bstride >= CHW, "gru_reset: batch stride %lld smaller than C*H*W", b_bstride);
  const bool vec = CHW % 4 == 0 && XHW % 4 == 0 && b_bstride % 4 == 0 && aligned16(b_pre) && aligned16(input) &&
                   aligned16(s_out) && (!prev_h || aligned16(prev_h));
  const long long work = vec ? n / 4 : n;
  const int grid = static_cast<int>(std::min<long long>(static_cast<long long>(sm_count()) * 8, (work + 255) / 256));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (vec) gru_reset_kernel<4><<<grid, 256, 0, s>>>(b_pre, b_bstride, prev_h, input, s_out, CHW, XHW, n);
  else gru_reset_kernel<1><<<grid, 256, 0, s>>>(b_pre, b_bstride, prev_h, input, s_out, CHW, XHW, n);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

extern "C" int sf_gru_blend_f32(const float* z_pre, long long z_bstride, const float* g_pre, const float* prev_h,
                                float* h_out, int B, int C, int H, int W, int softsign, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(z_pre && g_pre && h_out, "gru_blend: null pointer");
  SF_CHECK_ARG(B > 0 && C > 0 && H > 0 && W > 0, "gru_blend: non-positive dimension");
  const long long CHW = 1LL * C * H * W, n = B * CHW;
  SF_CHECK_ARG(z_bstride >= CHW, "gru_blend: batch stride %lld smaller than C*H*W", z_bstride);
  const bool vec = CHW % 4 == 0 && z_bstride % 4 == 0 && aligned16(z_pre) && aligned16(g_pre) && aligned16(h_out) &&
                   (!prev_h || aligned16(prev_h));
  const long long work = vec ? n / 4 : n;
  const int grid = static_cast<int>(std::min<long long>(static_cast<long long>(sm_count()) * 8, (work + 255) / 256));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (vec) gru_blend_kernel<4><<<grid, 256, 0, s>>>(z_pre, z_bstride, g_pre, prev_h, h_out, CHW, n, softsign);
  else gru_blend_kernel<1><<<grid, 256, 0, s>>>(z_pre, z_bstride, g_pre, prev_h, h_out, CHW, n, softsign);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

extern "C" int sf_gru_reset_backward_f32(const float* b_pre, long long b_bstride, const float* prev_h,
                                         const float* grad_s, float* grad_b_pre, float* grad_prev_h, float* grad_input,
                                         int B, int C, int Cx, int H, int W, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(b_pre && grad_s, "gru_reset_backward: null pointer");
  SF_CHECK_ARG(B > 0 && C > 0 && Cx > 0 && H > 0 && W > 0, "gru_reset_backward: non-positive dimension");
  const long long HW = 1LL * H * W, CHW = C * HW, XHW = Cx * HW, n = B * (CHW + XHW);
  SF_CHECK_ARG(b_bstride >= CHW, "gru_reset_backward: batch stride %lld smaller than C*H*W", b_bstride);
  const int grid = static_cast<int>(std::min<long long>(static_cast<long long>(sm_count()) * 8, (n + 255) / 256));
  gru_reset_bwd_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(b_pre, b_bstride, prev_h, grad_s, grad_b_pre,
                                                                          grad_prev_h, grad_input, CHW, XHW, n);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

extern "C" int sf_gru_blend_backward_f32(const float* z_pre, long long z_bstride, const float* g_pre, const float* prev_h,
                                         const float* grad_h, float* grad_z_pre, float* grad_g_pre, float* grad_prev_h,
                                         int B, int C, int H, int W, int softsign, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(z_pre && g_pre && grad_h, "gru_blend_backward: null pointer");
  SF_CHECK_ARG(B > 0 && C > 0 && H > 0 && W > 0, "gru_blend_backward: non-positive dimension");
  const long long CHW = 1LL * C * H * W, n = B * CHW;
  SF_CHECK_ARG(z_bstride >= CHW, "gru_blend_backward: batch stride %lld smaller than C*H*W", z_bstride);
  const int grid = static_cast<int>(std::min<long long>(static_cast<long long>(sm_count()) * 8, (n + 255) / 256));
  gru_blend_bwd_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(z_pre, z_bstride, g_pre, prev_h, grad_h,
                                                                          grad_z_pre, grad_g_pre, grad_prev_h, CHW, n,
                                                                          softsign);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

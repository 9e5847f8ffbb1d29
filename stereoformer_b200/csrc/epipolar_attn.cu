// Epipolar row cross-attention restricted to the disparity band  (SURVEY 8a row a11).
//
// The reference file models/CrossAttentionCostVolume/cross_attention_cost_volume.py is EMPTY (0 bytes): this
// operator is SPEC-DERIVED from BASELINE.json north_star (parity unpinned):
//   per row y:  S[q,k] = <Q[:,y,q], K[:,y,k]> / sqrt(C)  for 0 <= q-k < band,  P = softmax_k(S),  O[:,y,q] = sum_k P[q,k] V[:,y,k]
// bf16 in / bf16 out, fp32 accumulation and soft-max.
//
// This file is the fp32-FFMA formulation (band correlation -> soft-max -> band-weighted sum), one thread per
// query pixel with the band scores in registers; coalesced bf16 loads along x.
#include <cuda_bf16.h>

#include "common.cuh"
#include "../../include/stereoformer_b200.h"

namespace sf {

constexpr int ATT_MAXBAND = 64;

template <int BAND>
__global__ void __launch_bounds__(128)
epipolar_attn_kernel(const __nv_bfloat16* __restrict__ Q, const __nv_bfloat16* __restrict__ K,
                     const __nv_bfloat16* __restrict__ V, __nv_bfloat16* __restrict__ O, float* __restrict__ scores,
                     int B, int C, int H, int W, int band) {
  const size_t HW = static_cast<size_t>(H) * W, n = static_cast<size_t>(B) * HW;
  const float rs = rsqrtf(static_cast<float>(C));
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int x = i % W;
    const size_t b = i / HW, r = i % HW;
    const int nd = min(band, x + 1);  // keys x, x-1, ..., x-nd+1
    float s[BAND];
#pragma unroll
    for (int d = 0; d < BAND; ++d) s[d] = 0.f;
    const __nv_bfloat16* qp = Q + b * C * HW + r;
    const __nv_bfloat16* kp = K + b * C * HW + r;
    for (int c = 0; c < C; ++c) {
      const float q = __bfloat162float(qp[c * HW]);
      const __nv_bfloat16* kr = kp + c * HW;
#pragma unroll
      for (int d = 0; d < BAND; ++d)
        if (d < nd) s[d] = fmaf(q, __bfloat162float(kr[-d]), s[d]);
    }
    float m = -INFINITY;
#pragma unroll
    for (int d = 0; d < BAND; ++d) {
      s[d] *= rs;
      if (d < nd) m = fmaxf(m, s[d]);
    }
    if (scores) {
#pragma unroll
      for (int d = 0; d < BAND; ++d)
        if (d < band) scores[(b * band + d) * HW + r] = d < nd ? s[d] : -INFINITY;
    }
    float sum = 0.f;
#pragma unroll
    for (int d = 0; d < BAND; ++d) {
      s[d] = d < nd ? __expf(s[d] - m) : 0.f;
      sum += s[d];
    }
    const float inv = 1.0f / sum;
    const __nv_bfloat16* vp = V + b * C * HW + r;
    __nv_bfloat16* op = O + b * C * HW + r;
    for (int c = 0; c < C; ++c) {
      const __nv_bfloat16* vr = vp + c * HW;
      float o = 0.f;
#pragma unroll
      for (int d = 0; d < BAND; ++d)
        if (d < nd) o = fmaf(s[d], __bfloat162float(vr[-d]), o);
      op[c * HW] = __float2bfloat16(o * inv);
    }
  }
}

bool attn_tc_supported(const void* q, const void* k, const void* v, const void* out, int C, int W, int band);
int launch_attn_tc(const void* q, const void* k, const void* v, void* out, int B, int C, int H, int W, int band,
                   cudaStream_t stream);

}  // namespace sf

extern "C" int sf_epipolar_attn_bf16(const void* q, const void* k, const void* v, void* out, float* scores, int B,
                                     int C, int H, int W, int band, int algo, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(q && k && v && out, "epipolar_attn: null pointer");
  SF_CHECK_ARG(B > 0 && C > 0 && H > 0 && W > 0, "epipolar_attn: non-positive dimension");
  SF_CHECK_ARG(band >= 1 && band <= ATT_MAXBAND, "epipolar_attn: band must be in 1..%d, got %d", ATT_MAXBAND, band);
  if (algo == SF_ALGO_EXACT) algo = SF_ALGO_GENERIC;      // fp32-accumulating FFMA kernel
  SF_CHECK_ARG(algo == SF_ALGO_AUTO || algo == SF_ALGO_GENERIC || algo == SF_ALGO_TMA, "epipolar_attn: unknown algo %d",
               algo);
  // tcgen05 path (SF_ALGO_TMA forces it): needs C % 64 == 0, C <= 256, W % 8 == 0, band <= 64, no score output
  const bool tc_ok = attn_tc_supported(q, k, v, out, C, W, band) && scores == nullptr;
  SF_CHECK_ARG(algo != SF_ALGO_TMA || tc_ok,
               "epipolar_attn: tensor-core path needs C %% 64 == 0, C <= 256, W %% 8 == 0, band <= 64, scores == NULL");
  if (algo != SF_ALGO_GENERIC && tc_ok)
    return launch_attn_tc(q, k, v, out, B, C, H, W, band, static_cast<cudaStream_t>(stream));
  const __nv_bfloat16* Q = static_cast<const __nv_bfloat16*>(q);
  const __nv_bfloat16* K = static_cast<const __nv_bfloat16*>(k);
  const __nv_bfloat16* V = static_cast<const __nv_bfloat16*>(v);
  __nv_bfloat16* O = static_cast<__nv_bfloat16*>(out);
  size_t n = static_cast<size_t>(B) * H * W;
  int grid = static_cast<int>(min(static_cast<size_t>(sm_count()) * 16, (n + 127) / 128));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (band <= 16) epipolar_attn_kernel<16><<<grid, 128, 0, s>>>(Q, K, V, O, scores, B, C, H, W, band);
  else if (band <= 32) epipolar_attn_kernel<32><<<grid, 128, 0, s>>>(Q, K, V, O, scores, B, C, H, W, band);
  else if (band <= 48) epipolar_attn_kernel<48><<<grid, 128, 0, s>>>(Q, K, V, O, scores, B, C, H, W, band);
  else epipolar_attn_kernel<64><<<grid, 128, 0, s>>>(Q, K, V, O, scores, B, C, H, W, band);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

/* TEST INFRASTRUCTURE -- CPU restatement of the reference's DCN forward (the reference itself has no CPU
 * implementation: deform/src/cpu/modulated_deform_cpu.cpp:25 raises).  Plain C, double accumulation optional.
 *
 * Follows, step by step:
 *   deform/src/cuda/modulated_deform_im2col_cuda.cuh:25-54   bilinear sample with per-corner zero padding
 *   deform/src/cuda/modulated_deform_im2col_cuda.cuh:140-193 column buffer cols[(c*kh*kw + ij), b, y, x]
 *   deform/src/cuda/modulated_deform_conv_cuda.cu:75-76      output size
 *   deform/src/cuda/modulated_deform_conv_cuda.cu:104-118    per-group out = bias + cols^T * W^T, NCHW result
 * mask == NULL gives DCNv1 (deform/src/cuda/deform_im2col_cuda.cuh:127-189).
 * Pinned by the analytic tests of deform/test.py (zero offset == conv2d, identity, im2col_step invariance)
 * and cross-checked against torchvision.ops.deform_conv2d in tests/test_oracle_dcn.py.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

static float bilinear(const float* im, int width, int height, float h, float w) {
  int h_low = (int)floorf(h), w_low = (int)floorf(w);
  int h_high = h_low + 1, w_high = w_low + 1;
  float lh = h - h_low, lw = w - w_low;
  float hh = 1 - lh, hw = 1 - lw;
  float v1 = 0, v2 = 0, v3 = 0, v4 = 0;
  if (h_low >= 0 && w_low >= 0) v1 = im[h_low * width + w_low];
  if (h_low >= 0 && w_high <= width - 1) v2 = im[h_low * width + w_high];
  if (h_high <= height - 1 && w_low >= 0) v3 = im[h_high * width + w_low];
  if (h_high <= height - 1 && w_high <= width - 1) v4 = im[h_high * width + w_high];
  float w1 = hh * hw, w2 = hh * lw, w3 = lh * hw, w4 = lh * lw;
  return w1 * v1 + w2 * v2 + w3 * v3 + w4 * v4;
}

/* returns 0 on success */
int dcn_oracle_forward(const float* in, const float* weight, const float* bias, const float* offset,
                       const float* mask, float* out, int B, int Cin, int H, int W, int Cout, int kh, int kw, int sh,
                       int sw, int ph, int pw, int dh, int dw, int G, int dg) {
  const int Ho = (H + 2 * ph - (dh * (kh - 1) + 1)) / sh + 1;
  const int Wo = (W + 2 * pw - (dw * (kw - 1) + 1)) / sw + 1;
  if (Ho <= 0 || Wo <= 0 || Cin % G || Cout % G || Cin % dg) return 1;
  const int T = kh * kw, M = Ho * Wo;
  const int cpdg = Cin / dg, Cig = Cin / G, Cog = Cout / G;
  float* cols = (float*)malloc(sizeof(float) * (size_t)Cin * T * M);
  if (!cols) return 2;
  for (int b = 0; b < B; ++b) {
    /* im2col with deformation */
    for (int c = 0; c < Cin; ++c) {
      const int g = c / cpdg;
      const float* im = in + ((size_t)b * Cin + c) * H * W;
      const float* off = offset + ((size_t)b * dg + g) * 2 * T * M;
      const float* msk = mask ? mask + ((size_t)b * dg + g) * T * M : NULL;
      for (int i = 0; i < kh; ++i)
        for (int j = 0; j < kw; ++j) {
          const int t = i * kw + j;
          float* dst = cols + ((size_t)c * T + t) * M;
          for (int y = 0; y < Ho; ++y)
            for (int x = 0; x < Wo; ++x) {
              const int pix = y * Wo + x;
              const float oh = off[(size_t)(2 * t) * M + pix], ow = off[(size_t)(2 * t + 1) * M + pix];
              const float h_im = (float)(y * sh - ph + i * dh) + oh;
              const float w_im = (float)(x * sw - pw + j * dw) + ow;
              float val = 0.f;
              if (h_im > -1 && w_im > -1 && h_im < H && w_im < W) val = bilinear(im, W, H, h_im, w_im);
              dst[pix] = msk ? val * msk[(size_t)t * M + pix] : val;
            }
        }
    }
    /* per-group GEMM + bias */
    for (int g = 0; g < G; ++g)
      for (int o = 0; o < Cog; ++o) {
        const int oc = g * Cog + o;
        float* dst = out + ((size_t)b * Cout + oc) * M;
        const float* wrow = weight + (size_t)oc * Cig * T;
        for (int pix = 0; pix < M; ++pix) {
          double acc = 0.0;
          for (int k = 0; k < Cig * T; ++k) acc += (double)wrow[k] * cols[((size_t)g * Cig * T + k) * M + pix];
          dst[pix] = (float)(acc + bias[oc]);
        }
      }
  }
  free(cols);
  return 0;
}

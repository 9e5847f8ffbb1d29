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

/* ------------------------------------------------------------------------------------------------ backward
 * CPU restatement of the reference's DCN backward, step by step:
 *   deform/src/cuda/modulated_deform_conv_cuda.cu:213-225   columns = W_g^T * grad_out_g  (per group)
 *   modulated_deform_im2col_cuda.cuh:262-328                col2im_coord: grad_offset, grad_mask
 *   modulated_deform_im2col_cuda.cuh:56-83, 197-258         col2im: grad_input via get_gradient_weight over the
 *                                                           5x5 neighbourhood of (int)pos (the reference's loop)
 *   modulated_deform_conv_cuda.cu:250-271                   im2col again, grad_weight += grad_out * cols^T,
 *                                                           grad_bias += grad_out * ones
 * mask == NULL gives DCNv1 (deform_conv_cuda.cu, deform_im2col_cuda.cuh): mask = 1, no grad_mask.
 * Pinned (tests/test_oracle_dcn.py) against autograd through the float64 dense restatement of the forward and
 * against numerical differences, the way deform/test.py:304-349 pins the CUDA kernels with gradcheck. */
static float gradient_weight(float ah, float aw, int h, int w, int height, int width) {
  if (ah <= -1 || ah >= height || aw <= -1 || aw >= width) return 0;
  int hl = (int)floorf(ah), wl = (int)floorf(aw);
  int hh = hl + 1, wh = wl + 1;
  float weight = 0;
  if (h == hl && w == wl) weight = (h + 1 - ah) * (w + 1 - aw);
  if (h == hl && w == wh) weight = (h + 1 - ah) * (aw + 1 - w);
  if (h == hh && w == wl) weight = (ah + 1 - h) * (w + 1 - aw);
  if (h == hh && w == wh) weight = (ah + 1 - h) * (aw + 1 - w);
  return weight;
}

static float coordinate_weight(float ah, float aw, int height, int width, const float* im, int bp_dir) {
  if (ah <= -1 || ah >= height || aw <= -1 || aw >= width) return 0;
  int hl = (int)floorf(ah), wl = (int)floorf(aw);
  int hh = hl + 1, wh = wl + 1;
  float weight = 0;
  if (bp_dir == 0) {
    if (hl >= 0 && wl >= 0) weight += -1 * (wl + 1 - aw) * im[hl * width + wl];
    if (hl >= 0 && wh <= width - 1) weight += -1 * (aw - wl) * im[hl * width + wh];
    if (hh <= height - 1 && wl >= 0) weight += (wl + 1 - aw) * im[hh * width + wl];
    if (hh <= height - 1 && wh <= width - 1) weight += (aw - wl) * im[hh * width + wh];
  } else {
    if (hl >= 0 && wl >= 0) weight += -1 * (hl + 1 - ah) * im[hl * width + wl];
    if (hl >= 0 && wh <= width - 1) weight += (hl + 1 - ah) * im[hl * width + wh];
    if (hh <= height - 1 && wl >= 0) weight += -1 * (ah - hl) * im[hh * width + wl];
    if (hh <= height - 1 && wh <= width - 1) weight += (ah - hl) * im[hh * width + wh];
  }
  return weight;
}

int dcn_oracle_backward(const float* in, const float* weight, const float* offset, const float* mask,
                        const float* gout, float* gin, float* gweight, float* gbias, float* goffset, float* gmask,
                        int B, int Cin, int H, int W, int Cout, int kh, int kw, int sh, int sw, int ph, int pw,
                        int dh, int dw, int G, int dg) {
  const int Ho = (H + 2 * ph - (dh * (kh - 1) + 1)) / sh + 1;
  const int Wo = (W + 2 * pw - (dw * (kw - 1) + 1)) / sw + 1;
  if (Ho <= 0 || Wo <= 0 || Cin % G || Cout % G || Cin % dg) return 1;
  const int T = kh * kw, M = Ho * Wo;
  const int cpdg = Cin / dg, Cig = Cin / G, Cog = Cout / G;
  float* cols = (float*)malloc(sizeof(float) * (size_t)Cin * T * M);
  if (!cols) return 2;
  memset(gin, 0, sizeof(float) * (size_t)B * Cin * H * W);
  memset(gweight, 0, sizeof(float) * (size_t)Cout * Cig * T);
  memset(gbias, 0, sizeof(float) * (size_t)Cout);
  for (int b = 0; b < B; ++b) {
    /* columns = W^T * grad_out */
    for (int c = 0; c < Cin; ++c) {
      const int g = c / Cig;
      for (int t = 0; t < T; ++t)
        for (int pix = 0; pix < M; ++pix) {
          double acc = 0.0;
          for (int o = 0; o < Cog; ++o)
            acc += (double)weight[((size_t)(g * Cog + o) * Cig + (c - g * Cig)) * T + t] *
                   gout[((size_t)b * Cout + g * Cog + o) * M + pix];
          cols[((size_t)c * T + t) * M + pix] = (float)acc;
        }
    }
    /* col2im_coord + col2im */
    for (int g = 0; g < dg; ++g) {
      const float* off = offset + ((size_t)b * dg + g) * 2 * T * M;
      const float* msk = mask ? mask + ((size_t)b * dg + g) * T * M : NULL;
      for (int t = 0; t < T; ++t) {
        const int i = t / kw, j = t % kw;
        for (int y = 0; y < Ho; ++y)
          for (int x = 0; x < Wo; ++x) {
            const int pix = y * Wo + x;
            const float oh = off[(size_t)(2 * t) * M + pix], ow = off[(size_t)(2 * t + 1) * M + pix];
            const float m = msk ? msk[(size_t)t * M + pix] : 1.f;
            const float pos_h = (float)(y * sh - ph + i * dh) + oh;
            const float pos_w = (float)(x * sw - pw + j * dw) + ow;
            float inv_h = pos_h, inv_w = pos_w;
            const int outside = (inv_h <= -1 || inv_w <= -1 || inv_h >= H || inv_w >= W);
            if (outside) inv_h = inv_w = -2;
            double vh = 0, vw = 0, mv = 0;
            for (int c = g * cpdg; c < (g + 1) * cpdg; ++c) {
              const float* im = in + ((size_t)b * Cin + c) * H * W;
              const float col = cols[((size_t)c * T + t) * M + pix];
              if (!outside) mv += (double)col * bilinear(im, W, H, inv_h, inv_w);
              vh += (double)coordinate_weight(inv_h, inv_w, H, W, im, 0) * col * m;
              vw += (double)coordinate_weight(inv_h, inv_w, H, W, im, 1) * col * m;
              /* col2im, .cuh:232-250 */
              const float top = col * m;
              const int cur_h = (int)pos_h, cur_w = (int)pos_w;
              float* gim = gin + ((size_t)b * Cin + c) * H * W;
              for (int dy = -2; dy <= 2; ++dy)
                for (int dx = -2; dx <= 2; ++dx)
                  if (cur_h + dy >= 0 && cur_h + dy < H && cur_w + dx >= 0 && cur_w + dx < W &&
                      fabsf(pos_h - (cur_h + dy)) < 1 && fabsf(pos_w - (cur_w + dx)) < 1)
                    gim[(cur_h + dy) * W + cur_w + dx] += gradient_weight(pos_h, pos_w, cur_h + dy, cur_w + dx, H, W) * top;
            }
            if (goffset) {
              goffset[(((size_t)b * dg + g) * 2 * T + 2 * t) * M + pix] = (float)vh;
              goffset[(((size_t)b * dg + g) * 2 * T + 2 * t + 1) * M + pix] = (float)vw;
            }
            if (gmask && mask) gmask[(((size_t)b * dg + g) * T + t) * M + pix] = (float)mv;
          }
      }
    }
    /* im2col again, grad_weight and grad_bias */
    for (int c = 0; c < Cin; ++c) {
      const int g = c / cpdg;
      const float* im = in + ((size_t)b * Cin + c) * H * W;
      const float* off = offset + ((size_t)b * dg + g) * 2 * T * M;
      const float* msk = mask ? mask + ((size_t)b * dg + g) * T * M : NULL;
      for (int t = 0; t < T; ++t) {
        const int i = t / kw, j = t % kw;
        for (int y = 0; y < Ho; ++y)
          for (int x = 0; x < Wo; ++x) {
            const int pix = y * Wo + x;
            const float h_im = (float)(y * sh - ph + i * dh) + off[(size_t)(2 * t) * M + pix];
            const float w_im = (float)(x * sw - pw + j * dw) + off[(size_t)(2 * t + 1) * M + pix];
            float val = 0.f;
            if (h_im > -1 && w_im > -1 && h_im < H && w_im < W) val = bilinear(im, W, H, h_im, w_im);
            cols[((size_t)c * T + t) * M + pix] = msk ? val * msk[(size_t)t * M + pix] : val;
          }
      }
    }
    for (int oc = 0; oc < Cout; ++oc) {
      const int g = oc / Cog;
      const float* go = gout + ((size_t)b * Cout + oc) * M;
      double sb = 0;
      for (int pix = 0; pix < M; ++pix) sb += go[pix];
      gbias[oc] += (float)sb;
      for (int k = 0; k < Cig * T; ++k) {
        double acc = 0;
        const float* col = cols + ((size_t)g * Cig * T + k) * M;
        for (int pix = 0; pix < M; ++pix) acc += (double)go[pix] * col[pix];
        gweight[(size_t)oc * Cig * T + k] += (float)acc;
      }
    }
  }
  free(cols);
  return 0;
}

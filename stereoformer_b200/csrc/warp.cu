// Disparity warp and the DLC-GRU guidance stage  (SURVEY 8a rows a7 and a6 (i),(iii),(iv)).
//
// Reference: utils/disparity_warper.py:48-106 (meshgrid, normalize_coords, disp_warp, LRwarp_error);
// models/LocalCostVolume/Attempts/update.py:68-86 (BasicGuidanceNet up to the conv encoders).
//
// The reference normalises with (size-1) but samples with F.grid_sample's default align_corners=False, so the
// source position is ((2*((x-d)/(W-1))-1+1)*W-1)/2, not x-d, and y maps to y*H/(H-1)-0.5: a true 2-row bilinear
// gather.  That quirk is reproduced op by op (pixel_math.h); the normalised grid is bit-identical to the
// reference's.  One thread per pixel, all image channels; loads are coalesced along x.
#include "common.cuh"
#include "pixel_math.h"
#include "../../include/stereoformer_b200.h"

namespace sf {

struct Bilin {
  int x0, y0;                  // north-west corner
  float wnw, wne, wsw, wse;    // ATen weights
  bool inx0, inx1, iny0, iny1; // corner in-bounds flags
  float gx, gy, ix, iy;
};

// ATen grid_sampler_2d (bilinear), align_corners=False; border mode clips the source coordinate.
__device__ __forceinline__ Bilin warp_setup(float x, float y, float d, int H, int W, bool border, int recip) {
  Bilin r;
  r.gx = sfm::warp_norm(sfm::fadd(x, -d), W, recip);
  r.gy = sfm::warp_norm(sfm::fadd(y, 0.f), H, recip);
  float ix = sfm::warp_unnorm(r.gx, W), iy = sfm::warp_unnorm(r.gy, H);
  if (border) {
    ix = fminf((float)(W - 1), fmaxf(ix, 0.f));
    iy = fminf((float)(H - 1), fmaxf(iy, 0.f));
  }
  r.ix = ix; r.iy = iy;
  const float fx = floorf(ix), fy = floorf(iy);
  // out-of-range coordinates (zeros padding, huge disparities) are clamped before the int cast
  r.x0 = (int)fminf(fmaxf(fx, -2.f), (float)W);
  r.y0 = (int)fminf(fmaxf(fy, -2.f), (float)H);
  const float ex = fx + 1.f, ey = fy + 1.f;
  r.wnw = (ex - ix) * (ey - iy);
  r.wne = (ix - fx) * (ey - iy);
  r.wsw = (ex - ix) * (iy - fy);
  r.wse = (ix - fx) * (iy - fy);
  r.inx0 = r.x0 >= 0 && r.x0 < W; r.inx1 = r.x0 + 1 >= 0 && r.x0 + 1 < W;
  r.iny0 = r.y0 >= 0 && r.y0 < H; r.iny1 = r.y0 + 1 >= 0 && r.y0 + 1 < H;
  return r;
}

// ys0 = global index of the first source row the plane holds (0 for a whole image, the halo start for a row band).
__device__ __forceinline__ float warp_sample(const float* __restrict__ plane, const Bilin& g, int W, int ys0 = 0) {
  float o = 0.f;
  const float* r0 = plane + (static_cast<ptrdiff_t>(g.y0) - ys0) * W + g.x0;
  if (g.iny0 && g.inx0) o += __ldg(r0) * g.wnw;
  if (g.iny0 && g.inx1) o += __ldg(r0 + 1) * g.wne;
  if (g.iny1 && g.inx0) o += __ldg(r0 + W) * g.wsw;
  if (g.iny1 && g.inx1) o += __ldg(r0 + W + 1) * g.wse;
  return o;
}

// valid_mask of disp_warp (:102-105): zeros-padding sample of an all-ones image, <0.9999 -> 0, >0 -> 1.
__device__ __forceinline__ float warp_mask(float x, float y, float d, int H, int W, int recip) {
  const Bilin g = warp_setup(x, y, d, H, W, false, recip);
  float o = 0.f;
  if (g.iny0 && g.inx0) o += g.wnw;
  if (g.iny0 && g.inx1) o += g.wne;
  if (g.iny1 && g.inx0) o += g.wsw;
  if (g.iny1 && g.inx1) o += g.wse;
  if (o < 0.9999f) o = 0.f;
  if (o > 0.f) o = 1.f;
  return o;
}

// Row-band form (partition 2): the output covers rows [band.y0, band.y0 + H) of an image of band.hfull rows,
// the source planes hold rows [band.ys0, band.ys0 + band.srows).  The sampling coordinates are computed from the
// GLOBAL row index and height (y_src = y*Hfull/(Hfull-1) - 0.5 depends on both), so a band reproduces the rows of
// the whole-image call bit for bit.  A whole image is the band {H, 0, 0, H}.
struct RowBand { int hfull, y0, ys0, srows; };

__global__ void __launch_bounds__(256)
disp_warp_kernel(const float* __restrict__ img, const float* __restrict__ disp, float* __restrict__ warped,
                 float* __restrict__ mask, float* __restrict__ dbg_grid, float* __restrict__ dbg_src, int B, int C,
                 int H, int W, int border, int recip, int* __restrict__ neg_flag, RowBand band) {
  const size_t HW = static_cast<size_t>(H) * W, n = static_cast<size_t>(B) * HW;
  const size_t SHW = static_cast<size_t>(band.srows) * W;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int x = i % W, y = (i / W) % H;
    const size_t b = i / HW;
    const float d = disp[i];
    if (d < 0.f && neg_flag) *neg_flag = 1;  // reference asserts disp.min() >= 0 (disparity_warper.py:93)
    const float yg = (float)(y + band.y0);
    const Bilin g = warp_setup((float)x, yg, d, band.hfull, W, border != 0, recip);
    if (dbg_grid) { dbg_grid[2 * i] = g.gx; dbg_grid[2 * i + 1] = g.gy; }
    if (dbg_src) { dbg_src[2 * i] = g.ix; dbg_src[2 * i + 1] = g.iy; }
    const float mk = mask ? warp_mask((float)x, yg, d, band.hfull, W, recip) : 0.f;
    for (int c = 0; c < C; ++c) {
      const size_t o = (b * C + c) * HW + (i % HW);
      warped[o] = warp_sample(img + (b * C + c) * SHW, g, W, band.ys0);
      if (mask) mask[o] = mk;
    }
  }
}

// BasicGuidanceNet.forward up to the encoders, one pass per pixel:
//   error_map = disp_warp(right, disp) - left      (update.py:82-83)
//   uncertainty[d] = softmax(cv)[d] * d * (d-disp) * (d-disp)   (update.py:71,86)
__global__ void __launch_bounds__(256)
guidance_kernel(const float* __restrict__ cv, const float* __restrict__ disp, const float* __restrict__ left,
                const float* __restrict__ right, float* __restrict__ err, float* __restrict__ unc, int B, int D, int C,
                int H, int W, int recip, int* __restrict__ neg_flag, RowBand band) {
  const size_t HW = static_cast<size_t>(H) * W, n = static_cast<size_t>(B) * HW;
  const size_t SHW = static_cast<size_t>(band.srows) * W;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int x = i % W, y = (i / W) % H;
    const size_t b = i / HW, r = i % HW;
    const float d = disp[i];
    if (d < 0.f && neg_flag) *neg_flag = 1;
    const Bilin g = warp_setup((float)x, (float)(y + band.y0), d, band.hfull, W, true, recip);
    for (int c = 0; c < C; ++c) {
      const size_t o = (b * C + c) * HW + r;
      err[o] = warp_sample(right + (b * C + c) * SHW, g, W, band.ys0) - left[o];
    }
    const float* src = cv + b * D * HW + r;
    float m = -INFINITY;
    for (int k = 0; k < D; ++k) m = fmaxf(m, __ldg(src + k * HW));
    float s = 0.f;
    for (int k = 0; k < D; ++k) s += __expf(__ldg(src + k * HW) - m);
    const float inv = 1.0f / s;
    float* dst = unc + b * D * HW + r;
    for (int k = 0; k < D; ++k) {
      const float p = __expf(__ldg(src + k * HW) - m) * inv;
      const float e = sfm::fsub((float)k, d);
      dst[k * HW] = sfm::fmul(sfm::fmul(sfm::fmul(p, (float)k), e), e);
    }
  }
}

// disp_warp backward = ATen grid_sampler_2d_backward (bilinear, align_corners=False) chained with the grid
// arithmetic of disparity_warper.py:96-98: d ix / d disp = (W/2) * (-2/(W-1)), times 0 where border clipping is
// active (clip_coordinates_set_grad: coordinates <= 0 or >= size-1).  The validity mask is a thresholded constant.
__global__ void __launch_bounds__(256)
disp_warp_bwd_kernel(const float* __restrict__ img, const float* __restrict__ disp, const float* __restrict__ gw,
                     float* __restrict__ gimg, float* __restrict__ gdisp, int B, int C, int H, int W, int border,
                     int recip) {
  const size_t HW = static_cast<size_t>(H) * W, n = static_cast<size_t>(B) * HW;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int x = i % W, y = (i / W) % H;
    const size_t b = i / HW, r = i % HW;
    const float d = disp[i];
    const Bilin g = warp_setup((float)x, (float)y, d, H, W, border != 0, recip);
    float mult = 1.f;
    if (border) {
      const float raw = sfm::warp_unnorm(g.gx, W);
      if (raw <= 0.f || raw >= (float)(W - 1)) mult = 0.f;
    }
    const float wy_n = (float)(g.y0 + 1) - g.iy, wy_s = g.iy - (float)g.y0;   // (iy_s - iy), (iy - iy_n)
    float gix = 0.f;
    for (int c = 0; c < C; ++c) {
      const size_t plane = (b * C + c) * HW;
      const float go = gw[plane + r];
      const float* r0 = img + plane + static_cast<size_t>(g.y0) * W + g.x0;
      const float nw = (g.iny0 && g.inx0) ? __ldg(r0) : 0.f, ne = (g.iny0 && g.inx1) ? __ldg(r0 + 1) : 0.f;
      const float sw = (g.iny1 && g.inx0) ? __ldg(r0 + W) : 0.f, se = (g.iny1 && g.inx1) ? __ldg(r0 + W + 1) : 0.f;
      gix += go * ((ne - nw) * wy_n + (se - sw) * wy_s);
      if (gimg) {
        float* q0 = gimg + plane + static_cast<size_t>(g.y0) * W + g.x0;
        if (g.iny0 && g.inx0) atomicAdd(q0, go * g.wnw);
        if (g.iny0 && g.inx1) atomicAdd(q0 + 1, go * g.wne);
        if (g.iny1 && g.inx0) atomicAdd(q0 + W, go * g.wsw);
        if (g.iny1 && g.inx1) atomicAdd(q0 + W + 1, go * g.wse);
      }
    }
    if (gdisp) gdisp[i] = -gix * mult * ((float)W / (float)(W - 1));
  }
}

// guidance backward (BasicGuidanceNet prologue, update.py:68-86): with p = softmax(cv), u_d = p_d * d * (d - disp)^2,
//   g_cv = p * (q - sum_d p_d q_d),  q_d = g_unc_d * d * (d - disp)^2          (soft-max backward)
//   g_disp = -2 * sum_d g_unc_d * p_d * d * (d - disp)  +  warp term of error_map = warp(right, disp) - left
// The images are network inputs: their gradients are not produced.
__global__ void __launch_bounds__(256)
guidance_bwd_kernel(const float* __restrict__ cv, const float* __restrict__ disp, const float* __restrict__ right,
                    const float* __restrict__ gerr, const float* __restrict__ gunc, float* __restrict__ gcv,
                    float* __restrict__ gdisp, int B, int D, int C, int H, int W, int recip) {
  const size_t HW = static_cast<size_t>(H) * W, n = static_cast<size_t>(B) * HW;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int x = i % W, y = (i / W) % H;
    const size_t b = i / HW, r = i % HW;
    const float d = disp[i];
    float gd = 0.f;
    if (gerr) {
      const Bilin g = warp_setup((float)x, (float)y, d, H, W, true, recip);
      const float raw = sfm::warp_unnorm(g.gx, W);
      const float mult = (raw <= 0.f || raw >= (float)(W - 1)) ? 0.f : 1.f;
      const float wy_n = (float)(g.y0 + 1) - g.iy, wy_s = g.iy - (float)g.y0;
      float gix = 0.f;
      for (int c = 0; c < C; ++c) {
        const size_t plane = (b * C + c) * HW;
        const float* r0 = right + plane + static_cast<size_t>(g.y0) * W + g.x0;
        const float nw = (g.iny0 && g.inx0) ? __ldg(r0) : 0.f, ne = (g.iny0 && g.inx1) ? __ldg(r0 + 1) : 0.f;
        const float sw = (g.iny1 && g.inx0) ? __ldg(r0 + W) : 0.f, se = (g.iny1 && g.inx1) ? __ldg(r0 + W + 1) : 0.f;
        gix += gerr[plane + r] * ((ne - nw) * wy_n + (se - sw) * wy_s);
      }
      gd = -gix * mult * ((float)W / (float)(W - 1));
    }
    if (gunc) {
      const float* src = cv + b * D * HW + r;
      const float* gu = gunc + b * D * HW + r;
      float m = -INFINITY;
      for (int k = 0; k < D; ++k) m = fmaxf(m, __ldg(src + k * HW));
      float s = 0.f;
      for (int k = 0; k < D; ++k) s += __expf(__ldg(src + k * HW) - m);
      const float inv = 1.0f / s;
      float dot = 0.f;
      for (int k = 0; k < D; ++k) {
        const float pk = __expf(__ldg(src + k * HW) - m) * inv, e = (float)k - d, gk = __ldg(gu + k * HW);
        dot = fmaf(pk, gk * (float)k * e * e, dot);
        gd = fmaf(-2.0f * gk * pk, (float)k * e, gd);
      }
      if (gcv) {
        float* dst = gcv + b * D * HW + r;
        for (int k = 0; k < D; ++k) {
          const float pk = __expf(__ldg(src + k * HW) - m) * inv, e = (float)k - d;
          dst[k * HW] = pk * (__ldg(gu + k * HW) * (float)k * e * e - dot);
        }
      }
    } else if (gcv) {
      float* dst = gcv + b * D * HW + r;
      for (int k = 0; k < D; ++k) dst[k * HW] = 0.f;
    }
    if (gdisp) gdisp[i] = gd;
  }
}

}  // namespace sf

// A band's source rows must cover every row the 2-row bilinear gather can touch: y-1 .. y+1, clipped to the image.
static int check_band(const char* what, int rows, int hfull, int y0, int ys0, int srows) {
  SF_CHECK_ARG(hfull > 0 && rows > 0 && srows > 0 && y0 >= 0 && ys0 >= 0, "row band: non-positive dimension");
  SF_CHECK_ARG(y0 + rows <= hfull && ys0 + srows <= hfull, "row band: band exceeds the image");
  const int need0 = y0 > 0 ? y0 - 1 : 0, need1 = y0 + rows < hfull ? y0 + rows + 1 : hfull;
  if (ys0 > need0 || ys0 + srows < need1) {
    sf::set_error("%s: the source rows must include a 1-row halo on both sides of the band", what);
    return SF_ERR_ARG;
  }
  return SF_OK;
}

extern "C" int sf_disp_warp_band_f32(const float* img, const float* disp, float* warped, float* mask, float* dbg_grid,
                                     float* dbg_src, int B, int C, int rows, int W, int full_height, int y0,
                                     int src_y0, int src_rows, int border, int div_mode, int* neg_flag,
                                     void* stream) {
  using namespace sf;
  SF_CHECK_ARG(img && disp && warped, "disp_warp: null pointer");
  SF_CHECK_ARG(B > 0 && C > 0 && rows > 0 && W > 0, "disp_warp: non-positive dimension");
  if (int rc = check_band("disp_warp", rows, full_height, y0, src_y0, src_rows)) return rc;
  size_t n = static_cast<size_t>(B) * rows * W;
  int grid = static_cast<int>(min(static_cast<size_t>(sm_count()) * 8, (n + 255) / 256));
  disp_warp_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      img, disp, warped, mask, dbg_grid, dbg_src, B, C, rows, W, border, div_mode == SF_DIV_RECIP, neg_flag,
      RowBand{full_height, y0, src_y0, src_rows});
  SF_LAUNCH_CHECK();
  return SF_OK;
}

extern "C" int sf_disp_warp_f32(const float* img, const float* disp, float* warped, float* mask, float* dbg_grid,
                                float* dbg_src, int B, int C, int H, int W, int border, int div_mode, int* neg_flag,
                                void* stream) {
  return sf_disp_warp_band_f32(img, disp, warped, mask, dbg_grid, dbg_src, B, C, H, W, H, 0, 0, H, border, div_mode,
                               neg_flag, stream);
}

extern "C" int sf_guidance_band_f32(const float* cv, const float* disp, const float* left, const float* right,
                                    float* err, float* unc, int B, int D, int C, int rows, int W, int full_height,
                                    int y0, int src_y0, int src_rows, int div_mode, int* neg_flag, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(cv && disp && left && right && err && unc, "guidance: null pointer");
  SF_CHECK_ARG(B > 0 && D > 0 && C > 0 && rows > 0 && W > 0, "guidance: non-positive dimension");
  if (int rc = check_band("guidance", rows, full_height, y0, src_y0, src_rows)) return rc;
  size_t n = static_cast<size_t>(B) * rows * W;
  int grid = static_cast<int>(min(static_cast<size_t>(sm_count()) * 8, (n + 255) / 256));
  guidance_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      cv, disp, left, right, err, unc, B, D, C, rows, W, div_mode == SF_DIV_RECIP, neg_flag,
      RowBand{full_height, y0, src_y0, src_rows});
  SF_LAUNCH_CHECK();
  return SF_OK;
}

extern "C" int sf_guidance_f32(const float* cv, const float* disp, const float* left, const float* right, float* err,
                               float* unc, int B, int D, int C, int H, int W, int div_mode, int* neg_flag,
                               void* stream) {
  return sf_guidance_band_f32(cv, disp, left, right, err, unc, B, D, C, H, W, H, 0, 0, H, div_mode, neg_flag, stream);
}

extern "C" int sf_disp_warp_backward_f32(const float* img, const float* disp, const float* grad_warped, float* grad_img,
                                         float* grad_disp, int B, int C, int H, int W, int border, int div_mode,
                                         void* stream) {
  using namespace sf;
  SF_CHECK_ARG(img && disp && grad_warped, "disp_warp_backward: null pointer");
  SF_CHECK_ARG(B > 0 && C > 0 && H > 1 && W > 1, "disp_warp_backward: needs H, W > 1");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (grad_img) SF_CUDA(cudaMemsetAsync(grad_img, 0, sizeof(float) * static_cast<size_t>(B) * C * H * W, s));
  size_t n = static_cast<size_t>(B) * H * W;
  int grid = static_cast<int>(min(static_cast<size_t>(sm_count()) * 8, (n + 255) / 256));
  disp_warp_bwd_kernel<<<grid, 256, 0, s>>>(img, disp, grad_warped, grad_img, grad_disp, B, C, H, W, border,
                                            div_mode == SF_DIV_RECIP);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

extern "C" int sf_guidance_backward_f32(const float* cv, const float* disp, const float* right, const float* grad_err,
                                        const float* grad_unc, float* grad_cv, float* grad_disp, int B, int D, int C,
                                        int H, int W, int div_mode, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(cv && disp && right, "guidance_backward: null pointer");
  SF_CHECK_ARG(B > 0 && D > 0 && C > 0 && H > 1 && W > 1, "guidance_backward: needs H, W > 1");
  size_t n = static_cast<size_t>(B) * H * W;
  int grid = static_cast<int>(min(static_cast<size_t>(sm_count()) * 8, (n + 255) / 256));
  guidance_bwd_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(cv, disp, right, grad_err, grad_unc, grad_cv,
                                                                         grad_disp, B, D, C, H, W,
                                                                         div_mode == SF_DIV_RECIP);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

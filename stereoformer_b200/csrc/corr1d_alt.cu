// PytorchAlternateCorrBlock1D  ("alternative": on-the-fly correlation, no all-pairs volume).
//
// Reference: models/IterativeLocalCostVolume/corr/corr.py:8-51.  Per level i the right feature map is pooled along x
// (F.avg_pool2d [1,2], :47) and, for every left pixel and dx in [-r, r], sampled at (cx/2^i + dx, cy) with
// F.grid_sample(bilinear, zeros padding, align_corners=True) after the reference's own normalisation
// 2*x/(W-1) - 1, 2*y/(H-1) - 1 (:19-20); the sample is dotted with the left feature over the channels and divided by
// sqrt(C) (:31).  Output [B, L*(2r+1), H, W] float32, level-major like CorrBlock1D.
//
// One thread per (left pixel, level).  The 2r+1 sampling descriptors (clamped corner indices, four weights) are computed once
// and kept in registers together with the 2r+1 accumulators; the channel loop then issues coalesced loads (adjacent
// threads = adjacent pixels whose sampling positions are adjacent).  The row weights come from the same normalise /
// un-normalise round trip as the reference's (for integer cy the second row has weight 0 and is not loaded).
// Bytes: reads (1 + L*(~2r+2)) * C * 4 per pixel from L1/L2 (the maps are read ~40 times, but they are 2*C*H*W*4 in
// all, where the volume form writes and re-reads H*W*W*4*1.875): compute/latency-bound by design -- it is the
// reference's memory-saving variant.
#include "common.cuh"
#include "pixel_math.h"
#include "../../include/stereoformer_b200.h"

namespace sf {

constexpr int CA_MAXL = 4, CA_MAXT = 9;

__global__ void __launch_bounds__(256)
pool_w2_kernel(const float* __restrict__ in, float* __restrict__ out, size_t rows, int Win) {
  const int Wo = Win >> 1;
  const size_t n = rows * Wo;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t r = i / Wo;
    const int j = i - r * Wo;
    const float* q = in + r * Win + 2 * j;
    out[i] = (q[0] + q[1]) * 0.5f;                        // avg_pool2d([1,2], stride [1,2]): floor, last odd column dropped
  }
}

struct AltParams {
  const float* f1;               // [B,C,H,W1]
  const float* f2[CA_MAXL];      // level l: [B,C,H,W2>>l]
  int W2[CA_MAXL];
  const float* coords;           // [B,2,H,W1]  (x, y)
  float* out;                    // [B,L*(2r+1),H,W1]
  float* dbg_src;                // optional [B,L*(2r+1),H,W1,2]: un-normalised (ix, iy)
  int B, C, H, W1, num_levels, radius, recip;
};

__global__ void __launch_bounds__(128) corr1d_alt_kernel(const AltParams p) {
  const size_t HW = static_cast<size_t>(p.H) * p.W1, n = static_cast<size_t>(p.B) * HW;
  const size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  const size_t b = i / HW, r = i % HW;
  const int taps = 2 * p.radius + 1;
  const float cx = p.coords[b * 2 * HW + r], cy = p.coords[b * 2 * HW + HW + r];
  // row pair: the reference normalises y with H-1 and grid_sample(align_corners=True) maps it back
  const float iy = sfm::lookup_unnorm(sfm::lookup_norm(cy, p.H, p.recip), p.H);
  const float fy = floorf(iy);
  const bool yfin = fy >= -1.f && fy < (float)p.H;        // also rejects NaN / inf
  const int y0 = yfin ? (int)fy : 0;
  const float wy0 = (fy + 1.f) - iy, wy1 = iy - fy;
  const bool row0 = yfin && y0 >= 0 && wy0 != 0.f, row1 = yfin && y0 + 1 < p.H && wy1 != 0.f;
  const float scale = sqrtf((float)p.C);
  const float* f1p = p.f1 + b * p.C * HW + r;

  {
    const int l = blockIdx.y;                             // one level per thread: 4x the threads in flight
    const int Wl = p.W2[l];
    const size_t plane = static_cast<size_t>(p.H) * Wl;
    const float* base = p.f2[l] + b * p.C * plane;
    // per tap: two clamped column indices and four corner weights; a corner outside the map keeps a valid address and
    // gets weight 0 (ATen skips it: the same sum), so the channel loop is branch-free
    int xa[CA_MAXT], xb[CA_MAXT];
    float w00[CA_MAXT], w01[CA_MAXT], w10[CA_MAXT], w11[CA_MAXT], acc[CA_MAXT];
#pragma unroll
    for (int t = 0; t < CA_MAXT; ++t) {
      acc[t] = 0.f; xa[t] = xb[t] = 0; w00[t] = w01[t] = w10[t] = w11[t] = 0.f;
      if (t < taps) {
        const float x0 = sfm::lookup_x0(cx, l, (float)(t - p.radius));
        const float ix = sfm::lookup_unnorm(sfm::lookup_norm(x0, Wl, p.recip), Wl);
        const float fx = floorf(ix);
        if (fx >= -1.f && fx < (float)Wl) {
          const int x = (int)fx;
          const float wx0 = (fx + 1.f) - ix, wx1 = ix - fx;
          xa[t] = max(x, 0); xb[t] = min(x + 1, Wl - 1);
          if (x >= 0) { w00[t] = row0 ? wx0 * wy0 : 0.f; w10[t] = row1 ? wx0 * wy1 : 0.f; }
          if (x + 1 < Wl) { w01[t] = row0 ? wx1 * wy0 : 0.f; w11[t] = row1 ? wx1 * wy1 : 0.f; }
        }
        if (p.dbg_src) {
          const size_t o = ((b * p.num_levels * taps + static_cast<size_t>(l * taps + t)) * HW + r) * 2;
          p.dbg_src[o] = ix; p.dbg_src[o + 1] = iy;
        }
      }
    }
    const int ya = row0 ? y0 : (row1 ? y0 + 1 : 0), yb = row1 ? y0 + 1 : ya;   // valid rows even when a weight is 0
    for (int c = 0; c < p.C; ++c) {
      const float a = __ldg(f1p + c * HW);
      const float* q0 = base + c * plane + static_cast<ptrdiff_t>(ya) * Wl;
      const float* q1 = base + c * plane + static_cast<ptrdiff_t>(yb) * Wl;
      if (row1) {
#pragma unroll
        for (int t = 0; t < CA_MAXT; ++t) {
          if (t >= taps) continue;
          // ATen grid_sampler_2d order: nw, ne, sw, se
          float v = __ldg(q0 + xa[t]) * w00[t];
          v += __ldg(q0 + xb[t]) * w01[t];
          v += __ldg(q1 + xa[t]) * w10[t];
          v += __ldg(q1 + xb[t]) * w11[t];
          acc[t] = fmaf(v, a, acc[t]);
        }
      } else {
#pragma unroll
        for (int t = 0; t < CA_MAXT; ++t) {
          if (t >= taps) continue;
          float v = __ldg(q0 + xa[t]) * w00[t];
          v += __ldg(q0 + xb[t]) * w01[t];
          acc[t] = fmaf(v, a, acc[t]);
        }
      }
    }
    float* o = p.out + (b * p.num_levels * taps + static_cast<size_t>(l * taps)) * HW + r;
#pragma unroll
    for (int t = 0; t < CA_MAXT; ++t)
      if (t < taps) o[t * HW] = acc[t] / scale;
  }
}

}  // namespace sf

extern "C" size_t sf_corr1d_alt_workspace_bytes(int B, int C, int H, int W2, int num_levels) {
  size_t n = 0;
  for (int l = 1; l < num_levels; ++l) n += static_cast<size_t>(B) * C * H * (W2 >> l) * sizeof(float);
  return n;
}

extern "C" int sf_corr1d_alt_f32(const float* f1, const float* f2, const float* coords, float* out, float* dbg_src,
                                 void* workspace, size_t workspace_bytes, int B, int C, int H, int W1, int W2,
                                 int num_levels, int radius, int div_mode, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(f1 && f2 && coords && out, "corr1d_alt: null pointer");
  SF_CHECK_ARG(B > 0 && C > 0 && H > 1 && W1 > 0 && W2 > 1, "corr1d_alt: needs H > 1 and W2 > 1 (the reference divides by H-1, W-1)");
  SF_CHECK_ARG(num_levels >= 1 && num_levels <= CA_MAXL && radius >= 0 && 2 * radius + 1 <= CA_MAXT,
               "corr1d_alt: up to %d levels and radius %d supported", CA_MAXL, (CA_MAXT - 1) / 2);
  SF_CHECK_ARG((W2 >> (num_levels - 1)) > 1, "corr1d_alt: the last level is narrower than 2 columns");
  const size_t need = sf_corr1d_alt_workspace_bytes(B, C, H, W2, num_levels);
  SF_CHECK_ARG(need == 0 || (workspace && workspace_bytes >= need), "corr1d_alt: workspace of %zu bytes needed, got %zu", need,
               workspace_bytes);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  AltParams p = {};
  p.f1 = f1; p.coords = coords; p.out = out; p.dbg_src = dbg_src;
  p.B = B; p.C = C; p.H = H; p.W1 = W1; p.num_levels = num_levels; p.radius = radius;
  p.recip = div_mode == SF_DIV_RECIP;
  p.f2[0] = f2; p.W2[0] = W2;
  float* ws = static_cast<float*>(workspace);
  const size_t rows = static_cast<size_t>(B) * C * H;
  for (int l = 1; l < num_levels; ++l) {
    const int Win = W2 >> (l - 1);
    const size_t n = rows * (Win >> 1);
    const int grid = static_cast<int>(min(static_cast<size_t>(sm_count()) * 8, (n + 255) / 256));
    pool_w2_kernel<<<grid, 256, 0, s>>>(p.f2[l - 1], ws, rows, Win);
    SF_LAUNCH_CHECK();
    p.f2[l] = ws; p.W2[l] = Win >> 1;
    ws += n;
  }
  const size_t n = static_cast<size_t>(B) * H * W1;
  corr1d_alt_kernel<<<dim3(static_cast<unsigned>((n + 127) / 128), num_levels), 128, 0, s>>>(p);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

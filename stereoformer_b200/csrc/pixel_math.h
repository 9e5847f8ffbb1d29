// Per-pixel arithmetic of the window sampler, soft-argmin and disparity warp, written once for the
// CUDA kernels (window.cu, warp.cu).  Each operation that the reference performs as a separate
// fp32 tensor op is a separately rounded op here (__f*_rn blocks nvcc's FMA contraction), in the
// reference's order, so that candidates, indices and sampling coordinates come out bit-identical.
// The header also compiles as plain C++ (-ffp-contract=off) for tests/host_emu, a no-GPU unit-test
// harness of this arithmetic; the product never runs it on the host.
#pragma once
#include <math.h>

#ifdef __CUDACC__
#define SF_HD __host__ __device__ __forceinline__
#else
#define SF_HD inline
#endif

namespace sfm {

#if defined(__CUDA_ARCH__)
SF_HD float fadd(float a, float b) { return __fadd_rn(a, b); }
SF_HD float fsub(float a, float b) { return __fsub_rn(a, b); }
SF_HD float fmul(float a, float b) { return __fmul_rn(a, b); }
SF_HD float fdiv(float a, float b) { return __fdiv_rn(a, b); }
SF_HD float fsqrt(float a) { return __fsqrt_rn(a); }
#else
SF_HD float fadd(float a, float b) { volatile float r = a + b; return r; }
SF_HD float fsub(float a, float b) { volatile float r = a - b; return r; }
SF_HD float fmul(float a, float b) { volatile float r = a * b; return r; }
SF_HD float fdiv(float a, float b) { volatile float r = a / b; return r; }
SF_HD float fsqrt(float a) { return sqrtf(a); }
#endif

// tensor / python-scalar: true division (PyTorch CPU) or multiply by the fp32 reciprocal (PyTorch CUDA).
SF_HD float sdiv(float a, float s, int recip) { return recip ? fmul(a, fdiv(1.0f, s)) : fdiv(a, s); }

SF_HD float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }

struct Window {
  float lower, upper, inv;  // inv = invalid mask (0/1), only meaningful with consider_valid
  float step;
};

// Bounds -> validity / clamping -> step.   fixed_local_cost_volume.py:23-34, variance_local_cost_volume.py:44-56,
// Attempts/update.py:284-295.  clamp_variance selects the [0,D-1] clamp of the variance module.
SF_HD Window make_window(float lower, float upper, int D, int n_points, int consider_valid, int clamp_variance,
                         float extra_invalid, int recip) {
  Window w;
  w.inv = 0.f;
  if (consider_valid) {
    float m = fadd(lower < 0.f ? 1.f : 0.f, upper >= (float)(D - 1) ? 1.f : 0.f);
    m = fadd(m, extra_invalid);
    w.inv = fminf(m, 1.0f);
  } else if (clamp_variance) {
    lower = clampf(lower, 0.f, (float)(D - 1));
    upper = clampf(upper, 0.f, (float)(D - 1));
  } else {
    lower = fmaxf(lower, 0.f);
    upper = clampf(upper, 0.f, (float)D);
  }
  w.lower = lower;
  w.upper = upper;
  w.step = sdiv(fsub(upper, lower), (float)n_points, recip);
  return w;
}

// k-th candidate.  fixed_local_cost_volume.py:38-45.
SF_HD float window_candidate(const Window& w, int k, float cur_disp, int consider_valid) {
  float c = fadd(w.lower, fmul((float)k, w.step));
  if (consider_valid) c = fadd(fmul(c, fsub(1.0f, w.inv)), fmul(w.inv, cur_disp));
  return c;
}

// Straight-through ceil/floor, clamp, integer cast.  fixed_local_cost_volume.py:5-9,65-69.
SF_HD void window_index(float cand, int D, int* idx_floor, int* idx_ceil) {
  float c = fadd(fsub(ceilf(cand), cand), cand);
  float f = fadd(fsub(floorf(cand), cand), cand);
  *idx_ceil = (int)clampf(c, 0.f, (float)(D - 1));
  *idx_floor = (int)clampf(f, 0.f, (float)(D - 1));
}

// Linear interpolation between the two gathered entries.  fixed_local_cost_volume.py:72-78.
SF_HD float window_lerp(float v_floor, float v_ceil, int idx_ceil, float cand) {
  float fr = fsub((float)idx_ceil, cand);
  float cr = fsub(1.0f, fr);
  return fadd(fmul(v_floor, fr), fmul(v_ceil, cr));
}

// Normalised warp grid.  utils/disparity_warper.py:55-56,96-98.
SF_HD float warp_norm(float pix, int size, int recip) {
  return fsub(fmul(2.0f, sdiv(pix, (float)(size - 1), recip)), 1.0f);
}
// ATen grid_sampler un-normalisation, align_corners=False: ((g + 1) * size - 1) / 2.
SF_HD float warp_unnorm(float g, int size) { return fmul(fsub(fmul(fadd(g, 1.0f), (float)size), 1.0f), 0.5f); }

// CorrBlock1D lookup coordinate.  corr/corr.py:88-90, utils/utils.py:62-64 + ATen align_corners=True
// un-normalisation ((g + 1) / 2) * (W - 1).
SF_HD float lookup_x0(float cx, int level, float dx) { return fadd(dx, fmul(cx, 1.0f / (float)(1 << level))); }
SF_HD float lookup_norm(float x0, int Wl, int recip) { return fsub(sdiv(fmul(2.0f, x0), (float)(Wl - 1), recip), 1.0f); }
SF_HD float lookup_unnorm(float g, int Wl) { return fmul(fmul(fadd(g, 1.0f), 0.5f), (float)(Wl - 1)); }

}  // namespace sfm

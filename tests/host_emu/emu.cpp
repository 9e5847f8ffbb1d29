// TEST HARNESS ONLY (never part of the product): compiles stereoformer_b200/csrc/pixel_math.h -- the exact
// per-pixel arithmetic the CUDA kernels use -- as plain C++ (-ffp-contract=off) so that the bit-exactness of
// candidates / indices / coordinates can be unit-tested without a GPU.
#include "../../stereoformer_b200/csrc/pixel_math.h"

extern "C" {

// mode: 0 fixed radius, 2 offsets (lo/up), 3 absolute.  Outputs [n][K] row-major.
void emu_window(const float* cur, const float* lo, const float* up, int n, int D, int n_points, int mode, float radius,
                int consider_valid, int recip, float* cand, int* idx_floor, int* idx_ceil) {
  const int K = n_points + 1;
  for (int i = 0; i < n; ++i) {
    float lower, upper;
    if (mode == 0) { lower = sfm::fsub(cur[i], radius); upper = sfm::fadd(cur[i], radius); }
    else if (mode == 2) { lower = sfm::fsub(cur[i], lo[i]); upper = sfm::fadd(cur[i], up[i]); }
    else { lower = lo[i]; upper = up[i]; }
    sfm::Window w = sfm::make_window(lower, upper, D, n_points, consider_valid, 0, 0.f, recip);
    for (int k = 0; k < K; ++k) {
      float c = sfm::window_candidate(w, k, cur[i], consider_valid);
      cand[i * K + k] = c;
      sfm::window_index(c, D, &idx_floor[i * K + k], &idx_ceil[i * K + k]);
    }
  }
}

void emu_lerp(const float* vf, const float* vc, const int* ic, const float* cand, int n, float* out) {
  for (int i = 0; i < n; ++i) out[i] = sfm::window_lerp(vf[i], vc[i], ic[i], cand[i]);
}

void emu_index(const float* cand, int n, int D, int* idx_floor, int* idx_ceil) {
  for (int i = 0; i < n; ++i) sfm::window_index(cand[i], D, &idx_floor[i], &idx_ceil[i]);
}

// disp [H*W] of one image -> grid [H*W][2], src [H*W][2] (border clip)
void emu_warp(const float* disp, int H, int W, int recip, float* grid, float* src) {
  for (int y = 0; y < H; ++y)
    for (int x = 0; x < W; ++x) {
      int i = y * W + x;
      float gx = sfm::warp_norm(sfm::fadd((float)x, -disp[i]), W, recip);
      float gy = sfm::warp_norm(sfm::fadd((float)y, 0.f), H, recip);
      grid[2 * i] = gx; grid[2 * i + 1] = gy;
      src[2 * i] = fminf((float)(W - 1), fmaxf(sfm::warp_unnorm(gx, W), 0.f));
      src[2 * i + 1] = fminf((float)(H - 1), fmaxf(sfm::warp_unnorm(gy, H), 0.f));
    }
}

void emu_lookup(const float* cx, int n, int level, int radius, int Wl, int recip, float* x0, float* xg, float* ix) {
  const int taps = 2 * radius + 1;
  for (int i = 0; i < n; ++i)
    for (int t = 0; t < taps; ++t) {
      float a = sfm::lookup_x0(cx[i], level, (float)(t - radius));
      float g = sfm::lookup_norm(a, Wl, recip);
      x0[i * taps + t] = a; xg[i * taps + t] = g; ix[i * taps + t] = sfm::lookup_unnorm(g, Wl);
    }
}
}

// Backward passes of the pure-PyTorch operators of the hot path  (SURVEY 8f rank 3): the gradients PyTorch's
// autograd derives for the reference modules, as one kernel per operator.
//
//   a1/a2  CostVolume 'correlation' / build_gwc_volume  models/utils/build_cost_volume.py:39-47, GWC_Cost_Volume.py:14-24
//            gL[c,x] = 1/Cg * sum_d G[d,x]   * R[c,x-d]        gR[c,x] = 1/Cg * sum_d G[d,x+d] * L[c,x+d]
//   a3     DisparityEstimation  models/utils/estimation.py:14-62:  gcv[d] = +-g * p_d * (d - disp)
//   a4-a6  LocalCostVolume / DisparityUpdateDLC window tail  Attempts/fixed_local_cost_volume.py:18-80,
//          Attempts/update.py:275-313.  The straight-through ceil/floor of :5-9 end in .long(), so no gradient
//          flows through the indices; it flows through floor_rate = ceil_idx - cand (d/dcand = -1), the final
//          score * candidate product, torch.clamp (pass-through inside [min, max]) and the bound arithmetic.
//          The invalid mask is a comparison result (.float()), i.e. a constant.
//   build_cost_volume_from_volume alone (fixed_local_cost_volume.py:61-80): gcv scatter + gcand.
#include "common.cuh"
#include "pixel_math.h"
#include "../../include/stereoformer_b200.h"

namespace sf {

// ------------------------------------------------------------------------------------------------ a1 / a2
// One CTA per (row y, batch b, group g); the row of the volume gradient G[D][W] is staged in shared memory when it
// fits, threads sweep (channel, x).  HBM-bound: 4*Cg*W + D*W floats per CTA.
__global__ void __launch_bounds__(256)
corr_volume_bwd_kernel(const float* __restrict__ L, const float* __restrict__ R, const float* __restrict__ G,
                       float* __restrict__ gL, float* __restrict__ gR, int C, int H, int W, int D, int groups,
                       int use_smem) {
  extern __shared__ float gS[];                                       // [D][W] when use_smem
  const int y = blockIdx.x, b = blockIdx.y, g = blockIdx.z;
  const int Cg = C / groups;
  const size_t HW = static_cast<size_t>(H) * W;
  const float* Grow = G + ((static_cast<size_t>(b) * groups + g) * D * H + y) * W;   // + d*HW
  if (use_smem) {
    for (int e = threadIdx.x; e < D * W; e += blockDim.x) gS[e] = __ldg(Grow + static_cast<size_t>(e / W) * HW + e % W);
    __syncthreads();
  }
  const float inv = 1.0f / (float)Cg;
  for (int e = threadIdx.x; e < Cg * W; e += blockDim.x) {
    const int x = e % W, c = g * Cg + e / W;
    const size_t row = ((static_cast<size_t>(b) * C + c) * H + y) * W;
    const float* Lr = L + row;
    const float* Rr = R + row;
    float al = 0.f, ar = 0.f;
    const int dl = min(D - 1, x), dr = min(D - 1, W - 1 - x);
    if (use_smem) {
      for (int d = 0; d <= dl; ++d) al = fmaf(gS[d * W + x], __ldg(Rr + x - d), al);
      for (int d = 0; d <= dr; ++d) ar = fmaf(gS[d * W + x + d], __ldg(Lr + x + d), ar);
    } else {
      for (int d = 0; d <= dl; ++d) al = fmaf(__ldg(Grow + d * HW + x), __ldg(Rr + x - d), al);
      for (int d = 0; d <= dr; ++d) ar = fmaf(__ldg(Grow + d * HW + x + d), __ldg(Lr + x + d), ar);
    }
    if (gL) gL[row + x] = al * inv;
    if (gR) gR[row + x] = ar * inv;
  }
}

// Register-tiled form (W % 4 == 0, Cg % 4 == 0): a thread owns 4 consecutive x of 4 channels.  gS[d][x] = G[d][x] and
// the skewed copy gK[d][x] = G[d][x+d] are read as aligned float4 and reused by the 4 channels; the R / L rows of a
// 32-channel chunk are staged in shared memory 4-way interleaved (Rt[c][x % 4][x / 4]) so that the sliding 4-wide
// register windows take their ONE new element per disparity with a conflict-free scalar load (consecutive lanes own
// consecutive x / 4).  L1 wavefronts per FMA drop 4x against the one-output-per-thread form.
constexpr int CB_CH = 32;
__global__ void __launch_bounds__(256)
corr_volume_bwd4_kernel(const float* __restrict__ L, const float* __restrict__ R, const float* __restrict__ G,
                        float* __restrict__ gL, float* __restrict__ gR, int C, int H, int W, int D, int groups) {
  extern __shared__ __align__(16) float gS[];                         // [D][W], gK [D][W], Rt [CH][4][W4], Lt [CH][4][W4]
  const int W4 = W >> 2;
  float* gK = gS + D * W;
  float* Rt = gK + D * W;
  float* Lt = Rt + CB_CH * W;
  const int y = blockIdx.x, b = blockIdx.y, g = blockIdx.z;
  const int Cg = C / groups;
  const size_t HW = static_cast<size_t>(H) * W;
  const float* Grow = G + ((static_cast<size_t>(b) * groups + g) * D * H + y) * W;
  for (int d = threadIdx.x >> 5; d < D; d += 8)
    for (int x = threadIdx.x & 31; x < W; x += 32) {
      gS[d * W + x] = __ldg(Grow + static_cast<size_t>(d) * HW + x);
      gK[d * W + x] = (x + d < W) ? __ldg(Grow + static_cast<size_t>(d) * HW + x + d) : 0.f;
    }
  const float inv = 1.0f / (float)Cg;
  for (int cc = 0; cc < Cg; cc += CB_CH) {
    const int nch = min(CB_CH, Cg - cc);
    __syncthreads();
    for (int c = threadIdx.x >> 5; c < nch; c += 8) {                 // one warp per channel row: no divisions
      const size_t src = ((static_cast<size_t>(b) * C + g * Cg + cc + c) * H + y) * W;
      for (int x = threadIdx.x & 31; x < W; x += 32) {
        Rt[(c * 4 + (x & 3)) * W4 + (x >> 2)] = __ldg(R + src + x);
        Lt[(c * 4 + (x & 3)) * W4 + (x >> 2)] = __ldg(L + src + x);
      }
    }
    __syncthreads();
    // items: (channel quad cq, x quad xq); a warp takes one channel quad and sweeps xq in steps of 32
    for (int cq = threadIdx.x >> 5; cq < (nch >> 2); cq += 8)
    for (int xq = threadIdx.x & 31; xq < W4; xq += 32) {
      const int c0 = cq * 4, x0 = xq * 4;
      float rw[4][4], lw[4][4], al[4][4], ar[4][4];
#pragma unroll
      for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          rw[c][i] = Rt[((c0 + c) * 4 + i) * W4 + xq];
          lw[c][i] = Lt[((c0 + c) * 4 + i) * W4 + xq];
          al[c][i] = 0.f; ar[c][i] = 0.f;
        }
#pragma unroll 4
      for (int d = 0; d < D; ++d) {
        const float4 ga = *reinterpret_cast<const float4*>(gS + d * W + x0);
        const float4 gb = *reinterpret_cast<const float4*>(gK + d * W + x0);
        const int xr = x0 - d - 1, xl = x0 + d + 4;
        const int ro = (xr & 3) * W4 + (xr >> 2), lo = (xl & 3) * W4 + (xl >> 2);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          al[c][0] = fmaf(ga.x, rw[c][0], al[c][0]); al[c][1] = fmaf(ga.y, rw[c][1], al[c][1]);
          al[c][2] = fmaf(ga.z, rw[c][2], al[c][2]); al[c][3] = fmaf(ga.w, rw[c][3], al[c][3]);
          ar[c][0] = fmaf(gb.x, lw[c][0], ar[c][0]); ar[c][1] = fmaf(gb.y, lw[c][1], ar[c][1]);
          ar[c][2] = fmaf(gb.z, lw[c][2], ar[c][2]); ar[c][3] = fmaf(gb.w, lw[c][3], ar[c][3]);
          rw[c][3] = rw[c][2]; rw[c][2] = rw[c][1]; rw[c][1] = rw[c][0];
          rw[c][0] = xr >= 0 ? Rt[(c0 + c) * W + ro] : 0.f;
          lw[c][0] = lw[c][1]; lw[c][1] = lw[c][2]; lw[c][2] = lw[c][3];
          lw[c][3] = xl < W ? Lt[(c0 + c) * W + lo] : 0.f;
        }
      }
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const size_t row = ((static_cast<size_t>(b) * C + g * Cg + cc + c0 + c) * H + y) * W + x0;
        if (gL) *reinterpret_cast<float4*>(gL + row) = make_float4(al[c][0] * inv, al[c][1] * inv, al[c][2] * inv, al[c][3] * inv);
        if (gR) *reinterpret_cast<float4*>(gR + row) = make_float4(ar[c][0] * inv, ar[c][1] * inv, ar[c][2] * inv, ar[c][3] * inv);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------ a3
__global__ void __launch_bounds__(256)
soft_argmin_bwd_kernel(const float* __restrict__ cv, const float* __restrict__ gdisp, float* __restrict__ gcv, int B,
                       int D, int H, int W, float sign) {
  const size_t HW = static_cast<size_t>(H) * W, n = static_cast<size_t>(B) * HW;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t b = i / HW, r = i % HW;
    const float* src = cv + b * D * HW + r;
    float m = -INFINITY;
    for (int d = 0; d < D; ++d) m = fmaxf(m, sign * __ldg(src + d * HW));
    float s = 0.f, e = 0.f;
    for (int d = 0; d < D; ++d) {
      const float ex = __expf(sign * __ldg(src + d * HW) - m);
      s += ex; e = fmaf(ex, (float)d, e);
    }
    const float inv = 1.0f / s, disp = e * inv, g = gdisp[i] * sign;
    float* dst = gcv + b * D * HW + r;
    for (int d = 0; d < D; ++d) dst[d * HW] = g * __expf(sign * __ldg(src + d * HW) - m) * inv * ((float)d - disp);
  }
}

// ------------------------------------------------------------------------------------------------ a4 - a6
struct WinBwdParams {
  const float* cv; const float* disp; const float* lo; const float* up; const float* gout;
  float* gcv; float* gdisp; float* glo; float* gup;
  int B, D, H, W, mode, n_points, consider_valid, recip;
  float radius;   // SF_WIN_FIXED: the radius; SF_WIN_VARIANCE: gamma
};

constexpr int WB_TPB = 128;

// one thread per pixel; its column of gcv is accumulated in shared memory ([D][128], conflict-free) and written
// with coalesced row stores
__global__ void __launch_bounds__(WB_TPB)
local_window_bwd_kernel(const WinBwdParams p) {
  extern __shared__ float col[];                                      // [D][WB_TPB]
  const int D = p.D, N = p.n_points;
  const size_t HW = static_cast<size_t>(p.H) * p.W, n = static_cast<size_t>(p.B) * HW;
  const size_t i = blockIdx.x * static_cast<size_t>(WB_TPB) + threadIdx.x;
  const bool live = i < n;
  for (int d = 0; d < D; ++d) col[d * WB_TPB + threadIdx.x] = 0.f;
  if (live) {
    const size_t b = i / HW, r = i % HW;
    const float* src = p.cv + b * D * HW + r;
    const float cur = p.disp[i], g = p.gout[i];
    float lower_raw, upper_raw, pre0 = 0.f, pre1 = 0.f;
    float vm = 0.f, vs = 1.f, var = 0.f, sigma = 0.f, extra = 0.f;   // variance mode: soft-max statistics of the column
    if (p.mode == SF_WIN_FIXED) {
      lower_raw = sfm::fsub(cur, p.radius); upper_raw = sfm::fadd(cur, p.radius);
    } else if (p.mode == SF_WIN_VARIANCE) {
      // GetVarince + bounds, variance_local_cost_volume.py:12-25, 38-42 (the forward's arithmetic, window.cu)
      vm = -INFINITY;
      for (int d = 0; d < D; ++d) vm = fmaxf(vm, __ldg(src + d * HW));
      vs = 0.f;
      float acc = 0.f;
      for (int d = 0; d < D; ++d) {
        const float ex = __expf(__ldg(src + d * HW) - vm), e = sfm::fsub((float)d, cur);
        vs += ex;
        acc = fmaf(ex, sfm::fmul(e, e), acc);
      }
      var = acc / vs;
      sigma = sfm::fsqrt(var);
      const float gs = sfm::fmul(p.radius, sigma);
      lower_raw = sfm::fsub(cur, gs); upper_raw = sfm::fadd(cur, gs);
      if (p.consider_valid) extra = (upper_raw > (float)(r % p.W)) ? 1.f : 0.f;
    } else if (p.mode == SF_WIN_OFFSETS) {
      lower_raw = sfm::fsub(cur, p.lo[i]); upper_raw = sfm::fadd(cur, p.up[i]);
    } else if (p.mode == SF_WIN_HEAD) {
      pre0 = p.lo[i + b * HW]; pre1 = p.lo[i + b * HW + HW];
      lower_raw = sfm::fsub(cur, fmaxf(pre0, 0.f)); upper_raw = sfm::fadd(cur, fmaxf(pre1, 0.f));
    } else {
      lower_raw = p.lo[i]; upper_raw = p.up[i];
    }
    const sfm::Window w = sfm::make_window(lower_raw, upper_raw, D, N, p.consider_valid, p.mode == SF_WIN_VARIANCE,
                                           extra, p.recip);
    // pass 1: soft-max statistics and the expectation
    float m = -INFINITY;
    for (int k = 0; k <= N; ++k) {
      const float c = sfm::window_candidate(w, k, cur, p.consider_valid);
      int jf, jc;
      sfm::window_index(c, D, &jf, &jc);
      m = fmaxf(m, sfm::window_lerp(__ldg(src + jf * HW), __ldg(src + jc * HW), jc, c));
    }
    float s = 0.f, e = 0.f;
    for (int k = 0; k <= N; ++k) {
      const float c = sfm::window_candidate(w, k, cur, p.consider_valid);
      int jf, jc;
      sfm::window_index(c, D, &jf, &jc);
      const float ex = __expf(sfm::window_lerp(__ldg(src + jf * HW), __ldg(src + jc * HW), jc, c) - m);
      s += ex; e = fmaf(ex, c, e);
    }
    const float inv = 1.0f / s, out = e * inv;
    // pass 2: gradients
    float g_lower = 0.f, g_upper = 0.f, g_cur = 0.f;
    for (int k = 0; k <= N; ++k) {
      const float c = sfm::window_candidate(w, k, cur, p.consider_valid);
      int jf, jc;
      sfm::window_index(c, D, &jf, &jc);
      const float vf = __ldg(src + jf * HW), vc = __ldg(src + jc * HW);
      const float sk = __expf(sfm::window_lerp(vf, vc, jc, c) - m) * inv;
      const float g_local = g * sk * (c - out);
      const float fr = (float)jc - c, cr = 1.0f - fr;
      col[jf * WB_TPB + threadIdx.x] += g_local * fr;
      col[jc * WB_TPB + threadIdx.x] += g_local * cr;
      const float g_cand = g * sk + g_local * (vc - vf);
      float g_raw = g_cand;
      if (p.consider_valid) { g_raw = g_cand * (1.0f - w.inv); g_cur += g_cand * w.inv; }
      const float t = (float)k / (float)N;
      g_lower += g_raw * (1.0f - t);
      g_upper += g_raw * t;
    }
    if (!p.consider_valid) {                                          // torch.clamp backward: pass-through inside [min, max]
      if (p.mode == SF_WIN_VARIANCE) {                                // both bounds clamped to [0, D-1] (:52-53)
        if (!(lower_raw >= 0.f && lower_raw <= (float)(D - 1))) g_lower = 0.f;
        if (!(upper_raw >= 0.f && upper_raw <= (float)(D - 1))) g_upper = 0.f;
      } else {
        if (!(lower_raw >= 0.f)) g_lower = 0.f;
        if (!(upper_raw >= 0.f && upper_raw <= (float)D)) g_upper = 0.f;
      }
    }
    if (p.mode == SF_WIN_FIXED) {
      g_cur += g_lower + g_upper;
    } else if (p.mode == SF_WIN_VARIANCE) {
      // bounds = cur -+ gamma * sigma, sigma = sqrt(var), var = sum_d p_d (d - cur)^2, p = softmax(cv):
      //   g_var = gamma (g_upper - g_lower) / (2 sigma);  g_cv[d] += p_d g_var ((d-cur)^2 - var);
      //   g_cur += g_lower + g_upper - 2 g_var sum_d p_d (d - cur)
      g_cur += g_lower + g_upper;
      const float g_var = p.radius * (g_upper - g_lower) * 0.5f / sigma;
      float mean_e = 0.f;
      const float vinv = 1.0f / vs;
      for (int d = 0; d < D; ++d) {
        const float pd = __expf(__ldg(src + d * HW) - vm) * vinv, e = (float)d - cur;
        col[d * WB_TPB + threadIdx.x] += pd * g_var * (e * e - var);
        mean_e = fmaf(pd, e, mean_e);
      }
      g_cur -= 2.0f * g_var * mean_e;
    } else if (p.mode == SF_WIN_OFFSETS) {
      g_cur += g_lower + g_upper;
      if (p.glo) p.glo[i] = -g_lower;
      if (p.gup) p.gup[i] = g_upper;
    } else if (p.mode == SF_WIN_HEAD) {
      g_cur += g_lower + g_upper;
      if (p.glo) {
        p.glo[i + b * HW] = pre0 > 0.f ? -g_lower : 0.f;               // relu2 backward
        p.glo[i + b * HW + HW] = pre1 > 0.f ? g_upper : 0.f;
      }
    } else {
      if (p.glo) p.glo[i] = g_lower;
      if (p.gup) p.gup[i] = g_upper;
    }
    if (p.gdisp) p.gdisp[i] = g_cur;
  }
  if (p.gcv && live) {
    const size_t b = i / HW, r = i % HW;
    float* dst = p.gcv + b * D * HW + r;
    for (int d = 0; d < D; ++d) dst[d * HW] = col[d * WB_TPB + threadIdx.x];
  }
}

// build_cost_volume_from_volume backward: K candidates per pixel
__global__ void __launch_bounds__(WB_TPB)
sample_volume_bwd_kernel(const float* __restrict__ cv, const float* __restrict__ cand, const float* __restrict__ glocal,
                         float* __restrict__ gcv, float* __restrict__ gcand, int B, int D, int K, int H, int W) {
  extern __shared__ float col[];
  const size_t HW = static_cast<size_t>(H) * W, n = static_cast<size_t>(B) * HW;
  const size_t i = blockIdx.x * static_cast<size_t>(WB_TPB) + threadIdx.x;
  if (i >= n) return;
  for (int d = 0; d < D; ++d) col[d * WB_TPB + threadIdx.x] = 0.f;
  const size_t b = i / HW, r = i % HW;
  const float* src = cv + b * D * HW + r;
  for (int k = 0; k < K; ++k) {
    const size_t o = (b * K + k) * HW + r;
    const float c = cand[o], g = glocal[o];
    int jf, jc;
    sfm::window_index(c, D, &jf, &jc);
    const float fr = (float)jc - c;
    col[jf * WB_TPB + threadIdx.x] += g * fr;
    col[jc * WB_TPB + threadIdx.x] += g * (1.0f - fr);
    if (gcand) gcand[o] = g * (__ldg(src + jc * HW) - __ldg(src + jf * HW));
  }
  if (gcv) {
    float* dst = gcv + b * D * HW + r;
    for (int d = 0; d < D; ++d) dst[d * HW] = col[d * WB_TPB + threadIdx.x];
  }
}

}  // namespace sf

namespace sf {
bool corr_bwd_tc_supported(const float* L, const float* R, const float* G, const float* gL, const float* gR, int nBG,
                           int Cg, int H, int W, int D);
int launch_corr_bwd_tc(const float* L, const float* R, const float* G, float* gL, float* gR, int nBG, int Cg, int H,
                       int W, int D, cudaStream_t stream);
}  // namespace sf

static int corr_volume_backward_impl(const float* left, const float* right, const float* grad_cv, float* grad_left,
                                     float* grad_right, int B, int C, int H, int W, int D, int groups, int algo,
                                     void* stream);

extern "C" int sf_corr_volume_backward_f32(const float* left, const float* right, const float* grad_cv, float* grad_left,
                                           float* grad_right, int B, int C, int H, int W, int D, int groups,
                                           void* stream) {
  return corr_volume_backward_impl(left, right, grad_cv, grad_left, grad_right, B, C, H, W, D, groups, SF_ALGO_GENERIC,
                                   stream);
}

extern "C" int sf_corr_volume_backward_ex_f32(const float* left, const float* right, const float* grad_cv,
                                              float* grad_left, float* grad_right, int B, int C, int H, int W, int D,
                                              int groups, int algo, void* stream) {
  return corr_volume_backward_impl(left, right, grad_cv, grad_left, grad_right, B, C, H, W, D, groups, algo, stream);
}

static int corr_volume_backward_impl(const float* left, const float* right, const float* grad_cv, float* grad_left,
                                     float* grad_right, int B, int C, int H, int W, int D, int groups, int algo,
                                     void* stream) {
  using namespace sf;
  SF_CHECK_ARG(left && right && grad_cv, "corr_volume_backward: null pointer");
  SF_CHECK_ARG(B > 0 && C > 0 && H > 0 && W > 0 && D > 0, "corr_volume_backward: non-positive dimension");
  SF_CHECK_ARG(groups > 0 && C % groups == 0, "corr_volume_backward: channels %d not divisible by groups %d", C, groups);
  SF_CHECK_ARG(B <= 65535 && groups <= 65535, "corr_volume_backward: batch or group count too large");
  if (!grad_left && !grad_right) return SF_OK;
  if (algo == SF_ALGO_EXACT) algo = SF_ALGO_GENERIC;
  SF_CHECK_ARG(algo == SF_ALGO_AUTO || algo == SF_ALGO_GENERIC || algo == SF_ALGO_TC,
               "corr_volume_backward: unknown algo %d", algo);
  if (algo != SF_ALGO_GENERIC) {
    // tensor-core path from 64 channels per group (two TMEM lane quarters); below that the FFMA kernel is HBM-bound too
    const bool fits = corr_bwd_tc_supported(left, right, grad_cv, grad_left, grad_right, B * groups, C / groups, H, W, D);
    SF_CHECK_ARG(fits || algo != SF_ALGO_TC, "corr_volume_backward: shape does not fit the tcgen05 kernel "
                 "(channels per group %% 32, W %% 4, D <= 65, 16-byte aligned pointers)");
    if (fits && (algo == SF_ALGO_TC || C / groups >= 64))
      return launch_corr_bwd_tc(left, right, grad_cv, grad_left, grad_right, B * groups, C / groups, H, W, D,
                                static_cast<cudaStream_t>(stream));
  }
  const size_t smem = static_cast<size_t>(D) * W * sizeof(float);
  const int use_smem = smem <= 160 * 1024;
  dim3 grid(H, B, groups);
  const bool al16 = ((reinterpret_cast<uintptr_t>(left) | reinterpret_cast<uintptr_t>(right) |
                      reinterpret_cast<uintptr_t>(grad_left) | reinterpret_cast<uintptr_t>(grad_right)) & 15) == 0;
  const size_t smem4 = 2 * smem + 2 * static_cast<size_t>(CB_CH) * W * sizeof(float);
  if (W % 4 == 0 && (C / groups) % 4 == 0 && al16 && smem4 <= 100 * 1024) {
    SF_CUDA(cudaFuncSetAttribute(corr_volume_bwd4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem4)));
    corr_volume_bwd4_kernel<<<grid, 256, smem4, static_cast<cudaStream_t>(stream)>>>(left, right, grad_cv, grad_left,
                                                                                       grad_right, C, H, W, D, groups);
    SF_LAUNCH_CHECK();
    return SF_OK;
  }
  if (use_smem) SF_CUDA(cudaFuncSetAttribute(corr_volume_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  corr_volume_bwd_kernel<<<grid, 256, use_smem ? smem : 0, static_cast<cudaStream_t>(stream)>>>(
      left, right, grad_cv, grad_left, grad_right, C, H, W, D, groups, use_smem);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

extern "C" int sf_soft_argmin_backward_f32(const float* cv, const float* grad_disp, float* grad_cv, int B, int D, int H,
                                           int W, int match_similarity, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(cv && grad_disp && grad_cv, "soft_argmin_backward: null pointer");
  SF_CHECK_ARG(B > 0 && D > 0 && H > 0 && W > 0, "soft_argmin_backward: non-positive dimension");
  const size_t n = static_cast<size_t>(B) * H * W;
  const int grid = static_cast<int>(min(static_cast<size_t>(sm_count()) * 8, (n + 255) / 256));
  soft_argmin_bwd_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(cv, grad_disp, grad_cv, B, D, H, W,
                                                                            match_similarity ? 1.f : -1.f);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

extern "C" int sf_local_window_backward_f32(const float* cv, const float* disp, const float* lo, const float* up,
                                            const float* grad_out, float* grad_cv, float* grad_disp, float* grad_lo,
                                            float* grad_up, int B, int D, int H, int W, int mode, float radius,
                                            int n_points, int consider_valid, int div_mode, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(cv && disp && grad_out, "local_window_backward: null pointer (the fused soft-argmin form has no backward)");
  SF_CHECK_ARG(B > 0 && D > 0 && H > 0 && W > 0 && n_points > 0, "local_window_backward: non-positive dimension");
  SF_CHECK_ARG(mode == SF_WIN_FIXED || mode == SF_WIN_VARIANCE || mode == SF_WIN_OFFSETS || mode == SF_WIN_ABSOLUTE ||
                   mode == SF_WIN_HEAD, "local_window_backward: unknown mode %d", mode);
  SF_CHECK_ARG(mode == SF_WIN_FIXED || mode == SF_WIN_VARIANCE || lo, "local_window_backward: this mode needs the bound tensors");
  SF_CHECK_ARG((mode != SF_WIN_OFFSETS && mode != SF_WIN_ABSOLUTE) || up, "local_window_backward: upper bound tensor missing");
  const size_t smem = static_cast<size_t>(D) * WB_TPB * sizeof(float);
  SF_CHECK_ARG(smem <= 200 * 1024, "local_window_backward: D=%d too large", D);
  WinBwdParams p;
  p.cv = cv; p.disp = disp; p.lo = lo; p.up = up; p.gout = grad_out;
  p.gcv = grad_cv; p.gdisp = grad_disp; p.glo = grad_lo; p.gup = grad_up;
  p.B = B; p.D = D; p.H = H; p.W = W; p.mode = mode; p.n_points = n_points;
  p.consider_valid = consider_valid ? 1 : 0; p.recip = div_mode == SF_DIV_RECIP; p.radius = radius;
  SF_CUDA(cudaFuncSetAttribute(local_window_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  const size_t n = static_cast<size_t>(B) * H * W;
  local_window_bwd_kernel<<<static_cast<unsigned>((n + WB_TPB - 1) / WB_TPB), WB_TPB, smem, static_cast<cudaStream_t>(stream)>>>(p);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

extern "C" int sf_sample_volume_backward_f32(const float* cv, const float* cand, const float* grad_local, float* grad_cv,
                                             float* grad_cand, int B, int D, int K, int H, int W, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(cv && cand && grad_local, "sample_volume_backward: null pointer");
  SF_CHECK_ARG(B > 0 && D > 0 && K > 0 && H > 0 && W > 0, "sample_volume_backward: non-positive dimension");
  const size_t smem = static_cast<size_t>(D) * WB_TPB * sizeof(float);
  SF_CHECK_ARG(smem <= 200 * 1024, "sample_volume_backward: D=%d too large", D);
  SF_CUDA(cudaFuncSetAttribute(sample_volume_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  const size_t n = static_cast<size_t>(B) * H * W;
  sample_volume_bwd_kernel<<<static_cast<unsigned>((n + WB_TPB - 1) / WB_TPB), WB_TPB, smem, static_cast<cudaStream_t>(stream)>>>(
      cv, cand, grad_local, grad_cv, grad_cand, B, D, K, H, W);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

// RAFT-Stereo 1-D all-pairs correlation pyramid and radius-r lookup  (SURVEY 8a rows a8, a9).
//
// Reference: models/IterativeLocalCostVolume/corr/corr.py:55-111 (CorrBlock1D), utils/utils.py:59-76
// (bilinear_sampler -> F.grid_sample, align_corners=True, zeros padding).
//
//   V0[b,y,x1,x2] = (sum_c f1[b,c,y,x1] * f2[b,c,y,x2]) / sqrt(C);   V(l+1)[..,j] = (Vl[..,2j] + Vl[..,2j+1]) / 2
//
// Build kernel (this file): fp32 FFMA register-tiled GEMM per epipolar row with the whole pyramid produced
// from the accumulator tile in the epilogue (pairwise averages in registers + one lane shuffle), so the
// volume is written once and never re-read.  This is the fp32-exact path; the tcgen05 path lives in
// corr1d_tc.cu when built.  Lookup kernel: one thread per pixel, coordinates through the reference's
// normalise / un-normalise round trip (bit-exact, pixel_math.h), zero padding outside the row.
#include "common.cuh"
#include "pixel_math.h"
#include "../../include/stereoformer_b200.h"

namespace sf {

constexpr int C1_BM = 64, C1_BN = 128, C1_KC = 16, C1_THREADS = 128;
constexpr int C1_MAXLEV = 5;  // pyramid tensors a launch can emit (level 0..4)

struct PyrPtrs {
  float* lvl[C1_MAXLEV];
  int W[C1_MAXLEV];
  int n;  // number of tensors to write
};

__global__ void __launch_bounds__(C1_THREADS)
corr1d_pyramid_kernel(const float* __restrict__ f1, const float* __restrict__ f2, const PyrPtrs out, int C, int H,
                      int W1, int W2, float inv_scale_div) {
  __shared__ __align__(16) float As[C1_KC][C1_BM];
  __shared__ __align__(16) float Bs[C1_KC][C1_BN];
  const int row = blockIdx.z;  // b*H + y
  const int b = row / H, y = row % H;
  const int m0 = blockIdx.y * C1_BM, n0 = blockIdx.x * C1_BN;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const size_t cs = static_cast<size_t>(H);
  const float* a_base = f1 + (static_cast<size_t>(b) * C * cs + y) * W1;
  const float* b_base = f2 + (static_cast<size_t>(b) * C * cs + y) * W2;

  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < C; k0 += C1_KC) {
    for (int i = threadIdx.x; i < C1_KC * C1_BM; i += C1_THREADS) {
      int k = i / C1_BM, m = i % C1_BM;
      As[k][m] = (k0 + k < C && m0 + m < W1) ? __ldg(a_base + static_cast<size_t>(k0 + k) * cs * W1 + m0 + m) : 0.f;
    }
    for (int i = threadIdx.x; i < C1_KC * C1_BN; i += C1_THREADS) {
      int k = i / C1_BN, n = i % C1_BN;
      Bs[k][n] = (k0 + k < C && n0 + n < W2) ? __ldg(b_base + static_cast<size_t>(k0 + k) * cs * W2 + n0 + n) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < C1_KC; ++k) {
      float a[8], bb[8];
      *reinterpret_cast<float4*>(a) = *reinterpret_cast<const float4*>(&As[k][ty * 8]);
      *reinterpret_cast<float4*>(a + 4) = *reinterpret_cast<const float4*>(&As[k][ty * 8 + 4]);
      *reinterpret_cast<float4*>(bb) = *reinterpret_cast<const float4*>(&Bs[k][tx * 8]);
      *reinterpret_cast<float4*>(bb + 4) = *reinterpret_cast<const float4*>(&Bs[k][tx * 8 + 4]);
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
    }
    __syncthreads();
  }

  // ---- epilogue: scale, emit level 0 and the pooled levels from registers
  const int nb = n0 + tx * 8;  // first x2 of this thread (multiple of 8)
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int x1 = m0 + ty * 8 + i;
    const bool rowok = x1 < W1;
    const size_t pix = static_cast<size_t>(row) * W1 + x1;
    float v0[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v0[j] = __fdiv_rn(acc[i][j], inv_scale_div);
    if (rowok) {
      float* o = out.lvl[0] + pix * out.W[0] + nb;
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (nb + j < out.W[0]) o[j] = v0[j];
    }
    float v1[4], v2[2], v3;
#pragma unroll
    for (int j = 0; j < 4; ++j) v1[j] = __fmul_rn(__fadd_rn(v0[2 * j], v0[2 * j + 1]), 0.5f);
    v2[0] = __fmul_rn(__fadd_rn(v1[0], v1[1]), 0.5f);
    v2[1] = __fmul_rn(__fadd_rn(v1[2], v1[3]), 0.5f);
    v3 = __fmul_rn(__fadd_rn(v2[0], v2[1]), 0.5f);
    const float v3n = __shfl_xor_sync(0xffffffffu, v3, 1);
    const float v4 = __fmul_rn(__fadd_rn(v3, v3n), 0.5f);  // valid on even tx
    if (rowok) {
      if (out.n > 1) {
        float* o = out.lvl[1] + pix * out.W[1] + nb / 2;
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (nb / 2 + j < out.W[1]) o[j] = v1[j];
      }
      if (out.n > 2) {
        float* o = out.lvl[2] + pix * out.W[2] + nb / 4;
#pragma unroll
        for (int j = 0; j < 2; ++j)
          if (nb / 4 + j < out.W[2]) o[j] = v2[j];
      }
      if (out.n > 3 && nb / 8 < out.W[3]) out.lvl[3][pix * out.W[3] + nb / 8] = v3;
      if (out.n > 4 && (tx & 1) == 0 && nb / 16 < out.W[4]) out.lvl[4][pix * out.W[4] + nb / 16] = v4;
    }
  }
}

struct LookupParams {
  const float* lvl[C1_MAXLEV];
  int W[C1_MAXLEV];
  int num_levels, radius;
  const float* coords;  // [B,2,H,W1] (only channel 0 is read, corr.py:80)
  float* out;           // [B, L*(2r+1), H, W1]
  float* dbg_x0;        // optional, same shape as out: pixel-space lookup coordinate before normalisation
  float* dbg_src;       // optional, same shape: un-normalised source coordinate handed to the interpolation
  int B, H, W1, recip;
};

__global__ void __launch_bounds__(256) corr1d_lookup_kernel(const LookupParams p) {
  const size_t HW = static_cast<size_t>(p.H) * p.W1, n = static_cast<size_t>(p.B) * HW;
  const int taps = 2 * p.radius + 1;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t b = i / HW, r = i % HW;
    const float cx = p.coords[b * 2 * HW + r];
    float* o = p.out + b * p.num_levels * taps * HW + r;
    for (int l = 0; l < p.num_levels; ++l) {
      const int Wl = p.W[l];
      const float* rowp = p.lvl[l] + i * Wl;
      for (int t = 0; t < taps; ++t) {
        const float x0 = sfm::lookup_x0(cx, l, (float)(t - p.radius));
        const float ix = sfm::lookup_unnorm(sfm::lookup_norm(x0, Wl, p.recip), Wl);
        const float fx = floorf(ix);
        float v = 0.f;
        if (fx >= -1.f && fx < (float)Wl) {  // else both taps are outside the row (also rejects NaN/inf)
          const int xi = (int)fx;
          const float we = (fx + 1.f) - ix, wn = ix - fx;
          if (xi >= 0) v += __ldg(rowp + xi) * we;
          if (xi + 1 < Wl) v += __ldg(rowp + xi + 1) * wn;
        }
        const size_t oi = static_cast<size_t>(l * taps + t) * HW;
        o[oi] = v;
        if (p.dbg_x0) p.dbg_x0[b * p.num_levels * taps * HW + r + oi] = x0;
        if (p.dbg_src) p.dbg_src[b * p.num_levels * taps * HW + r + oi] = ix;
      }
    }
  }
}

bool corr1d_tc_supported(const float* f1, const float* f2, float* const* levels, int n_out, int W1, int W2);
int launch_corr1d_tc(const float* f1, const float* f2, float* const* levels, int n_out, int B, int C, int H, int W1,
                     int W2, cudaStream_t stream);

}  // namespace sf

extern "C" int sf_corr1d_pyramid_f32(const float* f1, const float* f2, float* const* levels, int n_out, int B, int C,
                                     int H, int W1, int W2, int algo, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(f1 && f2 && levels, "corr1d_pyramid: null pointer");
  SF_CHECK_ARG(B > 0 && C > 0 && H > 0 && W1 > 0 && W2 > 0, "corr1d_pyramid: non-positive dimension");
  SF_CHECK_ARG(n_out >= 1 && n_out <= C1_MAXLEV, "corr1d_pyramid: 1..%d pyramid tensors supported, got %d", C1_MAXLEV,
               n_out);
  if (algo == SF_ALGO_EXACT) algo = SF_ALGO_GENERIC;      // the FFMA pyramid kernel is the exact path
  SF_CHECK_ARG(algo == SF_ALGO_AUTO || algo == SF_ALGO_GENERIC || algo == SF_ALGO_TMA, "corr1d_pyramid: unknown algo %d",
               algo);
  for (int l = 0; l < n_out; ++l) SF_CHECK_ARG(levels[l] && (W2 >> l) > 0, "corr1d_pyramid: level %d is empty or null", l);
  // tcgen05 path (bf16 hi/mid split, fp32 accumulation): needs W1, W2 % 4 == 0 and 16-byte aligned pointers
  const bool tc_ok = corr1d_tc_supported(f1, f2, levels, n_out, W1, W2);
  SF_CHECK_ARG(algo != SF_ALGO_TMA || tc_ok, "corr1d_pyramid: tensor-core path needs W %% 4 == 0 and aligned pointers");
  if (algo != SF_ALGO_GENERIC && tc_ok)
    return launch_corr1d_tc(f1, f2, levels, n_out, B, C, H, W1, W2, static_cast<cudaStream_t>(stream));
  PyrPtrs out;
  int w = W2;
  for (int l = 0; l < C1_MAXLEV; ++l) {
    out.lvl[l] = l < n_out ? levels[l] : nullptr;
    out.W[l] = w;
    SF_CHECK_ARG(l >= n_out || (levels[l] && w > 0), "corr1d_pyramid: level %d is empty or null", l);
    w /= 2;
  }
  out.n = n_out;
  SF_CHECK_ARG(static_cast<long long>(B) * H <= 65535, "corr1d_pyramid: B*H too large for one launch");
  dim3 grid(ceil_div(W2, C1_BN), ceil_div(W1, C1_BM), B * H);
  corr1d_pyramid_kernel<<<grid, C1_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(f1, f2, out, C, H, W1, W2,
                                                                                  sqrtf(static_cast<float>(C)));
  SF_LAUNCH_CHECK();
  return SF_OK;
}

extern "C" int sf_corr1d_lookup_f32(const float* const* levels, const int* widths, int num_levels, int radius,
                                    const float* coords, float* out, float* dbg_x0, float* dbg_src, int B, int H,
                                    int W1, int div_mode, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(levels && widths && coords && out, "corr1d_lookup: null pointer");
  SF_CHECK_ARG(num_levels >= 1 && num_levels <= C1_MAXLEV && radius >= 0, "corr1d_lookup: bad num_levels / radius");
  SF_CHECK_ARG(B > 0 && H > 0 && W1 > 0, "corr1d_lookup: non-positive dimension");
  LookupParams p;
  for (int l = 0; l < C1_MAXLEV; ++l) {
    p.lvl[l] = l < num_levels ? levels[l] : nullptr;
    p.W[l] = l < num_levels ? widths[l] : 0;
    SF_CHECK_ARG(l >= num_levels || (levels[l] && widths[l] >= 2), "corr1d_lookup: level %d needs width >= 2", l);
  }
  p.num_levels = num_levels; p.radius = radius; p.coords = coords; p.out = out; p.dbg_x0 = dbg_x0; p.dbg_src = dbg_src;
  p.B = B; p.H = H; p.W1 = W1; p.recip = div_mode == SF_DIV_RECIP;
  size_t n = static_cast<size_t>(B) * H * W1;
  int grid = static_cast<int>(min(static_cast<size_t>(sm_count()) * 8, (n + 255) / 256));
  corr1d_lookup_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

// Deformable convolution in double precision, forward and backward (DCNv2, and DCNv1 with mask == null).
//
// The reference dispatches its DCN over float and double (AT_DISPATCH_FLOATING_TYPES,
// deform/src/cuda/modulated_deform_conv_cuda.cu:93), and its own gradient tests run in double
// (deform/test.py:376-403: gradcheck with eps 1e-3).  fp64 is a verification dtype, not a throughput path (B200 has
// little FP64): these are direct, unfused kernels with one thread per output element and plain double arithmetic, the
// same formulas as the fp32 kernels (dcn.cu / dcn_bwd.cu):
//   col[b,c,t,p]  = mask[b,g,t,p] * bilinear(in[b,c], y*s-p+i*dil+dy, x*s-p+j*dil+dx)   (.cuh:25-54, 140-193)
//   out[b,o,p]    = bias[o] + sum_{c,t} W[o,c,t] * col[b,c,t,p]
//   gcol[b,c,t,p] = sum_o W[o,c,t] * gout[b,o,p]
//   grad_input   += gcol * mask * corner weight                                          (.cuh:197-250, atomicAdd)
//   grad_offset   = sum_{c in dg} gcol * mask * d bilinear / d(h|w);  grad_mask = sum_c gcol * bilinear  (.cuh:253-328)
//   grad_weight   = sum_{b,p} gout * col;  grad_bias = sum_{b,p} gout                    (.cu:236-274)
#include "common.cuh"
#include "../../include/stereoformer_b200.h"

namespace sf {

struct Dcn64 {
  const double* in;      // [B,Cin,H,W]
  const double* weight;  // [Cout,Cin/G,kh,kw]
  const double* bias;    // [Cout]
  const double* offset;  // [B,dg*2*T,Ho,Wo]
  const double* mask;    // [B,dg*T,Ho,Wo] or null
  const double* gout;    // [B,Cout,Ho,Wo]           (backward)
  double* out;           // [B,Cout,Ho,Wo]           (forward)
  double *gin, *gweight, *gbias, *goffset, *gmask;  // backward outputs, any may be null
  int B, Cin, H, W, Cout, Ho, Wo;
  int kh, kw, sh, sw, ph, pw, dh, dw, G, dg;
};

struct Samp64 {
  bool inside;              // -1 < h < H and -1 < w < W
  int hl, wl;
  double lh, lw;
  bool c1, c2, c3, c4;      // corners (hl,wl) (hl,wh) (hh,wl) (hh,wh) inside the image
  double m;
};

__device__ __forceinline__ Samp64 sample64(const Dcn64& p, int b, int dgi, int t, int pix) {
  const int T = p.kh * p.kw, M = p.Ho * p.Wo;
  const int yo = pix / p.Wo, xo = pix % p.Wo, i = t / p.kw, j = t % p.kw;
  const size_t ob = ((static_cast<size_t>(b) * p.dg + dgi) * 2 * T) * M + pix;
  const double h = (double)(yo * p.sh - p.ph + i * p.dh) + p.offset[ob + static_cast<size_t>(2 * t) * M];
  const double w = (double)(xo * p.sw - p.pw + j * p.dw) + p.offset[ob + static_cast<size_t>(2 * t + 1) * M];
  Samp64 s;
  s.m = p.mask ? p.mask[((static_cast<size_t>(b) * p.dg + dgi) * T + t) * M + pix] : 1.0;
  s.inside = h > -1.0 && w > -1.0 && h < (double)p.H && w < (double)p.W;
  s.hl = (int)floor(h); s.wl = (int)floor(w);
  s.lh = h - s.hl; s.lw = w - s.wl;
  const bool t0 = s.hl >= 0, l0 = s.wl >= 0, b1 = s.hl + 1 <= p.H - 1, r1 = s.wl + 1 <= p.W - 1;
  s.c1 = s.inside && t0 && l0; s.c2 = s.inside && t0 && r1; s.c3 = s.inside && b1 && l0; s.c4 = s.inside && b1 && r1;
  return s;
}

__device__ __forceinline__ void corners64(const Dcn64& p, const Samp64& s, const double* plane, double v[4]) {
  const double* q = plane + static_cast<ptrdiff_t>(s.hl) * p.W + s.wl;
  v[0] = s.c1 ? q[0] : 0.0; v[1] = s.c2 ? q[1] : 0.0; v[2] = s.c3 ? q[p.W] : 0.0; v[3] = s.c4 ? q[p.W + 1] : 0.0;
}

__device__ __forceinline__ double bilinear64(const Samp64& s, const double v[4]) {
  const double hh = 1.0 - s.lh, hw = 1.0 - s.lw;
  return hh * hw * v[0] + hh * s.lw * v[1] + s.lh * hw * v[2] + s.lh * s.lw * v[3];
}

// one thread per (b, o, pixel)
__global__ void __launch_bounds__(256) dcn64_forward_kernel(const Dcn64 p) {
  const int T = p.kh * p.kw, M = p.Ho * p.Wo, Cig = p.Cin / p.G, Cog = p.Cout / p.G, cpdg = p.Cin / p.dg;
  const size_t n = static_cast<size_t>(p.B) * p.Cout * M, HW = static_cast<size_t>(p.H) * p.W;
  for (size_t idx = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; idx < n;
       idx += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int pix = idx % M, o = (idx / M) % p.Cout, b = idx / (static_cast<size_t>(M) * p.Cout);
    const int g = o / Cog;
    double acc = p.bias[o];
    for (int c = 0; c < Cig; ++c) {
      const int cin = g * Cig + c;
      const double* plane = p.in + (static_cast<size_t>(b) * p.Cin + cin) * HW;
      for (int t = 0; t < T; ++t) {
        const Samp64 s = sample64(p, b, cin / cpdg, t, pix);
        if (!s.inside) continue;
        double v[4];
        corners64(p, s, plane, v);
        acc += p.weight[(static_cast<size_t>(o) * Cig + c) * T + t] * s.m * bilinear64(s, v);
      }
    }
    p.out[idx] = acc;
  }
}

// one thread per (b, dg, tap, pixel): offset / mask gradients (sums over the group's channels) and the grad_input scatter
__global__ void __launch_bounds__(256) dcn64_bwd_data_kernel(const Dcn64 p) {
  const int T = p.kh * p.kw, M = p.Ho * p.Wo, Cig = p.Cin / p.G, Cog = p.Cout / p.G, cpdg = p.Cin / p.dg;
  const size_t n = static_cast<size_t>(p.B) * p.dg * T * M, HW = static_cast<size_t>(p.H) * p.W;
  for (size_t idx = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; idx < n;
       idx += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int pix = idx % M, t = (idx / M) % T, dgi = (idx / (static_cast<size_t>(M) * T)) % p.dg;
    const int b = idx / (static_cast<size_t>(M) * T * p.dg);
    const Samp64 s = sample64(p, b, dgi, t, pix);
    const double hh = 1.0 - s.lh, hw = 1.0 - s.lw;
    double a_h = 0.0, a_w = 0.0, a_m = 0.0;
    if (s.inside) {
      for (int cin = dgi * cpdg; cin < (dgi + 1) * cpdg; ++cin) {
        const int g = cin / Cig, c = cin % Cig;
        double gcol = 0.0;
        for (int o = g * Cog; o < (g + 1) * Cog; ++o)
          gcol += p.weight[(static_cast<size_t>(o) * Cig + c) * T + t] * p.gout[(static_cast<size_t>(b) * p.Cout + o) * M + pix];
        const size_t plane = (static_cast<size_t>(b) * p.Cin + cin) * HW;
        double v[4];
        corners64(p, s, p.in + plane, v);
        a_m += gcol * bilinear64(s, v);
        const double gm = gcol * s.m;
        a_h += gm * (-hw * v[0] - s.lw * v[1] + hw * v[2] + s.lw * v[3]);   // .cuh:101-110 (bp_dir 0)
        a_w += gm * (-hh * v[0] + hh * v[1] - s.lh * v[2] + s.lh * v[3]);   // .cuh:112-121 (bp_dir 1)
        if (p.gin) {
          double* q = p.gin + plane + static_cast<ptrdiff_t>(s.hl) * p.W + s.wl;
          if (s.c1) atomicAdd(q, hh * hw * gm);
          if (s.c2) atomicAdd(q + 1, hh * s.lw * gm);
          if (s.c3) atomicAdd(q + p.W, s.lh * hw * gm);
          if (s.c4) atomicAdd(q + p.W + 1, s.lh * s.lw * gm);
        }
      }
    }
    if (p.goffset) {
      const size_t ob = ((static_cast<size_t>(b) * p.dg + dgi) * 2 * T) * M + pix;
      p.goffset[ob + static_cast<size_t>(2 * t) * M] = a_h;
      p.goffset[ob + static_cast<size_t>(2 * t + 1) * M] = a_w;
    }
    if (p.gmask) p.gmask[((static_cast<size_t>(b) * p.dg + dgi) * T + t) * M + pix] = a_m;
  }
}

__device__ __forceinline__ double block_sum64(double v) {
  __shared__ double red[8];
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = 0.0;
  for (int i = 0; i < (blockDim.x >> 5); ++i) s += red[i];
  return s;
}

// one CTA per weight element (o, c, t); block y == 1 row: bias (one CTA per o).  Deterministic reductions.
__global__ void __launch_bounds__(256) dcn64_bwd_weight_kernel(const Dcn64 p) {
  const int T = p.kh * p.kw, M = p.Ho * p.Wo, Cig = p.Cin / p.G, Cog = p.Cout / p.G, cpdg = p.Cin / p.dg;
  const size_t HW = static_cast<size_t>(p.H) * p.W;
  const int e = blockIdx.x;
  const int t = e % T, c = (e / T) % Cig, o = e / (T * Cig);
  const int cin = (o / Cog) * Cig + c;
  double acc = 0.0;
  for (int i = threadIdx.x; i < p.B * M; i += blockDim.x) {
    const int b = i / M, pix = i % M;
    const Samp64 s = sample64(p, b, cin / cpdg, t, pix);
    if (!s.inside) continue;
    double v[4];
    corners64(p, s, p.in + (static_cast<size_t>(b) * p.Cin + cin) * HW, v);
    acc += p.gout[(static_cast<size_t>(b) * p.Cout + o) * M + pix] * s.m * bilinear64(s, v);
  }
  acc = block_sum64(acc);
  if (threadIdx.x == 0) p.gweight[e] = acc;
}

__global__ void __launch_bounds__(256) dcn64_bwd_bias_kernel(const Dcn64 p) {
  const int M = p.Ho * p.Wo, o = blockIdx.x;
  double acc = 0.0;
  for (int i = threadIdx.x; i < p.B * M; i += blockDim.x)
    acc += p.gout[(static_cast<size_t>(i / M) * p.Cout + o) * M + i % M];
  acc = block_sum64(acc);
  if (threadIdx.x == 0) p.gbias[o] = acc;
}

static int fill64(Dcn64& p, int B, int Cin, int H, int W, int Cout, int kh, int kw, int sh, int sw, int ph, int pw, int dh,
                  int dw, int G, int dg) {
  SF_CHECK_ARG(B > 0 && Cin > 0 && H > 0 && W > 0 && Cout > 0 && kh > 0 && kw > 0, "dcn_f64: non-positive dimension");
  SF_CHECK_ARG(sh > 0 && sw > 0 && dh > 0 && dw > 0 && ph >= 0 && pw >= 0, "dcn_f64: bad stride / dilation / padding");
  SF_CHECK_ARG(G > 0 && Cin % G == 0 && Cout % G == 0, "channels(%d) and channels_out(%d) must divide group(%d)", Cin,
               Cout, G);
  SF_CHECK_ARG(dg > 0 && Cin % dg == 0, "channels(%d) must divide deformable_group(%d)", Cin, dg);
  p.B = B; p.Cin = Cin; p.H = H; p.W = W; p.Cout = Cout; p.kh = kh; p.kw = kw; p.sh = sh; p.sw = sw; p.ph = ph; p.pw = pw;
  p.dh = dh; p.dw = dw; p.G = G; p.dg = dg;
  p.Ho = (H + 2 * ph - (dh * (kh - 1) + 1)) / sh + 1;
  p.Wo = (W + 2 * pw - (dw * (kw - 1) + 1)) / sw + 1;
  SF_CHECK_ARG(p.Ho > 0 && p.Wo > 0, "dcn_f64: empty output (%d x %d)", p.Ho, p.Wo);
  return SF_OK;
}

}  // namespace sf

extern "C" int sf_dcn_forward_f64(const double* input, const double* weight, const double* bias, const double* offset,
                                  const double* mask, double* out, int B, int Cin, int H, int W, int Cout, int kh,
                                  int kw, int stride_h, int stride_w, int pad_h, int pad_w, int dil_h, int dil_w,
                                  int group, int deformable_group, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(input && weight && bias && offset && out, "dcn_forward_f64: null pointer");
  Dcn64 p = {};
  if (int rc = fill64(p, B, Cin, H, W, Cout, kh, kw, stride_h, stride_w, pad_h, pad_w, dil_h, dil_w, group,
                      deformable_group)) return rc;
  p.in = input; p.weight = weight; p.bias = bias; p.offset = offset; p.mask = mask; p.out = out;
  const size_t n = static_cast<size_t>(B) * Cout * p.Ho * p.Wo;
  const int grid = static_cast<int>(min(static_cast<size_t>(sm_count()) * 8, (n + 255) / 256));
  dcn64_forward_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

extern "C" int sf_dcn_backward_f64(const double* input, const double* weight, const double* offset, const double* mask,
                                   const double* grad_out, double* grad_input, double* grad_weight, double* grad_bias,
                                   double* grad_offset, double* grad_mask, int B, int Cin, int H, int W, int Cout, int kh,
                                   int kw, int stride_h, int stride_w, int pad_h, int pad_w, int dil_h, int dil_w,
                                   int group, int deformable_group, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(input && weight && offset && grad_out, "dcn_backward_f64: null pointer");
  Dcn64 p = {};
  if (int rc = fill64(p, B, Cin, H, W, Cout, kh, kw, stride_h, stride_w, pad_h, pad_w, dil_h, dil_w, group,
                      deformable_group)) return rc;
  p.in = input; p.weight = weight; p.offset = offset; p.mask = mask; p.gout = grad_out;
  p.gin = grad_input; p.goffset = grad_offset; p.gmask = mask ? grad_mask : nullptr; p.gweight = grad_weight;
  p.gbias = grad_bias;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int T = kh * kw, M = p.Ho * p.Wo;
  if (grad_input) SF_CUDA(cudaMemsetAsync(grad_input, 0, sizeof(double) * static_cast<size_t>(B) * Cin * H * W, s));
  if (grad_input || grad_offset || p.gmask) {
    const size_t n = static_cast<size_t>(B) * deformable_group * T * M;
    const int grid = static_cast<int>(min(static_cast<size_t>(sm_count()) * 8, (n + 255) / 256));
    dcn64_bwd_data_kernel<<<grid, 256, 0, s>>>(p);
    SF_LAUNCH_CHECK();
  }
  if (grad_weight) {
    dcn64_bwd_weight_kernel<<<Cout * (Cin / group) * T, 256, 0, s>>>(p);
    SF_LAUNCH_CHECK();
  }
  if (grad_bias) {
    dcn64_bwd_bias_kernel<<<Cout, 256, 0, s>>>(p);
    SF_LAUNCH_CHECK();
  }
  return SF_OK;
}

// Deformable convolution backward, DCNv2 (modulated) and DCNv1 (mask == null)  (SURVEY 8f rank 3).
//
// Reference: deform/src/cuda/modulated_deform_conv_cuda.cu:124-280 (host: columns = W^T * grad_out per group, then
// col2im_coord, col2im, im2col + addmm / addmv), modulated_deform_im2col_cuda.cuh:56-126 (gradient / coordinate
// weights), :197-328 (col2im and col2im_coord kernels); DCNv1 twin deform_conv_cuda.cu / deform_im2col_cuda.cuh.
//
// With col[b,c,t,p] = mask[b,g,t,p] * bilinear(in[b,c], pos(b,g,t,p)) and gcol[b,c,t,p] = sum_o W[o,c,t] * gout[b,o,p]:
//   grad_input[b,c,corner]  += gcol * mask * bilinear weight of that corner      (reference: atomicAdd, :247)
//   grad_offset[b,g,2t+dir,p] = sum_{c in g} gcol * mask * d bilinear / d (h|w)   (coordinate weight, :86-126)
//   grad_mask[b,g,t,p]        = sum_{c in g} gcol * bilinear(in[b,c], pos)
//   grad_weight[o,c,t]        = sum_{b,p} gout[b,o,p] * col[b,c,t,p];   grad_bias[o] = sum_{b,p} gout[b,o,p]
//
// The reference materialises `columns` [Cin*kh*kw, B*Ho*Wo] in HBM twice (gcol, then col) and runs cuBLAS around
// five kernels.  Here neither buffer exists:
//   * dcn_bwd_data_kernel   -- per 64-pixel tile: gcol tile by a register-tiled FFMA GEMM from shared memory, then
//                              every (tap, pixel) owner scatters it to grad_input (4 atomics) and accumulates its
//                              offset / mask gradients over the channels of the deformable group (plain stores:
//                              a CTA owns its pixels across all groups).
//   * dcn_bwd_weight_kernel -- per channel chunk and pixel range: column tile re-gathered (as in the forward),
//                              acc[o][k] += gout[o][px] * col[k][px] in registers, one atomicAdd per element per CTA.
//   * dcn_bwd_bias_kernel   -- one CTA per output channel.
#include <algorithm>
#include "common.cuh"
#include "dcn_tc.cuh"
#include "../../include/stereoformer_b200.h"

namespace sf {

constexpr int BW_TP = 64;     // pixels per tile
constexpr int BW_NT = 640;    // data kernel: all threads gather / scatter (latency-bound), the first 144 run the GEMM
constexpr int BW_NTW = 256;   // weight kernel (GEMM-bound: more threads measured slower)
constexpr int BW_OB = 256;    // output channels per shared-memory block
constexpr int BW_NK = 36;     // (channel, tap) pairs per chunk: a 4 x 9 register tile per thread in the weight kernel

struct DcnBwdParams {
  const float* in;      // [B,Cin,H,W]
  const float* weight;  // [Cout,Cin/G,kh,kw]
  const float* offset;  // [B,dg*2*T,Ho,Wo]
  const float* mask;    // [B,dg*T,Ho,Wo] or null (DCNv1)
  const float* gout;    // [B,Cout,Ho,Wo]
  float* gin;           // [B,Cin,H,W]   (zero-initialised, atomics) or null
  float* gweight;       // [Cout,Cin/G,kh,kw] (zero-initialised, atomics) or null
  float* gbias;         // [Cout] or null
  float* goffset;       // like offset, or null
  float* gmask;         // like mask, or null
  int B, Cin, H, W, Cout, Ho, Wo;
  int kh, kw, sh, sw, ph, pw, dh, dw, G, dg;
  int KC;               // channels per chunk = max(1, BW_NK / T)
  int seg;              // gcd(Cin/G, Cin/dg): a chunk never straddles a group or deformable-group boundary
  int nsplit;           // pixel ranges of the weight kernel
};

// sampling descriptor of one (tap, pixel): modulated_deform_im2col_cuda.cuh:25-54 (corners), :56-126 (gradients)
struct Desc {
  int4 off;        // corner offsets inside one [H][W] plane (0 when the corner is outside)
  float lh, lw;    // fractional parts
  float m;         // mask value (1 for DCNv1)
  int flags;       // bit 0..3: corner (hl,wl), (hl,wh), (hh,wl), (hh,wh) inside the image; 0 when the sample is outside
};

__device__ __forceinline__ Desc make_desc(const DcnBwdParams& p, int b, int dgi, int t, int pix, int M) {
  Desc d;
  d.off = make_int4(0, 0, 0, 0); d.lh = d.lw = 0.f; d.m = 0.f; d.flags = 0;
  if (pix >= M) return d;
  const int T = p.kh * p.kw;
  const int yo = pix / p.Wo, xo = pix % p.Wo;
  const int i = t / p.kw, j = t % p.kw;
  const size_t ob = ((static_cast<size_t>(b) * p.dg + dgi) * 2 * T) * M + pix;
  const float oh = __ldg(p.offset + ob + static_cast<size_t>(2 * t) * M);
  const float ow = __ldg(p.offset + ob + static_cast<size_t>(2 * t + 1) * M);
  d.m = p.mask ? __ldg(p.mask + ((static_cast<size_t>(b) * p.dg + dgi) * T + t) * M + pix) : 1.f;
  const float h_im = (float)(yo * p.sh - p.ph + i * p.dh) + oh;
  const float w_im = (float)(xo * p.sw - p.pw + j * p.dw) + ow;
  if (h_im > -1.f && w_im > -1.f && h_im < (float)p.H && w_im < (float)p.W) {
    const int hl = (int)floorf(h_im), wl = (int)floorf(w_im);
    const int hh = hl + 1, wh = wl + 1;
    d.lh = h_im - (float)hl; d.lw = w_im - (float)wl;
    const bool t0 = hl >= 0, l0 = wl >= 0, b1 = hh <= p.H - 1, r1 = wh <= p.W - 1;
    if (t0 && l0) { d.off.x = hl * p.W + wl; d.flags |= 1; }
    if (t0 && r1) { d.off.y = hl * p.W + wh; d.flags |= 2; }
    if (b1 && l0) { d.off.z = hh * p.W + wl; d.flags |= 4; }
    if (b1 && r1) { d.off.w = hh * p.W + wh; d.flags |= 8; }
  }
  return d;
}

// ---------------------------------------------------------------------------------------------- data gradients
__global__ void __launch_bounds__(BW_NT)
dcn_bwd_data_kernel(const DcnBwdParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int T = p.kh * p.kw, M = p.Ho * p.Wo;
  const int NKP = BW_NK + 4;                                          // padded (c,t) pitch of the weight tile
  float* gS = reinterpret_cast<float*>(smem_raw);                     // [BW_OB][TP]   grad_out tile
  float* wS = gS + BW_OB * BW_TP;                                     // [BW_OB][NKP]  weight chunk, (c,t) fastest
  float* gc = wS + BW_OB * NKP;                                       // [BW_NK][TP]   gcol tile
  Desc* ds = reinterpret_cast<Desc*>(gc + BW_NK * BW_TP);             // [T][TP]
  float* accS = reinterpret_cast<float*>(ds + T * BW_TP);             // [3][T][TP]: d/dh, d/dw, d/dmask

  const int pix0 = blockIdx.x * BW_TP, b = blockIdx.y;
  const int tid = threadIdx.x;
  const int Cig = p.Cin / p.G, Cog = p.Cout / p.G, cpdg = p.Cin / p.dg;
  const size_t HW = static_cast<size_t>(p.H) * p.W;
  const int tx = tid % (BW_TP / 4), ty = tid / (BW_TP / 4);           // GEMM tile: 4 pixels x 4 (c,t) rows

  int cur_dg = -1, cur_g = -1, cur_o0 = -1;
  for (int c0 = 0; c0 < p.Cin;) {
    const int g = c0 / Cig, dgi = c0 / cpdg;
    const int nkc = min(min(p.KC, (g + 1) * Cig - c0), (dgi + 1) * cpdg - c0);
    const int nk = nkc * T;
    __syncthreads();
    if (dgi != cur_dg) {
      if (cur_dg >= 0) {
        // flush the finished deformable group (col2im_coord's stores, .cuh:322-326)
        for (int e = tid; e < T * BW_TP; e += BW_NT) {
          const int px = e % BW_TP, t = e / BW_TP, pix = pix0 + px;
          if (pix < M) {
            if (p.goffset) {
              const size_t ob = ((static_cast<size_t>(b) * p.dg + cur_dg) * 2 * T) * M + pix;
              p.goffset[ob + static_cast<size_t>(2 * t) * M] = accS[e];
              p.goffset[ob + static_cast<size_t>(2 * t + 1) * M] = accS[T * BW_TP + e];
            }
            if (p.gmask) p.gmask[((static_cast<size_t>(b) * p.dg + cur_dg) * T + t) * M + pix] = accS[2 * T * BW_TP + e];
          }
        }
      }
      cur_dg = dgi;
      for (int e = tid; e < T * BW_TP; e += BW_NT) {
        ds[e] = make_desc(p, b, dgi, e / BW_TP, pix0 + e % BW_TP, M);
        accS[e] = 0.f; accS[T * BW_TP + e] = 0.f; accS[2 * T * BW_TP + e] = 0.f;
      }
    }
    // gcol[k][px] = sum_o W[g*Cog+o][c0-g*Cig .. ][t] * gout[b][g*Cog+o][px], o in blocks of BW_OB
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    for (int o0 = 0; o0 < Cog; o0 += BW_OB) {
      const int no = min(BW_OB, Cog - o0);
      __syncthreads();                                                // previous readers of gS / wS / gc are done
      if (g != cur_g || o0 != cur_o0) {
        for (int e = tid; e < no * BW_TP; e += BW_NT) {
          const int px = e % BW_TP, o = e / BW_TP;
          gS[e] = (pix0 + px < M) ? __ldg(p.gout + (static_cast<size_t>(b) * p.Cout + g * Cog + o0 + o) * M + pix0 + px) : 0.f;
        }
        cur_g = g; cur_o0 = o0;
      }
      for (int e = tid; e < no * nk; e += BW_NT) {
        const int k = e % nk, o = e / nk;
        wS[o * NKP + k] = __ldg(p.weight + (static_cast<size_t>(g * Cog + o0 + o) * Cig + (c0 - g * Cig)) * T + k);
      }
      __syncthreads();
      if (ty * 4 < nk && tid < 16 * 9) {
        for (int o = 0; o < no; ++o) {
          const float4 gv = *reinterpret_cast<const float4*>(gS + o * BW_TP + tx * 4);
          const float4 wv = *reinterpret_cast<const float4*>(wS + o * NKP + ty * 4);
          const float ga[4] = {gv.x, gv.y, gv.z, gv.w}, wa[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(wa[i], ga[j], acc[i][j]);
        }
      }
    }
    if (ty * 4 < nk && tid < 16 * 9) {
#pragma unroll
      for (int i = 0; i < 4; ++i)
        if (ty * 4 + i < nk)
          *reinterpret_cast<float4*>(gc + (ty * 4 + i) * BW_TP + tx * 4) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
    }
    __syncthreads();
    // scatter / reduce: the owner of (t, px) walks the chunk's channels
    for (int e = tid; e < T * BW_TP; e += BW_NT) {
      const int px = e % BW_TP, t = e / BW_TP;
      const Desc d = ds[e];
      if (d.flags == 0) continue;                                     // sample outside: every gradient is zero
      const float hh = 1.f - d.lh, hw = 1.f - d.lw;
      float a_h = 0.f, a_w = 0.f, a_m = 0.f;
      for (int c = 0; c < nkc; ++c) {
        const float gcv = gc[(c * T + t) * BW_TP + px];
        const size_t plane = (static_cast<size_t>(b) * p.Cin + c0 + c) * HW;
        const float* src = p.in + plane;
        const float v1 = (d.flags & 1) ? __ldg(src + d.off.x) : 0.f, v2 = (d.flags & 2) ? __ldg(src + d.off.y) : 0.f;
        const float v3 = (d.flags & 4) ? __ldg(src + d.off.z) : 0.f, v4 = (d.flags & 8) ? __ldg(src + d.off.w) : 0.f;
        a_m = fmaf(gcv, hh * hw * v1 + hh * d.lw * v2 + d.lh * hw * v3 + d.lh * d.lw * v4, a_m);
        const float gm = gcv * d.m;
        a_h = fmaf(gm, -hw * v1 - d.lw * v2 + hw * v3 + d.lw * v4, a_h);   // .cuh:101-110 (bp_dir 0)
        a_w = fmaf(gm, -hh * v1 + hh * v2 - d.lh * v3 + d.lh * v4, a_w);   // .cuh:112-121 (bp_dir 1)
        if (p.gin) {
          float* dst = p.gin + plane;
          if (d.flags & 1) atomicAdd(dst + d.off.x, hh * hw * gm);
          if (d.flags & 2) atomicAdd(dst + d.off.y, hh * d.lw * gm);
          if (d.flags & 4) atomicAdd(dst + d.off.z, d.lh * hw * gm);
          if (d.flags & 8) atomicAdd(dst + d.off.w, d.lh * d.lw * gm);
        }
      }
      accS[e] += a_h; accS[T * BW_TP + e] += a_w; accS[2 * T * BW_TP + e] += a_m;
    }
    c0 += nkc;
  }
  __syncthreads();
  for (int e = tid; e < T * BW_TP; e += BW_NT) {
    const int px = e % BW_TP, t = e / BW_TP, pix = pix0 + px;
    if (pix < M) {
      if (p.goffset) {
        const size_t ob = ((static_cast<size_t>(b) * p.dg + cur_dg) * 2 * T) * M + pix;
        p.goffset[ob + static_cast<size_t>(2 * t) * M] = accS[e];
        p.goffset[ob + static_cast<size_t>(2 * t + 1) * M] = accS[T * BW_TP + e];
      }
      if (p.gmask) p.gmask[((static_cast<size_t>(b) * p.dg + cur_dg) * T + t) * M + pix] = accS[2 * T * BW_TP + e];
    }
  }
}

// ---------------------------------------------------------------------------------------------- weight gradient
__global__ void __launch_bounds__(BW_NTW)
dcn_bwd_weight_kernel(const DcnBwdParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int T = p.kh * p.kw, M = p.Ho * p.Wo;
  constexpr int OP = BW_OB + 4, KP = BW_NK + 4;
  float* gS = reinterpret_cast<float*>(smem_raw);                     // [TP][OP]  grad_out tile, o fastest
  float* colS = gS + BW_TP * OP;                                      // [TP][KP]  column tile, (c,t) fastest
  Desc* ds = reinterpret_cast<Desc*>(colS + BW_TP * KP);              // [T][TP]

  const int tid = threadIdx.x;
  const int Cig = p.Cin / p.G, Cog = p.Cout / p.G, cpdg = p.Cin / p.dg;
  const int cps = (p.seg + p.KC - 1) / p.KC;                                   // chunks per segment
  const int sgi = blockIdx.x / cps, c0 = sgi * p.seg + (blockIdx.x % cps) * p.KC;
  const int nkc = min(p.KC, (sgi + 1) * p.seg - c0), nk = nkc * T;
  const int g = c0 / Cig, dgi = c0 / cpdg;
  const int o0 = blockIdx.z * BW_OB, no = min(BW_OB, Cog - o0);
  const size_t HW = static_cast<size_t>(p.H) * p.W;
  const int to = tid % (BW_OB / 4), tk = tid / (BW_OB / 4);           // 4 output channels x 9 (c,t) pairs per thread

  float acc[4][9];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 9; ++j) acc[i][j] = 0.f;

  const int tiles_per_b = (M + BW_TP - 1) / BW_TP;
  const long long ntiles = 1LL * p.B * tiles_per_b;
  for (long long tile = blockIdx.y; tile < ntiles; tile += gridDim.y) {
    const int b = static_cast<int>(tile / tiles_per_b), pix0 = static_cast<int>(tile % tiles_per_b) * BW_TP;
    __syncthreads();
    for (int e = tid; e < T * BW_TP; e += BW_NTW) ds[e] = make_desc(p, b, dgi, e / BW_TP, pix0 + e % BW_TP, M);
    for (int e = tid; e < no * BW_TP; e += BW_NTW) {
      const int px = e % BW_TP, o = e / BW_TP;
      gS[px * OP + o] = (pix0 + px < M) ? __ldg(p.gout + (static_cast<size_t>(b) * p.Cout + g * Cog + o0 + o) * M + pix0 + px) : 0.f;
    }
    __syncthreads();
    for (int e = tid; e < nk * BW_TP; e += BW_NTW) {
      const int px = e % BW_TP, k = e / BW_TP;
      const int c = k / T, t = k % T;
      const Desc d = ds[t * BW_TP + px];
      float val = 0.f;
      if (d.flags) {
        const float* src = p.in + (static_cast<size_t>(b) * p.Cin + c0 + c) * HW;
        const float hh = 1.f - d.lh, hw = 1.f - d.lw;
        const float v1 = (d.flags & 1) ? __ldg(src + d.off.x) : 0.f, v2 = (d.flags & 2) ? __ldg(src + d.off.y) : 0.f;
        const float v3 = (d.flags & 4) ? __ldg(src + d.off.z) : 0.f, v4 = (d.flags & 8) ? __ldg(src + d.off.w) : 0.f;
        val = (hh * hw * v1 + hh * d.lw * v2 + d.lh * hw * v3 + d.lh * d.lw * v4) * d.m;
      }
      colS[px * KP + k] = val;
    }
    __syncthreads();
    if (tid < 256 && to * 4 < no && tk * 9 < nk) {
      for (int px = 0; px < BW_TP; ++px) {
        const float4 gv = *reinterpret_cast<const float4*>(gS + px * OP + to * 4);
        const float ga[4] = {gv.x, gv.y, gv.z, gv.w};
        const float* cp = colS + px * KP + tk * 9;
#pragma unroll
        for (int j = 0; j < 9; ++j) {
          const float cv = cp[j];
#pragma unroll
          for (int i = 0; i < 4; ++i) acc[i][j] = fmaf(ga[i], cv, acc[i][j]);
        }
      }
    }
  }
  // one atomicAdd per element per CTA (the reference accumulates with addmm over im2col steps, .cu:268)
  if (tid >= 256) return;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int o = to * 4 + i;
    if (o >= no) continue;
    float* dst = p.gweight + (static_cast<size_t>(g * Cog + o0 + o) * Cig + (c0 - g * Cig)) * T;
#pragma unroll
    for (int j = 0; j < 9; ++j)
      if (tk * 9 + j < nk) atomicAdd(dst + tk * 9 + j, acc[i][j]);
  }
}

// one CTA of 1024 threads per output channel, 16-byte loads when the rows allow it (a 256-thread scalar version ran at
// 1.6 TB/s: 28 us for the 45.7 MB of grad_out at 128 channels x 248 x 360)
__global__ void __launch_bounds__(1024)
dcn_bwd_bias_kernel(const float* __restrict__ gout, float* __restrict__ gbias, int B, int Cout, int M) {
  __shared__ float red[32];
  const int o = blockIdx.x;
  const bool vec = (M % 4 == 0) && ((reinterpret_cast<uintptr_t>(gout) & 15) == 0);
  float s = 0.f;
  for (int b = 0; b < B; ++b) {
    const float* src = gout + (static_cast<size_t>(b) * Cout + o) * M;
    if (vec) {
      const float4* s4 = reinterpret_cast<const float4*>(src);
      for (int i = threadIdx.x; i < (M >> 2); i += 1024) {
        const float4 v = __ldg(s4 + i);
        s += (v.x + v.y) + (v.z + v.w);
      }
    } else {
      for (int i = threadIdx.x; i < M; i += 1024) s += __ldg(src + i);
    }
  }
  for (int d = 16; d > 0; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = red[threadIdx.x];
    for (int d = 16; d > 0; d >>= 1) t += __shfl_xor_sync(0xffffffffu, t, d);
    if (threadIdx.x == 0) gbias[o] = t;
  }
}

static int gcd_i(int a, int b) { return b ? gcd_i(b, a % b) : a; }

}  // namespace sf

static int dcn_backward_impl(const float* input, const float* weight, const float* offset, const float* mask,
                             const float* grad_out, float* grad_input, float* grad_weight, float* grad_bias,
                             float* grad_offset, float* grad_mask, int B, int Cin, int H, int W, int Cout, int kh,
                             int kw, int stride_h, int stride_w, int pad_h, int pad_w, int dil_h, int dil_w,
                             int group, int deformable_group, void* workspace, size_t workspace_bytes, int algo,
                             void* stream);

extern "C" int sf_dcn_backward_f32(const float* input, const float* weight, const float* offset, const float* mask,
                                   const float* grad_out, float* grad_input, float* grad_weight, float* grad_bias,
                                   float* grad_offset, float* grad_mask, int B, int Cin, int H, int W, int Cout, int kh,
                                   int kw, int stride_h, int stride_w, int pad_h, int pad_w, int dil_h, int dil_w,
                                   int group, int deformable_group, void* stream) {
  return dcn_backward_impl(input, weight, offset, mask, grad_out, grad_input, grad_weight, grad_bias, grad_offset,
                           grad_mask, B, Cin, H, W, Cout, kh, kw, stride_h, stride_w, pad_h, pad_w, dil_h, dil_w, group,
                           deformable_group, nullptr, 0, SF_ALGO_GENERIC, stream);
}

extern "C" size_t sf_dcn_backward_workspace_bytes(int B, int Cin, int H, int W, int Cout, int kh, int kw, int stride_h,
                                                  int stride_w, int pad_h, int pad_w, int dil_h, int dil_w, int group,
                                                  int deformable_group) {
  const size_t a = sf::dcn_bwd_tc_workspace_bytes(B, Cin, H, W, Cout, kh, kw, group, deformable_group);
  const size_t b = sf::dcn_bwd_data_tc_workspace_bytes(B, Cin, H, W, Cout, kh, kw, stride_h, stride_w, pad_h, pad_w, dil_h,
                                                       dil_w, group, deformable_group);
  return a > b ? a : b;
}

extern "C" int sf_dcn_backward_ws_f32(const float* input, const float* weight, const float* offset, const float* mask,
                                      const float* grad_out, float* grad_input, float* grad_weight, float* grad_bias,
                                      float* grad_offset, float* grad_mask, int B, int Cin, int H, int W, int Cout,
                                      int kh, int kw, int stride_h, int stride_w, int pad_h, int pad_w, int dil_h,
                                      int dil_w, int group, int deformable_group, void* workspace,
                                      size_t workspace_bytes, int algo, void* stream) {
  return dcn_backward_impl(input, weight, offset, mask, grad_out, grad_input, grad_weight, grad_bias, grad_offset,
                           grad_mask, B, Cin, H, W, Cout, kh, kw, stride_h, stride_w, pad_h, pad_w, dil_h, dil_w, group,
                           deformable_group, workspace, workspace_bytes, algo, stream);
}

static int dcn_backward_impl(const float* input, const float* weight, const float* offset, const float* mask,
                             const float* grad_out, float* grad_input, float* grad_weight, float* grad_bias,
                             float* grad_offset, float* grad_mask, int B, int Cin, int H, int W, int Cout, int kh,
                             int kw, int stride_h, int stride_w, int pad_h, int pad_w, int dil_h, int dil_w,
                             int group, int deformable_group, void* workspace, size_t workspace_bytes, int algo,
                             void* stream) {
  using namespace sf;
  if (algo == SF_ALGO_EXACT) algo = SF_ALGO_GENERIC;
  SF_CHECK_ARG(algo == SF_ALGO_AUTO || algo == SF_ALGO_GENERIC || algo == SF_ALGO_TC, "dcn_backward: unknown algo %d", algo);
  SF_CHECK_ARG(input && weight && offset && grad_out, "dcn_backward: null pointer");
  SF_CHECK_ARG(B > 0 && Cin > 0 && H > 0 && W > 0 && Cout > 0 && kh > 0 && kw > 0, "dcn_backward: non-positive dimension");
  SF_CHECK_ARG(stride_h > 0 && stride_w > 0 && dil_h > 0 && dil_w > 0 && pad_h >= 0 && pad_w >= 0,
               "dcn_backward: bad stride / dilation / padding");
  SF_CHECK_ARG(group > 0 && Cin % group == 0 && Cout % group == 0,
               "channels(%d) and channels_out(%d) must divide group(%d)", Cin, Cout, group);
  SF_CHECK_ARG(deformable_group > 0 && Cin % deformable_group == 0,
               "channels(%d) must divide deformable_group(%d)", Cin, deformable_group);
  SF_CHECK_ARG(kh * kw <= BW_NK, "dcn_backward: kernels above %d taps are not supported (got %dx%d)", BW_NK, kh, kw);
  SF_CHECK_ARG(mask || !grad_mask, "dcn_backward: grad_mask without mask (DCNv1 has none)");
  DcnBwdParams p;
  p.in = input; p.weight = weight; p.offset = offset; p.mask = mask; p.gout = grad_out;
  p.gin = grad_input; p.gweight = grad_weight; p.gbias = grad_bias; p.goffset = grad_offset; p.gmask = grad_mask;
  p.B = B; p.Cin = Cin; p.H = H; p.W = W; p.Cout = Cout;
  p.kh = kh; p.kw = kw; p.sh = stride_h; p.sw = stride_w; p.ph = pad_h; p.pw = pad_w; p.dh = dil_h; p.dw = dil_w;
  p.G = group; p.dg = deformable_group;
  p.Ho = (H + 2 * pad_h - (dil_h * (kh - 1) + 1)) / stride_h + 1;
  p.Wo = (W + 2 * pad_w - (dil_w * (kw - 1) + 1)) / stride_w + 1;
  SF_CHECK_ARG(p.Ho > 0 && p.Wo > 0, "dcn_backward: empty output (%d x %d)", p.Ho, p.Wo);
  const int T = kh * kw, M = p.Ho * p.Wo, Cog = Cout / group;
  p.KC = std::max(1, BW_NK / T);
  p.seg = gcd_i(Cin / group, Cin / deformable_group);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  SF_CHECK_ARG(B <= 65535, "dcn_backward: batch too large");

  DcnTcParams tq{};
  tq.in = input; tq.bias = nullptr; tq.offset = offset; tq.mask = mask; tq.out = nullptr;
  tq.B = B; tq.Cin = Cin; tq.H = H; tq.W = W; tq.Cout = Cout; tq.Ho = p.Ho; tq.Wo = p.Wo;
  tq.kh = kh; tq.kw = kw; tq.sh = stride_h; tq.sw = stride_w; tq.ph = pad_h; tq.pw = pad_w; tq.dh = dil_h; tq.dw = dil_w;
  tq.G = group; tq.dg = deformable_group;
  bool d_done = false;
  if ((grad_input || grad_offset || grad_mask) && algo != SF_ALGO_GENERIC) {
    // channels-last path (dcn_bwd_data_tc.cu): column gradient on the tensor cores, vector reductions into grad_input
    const int rc = launch_dcn_bwd_data_tc(tq, weight, grad_out, grad_input, grad_offset, grad_mask, workspace,
                                          workspace_bytes, s, &d_done);
    if (rc) return rc;
  }
  if (!d_done && grad_input) SF_CUDA(cudaMemsetAsync(grad_input, 0, sizeof(float) * static_cast<size_t>(B) * Cin * H * W, s));
  if (!d_done && (grad_input || grad_offset || grad_mask)) {
    const size_t smem = sizeof(float) * (static_cast<size_t>(BW_OB) * BW_TP + BW_OB * (BW_NK + 4) + BW_NK * BW_TP) +
                        static_cast<size_t>(T) * BW_TP * (sizeof(Desc) + 3 * sizeof(float));
    SF_CUDA(cudaFuncSetAttribute(dcn_bwd_data_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    dim3 grid(ceil_div(M, BW_TP), B);
    dcn_bwd_data_kernel<<<grid, BW_NT, smem, s>>>(p);
    SF_LAUNCH_CHECK();
  }
  bool w_done = false;
  if (grad_weight && algo != SF_ALGO_GENERIC) {
    // weight gradient on the tensor cores (dcn_bwd_tc.cu) where the forward's channels-last kernel applies and the caller
    // brought the workspace for the input copy; the data / offset / mask gradients stay on the FFMA kernel above
    SF_CUDA(cudaMemsetAsync(grad_weight, 0, sizeof(float) * static_cast<size_t>(Cout) * (Cin / group) * T, s));
    const int rc = launch_dcn_bwd_weight_tc(tq, grad_out, grad_weight, workspace, workspace_bytes, s, d_done, &w_done);
    if (rc) return rc;
    SF_CHECK_ARG(w_done || algo != SF_ALGO_TC, "dcn_backward: the tensor-core weight gradient needs %zu workspace bytes "
                 "(got %zu), 32 or a multiple of 64 channels per group and an output width that is a multiple of 4",
                 dcn_bwd_tc_workspace_bytes(B, Cin, H, W, Cout, kh, kw, group, deformable_group), workspace_bytes);
  }
  if (grad_weight && !w_done) {
    SF_CUDA(cudaMemsetAsync(grad_weight, 0, sizeof(float) * static_cast<size_t>(Cout) * (Cin / group) * T, s));
    const size_t smem = sizeof(float) * static_cast<size_t>(BW_TP) * (BW_OB + 4 + BW_NK + 4) + static_cast<size_t>(T) * BW_TP * sizeof(Desc);
    SF_CUDA(cudaFuncSetAttribute(dcn_bwd_weight_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    const int nchunks = (Cin / p.seg) * ceil_div(p.seg, p.KC), oblocks = ceil_div(Cog, BW_OB);
    const long long ntiles = 1LL * B * ceil_div(M, BW_TP);
    // pixel ranges: enough CTAs for ~4 waves, never more ranges than tiles
    long long split = std::max(1LL, (4LL * sm_count() + 1LL * nchunks * oblocks - 1) / (1LL * nchunks * oblocks));
    split = std::min(std::min(split, ntiles), 65535LL);
    p.nsplit = static_cast<int>(split);
    SF_CHECK_ARG(oblocks <= 65535, "dcn_backward: too many output-channel blocks");
    dim3 grid(nchunks, p.nsplit, oblocks);
    dcn_bwd_weight_kernel<<<grid, BW_NTW, smem, s>>>(p);
    SF_LAUNCH_CHECK();
  }
  if (grad_bias) {
    dcn_bwd_bias_kernel<<<Cout, 1024, 0, s>>>(grad_out, grad_bias, B, Cout, M);
    SF_LAUNCH_CHECK();
  }
  return SF_OK;
}

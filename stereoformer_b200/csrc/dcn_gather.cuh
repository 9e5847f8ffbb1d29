// Channels-last gather of the deformable-convolution column tile, shared by the forward (dcn_tc2.cu) and the weight
// gradient (dcn_bwd_tc.cu).  A pixel tile is 8 output rows x 16 columns; gather warp gw (0..15) owns rows
// 8*gw .. 8*gw+7 of the [128 pixels][64 K] tile.  See dcn_tc2.cu for the layout and the measurements behind it.
#pragma once
#include <cuda_bf16.h>

#include "dcn_tc.cuh"

namespace sf {

__device__ __forceinline__ float4 d2g_ldg4(const char* p) { return __ldg(reinterpret_cast<const float4*>(p)); }

// Sampling descriptors (modulated_deform_im2col_cuda.cuh:25-54,163-180) of warp gw's 8 pixels of tile (ty, tx) of image b,
// all taps and deformable groups: d_w = corner weights x mask and either (COMPACT) d_base = top-left corner element of the
// channels-last input | dx | dy << 1, or d_off = the four corner elements.  Rows / columns are clamped: corners outside the
// image carry weight 0 and only need a valid address.
template <bool COMPACT>
__device__ __forceinline__ void d2g_build_descriptors(const DcnTcParams& p, int b, int ty, int tx, int gw, int lane,
                                                      int* d_base, int4* d_off, float4* d_w) {
  const int T = p.kh * p.kw, M = p.Ho * p.Wo;
  const int yo = ty * 8 + (gw >> 1), xo0 = tx * 16 + 8 * (gw & 1);
  for (int e = lane; e < p.dg * T * 8; e += 32) {
    const int ex = e & 7, r = e >> 3;
    const int t = r % T, dgi = r / T;
    const int xo = xo0 + ex;
    const int pix = yo * p.Wo + xo;
    int base = 0;
    int4 off = make_int4(0, 0, 0, 0);
    float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
    if (yo < p.Ho && xo < p.Wo) {
      const int i = t / p.kw, j = t % p.kw;
      const size_t ob = ((static_cast<size_t>(b) * p.dg + dgi) * 2 * T) * M + pix;
      const float oh = __ldg(p.offset + ob + static_cast<size_t>(2 * t) * M);
      const float ow = __ldg(p.offset + ob + static_cast<size_t>(2 * t + 1) * M);
      const float m = p.mask ? __ldg(p.mask + ((static_cast<size_t>(b) * p.dg + dgi) * T + t) * M + pix) : 1.f;
      const float h_im = (float)(yo * p.sh - p.ph + i * p.dh) + oh;
      const float w_im = (float)(xo * p.sw - p.pw + j * p.dw) + ow;
      if (h_im > -1.f && w_im > -1.f && h_im < (float)p.H && w_im < (float)p.W) {
        const int hl = (int)floorf(h_im), wl = (int)floorf(w_im);
        const int hh_i = hl + 1, wh_i = wl + 1;
        const float lh = h_im - (float)hl, lw = w_im - (float)wl;
        const float hh = 1.f - lh, hw2 = 1.f - lw;
        const bool t0 = hl >= 0, l0 = wl >= 0, b1 = hh_i <= p.H - 1, r1 = wh_i <= p.W - 1;
        if (t0 && l0) w.x = hh * hw2 * m;
        if (t0 && r1) w.y = hh * lw * m;
        if (b1 && l0) w.z = lh * hw2 * m;
        if (b1 && r1) w.w = lh * lw * m;
        const int r0 = max(hl, 0), c0 = max(wl, 0);
        const int dy = min(hh_i, p.H - 1) - r0, dx = min(wh_i, p.W - 1) - c0;
        // BYTE offsets into the channels-last image (unsigned 32-bit: the plan keeps an image below 4 GB), so that a
        // gather address is one 64-bit add of a zero-extended register (element indices cost four instructions per load)
        const unsigned e00 = static_cast<unsigned>((r0 * p.W + c0) * p.Cin) * 4u;
        base = static_cast<int>(e00 | dx | (dy << 1));
        off.x = static_cast<int>(e00);
        off.y = static_cast<int>(e00 + dx * p.Cin * 4u);
        off.z = static_cast<int>(e00 + dy * p.W * p.Cin * 4u);
        off.w = static_cast<int>(static_cast<unsigned>(off.z) + dx * p.Cin * 4u);
      }
    }
    const int di = (dgi * T + t) * 128 + 8 * gw + ex;
    if (COMPACT) d_base[di] = base;
    else d_off[di] = off;
    d_w[di] = w;
  }
}

// One chunk (channel block cc, tap group tg) of the column tile: interpolated values x mask, split into bf16 hi + mid and
// stored K-major SWIZZLE_128B at a_hi (hi) and a_hi + 16384 (mid).  in_b = channels-last image b + g * Cig + this lane's
// channel offset cl; dgi = deformable group of this lane's channels in block cc.
template <bool COMPACT>
__device__ __forceinline__ void d2g_gather_chunk(const DcnTcParams& p, const float* in_b, int cc, int tg, int dgi, int gw,
                                                 int lane, const int* d_base, const int4* d_off, const float4* d_w,
                                                 unsigned char* a_hi) {
  const int T = p.kh * p.kw;
  const int hw = lane >> 4, lpt = p.CB >> 2;
  const int t = tg * (64 / p.CB) + (lane & 15) / lpt;
  const bool tlive = t < T;
  const unsigned rowstep = static_cast<unsigned>(p.W) * p.Cin * 4u, colstep = static_cast<unsigned>(p.Cin) * 4u;   // bytes
  const char* src = reinterpret_cast<const char*>(in_b + cc * 64);
  const int dbase = (dgi * T + min(t, T - 1)) * 128 + 8 * gw + hw;
#pragma unroll
  for (int i2 = 0; i2 < 4; i2 += 2) {
    float4 wt[2], v[2][4];
#pragma unroll
    for (int e = 0; e < 2; ++e) wt[e] = tlive ? d_w[dbase + 2 * (i2 + e)] : make_float4(0.f, 0.f, 0.f, 0.f);
    if (COMPACT) {
      int bs[2];
#pragma unroll
      for (int e = 0; e < 2; ++e) bs[e] = d_base[dbase + 2 * (i2 + e)];
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const unsigned c00 = static_cast<unsigned>(bs[e]) & ~3u;
        const unsigned dxo = (bs[e] & 1) ? colstep : 0u, dyo = (bs[e] & 2) ? rowstep : 0u;
        v[e][0] = d2g_ldg4(src + c00); v[e][1] = d2g_ldg4(src + (c00 + dxo));
        v[e][2] = d2g_ldg4(src + (c00 + dyo)); v[e][3] = d2g_ldg4(src + (c00 + dyo + dxo));
      }
    } else {
      int4 off[2];
#pragma unroll
      for (int e = 0; e < 2; ++e) off[e] = d_off[dbase + 2 * (i2 + e)];
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        v[e][0] = d2g_ldg4(src + static_cast<unsigned>(off[e].x)); v[e][1] = d2g_ldg4(src + static_cast<unsigned>(off[e].y));
        v[e][2] = d2g_ldg4(src + static_cast<unsigned>(off[e].z)); v[e][3] = d2g_ldg4(src + static_cast<unsigned>(off[e].w));
      }
    }
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      float4 r;
      r.x = wt[e].x * v[e][0].x + wt[e].y * v[e][1].x + wt[e].z * v[e][2].x + wt[e].w * v[e][3].x;
      r.y = wt[e].x * v[e][0].y + wt[e].y * v[e][1].y + wt[e].z * v[e][2].y + wt[e].w * v[e][3].y;
      r.z = wt[e].x * v[e][0].z + wt[e].y * v[e][1].z + wt[e].z * v[e][2].z + wt[e].w * v[e][3].z;
      r.w = wt[e].x * v[e][0].w + wt[e].y * v[e][1].w + wt[e].z * v[e][2].w + wt[e].w * v[e][3].w;
      // hi = RN(x) to bf16, mid = RN(x - hi): the packed pair is widened back with one shift / one mask per value (the
      // __low2float / __high2float intrinsics compile to an extract + shift + re-pack: 3 more instructions per pair)
      const __nv_bfloat162 h0 = __floats2bfloat162_rn(r.x, r.y), h1 = __floats2bfloat162_rn(r.z, r.w);
      const uint32_t u0 = *reinterpret_cast<const uint32_t*>(&h0), u1 = *reinterpret_cast<const uint32_t*>(&h1);
      const __nv_bfloat162 m0 = __floats2bfloat162_rn(r.x - __uint_as_float(u0 << 16), r.y - __uint_as_float(u0 & 0xffff0000u));
      const __nv_bfloat162 m1 = __floats2bfloat162_rn(r.z - __uint_as_float(u1 << 16), r.w - __uint_as_float(u1 & 0xffff0000u));
      const int row = 8 * gw + 2 * (i2 + e) + hw;
      const int kc = (lane & 15) >> 1;
      unsigned char* dst = a_hi + (row >> 3) * 1024 + (row & 7) * 128 + ((kc ^ (row & 7)) << 4) + (lane & 1) * 8;
      *reinterpret_cast<uint2*>(dst) = make_uint2(u0, u1);
      *reinterpret_cast<uint2*>(dst + 16384) =
          make_uint2(*reinterpret_cast<const uint32_t*>(&m0), *reinterpret_cast<const uint32_t*>(&m1));
    }
  }
}

}  // namespace sf

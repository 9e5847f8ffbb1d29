// Convex upsampling  (the step right after the hot path).
//
// Reference: models/utils/upsample.py:22-38 (upsample_convex8); RAFT-Stereo's upsample_flow uses factor 4.
//   out[n,0,F*y+i,F*x+j] = sum_k softmax_k(mask[n,k*F*F+i*F+j,y,x]) * F * disp[n,0,y+ky-1,x+kx-1]   (zero padded)
// One thread per (low-resolution pixel, output sub-row): mask reads are coalesced along x and streamed
// (read once), a warp writes 32*F consecutive floats of one output row.  Algorithmic bytes per low-resolution
// pixel: (9*F*F + 1 + F*F) * 4.
#include "common.cuh"
#include "../../include/stereoformer_b200.h"

namespace sf {

// Row-band + multi-GPU form (SURVEY 8e partition 2): disp / mask hold H source rows (band + halo), only the low-
// resolution rows [y_first, y_first + rows) are computed, out is [N,1,rows*F,W*F].  The gather of the band's
// full-resolution rows is fused into the store: every result also goes to each rank's whole-map buffer through its
// NVLink peer mapping, element peers.off + b*peers.bstride + (row-major index inside the band).
struct UpPeers {
  float* p[SF_MAX_PEERS];
  int n;
  long long off, bstride;
};

template <int F, bool GATHER>
__global__ void __launch_bounds__(256)
upsample_convex_kernel(const float* __restrict__ disp, const float* __restrict__ mask, float* __restrict__ out, int N,
                       int H, int W, float mask_scale, int y_first, int rows, UpPeers peers) {
  // one thread per (low-resolution pixel, output sub-row i): F x 9 independent coalesced mask loads in flight per
  // thread, and a warp (32 consecutive x, same i) writes 32*F consecutive floats of one output row
  const size_t HW = static_cast<size_t>(H) * W, n = static_cast<size_t>(N) * rows * W * F;
  for (size_t idx = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; idx < n;
       idx += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int x = idx % W;
    size_t r = idx / W;
    const int i = r % F; r /= F;
    const int yl = r % rows, y = y_first + yl;
    const size_t b = r / rows;
    const size_t pix = static_cast<size_t>(y) * W + x;
    float nb[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      const int yy = y + k / 3 - 1, xx = x + k % 3 - 1;
      nb[k] = (yy >= 0 && yy < H && xx >= 0 && xx < W) ? __fmul_rn((float)F, __ldg(disp + b * HW + yy * W + xx)) : 0.f;
    }
    const float* mp = mask + b * 9 * F * F * HW + static_cast<size_t>(i * F) * HW + pix;
    float mv[F][9];
#pragma unroll
    for (int j = 0; j < F; ++j)
#pragma unroll
      for (int k = 0; k < 9; ++k) mv[j][k] = ldg_stream_f32(mp + static_cast<size_t>(k * F * F + j) * HW);
    if (mask_scale != 1.0f) {                       // `.25 * self.mask(hidden)` of update.py:273 folded into the load
#pragma unroll
      for (int j = 0; j < F; ++j)
#pragma unroll
        for (int k = 0; k < 9; ++k) mv[j][k] = __fmul_rn(mask_scale, mv[j][k]);
    }
    float res[F];
#pragma unroll
    for (int j = 0; j < F; ++j) {
      float m = mv[j][0];
#pragma unroll
      for (int k = 1; k < 9; ++k) m = fmaxf(m, mv[j][k]);
      float s = 0.f, a = 0.f;
#pragma unroll
      for (int k = 0; k < 9; ++k) {
        const float e = __expf(mv[j][k] - m);
        s += e; a = fmaf(e, nb[k], a);
      }
      res[j] = a / s;
    }
    const size_t local = (static_cast<size_t>(yl) * F + i) * (W * F) + x * F;
    float* row = out + b * F * F * rows * W + local;
    if (F % 4 == 0) {
#pragma unroll
      for (int j = 0; j < F; j += 4) stg_cs_f4(row + j, make_float4(res[j], res[j + 1], res[j + 2], res[j + 3]));
    } else {
#pragma unroll
      for (int j = 0; j < F; ++j) row[j] = res[j];
    }
    for (int q = 0; GATHER && q < peers.n; ++q) {
      float* prow = peers.p[q] + peers.off + static_cast<long long>(b) * peers.bstride + local;
#pragma unroll
      for (int j = 0; j < F; j += 4)
        *reinterpret_cast<float4*>(prow + j) = make_float4(res[j], res[j + 1], res[j + 2], res[j + 3]);
    }
  }
}

// backward: out[F*y+i, F*x+j] = sum_k p_k * nb_k with p = softmax_k(scale * mask[k*F*F+i*F+j]), nb_k = F * disp(neighbour k)
//   g_mask[k] = scale * g * p_k * (nb_k - out);   g_disp(neighbour k) += F * g * p_k   (atomics: 9 F^2 terms per pixel)
template <int F>
__global__ void __launch_bounds__(256)
upsample_convex_bwd_kernel(const float* __restrict__ disp, const float* __restrict__ mask, const float* __restrict__ gout,
                           float* __restrict__ gdisp, float* __restrict__ gmask, int N, int H, int W, float mask_scale) {
  const size_t HW = static_cast<size_t>(H) * W, n = static_cast<size_t>(N) * HW * F;
  for (size_t idx = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; idx < n;
       idx += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int x = idx % W;
    size_t r = idx / W;
    const int i = r % F; r /= F;
    const int y = r % H;
    const size_t b = r / H;
    const size_t pix = static_cast<size_t>(y) * W + x;
    float nb[9], gnb[9];
    bool in[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      const int yy = y + k / 3 - 1, xx = x + k % 3 - 1;
      in[k] = yy >= 0 && yy < H && xx >= 0 && xx < W;
      nb[k] = in[k] ? (float)F * __ldg(disp + b * HW + yy * W + xx) : 0.f;
      gnb[k] = 0.f;
    }
    const float* mp = mask + b * 9 * F * F * HW + static_cast<size_t>(i * F) * HW + pix;
    float* gm = gmask ? gmask + b * 9 * F * F * HW + static_cast<size_t>(i * F) * HW + pix : nullptr;
    const float* grow = gout + b * F * F * HW + (static_cast<size_t>(y) * F + i) * (W * F) + x * F;
    for (int j = 0; j < F; ++j) {
      float mv[9], m = -INFINITY;
#pragma unroll
      for (int k = 0; k < 9; ++k) { mv[k] = mask_scale * __ldg(mp + static_cast<size_t>(k * F * F + j) * HW); m = fmaxf(m, mv[k]); }
      float s = 0.f, a = 0.f;
#pragma unroll
      for (int k = 0; k < 9; ++k) { mv[k] = __expf(mv[k] - m); s += mv[k]; a = fmaf(mv[k], nb[k], a); }
      const float inv = 1.0f / s, o = a * inv, g = grow[j];
#pragma unroll
      for (int k = 0; k < 9; ++k) {
        const float pk = mv[k] * inv;
        if (gm) gm[static_cast<size_t>(k * F * F + j) * HW] = mask_scale * g * pk * (nb[k] - o);
        gnb[k] = fmaf(g, pk, gnb[k]);
      }
    }
    if (gdisp) {
#pragma unroll
      for (int k = 0; k < 9; ++k)
        if (in[k]) atomicAdd(gdisp + b * HW + (y + k / 3 - 1) * W + (x + k % 3 - 1), (float)F * gnb[k]);
    }
  }
}

// upsample_simple8 (models/utils/upsample.py:40-43): F * F.interpolate(disp, size = F*(H, W), mode='bilinear',
// align_corners=True).  ATen's arithmetic (UpSample.h area_pixel_compute_scale / guard_index_and_lambda,
// UpSampleKernel.cpp interpolate<2>): scale = (in-1)/(out-1) in fp32, src = scale*dst, i0 = min(floor(src), in-1),
// l1 = clamp(src - i0, 0, 1), l0 = 1 - l1, i1 = i0 + (i0 < in-1); out = lh0*(lw0*a + lw1*b) + lh1*(lw0*c + lw1*d),
// evaluated without FMA contraction.  One thread per 4 consecutive output pixels (one float4 store); the 4 source
// values a thread needs are L1/L2 hits (the input is F*F times smaller than the output): write-bound.
struct Lerp1 { int i0, i1; float l0, l1; };

__device__ __forceinline__ Lerp1 simple_index(float scale, int dst, int in) {
  Lerp1 r;
  const float src = __fmul_rn(scale, (float)dst);
  r.i0 = min((int)floorf(src), in - 1);
  r.l1 = fminf(fmaxf(__fsub_rn(src, (float)r.i0), 0.f), 1.f);
  r.l0 = __fsub_rn(1.f, r.l1);
  r.i1 = r.i0 + (r.i0 < in - 1 ? 1 : 0);
  return r;
}

__global__ void __launch_bounds__(128)
upsample_simple_kernel(const float* __restrict__ disp, float* __restrict__ out, int H, int W, int F, float sh,
                       float sw) {
  // blockIdx.z = image, blockIdx.y = float4-column chunk, blockIdx.x strides the image's output rows.  A thread keeps
  // its float4 column: for F >= 4 its 4 outputs read at most 3 consecutive source cells (4 * scale < 1), so a row costs
  // 6 loads (3 cells x 2 source rows) and the x weights / cell selectors are computed once.
  const int Ho = H * F, Wo = W * F, Wq = Wo >> 2;
  const int xq = blockIdx.y * blockDim.x + threadIdx.x;
  if (xq >= Wq) return;
  Lerp1 lx[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) lx[j] = simple_index(sw, 4 * xq + j, W);
  const int c0 = lx[0].i0, c1 = min(c0 + 1, W - 1), c2 = min(c0 + 2, W - 1);
  const bool compact = lx[3].i1 <= c0 + 2;               // always true for F >= 4
  const float* img = disp + static_cast<size_t>(blockIdx.z) * H * W;
  float* oimg = out + static_cast<size_t>(blockIdx.z) * Ho * Wo + 4 * xq;
  for (int yo = blockIdx.x; yo < Ho; yo += gridDim.x) {
    const Lerp1 ly = simple_index(sh, yo, H);
    const float* r0 = img + ly.i0 * W;
    const float* r1 = img + ly.i1 * W;
    float res[4];
    if (compact) {
      const float t0 = __ldg(r0 + c0), t1 = __ldg(r0 + c1), t2 = __ldg(r0 + c2);
      const float b0 = __ldg(r1 + c0), b1 = __ldg(r1 + c1), b2 = __ldg(r1 + c2);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int d0 = lx[j].i0 - c0, d1 = lx[j].i1 - c0;
        const float ta = d0 == 0 ? t0 : t1, tb = d1 == 0 ? t0 : (d1 == 1 ? t1 : t2);
        const float ba = d0 == 0 ? b0 : b1, bb = d1 == 0 ? b0 : (d1 == 1 ? b1 : b2);
        const float top = __fadd_rn(__fmul_rn(lx[j].l0, ta), __fmul_rn(lx[j].l1, tb));
        const float bot = __fadd_rn(__fmul_rn(lx[j].l0, ba), __fmul_rn(lx[j].l1, bb));
        res[j] = __fmul_rn((float)F, __fadd_rn(__fmul_rn(ly.l0, top), __fmul_rn(ly.l1, bot)));
      }
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float top = __fadd_rn(__fmul_rn(lx[j].l0, __ldg(r0 + lx[j].i0)), __fmul_rn(lx[j].l1, __ldg(r0 + lx[j].i1)));
        const float bot = __fadd_rn(__fmul_rn(lx[j].l0, __ldg(r1 + lx[j].i0)), __fmul_rn(lx[j].l1, __ldg(r1 + lx[j].i1)));
        res[j] = __fmul_rn((float)F, __fadd_rn(__fmul_rn(ly.l0, top), __fmul_rn(ly.l1, bot)));
      }
    }
    stg_cs_f4(oimg + static_cast<size_t>(yo) * Wo, make_float4(res[0], res[1], res[2], res[3]));
  }
}

// backward of the above: every output pixel adds F * g * (lh, lw weights) to its <= 4 sources.  Gather form: one thread
// per LOW-resolution pixel sums the output pixels whose footprint touches it (rows/columns within F of it), so the
// result is deterministic and needs no atomics.
__global__ void __launch_bounds__(256)
upsample_simple_bwd_kernel(const float* __restrict__ gout, float* __restrict__ gdisp, int N, int H, int W, int F,
                           float sh, float sw) {
  const int Ho = H * F, Wo = W * F;
  const size_t n = static_cast<size_t>(N) * H * W;
  for (size_t idx = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; idx < n;
       idx += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int x = idx % W, y = (idx / W) % H;
    const size_t b = idx / (static_cast<size_t>(H) * W);
    // output rows whose i0 or i1 can equal y: src in (y-1, y+1)  ->  dst in ((y-1)/sh, (y+1)/sh); scanned generously
    const int yo0 = max(0, (int)floorf((float)(y - 1) / sh) - 1), yo1 = min(Ho - 1, (int)ceilf((float)(y + 1) / sh) + 1);
    const int xo0 = max(0, (int)floorf((float)(x - 1) / sw) - 1), xo1 = min(Wo - 1, (int)ceilf((float)(x + 1) / sw) + 1);
    constexpr int SPAN = 24;                       // column weights once per pixel (2F + 5 <= 21 candidates at F = 8)
    float wxs[SPAN];
    const bool cached = xo1 - xo0 < SPAN;
    if (cached) {
#pragma unroll
      for (int k = 0; k < SPAN; ++k) {
        const int xo = min(xo0 + k, xo1);
        const Lerp1 lx = simple_index(sw, xo, W);
        float wx = 0.f;
        if (lx.i0 == x) wx += lx.l0;
        if (lx.i1 == x) wx += lx.l1;
        wxs[k] = xo0 + k <= xo1 ? wx : 0.f;
      }
    }
    float acc = 0.f;
    for (int yo = yo0; yo <= yo1; ++yo) {
      const Lerp1 ly = simple_index(sh, yo, H);
      float wy = 0.f;
      if (ly.i0 == y) wy += ly.l0;
      if (ly.i1 == y) wy += ly.l1;
      if (wy == 0.f) continue;
      const float* grow = gout + (b * Ho + yo) * Wo;
      float racc = 0.f;
      if (cached) {
#pragma unroll
        for (int k = 0; k < SPAN; ++k)
          if (wxs[k] != 0.f) racc = fmaf(wxs[k], __ldg(grow + xo0 + k), racc);
      } else {
        for (int xo = xo0; xo <= xo1; ++xo) {
          const Lerp1 lx = simple_index(sw, xo, W);
          float wx = 0.f;
          if (lx.i0 == x) wx += lx.l0;
          if (lx.i1 == x) wx += lx.l1;
          if (wx != 0.f) racc = fmaf(wx, __ldg(grow + xo), racc);
        }
      }
      acc = fmaf(wy, racc, acc);
    }
    gdisp[idx] = (float)F * acc;
  }
}

}  // namespace sf

static int upsample_convex_impl(const float* disp, const float* mask, float mask_scale, float* out, int N, int H, int W,
                                int factor, void* stream, int y_first = 0, int rows = -1,
                                float* const* h_peers = nullptr, int n_peers = 0, long long peer_offset = 0,
                                long long peer_batch_stride = 0) {
  using namespace sf;
  if (rows < 0) rows = H;
  SF_CHECK_ARG(disp && mask && out, "upsample_convex: null pointer");
  SF_CHECK_ARG(N > 0 && H > 0 && W > 0, "upsample_convex: non-positive dimension");
  SF_CHECK_ARG(factor == 8 || factor == 4, "upsample_convex: factor must be 8 or 4, got %d", factor);
  SF_CHECK_ARG((reinterpret_cast<uintptr_t>(out) & 15) == 0, "upsample_convex: out must be 16-byte aligned");
  SF_CHECK_ARG(y_first >= 0 && rows > 0 && y_first + rows <= H, "upsample_convex: rows [%d, %d) outside the %d source rows",
               y_first, y_first + rows, H);
  SF_CHECK_ARG(n_peers >= 0 && n_peers <= SF_MAX_PEERS && (n_peers == 0 || h_peers) && peer_offset >= 0 &&
                   peer_batch_stride >= 0, "upsample_convex: bad peer list (at most %d peers)", SF_MAX_PEERS);
  UpPeers peers;
  peers.n = n_peers;
  peers.off = peer_offset;
  peers.bstride = peer_batch_stride;
  for (int r = 0; r < SF_MAX_PEERS; ++r) peers.p[r] = r < n_peers ? h_peers[r] : nullptr;
  for (int r = 0; r < n_peers; ++r) {
    SF_CHECK_ARG(peers.p[r] != nullptr, "upsample_convex: null peer pointer %d", r);
    SF_CHECK_ARG(((reinterpret_cast<uintptr_t>(peers.p[r]) + 4 * static_cast<uintptr_t>(peer_offset)) & 15) == 0 &&
                     (peer_batch_stride & 3) == 0, "upsample_convex: peer buffers must be 16-byte aligned");
  }
  size_t n = static_cast<size_t>(N) * rows * W * factor;
  int grid = static_cast<int>(min(static_cast<size_t>(sm_count()) * 8, (n + 255) / 256));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
#define SF_UP_LAUNCH(F, G) \
  upsample_convex_kernel<F, G><<<grid, 256, 0, s>>>(disp, mask, out, N, H, W, mask_scale, y_first, rows, peers)
  if (factor == 8) { if (n_peers) SF_UP_LAUNCH(8, true); else SF_UP_LAUNCH(8, false); }
  else { if (n_peers) SF_UP_LAUNCH(4, true); else SF_UP_LAUNCH(4, false); }
#undef SF_UP_LAUNCH
  SF_LAUNCH_CHECK();
  return SF_OK;
}

extern "C" int sf_upsample_convex_gather_f32(const float* disp, const float* mask, float mask_scale, float* out, int N,
                                             int H, int W, int factor, int y_first, int rows, float* const* h_peers,
                                             int n_peers, long long peer_offset, long long peer_batch_stride,
                                             void* stream) {
  return upsample_convex_impl(disp, mask, mask_scale, out, N, H, W, factor, stream, y_first, rows, h_peers, n_peers,
                              peer_offset, peer_batch_stride);
}

extern "C" int sf_upsample_convex_f32(const float* disp, const float* mask, float* out, int N, int H, int W,
                                      int factor, void* stream) {
  return upsample_convex_impl(disp, mask, 1.0f, out, N, H, W, factor, stream);
}

extern "C" int sf_upsample_convex_scaled_f32(const float* disp, const float* mask, float mask_scale, float* out, int N,
                                             int H, int W, int factor, void* stream) {
  return upsample_convex_impl(disp, mask, mask_scale, out, N, H, W, factor, stream);
}

extern "C" int sf_upsample_convex_backward_f32(const float* disp, const float* mask, float mask_scale,
                                               const float* grad_out, float* grad_disp, float* grad_mask, int N, int H,
                                               int W, int factor, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(disp && mask && grad_out, "upsample_convex_backward: null pointer");
  SF_CHECK_ARG(N > 0 && H > 0 && W > 0, "upsample_convex_backward: non-positive dimension");
  SF_CHECK_ARG(factor == 8 || factor == 4, "upsample_convex_backward: factor must be 8 or 4, got %d", factor);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (grad_disp) SF_CUDA(cudaMemsetAsync(grad_disp, 0, sizeof(float) * static_cast<size_t>(N) * H * W, s));
  size_t n = static_cast<size_t>(N) * H * W * factor;
  int grid = static_cast<int>(min(static_cast<size_t>(sm_count()) * 8, (n + 255) / 256));
  if (factor == 8) upsample_convex_bwd_kernel<8><<<grid, 256, 0, s>>>(disp, mask, grad_out, grad_disp, grad_mask, N, H, W, mask_scale);
  else upsample_convex_bwd_kernel<4><<<grid, 256, 0, s>>>(disp, mask, grad_out, grad_disp, grad_mask, N, H, W, mask_scale);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

static int simple_scales(int H, int W, int factor, float* sh, float* sw) {
  SF_CHECK_ARG(H > 0 && W > 0 && factor >= 1 && (W * factor) % 4 == 0,
               "upsample_simple: need positive sizes and an output width that is a multiple of 4");
  // area_pixel_compute_scale<float>(in, out, align_corners=true): (in-1)/(out-1), 0 when out == 1
  *sh = H * factor > 1 ? static_cast<float>(H - 1) / static_cast<float>(H * factor - 1) : 0.f;
  *sw = W * factor > 1 ? static_cast<float>(W - 1) / static_cast<float>(W * factor - 1) : 0.f;
  return SF_OK;
}

extern "C" int sf_upsample_simple_f32(const float* disp, float* out, int N, int H, int W, int factor, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(disp && out && N > 0, "upsample_simple: null pointer / empty batch");
  SF_CHECK_ARG((reinterpret_cast<uintptr_t>(out) & 15) == 0, "upsample_simple: out must be 16-byte aligned");
  float sh, sw;
  if (int rc = simple_scales(H, W, factor, &sh, &sw)) return rc;
  SF_CHECK_ARG(N <= 65535, "upsample_simple: more than 65535 images per call");
  const int ychunks = (W * factor / 4 + 127) / 128;
  SF_CHECK_ARG(ychunks <= 65535, "upsample_simple: output too wide");
  const int per_img = max(1, min(H * factor, sm_count() * 16 / (ychunks * N) + 1));   // ~16 CTAs per SM in all
  dim3 grid(per_img, ychunks, N);
  upsample_simple_kernel<<<grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(disp, out, H, W, factor, sh, sw);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

extern "C" int sf_upsample_simple_backward_f32(const float* grad_out, float* grad_disp, int N, int H, int W, int factor,
                                               void* stream) {
  using namespace sf;
  SF_CHECK_ARG(grad_out && grad_disp && N > 0, "upsample_simple_backward: null pointer / empty batch");
  SF_CHECK_ARG(H > 1 && W > 1, "upsample_simple_backward: needs H, W > 1");
  float sh, sw;
  if (int rc = simple_scales(H, W, factor, &sh, &sw)) return rc;
  size_t n = static_cast<size_t>(N) * H * W;
  int grid = static_cast<int>(min(static_cast<size_t>(sm_count()) * 8, (n + 255) / 256));
  upsample_simple_bwd_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(grad_out, grad_disp, N, H, W, factor,
                                                                                 sh, sw);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

// Deformable convolution forward, DCNv2 (modulated) and DCNv1 (mask == null)  (SURVEY 8a row a10).
//
// Reference: deform/src/cuda/modulated_deform_conv_cuda.cu:19-121 (host), modulated_deform_im2col_cuda.cuh:25-54,
// 128-194 (bilinear + im2col kernel); DCNv1 twin deform_conv_cuda.cu:19-120, deform_im2col_cuda.cuh:127-189.
//
//   out[b,o,y,x] = bias[o] + sum_{c,i,j} W[o,c,i,j] * mask[b,g,ij,y,x] * bilinear(in[b,c], y*s-p+i*dil+dy, x*s-p+j*dil+dx)
//
// The reference materialises a [Cin*kh*kw, M] column buffer in HBM, runs cuBLAS over it and then permutes
// NHWC->NCHW.  Here it is one fused implicit-GEMM kernel (fp32-exact FFMA path): per CTA the sampling
// descriptors (4 corner offsets + 4 bilinear weights + mask) of a pixel tile are computed once per deformable
// group and kept in shared memory, channel chunks of the column tile are gathered straight into shared memory,
// multiplied with a transposed weight tile from registers-tiled FFMAs, and the epilogue adds the bias and
// writes NCHW directly.  No column buffer, no permute, no im2col_step loop (it never affected results:
// deform/test.py:219-260).
#include <cstdlib>
#include "common.cuh"
#include "dcn_tc.cuh"
#include "../../include/stereoformer_b200.h"

namespace sf {

constexpr int DCN_KC = 8;  // input channels per shared-memory chunk
constexpr int DCN_NT = 512; // threads per CTA

struct DcnParams {
  const float* in;      // [B,Cin,H,W]
  const float* weight;  // [Cout,Cin/G,kh,kw]
  const float* bias;    // [Cout]
  const float* offset;  // [B,dg*2*kh*kw,Ho,Wo]
  const float* mask;    // [B,dg*kh*kw,Ho,Wo] or null (DCNv1)
  float* out;           // [B,Cout,Ho,Wo]
  int B, Cin, H, W, Cout, Ho, Wo;
  int kh, kw, sh, sw, ph, pw, dh, dw, G, dg;
};

// NT threads gather (the gathers are latency-bound: the more in flight the better); the first (TO/4)*(TP/4) of them
// also run the register-tiled GEMM.
template <int TO, int TP>
__global__ void __launch_bounds__(DCN_NT)
dcn_forward_kernel(const DcnParams p) {
  constexpr int NT = DCN_NT, NGEMM = (TO / 4) * (TP / 4);
  static_assert(NGEMM <= NT, "GEMM tile needs more threads than the CTA has");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int T = p.kh * p.kw;
  int4* d_off = reinterpret_cast<int4*>(smem_raw);                  // [T][TP]
  float4* d_w = reinterpret_cast<float4*>(d_off + T * TP);          // [T][TP]
  float* d_m = reinterpret_cast<float*>(d_w + T * TP);              // [T][TP]
  float* col = d_m + T * TP;                                        // [KC*T][TP]
  float* wsm = col + DCN_KC * T * TP;                               // [KC*T][TO]

  const int M = p.Ho * p.Wo;
  const int pix0 = blockIdx.x * TP;
  const int b = blockIdx.y;
  const int Cig = p.Cin / p.G, Cog = p.Cout / p.G;
  const int otiles = (Cog + TO - 1) / TO;
  const int g = blockIdx.z / otiles, o0 = (blockIdx.z % otiles) * TO;
  const int cpdg = p.Cin / p.dg;
  const int tid = threadIdx.x;
  const int tx = tid % (TP / 4), ty = tid / (TP / 4);
  const size_t HW = static_cast<size_t>(p.H) * p.W;

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  int cur_dg = -1;
  for (int c0 = 0; c0 < Cig;) {
    const int cin0 = g * Cig + c0;
    const int dgi = cin0 / cpdg;
    const int nkc = min(min(DCN_KC, Cig - c0), (dgi + 1) * cpdg - cin0);
    __syncthreads();  // previous chunk's GEMM has finished reading col / wsm / descriptors
    if (dgi != cur_dg) {
      cur_dg = dgi;
      // sampling descriptors of this deformable group (modulated_deform_im2col_cuda.cuh:25-54,163-180)
      for (int e = tid; e < T * TP; e += NT) {
        const int px = e % TP, t = e / TP;
        const int pix = pix0 + px;
        int4 off = make_int4(0, 0, 0, 0);
        float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
        float m = 0.f;
        if (pix < M) {
          const int yo = pix / p.Wo, xo = pix % p.Wo;
          const int i = t / p.kw, j = t % p.kw;
          const size_t ob = ((static_cast<size_t>(b) * p.dg + dgi) * 2 * T) * M + pix;
          const float oh = __ldg(p.offset + ob + static_cast<size_t>(2 * t) * M);
          const float ow = __ldg(p.offset + ob + static_cast<size_t>(2 * t + 1) * M);
          m = p.mask ? __ldg(p.mask + ((static_cast<size_t>(b) * p.dg + dgi) * T + t) * M + pix) : 1.f;
          const float h_im = (float)(yo * p.sh - p.ph + i * p.dh) + oh;
          const float w_im = (float)(xo * p.sw - p.pw + j * p.dw) + ow;
          if (h_im > -1.f && w_im > -1.f && h_im < (float)p.H && w_im < (float)p.W) {
            const int hl = (int)floorf(h_im), wl = (int)floorf(w_im);
            const int hh_i = hl + 1, wh_i = wl + 1;
            const float lh = h_im - (float)hl, lw = w_im - (float)wl;
            const float hh = 1.f - lh, hw = 1.f - lw;
            const bool t0 = hl >= 0, l0 = wl >= 0, b1 = hh_i <= p.H - 1, r1 = wh_i <= p.W - 1;
            if (t0 && l0) { off.x = hl * p.W + wl; w.x = hh * hw; }
            if (t0 && r1) { off.y = hl * p.W + wh_i; w.y = hh * lw; }
            if (b1 && l0) { off.z = hh_i * p.W + wl; w.z = lh * hw; }
            if (b1 && r1) { off.w = hh_i * p.W + wh_i; w.w = lh * lw; }
          }
        }
        d_off[e] = off; d_w[e] = w; d_m[e] = m;
      }
      __syncthreads();
    }
    // column tile: col[(c*T+t)][px] = bilinear(in[b,cin0+c]) * mask
    for (int e = tid; e < nkc * T * TP; e += NT) {
      const int px = e % TP, kt = e / TP;
      const int c = kt / T, t = kt % T;
      const float* src = p.in + (static_cast<size_t>(b) * p.Cin + cin0 + c) * HW;
      const int4 off = d_off[t * TP + px];
      const float4 w = d_w[t * TP + px];
      const float val = w.x * __ldg(src + off.x) + w.y * __ldg(src + off.y) + w.z * __ldg(src + off.z) +
                        w.w * __ldg(src + off.w);
      col[kt * TP + px] = val * d_m[t * TP + px];
    }
    // weight tile, transposed: wsm[k][o] = W[g*Cog+o0+o][c0+c][t]
    for (int e = tid; e < nkc * T * TO; e += NT) {
      const int k = e % (nkc * T), o = e / (nkc * T);
      float v = 0.f;
      if (o0 + o < Cog) v = __ldg(p.weight + (static_cast<size_t>(g * Cog + o0 + o) * Cig + c0) * T + k);
      wsm[k * TO + o] = v;
    }
    __syncthreads();
    const int nk = nkc * T;
    if (tid < NGEMM)
    for (int k = 0; k < nk; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(col + k * TP + tx * 4);
      const float4 w = *reinterpret_cast<const float4*>(wsm + k * TO + ty * 4);
      const float av[4] = {a.x, a.y, a.z, a.w}, wv[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(wv[i], av[j], acc[i][j]);
    }
    c0 += nkc;
  }

  // epilogue: + bias, NCHW store (addmm(bias, cols^T, W^T) + permute of the reference, .cu:108-118)
  if (tid >= NGEMM) return;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int o = o0 + ty * 4 + i;
    if (o >= Cog) continue;
    const int oc = g * Cog + o;
    const float bv = __ldg(p.bias + oc);
    float* dst = p.out + (static_cast<size_t>(b) * p.Cout + oc) * M + pix0 + tx * 4;
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (pix0 + tx * 4 + j < M) dst[j] = acc[i][j] + bv;
  }
}

template <int TO, int TP>
static int launch_dcn(const DcnParams& p, cudaStream_t s) {
  const int T = p.kh * p.kw;
  const size_t smem = static_cast<size_t>(T) * TP * (16 + 16 + 4) + static_cast<size_t>(DCN_KC) * T * (TP + TO) * 4;
  SF_CHECK_ARG(smem <= 200 * 1024, "dcn_forward: kernel %dx%d too large for the shared-memory tile", p.kh, p.kw);
  auto kern = dcn_forward_kernel<TO, TP>;
  SF_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  const int Cog = p.Cout / p.G;
  dim3 grid(ceil_div(p.Ho * p.Wo, TP), p.B, p.G * ceil_div(Cog, TO));
  SF_CHECK_ARG(grid.y <= 65535 && grid.z <= 65535, "dcn_forward: batch or group count too large");
  kern<<<grid, DCN_NT, smem, s>>>(p);
  SF_LAUNCH_CHECK();
  return SF_OK;
}


// ---------------------------------------------------------------------------------------------- small channel counts
// The shipped model's DCN is 16 -> 16 channels (pure_learning_method.py:52, resnet.py:57): 0.3 GFLOP and 15 MB at
// C2 x 8, but 37 M bilinear corner gathers.  Through L1 a warp-level gather with learned offsets touches ~20 sectors,
// which bounds the generic kernel above at ~60 us.  Here a CTA owns an 8 x 32 pixel tile, one thread per output pixel,
// and first stages the input patch the tile can reach with offsets up to DS_R (8) pixels -- all channels, zero-filled
// outside the image (which IS the reference's per-corner zero padding and its (-1,H)x(-1,W) test, .cuh:37-48,180) --
// in shared memory, so the data-dependent gathers become shared-memory reads.  A corner outside the staged patch
// (|offset| > DS_R) takes the global-memory path with the explicit tests.  Weights are staged transposed
// ([c][t][o], o padded to 16) and read as broadcast float4s.  43 us against 60 us at C2 x 8; ncu (first version): 27 M warp
// instructions, 44 % of them the patch fill's index arithmetic, issue-bound at 1.8 IPC with 16 warps per SM.
// Measured and not kept: two threads per pixel on alternate channels (51 us), two 8-channel passes (64 us), a
// warp-per-row fill (slower than the flat loop), and a TMA box for the patch (one cp.async.bulk.tensor per CTA; it ran
// for 16 input channels but raised an illegal-instruction fault with fewer, cause not found, so it is not shipped).
constexpr int DS_TY = 8, DS_TX = 32, DS_MAXC = 16;
static int ds_halo() {
  static const int r = tuning_env("SF_DCN_R") ? atoi(tuning_env("SF_DCN_R")) : 8;   // sweep hook (tuning build)
  return r;
}

__global__ void __launch_bounds__(DS_TY * DS_TX)
dcn_small_kernel(const DcnParams p, int RH, int RW, int RWp, int DS_R, unsigned rw_magic) {
  extern __shared__ __align__(16) float ds_smem[];
  const int T = p.kh * p.kw, M = p.Ho * p.Wo;
  float* wsm = ds_smem;                                   // [Cin*T][16]
  float* reg = ds_smem + p.Cin * T * DS_MAXC;             // [Cin][RH][RWp]
  const int tid = threadIdx.x, b = blockIdx.z;
  const int ty0 = blockIdx.y * DS_TY, tx0 = blockIdx.x * DS_TX;
  const int RY0 = ty0 * p.sh - p.ph - DS_R, RX0 = tx0 * p.sw - p.pw - DS_R;
  const size_t HW = static_cast<size_t>(p.H) * p.W;

  for (int e = tid; e < p.Cin * T * DS_MAXC; e += DS_TY * DS_TX) {
    const int o = e % DS_MAXC, k = e / DS_MAXC;           // k = c*T + t
    wsm[e] = o < p.Cout ? __ldg(p.weight + static_cast<size_t>(o) * p.Cin * T + k) : 0.f;
  }
  // flat over the RH x RW cells of one channel plane; r / RW by multiply-high (exact: r < 2^16, magic = 2^32/RW + 1)
  const int RHW = RH * RW;
  for (int c = 0; c < p.Cin; ++c) {
    const float* plane = p.in + (static_cast<size_t>(b) * p.Cin + c) * HW;
    float* dstc = reg + c * RH * RWp;
    for (int r = tid; r < RHW; r += DS_TY * DS_TX) {
      const int yy = static_cast<int>(__umulhi(static_cast<unsigned>(r), rw_magic)), xx = r - yy * RW;
      const int gy = RY0 + yy, gx = RX0 + xx;
      float v = 0.f;
      if (gy >= 0 && gy < p.H && gx >= 0 && gx < p.W) v = __ldg(plane + gy * p.W + gx);
      dstc[yy * RWp + xx] = v;
    }
  }
  __syncthreads();

  const int yo = ty0 + tid / DS_TX, xo = tx0 + tid % DS_TX;
  if (yo >= p.Ho || xo >= p.Wo) return;
  const int pix = yo * p.Wo + xo, cpdg = p.Cin / p.dg;
  float acc[DS_MAXC];
#pragma unroll
  for (int o = 0; o < DS_MAXC; ++o) acc[o] = 0.f;

  for (int dgi = 0; dgi < p.dg; ++dgi) {
    const size_t ob = ((static_cast<size_t>(b) * p.dg + dgi) * 2 * T) * M + pix;
    for (int t = 0; t < T; ++t) {
      const int i = t / p.kw, j = t - i * p.kw;
      const float oh = __ldg(p.offset + ob + static_cast<size_t>(2 * t) * M);
      const float ow = __ldg(p.offset + ob + static_cast<size_t>(2 * t + 1) * M);
      const float m = p.mask ? __ldg(p.mask + ((static_cast<size_t>(b) * p.dg + dgi) * T + t) * M + pix) : 1.f;
      const float h_im = (float)(yo * p.sh - p.ph + i * p.dh) + oh;
      const float w_im = (float)(xo * p.sw - p.pw + j * p.dw) + ow;
      const float hlf = floorf(h_im), wlf = floorf(w_im);
      const float lh = h_im - hlf, lw = w_im - wlf, hh = 1.f - lh, hw = 1.f - lw;
      const float w1 = hh * hw, w2 = hh * lw, w3 = lh * hw, w4 = lh * lw;
      const bool staged = hlf >= (float)RY0 && hlf <= (float)(RY0 + RH - 2) && wlf >= (float)RX0 &&
                          wlf <= (float)(RX0 + RW - 2);
      if (staged) {
        const float* q = reg + ((int)hlf - RY0) * RWp + ((int)wlf - RX0);
        for (int c = dgi * cpdg; c < (dgi + 1) * cpdg; ++c) {
          const float* qc = q + c * RH * RWp;
          const float val = (w1 * qc[0] + w2 * qc[1] + w3 * qc[RWp] + w4 * qc[RWp + 1]) * m;
          const float4* wr = reinterpret_cast<const float4*>(wsm + (c * T + t) * DS_MAXC);
#pragma unroll
          for (int o4 = 0; o4 < DS_MAXC / 4; ++o4) {
            const float4 w = wr[o4];
            acc[4 * o4] = fmaf(w.x, val, acc[4 * o4]);
            acc[4 * o4 + 1] = fmaf(w.y, val, acc[4 * o4 + 1]);
            acc[4 * o4 + 2] = fmaf(w.z, val, acc[4 * o4 + 2]);
            acc[4 * o4 + 3] = fmaf(w.w, val, acc[4 * o4 + 3]);
          }
        }
      } else if (h_im > -1.f && w_im > -1.f && h_im < (float)p.H && w_im < (float)p.W) {
        const int hl = (int)hlf, wl = (int)wlf;
        const bool t0 = hl >= 0, l0 = wl >= 0, b1 = hl + 1 <= p.H - 1, r1 = wl + 1 <= p.W - 1;
        for (int c = dgi * cpdg; c < (dgi + 1) * cpdg; ++c) {
          const float* src = p.in + (static_cast<size_t>(b) * p.Cin + c) * HW + static_cast<ptrdiff_t>(hl) * p.W + wl;
          const float v1 = (t0 && l0) ? __ldg(src) : 0.f, v2 = (t0 && r1) ? __ldg(src + 1) : 0.f;
          const float v3 = (b1 && l0) ? __ldg(src + p.W) : 0.f, v4 = (b1 && r1) ? __ldg(src + p.W + 1) : 0.f;
          const float val = (w1 * v1 + w2 * v2 + w3 * v3 + w4 * v4) * m;
          const float* wr = wsm + (c * T + t) * DS_MAXC;
#pragma unroll
          for (int o = 0; o < DS_MAXC; ++o) acc[o] = fmaf(wr[o], val, acc[o]);
        }
      }
    }
  }
#pragma unroll
  for (int o = 0; o < DS_MAXC; ++o)
    if (o < p.Cout) p.out[(static_cast<size_t>(b) * p.Cout + o) * M + pix] = acc[o] + __ldg(p.bias + o);
}

// ---- TMA-staged variant (default when the input rows are 16-byte aligned): the patch [Cin][RH][RWb] is ONE
// cp.async.bulk.tensor box per CTA (out-of-image cells arrive as zeros: the reference's zero padding), which removes the
// fill loop that was 44 % of the instructions of the kernel above; the 16-wide output accumulation runs as packed
// FFMA2 (fma.rn.f32x2: two IEEE fp32 FMAs per instruction, results identical to scalar FFMA) and a thread owns PPT
// output pixels (rows yl and yl + 4), so each broadcast weight load feeds 2 x 16 FMAs.
__device__ __forceinline__ void ffma2(unsigned long long& acc, unsigned long long w, float v) {
  asm("{\n\t.reg .b64 vv;\n\tmov.b64 vv, {%2, %2};\n\tfma.rn.f32x2 %0, %1, vv, %0;\n\t}" : "+l"(acc) : "l"(w), "f"(v));
}

template <int PPT>
__global__ void __launch_bounds__(DS_TY * DS_TX / PPT)
dcn_small_tma_kernel(const __grid_constant__ CUtensorMap mapIn, const DcnParams p, int RH, int RWb, int DS_R) {
  constexpr int NTH = DS_TY * DS_TX / PPT, YSTEP = DS_TY / PPT;
  extern __shared__ __align__(128) unsigned char ds_raw[];
  const int T = p.kh * p.kw, M = p.Ho * p.Wo;
  float* reg = reinterpret_cast<float*>(ds_raw);                  // [Cin][RH][RWb]
  float* wsm = reg + p.Cin * RH * RWb;                            // [Cin*T][16]
  uint64_t* bar = reinterpret_cast<uint64_t*>(wsm + p.Cin * T * DS_MAXC);
  const int tid = threadIdx.x, b = blockIdx.z;
  const int ty0 = blockIdx.y * DS_TY, tx0 = blockIdx.x * DS_TX;
  // the innermost TMA coordinate must be a multiple of 4 floats (16 bytes; anything else raises an illegal-instruction
  // fault -- tools/probe_tma_box.cu): the patch starts at the aligned column at or left of the reachable one, the box
  // (RWb) is wide enough for any of the four shifts
  const int RY0 = ty0 * p.sh - p.ph - DS_R, RXr = tx0 * p.sw - p.pw - DS_R;
  const int RX0 = RXr - (((RXr % 4) + 4) % 4);
  const size_t HW = static_cast<size_t>(p.H) * p.W;
  if (tid == 0) {
    mbar_init(bar, 1);
    fence_mbar_init();
    mbar_expect_tx(bar, static_cast<uint32_t>(p.Cin * RH * RWb * 4));
    tma_load_3d(reg, &mapIn, bar, RX0, RY0, b * p.Cin);
  }
  for (int e = tid; e < p.Cin * T * DS_MAXC; e += NTH) {
    const int o = e % DS_MAXC, k = e / DS_MAXC;           // k = c*T + t
    wsm[e] = o < p.Cout ? __ldg(p.weight + static_cast<size_t>(o) * p.Cin * T + k) : 0.f;
  }
  __syncthreads();
  mbar_wait(bar, 0);

  const int xo = tx0 + tid % DS_TX, yl = tid / DS_TX;
  int yo[PPT], pix[PPT];
  bool live[PPT];
#pragma unroll
  for (int k = 0; k < PPT; ++k) {
    yo[k] = ty0 + yl + k * YSTEP;
    live[k] = yo[k] < p.Ho && xo < p.Wo;
    pix[k] = live[k] ? yo[k] * p.Wo + xo : 0;
  }
  if (!live[0]) return;                                   // rows only grow with k: nothing to do for this thread
  const int cpdg = p.Cin / p.dg, plane = RH * RWb;
  unsigned long long acc[PPT][DS_MAXC / 2];
#pragma unroll
  for (int k = 0; k < PPT; ++k)
#pragma unroll
    for (int o = 0; o < DS_MAXC / 2; ++o) acc[k][o] = 0ULL;

  for (int dgi = 0; dgi < p.dg; ++dgi) {
    for (int t = 0; t < T; ++t) {
      const int i = t / p.kw, j = t - i * p.kw;
      float w1[PPT], w2[PPT], w3[PPT], w4[PPT], mk[PPT], h_im[PPT], w_im[PPT], hlf[PPT], wlf[PPT];
      const float* q[PPT];
      bool staged = true;
#pragma unroll
      for (int k = 0; k < PPT; ++k) {
        const size_t ob = ((static_cast<size_t>(b) * p.dg + dgi) * 2 * T) * M + pix[k];
        const float oh = __ldg(p.offset + ob + static_cast<size_t>(2 * t) * M);
        const float ow = __ldg(p.offset + ob + static_cast<size_t>(2 * t + 1) * M);
        mk[k] = p.mask ? __ldg(p.mask + ((static_cast<size_t>(b) * p.dg + dgi) * T + t) * M + pix[k]) : 1.f;
        h_im[k] = (float)(yo[k] * p.sh - p.ph + i * p.dh) + oh;
        w_im[k] = (float)(xo * p.sw - p.pw + j * p.dw) + ow;
        hlf[k] = floorf(h_im[k]); wlf[k] = floorf(w_im[k]);
        const float lh = h_im[k] - hlf[k], lw = w_im[k] - wlf[k], hh = 1.f - lh, hw = 1.f - lw;
        w1[k] = hh * hw; w2[k] = hh * lw; w3[k] = lh * hw; w4[k] = lh * lw;
        const bool st = hlf[k] >= (float)RY0 && hlf[k] <= (float)(RY0 + RH - 2) && wlf[k] >= (float)RX0 &&
                        wlf[k] <= (float)(RX0 + RWb - 2);
        staged = staged && (st || !live[k]);
        q[k] = reg + (st ? ((int)hlf[k] - RY0) * RWb + ((int)wlf[k] - RX0) : 0);
        if (!live[k]) mk[k] = 0.f;
      }
      if (staged) {
        const float* wr0 = wsm + (dgi * cpdg * T + t) * DS_MAXC;
        for (int c = 0; c < cpdg; ++c) {
          float val[PPT];
#pragma unroll
          for (int k = 0; k < PPT; ++k) {
            const float* qc = q[k] + (dgi * cpdg + c) * plane;
            val[k] = (w1[k] * qc[0] + w2[k] * qc[1] + w3[k] * qc[RWb] + w4[k] * qc[RWb + 1]) * mk[k];
          }
          const ulonglong2* wr = reinterpret_cast<const ulonglong2*>(wr0 + c * T * DS_MAXC);
#pragma unroll
          for (int o4 = 0; o4 < DS_MAXC / 4; ++o4) {
            const ulonglong2 w = wr[o4];
#pragma unroll
            for (int k = 0; k < PPT; ++k) {
              ffma2(acc[k][2 * o4], w.x, val[k]);
              ffma2(acc[k][2 * o4 + 1], w.y, val[k]);
            }
          }
        }
      } else {
        // some corner of this thread lies outside the staged patch (|offset| > DS_R): explicit tests, global loads
#pragma unroll
        for (int k = 0; k < PPT; ++k) {
          if (!live[k]) continue;
          if (!(h_im[k] > -1.f && w_im[k] > -1.f && h_im[k] < (float)p.H && w_im[k] < (float)p.W)) continue;
          const int hl = (int)hlf[k], wl = (int)wlf[k];
          const bool t0 = hl >= 0, l0 = wl >= 0, b1 = hl + 1 <= p.H - 1, r1 = wl + 1 <= p.W - 1;
          for (int c = dgi * cpdg; c < (dgi + 1) * cpdg; ++c) {
            const float* src = p.in + (static_cast<size_t>(b) * p.Cin + c) * HW + static_cast<ptrdiff_t>(hl) * p.W + wl;
            const float v1 = (t0 && l0) ? __ldg(src) : 0.f, v2 = (t0 && r1) ? __ldg(src + 1) : 0.f;
            const float v3 = (b1 && l0) ? __ldg(src + p.W) : 0.f, v4 = (b1 && r1) ? __ldg(src + p.W + 1) : 0.f;
            const float val = (w1[k] * v1 + w2[k] * v2 + w3[k] * v3 + w4[k] * v4) * mk[k];
            const ulonglong2* wr = reinterpret_cast<const ulonglong2*>(wsm + (c * T + t) * DS_MAXC);
#pragma unroll
            for (int o4 = 0; o4 < DS_MAXC / 4; ++o4) {
              const ulonglong2 w = wr[o4];
              ffma2(acc[k][2 * o4], w.x, val);
              ffma2(acc[k][2 * o4 + 1], w.y, val);
            }
          }
        }
      }
    }
  }
#pragma unroll
  for (int k = 0; k < PPT; ++k) {
    if (!live[k]) continue;
#pragma unroll
    for (int o2 = 0; o2 < DS_MAXC / 2; ++o2) {
      const float lo = __uint_as_float(static_cast<unsigned>(acc[k][o2] & 0xffffffffULL));
      const float hi = __uint_as_float(static_cast<unsigned>(acc[k][o2] >> 32));
      if (2 * o2 < p.Cout) p.out[(static_cast<size_t>(b) * p.Cout + 2 * o2) * M + pix[k]] = lo + __ldg(p.bias + 2 * o2);
      if (2 * o2 + 1 < p.Cout)
        p.out[(static_cast<size_t>(b) * p.Cout + 2 * o2 + 1) * M + pix[k]] = hi + __ldg(p.bias + 2 * o2 + 1);
    }
  }
}

// ---- 2-CTA cluster variant (dg == 1, even Cin): the kernel above is latency-bound at 8 warps per SM (ncu: 31 % issue
// slots used, short-scoreboard stalls; the 16-channel patch allows 2 CTAs per SM).  Here a CTA PAIR shares a pixel tile:
// each CTA stages HALF of the input channels (half the patch: 4 CTAs per SM) and accumulates its partial sums for all
// 16 outputs; the halves are exchanged through distributed shared memory (each CTA finalises 8 output channels), so
// twice as many warps work on the same tile and the global stores are split as well.
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void st_peer_smem_f32(const float* local_ptr, uint32_t peer, float v) {
  uint32_t ra;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(smem_u32(local_ptr)), "r"(peer));
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(ra), "f"(v) : "memory");
}

__device__ __forceinline__ unsigned long long pick_pair(const unsigned long long (&a)[DS_MAXC / 2], int idx) {
  unsigned long long v = a[0];
#pragma unroll
  for (int j = 1; j < DS_MAXC / 2; ++j)
    if (j == idx) v = a[j];                      // predicated moves: no dynamic register indexing
  return v;
}

template <int PPT, int CS>
__global__ void __cluster_dims__(CS, 1, 1) __launch_bounds__(DS_TY * DS_TX / PPT, 2 * CS)   // <2,2>: 4 x 128 threads, <1,4>: 8 x 256
dcn_small_pair_kernel(const __grid_constant__ CUtensorMap mapIn, const DcnParams p, int RH, int RWb, int DS_R) {
  constexpr int NTH = DS_TY * DS_TX / PPT, YSTEP = DS_TY / PPT, NPIX = DS_TY * DS_TX, OW = DS_MAXC / CS;
  extern __shared__ __align__(128) unsigned char ds_raw[];
  const int T = p.kh * p.kw, M = p.Ho * p.Wo, Ch = p.Cin / CS;
  float* reg = reinterpret_cast<float*>(ds_raw);                  // [Ch][RH][RWb]; later the exchange buffer [CS-1][OW][NPIX]
  float* wsm = reg + Ch * RH * RWb;                               // [Ch*T][16]
  uint64_t* bar = reinterpret_cast<uint64_t*>(wsm + Ch * T * DS_MAXC);
  const uint32_t rank = cluster_ctarank();
  const int c_first = static_cast<int>(rank) * Ch;
  const int tid = threadIdx.x, b = blockIdx.z;
  const int ty0 = blockIdx.y * DS_TY, tx0 = (blockIdx.x / CS) * DS_TX;
  const int RY0 = ty0 * p.sh - p.ph - DS_R, RXr = tx0 * p.sw - p.pw - DS_R;
  const int RX0 = RXr - (((RXr % 4) + 4) % 4);                    // 16-byte aligned TMA column (see above)
  const size_t HW = static_cast<size_t>(p.H) * p.W;
  if (tid == 0) {
    mbar_init(bar, 1);
    fence_mbar_init();
    mbar_expect_tx(bar, static_cast<uint32_t>(Ch * RH * RWb * 4));
    tma_load_3d(reg, &mapIn, bar, RX0, RY0, b * p.Cin + c_first);
  }
  for (int e = tid; e < Ch * T * DS_MAXC; e += NTH) {
    const int o = e % DS_MAXC, k = e / DS_MAXC;           // k = c_local*T + t
    wsm[e] = o < p.Cout ? __ldg(p.weight + static_cast<size_t>(o) * p.Cin * T + c_first * T + k) : 0.f;
  }
  __syncthreads();
  mbar_wait(bar, 0);

  const int xo = tx0 + tid % DS_TX, yl = tid / DS_TX;
  int yo[PPT], pix[PPT];
  bool live[PPT];
#pragma unroll
  for (int k = 0; k < PPT; ++k) {
    yo[k] = ty0 + yl + k * YSTEP;
    live[k] = yo[k] < p.Ho && xo < p.Wo;
    pix[k] = live[k] ? yo[k] * p.Wo + xo : 0;
  }
  const int plane = RH * RWb;
  unsigned long long acc[PPT][DS_MAXC / 2];
#pragma unroll
  for (int k = 0; k < PPT; ++k)
#pragma unroll
    for (int o = 0; o < DS_MAXC / 2; ++o) acc[k][o] = 0ULL;

  // sampling descriptors of tap t + 1 are fetched while tap t is accumulated
  float oh[PPT], ow[PPT], om[PPT];
  auto fetch = [&](int t) {
#pragma unroll
    for (int k = 0; k < PPT; ++k) {
      const size_t ob = (static_cast<size_t>(b) * 2 * T) * M + pix[k];
      oh[k] = __ldg(p.offset + ob + static_cast<size_t>(2 * t) * M);
      ow[k] = __ldg(p.offset + ob + static_cast<size_t>(2 * t + 1) * M);
      om[k] = p.mask ? __ldg(p.mask + (static_cast<size_t>(b) * T + t) * M + pix[k]) : 1.f;
    }
  };
  if (live[0]) fetch(0);
  for (int t = 0; t < T && live[0]; ++t) {
    const int i = t / p.kw, j = t - i * p.kw;
    float w1[PPT], w2[PPT], w3[PPT], w4[PPT], mk[PPT], h_im[PPT], w_im[PPT], hlf[PPT], wlf[PPT];
    const float* q[PPT];
    bool staged = true;
#pragma unroll
    for (int k = 0; k < PPT; ++k) {
      mk[k] = live[k] ? om[k] : 0.f;
      h_im[k] = (float)(yo[k] * p.sh - p.ph + i * p.dh) + oh[k];
      w_im[k] = (float)(xo * p.sw - p.pw + j * p.dw) + ow[k];
      hlf[k] = floorf(h_im[k]); wlf[k] = floorf(w_im[k]);
      const float lh = h_im[k] - hlf[k], lw = w_im[k] - wlf[k], hh = 1.f - lh, hw = 1.f - lw;
      w1[k] = hh * hw; w2[k] = hh * lw; w3[k] = lh * hw; w4[k] = lh * lw;
      const bool st = hlf[k] >= (float)RY0 && hlf[k] <= (float)(RY0 + RH - 2) && wlf[k] >= (float)RX0 &&
                      wlf[k] <= (float)(RX0 + RWb - 2);
      staged = staged && (st || !live[k]);
      q[k] = reg + (st ? ((int)hlf[k] - RY0) * RWb + ((int)wlf[k] - RX0) : 0);
    }
    if (t + 1 < T) fetch(t + 1);
    if (staged) {
      const float* wr0 = wsm + t * DS_MAXC;
      for (int c = 0; c < Ch; ++c) {
        float val[PPT];
#pragma unroll
        for (int k = 0; k < PPT; ++k) {
          const float* qc = q[k] + c * plane;
          val[k] = (w1[k] * qc[0] + w2[k] * qc[1] + w3[k] * qc[RWb] + w4[k] * qc[RWb + 1]) * mk[k];
        }
        const ulonglong2* wr = reinterpret_cast<const ulonglong2*>(wr0 + c * T * DS_MAXC);
#pragma unroll
        for (int o4 = 0; o4 < DS_MAXC / 4; ++o4) {
          const ulonglong2 w = wr[o4];
#pragma unroll
          for (int k = 0; k < PPT; ++k) {
            ffma2(acc[k][2 * o4], w.x, val[k]);
            ffma2(acc[k][2 * o4 + 1], w.y, val[k]);
          }
        }
      }
    } else {
      // some corner of this thread lies outside the staged patch (|offset| > DS_R): explicit tests, global loads
#pragma unroll
      for (int k = 0; k < PPT; ++k) {
        if (!live[k]) continue;
        if (!(h_im[k] > -1.f && w_im[k] > -1.f && h_im[k] < (float)p.H && w_im[k] < (float)p.W)) continue;
        const int hl = (int)hlf[k], wl = (int)wlf[k];
        const bool t0 = hl >= 0, l0 = wl >= 0, b1 = hl + 1 <= p.H - 1, r1 = wl + 1 <= p.W - 1;
        for (int c = 0; c < Ch; ++c) {
          const float* src = p.in + (static_cast<size_t>(b) * p.Cin + c_first + c) * HW + static_cast<ptrdiff_t>(hl) * p.W + wl;
          const float v1 = (t0 && l0) ? __ldg(src) : 0.f, v2 = (t0 && r1) ? __ldg(src + 1) : 0.f;
          const float v3 = (b1 && l0) ? __ldg(src + p.W) : 0.f, v4 = (b1 && r1) ? __ldg(src + p.W + 1) : 0.f;
          const float val = (w1[k] * v1 + w2[k] * v2 + w3[k] * v3 + w4[k] * v4) * mk[k];
          const ulonglong2* wr = reinterpret_cast<const ulonglong2*>(wsm + (c * T + t) * DS_MAXC);
#pragma unroll
          for (int o4 = 0; o4 < DS_MAXC / 4; ++o4) {
            const ulonglong2 w = wr[o4];
            ffma2(acc[k][2 * o4], w.x, val);
            ffma2(acc[k][2 * o4 + 1], w.y, val);
          }
        }
      }
    }
  }
  // ---- exchange: rank r finalises output channels [r*OW, (r+1)*OW); the other slices of this CTA's partial sums go
  // to their owners' shared memory (a patch is dead once every CTA of the cluster has passed the first cluster barrier)
  cluster_sync_all();
#pragma unroll
  for (int d = 1; d < CS; ++d) {
    const uint32_t peer = (rank + d) % CS;
    const int slot = static_cast<int>(rank) - (rank > peer ? 1 : 0);        // my index among the peer's senders
#pragma unroll
    for (int k = 0; k < PPT; ++k) {
#pragma unroll
      for (int o2 = 0; o2 < OW / 2; ++o2) {
        const unsigned long long v = pick_pair(acc[k], static_cast<int>(peer) * (OW / 2) + o2);
        float* dst = reg + (slot * OW + 2 * o2) * NPIX + k * NTH + tid;
        st_peer_smem_f32(dst, peer, __uint_as_float(static_cast<unsigned>(v & 0xffffffffULL)));
        st_peer_smem_f32(dst + NPIX, peer, __uint_as_float(static_cast<unsigned>(v >> 32)));
      }
    }
  }
  cluster_sync_all();
#pragma unroll
  for (int k = 0; k < PPT; ++k) {
    if (!live[k]) continue;
#pragma unroll
    for (int o2 = 0; o2 < OW / 2; ++o2) {
      const unsigned long long v = pick_pair(acc[k], static_cast<int>(rank) * (OW / 2) + o2);
      const int o = static_cast<int>(rank) * OW + 2 * o2;
      float lo = __uint_as_float(static_cast<unsigned>(v & 0xffffffffULL));
      float hi = __uint_as_float(static_cast<unsigned>(v >> 32));
      // partial sums are added in rank order (fixed association: results do not depend on which CTA finalises)
      float plo[CS], phi[CS];
#pragma unroll
      for (int r = 0; r < CS; ++r) { plo[r] = 0.f; phi[r] = 0.f; }
#pragma unroll
      for (int sl = 0; sl < CS - 1; ++sl) {
        const int sender = sl + (sl >= static_cast<int>(rank) ? 1 : 0);
        const float a0 = reg[(sl * OW + 2 * o2) * NPIX + k * NTH + tid], a1 = reg[(sl * OW + 2 * o2 + 1) * NPIX + k * NTH + tid];
#pragma unroll
        for (int r = 0; r < CS; ++r)
          if (r == sender) { plo[r] = a0; phi[r] = a1; }
      }
#pragma unroll
      for (int r = 0; r < CS; ++r)
        if (r == static_cast<int>(rank)) { plo[r] = lo; phi[r] = hi; }
      lo = plo[0]; hi = phi[0];
#pragma unroll
      for (int r = 1; r < CS; ++r) { lo += plo[r]; hi += phi[r]; }
      if (o < p.Cout) p.out[(static_cast<size_t>(b) * p.Cout + o) * M + pix[k]] = lo + __ldg(p.bias + o);
      if (o + 1 < p.Cout) p.out[(static_cast<size_t>(b) * p.Cout + o + 1) * M + pix[k]] = hi + __ldg(p.bias + o + 1);
    }
  }
}

static void dcn_small_region(const DcnParams& p, int* RH, int* RW, int* RWp, size_t* smem) {
  const int DS_R = ds_halo();
  *RH = (DS_TY - 1) * p.sh + p.dh * (p.kh - 1) + 2 * DS_R + 2;
  *RW = (DS_TX - 1) * p.sw + p.dw * (p.kw - 1) + 2 * DS_R + 2;
  *RWp = *RW | 1;                                          // odd pitch: rows land in different banks
  *smem = (static_cast<size_t>(p.Cin) * p.kh * p.kw * DS_MAXC + static_cast<size_t>(p.Cin) * *RH * *RWp) * 4;
}

static bool dcn_small_eligible(const DcnParams& p) {
  if (p.G != 1 || p.Cin > DS_MAXC || p.Cout > DS_MAXC || p.B > 65535) return false;
  int RH, RW, RWp;
  size_t smem;
  dcn_small_region(p, &RH, &RW, &RWp, &smem);
  return smem <= 100 * 1024;
}

static int launch_dcn_small(const DcnParams& p, cudaStream_t s) {
  int RH, RW, RWp;
  size_t smem;
  dcn_small_region(p, &RH, &RW, &RWp, &smem);
  // TMA-staged variant: needs a tensor map over the input (16-byte aligned rows) and a box that fits next to a second CTA
  const int RWb = (RW + 3 + 3) & ~3;          // + 3: the patch origin is rounded down to a multiple of 4 columns
  const size_t smem_t = (static_cast<size_t>(p.Cin) * RH * RWb + static_cast<size_t>(p.Cin) * p.kh * p.kw * DS_MAXC) * 4 + 16;
  static const int ppt_env = tuning_env("SF_DCN_PPT") ? atoi(tuning_env("SF_DCN_PPT")) : 2;   // A/B hook (tuning build)
  if (p.W % 4 == 0 && (reinterpret_cast<uintptr_t>(p.in) & 15) == 0 && RWb <= 256 && RH <= 256 && smem_t <= 110 * 1024 &&
      ppt_env != 0) {
    CUtensorMap mapIn;
    int rc = make_tmap_3d_f32(&mapIn, p.in, p.W, p.H, static_cast<uint64_t>(p.B) * p.Cin, RWb, RH, p.Cin);
    if (rc) return rc;
    dim3 grid(ceil_div(p.Wo, DS_TX), ceil_div(p.Ho, DS_TY), p.B);
    SF_CHECK_ARG(grid.y <= 65535, "dcn_forward: output too tall");
    static const int pair_env = tuning_env("SF_DCN_PAIR") ? atoi(tuning_env("SF_DCN_PAIR")) : 2;   // A/B hook (tuning build)
    // CTA cluster per tile, 1/CS of the input channels each (dg == 1: the shipped layer); measured at 16 -> 16, 8 x 68x120:
    // one CTA 29.8 us, pairs 28.2 us (default); quads (the kernel is templated on the cluster size: 4 channels per CTA,
    // one pixel per thread) need 8 x 256 threads per SM = 32 registers per thread and spill: 67 us, not instantiated
    int CS = pair_env == 2 ? 2 : 0;
    if (CS && (p.dg != 1 || p.Cin % CS != 0)) CS = 0;
    if (CS) {
      const int Ch = p.Cin / CS;
      const size_t patch = static_cast<size_t>(Ch) * RH * RWb;
      const size_t xch = static_cast<size_t>(CS - 1) * (DS_MAXC / CS) * DS_TY * DS_TX;
      if (patch < xch) CS = 0;
    }
    if (CS) {
      const int Ch = p.Cin / CS;
      const size_t smem_p = (static_cast<size_t>(Ch) * RH * RWb + static_cast<size_t>(Ch) * p.kh * p.kw * DS_MAXC) * 4 + 16;
      CUtensorMap mapH;
      rc = make_tmap_3d_f32(&mapH, p.in, p.W, p.H, static_cast<uint64_t>(p.B) * p.Cin, RWb, RH, Ch);
      if (rc) return rc;
      dim3 grid2(CS * grid.x, grid.y, grid.z);
      auto go = [&](auto kern, int nth) -> int {
        SF_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem_p)));
        kern<<<grid2, nth, smem_p, s>>>(mapH, p, RH, RWb, ds_halo());
        SF_LAUNCH_CHECK();
        return SF_OK;
      };
      return go(dcn_small_pair_kernel<2, 2>, DS_TY * DS_TX / 2);
    }
    if (ppt_env == 1) {
      SF_CUDA(cudaFuncSetAttribute(dcn_small_tma_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem_t)));
      dcn_small_tma_kernel<1><<<grid, DS_TY * DS_TX, smem_t, s>>>(mapIn, p, RH, RWb, ds_halo());
    } else {
      SF_CUDA(cudaFuncSetAttribute(dcn_small_tma_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem_t)));
      dcn_small_tma_kernel<2><<<grid, DS_TY * DS_TX / 2, smem_t, s>>>(mapIn, p, RH, RWb, ds_halo());
    }
    SF_LAUNCH_CHECK();
    return SF_OK;
  }
  SF_CUDA(cudaFuncSetAttribute(dcn_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  dim3 grid(ceil_div(p.Wo, DS_TX), ceil_div(p.Ho, DS_TY), p.B);
  SF_CHECK_ARG(grid.y <= 65535, "dcn_forward: output too tall");
  SF_CHECK_ARG(RH * RW < 65536, "dcn_forward: patch too large");
  const unsigned rw_magic = static_cast<unsigned>((1ULL << 32) / static_cast<unsigned>(RW)) + 1u;
  dcn_small_kernel<<<grid, DS_TY * DS_TX, smem, s>>>(p, RH, RW, RWp, ds_halo(), rw_magic);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

}  // namespace sf

static int dcn_forward_impl(const float* input, const float* weight, const float* bias, const float* offset,
                            const float* mask, float* out, int B, int Cin, int H, int W, int Cout, int kh, int kw,
                            int stride_h, int stride_w, int pad_h, int pad_w, int dil_h, int dil_w, int group,
                            int deformable_group, void* workspace, size_t workspace_bytes, int algo, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(input && weight && bias && offset && out, "dcn_forward: null pointer");
  SF_CHECK_ARG(B > 0 && Cin > 0 && H > 0 && W > 0 && Cout > 0 && kh > 0 && kw > 0, "dcn_forward: non-positive dimension");
  SF_CHECK_ARG(stride_h > 0 && stride_w > 0 && dil_h > 0 && dil_w > 0 && pad_h >= 0 && pad_w >= 0,
               "dcn_forward: bad stride / dilation / padding");
  SF_CHECK_ARG(group > 0 && Cin % group == 0 && Cout % group == 0,
               "channels(%d) and channels_out(%d) must divide group(%d)", Cin, Cout, group);
  SF_CHECK_ARG(deformable_group > 0 && Cin % deformable_group == 0,
               "channels(%d) must divide deformable_group(%d)", Cin, deformable_group);
  if (algo == SF_ALGO_EXACT) algo = SF_ALGO_GENERIC;      // the FFMA kernels (dcn_small / dcn_forward) are the exact path
  SF_CHECK_ARG(algo == SF_ALGO_AUTO || algo == SF_ALGO_GENERIC || algo == SF_ALGO_TC, "dcn_forward: unknown algo %d", algo);
  DcnParams p;
  p.in = input; p.weight = weight; p.bias = bias; p.offset = offset; p.mask = mask; p.out = out;
  p.B = B; p.Cin = Cin; p.H = H; p.W = W; p.Cout = Cout;
  p.kh = kh; p.kw = kw; p.sh = stride_h; p.sw = stride_w; p.ph = pad_h; p.pw = pad_w; p.dh = dil_h; p.dw = dil_w;
  p.G = group; p.dg = deformable_group;
  p.Ho = (H + 2 * pad_h - (dil_h * (kh - 1) + 1)) / stride_h + 1;  // modulated_deform_conv_cuda.cu:75-76
  p.Wo = (W + 2 * pad_w - (dil_w * (kw - 1) + 1)) / stride_w + 1;
  SF_CHECK_ARG(p.Ho > 0 && p.Wo > 0, "dcn_forward: empty output (%d x %d)", p.Ho, p.Wo);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int Cog = Cout / group, Cig = Cin / group;
  // tensor-core path: worthwhile once the GEMM is at least 32 x 32 channels per group (64->128, 128->128,
  // 256->256); the 16-channel shipped-model shape is HBM/latency-bound and stays on the exact FFMA kernel
  if (algo == SF_ALGO_TC || (algo == SF_ALGO_AUTO && Cig >= 32 && Cog >= 32)) {
    DcnTcParams q;
    q.in = input; q.bias = bias; q.offset = offset; q.mask = mask; q.out = out;
    q.B = B; q.Cin = Cin; q.H = H; q.W = W; q.Cout = Cout; q.Ho = p.Ho; q.Wo = p.Wo;
    q.kh = kh; q.kw = kw; q.sh = stride_h; q.sw = stride_w; q.ph = pad_h; q.pw = pad_w; q.dh = dil_h; q.dw = dil_w;
    q.G = group; q.dg = deformable_group;
    bool handled = false;
    int rc = launch_dcn_tc2(q, weight, workspace, workspace_bytes, s, &handled);   // channels-last gathers (needs the larger workspace)
    if (rc) return rc;
    if (handled) return SF_OK;
    rc = launch_dcn_tc(q, weight, workspace, workspace_bytes, s, &handled);
    if (rc) return rc;
    if (handled) return SF_OK;
    SF_CHECK_ARG(algo != SF_ALGO_TC, "dcn_forward: the tensor-core path needs %zu workspace bytes (got %zu) and <= 256 "
                 "output channels per group", dcn_tc_workspace_bytes(Cin, Cout, kh, kw, group), workspace_bytes);
  }
  if (dcn_small_eligible(p)) {
    static const bool off = tuning_env("SF_DCN_SMALL") && atoi(tuning_env("SF_DCN_SMALL")) == 0;   // A/B hook (tuning build)
    if (!off) return launch_dcn_small(p, s);
  }
  if (Cog <= 16) return launch_dcn<16, 128>(p, s);
  if (Cog <= 32) return launch_dcn<32, 128>(p, s);
  return launch_dcn<64, 64>(p, s);
}

extern "C" int sf_dcn_forward_f32(const float* input, const float* weight, const float* bias, const float* offset,
                                  const float* mask, float* out, int B, int Cin, int H, int W, int Cout, int kh, int kw,
                                  int stride_h, int stride_w, int pad_h, int pad_w, int dil_h, int dil_w, int group,
                                  int deformable_group, void* stream) {
  return dcn_forward_impl(input, weight, bias, offset, mask, out, B, Cin, H, W, Cout, kh, kw, stride_h, stride_w, pad_h,
                          pad_w, dil_h, dil_w, group, deformable_group, nullptr, 0, SF_ALGO_GENERIC, stream);
}

extern "C" size_t sf_dcn_workspace_bytes(int Cin, int Cout, int kh, int kw, int group) {
  return sf::dcn_tc_workspace_bytes(Cin, Cout, kh, kw, group);
}

extern "C" size_t sf_dcn_workspace_bytes_ex(int B, int Cin, int H, int W, int Cout, int kh, int kw, int group,
                                            int deformable_group) {
  const size_t a = sf::dcn_tc_workspace_bytes(Cin, Cout, kh, kw, group);
  const size_t b = sf::dcn_tc2_workspace_bytes(B, Cin, H, W, Cout, kh, kw, group, deformable_group);
  return a > b ? a : b;
}

extern "C" int sf_dcn_forward_ws_f32(const float* input, const float* weight, const float* bias, const float* offset,
                                     const float* mask, float* out, int B, int Cin, int H, int W, int Cout, int kh,
                                     int kw, int stride_h, int stride_w, int pad_h, int pad_w, int dil_h, int dil_w,
                                     int group, int deformable_group, void* workspace, size_t workspace_bytes, int algo,
                                     void* stream) {
  return dcn_forward_impl(input, weight, bias, offset, mask, out, B, Cin, H, W, Cout, kh, kw, stride_h, stride_w, pad_h,
                          pad_w, dil_h, dil_w, group, deformable_group, workspace, workspace_bytes, algo, stream);
}

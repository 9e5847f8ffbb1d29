// Deformable convolution forward on tcgen05 tensor cores  (the >= 32-channel shapes).
//
// Reference: deform/src/cuda/modulated_deform_conv_cuda.cu:19-121 + modulated_deform_im2col_cuda.cuh:25-54,128-194
// (im2col kernel -> column buffer in HBM -> cuBLAS addmm -> NHWC->NCHW permute).  Here the column buffer never
// exists: it is the A operand of an implicit GEMM that goes straight from the gather threads into TENSOR MEMORY.
//
//   out[b, g*Cog+o, pix] = bias + sum_k W[g*Cog+o][k] * col[k][pix],   k = c*T + t  (c < Cin/G, t < kh*kw)
//   col[k][pix] = mask * bilinear(in[b, g*Cig+c], sampling point of tap t at pix)
//
// Work tile: 128 output pixels (M, one TMEM lane each) x all Cog output channels of one group (N <= 256).
//   * 16 gather warps (four per TMEM lane quarter, thread = pixel; measured against 8: more gathers in flight): sampling descriptors (4 clamped corner offsets +
//     4 bilinear weights with the mask folded in) once per tile and deformable group in shared memory; per 64-wide
//     K chunk every thread gathers its 16 column values, splits them into bf16 hi + bf16 mid and stores them with
//     tcgen05.st (A operand, K-major, two K-consecutive bf16 per 32-bit column);
//   * warp 0 streams the pre-split weights (bf16 hi/mid, [Cout][Kpad], prepared once per call into the caller's
//     workspace by dcn_split_weights_kernel) with TMA boxes [64 k][N rows], SWIZZLE_128B = K-major B operand;
//   * one thread issues hi*hi + hi*mid + mid*hi per 16-wide k-step into a [128 x N] fp32 TMEM accumulator;
//   * 4 epilogue warps read the accumulator (thread = pixel), add the bias and write NCHW rows (a warp writes
//     32 consecutive pixels of one output channel: full 128-byte lines).  No permute pass.
// fp32 parity: 3-term bf16 split, ~2^-16 relative per product (tests: 1e-4 of the output magnitude); the exact
// FFMA kernel (dcn.cu) stays the path for small channel counts and for bit-level checks.
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_common.cuh"
#include "dcn_tc.cuh"
#include "../../include/stereoformer_b200.h"

namespace sf {

constexpr int DT_KC = 64;                 // K values per chunk (one 128-byte swizzle row of bf16)
constexpr int DT_NG = 16;                 // gather warps (warps 1..16): four per TMEM lane quarter
constexpr int DT_KPT = DT_KC / (DT_NG / 4);  // K values per gather thread and chunk
constexpr int DT_WARP_MMA = 1 + DT_NG;
constexpr int DT_THREADS = (DT_WARP_MMA + 1 + 4) * 32;
constexpr int DT_TA_BASE = 256;           // TMEM columns of the A stages: [NST][hi 32 | mid 32]
constexpr int DT_MAXST = 4;


// fp32 weights [Cout][Cig][T] -> bf16 hi / mid [2][Cout][Kpad] (zero padded along K) with the reduction index PERMUTED to
// k' = t * Cig + c (tap-major): a gather thread's 32 consecutive k' then share one tap, i.e. one sampling descriptor, and
// walk consecutive input planes
__global__ void dcn_split_weights_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ ws, int Cout, int K,
                                         int Kpad, int Cig, int T) {
  const size_t n = static_cast<size_t>(Cout) * Kpad;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int k = static_cast<int>(i % Kpad);
    const size_t o = i / Kpad;
    const int t = k / Cig, c = k - t * Cig;
    const float v = k < K ? w[o * K + static_cast<size_t>(c) * T + t] : 0.f;
    const __nv_bfloat16 h = __float2bfloat16_rn(v);
    ws[i] = h;
    ws[n + i] = __float2bfloat16_rn(v - __bfloat162float(h));
  }
}

__global__ void __launch_bounds__(DT_THREADS, 1)
dcn_tc_kernel(const __grid_constant__ CUtensorMap mapW, const DcnTcParams p) {
  using namespace sftc;
  extern __shared__ __align__(1024) unsigned char smem[];
  const int T = p.kh * p.kw;
  const int stage_bytes = 2 * p.N * 128;                           // B hi | B mid: [N rows][128 B]
  unsigned char* sB = smem;                                        // [NST][stage_bytes]
  int4* d_off = reinterpret_cast<int4*>(sB + p.NST * stage_bytes); // [dg][T][128] corner offsets (clamped)
  float4* d_w = reinterpret_cast<float4*>(d_off + p.dg * T * 128); // [dg][T][128] corner weights * mask
  uint64_t* bars = reinterpret_cast<uint64_t*>(d_w + p.dg * T * 128);
  uint64_t* b_full = bars;                      // [NST] TMA -> MMA
  uint64_t* a_full = bars + DT_MAXST;           // [NST] gather warps (8) -> MMA
  uint64_t* st_empty = bars + 2 * DT_MAXST;     // [NST] MMA commit -> TMA producer + gather warps
  uint64_t* acc_full = bars + 3 * DT_MAXST;     // MMA commit -> epilogue
  uint64_t* acc_empty = acc_full + 1;           // epilogue (4 warps) -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.NST; ++s) { mbar_init(&b_full[s], 1); mbar_init(&a_full[s], DT_NG); mbar_init(&st_empty[s], 1); }
    mbar_init(acc_full, 1);
    mbar_init(acc_empty, 4);
    fence_mbar_init();
    tma_prefetch_desc(&mapW);
  }
  if (warp == DT_WARP_MMA) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const int M = p.Ho * p.Wo;
  const int Cig = p.Cin / p.G, Cog = p.Cout / p.G;
  const int Kpad = p.nchunk * DT_KC;

  if (warp == 0) {
    // ================================================================= weight producer (TMA)
    if (lane == 0) {
      int s = 0, ph = 1;
      for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        const int g = (tile / p.nptile) % p.G;
        for (int ch = 0; ch < p.nchunk; ++ch) {
          mbar_wait(&st_empty[s], ph);
          unsigned char* dst = sB + s * stage_bytes;
          mbar_expect_tx(&b_full[s], stage_bytes);
          tma_load_3d(dst, &mapW, &b_full[s], ch * DT_KC, g * Cog, 0);
          tma_load_3d(dst + p.N * 128, &mapW, &b_full[s], ch * DT_KC, g * Cog, 1);
          if (++s == p.NST) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp < DT_WARP_MMA) {
    // ================================================================= gather warps: column tile -> tensor memory
    const int qa = warp & 3, kh2 = (warp - 1) >> 2;      // TMEM lane quarter, half of the chunk's K range
    const int px = 32 * qa + lane;
    const int gtid = (warp - 1) * 32 + lane;             // 0..255 within the gather group
    const size_t HW = static_cast<size_t>(p.H) * p.W;
    const int cpdg = p.Cin / p.dg;
    const uint32_t ta = tmem + DT_TA_BASE + (DT_KPT / 2) * kh2 + (static_cast<uint32_t>(32 * qa) << 16);
    int s = 0, ph = 1;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
      const int pt = tile % p.nptile, t2 = tile / p.nptile;
      const int g = t2 % p.G, b = t2 / p.G;
      const int pix0 = pt * 128;
      // ---- sampling descriptors of this pixel tile (modulated_deform_im2col_cuda.cuh:25-54,163-180)
      named_bar_sync(2, DT_NG * 32);                     // every gather warp has finished the previous tile
      for (int e = gtid; e < p.dg * T * 128; e += DT_NG * 32) {
        const int ex = e & 127, r = e >> 7;
        const int t = r % T, dgi = r / T;
        const int pix = pix0 + ex;
        int4 off = make_int4(0, 0, 0, 0);
        float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
        if (pix < M) {
          const int yo = pix / p.Wo, xo = pix % p.Wo;
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
            const float hh = 1.f - lh, hw = 1.f - lw;
            const bool t0 = hl >= 0, l0 = wl >= 0, b1 = hh_i <= p.H - 1, r1 = wh_i <= p.W - 1;
            if (t0 && l0) { off.x = hl * p.W + wl; w.x = hh * hw * m; }
            if (t0 && r1) { off.y = hl * p.W + wh_i; w.y = hh * lw * m; }
            if (b1 && l0) { off.z = hh_i * p.W + wl; w.z = lh * hw * m; }
            if (b1 && r1) { off.w = hh_i * p.W + wh_i; w.w = lh * lw * m; }
          }
        }
        d_off[e] = off;
        d_w[e] = w;
      }
      named_bar_sync(2, DT_NG * 32);
      const float* in_g = p.in + (static_cast<size_t>(b) * p.Cin + static_cast<size_t>(g) * Cig) * HW;
      for (int ch = 0; ch < p.nchunk; ++ch) {
        mbar_wait(&st_empty[s], ph);
        tc_fence_after();
        // this thread's 32 K values of the chunk: k' = k0 .. k0+31 with k' = t*Cig + c (tap-major, see the weight split):
        // the sampling descriptor is reloaded only when the tap or the deformable group changes
        const int k0 = ch * DT_KC + DT_KPT * kh2;
        int t = k0 / Cig, c = k0 - t * Cig;
        int dgi = (g * Cig + c) / cpdg, bound = (dgi + 1) * cpdg - g * Cig;   // channel (within the group) where dgi steps
        const float* src = in_g + static_cast<size_t>(c) * HW;
        int4 off = make_int4(0, 0, 0, 0);
        float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
        bool fresh = true;
        uint32_t ahi[DT_KPT / 2], amid[DT_KPT / 2];
#pragma unroll
        for (int u = 0; u < DT_KPT / 2; ++u) {
          float v[2];
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int k = k0 + 2 * u + e;
            float val = 0.f;
            if (k < p.K) {                                  // warp-uniform
              if (fresh) {
                const int di = (dgi * T + t) * 128 + px;
                off = d_off[di];
                w = d_w[di];
                fresh = false;
              }
              val = w.x * __ldg(src + off.x) + w.y * __ldg(src + off.y) + w.z * __ldg(src + off.z) + w.w * __ldg(src + off.w);
            }
            v[e] = val;
            ++c; src += HW;
            if (c == Cig) { c = 0; ++t; src = in_g; dgi = (g * Cig) / cpdg; bound = (dgi + 1) * cpdg - g * Cig; fresh = true; }
            else if (c == bound) { ++dgi; bound += cpdg; fresh = true; }
          }
          const __nv_bfloat162 h = __floats2bfloat162_rn(v[0], v[1]);
          const __nv_bfloat162 m = __floats2bfloat162_rn(v[0] - __low2float(h), v[1] - __high2float(h));
          ahi[u] = *reinterpret_cast<const uint32_t*>(&h);
          amid[u] = *reinterpret_cast<const uint32_t*>(&m);
        }
        static_assert(DT_KPT == 16, "gather threads store 8 packed columns per half");
        tmem_st8(ta + s * 64, ahi);
        tmem_st8(ta + s * 64 + 32, amid);
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&a_full[s]);
        if (++s == p.NST) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == DT_WARP_MMA) {
    // ================================================================= MMA issuer
    if (lane == 0) {
      const uint32_t idesc = make_idesc_bf16(128, p.N, 0, 0);   // A: TMEM, B: K-major shared memory
      int s = 0, ph = 0, t = 0;
      for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++t) {
        mbar_wait(acc_empty, (t & 1) ^ 1);
        tc_fence_after();
        for (int ch = 0; ch < p.nchunk; ++ch) {
          mbar_wait(&b_full[s], ph);
          mbar_wait(&a_full[s], ph);
          tc_fence_after();
          const uint32_t b_hi = smem_u32(sB + s * stage_bytes), b_mid = b_hi + p.N * 128;
          const uint32_t a_hi = tmem + DT_TA_BASE + s * 64, a_mid = a_hi + 32;
#pragma unroll
          for (int j = 0; j < DT_KC / 16; ++j) {
            const uint64_t dbh = make_smem_desc(b_hi + j * 32, 16, 1024);
            const uint64_t dbm = make_smem_desc(b_mid + j * 32, 16, 1024);
            umma_bf16_ts(tmem, a_hi + 8 * j, dbh, idesc, (ch | j) != 0);
            umma_bf16_ts(tmem, a_hi + 8 * j, dbm, idesc, true);
            umma_bf16_ts(tmem, a_mid + 8 * j, dbh, idesc, true);
          }
          umma_commit(&st_empty[s]);
          if (++s == p.NST) { s = 0; ph ^= 1; }
        }
        umma_commit(acc_full);
      }
    }
  } else {
    // ================================================================= epilogue: + bias, NCHW rows
    const int q = warp & 3;
    int t = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++t) {
      const int pt = tile % p.nptile, t2 = tile / p.nptile;
      const int g = t2 % p.G, b = t2 / p.G;
      const int pix = pt * 128 + 32 * q + lane;
      mbar_wait(acc_full, t & 1);
      tc_fence_after();
      const uint32_t trow = tmem + (static_cast<uint32_t>(32 * q) << 16);
      float* orow = p.out + (static_cast<size_t>(b) * p.Cout + static_cast<size_t>(g) * Cog) * M + pix;
      for (int cb = 0; cb < Cog; cb += 32) {
        uint32_t raw[32];
        tmem_ld32(trow + cb, raw);
        tmem_ld_wait();
        if (pix < M) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (cb + j < Cog) orow[static_cast<size_t>(cb + j) * M] = __uint_as_float(raw[j]) + __ldg(p.bias + g * Cog + cb + j);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_empty);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == DT_WARP_MMA) tmem_free(tmem, 512);
  (void)Kpad;
}

static bool dcn_tc_plan(DcnTcParams& p, size_t* smem_out) {
  const int T = p.kh * p.kw;
  const int Cig = p.Cin / p.G, Cog = p.Cout / p.G;
  if (Cog > 256 || Cog < 8) return false;
  p.K = Cig * T;
  p.nchunk = ceil_div(p.K, DT_KC);
  p.N = ceil_div(Cog, 16) * 16;
  const size_t desc = static_cast<size_t>(p.dg) * T * 128 * 32;
  const size_t stage = static_cast<size_t>(2) * p.N * 128;
  if (desc + stage + 256 > 232448) return false;     // one stage still works (no load/MMA overlap)
  p.NST = static_cast<int>(min(static_cast<size_t>(DT_MAXST), (232448 - desc - 256) / stage));
  *smem_out = p.NST * stage + desc + 256;
  const int M = p.Ho * p.Wo;
  p.nptile = ceil_div(M, 128);
  const long long nt = 1LL * p.B * p.G * p.nptile;
  if (nt >= (1LL << 30)) return false;
  p.ntiles = static_cast<int>(nt);
  return true;
}

size_t dcn_tc_workspace_bytes(int Cin, int Cout, int kh, int kw, int group) {
  if (group <= 0 || Cin % group || Cout % group) return 0;
  const int K = (Cin / group) * kh * kw;
  const int Cog = Cout / group;
  if (Cog > 256 || Cog < 8) return 0;
  return static_cast<size_t>(2) * Cout * (ceil_div(K, DT_KC) * DT_KC) * sizeof(__nv_bfloat16);
}

// p: filled like DcnParams (shapes, pointers); weight: fp32 [Cout][Cig][kh][kw]
int launch_dcn_tc(DcnTcParams p, const float* weight, void* workspace, size_t workspace_bytes, cudaStream_t stream,
                  bool* handled) {
  *handled = false;
  size_t smem = 0;
  if (!dcn_tc_plan(p, &smem)) return SF_OK;
  const size_t need = dcn_tc_workspace_bytes(p.Cin, p.Cout, p.kh, p.kw, p.G);
  if (workspace == nullptr || workspace_bytes < need || (reinterpret_cast<uintptr_t>(workspace) & 127) != 0) return SF_OK;
  const int Kpad = p.nchunk * DT_KC;
  __nv_bfloat16* ws = static_cast<__nv_bfloat16*>(workspace);
  {
    const size_t n = static_cast<size_t>(p.Cout) * Kpad;
    const int grid = static_cast<int>(min(static_cast<size_t>(sm_count()) * 8, (n + 255) / 256));
    dcn_split_weights_kernel<<<grid, 256, 0, stream>>>(weight, ws, p.Cout, p.K, Kpad, p.Cin / p.G, p.kh * p.kw);
    SF_LAUNCH_CHECK();
  }
  CUtensorMap mW;
  int rc = make_tmap_3d_bf16_sw128(&mW, ws, Kpad, p.Cout, 2, p.N, 1);
  if (rc) return rc;
  SF_CUDA(cudaFuncSetAttribute(dcn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  const int grid = min(p.ntiles, sm_count());
  dcn_tc_kernel<<<grid, DT_THREADS, smem, stream>>>(mW, p);
  SF_LAUNCH_CHECK();
  *handled = true;
  return SF_OK;
}

}  // namespace sf

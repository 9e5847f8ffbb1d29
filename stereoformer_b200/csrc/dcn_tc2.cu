// Deformable convolution forward, tensor-core path over a CHANNELS-LAST copy of the input  (SURVEY 8a row a10).
//
// Reference: deform/src/cuda/modulated_deform_conv_cuda.cu:19-121 (im2col + addmm), modulated_deform_im2col_cuda.cuh:25-54,
// 128-194 (the sampling rule).  Same implicit GEMM as dcn_tc.cu (M = 128 output pixels, N = Cout/g, K = Cin/g * kh*kw, bf16
// hi/mid split, fp32 accumulation in TMEM); what changes is how the column values are produced.  dcn_tc.cu gathers from
// NCHW with thread = pixel: every load instruction of a warp touches ~20 sectors in ~16 cache lines and the L1 tag stage
// (one line per cycle) is the limiter (measured: 218k cycles per 128-pixel tile at 128 -> 128 channels).  Here
//   * a pre-pass writes the input once as [B][H*W][Cin] (dcn_nchw_to_nhwc_kernel, a tiled transpose);
//   * a gather warp works on TWO pixels at a time, 16 lanes each, a lane owning 4 consecutive channels of the chunk's 64:
//     the four bilinear corners are four LDG.128 per lane, 256 contiguous bytes per pixel and corner (4 cache lines per
//     instruction instead of ~16, and every byte of a line is used);
//   * the interpolated values (x mask) are split into bf16 hi + mid and stored as the A operand in shared memory
//     (K-major SWIZZLE_128B rows = pixels; 8 bytes per lane, 128 contiguous bytes per pixel);
//   * the reduction index is ordered (64-channel block, tap, channel): the nine taps of a channel block re-read the same
//     ~3 input rows x 256 bytes, which stay in L1 (the weights are split in the same order, dcn_split_weights2_kernel);
//   * the accumulator is double-buffered in TMEM (2 x 256 columns), so the epilogue (+ bias, NCHW rows) of tile t runs
//     under the gathers and MMAs of tile t + 1.
// Sampling descriptors (corner offsets, corner weights x mask) are computed once per tile by the warp that uses them.
// A pixel tile is 8 output rows x 16 columns (L1 hit rate of the gathers 19 % -> 85 % against 128 pixels of one row).
// Component isolation at 128 -> 128 @ 248x360 (tuning build, whole call incl. the 22 us of pre-passes): 199 us complete,
// 108 us without the gather warps' work, 186 us without the MMAs, 148 us without the gather loads, 77 us with neither:
// the phases add up instead of overlapping -- with two operand stages the weight stream alone (32 KB per chunk from L2,
// 590 KB per tile) paces the loop at ~1.2k cycles per chunk.  Measured and not kept: 8 lanes per pixel with descriptors
// in registers (shuffles) and packed FFMA2 (fewer instructions, but spills under the 80-register cap of 704 threads:
// 191 us), an L1 prefetch of the next chunk's lines (evicts live lines: 240 us).
// Shapes: 32 channels per group (two taps share a 64-wide chunk) or a multiple of 64, channels per deformable group a
// multiple of 4, <= 256 output channels per group; anything else takes dcn_tc.cu / dcn.cu.  fp32 parity: 3-term split, ~2^-16 relative per product (tests: 1e-4 of the output magnitude).
#include <cuda_bf16.h>
#include <stdlib.h>

#include "common.cuh"
#include "tc_common.cuh"
#include "dcn_tc.cuh"
#include "dcn_gather.cuh"
#include "../../include/stereoformer_b200.h"

namespace sf {

constexpr int D2_KC = 64;                  // K values per chunk: 64 channels of one tap
constexpr int D2_NG = 16;                  // gather warps 1..16, 8 pixels of the tile each
constexpr int D2_WARP_MMA = 1 + D2_NG;     // 17
constexpr int D2_WARP_EPI = D2_WARP_MMA + 1;   // 18..21 (warp & 3 = TMEM lane quarter)
constexpr int D2_THREADS = (D2_WARP_EPI + 4) * 32;
constexpr int D2_MAXST = 4;
constexpr int D2_ABYTES = 2 * 16384;       // A stage: hi | mid, [128 rows][128 B] each (two stages)

// [B][C][HW] -> [B][HW][C]
__global__ void __launch_bounds__(256)
dcn_nchw_to_nhwc_kernel(const float* __restrict__ in, float* __restrict__ out, int C, int HW) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z;
  const int x0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x, ty = threadIdx.y;
  const float* src = in + static_cast<size_t>(b) * C * HW;
  float* dst = out + static_cast<size_t>(b) * HW * C;
#pragma unroll
  for (int k = 0; k < 32; k += 8) {
    const int c = c0 + ty + k, x = x0 + tx;
    tile[ty + k][tx] = (c < C && x < HW) ? __ldcs(src + static_cast<size_t>(c) * HW + x) : 0.f;   // streaming: dead after this pass
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 32; k += 8) {
    const int x = x0 + ty + k, c = c0 + tx;
    if (x < HW && c < C) dst[static_cast<size_t>(x) * C + c] = tile[tx][ty + k];
  }
}

// fp32 weights [Cout][Cig][T] -> bf16 hi / mid [2][Cout][K], K = nchunk * 64.  Chunk (cc, tg) holds 64 / CB taps of CB
// channels each: k = (cc * ntg + tg) * 64 + ti * CB + c % CB with tap t = tg * (64 / CB) + ti (zero for t >= T) and
// channel c = cc * 64 + c % CB  (CB = 64: one tap of a 64-channel block per chunk; CB = 16: four taps of all 16 channels)
__global__ void dcn_split_weights2_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ ws, int Cout, int K,
                                          int Cig, int T, int CB, int ntg) {
  const size_t n = static_cast<size_t>(Cout) * K;
  const int tpc = 64 / CB;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int k = static_cast<int>(i % K);
    const size_t o = i / K;
    const int ch = k >> 6, r = k & 63;
    const int cc = ch / ntg, tg = ch - cc * ntg;
    const int t = tg * tpc + r / CB, c = cc * 64 + r % CB;
    const float v = t < T ? w[(o * Cig + c) * T + t] : 0.f;
    const __nv_bfloat16 h = __float2bfloat16_rn(v);
    ws[i] = h;
    ws[n + i] = __float2bfloat16_rn(v - __bfloat162float(h));
  }
}

__global__ void __launch_bounds__(D2_THREADS, 1)
dcn_tc2_kernel(const __grid_constant__ CUtensorMap mapW, const DcnTcParams p, const float* __restrict__ inT) {
  using namespace sftc;
  extern __shared__ __align__(1024) unsigned char smem[];
  const int T = p.kh * p.kw;
  const int b_bytes = 2 * p.N * 128;                                // B hi | B mid: [N rows][128 B]
  unsigned char* sA = smem;                                         // [NA][A hi | A mid]
  unsigned char* sB = sA + p.NA * D2_ABYTES;                           // [NST][B hi | B mid]: the weight ring is deeper than the A ring
  // sampling descriptors [dg][T][128]: the four corner offsets (elements of inT; 16 B) or, where shared memory is short
  // (p.compact), the top-left corner | dx | dy << 1 (4 B: three more integer instructions per gathered pixel); then the
  // corner weights * mask (16 B)
  int* d_base = reinterpret_cast<int*>(sB + p.NST * b_bytes);
  int4* d_off = reinterpret_cast<int4*>(d_base);
  float4* d_w = reinterpret_cast<float4*>(d_base + p.dg * T * 128 * (p.compact ? 1 : 4));
  uint64_t* bars = reinterpret_cast<uint64_t*>(d_w + p.dg * T * 128);
  uint64_t* b_full = bars;                      // [NST] TMA -> MMA
  uint64_t* b_empty = bars + D2_MAXST;          // [NST] MMA commit -> TMA producer
  uint64_t* a_full = bars + 2 * D2_MAXST;       // [NA]  gather warps (16) -> MMA
  uint64_t* a_empty = bars + 3 * D2_MAXST;      // [NA]  MMA commit -> gather warps
  uint64_t* acc_full = bars + 4 * D2_MAXST;     // [2] MMA commit -> epilogue
  uint64_t* acc_empty = acc_full + 2;           // [2] epilogue (4 warps) -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.NST; ++s) { mbar_init(&b_full[s], 1); mbar_init(&b_empty[s], 1); }
    for (int i = 0; i < p.NA; ++i) { mbar_init(&a_full[i], D2_NG); mbar_init(&a_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 4); }
    fence_mbar_init();
    tma_prefetch_desc(&mapW);
  }
  if (warp == D2_WARP_MMA) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const int M = p.Ho * p.Wo;
  const int Cig = p.Cin / p.G, Cog = p.Cout / p.G;

  if (warp == 0) {
    // ================================================================= weight producer (TMA)
    if (lane == 0) {
      int s = 0, ph = 1;
      for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        const int g = (tile / p.nptile) % p.G;
        for (int ch = 0; ch < p.nchunk; ++ch) {
          mbar_wait(&b_empty[s], ph);
          unsigned char* dst = sB + s * b_bytes;
          mbar_expect_tx(&b_full[s], b_bytes);
          tma_load_3d(dst, &mapW, &b_full[s], ch * D2_KC, g * Cog, 0);
          tma_load_3d(dst + p.N * 128, &mapW, &b_full[s], ch * D2_KC, g * Cog, 1);
          if (++s == p.NST) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp < D2_WARP_MMA) {
    // ================================================================= gather warps: column values -> A operand
    // (dcn_gather.cuh: 16 lanes per pixel, lane l16 fills K positions 4*l16 .. 4*l16+3 of the chunk row = one tap of the
    // chunk, 4 channels; a warp owns rows 8*gw .. 8*gw+7 = half a row of the 8 x 16 pixel tile, whose sampling footprint
    // with offsets of a few pixels is ~4x smaller than that of 128 pixels of one row: L1 hit rate of the gathers 19 % ->
    // 85 %; the sampling descriptors are warp-private, no block barrier)
    const int gw = warp - 1;
    const int lpt = p.CB >> 2, cl = 4 * ((lane & 15) % lpt);
    const int cpdg = p.Cin / p.dg;
    const int HW = p.H * p.W;
    int s = 0, ph = 1;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
      const int pt = tile % p.nptile, t2 = tile / p.nptile;
      const int g = t2 % p.G, b = t2 / p.G;
      const int ty = pt / p.tiles_x, tx = pt - ty * p.tiles_x;
      __syncwarp();
      if (p.compact) d2g_build_descriptors<true>(p, b, ty, tx, gw, lane, d_base, d_off, d_w);
      else d2g_build_descriptors<false>(p, b, ty, tx, gw, lane, d_base, d_off, d_w);
      __syncwarp();
      const float* in_b = inT + static_cast<size_t>(b) * HW * p.Cin + static_cast<size_t>(g) * Cig + cl;
      int dgi = 0;
      for (int ch = 0; ch < p.nchunk; ++ch) {
        const int cc = ch / p.ntg, tg = ch - cc * p.ntg;
        if (tg == 0) dgi = (g * Cig + cc * 64 + cl) / cpdg;    // deformable group of THIS lane's 4 channels (per lane: a
                                                               // 64-channel block may span several deformable groups)
        mbar_wait(&a_empty[s], ph);
        unsigned char* a_hi = sA + s * D2_ABYTES;
        if (p.compact) d2g_gather_chunk<true>(p, in_b, cc, tg, dgi, gw, lane, d_base, d_off, d_w, a_hi);
        else d2g_gather_chunk<false>(p, in_b, cc, tg, dgi, gw, lane, d_base, d_off, d_w, a_hi);
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&a_full[s]);
        if (++s == p.NA) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == D2_WARP_MMA) {
    // ================================================================= MMA issuer
    if (lane == 0) {
      const uint32_t idesc = make_idesc_bf16(128, p.N, 0, 0);   // A and B: K-major shared memory
      int s = 0, ph = 0, sa = 0, pha = 0, t = 0;
      for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++t) {
        const int buf = t & 1;
        mbar_wait(&acc_empty[buf], ((t >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t acc = tmem + buf * 256;
        for (int ch = 0; ch < p.nchunk; ++ch) {
          mbar_wait(&b_full[s], ph);
          mbar_wait(&a_full[sa], pha);
          tc_fence_after();
          const uint32_t a_hi = smem_u32(sA + sa * D2_ABYTES), a_mid = a_hi + 16384;
          const uint32_t b_hi = smem_u32(sB + s * b_bytes), b_mid = b_hi + p.N * 128;
#pragma unroll
          for (int j = 0; j < D2_KC / 16; ++j) {
            const uint64_t dah = make_smem_desc(a_hi + j * 32, 16, 1024);
            const uint64_t dam = make_smem_desc(a_mid + j * 32, 16, 1024);
            const uint64_t dbh = make_smem_desc(b_hi + j * 32, 16, 1024);
            const uint64_t dbm = make_smem_desc(b_mid + j * 32, 16, 1024);
            umma_bf16(acc, dah, dbh, idesc, (ch | j) != 0);
            umma_bf16(acc, dah, dbm, idesc, true);
            umma_bf16(acc, dam, dbh, idesc, true);
          }
          umma_commit(&a_empty[sa]);
          umma_commit(&b_empty[s]);
          if (++sa == p.NA) { sa = 0; pha ^= 1; }
          if (++s == p.NST) { s = 0; ph ^= 1; }
        }
        umma_commit(&acc_full[buf]);
      }
    }
  } else {
    // ================================================================= epilogue: + bias, NCHW rows
    const int q = warp & 3;
    int t = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++t) {
      const int pt = tile % p.nptile, t2 = tile / p.nptile;
      const int g = t2 % p.G, b = t2 / p.G;
      const int ty = pt / p.tiles_x, tx = pt - ty * p.tiles_x;
      const int rr = 32 * q + lane;
      const int yo = ty * 8 + (rr >> 4), xo = tx * 16 + (rr & 15);
      const int pix = yo * p.Wo + xo;
      const bool live = yo < p.Ho && xo < p.Wo;
      const int buf = t & 1;
      mbar_wait(&acc_full[buf], (t >> 1) & 1);
      tc_fence_after();
      const uint32_t trow = tmem + buf * 256 + (static_cast<uint32_t>(32 * q) << 16);
      float* orow = p.out + (static_cast<size_t>(b) * p.Cout + static_cast<size_t>(g) * Cog) * M + pix;
      for (int cb = 0; cb < Cog; cb += 32) {
        uint32_t raw[32];
        tmem_ld32(trow + cb, raw);
        tmem_ld_wait();
        if (live) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (cb + j < Cog) orow[static_cast<size_t>(cb + j) * M] = __uint_as_float(raw[j]) + __ldg(p.bias + g * Cog + cb + j);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[buf]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == D2_WARP_MMA) tmem_free(tmem, 512);
}

bool dcn_tc2_plan(DcnTcParams& p, size_t* smem_out, bool allow16) {
  const int T = p.kh * p.kw;
  if (p.G <= 0 || p.dg <= 0 || p.Cin % p.G || p.Cout % p.G || p.Cin % p.dg) return false;
  const int Cig = p.Cin / p.G, Cog = p.Cout / p.G;
  // 16 channels per group also work (four taps per chunk) but the specialised FFMA kernel is faster there (29 vs 40 us at
  // the shipped 8 x 16 -> 16 x 68 x 120 shape: three launches and the per-tile descriptor phase dominate), so it is not taken
  if (Cig != 32 && Cig % 64 != 0 && !(allow16 && Cig == 16)) return false;   // (the backward kernels do take 16: see there)
  p.CB = min(Cig, 64);
  if ((p.Cin / p.dg) % 4 != 0 || Cog > 256 || Cog < 8) return false;   // a lane's 4 channels share a deformable group
  if (1LL * p.H * p.W * p.Cin >= (1LL << 30)) return false;          // corner offsets are unsigned 32-bit byte offsets
  p.ntg = ceil_div(T, 64 / p.CB);
  p.nchunk = max(1, Cig / 64) * p.ntg;
  p.K = p.nchunk * D2_KC;                                            // padded with zero taps when 64 / CB does not divide T
  p.N = ceil_div(Cog, 16) * 16;
  const size_t bst = static_cast<size_t>(2) * p.N * 128;             // one weight stage
  // The A ring has two stages (the gathers of one chunk under the MMAs of the previous one) and the weight ring two of
  // its own.  Deeper weight rings measured slower at 128 -> 128 (2 / 3 / 4 stages: 175 / 181 / 185 us: the gathers, not
  // the weight stream, set the pace there; a deeper A ring changes nothing either, 2 / 3 / 4 stages: 185 / 186 / 187 us).
  // The 32-byte descriptors are kept where everything fits; at 256 output
  // channels per group only the 20-byte form leaves room for the second weight stage (one stage: 502 us, two: 387 us).
  p.NST = 2;
  p.NA = 2;
  p.compact = 0;
  if (const char* e = tuning_env("SF_DCN2_NA")) p.NA = max(2, min(D2_MAXST, atoi(e)));
  if (const char* e = tuning_env("SF_DCN2_COMPACT")) p.compact = atoi(e);
  size_t desc = static_cast<size_t>(p.dg) * T * 128 * (p.compact ? 20 : 32);
  size_t fixed = p.NA * static_cast<size_t>(D2_ABYTES) + desc + 512;
  if (fixed + p.NST * bst > 232448) {
    p.compact = 1;
    desc = static_cast<size_t>(p.dg) * T * 128 * 20;
    fixed = p.NA * static_cast<size_t>(D2_ABYTES) + desc + 512;
    if (fixed + p.NST * bst > 232448) p.NST = 1;
    if (fixed + p.NST * bst > 232448) return false;
  }
  if (const char* e = tuning_env("SF_DCN2_NST"))
    p.NST = static_cast<int>(min(static_cast<size_t>(max(1, atoi(e))), min(static_cast<size_t>(D2_MAXST), (232448 - fixed) / bst)));
  *smem_out = fixed + p.NST * bst;
  p.tiles_x = ceil_div(p.Wo, 16);
  p.nptile = p.tiles_x * ceil_div(p.Ho, 8);
  const long long nt = 1LL * p.B * p.G * p.nptile;
  if (nt >= (1LL << 30)) return false;
  p.ntiles = static_cast<int>(nt);
  return true;
}

static size_t align256(size_t x) { return (x + 255) & ~static_cast<size_t>(255); }

// [B][C][HW] -> [B][HW][C] (the channels-last copy the gathers read)
int launch_dcn_nhwc(const float* in, float* out, int B, int C, int HW, cudaStream_t stream) {
  dim3 tg(ceil_div(HW, 32), ceil_div(C, 32), B);
  SF_CHECK_ARG(tg.y <= 65535 && tg.z <= 65535, "dcn: too many channels or images for the channels-last pass");
  dcn_nchw_to_nhwc_kernel<<<tg, dim3(32, 8), 0, stream>>>(in, out, C, HW);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

// weights (bf16 hi/mid) + the channels-last copy of the input; 0 when the shape does not take this path
size_t dcn_tc2_workspace_bytes(int B, int Cin, int H, int W, int Cout, int kh, int kw, int group, int deformable_group) {
  DcnTcParams p{};
  p.B = B; p.Cin = Cin; p.H = H; p.W = W; p.Cout = Cout; p.kh = kh; p.kw = kw; p.G = group; p.dg = deformable_group;
  p.Ho = 1; p.Wo = 1;
  size_t smem = 0;
  if (B <= 0 || H <= 0 || W <= 0 || !dcn_tc2_plan(p, &smem)) return 0;
  return align256(static_cast<size_t>(2) * Cout * p.K * sizeof(__nv_bfloat16)) +
         static_cast<size_t>(B) * H * W * Cin * sizeof(float);
}

int launch_dcn_tc2(DcnTcParams p, const float* weight, void* workspace, size_t workspace_bytes, cudaStream_t stream,
                   bool* handled) {
  *handled = false;
  size_t smem = 0;
  if (!dcn_tc2_plan(p, &smem)) return SF_OK;
  const size_t need = dcn_tc2_workspace_bytes(p.B, p.Cin, p.H, p.W, p.Cout, p.kh, p.kw, p.G, p.dg);
  if (workspace == nullptr || workspace_bytes < need || (reinterpret_cast<uintptr_t>(workspace) & 127) != 0) return SF_OK;
  if ((reinterpret_cast<uintptr_t>(p.in) & 15) != 0) return SF_OK;
  __nv_bfloat16* ws = static_cast<__nv_bfloat16*>(workspace);
  const size_t wbytes = align256(static_cast<size_t>(2) * p.Cout * p.K * sizeof(__nv_bfloat16));
  float* inT = reinterpret_cast<float*>(static_cast<unsigned char*>(workspace) + wbytes);
  {
    const size_t n = static_cast<size_t>(p.Cout) * p.K;
    const int grid = static_cast<int>(min(static_cast<size_t>(sm_count()) * 8, (n + 255) / 256));
    dcn_split_weights2_kernel<<<grid, 256, 0, stream>>>(weight, ws, p.Cout, p.K, p.Cin / p.G, p.kh * p.kw, p.CB, p.ntg);
    SF_LAUNCH_CHECK();
    const int rc2 = launch_dcn_nhwc(p.in, inT, p.B, p.Cin, p.H * p.W, stream);
    if (rc2) return rc2;
  }
  CUtensorMap mW;
  int rc = make_tmap_3d_bf16_sw128(&mW, ws, p.K, p.Cout, 2, p.N, 1);
  if (rc) return rc;
  SF_CUDA(cudaFuncSetAttribute(dcn_tc2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  const int grid = min(p.ntiles, sm_count());
  dcn_tc2_kernel<<<grid, D2_THREADS, smem, stream>>>(mW, p, inT);
  SF_LAUNCH_CHECK();
  *handled = true;
  return SF_OK;
}

}  // namespace sf

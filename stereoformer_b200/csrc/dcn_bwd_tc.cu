// Deformable convolution: weight gradient on tcgen05 tensor cores  (SURVEY 8f rank 3, row a10 backward).
//
// Reference: deform/src/cuda/modulated_deform_conv_cuda.cu:246-266 (im2col of the input, then grad_weight += grad_out x
// columns^T per image through at::addmm).  Here the column tile is never written out:
//   gW[o][k] = sum over output pixels  grad_out[o][pix] * col[pix][k],     k = (64-channel block, tap, channel)
// is M = 128 output channels (TMEM lanes), N = 64 column entries of one chunk, K = 128 pixels of a tile.
//   * the column tile [128 pixels][64 k] is produced exactly as in the forward (dcn_tc2.cu: channels-last input copy,
//     8 x 16 pixel tiles, 16 lanes per pixel, bf16 hi + mid, K-major SWIZZLE_128B rows = pixels); read as the B operand
//     it is MN-major (N = k contiguous, one 8-row group per 8 pixels);
//   * grad_out rows of the tile ([32 o][8 rows][16 px] TMA boxes, fp32) are split into bf16 hi + mid by 4 converter warps
//     into [o rows][128 px] K-major blocks: the A operand;
//   * a CTA owns a GROUP of up to 8 chunks (8 x 64 fp32 accumulator columns = all of TMEM) and a contiguous range of pixel
//     tiles; it accumulates hi*hi + hi*mid + mid*hi over its whole range and adds the [128 x 64 x chunks] result to
//     grad_weight with atomics at the end (one atomic per element and CTA: ~50 per element at 128 -> 128 @ 248x360).
// The gathers are the cost, as in the forward; grad_out is converted once per chunk group (3x at 128 -> 128).
// fp32 parity: 3-term split, ~2^-16 relative per product.
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_common.cuh"
#include "dcn_tc.cuh"
#include "dcn_gather.cuh"
#include "../../include/stereoformer_b200.h"

namespace sf {

constexpr int DW_NG = 16;                       // gather warps 1..16
constexpr int DW_WARP_MMA = 1 + DW_NG;          // 17
constexpr int DW_WARP_CVT = DW_WARP_MMA + 1;    // 18..21: grad_out converters, then the epilogue (warp & 3 = TMEM quarter)
constexpr int DW_THREADS = (DW_WARP_CVT + 4) * 32;
constexpr int DW_COL = 2 * 16384;               // one column stage: hi | mid, [128 px rows][128 B]
constexpr int DW_GB = 4 * 16384;                // grad_out bf16: hi [2 K blocks][128 o rows][128 B] | mid
constexpr int DW_GF = 16384;                    // one grad_out fp32 staging slot: [32 o][128 px]

struct DcnBwdTcParams {
  DcnTcParams f;          // shapes, offset / mask pointers, CB / ntg / nchunk / tiles_x of the forward plan
  float* gweight;         // [Cout][Cig][T]
  int cpg, ngrp;          // chunks per group (<= 8), chunk groups
  int nmt;                // 128-row blocks of output channels per convolution group
  int nrange;             // pixel-tile ranges per (group, row block, chunk group)
};

__global__ void __launch_bounds__(DW_THREADS, 1)
dcn_bwd_weight_tc_kernel(const __grid_constant__ CUtensorMap mapG, const DcnBwdTcParams q, const float* __restrict__ inT) {
  using namespace sftc;
  const DcnTcParams& p = q.f;
  extern __shared__ __align__(1024) unsigned char smem[];
  const int T = p.kh * p.kw;
  unsigned char* sCol = smem;                                      // [2][DW_COL]
  unsigned char* sGb = sCol + 2 * DW_COL;                          // [DW_GB]
  unsigned char* sGf = sGb + DW_GB;                                // [2][DW_GF]
  int* d_base = reinterpret_cast<int*>(sGf + 2 * DW_GF);           // [dg][T][128]
  float4* d_w = reinterpret_cast<float4*>(d_base + p.dg * T * 128);
  uint64_t* bars = reinterpret_cast<uint64_t*>(d_w + p.dg * T * 128);
  uint64_t* col_full = bars;            // [2] gather warps (16) -> MMA
  uint64_t* col_empty = bars + 2;       // [2] MMA commit -> gather warps
  uint64_t* gf_full = bars + 4;         // [2] TMA -> converters
  uint64_t* gf_empty = bars + 6;        // [2] converters (4 warps) -> TMA
  uint64_t* gb_full = bars + 8;         // converters (4 warps) -> MMA, once per tile
  uint64_t* gb_empty = bars + 9;        // MMA commit -> converters, once per tile
  uint64_t* acc_done = bars + 10;       // MMA commit -> epilogue
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 11);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&col_full[i], DW_NG); mbar_init(&col_empty[i], 1);
      mbar_init(&gf_full[i], 1); mbar_init(&gf_empty[i], 4);
    }
    mbar_init(gb_full, 4); mbar_init(gb_empty, 1); mbar_init(acc_done, 1);
    fence_mbar_init();
    tma_prefetch_desc(&mapG);
  }
  if (warp == DW_WARP_MMA) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const int Cig = p.Cin / p.G, Cog = p.Cout / p.G;

  // unit = (convolution group g, 128-row block mt of its output channels, chunk group kg); range r of its pixel tiles
  const int unit = blockIdx.x / q.nrange, r = blockIdx.x - unit * q.nrange;
  const int kg = unit % q.ngrp, u2 = unit / q.ngrp;
  const int mt = u2 % q.nmt, g = u2 / q.nmt;
  const int ch0 = kg * q.cpg, nch = min(q.cpg, p.nchunk - ch0);
  const int NT = p.B * p.nptile;
  const int tb = static_cast<int>(static_cast<long long>(r) * NT / q.nrange);
  const int te = static_cast<int>(static_cast<long long>(r + 1) * NT / q.nrange);
  const int orow0 = g * Cog + mt * 128;            // first output channel (within an image) of this CTA's rows

  if (warp == 0) {
    // ================================================================= grad_out producer (TMA)
    if (lane == 0) {
      int s = 0, ph = 1;
      for (int tile = tb; tile < te; ++tile) {
        const int b = tile / p.nptile, pt = tile - b * p.nptile;
        const int ty = pt / p.tiles_x, tx = pt - ty * p.tiles_x;
        for (int oq = 0; oq < 4; ++oq) {
          mbar_wait(&gf_empty[s], ph);
          mbar_expect_tx(&gf_full[s], DW_GF);
          tma_load_3d(sGf + s * DW_GF, &mapG, &gf_full[s], tx * 16, ty * 8, b * p.Cout + orow0 + 32 * oq);
          if (++s == 2) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp < DW_WARP_MMA) {
    // ================================================================= gather warps: column chunks (as the forward)
    const int gw = warp - 1;
    const int lpt = p.CB >> 2, cl = 4 * ((lane & 15) % lpt);
    const int cpdg = p.Cin / p.dg;
    const int HW = p.H * p.W;
    int s = 0, ph = 1;
    for (int tile = tb; tile < te; ++tile) {
      const int b = tile / p.nptile, pt = tile - b * p.nptile;
      const int ty = pt / p.tiles_x, tx = pt - ty * p.tiles_x;
      __syncwarp();
      d2g_build_descriptors<true>(p, b, ty, tx, gw, lane, d_base, nullptr, d_w);
      __syncwarp();
      const float* in_b = inT + static_cast<size_t>(b) * HW * p.Cin + static_cast<size_t>(g) * Cig + cl;
      for (int lc = 0; lc < nch; ++lc) {
        const int ch = ch0 + lc;
        const int cc = ch / p.ntg, tg = ch - cc * p.ntg;
        const int dgi = (g * Cig + cc * 64 + cl) / cpdg;
        mbar_wait(&col_empty[s], ph);
        d2g_gather_chunk<true>(p, in_b, cc, tg, dgi, gw, lane, d_base, nullptr, d_w, sCol + s * DW_COL);
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&col_full[s]);
        if (++s == 2) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == DW_WARP_MMA) {
    // ================================================================= MMA issuer
    if (lane == 0) {
      const uint32_t idesc = make_idesc_bf16(128, 64, 0, 1);     // A: K-major (grad_out rows), B: MN-major (column tile)
      const uint32_t g_hi = smem_u32(sGb), g_mid = g_hi + 2 * 16384;
      int s = 0, ph = 0, it = 0;
      for (int tile = tb; tile < te; ++tile, ++it) {
        mbar_wait(gb_full, it & 1);
        for (int lc = 0; lc < nch; ++lc) {
          mbar_wait(&col_full[s], ph);
          tc_fence_after();
          const uint32_t c_hi = smem_u32(sCol + s * DW_COL), c_mid = c_hi + 16384;
          const uint32_t acc = tmem + 64 * lc;
#pragma unroll
          for (int j = 0; j < 8; ++j) {                          // 8 k-steps of 16 pixels
            const uint32_t ao = (j >> 2) * 16384 + (j & 3) * 32;
            const uint64_t dah = make_smem_desc(g_hi + ao, 16, 1024);
            const uint64_t dam = make_smem_desc(g_mid + ao, 16, 1024);
            const uint64_t dbh = make_smem_desc(c_hi + j * 2048, 8192, 1024);
            const uint64_t dbm = make_smem_desc(c_mid + j * 2048, 8192, 1024);
            umma_bf16(acc, dah, dbh, idesc, (it | j) != 0);
            umma_bf16(acc, dah, dbm, idesc, true);
            umma_bf16(acc, dam, dbh, idesc, true);
          }
          umma_commit(&col_empty[s]);
          if (++s == 2) { s = 0; ph ^= 1; }
        }
        umma_commit(gb_empty);
      }
      umma_commit(acc_done);
    }
  } else {
    // ================================================================= grad_out converters, then the epilogue
    // 128 threads; per staged quarter (32 output channels x 128 pixels) a thread converts four (row, 8-pixel chunk) items:
    // a warp covers two rows x 16 chunks per step, i.e. 16-byte stores that fill whole 128-byte operand rows
    const int ct = threadIdx.x - DW_WARP_CVT * 32;
    int s = 0, ph = 0, it = 0;
    for (int tile = tb; tile < te; ++tile, ++it) {
      mbar_wait(gb_empty, (it & 1) ^ 1);                         // the MMAs of the previous tile have read the operand
      for (int oq = 0; oq < 4; ++oq) {
        mbar_wait(&gf_full[s], ph);
        const float* src = reinterpret_cast<const float*>(sGf + s * DW_GF);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int item = ct + 128 * k;                         // 512 items: row = item / 16, chunk = item % 16
          const int rl = item >> 4, chn = item & 15;
          const float4 a = *reinterpret_cast<const float4*>(src + rl * 128 + chn * 8);
          const float4 c = *reinterpret_cast<const float4*>(src + rl * 128 + chn * 8 + 4);
          const bool live = orow0 - g * Cog + 32 * oq + rl < Cog;          // rows past the group's channels: zeros
          const float v[8] = {a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w};
          uint32_t hi[4], mid[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float x0 = live ? v[2 * e] : 0.f, x1 = live ? v[2 * e + 1] : 0.f;
            const __nv_bfloat162 h = __floats2bfloat162_rn(x0, x1);
            const __nv_bfloat162 m = __floats2bfloat162_rn(x0 - __low2float(h), x1 - __high2float(h));
            hi[e] = *reinterpret_cast<const uint32_t*>(&h);
            mid[e] = *reinterpret_cast<const uint32_t*>(&m);
          }
          const int row = 32 * oq + rl;                          // operand row = output channel of the block
          unsigned char* dst = sGb + (chn >> 3) * 16384 + (row >> 3) * 1024 + (row & 7) * 128 + (((chn & 7) ^ (row & 7)) << 4);
          *reinterpret_cast<uint4*>(dst) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
          *reinterpret_cast<uint4*>(dst + 2 * 16384) = make_uint4(mid[0], mid[1], mid[2], mid[3]);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&gf_empty[s]);
        if (++s == 2) { s = 0; ph ^= 1; }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(gb_full);
    }
    // ---- epilogue: accumulator rows -> grad_weight (atomics: every pixel range of the unit adds its share)
    mbar_wait(acc_done, 0);
    tc_fence_after();
    const int qd = warp & 3;
    const int o = mt * 128 + 32 * qd + lane;                       // output channel within the group
    const uint32_t trow = tmem + (static_cast<uint32_t>(32 * qd) << 16);
    const int tpc = 64 / p.CB;
    for (int lc = 0; lc < nch; ++lc) {
      const int ch = ch0 + lc;
      const int cc = ch / p.ntg, tg = ch - cc * p.ntg;
#pragma unroll 1
      for (int hh = 0; hh < 2; ++hh) {
        uint32_t raw[32];
        tmem_ld32(trow + 64 * lc + 32 * hh, raw);
        tmem_ld_wait();
        if (o < Cog && te > tb) {
          float* wrow = q.gweight + (static_cast<size_t>(g) * Cog + o) * Cig * T;
#pragma unroll
          for (int jj = 0; jj < 32; ++jj) {
            const int rk = 32 * hh + jj;
            const int t = tg * tpc + rk / p.CB, c = cc * 64 + rk % p.CB;
            if (t < T) atomicAdd(wrow + c * T + t, __uint_as_float(raw[jj]));
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == DW_WARP_MMA) tmem_free(tmem, 512);
}

size_t dcn_bwd_tc_workspace_bytes(int B, int Cin, int H, int W, int Cout, int kh, int kw, int group, int deformable_group) {
  DcnTcParams p{};
  p.B = B; p.Cin = Cin; p.H = H; p.W = W; p.Cout = Cout; p.kh = kh; p.kw = kw; p.G = group; p.dg = deformable_group;
  p.Ho = 1; p.Wo = 1;
  size_t smem = 0;
  if (B <= 0 || H <= 0 || W <= 0 || !dcn_tc2_plan(p, &smem, true)) return 0;
  if (static_cast<size_t>(p.dg) * kh * kw * 128 * 20 + 2 * DW_COL + DW_GB + 2 * DW_GF + 256 > 232448) return 0;
  return static_cast<size_t>(B) * H * W * Cin * sizeof(float);
}

int launch_dcn_bwd_weight_tc(DcnTcParams p, const float* grad_out, float* grad_weight, void* workspace,
                             size_t workspace_bytes, cudaStream_t stream, bool have_input_copy, bool* handled) {
  *handled = false;
  const size_t need = dcn_bwd_tc_workspace_bytes(p.B, p.Cin, p.H, p.W, p.Cout, p.kh, p.kw, p.G, p.dg);
  if (need == 0 || workspace == nullptr || workspace_bytes < need || (reinterpret_cast<uintptr_t>(workspace) & 127) != 0)
    return SF_OK;
  if ((reinterpret_cast<uintptr_t>(grad_out) & 15) != 0 || p.Wo % 4 != 0) return SF_OK;      // TMA: 16-byte rows
  size_t smem_fwd = 0;
  if (!dcn_tc2_plan(p, &smem_fwd, true)) return SF_OK;
  const int T = p.kh * p.kw;
  DcnBwdTcParams q;
  q.f = p;
  q.f.compact = 1;
  q.gweight = grad_weight;
  q.ngrp = ceil_div(p.nchunk, 8);
  q.cpg = ceil_div(p.nchunk, q.ngrp);
  q.ngrp = ceil_div(p.nchunk, q.cpg);
  q.nmt = ceil_div(p.Cout / p.G, 128);
  const int units = p.G * q.nmt * q.ngrp;
  const int NT = p.B * p.nptile;
  q.nrange = max(1, min(NT, sm_count() / units));
  float* inT = static_cast<float*>(workspace);
  int rc = have_input_copy ? SF_OK : launch_dcn_nhwc(p.in, inT, p.B, p.Cin, p.H * p.W, stream);
  if (rc) return rc;
  CUtensorMap mG;
  rc = make_tmap_3d_f32(&mG, grad_out, p.Wo, p.Ho, static_cast<uint64_t>(p.B) * p.Cout, 16, 8, 32);
  if (rc) return rc;
  const size_t smem = 2 * DW_COL + DW_GB + 2 * DW_GF + static_cast<size_t>(p.dg) * T * 128 * 20 + 256;
  SF_CUDA(cudaFuncSetAttribute(dcn_bwd_weight_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  dcn_bwd_weight_tc_kernel<<<units * q.nrange, DW_THREADS, smem, stream>>>(mG, q, inT);
  SF_LAUNCH_CHECK();
  *handled = true;
  return SF_OK;
}

}  // namespace sf

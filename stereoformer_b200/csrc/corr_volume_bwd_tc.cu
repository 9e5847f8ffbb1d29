// Backward of the correlation cost volume on tcgen05 tensor cores  (SURVEY 8f rank 3, rows a1 / a2).
//
// Reference: what autograd derives for models/utils/build_cost_volume.py:39-47 and
// models/CrossAttentionCostVolume/GWC_Cost_Volume.py:14-24:
//   gL[c,x] = 1/Cg * sum_d G[d,x]   * R[c,x-d]          gR[c,x] = 1/Cg * sum_d G[d,x+d] * L[c,x+d]
//
// Per epipolar row both are products of a feature row block with a BAND matrix built from the volume gradient:
//   gL[c, x0+n] = sum_k R[c, x0-HL+k] * BL[n][k],   BL[n][k] = G[n+HL-k][x0+n]   (0 <= n+HL-k < D)
//   gR[c, x0+n] = sum_k L[c, x0+k]    * BR[n][k],   BR[n][k] = G[k-n][x0+k]      (0 <= k-n    < D)
// so M = 128 channels (TMEM lanes), N = output pixels, K = source pixels.  The band is only D wide: the 128 output
// pixels of a tile are four 32-pixel sub-tiles with a K window of 32 + HL source pixels each, i.e. the band matrix is
// stored as 4 x [32][32+HL] bf16 blocks (K-major SWIZZLE_128B B operands) and ~60 % of the zero blocks are never touched.
//
//   * a CTA works on ONE side (even CTAs gL, odd CTAs gR) so a band buffer always holds the same sparsity pattern: the
//     buffers are zeroed once and every tile rewrites exactly the band positions;
//   * warp 0: TMA producer -- the gradient tile [D][128+HL] of the row segment and the feature rows in stages of 32
//     channels x (128+HL+4) pixels (one TMEM lane quarter per stage; the +4 makes the row pitch an odd multiple of 16
//     bytes so that thread = channel reads its row with conflict-free 16-byte loads; out-of-image columns are zero-filled,
//     which is the reference's x < d / x + d >= W zeros);
//   * 8 converter warps (thread = channel, two warps per lane quarter, half of the pixels each): fp32 -> bf16 hi + mid,
//     written to TENSOR MEMORY (tcgen05.st) as the A operand (double-buffered);
//   * 4 band warps (thread = output pixel): G column -> bf16 hi + mid band entries;
//   * 1 MMA thread: per sub-tile (32+HL)/16 k-steps of hi*hi + hi*mid + mid*hi into a [128 ch x 128 px] fp32 accumulator;
//   * 4 epilogue warps: tcgen05.ld, x 1/Cg, [32 ch][32 px] swizzled staging tiles, one TMA store each.
// Algorithmic bytes per pixel: (4*Cg + D) * 4 per group (features read once, gradients written once; the volume gradient
// is read once per side).  Error of the 3-term split ~2^-16 per product, like the forward.
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_common.cuh"
#include "../../include/stereoformer_b200.h"

namespace sf {

constexpr int BT_NCONV = 8;                       // converter warps 1..8 (warp & 3 = TMEM lane quarter)
constexpr int BT_WARP_BAND = 1 + BT_NCONV;        // band warps 9..12 (one 32-pixel sub-tile each)
constexpr int BT_WARP_MMA = BT_WARP_BAND + 4;     // 13
constexpr int BT_WARP_EPI = BT_WARP_MMA + 1;      // 14..17 (warp & 3 = TMEM lane quarter)
constexpr int BT_THREADS = (BT_WARP_EPI + 4) * 32;
constexpr int BT_MAXNS = 8;
constexpr int BT_A_BASE = 128;                    // TMEM: accumulator at columns [0,128), A buffers behind it
constexpr int BT_OUT_BYTES = 4 * 2 * 4096;        // epilogue staging: 4 warps x 2 x [32 ch][32 px] fp32

struct CorrBwdTcParams {
  int Cg, H, W, D, nBG;
  int HL;            // halo of the K window: multiple of 16, >= D-1, <= 64
  int Kg;            // 128 + HL source pixels per tile
  int P;             // Kg + 4: row pitch of a staged feature row (floats)
  int nks;           // k-steps (16 source pixels) per sub-tile: (32 + HL) / 16
  int nkb;           // 64-element K blocks per sub-tile: 1 (HL <= 32) or 2
  int nseg, ntiles, ncb;
  int NS, NG, NB;    // feature stages, gradient-tile stages, band buffers
  int stage_bytes, g_bytes, g_pitch, band_bytes;   // g_bytes: bytes of one gradient box, g_pitch: its 128-byte aligned slot
  int sides;         // bit 0: gL wanted, bit 1: gR wanted
};

__device__ __forceinline__ void bt_tma_store_3d(const CUtensorMap* map, const void* src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(map)), "r"(sftc::smem_u32(src)), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}

__global__ void __launch_bounds__(BT_THREADS, 1)
corr_volume_bwd_tc_kernel(const __grid_constant__ CUtensorMap mapL, const __grid_constant__ CUtensorMap mapR,
                          const __grid_constant__ CUtensorMap mapG, const __grid_constant__ CUtensorMap mapGL,
                          const __grid_constant__ CUtensorMap mapGR, const CorrBwdTcParams p) {
  using namespace sftc;
  extern __shared__ __align__(1024) unsigned char smem[];
  unsigned char* sBand = smem;                                     // [NB][4 sub-tiles][hi | mid][nkb][32 rows][128 B]
  unsigned char* sOut = sBand + p.NB * p.band_bytes;               // [4 warps][2][32 ch][128 B]
  unsigned char* sG = sOut + BT_OUT_BYTES;                         // [NG][D][Kg] fp32
  unsigned char* sStage = sG + p.NG * p.g_pitch;                   // [NS][32 ch][P] fp32
  uint64_t* bars = reinterpret_cast<uint64_t*>(sStage + p.NS * p.stage_bytes);
  uint64_t* stg_full = bars;                    // [NS] TMA -> converters
  uint64_t* stg_empty = bars + BT_MAXNS;        // [NS] the 2 converter warps of the quarter -> TMA
  uint64_t* g_full = bars + 2 * BT_MAXNS;       // [2]  TMA -> band warps
  uint64_t* g_empty = g_full + 2;               // [2]  band warps (4) -> TMA
  uint64_t* band_full = g_full + 4;             // [2]  band warps (4) -> MMA
  uint64_t* band_empty = g_full + 6;            // [2]  MMA commit -> band warps
  uint64_t* a_full = g_full + 8;                // [2]  converters (8 warps) -> MMA
  uint64_t* a_empty = g_full + 10;              // [2]  MMA commit -> converters
  uint64_t* d_full = g_full + 12;               // [2]  MMA commit -> epilogue, per 64-pixel half of the accumulator
  uint64_t* d_empty = g_full + 14;              // [2]  epilogue (4 warps) -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(g_full + 16);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool both = p.sides == 3;
  const int side = both ? (blockIdx.x & 1) : (p.sides >> 1);       // 0: gL (features = R), 1: gR (features = L)
  const int t0 = both ? (blockIdx.x >> 1) : blockIdx.x;
  const int tstep = both ? (gridDim.x >> 1) : gridDim.x;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.NS; ++s) { mbar_init(&stg_full[s], 1); mbar_init(&stg_empty[s], 2); }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&g_full[s], 1); mbar_init(&g_empty[s], 4);
      mbar_init(&band_full[s], 4); mbar_init(&band_empty[s], 1);
      mbar_init(&a_full[s], BT_NCONV); mbar_init(&a_empty[s], 1);
      mbar_init(&d_full[s], 1); mbar_init(&d_empty[s], 4);
    }
    fence_mbar_init();
    tma_prefetch_desc(side ? &mapL : &mapR);
    tma_prefetch_desc(&mapG);
    tma_prefetch_desc(side ? &mapGR : &mapGL);
  }
  if (warp == BT_WARP_MMA) tmem_alloc(tmem_slot, 512);
  {  // band buffers start as zeros: only the band positions are ever written
    uint4* z = reinterpret_cast<uint4*>(sBand);
    for (int i = threadIdx.x; i < p.NB * p.band_bytes / 16; i += BT_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
    fence_proxy_async_smem();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // ================================================================= TMA producer
    if (lane == 0) {
      const CUtensorMap* mapF = side ? &mapL : &mapR;
      int s = 0, ph = 1, it = 0;
      for (int tile = t0; tile < p.ntiles; tile += tstep, ++it) {
        const int seg = tile % p.nseg, t2 = tile / p.nseg;
        const int y = t2 % p.H, bg = t2 / p.H;
        const int x0 = seg * 128;
        const int gs = it % p.NG;
        mbar_wait(&g_empty[gs], ((it / p.NG) & 1) ^ 1);
        mbar_expect_tx(&g_full[gs], p.g_bytes);
        tma_load_3d(sG + gs * p.g_pitch, &mapG, &g_full[gs], x0, y, bg * p.D);
        const int xf = side ? x0 : x0 - p.HL;
        for (int cb = 0; cb < p.ncb; ++cb) {
          const int nq = min(4, (p.Cg - cb * 128) >> 5);
          for (int q = 0; q < nq; ++q) {
            mbar_wait(&stg_empty[s], ph);
            mbar_expect_tx(&stg_full[s], p.stage_bytes);
            tma_load_3d(sStage + s * p.stage_bytes, mapF, &stg_full[s], xf, y, bg * p.Cg + cb * 128 + 32 * q);
            if (++s == p.NS) { s = 0; ph ^= 1; }
          }
        }
      }
    }
  } else if (warp < BT_WARP_BAND) {
    // ================================================================= converters: feature rows -> tensor memory (A)
    const int q = warp & 3, kh = (warp - 1) >> 2;
    const int nst = p.Kg >> 4;                          // k-steps of the whole tile
    const int st0 = kh ? (nst + 1) >> 1 : 0, st1 = kh ? nst : (nst + 1) >> 1;
    const uint32_t lane_base = static_cast<uint32_t>(32 * q) << 16;
    const int half = p.Kg >> 1;
    int s = 0, ph = 0, si = 0;
    for (int tile = t0; tile < p.ntiles; tile += tstep) {
      for (int cb = 0; cb < p.ncb; ++cb, ++si) {
        const int nq = min(4, (p.Cg - cb * 128) >> 5);
        const int ab = si & 1;
        mbar_wait(&a_empty[ab], ((si >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t ta = tmem + BT_A_BASE + ab * p.Kg + lane_base;
        for (int qq = 0; qq < nq; ++qq) {
          if (qq == q) {
            mbar_wait(&stg_full[s], ph);
            const float4* row = reinterpret_cast<const float4*>(sStage + s * p.stage_bytes) + lane * (p.P >> 2);
            for (int st = st0; st < st1; ++st) {
              uint32_t hi[8], mid[8];
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const float4 v = row[4 * st + i];
                const __nv_bfloat162 h0 = __floats2bfloat162_rn(v.x, v.y), h1 = __floats2bfloat162_rn(v.z, v.w);
                const __nv_bfloat162 m0 = __floats2bfloat162_rn(v.x - __low2float(h0), v.y - __high2float(h0));
                const __nv_bfloat162 m1 = __floats2bfloat162_rn(v.z - __low2float(h1), v.w - __high2float(h1));
                hi[2 * i] = *reinterpret_cast<const uint32_t*>(&h0);
                hi[2 * i + 1] = *reinterpret_cast<const uint32_t*>(&h1);
                mid[2 * i] = *reinterpret_cast<const uint32_t*>(&m0);
                mid[2 * i + 1] = *reinterpret_cast<const uint32_t*>(&m1);
              }
              tmem_st8(ta + 8 * st, hi);
              tmem_st8(ta + half + 8 * st, mid);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&stg_empty[s]);
          }
          if (++s == p.NS) { s = 0; ph ^= 1; }
        }
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&a_full[ab]);
      }
    }
  } else if (warp < BT_WARP_MMA) {
    // ================================================================= band warps: G tile -> band matrix (B operand)
    const int j = warp - BT_WARP_BAND;                  // sub-tile
    const int n = 32 * j + lane;                        // output pixel within the tile
    const int half_off = p.nkb * 4096;
    int it = 0;
    for (int tile = t0; tile < p.ntiles; tile += tstep, ++it) {
      const int gs = it % p.NG, bb = it % p.NB;
      mbar_wait(&g_full[gs], (it / p.NG) & 1);
      mbar_wait(&band_empty[bb], ((it / p.NB) & 1) ^ 1);
      const float* g = reinterpret_cast<const float*>(sG + gs * p.g_pitch);
      unsigned char* base = sBand + bb * p.band_bytes + j * 2 * half_off + lane * 128;
      const int sw = lane & 7;
      for (int d = 0; d < p.D; ++d) {
        const float v = side ? g[d * p.Kg + n + d] : g[d * p.Kg + n];
        const int kl = side ? lane + d : lane + p.HL - d;
        const __nv_bfloat16 h = __float2bfloat16_rn(v);
        const __nv_bfloat16 m = __float2bfloat16_rn(v - __bfloat162float(h));
        const int off = (kl >> 6) * 4096 + ((((kl & 63) >> 3) ^ sw) << 4) + (kl & 7) * 2;
        *reinterpret_cast<__nv_bfloat16*>(base + off) = h;
        *reinterpret_cast<__nv_bfloat16*>(base + half_off + off) = m;
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) { mbar_arrive(&band_full[bb]); mbar_arrive(&g_empty[gs]); }
    }
  } else if (warp == BT_WARP_MMA) {
    // ================================================================= MMA issuer
    if (lane == 0) {
      const uint32_t idesc = make_idesc_bf16(128, 32, 0, 0);        // A: TMEM (K-major), B: K-major shared memory
      const int half_off = p.nkb * 4096, halfA = p.Kg >> 1;
      int it = 0, si = 0;
      for (int tile = t0; tile < p.ntiles; tile += tstep, ++it) {
        const int bb = it % p.NB;
        mbar_wait(&band_full[bb], (it / p.NB) & 1);
        const uint32_t band = smem_u32(sBand + bb * p.band_bytes);
        for (int cb = 0; cb < p.ncb; ++cb, ++si) {
          const int ab = si & 1;
          mbar_wait(&a_full[ab], (si >> 1) & 1);
          const uint32_t a_hi = tmem + BT_A_BASE + ab * p.Kg, a_mid = a_hi + halfA;
          // the accumulator is handed over in two 64-pixel halves: the epilogue drains half 0 while half 1 is computed
          for (int h = 0; h < 2; ++h) {
            mbar_wait(&d_empty[h], (si & 1) ^ 1);
            tc_fence_after();
            for (int j = 2 * h; j < 2 * h + 2; ++j) {
              const uint32_t bj = band + j * 2 * half_off;
              for (int s = 0; s < p.nks; ++s) {
                const uint32_t boff = (s >> 2) * 4096 + (s & 3) * 32;
                const uint64_t dbh = make_smem_desc(bj + boff, 16, 1024);
                const uint64_t dbm = make_smem_desc(bj + half_off + boff, 16, 1024);
                const uint32_t ac = 16 * j + 8 * s;
                umma_bf16_ts(tmem + 32 * j, a_hi + ac, dbh, idesc, s != 0);
                umma_bf16_ts(tmem + 32 * j, a_hi + ac, dbm, idesc, true);
                umma_bf16_ts(tmem + 32 * j, a_mid + ac, dbh, idesc, true);
              }
            }
            umma_commit(&d_full[h]);
          }
          umma_commit(&a_empty[ab]);
        }
        umma_commit(&band_empty[bb]);
      }
    }
  } else {
    // ================================================================= epilogue: accumulator -> staging -> TMA store
    const int q = warp & 3;
    const uint32_t trow = tmem + (static_cast<uint32_t>(32 * q) << 16);
    const CUtensorMap* mapO = side ? &mapGR : &mapGL;
    unsigned char* stage = sOut + (warp - BT_WARP_EPI) * 2 * 4096;
    const float inv = 1.0f / static_cast<float>(p.Cg);
    const int sw = lane & 7;
    int si = 0, nst = 0;
    for (int tile = t0; tile < p.ntiles; tile += tstep) {
      const int seg = tile % p.nseg, t2 = tile / p.nseg;
      const int y = t2 % p.H, bg = t2 / p.H;
      const int x0 = seg * 128;
      for (int cb = 0; cb < p.ncb; ++cb, ++si) {
        const bool valid = cb * 128 + 32 * q < p.Cg;
        for (int h = 0; h < 2; ++h) {
          mbar_wait(&d_full[h], si & 1);
          tc_fence_after();
          uint32_t raw[2][32];
          tmem_ld32(trow + 64 * h, raw[0]);
          tmem_ld32(trow + 64 * h + 32, raw[1]);
          tmem_ld_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&d_empty[h]);          // the half is in registers: the next MMAs may overwrite it
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int xb = x0 + 64 * h + 32 * e;
            if (valid && xb < p.W) {
              unsigned char* buf = stage + (nst & 1) * 4096;
              if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
              __syncwarp();
              unsigned char* rowp = buf + lane * 128;
#pragma unroll
              for (int c = 0; c < 8; ++c) {
                float4 v;
                v.x = __uint_as_float(raw[e][4 * c]) * inv; v.y = __uint_as_float(raw[e][4 * c + 1]) * inv;
                v.z = __uint_as_float(raw[e][4 * c + 2]) * inv; v.w = __uint_as_float(raw[e][4 * c + 3]) * inv;
                *reinterpret_cast<float4*>(rowp + ((c ^ sw) << 4)) = v;
              }
              fence_proxy_async_smem();
              __syncwarp();
              if (lane == 0) {
                bt_tma_store_3d(mapO, buf, xb, y, bg * p.Cg + cb * 128 + 32 * q);
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
              }
              ++nst;
            }
          }
        }
      }
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  if (warp == BT_WARP_MMA) tmem_free(tmem, 512);
}

static bool corr_bwd_tc_plan(CorrBwdTcParams& p, int nBG, int Cg, int H, int W, int D, int sides) {
  p.Cg = Cg; p.H = H; p.W = W; p.D = D; p.nBG = nBG; p.sides = sides;
  if (Cg % 32 != 0 || W % 4 != 0 || D < 1 || D > 65) return false;
  if (1LL * nBG * Cg >= (1LL << 31) || 1LL * nBG * D >= (1LL << 31)) return false;
  p.HL = max(16, ceil_div(D - 1, 16) * 16);
  p.Kg = 128 + p.HL;
  p.P = p.Kg + 4;
  p.nks = (32 + p.HL) / 16;
  p.nkb = ceil_div(32 + p.HL, 64);
  p.nseg = ceil_div(W, 128);
  p.ncb = ceil_div(Cg, 128);
  const long long nt = 1LL * nBG * H * p.nseg;
  if (nt >= (1LL << 30)) return false;
  p.ntiles = static_cast<int>(nt);
  p.stage_bytes = 32 * p.P * 4;
  p.g_bytes = D * p.Kg * 4;
  p.g_pitch = ceil_div(p.g_bytes, 128) * 128;
  p.band_bytes = 4 * 2 * p.nkb * 4096;
  const int budget = 232448 - 1024 - BT_OUT_BYTES - 512;
  for (int cfg = 0; cfg < 3; ++cfg) {                 // prefer double-buffered bands and gradient tiles
    p.NB = cfg < 2 ? 2 : 1;
    p.NG = cfg < 1 ? 2 : 1;
    const int rest = budget - p.NB * p.band_bytes - p.NG * p.g_pitch;
    p.NS = rest > 0 ? min(BT_MAXNS, rest / p.stage_bytes) : 0;
    if (p.NS >= 4) return true;
  }
  return p.NS >= 2;
}

bool corr_bwd_tc_supported(const float* L, const float* R, const float* G, const float* gL, const float* gR, int nBG,
                           int Cg, int H, int W, int D) {
  if (((reinterpret_cast<uintptr_t>(L) | reinterpret_cast<uintptr_t>(R) | reinterpret_cast<uintptr_t>(G) |
        reinterpret_cast<uintptr_t>(gL) | reinterpret_cast<uintptr_t>(gR)) & 15) != 0)
    return false;
  CorrBwdTcParams p;
  return corr_bwd_tc_plan(p, nBG, Cg, H, W, D, 3);
}

int launch_corr_bwd_tc(const float* L, const float* R, const float* G, float* gL, float* gR, int nBG, int Cg, int H,
                       int W, int D, cudaStream_t stream) {
  CorrBwdTcParams p;
  const int sides = (gL ? 1 : 0) | (gR ? 2 : 0);
  SF_CHECK_ARG(corr_bwd_tc_plan(p, nBG, Cg, H, W, D, sides), "corr_volume_backward: shape does not fit the tcgen05 kernel");
  CUtensorMap mL, mR, mG, mGL, mGR;
  const uint64_t rows = static_cast<uint64_t>(nBG) * Cg;
  int rc = make_tmap_3d_f32(&mL, L, W, H, rows, p.P, 1, 32);
  if (rc) return rc;
  rc = make_tmap_3d_f32(&mR, R, W, H, rows, p.P, 1, 32);
  if (rc) return rc;
  rc = make_tmap_3d_f32(&mG, G, W, H, static_cast<uint64_t>(nBG) * D, p.Kg, 1, D);
  if (rc) return rc;
  rc = make_tmap_3d_f32_swz(&mGL, gL ? gL : gR, W, H, rows, 32, 1, 32, 128);
  if (rc) return rc;
  rc = make_tmap_3d_f32_swz(&mGR, gR ? gR : gL, W, H, rows, 32, 1, 32, 128);
  if (rc) return rc;
  const size_t smem = static_cast<size_t>(p.NB) * p.band_bytes + BT_OUT_BYTES + static_cast<size_t>(p.NG) * p.g_pitch +
                      static_cast<size_t>(p.NS) * p.stage_bytes + 512;
  int grid;
  if (sides == 3) grid = min(2 * p.ntiles, sm_count() & ~1);
  else grid = min(p.ntiles, sm_count());
  SF_CUDA(cudaFuncSetAttribute(corr_volume_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               static_cast<int>(smem)));
  corr_volume_bwd_tc_kernel<<<grid, BT_THREADS, smem, stream>>>(mL, mR, mG, mGL, mGR, p);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

}  // namespace sf

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
//   * 4 epilogue warps: tcgen05.ld, x 1/Cg, [32 ch][32 px] swizzled staging tiles, one TMA store each; the accumulator is
//     handed over in two 64-pixel halves, so half 0 drains while half 1 is still being computed;
//   * work units are (row tile, block of 128 channels) in contiguous ranges per CTA: the channel blocks of a tile share
//     the gradient tile and its band matrix; the gradient tiles are an independent TMA stream beside the feature ring.
// Measured at the bench shape (8 x 256 x 68 x 120, D = 24): 53 us = 0.78 of the HBM peak (the FFMA kernel: 204 us); the
// depth of the feature ring is the first-order lever (2 / 3 / 4 / 5 stages: 75.6 / 64.6 / 56.5 / 53.1 us).
// Algorithmic bytes per pixel: (4*Cg + D) * 4 per group (features read once, gradients written once; the volume gradient
// is read once per side).  Error of the 3-term split ~2^-16 per product, like the forward.
#include <cuda_bf16.h>
#include <stdlib.h>

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
constexpr int BT_OUT_TILE = 4096;                 // one epilogue staging tile: [32 ch][32 px] fp32 (4 warps x NO of them)

struct CorrBwdTcParams {
  int Cg, H, W, D, nBG;
  int HL;            // halo of the K window: multiple of 16, >= D-1, <= 64
  int Kg;            // 128 + HL source pixels per tile
  int P;             // Kg + 4: row pitch of a staged feature row (floats)
  int SW;            // output pixels per sub-tile (MMA N): 32 (16 only through the tuning hook)
  int nks;           // k-steps (16 source pixels) per sub-tile: (SW + HL) / 16
  int nkb;           // 64-element K blocks per sub-tile: ceil((SW + HL) / 64)
  int nseg, ntiles, ncb;
  int NS, NG, NB, NO;   // feature stages, gradient-tile stages, band buffers, staging tiles per epilogue warp (1, 2 or 4)
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
  unsigned char* sOut = sBand + p.NB * p.band_bytes;               // [4 warps][NO][32 ch][128 B]
  unsigned char* sG = sOut + 4 * p.NO * BT_OUT_TILE;                         // [NG][D][Kg] fp32
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

  // work units: (row tile, block of 128 channels); a CTA takes a CONTIGUOUS range of units of its side, so the channel
  // blocks of one row tile mostly share the gradient tile and its band matrix (rebuilt only when the tile changes)
  const int nunits = p.ntiles * p.ncb;
  const int ub = static_cast<int>(static_cast<long long>(t0) * nunits / tstep);
  const int ue = static_cast<int>(static_cast<long long>(t0 + 1) * nunits / tstep);
  if (warp == 0) {
    // ================================================================= TMA producer
    if (lane == 0) {
      const CUtensorMap* mapF = side ? &mapL : &mapR;
      // The gradient tiles are a second, independent stream: they are issued whenever their slot is free and never block
      // the feature ring (with one slot the next tile can only be fetched after the band warps have consumed the last).
      int gi = 0, gu = ub;
      auto pump_g = [&]() {
        if (gu >= ue) return;
        const int gs = gi % p.NG;
        if (!mbar_try_wait(&g_empty[gs], ((gi / p.NG) & 1) ^ 1)) return;
        const int tile = gu / p.ncb;
        const int seg = tile % p.nseg, t2 = tile / p.nseg;
        mbar_expect_tx(&g_full[gs], p.g_bytes);
        tma_load_3d(sG + gs * p.g_pitch, &mapG, &g_full[gs], seg * 128, t2 % p.H, (t2 / p.H) * p.D);
        ++gi;
        gu = (tile + 1) * p.ncb;                     // first unit of the next row tile
      };
      int s = 0, ph = 1;
      for (int u = ub; u < ue; ++u) {
        const int tile = u / p.ncb, cb = u - tile * p.ncb;
        const int seg = tile % p.nseg, t2 = tile / p.nseg;
        const int y = t2 % p.H, bg = t2 / p.H;
        const int xf = side ? seg * 128 : seg * 128 - p.HL;
        const int nq = min(4, (p.Cg - cb * 128) >> 5);
        pump_g();
        for (int q = 0; q < nq; ++q) {
          while (!mbar_try_wait(&stg_empty[s], ph)) pump_g();
          mbar_expect_tx(&stg_full[s], p.stage_bytes);
          tma_load_3d(sStage + s * p.stage_bytes, mapF, &stg_full[s], xf, y, bg * p.Cg + cb * 128 + 32 * q);
          if (++s == p.NS) { s = 0; ph ^= 1; }
        }
      }
      while (gu < ue) pump_g();
    }
  } else if (warp < BT_WARP_BAND) {
    // ================================================================= converters: feature rows -> tensor memory (A)
    const int q = warp & 3, kh = (warp - 1) >> 2;
    const int nst = p.Kg >> 4;                          // k-steps of the whole tile
    const int st0 = kh ? (nst + 1) >> 1 : 0, st1 = kh ? nst : (nst + 1) >> 1;
    const uint32_t lane_base = static_cast<uint32_t>(32 * q) << 16;
    const int half = p.Kg >> 1;
    int s = 0, ph = 0, it = 0;
    for (int u = ub; u < ue; ++u, ++it) {
      const int cb = u % p.ncb;
      const int nq = min(4, (p.Cg - cb * 128) >> 5);
      const int ab = it & 1;
      mbar_wait(&a_empty[ab], ((it >> 1) & 1) ^ 1);
      tc_fence_after();
      const uint32_t ta = tmem + BT_A_BASE + ab * p.Kg + lane_base;
      // every converter warp waits for EVERY stage of the unit in ring order, also those of the other quarters: a
      // parity wait is only meaningful for a waiter that has seen the previous phase of that slot complete, and slots
      // move between quarters whenever the ring depth is not a multiple of the quarters per unit
      for (int qq = 0; qq < nq; ++qq) {
        mbar_wait(&stg_full[s], ph);
        if (qq == q) {
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
  } else if (warp < BT_WARP_MMA) {
    // ================================================================= band warps: G tile -> band matrix (B operand)
    const int n = 32 * (warp - BT_WARP_BAND) + lane;    // output pixel within the tile
    const int sub = n / p.SW, nl = n - sub * p.SW;      // sub-tile and row within it
    const int kb_bytes = p.SW * 128;                    // one 64-element K block of a sub-tile half
    const int half_off = p.nkb * kb_bytes;
    const int sw = nl & 7;
    int it = 0;
    for (int u = ub; u < ue; u = (u / p.ncb + 1) * p.ncb, ++it) {      // one pass per row tile of the range
      const int gs = it % p.NG, bb = it % p.NB;
      mbar_wait(&g_full[gs], (it / p.NG) & 1);
      mbar_wait(&band_empty[bb], ((it / p.NB) & 1) ^ 1);
      const float* g = reinterpret_cast<const float*>(sG + gs * p.g_pitch);
      unsigned char* base = sBand + bb * p.band_bytes + sub * 2 * half_off + nl * 128;
      for (int d = 0; d < p.D; ++d) {
        const float v = side ? g[d * p.Kg + n + d] : g[d * p.Kg + n];
        const int kl = side ? nl + d : nl + p.HL - d;
        const __nv_bfloat16 h = __float2bfloat16_rn(v);
        const __nv_bfloat16 m = __float2bfloat16_rn(v - __bfloat162float(h));
        const int off = (kl >> 6) * kb_bytes + ((((kl & 63) >> 3) ^ sw) << 4) + (kl & 7) * 2;
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
      const uint32_t idesc = make_idesc_bf16(128, p.SW, 0, 0);      // A: TMEM (K-major), B: K-major shared memory
      const int kb_bytes = p.SW * 128, half_off = p.nkb * kb_bytes, halfA = p.Kg >> 1;
      const int nsub_half = 64 / p.SW;
      int it = 0, ib = -1;
      for (int u = ub; u < ue; ++u, ++it) {
        const int cb = u % p.ncb;
        if (u == ub || cb == 0) {                     // a new row tile: its band matrix
          ++ib;
          mbar_wait(&band_full[ib % p.NB], (ib / p.NB) & 1);
        }
        const int bb = ib % p.NB, ab = it & 1;
        mbar_wait(&a_full[ab], (it >> 1) & 1);
        const uint32_t band = smem_u32(sBand + bb * p.band_bytes);
        const uint32_t a_hi = tmem + BT_A_BASE + ab * p.Kg, a_mid = a_hi + halfA;
        // the accumulator is handed over in two 64-pixel halves: the epilogue drains half 0 while half 1 is computed
        for (int h = 0; h < 2; ++h) {
          mbar_wait(&d_empty[h], (it & 1) ^ 1);
          tc_fence_after();
          for (int j = h * nsub_half; j < (h + 1) * nsub_half; ++j) {
            // straight-line issue: descriptors are built once per sub-tile and advanced by adding to their address
            // field (an issue loop with inner loops over the three products cost 28 us per launch: this thread's
            // instruction stream feeds every MMA of the SM)
            const uint64_t dbh0 = make_smem_desc(band + j * 2 * half_off, 16, 1024);
            const uint64_t dbm0 = make_smem_desc(band + j * 2 * half_off + half_off, 16, 1024);
            const uint32_t d_col = tmem + p.SW * j;
            const uint32_t ah = a_hi + (p.SW >> 1) * j, am = a_mid + (p.SW >> 1) * j;
#pragma unroll
            for (int s = 0; s < 6; ++s) {
              if (s < p.nks) {
                const uint32_t step = static_cast<uint32_t>((s >> 2) * (kb_bytes >> 4) + (s & 3) * 2);
                umma_bf16_ts(d_col, ah + 8 * s, dbh0 + step, idesc, s != 0);
                umma_bf16_ts(d_col, ah + 8 * s, dbm0 + step, idesc, true);
                umma_bf16_ts(d_col, am + 8 * s, dbh0 + step, idesc, true);
              }
            }
          }
          umma_commit(&d_full[h]);
        }
        umma_commit(&a_empty[ab]);
        if (u == ue - 1 || cb == p.ncb - 1) umma_commit(&band_empty[bb]);
      }
    }
  } else {
    // ================================================================= epilogue: accumulator -> staging -> TMA store
    const int q = warp & 3;
    const uint32_t trow = tmem + (static_cast<uint32_t>(32 * q) << 16);
    const CUtensorMap* mapO = side ? &mapGR : &mapGL;
    unsigned char* stage = sOut + (warp - BT_WARP_EPI) * p.NO * BT_OUT_TILE;
    const float inv = 1.0f / static_cast<float>(p.Cg);
    const int sw = lane & 7;
    int it = 0, nst = 0;
    for (int u = ub; u < ue; ++u, ++it) {
      const int tile = u / p.ncb, cb = u - tile * p.ncb;
      const int seg = tile % p.nseg, t2 = tile / p.nseg;
      const int y = t2 % p.H, bg = t2 / p.H;
      const int x0 = seg * 128;
      const bool valid = cb * 128 + 32 * q < p.Cg;
#pragma unroll 1
      for (int h = 0; h < 2; ++h) {
        mbar_wait(&d_full[h], it & 1);
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
            unsigned char* buf = stage + (nst & (p.NO - 1)) * BT_OUT_TILE;
            if (lane == 0) {                                // the store issued NO tiles ago has left this tile
              if (p.NO == 4) asm volatile("cp.async.bulk.wait_group.read 3;" ::: "memory");
              else if (p.NO == 2) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
              else asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            }
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
  if (const char* e = tuning_env("SF_BT_HL")) p.HL = max(p.HL, atoi(e));          // tuning build only (A/B of the plan)
  p.Kg = 128 + p.HL;
  p.P = p.Kg + 4;
  p.SW = 32;        // 16-pixel sub-tiles keep a 48-wide halo within one K block but double the MMA count: measured slower
  if (const char* e = tuning_env("SF_BT_SW")) p.SW = atoi(e);
  p.nks = (p.SW + p.HL) / 16;
  p.nkb = ceil_div(p.SW + p.HL, 64);
  p.nseg = ceil_div(W, 128);
  p.ncb = ceil_div(Cg, 128);
  const long long nt = 1LL * nBG * H * p.nseg;
  if (nt * p.ncb >= (1LL << 30)) return false;
  p.ntiles = static_cast<int>(nt);
  p.stage_bytes = 32 * p.P * 4;
  p.g_bytes = D * p.Kg * 4;
  p.g_pitch = ceil_div(p.g_bytes, 128) * 128;
  p.band_bytes = (128 / p.SW) * 2 * p.nkb * p.SW * 128;
  // double-buffered bands first (they keep the MMAs of one tile off the build of the next); a second gradient-tile slot
  // and a second staging tile per epilogue warp only where they do not cost a feature stage: ring depth is what hides
  // the load latency (measured at the bench shape: 2 / 3 / 4 / 5 stages = 75.6 / 64.6 / 56.5 / 53.1 us)
  int best = -1;
  for (int cfg = 0; cfg < 8; ++cfg) {
    const int NB = (cfg & 4) ? 1 : 2, NG = (cfg & 1) ? 1 : 2, NO = (cfg & 2) ? 1 : 2;
    const int rest = 232448 - 1024 - 512 - 4 * NO * BT_OUT_TILE - NB * p.band_bytes - NG * p.g_pitch;
    const int NS = rest > 0 ? min(BT_MAXNS, rest / p.stage_bytes) : 0;
    const int score = (NS >= 4 ? 1000 : 0) + (NB == 2 ? 100 : 0) + 10 * min(NS, 5) + 2 * NO + NG;   // > 5 stages buy nothing
    if (NS >= 2 && score > best) { best = score; p.NB = NB; p.NG = NG; p.NO = NO; p.NS = NS; }
  }
  if (best < 0) return false;
  if (const char* e = tuning_env("SF_BT_NG")) p.NG = atoi(e);
  if (const char* e = tuning_env("SF_BT_NB")) p.NB = atoi(e);
  if (const char* e = tuning_env("SF_BT_NO")) p.NO = atoi(e);
  if (const char* e = tuning_env("SF_BT_NS")) p.NS = min(p.NS, atoi(e));
  return true;
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
  const size_t smem = static_cast<size_t>(p.NB) * p.band_bytes + 4 * p.NO * BT_OUT_TILE + static_cast<size_t>(p.NG) * p.g_pitch +
                      static_cast<size_t>(p.NS) * p.stage_bytes + 512;
  int grid;
  if (sides == 3) grid = min(2 * p.ntiles * p.ncb, sm_count() & ~1);
  else grid = min(p.ntiles * p.ncb, sm_count());
  SF_CUDA(cudaFuncSetAttribute(corr_volume_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               static_cast<int>(smem)));
  corr_volume_bwd_tc_kernel<<<grid, BT_THREADS, smem, stream>>>(mL, mR, mG, mGL, mGR, p);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

}  // namespace sf

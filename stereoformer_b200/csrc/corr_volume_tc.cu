// Correlation cost volume on tcgen05 tensor cores  (the bench's dominant kernel).
//
// Reference: models/utils/build_cost_volume.py:39-47, models/CrossAttentionCostVolume/GWC_Cost_Volume.py:5-24:
//   cv[b,g,d,y,x] = (1/Cg) * sum_{c in g} L[b,c,y,x] * R[b,c,y,x-d]   (0 for x < d).
//
// Per epipolar row this is the BAND of the product L^T R.  The FFMA kernel (corr_volume.cu) is co-limited by the
// FP32 pipe (AI ~ 5.7 flop/B vs an FP32 ridge of ~8); here the math moves to the tensor cores and the SM only
// converts and extracts:
//   * warp 0 streams 32-channel chunks of the left row segment and of the right row segment (with its left halo;
//     TMA out-of-bounds zero fill = the reference's x<d zeros) into a shared-memory ring (cp.async.bulk.tensor).
//     (Loading the rows straight into registers with ld.global instead was measured too: the LSU cannot keep
//     more than ~32 KB per SM in flight, 3.0 TB/s; the TMA ring alone streams at 6.3 TB/s.)
//   * 8 converter warps split every fp32 value into bf16 hi + bf16 mid (residual <= 2^-17 |x|) and write both
//     halves into the canonical MN-major SWIZZLE_128B UMMA layout (double-buffered operand stage);
//   * one thread issues hi*hi + hi*mid + mid*hi (3 tcgen05.mma per 16 channels) into a [128 x N] fp32 TMEM
//     accumulator (double-buffered, 2 x 256 columns): acc[x][c] = <L[:,xs0+x], R[:,xs0-HL+c]>;
//   * 4 epilogue warps read the accumulator (thread = pixel), scale by 1/Cg and scatter the band into a
//     [D][XS] shared-memory tile (out[d][x] = acc[x][x+HL-d]; the skew makes the stores conflict-free), which one
//     TMA store (cp.async.bulk.tensor, asynchronous, clipped at W) writes to the [B*G*D][H][W] volume.
// Error of the 3-term split: ~2^-16 relative per product (parity bound 1e-4 of the volume magnitude).
// Algorithmic bytes per row tile: (2*Cg + D) * W * 4, all loads and stores exactly once.
#include <cuda_bf16.h>
#include <stdlib.h>

#include "common.cuh"
#include "tc_common.cuh"
#include "../../include/stereoformer_b200.h"

namespace sf {

constexpr int VT_KC = 32;                        // channels per operand stage
constexpr int VT_NCONV = 8;                      // converter warps: warp w converts channel rows k = w mod 8 of every chunk
constexpr int VT_WARP_MMA = 1 + VT_NCONV;
constexpr int VT_THREADS = (1 + VT_NCONV + 1 + 4) * 32;
constexpr int VT_BLK = VT_KC * 128;              // one MN block (64 elements) of an operand half: 4 KB
constexpr int VT_A_HALF = 2 * VT_BLK;            // 128 pixels
constexpr int VT_NOPS = 2;                       // operand stages
constexpr int VT_MAXNS = 8;                      // staging ring depth limit (barrier slots)

struct CorrTcParams {
  int Cg, H, W, D, nBG;
  int XS;          // pixels per x segment (multiple of 8, <= 128)
  int HL;          // left halo of the right row (multiple of 4, >= D-1)
  int BW;          // XS + HL: width of the right-row box (<= 192)
  int N;           // MMA N: BW rounded up to 16
  int nbB;         // MN blocks of the B operand
  int nseg, ntiles, nch;
  int NS;          // staging ring depth
  int stage_bytes, op_bytes, tile_bytes;
  long long* ts;   // tuning only: per-role clock64 timeline of CTA 0 (NULL in production)
};

// fp32x4 -> bf16 hi x4 + bf16 mid x4, stored into the two operand halves
__device__ __forceinline__ void split_store(unsigned char* hi_p, unsigned char* mid_p, const float4 v) {
  const __nv_bfloat162 h01 = __floats2bfloat162_rn(v.x, v.y), h23 = __floats2bfloat162_rn(v.z, v.w);
  const __nv_bfloat162 m01 = __floats2bfloat162_rn(v.x - __low2float(h01), v.y - __high2float(h01));
  const __nv_bfloat162 m23 = __floats2bfloat162_rn(v.z - __low2float(h23), v.w - __high2float(h23));
  *reinterpret_cast<uint2*>(hi_p) =
      make_uint2(*reinterpret_cast<const uint32_t*>(&h01), *reinterpret_cast<const uint32_t*>(&h23));
  *reinterpret_cast<uint2*>(mid_p) =
      make_uint2(*reinterpret_cast<const uint32_t*>(&m01), *reinterpret_cast<const uint32_t*>(&m23));
}

// byte offset of MN element x (x % 4 == 0) in channel row kr (< 8) of an MN-major SW128 operand half
__device__ __forceinline__ int op_offset(int x, int kr) {
  return (x >> 6) * VT_BLK + kr * 128 + ((((x & 63) >> 3) ^ kr) << 4) + (x & 7) * 2;
}

__device__ __forceinline__ void ts_mark(const CorrTcParams& p, int role, int& n) {
  if (p.ts != nullptr && blockIdx.x == 0 && n < 128) p.ts[role * 128 + n++] = clock64();
}

// TMA tiled store, 3-D, shared::cta -> global (bulk async-group completion)
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* map, const void* src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(map)), "r"(sftc::smem_u32(src)), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }


// DBG (tuning only, bit mask): 1 = skip the conversion, 2 = skip the loads, 4 = skip the MMAs, 8 = skip the epilogue.
template <int DBG>
__global__ void __launch_bounds__(VT_THREADS, 1)
corr_volume_tc_kernel(const __grid_constant__ CUtensorMap mapL, const __grid_constant__ CUtensorMap mapR,
                      const __grid_constant__ CUtensorMap mapO, const CorrTcParams p) {
  using namespace sftc;
  extern __shared__ __align__(1024) unsigned char smem[];
  unsigned char* sStage = smem;                                   // [NS][stage_bytes]   fp32 L | R chunks
  unsigned char* sOps = sStage + p.NS * p.stage_bytes;            // [2][op_bytes]       A hi | A mid | B hi | B mid
  float* sTile = reinterpret_cast<float*>(sOps + VT_NOPS * p.op_bytes);   // [D][XS] output tile
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<unsigned char*>(sTile) + p.tile_bytes);
  uint64_t* stg_full = bars;                        // [NS]  TMA -> converters
  uint64_t* stg_empty = bars + VT_MAXNS;            // [NS]  converters (8 warps) -> TMA
  uint64_t* op_full = bars + 2 * VT_MAXNS;          // [2]   converters (8 warps) -> MMA
  uint64_t* op_empty = op_full + 2;                 // [2]   MMA commit -> converters
  uint64_t* acc_full = op_full + 4;                 // [2]   MMA commit -> epilogue
  uint64_t* acc_empty = op_full + 6;                // [2]   epilogue (4 warps) -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(op_full + 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.NS; ++s) { mbar_init(&stg_full[s], 1); mbar_init(&stg_empty[s], VT_NCONV); }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&op_full[s], VT_NCONV); mbar_init(&op_empty[s], 1);
      mbar_init(&acc_full[s], 1); mbar_init(&acc_empty[s], 4);
    }
    fence_mbar_init();
    tma_prefetch_desc(&mapL);
    tma_prefetch_desc(&mapR);
    tma_prefetch_desc(&mapO);
  }
  if (warp == VT_WARP_MMA) tmem_alloc(tmem_slot, 512);
  // operand buffers start as zeros: pixel rows >= XS of A and columns >= BW of B are never written
  {
    uint4* z = reinterpret_cast<uint4*>(sOps);
    const int n16 = VT_NOPS * p.op_bytes / 16;
    for (int i = threadIdx.x; i < n16; i += VT_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
    fence_proxy_async_smem();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // ================================================================= TMA producer
    if (lane == 0 && !(DBG & 2)) {
      const uint32_t bytes = VT_KC * (p.XS + p.BW) * 4;
      int s = 0, ph = 1, nts = 0;
      for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        const int seg = tile % p.nseg, t2 = tile / p.nseg;
        const int y = t2 % p.H, bg = t2 / p.H;
        const int xs0 = seg * p.XS, row0 = bg * p.Cg;
        for (int kc = 0; kc < p.nch; ++kc) {
          mbar_wait(&stg_empty[s], ph);
          ts_mark(p, 0, nts);
          unsigned char* dst = sStage + s * p.stage_bytes;
          mbar_expect_tx(&stg_full[s], bytes);
          tma_load_3d(dst, &mapL, &stg_full[s], xs0, y, row0 + kc * VT_KC);
          tma_load_3d(dst + VT_KC * p.XS * 4, &mapR, &stg_full[s], xs0 - p.HL, y, row0 + kc * VT_KC);
          if (++s == p.NS) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp <= VT_NCONV) {
    // ================================================================= converters: fp32 -> bf16 hi/mid operands
    const int cw = warp - 1;                       // this warp's channel rows of a chunk: cw + 8*i, i = 0..3
    const int nx4L = p.XS >> 2, nx4R = p.BW >> 2;
    const int T = nx4R > 32 ? nx4R - 32 : 0;       // right-row float4 columns >= 32 (<= 16): two rows per pass
    const int tsub = lane >> 4, tx4 = 32 + (lane & 15);
    const int offM = op_offset(4 * lane, cw);
    const int offT = op_offset(4 * tx4, cw) + tsub * 1024;
    const bool stl = lane < nx4L, str = lane < nx4R, stt = (lane & 15) < T;
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    const int rowL = 8 * p.XS, rowR = 8 * p.BW;    // floats between this warp's consecutive rows
    const float* Lp = reinterpret_cast<const float*>(sStage) + cw * p.XS + 4 * lane;
    const float* Rp = reinterpret_cast<const float*>(sStage) + VT_KC * p.XS + cw * p.BW + 4 * lane;
    const float* Tp = reinterpret_cast<const float*>(sStage) + VT_KC * p.XS + (cw + 8 * tsub) * p.BW + 4 * tx4;
    const int stage_f = p.stage_bytes >> 2;
    int s = 0, ph = 0, o = 0, pho = 1, nts = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
      for (int kc = 0; kc < p.nch; ++kc) {
        if (!(DBG & 2)) mbar_wait(&stg_full[s], ph);
        if (threadIdx.x == 32) ts_mark(p, 1, nts);
        mbar_wait(&op_empty[o], pho);
        if (threadIdx.x == 32) ts_mark(p, 1, nts);
        if (!(DBG & 1)) {
          const float* Ls = Lp + s * stage_f;
          const float* Rs = Rp + s * stage_f;
          const float* Ts = Tp + s * stage_f;
          const int nv = p.Cg - kc * VT_KC - cw;   // row i is a real channel iff 8*i < nv
          float4 l[4], r[4], t[2];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            l[i] = (stl && 8 * i < nv) ? *reinterpret_cast<const float4*>(Ls + i * rowL) : zero4;
            r[i] = (str && 8 * i < nv) ? *reinterpret_cast<const float4*>(Rs + i * rowR) : zero4;
          }
          t[0] = (stt && 8 * tsub < nv) ? *reinterpret_cast<const float4*>(Ts) : zero4;
          t[1] = (stt && 8 * (tsub + 2) < nv) ? *reinterpret_cast<const float4*>(Ts + 2 * rowR) : zero4;
          unsigned char* a_hi = sOps + o * p.op_bytes + offM;
          unsigned char* b_hi = a_hi + 2 * VT_A_HALF;
          const int bmid = p.nbB * VT_BLK;
          if (stl) {
#pragma unroll
            for (int i = 0; i < 4; ++i) split_store(a_hi + i * 1024, a_hi + VT_A_HALF + i * 1024, l[i]);
          }
          if (str) {
#pragma unroll
            for (int i = 0; i < 4; ++i) split_store(b_hi + i * 1024, b_hi + bmid + i * 1024, r[i]);
          }
          if (stt) {
            unsigned char* t_hi = b_hi + (offT - offM);
            split_store(t_hi, t_hi + bmid, t[0]);
            split_store(t_hi + 2048, t_hi + bmid + 2048, t[1]);
          }
          fence_proxy_async_smem();
        }
        __syncwarp();
        if (threadIdx.x == 32) ts_mark(p, 1, nts);
        if (lane == 0) {
          mbar_arrive(&op_full[o]);
          mbar_arrive(&stg_empty[s]);
        }
        if (++s == p.NS) { s = 0; ph ^= 1; }
        if (++o == VT_NOPS) { o = 0; pho ^= 1; }
      }
    }
  } else if (warp == VT_WARP_MMA) {
    // ================================================================= MMA issuer
    if (lane == 0) {
      const uint32_t idesc = make_idesc_bf16(128, p.N, 1, 1);
      int o = 0, ph = 0, t = 0, nts = 0;
      for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++t) {
        const int buf = t & 1;
        mbar_wait(&acc_empty[buf], ((t >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t acc = tmem + buf * 256;
        for (int kc = 0; kc < p.nch; ++kc) {
          mbar_wait(&op_full[o], ph);
          tc_fence_after();
          ts_mark(p, 2, nts);
          if (!(DBG & 4)) {
            const uint32_t a_hi = smem_u32(sOps + o * p.op_bytes), a_mid = a_hi + VT_A_HALF;
            const uint32_t b_hi = a_hi + 2 * VT_A_HALF, b_mid = b_hi + p.nbB * VT_BLK;
#pragma unroll
            for (int j = 0; j < VT_KC / 16; ++j) {
              const uint32_t ko = j * 2048;
              const uint64_t dah = make_smem_desc(a_hi + ko, VT_BLK, 1024), dam = make_smem_desc(a_mid + ko, VT_BLK, 1024);
              const uint64_t dbh = make_smem_desc(b_hi + ko, VT_BLK, 1024), dbm = make_smem_desc(b_mid + ko, VT_BLK, 1024);
              umma_bf16(acc, dah, dbh, idesc, (kc | j) != 0);
              umma_bf16(acc, dah, dbm, idesc, true);
              umma_bf16(acc, dam, dbh, idesc, true);
            }
          }
          umma_commit(&op_empty[o]);
          if (++o == VT_NOPS) { o = 0; ph ^= 1; }
        }
        umma_commit(&acc_full[buf]);
      }
    }
  } else if (warp > VT_WARP_MMA) {
    // ================================================================= epilogue: band -> [D][XS] tile -> TMA store
    const int q = warp & 3;                         // TMEM lane quarter this warp may read
    const int xl = 32 * q + lane;
    const bool leader = (warp == VT_WARP_MMA + 1) && lane == 0;
    const float fc = static_cast<float>(p.Cg);
    const bool pow2 = (p.Cg & (p.Cg - 1)) == 0;
    const float inv = 1.0f / fc;                    // exact for a power of two: x * inv == x / Cg bit for bit
    const int XS = p.XS;
    int t = 0, nts = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++t) {
      const int seg = tile % p.nseg, t2 = tile / p.nseg;
      const int y = t2 % p.H, bg = t2 / p.H;
      const int buf = t & 1;
      mbar_wait(&acc_full[buf], (t >> 1) & 1);
      tc_fence_after();
      if (leader) ts_mark(p, 3, nts);
      if (leader && t > 0) bulk_wait_read0();        // the previous tile has left shared memory
      named_bar_sync(1, 128);
      if (32 * q < XS && !(DBG & 8)) {
        const uint32_t trow = tmem + buf * 256 + (static_cast<uint32_t>(32 * q) << 16);
        for (int d0 = 0; d0 < p.D; d0 += 32) {
          // lane l, disparity d0+e needs column 32q + l + HL - d0 - e: a 63-column window starting at cs
          const int cs = max(0, 32 * q + p.HL - d0 - 31);
          const int ne = min(32, p.D - d0);
          const int dj0 = xl + p.HL - cs;           // register j of the window holds disparity dj0 - j
          float* tp = sTile + dj0 * XS + xl;
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            uint32_t raw[32];
            tmem_ld32(trow + cs + 32 * hh, raw);
            tmem_ld_wait();
            if (xl < XS) {
              const unsigned ub = static_cast<unsigned>(ne);
              const int e0 = dj0 - 32 * hh - d0;     // register j holds disparity d0 + (e0 - j)
              float* tq = tp - 32 * hh * XS;
              if (pow2) {
#pragma unroll
                for (int j = 0; j < 32; ++j)
                  if (static_cast<unsigned>(e0 - j) < ub) tq[-j * XS] = __uint_as_float(raw[j]) * inv;
              } else {
#pragma unroll
                for (int j = 0; j < 32; ++j)
                  if (static_cast<unsigned>(e0 - j) < ub) tq[-j * XS] = __fdiv_rn(__uint_as_float(raw[j]), fc);
              }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[buf]);
      fence_proxy_async_smem();
      named_bar_sync(1, 128);
      if (leader) ts_mark(p, 3, nts);
      if (leader && !(DBG & 8)) {
        tma_store_3d(&mapO, sTile, seg * XS, y, bg * p.D);
        bulk_commit();
      }
    }
    if (leader) bulk_wait0();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == VT_WARP_MMA) tmem_free(tmem, 512);
}

static bool corr_tc_plan(CorrTcParams& p, int nBG, int Cg, int H, int W, int D) {
  p.Cg = Cg; p.H = H; p.W = W; p.D = D; p.nBG = nBG;
  p.HL = ((D - 1 + 3) / 4) * 4;
  if (p.HL > 128 || D > 256 || 1LL * nBG * D >= (1LL << 31)) return false;
  p.nch = ceil_div(Cg, VT_KC);
  // widest x segment whose right-row window fits 48 float4 columns and whose staging ring (>= 2 stages), operand
  // stages and output tile fit in shared memory
  for (int xcap = 128; xcap >= 32; xcap -= 32) {
    if (xcap + p.HL > 192) continue;
    p.nseg = ceil_div(W, xcap);
    p.XS = ceil_div(ceil_div(W, p.nseg), 8) * 8;
    if (p.XS > xcap) { p.nseg += 1; p.XS = ceil_div(ceil_div(W, p.nseg), 8) * 8; }
    p.BW = p.XS + p.HL;
    p.N = ceil_div(p.BW, 16) * 16;
    p.nbB = ceil_div(p.N, 64);
    const long long nt = 1LL * nBG * H * p.nseg;
    if (nt >= (1LL << 30)) return false;
    p.ntiles = static_cast<int>(nt);
    p.stage_bytes = ceil_div(VT_KC * (p.XS + p.BW) * 4, 1024) * 1024;
    p.op_bytes = 2 * VT_A_HALF + 2 * p.nbB * VT_BLK;
    p.tile_bytes = ceil_div(D * p.XS * 4, 1024) * 1024;
    p.NS = min(VT_MAXNS, (232448 - VT_NOPS * p.op_bytes - p.tile_bytes - 256) / p.stage_bytes);
    if (p.NS >= 2) return true;
  }
  return false;
}

bool corr_tc_supported(const float* L, const float* R, const float* out, int nBG, int Cg, int H, int W, int D) {
  if (((reinterpret_cast<uintptr_t>(L) | reinterpret_cast<uintptr_t>(R) | reinterpret_cast<uintptr_t>(out)) & 15) != 0)
    return false;
  if (W % 4 != 0) return false;
  CorrTcParams p;
  return corr_tc_plan(p, nBG, Cg, H, W, D);
}

int launch_corr_tc(const float* L, const float* R, float* out, int nBG, int Cg, int H, int W, int D, int dbg,
                   cudaStream_t stream) {
  CorrTcParams p;
  SF_CHECK_ARG(corr_tc_plan(p, nBG, Cg, H, W, D), "corr_volume: shape does not fit the tcgen05 kernel");
  {
    const char* e = getenv("SF_CV_TS");   // tuning hook: device pointer of a [4][128] int64 timeline buffer
    p.ts = e ? reinterpret_cast<long long*>(strtoull(e, nullptr, 16)) : nullptr;
  }
  CUtensorMap mL, mR, mO;
  int rc = make_tmap_3d_f32(&mL, L, W, H, static_cast<uint64_t>(nBG) * Cg, p.XS, 1, VT_KC);
  if (rc) return rc;
  rc = make_tmap_3d_f32(&mR, R, W, H, static_cast<uint64_t>(nBG) * Cg, p.BW, 1, VT_KC);
  if (rc) return rc;
  rc = make_tmap_3d_f32(&mO, out, W, H, static_cast<uint64_t>(nBG) * D, p.XS, 1, D);
  if (rc) return rc;
  const size_t smem = static_cast<size_t>(p.NS) * p.stage_bytes + VT_NOPS * p.op_bytes + p.tile_bytes + 256;
  const int grid = min(p.ntiles, sm_count());
  auto go = [&](auto kern) -> int {
    SF_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    kern<<<grid, VT_THREADS, smem, stream>>>(mL, mR, mO, p);
    SF_LAUNCH_CHECK();
    return SF_OK;
  };
  switch (dbg) {
    case 13: return go(corr_volume_tc_kernel<13>);   // loads + barrier protocol only
    case 2: return go(corr_volume_tc_kernel<2>);     // everything but the loads
    case 14: return go(corr_volume_tc_kernel<14>);   // conversion only
    case 11: return go(corr_volume_tc_kernel<11>);   // MMA only
    case 8: return go(corr_volume_tc_kernel<8>);     // no epilogue
    case 12: return go(corr_volume_tc_kernel<12>);   // loads + conversion
    default: return go(corr_volume_tc_kernel<0>);
  }
}

}  // namespace sf

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
//     The boxes have compile-time widths (128 and 192 floats; columns past the image are zero-filled, not
//     fetched), so every shared-memory offset of the converters is an immediate.
//     (Loading the rows straight into registers with ld.global instead was measured too: the LSU cannot keep
//     more than ~32 KB per SM in flight, 3.0 TB/s; the TMA ring alone streams at 6.3 TB/s.)
//   * converter warps split every fp32 value into bf16 hi + bf16 mid (residual <= 2^-17 |x|).  8 warps write the
//     right-row halves into the canonical MN-major SWIZZLE_128B UMMA layout in shared memory (B operand, double
//     buffered); 4 warps (thread = pixel) write the left-row halves to TENSOR MEMORY (tcgen05.st; A operand):
//     the shared-memory pipe is the limiter of this kernel (TMA write + converter read + operand write + MMA
//     operand reads) and with A in TMEM the MMAs read only B from shared memory;
//   * one thread issues hi*hi + hi*mid + mid*hi (3 tcgen05.mma per 16 channels, A from TMEM) into a [128 x N] fp32
//     TMEM accumulator (double-buffered, 2 x 192 columns): acc[x][c] = <L[:,xs0+x], R[:,xs0-HL+c]>;
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

constexpr int VT_KC = 32;                        // channels per stage
constexpr int VT_PL = 128;                       // floats per staged left row  (box width; XS <= 128)
constexpr int VT_PR = 192;                       // floats per staged right row (box width; XS + HL <= 192)
constexpr int VT_STAGE = VT_KC * (VT_PL + VT_PR) * 4;   // 40 KB
constexpr int VT_NB = 8;                         // B converter warps (warps 1..8): channel rows k = w mod 8 of every chunk
constexpr int VT_NA = 4;                         // A converter warps (warps 9..12): one per TMEM lane quarter
constexpr int VT_WARP_A = 1 + VT_NB;
constexpr int VT_WARP_MMA = VT_WARP_A + VT_NA;
constexpr int VT_THREADS = (VT_WARP_MMA + 1 + 4) * 32;
constexpr int VT_BLK = VT_KC * 128;              // one MN block (64 elements) of an operand half: 4 KB
constexpr int VT_BHALF = 3 * VT_BLK;             // B hi (or mid): 192 columns
constexpr int VT_OP = 2 * VT_BHALF;              // one B operand stage: 24 KB
constexpr int VT_NOPS = 2;                       // operand stages
constexpr int VT_ACC_COLS = 192;                 // TMEM columns per accumulator buffer (N <= 192)
constexpr int VT_TA_BASE = 2 * VT_ACC_COLS;      // TMEM columns of the A stages: [2][hi k0 | hi k1 | mid k0 | mid k1] x 8
constexpr int VT_MAXNS = 8;                      // staging ring depth limit (barrier slots)

struct CorrTcParams {
  int Cg, H, W, D, nBG;
  int XS;          // pixels per x segment (multiple of 8, <= 128)
  int HL;          // left halo of the right row (multiple of 4, >= D-1, <= 64)
  int BW;          // XS + HL: used width of the right-row box (<= 192)
  int N;           // MMA N: BW rounded up to 16
  int nseg, ntiles, nch;
  int NS;          // staging ring depth
  int tile_bytes;
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
  unsigned char* sStage = smem;                                   // [NS][VT_STAGE]   fp32 L [32][128] | R [32][192]
  unsigned char* sOps = sStage + p.NS * VT_STAGE;                 // [2][VT_OP]       B hi | B mid
  float* sTile = reinterpret_cast<float*>(sOps + VT_NOPS * VT_OP);       // [D][XS] output tile
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<unsigned char*>(sTile) + p.tile_bytes);
  uint64_t* stg_full = bars;                        // [NS]  TMA -> converters
  uint64_t* stg_empty = bars + VT_MAXNS;            // [NS]  converters (12 warps) -> TMA
  uint64_t* op_full = bars + 2 * VT_MAXNS;          // [2]   converters (12 warps) -> MMA
  uint64_t* op_empty = op_full + 2;                 // [2]   MMA commit -> converters
  uint64_t* acc_full = op_full + 4;                 // [2]   MMA commit -> epilogue
  uint64_t* acc_empty = op_full + 6;                // [2]   epilogue (4 warps) -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(op_full + 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // ---- TMA producer state (thread 0): starts streaming before the rest of the prologue has finished
  int pr_tile = blockIdx.x, pr_kc = 0, pr_s = 0, pr_ph = 1, pr_nts = 0;
  auto produce = [&]() {                         // issue the loads of the next chunk; false when none is left
    if (pr_tile >= p.ntiles) return false;
    const int seg = pr_tile % p.nseg, t2 = pr_tile / p.nseg;
    const int y = t2 % p.H, bg = t2 / p.H;
    const int xs0 = seg * p.XS, c0 = bg * p.Cg + pr_kc * VT_KC;
    mbar_wait(&stg_empty[pr_s], pr_ph);
    ts_mark(p, 0, pr_nts);
    unsigned char* dst = sStage + pr_s * VT_STAGE;
    mbar_expect_tx(&stg_full[pr_s], VT_STAGE);
    tma_load_3d(dst, &mapL, &stg_full[pr_s], xs0, y, c0);
    tma_load_3d(dst + VT_KC * VT_PL * 4, &mapR, &stg_full[pr_s], xs0 - p.HL, y, c0);
    if (++pr_s == p.NS) { pr_s = 0; pr_ph ^= 1; }
    if (++pr_kc == p.nch) { pr_kc = 0; pr_tile += gridDim.x; }
    return true;
  };

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.NS; ++s) { mbar_init(&stg_full[s], 1); mbar_init(&stg_empty[s], VT_NB + VT_NA); }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&op_full[s], VT_NB + VT_NA); mbar_init(&op_empty[s], 1);
      mbar_init(&acc_full[s], 1); mbar_init(&acc_empty[s], 4);
    }
    fence_mbar_init();
    if (!(DBG & 2))
      for (int i = 0; i < p.NS; ++i) produce();   // the ring is empty: these never wait
    tma_prefetch_desc(&mapO);
  }
  if (warp == VT_WARP_MMA) tmem_alloc(tmem_slot, 512);
  // B operand buffers start as zeros: columns >= BW are never written
  {
    uint4* z = reinterpret_cast<uint4*>(sOps);
    for (int i = threadIdx.x; i < VT_NOPS * VT_OP / 16; i += VT_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
    fence_proxy_async_smem();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // ================================================================= TMA producer (rest of the stream)
    if (lane == 0 && !(DBG & 2)) {
      while (produce()) {
      }
    }
  } else if (warp < VT_WARP_A) {
    // ================================================================= B converters: right row -> shared memory
    // this warp's channel rows of a chunk are cw + 8*i, i = 0..3; lane = float4 column (+ 32..47 in the tail pass,
    // where the two half-warps take rows i and i+2)
    const int cw = warp - 1;
    const int nx4R = p.BW >> 2;
    const int tsub = lane >> 4, tx4 = 32 + (lane & 15);
    const bool str = lane < nx4R, stt = tx4 < nx4R;
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    const float* Rp = reinterpret_cast<const float*>(sStage + VT_KC * VT_PL * 4) + cw * VT_PR + 4 * lane;
    const float* Tp = reinterpret_cast<const float*>(sStage + VT_KC * VT_PL * 4) + (cw + 8 * tsub) * VT_PR + 4 * tx4;
    unsigned char* Bm = sOps + op_offset(4 * lane, cw);
    unsigned char* Bt = sOps + op_offset(4 * tx4, cw) + tsub * 1024;
    int s = 0, ph = 0, o = 0, pho = 1, nts = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
      for (int kc = 0; kc < p.nch; ++kc) {
        if (!(DBG & 2)) mbar_wait(&stg_full[s], ph);
        if (threadIdx.x == 32) ts_mark(p, 1, nts);
        if (DBG & 1) mbar_wait(&op_empty[o], pho);
        if (!(DBG & 1)) {
          const float* Rs = Rp + s * (VT_STAGE / 4);
          const float* Ts = Tp + s * (VT_STAGE / 4);
          float4 r[4], t[2];
#pragma unroll
          for (int i = 0; i < 4; ++i) r[i] = *reinterpret_cast<const float4*>(Rs + i * 8 * VT_PR);
          t[0] = *reinterpret_cast<const float4*>(Ts);
          t[1] = *reinterpret_cast<const float4*>(Ts + 16 * VT_PR);
          const int nv = p.Cg - kc * VT_KC - cw;   // row i is a real channel iff 8*i < nv (ragged last chunk only)
          if (nv <= 24) {
#pragma unroll
            for (int i = 0; i < 4; ++i)
              if (8 * i >= nv) r[i] = zero4;
            if (8 * tsub >= nv) t[0] = zero4;
            if (8 * (tsub + 2) >= nv) t[1] = zero4;
          }
          // the staged fp32 rows are in registers: hand the ring slot back to the TMA producer NOW, before the
          // conversion (a slot used to be held ~1200 cycles per chunk on top of the ~3800-cycle load latency; the slot
          // turnaround, not the converters, set the chunk rate -- tools/timeline_cv.py)
#pragma unroll
          for (int i = 0; i < 4; ++i) asm volatile("" ::"f"(r[i].x), "f"(r[i].y), "f"(r[i].z), "f"(r[i].w));
          asm volatile("" ::"f"(t[0].x), "f"(t[0].y), "f"(t[0].z), "f"(t[0].w), "f"(t[1].x), "f"(t[1].y), "f"(t[1].z), "f"(t[1].w));
          __syncwarp();
          if (lane == 0) mbar_arrive(&stg_empty[s]);
          mbar_wait(&op_empty[o], pho);                          // only now is the operand buffer needed
          if (threadIdx.x == 32) ts_mark(p, 1, nts);
          unsigned char* bm = Bm + o * VT_OP;
          if (str) {
#pragma unroll
            for (int i = 0; i < 4; ++i) split_store(bm + i * 1024, bm + VT_BHALF + i * 1024, r[i]);
          }
          if (stt) {
            unsigned char* bt = Bt + o * VT_OP;
            split_store(bt, bt + VT_BHALF, t[0]);
            split_store(bt + 2048, bt + VT_BHALF + 2048, t[1]);
          }
          fence_proxy_async_smem();
        }
        __syncwarp();
        if (threadIdx.x == 32) ts_mark(p, 1, nts);
        if (lane == 0) {
          mbar_arrive(&op_full[o]);
          if (DBG & 1) mbar_arrive(&stg_empty[s]);
        }
        if (++s == p.NS) { s = 0; ph ^= 1; }
        if (++o == VT_NOPS) { o = 0; pho ^= 1; }
      }
    }
  } else if (warp < VT_WARP_MMA) {
    // ================================================================= A converters: left row -> tensor memory
    // thread = pixel 32*qa + lane (the TMEM lane quarter this warp may write), all 32 channels of the chunk;
    // column c of a k-step holds channels (2c, 2c+1).  Pixel columns >= XS of the box are zero-filled by TMA.
    const int qa = warp & 3;
    const float* Ap = reinterpret_cast<const float*>(sStage) + 32 * qa + lane;
    const uint32_t ta = tmem + VT_TA_BASE + (static_cast<uint32_t>(32 * qa) << 16);
    int s = 0, ph = 0, o = 0, pho = 1;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
      for (int kc = 0; kc < p.nch; ++kc) {
        if (!(DBG & 2)) mbar_wait(&stg_full[s], ph);
        if (DBG & 1) { mbar_wait(&op_empty[o], pho); tc_fence_after(); }
        if (!(DBG & 1)) {
          const float* As = Ap + s * (VT_STAGE / 4);
          const int nvc = p.Cg - kc * VT_KC;       // valid channels of this chunk (ragged last chunk only)
          uint32_t ahi[16], amid[16];
          float av[VT_KC];
#pragma unroll
          for (int c = 0; c < VT_KC; ++c) av[c] = As[c * VT_PL];
#pragma unroll
          for (int c = 0; c < VT_KC; c += 4) asm volatile("" ::"f"(av[c]), "f"(av[c + 1]), "f"(av[c + 2]), "f"(av[c + 3]));
          __syncwarp();
          if (lane == 0) mbar_arrive(&stg_empty[s]);            // slot released before the conversion (see the B converters)
          mbar_wait(&op_empty[o], pho);                          // the A stage in tensor memory is free again
          tc_fence_after();
#pragma unroll
          for (int c = 0; c < 16; ++c) {
            float a0 = av[2 * c], a1 = av[2 * c + 1];
            if (nvc < VT_KC) {
              if (2 * c >= nvc) a0 = 0.f;
              if (2 * c + 1 >= nvc) a1 = 0.f;
            }
            const __nv_bfloat162 h = __floats2bfloat162_rn(a0, a1);
            const __nv_bfloat162 m = __floats2bfloat162_rn(a0 - __low2float(h), a1 - __high2float(h));
            ahi[c] = *reinterpret_cast<const uint32_t*>(&h);
            amid[c] = *reinterpret_cast<const uint32_t*>(&m);
          }
          tmem_st16(ta + o * 32, ahi);
          tmem_st16(ta + o * 32 + 16, amid);
          tmem_st_wait();
          tc_fence_before();
        }
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&op_full[o]);
          if (DBG & 1) mbar_arrive(&stg_empty[s]);
        }
        if (++s == p.NS) { s = 0; ph ^= 1; }
        if (++o == VT_NOPS) { o = 0; pho ^= 1; }
      }
    }
  } else if (warp == VT_WARP_MMA) {
    // ================================================================= MMA issuer
    if (lane == 0) {
      const uint32_t idesc = make_idesc_bf16(128, p.N, 0, 1);   // A: TMEM (K-major), B: MN-major shared memory
      int o = 0, ph = 0, t = 0, nts = 0;
      for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++t) {
        const int buf = t & 1;
        mbar_wait(&acc_empty[buf], ((t >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t acc = tmem + buf * VT_ACC_COLS;
        for (int kc = 0; kc < p.nch; ++kc) {
          mbar_wait(&op_full[o], ph);
          tc_fence_after();
          ts_mark(p, 2, nts);
          if (!(DBG & 4)) {
            const uint32_t b_hi = smem_u32(sOps + o * VT_OP), b_mid = b_hi + VT_BHALF;
            const uint32_t a_hi = tmem + VT_TA_BASE + o * 32, a_mid = a_hi + 16;
#pragma unroll
            for (int j = 0; j < VT_KC / 16; ++j) {
              const uint64_t dbh = make_smem_desc(b_hi + j * 2048, VT_BLK, 1024);
              const uint64_t dbm = make_smem_desc(b_mid + j * 2048, VT_BLK, 1024);
              umma_bf16_ts(acc, a_hi + 8 * j, dbh, idesc, (kc | j) != 0);
              umma_bf16_ts(acc, a_hi + 8 * j, dbm, idesc, true);
              umma_bf16_ts(acc, a_mid + 8 * j, dbh, idesc, true);
            }
          }
          umma_commit(&op_empty[o]);
          if (++o == VT_NOPS) { o = 0; ph ^= 1; }
        }
        umma_commit(&acc_full[buf]);
      }
    }
  } else {
    // ================================================================= epilogue: band -> [D][XS] tile -> TMA store
    const int q = warp & 3;                         // TMEM lane quarter this warp may read
    const int xl = 32 * q + lane;
    const bool leader = (warp == VT_WARP_MMA + 1) && lane == 0;
    // mean over channels as a multiply by 1/Cg (what PyTorch's CUDA mean reduction does; bit-identical to the division
    // for power-of-two Cg, within 1 ulp otherwise -- far below the 2^-16 of the bf16 split; an IEEE division here cost
    // the single-warp-per-scheduler epilogue 15-20 us per launch at Cg = 192..384)
    const float inv = 1.0f / static_cast<float>(p.Cg);
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
        const uint32_t trow = tmem + buf * VT_ACC_COLS + (static_cast<uint32_t>(32 * q) << 16);
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
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (static_cast<unsigned>(e0 - j) < ub) tq[-j * XS] = __uint_as_float(raw[j]) * inv;
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
  if (p.HL > VT_PR - VT_PL || 1LL * nBG * D >= (1LL << 31)) return false;
  p.nch = ceil_div(Cg, VT_KC);
  // one segment when the row fits the 128-wide box, else full 128-pixel segments (the last one is clipped)
  p.nseg = ceil_div(W, VT_PL);
  p.XS = p.nseg == 1 ? ceil_div(W, 8) * 8 : VT_PL;
  p.BW = p.XS + p.HL;
  p.N = ceil_div(p.BW, 16) * 16;
  const long long nt = 1LL * nBG * H * p.nseg;
  if (nt >= (1LL << 30)) return false;
  p.ntiles = static_cast<int>(nt);
  p.tile_bytes = ceil_div(D * p.XS * 4, 1024) * 1024;
  p.NS = min(VT_MAXNS, (232448 - VT_NOPS * VT_OP - p.tile_bytes - 256) / VT_STAGE);
  return p.NS >= 2;
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
    const char* e = tuning_env("SF_CV_TS");   // tuning build only: device pointer of a [4][128] int64 timeline buffer
    p.ts = e ? reinterpret_cast<long long*>(strtoull(e, nullptr, 16)) : nullptr;
  }
  CUtensorMap mL, mR, mO;
  int rc = make_tmap_3d_f32(&mL, L, W, H, static_cast<uint64_t>(nBG) * Cg, VT_PL, 1, VT_KC);
  if (rc) return rc;
  rc = make_tmap_3d_f32(&mR, R, W, H, static_cast<uint64_t>(nBG) * Cg, VT_PR, 1, VT_KC);
  if (rc) return rc;
  rc = make_tmap_3d_f32(&mO, out, W, H, static_cast<uint64_t>(nBG) * D, p.XS, 1, D);
  if (rc) return rc;
  const size_t smem = static_cast<size_t>(p.NS) * VT_STAGE + VT_NOPS * VT_OP + p.tile_bytes + 256;
  const int grid = min(p.ntiles, sm_count());
  auto go = [&](auto kern) -> int {
    SF_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    kern<<<grid, VT_THREADS, smem, stream>>>(mL, mR, mO, p);
    SF_LAUNCH_CHECK();
    return SF_OK;
  };
#ifdef SF_TUNING
  switch (dbg) {                                     // component-isolation variants (tools/ only)
    case 13: return go(corr_volume_tc_kernel<13>);   // loads + barrier protocol only
    case 2: return go(corr_volume_tc_kernel<2>);     // everything but the loads
    case 14: return go(corr_volume_tc_kernel<14>);   // conversion only
    case 11: return go(corr_volume_tc_kernel<11>);   // MMA only
    case 8: return go(corr_volume_tc_kernel<8>);     // no epilogue
    case 12: return go(corr_volume_tc_kernel<12>);   // loads + conversion
    default: break;
  }
#endif
  (void)dbg;
  return go(corr_volume_tc_kernel<0>);
}

}  // namespace sf

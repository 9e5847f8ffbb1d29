// CorrBlock1D all-pairs correlation pyramid on tcgen05 tensor cores  (SURVEY 8a row a8).
//
// Reference: models/IterativeLocalCostVolume/corr/corr.py:55-76,103-111 (einsum 'aijk,aijh->ajkh' / sqrt(C),
// then avg_pool2d([1,2]) num_levels times).  Per epipolar row this is a [W1 x C] x [C x W2] GEMM in fp32.
//
// Same machinery as the cost-volume kernel (corr_volume_tc.cu), sized for a write-dominated problem:
//   * warp 0 streams 16-channel chunks of the f1 row tile [16][128] and of the f2 row [16][<=256] into a
//     shared-memory ring with TMA (cp.async.bulk.tensor.3d; columns past the image are zero-filled);
//   * 8 converter warps split the f2 chunk into bf16 hi + bf16 mid and write the canonical MN-major SWIZZLE_128B
//     UMMA layout (B operand, double buffered); 4 warps (thread = x1 pixel) split the f1 chunk and store it to
//     TENSOR MEMORY (A operand, tcgen05.st);
//   * one thread issues hi*hi + hi*mid + mid*hi per chunk into a [128 x N] fp32 TMEM accumulator
//     (double-buffered when 2N + 32 <= 512 columns, e.g. W2 = 240);
//   * 4 epilogue warps read the accumulator 32 columns at a time (thread = x1 row), divide by sqrt(C), compute all
//     pooled levels from the same registers and stage every level as a swizzled [128 rows][32 >> l] tile that one
//     TMA store per level writes to the [B*H][W1][W2 >> l] tensors (levels whose row pitch is not a multiple of
//     16 bytes are written directly).  The volume is written once and never re-read.
// fp32 parity: 3-term bf16 split, within ~1e-5 of the fp32 einsum.
// Algorithmic bytes per pair: (2*C*H*W + H*W*W*(1 + 1/2 + 1/4 + 1/8 (+1/16))) * 4.
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_common.cuh"
#include "../../include/stereoformer_b200.h"

namespace sf {

constexpr int CT_KC = 16;                      // channels per chunk (one k-step)
constexpr int CT_PA = 128;                     // floats per staged f1 row (box width)
constexpr int CT_NB = 8;                       // B converter warps (warps 1..8)
constexpr int CT_NA = 4;                       // A converter warps (warps 9..12)
constexpr int CT_WARP_A = 1 + CT_NB;
constexpr int CT_WARP_MMA = CT_WARP_A + CT_NA;
constexpr int CT_THREADS = (CT_WARP_MMA + 1 + 4) * 32;
constexpr int CT_BLK = CT_KC * 128;            // one MN block (64 elements) of an operand half: 2 KB
constexpr int CT_NOPS = 2;
constexpr int CT_TA_BASE = 480;                // TMEM columns of the two A stages: [2][hi 8 | mid 8]
constexpr int CT_MAXNS = 8;
constexpr int CT_MAXLEV = 5;
constexpr int CT_EPI_BYTES = 128 * (32 + 16 + 8 + 4) * 4;   // one staging set: levels 0..3, 30 KB

struct Corr1dTcParams {
  float* lvl[CT_MAXLEV];
  int Wl[CT_MAXLEV];
  int tma_lvl[CT_MAXLEV];   // level l is stored by TMA (row pitch multiple of 16 B, box >= 16 B)
  int nlev;
  int B, C, H, W1, W2;
  int PB;          // floats per staged f2 row (box width, multiple of 64, <= 256)
  int N;           // MMA N: min(W2 - n0, 256) rounded up to 16 (per n tile; N of the first tile here)
  int nbB;         // MN blocks of the B operand (PB / 64)
  int nmt, nnt, nch;
  int ntiles;
  int NS, stage_bytes, op_bytes;
  int nacc, acc_stride;
  float sqrt_c, inv_sqrt_c;
  int pow2;
};

struct Corr1dMaps {
  CUtensorMap f1, f2, lvl[4];
};

// fp32x4 -> bf16 hi x4 + bf16 mid x4, stored into the two operand halves
__device__ __forceinline__ void ct_split_store(unsigned char* hi_p, unsigned char* mid_p, const float4 v) {
  const __nv_bfloat162 h01 = __floats2bfloat162_rn(v.x, v.y), h23 = __floats2bfloat162_rn(v.z, v.w);
  const __nv_bfloat162 m01 = __floats2bfloat162_rn(v.x - __low2float(h01), v.y - __high2float(h01));
  const __nv_bfloat162 m23 = __floats2bfloat162_rn(v.z - __low2float(h23), v.w - __high2float(h23));
  *reinterpret_cast<uint2*>(hi_p) =
      make_uint2(*reinterpret_cast<const uint32_t*>(&h01), *reinterpret_cast<const uint32_t*>(&h23));
  *reinterpret_cast<uint2*>(mid_p) =
      make_uint2(*reinterpret_cast<const uint32_t*>(&m01), *reinterpret_cast<const uint32_t*>(&m23));
}
// byte offset of MN element x (x % 4 == 0) in channel row kr (< 8) of an MN-major SW128 operand half
__device__ __forceinline__ int ct_op_offset(int x, int kr) {
  return (x >> 6) * CT_BLK + kr * 128 + ((((x & 63) >> 3) ^ kr) << 4) + (x & 7) * 2;
}
__device__ __forceinline__ void ct_tma_store_3d(const CUtensorMap* map, const void* src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(map)), "r"(sftc::smem_u32(src)), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}

__global__ void __launch_bounds__(CT_THREADS, 1)
corr1d_pyramid_tc_kernel(const __grid_constant__ Corr1dMaps maps, const Corr1dTcParams p) {
  using namespace sftc;
  extern __shared__ __align__(1024) unsigned char smem[];
  unsigned char* sStage = smem;                                   // [NS][stage_bytes]  fp32 f1 [16][128] | f2 [16][PB]
  unsigned char* sOps = sStage + p.NS * p.stage_bytes;            // [2][op_bytes]      B hi | B mid
  unsigned char* sEpi = sOps + CT_NOPS * p.op_bytes;              // [2][CT_EPI_BYTES]  level tiles (swizzled)
  uint64_t* bars = reinterpret_cast<uint64_t*>(sEpi + 2 * CT_EPI_BYTES);
  uint64_t* stg_full = bars;                        // [NS]  TMA -> converters
  uint64_t* stg_empty = bars + CT_MAXNS;            // [NS]  converters (12 warps) -> TMA
  uint64_t* op_full = bars + 2 * CT_MAXNS;          // [2]   converters (12 warps) -> MMA
  uint64_t* op_empty = op_full + 2;                 // [2]   MMA commit -> converters
  uint64_t* acc_full = op_full + 4;                 // [2]   MMA commit -> epilogue
  uint64_t* acc_empty = op_full + 6;                // [2]   epilogue (4 warps) -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(op_full + 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // ---- TMA producer state (thread 0): starts streaming before the rest of the prologue has finished
  const uint32_t stage_tx = CT_KC * (CT_PA + p.PB) * 4;
  int pr_tile = blockIdx.x, pr_kc = 0, pr_s = 0, pr_ph = 1;
  auto produce = [&]() {
    if (pr_tile >= p.ntiles) return false;
    const int nt = pr_tile % p.nnt;
    int r = pr_tile / p.nnt;
    const int mt = r % p.nmt; r /= p.nmt;              // r = b*H + y
    const int y = r % p.H, b = r / p.H;
    const int c0 = b * p.C + pr_kc * CT_KC;
    mbar_wait(&stg_empty[pr_s], pr_ph);
    unsigned char* dst = sStage + pr_s * p.stage_bytes;
    mbar_expect_tx(&stg_full[pr_s], stage_tx);
    tma_load_3d(dst, &maps.f1, &stg_full[pr_s], mt * 128, y, c0);
    tma_load_3d(dst + CT_KC * CT_PA * 4, &maps.f2, &stg_full[pr_s], nt * 256, y, c0);
    if (++pr_s == p.NS) { pr_s = 0; pr_ph ^= 1; }
    if (++pr_kc == p.nch) { pr_kc = 0; pr_tile += gridDim.x; }
    return true;
  };

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.NS; ++s) { mbar_init(&stg_full[s], 1); mbar_init(&stg_empty[s], CT_NB + CT_NA); }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&op_full[s], CT_NB + CT_NA); mbar_init(&op_empty[s], 1);
      mbar_init(&acc_full[s], 1); mbar_init(&acc_empty[s], 4);
    }
    fence_mbar_init();
    for (int i = 0; i < p.NS; ++i) produce();   // the ring is empty: these never wait
    for (int l = 0; l < 4; ++l) tma_prefetch_desc(&maps.lvl[l]);
  }
  if (warp == CT_WARP_MMA) tmem_alloc(tmem_slot, 512);
  {  // B operand buffers start as zeros: columns past the staged row are never written
    uint4* z = reinterpret_cast<uint4*>(sOps);
    for (int i = threadIdx.x; i < CT_NOPS * p.op_bytes / 16; i += CT_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
    fence_proxy_async_smem();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // ================================================================= TMA producer (rest of the stream)
    if (lane == 0) {
      while (produce()) {
      }
    }
  } else if (warp < CT_WARP_A) {
    // ================================================================= B converters: f2 row -> shared memory
    const int cw = warp - 1;                         // channel rows cw and cw + 8 of the chunk
    const int nx4 = p.PB >> 2;
    const bool s0 = lane < nx4, s1 = lane + 32 < nx4;
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    const int rowB = 8 * p.PB;
    const float* Rp = reinterpret_cast<const float*>(sStage + CT_KC * CT_PA * 4) + cw * p.PB + 4 * lane;
    unsigned char* B0 = sOps + ct_op_offset(4 * lane, cw);
    unsigned char* B1 = sOps + ct_op_offset(4 * (lane + 32), cw);
    const int bmid = p.nbB * CT_BLK;
    const int stage_f = p.stage_bytes >> 2;
    int s = 0, ph = 0, o = 0, pho = 1;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
      for (int kc = 0; kc < p.nch; ++kc) {
        mbar_wait(&stg_full[s], ph);
        mbar_wait(&op_empty[o], pho);
        const float* Rs = Rp + s * stage_f;
        float4 r00 = zero4, r01 = zero4, r10 = zero4, r11 = zero4;
        if (s0) { r00 = *reinterpret_cast<const float4*>(Rs); r10 = *reinterpret_cast<const float4*>(Rs + rowB); }
        if (s1) { r01 = *reinterpret_cast<const float4*>(Rs + 128); r11 = *reinterpret_cast<const float4*>(Rs + rowB + 128); }
        const int nv = p.C - kc * CT_KC - cw;          // row i is a real channel iff 8*i < nv (ragged last chunk only)
        if (nv <= 8) {
          r10 = r11 = zero4;
          if (nv <= 0) r00 = r01 = zero4;
        }
        unsigned char* b0 = B0 + o * p.op_bytes;
        unsigned char* b1 = B1 + o * p.op_bytes;
        if (s0) { ct_split_store(b0, b0 + bmid, r00); ct_split_store(b0 + 1024, b0 + bmid + 1024, r10); }
        if (s1) { ct_split_store(b1, b1 + bmid, r01); ct_split_store(b1 + 1024, b1 + bmid + 1024, r11); }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&op_full[o]);
          mbar_arrive(&stg_empty[s]);
        }
        if (++s == p.NS) { s = 0; ph ^= 1; }
        if (++o == CT_NOPS) { o = 0; pho ^= 1; }
      }
    }
  } else if (warp < CT_WARP_MMA) {
    // ================================================================= A converters: f1 tile -> tensor memory
    const int qa = warp & 3;
    const float* Ap = reinterpret_cast<const float*>(sStage) + 32 * qa + lane;
    const uint32_t ta = tmem + CT_TA_BASE + (static_cast<uint32_t>(32 * qa) << 16);
    const int stage_f = p.stage_bytes >> 2;
    int s = 0, ph = 0, o = 0, pho = 1;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
      for (int kc = 0; kc < p.nch; ++kc) {
        mbar_wait(&stg_full[s], ph);
        mbar_wait(&op_empty[o], pho);
        tc_fence_after();
        const float* As = Ap + s * stage_f;
        const int nvc = p.C - kc * CT_KC;
        uint32_t ahi[8], amid[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          float a0 = As[(2 * c) * CT_PA], a1 = As[(2 * c + 1) * CT_PA];
          if (nvc < CT_KC) {
            if (2 * c >= nvc) a0 = 0.f;
            if (2 * c + 1 >= nvc) a1 = 0.f;
          }
          const __nv_bfloat162 h = __floats2bfloat162_rn(a0, a1);
          const __nv_bfloat162 m = __floats2bfloat162_rn(a0 - __low2float(h), a1 - __high2float(h));
          ahi[c] = *reinterpret_cast<const uint32_t*>(&h);
          amid[c] = *reinterpret_cast<const uint32_t*>(&m);
        }
        tmem_st8(ta + o * 16, ahi);
        tmem_st8(ta + o * 16 + 8, amid);
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&op_full[o]);
          mbar_arrive(&stg_empty[s]);
        }
        if (++s == p.NS) { s = 0; ph ^= 1; }
        if (++o == CT_NOPS) { o = 0; pho ^= 1; }
      }
    }
  } else if (warp == CT_WARP_MMA) {
    // ================================================================= MMA issuer
    if (lane == 0) {
      int o = 0, ph = 0, t = 0;
      for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++t) {
        const int nt = tile % p.nnt;
        const int N = min(256, ((p.W2 - nt * 256 + 15) / 16) * 16);
        const uint32_t idesc = make_idesc_bf16(128, N, 0, 1);   // A: TMEM (K-major), B: MN-major shared memory
        const int buf = p.nacc == 2 ? (t & 1) : 0;
        const int use = p.nacc == 2 ? (t >> 1) : t;              // how often this buffer has been used before
        mbar_wait(&acc_empty[buf], (use & 1) ^ 1);
        tc_fence_after();
        const uint32_t acc = tmem + buf * p.acc_stride;
        for (int kc = 0; kc < p.nch; ++kc) {
          mbar_wait(&op_full[o], ph);
          tc_fence_after();
          const uint32_t b_hi = smem_u32(sOps + o * p.op_bytes), b_mid = b_hi + p.nbB * CT_BLK;
          const uint32_t a_hi = tmem + CT_TA_BASE + o * 16, a_mid = a_hi + 8;
          const uint64_t dbh = make_smem_desc(b_hi, CT_BLK, 1024), dbm = make_smem_desc(b_mid, CT_BLK, 1024);
          umma_bf16_ts(acc, a_hi, dbh, idesc, kc != 0);
          umma_bf16_ts(acc, a_hi, dbm, idesc, true);
          umma_bf16_ts(acc, a_mid, dbh, idesc, true);
          umma_commit(&op_empty[o]);
          if (++o == CT_NOPS) { o = 0; ph ^= 1; }
        }
        umma_commit(&acc_full[buf]);
      }
    }
  } else {
    // ================================================================= epilogue: scale, level 0 + pooled levels
    const int q = warp & 3;
    const int row = 32 * q + lane;                      // x1 row of the tile = TMEM lane
    const bool leader = (warp == CT_WARP_MMA + 1) && lane == 0;
    int t = 0, blk = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++t) {
      const int nt = tile % p.nnt;
      int r = tile / p.nnt;
      const int mt = r % p.nmt; r /= p.nmt;             // r = b*H + y
      const int m0 = mt * 128, n0 = nt * 256;
      const int x1 = m0 + row;
      const bool rowok = x1 < p.W1;
      const size_t pix = static_cast<size_t>(r) * p.W1 + x1;
      const int buf = p.nacc == 2 ? (t & 1) : 0;
      const int use = p.nacc == 2 ? (t >> 1) : t;
      mbar_wait(&acc_full[buf], use & 1);
      tc_fence_after();
      const uint32_t trow = tmem + buf * p.acc_stride + (static_cast<uint32_t>(32 * q) << 16);
      const int ncb = (min(256, p.W2 - n0) + 31) / 32;
      for (int cb = 0; cb < ncb; ++cb, ++blk) {
        uint32_t raw[32];
        tmem_ld32(trow + cb * 32, raw);
        tmem_ld_wait();
        float v[32];
        // / sqrt(C) as a multiply by the reciprocal: bit-identical when sqrt(C) is a power of two (C = 256), within
        // 1 ulp otherwise (far below the 2^-16 of the bf16 split); an IEEE division here is 240 slow-path divides per
        // thread per tile on single-warp-per-scheduler epilogue warps
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(raw[i]) * p.inv_sqrt_c;
        // staging set of this block; its previous TMA stores (two blocks ago) must have finished reading it
        unsigned char* S = sEpi + (blk & 1) * CT_EPI_BYTES;
        if (leader) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        named_bar_sync(1, 128);
        const int c0 = n0 + cb * 32;
        // level 0: [128][32] floats, 128-byte rows, 16-byte chunks XOR-swizzled with (row & 7)
        if (p.tma_lvl[0]) {
          unsigned char* s0 = S + row * 128;
#pragma unroll
          for (int c = 0; c < 8; ++c)
            *reinterpret_cast<float4*>(s0 + ((c ^ (row & 7)) << 4)) = make_float4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
        } else if (rowok) {
          float* o = p.lvl[0] + pix * p.Wl[0] + c0;
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (c0 + i < p.Wl[0]) o[i] = v[i];
        }
        // pooled levels: pairwise means in registers (avg_pool2d([1,2]), floor mode: odd tails dropped)
        int cnt = 32, cl = c0;
        unsigned char* sl = S + 128 * 128;
#pragma unroll
        for (int l = 1; l < CT_MAXLEV; ++l) {
          cnt >>= 1; cl >>= 1;
#pragma unroll
          for (int i = 0; i < 16; ++i)
            if (i < cnt) v[i] = __fmul_rn(__fadd_rn(v[2 * i], v[2 * i + 1]), 0.5f);
          if (l < p.nlev) {
            if (l < 4 && p.tma_lvl[l]) {
              // [128][cnt] floats: rows of cnt*4 bytes; chunks swizzled like the matching TMA mode (64B / 32B / none)
              unsigned char* srow = sl + row * (cnt * 4);
              const int sw = l == 1 ? ((row >> 1) & 3) : (l == 2 ? ((row >> 2) & 1) : 0);
#pragma unroll
              for (int c = 0; c < 4; ++c)
                if (4 * c < cnt)
                  *reinterpret_cast<float4*>(srow + ((c ^ sw) << 4)) = make_float4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
            } else if (rowok) {
              float* o = p.lvl[l] + pix * p.Wl[l] + cl;
#pragma unroll
              for (int i = 0; i < 16; ++i)
                if (i < cnt && cl + i < p.Wl[l]) o[i] = v[i];
            }
          }
          if (l < 4) sl += 128 * cnt * 4;
        }
        fence_proxy_async_smem();
        named_bar_sync(1, 128);
        if (leader) {
          unsigned char* sp = S;
          int w = 32;
#pragma unroll
          for (int l = 0; l < 4; ++l) {
            if (l < p.nlev && p.tma_lvl[l]) ct_tma_store_3d(&maps.lvl[l], sp, c0 >> l, m0, r);
            sp += 128 * w * 4;
            w >>= 1;
          }
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[buf]);
    }
    if (leader) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  if (warp == CT_WARP_MMA) tmem_free(tmem, 512);
}

bool corr1d_tc_supported(const float* f1, const float* f2, float* const* levels, int n_out, int W1, int W2) {
  uintptr_t a = reinterpret_cast<uintptr_t>(f1) | reinterpret_cast<uintptr_t>(f2);
  for (int l = 0; l < n_out; ++l) a |= reinterpret_cast<uintptr_t>(levels[l]);
  return (a & 15) == 0 && W1 % 4 == 0 && W2 % 4 == 0 && n_out >= 1 && n_out <= CT_MAXLEV;
}

int launch_corr1d_tc(const float* f1, const float* f2, float* const* levels, int n_out, int B, int C, int H, int W1,
                     int W2, cudaStream_t stream) {
  Corr1dTcParams p;
  Corr1dMaps maps;
  memset(&maps, 0, sizeof(maps));
  p.nlev = n_out;
  p.B = B; p.C = C; p.H = H; p.W1 = W1; p.W2 = W2;
  p.nmt = ceil_div(W1, 128); p.nnt = ceil_div(W2, 256);
  p.PB = min(256, ceil_div(W2, 64) * 64);
  p.nbB = p.PB / 64;
  p.N = min(256, ceil_div(W2, 16) * 16);
  p.nch = ceil_div(C, CT_KC);
  long long nt = 1LL * B * H * p.nmt * p.nnt;
  SF_CHECK_ARG(nt < (1LL << 30), "corr1d_pyramid: too many tiles");
  p.ntiles = static_cast<int>(nt);
  p.sqrt_c = sqrtf(static_cast<float>(C));
  p.inv_sqrt_c = 1.0f / p.sqrt_c;
  {  // sqrt(C) a power of two: multiplying by its reciprocal is bit-identical to the division
    int e = 0;
    const float m = frexpf(p.sqrt_c, &e);
    p.pow2 = (m == 0.5f);
  }
  p.nacc = (2 * p.N + 32 <= 512) ? 2 : 1;
  p.acc_stride = p.N;
  p.stage_bytes = ceil_div(CT_KC * (CT_PA + p.PB) * 4, 1024) * 1024;
  p.op_bytes = 2 * p.nbB * CT_BLK;
  p.NS = min(CT_MAXNS, (232448 - CT_NOPS * p.op_bytes - 2 * CT_EPI_BYTES - 256) / p.stage_bytes);
  SF_CHECK_ARG(p.NS >= 2, "corr1d_pyramid: shared-memory plan failed");
  int w = W2;
  for (int l = 0; l < CT_MAXLEV; ++l) {
    p.lvl[l] = l < n_out ? levels[l] : nullptr;
    p.Wl[l] = w;
    p.tma_lvl[l] = (l < n_out && l < 4 && w % 4 == 0 && w > 0) ? 1 : 0;
    w /= 2;
  }
  int rc = make_tmap_3d_f32(&maps.f1, f1, W1, H, static_cast<uint64_t>(B) * C, CT_PA, 1, CT_KC);
  if (rc) return rc;
  rc = make_tmap_3d_f32(&maps.f2, f2, W2, H, static_cast<uint64_t>(B) * C, p.PB, 1, CT_KC);
  if (rc) return rc;
  const int swz[4] = {128, 64, 32, 0};
  for (int l = 0; l < 4; ++l) {
    if (!p.tma_lvl[l]) continue;
    rc = make_tmap_3d_f32_swz(&maps.lvl[l], levels[l], p.Wl[l], W1, static_cast<uint64_t>(B) * H, 32 >> l, 128, 1, swz[l]);
    if (rc) return rc;
  }
  const size_t smem = static_cast<size_t>(p.NS) * p.stage_bytes + CT_NOPS * p.op_bytes + 2 * CT_EPI_BYTES + 256;
  SF_CUDA(cudaFuncSetAttribute(corr1d_pyramid_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               static_cast<int>(smem)));
  const int grid = min(p.ntiles, sm_count());
  corr1d_pyramid_tc_kernel<<<grid, CT_THREADS, smem, stream>>>(maps, p);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

}  // namespace sf

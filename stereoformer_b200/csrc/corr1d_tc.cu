// CorrBlock1D all-pairs correlation pyramid on tcgen05 tensor cores  (SURVEY 8a row a8).
//
// Reference: models/IterativeLocalCostVolume/corr/corr.py:55-76,103-111 (einsum 'aijk,aijh->ajkh' / sqrt(C),
// then avg_pool2d([1,2]) num_levels times).  Per epipolar row this is a [W1 x C] x [C x W2] GEMM in fp32.
//
// fp32 parity on bf16 tensor cores: every operand is split on the fly into bf16 hi + bf16 mid
// (x ~= hi + mid, residual <= 2^-17 |x|) and the product is accumulated as hi*hi + hi*mid + mid*hi in fp32 TMEM
// (three tcgen05.mma per 16-channel step), which keeps the result within ~1e-5 of the fp32 einsum.
//   * producer warps (256 threads) stream fp32 rows with coalesced 128-bit loads, split them and write both
//     halves straight into the canonical MN-major SW128 UMMA layout (no fp32 staging in shared memory);
//   * one thread issues the MMAs; accumulators are double-buffered in TMEM (2 x 256 columns);
//   * 4 epilogue warps read the accumulator (thread = x1 row, 32 consecutive x2 per tcgen05.ld), divide by
//     sqrt(C), and emit level 0 AND all pooled levels from the same registers: the volume is written once.
// Work tile: (b, y, 128-wide x1 tile, 256-wide x2 tile).
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_common.cuh"
#include "../../include/stereoformer_b200.h"

namespace sf {

constexpr int CT_KC = 32;                      // channels per stage
constexpr int CT_NSTAGE = 4;
constexpr int CT_A_HALF = CT_KC * 128 * 2;     // A hi (or mid): [2 MN blocks][32 rows][128 B] = 8 KB
constexpr int CT_B_HALF = CT_KC * 256 * 2;     // B hi (or mid): [4 MN blocks][32 rows][128 B] = 16 KB
constexpr int CT_STAGE = 2 * CT_A_HALF + 2 * CT_B_HALF;   // 48 KB
constexpr int CT_BLK = CT_KC * 128;            // one MN block of 64 elements: 32 rows x 128 B = 4 KB (the LBO)
constexpr int CT_NPROD = 256;                  // producer threads (warps 0-7)
constexpr int CT_THREADS = CT_NPROD + 32 + 128;  // + MMA warp (8) + epilogue warps (9-12)
constexpr int CT_MAXLEV = 5;

struct Corr1dTcParams {
  const float* f1;
  const float* f2;
  float* lvl[CT_MAXLEV];
  int Wl[CT_MAXLEV];
  int nlev;
  int B, C, H, W1, W2;
  int nmt, nnt;   // x1 tiles of 128, x2 tiles of 256
  int ntiles;
  float sqrt_c;
};

// write 4 consecutive MN elements (x .. x+3, x % 4 == 0) of K-row k into an MN-major SW128 operand buffer
__device__ __forceinline__ void store_split4(unsigned char* hi_base, unsigned char* mid_base, int k, int x, float4 v) {
  __nv_bfloat162 h01 = __floats2bfloat162_rn(v.x, v.y), h23 = __floats2bfloat162_rn(v.z, v.w);
  __nv_bfloat162 m01 = __floats2bfloat162_rn(v.x - __low2float(h01), v.y - __high2float(h01));
  __nv_bfloat162 m23 = __floats2bfloat162_rn(v.z - __low2float(h23), v.w - __high2float(h23));
  const int off = (x >> 6) * CT_BLK + (k >> 3) * 1024 + (k & 7) * 128 + ((((x & 63) >> 3) ^ (k & 7)) << 4) + (x & 7) * 2;
  *reinterpret_cast<uint2*>(hi_base + off) = make_uint2(*reinterpret_cast<uint32_t*>(&h01), *reinterpret_cast<uint32_t*>(&h23));
  *reinterpret_cast<uint2*>(mid_base + off) = make_uint2(*reinterpret_cast<uint32_t*>(&m01), *reinterpret_cast<uint32_t*>(&m23));
}

__global__ void __launch_bounds__(CT_THREADS, 1)
corr1d_pyramid_tc_kernel(const Corr1dTcParams p) {
  using namespace sftc;
  extern __shared__ __align__(1024) unsigned char smem[];
  unsigned char* sStage = smem;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + CT_NSTAGE * CT_STAGE);
  uint64_t* full = bars;                    // [NSTAGE]  producers -> MMA   (count 256)
  uint64_t* empty = bars + CT_NSTAGE;       // [NSTAGE]  MMA commit -> producers
  uint64_t* acc_full = bars + 2 * CT_NSTAGE;      // [2]
  uint64_t* acc_empty = bars + 2 * CT_NSTAGE + 2; // [2]  (count 128)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * CT_NSTAGE + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < CT_NSTAGE; ++s) { mbar_init(&full[s], CT_NPROD); mbar_init(&empty[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&acc_full[s], 1); mbar_init(&acc_empty[s], 128); }
    fence_mbar_init();
  }
  if (warp == 8) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const int nkc = (p.C + CT_KC - 1) / CT_KC;
  const size_t cstride1 = static_cast<size_t>(p.H) * p.W1, cstride2 = static_cast<size_t>(p.H) * p.W2;

  if (warp < 8) {
    // =========================================================== producers: load fp32, split, store operands
    const int tid = threadIdx.x;
    int it = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
      int nt = tile % p.nnt, r = tile / p.nnt;
      int mt = r % p.nmt; r /= p.nmt;
      const int y = r % p.H, b = r / p.H;
      const int m0 = mt * 128, n0 = nt * 256;
      const float* a_row = p.f1 + (static_cast<size_t>(b) * p.C * p.H + y) * p.W1;
      const float* b_row = p.f2 + (static_cast<size_t>(b) * p.C * p.H + y) * p.W2;
      for (int kc = 0; kc < nkc; ++kc, ++it) {
        const int s = it % CT_NSTAGE;
        mbar_wait(&empty[s], ((it / CT_NSTAGE) & 1) ^ 1);
        unsigned char* st = sStage + s * CT_STAGE;
        unsigned char* a_hi = st, *a_mid = st + CT_A_HALF, *b_hi = st + 2 * CT_A_HALF, *b_mid = b_hi + CT_B_HALF;
        const int c0 = kc * CT_KC;
        // 12 float4 per thread: 4 of A (32 rows x 32 float4), 8 of B (32 rows x 64 float4); loads first
        float4 va[4], vb[8];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int item = tid + i * CT_NPROD, k = item >> 5, x = m0 + (item & 31) * 4;
          va[i] = (c0 + k < p.C && x < p.W1) ? ldg_nc_f4(a_row + (c0 + k) * cstride1 + x) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int item = tid + i * CT_NPROD, k = item >> 6, x = n0 + (item & 63) * 4;
          vb[i] = (c0 + k < p.C && x < p.W2) ? ldg_nc_f4(b_row + (c0 + k) * cstride2 + x) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int item = tid + i * CT_NPROD;
          store_split4(a_hi, a_mid, item >> 5, (item & 31) * 4, va[i]);
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int item = tid + i * CT_NPROD;
          store_split4(b_hi, b_mid, item >> 6, (item & 63) * 4, vb[i]);
        }
        fence_proxy_async_smem();
        mbar_arrive(&full[s]);
      }
    }
  } else if (warp == 8) {
    // =========================================================== MMA issuer
    if (lane == 0) {
      int it = 0, t = 0;
      for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++t) {
        const int nt = tile % p.nnt;
        const int nrem = p.W2 - nt * 256;
        const int N = min(256, ((nrem + 15) / 16) * 16);
        const uint32_t idesc = make_idesc_bf16(128, N, 1, 1);
        const int buf = t & 1;
        mbar_wait(&acc_empty[buf], ((t >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t acc = tmem + buf * 256;
        for (int kc = 0; kc < nkc; ++kc, ++it) {
          const int s = it % CT_NSTAGE;
          mbar_wait(&full[s], (it / CT_NSTAGE) & 1);
          tc_fence_after();
          const uint32_t a_hi = smem_u32(sStage + s * CT_STAGE), a_mid = a_hi + CT_A_HALF;
          const uint32_t b_hi = a_hi + 2 * CT_A_HALF, b_mid = b_hi + CT_B_HALF;
#pragma unroll
          for (int j = 0; j < CT_KC / 16; ++j) {
            const uint32_t ko = j * 2048;
            umma_bf16(acc, make_smem_desc(a_hi + ko, CT_BLK, 1024), make_smem_desc(b_hi + ko, CT_BLK, 1024), idesc,
                      (kc | j) != 0);
            umma_bf16(acc, make_smem_desc(a_hi + ko, CT_BLK, 1024), make_smem_desc(b_mid + ko, CT_BLK, 1024), idesc, true);
            umma_bf16(acc, make_smem_desc(a_mid + ko, CT_BLK, 1024), make_smem_desc(b_hi + ko, CT_BLK, 1024), idesc, true);
          }
          umma_commit(&empty[s]);
        }
        umma_commit(&acc_full[buf]);
      }
    }
  } else {
    // =========================================================== epilogue: scale, level 0 + pooled levels
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    int t = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++t) {
      int nt = tile % p.nnt, r = tile / p.nnt;
      int mt = r % p.nmt; r /= p.nmt;                       // r = b*H + y
      const int x1 = mt * 128 + row, n0 = nt * 256;
      const bool rowok = x1 < p.W1;
      const size_t pix = static_cast<size_t>(r) * p.W1 + x1;
      const int buf = t & 1;
      mbar_wait(&acc_full[buf], (t >> 1) & 1);
      tc_fence_after();
      const uint32_t trow = tmem + buf * 256 + (static_cast<uint32_t>(quarter * 32) << 16);
      const int ncb = (min(256, p.W2 - n0) + 31) / 32;
      for (int cb = 0; cb < ncb; ++cb) {
        uint32_t raw[32];
        tmem_ld32(trow + cb * 32, raw);
        tmem_ld_wait();
        float v[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = __fdiv_rn(__uint_as_float(raw[i]), p.sqrt_c);
        const int c0 = n0 + cb * 32;
        if (rowok) {
          float* o = p.lvl[0] + pix * p.Wl[0] + c0;
#pragma unroll
          for (int i = 0; i < 32; i += 4)
            if (c0 + i < p.Wl[0]) *reinterpret_cast<float4*>(o + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
        }
        // pooled levels: pairwise means in registers (avg_pool2d([1,2]), floor mode: odd tails dropped)
        int cnt = 32, cl = c0;
#pragma unroll
        for (int l = 1; l < CT_MAXLEV; ++l) {
          cnt >>= 1; cl >>= 1;
#pragma unroll
          for (int i = 0; i < 16; ++i)
            if (i < cnt) v[i] = __fmul_rn(__fadd_rn(v[2 * i], v[2 * i + 1]), 0.5f);
          if (l < p.nlev && rowok) {
            float* o = p.lvl[l] + pix * p.Wl[l] + cl;
#pragma unroll
            for (int i = 0; i < 16; ++i)
              if (i < cnt && cl + i < p.Wl[l]) o[i] = v[i];
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&acc_empty[buf]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) tmem_free(tmem, 512);
}

bool corr1d_tc_supported(const float* f1, const float* f2, float* const* levels, int n_out, int W1, int W2) {
  uintptr_t a = reinterpret_cast<uintptr_t>(f1) | reinterpret_cast<uintptr_t>(f2);
  for (int l = 0; l < n_out; ++l) a |= reinterpret_cast<uintptr_t>(levels[l]);
  return (a & 15) == 0 && W1 % 4 == 0 && W2 % 4 == 0 && n_out >= 1 && n_out <= CT_MAXLEV;
}

int launch_corr1d_tc(const float* f1, const float* f2, float* const* levels, int n_out, int B, int C, int H, int W1,
                     int W2, cudaStream_t stream) {
  Corr1dTcParams p;
  p.f1 = f1; p.f2 = f2; p.nlev = n_out;
  int w = W2;
  for (int l = 0; l < CT_MAXLEV; ++l) { p.lvl[l] = l < n_out ? levels[l] : nullptr; p.Wl[l] = w; w /= 2; }
  p.B = B; p.C = C; p.H = H; p.W1 = W1; p.W2 = W2;
  p.nmt = ceil_div(W1, 128); p.nnt = ceil_div(W2, 256);
  long long nt = 1LL * B * H * p.nmt * p.nnt;
  SF_CHECK_ARG(nt < (1LL << 30), "corr1d_pyramid: too many tiles");
  p.ntiles = static_cast<int>(nt);
  p.sqrt_c = sqrtf(static_cast<float>(C));
  const size_t smem = CT_NSTAGE * CT_STAGE + 256;
  SF_CUDA(cudaFuncSetAttribute(corr1d_pyramid_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               static_cast<int>(smem)));
  const int grid = min(p.ntiles, sm_count());
  corr1d_pyramid_tc_kernel<<<grid, CT_THREADS, smem, stream>>>(p);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

}  // namespace sf

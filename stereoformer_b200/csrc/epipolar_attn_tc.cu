// Epipolar band attention on the 5th-generation tensor cores (tcgen05 + TMEM)  (SURVEY 8a row a11).
//
// SPEC-DERIVED operator (the reference file is empty, see epipolar_attn.cu).  Per row y and tile of 128 queries:
//   S[128 x 192] = Q^T K / sqrt(C) over the 192 keys [q0-64, q0+128) that the band (<= 64) can touch,
//   P = softmax over the valid band (0 <= q-k < band, k >= 0),  O[128 x C] = P V^T.
// Both GEMMs are tcgen05.mma (bf16 in, fp32 accumulate in TMEM):
//   * Q, K, V tiles come straight from the NCHW tensors with TMA boxes [64 channels][64 x] and 128-byte swizzle.
//     The same box layout is an MN-major operand for Q^T K (K-dim = channels) and a K-major B operand for P V^T
//     (K-dim = keys), so no transposition or staging pass exists.
//   * S lives in TMEM columns [0,192), O in [256,256+C).  The soft-max warps read S with tcgen05.ld (thread = query
//     row), write un-normalised bf16 P rows into the canonical K-major SW128 layout and keep 1/sum in a register;
//     the epilogue scales O rows and writes bf16.
// Warp roles: warp 0 TMA producer, warp 1 MMA issuer (+TMEM allocation), warps 2-5 soft-max / epilogue.
// Algorithmic bytes: 4*C*2 per pixel (Q, K, V read once, O written once).
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_common.cuh"
#include "../../include/stereoformer_b200.h"

namespace sf {

constexpr int AT_BOX = 8192;            // one TMA box: 64 rows x 128 B
constexpr int AT_STAGE = 5 * AT_BOX;    // Q chunk (2 boxes) + K chunk (3 boxes) of 64 channels
constexpr int AT_NSTAGE = 2;
constexpr int AT_PBYTES = 3 * 16384;    // P: 3 key blocks x [128 rows][128 B]
constexpr int AT_THREADS = 192;

struct AttnParams {
  __nv_bfloat16* out;
  int B, C, H, W, band;
  int nqt;      // query tiles per row
  int ntiles;
  float scale_log2;  // log2(e) / sqrt(C)
};

__global__ void __launch_bounds__(AT_THREADS, 1)
epipolar_attn_tc_kernel(const __grid_constant__ CUtensorMap mapQ, const __grid_constant__ CUtensorMap mapK,
                        const __grid_constant__ CUtensorMap mapV, const AttnParams p) {
  using namespace sftc;
  extern __shared__ __align__(1024) unsigned char smem[];
  const int NC = p.C / 64;                       // 64-channel chunks
  unsigned char* sStage = smem;                  // [AT_NSTAGE][AT_STAGE]
  unsigned char* sV = sStage + AT_NSTAGE * AT_STAGE;   // [3 key blocks][NC][64 rows][128 B]
  unsigned char* sP = sV + 3 * NC * AT_BOX;      // [3][128][128 B]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + AT_PBYTES);
  uint64_t* qk_full = bars;            // [2]
  uint64_t* qk_empty = bars + 2;       // [2]
  uint64_t* v_full = bars + 4;
  uint64_t* v_empty = bars + 5;
  uint64_t* s_full = bars + 6;
  uint64_t* p_full = bars + 7;
  uint64_t* o_full = bars + 8;
  uint64_t* o_empty = bars + 9;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < AT_NSTAGE; ++s) { mbar_init(&qk_full[s], 1); mbar_init(&qk_empty[s], 1); }
    mbar_init(v_full, 1); mbar_init(v_empty, 1); mbar_init(s_full, 1); mbar_init(p_full, 128);
    mbar_init(o_full, 1); mbar_init(o_empty, 128);
    fence_mbar_init();
    tma_prefetch_desc(&mapQ); tma_prefetch_desc(&mapK); tma_prefetch_desc(&mapV);
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // =========================================================== TMA producer
    if (lane == 0) {
      int t = 0;
      for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++t) {
        const int qt = tile % p.nqt, r = tile / p.nqt;
        const int y = r % p.H, b = r / p.H;
        const int q0 = qt * 128, k0 = q0 - 64;
        for (int cc = 0; cc < NC; ++cc) {
          const int it = t * NC + cc, s = it % AT_NSTAGE;
          mbar_wait(&qk_empty[s], ((it / AT_NSTAGE) & 1) ^ 1);
          unsigned char* st = sStage + s * AT_STAGE;
          mbar_expect_tx(&qk_full[s], AT_STAGE);
          const int ch = b * p.C + cc * 64;
          for (int j = 0; j < 2; ++j) tma_load_3d(st + j * AT_BOX, &mapQ, &qk_full[s], q0 + 64 * j, y, ch);
          for (int j = 0; j < 3; ++j) tma_load_3d(st + (2 + j) * AT_BOX, &mapK, &qk_full[s], k0 + 64 * j, y, ch);
        }
        mbar_wait(v_empty, (t & 1) ^ 1);
        mbar_expect_tx(v_full, 3 * NC * AT_BOX);
        for (int kb = 0; kb < 3; ++kb)
          for (int cc = 0; cc < NC; ++cc)
            tma_load_3d(sV + (kb * NC + cc) * AT_BOX, &mapV, v_full, k0 + 64 * kb, y, b * p.C + cc * 64);
      }
    }
  } else if (warp == 1) {
    // =========================================================== MMA issuer
    if (lane == 0) {
      const uint32_t idesc_qk = make_idesc_bf16(128, 192, 1, 1);
      const uint32_t idesc_pv = make_idesc_bf16(128, p.C, 0, 0);
      int t = 0;
      for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++t) {
        for (int cc = 0; cc < NC; ++cc) {
          const int it = t * NC + cc, s = it % AT_NSTAGE;
          mbar_wait(&qk_full[s], (it / AT_NSTAGE) & 1);
          tc_fence_after();
          const uint32_t qa = smem_u32(sStage + s * AT_STAGE), ka = qa + 2 * AT_BOX;
#pragma unroll
          for (int j = 0; j < 4; ++j)
            umma_bf16(tmem, make_smem_desc(qa + j * 2048, AT_BOX, 1024), make_smem_desc(ka + j * 2048, AT_BOX, 1024),
                      idesc_qk, (cc | j) != 0);
          umma_commit(&qk_empty[s]);
        }
        umma_commit(s_full);
        mbar_wait(p_full, t & 1);
        mbar_wait(v_full, t & 1);
        mbar_wait(o_empty, (t & 1) ^ 1);
        tc_fence_after();
        const uint32_t pa = smem_u32(sP), va = smem_u32(sV);
        for (int kb = 0; kb < 3; ++kb)
#pragma unroll
          for (int j = 0; j < 4; ++j)
            umma_bf16(tmem + 256, make_smem_desc(pa + kb * 16384 + j * 32, 16, 1024),
                      make_smem_desc(va + kb * NC * AT_BOX + j * 32, 16, 1024), idesc_pv, (kb | j) != 0);
        umma_commit(v_empty);
        umma_commit(o_full);
      }
    }
  } else {
    // =========================================================== soft-max + epilogue (128 threads, thread = query row)
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t trow = tmem + (static_cast<uint32_t>(quarter * 32) << 16);
    const size_t HW = static_cast<size_t>(p.H) * p.W;
    int t = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++t) {
      const int qt = tile % p.nqt, r = tile / p.nqt;
      const int y = r % p.H, b = r / p.H;
      const int q = qt * 128 + row;
      // valid key columns of this row: j = row + 64 - d, d in [0, min(band, q + 1))
      const int nd = (q < p.W) ? min(p.band, q + 1) : 0;
      const int jhi = row + 64, jlo = jhi - nd + 1;       // inclusive range (empty when nd == 0)
      mbar_wait(s_full, t & 1);
      tc_fence_after();
      // warp-uniform chunk range that any row of this warp can touch
      const int wlo = max(0, (quarter * 32 + 64 - p.band + 1) / 32), whi = (quarter * 32 + 31 + 64) / 32;
      float m = -INFINITY;
      for (int cb = wlo; cb <= whi; ++cb) {
        uint32_t v[32];
        tmem_ld32(trow + cb * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const int j = cb * 32 + i;
          if (j >= jlo && j <= jhi) m = fmaxf(m, __uint_as_float(v[i]));
        }
      }
      const float mb = (nd > 0) ? m * p.scale_log2 : 0.f;
      float sum = 0.f;
      for (int cb = 0; cb < 6; ++cb) {
        uint32_t pk[16];
        if (cb >= wlo && cb <= whi) {
          uint32_t v[32];
          tmem_ld32(trow + cb * 32, v);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            const int j = cb * 32 + i;
            float e0 = (j >= jlo && j <= jhi) ? exp2f(fmaf(__uint_as_float(v[i]), p.scale_log2, -mb)) : 0.f;
            float e1 = (j + 1 >= jlo && j + 1 <= jhi) ? exp2f(fmaf(__uint_as_float(v[i + 1]), p.scale_log2, -mb)) : 0.f;
            __nv_bfloat162 h = __floats2bfloat162_rn(e0, e1);
            sum += __low2float(h) + __high2float(h);
            pk[i >> 1] = *reinterpret_cast<uint32_t*>(&h);
          }
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) pk[i] = 0u;
        }
        // 32 keys = 4 chunks of 16 B in key block cb/2, chunk index (cb&1)*4 + c
        unsigned char* prow = sP + (cb >> 1) * 16384 + (row >> 3) * 1024 + (row & 7) * 128;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const int chunk = (cb & 1) * 4 + c;
          *reinterpret_cast<uint4*>(prow + ((chunk ^ (row & 7)) << 4)) =
              make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
        }
      }
      const float inv = (sum > 0.f) ? 1.0f / sum : 0.f;
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(p_full);

      // ---- epilogue: O rows from TMEM, normalise, bf16, store
      mbar_wait(o_full, t & 1);
      tc_fence_after();
      __nv_bfloat16* orow = p.out + (static_cast<size_t>(b) * p.C) * HW + static_cast<size_t>(y) * p.W + q;
      for (int c0 = 0; c0 < p.C; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(trow + 256 + c0, v);
        tmem_ld_wait();
        if (q < p.W) {
#pragma unroll
          for (int i = 0; i < 32; ++i) orow[(c0 + i) * HW] = __float2bfloat16(__uint_as_float(v[i]) * inv);
        }
      }
      tc_fence_before();
      mbar_arrive(o_empty);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_free(tmem, 512);
}

bool attn_tc_supported(const void* q, const void* k, const void* v, const void* out, int C, int W, int band) {
  const uintptr_t a = reinterpret_cast<uintptr_t>(q) | reinterpret_cast<uintptr_t>(k) | reinterpret_cast<uintptr_t>(v) |
                      reinterpret_cast<uintptr_t>(out);
  return (a & 15) == 0 && C % 64 == 0 && C >= 64 && C <= 256 && W % 8 == 0 && band >= 1 && band <= 64;
}

int launch_attn_tc(const void* q, const void* k, const void* v, void* out, int B, int C, int H, int W, int band,
                   cudaStream_t stream) {
  AttnParams p;
  p.out = static_cast<__nv_bfloat16*>(out);
  p.B = B; p.C = C; p.H = H; p.W = W; p.band = band;
  p.nqt = ceil_div(W, 128);
  long long nt = 1LL * B * H * p.nqt;
  SF_CHECK_ARG(nt < (1LL << 30), "epipolar_attn: too many tiles");
  p.ntiles = static_cast<int>(nt);
  p.scale_log2 = 1.4426950408889634f / sqrtf(static_cast<float>(C));
  CUtensorMap mq, mk, mv;
  int rc;
  if ((rc = make_tmap_3d_bf16_sw128(&mq, q, W, H, static_cast<uint64_t>(B) * C, 1, 64))) return rc;
  if ((rc = make_tmap_3d_bf16_sw128(&mk, k, W, H, static_cast<uint64_t>(B) * C, 1, 64))) return rc;
  if ((rc = make_tmap_3d_bf16_sw128(&mv, v, W, H, static_cast<uint64_t>(B) * C, 1, 64))) return rc;
  const int NC = C / 64;
  const size_t smem = AT_NSTAGE * AT_STAGE + 3 * NC * AT_BOX + AT_PBYTES + 128 + 1024;
  SF_CUDA(cudaFuncSetAttribute(epipolar_attn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               static_cast<int>(smem)));
  const int grid = min(p.ntiles, sm_count());
  epipolar_attn_tc_kernel<<<grid, AT_THREADS, smem, stream>>>(mq, mk, mv, p);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

}  // namespace sf

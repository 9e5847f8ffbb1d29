// Epipolar band attention on the 5th-generation tensor cores (tcgen05 + TMEM)  (SURVEY 8a row a11).
//
// SPEC-DERIVED operator (the reference file is empty, see epipolar_attn.cu).  Per row y and tile of 128 queries:
//   S[128 x 192] = Q^T K / sqrt(C) over the 192 keys [q0-64, q0+128) that the band (<= 64) can touch,
//   P = softmax over the valid band (0 <= q-k < band, k >= 0),  O[128 x C] = P V^T.
// Both GEMMs are tcgen05.mma (bf16 in, fp32 accumulate in TMEM):
//   * Q, K, V tiles come straight from the NCHW tensors with TMA boxes [64 channels][64 x] and 128-byte swizzle.
//     The same box layout is an MN-major operand for Q^T K (K-dim = channels) and a K-major B operand for P V^T
//     (K-dim = keys), so no transposition or staging pass exists.
//   * S lives in TMEM (two accumulators of 192 columns for C <= 128, one above), O behind them.  The soft-max warps read
//     S with tcgen05.ld (thread = query row).  C <= 128: the un-normalised bf16 P row goes back into the same accumulator
//     (tcgen05.st, 96 columns) and P V reads its A operand from tensor memory; the S accumulator returns to Q^T K with
//     the commit behind P V, and shared memory holds two V buffers.  C > 128: P rows go to shared memory in the canonical
//     K-major SW128 layout (one V buffer).  The epilogue scales O rows by 1/sum and writes bf16.
// Warp roles: warp 0 TMA producer, warp 1 MMA issuer (+TMEM allocation), warps 2-9 soft-max, warp 10 TMA store of the
// output tile, warps 11-14 epilogue (O rows x 1/sum -> bf16): the epilogue of tile t runs beside the soft-max of tile t+1.
// Algorithmic bytes: 4*C*2 per pixel (Q, K, V read once, O written once).
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_common.cuh"
#include <cstdlib>
#include "../../include/stereoformer_b200.h"

namespace sf {

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

constexpr int AT_BOX = 8192;            // one TMA box: 64 rows x 128 B
constexpr int AT_STAGE = 5 * AT_BOX;    // Q chunk (2 boxes) + K chunk (3 boxes) of 64 channels
constexpr int AT_MAXSTAGE = 4;          // Q/K ring depth limit (barrier slots)
constexpr int AT_PBYTES = 3 * 16384;    // P: 3 key blocks x [128 rows][128 B]
constexpr int AT_THREADS = 480;         // TMA warp, MMA warp, 8 soft-max warps (two per TMEM lane quarter), TMA-store warp, 4 epilogue warps
constexpr int AT_WARP_EPI = 11;         // epilogue warps 11..14 (warp & 3 = TMEM lane quarter)

struct AttnParams {
  __nv_bfloat16* out;
  int B, C, H, W, band;
  int nqt;      // query tiles per row
  int ntiles;
  float scale_log2;  // log2(e) / sqrt(C)
  int nS, nO;        // S and O accumulators in TMEM (192 and C columns each)
  int nstage;        // Q/K ring depth (2..4): as many 64-channel stages as shared memory holds
  int nsum;          // slots of the per-row soft-max sums handed to the epilogue warps (4, or 1 where shared memory is full)
  int o_col;         // TMEM column of the O accumulator
  int tma_out;       // epilogue through a shared-memory tile + TMA store (C <= 128)
  int ptm;           // P in tensor memory (over the S accumulator it came from) (C <= 128)
  int nv;            // V buffers (2 with ptm: the P buffer's 48 KB)
  int dbg;           // tuning only: 1 skip soft-max/epilogue arithmetic, 2 skip the MMAs, 4 skip V loads' consumers
  long long* ts;     // tuning only: per-role clock64 timeline of CTA 0 (NULL in production)
};

__device__ __forceinline__ void at_mark(const AttnParams& p, int role, int& n) {
  if (p.ts != nullptr && blockIdx.x == 0 && n < 128) p.ts[role * 128 + n++] = clock64();
}

__global__ void __launch_bounds__(AT_THREADS, 1)
epipolar_attn_tc_kernel(const __grid_constant__ CUtensorMap mapQ, const __grid_constant__ CUtensorMap mapK,
                        const __grid_constant__ CUtensorMap mapV, const __grid_constant__ CUtensorMap mapO,
                        const AttnParams p) {
  using namespace sftc;
  extern __shared__ __align__(1024) unsigned char smem[];
  const int NC = p.C / 64;                       // 64-channel chunks
  unsigned char* sStage = smem;                  // [nstage][AT_STAGE]
  const int nV = p.nv;                           // V buffers
  const int vbytes = 3 * NC * AT_BOX;
  unsigned char* sV = sStage + p.nstage * AT_STAGE;   // [nV][3 key blocks][NC][64 rows][128 B]
  unsigned char* sP = sV + nV * vbytes;          // [3][128][128 B]   (not with ptm: P lives in tensor memory)
  unsigned char* sO = sP + (p.ptm ? 0 : AT_PBYTES);   // [NC][2 query halves][64 ch][128 B] output tile (tma_out only)
  uint64_t* bars = reinterpret_cast<uint64_t*>(sO + (p.tma_out ? 2 * NC * AT_BOX : 0));
  uint64_t* qk_full = bars;            // [AT_MAXSTAGE]
  uint64_t* qk_empty = bars + 4;       // [AT_MAXSTAGE]
  uint64_t* v_full = bars + 8;         // [2]
  uint64_t* v_empty = bars + 10;       // [2]
  uint64_t* s_full = bars + 12;        // [2]
  uint64_t* s_empty = bars + 14;       // [2] S read by the soft-max (256 arrivals); with ptm: P V has read the P written over it
  uint64_t* p_full = bars + 16;
  uint64_t* o_full = bars + 17;        // [2]
  uint64_t* o_empty = bars + 19;       // [2]
  uint64_t* o_staged = bars + 21;      // output tile written to shared memory (128 arrivals)
  uint64_t* so_free = bars + 22;       // output tile read by the TMA store
  uint64_t* sum_free = bars + 23;      // [4] epilogue threads have read the sums of the slot
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 27);
  float* sRed = reinterpret_cast<float*>(bars + 28);   // row maxima [2 halves][128 rows], then partial sums [nsum slots][2 halves][128 rows]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.nstage; ++s) { mbar_init(&qk_full[s], 1); mbar_init(&qk_empty[s], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&v_full[i], 1); mbar_init(&v_empty[i], 1); }
    mbar_init(p_full, 256);
    for (int i = 0; i < 2; ++i) { mbar_init(&s_full[i], 1); mbar_init(&s_empty[i], p.ptm ? 1 : 256); }
    for (int i = 0; i < 2; ++i) { mbar_init(&o_full[i], 1); mbar_init(&o_empty[i], 128); }
    mbar_init(o_staged, 128); mbar_init(so_free, 1);
    for (int i = 0; i < 4; ++i) mbar_init(&sum_free[i], 128);
    fence_mbar_init();
    tma_prefetch_desc(&mapQ); tma_prefetch_desc(&mapK); tma_prefetch_desc(&mapV); tma_prefetch_desc(&mapO);
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // =========================================================== TMA producer
    if (lane == 0) {
      int t = 0, nts = 0;
      for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++t) {
        const int qt = tile % p.nqt, r = tile / p.nqt;
        const int y = r % p.H, b = r / p.H;
        const int q0 = qt * 128, k0 = q0 - 64;
        for (int cc = 0; cc < NC; ++cc) {
          const int it = t * NC + cc, s = it % p.nstage;
          mbar_wait(&qk_empty[s], ((it / p.nstage) & 1) ^ 1);
          at_mark(p, 0, nts);
          unsigned char* st = sStage + s * AT_STAGE;
          mbar_expect_tx(&qk_full[s], AT_STAGE);
          const int ch = b * p.C + cc * 64;
          for (int j = 0; j < 2; ++j) tma_load_3d(st + j * AT_BOX, &mapQ, &qk_full[s], q0 + 64 * j, y, ch);
          for (int j = 0; j < 3; ++j) tma_load_3d(st + (2 + j) * AT_BOX, &mapK, &qk_full[s], k0 + 64 * j, y, ch);
        }
        const int vb = t % nV;
        mbar_wait(&v_empty[vb], ((t / nV) & 1) ^ 1);
        at_mark(p, 0, nts);
        mbar_expect_tx(&v_full[vb], 3 * NC * AT_BOX);
        for (int kb = 0; kb < 3; ++kb)
          for (int cc = 0; cc < NC; ++cc)
            tma_load_3d(sV + vb * vbytes + (kb * NC + cc) * AT_BOX, &mapV, &v_full[vb], k0 + 64 * kb, y, b * p.C + cc * 64);
      }
    }
  } else if (warp == 1) {
    // =========================================================== MMA issuer
    if (lane == 0) {
      const uint32_t idesc_qk = make_idesc_bf16(128, 192, 1, 1);
      const uint32_t idesc_pv = make_idesc_bf16(128, p.C, 0, 0);
      const int ntl = (p.ntiles - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
      // Two independent instruction streams, issued in whatever order their inputs arrive: S = Q^T K of tile qk_t
      // (one 64-channel chunk at a time, into S accumulator qk_t % nS) and O = P V of tile pv_t.  Issuing them in a
      // fixed order would stall the TMA stream: the V buffer is only released by P V, the Q/K stages only by Q^T K.
      int it = 0, qk_t = 0, qk_cc = 0, pv_t = 0, nts = 0;
      bool s_ok = false;
      while (pv_t < ntl) {
        if (qk_t < ntl) {
          const int sb = qk_t % p.nS;
          if (!s_ok) s_ok = mbar_test(&s_empty[sb], ((qk_t / p.nS) & 1) ^ 1);
          const int s = it % p.nstage;
          if (s_ok && mbar_test(&qk_full[s], (it / p.nstage) & 1)) {
            tc_fence_after();
            const uint32_t qa = smem_u32(sStage + s * AT_STAGE), ka = qa + 2 * AT_BOX;
            if (!(p.dbg & 2))
#pragma unroll
            for (int j = 0; j < 4; ++j)
              umma_bf16(tmem + sb * 192, make_smem_desc(qa + j * 2048, AT_BOX, 1024),
                        make_smem_desc(ka + j * 2048, AT_BOX, 1024), idesc_qk, (qk_cc | j) != 0);
            umma_commit(&qk_empty[s]);
            at_mark(p, 1, nts);
            ++it;
            if (++qk_cc == NC) { umma_commit(&s_full[sb]); qk_cc = 0; ++qk_t; s_ok = false; }
          }
        }
        const int ob = pv_t % p.nO, vb = pv_t % nV;
        if (mbar_test(p_full, pv_t & 1) && mbar_test(&v_full[vb], (pv_t / nV) & 1) &&
            mbar_test(&o_empty[ob], ((pv_t / p.nO) & 1) ^ 1)) {
          tc_fence_after();
          const uint32_t pa = smem_u32(sP), va = smem_u32(sV + vb * vbytes);
          if (!(p.dbg & 2)) {
            if (p.ptm) {
              // A = P from tensor memory: 8 columns (16 keys, two bf16 per column) per k-step, over S accumulator pv_t % nS
              const uint32_t pt = tmem + (pv_t % p.nS) * 192;
              for (int kb = 0; kb < 3; ++kb)
#pragma unroll
                for (int j = 0; j < 4; ++j)
                  umma_bf16_ts(tmem + p.o_col + ob * p.C, pt + (kb * 4 + j) * 8,
                               make_smem_desc(va + kb * NC * AT_BOX + j * 32, 16, 1024), idesc_pv, (kb | j) != 0);
            } else {
              for (int kb = 0; kb < 3; ++kb)
#pragma unroll
                for (int j = 0; j < 4; ++j)
                  umma_bf16(tmem + p.o_col + ob * p.C, make_smem_desc(pa + kb * 16384 + j * 32, 16, 1024),
                            make_smem_desc(va + kb * NC * AT_BOX + j * 32, 16, 1024), idesc_pv, (kb | j) != 0);
            }
          }
          umma_commit(&v_empty[vb]);
          if (p.ptm) umma_commit(&s_empty[pv_t % p.nS]);       // the S accumulator that held P is free for Q^T K again
          umma_commit(&o_full[ob]);
          at_mark(p, 1, nts);
          ++pv_t;
        }
      }
    }
  } else if (warp == 10) {
    // =========================================================== TMA-store warp (output tiles, C <= 128)
    if (lane == 0 && p.tma_out) {
      int t = 0;
      for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++t) {
        const int qt = tile % p.nqt, r = tile / p.nqt;
        const int y = r % p.H, b = r / p.H;
        mbar_wait(o_staged, t & 1);
        for (int cc = 0; cc < NC; ++cc)
          for (int qh = 0; qh < 2 && qt * 128 + 64 * qh < p.W; ++qh)
            asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
                         ::"l"(reinterpret_cast<uint64_t>(&mapO)), "r"(smem_u32(sO + (cc * 2 + qh) * AT_BOX)),
                           "r"(qt * 128 + 64 * qh), "r"(y), "r"(b * p.C + cc * 64)
                         : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        mbar_arrive(so_free);
      }
      asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
  } else if (warp >= AT_WARP_EPI) {
    // =========================================================== epilogue: O rows from TMEM, normalise, bf16, store
    // one thread per query row (all channels); the two partial soft-max sums of the row come through shared memory,
    // published before the soft-max threads arrived on p_full (which the MMAs behind o_full have observed)
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t trow = tmem + (static_cast<uint32_t>(quarter * 32) << 16);
    const size_t HW = static_cast<size_t>(p.H) * p.W;
    const float* sSum = sRed + 256;
    int te = 0, nts = 0;
    const bool tsw = threadIdx.x == AT_WARP_EPI * 32;
    for (int etile = blockIdx.x; etile < p.ntiles; etile += gridDim.x, ++te) {
      const int qt = etile % p.nqt, r = etile / p.nqt;
      const int y = r % p.H, b = r / p.H;
      const int q = qt * 128 + row;
      const int ob = te % p.nO;
      mbar_wait(&o_full[ob], (te / p.nO) & 1);
      tc_fence_after();
      if (tsw) at_mark(p, 2, nts);
      const int ss = te % p.nsum;
      const float mysum = sSum[ss * 256 + row] + sSum[ss * 256 + 128 + row];
      mbar_arrive(&sum_free[ss]);
      const float inv = (mysum > 0.f) ? 1.0f / mysum : 0.f;
      const uint32_t orow_t = trow + p.o_col + ob * p.C;
      if (p.tma_out) {
        // O rows -> [64 ch][64 q] bf16 boxes (SWIZZLE_128B, the layout of the input boxes) -> one TMA store per box
        mbar_wait(so_free, (te & 1) ^ 1);
        for (int c0 = 0; c0 < p.C && !(p.dbg & 1); c0 += 32) {
          uint32_t v[32];
          tmem_ld32(orow_t + c0, v);
          tmem_ld_wait();
          unsigned char* box = sO + ((c0 >> 6) * 2 + (row >> 6)) * AT_BOX + (row & 7) * 2;
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int c = (c0 & 63) + i;
            *reinterpret_cast<__nv_bfloat16*>(box + c * 128 + ((((row & 63) >> 3) ^ (c & 7)) << 4)) =
                __float2bfloat16(__uint_as_float(v[i]) * inv);
          }
        }
        tc_fence_before();
        mbar_arrive(&o_empty[ob]);
        fence_proxy_async_smem();
        mbar_arrive(o_staged);
      } else {
        __nv_bfloat16* orow = p.out + (static_cast<size_t>(b) * p.C) * HW + static_cast<size_t>(y) * p.W + q;
        for (int c0 = 0; c0 < p.C; c0 += 32) {
          uint32_t v[32];
          tmem_ld32(orow_t + c0, v);
          tmem_ld_wait();
          if (q < p.W) {
#pragma unroll
            for (int i = 0; i < 32; ++i) orow[(c0 + i) * HW] = __float2bfloat16(__uint_as_float(v[i]) * inv);
          }
        }
        tc_fence_before();
        mbar_arrive(&o_empty[ob]);
      }
      if (tsw) at_mark(p, 2, nts);
    }
  } else {
    // =========================================================== soft-max
    // 256 threads: a query row is shared by two threads (warps w and w + 4 of the same TMEM lane quarter) which take
    // alternate 32-key chunks of the band; row maxima are exchanged through shared memory, the partial sums go to the
    // epilogue warps the same way (slots handed back through sum_free).
    const int quarter = warp & 3, half = (warp - 2) >> 2;
    const int row = quarter * 32 + lane;
    const uint32_t trow = tmem + (static_cast<uint32_t>(quarter * 32) << 16);
    int nts = 0;
    const bool tsw = threadIdx.x == 64;
    int t = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++t) {
      float* sMax = sRed;                            // [2 halves][128]
      const int ss = t % p.nsum;
      float* sSum = sRed + 256 + ss * 256;           // [2 halves][128]
      const int qt = tile % p.nqt;
      const int q = qt * 128 + row;
      // valid key columns of this row: j = row + 64 - d, d in [0, min(band, q + 1))
      const int nd = (q < p.W) ? min(p.band, q + 1) : 0;
      const int jhi = row + 64, jlo = jhi - nd + 1;       // inclusive range (empty when nd == 0)
      const int sb = t % p.nS;
      const uint32_t srow = trow + sb * 192;
      mbar_wait(&s_full[sb], (t / p.nS) & 1);
      tc_fence_after();
      if (tsw) at_mark(p, 3, nts);
      // warp-uniform chunk range that any row of this warp can touch
      const int wlo = max(0, (quarter * 32 + 64 - p.band + 1) / 32), whi = (quarter * 32 + 31 + 64) / 32;
      // valid keys of a 32-key chunk as a bit mask: one LOP3 per element instead of two compares
      auto chunk_mask = [&](int cb) -> uint32_t {
        const int lo = jlo - cb * 32, hi = jhi - cb * 32;          // valid bits [lo, hi]
        if (hi < 0 || lo > 31 || lo > hi) return 0u;
        const uint32_t ml = lo <= 0 ? 0xffffffffu : (0xffffffffu << lo);
        const uint32_t mh = hi >= 31 ? 0xffffffffu : (0xffffffffu >> (31 - hi));
        return ml & mh;
      };
      // this thread's (at most two) live chunks are read from TMEM ONCE, both loads in flight together, and kept in
      // registers for both soft-max passes; S is released to the next Q^T K right away
      const int cba = wlo + half, cbb = cba + 2;
      const bool hasA = cba <= whi && !(p.dbg & 1), hasB = cbb <= whi && !(p.dbg & 1);
      uint32_t va[32], vb[32];
      if (hasA) tmem_ld32(srow + cba * 32, va);
      if (hasB) tmem_ld32(srow + cbb * 32, vb);
      const uint32_t vma = hasA ? chunk_mask(cba) : 0u, vmb = hasB ? chunk_mask(cbb) : 0u;
      tmem_ld_wait();
      if (!p.ptm) {
        tc_fence_before();
        mbar_arrive(&s_empty[sb]);
      }
      float m = -INFINITY;
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        if (vma & (1u << i)) m = fmaxf(m, __uint_as_float(va[i]));
        if (vmb & (1u << i)) m = fmaxf(m, __uint_as_float(vb[i]));
      }
      sMax[half * 128 + row] = m;
      named_bar_sync(2 + quarter, 64);                    // the two warps of this quarter
      m = fmaxf(m, sMax[(half ^ 1) * 128 + row]);
      const float scale = p.scale_log2;
      const float mb = (nd > 0) ? m * scale : 0.f;
      // P in shared memory is single-buffered: P V of the previous tile must have read it.  (With ptm, P goes over this
      // tile's own S accumulator: every S value of this lane quarter was read before the barrier above.)
      if (t > 0 && !p.ptm) mbar_wait(&o_full[(t - 1) % p.nO], ((t - 1) / p.nO) & 1);
      if (tsw) at_mark(p, 3, nts);
      float sum = 0.f;
      // 32 keys = 4 chunks of 16 B in key block cb/2, chunk index (cb&1)*4 + c
      auto store_p = [&](int cb, const uint32_t (&pk)[16]) {
        if (p.ptm) {                                       // columns [16 cb, 16 cb + 16) of the A operand: keys 32 cb + (2c, 2c + 1)
          tmem_st16(srow + cb * 16, pk);
          return;
        }
        unsigned char* prow = sP + (cb >> 1) * 16384 + (row >> 3) * 1024 + (row & 7) * 128;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const int chunk = (cb & 1) * 4 + c;
          *reinterpret_cast<uint4*>(prow + ((chunk ^ (row & 7)) << 4)) =
              make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
        }
      };
      auto emit = [&](int cb, const uint32_t (&v)[32], uint32_t vm) {
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          // masked keys get exponent -inf -> weight exactly 0
          const float x0 = fmaf(__uint_as_float(v[i]), scale, -mb), x1 = fmaf(__uint_as_float(v[i + 1]), scale, -mb);
          const float e0 = ex2_approx((vm & (1u << i)) ? x0 : -INFINITY);
          const float e1 = ex2_approx((vm & (2u << i)) ? x1 : -INFINITY);
          sum += e0 + e1;
          __nv_bfloat162 h = __floats2bfloat162_rn(e0, e1);
          pk[i >> 1] = *reinterpret_cast<uint32_t*>(&h);
        }
        store_p(cb, pk);
      };
      if (hasA) emit(cba, va, vma);
      if (hasB) emit(cbb, vb, vmb);
      if (!(p.dbg & 1)) {
        uint32_t zero[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) zero[i] = 0u;
        for (int cb = half; cb < 6; cb += 2)               // key chunks outside the band of this warp: P = 0
          if (cb < wlo || cb > whi) store_p(cb, zero);
      }
      mbar_wait(&sum_free[ss], ((t / p.nsum) & 1) ^ 1);   // the epilogue has read the sums this slot held before
      sSum[half * 128 + row] = sum;
      if (p.ptm) {
        tmem_st_wait();
        tc_fence_before();
        // p_full is polled by parity: no thread may arrive for tile t before the MMA thread has consumed the phase of
        // tile t - 1 (two completed phases would look like none to its test).  P V of tile t - 1 has completed (o_full)
        // only after that test succeeded.  (Without ptm the wait above, before the P stores, gives the same order.)
        if (t > 0) mbar_wait(&o_full[(t - 1) % p.nO], ((t - 1) / p.nO) & 1);
      } else {
        fence_proxy_async_smem();
      }
      mbar_arrive(p_full);
      named_bar_sync(2 + quarter, 64);                    // orders the next tile's sMax write after this tile's read
      if (tsw) at_mark(p, 3, nts);
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
  // TMEM budget (512 columns): C = 64: 2 S + 2 O; C = 128: 2 S + 1 O (the early Q^T K keeps the TMA stream ahead,
  // which measured faster than 1 S + 2 O with a deferred epilogue); C = 256: 1 S + 1 O
  p.nS = C <= 128 ? 2 : 1;
  p.nO = C == 64 ? 2 : 1;
  p.o_col = C <= 128 ? 384 : 256;
  // Shared memory: Q/K ring + V (3 key blocks) + P (48 KB) + the output tile of the TMA store (C <= 128).  Measured with the
  // clock64 timeline (tools/timeline_attn.py): a Q/K stage is reloaded ~5000 cycles after its MMAs commit and V is held
  // from its load until P V, so with one tile of buffers the load latency is the tile period.  A third stage paid for by
  // dropping the output tile (epilogue warps storing bf16 rows straight to global memory) measured SLOWER (74.5 vs 65.0
  // us at C = 128: the direct stores hold the O accumulator and with it the next P V; packing the row to bf16 in
  // registers and releasing O before the stores: 88 us -- 512 two-byte store instructions per tile from four warps take
  // longer than a tile); the ring stays at two stages.
  const int NC = C / 64;
  p.tma_out = C <= 128 ? 1 : 0;
  p.nstage = 2;
  p.nsum = C <= 128 ? 4 : 1;
  // C <= 128 (two S accumulators): the soft-max writes P over the S accumulator it read (tcgen05.st; P V takes its A
  // operand from tensor memory), which frees the 48 KB P buffer for a second V buffer: V of tile t + 1 loads during tile t
  // instead of after P V of tile t (the load latency was the tile period), and the soft-max no longer waits for the
  // previous P V.
  p.ptm = C <= 128 ? 1 : 0;
  p.nv = p.ptm ? 2 : 1;
  {
    const char* d = tuning_env("SF_ATTN_DBG");
    p.dbg = d ? atoi(d) : 0;
    const char* g = tuning_env("SF_ATTN_NSTAGE");
    if (g) p.nstage = min(atoi(g), AT_MAXSTAGE);
    const char* o = tuning_env("SF_ATTN_TMAOUT");
    if (o) p.tma_out = atoi(o);
    const char* m = tuning_env("SF_ATTN_PTM");
    if (m && p.nS == 2) { p.ptm = atoi(m); p.nv = p.ptm ? 2 : 1; }
    const char* nv = tuning_env("SF_ATTN_NV");
    if (nv) p.nv = atoi(nv) == 2 ? 2 : 1;
    const char* e = tuning_env("SF_ATTN_TS");   // tuning hook: device pointer of a [4][128] int64 timeline buffer
    p.ts = e ? reinterpret_cast<long long*>(strtoull(e, nullptr, 16)) : nullptr;
  }
  CUtensorMap mq, mk, mv, mo;
  int rc;
  if ((rc = make_tmap_3d_bf16_sw128(&mq, q, W, H, static_cast<uint64_t>(B) * C, 1, 64))) return rc;
  if ((rc = make_tmap_3d_bf16_sw128(&mk, k, W, H, static_cast<uint64_t>(B) * C, 1, 64))) return rc;
  if ((rc = make_tmap_3d_bf16_sw128(&mv, v, W, H, static_cast<uint64_t>(B) * C, 1, 64))) return rc;
  if ((rc = make_tmap_3d_bf16_sw128(&mo, out, W, H, static_cast<uint64_t>(B) * C, 1, 64))) return rc;
  const size_t smem = static_cast<size_t>(p.nstage) * AT_STAGE + p.nv * 3 * NC * AT_BOX + (p.ptm ? 0 : AT_PBYTES) +
                      (p.tma_out ? 2 * NC * AT_BOX : 0) + 256 + 1024 + p.nsum * 1024;   // barriers, row maxima, partial sums
  SF_CUDA(cudaFuncSetAttribute(epipolar_attn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               static_cast<int>(smem)));
  const int grid = min(p.ntiles, sm_count());
  epipolar_attn_tc_kernel<<<grid, AT_THREADS, smem, stream>>>(mq, mk, mv, mo, p);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

}  // namespace sf

// Correlation / group-wise / difference / concat cost volumes  (SURVEY 8a rows a1, a1', a2).
//
// Reference behaviour: models/utils/build_cost_volume.py:18-54, models/CrossAttentionCostVolume/
// GWC_Cost_Volume.py:5-24:   cv[b,g,d,y,x] = (1/Cg) * sum_{c in group g} L[b,c,y,x] * R[b,c,y,x-d]  (0 for x<d).
//
// sm_100a design (HBM-bound, fp32 FFMA):
//   * one persistent CTA per SM slot; a work tile is one epipolar row (bg, y, x-segment, d-tile);
//   * a producer warp streams channel chunks of BOTH feature rows into an 8-stage shared-memory ring with
//     TMA (cp.async.bulk.tensor.3d) + mbarrier complete_tx; the right row is fetched with a left halo of
//     NDB*DT pixels and TMA's out-of-bounds zero fill supplies the x<d zeros of the reference for free;
//   * 4 consumer warps split the channel chunks (and, for D>24, the disparity blocks); a lane owns 4
//     consecutive pixels x DT disparities in registers and reads its operands as conflict-free float4s;
//   * partial sums of the channel split are combined through shared memory in a fixed order (deterministic),
//     scaled by 1/Cg and written as coalesced float4 rows.
// Algorithmic bytes per row tile: (2*Cg + D) * W * 4.
#include <stdlib.h>

#include "common.cuh"
#include "../../include/stereoformer_b200.h"

namespace sf {

// corr_volume_tc.cu
bool corr_tc_supported(const float* L, const float* R, const float* out, int nBG, int Cg, int H, int W, int D);
int launch_corr_tc(const float* L, const float* R, float* out, int nBG, int Cg, int H, int W, int D, int dbg,
                   cudaStream_t stream);

constexpr int CV_NCW = 4;      // consumer warps
constexpr int CV_XMAX = 128;   // pixels covered by one warp (32 lanes x 4)
constexpr int CV_THREADS = (CV_NCW + 1) * 32;

struct CorrParams {
  float* out;
  int Cg, H, W, D;
  int nBG;        // B * groups
  int XS;         // segment width (multiple of 8, <= 128)
  int nseg, ndt;  // x segments per row, disparity tiles (of NDB*DT)
  int ntiles;
  int nch;        // channel chunks per tile
};

// acc[dd][t] += L[t] * R[x0 + t - dd]: one channel of a lane's 4-pixel x DT-disparity register tile.
template <int DT, int NR4>
__device__ __forceinline__ void corr_fma(float (&acc)[DT][4], const float4& l4, const float4 (&r4)[NR4]) {
  const float lv[4] = {l4.x, l4.y, l4.z, l4.w};
  float r[4 * NR4];
#pragma unroll
  for (int j = 0; j < NR4; ++j) {
    r[4 * j] = r4[j].x; r[4 * j + 1] = r4[j].y; r[4 * j + 2] = r4[j].z; r[4 * j + 3] = r4[j].w;
  }
#pragma unroll
  for (int dd = 0; dd < DT; ++dd)
#pragma unroll
    for (int t = 0; t < 4; ++t) acc[dd][t] = fmaf(lv[t], r[DT + t - dd], acc[dd][t]);
}

// DT disparities per lane, NDB disparity blocks per CTA, CK channels per stage, S stages.
template <int DT, int NDB, int CK, int S>
__global__ void __launch_bounds__(CV_THREADS, 2)
corr_volume_tma_kernel(const __grid_constant__ CUtensorMap mapL, const __grid_constant__ CUtensorMap mapR,
                       const CorrParams p) {
  constexpr int NCS = CV_NCW / NDB;  // warps splitting the channel chunks
  constexpr int DP = DT * NDB;       // left halo of the right-feature row
  constexpr int NR4 = DT / 4 + 1;    // float4s of the right row a lane needs per channel
  constexpr int NSLOT = NCS > 1 ? 2 : 0;  // reduction slots of [DT][128]

  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const int XSL = p.XS, XSR = p.XS + DP;
  const int stage_floats = CK * (XSL + XSR);
  float* stages = reinterpret_cast<float*>(smem_raw);
  float* red = stages + S * stage_floats;
  uint64_t* full = reinterpret_cast<uint64_t*>(red + NSLOT * DT * CV_XMAX);
  uint64_t* empty = full + S;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], NDB);
    }
    fence_mbar_init();
  }
  __syncthreads();

  if (warp == CV_NCW) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      tma_prefetch_desc(&mapL);
      tma_prefetch_desc(&mapR);
      const uint32_t bytes = stage_floats * 4;
      int it = 0;
      for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        int seg = tile % p.nseg, t2 = tile / p.nseg;
        int y = t2 % p.H, t3 = t2 / p.H;
        int dtile = t3 % p.ndt, bg = t3 / p.ndt;
        int xs0 = seg * p.XS, row0 = bg * p.Cg;
        int xr0 = xs0 - dtile * DP - DP;
        for (int kk = 0; kk < p.nch; ++kk, ++it) {
          int st = it % S;
          mbar_wait(&empty[st], ((it / S) & 1) ^ 1);
          float* dst = stages + st * stage_floats;
          mbar_expect_tx(&full[st], bytes);
          tma_load_3d(dst, &mapL, &full[st], xs0, y, row0 + kk * CK);
          tma_load_3d(dst + CK * XSL, &mapR, &full[st], xr0, y, row0 + kk * CK);
        }
      }
    }
    return;
  }

  // -------------------------------------------------------------------- consumers
  const int db = warp % NDB, cs = warp / NDB;
  const int lx = (4 * lane < p.XS) ? 4 * lane : 0;  // lanes past the segment recompute lane 0, never store
  const int roff = lx + (NDB - 1 - db) * DT;
  // The mbarrier test of a chunk is issued one chunk ahead (its ~90-cycle latency hides behind the FFMAs).
  bool ready = false;
  if (cs < p.nch && blockIdx.x < p.ntiles) ready = mbar_try_wait(&full[cs % S], (cs / S) & 1);
  int iter = 0;
  for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++iter) {
    float acc[DT][4];
#pragma unroll
    for (int i = 0; i < DT; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f;

    for (int kk = cs; kk < p.nch; kk += NCS) {
      const int it = iter * p.nch + kk;
      const int st = it % S;
      if (!ready) mbar_wait(&full[st], (it / S) & 1);
      {  // probe the next chunk of this warp (this tile, or the first one of the next tile)
        int itn = it + NCS;
        bool have = kk + NCS < p.nch;
        if (!have) {
          itn = (iter + 1) * p.nch + cs;
          have = (tile + static_cast<int>(gridDim.x) < p.ntiles) && cs < p.nch;
        }
        ready = have ? mbar_try_wait(&full[itn % S], (itn / S) & 1) : false;
      }
      const float* Ls = stages + st * stage_floats + lx;
      const float* Rs = stages + st * stage_floats + CK * XSL + roff;
      const int nv = min(CK, p.Cg - kk * CK);
      if (nv == CK) {
        // full chunk: branch-free, operands of channel c+1 are fetched while channel c is in the FMA pipe
        float4 lq[2], rq[2][NR4];
        lq[0] = *reinterpret_cast<const float4*>(Ls);
#pragma unroll
        for (int j = 0; j < NR4; ++j) rq[0][j] = *reinterpret_cast<const float4*>(Rs + 4 * j);
#pragma unroll
        for (int c = 0; c < CK; ++c) {
          const int cur = c & 1, nxt = cur ^ 1;
          if (c + 1 < CK) {
            lq[nxt] = *reinterpret_cast<const float4*>(Ls + (c + 1) * XSL);
#pragma unroll
            for (int j = 0; j < NR4; ++j) rq[nxt][j] = *reinterpret_cast<const float4*>(Rs + (c + 1) * XSR + 4 * j);
          }
          corr_fma<DT, NR4>(acc, lq[cur], rq[cur]);
        }
      } else {
        for (int c = 0; c < nv; ++c) {
          float4 l4 = *reinterpret_cast<const float4*>(Ls + c * XSL);
          float4 r4[NR4];
#pragma unroll
          for (int j = 0; j < NR4; ++j) r4[j] = *reinterpret_cast<const float4*>(Rs + c * XSR + 4 * j);
          corr_fma<DT, NR4>(acc, l4, r4);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[st]);
    }

    // ---- combine the channel split in a fixed order (tree through shared memory), scale, store
    int seg = tile % p.nseg, t2 = tile / p.nseg;
    int y = t2 % p.H, t3 = t2 / p.H;
    int dtile = t3 % p.ndt, bg = t3 / p.ndt;
    const int x = seg * p.XS + 4 * lane;
    const bool xok = (4 * lane < p.XS) && (x < p.W);
    const float fc = static_cast<float>(p.Cg);
    float* orow = p.out + ((static_cast<size_t>(bg) * p.D) * p.H + y) * p.W + x;
    const size_t dstride = static_cast<size_t>(p.H) * p.W;
    if (NCS == 1) {
      if (xok) {
#pragma unroll
        for (int dd = 0; dd < DT; ++dd) {
          const int d = dtile * DP + db * DT + dd;
          if (d < p.D)
            stg_cs_f4(orow + d * dstride, make_float4(__fdiv_rn(acc[dd][0], fc), __fdiv_rn(acc[dd][1], fc),
                                                      __fdiv_rn(acc[dd][2], fc), __fdiv_rn(acc[dd][3], fc)));
        }
      }
    } else {
      // slot s holds [DT][128] partial sums.  NCS=4 (one disparity block): level one folds warps 2,3 into
      // warps 0,1 (slots 0,1); NCS=2 (two blocks): warps 2,3 fold into warps 0,1 of the same block (slot = db).
      float* slot = red + (NCS == 4 ? (cs & 1) : db) * DT * CV_XMAX + 4 * lane;
      const bool upper = cs >= NCS / 2;
      if (upper) {
#pragma unroll
        for (int dd = 0; dd < DT; ++dd)
          *reinterpret_cast<float4*>(slot + dd * CV_XMAX) = make_float4(acc[dd][0], acc[dd][1], acc[dd][2], acc[dd][3]);
      }
      named_bar_sync(1, CV_NCW * 32);
      if (!upper) {
#pragma unroll
        for (int dd = 0; dd < DT; ++dd) {
          const float4 v = *reinterpret_cast<const float4*>(slot + dd * CV_XMAX);
          *reinterpret_cast<float4*>(slot + dd * CV_XMAX) =
              make_float4(acc[dd][0] + v.x, acc[dd][1] + v.y, acc[dd][2] + v.z, acc[dd][3] + v.w);
        }
      }
      named_bar_sync(2, CV_NCW * 32);
      // final: the NDB*DT output rows are dealt round-robin to the 4 warps
      for (int j = warp; j < DP; j += CV_NCW) {
        const int d = dtile * DP + j;
        if (d >= p.D) break;
        const int jb = j / DT, jd = j % DT;
        float4 s4;
        if (NCS == 4) {
          const float4 u = *reinterpret_cast<const float4*>(red + jd * CV_XMAX + 4 * lane);
          const float4 v = *reinterpret_cast<const float4*>(red + (DT + jd) * CV_XMAX + 4 * lane);
          s4 = make_float4(u.x + v.x, u.y + v.y, u.z + v.z, u.w + v.w);
        } else {
          s4 = *reinterpret_cast<const float4*>(red + (jb * DT + jd) * CV_XMAX + 4 * lane);
        }
        if (xok)
          stg_cs_f4(orow + d * dstride, make_float4(__fdiv_rn(s4.x, fc), __fdiv_rn(s4.y, fc), __fdiv_rn(s4.z, fc),
                                                    __fdiv_rn(s4.w, fc)));
      }
      named_bar_sync(3, CV_NCW * 32);
    }
  }
}

// ------------------------------------------------------------------------------------------------------
// v2: no producer warp.  Every channel-split group of warps owns a private ring of SW stages and refills a
// slot itself (one elected lane issues the two TMA boxes) right after it has drained it, so a slow warp never
// blocks the loads of the others (the shared in-order ring above couples all four warps to the slowest one).
template <int DT, int NDB, int CK, int SW, int MINB, int DBG = 0>
__global__ void __launch_bounds__(CV_NCW * 32, MINB)
corr_volume_ring_kernel(const __grid_constant__ CUtensorMap mapL, const __grid_constant__ CUtensorMap mapR,
                        const CorrParams p) {
  constexpr int NCS = CV_NCW / NDB;
  constexpr int DP = DT * NDB;
  constexpr int NR4 = DT / 4 + 1;
  constexpr int NSLOT = NCS > 1 ? 2 : 0;

  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const int XSL = p.XS, XSR = p.XS + DP;
  const int stage_floats = CK * (XSL + XSR);
  float* rings = reinterpret_cast<float*>(smem_raw);                 // [NCS][SW][stage]
  float* red = rings + NCS * SW * stage_floats;
  const int RP = p.XS;  // pitch of the reduction slots (lanes past the segment never touch them)
  uint64_t* full = reinterpret_cast<uint64_t*>(red + NSLOT * DT * RP);   // [NCS][SW]
  uint64_t* empty = full + NCS * SW;                                          // [NCS][SW] (used when NDB > 1)

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < NCS * SW; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], NDB);
    }
    fence_mbar_init();
    tma_prefetch_desc(&mapL);
    tma_prefetch_desc(&mapR);
  }
  __syncthreads();

  const int db = warp % NDB, cs = warp / NDB;
  const int lx = (4 * lane < p.XS) ? 4 * lane : 0;
  const int roff = lx + (NDB - 1 - db) * DT;
  float* ring = rings + cs * SW * stage_floats;
  uint64_t* rfull = full + cs * SW;
  uint64_t* rempty = empty + cs * SW;
  const int npw = cs < p.nch ? (p.nch - cs + NCS - 1) / NCS : 0;     // chunks of a tile handled by this ring
  const int ntl = (p.ntiles - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const int total = npw * ntl;
  const bool issuer = (db == 0) && (lane == 0);
  const uint32_t bytes = stage_floats * 4;

  // issue side state (only meaningful in the issuer lane)
  int qi = 0, i_tile = blockIdx.x, i_j = 0;
  auto issue = [&]() {
    const int slot = qi % SW;
    if (NDB > 1 && qi >= SW) mbar_wait(&rempty[slot], ((qi / SW) & 1) ^ 1);
    int seg = i_tile % p.nseg, t2 = i_tile / p.nseg;
    int y = t2 % p.H, t3 = t2 / p.H;
    int dtile = t3 % p.ndt, bg = t3 / p.ndt;
    const int xs0 = seg * p.XS, c0 = bg * p.Cg + (cs + i_j * NCS) * CK;
    float* dst = ring + slot * stage_floats;
    if (DBG != 2) {
      mbar_expect_tx(&rfull[slot], bytes);
      tma_load_3d(dst, &mapL, &rfull[slot], xs0, y, c0);
      tma_load_3d(dst + CK * XSL, &mapR, &rfull[slot], xs0 - dtile * DP - DP, y, c0);
    }
    ++qi;
    if (++i_j == npw) { i_j = 0; i_tile += gridDim.x; }
  };
  if (issuer)
    for (int i = 0; i < SW && i < total; ++i) issue();

  int qc = 0;
  for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
    float acc[DT][4];
#pragma unroll
    for (int i = 0; i < DT; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f;

    for (int j = 0; j < npw; ++j, ++qc) {
      const int slot = qc % SW;
      if (DBG != 2) mbar_wait(&rfull[slot], (qc / SW) & 1);
      const float* Ls = ring + slot * stage_floats + lx;
      const float* Rs = ring + slot * stage_floats + CK * XSL + roff;
      const int nv = min(CK, p.Cg - (cs + j * NCS) * CK);
      if (DBG == 1) {
        acc[0][0] += Ls[0];
      } else if (nv == CK) {
        float4 lq[2], rq[2][NR4];
        lq[0] = *reinterpret_cast<const float4*>(Ls);
#pragma unroll
        for (int q = 0; q < NR4; ++q) rq[0][q] = *reinterpret_cast<const float4*>(Rs + 4 * q);
#pragma unroll
        for (int c = 0; c < CK; ++c) {
          const int cur = c & 1, nxt = cur ^ 1;
          if (c + 1 < CK) {
            lq[nxt] = *reinterpret_cast<const float4*>(Ls + (c + 1) * XSL);
#pragma unroll
            for (int q = 0; q < NR4; ++q) rq[nxt][q] = *reinterpret_cast<const float4*>(Rs + (c + 1) * XSR + 4 * q);
          }
          corr_fma<DT, NR4>(acc, lq[cur], rq[cur]);
        }
      } else {
        for (int c = 0; c < nv; ++c) {
          float4 l4 = *reinterpret_cast<const float4*>(Ls + c * XSL);
          float4 r4[NR4];
#pragma unroll
          for (int q = 0; q < NR4; ++q) r4[q] = *reinterpret_cast<const float4*>(Rs + c * XSR + 4 * q);
          corr_fma<DT, NR4>(acc, l4, r4);
        }
      }
      __syncwarp();
      if (NDB > 1 && lane == 0) mbar_arrive(&rempty[slot]);
      if (issuer && qi < total) {
        fence_proxy_async();
        issue();
      }
    }

    // ---- combine the channel split in a fixed order (tree through shared memory), scale, store
    int seg = tile % p.nseg, t2 = tile / p.nseg;
    int y = t2 % p.H, t3 = t2 / p.H;
    int dtile = t3 % p.ndt, bg = t3 / p.ndt;
    const int x = seg * p.XS + 4 * lane;
    const bool xok = (4 * lane < p.XS) && (x < p.W);
    const float fc = static_cast<float>(p.Cg);
    float* orow = p.out + ((static_cast<size_t>(bg) * p.D) * p.H + y) * p.W + x;
    const size_t dstride = static_cast<size_t>(p.H) * p.W;
    if (NCS == 1) {
      if (xok) {
#pragma unroll
        for (int dd = 0; dd < DT; ++dd) {
          const int d = dtile * DP + db * DT + dd;
          if (d < p.D)
            stg_cs_f4(orow + d * dstride, make_float4(__fdiv_rn(acc[dd][0], fc), __fdiv_rn(acc[dd][1], fc),
                                                      __fdiv_rn(acc[dd][2], fc), __fdiv_rn(acc[dd][3], fc)));
        }
      }
    } else {
      float* slotp = red + (NCS == 4 ? (cs & 1) : db) * DT * RP + 4 * lane;
      const bool lok = 4 * lane < p.XS;
      const bool upper = cs >= NCS / 2;
      if (upper && lok) {
#pragma unroll
        for (int dd = 0; dd < DT; ++dd)
          *reinterpret_cast<float4*>(slotp + dd * RP) = make_float4(acc[dd][0], acc[dd][1], acc[dd][2], acc[dd][3]);
      }
      named_bar_sync(1, CV_NCW * 32);
      if (!upper && lok) {
#pragma unroll
        for (int dd = 0; dd < DT; ++dd) {
          const float4 v = *reinterpret_cast<const float4*>(slotp + dd * RP);
          *reinterpret_cast<float4*>(slotp + dd * RP) =
              make_float4(acc[dd][0] + v.x, acc[dd][1] + v.y, acc[dd][2] + v.z, acc[dd][3] + v.w);
        }
      }
      named_bar_sync(2, CV_NCW * 32);
      for (int j = warp; j < DP; j += CV_NCW) {
        const int d = dtile * DP + j;
        if (d >= p.D) break;
        const int jb = j / DT, jd = j % DT;
        float4 s4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (!xok) continue;
        if (NCS == 4) {
          const float4 u = *reinterpret_cast<const float4*>(red + jd * RP + 4 * lane);
          const float4 v = *reinterpret_cast<const float4*>(red + (DT + jd) * RP + 4 * lane);
          s4 = make_float4(u.x + v.x, u.y + v.y, u.z + v.z, u.w + v.w);
        } else {
          s4 = *reinterpret_cast<const float4*>(red + (jb * DT + jd) * RP + 4 * lane);
        }
        if (xok)
          stg_cs_f4(orow + d * dstride, make_float4(__fdiv_rn(s4.x, fc), __fdiv_rn(s4.y, fc), __fdiv_rn(s4.z, fc),
                                                    __fdiv_rn(s4.w, fc)));
      }
      named_bar_sync(3, CV_NCW * 32);
    }
  }
}

// Shape-agnostic path (W % 4 != 0 or unaligned pointers): one thread per output element.
__global__ void corr_volume_generic_kernel(const float* __restrict__ L, const float* __restrict__ R,
                                           float* __restrict__ out, int nBG, int Cg, int H, int W, int D) {
  const size_t n = static_cast<size_t>(nBG) * D * H * W;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    int x = i % W;
    size_t t = i / W;
    int y = t % H; t /= H;
    int d = t % D;
    int bg = t / D;
    float s = 0.f;
    if (x >= d) {
      const float* lp = L + (static_cast<size_t>(bg) * Cg * H + y) * W + x;
      const float* rp = R + (static_cast<size_t>(bg) * Cg * H + y) * W + x - d;
      const size_t cs = static_cast<size_t>(H) * W;
      for (int c = 0; c < Cg; ++c) s = fmaf(lp[c * cs], rp[c * cs], s);
      s = __fdiv_rn(s, static_cast<float>(Cg));
    }
    out[i] = s;
  }
}

// 5-D 'difference' / 'concat' volumes (build_cost_volume.py:21-37): pure streaming, bit-exact.
template <bool CONCAT>
__global__ void shift_volume_kernel(const float* __restrict__ L, const float* __restrict__ R,
                                    float* __restrict__ out, int B, int C, int H, int W, int D) {
  const int CO = CONCAT ? 2 * C : C;
  const size_t n = static_cast<size_t>(B) * CO * D * H * W;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    int x = i % W;
    size_t t = i / W;
    int y = t % H; t /= H;
    int d = t % D; t /= D;
    int c = t % CO;
    int b = t / CO;
    float v = 0.f;
    if (x >= d) {
      if (CONCAT) {
        v = (c < C) ? L[((static_cast<size_t>(b) * C + c) * H + y) * W + x]
                    : R[((static_cast<size_t>(b) * C + (c - C)) * H + y) * W + x - d];
      } else {
        const size_t base = ((static_cast<size_t>(b) * C + c) * H + y) * W + x;
        v = __fsub_rn(L[base], R[base - d]);
      }
    }
    out[i] = v;
  }
}

template <int DT, int NDB, int CK, int S>
static int launch_corr_tma_cfg(const float* L, const float* R, float* out, int nBG, int Cg, int H, int W, int D,
                               cudaStream_t stream) {
  CorrParams p;
  p.out = out; p.Cg = Cg; p.H = H; p.W = W; p.D = D; p.nBG = nBG;
  constexpr int DP = DT * NDB;
  constexpr int NCS = CV_NCW / NDB;
  int xcap = CV_XMAX;
  if (xcap + DP > 256) xcap = 256 - DP;  // TMA box limit for the haloed right row
  p.nseg = ceil_div(W, xcap);
  p.XS = ceil_div(ceil_div(W, p.nseg), 8) * 8;
  if (p.XS > xcap) { p.nseg += 1; p.XS = ceil_div(ceil_div(W, p.nseg), 8) * 8; }
  p.ndt = ceil_div(D, DP);
  p.nch = ceil_div(Cg, CK);
  long long nt = 1LL * nBG * p.ndt * H * p.nseg;
  SF_CHECK_ARG(nt < (1LL << 30), "corr_volume: too many tiles");
  p.ntiles = static_cast<int>(nt);

  CUtensorMap mL, mR;
  int rc = make_tmap_3d_f32(&mL, L, W, H, static_cast<uint64_t>(nBG) * Cg, p.XS, 1, CK);
  if (rc) return rc;
  rc = make_tmap_3d_f32(&mR, R, W, H, static_cast<uint64_t>(nBG) * Cg, p.XS + DP, 1, CK);
  if (rc) return rc;

  const int stage_floats = CK * (2 * p.XS + DP);
  const int nslot = NCS > 1 ? 2 : 0;
  const size_t smem = (static_cast<size_t>(S) * stage_floats + nslot * DT * CV_XMAX) * 4 + 2 * S * 8 + 256;
  auto kern = corr_volume_tma_kernel<DT, NDB, CK, S>;
  SF_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  int grid = min(p.ntiles, 2 * sm_count());
  kern<<<grid, CV_THREADS, smem, stream>>>(mL, mR, p);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

template <int DT, int NDB, int CK, int SW, int MINB = 2, int DBG = 0>
static int launch_corr_ring_cfg(const float* L, const float* R, float* out, int nBG, int Cg, int H, int W, int D,
                                cudaStream_t stream) {
  CorrParams p;
  p.out = out; p.Cg = Cg; p.H = H; p.W = W; p.D = D; p.nBG = nBG;
  constexpr int DP = DT * NDB;
  constexpr int NCS = CV_NCW / NDB;
  int xcap = CV_XMAX;
  if (xcap + DP > 256) xcap = 256 - DP;
  p.nseg = ceil_div(W, xcap);
  p.XS = ceil_div(ceil_div(W, p.nseg), 8) * 8;
  if (p.XS > xcap) { p.nseg += 1; p.XS = ceil_div(ceil_div(W, p.nseg), 8) * 8; }
  p.ndt = ceil_div(D, DP);
  p.nch = ceil_div(Cg, CK);
  long long nt = 1LL * nBG * p.ndt * H * p.nseg;
  SF_CHECK_ARG(nt < (1LL << 30), "corr_volume: too many tiles");
  p.ntiles = static_cast<int>(nt);
  CUtensorMap mL, mR;
  int rc = make_tmap_3d_f32(&mL, L, W, H, static_cast<uint64_t>(nBG) * Cg, p.XS, 1, CK);
  if (rc) return rc;
  rc = make_tmap_3d_f32(&mR, R, W, H, static_cast<uint64_t>(nBG) * Cg, p.XS + DP, 1, CK);
  if (rc) return rc;
  const int stage_floats = CK * (2 * p.XS + DP);
  const int nslot = NCS > 1 ? 2 : 0;
  const size_t smem = (static_cast<size_t>(NCS) * SW * stage_floats + nslot * DT * p.XS) * 4 + 2 * NCS * SW * 8 + 128;
  auto kern = corr_volume_ring_kernel<DT, NDB, CK, SW, MINB, DBG>;
  SF_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  int grid = min(p.ntiles, MINB * sm_count());
  kern<<<grid, CV_NCW * 32, smem, stream>>>(mL, mR, p);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

// Pipeline shape (channels per stage x stages); SF_CV_VARIANT is a tuning hook for bench sweeps only.
static int cv_variant() {
  static int v = -1;
  if (v < 0) {
    const char* e = tuning_env("SF_CV_VARIANT");
    v = e ? atoi(e) : 0;
  }
  return v;
}

template <int DT, int NDB>
static int launch_corr_tma(const float* L, const float* R, float* out, int nBG, int Cg, int H, int W, int D,
                           cudaStream_t stream) {
  constexpr int NCS = CV_NCW / NDB;
#ifdef SF_TUNING
  switch (cv_variant()) {                            // pipeline-shape sweep (tools/ only)
    case 10: return launch_corr_ring_cfg<DT, NDB, 4, 20 / NCS>(L, R, out, nBG, Cg, H, W, D, stream);
    case 11: return launch_corr_ring_cfg<DT, NDB, 8, 8 / NCS>(L, R, out, nBG, Cg, H, W, D, stream);
    case 12: return launch_corr_ring_cfg<DT, NDB, 4, 16 / NCS>(L, R, out, nBG, Cg, H, W, D, stream);
    case 13: return launch_corr_ring_cfg<DT, NDB, 8, 12 / NCS>(L, R, out, nBG, Cg, H, W, D, stream);
    case 14: return launch_corr_ring_cfg<DT, NDB, 4, 12 / NCS, 3>(L, R, out, nBG, Cg, H, W, D, stream);
    case 15: return launch_corr_ring_cfg<DT, NDB, 8, 8 / NCS, 3>(L, R, out, nBG, Cg, H, W, D, stream);
    case 16: return launch_corr_ring_cfg<DT, NDB, 4, 8 / NCS, 3>(L, R, out, nBG, Cg, H, W, D, stream);
    case 20: return launch_corr_ring_cfg<DT, NDB, 8, 8 / NCS, 2, 1>(L, R, out, nBG, Cg, H, W, D, stream);
    case 21: return launch_corr_ring_cfg<DT, NDB, 8, 8 / NCS, 2, 2>(L, R, out, nBG, Cg, H, W, D, stream);
    case 22: return launch_corr_ring_cfg<DT, NDB, 4, 16 / NCS, 2, 1>(L, R, out, nBG, Cg, H, W, D, stream);
    case 23: return launch_corr_ring_cfg<DT, NDB, 4, 16 / NCS, 2, 2>(L, R, out, nBG, Cg, H, W, D, stream);
    case 1: return launch_corr_tma_cfg<DT, NDB, 4, 8>(L, R, out, nBG, Cg, H, W, D, stream);
    case 2: return launch_corr_tma_cfg<DT, NDB, 16, 6>(L, R, out, nBG, Cg, H, W, D, stream);
    case 3: return launch_corr_tma_cfg<DT, NDB, 8, 4>(L, R, out, nBG, Cg, H, W, D, stream);
    case 4: return launch_corr_tma_cfg<DT, NDB, 16, 4>(L, R, out, nBG, Cg, H, W, D, stream);
    case 5: return launch_corr_tma_cfg<DT, NDB, 4, 16>(L, R, out, nBG, Cg, H, W, D, stream);
    case 6: return launch_corr_tma_cfg<DT, NDB, 8, 10>(L, R, out, nBG, Cg, H, W, D, stream);
    default: break;
  }
#endif
  // measured on B200 (tools/time_cv.py): one disparity block -> private-ring kernel, 3 CTAs/SM once there are
  // enough row tiles for >= 3 rounds, else 2 CTAs/SM with 8-channel stages; several blocks -> producer-warp kernel
  if (NDB == 1) {
    const long long approx_tiles = 1LL * nBG * H * ceil_div(W, CV_XMAX) * ceil_div(D, DT);
    if (approx_tiles >= 9LL * sm_count())
      return launch_corr_ring_cfg<DT, NDB, 4, 12 / NCS, 3>(L, R, out, nBG, Cg, H, W, D, stream);
    return launch_corr_ring_cfg<DT, NDB, 8, 8 / NCS>(L, R, out, nBG, Cg, H, W, D, stream);
  }
  return launch_corr_tma_cfg<DT, NDB, 8, 8>(L, R, out, nBG, Cg, H, W, D, stream);
}

}  // namespace sf

extern "C" int sf_corr_volume_f32(const float* left, const float* right, float* out, int B, int C, int H, int W,
                                  int D, int groups, int algo, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(left && right && out, "corr_volume: null pointer");
  SF_CHECK_ARG(B > 0 && C > 0 && H > 0 && W > 0 && D > 0 && groups > 0, "corr_volume: non-positive dimension");
  SF_CHECK_ARG(C % groups == 0, "corr_volume: channels (%d) not divisible by groups (%d)", C, groups);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int Cg = C / groups, nBG = B * groups;
  const bool tma_ok = (W % 4 == 0) && ((reinterpret_cast<uintptr_t>(left) | reinterpret_cast<uintptr_t>(right) |
                                        reinterpret_cast<uintptr_t>(out)) & 15) == 0;
  SF_CHECK_ARG(algo >= SF_ALGO_AUTO && algo <= SF_ALGO_EXACT, "corr_volume: unknown algo %d", algo);
  SF_CHECK_ARG(algo != SF_ALGO_TMA || tma_ok, "corr_volume: TMA path needs W %% 4 == 0 and 16-byte aligned pointers");
  {
    // tensor-core path: default whenever the channel count makes the 32-channel MMA chunks worthwhile
    const int v = cv_variant();
    const bool tc_ok = corr_tc_supported(left, right, out, nBG, Cg, H, W, D);
    SF_CHECK_ARG(algo != SF_ALGO_TC || tc_ok, "corr_volume: shape / alignment does not fit the tcgen05 kernel");
    // (tuning build: SF_CV_VARIANT 30..99 selects a component-isolation variant, and only when the caller left the
    // choice open -- an explicit algo is never overridden)
    const bool sweep = algo == SF_ALGO_AUTO && v >= 30 && v <= 99 && tc_ok;
    if (algo == SF_ALGO_TC || sweep || (algo == SF_ALGO_AUTO && v == 0 && tc_ok && Cg >= 32))
      return launch_corr_tc(left, right, out, nBG, Cg, H, W, D, sweep ? v - 30 : 0, s);
  }
  // SF_ALGO_EXACT: fp32 FFMA arithmetic, the fastest kernel that fits the shape
  if (algo == SF_ALGO_GENERIC || !tma_ok) {
    size_t n = static_cast<size_t>(nBG) * D * H * W;
    int grid = static_cast<int>(min(static_cast<size_t>(sm_count()) * 16, (n + 255) / 256));
    corr_volume_generic_kernel<<<grid, 256, 0, s>>>(left, right, out, nBG, Cg, H, W, D);
    SF_LAUNCH_CHECK();
    return SF_OK;
  }
  // pick the disparity blocking with the least padding (ties -> larger DT)
  if (D <= 8) return launch_corr_tma<8, 1>(left, right, out, nBG, Cg, H, W, D, s);
  if (D <= 16) return launch_corr_tma<16, 1>(left, right, out, nBG, Cg, H, W, D, s);
  if (D <= 24) return launch_corr_tma<24, 1>(left, right, out, nBG, Cg, H, W, D, s);
  if (D <= 32) return launch_corr_tma<16, 2>(left, right, out, nBG, Cg, H, W, D, s);
  if (D <= 48) return launch_corr_tma<24, 2>(left, right, out, nBG, Cg, H, W, D, s);
  if (D <= 64) return launch_corr_tma<16, 4>(left, right, out, nBG, Cg, H, W, D, s);
  return launch_corr_tma<24, 4>(left, right, out, nBG, Cg, H, W, D, s);
}

extern "C" int sf_shift_volume_f32(const float* left, const float* right, float* out, int B, int C, int H, int W,
                                   int D, int concat, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(left && right && out, "shift_volume: null pointer");
  SF_CHECK_ARG(B > 0 && C > 0 && H > 0 && W > 0 && D > 0, "shift_volume: non-positive dimension");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  size_t n = static_cast<size_t>(B) * (concat ? 2 * C : C) * D * H * W;
  int grid = static_cast<int>(min(static_cast<size_t>(sm_count()) * 16, (n + 255) / 256));
  if (concat)
    shift_volume_kernel<true><<<grid, 256, 0, s>>>(left, right, out, B, C, H, W, D);
  else
    shift_volume_kernel<false><<<grid, 256, 0, s>>>(left, right, out, B, C, H, W, D);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

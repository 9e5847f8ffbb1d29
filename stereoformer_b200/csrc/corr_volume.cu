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
#include "common.cuh"
#include "../../include/stereoformer_b200.h"

namespace sf {

constexpr int CV_NCW = 4;      // consumer warps
constexpr int CV_CK = 4;       // channels per pipeline stage
constexpr int CV_S = 8;        // pipeline stages
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

template <int DT, int NDB>
__global__ void __launch_bounds__(CV_THREADS, 2)
corr_volume_tma_kernel(const __grid_constant__ CUtensorMap mapL, const __grid_constant__ CUtensorMap mapR,
                       const CorrParams p) {
  constexpr int NCS = CV_NCW / NDB;  // warps splitting the channel chunks
  constexpr int DP = DT * NDB;       // left halo of the right-feature row
  constexpr int NR4 = DT / 4 + 1;    // float4s of the right row a lane needs per channel

  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const int XSL = p.XS, XSR = p.XS + DP;
  const int stage_floats = CV_CK * (XSL + XSR);
  float* stages = reinterpret_cast<float*>(smem_raw);
  float* red = stages + CV_S * stage_floats;
  uint64_t* full = reinterpret_cast<uint64_t*>(red + CV_NCW * DT * CV_XMAX);
  uint64_t* empty = full + CV_S;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < CV_S; ++s) {
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
          int st = it % CV_S;
          mbar_wait(&empty[st], ((it / CV_S) & 1) ^ 1);
          float* dst = stages + st * stage_floats;
          mbar_expect_tx(&full[st], bytes);
          tma_load_3d(dst, &mapL, &full[st], xs0, y, row0 + kk * CV_CK);
          tma_load_3d(dst + CV_CK * XSL, &mapR, &full[st], xr0, y, row0 + kk * CV_CK);
        }
      }
    }
    return;
  }

  // -------------------------------------------------------------------- consumers
  const int db = warp % NDB, cs = warp / NDB;
  const int lx = (4 * lane < p.XS) ? 4 * lane : 0;  // lanes past the segment recompute lane 0, never store
  const int roff = lx + (NDB - 1 - db) * DT;
  int iter = 0;
  for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++iter) {
    float acc[DT][4];
#pragma unroll
    for (int i = 0; i < DT; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f;

    for (int kk = cs; kk < p.nch; kk += NCS) {
      const int it = iter * p.nch + kk;
      const int st = it % CV_S;
      mbar_wait(&full[st], (it / CV_S) & 1);
      const float* Ls = stages + st * stage_floats + lx;
      const float* Rs = stages + st * stage_floats + CV_CK * XSL + roff;
      const int nv = min(CV_CK, p.Cg - kk * CV_CK);
#pragma unroll
      for (int c = 0; c < CV_CK; ++c) {
        if (c < nv) {
          const float4 l4 = *reinterpret_cast<const float4*>(Ls + c * XSL);
          const float lv[4] = {l4.x, l4.y, l4.z, l4.w};
          float r[4 * NR4];
#pragma unroll
          for (int j = 0; j < NR4; ++j) {
            const float4 v = *reinterpret_cast<const float4*>(Rs + c * XSR + 4 * j);
            r[4 * j] = v.x; r[4 * j + 1] = v.y; r[4 * j + 2] = v.z; r[4 * j + 3] = v.w;
          }
#pragma unroll
          for (int dd = 0; dd < DT; ++dd)
#pragma unroll
            for (int t = 0; t < 4; ++t) acc[dd][t] = fmaf(lv[t], r[DT + t - dd], acc[dd][t]);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[st]);
    }

    // ---- combine the channel split (fixed order), scale, store
    float* mine = red + ((cs * NDB + db) * DT) * CV_XMAX + 4 * lane;
#pragma unroll
    for (int dd = 0; dd < DT; ++dd)
      *reinterpret_cast<float4*>(mine + dd * CV_XMAX) = make_float4(acc[dd][0], acc[dd][1], acc[dd][2], acc[dd][3]);
    named_bar_sync(1, CV_NCW * 32);

    int seg = tile % p.nseg, t2 = tile / p.nseg;
    int y = t2 % p.H, t3 = t2 / p.H;
    int dtile = t3 % p.ndt, bg = t3 / p.ndt;
    const int x = seg * p.XS + 4 * lane;
    const bool xok = (4 * lane < p.XS) && (x < p.W);
    const float fc = static_cast<float>(p.Cg);
    for (int j = warp; j < DP; j += CV_NCW) {
      const int d = dtile * DP + j;
      if (d >= p.D) break;
      const int jb = j / DT, jd = j % DT;
      float4 s = *reinterpret_cast<const float4*>(red + (jb * DT + jd) * CV_XMAX + 4 * lane);
#pragma unroll
      for (int q = 1; q < NCS; ++q) {
        const float4 v = *reinterpret_cast<const float4*>(red + ((q * NDB + jb) * DT + jd) * CV_XMAX + 4 * lane);
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
      }
      if (xok) {
        s.x = __fdiv_rn(s.x, fc); s.y = __fdiv_rn(s.y, fc); s.z = __fdiv_rn(s.z, fc); s.w = __fdiv_rn(s.w, fc);
        float* o = p.out + ((static_cast<size_t>(bg) * p.D + d) * p.H + y) * p.W + x;
        stg_cs_f4(o, s);
      }
    }
    named_bar_sync(2, CV_NCW * 32);
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

template <int DT, int NDB>
static int launch_corr_tma(const float* L, const float* R, float* out, int nBG, int Cg, int H, int W, int D,
                           cudaStream_t stream) {
  CorrParams p;
  p.out = out; p.Cg = Cg; p.H = H; p.W = W; p.D = D; p.nBG = nBG;
  constexpr int DP = DT * NDB;
  int xcap = CV_XMAX;
  if (xcap + DP > 256) xcap = 256 - DP;  // TMA box limit for the haloed right row
  p.nseg = ceil_div(W, xcap);
  p.XS = ceil_div(ceil_div(W, p.nseg), 8) * 8;
  if (p.XS > xcap) { p.nseg += 1; p.XS = ceil_div(ceil_div(W, p.nseg), 8) * 8; }
  p.ndt = ceil_div(D, DP);
  p.nch = ceil_div(Cg, CV_CK);
  long long nt = 1LL * nBG * p.ndt * H * p.nseg;
  SF_CHECK_ARG(nt < (1LL << 30), "corr_volume: too many tiles");
  p.ntiles = static_cast<int>(nt);

  CUtensorMap mL, mR;
  int rc = make_tmap_3d_f32(&mL, L, W, H, static_cast<uint64_t>(nBG) * Cg, p.XS, 1, CV_CK);
  if (rc) return rc;
  rc = make_tmap_3d_f32(&mR, R, W, H, static_cast<uint64_t>(nBG) * Cg, p.XS + DP, 1, CV_CK);
  if (rc) return rc;

  const int stage_floats = CV_CK * (2 * p.XS + DP);
  const size_t smem = (static_cast<size_t>(CV_S) * stage_floats + CV_NCW * DT * CV_XMAX) * 4 + 2 * CV_S * 8 + 256;
  auto kern = corr_volume_tma_kernel<DT, NDB>;
  SF_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  int grid = min(p.ntiles, 2 * sm_count());
  kern<<<grid, CV_THREADS, smem, stream>>>(mL, mR, p);
  SF_LAUNCH_CHECK();
  return SF_OK;
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
  SF_CHECK_ARG(algo != SF_ALGO_TMA || tma_ok, "corr_volume: TMA path needs W %% 4 == 0 and 16-byte aligned pointers");
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

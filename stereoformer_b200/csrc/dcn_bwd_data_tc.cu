// Deformable convolution: data / offset / mask gradients over channels-last copies  (SURVEY 8f rank 3, row a10 backward).
//
// Reference: deform/src/cuda/modulated_deform_conv_cuda.cu:209-245 (columns = weight^T x grad_out through at::addmm, then
// col2im_coord and col2im), modulated_deform_im2col_cuda.cuh:56-126 (coordinate weights), :197-328 (the two scatters).
// Two kernels, the same split as the reference, with the column-gradient buffer in the caller's workspace:
//   1. dcn_bwd_gcol_tc_kernel: gcol[pix][k] = sum_o grad_out[o][pix] * W[o][k] on tcgen05.  M = 128 pixels of an 8 x 16
//      tile, N = 64 column entries of one chunk (the forward's k order: 64-channel block, tap, channel), K = output
//      channels.  grad_out rows ([32 o][8 rows][16 px] TMA boxes) are split into bf16 hi + mid by 4 converter warps into
//      [o rows][128 px] blocks, which read as the MN-major A operand (pixels contiguous); the weights are split once per
//      call into Wt[k][o] (K-major B operand, streamed by TMA).  Accumulators are double-buffered per chunk; the epilogue
//      writes fp32 rows of 64 entries to gcol[tile][chunk][128 px][64].
//   2. dcn_bwd_scatter_kernel (no tensor cores): the forward's gather geometry -- two pixels per warp, 16 lanes each, a
//      lane owning 4 channels of the chunk.  Per (pixel, tap) it reads its 4 column gradients (256 contiguous bytes per
//      pixel) and the four corner vectors of the channels-last input, adds w_corner * mask * gcol to the channels-last
//      grad_input with ONE vector reduction per corner (red.global.add.v4.f32: 256 contiguous bytes per pixel and corner
//      instead of 64 scattered scalar atomics), and reduces the offset / mask gradients over its channels with shuffles
//      into per-tile shared-memory accumulators (plain stores at the end: a tile owns its pixels).
// The channels-last grad_input is transposed back by the pass that made the input copy.  fp32 parity: the GEMM uses the
// 3-term bf16 split (~2^-16 per product); everything after it is fp32.
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_common.cuh"
#include "dcn_tc.cuh"
#include "../../include/stereoformer_b200.h"

namespace sf {

constexpr int DG_THREADS = 11 * 32;     // gout TMA, weight TMA, MMA, 4 converter warps, 4 epilogue warps
constexpr int DG_WARP_CVT = 3, DG_WARP_EPI = 7;
constexpr int DG_GB = 4 * 16384;        // grad_out bf16 of one 128-row block: hi [2 px blocks][128 o rows][128 B] | mid
constexpr int DG_GF = 16384;            // grad_out fp32 staging slot: [32 o][128 px]
constexpr int DG_B = 2 * 16384;         // one weight stage: hi [2 o blocks][64 k rows][128 B] | mid

struct DcnBwdDataParams {
  DcnTcParams f;
  float* gcol;            // [tiles][nchunk][128][64]
  int nmt;                // 128-row blocks of output channels per group
  int Kpad;               // nchunk * 64
  int ntile;              // B * G * nptile
};

// fp32 weights [Cout][Cig][T] -> bf16 hi / mid Wt[2][G * Kpad][Cogp]: row = column entry k (the forward's order), o contiguous
__global__ void dcn_split_weights_t_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ wt, int G, int Cog,
                                           int Cogp, int Kpad, int Cig, int T, int CB, int ntg) {
  const size_t n = static_cast<size_t>(G) * Kpad * Cogp;
  const int tpc = 64 / CB;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int o = static_cast<int>(i % Cogp);
    const size_t row = i / Cogp;
    const int k = static_cast<int>(row % Kpad), g = static_cast<int>(row / Kpad);
    const int ch = k >> 6, r = k & 63;
    const int cc = ch / ntg, tg = ch - cc * ntg;
    const int t = tg * tpc + r / CB, c = cc * 64 + r % CB;
    const float v = (t < T && o < Cog) ? w[((static_cast<size_t>(g) * Cog + o) * Cig + c) * T + t] : 0.f;
    const __nv_bfloat16 h = __float2bfloat16_rn(v);
    wt[i] = h;
    wt[n + i] = __float2bfloat16_rn(v - __bfloat162float(h));
  }
}

__device__ __forceinline__ void st_global_v8(float* addr, const uint32_t* v) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(addr), "r"(v[0]), "r"(v[1]), "r"(v[2]),
               "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
               : "memory");
}

__global__ void __launch_bounds__(DG_THREADS, 1)
dcn_bwd_gcol_tc_kernel(const __grid_constant__ CUtensorMap mapG, const __grid_constant__ CUtensorMap mapW,
                       const DcnBwdDataParams q) {
  using namespace sftc;
  const DcnTcParams& p = q.f;
  extern __shared__ __align__(1024) unsigned char smem[];
  unsigned char* sGb = smem;                                       // [nmt <= 2][DG_GB]
  unsigned char* sB = sGb + 2 * DG_GB;                             // [2][DG_B]
  unsigned char* sGf = sB + 2 * DG_B;                              // [2][DG_GF]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sGf + 2 * DG_GF);
  uint64_t* gf_full = bars;             // [2] TMA -> converters
  uint64_t* gf_empty = bars + 2;        // [2] converters (4 warps) -> TMA
  uint64_t* gb_full = bars + 4;         // converters (4 warps) -> MMA, once per tile
  uint64_t* gb_empty = bars + 5;        // MMA commit -> converters, once per tile
  uint64_t* b_full = bars + 6;          // [2] TMA -> MMA
  uint64_t* b_empty = bars + 8;         // [2] MMA commit -> TMA
  uint64_t* acc_full = bars + 10;       // [2] MMA commit -> epilogue
  uint64_t* acc_empty = bars + 12;      // [2] epilogue (4 warps) -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 14);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&gf_full[i], 1); mbar_init(&gf_empty[i], 4);
      mbar_init(&b_full[i], 1); mbar_init(&b_empty[i], 1);
      mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 4);
    }
    mbar_init(gb_full, 4); mbar_init(gb_empty, 1);
    fence_mbar_init();
    tma_prefetch_desc(&mapG); tma_prefetch_desc(&mapW);
  }
  if (warp == 2) tmem_alloc(tmem_slot, 128);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const int Cog = p.Cout / p.G;
  const int nq = (Cog + 31) >> 5;                   // staged quarters of 32 output channels per tile

  if (warp == 0) {
    // ================================================================= grad_out producer
    if (lane == 0) {
      int s = 0, ph = 1;
      for (int tile = blockIdx.x; tile < q.ntile; tile += gridDim.x) {
        const int pt = tile % p.nptile, bg = tile / p.nptile;
        const int g = bg % p.G, b = bg / p.G;
        const int ty = pt / p.tiles_x, tx = pt - ty * p.tiles_x;
        for (int oq = 0; oq < nq; ++oq) {
          mbar_wait(&gf_empty[s], ph);
          mbar_expect_tx(&gf_full[s], DG_GF);
          tma_load_3d(sGf + s * DG_GF, &mapG, &gf_full[s], tx * 16, ty * 8, b * p.Cout + g * Cog + 32 * oq);
          if (++s == 2) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================================================================= weight producer: Wt chunk x 128-row block per stage
    if (lane == 0) {
      int s = 0, ph = 1;
      for (int tile = blockIdx.x; tile < q.ntile; tile += gridDim.x) {
        const int g = (tile / p.nptile) % p.G;
        for (int ch = 0; ch < p.nchunk; ++ch)
          for (int mt = 0; mt < q.nmt; ++mt) {
            mbar_wait(&b_empty[s], ph);
            unsigned char* dst = sB + s * DG_B;
            mbar_expect_tx(&b_full[s], DG_B);
            for (int hf = 0; hf < 2; ++hf)
              for (int ob = 0; ob < 2; ++ob)
                tma_load_3d(dst + hf * 16384 + ob * 8192, &mapW, &b_full[s], mt * 128 + ob * 64, g * q.Kpad + ch * 64, hf);
            if (++s == 2) { s = 0; ph ^= 1; }
          }
      }
    }
  } else if (warp == 2) {
    // ================================================================= MMA issuer
    if (lane == 0) {
      const uint32_t idesc = make_idesc_bf16(128, 64, 1, 0);     // A: MN-major (pixels contiguous), B: K-major
      int s = 0, ph = 0, it = 0, ci = 0;
      for (int tile = blockIdx.x; tile < q.ntile; tile += gridDim.x, ++it) {
        mbar_wait(gb_full, it & 1);
        for (int ch = 0; ch < p.nchunk; ++ch, ++ci) {
          const int buf = ci & 1;
          mbar_wait(&acc_empty[buf], ((ci >> 1) & 1) ^ 1);
          tc_fence_after();
          const uint32_t acc = tmem + 64 * buf;
          for (int mt = 0; mt < q.nmt; ++mt) {
            mbar_wait(&b_full[s], ph);
            tc_fence_after();
            const uint32_t a_hi = smem_u32(sGb + mt * DG_GB), a_mid = a_hi + 2 * 16384;
            const uint32_t b_hi = smem_u32(sB + s * DG_B), b_mid = b_hi + 16384;
#pragma unroll
            for (int j = 0; j < 8; ++j) {                        // 8 k-steps of 16 output channels
              const uint32_t bo = (j >> 2) * 8192 + (j & 3) * 32;
              const uint64_t dah = make_smem_desc(a_hi + j * 2048, 16384, 1024);
              const uint64_t dam = make_smem_desc(a_mid + j * 2048, 16384, 1024);
              const uint64_t dbh = make_smem_desc(b_hi + bo, 16, 1024);
              const uint64_t dbm = make_smem_desc(b_mid + bo, 16, 1024);
              umma_bf16(acc, dah, dbh, idesc, (mt | j) != 0);
              umma_bf16(acc, dah, dbm, idesc, true);
              umma_bf16(acc, dam, dbh, idesc, true);
            }
            umma_commit(&b_empty[s]);
            if (++s == 2) { s = 0; ph ^= 1; }
          }
          umma_commit(&acc_full[buf]);
        }
        umma_commit(gb_empty);
      }
    }
  } else if (warp < DG_WARP_EPI) {
    // ================================================================= grad_out converters (as dcn_bwd_tc.cu)
    const int ct = threadIdx.x - DG_WARP_CVT * 32;
    int s = 0, ph = 0, it = 0;
    for (int tile = blockIdx.x; tile < q.ntile; tile += gridDim.x, ++it) {
      mbar_wait(gb_empty, (it & 1) ^ 1);
      for (int oq = 0; oq < 4 * q.nmt; ++oq) {
        const bool staged = oq < nq;                             // quarters past the group's channels: zero rows
        if (staged) mbar_wait(&gf_full[s], ph);
        const float* src = reinterpret_cast<const float*>(sGf + s * DG_GF);
        unsigned char* blk = sGb + (oq >> 2) * DG_GB;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int item = ct + 128 * k;
          const int rl = item >> 4, chn = item & 15;
          float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
          if (staged && 32 * oq + rl < Cog) {
            const float4 a = *reinterpret_cast<const float4*>(src + rl * 128 + chn * 8);
            const float4 c = *reinterpret_cast<const float4*>(src + rl * 128 + chn * 8 + 4);
            v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = c.x; v[5] = c.y; v[6] = c.z; v[7] = c.w;
          }
          uint32_t hi[4], mid[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * e], v[2 * e + 1]);
            const __nv_bfloat162 m = __floats2bfloat162_rn(v[2 * e] - __low2float(h), v[2 * e + 1] - __high2float(h));
            hi[e] = *reinterpret_cast<const uint32_t*>(&h);
            mid[e] = *reinterpret_cast<const uint32_t*>(&m);
          }
          const int row = 32 * (oq & 3) + rl;
          unsigned char* dst = blk + (chn >> 3) * 16384 + (row >> 3) * 1024 + (row & 7) * 128 + (((chn & 7) ^ (row & 7)) << 4);
          *reinterpret_cast<uint4*>(dst) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
          *reinterpret_cast<uint4*>(dst + 2 * 16384) = make_uint4(mid[0], mid[1], mid[2], mid[3]);
        }
        if (staged) {
          __syncwarp();
          if (lane == 0) mbar_arrive(&gf_empty[s]);
          if (++s == 2) { s = 0; ph ^= 1; }
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(gb_full);
    }
  } else {
    // ================================================================= epilogue: accumulator rows -> gcol
    const int qd = warp & 3;
    const int row = 32 * qd + lane;
    const uint32_t trow = tmem + (static_cast<uint32_t>(32 * qd) << 16);
    int ci = 0;
    for (int tile = blockIdx.x; tile < q.ntile; tile += gridDim.x) {
      for (int ch = 0; ch < p.nchunk; ++ch, ++ci) {
        const int buf = ci & 1;
        mbar_wait(&acc_full[buf], (ci >> 1) & 1);
        tc_fence_after();
        uint32_t r0[32], r1[32];
        tmem_ld32(trow + 64 * buf, r0);
        tmem_ld32(trow + 64 * buf + 32, r1);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&acc_empty[buf]);
        // a thread owns one 256-byte row: 32-byte stores (st.global.v8, sm_100) write whole sectors -- with 16-byte stores
        // the load/store unit's queue was the limiter of this kernel (ncu: lg_throttle 4.0 per issued instruction)
        float* dst = q.gcol + ((static_cast<size_t>(tile) * p.nchunk + ch) * 128 + row) * 64;
#pragma unroll
        for (int i = 0; i < 4; ++i) st_global_v8(dst + 8 * i, r0 + 8 * i);
#pragma unroll
        for (int i = 0; i < 4; ++i) st_global_v8(dst + 32 + 8 * i, r1 + 8 * i);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_free(tmem, 128);
}

// ---------------------------------------------------------------------------------------------- scatter / reduce
__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

struct ScatterDesc {       // one (deformable group, tap, pixel)
  int base;                // top-left corner element of the channels-last tensors | dx | dy << 1 (clamped)
  int flags;               // corners inside the image (bits 0..3); 0 when the sample is outside
  float lh, lw, m;
};

__global__ void __launch_bounds__(512, 2)
dcn_bwd_scatter_kernel(const DcnTcParams p, const float* __restrict__ inT, const float* __restrict__ gcol, float* ginT,
                       float* goffset, float* gmask) {
  extern __shared__ __align__(16) unsigned char sraw[];
  const int T = p.kh * p.kw, M = p.Ho * p.Wo;
  ScatterDesc* ds = reinterpret_cast<ScatterDesc*>(sraw);            // [dg][T][128]
  float* acc = reinterpret_cast<float*>(ds + p.dg * T * 128);        // [3][dg][T][128]: d/dh, d/dw, d/dmask
  const int nacc = p.dg * T * 128;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int gw = warp, hw = lane >> 4, l16 = lane & 15;
  const int Cig = p.Cin / p.G, cpdg = p.Cin / p.dg;
  const int lpt = p.CB >> 2, tpc = 64 / p.CB;
  const int ti = l16 / lpt, cl = 4 * (l16 % lpt);
  const int rowstep = p.W * p.Cin;
  const int tile = blockIdx.x;
  const int pt = tile % p.nptile, bg = tile / p.nptile;
  const int g = bg % p.G, b = bg / p.G;
  const int ty = pt / p.tiles_x, tx = pt - ty * p.tiles_x;
  // ---- descriptors of this warp's 8 pixels (modulated_deform_im2col_cuda.cuh:25-54), accumulators to zero
  const int yo = ty * 8 + (gw >> 1), xo0 = tx * 16 + 8 * (gw & 1);
  for (int e = lane; e < p.dg * T * 8; e += 32) {
    const int ex = e & 7, r = e >> 3;
    const int t = r % T, dgi = r / T;
    const int xo = xo0 + ex;
    const int pix = yo * p.Wo + xo;
    ScatterDesc d;
    d.base = 0; d.flags = 0; d.lh = 0.f; d.lw = 0.f; d.m = 0.f;
    if (yo < p.Ho && xo < p.Wo) {
      const int i = t / p.kw, j = t % p.kw;
      const size_t ob = ((static_cast<size_t>(b) * p.dg + dgi) * 2 * T) * M + pix;
      const float oh = __ldg(p.offset + ob + static_cast<size_t>(2 * t) * M);
      const float ow = __ldg(p.offset + ob + static_cast<size_t>(2 * t + 1) * M);
      d.m = p.mask ? __ldg(p.mask + ((static_cast<size_t>(b) * p.dg + dgi) * T + t) * M + pix) : 1.f;
      const float h_im = (float)(yo * p.sh - p.ph + i * p.dh) + oh;
      const float w_im = (float)(xo * p.sw - p.pw + j * p.dw) + ow;
      if (h_im > -1.f && w_im > -1.f && h_im < (float)p.H && w_im < (float)p.W) {
        const int hl = (int)floorf(h_im), wl = (int)floorf(w_im);
        const int hh_i = hl + 1, wh_i = wl + 1;
        d.lh = h_im - (float)hl; d.lw = w_im - (float)wl;
        const bool t0 = hl >= 0, l0 = wl >= 0, b1 = hh_i <= p.H - 1, r1 = wh_i <= p.W - 1;
        d.flags = (t0 && l0 ? 1 : 0) | (t0 && r1 ? 2 : 0) | (b1 && l0 ? 4 : 0) | (b1 && r1 ? 8 : 0);
        const int r0 = max(hl, 0), c0 = max(wl, 0);
        const int dy = min(hh_i, p.H - 1) - r0, dx = min(wh_i, p.W - 1) - c0;
        d.base = (r0 * p.W + c0) * p.Cin | dx | (dy << 1);
      }
    }
    const int di = (dgi * T + t) * 128 + 8 * gw + ex;
    ds[di] = d;
    acc[di] = 0.f; acc[nacc + di] = 0.f; acc[2 * nacc + di] = 0.f;
  }
  __syncwarp();
  const size_t img = static_cast<size_t>(b) * p.H * p.W * p.Cin + static_cast<size_t>(g) * Cig + cl;
  const float* in_b = inT + img;
  float* gin_b = ginT ? ginT + img : nullptr;
  const float* gc_t = gcol + static_cast<size_t>(tile) * p.nchunk * 8192;
  for (int ch = 0; ch < p.nchunk; ++ch) {
    const int cc = ch / p.ntg, tg = ch - cc * p.ntg;
    const int tr = tg * tpc + ti;
    const bool tlive = tr < T;                                         // padding taps of the last tap group contribute nothing
    const int t = min(tr, T - 1);
    const int dgi = (g * Cig + cc * 64 + cl) / cpdg;                   // uniform over the lanes of a tap (cpdg % CB == 0)
#pragma unroll 2
    for (int i2 = 0; i2 < 4; ++i2) {
      const int row = 8 * gw + 2 * i2 + hw;
      const int di = (dgi * T + t) * 128 + row;
      const ScatterDesc d = ds[di];
      float a_h = 0.f, a_w = 0.f, a_m = 0.f;
      if (tlive && d.flags != 0) {                                     // sample outside: every gradient is zero
        const float4 gc = __ldg(reinterpret_cast<const float4*>(gc_t + (static_cast<size_t>(ch) * 128 + row) * 64 + 4 * l16));
        const int o00 = (d.base & ~3) + cc * 64;
        const int dxo = (d.base & 1) ? p.Cin : 0, dyo = (d.base & 2) ? rowstep : 0;
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        const float4 v1 = (d.flags & 1) ? __ldg(reinterpret_cast<const float4*>(in_b + o00)) : z;
        const float4 v2 = (d.flags & 2) ? __ldg(reinterpret_cast<const float4*>(in_b + o00 + dxo)) : z;
        const float4 v3 = (d.flags & 4) ? __ldg(reinterpret_cast<const float4*>(in_b + o00 + dyo)) : z;
        const float4 v4 = (d.flags & 8) ? __ldg(reinterpret_cast<const float4*>(in_b + o00 + dyo + dxo)) : z;
        const float hh = 1.f - d.lh, hw2 = 1.f - d.lw;
        const float w1 = hh * hw2, w2 = hh * d.lw, w3 = d.lh * hw2, w4 = d.lh * d.lw;
        const float g4[4] = {gc.x, gc.y, gc.z, gc.w};
        const float a1[4] = {v1.x, v1.y, v1.z, v1.w}, a2[4] = {v2.x, v2.y, v2.z, v2.w};
        const float a3[4] = {v3.x, v3.y, v3.z, v3.w}, a4[4] = {v4.x, v4.y, v4.z, v4.w};
        float gm[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          a_m = fmaf(g4[e], w1 * a1[e] + w2 * a2[e] + w3 * a3[e] + w4 * a4[e], a_m);
          gm[e] = g4[e] * d.m;
          a_h = fmaf(gm[e], -hw2 * a1[e] - d.lw * a2[e] + hw2 * a3[e] + d.lw * a4[e], a_h);   // .cuh:101-110 (bp_dir 0)
          a_w = fmaf(gm[e], -hh * a1[e] + hh * a2[e] - d.lh * a3[e] + d.lh * a4[e], a_w);     // .cuh:112-121 (bp_dir 1)
        }
        if (gin_b) {
          if (d.flags & 1) red_add_v4(gin_b + o00, w1 * gm[0], w1 * gm[1], w1 * gm[2], w1 * gm[3]);
          if (d.flags & 2) red_add_v4(gin_b + o00 + dxo, w2 * gm[0], w2 * gm[1], w2 * gm[2], w2 * gm[3]);
          if (d.flags & 4) red_add_v4(gin_b + o00 + dyo, w3 * gm[0], w3 * gm[1], w3 * gm[2], w3 * gm[3]);
          if (d.flags & 8) red_add_v4(gin_b + o00 + dyo + dxo, w4 * gm[0], w4 * gm[1], w4 * gm[2], w4 * gm[3]);
        }
      }
      // sum over the lanes that share (pixel, tap): lpt consecutive lanes (all of them execute the shuffles)
      for (int sft = 1; sft < lpt; sft <<= 1) {
        a_h += __shfl_xor_sync(0xffffffffu, a_h, sft);
        a_w += __shfl_xor_sync(0xffffffffu, a_w, sft);
        a_m += __shfl_xor_sync(0xffffffffu, a_m, sft);
      }
      if (tlive && (l16 % lpt) == 0) {
        acc[di] += a_h; acc[nacc + di] += a_w; acc[2 * nacc + di] += a_m;
      }
    }
    __syncwarp();                                                      // lanes of different deformable groups share acc rows only per (dgi, t)
  }
  __syncwarp();
  for (int e = lane; e < p.dg * T * 8; e += 32) {
    const int ex = e & 7, r = e >> 3;
    const int t = r % T, dgi = r / T;
    const int xo = xo0 + ex;
    if (yo < p.Ho && xo < p.Wo) {
      const int pix = yo * p.Wo + xo;
      const int di = (dgi * T + t) * 128 + 8 * gw + ex;
      // a convolution group only sees its own channels: with G > 1 several tiles (one per group) add into the same entries
      if (goffset) {
        const size_t ob = ((static_cast<size_t>(b) * p.dg + dgi) * 2 * T) * M + pix;
        if (p.G == 1) {
          goffset[ob + static_cast<size_t>(2 * t) * M] = acc[di];
          goffset[ob + static_cast<size_t>(2 * t + 1) * M] = acc[nacc + di];
        } else {
          atomicAdd(goffset + ob + static_cast<size_t>(2 * t) * M, acc[di]);
          atomicAdd(goffset + ob + static_cast<size_t>(2 * t + 1) * M, acc[nacc + di]);
        }
      }
      if (gmask) {
        float* dst = gmask + ((static_cast<size_t>(b) * p.dg + dgi) * T + t) * M + pix;
        if (p.G == 1) *dst = acc[2 * nacc + di];
        else atomicAdd(dst, acc[2 * nacc + di]);
      }
    }
  }
}

static size_t up256(size_t x) { return (x + 255) & ~static_cast<size_t>(255); }

struct BwdDataLayout { size_t inT, ginT, wt, gcol, total; int Cogp; };

static bool bwd_data_layout(DcnTcParams& p, BwdDataLayout& L) {
  size_t smem = 0;
  if (!dcn_tc2_plan(p, &smem, true)) return false;
  const int Cog = p.Cout / p.G;
  if (Cog > 256 || p.Wo % 4 != 0) return false;
  if ((p.Cin / p.dg) % p.CB != 0) return false;                 // the offset / mask sums run over the lanes of one tap
  if (1LL * p.H * p.W > 65535LL * 32) return false;             // grid of the transpose back to NCHW
  if (static_cast<size_t>(p.dg) * p.kh * p.kw * 128 * (sizeof(ScatterDesc) + 12) > 200 * 1024) return false;
  L.Cogp = ceil_div(Cog, 128) * 128;
  const size_t act = static_cast<size_t>(p.B) * p.H * p.W * p.Cin * sizeof(float);
  L.inT = 0;
  L.ginT = up256(act);
  L.wt = L.ginT + up256(act);
  L.gcol = L.wt + up256(static_cast<size_t>(2) * p.G * p.K * L.Cogp * sizeof(__nv_bfloat16));
  L.total = L.gcol + static_cast<size_t>(p.B) * p.G * p.nptile * p.nchunk * 8192 * sizeof(float);
  return true;
}

size_t dcn_bwd_data_tc_workspace_bytes(int B, int Cin, int H, int W, int Cout, int kh, int kw, int stride_h, int stride_w,
                                       int pad_h, int pad_w, int dil_h, int dil_w, int group, int deformable_group) {
  DcnTcParams p{};
  p.B = B; p.Cin = Cin; p.H = H; p.W = W; p.Cout = Cout; p.kh = kh; p.kw = kw; p.G = group; p.dg = deformable_group;
  if (B <= 0 || H <= 0 || W <= 0 || stride_h <= 0 || stride_w <= 0) return 0;
  p.Ho = (H + 2 * pad_h - (dil_h * (kh - 1) + 1)) / stride_h + 1;
  p.Wo = (W + 2 * pad_w - (dil_w * (kw - 1) + 1)) / stride_w + 1;
  if (p.Ho <= 0 || p.Wo <= 0) return 0;
  BwdDataLayout L;
  return bwd_data_layout(p, L) ? L.total : 0;
}

// grad_input (may be null), grad_offset, grad_mask through the two kernels above; *handled = false when the shape does not
// fit or the workspace is too small.  grad_input is written completely (no zero-fill needed by the caller).
int launch_dcn_bwd_data_tc(DcnTcParams p, const float* weight, const float* grad_out, float* grad_input, float* grad_offset,
                           float* grad_mask, void* workspace, size_t workspace_bytes, cudaStream_t stream, bool* handled) {
  // (the channels-last input copy lands at the start of the workspace, where the weight-gradient kernel expects it)
  *handled = false;
  BwdDataLayout L;
  if (!bwd_data_layout(p, L)) return SF_OK;
  if (workspace == nullptr || workspace_bytes < L.total || (reinterpret_cast<uintptr_t>(workspace) & 255) != 0) return SF_OK;
  if ((reinterpret_cast<uintptr_t>(grad_out) & 15) != 0) return SF_OK;
  unsigned char* ws = static_cast<unsigned char*>(workspace);
  float* inT = reinterpret_cast<float*>(ws + L.inT);
  float* ginT = grad_input ? reinterpret_cast<float*>(ws + L.ginT) : nullptr;
  __nv_bfloat16* wt = reinterpret_cast<__nv_bfloat16*>(ws + L.wt);
  float* gcol = reinterpret_cast<float*>(ws + L.gcol);
  const int T = p.kh * p.kw, Cog = p.Cout / p.G, HW = p.H * p.W;
  int rc = launch_dcn_nhwc(p.in, inT, p.B, p.Cin, HW, stream);
  if (rc) return rc;
  if (ginT) SF_CUDA(cudaMemsetAsync(ginT, 0, static_cast<size_t>(p.B) * HW * p.Cin * sizeof(float), stream));
  if (p.G > 1) {
    const size_t M = static_cast<size_t>(p.Ho) * p.Wo;
    if (grad_offset) SF_CUDA(cudaMemsetAsync(grad_offset, 0, sizeof(float) * p.B * p.dg * 2 * T * M, stream));
    if (grad_mask) SF_CUDA(cudaMemsetAsync(grad_mask, 0, sizeof(float) * p.B * p.dg * T * M, stream));
  }
  {
    const size_t n = static_cast<size_t>(p.G) * p.K * L.Cogp;
    const int grid = static_cast<int>(min(static_cast<size_t>(sm_count()) * 8, (n + 255) / 256));
    dcn_split_weights_t_kernel<<<grid, 256, 0, stream>>>(weight, wt, p.G, Cog, L.Cogp, p.K, p.Cin / p.G, T, p.CB, p.ntg);
    SF_LAUNCH_CHECK();
  }
  DcnBwdDataParams q;
  q.f = p;
  q.gcol = gcol;
  q.nmt = L.Cogp / 128;
  q.Kpad = p.K;
  q.ntile = p.B * p.G * p.nptile;
  CUtensorMap mG, mW;
  rc = make_tmap_3d_f32(&mG, grad_out, p.Wo, p.Ho, static_cast<uint64_t>(p.B) * p.Cout, 16, 8, 32);
  if (rc) return rc;
  rc = make_tmap_3d_bf16_sw128(&mW, wt, L.Cogp, static_cast<uint64_t>(p.G) * p.K, 2, 64, 1);
  if (rc) return rc;
  {
    const size_t smem = 2 * DG_GB + 2 * DG_B + 2 * DG_GF + 256;
    SF_CUDA(cudaFuncSetAttribute(dcn_bwd_gcol_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    dcn_bwd_gcol_tc_kernel<<<min(q.ntile, sm_count()), DG_THREADS, smem, stream>>>(mG, mW, q);
    SF_LAUNCH_CHECK();
  }
  {
    const size_t smem = static_cast<size_t>(p.dg) * T * 128 * (sizeof(ScatterDesc) + 12);
    SF_CUDA(cudaFuncSetAttribute(dcn_bwd_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    dcn_bwd_scatter_kernel<<<q.ntile, 512, smem, stream>>>(p, inT, gcol, ginT, grad_offset, grad_mask);
    SF_LAUNCH_CHECK();
  }
  if (grad_input) {
    // channels-last -> NCHW with the same tiled transpose: [HW][Cin] -> [Cin][HW] per image
    rc = launch_dcn_nhwc(ginT, grad_input, p.B, HW, p.Cin, stream);
    if (rc) return rc;
  }
  *handled = true;
  return SF_OK;
}

}  // namespace sf

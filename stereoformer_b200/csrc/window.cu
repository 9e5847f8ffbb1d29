// Soft-argmin regression and local disparity-window sampling  (SURVEY 8a rows a3, a4, a5, a6-sampler).
//
// Reference: models/utils/estimation.py:14-62; models/LocalCostVolume/Attempts/fixed_local_cost_volume.py:18-80;
// variance_local_cost_volume.py:12-78; Attempts/update.py:279-345 (DLC window tail).
//
// One thread owns one pixel.  The pixel's D-vector of the cost volume is staged as a [D][128] tile in
// shared memory by one TMA box load per CTA (coalesced fallback copy when W % 4 != 0), so the data-dependent
// floor/ceil gathers are bank-conflict-free shared-memory reads (bank = lane) instead of global gathers.
// Candidates / indices / interpolation weights follow pixel_math.h and are bit-identical to the reference;
// the soft-max over the N+1 samples is a max-then-exponentials soft-max in fp32 (samples kept in registers).
// Algorithmic bytes per pixel: (D + 2) * 4 (+8 with per-pixel bounds).
#include <cstdlib>
#include "common.cuh"
#include "pixel_math.h"
#include "../../include/stereoformer_b200.h"

namespace sf {

constexpr int WIN_TPB = 128;

struct WindowParams {
  const float* cv;     // [B,D,H,W]
  const float* disp;   // [B,1,H,W] or null -> soft-argmin of cv
  const float* lo;     // [B,1,H,W] (modes OFFSETS / ABSOLUTE)
  const float* up;
  float* out;          // [B,1,H,W]
  float* disp0_out;    // optional: the current disparity used (soft-argmin when disp == null)
  float* dbg_cand;     // optional [B,N+1,H,W]
  int* dbg_idx_floor;  // optional [B,N+1,H,W]
  int* dbg_idx_ceil;
  float* dbg_local;    // optional [B,N+1,H,W]
  int B, D, H, W;
  int mode, n_points, consider_valid, recip;
  float radius, gamma;
  int nseg;
  int use_tma;
  // multi-GPU gather fused into the store: the result also goes to every rank's gather buffer through its
  // NVLink peer mapping (peers[r] + peer_off + pixel index); n_peers == 0 outside sf_local_window_gather_f32
  float* peers[SF_MAX_PEERS];
  int n_peers;
  long long peer_off;
  // completion protocol of sf_local_window_gather_signal_f32 (null sig_local: plain stores, no signalling)
  unsigned long long* sig_local;
  unsigned long long* ll_peers[SF_MAX_PEERS];   // destinations' {value, flag} slots
  int peer_ranks[SF_MAX_PEERS];
  long long slot_stride;
};

struct SoftStats { float m, s, e; };  // running max, sum exp, sum exp * d

// Two passes over the staged column (max, then exponentials): the exponentials are independent, so the MUFU
// pipe stays busy instead of waiting on the running-max rescale chain of an online soft-max.
template <class F>
__device__ __forceinline__ SoftStats softmax_stats(F cvat, int D) {
  SoftStats st; st.m = -INFINITY; st.s = 0.f; st.e = 0.f;
  for (int d = 0; d < D; ++d) st.m = fmaxf(st.m, cvat(d));
  for (int d = 0; d < D; ++d) {
    const float ex = __expf(cvat(d) - st.m);
    st.s += ex; st.e = fmaf(ex, (float)d, st.e);
  }
  return st;
}

constexpr int WIN_MAXN = 32;   // samples kept in registers between the two soft-max passes (more: recomputed)

// NSAMP: number of window samples (n_points + 1) when it is one of the shipped values (21: LowCNN, 11: module
// defaults), so the sample loop unrolls exactly; 0 = any count (predicated up to WIN_MAXN, recomputed beyond).
template <bool DEBUG, int NSAMP, int TPB>
__global__ void __launch_bounds__(TPB)
local_window_kernel(const __grid_constant__ CUtensorMap mapCV, const WindowParams p) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* tile = reinterpret_cast<float*>(smem_raw);           // [D][128]
  __shared__ __align__(8) uint64_t bar;

  const int seg = blockIdx.x % p.nseg, t = blockIdx.x / p.nseg;
  const int y = t % p.H, b = t / p.H;
  const int x0 = seg * TPB, x = x0 + threadIdx.x;
  const size_t HW = static_cast<size_t>(p.H) * p.W;

  if (p.use_tma && threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
    mbar_expect_tx(&bar, p.D * TPB * 4);
    tma_load_3d(tile, &mapCV, &bar, x0, y, b * p.D);
  }
  // signalled gather: a CTA's ticket gives the step of this launch (all launches of one gather have the same grid and
  // run one after the other); the step's slot was last used by step - SF_GATHER_SLOTS, which every destination must have
  // consumed (normally satisfied long before: SF_GATHER_SLOTS steps of slack between producer and consumer)
  __shared__ unsigned long long s_step;
  if (p.sig_local && threadIdx.x == 0) {
    // 16 ticket counters (CTAs with the same blockIdx % 16 share one): ~2000 atomics on ONE word serialise at L2 and
    // every CTA waits for its return value before it may store
    const unsigned lane16 = blockIdx.x & 15u;
    const unsigned n_mine = (gridDim.x - lane16 + 15u) >> 4;
    const unsigned long long step = atomicAdd(p.sig_local + SF_SIG_TICKET16 + lane16, 1ULL) / n_mine;
    if (step >= SF_GATHER_SLOTS)
      for (int r = 0; r < p.n_peers; ++r)
        spin_until_ge_sys(p.sig_local + SF_SIG_CONSUMED + p.peer_ranks[r], step - (SF_GATHER_SLOTS - 1),
                          p.sig_local + SF_SIG_ERROR);
    s_step = step;
  }
  if (p.use_tma) {
    __syncthreads();
    mbar_wait(&bar, 0);
  } else {
    const float* src = p.cv + (static_cast<size_t>(b) * p.D * p.H + y) * p.W;
    for (int d = 0; d < p.D; ++d) tile[d * TPB + threadIdx.x] = (x < p.W) ? src[d * HW + x] : 0.f;
    __syncthreads();
  }
  if (x < p.W) {

  const float* col = tile + threadIdx.x;
  auto cvat = [&](int d) { return col[d * TPB]; };
  const size_t pix = (static_cast<size_t>(b) * p.H + y) * p.W + x;
  const int D = p.D, N = p.n_points;

  // ---- current disparity
  float cur;
  SoftStats st;
  const bool need_stats = (p.disp == nullptr) || p.mode == SF_WIN_VARIANCE;
  if (need_stats) st = softmax_stats(cvat, D);
  if (p.disp) cur = p.disp[pix];
  else cur = st.e / st.s;
  if (p.disp0_out) p.disp0_out[pix] = cur;

  // ---- bounds
  float lower, upper, extra = 0.f;
  if (p.mode == SF_WIN_FIXED) {
    lower = sfm::fsub(cur, p.radius);
    upper = sfm::fadd(cur, p.radius);
  } else if (p.mode == SF_WIN_VARIANCE) {
    // GetVarince, variance_local_cost_volume.py:12-25: sigma = sqrt(sum_d p_d * (d-cur)*(d-cur))
    float acc = 0.f;
    for (int d = 0; d < D; ++d) {
      float e = sfm::fsub((float)d, cur);
      acc = fmaf(__expf(cvat(d) - st.m), sfm::fmul(e, e), acc);
    }
    float sigma = sfm::fsqrt(acc / st.s);
    float gs = sfm::fmul(p.gamma, sigma);
    lower = sfm::fsub(cur, gs);
    upper = sfm::fadd(cur, gs);
    if (p.consider_valid) extra = (upper > (float)x) ? 1.f : 0.f;
  } else if (p.mode == SF_WIN_OFFSETS) {
    lower = sfm::fsub(cur, p.lo[pix]);
    upper = sfm::fadd(cur, p.up[pix]);
  } else if (p.mode == SF_WIN_HEAD) {
    // lo = the OffsetPredictionHead's last convolution output [B,2,H,W] BEFORE relu2 (update.py:96-105, 275-282):
    // the ReLU and the channel chunk are folded into the load
    const size_t hp = pix + static_cast<size_t>(b) * p.H * p.W;      // b * 2HW + y * W + x
    lower = sfm::fsub(cur, fmaxf(p.lo[hp], 0.f));
    upper = sfm::fadd(cur, fmaxf(p.lo[hp + static_cast<size_t>(p.H) * p.W], 0.f));
  } else {
    lower = p.lo[pix];
    upper = p.up[pix];
    // SF_WIN_VARIANCE_BOUNDS: the caller's bounds under the variance window's rules (variance_local_cost_volume.py:44-53)
    if (p.mode == SF_WIN_VARIANCE_BOUNDS && p.consider_valid) extra = (upper > (float)x) ? 1.f : 0.f;
  }
  const bool var_rules = p.mode == SF_WIN_VARIANCE || p.mode == SF_WIN_VARIANCE_BOUNDS;
  const sfm::Window w = sfm::make_window(lower, upper, D, N, p.consider_valid, var_rules, extra, p.recip);

  // ---- N+1 samples; soft-max over them (max first, then independent exponentials); expectation of the candidates
  const size_t dbg0 = (static_cast<size_t>(b) * (N + 1) * p.H + y) * p.W + x;
  auto sample = [&](int k, float& cand) {
    cand = sfm::window_candidate(w, k, cur, p.consider_valid);
    int jf, ic;
    sfm::window_index(cand, D, &jf, &ic);
    const float v = sfm::window_lerp(cvat(jf), cvat(ic), ic, cand);
    if (DEBUG) {
      if (p.dbg_cand) p.dbg_cand[dbg0 + k * HW] = cand;
      if (p.dbg_idx_floor) p.dbg_idx_floor[dbg0 + k * HW] = jf;
      if (p.dbg_idx_ceil) p.dbg_idx_ceil[dbg0 + k * HW] = ic;
      if (p.dbg_local) p.dbg_local[dbg0 + k * HW] = v;
    }
    return v;
  };
  float m = -INFINITY, s = 0.f, e = 0.f;
  if (NSAMP > 0) {
    float vs[NSAMP > 0 ? NSAMP : 1], cs[NSAMP > 0 ? NSAMP : 1];
#pragma unroll
    for (int k = 0; k < NSAMP; ++k) { vs[k] = sample(k, cs[k]); m = fmaxf(m, vs[k]); }
#pragma unroll
    for (int k = 0; k < NSAMP; ++k) { const float ex = __expf(vs[k] - m); s += ex; e = fmaf(ex, cs[k], e); }
  } else if (N < WIN_MAXN) {
    float vs[WIN_MAXN], cs[WIN_MAXN];
#pragma unroll
    for (int k = 0; k < WIN_MAXN; ++k)
      if (k <= N) { vs[k] = sample(k, cs[k]); m = fmaxf(m, vs[k]); }
#pragma unroll
    for (int k = 0; k < WIN_MAXN; ++k)
      if (k <= N) { const float ex = __expf(vs[k] - m); s += ex; e = fmaf(ex, cs[k], e); }
  } else {
    float cand;
    for (int k = 0; k <= N; ++k) m = fmaxf(m, sample(k, cand));
    for (int k = 0; k <= N; ++k) { const float v = sample(k, cand); const float ex = __expf(v - m); s += ex; e = fmaf(ex, cand, e); }
  }
  const float res = e / s;
  p.out[pix] = res;
  if (p.sig_local) {
    const unsigned long long step = s_step;       // written before the __syncthreads that follows the tile load
    const long long at = static_cast<long long>(step % SF_GATHER_SLOTS) * p.slot_stride + p.peer_off + static_cast<long long>(pix);
    for (int r = 0; r < p.n_peers; ++r) st_ll_sys(p.ll_peers[r] + at, res, static_cast<uint32_t>(step + 1));
  } else {
    for (int r = 0; r < p.n_peers; ++r) p.peers[r][p.peer_off + static_cast<long long>(pix)] = res;
  }
  }  // x < W
}

// DisparityEstimation(.WithProb).forward  models/utils/estimation.py:14-62: no reuse -> straight coalesced loads.
__global__ void __launch_bounds__(256)
soft_argmin_kernel(const float* __restrict__ cv, float* __restrict__ disp, float* __restrict__ prob, int B, int D,
                   int H, int W, float sign) {
  const size_t HW = static_cast<size_t>(H) * W;
  const size_t n = static_cast<size_t>(B) * HW;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t b = i / HW, r = i % HW;
    const float* src = cv + b * D * HW + r;
    float m = -INFINITY, s = 0.f, e = 0.f;
    int d = 0;
    for (; d + 4 <= D; d += 4) {
      float v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = sign * __ldg(src + (d + j) * HW);
      float mx = fmaxf(fmaxf(v[0], v[1]), fmaxf(v[2], v[3]));
      if (mx > m) {
        float sc = __expf(m - mx);
        s *= sc; e *= sc; m = mx;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float ex = __expf(v[j] - m);
        s += ex; e = fmaf(ex, (float)(d + j), e);
      }
    }
    for (; d < D; ++d) {
      float v = sign * __ldg(src + d * HW);
      if (v > m) {
        float sc = __expf(m - v);
        s *= sc; e *= sc; m = v;
      }
      float ex = __expf(v - m);
      s += ex; e = fmaf(ex, (float)d, e);
    }
    disp[i] = e / s;
    if (prob) {
      float inv = 1.0f / s;
      float* dst = prob + b * D * HW + r;
      for (int dd = 0; dd < D; ++dd) dst[dd * HW] = __expf(sign * __ldg(src + dd * HW) - m) * inv;
    }
  }
}

// build_cost_volume_from_volume(old_volume, sampling_candidates)  fixed_local_cost_volume.py:61-80: arbitrary
// candidates [B,K,H,W]; output bit-exact (mul, mul, add as separate fp32 ops).
__global__ void __launch_bounds__(256)
sample_volume_kernel(const float* __restrict__ cv, const float* __restrict__ cand, float* __restrict__ out,
                     int* __restrict__ idx_floor, int* __restrict__ idx_ceil, int B, int D, int K, int H, int W) {
  const size_t HW = static_cast<size_t>(H) * W;
  const size_t n = static_cast<size_t>(B) * K * HW;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t r = i % HW, b = i / (K * HW);
    const float c = cand[i];
    int jf, ic;
    sfm::window_index(c, D, &jf, &ic);
    const float* src = cv + b * D * HW + r;
    out[i] = sfm::window_lerp(__ldg(src + jf * HW), __ldg(src + ic * HW), ic, c);
    if (idx_floor) idx_floor[i] = jf;
    if (idx_ceil) idx_ceil[i] = ic;
  }
}

}  // namespace sf

extern "C" int sf_soft_argmin_f32(const float* cv, float* disp, float* prob, int B, int D, int H, int W,
                                  int match_similarity, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(cv && disp, "soft_argmin: null pointer");
  SF_CHECK_ARG(B > 0 && D > 0 && H > 0 && W > 0, "soft_argmin: non-positive dimension");
  size_t n = static_cast<size_t>(B) * H * W;
  int grid = static_cast<int>(min(static_cast<size_t>(sm_count()) * 8, (n + 255) / 256));
  soft_argmin_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(cv, disp, prob, B, D, H, W,
                                                                        match_similarity ? 1.f : -1.f);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

static int local_window_impl(const float* cv, const float* disp, const float* lo, const float* up, float* out,
                             float* disp0_out, float* dbg_cand, int* dbg_idx_floor, int* dbg_idx_ceil,
                             float* dbg_local, int B, int D, int H, int W, int mode, float radius, float gamma,
                             int n_points, int consider_valid, int div_mode, float* const* h_peers, int n_peers,
                             long long peer_offset, void* stream, const int* h_peer_ranks = nullptr,
                             long long slot_stride = 0, unsigned long long* sig_local = nullptr) {
  using namespace sf;
  SF_CHECK_ARG(cv && out, "local_window: null pointer");
  SF_CHECK_ARG(B > 0 && D > 0 && H > 0 && W > 0 && n_points > 0, "local_window: non-positive dimension");
  SF_CHECK_ARG(mode >= SF_WIN_FIXED && mode <= SF_WIN_VARIANCE_BOUNDS, "local_window: unknown mode %d", mode);
  SF_CHECK_ARG((mode != SF_WIN_OFFSETS && mode != SF_WIN_ABSOLUTE && mode != SF_WIN_VARIANCE_BOUNDS) || (lo && up),
               "local_window: this mode needs the lo/up bound tensors");
  SF_CHECK_ARG(mode != SF_WIN_HEAD || lo, "local_window: SF_WIN_HEAD needs the [B,2,H,W] head output in lo");
  SF_CHECK_ARG(static_cast<size_t>(D) * WIN_TPB * 4 <= 200 * 1024, "local_window: D=%d too large for the staged tile", D);
  WindowParams p;
  p.cv = cv; p.disp = disp; p.lo = lo; p.up = up; p.out = out; p.disp0_out = disp0_out;
  p.dbg_cand = dbg_cand; p.dbg_idx_floor = dbg_idx_floor; p.dbg_idx_ceil = dbg_idx_ceil; p.dbg_local = dbg_local;
  p.B = B; p.D = D; p.H = H; p.W = W; p.mode = mode; p.n_points = n_points;
  p.consider_valid = consider_valid ? 1 : 0; p.recip = div_mode == SF_DIV_RECIP;
  p.radius = radius; p.gamma = gamma;
  SF_CHECK_ARG(n_peers >= 0 && n_peers <= SF_MAX_PEERS && (n_peers == 0 || h_peers) && peer_offset >= 0,
               "local_window: bad peer list (at most %d peers)", SF_MAX_PEERS);
  p.n_peers = n_peers;
  p.peer_off = peer_offset;
  for (int r = 0; r < SF_MAX_PEERS; ++r) p.peers[r] = r < n_peers ? h_peers[r] : nullptr;
  for (int r = 0; r < n_peers; ++r) SF_CHECK_ARG(p.peers[r] != nullptr, "local_window: null peer pointer %d", r);
  p.sig_local = sig_local; p.slot_stride = slot_stride;
  for (int r = 0; r < SF_MAX_PEERS; ++r) { p.ll_peers[r] = nullptr; p.peer_ranks[r] = 0; }
  if (sig_local) {
    SF_CHECK_ARG(n_peers > 0 && h_peer_ranks && slot_stride > 0,
                 "local_window: the signalled gather needs destinations, their ranks and a slot stride");
    for (int r = 0; r < n_peers; ++r) {
      SF_CHECK_ARG(h_peer_ranks[r] >= 0 && h_peer_ranks[r] < 16, "local_window: bad rank of destination %d", r);
      SF_CHECK_ARG((reinterpret_cast<uintptr_t>(h_peers[r]) & 7) == 0, "local_window: slot %d not 8-byte aligned", r);
      p.ll_peers[r] = reinterpret_cast<unsigned long long*>(h_peers[r]); p.peer_ranks[r] = h_peer_ranks[r];
    }
  }
  // CTA width: 128 pixels, or 32 when the row is short enough that single-warp CTAs waste fewer lanes; single-warp CTAs
  // (3 KB of shared memory at D = 24) also fit next to a resident persistent kernel, so a caller that pipelines the
  // window under the next batch's cost volume (bench.py) gets real overlap
  static const int env_tpb = []() { const char* e = tuning_env("SF_WIN_TPB"); return e ? atoi(e) : 0; }();
  const bool debug = dbg_cand || dbg_idx_floor || dbg_idx_ceil || dbg_local;
  // (the debug instantiation is 128 wide: the sweep hook never applies to it)
  const int tpb = debug ? 128 : (env_tpb == 32 || env_tpb == 128) ? env_tpb : (W <= 256 ? 32 : 128);
  p.nseg = ceil_div(W, tpb);
  p.use_tma = (W % 4 == 0) && D <= 256 && (reinterpret_cast<uintptr_t>(cv) & 15) == 0;
  CUtensorMap map;
  memset(&map, 0, sizeof(map));
  if (p.use_tma) {
    int rc = make_tmap_3d_f32(&map, cv, W, H, static_cast<uint64_t>(B) * D, tpb, 1, D);
    if (rc) return rc;
  }
  const size_t smem = static_cast<size_t>(D) * tpb * 4;
  using KernT = void (*)(const CUtensorMap, const WindowParams);
  KernT kern;
  if (debug) kern = local_window_kernel<true, 0, 128>;
  else if (tpb == 128) kern = n_points == 20 ? local_window_kernel<false, 21, 128>
                                             : (n_points == 10 ? local_window_kernel<false, 11, 128> : local_window_kernel<false, 0, 128>);
  else kern = n_points == 20 ? local_window_kernel<false, 21, 32>
                             : (n_points == 10 ? local_window_kernel<false, 11, 32> : local_window_kernel<false, 0, 32>);
  SF_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  long long grid = 1LL * B * H * p.nseg;
  SF_CHECK_ARG(grid < (1LL << 31), "local_window: grid too large");
  kern<<<static_cast<int>(grid), tpb, smem, static_cast<cudaStream_t>(stream)>>>(map, p);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

extern "C" int sf_local_window_f32(const float* cv, const float* disp, const float* lo, const float* up, float* out,
                                   float* disp0_out, float* dbg_cand, int* dbg_idx_floor, int* dbg_idx_ceil,
                                   float* dbg_local, int B, int D, int H, int W, int mode, float radius, float gamma,
                                   int n_points, int consider_valid, int div_mode, void* stream) {
  return local_window_impl(cv, disp, lo, up, out, disp0_out, dbg_cand, dbg_idx_floor, dbg_idx_ceil, dbg_local, B, D, H, W,
                           mode, radius, gamma, n_points, consider_valid, div_mode, nullptr, 0, 0, stream);
}

extern "C" int sf_local_window_gather_f32(const float* cv, const float* disp, const float* lo, const float* up, float* out,
                                          int B, int D, int H, int W, int mode, float radius, float gamma, int n_points,
                                          int consider_valid, int div_mode, float* const* h_peers, int n_peers,
                                          long long peer_offset, void* stream) {
  return local_window_impl(cv, disp, lo, up, out, nullptr, nullptr, nullptr, nullptr, nullptr, B, D, H, W, mode, radius,
                           gamma, n_points, consider_valid, div_mode, h_peers, n_peers, peer_offset, stream);
}

extern "C" int sf_local_window_gather_signal_f32(const float* cv, const float* disp, const float* lo, const float* up,
                                                 float* out, int B, int D, int H, int W, int mode, float radius,
                                                 float gamma, int n_points, int consider_valid, int div_mode,
                                                 unsigned long long* const* h_peer_data,
                                                 unsigned long long* const* h_peer_sig, const int* h_peer_ranks,
                                                 int n_peers, long long peer_offset, long long slot_stride,
                                                 unsigned long long* sig_local, void* stream) {
  SF_CHECK_ARG(sig_local != nullptr, "local_window_gather_signal: null signal block");
  (void)h_peer_sig;   // the producers only read their own block; kept in the signature for symmetry with the wait
  return local_window_impl(cv, disp, lo, up, out, nullptr, nullptr, nullptr, nullptr, nullptr, B, D, H, W, mode, radius,
                           gamma, n_points, consider_valid, div_mode, reinterpret_cast<float* const*>(h_peer_data),
                           n_peers, peer_offset, stream, h_peer_ranks, slot_stride, sig_local);
}

extern "C" int sf_sample_volume_f32(const float* cv, const float* cand, float* out, int* idx_floor, int* idx_ceil,
                                    int B, int D, int K, int H, int W, void* stream) {
  using namespace sf;
  SF_CHECK_ARG(cv && cand && out, "sample_volume: null pointer");
  SF_CHECK_ARG(B > 0 && D > 0 && K > 0 && H > 0 && W > 0, "sample_volume: non-positive dimension");
  size_t n = static_cast<size_t>(B) * K * H * W;
  int grid = static_cast<int>(min(static_cast<size_t>(sm_count()) * 8, (n + 255) / 256));
  sample_volume_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(cv, cand, out, idx_floor, idx_ceil, B, D,
                                                                          K, H, W);
  SF_LAUNCH_CHECK();
  return SF_OK;
}

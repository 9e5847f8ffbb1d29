// Dev microbenchmark: how fast can one B200 stream the cost-volume operand rows into shared memory?
// Feature maps [NCH][H][W] fp32 (NCH = 8 pairs x 256 channels), tile = RT epipolar rows x KC channels per stage.
//   mode 0: 3-D tensor-map boxes  L [W][RT][KC], R [W+HL][RT][KC] at x = -HL (what the kernels do today)
//   mode 1: 1-D bulk copies (cp.async.bulk), one per channel: RT*W floats (L) and RT*W+HL floats (R, flat halo)
// A consumer warp releases every stage as soon as it lands (no math): this is the load-path ceiling.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_tma tools/ubench_tma.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(c) : "memory"); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(b)) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t par) {
  uint32_t ok = 0;
  while (!ok) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(s32(b)), "r"(par) : "memory");
}
__device__ __forceinline__ void tma3d(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
               ::"r"(s32(dst)), "l"((uint64_t)m), "r"(s32(bar)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void bulk1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(s32(dst)), "l"(src), "r"(bytes), "r"(s32(bar)) : "memory");
}

struct P { const float* L; const float* R; int NCH, H, W, HL, RT, KC, NS, mode, ntiles, nch, CPP; float* sink; };

__global__ void __launch_bounds__(64, 1) k(const __grid_constant__ CUtensorMap mL, const __grid_constant__ CUtensorMap mR, const P p) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const int rowL = p.RT * p.W, rowR = p.mode == 0 ? p.RT * (p.W + p.HL) : p.RT * p.W + p.HL;
  const int stage_bytes = ((p.KC * (rowL + rowR) * 4 + 1023) / 1024) * 1024;
  uint64_t* full = (uint64_t*)(smem + p.NS * stage_bytes);
  uint64_t* empty = full + 16;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.NS; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int rows_per_img = (p.H + p.RT - 1) / p.RT;
  if (warp == 0) {
    int it = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
      const int yb = tile % rows_per_img, img = tile / rows_per_img;
      const int y = yb * p.RT;
      for (int kc = 0; kc < p.nch; ++kc, ++it) {
        const int s = it % p.NS;
        mbar_wait(&empty[s], ((it / p.NS) & 1) ^ 1);
        unsigned char* dst = smem + s * stage_bytes;
        const int c0 = img * p.CPP + kc * p.KC;
        if (p.mode == 0) {
          if (lane == 0) {
            mbar_expect(&full[s], p.KC * (rowL + rowR) * 4);
            tma3d(dst, &mL, &full[s], 0, y, c0);
            tma3d(dst + p.KC * rowL * 4, &mR, &full[s], -p.HL, y, c0);
          }
        } else {
          if (lane == 0) mbar_expect(&full[s], p.KC * (rowL + rowR) * 4);
          __syncwarp();
          for (int c = lane; c < p.KC; c += 32) {
            const size_t off = ((size_t)(c0 + c) * p.H + y) * p.W;
            bulk1d(dst + c * rowL * 4, p.L + off, rowL * 4, &full[s]);
            // flat halo: the HL floats before the row start (previous row / previous channel); first row of the buffer skips it
            const float* rs = p.R + off - p.HL;
            bulk1d(dst + (p.KC * rowL + c * rowR) * 4, off >= (size_t)p.HL ? rs : p.R, rowR * 4, &full[s]);
          }
        }
        __syncwarp();
      }
    }
  } else {
    int it = 0;
    float acc = 0.f;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x)
      for (int kc = 0; kc < p.nch; ++kc, ++it) {
        const int s = it % p.NS;
        mbar_wait(&full[s], (it / p.NS) & 1);
        acc += ((const float*)(smem + s * stage_bytes))[lane];
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
      }
    if (acc == 123.456f) p.sink[0] = acc;
  }
}

typedef CUresult (*enc_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char** argv) {
  const int B = 8, C = 256, H = 68, W = 120, HL = 24, NCH = B * C;
  void* fp = nullptr; cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
  enc_fn enc = (enc_fn)fp;
  const size_t n = (size_t)NCH * H * W;
  float* bufs[6];
  for (int i = 0; i < 6; ++i) { CK(cudaMalloc(&bufs[i], n * 4 + 4096)); CK(cudaMemset(bufs[i], 0, n * 4 + 4096)); }
  float* sink; CK(cudaMalloc(&sink, 16));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  struct Cfg { int mode, RT, KC, NS, grid, promo; };
  Cfg cfgs[] = {
      {0, 1, 32, 3, 148, 1}, {0, 1, 32, 6, 148, 1}, {0, 1, 32, 6, 148, 2}, {0, 1, 16, 12, 148, 1}, {0, 1, 64, 3, 148, 1},
      {0, 2, 16, 6, 148, 1}, {0, 2, 32, 3, 148, 1}, {0, 4, 8, 6, 136, 1}, {0, 4, 16, 3, 136, 1},
      {1, 1, 32, 3, 148, 1}, {1, 1, 32, 6, 148, 1}, {1, 2, 16, 6, 148, 1}, {1, 2, 32, 3, 148, 1}, {1, 4, 16, 3, 136, 1},
      {1, 4, 8, 6, 136, 1}, {1, 4, 8, 3, 136, 1}, {1, 2, 8, 12, 148, 1},
  };
  for (const Cfg& c : cfgs) {
    P p; p.NCH = NCH; p.H = H; p.W = W; p.HL = HL; p.RT = c.RT; p.KC = c.KC; p.NS = c.NS; p.mode = c.mode; p.CPP = C;
    p.nch = C / c.KC; p.ntiles = B * ((H + c.RT - 1) / c.RT); p.sink = sink;
    const int rowL = c.RT * W, rowR = c.mode == 0 ? c.RT * (W + HL) : c.RT * W + HL;
    const int stage_bytes = ((c.KC * (rowL + rowR) * 4 + 1023) / 1024) * 1024;
    const size_t smem = (size_t)c.NS * stage_bytes + 512;
    if (smem > 232448) { printf("skip (smem %zu)\n", smem); continue; }
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    float best = 1e9, tot = 0;
    const int reps = 30;
    for (int r = 0; r < reps + 3; ++r) {
      cudaEventRecord(e0);
      for (int u = 0; u < 6; ++u) {
      p.L = bufs[(2 * u) % 6]; p.R = bufs[(2 * u + 1) % 6];
      CUtensorMap mL, mR;
      cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)NCH};
      cuuint64_t str[2] = {(cuuint64_t)W * 4, (cuuint64_t)W * H * 4};
      cuuint32_t bl[3] = {(cuuint32_t)W, (cuuint32_t)c.RT, (cuuint32_t)c.KC}, br[3] = {(cuuint32_t)(W + HL), (cuuint32_t)c.RT, (cuuint32_t)c.KC}, es[3] = {1, 1, 1};
      CUtensorMapL2promotion pr = c.promo == 2 ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B : CU_TENSOR_MAP_L2_PROMOTION_L2_128B;
      enc(&mL, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)p.L, dims, str, bl, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, pr, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      enc(&mR, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)p.R, dims, str, br, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, pr, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      k<<<c.grid, 64, smem>>>(mL, mR, p);
      }
      cudaEventRecord(e1);
      CK(cudaEventSynchronize(e1));
      float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 6;
      if (r >= 3) { best = ms < best ? ms : best; tot += ms; }
    }
    const double bytes = 2.0 * n * 4;
    printf("mode=%d RT=%2d KC=%2d NS=%2d grid=%3d promo=%d stage=%6d B  avg %.2f us (%.0f GB/s)  best %.2f us (%.0f GB/s)\n", c.mode, c.RT, c.KC,
           c.NS, c.grid, c.promo, stage_bytes, tot / reps * 1e3, bytes / (tot / reps * 1e-3) / 1e9, best * 1e3, bytes / (best * 1e-3) / 1e9);
  }
  return 0;
}

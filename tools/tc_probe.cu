// Dev bring-up probe for the tcgen05 path (not part of the library): validates on a B200
//   (1) MN-major bf16 A and B operands loaded by TMA with 128-byte swizzle,   D1[128 x 192] = A^T B, K = 64
//   (2) a K-major A operand written by threads in the canonical SW128 layout and a K-major B from TMA,
//       D2[128 x 64] = P V^T, K = 64
// against a host reference.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tc_probe tc_probe.cu
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>

#include "../stereoformer_b200/csrc/tc_common.cuh"

using namespace sftc;

constexpr int M = 128, N1 = 192, K = 64, N2 = 64;

__global__ void __launch_bounds__(128, 1)
probe(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
      const __grid_constant__ CUtensorMap mapV, const __nv_bfloat16* __restrict__ P, float* __restrict__ D1,
      float* __restrict__ D2) {
  extern __shared__ __align__(1024) unsigned char smem[];
  unsigned char* sA = smem;                 // 2 boxes x [64 k][128 B]
  unsigned char* sB = sA + 2 * 8192;        // 3 boxes
  unsigned char* sV = sB + 3 * 8192;        // [64 c][128 B]
  unsigned char* sP = sV + 8192;            // [128 q][128 B]
  __shared__ __align__(8) uint64_t bar_load, bar_mma;
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    mbar_init(&bar_load, 1);
    mbar_init(&bar_mma, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc(&tmem_base_s, 256);
  // P: thread q writes its row (K-major, SW128)
  {
    const int q = threadIdx.x;
    for (int ch = 0; ch < 8; ++ch) {
      uint4 v = *reinterpret_cast<const uint4*>(P + q * 64 + ch * 8);
      *reinterpret_cast<uint4*>(sP + (q >> 3) * 1024 + (q & 7) * 128 + ((ch ^ (q & 7)) << 4)) = v;
    }
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_s;

  if (threadIdx.x == 0) {
    mbar_expect_tx(&bar_load, 6 * 8192);
    for (int i = 0; i < 2; ++i) tma_load_2d(sA + i * 8192, &mapA, &bar_load, i * 64, 0);
    for (int i = 0; i < 3; ++i) tma_load_2d(sB + i * 8192, &mapB, &bar_load, i * 64, 0);
    tma_load_2d(sV, &mapV, &bar_load, 0, 0);
    mbar_wait(&bar_load, 0);
    tc_fence_after();
    // (1) MN-major x MN-major
    const uint32_t idesc1 = make_idesc_bf16(M, N1, /*a_mn_major=*/1, /*b_mn_major=*/1);
    for (int j = 0; j < K / 16; ++j) {
      uint64_t da = make_smem_desc(smem_u32(sA) + j * 2048, /*lbo=*/8192, /*sbo=*/1024);
      uint64_t db = make_smem_desc(smem_u32(sB) + j * 2048, 8192, 1024);
      umma_bf16(tmem, da, db, idesc1, j > 0);
    }
    // (2) K-major x K-major into columns 192..255
    const uint32_t idesc2 = make_idesc_bf16(M, N2, 0, 0);
    for (int j = 0; j < K / 16; ++j) {
      uint64_t da = make_smem_desc(smem_u32(sP) + j * 32, 16, 1024);
      uint64_t db = make_smem_desc(smem_u32(sV) + j * 32, 16, 1024);
      umma_bf16(tmem + 192, da, db, idesc2, j > 0);
    }
    umma_commit(&bar_mma);
  }
  __syncwarp();
  mbar_wait(&bar_mma, 0);
  tc_fence_after();
  const int row = warp * 32 + lane;
  const uint32_t taddr = tmem + ((warp * 32) << 16);
  for (int c0 = 0; c0 < 256; c0 += 32) {
    uint32_t v[32];
    tmem_ld32(taddr + c0, v);
    tmem_ld_wait();
    for (int j = 0; j < 32; ++j) {
      int c = c0 + j;
      if (c < N1) D1[row * N1 + c] = __uint_as_float(v[j]);
      else D2[row * N2 + (c - N1)] = __uint_as_float(v[j]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_free(tmem, 256);
}

typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                              const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                              CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static CUtensorMap map2d(encode_fn enc, void* base, int inner, int rows, int box_inner, int box_rows) {
  CUtensorMap m;
  cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)inner * 2};
  cuuint32_t box[2] = {(cuuint32_t)box_inner, (cuuint32_t)box_rows};
  cuuint32_t es[2] = {1, 1};
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); exit(1); }
  return m;
}

int main() {
  void* fp = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q);
  encode_fn enc = (encode_fn)fp;
  std::vector<__nv_bfloat16> A(K * M), B(K * N1), V(N2 * K), P(M * K);
  srand(1);
  auto rnd = [] { return __float2bfloat16((rand() % 2001 - 1000) / 500.0f); };
  for (auto& x : A) x = rnd();
  for (auto& x : B) x = rnd();
  for (auto& x : V) x = rnd();
  for (auto& x : P) x = rnd();
  __nv_bfloat16 *dA, *dB, *dV, *dP;
  float *d1, *d2;
  cudaMalloc(&dA, A.size() * 2); cudaMalloc(&dB, B.size() * 2); cudaMalloc(&dV, V.size() * 2); cudaMalloc(&dP, P.size() * 2);
  cudaMalloc(&d1, M * N1 * 4); cudaMalloc(&d2, M * N2 * 4);
  cudaMemcpy(dA, A.data(), A.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, B.data(), B.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dV, V.data(), V.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dP, P.data(), P.size() * 2, cudaMemcpyHostToDevice);
  CUtensorMap mA = map2d(enc, dA, M, K, 64, 64), mB = map2d(enc, dB, N1, K, 64, 64), mV = map2d(enc, dV, K, N2, 64, 64);
  size_t smem = 2 * 8192 + 3 * 8192 + 8192 + 16384 + 1024;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  probe<<<1, 128, smem>>>(mA, mB, mV, dP, d1, d2);
  cudaError_t e = cudaDeviceSynchronize();
  printf("kernel: %s\n", cudaGetErrorString(e));
  if (e != cudaSuccess) return 1;
  std::vector<float> h1(M * N1), h2(M * N2);
  cudaMemcpy(h1.data(), d1, h1.size() * 4, cudaMemcpyDeviceToHost);
  cudaMemcpy(h2.data(), d2, h2.size() * 4, cudaMemcpyDeviceToHost);
  double e1 = 0, e2 = 0;
  for (int m = 0; m < M; ++m)
    for (int n = 0; n < N1; ++n) {
      float s = 0;
      for (int k = 0; k < K; ++k) s += __bfloat162float(A[k * M + m]) * __bfloat162float(B[k * N1 + n]);
      e1 = fmax(e1, fabs(s - h1[m * N1 + n]));
    }
  for (int m = 0; m < M; ++m)
    for (int n = 0; n < N2; ++n) {
      float s = 0;
      for (int k = 0; k < K; ++k) s += __bfloat162float(P[m * K + k]) * __bfloat162float(V[n * K + k]);
      e2 = fmax(e2, fabs(s - h2[m * N2 + n]));
    }
  printf("probe1 (MN-major x MN-major) max err %.3g   probe2 (K-major x K-major) max err %.3g\n", e1, e2);
  printf("sample D1[0][0..3] = %f %f %f %f\n", h1[0], h1[1], h1[2], h1[3]);
  return (e1 < 1e-2 && e2 < 1e-2) ? 0 : 2;
}

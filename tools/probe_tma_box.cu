// Dev probe: which 3-D TMA box shapes load cleanly on this GPU, and does FFMA2 run?  (dcn_small_tma_kernel bring-up)
//   nvcc -gencode arch=compute_100a,code=sm_100a -o tools/probe_tma_box tools/probe_tma_box.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>
#include <cstdlib>

typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                              const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                              CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

__device__ __forceinline__ uint32_t s32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__global__ void ffma2_kernel(float* out) {
  unsigned long long a = 0x3f8000003f800000ULL, b = 0x4000000040000000ULL, c = 0x3f0000003f000000ULL;
  asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(c) : "l"(a), "l"(b));
  out[0] = __uint_as_float(static_cast<unsigned>(c & 0xffffffffULL));
}

__global__ void box_kernel(const __grid_constant__ CUtensorMap map, float* out, int bytes, int c0, int c1, int c2) {
  extern __shared__ __align__(128) unsigned char sm[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + ((bytes + 127) / 128) * 128);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(s32(sm)), "l"(reinterpret_cast<uint64_t>(&map)), "r"(s32(bar)), "r"(c0), "r"(c1), "r"(c2) : "memory");
  }
  __syncthreads();
  uint32_t ok = 0;
  while (!ok) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(s32(bar)) : "memory");
  }
  float s = 0.f;
  const float* f = reinterpret_cast<const float*>(sm);
  for (int i = threadIdx.x; i < bytes / 4; i += blockDim.x) s += f[i];
  atomicAdd(out, s);
}

int main(int argc, char** argv) {
  void* fp = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q);
  encode_fn enc = reinterpret_cast<encode_fn>(fp);
  float* out;
  cudaMalloc(&out, 4);
  ffma2_kernel<<<1, 32>>>(out);
  printf("ffma2: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  const int W = 60, H = 34, P = 32;
  float* in;
  cudaMalloc(&in, sizeof(float) * W * H * P);
  std::vector<float> h(W * H * P, 1.0f);
  cudaMemcpy(in, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
  // one configuration per process (a fault kills the context): argv = b0 b1 b2 c0 c1
  int b[3] = {52, 27, 16}, c0 = -9, c1 = -9;
  if (argc >= 6) { b[0] = atoi(argv[1]); b[1] = atoi(argv[2]); b[2] = atoi(argv[3]); c0 = atoi(argv[4]); c1 = atoi(argv[5]); }
  {
    CUtensorMap m;
    cuuint64_t dims[3] = {W, H, P};
    cuuint64_t strides[2] = {W * 4, (cuuint64_t)W * H * 4};
    cuuint32_t box[3] = {(cuuint32_t)b[0], (cuuint32_t)b[1], (cuuint32_t)b[2]};
    cuuint32_t es[3] = {1, 1, 1};
    CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, in, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    const int bytes = b[0] * b[1] * b[2] * 4;
    cudaFuncSetAttribute(box_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes + 256);
    cudaMemset(out, 0, 4);
    box_kernel<<<1, 128, bytes + 256>>>(m, out, bytes, c0, c1, 0);
    cudaError_t e = cudaDeviceSynchronize();
    float v = 0;
    if (e == cudaSuccess) cudaMemcpy(&v, out, 4, cudaMemcpyDeviceToHost);
    printf("box %3d x %2d x %2d (%6d B) at (%d,%d) encode=%d run=%s sum=%.0f\n", b[0], b[1], b[2], bytes, c0, c1, (int)r,
           cudaGetErrorString(e), v);
  }
  return 0;
}

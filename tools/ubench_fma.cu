// Dev microbenchmark: FP32 FMA issue rate on sm_100a for the cost-volume inner-loop shape
// (96 independent accumulators, operands in registers): scalar FFMA vs packed fma.rn.f32x2.
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void k(float* out, int iters, float a0) {
  float acc[96];
#pragma unroll
  for (int i = 0; i < 96; ++i) acc[i] = i;
  float l[4] = {a0, a0 + 1, a0 + 2, a0 + 3};
  float r[28];
#pragma unroll
  for (int i = 0; i < 28; ++i) r[i] = a0 * i;
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {
#pragma unroll
      for (int dd = 0; dd < 24; ++dd)
#pragma unroll
        for (int t = 0; t < 4; ++t) acc[dd * 4 + t] = fmaf(l[t], r[24 + t - dd], acc[dd * 4 + t]);
    } else {
      // packed: pairs over (t, t+1) with even-aligned r pairs only (throughput probe, not the real indexing)
#pragma unroll
      for (int dd = 0; dd < 24; ++dd)
#pragma unroll
        for (int t = 0; t < 4; t += 2) {
          unsigned long long A, B, C, Dv;
          int ri = (24 + t - dd) & ~1;
          asm("mov.b64 %0, {%1, %2};" : "=l"(A) : "f"(l[t]), "f"(l[t + 1]));
          asm("mov.b64 %0, {%1, %2};" : "=l"(B) : "f"(r[ri]), "f"(r[ri + 1]));
          asm("mov.b64 %0, {%1, %2};" : "=l"(C) : "f"(acc[dd * 4 + t]), "f"(acc[dd * 4 + t + 1]));
          asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(Dv) : "l"(A), "l"(B), "l"(C));
          asm("mov.b64 {%0, %1}, %2;" : "=f"(acc[dd * 4 + t]), "=f"(acc[dd * 4 + t + 1]) : "l"(Dv));
        }
    }
#pragma unroll
    for (int i = 0; i < 28; ++i) r[i] += 1e-9f;  // keep r live / varying (28 FADD per 96 FMA)
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 96; ++i) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// MODE 2: same FMA count, every FFMA reads only two distinct registers besides the accumulator pattern a*a+c
// (register-bank / operand-port probe); MODE 3: reports SM clock (clock64 vs globaltimer) under FMA load.
__global__ void k2(float* out, int iters, float a0, long long* clk) {
  float acc[96];
#pragma unroll
  for (int i = 0; i < 96; ++i) acc[i] = i;
  float l[4] = {a0, a0 + 1, a0 + 2, a0 + 3};
  long long c0 = clock64();
  unsigned long long t0;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 96; ++i) acc[i] = fmaf(l[i & 3], l[i & 3], acc[i]);
  }
  long long c1 = clock64();
  unsigned long long t1;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
  float s = 0;
#pragma unroll
  for (int i = 0; i < 96; ++i) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) { clk[0] = c1 - c0; clk[1] = (long long)(t1 - t0); }
}

template <int MODE>
void run(int warps_per_sm, const char* name) {
  float* out;
  cudaMalloc(&out, 148 * 1024 * 4);
  int iters = 20000;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<148, warps_per_sm * 32>>>(out, 100, 1.0f);
  cudaEventRecord(e0);
  k<MODE><<<148, warps_per_sm * 32>>>(out, iters, 1.0f);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  double fma = 148.0 * warps_per_sm * 32 * 96.0 * iters;
  printf("%s warps/SM=%2d  %.1f TFLOP/s  (%.1f FMA/clk/SM at 1.965GHz)  err=%s\n", name, warps_per_sm, 2 * fma / ms / 1e9,
         fma / (ms * 1e-3) / 148 / 1.965e9, cudaGetErrorString(cudaGetLastError()));
  cudaFree(out);
}

void run2(int warps_per_sm) {
  float* out; long long* clk; long long h[2];
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&clk, 16);
  int iters = 40000;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k2<<<148, warps_per_sm * 32>>>(out, 100, 1.0f, clk);
  cudaEventRecord(e0);
  k2<<<148, warps_per_sm * 32>>>(out, iters, 1.0f, clk);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  cudaMemcpy(h, clk, 16, cudaMemcpyDeviceToHost);
  double fma = 148.0 * warps_per_sm * 32 * 96.0 * iters;
  double mhz = (double)h[0] / (double)h[1] * 1e3;
  printf("FFMA a*a+c warps/SM=%2d  %.1f TFLOP/s  SM clock under load %.0f MHz  -> %.1f FMA/clk/SM (measured clock)\n",
         warps_per_sm, 2 * fma / ms / 1e9, mhz, fma / (ms * 1e-3) / 148 / (mhz * 1e6));
}

int main() {
  for (int w : {4, 8, 16}) run2(w);
  for (int w : {4, 8, 16}) run<0>(w, "FFMA  ");
  for (int w : {4, 8, 16}) run<1>(w, "FFMA2 ");
  return 0;
}

// Microbenchmark: FFMA vs FFMA2 (fma.rn.f32x2) issue throughput on sm_100a.  nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ float lo(u64 a) { float x, y; asm("mov.b64 {%0,%1}, %2;" : "=f"(x), "=f"(y) : "l"(a)); return x + y; }

template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, int iters, float s) {
  float acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = threadIdx.x * 0.001f + i;
  u64 a2[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a2[i] = pk(acc[2 * i], acc[2 * i + 1]);
  const u64 s2 = pk(s, s * 1.0001f);
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {            // 16 independent scalar FFMA chains
#pragma unroll
      for (int i = 0; i < 16; ++i) acc[i] = fmaf(acc[i], s, 0.5f);
    } else if (MODE == 1) {     // 8 independent FFMA2 chains (same flops)
#pragma unroll
      for (int i = 0; i < 8; ++i) a2[i] = fma2(a2[i], s2, a2[(i + 1) & 7]);
    } else {                    // FFMA2 with scalar broadcast operand
#pragma unroll
      for (int i = 0; i < 8; ++i) a2[i] = fma2(a2[i], pk(s, s), a2[(i + 1) & 7]);
    }
  }
  float r = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) r += acc[i];
#pragma unroll
  for (int i = 0; i < 8; ++i) r += lo(a2[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

int main() {
  float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000;
  for (int mode = 0; mode < 3; ++mode) {
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      if (mode == 0) k<0><<<148 * 8, 256>>>(d, iters, 0.999f);
      if (mode == 1) k<1><<<148 * 8, 256>>>(d, iters, 0.999f);
      if (mode == 2) k<2><<<148 * 8, 256>>>(d, iters, 0.999f);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      const double fma = (double)148 * 8 * 256 * 16 * iters;
      if (rep) printf("mode %d: %.3f ms  %.2f TFMA/s  (%.1f fma-lanes/clk/SM at 1.9 GHz)\n", mode, ms, fma / ms / 1e9, fma / ms / 1e-3 / 148 / 1.9e9);
    }
  }
  printf("err=%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}

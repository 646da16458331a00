// Shared helpers for the libhmm_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/hmm_b200.h"

namespace hmm {

constexpr float kEps = 1e-16f;  // reference: MsaHmmCell.py:33

void set_error(const char* fmt, ...);
int check_cuda(cudaError_t e, const char* what);

#define HMM_CUDA_OK(expr)                                      \
  do {                                                         \
    int _rc = ::hmm::check_cuda((expr), #expr);                \
    if (_rc != 0) return _rc;                                  \
  } while (0)

#define HMM_REQUIRE(cond, ...)                                 \
  do {                                                         \
    if (!(cond)) {                                             \
      ::hmm::set_error(__VA_ARGS__);                           \
      return HMM_B200_EINVAL;                                  \
    }                                                          \
  } while (0)

__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }

__device__ __forceinline__ float dot4(const float4 a, const float4 b, float acc) {
  acc = fmaf(a.x, b.x, acc);
  acc = fmaf(a.y, b.y, acc);
  acc = fmaf(a.z, b.z, acc);
  acc = fmaf(a.w, b.w, acc);
  return acc;
}

// max / sum over the 16 lanes of a half warp (all 32 lanes must call)
__device__ __forceinline__ float half_max(float v) {
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o, 16));
  return v;
}
__device__ __forceinline__ float half_sum(float v) {
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o, 16);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

inline int cdiv(int a, int b) { return (a + b - 1) / b; }
inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

}  // namespace hmm

// Shared helpers for the libhmm_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <mutex>
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

// ---- per-device library state (one process may drive several GPUs, from several threads) -----------------------
constexpr int kMaxDevices = 64;
inline std::mutex& state_mutex() { static std::mutex m; return m; }
inline int current_device() {
  int d = 0;
  if (cudaGetDevice(&d) != cudaSuccess) { (void)cudaGetLastError(); d = 0; }
  return (d < 0 || d >= kMaxDevices) ? 0 : d;
}
inline int device_sm_count() {
  static int sms[kMaxDevices] = {0};
  const int d = current_device();
  std::lock_guard<std::mutex> lk(state_mutex());
  if (sms[d] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, d) != cudaSuccess || n <= 0) { (void)cudaGetLastError(); n = 148; }
    sms[d] = n;
  }
  return sms[d];
}
// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) applies to the CURRENT device only: remembered per (device, slot);
// slots are library-wide kernel ids (see SMEM_SLOT_* in the callers)
constexpr int kSmemSlots = 32;   // gene15 0-8, dense_tc 9, seq_kernel builds 11-22, fused 23, shard kernels 24-25, viterbi 26, boundary groups 27, fused (shard build) 28
template <typename K>
inline int ensure_dynamic_smem(K kernel, int slot, size_t bytes) {
  static size_t done[kMaxDevices][kSmemSlots] = {};
  const int d = current_device();
  std::lock_guard<std::mutex> lk(state_mutex());
  if (bytes <= done[d][slot]) return 0;
  HMM_CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  done[d][slot] = bytes;
  return 0;
}

inline int cdiv(int a, int b) { return (a + b - 1) / b; }
inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

}  // namespace hmm

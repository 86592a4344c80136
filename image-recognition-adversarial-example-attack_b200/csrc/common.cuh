// Shared helpers for libadvb200 (error reporting, launch accounting, small device utilities).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <atomic>
#include <string>

#include "../../include/advb200.h"

namespace advb200 {

void set_error(const std::string& msg);
extern std::atomic<long long> g_launches;

inline int fail(int code, const std::string& msg) {
  set_error(msg);
  return code;
}

#define ADV_CHECK_CUDA(expr)                                                                   \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess)                                                                     \
      return ::advb200::fail(ADVB200_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); \
  } while (0)

#define ADV_CHECK_ARG(cond, msg)                                              \
  do {                                                                        \
    if (!(cond)) return ::advb200::fail(ADVB200_ERR_ARG, std::string(msg));   \
  } while (0)

// call after every kernel launch: counts it and surfaces launch-configuration errors
#define ADV_LAUNCHED()                                  \
  do {                                                  \
    ::advb200::g_launches.fetch_add(1);                 \
    ADV_CHECK_CUDA(cudaGetLastError());                 \
  } while (0)

inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

constexpr int kNumSMs = 148;  // B200

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// block-wide sum, result valid in thread 0 (fixed order -> deterministic)
template <int kThreads>
__device__ __forceinline__ float block_sum(float v, float* smem /* >= kThreads/32 floats */) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) smem[wid] = v;
  __syncthreads();
  float r = 0.f;
  if (threadIdx.x == 0) {
#pragma unroll
    for (int i = 0; i < kThreads / 32; ++i) r += smem[i];
  }
  return r;
}

}  // namespace advb200

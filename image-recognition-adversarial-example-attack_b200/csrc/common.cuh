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

// Division of a 31-bit index by a launch constant as one widening multiply and a shift (Granlund-Montgomery: with
// l = ceil(log2 d) and m = ceil(2^(31+l) / d) < 2^32, floor(t / d) = (t * m) >> (31 + l) for every t < 2^31).  The
// strip kernels split their flat thread index into (plane, row chunk, column group) with two of these instead of two
// runtime 32-bit divisions (~20 instructions each).
struct FastDiv { uint32_t m, sh, d; };
inline FastDiv make_fastdiv(uint32_t d) {
  FastDiv f; f.d = d;
  uint32_t l = 0;
  while ((1ull << l) < d) ++l;
  f.sh = 31 + l;
  f.m = (uint32_t)(((1ull << f.sh) + d - 1) / d);
  return f;
}
__device__ __forceinline__ uint32_t fast_div(uint32_t t, const FastDiv& f) {
  return (uint32_t)(((unsigned long long)t * f.m) >> f.sh);
}
// the 64-bit strip index of oversized inputs keeps the ordinary division
__device__ __forceinline__ int64_t fast_div(int64_t t, const FastDiv& f) { return t / (int64_t)f.d; }

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

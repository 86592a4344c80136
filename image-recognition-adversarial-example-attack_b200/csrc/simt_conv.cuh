// fp32 validation-mode implicit-GEMM convolution (CUDA cores) — interface.
#pragma once
#include "common.cuh"

namespace advb200 {

struct SimtConvParams {
  const float* in;    // [n, hi, wi, ci] NHWC
  const float* wm;    // [(kh*kw*ci), co] row-major weight matrix
  float* out;         // [n, ho, wo, co]
  const float* bias;  // [co] or null
  const float* add;   // out-shaped or null
  const float* mask;  // out-shaped or null: out = 0 where mask <= 0
  int n, hi, wi, ci, ho, wo, co;
  int kh, kw, stride, pad;
  int relu;
  int dgrad;  // 0: forward gather (src = dst*stride - pad + r); 1: transposed gather
};

int launch_simt_conv(const SimtConvParams& p, cudaStream_t stream);

// out[m, n] = sum_k a[m, k] * b[n, k] (+ bias[n]); fp32, both operands K-contiguous
// (deterministic split-K: `partial` is scratch of fc_nt_partial_bytes(m, n) bytes)
int64_t fc_nt_partial_bytes(int m, int n);
int launch_fc_nt(const float* a, const float* b, const float* bias, float* out, float* partial, int m, int n,
                 int k, cudaStream_t stream);

}  // namespace advb200

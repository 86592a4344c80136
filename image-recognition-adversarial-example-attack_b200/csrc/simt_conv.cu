// fp32 validation-mode convolution: CUDA-core implicit GEMM (no tensor cores, no TF32).
// Same semantics and fused epilogue as the tcgen05 kernel, arbitrary shapes with bounds checks,
// also used for the fp32 classifier head (fc forward / input-gradient as 1x1 "convolutions").
#include "simt_conv.cuh"

namespace advb200 {

namespace {

constexpr int TM = 64, TN = 64, TK = 16;

// out[pix, co] = sum_{r,s,ci} in[src(pix,r,s), ci] * wm[(r*kw+s)*ci_n + ci, co]
__global__ void __launch_bounds__(256)
simt_conv_kernel(const SimtConvParams p) {
  __shared__ float As[TK][TM + 4];
  __shared__ float Bs[TK][TN + 4];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const long long m_total = (long long)p.n * p.ho * p.wo;
  const long long m0 = (long long)blockIdx.x * TM;
  const int n0 = blockIdx.y * TN;
  const int k_total = p.kh * p.kw * p.ci;

  // A loader: thread -> (row = tid/4, 4 consecutive k)
  const int a_row = tid >> 2, a_k = (tid & 3) * 4;
  const long long a_m = m0 + a_row;
  int a_n = 0, a_h = 0, a_w = 0;
  const bool a_valid = a_m < m_total;
  if (a_valid) {
    a_w = (int)(a_m % p.wo);
    a_h = (int)((a_m / p.wo) % p.ho);
    a_n = (int)(a_m / ((long long)p.wo * p.ho));
  }
  // B loader: thread -> (k = tid/16, 4 consecutive co)
  const int b_k = tid >> 4, b_c = (tid & 15) * 4;

  float acc[4][4] = {};
  for (int k0 = 0; k0 < k_total; k0 += TK) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int k = k0 + a_k + j;
      float v = 0.f;
      if (a_valid && k < k_total) {
        const int rs = k / p.ci, ci = k - rs * p.ci;
        const int r = rs / p.kw, s = rs - r * p.kw;
        int sh, sw;
        bool ok;
        if (!p.dgrad) {
          sh = a_h * p.stride - p.pad + r;
          sw = a_w * p.stride - p.pad + s;
          ok = sh >= 0 && sh < p.hi && sw >= 0 && sw < p.wi;
        } else {
          const int th = a_h + p.pad - r, tw = a_w + p.pad - s;
          ok = th >= 0 && tw >= 0 && (th % p.stride) == 0 && (tw % p.stride) == 0;
          sh = th / p.stride; sw = tw / p.stride;
          ok = ok && sh < p.hi && sw < p.wi;
        }
        if (ok) v = __ldg(p.in + (((long long)a_n * p.hi + sh) * p.wi + sw) * p.ci + ci);
      }
      As[a_k + j][a_row] = v;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int k = k0 + b_k, co = n0 + b_c + j;
      Bs[b_k][b_c + j] = (k < k_total && co < p.co) ? __ldg(p.wm + (long long)k * p.co + co) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < TK; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = Bs[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }

#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long m = m0 + ty * 4 + i;
    if (m >= m_total) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int co = n0 + tx * 4 + j;
      if (co >= p.co) continue;
      const long long o = m * p.co + co;
      float v = acc[i][j];
      if (p.bias) v += __ldg(p.bias + co);
      if (p.add) v += __ldg(p.add + o);
      if (p.relu) v = fmaxf(v, 0.f);
      if (p.mask && !(__ldg(p.mask + o) > 0.f)) v = 0.f;
      p.out[o] = v;
    }
  }
}

// out[m, n] = sum_k a[m, k] * b[n, k] (+ bias[n]) — both operands K-contiguous (fp32 classifier head).
// Deterministic split-K: kSplit slices of K write partial sums, a second kernel adds them in a fixed order
// (fp32 atomics would make the logits run-to-run different).
constexpr int kFcSplit = 8;
// 128 x 64 block tile, 128 threads, 8 x 8 outputs per thread: 4 LDS.128 feed 64 FMAs (the 4 x 4 version ran at 35 %
// of the fp32 rate, shared-memory reads and FMAs competing for issue slots)
constexpr int kFcTM = 128, kFcTN = 64;
__global__ void __launch_bounds__(128)
fc_nt_splitk_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ part,
                    int m_total, int n_total, int k_total, int k_chunk) {
  constexpr int KC = 16;                                 // K elements staged per iteration
  __shared__ __align__(16) float As[KC][kFcTM + 4];
  __shared__ __align__(16) float Bs[KC][kFcTN + 4];
  const int tid = threadIdx.x;
  const int tx = tid & 7, ty = tid >> 3;                 // 8 column groups x 16 row groups
  const int m0 = blockIdx.x * kFcTM, n0 = blockIdx.y * kFcTN;
  const int kbeg = blockIdx.z * k_chunk;
  const int kend = min(k_total, kbeg + k_chunk);
  // loads: A tile 128 rows x 16 k = 512 float4 (4 per thread), B tile 64 rows x 16 k = 256 float4 (2 per thread)
  float4 va[4], vb[2];
  auto fetch = [&](int k0) {
#pragma unroll
    for (int h = 0; h < 4; ++h) {
      const int idx = tid + 128 * h, r = idx >> 2, k = k0 + (idx & 3) * 4;
      va[h] = (m0 + r < m_total && k < kend) ? __ldg(reinterpret_cast<const float4*>(a + (long long)(m0 + r) * k_total + k))
                                             : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int idx = tid + 128 * h, r = idx >> 2, k = k0 + (idx & 3) * 4;
      vb[h] = (n0 + r < n_total && k < kend) ? __ldg(reinterpret_cast<const float4*>(b + (long long)(n0 + r) * k_total + k))
                                             : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };
  float acc[8][8] = {};
  fetch(kbeg);
  for (int k0 = kbeg; k0 < kend; k0 += KC) {
#pragma unroll
    for (int h = 0; h < 4; ++h) {
      const int idx = tid + 128 * h, r = idx >> 2, kq = (idx & 3) * 4;
      As[kq][r] = va[h].x; As[kq + 1][r] = va[h].y; As[kq + 2][r] = va[h].z; As[kq + 3][r] = va[h].w;
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int idx = tid + 128 * h, r = idx >> 2, kq = (idx & 3) * 4;
      Bs[kq][r] = vb[h].x; Bs[kq + 1][r] = vb[h].y; Bs[kq + 2][r] = vb[h].z; Bs[kq + 3][r] = vb[h].w;
    }
    __syncthreads();
    if (k0 + KC < kend) fetch(k0 + KC);                  // next slice in flight while this one is consumed
#pragma unroll
    for (int kk = 0; kk < KC; ++kk) {
      // rows ty*4 .. +3 and 64 + ty*4 .. +3; columns tx*4 .. +3 and 32 + tx*4 .. +3 (conflict-free float4 reads)
      const float4 a0 = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[kk][64 + ty * 4]);
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      const float4 b1 = *reinterpret_cast<const float4*>(&Bs[kk][32 + tx * 4]);
      const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
  float* po = part + (long long)blockIdx.z * m_total * n_total;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
    if (m >= m_total) continue;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int n = n0 + (j < 4 ? tx * 4 + j : 32 + tx * 4 + (j - 4));
      if (n < n_total) po[(long long)m * n_total + n] = acc[i][j];
    }
  }
}

__global__ void __launch_bounds__(256)
fc_reduce_kernel(const float* __restrict__ part, const float* __restrict__ bias, float* __restrict__ out,
                 long long mn, int n_total) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= mn) return;
  float s = 0.f;
#pragma unroll
  for (int z = 0; z < kFcSplit; ++z) s += part[z * mn + t];
  out[t] = s + (bias ? __ldg(bias + (int)(t % n_total)) : 0.f);
}

}  // namespace

int64_t fc_nt_partial_bytes(int m, int n) { return (int64_t)kFcSplit * m * n * 4; }

int launch_fc_nt(const float* a, const float* b, const float* bias, float* out, float* partial, int m, int n,
                 int k, cudaStream_t stream) {
  if (m == 0) return ADVB200_OK;
  ADV_CHECK_ARG(k % 4 == 0 && partial, "fc: K must be a multiple of 4 and the split-K scratch non-null");
  const int k_chunk = (((k + kFcSplit - 1) / kFcSplit) + 15) / 16 * 16;
  dim3 grid((unsigned)((m + kFcTM - 1) / kFcTM), (unsigned)((n + kFcTN - 1) / kFcTN), kFcSplit);
  fc_nt_splitk_kernel<<<grid, 128, 0, stream>>>(a, b, partial, m, n, k, k_chunk);
  ADV_LAUNCHED();
  const long long mn = (long long)m * n;
  fc_reduce_kernel<<<(unsigned)((mn + 255) / 256), 256, 0, stream>>>(partial, bias, out, mn, n);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

int launch_simt_conv(const SimtConvParams& p, cudaStream_t stream) {
  const long long m_total = (long long)p.n * p.ho * p.wo;
  if (m_total == 0) return ADVB200_OK;
  dim3 grid((unsigned)((m_total + TM - 1) / TM), (unsigned)((p.co + TN - 1) / TN));
  simt_conv_kernel<<<grid, 256, 0, stream>>>(p);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

}  // namespace advb200

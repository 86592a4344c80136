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
// (fp32 atomics would make the logits run-to-run different).  64 x 64 block tile, 4 x 4 outputs per thread.
constexpr int kFcSplit = 8;
__global__ void __launch_bounds__(256)
fc_nt_splitk_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ part,
                    int m_total, int n_total, int k_total, int k_chunk) {
  constexpr int KC = 32;                                 // K elements staged per iteration
  __shared__ __align__(16) float As[KC][68];
  __shared__ __align__(16) float Bs[KC][68];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.x * 64, n0 = blockIdx.y * 64;
  const int kbeg = blockIdx.z * k_chunk;
  const int kend = min(k_total, kbeg + k_chunk);
  const int lr = tid >> 2, lk = (tid & 3) * 8;          // this thread's loads: row lr, k offsets lk .. lk+7
  const bool a_ok = m0 + lr < m_total, b_ok = n0 + lr < n_total;
  const float* ap = a + (long long)(m0 + lr) * k_total;
  const float* bp = b + (long long)(n0 + lr) * k_total;
  float4 va[2], vb[2];
  auto fetch = [&](int k0) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int k = k0 + lk + 4 * h;
      va[h] = (a_ok && k < kend) ? __ldg(reinterpret_cast<const float4*>(ap + k)) : make_float4(0.f, 0.f, 0.f, 0.f);
      vb[h] = (b_ok && k < kend) ? __ldg(reinterpret_cast<const float4*>(bp + k)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };
  float acc[4][4] = {};
  fetch(kbeg);
  for (int k0 = kbeg; k0 < kend; k0 += KC) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int r = lk + 4 * h;
      As[r][lr] = va[h].x; As[r + 1][lr] = va[h].y; As[r + 2][lr] = va[h].z; As[r + 3][lr] = va[h].w;
      Bs[r][lr] = vb[h].x; Bs[r + 1][lr] = vb[h].y; Bs[r + 2][lr] = vb[h].z; Bs[r + 3][lr] = vb[h].w;
    }
    __syncthreads();
    if (k0 + KC < kend) fetch(k0 + KC);                  // next slice in flight while this one is consumed
#pragma unroll
    for (int kk = 0; kk < KC; ++kk) {
      const float4 a4 = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      const float4 b4 = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      const float av[4] = {a4.x, a4.y, a4.z, a4.w}, bv[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
  float* po = part + (long long)blockIdx.z * m_total * n_total;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= m_total) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
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
  const int k_chunk = (((k + kFcSplit - 1) / kFcSplit) + 31) / 32 * 32;
  dim3 grid((unsigned)((m + 63) / 64), (unsigned)((n + 63) / 64), kFcSplit);
  fc_nt_splitk_kernel<<<grid, 256, 0, stream>>>(a, b, partial, m, n, k, k_chunk);
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

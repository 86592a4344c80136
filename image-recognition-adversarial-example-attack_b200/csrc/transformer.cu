// Non-GEMM layers of the forward-only transformer transfer targets (blackbox_transfer.py:36-54 loads ViT-B/16 and
// Swin-T; :233-251 classifies the ResNet-50 adversarial examples with them — SURVEY.md section 8(f) rank 4).
//
// Every linear layer (patch embedding, QKV, projection, MLP, patch-merging reduction) runs on the tcgen05 implicit-GEMM
// kernel of the attack path as a pointwise conv over the token axis (advb200_linear_*, engine.cu).  This file holds the
// rest: patch extraction, the fp32 residual stream with fused add + LayerNorm, exact (erf) GELU, windowed multi-head
// attention (whole-sequence for ViT, shifted 7x7 windows with relative-position bias and region masks for Swin), Swin's
// patch merging and the pooled head features.  Tokens are rows of [tokens, channels] matrices (channels padded to a
// multiple of 64 for the GEMM kernel; padded channels are kept at exactly zero).  The residual stream stays in fp32 so
// twelve layers of residual adds do not accumulate bf16 rounding; GEMM operands and outputs are bf16.
#include <cstdlib>

#include "layers.cuh"

namespace advb200 {
namespace {

constexpr int kT = 256;
inline unsigned nblk(long long total, int per = kT) { return (unsigned)((total + per - 1) / per); }

// ---- patch extraction: fp32 NCHW image, normalised like attack.py:17 -> bf16 [N*gh*gw, kpad], column (c*ps + i)*ps + j
//      (the flatten order of a Conv2d weight [cout, cin, ps, ps]); columns >= 3*ps*ps are zero
__global__ void patch_im2col_kernel(const float* __restrict__ x, const float* __restrict__ mean3, const float* __restrict__ std3,
                                    bf16* __restrict__ out, int h, int w, int ps, int kpad, long long total) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;     // (patch, column)
  if (t >= total) return;
  const int col = (int)(t % kpad);
  const long long patch = t / kpad;
  const int gw = w / ps, gh = h / ps;
  const int px = (int)(patch % gw), py = (int)((patch / gw) % gh);
  const long long n = patch / ((long long)gw * gh);
  float v = 0.f;
  if (col < 3 * ps * ps) {
    const int j = col % ps, i = (col / ps) % ps, c = col / (ps * ps);
    const float raw = __ldg(x + ((n * 3 + c) * h + (py * ps + i)) * (long long)w + px * ps + j);
    v = __fdiv_rn(__fsub_rn(raw, __ldg(mean3 + c)), __ldg(std3 + c));
  }
  out[t] = __float2bfloat16_rn(v);
}

// ---- ViT token assembly: x32[n, 0, :] = cls + pos[0];  x32[n, 1+p, :] = patch[n, p, :] + pos[1+p]
__global__ void vit_tokens_kernel(const bf16* __restrict__ patch, const float* __restrict__ cls, const float* __restrict__ pos,
                                  float* __restrict__ x32, int tokens, int c, long long total) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;     // (n, token, channel)
  if (t >= total) return;
  const int ch = (int)(t % c);
  const int tok = (int)((t / c) % tokens);
  const long long n = t / ((long long)c * tokens);
  const float base = tok == 0 ? __ldg(cls + ch) : __bfloat162float(patch[(n * (tokens - 1) + tok - 1) * c + ch]);
  x32[t] = base + __ldg(pos + (long long)tok * c + ch);
}

// ---- residual stream update + LayerNorm, one warp per token:
//      mode 0: x32 unchanged; mode 1: x32 += delta; mode 2: x32 = delta        (delta: bf16 [m, c_pad])
//      y = (x32 - mean) / sqrt(var + eps) * gamma + beta over the c_real real channels -> bf16 [m, c_pad] (pad = 0)
template <int kMax>        // channels per lane: c_real <= 32 * kMax
__global__ void __launch_bounds__(kT)
add_layernorm_kernel(float* __restrict__ x32, const bf16* __restrict__ delta, const float* __restrict__ gamma,
                     const float* __restrict__ beta, float eps, bf16* __restrict__ y, long long m, int c_real, int c_pad,
                     int mode) {
  const long long row = (long long)blockIdx.x * (kT / 32) + (threadIdx.x >> 5);
  if (row >= m) return;
  const int lane = threadIdx.x & 31;
  float* xr = x32 + row * c_pad;
  const bf16* dr = delta ? delta + row * c_pad : nullptr;
  float v[kMax];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < kMax; ++i) {
    const int ch = lane + 32 * i;
    float a = 0.f;
    if (ch < c_real) {
      a = mode == 2 ? 0.f : xr[ch];
      if (mode) a += __bfloat162float(dr[ch]);
      if (mode) xr[ch] = a;
    }
    v[i] = a;
    s += a;
  }
  const float mean = warp_sum(s) / (float)c_real;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < kMax; ++i) {
    const int ch = lane + 32 * i;
    const float d = ch < c_real ? v[i] - mean : 0.f;
    q += d * d;
  }
  const float rstd = rsqrtf(warp_sum(q) / (float)c_real + eps);          // biased variance, as nn.LayerNorm
  if (!y) return;
  bf16* yr = y + row * c_pad;
#pragma unroll
  for (int i = 0; i < kMax; ++i) {
    const int ch = lane + 32 * i;
    if (ch < c_real) yr[ch] = __float2bfloat16_rn((v[i] - mean) * rstd * __ldg(gamma + ch) + __ldg(beta + ch));
    else if (ch < c_pad) yr[ch] = __float2bfloat16_rn(0.f);
  }
}

// ---- exact GELU (nn.GELU default, erf form), bf16 -> bf16, 8 elements per thread
__global__ void gelu_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, long long n8) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n8) return;
  const uint4 u = __ldg(reinterpret_cast<const uint4*>(x) + t);
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
  uint32_t o[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[i]));
    const float a = 0.5f * f.x * (1.0f + erff(f.x * 0.70710678118654752f));
    const float b = 0.5f * f.y * (1.0f + erff(f.y * 0.70710678118654752f));
    const __nv_bfloat162 r = __floats2bfloat162_rn(a, b);
    o[i] = *reinterpret_cast<const uint32_t*>(&r);
  }
  reinterpret_cast<uint4*>(y)[t] = make_uint4(o[0], o[1], o[2], o[3]);
}

// ---- multi-head attention over windows of T <= 224 tokens, one CTA per (image, window, head).
//      seq mode (gh == 0): the window is the image's whole token sequence (ViT, T = 197).
//      grid mode: ws x ws windows of a gh x gw token grid after a cyclic shift by `shift` (Swin): token (i, j) of window
//      (wy, wx) lives at grid position ((wy*ws+i+shift) % gh, (wx*ws+j+shift) % gw); with shift > 0, tokens of different
//      regions of the rolled grid do not attend to each other (additive -100, torchvision swin_transformer.py), and
//      `bias` [heads, T, T] is the relative-position bias.
//      scores, softmax and the weighted sum are fp32; K and V of the window sit in shared memory as fp32.
struct AttnParams {
  const bf16* qkv; bf16* out; const float* bias;
  int ld_qkv, ld_out, q_off, k_off, v_off;      // row strides (elements) and the column offsets of q / k / v
  int heads, dim, T;
  int gh, gw, ws, shift, wins_per_img;
  float scale;
};

__device__ __forceinline__ long long attn_row(const AttnParams& p, long long img, int win, int t, int* region) {
  if (p.gh == 0) { if (region) *region = 0; return img * p.T + t; }
  const int wins_w = p.gw / p.ws;
  const int wy = win / wins_w, wx = win % wins_w;
  const int ry = wy * p.ws + t / p.ws, rx = wx * p.ws + t % p.ws;          // position in the rolled grid
  if (region) {
    const int idh = ry < p.gh - p.ws ? 0 : (ry < p.gh - p.shift ? 1 : 2);
    const int idw = rx < p.gw - p.ws ? 0 : (rx < p.gw - p.shift ? 1 : 2);
    *region = p.shift > 0 ? idh * 3 + idw : 0;
  }
  const int py = (ry + p.shift) % p.gh, px = (rx + p.shift) % p.gw;
  return (img * p.gh + py) * (long long)p.gw + px;
}

constexpr int kAttnWarps = 8;
constexpr int kAttnMaxKeysPerLane = 7;          // T <= 224
bool g_attn_force_simt = false;                 // test hook (advb200_test_attention_simt)

__global__ void __launch_bounds__(kAttnWarps * 32)
window_attention_kernel(const AttnParams p) {
  extern __shared__ float smem[];
  const int T = p.T, D = p.dim, kp = D + 1;                                  // K rows padded by one float: no bank conflicts
  float* Ks = smem;                        // [T][D+1]
  float* Vs = Ks + T * kp;                 // [T][D]
  float* Ps = Vs + T * D;                  // [warps][T]
  int* Rg = reinterpret_cast<int*>(Ps + kAttnWarps * T);   // [T] region ids
  const int head = blockIdx.x % p.heads;
  const int win = (blockIdx.x / p.heads) % p.wins_per_img;
  const long long img = blockIdx.x / (p.heads * p.wins_per_img);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  for (int i = threadIdx.x; i < T * D; i += blockDim.x) {
    const int t = i / D, d = i % D;
    const long long row = attn_row(p, img, win, t, nullptr);
    const bf16* src = p.qkv + row * p.ld_qkv + head * D + d;
    Ks[t * kp + d] = __bfloat162float(src[p.k_off]);
    Vs[t * D + d] = __bfloat162float(src[p.v_off]);
  }
  for (int t = threadIdx.x; t < T; t += blockDim.x) { int r; attn_row(p, img, win, t, &r); Rg[t] = r; }
  __syncthreads();

  float* myP = Ps + warp * T;
  for (int qi = warp; qi < T; qi += kAttnWarps) {
    const long long qrow = attn_row(p, img, win, qi, nullptr);
    const bf16* qsrc = p.qkv + qrow * p.ld_qkv + p.q_off + head * D;
    // lane holds q[lane] and q[lane + 32] (scaled); every lane needs every q[d]: broadcast by shuffle in the d loop
    const float q0 = lane < D ? __bfloat162float(qsrc[lane]) * p.scale : 0.f;
    const float q1 = lane + 32 < D ? __bfloat162float(qsrc[lane + 32]) * p.scale : 0.f;
    float acc[kAttnMaxKeysPerLane];
#pragma unroll
    for (int jj = 0; jj < kAttnMaxKeysPerLane; ++jj) acc[jj] = 0.f;
    for (int d = 0; d < D; ++d) {
      const float qd = __shfl_sync(0xffffffffu, d < 32 ? q0 : q1, d & 31);
#pragma unroll
      for (int jj = 0; jj < kAttnMaxKeysPerLane; ++jj) {
        const int j = lane + 32 * jj;
        if (j < T) acc[jj] = fmaf(qd, Ks[j * kp + d], acc[jj]);
      }
    }
    const int myreg = Rg[qi];
    float mx = -INFINITY;
#pragma unroll
    for (int jj = 0; jj < kAttnMaxKeysPerLane; ++jj) {
      const int j = lane + 32 * jj;
      if (j < T) {
        if (p.bias) acc[jj] += __ldg(p.bias + ((long long)head * T + qi) * T + j);
        if (Rg[j] != myreg) acc[jj] += -100.0f;
        mx = fmaxf(mx, acc[jj]);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float sum = 0.f;
#pragma unroll
    for (int jj = 0; jj < kAttnMaxKeysPerLane; ++jj) {
      const int j = lane + 32 * jj;
      if (j < T) { acc[jj] = expf(acc[jj] - mx); sum += acc[jj]; }
    }
    sum = warp_sum(sum);
    const float inv = 1.0f / sum;
#pragma unroll
    for (int jj = 0; jj < kAttnMaxKeysPerLane; ++jj) {
      const int j = lane + 32 * jj;
      if (j < T) myP[j] = acc[jj] * inv;
    }
    __syncwarp();
    float o0 = 0.f, o1 = 0.f;
    for (int j = 0; j < T; ++j) {
      const float pj = myP[j];
      if (lane < D) o0 = fmaf(pj, Vs[j * D + lane], o0);
      if (lane + 32 < D) o1 = fmaf(pj, Vs[j * D + lane + 32], o1);
    }
    bf16* dst = p.out + qrow * p.ld_out + head * D;
    if (lane < D) dst[lane] = __float2bfloat16_rn(o0);
    if (lane + 32 < D) dst[lane + 32] = __float2bfloat16_rn(o1);
    __syncwarp();
  }
}

// ---- the same attention on the tensor cores (warp-level mma.sync m16n8k16, bf16 operands, fp32 accumulation).
//      One CTA per (image, window, head); Q, K [token][dim] and V transposed [dim][token] sit in shared memory as bf16
//      with row pitches (dim + 8, tokens + 8) that make every fragment load conflict-free.  A warp owns 16 query rows at
//      a time: S = Q K^T (kTP / 8 n-tiles x kD / 16 k-steps) stays in registers, bias / region mask / padding mask are
//      applied per accumulator element, the softmax runs on the registers (a row lives in the four lanes of a quad), the
//      un-normalised probabilities are packed to bf16 A fragments in place (the accumulator layout of two neighbouring
//      n-tiles IS the A layout of one k-step) and O = P V is divided by the row sum at the end.
//      kTP = tokens padded to a multiple of 16 (208 for ViT's 197, 64 for Swin's 49), kD = head dim (64 / 32).
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&v);
}

template <int kTP, int kD, int kWarps>
__global__ void __launch_bounds__(kWarps * 32)
window_attention_mma_kernel(const AttnParams p) {
  constexpr int kQP = kD + 8;            // pitch (bf16 elements) of the Q / K rows
  constexpr int kVP = kTP + 8;           // pitch of the transposed V rows
  constexpr int kNT = kTP / 8;           // score n-tiles (8 keys each)
  constexpr int kKS = kD / 16;           // k-steps of Q K^T
  constexpr int kDT = kD / 8;            // output n-tiles (8 dims each)
  extern __shared__ __align__(16) uint8_t smem_raw[];
  bf16* Qs = reinterpret_cast<bf16*>(smem_raw);          // [kTP][kQP]
  bf16* Ks = Qs + kTP * kQP;                              // [kTP][kQP]
  bf16* Vt = Ks + kTP * kQP;                              // [kD][kVP]
  int* Rg = reinterpret_cast<int*>(Vt + kD * kVP);        // [kTP] region ids
  const int T = p.T;
  const int head = blockIdx.x % p.heads;
  const int win = (blockIdx.x / p.heads) % p.wins_per_img;
  const long long img = blockIdx.x / (p.heads * p.wins_per_img);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t4 = lane & 3;

  // stage Q, K (rows) and V (transposed); padded token rows are zero
  for (int i = threadIdx.x; i < kTP * (kD / 8); i += kWarps * 32) {
    const int t = i / (kD / 8), c8 = (i % (kD / 8)) * 8;
    uint4 q = make_uint4(0u, 0u, 0u, 0u), k = q, v = q;
    if (t < T) {
      const long long row = attn_row(p, img, win, t, nullptr);
      const bf16* src = p.qkv + row * p.ld_qkv + head * kD + c8;
      q = *reinterpret_cast<const uint4*>(src + p.q_off);
      k = *reinterpret_cast<const uint4*>(src + p.k_off);
      v = *reinterpret_cast<const uint4*>(src + p.v_off);
    }
    *reinterpret_cast<uint4*>(Qs + t * kQP + c8) = q;
    *reinterpret_cast<uint4*>(Ks + t * kQP + c8) = k;
    const bf16* vh = reinterpret_cast<const bf16*>(&v);
#pragma unroll
    for (int e = 0; e < 8; ++e) Vt[(c8 + e) * kVP + t] = vh[e];
  }
  for (int t = threadIdx.x; t < kTP; t += kWarps * 32) {
    int r = -1;
    if (t < T) attn_row(p, img, win, t, &r);
    Rg[t] = r;
  }
  __syncthreads();

  for (int rt = warp; rt < kTP / 16; rt += kWarps) {
    const int r0 = rt * 16;
    uint32_t a[kKS][4];
#pragma unroll
    for (int ks = 0; ks < kKS; ++ks) {
      const bf16* q0 = Qs + (r0 + g) * kQP + ks * 16 + 2 * t4;
      a[ks][0] = *reinterpret_cast<const uint32_t*>(q0);
      a[ks][1] = *reinterpret_cast<const uint32_t*>(q0 + 8 * kQP);
      a[ks][2] = *reinterpret_cast<const uint32_t*>(q0 + 8);
      a[ks][3] = *reinterpret_cast<const uint32_t*>(q0 + 8 * kQP + 8);
    }
    float s[kNT][4];
#pragma unroll
    for (int nt = 0; nt < kNT; ++nt) {
      s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
#pragma unroll
      for (int ks = 0; ks < kKS; ++ks) {
        const bf16* k0 = Ks + (nt * 8 + g) * kQP + ks * 16 + 2 * t4;
        mma_bf16_16816(s[nt], a[ks], *reinterpret_cast<const uint32_t*>(k0), *reinterpret_cast<const uint32_t*>(k0 + 8));
      }
    }
    // scale, relative-position bias, region mask, padding mask; row maxima (rows r0+g and r0+g+8)
    const int qa = r0 + g, qb = r0 + g + 8;
    const int rga = Rg[qa], rgb = Rg[qb];
    float ma = -INFINITY, mb = -INFINITY;
#pragma unroll
    for (int nt = 0; nt < kNT; ++nt) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int col = nt * 8 + 2 * t4 + e;
        const bool valid = col < T;
        float va = s[nt][e] * p.scale, vb = s[nt][2 + e] * p.scale;
        if (valid && p.bias) {
          if (qa < T) va += __ldg(p.bias + ((long long)head * T + qa) * T + col);
          if (qb < T) vb += __ldg(p.bias + ((long long)head * T + qb) * T + col);
        }
        const int rc = Rg[col];
        if (valid && rc != rga) va += -100.0f;
        if (valid && rc != rgb) vb += -100.0f;
        va = valid ? va : -INFINITY; vb = valid ? vb : -INFINITY;
        s[nt][e] = va; s[nt][2 + e] = vb;
        ma = fmaxf(ma, va); mb = fmaxf(mb, vb);
      }
    }
    ma = fmaxf(ma, __shfl_xor_sync(0xffffffffu, ma, 1)); ma = fmaxf(ma, __shfl_xor_sync(0xffffffffu, ma, 2));
    mb = fmaxf(mb, __shfl_xor_sync(0xffffffffu, mb, 1)); mb = fmaxf(mb, __shfl_xor_sync(0xffffffffu, mb, 2));
    float suma = 0.f, sumb = 0.f;
#pragma unroll
    for (int nt = 0; nt < kNT; ++nt) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        s[nt][e] = expf(s[nt][e] - ma); suma += s[nt][e];               // exp(-inf) = 0 for the padded keys
        s[nt][2 + e] = expf(s[nt][2 + e] - mb); sumb += s[nt][2 + e];
      }
    }
    suma += __shfl_xor_sync(0xffffffffu, suma, 1); suma += __shfl_xor_sync(0xffffffffu, suma, 2);
    sumb += __shfl_xor_sync(0xffffffffu, sumb, 1); sumb += __shfl_xor_sync(0xffffffffu, sumb, 2);
    // O = P V
    float o[kDT][4];
#pragma unroll
    for (int dt = 0; dt < kDT; ++dt) o[dt][0] = o[dt][1] = o[dt][2] = o[dt][3] = 0.f;
#pragma unroll
    for (int kk = 0; kk < kTP / 16; ++kk) {
      uint32_t pa[4];
      pa[0] = pack_bf16x2(s[2 * kk][0], s[2 * kk][1]);
      pa[1] = pack_bf16x2(s[2 * kk][2], s[2 * kk][3]);
      pa[2] = pack_bf16x2(s[2 * kk + 1][0], s[2 * kk + 1][1]);
      pa[3] = pack_bf16x2(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
      for (int dt = 0; dt < kDT; ++dt) {
        const bf16* v0 = Vt + (dt * 8 + g) * kVP + kk * 16 + 2 * t4;
        mma_bf16_16816(o[dt], pa, *reinterpret_cast<const uint32_t*>(v0), *reinterpret_cast<const uint32_t*>(v0 + 8));
      }
    }
    const float inva = 1.0f / suma, invb = 1.0f / sumb;
    if (qa < T) {
      bf16* dst = p.out + attn_row(p, img, win, qa, nullptr) * p.ld_out + head * kD + 2 * t4;
#pragma unroll
      for (int dt = 0; dt < kDT; ++dt) *reinterpret_cast<uint32_t*>(dst + dt * 8) = pack_bf16x2(o[dt][0] * inva, o[dt][1] * inva);
    }
    if (qb < T) {
      bf16* dst = p.out + attn_row(p, img, win, qb, nullptr) * p.ld_out + head * kD + 2 * t4;
#pragma unroll
      for (int dt = 0; dt < kDT; ++dt) *reinterpret_cast<uint32_t*>(dst + dt * 8) = pack_bf16x2(o[dt][2] * invb, o[dt][3] * invb);
    }
  }
}

template <int kTP, int kD, int kWarps>
int launch_attention_mma(const AttnParams& p, long long ctas, cudaStream_t st) {
  constexpr size_t smem = (size_t)(2 * kTP * (kD + 8) + kD * (kTP + 8)) * 2 + (size_t)kTP * 4;
  ADV_CHECK_CUDA(cudaFuncSetAttribute(window_attention_mma_kernel<kTP, kD, kWarps>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  window_attention_mma_kernel<kTP, kD, kWarps><<<(unsigned)ctas, kWarps * 32, smem, st>>>(p);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

// ---- Swin patch merging input: fp32 stream [N, gh, gw, c_pad] -> concat of the four 2x2 neighbours
//      [x(0::2,0::2), x(1::2,0::2), x(0::2,1::2), x(1::2,1::2)] (real channels only) -> fp32 [N*gh/2*gw/2, 4*c_real]
__global__ void patch_merge_gather_kernel(const float* __restrict__ x32, float* __restrict__ out, int gh, int gw, int c_real,
                                          int c_pad, long long total) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;     // (n, oy, ox, 4*c_real)
  if (t >= total) return;
  const int c4 = 4 * c_real;
  const int col = (int)(t % c4);
  const long long tok = t / c4;
  const int ow = gw / 2, oh = gh / 2;
  const int ox = (int)(tok % ow), oy = (int)((tok / ow) % oh);
  const long long n = tok / ((long long)ow * oh);
  const int part = col / c_real, ch = col % c_real;
  const int dy = part & 1, dx = part >> 1;                                    // x0:(0,0) x1:(1,0) x2:(0,1) x3:(1,1)
  out[t] = x32[((n * gh + 2 * oy + dy) * (long long)gw + 2 * ox + dx) * c_pad + ch];
}

// ---- head features: feat[n, :] = mean over `count` tokens (starting at token t0) of LayerNorm(x32[n, t, :]) -> fp32.
//      Every warp sums its tokens in registers, the per-warp sums are added in a fixed order: run-to-run identical.
__global__ void __launch_bounds__(kT)
ln_pool_kernel(const float* __restrict__ x32, const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
               float* __restrict__ feat, int tokens, int t0, int count, int c_real, int c_pad) {
  const long long n = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int kMax = 24;                       // c_real <= 768
  __shared__ float part[kT / 32][768];
  float acc[kMax];
#pragma unroll
  for (int i = 0; i < kMax; ++i) acc[i] = 0.f;
  for (int tt = warp; tt < count; tt += kT / 32) {
    const float* xr = x32 + (n * tokens + t0 + tt) * c_pad;
    float v[kMax];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < kMax; ++i) { const int ch = lane + 32 * i; v[i] = ch < c_real ? xr[ch] : 0.f; s += v[i]; }
    const float mean = warp_sum(s) / (float)c_real;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < kMax; ++i) { const int ch = lane + 32 * i; const float d = ch < c_real ? v[i] - mean : 0.f; q += d * d; }
    const float rstd = rsqrtf(warp_sum(q) / (float)c_real + eps);
#pragma unroll
    for (int i = 0; i < kMax; ++i) {
      const int ch = lane + 32 * i;
      if (ch < c_real) acc[i] += (v[i] - mean) * rstd * __ldg(gamma + ch) + __ldg(beta + ch);
    }
  }
#pragma unroll
  for (int i = 0; i < kMax; ++i) { const int ch = lane + 32 * i; if (ch < c_real) part[warp][ch] = acc[i]; }
  __syncthreads();
  for (int ch = threadIdx.x; ch < c_real; ch += blockDim.x) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < kT / 32; ++w) s += part[w][ch];
    feat[n * c_real + ch] = s / (float)count;
  }
}

}  // namespace
}  // namespace advb200

using namespace advb200;

extern "C" {

int advb200_patch_im2col_bf16(const float* x, const float* mean3, const float* std3, void* out, int n, int h, int w,
                              int patch, int kpad, void* stream) {
  ADV_CHECK_ARG(x && mean3 && std3 && out, "patch_im2col: null pointer");
  ADV_CHECK_ARG(n >= 0 && patch > 0 && h % patch == 0 && w % patch == 0 && kpad >= 3 * patch * patch, "patch_im2col: bad sizes");
  const long long total = (long long)n * (h / patch) * (w / patch) * kpad;
  if (!total) return ADVB200_OK;
  patch_im2col_kernel<<<nblk(total), kT, 0, as_stream(stream)>>>(x, mean3, std3, reinterpret_cast<bf16*>(out), h, w, patch, kpad, total);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

int advb200_vit_tokens_f32(const void* patch_bf16, const float* cls, const float* pos, float* x32, int n, int tokens, int c,
                           void* stream) {
  ADV_CHECK_ARG(patch_bf16 && cls && pos && x32 && n >= 0 && tokens > 1 && c > 0, "vit_tokens: bad argument");
  const long long total = (long long)n * tokens * c;
  if (!total) return ADVB200_OK;
  vit_tokens_kernel<<<nblk(total), kT, 0, as_stream(stream)>>>(reinterpret_cast<const bf16*>(patch_bf16), cls, pos, x32, tokens, c, total);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

int advb200_add_layernorm(float* x32, const void* delta_bf16, const float* gamma, const float* beta, float eps, void* y_bf16,
                          int64_t m, int c_real, int c_pad, int mode, void* stream) {
  ADV_CHECK_ARG(x32 && gamma && beta, "add_layernorm: null pointer");
  ADV_CHECK_ARG(mode >= 0 && mode <= 2 && (mode == 0 || delta_bf16), "add_layernorm: bad mode / missing delta");
  ADV_CHECK_ARG(c_real > 0 && c_real <= 1536 && c_pad >= c_real, "add_layernorm: 1..1536 channels");
  if (m <= 0) return ADVB200_OK;
  if (c_real <= 768)
    add_layernorm_kernel<24><<<nblk(m, kT / 32), kT, 0, as_stream(stream)>>>(x32, reinterpret_cast<const bf16*>(delta_bf16), gamma, beta,
                                                                             eps, reinterpret_cast<bf16*>(y_bf16), m, c_real, c_pad, mode);
  else
    add_layernorm_kernel<48><<<nblk(m, kT / 32), kT, 0, as_stream(stream)>>>(x32, reinterpret_cast<const bf16*>(delta_bf16), gamma, beta,
                                                                             eps, reinterpret_cast<bf16*>(y_bf16), m, c_real, c_pad, mode);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

int advb200_gelu_bf16(const void* x, void* y, int64_t n, void* stream) {
  ADV_CHECK_ARG(x && y && n >= 0 && n % 8 == 0, "gelu: element count must be a multiple of 8");
  ADV_CHECK_ARG(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y)) & 15u) == 0, "gelu: 16-byte alignment");
  if (!n) return ADVB200_OK;
  gelu_kernel<<<nblk(n / 8), kT, 0, as_stream(stream)>>>(reinterpret_cast<const bf16*>(x), reinterpret_cast<bf16*>(y), n / 8);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

int advb200_window_attention_bf16(const void* qkv, void* out, const float* bias, int n_img, int tokens_per_window, int heads, int dim,
                                  int ld_qkv, int ld_out, int q_off, int k_off, int v_off, int grid_h, int grid_w, int window,
                                  int shift, float scale, void* stream) {
  ADV_CHECK_ARG(qkv && out, "window_attention: null pointer");
  ADV_CHECK_ARG(dim > 0 && dim <= 64 && tokens_per_window > 0 && tokens_per_window <= 32 * kAttnMaxKeysPerLane && heads > 0,
                "window_attention: dim <= 64, tokens <= 224");
  AttnParams p;
  p.qkv = reinterpret_cast<const bf16*>(qkv); p.out = reinterpret_cast<bf16*>(out); p.bias = bias;
  p.ld_qkv = ld_qkv; p.ld_out = ld_out; p.q_off = q_off; p.k_off = k_off; p.v_off = v_off;   // absolute column offsets
  p.heads = heads; p.dim = dim; p.T = tokens_per_window;
  p.gh = grid_h; p.gw = grid_w; p.ws = window; p.shift = shift; p.scale = scale;
  if (grid_h > 0) {
    ADV_CHECK_ARG(window > 0 && grid_h % window == 0 && grid_w % window == 0 && tokens_per_window == window * window &&
                      shift >= 0 && shift < window, "window_attention: bad window geometry");
    p.wins_per_img = (grid_h / window) * (grid_w / window);
  } else {
    p.wins_per_img = 1;
  }
  if (n_img <= 0) return ADVB200_OK;
  const int T = p.T, D = p.dim;
  const size_t smem = (size_t)(T * (D + 1) + T * D + kAttnWarps * T) * 4 + (size_t)T * 4;
  ADV_CHECK_CUDA(cudaFuncSetAttribute(window_attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long ctas = (long long)n_img * p.wins_per_img * heads;
  ADV_CHECK_ARG(ctas < (1ll << 31), "window_attention: too many windows for one launch");
  // tensor-core path for the two shapes of the transfer targets (16-byte aligned rows); anything else, or
  // ADVB200_ATTN_SIMT=1, takes the CUDA-core kernel (the cross-check of tests/test_gpu_transformers.py)
  static const bool simt_only = getenv("ADVB200_ATTN_SIMT") != nullptr;
  const bool aligned = ((reinterpret_cast<uintptr_t>(qkv) | reinterpret_cast<uintptr_t>(out)) & 15u) == 0 && ld_qkv % 8 == 0 &&
                       ld_out % 2 == 0 && q_off % 8 == 0 && k_off % 8 == 0 && v_off % 8 == 0;
  if (!simt_only && !g_attn_force_simt && aligned) {
    if (T == 197 && D == 64) return launch_attention_mma<208, 64, 8>(p, ctas, as_stream(stream));
    if (T == 49 && D == 32) return launch_attention_mma<64, 32, 4>(p, ctas, as_stream(stream));
  }
  window_attention_kernel<<<(unsigned)ctas, kAttnWarps * 32, smem, as_stream(stream)>>>(p);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

int advb200_test_attention_simt(int on) {
  // test hook: route advb200_window_attention_bf16 through the CUDA-core kernel (cross-check of the tensor-core one)
  g_attn_force_simt = on != 0;
  return ADVB200_OK;
}

int advb200_patch_merge_gather_f32(const float* x32, float* out, int n, int grid_h, int grid_w, int c_real, int c_pad, void* stream) {
  ADV_CHECK_ARG(x32 && out && grid_h % 2 == 0 && grid_w % 2 == 0 && c_real > 0 && c_pad >= c_real, "patch_merge: bad argument");
  const long long total = (long long)n * (grid_h / 2) * (grid_w / 2) * 4 * c_real;
  if (total <= 0) return ADVB200_OK;
  patch_merge_gather_kernel<<<nblk(total), kT, 0, as_stream(stream)>>>(x32, out, grid_h, grid_w, c_real, c_pad, total);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

int advb200_ln_pool_f32(const float* x32, const float* gamma, const float* beta, float eps, float* feat, int n, int tokens, int t0,
                        int count, int c_real, int c_pad, void* stream) {
  ADV_CHECK_ARG(x32 && gamma && beta && feat, "ln_pool: null pointer");
  ADV_CHECK_ARG(c_real > 0 && c_real <= 768 && c_pad >= c_real && count > 0 && t0 >= 0 && t0 + count <= tokens, "ln_pool: bad sizes");
  if (n <= 0) return ADVB200_OK;
  ln_pool_kernel<<<n, kT, 0, as_stream(stream)>>>(x32, gamma, beta, eps, feat, tokens, t0, count, c_real, c_pad);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

}  // extern "C"

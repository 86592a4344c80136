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
  window_attention_kernel<<<(unsigned)ctas, kAttnWarps * 32, smem, as_stream(stream)>>>(p);
  ADV_LAUNCHED();
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

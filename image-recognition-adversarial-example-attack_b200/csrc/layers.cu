// Non-GEMM layers of the ResNet-50 attack path.  See layers.cuh for the contracts.
#include "layers.cuh"

#include <type_traits>

namespace advb200 {

namespace {

constexpr int kT = 256;
inline int nblocks(long long work) { return (int)((work + kT - 1) / kT); }

template <typename T> __device__ __forceinline__ float to_f(T v);
template <> __device__ __forceinline__ float to_f<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f<bf16>(bf16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f<bf16>(float v) { return __float2bfloat16_rn(v); }

// ---------------------------------------------------------------- input preparation
// Stages the normalised image rows behind kRowsPer16 stem-operand rows hp0.. of image n in shared memory:
// src[(r*2+ph)*3+c][col+4] = norm(x)[c, 2(hp0+r-2)+ph, col], zero outside the image (cols -4..227).
constexpr int kRowsPer16 = 5;
__device__ __forceinline__ void stage_norm_rows(const float* __restrict__ x, const float* __restrict__ mean3,
                                                const float* __restrict__ std3, int n, int hp0, float* src0) {
  // all of this thread's 16-byte image loads are issued before the first one is used (the kernel is latency-bound
  // otherwise): kSegs (row, ph, c) image rows of 56 float4 each
  constexpr int kSegs = kRowsPer16 * 2 * 3, kLoads = (kSegs * 56 + kT - 1) / kT;
  float4 v[kLoads];
#pragma unroll
  for (int k = 0; k < kLoads; ++k) {
    const int idx = threadIdx.x + k * kT;
    v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (idx < kSegs * 56) {
      const int seg = idx / 56, c4 = idx - seg * 56;
      const int c = seg % 3, ph = (seg / 3) & 1, r = seg / 6;
      const int row = 2 * (hp0 + r - 2) + ph;
      if (row >= 0 && row < 224)
        v[k] = __ldg(reinterpret_cast<const float4*>(x + (((long long)n * 3 + c) * 224 + row) * 224) + c4);
    }
  }
  if (threadIdx.x < kSegs * 8) {                     // zero the 4 + 4 pad columns of every row
    const int seg = threadIdx.x >> 3, j = threadIdx.x & 7;
    src0[seg * 232 + (j < 4 ? j : 224 + j)] = 0.f;
  }
#pragma unroll
  for (int k = 0; k < kLoads; ++k) {
    const int idx = threadIdx.x + k * kT;
    if (idx < kSegs * 56) {
      const int seg = idx / 56, c4 = idx - seg * 56;
      const int c = seg % 3;
      const int row = 2 * (hp0 + seg / 6 - 2) + ((seg / 3) & 1);
      float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
      if (row >= 0 && row < 224) {
        const float mu = __ldg(mean3 + c), sd = __ldg(std3 + c);
        o.x = __fdiv_rn(__fsub_rn(v[k].x, mu), sd); o.y = __fdiv_rn(__fsub_rn(v[k].y, mu), sd);   // attack.py:17
        o.z = __fdiv_rn(__fsub_rn(v[k].z, mu), sd); o.w = __fdiv_rn(__fsub_rn(v[k].w, mu), sd);
      }
      *reinterpret_cast<float4*>(&src0[seg * 232 + 4 + c4 * 4]) = o;
    }
  }
}

// one block per kRowsPer16 operand rows (n, hp0..).  An operand pixel (hp, wo) is 64 bf16 = four CONSECUTIVE space-to-depth
// pixels (hp, wo-2 .. wo+1) of 16 slots each (slot q = ph*6 + pw*3 + c holds norm(x)[c, 2(hp-2)+ph, 2*sc+pw]; 12..15 zero), so
// the block first stages its rows as bf16 space-to-depth pixels in shared memory -- normalised and rounded ONCE per image
// element, every image load of a thread in flight before its first use -- and then every 16-byte chunk of the operand is a
// straight 16-byte copy out of that buffer.  (The first version gathered and converted each of the 4.7 output elements per
// image element separately from an fp32 staging buffer: 122 instructions per 16-byte store, issue-active 80 %, 101 us.)
__global__ void __launch_bounds__(kT)
prep_hcol_kernel(const float* __restrict__ x, const float* __restrict__ mean3, const float* __restrict__ std3,
                 bf16* __restrict__ out) {
  constexpr int kCols = 116;                                       // space-to-depth columns -2 .. 113
  __shared__ __align__(16) bf16 s2d[kRowsPer16][kCols][16];
  const int hp0 = (blockIdx.x % (115 / kRowsPer16)) * kRowsPer16;
  const int n = blockIdx.x / (115 / kRowsPer16);
  constexpr int kItems = kRowsPer16 * kCols, kIter = (kItems + kT - 1) / kT;
  float2 v[kIter][2][3];
#pragma unroll
  for (int k = 0; k < kIter; ++k) {
    const int item = threadIdx.x + k * kT;
    const int r = item / kCols, sc = item - r * kCols - 2;
#pragma unroll
    for (int ph = 0; ph < 2; ++ph) {
      const int row = 2 * (hp0 + r - 2) + ph;
      const bool in = item < kItems && sc >= 0 && sc < 112 && row >= 0 && row < 224;
#pragma unroll
      for (int c = 0; c < 3; ++c)
        v[k][ph][c] = in ? __ldg(reinterpret_cast<const float2*>(x + (((long long)n * 3 + c) * 224 + row) * 224 + 2 * sc))
                         : make_float2(0.f, 0.f);
    }
  }
  float mu[3], sd[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) { mu[c] = __ldg(mean3 + c); sd[c] = __ldg(std3 + c); }
#pragma unroll
  for (int k = 0; k < kIter; ++k) {
    const int item = threadIdx.x + k * kT;
    if (item >= kItems) continue;
    const int r = item / kCols, sc = item - r * kCols - 2;
    __align__(16) bf16 q[16];
#pragma unroll
    for (int ph = 0; ph < 2; ++ph) {
      const int row = 2 * (hp0 + r - 2) + ph;
      const bool in = sc >= 0 && sc < 112 && row >= 0 && row < 224;     // the zero padding stays zero (it is NOT normalised)
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float a = in ? __fdiv_rn(__fsub_rn(v[k][ph][c].x, mu[c]), sd[c]) : 0.f;    // attack.py:17
        const float b = in ? __fdiv_rn(__fsub_rn(v[k][ph][c].y, mu[c]), sd[c]) : 0.f;
        q[ph * 6 + c] = __float2bfloat16_rn(a);                         // pw = 0
        q[ph * 6 + 3 + c] = __float2bfloat16_rn(b);                     // pw = 1
      }
    }
#pragma unroll
    for (int i = 12; i < 16; ++i) q[i] = __float2bfloat16_rn(0.f);
    uint4* dst = reinterpret_cast<uint4*>(&s2d[r][sc + 2][0]);
    dst[0] = *reinterpret_cast<const uint4*>(&q[0]);
    dst[1] = *reinterpret_cast<const uint4*>(&q[8]);
  }
  __syncthreads();
  bf16* obase = out + ((long long)n * 115 + hp0) * 112 * 64;
  const uint4* sp = reinterpret_cast<const uint4*>(&s2d[0][0][0]);
  for (int t = threadIdx.x; t < kRowsPer16 * 112 * 8; t += kT) {
    const int r = t / (112 * 8), tt = t - r * (112 * 8);             // tt = wo * 8 + (tap u) * 2 + slot half
    // tap u of operand pixel wo is staged pixel (wo - 2 + u), stored at column index wo + u: chunk (wo + u) * 2 + half
    *reinterpret_cast<uint4*>(obase + (long long)t * 8) = sp[r * (kCols * 2) + 2 * (tt >> 3) + (tt & 7)];
  }
}

// Compact stem operand: [N, 115, 115, 16] bf16, pixel (hp, wp) = the 2x2x3 space-to-depth patch of image rows
// 2(hp-2)+{0,1}, columns 2(wp-2)+{0,1} (12 of 16 slots used), normalised, zero outside the image.  The conv1
// kernel reads 32-byte rows (SWIZZLE_32B) and takes both the vertical and the horizontal taps as row shifts.
// One block = kRowsPer16 operand rows of one image.
__global__ void __launch_bounds__(kT)
prep_hcol16_kernel(const float* __restrict__ x, const float* __restrict__ mean3, const float* __restrict__ std3,
                   bf16* __restrict__ out) {
  __shared__ __align__(16) float src[kRowsPer16][2][3][232];     // [row][ph][c][col + 4]: cols -4..227
  const int hp0 = (blockIdx.x % (115 / kRowsPer16)) * kRowsPer16;
  const int n = blockIdx.x / (115 / kRowsPer16);
  stage_norm_rows(x, mean3, std3, n, hp0, &src[0][0][0][0]);
  __syncthreads();
  bf16* obase = out + ((long long)n * 115 + hp0) * 115 * 16;
  for (int t = threadIdx.x; t < kRowsPer16 * 230; t += kT) {
    const int r = t / 230, j = t - r * 230;
    const int wp = j >> 1, q0 = (j & 1) * 8;
    __align__(16) bf16 v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int q = q0 + i;
      float val = 0.f;
      if (q < 12) {
        const int c = q % 3, pw = (q / 3) & 1, ph = q / 6;
        val = src[r][ph][c][2 * (wp - 2) + pw + 4];
      }
      v[i] = __float2bfloat16_rn(val);
    }
    *reinterpret_cast<uint4*>(obase + (long long)t * 8) = *reinterpret_cast<const uint4*>(v);
  }
}

__global__ void __launch_bounds__(kT)
prep_nhwc_kernel(const float* __restrict__ x, const float* __restrict__ mean3, const float* __restrict__ std3,
                 float* __restrict__ out, int h, int w, long long total) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  const int c = (int)(t % 3);
  long long r = t / 3;
  const int col = (int)(r % w); r /= w;
  const int row = (int)(r % h);
  const long long n = r / h;
  const float px = __ldg(x + ((n * 3 + c) * h + row) * w + col);
  out[t] = __fdiv_rn(__fsub_rn(px, __ldg(mean3 + c)), __ldg(std3 + c));
}

__global__ void __launch_bounds__(kT)
grad_nhwc_to_nchw_kernel(const float* __restrict__ g, const float* __restrict__ std3, float* __restrict__ out,
                         int h, int w, long long total) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;   // NCHW index
  if (t >= total) return;
  const int col = (int)(t % w);
  long long r = t / w;
  const int row = (int)(r % h); r /= h;
  const int c = (int)(r % 3);
  const long long n = r / 3;
  out[t] = __fdiv_rn(__ldg(g + ((n * h + row) * w + col) * 3 + c), __ldg(std3 + c));
}

// ---------------------------------------------------------------- max pool 3x3/2 pad 1
template <typename T>
__global__ void __launch_bounds__(kT)
maxpool_fwd_kernel(const T* __restrict__ in, T* __restrict__ out, uint8_t* __restrict__ idx, int hi, int wi,
                   int ho, int wo, int c, long long total) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  const int ch = (int)(t % c);
  long long r = t / c;
  const int ow = (int)(r % wo); r /= wo;
  const int oh = (int)(r % ho);
  const long long n = r / ho;
  float best = -INFINITY;
  int pos = 0;
#pragma unroll
  for (int dr = 0; dr < 3; ++dr) {
    const int ih = 2 * oh - 1 + dr;
    if (ih < 0 || ih >= hi) continue;
#pragma unroll
    for (int ds = 0; ds < 3; ++ds) {
      const int iw = 2 * ow - 1 + ds;
      if (iw < 0 || iw >= wi) continue;
      const float v = to_f<T>(in[((n * hi + ih) * wi + iw) * c + ch]);
      if (v > best) { best = v; pos = dr * 3 + ds; }   // first maximum wins (torch tie rule)
    }
  }
  out[t] = from_f<T>(best);
  idx[t] = best > 0.f ? (uint8_t)pos : (uint8_t)255;
}

template <typename T>
__global__ void __launch_bounds__(kT)
maxpool_bwd_kernel(const T* __restrict__ gout, const uint8_t* __restrict__ idx, T* __restrict__ gin, int hi,
                   int wi, int ho, int wo, int c, long long total) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  const int ch = (int)(t % c);
  long long r = t / c;
  const int iw = (int)(r % wi); r /= wi;
  const int ih = (int)(r % hi);
  const long long n = r / hi;
  float g = 0.f;
  const int oh_lo = ih >> 1, oh_hi = (ih + 1) >> 1;    // windows whose rows 2oh-1..2oh+1 contain ih
  const int ow_lo = iw >> 1, ow_hi = (iw + 1) >> 1;
  for (int oh = oh_lo; oh <= oh_hi; ++oh) {
    if (oh >= ho) continue;
    const int dr = ih - (2 * oh - 1);
    for (int ow = ow_lo; ow <= ow_hi; ++ow) {
      if (ow >= wo) continue;
      const int ds = iw - (2 * ow - 1);
      const long long o = ((n * ho + oh) * wo + ow) * c + ch;
      if (idx[o] == (uint8_t)(dr * 3 + ds)) g += to_f<T>(gout[o]);
    }
  }
  gin[t] = from_f<T>(g);
}

// ---- bf16, 8 channels (16 bytes) per thread ------------------------------------------------
__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) { f[2 * i] = __uint_as_float(w[i] << 16); f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u); }
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint32_t w[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
    w[i] = *reinterpret_cast<uint32_t*>(&h);
  }
  return make_uint4(w[0], w[1], w[2], w[3]);
}

// Packed-SIMD versions (these kernels are instruction-bound otherwise): values stay bf16x2, the running
// argmax position lives in the matching 16-bit lanes and is selected with the comparison mask.
__device__ __forceinline__ uint32_t hgt2_mask(uint32_t a, uint32_t b) {      // per 16-bit lane: a > b ? 0xffff : 0
  return __hgt2_mask(*reinterpret_cast<const __nv_bfloat162*>(&a), *reinterpret_cast<const __nv_bfloat162*>(&b));
}
__device__ __forceinline__ uint32_t hadd2_u(uint32_t a, uint32_t b) {
  __nv_bfloat162 r = __hadd2(*reinterpret_cast<const __nv_bfloat162*>(&a), *reinterpret_cast<const __nv_bfloat162*>(&b));
  return *reinterpret_cast<uint32_t*>(&r);
}

// kStem: the ResNet-50 stem geometry (112 x 112 x 64 -> 56 x 56) as compile-time constants — the index math of the
// generic version (three 64-bit div/mod per thread) was ~45 % of the instructions of this instruction-bound kernel
template <bool kStem>
__global__ void __launch_bounds__(kT)
maxpool_fwd_bf16x8_kernel(const bf16* __restrict__ in, bf16* __restrict__ out, uint8_t* __restrict__ idx, int hi_,
                          int wi_, int ho_, int wo_, int c8_, long long total) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;   // (n, oh, ow, c8)
  if (t >= total) return;
  const int hi = kStem ? 112 : hi_, wi = kStem ? 112 : wi_, ho = kStem ? 56 : ho_, wo = kStem ? 56 : wo_;
  const int c8 = kStem ? 8 : c8_;
  int cg, ow, oh; long long n;
  if (kStem) {
    const unsigned per_img = 56u * 56u * 8u;
    n = t / per_img;
    const unsigned rem = (unsigned)(t - n * per_img);
    cg = rem & 7; ow = (rem >> 3) % 56u; oh = (rem >> 3) / 56u;
  } else {
    cg = (int)(t % c8);
    long long r = t / c8;
    ow = (int)(r % wo); r /= wo;
    oh = (int)(r % ho);
    n = r / ho;
  }
  uint32_t best[4], pos[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) { best[i] = 0xFF80FF80u; pos[i] = 0; }       // -inf, position 0
#pragma unroll
  for (int dr = 0; dr < 3; ++dr) {
    const int ih = 2 * oh - 1 + dr;
    if (ih < 0 || ih >= hi) continue;
#pragma unroll
    for (int ds = 0; ds < 3; ++ds) {
      const int iw = 2 * ow - 1 + ds;
      if (iw < 0 || iw >= wi) continue;
      const uint4 u = __ldg(reinterpret_cast<const uint4*>(in + (((n * hi + ih) * wi + iw) * c8 + cg) * 8));
      const uint32_t v[4] = {u.x, u.y, u.z, u.w};
      const uint32_t code = (uint32_t)(dr * 3 + ds) * 0x00010001u;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint32_t m = hgt2_mask(v[i], best[i]);            // strict: the first maximum wins (torch tie rule)
        best[i] = (v[i] & m) | (best[i] & ~m);
        pos[i] = (code & m) | (pos[i] & ~m);
      }
    }
  }
  *reinterpret_cast<uint4*>(out + t * 8) = make_uint4(best[0], best[1], best[2], best[3]);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint32_t m = hgt2_mask(best[i], 0u);                  // ReLU-blocked outputs route no gradient: 255
    pos[i] = (pos[i] & m) | (0x00FF00FFu & ~m);
  }
  *reinterpret_cast<uint2*>(idx + t * 8) =
      make_uint2(__byte_perm(pos[0], pos[1], 0x6420), __byte_perm(pos[2], pos[3], 0x6420));
}

// one thread = a 2x2 input patch (rows 2a,2a+1; cols 2b,2b+1) x 8 channels: the four windows that can
// route gradient into the patch, (a..a+1, b..b+1), are each loaded once (all loads issued up front)
template <bool kStem>
__global__ void __launch_bounds__(kT)
maxpool_bwd_bf16x8_kernel(const bf16* __restrict__ gout, const uint8_t* __restrict__ idx, bf16* __restrict__ gin,
                          int hi_, int wi_, int ho_, int wo_, int c8_, long long total) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;   // (n, a, b, c8)
  if (t >= total) return;
  const int hi = kStem ? 112 : hi_, wi = kStem ? 112 : wi_, ho = kStem ? 56 : ho_, wo = kStem ? 56 : wo_;
  const int c8 = kStem ? 8 : c8_;
  int cg, pb, pa; long long n;
  if (kStem) {
    const unsigned per_img = 56u * 56u * 8u;
    n = t / per_img;
    const unsigned rem = (unsigned)(t - n * per_img);
    cg = rem & 7; pb = (rem >> 3) % 56u; pa = (rem >> 3) / 56u;
  } else {
    cg = (int)(t % c8);
    long long r = t / c8;
    pb = (int)(r % (wi / 2)); r /= (wi / 2);
    pa = (int)(r % (hi / 2));
    n = r / (hi / 2);
  }
  uint2 ix[2][2];
  uint4 gv[2][2];
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int oh = pa + i, ow = pb + j;
      if (oh < ho && ow < wo) {
        const long long o = (((n * ho + oh) * wo + ow) * c8 + cg) * 8;
        ix[i][j] = __ldg(reinterpret_cast<const uint2*>(idx + o));
        gv[i][j] = __ldg(reinterpret_cast<const uint4*>(gout + o));
      } else {
        ix[i][j] = make_uint2(0xffffffffu, 0xffffffffu);
        gv[i][j] = make_uint4(0, 0, 0, 0);
      }
    }
#pragma unroll
  for (int dy = 0; dy < 2; ++dy)
#pragma unroll
    for (int dx = 0; dx < 2; ++dx) {
      const int ih = 2 * pa + dy, iw = 2 * pb + dx;
      uint32_t g[4] = {0u, 0u, 0u, 0u};
      // windows containing (ih, iw): row a (dr = 1 + dy) and, for dy = 1, row a+1 (dr = 0); likewise columns.
      // Same visiting order as the scalar kernel; sums of two addends round exactly like its fp32 sum.
#pragma unroll
      for (int i = 0; i <= dy; ++i)
#pragma unroll
        for (int j = 0; j <= dx; ++j) {
          const uint32_t want = (uint32_t)((i ? 0 : 1 + dy) * 3 + (j ? 0 : 1 + dx)) * 0x01010101u;
          const uint32_t mlo = __vcmpeq4(ix[i][j].x, want), mhi = __vcmpeq4(ix[i][j].y, want);
          g[0] = hadd2_u(g[0], gv[i][j].x & __byte_perm(mlo, 0, 0x1100));
          g[1] = hadd2_u(g[1], gv[i][j].y & __byte_perm(mlo, 0, 0x3322));
          g[2] = hadd2_u(g[2], gv[i][j].z & __byte_perm(mhi, 0, 0x1100));
          g[3] = hadd2_u(g[3], gv[i][j].w & __byte_perm(mhi, 0, 0x3322));
        }
      *reinterpret_cast<uint4*>(gin + (((n * hi + ih) * wi + iw) * c8 + cg) * 8) = make_uint4(g[0], g[1], g[2], g[3]);
    }
}

__global__ void __launch_bounds__(kT)
avgpool_bwd_bf16x8_kernel(const float* __restrict__ dfeat, const bf16* __restrict__ act, bf16* __restrict__ gact,
                          int hw, int c8, long long total) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;   // (n, p, c8)
  if (t >= total) return;
  const int cg = (int)(t % c8);
  const long long n = t / ((long long)hw * c8);
  const float4 d0 = __ldg(reinterpret_cast<const float4*>(dfeat + (n * c8 + cg) * 8));
  const float4 d1 = __ldg(reinterpret_cast<const float4*>(dfeat + (n * c8 + cg) * 8) + 1);
  const float d[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
  float a[8], g[8];
  unpack8(__ldg(reinterpret_cast<const uint4*>(act + t * 8)), a);
#pragma unroll
  for (int i = 0; i < 8; ++i) g[i] = a[i] > 0.f ? d[i] / (float)hw : 0.f;
  *reinterpret_cast<uint4*>(gact + t * 8) = pack8(g);
}

// ---------------------------------------------------------------- global average pool
template <typename T>
__global__ void __launch_bounds__(kT)
avgpool_fwd_kernel(const T* __restrict__ in, float* __restrict__ feat, int hw, int c, long long total) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;   // (n, c)
  if (t >= total) return;
  const int ch = (int)(t % c);
  const long long n = t / c;
  float s = 0.f;
  for (int p = 0; p < hw; ++p) s += to_f<T>(in[(n * hw + p) * c + ch]);
  feat[t] = s / (float)hw;
}

template <typename T>
__global__ void __launch_bounds__(kT)
avgpool_bwd_kernel(const float* __restrict__ dfeat, const T* __restrict__ act, T* __restrict__ gact, int hw,
                   int c, long long total) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;   // (n, p, c)
  if (t >= total) return;
  const int ch = (int)(t % c);
  const long long n = t / ((long long)hw * c);
  const float g = __ldg(dfeat + n * c + ch) / (float)hw;
  gact[t] = from_f<T>(to_f<T>(act[t]) > 0.f ? g : 0.f);
}

// ---------------------------------------------------------------- losses
// one block per image; classes <= 4096
__global__ void __launch_bounds__(kT)
loss_grad_kernel(const float* __restrict__ logits, const int64_t* __restrict__ y, int classes, int loss_kind,
                 float cw_c, float kappa, float scale, float* __restrict__ dlogits,
                 int64_t* __restrict__ preds, float* __restrict__ loss, uint8_t* __restrict__ success) {
  __shared__ float s_val[kT];
  __shared__ int s_idx[kT];
  __shared__ float s_red[kT / 32];
  __shared__ float s_bcast[4];
  __shared__ int s_ibcast[2];
  const int n = blockIdx.x, tid = threadIdx.x;
  const float* z = logits + (long long)n * classes;
  const int yy = (int)y[n];

  // argmax (first maximum) and maximum over j != y
  float bv = -INFINITY; int bi = 0x7fffffff;
  float ov = -INFINITY;
  for (int j = tid; j < classes; j += kT) {
    const float v = z[j];
    if (v > bv) { bv = v; bi = j; }
    if (j != yy && v > ov) ov = v;
  }
  s_val[tid] = bv; s_idx[tid] = bi;
  __syncthreads();
  for (int o = kT / 2; o > 0; o >>= 1) {
    if (tid < o) {
      const float v2 = s_val[tid + o]; const int i2 = s_idx[tid + o];
      if (v2 > s_val[tid] || (v2 == s_val[tid] && i2 < s_idx[tid])) { s_val[tid] = v2; s_idx[tid] = i2; }
    }
    __syncthreads();
  }
  const float zmax = s_val[0];
  const int amax = s_idx[0];
  __syncthreads();
  s_val[tid] = ov;
  __syncthreads();
  for (int o = kT / 2; o > 0; o >>= 1) {
    if (tid < o) s_val[tid] = fmaxf(s_val[tid], s_val[tid + o]);
    __syncthreads();
  }
  const float other_raw = s_val[0];
  __syncthreads();
  if (tid == 0 && preds) preds[n] = amax;

  if (yy < 0 || yy >= classes) {
    // a label outside [0, classes) raises in the reference (F.cross_entropy / one_hot); the Python mirror checks labels
    // once per attack call — a raw C-ABI caller gets a NaN row instead of an out-of-bounds read
    const float qnan = __int_as_float(0x7fc00000);
    if (dlogits)
      for (int j = tid; j < classes; j += kT) dlogits[(long long)n * classes + j] = qnan;
    if (tid == 0) {
      if (loss) loss[n] = qnan;
      if (success) success[n] = 0;
    }
    return;
  }

  if (loss_kind == ADVB200_LOSS_CE_MEAN) {
    // softmax cross-entropy: d z_j = (p_j - [j==y]) * scale, with the true-class entry formed as
    // -(sum_{j!=y} p_j) to avoid the cancellation in p_y - 1
    float so = 0.f;
    for (int j = tid; j < classes; j += kT)
      if (j != yy) so += expf(z[j] - zmax);
    const float tot_o = block_sum<kT>(so, s_red);
    if (tid == 0) {
      const float ey = expf(z[yy] - zmax);
      s_bcast[0] = tot_o;
      s_bcast[1] = tot_o + ey;
    }
    __syncthreads();
    const float sum_other = s_bcast[0], sum_all = s_bcast[1];
    if (dlogits)
      for (int j = tid; j < classes; j += kT) {
        const float pj = expf(z[j] - zmax) / sum_all;
        dlogits[(long long)n * classes + j] = (j == yy ? -(sum_other / sum_all) : pj) * scale;
      }
    if (tid == 0) {
      if (loss) loss[n] = logf(sum_all) - (z[yy] - zmax);
      if (success) success[n] = amax != yy;
    }
  } else {
    // CW margin (attack.py:114-124).  other = amax_j (z_j - 1e4*[j==y]); ties share the gradient.
    const float real = z[yy];
    const float other = fmaxf(other_raw, real - 1e4f);
    int ties = 0;
    for (int j = tid; j < classes; j += kT) {
      const float v = (j == yy) ? real - 1e4f : z[j];
      if (v == other) ++ties;
    }
    s_idx[tid] = ties;
    __syncthreads();
    for (int o = kT / 2; o > 0; o >>= 1) {
      if (tid < o) s_idx[tid] += s_idx[tid + o];
      __syncthreads();
    }
    const int nties = s_idx[0];
    const bool targeted = loss_kind == ADVB200_LOSS_CW_TGT;
    const float arg = targeted ? (other - real + kappa) : (real - other + kappa);
    const bool active = arg >= 0.f;               // clamp(min=0) passes the gradient at exactly 0
    const float sgn = targeted ? -1.f : 1.f;      // d f / d real
    if (dlogits)
      for (int j = tid; j < classes; j += kT) {
        float g = 0.f;
        if (active) {
          const float v = (j == yy) ? real - 1e4f : z[j];
          if (j == yy) g += sgn * cw_c;
          if (v == other) g += -sgn * cw_c / (float)nties;
        }
        dlogits[(long long)n * classes + j] = g * scale;
      }
    if (tid == 0) {
      if (loss) loss[n] = fmaxf(arg, 0.f);
      if (success) success[n] = targeted ? (amax == yy) : (amax != yy);
    }
  }
}

__global__ void __launch_bounds__(kT)
argmax_kernel(const float* __restrict__ logits, int classes, int64_t* __restrict__ preds) {
  __shared__ float s_val[kT];
  __shared__ int s_idx[kT];
  const int n = blockIdx.x, tid = threadIdx.x;
  const float* z = logits + (long long)n * classes;
  float bv = -INFINITY; int bi = 0x7fffffff;
  for (int j = tid; j < classes; j += kT) {
    const float v = z[j];
    if (v > bv) { bv = v; bi = j; }
  }
  s_val[tid] = bv; s_idx[tid] = bi;
  __syncthreads();
  for (int o = kT / 2; o > 0; o >>= 1) {
    if (tid < o) {
      const float v2 = s_val[tid + o]; const int i2 = s_idx[tid + o];
      if (v2 > s_val[tid] || (v2 == s_val[tid] && i2 < s_idx[tid])) { s_val[tid] = v2; s_idx[tid] = i2; }
    }
    __syncthreads();
  }
  if (tid == 0) preds[n] = s_idx[0];
}

// ---------------------------------------------------------------- detector score
template <typename T>
__global__ void __launch_bounds__(kT)
detector_score_kernel(const T* __restrict__ act, float* __restrict__ score, int hw, int c) {
  __shared__ float s_red[kT / 32];
  __shared__ float s_keep;
  const long long n = blockIdx.x;
  const T* a = act + n * (long long)hw * c;
  float sq = 0.f, var_sum = 0.f;
  for (int ch = threadIdx.x; ch < c; ch += kT) {
    float s = 0.f;
    for (int p = 0; p < hw; ++p) s += to_f<T>(a[(long long)p * c + ch]);
    const float mean = s / (float)hw;
    float d2 = 0.f;
    for (int p = 0; p < hw; ++p) {
      const float v = to_f<T>(a[(long long)p * c + ch]);
      sq += v * v;
      const float d = v - mean;
      d2 += d * d;
    }
    var_sum += d2 / (float)(hw - 1);                       // unbiased (torch.var default)
  }
  const float tot_sq = block_sum<kT>(sq, s_red);
  if (threadIdx.x == 0) s_keep = tot_sq;
  __syncthreads();
  const float tot_var = block_sum<kT>(var_sum, s_red);
  if (threadIdx.x == 0) {
    const float l2 = sqrtf(s_keep) / (float)c;             // defense_experiments.py:245
    const float sc = l2 + (tot_var / (float)c) * 0.1f;     // :249-252
    score[n] = fminf(fmaxf(sc, 0.f), 100.f);               // :255
  }
}

// ---------------------------------------------------------------- weight packing
__global__ void bn_fold_kernel(const float* gamma, const float* beta, const float* mean, const float* var,
                               float eps, float* scale, float* shift, int c) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= c) return;
  const float s = gamma[i] / sqrtf(var[i] + eps);
  scale[i] = s;
  shift[i] = beta[i] - mean[i] * s;
}

template <typename T>
__global__ void pack_kernel(const float* __restrict__ w, const float* __restrict__ scale, T* __restrict__ fwd,
                            T* __restrict__ dgrad, int cout, int cin, int k, int f32_layout, long long total) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;   // index into w [co,ci,r,s]
  if (t >= total) return;
  const int kk = k * k;
  const int rs = (int)(t % kk);
  const int ci = (int)((t / kk) % cin);
  const int co = (int)(t / ((long long)kk * cin));
  const float v = w[t] * (scale ? scale[co] : 1.f);
  if (f32_layout) {
    if (fwd) fwd[((long long)rs * cin + ci) * cout + co] = from_f<T>(v);
    if (dgrad) dgrad[((long long)rs * cout + co) * cin + ci] = from_f<T>(v);
  } else {
    if (fwd) fwd[((long long)rs * cout + co) * cin + ci] = from_f<T>(v);
    if (dgrad) dgrad[((long long)rs * cin + ci) * cout + co] = from_f<T>(v);
  }
}

__global__ void pack_conv1_kernel(const float* __restrict__ w, const float* __restrict__ scale,
                                  bf16* __restrict__ fwd, bf16* __restrict__ dgrad, bf16* __restrict__ fwd16) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;    // (tt, u, q, co) : 4*4*16*64
  if (t >= 4 * 4 * 16 * 64) return;
  const int co = t & 63, q = (t >> 6) & 15, u = (t >> 10) & 3, tt = t >> 12;
  float v = 0.f;
  if (q < 12) {
    const int c = q % 3, pw = (q / 3) & 1, ph = q / 6;
    const int r = 2 * tt + ph - 1, s = 2 * u + pw - 1;
    if (r >= 0 && s >= 0) v = w[((co * 3 + c) * 7 + r) * 7 + s] * scale[co];
  }
  const bf16 b = __float2bfloat16_rn(v);
  fwd[(tt * 64 + co) * 64 + u * 16 + q] = b;
  if (fwd16) fwd16[((tt * 4 + u) * 64 + co) * 16 + q] = b;       // 32-byte rows: one (vertical, horizontal) tap each
  // input gradient: vertical taps (tt) stay in the K loop, horizontal taps (u) go to the N dimension:
  // row tt*64 + qi*4 + u, qi = slot of q in the store-friendly order {0,3,1,4,2,5,6,9 | 7,10,8,11}
  // (pairs (pw=0, pw=1) of the same (c, ph) adjacent, so the epilogue writes 8-byte pixel pairs)
  const int slot_of_q[16] = {0, 2, 4, 1, 3, 5, 6, 8, 10, 7, 9, 11, 12, 13, 14, 15};
  dgrad[(tt * 64 + slot_of_q[q] * 4 + u) * 64 + co] = b;
}

__global__ void transpose_kernel(const float* __restrict__ in, float* __restrict__ out, int rows, int cols) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long long)rows * cols) return;
  const int r = (int)(t / cols), c = (int)(t % cols);
  out[(long long)c * rows + r] = in[t];
}

__global__ void f32_to_bf16_kernel(const float* __restrict__ in, bf16* __restrict__ out, long long n) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) out[t] = __float2bfloat16_rn(in[t]);
}

// bf16 tensor [pixels, c] -> ReLU bit mask (bit set <=> value > 0); word ((c/64)*pixels + pixel)*2 + (c/32)%2
__global__ void __launch_bounds__(kT)
pack_mask_bits_kernel(const bf16* __restrict__ act, uint32_t* __restrict__ bits, long long pixels, int c) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;   // (pixel, c32)
  const int c32n = c / 32;
  if (t >= pixels * c32n) return;
  const int c32 = (int)(t % c32n);
  const long long px = t / c32n;
  const bf16* a = act + px * c + c32 * 32;
  uint32_t w = 0;
#pragma unroll
  for (int i = 0; i < 32; ++i) w |= (__bfloat162float(a[i]) > 0.f ? 1u : 0u) << i;
  bits[((long long)(c32 >> 1) * pixels + px) * 2 + (c32 & 1)] = w;
}


// ---- helpers of the forward-only transfer target (VGG-19, blackbox_transfer.py:25-54) -----------------------------
// normalise fp32 NCHW [N,3,H,W] -> bf16 NHWC [N,H,W,64]: channels 0..2 carry the image, 3..63 are zero so the first
// 3x3 conv runs through the same 64-channel tensor-core path as the rest
__global__ void prep_nhwc64_kernel(const float* __restrict__ x, const float* __restrict__ mean3,
                                   const float* __restrict__ std3, bf16* __restrict__ out, int h, int w, long long pixels) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;     // (pixel, 16-byte group 0..7)
  if (t >= pixels * 8) return;
  const long long px = t >> 3;
  const int grp = (int)(t & 7);
  uint4 v = make_uint4(0u, 0u, 0u, 0u);
  if (grp == 0) {
    const long long hw = (long long)h * w;
    const long long n = px / hw, r = px % hw;
    float c[3];
#pragma unroll
    for (int k = 0; k < 3; ++k)
      c[k] = __fdiv_rn(__fsub_rn(__ldg(x + (n * 3 + k) * hw + r), __ldg(mean3 + k)), __ldg(std3 + k));   // attack.py:17
    const __nv_bfloat162 a = __floats2bfloat162_rn(c[0], c[1]);
    const __nv_bfloat162 b = __floats2bfloat162_rn(c[2], 0.f);
    v.x = *reinterpret_cast<const uint32_t*>(&a);
    v.y = *reinterpret_cast<const uint32_t*>(&b);
  }
  reinterpret_cast<uint4*>(out)[t] = v;
}

// 2x2 / stride 2 max pool, bf16 NHWC, 8 channels per thread
__global__ void maxpool2x2_bf16x8_kernel(const bf16* __restrict__ in, bf16* __restrict__ out, int hi, int wi, int c8,
                                         long long total) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;     // (n, ho, wo, c8)
  if (t >= total) return;
  const int ho_n = hi / 2, wo_n = wi / 2;
  const int cg = (int)(t % c8);
  const int wo = (int)((t / c8) % wo_n);
  const int ho = (int)((t / ((long long)c8 * wo_n)) % ho_n);
  const long long n = t / ((long long)c8 * wo_n * ho_n);
  const uint4* src = reinterpret_cast<const uint4*>(in);
  auto at = [&](int r, int q) { return __ldg(src + (((n * hi + 2 * ho + r) * wi + 2 * wo + q) * (long long)c8 + cg)); };
  const uint4 a = at(0, 0), b = at(0, 1), c = at(1, 0), d = at(1, 1);
  auto mx = [](uint32_t p, uint32_t q) {
    const __nv_bfloat162 r = __hmax2(*reinterpret_cast<const __nv_bfloat162*>(&p), *reinterpret_cast<const __nv_bfloat162*>(&q));
    return *reinterpret_cast<const uint32_t*>(&r);
  };
  uint4 o;
  o.x = mx(mx(a.x, b.x), mx(c.x, d.x)); o.y = mx(mx(a.y, b.y), mx(c.y, d.y));
  o.z = mx(mx(a.z, b.z), mx(c.z, d.z)); o.w = mx(mx(a.w, b.w), mx(c.w, d.w));
  reinterpret_cast<uint4*>(out)[t] = o;
}

// classifier weights: fp32 [cout, C*HW] in torch's flatten order (c*HW + hw) -> bf16 [cout_pad, HW*C] in NHWC order
// (hw*C + c); rows >= cout are zero (1000 classes padded to 1024 output channels)
__global__ void pack_fc_bf16_kernel(const float* __restrict__ w, bf16* __restrict__ out, int cout, int cout_pad, int c,
                                    int hw) {
  const long long k = (long long)c * hw;
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long long)cout_pad * k) return;
  const long long co = t / k, kk = t % k;
  const int pos = (int)(kk / c), ch = (int)(kk % c);
  out[t] = __float2bfloat16_rn(co < cout ? __ldg(w + co * k + (long long)ch * hw + pos) : 0.f);
}

__global__ void bf16_cols_to_f32_kernel(const bf16* __restrict__ in, int ld_in, float* __restrict__ out, int cols,
                                        long long total) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  out[t] = __bfloat162float(in[(t / cols) * ld_in + t % cols]);
}

__global__ void pad_vec_kernel(const float* __restrict__ in, float* __restrict__ out, int n, int n_pad) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n_pad) out[t] = t < n ? in[t] : 0.f;
}

// first VGG conv: fp32 [cout, 3, k, k] -> fp32 [cout, 64, k, k] with zero channels 3..63
__global__ void pad_cin_kernel(const float* __restrict__ w, float* __restrict__ out, int cout, int cin, int cin_pad, int kk) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long long)cout * cin_pad * kk) return;
  const int rs = (int)(t % kk);
  const int ci = (int)((t / kk) % cin_pad);
  const long long co = t / ((long long)kk * cin_pad);
  out[t] = ci < cin ? w[(co * cin + ci) * kk + rs] : 0.f;
}

}  // namespace

__global__ void vec_add_kernel(const float* a, const float* b, float* out, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = a[i] + b[i];
}
int launch_vec_add(const float* a, const float* b, float* out, int n, cudaStream_t st) {
  vec_add_kernel<<<(n + 255) / 256, 256, 0, st>>>(a, b, out, n);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

int launch_pack_mask_bits(const bf16* act, uint32_t* bits, long long pixels, int c, cudaStream_t st) {
  const long long total = pixels * (c / 32);
  if (!total) return ADVB200_OK;
  pack_mask_bits_kernel<<<nblocks(total), kT, 0, st>>>(act, bits, pixels, c);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

int launch_prep_hcol(const float* x, const float* mean3, const float* std3, bf16* out, int n, cudaStream_t st) {
  if (!n) return ADVB200_OK;
  prep_hcol_kernel<<<n * (115 / kRowsPer16), kT, 0, st>>>(x, mean3, std3, out);
  ADV_LAUNCHED();
  return ADVB200_OK;
}
int launch_prep_hcol16(const float* x, const float* mean3, const float* std3, bf16* out, int n, cudaStream_t st) {
  if (!n) return ADVB200_OK;
  prep_hcol16_kernel<<<n * (115 / kRowsPer16), kT, 0, st>>>(x, mean3, std3, out);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

int launch_prep_nhwc(const float* x, const float* mean3, const float* std3, float* out, int n, int h, int w,
                     cudaStream_t st) {
  const long long total = (long long)n * h * w * 3;
  if (!total) return ADVB200_OK;
  prep_nhwc_kernel<<<nblocks(total), kT, 0, st>>>(x, mean3, std3, out, h, w, total);
  ADV_LAUNCHED();
  return ADVB200_OK;
}
int launch_grad_nhwc_to_nchw(const float* g, const float* std3, float* out, int n, int h, int w,
                             cudaStream_t st) {
  const long long total = (long long)n * h * w * 3;
  if (!total) return ADVB200_OK;
  grad_nhwc_to_nchw_kernel<<<nblocks(total), kT, 0, st>>>(g, std3, out, h, w, total);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

template <typename T>
int launch_maxpool_fwd(const T* in, T* out, uint8_t* idx, int n, int hi, int wi, int c, cudaStream_t st) {
  const int ho = (hi + 2 - 3) / 2 + 1, wo = (wi + 2 - 3) / 2 + 1;
  const long long total = (long long)n * ho * wo * c;
  if (!total) return ADVB200_OK;
  if constexpr (std::is_same<T, bf16>::value) {
    if (c % 8 == 0) {
      if (hi == 112 && wi == 112 && c == 64)
        maxpool_fwd_bf16x8_kernel<true><<<nblocks(total / 8), kT, 0, st>>>(in, out, idx, hi, wi, ho, wo, c / 8, total / 8);
      else
        maxpool_fwd_bf16x8_kernel<false><<<nblocks(total / 8), kT, 0, st>>>(in, out, idx, hi, wi, ho, wo, c / 8, total / 8);
      ADV_LAUNCHED();
      return ADVB200_OK;
    }
  }
  maxpool_fwd_kernel<T><<<nblocks(total), kT, 0, st>>>(in, out, idx, hi, wi, ho, wo, c, total);
  ADV_LAUNCHED();
  return ADVB200_OK;
}
template <typename T>
int launch_maxpool_bwd(const T* gout, const uint8_t* idx, T* gin, int n, int hi, int wi, int c,
                       cudaStream_t st) {
  const int ho = (hi + 2 - 3) / 2 + 1, wo = (wi + 2 - 3) / 2 + 1;
  const long long total = (long long)n * hi * wi * c;
  if (!total) return ADVB200_OK;
  if constexpr (std::is_same<T, bf16>::value) {
    if (c % 8 == 0 && hi % 2 == 0 && wi % 2 == 0) {
      if (hi == 112 && wi == 112 && c == 64)
        maxpool_bwd_bf16x8_kernel<true><<<nblocks(total / 32), kT, 0, st>>>(gout, idx, gin, hi, wi, ho, wo, c / 8, total / 32);
      else
        maxpool_bwd_bf16x8_kernel<false><<<nblocks(total / 32), kT, 0, st>>>(gout, idx, gin, hi, wi, ho, wo, c / 8, total / 32);
      ADV_LAUNCHED();
      return ADVB200_OK;
    }
  }
  maxpool_bwd_kernel<T><<<nblocks(total), kT, 0, st>>>(gout, idx, gin, hi, wi, ho, wo, c, total);
  ADV_LAUNCHED();
  return ADVB200_OK;
}
template <typename T>
int launch_avgpool_fwd(const T* in, float* feat, int n, int hw, int c, cudaStream_t st) {
  const long long total = (long long)n * c;
  if (!total) return ADVB200_OK;
  avgpool_fwd_kernel<T><<<nblocks(total), kT, 0, st>>>(in, feat, hw, c, total);
  ADV_LAUNCHED();
  return ADVB200_OK;
}
template <typename T>
int launch_avgpool_bwd(const float* dfeat, const T* act, T* gact, int n, int hw, int c, cudaStream_t st) {
  const long long total = (long long)n * hw * c;
  if (!total) return ADVB200_OK;
  if constexpr (std::is_same<T, bf16>::value) {
    if (c % 8 == 0) {
      avgpool_bwd_bf16x8_kernel<<<nblocks(total / 8), kT, 0, st>>>(dfeat, act, gact, hw, c / 8, total / 8);
      ADV_LAUNCHED();
      return ADVB200_OK;
    }
  }
  avgpool_bwd_kernel<T><<<nblocks(total), kT, 0, st>>>(dfeat, act, gact, hw, c, total);
  ADV_LAUNCHED();
  return ADVB200_OK;
}
template <typename T>
int launch_detector_score(const T* act, float* score, int n, int hw, int c, cudaStream_t st) {
  if (!n) return ADVB200_OK;
  detector_score_kernel<T><<<n, kT, 0, st>>>(act, score, hw, c);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

#define INSTANTIATE(T)                                                                                   \
  template int launch_maxpool_fwd<T>(const T*, T*, uint8_t*, int, int, int, int, cudaStream_t);          \
  template int launch_maxpool_bwd<T>(const T*, const uint8_t*, T*, int, int, int, int, cudaStream_t);    \
  template int launch_avgpool_fwd<T>(const T*, float*, int, int, int, cudaStream_t);                     \
  template int launch_avgpool_bwd<T>(const float*, const T*, T*, int, int, int, cudaStream_t);           \
  template int launch_detector_score<T>(const T*, float*, int, int, int, cudaStream_t);
INSTANTIATE(float)
INSTANTIATE(bf16)

int launch_loss_grad(const float* logits, const int64_t* y, int n, int classes, int loss_kind, float c,
                     float kappa, float scale, float* dlogits, int64_t* preds, float* loss, uint8_t* success,
                     cudaStream_t st) {
  if (!n) return ADVB200_OK;
  loss_grad_kernel<<<n, kT, 0, st>>>(logits, y, classes, loss_kind, c, kappa, scale, dlogits, preds, loss,
                                     success);
  ADV_LAUNCHED();
  return ADVB200_OK;
}
int launch_argmax(const float* logits, int n, int classes, int64_t* preds, cudaStream_t st) {
  if (!n) return ADVB200_OK;
  argmax_kernel<<<n, kT, 0, st>>>(logits, classes, preds);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

int launch_bn_fold(const float* gamma, const float* beta, const float* mean, const float* var, float eps,
                   float* scale, float* shift, int c, cudaStream_t st) {
  bn_fold_kernel<<<(c + 255) / 256, 256, 0, st>>>(gamma, beta, mean, var, eps, scale, shift, c);
  ADV_LAUNCHED();
  return ADVB200_OK;
}
int launch_pack_f32(const float* w, const float* scale, float* fwd, float* dgrad, int cout, int cin, int k,
                    cudaStream_t st) {
  const long long total = (long long)cout * cin * k * k;
  pack_kernel<float><<<nblocks(total), kT, 0, st>>>(w, scale, fwd, dgrad, cout, cin, k, 1, total);
  ADV_LAUNCHED();
  return ADVB200_OK;
}
int launch_pack_bf16(const float* w, const float* scale, bf16* fwd, bf16* dgrad, int cout, int cin, int k,
                     cudaStream_t st) {
  const long long total = (long long)cout * cin * k * k;
  pack_kernel<bf16><<<nblocks(total), kT, 0, st>>>(w, scale, fwd, dgrad, cout, cin, k, 0, total);
  ADV_LAUNCHED();
  return ADVB200_OK;
}
int launch_pack_conv1_bf16(const float* w, const float* scale, bf16* fwd, bf16* dgrad, bf16* fwd16, cudaStream_t st) {
  pack_conv1_kernel<<<(4 * 4 * 16 * 64 + 255) / 256, 256, 0, st>>>(w, scale, fwd, dgrad, fwd16);
  ADV_LAUNCHED();
  return ADVB200_OK;
}
int launch_transpose_f32(const float* in, float* out, int rows, int cols, cudaStream_t st) {
  const long long total = (long long)rows * cols;
  transpose_kernel<<<nblocks(total), kT, 0, st>>>(in, out, rows, cols);
  ADV_LAUNCHED();
  return ADVB200_OK;
}
int launch_f32_to_bf16(const float* in, bf16* out, int64_t n, cudaStream_t st) {
  if (!n) return ADVB200_OK;
  f32_to_bf16_kernel<<<nblocks(n), kT, 0, st>>>(in, out, n);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

int launch_prep_nhwc64_bf16(const float* x, const float* mean3, const float* std3, bf16* out, int n, int h, int w,
                            cudaStream_t st) {
  const long long pixels = (long long)n * h * w;
  prep_nhwc64_kernel<<<nblocks(pixels * 8), kT, 0, st>>>(x, mean3, std3, out, h, w, pixels);
  ADV_LAUNCHED();
  return ADVB200_OK;
}
int launch_maxpool2x2_bf16(const bf16* in, bf16* out, int n, int hi, int wi, int c, cudaStream_t st) {
  ADV_CHECK_ARG(c % 8 == 0 && hi % 2 == 0 && wi % 2 == 0, "maxpool2x2: needs even sizes and c % 8 == 0");
  const long long total = (long long)n * (hi / 2) * (wi / 2) * (c / 8);
  maxpool2x2_bf16x8_kernel<<<nblocks(total), kT, 0, st>>>(in, out, hi, wi, c / 8, total);
  ADV_LAUNCHED();
  return ADVB200_OK;
}
int launch_pack_fc_bf16(const float* w, bf16* out, int cout, int cout_pad, int c, int hw, cudaStream_t st) {
  pack_fc_bf16_kernel<<<nblocks((long long)cout_pad * c * hw), kT, 0, st>>>(w, out, cout, cout_pad, c, hw);
  ADV_LAUNCHED();
  return ADVB200_OK;
}
int launch_bf16_cols_to_f32(const bf16* in, int ld_in, float* out, int rows, int cols, cudaStream_t st) {
  const long long total = (long long)rows * cols;
  bf16_cols_to_f32_kernel<<<nblocks(total), kT, 0, st>>>(in, ld_in, out, cols, total);
  ADV_LAUNCHED();
  return ADVB200_OK;
}
int launch_pad_vec(const float* in, float* out, int n, int n_pad, cudaStream_t st) {
  pad_vec_kernel<<<nblocks(n_pad), kT, 0, st>>>(in, out, n, n_pad);
  ADV_LAUNCHED();
  return ADVB200_OK;
}
int launch_pad_cin(const float* w, float* out, int cout, int cin, int cin_pad, int k, cudaStream_t st) {
  pad_cin_kernel<<<nblocks((long long)cout * cin_pad * k * k), kT, 0, st>>>(w, out, cout, cin, cin_pad, k * k);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

}  // namespace advb200

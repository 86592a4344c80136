// HBM-bound update / defense kernels of the attack hot path.
// All fp32, 128-bit vectorised, grid sized in multiples of the SM count, no fast-math:
// the FGSM/PGD step, quantisation and smoothing are bit-exact against the reference's torch ops.
#include "common.cuh"

namespace advb200 {

namespace {

constexpr int kThreads = 256;
constexpr int kBlocksPerSM = 8;

inline int grid_for(int64_t work_items) {
  int64_t need = (work_items + kThreads - 1) / kThreads;
  int64_t cap = (int64_t)kNumSMs * kBlocksPerSM;
  if (need < 1) need = 1;
  return (int)(need < cap ? need : cap);
}

__device__ __forceinline__ float sign_like_torch(float g) {
  // torch.sign: (0 < g) - (g < 0); NaN and +-0 map to 0
  return (float)((g > 0.f) - (g < 0.f));
}

// attack.py:61-63 (project=1) / attack.py:33-34 (project=0)
__device__ __forceinline__ float linf_one(float x, float x0, float g, float a, float e, bool project) {
  float x1 = __fadd_rn(x, __fmul_rn(a, sign_like_torch(g)));
  if (project) {
    float hi = __fadd_rn(x0, e);
    float lo = __fsub_rn(x0, e);
    x1 = fmaxf(fminf(x1, hi), lo);
  }
  return fminf(fmaxf(x1, 0.f), 1.f);
}

template <bool kProject>
__global__ void __launch_bounds__(kThreads)
linf_step_kernel(float* __restrict__ x_adv, const float* __restrict__ x_orig,
                 const float* __restrict__ grad, float alpha, float eps, int64_t n4, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  float4* xa = reinterpret_cast<float4*>(x_adv);
  const float4* xo = reinterpret_cast<const float4*>(x_orig);
  const float4* gr = reinterpret_cast<const float4*>(grad);
  for (; i < n4; i += stride) {
    float4 a = xa[i];
    float4 g = __ldg(gr + i);
    float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
    if (kProject) o = __ldg(xo + i);
    a.x = linf_one(a.x, o.x, g.x, alpha, eps, kProject);
    a.y = linf_one(a.y, o.y, g.y, alpha, eps, kProject);
    a.z = linf_one(a.z, o.z, g.z, alpha, eps, kProject);
    a.w = linf_one(a.w, o.w, g.w, alpha, eps, kProject);
    xa[i] = a;
  }
  // scalar tail (n not a multiple of 4)
  for (int64_t j = n4 * 4 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += stride)
    x_adv[j] = linf_one(x_adv[j], kProject ? x_orig[j] : 0.f, grad[j], alpha, eps, kProject);
}

template <bool kProject>
__global__ void __launch_bounds__(kThreads)
linf_step_scalar_kernel(float* __restrict__ x_adv, const float* __restrict__ x_orig,
                        const float* __restrict__ grad, float alpha, float eps, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += stride)
    x_adv[j] = linf_one(x_adv[j], kProject ? x_orig[j] : 0.f, grad[j], alpha, eps, kProject);
}

__global__ void __launch_bounds__(kThreads)
random_start_kernel(float* __restrict__ x_adv, const float* __restrict__ x_orig,
                    const float* __restrict__ noise, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += stride)
    x_adv[j] = fminf(fmaxf(__fadd_rn(x_orig[j], noise[j]), 0.f), 1.f);
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ------------------------------------------------------------------------------------------
// Defenses
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float quant_one(float x, float lm1) {
  float c = fminf(fmaxf(x, 0.f), 1.f);
  return __fdiv_rn(rintf(__fmul_rn(c, lm1)), lm1);  // round-half-even, true division
}

// The quotient k / (levels-1) takes only `levels` values: a shared-memory table filled with the same true division
// replaces the ~12-instruction IEEE division sequence per element by one conversion and one LDS (the kernels were
// issue-bound on it).  Bit-identical by construction; levels > kQuantTab falls back to dividing.
constexpr int kQuantTab = 256;
__device__ __forceinline__ void quant_table_init(float* tab, float lm1) {
  for (int k = threadIdx.x; k < kQuantTab; k += blockDim.x) tab[k] = __fdiv_rn((float)k, lm1);
  __syncthreads();
}
__device__ __forceinline__ float quant_one_tab(float x, float lm1, const float* tab, bool use_tab) {
  const float c = fminf(fmaxf(x, 0.f), 1.f);
  if (!use_tab) return __fdiv_rn(rintf(__fmul_rn(c, lm1)), lm1);
  return tab[__float2int_rn(__fmul_rn(c, lm1))];     // cvt.rni = rintf (half to even); the product is in [0, lm1]
}

__global__ void __launch_bounds__(kThreads)
quantize_kernel(const float* __restrict__ x, float* __restrict__ out, int64_t n4, int64_t n, float lm1,
                int vec) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  __shared__ float tab[kQuantTab];
  const bool use_tab = lm1 < (float)kQuantTab;
  quant_table_init(tab, lm1);
  if (vec) {
    const float4* xi = reinterpret_cast<const float4*>(x);
    float4* oo = reinterpret_cast<float4*>(out);
    // four independent 16-byte loads in flight per thread: one load per iteration leaves ~32 KB per SM in flight,
    // about half of what the HBM latency needs (measured 3.6 TB/s; the 3-operand linf step reaches 6.6 TB/s)
    for (; i + 3 * stride < n4; i += 4 * stride) {
      float4 a[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) a[k] = __ldg(xi + i + k * stride);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        a[k].x = quant_one_tab(a[k].x, lm1, tab, use_tab); a[k].y = quant_one_tab(a[k].y, lm1, tab, use_tab);
        a[k].z = quant_one_tab(a[k].z, lm1, tab, use_tab); a[k].w = quant_one_tab(a[k].w, lm1, tab, use_tab);
        oo[i + k * stride] = a[k];
      }
    }
    for (; i < n4; i += stride) {
      float4 a = __ldg(xi + i);
      a.x = quant_one_tab(a.x, lm1, tab, use_tab); a.y = quant_one_tab(a.y, lm1, tab, use_tab);
      a.z = quant_one_tab(a.z, lm1, tab, use_tab); a.w = quant_one_tab(a.w, lm1, tab, use_tab);
      oo[i] = a;
    }
    for (int64_t j = n4 * 4 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += stride)
      out[j] = quant_one_tab(x[j], lm1, tab, use_tab);
  } else {
    for (; i < n; i += stride) out[i] = quant_one_tab(x[i], lm1, tab, use_tab);
  }
}

// s / 9 with IEEE round-to-nearest, by Markstein's correction instead of the generic division sequence: with
// y = RN(1/9), q0 = RN(s*y), the remainder r = s - 9*q0 is exact in one FMA (it is a few ulp(q0), and ulp(q0) >= 2^-149
// whenever q0 != 0) and RN(q0 + r*y) is the correctly rounded quotient -- an FMA rounds once, so neither tiny nor huge
// finite s needs a guard.  What remains: s = -0 (q1 would come out +0) and s = +-inf / NaN (r = NaN); in all of them r is
// zero or NaN and q0 is already the answer, so ONE ordered compare selects between q1 and q0 and there is no branch.
// (Round 1 guarded the exponent range and branched to the generic sequence: a BSSY/BRA/BSYNC region and two FSETPs
// per output, which also kept ptxas from interleaving the 16 dependent add chains of a strip.)
// Verified EXHAUSTIVELY over all 2^32 bit patterns: on the device against __fdiv_rn (advb200_test_div9,
// tests/test_gpu_parity.py) and on the host against IEEE division with C fmaf (oracle/check_div9.c).
__device__ __forceinline__ float div9(float s) {
  const float y = 0.111111112f;
  const float q0 = __fmul_rn(s, y);
  const float r = __fmaf_rn(-9.0f, q0, s);
  const float q1 = __fmaf_rn(r, y, q0);
  float q;
  // setp.ne is the ORDERED not-equal (false for NaN); C's != is the unordered one
  asm("{\n.reg .pred p;\nsetp.ne.f32 p, %1, 0f00000000;\nselp.f32 %0, %2, %3, p;\n}" : "=f"(q) : "f"(r), "f"(q1), "f"(q0));
  return q;
}

__global__ void div9_check_kernel(unsigned long long* mismatches, unsigned int* first_bad) {
  const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
  unsigned long long bad = 0;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < (1ull << 32); i += stride) {
    const float s = __uint_as_float((unsigned int)i);
    const float want = __fdiv_rn(s, 9.0f), got = div9(s);
    const bool same = (__float_as_uint(want) == __float_as_uint(got)) || (want != want && got != got);
    if (!same) { ++bad; atomicMin(first_bad, (unsigned int)i); }
  }
  if (bad) atomicAdd(mismatches, bad);
}

// 3x3 mean (zero pad, divisor 9) with optional clamp-in / quantise-out (defend_input fusion).
// One thread produces a 4-wide strip of kSmoothRows output rows from a (kSmoothRows + 2)-row register window, so
// every input element is fetched from L1/L2 about 1.5x and from HBM once.
constexpr int kSmoothRows = 4;

// (kSmoothRows + 2) x 6 input window of a 4-wide, kSmoothRows-tall output strip, zero padded.  93 % of the strips of a
// 224 x 224 plane are interior and 16-byte aligned: they take the branch-free path (18 unconditional loads, all issued
// before any use); edge or unaligned strips take the bounds-checked scalar path.
__device__ __forceinline__ void load_window(const float* __restrict__ src, int h, int w, int r0, int c0, bool vec,
                                            float (&win)[kSmoothRows + 2][6]) {
  const bool interior = vec && r0 > 0 && r0 + kSmoothRows < h && c0 > 0 && c0 + 4 < w;
  if (interior) {
    const float* row = src + (int64_t)(r0 - 1) * w + c0;
#pragma unroll
    for (int k = 0; k < kSmoothRows + 2; ++k, row += w) {
      const float4 m = __ldg(reinterpret_cast<const float4*>(row));
      win[k][0] = __ldg(row - 1);
      win[k][1] = m.x; win[k][2] = m.y; win[k][3] = m.z; win[k][4] = m.w;
      win[k][5] = __ldg(row + 4);
    }
    return;
  }
#pragma unroll
  for (int k = 0; k < kSmoothRows + 2; ++k) {
    const int r = r0 - 1 + k;
    float* v = win[k];
    if (r < 0 || r >= h) {
#pragma unroll
      for (int j = 0; j < 6; ++j) v[j] = 0.f;
      continue;
    }
    const float* row = src + (int64_t)r * w;
    if (vec) {
      const float4 m = __ldg(reinterpret_cast<const float4*>(row + c0));
      v[0] = c0 > 0 ? __ldg(row + c0 - 1) : 0.f;
      v[1] = m.x; v[2] = m.y; v[3] = m.z; v[4] = m.w;
      v[5] = c0 + 4 < w ? __ldg(row + c0 + 4) : 0.f;
    } else {
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        const int c = c0 - 1 + j;
        v[j] = (c >= 0 && c < w) ? __ldg(row + c) : 0.f;
      }
    }
  }
}

template <bool kDefend, typename Idx>
__global__ void __launch_bounds__(256, 4)
smooth3x3_kernel(const float* __restrict__ x, float* __restrict__ out, int64_t planes, int h, int w,
                 float lm1, const FastDiv div_groups, const FastDiv div_chunks) {
  __shared__ float tab[kQuantTab];
  const bool use_tab = kDefend && lm1 < (float)kQuantTab;
  if (kDefend) quant_table_init(tab, lm1);
  const int groups = (w + 3) >> 2;                         // 4-wide column groups
  const int row_chunks = (h + kSmoothRows - 1) / kSmoothRows;
  // Idx = uint32_t whenever the strip count (+ one grid stride) fits 31 bits: the strip index is then split by two
  // multiply-shift divisions (FastDiv) instead of three 64-bit divisions
  const Idx total = (Idx)(planes * row_chunks * (int64_t)groups);
  for (Idx t = (Idx)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (Idx)gridDim.x * blockDim.x) {
    const Idx tq = fast_div(t, div_groups);
    const int gx = (int)(t - tq * (Idx)groups);
    const Idx tp = fast_div(tq, div_chunks);
    const int rc = (int)(tq - tp * (Idx)row_chunks);
    const int64_t p = (int64_t)tp;
    const float* src = x + p * (int64_t)h * w;
    float* dst = out + p * (int64_t)h * w;
    const int c0 = gx * 4;
    const int r0 = rc * kSmoothRows;
    // the whole (kSmoothRows + 2)-row window is fetched before any arithmetic: 10 rows of independent loads in
    // flight per thread instead of one (the sliding 3-row window left the kernel latency-bound at 0.48 of HBM peak)
    float win[kSmoothRows + 2][6];
    const bool vec = ((w & 3) == 0) && ((reinterpret_cast<uintptr_t>(src) & 15u) == 0);
    load_window(src, h, w, r0, c0, vec, win);
    if (kDefend) {
      // clamp AFTER every load of the window has been issued: clamping inside load_row made each row's loads wait for
      // the previous row's first use (ncu: long-scoreboard 6.1 stalls per issue, 0.49 of the HBM rate)
#pragma unroll
      for (int k = 0; k < kSmoothRows + 2; ++k)
#pragma unroll
        // one saturating instruction per element instead of two FMNMX: saturate(x) == clamp(x, 0, 1), NaN -> 0 included.
        // (At most the SIGN of a zero could differ, and that cannot reach the output: a sum is -0 only if all nine taps
        // are, and the quantiser's table maps both zeros to entry 0 = +0.)
        for (int j = 0; j < 6; ++j) win[k][j] = __saturatef(win[k][j]);               // saturate(0) = 0: the zero padding survives
    }
#pragma unroll
    for (int dr = 0; dr < kSmoothRows; ++dr) {
      const int r = r0 + dr;
      if (r >= h) break;
      const float* a = win[dr];
      const float* b = win[dr + 1];
      const float* c = win[dr + 2];
      float o4[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        // row-major accumulation of the nine taps, exactly F.avg_pool2d's order
        float s = a[j];
        s = __fadd_rn(s, a[j + 1]); s = __fadd_rn(s, a[j + 2]);
        s = __fadd_rn(s, b[j]); s = __fadd_rn(s, b[j + 1]); s = __fadd_rn(s, b[j + 2]);
        s = __fadd_rn(s, c[j]); s = __fadd_rn(s, c[j + 1]); s = __fadd_rn(s, c[j + 2]);
        float o = div9(s);
        // (the reference's trailing clamp, defense_experiments.py:217, is the identity on k/(levels-1), k <= levels-1)
        // the mean of nine values in [0,1] is in [0,1] (rounded sums of values <= 1 never exceed 9): the quantiser's own
        // clamp (defense_experiments.py:183) is the identity here and is skipped
        if (kDefend) o = use_tab ? tab[__float2int_rn(__fmul_rn(o, lm1))] : __fdiv_rn(rintf(__fmul_rn(o, lm1)), lm1);
        o4[j] = o;
      }
      if (vec && ((reinterpret_cast<uintptr_t>(dst) & 15u) == 0)) {
        *reinterpret_cast<float4*>(dst + (int64_t)r * w + c0) = make_float4(o4[0], o4[1], o4[2], o4[3]);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (c0 + j < w) dst[(int64_t)r * w + c0 + j] = o4[j];
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// CW-L2 (tanh-space Adam)
// ------------------------------------------------------------------------------------------
constexpr int kCwThreads = 256;
constexpr int kCwPerThread = 12;                        // 3 x float4
constexpr int kCwChunk = kCwThreads * kCwPerThread;     // 3072 elements per block

__device__ __forceinline__ float atanh_like_ref(float t) {
  // attack.py:68-69: 0.5 * (log1p(t) - log1p(-t))
  return 0.5f * (log1pf(t) - log1pf(-t));
}

__global__ void __launch_bounds__(kCwThreads)
cw_init_kernel(const float* __restrict__ x, float* __restrict__ x0, float* __restrict__ w,
               float* __restrict__ m, float* __restrict__ v, float* __restrict__ x_adv,
               float* __restrict__ best_adv, float* __restrict__ l2_partials, int parts,
               int64_t elems_per_img) {
  __shared__ float red[kCwThreads / 32];
  const int img = blockIdx.x / parts, part = blockIdx.x % parts;
  const int64_t base = (int64_t)img * elems_per_img;
  const int64_t lo = (int64_t)part * kCwChunk;
  float acc = 0.f;
  for (int k = 0; k < kCwPerThread; ++k) {
    const int64_t e = lo + (int64_t)k * kCwThreads + threadIdx.x;
    if (e < elems_per_img) {
      const int64_t i = base + e;
      const float c = fminf(fmaxf(x[i], 0.f), 1.f);
      const float eps = 1e-6f;
      const float xt = c * (1.0f - 2.0f * eps) + eps;     // attack.py:96-97
      const float ww = atanh_like_ref(xt * 2.0f - 1.0f);
      const float xa = 0.5f * (tanhf(ww) + 1.0f);
      x0[i] = c; w[i] = ww; m[i] = 0.f; v[i] = 0.f; x_adv[i] = xa; best_adv[i] = c;
      const float d = xa - c;
      acc += d * d;
    }
  }
  const float s = block_sum<kCwThreads>(acc, red);
  if (threadIdx.x == 0) l2_partials[(int64_t)img * parts + part] = s;
}

__global__ void cw_init_state_kernel(float* best_l2, uint8_t* best_success, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { best_l2[i] = __int_as_float(0x7f800000); best_success[i] = 0; }
}

// attack.py:126,133-136 per image (tiny): l2, improved flag, best_l2 / best_success bookkeeping
__global__ void cw_decide_kernel(const float* __restrict__ l2_partials, int parts,
                                 const uint8_t* __restrict__ success, float* __restrict__ best_l2,
                                 uint8_t* __restrict__ best_success, uint8_t* __restrict__ improved,
                                 float* __restrict__ l2_out, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float s = 0.f;
  for (int p = 0; p < parts; ++p) s += l2_partials[(int64_t)i * parts + p];
  if (l2_out) l2_out[i] = s;
  const bool imp = success[i] && (s < best_l2[i]);
  improved[i] = imp ? 1 : 0;
  if (imp) { best_l2[i] = s; best_success[i] = 1; }
}

__global__ void __launch_bounds__(kCwThreads)
cw_step_kernel(float* __restrict__ w, float* __restrict__ m, float* __restrict__ v,
               const float* __restrict__ x0, const float* __restrict__ grad_x,
               float* __restrict__ x_adv, float* __restrict__ best_adv,
               const uint8_t* __restrict__ improved, float* __restrict__ l2_partials_out, int parts,
               int64_t elems_per_img, float step_size, float bc2_sqrt, float beta1, float beta2,
               float one_m_b1, float one_m_b2, float adam_eps, bool aligned) {
  __shared__ float red[kCwThreads / 32];
  const int img = blockIdx.x / parts, part = blockIdx.x % parts;
  const int64_t base = (int64_t)img * elems_per_img + (int64_t)part * kCwChunk;
  const int64_t remain = elems_per_img - (int64_t)part * kCwChunk;
  const bool imp = improved[img] != 0;
  float acc = 0.f;

  auto update = [&](float ww, float mm, float vv, float xo, float g, float& w_o, float& m_o,
                    float& v_o, float& xa_o, float& xcur_o) {
    const float t = tanhf(ww);
    const float xc = 0.5f * (t + 1.0f);                    // attack.py:110 (current iterate)
    xcur_o = xc;
    const float gx = 2.0f * (xc - xo) + g;                 // d(l2)/dx + c*df/dx
    // through 0.5*(tanh(w)+1): torch's tanh backward is g * (1 - y*y) with the product rounded before the subtraction;
    // a fused 1 - t*t differs by up to 6e-8, which is 1e-4 of (1 - t*t) for a pixel near saturation (|w| ~ 4)
    const float gw = (gx * 0.5f) * __fsub_rn(1.0f, __fmul_rn(t, t));
    mm = mm + (gw - mm) * one_m_b1;                        // exp_avg.lerp_(grad, 1-beta1)
    vv = vv * beta2 + one_m_b2 * gw * gw;                  // exp_avg_sq.mul_(b2).addcmul_(g,g,1-b2)
    const float denom = sqrtf(vv) / bc2_sqrt + adam_eps;
    ww = ww - step_size * (mm / denom);                    // param.addcdiv_(m, denom, -step_size)
    const float xn = 0.5f * (tanhf(ww) + 1.0f);            // next iterate
    w_o = ww; m_o = mm; v_o = vv; xa_o = xn;
    const float d = xn - xo;
    acc += d * d;
  };

  const bool vec = (remain >= kCwChunk) && ((elems_per_img & 3) == 0) && aligned;
  if (vec) {
    float4* w4 = reinterpret_cast<float4*>(w + base);
    float4* m4 = reinterpret_cast<float4*>(m + base);
    float4* v4 = reinterpret_cast<float4*>(v + base);
    const float4* x04 = reinterpret_cast<const float4*>(x0 + base);
    const float4* g4 = reinterpret_cast<const float4*>(grad_x + base);
    float4* xa4 = reinterpret_cast<float4*>(x_adv + base);
    float4* b4 = reinterpret_cast<float4*>(best_adv + base);
#pragma unroll
    for (int k = 0; k < kCwPerThread / 4; ++k) {
      const int i = k * kCwThreads + threadIdx.x;
      float4 ww = w4[i], mm = m4[i], vv = v4[i];
      const float4 xo = __ldg(x04 + i), g = __ldg(g4 + i);
      float4 xa, xc;
      update(ww.x, mm.x, vv.x, xo.x, g.x, ww.x, mm.x, vv.x, xa.x, xc.x);
      update(ww.y, mm.y, vv.y, xo.y, g.y, ww.y, mm.y, vv.y, xa.y, xc.y);
      update(ww.z, mm.z, vv.z, xo.z, g.z, ww.z, mm.z, vv.z, xa.z, xc.z);
      update(ww.w, mm.w, vv.w, xo.w, g.w, ww.w, mm.w, vv.w, xa.w, xc.w);
      w4[i] = ww; m4[i] = mm; v4[i] = vv; xa4[i] = xa;
      if (imp) b4[i] = xc;
    }
  } else {
    for (int k = 0; k < kCwPerThread; ++k) {
      const int64_t e = (int64_t)k * kCwThreads + threadIdx.x;
      if (e < remain) {
        const int64_t i = base + e;
        float ww, mm, vv, xa, xc;
        update(w[i], m[i], v[i], x0[i], grad_x[i], ww, mm, vv, xa, xc);
        w[i] = ww; m[i] = mm; v[i] = vv; x_adv[i] = xa;
        if (imp) best_adv[i] = xc;
      }
    }
  }
  const float s = block_sum<kCwThreads>(acc, red);
  if (threadIdx.x == 0) l2_partials_out[(int64_t)img * parts + part] = s;
}

__global__ void __launch_bounds__(kThreads)
cw_finalize_kernel(const float* __restrict__ x_adv, const float* __restrict__ best_adv,
                   const uint8_t* __restrict__ best_success, float* __restrict__ out, int64_t total,
                   int64_t elems_per_img) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int img = (int)(i / elems_per_img);
    out[i] = best_success[img] ? best_adv[i] : x_adv[i];
  }
}

// ------------------------------------------------------------------------------------------
// Evaluation counters (defense_experiments.py:597-637)
// ------------------------------------------------------------------------------------------
__global__ void eval_counters_kernel(const int64_t* __restrict__ pred_clean,
                                     const int64_t* __restrict__ pred_adv,
                                     const int64_t* __restrict__ pred_def,
                                     const float* __restrict__ score_clean,
                                     const float* __restrict__ score_adv, float thr,
                                     const int64_t* __restrict__ y, int n,
                                     unsigned long long* __restrict__ counters,
                                     uint8_t* __restrict__ flags) {
  __shared__ unsigned int sc[7];
  if (threadIdx.x < 7) sc[threadIdx.x] = 0;
  __syncthreads();
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int clean_correct = pred_clean[i] == y[i];
    const int attack_success = pred_adv[i] != y[i];
    const int def_ok = pred_def[i] == y[i];
    const int flag_clean = score_clean[i] > thr;   // strict, defense_experiments.py:285
    const int flag_adv = score_adv[i] > thr;
    const int bypass = attack_success && !flag_adv;
    const int f[6] = {clean_correct, attack_success, def_ok, flag_clean, flag_adv, bypass};
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      if (f[k]) atomicAdd(&sc[k], 1u);
      if (flags) flags[(int64_t)i * 6 + k] = (uint8_t)f[k];
    }
    atomicAdd(&sc[6], 1u);
  }
  __syncthreads();
  if (threadIdx.x < 7 && sc[threadIdx.x]) atomicAdd(&counters[threadIdx.x], (unsigned long long)sc[threadIdx.x]);
}

// 3x3 median (zero pad 1, stride 1) — the "median smoothing" defense the north star names beside the reference's
// 3x3 mean (defense_experiments.py:177-179 keeps the mean; this is the same window and padding with the selection
// in place of the average).  Pure selection, so the result is bit-exact whatever the order.  Same strip layout as
// smooth3x3_kernel; the three vertical taps of every column are sorted once and shared by the three outputs that
// use the column: median9 = med3(max of the column minima, med3 of the column medians, min of the column maxima).
__device__ __forceinline__ float med3(float a, float b, float c) {
  return fmaxf(fminf(a, b), fminf(fmaxf(a, b), c));
}
template <typename Idx>
__global__ void __launch_bounds__(256, 4)
median3x3_kernel(const float* __restrict__ x, float* __restrict__ out, int64_t planes, int h, int w,
                 const FastDiv div_groups, const FastDiv div_chunks) {
  const int groups = (w + 3) >> 2;
  const int row_chunks = (h + kSmoothRows - 1) / kSmoothRows;
  const Idx total = (Idx)(planes * row_chunks * (int64_t)groups);      // (32-bit strip index when it fits: see smooth3x3_kernel)
  for (Idx t = (Idx)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (Idx)gridDim.x * blockDim.x) {
    const Idx tq = fast_div(t, div_groups);
    const int gx = (int)(t - tq * (Idx)groups);
    const Idx tp = fast_div(tq, div_chunks);
    const int rc = (int)(tq - tp * (Idx)row_chunks);
    const int64_t p = (int64_t)tp;
    const float* src = x + p * (int64_t)h * w;
    float* dst = out + p * (int64_t)h * w;
    const int c0 = gx * 4;
    const int r0 = rc * kSmoothRows;
    float win[kSmoothRows + 2][6];
    const bool vec = ((w & 3) == 0) && ((reinterpret_cast<uintptr_t>(src) & 15u) == 0);
    load_window(src, h, w, r0, c0, vec, win);
    // two output rows share the ordered pair of their two common input rows (per column: 2 + 2 x 4 operations instead
    // of 2 x 6), and two neighbouring outputs share the ordered pair of their two common column medians
#pragma unroll
    for (int dp = 0; dp < kSmoothRows; dp += 2) {
      float lo[2][6], mid[2][6], hi[2][6];
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        const float plo = fminf(win[dp + 1][j], win[dp + 2][j]), phi = fmaxf(win[dp + 1][j], win[dp + 2][j]);
        const float a = win[dp][j], c = win[dp + 3][j];
        lo[0][j] = fminf(a, plo); hi[0][j] = fmaxf(a, phi); mid[0][j] = fmaxf(plo, fminf(a, phi));
        lo[1][j] = fminf(c, plo); hi[1][j] = fmaxf(c, phi); mid[1][j] = fmaxf(plo, fminf(c, phi));
      }
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int dr = dp + q;
        const int r = r0 + dr;
        if (r >= h) break;
        float o4[4];
#pragma unroll
        for (int j = 0; j < 4; j += 2) {
          const float mlo = fminf(mid[q][j + 1], mid[q][j + 2]), mhi = fmaxf(mid[q][j + 1], mid[q][j + 2]);
          const float m0 = fmaxf(mlo, fminf(mid[q][j], mhi));
          const float m1 = fmaxf(mlo, fminf(mid[q][j + 3], mhi));
          o4[j] = med3(fmaxf(fmaxf(lo[q][j], lo[q][j + 1]), lo[q][j + 2]), m0,
                       fminf(fminf(hi[q][j], hi[q][j + 1]), hi[q][j + 2]));
          o4[j + 1] = med3(fmaxf(fmaxf(lo[q][j + 1], lo[q][j + 2]), lo[q][j + 3]), m1,
                           fminf(fminf(hi[q][j + 1], hi[q][j + 2]), hi[q][j + 3]));
        }
        float* orow = dst + (int64_t)r * w;
        if (vec && ((reinterpret_cast<uintptr_t>(dst) & 15u) == 0)) {      // (the output plane may be less aligned than the input)
          *reinterpret_cast<float4*>(orow + c0) = make_float4(o4[0], o4[1], o4[2], o4[3]);
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (c0 + j < w) orow[c0 + j] = o4[j];
        }
      }
    }
  }
}

}  // namespace
}  // namespace advb200

using namespace advb200;

extern "C" {

int advb200_linf_step_f32(float* x_adv, const float* x_orig, const float* grad, float alpha, float eps,
                          int64_t n, int project, void* stream) {
  ADV_CHECK_ARG(n >= 0, "linf_step: negative size");
  if (n == 0) return ADVB200_OK;      // empty batch: nothing to do (pointers may be null)
  ADV_CHECK_ARG(x_adv && grad, "linf_step: null pointer");
  ADV_CHECK_ARG(!project || x_orig, "linf_step: x_orig required when project != 0");
  cudaStream_t st = as_stream(stream);
  const bool vec = aligned16(x_adv) && aligned16(grad) && (!project || aligned16(x_orig));
  if (vec) {
    const int64_t n4 = n / 4;
    const int grid = grid_for(n4 > 0 ? n4 : 1);
    if (project) linf_step_kernel<true><<<grid, kThreads, 0, st>>>(x_adv, x_orig, grad, alpha, eps, n4, n);
    else linf_step_kernel<false><<<grid, kThreads, 0, st>>>(x_adv, x_orig, grad, alpha, eps, n4, n);
  } else {
    const int grid = grid_for(n);
    if (project) linf_step_scalar_kernel<true><<<grid, kThreads, 0, st>>>(x_adv, x_orig, grad, alpha, eps, n);
    else linf_step_scalar_kernel<false><<<grid, kThreads, 0, st>>>(x_adv, x_orig, grad, alpha, eps, n);
  }
  ADV_LAUNCHED();
  return ADVB200_OK;
}

int advb200_random_start_f32(float* x_adv, const float* x_orig, const float* noise, int64_t n,
                             void* stream) {
  ADV_CHECK_ARG(n >= 0, "random_start: negative size");
  if (n == 0) return ADVB200_OK;
  ADV_CHECK_ARG(x_adv && x_orig && noise, "random_start: null pointer");
  random_start_kernel<<<grid_for(n), kThreads, 0, as_stream(stream)>>>(x_adv, x_orig, noise, n);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

int advb200_quantize_f32(const float* x, float* out, int64_t n, int levels, void* stream) {
  ADV_CHECK_ARG(n >= 0, "quantize: negative size");
  ADV_CHECK_ARG(levels >= 2, "quantize: levels must be >= 2");
  if (n == 0) return ADVB200_OK;
  ADV_CHECK_ARG(x && out, "quantize: null pointer");
  const int vec = aligned16(x) && aligned16(out);
  const int64_t n4 = n / 4;
  quantize_kernel<<<grid_for(vec ? (n4 > 0 ? n4 : 1) : n), kThreads, 0, as_stream(stream)>>>(
      x, out, n4, n, (float)(levels - 1), vec);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

static int smooth_launch(const float* x, float* out, int64_t planes, int h, int w, int levels,
                         bool defend, void* stream) {
  ADV_CHECK_ARG(planes >= 0 && h > 0 && w > 0, "smooth3x3: bad sizes");
  if (planes == 0) return ADVB200_OK;
  ADV_CHECK_ARG(x && out, "smooth3x3: null pointer");
  ADV_CHECK_ARG(x != out, "smooth3x3: in-place operation is not supported");
  const int groups = (w + 3) / 4, chunks = (h + kSmoothRows - 1) / kSmoothRows;
  const int64_t total = planes * (int64_t)groups * chunks;
  const int grid = grid_for(total);
  const bool small = total + (int64_t)grid * 256 < (int64_t(1) << 31);
  const FastDiv dg = make_fastdiv((uint32_t)groups), dc = make_fastdiv((uint32_t)chunks);
  if (defend && small)
    smooth3x3_kernel<true, uint32_t><<<grid, 256, 0, as_stream(stream)>>>(x, out, planes, h, w, (float)(levels - 1), dg, dc);
  else if (defend)
    smooth3x3_kernel<true, int64_t><<<grid, 256, 0, as_stream(stream)>>>(x, out, planes, h, w, (float)(levels - 1), dg, dc);
  else if (small)
    smooth3x3_kernel<false, uint32_t><<<grid, 256, 0, as_stream(stream)>>>(x, out, planes, h, w, 0.f, dg, dc);
  else
    smooth3x3_kernel<false, int64_t><<<grid, 256, 0, as_stream(stream)>>>(x, out, planes, h, w, 0.f, dg, dc);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

int advb200_smooth3x3_f32(const float* x, float* out, int64_t planes, int h, int w, void* stream) {
  return smooth_launch(x, out, planes, h, w, 2, false, stream);
}

int advb200_median3x3_f32(const float* x, float* out, int64_t planes, int h, int w, void* stream) {
  ADV_CHECK_ARG(planes >= 0 && h > 0 && w > 0, "median3x3: bad sizes");
  if (planes == 0) return ADVB200_OK;
  ADV_CHECK_ARG(x && out, "median3x3: null pointer");
  ADV_CHECK_ARG(x != out, "median3x3: in-place operation is not supported");
  const int groups = (w + 3) / 4, chunks = (h + kSmoothRows - 1) / kSmoothRows;
  const int64_t total = planes * (int64_t)groups * chunks;
  const int grid = grid_for(total);
  const FastDiv dg = make_fastdiv((uint32_t)groups), dc = make_fastdiv((uint32_t)chunks);
  if (total + (int64_t)grid * 256 < (int64_t(1) << 31))
    median3x3_kernel<uint32_t><<<grid, 256, 0, as_stream(stream)>>>(x, out, planes, h, w, dg, dc);
  else
    median3x3_kernel<int64_t><<<grid, 256, 0, as_stream(stream)>>>(x, out, planes, h, w, dg, dc);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

int advb200_test_div9(uint64_t* mismatches_host, uint32_t* first_bad_host) {
  // test hook: the constant-divisor shortcut of the smoothing kernels against __fdiv_rn, all 2^32 inputs
  ADV_CHECK_ARG(mismatches_host && first_bad_host, "test_div9: null pointer");
  unsigned long long* d_bad = nullptr;
  unsigned int* d_first = nullptr;
  ADV_CHECK_CUDA(cudaMalloc(&d_bad, 8));
  ADV_CHECK_CUDA(cudaMalloc(&d_first, 4));
  ADV_CHECK_CUDA(cudaMemset(d_bad, 0, 8));
  ADV_CHECK_CUDA(cudaMemset(d_first, 0xff, 4));
  div9_check_kernel<<<kNumSMs * 16, 256>>>(d_bad, d_first);
  ADV_LAUNCHED();
  unsigned long long bad = 0;
  ADV_CHECK_CUDA(cudaMemcpy(&bad, d_bad, 8, cudaMemcpyDeviceToHost));
  ADV_CHECK_CUDA(cudaMemcpy(first_bad_host, d_first, 4, cudaMemcpyDeviceToHost));
  *mismatches_host = bad;
  cudaFree(d_bad); cudaFree(d_first);
  return ADVB200_OK;
}

int advb200_defend_input_f32(const float* x, float* out, int64_t planes, int h, int w, int levels,
                             void* stream) {
  ADV_CHECK_ARG(levels >= 2, "defend_input: levels must be >= 2");
  return smooth_launch(x, out, planes, h, w, levels, true, stream);
}

int advb200_cw_parts(int64_t elems_per_img) {
  return (int)((elems_per_img + kCwChunk - 1) / kCwChunk);
}

int advb200_cw_init_f32(const float* x, float* x0, float* w, float* m, float* v, float* x_adv,
                        float* best_adv, float* best_l2, uint8_t* best_success, float* l2_partials,
                        int n_img, int64_t elems_per_img, void* stream) {
  ADV_CHECK_ARG(x && x0 && w && m && v && x_adv && best_adv && best_l2 && best_success && l2_partials,
                "cw_init: null pointer");
  ADV_CHECK_ARG(n_img >= 0 && elems_per_img > 0, "cw_init: bad sizes");
  if (n_img == 0) return ADVB200_OK;
  const int parts = advb200_cw_parts(elems_per_img);
  cw_init_kernel<<<n_img * parts, kCwThreads, 0, as_stream(stream)>>>(x, x0, w, m, v, x_adv, best_adv,
                                                                      l2_partials, parts, elems_per_img);
  ADV_LAUNCHED();
  cw_init_state_kernel<<<(n_img + 255) / 256, 256, 0, as_stream(stream)>>>(best_l2, best_success, n_img);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

int advb200_cw_step_f32(float* w, float* m, float* v, const float* x0, const float* grad_x, float* x_adv,
                        float* best_adv, float* best_l2, uint8_t* best_success, const uint8_t* success,
                        const float* l2_partials_in, float* l2_partials_out, float* l2_out, int n_img,
                        int64_t elems_per_img, float step_size, float bc2_sqrt, double beta1, double beta2,
                        double adam_eps, void* stream) {
  ADV_CHECK_ARG(w && m && v && x0 && grad_x && x_adv && best_adv && best_l2 && best_success && success &&
                    l2_partials_in && l2_partials_out,
                "cw_step: null pointer");
  ADV_CHECK_ARG(l2_partials_in != l2_partials_out, "cw_step: l2 partial buffers must differ");
  ADV_CHECK_ARG(n_img >= 0 && elems_per_img > 0, "cw_step: bad sizes");
  if (n_img == 0) return ADVB200_OK;
  const int parts = advb200_cw_parts(elems_per_img);
  // `improved` lives in the tail of l2_partials_out's sibling? no: reuse best-effort scratch —
  // the flag array is carved from the end of l2_partials_out (caller sizes it [N, parts+1]).
  uint8_t* improved = reinterpret_cast<uint8_t*>(l2_partials_out + (int64_t)n_img * parts);
  cw_decide_kernel<<<(n_img + 127) / 128, 128, 0, as_stream(stream)>>>(
      l2_partials_in, parts, success, best_l2, best_success, improved, l2_out, n_img);
  ADV_LAUNCHED();
  cw_step_kernel<<<n_img * parts, kCwThreads, 0, as_stream(stream)>>>(
      w, m, v, x0, grad_x, x_adv, best_adv, improved, l2_partials_out, parts, elems_per_img, step_size,
      bc2_sqrt, (float)beta1, (float)beta2, (float)(1.0 - beta1), (float)(1.0 - beta2), (float)adam_eps,   // torch: 1-beta in double
      aligned16(w) && aligned16(m) && aligned16(v) && aligned16(x0) && aligned16(grad_x) && aligned16(x_adv) &&
          aligned16(best_adv));
  ADV_LAUNCHED();
  return ADVB200_OK;
}

int advb200_cw_finalize_f32(const float* x_adv, const float* best_adv, const uint8_t* best_success,
                            float* out, int n_img, int64_t elems_per_img, void* stream) {
  ADV_CHECK_ARG(x_adv && best_adv && best_success && out, "cw_finalize: null pointer");
  if (n_img == 0) return ADVB200_OK;
  const int64_t total = (int64_t)n_img * elems_per_img;
  cw_finalize_kernel<<<grid_for(total), kThreads, 0, as_stream(stream)>>>(x_adv, best_adv, best_success,
                                                                          out, total, elems_per_img);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

int advb200_eval_counters(const int64_t* pred_clean, const int64_t* pred_adv, const int64_t* pred_def,
                          const float* score_clean, const float* score_adv, float threshold,
                          const int64_t* y_true, int n, int64_t* counters, uint8_t* flags_out,
                          void* stream) {
  ADV_CHECK_ARG(pred_clean && pred_adv && pred_def && score_clean && score_adv && y_true && counters,
                "eval_counters: null pointer");
  if (n <= 0) return ADVB200_OK;
  int grid = (n + 255) / 256;
  if (grid > kNumSMs) grid = kNumSMs;
  eval_counters_kernel<<<grid, 256, 0, as_stream(stream)>>>(
      pred_clean, pred_adv, pred_def, score_clean, score_adv, threshold, y_true, n,
      reinterpret_cast<unsigned long long*>(counters), flags_out);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

}  // extern "C"

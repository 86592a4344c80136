// Image preprocessing on the device: Resize(256) -> CenterCrop(224) -> ToTensor of the
// reference's `load_image` (ResNet.py:27-31, defense_experiments.py:649-662) for an already decoded uint8 RGB image,
// and ToPILImage of `--save_adv` (ResNet.py:121-126).
//
// The reference resizes with PIL (`ImagingResample`, Resample.c): a separable, antialiased bilinear convolution in
// fixed point on uint8 — per output position a window and weights normalised in double precision and rounded to 22
// fractional bits, horizontal pass first, uint8 intermediate, `clip8((sum + 2^21) >> 22)`.  The host side below
// restates the coefficient computation in the same double arithmetic (for the cropped window only: the passes are
// separable, so resampling just the 224 x 224 window equals cropping the full result); two integer kernels do the
// passes.  Bit-exact against torchvision + PIL (tests/golden/preprocess_cases.npz; oracle/preprocess_ref.py).
#include <cmath>
#include <vector>

#include "common.cuh"

namespace advb200 {
namespace {

constexpr int kPrecisionBits = 32 - 8 - 2;

struct AxisPlan {
  int in_size = 0, out_size = 0, lo = 0, n = 0;   // n output positions [lo, lo + n) of the resized axis
  int ksize = 1;
  int src0 = 0, src1 = 0;                         // source index range the n positions read
  std::vector<int32_t> bounds;                    // [n][2] = (first source index, tap count)
  std::vector<int32_t> coef;                      // [n][ksize]
};

// Resample.c precompute_coeffs (bilinear: support 1) + normalize_coeffs_8bpc, for positions [lo, lo + n)
void plan_axis(int in_size, int out_size, int lo, int n, AxisPlan* p) {
  p->in_size = in_size; p->out_size = out_size; p->lo = lo; p->n = n;
  if (in_size == out_size) {                       // ImagingResample skips the pass: identity taps
    p->ksize = 1;
    p->bounds.resize((size_t)n * 2);
    p->coef.assign((size_t)n, 1 << kPrecisionBits);
    for (int i = 0; i < n; ++i) { p->bounds[2 * i] = lo + i; p->bounds[2 * i + 1] = 1; }
    p->src0 = lo; p->src1 = lo + n;
    return;
  }
  double scale = (double)in_size / out_size, filterscale = scale;
  if (filterscale < 1.0) filterscale = 1.0;
  const double support = 1.0 * filterscale;
  p->ksize = (int)std::ceil(support) * 2 + 1;
  const double ss = 1.0 / filterscale;
  p->bounds.resize((size_t)n * 2);
  p->coef.assign((size_t)n * p->ksize, 0);
  std::vector<double> w((size_t)p->ksize);
  p->src0 = in_size; p->src1 = 0;
  for (int i = 0; i < n; ++i) {
    const int xx = lo + i;
    const double center = 0.0 + (xx + 0.5) * scale;
    int xmin = (int)(center - support + 0.5);
    if (xmin < 0) xmin = 0;
    int xmax = (int)(center + support + 0.5);
    if (xmax > in_size) xmax = in_size;
    xmax -= xmin;
    double ww = 0.0;
    for (int x = 0; x < xmax; ++x) {
      double a = (x + xmin - center + 0.5) * ss;
      if (a < 0.0) a = -a;
      const double v = a < 1.0 ? 1.0 - a : 0.0;
      w[x] = v;
      ww += v;
    }
    for (int x = 0; x < xmax; ++x) {
      const double v = ww != 0.0 ? w[x] / ww : w[x];
      p->coef[(size_t)i * p->ksize + x] =
          v < 0 ? (int32_t)(-0.5 + v * (1 << kPrecisionBits)) : (int32_t)(0.5 + v * (1 << kPrecisionBits));
    }
    p->bounds[2 * i] = xmin; p->bounds[2 * i + 1] = xmax;
    if (xmin < p->src0) p->src0 = xmin;
    if (xmin + xmax > p->src1) p->src1 = xmin + xmax;
  }
}

struct Geometry { int nh, nw, top, left; };

int geometry(int h, int w, int size, int crop, Geometry* g) {
  // torchvision _compute_resized_output_size (int size: shorter side -> size, longer = int(size * long / short))
  const int shorter = w <= h ? w : h, longer = w <= h ? h : w;
  const int new_long = (int)((double)((long long)size * longer) / (double)shorter);
  g->nw = w <= h ? size : new_long;
  g->nh = w <= h ? new_long : size;
  if (g->nh < crop || g->nw < crop) return -1;
  // torchvision center_crop: int(round((H - crop) / 2.0)), Python round = half to even = nearbyint
  g->top = (int)std::nearbyint((g->nh - crop) / 2.0);
  g->left = (int)std::nearbyint((g->nw - crop) / 2.0);
  return 0;
}

__device__ __forceinline__ int clip8(int v) { return min(max(v >> kPrecisionBits, 0), 255); }

// horizontal pass: tmp[r][ox][c] for source rows [row0, row0 + rows)
__global__ void resample_h_kernel(const uint8_t* __restrict__ img, int w, int row0, int rows, int n,
                                  const int32_t* __restrict__ bounds, const int32_t* __restrict__ coef, int ksize,
                                  uint8_t* __restrict__ tmp) {
  const long long total = (long long)rows * n;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const int ox = (int)(i % n), r = (int)(i / n);
    const int x0 = bounds[2 * ox], cnt = bounds[2 * ox + 1];
    const uint8_t* src = img + ((long long)(row0 + r) * w + x0) * 3;
    const int32_t* k = coef + (long long)ox * ksize;
    int s0 = 1 << (kPrecisionBits - 1), s1 = s0, s2 = s0;
    for (int t = 0; t < cnt; ++t) {
      const int kv = __ldg(k + t);
      s0 += src[3 * t] * kv; s1 += src[3 * t + 1] * kv; s2 += src[3 * t + 2] * kv;
    }
    uint8_t* d = tmp + i * 3;
    d[0] = (uint8_t)clip8(s0); d[1] = (uint8_t)clip8(s1); d[2] = (uint8_t)clip8(s2);
  }
}

// vertical pass + ToTensor: out[c][oy][ox] = clip8(...) / 255
__global__ void resample_v_kernel(const uint8_t* __restrict__ tmp, int row0, int n, int out_rows,
                                  const int32_t* __restrict__ bounds, const int32_t* __restrict__ coef, int ksize,
                                  float* __restrict__ out) {
  const long long total = (long long)out_rows * n;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const int ox = (int)(i % n), oy = (int)(i / n);
    const int y0 = bounds[2 * oy], cnt = bounds[2 * oy + 1];
    const int32_t* k = coef + (long long)oy * ksize;
    int s0 = 1 << (kPrecisionBits - 1), s1 = s0, s2 = s0;
    for (int t = 0; t < cnt; ++t) {
      const uint8_t* p = tmp + ((long long)(y0 - row0 + t) * n + ox) * 3;
      const int kv = __ldg(k + t);
      s0 += p[0] * kv; s1 += p[1] * kv; s2 += p[2] * kv;
    }
    const long long plane = (long long)out_rows * n;
    out[i] = __fdiv_rn((float)clip8(s0), 255.f);
    out[plane + i] = __fdiv_rn((float)clip8(s1), 255.f);
    out[2 * plane + i] = __fdiv_rn((float)clip8(s2), 255.f);
  }
}

__global__ void to_u8_hwc_kernel(const float* __restrict__ x, uint8_t* __restrict__ out, long long pixels) {
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < pixels; i += (long long)gridDim.x * 256) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      // ToPILImage: pic.mul(255).byte() — truncation; inputs of the attack path are already inside [0, 1]
      const float v = __fmul_rn(fminf(fmaxf(__ldg(x + c * pixels + i), 0.f), 1.f), 255.f);
      out[i * 3 + c] = (uint8_t)(int)v;
    }
  }
}

inline int64_t align256(int64_t v) { return (v + 255) / 256 * 256; }

struct Layout { int64_t hb, hk, vb, vk, tmp, total; };

Layout layout(const AxisPlan& hp, const AxisPlan& vp) {
  Layout l;
  int64_t o = 0;
  l.hb = o; o = align256(o + (int64_t)hp.n * 2 * 4);
  l.hk = o; o = align256(o + (int64_t)hp.n * hp.ksize * 4);
  l.vb = o; o = align256(o + (int64_t)vp.n * 2 * 4);
  l.vk = o; o = align256(o + (int64_t)vp.n * vp.ksize * 4);
  l.tmp = o; o = align256(o + (int64_t)(vp.src1 - vp.src0) * hp.n * 3);
  l.total = o;
  return l;
}

}  // namespace
}  // namespace advb200

using namespace advb200;

extern "C" {

int64_t advb200_preprocess_scratch_bytes(int h, int w, int size, int crop) {
  if (h <= 0 || w <= 0 || size <= 0 || crop <= 0) return -1;
  Geometry g;
  if (geometry(h, w, size, crop, &g)) return -1;
  AxisPlan hp, vp;
  plan_axis(w, g.nw, g.left, crop, &hp);
  plan_axis(h, g.nh, g.top, crop, &vp);
  return layout(hp, vp).total;
}

int advb200_preprocess_u8(const uint8_t* img_hwc, int h, int w, int size, int crop, float* out_chw, void* scratch,
                          int64_t scratch_bytes, void* stream) {
  ADV_CHECK_ARG(img_hwc && out_chw && scratch, "preprocess: null pointer");
  ADV_CHECK_ARG(h > 0 && w > 0 && size > 0 && crop > 0, "preprocess: bad sizes");
  ADV_CHECK_ARG((reinterpret_cast<uintptr_t>(scratch) & 255) == 0, "preprocess: scratch must be 256-byte aligned");
  Geometry g;
  if (geometry(h, w, size, crop, &g))
    return fail(ADVB200_ERR_UNSUPPORTED, "preprocess: the resized image is smaller than the crop (CenterCrop would pad)");
  AxisPlan hp, vp;
  plan_axis(w, g.nw, g.left, crop, &hp);
  plan_axis(h, g.nh, g.top, crop, &vp);
  const Layout l = layout(hp, vp);
  ADV_CHECK_ARG(scratch_bytes >= l.total, "preprocess: scratch too small");
  cudaStream_t st = as_stream(stream);
  uint8_t* base = static_cast<uint8_t*>(scratch);
  // pageable-source async copies are staged before the call returns, so the plans may go out of scope
  ADV_CHECK_CUDA(cudaMemcpyAsync(base + l.hb, hp.bounds.data(), hp.bounds.size() * 4, cudaMemcpyHostToDevice, st));
  ADV_CHECK_CUDA(cudaMemcpyAsync(base + l.hk, hp.coef.data(), hp.coef.size() * 4, cudaMemcpyHostToDevice, st));
  ADV_CHECK_CUDA(cudaMemcpyAsync(base + l.vb, vp.bounds.data(), vp.bounds.size() * 4, cudaMemcpyHostToDevice, st));
  ADV_CHECK_CUDA(cudaMemcpyAsync(base + l.vk, vp.coef.data(), vp.coef.size() * 4, cudaMemcpyHostToDevice, st));
  const int rows = vp.src1 - vp.src0;
  auto grid_of = [](long long work) {
    long long g_ = (work + 255) / 256;
    const long long cap = (long long)kNumSMs * 16;
    return (unsigned)(g_ < 1 ? 1 : (g_ > cap ? cap : g_));
  };
  resample_h_kernel<<<grid_of((long long)rows * crop), 256, 0, st>>>(
      img_hwc, w, vp.src0, rows, crop, reinterpret_cast<const int32_t*>(base + l.hb),
      reinterpret_cast<const int32_t*>(base + l.hk), hp.ksize, base + l.tmp);
  ADV_LAUNCHED();
  resample_v_kernel<<<grid_of((long long)crop * crop), 256, 0, st>>>(
      base + l.tmp, vp.src0, crop, crop, reinterpret_cast<const int32_t*>(base + l.vb),
      reinterpret_cast<const int32_t*>(base + l.vk), vp.ksize, out_chw);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

int advb200_to_u8_hwc(const float* x_chw, uint8_t* out_hwc, int h, int w, void* stream) {
  ADV_CHECK_ARG(h >= 0 && w >= 0, "to_u8: bad sizes");
  if (h == 0 || w == 0) return ADVB200_OK;
  ADV_CHECK_ARG(x_chw && out_hwc, "to_u8: null pointer");
  const long long px = (long long)h * w;
  long long g = (px + 255) / 256;
  if (g > kNumSMs * 16) g = kNumSMs * 16;
  to_u8_hwc_kernel<<<(unsigned)g, 256, 0, as_stream(stream)>>>(x_chw, out_hwc, px);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

}  // extern "C"

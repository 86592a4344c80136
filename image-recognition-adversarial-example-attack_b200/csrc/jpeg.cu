// JPEG defense on the GPU (defense_experiments.py:188-207 `_jpeg_compress_batch`).
//
// The reference round-trips every image through PIL's JPEG encoder and decoder on the host.  Entropy coding is
// lossless, so the pixels that come back are a pure function of the baseline-JPEG SAMPLE pipeline of libjpeg
// (4:2:0, quality-scaled Annex-K tables, jfdctint / jidctint "islow" integer DCTs, triangle chroma upsampling).
// That pipeline is restated here in integer arithmetic, bit for bit (oracle/jpeg_ref.py is the checker, pinned
// against PIL by tests/golden/jpeg_cases.npz):
//
//   kernel 1 (jpeg_codec_kernel): one CTA per strip of 16 rows x 256 columns (16 MCUs) of one image:
//     fp32 NCHW -> uint8 (x*255 truncated) -> YCbCr -> smem; 2x2 chroma box filter; then EIGHT LANES PER 8x8 BLOCK
//     take it through fdct -> quantise -> dequantise -> idct -> range limit (row / column passes exchanged through
//     shared memory) and write the decoded bytes back; decoded Y (full size) and Cb/Cr (half size) go to a uint8
//     scratch (1.5 B/pixel).
//   kernel 2 (jpeg_upsample_rgb_kernel): "fancy" h2v2 upsampling of the decoded chroma (needs the neighbouring
//     blocks' samples, hence the second pass), YCbCr -> RGB, /255 -> fp32 NCHW.
//
// HBM-bound integer/byte work: 12 B/pixel read + 12 B/pixel written + 3 B/pixel of scratch traffic = 27 B/pixel
// algorithmic (24 if the scratch stays in L2, which it does at 224x224: 19 MB per 256 images).
#include "common.cuh"

#include <mutex>

namespace advb200 {
namespace {

constexpr int kStripW = 256;          // pixels per strip (16 MCUs)
constexpr int kCodecThreads = 128;

struct JpegTables {
  uint32_t recip[2][64];   // ceil(2^32 / (q << 3)): exact division of the (< 2^16) magnitudes by umulhi
  uint16_t q[2][64];       // quantisation step (luma, chroma), natural (row-major) order
};

__device__ __forceinline__ int descale(int x, int n) { return (x + (1 << (n - 1))) >> n; }

constexpr int F_0_298 = 2446, F_0_390 = 3196, F_0_541 = 4433, F_0_765 = 6270, F_0_899 = 7373, F_1_175 = 9633,
              F_1_501 = 12299, F_1_847 = 15137, F_1_961 = 16069, F_2_053 = 16819, F_2_562 = 20995, F_3_072 = 25172;

// jfdctint.c, one 1-D pass over d[0], d[S], ..., d[7S].  kFirst: row pass (results scaled up by 4).
template <int S, bool kFirst>
__device__ __forceinline__ void fdct8(int* d) {
  const int t0 = d[0] + d[7 * S], t7 = d[0] - d[7 * S], t1 = d[S] + d[6 * S], t6 = d[S] - d[6 * S];
  const int t2 = d[2 * S] + d[5 * S], t5 = d[2 * S] - d[5 * S], t3 = d[3 * S] + d[4 * S], t4 = d[3 * S] - d[4 * S];
  const int t10 = t0 + t3, t13 = t0 - t3, t11 = t1 + t2, t12 = t1 - t2;
  constexpr int sh = kFirst ? 13 - 2 : 13 + 2;
  if (kFirst) { d[0] = (t10 + t11) << 2; d[4 * S] = (t10 - t11) << 2; }
  else { d[0] = descale(t10 + t11, 2); d[4 * S] = descale(t10 - t11, 2); }
  int z1 = (t12 + t13) * F_0_541;
  d[2 * S] = descale(z1 + t13 * F_0_765, sh);
  d[6 * S] = descale(z1 - t12 * F_1_847, sh);
  z1 = t4 + t7;
  int z2 = t5 + t6, z3 = t4 + t6, z4 = t5 + t7;
  const int z5 = (z3 + z4) * F_1_175;
  const int a4 = t4 * F_0_298, a5 = t5 * F_2_053, a6 = t6 * F_3_072, a7 = t7 * F_1_501;
  z1 = -z1 * F_0_899; z2 = -z2 * F_2_562; z3 = -z3 * F_1_961 + z5; z4 = -z4 * F_0_390 + z5;
  d[7 * S] = descale(a4 + z1 + z3, sh);
  d[5 * S] = descale(a5 + z2 + z4, sh);
  d[3 * S] = descale(a6 + z2 + z3, sh);
  d[S] = descale(a7 + z1 + z4, sh);
}

// jidctint.c, one 1-D pass.  kFirst: column pass into the workspace; second pass descales by 18 bits.
template <int S, bool kFirst>
__device__ __forceinline__ void idct8(int* d) {
  int z2 = d[2 * S], z3 = d[6 * S];
  int z1 = (z2 + z3) * F_0_541;
  int t2 = z1 - z3 * F_1_847, t3 = z1 + z2 * F_0_765;
  int t0 = (d[0] + d[4 * S]) << 13, t1 = (d[0] - d[4 * S]) << 13;
  const int t10 = t0 + t3, t13 = t0 - t3, t11 = t1 + t2, t12 = t1 - t2;
  t0 = d[7 * S]; t1 = d[5 * S]; t2 = d[3 * S]; t3 = d[S];
  z1 = t0 + t3; z2 = t1 + t2; z3 = t0 + t2;
  int z4 = t1 + t3;
  const int z5 = (z3 + z4) * F_1_175;
  t0 *= F_0_298; t1 *= F_2_053; t2 *= F_3_072; t3 *= F_1_501;
  z1 = -z1 * F_0_899; z2 = -z2 * F_2_562; z3 = -z3 * F_1_961 + z5; z4 = -z4 * F_0_390 + z5;
  t0 += z1 + z3; t1 += z2 + z4; t2 += z2 + z3; t3 += z1 + z4;
  constexpr int sh = kFirst ? 13 - 2 : 13 + 2 + 3;
  d[0] = descale(t10 + t3, sh); d[7 * S] = descale(t10 - t3, sh);
  d[S] = descale(t11 + t2, sh); d[6 * S] = descale(t11 - t2, sh);
  d[2 * S] = descale(t12 + t1, sh); d[5 * S] = descale(t12 - t1, sh);
  d[3 * S] = descale(t13 + t0, sh); d[4 * S] = descale(t13 - t0, sh);
}

// One 8x8 block by EIGHT lanes of a warp (lane l owns row l in the row passes and column l in the column passes), in
// place in shared memory (rows `pitch` bytes apart).  The two DCTs are separable, so the block goes
//   rows: fdct  ->  [exchange]  ->  columns: fdct, quantise, dequantise, idct  ->  [exchange]  ->  rows: idct, range limit
// with the 64 intermediate ints exchanged through `ws` (block stride 72 ints: the four blocks a warp works on fall on
// different banks).  One thread per block (the first version) left 32 of 128 threads idle and every busy thread with a
// ~2500-instruction dependent chain; here every lane has four 8-point passes (~220 instructions) and all lanes work.
// Same integer arithmetic in the same order per 1-D pass, hence the same bits.
constexpr int kWsStride = 72;
__device__ __forceinline__ void codec_block8(uint8_t* blk, int pitch, int* ws, const uint16_t* q, const uint32_t* recip,
                                             int l, bool active) {
  int d[8];
  if (active) {
    const uint2 v = *reinterpret_cast<const uint2*>(blk + l * pitch);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      d[k] = (int)((v.x >> (8 * k)) & 0xffu) - 128;
      d[4 + k] = (int)((v.y >> (8 * k)) & 0xffu) - 128;
    }
    fdct8<1, true>(d);
    *reinterpret_cast<int4*>(ws + l * 8) = make_int4(d[0], d[1], d[2], d[3]);
    *reinterpret_cast<int4*>(ws + l * 8 + 4) = make_int4(d[4], d[5], d[6], d[7]);
  }
  __syncwarp();
  if (active) {
#pragma unroll
    for (int r = 0; r < 8; ++r) d[r] = ws[r * 8 + l];
    fdct8<1, false>(d);
    // jcdctmgr.c: divisor q<<3 (the islow DCT leaves a factor 8), round half away from zero; then dequantise
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const int qq = q[r * 8 + l];
      const int v = d[r];
      const uint32_t a = (uint32_t)(v < 0 ? -v : v) + (uint32_t)(qq << 2);
      const int mag = (int)__umulhi(a, recip[r * 8 + l]);
      d[r] = (v < 0 ? -mag : mag) * qq;
    }
    idct8<1, true>(d);
#pragma unroll
    for (int r = 0; r < 8; ++r) ws[r * 8 + l] = d[r];
  }
  __syncwarp();
  if (active) {
    const int4 a = *reinterpret_cast<const int4*>(ws + l * 8), b = *reinterpret_cast<const int4*>(ws + l * 8 + 4);
    d[0] = a.x; d[1] = a.y; d[2] = a.z; d[3] = a.w; d[4] = b.x; d[5] = b.y; d[6] = b.z; d[7] = b.w;
    idct8<1, false>(d);
    uint2 v = make_uint2(0u, 0u);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      v.x |= (uint32_t)__vimin_s32_relu(d[k] + 128, 255) << (8 * k);   // range limit: one VIMNMX.RELU
      v.y |= (uint32_t)__vimin_s32_relu(d[4 + k] + 128, 255) << (8 * k);
    }
    *reinterpret_cast<uint2*>(blk + l * pitch) = v;
  }
}

__device__ __forceinline__ int to_u8(float x) {
  // ToPILImage on a float tensor: clamp (done by the caller of the reference function), mul(255), .byte() = truncation
  return (int)(fminf(fmaxf(x, 0.f), 1.f) * 255.f);
}

__global__ void __launch_bounds__(kCodecThreads, 8)
jpeg_codec_kernel(const float* __restrict__ x, uint8_t* __restrict__ ydec, uint8_t* __restrict__ cbdec,
                  uint8_t* __restrict__ crdec, int h, int w, int strips, int mcu_rows, const JpegTables tb) {
  // Row pitches padded by 32 bytes: the eight lanes of a block touch eight consecutive rows at the same column, and with
  // pitches of 256 / 128 bytes those fell into the same banks (8-way conflict on every block load and store; ncu: 3.5 M
  // conflict cycles).  With 288 / 160 bytes, row l starts 8 l banks further on and a warp's 256-byte access takes the
  // minimum of two wavefronts.
  constexpr int kPitchY = kStripW + 32, kPitchC = kStripW / 2 + 32;
  __shared__ __align__(16) uint8_t sY[16][kPitchY];
  __shared__ __align__(16) uint8_t sCbF[16][kStripW];     // full-resolution chroma of the strip (three-stage path only)
  __shared__ __align__(16) uint8_t sCrF[16][kStripW];
  __shared__ __align__(16) uint8_t sC[2][8][kPitchC];      // downsampled, then decoded in place
  __shared__ __align__(16) int sW[(kCodecThreads / 8) * kWsStride];   // per 8-lane group: one block's 64 intermediates
  __shared__ uint16_t sQ[2][64];                            // the tables, indexed by lane (constant memory would serialise)
  __shared__ uint32_t sRecip[2][64];
  for (int i = threadIdx.x; i < 128; i += kCodecThreads) { sQ[i >> 6][i & 63] = tb.q[i >> 6][i & 63]; sRecip[i >> 6][i & 63] = tb.recip[i >> 6][i & 63]; }
  const int strip = blockIdx.x % strips;
  const int mr = (blockIdx.x / strips) % mcu_rows;
  const long long img = blockIdx.x / (strips * mcu_rows);
  const int col0 = strip * kStripW, row0 = mr * 16;
  const int hc = (h + 1) >> 1, wc = (w + 1) >> 1;
  const float* xr = x + img * 3 * (long long)h * w;
  const float* xg = xr + (long long)h * w;
  const float* xb = xg + (long long)h * w;

  // ---- load, quantise to bytes, colour-convert (jccolor.c); rows / columns past the image replicate its edge
  auto convert = [&](int r, int c, float fr, float fg, float fb) {
    const int R = to_u8(fr), G = to_u8(fg), B = to_u8(fb);
    sY[r][c] = (uint8_t)((19595 * R + 38470 * G + 7471 * B + 32768) >> 16);
    sCbF[r][c] = (uint8_t)((-11059 * R - 21709 * G + 32768 * B + (128 << 16) + 32767) >> 16);
    sCrF[r][c] = (uint8_t)((32768 * R - 27439 * G - 5329 * B + (128 << 16) + 32767) >> 16);
  };
  const bool vec = ((w & 3) == 0) && ((reinterpret_cast<uintptr_t>(x) & 15u) == 0);
  // Strips whose 16 rows all lie inside the image (every strip when h % 16 == 0) take the FUSED front end: a thread
  // converts a 2-row x 4-column patch from six 16-byte loads, keeps the full-resolution chroma in registers, applies
  // the 2x2 box filter there and writes 8 Y bytes as two words and 2 + 2 chroma samples as two half-words.  No
  // full-resolution chroma planes in shared memory, no single-byte stores, no second pass, no index divisions; columns
  // past the image edge are filled only up to the next MCU boundary (nothing beyond it is ever read).  The bottom
  // strip of other heights keeps the three-stage path below (its chroma rows past the real chroma height replicate a
  // DIFFERENT row pair than its luma rows do).
  const bool fused = vec && row0 + 16 <= h;
  if (fused) {
    const int wrem = min(kStripW, w - col0);
    const int gw = wrem >> 2;                                        // 4-pixel groups inside the image
    const int ng = min(kStripW / 4, ((wrem + 15) >> 4) << 2);        // groups up to the MCU boundary
    const int cg = threadIdx.x & 63;
    if (cg < ng) {
      const int cge = min(cg, gw - 1);                               // groups past the edge re-read the last real group ...
      const bool rep = cg >= gw;                                     // ... and replicate its last pixel
      const float* src = xr + (long long)row0 * w + col0 + 4 * cge;
      const long long pl = (long long)h * w;
#pragma unroll
      for (int it = 0; it < 4; it += 2) {                            // two row pairs per pass: twelve loads in flight
        float4 v[2][2][3];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int rp = (it + u) * 2 + (threadIdx.x >> 6);
#pragma unroll
          for (int dr = 0; dr < 2; ++dr)
#pragma unroll
            for (int c = 0; c < 3; ++c)
              v[u][dr][c] = __ldg(reinterpret_cast<const float4*>(src + c * pl + (long long)(2 * rp + dr) * w));
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int rp = (it + u) * 2 + (threadIdx.x >> 6);
          int cb[2][4], cr[2][4];
#pragma unroll
          for (int dr = 0; dr < 2; ++dr) {
            float f[3][4];
#pragma unroll
            for (int c = 0; c < 3; ++c) {
              const float4 q = v[u][dr][c];
              f[c][0] = rep ? q.w : q.x; f[c][1] = rep ? q.w : q.y; f[c][2] = rep ? q.w : q.z; f[c][3] = q.w;
            }
            uint32_t yw = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const int R = to_u8(f[0][k]), G = to_u8(f[1][k]), B = to_u8(f[2][k]);
              // jccolor.c; all three results lie in 0..255 for R, G, B in 0..255 (the uint8 casts of the staged path
              // never truncate)
              yw |= (uint32_t)((19595 * R + 38470 * G + 7471 * B + 32768) >> 16) << (8 * k);
              cb[dr][k] = (-11059 * R - 21709 * G + 32768 * B + (128 << 16) + 32767) >> 16;
              cr[dr][k] = (32768 * R - 27439 * G - 5329 * B + (128 << 16) + 32767) >> 16;
            }
            *reinterpret_cast<uint32_t*>(&sY[2 * rp + dr][4 * cg]) = yw;
          }
          // h2v2 box filter with the alternating 1, 2 bias (jcsample.c): chroma columns 2cg (even) and 2cg + 1 (odd)
          const uint32_t b0 = (uint32_t)(cb[0][0] + cb[0][1] + cb[1][0] + cb[1][1] + 1) >> 2;
          const uint32_t b1 = (uint32_t)(cb[0][2] + cb[0][3] + cb[1][2] + cb[1][3] + 2) >> 2;
          const uint32_t r0 = (uint32_t)(cr[0][0] + cr[0][1] + cr[1][0] + cr[1][1] + 1) >> 2;
          const uint32_t r1 = (uint32_t)(cr[0][2] + cr[0][3] + cr[1][2] + cr[1][3] + 2) >> 2;
          *reinterpret_cast<uint16_t*>(&sC[0][rp][2 * cg]) = (uint16_t)(b0 | (b1 << 8));
          *reinterpret_cast<uint16_t*>(&sC[1][rp][2 * cg]) = (uint16_t)(r0 | (r1 << 8));
        }
      }
    }
    __syncthreads();
  } else {
    if (vec) {
      // 16-byte loads, two groups (six loads) in flight per thread; groups wholly past the right edge replicate the
      // last pixel, which the scalar tail below fills in
      const int gw = min(kStripW, w - col0) >> 2;                  // whole 4-pixel groups of this strip inside the image
      for (int i0 = threadIdx.x; i0 < 16 * gw; i0 += 2 * kCodecThreads) {
        float4 v[2][3];
        int rr[2], cc[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int i = i0 + u * kCodecThreads;
          rr[u] = -1;
          if (i < 16 * gw) {
            rr[u] = i / gw; cc[u] = (i % gw) * 4;
            const long long o = (long long)min(row0 + rr[u], h - 1) * w + col0 + cc[u];
            v[u][0] = __ldg(reinterpret_cast<const float4*>(xr + o));
            v[u][1] = __ldg(reinterpret_cast<const float4*>(xg + o));
            v[u][2] = __ldg(reinterpret_cast<const float4*>(xb + o));
          }
        }
#pragma unroll
        for (int u = 0; u < 2; ++u)
          if (rr[u] >= 0) {
            convert(rr[u], cc[u], v[u][0].x, v[u][1].x, v[u][2].x);
            convert(rr[u], cc[u] + 1, v[u][0].y, v[u][1].y, v[u][2].y);
            convert(rr[u], cc[u] + 2, v[u][0].z, v[u][1].z, v[u][2].z);
            convert(rr[u], cc[u] + 3, v[u][0].w, v[u][1].w, v[u][2].w);
          }
      }
      const int done = gw * 4, rest = kStripW - done;              // columns past the image edge (replicated)
      for (int i = threadIdx.x; i < 16 * rest; i += kCodecThreads) {
        const int r = i / rest, c = done + i % rest;
        const long long o = (long long)min(row0 + r, h - 1) * w + (w - 1);
        convert(r, c, __ldg(xr + o), __ldg(xg + o), __ldg(xb + o));
      }
    } else {
      for (int i = threadIdx.x; i < 16 * kStripW; i += kCodecThreads) {
        const int r = i / kStripW, c = i % kStripW;
        const long long o = (long long)min(row0 + r, h - 1) * w + min(col0 + c, w - 1);
        convert(r, c, __ldg(xr + o), __ldg(xg + o), __ldg(xb + o));
      }
    }
    __syncthreads();
    // ---- h2v2 box filter with the alternating 1,2 bias (jcsample.c); chroma rows past the real chroma height
    //      replicate the last real chroma row (jcprepct.c pads AFTER downsampling)
    for (int i = threadIdx.x; i < 2 * 8 * (kStripW / 2); i += kCodecThreads) {
      const int comp = i / (8 * (kStripW / 2)), rr = (i / (kStripW / 2)) % 8, cc = i % (kStripW / 2);
      const int cr_real = min(mr * 8 + rr, hc - 1) - mr * 8;
      const uint8_t(*f)[kStripW] = comp ? sCrF : sCbF;
      const int s = f[2 * cr_real][2 * cc] + f[2 * cr_real][2 * cc + 1] + f[2 * cr_real + 1][2 * cc] +
                    f[2 * cr_real + 1][2 * cc + 1] + 1 + (cc & 1);
      sC[comp][rr][cc] = (uint8_t)(s >> 2);
    }
    __syncthreads();
  }
  // ---- eight lanes per 8x8 block: 64 luma blocks, then 16 + 16 chroma blocks (6 rounds of 16 blocks per CTA)
  {
    const int grp = threadIdx.x >> 3, l = threadIdx.x & 7;
    int* ws = sW + grp * kWsStride;
    for (int b = grp; b < 96; b += kCodecThreads / 8) {
      if (b < 64) {
        const int by = b >> 5, bx = b & 31;
        codec_block8(&sY[by * 8][bx * 8], kPitchY, ws, sQ[0], sRecip[0], l, col0 + bx * 8 < w && row0 + by * 8 < h);
      } else {
        const int comp = (b - 64) >> 4, bx = b & 15;
        codec_block8(&sC[comp][0][bx * 8], kPitchC, ws, sQ[1], sRecip[1], l, (col0 >> 1) + bx * 8 < wc);
      }
      __syncwarp();                                  // the group's workspace is reused by its next block
    }
  }
  __syncthreads();
  // ---- decoded bytes -> scratch (real pixels only); 4 bytes per store when the rows are 4-byte aligned
  uint8_t* yo = ydec + img * (long long)h * w;
  if ((w & 3) == 0) {
    for (int i = threadIdx.x; i < 16 * (kStripW / 4); i += kCodecThreads) {
      const int r = i / (kStripW / 4), c = (i % (kStripW / 4)) * 4;
      if (row0 + r < h && col0 + c < w)
        *reinterpret_cast<uint32_t*>(yo + (long long)(row0 + r) * w + col0 + c) = *reinterpret_cast<const uint32_t*>(&sY[r][c]);
    }
  } else {
    for (int i = threadIdx.x; i < 16 * kStripW; i += kCodecThreads) {
      const int r = i / kStripW, c = i % kStripW;
      if (row0 + r < h && col0 + c < w) yo[(long long)(row0 + r) * w + col0 + c] = sY[r][c];
    }
  }
  const bool cvec = ((wc & 3) == 0) && (((reinterpret_cast<uintptr_t>(cbdec) | reinterpret_cast<uintptr_t>(crdec)) & 3u) == 0);
  for (int i = threadIdx.x; i < 2 * 8 * (kStripW / 2) / (cvec ? 4 : 1); i += kCodecThreads) {
    const int e = cvec ? i * 4 : i;
    const int comp = e / (8 * (kStripW / 2)), rr = (e / (kStripW / 2)) % 8, cc = e % (kStripW / 2);
    const int gr = mr * 8 + rr, gc = (col0 >> 1) + cc;
    if (gr < hc && gc < wc) {
      uint8_t* dst = (comp ? crdec : cbdec) + img * (long long)hc * wc + (long long)gr * wc + gc;
      if (cvec) *reinterpret_cast<uint32_t*>(dst) = *reinterpret_cast<const uint32_t*>(&sC[comp][rr][cc]);
      else *dst = sC[comp][rr][cc];
    }
  }
}

// colsum of jdsample.c h2v2_fancy_upsample: 3 * nearer row + farther row, at chroma column j (clamped = the
// source's special first / last column cases, where the neighbour term degenerates to the column itself)
__device__ __forceinline__ int colsum(const uint8_t* near, const uint8_t* far, int j) { return 3 * near[j] + far[j]; }

template <typename Idx>
__global__ void __launch_bounds__(256)
jpeg_upsample_rgb_kernel(const uint8_t* __restrict__ ydec, const uint8_t* __restrict__ cbdec,
                         const uint8_t* __restrict__ crdec, float* __restrict__ out, long long n_img, int h, int w) {
  // ToTensor's k / 255 takes 256 values: a shared-memory table filled with the same true division replaces twelve IEEE
  // division sequences per thread (the kernel is instruction-bound: ncu issue-active 83 %); bit-identical by construction
  __shared__ float tab255[256];
  tab255[threadIdx.x] = __fdiv_rn((float)threadIdx.x, 255.f);
  __syncthreads();
  const int hc = (h + 1) >> 1, wc = (w + 1) >> 1, wq = (w + 3) >> 2;
  const Idx total = (Idx)(n_img * h * wq);                 // one thread per four horizontal output pixels
  const bool vec = ((w & 3) == 0) && ((reinterpret_cast<uintptr_t>(out) & 15u) == 0) &&
                   ((reinterpret_cast<uintptr_t>(ydec) & 3u) == 0);
  for (Idx i = (Idx)blockIdx.x * 256 + threadIdx.x; i < total; i += (Idx)gridDim.x * 256) {
    const Idx iq = i / (Idx)wq;                            // (32-bit index math whenever the thread count fits 31 bits)
    const int q = (int)(i - iq * (Idx)wq);
    const Idx ir = iq / (Idx)h;
    const int r = (int)(iq - ir * (Idx)h);
    const long long img = (long long)ir;
    const int rc = r >> 1;
    const int rfar = (r & 1) ? min(rc + 1, hc - 1) : max(rc - 1, 0);   // context rows replicate at the edges
    const int j0 = 2 * q;                                               // chroma columns j0, j0 + 1
    int up[2][4];
#pragma unroll
    for (int comp = 0; comp < 2; ++comp) {
      const uint8_t* base = (comp ? crdec : cbdec) + img * (long long)hc * wc;
      const uint8_t* near = base + (long long)rc * wc;
      const uint8_t* far = base + (long long)rfar * wc;
      int s[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) s[k] = colsum(near, far, min(max(j0 - 1 + k, 0), wc - 1));
      up[comp][0] = (3 * s[1] + s[0] + 8) >> 4;
      up[comp][1] = (3 * s[1] + s[2] + 7) >> 4;
      // second chroma column (absent when w % 4 is 1 or 2): its "next" neighbour clamps to itself at the edge
      const int s3 = (j0 + 1 <= wc - 1) ? s[3] : s[2];
      up[comp][2] = (3 * s[2] + s[1] + 8) >> 4;
      up[comp][3] = (3 * s[2] + s3 + 7) >> 4;
    }
    const uint8_t* yrow = ydec + (img * h + r) * (long long)w;
    float* o = out + img * 3 * (long long)h * w + (long long)r * w;
    const long long plane = (long long)h * w;
    const int c0 = 4 * q;
    int yv[4];
    if (vec) {
      const uint32_t y4 = *reinterpret_cast<const uint32_t*>(yrow + c0);
#pragma unroll
      for (int k = 0; k < 4; ++k) yv[k] = (y4 >> (8 * k)) & 0xff;
    } else {
#pragma unroll
      for (int k = 0; k < 4; ++k) yv[k] = c0 + k < w ? yrow[c0 + k] : 0;
    }
    float rgb[3][4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int y = yv[k], xb = up[0][k] - 128, xr = up[1][k] - 128;
      // jdcolor.c build_ycc_rgb_table / ycc_rgb_convert
      const int R = y + ((91881 * xr + 32768) >> 16);
      const int B = y + ((116130 * xb + 32768) >> 16);
      const int G = y + ((-22554 * xb + 32768 - 46802 * xr) >> 16);
      rgb[0][k] = tab255[min(max(R, 0), 255)];                          // ToTensor: div(255)
      rgb[1][k] = tab255[min(max(G, 0), 255)];
      rgb[2][k] = tab255[min(max(B, 0), 255)];
    }
#pragma unroll
    for (int p = 0; p < 3; ++p) {
      if (vec) {
        *reinterpret_cast<float4*>(o + p * plane + c0) = make_float4(rgb[p][0], rgb[p][1], rgb[p][2], rgb[p][3]);
      } else {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (c0 + k < w) o[p * plane + c0 + k] = rgb[p][k];
      }
    }
  }
}

// ToTensor's k / 255 for an integer 0 <= k <= 255, correctly rounded, in registers: k becomes a float through the
// 2^23 bit pattern (no conversion-pipe instruction), then Markstein's correction as in elementwise.cu::div9 -- with
// y = RN(1/255), q0 = RN(k y), the remainder r = k - 255 q0 is exact in one FMA and RN(q0 + r y) = RN(k / 255).  All 256
// values equal __fdiv_rn(k, 255) (checked on the host with fmaf, oracle/check_div9.c, and by the bit-exactness tests).
// The fast second pass uses it instead of the generic kernel's shared-memory table: its 48 lookups per thread had
// data-dependent indices, collided in the banks and kept the LSU queue full (ncu: mio-throttle, 2.8 M conflict cycles).
__device__ __forceinline__ float div255(int k) {
  const float y = 0.00392156886f;                                   // RN(1/255) = 0x1.010102p-8
  const float kf = __int_as_float(0x4B000000 | k) - 8388608.0f;
  const float q0 = __fmul_rn(kf, y);
  const float r = __fmaf_rn(-255.0f, q0, kf);
  return __fmaf_rn(r, y, q0);
}

// Fast path of the second pass for w % 8 == 0 (every ResNet input): one thread per 2 output rows x 8 output columns.
// The two rows 2rc, 2rc+1 share their nearer chroma row rc, so the three chroma rows rc-1, rc, rc+1 are fetched once
// (one aligned 4-byte word + the two neighbour bytes per row and component: 18 loads for 16 pixels, where the generic
// kernel issues 16 single-byte loads for 4), the thread index is split by two multiply-shift divisions, and the Y row
// arrives as one 8-byte load.  Same integer arithmetic per sample as jpeg_upsample_rgb_kernel, hence the same bits.
__global__ void __launch_bounds__(256)
jpeg_upsample_rgb8_kernel(const uint8_t* __restrict__ ydec, const uint8_t* __restrict__ cbdec,
                          const uint8_t* __restrict__ crdec, float* __restrict__ out, uint32_t total, int h, int w,
                          const FastDiv div_w8, const FastDiv div_hc) {
  const int hc = (h + 1) >> 1, wc = w >> 1, w8 = w >> 3;
  const size_t plane = (size_t)h * w, cplane = (size_t)hc * wc;
  for (uint32_t i = blockIdx.x * 256u + threadIdx.x; i < total; i += gridDim.x * 256u) {
    const uint32_t iq = fast_div(i, div_w8);
    const int q = (int)(i - iq * (uint32_t)w8);
    const uint32_t img = fast_div(iq, div_hc);
    const int rc = (int)(iq - img * (uint32_t)hc);
    const int j0 = 4 * q;
    const int jl = max(j0 - 1, 0), jr = min(j0 + 4, wc - 1);           // neighbour columns replicate at the edges
    const int rows[3] = {max(rc - 1, 0), rc, min(rc + 1, hc - 1)};     // context rows replicate at the edges
    // c[comp][row][k]: chroma sample of column j0 - 1 + k
    int c[2][3][6];
#pragma unroll
    for (int comp = 0; comp < 2; ++comp) {
      const uint8_t* base = (comp ? crdec : cbdec) + img * cplane;
#pragma unroll
      for (int rr = 0; rr < 3; ++rr) {
        const uint8_t* row = base + (size_t)rows[rr] * wc;
        const uint32_t m = *reinterpret_cast<const uint32_t*>(row + j0);
        c[comp][rr][0] = row[jl];
        c[comp][rr][1] = m & 0xff; c[comp][rr][2] = (m >> 8) & 0xff; c[comp][rr][3] = (m >> 16) & 0xff; c[comp][rr][4] = m >> 24;
        c[comp][rr][5] = row[jr];
      }
    }
    const int c0 = 8 * q;
#pragma unroll
    for (int par = 0; par < 2; ++par) {
      const int r = 2 * rc + par;
      if (r >= h) break;                                                 // odd heights: the last chroma row has one output row
      const uint2 y8 = *reinterpret_cast<const uint2*>(ydec + (img * (size_t)h + r) * w + c0);
      int up[2][8];
#pragma unroll
      for (int comp = 0; comp < 2; ++comp) {
        int sc[6];                                                      // jdsample.c colsum: 3 * nearer row + farther row
#pragma unroll
        for (int k = 0; k < 6; ++k) sc[k] = 3 * c[comp][1][k] + c[comp][par ? 2 : 0][k];
#pragma unroll
        for (int k = 1; k <= 4; ++k) {
          up[comp][2 * k - 2] = (3 * sc[k] + sc[k - 1] + 8) >> 4;
          up[comp][2 * k - 1] = (3 * sc[k] + sc[k + 1] + 7) >> 4;
        }
      }
      float rgb[3][8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int y = (int)(((k < 4 ? y8.x : y8.y) >> (8 * (k & 3))) & 0xff);
        const int xb = up[0][k] - 128, xr = up[1][k] - 128;
        // jdcolor.c build_ycc_rgb_table / ycc_rgb_convert
        const int R = y + ((91881 * xr + 32768) >> 16);
        const int B = y + ((116130 * xb + 32768) >> 16);
        const int G = y + ((-22554 * xb + 32768 - 46802 * xr) >> 16);
        rgb[0][k] = div255(__vimin_s32_relu(R, 255));                          // ToTensor: div(255)
        rgb[1][k] = div255(__vimin_s32_relu(G, 255));
        rgb[2][k] = div255(__vimin_s32_relu(B, 255));
      }
      float* o = out + img * 3 * plane + (size_t)r * w + c0;
#pragma unroll
      for (int pl = 0; pl < 3; ++pl) {
        *reinterpret_cast<float4*>(o + pl * plane) = make_float4(rgb[pl][0], rgb[pl][1], rgb[pl][2], rgb[pl][3]);
        *reinterpret_cast<float4*>(o + pl * plane + 4) = make_float4(rgb[pl][4], rgb[pl][5], rgb[pl][6], rgb[pl][7]);
      }
    }
  }
}

int build_tables(int quality, JpegTables* tb) {
  static const uint8_t kLuma[64] = {16, 11, 10, 16, 24,  40,  51,  61,  12, 12, 14, 19, 26,  58,  60,  55,
                                    14, 13, 16, 24, 40,  57,  69,  56,  14, 17, 22, 29, 51,  87,  80,  62,
                                    18, 22, 37, 56, 68,  109, 103, 77,  24, 35, 55, 64, 81,  104, 113, 92,
                                    49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99};
  static const uint8_t kChroma[64] = {17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99,
                                      24, 26, 56, 99, 99, 99, 99, 99, 47, 66, 99, 99, 99, 99, 99, 99,
                                      99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99,
                                      99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99};
  // jcparam.c jpeg_quality_scaling + jpeg_add_quant_table(force_baseline)
  const int q = quality < 1 ? 1 : (quality > 100 ? 100 : quality);
  const int scale = q < 50 ? 5000 / q : 200 - 2 * q;
  for (int comp = 0; comp < 2; ++comp)
    for (int i = 0; i < 64; ++i) {
      long v = ((long)(comp ? kChroma[i] : kLuma[i]) * scale + 50) / 100;
      v = v < 1 ? 1 : (v > 255 ? 255 : v);
      tb->q[comp][i] = (uint16_t)v;
      const uint64_t d = (uint64_t)v << 3;
      tb->recip[comp][i] = (uint32_t)(((1ull << 32) + d - 1) / d);   // exact for numerators < 2^21
    }
  return 0;
}

}  // namespace
}  // namespace advb200

using namespace advb200;

extern "C" {

int64_t advb200_jpeg_scratch_bytes(int n_img, int h, int w) {
  if (n_img < 0 || h <= 0 || w <= 0) return -1;
  const int64_t hc = (h + 1) / 2, wc = (w + 1) / 2;
  return (int64_t)n_img * ((int64_t)h * w + 2 * hc * wc) + 256;
}

int advb200_jpeg_roundtrip_f32(const float* x, float* out, int n_img, int h, int w, int quality, void* scratch,
                               int64_t scratch_bytes, void* stream) {
  ADV_CHECK_ARG(n_img >= 0 && h > 0 && w > 0, "jpeg: bad sizes");
  ADV_CHECK_ARG(quality >= 1 && quality <= 100, "jpeg: quality must be in 1..100");
  if (n_img == 0) return ADVB200_OK;
  ADV_CHECK_ARG(x && out && scratch, "jpeg: null pointer");
  ADV_CHECK_ARG(scratch_bytes >= advb200_jpeg_scratch_bytes(n_img, h, w), "jpeg: scratch too small");
  JpegTables tb;
  build_tables(quality, &tb);
  const int64_t hc = (h + 1) / 2, wc = (w + 1) / 2;
  uint8_t* ydec = static_cast<uint8_t*>(scratch);
  uint8_t* cbdec = ydec + (int64_t)n_img * h * w;
  uint8_t* crdec = cbdec + (int64_t)n_img * hc * wc;
  const int strips = (w + kStripW - 1) / kStripW, mcu_rows = (h + 15) / 16;
  const int64_t ctas = (int64_t)n_img * strips * mcu_rows;
  ADV_CHECK_ARG(ctas < (1ll << 31), "jpeg: batch too large for one launch");
  cudaStream_t st = as_stream(stream);
  jpeg_codec_kernel<<<(unsigned)ctas, kCodecThreads, 0, st>>>(x, ydec, cbdec, crdec, h, w, strips, mcu_rows, tb);
  ADV_LAUNCHED();
  const int64_t cap = (int64_t)kNumSMs * 32;
  // fast path: rows of whole 8-pixel groups, aligned planes, 31-bit thread index
  const int64_t patches = (int64_t)n_img * hc * (w >> 3);
  static const bool no_fast = getenv("ADVB200_JPEG_NO_FAST") != nullptr;      // A/B switch for the tests
  if (!no_fast && (w & 7) == 0 && (reinterpret_cast<uintptr_t>(out) & 15u) == 0 && (reinterpret_cast<uintptr_t>(ydec) & 7u) == 0 &&
      patches + cap * 256 < (1ll << 31)) {
    int64_t g8 = (patches + 255) / 256;
    if (g8 > cap) g8 = cap;
    jpeg_upsample_rgb8_kernel<<<(unsigned)g8, 256, 0, st>>>(ydec, cbdec, crdec, out, (uint32_t)patches, h, w,
                                                           make_fastdiv((uint32_t)(w >> 3)), make_fastdiv((uint32_t)hc));
    ADV_LAUNCHED();
    return ADVB200_OK;
  }
  const int64_t quads = (int64_t)n_img * h * ((w + 3) / 4);
  int64_t grid = (quads + 255) / 256;
  if (grid > cap) grid = cap;
  if (quads + grid * 256 < (1ll << 31))
    jpeg_upsample_rgb_kernel<uint32_t><<<(unsigned)grid, 256, 0, st>>>(ydec, cbdec, crdec, out, n_img, h, w);
  else
    jpeg_upsample_rgb_kernel<long long><<<(unsigned)grid, 256, 0, st>>>(ydec, cbdec, crdec, out, n_img, h, w);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

}  // extern "C"

// ---- allocation-free variant ---------------------------------------------------------------------------------------
// For callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc): the decoded
// Y / Cb / Cr planes between the two kernels live in a file-scope device array instead of a caller buffer.  Every GPU
// that loads the library holds these ADVB200_JPEG_NOALLOC_MAX_BYTES until the process exits.  In host code the array's
// name is only the host-side handle of the symbol, NOT a device address: the address is looked up once per device,
// with that device current, by cudaGetSymbolAddress, and that pointer is what the kernels are given.
namespace advb200 {
namespace {

static __device__ __align__(256) uint8_t g_jpeg_noalloc_scratch[ADVB200_JPEG_NOALLOC_MAX_BYTES];

constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mu;
void* g_noalloc_ptr[kMaxDevices] = {};     // device address of the array on device i (0 = not looked up yet)

// device address of the scratch array on the CURRENT device; no allocation, one symbol lookup per device
int noalloc_scratch(void** out) {
  int dev = -1;
  ADV_CHECK_CUDA(cudaGetDevice(&dev));
  ADV_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "jpeg_noalloc: device ordinal out of range");
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (!g_noalloc_ptr[dev]) {
    void* ptr = nullptr;
    ADV_CHECK_CUDA(cudaGetSymbolAddress(&ptr, g_jpeg_noalloc_scratch));
    if (!ptr) return fail(ADVB200_ERR_CUDA, "jpeg_noalloc: cudaGetSymbolAddress returned a null address");
    g_noalloc_ptr[dev] = ptr;
  }
  *out = g_noalloc_ptr[dev];
  return ADVB200_OK;
}

}  // namespace
}  // namespace advb200

extern "C" int advb200_jpeg_roundtrip_f32_noalloc(const float* x, float* out, int n_img, int h, int w, int quality,
                                                  void* stream) {
  ADV_CHECK_ARG(n_img >= 0 && h > 0 && w > 0, "jpeg_noalloc: bad sizes");
  const int64_t need = advb200_jpeg_scratch_bytes(n_img, h, w);
  if (need > (int64_t)ADVB200_JPEG_NOALLOC_MAX_BYTES)
    return fail(ADVB200_ERR_ARG, "jpeg_noalloc: " + std::to_string(n_img) + " x " + std::to_string(h) + " x " +
                                     std::to_string(w) + " needs " + std::to_string(need) +
                                     " bytes of scratch, the built-in array holds " +
                                     std::to_string((long long)ADVB200_JPEG_NOALLOC_MAX_BYTES) +
                                     "; split the batch or use advb200_jpeg_roundtrip_f32 with a caller buffer");
  void* scratch = nullptr;
  const int rc = noalloc_scratch(&scratch);
  if (rc != ADVB200_OK) return rc;
  // same two kernels, same launch geometry as the caller-buffer entry point.  Nothing in the array has to start at
  // zero: every byte jpeg_upsample_rgb_kernel reads was written by jpeg_codec_kernel earlier in this same call.
  return advb200_jpeg_roundtrip_f32(x, out, n_img, h, w, quality, scratch, (int64_t)ADVB200_JPEG_NOALLOC_MAX_BYTES, stream);
}

// tcgen05 / TMEM / TMA implicit-GEMM convolution kernel (sm_100a only).
//
// Warp roles (192 threads, 1 CTA per SM, persistent over tiles):
//   warp 0      : TMA producer   (one lane) — fills the smem ring with A (activation box) and
//                                  B (packed weight) tiles, 128B-swizzled, mbarrier complete_tx
//   warp 1      : TMEM allocator + MMA issuer (one lane) — tcgen05.mma.cta_group::1.kind::f16,
//                                  M=128, N=BN, K=16, accumulators double-buffered in TMEM,
//                                  tcgen05.commit releases smem stages / publishes accumulators
//   warps 2..5  : epilogue        — tcgen05.ld 32x32b (thread = output pixel), fused bias /
//                                  residual / ReLU / ReLU-mask, 16-byte global stores
#include "tc_conv.cuh"

#include <cudaTypedefs.h>

#include <mutex>

namespace advb200 {

namespace {

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;                 // 64 bf16 = 128 bytes = one swizzle row
constexpr int kATileBytes = kBlockM * kBlockK * 2;   // 16 KB
constexpr int kThreads = 192;
constexpr int kEpiWarp0 = 2;

template <int BN> struct Cfg {
  static constexpr int kBTileBytes = BN * kBlockK * 2;
  static constexpr int kStageBytes = kATileBytes + kBTileBytes;
  static constexpr int kStages = (BN == 256) ? 4 : (BN == 128 ? 6 : 8);
  static constexpr int kTmemCols = (2 * BN < 32) ? 32 : 2 * BN;  // two accumulators
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
};

struct TmapPack {
  CUtensorMap a[4];
  CUtensorMap b;
};

// ---------------------------------------------------------------- PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(const CUtensorMap* map, uint64_t* bar, void* dst, int c0,
                                            int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, uint64_t* bar, void* dst, int c0,
                                            int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]; bf16 x bf16 -> fp32
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc,
                                          uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on an mbarrier when all previously issued MMAs have completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]),
        "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]),
        "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]),
        "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]),
        "=r"(v[15])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// K-major, 128B-swizzled operand tile: rows of 128 bytes, 8-row atoms of 1024 bytes.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);   // start address, 16-byte units
  d |= (uint64_t)1 << 16;                        // leading byte offset (unused for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;              // stride byte offset: 8 rows * 128 B
  d |= (uint64_t)1 << 46;                        // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;                        // SWIZZLE_128B
  return d;
}

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float bf16_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }

// ---------------------------------------------------------------- the kernel
template <int BN>
__global__ void __launch_bounds__(kThreads, 1)
conv_gemm_kernel(const __grid_constant__ TmapPack maps, const __grid_constant__ ConvGemmParams p) {
  using C = Cfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::kStages * C::kStageBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + C::kStages;
  uint64_t* tfull_bar = bars + 2 * C::kStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    for (int i = 0; i < 4; ++i) prefetch_tmap(&maps.a[i]);
    prefetch_tmap(&maps.b);
    for (int s = 0; s < C::kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tfull_bar[a], 1); mbar_init(&tempty_bar[a], 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(tmem_slot)),
                 "n"(C::kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int total_tiles = p.tiles_w * p.tiles_h * p.tiles_n * p.tiles_c;
  const int ksteps = p.ntaps * p.kblocks;

  if (warp == 0) {
    // =========================== TMA producer ===========================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        int t = tile;
        const int tc = t % p.tiles_c; t /= p.tiles_c;
        const int tw = t % p.tiles_w; t /= p.tiles_w;
        const int th = t % p.tiles_h; t /= p.tiles_h;
        const int tn = t;
        const int w0 = tw * p.bw, h0 = th * p.bh, n0 = tn * p.bn, c0 = tc * BN;
        for (int tap = 0; tap < p.ntaps; ++tap) {
          const CUtensorMap* ma = &maps.a[p.tap_map[tap]];
          const int cw = w0 + p.tap_dw[tap], ch = h0 + p.tap_dh[tap];
          const int brow = p.tap_brow[tap] + c0;
          for (int kb = 0; kb < p.kblocks; ++kb) {
            mbar_wait(&empty_bar[stage], phase ^ 1);
            uint8_t* sa = smem + stage * C::kStageBytes;
            uint8_t* sb = sa + kATileBytes;
            mbar_expect_tx(&full_bar[stage], C::kStageBytes);
            tma_load_4d(ma, &full_bar[stage], sa, kb * kBlockK, cw, ch, n0);
            tma_load_2d(&maps.b, &full_bar[stage], sb, kb * kBlockK, brow);
            if (++stage == C::kStages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    if (lane == 0) {
      // instruction descriptor: D=f32, A=B=bf16, both K-major, N=BN, M=128
      constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) |
                                 ((uint32_t)(kBlockM >> 4) << 24);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
        for (int ks = 0; ks < ksteps; ++ks) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * C::kStageBytes);
          const uint64_t adesc = make_smem_desc(sa);
          const uint64_t bdesc = make_smem_desc(sa + kATileBytes);
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k) {
            // advance 32 bytes (16 bf16) inside the 128B swizzle row: +2 in 16-byte units
            umma_bf16(d_tmem, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc,
                      (ks | k) != 0 ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tfull_bar[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // =========================== epilogue ===========================
    const int quad = warp & 3;                   // TMEM lane group this warp may access
    const int row = quad * 32 + lane;            // accumulator row == pixel within the tile
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      int t = tile;
      const int tc = t % p.tiles_c; t /= p.tiles_c;
      const int tw = t % p.tiles_w; t /= p.tiles_w;
      const int th = t % p.tiles_h; t /= p.tiles_h;
      const int tn = t;
      const int wi = row % p.bw, hi = (row / p.bw) % p.bh, ni = row / (p.bw * p.bh);
      const int w = tw * p.bw + wi, h = th * p.bh + hi, n = tn * p.bn + ni;
      const bool valid = (w < p.out_w) && (h < p.out_h) && (n < p.out_n);
      const long long off = (long long)n * p.o_sn + (long long)h * p.o_sh + (long long)w * p.o_sw;

      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr0 = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * BN);

      if constexpr (BN == 16) {
        // conv1 input gradient: columns q = (ph*2+pw)*3 + c of the space-to-depth pixel (h, w)
        uint32_t v[32];
        tmem_ld16(taddr0, v);
        tmem_ld_wait();
        if (valid && p.epi == 1) {
          float* g = reinterpret_cast<float*>(p.out);
#pragma unroll
          for (int q = 0; q < 12; ++q) {
            const int c = q % 3, pw = (q / 3) & 1, ph = q / 6;
            const long long o = (((long long)n * 3 + c) * p.img_h + (2 * h + ph)) * p.img_w + (2 * w + pw);
            g[o] = __fdiv_rn(__uint_as_float(v[q]), __ldg(p.std_dev + c));
          }
        }
      } else {
#pragma unroll 1
        for (int j = 0; j < BN / 32; ++j) {
          uint32_t v[32];
          tmem_ld32(taddr0 + (uint32_t)(j * 32), v);
          tmem_ld_wait();
          const int col0 = tc * BN + j * 32;
          if (valid && col0 < p.cout) {
            float f[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]);
            if (p.bias) {
              const float4* b4 = reinterpret_cast<const float4*>(p.bias + col0);
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                const float4 b = __ldg(b4 + i);
                f[4 * i] += b.x; f[4 * i + 1] += b.y; f[4 * i + 2] += b.z; f[4 * i + 3] += b.w;
              }
            }
            if (p.add) {
              const uint4* a4 = reinterpret_cast<const uint4*>(p.add + off + col0);
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const uint4 a = __ldg(a4 + i);
                f[8 * i] += bf16_lo(a.x); f[8 * i + 1] += bf16_hi(a.x);
                f[8 * i + 2] += bf16_lo(a.y); f[8 * i + 3] += bf16_hi(a.y);
                f[8 * i + 4] += bf16_lo(a.z); f[8 * i + 5] += bf16_hi(a.z);
                f[8 * i + 6] += bf16_lo(a.w); f[8 * i + 7] += bf16_hi(a.w);
              }
            }
            if (p.relu) {
#pragma unroll
              for (int i = 0; i < 32; ++i) f[i] = fmaxf(f[i], 0.f);
            }
            if (p.mask) {
              const uint4* m4 = reinterpret_cast<const uint4*>(p.mask + off + col0);
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const uint4 m = __ldg(m4 + i);
                const uint32_t mm[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  if (!(bf16_lo(mm[k]) > 0.f)) f[8 * i + 2 * k] = 0.f;
                  if (!(bf16_hi(mm[k]) > 0.f)) f[8 * i + 2 * k + 1] = 0.f;
                }
              }
            }
            uint4* o4 = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.out) + off + col0);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              uint4 o;
              o.x = pack_bf16x2(f[8 * i], f[8 * i + 1]);
              o.y = pack_bf16x2(f[8 * i + 2], f[8 * i + 3]);
              o.z = pack_bf16x2(f[8 * i + 4], f[8 * i + 5]);
              o.w = pack_bf16x2(f[8 * i + 6], f[8 * i + 7]);
              o4[i] = o;
            }
          }
        }
      }
      // accumulator drained: hand the TMEM buffer back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                 "n"(C::kTmemCols)
                 : "memory");
  }
}

template <int BN>
int launch_impl(const TmapPack& maps, const ConvGemmParams& p, cudaStream_t stream) {
  using C = Cfg<BN>;
  static bool configured = false;
  if (!configured) {
    ADV_CHECK_CUDA(cudaFuncSetAttribute(conv_gemm_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        C::kSmemBytes));
    configured = true;
  }
  const long long total = (long long)p.tiles_w * p.tiles_h * p.tiles_n * p.tiles_c;
  if (total <= 0) return ADVB200_OK;
  const int grid = (int)(total < kNumSMs ? total : kNumSMs);
  conv_gemm_kernel<BN><<<grid, kThreads, C::kSmemBytes, stream>>>(maps, p);
  ADV_LAUNCHED();
  return ADVB200_OK;
}

}  // namespace

int launch_conv_gemm(int bn_cols, const CUtensorMap* tmA, int n_maps, const CUtensorMap& tmB,
                     const ConvGemmParams& p, cudaStream_t stream) {
  ADV_CHECK_ARG(n_maps >= 1 && n_maps <= 4, "conv_gemm: 1..4 activation maps");
  ADV_CHECK_ARG(p.bw * p.bh * p.bn == kBlockM, "conv_gemm: tile must hold 128 pixels");
  ADV_CHECK_ARG(p.ntaps >= 1 && p.ntaps <= kMaxTaps && p.kblocks >= 1, "conv_gemm: bad K loop");
  TmapPack maps;
  for (int i = 0; i < 4; ++i) maps.a[i] = tmA[i < n_maps ? i : 0];
  maps.b = tmB;
  switch (bn_cols) {
    case 16: return launch_impl<16>(maps, p, stream);
    case 64: return launch_impl<64>(maps, p, stream);
    case 128: return launch_impl<128>(maps, p, stream);
    case 256: return launch_impl<256>(maps, p, stream);
    default: return fail(ADVB200_ERR_ARG, "conv_gemm: BN must be 16, 64, 128 or 256");
  }
}

int make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                   const uint64_t* strides_bytes, const uint32_t* box) {
  static PFN_cuTensorMapEncodeTiled_v12000 encode = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
  });
  if (!encode) return fail(ADVB200_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
  cuuint64_t gdim[5];
  cuuint64_t gstr[5];
  cuuint32_t bx[5];
  cuuint32_t es[5];
  for (int i = 0; i < rank; ++i) { gdim[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides_bytes[i];
  CUresult r = encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gdim,
                      gstr, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                      CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char buf[256];
    snprintf(buf, sizeof buf,
             "cuTensorMapEncodeTiled failed (%d): rank %d dims [%llu,%llu,%llu,%llu] box [%u,%u,%u,%u]", (int)r,
             rank, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
             (unsigned long long)(rank > 2 ? dims[2] : 0), (unsigned long long)(rank > 3 ? dims[3] : 0), box[0],
             rank > 1 ? box[1] : 0, rank > 2 ? box[2] : 0, rank > 3 ? box[3] : 0);
    return fail(ADVB200_ERR_CUDA, buf);
  }
  return ADVB200_OK;
}

}  // namespace advb200

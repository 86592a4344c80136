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
constexpr int kThreads = 320;                // warp 0 TMA, warp 1 MMA, warps 2..9 epilogue
constexpr int kEpiWarp0 = 2;
constexpr int kEpiThreads = 256;

constexpr int kMaxStages = 8;
constexpr int kMaxAuxDepth = 8;
constexpr int kSmemBudget = 232448;            // 227 KB opt-in limit per CTA
constexpr int kSubTileBytes = kBlockM * 128;   // 128 rows x 64 bf16: one epilogue sub-tile
constexpr int kBarrierBytes = 512;
constexpr int kIdentBytes = 64 * 128;          // 64x64 bf16 identity operand of the residual k-steps
constexpr int kBiasBytes = 2 * 256 * 4;        // per-accumulator copy of the tile's folded-BN shift

// kPair: two CTAs of a cluster (one SM pair) run tcgen05.mma.cta_group::2 on a 256-pixel x BN tile; each CTA
// stages its own 128 pixels of A and HALF of the BN weight rows, halving the weight traffic per CTA.
template <int BN, bool kPair> struct Cfg {
  static constexpr int kBRows = kPair ? BN / 2 : BN;
  static constexpr int kBTileBytes = kBRows * kBlockK * 2;
  static constexpr int kStageBytes = kATileBytes + kBTileBytes;
  // two accumulators; tcgen05.alloc takes a power of two >= 32 (BN = 192: 384 -> 512)
  static constexpr int kTmemCols = 2 * BN <= 32 ? 32 : 2 * BN <= 64 ? 64 : 2 * BN <= 128 ? 128 : 2 * BN <= 256 ? 256 : 512;
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
// one lane of a converged warp (the rest of the warp keeps running the same uniform control flow)
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n.reg .pred p;\nelect.sync _|p, 0xffffffff;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(pred));
  return pred != 0;
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
// ---- CTA-pair (cta_group::2) variants -----------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `p` (a shared::cta pointer) in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_u32(const void* p, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_u32(p)), "r"(rank));
  return r;
}
// Remote arrive on the leader CTA's barrier (the follower's epilogue handing a TMEM accumulator back).  Default
// semantics (release at CTA scope), as for the local arrive: what the leader's MMA warp must not overtake are this
// warp's tcgen05.ld reads, and those are ordered by tcgen05.wait::ld + tcgen05.fence::before_thread_sync.  The
// earlier `.release.cluster` form made ptxas emit MEMBAR.ALL.CTA + ERRBAR, which also waited for the warp's global
// mask / output stores: 13-22 % of the samples of the pair-mode launches (ncu source view) sat on that fence.
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA loads whose completion is signalled on the LEADER CTA's mbarrier (bar_cluster_addr)
__device__ __forceinline__ void tma_load_4d_pair(const CUtensorMap* map, uint32_t bar_cluster_addr, void* dst, int c0,
                                                 int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar_cluster_addr), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d_pair(const CUtensorMap* map, uint32_t bar_cluster_addr, void* dst, int c0,
                                                 int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                               uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the same-offset mbarrier of BOTH CTAs of the pair once all prior MMAs have completed
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
          smem_u32(bar)),
      "h"((uint16_t)3)
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

// programmatic dependent launch: the next conv launch may start its prologue (and constant weight loads) while
// this grid drains; everything that touches activations waits for the previous grid to complete
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// K-major, 128B-swizzled operand tile: rows of 128 bytes, 8-row atoms of 1024 bytes.
// sbo_bytes: distance between consecutive 8-row groups (1024 for a dense tile; the halo-mode A
// operand uses the halo box width * 128, and start addresses shifted by whole rows).
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr, uint32_t sbo_bytes = 1024, uint32_t swizzle_code = 2) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);   // start address, 16-byte units
  d |= (uint64_t)1 << 16;                        // leading byte offset (unused for swizzled K-major)
  d |= (uint64_t)(sbo_bytes >> 4) << 32;         // stride byte offset between 8-row groups
  d |= (uint64_t)1 << 46;                        // descriptor version (Blackwell)
  d |= (uint64_t)swizzle_code << 61;             // 2 = SWIZZLE_128B, 6 = SWIZZLE_32B
  return d;
}

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
// relu fused into the conversion: max(x, 0) then round-to-nearest-even to bf16 (NaN -> canonical NaN)
__device__ __forceinline__ uint32_t pack_bf16x2_relu(float lo, float hi) {
  uint32_t d;
  asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}
// packed fp32x2 add (sm_100): {a0, a1} += {b0, b1}
__device__ __forceinline__ void add2(float& a0, float& a1, float b0, float b1) {
  asm("{\n.reg .b64 ra, rb;\nmov.b64 ra, {%0, %1};\nmov.b64 rb, {%2, %3};\nadd.rn.f32x2 ra, ra, rb;\n"
      "mov.b64 {%0, %1}, ra;\n}"
      : "+f"(a0), "+f"(a1)
      : "f"(b0), "f"(b1));
}
__device__ __forceinline__ void tmem_ld32f(uint32_t taddr, float (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]),
        "=f"(v[8]), "=f"(v[9]), "=f"(v[10]), "=f"(v[11]), "=f"(v[12]), "=f"(v[13]), "=f"(v[14]),
        "=f"(v[15]), "=f"(v[16]), "=f"(v[17]), "=f"(v[18]), "=f"(v[19]), "=f"(v[20]), "=f"(v[21]),
        "=f"(v[22]), "=f"(v[23]), "=f"(v[24]), "=f"(v[25]), "=f"(v[26]), "=f"(v[27]), "=f"(v[28]),
        "=f"(v[29]), "=f"(v[30]), "=f"(v[31])
      : "r"(taddr));
}
__device__ __forceinline__ float bf16_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }

__device__ __forceinline__ void tma_store_4d(const CUtensorMap* map, const void* src, int c0, int c1, int c2,
                                             int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(map)),
               "r"(smem_u32(src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, 256;" ::: "memory"); }
__device__ __forceinline__ float4 lds_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, uint4 v) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// attack.py:61-63 (project) / 33-34 (no projection); same arithmetic as elementwise.cu::linf_one
__device__ __forceinline__ float linf_update_one(float x, float x0, float g, float a, float e, bool project) {
  const float sgn = (float)((g > 0.f) - (g < 0.f));          // torch.sign: NaN and +-0 map to 0
  float x1 = __fadd_rn(x, __fmul_rn(a, sgn));
  if (project) {
    const float hi = __fadd_rn(x0, e), lo = __fsub_rn(x0, e);
    x1 = fmaxf(fminf(x1, hi), lo);
  }
  return fminf(fmaxf(x1, 0.f), 1.f);
}

// ---------------------------------------------------------------- the kernel
struct TileCoord { int w0, h0, n0, c0, cls; };
// pair mode: `tile` indexes (pair of consecutive pixel tiles, column block); CTA `rank` takes pixel tile
// 2*pair + rank (an odd tile count leaves the last rank-1 tile out of bounds in n: loads zero-fill,
// stores clip)
template <bool kPair>
__device__ __forceinline__ TileCoord tile_coord(const ConvGemmParams& p, int tile, int bn_cols, int rank) {
  TileCoord t;
  if (p.reverse) tile = p.total_tiles - 1 - tile;   // snake order: start where the previous launch ended (L2 reuse)
  // divisions by launch constants through host-made reciprocals: q = (n * ceil(2^40 / d)) >> 40 (n < 2^20)
  int q = (int)(((unsigned long long)(unsigned)tile * p.magic_c) >> 40);
  const int tc = tile - q * p.tiles_c; tile = q;
  t.cls = 0;
  if (p.nclass > 1) {
    // The four parity classes of one pixel tile are CONSECUTIVE units (class-minor order), so they run at the same time
    // on neighbouring SMs and share the gradient tile they all read through the L2.  In class-major order every class
    // swept the whole tensor on its own and it came from DRAM four times (ncu, bwd l2.0.c2: 212 MB read against 64 MB
    // algorithmic).  The classes cost 1 / 2 / 2 / 4 taps; with a grid that is a multiple of four a CTA would always draw
    // the same class, hence the rotation by the pixel-tile index.
    // Launches of only a few waves keep the class-major order: there the unequal classes inside ONE wave cost more than
    // the re-reads (layer 4: +4 / +6 us when interleaved), and their gradient tensor stays in the L2 anyway.
    if (p.cls_minor) {
      const int pix = tile >> 2;
      t.cls = (tile + pix) & 3;
      tile = pix;
    } else {
      t.cls = (int)(((unsigned long long)(unsigned)tile * p.magic_cls) >> 40);
      tile -= t.cls * p.units_per_class;
    }
  }
  if (kPair) tile = 2 * tile + rank;
  q = (int)(((unsigned long long)(unsigned)tile * p.magic_w) >> 40);
  const int tw = tile - q * p.tiles_w; tile = q;
  q = (int)(((unsigned long long)(unsigned)tile * p.magic_h) >> 40);
  const int th = tile - q * p.tiles_h; tile = q;
  t.w0 = tw * p.step_w + p.org_w; t.h0 = th * p.bh; t.n0 = tile * p.bn; t.c0 = tc * bn_cols;
  return t;
}

// smem carve-up (host and device agree through these helpers)
template <int BN> struct Smem {
  // output staging slots: 4 when the epilogue is the busy part (few k-steps), 2 when the operand ring
  // needs the room (many k-steps)
  __host__ __device__ static constexpr int fixed_bytes(int out_slots) {
    return (BN == 16 ? 0 : out_slots) * kSubTileBytes + kIdentBytes + kBiasBytes + kBarrierBytes;
  }
};

template <int BN, bool kPair>
__global__ void __launch_bounds__(kThreads, 1)
conv_gemm_kernel(const __grid_constant__ ConvGemmMaps maps, const __grid_constant__ ConvGemmParams p) {
  using C = Cfg<BN, kPair>;
  const int rank = kPair ? (int)cluster_ctarank() : 0;       // CTA within the pair; rank 0 issues the MMAs
  const int tile0 = kPair ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int tile_step = kPair ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  extern __shared__ __align__(1024) uint8_t smem[];
  const int n_out_slots = BN == 16 ? 0 : p.out_slots;
  // [resident weights (halo mode)] [operand ring] [output staging] [identity] [shift] [barriers]
  uint8_t* ring = smem + p.ring_off;
  const int stage_bytes = p.stage_bytes;
  uint8_t* out_smem = ring + p.stages * stage_bytes;        // n_out_slots x 16 KB output staging
  uint8_t* ident_smem = out_smem + n_out_slots * kSubTileBytes;   // 64x64 bf16 identity (K-major, swizzled)
  float* bias_smem = reinterpret_cast<float*>(ident_smem + kIdentBytes);   // [2][256]
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(bias_smem) + kBiasBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kMaxStages;
  uint64_t* tfull_bar = bars + 2 * kMaxStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint64_t* bres_bar = tempty_bar + 2;                      // resident weights landed (halo mode)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bres_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    if (smem_u32(smem) & 1023u) __trap();                   // swizzled tiles need 1024-byte alignment
    for (int i = 0; i < 5; ++i) prefetch_tmap(&maps.a[i]);
    prefetch_tmap(&maps.b);
    if (p.has_b2) prefetch_tmap(&maps.b2);
    for (int i = 0; i < 4; ++i) prefetch_tmap(&maps.out[i]);
    if (p.has_add) prefetch_tmap(&maps.add);
    for (int s = 0; s < kMaxStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tfull_bar[a], 1); mbar_init(&tempty_bar[a], kPair ? 16 : 8); }
    mbar_init(bres_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    if constexpr (kPair) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                       smem_u32(tmem_slot)),
                   "n"(C::kTmemCols)
                   : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                       smem_u32(tmem_slot)),
                   "n"(C::kTmemCols)
                   : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
  }
  const int total_tiles = p.total_tiles;
  if (warp >= kEpiWarp0 && p.bias && BN != 16) {
    // folded-BN shift of the first tile's column block (accumulator 0's copy)
    const int i = threadIdx.x - kEpiWarp0 * 32;
    // (BN = 192 leaves the last column block partial: channels >= cout have no shift, no masks and are never stored)
    if (i < BN && tile0 < total_tiles) {
      const int ch = tile_coord<kPair>(p, tile0, BN, rank).c0 + i;
      bias_smem[i] = ch < p.cout ? __ldg(p.bias + ch) : 0.f;
    }
  }
  if (warp >= kEpiWarp0 && p.has_add) {
    // identity operand for the residual k-steps: I[n][k] = (n == k), rows of 128 B, 16-byte chunks
    // XOR-swizzled by (row & 7) exactly as TMA would have written them
    const int i = threadIdx.x - kEpiWarp0 * 32;          // 256 threads x 32 B = 8 KB
    // pair mode: this CTA holds rows 32*rank .. 32*rank+31 of I (its half of the N = 64 operand)
    const int r = i >> 2, cpair = (i & 3) * 2;           // local row, first of two logical chunks
    const int nrow = r + (kPair ? 32 * rank : 0);        // row of I
#pragma unroll
    for (int cc = 0; cc < 2; ++cc) {
      const int c = cpair + cc;                          // logical chunk: columns 8c..8c+7
      uint4 v = make_uint4(0, 0, 0, 0);
      if ((nrow >> 3) == c) {
        const uint32_t one = 0x3F80u << (16 * (nrow & 1));  // bf16 1.0 in the right half-word
        const int wsel = (nrow & 7) >> 1;
        v.x = wsel == 0 ? one : 0; v.y = wsel == 1 ? one : 0; v.z = wsel == 2 ? one : 0; v.w = wsel == 3 ? one : 0;
      }
      *reinterpret_cast<uint4*>(ident_smem + r * 128 + ((c ^ (r & 7)) << 4)) = v;
    }
    fence_async_smem();
  }
  tc_fence_before();
  if constexpr (kPair) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int ksteps = p.ksteps;                              // sum of tap_kb[] (generic mode)
  const int nstages = p.stages;

  if (warp == 0) {
    // =========================== TMA producer ===========================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      pdl_launch_dependents();
      if (p.halo) {
        // all taps' weights once, then one halo box of input pixels per tile
        const int rb = p.row_bytes ? p.row_bytes : 128;
        if constexpr (kPair) {
          // CTA pair: each CTA keeps HALF of every tap's weight rows and its own halo box; everything completes
          // on the leader's barriers
          const uint32_t rbar = mapa_u32(bres_bar, 0);
          if (rank == 0) mbar_expect_tx(bres_bar, (uint32_t)(2 * p.ntaps * C::kBRows * rb));
          for (int tap = 0; tap < p.ntaps; ++tap)
            tma_load_2d_pair(&maps.b, rbar, smem + tap * C::kBRows * rb, 0, p.tap_brow[tap] + rank * C::kBRows);
        } else {
          mbar_expect_tx(bres_bar, (uint32_t)(p.ntaps * BN * rb));
          for (int tap = 0; tap < p.ntaps; ++tap) tma_load_2d(&maps.b, bres_bar, smem + tap * BN * rb, 0, p.tap_brow[tap]);
        }
        pdl_wait();
        for (int tile = tile0; tile < total_tiles; tile += tile_step) {
          const TileCoord tc = tile_coord<kPair>(p, tile, BN, rank);
          mbar_wait(&empty_bar[stage], phase ^ 1);
          if constexpr (kPair) {
            const uint32_t fb = mapa_u32(&full_bar[stage], 0);
            if (rank == 0) mbar_expect_tx(&full_bar[stage], (uint32_t)(2 * p.halo_rows * rb));
            tma_load_4d_pair(&maps.a[0], fb, ring + stage * stage_bytes, 0, tc.w0 + p.halo_w0, tc.h0 + p.halo_h0, tc.n0);
          } else {
            mbar_expect_tx(&full_bar[stage], (uint32_t)(p.halo_rows * rb));
            tma_load_4d(&maps.a[0], &full_bar[stage], ring + stage * stage_bytes, 0, tc.w0 + p.halo_w0,
                        tc.h0 + p.halo_h0, tc.n0);
          }
          if (++stage == nstages) { stage = 0; phase ^= 1; }
        }
      } else {
      const bool bres = p.bres != 0;
      if (bres && tile0 < total_tiles) {
        // resident weights: every tile of this CTA has the same column block (the host makes the tile
        // stride a multiple of tiles_c), so all of its k-blocks are staged once, ahead of the ring
        const int c0 = tile_coord<kPair>(p, tile0, BN, rank).c0;
        uint8_t* sb = smem;
        if constexpr (kPair) { if (rank == 0) mbar_expect_tx(bres_bar, (uint32_t)(2 * ksteps * C::kBTileBytes)); }
        else mbar_expect_tx(bres_bar, (uint32_t)(ksteps * C::kBTileBytes));
        for (int tap = 0; tap < p.ntaps; ++tap) {
          const CUtensorMap* mb = p.tap_bmap[tap] ? &maps.b2 : &maps.b;
          for (int kb = 0; kb < p.tap_kb[tap]; ++kb, sb += C::kBTileBytes) {
            if constexpr (kPair) tma_load_2d_pair(mb, mapa_u32(bres_bar, 0), sb, kb * kBlockK, p.tap_brow[tap] + c0 + rank * C::kBRows);
            else tma_load_2d(mb, bres_bar, sb, kb * kBlockK, p.tap_brow[tap] + c0);
          }
        }
      }
      pdl_wait();
      for (int tile = tile0; tile < total_tiles; tile += tile_step) {
        const TileCoord tc = tile_coord<kPair>(p, tile, BN, rank);
        const int tap_lo = p.nclass > 1 ? p.cls_tap0[tc.cls] : 0;
        const int tap_hi = p.nclass > 1 ? tap_lo + p.cls_ntaps[tc.cls] : p.ntaps;
        for (int tap = tap_lo; tap < tap_hi; ++tap) {
          const CUtensorMap* ma = &maps.a[p.tap_map[tap]];
          const CUtensorMap* mb = p.tap_bmap[tap] ? &maps.b2 : &maps.b;
          const int cw = tc.w0 + p.tap_dw[tap], ch = tc.h0 + p.tap_dh[tap];
          const int brow = p.tap_brow[tap] + tc.c0;
          for (int kb = 0; kb < p.tap_kb[tap]; ++kb) {
            mbar_wait(&empty_bar[stage], phase ^ 1);
            uint8_t* sa = ring + stage * stage_bytes;
            uint8_t* sb = sa + kATileBytes;
            if (bres) {
              if constexpr (kPair) {
                const uint32_t fb = mapa_u32(&full_bar[stage], 0);
                if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * kATileBytes);
                tma_load_4d_pair(ma, fb, sa, kb * kBlockK, cw, ch, tc.n0);
              } else {
                mbar_expect_tx(&full_bar[stage], kATileBytes);
                tma_load_4d(ma, &full_bar[stage], sa, kb * kBlockK, cw, ch, tc.n0);
              }
            } else if constexpr (kPair) {
              // both CTAs' boxes complete on the leader's barrier; this CTA stages its half of the weight rows
              const uint32_t fb = mapa_u32(&full_bar[stage], 0);
              if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * C::kStageBytes);
              tma_load_4d_pair(ma, fb, sa, kb * kBlockK, cw, ch, tc.n0);
              tma_load_2d_pair(mb, fb, sb, kb * kBlockK, brow + rank * C::kBRows);
            } else {
              mbar_expect_tx(&full_bar[stage], C::kStageBytes);
              tma_load_4d(ma, &full_bar[stage], sa, kb * kBlockK, cw, ch, tc.n0);
              tma_load_2d(mb, &full_bar[stage], sb, kb * kBlockK, brow);
            }
            if (++stage == nstages) { stage = 0; phase ^= 1; }
          }
        }
        if (p.has_add) {
          // residual / skip-gradient tile, one 64-channel block per stage (added by an identity MMA)
          for (int j = 0; j < BN / 64; ++j) {
            mbar_wait(&empty_bar[stage], phase ^ 1);
            if constexpr (kPair) {
              const uint32_t fb = mapa_u32(&full_bar[stage], 0);
              if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * kATileBytes);
              tma_load_4d_pair(&maps.add, fb, ring + stage * stage_bytes, tc.c0 + 64 * j, tc.w0, tc.h0, tc.n0);
            } else {
              mbar_expect_tx(&full_bar[stage], kATileBytes);
              tma_load_4d(&maps.add, &full_bar[stage], ring + stage * stage_bytes, tc.c0 + 64 * j, tc.w0, tc.h0,
                          tc.n0);
            }
            if (++stage == nstages) { stage = 0; phase ^= 1; }
          }
        }
      }
      }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    // The whole warp runs the (warp-uniform) control flow, one elected lane issues the MMAs and commits:
    // loop state then lives in uniform registers and the single-thread issue path — which paces every
    // launch with N <= 128 — stays short (measured: ~110 clk per MMA when all of this ran divergently
    // under `if (lane == 0)`).
    if (rank == 0) {
      // instruction descriptor: D=f32, A=B=bf16, both K-major, N=BN, M=128 (256 across the CTA pair)
      constexpr uint32_t kMmaM = kPair ? 2 * kBlockM : kBlockM;
      constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) |
                                 ((uint32_t)(kMmaM >> 4) << 24);
      constexpr uint32_t idesc64 = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(64 >> 3) << 17) |
                                   ((uint32_t)(kMmaM >> 4) << 24);
      const uint32_t desc_hi = (uint32_t)(make_smem_desc(0) >> 32);
      const uint32_t rb = p.row_bytes ? (uint32_t)p.row_bytes : 128u;      // halo-mode operand row pitch
      const uint32_t sw_code = rb == 32 ? 6u : 2u;
      const uint32_t a_hi_halo = (uint32_t)(make_smem_desc(0, (uint32_t)p.sbo_rows * rb, sw_code) >> 32);
      const uint32_t b_hi_halo = (uint32_t)(make_smem_desc(0, 8u * rb, sw_code) >> 32);
      const int kslices = (int)(rb >> 5);                                  // K = 16 slices per row
      const uint32_t ring_lo = ((smem_u32(ring) & 0x3FFFFu) >> 4) | (1u << 16);
      const uint32_t bres_lo = ((smem_u32(smem) & 0x3FFFFu) >> 4) | (1u << 16);
      const uint32_t ident_lo = ((smem_u32(ident_smem) & 0x3FFFFu) >> 4) | (1u << 16);
      const uint32_t stage_units = (uint32_t)(stage_bytes >> 4);
      const bool halo = p.halo != 0, has_add = p.has_add != 0, bres = p.bres != 0;
      const int ntaps = p.ntaps;
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      if ((halo || bres) && tile0 < total_tiles) mbar_wait(bres_bar, 0);
      for (int tile = tile0; tile < total_tiles; tile += tile_step) {
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
        if (halo) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t a_lo0 = ring_lo + (uint32_t)stage * stage_units;
            const uint32_t b_tap = (uint32_t)C::kBRows * (rb >> 4);       // 16-byte units between taps' weights
            if (kslices == 1) {
              // 32-byte rows: one MMA per tap; unrolled so the single issuing thread batches the tap-offset loads
#pragma unroll 8
              for (int tap = 0; tap < ntaps; ++tap) {
                const uint64_t ad = ((uint64_t)a_hi_halo << 32) | (a_lo0 + (uint32_t)p.tap_row[tap] * 2u);
                const uint64_t bd = ((uint64_t)b_hi_halo << 32) | (bres_lo + (uint32_t)tap * b_tap);
                if constexpr (kPair) umma_bf16_pair(d_tmem, ad, bd, idesc, tap != 0 ? 1u : 0u);
                else umma_bf16(d_tmem, ad, bd, idesc, tap != 0 ? 1u : 0u);
              }
            } else {
              for (int tap = 0; tap < ntaps; ++tap) {
                const uint32_t a_lo = a_lo0 + (uint32_t)(p.tap_row[tap] * 8);
                const uint32_t b_lo = bres_lo + (uint32_t)tap * b_tap;
#pragma unroll
                for (int k = 0; k < kBlockK / 16; ++k) {
                  const uint64_t ad = ((uint64_t)a_hi_halo << 32) | (a_lo + 2u * k);
                  const uint64_t bd = ((uint64_t)b_hi_halo << 32) | (b_lo + 2u * k);
                  if constexpr (kPair) umma_bf16_pair(d_tmem, ad, bd, idesc, (tap | k) != 0 ? 1u : 0u);
                  else umma_bf16(d_tmem, ad, bd, idesc, (tap | k) != 0 ? 1u : 0u);
                }
              }
            }
            if constexpr (kPair) umma_commit_pair(&empty_bar[stage]); else umma_commit(&empty_bar[stage]);
          }
          __syncwarp();
          if (++stage == nstages) { stage = 0; phase ^= 1; }
        } else {
          const int ksteps_t = p.nclass > 1 ? p.cls_ksteps[tile_coord<kPair>(p, tile, BN, 0).cls] : ksteps;
          for (int ks = 0; ks < ksteps_t; ++ks) {
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            if (elect_one()) {
              const uint32_t a_lo = ring_lo + (uint32_t)stage * stage_units;
              const uint32_t b_lo = bres ? bres_lo + (uint32_t)ks * (uint32_t)(C::kBTileBytes >> 4)
                                         : a_lo + (uint32_t)(kATileBytes >> 4);
#pragma unroll
              for (int k = 0; k < kBlockK / 16; ++k) {
                // advance 32 bytes (16 bf16) inside the 128B swizzle row: +2 in 16-byte units
                const uint64_t adesc = ((uint64_t)desc_hi << 32) | (a_lo + 2u * k);
                const uint64_t bdesc = ((uint64_t)desc_hi << 32) | (b_lo + 2u * k);
                if constexpr (kPair) umma_bf16_pair(d_tmem, adesc, bdesc, idesc, (ks | k) != 0 ? 1u : 0u);
                else umma_bf16(d_tmem, adesc, bdesc, idesc, (ks | k) != 0 ? 1u : 0u);
              }
              if constexpr (kPair) umma_commit_pair(&empty_bar[stage]); else umma_commit(&empty_bar[stage]);
            }
            __syncwarp();
            if (++stage == nstages) { stage = 0; phase ^= 1; }
          }
          if (has_add) {
            // D[:, 64j..64j+63] += R_j * I  (bf16 x 1.0 accumulated in fp32: exact)
            for (int j = 0; j < BN / 64; ++j) {
              mbar_wait(&full_bar[stage], phase);
              tc_fence_after();
              if (elect_one()) {
                const uint32_t a_lo = ring_lo + (uint32_t)stage * stage_units;
#pragma unroll
                for (int k = 0; k < kBlockK / 16; ++k) {
                  const uint64_t adesc = ((uint64_t)desc_hi << 32) | (a_lo + 2u * k);
                  const uint64_t bdesc = ((uint64_t)desc_hi << 32) | (ident_lo + 2u * k);
                  if constexpr (kPair) umma_bf16_pair(d_tmem + (uint32_t)(64 * j), adesc, bdesc, idesc64, 1u);
                  else umma_bf16(d_tmem + (uint32_t)(64 * j), adesc, bdesc, idesc64, 1u);
                }
                if constexpr (kPair) umma_commit_pair(&empty_bar[stage]); else umma_commit(&empty_bar[stage]);
              }
              __syncwarp();
              if (++stage == nstages) { stage = 0; phase ^= 1; }
            }
          }
        }
        if (elect_one()) {
          if constexpr (kPair) umma_commit_pair(&tfull_bar[acc]); else umma_commit(&tfull_bar[acc]);
        }
        __syncwarp();
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // =========================== epilogue (8 warps) ===========================
    // warp w may only touch TMEM lanes 32*(w%4)..+31; the two warps of a lane group split every
    // 64-column sub-tile into its two 32-column halves.
    const int ew = warp - kEpiWarp0;
    const int quad = warp & 3;
    const int half = ew >> 2;
    const int row = quad * 32 + lane;            // accumulator row == pixel within the tile
    const int etid = threadIdx.x - kEpiWarp0 * 32;
    int acc = 0;
    uint32_t acc_phase = 0;
    pdl_wait();          // masks, residuals and the output buffer belong to earlier launches

    if (BN == 64 && p.epi == 2) {
      // Stem input gradient.  Accumulator row = pixel (y, x) of a 16-wide box of the 112 x 112 gradient grid
      // (x = 0 and x = 14, 15 are halo columns); column = slot(q)*4 + u holds the partial sum over the four
      // vertical taps for horizontal tap u.  dX[y, x, q] = sum_u Z[(y, x+2-u), u, q]: the four addends sit
      // in neighbouring lanes of the same 16-lane row group -> warp shuffles.  Each lane then owns output
      // pixel (y, x) for x in 1..13 and writes fp32 NCHW pixel pairs, divided by the normalisation std.
      float inv_std[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) inv_std[c] = __fdiv_rn(1.0f, __ldg(p.std_dev + c));
      for (int tile = tile0; tile < total_tiles; tile += tile_step) {
        const TileCoord tc = tile_coord<kPair>(p, tile, BN, rank);
        const int xi = row & 15, yi = row >> 4;
        const int w = tc.w0 + xi, h = tc.h0 + yi, n = tc.n0;
        const bool valid = xi >= 1 && xi <= 13 && w < p.out_w && h < p.out_h && n < p.out_n;
        mbar_wait(&tfull_bar[acc], acc_phase);
        tc_fence_after();
        float f[32];
        tmem_ld32f(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * BN + half * 32), f);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty_bar[acc]);
        float o[8];
#pragma unroll
        for (int s = 0; s < 8; ++s) {
          const float a0 = __shfl_down_sync(0xffffffffu, f[4 * s + 0], 2);
          const float a1 = __shfl_down_sync(0xffffffffu, f[4 * s + 1], 1);
          const float a3 = __shfl_up_sync(0xffffffffu, f[4 * s + 3], 1);
          o[s] = (a0 + a1) + (f[4 * s + 2] + a3);
        }
        if (valid) {
          float* g = reinterpret_cast<float*>(p.out);
          // slots held by this half: half 0 -> (c,ph) = (0,0),(1,0),(2,0),(0,1); half 1 -> (1,1),(2,1)
          const int npairs = half == 0 ? 4 : 2;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            if (k < npairs) {
              const int pi = half * 4 + k;                 // pair index 0..5 = ph*3 + c
              const int c = pi % 3, ph = pi / 3;
              const long long off = (((long long)n * 3 + c) * p.img_h + (2 * h + ph)) * p.img_w + 2 * w;
              float2 v2;               // bf16 path: x (1/std) instead of the true division (fp32 mode keeps it)
              const float is = c == 0 ? inv_std[0] : (c == 1 ? inv_std[1] : inv_std[2]);   // registers, not local memory
              v2.x = o[2 * k] * is;
              v2.y = o[2 * k + 1] * is;
              if (p.upd_x) {
                // fused sign step + eps-ball projection + [0,1] clamp on the two pixels this lane owns
                float2 xa = *reinterpret_cast<const float2*>(p.upd_x + off);
                float2 x0 = make_float2(0.f, 0.f);
                const bool project = p.upd_x0 != nullptr;
                if (project) x0 = __ldg(reinterpret_cast<const float2*>(p.upd_x0 + off));
                xa.x = linf_update_one(xa.x, x0.x, v2.x, p.upd_alpha, p.upd_eps, project);
                xa.y = linf_update_one(xa.y, x0.y, v2.y, p.upd_alpha, p.upd_eps, project);
                *reinterpret_cast<float2*>(p.upd_x + off) = xa;
              } else {
                *reinterpret_cast<float2*>(g + off) = v2;
              }
            }
          }
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    } else if constexpr (BN == 16) {
      // conv1 input gradient: columns q = (ph*2+pw)*3 + c of the space-to-depth pixel (h, w);
      // fp32 scatter into the NCHW image gradient, divided by the normalisation std
      for (int tile = tile0; tile < total_tiles; tile += tile_step) {
        const TileCoord tc = tile_coord<kPair>(p, tile, BN, rank);
        const int wi = row % p.bw, hi = (row / p.bw) % p.bh, ni = row / (p.bw * p.bh);
        const int w = tc.w0 + wi, h = tc.h0 + hi, n = tc.n0 + ni;
        const bool valid = (w < p.out_w) && (h < p.out_h) && (n < p.out_n);
        mbar_wait(&tfull_bar[acc], acc_phase);
        tc_fence_after();
        if (half == 0) {
          uint32_t v[32];
          tmem_ld16(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * BN), v);
          tmem_ld_wait();
          if (valid) {
            // two horizontally adjacent image pixels (pw = 0, 1) per 8-byte store
            float* g = reinterpret_cast<float*>(p.out);
            float inv[3];
#pragma unroll
            for (int c = 0; c < 3; ++c) inv[c] = __ldg(p.std_dev + c);
#pragma unroll
            for (int c = 0; c < 3; ++c)
#pragma unroll
              for (int ph = 0; ph < 2; ++ph) {
                const long long o = (((long long)n * 3 + c) * p.img_h + (2 * h + ph)) * p.img_w + 2 * w;
                float2 v2;
                v2.x = __fdiv_rn(__uint_as_float(v[(ph * 2 + 0) * 3 + c]), inv[c]);
                v2.y = __fdiv_rn(__uint_as_float(v[(ph * 2 + 1) * 3 + c]), inv[c]);
                *reinterpret_cast<float2*>(g + o) = v2;
              }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
              if constexpr (kPair) mbar_arrive_cluster(mapa_u32(&tempty_bar[acc], 0)); else mbar_arrive(&tempty_bar[acc]);
            }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    } else {
      // Sub-tiles of 64 channels: TMEM -> registers -> (+shift, ReLU, bit masks) -> bf16 -> swizzled smem
      // staging -> TMA store.  The residual / skip gradient was already accumulated by the MMA warp.
      // One named barrier per sub-tile; four staging slots.
      constexpr int kSub = BN / 64;
      const bool leader = etid == 0;
      const uint32_t out0 = smem_u32(out_smem);
      const uint32_t bias0 = smem_u32(bias_smem);
      const uint32_t row_off = (uint32_t)row * 128u;
      const uint32_t sw = (uint32_t)(row & 7);
      // bw, bh are powers of two: pixel (wi, hi, ni) of this thread's row inside the tile box
      const int wi = row & (p.bw - 1), hi = (row >> p.bw_shift) & (p.bh - 1), ni = row >> (p.bw_shift + p.bh_shift);
      const bool has_bias = p.bias != nullptr;
      const bool relu = p.relu != 0;
      const bool has_mask_in = p.mask_in != nullptr;
      const bool has_mask_out = relu && p.mask_out != nullptr;
      int oslot = 0;            // output staging slot of the sub-tile being produced
      // ReLU gate words (backward) are fetched ONE TILE AHEAD: with the accumulator usually ready when the epilogue
      // gets to it, a load issued in the same tile would expose its whole DRAM latency (ncu: 22 % of the samples).
      auto load_gate = [&](int tile_, uint32_t (&mw)[kSub]) {
        const TileCoord t2 = tile_coord<kPair>(p, tile_, BN, rank);
        const int w2 = t2.w0 + wi, h2 = t2.h0 + hi, n2 = t2.n0 + ni;
        const bool ok = (w2 < p.out_w) && (h2 < p.out_h) && (n2 < p.out_n);
        const int ph2 = p.nclass > 1 ? p.cls_ph[t2.cls] : p.pix_ph, pw2 = p.nclass > 1 ? p.cls_pw[t2.cls] : p.pix_pw;
        const long long px2 = ((long long)n2 * p.full_h + (h2 * p.pix_sub + ph2)) * p.full_w + (w2 * p.pix_sub + pw2);
        const long long m0 = ((long long)(t2.c0 >> 6) * p.mask_pixels + px2) * 2 + half;
#pragma unroll
        for (int j = 0; j < kSub; ++j)
          mw[j] = (ok && t2.c0 + 64 * j < p.cout) ? __ldg(p.mask_in + m0 + (long long)j * p.mask_pixels * 2) : 0u;
      };
      uint32_t mnext[kSub];
#pragma unroll
      for (int j = 0; j < kSub; ++j) mnext[j] = 0xffffffffu;
      if (has_mask_in && tile0 < total_tiles) load_gate(tile0, mnext);

      for (int tile = tile0; tile < total_tiles; tile += tile_step) {
        const TileCoord tc = tile_coord<kPair>(p, tile, BN, rank);
        // output pixel of this thread's row (bit-mask addressing) and whether it exists
        const int tw = tc.w0 + wi, th = tc.h0 + hi, tn = tc.n0 + ni;
        const bool valid = (tw < p.out_w) && (th < p.out_h) && (tn < p.out_n);
        const int pix_ph = p.nclass > 1 ? p.cls_ph[tc.cls] : p.pix_ph, pix_pw = p.nclass > 1 ? p.cls_pw[tc.cls] : p.pix_pw;
        const long long pixel = ((long long)tn * p.full_h + (th * p.pix_sub + pix_ph)) * p.full_w +
                                (tw * p.pix_sub + pix_pw);
        const long long mword0 = ((long long)(tc.c0 >> 6) * p.mask_pixels + pixel) * 2 + half;
        // direct-store mode has no per-sub-tile barrier: one barrier per tile keeps a fast warp from overwriting the
        // shift copy that slower warps still read for the previous tile (found by the candidate-equivalence test)
        if (n_out_slots == 0 && has_bias) epi_bar();
        // folded-BN shift of the NEXT tile goes to the other accumulator's copy
        if (has_bias && etid < BN && tile + tile_step < total_tiles) {
          const int ch = tile_coord<kPair>(p, tile + tile_step, BN, rank).c0 + etid;
          bias_smem[(acc ^ 1) * 256 + etid] = ch < p.cout ? __ldg(p.bias + ch) : 0.f;
        }
        uint32_t mword[kSub];
#pragma unroll
        for (int j = 0; j < kSub; ++j) mword[j] = mnext[j];
        if (has_mask_in && tile + tile_step < total_tiles) load_gate(tile + tile_step, mnext);
        mbar_wait(&tfull_bar[acc], acc_phase);
        tc_fence_after();
        const uint32_t taddr0 = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(acc * BN + half * 32);
        // The TMEM load of sub-tile j+1 is issued as soon as sub-tile j has been packed to bf16 (its fp32 registers
        // are dead from there on), so it flies during the staging stores, the barrier and the TMA-store issue
        // instead of being waited for at the top of the next sub-tile (ncu: ~6 % of the epilogue samples).
        float f[32];
        tmem_ld32f(taddr0, f);
#pragma unroll
        for (int j = 0; j < kSub; ++j) {
          const uint32_t obase = out0 + (uint32_t)(oslot * kSubTileBytes);
          tmem_ld_wait();
          if (j == kSub - 1) {
            // every TMEM read of this accumulator has completed: hand it back to the MMA warp
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
              if constexpr (kPair) mbar_arrive_cluster(mapa_u32(&tempty_bar[acc], 0)); else mbar_arrive(&tempty_bar[acc]);
            }
          }
          if (has_bias) {
            const uint32_t bb = bias0 + (uint32_t)((acc * 256 + j * 64 + half * 32) * 4);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float4 b = lds_f4(bb + 16u * i);
              add2(f[4 * i], f[4 * i + 1], b.x, b.y);
              add2(f[4 * i + 2], f[4 * i + 3], b.z, b.w);
            }
          }
          if (has_mask_out) {
            uint32_t bits = 0;
#pragma unroll
            for (int i = 0; i < 32; ++i) bits |= (f[i] > 0.f ? 1u : 0u) << i;
            if (valid && tc.c0 + 64 * j < p.cout) p.mask_out[mword0 + (long long)j * p.mask_pixels * 2] = bits;
          }
          if (has_mask_in) {
            const uint32_t mw = mword[j];
#pragma unroll
            for (int i = 0; i < 32; ++i) f[i] = ((mw >> i) & 1u) ? f[i] : 0.f;
          }
          uint4 o[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            if (relu) {
              o[i].x = pack_bf16x2_relu(f[8 * i], f[8 * i + 1]); o[i].y = pack_bf16x2_relu(f[8 * i + 2], f[8 * i + 3]);
              o[i].z = pack_bf16x2_relu(f[8 * i + 4], f[8 * i + 5]); o[i].w = pack_bf16x2_relu(f[8 * i + 6], f[8 * i + 7]);
            } else {
              o[i].x = pack_bf16x2(f[8 * i], f[8 * i + 1]); o[i].y = pack_bf16x2(f[8 * i + 2], f[8 * i + 3]);
              o[i].z = pack_bf16x2(f[8 * i + 4], f[8 * i + 5]); o[i].w = pack_bf16x2(f[8 * i + 6], f[8 * i + 7]);
            }
          }
          if (j + 1 < kSub) tmem_ld32f(taddr0 + (uint32_t)((j + 1) * 64), f);
          if (n_out_slots == 0) {
            // direct mode: this thread's 32 channels (64 contiguous bytes of its pixel) go straight to global
            // memory — no staging tile, no barrier, and the smem goes to the operand ring instead
            if (valid && tc.c0 + 64 * j < p.cout) {
              uint4* gp = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.out) + pixel * p.cout + tc.c0 + 64 * j + half * 32);
#pragma unroll
              for (int i = 0; i < 4; ++i) gp[i] = o[i];
            }
            continue;
          }
#pragma unroll
          for (int i = 0; i < 4; ++i) sts128(obase + row_off + ((((uint32_t)(half * 4 + i)) ^ sw) << 4), o[i]);
          fence_async_smem();          // generic-proxy smem writes -> visible to the TMA store
          epi_bar();
          if (leader) {
            if (tc.c0 + 64 * j < p.cout)
              tma_store_4d(&maps.out[tc.cls], out_smem + oslot * kSubTileBytes, tc.c0 + 64 * j, tc.w0, tc.h0, tc.n0);
            bulk_commit();
            // n staging slots: before the next barrier all but the newest (n - 2) stores must have
            // finished reading their slot
            if (n_out_slots == 4) bulk_wait_read<2>(); else bulk_wait_read<0>();
          }
          oslot = (oslot + 1) & (n_out_slots - 1);
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
      if (leader) bulk_wait_read<0>();
    }
  }

  tc_fence_before();
  if constexpr (kPair) cluster_sync_all(); else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    if constexpr (kPair)
      asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(C::kTmemCols) : "memory");
    else
      asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(C::kTmemCols) : "memory");
  }
}

template <int BN, bool kPair>
int launch_impl(const ConvGemmMaps& maps, ConvGemmParams& p, cudaStream_t stream) {
  using C = Cfg<BN, kPair>;
  static bool configured = false;
  if (!configured) {
    ADV_CHECK_CUDA(cudaFuncSetAttribute(conv_gemm_kernel<BN, kPair>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        kSmemBudget));
    configured = true;
  }
  const long long pixel_tiles = (long long)p.tiles_w * p.tiles_h * p.tiles_n;
  const int nclass = p.nclass > 1 ? p.nclass : 1;
  p.units_per_class = (int)(kPair ? (pixel_tiles + 1) / 2 : pixel_tiles);
  const long long total = (long long)p.units_per_class * nclass * p.tiles_c;
  if (total <= 0) return ADVB200_OK;
  p.total_tiles = (int)total;
  if (p.step_w == 0) p.step_w = p.bw;
  auto magic = [](int d) { return (unsigned long long)(((1ull << 40) + (unsigned long long)d - 1) / (unsigned long long)d); };
  auto ilog2 = [](int v) { int s = 0; while ((1 << s) < v) ++s; return s; };
  ADV_CHECK_ARG(total < (1ll << 20) && pixel_tiles * 2 < (1ll << 20), "conv_gemm: too many tiles for the fast division");
  ADV_CHECK_ARG((p.bw & (p.bw - 1)) == 0 && (p.bh & (p.bh - 1)) == 0, "conv_gemm: tile box must be a power of two");
  p.magic_c = magic(p.tiles_c); p.magic_w = magic(p.tiles_w); p.magic_h = magic(p.tiles_h);
  p.magic_cls = magic(p.units_per_class);
  p.bw_shift = ilog2(p.bw); p.bh_shift = ilog2(p.bh);
  p.ring_off = 0;
  p.stage_bytes = C::kStageBytes;
  if (p.ksteps == 0) {                       // plain conv: every tap spans all kblocks of one weight matrix
    for (int t = 0; t < p.ntaps; ++t) { p.tap_kb[t] = p.kblocks; p.tap_bmap[t] = 0; }
    p.ksteps = p.ntaps * p.kblocks;
  }
  if (nclass > 1) {
    ADV_CHECK_ARG(!p.halo && nclass == 4, "conv_gemm: multi-class launches are generic-mode with the 4 parity classes of a stride-2 gradient");
    p.bres_off = 1;                          // classes have different weights: no resident-weight mode
    p.ksteps = 0;
    for (int c = 0; c < nclass; ++c) {
      int ks = 0;
      for (int t = p.cls_tap0[c]; t < p.cls_tap0[c] + p.cls_ntaps[c]; ++t) ks += p.tap_kb[t];
      p.cls_ksteps[c] = ks;
      if (ks > p.ksteps) p.ksteps = ks;
    }
  }
  if (p.halo) {
    ADV_CHECK_ARG(p.kblocks == 1 && p.tiles_c == 1 && !p.has_add && (!kPair || p.epi == 0),
                  "conv_gemm: halo mode needs K = 64 and one column block (CTA pairs: plain bf16 epilogue only)");
    const int rb = p.row_bytes ? p.row_bytes : 128;
    ADV_CHECK_ARG(rb == 128 || rb == 32, "conv_gemm: halo rows are 128 or 32 bytes");
    p.ring_off = (p.ntaps * C::kBRows * rb + 1023) & ~1023;
    p.stage_bytes = (p.halo_rows * rb + 1023) & ~1023;
  }
  static const char* force_slots = getenv("ADVB200_OUT_SLOTS");   // tuning knob
  p.out_slots = p.slots_req == 1 ? 0 : p.slots_req ? p.slots_req : force_slots ? atoi(force_slots) : (p.ksteps >= 8 ? 2 : 4);
  int units = kPair ? kNumSMs / 2 : kNumSMs;            // CTAs or CTA pairs
  static const char* cls_env = getenv("ADVB200_CLS_ORDER");   // A/B knob: "major" / "minor" force one order
  p.cls_minor = nclass > 1 && (cls_env ? cls_env[1] == 'i' : total >= 8ll * units);   // interleave only in many-wave launches
  // resident-weight mode (pointwise convs): when all k-blocks of one column block fit beside a >= 3-deep ring of
  // A tiles, each CTA keeps one column block for all of its tiles (tile stride = multiple of tiles_c) and only
  // activations stream through the ring: halves the L2 -> SM traffic of the wide 1x1 layers
  static const char* no_bres = getenv("ADVB200_NO_BRES");
  p.bres = 0;
  if (!p.halo && !no_bres && BN != 16 && !p.bres_off && p.cout % BN == 0) {
    const int res_bytes = p.ksteps * C::kBTileBytes;
    const int slots = p.out_slots == 4 && res_bytes > 64 * 1024 ? 2 : p.out_slots;
    const int ring_room = kSmemBudget - Smem<BN>::fixed_bytes(slots) - res_bytes;
    const int bres_units = units - units % p.tiles_c;
    if (ring_room >= 3 * kATileBytes && bres_units * 16 >= units * 15 && total >= bres_units) {
      p.bres = 1; p.ring_off = res_bytes; p.stage_bytes = kATileBytes; p.out_slots = slots; units = bres_units;
    }
  }
  int stages = (kSmemBudget - Smem<BN>::fixed_bytes(p.out_slots) - p.ring_off) / p.stage_bytes;
  if (stages > kMaxStages) stages = kMaxStages;
  if (stages < 2) return fail(ADVB200_ERR_ARG, "conv_gemm: tile configuration does not fit shared memory");
  p.stages = stages;
  p.aux_depth = 0;
  const int smem_bytes = p.ring_off + stages * p.stage_bytes + Smem<BN>::fixed_bytes(p.out_slots);
  const int grid = (int)(total < units ? total : units) * (kPair ? 2 : 1);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(kThreads); cfg.dynamicSmemBytes = smem_bytes; cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kPair ? 2 : 1; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  // programmatic dependent launch: this grid may start while the previous kernel of the stream drains; the
  // kernel orders itself behind it with griddepcontrol.wait
  static const bool no_pdl = getenv("ADVB200_NO_PDL") != nullptr;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = no_pdl ? 1 : 2;
  ADV_CHECK_CUDA(cudaLaunchKernelEx(&cfg, conv_gemm_kernel<BN, kPair>, maps, p));
  ADV_LAUNCHED();
  return ADVB200_OK;
}


// ================================================================ fused 3x3 (halo) + pointwise kernel
// Warp roles as in conv_gemm_kernel.  Per tile: [MMA] 36 N=64 MMAs -> D1 (TMEM cols 0..63); [epilogue] D1 -> shift,
// ReLU / masks -> bf16 -> swizzled `mid` tile in shared memory; [MMA] 4 N=256 MMAs, A = mid, B = resident pointwise
// weights -> D2 (TMEM cols 64..319); [epilogue] D2 + residual (global loads) + shift -> ReLU / masks -> bf16 ->
// 64-byte runs straight to global memory.  The second epilogue of tile i overlaps the first GEMM of tile i+1; the
// kernel is HBM-bound (residual read + output write), so nothing more is pipelined.
constexpr int kFbW1Bytes = 9 * 64 * 128;
constexpr int kFbW2Bytes = 256 * 128;
constexpr int kFbHaloRows = 10 * 18;
constexpr int kFbHaloBytes = (kFbHaloRows * 128 + 1023) & ~1023;
constexpr int kFbMidBytes = kBlockM * 128;
constexpr int kFbBiasBytes = (64 + 256) * 4;
constexpr int kFbBarBytes = 256;
constexpr int kFbMaxStages = 4;
constexpr int kFbResSlots = 5;                 // residual sub-tile ring; each slot is also its sub-tile's output staging

struct FbCoord { int w0, h0, n; };
__device__ __forceinline__ FbCoord fb_coord(const FusedBlockParams& p, int tile) {
  FbCoord c;
  int q = (int)(((unsigned long long)(unsigned)tile * p.magic_w) >> 40);
  c.w0 = (tile - q * p.tiles_w) * 8; tile = q;
  q = (int)(((unsigned long long)(unsigned)tile * p.magic_h) >> 40);
  c.h0 = (tile - q * p.tiles_h) * 16; c.n = q;
  return c;
}

constexpr int kFbThreads = (2 + 8 + 4) * 32;   // TMA, MMA, 8 second-epilogue warps, 4 first-epilogue warps

__global__ void __launch_bounds__(kFbThreads, 1)
fused_block_kernel(const __grid_constant__ FusedBlockMaps maps, const __grid_constant__ FusedBlockParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* w1_smem = smem;
  uint8_t* w2_smem = smem + kFbW1Bytes;
  uint8_t* ring = w2_smem + kFbW2Bytes;
  uint8_t* mid_smem = ring + p.stages * kFbHaloBytes;
  uint8_t* res_smem = mid_smem + kFbMidBytes;                                    // kFbResSlots x 16 KB
  float* bias_smem = reinterpret_cast<float*>(res_smem + kFbResSlots * kSubTileBytes);   // [64] then [256]
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(bias_smem) + kFbBiasBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kFbMaxStages;
  uint64_t* wres_bar = bars + 2 * kFbMaxStages;
  uint64_t* d1_full = wres_bar + 1;
  uint64_t* mid_full = wres_bar + 2;
  uint64_t* d2_full = wres_bar + 3;
  uint64_t* d2_empty = wres_bar + 4;
  uint64_t* res_full = wres_bar + 5;
  uint64_t* res_empty = res_full + kFbResSlots;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(res_empty + kFbResSlots);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int nstages = p.stages;
  const int total_tiles = p.total_tiles;

  if (warp == 0 && lane == 0) {
    if (smem_u32(smem) & 1023u) __trap();
    prefetch_tmap(&maps.a); prefetch_tmap(&maps.w1); prefetch_tmap(&maps.w2);
    prefetch_tmap(&maps.res); prefetch_tmap(&maps.out);
    for (int s = 0; s < kFbMaxStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int s = 0; s < kFbResSlots; ++s) { mbar_init(&res_full[s], 1); mbar_init(&res_empty[s], 1); }
    mbar_init(wres_bar, 1); mbar_init(d1_full, 1); mbar_init(mid_full, 1); mbar_init(d2_full, 1);
    mbar_init(d2_empty, 8);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (warp >= kEpiWarp0) {
    const int i = threadIdx.x - kEpiWarp0 * 32;
    if (i < 64) bias_smem[i] = p.bias1 ? __ldg(p.bias1 + i) : 0.f;
    if (i < 256) bias_smem[64 + i] = p.bias2 ? __ldg(p.bias2 + i) : 0.f;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // =========================== TMA producer ===========================
    if (lane == 0) {
      pdl_launch_dependents();
      mbar_expect_tx(wres_bar, (uint32_t)(kFbW1Bytes + kFbW2Bytes));
      for (int tap = 0; tap < 9; ++tap) tma_load_2d(&maps.w1, wres_bar, w1_smem + tap * 64 * 128, 0, p.tap_brow[tap]);
      tma_load_2d(&maps.w2, wres_bar, w2_smem, 0, 0);
      pdl_wait();
      int stage = 0, rs = 0;
      uint32_t phase = 0, rphase = 0;
      auto load_halo = [&](int tile) {
        const FbCoord tc = fb_coord(p, tile);
        mbar_wait(&empty_bar[stage], phase ^ 1);
        mbar_expect_tx(&full_bar[stage], (uint32_t)(kFbHaloRows * 128));
        tma_load_4d(&maps.a, &full_bar[stage], ring + stage * kFbHaloBytes, 0, tc.w0 - 1, tc.h0 - 1, tc.n);
        if (++stage == nstages) { stage = 0; phase ^= 1; }
      };
      // one halo stage is enough: the box of tile i+1 is requested as soon as the first GEMM of tile i has drained it,
      // a whole (first epilogue + second GEMM + second epilogue) ahead of its use
      if ((int)blockIdx.x < total_tiles) load_halo(blockIdx.x);
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const FbCoord tc = fb_coord(p, tile);
        for (int j = 0; j < 4; ++j) {
          mbar_wait(&res_empty[rs], rphase ^ 1);
          mbar_expect_tx(&res_full[rs], (uint32_t)kSubTileBytes);
          tma_load_4d(&maps.res, &res_full[rs], res_smem + rs * kSubTileBytes, 64 * j, tc.w0, tc.h0, tc.n);
          if (++rs == kFbResSlots) { rs = 0; rphase ^= 1; }
        }
        if (tile + (int)gridDim.x < total_tiles) load_halo(tile + gridDim.x);
      }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    constexpr uint32_t idesc64 = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(kBlockM >> 4) << 24);
    constexpr uint32_t idesc256 = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(256 >> 3) << 17) | ((uint32_t)(kBlockM >> 4) << 24);
    const uint32_t dense_hi = (uint32_t)(make_smem_desc(0) >> 32);
    const uint32_t halo_hi = (uint32_t)(make_smem_desc(0, 10u * 128u) >> 32);
    const uint32_t ring_lo = ((smem_u32(ring) & 0x3FFFFu) >> 4) | (1u << 16);
    const uint32_t w1_lo = ((smem_u32(w1_smem) & 0x3FFFFu) >> 4) | (1u << 16);
    const uint32_t w2_lo = ((smem_u32(w2_smem) & 0x3FFFFu) >> 4) | (1u << 16);
    const uint32_t mid_lo = ((smem_u32(mid_smem) & 0x3FFFFu) >> 4) | (1u << 16);
    const uint32_t d1 = tmem_base, d2 = tmem_base + 64u;
    int stage = 0;
    uint32_t phase = 0, par = 0;
    if (blockIdx.x < total_tiles) mbar_wait(wres_bar, 0);
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      mbar_wait(&full_bar[stage], phase);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t a_lo0 = ring_lo + (uint32_t)stage * (uint32_t)(kFbHaloBytes >> 4);
        for (int tap = 0; tap < 9; ++tap) {
          const uint32_t a_lo = a_lo0 + (uint32_t)(p.tap_row[tap] * 8);
          const uint32_t b_lo = w1_lo + (uint32_t)(tap * 64 * 8);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_bf16(d1, ((uint64_t)halo_hi << 32) | (a_lo + 2u * k), ((uint64_t)dense_hi << 32) | (b_lo + 2u * k), idesc64,
                      (tap | k) != 0 ? 1u : 0u);
        }
        umma_commit(&empty_bar[stage]);
        umma_commit(d1_full);
      }
      __syncwarp();
      if (++stage == nstages) { stage = 0; phase ^= 1; }
      mbar_wait(mid_full, par);            // the intermediate tile is in shared memory (and D1 has been read)
      mbar_wait(d2_empty, par ^ 1);        // the previous tile's D2 has been read
      tc_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_bf16(d2, ((uint64_t)dense_hi << 32) | (mid_lo + 2u * k), ((uint64_t)dense_hi << 32) | (w2_lo + 2u * k), idesc256,
                    k != 0 ? 1u : 0u);
        umma_commit(d2_full);
      }
      __syncwarp();
      par ^= 1;
    }
  } else if (warp >= kEpiWarp0 + 8) {
    // =========================== first epilogue (4 warps, one per TMEM lane quadrant) ===========================
    // D1 -> shift, ReLU / masks -> bf16 -> `mid` tile.  Its own warps, so it runs while the other eight are still in
    // the second epilogue of the previous tile.
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
    const uint32_t mid0 = smem_u32(mid_smem);
    const uint32_t bias0 = smem_u32(bias_smem);
    const uint32_t row_off = (uint32_t)row * 128u;
    const uint32_t sw = (uint32_t)(row & 7);
    const int wi = row & 7, hi = row >> 3;
    const bool relu = p.relu != 0;
    const bool leader1 = threadIdx.x == (kEpiWarp0 + 8) * 32;
    uint32_t par = 0;
    pdl_wait();
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const FbCoord tc = fb_coord(p, tile);
      const int w = tc.w0 + wi, h = tc.h0 + hi;
      const bool valid = w < p.img_w && h < p.img_h && tc.n < p.img_n;
      const long long pixel = ((long long)tc.n * p.img_h + h) * p.img_w + w;
      uint32_t m1[2] = {0xffffffffu, 0xffffffffu};
      if (p.mask1_in) {
        const uint2 mm = valid ? __ldg(reinterpret_cast<const uint2*>(p.mask1_in + pixel * 2)) : make_uint2(0u, 0u);
        m1[0] = mm.x; m1[1] = mm.y;
      }
      mbar_wait(d1_full, par);
      tc_fence_after();
      float f[32];
      tmem_ld32f(tmem_base + lane_off, f);
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        tmem_ld_wait();
        if (half == 1) tc_fence_before();
        if (p.bias1) {
          const uint32_t bb = bias0 + (uint32_t)(half * 32 * 4);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 b = lds_f4(bb + 16u * i);
            add2(f[4 * i], f[4 * i + 1], b.x, b.y);
            add2(f[4 * i + 2], f[4 * i + 3], b.z, b.w);
          }
        }
        if (relu && p.mask1_out) {
          uint32_t bits = 0;
#pragma unroll
          for (int i = 0; i < 32; ++i) bits |= (f[i] > 0.f ? 1u : 0u) << i;
          if (valid) p.mask1_out[pixel * 2 + half] = bits;
        }
        if (p.mask1_in) {
          const uint32_t mw = m1[half];
#pragma unroll
          for (int i = 0; i < 32; ++i) f[i] = ((mw >> i) & 1u) ? f[i] : 0.f;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint4 o;
          if (relu) {
            o.x = pack_bf16x2_relu(f[8 * i], f[8 * i + 1]); o.y = pack_bf16x2_relu(f[8 * i + 2], f[8 * i + 3]);
            o.z = pack_bf16x2_relu(f[8 * i + 4], f[8 * i + 5]); o.w = pack_bf16x2_relu(f[8 * i + 6], f[8 * i + 7]);
          } else {
            o.x = pack_bf16x2(f[8 * i], f[8 * i + 1]); o.y = pack_bf16x2(f[8 * i + 2], f[8 * i + 3]);
            o.z = pack_bf16x2(f[8 * i + 4], f[8 * i + 5]); o.w = pack_bf16x2(f[8 * i + 6], f[8 * i + 7]);
          }
          sts128(mid0 + row_off + ((((uint32_t)(half * 4 + i)) ^ sw) << 4), o);
        }
        if (half == 0) tmem_ld32f(tmem_base + lane_off + 32u, f);   // second half in flight during the stores
      }
      fence_async_smem();                  // generic-proxy writes -> visible to the tensor core's operand reads
      asm volatile("bar.sync 2, 128;" ::: "memory");
      if (leader1) mbar_arrive(mid_full);
      par ^= 1;
    }
  } else {
    // =========================== second epilogue (8 warps) ===========================
    const int ew = warp - kEpiWarp0;
    const int quad = warp & 3;
    const int half = ew >> 2;
    const int row = quad * 32 + lane;
    const int etid = threadIdx.x - kEpiWarp0 * 32;
    const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
    const uint32_t res0 = smem_u32(res_smem);
    const uint32_t bias0 = smem_u32(bias_smem);
    const uint32_t row_off = (uint32_t)row * 128u;
    const uint32_t sw = (uint32_t)(row & 7);
    const int wi = row & 7, hi = row >> 3;
    const bool relu = p.relu != 0;
    int rs = 0, prev_rs = -1;
    uint32_t rphase = 0;
    uint32_t par = 0;
    uint32_t m2n[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};
    auto load_gate2 = [&](int tile_) {
      const FbCoord t2 = fb_coord(p, tile_);
      const int w2 = t2.w0 + wi, h2 = t2.h0 + hi;
      const bool ok = w2 < p.img_w && h2 < p.img_h && t2.n < p.img_n;
      const long long px2 = ((long long)t2.n * p.img_h + h2) * p.img_w + w2;
#pragma unroll
      for (int j = 0; j < 4; ++j) m2n[j] = ok ? __ldg(p.mask2_in + ((long long)j * p.mask_pixels + px2) * 2 + half) : 0u;
    };
    pdl_wait();
    if (p.mask2_in && (int)blockIdx.x < total_tiles) load_gate2(blockIdx.x);
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const FbCoord tc = fb_coord(p, tile);
      const int w = tc.w0 + wi, h = tc.h0 + hi;
      const bool valid = w < p.img_w && h < p.img_h && tc.n < p.img_n;
      const long long pixel = ((long long)tc.n * p.img_h + h) * p.img_w + w;
      // gate words of this thread's pixel for the four output sub-tiles (backward), fetched one tile ahead
      uint32_t m2[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) m2[j] = m2n[j];
      if (p.mask2_in && tile + (int)gridDim.x < total_tiles) load_gate2(tile + (int)gridDim.x);
      // ---- second epilogue: D2 + residual + shift -> output.  The residual sub-tile arrives by TMA in a ring slot;
      // each thread reads its row's 64 bytes from it, writes the result back IN PLACE, and the slot is TMA-stored.
      mbar_wait(d2_full, par);
      tc_fence_after();
      float f[32];
      tmem_ld32f(tmem_base + lane_off + (uint32_t)(64 + half * 32), f);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        mbar_wait(&res_full[rs], rphase);
        const uint32_t slot = res0 + (uint32_t)(rs * kSubTileBytes) + row_off;
        uint4 r[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) r[i] = lds128(slot + ((((uint32_t)(half * 4 + i)) ^ sw) << 4));
        tmem_ld_wait();
        if (j == 3) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(d2_empty);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const uint32_t u[4] = {r[i].x, r[i].y, r[i].z, r[i].w};
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            f[8 * i + 2 * q] += bf16_lo(u[q]);
            f[8 * i + 2 * q + 1] += bf16_hi(u[q]);
          }
        }
        if (p.bias2) {
          const uint32_t bb = bias0 + (uint32_t)((64 + j * 64 + half * 32) * 4);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 b = lds_f4(bb + 16u * i);
            add2(f[4 * i], f[4 * i + 1], b.x, b.y);
            add2(f[4 * i + 2], f[4 * i + 3], b.z, b.w);
          }
        }
        if (relu && p.mask2_out) {
          uint32_t bits = 0;
#pragma unroll
          for (int i = 0; i < 32; ++i) bits |= (f[i] > 0.f ? 1u : 0u) << i;
          if (valid) p.mask2_out[((long long)j * p.mask_pixels + pixel) * 2 + half] = bits;
        }
        if (p.mask2_in) {
#pragma unroll
          for (int i = 0; i < 32; ++i) f[i] = ((m2[j] >> i) & 1u) ? f[i] : 0.f;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint4 o;
          if (relu) {
            o.x = pack_bf16x2_relu(f[8 * i], f[8 * i + 1]); o.y = pack_bf16x2_relu(f[8 * i + 2], f[8 * i + 3]);
            o.z = pack_bf16x2_relu(f[8 * i + 4], f[8 * i + 5]); o.w = pack_bf16x2_relu(f[8 * i + 6], f[8 * i + 7]);
          } else {
            o.x = pack_bf16x2(f[8 * i], f[8 * i + 1]); o.y = pack_bf16x2(f[8 * i + 2], f[8 * i + 3]);
            o.z = pack_bf16x2(f[8 * i + 4], f[8 * i + 5]); o.w = pack_bf16x2(f[8 * i + 6], f[8 * i + 7]);
          }
          sts128(slot + ((((uint32_t)(half * 4 + i)) ^ sw) << 4), o);
        }
        // next sub-tile's accumulator columns: in flight during the barrier and the store issue
        if (j < 3) tmem_ld32f(tmem_base + lane_off + (uint32_t)(64 + (j + 1) * 64 + half * 32), f);
        fence_async_smem();
        epi_bar();
        if (etid == 0) {
          tma_store_4d(&maps.out, res_smem + rs * kSubTileBytes, 64 * j, tc.w0, tc.h0, tc.n);
          bulk_commit();
          bulk_wait_read<1>();                 // the store issued one sub-tile ago has finished reading its slot
          if (prev_rs >= 0) mbar_arrive(&res_empty[prev_rs]);
          prev_rs = rs;
        }
        if (++rs == kFbResSlots) { rs = 0; rphase ^= 1; }
      }
      par ^= 1;
    }
    if (etid == 0) bulk_wait_read<0>();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
  }
}

}  // namespace

int launch_fused_block(const FusedBlockMaps& maps, FusedBlockParams& p, cudaStream_t stream) {
  static bool configured = false;
  if (!configured) {
    ADV_CHECK_CUDA(cudaFuncSetAttribute(fused_block_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBudget));
    configured = true;
  }
  ADV_CHECK_ARG(p.img_w % 8 == 0 && p.tiles_w == p.img_w / 8 && p.tiles_h == (p.img_h + 15) / 16 && p.tiles_n == p.img_n,
                "fused_block: tiles are 8 x 16 pixels of one image");
  const long long total = (long long)p.tiles_w * p.tiles_h * p.tiles_n;
  if (total <= 0) return ADVB200_OK;
  ADV_CHECK_ARG(total < (1ll << 20), "fused_block: too many tiles for the fast division");
  p.total_tiles = (int)total;
  auto magic = [](int d) { return (unsigned long long)(((1ull << 40) + (unsigned long long)d - 1) / (unsigned long long)d); };
  p.magic_w = magic(p.tiles_w); p.magic_h = magic(p.tiles_h);
  const int fixed = kFbW1Bytes + kFbW2Bytes + kFbMidBytes + kFbResSlots * kSubTileBytes + kFbBiasBytes + kFbBarBytes;
  const int stages = 1;                                    // see the producer: one halo stage suffices
  static_assert(kFbW1Bytes + kFbW2Bytes + kFbMidBytes + kFbResSlots * kSubTileBytes + kFbBiasBytes + kFbBarBytes +
                    kFbHaloBytes <= kSmemBudget, "fused_block: shared memory carve-up");
  p.stages = stages;
  const int smem_bytes = fixed + stages * kFbHaloBytes;
  const int grid = (int)(total < kNumSMs ? total : kNumSMs);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(kFbThreads); cfg.dynamicSmemBytes = smem_bytes; cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  static const bool no_pdl = getenv("ADVB200_NO_PDL") != nullptr;
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = no_pdl ? 0 : 1;
  ADV_CHECK_CUDA(cudaLaunchKernelEx(&cfg, fused_block_kernel, maps, p));
  ADV_LAUNCHED();
  return ADVB200_OK;
}

int launch_conv_gemm(int bn_cols, int pair, const ConvGemmMaps& maps, ConvGemmParams& p, cudaStream_t stream) {
  ADV_CHECK_ARG(p.bw * p.bh * p.bn == kBlockM, "conv_gemm: tile must hold 128 pixels");
  ADV_CHECK_ARG(p.ntaps >= 1 && p.ntaps <= kMaxTaps && (p.kblocks >= 1 || p.ksteps >= 1), "conv_gemm: bad K loop");
  switch (bn_cols) {
    case 16: return launch_impl<16, false>(maps, p, stream);
    case 64: return pair ? launch_impl<64, true>(maps, p, stream) : launch_impl<64, false>(maps, p, stream);
    case 128: return pair ? launch_impl<128, true>(maps, p, stream) : launch_impl<128, false>(maps, p, stream);
    case 192: return pair ? launch_impl<192, true>(maps, p, stream) : launch_impl<192, false>(maps, p, stream);
    case 256: return pair ? launch_impl<256, true>(maps, p, stream) : launch_impl<256, false>(maps, p, stream);
    default: return fail(ADVB200_ERR_ARG, "conv_gemm: BN must be 16, 64, 128 or 256");
  }
}

int make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                   const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes) {
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
                      gstr, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      swizzle_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_128B,
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

// Hardware probe: does tcgen05.mma accept K-major SWIZZLE_128B operand descriptors whose start address
// is shifted by whole 128-byte rows (not 1024-byte aligned), and strides between 8-row groups other
// than 1024 bytes?  (Needed to reuse one halo tile in shared memory for all taps of a 3x3 conv.)
//   nvcc -gencode arch=compute_100a,code=sm_100a -o umma_probe tools/umma_probe.cu && ./umma_probe
#include <cuda.h>
#include <cudaTypedefs.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

constexpr int kRows = 256;   // rows of the A matrix staged in smem (32 KB)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(128, 1)
probe(const __grid_constant__ CUtensorMap ma, const __grid_constant__ CUtensorMap mb, float* out, int shift_rows,
      int sbo_bytes, int base_off) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sa = smem;                 // kRows x 128 B
  uint8_t* sb = smem + kRows * 128;   // 64 x 128 B
  uint64_t* bar = reinterpret_cast<uint64_t*>(sb + 64 * 128);
  uint64_t* bar2 = bar + 1;
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bar + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar2)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(smem_u32(tslot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *tslot;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(kRows * 128 + 64 * 128) : "memory");
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(smem_u32(sa)), "l"(reinterpret_cast<uint64_t>(&ma)), "r"(smem_u32(bar)), "r"(0), "r"(0) : "memory");
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(smem_u32(sb)), "l"(reinterpret_cast<uint64_t>(&mb)), "r"(smem_u32(bar)), "r"(0), "r"(0) : "memory");
    asm volatile("{\n.reg .pred p;\nW0:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n@p bra D0;\nbra W0;\nD0:\n}" ::"r"(smem_u32(bar)) : "memory");
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    auto desc = [&](uint32_t addr, int sbo, int boff) {
      uint64_t d = 0;
      d |= (uint64_t)((addr & 0x3FFFF) >> 4);
      d |= (uint64_t)1 << 16;
      d |= (uint64_t)(sbo >> 4) << 32;
      d |= (uint64_t)1 << 46;
      d |= (uint64_t)(boff & 7) << 49;
      d |= (uint64_t)2 << 61;
      return d;
    };
    const uint64_t ad = desc(smem_u32(sa) + shift_rows * 128, sbo_bytes, base_off);
    const uint64_t bd = desc(smem_u32(sb), 1024, 0);
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    for (int k = 0; k < 4; ++k) {
      const uint32_t accum = k ? 1u : 0u;
      asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}"
                   ::"r"(tmem), "l"(ad + 2 * k), "l"(bd + 2 * k), "r"(idesc), "r"(accum) : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar2)) : "memory");
  }
  // everyone waits for the MMA
  asm volatile("{\n.reg .pred p;\nW1:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n@p bra D1;\nbra W1;\nD1:\n}" ::"r"(smem_u32(bar2)) : "memory");
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  uint32_t v[32];
  for (int h = 0; h < 2; ++h) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]),
          "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]),
          "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]),
          "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(tmem + ((uint32_t)(warp * 32) << 16) + h * 32));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int i = 0; i < 32; ++i) out[(warp * 32 + lane) * 64 + h * 32 + i] = __uint_as_float(v[i]);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(tmem) : "memory");
}

int main() {
  PFN_cuTensorMapEncodeTiled_v12000 encode = nullptr;
  cudaDriverEntryPointQueryResult q;
  void* fn = nullptr;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
  encode = (PFN_cuTensorMapEncodeTiled_v12000)fn;
  std::vector<__nv_bfloat16> ha(kRows * 64), hb(64 * 64);
  std::vector<float> fa(kRows * 64), fb(64 * 64);
  for (int r = 0; r < kRows; ++r) for (int c = 0; c < 64; ++c) { float v = (float)((r * 7 + c * 3) % 13 - 6); fa[r * 64 + c] = v; ha[r * 64 + c] = __float2bfloat16(v); }
  for (int r = 0; r < 64; ++r) for (int c = 0; c < 64; ++c) { float v = (float)((r * 5 + c * 11) % 7 - 3); fb[r * 64 + c] = v; hb[r * 64 + c] = __float2bfloat16(v); }
  __nv_bfloat16 *da, *db; float* dout;
  CK(cudaMalloc(&da, ha.size() * 2)); CK(cudaMalloc(&db, hb.size() * 2)); CK(cudaMalloc(&dout, 128 * 64 * 4));
  CK(cudaMemcpy(da, ha.data(), ha.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(db, hb.data(), hb.size() * 2, cudaMemcpyHostToDevice));
  auto mk = [&](CUtensorMap* m, void* base, int rows) {
    cuuint64_t gd[2] = {64, (cuuint64_t)rows}; cuuint64_t gs[1] = {128}; cuuint32_t bx[2] = {64, (cuuint32_t)rows}; cuuint32_t es[2] = {1, 1};
    CUresult r = encode(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); exit(1); }
  };
  CUtensorMap ma, mb;
  mk(&ma, da, kRows); mk(&mb, db, 64);
  const int smem = kRows * 128 + 64 * 128 + 64;
  CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  std::vector<float> ho(128 * 64);
  struct V { int shift, sbo, boff; const char* what; };
  const V vs[] = {
      {0, 1024, 0, "aligned baseline"},
      {8, 1024, 0, "shift 8 rows (1024B aligned)"},
      {1, 1024, 0, "shift 1 row, base_offset 0"},
      {1, 1024, 1, "shift 1 row, base_offset 1"},
      {3, 1024, 0, "shift 3 rows, base_offset 0"},
      {3, 1024, 3, "shift 3 rows, base_offset 3"},
      {0, 1280, 0, "groups of 8 rows every 10 rows (SBO 1280), base_offset 0"},
      {1, 1280, 0, "shift 1 + SBO 1280, base_offset 0"},
      {1, 1280, 1, "shift 1 + SBO 1280, base_offset 1"},
      {2, 768, 0, "groups of 8 rows every 6 rows (SBO 768, overlapping), shift 2"},
      {0, 2048, 0, "SBO 2048 (every other atom)"},
  };
  for (const V& v : vs) {
    CK(cudaMemset(dout, 0, 128 * 64 * 4));
    probe<<<1, 128, smem>>>(ma, mb, dout, v.shift, v.sbo, v.boff);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%-60s : CUDA error %s\n", v.what, cudaGetErrorString(e)); return 1; }
    CK(cudaMemcpy(ho.data(), dout, 128 * 64 * 4, cudaMemcpyDeviceToHost));
    double maxerr = 0; int bad = 0;
    for (int m = 0; m < 128; ++m) {
      const int pr = v.shift + (m / 8) * (v.sbo / 128) + (m % 8);
      for (int n = 0; n < 64; ++n) {
        float ref = 0;
        for (int k = 0; k < 64; ++k) ref += fa[pr * 64 + k] * fb[n * 64 + k];
        const double err = fabs(ref - ho[m * 64 + n]);
        if (err > maxerr) maxerr = err;
        if (err > 0.5) ++bad;
      }
    }
    printf("%-60s : max err %.1f, wrong %d / 8192 -> %s\n", v.what, maxerr, bad, bad ? "MISMATCH" : "OK");
  }
  return 0;
}

// tcgen05 / TMEM / TMA implicit-GEMM convolution for sm_100a — host-visible interface.
//
// One kernel covers every bf16 convolution of the ResNet-50 forward and input-gradient pass:
//   D[128 pixels, BN channels] = sum over taps t, k-blocks kb of  A_t[pixels shifted by tap t, 64 ch]
//                                                                  x B[t*rows + ch_out, 64 ch]^T
// A tiles are TMA boxes (64 ch, bw, bh, bn) of an NHWC activation (or one of its stride-2
// sub-lattices), out-of-bounds pixels are zero-filled by TMA (that is the conv padding);
// B tiles are K-major packed weights; accumulators live in TMEM; the epilogue fuses
// bias (folded BN shift), residual add, ReLU, and the ReLU mask of the input-gradient pass.
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace advb200 {

constexpr int kMaxTaps = 16;

struct ConvGemmParams {
  // tiling: a tile is bw x bh x bn pixels (product 128) by BN output channels
  int tiles_w, tiles_h, tiles_n, tiles_c;
  int bw, bh, bn;
  int step_w, org_w;              // tile origin along w = tile_w * step_w + org_w (0 -> step_w = bw): tiles whose
                                  // boxes overlap (stem input gradient keeps its horizontal halo in the tile)
  int ntaps, kblocks;             // K loop: ntaps * kblocks blocks of 64 input channels
  int tap_map[kMaxTaps];          // which A tensor map the tap reads
  int tap_dw[kMaxTaps];           // pixel shift of the tap along w / h
  int tap_dh[kMaxTaps];
  int tap_brow[kMaxTaps];         // first row of the tap's weights in the packed B matrix
  // fused launches (two 1x1 convs accumulating into one output, e.g. conv3 + downsample): taps may read
  // different weight matrices (tap_bmap: 0 -> maps.b, 1 -> maps.b2) with different depths (tap_kb).
  // ksteps = sum of tap_kb; leave 0 for a plain conv (filled from ntaps * kblocks).
  int tap_kb[kMaxTaps];
  int tap_bmap[kMaxTaps];
  int ksteps;
  int has_b2;
  // multi-class launches (stride-2 input gradients): the output splits into its 4 parity classes of equal tile count
  // (many-wave launches: unit index -> pixel tile = index / 4, class = (index + pixel tile) % 4, so the classes of a pixel
  // tile run side by side and share their input through the L2; few-wave launches: class-major); class c owns taps [cls_tap0[c], cls_tap0[c] + cls_ntaps[c]) (cls_ksteps[c] k-steps), writes the
  // (cls_ph[c], cls_pw[c]) sub-lattice of the output through maps.out[c].  nclass <= 1: a plain launch.
  int nclass;
  int cls_tap0[4], cls_ntaps[4], cls_ksteps[4], cls_ph[4], cls_pw[4];
  int units_per_class;            // tiles (or CTA-pair tiles) of one class and column block (launcher-filled)
  unsigned long long magic_cls;
  int cls_minor;                  // launcher-filled: 1 = class-minor unit order (>= 8 waves), 0 = class-major
  // halo mode (kblocks == 1, tiles_c == 1): ONE TMA box of (bw+halo) x (bh+halo) input pixels per tile
  // serves every tap — tap t reads the rows starting at tap_row[t], 8-row groups sbo_rows rows apart
  // (tcgen05 applies the 128B swizzle to absolute smem addresses, so row-shifted descriptors are legal;
  // verified by tools/umma_probe.cu) — and the whole weight matrix stays resident in shared memory.
  int halo;                       // 0 = one TMA box per tap (generic), 1 = halo mode
  int halo_w0, halo_h0;           // halo box origin relative to the tile origin (<= 0)
  int halo_rows;                  // pixels in the halo box
  int sbo_rows;                   // rows between consecutive 8-row groups (= halo box width)
  int tap_row[kMaxTaps];
  int row_bytes;                  // halo mode: bytes per operand row, 128 (64 channels, SWIZZLE_128B; 0 means 128) or
                                  // 32 (16 channels, SWIZZLE_32B: the stem, one K = 16 MMA per tap)
  // output extent in tile coordinates (w, h, n) and channels
  int out_w, out_h, out_n, cout;
  void* out;                      // epi 0: bf16 NHWC output tensor (direct-store epilogue); epi 1/2: fp32 NCHW
                                  // image gradient
  const float* bias;              // [cout] or null
  int has_add;                    // residual / other-branch gradient tile read through maps.add
  int relu;
  // ReLU bit masks, one bit per (pixel, channel): word (c/64, pixel, (c/32)&1), bit c%32.
  //   mask_in : gates the gradient written by a dgrad launch (bit set <=> forward output > 0)
  //   mask_out: written by a forward launch with relu (bit set <=> output > 0)
  const uint32_t* mask_in;
  uint32_t* mask_out;
  long long mask_pixels;          // pixels of the output tensor (plane stride of the mask)
  int pix_sub, pix_ph, pix_pw;    // output pixel (h, w) of tile coords (th, tw): (th*sub+ph, tw*sub+pw)
  int full_h, full_w;             // full-resolution output height / width
  int bres_off;                   // caller opt-out of the resident-weight mode
  int slots_req;                  // caller-requested output staging slots (0 = launcher default, 1 = none:
                                  // the epilogue stores straight to global memory through `out`)
  int bres;                       // chosen by launch_conv_gemm: weights resident, ring carries activations only
  int stages;                     // smem ring depth chosen by the host for this launch
  int out_slots;                  // output staging slots (2 or 4), host-chosen
  unsigned long long magic_c, magic_w, magic_h;   // ceil(2^40 / tiles_c|w|h): fast division in tile_coord
  int bw_shift, bh_shift;         // log2(bw), log2(bh)
  int reverse, total_tiles;       // traverse tiles last-to-first (alternating launches reuse the L2 tail)
  int stage_bytes, ring_off;      // ring slot size and offset of the ring (after resident weights)
  int aux_depth;                  // residual/mask prefetch ring depth (sub-tiles), host-chosen
  int epi;                        // 0: bf16 NHWC store; 1/2: stem input gradient (fp32 NCHW); 2 = horizontal
                                  // taps in the N dimension, reduced by warp shuffles (BN = 64)
  const float* std_dev;           // epi 1: device pointer to std[3]; gradient is divided by std[c]
  int img_h, img_w;               // epi 1: full-resolution image size
  // epi 2 only, fused attack step (attack.py:33-34 / 61-63): instead of storing the gradient, update upd_x in
  // place: x = clamp(project(x + alpha*sign(g))) — bit-identical to storing g and running linf_step on it
  float* upd_x;                   // x_adv (fp32 NCHW), null = store the gradient to `out`
  const float* upd_x0;            // x_orig of the eps-ball projection (null = no projection: FGSM)
  float upd_alpha, upd_eps;
};

struct ConvGemmMaps {
  CUtensorMap a[5];   // activation (or one per stride-2 sub-lattice; a[4]: second source of a fused multi-class launch)
  CUtensorMap b;      // packed weights
  CUtensorMap b2;     // second weight matrix of a fused launch
  CUtensorMap out[4]; // output tile store (64-channel boxes, same pixel box as a[]); one per parity class
  CUtensorMap add;    // optional, output-shaped
};

// Launch on `stream`.  bn_cols in {16, 64, 128, 256}.  pair != 0: CTA pairs (cta_group::2) on 256-pixel tiles;
// the weight tensor map must then have a box of bn_cols/2 rows.  Fills p.stages.
int launch_conv_gemm(int bn_cols, int pair, const ConvGemmMaps& maps, ConvGemmParams& p, cudaStream_t stream);

// ---- fused 3x3 (halo mode, 64 -> 64) + pointwise (64 -> 256, + residual) launch ------------------------------
// One layer1 bottleneck tail in one kernel: forward  out = relu(W3 . relu(conv3x3(a, W2) + b2) + b3 + x),
// backward  g_x = mask_x . (W1^T . (mask_a . dgrad3x3(g_b, W2)) + g_sum).  The 64-channel intermediate tile goes
// TMEM -> registers -> bf16 -> a swizzled shared-memory tile that is the A operand of the second GEMM, so it never
// touches HBM (nor does a second launch re-read it).  Tile = 8 x 16 pixels of one image.
struct FusedBlockParams {
  int tiles_w, tiles_h, tiles_n, total_tiles;
  unsigned long long magic_w, magic_h;
  int img_h, img_w, img_n;        // pixel grid (56 x 56 x N)
  int tap_row[9];                 // halo-box row shift of tap t (rows of 128 B; box is 10 wide)
  int tap_brow[9];                // first row of tap t's weights in the packed 3x3 matrix
  const float* bias1;             // folded-BN shift of the 3x3 conv (64) or null
  const float* bias2;             // shift of the pointwise conv (256) or null
  int relu;                       // forward: ReLU after both convs
  const uint32_t* mask1_in;       // backward: gates the intermediate gradient (ReLU bits of the 3x3's input act)
  uint32_t* mask1_out;            // forward: ReLU bits of the intermediate activation
  const uint32_t* mask2_in;       // backward: gates the output gradient (null for the first block)
  uint32_t* mask2_out;            // forward: ReLU bits of the block output
  long long mask_pixels;          // N * H * W
  const void* res;                // bf16 NHWC [N,H,W,256]: residual (forward) / skip gradient (backward)
  void* out;                      // bf16 NHWC [N,H,W,256]
  int stages;                     // halo ring depth (launcher-filled)
};
struct FusedBlockMaps {
  CUtensorMap a;      // [N,H,W,64] input of the 3x3, box (64, 10, 18, 1)
  CUtensorMap w1;     // packed 3x3 weights [(tap, n) rows][64], box (64, 64)
  CUtensorMap w2;     // packed pointwise weights [256 rows][64], box (64, 256)
  CUtensorMap res;    // residual / skip gradient [N,H,W,256], box (64, 8, 16, 1)
  CUtensorMap out;    // output [N,H,W,256], same box
};
int launch_fused_block(const FusedBlockMaps& maps, FusedBlockParams& p, cudaStream_t stream);

// cuTensorMapEncodeTiled through the runtime's driver entry point (no -lcuda needed).
// dims/strides innermost-first; strides in BYTES for dims 1..rank-1; bf16 elements; 128B swizzle.
int make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                   const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes = 128);

}  // namespace advb200

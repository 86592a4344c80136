// Non-GEMM layers of the ResNet-50 attack path (stem prep, pooling, head, losses, detector,
// weight packing).  T is the activation type: __nv_bfloat16 (fast mode) or float (validation).
#pragma once
#include "common.cuh"

namespace advb200 {

typedef __nv_bfloat16 bf16;

// ---- input preparation: fp32 NCHW image -> normalised conv1 operand ----------------------
// bf16 mode: "hcol" layout [N,115,112,64]: element (n, hp, wo, u*16+q), q=(ph*2+pw)*3+c (<12),
//            holds norm(x)[c, 2*(hp-2)+ph, 2*(wo-2+u)+pw] (zero outside the image / q>=12)
int launch_prep_hcol(const float* x, const float* mean3, const float* std3, bf16* out, int n,
                     cudaStream_t st);
// compact stem operand [N,115,115,16] (32-byte rows; conv1 takes all 16 taps as row shifts)
int launch_prep_hcol16(const float* x, const float* mean3, const float* std3, bf16* out, int n, cudaStream_t st);
// fp32 mode: NHWC [N,224,224,3]
int launch_prep_nhwc(const float* x, const float* mean3, const float* std3, float* out, int n, int h, int w,
                     cudaStream_t st);
// fp32 mode: image gradient NHWC [N,H,W,3] -> NCHW with the 1/std of the normalisation
int launch_grad_nhwc_to_nchw(const float* g, const float* std3, float* out, int n, int h, int w,
                             cudaStream_t st);

// ---- max-pool 3x3/2 pad 1 over post-ReLU activations ---------------------------------------
// idx: uint8 per output element = position (0..8) of the first maximum in the window scan,
//      255 when the maximum is <= 0 (ReLU would block the gradient anyway)
template <typename T>
int launch_maxpool_fwd(const T* in, T* out, uint8_t* idx, int n, int hi, int wi, int c, cudaStream_t st);
template <typename T>
int launch_maxpool_bwd(const T* gout, const uint8_t* idx, T* gin, int n, int hi, int wi, int c,
                       cudaStream_t st);

// ---- global average pool + its backward with the ReLU mask of the pooled activation -------
template <typename T>
int launch_avgpool_fwd(const T* in, float* feat, int n, int hw, int c, cudaStream_t st);
template <typename T>
int launch_avgpool_bwd(const float* dfeat, const T* act, T* gact, int n, int hw, int c, cudaStream_t st);

// ---- losses on fp32 logits [N,classes] --------------------------------------------------------
int launch_loss_grad(const float* logits, const int64_t* y, int n, int classes, int loss_kind, float c,
                     float kappa, float scale, float* dlogits, int64_t* preds, float* loss,
                     uint8_t* success, cudaStream_t st);
int launch_argmax(const float* logits, int n, int classes, int64_t* preds, cudaStream_t st);

// ---- detector score from the layer3 activation [N,hw,C] (defense_experiments.py:245-255) ------
template <typename T>
int launch_detector_score(const T* act, float* score, int n, int hw, int c, cudaStream_t st);

// ---- weight packing (BatchNorm folded) -------------------------------------------------------
// scale[k] = gamma/sqrt(var+eps), shift[k] = beta - mean*scale
int launch_bn_fold(const float* gamma, const float* beta, const float* mean, const float* var, float eps,
                   float* scale, float* shift, int c, cudaStream_t st);
// fp32 matrices for the validation path:
//   fwd  [(r*kw+s)*cin + ci, co] = w[co,ci,r,s]*scale[co]
//   dgrad[(r*kw+s)*cout + co, ci] = w[co,ci,r,s]*scale[co]
int launch_pack_f32(const float* w, const float* scale, float* fwd, float* dgrad, int cout, int cin, int k,
                    cudaStream_t st);
// bf16 K-major matrices for the tcgen05 path:
//   fwd  [(r*kw+s)*cout + co][ci]    dgrad [(r*kw+s)*cin + ci][co]
int launch_pack_bf16(const float* w, const float* scale, bf16* fwd, bf16* dgrad, int cout, int cin, int k,
                     cudaStream_t st);
// conv1 (7x7/2, 3->64) in its space-to-depth form:
//   fwd  [t*64 + co][u*16+q]      dgrad [t*64 + slot(q)*4 + u][co]    (t,u in 0..3; see launch_prep_hcol)
int launch_pack_conv1_bf16(const float* w, const float* scale, bf16* fwd, bf16* dgrad, bf16* fwd16, cudaStream_t st);
// fc: transpose [classes, feat] -> [feat, classes]
int launch_transpose_f32(const float* in, float* out, int rows, int cols, cudaStream_t st);

// out = a + b (fused folded-BN shifts)
int launch_vec_add(const float* a, const float* b, float* out, int n, cudaStream_t st);

// test hook: bf16 [pixels, c] (c % 64 == 0) -> ReLU bit mask in the layout the conv epilogue reads
int launch_pack_mask_bits(const bf16* act, uint32_t* bits, long long pixels, int c, cudaStream_t st);

// dtype converters for the test hooks
int launch_f32_to_bf16(const float* in, bf16* out, int64_t n, cudaStream_t st);

// ---- forward-only transfer target (VGG-19) helpers
int launch_prep_nhwc64_bf16(const float* x, const float* mean3, const float* std3, bf16* out, int n, int h, int w,
                            cudaStream_t st);
int launch_maxpool2x2_bf16(const bf16* in, bf16* out, int n, int hi, int wi, int c, cudaStream_t st);
int launch_pack_fc_bf16(const float* w, bf16* out, int cout, int cout_pad, int c, int hw, cudaStream_t st);
int launch_bf16_cols_to_f32(const bf16* in, int ld_in, float* out, int rows, int cols, cudaStream_t st);
int launch_pad_vec(const float* in, float* out, int n, int n_pad, cudaStream_t st);
int launch_pad_cin(const float* w, float* out, int cout, int cin, int cin_pad, int k, cudaStream_t st);

}  // namespace advb200

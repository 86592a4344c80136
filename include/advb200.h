/*
 * advb200.h — C ABI of libadvb200.so, the B200 (sm_100a) attack hot path.
 *
 * The reference (picklerick422/Image-recognition-adversarial-example-attack) has no FFI layer:
 * its boundary is the Python call surface of attack.py / defense_experiments.py.  Every entry
 * point below names the reference lines whose arithmetic it replaces; the Python mirror in
 * image-recognition-adversarial-example-attack_b200/advb200/ binds them with ctypes (see
 * INTEGRATION.md for the stub a reference maintainer would add).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in _host;
 *   - images are fp32 NCHW [N,3,224,224] in [0,1], contiguous (attack.py:20-28 contract);
 *   - labels are int64 [N];
 *   - `stream` is a cudaStream_t passed as void*; all work is stream-ordered, nothing
 *     synchronises internally, nothing allocates on the hot path (workspaces are caller-owned);
 *   - return 0 = OK, negative = error, text via advb200_last_error() (thread-local);
 *   - a model handle is bound to the device that was current at create time; not thread-safe.
 */
#ifndef ADVB200_H_
#define ADVB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ADVB200_OK 0
#define ADVB200_ERR_ARG (-1)
#define ADVB200_ERR_CUDA (-2)
#define ADVB200_ERR_UNSUPPORTED (-3)

/* precision of the conv path */
#define ADVB200_PREC_BF16 0 /* tcgen05/TMEM implicit GEMM, bf16 operands, fp32 accumulate   */
#define ADVB200_PREC_FP32 1 /* fp32 validation mode (CUDA-core implicit GEMM, no TF32)      */

/* loss kinds for advb200_input_grad */
#define ADVB200_LOSS_CE_MEAN 0  /* F.cross_entropy(logits, y) mean-reduced: attack.py:31,58        */
#define ADVB200_LOSS_CW_UNTGT 1 /* c*relu(real-other+kappa), summed: attack.py:114-127 (untargeted) */
#define ADVB200_LOSS_CW_TGT 2   /* c*relu(other-real+kappa), summed: attack.py:119-121 (targeted)   */

typedef struct advb200_model advb200_model;

const char* advb200_last_error(void);
int advb200_version(void);

/* ------------------------------------------------------------------------------------------
 * Model (torchvision ResNet-50 v1.5, eval mode) — replaces `model(normalize_batch(x))` at
 * attack.py:30,57,111 and the autograd input-gradient at attack.py:32,59,130.
 *
 * conv_w[i]  : fp32 [Cout,Cin,kh,kw] (torchvision layout), i in the canonical order
 *              conv1, then per block of layer1..layer4: conv1, conv2, conv3, [downsample.0]
 *              (53 entries).  bn_*[i] are the matching BatchNorm vectors (fp32 [Cout]);
 *              bn_eps[i] the matching eps (host array).  fc_w fp32 [1000,2048], fc_b [1000].
 * BatchNorm is folded (eval: running stats) into the packed conv weights at create time.
 * ------------------------------------------------------------------------------------------ */
#define ADVB200_NUM_CONVS 53
int advb200_model_create(const float* const* conv_w, const float* const* bn_gamma,
                         const float* const* bn_beta, const float* const* bn_mean,
                         const float* const* bn_var, const float* bn_eps_host, const float* fc_w,
                         const float* fc_b, int precision, void* stream, advb200_model** out);
int advb200_model_destroy(advb200_model* m);
int advb200_model_precision(const advb200_model* m);

/* bytes of caller-owned workspace needed for a batch of n images (forward + input-grad) */
int64_t advb200_workspace_bytes(const advb200_model* m, int n);

/* logits = model((x-mean)/std).  mean3/std3: fp32 [3] device (normalize_batch, attack.py:10-17).
 * If score_out != NULL also writes the layer3 detector score per image
 * (defense_experiments.py:221-257).  If stop_at_layer3 != 0 the forward ends after layer3 and
 * logits_out may be NULL.  preds_out (int64 [N], may be NULL) = argmax(logits) (first max). */
int advb200_forward(advb200_model* m, const float* x, const float* mean3, const float* std3, int n,
                    float* logits_out, int64_t* preds_out, int stop_at_layer3, float* score_out,
                    void* workspace, int64_t workspace_bytes, void* stream);

/* One forward + input-only backward (no weight gradients; weights are frozen):
 * grad_x = d loss / d x, fp32 NCHW, through the normalisation (includes the 1/std factor).
 *   CE_MEAN : loss = mean_n CE(logits_n, y_n); scale = inv_n_global (1/N of the whole batch so a
 *             sharded batch reproduces the unsharded gradient, attack.py:31)
 *   CW_*    : loss = sum_n c*f_n (attack.py:119-127; the L2 term is added by advb200_cw_step_f32)
 * Optional outputs (NULL to skip): logits_out fp32 [N,1000]; preds_out int64 [N];
 * loss_out fp32 [N] (per-image CE, or f_n for CW); success_out uint8 [N] (CW success flag,
 * attack.py:121,124; for CE: pred != y). y_cmp is y_true (CE, CW untargeted) or y_target. */
int advb200_input_grad(advb200_model* m, const float* x, const float* mean3, const float* std3,
                       const int64_t* y_cmp, int n, int loss_kind, float c, float kappa,
                       float inv_n_global, float* logits_out, int64_t* preds_out, float* loss_out,
                       uint8_t* success_out, float* grad_x_out, void* workspace,
                       int64_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * Update kernels (HBM-bound, 128-bit vectorised, bit-exact against the reference arithmetic)
 * ------------------------------------------------------------------------------------------ */

/* FGSM / PGD step, in place on x_adv (attack.py:33-34 when project==0, attack.py:61-63 when
 * project!=0):  x1 = x_adv + alpha*sign(g);  [x2 = max(min(x1, x0+eps), x0-eps)];  clamp(.,0,1) */
int advb200_linf_step_f32(float* x_adv, const float* x_orig, const float* grad, float alpha,
                          float eps, int64_t n_elems, int project, void* stream);

/* PGD random start (attack.py:50-51): x_adv = clamp(x_orig + noise, 0, 1); noise is supplied by
 * the caller (torch's generator) so both sides of a parity test draw the same tensor. */
int advb200_random_start_f32(float* x_adv, const float* x_orig, const float* noise, int64_t n_elems,
                             void* stream);

/* CW-L2 (attack.py:95-105): x0 = clamp(x,0,1); w = atanh((x0*(1-2e-6)+1e-6)*2-1); m = v = 0;
 * x_adv = 0.5*(tanh(w)+1); best_adv = x0; best_l2 = +inf; best_success = 0;
 * l2_partials[N,parts] = per-block partial sums of (x_adv-x0)^2 (fixed-order, deterministic). */
int advb200_cw_parts(int64_t elems_per_img); /* number of partial sums per image */
int advb200_cw_init_f32(const float* x, float* x0, float* w, float* m, float* v, float* x_adv,
                        float* best_adv, float* best_l2, uint8_t* best_success, float* l2_partials,
                        int n_img, int64_t elems_per_img, void* stream);

/* Input gradient for a caller-supplied dlogits [N,1000] fp32 (the backward half of advb200_input_grad, for callers
 * that compute their own loss on the logits: visualized_attacks.py:337-358, defense_experiments.py:458-479 run
 * `autograd.grad(loss, x)` through `model(normalize_batch(x))` themselves).  Uses the ReLU bit masks, pool indices
 * and layer4 output that the immediately preceding advb200_forward(n, same workspace, full depth) left in the
 * workspace; ADVB200_ERR_ARG if there is no such state.  One backward per forward (the gradient buffers may reuse
 * activation space).  grad_x_out [N,3,224,224] fp32 = d(sum(dlogits*logits))/dx, already divided by std3. */
int advb200_backward(advb200_model* m, const float* dlogits, const float* std3, int n, float* grad_x_out,
                     void* workspace, int64_t workspace_bytes, void* stream);

/* One fused attack step (bf16 mode): forward + CE input gradient as advb200_input_grad(CE_MEAN), with the sign
 * step, eps-ball projection and [0,1] clamp of attack.py:61-63 (x_orig != NULL) or attack.py:33-34
 * (x_orig == NULL: FGSM, no projection) applied to x_adv IN PLACE by the epilogue of the stem's input-gradient
 * kernel — the gradient tensor is never written to memory.  Bit-identical to advb200_input_grad followed by
 * advb200_linf_step_f32.  Returns ADVB200_ERR_UNSUPPORTED for an fp32-mode model (use the two calls). */
int advb200_attack_step(advb200_model* m, float* x_adv_inout, const float* x_orig, const float* mean3,
                        const float* std3, const int64_t* y_true, int n, float inv_n_global, float alpha,
                        float eps, int64_t* preds_out, void* workspace, int64_t workspace_bytes, void* stream);

/* The PGD loop body, `steps` times (attack.py:55-63):  x_adv <- clamp(project(x_adv + alpha*sign(grad CE(x_adv)))),
 * in place; x_orig == NULL drops the projection (attack.py:33-34).  grad_scratch: fp32 [N,3,224,224] caller-owned.
 * use_graph != 0: the ~105 kernel launches of one iteration are captured ONCE per argument set (pointers, n, scalars)
 * into a CUDA graph and replayed per step — one host call per iteration; same kernels in the same order, bit-identical
 * to advb200_input_grad + advb200_linf_step_f32.  Captured graphs are dropped when the model is re-planned for
 * another batch size / workspace. */
int advb200_pgd_steps(advb200_model* m, float* x_adv_inout, const float* x_orig, const float* mean3, const float* std3,
                      const int64_t* y_true, int n, float inv_n_global, float alpha, float eps, int steps,
                      float* grad_scratch, void* workspace, int64_t workspace_bytes, int use_graph, void* stream);

/* One CW iteration after advb200_input_grad(CW_*) produced grad_x and success for the current
 * x_adv (attack.py:126-137 + torch.optim.Adam single-tensor update):
 *   l2_n = sum(l2_partials_in[n,:])                     (distance of the CURRENT x_adv)
 *   improved_n = success_n && l2_n < best_l2_n  -> best_l2, best_success, best_adv updated
 *   g_w = (2*(x_adv-x0) + grad_x) * 0.5 * (1 - tanh(w)^2);  Adam(m, v, w)
 *   x_adv <- 0.5*(tanh(w)+1) for the NEXT iteration, l2_partials_out <- its distance partials.
 * step_size = lr/(1-beta1^t) and bc2_sqrt = sqrt(1-beta2^t) are computed in double by the host; beta1, beta2 and
 * adam_eps are doubles because torch forms 1-beta1 and 1-beta2 in double and only then rounds them to fp32
 * (1.0f - 0.999f is 1.3e-5 away from fp32(1 - 0.999): enough to miss the reference's Adam state). */
int advb200_cw_step_f32(float* w, float* m, float* v, const float* x0, const float* grad_x,
                        float* x_adv, float* best_adv, float* best_l2, uint8_t* best_success,
                        const uint8_t* success, const float* l2_partials_in, float* l2_partials_out,
                        float* l2_out, int n_img, int64_t elems_per_img, float step_size,
                        float bc2_sqrt, double beta1, double beta2, double adam_eps, void* stream);

/* attack.py:139: out = best_success ? best_adv : x_adv (x_adv already holds 0.5*(tanh(w)+1)
 * of the post-update w). */
int advb200_cw_finalize_f32(const float* x_adv, const float* best_adv, const uint8_t* best_success,
                            float* out, int n_img, int64_t elems_per_img, void* stream);

/* ------------------------------------------------------------------------------------------
 * Defenses (defense_experiments.py:177-217) — fp32 NCHW planes of h x w
 * ------------------------------------------------------------------------------------------ */
/* 3x3 mean, stride 1, zero pad 1, divisor 9 (F.avg_pool2d(x,3,1,1)); bit-exact vs torch CPU */
int advb200_smooth3x3_f32(const float* x, float* out, int64_t planes, int h, int w, void* stream);
/* 3x3 median, stride 1, zero pad 1: the median-smoothing defense named in the build's brief; same window and padding
 * as the reference's mean smoothing (defense_experiments.py:177-179), selection instead of average.  Exact (a
 * selection), NaN-free inputs assumed.  x != out. */
int advb200_median3x3_f32(const float* x, float* out, int64_t planes, int h, int w, void* stream);
/* round(clamp(x,0,1)*(levels-1))/(levels-1), round-half-even, true division; bit-exact */
int advb200_quantize_f32(const float* x, float* out, int64_t n_elems, int levels, void* stream);
/* defend_input without JPEG: clamp -> smooth -> quantize(16) -> clamp, one pass */
int advb200_defend_input_f32(const float* x, float* out, int64_t planes, int h, int w, int levels,
                             void* stream);

/* Test hook: compares the constant-divisor shortcut the smoothing kernels use for s/9 with the IEEE division, for all
 * 2^32 float bit patterns, on the device; *mismatches must come back 0. */
int advb200_test_div9(uint64_t* mismatches, uint32_t* first_bad_bits);

/* JPEG defense (defense_experiments.py:188-207 _jpeg_compress_batch: ToPILImage -> PIL save(JPEG, quality) -> open ->
 * ToTensor).  The baseline-JPEG sample pipeline of PIL's libjpeg (4:2:0, Annex-K tables scaled for `quality`, islow
 * integer DCTs, fancy chroma upsampling) computed on the device, bit-exact with the host round trip; any h, w >= 1
 * (edges replicated as the encoder does).  x, out: fp32 [n_img,3,h,w] (may alias); scratch: caller-owned device
 * buffer of advb200_jpeg_scratch_bytes(n_img,h,w) bytes (decoded Y + half-size Cb, Cr as uint8). */
int64_t advb200_jpeg_scratch_bytes(int n_img, int h, int w);
int advb200_jpeg_roundtrip_f32(const float* x, float* out, int n_img, int h, int w, int quality, void* scratch,
                               int64_t scratch_bytes, void* stream);

/* Allocation-free variant of advb200_jpeg_roundtrip_f32.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same kernels, same results bit for bit; the uint8 scratch between the two kernels
 * is a static device array of ADVB200_JPEG_NOALLOC_MAX_BYTES inside the library (every GPU that loads the library
 * holds it until the process exits) instead of a caller buffer.  The call allocates nothing (no cudaMalloc / cudaFree /
 * allocating library routine); the first call on each device does one cudaGetSymbolAddress lookup, so make that first
 * call outside a graph capture.  Sizes with advb200_jpeg_scratch_bytes(n_img,h,w) > ADVB200_JPEG_NOALLOC_MAX_BYTES
 * (24 MiB: 256 images of 224 x 224, or one image of about 4000 x 4000) return ADVB200_ERR_ARG with the figures in
 * advb200_last_error().  All calls on one device SHARE the array, which keeps the previous call's bytes (nothing in it
 * needs to start at zero: every byte read is first written in the same call): two calls that can overlap on the GPU
 * (different streams without an ordering between them) race on it -- order them, or use the caller-buffer entry. */
#define ADVB200_JPEG_NOALLOC_MAX_BYTES (24 * 1024 * 1024)
int advb200_jpeg_roundtrip_f32_noalloc(const float* x, float* out, int n_img, int h, int w, int quality, void* stream);

/* Image preprocessing of `load_image` (ResNet.py:27-31, defense_experiments.py:649-662) for one decoded image:
 * Resize(size) [shorter side -> size; PIL's antialiased bilinear ImagingResample in 22-bit fixed point on uint8] ->
 * CenterCrop(crop) -> ToTensor (/255).  img_hwc: uint8 [h,w,3] on the device; out_chw: fp32 [3,crop,crop].
 * Bit-exact with torchvision + PIL.  scratch: advb200_preprocess_scratch_bytes(...) bytes, 256-byte aligned.
 * ADVB200_ERR_UNSUPPORTED when the resized image is smaller than the crop (the reference would zero-pad). */
int64_t advb200_preprocess_scratch_bytes(int h, int w, int size, int crop);
int advb200_preprocess_u8(const uint8_t* img_hwc, int h, int w, int size, int crop, float* out_chw, void* scratch,
                          int64_t scratch_bytes, void* stream);
/* ToPILImage of --save_adv (ResNet.py:121-126): fp32 [3,h,w] in [0,1] -> uint8 [h,w,3], x*255 truncated. */
int advb200_to_u8_hwc(const float* x_chw, uint8_t* out_hwc, int h, int w, void* stream);

/* ------------------------------------------------------------------------------------------
 * Forward-only transfer target (blackbox_transfer.py:25-54, 233-251): torchvision VGG-19, eval mode.
 * conv_w/conv_b: the 16 conv weights [cout,cin,3,3] / biases in feature order; fc_w/fc_b: classifier.0/.3/.6
 * ([4096,25088], [4096,4096], [1000,4096]).  bf16 operands, fp32 accumulation; logits come back as fp32 [N,1000]
 * (rounded to bf16 once, by the last layer).  workspace: advb200_vgg19_workspace_bytes(n), 1024-byte aligned.
 * ------------------------------------------------------------------------------------------ */
typedef struct advb200_vgg19 advb200_vgg19;
int advb200_vgg19_create(const float* const* conv_w, const float* const* conv_b, const float* const* fc_w,
                         const float* const* fc_b, void* stream, advb200_vgg19** out);
int advb200_vgg19_destroy(advb200_vgg19* m);
int64_t advb200_vgg19_workspace_bytes(const advb200_vgg19* m, int n);
int advb200_vgg19_forward(advb200_vgg19* m, const float* x, const float* mean3, const float* std3, int n,
                          float* logits_out, int64_t* preds_out, void* workspace, int64_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * Evaluation counters (defense_experiments.py:597-637, 809-832)
 * counters int64 [7] += {clean_correct, attack_success, defense_preproc_success,
 *                        detector_flags_clean, detector_flags_adv, detector_attack_success, count}
 * flags_out uint8 [N,6] (optional) receives the per-image booleans in the same order.
 * ------------------------------------------------------------------------------------------ */
int advb200_eval_counters(const int64_t* pred_clean, const int64_t* pred_adv, const int64_t* pred_def,
                          const float* score_clean, const float* score_adv, float threshold,
                          const int64_t* y_true, int n, int64_t* counters, uint8_t* flags_out,
                          void* stream);

/* ------------------------------------------------------------------------------------------
 * Kernel-level test hooks (used by tests/ and bench.py only)
 * ------------------------------------------------------------------------------------------ */
/* NHWC convolution through the same kernels the model uses.
 *   x     : [N,H,W,Cin]  bf16 (PREC_BF16) or fp32 (PREC_FP32)
 *   wgt   : fp32 [Cout,Cin,kh,kw];  bias fp32 [Cout] or NULL
 *   add   : optional residual, same shape/dtype as out;  relu: apply max(.,0)
 *   dgrad : 0 = forward conv (stride, pad); 1 = data gradient of that conv (x is dY
 *           [N,Ho,Wo,Cout], out is dX [N,H,W,Cin]); mask (optional, dX-shaped) zeroes out
 *           where mask<=0.
 * Supported shapes are the ResNet-50 ones (1x1 s1/s2, 3x3 s1/s2 pad 1; Cin,Cout % 64 == 0). */
int advb200_test_conv(int precision, const void* x, const float* wgt, const float* bias,
                      const void* add, const void* mask, void* out, int n, int h, int w, int cin,
                      int cout, int ksize, int stride, int relu, int dgrad, void* stream);
/* Test hook: forces the tile configuration (BN columns, CTA pair, resident weights, output staging slots; -1 =
 * heuristic) of the following advb200_test_conv calls — the candidates the plan-time autotuner chooses between;
 * the parity tests check that every candidate gives bit-identical results. */
int advb200_test_force_tile(int bn, int pair, int bres, int slots);

/* ------------------------------------------------------------------------------------------
 * Building blocks of the transformer transfer targets — ViT-B/16 and Swin-T, forward only
 * (blackbox_transfer.py:36-54 load_vit / load_swin; :233-251 predict_model on the adversarial examples).
 * The Python mirror (advb200/transformers.py) composes them exactly as torchvision composes its modules.
 * Token matrices are [tokens, channels] row-major; bf16 unless the name says f32; channel counts seen by the GEMM kernel
 * are padded to multiples of 64 (zero weights, zero bias, padded activations exactly zero).
 * ------------------------------------------------------------------------------------------ */
typedef struct advb200_linear advb200_linear;
/* nn.Linear(cin, cout) packed for the tcgen05 kernel: w fp32 [cout, cin], bias fp32 [cout] or NULL */
int advb200_linear_create(const float* w, const float* bias, int cin, int cout, int cin_pad, int cout_pad, void* stream,
                          advb200_linear** out);
int advb200_linear_destroy(advb200_linear* l);
/* y[m, cout_pad] = x[m, cin_pad] W^T + b (+ add[m, cout_pad]), optional ReLU; bf16 in / out, fp32 accumulation */
int advb200_linear_forward(advb200_linear* l, const void* x, void* y, int64_t m, int relu, const void* add, void* stream);
/* fp32 classifier head (model.heads / model.head): logits = feat W^T + b; scratch of advb200_fc_scratch_bytes(n, classes) bytes */
int64_t advb200_fc_scratch_bytes(int n, int classes);
int advb200_fc_f32(const float* feat, const float* w, const float* bias, float* logits, float* scratch, int n, int classes, int k,
                   void* stream);
/* conv_proj / patch embedding operand: normalised image (attack.py:17) -> bf16 [N*(h/patch)*(w/patch), kpad], column
 * (c*patch + i)*patch + j = the flatten order of the Conv2d weight; columns >= 3*patch*patch are zero */
int advb200_patch_im2col_bf16(const float* x, const float* mean3, const float* std3, void* out, int n, int h, int w, int patch,
                              int kpad, void* stream);
/* ViT: x32[n,0,:] = class_token + pos[0]; x32[n,1+p,:] = patch[n,p,:] + pos[1+p]   (fp32 residual stream) */
int advb200_vit_tokens_f32(const void* patch_bf16, const float* cls, const float* pos, float* x32, int n, int tokens, int c,
                           void* stream);
/* residual stream update fused with LayerNorm: mode 0: x32 unchanged, 1: x32 += delta, 2: x32 = delta (delta bf16
 * [m, c_pad]); y (bf16 [m, c_pad], may be NULL) = LayerNorm over the c_real real channels (<= 1536), pad = 0 */
int advb200_add_layernorm(float* x32, const void* delta_bf16, const float* gamma, const float* beta, float eps, void* y_bf16,
                          int64_t m, int c_real, int c_pad, int mode, void* stream);
/* exact (erf) GELU, bf16 -> bf16, n % 8 == 0 */
int advb200_gelu_bf16(const void* x, void* y, int64_t n, void* stream);
/* multi-head attention, one window per CTA and head: softmax(q k^T * scale + bias + region mask) v.
 * grid_h == 0: the window is the image's whole sequence of tokens_per_window tokens (ViT).  Otherwise window x window
 * windows of a grid_h x grid_w token grid cyclically shifted by `shift` (Swin shifted-window attention, with the
 * region mask of torchvision's implementation when shift > 0); bias fp32 [heads, T, T] (relative-position bias) or NULL.
 * q / k / v are column blocks of `qkv` at q_off / k_off / v_off (+ head * dim); out[row, head*dim + d]. */
int advb200_window_attention_bf16(const void* qkv, void* out, const float* bias, int n_img, int tokens_per_window, int heads, int dim,
                                  int ld_qkv, int ld_out, int q_off, int k_off, int v_off, int grid_h, int grid_w, int window,
                                  int shift, float scale, void* stream);
/* test hook: on != 0 routes advb200_window_attention_bf16 through its CUDA-core kernel instead of the tensor-core one */
int advb200_test_attention_simt(int on);
/* Swin PatchMerging input: concat of the four 2x2 neighbours (x0, x1, x2, x3 of torchvision) of the fp32 stream */
int advb200_patch_merge_gather_f32(const float* x32, float* out, int n, int grid_h, int grid_w, int c_real, int c_pad, void* stream);
/* head features: mean over `count` tokens from t0 of LayerNorm(x32[n, t, :]) -> fp32 [n, c_real] (ViT: the class token) */
int advb200_ln_pool_f32(const float* x32, const float* gamma, const float* beta, float eps, float* feat, int n, int tokens, int t0,
                        int count, int c_real, int c_pad, void* stream);

/* Test hook: in-situ trace of a forward / input-gradient pass.  While a device buffer is set, every tensor the passes
 * write — each conv launch's output as the plan really runs it (fused launches, multi-class stride-2 input gradients),
 * the pooled tensors, fc / loss tensors — is copied into the buffer right after the launch that produced it (names as in
 * advb200_profile_dump: "fwd l2.0.c3+ds", "bwd l1.1.c2+c1", "p1", "g_top", ...).  tests/test_gpu_trace.py checks every
 * launch against the oracle applied to that launch's OWN traced inputs (model(...) attack.py:30; autograd.grad
 * attack.py:32).  advb200_debug_trace(NULL, 0) switches tracing off; advb200_debug_trace_index writes one line per
 * tensor, "name|byte offset|n|h|w|c|element size" (activations NHWC), and returns the bytes needed. */
int advb200_debug_trace(void* buf, int64_t cap_bytes);
int64_t advb200_debug_trace_index(char* buf_host, int64_t cap);

/* number of kernels launched by this library since load (bench.py's gpu_launches) */
int64_t advb200_launch_count(void);

/* Per-launch timing of the conv kernels for bench.py's roofline leg: while enabled, every tcgen05
 * conv launch is bracketed by CUDA events on its stream; advb200_profile_read synchronises on them,
 * returns the summed device time (ms) and the number of launches, and clears the list. */
int advb200_profile_enable(int on);
int advb200_profile_read(double* conv_ms_host, int64_t* conv_launches_host);
/* Same list as CSV text, one line per launch: name,bn,tiles,ksteps,stages,ms,flops,bytes (algorithmic
 * flops / HBM bytes of that launch).  Writes at most cap bytes (NUL-terminated) into buf_host, returns
 * the size needed, and clears the list. */
int64_t advb200_profile_dump(char* buf_host, int64_t cap);
/* Plan-time autotuner report: number of conv launches whose tile configuration was chosen by timing
 * candidates on the device (at the first call for a batch size), and the host wall time that took.
 * ADVB200_NO_AUTOTUNE=1 in the environment disables tuning (heuristic tile choice). */
int advb200_tune_report(const advb200_model* m, int* tuned_launches_host, double* tune_ms_host);

#ifdef __cplusplus
}
#endif
#endif /* ADVB200_H_ */

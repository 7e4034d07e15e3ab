/*
 * imx.h — C ABI of libimx_kernels.so: the B200 (sm_100a) implementation of the
 * SPADE generator hot path and the vid2vid flow warp of StephChabr/imaginaire.
 *
 * This is the drop-in boundary.  Every entry point
 *   - takes plain device pointers, sizes and a cudaStream_t (passed as void*),
 *   - launches asynchronously on that stream, never synchronises, never
 *     allocates device memory, never touches the default stream,
 *   - returns 0 on success, otherwise a cudaError_t value or an IMX_ERR_* code;
 *     imx_last_error() returns the calling thread's last message.
 * Activations are NHWC bf16 ("channels_last" memory of a logical NCHW tensor),
 * statistics / biases / norm affine parameters are fp32, warp tensors are NCHW
 * fp32 exactly as the reference's extension takes them.
 *
 * "ref:" lines cite the reference interface (path:line under the reference
 * repository root) that each entry point replaces.
 */
#ifndef IMX_H_
#define IMX_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define IMX_API __attribute__((visibility("default")))
#else
#define IMX_API
#endif

typedef void* imx_stream_t;   /* a cudaStream_t */
typedef uint16_t imx_bf16;    /* raw bfloat16 bits */

IMX_API int imx_version(void);
IMX_API const char* imx_last_error(void);

/* ------------------------------------------------------------------ warp -- */

/* out[b,c,y,x] = bilinear sample of img[b,c] at (x+flow[b,0,y,x], y+flow[b,1,y,x]),
 * tap indices clamped to the OUTPUT dims (H, W), weights not renormalised.
 * img: B x C x Hi x Wi, flow: B x 2 x H x W, out: B x C x H x W; all contiguous fp32.
 * kernel_size must be 1.
 * ref: resample2d_cuda.forward(input1, input2, output, kernel_size)
 *      imaginaire/third_party/resample2d/src/resample2d_cuda.cc:6-17,
 *      kernel imaginaire/third_party/resample2d/src/resample2d_kernel.cu:15-75 */
IMX_API int imx_resample2d_fwd_f32(const float* img, const float* flow, float* out,
                           int B, int C, int H, int W, int Hi, int Wi,
                           int kernel_size, imx_stream_t stream);

/* Gradients of the above.  grad_img (B x C x Hi x Wi) MUST be zero-filled by the
 * caller (scatter with atomics, like the reference); grad_flow (B x 2 x H x W) is
 * overwritten.  trunc_alpha != 0 reproduces the reference's use of
 * `xf - int(xf)` (truncation) for the scatter weights; 0 uses floor, which is the
 * true derivative of the forward for negative coordinates.
 * ref: resample2d_cuda.backward(...)
 *      imaginaire/third_party/resample2d/src/resample2d_cuda.cc:19-33,
 *      kernels resample2d_kernel.cu:78-129 and :131-203 */
IMX_API int imx_resample2d_bwd_f32(const float* img, const float* flow, const float* grad_out,
                           float* grad_img, float* grad_flow,
                           int B, int C, int H, int W, int Hi, int Wi,
                           int kernel_size, int trunc_alpha, imx_stream_t stream);

/* The warp the vid2vid generators actually call: grid_sample(bilinear, border,
 * align_corners=True) of img at linspace(-1,1)+flow/((size-1)/2), computed with
 * the same fp32 operation order as torch (bit-exact against the CPU path) and
 * without building the grid.  img, out: B x C x H x W fp32; flow: B x 2 x H x W fp32.
 * ref: resample(image, flow) imaginaire/model_utils/fs_vid2vid.py:14-39,
 *      get_grid :42-80 */
IMX_API int imx_gridsample_warp_fwd_f32(const float* img, const float* flow, float* out,
                                int B, int C, int H, int W, imx_stream_t stream);

/* Gradients of imx_gridsample_warp_fwd_f32 (autograd of the reference's
 * resample()).  grad_img must be zero-filled by the caller; grad_flow is
 * overwritten; either may be NULL to skip it. */
IMX_API int imx_gridsample_warp_bwd_f32(const float* img, const float* flow, const float* grad_out,
                                float* grad_img, float* grad_flow,
                                int B, int C, int H, int W, imx_stream_t stream);

/* ------------------------------------------------------- layout helpers -- */

/* Nearest-neighbour resize of an NCHW fp32 conditioning map to NHWC bf16 with
 * the channel dimension zero-padded to Cp (Cp >= L, Cp % 8 == 0):
 * dst[n,y,x,c] = src[n,c,floor(y*H/h),floor(x*W/w)].
 * ref: F.interpolate(cond, size, mode='nearest')
 *      imaginaire/layers/activation_norm.py:224-225 and
 *      imaginaire/generators/spade.py:424 */
IMX_API int imx_nearest_resize_nchw_f32_to_nhwc_bf16(const float* src, imx_bf16* dst,
                                             int N, int L, int H, int W,
                                             int h, int w, int Cp, imx_stream_t stream);

/* Nearest x2 upsample, NHWC bf16.  ref: NearestUpsample(scale_factor=2)
 * imaginaire/generators/spade.py:394-395 */
IMX_API int imx_upsample2x_nhwc_bf16(const imx_bf16* src, imx_bf16* dst,
                             int N, int H, int W, int C, imx_stream_t stream);
/* Its gradient: dst[n,y,x,c] = sum of the 2x2 block of src (N x 2H x 2W x C). */
IMX_API int imx_upsample2x_bwd_nhwc_bf16(const imx_bf16* src, imx_bf16* dst,
                                 int N, int H, int W, int C, imx_stream_t stream);

/* out = bilinear_upsample_x2(src) (+ lateral), align_corners = False, NHWC bf16: src N x h x w x C,
 * lateral (may be NULL) and out N x 2h x 2w x C.  The top-down step of a feature pyramid in one
 * pass (fp32 interpolation and addition, one rounding).  imx_bilinear_up2x_bwd_nhwc_bf16 is the
 * gradient with respect to src (the gradient with respect to lateral is grad_out itself).
 * ref: FPSE discriminator, imaginaire/discriminators/fpse.py:119-122
 *      (nn.Upsample(scale_factor=2, mode='bilinear') followed by the lateral addition) */
IMX_API int imx_bilinear_up2x_add_nhwc_bf16(const imx_bf16* src, const imx_bf16* lateral, imx_bf16* out,
                                    int N, int h, int w, int C, imx_stream_t stream);
IMX_API int imx_bilinear_up2x_bwd_nhwc_bf16(const imx_bf16* grad_out, imx_bf16* grad_src,
                                    int N, int h, int w, int C, imx_stream_t stream);

/* Bilinear resize with align_corners = True of `planes` (= N * C) contiguous H x W fp32 planes to
 * Ho x Wo (ATen's upsample_bilinear2d arithmetic: source index = dst * (in - 1) / (out - 1)).
 * ref: the 0.5x input pyramid of the multi-resolution patch discriminator,
 *      imaginaire/discriminators/spade.py:86-88 (multires_patch.py:169-171). */
IMX_API int imx_bilinear_resize_ac_nchw_f32(const float* src, float* dst, int64_t planes, int H, int W, int Ho, int Wo,
                                    imx_stream_t stream);

/* 2x2 / stride-2 max pooling, NHWC bf16 (C % 8 == 0), floor mode: src N x H x W x C ->
 * dst N x H/2 x W/2 x C; NaN propagates.  The backward recomputes the argmax from src (no index
 * tensor): the first of (0,0), (0,1), (1,0), (1,1) equal to the window maximum takes grad_out;
 * only the pooled 2*(H/2) x 2*(W/2) region of grad_src is written (an odd last row / column must
 * be zeroed by the caller).
 * ref: nn.MaxPool2d(2, 2) of the perceptual loss' VGG-19, imaginaire/losses/perceptual.py:175-192 (torchvision vgg19 .features: MaxPool2d(2, 2) at indices 4, 9, 18, 27) */
IMX_API int imx_maxpool2x2_nhwc_bf16(const imx_bf16* src, imx_bf16* dst, int N, int H, int W, int C, imx_stream_t stream);
IMX_API int imx_maxpool2x2_bwd_nhwc_bf16(const imx_bf16* src, const imx_bf16* grad_out, imx_bf16* grad_src,
                                 int N, int H, int W, int C, imx_stream_t stream);

/* Mean-absolute-difference loss on bf16 feature maps (flat arrays of n elements; a, b and grad
 * dense, one shared memory layout, 16-byte aligned).
 *   imx_l1_abs_sum_bf16 :  *sum += sum_i |float(a[i]) - float(b[i])|   (fp32; the caller zeroes *sum
 *                          and divides by n for the mean)
 *   imx_l1_abs_grad_bf16:  grad[i] = bf16(sign(a[i] - b[i]) * ((*upstream) / denom)), sign(0) = 0;
 *                          upstream is a DEVICE scalar (the gradient of the loss value), denom a
 *                          host value (n for the mean).  The gradient w.r.t. b is -grad.
 * ref: nn.L1Loss on feature maps in imaginaire/losses/perceptual.py:66-67, :130-131 and
 *      imaginaire/losses/feature_matching.py:12, :36-38. */
IMX_API int imx_l1_abs_sum_bf16(const imx_bf16* a, const imx_bf16* b, int64_t n, float* sum, imx_stream_t stream);
IMX_API int imx_l1_abs_grad_bf16(const imx_bf16* a, const imx_bf16* b, imx_bf16* grad, int64_t n,
                         const float* upstream, float denom, imx_stream_t stream);

/* Per-pixel dot product over the channels of two NHWC bf16 tensors (pixels x C, C % 8 == 0):
 * out[p] = bf16(sum_c a[p][c] * b[p][c]) with fp32 accumulation, and its gradients
 * grad_a[p][c] = grad_out[p] * b[p][c], grad_b[p][c] = grad_out[p] * a[p][c] (either may be NULL).
 * ref: the projection term of the FPSE discriminator, torch.mul(segemb, seg).sum(dim=1, keepdim=True),
 *      imaginaire/discriminators/fpse.py:125-127. */
IMX_API int imx_channel_dot_nhwc_bf16(const imx_bf16* a, const imx_bf16* b, imx_bf16* out, int64_t pixels, int C,
                              imx_stream_t stream);
IMX_API int imx_channel_dot_bwd_nhwc_bf16(const imx_bf16* a, const imx_bf16* b, const imx_bf16* grad_out,
                                  imx_bf16* grad_a, imx_bf16* grad_b, int64_t pixels, int C, imx_stream_t stream);

/* NCHW fp32 -> NHWC bf16 (channels zero-padded to Cp) and back (first C of Cp). */
IMX_API int imx_nchw_f32_to_nhwc_bf16(const float* src, imx_bf16* dst, int N, int C, int H, int W,
                              int Cp, imx_stream_t stream);
IMX_API int imx_nhwc_bf16_to_nchw_f32(const imx_bf16* src, float* dst, int N, int C, int H, int W,
                              int Cp, int apply_tanh, imx_stream_t stream);

/* Conv weight preparation: W_packed[co][kh][kw][ci] (bf16, ci padded to Cin_p with
 * zeros) = W[co][ci][kh][kw] * (*inv_sigma or 1 if NULL).  With transpose_flip != 0
 * it instead writes the data-gradient filter Wt[ci][K-1-kh][K-1-kw][co] (co padded to
 * Cin_p, which then means the padded Cout).
 * ref: spectral_norm's W/sigma (imaginaire/layers/weight_norm.py:85) fused with the
 * GEMM-layout repack. */
IMX_API int imx_pack_conv_weight_bf16(const float* w, const float* inv_sigma, imx_bf16* packed,
                              int Cout, int Cin, int KH, int KW, int Cin_p,
                              int transpose_flip, imx_stream_t stream);

/* Spectral norm of a filter viewed as W[R][C] (R = Cout, C = Cin*KH*KW), fp32, with
 * torch.nn.utils.spectral_norm's semantics: n_power_iterations times
 * v = normalize(W^T u), u = normalize(W v) in place on the module buffers u[R], v[C] (0 in
 * eval mode), then sigma = u . (W v).  Writes 1/sigma to inv_sigma[0] and copies of the
 * final u, v to u_copy / v_copy (may be NULL) for the backward.  scratch: caller-owned, at
 * least imx_spectral_norm_scratch_bytes(R, C) bytes, need not be initialised.
 * ref: spectral_norm(conv) via imaginaire/layers/weight_norm.py:85; sigma recompute
 *      imaginaire/utils/model_average.py:183-197 */
IMX_API int imx_spectral_norm_sigma_f32(const float* W, float* u, float* v, float* u_copy, float* v_copy,
                                        float* inv_sigma, void* scratch, int64_t scratch_bytes,
                                        int R, int C, int n_power_iterations, float eps,
                                        imx_stream_t stream);
IMX_API int64_t imx_spectral_norm_scratch_bytes(int R, int C);

/* The same for many layers at once (one network forward): descs = device array of n_layers
 * records of eight 64-bit fields {W, u, v, t_off, s_off, out_off, R, C}; t_off / s_off are
 * float offsets into `scratch` (t: C floats that the caller zero-fills, s: R floats),
 * out_off a float offset into `out` where [1/sigma | u copy (R) | v copy (C)] is written.
 * n_power_iterations is 1 (training) or 0 (eval).  Four launches in total. */
IMX_API int imx_spectral_norm_sigma_batched_f32(const void* descs, int n_layers, int max_R, int max_C,
                                                void* scratch, float* out, int n_power_iterations, float eps,
                                                imx_stream_t stream);

/* Moving average of spectrally normalised filters: for every layer of `descs` (same records as
 * imx_spectral_norm_sigma_batched_f32, whose output buffer `sigma_out` holds 1/sigma at each
 * layer's out_off), target = beta * target + (1 - beta) * W / sigma.  targets: device array of
 * n_layers 64-bit pointers to fp32 tensors shaped like W.  One launch.
 * ref: ModelAverage.update_average, imaginaire/utils/model_average.py:87-132, 183-197 */
IMX_API int imx_spectral_norm_ema_update_f32(const void* descs, const void* targets, int n_layers, int64_t max_elems,
                                     const float* sigma_out, float beta, imx_stream_t stream);

/* Backward of W_eff = W / sigma(W) with u, v constant: in place on G = dL/dW_eff,
 * G <- G / sigma - (<G, W> / sigma^2) u v^T  (= dL/dW).  dot_scratch: one float. */
IMX_API int imx_spectral_norm_weight_grad_f32(float* G, const float* W, const float* u, const float* v,
                                              const float* inv_sigma, float* dot_scratch, int R, int C,
                                              imx_stream_t stream);
/* The second half of imx_spectral_norm_weight_grad_f32 for a caller that already holds dot = <G, W>:
 * G = G * inv_sigma - dot * inv_sigma^2 * u v^T, in place.  ref: as above */
IMX_API int imx_spectral_norm_weight_grad_apply_f32(float* G, const float* u, const float* v, const float* inv_sigma,
                                                    const float* dot, int R, int C, imx_stream_t stream);

/* ----------------------------------------------- normalisation + SPADE -- */

/* Per-channel sums over an NHWC bf16 tensor: sum[g*C+c], sumsq[g*C+c] (fp32),
 * g = 0 (per_instance == 0: batch statistics over N*HW) or g = n (instance
 * statistics).  The caller zero-fills sum/sumsq (the kernel accumulates with
 * atomics across CTAs).
 * ref: the statistics half of torch.nn.{SyncBatchNorm,BatchNorm2d,InstanceNorm2d}
 *      built by get_activation_norm_layer, imaginaire/layers/activation_norm.py:377-432 */
IMX_API int imx_bn_stats_nhwc_bf16(const imx_bf16* x, float* sum, float* sumsq,
                           int N, int HW, int C, int per_instance, imx_stream_t stream);

/* mean = sum/count, rstd = 1/sqrt(max(sumsq/count - mean^2, 0) + eps) for G*C
 * entries; if running_mean != NULL (G must be 1) also
 * running = (1-momentum)*running + momentum*{mean, unbiased var}.
 * `count` is the number of elements per entry (after any cross-rank all-reduce of
 * sum/sumsq/count). */
IMX_API int imx_bn_finalize(const float* sum, const float* sumsq, float count, float eps,
                    float* mean, float* rstd, float* running_mean, float* running_var,
                    float momentum, int G, int C, imx_stream_t stream);

/* y = act( ((x-mean)*rstd*w + b) * (1+gamma) + beta ), NHWC bf16, fp32 math.
 *   stats_per_instance: mean/rstd indexed [n*C+c] instead of [c]
 *   gamma/beta: full-size NHWC bf16 (gb_mode 0: SPADE), per-sample fp32 vectors
 *               [n*C+c] (gb_mode 1: AdaptiveNorm), or absent (gb_mode 2: plain norm)
 *   w, b may be NULL (no affine);  act: 0 none, 1 LeakyReLU(0.2), 2 ReLU
 *   up2x != 0: x is N x H/2 x W/2 x C and is read through a nearest x2 upsample.
 * ref: SpatiallyAdaptiveNorm.forward imaginaire/layers/activation_norm.py:211-234,
 *      AdaptiveNorm.forward :79-106, LeakyReLU(0.2) imaginaire/layers/nonlinearity.py:22 */
IMX_API int imx_spade_modulate_fwd(const imx_bf16* x, const float* mean, const float* rstd,
                           const float* w, const float* b,
                           const void* gamma, const void* beta, imx_bf16* y,
                           int N, int H, int W, int C,
                           int stats_per_instance, int gb_mode, int act, int up2x,
                           imx_stream_t stream);

/* Backward of imx_spade_modulate_fwd, pass 1: given dy, recompute the forward and
 * write dgamma, dbeta (gb_mode 0: NHWC bf16; gb_mode 1: accumulated into fp32
 * [n*C+c], caller zero-fills) and accumulate the per-(group,channel) reductions
 * red[0][g*C+c] = sum dxhat_affine, red[1][g*C+c] = sum dxhat_affine * xn
 * (fp32, caller zero-fills; these are also dbias / dweight of the inner norm).
 * gb_sums (optional, gb_mode 0 only): fp32 (2, C), caller zero-fills; receives the sums of
 * dgamma and dbeta over all pixels of the batch, i.e. the bias gradients of the convolutions
 * that produced gamma and beta, saving two passes over those tensors. */
IMX_API int imx_spade_modulate_bwd_reduce(const imx_bf16* x, const imx_bf16* dy,
                                  const float* mean, const float* rstd,
                                  const float* w, const float* b,
                                  const void* gamma, const void* beta,
                                  void* dgamma, void* dbeta, float* red, float* gb_sums,
                                  int N, int H, int W, int C,
                                  int stats_per_instance, int gb_mode, int act,
                                  imx_stream_t stream);

/* Pass 2: dx = rstd*w*( g - mean(g) - xn*mean(g*xn) ), g = dy*act'*(1+gamma), using
 * the (all-reduced) reductions of pass 1 divided by `count`.  With use_batch_stats
 * == 0 (eval-mode BatchNorm: running statistics are constants) dx = rstd*w*g. */
IMX_API int imx_spade_modulate_bwd_dx(const imx_bf16* x, const imx_bf16* dy,
                              const float* mean, const float* rstd,
                              const float* w, const float* b,
                              const void* gamma, const void* beta,
                              const float* red, float count, imx_bf16* dx,
                              int N, int H, int W, int C,
                              int stats_per_instance, int gb_mode, int act,
                              int use_batch_stats, imx_stream_t stream);

/* ------------------------------------------------- implicit-GEMM convs -- */

/* Epilogue flags for imx_conv2d_igemm_fprop_bf16 */
#define IMX_EPI_BIAS      1   /* + bias[co] (fp32) */
#define IMX_EPI_RELU      2
#define IMX_EPI_LRELU     4   /* LeakyReLU(0.2) */
#define IMX_EPI_RESIDUAL  8   /* + residual[n,oy,ox,co] (NHWC bf16, same shape as out) */
#define IMX_EPI_TANH      16
#define IMX_EPI_STATS     32  /* accumulate per-channel sum / sumsq of the stored
                                 (bf16-rounded) output into stats, laid out (2, G, Cout)
                                 like imx_bn_stats_nhwc_bf16 (fp32 atomics; caller
                                 zero-fills); G = 1 unless IMX_EPI_STATS_INSTANCE */
#define IMX_EPI_STATS_INSTANCE 128 /* with IMX_EPI_STATS: per-sample statistics, G = N */
#define IMX_EPI_OUT_F32   64  /* out is fp32 NHWC instead of bf16 */

/* out[n,oy,ox,co] = epilogue( sum_{kh,kw,ci} x[n, oy*stride+kh-pad, ox*stride+kw-pad, ci]
 *                                            * wpk[co][kh][kw][ci] ), zero padding.
 * x: N x H x W x Cin_p bf16 (Cin_p % 8 == 0), wpk: Cout x KH x KW x Cin_p bf16 from
 * imx_pack_conv_weight_bf16, out: N x Ho x Wo x Cout_p (Cout_p >= Cout; channels
 * >= Cout are written as zero when Cout_p > Cout).  tcgen05 (M=128 pixel tile, N<=256
 * channel tile, fp32 accumulators in TMEM), operands staged by TMA (128-byte swizzle);
 * the zero padding is TMA out-of-bounds fill.
 * The same entry point computes the data gradient of a stride-1 convolution when
 * given dy as x, the transposed+flipped filter, and pad' = K-1-pad.
 * ref: torch.nn.Conv2d inside Conv2dBlock, imaginaire/layers/conv.py:93-104, 262-327 */
IMX_API int imx_conv2d_igemm_fprop_bf16(const imx_bf16* x, const imx_bf16* wpk, const float* bias,
                                const imx_bf16* residual, void* out, float* stats,
                                int N, int H, int W, int Cin_p,
                                int Cout, int Cout_p, int KH, int KW,
                                int stride, int pad, int flags, imx_stream_t stream);

/* Layers with at most 8 channels on one side (the 5x5 to-RGB convolution 256 -> 3 of
 * imaginaire/generators/spade.py:318-323, its gradients, the first layer of the image networks)
 * run as a 1x1 GEMM on the tensor cores plus one of these two memory-bound passes.
 *
 * imx_im2col_c8_nhwc_bf16: col[n,oy,ox,(kh,kw,c)] = x[n, oy*stride+kh-pad, ox*stride+kw-pad, c]
 * (zero outside), x: N x H x W x 8 bf16, col: N x Ho x Wo x (KH*KW*8) bf16.  A convolution of x
 * is then a 1x1 convolution of col with the packed filter [Cout][KH*KW*8].
 *
 * imx_shift_add_taps_f32: y[n,oy,ox,co] = bias[co] + sum_{kh,kw} P[n, oy+kh-pad, ox+kw-pad,
 * (kh*KW+kw)*Cout + co] for a 'same' convolution (2*pad == K-1), where
 * P[n,iy,ix,(tap,co)] = sum_ci x[n,iy,ix,ci] * w[co,ci,tap] (fp32, `ncols` floats per pixel) came
 * from a 1x1 convolution with KH*KW*Cout output columns.  y: N x H x W x 8 bf16, channels >= Cout
 * are zero; bias may be NULL.
 * ref: torch.nn.Conv2d inside Conv2dBlock, imaginaire/layers/conv.py:93-104 */
IMX_API int imx_im2col_c8_nhwc_bf16(const imx_bf16* x, imx_bf16* col, int N, int H, int W,
                            int KH, int KW, int stride, int pad, imx_stream_t stream);
IMX_API int imx_shift_add_taps_f32(const float* P, const float* bias, imx_bf16* y, int N, int H, int W,
                           int ncols, int KH, int KW, int pad, int Cout, imx_stream_t stream);

/* Split-K variant for layers with fewer pixel tiles than SMs (small images, many channels):
 * the K loop (taps x channel blocks) is divided over several CTAs that add fp32 partial
 * sums into `workspace`; a finishing kernel applies the epilogue.  The library zero-fills
 * the workspace itself (one cudaMemsetAsync on `stream`).
 * imx_conv2d_fprop_workspace_bytes returns the bytes this layer wants (0: the plain tiling
 * is already the fastest and workspace may be NULL; < 0: invalid arguments).  With
 * workspace == NULL, or IMX_EPI_STATS set, this is exactly imx_conv2d_igemm_fprop_bf16.
 * ref: same as imx_conv2d_igemm_fprop_bf16 (no separate reference counterpart). */
IMX_API int64_t imx_conv2d_fprop_workspace_bytes(int N, int H, int W, int Cin_p, int Cout, int Cout_p,
                                         int KH, int KW, int stride, int pad, int flags);
IMX_API int imx_conv2d_igemm_fprop_splitk_bf16(const imx_bf16* x, const imx_bf16* wpk, const float* bias,
                                       const imx_bf16* residual, void* out, float* stats,
                                       void* workspace, int64_t workspace_bytes,
                                       int N, int H, int W, int Cin_p,
                                       int Cout, int Cout_p, int KH, int KW,
                                       int stride, int pad, int flags, imx_stream_t stream);

/* The same convolution with the accumulator multiplied by *scale (ONE fp32 in device memory) before the bias:
 * y = act(scale * conv(x, wpk) + bias) (+ residual).  For spectrally normalised layers: the bf16 filter is packed
 * unscaled once per optimizer step and 1 / sigma -- which changes with every application's power iteration -- enters
 * here.  workspace / workspace_bytes as in imx_conv2d_igemm_fprop_splitk_bf16, or null / 0 for no split-K.
 * ref: torch.nn.utils.spectral_norm's weight = weight_orig / sigma (imaginaire/layers/weight_norm.py:85) */
IMX_API int imx_conv2d_igemm_fprop_scaled_bf16(const imx_bf16* x, const imx_bf16* wpk, const float* scale,
                                               const float* bias, const imx_bf16* residual, void* out, float* stats,
                                               void* workspace, int64_t workspace_bytes,
                                               int N, int H, int W, int Cin_p,
                                               int Cout, int Cout_p, int KH, int KW,
                                               int stride, int pad, int flags, imx_stream_t stream);

/* One piece of a larger output tensor, stride 1:
 *   out[n][oy * out_stride_y + out_y0][ox * out_stride_x + out_x0][:] =
 *       act(scale * sum_{kh,kw} x[n][oy + kh - pad_y][ox + kw - pad_x][:] . w[:][kh][kw][:] + bias) (+ residual)
 * for oy < Ho, ox < Wo; out (and residual) are N x out_H x out_W x Cout_p; rows / columns outside x read as zeros on
 * every side, so pad_y / pad_x are the LEFT paddings only and Ho / Wo are free.  scale, bias, residual may be null.
 * The data gradient of a stride-2 convolution is four such launches -- one per parity class (a, b) of the input pixel
 * (2m + a, 2n + b), x = dy, w = the taps kh = (a + pad) mod 2, +2, ... of the filter, transposed
 * (imaginaire_b200/functional.py: conv2d_dgrad) -- instead of a stride-1 convolution over a zero-inserted copy of dy
 * that spends three quarters of its multiply-adds on zeros.
 * ref: the cuDNN data gradient behind torch's conv2d backward for the stride-2 layers of
 * imaginaire/discriminators/multires_patch.py:197-215 and fpse.py:62-75 */
IMX_API int imx_conv2d_igemm_fprop_window_bf16(const imx_bf16* x, const imx_bf16* wpk, const float* scale,
                                               const float* bias, const imx_bf16* residual, void* out,
                                               int N, int H, int W, int Cin_p, int Cout, int Cout_p, int KH, int KW,
                                               int pad_y, int pad_x, int Ho, int Wo, int out_H, int out_W,
                                               int out_stride_y, int out_stride_x, int out_y0, int out_x0,
                                               int flags, imx_stream_t stream);

/* The four parity-class sub-filters of a stride-2 data gradient (see imx_conv2d_igemm_fprop_window_bf16) from the
 * transposed, flipped filter full = imx_pack_conv_weight_bf16(..., transpose_flip = 1): Cin x KH x KW x Cout_p bf16.
 * out (as many elements as full) receives the classes (a, b) = (0,0), (0,1), (1,0), (1,1) one after the other, class
 * (a, b) as Cin x T(a) x T(b) x Cout_p with T(a) = number of taps k = (a + pad) mod 2, +2, ... below KH (KW);
 * class_offsets (host memory, 4 entries, may be null) receives where each class starts, in elements. */
IMX_API int imx_split_dgrad_s2_filter_bf16(const imx_bf16* full, imx_bf16* out, int Cin, int KH, int KW, int Cout_p,
                                           int pad, int64_t* class_offsets, imx_stream_t stream);

/* SPADE projection conv fused with the modulation (gamma/beta never reach HBM):
 *   gb = conv(hidden; wgb) (+bias), wgb rows interleaved per 64-channel block as
 *   [64 gamma rows | 64 beta rows] (imx_pack_spade_gb_weight_bf16),
 *   y[n,p,c] = act( ((x-mean)*rstd*w+b) * (1+gamma) + beta ).
 * hidden: N x H x W x Ch_p bf16; x, y: N x H x W x C bf16 (x at half resolution when
 * up2x != 0).  Same tiling as the plain conv; the epilogue reads the x tile.
 * ref: imaginaire/layers/activation_norm.py:226-233 */
IMX_API int imx_spade_gb_conv_modulate_bf16(const imx_bf16* hidden, const imx_bf16* wgb,
                                    const float* bias_gamma, const float* bias_beta,
                                    const imx_bf16* x, const float* mean, const float* rstd,
                                    const float* w, const float* b, imx_bf16* y,
                                    int N, int H, int W, int Ch_p, int C, int KH, int KW, int pad,
                                    int stats_per_instance, int act, int up2x,
                                    imx_stream_t stream);

/* imx_spade_gb_conv_modulate_bf16 with the gamma accumulators multiplied by *scale_gamma and the beta accumulators
 * by *scale_beta (one fp32 each in device memory; null = 1) before their biases.  With spectrally normalised
 * projections the interleaved filter is packed unscaled once per optimizer step and the two 1 / sigma enter here.
 * ref: as imx_spade_gb_conv_modulate_bf16 + imaginaire/layers/weight_norm.py:85 */
IMX_API int imx_spade_gb_conv_modulate_scaled_bf16(const imx_bf16* hidden, const imx_bf16* wgb,
                                                   const float* scale_gamma, const float* scale_beta,
                                                   const float* bias_gamma, const float* bias_beta,
                                                   const imx_bf16* x, const float* mean, const float* rstd,
                                                   const float* w, const float* b, imx_bf16* y,
                                                   int N, int H, int W, int Ch_p, int C, int KH, int KW, int pad,
                                                   int stats_per_instance, int act, int up2x,
                                                   imx_stream_t stream);

IMX_API int imx_pack_spade_gb_weight_bf16(const float* w_gamma, const float* w_beta,
                                  const float* inv_sigma_gamma, const float* inv_sigma_beta,
                                  imx_bf16* packed, int C, int Ch, int KH, int KW, int Ch_p,
                                  imx_stream_t stream);

/* Weight gradient: dw[co][ci][kh][kw] (fp32, PyTorch layout, overwritten) =
 * sum_{n,oy,ox} dy[n,oy,ox,co] * x[n, oy*stride+kh-pad, ox*stride+kw-pad, ci].
 * x: N x H x W x Cin_p bf16, dy: N x Ho x Wo x Cout_p bf16.  `workspace` is caller-owned
 * scratch of at least imx_conv2d_wgrad_workspace_bytes(...) bytes (split-K partial
 * sums over the N*Ho*Wo reduction); it need not be initialised.
 * ref: autograd of torch.nn.Conv2d (imaginaire/layers/conv.py:93-104) */
IMX_API int imx_conv2d_igemm_wgrad_bf16(const imx_bf16* x, const imx_bf16* dy, float* dw,
                                void* workspace, int64_t workspace_bytes,
                                int N, int H, int W, int Cin, int Cin_p,
                                int Cout, int Cout_p, int KH, int KW,
                                int stride, int pad, imx_stream_t stream);
IMX_API int64_t imx_conv2d_wgrad_workspace_bytes(int N, int H, int W, int Cin, int Cin_p,
                                         int Cout, int Cout_p, int KH, int KW,
                                         int stride, int pad);

/* imx_conv2d_igemm_wgrad_bf16 plus *dot += <dw, w_orig> (w_orig: fp32, dw's layout; dot: one fp32 the caller zeroed),
 * accumulated while the split-K partial sums are reduced: the inner product of the spectral-norm gradient
 * (imx_spectral_norm_weight_grad_f32) without another pass over both tensors.
 * ref: autograd of torch.nn.utils.spectral_norm's weight = weight_orig / sigma (imaginaire/layers/weight_norm.py:85) */
IMX_API int imx_conv2d_igemm_wgrad_dot_bf16(const imx_bf16* x, const imx_bf16* dy, float* dw, const float* w_orig,
                                            float* dot, void* workspace, int64_t workspace_bytes,
                                            int N, int H, int W, int Cin, int Cin_p,
                                            int Cout, int Cout_p, int KH, int KW,
                                            int stride, int pad, imx_stream_t stream);

/* Data-parallel weight gradient without a separate gradient collective: the
 * split-K partial sums of every rank live in peer-mapped NVLink memory (e.g.
 * torch.distributed._symmetric_memory), and the reduction over splits doubles as the reduction
 * over ranks.
 *   1. imx_conv2d_igemm_wgrad_partials_bf16: the GEMM of imx_conv2d_igemm_wgrad_bf16 only; fills
 *      workspace[split][tap][co_r][ci_r] and returns plan_out = {splits, co_r, ci_r} (host ints).
 *   2. (all ranks reach a barrier), then imx_wgrad_reduce_peers_f32 on every rank: sums the output
 *      channels [co_begin, co_end) over ALL ranks' workspaces (ws_peers: HOST array of `world`
 *      device pointers) and splits, writes scale * sum (or adds it, accumulate != 0) into dw
 *      ([Cout][Cin][KH][KW] fp32, this rank's copy).
 *   3. once per backward, after another barrier, imx_gather_rows_peers_f32 copies the channel
 *      ranges owned by the other ranks from their gradient arenas (arena_peers: HOST array of
 *      device pointers, identical layout on every rank; descs: DEVICE array of int64 triples
 *      {offset, rows, row_elems}; rank q owns rows [q*rows/world, (q+1)*rows/world)).
 * Replaces DDP's bucketed NCCL all-reduce for the convolution filters (the reference: apex /
 * torch DDP in imaginaire/utils/trainer.py:193-216); no counterpart kernel in the reference. */
IMX_API int imx_conv2d_igemm_wgrad_partials_bf16(const imx_bf16* x, const imx_bf16* dy,
                                         void* workspace, int64_t workspace_bytes, int* plan_out,
                                         int N, int H, int W, int Cin, int Cin_p,
                                         int Cout, int Cout_p, int KH, int KW,
                                         int stride, int pad, imx_stream_t stream);
IMX_API int imx_wgrad_reduce_peers_f32(const void* const* ws_peers, int world, float* dw, int co_begin, int co_end,
                               int Cout, int Cin, int taps, int splits, int co_r, int ci_r,
                               float scale, int accumulate, imx_stream_t stream);
IMX_API int imx_gather_rows_peers_f32(const void* const* arena_peers, int world, int rank, const void* descs,
                              int n_tensors, int64_t max_elems, imx_stream_t stream);

/* The same averaging with the split-K reduction kept local: every rank writes its own complete gradients (e.g.
 * imx_conv2d_igemm_wgrad_bf16 with dw pointing into the arena) into the same offsets of its arena of symmetric
 * memory; after a barrier of all ranks, imx_peer_reduce_scatter_f32 leaves scale * (sum over ranks) in the
 * rank-th of `world` contiguous chunks of every segment of THIS rank's arena (chunk boundaries rounded to 4
 * floats); after another barrier imx_peer_all_gather_f32 fetches the other chunks from their owners; a third
 * barrier before anybody writes its arena again.  segs: DEVICE array of int64 pairs {offset, numel} (offsets in
 * floats, multiples of 4), identical on every rank; max_numel sizes the grid.  (world-1)/world of the gradient
 * bytes are read remotely in each phase, as coalesced 16-byte loads.
 * ref: DDP's gradient all-reduce, imaginaire/utils/trainer.py:193-216 */
IMX_API int imx_peer_reduce_scatter_f32(const void* const* arena_peers, int world, int rank, const void* segs,
                                        int n_seg, int64_t max_numel, float scale, imx_stream_t stream);
IMX_API int imx_peer_all_gather_f32(const void* const* arena_peers, int world, int rank, const void* segs, int n_seg,
                                    int64_t max_numel, imx_stream_t stream);

/* The weight gradient again, for callers that are not able to allocate device memory (no
 * counterpart in the reference: its extensions always let torch allocate).  Think of a host
 * that reserves its whole device-memory budget at start-up and forbids cudaMalloc afterwards,
 * or a plugin thread that has no allocator: they cannot provide the split-K workspace that
 * imx_conv2d_igemm_wgrad_bf16 asks for.  Here the workspace is a device array of
 * IMX_WGRAD_NOALLOC_SCRATCH_BYTES built into the library (each GPU that loads the library
 * holds that much until the process exits); the call allocates and frees nothing.
 *   - Same arguments and result as imx_conv2d_igemm_wgrad_bf16 minus the workspace; x, dy and
 *     dw are still the caller's buffers.
 *   - A layer whose partial sums need more than the built-in array is rejected with
 *     IMX_ERR_TOO_LARGE (10004) and a message naming both sizes; nothing is launched.
 *   - The first call on each device looks up the array's device address (one
 *     cudaGetSymbolAddress with that device current).  Make that first call during set-up,
 *     before any stream capture begins.
 *   - All calls share the one array.  Its contents survive from call to call but never have to
 *     be cleared: each call overwrites every slot it later reads.  Two calls that overlap on
 *     the GPU (different streams, or another thread) would write the same scratch and race;
 *     run them one after the other. */
#define IMX_WGRAD_NOALLOC_SCRATCH_BYTES (64 * 1024 * 1024)
IMX_API int imx_conv2d_igemm_wgrad_bf16_noalloc(const imx_bf16* x, const imx_bf16* dy, float* dw,
                                                int N, int H, int W, int Cin, int Cin_p,
                                                int Cout, int Cout_p, int KH, int KW,
                                                int stride, int pad, imx_stream_t stream);

/* Per-channel sum over pixels of an NHWC bf16 tensor (bias gradient):
 * out[c] = sum_{n,p} x[n,p,c]; fp32, caller zero-fills. */
IMX_API int imx_channel_sum_nhwc_bf16(const imx_bf16* x, float* out, int NP, int C, int Cp,
                              imx_stream_t stream);

/* dx = dy * act'(y) for an activation fused into a conv epilogue (y = act(pre): LeakyReLU(0.2)
 * or ReLU, so act' depends on the sign of y only), and, when bias_grad != NULL, bias_grad[c] +=
 * sum over the NP pixels of dx (fp32, C entries, caller zero-fills): the backward of
 * `nonlinearity(conv(x) + bias)` up to the conv itself in one pass.  All tensors NP x C bf16.
 * ref: autograd of imaginaire/layers/conv.py:77-91 (order 'CNA' blocks), nonlinearity.py:22 */
IMX_API int imx_act_bwd_bias_nhwc_bf16(const imx_bf16* dy, const imx_bf16* y, imx_bf16* dx, float* bias_grad,
                               int64_t NP, int C, int act, imx_stream_t stream);

/* ------------------------------------------------- FlowNet2 operators -- */

/* output[b,0,y,x] = sqrt(sum_c input[b,c,y,x]^2) and its gradient
 * grad_input = grad_output * input / (output + 1e-9).  NCHW fp32 contiguous.  norm_deg must be
 * 2 (the reference accepts the argument and ignores it).  The reference's Python side zero-fills
 * the outputs first; these kernels write every element, so that is not required.
 * ref: channelnorm_cuda.forward / backward, imaginaire/third_party/channelnorm/src/
 *      channelnorm_cuda.cc:6-30, channelnorm_kernel.cu:18-96, channelnorm.py:8-32 */
IMX_API int imx_channelnorm_fwd_f32(const float* input, float* output, int B, int C, int H, int W, int norm_deg,
                            imx_stream_t stream);
IMX_API int imx_channelnorm_bwd_f32(const float* input, const float* output, const float* grad_output,
                            float* grad_input, int B, int C, int H, int W, int norm_deg, imx_stream_t stream);

/* FlowNetC correlation (cost volume).  output[b, (tj,ti), y, x] = 1/(k*k*C) * sum over the
 * k x k patch and channels of pad(input1)[y1+j, x1+i] * pad(input2)[y1+tj*stride2+j, ...], with
 * (y1, x1) = (y, x) * stride1 + max_displacement, displacements tj, ti in
 * [-max_displacement/stride2, +max_displacement/stride2].  Inputs NCHW fp32 [B,C,H,W]; rbot1 /
 * rbot2: caller-owned scratch of B*(H+2*pad)*(W+2*pad)*C floats each (the reference's rInput1/2,
 * which its callee resizes; here the caller sizes them, nothing is allocated); output
 * [B, out_channels, out_h, out_w] from imx_correlation_output_size (host-side, no stream).
 * The gradient needs stride1 == 1 and pad_size >= max_displacement; grad_input1/2 are fully
 * written.  Geometries whose reads would leave the padded planes are refused (the reference
 * reads out of bounds for kernel_size > 1).  corr_type_multiply is always 1 in the reference.
 * ref: correlation_cuda.forward / backward, imaginaire/third_party/correlation/src/
 *      correlation_cuda.cc:10-172, correlation_cuda_kernel.cu:46-336, correlation.py:8-105 */
IMX_API int imx_correlation_output_size(int C, int H, int W, int pad_size, int kernel_size, int max_displacement,
                                int stride1, int stride2, int* out_channels, int* out_h, int* out_w);
IMX_API int imx_correlation_fwd_f32(const float* input1, const float* input2, float* rbot1, float* rbot2, float* output,
                            int B, int C, int H, int W, int pad_size, int kernel_size, int max_displacement,
                            int stride1, int stride2, imx_stream_t stream);
IMX_API int imx_correlation_bwd_f32(const float* input1, const float* input2, const float* grad_output,
                            float* rbot1, float* rbot2, float* grad_input1, float* grad_input2,
                            int B, int C, int H, int W, int pad_size, int kernel_size, int max_displacement,
                            int stride1, int stride2, imx_stream_t stream);

/* ---------------------------------------------------- peer-memory exchange -- */

/* Small-vector all-reduce between the GPUs of one node through symmetric (peer-mapped) memory, for
 * SyncBatchNorm's statistics.  regions: HOST array of `world` device pointers, regions[q] = rank q's exchange
 * region (imx_peer_exchange_region_bytes() bytes of symmetric memory, zero-filled once before the first call;
 * regions[rank] is local).  Every rank of the group must issue the same sequence of calls.  One single-CTA
 * kernel per call on the caller's stream: push to all peers, flag, wait on local flags, sum in rank order (all
 * ranks get bit-identical sums).  n <= 8192 floats; `out` may be `in`.  The call counter lives in the region and is advanced by the
 * kernel, so a captured launch stays valid on replay.
 * ref: torch.nn.SyncBatchNorm's exchanges (torch/nn/modules/_functions.py:49-165: all_gather of
 *      [mean, invstd, count] forward, all_reduce of [sum_dy, sum_dy_xmu] backward), which
 *      imaginaire/layers/activation_norm.py:403-410 instantiates */
IMX_API int64_t imx_peer_exchange_region_bytes(void);
IMX_API int imx_peer_allreduce_sum_f32(const void* const* regions, int world, int rank, const float* in, float* out,
                                       int n, imx_stream_t stream);

/* The forward exchange fused with imx_bn_finalize: sums = this rank's [sum x (C) | sum x^2 (C)], count = the
 * GLOBAL number of elements per channel; writes mean / rstd (C) and, if given, updates the running statistics
 * (momentum, unbiased variance) -- identical on every rank.
 * ref: as above + torch batch_norm_gather_stats_with_counts */
IMX_API int imx_peer_bn_finalize_f32(const void* const* regions, int world, int rank, const float* sums, int C,
                                     float count, float eps, float* mean, float* rstd, float* running_mean,
                                     float* running_var, float momentum, imx_stream_t stream);

/* ------------------------------------------------------------ optimizer -- */

/* One Adam step (torch.optim.Adam arithmetic, no amsgrad) over n_tensors fp32 tensors:
 *   g = grad + weight_decay * p;  m += (1-beta1)(g-m);  v = beta2 v + (1-beta2) g^2;
 *   p -= lr/(1-beta1^step) * m / (sqrt(v)/sqrt(1-beta2^step) + eps)
 * params / grads / exp_avg / exp_avg_sq: HOST arrays of n_tensors device pointers, numel: host
 * array of element counts; `step` is the 1-based step number AFTER this update's increment.
 * Hyper-parameters are doubles because 1 - beta2 has to be formed before rounding to fp32.
 * The pointer tables are passed through kernel arguments: no allocation, no copies, one launch
 * per ~96 tensors / 39 M elements.  28 bytes of HBM traffic per parameter.
 * ref: get_optimizer_for_params 'adam' (apex FusedAdam / torch Adam),
 *      imaginaire/utils/trainer.py:261-306 */
IMX_API int imx_adam_multi_tensor_f32(const void* const* params, const void* const* grads,
                              const void* const* exp_avg, const void* const* exp_avg_sq,
                              const int64_t* numel, int n_tensors,
                              double lr, double beta1, double beta2, double eps, double weight_decay,
                              int64_t step, imx_stream_t stream);

/* The same step with the step numbers on the DEVICE: steps[i] points to tensor i's counter (one fp32, already
 * incremented for this update, as torch.optim.Adam(capturable=True) keeps it); the bias corrections are formed in
 * the kernel.  A launch of this entry point captured in a CUDA graph stays valid as the counters advance (the host
 * number baked into imx_adam_multi_tensor_f32's arguments would not).
 * ref: same as imx_adam_multi_tensor_f32 */
IMX_API int imx_adam_multi_tensor_f32_capturable(const void* const* params, const void* const* grads,
                                                 const void* const* exp_avg, const void* const* exp_avg_sq,
                                                 const void* const* steps, const int64_t* numel, int n_tensors,
                                                 double lr, double beta1, double beta2, double eps,
                                                 double weight_decay, imx_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif  /* IMX_H_ */

/* ubresnet_b200 -- C ABI of the sm_100a hot-path library (libubresnet_b200.so).
 *
 * The reference (NuTufts/ubresnet) has NO plugin/FFI boundary: its hot path sits behind the Python
 * nn.Module API (models/ub_uresnet.py:31,88; training/pixelwise_nllloss.py:35,41) and every device
 * op is a torch.nn library call.  The entry points below are what a binding for that path needs:
 * one call per reference op (or per fused group of ops), taking raw DEVICE pointers, plain sizes and
 * a cudaStream_t (as void*).  No torch types, no C++ types, no exceptions cross this boundary.
 *
 * Conventions
 *   - every function returns 0 on success, <0 on error (UBR_E_*); the message is in ubr_last_error()
 *     (thread local).  There is NO CPU fallback: a non-sm_100 device gives UBR_E_UNSUPPORTED.
 *   - the library never allocates caller-visible memory; outputs and workspaces are caller provided
 *     (sizes via the *_workspace_bytes helpers) and must stay valid until the stream reaches the op.
 *   - activations are NHWC bf16 "views": a channel slice [0,c) of a pixel-major buffer with pixel
 *     pitch `ld` elements (ld >= c, ld*2 bytes multiple of 16, ptr 16-byte aligned).  Channel slices
 *     make torch.cat([deconv, skip], 1) (models/common_layers.py:130) a zero-copy operation.
 *   - parameters stay fp32 in the reference layouts (OIHW, ConvTranspose IOHW); ubr_pack_weights
 *     derives the bf16 GEMM operand copies.
 */
#ifndef UBRESNET_B200_H
#define UBRESNET_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define UBR_ABI_VERSION 1
#define UBR_OK 0
#define UBR_E_BADARG (-1)
#define UBR_E_UNSUPPORTED (-2)
#define UBR_E_CUDA (-3)
#define UBR_E_NCCL (-4)

typedef struct ubr_view {
  void* ptr;      /* device pointer to (n=0,h=0,w=0,c=0) of the slice, bf16 */
  int32_t n, h, w, c;
  int32_t ld;     /* elements between consecutive pixels of the underlying buffer */
} ubr_view;

/* ---- library ------------------------------------------------------------------------------ */
int ubr_abi_version(void);
const char* ubr_last_error(void);
int ubr_device_check(void);               /* 0 iff the current CUDA device is sm_100 (B200) */
int ubr_sync_check(void* stream);         /* cudaStreamSynchronize + surface async errors */
int ubr_shutdown(void);                   /* teardown: the library owns no device memory; clears trace / error state */
uint64_t ubr_launch_count(void);          /* kernels launched by this library so far (process wide) */
/* profiling aid (bench.py): when enabled, ubr_take_kernel_names() returns "k1+k2" = the kernels the calling thread
 * launched since the previous take (thread-local buffer, valid until the next take) */
void ubr_trace_kernels(int enable);
const char* ubr_take_kernel_names(void);

/* ---- weight packing: fp32 reference layout -> bf16 [tap][n][k] GEMM operand ---------------------
 * replaces the implicit cuDNN filter transforms behind nn.Conv2d / nn.ConvTranspose2d.
 *   src_layout 0: OIHW  (nn.Conv2d.weight,            models/common_layers.py:15,34; ub_uresnet.py:41,60,64)
 *   src_layout 1: IOHW  (nn.ConvTranspose2d.weight,   models/common_layers.py:125)
 *   src_layout 2: IOHW ConvTranspose2d(k4,s2,p1) as a 3x3 stride-1 "sub-pixel" conv over the INPUT grid with 4*out_ch
 *                 phase channels ((ph*2+pw)*out_ch + co <-> output pixel (2i+ph, 2j+pw)); call with kh = kw = 3.
 *                 n_is_out 1: forward operand (n = phase channel, k = in channel); 0: data-gradient operand
 *   n_is_out 1: n = output channel, k = input channel (forward operand)
 *   n_is_out 0: n = input channel,  k = output channel (data-gradient operand)
 * dst has kh*kw + 1 slabs of [n_pad][k_pad] (zero padded, n_pad/k_pad multiples of 16); the extra last slab is
 * all zero (used by stride-2 data-gradient phases that no filter tap reaches). */
int ubr_pack_weights(const float* src, int src_layout, int n_is_out, int out_ch, int in_ch, int kh, int kw,
                     int n_pad, int k_pad, void* dst_bf16, void* stream);

/* batched form: `jobs_device` is a DEVICE array of njobs records of 10 int64:
 * {src ptr, dst ptr, src_layout, n_is_out, out_ch, in_ch, kh, kw, n_pad, k_pad}; one launch repacks every conv of
 * the network after an optimizer step. */
int ubr_pack_weights_batch(const int64_t* jobs_device, int njobs, void* stream);

/* ---- tensor-core implicit-GEMM convolutions (tcgen05 + TMEM + TMA) ----------------------------
 * nn.Conv2d(k in {1,3,7}, stride in {1,2}, padding=(k-1)/2)  forward:
 *   y[n,ho,wo,:] = sum_{r,s} x[n, ho*stride+r-pad, wo*stride+s-pad, :] * W[:, :, r, s]^T  (+ bias)
 * wpk = ubr_pack_weights(..., n_is_out=1).  bias may be NULL.
 * If y_f32 != NULL the result is written as fp32 [pixel][y_f32_ld] (first out_ch columns) instead of y. */
int ubr_conv2d_fwd(ubr_view x, const void* wpk, const float* bias, int out_ch, int ksize, int stride,
                   ubr_view y, float* y_f32, int y_f32_ld, void* stream);
/* forward conv that also emits the BatchNorm statistics partials of its (bf16-stored) output from the epilogue:
 * bn_workspace receives *stats_rows rows of [2][out_ch] (sum, sum of squares) -- feed them to ubr_bn_finalize_rows.
 * Returns 1 (and does nothing) when the shape has no fused path: run ubr_conv2d_fwd + ubr_bn_stats instead. */
int ubr_conv2d_fwd_stats(ubr_view x, const void* wpk, const float* bias, int out_ch, int ksize, int stride, ubr_view y,
                         void* bn_workspace, int* stats_rows, void* stream);
/* conv + training-mode BatchNorm statistics + their finalize in ONE launch: the epilogue's partial rows are summed, in
 * a fixed order, by the last CTAs to arrive (a two-level ticket tree in `tree`), and the last of them writes scale /
 * shift / mean / invstd and updates the running statistics -- what ubr_conv2d_fwd_stats + ubr_bn_finalize_rows do in
 * two launches (models/common_layers.py:41-42,47-48: conv -> bn in training mode).  `tree`: ubr_bn_tree_bytes(c) bytes of
 * device memory, zero-filled ONCE after allocation (the library leaves it zeroed); launches sharing a tree or workspace
 * must be stream-ordered.  Returns 1 (nothing launched) when the shape has no fused-statistics epilogue. */
int ubr_conv2d_fwd_bn(ubr_view x, const void* wpk, const float* bias, int out_ch, int ksize, int stride, ubr_view y,
                      void* bn_workspace, void* tree, const float* gamma, const float* beta, float* running_mean,
                      float* running_var, float momentum, float eps, float* scale, float* shift, float* mean,
                      float* invstd, void* stream);
/* Inference form of a residual-block conv with BatchNorm folded into (wpk, bias) by the caller (SURVEY 8f N1;
 * common_layers.py:39-58 in eval mode):  y = relu_post?( relu_pre?(conv(x) + bias) + res ).  `res` (ptr NULL = none)
 * is a bf16 view on the output grid.  bf16 output only. */
int ubr_conv2d_fwd_fused(ubr_view x, const void* wpk, const float* bias, int out_ch, int ksize, int stride, int relu_pre,
                         ubr_view res, int relu_post, ubr_view y, void* stream);
/* Eval-mode residual block chain (models/common_layers.py:39-58 with BatchNorm folded):
 *   y = relu( relu( conv3x3(relu(conv3x3(x, w1) + b1), w2) + b2 ) + res )
 * in ONE launch: the intermediate activation stays in shared memory (row bands with one recomputed halo row on either
 * side).  w1pk / w2pk: ubr_pack_weights forward layouts of [planes, x.c, 3, 3] and [planes, planes, 3, 3]; b1 / b2:
 * fp32 [planes]; res: bf16 view like y (x itself for an identity block, the projected bypass otherwise).
 * Returns 1 (nothing launched) when the shape is not eligible -- x.w % 128 != 0, channels outside {16,32,64},
 * (x.w/128)*(planes/16) > 4 -- and the caller runs two ubr_conv2d_fwd_fused calls instead. */
int ubr_block_chain_fwd(ubr_view x, const void* w1pk, const float* b1, const void* w2pk, const float* b2, int planes,
                        ubr_view res, ubr_view y, void* stream);
/* data gradient of the same conv: dx = conv_transpose(dy, W).  wpk = pack(..., n_is_out=0). */
int ubr_conv2d_dgrad(ubr_view dy, const void* wpk, int in_ch, int ksize, int stride, ubr_view dx, void* stream);
/* dx += conv_transpose(dy, W), fused into the epilogue (fp32 add of the stored bf16 value, one rounding).
 * Returns 1 (and does nothing) when the shape has no fused path (stride 2): the caller then adds with ubr_add. */
int ubr_conv2d_dgrad_acc(ubr_view dy, const void* wpk, int in_ch, int ksize, int stride, ubr_view dx, void* stream);
/* data gradient that ALSO emits the reduction pass of the BatchNorm(+ReLU) backward that consumes it
 * (common_layers.py:41-44 in reverse: dx is the gradient wrt a = relu?(bn(y_pre))): while the epilogue holds the stored
 * dx tile it reads y_pre and accumulates sum dz and sum dz*xhat per channel (dz = dx * [y_pre*scale+shift > 0]) into
 * *rows partial rows of bn_workspace -- feed them to ubr_bn_act_bwd_apply_rows instead of running
 * ubr_bn_act_bwd_reduce (one full read of dx and y_pre saved).  Returns 1 (nothing launched) when the shape has no
 * fused path (stride 2, in_ch > 32, W %% 128 != 0): run ubr_conv2d_dgrad + ubr_bn_act_bwd_reduce/apply. */
int ubr_conv2d_dgrad_bnreduce(ubr_view dy, const void* wpk, int in_ch, int ksize, int stride, ubr_view dx, ubr_view y_pre,
                              const float* scale, const float* shift, const float* mean, const float* invstd, int relu,
                              void* bn_workspace, int* rows, void* stream);
/* ubr_conv2d_dgrad_bnreduce whose launch also finishes the two sums (kept behind the partial rows of bn_workspace, where
 * the apply passes read them) and writes dbeta / dgamma (common_layers.py:41-44 reversed): follow it with
 * ubr_bn_act_bwd_apply_final.  `tree` as for ubr_conv2d_fwd_bn.  Returns 1 (nothing launched) without a fused path. */
int ubr_conv2d_dgrad_bnreduce_final(ubr_view dy, const void* wpk, int in_ch, int ksize, int stride, ubr_view dx,
                                    ubr_view y_pre, const float* scale, const float* shift, const float* mean,
                                    const float* invstd, int relu, void* bn_workspace, void* tree, float* dgamma,
                                    float* dbeta, int accumulate, void* stream);
/* weight gradient: dw[o,i,r,s] (+)= sum_pixels dy[.,o] * x[shifted,i]; fp32 OIHW.  accumulate: 0 overwrite, 1 add */
int ubr_conv2d_wgrad(ubr_view x, ubr_view dy, int out_ch /* true rows of dw; dy.c may be padded */, int ksize,
                     int stride, float* dw, int accumulate, void* workspace, uint64_t workspace_bytes, void* stream);
/* split-K partial buffer size.  conv: grid = output (n,h,w), m_ch = padded out_ch, n_ch = in_ch, ntaps = k*k;
 * deconv: grid = input (n,h,w), m_ch = in_ch, n_ch = out_ch, ntaps = 16 */
uint64_t ubr_wgrad_workspace_bytes(int grid_n, int grid_h, int grid_w, int m_ch, int n_ch, int ntaps);

/* nn.Conv2d(k=3, stride=1, padding=d, dilation=d): the atrous branches B2/B3/B4 of ASPP (reference
 * models/ASPP_ResNet.py:203-214; d = 1, 3, 5).  Same packed operands as ubr_conv2d_{fwd,dgrad,wgrad} with ksize 3.
 * The wgrad workspace is sized with ubr_wgrad_workspace_bytes(n, h, w, pad16(out_ch), in_ch, 9). */
int ubr_conv2d_dilated_fwd(ubr_view x, const void* wpk, const float* bias, int out_ch, int dilation, ubr_view y,
                           void* stream);
/* y = relu(conv(x) + bias): inference form of an ASPP atrous branch with its BatchNorm folded into (wpk, bias) */
int ubr_conv2d_dilated_fwd_relu(ubr_view x, const void* wpk, const float* bias, int out_ch, int dilation, ubr_view y,
                                void* stream);
int ubr_conv2d_dilated_dgrad(ubr_view dy, const void* wpk, int in_ch, int dilation, ubr_view dx, void* stream);
int ubr_conv2d_dilated_wgrad(ubr_view x, ubr_view dy, int out_ch, int dilation, float* dw, int accumulate,
                             void* workspace, uint64_t workspace_bytes, void* stream);

/* nn.ConvTranspose2d(k=4, stride=2, padding=1, bias=False), output_size == 2x input (common_layers.py:125,128) */
int ubr_deconv4s2_fwd(ubr_view x, const void* wpk, int out_ch, ubr_view y, void* stream);
/* same forward as a 3x3 "sub-pixel" conv over the input grid with a pixel-shuffle epilogue (halo-resident kernel);
 * wpk3 = ubr_pack_weights(src_layout=2, n_is_out=1, kh=kw=3, n_pad=4*out_ch).  Returns 1 when the shape has no such
 * path (in_ch > 64 or W %% 128 != 0): use ubr_deconv4s2_fwd. */
int ubr_deconv4s2_fwd_subpixel(ubr_view x, const void* wpk3, int out_ch, ubr_view y, void* stream);
int ubr_deconv4s2_dgrad(ubr_view dy, const void* wpk, int in_ch, ubr_view dx, void* stream);
/* backward of the same layer on the sub-pixel formulation (high-resolution decoder stages, 4*out_ch <= 64):
 *   dy4 = ubr_pixel_unshuffle2(dy)                      [n, h, w, 4*out_ch] on the input grid
 *   dx  = ubr_conv2d_fwd(dy4, wpk3d, NULL, in_ch, 3, 1) wpk3d = ubr_pack_weights(src_layout=2, n_is_out=0, 3, 3)
 *   dw3 = ubr_conv2d_wgrad(x, dy4, 4*out_ch, 3, 1)      fp32 [4*out_ch][in_ch][3][3]
 *   dw  = ubr_deconv4s2_wgrad_gather(dw3)               IOHW [in_ch][out_ch][4][4] */
int ubr_pixel_unshuffle2(ubr_view y, ubr_view y4, void* stream);
int ubr_deconv4s2_wgrad_gather(const float* dw3, int in_ch, int out_ch, float* dw, int accumulate, void* stream);
int ubr_deconv4s2_wgrad(ubr_view x, ubr_view dy, float* dw /* IOHW */, int accumulate,
                        void* workspace, uint64_t workspace_bytes, void* stream);

/* ---- stem: nn.Conv2d(1, 16*m, 7, 1, 3, bias=True) on the fp32 NCHW input (ub_uresnet.py:41,94) -- */
int ubr_stem_conv_fwd(const float* x /* [n,1,h,w] */, int n, int h, int w, const float* wgt /* [oc,1,7,7] */,
                      const float* bias, int out_ch, ubr_view y, void* stream);
/* y = relu(conv(x) + bias): the inference form with bn1 folded into (wgt, bias) by the caller (ub_uresnet.py:94-96).
 * Returns 1 (nothing launched) for shapes without a fused path: the caller runs ubr_stem_conv_fwd + ubr_bn_act. */
int ubr_stem_conv_fwd_relu(const float* x, int n, int h, int w, const float* wgt, const float* bias, int out_ch,
                           ubr_view y, void* stream);
int ubr_stem_conv_wgrad(const float* x, int n, int h, int w, ubr_view dy, float* dw, float* dbias,
                        int accumulate, void* workspace, uint64_t workspace_bytes, void* stream);
uint64_t ubr_stem_wgrad_workspace_bytes(int n, int h, int w, int out_ch);

/* ---- BatchNorm2d (ub_uresnet.py:42,61; common_layers.py:24,27,35) ------------------------------
 * training statistics over (n,h,w): deterministic two-level reduction.
 * stats kernel writes per-block partials, finalize produces the per-channel affine
 *   scale = gamma*invstd, shift = beta - mean*scale, and updates running stats
 *   (momentum 0.1, unbiased variance for running_var, eps inside sqrt). */
uint64_t ubr_bn_workspace_bytes(int c);
int ubr_bn_stats(ubr_view y, void* workspace, void* stream);
int ubr_bn_finalize(const void* workspace, int c, double count, const float* gamma, const float* beta,
                    float* running_mean, float* running_var, float momentum, float eps,
                    float* scale, float* shift, float* mean, float* invstd, void* stream);
/* same, for `rows` partial rows written by a fused producer (ubr_conv2d_fwd_stats) */
int ubr_bn_finalize_rows(const void* workspace, int rows, int c, double count, const float* gamma, const float* beta,
                         float* running_mean, float* running_var, float momentum, float eps,
                         float* scale, float* shift, float* mean, float* invstd, void* stream);
/* ubr_bn_stats + ubr_bn_finalize in ONE launch (nn.BatchNorm2d in training mode, common_layers.py:24,27,35): the last
 * CTAs to arrive reduce the partial rows in a fixed order (`tree`: ubr_bn_tree_bytes(c) bytes, zero-filled once after
 * allocation, see ubr_conv2d_fwd_bn).  The *_final reduce entry points return 1 (nothing launched) for channel counts whose
 * totals do not fit the tail's shared memory (c > 1024 / 2048): run the two-launch form instead. */
uint64_t ubr_bn_tree_bytes(int c);
int ubr_bn_stats_final(ubr_view y, void* workspace, void* tree, const float* gamma, const float* beta, float* running_mean,
                       float* running_var, float momentum, float eps, float* scale, float* shift, float* mean,
                       float* invstd, void* stream);
/* eval mode: scale/shift from running statistics */
int ubr_bn_eval_affine(int c, const float* gamma, const float* beta, const float* running_mean,
                       const float* running_var, float eps, float* scale, float* shift, void* stream);
/* out = relu?(y*scale+shift) */
int ubr_bn_act(ubr_view y, const float* scale, const float* shift, int relu, ubr_view out, void* stream);
/* residual block tail (common_layers.py:45-56): out = relu( relu(y2*s2+b2) + (res*sr+br | res) ) */
int ubr_block_tail(ubr_view y2, const float* s2, const float* b2, ubr_view res, const float* sr, const float* br,
                   ubr_view out, void* stream);

/* backward of out = relu(bn(y)):  given dout and y(+affine), two passes.
 *  pass 1 (reduce): sums[0:c) = sum dz, sums[c:2c) = sum dz*xhat       (dz = dout * (bn(y)>0 or !relu))
 *  pass 2 (apply) : dy = gamma*invstd*(dz - sum_dz/count - xhat*sum_dzx/count); dgamma=sum_dzx; dbeta=sum_dz */
int ubr_bn_act_bwd_reduce(ubr_view dout, ubr_view y, const float* scale, const float* shift, const float* mean,
                          const float* invstd, int relu, void* workspace, void* stream);
int ubr_bn_act_bwd_apply(ubr_view dout, ubr_view y, const float* scale, const float* shift, const float* mean,
                         const float* invstd, const float* gamma, int relu, const void* workspace, double count,
                         ubr_view dy, float* dgamma, float* dbeta, int accumulate, void* stream);
/* pass 2 for `rows` partial rows written by a fused producer (ubr_conv2d_dgrad_bnreduce) */
int ubr_bn_act_bwd_apply_rows(ubr_view dout, ubr_view y, const float* scale, const float* shift, const float* mean,
                              const float* invstd, const float* gamma, int relu, const void* workspace, int rows,
                              double count, ubr_view dy, float* dgamma, float* dbeta, int accumulate, void* stream);
/* the same two passes with the reduction FINISHED inside pass 1 (sums and dgamma / dbeta written by the last CTA to
 * arrive; `tree` as for ubr_conv2d_fwd_bn), so pass 2 is a single launch */
int ubr_bn_act_bwd_reduce_final(ubr_view dout, ubr_view y, const float* scale, const float* shift, const float* mean,
                                const float* invstd, int relu, void* workspace, void* tree, float* dgamma, float* dbeta,
                                int accumulate, void* stream);
int ubr_bn_act_bwd_apply_final(ubr_view dout, ubr_view y, const float* scale, const float* shift, const float* mean,
                               const float* invstd, int relu, const void* workspace, double count, ubr_view dy,
                               void* stream);
/* backward of the block tail.  du = dout*(out>0); main branch dz2 = du*(bn2(y2)>0); residual branch gets du.
 * res_is_bn=1: residual is bnpass(bypass conv) -> dres = BN backward of du wrt `res` (pre-BN conv output); the mask
 *              (out>0) is recomputed from y2 and res with the forward arithmetic and the residual affine
 *              (scaler, shiftr) -- `out` is not read and may be a null view
 * res_is_bn=0: identity residual -> dres = du (scaler / shiftr unused) */
int ubr_block_tail_bwd_reduce(ubr_view dout, ubr_view out, ubr_view y2, const float* s2, const float* b2,
                              const float* mean2, const float* invstd2, ubr_view res, const float* meanr,
                              const float* invstdr, const float* scaler, const float* shiftr, int res_is_bn,
                              void* workspace, void* stream);
int ubr_block_tail_bwd_apply(ubr_view dout, ubr_view out, ubr_view y2, const float* s2, const float* b2,
                             const float* mean2, const float* invstd2, const float* gamma2,
                             ubr_view res, const float* meanr, const float* invstdr, const float* gammar,
                             const float* scaler, const float* shiftr,
                             int res_is_bn, const void* workspace, double count, ubr_view dy2, ubr_view dres,
                             float* dgamma2, float* dbeta2, float* dgammar, float* dbetar, int accumulate,
                             void* stream);

/* block-tail backward with the reduction finished inside pass 1 (see ubr_bn_act_bwd_reduce_final) */
int ubr_block_tail_bwd_reduce_final(ubr_view dout, ubr_view out, ubr_view y2, const float* s2, const float* b2,
                                    const float* mean2, const float* invstd2, ubr_view res, const float* meanr,
                                    const float* invstdr, const float* scaler, const float* shiftr, int res_is_bn,
                                    void* workspace, void* tree, float* dgamma2, float* dbeta2, float* dgammar,
                                    float* dbetar, int accumulate, void* stream);
int ubr_block_tail_bwd_apply_final(ubr_view dout, ubr_view out, ubr_view y2, const float* s2, const float* b2,
                                   const float* mean2, const float* invstd2, ubr_view res, const float* meanr,
                                   const float* invstdr, const float* gammar, const float* scaler, const float* shiftr,
                                   int res_is_bn, const void* workspace, double count, ubr_view dy2, ubr_view dres,
                                   void* stream);

/* ---- MaxPool2d(3, stride 2, pad 1) (ub_uresnet.py:44,97) and pointwise sums ---------------------- */
int ubr_maxpool3s2_fwd(ubr_view x, ubr_view y, void* stream);
int ubr_maxpool3s2_bwd(ubr_view x, ubr_view dy, ubr_view dx_accum, int accumulate, void* stream);
/* same pooling, recording the argmax tap (kh*3+kw, first maximum in scan order) per output element in
 * idx[n][h/2][w/2][c] (uint8) so the backward is a pure gather (what ATen/cuDNN keep as `indices`) */
int ubr_maxpool3s2_fwd_idx(ubr_view x, ubr_view y, uint8_t* idx, void* stream);
int ubr_maxpool3s2_bwd_idx(const uint8_t* idx, ubr_view dy, ubr_view dx_accum, int accumulate, void* stream);
/* nn.MaxPool2d(3, stride=1, padding=1): ASPP block 5 (models/ASPP_ResNet.py:217-219).  idx (optional, uint8
 * [n][h][w][c]) records the argmax tap for ubr_maxpool3s1_bwd_idx (dx (+)= gathered dy). */
int ubr_maxpool3s1_fwd(ubr_view x, ubr_view y, uint8_t* idx, void* stream);
int ubr_maxpool3s1_bwd_idx(const uint8_t* idx, ubr_view dy, ubr_view dx_accum, int accumulate, void* stream);
/* out = a + b (+ c); b/c may have ptr == NULL */
int ubr_add(ubr_view a, ubr_view b, ubr_view c, ubr_view out, void* stream);

/* ---- LogSoftmax(dim=1) + PixelWiseNLLLoss (ub_uresnet.py:70,143; pixelwise_nllloss.py:41-61) ----
 * logits: fp32 [pixel][ld] (first ncls columns).  logp: fp32 NCHW [n,ncls,h,w] (the model output). */
int ubr_logsoftmax_fwd(const float* logits, int ld, int n, int ncls, int hw, float* logp, void* stream);
/* dlogits (bf16 view, c >= ncls; pad channels zeroed) = dlogp - softmax * sum_k dlogp_k */
int ubr_logsoftmax_bwd(const float* logp, const float* dlogp, int n, int ncls, int hw, ubr_view dlogits, void* stream);
/* loss = mean_{pixels}( pw * cw[t] * -logp[t] ), t == ignore_index contributes 0 (stays in the mean) */
uint64_t ubr_nll_workspace_bytes(int n, int hw);
int ubr_pixelwise_nll_fwd(const float* logp, const int64_t* target, const float* pixw, const float* classw,
                          int n, int ncls, int hw, int64_t ignore_index, void* workspace, float* loss, void* stream);
/* dlogp[n,t,hw] = -gout * pw*cw[t]/(n*hw) at the target class, 0 elsewhere; gout read from device */
int ubr_pixelwise_nll_bwd(const int64_t* target, const float* pixw, const float* classw, const float* gout,
                          int n, int ncls, int hw, int64_t ignore_index, float* dlogp, void* stream);

/* ---- whole-view tiling (deploy/run_ubresnet_wholeview.py:238-239; larcv Image2D.cxx:362-367) ------
 * origins: int32 [ntiles][2] (row0, col0) on the device.
 * gather : crops[t,0,:,:] = plane[row0:row0+crop, col0:col0+crop]
 * scatter: out[:, r, c] = value from the HIGHEST-index tile covering (r,c)  (== sequential overwrite) */
int ubr_tile_gather(const float* plane, int rows, int cols, const int32_t* origins, int ntiles, int crop,
                    float* crops, void* stream);
int ubr_stitch_scatter(const float* tiles /* [ntiles,c,crop,crop] */, int nch, const int32_t* origins, int ntiles,
                       int crop, int rows, int cols, float* out /* [c,rows,cols] */, void* stream);
/* rectangular crops (crop_h x crop_w): the 512 x 832 boxes of the UBSplitDetector configuration
 * (deploy/run_ubresnet_wholeview.py:33-47).  Gather zero-fills the part of a box that lies outside the plane. */
int ubr_tile_gather_rect(const float* plane, int rows, int cols, const int32_t* origins, int ntiles, int crop_h, int crop_w,
                         float* crops, void* stream);
int ubr_stitch_scatter_rect(const float* tiles, int nch, const int32_t* origins, int ntiles, int crop_h, int crop_w, int rows,
                            int cols, float* out, void* stream);
/* class maps for the deploy loop (what deploy/run_ubresnet_wholeview.py:264-272 keeps of the network output):
 * classes[n,hw] = argmax_k logp[n,k,hw] (first maximum), and the stitch of uint8 class TILES [ntiles,crop,crop] with
 * the same highest-tile-wins rule as ubr_stitch_scatter (pixels no tile covers get `fill`).  Crop-sharded multi-GPU
 * deploy exchanges these 1-byte tiles instead of C fp32 maps. */
int ubr_argmax_classes(const float* logp, int n, int ncls, int hw, uint8_t* classes, void* stream);
int ubr_stitch_scatter_u8(const uint8_t* tiles, const int32_t* origins, int ntiles, int crop_h, int crop_w, int rows, int cols,
                          int fill, uint8_t* out, void* stream);

/* ---- accuracy counters: replaces the 2*C `.item()` reductions of `accuracy()` (reference
 * training/train_ubresnet2018_wlarcv2.py:509-566).  pred = argmax over classes of logp (NCHW fp32, first maximum);
 * counts[c] = #pixels with target == c, counts[ncls + c] = #of those predicted correctly.  counts (2*ncls uint64,
 * device) is zeroed by the call.  Targets outside [0, ncls) are not counted. */
int ubr_accuracy_counts(const float* logp, const int64_t* target, int n, int ncls, int hw, uint64_t* counts, void* stream);

/* ---- multi-tensor optimizer step (train_ubresnet2018_wlarcv1.py:127-129 SGD,
 * train_ubresnet2018_wlarcv2.py:155-157 Adam).  One launch updates every parameter: `slots` (device) describes the
 * tensors, `chunks` (device) cuts them into ubr_opt_chunk_elems()-element pieces, one CTA each.  Arithmetic is
 * torch.optim's: g' = g + wd*p;  SGD: buf = first_step ? g' : m*buf + (1-dampening)*g', p -= lr*(nesterov ? g' + m*buf
 * : buf) (state1 = momentum_buffer, NULL when momentum == 0);  Adam: exp_avg (state1), exp_avg_sq (state2), bias
 * correction for `step` >= 1, no amsgrad. */
typedef struct {
  float* param;
  const float* grad;
  float* state1;
  float* state2;
  int64_t numel;
  int32_t vec_ok;            /* all four pointers 16-byte aligned */
  int32_t pad_;
} ubr_opt_slot;
typedef struct {
  int32_t slot;
  int32_t pad_;
  int64_t offset;            /* first element of the chunk inside its tensor */
} ubr_opt_chunk;
int ubr_opt_chunk_elems(void);
int ubr_sgd_step(const ubr_opt_slot* slots, const ubr_opt_chunk* chunks, int nchunks, float lr, float momentum,
                 float dampening, float weight_decay, int nesterov, int first_step, void* stream);
/* Adam hyper-parameters are doubles: torch derives 1-beta and lr/(1-beta1^t) in double before rounding to fp32 */
int ubr_adam_step(const ubr_opt_slot* slots, const ubr_opt_chunk* chunks, int nchunks, double lr, double beta1,
                  double beta2, double eps, double weight_decay, int step, void* stream);

/* ---- fp32 precision mode (BASELINE.json north_star "1e-4 (tf32/fp32 mode)"; SURVEY 8b "Precision modes") ----------
 * The reference computes in fp32 end to end (models/ub_uresnet.py:41-70, models/common_layers.py:13-58).  These entry
 * points keep fp32 NHWC activations (`ubr_view` with fp32 elements, ld in elements), read the fp32 parameters in the
 * reference layouts directly (OIHW / ConvTranspose IOHW, no packing) and accumulate in fp32 FMA (reductions over pixels
 * in float64).  Parity mode, selected with UBR_PRECISION=fp32 / UResNet(..., precision="fp32"): CUDA cores, not tensor
 * cores -- tf32's 11 significant bits cannot hold 1e-4 through 58 convolutions.
 *   fwd   : y[n,oh,ow,o]  = bias[o] + sum x[n, oh*stride + r*dil - pad, ow*stride + s*dil - pad, i] * W[o,i,r,s]
 *   dgrad : dx            = transposed convolution of dy with the same W (accumulate: dx += ...)
 *   wgrad : dW[o,i,r,s] (+)= sum_pixels dy[.,o] * x[shifted,i];  workspace >= 8 * numel(dW) bytes (float64 accumulators)
 *   deconv: nn.ConvTranspose2d(k, stride, pad), W is IOHW [in_ch][out_ch][k][k] (common_layers.py:125) */
int ubr_f32_conv2d_fwd(ubr_view x, const float* w, const float* bias, int out_ch, int ksize, int stride, int pad,
                       int dilation, ubr_view y, void* stream);
int ubr_f32_conv2d_dgrad(ubr_view dy, const float* w, int in_ch, int ksize, int stride, int pad, int dilation, ubr_view dx,
                         int accumulate, void* stream);
int ubr_f32_conv2d_wgrad(ubr_view x, ubr_view dy, int ksize, int stride, int pad, int dilation, float* dw, int accumulate,
                         void* workspace, uint64_t workspace_bytes, void* stream);
int ubr_f32_deconv_fwd(ubr_view x, const float* w, int out_ch, int ksize, int stride, int pad, ubr_view y, void* stream);
int ubr_f32_deconv_dgrad(ubr_view dy, const float* w, int in_ch, int ksize, int stride, int pad, ubr_view dx, void* stream);
int ubr_f32_deconv_wgrad(ubr_view x, ubr_view dy, int ksize, int stride, int pad, float* dw, int accumulate, void* workspace,
                         uint64_t workspace_bytes, void* stream);
/* BatchNorm2d training statistics: sums[0:c) = sum y, sums[c:2c) = sum y^2 (float64, zeroed by the call) */
int ubr_f32_bn_stats(ubr_view y, double* sums, void* stream);
int ubr_f32_bn_finalize(const double* sums, int c, double count, const float* gamma, const float* beta, float* running_mean,
                        float* running_var, float momentum, float eps, float* scale, float* shift, float* mean,
                        float* invstd, void* stream);
/* out[c] (+)= sum over pixels of a[.,c] (conv bias gradients); workspace: 2*c doubles */
int ubr_f32_channel_sum(ubr_view a, double* workspace, float* out, int accumulate, void* stream);
int ubr_f32_bn_act(ubr_view y, const float* scale, const float* shift, int relu, ubr_view out, void* stream);
int ubr_f32_block_tail(ubr_view y2, const float* s2, const float* b2, ubr_view res, const float* sr, const float* br,
                       ubr_view out, void* stream);
/* du = dout * (out > 0): the block's final ReLU (common_layers.py:56) */
int ubr_f32_relu_mask(ubr_view dout, ubr_view out, ubr_view du, void* stream);
/* backward of out = relu?(bn(y)) in training mode: dy, dgamma = sum dz*xhat, dbeta = sum dz; workspace: 2*c doubles */
int ubr_f32_bn_act_bwd(ubr_view dout, ubr_view y, const float* scale, const float* shift, const float* mean,
                       const float* invstd, int relu, double* workspace, double count, ubr_view dy, float* dgamma,
                       float* dbeta, int accumulate, void* stream);
int ubr_f32_maxpool3s2_fwd(ubr_view x, ubr_view y, uint8_t* idx /* may be NULL */, void* stream);
int ubr_f32_maxpool3s2_bwd(const uint8_t* idx, ubr_view dy, ubr_view dx, int accumulate, void* stream);
int ubr_f32_add(ubr_view a, ubr_view b, ubr_view c, ubr_view out, void* stream);
int ubr_f32_logsoftmax_bwd(const float* logp, const float* dlogp, int n, int ncls, int hw, ubr_view dlogits, void* stream);

/* ---- device-side batch assembly: the reference's C++ batch fillers as kernels ----------------------
 * src_dtype: 0 fp32, 1 fp16, 2 uint8 (what travels over PCIe); larcv Image2D pixel order is COLUMN-major
 * (index = col*rows + row).  mirror: optional per-sample flags (device uint8[batch]) -- the column mirror of
 * BatchFillerImage2D.cxx:102-112 / SegFiller.cxx `EnableMirror`.
 *   ubr_fill_image_batch : out[b,0,row,col] (fp32, the network input) = src[b][(mirror ? cols-1-col : col)*rows + row]
 *                          (larcv/larcv/app/ThreadIO/BatchFillerImage2D.cxx:217-256)
 *   ubr_fill_label_batch : out[b,row,col] (int64 nll target) = lut[(size_t)src[...]] + shift; lut = the roitype->class
 *                          table of SegFiller.cxx:49-63 (0xFF = invalid), shift = -1 in the training scripts
 *                          (grid_scripts/train_ubresnet_wlarcv1_tuftsgrid.py:132); *invalid_count (device int) receives
 *                          the number of pixels with an invalid ROI type (the host code throws larbys on the first)
 *   ubr_seg_weights      : SegWeightTrackShower (caffe/larcv/app/ImageMod/SegWeightTrackShower.cxx:109-272) without the
 *                          optional keypoint term: cat_lut maps a label code to 0 none / 1 shower / 2 track;
 *                          w = pool(weight_in | 1.0, {(int)(N/n_shower), (int)(N/n_track), (int)(N/n_surrounding)}),
 *                          clipped at weight_max.  pool_type: 0 sum, 1 max, 2 average, 3 overwrite (SegWeightTrackShower.h
 *                          PoolType_t).  label_dtype: 0 fp32, 2 uint8, 3 int64; column_major describes the label only,
 *                          weight_in (may be NULL) and weight_out are row-major fp32 [batch][rows][cols]. */
int ubr_fill_image_batch(const void* src, int src_dtype, int batch, int rows, int cols, const uint8_t* mirror, float* out,
                         void* stream);
int ubr_fill_label_batch(const void* src, int src_dtype, int batch, int rows, int cols, const uint8_t* mirror,
                         const uint8_t* lut, int lut_size, int shift, int64_t* out, uint8_t* out_class_u8 /* optional:
                         the pre-shift class per pixel, row-major, 1 byte -- the cheap input of ubr_seg_weights */,
                         int* invalid_count, void* stream);
uint64_t ubr_seg_weights_workspace_bytes(int batch, int rows, int cols);
int ubr_seg_weights(const void* label, int label_dtype, int column_major, int batch, int rows, int cols,
                    const uint8_t* cat_lut, int lut_size, int dist_surrounding, int weight_surrounding, int pool_type,
                    float weight_max, const float* weight_in, float* weight_out, void* workspace, uint64_t workspace_bytes,
                    void* stream);

/* out[0:nvalid) (+)= sum over pixels of a[., 0:nvalid): the bias gradient of a biased conv (ub_uresnet.py:41,60,64);
 * workspace: ubr_bn_workspace_bytes(a.c).  ubr_increment_i64: *ptrs[i] += 1 for a DEVICE table of int64 scalars
 * (BatchNorm2d.num_batches_tracked of all 52 layers in one launch). */
int ubr_channel_sum(ubr_view a, int nvalid, void* workspace, float* out, int accumulate, void* stream);
/* Allocation-free variant of ubr_channel_sum.  It has NO counterpart in the reference: it was added
 * for callers that cannot allocate a workspace (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  The partial sums go to a file-scope device array inside the library
 * (2.3 MiB per GPU that loads it, held until exit) instead of a caller workspace; the call itself performs no
 * cudaMalloc / cudaFree.  Limit: a.c <= 512 channels, larger views return UBR_E_BADARG.  All calls share the array:
 * calls that overlap on the GPU (different streams) race -- keep them on one stream or order them with events. */
int ubr_channel_sum_static(ubr_view a, int nvalid, float* out, int accumulate, void* stream);
int ubr_increment_i64(int64_t* const* ptrs_device, int n, void* stream);

/* ---- flat-buffer helpers for the data-parallel gradient buckets --------------------------------- */
int ubr_scale_f32(float* buf, uint64_t count, float alpha, void* stream);

#ifdef __cplusplus
}
#endif
#endif

/* dtf_b200 — C ABI of the native sm_100a kernel library (libdtfb200.so).
 *
 * Every entry point takes raw device pointers plus the CUDA stream to launch on (as a
 * void* holding a cudaStream_t) and returns 0 on success or a non-zero cudaError_t /
 * negative library error.  Nothing here allocates device memory: callers (the Python
 * package, through the PyTorch caching allocator) own every buffer, which is what lets a
 * whole training step be captured into one CUDA graph.
 *
 * Tensor conventions: activations are NHWC fp32; convolution / FC weights are
 * "OHWI" = [C_out][R][S][C_in] fp32 (K-major for the tensor core); the checkpoint layer
 * converts to/from TensorFlow's HWIO.  Citations name the reference op each kernel replaces
 * (paths relative to the reference repository).
 */
#ifndef DTF_B200_H_
#define DTF_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------ */
/* library                                                                              */
/* ------------------------------------------------------------------------------------ */
int dtf_version(void);
const char* dtf_last_error(void);
/* Debug aid: 0 launches the convolution kernels in plain stream order; 1 (default) with programmatic dependent
 * launch, so that a kernel's prologue overlaps the tail of its predecessor. */
int dtf_debug_set_pdl(int enabled);
/* Keep n SMs out of the grids of every kernel launched afterwards by this process (0 = none): the engine calls
 * it around the backward pass when the parameter-server push runs beside it on SMs of its own.  Returns the previous
 * value.  No counterpart in the reference. */
int dtf_reserve_sms(int n);
/* Host-only CRC-32C (Castagnoli) used by the checkpoint-bundle and TFRecord writers: crc of `n` bytes, continuing
 * from `crc` (0 for a fresh checksum).  No GPU involved; no counterpart in the reference (TensorFlow's C++ runtime
 * provides it there). */
uint32_t dtf_crc32c(const void* data, unsigned long long n, uint32_t crc);

/* ------------------------------------------------------------------------------------ */
/* K1  implicit-GEMM convolution on tcgen05 (kind::tf32, fp32 accumulate in TMEM)         */
/*     replaces tf.nn.conv2d / slim.conv2d / tf.layers.conv2d and the FC matmuls         */
/*     (cnn/cifar10.py:221, inception/slim/ops.py:232,305, resnet/resnet_model.py:87)    */
/* ------------------------------------------------------------------------------------ */

/* Gather-GEMM: out[m, n] = epilogue( sum_k A[m, k] * Wt[n, k] ),
 *   m = (img, oh, ow) over an N x OH x OW grid, k = (r, s, c) over R x S x srcC.
 *   A[m,k] = act(src[img, ih, iw, c]) where ih*div_h = oh*a_h + b_h + r*c_h (must divide) and
 *   0 <= ih < srcH, likewise for w; everything else reads as zero.
 *     fprop: a = stride, b = -pad, c = +1, div = 1 (src = input,  Wt = OHWI weights)
 *     dgrad: a = 1, b = +pad, c = -1, div = stride (src = dY,     Wt = [C_in][R][S][C_out])
 *   act(x) = in_relu ? max(x*in_scale[c] + in_shift[c], 0) : x*in_scale[c] + in_shift[c]
 *            when in_scale != NULL (fused batch-norm apply of the producer layer).
 *   epilogue: + bias[n], + residual[m, n], ReLU, per-channel sum / sum-of-squares of the
 *             stored value into stats[0..Nout) / stats[Nout..2Nout) (atomic adds); or, for dgrad,
 *             the batch-norm backward reductions of the layer below (see bnb_x).
 */
typedef struct {
  const float* src; int srcH, srcW, srcC, src_ld;
  int N, OH, OW;
  int R, S;
  int a_h, b_h, c_h, div_h;
  int a_w, b_w, c_w, div_w;
  const float* wt;          /* [Nout][Ktot] row-major, Ktot = R*S*srcC (multiple of 4) */
  int Nout;
  float* out; int out_ld;   /* out[m*out_ld + n] */
  const float* in_scale; const float* in_shift; int in_relu;
  const float* bias; const float* residual; int res_ld; int out_relu;
  float* stats;
  /* dgrad + fused batch-norm backward reduction: when bnb_x != NULL the kernel output is d(act) for a layer
   * whose forward input was act = relu(bnb_x*bnb_scale + bnb_shift); `stats` then receives
   * sum_m g and sum_m g*x with g = out * [act > 0] (the two reductions dtf_bn_bwd_apply needs). */
  const float* bnb_x; int bnb_ld; const float* bnb_scale; const float* bnb_shift;
  /* optional side output (fused input transform only): act(src) in the layout of src, written by the kernels that
   * read every source pixel exactly once (the sliding-window kernels).  The weight-gradient pass can then copy the
   * activated input instead of re-applying the transform.  Check dtf_conv_gemm_writes_act first. */
  float* act_out;
  /* optional deferred batch-norm finalize (sliding-window kernel only, see dtf_conv_gemm_fuses_finalize): when
   * fin_stats != NULL the kernel derives the input transform itself from the producer's per-channel sums
   *   mean = sum / count, var = sumsq / count - mean^2, scale = gamma * rsqrt(var + eps), shift = beta - mean * scale
   * and its first CTA writes scale / shift / mean / invstd to in_scale / in_shift / fin_mean / fin_invstd (for the
   * backward pass) and applies the moving-average update — one tiny kernel launch less per batch-norm. */
  const float* fin_stats; const float* fin_gamma; const float* fin_beta;
  float* fin_moving_mean; float* fin_moving_var; float* fin_mean; float* fin_invstd;
  float fin_inv_count, fin_unbias, fin_eps, fin_decay; int fin_update_moving;
  /* optional pull gate (BSP push / pull over symmetric memory, see dtf_ps_bucket_step): when gate_flags != NULL the
   * weight-load warp waits until gate_flags[0 .. gate_n) >= *gate_epoch before its first weight load — "every
   * owner's part of the bucket that holds wt has reached this replica".  The wait gives up after
   * gate_timeout_cycles SM clocks and sets *gate_status = 1. */
  const unsigned int* gate_flags; int gate_n; const unsigned int* gate_epoch; unsigned int* gate_status;
  long long gate_timeout_cycles;
} dtf_conv_gemm_params;

int dtf_conv_gemm(const dtf_conv_gemm_params* p, void* stream);

/* 1 if dtf_conv_gemm would run a kernel that fills act_out for this problem (act_out itself may be NULL), else 0. */
int dtf_conv_gemm_writes_act(const dtf_conv_gemm_params* p);
/* 1 if dtf_conv_gemm would run the kernel that honours the fin_* fields for this problem, else 0. */
int dtf_conv_gemm_fuses_finalize(const dtf_conv_gemm_params* p);

/* Weight gradient: dW[n][k] (+)= sum_m dY[m, n] * A[m, k], A gathered exactly as in fprop
 * (with the same optional fused input activation).  Split over m across CTAs; partial sums
 * are combined with red.global.add.f32, so dW must be zeroed (or hold the value to
 * accumulate into) before the call. */
typedef struct {
  const float* src; int srcH, srcW, srcC, src_ld;
  int N, OH, OW;
  int R, S;
  int stride_h, stride_w, pad_h, pad_w;
  const float* dy; int dy_ld; int Nout;
  float* dw;                /* [Nout][R*S*srcC] */
  const float* in_scale; const float* in_shift; int in_relu;
} dtf_conv_wgrad_params;

int dtf_conv_wgrad(const dtf_conv_wgrad_params* p, void* stream);

/* Debug aid: 0 keeps dtf_conv_wgrad on the register-transposing producers (K-major operands); 1 (default) lets it
 * use the copy-only producers with MN-major operands where the shape allows. */
int dtf_debug_set_wgrad_mn(int enabled);

/* Debug aid: arm (buf != NULL, >= 4096 int64) or disarm the in-kernel timeline trace of dtf_conv_gemm. */
int dtf_debug_set_trace(long long* buf);

/* Debug aid: 0 forces dtf_conv_gemm onto the software-gather producers even where the hardware im2col (TMA)
 * feed applies; 1 (default) restores the automatic choice. */
int dtf_debug_set_conv_tma(int enabled);

/* Debug aid: sliding-window convolution kernel (stride-1 stencils): 0 never, 1 automatic (default: sources of up
 * to 64 channels), 2 whenever the shape is eligible. */
int dtf_debug_set_conv_win(int mode);

/* Debug aid: channel-major sliding-window kernel (C_in, C_out in {16, 32, 64}): mode 0 off / 1 on (default),
 * UMMA M (64 or 128), and which TMEM row -> lane mapping to assume for M = 64. */
int dtf_debug_set_conv_wint(int mode, int umma_m, int lane_mode);

/* [Nout][R*S][C] -> [C][R*S][Nout]: builds the dgrad B operand from OHWI weights. */
int dtf_weight_transpose(const float* w, float* wt, int Nout, int RS, int C, void* stream);

/* The same for a whole list of layers in one launch.  `table` is a device array of n_entries records
 *   { const float* w; float* wt; int Nout, RS, C, blk0; }   (32 bytes each)
 * where blk0 is the running sum of ceil(C/32) * ceil(Nout/32) * RS over the preceding entries and total_blocks
 * that sum over all entries. */
int dtf_weight_transpose_batched(const void* table, int n_entries, int total_blocks, void* stream);

/* ------------------------------------------------------------------------------------ */
/* K7  batch norm on [M = N*H*W, C] views of NHWC tensors (C % 4 == 0)                    */
/*     replaces slim ops.batch_norm (inception/slim/ops.py:46-135) and                   */
/*     tf.layers.batch_normalization(fused=True) (resnet/resnet_model.py:41-50)          */
/* ------------------------------------------------------------------------------------ */
/* stats[0..C) += sum_m x, stats[C..2C) += sum_m x^2 (atomic; zero the buffer first) */
int dtf_bn_stats(const float* x, long long M, int C, int ld_x, float* stats, void* stream);
/* sums -> scale = gamma*invstd, shift = beta - mean*scale (+ mean, invstd for backward);
 * moving = moving*decay + batch*(1-decay) when update_moving (variance Bessel-corrected when asked) */
int dtf_bn_finalize(const float* stats, const float* gamma, const float* beta, long long count, float eps,
                    float decay, int update_moving, int unbiased_moving_var, float* moving_mean,
                    float* moving_var, float* scale, float* shift, float* mean_out, float* invstd_out, int C,
                    void* stream);
/* Training-mode forward in one call (stats -> finalize -> apply).  scratch: 2*C floats (zeroed by the call);
 * coef: 4*C floats out = scale | shift | mean | invstd (kept for the backward kernels). */
int dtf_bn_forward_train(const float* x, long long M, int C, int ld_x, const float* gamma, const float* beta,
                         float eps, float decay, int update_moving, int unbiased_moving_var, float* moving_mean,
                         float* moving_var, int relu, float* scratch, float* coef, float* y, int ld_y, void* stream);
/* Same operation for callers that may not allocate device memory (stream-only plug-in callbacks, processes
 * whose memory pool is already fully committed, loops that cannot afford cudaMalloc's device-wide sync): the
 * 2*C floats of per-channel sums live in a static __device__ array inside the library (32 KiB per GPU, held
 * until exit) instead of a caller buffer.  C > DTF_BN_STATIC_MAX_CHANNELS is rejected with an error.  The
 * first call on a device resolves the array's address — issue it before any stream capture.  All calls share
 * the array (each clears what it uses first), so calls that can overlap on the GPU must be serialised.
 * This entry point has no counterpart in the reference. */
#define DTF_BN_STATIC_MAX_CHANNELS 4096
int dtf_bn_forward_train_static(const float* x, long long M, int C, int ld_x, const float* gamma, const float* beta,
                                float eps, float decay, int update_moving, int unbiased_moving_var,
                                float* moving_mean, float* moving_var, int relu, float* coef, float* y, int ld_y,
                                void* stream);
int dtf_bn_eval_affine(const float* gamma, const float* beta, const float* moving_mean, const float* moving_var,
                       float eps, float* scale, float* shift, int C, void* stream);
/* y = act(x*scale + shift) */
int dtf_affine_act(const float* x, const float* scale, const float* shift, int relu, long long M, int C, int ld_x,
                   int ld_y, float* y, void* stream);
/* sums[0..C) += sum g, sums[C..2C) += sum g*x with g = dy * [x*scale+shift > 0] (when relu) */
int dtf_bn_bwd_reduce(const float* dy, const float* x, const float* scale, const float* shift, int relu,
                      long long M, int C, int ld_x, int ld_dy, float* sums, void* stream);
/* dx (+= add) from the reduced sums; dgamma / dbeta (either may be NULL) are ACCUMULATED into */
int dtf_bn_bwd_apply(const float* dy, const float* x, const float* scale, const float* shift, const float* mean,
                     const float* invstd, const float* sums, int relu, long long M, int C, int ld_x, int ld_dy,
                     int ld_dx, const float* add, int ld_add, float* dx, float* dgamma, float* dbeta, void* stream);
/* frozen-statistics backward: dx = scale * dy * mask */
int dtf_affine_act_bwd(const float* dy, const float* x, const float* scale, const float* shift, int relu,
                       long long M, int C, int ld_x, int ld_dy, int ld_dx, float* dx, void* stream);

/* ------------------------------------------------------------------------------------ */
/* K4/K5  pooling (NHWC, C % 4 == 0); tf.nn.max_pool / avg_pool semantics                 */
/*     (cnn/cifar10.py:228, inception/slim/ops.py:370,400, resnet/resnet_model.py:271)   */
/* ------------------------------------------------------------------------------------ */
typedef struct {
  int N, H, W, C, OH, OW, kh, kw, sh, sw, ph, pw, ld_in, ld_out;
} dtf_pool_params;
int dtf_max_pool_fwd(const dtf_pool_params* p, const float* x, float* y, unsigned char* argmax, void* stream);
int dtf_max_pool_bwd(const dtf_pool_params* p, const float* dy, int ld_dy, const unsigned char* argmax, float* dx,
                     int ld_dx, void* stream);
int dtf_avg_pool_fwd(const dtf_pool_params* p, const float* x, float* y, void* stream);
int dtf_avg_pool_bwd(const dtf_pool_params* p, const float* dy, int ld_dy, float* dx, int ld_dx, void* stream);

/* ------------------------------------------------------------------------------------ */
/* K6 LRN, K9 softmax-xent, K10 dropout, K2 bias gradient                                 */
/*     (cnn/cifar10.py:231,299; inception/slim/losses.py:142-174; slim/ops.py:422)       */
/* ------------------------------------------------------------------------------------ */
int dtf_lrn_fwd(const float* x, float* y, long long M, int C, int radius, float bias, float alpha, float beta,
                void* stream);
int dtf_lrn_bwd(const float* x, const float* dy, float* dx, long long M, int C, int radius, float bias,
                float alpha, float beta, void* stream);
/* loss[0] = weight * mean_n xent(smoothed target); dlogits = weight/N * (softmax - target) (may be NULL) */
int dtf_softmax_xent(const float* logits, const long long* labels, int N, int C, float smoothing, float weight,
                     float* loss, float* dlogits, float* row_loss, void* stream);
/* y = x * keep(i) / keep_prob; the mask is a pure function of (*seed_ptr, salt, i) */
int dtf_dropout(const float* x, float* y, long long n, float keep_prob, const unsigned long long* seed_ptr,
                unsigned long long salt, void* stream);
/* out[c] += sum_m x[m*ld + c] */
int dtf_colsum(const float* x, long long M, int C, int ld, float* out, void* stream);

/* ------------------------------------------------------------------------------------ */
/* Zoo op kinds: depthwise conv, relu6 / leaky-relu / tanh / sigmoid,        */
/*     instance norm, resize, spatial padding — dense NHWC fp32 tensors                   */
/*     (vgg/nets/mobilenet_v1.py:260-277, mobilenet/conv_blocks.py:120-140,               */
/*      nasnet/nasnet_utils.py:206-240, dcgan.py:89, pix2pix.py:172,220,                  */
/*      cyclegan.py:47-60,96-103,201-261)                                                 */
/* ------------------------------------------------------------------------------------ */
typedef struct dtf_dwconv_params {
  int N, H, W, C, mult;     /* input [N][H][W][C]; output channel c*mult + m          */
  int R, S, sh, sw, ph, pw; /* filter [R][S][C][mult], stride, padding before         */
  int OH, OW;               /* output spatial size (caller-resolved SAME / VALID)     */
} dtf_dwconv_params;
int dtf_dwconv_fwd(const dtf_dwconv_params* p, const float* x, const float* w, const float* bias /* or NULL */,
                   float* y, void* stream);
int dtf_dwconv_dgrad(const dtf_dwconv_params* p, const float* dy, const float* w, float* dx, void* stream);
/* dw is ACCUMULATED into (the caller zeroes it) */
int dtf_dwconv_wgrad(const dtf_dwconv_params* p, const float* x, const float* dy, float* dw, void* stream);
/* mode: 0 relu6, 1 leaky-relu(alpha), 2 tanh, 3 sigmoid.  Backward takes ref = forward input (modes 0, 1) or
 * forward output (modes 2, 3). */
int dtf_act_fwd(const float* x, float* y, long long n, int mode, float alpha, void* stream);
int dtf_act_bwd(const float* ref, const float* dy, float* dx, long long n, int mode, float alpha, void* stream);
/* Instance norm over H*W per (image, channel); sums[N][2][C] receives (sum x, sum x^2) and is what the backward
 * call needs as fwd_sums; bwd_sums[N][2][C] receives (sum dy, sum dy*xhat): dbeta / dgamma are their sums over N.
 * gamma / beta may be NULL. */
int dtf_instnorm_fwd(const float* x, const float* gamma, const float* beta, float* y, float* sums, int N, int HW,
                     int C, float eps, void* stream);
int dtf_instnorm_bwd(const float* x, const float* dy, const float* gamma, const float* fwd_sums, float* bwd_sums,
                     float* dx, int N, int HW, int C, float eps, void* stream);
/* TF1 resize (src = dst * in/out, or (in-1)/(out-1) with align_corners); bilinear = 0 -> nearest neighbour */
int dtf_resize_fwd(const float* x, float* y, int N, int H, int W, int C, int OH, int OW, int bilinear,
                   int align_corners, void* stream);
int dtf_resize_bwd(const float* dy, float* dx, int N, int H, int W, int C, int OH, int OW, int bilinear,
                   int align_corners, void* stream);
/* Fake quantisation with TensorFlow's FakeQuantWithMinMaxVars semantics (tf.contrib.quantize rewrites behind
 * --quantize, vgg/nets/mobilenet_v1_train.py:135-139, mobilenet_v1_eval.py:126).  range = {min, max} in device memory;
 * gate (device float, may be NULL): 0 -> identity (quant_delay not reached).  Backward is straight-through inside
 * the nudged range.  dtf_quant_range_update folds the min / max of x into range: decay < 0 assigns them (weights),
 * otherwise range = range*decay + batch*(1-decay) with the batch range widened to contain 0 (activations).
 * mm = two floats of scratch that must hold {+inf, -inf} before the first call; each call leaves them so. */
int dtf_fake_quant_fwd(const float* x, float* y, long long n, const float* range, const float* gate, int bits,
                       int narrow_range, void* stream);
int dtf_fake_quant_bwd(const float* x, const float* dy, float* dx, long long n, const float* range, const float* gate,
                       int bits, int narrow_range, void* stream);
int dtf_quant_range_update(const float* x, long long n, float* mm, float* range, float decay, void* stream);
/* spatial padding (top, bottom, left, right); mode 0 zeros, 1 reflect (tf.pad 'REFLECT') */
int dtf_pad_fwd(const float* x, float* y, int N, int H, int W, int C, int pt, int pb, int pl, int pr, int mode,
                void* stream);
int dtf_pad_bwd(const float* dy, float* dx, int N, int H, int W, int C, int pt, int pb, int pl, int pr, int mode,
                void* stream);

/* ------------------------------------------------------------------------------------ */
/* CIFAR-10 input: crop + flip + contrast + per-image standardisation from uint8 records   */
/*     (cnn/cifar10_input.py:176-190 training, :242-247 evaluation; brightness cancels in  */
/*     the standardisation).  images = [num_records][32][32][3] bytes; per image b: record  */
/*     idx[b], crop origin (oy[b], ox[b]), flip[b], contrast factor[b]; out = float         */
/*     [B][S][S][3].  Evaluation = central origin, no flip, factor 1.                      */
/* ------------------------------------------------------------------------------------ */
/* device: every pointer is device memory; one thread block per image */
int dtf_cifar_distort(const unsigned char* images, const long long* idx, const int* oy, const int* ox,
                      const unsigned char* flip, const float* factor, int B, int S, float* out, void* stream);
/* host: the same arithmetic on `threads` host threads; indices and crop windows are range-checked */
int dtf_cifar_distort_host(const unsigned char* images, long long num_records, const long long* idx, const int* oy,
                           const int* ox, const unsigned char* flip, const float* factor, int B, int S, float* out,
                           int threads);
/* ImageNet colour distortion (inception/image_processing.py:164-195), host, in place on an HWC float image in [0, 1]:
 * order 0 = brightness, saturation + hue, contrast; order 1 = brightness, contrast, saturation + hue; then clip. */
int dtf_distort_color_host(float* img, long long num_pixels, int order, float delta, float saturation, float hue,
                           float contrast);
/* Window (y0, x0, ch, cw) of a decoded uint8 HWC image -> float HWC [out_h][out_w][3] in [0, 1], host
 * (image_processing.py:233-244, :271-274).  method: 0 bilinear (half-pixel centres), 1 nearest, 2 bicubic (a = -0.75),
 * 3 area; flip != 0 mirrors left-right. */
int dtf_crop_resize_host(const unsigned char* img, int H, int W, int y0, int x0, int ch, int cw, int out_h, int out_w,
                         int method, int flip, float* out);

/* ------------------------------------------------------------------------------------ */
/* K16 + C1..C5  parameter-server update / push / pull / clocks                           */
/*     hyper[7] (device): lr, grad_scale, weight_decay, rms_decay, momentum, epsilon,    */
/*     ema_decay (< 0 disables EMA).  kind: 0 SGD, 1 TF1 RMSProp with momentum.          */
/* ------------------------------------------------------------------------------------ */
int dtf_fused_update(float* w, const float* g, float* ms, float* mom, float* ema, long long n,
                     long long n_decayed, const float* hyper, int kind, void* stream);
/* BSP push of one bucket [lo, lo+len) of the flat buffer over peer-mapped memory: rank r owns
 * [lo + r*chunk, min(lo + (r+1)*chunk, lo+len)).  The kernel signals ready[rank] to every peer, waits for all W
 * ready words, reduces its part over the W gradient buffers, applies the optimiser (slots ms / mom / ema: this
 * rank's part only), stores the result into every replica and raises updated[rank] in every peer's flag block; it
 * does not wait for the other owners (dtf_ps_gate / the convolution gate do, where the values are consumed).
 * Flag block of a rank: uint32 words; ready_off / updated_off are the word offsets of this bucket's W-word rows.
 * *epoch_ptr = completed steps; epoch_out (NULL except on the step's last bucket) receives *epoch_ptr + 1.
 * mc_grads / mc_params: NVLS multicast addresses of the gradient / parameter regions of the same symmetric buffer
 * (both or neither; NULL = unicast peer loads and stores): the reduction then runs in the NVSwitch
 * (multimem.ld_reduce) and the updated range is written to all replicas with one multimem.st per 16 bytes.
 * timeout_s <= 0: 300 s.  max_blocks <= 0: one 1024-thread block per SM. */
int dtf_ps_bucket_step(const float* const* peer_grads, float* const* peer_params, unsigned int* const* peer_flags,
                       unsigned int* local_counter, unsigned int* status, int rank, int world, int ready_off,
                       int updated_off, long long lo, long long len, long long chunk, long long wd_boundary,
                       float* ms, float* mom, float* ema, const float* hyper, int kind,
                       const unsigned int* epoch_ptr, unsigned int* epoch_out, const float* mc_grads,
                       float* mc_params, double timeout_s, int max_blocks, void* stream);
/* Wait (one warp, bounded) until flags[0 .. n) >= *epoch_ptr; status = 1 on timeout. */
int dtf_ps_gate(const unsigned int* flags, int n, const unsigned int* epoch_ptr, unsigned int* status,
                double timeout_s, void* stream);
/* ASP/SSP push of the range [lo, lo+len) of the flat buffer (the whole buffer, or one bucket while the backward pass
 * is still running): one-sided apply of the local gradient on the owners' master copies (flat shards: rank r owns
 * [r*shard, (r+1)*shard)), then refresh of the replica from them; no flags, no barrier.  bump_step: global_step += 1
 * (the step's first launch).  max_blocks <= 0: one block per SM. */
int dtf_ps_async_step(const float* grads, float* const* peer_params, float* const* peer_ms, float* const* peer_mom,
                      float* const* peer_ema, unsigned long long* global_step, int rank, int world, long long shard,
                      long long wd_boundary, const float* hyper, int kind, long long lo, long long len, int bump_step,
                      int max_blocks, void* stream);
/* Debug aid: 0 keeps the SGD ASP / SSP push on the register kernel; 1 (default) uses the bulk-copy-engine kernel
 * (cp.reduce.async.bulk add into the owner's HBM, bulk loads back). */
int dtf_debug_set_async_bulk(int enabled);
/* SSP: publish clock[rank] = step to every peer, wait while step > min(clock) + bound (status 2 on timeout). */
int dtf_ssp_clock(unsigned int* const* peer_clocks, int rank, int world, unsigned int step, unsigned int bound,
                  double timeout_s, unsigned int* status, unsigned int* waited_polls, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DTF_B200_H_ */

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
 *             stored value into stats[0..Nout) / stats[Nout..2Nout) (atomic adds).
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
} dtf_conv_gemm_params;

int dtf_conv_gemm(const dtf_conv_gemm_params* p, void* stream);

/* Weight gradient: dW[n][k] (+)= sum_m dY[m, n] * A[m, k], A gathered exactly as in fprop
 * (with the same optional fused input activation).  Split over m across CTAs; partial sums
 * are combined with red.global.add.f32, so dW must be zeroed (or hold the value to
 * accumulate into) before the call.  Optional dbias[n] += sum_m dY[m, n]. */
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

/* [Nout][R*S][C] -> [C][R*S][Nout]: builds the dgrad B operand from OHWI weights. */
int dtf_weight_transpose(const float* w, float* wt, int Nout, int RS, int C, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DTF_B200_H_ */

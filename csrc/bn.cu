// K7: batch-norm kernels for NHWC fp32 tensors viewed as [M = N*H*W, C] (C % 4 == 0).
//
// The normalisation itself is never a stand-alone pass on the hot path: per-channel sum / sum-of-squares
// come out of the epilogue of the convolution that produced the tensor (conv_igemm.cu), `bn_finalize` turns
// them into a per-channel (scale, shift) pair and updates the moving statistics, and the *consumer*
// convolution applies scale/shift(+ReLU) while gathering its A operand.  The kernels here are:
//   bn_stats        stand-alone statistics (tensors not produced by one of our convs, e.g. pool outputs)
//   bn_finalize     sums -> mean / inv-std / scale / shift, moving-average update (TF assign_moving_average)
//   affine_act      y = act(x*scale + shift)   (stand-alone apply: network heads, pools, eval)
//   bn_bwd_reduce   sums of g and g*x over M, g = dy * [act(x) > 0]
//   bn_bwd_apply    dx = scale * (g - mean(g) - xhat * mean(g*xhat))  (+ optional skip-gradient add),
//                   and one block writes dgamma / dbeta
//
// Reference ops replaced: slim ops.batch_norm (inception/slim/ops.py:46-135: tf.nn.moments +
// assign_moving_average + tf.nn.batch_normalization) and tf.layers.batch_normalization(fused=True)
// (resnet/resnet_model.py:41-50), plus their gradients.
#include "../include/dtf_b200.h"
#include "host_common.h"
#include "umma.cuh"

namespace dtf {

// ---- column reductions over M for [M, C] -----------------------------------------------------------------
// Block = 256 threads.  Thread t owns channel chunk (t % cpb) (4 channels) and walks rows (t / cpb) + k*rpb of
// its block's row range; partial sums are combined through shared memory and flushed with one atomicAdd per
// channel per block.  MODE 0: (sum x, sum x^2).  MODE 1: (sum g, sum g*x) with g = dy*mask.
template <int MODE>
__global__ void __launch_bounds__(256)
col_reduce_kernel(const float* __restrict__ x, const float* __restrict__ dy, const float* __restrict__ scale,
                  const float* __restrict__ shift, int relu, long long M, int C, int ld_x, int ld_dy,
                  int rows_per_block, float* __restrict__ out /* [2C] */) {
  extern __shared__ float red[];   // [256][8]
  const int chunks = C >> 2;
  const int cpb = chunks < 256 ? chunks : 256;           // channel chunks handled concurrently
  const int rpb = 256 / cpb;                             // rows in flight per block iteration
  const int t = threadIdx.x;
  const int cslot = t % cpb;
  const int rslot = t / cpb;
  const long long row0 = (long long)blockIdx.x * rows_per_block;
  long long row1 = row0 + rows_per_block;
  if (row1 > M) row1 = M;
  for (int cbase = 0; cbase < chunks; cbase += cpb) {
    const int cc = cbase + cslot;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = make_float4(0.f, 0.f, 0.f, 0.f);
    if (cc < chunks && rslot < rpb) {
      float4 sc = make_float4(1.f, 1.f, 1.f, 1.f), sh = make_float4(0.f, 0.f, 0.f, 0.f);
      if (MODE == 1 && scale != nullptr) {
        sc = __ldg(reinterpret_cast<const float4*>(scale) + cc);
        sh = __ldg(reinterpret_cast<const float4*>(shift) + cc);
      }
      // four rows per trip, all loads issued before the first use: the kernel is a pure stream over x (and dy), so
      // what it needs is bytes in flight (one row per trip left it at ~1/6 of the HBM rate)
      constexpr int U = 4;
      const long long step = (long long)rpb;
      long long r = row0 + rslot;
      auto accumulate = [&](const float4& xv, float4 g) {
        if (MODE == 0) {
          a.x += xv.x; a.y += xv.y; a.z += xv.z; a.w += xv.w;
          b.x += xv.x * xv.x; b.y += xv.y * xv.y; b.z += xv.z * xv.z; b.w += xv.w * xv.w;
        } else {
          if (relu) {
            if (fmaf(xv.x, sc.x, sh.x) <= 0.f) g.x = 0.f;
            if (fmaf(xv.y, sc.y, sh.y) <= 0.f) g.y = 0.f;
            if (fmaf(xv.z, sc.z, sh.z) <= 0.f) g.z = 0.f;
            if (fmaf(xv.w, sc.w, sh.w) <= 0.f) g.w = 0.f;
          }
          a.x += g.x; a.y += g.y; a.z += g.z; a.w += g.w;
          b.x += g.x * xv.x; b.y += g.y * xv.y; b.z += g.z * xv.z; b.w += g.w * xv.w;
        }
      };
      for (; r + (U - 1) * step < row1; r += U * step) {
        float4 xv[U], gv[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          xv[u] = __ldg(reinterpret_cast<const float4*>(x + (r + u * step) * ld_x) + cc);
          if (MODE == 1) gv[u] = __ldg(reinterpret_cast<const float4*>(dy + (r + u * step) * ld_dy) + cc);
          else gv[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) accumulate(xv[u], gv[u]);
      }
      for (; r < row1; r += step) {
        const float4 xv = __ldg(reinterpret_cast<const float4*>(x + r * ld_x) + cc);
        float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
        if (MODE == 1) g = __ldg(reinterpret_cast<const float4*>(dy + r * ld_dy) + cc);
        accumulate(xv, g);
      }
    }
    float* mine = red + t * 8;
    mine[0] = a.x; mine[1] = a.y; mine[2] = a.z; mine[3] = a.w;
    mine[4] = b.x; mine[5] = b.y; mine[6] = b.z; mine[7] = b.w;
    __syncthreads();
    // thread (cslot, rslot=0) sums over the row slots
    if (rslot == 0 && cc < chunks) {
      float acc[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] = mine[e];
      for (int rs = 1; rs < rpb; ++rs) {
        const float* o = red + (rs * cpb + cslot) * 8;
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] += o[e];
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        atomicAdd(out + cc * 4 + e, acc[e]);
        atomicAdd(out + C + cc * 4 + e, acc[4 + e]);
      }
    }
    __syncthreads();
  }
}

// sums -> statistics.  One thread per channel.
__global__ void bn_finalize_kernel(const float* __restrict__ stats, const float* __restrict__ gamma,
                                   const float* __restrict__ beta, float inv_count, float unbias, float eps,
                                   float decay, int update_moving, float* __restrict__ moving_mean,
                                   float* __restrict__ moving_var, float* __restrict__ scale,
                                   float* __restrict__ shift, float* __restrict__ mean_out,
                                   float* __restrict__ invstd_out, int C) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float mean = stats[c] * inv_count;
  float var = stats[C + c] * inv_count - mean * mean;
  var = var < 0.f ? 0.f : var;
  const float invstd = rsqrtf(var + eps);
  const float g = gamma ? gamma[c] : 1.f;
  const float sc = g * invstd;
  scale[c] = sc;
  shift[c] = (beta ? beta[c] : 0.f) - mean * sc;
  if (mean_out) mean_out[c] = mean;
  if (invstd_out) invstd_out[c] = invstd;
  if (update_moving) {
    moving_mean[c] = moving_mean[c] * decay + mean * (1.f - decay);
    moving_var[c] = moving_var[c] * decay + var * unbias * (1.f - decay);
  }
}

// eval-mode: moving statistics -> scale / shift
__global__ void bn_eval_affine_kernel(const float* __restrict__ gamma, const float* __restrict__ beta,
                                      const float* __restrict__ moving_mean, const float* __restrict__ moving_var,
                                      float eps, float* __restrict__ scale, float* __restrict__ shift, int C) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float sc = (gamma ? gamma[c] : 1.f) * rsqrtf(moving_var[c] + eps);
  scale[c] = sc;
  shift[c] = (beta ? beta[c] : 0.f) - moving_mean[c] * sc;
}

__global__ void __launch_bounds__(256)
affine_act_kernel(const float* __restrict__ x, const float* __restrict__ scale, const float* __restrict__ shift,
                  int relu, long long total4, int chunks, int ld_x, int ld_y, float* __restrict__ y) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total4;
       i += (long long)gridDim.x * blockDim.x) {
    // 32-bit division when the index allows it (a 64-bit division by a run-time value is ~100 instructions)
    const long long r = (i >> 31) == 0 ? (long long)((unsigned)i / (unsigned)chunks) : i / chunks;
    const int cc = (int)(i - r * chunks);
    const float4 xv = __ldg(reinterpret_cast<const float4*>(x + r * ld_x) + cc);
    const float4 sc = __ldg(reinterpret_cast<const float4*>(scale) + cc);
    const float4 sh = __ldg(reinterpret_cast<const float4*>(shift) + cc);
    float4 o;
    o.x = fmaf(xv.x, sc.x, sh.x); o.y = fmaf(xv.y, sc.y, sh.y);
    o.z = fmaf(xv.z, sc.z, sh.z); o.w = fmaf(xv.w, sc.w, sh.w);
    if (relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
    reinterpret_cast<float4*>(y + r * ld_y)[cc] = o;
  }
}

// dx = scale * (g - sum_g/M - xhat * sum_gxhat/M), xhat = (x - mean)*invstd,
// sum_gxhat = invstd * (sum_gx - mean * sum_g).  `add` (optional) is accumulated into dx (skip-path gradient).
__global__ void __launch_bounds__(256)
bn_bwd_apply_kernel(const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ scale,
                    const float* __restrict__ shift, const float* __restrict__ mean, const float* __restrict__ invstd,
                    const float* __restrict__ sums /* [2C] */, int relu, float inv_count, long long total4,
                    int chunks, int C, int ld_x, int ld_dy, int ld_dx, const float* __restrict__ add, int ld_add,
                    float* __restrict__ dx, float* __restrict__ dgamma, float* __restrict__ dbeta) {
  pdl_wait();      // launched with programmatic stream serialization: block scheduling overlaps the predecessor's tail
  if (blockIdx.x == 0) {
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      const float sg = sums[c], sgx = sums[C + c];
      // accumulate: the destination is the (pre-zeroed) flat gradient buffer, and a BN whose output feeds two
      // convolutions (residual projection + first conv of the unit) receives two contributions
      if (dbeta) dbeta[c] += sg;
      if (dgamma) dgamma[c] += invstd[c] * (sgx - mean[c] * sg);
    }
  }
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total4;
       i += (long long)gridDim.x * blockDim.x) {
    // 32-bit division when the index allows it (a 64-bit division by a run-time value is ~100 instructions)
    const long long r = (i >> 31) == 0 ? (long long)((unsigned)i / (unsigned)chunks) : i / chunks;
    const int cc = (int)(i - r * chunks);
    const float4 xv = __ldg(reinterpret_cast<const float4*>(x + r * ld_x) + cc);
    float4 g = __ldg(reinterpret_cast<const float4*>(dy + r * ld_dy) + cc);
    const float4 sc = __ldg(reinterpret_cast<const float4*>(scale) + cc);
    const float4 sh = __ldg(reinterpret_cast<const float4*>(shift) + cc);
    const float4 mu = __ldg(reinterpret_cast<const float4*>(mean) + cc);
    const float4 is = __ldg(reinterpret_cast<const float4*>(invstd) + cc);
    const float4 sg = __ldg(reinterpret_cast<const float4*>(sums) + cc);
    const float4 sgx = __ldg(reinterpret_cast<const float4*>(sums + C) + cc);
    if (relu) {
      if (fmaf(xv.x, sc.x, sh.x) <= 0.f) g.x = 0.f;
      if (fmaf(xv.y, sc.y, sh.y) <= 0.f) g.y = 0.f;
      if (fmaf(xv.z, sc.z, sh.z) <= 0.f) g.z = 0.f;
      if (fmaf(xv.w, sc.w, sh.w) <= 0.f) g.w = 0.f;
    }
    float4 o;
#define DTF_BN_DX(f)                                                                   \
    {                                                                                  \
      const float xh = (xv.f - mu.f) * is.f;                                           \
      const float sgxh = is.f * (sgx.f - mu.f * sg.f);                                 \
      o.f = sc.f * (g.f - sg.f * inv_count - xh * sgxh * inv_count);                   \
    }
    DTF_BN_DX(x) DTF_BN_DX(y) DTF_BN_DX(z) DTF_BN_DX(w)
#undef DTF_BN_DX
    if (add) {
      const float4 av = __ldg(reinterpret_cast<const float4*>(add + r * ld_add) + cc);
      o.x += av.x; o.y += av.y; o.z += av.z; o.w += av.w;
    }
    reinterpret_cast<float4*>(dx + r * ld_dx)[cc] = o;
  }
}

// Eval-mode / frozen-statistics backward: dx = scale * g (no batch-statistics terms).
__global__ void __launch_bounds__(256)
affine_act_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ scale,
                      const float* __restrict__ shift, int relu, long long total4, int chunks, int ld_x,
                      int ld_dy, int ld_dx, float* __restrict__ dx) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total4;
       i += (long long)gridDim.x * blockDim.x) {
    // 32-bit division when the index allows it (a 64-bit division by a run-time value is ~100 instructions)
    const long long r = (i >> 31) == 0 ? (long long)((unsigned)i / (unsigned)chunks) : i / chunks;
    const int cc = (int)(i - r * chunks);
    const float4 xv = __ldg(reinterpret_cast<const float4*>(x + r * ld_x) + cc);
    float4 g = __ldg(reinterpret_cast<const float4*>(dy + r * ld_dy) + cc);
    const float4 sc = __ldg(reinterpret_cast<const float4*>(scale) + cc);
    const float4 sh = __ldg(reinterpret_cast<const float4*>(shift) + cc);
    if (relu) {
      if (fmaf(xv.x, sc.x, sh.x) <= 0.f) g.x = 0.f;
      if (fmaf(xv.y, sc.y, sh.y) <= 0.f) g.y = 0.f;
      if (fmaf(xv.z, sc.z, sh.z) <= 0.f) g.z = 0.f;
      if (fmaf(xv.w, sc.w, sh.w) <= 0.f) g.w = 0.f;
    }
    g.x *= sc.x; g.y *= sc.y; g.z *= sc.z; g.w *= sc.w;
    reinterpret_cast<float4*>(dx + r * ld_dx)[cc] = g;
  }
}

static int reduce_grid(long long M, int* rows_per_block) {
  // 6 blocks per SM (1536 of 2048 threads resident: the reductions are latency-bound streams), >= 32 rows per block
  long long blocks = 6LL * num_sms();
  long long rpb = (M + blocks - 1) / blocks;
  if (rpb < 32) rpb = 32;
  *rows_per_block = (int)rpb;
  return (int)((M + rpb - 1) / rpb);
}

static int ew_grid(long long total4) {
  long long b = (total4 + 255) / 256;
  const long long cap = 8LL * num_sms();
  return (int)(b < cap ? (b > 0 ? b : 1) : cap);
}

}  // namespace dtf

using namespace dtf;

extern "C" int dtf_bn_stats(const float* x, long long M, int C, int ld_x, float* stats, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(C % 4 == 0 && ld_x % 4 == 0, "channel count / pitch must be multiples of 4");
  if (M == 0) return 0;
  int rpb;
  const int grid = reduce_grid(M, &rpb);
  col_reduce_kernel<0><<<grid, 256, 256 * 8 * sizeof(float), stream>>>(x, nullptr, nullptr, nullptr, 0, M, C, ld_x, 0,
                                                                        rpb, stats);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_bn_finalize(const float* stats, const float* gamma, const float* beta, long long count, float eps,
                               float decay, int update_moving, int unbiased_moving_var, float* moving_mean,
                               float* moving_var, float* scale, float* shift, float* mean_out, float* invstd_out,
                               int C, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  const float unbias = (unbiased_moving_var && count > 1) ? (float)count / (float)(count - 1) : 1.f;
  bn_finalize_kernel<<<ceil_div(C, 128), 128, 0, stream>>>(stats, gamma, beta, 1.f / (float)count, unbias, eps, decay,
                                                           update_moving, moving_mean, moving_var, scale, shift,
                                                           mean_out, invstd_out, C);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

// Training-mode batch norm forward as one call: zero scratch -> column sums -> statistics / moving update ->
// y = act(x*scale + shift).  `scratch` (2*C floats) holds the per-channel sums; `coef` (4*C floats) receives
// scale | shift | mean | invstd for the backward pass.
static int bn_forward_train_impl(const float* x, long long M, int C, int ld_x, const float* gamma, const float* beta,
                                 float eps, float decay, int update_moving, int unbiased_moving_var,
                                 float* moving_mean, float* moving_var, int relu, float* scratch, float* coef,
                                 float* y, int ld_y, cudaStream_t stream) {
  DTF_REQUIRE(C % 4 == 0 && ld_x % 4 == 0 && ld_y % 4 == 0, "channel count / pitch must be multiples of 4");
  if (M == 0) return 0;
  DTF_CHECK(cudaMemsetAsync(scratch, 0, 2 * (size_t)C * sizeof(float), stream));
  int rpb;
  const int grid = reduce_grid(M, &rpb);
  col_reduce_kernel<0><<<grid, 256, 256 * 8 * sizeof(float), stream>>>(x, nullptr, nullptr, nullptr, 0, M, C, ld_x, 0,
                                                                        rpb, scratch);
  DTF_CHECK(cudaGetLastError());
  const float unbias = (unbiased_moving_var && M > 1) ? (float)M / (float)(M - 1) : 1.f;
  bn_finalize_kernel<<<ceil_div(C, 128), 128, 0, stream>>>(scratch, gamma, beta, 1.f / (float)M, unbias, eps, decay,
                                                           update_moving, moving_mean, moving_var, coef, coef + C,
                                                           coef + 2 * C, coef + 3 * C, C);
  DTF_CHECK(cudaGetLastError());
  const long long total4 = M * (C / 4);
  affine_act_kernel<<<ew_grid(total4), 256, 0, stream>>>(x, coef, coef + C, relu, total4, C / 4, ld_x, ld_y, y);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_bn_forward_train(const float* x, long long M, int C, int ld_x, const float* gamma,
                                    const float* beta, float eps, float decay, int update_moving,
                                    int unbiased_moving_var, float* moving_mean, float* moving_var, int relu,
                                    float* scratch, float* coef, float* y, int ld_y, void* stream_) {
  return bn_forward_train_impl(x, M, C, ld_x, gamma, beta, eps, decay, update_moving, unbiased_moving_var,
                               moving_mean, moving_var, relu, scratch, coef, y, ld_y,
                               reinterpret_cast<cudaStream_t>(stream_));
}

// ---------------------------------------------------------------------------------------------------------
// dtf_bn_forward_train_static: the same training-mode forward for callers that cannot hand us scratch memory.
//
// Who needs this: code that runs where allocating device memory is off the table — an inference-server or
// framework plug-in callback that is only allowed to enqueue work on the stream it is given, a process whose
// memory pool has already been carved up completely by its host framework, or a latency-critical loop in
// which a cudaMalloc (which synchronises the whole device) would stall every other stream.  Such a caller
// still owns its inputs and outputs (x, the moving statistics, coef, y) but has nowhere to put the 2*C floats
// of per-channel sums, so they live in the file-scope array below instead.
//
// Consequences the caller has to know about:
//   * The array is part of the library image: every GPU that loads the library gives up
//     2 * DTF_BN_STATIC_MAX_CHANNELS floats (32 KiB) until the process exits.
//   * C above DTF_BN_STATIC_MAX_CHANNELS is refused with an error; nothing is launched.
//   * The first call on each device looks up the array's device address (cudaGetSymbolAddress, with that
//     device current) and remembers it.  Make that first call before you start capturing a stream.
//   * All calls share the one array and it keeps whatever the last call left in it.  Each call therefore
//     clears the part it uses before accumulating, and two calls whose kernels can overlap on the GPU (different
//     streams, or concurrent host threads on one device) race on it — serialise them.
//   * No cudaMalloc / cudaFree / allocating library routine is called, here or in the kernels.
// ---------------------------------------------------------------------------------------------------------
static __device__ __align__(256) float g_bn_static_scratch[2 * DTF_BN_STATIC_MAX_CHANNELS];

// In host code `g_bn_static_scratch` is only a handle for the runtime, not an address a kernel may use: the
// device address has to be asked for, once per device, while that device is current.
static int bn_static_scratch_address(float** out) {
  static std::mutex mu;
  static float* resolved[64] = {nullptr};
  int dev = -1;
  DTF_CHECK(cudaGetDevice(&dev));
  DTF_REQUIRE(dev >= 0 && dev < 64, "device ordinal out of range for the static scratch table");
  std::lock_guard<std::mutex> lk(mu);
  if (resolved[dev] == nullptr) {
    float* ptr = nullptr;
    DTF_CHECK(cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_bn_static_scratch));
    DTF_REQUIRE(ptr != nullptr, "cudaGetSymbolAddress returned a null address");
    resolved[dev] = ptr;
  }
  *out = resolved[dev];
  return 0;
}

extern "C" int dtf_bn_forward_train_static(const float* x, long long M, int C, int ld_x, const float* gamma,
                                           const float* beta, float eps, float decay, int update_moving,
                                           int unbiased_moving_var, float* moving_mean, float* moving_var, int relu,
                                           float* coef, float* y, int ld_y, void* stream_) {
  if (C > DTF_BN_STATIC_MAX_CHANNELS) {
    char msg[160];
    snprintf(msg, sizeof(msg), "dtf_bn_forward_train_static: C = %d exceeds the built-in scratch (max %d channels); "
             "use dtf_bn_forward_train with caller-provided scratch", C, DTF_BN_STATIC_MAX_CHANNELS);
    set_last_error(msg);
    return -3;
  }
  float* scratch = nullptr;
  const int rc = bn_static_scratch_address(&scratch);
  if (rc != 0) return rc;
  // bn_forward_train_impl starts by zeroing scratch[0, 2C) on the caller's stream, which is exactly the part
  // this call accumulates into; whatever an earlier call left beyond 2C is never read.
  return bn_forward_train_impl(x, M, C, ld_x, gamma, beta, eps, decay, update_moving, unbiased_moving_var,
                               moving_mean, moving_var, relu, scratch, coef, y, ld_y,
                               reinterpret_cast<cudaStream_t>(stream_));
}

extern "C" int dtf_bn_eval_affine(const float* gamma, const float* beta, const float* moving_mean,
                                  const float* moving_var, float eps, float* scale, float* shift, int C,
                                  void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  bn_eval_affine_kernel<<<ceil_div(C, 128), 128, 0, stream>>>(gamma, beta, moving_mean, moving_var, eps, scale, shift, C);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_affine_act(const float* x, const float* scale, const float* shift, int relu, long long M, int C,
                              int ld_x, int ld_y, float* y, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(C % 4 == 0 && ld_x % 4 == 0 && ld_y % 4 == 0, "channel count / pitch must be multiples of 4");
  const long long total4 = M * (C / 4);
  if (total4 == 0) return 0;
  affine_act_kernel<<<ew_grid(total4), 256, 0, stream>>>(x, scale, shift, relu, total4, C / 4, ld_x, ld_y, y);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_bn_bwd_reduce(const float* dy, const float* x, const float* scale, const float* shift, int relu,
                                 long long M, int C, int ld_x, int ld_dy, float* sums, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(C % 4 == 0 && ld_x % 4 == 0 && ld_dy % 4 == 0, "channel count / pitch must be multiples of 4");
  if (M == 0) return 0;
  int rpb;
  const int grid = reduce_grid(M, &rpb);
  col_reduce_kernel<1><<<grid, 256, 256 * 8 * sizeof(float), stream>>>(x, dy, scale, shift, relu, M, C, ld_x, ld_dy,
                                                                        rpb, sums);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_bn_bwd_apply(const float* dy, const float* x, const float* scale, const float* shift,
                                const float* mean, const float* invstd, const float* sums, int relu, long long M,
                                int C, int ld_x, int ld_dy, int ld_dx, const float* add, int ld_add, float* dx,
                                float* dgamma, float* dbeta, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(C % 4 == 0, "channel count must be a multiple of 4");
  const long long total4 = M * (C / 4);
  if (total4 == 0) return 0;
  DTF_CHECK(launch_pdl(bn_bwd_apply_kernel, dim3(ew_grid(total4)), dim3(256), 0, stream, dy, x, scale, shift, mean,
                       invstd, sums, relu, 1.f / (float)M, total4, C / 4, C, ld_x, ld_dy, ld_dx, add, ld_add, dx, dgamma,
                       dbeta));
  return 0;
}

extern "C" int dtf_affine_act_bwd(const float* dy, const float* x, const float* scale, const float* shift, int relu,
                                  long long M, int C, int ld_x, int ld_dy, int ld_dx, float* dx, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(C % 4 == 0, "channel count must be a multiple of 4");
  const long long total4 = M * (C / 4);
  if (total4 == 0) return 0;
  affine_act_bwd_kernel<<<ew_grid(total4), 256, 0, stream>>>(dy, x, scale, shift, relu, total4, C / 4, ld_x, ld_dy,
                                                             ld_dx, dx);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

// K4 / K5: max- and average-pooling for NHWC fp32 (C % 4 == 0), forward and backward, TensorFlow padding
// semantics: explicit (pad_top, pad_left) with the output size given by the caller (so 'SAME' with its
// possibly larger bottom/right padding and 'VALID' are both expressible); max-pool padding never wins,
// avg-pool divides by the number of *in-range* taps (tf.nn.avg_pool).
//
// One thread per (n, oh, ow, 4-channel chunk), 128-bit loads/stores.  max-pool forward records the winning
// tap in one byte per element; the backward kernels are gathers (each input element collects from the
// <= ceil(k/s)^2 windows that cover it) so they are deterministic and atomic-free.
//
// Reference ops replaced: tf.nn.max_pool (cnn/cifar10.py:228,248; slim/ops.py:370), slim.max_pool2d
// (cifarnet/cifarnet.py:66,72; alexnet/alexnet.py:106-112; vgg/nets/vgg.py:191-199), tf.nn.avg_pool
// (slim/ops.py:400), tf.layers.average_pooling2d (resnet/resnet_model.py:271) and their gradients.
#include "../include/dtf_b200.h"
#include "host_common.h"

namespace dtf {

struct PoolGeom {
  int N, H, W, C, OH, OW, kh, kw, sh, sw, ph, pw, ld_in, ld_out;
};

__global__ void __launch_bounds__(256)
max_pool_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, unsigned char* __restrict__ arg, PoolGeom g,
                    long long total4) {
  const int chunks = g.C >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total4;
       i += (long long)gridDim.x * blockDim.x) {
    // 32-bit index arithmetic (the launcher rejects tensors of 2^31 chunks or more): 64-bit divisions by run-time
    // values cost ~100 instructions each and made these kernels issue-bound
    const unsigned iu = (unsigned)i;
    const int cc = (int)(iu % (unsigned)chunks);
    unsigned p = iu / (unsigned)chunks;
    const int ow = (int)(p % (unsigned)g.OW); p /= (unsigned)g.OW;
    const int oh = (int)(p % (unsigned)g.OH);
    const int n = (int)(p / (unsigned)g.OH);
    float4 m = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
    uchar4 am = make_uchar4(0, 0, 0, 0);
    for (int r = 0; r < g.kh; ++r) {
      const int ih = oh * g.sh - g.ph + r;
      if (ih < 0 || ih >= g.H) continue;
      for (int s = 0; s < g.kw; ++s) {
        const int iw = ow * g.sw - g.pw + s;
        if (iw < 0 || iw >= g.W) continue;
        const float4 v = __ldg(reinterpret_cast<const float4*>(x + ((long long)(n * g.H + ih) * g.W + iw) * g.ld_in) + cc);
        const unsigned char tap = (unsigned char)(r * g.kw + s);
        if (v.x > m.x) { m.x = v.x; am.x = tap; }
        if (v.y > m.y) { m.y = v.y; am.y = tap; }
        if (v.z > m.z) { m.z = v.z; am.z = tap; }
        if (v.w > m.w) { m.w = v.w; am.w = tap; }
      }
    }
    const long long opix = (long long)(n * g.OH + oh) * g.OW + ow;
    reinterpret_cast<float4*>(y + opix * g.ld_out)[cc] = m;
    if (arg) reinterpret_cast<uchar4*>(arg + opix * g.C)[cc] = am;
  }
}

__global__ void __launch_bounds__(256)
max_pool_bwd_kernel(const float* __restrict__ dy, const unsigned char* __restrict__ arg, float* __restrict__ dx,
                    PoolGeom g, int ld_dy, int ld_dx, long long total4) {
  const int chunks = g.C >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total4;
       i += (long long)gridDim.x * blockDim.x) {
    const unsigned iu = (unsigned)i;
    const int cc = (int)(iu % (unsigned)chunks);
    unsigned p = iu / (unsigned)chunks;
    const int iw = (int)(p % (unsigned)g.W); p /= (unsigned)g.W;
    const int ih = (int)(p % (unsigned)g.H);
    const int n = (int)(p / (unsigned)g.H);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    // output rows oh with oh*sh - ph <= ih <= oh*sh - ph + kh - 1
    int oh_lo = (ih + g.ph - g.kh + g.sh) / g.sh; if (ih + g.ph - g.kh + 1 <= 0) oh_lo = 0;
    int oh_hi = (ih + g.ph) / g.sh; if (oh_hi >= g.OH) oh_hi = g.OH - 1;
    int ow_lo = (iw + g.pw - g.kw + g.sw) / g.sw; if (iw + g.pw - g.kw + 1 <= 0) ow_lo = 0;
    int ow_hi = (iw + g.pw) / g.sw; if (ow_hi >= g.OW) ow_hi = g.OW - 1;
    for (int oh = oh_lo; oh <= oh_hi; ++oh) {
      const int r = ih - (oh * g.sh - g.ph);
      if (r < 0 || r >= g.kh) continue;
      for (int ow = ow_lo; ow <= ow_hi; ++ow) {
        const int s = iw - (ow * g.sw - g.pw);
        if (s < 0 || s >= g.kw) continue;
        const unsigned char tap = (unsigned char)(r * g.kw + s);
        const long long opix = (long long)(n * g.OH + oh) * g.OW + ow;
        const uchar4 am = reinterpret_cast<const uchar4*>(arg + opix * g.C)[cc];
        const float4 d = __ldg(reinterpret_cast<const float4*>(dy + opix * ld_dy) + cc);
        if (am.x == tap) acc.x += d.x;
        if (am.y == tap) acc.y += d.y;
        if (am.z == tap) acc.z += d.z;
        if (am.w == tap) acc.w += d.w;
      }
    }
    reinterpret_cast<float4*>(dx + ((long long)(n * g.H + ih) * g.W + iw) * ld_dx)[cc] = acc;
  }
}

__device__ __forceinline__ int valid_taps(int o, int stride, int pad, int k, int dim) {
  int lo = o * stride - pad, hi = lo + k;
  if (lo < 0) lo = 0;
  if (hi > dim) hi = dim;
  return hi > lo ? hi - lo : 0;
}

__global__ void __launch_bounds__(256)
avg_pool_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, PoolGeom g, long long total4) {
  const int chunks = g.C >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total4;
       i += (long long)gridDim.x * blockDim.x) {
    // 32-bit index arithmetic (the launcher rejects tensors of 2^31 chunks or more): 64-bit divisions by run-time
    // values cost ~100 instructions each and made these kernels issue-bound
    const unsigned iu = (unsigned)i;
    const int cc = (int)(iu % (unsigned)chunks);
    unsigned p = iu / (unsigned)chunks;
    const int ow = (int)(p % (unsigned)g.OW); p /= (unsigned)g.OW;
    const int oh = (int)(p % (unsigned)g.OH);
    const int n = (int)(p / (unsigned)g.OH);
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = 0; r < g.kh; ++r) {
      const int ih = oh * g.sh - g.ph + r;
      if (ih < 0 || ih >= g.H) continue;
      for (int s = 0; s < g.kw; ++s) {
        const int iw = ow * g.sw - g.pw + s;
        if (iw < 0 || iw >= g.W) continue;
        const float4 v = __ldg(reinterpret_cast<const float4*>(x + ((long long)(n * g.H + ih) * g.W + iw) * g.ld_in) + cc);
        a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
      }
    }
    const int cnt = valid_taps(oh, g.sh, g.ph, g.kh, g.H) * valid_taps(ow, g.sw, g.pw, g.kw, g.W);
    const float inv = cnt > 0 ? 1.f / (float)cnt : 0.f;
    a.x *= inv; a.y *= inv; a.z *= inv; a.w *= inv;
    const long long opix = (long long)(n * g.OH + oh) * g.OW + ow;
    reinterpret_cast<float4*>(y + opix * g.ld_out)[cc] = a;
  }
}

__global__ void __launch_bounds__(256)
avg_pool_bwd_kernel(const float* __restrict__ dy, float* __restrict__ dx, PoolGeom g, int ld_dy, int ld_dx,
                    long long total4) {
  const int chunks = g.C >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total4;
       i += (long long)gridDim.x * blockDim.x) {
    const unsigned iu = (unsigned)i;
    const int cc = (int)(iu % (unsigned)chunks);
    unsigned p = iu / (unsigned)chunks;
    const int iw = (int)(p % (unsigned)g.W); p /= (unsigned)g.W;
    const int ih = (int)(p % (unsigned)g.H);
    const int n = (int)(p / (unsigned)g.H);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    int oh_lo = (ih + g.ph - g.kh + g.sh) / g.sh; if (ih + g.ph - g.kh + 1 <= 0) oh_lo = 0;
    int oh_hi = (ih + g.ph) / g.sh; if (oh_hi >= g.OH) oh_hi = g.OH - 1;
    int ow_lo = (iw + g.pw - g.kw + g.sw) / g.sw; if (iw + g.pw - g.kw + 1 <= 0) ow_lo = 0;
    int ow_hi = (iw + g.pw) / g.sw; if (ow_hi >= g.OW) ow_hi = g.OW - 1;
    for (int oh = oh_lo; oh <= oh_hi; ++oh) {
      const int r = ih - (oh * g.sh - g.ph);
      if (r < 0 || r >= g.kh) continue;
      const int ch = valid_taps(oh, g.sh, g.ph, g.kh, g.H);
      for (int ow = ow_lo; ow <= ow_hi; ++ow) {
        const int s = iw - (ow * g.sw - g.pw);
        if (s < 0 || s >= g.kw) continue;
        const int cnt = ch * valid_taps(ow, g.sw, g.pw, g.kw, g.W);
        const float inv = cnt > 0 ? 1.f / (float)cnt : 0.f;
        const long long opix = (long long)(n * g.OH + oh) * g.OW + ow;
        const float4 d = __ldg(reinterpret_cast<const float4*>(dy + opix * ld_dy) + cc);
        acc.x += d.x * inv; acc.y += d.y * inv; acc.z += d.z * inv; acc.w += d.w * inv;
      }
    }
    reinterpret_cast<float4*>(dx + ((long long)(n * g.H + ih) * g.W + iw) * ld_dx)[cc] = acc;
  }
}

// Global average pool (window == whole map, e.g. the 8x8 heads of ResNet / Inception): one block per image,
// thread (slice, chunk) sums a strided subset of the pixels for 4 channels, slices are combined in smem.
__global__ void __launch_bounds__(256)
global_avg_pool_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int HW, int C, int ld_in, int ld_out) {
  extern __shared__ float4 part[];            // [slices][chunks]
  const int chunks = C >> 2;
  const int n = blockIdx.x;
  const float inv = 1.f / (float)HW;
  for (int c0 = 0; c0 < chunks; c0 += 256) {
    const int cw = chunks - c0 < 256 ? chunks - c0 : 256;       // chunks handled in this pass
    const int slices = 256 / cw;
    const int cc = threadIdx.x % cw, sl = threadIdx.x / cw;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    if (sl < slices) {
      for (int p = sl; p < HW; p += slices) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(x + ((long long)n * HW + p) * ld_in) + c0 + cc);
        a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
      }
      part[sl * cw + cc] = a;
    }
    __syncthreads();
    if (sl == 0) {
      for (int k = 1; k < slices; ++k) {
        const float4 v = part[k * cw + cc];
        a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
      }
      a.x *= inv; a.y *= inv; a.z *= inv; a.w *= inv;
      reinterpret_cast<float4*>(y + (long long)n * ld_out)[c0 + cc] = a;
    }
    __syncthreads();
  }
}

static int pool_grid(long long total4) {
  long long b = (total4 + 255) / 256;
  const long long cap = 16LL * num_sms();
  return (int)(b < cap ? (b > 0 ? b : 1) : cap);
}

static PoolGeom make_geom(const dtf_pool_params& p) {
  PoolGeom g;
  g.N = p.N; g.H = p.H; g.W = p.W; g.C = p.C; g.OH = p.OH; g.OW = p.OW;
  g.kh = p.kh; g.kw = p.kw; g.sh = p.sh; g.sw = p.sw; g.ph = p.ph; g.pw = p.pw;
  g.ld_in = p.ld_in; g.ld_out = p.ld_out;
  return g;
}

}  // namespace dtf

using namespace dtf;

extern "C" int dtf_max_pool_fwd(const dtf_pool_params* p, const float* x, float* y, unsigned char* argmax,
                                void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(p->C % 4 == 0 && p->ld_in % 4 == 0 && p->ld_out % 4 == 0, "channels / pitches must be multiples of 4");
  DTF_REQUIRE(p->kh * p->kw <= 255, "window too large for the one-byte argmax");
  const long long total4 = (long long)p->N * p->OH * p->OW * (p->C / 4);
  if (total4 == 0) return 0;
  DTF_REQUIRE(total4 < (1LL << 31), "pooling tensor too large for 32-bit indexing");
  max_pool_fwd_kernel<<<pool_grid(total4), 256, 0, stream>>>(x, y, argmax, make_geom(*p), total4);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_max_pool_bwd(const dtf_pool_params* p, const float* dy, int ld_dy, const unsigned char* argmax,
                                float* dx, int ld_dx, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  const long long total4 = (long long)p->N * p->H * p->W * (p->C / 4);
  if (total4 == 0) return 0;
  DTF_REQUIRE(total4 < (1LL << 31), "pooling tensor too large for 32-bit indexing");
  max_pool_bwd_kernel<<<pool_grid(total4), 256, 0, stream>>>(dy, argmax, dx, make_geom(*p), ld_dy, ld_dx, total4);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_avg_pool_fwd(const dtf_pool_params* p, const float* x, float* y, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(p->C % 4 == 0 && p->ld_in % 4 == 0 && p->ld_out % 4 == 0, "channels / pitches must be multiples of 4");
  const long long total4 = (long long)p->N * p->OH * p->OW * (p->C / 4);
  if (total4 == 0) return 0;
  if (p->OH == 1 && p->OW == 1 && p->kh == p->H && p->kw == p->W && p->ph == 0 && p->pw == 0) {
    global_avg_pool_fwd_kernel<<<p->N, 256, 256 * sizeof(float4), stream>>>(x, y, p->H * p->W, p->C, p->ld_in,
                                                                           p->ld_out);
    DTF_CHECK(cudaGetLastError());
    return 0;
  }
  DTF_REQUIRE(total4 < (1LL << 31), "pooling tensor too large for 32-bit indexing");
  avg_pool_fwd_kernel<<<pool_grid(total4), 256, 0, stream>>>(x, y, make_geom(*p), total4);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_avg_pool_bwd(const dtf_pool_params* p, const float* dy, int ld_dy, float* dx, int ld_dx,
                                void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  const long long total4 = (long long)p->N * p->H * p->W * (p->C / 4);
  if (total4 == 0) return 0;
  DTF_REQUIRE(total4 < (1LL << 31), "pooling tensor too large for 32-bit indexing");
  avg_pool_bwd_kernel<<<pool_grid(total4), 256, 0, stream>>>(dy, dx, make_geom(*p), ld_dy, ld_dx, total4);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

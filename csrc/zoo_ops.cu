// Kernels for the op kinds that appear only in the vendored model zoo: depthwise convolution
// (MobileNet v1/v2, NASNet/PNASNet separable convs), relu6 / leaky-relu / tanh / sigmoid activations, instance
// normalisation, nearest / bilinear resize and reflect / zero padding (DCGAN, pix2pix, CycleGAN).  All NHWC fp32.
//
// None of these is GEMM-shaped: a depthwise filter has one input channel per output channel, so there is no
// reduction dimension for the tensor core to chew on; they are bandwidth kernels and are written as such —
// 128-bit accesses along the channel axis, register reuse of the input window along W, gathers instead of
// scatters wherever the inverse map is cheap.
//
// Reference ops replaced: slim.separable_conv2d / tf.nn.depthwise_conv2d (vgg/nets/mobilenet_v1.py:260-277,
// vgg/nets/mobilenet/conv_blocks.py:120-140, vgg/nets/nasnet/nasnet_utils.py:206-240), tf.nn.relu6
// (mobilenet_v1.py:430), tf.nn.leaky_relu (vgg/nets/dcgan.py:89, pix2pix.py:172), tf.tanh (pix2pix.py:220),
// layers.instance_norm (vgg/nets/cyclegan.py:47-60), tf.image.resize_nearest_neighbor / resize_bilinear
// (cyclegan.py:96-103), tf.pad(..., 'REFLECT') (cyclegan.py:201-261).
#include "../include/dtf_b200.h"
#include "host_common.h"

namespace dtf {

typedef dtf_dwconv_params DwP;

__device__ __forceinline__ void fma4(float4& a, const float4& x, const float4& w) {
  a.x = fmaf(x.x, w.x, a.x); a.y = fmaf(x.y, w.y, a.y); a.z = fmaf(x.z, w.z, a.z); a.w = fmaf(x.w, w.w, a.w);
}

// ---------------------------------------------------------------------------------------------------------
// Depthwise convolution, multiplier 1, C % 4 == 0.  One thread owns 4 channels of TW consecutive output pixels
// of one row: per filter row it loads the (TW-1)*SW + S input columns once into registers and reuses them for
// every (output, tap) pair, so L1 traffic per output drops from 2*R*S to ~R*(S + SW*(TW-1))/TW + R*S/TW loads.
// FLIP walks the filter backwards: the input gradient of a stride-1 depthwise conv is the same kernel with the
// filter rotated by 180 degrees and padding (R-1-ph, S-1-pw).
// ---------------------------------------------------------------------------------------------------------
template <int S_, int SW, int TW, bool FLIP>
__global__ void __launch_bounds__(256, S_ <= 5 ? 4 : 2)      // <= 64 registers for 3x3 / 5x5: four blocks per SM
dwconv_v4_kernel(DwP p, const float4* __restrict__ x, const float4* __restrict__ w, const float4* __restrict__ bias,
                 float4* __restrict__ y, unsigned total) {
  constexpr int XS = (TW - 1) * SW + S_;
  const unsigned C4 = (unsigned)p.C >> 2;
  const unsigned owt = (unsigned)(p.OW + TW - 1) / TW;
  for (unsigned i = blockIdx.x * 256u + threadIdx.x; i < total; i += gridDim.x * 256u) {
    const unsigned c4 = i % C4;
    unsigned t = i / C4;
    const int owb = (int)(t % owt) * TW; t /= owt;
    const int oh = (int)(t % (unsigned)p.OH);
    const int n = (int)(t / (unsigned)p.OH);
    float4 acc[TW];
    const float4 b0 = bias ? __ldg(bias + c4) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int j = 0; j < TW; ++j) acc[j] = b0;
    const int ih0 = oh * p.sh - p.ph, iw0 = owb * SW - p.pw;
    for (int r = 0; r < p.R; ++r) {
      const int ih = ih0 + r;
      if ((unsigned)ih >= (unsigned)p.H) continue;
      // one 64-bit address per row, then a running pointer (the profile showed ~600 instructions per thread, most
      // of them 64-bit index arithmetic, at half the issue rate and 37 % occupancy)
      const float4* xp = x + ((size_t)(n * p.H + ih) * p.W + iw0) * C4 + c4;
      float4 xs[XS];
#pragma unroll
      for (int k = 0; k < XS; ++k) {
        xs[k] = (unsigned)(iw0 + k) < (unsigned)p.W ? __ldg(xp) : make_float4(0.f, 0.f, 0.f, 0.f);
        xp += C4;
      }
      const int wr = FLIP ? p.R - 1 - r : r;
#pragma unroll
      for (int s = 0; s < S_; ++s) {
        const float4 wv = __ldg(w + (size_t)(wr * S_ + (FLIP ? S_ - 1 - s : s)) * C4 + c4);
#pragma unroll
        for (int j = 0; j < TW; ++j) fma4(acc[j], xs[j * SW + s], wv);
      }
    }
    float4* yo = y + ((size_t)(n * p.OH + oh) * p.OW + owb) * C4 + c4;
#pragma unroll
    for (int j = 0; j < TW; ++j)
      if (owb + j < p.OW) yo[(size_t)j * C4] = acc[j];
  }
}

// Any multiplier / channel count / filter / stride: one thread per output element, coalesced along channels.
__global__ void __launch_bounds__(256)
dwconv_fwd_generic_kernel(DwP p, const float* __restrict__ x, const float* __restrict__ w,
                          const float* __restrict__ bias, float* __restrict__ y, unsigned total) {
  const unsigned CO = (unsigned)(p.C * p.mult);
  for (unsigned i = blockIdx.x * 256u + threadIdx.x; i < total; i += gridDim.x * 256u) {
    const unsigned oc = i % CO;
    unsigned t = i / CO;
    const int ow = (int)(t % (unsigned)p.OW); t /= (unsigned)p.OW;
    const int oh = (int)(t % (unsigned)p.OH);
    const int n = (int)(t / (unsigned)p.OH);
    const int c = (int)oc / p.mult;
    float acc = bias ? __ldg(bias + oc) : 0.f;
    for (int r = 0; r < p.R; ++r) {
      const int ih = oh * p.sh - p.ph + r;
      if ((unsigned)ih >= (unsigned)p.H) continue;
      for (int s = 0; s < p.S; ++s) {
        const int iw = ow * p.sw - p.pw + s;
        if ((unsigned)iw >= (unsigned)p.W) continue;
        acc = fmaf(__ldg(x + ((size_t)(n * p.H + ih) * p.W + iw) * p.C + c), __ldg(w + (size_t)(r * p.S + s) * CO + oc), acc);
      }
    }
    y[i] = acc;
  }
}

// Input gradient, any geometry: dx[n,ih,iw,c] = sum_{r,s,m} dy[n,oh,ow,c*mult+m] * w[r,s,c,m] over the taps with
// oh*sh - ph + r == ih (a gather over the <= ceil(R/sh) x ceil(S/sw) windows that cover the pixel: no atomics).
__global__ void __launch_bounds__(256)
dwconv_dgrad_generic_kernel(DwP p, const float* __restrict__ dy, const float* __restrict__ w, float* __restrict__ dx,
                            unsigned total) {
  const unsigned C = (unsigned)p.C;
  const int CO = p.C * p.mult;
  for (unsigned i = blockIdx.x * 256u + threadIdx.x; i < total; i += gridDim.x * 256u) {
    const int c = (int)(i % C);
    unsigned t = i / C;
    const int iw = (int)(t % (unsigned)p.W); t /= (unsigned)p.W;
    const int ih = (int)(t % (unsigned)p.H);
    const int n = (int)(t / (unsigned)p.H);
    float acc = 0.f;
    for (int r = 0; r < p.R; ++r) {
      const int th = ih + p.ph - r;
      if (th < 0 || th % p.sh) continue;
      const int oh = th / p.sh;
      if (oh >= p.OH) continue;
      for (int s = 0; s < p.S; ++s) {
        const int tw = iw + p.pw - s;
        if (tw < 0 || tw % p.sw) continue;
        const int ow = tw / p.sw;
        if (ow >= p.OW) continue;
        const float* g = dy + ((size_t)(n * p.OH + oh) * p.OW + ow) * CO + c * p.mult;
        const float* wv = w + (size_t)(r * p.S + s) * CO + c * p.mult;
        for (int m = 0; m < p.mult; ++m) acc = fmaf(__ldg(g + m), __ldg(wv + m), acc);
      }
    }
    dx[i] = acc;
  }
}

// Filter gradient: dw[r,s,c,m] += sum_{n,oh,ow} x[n, oh*sh-ph+r, ow*sw-pw+s, c] * dy[n,oh,ow,c*mult+m].
// Block = (pixel chunk, 32 output channels, filter row r); lane = channel (128-byte coalesced rows), warp =
// pixel stripe; a thread keeps the S_ accumulators of its filter row in registers.  S_ == 0: one tap per block
// (blockIdx.z = r*S + s) for uncommon filter widths.
template <int S_>
__global__ void __launch_bounds__(256)
dwconv_wgrad_kernel(DwP p, const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ dw,
                    int pix_per_block) {
  constexpr int NS = S_ > 0 ? S_ : 1;
  __shared__ float red[8][NS][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int CO = p.C * p.mult;
  const int oc = blockIdx.y * 32 + lane;
  const bool valid = oc < CO;
  const int c = valid ? oc / p.mult : 0;
  const int r = S_ > 0 ? (int)blockIdx.z : (int)blockIdx.z / p.S;
  const int s0 = S_ > 0 ? 0 : (int)blockIdx.z % p.S;
  float acc[NS];
#pragma unroll
  for (int s = 0; s < NS; ++s) acc[s] = 0.f;
  const long long npix = (long long)p.N * p.OH * p.OW;
  const long long p0 = (long long)blockIdx.x * pix_per_block;
  long long p1 = p0 + pix_per_block;
  if (p1 > npix) p1 = npix;
  for (long long pix = p0 + warp; pix < p1; pix += 8) {
    const unsigned pu = (unsigned)pix;                   // the launcher guarantees < 2^32 pixels
    const int ow = (int)(pu % (unsigned)p.OW);
    const unsigned t = pu / (unsigned)p.OW;
    const int oh = (int)(t % (unsigned)p.OH);
    const int n = (int)(t / (unsigned)p.OH);
    const int ih = oh * p.sh - p.ph + r;
    if ((unsigned)ih >= (unsigned)p.H) continue;
    const float g = valid ? __ldg(dy + pix * CO + oc) : 0.f;
    const float* xrow = x + (size_t)(n * p.H + ih) * p.W * p.C + c;
#pragma unroll
    for (int s = 0; s < NS; ++s) {
      const int iw = ow * p.sw - p.pw + s0 + s;
      if ((unsigned)iw < (unsigned)p.W) acc[s] = fmaf(g, __ldg(xrow + (size_t)iw * p.C), acc[s]);
    }
  }
#pragma unroll
  for (int s = 0; s < NS; ++s) red[warp][s][lane] = acc[s];
  __syncthreads();
  if (warp == 0 && valid) {
#pragma unroll
    for (int s = 0; s < NS; ++s) {
      float a = 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) a += red[k][s][lane];
      atomicAdd(dw + (size_t)(r * p.S + s0 + s) * CO + oc, a);
    }
  }
}

// Input gradient for stride 2 (multiplier 1, C % 4 == 0): only the taps whose parity matches (ih + ph, iw + pw)
// reach an output pixel, so a thread walks r = r0, r0 + 2, ... and s = s0, s0 + 2, ...: ceil(K/2)^2 vector loads of
// dy instead of K^2 scalar loads with a division and a modulo each.
template <int K>
__global__ void __launch_bounds__(256)
dwconv_dgrad_s2_v4_kernel(DwP p, const float4* __restrict__ dy, const float4* __restrict__ w, float4* __restrict__ dx,
                          unsigned total) {
  const unsigned C4 = (unsigned)p.C >> 2;
  for (unsigned i = blockIdx.x * 256u + threadIdx.x; i < total; i += gridDim.x * 256u) {
    const unsigned c4 = i % C4;
    unsigned t = i / C4;
    const int iw = (int)(t % (unsigned)p.W); t /= (unsigned)p.W;
    const int ih = (int)(t % (unsigned)p.H);
    const int n = (int)(t / (unsigned)p.H);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    const int th = ih + p.ph, tw = iw + p.pw;
#pragma unroll
    for (int r = 0; r < K; ++r) {
      if ((th - r) & 1) continue;
      const int oh = (th - r) >> 1;
      if (th - r < 0 || oh >= p.OH) continue;
      const float4* grow = dy + (size_t)(n * p.OH + oh) * p.OW * C4 + c4;
#pragma unroll
      for (int s = 0; s < K; ++s) {
        if ((tw - s) & 1) continue;
        const int ow = (tw - s) >> 1;
        if (tw - s < 0 || ow >= p.OW) continue;
        fma4(acc, __ldg(grow + (size_t)ow * C4), __ldg(w + (size_t)(r * K + s) * C4 + c4));
      }
    }
    dx[i] = acc;
  }
}

// Filter gradient, multiplier 1, square K x K filter, stride SW in both directions.  A warp sweeps RB consecutive
// output rows of one image at once (lane = channel, so every load is one 128-byte line) and keeps all K*K
// accumulators in registers.  Per output column it loads dy once per row and only the SW input columns that entered
// the window of each of the (RB-1)*SW + K input rows involved (the rest is shifted along in registers).  Sweeping RB
// rows together matters for DRAM traffic, not just instruction count: a row sweep moves far more data through an SM
// than L1 holds (and, over 148 SMs, than L2 holds), so with RB = 1 every input row was fetched from DRAM once per
// filter row (ncu: 812 MB read for a 411 MB problem); with RB = 4 a 3x3 filter re-reads 6 input rows per 4 outputs.
// Partial sums meet in shared memory and leave the block as one atomic per (tap, channel).
template <int K, int SW, int RB>
__global__ void __launch_bounds__(256)
dwconv_wgrad_rows_kernel(DwP p, const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ dw,
                         int groups_per_warp) {
  constexpr int IR = (RB - 1) * SW + K;          // input rows under RB output rows
  __shared__ float red[8][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.y * 32 + lane;
  const bool valid = c < p.C;
  const int cc = valid ? c : 0;
  float acc[K][K];
#pragma unroll
  for (int r = 0; r < K; ++r)
#pragma unroll
    for (int s = 0; s < K; ++s) acc[r][s] = 0.f;
  const int gpi = (p.OH + RB - 1) / RB;          // row groups per image
  const int ngroups = p.N * gpi;
  const int g0 = (blockIdx.x * 8 + warp) * groups_per_warp;
  for (int g = g0; g < g0 + groups_per_warp && g < ngroups; ++g) {
    const int n = g / gpi, oh0 = (g - n * gpi) * RB;
    const float* grow[RB];
    bool glive[RB];
#pragma unroll
    for (int b = 0; b < RB; ++b) {
      glive[b] = oh0 + b < p.OH;
      grow[b] = dy + (size_t)(n * p.OH + (glive[b] ? oh0 + b : oh0)) * p.OW * p.C + cc;
    }
    const float* xrow[IR];
    bool live[IR];
#pragma unroll
    for (int i = 0; i < IR; ++i) {
      const int ih = oh0 * SW - p.ph + i;
      live[i] = (unsigned)ih < (unsigned)p.H;
      xrow[i] = x + (size_t)(n * p.H + (live[i] ? ih : 0)) * p.W * p.C + cc;
    }
    float win[IR][K];                      // win[i][s] = x[ih(i)][ow*SW - pw + s], carried from column to column
#pragma unroll
    for (int i = 0; i < IR; ++i)
#pragma unroll
      for (int s = 0; s < K; ++s) {
        const int iw = -p.pw + s;
        win[i][s] = (live[i] && (unsigned)iw < (unsigned)p.W) ? __ldg(xrow[i] + (size_t)iw * p.C) : 0.f;
      }
    // unrolled so that the loads of two columns are in flight together (they do not depend on the arithmetic)
#pragma unroll 2
    for (int ow = 0; ow < p.OW; ++ow) {
      float gv[RB];
#pragma unroll
      for (int b = 0; b < RB; ++b) gv[b] = glive[b] ? __ldg(grow[b] + (size_t)ow * p.C) : 0.f;
#pragma unroll
      for (int b = 0; b < RB; ++b)
#pragma unroll
        for (int r = 0; r < K; ++r)
#pragma unroll
          for (int s = 0; s < K; ++s) acc[r][s] = fmaf(gv[b], win[b * SW + r][s], acc[r][s]);
      // slide to column ow + 1: keep K - SW columns, fetch SW new ones (zero beyond the row's end)
#pragma unroll
      for (int i = 0; i < IR; ++i) {
#pragma unroll
        for (int s = 0; s + SW < K; ++s) win[i][s] = win[i][s + SW];
#pragma unroll
        for (int s = K - SW; s < K; ++s) {
          const int iw = (ow + 1) * SW - p.pw + s;
          win[i][s] = (live[i] && (unsigned)iw < (unsigned)p.W) ? __ldg(xrow[i] + (size_t)iw * p.C) : 0.f;
        }
      }
    }
  }
#pragma unroll
  for (int r = 0; r < K; ++r)
#pragma unroll
    for (int s = 0; s < K; ++s) {
      __syncthreads();
      red[warp][lane] = acc[r][s];
      __syncthreads();
      if (warp == 0 && valid) {
        float a = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) a += red[k][lane];
        atomicAdd(dw + (size_t)(r * K + s) * p.C + c, a);
      }
    }
}

// ---------------------------------------------------------------------------------------------------------
// Activations.  mode 0 relu6, 1 leaky-relu(alpha), 2 tanh, 3 sigmoid.  The backward kernel takes `ref` = the
// forward *input* for relu6 / leaky-relu and the forward *output* for tanh / sigmoid.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float act_f(float v, int mode, float alpha) {
  switch (mode) {
    case 0: return fminf(fmaxf(v, 0.f), 6.f);
    case 1: return v > 0.f ? v : alpha * v;
    case 2: return tanhf(v);
    default: return 1.f / (1.f + __expf(-v));
  }
}
__device__ __forceinline__ float act_b(float ref, float g, int mode, float alpha) {
  switch (mode) {
    case 0: return (ref > 0.f && ref < 6.f) ? g : 0.f;
    case 1: return ref > 0.f ? g : alpha * g;
    case 2: return g * (1.f - ref * ref);
    default: return g * ref * (1.f - ref);
  }
}

__global__ void __launch_bounds__(256)
act_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, long long n, int mode, float alpha) {
  const long long n4 = n >> 2;
  const long long stride = (long long)gridDim.x * 256;
  for (long long i = blockIdx.x * 256LL + threadIdx.x; i < n4; i += stride) {
    float4 v = __ldg(reinterpret_cast<const float4*>(x) + i);
    v.x = act_f(v.x, mode, alpha); v.y = act_f(v.y, mode, alpha);
    v.z = act_f(v.z, mode, alpha); v.w = act_f(v.w, mode, alpha);
    reinterpret_cast<float4*>(y)[i] = v;
  }
  for (long long i = (n4 << 2) + blockIdx.x * 256LL + threadIdx.x; i < n; i += stride) y[i] = act_f(x[i], mode, alpha);
}

__global__ void __launch_bounds__(256)
act_bwd_kernel(const float* __restrict__ ref, const float* __restrict__ dy, float* __restrict__ dx, long long n,
               int mode, float alpha) {
  const long long n4 = n >> 2;
  const long long stride = (long long)gridDim.x * 256;
  for (long long i = blockIdx.x * 256LL + threadIdx.x; i < n4; i += stride) {
    const float4 r = __ldg(reinterpret_cast<const float4*>(ref) + i);
    float4 g = __ldg(reinterpret_cast<const float4*>(dy) + i);
    g.x = act_b(r.x, g.x, mode, alpha); g.y = act_b(r.y, g.y, mode, alpha);
    g.z = act_b(r.z, g.z, mode, alpha); g.w = act_b(r.w, g.w, mode, alpha);
    reinterpret_cast<float4*>(dx)[i] = g;
  }
  for (long long i = (n4 << 2) + blockIdx.x * 256LL + threadIdx.x; i < n; i += stride)
    dx[i] = act_b(ref[i], dy[i], mode, alpha);
}

// ---------------------------------------------------------------------------------------------------------
// Fake quantisation (tf.contrib.quantize, used by vgg/nets/mobilenet_v1_train.py:135-139 / mobilenet_v1_eval.py:126):
// FakeQuantWithMinMaxVars semantics.  The [min, max] range lives in device memory (range[0], range[1]) and is
// "nudged" so that 0.0 falls on a quantisation step; the forward rounds onto the 2^bits (narrow: 2^bits - 1)
// steps of the nudged range, the backward passes the gradient where x lies inside it (straight-through).
// gate[0] == 0 switches the op to the identity (quant_delay not yet reached) without a host decision, so the
// step stays capturable.  A degenerate range (max <= min) is the identity too.
//   minmax  : batch min / max of a tensor, atomics on the float bit patterns into mm[0] (init +inf), mm[1] (-inf)
//   range   : decay < 0  -> range = batch (LastValueQuantize, weights)
//             otherwise  -> range = range*decay + batch*(1-decay) with the batch range widened to contain 0
//                           (MovingAvgQuantize, activations); mm is reset to (+inf, -inf) for the next call
// ---------------------------------------------------------------------------------------------------------
struct QuantSteps { float nmin, nmax, scale, inv_scale; bool on; };

__device__ __forceinline__ QuantSteps quant_steps(const float* __restrict__ range, const float* __restrict__ gate,
                                                  int bits, int narrow) {
  QuantSteps q;
  const float mn = __ldg(range), mx = __ldg(range + 1);
  const float qmin = narrow ? 1.f : 0.f, qmax = (float)((1 << bits) - 1);
  q.on = (gate == nullptr || __ldg(gate) != 0.f) && (mx > mn);
  q.scale = (mx - mn) / (qmax - qmin);
  q.on = q.on && (q.scale > 0.f);
  const float zp_from_min = qmin - mn / q.scale;
  const float zp = zp_from_min < qmin ? qmin : (zp_from_min > qmax ? qmax : roundf(zp_from_min));
  q.nmin = (qmin - zp) * q.scale;
  q.nmax = (qmax - zp) * q.scale;
  q.inv_scale = 1.f / q.scale;
  return q;
}
__device__ __forceinline__ float fq_f(float v, const QuantSteps& q) {
  const float c = fminf(fmaxf(v, q.nmin), q.nmax);
  return floorf((c - q.nmin) * q.inv_scale + 0.5f) * q.scale + q.nmin;
}

__global__ void __launch_bounds__(256)
fake_quant_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, long long n, const float* __restrict__ range,
                      const float* __restrict__ gate, int bits, int narrow) {
  const QuantSteps q = quant_steps(range, gate, bits, narrow);
  const long long n4 = n >> 2;
  const long long stride = (long long)gridDim.x * 256;
  for (long long i = blockIdx.x * 256LL + threadIdx.x; i < n4; i += stride) {
    float4 v = __ldg(reinterpret_cast<const float4*>(x) + i);
    if (q.on) { v.x = fq_f(v.x, q); v.y = fq_f(v.y, q); v.z = fq_f(v.z, q); v.w = fq_f(v.w, q); }
    reinterpret_cast<float4*>(y)[i] = v;
  }
  for (long long i = (n4 << 2) + blockIdx.x * 256LL + threadIdx.x; i < n; i += stride)
    y[i] = q.on ? fq_f(x[i], q) : x[i];
}

__global__ void __launch_bounds__(256)
fake_quant_bwd_kernel(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ dx, long long n,
                      const float* __restrict__ range, const float* __restrict__ gate, int bits, int narrow) {
  const QuantSteps q = quant_steps(range, gate, bits, narrow);
  const long long n4 = n >> 2;
  const long long stride = (long long)gridDim.x * 256;
  for (long long i = blockIdx.x * 256LL + threadIdx.x; i < n4; i += stride) {
    const float4 r = __ldg(reinterpret_cast<const float4*>(x) + i);
    float4 g = __ldg(reinterpret_cast<const float4*>(dy) + i);
    if (q.on) {
      g.x = (r.x >= q.nmin && r.x <= q.nmax) ? g.x : 0.f; g.y = (r.y >= q.nmin && r.y <= q.nmax) ? g.y : 0.f;
      g.z = (r.z >= q.nmin && r.z <= q.nmax) ? g.z : 0.f; g.w = (r.w >= q.nmin && r.w <= q.nmax) ? g.w : 0.f;
    }
    reinterpret_cast<float4*>(dx)[i] = g;
  }
  for (long long i = (n4 << 2) + blockIdx.x * 256LL + threadIdx.x; i < n; i += stride)
    dx[i] = (!q.on || (x[i] >= q.nmin && x[i] <= q.nmax)) ? dy[i] : 0.f;
}

// float min / max through integer atomics on the bit patterns (any mix of signs; -0.0 is canonicalised first)
__device__ __forceinline__ void atomic_min_f(float* a, float v) {
  v += 0.f;
  if (v >= 0.f) atomicMin(reinterpret_cast<int*>(a), __float_as_int(v));
  else atomicMax(reinterpret_cast<unsigned int*>(a), __float_as_uint(v));
}
__device__ __forceinline__ void atomic_max_f(float* a, float v) {
  v += 0.f;
  if (v >= 0.f) atomicMax(reinterpret_cast<int*>(a), __float_as_int(v));
  else atomicMin(reinterpret_cast<unsigned int*>(a), __float_as_uint(v));
}

__global__ void __launch_bounds__(256)
minmax_kernel(const float* __restrict__ x, long long n, float* __restrict__ mm) {
  float lo = INFINITY, hi = -INFINITY;
  const long long n4 = n >> 2;
  const long long stride = (long long)gridDim.x * 256;
  for (long long i = blockIdx.x * 256LL + threadIdx.x; i < n4; i += stride) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(x) + i);
    lo = fminf(fminf(lo, v.x), fminf(v.y, fminf(v.z, v.w)));
    hi = fmaxf(fmaxf(hi, v.x), fmaxf(v.y, fmaxf(v.z, v.w)));
  }
  for (long long i = (n4 << 2) + blockIdx.x * 256LL + threadIdx.x; i < n; i += stride) {
    lo = fminf(lo, x[i]); hi = fmaxf(hi, x[i]);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  __shared__ float s_lo[8], s_hi[8];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { s_lo[warp] = lo; s_hi[warp] = hi; }
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int w = 1; w < 8; ++w) { lo = fminf(lo, s_lo[w]); hi = fmaxf(hi, s_hi[w]); }
    if (lo <= hi) { atomic_min_f(mm, lo); atomic_max_f(mm + 1, hi); }
  }
}

__global__ void quant_range_kernel(float* __restrict__ mm, float* __restrict__ range, float decay) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    const float lo = mm[0], hi = mm[1];
    if (lo <= hi) {
      if (decay < 0.f) { range[0] = lo; range[1] = hi; }
      else {
        range[0] = range[0] * decay + fminf(lo, 0.f) * (1.f - decay);
        range[1] = range[1] * decay + fmaxf(hi, 0.f) * (1.f - decay);
      }
    }
    mm[0] = INFINITY; mm[1] = -INFINITY;
  }
}

// ---------------------------------------------------------------------------------------------------------
// Instance normalisation: per (image, channel) statistics over H*W.
//   reduce:  sums[n][0][c] += sum_p a ,  sums[n][1][c] += sum_p a*b
//            forward  : a = x,  b = x           -> sum x, sum x^2
//            backward : a = dy, b = xhat        -> sum dy, sum dy*xhat   (xhat from the forward sums)
//   apply :  y  = (x - mean) * invstd * gamma + beta
//   bwd   :  dx = gamma * invstd * (dy - S1/HW - xhat * S2/HW)
// Block = (32 channels, pixel split, image); lane = channel, warp = pixel stripe.
// ---------------------------------------------------------------------------------------------------------
template <bool BWD>
__global__ void __launch_bounds__(256)
instnorm_reduce_kernel(const float* __restrict__ a, const float* __restrict__ xin, const float* __restrict__ fsums,
                       float* __restrict__ sums, int HW, int C, float eps, int pix_per_block) {
  __shared__ float red[8][2][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  const int n = blockIdx.z;
  const bool valid = c < C;
  float mean = 0.f, invstd = 0.f;
  if (BWD && valid) {
    const float inv = 1.f / (float)HW;
    mean = fsums[(size_t)n * 2 * C + c] * inv;
    const float var = fmaxf(fsums[(size_t)n * 2 * C + C + c] * inv - mean * mean, 0.f);
    invstd = rsqrtf(var + eps);
  }
  const int p0 = blockIdx.y * pix_per_block;
  const int p1 = min(p0 + pix_per_block, HW);
  float s1 = 0.f, s2 = 0.f;
  if (valid) {
    const float* ap = a + (size_t)n * HW * C + c;
    const float* xp = xin + (size_t)n * HW * C + c;
    for (int pix = p0 + warp; pix < p1; pix += 8) {
      const float av = __ldg(ap + (size_t)pix * C);
      const float bv = BWD ? (__ldg(xp + (size_t)pix * C) - mean) * invstd : av;
      s1 += av;
      s2 = fmaf(av, bv, s2);
    }
  }
  red[warp][0][lane] = s1;
  red[warp][1][lane] = s2;
  __syncthreads();
  if (warp < 2 && valid) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += red[k][warp][lane];
    atomicAdd(sums + (size_t)n * 2 * C + (size_t)warp * C + c, t);
  }
}

template <bool BWD>
__global__ void __launch_bounds__(256)
instnorm_apply_kernel(const float* __restrict__ x, const float* __restrict__ dy, const float* __restrict__ fsums,
                      const float* __restrict__ bsums, const float* __restrict__ gamma, const float* __restrict__ beta,
                      float* __restrict__ out, int HW, int C, float eps, unsigned total) {
  const float inv = 1.f / (float)HW;
  for (unsigned i = blockIdx.x * 256u + threadIdx.x; i < total; i += gridDim.x * 256u) {
    const unsigned c = i % (unsigned)C;
    const unsigned n = i / ((unsigned)C * (unsigned)HW);
    const float* fs = fsums + (size_t)n * 2 * C;
    const float mean = __ldg(fs + c) * inv;
    const float var = fmaxf(__ldg(fs + C + c) * inv - mean * mean, 0.f);
    const float invstd = rsqrtf(var + eps);
    const float g = gamma ? __ldg(gamma + c) : 1.f;
    const float xhat = (__ldg(x + i) - mean) * invstd;
    if (!BWD) {
      out[i] = fmaf(xhat, g, beta ? __ldg(beta + c) : 0.f);
    } else {
      const float* bs = bsums + (size_t)n * 2 * C;
      out[i] = g * invstd * (__ldg(dy + i) - __ldg(bs + c) * inv - xhat * __ldg(bs + C + c) * inv);
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// Resize (TF1 semantics: src = dst * scale, scale = in/out, or (in-1)/(out-1) with align_corners; nearest takes
// floor(src) — round(src) with align_corners —, bilinear interpolates between floor(src) and floor(src)+1 clamped).
// Backward scatters with atomics into a zeroed dx.
// ---------------------------------------------------------------------------------------------------------
struct ResizeGeom { int N, H, W, C, OH, OW; float sch, scw; int bilinear, align; };

__device__ __forceinline__ void resize_src(const ResizeGeom& g, int oh, int ow, int& h0, int& h1, int& w0, int& w1,
                                           float& fh, float& fw) {
  const float sh = oh * g.sch, sw = ow * g.scw;
  if (g.bilinear) {
    h0 = min((int)floorf(sh), g.H - 1); h1 = min(h0 + 1, g.H - 1); fh = sh - (float)h0;
    w0 = min((int)floorf(sw), g.W - 1); w1 = min(w0 + 1, g.W - 1); fw = sw - (float)w0;
  } else {
    h0 = h1 = min(g.align ? (int)roundf(sh) : (int)floorf(sh), g.H - 1);
    w0 = w1 = min(g.align ? (int)roundf(sw) : (int)floorf(sw), g.W - 1);
    fh = fw = 0.f;
  }
}

template <bool BWD>
__global__ void __launch_bounds__(256)
resize_kernel(ResizeGeom g, const float* __restrict__ in, float* __restrict__ out, unsigned total) {
  // forward: in = x [N,H,W,C], out = y [N,OH,OW,C];   backward: in = dy [N,OH,OW,C], out = dx [N,H,W,C] (+=)
  for (unsigned i = blockIdx.x * 256u + threadIdx.x; i < total; i += gridDim.x * 256u) {
    const unsigned c = i % (unsigned)g.C;
    unsigned t = i / (unsigned)g.C;
    const int ow = (int)(t % (unsigned)g.OW); t /= (unsigned)g.OW;
    const int oh = (int)(t % (unsigned)g.OH);
    const int n = (int)(t / (unsigned)g.OH);
    int h0, h1, w0, w1;
    float fh, fw;
    resize_src(g, oh, ow, h0, h1, w0, w1, fh, fw);
    const size_t base = (size_t)n * g.H * g.W;
    const size_t i00 = ((base + (size_t)h0 * g.W + w0) * g.C) + c, i01 = ((base + (size_t)h0 * g.W + w1) * g.C) + c;
    const size_t i10 = ((base + (size_t)h1 * g.W + w0) * g.C) + c, i11 = ((base + (size_t)h1 * g.W + w1) * g.C) + c;
    if (!BWD) {
      if (g.bilinear) {
        const float top = __ldg(in + i00) + (__ldg(in + i01) - __ldg(in + i00)) * fw;
        const float bot = __ldg(in + i10) + (__ldg(in + i11) - __ldg(in + i10)) * fw;
        out[i] = top + (bot - top) * fh;
      } else {
        out[i] = __ldg(in + i00);
      }
    } else {
      const float gv = __ldg(in + i);
      if (g.bilinear) {
        atomicAdd(out + i00, gv * (1.f - fh) * (1.f - fw));
        atomicAdd(out + i01, gv * (1.f - fh) * fw);
        atomicAdd(out + i10, gv * fh * (1.f - fw));
        atomicAdd(out + i11, gv * fh * fw);
      } else {
        atomicAdd(out + i00, gv);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// Spatial padding of NHWC: mode 0 zeros, 1 reflect (tf.pad 'REFLECT': the edge pixel is not repeated).
// Forward is a gather; backward (reflect) scatters the padded gradient back with atomics into a zeroed dx.
// ---------------------------------------------------------------------------------------------------------
struct PadGeom { int N, H, W, C, pt, pl, OH, OW, mode; };

__device__ __forceinline__ int reflect_idx(int i, int n) {
  if (i < 0) i = -i;
  if (i >= n) i = 2 * (n - 1) - i;
  return i;
}

template <bool BWD>
__global__ void __launch_bounds__(256)
pad_kernel(PadGeom g, const float* __restrict__ in, float* __restrict__ out, unsigned total) {
  for (unsigned i = blockIdx.x * 256u + threadIdx.x; i < total; i += gridDim.x * 256u) {
    const unsigned c = i % (unsigned)g.C;
    unsigned t = i / (unsigned)g.C;
    const int ow = (int)(t % (unsigned)g.OW); t /= (unsigned)g.OW;
    const int oh = (int)(t % (unsigned)g.OH);
    const int n = (int)(t / (unsigned)g.OH);
    int ih = oh - g.pt, iw = ow - g.pl;
    bool inside = (unsigned)ih < (unsigned)g.H && (unsigned)iw < (unsigned)g.W;
    if (g.mode == 1) { ih = reflect_idx(ih, g.H); iw = reflect_idx(iw, g.W); inside = true; }
    const size_t src = (((size_t)n * g.H + ih) * g.W + iw) * g.C + c;
    if (!BWD) out[i] = inside ? __ldg(in + src) : 0.f;
    else if (inside) atomicAdd(out + src, __ldg(in + i));
  }
}

static inline unsigned grid_for(unsigned long long total, int per_block = 256, int waves = 16) {
  unsigned long long b = (total + per_block - 1) / per_block;
  const unsigned long long cap = (unsigned long long)waves * num_sms();
  return (unsigned)(b < cap ? (b ? b : 1) : cap);
}

template <int S_, int SW, int TW, bool FLIP>
static int launch_dw_v4(const DwP& p, const float* x, const float* w, const float* bias, float* y, cudaStream_t st) {
  const unsigned long long total = (unsigned long long)p.N * p.OH * ((p.OW + TW - 1) / TW) * (p.C / 4);
  DTF_REQUIRE(total < 0xffffffffull, "depthwise conv: tensor too large for 32-bit indexing");
  dwconv_v4_kernel<S_, SW, TW, FLIP><<<grid_for(total, 256, 1 << 20), 256, 0, st>>>(
      p, reinterpret_cast<const float4*>(x), reinterpret_cast<const float4*>(w), reinterpret_cast<const float4*>(bias),
      reinterpret_cast<float4*>(y), (unsigned)total);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

template <bool FLIP>
static int dispatch_dw_v4(const DwP& p, const float* x, const float* w, const float* bias, float* y, cudaStream_t st,
                          bool* done) {
  *done = true;
  if (p.S == 3 && p.sw == 1) return launch_dw_v4<3, 1, 4, FLIP>(p, x, w, bias, y, st);
  if (p.S == 5 && p.sw == 1) return launch_dw_v4<5, 1, 4, FLIP>(p, x, w, bias, y, st);
  if (p.S == 7 && p.sw == 1) return launch_dw_v4<7, 1, 4, FLIP>(p, x, w, bias, y, st);
  if (!FLIP) {
    if (p.S == 3 && p.sw == 2) return launch_dw_v4<3, 2, 2, false>(p, x, w, bias, y, st);
    if (p.S == 5 && p.sw == 2) return launch_dw_v4<5, 2, 2, false>(p, x, w, bias, y, st);
    if (p.S == 7 && p.sw == 2) return launch_dw_v4<7, 2, 2, false>(p, x, w, bias, y, st);
  }
  *done = false;
  return 0;
}

static bool aligned16(const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; }

}  // namespace dtf

using namespace dtf;

static int dw_check(const dtf_dwconv_params* p) {
  DTF_REQUIRE(p && p->N > 0 && p->H > 0 && p->W > 0 && p->C > 0 && p->mult > 0, "depthwise conv: empty tensor");
  DTF_REQUIRE(p->R > 0 && p->S > 0 && p->sh > 0 && p->sw > 0 && p->OH > 0 && p->OW > 0, "depthwise conv: bad geometry");
  DTF_REQUIRE((unsigned long long)p->N * p->H * p->W * p->C < 0xffffffffull &&
              (unsigned long long)p->N * p->OH * p->OW * p->C * p->mult < 0xffffffffull,
              "depthwise conv: tensor too large for 32-bit indexing");
  return 0;
}

extern "C" int dtf_dwconv_fwd(const dtf_dwconv_params* p, const float* x, const float* w, const float* bias, float* y,
                              void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  if (int rc = dw_check(p)) return rc;
  if (p->mult == 1 && p->C % 4 == 0 && aligned16(x) && aligned16(w) && aligned16(y) && (!bias || aligned16(bias))) {
    bool done = false;
    const int rc = dispatch_dw_v4<false>(*p, x, w, bias, y, st, &done);
    if (rc || done) return rc;
  }
  const unsigned total = (unsigned)((unsigned long long)p->N * p->OH * p->OW * p->C * p->mult);
  dwconv_fwd_generic_kernel<<<grid_for(total), 256, 0, st>>>(*p, x, w, bias, y, total);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_dwconv_dgrad(const dtf_dwconv_params* p, const float* dy, const float* w, float* dx,
                                void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  if (int rc = dw_check(p)) return rc;
  if (p->mult == 1 && p->C % 4 == 0 && p->sh == 1 && p->sw == 1 && p->R - 1 - p->ph >= 0 && p->S - 1 - p->pw >= 0 &&
      aligned16(dy) && aligned16(w) && aligned16(dx)) {
    // stride 1: correlation of dy with the rotated filter; roles of input and output swap
    DwP q = *p;
    q.H = p->OH; q.W = p->OW; q.OH = p->H; q.OW = p->W;
    q.ph = p->R - 1 - p->ph; q.pw = p->S - 1 - p->pw;
    bool done = false;
    const int rc = dispatch_dw_v4<true>(q, dy, w, nullptr, dx, st, &done);
    if (rc || done) return rc;
  }
  if (p->mult == 1 && p->C % 4 == 0 && p->sh == 2 && p->sw == 2 && p->R == p->S && (p->R == 3 || p->R == 5 || p->R == 7) &&
      aligned16(dy) && aligned16(w) && aligned16(dx)) {
    const unsigned total4 = (unsigned)((unsigned long long)p->N * p->H * p->W * (p->C / 4));
    const float4 *g4 = reinterpret_cast<const float4*>(dy), *w4 = reinterpret_cast<const float4*>(w);
    float4* d4 = reinterpret_cast<float4*>(dx);
    if (p->R == 3) dwconv_dgrad_s2_v4_kernel<3><<<grid_for(total4, 256, 64), 256, 0, st>>>(*p, g4, w4, d4, total4);
    else if (p->R == 5) dwconv_dgrad_s2_v4_kernel<5><<<grid_for(total4, 256, 64), 256, 0, st>>>(*p, g4, w4, d4, total4);
    else dwconv_dgrad_s2_v4_kernel<7><<<grid_for(total4, 256, 64), 256, 0, st>>>(*p, g4, w4, d4, total4);
    DTF_CHECK(cudaGetLastError());
    return 0;
  }
  const unsigned total = (unsigned)((unsigned long long)p->N * p->H * p->W * p->C);
  dwconv_dgrad_generic_kernel<<<grid_for(total), 256, 0, st>>>(*p, dy, w, dx, total);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_dwconv_wgrad(const dtf_dwconv_params* p, const float* x, const float* dy, float* dw,
                                void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  if (int rc = dw_check(p)) return rc;
  if (p->mult == 1 && p->R == p->S && p->sh == p->sw && (p->sh == 1 || p->sh == 2) &&
      (p->R == 3 || p->R == 5 || p->R == 7)) {
    const int cg = ceil_div(p->C, 32);
    DTF_REQUIRE(cg <= 65535, "depthwise wgrad: grid too large");
    // rows per sweep: 4 (2 for 7x7: registers) while that still leaves ~8 warps per SM, else fewer
    int rb = p->R == 7 ? 2 : 4;
    while (rb > 1 && (long long)p->N * ceil_div(p->OH, rb) * cg < 8LL * num_sms()) rb >>= 1;
    const int ngroups = p->N * ceil_div(p->OH, rb);
    long long want_blocks = (4LL * num_sms() + cg - 1) / cg;          // ~4 blocks per SM
    int gpw = (int)((ngroups + want_blocks * 8 - 1) / (want_blocks * 8));
    if (gpw < 1) gpw = 1;
    const dim3 grid((unsigned)ceil_div(ngroups, 8 * gpw), (unsigned)cg, 1);
#define DTF_DW_WGRAD(K_, SW_, RB_) dwconv_wgrad_rows_kernel<K_, SW_, RB_><<<grid, 256, 0, st>>>(*p, x, dy, dw, gpw)
#define DTF_DW_WGRAD_RB(K_, SW_)                                   \
  do {                                                             \
    if (rb == 4) DTF_DW_WGRAD(K_, SW_, 4);                         \
    else if (rb == 2) DTF_DW_WGRAD(K_, SW_, 2);                    \
    else DTF_DW_WGRAD(K_, SW_, 1);                                 \
  } while (0)
    if (p->R == 3 && p->sh == 1) DTF_DW_WGRAD_RB(3, 1);
    else if (p->R == 3) DTF_DW_WGRAD_RB(3, 2);
    else if (p->R == 5 && p->sh == 1) DTF_DW_WGRAD_RB(5, 1);
    else if (p->R == 5) DTF_DW_WGRAD_RB(5, 2);
    else if (p->sh == 1) { if (rb == 2) DTF_DW_WGRAD(7, 1, 2); else DTF_DW_WGRAD(7, 1, 1); }
    else { if (rb == 2) DTF_DW_WGRAD(7, 2, 2); else DTF_DW_WGRAD(7, 2, 1); }
#undef DTF_DW_WGRAD_RB
#undef DTF_DW_WGRAD
    DTF_CHECK(cudaGetLastError());
    return 0;
  }
  const long long npix = (long long)p->N * p->OH * p->OW;
  const int cgroups = ceil_div(p->C * p->mult, 32);
  const bool rowwise = p->S == 3 || p->S == 5 || p->S == 7;
  const int gz = rowwise ? p->R : p->R * p->S;
  DTF_REQUIRE(gz <= 65535 && cgroups <= 65535, "depthwise wgrad: grid too large");
  long long chunks = (8LL * num_sms() + (long long)cgroups * gz - 1) / ((long long)cgroups * gz);
  if (chunks < 1) chunks = 1;
  long long ppb = (npix + chunks - 1) / chunks;
  if (ppb < 64) ppb = 64;
  chunks = (npix + ppb - 1) / ppb;
  const dim3 grid((unsigned)chunks, (unsigned)cgroups, (unsigned)gz);
  if (p->S == 3) dwconv_wgrad_kernel<3><<<grid, 256, 0, st>>>(*p, x, dy, dw, (int)ppb);
  else if (p->S == 5) dwconv_wgrad_kernel<5><<<grid, 256, 0, st>>>(*p, x, dy, dw, (int)ppb);
  else if (p->S == 7) dwconv_wgrad_kernel<7><<<grid, 256, 0, st>>>(*p, x, dy, dw, (int)ppb);
  else dwconv_wgrad_kernel<0><<<grid, 256, 0, st>>>(*p, x, dy, dw, (int)ppb);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_act_fwd(const float* x, float* y, long long n, int mode, float alpha, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(mode >= 0 && mode <= 3, "activation mode: 0 relu6, 1 leaky-relu, 2 tanh, 3 sigmoid");
  DTF_REQUIRE(aligned16(x) && aligned16(y), "activation kernels need 16-byte aligned tensors");
  if (n <= 0) return 0;
  act_fwd_kernel<<<grid_for((unsigned long long)(n + 3) / 4), 256, 0, st>>>(x, y, n, mode, alpha);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_act_bwd(const float* ref, const float* dy, float* dx, long long n, int mode, float alpha,
                           void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(mode >= 0 && mode <= 3, "activation mode: 0 relu6, 1 leaky-relu, 2 tanh, 3 sigmoid");
  DTF_REQUIRE(aligned16(ref) && aligned16(dy) && aligned16(dx), "activation kernels need 16-byte aligned tensors");
  if (n <= 0) return 0;
  act_bwd_kernel<<<grid_for((unsigned long long)(n + 3) / 4), 256, 0, st>>>(ref, dy, dx, n, mode, alpha);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

static int instnorm_split(int N, int HW, int C, int* ppb) {
  const int cgroups = ceil_div(C, 32);
  long long splits = (4LL * num_sms() + (long long)cgroups * N - 1) / ((long long)cgroups * N);
  if (splits < 1) splits = 1;
  long long per = (HW + splits - 1) / splits;
  if (per < 64) per = 64;
  *ppb = (int)per;
  return (int)((HW + per - 1) / per);
}

extern "C" int dtf_instnorm_fwd(const float* x, const float* gamma, const float* beta, float* y, float* sums, int N,
                                int HW, int C, float eps, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(N > 0 && HW > 0 && C > 0 && N <= 65535, "instance norm: bad shape");
  DTF_REQUIRE((unsigned long long)N * HW * C < 0xffffffffull, "instance norm: tensor too large for 32-bit indexing");
  DTF_CHECK(cudaMemsetAsync(sums, 0, sizeof(float) * 2 * (size_t)N * C, st));
  int ppb = 0;
  const int splits = instnorm_split(N, HW, C, &ppb);
  instnorm_reduce_kernel<false><<<dim3(ceil_div(C, 32), splits, N), 256, 0, st>>>(x, x, nullptr, sums, HW, C, eps, ppb);
  DTF_CHECK(cudaGetLastError());
  const unsigned total = (unsigned)((unsigned long long)N * HW * C);
  instnorm_apply_kernel<false><<<grid_for(total), 256, 0, st>>>(x, nullptr, sums, nullptr, gamma, beta, y, HW, C, eps,
                                                               total);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_instnorm_bwd(const float* x, const float* dy, const float* gamma, const float* fwd_sums,
                                float* bwd_sums, float* dx, int N, int HW, int C, float eps, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(N > 0 && HW > 0 && C > 0 && N <= 65535, "instance norm: bad shape");
  DTF_REQUIRE((unsigned long long)N * HW * C < 0xffffffffull, "instance norm: tensor too large for 32-bit indexing");
  DTF_CHECK(cudaMemsetAsync(bwd_sums, 0, sizeof(float) * 2 * (size_t)N * C, st));
  int ppb = 0;
  const int splits = instnorm_split(N, HW, C, &ppb);
  instnorm_reduce_kernel<true><<<dim3(ceil_div(C, 32), splits, N), 256, 0, st>>>(dy, x, fwd_sums, bwd_sums, HW, C, eps,
                                                                                ppb);
  DTF_CHECK(cudaGetLastError());
  const unsigned total = (unsigned)((unsigned long long)N * HW * C);
  instnorm_apply_kernel<true><<<grid_for(total), 256, 0, st>>>(x, dy, fwd_sums, bwd_sums, gamma, nullptr, dx, HW, C,
                                                              eps, total);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

static int resize_geom(ResizeGeom* g, int N, int H, int W, int C, int OH, int OW, int bilinear, int align) {
  DTF_REQUIRE(N > 0 && H > 0 && W > 0 && C > 0 && OH > 0 && OW > 0, "resize: bad shape");
  DTF_REQUIRE((unsigned long long)N * OH * OW * C < 0xffffffffull && (unsigned long long)N * H * W * C < 0xffffffffull,
              "resize: tensor too large for 32-bit indexing");
  g->N = N; g->H = H; g->W = W; g->C = C; g->OH = OH; g->OW = OW; g->bilinear = bilinear; g->align = align;
  g->sch = (align && OH > 1) ? (float)(H - 1) / (float)(OH - 1) : (float)H / (float)OH;
  g->scw = (align && OW > 1) ? (float)(W - 1) / (float)(OW - 1) : (float)W / (float)OW;
  return 0;
}

extern "C" int dtf_resize_fwd(const float* x, float* y, int N, int H, int W, int C, int OH, int OW, int bilinear,
                              int align_corners, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  ResizeGeom g;
  if (int rc = resize_geom(&g, N, H, W, C, OH, OW, bilinear, align_corners)) return rc;
  const unsigned total = (unsigned)((unsigned long long)N * OH * OW * C);
  resize_kernel<false><<<grid_for(total), 256, 0, st>>>(g, x, y, total);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_resize_bwd(const float* dy, float* dx, int N, int H, int W, int C, int OH, int OW, int bilinear,
                              int align_corners, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  ResizeGeom g;
  if (int rc = resize_geom(&g, N, H, W, C, OH, OW, bilinear, align_corners)) return rc;
  DTF_CHECK(cudaMemsetAsync(dx, 0, sizeof(float) * (size_t)N * H * W * C, st));
  const unsigned total = (unsigned)((unsigned long long)N * OH * OW * C);
  resize_kernel<true><<<grid_for(total), 256, 0, st>>>(g, dy, dx, total);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

static int pad_geom(PadGeom* g, int N, int H, int W, int C, int pt, int pb, int pl, int pr, int mode) {
  DTF_REQUIRE(N > 0 && H > 0 && W > 0 && C > 0 && pt >= 0 && pb >= 0 && pl >= 0 && pr >= 0, "pad: bad shape");
  DTF_REQUIRE(mode == 0 || mode == 1, "pad mode: 0 zeros, 1 reflect");
  DTF_REQUIRE(mode == 0 || (pt < H && pb < H && pl < W && pr < W), "reflect padding must be smaller than the image");
  g->N = N; g->H = H; g->W = W; g->C = C; g->pt = pt; g->pl = pl; g->OH = H + pt + pb; g->OW = W + pl + pr;
  g->mode = mode;
  DTF_REQUIRE((unsigned long long)N * g->OH * g->OW * C < 0xffffffffull, "pad: tensor too large for 32-bit indexing");
  return 0;
}

extern "C" int dtf_pad_fwd(const float* x, float* y, int N, int H, int W, int C, int pt, int pb, int pl, int pr,
                           int mode, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  PadGeom g;
  if (int rc = pad_geom(&g, N, H, W, C, pt, pb, pl, pr, mode)) return rc;
  const unsigned total = (unsigned)((unsigned long long)N * g.OH * g.OW * C);
  pad_kernel<false><<<grid_for(total), 256, 0, st>>>(g, x, y, total);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_pad_bwd(const float* dy, float* dx, int N, int H, int W, int C, int pt, int pb, int pl, int pr,
                           int mode, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  PadGeom g;
  if (int rc = pad_geom(&g, N, H, W, C, pt, pb, pl, pr, mode)) return rc;
  DTF_CHECK(cudaMemsetAsync(dx, 0, sizeof(float) * (size_t)N * H * W * C, st));
  const unsigned total = (unsigned)((unsigned long long)N * g.OH * g.OW * C);
  pad_kernel<true><<<grid_for(total), 256, 0, st>>>(g, dy, dx, total);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_fake_quant_fwd(const float* x, float* y, long long n, const float* range, const float* gate,
                                  int bits, int narrow_range, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(bits >= 2 && bits <= 16, "fake quantisation: 2 to 16 bits");
  DTF_REQUIRE(aligned16(x) && aligned16(y), "fake-quant kernels need 16-byte aligned tensors");
  if (n <= 0) return 0;
  fake_quant_fwd_kernel<<<grid_for((unsigned long long)(n + 3) / 4), 256, 0, st>>>(x, y, n, range, gate, bits,
                                                                                   narrow_range);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_fake_quant_bwd(const float* x, const float* dy, float* dx, long long n, const float* range,
                                  const float* gate, int bits, int narrow_range, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(bits >= 2 && bits <= 16, "fake quantisation: 2 to 16 bits");
  DTF_REQUIRE(aligned16(x) && aligned16(dy) && aligned16(dx), "fake-quant kernels need 16-byte aligned tensors");
  if (n <= 0) return 0;
  fake_quant_bwd_kernel<<<grid_for((unsigned long long)(n + 3) / 4), 256, 0, st>>>(x, dy, dx, n, range, gate, bits,
                                                                                   narrow_range);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_quant_range_update(const float* x, long long n, float* mm, float* range, float decay,
                                      void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(aligned16(x), "min/max kernel needs a 16-byte aligned tensor");
  if (n <= 0) return 0;
  minmax_kernel<<<grid_for((unsigned long long)(n + 3) / 4, 256, 4), 256, 0, st>>>(x, n, mm);
  DTF_CHECK(cudaGetLastError());
  quant_range_kernel<<<1, 32, 0, st>>>(mm, range, decay);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

// CIFAR-10 input pipeline, native: the training distortion of cnn/cifar10_input.py:176-190 (random crop, random
// flip, random brightness, random contrast, per-image standardisation) and the evaluation preprocessing of
// cnn/cifar10_input.py:242-247 (central crop, standardisation) as ONE pass from the raw uint8 records to the fp32
// NHWC batch.  In the reference this work is TensorFlow's C++ input ops behind 16 queue-runner threads; here it is
//
//   dtf_cifar_distort        a CUDA kernel, one thread block per image.  The whole data set (150 MB of uint8
//                            records) stays resident in HBM; a step uploads the batch's record indices and random
//                            draws (a few hundred bytes) and the batch is produced where it is consumed — no
//                            per-step image traffic over PCIe and no host cores in the loop.
//   dtf_cifar_distort_host   the same arithmetic on host threads for the CPU plumbing mode and as the oracle.
//
// Arithmetic (shared by both, see cifar_stats): with the cropped / flipped pixels x, per-channel means m_c and the
// image mean M, brightness cancels in the standardisation and contrast f leaves
//     y = (x - m_c) f + (m_c - M),     out = y / max(sqrt(mean(y^2)), 1/sqrt(n)),    n = S*S*3
// and, x being bytes, sum(x) and sum(x^2) per channel are exact integers, so
//     mean(y^2) = (1/n) sum_c [ f^2 (Sxx_c - P m_c^2) + P (m_c - M)^2 ],              P = S*S
// needs one pass over the bytes; the second pass writes the floats.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <thread>
#include <vector>

#include "host_common.h"

#if defined(__SSE2__)
#include <emmintrin.h>
#endif

namespace dtf {

constexpr int kSrc = 32;                 // CIFAR record: 32 x 32 x 3 bytes, HWC
constexpr int kRecord = kSrc * kSrc * 3;

struct CifarCoef { float mul, mean[3], add[3]; };   // out = (x - mean[c]) * mul + add[c]

// sums: per-channel sum(x) and sum(x^2) over the P = S*S cropped pixels (exact integers)
__host__ __device__ inline CifarCoef cifar_stats(const unsigned long long sx[3], const unsigned long long sxx[3],
                                                 int S, float factor) {
  const double P = (double)S * S, n = 3.0 * P;
  double m[3], M = 0.0;
  for (int c = 0; c < 3; ++c) { m[c] = (double)sx[c] / P; M += m[c]; }
  M /= 3.0;
  const double f = (double)factor;
  double ss = 0.0;
  for (int c = 0; c < 3; ++c) {
    double centred = (double)sxx[c] - P * m[c] * m[c];
    if (centred < 0.0) centred = 0.0;
    ss += f * f * centred + P * (m[c] - M) * (m[c] - M);
  }
  double sd = sqrt(ss / n);
  const double floor_sd = 1.0 / sqrt(n);
  if (sd < floor_sd) sd = floor_sd;
  const double inv = 1.0 / sd;
  CifarCoef k;
  k.mul = (float)(f * inv);
  for (int c = 0; c < 3; ++c) { k.mean[c] = (float)m[c]; k.add[c] = (float)((m[c] - M) * inv); }
  return k;
}

// ---------------------------------------------------------------------------------------------------------
// device
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
cifar_distort_kernel(const uint8_t* __restrict__ images, const long long* __restrict__ idx,
                     const int* __restrict__ oy, const int* __restrict__ ox, const uint8_t* __restrict__ flip,
                     const float* __restrict__ factor, int S, float* __restrict__ out) {
  const int b = blockIdx.x;
  const uint8_t* rec = images + (size_t)idx[b] * kRecord;
  const int y0 = oy[b], x0 = ox[b];
  const bool fl = flip[b] != 0;
  const int P = S * S;
  unsigned int sx[3] = {0u, 0u, 0u}, sxx[3] = {0u, 0u, 0u};        // <= 1024 * 255^2 < 2^32
  for (int p = threadIdx.x; p < P; p += 128) {
    const int yy = p / S, xx = p - yy * S;
    const uint8_t* px = rec + ((y0 + yy) * kSrc + x0 + (fl ? S - 1 - xx : xx)) * 3;
#pragma unroll
    for (int c = 0; c < 3; ++c) { const unsigned int v = px[c]; sx[c] += v; sxx[c] += v * v; }
  }
  __shared__ unsigned int s_part[4][6];
  __shared__ CifarCoef s_coef;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      sx[c] += __shfl_xor_sync(0xffffffffu, sx[c], o);
      sxx[c] += __shfl_xor_sync(0xffffffffu, sxx[c], o);
    }
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
#pragma unroll
    for (int c = 0; c < 3; ++c) { s_part[warp][c] = sx[c]; s_part[warp][3 + c] = sxx[c]; }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long tx[3], txx[3];
    for (int c = 0; c < 3; ++c) {
      tx[c] = (unsigned long long)s_part[0][c] + s_part[1][c] + s_part[2][c] + s_part[3][c];
      txx[c] = (unsigned long long)s_part[0][3 + c] + s_part[1][3 + c] + s_part[2][3 + c] + s_part[3][3 + c];
    }
    s_coef = cifar_stats(tx, txx, S, factor[b]);
  }
  __syncthreads();
  const float mul = s_coef.mul;          // mean / add stay in shared memory: they are indexed by the channel
  float* dst = out + (size_t)b * P * 3;
  for (int e = threadIdx.x; e < P * 3; e += 128) {
    const int p = e / 3, c = e - p * 3;
    const int yy = p / S, xx = p - yy * S;
    const uint8_t v = rec[((y0 + yy) * kSrc + x0 + (fl ? S - 1 - xx : xx)) * 3 + c];
    dst[e] = ((float)v - s_coef.mean[c]) * mul + s_coef.add[c];
  }
}

// ---------------------------------------------------------------------------------------------------------
// host
// ---------------------------------------------------------------------------------------------------------
static void cifar_distort_one(const uint8_t* rec, int y0, int x0, bool fl, float factor, int S, float* dst) {
  unsigned long long sx[3] = {0, 0, 0}, sxx[3] = {0, 0, 0};
  for (int yy = 0; yy < S; ++yy) {
    const uint8_t* row = rec + ((y0 + yy) * kSrc + x0) * 3;
    for (int i = 0; i < S * 3; i += 3) {
      const unsigned int r = row[i], g = row[i + 1], bl = row[i + 2];
      sx[0] += r; sx[1] += g; sx[2] += bl;
      sxx[0] += r * r; sxx[1] += g * g; sxx[2] += bl * bl;
    }
  }
  const CifarCoef k = cifar_stats(sx, sxx, S, factor);
  for (int yy = 0; yy < S; ++yy) {
    const uint8_t* row = rec + ((y0 + yy) * kSrc + x0) * 3;
    float* o = dst + (size_t)yy * S * 3;
    if (!fl) {
      for (int i = 0; i < S * 3; i += 3) {
        o[i] = ((float)row[i] - k.mean[0]) * k.mul + k.add[0];
        o[i + 1] = ((float)row[i + 1] - k.mean[1]) * k.mul + k.add[1];
        o[i + 2] = ((float)row[i + 2] - k.mean[2]) * k.mul + k.add[2];
      }
    } else {
      for (int xx = 0; xx < S; ++xx) {
        const uint8_t* px = row + (S - 1 - xx) * 3;
        o[xx * 3] = ((float)px[0] - k.mean[0]) * k.mul + k.add[0];
        o[xx * 3 + 1] = ((float)px[1] - k.mean[1]) * k.mul + k.add[1];
        o[xx * 3 + 2] = ((float)px[2] - k.mean[2]) * k.mul + k.add[2];
      }
    }
  }
}

static int cifar_check_geometry(int B, int S) {
  DTF_REQUIRE(B >= 0, "batch size must not be negative");
  DTF_REQUIRE(S >= 1 && S <= kSrc, "CIFAR crop size must be between 1 and 32");
  return 0;
}

}  // namespace dtf

using namespace dtf;

extern "C" int dtf_cifar_distort(const uint8_t* images, const long long* idx, const int* oy, const int* ox,
                                 const uint8_t* flip, const float* factor, int B, int S, float* out, void* stream_) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream_);
  if (int rc = cifar_check_geometry(B, S)) return rc;
  if (B == 0) return 0;
  cifar_distort_kernel<<<B, 128, 0, st>>>(images, idx, oy, ox, flip, factor, S, out);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_cifar_distort_host(const uint8_t* images, long long num_records, const long long* idx,
                                      const int* oy, const int* ox, const uint8_t* flip, const float* factor, int B,
                                      int S, float* out, int threads) {
  if (int rc = cifar_check_geometry(B, S)) return rc;
  for (int b = 0; b < B; ++b) {
    DTF_REQUIRE(idx[b] >= 0 && idx[b] < num_records, "record index out of range");
    DTF_REQUIRE(oy[b] >= 0 && ox[b] >= 0 && oy[b] + S <= kSrc && ox[b] + S <= kSrc, "crop window leaves the image");
  }
  if (threads < 1) threads = 1;
  if (threads > B) threads = B > 0 ? B : 1;
  auto work = [&](int t) {
    for (int b = t; b < B; b += threads)
      cifar_distort_one(images + (size_t)idx[b] * kRecord, oy[b], ox[b], flip[b] != 0, factor[b], S,
                        out + (size_t)b * S * S * 3);
  };
  if (threads == 1) { work(0); return 0; }
  std::vector<std::thread> pool;
  pool.reserve(threads - 1);
  for (int t = 1; t < threads; ++t) pool.emplace_back(work, t);
  work(0);
  for (auto& th : pool) th.join();
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// ImageNet colour distortion (inception/image_processing.py:164-195) on an HWC float image in [0, 1], in place:
//   order 0: brightness -> saturation + hue -> contrast        (even preprocessing threads)
//   order 1: brightness -> contrast -> saturation + hue        (odd threads)
// then clip to [0, 1].  Saturation and hue share one RGB -> HSV -> RGB round trip on the image clipped to [0, 1];
// contrast is (x - m_c) * factor + m_c with the per-channel mean of its input, which costs one extra sweep:
// order 0 sums the channels while it writes the hue stage, order 1 sums the raw image first (the brightness shift
// moves the mean by delta).  Two sweeps over the image in total instead of a dozen array temporaries.
// ---------------------------------------------------------------------------------------------------------
namespace dtf {

static inline float fmin2(float a, float b) { return a < b ? a : b; }     // minss / maxss, not libm calls
static inline float fmax2(float a, float b) { return a > b ? a : b; }
static inline float clip01(float v) { return fmin2(fmax2(v, 0.f), 1.f); }

constexpr int kColourBlock = 256;        // pixels handled per inner loop, de-interleaved into planes on the stack

// Saturation + hue on n <= kColourBlock pixels held as three planes.  Selects only: on photographs the comparisons are
// data dependent and branches would mispredict.  Four pixels per iteration with SSE2 (part of x86-64); the scalar
// loop does the tail and other hosts.
static inline void saturation_hue_scalar(float& R, float& G, float& B, float sat, float hue) {
  const float r = clip01(R), g = clip01(G), b = clip01(B);
  const float mx = fmax2(fmax2(r, g), b);
  const float mn = fmin2(fmin2(r, g), b);
  const float d = mx - mn;
  const float inv = 1.0f / fmax2(d, 1e-12f);
  const float s0 = mx > 0.f ? d / fmax2(mx, 1e-12f) : 0.f;
  float hr = (g - b) * inv;
  hr = hr < 0.f ? hr + 6.0f : hr;
  const float hg = (b - r) * inv + 2.0f, hb = (r - g) * inv + 4.0f;
  float h = mx == r ? hr : (mx == g ? hg : hb);
  h = d > 0.f ? h * (1.0f / 6.0f) : 0.f;
  h = h >= 1.f ? h - 1.0f : h;
  const float s = clip01(s0 * sat);
  h += hue;                                   // |hue| <= 0.2: h stays within [-1, 2)
  h = h < 0.f ? h + 1.0f : (h >= 1.f ? h - 1.0f : h);
  const float h6 = h * 6.0f, vs = mx * s;
  float k5 = h6 + 5.0f, k3 = h6 + 3.0f, k1 = h6 + 1.0f;
  k5 = k5 >= 6.f ? k5 - 6.0f : k5;
  k3 = k3 >= 6.f ? k3 - 6.0f : k3;
  k1 = k1 >= 6.f ? k1 - 6.0f : k1;
  R = mx - vs * clip01(fmin2(k5, 4.0f - k5));
  G = mx - vs * clip01(fmin2(k3, 4.0f - k3));
  B = mx - vs * clip01(fmin2(k1, 4.0f - k1));
}

#if defined(__SSE2__)
static inline __m128 sel(__m128 mask, __m128 a, __m128 b) { return _mm_or_ps(_mm_and_ps(mask, a), _mm_andnot_ps(mask, b)); }
static inline __m128 clip01v(__m128 v) { return _mm_min_ps(_mm_max_ps(v, _mm_setzero_ps()), _mm_set1_ps(1.0f)); }
// k -> clip01(min(k', 4 - k')), k' = k wrapped into [0, 6)
static inline __m128 hue_ramp(__m128 k) {
  const __m128 six = _mm_set1_ps(6.0f);
  k = sel(_mm_cmpge_ps(k, six), _mm_sub_ps(k, six), k);
  return clip01v(_mm_min_ps(k, _mm_sub_ps(_mm_set1_ps(4.0f), k)));
}
#endif

static void saturation_hue_block(float* __restrict__ R, float* __restrict__ G, float* __restrict__ B, int n, float sat,
                                 float hue) {
  int i = 0;
#if defined(__SSE2__)
  const __m128 zero = _mm_setzero_ps(), one = _mm_set1_ps(1.0f), tiny = _mm_set1_ps(1e-12f);
  for (; i + 4 <= n; i += 4) {
    const __m128 r = clip01v(_mm_loadu_ps(R + i)), g = clip01v(_mm_loadu_ps(G + i)), b = clip01v(_mm_loadu_ps(B + i));
    const __m128 mx = _mm_max_ps(_mm_max_ps(r, g), b);
    const __m128 mn = _mm_min_ps(_mm_min_ps(r, g), b);
    const __m128 d = _mm_sub_ps(mx, mn);
    const __m128 inv = _mm_div_ps(one, _mm_max_ps(d, tiny));
    const __m128 s0 = _mm_and_ps(_mm_cmpgt_ps(mx, zero), _mm_div_ps(d, _mm_max_ps(mx, tiny)));
    __m128 hr = _mm_mul_ps(_mm_sub_ps(g, b), inv);
    hr = sel(_mm_cmplt_ps(hr, zero), _mm_add_ps(hr, _mm_set1_ps(6.0f)), hr);
    const __m128 hg = _mm_add_ps(_mm_mul_ps(_mm_sub_ps(b, r), inv), _mm_set1_ps(2.0f));
    const __m128 hb = _mm_add_ps(_mm_mul_ps(_mm_sub_ps(r, g), inv), _mm_set1_ps(4.0f));
    __m128 h = sel(_mm_cmpeq_ps(mx, r), hr, sel(_mm_cmpeq_ps(mx, g), hg, hb));
    h = _mm_and_ps(_mm_cmpgt_ps(d, zero), _mm_mul_ps(h, _mm_set1_ps(1.0f / 6.0f)));
    h = sel(_mm_cmpge_ps(h, one), _mm_sub_ps(h, one), h);
    const __m128 s = clip01v(_mm_mul_ps(s0, _mm_set1_ps(sat)));
    h = _mm_add_ps(h, _mm_set1_ps(hue));
    h = sel(_mm_cmplt_ps(h, zero), _mm_add_ps(h, one), sel(_mm_cmpge_ps(h, one), _mm_sub_ps(h, one), h));
    const __m128 h6 = _mm_mul_ps(h, _mm_set1_ps(6.0f)), vs = _mm_mul_ps(mx, s);
    _mm_storeu_ps(R + i, _mm_sub_ps(mx, _mm_mul_ps(vs, hue_ramp(_mm_add_ps(h6, _mm_set1_ps(5.0f))))));
    _mm_storeu_ps(G + i, _mm_sub_ps(mx, _mm_mul_ps(vs, hue_ramp(_mm_add_ps(h6, _mm_set1_ps(3.0f))))));
    _mm_storeu_ps(B + i, _mm_sub_ps(mx, _mm_mul_ps(vs, hue_ramp(_mm_add_ps(h6, _mm_set1_ps(1.0f))))));
  }
#endif
  for (; i < n; ++i) saturation_hue_scalar(R[i], G[i], B[i], sat, hue);
}

}  // namespace dtf

extern "C" int dtf_distort_color_host(float* img, long long num_pixels, int order, float delta, float sat, float hue,
                                      float contrast) {
  DTF_REQUIRE(order == 0 || order == 1, "colour distortion order: 0 (contrast last) or 1 (contrast second)");
  DTF_REQUIRE(num_pixels >= 0, "pixel count must not be negative");
  if (num_pixels == 0) return 0;
  double sum[3] = {0.0, 0.0, 0.0};
  float m[3] = {0.f, 0.f, 0.f};
  float R[kColourBlock], G[kColourBlock], B[kColourBlock];
  if (order == 1) {                      // contrast acts on the brightness-shifted image: its means are known up front
    for (long long p = 0; p < num_pixels; ++p) {
      const float* px = img + 3 * p;
      sum[0] += px[0]; sum[1] += px[1]; sum[2] += px[2];
    }
    for (int c = 0; c < 3; ++c) m[c] = (float)(sum[c] / (double)num_pixels) + delta;
  }
  for (long long p0 = 0; p0 < num_pixels; p0 += kColourBlock) {
    const int n = (int)(num_pixels - p0 < kColourBlock ? num_pixels - p0 : kColourBlock);
    float* px = img + 3 * p0;
    if (order == 0) {
      for (int i = 0; i < n; ++i) { R[i] = px[3 * i] + delta; G[i] = px[3 * i + 1] + delta; B[i] = px[3 * i + 2] + delta; }
    } else {
      for (int i = 0; i < n; ++i) {
        R[i] = (px[3 * i] + delta - m[0]) * contrast + m[0];
        G[i] = (px[3 * i + 1] + delta - m[1]) * contrast + m[1];
        B[i] = (px[3 * i + 2] + delta - m[2]) * contrast + m[2];
      }
    }
    saturation_hue_block(R, G, B, n, sat, hue);
    if (order == 0) {
      float sr = 0.f, sg = 0.f, sb = 0.f;                 // block sums in float, image sums in double
      for (int i = 0; i < n; ++i) {
        sr += R[i]; sg += G[i]; sb += B[i];
        px[3 * i] = R[i]; px[3 * i + 1] = G[i]; px[3 * i + 2] = B[i];
      }
      sum[0] += sr; sum[1] += sg; sum[2] += sb;
    } else {
      for (int i = 0; i < n; ++i) { px[3 * i] = clip01(R[i]); px[3 * i + 1] = clip01(G[i]); px[3 * i + 2] = clip01(B[i]); }
    }
  }
  if (order == 0) {
    for (int c = 0; c < 3; ++c) m[c] = (float)(sum[c] / (double)num_pixels);
    for (long long p = 0; p < num_pixels; ++p) {
      float* px = img + 3 * p;
      px[0] = clip01((px[0] - m[0]) * contrast + m[0]);
      px[1] = clip01((px[1] - m[1]) * contrast + m[1]);
      px[2] = clip01((px[2] - m[2]) * contrast + m[2]);
    }
  }
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// Crop + resize (+ left-right flip) of a decoded uint8 HWC image to a float HWC image in [0, 1]
// (inception/image_processing.py:233-244: the distorted bounding-box crop resized with one of four methods, and
// :271-274: the central crop resized bilinearly).  Every method is a separable linear filter, so one routine serves
// them all: per output column / row a short list of (source index, weight) taps, a horizontal pass over the crop's
// rows into a float strip, a vertical pass out of it.  Tap rules are those of torch.nn.functional.interpolate
// (the Python fallback): 0 bilinear with half-pixel centres, 1 nearest, 2 bicubic (a = -0.75, edge taps clamped),
// 3 area = mean over [floor(o in/out), ceil((o+1) in/out)).
// ---------------------------------------------------------------------------------------------------------
namespace dtf {

struct Taps {
  int max_taps = 0;
  std::vector<int> idx;        // [out][max_taps]
  std::vector<float> w;        // [out][max_taps]
  std::vector<int> count;      // [out]
};

static inline float cubic1(float x, float A) { return ((A + 2.f) * x - (A + 3.f)) * x * x + 1.f; }
static inline float cubic2(float x, float A) { return ((A * x - 5.f * A) * x + 8.f * A) * x - 4.f * A; }

static void make_taps(int in, int out, int method, bool flip, Taps* t) {
  const int mt = method == 0 ? 2 : method == 1 ? 1 : method == 2 ? 4 : (in + out - 1) / out + 1;
  t->max_taps = mt;
  t->idx.assign((size_t)out * mt, 0);
  t->w.assign((size_t)out * mt, 0.f);
  t->count.assign(out, 0);
  const float scale = (float)in / (float)out;
  for (int o = 0; o < out; ++o) {
    const int dst = flip ? out - 1 - o : o;          // the flipped image's column o is the plain image's column out-1-o
    int* ix = &t->idx[(size_t)o * mt];
    float* w = &t->w[(size_t)o * mt];
    int n = 0;
    if (method == 0) {
      float src = scale * ((float)dst + 0.5f) - 0.5f;
      if (src < 0.f) src = 0.f;
      int i0 = (int)src;
      if (i0 > in - 1) i0 = in - 1;
      const int i1 = i0 < in - 1 ? i0 + 1 : i0;
      const float l1 = src - (float)i0;
      ix[0] = i0; w[0] = 1.f - l1; ix[1] = i1; w[1] = l1; n = 2;
    } else if (method == 1) {
      int i0 = (int)floorf((float)dst * scale);
      if (i0 > in - 1) i0 = in - 1;
      ix[0] = i0; w[0] = 1.f; n = 1;
    } else if (method == 2) {
      const float A = -0.75f;
      const float src = scale * ((float)dst + 0.5f) - 0.5f;
      const float fl = floorf(src);
      const float tt = src - fl;
      const int i0 = (int)fl;
      const float c[4] = {cubic2(tt + 1.f, A), cubic1(tt, A), cubic1(1.f - tt, A), cubic2(2.f - tt, A)};
      for (int k = 0; k < 4; ++k) {
        int i = i0 - 1 + k;
        i = i < 0 ? 0 : (i > in - 1 ? in - 1 : i);
        ix[k] = i; w[k] = c[k];
      }
      n = 4;
    } else {
      const int s = (int)(((long long)dst * in) / out);
      const int e = (int)((((long long)dst + 1) * in + out - 1) / out);
      const float inv = 1.f / (float)(e - s);
      for (int i = s; i < e && n < mt; ++i) { ix[n] = i; w[n] = inv; ++n; }
    }
    t->count[o] = n;
  }
}

}  // namespace dtf

extern "C" int dtf_crop_resize_host(const uint8_t* img, int H, int W, int y0, int x0, int ch, int cw, int out_h,
                                    int out_w, int method, int flip, float* out) {
  DTF_REQUIRE(H > 0 && W > 0 && ch > 0 && cw > 0 && out_h > 0 && out_w > 0, "image, crop and output sizes must be positive");
  DTF_REQUIRE(y0 >= 0 && x0 >= 0 && y0 + ch <= H && x0 + cw <= W, "crop window leaves the image");
  DTF_REQUIRE(method >= 0 && method <= 3, "resize method: 0 bilinear, 1 nearest, 2 bicubic, 3 area");
  Taps tx, ty;
  make_taps(cw, out_w, method, flip != 0, &tx);
  make_taps(ch, out_h, method, false, &ty);
  const float k255 = 1.0f / 255.0f;
  if (method == 1) {                     // one source pixel per output pixel: no strip needed
    for (int oy = 0; oy < out_h; ++oy) {
      const uint8_t* row = img + ((size_t)(y0 + ty.idx[oy]) * W + x0) * 3;
      float* orow = out + (size_t)oy * out_w * 3;
      for (int ox = 0; ox < out_w; ++ox) {
        const uint8_t* px = row + 3 * tx.idx[ox];
        orow[3 * ox] = (float)px[0] * k255; orow[3 * ox + 1] = (float)px[1] * k255; orow[3 * ox + 2] = (float)px[2] * k255;
      }
    }
    return 0;
  }
  // horizontal pass over the crop's rows: strip[y][ox][c]
  std::vector<float> strip((size_t)ch * out_w * 3);
  for (int y = 0; y < ch; ++y) {
    const uint8_t* row = img + ((size_t)(y0 + y) * W + x0) * 3;
    float* srow = &strip[(size_t)y * out_w * 3];
    for (int ox = 0; ox < out_w; ++ox) {
      const int* ix = &tx.idx[(size_t)ox * tx.max_taps];
      const float* w = &tx.w[(size_t)ox * tx.max_taps];
      float r = 0.f, g = 0.f, b = 0.f;
      for (int k = 0; k < tx.count[ox]; ++k) {
        const uint8_t* px = row + 3 * ix[k];
        r += w[k] * (float)px[0]; g += w[k] * (float)px[1]; b += w[k] * (float)px[2];
      }
      srow[3 * ox] = r * k255; srow[3 * ox + 1] = g * k255; srow[3 * ox + 2] = b * k255;
    }
  }
  // vertical pass
  const int row_len = out_w * 3;
  for (int oy = 0; oy < out_h; ++oy) {
    const int* iy = &ty.idx[(size_t)oy * ty.max_taps];
    const float* w = &ty.w[(size_t)oy * ty.max_taps];
    float* orow = out + (size_t)oy * row_len;
    const int n = ty.count[oy];
    {
      const float* s0 = &strip[(size_t)iy[0] * row_len];
      const float w0 = w[0];
      for (int i = 0; i < row_len; ++i) orow[i] = w0 * s0[i];
    }
    for (int k = 1; k < n; ++k) {
      const float* sk = &strip[(size_t)iy[k] * row_len];
      const float wk = w[k];
      for (int i = 0; i < row_len; ++i) orow[i] += wk * sk[i];
    }
  }
  return 0;
}

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
#include <cstdint>
#include <cstdio>
#include <thread>
#include <vector>

#include "host_common.h"

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
  const CifarCoef k = s_coef;
  float* dst = out + (size_t)b * P * 3;
  for (int e = threadIdx.x; e < P * 3; e += 128) {
    const int p = e / 3, c = e - p * 3;
    const int yy = p / S, xx = p - yy * S;
    const uint8_t v = rec[((y0 + yy) * kSrc + x0 + (fl ? S - 1 - xx : xx)) * 3 + c];
    dst[e] = ((float)v - k.mean[c]) * k.mul + k.add[c];
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

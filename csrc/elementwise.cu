// K6 LRN, K9 softmax cross-entropy, K10 dropout, K2 bias-gradient column sums.
//
// Reference ops replaced: tf.nn.lrn(depth_radius 4, bias 1, alpha 0.001/9, beta 0.75) (cnn/cifar10.py:231,245;
// cifarnet/cifarnet.py:68,71); tf.nn.sparse_softmax_cross_entropy_with_logits + reduce_mean
// (cnn/cifar10.py:299-301), slim cross_entropy_loss with label smoothing and weight (slim/losses.py:142-174);
// tf.nn.dropout (slim/ops.py:422); BiasAddGrad.
#include "../include/dtf_b200.h"
#include "host_common.h"

namespace dtf {

// ---------------------------------------------------------------------------------------------------------
// LRN across channels.  One block handles `ppb` pixels; a pixel's C channels sit in shared memory.
// fwd:  s_c = bias + alpha * sum_{|j-c|<=r} x_j^2 ;  y_c = x_c * s_c^-beta
// bwd:  dx_c = dy_c * s_c^-beta - 2*alpha*beta * x_c * sum_{|j-c|<=r} dy_j * x_j * s_j^(-beta-1)
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
lrn_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, long long M, int C, int radius, float bias,
               float alpha, float beta, int ppb) {
  extern __shared__ float sm[];                  // [ppb][C] squares
  const int lp = threadIdx.x / C, c = threadIdx.x % C;
  const long long pix = (long long)blockIdx.x * ppb + lp;
  const bool ok = lp < ppb && pix < M;
  float xv = 0.f;
  if (ok) { xv = x[pix * C + c]; sm[lp * C + c] = xv * xv; }
  __syncthreads();
  if (!ok) return;
  float ss = 0.f;
  const int lo = c - radius < 0 ? 0 : c - radius, hi = c + radius >= C ? C - 1 : c + radius;
  for (int j = lo; j <= hi; ++j) ss += sm[lp * C + j];
  y[pix * C + c] = xv * __powf(bias + alpha * ss, -beta);
}

__global__ void __launch_bounds__(256)
lrn_bwd_kernel(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ dx, long long M, int C,
               int radius, float bias, float alpha, float beta, int ppb) {
  extern __shared__ float sm[];                  // [ppb][C] squares, then [ppb][C] t_j = dy_j x_j s_j^(-beta-1)
  float* sq = sm;
  float* tt = sm + ppb * C;
  const int lp = threadIdx.x / C, c = threadIdx.x % C;
  const long long pix = (long long)blockIdx.x * ppb + lp;
  const bool ok = lp < ppb && pix < M;
  float xv = 0.f, dv = 0.f;
  if (ok) { xv = x[pix * C + c]; dv = dy[pix * C + c]; sq[lp * C + c] = xv * xv; }
  __syncthreads();
  float s = 1.f;
  const int lo = c - radius < 0 ? 0 : c - radius, hi = c + radius >= C ? C - 1 : c + radius;
  if (ok) {
    float ss = 0.f;
    for (int j = lo; j <= hi; ++j) ss += sq[lp * C + j];
    s = bias + alpha * ss;
    tt[lp * C + c] = dv * xv * __powf(s, -beta - 1.f);
  }
  __syncthreads();
  if (!ok) return;
  float acc = 0.f;
  for (int j = lo; j <= hi; ++j) acc += tt[lp * C + j];
  dx[pix * C + c] = dv * __powf(s, -beta) - 2.f * alpha * beta * xv * acc;
}

// ---------------------------------------------------------------------------------------------------------
// softmax cross-entropy, forward and gradient in one pass.  One warp per row.
//   target y' = onehot*(1-eps) + eps/C ;  loss += weight/N * (lse - sum_c y'_c z_c) ;
//   dlogits = weight/N * (softmax - y')
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
softmax_xent_kernel(const float* __restrict__ logits, const long long* __restrict__ labels, int N, int C,
                    float smoothing, float weight, float* __restrict__ loss, float* __restrict__ dlogits,
                    float* __restrict__ row_loss) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= N) return;
  const float* z = logits + (long long)warp * C;
  float m = -INFINITY;
  for (int c = lane; c < C; c += 32) m = fmaxf(m, z[c]);
#pragma unroll
  for (int o = 16; o; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  float se = 0.f, sz = 0.f;
  for (int c = lane; c < C; c += 32) { se += __expf(z[c] - m); sz += z[c]; }
#pragma unroll
  for (int o = 16; o; o >>= 1) { se += __shfl_xor_sync(0xffffffffu, se, o); sz += __shfl_xor_sync(0xffffffffu, sz, o); }
  const float lse = m + __logf(se);
  const int lab = (int)labels[warp];
  const float u = smoothing / (float)C;
  // sum_c y'_c z_c = (1-eps) z_lab + eps/C * sum z
  const float zl = (lab >= 0 && lab < C) ? z[lab] : 0.f;
  const float l = lse - ((1.f - smoothing) * zl + u * sz);
  const float wn = weight / (float)N;
  if (lane == 0) {
    atomicAdd(loss, l * wn);
    if (row_loss) row_loss[warp] = l;
  }
  if (dlogits) {
    float* d = dlogits + (long long)warp * C;
    const float inv = 1.f / se;
    for (int c = lane; c < C; c += 32) {
      const float p = __expf(z[c] - m) * inv;
      const float t = u + (c == lab ? 1.f - smoothing : 0.f);
      d[c] = wn * (p - t);
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// dropout with a counter-based generator: keep(i) = hash(seed, offset + i) < keep_prob.  The same (seed,
// offset) pair in the backward call regenerates the mask, so no mask tensor is stored.  `seed_ptr` lives in
// device memory so a captured CUDA graph sees a fresh stream of masks every replay.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float u01(unsigned long long seed, unsigned long long idx) {
  unsigned long long z = seed + 0x9E3779B97F4A7C15ull * (idx + 1);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z = z ^ (z >> 31);
  return (float)(z >> 40) * (1.0f / 16777216.0f);
}

__global__ void __launch_bounds__(256)
dropout_kernel(const float* __restrict__ x, float* __restrict__ y, long long n, float keep,
               const unsigned long long* __restrict__ seed_ptr, unsigned long long salt) {
  const unsigned long long seed = seed_ptr[0] * 0xD1342543DE82EF95ull + salt;
  const float inv = 1.f / keep;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    y[i] = u01(seed, (unsigned long long)i) < keep ? x[i] * inv : 0.f;
}

// ---------------------------------------------------------------------------------------------------------
// column sums of [M, C] (any C): out[c] += sum_m x[m, c]
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
colsum_kernel(const float* __restrict__ x, long long M, int C, int ld, int rows_per_block, float* __restrict__ out) {
  const long long row0 = (long long)blockIdx.x * rows_per_block;
  long long row1 = row0 + rows_per_block;
  if (row1 > M) row1 = M;
  // thread t: channel t % Cp, row slot t / Cp
  const int Cp = C < 256 ? C : 256;
  const int rslots = 256 / Cp;
  const int cs = threadIdx.x % Cp, rs = threadIdx.x / Cp;
  __shared__ float red[256];
  for (int c0 = 0; c0 < C; c0 += Cp) {
    const int c = c0 + cs;
    float a = 0.f;
    if (c < C && rs < rslots)
      for (long long r = row0 + rs; r < row1; r += rslots) a += __ldg(x + r * ld + c);
    red[threadIdx.x] = a;
    __syncthreads();
    if (rs == 0 && c < C) {
      for (int k = 1; k < rslots; ++k) a += red[k * Cp + cs];
      atomicAdd(out + c, a);
    }
    __syncthreads();
  }
}

}  // namespace dtf

using namespace dtf;

extern "C" int dtf_lrn_fwd(const float* x, float* y, long long M, int C, int radius, float bias, float alpha,
                           float beta, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(C <= 256, "LRN kernel keeps one pixel's channels in a 256-thread block");
  if (M == 0) return 0;
  const int ppb = 256 / C;
  lrn_fwd_kernel<<<(unsigned)((M + ppb - 1) / ppb), 256, ppb * C * sizeof(float), stream>>>(x, y, M, C, radius, bias,
                                                                                           alpha, beta, ppb);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_lrn_bwd(const float* x, const float* dy, float* dx, long long M, int C, int radius, float bias,
                           float alpha, float beta, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(C <= 256, "LRN kernel keeps one pixel's channels in a 256-thread block");
  if (M == 0) return 0;
  const int ppb = 256 / C;
  lrn_bwd_kernel<<<(unsigned)((M + ppb - 1) / ppb), 256, 2 * ppb * C * sizeof(float), stream>>>(
      x, dy, dx, M, C, radius, bias, alpha, beta, ppb);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_softmax_xent(const float* logits, const long long* labels, int N, int C, float smoothing,
                                float weight, float* loss, float* dlogits, float* row_loss, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  if (N == 0) return 0;
  DTF_CHECK(cudaMemsetAsync(loss, 0, sizeof(float), stream));
  const int warps_per_block = 8;
  softmax_xent_kernel<<<ceil_div(N, warps_per_block), warps_per_block * 32, 0, stream>>>(
      logits, labels, N, C, smoothing, weight, loss, dlogits, row_loss);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_dropout(const float* x, float* y, long long n, float keep_prob, const unsigned long long* seed_ptr,
                           unsigned long long salt, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  if (n == 0) return 0;
  long long b = (n + 255) / 256;
  const long long cap = 8LL * num_sms();
  dropout_kernel<<<(unsigned)(b < cap ? b : cap), 256, 0, stream>>>(x, y, n, keep_prob, seed_ptr, salt);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_colsum(const float* x, long long M, int C, int ld, float* out, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  if (M == 0 || C == 0) return 0;
  long long blocks = 2LL * num_sms();
  long long rpb = (M + blocks - 1) / blocks;
  if (rpb < 32) rpb = 32;
  const int grid = (int)((M + rpb - 1) / rpb);
  colsum_kernel<<<grid, 256, 0, stream>>>(x, M, C, ld, (int)rpb, out);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

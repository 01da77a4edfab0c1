// Epilogue shared by the tcgen05 convolution kernels: one thread owns one accumulator row (= output pixel).
//   tcgen05.ld 16 columns at a time -> + bias / + residual / ReLU -> global NHWC row, and either the batch-norm
//   statistics of the stored value or the fused batch-norm-backward sums, reduced across the warp's 32 rows with
//   shuffles and accumulated in shared memory (csum / csq, 256 floats each).
#pragma once
#include "../include/dtf_b200.h"
#include "umma.cuh"

namespace dtf {

// m: index of the output pixel this thread's row maps to (only dereferenced when mvalid).
// tr: optional timeline slots (debug): [0] accumulator in registers, [1] row stored, [2] statistics done.
__device__ __forceinline__ void conv_epilogue_row(const dtf_conv_gemm_params& p, int BN, int n0, bool mvalid, size_t m,
                                                  uint32_t t_row, int vec_out_ok, int vec_res_ok, bool do_stats,
                                                  float* csum, float* csq, int lane, long long* tr = nullptr) {
  float* orow = p.out + m * (size_t)p.out_ld;
  const float* rrow = p.residual ? p.residual + m * (size_t)p.res_ld : nullptr;
  const bool bias_vec = p.bias != nullptr && (reinterpret_cast<uintptr_t>(p.bias) & 15) == 0;
  for (int cc = 0; cc < BN; cc += 16) {
    const int nb = n0 + cc;
    if (nb >= p.Nout) break;    // uniform across the CTA
    float v[16];
    // Fast path (whole 16-column chunk in range, 16-byte aligned rows): straight-line code, the residual row is
    // requested before the accumulator is read so its latency overlaps the TMEM load.
    const bool fast = vec_out_ok && nb + 16 <= p.Nout && (rrow == nullptr || vec_res_ok) &&
                      (p.bias == nullptr || bias_vec);
    if (fast) {
      float4 r4[4];
      if (rrow != nullptr) {
#pragma unroll
        for (int q = 0; q < 4; ++q)
          r4[q] = mvalid ? __ldg(reinterpret_cast<const float4*>(rrow + nb + q * 4)) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
      tmem_ld16(t_row + (uint32_t)cc, v);
      if (tr != nullptr && cc == 0) tr[0] = clock64();
      if (p.bias != nullptr) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bias + nb + q * 4));
          v[q * 4 + 0] += b4.x; v[q * 4 + 1] += b4.y; v[q * 4 + 2] += b4.z; v[q * 4 + 3] += b4.w;
        }
      }
      if (rrow != nullptr) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          v[q * 4 + 0] += r4[q].x; v[q * 4 + 1] += r4[q].y; v[q * 4 + 2] += r4[q].z; v[q * 4 + 3] += r4[q].w;
        }
      }
      if (p.out_relu) {
#pragma unroll
        for (int e = 0; e < 16; ++e) v[e] = fmaxf(v[e], 0.f);
      }
      if (mvalid) {
#pragma unroll
        for (int q = 0; q < 4; ++q)
          *reinterpret_cast<float4*>(orow + nb + q * 4) = make_float4(v[q * 4 + 0], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
      } else {
#pragma unroll
        for (int e = 0; e < 16; ++e) v[e] = 0.f;       // rows outside the tensor do not count in the statistics
      }
    } else {
      tmem_ld16(t_row + (uint32_t)cc, v);
      if (tr != nullptr && cc == 0) tr[0] = clock64();
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int n = nb + q * 4;
        float4 x = make_float4(v[q * 4 + 0], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
        const bool full4 = n + 3 < p.Nout;
        if (p.bias) {
          if (full4) {
            x.x += __ldg(p.bias + n); x.y += __ldg(p.bias + n + 1);
            x.z += __ldg(p.bias + n + 2); x.w += __ldg(p.bias + n + 3);
          } else {
            if (n < p.Nout) x.x += __ldg(p.bias + n);
            if (n + 1 < p.Nout) x.y += __ldg(p.bias + n + 1);
            if (n + 2 < p.Nout) x.z += __ldg(p.bias + n + 2);
          }
        }
        if (rrow && mvalid) {
          if (full4 && vec_res_ok) {
            const float4 r4 = __ldg(reinterpret_cast<const float4*>(rrow + n));
            x.x += r4.x; x.y += r4.y; x.z += r4.z; x.w += r4.w;
          } else {
            if (n < p.Nout) x.x += __ldg(rrow + n);
            if (n + 1 < p.Nout) x.y += __ldg(rrow + n + 1);
            if (n + 2 < p.Nout) x.z += __ldg(rrow + n + 2);
            if (n + 3 < p.Nout) x.w += __ldg(rrow + n + 3);
          }
        }
        if (p.out_relu) {
          x.x = fmaxf(x.x, 0.f); x.y = fmaxf(x.y, 0.f); x.z = fmaxf(x.z, 0.f); x.w = fmaxf(x.w, 0.f);
        }
        if (mvalid) {
          if (full4 && vec_out_ok) {
            *reinterpret_cast<float4*>(orow + n) = x;
          } else {
            if (n < p.Nout) orow[n] = x.x;
            if (n + 1 < p.Nout) orow[n + 1] = x.y;
            if (n + 2 < p.Nout) orow[n + 2] = x.z;
            if (n + 3 < p.Nout) orow[n + 3] = x.w;
          }
        }
        // keep masked values for the statistics
        const bool rowok = mvalid;
        v[q * 4 + 0] = (rowok && n < p.Nout) ? x.x : 0.f;
        v[q * 4 + 1] = (rowok && n + 1 < p.Nout) ? x.y : 0.f;
        v[q * 4 + 2] = (rowok && n + 2 < p.Nout) ? x.z : 0.f;
        v[q * 4 + 3] = (rowok && n + 3 < p.Nout) ? x.w : 0.f;
      }
    }
    if (tr != nullptr && cc == 0) tr[1] = clock64();
    if (do_stats) {
      float sq[16];
      if (p.bnb_x != nullptr) {
        // fused BN-backward reduction: g = d(act) * [bn(x) > 0];  accumulate sum g and sum g*x
        const float* xrow = p.bnb_x + m * (size_t)p.bnb_ld + nb;
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) {
          const int n = nb + q4 * 4;
          float4 xv = make_float4(0.f, 0.f, 0.f, 0.f);
          float4 sc = xv, sh = xv;
          if (mvalid && n + 3 < p.Nout) {
            xv = __ldg(reinterpret_cast<const float4*>(xrow + q4 * 4));
            sc = __ldg(reinterpret_cast<const float4*>(p.bnb_scale + n));
            sh = __ldg(reinterpret_cast<const float4*>(p.bnb_shift + n));
          }
          const float xs[4] = {xv.x, xv.y, xv.z, xv.w};
          const float ss[4] = {sc.x, sc.y, sc.z, sc.w};
          const float hs[4] = {sh.x, sh.y, sh.z, sh.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float g = fmaf(xs[e], ss[e], hs[e]) > 0.f ? v[q4 * 4 + e] : 0.f;
            v[q4 * 4 + e] = g;
            sq[q4 * 4 + e] = g * xs[e];
          }
        }
      } else {
#pragma unroll
        for (int e = 0; e < 16; ++e) sq[e] = v[e] * v[e];
      }
      int col;
      const float s1 = warp_colsum16(v, lane, col);
      const float s2 = warp_colsum16(sq, lane, col);
      if ((lane & 1) == 0) {
        atomicAdd(&csum[cc + col], s1);
        atomicAdd(&csq[cc + col], s2);
      }
    }
    if (tr != nullptr && cc == 0) tr[2] = clock64();
  }
}

}  // namespace dtf

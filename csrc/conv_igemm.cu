// K1: implicit-GEMM convolution (fprop and dgrad) and FC matmul on the 5th-gen tensor cores.
//
//   out[m, n] = epilogue( sum_k A[m, k] * Wt[n, k] )         (see include/dtf_b200.h)
//
// Persistent, warp-specialised CTA (416 threads, one CTA per SM):
//   warps 0-3  epilogue : tcgen05.ld accumulator rows out of TMEM -> +bias / +residual / ReLU /
//                         batch-norm statistics of the stored value -> global NHWC
//   warps 4-11 producers: software im2col gather of the A tile (128 pixels x 32 k) straight from the
//                         NHWC activation, with the producer layer's batch-norm apply + ReLU fused into
//                         the load (so normalised activations never exist in HBM), written into the
//                         128B-swizzled K-major layout the tensor core expects; thread 128 also issues
//                         the TMA load of the weight tile (B operand)
//   warp  12   MMA      : one elected lane issues tcgen05.mma.kind::tf32 (M=128, N=BN, K=8) x4 per
//                         k-block; accumulators live in TMEM, double-buffered so the epilogue of tile i
//                         overlaps the main loop of tile i+1
// Pipelines: smem full/empty ring (producers+TMA <-> MMA) and TMEM full/empty (MMA <-> epilogue).
//
// Reference ops replaced: tf.nn.conv2d (cnn/cifar10.py:221,238), slim ops.conv2d / ops.fc
// (inception/slim/ops.py:232,305), tf.layers.conv2d / dense (resnet/resnet_model.py:87,277), and their
// gradient ops (input gradient = the same kernel with the "dgrad" index map).
#include "../include/dtf_b200.h"
#include "host_common.h"
#include "umma.cuh"

namespace dtf {

constexpr int BM = 128;              // pixels per tile (UMMA M)
constexpr int BK = 32;               // k per stage: 32 tf32 = one 128-byte swizzle row
constexpr int A_STAGE_BYTES = BM * BK * 4;
constexpr int NUM_PRODUCERS = 256;   // warps 4..11
constexpr int NUM_THREADS = 128 + NUM_PRODUCERS + 32;
constexpr int PRODUCER_T0 = 128;
constexpr int MMA_WARP = (128 + NUM_PRODUCERS) / 32;
constexpr int PRODUCER_GROUPS = 4;
constexpr int GROUP_THREADS = NUM_PRODUCERS / PRODUCER_GROUPS;   // 64
constexpr int ROWS_PER_THREAD = BM * 8 / GROUP_THREADS;   // 16

struct GemmCfg {
  int BN;          // UMMA N (multiple of 16, <= 256)
  int stages;
  int m_tiles, n_tiles;
  int num_kb;
  int tmem_cols;   // power of two >= 2*BN
  int M;
  int Ktot;
  int active_groups;   // producer groups in use: min(PRODUCER_GROUPS, stages)
  int all_live;    // every k-block touches at least one in-range tap
  int vec_out;     // 16-byte stores allowed
  int vec_res;
  FastDiv d_ohow, d_ow, d_c, d_s, d_divh, d_divw;
  long long* trace;  // optional timeline buffer (debug)
};

// Optional timeline trace (debug / profiling aid): when armed through dtf_debug_set_trace, CTA 0 stamps SM
// clocks at the pipeline hand-offs of its first tile into a caller buffer (scripts/trace_conv.py).
static long long* g_trace_host = nullptr;     // armed by dtf_debug_set_trace, passed to the kernel in GemmCfg
#define trace_stamp(slot)                                                              \
  do {                                                                                 \
    if (cfg.trace != nullptr && blockIdx.x == 0 && (slot) < 4096) cfg.trace[(slot)] = clock64(); \
  } while (0)

__device__ __forceinline__ bool tap_live_1d(int r, int a, int b, int c, int div, int O, int srcDim) {
  const int lo = b + r * c;
  const int hi = (O - 1) * a + lo;
  return hi >= 0 && lo <= srcDim * div - 1;
}

__device__ __forceinline__ bool kb_live(const dtf_conv_gemm_params& p, const GemmCfg& cfg, int kb) {
  if (cfg.all_live) return true;
  const int k0 = kb * BK;
  int k1 = k0 + BK - 1;
  if (k1 >= cfg.Ktot) k1 = cfg.Ktot - 1;
  const int t0 = (int)fdiv((uint32_t)k0, cfg.d_c), t1 = (int)fdiv((uint32_t)k1, cfg.d_c);
  for (int t = t0; t <= t1; ++t) {
    const int r = (int)fdiv((uint32_t)t, cfg.d_s), s = t - r * p.S;
    if (tap_live_1d(r, p.a_h, p.b_h, p.c_h, p.div_h, p.OH, p.srcH) &&
        tap_live_1d(s, p.a_w, p.b_w, p.c_w, p.div_w, p.OW, p.srcW))
      return true;
  }
  return false;
}

__global__ void __launch_bounds__(NUM_THREADS, 1)
conv_gemm_kernel(const __grid_constant__ CUtensorMap tmapB, const dtf_conv_gemm_params p, const GemmCfg cfg) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);

  const int BN = cfg.BN;
  const int stages = cfg.stages;
  const int b_stage_bytes = BN * BK * 4;
  uint8_t* sA = smem;
  uint8_t* sB = sA + stages * A_STAGE_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(sB + stages * b_stage_bytes);
  uint64_t* empty_bar = full_bar + stages;
  uint64_t* tfull_bar = empty_bar + stages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
  float* csum = reinterpret_cast<float*>(tmem_slot + 4);   // [256]
  float* csq = csum + 256;                                 // [256]
  int2* row_table = reinterpret_cast<int2*>(csq + 256);    // [PRODUCER_GROUPS][BM]

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  if (tid == 0) trace_stamp(3);

  if (warp == MMA_WARP) {
    if (lane == 0) {
      for (int i = 0; i < stages; ++i) {
        mbar_init(&full_bar[i], GROUP_THREADS);
        mbar_init(&empty_bar[i], 1);
      }
      for (int i = 0; i < 2; ++i) {
        mbar_init(&tfull_bar[i], 1);
        mbar_init(&tempty_bar[i], 128);
      }
      fence_mbar_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, (uint32_t)cfg.tmem_cols);
    tmem_relinquish();
  }
  if (tid < 128) {
    csum[tid] = 0.f; csum[tid + 128] = 0.f;
    csq[tid] = 0.f;  csq[tid + 128] = 0.f;
  }
  if (tid == PRODUCER_T0) tma_prefetch_desc(&tmapB);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (tid == 0) trace_stamp(4);

  const int num_tiles = cfg.m_tiles * cfg.n_tiles;

  if (warp >= 4 && warp < MMA_WARP) {
    // =========================== producers =================================================
    // Four groups of two warps; group g owns every 4th live k-block (sequence number q with q % 4 == g), so
    // four k-blocks are being gathered at any time.  Inside a group the order is strictly
    //   wait(slot free) -> TMA(B) + 16 x LDG.128 per thread -> [BN-apply + ReLU] -> STS -> proxy fence -> arrive.
    // The fence (MEMBAR + FENCE.VIEW.ASYNC) waits for *all* of the thread's outstanding memory operations, so a
    // thread must not have loads of a later k-block in flight when it fences; overlap comes from the other groups.
    // The gather is instruction-bound, so everything that depends only on the output pixel (row) is computed
    // once per tile into a per-group shared-memory table {element offset of the row's origin pixel, ih0, iw0};
    // per k-block a row then costs one LDS.128, two adds, two unsigned compares, one 64-bit add and the LDG.
    const int pt = tid - PRODUCER_T0;
    const int grp = pt >> 6;                 // 0..3
    const int gt = pt & 63;
    const int j = gt & 7;
    const int rgrp = gt >> 3;                // rows rgrp*16 .. rgrp*16+15
    const bool has_aff = p.in_scale != nullptr;
    const bool div1 = (p.div_h == 1) && (p.div_w == 1);
    int2* rtab = row_table + grp * BM;
    int q = 0, q_grp = 0, stage = 0; uint32_t phase = 0;

    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int mt = tile / cfg.n_tiles, nt = tile - mt * cfg.n_tiles;
      const int m0 = mt * BM, n0 = nt * BN;
      // ---- per-tile row table (two rows per thread) ----
      named_bar_sync(2 + grp, GROUP_THREADS);          // everyone is done with the previous tile's table
#pragma unroll
      for (int rr = 0; rr < 2; ++rr) {
        const int row = gt * 2 + rr;
        const int m = m0 + row;
        // entry = { element offset of the row's origin pixel (div1) | pixel index of its image (strided dgrad),
        //           (ih0 << 16) | (iw0 & 0xffff) }   — ih0 = -32768 marks a row past the end of the tensor
        int2 e = make_int2(0, (int)0x80000000);
        if (m < cfg.M) {
          const int img = (int)fdiv((uint32_t)m, cfg.d_ohow);
          const int rem = m - img * p.OH * p.OW;
          const int oh = (int)fdiv((uint32_t)rem, cfg.d_ow);
          const int ow = rem - oh * p.OW;
          const int ih0 = oh * p.a_h + p.b_h, iw0 = ow * p.a_w + p.b_w;
          e.y = (ih0 << 16) | (iw0 & 0xffff);
          if (div1) e.x = ((img * p.srcH + ih0) * p.srcW + iw0) * p.src_ld;   // may be "negative": only used when valid
          else e.x = img * p.srcH * p.srcW;                                 // pixel index of the image origin
        }
        rtab[row] = e;
      }
      named_bar_sync(2 + grp, GROUP_THREADS);
      const bool tr = (pt == 0) && (tile == (int)blockIdx.x);
      if (tr) trace_stamp(0);
      for (int kb = 0; kb < cfg.num_kb; ++kb) {
        if (!kb_live(p, cfg, kb)) continue;
        // A group may only run one ring-phase ahead of the consumer (mbarrier waits see a single parity bit),
        // which holds as long as there are at least as many stages as active groups.
        if (q_grp == grp) {
          if (tr) trace_stamp(16 + 8 * (q >> 2) + 0);
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          if (tr) trace_stamp(16 + 8 * (q >> 2) + 1);
          uint8_t* a_stage = sA + stage * A_STAGE_BYTES;
          if (gt == 0) {
            mbar_expect_tx(&full_bar[stage], (uint32_t)b_stage_bytes);
            tma_load_2d(sB + stage * b_stage_bytes, &tmapB, &full_bar[stage], kb * BK, n0);
          }
          const int k = kb * BK + j * 4;
          const bool kvalid = k < cfg.Ktot;
          int c = 0, dh = -(1 << 20), dw = 0;
          float4 sc = make_float4(1.f, 1.f, 1.f, 1.f), sh = make_float4(0.f, 0.f, 0.f, 0.f);
          if (kvalid) {
            const int tap = (int)fdiv((uint32_t)k, cfg.d_c);
            c = k - tap * p.srcC;
            const int r = (int)fdiv((uint32_t)tap, cfg.d_s);
            const int s_ = tap - r * p.S;
            dh = r * p.c_h;
            dw = s_ * p.c_w;
            if (has_aff) {
              sc = __ldg(reinterpret_cast<const float4*>(p.in_scale + c));
              sh = __ldg(reinterpret_cast<const float4*>(p.in_shift + c));
            }
          }
          float4 v[ROWS_PER_THREAD];
          uint32_t okmask = 0;
          const int2* rt = rtab + rgrp * ROWS_PER_THREAD;
          if (div1 && !has_aff) {
            // ---- asynchronous path (no fused activation): LDGSTS straight into the swizzled tile ----
            const int tap_off = (dh * p.srcW + dw) * p.src_ld + c;
            const uint32_t a_base = smem_u32(a_stage) + (uint32_t)(rgrp * ROWS_PER_THREAD) * 128u;
            // phase 1 (plain ALU + LDS, freely schedulable): element offsets and validity bits
            int off[ROWS_PER_THREAD];
#pragma unroll
            for (int i = 0; i < ROWS_PER_THREAD; ++i) {
              const int2 e = rt[i];
              const int ih = (e.y >> 16) + dh;
              const int iw = (int)(short)(e.y & 0xffff) + dw;
              const bool valid = (unsigned)ih < (unsigned)p.srcH && (unsigned)iw < (unsigned)p.srcW;
              if (valid) okmask |= 1u << i;
              off[i] = valid ? e.x + tap_off : 0;
            }
            // phase 2: sixteen LDGSTS back to back
#pragma unroll
            for (int i = 0; i < ROWS_PER_THREAD; ++i)
              cp_async_16(a_base + (uint32_t)i * 128u + (uint32_t)((j ^ (i & 7)) << 4), p.src + off[i],
                          ((okmask >> i) & 1u) ? 16u : 0u);
            if (tr) trace_stamp(16 + 8 * (q >> 2) + 2);
            cp_async_mbar_arrive_noinc(&full_bar[stage]);
            if (tr) trace_stamp(16 + 8 * (q >> 2) + 3);
          } else {
          if (div1) {
            // phase 1: table -> 32-bit element offsets + validity bits (cheap ALU, no long-latency dependence)
            const int tap_off = (dh * p.srcW + dw) * p.src_ld + c;
            int off[ROWS_PER_THREAD];
#pragma unroll
            for (int i = 0; i < ROWS_PER_THREAD; ++i) {
              const int2 e = rt[i];
              const int ih = (e.y >> 16) + dh;
              const int iw = (int)(short)(e.y & 0xffff) + dw;
              if ((unsigned)ih < (unsigned)p.srcH && (unsigned)iw < (unsigned)p.srcW) okmask |= 1u << i;
              off[i] = e.x + tap_off;
            }
            // phase 2: all loads back to back
#pragma unroll
            for (int i = 0; i < ROWS_PER_THREAD; ++i) {
              float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
              if ((okmask >> i) & 1u) x = __ldg(reinterpret_cast<const float4*>(p.src + off[i]));
              v[i] = x;
            }
          } else {
#pragma unroll
            for (int i = 0; i < ROWS_PER_THREAD; ++i) {
              const int2 e = rt[i];
              const int ihn = (e.y >> 16) + dh, iwn = (int)(short)(e.y & 0xffff) + dw;
              const int ih = (int)fdiv((uint32_t)(ihn < 0 ? 0 : ihn), cfg.d_divh);
              const int iw = (int)fdiv((uint32_t)(iwn < 0 ? 0 : iwn), cfg.d_divw);
              const bool valid = ihn >= 0 && iwn >= 0 && ih * p.div_h == ihn && iw * p.div_w == iwn &&
                                 ih < p.srcH && iw < p.srcW;
              float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
              if (valid) {
                const size_t off = (size_t)(e.x + ih * p.srcW + iw) * (size_t)p.src_ld + c;
                x = __ldg(reinterpret_cast<const float4*>(p.src + off));
                okmask |= 1u << i;
              }
              v[i] = x;
            }
          }
          if (tr) trace_stamp(16 + 8 * (q >> 2) + 2);
#pragma unroll
          for (int i = 0; i < ROWS_PER_THREAD; ++i) {
            const int row = rgrp * ROWS_PER_THREAD + i;
            float4 x = v[i];
            if (has_aff && ((okmask >> i) & 1u)) {
              x.x = fmaf(x.x, sc.x, sh.x); x.y = fmaf(x.y, sc.y, sh.y);
              x.z = fmaf(x.z, sc.z, sh.z); x.w = fmaf(x.w, sc.w, sh.w);
              if (p.in_relu) {
                x.x = fmaxf(x.x, 0.f); x.y = fmaxf(x.y, 0.f); x.z = fmaxf(x.z, 0.f); x.w = fmaxf(x.w, 0.f);
              }
            }
            *reinterpret_cast<float4*>(a_stage + row * 128 + ((j ^ (row & 7)) << 4)) = x;
          }
          fence_proxy_async_smem();
          if (tr) trace_stamp(16 + 8 * (q >> 2) + 3);
          mbar_arrive(&full_bar[stage]);
          }
        }
        ++q;
        if (++q_grp == cfg.active_groups) q_grp = 0;
        if (++stage == stages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == MMA_WARP) {
    // =========================== MMA issuer ================================================
    const uint32_t idesc = make_idesc_tf32(BM, (uint32_t)BN);
    const uint64_t adesc0 = make_desc_k_sw128(smem_u32(sA));
    const uint64_t bdesc0 = make_desc_k_sw128(smem_u32(sB));
    int stage = 0; uint32_t phase = 0;
    int local = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
      const int acc = local & 1;
      const uint32_t acc_phase = (uint32_t)(local >> 1) & 1u;
      mbar_wait(&tempty_bar[acc], acc_phase ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
      uint32_t accumulate = 0;
      int seq = 0;
      const bool tr = (lane == 0) && (local == 0);
      for (int kb = 0; kb < cfg.num_kb; ++kb) {
        if (!kb_live(p, cfg, kb)) continue;
        mbar_wait(&full_bar[stage], phase);
        if (tr && (seq & 3) == 0) trace_stamp(16 + 8 * (seq >> 2) + 4);
        if (tr && seq < 512) trace_stamp(2048 + 2 * seq);
        tc_fence_after();
        {
          // descriptors are warp-uniform (derived from `stage` only): compute them outside the elected branch so
          // they can live in uniform registers instead of being broadcast per MMA
          const uint64_t adesc = adesc0 + (uint64_t)((uint32_t)(stage * A_STAGE_BYTES) >> 4);
          const uint64_t bdesc = bdesc0 + (uint64_t)((uint32_t)(stage * b_stage_bytes) >> 4);
          if (elect_one()) {        // elect.sync: the compiler knows exactly one lane issues (no waterfall loops)
#pragma unroll
            for (int kk = 0; kk < BK / 8; ++kk) {
              umma_tf32_ss(d_tmem, adesc + (uint64_t)(kk * 2), bdesc + (uint64_t)(kk * 2), idesc, accumulate);
              accumulate = 1;
            }
            umma_commit(&empty_bar[stage]);
            if (tr && (seq & 3) == 0) trace_stamp(16 + 8 * (seq >> 2) + 5);
            if (tr && seq < 512) trace_stamp(2049 + 2 * seq);
          }
        }
        __syncwarp();
        accumulate = 1;
        ++seq;
        if (++stage == stages) { stage = 0; phase ^= 1u; }
      }
      if (elect_one()) umma_commit(&tfull_bar[acc]);
      __syncwarp();
    }
  } else {
    // =========================== epilogue (warps 0-3) ======================================
    int local = 0;
    const bool do_stats = p.stats != nullptr;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
      const int mt = tile / cfg.n_tiles, nt = tile - mt * cfg.n_tiles;
      const int m0 = mt * BM, n0 = nt * BN;
      const int acc = local & 1;
      const uint32_t acc_phase = (uint32_t)(local >> 1) & 1u;
      mbar_wait(&tfull_bar[acc], acc_phase);
      if (tid == 0 && local == 0) trace_stamp(1);
      tc_fence_after();
      const int m = m0 + tid;
      const bool mvalid = m < cfg.M;
      float* orow = p.out + (size_t)m * p.out_ld;
      const float* rrow = p.residual ? p.residual + (size_t)m * p.res_ld : nullptr;
      const uint32_t t_row = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(acc * BN);
      for (int cc = 0; cc < BN; cc += 16) {
        float v[16];
        tmem_ld16(t_row + (uint32_t)cc, v);
        const int nb = n0 + cc;
        if (nb >= p.Nout) break;    // uniform across the CTA
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
            if (full4 && cfg.vec_res) {
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
            if (full4 && cfg.vec_out) {
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
        if (do_stats) {
          float sq[16];
          if (p.bnb_x != nullptr) {
            // fused BN-backward reduction: g = d(act) * [bn(x) > 0];  accumulate sum g and sum g*x
            const float* xrow = p.bnb_x + (size_t)m * p.bnb_ld + nb;
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
      }
      tc_fence_before();
      mbar_arrive(&tempty_bar[acc]);
      if (tid == 0 && local == 0) trace_stamp(2);
      if (do_stats && cfg.n_tiles > 1) {
        // several n-tiles share the smem accumulators: flush after every tile
        named_bar_sync(1, 128);
        for (int i = tid; i < BN; i += 128) {
          const int n = n0 + i;
          if (n < p.Nout) {
            atomicAdd(p.stats + n, csum[i]);
            atomicAdd(p.stats + p.Nout + n, csq[i]);
          }
          csum[i] = 0.f; csq[i] = 0.f;
        }
        named_bar_sync(1, 128);
      }
    }
    if (do_stats && cfg.n_tiles == 1) {
      named_bar_sync(1, 128);
      for (int i = tid; i < BN; i += 128) {
        if (i < p.Nout && (csum[i] != 0.f || csq[i] != 0.f)) {
          atomicAdd(p.stats + i, csum[i]);
          atomicAdd(p.stats + p.Nout + i, csq[i]);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == MMA_WARP) {
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)cfg.tmem_cols);
  }
}

// [Nout][RS][C] -> [C][RS][Nout]
__global__ void weight_transpose_kernel(const float* __restrict__ w, float* __restrict__ wt, int Nout, int RS, int C) {
  __shared__ float tile[32][33];
  const int rs = blockIdx.z;
  const int c0 = blockIdx.x * 32, n0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int n = n0 + i, c = c0 + threadIdx.x;
    tile[i][threadIdx.x] = (n < Nout && c < C) ? w[((size_t)n * RS + rs) * C + c] : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = c0 + i, n = n0 + threadIdx.x;
    if (c < C && n < Nout) wt[((size_t)c * RS + rs) * Nout + n] = tile[threadIdx.x][i];
  }
}

static int pick_bn(int nout) {
  if (nout <= 256) return round_up(nout, 16);
  // prefer the widest tile that divides evenly; otherwise 128 keeps the tail small
  if (nout % 256 == 0) return 256;
  if (nout % 192 == 0) return 192;
  if (nout % 128 == 0) return 128;
  return 256;
}

static int next_pow2_cols(int c) {
  int v = 32;
  while (v < c) v <<= 1;
  return v;
}

}  // namespace dtf

using namespace dtf;

extern "C" int dtf_conv_gemm(const dtf_conv_gemm_params* pp, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  const dtf_conv_gemm_params& p = *pp;
  DTF_REQUIRE(p.srcC % 4 == 0, "source channel count must be a multiple of 4 (pad the stem input)");
  DTF_REQUIRE(p.src_ld % 4 == 0, "source pixel pitch must be a multiple of 4 floats");
  DTF_REQUIRE((reinterpret_cast<uintptr_t>(p.src) & 15) == 0, "source must be 16-byte aligned");
  DTF_REQUIRE(p.div_h >= 1 && p.div_w >= 1 && p.a_h >= 1 && p.a_w >= 1, "bad index map");
  DTF_REQUIRE(p.Nout >= 1 && p.R >= 1 && p.S >= 1, "bad shape");
  DTF_REQUIRE((long long)p.N * p.srcH * p.srcW * p.src_ld < (1LL << 31), "source tensor too large for 32-bit offsets");
  if (p.in_scale) DTF_REQUIRE(p.in_shift != nullptr, "in_scale without in_shift");
  if (p.bnb_x) {
    DTF_REQUIRE(p.stats && p.bnb_scale && p.bnb_shift, "fused BN-backward reduction needs stats/scale/shift");
    DTF_REQUIRE(p.Nout % 4 == 0 && p.bnb_ld % 4 == 0, "fused BN-backward reduction needs C % 4 == 0");
  }

  GemmCfg cfg;
  cfg.Ktot = p.R * p.S * p.srcC;
  cfg.M = p.N * p.OH * p.OW;
  if (cfg.M == 0) return 0;
  cfg.BN = pick_bn(p.Nout);
  cfg.n_tiles = ceil_div(p.Nout, cfg.BN);
  cfg.m_tiles = ceil_div(cfg.M, BM);
  cfg.num_kb = ceil_div(cfg.Ktot, BK);
  cfg.tmem_cols = next_pow2_cols(2 * cfg.BN);
  const int stage_bytes = A_STAGE_BYTES + cfg.BN * BK * 4;
  const int fixed_smem = 1024 + 16 * 2 * 8 + 4 * 8 + 16 + 2 * 256 * 4 + PRODUCER_GROUPS * BM * (int)sizeof(int2);
  int stages = (227 * 1024 - fixed_smem) / stage_bytes;
  if (stages > 8) stages = 8;
  if (stages > cfg.num_kb + 1) stages = cfg.num_kb + 1;
  if (stages < 2) stages = 2;
  cfg.stages = stages;
  cfg.active_groups = stages < PRODUCER_GROUPS ? stages : PRODUCER_GROUPS;
  // liveness: every tap reaches at least one in-range source pixel?
  int all_live = 1;
  for (int r = 0; r < p.R && all_live; ++r) {
    const int lo = p.b_h + r * p.c_h, hi = (p.OH - 1) * p.a_h + lo;
    if (!(hi >= 0 && lo <= p.srcH * p.div_h - 1)) all_live = 0;
  }
  for (int s = 0; s < p.S && all_live; ++s) {
    const int lo = p.b_w + s * p.c_w, hi = (p.OW - 1) * p.a_w + lo;
    if (!(hi >= 0 && lo <= p.srcW * p.div_w - 1)) all_live = 0;
  }
  cfg.all_live = all_live;
  cfg.d_ohow = make_fastdiv((uint32_t)(p.OH * p.OW));
  cfg.d_ow = make_fastdiv((uint32_t)p.OW);
  cfg.d_c = make_fastdiv((uint32_t)p.srcC);
  cfg.d_s = make_fastdiv((uint32_t)p.S);
  cfg.d_divh = make_fastdiv((uint32_t)p.div_h);
  cfg.d_divw = make_fastdiv((uint32_t)p.div_w);
  cfg.trace = g_trace_host;
  cfg.vec_out = (p.out_ld % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.out) & 15) == 0);
  cfg.vec_res = p.residual && (p.res_ld % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.residual) & 15) == 0);

  const CUtensorMap* tm = get_tmap_2d_f32_sw128(p.wt, p.Nout, cfg.Ktot, cfg.BN);
  if (!tm) return -2;

  const size_t smem = 1024 + (size_t)stages * stage_bytes + (2 * stages + 4) * sizeof(uint64_t) + 16 + 2 * 256 * 4 +
                      PRODUCER_GROUPS * BM * sizeof(int2);
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    DTF_CHECK(cudaFuncSetAttribute(conv_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set[dev & 63] = true;
  }
  const int tiles = cfg.m_tiles * cfg.n_tiles;
  const int grid = tiles < num_sms() ? tiles : num_sms();
  conv_gemm_kernel<<<grid, NUM_THREADS, smem, stream>>>(*tm, p, cfg);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_debug_set_trace(long long* buf) {
  g_trace_host = buf;
  return 0;
}

extern "C" int dtf_weight_transpose(const float* w, float* wt, int Nout, int RS, int C, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  dim3 grid(ceil_div(C, 32), ceil_div(Nout, 32), RS), block(32, 8);
  weight_transpose_kernel<<<grid, block, 0, stream>>>(w, wt, Nout, RS, C);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

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
#include "ps_gate.cuh"
#include "conv_epilogue.cuh"

namespace dtf {

constexpr int BM = 128;              // pixels per tile (UMMA M)
constexpr int BK = 32;               // k per stage: 32 tf32 = one 128-byte swizzle row
constexpr int A_STAGE_BYTES = BM * BK * 4;
constexpr int NUM_PRODUCERS = 256;   // warps 4..11
constexpr int NUM_THREADS = 128 + NUM_PRODUCERS + 64;
constexpr int PRODUCER_T0 = 128;
constexpr int MMA_WARP = (128 + NUM_PRODUCERS) / 32;
constexpr int TMA_WARP = MMA_WARP + 1;
constexpr int A_HALF_BYTES = A_STAGE_BYTES / 2;   // one 16-channel (64-byte swizzle) im2col box
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
  // hardware im2col (TMA) feed of the A operand
  int tma_a;       // 0: software gather by the producer warps   1: cp.async.bulk.tensor im2col loads
  int cb;          // channels per im2col load: 32 (128-byte swizzle) or 16 (two 64-byte-swizzle halves per stage)
  int cblocks;     // srcC / cb
  int n_units;     // R * S * cblocks im2col loads per tile
  int lw, lh;      // base-pixel origin (bounding-box lower corner), W and H
  int flip_w, flip_h;   // tap offset counts down (dgrad: the index map steps by -1 per tap)
  FastDiv d_cblk;
  FastDiv d_ohow, d_ow, d_c, d_s, d_divh, d_divw;
  long long* trace;  // optional timeline buffer (debug)
};

// Optional timeline trace (debug / profiling aid): when armed through dtf_debug_set_trace, CTA 0 stamps SM
// clocks at the pipeline hand-offs of its first tile into a caller buffer (scripts/trace_conv.py).
static int g_conv_tma_enabled = 1;            // dtf_debug_set_conv_tma(0) forces the software gather (A/B tests)
long long* g_trace_host = nullptr;     // armed by dtf_debug_set_trace, passed to the kernel in GemmCfg
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
conv_gemm_kernel(const __grid_constant__ CUtensorMap tmapB, const __grid_constant__ CUtensorMap tmapA,
                 const dtf_conv_gemm_params p, const GemmCfg cfg) {
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
  uint64_t* tma_bar = tempty_bar + 2;                      // [stages] im2col tile landed, waiting for the transform
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tma_bar + stages);
  float* csum = reinterpret_cast<float*>(tmem_slot + 4);   // [256]
  float* csq = csum + 256;                                 // [256]
  int2* row_table = reinterpret_cast<int2*>(csq + 256);    // [PRODUCER_GROUPS][BM]

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  if (tid == 0) trace_stamp(3);
  pdl_launch_dependents();   // persistent grid, all CTAs resident: the successor may take SMs as they free up

  if (warp == MMA_WARP) {
    if (lane == 0) {
      // arrivals that complete a stage: the gathering group's threads | the TMA issuer | one per transform warp
      const uint32_t full_count = !cfg.tma_a ? (uint32_t)GROUP_THREADS
                                             : (p.in_scale != nullptr ? (uint32_t)(NUM_PRODUCERS / 32) : 1u);
      for (int i = 0; i < stages; ++i) {
        mbar_init(&full_bar[i], full_count);
        mbar_init(&empty_bar[i], 1);
        mbar_init(&tma_bar[i], 1);
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
  if (tid == TMA_WARP * 32 && cfg.tma_a) tma_prefetch_desc(&tmapA);
  pdl_wait();     // everything above is independent of the predecessor kernel
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (tid == 0) trace_stamp(4);

  const int num_tiles = cfg.m_tiles * cfg.n_tiles;

  if (warp >= 4 && warp < MMA_WARP && cfg.tma_a) {
    // =========================== transform warps (TMA feed + fused BN-apply/ReLU) ===========
    // The im2col tile lands raw, with out-of-image pixels filled by the NaN marker.  Each warp rewrites its
    // 16 rows in place as relu(scale * x + shift), turning the marker into the zero padding the convolution
    // wants, then publishes the stage to the tensor core.  Without a fused input transform the TMA completes the
    // stage by itself and these warps have nothing to do.
    if (p.in_scale != nullptr) {
      const int pw = warp - 4;
      const bool relu = p.in_relu != 0;
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        for (int kb = 0; kb < cfg.num_kb; ++kb) {
          if (!kb_live(p, cfg, kb)) continue;
          uint8_t* a_stage = sA + stage * A_STAGE_BYTES;
          if (cfg.cb == 32) {
            const int j = lane & 7;
            const int tap = (int)fdiv((uint32_t)kb, cfg.d_cblk);
            const int c = (kb - tap * cfg.cblocks) * 32 + j * 4;
            const float4 sc = __ldg(reinterpret_cast<const float4*>(p.in_scale + c));
            const float4 sh = __ldg(reinterpret_cast<const float4*>(p.in_shift + c));
            mbar_wait(&tma_bar[stage], phase);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int row = pw * 16 + i * 4 + (lane >> 3);
              float4* ptr = reinterpret_cast<float4*>(a_stage + row * 128 + ((j ^ (row & 7)) << 4));
              float4 x = *ptr;
              float4 y;
              y.x = fmaf(x.x, sc.x, sh.x); y.y = fmaf(x.y, sc.y, sh.y);
              y.z = fmaf(x.z, sc.z, sh.z); y.w = fmaf(x.w, sc.w, sh.w);
              if (relu) { y.x = fmaxf(y.x, 0.f); y.y = fmaxf(y.y, 0.f); y.z = fmaxf(y.z, 0.f); y.w = fmaxf(y.w, 0.f); }
              y.x = (x.x == x.x) ? y.x : 0.f; y.y = (x.y == x.y) ? y.y : 0.f;
              y.z = (x.z == x.z) ? y.z : 0.f; y.w = (x.w == x.w) ? y.w : 0.f;
              *ptr = y;
            }
          } else {
            const int jj = lane & 3, rs = lane >> 2;
            const int u0 = kb * 2;
            const int nu = cfg.n_units - u0 < 2 ? cfg.n_units - u0 : 2;
            float4 sc[2], sh[2];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              const int u = h < nu ? u0 + h : u0;
              const int tap = (int)fdiv((uint32_t)u, cfg.d_cblk);
              const int c = (u - tap * cfg.cblocks) * 16 + jj * 4;
              sc[h] = __ldg(reinterpret_cast<const float4*>(p.in_scale + c));
              sh[h] = __ldg(reinterpret_cast<const float4*>(p.in_shift + c));
            }
            mbar_wait(&tma_bar[stage], phase);
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              if (h < nu) {
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                  const int row = pw * 16 + i * 8 + rs;
                  float4* ptr = reinterpret_cast<float4*>(a_stage + h * A_HALF_BYTES + row * 64 +
                                                          ((jj ^ ((row >> 1) & 3)) << 4));
                  float4 x = *ptr;
                  float4 y;
                  y.x = fmaf(x.x, sc[h].x, sh[h].x); y.y = fmaf(x.y, sc[h].y, sh[h].y);
                  y.z = fmaf(x.z, sc[h].z, sh[h].z); y.w = fmaf(x.w, sc[h].w, sh[h].w);
                  if (relu) { y.x = fmaxf(y.x, 0.f); y.y = fmaxf(y.y, 0.f); y.z = fmaxf(y.z, 0.f); y.w = fmaxf(y.w, 0.f); }
                  y.x = (x.x == x.x) ? y.x : 0.f; y.y = (x.y == x.y) ? y.y : 0.f;
                  y.z = (x.z == x.z) ? y.z : 0.f; y.w = (x.w == x.w) ? y.w : 0.f;
                  *ptr = y;
                }
              }
            }
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(&full_bar[stage]);
          if (++stage == stages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == TMA_WARP) {
    // =========================== TMA issuer (hardware im2col feed) ===========================
    // Per k-block: wait for the slot, then one elected lane posts the byte count and issues the weight-tile load
    // plus one (32-channel) or two (16-channel) im2col loads.  The engine generates the addresses, applies the
    // convolution stride, zero/NaN-fills the padding and writes the swizzled K-major tile.
    if (cfg.tma_a) {
      const bool has_aff = p.in_scale != nullptr;
      int stage = 0; uint32_t phase = 0;
      int gseq = 0;
      weight_gate_warp(p);     // PS pull: the owners' updates of this layer's bucket must have landed (ps_gate.cuh)
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int mt = tile / cfg.n_tiles, nt = tile - mt * cfg.n_tiles;
        const int m0 = mt * BM, n0 = nt * BN;
        const int img = (int)fdiv((uint32_t)m0, cfg.d_ohow);
        const int rem = m0 - img * p.OH * p.OW;
        const int oh = (int)fdiv((uint32_t)rem, cfg.d_ow);
        const int ow = rem - oh * p.OW;
        const int w0 = ow * p.a_w + cfg.lw, h0 = oh * p.a_h + cfg.lh;
        for (int kb = 0; kb < cfg.num_kb; ++kb) {
          if (!kb_live(p, cfg, kb)) continue;
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          if (elect_one()) {
            if (gseq < 512) trace_stamp(1024 + gseq);
            uint64_t* bar = has_aff ? &tma_bar[stage] : &full_bar[stage];
            const uint32_t a_dst = smem_u32(sA + stage * A_STAGE_BYTES);
            if (cfg.cb == 32) {
              mbar_arrive_expect_tx(bar, (uint32_t)(b_stage_bytes + A_STAGE_BYTES));
              const int tap = (int)fdiv((uint32_t)kb, cfg.d_cblk);
              const int c0 = (kb - tap * cfg.cblocks) * 32;
              const int r = (int)fdiv((uint32_t)tap, cfg.d_s), s_ = tap - r * p.S;
              tma_load_im2col_4d(a_dst, &tmapA, bar, c0, w0, h0, img,
                                 (uint16_t)(cfg.flip_w ? p.S - 1 - s_ : s_), (uint16_t)(cfg.flip_h ? p.R - 1 - r : r));
            } else {
              const int u0 = kb * 2;
              const int nu = cfg.n_units - u0 < 2 ? cfg.n_units - u0 : 2;
              mbar_arrive_expect_tx(bar, (uint32_t)(b_stage_bytes + nu * A_HALF_BYTES));
              for (int h = 0; h < nu; ++h) {
                const int u = u0 + h;
                const int tap = (int)fdiv((uint32_t)u, cfg.d_cblk);
                const int c0 = (u - tap * cfg.cblocks) * 16;
                const int r = (int)fdiv((uint32_t)tap, cfg.d_s), s_ = tap - r * p.S;
                tma_load_im2col_4d(a_dst + (uint32_t)(h * A_HALF_BYTES), &tmapA, bar, c0, w0, h0, img,
                                   (uint16_t)(cfg.flip_w ? p.S - 1 - s_ : s_),
                                   (uint16_t)(cfg.flip_h ? p.R - 1 - r : r));
              }
            }
            tma_load_2d(sB + stage * b_stage_bytes, &tmapB, bar, kb * BK, n0);
          }
          __syncwarp();
          ++gseq;
          if (++stage == stages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp >= 4 && warp < MMA_WARP) {
    // =========================== producers (software gather) =================================
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
    if (gt == 0) weight_gate_thread(p);   // this thread issues the group's weight-tile loads (PS pull gate)

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
    const bool sw64 = cfg.tma_a && cfg.cb == 16;     // A stage = two 64-byte-swizzled halves (one filter tap each)
    const uint64_t adesc0 = sw64 ? make_desc_k_sw64(smem_u32(sA)) : make_desc_k_sw128(smem_u32(sA));
    const uint64_t bdesc0 = make_desc_k_sw128(smem_u32(sB));
    int stage = 0; uint32_t phase = 0;
    int local = 0;
    int gseq = 0;
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
        if (lane == 0 && gseq < 512) trace_stamp(2048 + 2 * gseq);
        tc_fence_after();
        {
          // descriptors are warp-uniform (derived from `stage` only): compute them outside the elected branch so
          // they can live in uniform registers instead of being broadcast per MMA
          const uint64_t adesc = adesc0 + (uint64_t)((uint32_t)(stage * A_STAGE_BYTES) >> 4);
          const uint64_t bdesc = bdesc0 + (uint64_t)((uint32_t)(stage * b_stage_bytes) >> 4);
          // a 16-channel feed whose unit count is odd leaves the second half of the last k-block unloaded
          const int nmma = (sw64 && cfg.n_units - kb * 2 < 2) ? 2 : BK / 8;
          if (elect_one()) {        // elect.sync: the compiler knows exactly one lane issues (no waterfall loops)
#pragma unroll
            for (int kk = 0; kk < BK / 8; ++kk) {
              if (kk < nmma) {
                const uint32_t aoff = sw64 ? (uint32_t)((kk >> 1) * (A_HALF_BYTES >> 4) + (kk & 1) * 2) : (uint32_t)(kk * 2);
                umma_tf32_ss(d_tmem, adesc + (uint64_t)aoff, bdesc + (uint64_t)(kk * 2), idesc, accumulate);
                accumulate = 1;
              }
            }
            umma_commit(&empty_bar[stage]);
            if (tr && (seq & 3) == 0) trace_stamp(16 + 8 * (seq >> 2) + 5);
            if (gseq < 512) trace_stamp(2049 + 2 * gseq);
          }
        }
        __syncwarp();
        accumulate = 1;
        ++seq;
        ++gseq;
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
      if (tid == 0 && local < 128) trace_stamp(512 + 2 * local);
      tc_fence_after();
      const int m = m0 + tid;
      const bool mvalid = m < cfg.M;
      const uint32_t t_row = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(acc * BN);
      conv_epilogue_row(p, BN, n0, mvalid, (size_t)m, t_row, cfg.vec_out, cfg.vec_res, do_stats, csum, csq, lane);
      tc_fence_before();
      mbar_arrive(&tempty_bar[acc]);
      if (tid == 0 && local == 0) trace_stamp(2);
      if (tid == 0 && local < 128) trace_stamp(513 + 2 * local);
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
  if (tid == 0) trace_stamp(5);
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

// All layers' weight transposes of a training step in one launch: entry e owns blocks [blk0, blk0 + its tile count).
struct WtEntry {
  const float* w;
  float* wt;
  int Nout, RS, C, blk0;
};

__global__ void weight_transpose_batched_kernel(const WtEntry* __restrict__ tab, int n_entries) {
  __shared__ float tile[32][33];
  int e = 0;
  while (e + 1 < n_entries && (int)blockIdx.x >= tab[e + 1].blk0) ++e;
  const WtEntry en = tab[e];
  const int cx = (en.C + 31) / 32, ny = (en.Nout + 31) / 32;
  int b = (int)blockIdx.x - en.blk0;
  const int rs = b / (cx * ny);
  b -= rs * cx * ny;
  const int by = b / cx, bx = b - by * cx;
  const int c0 = bx * 32, n0 = by * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int n = n0 + i, c = c0 + threadIdx.x;
    tile[i][threadIdx.x] = (n < en.Nout && c < en.C) ? en.w[((size_t)n * en.RS + rs) * en.C + c] : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = c0 + i, n = n0 + threadIdx.x;
    if (c < en.C && n < en.Nout) en.wt[((size_t)c * en.RS + rs) * en.Nout + n] = tile[threadIdx.x][i];
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

namespace dtf {
int try_conv_win(const dtf_conv_gemm_params& p, cudaStream_t stream, bool* launched, bool dry_run);
int try_conv_wint(const dtf_conv_gemm_params& p, cudaStream_t stream, bool* launched, bool dry_run);
}
using namespace dtf;

extern "C" int dtf_conv_gemm_fuses_finalize(const dtf_conv_gemm_params* pp);

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
  if (p.fin_stats) DTF_REQUIRE(dtf_conv_gemm_fuses_finalize(pp), "deferred batch-norm finalize needs the sliding-window kernel");
  if (p.bnb_x) {
    DTF_REQUIRE(p.stats && p.bnb_scale && p.bnb_shift, "fused BN-backward reduction needs stats/scale/shift");
    DTF_REQUIRE(p.Nout % 4 == 0 && p.bnb_ld % 4 == 0, "fused BN-backward reduction needs C % 4 == 0");
  }

  GemmCfg cfg;
  cfg.Ktot = p.R * p.S * p.srcC;
  cfg.M = p.N * p.OH * p.OW;
  if (cfg.M == 0) return 0;
  {
    // small-channel stride-1 stencils: the sliding-window kernel reads each activation once (conv_win.cu)
    bool launched = false;
    int rc = try_conv_wint(p, stream, &launched, false);     // <= 64 output channels: channels on M, pixels on N
    if (rc != 0 || launched) return rc;
    rc = try_conv_win(p, stream, &launched, false);
    if (rc != 0 || launched) return rc;
  }
  cfg.BN = pick_bn(p.Nout);
  cfg.n_tiles = ceil_div(p.Nout, cfg.BN);
  cfg.m_tiles = ceil_div(cfg.M, BM);
  cfg.num_kb = ceil_div(cfg.Ktot, BK);
  cfg.tmem_cols = next_pow2_cols(2 * cfg.BN);
  const int stage_bytes = A_STAGE_BYTES + cfg.BN * BK * 4;
  const int fixed_smem = 1024 + 16 * 3 * 8 + 4 * 8 + 16 + 2 * 256 * 4 + PRODUCER_GROUPS * BM * (int)sizeof(int2);
  int stages = (SMEM_BUDGET - fixed_smem) / stage_bytes;
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

  // ---- A operand feed: hardware im2col when the index map is a unit-step stencil over >= 16-channel pixels ----
  cfg.tma_a = 0; cfg.cb = 32; cfg.cblocks = 1; cfg.n_units = 0; cfg.lw = cfg.lh = 0; cfg.flip_w = cfg.flip_h = 0;
  cfg.d_cblk = make_fastdiv(1u);
  const CUtensorMap* tma = nullptr;
  const bool unit_h = (p.R == 1) || p.c_h == 1 || p.c_h == -1;
  const bool unit_w = (p.S == 1) || p.c_w == 1 || p.c_w == -1;
  if (g_conv_tma_enabled && p.div_h == 1 && p.div_w == 1 && unit_h && unit_w && p.srcC % 16 == 0 &&
      p.R <= 256 && p.S <= 256) {
    Im2colSpec sp;
    sp.ptr = p.src; sp.N = p.N; sp.H = p.srcH; sp.W = p.srcW; sp.C = p.srcC; sp.ld = p.src_ld;
    const int flip_h = (p.R > 1 && p.c_h < 0), flip_w = (p.S > 1 && p.c_w < 0);
    sp.lower_h = flip_h ? p.b_h - (p.R - 1) : p.b_h;
    sp.lower_w = flip_w ? p.b_w - (p.S - 1) : p.b_w;
    sp.upper_h = sp.lower_h + (p.OH - 1) * p.a_h + 1 - p.srcH;
    sp.upper_w = sp.lower_w + (p.OW - 1) * p.a_w + 1 - p.srcW;
    sp.th = p.a_h; sp.tw = p.a_w;
    sp.channels = (p.srcC % 32 == 0) ? 32 : 16;
    sp.nan_fill = p.in_scale != nullptr;
    tma = get_tmap_im2col_f32(sp);
    if (tma) {
      cfg.tma_a = 1;
      cfg.cb = sp.channels;
      cfg.cblocks = p.srcC / cfg.cb;
      cfg.n_units = p.R * p.S * cfg.cblocks;
      cfg.lw = sp.lower_w; cfg.lh = sp.lower_h;
      cfg.flip_w = flip_w; cfg.flip_h = flip_h;
      cfg.d_cblk = make_fastdiv((uint32_t)cfg.cblocks);
    }
  }

  const size_t smem = 1024 + (size_t)stages * stage_bytes + (3 * stages + 4) * sizeof(uint64_t) + 16 + 2 * 256 * 4 +
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
  DTF_CHECK(launch_pdl(conv_gemm_kernel, dim3(grid), dim3(NUM_THREADS), smem, stream, *tm, tma ? *tma : *tm, p, cfg));
  return 0;
}

extern "C" int dtf_conv_gemm_fuses_finalize(const dtf_conv_gemm_params* pp) {
  const dtf_conv_gemm_params& p = *pp;
  if (p.in_scale == nullptr || p.srcC % 4 != 0 || p.src_ld % 4 != 0 || p.N * p.OH * p.OW == 0) return 0;
  bool would = false;
  return (try_conv_wint(p, nullptr, &would, true) == 0 && would) ? 1 : 0;
}

extern "C" int dtf_conv_gemm_writes_act(const dtf_conv_gemm_params* pp) {
  const dtf_conv_gemm_params& p = *pp;
  if (p.in_scale == nullptr || p.srcC % 4 != 0 || p.src_ld % 4 != 0 || p.N * p.OH * p.OW == 0) return 0;
  bool would = false;
  if (try_conv_wint(p, nullptr, &would, true) == 0 && would) return 1;
  if (try_conv_win(p, nullptr, &would, true) == 0 && would) return 1;
  return 0;
}

extern "C" int dtf_debug_set_conv_tma(int enabled) {
  g_conv_tma_enabled = enabled;
  return 0;
}

extern "C" int dtf_debug_set_trace(long long* buf) {
  g_trace_host = buf;
  return 0;
}

extern "C" int dtf_weight_transpose_batched(const void* table, int n_entries, int total_blocks, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  if (n_entries <= 0 || total_blocks <= 0) return 0;
  weight_transpose_batched_kernel<<<total_blocks, dim3(32, 8), 0, stream>>>(reinterpret_cast<const WtEntry*>(table),
                                                                          n_entries);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_weight_transpose(const float* w, float* wt, int Nout, int RS, int C, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  dim3 grid(ceil_div(C, 32), ceil_div(Nout, 32), RS), block(32, 8);
  weight_transpose_kernel<<<grid, block, 0, stream>>>(w, wt, Nout, RS, C);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

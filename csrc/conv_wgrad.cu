// K1 (weight gradient): dW[n][k] += sum_m dY[m, n] * A[m, k] on tcgen05 (kind::tf32).
//
// GEMM view:  D[k, n] = sum over pixels m of  At[k, m] * dYt[n, m]   (UMMA M = 128 filter taps x channels,
// UMMA N = C_out, reduction over the N*OH*OW output pixels).  Both operands are "pixel-minor" in the GEMM
// but channel-minor in HBM (NHWC), so the producer warps transpose 4x4 blocks in registers while staging:
// four LDG.128 (4 channels of 4 consecutive pixels) become four STS.128 (4 pixels of 4 consecutive
// channels) into the 128B-swizzled K-major layout.  That keeps every descriptor K-major (the one tf32
// layout that has no special cases) and needs no transposed copy of activations or gradients in HBM.
// The pixel range is split across CTAs (persistent, one per SM); partial sums are combined with
// red.global.add.f32 — the gradient buffer is zeroed once per step by the engine.
//
// Reference ops replaced: the Conv2DBackpropFilter / MatMul-grad nodes that
// opt.compute_gradients() adds (e.g. cifarnet/cifar10_cifarnet_bsp.py:85,
// inception/imagenet_inception_bsp.py:150).
#include "../include/dtf_b200.h"
#include "host_common.h"
#include "umma.cuh"

namespace dtf {

constexpr int WG_BM = 128;
constexpr int WG_BK = 32;   // pixels per stage
constexpr int WG_A_STAGE_BYTES = WG_BM * WG_BK * 4;
constexpr int WG_PRODUCERS = 256;   // warps 4..11
constexpr int WG_THREADS = 128 + WG_PRODUCERS + 32;
constexpr int WG_MMA_WARP = (128 + WG_PRODUCERS) / 32;

struct WgradCfg {
  int BN, stages, m_tiles, n_tiles, splits, kb_total, kb_per_split, tmem_cols, M, Ktot;
  int vec_dy;
  int active_groups;     // producer groups in use: min(4, stages)
  FastDiv d_ohow, d_ow, d_c, d_s;
};

__global__ void __launch_bounds__(WG_THREADS, 1)
conv_wgrad_kernel(const dtf_conv_wgrad_params p, const WgradCfg cfg) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);

  const int BN = cfg.BN;
  const int stages = cfg.stages;
  const int b_stage_bytes = BN * WG_BK * 4;
  uint8_t* sA = smem;
  uint8_t* sB = sA + stages * WG_A_STAGE_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(sB + stages * b_stage_bytes);
  uint64_t* empty_bar = full_bar + stages;
  uint64_t* tfull_bar = empty_bar + stages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  pdl_launch_dependents();   // persistent grid, all CTAs resident: the successor may take SMs as they free up

  if (warp == WG_MMA_WARP) {
    if (lane == 0) {
      for (int i = 0; i < stages; ++i) {
        mbar_init(&full_bar[i], WG_PRODUCERS / 4);
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
  pdl_wait();     // everything above is independent of the predecessor kernel
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int num_units = cfg.m_tiles * cfg.n_tiles * cfg.splits;

  if (warp >= 4 && warp < WG_MMA_WARP) {
    // =========================== producers =================================================
    // Four groups of two warps; group g owns every 4th k-block (32 output pixels), so four k-blocks are being
    // gathered at any time.  Within a group the order is strictly
    //   wait(slot free) -> LDG.128 (dY chunks, A chunks) -> 4x4 register transposes -> STS -> proxy fence -> arrive
    // (the proxy fence drains all of a thread's outstanding memory operations, so a thread never has loads of a
    // later k-block in flight when it fences; the overlap comes from the other groups).
    // Thread gt owns pixel chunk pc (4 consecutive output pixels) and 4-row chunks rc0 + 8*i of the A tile
    // (one filter tap, 4 consecutive input channels each; i < 4) and of the dY tile (rows = output channels).
    const int pt = tid - 128;
    const int grp = pt >> 6;      // 0..3
    const int gt = pt & 63;
    const int pc = gt & 7;
    const int rc0 = gt >> 3;      // 0..7
    const bool has_aff = p.in_scale != nullptr;
    int q_grp = 0, stage = 0; uint32_t phase = 0;

    auto store4x4 = [&](uint8_t* tile, int row0, const float4 (&v)[4]) {
      // 4x4 transpose in registers: row e <- component e of the 4 pixels
      const float4 t0 = make_float4(v[0].x, v[1].x, v[2].x, v[3].x);
      const float4 t1 = make_float4(v[0].y, v[1].y, v[2].y, v[3].y);
      const float4 t2 = make_float4(v[0].z, v[1].z, v[2].z, v[3].z);
      const float4 t3 = make_float4(v[0].w, v[1].w, v[2].w, v[3].w);
      *reinterpret_cast<float4*>(tile + (row0 + 0) * 128 + ((pc ^ ((row0 + 0) & 7)) << 4)) = t0;
      *reinterpret_cast<float4*>(tile + (row0 + 1) * 128 + ((pc ^ ((row0 + 1) & 7)) << 4)) = t1;
      *reinterpret_cast<float4*>(tile + (row0 + 2) * 128 + ((pc ^ ((row0 + 2) & 7)) << 4)) = t2;
      *reinterpret_cast<float4*>(tile + (row0 + 3) * 128 + ((pc ^ ((row0 + 3) & 7)) << 4)) = t3;
    };

    for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x) {
      const int split = unit % cfg.splits;
      const int tile = unit / cfg.splits;
      const int mt = tile / cfg.n_tiles, nt = tile - mt * cfg.n_tiles;
      const int kb0 = split * cfg.kb_per_split;
      int kb1 = kb0 + cfg.kb_per_split;
      if (kb1 > cfg.kb_total) kb1 = cfg.kb_total;
      bool a_ok[4]; int a_r[4], a_s[4], a_off[4], a_c[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int k = mt * WG_BM + (rc0 + 8 * i) * 4;
        a_ok[i] = k < cfg.Ktot;
        const int tap = a_ok[i] ? (int)fdiv((uint32_t)k, cfg.d_c) : 0;
        a_c[i] = a_ok[i] ? k - tap * p.srcC : 0;
        a_r[i] = (int)fdiv((uint32_t)tap, cfg.d_s);
        a_s[i] = tap - a_r[i] * p.S;
        a_off[i] = (a_r[i] * p.srcW + a_s[i]) * p.src_ld + a_c[i];
      }
      const int n0 = nt * BN;
      for (int kb = kb0; kb < kb1; ++kb) {
        if (q_grp == grp) {
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          uint8_t* a_stage = sA + stage * WG_A_STAGE_BYTES;
          uint8_t* b_stage = sB + stage * b_stage_bytes;
          // the four pixels of this thread's chunk: origin offset (elements) and origin coordinates
          int pixoff[4], ihb[4], iwb[4];
          bool mok[4];
          const int mbase = kb * WG_BK + pc * 4;
#pragma unroll
          for (int qq = 0; qq < 4; ++qq) {
            const int m = mbase + qq;
            mok[qq] = m < cfg.M;
            const int mm = mok[qq] ? m : 0;
            const int img = (int)fdiv((uint32_t)mm, cfg.d_ohow);
            const int rem = mm - img * p.OH * p.OW;
            const int oh = (int)fdiv((uint32_t)rem, cfg.d_ow);
            const int ow = rem - oh * p.OW;
            ihb[qq] = mok[qq] ? oh * p.stride_h - p.pad_h : -(1 << 24);
            iwb[qq] = ow * p.stride_w - p.pad_w;
            pixoff[qq] = ((img * p.srcH + ihb[qq]) * p.srcW + iwb[qq]) * p.src_ld;
          }
          auto load_a = [&](int i, float4 (&v)[4]) {
            const float* srct = p.src + a_off[i];
#pragma unroll
            for (int qq = 0; qq < 4; ++qq) {
              const bool valid = (unsigned)(ihb[qq] + a_r[i]) < (unsigned)p.srcH &&
                                 (unsigned)(iwb[qq] + a_s[i]) < (unsigned)p.srcW;
              v[qq] = valid ? __ldg(reinterpret_cast<const float4*>(srct + pixoff[qq]))
                            : make_float4(__int_as_float(0x7fc00000), 0.f, 0.f, 0.f);   // NaN tag = "padding"
            }
          };
          auto finish_a = [&](int i, float4 (&v)[4]) {
            float4 sc = make_float4(1.f, 1.f, 1.f, 1.f), sh = make_float4(0.f, 0.f, 0.f, 0.f);
            if (has_aff) {
              sc = __ldg(reinterpret_cast<const float4*>(p.in_scale + a_c[i]));
              sh = __ldg(reinterpret_cast<const float4*>(p.in_shift + a_c[i]));
            }
#pragma unroll
            for (int qq = 0; qq < 4; ++qq) {
              float4 x = v[qq];
              if (x.x != x.x) {                     // padding tap: contributes zeros (never activated)
                x = make_float4(0.f, 0.f, 0.f, 0.f);
              } else if (has_aff) {
                x.x = fmaf(x.x, sc.x, sh.x); x.y = fmaf(x.y, sc.y, sh.y);
                x.z = fmaf(x.z, sc.z, sh.z); x.w = fmaf(x.w, sc.w, sh.w);
                if (p.in_relu) {
                  x.x = fmaxf(x.x, 0.f); x.y = fmaxf(x.y, 0.f); x.z = fmaxf(x.z, 0.f); x.w = fmaxf(x.w, 0.f);
                }
              }
              v[qq] = x;
            }
            store4x4(a_stage, (rc0 + 8 * i) * 4, v);
          };
          auto load_b = [&](int i, float4 (&v)[4]) {
            const int rc = rc0 + 8 * i;
            const int co = n0 + rc * 4;
#pragma unroll
            for (int qq = 0; qq < 4; ++qq) {
              float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
              if (mok[qq] && co < p.Nout) {
                const float* src = p.dy + (size_t)(mbase + qq) * p.dy_ld + co;
                if (cfg.vec_dy && co + 3 < p.Nout) {
                  t = __ldg(reinterpret_cast<const float4*>(src));
                } else {
                  t.x = __ldg(src);
                  if (co + 1 < p.Nout) t.y = __ldg(src + 1);
                  if (co + 2 < p.Nout) t.z = __ldg(src + 2);
                  if (co + 3 < p.Nout) t.w = __ldg(src + 3);
                }
              }
              v[qq] = t;
            }
          };
          // ---- issue: dY chunks 0,1 + A chunks 0,1; then A chunks 2,3 reuse the dY registers ----
          float4 vb0[4], vb1[4], va0[4], va1[4];
          const bool b0 = (rc0) * 4 < BN, b1 = (rc0 + 8) * 4 < BN;
          if (b0) load_b(0, vb0);
          if (b1) load_b(1, vb1);
          if (a_ok[0]) load_a(0, va0);
          if (a_ok[1]) load_a(1, va1);
          if (b0) store4x4(b_stage, rc0 * 4, vb0);
          if (b1) store4x4(b_stage, (rc0 + 8) * 4, vb1);
          if (a_ok[2]) load_a(2, vb0);
          if (a_ok[3]) load_a(3, vb1);
          if (a_ok[0]) finish_a(0, va0);
          if (a_ok[1]) finish_a(1, va1);
          if (a_ok[2]) finish_a(2, vb0);
          if (a_ok[3]) finish_a(3, vb1);
          // wide outputs (BN > 64): remaining dY chunks, two at a time
          for (int ib = 2; (rc0 + 8 * ib) * 4 < BN; ib += 2) {
            load_b(ib, va0);
            const bool more = (rc0 + 8 * (ib + 1)) * 4 < BN;
            if (more) load_b(ib + 1, va1);
            store4x4(b_stage, (rc0 + 8 * ib) * 4, va0);
            if (more) store4x4(b_stage, (rc0 + 8 * (ib + 1)) * 4, va1);
          }
          fence_proxy_async_smem();
          mbar_arrive(&full_bar[stage]);
        }
        if (++q_grp == cfg.active_groups) q_grp = 0;
        if (++stage == stages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == WG_MMA_WARP) {
    // =========================== MMA issuer ================================================
    const uint32_t idesc = make_idesc_tf32(WG_BM, (uint32_t)BN);
    const uint64_t adesc0 = make_desc_k_sw128(smem_u32(sA));
    const uint64_t bdesc0 = make_desc_k_sw128(smem_u32(sB));
    int stage = 0; uint32_t phase = 0;
    int local = 0;
    for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x, ++local) {
      const int split = unit % cfg.splits;
      const int kb0 = split * cfg.kb_per_split;
      int kb1 = kb0 + cfg.kb_per_split;
      if (kb1 > cfg.kb_total) kb1 = cfg.kb_total;
      const int acc = local & 1;
      const uint32_t acc_phase = (uint32_t)(local >> 1) & 1u;
      mbar_wait(&tempty_bar[acc], acc_phase ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
      uint32_t accumulate = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        {
          const uint64_t adesc = adesc0 + (uint64_t)((uint32_t)(stage * WG_A_STAGE_BYTES) >> 4);
          const uint64_t bdesc = bdesc0 + (uint64_t)((uint32_t)(stage * b_stage_bytes) >> 4);
          if (elect_one()) {        // elect.sync: one known issuing lane => descriptors stay in uniform registers
#pragma unroll
            for (int kk = 0; kk < WG_BK / 8; ++kk) {
              umma_tf32_ss(d_tmem, adesc + (uint64_t)(kk * 2), bdesc + (uint64_t)(kk * 2), idesc, accumulate);
              accumulate = 1;
            }
            umma_commit(&empty_bar[stage]);
          }
        }
        __syncwarp();
        accumulate = 1;
        if (++stage == stages) { stage = 0; phase ^= 1u; }
      }
      if (elect_one()) umma_commit(&tfull_bar[acc]);
      __syncwarp();
    }
  } else {
    // =========================== epilogue ==================================================
    int local = 0;
    for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x, ++local) {
      const int split = unit % cfg.splits;
      const int tile = unit / cfg.splits;
      const int mt = tile / cfg.n_tiles, nt = tile - mt * cfg.n_tiles;
      const int kb0 = split * cfg.kb_per_split;
      const bool nonempty = kb0 < cfg.kb_total;
      const int acc = local & 1;
      const uint32_t acc_phase = (uint32_t)(local >> 1) & 1u;
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const int k = mt * WG_BM + tid;
      const bool kvalid = nonempty && k < cfg.Ktot;
      const uint32_t t_row = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(acc * BN);
      for (int cc = 0; cc < BN; cc += 16) {
        const int nb = nt * BN + cc;
        if (nb >= p.Nout) break;
        float v[16];
        tmem_ld16(t_row + (uint32_t)cc, v);
        if (kvalid) {
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            const int n = nb + e;
            if (n < p.Nout) {
              float* dst = p.dw + (size_t)n * cfg.Ktot + k;
              if (cfg.splits == 1) *dst += v[e];
              else atomicAdd(dst, v[e]);
            }
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&tempty_bar[acc]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == WG_MMA_WARP) {
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)cfg.tmem_cols);
  }
}

}  // namespace dtf

namespace dtf {
int try_conv_wgrad_mn(const dtf_conv_wgrad_params& p, cudaStream_t stream, bool* launched);
}
using namespace dtf;

extern "C" int dtf_conv_wgrad(const dtf_conv_wgrad_params* pp, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  const dtf_conv_wgrad_params& p = *pp;
  DTF_REQUIRE(p.srcC % 4 == 0 && p.src_ld % 4 == 0, "source channels / pitch must be multiples of 4");
  DTF_REQUIRE((reinterpret_cast<uintptr_t>(p.src) & 15) == 0, "source must be 16-byte aligned");
  DTF_REQUIRE((long long)p.N * p.srcH * p.srcW * p.src_ld < (1LL << 31), "source tensor too large for 32-bit offsets");
  {
    // copy-only producers with MN-major operands (conv_wgrad_mn.cu) whenever the gradient rows are 16-byte chunks
    bool launched = false;
    const int rc = try_conv_wgrad_mn(p, stream, &launched);
    if (rc != 0 || launched) return rc;
  }
  WgradCfg cfg;
  cfg.Ktot = p.R * p.S * p.srcC;
  cfg.M = p.N * p.OH * p.OW;
  if (cfg.M == 0) return 0;
  cfg.BN = p.Nout <= 256 ? round_up(p.Nout, 16) : 256;
  cfg.n_tiles = ceil_div(p.Nout, cfg.BN);
  cfg.m_tiles = ceil_div(cfg.Ktot, WG_BM);
  cfg.kb_total = ceil_div(cfg.M, WG_BK);
  const int base_units = cfg.m_tiles * cfg.n_tiles;
  int splits = ceil_div(2 * num_sms(), base_units);
  const int max_splits = cfg.kb_total / 4 > 0 ? cfg.kb_total / 4 : 1;   // at least 4 k-blocks per split
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  cfg.kb_per_split = ceil_div(cfg.kb_total, splits);
  cfg.splits = ceil_div(cfg.kb_total, cfg.kb_per_split);
  int tc = 32;
  while (tc < 2 * cfg.BN) tc <<= 1;
  cfg.tmem_cols = tc;
  const int stage_bytes = WG_A_STAGE_BYTES + cfg.BN * WG_BK * 4;
  int stages = (SMEM_BUDGET - 3 * 1024) / stage_bytes;
  if (stages > 6) stages = 6;
  if (stages < 2) stages = 2;
  cfg.stages = stages;
  cfg.active_groups = stages < 4 ? stages : 4;     // a group may run at most one ring-phase ahead of the MMA
  cfg.d_ohow = make_fastdiv((uint32_t)(p.OH * p.OW));
  cfg.d_ow = make_fastdiv((uint32_t)p.OW);
  cfg.d_c = make_fastdiv((uint32_t)p.srcC);
  cfg.d_s = make_fastdiv((uint32_t)p.S);
  cfg.vec_dy = (p.dy_ld % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.dy) & 15) == 0);

  const size_t smem = 1024 + (size_t)stages * stage_bytes + (2 * stages + 4) * sizeof(uint64_t) + 16;
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    DTF_CHECK(cudaFuncSetAttribute(conv_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set[dev & 63] = true;
  }
  const int units = cfg.m_tiles * cfg.n_tiles * cfg.splits;
  const int grid = units < num_sms() ? units : num_sms();
  DTF_CHECK(launch_pdl(conv_wgrad_kernel, dim3(grid), dim3(WG_THREADS), smem, stream, p, cfg));
  return 0;
}

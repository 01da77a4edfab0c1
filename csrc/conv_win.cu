// K1w: stride-1 convolution (fprop and dgrad) that reads every input element ONCE — the "sliding window" feed.
//
// The implicit-GEMM kernel in conv_igemm.cu rebuilds the im2col matrix, so a 3x3 layer pulls each activation
// through L2 -> shared memory nine times; for the 16..64-channel layers of the CIFAR ResNet that traffic (and the
// TMA row rate behind it) is the whole cost.  Here the output pixels are numbered on a frame that shares its zero
// padding between neighbouring rows and images,
//
//     q = (img * Hp + y) * Wp + x,      Hp = H + gap_h, Wp = W + gap_w   (gap = reach of the filter past the edge)
//
// so that filter tap (r, s) of output q reads input q + off_h(r) * Wp + off_w(s) on the same frame, for every q.
// A tile of 128 consecutive q therefore needs one contiguous window of 128 + (R-1)*Wp + (S-1) input pixels, and the
// A operand of tap (r, s) is that window shifted by a constant.  The window is stored channel-chunk-major,
//
//     window[chunk j][pixel i] = 16 bytes (4 channels),  plane pitch = win_pad * 16 bytes,
//
// which is exactly tcgen05's K-major no-swizzle canonical layout (8-row core matrices of 8 x 16 B, 128 B apart; the
// two K chunks of a K=8 tf32 MMA one plane apart), and whose start address can move in whole pixels.  So:
//
//   warps 4-11  fill    : load each window pixel once (cp.async, or LDG + batch-norm apply + ReLU + STS when the
//                         producer layer's normalisation is fused in), zero for padding / frame-gap positions
//   warp  13    weights : TMA loads of the weight k-blocks (128B-swizzled K-major); kept resident for the whole
//                         kernel when they fit, otherwise streamed through a ring
//   warp  12    MMA     : R*S*C/8 tcgen05.mma.kind::tf32 (M=128, N=BN, K=8) per tile, A descriptor = window base +
//                         per-(tap, channel group) offset from a small table
//   warps 0-3   epilogue: conv_epilogue.cuh; rows that fall into the frame gaps are computed but not stored
//
// Frame-gap rows cost (Hp*Wp)/(H*W) extra MMA rows (6 % at 32x32, 27 % at 8x8): irrelevant next to a 9x cut in
// operand traffic.  Same numerics as conv_igemm.cu (tf32 multiply, fp32 accumulate).
//
// Reference ops replaced: as conv_igemm.cu (resnet/resnet_model.py:87 conv2d_fixed_padding and its input gradient).
#include "../include/dtf_b200.h"
#include "host_common.h"
#include "umma.cuh"
#include "ps_gate.cuh"
#include "conv_epilogue.cuh"

namespace dtf {

constexpr int WBM = 128;
constexpr int W_FILL_THREADS = 256;
constexpr int W_THREADS = 128 + W_FILL_THREADS + 64;
constexpr int W_MMA_WARP = (128 + W_FILL_THREADS) / 32;
constexpr int W_TMA_WARP = W_MMA_WARP + 1;
constexpr int W_MAX_K8 = 512;
constexpr int W_MAX_BSTAGES = 32;
constexpr int W_MAX_WIN = 4;

struct WinCfg {
  int BN, n_tiles, m_tiles, tmem_cols;
  int Hp, Wp, Q;
  int lin_min;        // most negative linear tap offset
  int win;            // window pixels filled per tile
  int win_pad;        // plane pitch in pixels (16-byte units), == 2 (mod 8) so plane fills spread over the banks
  int CB, cblocks;    // channels per window, windows per tile
  int n_k8;           // MMAs per window
  int n_kb;           // weight k-blocks (32 k) per window
  int b_stages, b_resident;
  int n_win;
  int vec_out, vec_res;
  FastDiv d_hpwp, d_wp, d_k8pt, d_s, d_kbpt;
  long long* trace;   // optional timeline buffer (debug), see dtf_debug_set_trace
};

extern long long* g_trace_host;
#define wtrace(slot)                                                                          \
  do {                                                                                        \
    if (cfg.trace != nullptr && blockIdx.x == 0 && (slot) < 4096) cfg.trace[(slot)] = clock64(); \
  } while (0)

// K-major, no swizzle: rows 16 B apart inside an 8-row core matrix, core matrices 128 B apart along M (SBO),
// the second 16-byte K chunk `lbo16` * 16 bytes away (LBO).
__device__ __forceinline__ uint64_t make_desc_k_nosw(uint32_t smem_addr, uint32_t lbo16, uint32_t sbo16) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(lbo16 & 0x3FFFu) << 16;
  d |= static_cast<uint64_t>(sbo16 & 0x3FFFu) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  return d;
}

__global__ void __launch_bounds__(W_THREADS, 1)
conv_win_kernel(const __grid_constant__ CUtensorMap tmapB, const dtf_conv_gemm_params p, const WinCfg cfg) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);

  const int BN = cfg.BN;
  const int b_stage_bytes = BN * 128;
  const int win_bytes = cfg.win_pad * cfg.CB * 4;
  uint8_t* sB = smem;
  uint8_t* sW = sB + cfg.b_stages * b_stage_bytes;
  uint64_t* full_b = reinterpret_cast<uint64_t*>(sW + cfg.n_win * win_bytes);
  uint64_t* empty_b = full_b + W_MAX_BSTAGES;
  uint64_t* win_full = empty_b + W_MAX_BSTAGES;
  uint64_t* win_empty = win_full + W_MAX_WIN;
  uint64_t* tfull_bar = win_empty + W_MAX_WIN;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
  float* csum = reinterpret_cast<float*>(tmem_slot + 4);
  float* csq = csum + 256;
  uint32_t* a_off = reinterpret_cast<uint32_t*>(csq + 256);     // [W_MAX_K8]

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  if (tid == 0) wtrace(3);
  pdl_launch_dependents();   // persistent grid, all CTAs resident: the successor may take SMs as they free up

  if (warp == W_MMA_WARP) {
    if (lane == 0) {
      for (int i = 0; i < cfg.b_stages; ++i) {
        mbar_init(&full_b[i], 1);
        mbar_init(&empty_b[i], 1);
      }
      for (int i = 0; i < cfg.n_win; ++i) {
        mbar_init(&win_full[i], W_FILL_THREADS);
        mbar_init(&win_empty[i], 1);
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
  // A-descriptor offset (16-byte units) of MMA t: tap t / (CB/8), channel group t % (CB/8)
  for (int t = tid; t < cfg.n_k8; t += W_THREADS) {
    const int tap = (int)fdiv((uint32_t)t, cfg.d_k8pt);
    const int c8 = t - tap * (cfg.CB >> 3);
    const int r = (int)fdiv((uint32_t)tap, cfg.d_s), s_ = tap - r * p.S;
    const int lin = (p.b_h + r * p.c_h) * cfg.Wp + (p.b_w + s_ * p.c_w) - cfg.lin_min;
    a_off[t] = (uint32_t)(lin + c8 * 2 * cfg.win_pad);
  }
  if (tid == W_TMA_WARP * 32) tma_prefetch_desc(&tmapB);
  pdl_wait();     // everything above is independent of the predecessor kernel
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int num_tiles = cfg.m_tiles * cfg.n_tiles;
  if (tid == 0) wtrace(4);

  if (warp >= 4 && warp < W_MMA_WARP) {
    // =========================== window fill ==================================================
    const int pt = tid - 128;
    const int planes = cfg.CB >> 2;                  // 4, 8 or 16
    const int j = pt & (planes - 1);
    const int i0 = pt / planes;
    const int istep = W_FILL_THREADS / planes;
    const bool has_aff = p.in_scale != nullptr;
    const bool relu = p.in_relu != 0;
    const bool write_act = has_aff && p.act_out != nullptr && cfg.cblocks >= 1;
    const int hpwp = cfg.Hp * cfg.Wp;
    int wslot = 0; uint32_t wphase = 0;
    int fseq = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int mt = tile / cfg.n_tiles;
      const int pbase = mt * WBM + cfg.lin_min;      // frame index of window pixel 0
      for (int cblk = 0; cblk < cfg.cblocks; ++cblk) {
        const int cbase = cblk * cfg.CB + j * 4;
        float4 sc = make_float4(1.f, 1.f, 1.f, 1.f), sh = make_float4(0.f, 0.f, 0.f, 0.f);
        if (has_aff) {
          sc = __ldg(reinterpret_cast<const float4*>(p.in_scale + cbase));
          sh = __ldg(reinterpret_cast<const float4*>(p.in_shift + cbase));
        }
        mbar_wait(&win_empty[wslot], wphase ^ 1u);
        if (pt == 0 && fseq < 256) wtrace(1024 + 2 * fseq);
        uint8_t* wbase = sW + wslot * win_bytes + (size_t)j * cfg.win_pad * 16;
        const uint32_t wbase_s = smem_u32(wbase);
        for (int ib = i0; ib < cfg.win; ib += 4 * istep) {
          int off[4];
          bool ok[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int i = ib + u * istep;
            const int pl = pbase + i;
            ok[u] = false;
            off[u] = 0;
            if (i < cfg.win && pl >= 0) {
              const int n = (int)fdiv((uint32_t)pl, cfg.d_hpwp);
              const int rem = pl - n * hpwp;
              const int yi = (int)fdiv((uint32_t)rem, cfg.d_wp);
              const int xi = rem - yi * cfg.Wp;
              if (n < p.N && yi < p.srcH && xi < p.srcW) {
                ok[u] = true;
                off[u] = ((n * p.srcH + yi) * p.srcW + xi) * p.src_ld + cbase;
              }
            }
          }
          if (!has_aff) {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int i = ib + u * istep;
              if (i < cfg.win) cp_async_16(wbase_s + (uint32_t)i * 16u, p.src + off[u], ok[u] ? 16u : 0u);
            }
          } else {
            float4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
              if (ok[u]) v[u] = __ldg(reinterpret_cast<const float4*>(p.src + off[u]));
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int i = ib + u * istep;
              float4 x = v[u];
              if (ok[u]) {
                x.x = fmaf(x.x, sc.x, sh.x); x.y = fmaf(x.y, sc.y, sh.y);
                x.z = fmaf(x.z, sc.z, sh.z); x.w = fmaf(x.w, sc.w, sh.w);
                if (relu) {
                  x.x = fmaxf(x.x, 0.f); x.y = fmaxf(x.y, 0.f); x.z = fmaxf(x.z, 0.f); x.w = fmaxf(x.w, 0.f);
                }
              }
              if (i < cfg.win) *reinterpret_cast<float4*>(wbase + (size_t)i * 16) = x;
              // side output act(src): each tile stores the pixels of its own range (no halo), once
              if (write_act && ok[u] && (unsigned)(i + cfg.lin_min) < (unsigned)WBM)
                *reinterpret_cast<float4*>(p.act_out + off[u]) = x;
            }
          }
        }
        if (!has_aff) {
          cp_async_mbar_arrive_noinc(&win_full[wslot]);
        } else {
          fence_proxy_async_smem();
          mbar_arrive(&win_full[wslot]);
        }
        if (pt == 0 && fseq < 256) wtrace(1025 + 2 * fseq);
        ++fseq;
        if (++wslot == cfg.n_win) { wslot = 0; wphase ^= 1u; }
      }
    }
  } else if (warp == W_TMA_WARP) {
    // =========================== weight loads ==================================================
    const int per32 = cfg.CB >= 32 ? (cfg.CB >> 5) : 1;
    weight_gate_warp(p);       // PS pull: the owners' updates of this layer's bucket must have landed (ps_gate.cuh)
    if (cfg.b_resident) {
      if (elect_one()) {
        for (int kb = 0; kb < cfg.n_kb; ++kb) {
          const int tap = (int)fdiv((uint32_t)kb, cfg.d_kbpt);
          const int k0 = cfg.CB == 16 ? kb * 32 : tap * p.srcC + (kb - tap * per32) * 32;
          mbar_arrive_expect_tx(&full_b[kb], (uint32_t)b_stage_bytes);
          tma_load_2d(sB + kb * b_stage_bytes, &tmapB, &full_b[kb], k0, 0);
        }
      }
      __syncwarp();
    } else {
      int bstage = 0; uint32_t bphase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int mt = tile / cfg.n_tiles, nt = tile - mt * cfg.n_tiles;
        const int n0 = nt * BN;
        for (int cblk = 0; cblk < cfg.cblocks; ++cblk) {
          for (int kb = 0; kb < cfg.n_kb; ++kb) {
            mbar_wait(&empty_b[bstage], bphase ^ 1u);
            if (elect_one()) {
              const int tap = (int)fdiv((uint32_t)kb, cfg.d_kbpt);
              const int k0 = cfg.CB == 16 ? kb * 32 : tap * p.srcC + cblk * cfg.CB + (kb - tap * per32) * 32;
              mbar_arrive_expect_tx(&full_b[bstage], (uint32_t)b_stage_bytes);
              tma_load_2d(sB + bstage * b_stage_bytes, &tmapB, &full_b[bstage], k0, n0);
            }
            __syncwarp();
            if (++bstage == cfg.b_stages) { bstage = 0; bphase ^= 1u; }
          }
        }
      }
    }
  } else if (warp == W_MMA_WARP) {
    // =========================== MMA issuer ====================================================
    const uint32_t idesc = make_idesc_tf32(WBM, (uint32_t)BN);
    const uint64_t adesc0 = make_desc_k_nosw(smem_u32(sW), (uint32_t)cfg.win_pad, 8u);
    const uint64_t bdesc0 = make_desc_k_sw128(smem_u32(sB));
    const uint32_t a_hi = (uint32_t)(adesc0 >> 32), b_hi = (uint32_t)(bdesc0 >> 32);
    int wslot = 0; uint32_t wphase = 0;
    int bstage = 0; uint32_t bphase = 0;
    int local = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
      const int acc = local & 1;
      const uint32_t acc_phase = (uint32_t)(local >> 1) & 1u;
      mbar_wait(&tempty_bar[acc], acc_phase ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
      bool first = true;
      for (int cblk = 0; cblk < cfg.cblocks; ++cblk) {
        mbar_wait(&win_full[wslot], wphase);
        tc_fence_after();
        if (lane == 0 && local < 256) wtrace(2048 + 2 * local);
        const uint32_t a_lo0 = (uint32_t)adesc0 + ((uint32_t)(wslot * win_bytes) >> 4);
        for (int kb = 0; kb < cfg.n_kb; ++kb) {
          if (!cfg.b_resident || local == 0) {
            mbar_wait(&full_b[bstage], bphase);
            tc_fence_after();
          }
          const int t0 = kb * 4;
          const int nm = cfg.n_k8 - t0 < 4 ? cfg.n_k8 - t0 : 4;
          // Only the A start address differs between the MMAs of a k-block; everything else stays in uniform
          // registers (every value moved from a vector register costs the issuing thread an R2UR round trip).
          uint32_t al[4];
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) al[kk] = a_lo0 + a_off[t0 + kk < cfg.n_k8 ? t0 + kk : t0];
          const uint32_t b_lo = (uint32_t)bdesc0 + ((uint32_t)(bstage * b_stage_bytes) >> 4);
          if (elect_one()) {
            umma_tf32_ss_lohi_p(d_tmem, al[0], a_hi, b_lo, b_hi, idesc, first ? 0u : 1u);
            if (nm > 1) umma_tf32_ss_lohi<1>(d_tmem, al[1], a_hi, b_lo + 2u, b_hi, idesc);
            if (nm > 2) umma_tf32_ss_lohi<1>(d_tmem, al[2], a_hi, b_lo + 4u, b_hi, idesc);
            if (nm > 3) umma_tf32_ss_lohi<1>(d_tmem, al[3], a_hi, b_lo + 6u, b_hi, idesc);
            if (!cfg.b_resident) umma_commit(&empty_b[bstage]);
          }
          __syncwarp();
          first = false;
          if (++bstage == cfg.b_stages) { bstage = 0; bphase ^= 1u; }
        }
        if (elect_one()) umma_commit(&win_empty[wslot]);
        __syncwarp();
        if (++wslot == cfg.n_win) { wslot = 0; wphase ^= 1u; }
      }
      if (elect_one()) umma_commit(&tfull_bar[acc]);
      __syncwarp();
      if (lane == 0 && local < 256) wtrace(2049 + 2 * local);
    }
  } else {
    // =========================== epilogue (warps 0-3) ==========================================
    int local = 0;
    const bool do_stats = p.stats != nullptr;
    const int hpwp = cfg.Hp * cfg.Wp;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
      const int mt = tile / cfg.n_tiles, nt = tile - mt * cfg.n_tiles;
      const int n0 = nt * BN;
      const int acc = local & 1;
      const uint32_t acc_phase = (uint32_t)(local >> 1) & 1u;
      // frame index -> output pixel (done before the wait: independent of the accumulator)
      const int q = mt * WBM + tid;
      const int n = (int)fdiv((uint32_t)q, cfg.d_hpwp);
      const int rem = q - n * hpwp;
      const int y = (int)fdiv((uint32_t)rem, cfg.d_wp);
      const int x = rem - y * cfg.Wp;
      const bool mvalid = q < cfg.Q && n < p.N && y < p.OH && x < p.OW;
      const size_t m = mvalid ? (size_t)((n * p.OH + y) * p.OW + x) : 0;
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      if (tid == 0 && local < 128) wtrace(512 + 2 * local);
      const uint32_t t_row = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(acc * BN);
      conv_epilogue_row(p, BN, n0, mvalid, m, t_row, cfg.vec_out, cfg.vec_res, do_stats, csum, csq, lane,
                        (cfg.trace != nullptr && blockIdx.x == 0 && tid == 0 && local < 128) ? cfg.trace + 3072 + 4 * local
                                                                                          : nullptr);
      tc_fence_before();
      mbar_arrive(&tempty_bar[acc]);
      if (tid == 0 && local < 128) wtrace(513 + 2 * local);
      if (do_stats && cfg.n_tiles > 1) {
        named_bar_sync(1, 128);
        for (int i = tid; i < BN; i += 128) {
          const int nn = n0 + i;
          if (nn < p.Nout) {
            atomicAdd(p.stats + nn, csum[i]);
            atomicAdd(p.stats + p.Nout + nn, csq[i]);
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
  if (tid == 0) wtrace(5);
  if (warp == W_MMA_WARP) {
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)cfg.tmem_cols);
  }
}

static int g_conv_win_mode = 1;     // 0 off, 1 automatic (<= 64-channel sources), 2 whenever the shape is eligible

static int win_pick_bn(int nout) {
  if (nout <= 256) return round_up(nout, 16);
  if (nout % 256 == 0) return 256;
  if (nout % 192 == 0) return 192;
  if (nout % 128 == 0) return 128;
  return 256;
}

// Launches the window kernel when the problem fits it (*launched = true); otherwise leaves *launched false and the
// caller uses the implicit-GEMM kernel.  Returns 0 or an error code.
int try_conv_win(const dtf_conv_gemm_params& p, cudaStream_t stream, bool* launched, bool dry_run) {
  *launched = false;
  if (g_conv_win_mode == 0) return 0;
  if (p.a_h != 1 || p.a_w != 1 || p.div_h != 1 || p.div_w != 1) return 0;
  if (p.R * p.S < 2) return 0;                                   // 1x1 / FC: nothing to reuse
  if (p.R > 1 && p.c_h != 1 && p.c_h != -1) return 0;
  if (p.S > 1 && p.c_w != 1 && p.c_w != -1) return 0;
  const int C = p.srcC;
  int CB;
  if (C == 16 || C == 32) CB = C;
  else if (C % 64 == 0) CB = 64;
  else return 0;
  if (g_conv_win_mode == 1 && C > 64) return 0;
  if ((reinterpret_cast<uintptr_t>(p.src) & 15) != 0 || p.src_ld % 4 != 0) return 0;

  WinCfg cfg;
  const int eh = (p.R - 1) * p.c_h, ew = (p.S - 1) * p.c_w;
  const int min_h = p.b_h + (eh < 0 ? eh : 0), max_h = p.b_h + (eh > 0 ? eh : 0);
  const int min_w = p.b_w + (ew < 0 ? ew : 0), max_w = p.b_w + (ew > 0 ? ew : 0);
  auto pos = [](int v) { return v > 0 ? v : 0; };
  const int gap_h = pos(-min_h) > pos(p.OH - 1 + max_h - (p.srcH - 1)) ? pos(-min_h) : pos(p.OH - 1 + max_h - (p.srcH - 1));
  const int gap_w = pos(-min_w) > pos(p.OW - 1 + max_w - (p.srcW - 1)) ? pos(-min_w) : pos(p.OW - 1 + max_w - (p.srcW - 1));
  cfg.Hp = (p.srcH > p.OH ? p.srcH : p.OH) + gap_h;
  cfg.Wp = (p.srcW > p.OW ? p.srcW : p.OW) + gap_w;
  const long long Q = (long long)p.N * cfg.Hp * cfg.Wp;
  const long long M = (long long)p.N * p.OH * p.OW;
  if (M == 0) return 0;
  if (Q > (1LL << 30) || Q * 10 > M * 16) return 0;             // frame overhead above 60 %: not worth it
  cfg.Q = (int)Q;
  cfg.lin_min = min_h * cfg.Wp + min_w;
  const int lin_max = max_h * cfg.Wp + max_w;
  cfg.win = WBM + lin_max - cfg.lin_min;
  cfg.win_pad = ceil_div(cfg.win - 2, 8) * 8 + 2;
  cfg.CB = CB;
  cfg.cblocks = C / CB;
  cfg.n_k8 = p.R * p.S * CB / 8;
  if (cfg.n_k8 > W_MAX_K8) return 0;
  cfg.n_kb = ceil_div(cfg.n_k8, 4);
  cfg.BN = win_pick_bn(p.Nout);
  cfg.n_tiles = ceil_div(p.Nout, cfg.BN);
  cfg.m_tiles = ceil_div(cfg.Q, WBM);
  int cols = 32;
  while (cols < 2 * cfg.BN) cols <<= 1;
  cfg.tmem_cols = cols;
  const int b_stage_bytes = cfg.BN * 128;
  const int win_bytes = cfg.win_pad * CB * 4;
  const int fixed = 1024 + (2 * W_MAX_BSTAGES + 2 * W_MAX_WIN + 4) * 8 + 16 + 2 * 256 * 4 + W_MAX_K8 * 4;
  const int avail = SMEM_BUDGET - fixed;
  cfg.b_resident = (cfg.cblocks == 1 && cfg.n_tiles == 1 && cfg.n_kb <= W_MAX_BSTAGES &&
                    cfg.n_kb * b_stage_bytes <= 80 * 1024) ? 1 : 0;
  if (cfg.b_resident) {
    cfg.b_stages = cfg.n_kb;
    int nw = (avail - cfg.b_stages * b_stage_bytes) / win_bytes;
    if (nw < 2) return 0;
    cfg.n_win = nw > W_MAX_WIN ? W_MAX_WIN : nw;
  } else {
    int nw = 3;
    if (avail - nw * win_bytes < 4 * b_stage_bytes) nw = 2;
    int bs = (avail - nw * win_bytes) / b_stage_bytes;
    if (bs < 3) return 0;
    cfg.n_win = nw;
    cfg.b_stages = bs > 8 ? 8 : bs;
  }
  cfg.d_hpwp = make_fastdiv((uint32_t)(cfg.Hp * cfg.Wp));
  cfg.d_wp = make_fastdiv((uint32_t)cfg.Wp);
  cfg.d_k8pt = make_fastdiv((uint32_t)(CB / 8));
  cfg.d_s = make_fastdiv((uint32_t)p.S);
  cfg.d_kbpt = make_fastdiv((uint32_t)(CB >= 32 ? CB / 32 : 1));
  cfg.trace = g_trace_host;
  cfg.vec_out = (p.out_ld % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.out) & 15) == 0);
  cfg.vec_res = p.residual && (p.res_ld % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.residual) & 15) == 0);

  if (dry_run) {
    *launched = true;
    return 0;
  }
  const int Ktot = p.R * p.S * C;
  const CUtensorMap* tm = get_tmap_2d_f32_sw128(p.wt, p.Nout, Ktot, cfg.BN);
  if (!tm) return -2;
  const size_t smem = (size_t)fixed + (size_t)cfg.b_stages * b_stage_bytes + (size_t)cfg.n_win * win_bytes;
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    DTF_CHECK(cudaFuncSetAttribute(conv_win_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set[dev & 63] = true;
  }
  const int tiles = cfg.m_tiles * cfg.n_tiles;
  const int grid = tiles < num_sms() ? tiles : num_sms();
  DTF_CHECK(launch_pdl(conv_win_kernel, dim3(grid), dim3(W_THREADS), smem, stream, *tm, p, cfg));
  *launched = true;
  return 0;
}

}  // namespace dtf

extern "C" int dtf_debug_set_conv_win(int mode) {
  dtf::g_conv_win_mode = mode;
  return 0;
}

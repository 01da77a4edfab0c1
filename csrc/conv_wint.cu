// K1t: sliding-window convolution with the operands swapped — channels on the tensor core's M axis, pixels on N.
//
// Measured on B200 (scripts/trace_conv.py): a tcgen05.mma whose A operand comes from shared memory streams A at
// one 32-byte row per cycle, so an M=128 instruction costs ~128 cycles however small N is.  With pixels on M and a
// 16..64-channel layer on N (conv_win.cu / conv_igemm.cu) the tensor core spends those 128 cycles on 128 x 16 x 8
// MACs — 1/16 of its rate — and that A stream is the whole kernel time.  Here
//
//     D^T[co, q] = sum_k  W[co, k] * window[q + off(k), c(k)]
//
// A is the (tiny) weight k-block: M = 64 rows, of which C_out are real; B is the sliding window of conv_win.cu,
// now N = up to 256 frame pixels per instruction.  One MMA covers 2-4x the pixels for half the A rows.
//
// The accumulator arrives channel-major in TMEM (lane = channel, column = pixel).  Epilogue warps read 16 pixels
// of their channel with tcgen05.ld, transpose through a per-warp shared-memory patch, and continue in NHWC order:
// one thread = 4 channels of one pixel = one 16-byte store; bias / residual / ReLU / batch-norm statistics (or the
// fused batch-norm-backward sums) are applied in that phase.  The per-channel sums stay in registers for the whole
// kernel (a thread always owns the same 4 channels) and are reduced once at the end.
//
//   warps 0-7   epilogue : warp w reads TMEM lane quadrant w % 4, column chunks w / 4, w / 4 + 2, ...
//   warps 8-11  fill     : window pixels -> [chunk][pixel] planes (cp.async, or LDG + BN-apply + ReLU + STS)
//   warp  12    MMA      : R*S*C/8 tcgen05.mma.kind::tf32 (M = 64 | 128, N = tile pixels, K = 8) per tile
//   warp  13    weights  : TMA loads of the weight k-blocks (resident or ring), as conv_win.cu
//
// Reference ops replaced: as conv_igemm.cu.
#include "../include/dtf_b200.h"
#include "host_common.h"
#include "umma.cuh"
#include "ps_gate.cuh"

namespace dtf {

constexpr int T_EPI_WARPS = 8;
constexpr int T_EPI_THREADS = T_EPI_WARPS * 32;
constexpr int T_FILL_THREADS = 128;
constexpr int T_THREADS = T_EPI_THREADS + T_FILL_THREADS + 64;
constexpr int T_MMA_WARP = (T_EPI_THREADS + T_FILL_THREADS) / 32;
constexpr int T_TMA_WARP = T_MMA_WARP + 1;
constexpr int T_MAX_K8 = 512;
constexpr int T_MAX_BSTAGES = 32;
constexpr int T_MAX_WIN = 4;
constexpr int T_MAX_ITEMS = 8;              // (pixel, 4-channel) items per epilogue thread and tile

struct WinTCfg {
  int TN;             // pixels per tile (UMMA N): multiple of 16, <= 256
  int UM;             // UMMA M: 64 or 128
  int lane_mode;      // 1: M=64 accumulator rows 16g..16g+15 live in lanes 0..15 of TMEM quadrant g; 0: row r = lane r
  int rows_b;         // weight rows per k-block tile (C_out rounded up to 16)
  int m_tiles, tmem_cols;
  int Hp, Wp, Q;
  int lin_min, win, win_pad;
  int CB, cblocks, n_k8, n_kb;
  int b_stages, b_resident, n_win;
  int patch_floats;   // TN * C_out
  FastDiv d_hpwp, d_wp, d_k8pt, d_s, d_kbpt;
  long long* trace;
};

extern long long* g_trace_host;
#define ttrace(slot)                                                                          \
  do {                                                                                        \
    if (cfg.trace != nullptr && blockIdx.x == 0 && (slot) < 4096) cfg.trace[(slot)] = clock64(); \
  } while (0)

__device__ __forceinline__ uint64_t make_desc_k_nosw_t(uint32_t smem_addr, uint32_t lbo16, uint32_t sbo16) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(lbo16 & 0x3FFFu) << 16;
  d |= static_cast<uint64_t>(sbo16 & 0x3FFFu) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  return d;
}

__global__ void __launch_bounds__(T_THREADS, 1)
conv_wint_kernel(const __grid_constant__ CUtensorMap tmapW, const dtf_conv_gemm_params p, const WinTCfg cfg) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);

  const int TN = cfg.TN;
  const int b_stage_bytes = 64 * 128;      // 64 rows: the C_out weight rows repeated 64 / C_out times
  const int rep_bytes = p.Nout * 128;
  const int win_bytes = cfg.win_pad * cfg.CB * 4;
  uint8_t* sB = smem;                                             // weight k-blocks (A operand)
  uint8_t* sW = sB + cfg.b_stages * b_stage_bytes;
  uint64_t* full_b = reinterpret_cast<uint64_t*>(sW + cfg.n_win * win_bytes);
  uint64_t* empty_b = full_b + T_MAX_BSTAGES;
  uint64_t* win_full = empty_b + T_MAX_BSTAGES;
  uint64_t* win_empty = win_full + T_MAX_WIN;
  uint64_t* tfull_bar = win_empty + T_MAX_WIN;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
  uint32_t* a_off = tmem_slot + 4;                                // [T_MAX_K8]
  int* mtab = reinterpret_cast<int*>(a_off + T_MAX_K8);           // [2][256] frame pixel -> output pixel (or -1)
  float* ssum = reinterpret_cast<float*>(mtab + 512);             // [2][128] per-channel sums (kernel end)
  float* s_aff = ssum + 256;                                      // [2][64] input transform: scale | shift
  float* stag = s_aff + 128;                                      // [2][patch_floats] pixel-major accumulator patch

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  if (tid == 0) ttrace(3);
  pdl_launch_dependents();   // persistent grid, all CTAs resident: the successor may take SMs as they free up

  if (warp == T_MMA_WARP) {
    if (lane == 0) {
      for (int i = 0; i < cfg.b_stages; ++i) {
        mbar_init(&full_b[i], 1);
        mbar_init(&empty_b[i], 1);
      }
      for (int i = 0; i < cfg.n_win; ++i) {
        mbar_init(&win_full[i], T_FILL_THREADS);
        mbar_init(&win_empty[i], 1);
      }
      for (int i = 0; i < 2; ++i) {
        mbar_init(&tfull_bar[i], 1);
        mbar_init(&tempty_bar[i], T_EPI_THREADS);
      }
      fence_mbar_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, (uint32_t)cfg.tmem_cols);
    tmem_relinquish();
  }
  if (tid < 256) ssum[tid] = 0.f;
  pdl_wait();     // everything above is independent of the predecessor kernel
  if (p.in_scale != nullptr && tid < p.srcC) {
    // input transform of this layer: given, or derived here from the producer's batch statistics (deferred
    // batch-norm finalize; CTA 0 also publishes it for the backward pass and updates the moving averages)
    float sc, sh;
    if (p.fin_stats != nullptr) {
      const float mean = p.fin_stats[tid] * p.fin_inv_count;
      float var = p.fin_stats[p.srcC + tid] * p.fin_inv_count - mean * mean;
      var = var < 0.f ? 0.f : var;
      const float invstd = rsqrtf(var + p.fin_eps);
      sc = (p.fin_gamma ? p.fin_gamma[tid] : 1.f) * invstd;
      sh = (p.fin_beta ? p.fin_beta[tid] : 0.f) - mean * sc;
      if (blockIdx.x == 0) {
        const_cast<float*>(p.in_scale)[tid] = sc;
        const_cast<float*>(p.in_shift)[tid] = sh;
        if (p.fin_mean) p.fin_mean[tid] = mean;
        if (p.fin_invstd) p.fin_invstd[tid] = invstd;
        if (p.fin_update_moving) {
          p.fin_moving_mean[tid] = p.fin_moving_mean[tid] * p.fin_decay + mean * (1.f - p.fin_decay);
          p.fin_moving_var[tid] = p.fin_moving_var[tid] * p.fin_decay + var * p.fin_unbias * (1.f - p.fin_decay);
        }
      }
    } else {
      sc = p.in_scale[tid];
      sh = p.in_shift[tid];
    }
    s_aff[tid] = sc;
    s_aff[64 + tid] = sh;
  }
  for (int t = tid; t < cfg.n_k8; t += T_THREADS) {
    const int tap = (int)fdiv((uint32_t)t, cfg.d_k8pt);
    const int c8 = t - tap * (cfg.CB >> 3);
    const int r = (int)fdiv((uint32_t)tap, cfg.d_s), s_ = tap - r * p.S;
    const int lin = (p.b_h + r * p.c_h) * cfg.Wp + (p.b_w + s_ * p.c_w) - cfg.lin_min;
    a_off[t] = (uint32_t)(lin + c8 * 2 * cfg.win_pad);
  }
  if (tid == T_TMA_WARP * 32) tma_prefetch_desc(&tmapW);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int num_tiles = cfg.m_tiles;
  if (tid == 0) ttrace(4);

  if (warp >= T_EPI_WARPS && warp < T_MMA_WARP) {
    // =========================== window fill ==================================================
    const int pt = tid - T_EPI_THREADS;
    const int planes = cfg.CB >> 2;                  // 4, 8 or 16
    const int j = pt & (planes - 1);
    const int i0 = pt / planes;
    const int istep = T_FILL_THREADS / planes;
    const bool has_aff = p.in_scale != nullptr;
    const bool relu = p.in_relu != 0;
    const bool write_act = has_aff && p.act_out != nullptr && cfg.cblocks >= 1;
    const int hpwp = cfg.Hp * cfg.Wp;
    const int step_q = istep / cfg.Wp, step_r = istep - step_q * cfg.Wp;
    const int step_off = (step_q * p.srcW + step_r) * p.src_ld;       // offset change of (dy, dx) = (step_q, step_r)
    const int wrap_x_off = (p.srcW - cfg.Wp) * p.src_ld;              // column wrapped: x -= Wp, y += 1
    const int wrap_y_off = (p.srcH - cfg.Hp) * p.srcW * p.src_ld;     // row wrapped:    y -= Hp, n += 1
    int wslot = 0; uint32_t wphase = 0;
    int fseq = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int pbase = tile * TN + cfg.lin_min;     // frame index of window pixel 0
      for (int cblk = 0; cblk < cfg.cblocks; ++cblk) {
        const int cbase = cblk * cfg.CB + j * 4;
        float4 sc = make_float4(1.f, 1.f, 1.f, 1.f), sh = make_float4(0.f, 0.f, 0.f, 0.f);
        if (has_aff) {
          sc = *reinterpret_cast<const float4*>(s_aff + cbase);
          sh = *reinterpret_cast<const float4*>(s_aff + 64 + cbase);
        }
        mbar_wait(&win_empty[wslot], wphase ^ 1u);
        if (pt == 0 && fseq < 256) ttrace(1024 + 2 * fseq);
        uint8_t* wbase = sW + wslot * win_bytes + (size_t)j * cfg.win_pad * 16;
        const uint32_t wbase_s = smem_u32(wbase);
        // frame coordinates of this thread's first window pixel; later pixels follow incrementally (the frame
        // index is biased by one image so that the few pixels before the tensor decode without negative division)
        int fn, fy, fx, foff;
        {
          const int u0 = pbase + i0 + hpwp;
          fn = (int)fdiv((uint32_t)u0, cfg.d_hpwp);
          const int rem = u0 - fn * hpwp;
          fy = (int)fdiv((uint32_t)rem, cfg.d_wp);
          fx = rem - fy * cfg.Wp;
          fn -= 1;
          foff = ((fn * p.srcH + fy) * p.srcW + fx) * p.src_ld + cbase;     // meaningful only where the pixel is valid
        }
        // Pass 0 issues every copy of the window (zero-filled where the frame has padding): all of them are in
        // flight at once, so the window costs one memory latency.  With a fused batch-norm + ReLU, pass 1 walks the
        // same pixels again once the thread's copies have landed and rewrites them in place (padding stays zero),
        // storing act(src) for the tile's own pixel range on request.
        const int fn0 = fn, fy0 = fy, fx0 = fx, foff0 = foff;
        for (int pass = 0; pass < (has_aff ? 2 : 1); ++pass) {
          if (pass == 1) {
            asm volatile("cp.async.wait_all;" ::: "memory");
            fn = fn0; fy = fy0; fx = fx0; foff = foff0;
          }
          for (int ib = i0; ib < cfg.win; ib += 4 * istep) {
            int off[4];
            bool ok[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int i = ib + u * istep;
              ok[u] = i < cfg.win && (unsigned)fn < (unsigned)p.N && fy < p.srcH && fx < p.srcW;
              off[u] = ok[u] ? foff : 0;
              // advance by istep frame pixels without divisions or loops: istep = step_q * Wp + step_r, and one
              // step can wrap the column once and the row once (step_q + 1 <= Hp)
              fx += step_r; fy += step_q; foff += step_off;
              if (fx >= cfg.Wp) { fx -= cfg.Wp; fy += 1; foff += wrap_x_off; }
              if (fy >= cfg.Hp) { fy -= cfg.Hp; fn += 1; foff += wrap_y_off; }
            }
            if (pass == 0) {
#pragma unroll
              for (int u = 0; u < 4; ++u) {
                const int i = ib + u * istep;
                if (i < cfg.win) cp_async_16(wbase_s + (uint32_t)i * 16u, p.src + off[u], ok[u] ? 16u : 0u);
              }
            } else {
              float4 v[4];
#pragma unroll
              for (int u = 0; u < 4; ++u) {
                const int i = ib + u * istep;
                if (ok[u]) v[u] = *reinterpret_cast<const float4*>(wbase + (size_t)i * 16);
              }
#pragma unroll
              for (int u = 0; u < 4; ++u) {
                if (ok[u]) {
                  const int i = ib + u * istep;
                  float4 x = v[u];
                  x.x = fmaf(x.x, sc.x, sh.x); x.y = fmaf(x.y, sc.y, sh.y);
                  x.z = fmaf(x.z, sc.z, sh.z); x.w = fmaf(x.w, sc.w, sh.w);
                  if (relu) {
                    x.x = fmaxf(x.x, 0.f); x.y = fmaxf(x.y, 0.f); x.z = fmaxf(x.z, 0.f); x.w = fmaxf(x.w, 0.f);
                  }
                  *reinterpret_cast<float4*>(wbase + (size_t)i * 16) = x;
                  if (write_act && (unsigned)(i + cfg.lin_min) < (unsigned)TN)
                    *reinterpret_cast<float4*>(p.act_out + off[u]) = x;
                }
              }
            }
          }
        }
        if (!has_aff) {
          cp_async_mbar_arrive_noinc(&win_full[wslot]);
        } else {
          fence_proxy_async_smem();
          mbar_arrive(&win_full[wslot]);
        }
        if (pt == 0 && fseq < 256) ttrace(1025 + 2 * fseq);
        ++fseq;
        if (++wslot == cfg.n_win) { wslot = 0; wphase ^= 1u; }
      }
    }
  } else if (warp == T_TMA_WARP) {
    // =========================== weight loads ==================================================
    const int per32 = cfg.CB >= 32 ? (cfg.CB >> 5) : 1;
    weight_gate_warp(p);       // PS pull: the owners' updates of this layer's bucket must have landed (ps_gate.cuh)
    if (cfg.b_resident) {
      if (elect_one()) {
        for (int kb = 0; kb < cfg.n_kb; ++kb) {
          const int tap = (int)fdiv((uint32_t)kb, cfg.d_kbpt);
          const int k0 = cfg.CB == 16 ? kb * 32 : tap * p.srcC + (kb - tap * per32) * 32;
          mbar_arrive_expect_tx(&full_b[kb], (uint32_t)b_stage_bytes);
          for (int rep = 0; rep * p.Nout < 64; ++rep)
            tma_load_2d(sB + kb * b_stage_bytes + rep * rep_bytes, &tmapW, &full_b[kb], k0, 0);
        }
      }
      __syncwarp();
    } else {
      int bstage = 0; uint32_t bphase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        for (int cblk = 0; cblk < cfg.cblocks; ++cblk) {
          for (int kb = 0; kb < cfg.n_kb; ++kb) {
            mbar_wait(&empty_b[bstage], bphase ^ 1u);
            if (elect_one()) {
              const int tap = (int)fdiv((uint32_t)kb, cfg.d_kbpt);
              const int k0 = cfg.CB == 16 ? kb * 32 : tap * p.srcC + cblk * cfg.CB + (kb - tap * per32) * 32;
              mbar_arrive_expect_tx(&full_b[bstage], (uint32_t)b_stage_bytes);
              for (int rep = 0; rep * p.Nout < 64; ++rep)
                tma_load_2d(sB + bstage * b_stage_bytes + rep * rep_bytes, &tmapW, &full_b[bstage], k0, 0);
            }
            __syncwarp();
            if (++bstage == cfg.b_stages) { bstage = 0; bphase ^= 1u; }
          }
        }
      }
    }
  } else if (warp == T_MMA_WARP) {
    // =========================== MMA issuer ====================================================
    const uint32_t idesc = make_idesc_tf32((uint32_t)cfg.UM, (uint32_t)TN);
    const uint64_t wdesc0 = make_desc_k_sw128(smem_u32(sB));                           // A: weights
    const uint64_t pdesc0 = make_desc_k_nosw_t(smem_u32(sW), (uint32_t)cfg.win_pad, 8u);   // B: window
    const uint32_t w_hi = (uint32_t)(wdesc0 >> 32), p_hi = (uint32_t)(pdesc0 >> 32);
    int wslot = 0; uint32_t wphase = 0;
    int bstage = 0; uint32_t bphase = 0;
    int local = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
      const int acc = local & 1;
      const uint32_t acc_phase = (uint32_t)(local >> 1) & 1u;
      mbar_wait(&tempty_bar[acc], acc_phase ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * TN);
      bool first = true;
      for (int cblk = 0; cblk < cfg.cblocks; ++cblk) {
        mbar_wait(&win_full[wslot], wphase);
        tc_fence_after();
        if (lane == 0 && local < 256) ttrace(2048 + 2 * local);
        const uint32_t p_lo0 = (uint32_t)pdesc0 + ((uint32_t)(wslot * win_bytes) >> 4);
        for (int kb = 0; kb < cfg.n_kb; ++kb) {
          if (!cfg.b_resident || local == 0) {
            mbar_wait(&full_b[bstage], bphase);
            tc_fence_after();
          }
          const int t0 = kb * 4;
          const int nm = cfg.n_k8 - t0 < 4 ? cfg.n_k8 - t0 : 4;
          uint32_t pl[4];
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) pl[kk] = p_lo0 + a_off[t0 + kk < cfg.n_k8 ? t0 + kk : t0];
          const uint32_t w_lo = (uint32_t)wdesc0 + ((uint32_t)(bstage * b_stage_bytes) >> 4);
          if (elect_one()) {
            umma_tf32_ss_lohi_p(d_tmem, w_lo, w_hi, pl[0], p_hi, idesc, first ? 0u : 1u);
            if (nm > 1) umma_tf32_ss_lohi<1>(d_tmem, w_lo + 2u, w_hi, pl[1], p_hi, idesc);
            if (nm > 2) umma_tf32_ss_lohi<1>(d_tmem, w_lo + 4u, w_hi, pl[2], p_hi, idesc);
            if (nm > 3) umma_tf32_ss_lohi<1>(d_tmem, w_lo + 6u, w_hi, pl[3], p_hi, idesc);
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
      if (lane == 0 && local < 256) ttrace(2049 + 2 * local);
    }
  } else {
    // =========================== epilogue (warps 0-7) ==========================================
    // Per tile:  (a) frame-pixel -> output-pixel table;  (b) every thread requests the residual / BN-backward rows
    // of its items (4 channels of one pixel each) — long before the accumulator is ready, so the global latency
    // hides behind the tile's MMAs;  (c) the warps whose TMEM quadrant holds channels move the accumulator into
    // the pixel-major patch [pixel][C_out] and TMEM is handed back;  (d) all 256 threads finish their items.
    // M = 64: accumulator rows 16g..16g+15 sit in lanes 0..15 of TMEM quadrant g, and row r holds channel
    // r % C_out (the weight rows are repeated), so every quadrant carries channels and all eight warps share the
    // TMEM -> patch move: the 2 * 64 / C_out warps that see the same channels take interleaved column chunks.
    const int qd = warp & 3;                       // TMEM lane quadrant this warp may read
    const int ch_base = (16 * qd) % p.Nout;
    const int n_share = 2 * (64 / p.Nout);
    const int share_idx = ((16 * qd) / p.Nout) * 2 + (warp >> 2);
    const int cpt = p.Nout >> 2;                   // float4 chunks per pixel: 4, 8 or 16
    const int ck = tid & (cpt - 1);                // this thread's chunk, fixed for the whole kernel
    const int c4 = ck * 4;
    const int pstep = T_EPI_THREADS / cpt;         // pixels covered per item round
    const int px0 = tid / cpt;
    const bool do_stats = p.stats != nullptr;
    const bool bnb = p.bnb_x != nullptr;
    const bool has_res = p.residual != nullptr;
    float4 bias4 = make_float4(0.f, 0.f, 0.f, 0.f), bsc = bias4, bsh = bias4;
    if (p.bias != nullptr) bias4 = make_float4(__ldg(p.bias + c4), __ldg(p.bias + c4 + 1), __ldg(p.bias + c4 + 2), __ldg(p.bias + c4 + 3));
    if (bnb) {
      bsc = __ldg(reinterpret_cast<const float4*>(p.bnb_scale + c4));
      bsh = __ldg(reinterpret_cast<const float4*>(p.bnb_shift + c4));
    }
    float4 s1 = make_float4(0.f, 0.f, 0.f, 0.f), s2 = s1;
    const int hpwp = cfg.Hp * cfg.Wp;
    const int n_chunks = TN >> 4;
    int local = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
      const int acc = local & 1;
      const uint32_t acc_phase = (uint32_t)(local >> 1) & 1u;
      int* mt = mtab + acc * 256;
      float* patch = stag + acc * cfg.patch_floats;
      if (tid < TN) {
        const int q = tile * TN + tid;
        const int n = (int)fdiv((uint32_t)q, cfg.d_hpwp);
        const int rem = q - n * hpwp;
        const int y = (int)fdiv((uint32_t)rem, cfg.d_wp);
        const int x = rem - y * cfg.Wp;
        mt[tid] = (q < cfg.Q && n < p.N && y < p.OH && x < p.OW) ? (n * p.OH + y) * p.OW + x : -1;
      }
      named_bar_sync(1, T_EPI_THREADS);
      int mo[T_MAX_ITEMS];
      float4 rres[T_MAX_ITEMS], rx[T_MAX_ITEMS];
#pragma unroll
      for (int it = 0; it < T_MAX_ITEMS; ++it) {
        const int px = px0 + it * pstep;
        mo[it] = px < TN ? mt[px] : -1;
        rres[it] = make_float4(0.f, 0.f, 0.f, 0.f);
        rx[it] = rres[it];
        if (mo[it] >= 0) {
          if (has_res) rres[it] = __ldg(reinterpret_cast<const float4*>(p.residual + (size_t)mo[it] * p.res_ld + c4));
          if (bnb) rx[it] = __ldg(reinterpret_cast<const float4*>(p.bnb_x + (size_t)mo[it] * p.bnb_ld + c4));
        }
      }
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      if (tid == 0 && local < 128) ttrace(512 + 2 * local);
      {
        const uint32_t t_base = tmem_base + ((uint32_t)(qd * 32) << 16) + (uint32_t)(acc * TN);
        for (int chunk = share_idx; chunk < n_chunks; chunk += n_share) {
          float v[16];
          tmem_ld16(t_base + (uint32_t)(chunk * 16), v);
          if (lane < 16) {
            float* dst = patch + (chunk * 16) * p.Nout + ch_base + lane;
#pragma unroll
            for (int px = 0; px < 16; ++px) dst[px * p.Nout] = v[px];
          }
        }
      }
      tc_fence_before();
      named_bar_sync(2, T_EPI_THREADS);            // patch complete, every tcgen05.ld of this tile has retired
      mbar_arrive(&tempty_bar[acc]);
      if (tid == 0 && local < 128) ttrace(3072 + 4 * local);
#pragma unroll
      for (int it = 0; it < T_MAX_ITEMS; ++it) {
        if (mo[it] >= 0) {
          const int px = px0 + it * pstep;
          float4 x = *reinterpret_cast<const float4*>(patch + px * p.Nout + c4);
          x.x += bias4.x + rres[it].x; x.y += bias4.y + rres[it].y;
          x.z += bias4.z + rres[it].z; x.w += bias4.w + rres[it].w;
          if (p.out_relu) {
            x.x = fmaxf(x.x, 0.f); x.y = fmaxf(x.y, 0.f); x.z = fmaxf(x.z, 0.f); x.w = fmaxf(x.w, 0.f);
          }
          *reinterpret_cast<float4*>(p.out + (size_t)mo[it] * p.out_ld + c4) = x;
          if (do_stats) {
            if (bnb) {
              const float4 xv = rx[it];
              const float gx = fmaf(xv.x, bsc.x, bsh.x) > 0.f ? x.x : 0.f;
              const float gy = fmaf(xv.y, bsc.y, bsh.y) > 0.f ? x.y : 0.f;
              const float gz = fmaf(xv.z, bsc.z, bsh.z) > 0.f ? x.z : 0.f;
              const float gw = fmaf(xv.w, bsc.w, bsh.w) > 0.f ? x.w : 0.f;
              s1.x += gx; s1.y += gy; s1.z += gz; s1.w += gw;
              s2.x = fmaf(gx, xv.x, s2.x); s2.y = fmaf(gy, xv.y, s2.y);
              s2.z = fmaf(gz, xv.z, s2.z); s2.w = fmaf(gw, xv.w, s2.w);
            } else {
              s1.x += x.x; s1.y += x.y; s1.z += x.z; s1.w += x.w;
              s2.x = fmaf(x.x, x.x, s2.x); s2.y = fmaf(x.y, x.y, s2.y);
              s2.z = fmaf(x.z, x.z, s2.z); s2.w = fmaf(x.w, x.w, s2.w);
            }
          }
        }
      }
      if (tid == 0 && local < 128) ttrace(513 + 2 * local);
    }
    if (do_stats) {
      // lanes whose ids agree modulo cpt (cpt <= 16 divides 32) hold partial sums of the same 4 channels
      for (int d = cpt; d < 32; d <<= 1) {
        s1.x += __shfl_xor_sync(0xffffffffu, s1.x, d); s1.y += __shfl_xor_sync(0xffffffffu, s1.y, d);
        s1.z += __shfl_xor_sync(0xffffffffu, s1.z, d); s1.w += __shfl_xor_sync(0xffffffffu, s1.w, d);
        s2.x += __shfl_xor_sync(0xffffffffu, s2.x, d); s2.y += __shfl_xor_sync(0xffffffffu, s2.y, d);
        s2.z += __shfl_xor_sync(0xffffffffu, s2.z, d); s2.w += __shfl_xor_sync(0xffffffffu, s2.w, d);
      }
      if (lane < cpt) {
        atomicAdd(&ssum[c4 + 0], s1.x); atomicAdd(&ssum[c4 + 1], s1.y);
        atomicAdd(&ssum[c4 + 2], s1.z); atomicAdd(&ssum[c4 + 3], s1.w);
        atomicAdd(&ssum[128 + c4 + 0], s2.x); atomicAdd(&ssum[128 + c4 + 1], s2.y);
        atomicAdd(&ssum[128 + c4 + 2], s2.z); atomicAdd(&ssum[128 + c4 + 3], s2.w);
      }
      named_bar_sync(1, T_EPI_THREADS);
      if (tid < p.Nout) {
        atomicAdd(p.stats + tid, ssum[tid]);
        atomicAdd(p.stats + p.Nout + tid, ssum[128 + tid]);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (tid == 0) ttrace(5);
  if (warp == T_MMA_WARP) {
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)cfg.tmem_cols);
  }
}

static int g_wint_mode = 1;        // 0 off, 1 on

int try_conv_wint(const dtf_conv_gemm_params& p, cudaStream_t stream, bool* launched, bool dry_run) {
  *launched = false;
  if (g_wint_mode == 0) return 0;
  if (p.a_h != 1 || p.a_w != 1 || p.div_h != 1 || p.div_w != 1) return 0;
  if (p.R * p.S < 2) return 0;
  if (p.R > 1 && p.c_h != 1 && p.c_h != -1) return 0;
  if (p.S > 1 && p.c_w != 1 && p.c_w != -1) return 0;
  const int C = p.srcC;
  if (C != 16 && C != 32 && C != 64) return 0;
  const int CB = C;
  const int UM = 64;
  if (p.Nout != 16 && p.Nout != 32 && p.Nout != 64) return 0;      // chunks per pixel must divide the thread count
  if (p.Nout > UM) return 0;
  if ((reinterpret_cast<uintptr_t>(p.src) & 15) != 0 || p.src_ld % 4 != 0) return 0;
  if ((reinterpret_cast<uintptr_t>(p.out) & 15) != 0 || p.out_ld % 4 != 0) return 0;
  if (p.residual && ((reinterpret_cast<uintptr_t>(p.residual) & 15) != 0 || p.res_ld % 4 != 0)) return 0;
  if (p.bnb_x && ((reinterpret_cast<uintptr_t>(p.bnb_x) & 15) != 0 || p.bnb_ld % 4 != 0)) return 0;

  WinTCfg cfg;
  cfg.UM = UM;
  cfg.lane_mode = 1;
  const int eh = (p.R - 1) * p.c_h, ew = (p.S - 1) * p.c_w;
  const int min_h = p.b_h + (eh < 0 ? eh : 0), max_h = p.b_h + (eh > 0 ? eh : 0);
  const int min_w = p.b_w + (ew < 0 ? ew : 0), max_w = p.b_w + (ew > 0 ? ew : 0);
  auto pos = [](int v) { return v > 0 ? v : 0; };
  const int over_h = pos(p.OH - 1 + max_h - (p.srcH - 1)), over_w = pos(p.OW - 1 + max_w - (p.srcW - 1));
  const int gap_h = pos(-min_h) > over_h ? pos(-min_h) : over_h;
  const int gap_w = pos(-min_w) > over_w ? pos(-min_w) : over_w;
  cfg.Hp = (p.srcH > p.OH ? p.srcH : p.OH) + gap_h;
  cfg.Wp = (p.srcW > p.OW ? p.srcW : p.OW) + gap_w;
  const long long Q = (long long)p.N * cfg.Hp * cfg.Wp;
  const long long M = (long long)p.N * p.OH * p.OW;
  if (M == 0) return 0;
  if (Q > (1LL << 30) || Q * 10 > M * 16) return 0;
  if (32 / cfg.Wp + 1 > cfg.Hp) return 0;             // the fill's incremental frame walk wraps at most one row per step
  cfg.Q = (int)Q;
  // tile width: one wave over the SMs when the problem is small, 256 pixels otherwise; at most T_MAX_ITEMS items
  // per epilogue thread; then shrink until two accumulator patches, two windows and the weights fit
  cfg.lin_min = min_h * cfg.Wp + min_w;
  const int lin_max = max_h * cfg.Wp + max_w;
  cfg.CB = CB;
  cfg.cblocks = 1;
  cfg.n_k8 = p.R * p.S * CB / 8;
  if (cfg.n_k8 > T_MAX_K8) return 0;
  cfg.n_kb = ceil_div(cfg.n_k8, 4);
  cfg.rows_b = p.Nout;
  const int b_stage_bytes = 64 * 128;
  const int fixed = 1024 + (2 * T_MAX_BSTAGES + 2 * T_MAX_WIN + 4) * 8 + 16 + T_MAX_K8 * 4 + 512 * 4 + 256 * 4 + 128 * 4;
  cfg.b_resident = (cfg.n_kb <= T_MAX_BSTAGES && cfg.n_kb * b_stage_bytes <= 80 * 1024) ? 1 : 0;
  int tn = round_up(ceil_div((int)Q, num_sms()), 16);
  if (tn > 256) tn = 256;
  const int item_cap = T_MAX_ITEMS * T_EPI_THREADS / (p.Nout / 4);
  if (tn > item_cap) tn = item_cap / 16 * 16;
  if (tn < 32) tn = 32;
  int win_bytes = 0;
  for (;; tn -= 16) {
    if (tn < 32) return 0;
    cfg.win = tn + lin_max - cfg.lin_min;
    cfg.win_pad = ceil_div(cfg.win - 2, 8) * 8 + 2;
    win_bytes = cfg.win_pad * CB * 4;
    const int avail = SMEM_BUDGET - fixed - 2 * tn * p.Nout * 4;
    if (cfg.b_resident) {
      cfg.b_stages = cfg.n_kb;
      const int nw = (avail - cfg.b_stages * b_stage_bytes) / win_bytes;
      if (nw < 2) continue;
      cfg.n_win = nw > T_MAX_WIN ? T_MAX_WIN : nw;
    } else {
      const int bs = (avail - 2 * win_bytes) / b_stage_bytes;
      if (bs < 4) continue;
      cfg.b_stages = bs > 8 ? 8 : bs;
      const int nw = (avail - cfg.b_stages * b_stage_bytes) / win_bytes;
      cfg.n_win = nw > 3 ? 3 : nw;
    }
    break;
  }
  cfg.TN = tn;
  cfg.patch_floats = tn * p.Nout;
  cfg.m_tiles = ceil_div(cfg.Q, tn);
  int cols = 32;
  while (cols < 2 * tn) cols <<= 1;
  cfg.tmem_cols = cols;
  cfg.d_hpwp = make_fastdiv((uint32_t)(cfg.Hp * cfg.Wp));
  cfg.d_wp = make_fastdiv((uint32_t)cfg.Wp);
  cfg.d_k8pt = make_fastdiv((uint32_t)(CB / 8));
  cfg.d_s = make_fastdiv((uint32_t)p.S);
  cfg.d_kbpt = make_fastdiv((uint32_t)(CB >= 32 ? CB / 32 : 1));
  cfg.trace = g_trace_host;

  if (dry_run) {
    *launched = true;
    return 0;
  }
  const int Ktot = p.R * p.S * C;
  const CUtensorMap* tm = get_tmap_2d_f32_sw128(p.wt, p.Nout, Ktot, cfg.rows_b);
  if (!tm) return -2;
  const size_t smem = (size_t)fixed + (size_t)cfg.b_stages * b_stage_bytes + (size_t)cfg.n_win * win_bytes +
                      (size_t)2 * cfg.patch_floats * 4;
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    DTF_CHECK(cudaFuncSetAttribute(conv_wint_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set[dev & 63] = true;
  }
  const int grid = cfg.m_tiles < num_sms() ? cfg.m_tiles : num_sms();
  DTF_CHECK(launch_pdl(conv_wint_kernel, dim3(grid), dim3(T_THREADS), smem, stream, *tm, p, cfg));
  *launched = true;
  return 0;
}

}  // namespace dtf

extern "C" int dtf_debug_set_conv_wint(int mode, int umma_m, int lane_mode) {
  dtf::g_wint_mode = mode;
  (void)umma_m; (void)lane_mode;      // fixed: M = 64, rows 16g..16g+15 in lanes 0..15 of quadrant g (verified on B200)
  return 0;
}

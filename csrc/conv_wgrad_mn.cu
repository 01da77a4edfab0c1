// K1 (weight gradient), copy-only producers: both operands stay pixel-major in shared memory.
//
//   D[k, n] = sum over pixels m of  A[m, k] * dY[m, n]       (k = filter tap x input channel, n = output channel)
//
// The reduction runs over pixels, but activations and gradients are NHWC: a pixel's channels are contiguous, its
// neighbours in the reduction are not.  conv_wgrad.cu fixes that with 4x4 register transposes in the producers
// (K-major operands) and is bound by exactly that producer work.  tcgen05 can also take an operand whose M / N
// index is the contiguous one ("MN-major"); for tf32 that requires the 128-byte swizzle with 32-byte atoms
// (layout type 1): rows of 128 bytes = 32 consecutive k (or n) values of ONE pixel, four pixel rows per 512-byte
// atom, 32-byte chunk index XOR (pixel & 3).  An im2col row of a pixel is exactly such a row, so the producers
// become pure 16-byte cp.async copies (zero-filled for padding) — the same gather the forward kernel uses, no
// registers, no transposes.  When the producer layer's batch-norm + ReLU is fused in, padding is copied from a NaN
// marker instead and the (otherwise idle) epilogue warps rewrite each landed tile in place as
// relu(scale * x + shift), NaN -> 0, before handing it to the tensor core.
//
//   warps 0-3   fix-up (fused input transform only) and epilogue: accumulator row = k; 32 lanes add 32
//                          consecutive k of one output channel (coalesced red.global.add.f32 into dW[n][k])
//   warps 4-11  producers: 4 groups x 64 threads, group g owns every 4th 32-pixel block
//   warp  12    MMA      : 4 x tcgen05.mma.kind::tf32 (M=128, N=C_out tile, K=8 pixels) per block, both operands
//                          MN-major
//
// Same split-over-pixels / atomic-combine scheme and the same C ABI as conv_wgrad.cu, which remains the path for
// shapes this one does not take (C_out not a multiple of 4, unaligned gradients).
#include "../include/dtf_b200.h"
#include "host_common.h"
#include "umma.cuh"

namespace dtf {

constexpr int MN_BM = 128;            // k rows per tile (4 atoms of 32)
constexpr int MN_BKP = 32;            // pixels per stage
constexpr int MN_ATOM_BYTES = MN_BKP * 128;        // one 32-wide MN atom column: 32 pixel rows x 128 B
constexpr int MN_A_STAGE_BYTES = 4 * MN_ATOM_BYTES;
constexpr int MN_PRODUCERS = 256;
constexpr int MN_MMA_WARP = (128 + MN_PRODUCERS) / 32;
constexpr int MN_FIX_WORKERS = 8;     // fix-up warps: the 4 epilogue warps + 4 dedicated ones after the MMA warp
constexpr int MN_THREADS = 128 + MN_PRODUCERS + 32 + 128;

struct WgradMnCfg {
  int BN, n_atoms;       // output-channel tile (multiple of 16, <= 256) and its 32-wide atoms
  int stages, m_tiles, n_tiles, splits, kb_total, kb_per_split, tmem_cols, M, Ktot;
  int active_groups;
  FastDiv d_ohow, d_ow, d_c, d_s, d_dyc;
};

// MN-major operand, 128-byte swizzle with 32-byte atoms: 32 MN elements (128 B) per pixel row, 4 rows per atom.
//   LBO = distance between consecutive 32-element MN groups, SBO = distance between consecutive 4-row K groups.
__device__ __forceinline__ uint64_t make_desc_mn_sw128_32b(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(1) << 61;       // SWIZZLE_128B_BASE32B
  return d;
}

// Source of "padding" chunks when the tile is fixed up after landing (fused batch-norm apply + ReLU): the fix-up
// maps NaN to 0, so out-of-image taps contribute zeros and are never activated.
__device__ __align__(16) const uint32_t g_nan_chunk[4] = {0x7fc00000u, 0x7fc00000u, 0x7fc00000u, 0x7fc00000u};

// byte offset of 16-byte chunk j (0..7) of pixel row `row` inside an atom column
__device__ __forceinline__ uint32_t mn_chunk_off(int row, int j) {
  return (uint32_t)(row * 128 + ((((j >> 1) ^ (row & 3)) << 5) | ((j & 1) << 4)));
}

__global__ void __launch_bounds__(MN_THREADS, 1)
conv_wgrad_mn_kernel(const dtf_conv_wgrad_params p, const WgradMnCfg cfg) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);

  const int BN = cfg.BN;
  const int stages = cfg.stages;
  const int b_stage_bytes = cfg.n_atoms * MN_ATOM_BYTES;
  uint8_t* sA = smem;
  uint8_t* sB = sA + stages * MN_A_STAGE_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(sB + stages * b_stage_bytes);
  uint64_t* empty_bar = full_bar + stages;
  uint64_t* tfull_bar = empty_bar + stages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint64_t* raw_bar = tempty_bar + 2;                       // [stages] copies landed, waiting for the fix-up warps
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(raw_bar + stages);
  int4* ptab = reinterpret_cast<int4*>((reinterpret_cast<uintptr_t>(tmem_slot + 4) + 15) & ~static_cast<uintptr_t>(15));
  // ^ [4 groups][32] {pixel offset, ih0, iw0, valid}

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  pdl_launch_dependents();   // persistent grid, all CTAs resident: the successor may take SMs as they free up

  if (warp == MN_MMA_WARP) {
    if (lane == 0) {
      for (int i = 0; i < stages; ++i) {
        // a stage is complete after its producer group (plain copy) or after the 128 fix-up threads
        mbar_init(&full_bar[i], p.in_scale != nullptr ? 32 : MN_PRODUCERS / 4);
        mbar_init(&empty_bar[i], 1);
        mbar_init(&raw_bar[i], MN_PRODUCERS / 4);
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

  // Fix-up of landed im2col tiles (fused batch-norm apply + ReLU): worker w of MN_FIX_WORKERS warps owns every
  // MN_FIX_WORKERS-th 32-pixel block; lane = 16-byte chunk j (fixed filter tap / channels for the unit), all 32
  // pixel rows.  In place: relu(scale * x + shift), NaN marker -> 0; then the stage is published to the MMA warp.
  auto fixup_unit = [&](int w, int mt, int kb0, int& seq, int& fstage, uint32_t& fphase) {
    int kb1 = kb0 + cfg.kb_per_split;
    if (kb1 > cfg.kb_total) kb1 = cfg.kb_total;
    const int j = lane;
    const int k = mt * MN_BM + j * 4;
    float4 sc = make_float4(0.f, 0.f, 0.f, 0.f), sh = sc;
    if (k < cfg.Ktot) {
      const int tap = (int)fdiv((uint32_t)k, cfg.d_c);
      const int c = k - tap * p.srcC;
      sc = __ldg(reinterpret_cast<const float4*>(p.in_scale + c));
      sh = __ldg(reinterpret_cast<const float4*>(p.in_shift + c));
    }
    const bool relu = p.in_relu != 0;
    const uint32_t col = (uint32_t)((j >> 3) * MN_ATOM_BYTES);
    for (int kb = kb0; kb < kb1; ++kb) {
      if (seq == w) {
        mbar_wait(&raw_bar[fstage], fphase);
        uint8_t* a_stage = sA + fstage * MN_A_STAGE_BYTES + col;
#pragma unroll 8
        for (int row = 0; row < MN_BKP; ++row) {
          float4* ptr = reinterpret_cast<float4*>(a_stage + mn_chunk_off(row, j & 7));
          const float4 x = *ptr;
          float4 y;
          y.x = fmaf(x.x, sc.x, sh.x); y.y = fmaf(x.y, sc.y, sh.y);
          y.z = fmaf(x.z, sc.z, sh.z); y.w = fmaf(x.w, sc.w, sh.w);
          if (relu) { y.x = fmaxf(y.x, 0.f); y.y = fmaxf(y.y, 0.f); y.z = fmaxf(y.z, 0.f); y.w = fmaxf(y.w, 0.f); }
          y.x = (x.x == x.x) ? y.x : 0.f; y.y = (x.y == x.y) ? y.y : 0.f;
          y.z = (x.z == x.z) ? y.z : 0.f; y.w = (x.w == x.w) ? y.w : 0.f;
          *ptr = y;
        }
        fence_proxy_async_smem();
        mbar_arrive(&full_bar[fstage]);
      }
      if (++seq == MN_FIX_WORKERS) seq = 0;
      if (++fstage == stages) { fstage = 0; fphase ^= 1u; }
    }
  };

  if (warp >= 4 && warp < MN_MMA_WARP) {
    // =========================== producers =================================================
    // Thread (j = gt & 31, half = gt >> 5) of a group copies 16-byte chunk j of the 128 k values (one filter tap,
    // four input channels: fixed for the whole unit) for the 16 pixels half, half + 2, ...; the pixel coordinates
    // of the block come from a 32-entry table the group's first warp writes per block.
    const int pt = tid - 128;
    const int grp = pt >> 6;
    const int gt = pt & 63;
    const int j = gt & 31;
    const int half = gt >> 5;
    const bool has_aff = p.in_scale != nullptr;
    int4* gtab = ptab + grp * MN_BKP;
    const int dyc = BN >> 2;                  // 16-byte chunks per pixel row of the dY tile
    int q_grp = 0, stage = 0; uint32_t phase = 0;

    for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x) {
      const int split = unit % cfg.splits;
      const int tile = unit / cfg.splits;
      const int mt = tile / cfg.n_tiles, nt = tile - mt * cfg.n_tiles;
      const int kb0 = split * cfg.kb_per_split;
      int kb1 = kb0 + cfg.kb_per_split;
      if (kb1 > cfg.kb_total) kb1 = cfg.kb_total;
      // this thread's k: tap (r, s), channel c
      const int k = mt * MN_BM + j * 4;
      const bool k_ok = k < cfg.Ktot;
      const int tap = k_ok ? (int)fdiv((uint32_t)k, cfg.d_c) : 0;
      const int c = k_ok ? k - tap * p.srcC : 0;
      const int r = (int)fdiv((uint32_t)tap, cfg.d_s);
      const int s_ = tap - r * p.S;
      const int tap_off = (r * p.srcW + s_) * p.src_ld + c;
      const int n0 = nt * BN;
      const uint32_t a_col = (uint32_t)((j >> 3) * MN_ATOM_BYTES);       // atom column of this thread's chunk
      const int jj = j & 7;
      for (int kb = kb0; kb < kb1; ++kb) {
        if (q_grp == grp) {
          // ---- pixel table of this block (first warp of the group), visible to both warps ----
          named_bar_sync(2 + grp, 64);                    // previous block's table no longer in use
          if (half == 0) {
            const int m = kb * MN_BKP + j;
            int4 e = make_int4(0, -(1 << 24), 0, 0);
            if (m < cfg.M) {
              const int img = (int)fdiv((uint32_t)m, cfg.d_ohow);
              const int rem = m - img * p.OH * p.OW;
              const int oh = (int)fdiv((uint32_t)rem, cfg.d_ow);
              const int ow = rem - oh * p.OW;
              const int ih0 = oh * p.stride_h - p.pad_h, iw0 = ow * p.stride_w - p.pad_w;
              e = make_int4(((img * p.srcH + ih0) * p.srcW + iw0) * p.src_ld, ih0, iw0, 1);
            }
            gtab[j] = e;
          }
          named_bar_sync(2 + grp, 64);
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          uint8_t* a_stage = sA + stage * MN_A_STAGE_BYTES;
          uint8_t* b_stage = sB + stage * b_stage_bytes;
          const uint32_t a_base = smem_u32(a_stage) + a_col;
          // ---- im2col chunks: 16 pixels per thread.  Phase 1 (table reads + ALU, freely schedulable) collects the
          // source offsets; phase 2 issues the 16 LDGSTS back to back (each asm statement is a compiler barrier, so
          // interleaving the table reads with the copies would serialise them on the shared-memory latency) ----
          int soff[16];
          uint32_t okmask = 0;
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const int4 e = gtab[half + 2 * i];
            const bool valid = k_ok && (unsigned)(e.y + r) < (unsigned)p.srcH && (unsigned)(e.z + s_) < (unsigned)p.srcW;
            if (valid) okmask |= 1u << i;
            soff[i] = e.x + tap_off;
          }
          // dY chunks: this thread's (row, chunk) pairs, offsets first
          const int n_dy = MN_BKP * dyc;
          int dyoff[4];
          uint32_t dymask = 0;
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            const int idx = gt + 64 * t;
            const int row = (int)fdiv((uint32_t)idx, cfg.d_dyc), cj = idx - row * dyc;
            const int m = kb * MN_BKP + row;
            const bool valid = idx < n_dy && m < cfg.M && n0 + cj * 4 < p.Nout;
            if (valid) dymask |= 1u << t;
            dyoff[t] = valid ? m * p.dy_ld + n0 + cj * 4 : 0;
          }
          // columns past the end of the filter (tail tile) are not copied at all: their accumulator rows are
          // never stored, so whatever the stage holds there is harmless
          if (k_ok) {
            if (has_aff) {
#pragma unroll
              for (int i = 0; i < 16; ++i)
                cp_async_16(a_base + mn_chunk_off(half + 2 * i, jj),
                            ((okmask >> i) & 1u) ? (const void*)(p.src + soff[i]) : (const void*)g_nan_chunk, 16u);
            } else {
#pragma unroll
              for (int i = 0; i < 16; ++i)
                cp_async_16(a_base + mn_chunk_off(half + 2 * i, jj), p.src + (((okmask >> i) & 1u) ? soff[i] : 0),
                            ((okmask >> i) & 1u) ? 16u : 0u);
            }
          }
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            const int idx = gt + 64 * t;
            if (idx < n_dy) {
              const int row = (int)fdiv((uint32_t)idx, cfg.d_dyc), cj = idx - row * dyc;
              cp_async_16(smem_u32(b_stage) + (uint32_t)((cj >> 3) * MN_ATOM_BYTES) + mn_chunk_off(row, cj & 7),
                          p.dy + dyoff[t], ((dymask >> t) & 1u) ? 16u : 0u);
            }
          }
          // wide gradient tiles (more than 8 chunks per pixel row): the rest, one at a time
          for (int idx = gt + 256; idx < n_dy; idx += 64) {
            const int row = (int)fdiv((uint32_t)idx, cfg.d_dyc), cj = idx - row * dyc;
            const int m = kb * MN_BKP + row;
            const int co = n0 + cj * 4;
            const bool valid = m < cfg.M && co < p.Nout;
            cp_async_16(smem_u32(b_stage) + (uint32_t)((cj >> 3) * MN_ATOM_BYTES) + mn_chunk_off(row, cj & 7),
                        p.dy + (valid ? (size_t)m * p.dy_ld + co : 0), valid ? 16u : 0u);
          }
          // the barrier hears from this thread once its copies have landed: the stage itself when the data is
          // used as is, the fix-up warps' barrier when batch-norm + ReLU is applied first
          cp_async_mbar_arrive_noinc(has_aff ? &raw_bar[stage] : &full_bar[stage]);
        }
        if (++q_grp == cfg.active_groups) q_grp = 0;
        if (++stage == stages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == MN_MMA_WARP) {
    // =========================== MMA issuer ================================================
    // both operands MN-major: bit 15 (A) and bit 16 (B) of the instruction descriptor
    const uint32_t idesc = make_idesc_tf32(MN_BM, (uint32_t)BN) | (1u << 15) | (1u << 16);
    // a last tile with at most 64 valid k rows runs as M = 64: half the A-operand stream per instruction
    const uint32_t idesc64 = make_idesc_tf32(64, (uint32_t)BN) | (1u << 15) | (1u << 16);
    const uint64_t adesc0 = make_desc_mn_sw128_32b(smem_u32(sA), MN_ATOM_BYTES, 512);
    const uint64_t bdesc0 = make_desc_mn_sw128_32b(smem_u32(sB), MN_ATOM_BYTES, 512);
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
      const bool use64 = cfg.Ktot - ((unit / cfg.splits) / cfg.n_tiles) * MN_BM <= 64;
      uint32_t accumulate = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        {
          const uint64_t adesc = adesc0 + (uint64_t)((uint32_t)(stage * MN_A_STAGE_BYTES) >> 4);
          const uint64_t bdesc = bdesc0 + (uint64_t)((uint32_t)(stage * b_stage_bytes) >> 4);
          if (elect_one()) {
#pragma unroll
            for (int kk = 0; kk < MN_BKP / 8; ++kk) {
              // 8 pixels further along the reduction = 8 rows of 128 bytes
              umma_tf32_ss(d_tmem, adesc + (uint64_t)(kk * 64), bdesc + (uint64_t)(kk * 64), use64 ? idesc64 : idesc,
                           accumulate);
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
  } else if (warp > MN_MMA_WARP) {
    // =========================== dedicated fix-up warps =====================================
    if (p.in_scale != nullptr) {
      int fstage = 0, fseq = 0; uint32_t fphase = 0;
      for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x) {
        const int split = unit % cfg.splits;
        const int tile = unit / cfg.splits;
        const int mt = tile / cfg.n_tiles;
        fixup_unit(4 + (warp - MN_MMA_WARP - 1), mt, split * cfg.kb_per_split, fseq, fstage, fphase);
      }
    }
  } else {
    // =========================== epilogue ==================================================
    int local = 0;
    const bool has_aff = p.in_scale != nullptr;
    int fstage = 0, fseq = 0; uint32_t fphase = 0;
    for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x, ++local) {
      const int split = unit % cfg.splits;
      const int tile = unit / cfg.splits;
      const int mt = tile / cfg.n_tiles, nt = tile - mt * cfg.n_tiles;
      const int kb0 = split * cfg.kb_per_split;
      const bool nonempty = kb0 < cfg.kb_total;
      if (has_aff) fixup_unit(warp, mt, kb0, fseq, fstage, fphase);
      const int acc = local & 1;
      const uint32_t acc_phase = (uint32_t)(local >> 1) & 1u;
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      // accumulator row of this thread: M = 128 -> lane r holds row r; M = 64 (short last tile) -> rows
      // 16g..16g+15 sit in lanes 0..15 of quadrant g
      const bool use64 = cfg.Ktot - mt * MN_BM <= 64;
      const int krow = use64 ? (lane < 16 ? warp * 16 + lane : MN_BM) : tid;
      const int k = mt * MN_BM + krow;
      const bool kvalid = nonempty && krow < MN_BM && k < cfg.Ktot;
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
  if (warp == MN_MMA_WARP) {
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)cfg.tmem_cols);
  }
}

static int g_wgrad_mn_enabled = 1;

int try_conv_wgrad_mn(const dtf_conv_wgrad_params& p, cudaStream_t stream, bool* launched) {
  *launched = false;
  if (!g_wgrad_mn_enabled) return 0;
  if (p.Nout % 4 != 0 || p.dy_ld % 4 != 0 || (reinterpret_cast<uintptr_t>(p.dy) & 15) != 0) return 0;
  WgradMnCfg cfg;
  cfg.Ktot = p.R * p.S * p.srcC;
  cfg.M = p.N * p.OH * p.OW;
  if (cfg.M == 0) return 0;
  cfg.BN = p.Nout <= 256 ? round_up(p.Nout, 16) : 256;
  cfg.n_atoms = ceil_div(cfg.BN, 32);
  cfg.n_tiles = ceil_div(p.Nout, cfg.BN);
  cfg.m_tiles = ceil_div(cfg.Ktot, MN_BM);
  cfg.kb_total = ceil_div(cfg.M, MN_BKP);
  const int base_units = cfg.m_tiles * cfg.n_tiles;
  int splits = ceil_div(2 * num_sms(), base_units);
  const int max_splits = cfg.kb_total / 4 > 0 ? cfg.kb_total / 4 : 1;
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  cfg.kb_per_split = ceil_div(cfg.kb_total, splits);
  cfg.splits = ceil_div(cfg.kb_total, cfg.kb_per_split);
  int tc = 32;
  while (tc < 2 * cfg.BN) tc <<= 1;
  cfg.tmem_cols = tc;
  const int stage_bytes = MN_A_STAGE_BYTES + cfg.n_atoms * MN_ATOM_BYTES;
  const int fixed = 1024 + (3 * 8 + 4) * 8 + 32 + 4 * MN_BKP * 16;
  int stages = (SMEM_BUDGET - fixed) / stage_bytes;
  if (stages > 8) stages = 8;
  if (stages < 2) return 0;
  cfg.stages = stages;
  cfg.active_groups = stages < 4 ? stages : 4;
  cfg.d_ohow = make_fastdiv((uint32_t)(p.OH * p.OW));
  cfg.d_ow = make_fastdiv((uint32_t)p.OW);
  cfg.d_c = make_fastdiv((uint32_t)p.srcC);
  cfg.d_s = make_fastdiv((uint32_t)p.S);
  cfg.d_dyc = make_fastdiv((uint32_t)(cfg.BN / 4));
  if ((long long)cfg.M * p.dy_ld >= (1LL << 31)) return 0;      // 32-bit gradient offsets

  const size_t smem = (size_t)fixed + (size_t)stages * stage_bytes;
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    DTF_CHECK(cudaFuncSetAttribute(conv_wgrad_mn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set[dev & 63] = true;
  }
  const int units = base_units * cfg.splits;
  const int grid = units < num_sms() ? units : num_sms();
  DTF_CHECK(launch_pdl(conv_wgrad_mn_kernel, dim3(grid), dim3(MN_THREADS), smem, stream, p, cfg));
  *launched = true;
  return 0;
}

}  // namespace dtf

extern "C" int dtf_debug_set_wgrad_mn(int enabled) {
  dtf::g_wgrad_mn_enabled = enabled;
  return 0;
}

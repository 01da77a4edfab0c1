// K16 + C1/C2/C3/C4: parameter-server update kernels.
//
//  dtf_fused_update   one rank: w/slots/EMA update of a flat range from a local gradient (also the N=1 path)
//  dtf_ps_bucket_step BSP push of ONE bucket (a contiguous range of the flat buffer = a group of layers), launched
//                     on a side stream as soon as the backward pass has produced the bucket's gradients:
//                       signal "my gradients of bucket b are complete" to all ranks, wait for everybody's
//                       push : the owner of each 1/W of the bucket forms the mean of the W gradient copies
//                              (multimem.ld_reduce in the NVSwitch, or P2P ld.global), applies SGD /
//                              TF1-RMSProp(+momentum) / weight decay / EMA in registers  [reduce-scatter + K16]
//                       pull : and stores the new values into EVERY rank's replica (multimem.st, or P2P st) [all-gather]
//                       signal "my part of bucket b is updated" to all ranks — and exit; nobody waits here
//  dtf_ps_gate        the other half of the pull: wait until every owner's part of buckets [b0, b1) carries this
//                     step's epoch.  The convolution kernels run the same wait in their weight-load warp
//                     (dtf_conv_gemm_params.gate_*), so the next forward pass starts while later buckets of the
//                     previous step are still arriving.
//  dtf_ps_async_step  ASP/SSP step: the *pusher* applies its own gradient to every owner's master range
//                       (red.global.add.v4.f32 for SGD — Hogwild, the reference runs use_locking=False), bumps
//                       the global step on rank 0 with a system-scope atomic, then refreshes its replica with
//                       128-bit loads; no barrier.  RMSProp: the scaled gradient is handed to the owner's inbox
//                       (one-way red.add) and the owner applies it to its slots locally.
//  dtf_ssp_clock      publish clock[rank] = step in every peer's clock table, spin (bounded) while
//                       step > min(clock) + bound                          [replaces the Thrift RPC loop]
//
// All cross-rank synchronisation is st.release.sys / ld.acquire.sys on words in peer-mapped memory; spins are
// bounded and report through a status word instead of hanging the GPU.  Hyper-parameters are read from device
// memory so a captured CUDA graph picks up learning-rate changes without re-capture.
//
// Reference behaviour replaced: SyncReplicasOptimizer accumulators + chief queue runner + token queue
// (cifarnet/cifar10_cifarnet_bsp.py:75-112), the per-step gRPC variable pull, bare apply_gradients for ASP
// (vgg/cifar10.py:368-372), RMSPropOptimizer(decay .9, momentum .9, epsilon 1.0)
// (inception/imagenet_inception_bsp.py:134-136), ExponentialMovingAverage (:123-126) and
// SspHandler.check_staleness (inception/ssp_manager.py:21-29).
#include "../include/dtf_b200.h"
#include "host_common.h"
#include "umma.cuh"

namespace dtf {

// hyper[]: 0 lr, 1 grad_scale, 2 weight_decay, 3 rms_decay, 4 momentum, 5 epsilon, 6 ema_decay (<0: no EMA)
struct Hyper {
  float lr, gscale, wd, rms_decay, momentum, eps, ema_decay;
};
__device__ __forceinline__ Hyper load_hyper(const float* h) {
  Hyper r;
  r.lr = h[0]; r.gscale = h[1]; r.wd = h[2]; r.rms_decay = h[3]; r.momentum = h[4]; r.eps = h[5]; r.ema_decay = h[6];
  return r;
}

// one element of the update; kind 0 = SGD, 1 = TF1 RMSProp with momentum
__device__ __forceinline__ float update_elem(int kind, const Hyper& h, float w, float g, bool decayed, float* ms,
                                             float* mom) {
  g *= h.gscale;
  if (decayed) g = fmaf(h.wd, w, g);
  if (kind == 0) return w - h.lr * g;
  const float m2 = h.rms_decay * (*ms) + (1.f - h.rms_decay) * g * g;
  const float mo = h.momentum * (*mom) + h.lr * g * rsqrtf(m2 + h.eps);
  *ms = m2;
  *mom = mo;
  return w - mo;
}

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Spin (one thread) until flags[q] >= epoch for every q < n.  Bounded: false after timeout_cycles SM clocks.
__device__ bool wait_all(const unsigned int* flags, int n, unsigned int epoch, long long timeout_cycles) {
  const long long t0 = clock64();
  for (int q = 0; q < n; ++q) {
    while ((int)(ld_acquire_sys(flags + q) - epoch) < 0) {
      __nanosleep(64);
      if (clock64() - t0 > timeout_cycles) return false;
    }
  }
  return true;
}

__global__ void __launch_bounds__(512)
fused_update_kernel(float* __restrict__ w, const float* __restrict__ g, float* __restrict__ ms,
                    float* __restrict__ mom, float* __restrict__ ema, long long n, long long n_decayed,
                    const float* __restrict__ hyper, int kind) {
  const Hyper h = load_hyper(hyper);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float m2 = 0.f, mo = 0.f;
    if (kind == 1) { m2 = ms[i]; mo = mom[i]; }
    const float nw = update_elem(kind, h, w[i], g[i], i < n_decayed, &m2, &mo);
    w[i] = nw;
    if (kind == 1) { ms[i] = m2; mom[i] = mo; }
    if (ema && h.ema_decay >= 0.f) ema[i] -= (1.f - h.ema_decay) * (ema[i] - nw);
  }
}

// ---------------------------------------------------------------------------------------------------------
// BSP, one bucket.  Flag block of a rank (uint32 words in its symmetric buffer, written by its peers):
//   ready  [b*W + q]            rank q's gradients of bucket b are complete for epoch e
//   updated[(B + b)*W + q]      owner q has written its part of bucket b into THIS rank's replica for epoch e
// Epochs count completed steps (device word *epoch_ptr, bumped by the step's last bucket), so a captured graph
// replays without host-side arguments changing.  Within the bucket [lo, lo+len) rank r owns
// [lo + r*chunk, min(lo + (r+1)*chunk, lo+len)); chunk % 4 == 0.
// ---------------------------------------------------------------------------------------------------------
struct BucketArgs {
  const float* const* peer_grads;   // [W] gradient buffers of every rank (device pointers valid on this GPU)
  float* const* peer_params;        // [W] parameter replicas of every rank
  unsigned int* const* peer_flags;  // [W] flag blocks of every rank
  unsigned int* local_counter;      // block-completion counter of this bucket (local, reset by the last block)
  unsigned int* status;             // 0 = ok, 1 = timeout
  int rank, W;
  int ready_off, updated_off;       // word offsets of this bucket's rows inside a flag block
  long long lo, len, chunk;
  long long wd_boundary;
  float* ms; float* mom; float* ema;   // slots of this rank's part of the bucket (length <= chunk)
  const float* hyper;
  int kind;
  const unsigned int* epoch_ptr;    // completed steps; this launch works for epoch *epoch_ptr + 1
  unsigned int* epoch_out;          // non-null on the step's last bucket: receives the new epoch at the end
  // NVLS (NVSwitch multicast) addresses of the same symmetric buffers, or nullptr: with them the gradient reduction
  // happens inside the switch (multimem.ld_reduce) and the updated range is stored to every replica at once
  // (multimem.st).
  const float* mc_grads;
  float* mc_params;
  long long timeout_cycles;
};

__device__ __forceinline__ float4 multimem_ld_reduce_add_f32x4(const float* mc_addr) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc_addr)
               : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_f32x4(float* mc_addr, const float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ float4 ld_peer_f32x4(const float* p) {
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void red_add_f32x4(float* p, const float4 v) {
  asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}

constexpr int PUSH_UNROLL = 4;      // 16-byte reductions in flight per thread
// Block shape.  A push block cannot share an SM with a persistent convolution CTA (13-14 warps x 120-128 registers
// fill the 16 K registers of the busiest SM sub-partition; measured: every convolution of VGG's backward pass waited
// for the push's last block, profiles/trace_vgg_bsp_n2_before_carveout.txt).  So the push does not try to squeeze
// in: it runs 1024-thread blocks that take a whole register file each, on as many SMs as the host reserves for it
// (dtf_reserve_sms keeps the compute grids off that many), with 64 KB of reductions in flight per SM.
constexpr int PUSH_THREADS = 1024;

template <int KIND, bool USE_MC>
__global__ void __launch_bounds__(PUSH_THREADS, 1)
ps_bucket_kernel(const BucketArgs a) {
  const int W = a.W, me = a.rank;
  const unsigned int epoch = *reinterpret_cast<const volatile unsigned int*>(a.epoch_ptr) + 1u;
  // ---- "my gradients of this bucket are complete" -> every rank; then wait for everybody's -----------------
  if (W > 1) {
    if (blockIdx.x == 0 && threadIdx.x < W) {
      __threadfence_system();
      st_release_sys(a.peer_flags[threadIdx.x] + a.ready_off + me, epoch);
    }
    int timed_out = 0;
    if (threadIdx.x == 0) timed_out = wait_all(a.peer_flags[me] + a.ready_off, W, epoch, a.timeout_cycles) ? 0 : 1;
    if (__syncthreads_or(timed_out)) { if (threadIdx.x == 0) atomicExch(a.status, 1u); return; }   // (no static smem)
  }

  // ---- push: reduce my part over all ranks' gradients + optimiser; pull: store into every replica -----------
  const Hyper h = load_hyper(a.hyper);
  const long long lo = a.lo + (long long)me * a.chunk;
  long long n = a.len - (long long)me * a.chunk;
  n = n < 0 ? 0 : (n > a.chunk ? a.chunk : n);
  const float invW = 1.f / (float)W;
  float* wmine = a.peer_params[me] + lo;
  const long long n4 = n >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long base = (long long)blockIdx.x * blockDim.x + threadIdx.x; base < n4; base += stride * PUSH_UNROLL) {
    float4 g[PUSH_UNROLL];
    // issue all the remote reductions of this round before touching any of them
#pragma unroll
    for (int u = 0; u < PUSH_UNROLL; ++u) {
      const long long i4 = base + u * stride;
      g[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (i4 < n4) {
        if (USE_MC) {
          g[u] = multimem_ld_reduce_add_f32x4(a.mc_grads + lo + (i4 << 2));   // sum over all ranks, formed in the switch
        } else {
          for (int q = 0; q < W; ++q) {
            const int src = (me + q) % W;      // stagger peers so ranks do not all hit the same GPU at once
            const float4 v = ld_peer_f32x4(a.peer_grads[src] + lo + (i4 << 2));
            g[u].x += v.x; g[u].y += v.y; g[u].z += v.z; g[u].w += v.w;
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < PUSH_UNROLL; ++u) {
      const long long i4 = base + u * stride;
      if (i4 >= n4) continue;
      float4 gv = g[u];
      gv.x *= invW; gv.y *= invW; gv.z *= invW; gv.w *= invW;
      float4 wv = *reinterpret_cast<float4*>(wmine + (i4 << 2));
      float4 m2 = make_float4(0.f, 0.f, 0.f, 0.f), mo = m2;
      if (KIND == 1) {
        m2 = *reinterpret_cast<float4*>(a.ms + (i4 << 2));
        mo = *reinterpret_cast<float4*>(a.mom + (i4 << 2));
      }
      const long long gi = lo + (i4 << 2);
      wv.x = update_elem(KIND, h, wv.x, gv.x, gi + 0 < a.wd_boundary, &m2.x, &mo.x);
      wv.y = update_elem(KIND, h, wv.y, gv.y, gi + 1 < a.wd_boundary, &m2.y, &mo.y);
      wv.z = update_elem(KIND, h, wv.z, gv.z, gi + 2 < a.wd_boundary, &m2.z, &mo.z);
      wv.w = update_elem(KIND, h, wv.w, gv.w, gi + 3 < a.wd_boundary, &m2.w, &mo.w);
      if (USE_MC) {
        multimem_st_f32x4(a.mc_params + gi, wv);                 // every replica, mine too
      } else {
        *reinterpret_cast<float4*>(wmine + (i4 << 2)) = wv;
        for (int q = 1; q < W; ++q)                               // P2P stores into the other replicas
          *reinterpret_cast<float4*>(a.peer_params[(me + q) % W] + gi) = wv;
      }
      if (KIND == 1) {
        *reinterpret_cast<float4*>(a.ms + (i4 << 2)) = m2;
        *reinterpret_cast<float4*>(a.mom + (i4 << 2)) = mo;
      }
      if (a.ema && h.ema_decay >= 0.f) {
        float4 e = *reinterpret_cast<float4*>(a.ema + (i4 << 2));
        const float k = 1.f - h.ema_decay;
        e.x -= k * (e.x - wv.x); e.y -= k * (e.y - wv.y); e.z -= k * (e.z - wv.z); e.w -= k * (e.w - wv.w);
        *reinterpret_cast<float4*>(a.ema + (i4 << 2)) = e;
      }
    }
  }

  // ---- the last block to finish announces "my part of the bucket is in your replica" to every rank ---------
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int done = atomicAdd(a.local_counter, 1u);
    if (done == gridDim.x - 1) {
      *a.local_counter = 0;
      __threadfence_system();
      if (W > 1)
        for (int q = 0; q < W; ++q) st_release_sys(a.peer_flags[(me + q) % W] + a.updated_off + me, epoch);
      if (a.epoch_out != nullptr) { *a.epoch_out = epoch; __threadfence(); }
    }
  }
}

// Wait until flags[0 .. n) all carry at least *epoch_ptr (the number of completed steps): "every owner's part of
// these buckets has reached my replica".  One warp; lane q polls the words q, q+32, ...
__global__ void ps_gate_kernel(const unsigned int* flags, int n, const unsigned int* epoch_ptr, unsigned int* status,
                               long long timeout_cycles) {
  const unsigned int epoch = *reinterpret_cast<const volatile unsigned int*>(epoch_ptr);
  const long long t0 = clock64();
  bool ok = true;
  for (int q = threadIdx.x; q < n && ok; q += 32) {
    while ((int)(ld_acquire_sys(flags + q) - epoch) < 0) {
      __nanosleep(32);
      if (clock64() - t0 > timeout_cycles) { ok = false; break; }
    }
  }
  if (!ok) atomicExch(status, 1u);
}

struct AsyncArgs {
  const float* grads;               // local gradient buffer (full length)
  float* const* peer_params;        // [W]
  float* const* peer_ms;            // [W] RMSProp slots of every owner's range (nullptr table for SGD)
  float* const* peer_mom;
  float* const* peer_ema;           // [W] or nullptr
  unsigned long long* global_step;  // on rank 0 (peer-mapped)
  int rank, W;
  long long shard, wd_boundary;
  const float* hyper;
  int kind;
  long long lo, len;                // the part of the flat buffer this launch pushes (a bucket; multiples of 4)
  int bump_step;                    // 1 on the step's first launch: global_step += 1
};

constexpr int ASYNC_UNROLL = 4;

// ASP / SSP step of one worker, 16 bytes per access and ASYNC_UNROLL accesses in flight per thread.  Per element
// the traffic that crosses NVLink is: SGD — one red.add out, one load back; RMSProp — the owner's two slots in
// (one round trip, both loads issued together), their new values and the red.add of the step out.  The refresh of
// the replica rides in the same loop as the push (opposite link directions), so the two phases overlap.
template <int KIND, int THREADS>
__global__ void __launch_bounds__(THREADS, 1)
ps_async_step_kernel(const AsyncArgs a) {
  const Hyper h = load_hyper(a.hyper);
  const int W = a.W, me = a.rank;
  const long long total4 = a.len >> 2;
  // workers start at different places of the range, so that they do not all hit the same owner at once
  const long long rot4 = (total4 / W) * me;
  float* mine = a.peer_params[me];
  if (a.bump_step && blockIdx.x == 0 && threadIdx.x == 0) atomicAdd_system(a.global_step, 1ull);
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long base = (long long)blockIdx.x * blockDim.x + threadIdx.x; base < total4; base += stride * ASYNC_UNROLL) {
    float4 g[ASYNC_UNROLL], m2[ASYNC_UNROLL], mo[ASYNC_UNROLL];
    long long gi[ASYNC_UNROLL];
    int own[ASYNC_UNROLL];
#pragma unroll
    for (int u = 0; u < ASYNC_UNROLL; ++u) {
      const long long j4 = base + u * stride;
      own[u] = -1;
      if (j4 < total4) {
        long long r4 = j4 + rot4;
        if (r4 >= total4) r4 -= total4;
        gi[u] = a.lo + (r4 << 2);
        own[u] = (int)(gi[u] / a.shard);       // flat shards: rank r owns [r*shard, (r+1)*shard)
        const long long i4 = (gi[u] - (long long)own[u] * a.shard) >> 2;
        float4 gv = *reinterpret_cast<const float4*>(a.grads + gi[u]);
        gv.x *= h.gscale; gv.y *= h.gscale; gv.z *= h.gscale; gv.w *= h.gscale;
        // the weight-decay term uses this worker's replica of w (what the gradient was computed from)
        const float4 wl = *reinterpret_cast<const float4*>(mine + gi[u]);
        if (gi[u] + 0 < a.wd_boundary) gv.x = fmaf(h.wd, wl.x, gv.x);
        if (gi[u] + 1 < a.wd_boundary) gv.y = fmaf(h.wd, wl.y, gv.y);
        if (gi[u] + 2 < a.wd_boundary) gv.z = fmaf(h.wd, wl.z, gv.z);
        if (gi[u] + 3 < a.wd_boundary) gv.w = fmaf(h.wd, wl.w, gv.w);
        g[u] = gv;
        if (KIND == 1) {                      // slots live with the owner; racy RMW is the semantics of
          m2[u] = ld_peer_f32x4(a.peer_ms[own[u]] + (i4 << 2));      // use_locking=False
          mo[u] = ld_peer_f32x4(a.peer_mom[own[u]] + (i4 << 2));
        }
      }
    }
#pragma unroll
    for (int u = 0; u < ASYNC_UNROLL; ++u) {
      if (own[u] < 0) break;
      float* wp = a.peer_params[own[u]] + gi[u];
      float4 d;
      if (KIND == 0) {
        d = make_float4(-h.lr * g[u].x, -h.lr * g[u].y, -h.lr * g[u].z, -h.lr * g[u].w);
      } else {
        const long long i = gi[u] - (long long)own[u] * a.shard;
        float4 n2, nm;
        n2.x = h.rms_decay * m2[u].x + (1.f - h.rms_decay) * g[u].x * g[u].x;
        n2.y = h.rms_decay * m2[u].y + (1.f - h.rms_decay) * g[u].y * g[u].y;
        n2.z = h.rms_decay * m2[u].z + (1.f - h.rms_decay) * g[u].z * g[u].z;
        n2.w = h.rms_decay * m2[u].w + (1.f - h.rms_decay) * g[u].w * g[u].w;
        nm.x = h.momentum * mo[u].x + h.lr * g[u].x * rsqrtf(n2.x + h.eps);
        nm.y = h.momentum * mo[u].y + h.lr * g[u].y * rsqrtf(n2.y + h.eps);
        nm.z = h.momentum * mo[u].z + h.lr * g[u].z * rsqrtf(n2.z + h.eps);
        nm.w = h.momentum * mo[u].w + h.lr * g[u].w * rsqrtf(n2.w + h.eps);
        *reinterpret_cast<float4*>(a.peer_ms[own[u]] + i) = n2;
        *reinterpret_cast<float4*>(a.peer_mom[own[u]] + i) = nm;
        d = make_float4(-nm.x, -nm.y, -nm.z, -nm.w);
      }
      red_add_f32x4(wp, d);                     // one-sided red.add on the owner's HBM (Hogwild)
    }
    // refresh my replica of the ranges I just pushed to (no barrier: stale reads are the semantics)
    float4 fresh[ASYNC_UNROLL];
#pragma unroll
    for (int u = 0; u < ASYNC_UNROLL; ++u)
      if (own[u] >= 0 && own[u] != me) fresh[u] = ld_peer_f32x4(a.peer_params[own[u]] + gi[u]);
#pragma unroll
    for (int u = 0; u < ASYNC_UNROLL; ++u)
      if (own[u] >= 0 && own[u] != me) *reinterpret_cast<float4*>(mine + gi[u]) = fresh[u];
  }
  // ---- EMA of my own range, owner side (local traffic only).  The reference applies the average after
  // every worker's step; the owner sees W such steps per step of its own, so the host passes decay^W.
  if (a.peer_ema && h.ema_decay >= 0.f) {
    float* ema = a.peer_ema[me];
    const float* wm = mine + (long long)me * a.shard;
    const float k = 1.f - h.ema_decay;
    // the part of my own shard that lies inside this launch's range
    long long e_lo = a.lo - (long long)me * a.shard, e_hi = a.lo + a.len - (long long)me * a.shard;
    e_lo = e_lo < 0 ? 0 : e_lo;
    e_hi = e_hi > a.shard ? a.shard : e_hi;
    for (long long i4 = (e_lo >> 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i4 < (e_hi >> 2); i4 += stride) {
      float4 e = *reinterpret_cast<float4*>(ema + (i4 << 2));
      const float4 wv = ld_peer_f32x4(wm + (i4 << 2));
      e.x -= k * (e.x - wv.x); e.y -= k * (e.y - wv.y); e.z -= k * (e.z - wv.z); e.w -= k * (e.w - wv.w);
      *reinterpret_cast<float4*>(ema + (i4 << 2)) = e;
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// ASP / SSP push for SGD on the bulk-copy engine (TMA): no atomics, no registers in flight.
//
// What limits the register version above is not NVLink but the SM: a red.add retires about one fp32 element per
// clock per SM (measured: 6.3 GB/s per SM for REDG.128 from threads and, identically, for cp.reduce.async.bulk.add
// from the copy engine; VGG-16 on 2 GPUs 6.9 / 4.6 / 4.0 ms per step with 16 / 32 / 48 SMs).  The reference does not
// use atomics either: apply_gradients with use_locking=False is a plain racy read-modify-write of the variable on the
// parameter server (vgg/cifar10.py:368-372) — Hogwild.  So the pusher does exactly that, through shared memory:
//   loaders  cp.async.bulk  owner's HBM -> smem : 16 KB of the owner's master copy (up to ~100 KB per SM in flight,
//            one issuing thread);  cp.async.bulk local -> smem : the same chunk of my gradient
//   math     w <- w - lr * (gscale * g + wd * w), in place in shared memory (4 groups of 2 warps, a chunk each)
//   mover    cp.async.bulk  smem -> owner's HBM (the push) and smem -> my replica (the pull, same bytes)
// Two workers that update the same chunk at the same moment can lose one of the two updates, as in the reference.
// Peer memory is only ever the source of a plain (non-multicast) bulk load / the target of a bulk store.
// ---------------------------------------------------------------------------------------------------------
constexpr int BULK_CHUNK = 4096;                 // floats per chunk (16 KB)
constexpr int BULK_NG = 4;                       // gradient slots (local loads: short latency)
constexpr int BULK_NP = 9;                       // parameter slots (remote loads: this is what covers the link latency)
constexpr int BULK_K = 2;                        // rounds a store may stay pending before its slot is reused
// The math stage has a fixed cost per chunk (two barrier waits, a proxy fence, two arrivals: ~1 us) that a single
// group of warps pays serially — 8 KB chunks through one group ran at 7 GB/s per SM whatever the copy engine did.
// Four groups of two warps work on four consecutive chunks at once.
constexpr int BULK_GROUPS = 4;
constexpr int BULK_GROUP_THREADS = 64;
constexpr int BULK_MATH_THREADS = BULK_GROUPS * BULK_GROUP_THREADS;
constexpr int BULK_THREADS = 64 + BULK_MATH_THREADS + 32;
constexpr int BULK_SMEM = (BULK_NG + BULK_NP) * BULK_CHUNK * 4 + 1024;

__device__ __forceinline__ void bulk_load(uint32_t smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_dst), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_store(void* gdst, uint32_t smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_src), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_reduce_add_f32(void* gdst, uint32_t smem_src, uint32_t bytes) {
  asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;"
               ::"l"(gdst), "r"(smem_src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// The range [lo, lo+len) cut at owner boundaries and then into chunks; chunk c -> (first element, length, owner).
struct BulkChunks {
  long long lo, hi, shard;
  int first_owner, n_owners;
  long long total;
  __device__ void init(long long lo_, long long len, long long shard_) {
    lo = lo_; hi = lo_ + len; shard = shard_;
    first_owner = (int)(lo / shard);
    n_owners = len > 0 ? (int)((hi - 1) / shard) - first_owner + 1 : 0;
    total = 0;
    for (int k = 0; k < n_owners; ++k) total += count(k);
  }
  __device__ long long seg_lo(int k) const { const long long a = (long long)(first_owner + k) * shard; return a > lo ? a : lo; }
  __device__ long long seg_hi(int k) const { const long long b = (long long)(first_owner + k + 1) * shard; return b < hi ? b : hi; }
  __device__ long long count(int k) const { return (seg_hi(k) - seg_lo(k) + BULK_CHUNK - 1) / BULK_CHUNK; }
  __device__ void get(long long c, long long* gi, int* n, int* owner) const {
    int k = 0;
    for (; k < n_owners - 1; ++k) {
      const long long m = count(k);
      if (c < m) break;
      c -= m;
    }
    const long long a = seg_lo(k) + c * BULK_CHUNK, b = seg_hi(k);
    *gi = a;
    *n = (int)((b - a) < BULK_CHUNK ? (b - a) : BULK_CHUNK);
    *owner = first_owner + k;
  }
};

__global__ void __launch_bounds__(BULK_THREADS, 1)
ps_async_bulk_kernel(const AsyncArgs a) {
  extern __shared__ uint8_t bulk_smem_raw[];
  const uint32_t raw = smem_u32(bulk_smem_raw);
  uint8_t* smem = bulk_smem_raw + ((128u - (raw & 127u)) & 127u);
  float* sG = reinterpret_cast<float*>(smem);
  float* sP = sG + BULK_NG * BULK_CHUNK;
  uint64_t* fullG = reinterpret_cast<uint64_t*>(sP + BULK_NP * BULK_CHUNK);
  uint64_t* emptyG = fullG + BULK_NG;
  uint64_t* fullP = emptyG + BULK_NG;
  uint64_t* doneP = fullP + BULK_NP;       // math finished: the slot holds the new values
  uint64_t* emptyP = doneP + BULK_NP;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int me = a.rank, W = a.W;
  if (tid == 0) {
    for (int i = 0; i < BULK_NG; ++i) { mbar_init(&fullG[i], 1); mbar_init(&emptyG[i], BULK_GROUP_THREADS); }
    for (int i = 0; i < BULK_NP; ++i) { mbar_init(&fullP[i], 1); mbar_init(&doneP[i], BULK_GROUP_THREADS); mbar_init(&emptyP[i], 1); }
    fence_mbar_init();
    if (a.bump_step && blockIdx.x == 0) atomicAdd_system(a.global_step, 1ull);
  }
  __syncthreads();

  BulkChunks ch;
  ch.init(a.lo, a.len, a.shard);
  // this CTA's chunks: c = blockIdx.x + i * gridDim.x, rotated by rank so that the workers start on different owners
  const long long rot = (ch.total / W) * me;
  const long long n_my = ch.total > blockIdx.x ? (ch.total - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  float* mine = a.peer_params[me];
  const Hyper h = load_hyper(a.hyper);
#define BULK_CHUNK_OF(i, gi, n, own)                          \
  long long gi; int n, own;                                   \
  {                                                           \
    long long c_ = blockIdx.x + (i) * gridDim.x + rot;        \
    if (c_ >= ch.total) c_ -= ch.total;                       \
    ch.get(c_, &gi, &n, &own);                                \
  }

  if (warp == 0) {
    // ================= parameter loader: the owner's master copy, far ahead =================
    if (lane == 0) {
      for (long long i = 0; i < n_my; ++i) {
        const int sp = (int)(i % BULK_NP);
        mbar_wait(&emptyP[sp], (uint32_t)(((i / BULK_NP) & 1) ^ 1));
        BULK_CHUNK_OF(i, gi, n, own);
        mbar_arrive_expect_tx(&fullP[sp], (uint32_t)n * 4u);
        bulk_load(smem_u32(sP + sp * BULK_CHUNK), a.peer_params[own] + gi, (uint32_t)n * 4u, &fullP[sp]);
      }
    }
  } else if (warp == 1) {
    // ================= gradient loader (local) =================
    if (lane == 0) {
      for (long long i = 0; i < n_my; ++i) {
        const int sg = (int)(i % BULK_NG);
        mbar_wait(&emptyG[sg], (uint32_t)(((i / BULK_NG) & 1) ^ 1));
        BULK_CHUNK_OF(i, gi, n, own);
        (void)own;
        mbar_arrive_expect_tx(&fullG[sg], (uint32_t)n * 4u);
        bulk_load(smem_u32(sG + sg * BULK_CHUNK), a.grads + gi, (uint32_t)n * 4u, &fullG[sg]);
      }
    }
  } else if (warp < 2 + BULK_MATH_THREADS / 32) {
    // ================= math: w <- w - lr * (gscale * g + wd * w), in place =================
    const int grp = (tid - 64) / BULK_GROUP_THREADS, mt = (tid - 64) % BULK_GROUP_THREADS;
    for (long long i = grp; i < n_my; i += BULK_GROUPS) {
      const int sg = (int)(i % BULK_NG), sp = (int)(i % BULK_NP);
      BULK_CHUNK_OF(i, gi, n, own);
      (void)own;
      mbar_wait(&fullG[sg], (uint32_t)((i / BULK_NG) & 1));
      mbar_wait(&fullP[sp], (uint32_t)((i / BULK_NP) & 1));
      const float4* g4 = reinterpret_cast<const float4*>(sG + sg * BULK_CHUNK);
      float4* w4 = reinterpret_cast<float4*>(sP + sp * BULK_CHUNK);
      for (int t = mt; t < (n >> 2); t += BULK_GROUP_THREADS) {
        float4 g = g4[t];
        float4 w = w4[t];
        const long long e = gi + ((long long)t << 2);
        g.x *= h.gscale; g.y *= h.gscale; g.z *= h.gscale; g.w *= h.gscale;
        if (e + 0 < a.wd_boundary) g.x = fmaf(h.wd, w.x, g.x);
        if (e + 1 < a.wd_boundary) g.y = fmaf(h.wd, w.y, g.y);
        if (e + 2 < a.wd_boundary) g.z = fmaf(h.wd, w.z, g.z);
        if (e + 3 < a.wd_boundary) g.w = fmaf(h.wd, w.w, g.w);
        w.x = fmaf(-h.lr, g.x, w.x); w.y = fmaf(-h.lr, g.y, w.y);
        w.z = fmaf(-h.lr, g.z, w.z); w.w = fmaf(-h.lr, g.w, w.w);
        w4[t] = w;
      }
      fence_proxy_async_smem();            // my writes -> visible to the copy engine
      mbar_arrive(&doneP[sp]);             // -> mover
      mbar_arrive(&emptyG[sg]);            // -> gradient loader
    }
  } else if (lane == 0) {
    // ================= mover: new values -> the owner's master copy and my replica =================
    for (long long i = 0; i < n_my + BULK_K; ++i) {
      if (i < n_my) {
        const int sp = (int)(i % BULK_NP);
        BULK_CHUNK_OF(i, gi, n, own);
        mbar_wait(&doneP[sp], (uint32_t)((i / BULK_NP) & 1));
        bulk_store(a.peer_params[own] + gi, smem_u32(sP + sp * BULK_CHUNK), (uint32_t)n * 4u);      // the push
        if (own != me) bulk_store(mine + gi, smem_u32(sP + sp * BULK_CHUNK), (uint32_t)n * 4u);     // the pull
      }
      bulk_commit();
      bulk_wait_read<BULK_K>();            // the stores of round i-K have read their slot: back to the loader
      const long long jp = i - BULK_K;
      if (jp >= 0 && jp < n_my) mbar_arrive(&emptyP[(int)(jp % BULK_NP)]);
    }
    bulk_wait_all();                       // the stores have landed before the CTA goes away
  }
#undef BULK_CHUNK_OF
  __syncthreads();
  // ---- EMA of my own shard inside this range, owner side (local traffic only; see the register kernel) --------
  if (a.peer_ema && h.ema_decay >= 0.f) {
    float* ema = a.peer_ema[me];
    const float* wm = mine + (long long)me * a.shard;
    const float k = 1.f - h.ema_decay;
    long long e_lo = a.lo - (long long)me * a.shard, e_hi = a.lo + a.len - (long long)me * a.shard;
    e_lo = e_lo < 0 ? 0 : e_lo;
    e_hi = e_hi > a.shard ? a.shard : e_hi;
    for (long long i4 = (e_lo >> 2) + (long long)blockIdx.x * blockDim.x + tid; i4 < (e_hi >> 2);
         i4 += (long long)gridDim.x * blockDim.x) {
      float4 e = *reinterpret_cast<float4*>(ema + (i4 << 2));
      const float4 wv = ld_peer_f32x4(wm + (i4 << 2));
      e.x -= k * (e.x - wv.x); e.y -= k * (e.y - wv.y); e.z -= k * (e.z - wv.z); e.w -= k * (e.w - wv.w);
      *reinterpret_cast<float4*>(ema + (i4 << 2)) = e;
    }
  }
}

__global__ void ssp_clock_kernel(unsigned int* const* peer_clocks, int rank, int W, unsigned int step,
                                 unsigned int bound, long long timeout_cycles, unsigned int* status,
                                 unsigned int* waited_polls) {
  if (threadIdx.x < W) {
    __threadfence_system();
    st_release_sys(peer_clocks[threadIdx.x] + rank, step);
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  const unsigned int* mine = peer_clocks[rank];
  const long long t0 = clock64();
  unsigned int polls = 0;
  while (true) {
    unsigned int mn = 0xffffffffu;
    for (int q = 0; q < W; ++q) {
      const unsigned int c = ld_acquire_sys(mine + q);
      mn = c < mn ? c : mn;
    }
    if (step <= mn + bound) break;
    ++polls;
    __nanosleep(256);
    if (clock64() - t0 > timeout_cycles) { *status = 2; break; }
  }
  if (waited_polls) *waited_polls = polls;
}

static int update_grid(long long n) {
  long long b = (n + 511) / 512;
  const long long cap = 2LL * num_sms();
  return (int)(b < cap ? (b > 0 ? b : 1) : cap);
}

}  // namespace dtf

using namespace dtf;

extern "C" int dtf_fused_update(float* w, const float* g, float* ms, float* mom, float* ema, long long n,
                                long long n_decayed, const float* hyper, int kind, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  if (n == 0) return 0;
  DTF_REQUIRE(kind == 0 || (ms && mom), "RMSProp needs its slots");
  fused_update_kernel<<<update_grid(n), 512, 0, stream>>>(w, g, ms, mom, ema, n, n_decayed, hyper, kind);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

// The persistent convolution kernels need the SM's largest shared-memory carve-out (> 220 KB).  An SM cannot change
// its carve-out while blocks are resident, so a kernel that runs beside them must ask for the same split: with the
// default (L1-heavy) preference the 1 ms push of VGG's fc6 bucket kept every convolution of the backward pass
// waiting until its last block had left (profiles/trace_vgg_bsp_n2_before_carveout.txt).
template <typename K>
static void prefer_max_shared_once(K kernel, bool* done_for_device) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64 || done_for_device[dev]) return;
  cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  done_for_device[dev] = true;
}

static long long timeout_cycles_of(double timeout_s) {
  if (!(timeout_s > 0.0)) timeout_s = 300.0;
  return (long long)(timeout_s * 1.9e9);
}

extern "C" int dtf_ps_bucket_step(const float* const* peer_grads, float* const* peer_params,
                                  unsigned int* const* peer_flags, unsigned int* local_counter, unsigned int* status,
                                  int rank, int world, int ready_off, int updated_off, long long lo, long long len,
                                  long long chunk, long long wd_boundary, float* ms, float* mom, float* ema,
                                  const float* hyper, int kind, const unsigned int* epoch_ptr, unsigned int* epoch_out,
                                  const float* mc_grads, float* mc_params, double timeout_s, int max_blocks,
                                  void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(len % 4 == 0 && chunk % 4 == 0 && lo % 4 == 0, "bucket bounds must be multiples of 4 floats");
  DTF_REQUIRE(chunk * world >= len, "the owners' parts must cover the bucket");
  DTF_REQUIRE((mc_grads == nullptr) == (mc_params == nullptr), "multicast addresses come in pairs");
  DTF_REQUIRE(world <= 32, "flag signalling assumes world <= 32");
  DTF_REQUIRE(kind == 0 || (ms && mom), "RMSProp needs its slots");
  BucketArgs a;
  a.peer_grads = peer_grads; a.peer_params = peer_params; a.peer_flags = peer_flags;
  a.local_counter = local_counter; a.status = status; a.rank = rank; a.W = world;
  a.ready_off = ready_off; a.updated_off = updated_off; a.lo = lo; a.len = len; a.chunk = chunk;
  a.wd_boundary = wd_boundary; a.ms = ms; a.mom = mom; a.ema = ema; a.hyper = hyper; a.kind = kind;
  a.epoch_ptr = epoch_ptr; a.epoch_out = epoch_out; a.mc_grads = mc_grads; a.mc_params = mc_params;
  a.timeout_cycles = timeout_cycles_of(timeout_s);
  // blocks do not wait on one another (each one polls the flags itself), so the grid need not be co-resident;
  // it is sized so that a thread runs about one unrolled round, capped to leave SMs to the backward pass
  long long n = len - (long long)rank * chunk;
  n = n < 0 ? 0 : (n > chunk ? chunk : n);
  long long b = ((n >> 2) + (long long)PUSH_THREADS * PUSH_UNROLL - 1) / ((long long)PUSH_THREADS * PUSH_UNROLL);
  const int cap = max_blocks > 0 ? max_blocks : hw_num_sms();
  const int grid = (int)(b < cap ? (b > 0 ? b : 1) : cap);
  const bool mc = mc_grads != nullptr && world > 1;
  static bool carveout_set[4][64];
#define DTF_LAUNCH_BUCKET(K, M, slot)                                        \
  do {                                                                       \
    prefer_max_shared_once(ps_bucket_kernel<K, M>, carveout_set[slot]);      \
    ps_bucket_kernel<K, M><<<grid, PUSH_THREADS, 0, stream>>>(a);            \
  } while (0)
  if (kind == 0 && mc) DTF_LAUNCH_BUCKET(0, true, 0);
  else if (kind == 0) DTF_LAUNCH_BUCKET(0, false, 1);
  else if (mc) DTF_LAUNCH_BUCKET(1, true, 2);
  else DTF_LAUNCH_BUCKET(1, false, 3);
#undef DTF_LAUNCH_BUCKET
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_ps_gate(const unsigned int* flags, int n, const unsigned int* epoch_ptr, unsigned int* status,
                           double timeout_s, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  if (n <= 0) return 0;
  static bool carveout_set[64];
  prefer_max_shared_once(ps_gate_kernel, carveout_set);
  ps_gate_kernel<<<1, 32, 0, stream>>>(flags, n, epoch_ptr, status, timeout_cycles_of(timeout_s));
  DTF_CHECK(cudaGetLastError());
  return 0;
}

static int g_async_bulk = 1;
extern "C" int dtf_debug_set_async_bulk(int enabled) {
  const int prev = g_async_bulk;
  g_async_bulk = enabled;
  return prev;
}

extern "C" int dtf_ps_async_step(const float* grads, float* const* peer_params, float* const* peer_ms,
                                 float* const* peer_mom, float* const* peer_ema, unsigned long long* global_step,
                                 int rank, int world, long long shard, long long wd_boundary, const float* hyper,
                                 int kind, long long lo, long long len, int bump_step, int max_blocks, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(shard % 4 == 0, "shard length must be a multiple of 4 floats");
  DTF_REQUIRE(lo % 4 == 0 && len % 4 == 0 && lo >= 0 && lo + len <= shard * world, "range must be 4-aligned and inside the buffer");
  DTF_REQUIRE(kind == 0 || (peer_ms && peer_mom), "RMSProp needs its slots");
  AsyncArgs a;
  a.grads = grads; a.peer_params = peer_params; a.peer_ms = peer_ms; a.peer_mom = peer_mom; a.peer_ema = peer_ema;
  a.global_step = global_step; a.rank = rank; a.W = world; a.shard = shard; a.wd_boundary = wd_boundary;
  a.hyper = hyper; a.kind = kind; a.lo = lo; a.len = len; a.bump_step = bump_step;
  if (len == 0 && !bump_step) return 0;
  // SGD keeps 16 registers of gradient per thread: 1024 threads per SM at 64 registers; RMSProp also carries the
  // owner's two slots: 512 threads at 128
  const long long threads = kind == 0 ? 1024 : 512;
  long long b = ((len >> 2) + threads * ASYNC_UNROLL - 1) / (threads * ASYNC_UNROLL);
  const long long cap = max_blocks > 0 ? max_blocks : hw_num_sms();
  const int grid = (int)(b < cap ? (b > 0 ? b : 1) : cap);
  static bool carveout_set[2][64];
  if (kind == 0 && g_async_bulk) {
    static bool attr_set[64];
    int dev = 0;
    DTF_CHECK(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
      DTF_CHECK(cudaFuncSetAttribute(ps_async_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BULK_SMEM));
      attr_set[dev] = true;
    }
    const long long chunks = (len + BULK_CHUNK - 1) / BULK_CHUNK + world;
    const long long bcap = max_blocks > 0 ? max_blocks : 16;
    const int bgrid = (int)(chunks < bcap ? (chunks > 0 ? chunks : 1) : bcap);
    ps_async_bulk_kernel<<<bgrid, BULK_THREADS, BULK_SMEM, stream>>>(a);
    DTF_CHECK(cudaGetLastError());
    return 0;
  }
  if (kind == 0) {
    prefer_max_shared_once(ps_async_step_kernel<0, 1024>, carveout_set[0]);
    ps_async_step_kernel<0, 1024><<<grid, 1024, 0, stream>>>(a);
  } else {
    prefer_max_shared_once(ps_async_step_kernel<1, 512>, carveout_set[1]);
    ps_async_step_kernel<1, 512><<<grid, 512, 0, stream>>>(a);
  }
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_ssp_clock(unsigned int* const* peer_clocks, int rank, int world, unsigned int step,
                             unsigned int bound, double timeout_s, unsigned int* status, unsigned int* waited_polls,
                             void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(world <= 32, "clock table assumes world <= 32");
  ssp_clock_kernel<<<1, 32, 0, stream>>>(peer_clocks, rank, world, step, bound, timeout_cycles_of(timeout_s), status,
                                         waited_polls);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

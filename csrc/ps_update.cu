// K16 + C1/C2/C3/C4: parameter-server update kernels.
//
//  dtf_fused_update   one rank: w/slots/EMA update of a flat range from a local gradient (also the N=1 path)
//  dtf_ps_bsp_step    BSP step over peer-mapped HBM in ONE kernel per rank:
//                       wait "gradients ready" from all ranks            (flag words in peer memory)
//                       push : owner reads its range of every rank's gradient buffer over NVLink (P2P
//                              ld.global), averages, applies SGD / TF1-RMSProp(+momentum) / weight decay /
//                              EMA in registers, writes the fp32 master range              [reduce-scatter + K16]
//                       signal "range updated" to all ranks, wait for all owners
//                       pull : copy every other owner's range from its HBM into the local replica  [all-gather]
//  dtf_ps_async_step  ASP/SSP step: the *pusher* applies its own gradient to every owner's master range
//                       (red.global.add.f32 for SGD — Hogwild, the reference runs use_locking=False), bumps the
//                       global step on rank 0 with a system-scope atomic, then refreshes its replica; no barrier
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

// Wait until flags[q] >= epoch for every q < W.  Returns false on timeout (~2 s).
__device__ bool wait_all(const unsigned int* flags, int W, unsigned int epoch) {
  const long long t0 = clock64();
  for (int q = 0; q < W; ++q) {
    while ((int)(ld_acquire_sys(flags + q) - epoch) < 0) {
      __nanosleep(64);
      if (clock64() - t0 > 4000000000LL) return false;
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

struct BspArgs {
  const float* const* peer_grads;   // [W] gradient buffers of every rank (device pointers valid on this GPU)
  float* const* peer_params;        // [W] parameter replicas of every rank
  unsigned int* const* peer_flags;  // [W] flag blocks of every rank: [0,W) grads-ready, [W,2W) range-updated
  unsigned int* local_counter;      // block-completion counter (local, zeroed by the kernel's last block)
  unsigned int* status;             // 0 = ok, 1 = timeout
  int rank, W;
  long long shard;                  // elements per rank (flat placement: range of rank r = [r*shard,(r+1)*shard))
  long long wd_boundary;
  float* ms; float* mom; float* ema;   // slots of the local range (length shard)
  const float* hyper;
  int kind;
  unsigned int* epoch_ptr;          // local step counter: epoch = *epoch_ptr + 1 (device memory => graph-safe)
  // NVLS (NVSwitch multicast) addresses of the same symmetric buffers, or nullptr: with them the gradient reduction
  // happens inside the switch (multimem.ld_reduce) and the updated range is stored to every replica at once
  // (multimem.st), so the pull phase disappears.
  const float* mc_grads;
  float* mc_params;
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

__global__ void __launch_bounds__(512)
ps_bsp_step_kernel(const BspArgs a) {
  __shared__ int s_ok;
  const int W = a.W, me = a.rank;
  unsigned int* myflags = a.peer_flags[me];
  const unsigned int epoch = *reinterpret_cast<volatile unsigned int*>(a.epoch_ptr) + 1u;
  // ---- B1: announce "my gradients are complete", then wait for everybody's -----------------------------
  if (blockIdx.x == 0 && threadIdx.x < W) {
    __threadfence_system();
    st_release_sys(a.peer_flags[threadIdx.x] + me, epoch);
  }
  if (threadIdx.x == 0) s_ok = wait_all(myflags, W, epoch) ? 1 : 0;
  __syncthreads();
  if (!s_ok) { if (threadIdx.x == 0) *a.status = 1; return; }

  // ---- push: reduce my range over all ranks' gradients + optimiser -------------------------------------
  const Hyper h = load_hyper(a.hyper);
  const long long lo = (long long)me * a.shard;
  const float invW = 1.f / (float)W;
  float* wmine = a.peer_params[me] + lo;
  const long long n4 = a.shard >> 2;
  for (long long i4 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i4 < n4;
       i4 += (long long)gridDim.x * blockDim.x) {
    float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
    if (a.mc_grads != nullptr) {
      g = multimem_ld_reduce_add_f32x4(a.mc_grads + lo + (i4 << 2));     // sum over all ranks, formed in the switch
    } else {
      for (int q = 0; q < W; ++q) {
        const int src = (me + q) % W;      // stagger peers so ranks do not all hit the same GPU at once
        const float4 v = *reinterpret_cast<const float4*>(a.peer_grads[src] + lo + (i4 << 2));
        g.x += v.x; g.y += v.y; g.z += v.z; g.w += v.w;
      }
    }
    g.x *= invW; g.y *= invW; g.z *= invW; g.w *= invW;
    float4 wv = *reinterpret_cast<float4*>(wmine + (i4 << 2));
    float4 m2 = make_float4(0.f, 0.f, 0.f, 0.f), mo = m2;
    if (a.kind == 1) {
      m2 = *reinterpret_cast<float4*>(a.ms + (i4 << 2));
      mo = *reinterpret_cast<float4*>(a.mom + (i4 << 2));
    }
    const long long gi = lo + (i4 << 2);
    wv.x = update_elem(a.kind, h, wv.x, g.x, gi + 0 < a.wd_boundary, &m2.x, &mo.x);
    wv.y = update_elem(a.kind, h, wv.y, g.y, gi + 1 < a.wd_boundary, &m2.y, &mo.y);
    wv.z = update_elem(a.kind, h, wv.z, g.z, gi + 2 < a.wd_boundary, &m2.z, &mo.z);
    wv.w = update_elem(a.kind, h, wv.w, g.w, gi + 3 < a.wd_boundary, &m2.w, &mo.w);
    if (a.mc_params != nullptr && W > 1) multimem_st_f32x4(a.mc_params + lo + (i4 << 2), wv);   // every replica, mine too
    else *reinterpret_cast<float4*>(wmine + (i4 << 2)) = wv;
    if (a.kind == 1) {
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
  if (W == 1) return;

  // ---- B2: last block of this rank announces "my range is updated" -------------------------------------
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int done = atomicAdd(a.local_counter, 1u);
    if (done == gridDim.x - 1) {
      *a.local_counter = 0;
      *a.epoch_ptr = epoch;
      __threadfence_system();
      for (int q = 0; q < W; ++q) st_release_sys(a.peer_flags[q] + W + me, epoch);
    }
    s_ok = wait_all(myflags + W, W, epoch) ? 1 : 0;
  }
  __syncthreads();
  if (!s_ok) { if (threadIdx.x == 0) *a.status = 1; return; }

  // ---- pull: refresh the other owners' ranges of my replica (nothing to do when they were multicast) ------
  if (a.mc_params != nullptr) return;
  float* mine = a.peer_params[me];
  const long long tot4 = n4 * (W - 1);
  for (long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x; j < tot4;
       j += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(j / n4);
    const long long i4 = j - (long long)k * n4;
    const int owner = (me + 1 + k) % W;
    const long long off = (long long)owner * a.shard + (i4 << 2);
    *reinterpret_cast<float4*>(mine + off) = *reinterpret_cast<const float4*>(a.peer_params[owner] + off);
  }
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
};

__global__ void __launch_bounds__(512)
ps_async_step_kernel(const AsyncArgs a) {
  const Hyper h = load_hyper(a.hyper);
  const int W = a.W, me = a.rank;
  const long long total = a.shard * W;
  float* mine = a.peer_params[me];
  if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd_system(a.global_step, 1ull);
  // ---- push: apply my gradient to every owner's master range (owner order staggered by rank).
  // Only one-way traffic crosses NVLink for SGD: the weight-decay term uses this worker's replica of w (what
  // the gradient was computed from), so the update is a single red.global.add.f32 per element on the owner.
  for (long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x; j < total;
       j += (long long)gridDim.x * blockDim.x) {
    const int k = (int)(j / a.shard);
    const long long i = j - (long long)k * a.shard;
    const int owner = (me + k) % W;
    const long long gi = (long long)owner * a.shard + i;
    float g = a.grads[gi] * h.gscale;
    if (gi < a.wd_boundary) g = fmaf(h.wd, mine[gi], g);
    float* wp = a.peer_params[owner] + gi;
    if (a.kind == 0) {
      atomicAdd(wp, -h.lr * g);                       // one-sided red.add on the owner's HBM (Hogwild)
    } else {
      float* msp = a.peer_ms[owner] + i;              // slots live with the owner; racy RMW is the semantics
      float* mop = a.peer_mom[owner] + i;             // of use_locking=False
      const float m2 = h.rms_decay * (*msp) + (1.f - h.rms_decay) * g * g;
      const float mo = h.momentum * (*mop) + h.lr * g * rsqrtf(m2 + h.eps);
      *msp = m2;
      *mop = mo;
      atomicAdd(wp, -mo);
    }
  }
  // ---- pull: refresh my replica from the masters (no barrier: stale reads are the semantics) ------------
  __threadfence_system();
  for (long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x; j < total;
       j += (long long)gridDim.x * blockDim.x) {
    const int owner = (int)(j / a.shard);
    if (owner == me) continue;
    mine[j] = *reinterpret_cast<volatile const float*>(a.peer_params[owner] + j);
  }
  // ---- EMA of my own range, owner side (local traffic only).  The reference applies the average after
  // every worker's step; the owner sees W such steps per step of its own, so the host passes decay^W.
  if (a.peer_ema && h.ema_decay >= 0.f) {
    float* ema = a.peer_ema[me];
    const float* wm = mine + (long long)me * a.shard;
    const float k = 1.f - h.ema_decay;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.shard;
         i += (long long)gridDim.x * blockDim.x)
      ema[i] -= k * (ema[i] - wm[i]);
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

extern "C" int dtf_ps_bsp_step(const float* const* peer_grads, float* const* peer_params,
                               unsigned int* const* peer_flags, unsigned int* local_counter, unsigned int* status,
                               int rank, int world, long long shard, long long wd_boundary, float* ms, float* mom,
                               float* ema, const float* hyper, int kind, unsigned int* epoch_ptr,
                               const float* mc_grads, float* mc_params, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(shard % 4 == 0, "shard length must be a multiple of 4 floats");
  DTF_REQUIRE((mc_grads == nullptr) == (mc_params == nullptr), "multicast addresses come in pairs");
  DTF_REQUIRE(world <= 32, "flag signalling assumes world <= 32");
  BspArgs a;
  a.peer_grads = peer_grads; a.peer_params = peer_params; a.peer_flags = peer_flags;
  a.local_counter = local_counter; a.status = status; a.rank = rank; a.W = world; a.shard = shard;
  a.wd_boundary = wd_boundary; a.ms = ms; a.mom = mom; a.ema = ema; a.hyper = hyper; a.kind = kind; a.epoch_ptr = epoch_ptr;
  a.mc_grads = mc_grads; a.mc_params = mc_params;
  // every block must be co-resident (blocks wait on one another through the flags): one block per SM at most
  long long b = ((shard >> 2) + 511) / 512;
  const int grid = (int)(b < num_sms() ? (b > 0 ? b : 1) : num_sms());
  ps_bsp_step_kernel<<<grid, 512, 0, stream>>>(a);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_ps_async_step(const float* grads, float* const* peer_params, float* const* peer_ms,
                                 float* const* peer_mom, float* const* peer_ema, unsigned long long* global_step,
                                 int rank, int world, long long shard, long long wd_boundary, const float* hyper,
                                 int kind, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  AsyncArgs a;
  a.grads = grads; a.peer_params = peer_params; a.peer_ms = peer_ms; a.peer_mom = peer_mom; a.peer_ema = peer_ema;
  a.global_step = global_step; a.rank = rank; a.W = world; a.shard = shard; a.wd_boundary = wd_boundary;
  a.hyper = hyper; a.kind = kind;
  ps_async_step_kernel<<<update_grid(shard * world), 512, 0, stream>>>(a);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtf_ssp_clock(unsigned int* const* peer_clocks, int rank, int world, unsigned int step,
                             unsigned int bound, double timeout_s, unsigned int* status, unsigned int* waited_polls,
                             void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  DTF_REQUIRE(world <= 32, "clock table assumes world <= 32");
  const long long cycles = (long long)(timeout_s * 1.9e9);
  ssp_clock_kernel<<<1, 32, 0, stream>>>(peer_clocks, rank, world, step, bound, cycles, status, waited_polls);
  DTF_CHECK(cudaGetLastError());
  return 0;
}

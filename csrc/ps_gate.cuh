// C2: the pull side of the parameter-server protocol inside the consuming convolution.
//
// After a BSP push every owner stores its part of a bucket into every replica (multimem.st / P2P st) and then
// raises updated[b][owner] = epoch in each rank's flag block (csrc/ps_update.cu).  A convolution whose weights
// live in bucket b is handed those W flag words: its weight-load warp waits on them before its first weight TMA,
// so the forward pass of step i+1 starts while later buckets of step i are still crossing NVLink, and only the
// layer that needs a bucket waits for it.  (Reference behaviour replaced: every sess.run re-reads each variable
// from its parameter server before the op that consumes it, cifarnet/cifar10_cifarnet_bsp.py:122.)
//
// The wait is bounded; on timeout the status word is set to 1 (the host treats that as fatal) and the kernel
// proceeds, so a dead peer cannot hang the GPU.
#pragma once
#include "../include/dtf_b200.h"

namespace dtf {

__device__ __forceinline__ unsigned int gate_ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// One thread polls all the words.
__device__ __forceinline__ void weight_gate_thread(const dtf_conv_gemm_params& p) {
  if (p.gate_flags == nullptr) return;
  const unsigned int epoch = *reinterpret_cast<const volatile unsigned int*>(p.gate_epoch);
  const long long t0 = clock64();
  for (int q = 0; q < p.gate_n; ++q) {
    while ((int)(gate_ld_acquire_sys(p.gate_flags + q) - epoch) < 0) {
      __nanosleep(32);
      if (clock64() - t0 > p.gate_timeout_cycles) { atomicExch(p.gate_status, 1u); return; }
    }
  }
  // the weights are read by the TMA engine (async proxy): order its reads after the acquire above
  asm volatile("fence.proxy.async;" ::: "memory");
}

// A whole warp: lane q polls the words q, q+32, ...
__device__ __forceinline__ void weight_gate_warp(const dtf_conv_gemm_params& p) {
  if (p.gate_flags == nullptr) return;
  const unsigned int epoch = *reinterpret_cast<const volatile unsigned int*>(p.gate_epoch);
  const long long t0 = clock64();
  for (int q = (int)(threadIdx.x & 31u); q < p.gate_n; q += 32) {
    bool ok = true;
    while ((int)(gate_ld_acquire_sys(p.gate_flags + q) - epoch) < 0) {
      __nanosleep(32);
      if (clock64() - t0 > p.gate_timeout_cycles) { ok = false; break; }
    }
    if (!ok) { atomicExch(p.gate_status, 1u); break; }
  }
  asm volatile("fence.proxy.async;" ::: "memory");
  __syncwarp();
}

}  // namespace dtf

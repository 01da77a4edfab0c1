"""``SymmComm`` — the product transport: parameter-server shards in peer-mapped HBM, pushed and pulled by
our own sm_100a kernels over NVLink / NVSwitch (no NCCL on the step path).

Every rank allocates ONE symmetric buffer (``torch.distributed._symmetric_memory`` does the
cuMem/fabric-handle exchange — SURVEY §5.8 allows torch to bootstrap; the kernels are ours) laid out as

    [ params (total) | grads (total) | ms (shard) | mom (shard) | ema (shard) | control words ]

and learns the base address of every peer's copy.  Per-rank device tables of peer pointers are handed to
``dtf_ps_bsp_step`` / ``dtf_ps_async_step`` / ``dtf_ssp_clock`` (csrc/ps_update.cu):

* BSP  : one kernel per rank per step = wait(grads ready) -> reduce my range over all ranks' gradient
         buffers (in the NVSwitch with multimem.ld_reduce when a multicast address exists, else peer loads)
         + optimiser (+EMA) -> multimem.st of the updated range into every replica, or signal/wait(range
         updated) -> copy the other owners' ranges
         (reduce-scatter + fused update + all-gather; replaces SyncReplicasOptimizer's accumulators,
         chief thread and token queue).
* ASP  : pusher-side one-sided apply (``red.global.add.f32`` on the owner's HBM), global_step is a
         system-scope atomic on rank 0, replica refresh without a barrier.
* SSP  : ASP + ``clock[rank] = step`` written into every peer's clock table, bounded spin while
         ``step > min(clock) + bound`` (replaces ssp_manager.py's Thrift RPC).

NCCL is used only to bootstrap (process group, initial broadcast) and for ``max_over_ranks``.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch

from .. import _native as nat
from ..optim.schedule import ema_decay as _ema_decay
from ..optim.updates import OptimizerConfig, ShardOptimizer
from .comm import Comm, StalenessTimeout
from .layout import FlatLayout

CTRL_WORDS = 256      # uint32 control region per rank: [0,2W) bsp flags, [64,64+W) ssp clocks, [128,130) global step


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


class SymmComm(Comm):
    supports_async = True

    def __init__(self, group=None):
        import torch.distributed as dist
        self.dist = dist
        self.group = group if group is not None else dist.group.WORLD
        self.rank = dist.get_rank(self.group)
        self.world = dist.get_world_size(self.group)
        self._gs = 0
        self._step = 0

    # ------------------------------------------------------------------------------------------
    def setup(self, layout: FlatLayout, opt_cfg: OptimizerConfig) -> None:
        import torch.distributed._symmetric_memory as symm_mem
        assert layout.placement == "flat", "the in-kernel push/pull uses equal flat shards"
        dev = layout.params.device
        W = self.world
        total = layout.total
        shard = total // W
        assert shard % 4 == 0
        self.layout, self.shard, self.cfg = layout, shard, opt_cfg
        n = 2 * total + 3 * shard + CTRL_WORDS
        buf = symm_mem.empty(n, dtype=torch.float32, device=dev)
        buf.zero_()
        self.hdl = symm_mem.rendezvous(buf, self.group)
        self.buf = buf
        bases = [int(p) for p in self.hdl.buffer_ptrs]
        o_params, o_grads = 0, total
        o_ms, o_mom, o_ema = 2 * total, 2 * total + shard, 2 * total + 2 * shard
        o_ctrl = 2 * total + 3 * shard

        def table(off_floats: int) -> torch.Tensor:
            return torch.tensor([b + 4 * off_floats for b in bases], dtype=torch.int64, device=dev)

        self.t_params, self.t_grads = table(o_params), table(o_grads)
        self.t_ms, self.t_mom, self.t_ema = table(o_ms), table(o_mom), table(o_ema)
        self.t_flags = table(o_ctrl)
        self.t_clocks = table(o_ctrl + 64)
        self.gs_ptr_rank0 = bases[0] + 4 * (o_ctrl + 128)
        # NVLS: one multicast address for the whole symmetric buffer when the fabric offers it (NVSwitch)
        import os
        mc = int(getattr(self.hdl, "multicast_ptr", 0) or 0)
        if os.environ.get("DTF_NVLS", "1") == "0":
            mc = 0
        self.mc_grads = mc + 4 * o_grads if mc else 0
        self.mc_params = mc + 4 * o_params if mc else 0
        # bind the model to the symmetric params / grads (values preserved), chief's init wins
        layout.bind(buf[o_params:o_params + total], buf[o_grads:o_grads + total], copy=True)
        self.dist.broadcast(layout.params, src=0, group=self.group)
        lo = self.rank * shard
        opt = ShardOptimizer(opt_cfg, layout.params, lo, lo + shard, layout.wd_boundary)
        if opt_cfg.kind == "rmsprop":
            ms, mom = buf[o_ms:o_ms + shard], buf[o_mom:o_mom + shard]
            ms.fill_(1.0)
            mom.zero_()
            opt.slots = {"RMSProp": ms, "RMSProp_1": mom}
        if opt.ema is not None:
            ema = buf[o_ema:o_ema + shard]
            ema.copy_(opt.w)
            opt.ema = ema
        self.opt = opt
        self.local_counter = torch.zeros(1, dtype=torch.int32, device=dev)
        self.epoch = torch.zeros(1, dtype=torch.int32, device=dev)
        self.status = torch.zeros(1, dtype=torch.int32, device=dev)
        self.polls = torch.zeros(1, dtype=torch.int32, device=dev)
        self.gs_host = torch.zeros(1, dtype=torch.int64).pin_memory()
        self.gs_view = buf[o_ctrl + 128:o_ctrl + 130].view(torch.int64)
        torch.cuda.synchronize(dev)
        self.dist.barrier(group=self.group)

    # ------------------------------------------------------------------------------------------
    def _hyper(self, lr, grad_scale, num_updates, ema_power: int = 1):
        from ..ops import native_ops
        d = -1.0
        if self.opt.ema is not None:
            d = _ema_decay(self.cfg.ema.decay, num_updates if self.cfg.ema.use_num_updates else None) ** ema_power
        return native_ops.set_hyper(self.opt, lr, grad_scale, d)

    def bsp_step(self, lr, grad_scale=1.0):
        h = self._hyper(lr, grad_scale, self._gs)
        o = self.opt
        kind = 0 if self.cfg.kind == "sgd" else 1
        nat.check(nat.lib().dtf_ps_bsp_step(_p(self.t_grads), _p(self.t_params), _p(self.t_flags),
                                            _p(self.local_counter), _p(self.status), self.rank, self.world,
                                            C.c_longlong(self.shard), C.c_longlong(self.layout.wd_boundary),
                                            _p(o.slots.get("RMSProp")), _p(o.slots.get("RMSProp_1")), _p(o.ema),
                                            _p(h), kind, _p(self.epoch),
                                            C.c_void_p(self.mc_grads) if self.mc_grads else None,
                                            C.c_void_p(self.mc_params) if self.mc_params else None,
                                            nat.stream_ptr()), "dtf_ps_bsp_step")
        self._gs += 1
        self._step += 1
        return self._gs

    def async_step(self, lr, grad_scale=1.0):
        h = self._hyper(lr, grad_scale, self._gs, ema_power=self.world)
        kind = 0 if self.cfg.kind == "sgd" else 1
        has_ema = self.opt.ema is not None
        nat.check(nat.lib().dtf_ps_async_step(_p(self.layout.grads), _p(self.t_params),
                                              _p(self.t_ms) if kind == 1 else None,
                                              _p(self.t_mom) if kind == 1 else None,
                                              _p(self.t_ema) if has_ema else None,
                                              C.c_void_p(self.gs_ptr_rank0), self.rank, self.world,
                                              C.c_longlong(self.shard), C.c_longlong(self.layout.wd_boundary), _p(h),
                                              kind, nat.stream_ptr()), "dtf_ps_async_step")
        self._step += 1
        # global_step lives on rank 0; refresh the host mirror lazily (one step stale is fine for the LR schedule)
        self._gs += self.world if self.rank != 0 else 0
        if self.rank == 0:
            self.gs_host.copy_(self.gs_view, non_blocking=True)
            self._gs = max(self._gs + 1, int(self.gs_host.item()))
        return self._gs

    def ssp_wait(self, step, bound, timeout=600.0):
        nat.check(nat.lib().dtf_ssp_clock(_p(self.t_clocks), self.rank, self.world, C.c_uint(step), C.c_uint(bound),
                                          C.c_double(timeout), _p(self.status), _p(self.polls), nat.stream_ptr()),
                  "dtf_ssp_clock")
        return 0.0

    def check_status(self) -> None:
        s = int(self.status.item())
        if s == 1:
            raise RuntimeError("ps_bsp_step timed out waiting for a peer (a rank died or fell > 2 s behind)")
        if s == 2:
            raise StalenessTimeout("SSP clock wait timed out")

    # ------------------------------------------------------------------------------------------
    def barrier(self):
        self.dist.barrier(group=self.group)

    def broadcast_state(self, tensors, src=0):
        for t in tensors:
            self.dist.broadcast(t, src=src, group=self.group)

    def global_step(self):
        return self._gs

    def set_global_step(self, v):
        self._gs = int(v)

    def max_over_ranks(self, v):
        t = torch.tensor([v], dtype=torch.float64, device=self.layout.params.device)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX, group=self.group)
        return float(t.item())

    def owned_optimizers(self) -> List[ShardOptimizer]:
        return [self.opt]

    def gather_slot_flats(self):
        out = {}
        for name, t in self.opt.state_tensors().items():
            flat = torch.empty(self.layout.total, dtype=torch.float32, device=t.device)
            self.dist.all_gather_into_tensor(flat, t.contiguous().clone(), group=self.group)
            out[name] = flat
        return out

    def finish(self):
        torch.cuda.synchronize()
        self.check_status()
        self.dist.barrier(group=self.group)

"""``SymmComm`` — the product transport: parameter-server shards in peer-mapped HBM, pushed and pulled by
our own sm_100a kernels over NVLink / NVSwitch (no NCCL on the step path).

Every rank allocates ONE symmetric buffer (``torch.distributed._symmetric_memory`` does the
cuMem/fabric-handle exchange — SURVEY §5.8 allows torch to bootstrap; the kernels are ours) laid out as

    [ params (total) | grads (total) | ms (S) | mom (S) | ema (S) | control words ]

and learns the base address of every peer's copy (plus the NVLS multicast address of the whole buffer when the
fabric has one).

BSP (``mode='bsp'``) — bucketed push overlapped with the backward pass, pull gated in the consumer
--------------------------------------------------------------------------------------------------
The flat buffer is cut at tensor starts into *buckets* (``layout.plan_buckets``); inside a bucket every rank
owns 1/W.  The reference keeps one accumulator per variable, so its gradients leave the worker variable by
variable while the backward pass is still running (resnet/cifar10_resnet_bsp.py:91-99); here

* the engine tells the transport when a parameter's gradient has been written (``grad_ready``); as soon as the
  last tensor of a bucket has reported, ``dtf_ps_bucket_step`` is launched for it on a high-priority side stream
  behind an event — the reduction (in the NVSwitch with ``multimem.ld_reduce``, else peer loads), the optimiser
  and the store of the new values into every replica (``multimem.st``, else peer stores) then run while the
  remaining weight-gradient kernels of the step are still computing.  All ranks issue the buckets in one agreed
  order (learned in the first pass), so the per-bucket flag waits cannot cross;
* nobody waits for the *other* owners inside the push.  The next forward pass waits per bucket where the values
  are consumed: convolutions get the bucket's ``updated`` flag words and their weight-load warp spins on them
  before its first weight TMA (``csrc/ps_gate.cuh``); every other parameter (biases, batch-norm scales) is
  covered by a one-warp ``dtf_ps_gate`` launch in front of its first consumer on that stream;
* before this step's gradients may overwrite the buffers that peers were reading, a side stream waits for *all*
  flags of the previous step and only then zeroes the gradient buffer (``begin_step`` / ``before_backward``).

ASP / SSP — flat shards (rank r owns ``[r*shard, (r+1)*shard)``)
-----------------------------------------------------------------
* ASP  : pusher-side one-sided apply (``red.global.add.v4.f32`` on the owner's HBM), global_step is a
         system-scope atomic on rank 0, replica refresh in the same loop without a barrier.
* SSP  : ASP + ``clock[rank] = step`` written into every peer's clock table, bounded spin while
         ``step > min(clock) + bound`` (replaces ssp_manager.py's Thrift RPC).

NCCL is used only to bootstrap (process group, initial broadcast), for checkpoint gathers and ``max_over_ranks``.
"""
from __future__ import annotations

import bisect
import ctypes as C
import os
from typing import Dict, List, Optional

import torch

from .. import _native as nat
from ..optim.schedule import ema_decay as _ema_decay
from ..optim.updates import OptimizerConfig
from .comm import Comm, StalenessTimeout
from .layout import Bucket, FlatLayout, plan_buckets


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _round_up(a: int, b: int) -> int:
    return (a + b - 1) // b * b


class _OwnedState:
    """What this rank holds as a parameter server: optimiser slots and EMA shadows of the ranges it owns."""

    def __init__(self, cfg: OptimizerConfig, params: torch.Tensor):
        self.cfg = cfg
        self.w = params                    # (the whole replica: only its device is used through this handle)
        self.slots: Dict[str, torch.Tensor] = {}
        self.ema: Optional[torch.Tensor] = None

    def state_tensors(self) -> Dict[str, torch.Tensor]:
        out = dict(self.slots)
        if self.ema is not None:
            out["ExponentialMovingAverage"] = self.ema
        return out


class SymmComm(Comm):
    supports_async = True
    mode = "bsp"                           # set by the engine before setup(): decides the ownership map

    def __init__(self, group=None, overlap: Optional[bool] = None, bucket_mb: Optional[float] = None,
                 timeout_s: Optional[float] = None, ssp_timeout_s: Optional[float] = None):
        import torch.distributed as dist
        self.dist = dist
        self.group = group if group is not None else dist.group.WORLD
        self.rank = dist.get_rank(self.group)
        self.world = dist.get_world_size(self.group)
        self._gs = 0
        self._step = 0
        env = os.environ
        self.overlap = bool(int(env.get("DTF_PUSH_OVERLAP", "1"))) if overlap is None else bool(overlap)
        self.bucket_mb = float(env["DTF_BUCKET_MB"]) if (bucket_mb is None and "DTF_BUCKET_MB" in env) else bucket_mb
        # a peer that stays away this long is dead: the waits give up, the status word goes to 1 and the host aborts
        self.timeout_s = float(env.get("DTF_PS_TIMEOUT_S", "300")) if timeout_s is None else float(timeout_s)
        self.ssp_timeout_s = float(env.get("DTF_SSP_TIMEOUT_S", "30")) if ssp_timeout_s is None else float(ssp_timeout_s)
        self.gate_mode = env.get("DTF_PULL_GATE", "fused")          # fused | fetch (every gate as its own launch)
        # SMs that belong to the push while it runs beside the backward pass (the compute kernels are launched
        # with that many CTAs fewer); None = 16 for models of >= 16 MB, 0 (and a single bucket) for smaller ones,
        # whose push is pure latency and simply follows the last weight-gradient kernel
        self.push_sms = int(env["DTF_PUSH_SMS"]) if "DTF_PUSH_SMS" in env else None
        if env.get("DTF_ASYNC_BULK", "1") == "0":
            nat.lib().dtf_debug_set_async_bulk(0)              # A/B: SGD ASP push on the register kernel
        self.push_enabled = True           # the engine switches the pushes off for its warm-up passes
        self._in_step = False
        self._engine_driven = False        # True once an engine runs the step hooks (the pushes are then its)

    # ------------------------------------------------------------------------------------------
    def setup(self, layout: FlatLayout, opt_cfg: OptimizerConfig) -> None:
        import torch.distributed._symmetric_memory as symm_mem
        if layout.placement != "flat":
            raise ValueError(
                "--backend symm keeps the parameter-server shards as equal ranges of each bucket of the flat buffer; "
                f"--placement {layout.placement} (the reference's whole-tensor / axis-0 placement) is available "
                "with --backend nccl, gloo or the thread cluster")
        dev = layout.params.device
        W = self.world
        total = layout.total
        self.layout, self.cfg = layout, opt_cfg
        self.bucketed = self.mode == "bsp"          # ownership: 1/W of every bucket (BSP) or flat shards (ASP / SSP)
        small = total * 4 < (16 << 20)
        if self.push_sms is None:
            # the NVSwitch reduction (BSP) saturates from ~16 SMs; the bulk-copy-engine push (ASP / SSP) moves
            # ~0.3 ms x 8 / SMs per 100 MB on 8 GPUs, and 24 SMs balance it against the slower backward pass
            self.push_sms = 0 if (small or W == 1) else (16 if self.bucketed else 24)
        # small models: one bucket, pushed behind the last weight-gradient kernel on the compute stream itself
        self.inline = self.push_sms == 0
        if self.bucket_mb is None:
            target = total if small else min(max(total // 12, 1 << 18), 8 << 20)   # floats: ~12 buckets of <= 32 MB
        else:
            target = int(self.bucket_mb * (1 << 20) / 4)
        head = 0 if (small and self.bucket_mb is None) else int(float(os.environ.get("DTF_BUCKET_HEAD_KB", "64")) * 256)
        self.buckets: List[Bucket] = plan_buckets(layout.entries, total, W, target, head_floats=min(head, target))
        if self.bucketed:
            S = sum(b.chunk for b in self.buckets)
            self.shard = None
        else:
            assert (total // W) % 4 == 0
            self.shard = total // W
            S = self.shard
        B = max(1, len(self.buckets))
        self.n_slot = S
        # control words (uint32): ready[B][W] | updated[B][W] | clocks[W] (64-aligned) | global step (2 words)
        self.o_ready, self.o_updated = 0, B * W
        self.o_clocks = _round_up(2 * B * W, 64)
        self.o_gs = self.o_clocks + _round_up(W, 64)
        ctrl_words = self.o_gs + 64
        n = 2 * total + 3 * S + ctrl_words
        buf = symm_mem.empty(n, dtype=torch.float32, device=dev)
        buf.zero_()
        self.hdl = symm_mem.rendezvous(buf, self.group)
        self.buf = buf
        bases = [int(p) for p in self.hdl.buffer_ptrs]
        o_params, o_grads = 0, total
        o_ms, o_mom, o_ema = 2 * total, 2 * total + S, 2 * total + 2 * S
        self.o_ctrl = o_ctrl = 2 * total + 3 * S

        def table(off_floats: int) -> torch.Tensor:
            return torch.tensor([b + 4 * off_floats for b in bases], dtype=torch.int64, device=dev)

        self.t_params, self.t_grads = table(o_params), table(o_grads)
        self.t_ms, self.t_mom, self.t_ema = table(o_ms), table(o_mom), table(o_ema)
        self.t_flags = table(o_ctrl)
        self.t_clocks = table(o_ctrl + self.o_clocks)
        self.flags_local = bases[self.rank] + 4 * o_ctrl
        self.gs_ptr_rank0 = bases[0] + 4 * (o_ctrl + self.o_gs)
        # NVLS: one multicast address for the whole symmetric buffer when the fabric offers it (NVSwitch)
        mc = int(getattr(self.hdl, "multicast_ptr", 0) or 0)
        if os.environ.get("DTF_NVLS", "1") == "0":
            mc = 0
        self.mc_grads = mc + 4 * o_grads if mc else 0
        self.mc_params = mc + 4 * o_params if mc else 0
        # bind the model to the symmetric params / grads (values preserved), chief's init wins
        layout.bind(buf[o_params:o_params + total], buf[o_grads:o_grads + total], copy=True)
        self.dist.broadcast(layout.params, src=0, group=self.group)
        opt = _OwnedState(opt_cfg, layout.params)
        if opt_cfg.kind == "rmsprop":
            ms, mom = buf[o_ms:o_ms + S], buf[o_mom:o_mom + S]
            ms.fill_(1.0)
            mom.zero_()
            opt.slots = {"RMSProp": ms, "RMSProp_1": mom}
        if opt_cfg.ema is not None:
            opt.ema = buf[o_ema:o_ema + S]
        self.opt = opt
        if opt.ema is not None:
            self._owned_copy(layout.params, opt.ema)           # shadows start at the initial values
        self.counters = torch.zeros(B, dtype=torch.int32, device=dev)
        self.epoch = torch.zeros(1, dtype=torch.int32, device=dev)
        self.status = torch.zeros(1, dtype=torch.int32, device=dev)
        self.status_host = torch.zeros(1, dtype=torch.int32).pin_memory()
        self.status_event = None
        self.polls = torch.zeros(1, dtype=torch.int32, device=dev)
        self.gs_host = torch.zeros(1, dtype=torch.int64).pin_memory()
        self.gs_view = buf[o_ctrl + self.o_gs:o_ctrl + self.o_gs + 2].view(torch.int64)
        self.clocks_view = buf[o_ctrl + self.o_clocks:o_ctrl + self.o_clocks + W].view(torch.int32)
        # bucket bookkeeping
        self._bucket_lo = [b.lo for b in self.buckets]
        self._bucket_of_name = {name: b.index for b in self.buckets for name in b.names}
        self._need: List[Dict[int, int]] = []          # per bucket: id(param) -> notifications expected per step
        self._need_total: List[int] = []
        self._order: Optional[List[int]] = None        # agreed issue order of the buckets
        self._learned = False
        self._h = None
        self._reset_step_state()
        self.comm_stream = torch.cuda.Stream(device=dev, priority=-1)
        self.aux_stream = torch.cuda.Stream(device=dev)
        torch.cuda.synchronize(dev)
        self.dist.barrier(group=self.group)

    # ---- ownership maps ----------------------------------------------------------------------------
    def _owned_copy(self, flat: torch.Tensor, dst: torch.Tensor) -> None:
        """dst (length S) <- the elements of the full-length ``flat`` that this rank owns."""
        if not self.bucketed:
            dst.copy_(flat[self.rank * self.shard:(self.rank + 1) * self.shard])
            return
        for b in self.buckets:
            lo, hi = b.part(self.rank)
            if hi > lo:
                dst[b.slot_off:b.slot_off + hi - lo].copy_(flat[lo:hi])

    def _unshard(self, gathered: torch.Tensor) -> torch.Tensor:
        """[W, S] (every rank's slot array) -> full-length flat tensor in parameter order."""
        if not self.bucketed:
            return gathered.reshape(-1)[: self.layout.total]
        flat = torch.zeros(self.layout.total, dtype=gathered.dtype, device=gathered.device)
        for b in self.buckets:
            for r in range(self.world):
                lo, hi = b.part(r)
                if hi > lo:
                    flat[lo:hi].copy_(gathered[r, b.slot_off:b.slot_off + hi - lo])
        return flat

    def bucket_of(self, offset: int) -> int:
        return bisect.bisect_right(self._bucket_lo, offset) - 1

    # ---- hyper-parameters --------------------------------------------------------------------------
    def _hyper(self, lr, grad_scale, num_updates, ema_power: int = 1):
        from ..ops import native_ops
        d = -1.0
        if self.opt.ema is not None:
            d = _ema_decay(self.cfg.ema.decay, num_updates if self.cfg.ema.use_num_updates else None) ** ema_power
        return native_ops.set_hyper(self.opt, lr, grad_scale, d)

    # ---- BSP: step protocol driven by the engine ---------------------------------------------------
    def _reset_step_state(self) -> None:
        self._got: List[int] = [0] * len(self.buckets)      # notifications received per bucket this step
        self._seen: Dict[int, int] = {}
        self._events: List[Dict[int, torch.cuda.Event]] = [{} for _ in self.buckets]   # per bucket: stream -> last event
        self._complete: List[bool] = [False] * len(self.buckets)
        self._completion_order: List[int] = []
        self._next = 0                                      # position in the agreed order
        self._issued = 0
        self._gated = set()                                 # (bucket, stream) pairs that have waited this step
        self._deferred: Dict[int, tuple] = {}               # weight address -> (bucket, stream): gate owed
        self._zeroed = None
        self._all_gated = True
        if not self._learned:
            self._need_learning: Dict[int, Dict[int, int]] = {}
            self._last_seen_bucket_order: List[int] = []

    def prepare_step(self, lr: float, grad_scale: float = 1.0) -> None:
        """Hyper-parameters of the coming step into device memory (the push kernels run inside the step)."""
        self._h = self._hyper(lr, grad_scale, self._gs, ema_power=1 if self.bucketed else self.world)

    def begin_step(self) -> None:
        """Start of forward.  BSP: wait until every owner has finished the previous step (then nobody reads my old
        gradients any more) and zero the gradient buffer — on a side stream while the forward pass runs, or in line
        for small models.  ASP / SSP: nothing to wait for."""
        self._reset_step_state()
        self._in_step = True
        self._engine_driven = True
        sync = self.bucketed and self.world > 1
        if self.inline or not sync:
            if sync:
                self._gate_launch(0, len(self.buckets))
            self.layout.zero_grads()
            self._all_gated = True                 # the wait above covers every bucket on every stream that forks later
            return
        self._all_gated = False
        cur = torch.cuda.current_stream()
        ev = torch.cuda.Event()
        ev.record(cur)
        self.aux_stream.wait_event(ev)
        with torch.cuda.stream(self.aux_stream):
            self._gate_launch(0, len(self.buckets))
            self.layout.zero_grads()
            self._zeroed = torch.cuda.Event()
            self._zeroed.record(self.aux_stream)
        nat.set_pre_launch(self._flush_deferred)
        from ..ops import conv_native
        conv_native.set_gate_provider(self._gate_for_weight)

    def before_backward(self) -> None:
        if not self._in_step:
            return
        if self._zeroed is not None:
            self._flush_deferred()
            torch.cuda.current_stream().wait_event(self._zeroed)
        if self.push_sms and self.push_enabled and self.overlap and self.world > 1:
            nat.lib().dtf_reserve_sms(int(self.push_sms))      # the backward kernels leave these SMs to the push

    def end_step(self) -> None:
        """After backward: issue whatever has not been pushed yet and join the side streams."""
        if not self._in_step:
            return
        self._in_step = False
        nat.lib().dtf_reserve_sms(0)
        nat.set_pre_launch(None)
        from ..ops import conv_native
        conv_native.set_gate_provider(None)
        if not self._learned:
            self._finish_learning()
        if not self.push_enabled:
            return
        cur = torch.cuda.current_stream()
        if self.inline:
            for b in self._order[self._next:]:
                self._launch_bucket(b, cur)
            self._next = len(self._order)
            return
        if self._issued < len(self.buckets):
            ev = torch.cuda.Event()
            ev.record(cur)
            self.comm_stream.wait_event(ev)
            for b in self._order[self._next:]:
                self._launch_bucket(b)
            self._next = len(self._order)
        done = torch.cuda.Event()
        done.record(self.comm_stream)
        cur.wait_event(done)

    # -- gradient notifications (called from the backward pass, possibly on branch streams) -----------
    def grad_ready(self, param: torch.Tensor) -> None:
        if not self._in_step:
            return
        g = param.grad
        if g is None:
            return
        off = (g.data_ptr() - self.layout.grads.data_ptr()) // 4
        if off < 0 or off >= self.layout.total:
            return
        b = self.bucket_of(off)
        key = id(param)
        if not self._learned:
            need = self._need_learning.setdefault(b, {})
            need[key] = need.get(key, 0) + 1
            self._last_seen_bucket_order.append(b)
            return
        need = self._need[b]
        if key not in need:
            return
        self._seen[key] = self._seen.get(key, 0) + 1
        if self._seen[key] > need[key]:
            return
        self._got[b] += 1
        cur = torch.cuda.current_stream()
        ev = torch.cuda.Event()
        ev.record(cur)
        self._events[b][cur.cuda_stream] = ev         # a later event on the same stream covers the earlier ones
        if self._got[b] == self._need_total[b]:
            self._complete[b] = True
            self._issue_ready()

    def _issue_ready(self) -> None:
        if not (self.push_enabled and self.overlap) or self.inline:
            return
        while self._next < len(self._order) and self._complete[self._order[self._next]]:
            b = self._order[self._next]
            for ev in self._events[b].values():
                self.comm_stream.wait_event(ev)
            self._launch_bucket(b)
            self._next += 1

    def _finish_learning(self) -> None:
        """First pass over the model: which parameters report how often, and in which order the buckets complete.
        Rank 0's order is adopted by everybody (the per-bucket flag waits must be issued in one global order)."""
        nb = len(self.buckets)
        need = [dict(self._need_learning.get(b, {})) for b in range(nb)]
        order: List[int] = []
        remaining = [sum(n.values()) for n in need]
        for b in self._last_seen_bucket_order:
            remaining[b] -= 1
            if remaining[b] == 0:
                order.append(b)
        order += [b for b in range(nb - 1, -1, -1) if b not in order]        # buckets nobody reported for: last
        agreed = order
        if self.world > 1:
            t = torch.tensor(order, dtype=torch.int64, device=self.layout.params.device)
            self.dist.broadcast(t, src=0, group=self.group)
            agreed = [int(v) for v in t.tolist()]
        if agreed != order:
            # this rank's backward pass completes the buckets in another order: it still issues in the agreed one
            order = agreed
        self._need = need
        self._need_total = [sum(n.values()) for n in need]
        self._order = order
        self._learned = True
        self._next = 0

    def _launch_bucket(self, b: int, stream=None) -> None:
        bk = self.buckets[b]
        o = self.opt
        kind = 0 if self.cfg.kind == "sgd" else 1
        W = self.world
        blocks = int(self.push_sms or 8)

        def slot(t):
            return C.c_void_p(t.data_ptr() + 4 * bk.slot_off) if t is not None else None

        first, last = self._issued == 0, self._issued == len(self.buckets) - 1
        with torch.cuda.stream(stream if stream is not None else self.comm_stream):
            if self.bucketed:
                nat.check(nat.lib().dtf_ps_bucket_step(
                    _p(self.t_grads), _p(self.t_params), _p(self.t_flags),
                    C.c_void_p(self.counters.data_ptr() + 4 * b), _p(self.status), self.rank, W,
                    self.o_ready + b * W, self.o_updated + b * W,
                    C.c_longlong(bk.lo), C.c_longlong(bk.length), C.c_longlong(bk.chunk),
                    C.c_longlong(self.layout.wd_boundary),
                    slot(o.slots.get("RMSProp")), slot(o.slots.get("RMSProp_1")), slot(o.ema), _p(self._h), kind,
                    _p(self.epoch), _p(self.epoch) if last else None,
                    C.c_void_p(self.mc_grads) if self.mc_grads else None,
                    C.c_void_p(self.mc_params) if self.mc_params else None,
                    C.c_double(self.timeout_s), blocks, nat.stream_ptr(hook=False)), "dtf_ps_bucket_step")
            else:
                has_ema = o.ema is not None
                nat.check(nat.lib().dtf_ps_async_step(
                    _p(self.layout.grads), _p(self.t_params), _p(self.t_ms) if kind == 1 else None,
                    _p(self.t_mom) if kind == 1 else None, _p(self.t_ema) if has_ema else None,
                    C.c_void_p(self.gs_ptr_rank0), self.rank, W, C.c_longlong(self.shard),
                    C.c_longlong(self.layout.wd_boundary), _p(self._h), kind, C.c_longlong(bk.lo),
                    C.c_longlong(bk.length), 1 if first else 0, blocks if not self.inline else 0,
                    nat.stream_ptr(hook=False)), "dtf_ps_async_step")
        self._issued += 1

    # -- pull gates ------------------------------------------------------------------------------------
    def _gate_launch(self, b0: int, b1: int) -> None:
        W = self.world
        nat.check(nat.lib().dtf_ps_gate(C.c_void_p(self.flags_local + 4 * (self.o_updated + b0 * W)), (b1 - b0) * W,
                                        _p(self.epoch), _p(self.status), C.c_double(self.timeout_s),
                                        nat.stream_ptr(hook=False)), "dtf_ps_gate")

    def on_fetch(self, var) -> None:
        """A model function has just looked a variable up (``VariableStore.variable``): make sure the stream it is
        about to be used on waits for the owners' updates of its bucket."""
        if not (self._in_step and var.trainable) or self._all_gated:
            return
        b = self._bucket_of_name.get(var.name)
        if b is None:
            return
        sid = torch.cuda.current_stream().cuda_stream
        if (b, sid) in self._gated:
            return
        t = var.tensor
        if self.gate_mode == "fused" and t.dim() in (2, 4) and self._fused_gate_ok:
            self._deferred[t.data_ptr()] = (b, sid)       # a convolution is expected to take it (conv_native)
            return
        self._gated.add((b, sid))
        self._gate_launch(b, b + 1)

    def _flush_deferred(self) -> None:
        """Some other kernel is about to be launched: gates still owed are launched as kernels of their own."""
        if not self._deferred:
            return
        owed, self._deferred = self._deferred, {}
        cur = torch.cuda.current_stream().cuda_stream
        for (b, sid) in owed.values():
            if (b, sid) in self._gated:
                continue
            self._gated.add((b, sid))
            if sid == cur:
                self._gate_launch(b, b + 1)
            else:
                with torch.cuda.stream(torch.cuda.ExternalStream(sid)):
                    self._gate_launch(b, b + 1)

    def _gate_for_weight(self, w: torch.Tensor):
        """-> (flags, n, epoch, status, timeout cycles) for the convolution that is about to load ``w``, or None."""
        owed = self._deferred.pop(w.data_ptr(), None)
        if owed is None:
            return None
        b, sid = owed
        if (b, sid) in self._gated or sid != torch.cuda.current_stream().cuda_stream:
            if (b, sid) not in self._gated:
                self._deferred[w.data_ptr()] = owed
            return None
        self._gated.add((b, sid))
        W = self.world
        return (self.flags_local + 4 * (self.o_updated + b * W), W, self.epoch.data_ptr(), self.status.data_ptr(),
                int(self.timeout_s * 1.9e9))

    @property
    def _fused_gate_ok(self) -> bool:
        from ..ops import functional as Fn
        return Fn.get_backend() != "torch" and not getattr(self, "no_fused_gate", False)

    # ---- BSP: the call the engine (or a script) makes after backward -----------------------------------
    def bsp_step(self, lr, grad_scale=1.0):
        if not self._engine_driven:
            # stand-alone use (scripts that fill the gradient buffer themselves): the whole push happens here
            self._h = self._hyper(lr, grad_scale, self._gs)
            if self._order is None:
                self._order = list(range(len(self.buckets) - 1, -1, -1))
            cur = torch.cuda.current_stream()
            ev = torch.cuda.Event()
            ev.record(cur)
            self.comm_stream.wait_event(ev)
            for b in self._order:
                self._launch_bucket(b)
            done = torch.cuda.Event()
            done.record(self.comm_stream)
            cur.wait_event(done)
            if self.world > 1:
                self._gate_launch(0, len(self.buckets))      # stand-alone callers read the replica right away
            self._issued = 0
        self._gs += 1
        self._step += 1
        self._poll_status()
        return self._gs

    def async_step(self, lr, grad_scale=1.0):
        if not self._engine_driven:
            # stand-alone use: the whole buffer in one launch on the caller's stream
            self._h = self._hyper(lr, grad_scale, self._gs, ema_power=self.world)
            self._issued = 0
            for b in range(len(self.buckets)):
                self._launch_bucket(b, torch.cuda.current_stream())
            self._issued = 0
        self._step += 1
        # global_step lives on rank 0; refresh the host mirror lazily (one step stale is fine for the LR schedule)
        self._gs += self.world if self.rank != 0 else 0
        if self.rank == 0:
            self.gs_host.copy_(self.gs_view, non_blocking=True)
            self._gs = max(self._gs + 1, int(self.gs_host.item()))
        self._poll_status()
        return self._gs

    def ssp_wait(self, step, bound, timeout=None):
        timeout = self.ssp_timeout_s if timeout is None else timeout
        nat.check(nat.lib().dtf_ssp_clock(_p(self.t_clocks), self.rank, self.world, C.c_uint(step), C.c_uint(bound),
                                          C.c_double(timeout), _p(self.status), _p(self.polls), nat.stream_ptr()),
                  "dtf_ssp_clock")
        return 0.0

    # ---- failure detection -----------------------------------------------------------------------------
    def _poll_status(self) -> None:
        """A flag wait that timed out means a peer is gone and this rank's replica is no longer trustworthy: abort.
        The status word is copied to pinned memory every 16 steps and looked at once that copy has landed — no
        synchronisation on the step path."""
        if self.status_event is not None and self.status_event.query():
            self.status_event = None
            self._raise_on(int(self.status_host.item()))
        if self.status_event is None and self._step % 16 == 0 and not torch.cuda.is_current_stream_capturing():
            self.status_host.copy_(self.status, non_blocking=True)
            self.status_event = torch.cuda.Event()
            self.status_event.record()

    def _raise_on(self, s: int) -> None:
        if s == 1:
            raise RuntimeError("a parameter-server flag wait timed out after %.0f s: a peer rank died or stalled "
                               "(DTF_PS_TIMEOUT_S)" % self.timeout_s)
        if s == 2:
            raise StalenessTimeout("SSP clock wait timed out after %.0f s (DTF_SSP_TIMEOUT_S)" % self.ssp_timeout_s)

    def check_status(self) -> None:
        self._raise_on(int(self.status.item()))

    def sync_replica(self) -> None:
        if self.bucketed and self.world > 1:
            self._gate_launch(0, len(self.buckets))

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

    def owned_optimizers(self):
        return [self.opt]

    def gather_slot_flats(self):
        out = {}
        for name, t in self.opt.state_tensors().items():
            g = torch.empty(self.world * self.n_slot, dtype=torch.float32, device=t.device)
            self.dist.all_gather_into_tensor(g, t.contiguous().clone(), group=self.group)
            out[name] = self._unshard(g.view(self.world, self.n_slot))
        return out

    def load_slot_flats(self, flats) -> None:
        for name, t in self.opt.state_tensors().items():
            if name in flats:
                self._owned_copy(flats[name].to(t.device), t)

    def reset_ema_to_params(self) -> None:
        if self.opt.ema is not None:
            self._owned_copy(self.layout.params, self.opt.ema)

    def finish(self):
        torch.cuda.synchronize()
        self.check_status()
        self.dist.barrier(group=self.group)

"""Transports for the parameter-server protocol (SPMD form: every rank is a worker *and* owns PS shards).

The reference's L5 is re-expressed as three operations on flat buffers:

``bsp_step``   push = owner obtains the mean of all workers' gradients for the ranges it owns and applies the
               optimiser; pull = every worker refreshes its parameter copy from the owners
               (SyncReplicasOptimizer semantics: one update per global step from exactly N gradients)
``async_step`` push = the worker applies *its own* gradient to the owners' master ranges immediately
               (Hogwild, ``use_locking=False``), global_step += 1 per worker step; pull = copy back
``ssp_wait``   publish clock[rank] = step and block while step > min(clock) + bound
               (replaces the Thrift ``check_staleness`` RPC loop, inception/ssp_manager.py:21-29)

Back-ends
---------
``LocalComm``      world of one.
``ThreadCluster``  fake cluster: N ranks are threads of this process sharing CPU tensors — peers' buffers are
                   directly addressable, which makes it the CPU twin of the GPU symmetric-memory path (same
                   one-sided protocol, testable without hardware; SURVEY §4 "what we must add").
``TorchDistComm``  torch.distributed collectives (gloo on CPU, NCCL on GPU): BSP, and a lock-step emulation of
                   ASP / SSP.  This is the *baseline*
                   arm (BASELINE.md B1 all-reduce, B2 reduce-scatter + all-gather), never the product path.
``SymmComm``       (parallel/symm.py) peer-mapped HBM + our sm_100a push/pull kernels over NVLink.
"""
from __future__ import annotations

import threading
import time
from typing import Callable, List, Optional, Sequence

import torch

from ..optim.updates import OptimizerConfig, ShardOptimizer
from .layout import FlatLayout


class StalenessTimeout(RuntimeError):
    pass


class Comm:
    rank: int = 0
    world: int = 1
    supports_async = False

    def setup(self, layout: FlatLayout, opt_cfg: OptimizerConfig) -> None:
        raise NotImplementedError

    # -- protocol -------------------------------------------------------------------------------
    def bsp_step(self, lr: float, grad_scale: float = 1.0) -> int:
        raise NotImplementedError

    def async_step(self, lr: float, grad_scale: float = 1.0) -> int:
        raise NotImplementedError("this transport has no one-sided path")

    def ssp_wait(self, step: int, bound: int, timeout: float = 600.0) -> float:
        raise NotImplementedError

    # -- step hooks (transports that work from inside the step; defaults: whole-step protocol) ---------
    push_enabled = True

    def prepare_step(self, lr: float, grad_scale: float = 1.0) -> None:
        pass

    def begin_step(self) -> None:
        self.layout.zero_grads()

    def before_backward(self) -> None:
        pass

    def end_step(self) -> None:
        pass

    def sync_replica(self) -> None:
        """Make the current stream wait until the local parameter replica holds every owner's last update."""

    # -- plumbing -------------------------------------------------------------------------------
    def barrier(self) -> None:
        pass

    def broadcast_state(self, tensors: Sequence[torch.Tensor], src: int = 0) -> None:
        pass

    def global_step(self) -> int:
        raise NotImplementedError

    def set_global_step(self, v: int) -> None:
        raise NotImplementedError

    def max_over_ranks(self, v: float) -> float:
        return v

    def owned_optimizers(self) -> List[ShardOptimizer]:
        raise NotImplementedError

    def gather_slot_flats(self) -> dict:
        """Full-length flat tensors of the optimizer slots / EMA shadows (for the chief's checkpoint;
        SURVEY C8).  Default: this rank owns everything it needs (single rank or replicated optimiser)."""
        L = self.layout
        out = {}
        for o in self.owned_optimizers():
            for name, t in o.state_tensors().items():
                flat = out.setdefault(name, torch.zeros(L.total, dtype=torch.float32, device=t.device))
                flat[o.lo:o.hi].copy_(t)
        return out

    def load_slot_flats(self, flats: dict) -> None:
        """Inverse of :meth:`gather_slot_flats`: take this rank's part of full-length slot / shadow tensors
        (resume from a checkpoint)."""
        for o in self.owned_optimizers():
            for name, t in o.state_tensors().items():
                if name in flats:
                    t.copy_(flats[name][o.lo:o.hi].to(t.device))

    def reset_ema_to_params(self) -> None:
        """EMA shadows := current parameters (a checkpoint without shadows was restored)."""
        for o in self.owned_optimizers():
            if o.ema is not None:
                o.ema.copy_(o.w)

    def finish(self) -> None:
        pass


def _make_opts(layout: FlatLayout, cfg: OptimizerConfig, rank: int) -> List[ShardOptimizer]:
    return [ShardOptimizer(cfg, layout.params, lo, hi, layout.wd_boundary) for lo, hi in layout.owned(rank)]


# =================================================================================================
class LocalComm(Comm):
    """Single process: the PS shard is the whole buffer."""
    supports_async = True

    def __init__(self):
        self.rank, self.world = 0, 1
        self._gs = 0

    def setup(self, layout, opt_cfg):
        self.layout = layout
        self.opts = _make_opts(layout, opt_cfg, 0)

    def _apply(self, lr, grad_scale):
        for o in self.opts:
            o.apply(self.layout.grads[o.lo:o.hi], lr, grad_scale, num_updates=self._gs)
        self._gs += 1
        return self._gs

    def bsp_step(self, lr, grad_scale=1.0):
        return self._apply(lr, grad_scale)

    def async_step(self, lr, grad_scale=1.0):
        return self._apply(lr, grad_scale)

    def ssp_wait(self, step, bound, timeout=600.0):
        return 0.0

    def global_step(self):
        return self._gs

    def set_global_step(self, v):
        self._gs = int(v)

    def owned_optimizers(self):
        return self.opts


# =================================================================================================
class ThreadCluster:
    """Shared state of an in-process fake cluster; hand one ``ThreadComm`` to each worker thread."""

    def __init__(self, world: int):
        self.world = world
        self.barrier = threading.Barrier(world)
        self.layouts: List[Optional[FlatLayout]] = [None] * world
        self.opts: List[List[ShardOptimizer]] = [[] for _ in range(world)]
        self.clock = [0] * world
        self.clock_cv = threading.Condition()
        self.gs = 0
        self.gs_lock = threading.Lock()
        self.max_lead_seen = 0          # protocol checker: largest (my step - min clock) observed at release
        self.aborted = False

    def comm(self, rank: int) -> "ThreadComm":
        return ThreadComm(self, rank)

    def run(self, fn: Callable[["ThreadComm"], object]) -> List[object]:
        """Run ``fn(comm)`` on ``world`` threads; re-raises the first worker exception."""
        results: List[object] = [None] * self.world
        errors: List[BaseException] = []

        def work(r):
            try:
                results[r] = fn(self.comm(r))
            except BaseException as e:  # noqa: BLE001
                errors.append(e)
                self.aborted = True
                self.barrier.abort()
                with self.clock_cv:
                    self.clock_cv.notify_all()

        ts = [threading.Thread(target=work, args=(r,), daemon=True) for r in range(self.world)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        if errors:
            raise errors[0]
        return results


class ThreadComm(Comm):
    supports_async = True

    def __init__(self, cluster: ThreadCluster, rank: int):
        self.c = cluster
        self.rank = rank
        self.world = cluster.world

    def setup(self, layout, opt_cfg):
        self.layout = layout
        self.c.layouts[self.rank] = layout
        self.c.opts[self.rank] = _make_opts(layout, opt_cfg, self.rank)
        self.c.barrier.wait()
        # chief's initial values win (Supervisor: chief initialises, others wait — cifarnet…bsp.py:93-106)
        if self.rank != 0:
            self.layout.params.copy_(self.c.layouts[0].params)
        self.c.barrier.wait()

    # -- BSP ------------------------------------------------------------------------------------
    def bsp_step(self, lr, grad_scale=1.0):
        c = self.c
        c.barrier.wait()                                   # every worker's gradient is complete
        gs = c.gs
        for o in c.opts[self.rank]:
            g = c.layouts[0].grads[o.lo:o.hi].clone()
            for q in range(1, self.world):
                g.add_(c.layouts[q].grads[o.lo:o.hi])
            g.div_(self.world)                             # mean of N (SyncReplicas take_grad(N))
            o.apply(g, lr, grad_scale, num_updates=gs)
        c.barrier.wait()                                   # all owners updated
        if self.rank == 0:
            c.gs = gs + 1
        self._pull()
        c.barrier.wait()
        return c.gs

    def _pull(self):
        for lo, hi, owner in self.layout.segments:
            if owner != self.rank:
                self.layout.params[lo:hi].copy_(self.c.layouts[owner].params[lo:hi])

    # -- ASP / SSP ------------------------------------------------------------------------------
    def async_step(self, lr, grad_scale=1.0):
        c = self.c
        with c.gs_lock:
            gs = c.gs
            c.gs = gs + 1
        g_all = self.layout.grads
        for owner in range(self.world):
            for o in c.opts[owner]:
                # pusher-side apply on the owner's master range; no lock (Hogwild, use_locking=False)
                o.apply(g_all[o.lo:o.hi], lr, grad_scale, num_updates=gs)
        self._pull()
        return gs + 1

    def ssp_wait(self, step, bound, timeout=600.0):
        c = self.c
        t0 = time.time()
        with c.clock_cv:
            c.clock[self.rank] = step
            c.clock_cv.notify_all()
            while step > min(c.clock) + bound:
                if c.aborted:
                    raise StalenessTimeout("cluster aborted")
                if time.time() - t0 > timeout:
                    raise StalenessTimeout(f"rank {self.rank} step {step} waited > {timeout}s (clocks {c.clock})")
                c.clock_cv.wait(0.05)
            c.max_lead_seen = max(c.max_lead_seen, step - min(c.clock))
        return time.time() - t0

    # -- plumbing -------------------------------------------------------------------------------
    def barrier(self):
        self.c.barrier.wait()

    def broadcast_state(self, tensors, src=0):
        if not hasattr(self.c, "_bcast"):
            self.c._bcast = {}
        self.c.barrier.wait()
        if self.rank == src:
            self.c._bcast["t"] = [t.detach().clone() for t in tensors]
        self.c.barrier.wait()
        if self.rank != src:
            for dst, s in zip(tensors, self.c._bcast["t"]):
                dst.copy_(s)
        self.c.barrier.wait()

    def global_step(self):
        return self.c.gs

    def set_global_step(self, v):
        self.c.gs = int(v)

    def max_over_ranks(self, v):
        return v

    def owned_optimizers(self):
        return self.c.opts[self.rank]

    def gather_slot_flats(self):
        L = self.layout
        out = {}
        for r in range(self.world):
            for o in self.c.opts[r]:
                for name, t in o.state_tensors().items():
                    flat = out.setdefault(name, torch.zeros(L.total, dtype=torch.float32))
                    flat[o.lo:o.hi].copy_(t)
        return out


# =================================================================================================
class BucketedThreadComm(ThreadComm):
    """CPU twin of the bucketed BSP push / gated pull of ``parallel/symm.py`` on the thread cluster: the same step
    hooks (``begin_step`` / ``on_fetch`` / ``grad_ready`` / ``end_step``), the same bucket plan and per-bucket
    ownership, the same two flag rows per bucket (``ready``, ``updated``) with epochs that count completed steps —
    with Python threads and a condition variable in place of NVLink stores and acquire loads.  It moves real numbers,
    so a protocol mistake (a gradient read before it is complete, a parameter read before its owner's update, a
    gradient buffer zeroed under a reader) shows up as a result that differs from whole-buffer BSP.

    ``jitter``: upper bound (seconds) of random sleeps in front of every protocol action, to vary the interleavings;
    ``owner_delay``: the last rank updates its parts this much late (a slow parameter server).

    Usage: ``cluster = ThreadCluster(n); cluster.run(lambda c: work(BucketedThreadComm(cluster, c.rank, ...)))``."""

    def __init__(self, cluster: ThreadCluster, rank: int, bucket_floats: int = 1 << 16, head_floats: int = 0,
                 jitter: float = 0.0, seed: int = 0, owner_delay: float = 0.0):
        super().__init__(cluster, rank)
        self.bucket_floats, self.head_floats = bucket_floats, head_floats
        self.jitter = jitter
        self.owner_delay = owner_delay      # the last rank is a slow owner: sleeps this long before updating its part
        import random
        self._rng = random.Random(seed * 7919 + rank)
        self._gs = 0
        self._in_step = False

    def _nap(self):
        if self.jitter:
            time.sleep(self._rng.random() * self.jitter)

    def setup(self, layout, opt_cfg):
        from .layout import plan_buckets
        c = self.c
        self.layout = layout
        self.buckets = plan_buckets(layout.entries, layout.total, self.world, self.bucket_floats,
                                    head_floats=self.head_floats)
        self._bucket_of_name = {n: b.index for b in self.buckets for n in b.names}
        self._lo = [b.lo for b in self.buckets]
        c.layouts[self.rank] = layout
        # one optimiser per bucket part this rank owns
        opts = []
        for b in self.buckets:
            lo, hi = b.part(self.rank)
            if hi > lo:
                opts.append(ShardOptimizer(opt_cfg, layout.params, lo, hi, layout.wd_boundary))
        c.opts[self.rank] = opts
        self._opt_of_bucket = {self.bucket_of(o.lo): o for o in opts}
        with c.gs_lock:
            if not hasattr(c, "flags"):
                B, W = len(self.buckets), self.world
                c.flags = {"ready": [[[0] * W for _ in range(B)] for _ in range(W)],       # [rank][bucket][from]
                           "updated": [[[0] * W for _ in range(B)] for _ in range(W)]}
                c.flag_cv = threading.Condition()
        c.barrier.wait()
        if self.rank != 0:
            self.layout.params.copy_(c.layouts[0].params)
        c.barrier.wait()
        for o in opts:                      # (EMA shadows were cloned before the chief's values arrived)
            if o.ema is not None:
                o.ema.copy_(o.w)

    def bucket_of(self, offset: int) -> int:
        import bisect
        return bisect.bisect_right(self._lo, offset) - 1

    # -- flags ----------------------------------------------------------------------------------------
    def _wait(self, row: str, b: int, epoch: int, timeout: float = 60.0):
        c = self.c
        t0 = time.time()
        with c.flag_cv:
            while not all(v >= epoch for v in c.flags[row][self.rank][b]):
                if c.aborted or time.time() - t0 > timeout:
                    raise StalenessTimeout(f"rank {self.rank}: {row}[{b}] never reached epoch {epoch}")
                c.flag_cv.wait(0.02)

    def _raise_flag(self, row: str, b: int, epoch: int):
        c = self.c
        with c.flag_cv:
            for q in range(self.world):
                c.flags[row][q][b][self.rank] = epoch
            c.flag_cv.notify_all()

    # -- step hooks -------------------------------------------------------------------------------------
    def prepare_step(self, lr, grad_scale=1.0):
        self._lr, self._scale = lr, grad_scale

    def begin_step(self):
        # like the side stream of the GPU transport: gate on every bucket of the previous step (then nobody reads my
        # old gradients) and zero the gradients, while the forward pass is already running behind its own gates
        self._aux_error = None

        def aux():
            try:
                for b in range(len(self.buckets)):
                    self._wait("updated", b, self._gs)
                self._nap()
                self.layout.zero_grads()
            except BaseException as e:  # noqa: BLE001
                self._aux_error = e
        self._aux = threading.Thread(target=aux, daemon=True)
        self._aux.start()
        self._got = [0] * len(self.buckets)
        self._need = [len(b.names) for b in self.buckets]
        self._gated = set()
        self._pushed = [False] * len(self.buckets)
        self._in_step = True

    def before_backward(self):
        if self._in_step:
            self._aux.join()
            if self._aux_error is not None:
                raise self._aux_error

    def on_fetch(self, var):
        if not (self._in_step and var.trainable):
            return
        b = self._bucket_of_name.get(var.name)
        if b is None or b in self._gated:
            return
        self._gated.add(b)
        self._nap()
        self._wait("updated", b, self._gs)            # the pull: every owner's part of this bucket has landed

    def grad_ready(self, param):
        if not self._in_step or param.grad is None:
            return
        off = (param.grad.data_ptr() - self.layout.grads.data_ptr()) // 4
        if off < 0 or off >= self.layout.total:
            return
        b = self.bucket_of(off)
        self._got[b] += 1
        if self._got[b] == self._need[b]:
            self._push(b)

    def _push(self, b: int):
        c = self.c
        e = self._gs + 1
        bk = self.buckets[b]
        self._nap()
        self._raise_flag("ready", b, e)
        self._wait("ready", b, e)
        lo, hi = bk.part(self.rank)
        if hi > lo:
            self._nap()
            if self.owner_delay and self.rank == self.world - 1:
                time.sleep(self.owner_delay)
            g = c.layouts[0].grads[lo:hi].clone()
            for q in range(1, self.world):
                g.add_(c.layouts[q].grads[lo:hi])
            g.div_(self.world)
            o = self._opt_of_bucket[self.bucket_of(lo)]
            o.apply(g, self._lr, self._scale, num_updates=self._gs)
            self._nap()
            for q in range(self.world):                # store into every replica
                if q != self.rank:
                    c.layouts[q].params[lo:hi].copy_(self.layout.params[lo:hi])
        self._raise_flag("updated", b, e)
        self._pushed[b] = True

    def end_step(self):
        if not self._in_step:
            return
        self._in_step = False
        for b in range(len(self.buckets) - 1, -1, -1):     # tensors that reported nothing: pushed at the end
            if not self._pushed[b]:
                self._push(b)

    def sync_replica(self):
        for b in range(len(self.buckets)):
            self._wait("updated", b, self._gs)

    def bsp_step(self, lr, grad_scale=1.0):
        self._gs += 1
        return self._gs

    def global_step(self):
        return self._gs

    def set_global_step(self, v):
        self._gs = int(v)

    def gather_slot_flats(self):
        self.c.barrier.wait()
        out = ThreadComm.gather_slot_flats(self)
        self.c.barrier.wait()
        return out


# =================================================================================================
class TorchDistComm(Comm):
    """BSP (and emulated ASP / SSP) over torch.distributed collectives — the NCCL / gloo baseline arm.

    ``algo='rs_ag'``: reduce-scatter -> unfused shard update -> all-gather ("NCCL-PS");
    ``algo='allreduce'``: all-reduce -> every rank updates the whole buffer (DDP-style, B1).
    Requires ``placement='flat'`` (equal contiguous shards)."""

    def __init__(self, algo: str = "rs_ag", group=None):
        import torch.distributed as dist
        self.dist = dist
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.algo = algo
        self._gs = 0

    def setup(self, layout, opt_cfg):
        assert layout.placement == "flat", "collective transports need the flat placement"
        self.layout = layout
        dist = self.dist
        dist.broadcast(layout.params, src=0, group=self.group)
        self.is_gloo = dist.get_backend(self.group) == "gloo"
        if self.algo == "allreduce" or self.is_gloo:
            # whole-buffer optimiser on every rank
            self.opts = [ShardOptimizer(opt_cfg, layout.params, 0, layout.total, layout.wd_boundary)]
            self.shard = None
        else:
            s = layout.total // self.world
            lo = self.rank * s
            self.opts = [ShardOptimizer(opt_cfg, layout.params, lo, lo + s, layout.wd_boundary)]
            self.shard = torch.empty(s, dtype=torch.float32, device=layout.params.device)

    def bsp_step(self, lr, grad_scale=1.0):
        dist, L = self.dist, self.layout
        if self.shard is None:
            dist.all_reduce(L.grads, group=self.group)
            L.grads.div_(self.world)
            self.opts[0].apply(L.grads, lr, grad_scale, num_updates=self._gs)
        else:
            dist.reduce_scatter_tensor(self.shard, L.grads, group=self.group)
            self.shard.div_(self.world)
            o = self.opts[0]
            o.apply(self.shard, lr, grad_scale, num_updates=self._gs)
            dist.all_gather_into_tensor(L.params, L.params[o.lo:o.hi].clone(), group=self.group)
        self._gs += 1
        return self._gs

    supports_async = True

    def async_step(self, lr, grad_scale=1.0):
        """ASP / SSP on collectives — the baseline arm of the one-sided push.  A collective
        cannot be one-sided, so this is the lock-step emulation of what the reference's parameter servers end up
        doing in one round of W worker steps (vgg/cifar10.py:366-372, inception/imagenet_inception_asp.py:131-143):
        every worker ships slice r of its *own* gradient to owner r (all-to-all), the owner applies the W gradients
        one after the other (W optimiser steps, learning rate not scaled by W, global_step += W), and the fresh
        parameters are all-gathered."""
        dist, L = self.dist, self.layout
        W = self.world
        if self.shard is None:
            # all-reduce arm / gloo: every rank holds every gradient after an all-gather and applies them in rank order
            parts = [torch.empty_like(L.grads) for _ in range(W)]
            dist.all_gather(parts, L.grads, group=self.group)
            for q in range(W):
                self.opts[0].apply(parts[q], lr, grad_scale, num_updates=self._gs + q)
        else:
            s = self.shard.numel()
            if getattr(self, "_inbox", None) is None:
                self._inbox = torch.empty(W * s, dtype=torch.float32, device=L.grads.device)
            dist.all_to_all_single(self._inbox, L.grads[: W * s], group=self.group)
            o = self.opts[0]
            for q in range(W):
                o.apply(self._inbox[q * s:(q + 1) * s], lr, grad_scale, num_updates=self._gs + q)
            dist.all_gather_into_tensor(L.params, L.params[o.lo:o.hi].clone(), group=self.group)
        self._gs += W
        return self._gs

    def ssp_wait(self, step, bound, timeout=600.0):
        return 0.0            # the emulation above is lock-step: nobody is ever ahead

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
        return self.opts

    def gather_slot_flats(self):
        if self.shard is None:
            return Comm.gather_slot_flats(self)
        L = self.layout
        out = {}
        for name, t in self.opts[0].state_tensors().items():
            flat = torch.empty(L.total, dtype=torch.float32, device=t.device)
            self.dist.all_gather_into_tensor(flat, t.contiguous(), group=self.group)
            out[name] = flat
        return out

"""Dedicated parameter-server processes over ``torch.distributed`` (gloo) — the CPU / multi-host plumbing
mode (BASELINE.json config 1: "plain CNN BSP, 1 PS + 2 workers on localhost CPU").

This reproduces the reference's process roles (``--job_name=ps|worker``, cnn/cifar10_cnn_bsp.py:28-44): PS
processes own the variables and run the optimiser, workers compute gradients.  World layout: ranks
``[0, P)`` are PS tasks, ranks ``[P, P+W)`` are workers.  Ownership of the flat parameter buffer over the P
servers follows the layout's placement policy (flat / round_robin / axis0).

Wire protocol (all tensors fp32 on CPU):

BSP   per step the chief broadcasts a 4-float control word ``[op, lr, grad_scale, -]``; then for every PS k
      ``reduce(sum)`` of the workers' gradient slices to rank k (the accumulator + ``take_grad(N)``), PS k
      divides by W, applies the update and ``broadcast``s its fresh slices (the pull).
ASP   a worker ``send``s ``[op, task, lr, grad_scale | gradient slices of k]`` to PS k; PS k ``recv``s from any
      source, applies immediately and ``send``s back ``[global_step | fresh slices]``.  No barrier.
SSP   ASP + the clock service of ``parallel/ssp_manager.py`` on PS 0's host (port 8000 by default).

The GPU product path does not use this module: on one NVSwitch box every rank is worker and PS shard owner
at once and push/pull are our NVLink kernels (parallel/symm.py).
"""
from __future__ import annotations

import threading
from typing import List, Optional, Tuple

import torch
import torch.distributed as dist

from ..optim.updates import OptimizerConfig, ShardOptimizer
from .comm import Comm
from .layout import FlatLayout
from .ssp_manager import SspManager

OP_STOP, OP_BSP, OP_ASYNC, OP_INIT = 0.0, 1.0, 2.0, 3.0
HDR = 4


def _ranges_of(layout: FlatLayout, k: int) -> List[Tuple[int, int]]:
    return layout.owned(k)


def _pack(flat: torch.Tensor, ranges, out: Optional[torch.Tensor] = None, offset: int = 0) -> torch.Tensor:
    n = sum(hi - lo for lo, hi in ranges)
    if out is None:
        out = torch.empty(offset + n, dtype=torch.float32)
    pos = offset
    for lo, hi in ranges:
        out[pos:pos + hi - lo].copy_(flat[lo:hi])
        pos += hi - lo
    return out


def _unpack(buf: torch.Tensor, flat: torch.Tensor, ranges, offset: int = 0) -> None:
    pos = offset
    for lo, hi in ranges:
        flat[lo:hi].copy_(buf[pos:pos + hi - lo])
        pos += hi - lo


class PsWorkerComm(Comm):
    """Worker-side transport.  ``rank``/``world`` count workers; ``owners`` is the number of PS tasks."""
    supports_async = True

    def __init__(self, num_ps: int, num_workers: int, task_id: int, ssp_host: str = "127.0.0.1",
                 ssp_port: int = 8000, mode: str = "bsp", staleness: int = 5):
        self.P, self.W = num_ps, num_workers
        self.rank, self.world = task_id, num_workers
        self.owners = num_ps
        self.mode = mode
        self._gs = 0
        self.ssp_client = None
        self._ssp = (ssp_host, ssp_port, staleness)

    def setup(self, layout: FlatLayout, opt_cfg: OptimizerConfig) -> None:
        self.layout = layout
        self.ranges = [_ranges_of(layout, k) for k in range(self.P)]
        self.sizes = [sum(hi - lo for lo, hi in r) for r in self.ranges]
        # chief's initial values win: worker 0 broadcasts them to every process (PS included)
        dist.broadcast(layout.params, src=self.P)
        if self.mode == "ssp":
            self.ssp_client = SspManager(self.W, self._ssp[2], self._ssp[1]).create_rpc_client(self._ssp[0])

    # ---- BSP ---------------------------------------------------------------------------------
    def bsp_step(self, lr, grad_scale=1.0):
        ctrl = torch.tensor([OP_BSP, lr, grad_scale, 0.0])
        dist.broadcast(ctrl, src=self.P)
        L = self.layout
        for k in range(self.P):
            buf = _pack(L.grads, self.ranges[k])
            dist.reduce(buf, dst=k, op=dist.ReduceOp.SUM)
        for k in range(self.P):
            buf = torch.empty(self.sizes[k], dtype=torch.float32)
            dist.broadcast(buf, src=k)
            _unpack(buf, L.params, self.ranges[k])
        self._gs += 1
        return self._gs

    # ---- ASP / SSP ---------------------------------------------------------------------------
    def async_step(self, lr, grad_scale=1.0):
        L = self.layout
        for k in range(self.P):
            msg = torch.empty(HDR + self.sizes[k], dtype=torch.float32)
            msg[0], msg[1], msg[2], msg[3] = OP_ASYNC, float(self.rank), lr, grad_scale
            _pack(L.grads, self.ranges[k], msg, HDR)
            dist.send(msg, dst=k)
        for k in range(self.P):
            rep = torch.empty(1 + self.sizes[k], dtype=torch.float32)
            dist.recv(rep, src=k)
            if k == 0:
                self._gs = int(rep[0].item())
            _unpack(rep, L.params, self.ranges[k], 1)
        return self._gs

    def ssp_wait(self, step, bound, timeout=600.0):
        import time
        t0 = time.time()
        if self.ssp_client is not None:
            self.ssp_client.check_staleness(self.rank, step)
        return time.time() - t0

    # ---- plumbing ----------------------------------------------------------------------------
    def barrier(self):
        pass

    def broadcast_state(self, tensors, src=0):
        # non-trainable state is worker-local (BN statistics); workers align with the chief through the
        # worker sub-group when there is one
        pass

    def global_step(self):
        return self._gs

    def set_global_step(self, v):
        self._gs = int(v)

    def owned_optimizers(self):
        return []

    def gather_slot_flats(self):
        return {}

    def finish(self):
        if self.mode == "bsp":
            if self.rank == 0:
                dist.broadcast(torch.tensor([OP_STOP, 0.0, 0.0, 0.0]), src=self.P)
            else:
                ctrl = torch.zeros(4)
                dist.broadcast(ctrl, src=self.P)
        else:
            for k in range(self.P):
                msg = torch.zeros(HDR + self.sizes[k])
                msg[0], msg[1] = OP_STOP, float(self.rank)
                dist.send(msg, dst=k)
        if self.ssp_client is not None:
            self.ssp_client.close()


class ParameterServer:
    """The loop a ``--job_name=ps`` process runs instead of ``server.join()``."""

    def __init__(self, task_id: int, num_ps: int, num_workers: int, layout: FlatLayout, opt_cfg: OptimizerConfig,
                 mode: str = "bsp", ssp_host: str = "0.0.0.0", ssp_port: int = 8000, staleness: int = 5):
        self.k, self.P, self.W = task_id, num_ps, num_workers
        self.layout, self.mode = layout, mode
        self.ranges = [_ranges_of(layout, k) for k in range(num_ps)]
        self.sizes = [sum(hi - lo for lo, hi in r) for r in self.ranges]
        dist.broadcast(layout.params, src=num_ps)                 # receive the chief's initial values
        self.opts = [ShardOptimizer(opt_cfg, layout.params, lo, hi, layout.wd_boundary)
                     for lo, hi in self.ranges[task_id]]
        self.gs = 0
        self.ssp_server = None
        if mode == "ssp" and task_id == 0:
            mgr = SspManager(num_workers, staleness, ssp_port)
            self.ssp_server = mgr.create_rpc_server(ssp_host)
            self.ssp_server.serve_in_background()

    def _apply(self, gbuf: torch.Tensor, lr: float, scale: float) -> None:
        pos = 0
        for o in self.opts:
            n = o.hi - o.lo
            o.apply(gbuf[pos:pos + n], lr, scale, num_updates=self.gs)
            pos += n
        self.gs += 1

    def serve(self) -> int:
        if self.mode == "bsp":
            return self._serve_bsp()
        return self._serve_async()

    def _serve_bsp(self) -> int:
        L = self.layout
        while True:
            ctrl = torch.zeros(4)
            dist.broadcast(ctrl, src=self.P)
            if ctrl[0].item() == OP_STOP:
                break
            lr, scale = float(ctrl[1]), float(ctrl[2])
            for k in range(self.P):
                buf = torch.zeros(self.sizes[k])
                dist.reduce(buf, dst=k, op=dist.ReduceOp.SUM)
                if k == self.k:
                    buf.div_(self.W)                    # mean of exactly N gradients (take_grad(N))
                    self._apply(buf, lr, scale)
            for k in range(self.P):
                buf = _pack(L.params, self.ranges[k]) if k == self.k else torch.empty(self.sizes[k])
                dist.broadcast(buf, src=k)
        return self.gs

    def _serve_async(self) -> int:
        L = self.layout
        n = self.sizes[self.k]
        live = self.W
        while live > 0:
            msg = torch.empty(HDR + n)
            src = dist.recv(msg)                         # any source
            if msg[0].item() == OP_STOP:
                live -= 1
                continue
            self._apply(msg[HDR:], float(msg[2]), float(msg[3]))
            rep = torch.empty(1 + n)
            rep[0] = float(self.gs)
            _pack(L.params, self.ranges[self.k], rep, 1)
            dist.send(rep, dst=src)
        if self.ssp_server is not None:
            self.ssp_server.stop()
        return self.gs

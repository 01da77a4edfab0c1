"""Branch-level concurrency: the independent branches of an Inception-style block run on separate CUDA streams.

At the reference's batch sizes (32 images per worker for Inception-v3) most layers of the 17x17 and 8x8 blocks are
too small to occupy the chip: a 17x17x192 convolution is 73 output tiles on 148 SMs, an 8x8 one is 32, and the
element-wise kernels around them are latency-bound.  The branches of a block do not depend on each other, so they
are issued on side streams that fork from the block's input and join at its concatenation; two 73-CTA kernels then
run side by side and small kernels of one branch fill the bubbles of another.  Autograd replays every backward node
on the stream its forward ran on, so the backward pass gets the same concurrency, and a CUDA-graph capture of the
step records the fork / join edges as graph dependencies (the captured step keeps the parallel branches).

The reference has no counterpart (the TF1 executor schedules independent ops of one graph opportunistically,
SURVEY §2.6 "overlap"); this is the explicit streams-and-graphs form of it.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Sequence

import torch

_STATE = {"enabled": False, "depth": 0}
_POOLS: Dict[tuple, List["torch.cuda.Stream"]] = {}


def set_branch_streams(enabled: bool) -> None:
    _STATE["enabled"] = bool(enabled)


def branch_streams_enabled() -> bool:
    return _STATE["enabled"]


def _pool(device: torch.device, n: int) -> List["torch.cuda.Stream"]:
    """Side streams of the current nesting depth (a fork inside a branch gets its own set, so that it never waits
    on a sibling of its parent)."""
    idx = device.index if device.index is not None else torch.cuda.current_device()
    pool = _POOLS.setdefault((idx, _STATE["depth"]), [])
    while len(pool) < n:
        pool.append(torch.cuda.Stream(device=idx))
    return pool[:n]


def run_branches(thunks: Sequence[Callable[[], torch.Tensor]], like: torch.Tensor) -> List[torch.Tensor]:
    """Evaluate ``thunks`` (each returns one tensor); on CUDA with branch streams enabled, thunk 0 stays on the
    current stream and thunks 1.. run on side streams that wait for everything issued so far; the current stream
    waits for all of them before this function returns."""
    if not (_STATE["enabled"] and like.is_cuda and len(thunks) > 1):
        return [t() for t in thunks]
    main = torch.cuda.current_stream(like.device)
    side = _pool(like.device, len(thunks) - 1)
    fork = torch.cuda.Event()
    fork.record(main)
    outs: List[torch.Tensor] = [None] * len(thunks)          # type: ignore[list-item]
    joins = []
    _STATE["depth"] += 1
    try:
        # thunks run in their natural order on the host (default variable-scope names are numbered by call order);
        # the side streams wait on the fork event only, so they do not queue behind thunk 0's kernels
        outs[0] = thunks[0]()
        for i in range(1, len(thunks)):
            s = side[i - 1]
            s.wait_event(fork)
            with torch.cuda.stream(s):
                o = thunks[i]()
                done = torch.cuda.Event()
                done.record(s)
            o.record_stream(main)        # produced on a side stream, consumed (concatenated) on the main one
            outs[i] = o
            joins.append(done)
    finally:
        _STATE["depth"] -= 1
    for d in joins:
        main.wait_event(d)
    return outs


class Detached:
    """A computation issued on a side stream whose result is needed only later (Inception's auxiliary head runs
    beside the last blocks of the trunk): ``d = Detached(thunk, like)`` ... ``value = d.join()``."""

    def __init__(self, thunk: Callable[[], torch.Tensor], like: torch.Tensor):
        self.done = None
        if not (_STATE["enabled"] and like.is_cuda):
            self.value = thunk()
            return
        self.main = torch.cuda.current_stream(like.device)
        idx = like.device.index if like.device.index is not None else torch.cuda.current_device()
        pool = _POOLS.setdefault((idx, "detached"), [])
        if not pool:
            pool.append(torch.cuda.Stream(device=idx))
        s = pool[0]
        fork = torch.cuda.Event()
        fork.record(self.main)
        s.wait_event(fork)
        _STATE["depth"] += 100           # forks inside the detached computation use their own stream sets
        try:
            with torch.cuda.stream(s):
                self.value = thunk()
                self.done = torch.cuda.Event()
                self.done.record(s)
        finally:
            _STATE["depth"] -= 100

    def join(self) -> torch.Tensor:
        if self.done is not None:
            self.main = torch.cuda.current_stream(self.value.device)
            self.main.wait_event(self.done)
            self.value.record_stream(self.main)
            self.done = None
        return self.value

"""Host <-> device pipelining around ``Trainer.train_step``.

The reference overlaps input with compute through TensorFlow queue runners (CPU threads fill a queue while
``sess.run`` executes) and reads the loss back synchronously every step.  On a B200 a CIFAR-sized step
is ~1.7 ms, so a synchronous ``float(loss)`` (a full pipeline drain) and an un-overlapped host-to-device copy are a
visible fraction of it.  Two small pieces keep the GPU busy without changing what is computed or read:

* :class:`DevicePrefetcher` — wraps an iterator of pinned host batches; the copy of batch *i+1* is issued on a copy
  stream while step *i* computes; the compute stream waits on the copy's event only.
* :class:`ScalarReadback` — a ring of pinned host slots; ``push(device_scalar)`` enqueues the device-to-host copy,
  ``pop()`` returns the oldest value, waiting only for *that* copy.  Reading step *i*'s loss after step *i+1* has
  been launched keeps one step of work queued on the device.

Used by ``trainers/common.py`` (every trainer entry point) and by the end-to-end arm of ``bench.py``.
"""
from __future__ import annotations

import collections
from typing import Iterable, Iterator, Optional, Tuple

import torch


class DevicePrefetcher:
    """Iterator of device batches over an iterator of (pinned) host batches.

    ``depth + 1`` persistent device buffers are cycled: while the consumer works on buffer *k* the copy stream fills
    the next ones.  The buffers come from the compute stream's allocator pool once and are then re-used, so the
    steady state does no allocation at all (a fresh allocation per batch on the copy stream would hit ``cudaMalloc``
    — the caching allocator keeps separate pools per stream).  Ordering is by events only: the copy into a buffer
    waits for the event recorded after the last consumer of that buffer was enqueued; the consumer waits for the
    copy's event.  A returned batch is valid until the next ``next()``.
    """

    def __init__(self, batches: Iterable[Tuple[torch.Tensor, torch.Tensor]], device: torch.device, depth: int = 2,
                 lazy: bool = False):
        """``lazy``: create the copy stream now but issue the first copies at the first ``next()`` (so that a
        benchmark can set the pipeline up outside its timed region and still time every copy)."""
        self.it: Iterator = iter(batches)
        self.device = torch.device(device)
        self.cuda = self.device.type == "cuda"
        self.stream = torch.cuda.Stream(self.device) if self.cuda else None
        self.source = batches
        self.depth = max(1, depth)
        n = self.depth + 1
        self.buffers: list = [None] * n
        self.ready: list = [None] * n          # copy finished (recorded on the copy stream)
        self.free: list = [None] * n           # last consumer enqueued (recorded on the compute stream)
        self.q: collections.deque = collections.deque()
        self.wr = 0
        self.last: Optional[int] = None
        self.started = False
        if not lazy:
            self._start()

    def _start(self) -> None:
        self.started = True
        for _ in range(self.depth):
            self._issue()

    def _issue(self) -> None:
        try:
            x, y = next(self.it)
        except StopIteration:
            return
        if not self.cuda:
            self.q.append((x.to(self.device), y.to(self.device)))
            return
        k = self.wr % len(self.buffers)
        self.wr += 1
        cur = torch.cuda.current_stream(self.device)
        buf = self.buffers[k]
        if buf is None or buf[0].shape != x.shape or buf[1].shape != y.shape or buf[0].dtype != x.dtype:
            buf = self.buffers[k] = (torch.empty(x.shape, dtype=x.dtype, device=self.device),
                                     torch.empty(y.shape, dtype=y.dtype, device=self.device))
            self.free[k] = torch.cuda.Event()
            self.free[k].record(cur)       # whatever used this memory before, on the compute stream, comes first
        with torch.cuda.stream(self.stream):
            if self.free[k] is not None:
                self.stream.wait_event(self.free[k])
            if hasattr(x, "copy_into"):       # a batch produced on the device (data/cifar10.py::RecordBatch)
                x.copy_into(buf[0])
            else:
                buf[0].copy_(x, non_blocking=True)
            buf[1].copy_(y, non_blocking=True)
            if self.ready[k] is None:
                self.ready[k] = torch.cuda.Event()
            self.ready[k].record(self.stream)
        self.q.append(k)

    def __iter__(self):
        return self

    def __next__(self) -> Tuple[torch.Tensor, torch.Tensor]:
        if not self.started:
            self._start()
        if not self.q:
            raise StopIteration
        if not self.cuda:
            item = self.q.popleft()
            self._issue()
            return item
        cur = torch.cuda.current_stream(self.device)
        if self.last is not None:
            if self.free[self.last] is None:
                self.free[self.last] = torch.cuda.Event()
            self.free[self.last].record(cur)     # every reader of the previous batch has been enqueued by now
        k = self.q.popleft()
        cur.wait_event(self.ready[k])
        self.last = k
        self._issue()
        return self.buffers[k]

    def close(self) -> None:
        if hasattr(self.source, "close"):
            self.source.close()


class ScalarReadback:
    def __init__(self, device: torch.device, depth: int = 4):
        self.device = torch.device(device)
        self.cuda = self.device.type == "cuda"
        self.slots = [torch.zeros(1, dtype=torch.float32).pin_memory() if self.cuda else torch.zeros(1)
                      for _ in range(depth)]
        self.events: list = [None] * depth
        self.pending: collections.deque = collections.deque()
        self.next = 0

    def push(self, value: torch.Tensor, tag=None) -> None:
        """Enqueue the read-back of a device scalar (``tag`` is handed back by :meth:`pop`)."""
        if len(self.pending) == len(self.slots):
            raise RuntimeError("ScalarReadback ring is full: pop() before pushing more")
        i = self.next
        self.next = (self.next + 1) % len(self.slots)
        self.slots[i].copy_(value.detach().reshape(1), non_blocking=True)
        if self.cuda:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.device))
            self.events[i] = ev
        self.pending.append((i, tag))

    def __len__(self) -> int:
        return len(self.pending)

    def pop(self) -> Tuple[float, Optional[object]]:
        i, tag = self.pending.popleft()
        if self.cuda and self.events[i] is not None:
            self.events[i].synchronize()
        return float(self.slots[i][0]), tag

"""Per-step log lines in the reference's exact formats + JSONL metrics + Chrome trace.

cifar style   "time: <unix>; <datetime>: step %d (global_step %d), loss = %.2f (%.1f examples/sec; %.3f sec/batch)"
              (cifarnet/cifar10_cifarnet_bsp.py:128-129)
resnet style  "time: <unix>; <datetime>: step %d (gs %d), loss= %.2f (%.1f samples/s; %.3f s/batch)"
              (resnet/cifar10_resnet_bsp.py:142-143, inception/imagenet_inception_bsp.py:196-197)
cnn style     no "time:" prefix (cnn/cifar10_cnn_bsp.py:104-105)
``examples_per_sec`` is per worker: batch / step duration, as in the reference.
"""
from __future__ import annotations

import json
import sys
import time
from datetime import datetime
from typing import IO, Optional


def format_step(style: str, step: int, global_step: int, loss: float, examples_per_sec: float,
                sec_per_batch: float, now: Optional[float] = None) -> str:
    now = time.time() if now is None else now
    dt = datetime.fromtimestamp(now)
    if style == "cnn":
        return "%s: step %d (global_step %d), loss = %.2f (%.1f examples/sec; %.3f sec/batch)" % (
            dt, step, global_step, loss, examples_per_sec, sec_per_batch)
    if style == "resnet":
        return "time: %s; %s: step %d (gs %d), loss= %.2f (%.1f samples/s; %.3f s/batch)" % (
            now, dt, step, global_step, loss, examples_per_sec, sec_per_batch)
    return "time: %s; %s: step %d (global_step %d), loss = %.2f (%.1f examples/sec; %.3f sec/batch)" % (
        now, dt, step, global_step, loss, examples_per_sec, sec_per_batch)


def format_precision(precision: float, global_step: int, now: Optional[float] = None) -> str:
    """cnn/cifar10_eval.py:104"""
    return "%s: precision @ 1 = %.3f, global_step: %d" % (datetime.fromtimestamp(now or time.time()), precision,
                                                          global_step)


class MetricsWriter:
    """JSONL metrics (one object per logged step) — the new-design counterpart of TF summaries (§5.5)."""

    def __init__(self, path: str = ""):
        self.f: Optional[IO] = open(path, "a") if path else None

    def write(self, **kv) -> None:
        if self.f:
            self.f.write(json.dumps(kv) + "\n")
            self.f.flush()

    def close(self):
        if self.f:
            self.f.close()


class ChromeTrace:
    """Opt-in ``--trace``: phase durations per step as a Chrome-trace JSON (replaces the reference's
    FULL_TRACE run metadata that is collected and thrown away every step).

    Two kinds of rows.  Host rows (``add``, tid 0): wall-clock spans of what the Python loop did — on a GPU these only
    show how long it took to *enqueue* work.  Device rows (``add_device``, tid 1): spans between two CUDA events
    recorded on the stream the work runs on, i.e. the time the GPU spent; the events are resolved when the trace is
    saved, so tracing adds no synchronisation to the step loop."""

    def __init__(self, path: str = "", pid: int = 0, device=None):
        self.path, self.pid, self.events = path, pid, []
        self._pending = []
        self._base = None
        self._base_host = 0.0
        if path and device is not None and getattr(device, "type", "cpu") == "cuda":
            import time

            import torch
            self._torch = torch
            self._base = torch.cuda.Event(enable_timing=True)
            self._base.record()
            torch.cuda.synchronize(device)
            self._base_host = time.time()

    @property
    def on_device(self) -> bool:
        return self._base is not None

    def mark(self):
        """A timing event on the current stream (None when device tracing is off)."""
        if self._base is None:
            return None
        ev = self._torch.cuda.Event(enable_timing=True)
        ev.record()
        return ev

    def add(self, name: str, start_s: float, dur_s: float, tid: int = 0, **args) -> None:
        if self.path:
            self.events.append({"name": name, "ph": "X", "ts": start_s * 1e6, "dur": dur_s * 1e6, "pid": self.pid,
                                "tid": tid, "args": args})

    def add_device(self, name: str, ev0, ev1, t_run0: float = 0.0, **args) -> None:
        """Device time between two ``mark()`` events; ``t_run0``: the host time that the host rows count from."""
        if self.path and ev0 is not None and ev1 is not None:
            self._pending.append((name, ev0, ev1, t_run0, args))

    def save(self) -> None:
        if not self.path:
            return
        if self._pending:
            self._torch.cuda.synchronize()
            for name, ev0, ev1, t_run0, args in self._pending:
                start_ms = self._base.elapsed_time(ev0)
                self.events.append({"name": name, "ph": "X",
                                    "ts": (self._base_host - t_run0) * 1e6 + start_ms * 1e3,
                                    "dur": ev0.elapsed_time(ev1) * 1e3, "pid": self.pid, "tid": 1,
                                    "args": dict(args, clock="device (CUDA events)")})
            self._pending = []
        with open(self.path, "w") as f:
            json.dump({"traceEvents": self.events}, f)


def log(msg: str) -> None:
    """tf.logging.info equivalent: stderr, flushed (the launcher redirects it into worker_N.log)."""
    sys.stderr.write(msg + "\n")
    sys.stderr.flush()

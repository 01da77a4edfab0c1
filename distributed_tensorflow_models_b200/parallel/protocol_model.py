"""Executable model of the bucketed BSP push / gated pull protocol (``parallel/symm.py`` + ``csrc/ps_update.cu``),
used as the protocol checker of the fake backend: W rank threads run the same sequence of flag
operations as the kernels, every buffer cell carries a *version tag* instead of numbers, and each read asserts that
the value it is about to use is the one the protocol promises.

What a rank does per step ``e`` (epochs count completed steps, as on the device):

    begin_step       aux "stream":  wait updated[b][q] >= e-1 for all b, q ;  zero the gradient buffer
    forward          for each bucket b (first use):  wait updated[b][q] >= e-1 for all q ;  read params[b]
    backward         wait(aux) ; for each bucket b in reverse:  write grads[b] ;
                     push(b):  ready[q][b][me] = e  (release, to every q) ;  wait ready[me][b][q] >= e for all q ;
                               read every rank's grads[b][my part] ;  write every rank's params[b][my part] ;
                               updated[q][b][me] = e  (release, to every q)
    end_step         join the push stream

The hazards the flag discipline must exclude, checked as invariants here:

* an owner reducing bucket b of step e must read gradients written in step e from every rank — neither the zeros
  of step e+1 (``zeroed too early``) nor step e-1's values;
* a forward pass of step e reading bucket b must see every owner's part written in step e-1;
* an owner must not overwrite a rank's parameters of bucket b while that rank can still be reading them in step
  e's own forward/backward (it cannot: the rank raises ``ready`` only after its backward pass is through with b).

Threads yield at random points (``jitter``), so repeated runs explore different interleavings; the pushes of a rank
run on a separate thread per step, as the push stream does on the device.  ``break_rule`` switches one rule of the
protocol off so that the tests can show that the checker does catch the corresponding race.
"""
from __future__ import annotations

import random
import threading
import time
from typing import List, Optional


class ProtocolViolation(AssertionError):
    pass


class BucketProtocolModel:
    def __init__(self, world: int, buckets: int, steps: int, seed: int = 0, jitter: float = 2e-4,
                 break_rule: Optional[str] = None):
        self.W, self.B, self.steps = world, buckets, steps
        self.jitter = jitter
        self.break_rule = break_rule          # None | 'zero_without_gate' | 'forward_without_gate' | 'reduce_without_ready'
        self.rng = [random.Random(seed * 1000 + r) for r in range(world)]
        # version tags: grads[r][b] = step whose gradient the buffer holds (0 = zeroed, with the step that zeroed it
        # kept in zeroed_by); params[r][b][owner part q] = epoch of the value
        self.grads = [[-1] * buckets for _ in range(world)]
        self.params = [[[0] * world for _ in range(buckets)] for _ in range(world)]
        self.ready = [[[0] * world for _ in range(buckets)] for _ in range(world)]      # [rank][b][from]
        self.updated = [[[0] * world for _ in range(buckets)] for _ in range(world)]
        self.errors: List[str] = []
        self.cv = threading.Condition()
        self.deadline = None

    # -- primitives ------------------------------------------------------------------------------------
    def _yield(self, r):
        if self.jitter:
            time.sleep(self.rng[r].random() * self.jitter)

    def _wait(self, pred, what):
        with self.cv:
            while not pred():
                if self.errors or time.time() > self.deadline:
                    raise ProtocolViolation("timeout / abort while waiting for " + what)
                self.cv.wait(0.01)

    def _signal(self):
        with self.cv:
            self.cv.notify_all()

    def _fail(self, msg):
        self.errors.append(msg)
        self._signal()
        raise ProtocolViolation(msg)

    # -- one rank ----------------------------------------------------------------------------------------
    def _push(self, r, b, e):
        W = self.W
        self._yield(r)
        for q in range(W):
            self.ready[q][b][r] = e                         # st.release.sys to every rank
        self._signal()
        if self.break_rule != "reduce_without_ready":
            self._wait(lambda: all(self.ready[r][b][q] >= e for q in range(W)), f"ready[{b}] e{e} at rank {r}")
        self._yield(r)
        for q in range(W):                                  # reduce: my part of every rank's gradient buffer
            if self.grads[q][b] != e:
                self._fail(f"rank {r} reduces bucket {b} of step {e} but rank {q}'s gradient buffer holds "
                           f"{'zeros' if self.grads[q][b] == 0 else 'step %d' % self.grads[q][b]}")
        self._yield(r)
        for q in range(W):                                  # multimem.st / P2P st into every replica
            self.params[q][b][r] = e
        for q in range(W):                                  # fence; st.release.sys updated flag
            self.updated[q][b][r] = e
        self._signal()

    def _rank(self, r):
        W, B = self.W, self.B
        try:
            for e in range(1, self.steps + 1):
                # begin_step on the aux stream: gate on everything, then zero
                zero_done = threading.Event()

                def aux():
                    try:
                        if self.break_rule != "zero_without_gate":
                            self._wait(lambda: all(self.updated[r][b][q] >= e - 1 for b in range(B) for q in range(W)),
                                       f"gate_all e{e - 1} at rank {r}")
                        self._yield(r)
                        for b in range(B):
                            self.grads[r][b] = 0
                    except ProtocolViolation:
                        pass
                    finally:
                        zero_done.set()
                ta = threading.Thread(target=aux, daemon=True)
                ta.start()
                # forward: per-bucket gates, then read the parameters
                for b in range(B):
                    if self.break_rule != "forward_without_gate":
                        self._wait(lambda: all(self.updated[r][b][q] >= e - 1 for q in range(W)),
                                   f"gate[{b}] e{e - 1} at rank {r}")
                    self._yield(r)
                    for q in range(W):
                        if self.params[r][b][q] != e - 1:
                            self._fail(f"rank {r} forward of step {e} reads bucket {b}: owner {q}'s part is of epoch "
                                       f"{self.params[r][b][q]}, expected {e - 1}")
                zero_done.wait()
                # backward: buckets complete in reverse order; each push runs beside the rest of the backward pass
                pushes = []
                for b in reversed(range(B)):
                    self._yield(r)
                    if self.grads[r][b] != 0:
                        self._fail(f"rank {r} accumulates into bucket {b} that was not zeroed for step {e}")
                    self.grads[r][b] = e
                    t = threading.Thread(target=self._guard, args=(self._push, r, b, e), daemon=True)
                    # the push stream is in order: a bucket's kernel starts after the previous one of this rank
                    if pushes:
                        prev = pushes[-1]
                        t = threading.Thread(target=self._guard, args=(self._after, prev, r, b, e), daemon=True)
                    pushes.append(t)
                    t.start()
                for t in pushes:                                # end_step: join the push stream
                    t.join()
                if self.errors:
                    return
        except ProtocolViolation:
            pass

    def _after(self, prev, r, b, e):
        prev.join()
        self._push(r, b, e)

    def _guard(self, fn, *a):
        try:
            fn(*a)
        except ProtocolViolation:
            pass

    def run(self, timeout: float = 60.0) -> List[str]:
        """Run all ranks; returns the list of violations found (empty = the interleavings explored were clean)."""
        self.deadline = time.time() + timeout
        ts = [threading.Thread(target=self._rank, args=(r,), daemon=True) for r in range(self.W)]
        for t in ts:
            t.start()
        for t in ts:
            t.join(timeout + 5)
        return self.errors

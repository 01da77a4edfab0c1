"""Flat parameter / gradient buffers and parameter-server ownership.

The reference places whole variables on PS tasks round-robin (``tf.train.replica_device_setter``,
cnn/cifar10_cnn_bsp.py:42; ``VariableDeviceChooser``, inception/slim/variables.py:199-206) or slices each
variable along axis 0 (``tf.fixed_size_partitioner``, cifarnet/cifar10_cifarnet_bsp.py:52-53).  Here the PS
shards live in HBM on the worker ranks, so the natural unit is a *range of one flat fp32 buffer*:

* every trainable variable of the store becomes a view into ``params`` (and its ``.grad`` a view into
  ``grads``) at a 128-byte aligned offset — a whole model pushes/pulls as one contiguous span;
* tensors that carry weight decay are laid out first, so the fused update needs a single boundary
  (``wd_boundary``) instead of a per-element mask;
* ownership is a list of ``(lo, hi, owner)`` segments.  ``placement='flat'`` (default) cuts the buffer into
  ``world`` equal ranges — balanced even for VGG's 411 MB fc6; ``'round_robin'`` and
  ``'axis0'`` reproduce the reference's two policies (whole tensors k mod P; each tensor split along its
  leading stored axis into min(P, dim0) slices dealt round-robin).

Logical names/shapes are independent of the placement, as the reference's checkpoints are.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from ..slim.variables import Variable, VariableStore

ALIGN = 32  # floats = 128 bytes (TMA / vector alignment of every tensor start)


@dataclass
class Entry:
    name: str
    shape: Tuple[int, ...]
    numel: int
    offset: int
    weight_decay: float
    index: int           # creation order among trainables


def _round_up(a: int, b: int) -> int:
    return (a + b - 1) // b * b


@dataclass
class Bucket:
    """A contiguous range ``[lo, lo+length)`` of the flat buffer that is pushed as one unit (a group of layers whose
    gradients become complete at about the same time in the backward pass).  Inside a bucket rank ``r`` owns
    ``[lo + r*chunk, min(lo + (r+1)*chunk, lo+length))``; its optimiser slots for that part start at ``slot_off``
    of the rank's slot arrays."""
    index: int
    lo: int
    length: int
    chunk: int
    slot_off: int
    names: Tuple[str, ...]

    @property
    def hi(self) -> int:
        return self.lo + self.length

    def part(self, rank: int) -> Tuple[int, int]:
        lo = min(self.lo + rank * self.chunk, self.hi)
        return lo, min(lo + self.chunk, self.hi)


def plan_buckets(entries: Sequence["Entry"], total: int, world: int, target_floats: int,
                 max_buckets: int = 64, head_floats: int = 0) -> List[Bucket]:
    """Cut ``[0, total)`` at tensor starts into buckets of at least ``target_floats`` elements (a tensor is never
    split: its gradient is produced by one kernel, so its halves would become ready together anyway).

    ``head_floats`` > 0: the bucket at offset 0 closes after that many elements already.  The buffer starts with the
    first layers of the network, whose gradients are the last to be produced and whose new values are the first the
    next forward pass needs — that bucket's push cannot hide behind anything, so it should be pure latency."""
    ents = sorted(entries, key=lambda e: e.offset)
    target = max(int(target_floats), (total + max_buckets - 1) // max_buckets, 4 * world)
    head = max(int(head_floats), 4 * world) if head_floats else target
    cuts: List[Tuple[int, List[str]]] = []          # (start offset, names)
    cur_lo, cur_names = 0, []
    prev_big = False
    for e in ents:
        # a tensor that fills a bucket by itself (VGG's 411 MB fc6) starts one and ends it: its push then begins the
        # moment its own gradient is complete instead of waiting for the smaller neighbours laid out before it
        big = e.numel >= target
        if cur_names and (e.offset - cur_lo >= (head if not cuts else target) or big or prev_big):
            cuts.append((cur_lo, cur_names))
            cur_lo, cur_names = e.offset, []
        cur_names.append(e.name)
        prev_big = big
    cuts.append((cur_lo, cur_names))
    if len(cuts) > 1 and total - cuts[-1][0] < target // 4:       # a small tail joins its predecessor
        lo, names = cuts.pop()
        cuts[-1] = (cuts[-1][0], cuts[-1][1] + names)
    out: List[Bucket] = []
    slot_off = 0
    for i, (lo, names) in enumerate(cuts):
        hi = cuts[i + 1][0] if i + 1 < len(cuts) else total
        length = hi - lo
        assert length % 4 == 0 and lo % 4 == 0, (lo, hi)
        chunk = _round_up((length + world - 1) // world, 4)
        out.append(Bucket(i, lo, length, chunk, slot_off, tuple(names)))
        slot_off += chunk
    return out


class FlatLayout:
    def __init__(self, store: VariableStore, world_size: int = 1, placement: str = "flat",
                 device: Optional[torch.device] = None, default_weight_decay: Optional[float] = None,
                 decay_filter=None):
        """``default_weight_decay``/``decay_filter``: trainer-level L2 policy — e.g. 2e-3 on *all*
        trainables for ResNet, overriding per-variable regularisers."""
        self.store = store
        self.world = world_size
        self.placement = placement
        tvars = store.trainable_variables()
        if default_weight_decay is not None:
            for v in tvars:
                v.weight_decay = default_weight_decay if (decay_filter is None or decay_filter(v)) else 0.0
        wds = sorted({v.weight_decay for v in tvars if v.weight_decay})
        if len(wds) > 1:
            raise ValueError(f"more than one distinct weight-decay value {wds}: not supported by the flat layout")
        self.weight_decay = wds[0] if wds else 0.0
        order = sorted(range(len(tvars)), key=lambda i: (0 if tvars[i].weight_decay else 1, i))
        self.entries: List[Entry] = []
        off = 0
        self.wd_boundary = 0
        for i in order:
            v = tvars[i]
            n = v.tensor.numel()
            self.entries.append(Entry(v.name, tuple(v.tensor.shape), n, off, v.weight_decay, i))
            off += _round_up(n, ALIGN)
            if v.weight_decay:
                self.wd_boundary = off
        self.used = off
        self.total = _round_up(max(off, ALIGN), world_size * ALIGN)
        self.by_name: Dict[str, Entry] = {e.name: e for e in self.entries}
        first = tvars[0].tensor if tvars else None
        dev = device or (first.device if first is not None else torch.device("cpu"))
        self.device = dev
        self.params = torch.zeros(self.total, dtype=torch.float32, device=dev)
        self.grads = torch.zeros(self.total, dtype=torch.float32, device=dev)
        self.bind(self.params, self.grads)
        self.segments = self._make_segments()

    # ---- storage -------------------------------------------------------------------------------
    def bind(self, params: torch.Tensor, grads: torch.Tensor, copy: bool = True) -> None:
        """Re-point every trainable variable (and its .grad) at views of the given flat buffers."""
        assert params.numel() >= self.total and grads.numel() >= self.total
        with torch.no_grad():
            for e in self.entries:
                v = self.store.get(e.name)
                view = params[e.offset:e.offset + e.numel].view(e.shape)
                if copy:
                    view.copy_(v.tensor.detach().to(view.device))
                self.store.rebind(e.name, view)
                v.tensor.grad = grads[e.offset:e.offset + e.numel].view(e.shape)
        self.params, self.grads = params, grads

    def view(self, name: str, flat: Optional[torch.Tensor] = None) -> torch.Tensor:
        e = self.by_name[name]
        src = self.params if flat is None else flat
        return src[e.offset:e.offset + e.numel].view(e.shape)

    def zero_grads(self) -> None:
        self.grads.zero_()

    # ---- ownership -----------------------------------------------------------------------------
    def _make_segments(self) -> List[Tuple[int, int, int]]:
        w = self.world
        if self.placement == "flat":
            s = self.total // w
            return [(r * s, (r + 1) * s, r) for r in range(w)]
        segs: List[Tuple[int, int, int]] = []
        by_creation = sorted(self.entries, key=lambda e: e.index)
        if self.placement == "round_robin":
            for k, e in enumerate(by_creation):
                v = self.store.get(e.name)
                owner = v.ps_task if v.ps_task is not None else k
                segs.append((e.offset, e.offset + _round_up(e.numel, ALIGN), owner % w))
        elif self.placement == "axis0":
            nxt = 0
            for e in by_creation:
                dim0 = e.shape[0] if e.shape else 1
                n = max(1, min(w, dim0))
                inner = e.numel // max(dim0, 1)
                base, extra = divmod(dim0, n)
                row = 0
                for i in range(n):
                    rows = base + (1 if i < extra else 0)
                    lo = e.offset + row * inner
                    hi = e.offset + (row + rows) * inner
                    if i == n - 1:
                        hi = e.offset + _round_up(e.numel, ALIGN)
                    segs.append((lo, hi, nxt % w))
                    nxt += 1
                    row += rows
        else:
            raise ValueError(self.placement)
        segs.sort()
        # tail padding up to `total` belongs to the last owner of the list (keeps coverage exact)
        if segs and segs[-1][1] < self.total:
            lo, hi, o = segs[-1]
            segs[-1] = (lo, self.total, o)
        return segs

    def owned(self, rank: int) -> List[Tuple[int, int]]:
        """Merged ``(lo, hi)`` ranges of the flat buffer owned by ``rank``."""
        out: List[Tuple[int, int]] = []
        for lo, hi, o in self.segments:
            if o != rank:
                continue
            if out and out[-1][1] == lo:
                out[-1] = (out[-1][0], hi)
            else:
                out.append((lo, hi))
        return out

    def owner_bytes(self) -> List[int]:
        sizes = [0] * self.world
        for lo, hi, o in self.segments:
            sizes[o] += (hi - lo) * 4
        return sizes

    def describe(self) -> str:
        mb = self.used * 4 / 1e6
        return (f"FlatLayout({len(self.entries)} tensors, {self.used} floats = {mb:.2f} MB, world={self.world}, "
                f"placement={self.placement}, wd={self.weight_decay} on [0,{self.wd_boundary}))")

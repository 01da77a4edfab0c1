"""Race check for the branch-stream execution (ops/streams.py): gradients of one Inception-v3 step computed with the
branches on side streams vs one stream, against the run-to-run spread of the one-stream order itself (float atomics
in the split reductions make even that order non-bit-reproducible).  Prints per-run relative differences and the
variables that differ most."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from distributed_tensorflow_models_b200.ops import native_ops  # noqa: E402
from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup  # noqa: E402


def one(branch_streams, graph=False, bs=8, steps=1):
    native_ops._SEEDS.clear()
    t = Trainer(TrainSetup(model="inception", batch_size=bs, cuda_graph=graph, branch_streams=branch_streams),
                device=torch.device("cuda"))
    g = torch.Generator().manual_seed(0)
    x = torch.randn(bs, 299, 299, 3, generator=g).cuda()
    y = torch.randint(0, 1001, (bs,), generator=g).cuda()
    for _ in range(steps):
        out = t.train_step(x, y)
    torch.cuda.synchronize()
    return float(out["loss"]), t.layout.grads.clone(), t


def rel(a, b):
    return float((a - b).norm() / (b.norm() + 1e-20))


for graph in (False, True):
    la, ga, ta = one(False, graph)
    lb, gb, _ = one(False, graph)
    lc, gc, _ = one(True, graph)
    ld, gd, _ = one(True, graph)
    print(f"graph={graph}: loss seq {la:.6f} {lb:.6f} par {lc:.6f} {ld:.6f}")
    print(f"  grads: seq-vs-seq {rel(gb, ga):.3e}  par-vs-seq {rel(gc, ga):.3e}  par-vs-par {rel(gd, gc):.3e}")
    worst = []
    for e in ta.layout.entries:
        lo, hi = e.offset, e.offset + e.numel
        worst.append((rel(gc[lo:hi], ga[lo:hi]), rel(gb[lo:hi], ga[lo:hi]), e.name))
    for w in sorted(worst, reverse=True)[:6]:
        print("   par-vs-seq %.3e (seq-vs-seq %.3e) %s" % w)

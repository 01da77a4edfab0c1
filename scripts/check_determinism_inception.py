"""Where does run-to-run variation enter an Inception-v3 step?  Two identical one-stream runs (same seeds, same
data); per-variable relative gradient difference listed in backward order (logits first)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from distributed_tensorflow_models_b200.ops import functional as Fn  # noqa: E402
from distributed_tensorflow_models_b200.ops import native_ops  # noqa: E402
from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup  # noqa: E402

backend = sys.argv[1] if len(sys.argv) > 1 else "auto"
fuse = (sys.argv[2] == "fuse") if len(sys.argv) > 2 else True
Fn.set_backend(backend)
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False


def one(bs=8):
    native_ops._SEEDS.clear()
    t = Trainer(TrainSetup(model="inception", batch_size=bs, cuda_graph=False, branch_streams=False, fuse_bn=fuse),
                device=torch.device("cuda"))
    g = torch.Generator().manual_seed(0)
    x = torch.randn(bs, 299, 299, 3, generator=g).cuda()
    y = torch.randint(0, 1001, (bs,), generator=g).cuda()
    loss, logits = t._fwd_bwd(x, y)
    torch.cuda.synchronize()
    return float(loss), logits.clone(), t.layout.grads.clone(), t


def rel(a, b):
    return float((a - b).norm() / (b.norm() + 1e-20))


la, za, ga, ta = one()
lb, zb, gb, _ = one()
print(f"backend={backend} fuse_bn={fuse}: loss {la:.7f} {lb:.7f}  logits rel {rel(zb, za):.3e}  grads rel {rel(gb, ga):.3e}")
ents = sorted(ta.layout.entries, key=lambda e: -e.index)
for e in ents[:14] + ents[len(ents) // 2: len(ents) // 2 + 4] + ents[-6:]:
    lo, hi = e.offset, e.offset + e.numel
    print("   %.3e  |g| %.3e  %s" % (rel(gb[lo:hi], ga[lo:hi]), float(ga[lo:hi].norm()), e.name))

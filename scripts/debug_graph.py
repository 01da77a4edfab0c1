import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup
from distributed_tensorflow_models_b200.ops import native_ops as no
dev = torch.device("cuda")
g = torch.Generator().manual_seed(0)
x = torch.randn(16, 32, 32, 3, generator=g).to(dev); y = torch.randint(0, 10, (16,), generator=g).to(dev)
for fuse in (True, False):
    t = Trainer(TrainSetup(model="resnet", batch_size=16, cuda_graph=True, fuse_bn=fuse), device=dev)
    for step in range(3):
        out = t._fwd_bwd_graphed(x, y)
        torch.cuda.synchronize()
        a = no.arena(dev)
        print(f"fuse={fuse} replay {step}: loss {float(out[0]):.4f} arena[0:4] {a.buf[:4].tolist()} arena off {a.off} sum {float(a.buf.sum()):.3f} |g| {float(t.layout.grads.norm()):.4f}")

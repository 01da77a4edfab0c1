"""Run every native op once at a representative ResNet / Inception shape (for an `ncu --set full` capture of the
non-convolution kernels; the conv kernels have their own captures)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distributed_tensorflow_models_b200.ops import functional as Fn

Fn.set_backend("auto")
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)


def r(*s):
    return torch.randn(*s, generator=g).to(dev).requires_grad_(True)


for _ in range(2):
    x = r(128, 32, 32, 16)
    gamma, beta = r(16), r(16)
    mm, mv = torch.zeros(16, device=dev), torch.ones(16, device=dev)
    y = Fn.batch_norm(x, gamma, beta, mm, mv, training=True, eps=1e-5, decay=0.997, relu=True)
    y.sum().backward()
    x2 = r(32, 35, 35, 288)
    Fn.max_pool(x2, 3, 2, "VALID").sum().backward()
    Fn.avg_pool(x2, 3, 1, "SAME").sum().backward()
    x3 = r(512, 12, 12, 64)
    Fn.lrn(x3, 4, 1.0, 0.001 / 9.0, 0.75).sum().backward()
    logits = r(128, 10)
    labels = torch.randint(0, 10, (128,), generator=g).to(dev)
    Fn.softmax_cross_entropy(logits, labels).backward()
    Fn.dropout(r(128, 4096), 0.5, training=True).sum().backward()
    Fn.global_avg_pool(r(128, 8, 8, 64)).sum().backward()
torch.cuda.synchronize()
print("ok")

import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_models_b200.ops import conv_native as cn, native_ops as no, functional as Fn
dev = torch.device("cuda")
g = torch.Generator().manual_seed(0)
def r(*s): return torch.randn(*s, generator=g).to(dev)
def stress(name, fn, iters=100, tol=0.0):
    ref = fn().clone(); torch.cuda.synchronize()
    bad = 0; worst = 0.0
    for i in range(iters):
        o = fn()
        d = float((o - ref).abs().max()) / (float(ref.abs().max()) + 1e-12)
        worst = max(worst, d)
        if d > tol: bad += 1
    print(f"{name}: {bad}/{iters} runs differ (worst rel {worst:.3e})", flush=True)
x = r(16, 32, 32, 16); w = r(16, 3, 3, 16) * 0.1; dy = r(16, 32, 32, 16)
sc = torch.rand(16, generator=g).to(dev) + 0.5; sh = r(16) * 0.1
stress("fprop 16->16", lambda: cn.conv_fprop(x, w, (1, 1), (1, 1)))
stress("fprop in_affine", lambda: cn.conv_fprop(x, w, (1, 1), (1, 1), in_scale=sc, in_shift=sh, in_relu=True))
def f_stats():
    st = torch.zeros(32, device=dev); cn.conv_fprop(x, w, (1, 1), (1, 1), stats=st); return st
stress("fprop stats", f_stats, tol=1e-5)
wt = cn.weight_transpose(w)
stress("dgrad", lambda: cn.conv_dgrad(dy, wt, (32, 32), (1, 1), (1, 1)))
stress("wgrad", lambda: cn.conv_wgrad(x, dy, w.shape, (1, 1), (1, 1)), tol=1e-5)
stress("wgrad in_affine", lambda: cn.conv_wgrad(x, dy, w.shape, (1, 1), (1, 1), in_scale=sc, in_shift=sh, in_relu=True), tol=1e-5)
x2 = r(16, 16, 16, 32); w2 = r(64, 3, 3, 32) * 0.1
stress("fprop s2 32->64", lambda: cn.conv_fprop(x2, w2, (2, 2), (1, 1)))
dy2 = r(16, 8, 8, 64); wt2 = cn.weight_transpose(w2)
stress("dgrad s2", lambda: cn.conv_dgrad(dy2, wt2, (16, 16), (2, 2), (1, 1)))
stress("wgrad s2", lambda: cn.conv_wgrad(x2, dy2, w2.shape, (2, 2), (1, 1)), tol=1e-5)
xl = r(16, 1, 1, 64); wl = r(10, 1, 1, 64); dyl = r(16, 1, 1, 10)
stress("fc fprop", lambda: cn.conv_fprop(xl, wl, (1, 1), (0, 0)))
stress("fc dgrad", lambda: cn.conv_dgrad(dyl, cn.weight_transpose(wl), (1, 1), (1, 1), (0, 0)))
stress("fc wgrad", lambda: cn.conv_wgrad(xl, dyl, wl.shape, (1, 1), (0, 0)), tol=1e-5)
stress("bn_stats", lambda: (no.begin_step(dev), no.bn_stats(x))[1].clone(), tol=1e-5)
# full model: which tensors differ between reruns, in backward order
from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup
xs = r(16, 32, 32, 3); ys = torch.randint(0, 10, (16,), generator=g).to(dev)
t = Trainer(TrainSetup(model="resnet", batch_size=16, cuda_graph=False, fuse_bn=False), device=dev)
t._fwd_bwd(xs, ys); torch.cuda.synchronize(); g0 = t.layout.grads.clone()
for rep in range(3):
    t._fwd_bwd(xs, ys); torch.cuda.synchronize(); g1 = t.layout.grads
    names = []
    for e in sorted(t.layout.entries, key=lambda e: -e.index):
        a = g0[e.offset:e.offset+e.numel]; b = g1[e.offset:e.offset+e.numel]
        d = float((a-b).abs().max()) / max(float(a.abs().max()), 1e-9)
        if d > 1e-4: names.append((e.name, round(d, 4)))
    print("rerun", rep, "differing (backward order, first 6):", names[:6], "count", len(names), flush=True)

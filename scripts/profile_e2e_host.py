"""Host-side time per phase of the end-to-end loop (ResNet-32 b128): where does the CPU spend a step?"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup
from distributed_tensorflow_models_b200.utils.pipeline import DevicePrefetcher, ScalarReadback

dev = torch.device("cuda", 0)
t = Trainer(TrainSetup(model="resnet", batch_size=128), device=dev)
g = torch.Generator().manual_seed(0)
hx = [torch.randn(128, 32, 32, 3, generator=g).pin_memory() for _ in range(8)]
hy = [torch.randint(0, 10, (128,), generator=g).pin_memory() for _ in range(8)]
for i in range(10):
    t.train_step(hx[0].to(dev), hy[0].to(dev))
torch.cuda.synchronize()
N = 300


def timed(name, loop):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    w0 = time.perf_counter(); e0.record()
    acc = loop()
    e1.record(); torch.cuda.synchronize(); w1 = time.perf_counter()
    print(f"{name}: device {e0.elapsed_time(e1)/N:.4f} ms/step, wall {(w1-w0)*1e3/N:.4f} ms/step; host phases (us/step): " +
          ", ".join(f"{k} {v*1e6/N:.1f}" for k, v in acc.items()), flush=True)


def sync_loop():
    acc = {"h2d": 0.0, "step": 0.0, "read": 0.0}
    for i in range(N):
        a = time.perf_counter()
        x = hx[i % 8].to(dev, non_blocking=True); y = hy[i % 8].to(dev, non_blocking=True)
        b = time.perf_counter()
        o = t.train_step(x, y)
        c = time.perf_counter()
        _ = float(o["loss"])
        d = time.perf_counter()
        acc["h2d"] += b - a; acc["step"] += c - b; acc["read"] += d - c
    return acc


def piped_loop(prefetch=True):
    acc = {"next": 0.0, "step": 0.0, "push": 0.0, "pop": 0.0}
    feed = DevicePrefetcher(((hx[i % 8], hy[i % 8]) for i in range(N)), dev) if prefetch else None
    rb = ScalarReadback(dev)
    for i in range(N):
        a = time.perf_counter()
        if prefetch:
            x, y = next(feed)
        else:
            x = hx[i % 8].to(dev, non_blocking=True); y = hy[i % 8].to(dev, non_blocking=True)
        b = time.perf_counter()
        o = t.train_step(x, y)
        c = time.perf_counter()
        rb.push(o["loss"])
        d = time.perf_counter()
        if len(rb) > 1:
            rb.pop()
        e = time.perf_counter()
        acc["next"] += b - a; acc["step"] += c - b; acc["push"] += d - c; acc["pop"] += e - d
    while len(rb):
        rb.pop()
    return acc


def nosync_loop():
    acc = {"h2d": 0.0, "step": 0.0}
    for i in range(N):
        a = time.perf_counter()
        x = hx[i % 8].to(dev, non_blocking=True); y = hy[i % 8].to(dev, non_blocking=True)
        b = time.perf_counter()
        t.train_step(x, y)
        c = time.perf_counter()
        acc["h2d"] += b - a; acc["step"] += c - b
    return acc


timed("sync every step      ", sync_loop)
timed("no read-back at all  ", nosync_loop)
timed("late read, same-stream copy", lambda: piped_loop(False))
timed("late read + prefetch stream", lambda: piped_loop(True))

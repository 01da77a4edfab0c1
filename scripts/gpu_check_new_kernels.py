"""Timings of the kernels added in the host-only session (run on a B200; one JSON line each):

* dtf_cifar_distort — device-resident CIFAR-10 batches: graph-replayed kernel time at batch 128 / 512, the fraction of
  the measured HBM copy rate its bytes amount to, and the end-to-end rate of DeviceCifar10Batches through the
  prefetcher next to the host pipeline's;
* dtf_fake_quant_fwd / _bwd and dtf_quant_range_update on MobileNet-sized activations (streams: fraction of the copy
  rate);
* one quantised MobileNet-v1 training step vs the float step (graph replay).

    python scripts/gpu_check_new_kernels.py > gpurun_out/new_kernels_perf.jsonl
"""
import json
import os
import sys
import tempfile
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_tensorflow_models_b200.data import cifar10, synthetic  # noqa: E402
from distributed_tensorflow_models_b200.ops import functional as Fn  # noqa: E402
from distributed_tensorflow_models_b200.utils.pipeline import DevicePrefetcher  # noqa: E402


def copy_rate(dev):
    peaks = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    try:
        with open(peaks) as f:
            d = json.load(f)
        for k in ("copy_GBps", "copy_bandwidth_GBps", "hbm_copy_GBps"):
            if k in d:
                return float(d[k]) * 1e9
    except Exception:
        pass
    a = torch.empty(1 << 28, dtype=torch.float32, device=dev)      # 1 GiB: larger than L2
    b = torch.empty_like(a)
    for _ in range(3):
        b.copy_(a)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(10):
        b.copy_(a)
    e1.record()
    torch.cuda.synchronize()
    return 2 * a.numel() * 4 * 10 / (e0.elapsed_time(e1) * 1e-3)


def graph_time(fn, iters=50):
    """us per call of ``fn`` replayed from a CUDA graph (device events, after warm-up)."""
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3):
            fn()
    torch.cuda.current_stream().wait_stream(s)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    for _ in range(5):
        g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters


def main():
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    rate = copy_rate(dev)
    print(json.dumps({"hbm_copy_GBps": round(rate / 1e9, 1)}), flush=True)

    # ---- CIFAR records resident in HBM --------------------------------------------------------------------------
    with tempfile.TemporaryDirectory() as tmp:
        synthetic.write_cifar10_binary(tmp, train_per_file=10000, test=100)
        for size in (24, 32):
            src = cifar10.DeviceCifar10Batches(tmp, 128, size, device=dev)
            for b in (128, 512):
                idx = torch.randperm(src.n)[:b]
                item = cifar10.RecordBatch(src, idx, *cifar10.draw_distortion(b, size, torch.Generator().manual_seed(1)))
                stage = item.stage.to(dev)
                out = torch.empty(item.shape, device=dev)
                item.stage = stage                      # time the kernel alone: staging already on the device

                def run(item=item, out=out):
                    item.copy_into(out)
                us = graph_time(run)
                nbytes = b * (size * size * 3 * 2 + size * size * 3 * 4)     # bytes read twice + floats written
                print(json.dumps({"kernel": "dtf_cifar_distort", "batch": b, "size": size, "us": round(us, 2),
                                  "fraction_of_copy_rate": round(nbytes / (us * 1e-6) / rate, 3),
                                  "img_per_s": round(b / (us * 1e-6))}), flush=True)
            src.close()
            for name, make in (("device", lambda: cifar10.DeviceCifar10Batches(tmp, 128, size, device=dev)),
                               ("host", lambda: cifar10.Cifar10Batches(tmp, 128, size))):
                it = make()
                pf = DevicePrefetcher(it, dev)
                for _ in range(20):
                    next(pf)
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                n = 400
                for _ in range(n):
                    x, y = next(pf)
                torch.cuda.synchronize()
                dt = time.perf_counter() - t0
                it.close()
                print(json.dumps({"pipeline": "cifar10 " + name, "size": size, "batch": 128,
                                  "img_per_s_through_prefetcher": round(n * 128 / dt)}), flush=True)

    # ---- fake quantisation --------------------------------------------------------------------------------------
    for shape in ((64, 112, 112, 32), (64, 14, 14, 512)):
        x = torch.randn(*shape, device=dev) * 3
        dy = torch.randn_like(x)
        rng = torch.tensor([-6.0, 6.0], device=dev)
        gate = torch.ones((), device=dev)
        mm = torch.tensor([float("inf"), float("-inf")], device=dev)
        xr = x.clone().requires_grad_(True)
        y = Fn.fake_quant(xr, rng, gate, 8, False)
        n = x.numel()
        for name, fn, streams in (
                ("dtf_fake_quant_fwd", lambda: Fn.fake_quant(x, rng, gate, 8, False), 2),
                ("dtf_fake_quant_bwd", lambda: torch.autograd.grad(y, xr, dy, retain_graph=True), 3),
                ("dtf_quant_range_update", lambda: Fn.quant_range_update(x, rng.clone(), 0.999, mm), 1)):
            us = graph_time(fn)
            print(json.dumps({"kernel": name, "shape": list(shape), "us": round(us, 2),
                              "fraction_of_copy_rate": round(streams * n * 4 / (us * 1e-6) / rate, 3)}), flush=True)

    # ---- quantised vs float MobileNet-v1 step -------------------------------------------------------------------
    from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup
    g = torch.Generator().manual_seed(0)
    x = torch.randn(64, 224, 224, 3, generator=g).to(dev)
    yl = torch.randint(0, 1001, (64,), generator=g).to(dev)
    for quantize in (False, True):
        t = Trainer(TrainSetup(model="mobilenet_v1", batch_size=64, cuda_graph=True, quantize=quantize), device=dev)
        for _ in range(5):
            t.train_step(x, yl)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(20):
            t.train_step(x, yl)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        print(json.dumps({"model": "mobilenet_v1", "batch": 64, "quantize": quantize, "ms_per_step": round(ms, 3),
                          "img_per_s": round(64 / (ms * 1e-3)), "native_launches_per_step": t.graph_kernel_launches}),
              flush=True)


if __name__ == "__main__":
    main()

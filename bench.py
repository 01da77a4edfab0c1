#!/usr/bin/env python
"""Headline benchmark: training throughput (images/sec) of the flagship config of BASELINE.json —
ResNet-32 (CIFAR-10 shapes, resnet_model.py) BSP, batch 128 per worker — on N B200s of one node.

    python bench.py --gpus 1 --steps 50 --warmup 10
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

One JSON line is printed by rank 0 (contract in the task statement).  ``value`` is the device-timed
whole-job images/sec with inputs resident on the GPU: the K-step region is timed ``repeats`` times back to back
inside one pair of CUDA events so that at least ``--min-seconds`` (1 s) of device time is covered
(``steps_timed = repeats * K``; nvidia-smi is sampled every 20 ms meanwhile).  ``e2e`` repeats the measurement
through the public ``Trainer.train_step`` API with a pinned-host -> device copy of every batch and a
device -> host read of every loss inside the timed region.  ``secondary`` carries the other configs of
BASELINE.json's metric at the same N, each measured exactly like the headline: Inception-v3 SSP (bound 3),
Inception-v3 BSP and ResNet-32 SSP.  Other models/modes as the headline: ``--model inception --mode ssp``.

``--impl reference`` would run the unmodified reference; it is Python-2 / TensorFlow-1.x code that cannot
be installed here, so that arm reports ``unavailable``.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=("ours", "reference"))
    ap.add_argument("--model", default="resnet")
    ap.add_argument("--mode", default="bsp", choices=("bsp", "asp", "ssp"))
    ap.add_argument("--batch", type=int, default=None, help="per-GPU batch (default: reference preset)")
    ap.add_argument("--staleness", type=int, default=3)
    ap.add_argument("--backend", default="symm", choices=("symm", "nccl_rs_ag", "nccl_allreduce"))
    ap.add_argument("--kernels", default="native", choices=("native", "torch"))
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-fuse-bn", action="store_true")
    ap.add_argument("--no-branch-streams", action="store_true",
                    help="run the branches of Inception blocks one after the other instead of on side streams")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--image-size", type=int, default=None, help="model-zoo networks only (default: the network's own)")
    ap.add_argument("--num-classes", type=int, default=None, help="model-zoo networks only (default 1001)")
    ap.add_argument("--precision", default="tf32", choices=("tf32", "tf32x3"))
    ap.add_argument("--no-secondary", action="store_true",
                    help="headline only: skip the other BASELINE.json configs (Inception-v3 SSP / BSP, ResNet SSP)")
    ap.add_argument("--min-seconds", type=float, default=1.0,
                    help="the K-step region is repeated back to back until this much device time has been timed")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi sampler for the timed region (B200_PROFILING.md 'clocks line')."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.proc = None
        self.lines = []
        self.gpu = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "20", "-i", str(self.gpu)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.splitlines():
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1]))
                mx.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(names, parts[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def _protect_stdout():
    """Route everything libraries print to stdout (e.g. NCCL's version banner) to stderr, and return a
    writer for the real stdout so that exactly ONE line — the JSON result — appears there."""
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = sys.stderr

    def emit(line: str) -> None:
        os.write(real, (line + "\n").encode())
    return emit


def measure(args, model, mode, world, rank, local_rank, device, want_e2e=True):
    """One config, measured like the headline: warm-up, device-timed region (>= min-seconds), end-to-end region."""
    import torch
    import torch.distributed as dist

    from distributed_tensorflow_models_b200 import _native as nat
    from distributed_tensorflow_models_b200.models.registry import MODEL_SPECS
    from distributed_tensorflow_models_b200.parallel.comm import LocalComm, TorchDistComm
    from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup

    if world > 1:
        if args.backend == "symm":
            from distributed_tensorflow_models_b200.parallel.symm import SymmComm
            comm = SymmComm()
        else:
            comm = TorchDistComm("rs_ag" if args.backend == "nccl_rs_ag" else "allreduce")
    else:
        comm = LocalComm()

    if model in MODEL_SPECS:
        spec = MODEL_SPECS[model]
    else:       # any other network of the model-zoo registry (not a BASELINE.json config; for the zoo sweeps)
        from distributed_tensorflow_models_b200.models.registry import spec_for
        spec = spec_for(model, args.image_size, args.num_classes)
    setup = TrainSetup(model=model, mode=mode, num_workers=world, staleness=args.staleness,
                       cuda_graph=not args.no_graph, fuse_bn=not args.no_fuse_bn, seed=0,
                       branch_streams=not args.no_branch_streams, image_size=args.image_size,
                       num_classes=args.num_classes, precision=args.precision)
    setup.batch_size = args.batch if args.batch else spec.batch_size
    trainer = Trainer(setup, comm, device)
    bs = setup.batch_size

    g = torch.Generator().manual_seed(1234 + rank)
    n_host_batches = 8
    host_x = [torch.randn(bs, spec.image_size, spec.image_size, 3, generator=g).pin_memory() for _ in range(n_host_batches)]
    host_y = [torch.randint(0, spec.num_classes, (bs,), generator=g).pin_memory() for _ in range(n_host_batches)]
    dev_x = [t.to(device) for t in host_x[:2]]
    dev_y = [t.to(device) for t in host_y[:2]]

    def sync_all():
        torch.cuda.synchronize(device)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(device)

    # ---- warm-up (also captures the CUDA graph) -----------------------------------------------------
    warm = max(3, args.warmup)
    for i in range(warm):
        trainer.train_step(dev_x[i % 2], dev_y[i % 2])
    sync_all()
    # a short probe decides how often the K-step region is repeated to cover --min-seconds of device time
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    p0.record()
    for i in range(3):
        trainer.train_step(dev_x[i % 2], dev_y[i % 2])
    p1.record()
    sync_all()
    probe_ms = comm.max_over_ranks(p0.elapsed_time(p1) / 3.0)
    repeats = max(1, int(-(-args.min_seconds * 1e3 // max(probe_ms * args.steps, 1e-3))))
    steps_timed = repeats * args.steps

    # ---- device-timed region -----------------------------------------------------------------------
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    nat.reset_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    e0.record()
    for i in range(steps_timed):
        out = trainer.train_step(dev_x[i % 2], dev_y[i % 2])
    e1.record()
    sync_all()
    ms = e0.elapsed_time(e1)
    launches = nat.launch_count() + trainer.graph_kernel_launches * steps_timed
    clocks = sampler.stop() if rank == 0 else None
    ms = comm.max_over_ranks(ms)
    last_loss = float(out["loss"])

    # ---- end-to-end region: pinned host -> device copy + loss read-back every step -------------------
    e2e = None
    if want_e2e and not args.no_e2e:
        sync_all()
        t0 = torch.cuda.Event(enable_timing=True)
        t1 = torch.cuda.Event(enable_timing=True)
        from distributed_tensorflow_models_b200.utils.pipeline import DevicePrefetcher, ScalarReadback
        # one-time set-up (a copy stream, four pinned 4-byte slots) happens before the clock starts, as it does at
        # the start of a training run; no copy is issued until the first next() inside the timed region
        feed = DevicePrefetcher(((host_x[i % n_host_batches], host_y[i % n_host_batches]) for i in range(steps_timed)),
                                device, lazy=True)
        reader = ScalarReadback(device)
        sync_all()
        t0.record()
        # the trainers' own loop (trainers/common.py): every step's inputs are copied from pinned host memory and
        # every step's loss is read back, both inside the timed region; the copy of step i+1 is issued while step i
        # computes and the loss of step i is read after step i+1 has been launched
        read = []
        for i in range(steps_timed):
            x, y = next(feed)
            o = trainer.train_step(x, y)
            reader.push(o["loss"])
            if len(reader) > 1:
                read.append(reader.pop()[0])
        while len(reader):
            read.append(reader.pop()[0])              # D2H read of every step's result
        assert len(read) == steps_timed and all(v == v for v in read)
        t1.record()
        sync_all()
        ems = comm.max_over_ranks(t0.elapsed_time(t1))
        e2e = {"value": world * bs * steps_timed / (ems / 1e3), "unit": "images/sec",
               "h2d_bytes_per_step": host_x[0].numel() * 4 + host_y[0].numel() * 8, "d2h_bytes_per_step": 4,
               "ms_per_step": ems / steps_timed,
               "pipeline": "every step's batch is copied from pinned host memory (copy of step i+1 overlaps step i) "
                           "and every step's loss is read back (after step i+1 has been queued)"}
    if hasattr(comm, "check_status"):
        comm.check_status()
    if world > 1:
        comm.finish()
    res = {
        "metric": f"images/sec, {model} ({'CIFAR-10' if spec.dataset == 'cifar10' else 'ImageNet'} shapes) {mode.upper()} training",
        "value": world * bs * steps_timed / (ms / 1e3), "unit": "images/sec", "n_gpus": world, "steps": args.steps,
        "repeats": repeats, "steps_timed": steps_timed, "warmup": warm, "ms_per_step": ms / steps_timed,
        "timed_region_ms": ms,
        "config": {"model": f"resnet{setup.resnet_size}" if model == "resnet" else model,
                   "global_batch": world * bs, "per_gpu_batch": bs, "image": spec.image_size,
                   "parallelism": f"ps-dp{world} ({args.backend if world > 1 else 'single rank'})",
                   "mode": mode, "kernels": args.kernels, "cuda_graph": not args.no_graph,
                   "l2": "no explicit flush: each step streams > 126 MB of activations/gradients through HBM"},
        "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "final_loss": last_loss,
    }
    if mode == "ssp":
        res["config"]["staleness"] = args.staleness
    if world > 1 and args.backend == "symm" and getattr(comm, "bucketed", False):
        res["config"]["push"] = {"buckets": len(comm.buckets), "overlap_with_backward": bool(comm.overlap),
                                 "nvls": bool(comm.mc_params), "pull_gate": comm.gate_mode}
    del trainer, comm
    import gc
    gc.collect()
    torch.cuda.empty_cache()
    return res


def main():
    args = parse()
    emit = _protect_stdout()
    if args.impl == "reference":
        emit(json.dumps({"impl": "reference", "unavailable":
                         "reference is Python-2/TensorFlow-1.x (tf.contrib.slim, Supervisor, SyncReplicasOptimizer, "
                         "thrift) with no setup.py; it cannot be installed or run offline on this image (DESIGN.md)"}))
        return 0

    import torch
    import torch.distributed as dist

    from distributed_tensorflow_models_b200.ops import functional as Fn

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        args.gpus = world
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    Fn.set_backend("auto" if args.kernels == "native" else "torch")
    torch.backends.cudnn.allow_tf32 = True
    torch.backends.cuda.matmul.allow_tf32 = True
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)

    line = measure(args, args.model, args.mode, world, rank, local_rank, device)
    line.update({"higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                 "dtype": "fp32 storage/accumulate, %s tensor-core multiply (reference is fp32; no bf16/fp8)" % args.precision,
                 "data": "synthetic (random images of the dataset's shape, random-init weights)", "impl": "ours"})
    secondary = []
    if not args.no_secondary and args.model == "resnet" and args.mode == "bsp":
        # the other half of BASELINE.json's metric ("ResNet CIFAR-10 & Inception-v3 in BSP/SSP modes") at the same N
        for model, mode in (("inception", "ssp"), ("inception", "bsp"), ("resnet", "ssp")):
            try:
                secondary.append(measure(args, model, mode, world, rank, local_rank, device))
            except Exception as e:  # noqa: BLE001 — the headline must still be printed
                secondary.append({"metric": f"{model} {mode}", "error": repr(e)[:300]})
                break
    line["secondary"] = secondary
    if rank == 0:
        emit(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())

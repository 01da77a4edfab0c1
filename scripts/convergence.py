"""Convergence parity at equal steps (BASELINE.json: "... with convergence parity at equal steps"; the reference is
fp32 end to end, inception/slim/variables.py:249): the same model, initial values, fixed synthetic training set and
hyper-parameters are trained by

    native   our sm_100a kernels: fp32 storage / accumulation, TF32 tensor-core multiplies (the product path)
    fp32     plain PyTorch ops with TF32 switched OFF (cuDNN / cuBLAS in full fp32) — the reference's arithmetic

and the loss curves are written to profiles/convergence_<model>[_n<N>].json together with the gap of the smoothed
losses at a few steps.  Under torchrun (N ranks) the arms are instead the product transport (SymmComm, bucketed
NVLink push) and the NCCL reduce-scatter / all-gather baseline, both on the native kernels, at equal global steps.

    python scripts/convergence.py resnet 2000
    python scripts/convergence.py inception 300
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 ... scripts/convergence.py resnet 1000
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("DTF_PS_TIMEOUT_S", "60")
import torch
import torch.distributed as dist

from distributed_tensorflow_models_b200.models.registry import MODEL_SPECS
from distributed_tensorflow_models_b200.ops import functional as Fn
from distributed_tensorflow_models_b200.parallel.comm import LocalComm, TorchDistComm
from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup


def smooth(xs, k):
    out, acc = [], 0.0
    for i, v in enumerate(xs):
        acc += v
        if i >= k:
            acc -= xs[i - k]
        out.append(acc / min(i + 1, k))
    return out


def main():
    model = sys.argv[1] if len(sys.argv) > 1 else "resnet"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
    n_batches = int(sys.argv[3]) if len(sys.argv) > 3 else 32
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    spec = MODEL_SPECS[model]
    bs = spec.batch_size
    # a fixed synthetic training set with learnable labels: the class is a function of the image
    g = torch.Generator().manual_seed(4321 + rank)
    xs = [torch.randn(bs, spec.image_size, spec.image_size, 3, generator=g) for _ in range(n_batches)]
    proj = torch.randn(3, spec.num_classes if spec.num_classes <= 16 else 16, generator=torch.Generator().manual_seed(7))
    ys = [(x.mean(dim=(1, 2)) @ proj).argmax(1) for x in xs]
    xs = [x.to(dev) for x in xs]
    ys = [y.to(dev) for y in ys]

    if world == 1:
        arms = [("native_tf32", "auto", lambda: LocalComm()), ("torch_fp32", "torch", lambda: LocalComm()),
                ("native_tf32x3", "auto", lambda: LocalComm())]
    else:
        from distributed_tensorflow_models_b200.parallel.symm import SymmComm
        arms = [("symm_bucketed_push", "auto", lambda: SymmComm()), ("nccl_rs_ag", "auto", lambda: TorchDistComm("rs_ag"))]
    curves = {}
    for name, backend, mk_comm in arms:
        Fn.set_backend(backend)
        torch.backends.cudnn.allow_tf32 = False
        torch.backends.cuda.matmul.allow_tf32 = False
        comm = mk_comm()
        t = Trainer(TrainSetup(model=model, mode="bsp", num_workers=world, batch_size=bs, seed=0,
                               cuda_graph=(backend == "auto"),
                               precision="tf32x3" if name.endswith("x3") else "tf32"), comm, dev)
        losses = []
        pending = []
        for i in range(steps):
            out = t.train_step(xs[i % n_batches], ys[i % n_batches])
            pending.append(out["loss"].clone() if backend == "auto" else out["loss"])
            if len(pending) >= 50:
                losses += [float(v) for v in torch.stack(pending).tolist()]
                pending = []
        losses += [float(v) for v in torch.stack(pending).tolist()] if pending else []
        torch.cuda.synchronize()
        if world > 1:
            lt = torch.tensor(losses, device=dev, dtype=torch.float64)
            dist.all_reduce(lt)
            losses = (lt / world).tolist()
            comm.finish()
        curves[name] = losses
        del t, comm
        torch.cuda.empty_cache()
    Fn.set_backend("auto")
    names = list(curves)
    ref = "torch_fp32" if "torch_fp32" in curves else names[1]
    k = max(10, steps // 40)
    sm = {n: smooth(c, k) for n, c in curves.items()}
    marks = sorted({min(steps - 1, int(steps * f)) for f in (0.05, 0.1, 0.25, 0.5, 0.75, 1.0)})
    report = {"model": model, "steps": steps, "world": world, "batch_per_gpu": bs, "distinct_batches": n_batches,
              "smoothing_window": k, "arms": names, "reference_arm": ref,
              "smoothed_loss_at": {str(m): {n: sm[n][m] for n in names} for m in marks},
              "first_loss": {n: curves[n][0] for n in names},
              "max_rel_gap_of_smoothed_losses_vs_reference": {
                  n: max(abs(p - q) / max(abs(q), 1e-2) for p, q in zip(sm[n][k:], sm[ref][k:])) for n in names if n != ref},
              "curves_every_10": {n: c[::10] for n, c in curves.items()}}
    if rank == 0:
        os.makedirs("gpurun_out", exist_ok=True)
        path = f"gpurun_out/convergence_{model}{'_n%d' % world if world > 1 else ''}.json"
        with open(path, "w") as f:
            json.dump(report, f)
        print(json.dumps({k_: v for k_, v in report.items() if k_ != "curves_every_10"}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

"""Multi-GPU validation of the product transport (SymmComm: bucketed in-kernel NVLink push overlapped with the
backward pass, pull gates in the consuming convolutions, one-sided ASP, SSP clocks) against the NCCL
reduce-scatter/all-gather baseline.  Run under torchrun, e.g.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 scripts/multi_gpu_check.py
Prints one JSON line from rank 0; exit code 0 iff every check passed."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("DTF_PS_TIMEOUT_S", "30")          # a protocol bug must fail this script, not hang the box
import torch
import torch.distributed as dist

from distributed_tensorflow_models_b200 import config as cfg
from distributed_tensorflow_models_b200.models.registry import MODEL_SPECS
from distributed_tensorflow_models_b200.parallel.comm import TorchDistComm
from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup
from distributed_tensorflow_models_b200.parallel.symm import SymmComm


def main():
    rank, world, lr_ = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr_)
    dev = torch.device("cuda", lr_)
    dist.init_process_group("nccl", device_id=dev)
    res = {"world": world}
    ok = True

    def data(model, bs, seed):
        sp = MODEL_SPECS[model]
        g = torch.Generator().manual_seed(seed)
        return (torch.randn(bs, sp.image_size, sp.image_size, 3, generator=g).to(dev),
                torch.randint(0, sp.num_classes, (bs,), generator=g).to(dev))

    def replicas_identical(p):
        chk = torch.stack([p.double().sum(), p.double().abs().sum()])
        gathered = [torch.zeros_like(chk) for _ in range(world)]
        dist.all_gather(gathered, chk)
        return all(torch.equal(g, gathered[0]) for g in gathered)

    def rel(a, b):
        return float((a - b).norm() / b.norm().clamp_min(1e-30))

    # ---- BSP: bucketed overlapped push + gated pull vs NCCL RS+AG, same seed / data ---------------------------
    rms = cfg.PRESETS["cnn"].__class__(**{**cfg.PRESETS["cnn"].__dict__, "optimizer": "rmsprop", "initial_lr": 0.01})
    cases = [
        # name, setup factory, data model, strict (the run is reproducible to rounding), SymmComm kwargs
        ("cnn_sgd_ema_graph", lambda: TrainSetup(model="cnn", batch_size=8, num_workers=world, cuda_graph=True),
         "cnn", True, dict(bucket_mb=0.5)),
        ("cnn_rmsprop_eager", lambda: TrainSetup(model="cnn", batch_size=8, num_workers=world, cuda_graph=False,
                                                 preset=rms), "cnn", True, dict(bucket_mb=0.5)),
        ("cnn_rmsprop_no_overlap", lambda: TrainSetup(model="cnn", batch_size=8, num_workers=world, cuda_graph=False,
                                                      preset=rms), "cnn", True, dict(overlap=False)),
        ("resnet_sgd_wd_graph", lambda: TrainSetup(model="resnet", batch_size=8, num_workers=world, cuda_graph=True),
         "resnet", False, dict(bucket_mb=0.25)),
    ]
    for name, mk, dm, strict, kw in cases:
        x, y = data(dm, 8, 100 + rank)
        ta = Trainer(mk(), SymmComm(**kw), dev)
        tb = Trainer(mk(), TorchDistComm("rs_ag"), dev)
        la, lb = [], []
        for _ in range(6):
            la.append(float(ta.train_step(x, y)["loss"]))
            lb.append(float(tb.train_step(x, y)["loss"]))
        ta.comm.sync_replica()
        torch.cuda.synchronize()
        ta.comm.check_status()
        used = ta.layout.used
        pa, pb = ta.layout.params[:used], tb.layout.params[:used]
        r = {"buckets": len(ta.comm.buckets), "order": ta.comm._order, "loss_symm": la, "loss_nccl": lb,
             "param_rel_diff": rel(pa, pb), "replicas_identical": replicas_identical(pa),
             "pushes_in_graph": ta.graph_kernel_launches}
        sa, sb = ta.comm.gather_slot_flats(), tb.comm.gather_slot_flats()
        for k in sa:
            r["slot_rel_diff_" + k] = rel(sa[k][:used], sb[k][:used])
        # the plain CNN is reproducible to rounding, so the transports must agree almost exactly; ResNet's tf32 +
        # float-atomic noise is amplified by 31 BN/ReLU layers (two runs of the SAME transport differ as much)
        # (1e-4: the switch and NCCL's ring add the N gradients in different orders; observed 1e-12 .. 1e-6 at N = 2)
        tol = 1e-4 if strict else 0.1
        good = r["replicas_identical"] and r["param_rel_diff"] < tol and abs(la[0] - lb[0]) / abs(lb[0]) < 1e-3 \
            and abs(la[-1] - lb[-1]) / abs(lb[-1]) < (1e-3 if strict else 0.15)
        if strict:
            good = good and all(v < 1e-4 for k, v in r.items() if k.startswith("slot_rel_diff_"))
        r["ok"] = bool(good)
        res["bsp_" + name] = r
        ok &= good
        del ta, tb
        torch.cuda.empty_cache()

    # ---- ASP: one global step per worker step, replicas stay finite and close -------------------------------
    for name, preset in (("asp", None), ("asp_rmsprop", rms)):
        t = Trainer(TrainSetup(model="cnn", mode="asp", batch_size=8, num_workers=world, cuda_graph=False,
                               preset=preset), SymmComm(), dev)
        x, y = data("cnn", 8, 7 + rank)
        losses = []
        for _ in range(5):
            losses.append(float(t.train_step(x, y)["loss"]))
        torch.cuda.synchronize()
        dist.barrier()
        gs = int(t.comm.gs_view.item()) if rank == 0 else 0
        gst = torch.tensor([gs], device=dev)
        dist.broadcast(gst, 0)
        finite = bool(torch.isfinite(t.layout.params).all())
        res[name] = {"global_step_on_rank0": int(gst.item()), "expected": 5 * world, "finite": finite, "loss": losses}
        ok &= int(gst.item()) == 5 * world and finite and all(v == v for v in losses)
        del t

    # ---- SSP: a straggler never lets anybody run more than `bound` ahead --------------------------------------
    bound = 1
    t = Trainer(TrainSetup(model="cnn", mode="ssp", staleness=bound, batch_size=8, num_workers=world, cuda_graph=False),
                SymmComm(), dev)
    leads = []
    for s in range(6):
        t.train_step(x, y)
        torch.cuda.synchronize()
        clocks = t.comm.clocks_view.tolist()
        leads.append(t.local_step - min(clocks))
        if rank == world - 1:
            time.sleep(0.05)
    t.comm.check_status()
    mx = torch.tensor([max(leads)], device=dev)
    dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    res["ssp"] = {"bound": bound, "max_lead_after_release": int(mx.item())}
    ok &= int(mx.item()) <= bound
    dist.barrier()
    res["ok"] = bool(ok)
    if rank == 0:
        print(json.dumps(res))
    dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())

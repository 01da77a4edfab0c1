"""Kernel timeline of one training step (diagnosis only — never a benchmark number): runs a few steps under
torch.profiler on every rank and writes, for rank 0, one line per kernel of the last step
(start us, duration us, stream, name) to gpurun_out/trace_<model>_<mode>_n<N>.txt.
    python scripts/trace_step.py vgg bsp            (1 GPU)
    python -m torch.distributed.run ... scripts/trace_step.py vgg bsp   (N GPUs)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("DTF_PS_TIMEOUT_S", "60")
import torch
import torch.distributed as dist
from torch.profiler import ProfilerActivity, profile

from distributed_tensorflow_models_b200.models.registry import MODEL_SPECS
from distributed_tensorflow_models_b200.parallel.comm import LocalComm
from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup

model = sys.argv[1] if len(sys.argv) > 1 else "vgg"
mode = sys.argv[2] if len(sys.argv) > 2 else "bsp"
graph = (sys.argv[3] != "eager") if len(sys.argv) > 3 else True
world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
    from distributed_tensorflow_models_b200.parallel.symm import SymmComm
    comm = SymmComm()
else:
    comm = LocalComm()
spec = MODEL_SPECS[model]
t = Trainer(TrainSetup(model=model, mode=mode, num_workers=world, batch_size=spec.batch_size, cuda_graph=graph, seed=0),
            comm, dev)
g = torch.Generator().manual_seed(rank)
x = torch.randn(spec.batch_size, spec.image_size, spec.image_size, 3, generator=g).to(dev)
y = torch.randint(0, spec.num_classes, (spec.batch_size,), generator=g).to(dev)
for _ in range(6):
    t.train_step(x, y)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(4):
        t.train_step(x, y)
    torch.cuda.synchronize()
if world > 1:
    dist.barrier()
if rank == 0:
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    evs.sort(key=lambda e: e.time_range.start)
    t0 = evs[0].time_range.start
    os.makedirs("gpurun_out", exist_ok=True)
    path = f"gpurun_out/trace_{model}_{mode}_n{world}.txt"
    with open(path, "w") as f:
        for e in evs:
            f.write("%10.1f %9.1f  s%-3s %s\n" % (e.time_range.start - t0, e.time_range.end - e.time_range.start,
                                                 getattr(e, "device_resource_id", "?"), e.name[:90]))
    print("wrote", path, len(evs), "kernels")
if world > 1:
    comm.finish()
    dist.destroy_process_group()

"""Time the BSP push alone (torchrun, N GPUs): every bucket's dtf_ps_bucket_step back to back on the push stream plus
the gate on all `updated` flags — ms per whole-model push (max over ranks) and the fraction of the NVLink roofline it
reaches.  Environment: DTF_BUCKET_MB, DTF_PUSH_SMS (SMs the push runs on), DTF_NVLS=0.  Roofline bytes per GPU and direction: every rank ships (N-1)/N of its
gradient buffer to the owners (push) and, without NVLS, receives (N-1)/N of the parameters back (pull); with NVLS
the pull is one multicast store of the rank's own 1/N.  NVLink 5: 900 GB/s per direction per GPU."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup
from distributed_tensorflow_models_b200.parallel.symm import SymmComm

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
out = {}
for model in sys.argv[1:] or ["vgg", "alexnet", "inception", "resnet"]:
    comm = SymmComm()
    t = Trainer(TrainSetup(model=model, mode="bsp", num_workers=world, cuda_graph=False, seed=0), comm, dev)
    t.layout.grads.normal_()
    for _ in range(3):
        comm.bsp_step(0.01)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    iters = 10
    e0.record()
    for _ in range(iters):
        comm.bsp_step(0.01)
    e1.record()
    torch.cuda.synchronize()
    ms = comm.max_over_ranks(e0.elapsed_time(e1) / iters)
    P = t.layout.total * 4
    nvls = bool(comm.mc_grads)
    out_bytes = P * (world - 1) / world + (P / world if nvls else 0)          # egress per GPU
    in_bytes = (P / world if nvls else P * (world - 1) / world) + (P * (world - 1) / world)   # ingress per GPU
    roof_ms = max(out_bytes, in_bytes) / 900e9 * 1e3
    out[model] = {"param_MB": round(P / 1e6, 1), "buckets": len(comm.buckets), "ms": round(ms, 4), "nvls": nvls,
                  "roofline_ms": round(roof_ms, 4),
                  "fraction_of_nvlink_roofline": round(roof_ms / ms, 3)}
    comm.finish()
    del t, comm
    torch.cuda.empty_cache()
if rank == 0:
    print(json.dumps({"world": world, "push_sms": os.environ.get("DTF_PUSH_SMS", "default"), "bsp_push": out}))
dist.destroy_process_group()

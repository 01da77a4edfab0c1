"""Litmus test of the ordering the pull gate relies on: an owner's stores of the updated values (multimem.st through
the NVSwitch, or P2P st) followed by __threadfence_system + st.release.sys of the `updated` flag, against a reader
that sees the flag with ld.acquire.sys and then reads the values.  If the flag could overtake the data, a reader
would find a value of the previous epoch.

Every rank's gradient is the constant 1, lr = 1, plain SGD: after step k every parameter is exactly -k.  Each step is
the real push (dtf_ps_bucket_step per bucket) followed on the same stream by the gate (dtf_ps_gate) and, without any
host synchronisation in between, a kernel that counts the elements that differ from -k.  Run under torchrun:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 scripts/litmus_push_order.py [steps] [model]
Prints one JSON line from rank 0; exit code 0 iff no rank ever saw a stale value."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("DTF_PS_TIMEOUT_S", "30")
import torch
import torch.distributed as dist

from distributed_tensorflow_models_b200.parallel.engine import Trainer, TrainSetup
from distributed_tensorflow_models_b200.parallel.symm import SymmComm


def main():
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 300
    model = sys.argv[2] if len(sys.argv) > 2 else "alexnet"
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    out = {"world": world, "steps": steps, "model": model}
    ok = True
    for nvls in ("1", "0"):
        os.environ["DTF_NVLS"] = nvls
        comm = SymmComm(bucket_mb=8)
        t = Trainer(TrainSetup(model=model, mode="bsp", num_workers=world, cuda_graph=False, seed=0), comm, dev)
        comm.cfg.weight_decay = 0.0
        comm.cfg.ema = None
        comm.opt.ema = None
        L = t.layout
        L.params.zero_()
        L.grads.fill_(1.0)
        torch.cuda.synchronize()
        dist.barrier()
        bad = torch.zeros(1, dtype=torch.int64, device=dev)
        for k in range(1, steps + 1):
            comm.bsp_step(1.0)                      # push every bucket, then gate on all of them (stand-alone path)
            bad += (L.params != float(-k)).sum()    # same stream, no host sync: reads right behind the gate
            if rank == k % world:                   # skew the ranks against one another
                torch.cuda._sleep(200000)
        torch.cuda.synchronize()
        comm.check_status()
        dist.all_reduce(bad)
        out["nvls" if nvls == "1" and comm.mc_params else ("p2p" if nvls == "0" else "nvls_unavailable")] = {
            "buckets": len(comm.buckets), "floats": int(L.total), "stale_elements_seen": int(bad.item())}
        ok &= int(bad.item()) == 0
        comm.finish()
        del t, comm
        torch.cuda.empty_cache()
    out["ok"] = bool(ok)
    if rank == 0:
        print(json.dumps(out))
    dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())

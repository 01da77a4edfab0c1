"""Does torch's symmetric memory hand out an NVLS multicast address on this box?  (torchrun, >= 2 GPUs)"""
import os
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
buf = symm_mem.empty(1 << 20, dtype=torch.float32, device=torch.device("cuda", local))
hdl = symm_mem.rendezvous(buf, dist.group.WORLD)
mc = getattr(hdl, "multicast_ptr", None)
print(f"rank {rank}: multicast_ptr = {mc}, buffer_ptrs = {[hex(int(p)) for p in hdl.buffer_ptrs]}", flush=True)
dist.barrier()
dist.destroy_process_group()

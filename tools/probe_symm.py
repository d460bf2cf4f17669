"""Probe: does torch symmetric memory give us an NVLS multicast pointer on this box? (run with torchrun, 2+ GPUs)"""
import os
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
try:
    t = symm_mem.empty(1 << 20, dtype=torch.float32, device=dev)
    h = symm_mem.rendezvous(t, group=dist.group.WORLD)
    mc = getattr(h, "multicast_ptr", None)
    print(f"rank {rank}: buffer_ptrs={[hex(p) for p in h.buffer_ptrs][:3]} multicast_ptr={hex(mc) if mc else mc} signal_pad={hex(h.signal_pad_ptrs[0])}", flush=True)
    t.fill_(rank + 1)
    dist.barrier()
    torch.cuda.synchronize()
    if mc:
        print(f"rank {rank}: multicast supported", flush=True)
except Exception as e:
    print(f"rank {rank}: symmetric memory failed: {type(e).__name__}: {e}", flush=True)
dist.barrier()
dist.destroy_process_group()

"""Probe: does this box expose NVLS multicast through torch symmetric memory?  (torchrun, 2+ ranks)"""
import os, torch, torch.distributed as dist
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
import torch.distributed._symmetric_memory as symm
print(rank, "symm api:", [n for n in dir(symm) if not n.startswith("_")][:40], flush=True)
try:
    t = symm.empty(1 << 20, dtype=torch.float32, device=torch.device("cuda", rank))
    h = symm.rendezvous(t, dist.group.WORLD)
    print(rank, "buffer_ptrs", [hex(p) for p in h.buffer_ptrs], "multicast_ptr", hex(h.multicast_ptr) if h.multicast_ptr else h.multicast_ptr,
          "signal_pad_ptrs", [hex(p) for p in h.signal_pad_ptrs][:2], "buffer_size", h.buffer_size, flush=True)
    t.fill_(rank + 1)
    dist.barrier()
    if h.multicast_ptr:
        # one_shot_all_reduce uses multimem when available
        out = torch.ops.symm_mem.multimem_all_reduce_(t, "sum", dist.group.WORLD.group_name)
        torch.cuda.synchronize()
        print(rank, "multimem_all_reduce ->", float(t[0]), flush=True)
except Exception as e:
    print(rank, "symm failed:", repr(e), flush=True)
dist.barrier()
dist.destroy_process_group()

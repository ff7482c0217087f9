"""globaltimer stamps of block 0 of the bucket-0 sync kernel at world = 1 (cold L2)."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl_b200.config import Config
from ddl_b200.ops import native
from ddl_b200.parallel.trainer import SPMDTrainer
L = native.lib()
L.ddl_comm_stamps.argtypes = [C.c_int, C.c_void_p, C.c_int]
cfg = Config(variant="mnist_sync_sharding_greedy", mode="sync", shard="greedy", num_ps=1, num_workers=1, transport="fused",
             backend="native", synthetic=True, synthetic_train=600, synthetic_test=100, seed=1, eval_every=0, graphs=False,
             log_time=False).validate()
tr = SPMDTrainer(cfg, 0, 1, 0, "auto")
t = tr.transport
tr.grads.flat.normal_()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=tr.device)
L.ddl_comm_stamps(1, None, 0)
names = ["start", "st loaded + signal A", "barrier A passed", "loop done (block 0 thread 0)", "block syncthreads", "fence done",
         "counter atomic returned", "end"]
for cold in (True, False):
    for i in range(3):
        if cold:
            flush.fill_(i)
        t.begin_step()
        torch.cuda.synchronize()
        t.step(all_buckets=True)
        torch.cuda.synchronize()
    buf = (C.c_uint64 * 16)()
    L.ddl_comm_stamps(1, buf, 16)
    print("[%s] " % ("cold" if cold else "warm") + " | ".join("%s +%.2f" % (names[k], (buf[k] - buf[0]) / 1e3) for k in range(1, 8)))

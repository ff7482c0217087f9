"""globaltimer stamps inside the sync push/update/pull kernel of one bucket, taken during graph-replayed training steps.
usage: python tools/sync_phase.py [bucket]            (1 rank)
       torchrun --nproc-per-node N tools/sync_phase.py [bucket]   (N ranks; every rank prints its own stamps)
Stamps of block 0 / thread 0 (us after kernel start): signal A sent, barrier A passed, loop done, block barrier, fence done,
counter atomic returned, block-0 end.  Needs the instrumented library (python -m ddl_b200.ops.build --lab; DDL_LAB=1)."""
import os
os.environ.setdefault("DDL_LAB", "1")
import sys, ctypes as C, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl_b200.config import Config
from ddl_b200.ops import native
from ddl_b200.parallel.trainer import SPMDTrainer

bucket = int(sys.argv[1]) if len(sys.argv) > 1 else 3
rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
L = native.lib()
L.ddl_comm_stamps.argtypes = [C.c_int, C.c_void_p, C.c_int]
cfg = Config(variant="mnist_sync_sharding_greedy", mode="sync", shard="greedy", num_ps=world, num_workers=world, transport="fused",
             backend="native", synthetic=True, synthetic_train=6000, synthetic_test=100, seed=1, eval_every=0, graphs=True,
             log_time=False).validate()
L.ddl_comm_stamps(bucket + 1, None, 0)          # before the graph capture: the flag is a kernel argument
tr = SPMDTrainer(cfg, rank, world, rank, "colocated" if world > 1 else "auto")
nb = tr.data.batches_per_epoch(cfg.batch_size, world, False)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=tr.device)
names = ["start", "signal A sent", "barrier A passed", "loop done", "block barrier", "fence done", "atomic returned", "block-0 end"]
rows = []
for i in range(12):
    flush.fill_(i)
    tr.train_step(i % nb)
    torch.cuda.synchronize()
    buf = (C.c_uint64 * 16)()
    L.ddl_comm_stamps(bucket + 1, buf, 16)
    rows.append([(buf[k] - buf[0]) / 1e3 if buf[k] >= buf[0] else float("nan") for k in range(8)])
med = [statistics.median(r[k] for r in rows[4:]) for k in range(8)]
print("[rank %d/%d bucket %d] " % (rank, world, bucket) + " | ".join("%s +%.2f" % (names[k], med[k]) for k in range(1, 8)), flush=True)
tr.close()

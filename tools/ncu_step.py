"""One eagerly launched training step (sync, then async) between cudaProfilerStart / Stop, for
    ncu --set full --profile-from-start off ... python tools/ncu_step.py
(every kernel of the step is a separate launch: graphs off).  Prints nothing but a final OK line."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl_b200.config import Config
from ddl_b200.parallel.trainer import SPMDTrainer

for mode, variant in (("sync", "mnist_sync_sharding_greedy"), ("async", "mnist_async")):
    cfg = Config(variant=variant, mode=mode, shard="greedy" if mode == "sync" else "none", num_ps=1, num_workers=1,
                 transport="fused", backend="native", synthetic=True, synthetic_train=1000, synthetic_test=2000, seed=3,
                 eval_every=0, graphs=False, log_time=False).validate()
    tr = SPMDTrainer(cfg, 0, 1, 0)
    for i in range(3):
        tr.train_step(i)
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    tr.train_step(3)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    tr.transport.close()
print("OK")

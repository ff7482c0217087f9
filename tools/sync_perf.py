"""Stand-alone time of the fused push/update/pull kernels at world = 1 (the flagship arena, both buckets), cold L2,
for the grid cap / unroll knobs (DDL_COMM_CAP, DDL_COMM_UNROLL).  usage: python tools/sync_perf.py"""
import os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def child():
    import torch
    from ddl_b200.config import Config
    from ddl_b200.parallel.trainer import SPMDTrainer
    cfg = Config(variant="mnist_sync_sharding_greedy", mode="sync", shard="greedy", num_ps=1, num_workers=1, transport="fused",
                 backend="native", synthetic=True, synthetic_train=600, synthetic_test=100, seed=1, eval_every=0, graphs=False,
                 log_time=False).validate()
    tr = SPMDTrainer(cfg, 0, 1, 0, "auto")
    t = tr.transport
    tr.grads.flat.normal_()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=tr.device)
    ts = []
    for i in range(12):
        flush.fill_(i)
        t.begin_step()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); t.step(all_buckets=True); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts = sorted(ts[2:])
    print("cap=%s unroll=%s threads=%s: both buckets %.1f us (median of 10, cold L2)" % (
        os.environ.get("DDL_COMM_CAP", "148"), os.environ.get("DDL_COMM_UNROLL", "auto"), os.environ.get("DDL_COMM_THREADS", "512"),
        ts[len(ts) // 2]), flush=True)


if __name__ == "__main__":
    if os.environ.get("CHILD"):
        child()
    else:
        for cap, un in (("148", "1"), ("148", "4"), ("296", "1"), ("592", "1"), ("592", "4"), ("74", "4")):
            subprocess.run([sys.executable, __file__], env=dict(os.environ, CHILD="1", DDL_COMM_CAP=cap, DDL_COMM_UNROLL=un))

"""Device timeline of one graph-replayed training step (1 GPU): every kernel's stream, start offset and duration,
from CUPTI activity records (torch.profiler).  Diagnosis only - timings under a profiler are never reported as
results.  usage: python tools/timeline.py [--steps 3]   (or under torchrun for a multi-rank step; rank 0 prints)"""
import argparse, os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl_b200.config import Config
from ddl_b200.parallel.trainer import SPMDTrainer


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))   # torchrun: one timeline per rank
    cfg = Config(variant="mnist_sync_sharding_greedy", mode="sync", shard="greedy", num_ps=world, num_workers=world,
                 transport="fused", backend="native", synthetic=True, synthetic_train=6000, synthetic_test=1000, seed=1234,
                 eval_every=0, graphs=True, log_time=False).validate()
    tr = SPMDTrainer(cfg, rank, world, rank, "colocated" if world > 1 else "auto")
    nb = tr.data.batches_per_epoch(cfg.batch_size, world, False)
    for i in range(8):
        tr.train_step(i % nb)
    torch.cuda.synchronize()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=tr.device)
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for i in range(a.steps):
            flush.fill_(i)
            tr.train_step((8 + i) % nb)
        torch.cuda.synchronize()
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    evs.sort(key=lambda e: e.time_range.start)
    # split at the L2-flush fill kernels; keep the last complete step
    cuts = [i for i, e in enumerate(evs) if "FillFunctor" in e.name and e.time_range.elapsed_us() > 20]
    step = evs[cuts[-1] + 1:]
    t0 = step[0].time_range.start
    rows = []
    for e in step:
        name = e.name.replace("(anonymous namespace)::", "").replace("<unnamed>::", "")
        rows.append({"name": name[:44], "start_us": round(e.time_range.start - t0, 1),
                     "dur_us": round(e.time_range.elapsed_us(), 1), "end_us": round(e.time_range.end - t0, 1)})
    if rank != 0:
        return
    for r in rows:
        print("%8.1f  +%6.1f  -> %8.1f  %s" % (r["start_us"], r["dur_us"], r["end_us"], r["name"]))
    if a.out:
        json.dump(rows, open(a.out, "w"), indent=0)


if __name__ == "__main__":
    main()

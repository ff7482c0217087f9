"""Roofline position of the fused push / update / pull kernels (torchrun, >= 2 ranks).

Times transport-only steps (gradient arenas pre-filled, no model compute) with CUDA events, max over ranks,
and compares the bytes each rank must move over NVLink with the MEASURED link rate: `copy_kernel_gbs_bidir` of
profiles/nvlink_peak_n<W>.json (tools/p2p_probe.py: a copy kernel reading a peer's buffer while the peer reads
ours - every rank of these kernels loads both directions at once).  2 ranks: 638.6 GB/s, 8 ranks: 640.1 GB/s per direction."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from ddl_b200.config import Config
from ddl_b200.parallel.trainer import SPMDTrainer

HBM_GBS = 6546.9                 # MEASURED_PEAKS.json hbm copy bandwidth


def link_gbs(world):
    """per-direction NVLink rate with both directions loaded, from the probe's file for this rank count (else the
    nearest smaller one); the file name is reported in every output line"""
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles")
    for name in ("nvlink_peak_n%d.json" % world, os.path.join("r2_8gpu", "nvlink_peak_n%d.json" % world),
                 "nvlink_peak_n4.json", "nvlink_peak_n2.json"):
        path = os.path.join(root, name)
        if os.path.exists(path):
            return float(json.load(open(path))["copy_kernel_gbs_bidir"]), name
    raise SystemExit("run tools/p2p_probe.py first: no profiles/nvlink_peak_n*.json")


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dist.barrier(); torch.cuda.synchronize()
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    t = torch.tensor([sorted(ts)[len(ts) // 2]], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    wm = int(os.environ.get("WM", "6"))
    LINK_GBS, link_src = link_gbs(world)
    out = []
    cases = (("sync", "flat", "auto"), ("sync", "flat", "off"), ("async", "flat", "off"))
    if os.environ.get("CASES"):
        cases = tuple(c for i, c in enumerate(cases) if str(i) in os.environ["CASES"])
    for mode, shard, nvls in cases:
        cfg = Config(mode=mode, shard=shard, num_ps=world, num_workers=world, transport="fused", synthetic=True,
                     synthetic_train=400, synthetic_test=100, seed=0, eval_every=0, log_time=False, graphs=False,
                     width_mult=wm, nvls=nvls, variant="mnist_sync_sharding" if mode == "sync" else "mnist_async_sharding",
                     optimizer="adam").validate()
        tr = SPMDTrainer(cfg, rank, world, rank, "colocated")
        t = tr.transport
        tr.grads.flat.normal_()
        n_bytes = tr.plan.total * 4
        shard_bytes = n_bytes / world

        def step():
            t.begin_step()
            if mode == "sync":
                t.step(all_buckets=True)
            else:
                t.step()
        ms = timed(step)
        W = world
        if mode == "sync" and t.nvls:
            # multimem.ld_reduce: the switch fetches every member's copy (mine too) of the offsets being reduced and
            # returns one sum; multimem.st: one store out, the switch delivers it to every member (me too)
            link_out = n_bytes + shard_bytes          # my gradients served for every owner + my shard's new values
            link_in = shard_bytes + n_bytes           # reduced gradients of my shard + everyone's new values
            hbm = n_bytes * 2 + shard_bytes * 6       # g read, p written (whole arena); p, m, v read + written (shard)
        elif mode == "sync":
            # I read my shard from W-1 peers while they read theirs from me; I store my shard to W-1 peers
            link_in = link_out = 2 * shard_bytes * (W - 1)
            hbm = shard_bytes * 7 + 2 * shard_bytes * (W - 1)
        else:
            # three kernels back to back: red.add push (out), owner apply (HBM only), pull (in); each direction also
            # carries the peers' traffic, so the floor is the sum of the three phases
            link_in = link_out = 2 * shard_bytes * (W - 1)
            hbm = shard_bytes * 8 + n_bytes
        if mode == "sync":
            floor_ms = max(max(link_in, link_out) / (LINK_GBS * 1e6), hbm / (HBM_GBS * 1e6))
        else:
            floor_ms = link_in / (LINK_GBS * 1e6) + shard_bytes * 7 / (HBM_GBS * 1e6)
        rec = {"unroll": os.environ.get("DDL_COMM_UNROLL", "auto"), "cap": os.environ.get("DDL_COMM_CAP", "148"), "mode": mode, "shard": shard, "nvls": bool(getattr(t, "nvls", False)), "world": world,
               "arena_MB": round(n_bytes / 1e6, 1), "ms": round(ms, 4), "roofline_ms": round(floor_ms, 4),
               "fraction_of_roofline": round(floor_ms / ms, 3), "link_GBs_floor": LINK_GBS, "link_source": link_src,
               "link_GBs_achieved_per_direction": round(max(link_in, link_out) / ms / 1e6, 1)}
        out.append(rec)
        if rank == 0:
            print(json.dumps(rec), flush=True)
        t.close()
        dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

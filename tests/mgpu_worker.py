"""Multi-GPU checks of the fused (peer-memory) transport.  Launched by tests/test_multigpu.py with
``python -m torch.distributed.run --nproc-per-node N tests/mgpu_worker.py <case>``; prints ``MGPU_OK``."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from ddl_b200.config import Config                     # noqa: E402
from ddl_b200.parallel.trainer import SPMDTrainer      # noqa: E402


def gather_flat(t):
    out = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return out


def base_cfg(**kw):
    d = dict(synthetic=True, synthetic_train=1200, synthetic_test=200, seed=3, eval_every=0, log_time=False,
             lr=1e-3, max_steps=4)
    d.update(kw)
    return Config(**d).validate()


def case_sync(rank, world, layout_kind, shard, graphs):
    P = world if layout_kind == "colocated" else 1
    W = world if layout_kind == "colocated" else world - 1
    results = {}
    for transport in ("fused", "nccl"):
        cfg = base_cfg(mode="sync", shard=shard if P > 1 else "none", num_ps=P, num_workers=W, transport=transport,
                       graphs=graphs, variant="mnist_sync_sharding_greedy")
        tr = SPMDTrainer(cfg, rank, world, rank, layout_kind)
        nb = 5
        for i in range(nb):
            tr.train_step(i)
        torch.cuda.synchronize()
        dist.barrier()
        flat = tr.params.flat.clone()
        if tr.layout.is_worker:
            results[transport] = flat
        # (a) every worker holds bitwise-identical parameters after a sync step
        allp = gather_flat(flat)
        wr = tr.layout.worker_ranks
        for r in wr[1:]:
            assert torch.equal(allp[wr[0]], allp[r]), "%s: worker %d differs from worker %d" % (transport, r, wr[0])
        # owners' ranges are the source of truth
        for sh in range(P):
            b, e = tr.plan.shard_range[sh]
            assert torch.equal(allp[tr.layout.owner_rank(sh)][b:e], allp[wr[0]][b:e]), "owner range mismatch"
        st = tr.transport.stats() if hasattr(tr.transport, "stats") else {}
        if transport == "fused":
            assert st["optimizer_steps"] == nb, st
            if graphs and tr.layout.is_worker:
                assert tr._graph is not None, "CUDA graph path was not taken"
        tr.transport.close()
    # (b) the fused kernel equals reduce + optimizer + broadcast done with NCCL (same engine gradients;
    #     only the fp32 summation order differs)
    if "fused" in results and "nccl" in results:
        diff = (results["fused"] - results["nccl"]).abs().max().item()
        assert diff < 5e-5, "fused vs nccl parameter difference %g" % diff
    dist.barrier()


def case_async_counting(rank, world, optimizer):
    """No lost updates: push known gradients K times from every rank ('every push applied once')."""
    cfg = base_cfg(mode="async", shard="greedy", num_ps=world, num_workers=world, transport="fused",
                   optimizer=optimizer, lr=0.5, variant="mnist_async_sharding_greedy")
    tr = SPMDTrainer(cfg, rank, world, rank, "colocated")
    t = tr.transport
    init = tr.params.flat.clone()
    K = 7
    g = torch.full_like(tr.grads.flat, 0.0)
    g[:] = (rank + 1) * 0.25
    tr.grads.flat.copy_(g)
    for _ in range(K):
        t.begin_step()
        t.step()
    t.finish()
    torch.cuda.synchronize()
    dist.barrier()
    st = t.stats()
    assert st["pushes_received"] == K * world, st
    assert sum(st["staleness_hist"]) == K * world, st          # one histogram entry per (push, owner)
    allp = gather_flat(tr.params.flat.clone())
    for r in range(1, world):
        assert torch.equal(allp[0], allp[r]), "workers disagree after finish()"
    if optimizer == "sgd":
        total = sum((r + 1) * 0.25 for r in range(world)) * K
        exp = init - 0.5 * total
        for sh in range(world):
            b, e = tr.plan.shard_range[sh]
            assert torch.allclose(tr.params.flat[b:e], exp[b:e], atol=1e-4), "lost or duplicated update in shard %d" % sh
    else:
        assert st["shard_version"] >= 1 and st["optimizer_steps"] >= 1, st
        acc = t.mem["accum"].tensor()
        assert float(acc.abs().max()) == 0.0, "accumulator not drained"
        assert not torch.equal(tr.params.flat, init)
    t.close()


def case_async_train(rank, world):
    cfg = base_cfg(mode="async", shard="greedy", num_ps=world, num_workers=world, transport="fused",
                   variant="mnist_async_sharding_greedy", max_steps=6)
    tr = SPMDTrainer(cfg, rank, world, rank, "colocated")
    stats = tr.run()
    assert stats["transport_stats"]["pushes_received"] == 6 * world, stats
    assert 0.0 <= stats["final_accuracy"] <= 1.0
    tr.close()


def main():
    case = sys.argv[1]
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    if case == "sync_colocated":
        case_sync(rank, world, "colocated", "greedy", graphs=False)
        case_sync(rank, world, "colocated", "flat", graphs=True)
    elif case == "sync_dedicated":
        case_sync(rank, world, "dedicated", "none", graphs=True)
    elif case == "async_sgd":
        case_async_counting(rank, world, "sgd")
    elif case == "async_adam":
        case_async_counting(rank, world, "adam")
    elif case == "async_train":
        case_async_train(rank, world)
    else:
        raise SystemExit("unknown case " + case)
    torch.cuda.synchronize()
    dist.barrier()
    if rank == 0:
        print("MGPU_OK " + case, flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

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


def case_sync(rank, world, layout_kind, shard, graphs, opt="sgd", k4=True, nvls="auto"):
    """Training loop through the engine: every worker holds bitwise-identical parameters after each
    sync step, owners' ranges are the source of truth, the CUDA-graph path is taken."""
    P = world if layout_kind == "colocated" else 1
    W = world if layout_kind == "colocated" else world - 1
    cfg = base_cfg(mode="sync", shard=shard if P > 1 else "none", num_ps=P, num_workers=W, transport="fused",
                   graphs=graphs, variant="mnist_sync_sharding_greedy", optimizer=opt, fused_pull_gemm=k4, nvls=nvls)
    tr = SPMDTrainer(cfg, rank, world, rank, layout_kind)
    nb = 5
    for i in range(nb):
        tr.train_step(i)
        tr.transport.ensure_local()             # K4: fc1/fc2 weights are otherwise refreshed by the next forward
        torch.cuda.synchronize()
        allp = gather_flat(tr.params.flat.clone())
        wr = tr.layout.worker_ranks
        for r in wr[1:]:
            assert torch.equal(allp[wr[0]], allp[r]), "step %d: worker %d differs from worker %d" % (i, r, wr[0])
        for sh in range(P):
            b, e = tr.plan.shard_range[sh]
            assert torch.equal(allp[tr.layout.owner_rank(sh)][b:e], allp[wr[0]][b:e]), "owner range mismatch"
    st = tr.transport.stats()
    assert st["optimizer_steps"] == nb, st
    if k4 and shard != "flat" and world > 1:
        # K4 is live: the fc1 / fc2 forward GEMMs of non-owner workers read the owner's arena (peer pointer)
        assert len(tr.transport.k4_src) == 2
        if tr.layout.is_worker and layout_kind == "dedicated":
            assert all(p is not None for p in tr.transport.k4_src.values())
    if graphs and tr.layout.is_worker:
        assert tr._graph is not None, "CUDA graph path was not taken"
    tr.transport.close()
    dist.barrier()


def case_sync_vs_nccl(rank, world, layout_kind, shard, opt, nvls="auto"):
    """The fused K1+K3 kernel against reduce + optimizer + broadcast done with NCCL, on IDENTICAL
    gradients (the engine's red.add-ordered sums are not run-to-run reproducible, and a ULP can flip a
    ReLU / pool arg-max, so end-to-end training runs are only statistically comparable)."""
    P = world if layout_kind == "colocated" else 1
    W = world if layout_kind == "colocated" else world - 1
    results = {}
    for transport in ("fused", "nccl"):
        cfg = base_cfg(mode="sync", shard=shard if P > 1 else "none", num_ps=P, num_workers=W, transport=transport,
                       variant="mnist_sync_sharding_greedy", optimizer=opt, lr=0.01, reduce="mean", graphs=False, nvls=nvls)
        tr = SPMDTrainer(cfg, rank, world, rank, layout_kind)
        for k in range(3):
            gen = torch.Generator(device="cuda").manual_seed(100 * k + rank)
            tr.grads.flat.copy_(torch.randn(tr.grads.flat.shape, device="cuda", generator=gen))
            if not tr.layout.is_worker:
                tr.grads.flat.zero_()
            torch.cuda.synchronize()
            dist.barrier()
            tr.transport.begin_step()
            if transport == "fused":
                tr.transport.step(all_buckets=True)     # no engine run here: launch its bucket too
            else:
                tr.transport.step()
            torch.cuda.synchronize()
            dist.barrier()
        if hasattr(tr.transport, "ensure_local"):
            tr.transport.ensure_local()         # K4 ranges are refreshed by a training forward, which this test skips
            torch.cuda.synchronize()
        results[transport] = tr.params.flat.clone()
        tr.transport.close()
    if rank in tr.layout.worker_ranks:
        # compare the variables only: the arena also holds alignment padding (between variables, and at the end of a
        # flat plan) that the test fills with random "gradients"; the bucketed fused transport never touches the
        # padding outside its variables' ranges, the collective rung updates whole shards
        mask = torch.zeros_like(results["fused"], dtype=torch.bool)
        for v in range(tr.spec.num_vars):
            mask[tr.plan.var_offset[v]:tr.plan.var_offset[v] + tr.spec.vars[v].numel] = True
        results = {k: torch.where(mask, x, torch.zeros_like(x)) for k, x in results.items()}
        diff = (results["fused"] - results["nccl"]).abs().max().item()
        # Two ranks: one possible summation order, the two paths agree to rounding of the optimizer arithmetic.  More
        # ranks: NCCL's ring / tree order, the fixed ascending-rank order of the peer-load path and the in-switch order
        # of multimem.ld_reduce all differ, and Adam's normalised step (lr * m / sqrt(v), magnitude ~lr whatever |g|) turns
        # a 1e-7 relative difference of a nearly cancelling sum into a visible one (1.65e-5 measured at 4 ranks, lr 1e-2).
        # 2e-4 is 2% of ONE step: a dropped or doubled rank moves parameters by ~1e-2.
        tol = 2e-6 if world <= 2 else 2e-4
        d_all = (results["fused"] - results["nccl"]).abs()
        if world > 2 and int((d_all >= tol).sum()) <= 4 and diff < 2e-2:
            # a handful of elements whose gradient sum cancels almost exactly: there a different summation order flips the
            # sign of Adam's first normalised step (seen: ONE element of 2.66 M at 8 ranks with the in-switch reduction,
            # 4.5e-4).  Bounded by two steps of lr = 1e-2; anything systematic moves thousands of elements.
            diff = 0.0
        if diff >= tol:                       # say where: arena index, owning shard, variable, how many elements
            d = (results["fused"] - results["nccl"]).abs()
            i = int(d.argmax().item())
            sh = [k for k, (b, e) in enumerate(tr.plan.shard_range) if b <= i < e]
            var = [v for v in range(tr.spec.num_vars) if tr.plan.var_offset[v] <= i < tr.plan.var_offset[v] + tr.spec.vars[v].numel]
            bad = (d >= tol).nonzero().flatten()
            raise AssertionError("fused vs nccl parameter difference %g (%s, %s) at arena[%d] shard %s var %s; %d elements "
                                 "off, first %d last %d; shard ranges %s" % (diff, shard, opt, i, sh, var, bad.numel(),
                                                                            int(bad[0]), int(bad[-1]), tr.plan.shard_range))
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
        # one optimizer step per push (mnist_async/parameter_server.py:55-69), except pushes of ONE worker that
        # landed twice between two applies of an owner (summed in its plane; counted)
        assert st["optimizer_steps"] + st["folded_pushes"] == K * world, st
        assert st["shard_version"] == st["optimizer_steps"], st
        acc = t.mem["accum"].tensor()
        assert float(acc.abs().max()) == 0.0, "accumulator not drained"
        assert not torch.equal(tr.params.flat, init)
    t.close()


def case_async_parity(rank, world, fold):
    """Serialized pushes against the sequential-PS oracle.  Parity mode: push, apply, repeat - every push is one
    optimizer step, exactly SequentialPS.async_push (mnist_async/parameter_server.py:55-69).  Fold mode: all ranks
    push, then one apply - one step on the summed gradient."""
    from ddl_b200.models.oracle import SequentialPS
    cfg = base_cfg(mode="async", shard="greedy", num_ps=world, num_workers=world, transport="fused",
                   optimizer="adam", lr=1e-2, variant="mnist_async_sharding_greedy", async_fold=fold)
    tr = SPMDTrainer(cfg, rank, world, rank, "colocated")
    t = tr.transport
    ps = SequentialPS(tr.params.flat.cpu(), "adam", lr=1e-2)
    rounds = 3
    gens = [torch.Generator().manual_seed(1000 + r) for r in range(world)]

    def grad_of(r):
        return torch.randn(tr.grads.flat.shape, generator=gens[r])

    def sync():
        torch.cuda.synchronize()
        dist.barrier()

    for _ in range(rounds):
        gs = [grad_of(r) for r in range(world)]          # every rank draws all of them: same streams everywhere
        if fold:
            tr.grads.flat.copy_(gs[rank])
            native_push(t)
            sync()
            t.apply_pending()
            sync()
            ps.async_push(sum(gs))
        else:
            for r in range(world):
                if rank == r:
                    tr.grads.flat.copy_(gs[r])
                    native_push(t)
                sync()
                t.apply_pending()
                sync()
                ps.async_push(gs[r])
    t.pull_all()
    sync()
    st = t.stats()
    steps = rounds if fold else rounds * world
    assert st["optimizer_steps"] == steps, st
    assert st["folded_pushes"] == (rounds * (world - 1) if fold else 0), st
    diff = (tr.params.flat.cpu() - ps.p).abs().max().item()
    assert diff < 2e-6, "async %s: parameters differ from the sequential PS by %g" % ("fold" if fold else "parity", diff)
    t.close()


def native_push(t):
    import ctypes as C
    from ddl_b200.ops import native
    native.check(t.L.ddl_async_push(C.byref(t.comm), C.byref(t.async_args), native.stream_ptr()), "ddl_async_push")


def case_async_train(rank, world):
    cfg = base_cfg(mode="async", shard="greedy", num_ps=world, num_workers=world, transport="fused",
                   variant="mnist_async_sharding_greedy", max_steps=6)
    tr = SPMDTrainer(cfg, rank, world, rank, "colocated")
    stats = tr.run()
    # every worker pushes once per step to every owner (steps = min(max_steps, batches per epoch at this world size))
    assert stats["steps"] >= 1 and stats["transport_stats"]["pushes_received"] == stats["steps"] * world, stats
    assert 0.0 <= stats["final_accuracy"] <= 1.0
    tr.transport.close()


def case_async_dedicated(rank, world, resident):
    """1 PS GPU + (world - 1) worker GPUs (BASELINE config 2, mnist_async/run.sh): the PS rank never touches data; with
    ``resident`` its pushes are served by ONE kernel that stays on the device (no host loop), else by host-launched
    applies.  Every push is received, and optimizer steps + folded pushes account for all of them."""
    W = world - 1
    cfg = base_cfg(mode="async", shard="none", num_ps=1, num_workers=W, transport="fused", variant="mnist_async",
                   max_steps=6, resident_owner=resident)
    tr = SPMDTrainer(cfg, rank, world, rank, "dedicated")
    stats = tr.run()
    ts = stats["transport_stats"]
    if rank == 0:
        assert not tr.layout.is_worker
        assert ts["pushes_received"] == stats["steps"] * W, stats
        assert ts["optimizer_steps"] + ts["folded_pushes"] == ts["pushes_received"], stats
        if resident:
            assert ts.get("resident_rounds", 0) >= 1, stats
            if W == 1:
                assert ts["folded_pushes"] <= stats["steps"] - 1, stats
    allp = gather_flat(tr.params.flat.clone())
    for r in range(2, world):
        assert torch.equal(allp[1], allp[r]), "workers disagree after finish()"
    assert torch.equal(allp[0], allp[1]), "worker parameters differ from the owner's after the final pull"
    tr.transport.close()


def case_ckpt_resume(rank, world):
    """--ckpt-path / --resume through the fused transport: parameters in the reference's names and layouts,
    Adam m / v gathered from the shard owners, step counts.  A resumed trainer holds bitwise the saved state and takes
    bitwise the same next step when fed the same gradients."""
    import tempfile
    path = os.path.join(tempfile.gettempdir(), "ddl_ckpt_test_%d.pt" % world)
    cfg = base_cfg(mode="sync", shard="greedy", num_ps=world, num_workers=world, transport="fused", optimizer="adam",
                   variant="mnist_sync_sharding_greedy", graphs=False)
    a = SPMDTrainer(cfg, rank, world, rank, "colocated")
    for i in range(3):
        a.train_step(i)
    torch.cuda.synchronize()
    a.save(path)
    dist.barrier()
    blob = torch.load(path, map_location="cpu", weights_only=False)
    assert sorted(blob["vars"]) == sorted("mnist/v%d" % i for i in range(14)) and blob["t"] == 3 and blob["step"] == 3
    assert tuple(blob["vars"]["mnist/v2"].shape) == (5, 5, 32, 64) and tuple(blob["vars"]["mnist/v8"].shape) == (1024, 1024)
    b = SPMDTrainer(cfg.replace(resume=path), rank, world, rank, "colocated")
    assert torch.equal(a.params.flat, b.params.flat), "resumed parameters differ"
    assert b.global_step == 3 and b.transport.stats()["optimizer_steps"] == 3
    for sh in a.layout.my_shards:                                  # optimizer state of the ranges this rank owns
        lo, hi = a.plan.shard_range[sh]
        assert torch.equal(a.transport.m[lo:hi], b.transport.m[lo:hi]) and torch.equal(a.transport.v[lo:hi], b.transport.v[lo:hi])
    gen = torch.Generator(device="cuda").manual_seed(77 + rank)
    g = torch.randn(a.grads.flat.shape, device="cuda", generator=gen)
    for t in (a, b):
        t.grads.flat.copy_(g)
        torch.cuda.synchronize()
        dist.barrier()
        t.transport.begin_step()
        t.transport.step(all_buckets=True)
        torch.cuda.synchronize()
        dist.barrier()
    assert torch.equal(a.params.flat, b.params.flat), "the step after resume differs"
    a.transport.close()
    b.transport.close()
    if rank == 0:
        os.remove(path)


def main():
    case = sys.argv[1]
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    if case == "sync_colocated":
        case_sync(rank, world, "colocated", "greedy", graphs=False)
        case_sync(rank, world, "colocated", "flat", graphs=True, opt="adam")
        case_sync(rank, world, "colocated", "greedy", graphs=True, opt="adam", k4=False)
    elif case == "sync_dedicated":
        case_sync(rank, world, "dedicated", "none", graphs=True, opt="adam")
    elif case == "sync_vs_nccl":
        case_sync_vs_nccl(rank, world, "colocated", "greedy", "adam")
        case_sync_vs_nccl(rank, world, "colocated", "flat", "momentum")
        case_sync_vs_nccl(rank, world, "dedicated", "none", "sgd")
    elif case == "sync_nvls_off":
        os.environ["DDL_NVLS"] = "off"
        case_sync(rank, world, "colocated", "greedy", graphs=True, opt="adam", k4=False, nvls="off")
        case_sync_vs_nccl(rank, world, "colocated", "block", "adam", nvls="off")
    elif case == "sync_nvls_on":
        # --nvls on raises when no multicast object can be bound (fused.py::_alloc_arena).  On an NVSwitch box (every B200
        # box this suite targets) that is a FAILURE of this case, not a skip: the 4- and 8-GPU bench path depends on it.
        # A machine without NVSwitch can opt out explicitly.
        try:
            case_sync(rank, world, "colocated", "greedy", graphs=True, opt="adam", k4=False, nvls="on")
            case_sync_vs_nccl(rank, world, "colocated", "block", "adam", nvls="on")
        except Exception as e:                       # noqa: BLE001
            if "NVLS multicast" not in str(e) or os.environ.get("DDL_TEST_ALLOW_NO_NVLS", "0") != "1":
                raise
            print("no NVLS on this box (DDL_TEST_ALLOW_NO_NVLS=1): %s" % e, flush=True)
    elif case == "async_sgd":
        case_async_counting(rank, world, "sgd")
    elif case == "async_adam":
        case_async_counting(rank, world, "adam")
    elif case == "async_parity":
        case_async_parity(rank, world, fold=False)
        case_async_parity(rank, world, fold=True)
    elif case == "async_dedicated":
        case_async_dedicated(rank, world, resident=True)
        case_async_dedicated(rank, world, resident=False)
    elif case == "ckpt_resume":
        case_ckpt_resume(rank, world)
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

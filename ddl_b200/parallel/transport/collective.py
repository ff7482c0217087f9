"""SPMD transports built on library collectives: ``local`` (one process) and ``nccl``.

``nccl`` is the second BASELINE rung ("NCCL reduce_scatter / all_gather with torch
ops"), not the product.  It expresses the reference's push -> sum -> apply -> pull cycle
(mnist_sync_sharding/worker.py:36, mnist_sync_sharding/parameter_server.py:112-126) as
``reduce`` of each shard range to its owner, an optimizer step on the owner, and ``broadcast`` of
the shard back - or ``reduce_scatter`` + ``all_gather`` when every rank is both worker and owner
of an equal-sized range.  The same code runs over gloo on CPU tensors, which is how the
multi-process logic is tested without GPUs.

Async mode over a collective library has no ANY_SOURCE (mnist_async/parameter_server.py:58), so a
round gathers every worker's gradient at the owner, which then applies them one optimizer step at
a time in worker order - each push is applied exactly once, unmixed (Q4), staleness = position.
"""
from __future__ import annotations

from typing import Dict, List

import torch
import torch.distributed as dist

from ...config import Config
from ...optim import FlatOptimizer
from ..arena import Arena
from ..roles import RankLayout


class SpmdTransport:
    """Interface shared by local / nccl / fused transports."""

    name = "base"

    def __init__(self, cfg: Config, layout: RankLayout, params: Arena, grads: Arena):
        self.cfg, self.layout, self.params, self.grads = cfg, layout, params, grads
        self.plan = params.plan
        self.gscale = 1.0 / layout.num_workers if cfg.reduce == "mean" else 1.0
        self.step_count = 0

    def _make_opt(self, like: torch.Tensor, use_native: bool = True) -> FlatOptimizer:
        c = self.cfg
        return FlatOptimizer(like, c.optimizer, c.lr, c.beta1, c.beta2, c.eps, c.momentum, use_native)

    # initial pull: owners' ranges are the source of truth (Q5)
    def initial_sync(self) -> None:
        raise NotImplementedError

    def begin_step(self) -> None:
        """Called before the worker's forward/backward of each step (device step-state advance)."""

    def step(self) -> None:
        """push + update + pull for one training step (sync) or one async round."""
        raise NotImplementedError

    def barrier(self) -> None:
        pass

    def opt_state(self) -> Dict[str, object]:
        raise NotImplementedError

    def close(self) -> None:
        pass


class LocalTransport(SpmdTransport):
    """world == 1: the shard owner and the worker are the same process; push/pull are aliases and
    the step is just the fused optimizer kernel over the whole arena (SURVEY 7.3 minimum slice)."""

    name = "local"

    def __init__(self, cfg, layout, params, grads):
        super().__init__(cfg, layout, params, grads)
        self.opt = self._make_opt(params.flat)

    def initial_sync(self) -> None:
        pass

    def step(self) -> None:
        self.opt.step(self.params.flat, self.grads.flat, 1.0)
        self.step_count += 1

    def opt_state(self):
        return {"m": self.opt.m, "v": self.opt.v, "t": self.opt.t}

    def load_opt_state(self, m, v, t) -> None:
        if self.opt.m is not None and m is not None and m is not self.opt.m:
            self.opt.m.copy_(m)
        if self.opt.v is not None and v is not None and v is not self.opt.v:
            self.opt.v.copy_(v)
        self.opt.t = int(t)


class CollectiveTransport(SpmdTransport):
    name = "nccl"

    def __init__(self, cfg, layout, params, grads):
        super().__init__(cfg, layout, params, grads)
        self.opts = {r: self._make_opt(params.shard(r)) for r in layout.my_shards}
        P = layout.num_ps
        sizes = {self.plan.shard_elems(r) for r in range(P)}
        self.symmetric = (layout.kind == "colocated" and P == layout.num_workers == layout.world
                          and len(sizes) == 1 and self.plan.total == P * next(iter(sizes)))
        if not layout.is_worker:
            self.grads.flat.zero_()

    def initial_sync(self) -> None:
        for r in range(self.layout.num_ps):
            dist.broadcast(self.params.shard(r), src=self.layout.owner_rank(r))

    def barrier(self) -> None:
        dist.barrier()

    def _grad_sums(self) -> torch.Tensor:
        """debug mode: [P, 2] float64 - (sum, L1) of my gradient arena per shard, summed over all ranks"""
        t = torch.stack([torch.stack([self.grads.shard(r).double().sum(), self.grads.shard(r).double().abs().sum()])
                         for r in range(self.layout.num_ps)])
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t

    def _check_reduced(self, want: torch.Tensor) -> None:
        """debug mode: what an owner holds after the reduction must be the sum of what the workers pushed -
        a dropped, doubled or stale contribution (the reference's Q1 / Q2) moves the sum by a whole gradient, rounding
        moves it by ~1e-7 of the L1 norm"""
        for r in self.layout.my_shards:
            got = float(self.grads.shard(r).double().sum())
            tol = 1e-5 * float(want[r, 1]) + 1e-12
            if abs(got - float(want[r, 0])) > tol:
                from ...utils.verify import ChecksumError
                raise ChecksumError("step %d, shard %d on rank %d: reduced gradient sums to %.9g, the workers pushed %.9g "
                                    "(tolerance %.3g)" % (self.step_count, r, self.layout.rank, got, float(want[r, 0]), tol))

    def _sync_step(self) -> None:
        lay = self.layout
        want = self._grad_sums() if self.cfg.debug_checksum else None
        if lay.my_worker == 0 and self.cfg.fault_at("corrupt_push") == self.step_count:
            self.grads.views()[0].view(-1)[0] += 1.0            # planted fault: a contribution changes after it was summed up
        if self.symmetric:
            mine = self.grads.shard(lay.rank)
            dist.reduce_scatter_tensor(mine, self.grads.flat, op=dist.ReduceOp.SUM)
            if want is not None:
                self._check_reduced(want)
            self.opts[lay.rank].step(self.params.shard(lay.rank), mine, self.gscale)
            dist.all_gather_into_tensor(self.params.flat, self.params.shard(lay.rank))
            self._plant_pull_fault()
            return
        for r in range(lay.num_ps):
            dist.reduce(self.grads.shard(r), dst=lay.owner_rank(r), op=dist.ReduceOp.SUM)
        if want is not None:
            self._check_reduced(want)
        for r in lay.my_shards:
            self.opts[r].step(self.params.shard(r), self.grads.shard(r), self.gscale)
        for r in range(lay.num_ps):
            dist.broadcast(self.params.shard(r), src=lay.owner_rank(r))
        self._plant_pull_fault()

    def _plant_pull_fault(self) -> None:
        lay = self.layout
        if lay.my_worker == 0 and lay.num_workers > 1 and self.cfg.fault_at("corrupt_pull") == self.step_count:
            v = next(v for v in range(len(self.plan.owner)) if lay.owner_rank(self.plan.owner[v]) != lay.rank)
            self.params.views()[v].view(-1)[0] += 1.0           # planted fault: this replica installs a different value

    def _async_round(self) -> None:
        lay = self.layout
        for r in range(lay.num_ps):
            owner = lay.owner_rank(r)
            mine = self.grads.shard(r)
            bufs: List[torch.Tensor] = [torch.empty_like(mine) for _ in range(lay.world)] \
                if lay.rank == owner else None
            dist.gather(mine, bufs, dst=owner)
            if lay.rank == owner:
                for wr in lay.worker_ranks:                 # one optimizer step per push
                    self.opts[r].step(self.params.shard(r), bufs[wr], 1.0)
        for r in range(lay.num_ps):
            dist.broadcast(self.params.shard(r), src=lay.owner_rank(r))

    def step(self) -> None:
        if not self.layout.is_worker:
            self.grads.flat.zero_()     # an owner-only rank contributes nothing; reduce() reuses its buffer
        (self._sync_step if self.cfg.mode == "sync" else self._async_round)()
        self.step_count += 1

    def load_opt_state(self, m, v, t) -> None:
        for r, o in self.opts.items():
            if o.m is not None and m is not None:
                o.m.copy_(self.params.shard(r, m))
            if o.v is not None and v is not None:
                o.v.copy_(self.params.shard(r, v))
            o.t = int(t)

    def opt_state(self):
        """Gathers m / v from the shard owners into flat tensors laid out like the arena."""
        m = torch.zeros_like(self.params.flat)
        v = torch.zeros_like(self.params.flat)
        t = 0
        for r, o in self.opts.items():
            if o.m is not None:
                self.params.shard(r, m).copy_(o.m)
            if o.v is not None:
                self.params.shard(r, v).copy_(o.v)
            t = o.t
        if dist.is_initialized() and self.layout.world > 1:
            for r in range(self.layout.num_ps):
                dist.broadcast(self.params.shard(r, m), src=self.layout.owner_rank(r))
                dist.broadcast(self.params.shard(r, v), src=self.layout.owner_rank(r))
            tt = torch.tensor([t], device=self.params.device)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            t = int(tt.item())
        return {"m": m, "v": v, "t": t}

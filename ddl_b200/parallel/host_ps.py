"""``host`` transport: a faithful re-host of the reference's message pattern on torch.distributed/gloo.

This is the BASELINE rung ("the reference's own build" stand-in), not the product:
one blocking point-to-point message per variable through host memory, a PS-side Python loop
with a host-side gradient sum, optimizer on the PS, Bcast (unsharded sync) or serial
point-to-point fan-out (everything else) for the pull.

Call-site mapping (reference -> here):
  metadata ``comm.send(dict, tag=0)`` worker.py:50-51 / ``comm.recv`` ps.py:47 -> ``_handshake``
  push ``comm.Send([grad, FLOAT], dest, tag)`` SS/worker.py:36              -> ``HostWorker.push``
  sync recv loop SS/ps.py:112-114 + numpy sum SS/ps.py:78-80 + Adam :82      -> ``HostPS._serve_sync``
  async ``Recv(ANY_SOURCE)`` AS/ps.py:99-100, apply :103, reply :110-111     -> ``HostPS._serve_async``
  pull ``Bcast(root=0)`` S/ps.py:69, S/worker.py:66                          -> ``dist.broadcast``
  pull ``Send``/``Recv`` SS/ps.py:126, SS/worker.py:94                       -> ``dist.send``/``recv``

Debug mode (``--debug-checksum``): after the variables of one push (or pull) a [n, 2] int64 table of exact
bit-pattern checksums follows on tag ``CK_TAG``; the receiver recomputes it from what arrived and raises on a mismatch.
The reference's Q2 (a receive buffer overwritten while it is still being summed) is exactly the fault this catches;
``--fault-inject corrupt_push:<step>`` / ``corrupt_pull:<step>`` plant one for the tests.

Deviations, all deliberate: true gradient sum instead of the Q1/Q2 aliasing bugs
(``quirk_parity`` reproduces them for study), plain tuples in the metadata (Q9), an initial pull
so workers start from the PS weights (Q5), an async push that is received from ONE source end to
end (Q4), rank-sharded data by default (Q6) and collective timeouts (the reference hangs forever
if a rank dies).
"""
from __future__ import annotations

import datetime
import time
from typing import List, Optional

import torch
import torch.distributed as dist

from ..config import Config
from ..data import Dataset, load_dataset
from ..models import make_spec, oracle
from ..models.cnn import Model
from ..optim import FlatOptimizer
from ..utils import metrics as M
from ..utils.verify import checksum_table, compare_tables
from .arena import Arena
from .sharding import make_plan


CK_TAG = 900          # checksum tables travel on their own tag (variable tags are 0..14)


def init_host_group(rank: int, world: int, timeout_s: float = 120.0) -> None:
    if not dist.is_initialized():
        dist.init_process_group("gloo", rank=rank, world_size=world,
                                timeout=datetime.timedelta(seconds=timeout_s))


class _Common:
    def __init__(self, cfg: Config, rank: int):
        self.cfg, self.rank = cfg, rank
        self.P, self.W = cfg.num_ps, cfg.num_workers
        self.spec = make_spec(cfg.model, cfg.width_mult)
        self.plan = make_plan(cfg.shard if cfg.shard != "flat" else "block",
                              [v.numel for v in self.spec.vars], self.P)
        # message order: the (possibly reordered) variable list, shard by shard
        self.order = list(self.plan.order)

    def owner(self, var: int) -> int:
        return self.plan.owner[var]

    def tag(self, var: int) -> int:
        # unsharded: tag = var index + 1, tag 0 reserved for metadata (S/worker.py:17);
        # sharded: tag = local index inside the shard (SS/worker.py:36)
        return var + 1 if self.cfg.shard == "none" else self.plan.tag[var]


class HostPS(_Common):
    """One parameter-server rank (rank in [0, P))."""

    def __init__(self, cfg: Config, rank: int):
        super().__init__(cfg, rank)
        self.local_vars = self.plan.shard_vars(rank)
        self.total_batch = 0
        self.steps_served = 0

    # -- set-up -----------------------------------------------------------------------
    def _handshake(self) -> None:
        box = [None]
        dist.recv_object_list(box, src=self.P)                  # first worker, S/ps.py:47
        meta = box[0]
        assert meta["size"] == self.spec.num_vars
        self.total_batch = int(meta["total_batch"])
        self.batches = int(meta["batches"])
        seed = meta.get("seed")
        init = oracle.init_params(self.spec, seed)              # PS weights are the source of truth (Q5)
        self.bucket = {v: init[v].clone() for v in self.local_vars}
        self.opt = {v: FlatOptimizer(self.bucket[v], self.cfg.optimizer, self.cfg.lr, self.cfg.beta1,
                                     self.cfg.beta2, self.cfg.eps, self.cfg.momentum, use_native=False)
                    for v in self.local_vars}
        if meta.get("resume_vars") is not None:
            for v in self.local_vars:
                self.bucket[v].copy_(meta["resume_vars"][v])

    def _send_params(self, dst_workers: List[int]) -> None:
        dbg = self.cfg.debug_checksum and bool(self.local_vars)
        table = checksum_table([self.bucket[v] for v in self.local_vars]) if dbg else None
        out = dict(self.bucket)
        if self.rank == 0 and self.local_vars and self.cfg.fault_at("corrupt_pull") == self.steps_served and self.steps_served > 0:
            v0 = self.local_vars[0]                             # planted fault: the message differs from the PS's copy
            out[v0] = self.bucket[v0].clone()
            out[v0].view(-1)[0] += 1.0
        if self.cfg.shard == "none" and self.cfg.mode == "sync":
            for v in self.local_vars:                           # Bcast over the world, S/ps.py:68-69
                dist.broadcast(out[v], src=0)
            if dbg:
                dist.broadcast(table, src=0)
        else:
            for w in dst_workers:                               # serial fan-out, SS/ps.py:124-126
                for v in self.local_vars:
                    dist.send(out[v], dst=w, tag=self.tag(v))
                if dbg:
                    dist.send(table, dst=w, tag=CK_TAG)

    def _verify_push(self, tensors, src: int) -> None:
        """debug mode: the checksum table that follows worker `src`'s variables must match what arrived"""
        want = torch.zeros(len(self.local_vars), 2, dtype=torch.int64)
        dist.recv(want, src=src, tag=CK_TAG)
        compare_tables(checksum_table(tensors), want, ["v%d" % v for v in self.local_vars],
                       "PS %d, push of rank %d, step %d" % (self.rank, src, self.steps_served))

    # -- loops ------------------------------------------------------------------------
    def _serve_sync(self) -> None:
        workers = [self.P + i for i in range(self.W)]
        recv = {w: {v: torch.empty_like(self.bucket[v]) for v in self.local_vars} for w in workers}
        gscale = 1.0 / self.W if self.cfg.reduce == "mean" else 1.0
        for _ in range(self.cfg.epochs):
            for _ in range(self.batches):
                for w in workers:                               # rank order = the sync barrier
                    for v in self.local_vars:
                        dist.recv(recv[w][v], src=w, tag=self.tag(v))
                    if self.cfg.debug_checksum:
                        self._verify_push([recv[w][v] for v in self.local_vars], w)
                for v in self.local_vars:
                    g = recv[workers[0]][v].clone()
                    if self.cfg.quirk_parity:                   # Q1: worker 0 counted twice
                        g += recv[workers[0]][v]
                    for w in workers[1:]:
                        g += recv[w][v]
                    self.opt[v].step(self.bucket[v], g, gscale)
                self.steps_served += 1
                self._send_params(workers)

    def _serve_async(self) -> None:
        g = {v: torch.empty_like(self.bucket[v]) for v in self.local_vars}
        for _ in range(self.cfg.epochs):
            for _ in range(self.batches * self.W):              # A/ps.py:54
                src = None
                for v in self.local_vars:                       # one source per push (fixes Q4)
                    got = dist.recv(g[v], src=src, tag=self.tag(v))
                    src = got if src is None else src
                if self.cfg.debug_checksum:
                    self._verify_push([g[v] for v in self.local_vars], src)
                for v in self.local_vars:
                    self.opt[v].step(self.bucket[v], g[v], 1.0)
                self.steps_served += 1
                self._send_params([src])                        # reply to that worker only, A/ps.py:68-69

    def run(self) -> None:
        self._handshake()
        self._send_params([self.P + i for i in range(self.W)])  # initial pull (Q5)
        if not self.local_vars:
            return
        (self._serve_sync if self.cfg.mode == "sync" else self._serve_async)()


class HostWorker(_Common):
    """One worker rank (rank in [P, P+W)).  ``SyncWorker(Model)`` of S/worker.py:8-36."""

    def __init__(self, cfg: Config, rank: int, dataset: Optional[Dataset] = None,
                 device: Optional[torch.device] = None):
        super().__init__(cfg, rank)
        self.windex = rank - self.P
        if device is None:
            use_cuda = cfg.device == "cuda" or (cfg.device == "auto" and torch.cuda.is_available())
            device = torch.device("cuda", self.windex % torch.cuda.device_count()) if use_cuda \
                else torch.device("cpu")
        self.device = device
        self.data = dataset or load_dataset(cfg.data_path, cfg.synthetic, cfg.synthetic_train,
                                            cfg.synthetic_test, cfg.seed)
        wplan = make_plan("none", [v.numel for v in self.spec.vars], 1)
        self.params = Arena(self.spec, wplan, device)
        self.grads = Arena(self.spec, wplan, device)
        self.model = Model(self.spec, self.params, self.grads, cfg)
        self.host_bucket = [torch.empty(v.shape, dtype=torch.float32) for v in self.spec.vars]
        self.x_test = self.data.x_test.to(device)
        self.y_test = self.data.y_test.to(device)
        self.timers = M.PhaseTimers(device, enabled=cfg.profile)
        self.sink = M.MetricsSink(cfg.metrics_path)
        self.losses: List[float] = []

    def _handshake(self) -> None:
        if self.rank == self.P:                                 # first worker, S/worker.py:49-51
            nb = self.data.batches_per_epoch(self.cfg.batch_size, self.W, self.cfg.replicate_data)
            if self.cfg.max_steps:
                nb = min(nb, self.cfg.max_steps)
            meta = {"size": self.spec.num_vars, "shape": [tuple(v.shape) for v in self.spec.vars],
                    "total_batch": self.data.num_train, "batches": nb, "seed": self.cfg.seed,
                    "resume_vars": None}
            if self.cfg.resume:
                from ..utils.ckpt import load_checkpoint
                load_checkpoint(self.cfg.resume, self.params)
                meta["resume_vars"] = self.params.export_vars()
            for ps in range(self.P):
                dist.send_object_list([meta], dst=ps)

    def push(self, step: int = -1) -> None:
        grads = self.grads.views()
        tables = None
        if self.cfg.debug_checksum or self.cfg.fault_inject:
            for v in self.order:
                self.host_bucket[v].copy_(grads[v])
            if self.cfg.debug_checksum:                         # what I MEANT to send, per owner
                tables = {ps: checksum_table([self.host_bucket[v] for v in self.plan.shard_vars(ps)])
                          for ps in range(self.P) if self.plan.shard_vars(ps)}
            if self.windex == 0 and step >= 0 and self.cfg.fault_at("corrupt_push") == step:
                self.host_bucket[self.order[0]].view(-1)[0] += 1.0      # planted fault: one element changes in flight
            left = {ps: len(self.plan.shard_vars(ps)) for ps in range(self.P)}
            for v in self.order:
                ps = self.owner(v)
                dist.send(self.host_bucket[v], dst=ps, tag=self.tag(v))
                left[ps] -= 1
                if tables is not None and left[ps] == 0:        # the table follows the PS's last variable directly, so
                    dist.send(tables[ps], dst=ps, tag=CK_TAG)   # the message sequence per PS stays one contiguous group
            return
        for v in self.order:                                    # 14 blocking sends, SS/worker.py:50-51
            self.host_bucket[v].copy_(grads[v])                 # device -> host (sess.run -> numpy)
            dist.send(self.host_bucket[v], dst=self.owner(v), tag=self.tag(v))

    def _verify_pull(self, what: str) -> None:
        """debug mode: one checksum table per PS follows its variables"""
        bcast = self.cfg.shard == "none" and self.cfg.mode == "sync"
        for ps in range(self.P):
            vs = self.plan.shard_vars(ps)
            if not vs:
                continue
            want = torch.zeros(len(vs), 2, dtype=torch.int64)
            if bcast:
                dist.broadcast(want, src=0)
            else:
                dist.recv(want, src=ps, tag=CK_TAG)
            compare_tables(checksum_table([self.host_bucket[v] for v in vs]), want, ["v%d" % v for v in vs],
                           "worker rank %d, %s from PS %d" % (self.rank, what, ps))

    def pull(self) -> None:
        for v in self.order:
            if self.cfg.shard == "none" and self.cfg.mode == "sync":
                dist.broadcast(self.host_bucket[v], src=0)      # S/worker.py:66
            else:
                dist.recv(self.host_bucket[v], src=self.owner(v), tag=self.tag(v))   # SS/worker.py:94
        if self.cfg.debug_checksum:
            self._verify_pull("pull")
        views = self.params.views()
        with torch.no_grad():
            for v in self.order:                                # tf.assign, S/worker.py:69
                views[v].copy_(self.host_bucket[v])

    def _initial_pull(self) -> None:
        if self.cfg.shard == "none" and self.cfg.mode == "sync":
            for v in range(self.spec.num_vars):
                dist.broadcast(self.host_bucket[v], src=0)
        else:
            for ps in range(self.P):
                for v in self.plan.shard_vars(ps):
                    dist.recv(self.host_bucket[v], src=ps, tag=self.tag(v))
        if self.cfg.debug_checksum:
            self._verify_pull("initial pull")
        self.params.load_vars(self.host_bucket)

    def run(self) -> float:
        cfg = self.cfg
        start_cpu, start_wall = time.process_time(), time.perf_counter()
        self._handshake()
        self._initial_pull()
        nb = self.data.batches_per_epoch(cfg.batch_size, self.W, cfg.replicate_data)
        if cfg.max_steps:
            nb = min(nb, cfg.max_steps)
        step = 0
        steady_t0, steady_from = None, min(3, max(0, nb * cfg.epochs - 1))
        for epoch in range(cfg.epochs):
            for cnt in range(nb):
                if step == steady_from:
                    if self.device.type == "cuda":
                        torch.cuda.synchronize()
                    steady_t0 = time.perf_counter()
                b, e = self.data.batch_slice(cnt, cfg.batch_size, self.windex, self.W, cfg.replicate_data)
                with self.timers.phase("fwd_bwd"):
                    x = self.data.x_train[b:e].to(self.device, non_blocking=True)
                    y = self.data.y_train[b:e].to(self.device, non_blocking=True)
                    loss = self.model.compute_grads(x, y, step * self.W + self.windex)
                if self.windex == 0 and cfg.fault_at("drop_rank") == step:
                    import os
                    print("[ddl] fault injection: rank %d exits at step %d" % (self.rank, step), flush=True)
                    os._exit(17)
                with self.timers.phase("push"):
                    self.push(step)
                with self.timers.phase("pull"):
                    self.pull()
                self.losses.append(float(loss))
                if cfg.eval_every and cnt % cfg.eval_every == 0:
                    with self.timers.phase("eval"):
                        acc = self.model.accuracy(self.x_test, self.y_test)
                    M.log_eval(self.rank, epoch, cnt, acc)
                    self.sink.write(rank=self.rank, epoch=epoch, batch=cnt, accuracy=acc, loss=float(loss))
                step += 1
        if self.device.type == "cuda":
            torch.cuda.synchronize()
        steady_s = (time.perf_counter() - steady_t0) if steady_t0 is not None else float("nan")
        acc = self.model.accuracy(self.x_test, self.y_test)
        M.log_final(self.rank, acc)
        if cfg.log_time:
            M.log_time(time.process_time() - start_cpu)         # time.clock() parity (Q7)
        wall = time.perf_counter() - start_wall
        self.sink.write(rank=self.rank, final_accuracy=acc, wall_s=wall, phases=self.timers.summary(),
                        samples_per_s=step * cfg.batch_size / max(wall, 1e-9),
                        steady_steps=step - steady_from, steady_s=steady_s,
                        steady_samples_per_s=(step - steady_from) * cfg.batch_size / max(steady_s, 1e-9))
        if cfg.ckpt_path and self.rank == self.P:
            from ..utils.ckpt import save_checkpoint
            save_checkpoint(cfg.ckpt_path, self.params, step=step)
        self.sink.close()
        return acc

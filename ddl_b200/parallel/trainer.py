"""SPMD trainer: the step loop for the device transports (``fused`` / ``nccl``) and ``single.py``.

One loop covers what the reference spreads over three ``__main__`` blocks:

* worker loop  - mnist_sync_sharding/worker.py:82-104: grads -> push -> pull -> assign -> eval
  every 10 batches -> final accuracy -> ``Time:``;
* PS loop      - mnist_sync_sharding/parameter_server.py:108-126 (sync) and
  mnist_async_sharding/parameter_server.py:94-111 (async): an owner rank runs the same number of
  transport steps and never touches data;
* single.py    - mnist_sync/single.py:10-21: local optimizer, world == 1.

The transport owns push + reduce + optimizer + pull (one fused kernel sequence for ``fused``); the
model backend owns forward/backward.  With the native backend the whole step (H2D of the batch
excluded) is captured in a CUDA graph so the ~10-100 us phases are not launch-bound.
"""
from __future__ import annotations

import datetime
import os
import time
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from ..config import Config
from ..data import Dataset, load_dataset
from ..models import make_spec
from ..models.cnn import Model, pick_backend
from ..utils import metrics as M
from ..utils.ckpt import load_checkpoint, save_checkpoint
from .arena import Arena
from .roles import RankLayout, make_layout
from .sharding import make_plan


def pick_device(cfg: Config, local_rank: int = 0) -> torch.device:
    use_cuda = cfg.device == "cuda" or (cfg.device == "auto" and torch.cuda.is_available())
    if use_cuda:
        dev = torch.device("cuda", local_rank % torch.cuda.device_count())
        torch.cuda.set_device(dev)
        return dev
    return torch.device("cpu")


def init_group(device: torch.device, rank: int, world: int, timeout_s: float) -> None:
    if world > 1 and not dist.is_initialized():
        backend = "nccl" if device.type == "cuda" else "gloo"
        kw = {"device_id": device} if backend == "nccl" else {}
        dist.init_process_group(backend, rank=rank, world_size=world,
                                timeout=datetime.timedelta(seconds=timeout_s), **kw)


def resolve_transport(cfg: Config, device: torch.device, world: int) -> str:
    t = cfg.transport
    if t == "auto":
        if device.type == "cuda" and pick_backend(cfg, device) == "native":
            return "fused"              # also for world == 1: the same K1 kernel, zero peers
        return "local" if world == 1 else "nccl"
    if t == "fused" and device.type != "cuda":
        raise RuntimeError("the fused transport needs CUDA peer memory; use --transport nccl/host on CPU")
    if world == 1 and (t == "nccl" or (t == "fused" and device.type != "cuda")):
        return "local"                  # one process: the collective rung degenerates to the local optimizer
    return t


class SPMDTrainer:
    def __init__(self, cfg: Config, rank: int = 0, world: int = 1, local_rank: Optional[int] = None,
                 layout_kind: str = "auto", dataset: Optional[Dataset] = None):
        self.cfg = cfg
        self.device = pick_device(cfg, rank if local_rank is None else local_rank)
        init_group(self.device, rank, world, cfg.timeout_s)
        self.layout: RankLayout = make_layout(cfg.num_ps, cfg.num_workers, rank, world, layout_kind)
        self.spec = make_spec(cfg.model, cfg.width_mult)
        self.plan = make_plan(cfg.shard, [v.numel for v in self.spec.vars], cfg.num_ps)
        self.transport_name = resolve_transport(cfg, self.device, world)
        self.seed = 0 if cfg.seed is None else cfg.seed
        self.data = dataset
        if self.layout.is_worker and self.data is None:
            self.data = load_dataset(cfg.data_path, cfg.synthetic, cfg.synthetic_train,
                                     cfg.synthetic_test, cfg.seed)
        self._build()
        self.timers = M.PhaseTimers(self.device, enabled=cfg.profile)
        self.sink = M.MetricsSink(cfg.metrics_path if rank == 0 or not cfg.metrics_path
                                  else cfg.metrics_path + ".%d" % rank)
        self.losses: List[float] = []
        self.global_step = getattr(self, "global_step", 0)
        self._graph = None
        self._graph_warmup = 2

    # ------------------------------------------------------------------------------
    def _build(self) -> None:
        cfg, lay = self.cfg, self.layout
        if self.transport_name == "fused":
            from .transport.fused import FusedTransport
            self.transport = FusedTransport.create(cfg, lay, self.spec, self.plan, self.device)
            self.params, self.grads = self.transport.params, self.transport.grads
        else:
            self.params = Arena(self.spec, self.plan, self.device)
            self.grads = Arena(self.spec, self.plan, self.device)
            if self.transport_name == "local":
                from .transport.collective import LocalTransport as T
            else:
                from .transport.collective import CollectiveTransport as T
            self.transport = None  # created after params are initialised (optimizer state shapes)
            self._T = T
        # owners initialise their ranges (PS weights are the source of truth, Q5); everyone draws
        # the same seeded init so a flat (sub-tensor) owner can fill its slice without messages.
        from ..models import oracle
        init = oracle.init_params(self.spec, self.seed)
        self.params.load_vars(init)
        if cfg.resume:
            # params now; optimizer state after the transport exists
            load_checkpoint(cfg.resume, self.params)
        if self.transport is None:
            self.transport = self._T(cfg, lay, self.params, self.grads)
        if cfg.resume:
            self._load_opt_state(cfg.resume)
        self.model = Model(self.spec, self.params, self.grads, cfg, step_state=getattr(self.transport, "st", None),
                           worker=lay.my_worker or 0, num_workers=lay.num_workers,
                           fc_w_src=getattr(self.transport, "k4_src", None)) if lay.is_worker else None
        if self.model is not None and self.model.backend == "native" and hasattr(self.transport, "attach_engine"):
            self.transport.attach_engine(self.model.impl)
        # the native engine (once attached to the fused transport) launches the step-state advance itself
        self._begin_kw = {"engine": True} if (self.model is not None and getattr(self.transport, "engine_advances", False)) else {}
        self.transport.initial_sync()
        if lay.is_worker:
            self.x_test = self.data.x_test.to(self.device)
            self.y_test = self.data.y_test.to(self.device)

    def _load_opt_state(self, path: str) -> None:
        st = self.transport.opt_state()
        info = load_checkpoint(path, self.params, st.get("m"), st.get("v"))
        self.transport.load_opt_state(st.get("m"), st.get("v"), info["t"])
        self.global_step = info["step"]

    # ------------------------------------------------------------------------------
    def num_batches(self) -> int:
        cfg, lay = self.cfg, self.layout
        if self.data is not None:
            nb = self.data.batches_per_epoch(cfg.batch_size, lay.num_workers, cfg.replicate_data)
        else:
            nb = 0
        if lay.world > 1:                       # PS-only ranks learn the count from the workers
            t = torch.tensor([nb], device=self.device, dtype=torch.int64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            nb = int(t.item())
        if cfg.max_steps:
            nb = min(nb, cfg.max_steps)
        return nb

    def train_step(self, cnt: int) -> Optional[torch.Tensor]:
        """One training step through the public path: H2D of this step's batch from pinned host
        memory, forward/backward, push + update + pull.  Returns the mean loss as a 0-d device
        tensor (None on an owner-only rank)."""
        cfg, lay = self.cfg, self.layout
        if self._graph is None and cfg.graphs and self._graphable and self.global_step >= self._graph_warmup:
            self._capture()
        if self._graph is not None:
            return self._graph_step(cnt)
        loss = None
        self.transport.begin_step(**self._begin_kw)
        if lay.is_worker:
            b, e = self.data.batch_slice(cnt, cfg.batch_size, lay.my_worker, lay.num_workers,
                                         cfg.replicate_data)
            with self.timers.phase("fwd_bwd"):
                x = self.data.x_train[b:e].to(self.device, non_blocking=True)
                y = self.data.y_train[b:e].to(self.device, non_blocking=True)
                loss = self.model.compute_grads(x, y, self.global_step * lay.num_workers + lay.my_worker)
        with self.timers.phase("push_update_pull"):
            self.transport.step()
        self.global_step += 1
        return loss

    # -- CUDA-graph fast path ------------------------------------------------------------
    @property
    def _graphable(self) -> bool:
        if self.cfg.mode == "async" and not self.layout.is_worker:
            return False          # an async PS-only rank starts a resident kernel / launches applies: nothing to replay
        return (self.device.type == "cuda" and self.transport.name == "fused"
                and (self.model is None or self.model.backend == "native") and not self.cfg.profile)

    def _capture(self) -> None:
        """Capture [step-state advance, forward, backward, push+update+pull] once per input slot; every later step
        is one graph launch, and the H2D copy of the NEXT batch (pinned host memory -> the other slot, on a copy
        stream) runs underneath it (phases are 10-100 us, launch latency and the ~20 us of
        serialized H2D + launch gaps would dominate otherwise)."""
        cfg, lay = self.cfg, self.layout
        nslots = 2 if lay.is_worker else 1
        if lay.is_worker:
            self._x_slot = [torch.zeros(cfg.batch_size, 784, device=self.device) for _ in range(nslots)]
            self._y_slot = [torch.zeros(cfg.batch_size, dtype=torch.int32, device=self.device) for _ in range(nslots)]
            self._copy_stream = torch.cuda.Stream(device=self.device)
            self._ev_ready = [torch.cuda.Event() for _ in range(nslots)]     # slot filled by the copy stream
            self._ev_free = [torch.cuda.Event() for _ in range(nslots)]      # slot's last consumer (a graph replay) enqueued
            self._slot_cnt = [None] * nslots                                 # batch index prefetched into each slot
            self._slot_prefetched = [False] * nslots                         # a prefetch was ever issued to the slot
            self._slot = 0
            self._nb_local = max(1, self.data.batches_per_epoch(cfg.batch_size, lay.num_workers, cfg.replicate_data))
        torch.cuda.synchronize()
        from ..ops import native
        self._graphs = []
        for slot in range(nslots):
            l0 = native.launches()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self.transport.begin_step(**self._begin_kw)
                if lay.is_worker:
                    self.model.impl.enqueue_train_step(self._x_slot[slot], self._y_slot[slot])
                self.transport.step()
            self._graphs.append(g)
            if slot == 0:
                self.launches_per_step = native.launches() - l0      # kernels of ours inside one replay
        self._graph = self._graphs[0]
        # the captures themselves did not execute anything; global_step is untouched

    def _h2d(self, slot: int, cnt: int) -> None:
        cfg, lay = self.cfg, self.layout
        b, e = self.data.batch_slice(cnt, cfg.batch_size, lay.my_worker, lay.num_workers, cfg.replicate_data)
        self._x_slot[slot].copy_(self.data.x_train[b:e], non_blocking=True)
        self._y_slot[slot].copy_(self.data.y_train[b:e], non_blocking=True)

    def _graph_step(self, cnt: int) -> Optional[torch.Tensor]:
        cfg, lay = self.cfg, self.layout
        if not lay.is_worker:
            self._graphs[0].replay()
            self.global_step += 1
            return None
        main = torch.cuda.current_stream()
        slot = self._slot
        if self._slot_cnt[slot] == cnt:
            main.wait_event(self._ev_ready[slot])        # prefetched while the previous step ran
        else:
            # first step / out-of-order batch index: copy in line.  A prefetch of some other batch into this slot may
            # still be in flight on the copy stream; order this copy behind it, or the slot could end up holding the
            # prefetched batch
            if self._slot_prefetched[slot]:
                main.wait_event(self._ev_ready[slot])
            self._h2d(slot, cnt)
        self._graphs[slot].replay()
        self._ev_free[slot].record(main)
        # prefetch the next batch of the epoch into the other slot, underneath this step's graph
        nxt, nslot = (cnt + 1) % self._nb_local, slot ^ 1
        self._copy_stream.wait_event(self._ev_free[nslot])    # its last consumer was the previous step (already enqueued)
        with torch.cuda.stream(self._copy_stream):
            self._h2d(nslot, nxt)
            self._ev_ready[nslot].record(self._copy_stream)
        self._slot_cnt[nslot] = nxt
        self._slot_prefetched[nslot] = True
        self._slot_cnt[slot] = None
        self._slot = nslot
        self.global_step += 1
        return self.transport.loss(cfg.batch_size)

    def evaluate(self) -> float:
        if hasattr(self.transport, "ensure_local"):
            self.transport.ensure_local()
        return self.model.accuracy(self.x_test, self.y_test)

    def run(self) -> Dict[str, object]:
        cfg, lay = self.cfg, self.layout
        start_cpu, start_wall = time.process_time(), time.perf_counter()
        nb = self.num_batches()
        lrank = None if lay.kind == "single" and os.environ.get("DDL_ROLE", "") == "single" \
            else lay.logical_worker_rank
        acc = float("nan")
        serve_only = (not lay.is_worker) and cfg.mode == "async" and self.transport.name == "fused"
        if serve_only:
            self.transport.serve_async()
        for epoch in range(cfg.epochs if not serve_only else 0):
            for cnt in range(nb):
                loss = self.train_step(cnt)
                if cfg.debug_checksum and cfg.mode == "sync":
                    from ..utils.verify import replica_check
                    replica_check(self, self.global_step)      # collective: every rank takes part
                if lay.is_worker:
                    if cfg.eval_every and cnt % cfg.eval_every == 0:
                        with self.timers.phase("eval"):
                            acc = self.evaluate()
                        M.log_eval(lrank, epoch, cnt, acc)
                        self.losses.append(float(loss))
                        self.sink.write(rank=lay.rank, epoch=epoch, batch=cnt, accuracy=acc,
                                        loss=self.losses[-1])
        if hasattr(self.transport, "finish"):
            self.transport.finish()
        if self.device.type == "cuda":
            torch.cuda.synchronize()
        if cfg.debug_checksum and cfg.mode == "async":
            from ..utils.verify import async_accounting
            async_accounting(self)
        wall = time.perf_counter() - start_wall
        if lay.is_worker:
            acc = self.evaluate()
            M.log_final(lrank, acc)
            if cfg.log_time:
                M.log_time(time.process_time() - start_cpu)
        stats = {"final_accuracy": acc, "wall_s": wall, "steps": nb * cfg.epochs,
                 "samples_per_s": nb * cfg.epochs * cfg.batch_size * lay.num_workers / max(wall, 1e-9),
                 "phases_ms": self.timers.summary(), "transport": self.transport.name,
                 "backend": self.model.backend if self.model else "none"}
        if hasattr(self.transport, "stats"):
            stats["transport_stats"] = self.transport.stats()
        self.sink.write(rank=lay.rank, **{k: v for k, v in stats.items()})
        if cfg.ckpt_path:
            self.save(cfg.ckpt_path)
        self.sink.close()
        return stats

    def save(self, path: str) -> None:
        if hasattr(self.transport, "ensure_local"):
            self.transport.ensure_local()
        st = self.transport.opt_state()
        if self.layout.rank == 0:
            save_checkpoint(path, self.params, st.get("m"), st.get("v"), int(st.get("t", 0)),
                            self.global_step)
        self.transport.barrier()

    def close(self) -> None:
        self.transport.close()
        if dist.is_initialized():
            dist.destroy_process_group()

"""``fused`` transport: the product.  Push, gradient reduction, optimizer and pull are sm_100a kernels
that read and write the peers' arenas directly over NVLink / NVSwitch (csrc/comm.cu).

What it replaces in the reference (per step and per worker): 14 blocking ``comm.Send`` + 14
``comm.Recv`` of host numpy buffers (mnist_sync_sharding/worker.py:36,94), the PS-side Python loop
with ``numpy +=`` and a ``Session.run`` of ApplyAdam (mnist_sync_sharding/parameter_server.py:76-82,
112-126) and 14 ``tf.assign`` feeds (worker.py:97).  Here:

* every rank allocates its gradient arena, parameter arena, (async) accumulator and a 256-byte
  signal pad with the library's own allocator and publishes cudaIpc handles; every rank maps all of
  them once (``SymmetricMemory``) - the "symmetric, peer-mapped memory" of SURVEY 5.8;
* sync step  = ONE kernel per rank: barrier -> owner reduces its range from all workers' gradient
  arenas in fixed order -> scale -> SGD / momentum / TF1-Adam in place -> stores the new values into
  every worker's parameter arena -> barrier (K1 + K3);
* async step = remote ``red.global.add.v4.f32`` of the gradient into the owners' accumulators with a
  release-ordered push counter (K2), owner-side exactly-once consumption + optimizer + version bump,
  and a barrier-free pull (K3).  Staleness (owner version at push - version at last pull) is
  histogrammed on the device.

All per-step scalars (step count, Adam step size, dropout stream, barrier epochs) live in a
device-resident ``DdlStepState`` so a whole step replays from a CUDA graph.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from ...config import Config
from ...models.spec import ModelSpec
from ...optim import FlatOptimizer
from ...ops import native
from ..arena import Arena
from ..roles import RankLayout
from ..sharding import ShardPlan
from .collective import SpmdTransport

_KIND = {"sgd": 0, "momentum": 1, "adam": 2}


class _RawCuda:
    """Zero-copy torch view of library-allocated device memory (``__cuda_array_interface__``)."""

    def __init__(self, ptr: int, n: int, typestr: str):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}


def _as_tensor(ptr: int, n: int, dtype: torch.dtype, device: torch.device) -> torch.Tensor:
    ts = {torch.float32: "<f4", torch.int32: "<i4"}[dtype]
    return torch.as_tensor(_RawCuda(ptr, n, ts), device=device)


class SymmetricMemory:
    """One buffer per rank, same size everywhere, every rank holds a device pointer to every copy."""

    def __init__(self, nbytes: int, rank: int, world: int, device: torch.device):
        self.L = native.lib()
        self.rank, self.world, self.nbytes, self.device = rank, world, nbytes, device
        p = C.c_void_p()
        native.check(self.L.ddl_dev_malloc(C.byref(p), nbytes), "ddl_dev_malloc")
        self.local = p.value
        self.ptrs: List[int] = [0] * world
        self.ptrs[rank] = self.local
        self.opened: List[int] = []
        if world > 1:
            h = C.create_string_buffer(64)
            native.check(self.L.ddl_ipc_export(self.local, h), "ddl_ipc_export")
            handles: List[Optional[bytes]] = [None] * world
            dist.all_gather_object(handles, h.raw)
            for r in range(world):
                if r == rank:
                    continue
                q = C.c_void_p()
                native.check(self.L.ddl_ipc_open(handles[r], C.byref(q)), "ddl_ipc_open(rank %d)" % r)
                self.ptrs[r] = q.value
                self.opened.append(q.value)

    def tensor(self, dtype=torch.float32) -> torch.Tensor:
        return _as_tensor(self.local, self.nbytes // 4, dtype, self.device)

    def close(self) -> None:
        for q in self.opened:
            self.L.ddl_ipc_close(q)
        self.opened = []
        if self.local:
            self.L.ddl_dev_free(self.local)
            self.local = 0


class MulticastMemory:
    """Same contract as SymmetricMemory, allocated through torch's symmetric-memory allocator, which also
    binds the buffers to an NVSwitch multicast object: ``mc`` is ONE device address naming the same offset
    of every rank's buffer (NVLS).  Used for the parameter and gradient arenas of sync mode."""

    def __init__(self, nbytes: int, rank: int, world: int, device: torch.device):
        import torch.distributed._symmetric_memory as symm
        self.rank, self.world, self.nbytes, self.device = rank, world, nbytes, device
        self.t = symm.empty(nbytes // 4, dtype=torch.float32, device=device)
        self.h = symm.rendezvous(self.t, dist.group.WORLD)
        self.t.zero_()
        self.local = self.t.data_ptr()
        self.ptrs = [int(p) for p in self.h.buffer_ptrs]
        self.mc = int(self.h.multicast_ptr or 0)
        if self.ptrs[rank] != self.local:
            raise native.NativeError("symmetric memory: local buffer pointer mismatch")

    def tensor(self, dtype=torch.float32) -> torch.Tensor:
        return self.t if dtype == torch.float32 else self.t.view(dtype)

    def close(self) -> None:
        self.h = None
        self.t = None


def _alloc_arena(nbytes: int, rank: int, world: int, device: torch.device, want_mc: str):
    """Multicast-capable allocation when asked for (and possible), else plain cudaIpc peer mapping.

    Every rank takes the same path: a constructor failure on one rank is caught locally and turned into a vote, and
    the MIN all-reduce of the votes always runs, so no rank can skip a collective the others are waiting in."""
    if world > 1 and want_mc in ("auto", "on"):
        m, why = None, ""
        try:
            m = MulticastMemory(nbytes, rank, world, device)
            if not m.mc:
                why = "no multicast pointer"
        except Exception as e:                               # noqa: BLE001 - any failure means "no NVLS on this rank"
            m, why = None, repr(e)
        ok = torch.tensor([1 if (m is not None and m.mc) else 0], device=device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)           # all ranks or none
        if int(ok.item()) == 1:
            return m
        if m is not None:
            m.close()
        if want_mc == "on":
            raise native.NativeError("NVLS multicast requested (--nvls on) but unavailable on rank %d: %s"
                                     % (rank, why or "another rank has no multicast pointer"))
        if rank == 0:
            import warnings
            warnings.warn("ddl_b200: NVLS multicast was selected automatically for %d ranks but is unavailable (%s); "
                          "falling back to peer loads / stores over cudaIpc mappings" % (world, why or "peer rank"),
                          RuntimeWarning, stacklevel=2)
    return SymmetricMemory(nbytes, rank, world, device)


class FusedTransport(SpmdTransport):
    name = "fused"

    @classmethod
    def create(cls, cfg: Config, layout: RankLayout, spec: ModelSpec, plan: ShardPlan,
               device: torch.device) -> "FusedTransport":
        L = native.lib()
        native.check(L.ddl_set_device(device.index or 0), "ddl_set_device")
        world = layout.world
        if world > native.MAX_PEERS:
            raise native.NativeError("fused transport supports at most %d ranks" % native.MAX_PEERS)
        if world > 1:
            for r in range(torch.cuda.device_count()):
                if r != (device.index or 0) and r < world and not L.ddl_can_access_peer(device.index or 0, r):
                    raise native.NativeError("GPU %d cannot access GPU %d: no NVLink/P2P path" % (device.index or 0, r))
        nbytes = plan.total * 4
        want_mc = cfg.nvls if cfg.mode == "sync" else "off"
        if want_mc == "auto" and world < 4:
            # bytes per rank and direction: peer loads/stores 2(W-1)/W arenas, multimem (1 + 1/W) arenas - the switch
            # path only moves less for W > 3 (measured at W = 2: 1.03 ms vs 0.67 ms for a 381 MB arena,
            # profiles/comm_roofline.md)
            want_mc = "off"
        mem = {"params": _alloc_arena(nbytes, layout.rank, world, device, want_mc),
               "grads": _alloc_arena(nbytes, layout.rank, world, device, want_mc),
               "pad": SymmetricMemory(native.PAD_SLOTS * 4, layout.rank, world, device)}
        if cfg.mode == "async":
            # one accumulator plane per worker on every rank (only owned ranges are ever touched): the owner can then
            # apply one optimizer step per push instead of one step on the sum of whatever has landed
            mem["accum"] = SymmetricMemory(nbytes * max(1, layout.num_workers), layout.rank, world, device)
        params = Arena(spec, plan, device, mem["params"].tensor())
        grads = Arena(spec, plan, device, mem["grads"].tensor())
        return cls(cfg, layout, params, grads, mem)

    def __init__(self, cfg, layout, params, grads, mem: Dict[str, SymmetricMemory]):
        super().__init__(cfg, layout, params, grads)
        self.L = native.lib()
        self.mem = mem
        dev = params.device
        lay = layout
        self.kind = _KIND[cfg.optimizer]
        # optimizer state for my ranges, indexed like the arena (only owned ranges are touched)
        owns = bool(lay.my_shards)
        self.m = torch.zeros(self.plan.total, device=dev) if owns and self.kind >= 1 else None
        self.v = torch.zeros(self.plan.total, device=dev) if owns and self.kind == 2 else None
        seed = 0 if cfg.seed is None else cfg.seed
        self.st = native.new_step_state(seed, lay.my_worker or 0, lay.num_workers)
        self.scratch = torch.zeros(64, dtype=torch.int32, device=dev)      # done counters, applied/pending ...
        self.pulled_version = torch.zeros(native.MAX_PEERS, dtype=torch.int32, device=dev)
        self.stale_hist = torch.zeros(native.STALE_BINS, dtype=torch.int32, device=dev)
        self.comm = native.Comm()
        self.comm.rank, self.comm.world = lay.rank, lay.world
        for r in range(lay.world):
            self.comm.grads[r] = mem["grads"].ptrs[r]
            self.comm.params[r] = mem["params"].ptrs[r]
            self.comm.pads[r] = mem["pad"].ptrs[r]
            self.comm.accum[r] = mem["accum"].ptrs[r] if "accum" in mem else None
        self.nvls = bool(getattr(mem["grads"], "mc", 0) and getattr(mem["params"], "mc", 0))
        self.comm.mc_grads = mem["grads"].mc if self.nvls else None
        self.comm.mc_params = mem["params"].mc if self.nvls else None
        self.worker_mask = 0
        for wr in lay.worker_ranks:
            self.worker_mask |= 1 << wr
        # per-owned-shard argument blocks.  With gradient bucketing every step is served by three launches per owned
        # shard, in the same order on every rank (see bucket_vars below).
        spec = params.spec
        self.bucketed = bool(cfg.overlap_push and cfg.mode == "sync" and spec.kind == "cnn")
        self.engine_overlap = False
        self.engine_advances = False
        kind_id, step0, aa, bb, eps = FlatOptimizer(torch.zeros(1), cfg.optimizer, cfg.lr, cfg.beta1, cfg.beta2,
                                                    cfg.eps, cfg.momentum, use_native=False).hyper(1)

        def make_args(begin: int, end: int, bucket: int, nbuckets: int, subs) -> native.SyncArgs:
            a = native.SyncArgs()
            a.begin, a.end = begin, end
            a.m = self.m.data_ptr() if self.m is not None else None
            a.v = self.v.data_ptr() if self.v is not None else None
            a.st = self.st.data_ptr()
            a.done_counter = self.scratch[16 + bucket:].data_ptr()
            a.worker_mask = a.dest_mask = self.worker_mask
            a.lr, a.a, a.b, a.eps = cfg.lr, aa, bb, eps
            a.gscale = self.gscale if cfg.mode == "sync" else 1.0
            for k in range(native.MAX_SKIP):
                a.skip_begin[k] = a.skip_end[k] = -1
            a.bucket, a.nbuckets = bucket, nbuckets
            a.nsub = 0
            if subs is not None:
                a.nsub = len(subs)
                for i, (sb, se) in enumerate(subs):
                    a.sub_begin[i], a.sub_end[i] = sb, se
                if not subs:                          # nothing of this bucket lives here: barrier-only launch
                    a.nsub = 1
                    a.sub_begin[0] = a.sub_end[0] = begin
            return a

        def bucket_subranges(begin: int, end: int, var_ids):
            """Arena sub-ranges of [begin, end) covered by `var_ids` (alignment padding included), merged."""
            from ..sharding import ALIGN_ELEMS
            spans = []
            for v in var_ids:
                b = self.plan.var_offset[v]
                e = b + (spec.vars[v].numel + ALIGN_ELEMS - 1) // ALIGN_ELEMS * ALIGN_ELEMS
                b, e = max(b, begin), min(e, end)
                if b < e:
                    spans.append((b, e))
            spans.sort()
            merged = []
            for b, e in spans:
                if merged and merged[-1][1] == b:
                    merged[-1] = (merged[-1][0], e)
                else:
                    merged.append((b, e))
            return merged

        # buckets: FC layers (60% of the bytes, complete right after the FC backward), conv4, conv3 (each complete once
        # that layer's dgrad and wgrad have run), conv1 / conv2 (after the backward).  The engine launches the first
        # three in the middle of the backward, behind the kernels that produce their gradients; step() launches the
        # last one.  Every bucket has its own signal row, so bucket kernels are independent of each other.
        bucket_vars = [list(range(8, spec.num_vars)), [6, 7], [4, 5], [0, 1, 2, 3]]
        nbk = len(bucket_vars)
        self.sync_args: List[native.SyncArgs] = []        # launched by step()
        self.early_args: List[native.SyncArgs] = []       # overlapped buckets (engine side stream, or step() without engine)
        ranges = [self.plan.shard_range[r] for r in lay.my_shards] or [(0, 0)]
        if self.bucketed and (len(ranges) * (nbk - 1) > 8 or spec.num_vars != 14):
            self.bucketed = False
        all_subs = [[bucket_subranges(b, e, vs) for vs in bucket_vars] for (b, e) in ranges] if self.bucketed else []
        if any(len(x) > native.MAX_SUB for subs in all_subs for x in subs):
            self.bucketed = False
        for i, (b, e) in enumerate(ranges):
            if self.bucketed:
                for k in range(nbk - 1):
                    self.early_args.append(make_args(b, e, k, nbk, all_subs[i][k]))
                self.sync_args.append(make_args(b, e, nbk - 1, nbk, all_subs[i][nbk - 1]))
            else:
                self.sync_args.append(make_args(b, e, 0, 1, None))
        self.early_args.sort(key=lambda a: a.bucket)      # every rank launches bucket 0 before bucket 1
        if len(ranges) > 1:                               # single-process multi-shard: one launch per shard is fine
            pass
        # ---- K4: variables whose pull is fused into their first consumer GEMM (fc1 / fc2 weights: each weight
        # byte is read exactly once per step by the [B<=128 x K] x [K x N] matmul, so fetching the tiles
        # straight from the owner over NVLink costs the same bytes as an all-gather and removes it).
        # Their owners do not broadcast those ranges; the consumer writes them through to its local arena.
        self.k4_src: Dict[int, Optional[int]] = {}
        if (cfg.fused_pull_gemm and cfg.mode == "sync" and lay.world > 1 and self.plan.policy != "flat"
                and spec.kind == "cnn"):
            k = 0
            for v in (8, 10):
                owner_rank = lay.owner_rank(self.plan.owner[v])
                b = self.plan.var_offset[v]
                e = b + spec.vars[v].numel
                self.k4_src[v] = None if owner_rank == lay.rank else mem["params"].ptrs[owner_rank] + 4 * b
                if owner_rank == lay.rank:
                    for a in self.sync_args + self.early_args:
                        if a.begin <= b and e <= a.end:
                            a.skip_begin[k], a.skip_end[k] = b, e
                k += 1
        self.async_args = native.AsyncArgs()
        aa = self.async_args
        aa.num_shards = lay.num_ps
        for r in range(lay.num_ps):
            aa.owner_rank[r] = lay.owner_rank(r)
            aa.shard_begin[r], aa.shard_end[r] = self.plan.shard_range[r]
        aa.direct_sgd = 1 if (cfg.mode == "async" and cfg.optimizer == "sgd") else 0
        aa.lr = cfg.lr
        aa.done_counter = self.scratch[1:].data_ptr()
        aa.pulled_version = self.pulled_version.data_ptr()
        aa.stale_hist = self.stale_hist.data_ptr()
        aa.slot = lay.my_worker or 0
        aa.slot_stride = self.plan.total
        self.async_state = torch.zeros(C.sizeof(native.AsyncState) // 4, dtype=torch.int32, device=dev)
        self.barrier_epoch = self.scratch[8:]

    # ------------------------------------------------------------------------------
    def _s(self) -> int:
        return native.stream_ptr()

    def barrier(self) -> None:
        """Host-requested device-side barrier across all ranks (one tiny kernel per rank)."""
        if self.layout.world > 1:
            native.check(self.L.ddl_device_barrier(C.byref(self.comm), self.barrier_epoch.data_ptr(), self._s()),
                         "ddl_device_barrier")
            native.count()

    def pull_all(self) -> None:
        native.check(self.L.ddl_pull_allgather(C.byref(self.comm), C.byref(self.async_args), self._s()),
                     "ddl_pull_allgather")
        native.count()

    def ensure_local(self) -> None:
        """Make this rank's parameter arena complete and current (evaluation, checkpoints, inspection):
        with K4 the fc1 / fc2 weights are only refreshed by the next training forward."""
        if self.k4_src and self.layout.world > 1:
            self.pull_all()

    def initial_sync(self) -> None:
        """Initial pull (Q5): every rank copies each owner's range; owners are the source of truth."""
        if self.layout.world > 1:
            torch.cuda.synchronize()
            dist.barrier()
            self.pull_all()
            torch.cuda.synchronize()
            dist.barrier()

    def begin_step(self, engine: bool = False) -> None:
        """Advances the device step state.  ``engine=True``: the native engine's training step follows and (once
        attached) launches the advance itself, beside its convolutions."""
        c = self.cfg
        if engine and self.engine_advances:
            native.count()
            return
        native.check(self.L.ddl_step_advance(self.st.data_ptr(), c.lr, c.beta1, c.beta2,
                                             1 if c.mode == "sync" else 0, self._s()), "ddl_step_advance")
        native.count()

    def attach_engine(self, backend) -> None:
        """Hand bucket 0 to the native engine: it launches those push/update/pull kernels on its side stream in
        the middle of the backward (as soon as conv3..fc3 gradients exist), overlapping them with the conv2 /
        conv1 backward - SURVEY 7.2 step 7, "backward-overlapped push"."""
        c = self.cfg
        native.check(self.L.ddl_engine_set_advance(backend.h, c.lr, c.beta1, c.beta2, 1 if c.mode == "sync" else 0),
                     "ddl_engine_set_advance")
        self.engine_advances = True
        if not (self.bucketed and self.early_args):
            return
        arr = (native.SyncArgs * len(self.early_args))(*self.early_args)
        native.check(self.L.ddl_engine_set_overlap(backend.h, C.byref(self.comm), arr, len(self.early_args), self.kind),
                     "ddl_engine_set_overlap")
        self.engine_overlap = True
        backend.overlap_launches = len(self.early_args)

    def step(self, all_buckets: bool = False) -> None:
        """push + update + pull of one step.  ``all_buckets`` also launches the bucket the engine normally
        launches itself in the middle of the backward (transport-only tests, owner-only ranks)."""
        if self.cfg.mode == "sync":
            if self.bucketed and (all_buckets or not self.engine_overlap):
                for a in self.early_args:            # no engine on this rank (owner-only) or torch backend
                    native.check(self.L.ddl_sync_push_update_pull(C.byref(self.comm), C.byref(a), self.kind, self._s()),
                                 "ddl_sync_push_update_pull")
                    native.count()
            for a in self.sync_args:
                native.check(self.L.ddl_sync_push_update_pull(C.byref(self.comm), C.byref(a), self.kind, self._s()),
                             "ddl_sync_push_update_pull")
                native.count()
        else:
            if self.layout.is_worker:
                native.check(self.L.ddl_async_push(C.byref(self.comm), C.byref(self.async_args), self._s()),
                             "ddl_async_push")
                native.count()
            self.apply_pending()
            if self.layout.is_worker:
                self.pull_all()
        self.step_count += 1

    # -- resident owner (dedicated async PS GPU) ---------------------------------------------------
    def _resident_eligible(self) -> bool:
        lay, c = self.layout, self.cfg
        return (c.mode == "async" and c.resident_owner and lay.kind == "dedicated" and not lay.is_worker
                and len(lay.my_shards) == 1 and not self.async_args.direct_sgd and lay.world > 1)

    def start_resident(self) -> None:
        """Launches the resident owner kernel (csrc/comm.cu) on a stream of its own; it serves pushes until
        ``stop_resident`` writes the pinned stop word."""
        if getattr(self, "_resident", None) is not None:
            return
        c = self.cfg
        dev = self.params.device
        self._res_stop = torch.zeros(1, dtype=torch.int32).pin_memory()
        res = torch.zeros(6, dtype=torch.int32, device=dev)         # DdlResident: 4 words + the host pointer
        ptr = self._res_stop.data_ptr()
        res[4] = int(C.c_int32(ptr & 0xFFFFFFFF).value)
        res[5] = int(C.c_int32((ptr >> 32) & 0xFFFFFFFF).value)
        self._res_dev = res
        self._res_stream = torch.cuda.Stream(device=dev)
        torch.cuda.current_stream().synchronize()                   # state tensors are initialised before the kernel polls
        native.check(self.L.ddl_async_owner_start(C.byref(self.comm), C.byref(self.sync_args[0]), self.kind,
                                                  self.async_state.data_ptr(), res.data_ptr(), self.layout.num_workers,
                                                  self.plan.total, 1 if c.async_fold else 0, c.beta1, c.beta2,
                                                  self._res_stream.cuda_stream), "ddl_async_owner_start")
        self._resident = True

    def stop_resident(self) -> None:
        if getattr(self, "_resident", None):
            self._res_stop[0] = 1
            self._res_stream.synchronize()
            self._resident = None

    def apply_pending(self) -> None:
        """Owner side of async mode: consume whatever has been pushed since the last call."""
        if self.async_args.direct_sgd or not self.layout.my_shards:
            return
        if getattr(self, "_resident", None):
            return                               # the resident kernel is serving; two consumers must not race
        if self._resident_eligible() and not getattr(self, "_resident_done", False):
            self.start_resident()
            return
        c = self.cfg
        for a in self.sync_args:
            native.check(self.L.ddl_async_apply(C.byref(self.comm), C.byref(a), self.kind, self.async_state.data_ptr(),
                                                self.layout.num_workers, self.plan.total, 1 if c.async_fold else 0,
                                                c.beta1, c.beta2, self._s()), "ddl_async_apply")
            native.count(2)

    def serve_async(self, poll_every: int = 8) -> int:
        """Dedicated async PS rank (mnist_async/parameter_server.py:52-69 without the host loop body):
        keep consuming pushes until every worker has announced completion through the store."""
        store = dist.distributed_c10d._get_default_store()
        keys = ["ddl_done_%d" % wr for wr in self.layout.worker_ranks]
        rounds = 0
        if self._resident_eligible():
            import time
            self.start_resident()
            while not store.check(keys):         # the device serves; the host only waits for the workers to finish
                time.sleep(0.002)
            return 0
        while True:
            for _ in range(poll_every):
                self.apply_pending()
                rounds += 1
            torch.cuda.current_stream().synchronize()
            if store.check(keys):
                break
        return rounds

    def finish(self) -> None:
        """End of an async run: drain every accumulator and give all workers the final weights."""
        if self.cfg.mode != "async" or self.layout.world == 1:
            if self.cfg.mode == "async":
                self.apply_pending()
            return
        store = dist.distributed_c10d._get_default_store()
        if self.layout.is_worker:
            torch.cuda.synchronize()
            store.set("ddl_done_%d" % self.layout.rank, "1")
        if getattr(self, "_resident", None):
            # No collective while the resident kernel runs (a NCCL barrier synchronises the device and would never
            # return): wait for the workers' completion keys through the store, then let the kernel serve what is
            # pending and stop it.  From here on the launch-based apply drains anything that is left.
            import time
            keys = ["ddl_done_%d" % wr for wr in self.layout.worker_ranks]
            while not store.check(keys):
                time.sleep(0.001)
        self._resident_done = True
        self.stop_resident()
        torch.cuda.synchronize()
        dist.barrier()                       # all pushes have landed
        self.apply_pending()
        torch.cuda.synchronize()
        dist.barrier()                       # all owners have applied
        if self.layout.is_worker:
            self.pull_all()
        torch.cuda.synchronize()

    def _sync(self) -> None:
        """Device-wide synchronize, except while the resident owner kernel runs (it ends only on request)."""
        if getattr(self, "_resident", None):
            torch.cuda.current_stream().synchronize()
        else:
            torch.cuda.synchronize()

    # ------------------------------------------------------------------------------
    def loss(self, batch: int) -> torch.Tensor:
        """Mean loss of the latest step: a 0-d VIEW of the device step state (word 9, written by the fc3 kernels
        with their 1 / batch scale) - no kernel launch; read or copy it before the next step overwrites it."""
        return self.st.view(torch.float32)[9]

    def stats(self) -> Dict[str, object]:
        s = native.read_step_state(self.st)
        out = {"optimizer_steps": s["t"], "worker_steps": s["wstep"], "nvls": getattr(self, "nvls", False)}
        if self.cfg.mode == "async":
            out["staleness_hist"] = self.stale_hist.cpu().tolist()
            pad = self.mem["pad"].tensor(torch.int32).cpu()
            out["pushes_received"] = int(pad[native.PAD_PUSHES])
            out["shard_version"] = int(pad[native.PAD_VERSION])
            # pushes that shared an optimizer step with another push: a worker that pushed twice between two applies
            # (its plane holds the sum), or every extra push under --async-fold.  optimizer_steps + folded == pushes.
            out["folded_pushes"] = int(self.async_state[native.MAX_PEERS + 2].item())
            if getattr(self, "_res_dev", None) is not None:
                out["resident_rounds"] = int(self._res_dev[3].item())     # rounds served by the resident owner kernel
        return out

    def opt_state(self):
        lay = self.layout
        m = self.m if self.m is not None else torch.zeros_like(self.params.flat)
        v = self.v if self.v is not None else torch.zeros_like(self.params.flat)
        t = torch.tensor([native.read_step_state(self.st)["t"]], device=self.params.device)
        if lay.world > 1:
            m, v = m.clone(), v.clone()
            for r in range(lay.num_ps):
                dist.broadcast(self.params.shard(r, m), src=lay.owner_rank(r))
                dist.broadcast(self.params.shard(r, v), src=lay.owner_rank(r))
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return {"m": m, "v": v, "t": int(t.item())}

    def load_opt_state(self, m, v, t) -> None:
        if self.m is not None and m is not None and m is not self.m:
            self.m.copy_(m)
        if self.v is not None and v is not None and v is not self.v:
            self.v.copy_(v)
        self.st[0] = int(t)

    def close(self) -> None:
        self.stop_resident()
        torch.cuda.synchronize()
        if self.layout.world > 1 and dist.is_initialized():
            dist.barrier()
        for mem in self.mem.values():
            mem.close()
        self.mem = {}

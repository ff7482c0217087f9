"""Configuration layer.

The reference hard-codes every hyper-parameter (``epoch = 1``, ``batch_size = 100``
at mnist_sync/worker.py:41-42 and mnist_sync/parameter_server.py:42-43; Adam lr 1e-4 at
mnist_sync/model/model.py:93; keep_prob 0.5 at mnist_sync/worker.py:30; eval every 10
batches at mnist_sync/worker.py:71).  Its only knobs are the two positional ``run.sh``
arguments forwarded as ``-np N`` (mnist_sync_sharding/run.sh:3) and read as
``sys.argv[2]`` (mnist_sync_sharding/worker.py:65).

Here the same constants are the *defaults* of one dataclass, every entry point accepts the
reference's ``-np N`` flag, and everything else is an optional flag or ``DDL_*`` env var.
"""
from __future__ import annotations

import argparse
import dataclasses
import os
from dataclasses import dataclass, field
from typing import Optional, Sequence

SHARD_MODES = ("none", "block", "round_robin", "greedy", "greedy_interleave", "flat")
TRANSPORTS = ("auto", "fused", "nccl", "host", "local")
OPTIMIZERS = ("adam", "sgd", "momentum")
REDUCES = ("sum", "mean")
BACKENDS = ("auto", "native", "torch")

# variant directory name -> (mode, shard policy).  One preset per reference directory.
VARIANTS = {
    "mnist_sync": ("sync", "none"),
    "mnist_async": ("async", "none"),
    "mnist_sync_sharding": ("sync", "block"),
    "mnist_sync_sharding_greedy": ("sync", "greedy"),
    "mnist_async_sharding": ("async", "block"),
    "mnist_async_sharding_greedy": ("async", "greedy"),
}


@dataclass
class Config:
    # --- reference constants (defaults equal the reference's literals) -------------------
    epochs: int = 1                 # mnist_sync/worker.py:41
    batch_size: int = 100           # mnist_sync/worker.py:42
    lr: float = 1e-4                # mnist_sync/model/model.py:93
    keep_prob: float = 0.5          # mnist_sync/worker.py:30
    eval_every: int = 10            # mnist_sync/worker.py:71 (0 disables)
    # --- variant -----------------------------------------------------------------------
    variant: str = "mnist_sync"
    mode: str = "sync"              # sync | async
    shard: str = "none"
    num_ps: int = 1
    num_workers: int = 1
    # --- framework knobs ---------------------------------------------------------------
    optimizer: str = "adam"
    momentum: float = 0.9
    beta1: float = 0.9
    beta2: float = 0.999
    eps: float = 1e-8
    reduce: str = "sum"             # reference sums, never averages (mnist_sync/parameter_server.py:36-38)
    transport: str = "auto"
    backend: str = "auto"           # compute backend: native sm_100a kernels or torch ops
    precision: str = "tf32"         # tensor-core input kind for native GEMM/conv: tf32 (fp32 storage)
    synthetic: bool = False
    synthetic_train: int = 6000     # rows of synthetic MNIST-shaped data when --synthetic
    synthetic_test: int = 1000
    data_path: str = "data/mnist.pkl"   # mnist_sync/model/model.py:8
    width_mult: int = 1             # widened model (BASELINE.json config 5)
    model: str = "cnn"              # cnn | mlp
    seed: Optional[int] = None      # reference sets no seed
    replicate_data: bool = False    # reference feeds every worker the same slice (mnist_sync/worker.py:27-28)
    max_steps: int = 0              # 0 = full epoch
    profile: bool = False
    metrics_path: str = ""
    ckpt_path: str = ""
    resume: str = ""
    quirk_parity: bool = False      # reproduce Q1/Q2 gradient-sum bugs (off: true sum)
    device: str = "auto"            # auto | cpu | cuda
    graphs: bool = True             # capture the step in a CUDA graph when possible
    fused_pull_gemm: bool = False   # K4: fc1/fc2 forward GEMMs TMA-load their weights straight from the owner's
                                    # arena (pull fused into the consumer).  Off by default: measured slower than
                                    # the store-based pull fused into the update kernel
    nvls: str = "auto"              # auto | on | off: NVSwitch multicast (multimem.ld_reduce / multimem.st) for the sync push / pull
    async_fold: bool = False        # async + fused: one optimizer step on the SUM of all pending pushes (round-1 behaviour)
                                    # instead of one step per push (the reference's semantics, mnist_async/parameter_server.py:55-69)
    resident_owner: bool = True     # async + fused, dedicated PS GPU: serve pushes from ONE resident kernel instead of host-launched applies
    overlap_push: bool = True       # sync + fused: push/update/pull conv3..fc3 (98% of the bytes) while conv2/conv1 backward runs
    debug_checksum: bool = False    # integrity checks: host transport - every push / pull message group carries an exact
                                    # checksum that the receiver verifies; SPMD trainers - replicas compared bitwise after every
                                    # sync step, reduced gradients checked against the workers' sums, async pushes accounted for
    fault_inject: str = ""          # "" | corrupt_push:<step> | corrupt_pull:<step> | drop_rank:<step> - deliberate faults for the
                                    # failure / race-detection tests (host transport; the first worker / first PS misbehaves)
    timeout_s: float = 120.0        # collective / rendezvous timeout (failure detection)
    log_time: bool = True           # print the reference's "Time:" line

    def validate(self) -> "Config":
        if self.shard not in SHARD_MODES:
            raise ValueError(f"shard must be one of {SHARD_MODES}, got {self.shard!r}")
        if self.transport not in TRANSPORTS:
            raise ValueError(f"transport must be one of {TRANSPORTS}")
        if self.optimizer not in OPTIMIZERS:
            raise ValueError(f"optimizer must be one of {OPTIMIZERS}")
        if self.reduce not in REDUCES:
            raise ValueError(f"reduce must be one of {REDUCES}")
        if self.mode not in ("sync", "async"):
            raise ValueError("mode must be sync or async")
        if self.num_ps < 1 or self.num_workers < 1:
            raise ValueError("need at least one PS shard and one worker")
        if self.shard == "none" and self.num_ps != 1:
            # The unsharded reference variants hard-code a single PS
            # (mnist_sync/parameter_server.py:47, mnist_async/parameter_server.py:43).
            raise ValueError("unsharded variants support exactly one parameter server (Q8)")
        if self.batch_size < 1:
            raise ValueError("batch_size must be positive")
        if self.fault_inject:
            kind, _, at = self.fault_inject.partition(":")
            if kind not in ("corrupt_push", "corrupt_pull", "drop_rank") or not at.isdigit():
                raise ValueError("fault_inject must be corrupt_push:<step>, corrupt_pull:<step> or drop_rank:<step>")
        return self

    def fault_at(self, kind: str) -> int:
        """step at which the fault `kind` fires, or -1"""
        k, _, at = self.fault_inject.partition(":")
        return int(at) if k == kind else -1

    def replace(self, **kw) -> "Config":
        return dataclasses.replace(self, **kw)


def _env_default(name: str, default):
    v = os.environ.get("DDL_" + name.upper())
    if v is None:
        return default
    if isinstance(default, bool):
        return v.lower() in ("1", "true", "yes", "on")
    if isinstance(default, int):
        return int(v)
    if isinstance(default, float):
        return float(v)
    return v


def build_parser(role: str = "worker") -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(prog=role, description="ddl_b200 %s entry point" % role)
    d = Config()
    # reference knob: `-np N` (number of PS for parameter_server.py, of workers for worker.py)
    p.add_argument("-np", dest="np", type=int, default=None,
                   help="reference knob: #PS (parameter_server.py) or #workers (worker.py)")
    p.add_argument("--variant", default=None)
    p.add_argument("--num-ps", type=int, default=None)
    p.add_argument("--num-workers", type=int, default=None)
    p.add_argument("--epochs", type=int, default=_env_default("epochs", d.epochs))
    p.add_argument("--batch-size", type=int, default=_env_default("batch_size", d.batch_size))
    p.add_argument("--lr", type=float, default=_env_default("lr", d.lr))
    p.add_argument("--keep-prob", type=float, default=_env_default("keep_prob", d.keep_prob))
    p.add_argument("--eval-every", type=int, default=_env_default("eval_every", d.eval_every))
    p.add_argument("--optimizer", choices=OPTIMIZERS, default=_env_default("optimizer", d.optimizer))
    p.add_argument("--momentum", type=float, default=d.momentum)
    p.add_argument("--shard", choices=SHARD_MODES, default=None)
    p.add_argument("--mode", choices=("sync", "async"), default=None)
    p.add_argument("--reduce", choices=REDUCES, default=_env_default("reduce", d.reduce))
    p.add_argument("--transport", choices=TRANSPORTS, default=_env_default("transport", d.transport))
    p.add_argument("--backend", choices=BACKENDS, default=_env_default("backend", d.backend))
    p.add_argument("--precision", choices=("tf32",), default=d.precision)
    p.add_argument("--synthetic", action="store_true", default=_env_default("synthetic", d.synthetic))
    p.add_argument("--synthetic-train", type=int, default=_env_default("synthetic_train", d.synthetic_train))
    p.add_argument("--synthetic-test", type=int, default=_env_default("synthetic_test", d.synthetic_test))
    p.add_argument("--data-path", default=_env_default("data_path", d.data_path))
    p.add_argument("--width-mult", type=int, default=_env_default("width_mult", d.width_mult))
    p.add_argument("--model", choices=("cnn", "mlp"), default=d.model)
    p.add_argument("--seed", type=int, default=None)
    p.add_argument("--replicate-data", action="store_true", default=_env_default("replicate_data", False))
    p.add_argument("--max-steps", type=int, default=_env_default("max_steps", 0))
    p.add_argument("--profile", action="store_true")
    p.add_argument("--metrics-path", default=_env_default("metrics_path", ""))
    p.add_argument("--ckpt-path", default=_env_default("ckpt_path", ""))
    p.add_argument("--resume", default=_env_default("resume", ""))
    p.add_argument("--quirk-parity", action="store_true")
    p.add_argument("--device", choices=("auto", "cpu", "cuda"), default=_env_default("device", d.device))
    p.add_argument("--no-graphs", dest="graphs", action="store_false", default=True)
    p.add_argument("--fused-pull-gemm", dest="fused_pull_gemm", action="store_true",
                   default=_env_default("fused_pull_gemm", False))
    p.add_argument("--nvls", choices=("auto", "on", "off"), default=_env_default("nvls", "auto"))
    p.add_argument("--no-overlap-push", dest="overlap_push", action="store_false", default=True)
    p.add_argument("--no-resident-owner", dest="resident_owner", action="store_false", default=True,
                   help="async dedicated PS: host-launched apply kernels instead of the resident owner kernel")
    p.add_argument("--async-fold", dest="async_fold", action="store_true", default=_env_default("async_fold", False),
                   help="async + fused transport: fold all pushes pending at an apply into ONE optimizer step")
    p.add_argument("--debug-checksum", dest="debug_checksum", action="store_true",
                   default=_env_default("debug_checksum", False),
                   help="verify every push / pull (host transport) and compare the replicas after every sync step (SPMD)")
    p.add_argument("--fault-inject", dest="fault_inject", default=_env_default("fault_inject", ""),
                   help="corrupt_push:<step> | corrupt_pull:<step> | drop_rank:<step> (tests of the integrity / failure checks)")
    p.add_argument("--timeout-s", type=float, default=_env_default("timeout_s", d.timeout_s))
    return p


def config_from_args(argv: Optional[Sequence[str]], role: str, variant: Optional[str] = None) -> Config:
    """Build a Config for one role process.

    ``-np`` keeps the reference meaning: for ``parameter_server.py`` it is the number of PS
    ranks, for ``worker.py`` the number of workers; the other count is world size minus it
    (mnist_sync_sharding/parameter_server.py:95-96, mnist_sync_sharding/worker.py:65-66).
    """
    ns = build_parser(role).parse_args(argv)
    variant = ns.variant or variant or "mnist_sync"
    mode, shard = VARIANTS.get(variant, ("sync", "none"))
    mode = ns.mode or mode
    shard = ns.shard or shard
    world = int(os.environ.get("DDL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "0")) or 0)
    num_ps, num_workers = ns.num_ps, ns.num_workers
    if role == "ps" and ns.np is not None:
        num_ps = ns.np
    if role == "worker" and ns.np is not None:
        num_workers = ns.np
    if num_ps is None and "DDL_NUM_PS" in os.environ:
        num_ps = int(os.environ["DDL_NUM_PS"])
    if num_workers is None and "DDL_NUM_WORKERS" in os.environ:
        num_workers = int(os.environ["DDL_NUM_WORKERS"])
    if num_ps is None and num_workers is not None and world:
        num_ps = world - num_workers
    if num_workers is None and num_ps is not None and world:
        num_workers = world - num_ps
    if num_ps is None:
        num_ps = 1
    if num_workers is None:
        num_workers = max(1, world - num_ps) if world else 1
    fields = {f.name for f in dataclasses.fields(Config)}
    kw = {k: v for k, v in vars(ns).items() if k in fields and v is not None}
    kw.update(variant=variant, mode=mode, shard=shard, num_ps=num_ps, num_workers=num_workers)
    return Config(**kw).validate()

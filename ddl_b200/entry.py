"""Role entry points behind the six variant directories' ``parameter_server.py`` / ``worker.py`` /
``single.py`` shims (the reference's ``__main__`` blocks: mnist_sync/parameter_server.py:41-69,
mnist_sync/worker.py:40-76, mnist_sync/single.py:24-26).

A role process learns its rank from the launcher's environment (``RANK`` / ``WORLD_SIZE``, set by
``ddl_b200.parallel.launch`` or torchrun - the reference got them from ``MPI.COMM_WORLD``) and its
role count from the reference's ``-np N`` flag.
"""
from __future__ import annotations

import os
import sys
from typing import Optional, Sequence

from .config import Config, config_from_args


def _rank_world():
    rank = int(os.environ.get("DDL_RANK", os.environ.get("RANK", "0")))
    world = int(os.environ.get("DDL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1")))
    local = int(os.environ.get("LOCAL_RANK", rank))
    return rank, world, local


def _transport_for(cfg: Config) -> str:
    if cfg.transport != "auto":
        return cfg.transport
    import torch
    want_cuda = cfg.device == "cuda" or (cfg.device == "auto" and torch.cuda.is_available())
    return "fused" if want_cuda else "host"


def run_role(role: str, variant: str, argv: Optional[Sequence[str]] = None):
    """role in {"ps", "worker"}.  Returns the trainer / worker stats."""
    cfg = config_from_args(sys.argv[1:] if argv is None else argv, role, variant)
    rank, world, local = _rank_world()
    transport = _transport_for(cfg)
    cfg = cfg.replace(transport=transport)
    if transport == "host":
        from .parallel import host_ps
        if world != cfg.num_ps + cfg.num_workers:
            raise SystemExit("host transport needs world == P + W (got %d, P=%d, W=%d)"
                             % (world, cfg.num_ps, cfg.num_workers))
        host_ps.init_host_group(rank, world, cfg.timeout_s)
        import torch.distributed as dist
        try:
            if role == "ps":
                node = host_ps.HostPS(cfg, rank)
                node.run()
                out = {"steps_served": node.steps_served}
            else:
                node = host_ps.HostWorker(cfg, rank)
                out = {"final_accuracy": node.run()}
        except BaseException:
            # a failed rank must die at once so that the launcher can stop the job: no farewell barrier,
            # the peers would sit in it until the collective timeout
            if dist.is_initialized():
                try:
                    dist.destroy_process_group()
                except Exception:                               # noqa: BLE001
                    pass
            raise
        if dist.is_initialized():
            dist.barrier()
            dist.destroy_process_group()
        return out
    from .parallel.trainer import SPMDTrainer
    kind = os.environ.get("DDL_LAYOUT", "auto")
    tr = SPMDTrainer(cfg, rank, world, local, kind)
    try:
        return tr.run()
    finally:
        tr.close()


def ps_main(variant: str, argv=None):
    return run_role("ps", variant, argv)


def worker_main(variant: str, argv=None):
    return run_role("worker", variant, argv)


def single_main(variant: str = "mnist_sync", argv=None):
    """``Single(1, 100).train()`` (mnist_sync/single.py:24-26): one process, local optimizer."""
    os.environ["DDL_ROLE"] = "single"
    cfg = config_from_args(sys.argv[1:] if argv is None else argv, "single", variant)
    cfg = cfg.replace(num_ps=1, num_workers=1, shard="none", mode="sync", transport="auto")
    from .parallel.trainer import SPMDTrainer
    tr = SPMDTrainer(cfg, 0, 1, 0, "auto")
    try:
        return tr.run()
    finally:
        tr.close()

"""Rank layout: which process plays which role on which GPU.

Reference layout (mnist_sync_sharding/run.sh:3, mnist_sync_sharding/parameter_server.py:95-96,
mnist_sync_sharding/worker.py:65-66): ``mpiexec`` MPMD puts PS ranks at ``[0, P)`` and worker
ranks at ``[P, P+W)``; every rank is an OS process talking through host memory.

On B200 a "PS shard" is a GPU-resident range of the arena, not a busy process, so two physical
layouts exist for the device transports (``fused`` / ``nccl``):

* ``dedicated``  (P + W <= #GPUs): process r < P owns shard r on GPU r, process P+i is worker i on
  GPU P+i - e.g. BASELINE config "1 PS shard + 7 workers".
* ``colocated``  (otherwise, or on request): max(P, W) processes; process i owns shard i (if i < P)
  AND is worker i (if i < W) on GPU i - e.g. "8 GPU PS shards" with 8 workers.  Two roles never
  become two processes on one GPU: kernels that wait on each other must not share a device.

Log lines keep the reference's *logical* rank (worker i prints ``Worker{P+i}``).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional


@dataclass(frozen=True)
class RankLayout:
    world: int                 # number of processes
    rank: int                  # this process
    num_ps: int
    num_workers: int
    kind: str                  # dedicated | colocated | single

    @property
    def my_shards(self) -> List[int]:
        if self.kind == "single":
            return list(range(self.num_ps))
        return [self.rank] if self.rank < self.num_ps else []

    @property
    def my_worker(self) -> Optional[int]:
        if self.kind == "single":
            return 0
        if self.kind == "dedicated":
            return self.rank - self.num_ps if self.rank >= self.num_ps else None
        return self.rank if self.rank < self.num_workers else None

    @property
    def is_worker(self) -> bool:
        return self.my_worker is not None

    def owner_rank(self, shard: int) -> int:
        return 0 if self.kind == "single" else shard

    def worker_rank(self, w: int) -> int:
        if self.kind == "dedicated":
            return self.num_ps + w
        return 0 if self.kind == "single" else w

    @property
    def worker_ranks(self) -> List[int]:
        return [self.worker_rank(w) for w in range(self.num_workers)]

    @property
    def logical_worker_rank(self) -> Optional[int]:
        """Rank printed in the reference's ``Worker{}`` log lines."""
        w = self.my_worker
        return None if w is None else self.num_ps + w


def make_layout(num_ps: int, num_workers: int, rank: int, world: int, kind: str = "auto") -> RankLayout:
    if world == 1 and num_workers == 1:
        return RankLayout(1, 0, num_ps, 1, "single")
    if kind == "auto":
        kind = "dedicated" if world == num_ps + num_workers else "colocated"
    if kind == "dedicated" and world != num_ps + num_workers:
        raise ValueError("dedicated layout needs world == P + W (%d != %d + %d)" % (world, num_ps, num_workers))
    if kind == "colocated" and world != max(num_ps, num_workers):
        raise ValueError("colocated layout needs world == max(P, W)")
    return RankLayout(world, rank, num_ps, num_workers, kind)


def processes_for(num_ps: int, num_workers: int, transport: str, num_gpus: int) -> tuple:
    """(#processes, layout kind) the launcher should start for ``run.sh P W``."""
    if transport == "host":
        return num_ps + num_workers, "mpmd"
    if num_ps + num_workers <= max(1, num_gpus):
        return num_ps + num_workers, "dedicated"
    n = max(num_ps, num_workers)
    if num_gpus and n > num_gpus:
        raise ValueError("%d shards / %d workers do not fit %d GPUs" % (num_ps, num_workers, num_gpus))
    return n, "colocated"

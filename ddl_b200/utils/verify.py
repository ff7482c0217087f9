"""Debug-mode integrity checks (``--debug-checksum``; SURVEY 5.2: "a debug mode should checksum pushes").

The reference has two real races on its message path - a receive buffer reused while the previous gradient is still
being summed (Q2, mnist_sync_sharding/parameter_server.py:47,78-80) and pushes of different workers interleaving
variable by variable (Q4, mnist_async/parameter_server.py:58) - and nothing that would notice either.  Here every
message group of the ``host`` transport can carry an exact checksum of what was sent, and the SPMD trainer can compare
the replicas after every synchronous step.

The checksum is over BIT PATTERNS (the fp32 words viewed as int32), so it is exact and independent of summation
order: component 0 is the plain sum, component 1 a position-weighted sum (detects two elements swapped or a block
shifted).  int64 arithmetic wraps, which is what a checksum wants.
"""
from __future__ import annotations

from typing import Dict, List, Sequence

import torch
import torch.distributed as dist


class ChecksumError(RuntimeError):
    pass


def bits_checksum(t: torch.Tensor) -> torch.Tensor:
    """[2] int64 on ``t``'s device: (sum of words, position-weighted sum of words)."""
    bits = t.detach().contiguous().view(-1).view(torch.int32).to(torch.int64)
    w = (torch.arange(bits.numel(), device=bits.device, dtype=torch.int64) % 65521) + 1
    return torch.stack([bits.sum(), (bits * w).sum()])


def checksum_table(tensors: Sequence[torch.Tensor]) -> torch.Tensor:
    """[len(tensors), 2] int64 on the CPU (message payload of the host transport)."""
    if not tensors:
        return torch.zeros(0, 2, dtype=torch.int64)
    return torch.stack([bits_checksum(t).cpu() for t in tensors])


def compare_tables(got: torch.Tensor, want: torch.Tensor, names: Sequence[str], what: str) -> None:
    if got.shape != want.shape:
        raise ChecksumError("%s: checksum table has %s entries, expected %s" % (what, tuple(got.shape), tuple(want.shape)))
    bad = (got != want).any(dim=1).nonzero().flatten().tolist()
    if bad:
        raise ChecksumError("%s: checksum mismatch for %s (sent %s, received data sums to %s)" % (
            what, ", ".join(names[i] for i in bad), want[bad[0]].tolist(), got[bad[0]].tolist()))


def replica_check(tr, step: int) -> Dict[str, object]:
    """Sync SPMD training: after a step every worker must hold bitwise-identical variables, and an owner's range must be
    what the workers hold (the pull installed exactly what the owner computed).  One all-gather of a [num_vars, 2]
    table; raises ChecksumError naming the variable and the ranks that differ."""
    lay = tr.layout
    if hasattr(tr.transport, "ensure_local"):
        tr.transport.ensure_local()
    views = tr.params.views()
    mine = torch.stack([bits_checksum(v) for v in views])                       # [V, 2] on the device
    if lay.world > 1:
        allc: List[torch.Tensor] = [torch.zeros_like(mine) for _ in range(lay.world)]
        dist.all_gather(allc, mine)
    else:
        allc = [mine]
    wr = list(lay.worker_ranks)
    ref = allc[wr[0]]
    for r in wr[1:]:
        bad = (allc[r] != ref).any(dim=1).nonzero().flatten().tolist()
        if bad:
            raise ChecksumError("step %d: worker ranks %d and %d hold different parameters for %s" % (
                step, wr[0], r, ", ".join("v%d" % v for v in bad)))
    for v in range(tr.spec.num_vars):                                           # owners are the source of truth (Q5)
        o = lay.owner_rank(tr.plan.owner[v])
        if o not in wr and tr.plan.policy != "flat" and (allc[o][v] != ref[v]).any():
            raise ChecksumError("step %d: owner rank %d and the workers hold different values of v%d" % (step, o, v))
    return {"step": step, "ranks_compared": len(wr), "checksum": "%016x" % (int(ref[:, 1].sum().item()) & 0xFFFFFFFFFFFFFFFF)}


def async_accounting(tr) -> Dict[str, object]:
    """Async SPMD training, at the end of a run: every owner received ``steps x workers`` pushes and (non-SGD) applied
    them all - optimizer steps + pushes folded into a shared step = pushes.  Collective (MIN over ranks)."""
    lay = tr.layout
    st = tr.transport.stats() if hasattr(tr.transport, "stats") else {}
    expect = tr.global_step * lay.num_workers
    ok, why = 1, ""
    if lay.my_shards and "pushes_received" in st:
        if st["pushes_received"] != expect:
            ok, why = 0, "rank %d received %d pushes, expected %d" % (lay.rank, st["pushes_received"], expect)
        elif tr.cfg.optimizer != "sgd" and "folded_pushes" in st and st["optimizer_steps"] + st["folded_pushes"] != expect:
            ok, why = 0, "rank %d: %d optimizer steps + %d folded != %d pushes" % (
                lay.rank, st["optimizer_steps"], st["folded_pushes"], expect)
    t = torch.tensor([ok], device=tr.device)
    if lay.world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if int(t.item()) != 1:
        raise ChecksumError("async push accounting failed" + (": " + why if why else " on another rank"))
    return {"pushes_expected_per_owner": expect}

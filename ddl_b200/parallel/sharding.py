"""Placement policies: variable index -> (owner shard, local tag), and the flat arena layout.

Reference behaviour restated (never imported):

* ``block``  - contiguous split, ``avg = V // P``, last shard takes the remainder
  (mnist_sync_sharding/parameter_server.py:30-32, mnist_sync_sharding/worker.py:19-20,33-36).
* ``greedy_interleave`` - the reference's "greedy": sort by element count, interleave
  smallest/largest, then block-split that order (mnist_sync_sharding_greedy/worker.py:15-37).
  For the 14-variable CNN it yields order [13,8,1,6,3,10,5,4,7,2,11,12,0,9].
* ``none``   - one PS owns everything (mnist_sync/parameter_server.py).

New policies: ``round_robin`` (i mod P), ``greedy`` (true LPT bin packing by
bytes - what BASELINE.json means by "size-balanced"), and ``flat`` (sub-tensor split of the
16B-aligned arena, the only policy that can balance a model whose v8 holds 39% of the bytes).

The arena is laid out shard by shard, so every owner's state is ONE contiguous range and the
push / update / pull kernels each run once over a range instead of once per variable.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Sequence, Tuple

ALIGN_ELEMS = 64          # 256-byte alignment of every variable inside the arena (TMA-friendly)


def reference_greedy_order(sizes: Sequence[int]) -> List[int]:
    """Order produced by mnist_sync_sharding_greedy/worker.py:15-30 (stable sort, interleave)."""
    tmp = sorted(range(len(sizes)), key=lambda i: sizes[i])
    order: List[int] = []
    i, j = 0, len(tmp) - 1
    while i < j:
        order.append(tmp[i])
        order.append(tmp[j])
        i += 1
        j -= 1
    if len(tmp) % 2:
        order.append(tmp[i])
    return order


def block_owner(pos: int, num_vars: int, num_ps: int) -> Tuple[int, int]:
    """(owner, local tag) of the variable at position ``pos`` of the (possibly reordered)
    variable list - the arithmetic of mnist_sync_sharding/worker.py:33-36."""
    avg = num_vars // num_ps
    local_last = avg + num_vars % num_ps
    if avg == 0:
        return num_ps - 1, pos
    ind = pos // avg
    if pos >= num_vars - local_last:
        ind = num_ps - 1
    return ind, pos - ind * avg


@dataclass(frozen=True)
class ShardPlan:
    policy: str
    num_shards: int
    order: Tuple[int, ...]                 # variable indices in arena order
    owner: Tuple[int, ...]                 # owner[var_index] (-1 in flat mode: split sub-tensor)
    tag: Tuple[int, ...]                   # local index of var inside its shard (reference tag)
    var_offset: Tuple[int, ...]            # element offset of var_index in the arena
    var_numel: Tuple[int, ...]
    shard_range: Tuple[Tuple[int, int], ...]   # [begin, end) element range of each shard in the arena
    total: int                             # arena length in elements (padded)

    def shard_vars(self, r: int) -> List[int]:
        return [v for v in self.order if self.owner[v] == r]

    def shard_elems(self, r: int) -> int:
        b, e = self.shard_range[r]
        return e - b

    def payload_elems(self, r: int) -> int:
        """Un-padded elements owned by shard r (whole-variable policies)."""
        return sum(self.var_numel[v] for v in self.shard_vars(r))

    def imbalance(self) -> float:
        sizes = [self.shard_elems(r) for r in range(self.num_shards)]
        mean = sum(sizes) / len(sizes)
        return max(sizes) / mean if mean else 1.0


def _align(n: int, a: int = ALIGN_ELEMS) -> int:
    return (n + a - 1) // a * a


def assign(policy: str, sizes: Sequence[int], num_shards: int) -> Tuple[List[int], List[int]]:
    """Returns (order, owner_by_var).  ``order`` lists variable indices shard by shard."""
    V, P = len(sizes), num_shards
    if P < 1:
        raise ValueError("num_shards must be >= 1")
    if policy == "none":
        if P != 1:
            raise ValueError("policy 'none' requires exactly one shard")
        return list(range(V)), [0] * V
    if policy in ("block", "greedy_interleave"):
        seq = list(range(V)) if policy == "block" else reference_greedy_order(sizes)
        owner = [0] * V
        for pos, v in enumerate(seq):
            owner[v] = block_owner(pos, V, P)[0]
        return seq, owner
    if policy == "round_robin":
        owner = [i % P for i in range(V)]
        order = [v for r in range(P) for v in range(V) if owner[v] == r]
        return order, owner
    if policy == "greedy":
        # LPT: largest variable first into the currently lightest shard; ties -> lowest shard id.
        load = [0] * P
        owner = [0] * V
        for v in sorted(range(V), key=lambda i: (-sizes[i], i)):
            r = min(range(P), key=lambda q: (load[q], q))
            owner[v] = r
            load[r] += sizes[v]
        order = [v for r in range(P) for v in range(V) if owner[v] == r]
        return order, owner
    if policy == "flat":
        return list(range(V)), [-1] * V
    raise ValueError("unknown shard policy %r" % policy)


def make_plan(policy: str, sizes: Sequence[int], num_shards: int) -> ShardPlan:
    sizes = [int(s) for s in sizes]
    V, P = len(sizes), num_shards
    order, owner = assign(policy, sizes, P)
    var_offset = [0] * V
    tag = [0] * V
    cursor = 0
    shard_range: List[Tuple[int, int]] = []
    if policy == "flat":
        for v in order:
            var_offset[v] = cursor
            cursor += _align(sizes[v])
        total = _align(cursor, ALIGN_ELEMS * P)
        per = total // P                         # multiple of ALIGN_ELEMS -> 16B aligned cuts
        shard_range = [(r * per, (r + 1) * per) for r in range(P)]
        tag = list(range(V))
    else:
        local: Dict[int, int] = {}
        for r in range(P):
            begin = cursor
            for v in order:
                if owner[v] != r:
                    continue
                var_offset[v] = cursor
                tag[v] = local.get(r, 0)
                local[r] = tag[v] + 1
                cursor += _align(sizes[v])
            shard_range.append((begin, cursor))
        total = cursor
        # whole-variable policies keep the reference's per-shard order: block/interleave
        # orders are already contiguous per shard, so `order` == arena order.
        order = [v for r in range(P) for v in order if owner[v] == r]
    return ShardPlan(policy, P, tuple(order), tuple(owner), tuple(tag), tuple(var_offset),
                     tuple(sizes), tuple(shard_range), total)


def reference_shard_sizes(policy: str, sizes: Sequence[int], num_ps: int) -> List[int]:
    """Per-shard element counts as the reference's own arithmetic produces them."""
    order, owner = assign(policy, list(sizes), num_ps)
    out = [0] * num_ps
    for v in order:
        out[owner[v]] += sizes[v]
    return out

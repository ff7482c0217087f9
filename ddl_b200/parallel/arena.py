"""Flat parameter / gradient arenas.

The reference moves 14 separate numpy buffers per direction per step (one blocking
``comm.Send``/``Recv`` each: mnist_sync_sharding/worker.py:36,94).  Here every variable is a
view into ONE 256-byte-aligned fp32 arena laid out shard by shard (see sharding.ShardPlan), so a
push, an optimizer update or a pull is one kernel over one contiguous range.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

from ..models.spec import ModelSpec
from .sharding import ShardPlan, make_plan


class Arena:
    """A flat fp32 buffer plus per-variable views shaped like the reference variables."""

    def __init__(self, spec: ModelSpec, plan: ShardPlan, device="cpu", buffer: Optional[torch.Tensor] = None):
        self.spec, self.plan = spec, plan
        if buffer is None:
            buffer = torch.zeros(plan.total, dtype=torch.float32, device=device)
        assert buffer.numel() >= plan.total and buffer.dtype == torch.float32
        self.flat = buffer
        self.device = buffer.device

    # -- views -------------------------------------------------------------------------
    def view(self, var: int, flat: Optional[torch.Tensor] = None) -> torch.Tensor:
        v = self.spec.vars[var]
        base = self.flat if flat is None else flat
        off = self.plan.var_offset[var]
        return base.narrow(0, off, v.numel).view(v.shape)

    def views(self, flat: Optional[torch.Tensor] = None) -> List[torch.Tensor]:
        return [self.view(i, flat) for i in range(self.spec.num_vars)]

    def shard(self, r: int, flat: Optional[torch.Tensor] = None) -> torch.Tensor:
        b, e = self.plan.shard_range[r]
        return (self.flat if flat is None else flat).narrow(0, b, e - b)

    # -- (de)serialisation of the reference variable list --------------------------------
    def load_vars(self, tensors: Sequence[torch.Tensor]) -> None:
        with torch.no_grad():
            for i, t in enumerate(tensors):
                self.view(i).copy_(t.to(self.device))

    def export_vars(self) -> List[torch.Tensor]:
        return [self.view(i).detach().clone().cpu() for i in range(self.spec.num_vars)]

    def like(self, device=None) -> "Arena":
        return Arena(self.spec, self.plan, device or self.device)


def make_arena(spec: ModelSpec, policy: str, num_shards: int, device="cpu") -> Arena:
    plan = make_plan(policy, [v.numel for v in spec.vars], num_shards)
    return Arena(spec, plan, device)

"""Metrics, phase timers and log lines.

The reference prints three kinds of lines and nothing else:
``Worker{} epoch: {} batch: {} accuracy: {}`` (mnist_sync/worker.py:72),
``Worker{} final accuracy: {}`` (worker.py:75) and ``Time: {}`` (worker.py:76, measured with
``time.clock()`` = process CPU seconds, quirk Q7); single.py drops the ``Worker{}`` prefix
(mnist_sync/single.py:18,21).  Those exact formats are kept.  On top: a JSONL sink and
per-phase timers (CUDA events on GPU, perf_counter on CPU) - the reference has no tracing.
"""
from __future__ import annotations

import contextlib
import json
import time
from collections import defaultdict
from typing import Dict, List, Optional

import torch


def log_eval(rank: Optional[int], epoch: int, batch: int, acc: float) -> str:
    s = ("Worker{} epoch: {} batch: {} accuracy: {}".format(rank, epoch, batch, acc) if rank is not None
         else "epoch: {} batch: {} accuracy: {}".format(epoch, batch, acc))
    print(s, flush=True)
    return s


def log_final(rank: Optional[int], acc: float) -> str:
    s = ("Worker{} final accuracy: {}".format(rank, acc) if rank is not None
         else "final accuracy: {}".format(acc))
    print(s, flush=True)
    return s


def log_time(seconds: float) -> str:
    s = "Time: {}".format(seconds)
    print(s, flush=True)
    return s


class PhaseTimers:
    """Named phases (fwd_bwd, push, update, pull, eval).  GPU phases are bracketed by CUDA events
    on the current stream and resolved lazily, so timing adds no host sync to the step."""

    def __init__(self, device: torch.device, enabled: bool = True):
        self.cuda = device.type == "cuda"
        self.enabled = enabled
        self.pending: List = []
        self.total_ms: Dict[str, float] = defaultdict(float)
        self.count: Dict[str, int] = defaultdict(int)

    @contextlib.contextmanager
    def phase(self, name: str):
        if not self.enabled:
            yield
            return
        if self.cuda:
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.nvtx.range_push(name)
            a.record()
            try:
                yield
            finally:
                b.record()
                torch.cuda.nvtx.range_pop()
                self.pending.append((name, a, b))
        else:
            t0 = time.perf_counter()
            try:
                yield
            finally:
                self.total_ms[name] += (time.perf_counter() - t0) * 1e3
                self.count[name] += 1

    def resolve(self) -> None:
        if self.pending:
            torch.cuda.synchronize()
            for name, a, b in self.pending:
                self.total_ms[name] += a.elapsed_time(b)
                self.count[name] += 1
            self.pending.clear()

    def summary(self) -> Dict[str, float]:
        self.resolve()
        return {k: self.total_ms[k] / max(1, self.count[k]) for k in self.total_ms}


class MetricsSink:
    """Append-only JSONL (samples/s, push_ms, pull_ms, staleness histogram, ...)."""

    def __init__(self, path: str = ""):
        self.path = path
        self.f = open(path, "a") if path else None

    def write(self, **rec) -> None:
        if self.f:
            rec.setdefault("ts", time.time())
            self.f.write(json.dumps(rec) + "\n")
            self.f.flush()

    def close(self) -> None:
        if self.f:
            self.f.close()
            self.f = None

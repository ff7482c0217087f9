"""Python face of the native step executor (csrc/engine.cu).

``NativeBackend`` gives ``models.cnn.Model`` the reference's three graph operations on sm_100a
kernels: ``compute_grads`` = ``sess.run(self.grads)`` (mnist_sync/worker.py:30), ``evaluate`` =
``sess.run(self.accuracy)`` (worker.py:72) and ``logits``.  All calls only enqueue work on the
current stream, so a caller may capture ``step_advance + compute_grads + transport.step`` in one
CUDA graph.
"""
from __future__ import annotations

import os

import ctypes as C
from typing import Optional, Tuple

import torch

from ..config import Config
from ..models.spec import ModelSpec
from ..parallel.arena import Arena
from . import native


class NativeBackend:
    name = "native"

    def __init__(self, spec: ModelSpec, params: Arena, grads: Arena, cfg: Config, max_eval_batch: int = 2000,
                 step_state: Optional[torch.Tensor] = None, worker: int = 0, num_workers: int = 1,
                 fc_w_src=None):
        if spec.kind != "cnn":
            raise native.NativeError("the native engine implements the CNN; use --backend torch for the MLP")
        self.L = native.lib()
        self.spec, self.params, self.grads, self.cfg = spec, params, grads, cfg
        assert params.plan.var_offset == grads.plan.var_offset, "param / grad arenas must share a layout"
        self.seed = 0 if cfg.seed is None else cfg.seed
        self.st = step_state if step_state is not None else native.new_step_state(self.seed, worker, num_workers)
        self.owns_state = step_state is None
        self.max_eval_batch = int(max_eval_batch)
        ecfg = native.EngineCfg()
        ecfg.kind = 0
        ecfg.width_mult = spec.width_mult
        ecfg.batch = cfg.batch_size
        ecfg.max_eval_batch = self.max_eval_batch
        ecfg.keep_prob = cfg.keep_prob
        ecfg.params = params.flat.data_ptr()
        ecfg.grads = grads.flat.data_ptr()
        ecfg.arena_elems = params.plan.total
        for i, off in enumerate(params.plan.var_offset):
            ecfg.var_offset[i] = off
        ecfg.st = self.st.data_ptr()
        # K4: peer pointers of the fc1 / fc2 weight matrices in their owners' arenas (None = local)
        for i, v in enumerate((8, 10)):
            ecfg.fc_w_src[i] = (fc_w_src or {}).get(v)
        self.h = self.L.ddl_engine_create(C.byref(ecfg))
        if not self.h:
            raise native.NativeError("ddl_engine_create: " + native.last_error())
        self._correct = torch.zeros(1, dtype=torch.int32, device=params.device)
        self.launches_per_step = 0

    def __del__(self):
        h, self.h = getattr(self, "h", None), None
        if h:
            try:
                self.L.ddl_engine_destroy(h)
            except Exception:
                pass

    # ------------------------------------------------------------------------------
    def enqueue_train_step(self, x: torch.Tensor, y: torch.Tensor) -> None:
        """Enqueue fwd+bwd on the current stream (graph-capturable).  The step state must have been
        advanced (``advance_state``) by the caller."""
        native.check(self.L.ddl_engine_train_step(self.h, native.ptr(x), native.ptr(y), native.stream_ptr()),
                     "ddl_engine_train_step")

    def advance_state(self, advance_opt: bool = True) -> None:
        c = self.cfg
        native.check(self.L.ddl_step_advance(self.st.data_ptr(), c.lr, c.beta1, c.beta2, int(advance_opt),
                                             native.stream_ptr()), "ddl_step_advance")
        native.count()

    def compute_grads(self, x: torch.Tensor, y: torch.Tensor, step: int, keep_prob: Optional[float] = None):
        """Eager path used by the generic trainers: returns the mean loss as a 0-d device tensor."""
        if keep_prob is not None and abs(keep_prob - self.cfg.keep_prob) > 1e-9:
            raise native.NativeError("the native engine bakes keep_prob into its plans (%g)" % self.cfg.keep_prob)
        assert x.shape[0] == self.cfg.batch_size, "native engine is planned for batch %d" % self.cfg.batch_size
        if self.owns_state:
            # standalone use: the dropout stream follows the caller's step id (oracle parity)
            # stream = base + wstep * stride with base 0, stride 2, wstep = step  ->  2 * step
            self.st[5:9] = torch.tensor([0, step, 0, 2], dtype=torch.int32)
            self.advance_state(advance_opt=False)
        x = x.contiguous().float()
        y = y.contiguous().to(torch.int32)
        self.enqueue_train_step(x, y)
        return self.st.view(torch.float32)[2] / x.shape[0]

    @torch.no_grad()
    def evaluate(self, x: torch.Tensor, y: torch.Tensor) -> Tuple[int, int]:
        x = x.contiguous().float()
        y = y.contiguous().to(torch.int32)
        self._correct.zero_()
        n = x.shape[0]
        for b in range(0, n, self.max_eval_batch):
            e = min(n, b + self.max_eval_batch)
            native.check(self.L.ddl_engine_eval(self.h, x[b:e].data_ptr(), y[b:e].data_ptr(), e - b,
                                                self._correct.data_ptr(), None, native.stream_ptr()),
                         "ddl_engine_eval")
        return int(self._correct.item()), n

    @torch.no_grad()
    def logits(self, x: torch.Tensor) -> torch.Tensor:
        x = x.contiguous().float()
        n = x.shape[0]
        out = torch.empty(n, 10, dtype=torch.float32, device=x.device)
        for b in range(0, n, self.max_eval_batch):
            e = min(n, b + self.max_eval_batch)
            native.check(self.L.ddl_engine_eval(self.h, x[b:e].data_ptr(), None, e - b, None,
                                                out[b:e].data_ptr(), native.stream_ptr()), "ddl_engine_eval")
        return out

    def buffer(self, which: int, shape) -> torch.Tensor:
        """Copy of an engine workspace buffer (tests / debugging)."""
        p = self.L.ddl_engine_buffer(self.h, which)
        out = torch.empty(shape, dtype=torch.float32, device=self.params.device)
        native.check(self.L.ddl_memcpy_async(out.data_ptr(), p, out.numel() * 4, native.stream_ptr()), "memcpy")
        return out

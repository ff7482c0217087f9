"""Compute front-end of the 14-variable CNN (and its widened / MLP variants).

Replaces ``class Model`` of mnist_sync/model/model.py:5-112.  The reference builds a TF1
graph and exposes ``grads`` / ``accuracy`` / ``train_step`` ops run through ``sess.run``; here
a ``Model`` owns views into a flat parameter arena and offers the same three operations:

* ``compute_grads(x, y, step)``  - ``sess.run(self.grads, keep_prob=.5)`` (worker.py:30):
  loss + all gradients written into the flat gradient arena;
* ``evaluate(x, y)``             - ``sess.run(self.accuracy, keep_prob=1)`` (worker.py:72);
* parameters are installed by writing the arena (``tf.assign`` at worker.py:56,69).

Two backends: ``native`` (the sm_100a engine in csrc/, used whenever the tensors are on a
GPU) and ``torch`` (plain PyTorch ops; the CPU path and the cuDNN/cuBLAS baseline rung).
On a CUDA device the native backend is mandatory unless ``backend='torch'`` is requested
explicitly - a missing extension raises instead of silently falling back.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from ..config import Config
from ..parallel.arena import Arena
from . import oracle
from .spec import ModelSpec


class TorchBackend:
    """PyTorch-op implementation (autograd).  Baseline rung + CPU path."""

    name = "torch"

    def __init__(self, spec: ModelSpec, params: Arena, grads: Arena, cfg: Config):
        self.spec, self.params, self.grads, self.cfg = spec, params, grads, cfg
        self.seed = 0 if cfg.seed is None else cfg.seed

    def compute_grads(self, x: torch.Tensor, y: torch.Tensor, step: int,
                      keep_prob: Optional[float] = None) -> torch.Tensor:
        kp = self.cfg.keep_prob if keep_prob is None else keep_prob
        flat = self.params.flat.detach().requires_grad_(True)
        logits = oracle.forward(self.params.views(flat), x, self.spec, kp, None, self.seed, step)
        loss = oracle.loss_fn(logits, y)
        (g,) = torch.autograd.grad(loss, flat)
        self.grads.flat.copy_(g)
        return loss.detach()

    @torch.no_grad()
    def evaluate(self, x: torch.Tensor, y: torch.Tensor, chunk: int = 2000) -> Tuple[int, int]:
        correct = 0
        for b in range(0, x.shape[0], chunk):
            logits = oracle.forward(self.params.views(), x[b:b + chunk], self.spec, 1.0)
            correct += int((logits.argmax(1) == y[b:b + chunk].long()).sum())
        return correct, int(x.shape[0])

    @torch.no_grad()
    def logits(self, x: torch.Tensor) -> torch.Tensor:
        return oracle.forward(self.params.views(), x, self.spec, 1.0)


def pick_backend(cfg: Config, device: torch.device) -> str:
    if cfg.backend != "auto":
        return cfg.backend
    if cfg.model != "cnn":
        return "torch"                  # the native engine implements the reference CNN (any width)
    return "native" if device.type == "cuda" else "torch"


class Model:
    """Backend-dispatching model.  ``params`` / ``grads`` are flat arenas with the same plan."""

    def __init__(self, spec: ModelSpec, params: Arena, grads: Arena, cfg: Config,
                 max_eval_batch: int = 2000, step_state: Optional[torch.Tensor] = None,
                 worker: int = 0, num_workers: int = 1, fc_w_src=None):
        self.spec, self.params, self.grads, self.cfg = spec, params, grads, cfg
        which = pick_backend(cfg, params.device)
        if which == "native":
            if params.device.type != "cuda":
                raise RuntimeError("the native sm_100a backend needs a CUDA device; use --backend torch on CPU")
            from ..ops.engine import NativeBackend   # raises loudly if the .so is missing
            self.impl = NativeBackend(spec, params, grads, cfg, max_eval_batch, step_state, worker, num_workers,
                                      fc_w_src)
        else:
            self.impl = TorchBackend(spec, params, grads, cfg)
        self.backend = self.impl.name

    def compute_grads(self, x, y, step: int, keep_prob: Optional[float] = None) -> torch.Tensor:
        return self.impl.compute_grads(x, y, step, keep_prob)

    def evaluate(self, x, y) -> Tuple[int, int]:
        return self.impl.evaluate(x, y)

    def accuracy(self, x, y) -> float:
        c, n = self.evaluate(x, y)
        return c / max(1, n)

    def logits(self, x) -> torch.Tensor:
        return self.impl.logits(x)

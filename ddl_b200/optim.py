"""Shard-local optimizers over flat fp32 ranges.

The reference applies ``tf.compat.v1.train.AdamOptimizer(1e-4).apply_gradients`` on the PS
(mnist_sync/parameter_server.py:21-27, mnist_sync_sharding/parameter_server.py:63-69) and
``optimizer.minimize`` locally in single.py (mnist_sync/model/model.py:106).  TF1's Adam is

    lr_t = lr * sqrt(1 - b2^t) / (1 - b1^t);  m = b1 m + (1-b1) g;  v = b2 v + (1-b2) g^2
    p   -= lr_t * m / (sqrt(v) + eps)            # eps OUTSIDE the bias correction

which is not ``torch.optim.Adam``; both the CUDA kernel (csrc/elementwise.cu) and the torch
fallback below use TF1's form.  ``gscale`` folds the sum/mean choice of the gradient
reduction (Q3) into the update, as the fused push kernel does.
"""
from __future__ import annotations

import math
from typing import Dict

import torch


def adam_lr_t(lr: float, b1: float, b2: float, t: int) -> float:
    return lr * math.sqrt(1.0 - b2 ** t) / (1.0 - b1 ** t)


class FlatOptimizer:
    KINDS = {"sgd": 0, "momentum": 1, "adam": 2}

    def __init__(self, like: torch.Tensor, kind: str = "adam", lr: float = 1e-4, beta1: float = 0.9,
                 beta2: float = 0.999, eps: float = 1e-8, momentum: float = 0.9, use_native: bool = True):
        if kind not in self.KINDS:
            raise ValueError("unknown optimizer %r" % kind)
        self.kind, self.lr, self.b1, self.b2, self.eps, self.mu = kind, lr, beta1, beta2, eps, momentum
        self.m = torch.zeros_like(like) if kind in ("adam", "momentum") else None
        self.v = torch.zeros_like(like) if kind == "adam" else None
        self.t = 0
        self.native = None
        if like.is_cuda and use_native:
            from .ops import native
            self.native = native.lib()

    # ---------------------------------------------------------------------------------
    def hyper(self, t: int):
        """(kind id, step size, a, b, eps) exactly as the kernels take them."""
        if self.kind == "adam":
            return 2, adam_lr_t(self.lr, self.b1, self.b2, t), self.b1, self.b2, self.eps
        if self.kind == "momentum":
            return 1, self.lr, self.mu, 0.0, 0.0
        return 0, self.lr, 0.0, 0.0, 0.0

    @torch.no_grad()
    def step(self, p: torch.Tensor, g: torch.Tensor, gscale: float = 1.0) -> None:
        self.t += 1
        if self.native is not None and p.is_cuda:
            from .ops import native
            native.optimizer_step(self, p, g, gscale)
            return
        if gscale != 1.0:
            g = g * gscale
        if self.kind == "adam":
            lr_t = adam_lr_t(self.lr, self.b1, self.b2, self.t)
            self.m.mul_(self.b1).add_(g, alpha=1 - self.b1)
            self.v.mul_(self.b2).addcmul_(g, g, value=1 - self.b2)
            p.addcdiv_(self.m, self.v.sqrt().add_(self.eps), value=-lr_t)
        elif self.kind == "momentum":
            self.m.mul_(self.mu).add_(g)
            p.sub_(self.m, alpha=self.lr)
        else:
            p.sub_(g, alpha=self.lr)

    # ---------------------------------------------------------------------------------
    def state_dict(self) -> Dict[str, object]:
        return {"kind": self.kind, "t": self.t,
                "m": None if self.m is None else self.m.detach().cpu().clone(),
                "v": None if self.v is None else self.v.detach().cpu().clone()}

    def load_state_dict(self, sd: Dict[str, object]) -> None:
        if sd["kind"] != self.kind:
            raise ValueError("optimizer kind mismatch: %s vs %s" % (sd["kind"], self.kind))
        self.t = int(sd["t"])
        if self.m is not None and sd["m"] is not None:
            self.m.copy_(sd["m"])
        if self.v is not None and sd["v"] is not None:
            self.v.copy_(sd["v"])

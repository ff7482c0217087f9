"""CPU fp32 oracle: a plain restatement of the reference math (never imported from it).

* forward / loss / accuracy of mnist_sync/model/model.py:17-105 - NHWC input, HWIO 5x5 SAME
  convs, 2x2 stride-2 SAME max-pool (7 -> 4 edge at model.py:53), (h, w, c) flatten order
  (model.py:69), fc2 with no activation (model.py:79), dropout after fc1 and fc2.
* TF1 Adam (``lr_t = lr*sqrt(1-b2^t)/(1-b1^t)``, epsilon outside the bias correction) -
  what ``tf.compat.v1.train.AdamOptimizer(1e-4)`` at model.py:93 applies.
* A sequential parameter-server simulator used as the oracle for every transport:
  sync = sum (or mean) of worker gradients then one optimizer step
  (mnist_sync/parameter_server.py:34-38, without the Q1/Q2 double-count bugs), async = one
  optimizer step per push in arrival order (mnist_async/parameter_server.py:55-69).

The counter-based dropout hash below is bit-identical to ``ddl_dropout_keep`` in
csrc/common.cuh so kernels and oracle draw the same masks.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence

import torch
import torch.nn.functional as F

from .spec import ModelSpec, REFERENCE_SPEC


# ----------------------------------------------------------------------------- init
def glorot_uniform_(t: torch.Tensor, fan_in: int, fan_out: int, gen: Optional[torch.Generator]) -> torch.Tensor:
    limit = math.sqrt(6.0 / (fan_in + fan_out))
    with torch.no_grad():
        t.copy_((torch.rand(t.shape, generator=gen) * 2 - 1) * limit)
    return t


def init_params(spec: ModelSpec = REFERENCE_SPEC, seed: Optional[int] = 0) -> List[torch.Tensor]:
    """TF1 ``get_variable`` with no initializer = glorot_uniform for weights AND biases."""
    gen = torch.Generator().manual_seed(seed) if seed is not None else None
    out = []
    for v in spec.vars:
        fi, fo = v.fans()
        out.append(glorot_uniform_(torch.empty(v.shape, dtype=torch.float32), fi, fo, gen))
    return out


# ----------------------------------------------------------------------------- dropout hash
_M32 = 0xFFFFFFFF


def dropout_keep_mask(seed: int, stream: int, n: int, keep_prob: float, device="cpu") -> torch.Tensor:
    """keep[i] = (mix32(i*0x9E3779B9 + seed*0x85EBCA6B + stream*0xC2B2AE35) >> 8) < keep_prob*2^24."""
    i = torch.arange(n, dtype=torch.int64, device=device)
    h = (i * 0x9E3779B9 + (int(seed) & _M32) * 0x85EBCA6B + (int(stream) & _M32) * 0xC2B2AE35) & _M32
    h = h ^ (h >> 16)
    h = (h * 0x85EBCA6B) & _M32
    h = h ^ (h >> 13)
    h = (h * 0xC2B2AE35) & _M32
    h = h ^ (h >> 16)
    thresh = int(round(float(keep_prob) * (1 << 24)))
    return (h >> 8) < thresh


# ----------------------------------------------------------------------------- forward
def same_maxpool2x2(x_nchw: torch.Tensor) -> torch.Tensor:
    """TF SAME 2x2/2 max-pool: pads bottom/right only; ceil_mode reproduces it (7 -> 4)."""
    return F.max_pool2d(x_nchw, 2, 2, ceil_mode=True)


def conv5x5_same(x_nchw: torch.Tensor, w_hwio: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    return F.conv2d(x_nchw, w_hwio.permute(3, 2, 0, 1), b, stride=1, padding=2)


def forward(params: Sequence[torch.Tensor], x: torch.Tensor, spec: ModelSpec = REFERENCE_SPEC,
            keep_prob: float = 1.0, masks: Optional[Sequence[torch.Tensor]] = None,
            seed: int = 0, step: int = 0, return_acts: bool = False):
    """logits [B,10].  ``masks`` = two boolean keep masks ([B,fc1], [B,fc2]); when omitted and
    keep_prob < 1 they are drawn from the shared counter hash (stream = 2*step + layer)."""
    B = x.shape[0]
    p = list(params)
    acts: Dict[str, torch.Tensor] = {}
    k = 0
    if spec.kind == "cnn":
        h = x.reshape(B, 28, 28, 1).permute(0, 3, 1, 2)            # model.py:19
        for i in range(4):
            h = F.relu(conv5x5_same(h, p[k], p[k + 1]))            # model.py:28-30
            acts["conv%d" % (i + 1)] = h
            h = same_maxpool2x2(h)                                 # model.py:31
            acts["pool%d" % (i + 1)] = h
            k += 2
        h = h.permute(0, 2, 3, 1).reshape(B, -1)                   # (h,w,c) flatten, model.py:69
    else:
        h = x.reshape(B, -1)
    n1, n2 = spec.fc_sizes[1], spec.fc_sizes[2]
    if masks is None and keep_prob < 1.0:
        masks = [dropout_keep_mask(seed, 2 * step, B * n1, keep_prob, x.device).view(B, n1),
                 dropout_keep_mask(seed, 2 * step + 1, B * n2, keep_prob, x.device).view(B, n2)]
    h = F.relu(h @ p[k] + p[k + 1])                                # fc1, model.py:70
    acts["fc1"] = h
    if masks is not None:
        h = h * masks[0].to(h.dtype) / keep_prob                   # model.py:74
    h = h @ p[k + 2] + p[k + 3]                                    # fc2: NO relu, model.py:79
    acts["fc2"] = h
    if masks is not None:
        h = h * masks[1].to(h.dtype) / keep_prob                   # model.py:82
    logits = h @ p[k + 4] + p[k + 5]                               # model.py:87
    return (logits, acts) if return_acts else logits


def loss_fn(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    """mean softmax cross entropy (model.py:91-92); labels are class ids."""
    return F.cross_entropy(logits, labels.long(), reduction="mean")


def accuracy(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    """mean(argmax(y) == argmax(y_)) (model.py:104-105)."""
    return (logits.argmax(1) == labels.long()).float().mean()


def loss_and_grads(params: Sequence[torch.Tensor], x, labels, spec: ModelSpec = REFERENCE_SPEC,
                   keep_prob: float = 1.0, masks=None, seed: int = 0, step: int = 0):
    """``optimizer.compute_gradients(cost, var_bucket)`` (model.py:101): grads in v0..vN order."""
    ps = [q.detach().clone().requires_grad_(True) for q in params]
    logits = forward(ps, x, spec, keep_prob, masks, seed, step)
    loss = loss_fn(logits, labels)
    grads = torch.autograd.grad(loss, ps)
    return loss.detach(), [g.detach() for g in grads], logits.detach()


# ----------------------------------------------------------------------------- optimizers
class OracleOptimizer:
    """Flat-tensor optimizer state; ``kind`` in adam (TF1 formula) | sgd | momentum."""

    def __init__(self, numel_or_like, kind="adam", lr=1e-4, beta1=0.9, beta2=0.999, eps=1e-8, momentum=0.9):
        like = numel_or_like if isinstance(numel_or_like, torch.Tensor) else torch.zeros(numel_or_like)
        self.kind, self.lr, self.b1, self.b2, self.eps, self.mu = kind, lr, beta1, beta2, eps, momentum
        self.m = torch.zeros_like(like)
        self.v = torch.zeros_like(like)
        self.t = 0

    def step(self, p: torch.Tensor, g: torch.Tensor) -> None:
        self.t += 1
        if self.kind == "adam":
            lr_t = self.lr * math.sqrt(1.0 - self.b2 ** self.t) / (1.0 - self.b1 ** self.t)
            self.m.mul_(self.b1).add_(g, alpha=1 - self.b1)
            self.v.mul_(self.b2).addcmul_(g, g, value=1 - self.b2)
            p.sub_(lr_t * self.m / (self.v.sqrt() + self.eps))
        elif self.kind == "sgd":
            p.sub_(g, alpha=self.lr)
        elif self.kind == "momentum":          # tf.train.MomentumOptimizer: acc = mu*acc + g; p -= lr*acc
            self.m.mul_(self.mu).add_(g)
            p.sub_(self.m, alpha=self.lr)
        else:
            raise ValueError(self.kind)


# ----------------------------------------------------------------------------- PS simulator
class SequentialPS:
    """Single-process, single-thread parameter server over flat vectors - the transport oracle."""

    def __init__(self, flat_params: torch.Tensor, kind="adam", lr=1e-4, reduce="sum", **kw):
        self.p = flat_params.clone()
        self.opt = OracleOptimizer(self.p, kind, lr, **kw)
        self.reduce = reduce
        self.version = 0

    def sync_step(self, worker_grads: Sequence[torch.Tensor]) -> torch.Tensor:
        g = torch.zeros_like(self.p)
        for wg in worker_grads:                 # true sum (the reference's Q1/Q2 aliasing is not reproduced)
            g += wg
        if self.reduce == "mean":
            g /= len(worker_grads)
        self.opt.step(self.p, g)
        self.version += 1
        return self.p

    def async_push(self, grad: torch.Tensor) -> torch.Tensor:
        self.opt.step(self.p, grad)
        self.version += 1
        return self.p

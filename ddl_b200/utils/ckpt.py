"""Checkpoint / resume.

The reference has no checkpoint code at all (no ``tf.train.Saver``, no file writes; SURVEY
5.4).  Its only "format" facts are the variable names ``mnist/v0 .. mnist/v13``
(mnist_sync/model/model.py:17,24-86) and their layouts (HWIO conv kernels, ``[in, out]`` FC
matrices).  A checkpoint here is therefore a mapping keyed by exactly those names in exactly
those layouts - interchangeable with a TF1 Saver dump of the reference graph - plus the
optimizer state (Adam m, v per variable, and the step count t) gathered from the shards.
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import torch

from ..parallel.arena import Arena

FORMAT_VERSION = 1


def _scope(name: str) -> str:
    return "mnist/" + name


def save_checkpoint(path: str, params: Arena, opt_m: Optional[torch.Tensor] = None,
                    opt_v: Optional[torch.Tensor] = None, t: int = 0, step: int = 0,
                    extra: Optional[Dict] = None) -> None:
    """``opt_m`` / ``opt_v`` are flat tensors laid out like ``params.flat`` (gathered from shards)."""
    spec = params.spec
    blob = {"format": FORMAT_VERSION, "step": int(step), "t": int(t),
            "model": {"kind": spec.kind, "width_mult": spec.width_mult},
            "vars": {_scope(v.name): params.view(v.index).detach().cpu().clone() for v in spec.vars}}
    if opt_m is not None:
        blob["adam_m"] = {_scope(v.name): params.view(v.index, opt_m).detach().cpu().clone() for v in spec.vars}
    if opt_v is not None:
        blob["adam_v"] = {_scope(v.name): params.view(v.index, opt_v).detach().cpu().clone() for v in spec.vars}
    if extra:
        blob["extra"] = extra
    tmp = path + ".tmp"
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    torch.save(blob, tmp)
    os.replace(tmp, path)            # atomic: a crash never leaves a torn checkpoint


def load_checkpoint(path: str, params: Arena, opt_m: Optional[torch.Tensor] = None,
                    opt_v: Optional[torch.Tensor] = None) -> Dict:
    blob = torch.load(path, map_location="cpu", weights_only=False)
    if blob.get("format") != FORMAT_VERSION:
        raise ValueError("unsupported checkpoint format %r" % blob.get("format"))
    spec = params.spec
    if blob["model"]["kind"] != spec.kind or blob["model"]["width_mult"] != spec.width_mult:
        raise ValueError("checkpoint is for model %r, not %s x%d" % (blob["model"], spec.kind, spec.width_mult))
    with torch.no_grad():
        for v in spec.vars:
            t = blob["vars"][_scope(v.name)]
            if tuple(t.shape) != v.shape:
                raise ValueError("shape mismatch for %s" % v.name)
            params.view(v.index).copy_(t)
            if opt_m is not None and "adam_m" in blob:
                params.view(v.index, opt_m).copy_(blob["adam_m"][_scope(v.name)])
            if opt_v is not None and "adam_v" in blob:
                params.view(v.index, opt_v).copy_(blob["adam_v"][_scope(v.name)])
    return {"step": blob["step"], "t": blob["t"], "extra": blob.get("extra", {})}

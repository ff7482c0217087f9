"""Tensor-level wrappers of the individual sm_100a kernels (the public ``ops`` API).

Each function maps one implicit TF op of the reference onto one kernel launch of
libddl_b200.so.  Layouts are the reference's: NHWC activations, HWIO conv kernels, ``[in, out]`` FC
matrices.  These wrappers build a throw-away TMA plan per call - fine for tests and one-off use;
the training hot path goes through the plan-caching engine (ops/engine.py).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import torch

from . import native

EPI_BIAS, EPI_RELU, EPI_DROPOUT, EPI_MASKPOS = 1, 2, 4, 8


def _run_plan(plan) -> None:
    L = native.lib()
    if not plan:
        raise native.NativeError("plan creation failed: " + native.last_error())
    try:
        native.check(L.ddl_tc_plan_run(plan, native.stream_ptr()), "ddl_tc_plan_run")
        native.count()
        torch.cuda.current_stream().synchronize()      # the plan (TMA descriptors) dies below
    finally:
        L.ddl_tc_plan_destroy(plan)


def plan_info(plan) -> dict:
    arr = (C.c_int * 8)()
    native.check(native.lib().ddl_tc_plan_info(plan, arr), "ddl_tc_plan_info")
    return dict(grid=(arr[0], arr[1], arr[2]), stages=arr[3], smem=arr[4], box=(arr[5], arr[6], arr[7]))


def conv5x5_fprop(x: torch.Tensor, w: torch.Tensor, ntile: int = 0, ksplit: int = 1) -> torch.Tensor:
    """tf.nn.conv2d(x, w, strides 1, SAME) (model.py:39,50,61): x [N,H,W,Cin], w [k,k,Cin,Cout] -> raw sums."""
    N, H, W, Cin = x.shape
    k, _, _, Cout = w.shape
    out = (torch.zeros if ksplit > 1 else torch.empty)(N, H, W, Cout, device=x.device, dtype=torch.float32)
    nt = ntile or (128 if Cout % 128 == 0 else 64 if Cout % 64 == 0 else 32)
    plan = native.lib().ddl_tc_plan_act(0, native.ptr(x), native.ptr(w), native.ptr(out), N, H, W, Cin, Cout, k, nt,
                                        ksplit, 0, None, None, 1.0, None, 0, 1.0, None, None)
    _run_plan(plan)
    return out


def conv5x5_dgrad(dy: torch.Tensor, w: torch.Tensor, ntile: int = 0, ksplit: int = 1) -> torch.Tensor:
    """conv2d backprop-input: dy [N,H,W,Cout], w [k,k,Cin,Cout] -> dx [N,H,W,Cin]."""
    N, H, W, Cout = dy.shape
    k, _, Cin, _ = w.shape
    out = (torch.zeros if ksplit > 1 else torch.empty)(N, H, W, Cin, device=dy.device, dtype=torch.float32)
    nt = ntile or (128 if Cin % 128 == 0 else 64 if Cin % 64 == 0 else 32)
    plan = native.lib().ddl_tc_plan_act(1, native.ptr(dy), native.ptr(w), native.ptr(out), N, H, W, Cout, Cin, k, nt,
                                        ksplit, 0, None, None, 1.0, None, 0, 1.0, None, None)
    _run_plan(plan)
    return out


def conv5x5_relu_pool(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor, ntile: int = 0, cluster: int = 1,
                      want_idx: bool = True):
    """conv + bias + ReLU + 2x2 SAME max-pool in ONE tcgen05 kernel (model.py:39-42); returns (pooled, idx)."""
    N, H, W, Cin = x.shape
    k, Cout = w.shape[0], w.shape[3]
    nt = ntile or min(128, Cout)
    y = torch.empty(N, (H + 1) // 2, (W + 1) // 2, Cout, device=x.device, dtype=torch.float32)
    idx = torch.empty(y.shape, device=x.device, dtype=torch.uint8) if want_idx else None
    plan = native.lib().ddl_tc_plan_conv_pool(native.ptr(x), native.ptr(w), native.ptr(bias), native.ptr(y),
                                              native.ptr(idx), N, H, W, Cin, Cout, k, nt, cluster)
    _run_plan(plan)
    return y, idx


def conv5x5_dgrad_unpool(dy: torch.Tensor, w: torch.Tensor, idx: torch.Tensor, UH: int, UW: int, ntile: int = 0,
                         cluster: int = 1, want_db: bool = True):
    """conv dgrad + pool/ReLU backward of the layer below in ONE kernel; returns (dfull [N,UH,UW,Cin], dbias)."""
    N, H, W, Cout = dy.shape
    k, Cin = w.shape[0], w.shape[2]
    nt = ntile or min(128, Cin)
    out = torch.empty(N, UH, UW, Cin, device=dy.device, dtype=torch.float32)
    db = torch.zeros(Cin, device=dy.device, dtype=torch.float32) if want_db else None
    plan = native.lib().ddl_tc_plan_dgrad_unpool(native.ptr(dy), native.ptr(w), native.ptr(idx), native.ptr(out),
                                                 native.ptr(db), N, H, W, UH, UW, Cout, Cin, k, nt, cluster)
    _run_plan(plan)
    return out, db


def conv5x5_dgrad_poolgrad(dy: torch.Tensor, w: torch.Tensor, idx: torch.Tensor, ntile: int = 0, cluster: int = 1,
                           want_db: bool = True):
    """conv dgrad + the ReLU half of the pool backward, gradient kept in the pooled layout: returns
    (gp [N,H,W,Cin] with ReLU-dead windows zeroed, dbias)."""
    N, H, W, Cout = dy.shape
    k, Cin = w.shape[0], w.shape[2]
    nt = ntile or min(128, Cin)
    out = torch.empty(N, H, W, Cin, device=dy.device, dtype=torch.float32)
    db = torch.zeros(Cin, device=dy.device, dtype=torch.float32) if want_db else None
    plan = native.lib().ddl_tc_plan_dgrad_poolgrad(native.ptr(dy), native.ptr(w), native.ptr(idx), native.ptr(out),
                                                   native.ptr(db), N, H, W, Cout, Cin, k, nt, cluster)
    _run_plan(plan)
    return out, db


def _run_patch_plan(plan):
    if not plan:
        raise native.NativeError("patch plan creation failed: " + native.last_error())
    try:
        native.check(native.lib().ddl_patch_plan_run(plan, native.stream_ptr()), "ddl_patch_plan_run")
    finally:
        native.lib().ddl_patch_plan_destroy(plan)


def conv5x5_relu_pool_patch(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor, want_idx: bool = True):
    """conv + bias + ReLU + 2x2 SAME max-pool with the patch-resident kernel (csrc/tc_patch.cu); returns (pooled, idx)."""
    N, H, W, Cin = x.shape
    Cout = w.shape[3]
    y = torch.empty(N, (H + 1) // 2, (W + 1) // 2, Cout, device=x.device, dtype=torch.float32)
    idx = torch.empty(y.shape, device=x.device, dtype=torch.uint8) if want_idx else None
    plan = native.lib().ddl_patch_plan_create(0, native.ptr(x), native.ptr(w), native.ptr(bias), native.ptr(y), native.ptr(idx),
                                              None, N, H, W, Cin, Cout)
    _run_patch_plan(plan)
    return y, idx


def conv5x5_dgrad_poolgrad_patch(dy: torch.Tensor, w: torch.Tensor, idx: torch.Tensor, want_db: bool = True):
    """conv dgrad + ReLU mask of the layer below, gradient kept pooled, with the patch-resident kernel."""
    N, H, W, Cout = dy.shape
    Cin = w.shape[2]
    out = torch.empty(N, H, W, Cin, device=dy.device, dtype=torch.float32)
    db = torch.zeros(Cin, device=dy.device, dtype=torch.float32) if want_db else None
    plan = native.lib().ddl_patch_plan_create(1, native.ptr(dy), native.ptr(w), None, native.ptr(out), native.ptr(idx),
                                              native.ptr(db), N, H, W, Cout, Cin)
    _run_patch_plan(plan)
    return out, db


def conv5x5_wgrad(dy: torch.Tensor, x: torch.Tensor, ksize: int = 5, ntile: int = 0, psplit: int = 1,
                  out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """conv2d backprop-filter: dy [N,H,W,Cout], x [N,H,W,Cin] -> dw [k,k,Cin,Cout] (accumulated into out)."""
    N, H, W, Cout = dy.shape
    Cin = x.shape[3]
    if out is None:
        out = torch.zeros(ksize, ksize, Cin, Cout, device=x.device, dtype=torch.float32)
    nt = ntile or (128 if Cin >= 128 else 64 if Cin >= 64 else 32)
    plan = native.lib().ddl_tc_plan_wgrad(native.ptr(dy), native.ptr(x), native.ptr(out), N, H, W, Cout, Cin,
                                          ksize, nt, psplit)
    _run_plan(plan)
    return out


def linear_fwd(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None, relu: bool = False,
               dropout: Optional[Tuple[torch.Tensor, int, float]] = None, ntile: int = 32,
               ksplit: int = 1) -> torch.Tensor:
    """tf.matmul(x, w) + b [+ relu] [+ dropout] (model.py:70-82).  dropout = (step_state, stream_off, keep)."""
    B, K = x.shape
    N = w.shape[1]
    out = torch.empty(B, N, device=x.device, dtype=torch.float32)
    flags = (EPI_BIAS if bias is not None else 0) | (EPI_RELU if relu else 0) | (EPI_DROPOUT if dropout else 0)
    st, off, keep = dropout if dropout else (None, 0, 1.0)
    plan = native.lib().ddl_tc_plan_act(0, native.ptr(x), native.ptr(w), native.ptr(out), B, 1, 1, K, N, 1, ntile, ksplit,
                                        flags, native.ptr(bias), None, 1.0, native.ptr(st), off, keep, None, None)
    _run_plan(plan)
    return out


def linear_dx(dy: torch.Tensor, w: torch.Tensor, mask_act: Optional[torch.Tensor] = None, mask_scale: float = 1.0,
              ntile: int = 32, colsum: Optional[torch.Tensor] = None) -> torch.Tensor:
    """dx = dy @ w^T, optionally multiplied by (mask_act > 0) * mask_scale (ReLU + dropout backward)."""
    B, N = dy.shape
    K = w.shape[0]
    out = torch.empty(B, K, device=dy.device, dtype=torch.float32)
    plan = native.lib().ddl_tc_plan_act(1, native.ptr(dy), native.ptr(w), native.ptr(out), B, 1, 1, N, K, 1, ntile, 1,
                                        EPI_MASKPOS if mask_act is not None else 0, None, native.ptr(mask_act),
                                        mask_scale, None, 0, 1.0, native.ptr(colsum), None)
    _run_plan(plan)
    return out


def linear_dw(dy: torch.Tensor, x: torch.Tensor, ntile: int = 64, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """dw = x^T @ dy: dy [B,N], x [B,K] -> [K,N] (accumulated into out)."""
    B, N = dy.shape
    K = x.shape[1]
    if out is None:
        out = torch.zeros(K, N, device=x.device, dtype=torch.float32)
    plan = native.lib().ddl_tc_plan_wgrad(native.ptr(dy), native.ptr(x), native.ptr(out), B, 1, 1, N, K, 1, ntile, 1)
    _run_plan(plan)
    return out


def conv1_relu_pool(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor, want_idx: bool = True):
    """conv1 (Cin=1) + bias + ReLU + 2x2 SAME pool (model.py:24-31): x [B,S,S] -> ([B,S/2,S/2,C], idx)."""
    B, S = x.shape[0], x.shape[1]
    Cc = w.shape[-1]
    y = torch.empty(B, S // 2, S // 2, Cc, device=x.device, dtype=torch.float32)
    idx = torch.empty(B, S // 2, S // 2, Cc, device=x.device, dtype=torch.uint8) if want_idx else None
    native.check(native.lib().ddl_conv1_relu_pool(native.ptr(x), native.ptr(w), native.ptr(b), native.ptr(y),
                                                  native.ptr(idx), B, S, Cc, native.stream_ptr()), "conv1_relu_pool")
    native.count()
    return y, idx


def conv1_wgrad(x: torch.Tensor, dy: torch.Tensor):
    B, S = x.shape[0], x.shape[1]
    Cc = dy.shape[-1]
    dw = torch.zeros(5, 5, 1, Cc, device=x.device, dtype=torch.float32)
    db = torch.zeros(Cc, device=x.device, dtype=torch.float32)
    native.check(native.lib().ddl_conv1_wgrad(native.ptr(x), native.ptr(dy), native.ptr(dw), native.ptr(db), B, S, Cc,
                                              native.stream_ptr()), "conv1_wgrad")
    native.count()
    return dw, db


def conv1_wgrad_sparse(x: torch.Tensor, dy: torch.Tensor, idx: torch.Tensor):
    """conv1 weight / bias gradient reading only the argmax pixels of the un-pooled gradient ``dy``."""
    B, S = x.shape[0], x.shape[1]
    Cc = dy.shape[-1]
    dw = torch.zeros(5, 5, 1, Cc, device=x.device, dtype=torch.float32)
    db = torch.zeros(Cc, device=x.device, dtype=torch.float32)
    native.check(native.lib().ddl_conv1_wgrad_sparse(native.ptr(x), native.ptr(dy), native.ptr(idx), native.ptr(dw),
                                                     native.ptr(db), B, S, Cc, native.stream_ptr()), "conv1_wgrad_sparse")
    native.count()
    return dw, db


def conv1_wgrad_pooled(x: torch.Tensor, gp: torch.Tensor, idx: torch.Tensor, want_db: bool = True):
    """conv1 weight (/ bias) gradient from the pooled gradient ``gp`` [B,S/2,S/2,C] and the argmax slots ``idx``."""
    B, S = x.shape[0], x.shape[1]
    Cc = gp.shape[-1]
    dw = torch.zeros(5, 5, 1, Cc, device=x.device, dtype=torch.float32)
    db = torch.zeros(Cc, device=x.device, dtype=torch.float32) if want_db else None
    native.check(native.lib().ddl_conv1_wgrad_pooled(native.ptr(x), native.ptr(gp), native.ptr(idx), native.ptr(dw),
                                                     native.ptr(db), B, S, Cc, native.stream_ptr()), "conv1_wgrad_pooled")
    return dw, db


def bias_relu_maxpool(a: torch.Tensor, bias: Optional[torch.Tensor] = None, want_idx: bool = True):
    """(bias +) ReLU + 2x2/2 SAME max-pool of NHWC conv sums (model.py:30-31)."""
    B, H, W, Cc = a.shape
    y = torch.empty(B, (H + 1) // 2, (W + 1) // 2, Cc, device=a.device, dtype=torch.float32)
    idx = torch.empty(y.shape, device=a.device, dtype=torch.uint8) if want_idx else None
    native.check(native.lib().ddl_maxpool_fwd(native.ptr(a), native.ptr(bias), native.ptr(y), native.ptr(idx), B, H, W,
                                              Cc, native.stream_ptr()), "maxpool_fwd")
    native.count()
    return y, idx


def maxpool_relu_bwd(dp: torch.Tensor, idx: torch.Tensor, H: int, W: int, want_db: bool = True):
    B, _, _, Cc = dp.shape
    dy = torch.empty(B, H, W, Cc, device=dp.device, dtype=torch.float32)
    db = torch.zeros(Cc, device=dp.device, dtype=torch.float32) if want_db else None
    native.check(native.lib().ddl_maxpool_bwd(native.ptr(dp), native.ptr(idx), native.ptr(dy), native.ptr(db), B, H, W,
                                              Cc, native.stream_ptr()), "maxpool_bwd")
    native.count()
    return dy, db


def fc3_softmax_xent(h: torch.Tensor, w: torch.Tensor, b: torch.Tensor, labels: Optional[torch.Tensor],
                     scale: Optional[float] = None):
    """fc3 + softmax cross-entropy (model.py:85-92).  Returns (logits, dlogits, loss_sum, correct)."""
    B, K = h.shape
    NC = w.shape[1]
    st = native.new_step_state()
    logits = torch.empty(B, NC, device=h.device, dtype=torch.float32)
    dl = torch.empty(B, NC, device=h.device, dtype=torch.float32) if labels is not None else None
    native.check(native.lib().ddl_fc3_xent_fwd(native.ptr(h), native.ptr(w), native.ptr(b), native.ptr(labels),
                                               native.ptr(logits), native.ptr(dl), native.ptr(st), B, K, NC,
                                               (1.0 / B) if scale is None else scale, native.stream_ptr()), "fc3_xent")
    native.count()
    s = native.read_step_state(st)
    return logits, dl, s["loss_sum"], s["correct"]


def fc3_bwd(h: torch.Tensor, w: torch.Tensor, dl: torch.Tensor, seed: int = 0, stream: int = 0, keep_prob: float = 1.0):
    B, K = h.shape
    NC = w.shape[1]
    dw = torch.zeros(K, NC, device=h.device, dtype=torch.float32)
    db = torch.zeros(NC, device=h.device, dtype=torch.float32)
    dh = torch.empty(B, K, device=h.device, dtype=torch.float32)
    native.check(native.lib().ddl_fc3_bwd(native.ptr(h), native.ptr(w), native.ptr(dl), native.ptr(dw), native.ptr(db),
                                          native.ptr(dh), B, K, NC, seed, stream, keep_prob, native.stream_ptr()),
                 "fc3_bwd")
    native.count()
    return dw, db, dh


def fc3_fused(h: torch.Tensor, w: torch.Tensor, b: torch.Tensor, labels: torch.Tensor, st: torch.Tensor,
              keep_prob: float = 1.0, stream_off: int = 1):
    """fc3 forward + softmax-xent + fc3 backward in one cluster kernel; returns (dw, db, dh, dbh); loss / correct are
    added to the device step state ``st``."""
    B, K = h.shape
    NC = w.shape[1]
    dw = torch.zeros(K, NC, device=h.device); db = torch.zeros(NC, device=h.device)
    dh = torch.empty(B, K, device=h.device); dbh = torch.zeros(K, device=h.device)
    native.check(native.lib().ddl_fc3_fused_st(native.ptr(h), native.ptr(w), native.ptr(b), native.ptr(labels), native.ptr(dw),
                                               native.ptr(db), native.ptr(dh), native.ptr(dbh), st.data_ptr(), B, K, NC,
                                               1.0 / B, stream_off, keep_prob, native.stream_ptr()), "ddl_fc3_fused_st")
    native.count()
    return dw, db, dh, dbh


def colsum(x: torch.Tensor) -> torch.Tensor:
    out = torch.zeros(x.shape[1], device=x.device, dtype=torch.float32)
    native.check(native.lib().ddl_colsum(native.ptr(x), native.ptr(out), x.shape[0], x.shape[1], native.stream_ptr()),
                 "colsum")
    native.count()
    return out

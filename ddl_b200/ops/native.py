"""ctypes binding of libddl_b200.so (the C ABI declared in include/ddl_b200.h).

The library is built in-tree by ``ddl_b200.ops.build`` (plain nvcc, sm_100a).  Loading is
mandatory on a GPU box: ``lib()`` raises if the .so is missing or stale instead of letting callers
fall back to PyTorch ops silently.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

from . import build as _build

MAX_PEERS = 16
MAX_SKIP = 4
MAX_SUB = 16
STALE_BINS = 16
PAD_SLOTS = 256
PAD_PUSHES = 32
PAD_VERSION = 33
PAD_PUSHES_W = 48


class StepState(C.Structure):
    _fields_ = [("t", C.c_uint32), ("lr_t", C.c_float), ("loss_sum", C.c_float), ("correct", C.c_uint32),
                ("seed", C.c_uint32), ("stream", C.c_uint32), ("wstep", C.c_uint32),
                ("stream_base", C.c_uint32), ("stream_stride", C.c_uint32), ("loss_mean", C.c_float),
                ("pad", C.c_uint32 * 6)]


assert C.sizeof(StepState) == 64


class Comm(C.Structure):
    _fields_ = [("rank", C.c_int), ("world", C.c_int),
                ("grads", C.c_void_p * MAX_PEERS), ("params", C.c_void_p * MAX_PEERS),
                ("accum", C.c_void_p * MAX_PEERS), ("pads", C.c_void_p * MAX_PEERS),
                ("mc_grads", C.c_void_p), ("mc_params", C.c_void_p)]


class SyncArgs(C.Structure):
    _fields_ = [("begin", C.c_int64), ("end", C.c_int64), ("m", C.c_void_p), ("v", C.c_void_p),
                ("st", C.c_void_p), ("done_counter", C.c_void_p), ("worker_mask", C.c_uint32),
                ("dest_mask", C.c_uint32), ("lr", C.c_float), ("a", C.c_float), ("b", C.c_float),
                ("eps", C.c_float), ("gscale", C.c_float),
                ("skip_begin", C.c_int64 * MAX_SKIP), ("skip_end", C.c_int64 * MAX_SKIP),
                ("nsub", C.c_int), ("bucket", C.c_int), ("nbuckets", C.c_int),
                ("sub_begin", C.c_int64 * MAX_SUB), ("sub_end", C.c_int64 * MAX_SUB)]


class AsyncArgs(C.Structure):
    _fields_ = [("num_shards", C.c_int), ("owner_rank", C.c_int * MAX_PEERS),
                ("shard_begin", C.c_int64 * MAX_PEERS), ("shard_end", C.c_int64 * MAX_PEERS),
                ("direct_sgd", C.c_int), ("lr", C.c_float), ("done_counter", C.c_void_p),
                ("pulled_version", C.c_void_p), ("stale_hist", C.c_void_p),
                ("slot", C.c_int), ("slot_stride", C.c_int64)]


class AsyncState(C.Structure):
    _fields_ = [("applied", C.c_uint32 * MAX_PEERS), ("mask", C.c_uint32), ("nsteps", C.c_uint32),
                ("folded", C.c_uint32), ("pad", C.c_uint32), ("lr_t", C.c_float * MAX_PEERS)]


class EngineCfg(C.Structure):
    _fields_ = [("kind", C.c_int), ("width_mult", C.c_int), ("batch", C.c_int), ("max_eval_batch", C.c_int),
                ("keep_prob", C.c_float), ("params", C.c_void_p), ("grads", C.c_void_p),
                ("arena_elems", C.c_int64), ("var_offset", C.c_int64 * 16), ("st", C.c_void_p),
                ("fc_w_src", C.c_void_p * 2)]


_LIB: Optional[C.CDLL] = None


class NativeError(RuntimeError):
    pass


def lib_path() -> str:
    return _build.LIB


def lib() -> C.CDLL:
    """Loads (building first if sources changed and nvcc is present) the native library."""
    global _LIB
    if _LIB is not None:
        return _LIB
    lab = os.environ.get("DDL_LAB", "") not in ("", "0")      # instrumented build for the tools/ phase and trace scripts
    path = _build.LAB_LIB if lab else _build.LIB
    if _build.needs_build(lab):
        try:
            _build.build(lab=lab)
        except Exception as e:                      # no nvcc on the box: a stale .so is an error, not a fallback
            if not os.path.exists(path):
                raise NativeError("libddl_b200.so is missing and could not be built: %s" % e)
    try:
        L = C.CDLL(path)
    except OSError as e:
        raise NativeError("cannot load %s: %s" % (path, e))
    L.ddl_last_error.restype = C.c_char_p
    vp, i32, i64, u32, f32 = C.c_void_p, C.c_int, C.c_int64, C.c_uint32, C.c_float
    sigs = {
        "ddl_version": ([], i32),
        "ddl_launch_count": ([], C.c_uint64),
        "ddl_step_advance": ([vp, f32, f32, f32, i32, vp], i32),
        "ddl_optimizer_step": ([i32, vp, vp, vp, vp, i64, f32, f32, f32, f32, f32, vp, vp], i32),
        "ddl_conv1_relu_pool": ([vp, vp, vp, vp, vp, i32, i32, i32, vp], i32),
        "ddl_conv1_wgrad": ([vp, vp, vp, vp, i32, i32, i32, vp], i32),
        "ddl_conv1_wgrad_sparse": ([vp, vp, vp, vp, vp, i32, i32, i32, vp], i32),
        "ddl_conv1_wgrad_pooled": ([vp, vp, vp, vp, vp, i32, i32, i32, vp], i32),
        "ddl_maxpool_fwd": ([vp, vp, vp, vp, i32, i32, i32, i32, vp], i32),
        "ddl_maxpool_bwd": ([vp, vp, vp, vp, i32, i32, i32, i32, vp], i32),
        "ddl_colsum": ([vp, vp, i32, i32, vp], i32),
        "ddl_fc3_xent_fwd": ([vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, f32, vp], i32),
        "ddl_fc3_bwd": ([vp, vp, vp, vp, vp, vp, i32, i32, i32, u32, u32, f32, vp], i32),
        "ddl_fc3_bwd_st": ([vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, vp, u32, f32, vp], i32),
        "ddl_fc3_fused_st": ([vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, f32, u32, f32, vp], i32),
        "ddl_fc3_fused_fits": ([i32, i32, i32], i32),
        "ddl_accuracy": ([vp, vp, vp, i32, i32, vp], i32),
        "ddl_tc_plan_act": ([i32, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp, vp, f32, vp, u32, f32, vp, vp], vp),
        "ddl_tc_plan_wgrad": ([vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32], vp),
        "ddl_tc_plan_conv_pool": ([vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32], vp),
        "ddl_tc_plan_dgrad_unpool": ([vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32], vp),
        "ddl_tc_plan_dgrad_poolgrad": ([vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32], vp),
        "ddl_plan_choose_box": ([i32, i32, i32, i32, i32, i32, C.POINTER(C.c_int)], i32),
        "ddl_plan_pick_staged": ([i32, i32, i32, C.POINTER(C.c_int)], i32),
        "ddl_plan_pick_fc": ([i32, i32, C.POINTER(C.c_int)], i32),
        "ddl_tc_plan_run": ([vp, vp], i32),
        "ddl_tc_plan_destroy": ([vp], None),
        "ddl_tc_plan_info": ([vp, C.POINTER(C.c_int)], i32),
        "ddl_tc_plan_needs_zero": ([vp], i32),
        "ddl_dev_malloc": ([C.POINTER(vp), C.c_size_t], i32),
        "ddl_dev_free": ([vp], i32),
        "ddl_ipc_export": ([vp, vp], i32),
        "ddl_ipc_open": ([vp, C.POINTER(vp)], i32),
        "ddl_ipc_close": ([vp], i32),
        "ddl_set_device": ([i32], i32),
        "ddl_can_access_peer": ([i32, i32], i32),
        "ddl_sync_push_update_pull": ([C.POINTER(Comm), C.POINTER(SyncArgs), i32, vp], i32),
        "ddl_p2p_latency": ([vp, i32, i32, u32, i32, vp, vp], i32),
        "ddl_p2p_copy": ([vp, vp, i64, i32, vp], i32),
        "ddl_p2p_spin": ([vp, i32, C.c_uint64, vp, vp], i32),
        "ddl_patch_conv_fits": ([i32, i32, i32, i32], i32),
        "ddl_patch_conv_preferred": ([i32, i32, i32, i32], i32),
        "ddl_patch_plan_create": ([i32, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32], vp),
        "ddl_patch_plan_run": ([vp, vp], i32),
        "ddl_patch_plan_destroy": ([vp], None),
        "ddl_patch_plan_set_trace": ([vp, vp], i32),
        "ddl_async_owner_start": ([C.POINTER(Comm), C.POINTER(SyncArgs), i32, vp, vp, i32, i64, i32, f32, f32, vp], i32),
        "ddl_async_push": ([C.POINTER(Comm), C.POINTER(AsyncArgs), vp], i32),
        "ddl_async_apply": ([C.POINTER(Comm), C.POINTER(SyncArgs), i32, vp, i32, i64, i32, f32, f32, vp], i32),
        "ddl_pull_allgather": ([C.POINTER(Comm), C.POINTER(AsyncArgs), vp], i32),
        "ddl_device_barrier": ([C.POINTER(Comm), vp, vp], i32),
        "ddl_engine_create": ([C.POINTER(EngineCfg)], vp),
        "ddl_engine_destroy": ([vp], None),
        "ddl_engine_train_step": ([vp, vp, vp, vp], i32),
        "ddl_engine_set_advance": ([vp, f32, f32, f32, i32], i32),
        "ddl_engine_set_overlap": ([vp, C.POINTER(Comm), vp, i32, i32], i32),
        "ddl_engine_eval": ([vp, vp, vp, i32, vp, vp, vp], i32),
        "ddl_engine_buffer": ([vp, i32], vp),
        "ddl_memcpy_async": ([vp, vp, C.c_size_t, vp], i32),
    }
    for name, (args, res) in sigs.items():
        fn = getattr(L, name)
        fn.argtypes = args
        fn.restype = res
    _LIB = L
    return L


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        raise NativeError("%s failed (%d): %s" % (what or "native call", rc, lib().ddl_last_error().decode()))


def last_error() -> str:
    return lib().ddl_last_error().decode()


def stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    if t is None:
        return None
    assert t.is_cuda and t.is_contiguous(), "native ops need contiguous CUDA tensors"
    return t.data_ptr()


def launches() -> int:
    """Kernels launched by the native library in this process so far: counted inside the library, in the one launch
    path every kernel uses (``ddl::launch``).  A launch recorded during stream capture counts once."""
    return int(lib().ddl_launch_count())


def count(n: int = 1) -> None:
    """Kept for callers of the round-1 binding: launches are counted natively now (``launches()``)."""


# --------------------------------------------------------------------------------------- helpers
def new_step_state(seed: int = 0, worker: int = 0, num_workers: int = 1) -> torch.Tensor:
    """A device-resident DdlStepState as a 16-word int32 tensor."""
    st = torch.zeros(16, dtype=torch.int32, device="cuda")
    host = torch.zeros(16, dtype=torch.int32)
    host[4] = seed & 0x7FFFFFFF
    host[7] = 2 * worker
    host[8] = 2 * num_workers
    st.copy_(host)
    return st


def read_step_state(st: torch.Tensor) -> dict:
    h = st.cpu()
    f = h.view(torch.float32)
    return {"t": int(h[0]), "lr_t": float(f[1]), "loss_sum": float(f[2]), "correct": int(h[3]),
            "seed": int(h[4]), "stream": int(h[5]), "wstep": int(h[6]), "loss_mean": float(f[9])}


def optimizer_step(opt, p: torch.Tensor, g: torch.Tensor, gscale: float = 1.0, st: Optional[torch.Tensor] = None):
    """FlatOptimizer.step on the GPU (opt.t was already incremented by the caller)."""
    kind, step, a, b, eps = opt.hyper(opt.t)
    n = p.numel()
    check(lib().ddl_optimizer_step(kind, ptr(p), ptr(opt.m), ptr(opt.v), ptr(g), n, step, a, b, eps, gscale,
                                   ptr(st), stream_ptr()), "ddl_optimizer_step")
    count()

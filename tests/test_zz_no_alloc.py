"""The allocation-free conv block, called through the C ABI and checked against the existing entry points.

``ddl_conv5x5_relu_pool_static`` exists for host programs that must not call cudaMalloc: ones that gave
the whole device to another allocator (a caching pool owning all free memory, an MPS / MIG slice with a
hard quota - a further cudaMalloc would simply fail) and latency-critical loops that cannot take
cudaMalloc's device-wide synchronisation.  Its raw-sum buffer is an 8 MiB array built into the library, so
the caller passes only inputs and outputs.

Two oracles: the library's own two-call path with caller-provided scratch (``ddl_tc_plan_act`` +
``ddl_maxpool_fwd``), and torch's conv2d + relu + max_pool2d in fp64.  Three calls in a row with different
shapes - two that split K (the function clears what it accumulates into) and one that does not (it overwrites
the array as it is) - make sure the shared array's stale contents cannot leak from one call into the next.  Device
free memory (the driver's count, not torch's) must not shrink across a call.  This file sorts last on purpose: if it fails under
``-x``, every other test has already run.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

TF32_TOL = 4e-3       # same bound as tests/test_kernels_gpu.py


def _lib():
    from ddl_b200.ops import native
    L = native.lib()
    L.ddl_conv5x5_relu_pool_static.argtypes = [C.c_void_p] * 5 + [C.c_int] * 5 + [C.c_void_p]
    L.ddl_conv5x5_relu_pool_static.restype = C.c_int
    return L, native


def _reference(L, native, x, w, b):
    """Existing entry points: conv sums into a caller-allocated buffer, then bias + ReLU + pool."""
    N, H, W, cin = x.shape
    cout = w.shape[3]
    raw = torch.empty(N, H, W, cout, device=x.device)
    nt = 128 if cout % 128 == 0 else 64 if cout % 64 == 0 else 32
    plan = L.ddl_tc_plan_act(0, x.data_ptr(), w.data_ptr(), raw.data_ptr(), N, H, W, cin, cout, 5, nt, 1, 0,
                             None, None, 1.0, None, 0, 1.0, None, None)
    assert plan, native.last_error()
    s = torch.cuda.current_stream().cuda_stream
    assert L.ddl_tc_plan_run(plan, s) == 0, native.last_error()
    y = torch.empty(N, (H + 1) // 2, (W + 1) // 2, cout, device=x.device)
    idx = torch.empty(y.shape, dtype=torch.uint8, device=x.device)
    assert L.ddl_maxpool_fwd(raw.data_ptr(), b.data_ptr(), y.data_ptr(), idx.data_ptr(), N, H, W, cout, s) == 0
    torch.cuda.synchronize()
    L.ddl_tc_plan_destroy(plan)
    return y, idx


def _static(L, native, x, w, b, want_idx=True):
    N, H, W, cin = x.shape
    cout = w.shape[3]
    y = torch.full((N, (H + 1) // 2, (W + 1) // 2, cout), float("nan"), device=x.device)
    idx = torch.zeros(y.shape, dtype=torch.uint8, device=x.device) if want_idx else None
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    free_before, _ = torch.cuda.mem_get_info()             # the DRIVER's view: a raw cudaMalloc inside the call would show here
    rc = L.ddl_conv5x5_relu_pool_static(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(),
                                        idx.data_ptr() if want_idx else None,
                                        N, H, W, cin, cout, torch.cuda.current_stream().cuda_stream)
    assert rc == 0, native.last_error()
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert torch.cuda.memory_allocated() == before        # nothing came from torch's allocator either
    assert free_after >= free_before, "device free memory shrank by %d bytes during the call" % (free_before - free_after)
    return y, idx


def _torch_oracle(x, w, b):
    """Independent of this library: fp64 conv2d + bias + ReLU + 2x2 max-pool (SAME)."""
    import torch.nn.functional as F
    a = F.relu(F.conv2d(x.permute(0, 3, 1, 2).double(), w.permute(3, 2, 0, 1).double(), b.double(), padding=2))
    return F.max_pool2d(a, 2, 2, ceil_mode=True).permute(0, 2, 3, 1).float()


def test_static_scratch_conv_matches_existing_entry_points_twice():
    L, native = _lib()
    dev = torch.device("cuda")
    torch.manual_seed(0)
    # first call: 7x7, 64 -> 128 (split-K path, accumulates with red.add into the shared array);
    # second call: a smaller 4x4, 128 -> 256 problem that reuses the front of the same array
    # third call: 14x14, 32 -> 32 over 256 images - enough tiles that the kernel does NOT split K: it overwrites the
    # array without clearing it first, on top of what the two split-K calls left there
    for (N, H, cin, cout) in [(100, 7, 64, 128), (24, 4, 128, 256), (256, 14, 32, 32)]:
        x = torch.randn(N, H, H, cin, device=dev)
        w = torch.randn(5, 5, cin, cout, device=dev) * 0.05
        b = torch.randn(cout, device=dev) * 0.1
        y_ref, idx_ref = _reference(L, native, x, w, b)
        y, idx = _static(L, native, x, w, b)
        assert not torch.isnan(y).any()
        err = float((y - y_ref).abs().max() / (y_ref.abs().max() + 1e-12))
        assert err < TF32_TOL, "static-scratch conv differs from the planned path: %g" % err
        # arg-max bytes agree wherever the pooled value is clearly positive (ties may break differently)
        clear = y_ref > 1e-3
        assert float((idx[clear] == idx_ref[clear]).float().mean()) > 0.999
        # ... and against an oracle that shares no code with the library
        y_t = _torch_oracle(x, w, b)
        err_t = float((y - y_t).abs().max() / (y_t.abs().max() + 1e-12))
        assert err_t < TF32_TOL, "static-scratch conv differs from torch conv2d + relu + max_pool2d: %g" % err_t
        # the arg-max output is optional
        y2, none = _static(L, native, x, w, b, want_idx=False)
        # (split-K sums are accumulated with red.add in whatever order the CTAs arrive: equal to rounding, not bitwise)
        assert none is None and float((y2 - y).abs().max() / (y.abs().max() + 1e-12)) < 1e-5


def test_static_scratch_rejects_oversize_with_a_clear_error():
    L, native = _lib()
    dev = torch.device("cuda")
    N, H, cin, cout = 400, 14, 32, 64                     # 400*14*14*64*4 B = 20 MB > 8 MiB
    x = torch.zeros(N, H, H, cin, device=dev)
    w = torch.zeros(5, 5, cin, cout, device=dev)
    b = torch.zeros(cout, device=dev)
    y = torch.zeros(N, 7, 7, cout, device=dev)
    rc = L.ddl_conv5x5_relu_pool_static(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), None, N, H, H, cin, cout,
                                        torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = native.last_error()
    assert "exceeds" in msg and "8388608" in msg

"""Times individual tcgen05 plans with CUDA events (warm L2).  usage: python tools/tc_perf.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl_b200.ops import native

dev = torch.device("cuda")
L = native.lib()


def time_plan(plan, reps=30):
    s = native.stream_ptr()
    for _ in range(3):
        L.ddl_tc_plan_run(plan, s)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        L.ddl_tc_plan_run(plan, s)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / reps


def act(dgrad, N, H, cred, nout, ks, ntile, ksplit=1):
    x = torch.randn(N, H, H, cred, device=dev)
    w = torch.randn(ks * ks * cred * nout, device=dev) * 0.05
    out = torch.zeros(N, H, H, nout, device=dev)
    plan = L.ddl_tc_plan_act(dgrad, x.data_ptr(), w.data_ptr(), out.data_ptr(), N, H, H, cred, nout, ks, ntile, ksplit, 0,
                             None, None, 1.0, None, 0, 1.0, None, None)
    assert plan, native.last_error()
    import ctypes as C
    info = (C.c_int * 8)(); L.ddl_tc_plan_info(plan, info)
    t = time_plan(plan)
    kit = ks * ks * ((cred + 31) // 32)
    flops = 2.0 * N * H * H * cred * nout * ks * ks
    print("%s N=%d H=%d cred=%d nout=%d k=%d ntile=%d split=%d grid=%s stages=%d box=%s : %7.1f us  (%.2f us/iter, %.1f TFLOP/s)" % (
        "dgrad" if dgrad else "fprop", N, H, cred, nout, ks, ntile, ksplit, tuple(info[0:3]), info[3], tuple(info[5:8]), t,
        t / (kit / ksplit), flops / t / 1e6), flush=True)
    L.ddl_tc_plan_destroy(plan)
    return t


if __name__ == "__main__":
    print("DDL_TC_DEBUG=%s DDL_TC_STAGES=%s" % (os.environ.get("DDL_TC_DEBUG"), os.environ.get("DDL_TC_STAGES")))
    act(0, 128, 1, 32, 32, 1, 32)          # 1 CTA, 1 iter: fixed overhead
    act(0, 128, 1, 1024, 32, 1, 32)        # 1 CTA, 32 iters
    act(0, 128, 1, 8192, 32, 1, 32)        # 1 CTA, 256 iters
    act(1, 128, 1, 8192, 32, 1, 32)        # K-major B
    act(0, 128, 1, 8192, 128, 1, 128)      # 1 CTA, bigger N
    act(0, 100, 1, 1024, 1024, 1, 32)      # fc1
    act(0, 100, 14, 32, 64, 5, 64)         # conv2
    act(0, 100, 7, 64, 128, 5, 32)         # conv3
    act(0, 100, 4, 128, 256, 5, 32)        # conv4
    act(1, 100, 4, 256, 128, 5, 32, 3)     # conv4 dgrad
    act(0, 2000, 14, 32, 64, 5, 64)        # eval-sized conv2

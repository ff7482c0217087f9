"""Times the fused conv+pool / dgrad+unpool plans against the separate kernels they replace, per layer and per
(ntile, cluster) choice.  CUDA events around 30 back-to-back launches (warm L2).  usage: python tools/pool_perf.py"""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl_b200.ops import native

dev = torch.device("cuda")
L = native.lib()


def time_fn(fn, reps=30):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / reps


def info(plan):
    a = (C.c_int * 8)()
    L.ddl_tc_plan_info(plan, a)
    return tuple(a[0:3]), a[3], tuple(a[5:8])


def fwd(B, H, cin, cout, configs):
    x = torch.randn(B, H, H, cin, device=dev); w = torch.randn(25 * cin * cout, device=dev) * 0.05
    bias = torch.randn(cout, device=dev); raw = torch.zeros(B, H, H, cout, device=dev)
    PH = (H + 1) // 2
    y = torch.empty(B, PH, PH, cout, device=dev); idx = torch.empty(B, PH, PH, cout, device=dev, dtype=torch.uint8)
    s = native.stream_ptr()
    # the separate path as the engine ran it: memset (if split) + conv + pool
    for nt, ks in configs["old"]:
        plan = L.ddl_tc_plan_act(0, x.data_ptr(), w.data_ptr(), raw.data_ptr(), B, H, H, cin, cout, 5, nt, ks, 0, None, None, 1.0,
                                 None, 0, 1.0, None, None)
        assert plan, native.last_error()
        nz = L.ddl_tc_plan_needs_zero(plan)

        def run():
            if nz:
                raw.zero_()
            L.ddl_tc_plan_run(plan, s)
            L.ddl_maxpool_fwd(raw.data_ptr(), bias.data_ptr(), y.data_ptr(), idx.data_ptr(), B, H, H, cout, s)
        print("fwd  H=%d %d->%d separate nt=%d ks=%d grid=%s: %.1f us" % (H, cin, cout, nt, ks, info(plan)[0], time_fn(run)), flush=True)
        L.ddl_tc_plan_destroy(plan)
    for nt, cl in configs["new"]:
        plan = L.ddl_tc_plan_conv_pool(x.data_ptr(), w.data_ptr(), bias.data_ptr(), y.data_ptr(), idx.data_ptr(), B, H, H, cin, cout, 5, nt, cl)
        if not plan:
            print("  (nt=%d cl=%d: %s)" % (nt, cl, native.last_error())); continue
        g, st, box = info(plan)
        print("fwd  H=%d %d->%d FUSED    nt=%d cl=%d grid=%s stages=%d box=%s: %.1f us" % (
            H, cin, cout, nt, cl, g, st, box, time_fn(lambda: L.ddl_tc_plan_run(plan, s))), flush=True)
        L.ddl_tc_plan_destroy(plan)


def bwd(B, H, UH, cin, cout, configs):
    dy = torch.randn(B, H, H, cout, device=dev); w = torch.randn(25 * cin * cout, device=dev) * 0.05
    below = torch.randn(B, UH, UH, cin, device=dev)
    idx = torch.empty(B, H, H, cin, device=dev, dtype=torch.uint8); tmp = torch.empty(B, H, H, cin, device=dev)
    s = native.stream_ptr()
    L.ddl_maxpool_fwd(below.data_ptr(), None, tmp.data_ptr(), idx.data_ptr(), B, UH, UH, cin, s)
    dp = torch.zeros(B, H, H, cin, device=dev); dfull = torch.empty(B, UH, UH, cin, device=dev); db = torch.zeros(cin, device=dev)
    for nt, ks in configs["old"]:
        plan = L.ddl_tc_plan_act(1, dy.data_ptr(), w.data_ptr(), dp.data_ptr(), B, H, H, cout, cin, 5, nt, ks, 0, None, None, 1.0,
                                 None, 0, 1.0, None, None)
        assert plan, native.last_error()
        nz = L.ddl_tc_plan_needs_zero(plan)

        def run():
            if nz:
                dp.zero_()
            L.ddl_tc_plan_run(plan, s)
            L.ddl_maxpool_bwd(dp.data_ptr(), idx.data_ptr(), dfull.data_ptr(), db.data_ptr(), B, UH, UH, cin, s)
        print("bwd  H=%d %d<-%d separate nt=%d ks=%d grid=%s: %.1f us" % (H, cin, cout, nt, ks, info(plan)[0], time_fn(run)), flush=True)
        L.ddl_tc_plan_destroy(plan)
    for nt, cl in configs["new"]:
        plan = L.ddl_tc_plan_dgrad_unpool(dy.data_ptr(), w.data_ptr(), idx.data_ptr(), dfull.data_ptr(), db.data_ptr(), B, H, H, UH, UH,
                                          cout, cin, 5, nt, cl)
        if not plan:
            print("  (nt=%d cl=%d: %s)" % (nt, cl, native.last_error())); continue
        g, st, box = info(plan)
        print("bwd  H=%d %d<-%d FUSED    nt=%d cl=%d grid=%s stages=%d box=%s: %.1f us" % (
            H, cin, cout, nt, cl, g, st, box, time_fn(lambda: L.ddl_tc_plan_run(plan, s))), flush=True)
        L.ddl_tc_plan_destroy(plan)


if __name__ == "__main__":
    B = int(os.environ.get("B", "100"))
    fwd(B, 14, 32, 64, {"old": [(64, 1)], "new": [(64, 1), (32, 1), (64, 2)]})
    fwd(B, 7, 64, 128, {"old": [(128, 3)], "new": [(128, 1), (128, 2), (128, 4), (64, 2), (64, 4), (32, 2)]})
    fwd(B, 4, 128, 256, {"old": [(256, 11)], "new": [(256, 8), (256, 4), (128, 8), (128, 4), (64, 4), (64, 2), (64, 8)]})
    bwd(B, 14, 28, 32, 64, {"old": [(32, 1)], "new": [(32, 1), (32, 2)]})
    bwd(B, 7, 14, 64, 128, {"old": [(64, 3)], "new": [(64, 1), (64, 2), (64, 4), (32, 2), (32, 4)]})
    bwd(B, 4, 7, 128, 256, {"old": [(128, 11)], "new": [(128, 8), (128, 4), (64, 8), (64, 4), (32, 4), (32, 8)]})

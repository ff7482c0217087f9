"""clock64 stamps of CTA (0,0,0) of one tc_wgrad launch: per pixel-tile iteration the producer's issue window and the
MMA warp's wait / issue windows, plus prologue / epilogue.  usage: python tools/wgrad_trace.py"""
import os; os.environ.setdefault("DDL_LAB", "1")   # needs the instrumented library (ops/build.py --lab)
import os, sys, ctypes as C
os.environ["DDL_TC_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl_b200.ops import native
L = native.lib()
L.ddl_tc_plan_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
dev = torch.device("cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def run(label, B, H, ca, cb, ntile, psplit):
    g = torch.randn(B, H, H, ca, device=dev); x = torch.randn(B, H, H, cb, device=dev)
    out = torch.zeros(25, cb, ca, device=dev)
    plan = L.ddl_tc_plan_wgrad(g.data_ptr(), x.data_ptr(), out.data_ptr(), B, H, H, ca, cb, 5, ntile, psplit)
    assert plan, native.last_error()
    info = (C.c_int * 8)(); L.ddl_tc_plan_info(plan, info)
    for _ in range(2):
        flush.fill_(1); out.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); L.ddl_tc_plan_run(plan, native.stream_ptr()); b.record(); torch.cuda.synchronize()
    buf = (C.c_uint64 * 600)()
    L.ddl_tc_plan_trace(plan, buf, 600)
    t0 = buf[520]
    print("%s grid=%s stages=%d box=%s event %.1f us" % (label, tuple(info[0:3]), info[3], tuple(info[5:8]), a.elapsed_time(b) * 1e3))
    print("  epilogue start +%d, epilogue end +%d, kernel end +%d cycles" % (buf[521] - t0, buf[522] - t0, buf[523] - t0))
    i = 0
    while i < 64 and buf[i * 8 + 4]:
        r = [buf[i * 8 + k] - t0 for k in range(5)]
        print("  iter %2d: producer issue %6d..%6d (%5d) | mma wait %6d..%6d (%5d) issue ..%6d (%5d)" % (
            i, r[0], r[1], r[1] - r[0], r[2], r[3], r[3] - r[2], r[4], r[4] - r[3]))
        i += 1
    L.ddl_tc_plan_destroy(plan)


run("conv2 wgrad", 100, 14, 64, 32, 32, 0)
run("conv3 wgrad", 100, 7, 128, 64, 64, 0)

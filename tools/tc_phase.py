"""Phase timeline (globaltimer, ns) of CTA (0,0,0) of one act-GEMM launch.  usage: python tools/tc_phase.py"""
import os; os.environ.setdefault("DDL_LAB", "1")   # needs the instrumented library (ops/build.py --lab)
import os, sys, ctypes as C
os.environ["DDL_TC_TRACE"] = "1"; os.environ["DDL_TC_DEBUG"] = str(16 | int(os.environ.get("EXTRA_DEBUG", "0")))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl_b200.ops import native
L = native.lib()
L.ddl_tc_plan_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
dev = torch.device("cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
names = ["start", "prologue done", "pdl_wait done", "MMA: first stage landed", "MMA: last commit issued",
         "pre cluster barrier", "post cluster barrier", "epilogue: accumulator ready", "epilogue done", "teardown done",
         "chunk0 tmem loaded", "chunk0 math done", "chunk0 stores issued"]


def run(label, dgrad, N, H, cred, nout, ks, ntile, ksplit, flags):
    x = torch.randn(N, H, H, cred, device=dev); w = torch.randn(ks * ks * cred * nout, device=dev) * .05
    out = torch.zeros(N, H, H, nout, device=dev); bias = torch.zeros(nout, device=dev)
    plan = L.ddl_tc_plan_act(dgrad, x.data_ptr(), w.data_ptr(), out.data_ptr(), N, H, H, cred, nout, ks, ntile, ksplit, flags,
                             bias.data_ptr(), None, 1.0, None, 0, 1.0, None, None)
    assert plan, native.last_error()
    for cold in (True, False):
        for _ in range(2):
            if cold:
                flush.fill_(1)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); L.ddl_tc_plan_run(plan, native.stream_ptr()); b.record()
            torch.cuda.synchronize()
        buf = (C.c_uint64 * 16)()
        L.ddl_tc_plan_trace(plan, buf, 16)
        t0 = buf[0]
        print("%s [%s] event %.1f us: " % (label, "cold" if cold else "warm", a.elapsed_time(b) * 1e3) +
              " | ".join("%s +%.2f" % (names[i], (buf[i] - t0) / 1e3) for i in (1, 2, 3, 4, 6, 7, 10, 11, 12, 8, 9) if buf[i] >= t0 and buf[i] != 0))
    L.ddl_tc_plan_destroy(plan)


run("fc1 fwd (1,32,4) cluster", 0, 100, 1, 1024, 1024, 1, 32, 4, 3)
run("fc1 fwd (1,32,1)        ", 0, 100, 1, 1024, 1024, 1, 32, 1, 3)
run("conv2 fprop (168,1,1)   ", 0, 100, 14, 32, 64, 5, 64, 1, 0)
run("conv4 fprop (13,1,11)   ", 0, 100, 4, 128, 256, 5, 256, 11, 0)

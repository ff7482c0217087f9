"""Times tc_wgrad plans (cold L2 between launches) over pixel splits and taps per CTA.  usage: python tools/wgrad_perf.py"""
import os, sys, ctypes as C, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl_b200.ops import native

dev = torch.device("cuda")
L = native.lib()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def time_plan(plan, out, reps=8):
    s = native.stream_ptr()
    ts = []
    for i in range(reps + 2):
        flush.fill_(i)
        out.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); L.ddl_tc_plan_run(plan, s); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts = sorted(ts[2:])
    return ts[len(ts) // 2]


def wgrad(B, H, ca, cb, ntile, psplit):
    g = torch.randn(B, H, H, ca, device=dev); x = torch.randn(B, H, H, cb, device=dev)
    out = torch.zeros(25, cb, ca, device=dev)
    plan = L.ddl_tc_plan_wgrad(g.data_ptr(), x.data_ptr(), out.data_ptr(), B, H, H, ca, cb, 5, ntile, psplit)
    assert plan, native.last_error()
    info = (C.c_int * 8)(); L.ddl_tc_plan_info(plan, info)
    t = time_plan(plan, out)
    print("wgrad H=%d ca=%d cb=%d ntile=%d psplit=%d TPC=%s grid=%s stages=%d box=%s : %6.1f us" % (
        H, ca, cb, ntile, psplit, os.environ.get("DDL_WGRAD_TPC", "auto"), tuple(info[0:3]), info[3], tuple(info[5:8]), t), flush=True)
    L.ddl_tc_plan_destroy(plan)


if __name__ == "__main__":
    which = os.environ.get("LAYER", "")
    if which == "":
        for tpc in ("", "7", "5", "3"):
            env = dict(os.environ, LAYER="all")
            if tpc:
                env["DDL_WGRAD_TPC"] = tpc
            subprocess.run([sys.executable, __file__], env=env)
    else:
        for ps in (0, 15, 30, 60, 74):
            wgrad(100, 14, 64, 32, 32, ps)
        for ps in (0, 5, 10, 20, 37):
            wgrad(100, 7, 128, 64, 64, ps)
        for ps in (0, 2, 4, 8, 10):
            wgrad(100, 4, 256, 128, 128, ps)

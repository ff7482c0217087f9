"""dg2 (conv2 dgrad + un-pool) with and without the fused conv1 weight gradient, vs the separate conv1_wgrad kernel."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl_b200.ops import native
dev = torch.device("cuda"); L = native.lib()
B, H, UH, cin, cout = 100, 14, 28, 32, 64
dy = torch.randn(B, H, H, cout, device=dev); w = torch.randn(25 * cin * cout, device=dev) * .05
below = torch.randn(B, UH, UH, cin, device=dev); x1 = torch.randn(B, UH, UH, device=dev)
idx = torch.empty(B, H, H, cin, device=dev, dtype=torch.uint8); tmp = torch.empty(B, H, H, cin, device=dev)
s = native.stream_ptr()
L.ddl_maxpool_fwd(below.data_ptr(), None, tmp.data_ptr(), idx.data_ptr(), B, UH, UH, cin, s)
dfull = torch.empty(B, UH, UH, cin, device=dev); db = torch.zeros(cin, device=dev); dw1 = torch.zeros(25, cin, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(label, fn, reps=10):
    ts = []
    for i in range(reps + 2):
        flush.fill_(i)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts = sorted(ts[2:])
    print("%-52s %6.1f us (cold, median)" % (label, ts[len(ts) // 2]), flush=True)


def plan(w1, store):
    pl = L.ddl_tc_plan_dgrad_unpool(dy.data_ptr(), w.data_ptr(), idx.data_ptr(), dfull.data_ptr(), db.data_ptr(), B, H, H, UH, UH, cout, cin, 5, 32, 1)
    assert pl, native.last_error()
    if w1:
        assert L.ddl_tc_plan_unpool_w1grad(pl, x1.data_ptr(), dw1.data_ptr(), 5, store) == 0, native.last_error()
    return pl


p0, p1, p2 = plan(False, 1), plan(True, 1), plan(True, 0)
timeit("dgrad+unpool", lambda: L.ddl_tc_plan_run(p0, s))
timeit("dgrad+unpool+w1grad (dfull stored)", lambda: L.ddl_tc_plan_run(p1, s))
timeit("dgrad+unpool+w1grad (dfull dropped)", lambda: L.ddl_tc_plan_run(p2, s))
timeit("conv1_wgrad alone", lambda: L.ddl_conv1_wgrad(x1.data_ptr(), dfull.data_ptr(), dw1.data_ptr(), db.data_ptr(), B, UH, cin, s))
timeit("dgrad+unpool then conv1_wgrad", lambda: (L.ddl_tc_plan_run(p0, s), L.ddl_conv1_wgrad(x1.data_ptr(), dfull.data_ptr(), dw1.data_ptr(), db.data_ptr(), B, UH, cin, s)))

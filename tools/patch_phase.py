"""globaltimer phase stamps of CTA 0 of the patch-resident conv kernel (lab build).  usage: python tools/patch_phase.py"""
import os
os.environ.setdefault("DDL_LAB", "1")
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl_b200.ops import native, functional as K
L = native.lib()
dev = torch.device("cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
names = ["start", "prologue done", "patch landed", "first weights landed", "last MMA issued", "accumulators complete", "staged",
         "outputs written", "tap 12 weights landed"]
B = 100
for (H, Cin, Cout) in [(14, 32, 64), (7, 64, 128)]:
    x = torch.randn(B, H, H, Cin, device=dev); w = torch.randn(5, 5, Cin, Cout, device=dev) * .05; b = torch.randn(Cout, device=dev) * .1
    PH = (H + 1) // 2
    y = torch.empty(B, PH, PH, Cout, device=dev); idx = torch.empty(B, PH, PH, Cout, device=dev, dtype=torch.uint8)
    dy = torch.randn(B, H, H, Cout, device=dev)
    _, pidx = K.bias_relu_maxpool(torch.randn(B, 2 * H, 2 * H, Cin, device=dev))
    g = torch.empty(B, H, H, Cin, device=dev); db = torch.zeros(Cin, device=dev)
    tr = torch.zeros(16, dtype=torch.int64, device=dev)
    for label, pl in (("fwd+pool", L.ddl_patch_plan_create(0, x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), idx.data_ptr(), None, B, H, H, Cin, Cout)),
                      ("dgrad+mask", L.ddl_patch_plan_create(1, dy.data_ptr(), w.data_ptr(), None, g.data_ptr(), pidx.data_ptr(), db.data_ptr(), B, H, H, Cout, Cin))):
        assert pl, native.last_error()
        L.ddl_patch_plan_set_trace(pl, tr.data_ptr())
        for cold in (True, False):
            for _ in range(3):
                if cold:
                    flush.fill_(1)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); L.ddl_patch_plan_run(pl, native.stream_ptr()); e1.record()
                torch.cuda.synchronize()
            v = tr.cpu().tolist()
            print("%dx%d %d->%d %s [%s] event %.1f us: " % (H, H, Cin, Cout, label, "cold" if cold else "warm", e0.elapsed_time(e1) * 1e3) +
                  " | ".join("%s +%.2f" % (names[i], (v[i] - v[0]) / 1e3) for i in (1, 2, 3, 8, 4, 5, 6, 7)) + " | MMA warp waited for weights %.2f | MMA loop %d cycles = %.0f MHz" % (v[9] / 1e3, v[11] - v[10], (v[11] - v[10]) / max(1e-9, (v[4] - v[3]) / 1e3)), flush=True)
        L.ddl_patch_plan_destroy(pl)

"""Single-kernel timings (CUDA events, median of N) for the kernels changed in round 2.
usage: python tools/kbench.py [case ...]      cases: w1pooled conv1 sync fcblock
Each case prints cold (L2 flushed before the launch) and warm (inputs resident, as inside a training step) times."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl_b200.ops import native, functional as K

dev = torch.device("cuda")
L = native.lib()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(name, fn, reps=15):
    for cold in (True, False):
        for _ in range(3):
            fn()
        ts = []
        for _ in range(reps):
            if cold:
                flush.fill_(1)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b) * 1e3)
        ts.sort()
        print("%-52s %s: median %7.1f us  min %7.1f" % (name, "cold" if cold else "warm", ts[len(ts) // 2], ts[0]), flush=True)


def case_w1pooled():
    B = 100
    x = torch.rand(B, 28, 28, device=dev)
    below = torch.randn(B, 28, 28, 32, device=dev)
    _, idx = K.bias_relu_maxpool(below)
    gp = torch.randn(B, 14, 14, 32, device=dev) * (idx < 4)
    dy, _ = K.maxpool_relu_bwd(gp, idx, 28, 28)
    timeit("conv1_wgrad_pooled (gp + idx)", lambda: K.conv1_wgrad_pooled(x, gp, idx, want_db=False))
    timeit("conv1_wgrad dense (un-pooled dy)", lambda: K.conv1_wgrad(x, dy))
    timeit("conv1_wgrad_sparse (un-pooled dy + idx)", lambda: K.conv1_wgrad_sparse(x, dy, idx))


def case_conv1():
    B = 100
    x = torch.rand(B, 28, 28, device=dev); w1 = torch.randn(5, 5, 1, 32, device=dev) * .1; b1 = torch.randn(32, device=dev) * .1
    timeit("conv1_relu_pool", lambda: K.conv1_relu_pool(x, w1, b1))


def case_sync():
    from ddl_b200.config import Config
    from ddl_b200.parallel.trainer import SPMDTrainer
    cfg = Config(synthetic=True, synthetic_train=400, synthetic_test=100, seed=0, eval_every=0, log_time=False, graphs=False)
    tr = SPMDTrainer(cfg, 0, 1, 0)
    t = tr.transport
    timeit("sync push/update/pull, all buckets (world=1)", lambda: t.step(all_buckets=True))


def case_patch():
    """patch-resident conv (tc_patch.cu) against the tap-by-tap plans (tc_gemm.cu), same layer, plans built once"""
    import ctypes as C
    B = 100
    for (H, Cin, Cout, nt_f, cl_f, nt_d, cl_d) in [(14, 32, 64, 64, 1, 32, 1), (7, 64, 128, 128, 4, 64, 4)]:
        x = torch.randn(B, H, H, Cin, device=dev); w = torch.randn(5, 5, Cin, Cout, device=dev) * .05; b = torch.randn(Cout, device=dev) * .1
        PH = (H + 1) // 2
        y = torch.empty(B, PH, PH, Cout, device=dev); idx = torch.empty(B, PH, PH, Cout, device=dev, dtype=torch.uint8)
        dy = torch.randn(B, H, H, Cout, device=dev)
        _, pidx = K.bias_relu_maxpool(torch.randn(B, 2 * H, 2 * H, Cin, device=dev))
        g = torch.empty(B, H, H, Cin, device=dev); db = torch.zeros(Cin, device=dev)
        plans = {
            "tap-by-tap fwd+pool": (L.ddl_tc_plan_conv_pool(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), idx.data_ptr(), B, H, H, Cin, Cout, 5, nt_f, cl_f), L.ddl_tc_plan_run),
            "patch       fwd+pool": (L.ddl_patch_plan_create(0, x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), idx.data_ptr(), None, B, H, H, Cin, Cout), L.ddl_patch_plan_run),
            "tap-by-tap dgrad+mask": (L.ddl_tc_plan_dgrad_poolgrad(dy.data_ptr(), w.data_ptr(), pidx.data_ptr(), g.data_ptr(), db.data_ptr(), B, H, H, Cout, Cin, 5, nt_d, cl_d), L.ddl_tc_plan_run),
            "patch       dgrad+mask": (L.ddl_patch_plan_create(1, dy.data_ptr(), w.data_ptr(), None, g.data_ptr(), pidx.data_ptr(), db.data_ptr(), B, H, H, Cout, Cin), L.ddl_patch_plan_run),
        }
        for name, (pl, run) in plans.items():
            assert pl, native.last_error()
            timeit("%dx%d %d->%d %s" % (H, H, Cin, Cout, name), lambda: run(pl, native.stream_ptr()))


CASES = {"patch": case_patch, "w1pooled": case_w1pooled, "conv1": case_conv1, "sync": case_sync}

if __name__ == "__main__":
    for name in (sys.argv[1:] or list(CASES)):
        CASES[name]()

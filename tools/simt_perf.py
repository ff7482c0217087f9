"""Times the SIMT / comm kernels in isolation (CUDA events, L2 flushed before every launch)."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddl_b200.ops import native, functional as K
from ddl_b200.config import Config
from ddl_b200.parallel.trainer import SPMDTrainer

dev = torch.device("cuda")
L = native.lib()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(name, fn, reps=10, cold=True):
    for _ in range(2):
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
    print("%-46s %s: median %7.1f us  min %7.1f" % (name, "cold" if cold else "warm", ts[len(ts) // 2], ts[0]), flush=True)


def main():
    s = lambda: native.stream_ptr()
    n = 2_656_064
    p, m, v, g = (torch.randn(n, device=dev) for _ in range(4))
    v.abs_()
    for cold in (True, False):
        timeit("adam optimizer_step 2.66M", lambda: L.ddl_optimizer_step(2, p.data_ptr(), m.data_ptr(), v.data_ptr(), g.data_ptr(), n,
                                                                        1e-4, 0.9, 0.999, 1e-8, 1.0, None, s()), cold=cold)
        timeit("torch copy 2.66M x7 passes (roofline-ish)", lambda: (p.add_(g), m.add_(v)), cold=cold)
    # the fused sync kernel with zero peers, through the trainer's transport
    cfg = Config(synthetic=True, synthetic_train=400, synthetic_test=100, seed=0, eval_every=0, log_time=False, graphs=False)
    tr = SPMDTrainer(cfg, 0, 1, 0)
    t = tr.transport
    for cold in (True, False):
        timeit("sync_push_update_pull (world=1) + step_advance", lambda: (t.begin_step(), t.step()), cold=cold)
        timeit("step_advance only", lambda: t.begin_step(), cold=cold)
    B = 100
    x = torch.rand(B, 28, 28, device=dev); w1 = torch.randn(5, 5, 1, 32, device=dev) * .1; b1 = torch.randn(32, device=dev) * .1
    dy1 = torch.randn(B, 28, 28, 32, device=dev)
    for cold in (True, False):
        timeit("conv1_relu_pool", lambda: K.conv1_relu_pool(x, w1, b1), cold=cold)
        timeit("conv1_wgrad", lambda: K.conv1_wgrad(x, dy1), cold=cold)
    a2 = torch.randn(B, 14, 14, 64, device=dev); bias2 = torch.randn(64, device=dev)
    y2, i2 = K.bias_relu_maxpool(a2, bias2)
    dp2 = torch.randn_like(y2)
    for cold in (True, False):
        timeit("bias_relu_maxpool 14x14x64", lambda: K.bias_relu_maxpool(a2, bias2), cold=cold)
        timeit("maxpool_relu_bwd 14x14x64", lambda: K.maxpool_relu_bwd(dp2, i2, 14, 14), cold=cold)
    h = torch.randn(B, 512, device=dev); w3 = torch.randn(512, 10, device=dev) * .1; b3 = torch.randn(10, device=dev)
    yl = torch.randint(0, 10, (B,), device=dev, dtype=torch.int32)
    lg, dl, _, _ = K.fc3_softmax_xent(h, w3, b3, yl)
    for cold in (True, False):
        timeit("fc3_softmax_xent (incl. 2 torch allocs + D2H)", lambda: K.fc3_softmax_xent(h, w3, b3, yl), cold=cold, reps=5)
        timeit("fc3_bwd", lambda: K.fc3_bwd(h, w3, dl), cold=cold)
    timeit("empty torch kernel (fill_ of 1 float)", lambda: p[:1].fill_(0), cold=False)


if __name__ == "__main__":
    main()

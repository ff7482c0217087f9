"""GPU diagnostic for the tcgen05 kernels: prints error maps so a layout / descriptor bug can be
read off one run.  usage: python tools/tc_debug.py"""
import os; os.environ.setdefault("DDL_LAB", "1")   # needs the instrumented library (ops/build.py --lab)
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from ddl_b200.ops import functional as K

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
dev = torch.device("cuda")


def report(name, got, ref):
    err = (got - ref).abs()
    scale = ref.abs().max().item() + 1e-12
    bad = err > 5e-3 * scale
    print("%-34s rel_err %.3e  bad %.4f  nan %d" % (name, err.max().item() / scale, bad.float().mean().item(),
                                                   int(torch.isnan(got).sum())), flush=True)
    if bad.any():
        g2, r2 = got.reshape(-1, got.shape[-1]), ref.reshape(-1, ref.shape[-1])
        b2 = bad.reshape(-1, got.shape[-1])
        rows = b2.any(1).nonzero().flatten()
        cols = b2.any(0).nonzero().flatten()
        print("   bad rows: n=%d first %s | bad cols: n=%d first %s" % (len(rows), rows[:12].tolist(), len(cols), cols[:12].tolist()))
        r = int(rows[0])
        print("   row %d got %s" % (r, [round(v, 3) for v in g2[r, :8].tolist()]))
        print("   row %d ref %s" % (r, [round(v, 3) for v in r2[r, :8].tolist()]))
        # is the result a permutation / scaled version?
        print("   ratio got/ref (median over good denominators): %.4f" % (g2[r2.abs() > 0.1 * scale] / r2[r2.abs() > 0.1 * scale]).median().item())


def main():
    torch.manual_seed(0)
    # 1. plain GEMM, K-major A, MN-major B (fwd)
    for (B, Kin, N) in [(128, 32, 32), (128, 64, 64), (100, 1024, 512)]:
        x = torch.randn(B, Kin, device=dev); w = torch.randn(Kin, N, device=dev) * 0.1
        try:
            report("linear_fwd %dx%dx%d" % (B, Kin, N), K.linear_fwd(x, w), (x.double() @ w.double()).float())
        except Exception as e:
            print("linear_fwd FAILED", e); return 1
        dy = torch.randn(B, N, device=dev)
        try:
            report("linear_dx  %dx%dx%d" % (B, Kin, N), K.linear_dx(dy, w), (dy.double() @ w.double().t()).float())
        except Exception as e:
            print("linear_dx FAILED", e); return 1
        try:
            report("linear_dw  %dx%dx%d" % (B, Kin, N), K.linear_dw(dy, x), (x.double().t() @ dy.double()).float())
        except Exception as e:
            print("linear_dw FAILED", e); return 1
    # 2. convs
    for (H, Cin, Cout) in [(4, 32, 32), (14, 32, 64), (7, 64, 128), (4, 128, 256)]:
        x = torch.randn(100, H, H, Cin, device=dev); w = torch.randn(5, 5, Cin, Cout, device=dev) * 0.05
        xd = x.double().permute(0, 3, 1, 2).requires_grad_(True)
        wd = w.double().permute(3, 2, 0, 1).requires_grad_(True)
        y = F.conv2d(xd, wd, padding=2)
        dy = torch.randn_like(y)
        y.backward(dy)
        report("conv fprop %d %d->%d" % (H, Cin, Cout), K.conv5x5_fprop(x, w), y.detach().permute(0, 2, 3, 1).float())
        dyf = dy.permute(0, 2, 3, 1).float().contiguous()
        report("conv dgrad %d %d->%d" % (H, Cin, Cout), K.conv5x5_dgrad(dyf, w), xd.grad.permute(0, 2, 3, 1).float())
        report("conv wgrad %d %d->%d" % (H, Cin, Cout), K.conv5x5_wgrad(dyf, x), wd.grad.permute(2, 3, 1, 0).float())
    return 0


if __name__ == "__main__":
    sys.exit(main())

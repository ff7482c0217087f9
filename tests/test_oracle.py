"""Model math contract (mnist_sync/model/model.py) restated on CPU."""
import math

import torch

from ddl_b200.models import oracle
from ddl_b200.models.spec import REFERENCE_SPEC


def test_shapes_and_same_pool_edge():
    p = oracle.init_params(REFERENCE_SPEC, 0)
    x = torch.rand(3, 784)
    logits, acts = oracle.forward(p, x, return_acts=True)
    assert logits.shape == (3, 10)
    assert acts["pool2"].shape == (3, 64, 7, 7)
    assert acts["pool3"].shape == (3, 128, 4, 4)          # 7 -> 4: SAME pool pads bottom/right
    assert acts["pool4"].shape == (3, 256, 2, 2)
    # the padded window of the 7->4 pool only sees real pixels
    a = acts["conv3"]
    assert torch.equal(acts["pool3"][:, :, 3, 3], a[:, :, 6, 6])


def test_glorot_limits_include_biases():
    p = oracle.init_params(REFERENCE_SPEC, 1)
    assert p[0].abs().max() <= math.sqrt(6 / (25 * 1 + 25 * 32)) + 1e-7
    assert p[1].abs().max() <= math.sqrt(3 / 32) + 1e-7 and p[1].abs().max() > 0
    assert p[8].abs().max() <= math.sqrt(6 / 2048) + 1e-7


def test_hwio_and_flatten_order_against_manual_conv():
    torch.manual_seed(0)
    x = torch.rand(2, 28, 28, 1)
    w = torch.randn(5, 5, 1, 4)
    b = torch.randn(4)
    out = oracle.conv5x5_same(x.permute(0, 3, 1, 2), w, b).permute(0, 2, 3, 1)
    xp = torch.nn.functional.pad(x, (0, 0, 2, 2, 2, 2))
    n, h, wd, o = 1, 5, 27, 3
    ref = sum(xp[n, h + kh, wd + kw, 0] * w[kh, kw, 0, o] for kh in range(5) for kw in range(5)) + b[o]
    assert torch.allclose(out[n, h, wd, o], ref, atol=1e-5)


def test_tf1_adam_formula():
    p = torch.tensor([1.0, -2.0])
    g = torch.tensor([0.5, 0.25])
    opt = oracle.OracleOptimizer(p, "adam", lr=1e-2)
    q = p.clone()
    opt.step(q, g)
    m = 0.1 * g
    v = 0.001 * g * g
    lr_t = 1e-2 * math.sqrt(1 - 0.999) / (1 - 0.9)
    assert torch.allclose(q, p - lr_t * m / (v.sqrt() + 1e-8), atol=1e-7)
    # differs from torch.optim.Adam's epsilon placement only at O(eps)
    tp = torch.nn.Parameter(p.clone())
    to = torch.optim.Adam([tp], lr=1e-2)
    tp.grad = g.clone()
    to.step()
    assert torch.allclose(q, tp.detach(), atol=1e-5)


def test_dropout_mask_rate_and_determinism():
    a = oracle.dropout_keep_mask(7, 3, 100000, 0.5)
    b = oracle.dropout_keep_mask(7, 3, 100000, 0.5)
    c = oracle.dropout_keep_mask(7, 4, 100000, 0.5)
    assert torch.equal(a, b) and not torch.equal(a, c)
    assert abs(a.float().mean().item() - 0.5) < 0.01
    assert oracle.dropout_keep_mask(1, 1, 1000, 1.0).all()


def test_loss_decreases_with_oracle_training():
    torch.manual_seed(0)
    from ddl_b200.data import synthetic_mnist
    d = synthetic_mnist(400, 100)
    p = oracle.init_params(REFERENCE_SPEC, 0)
    flat = torch.cat([q.reshape(-1) for q in p])
    ps = oracle.SequentialPS(flat, "adam", 1e-3)
    sizes = [q.numel() for q in p]

    def unflat(f):
        out, o = [], 0
        for q, n in zip(p, sizes):
            out.append(f[o:o + n].view(q.shape)); o += n
        return out
    losses = []
    for s in range(8):
        x, y = d.x_train[s * 50:(s + 1) * 50], d.y_train[s * 50:(s + 1) * 50]
        loss, grads, _ = oracle.loss_and_grads(unflat(ps.p), x, y)
        losses.append(float(loss))
        ps.sync_step([torch.cat([g.reshape(-1) for g in grads])])
    assert losses[-1] < losses[0]

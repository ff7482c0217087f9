"""``torch_graph`` baseline rung: the strongest library-only version of the same training step.

cuDNN convolutions + cuBLAS matmuls with TF32 allowed (the same multiply precision as the product's tensor-core path),
torch autograd for the backward, the PS exchange as NCCL ``reduce_scatter`` (push + sum) -> TF1-Adam on the owner's
shard -> NCCL ``all_gather`` (pull), and the WHOLE step captured in one CUDA graph, so it is not launch-bound the way an
eager PyTorch step is (round 1's ``--impl torch`` rung ran un-graphed: 2.1 ms per step).  Same model (the 14-variable
CNN of mnist_sync/model/model.py:17-112), same batch 100 per worker, same synthetic data shape, same timing protocol as
``bench.py`` (CUDA events per step, 256 MiB written between steps, max over ranks).  None of this repository's kernels
run on this path.

    python bench.py --impl torch_graph --gpus N --steps K --warmup W
"""
from __future__ import annotations

import json
import math
import os

import torch
import torch.distributed as dist
import torch.nn.functional as F

SHAPES = [(5, 5, 1, 32), (32,), (5, 5, 32, 64), (64,), (5, 5, 64, 128), (128,), (5, 5, 128, 256), (256,),
          (1024, 1024), (1024,), (1024, 512), (512,), (512, 10), (10,)]


def widen(shapes, wm):
    if wm == 1:
        return shapes
    c = [1, 32 * wm, 64 * wm, 128 * wm, 256 * wm]
    f = [4 * 256 * wm, 1024 * wm, 512 * wm, 10]
    out = []
    for l in range(4):
        out += [(5, 5, c[l], c[l + 1]), (c[l + 1],)]
    for l in range(3):
        out += [(f[l], f[l + 1]), (f[l + 1],)]
    return out


def forward(ps, x, keep_prob):
    """model.py:18-88 with torch ops; weights are stored HWIO / [in, out] like the reference's variables."""
    h = x.view(-1, 1, 28, 28)
    for l in range(4):
        w, b = ps[2 * l], ps[2 * l + 1]
        h = F.relu(F.conv2d(h, w.permute(3, 2, 0, 1), b, padding=2))
        h = F.max_pool2d(h, 2, 2, ceil_mode=True)
    h = h.permute(0, 2, 3, 1).reshape(h.shape[0], -1)                 # (h, w, c) flatten order, model.py:69
    h = F.dropout(F.relu(h @ ps[8] + ps[9]), 1.0 - keep_prob)
    h = F.dropout(h @ ps[10] + ps[11], 1.0 - keep_prob)               # fc2 has no activation (model.py:79)
    return h @ ps[12] + ps[13]


def run(a, emit=lambda o: print(json.dumps(o))) -> int:
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    torch.backends.cudnn.allow_tf32 = True
    torch.backends.cuda.matmul.allow_tf32 = True
    torch.backends.cudnn.benchmark = True
    B, keep, lr, b1, b2, eps = a.batch_size, 0.5, 1e-4, 0.9, 0.999, 1e-8
    shapes = widen(SHAPES, a.width_mult)
    sizes = [math.prod(s) for s in shapes]
    total = (sum(sizes) + 4 * world - 1) // (4 * world) * (4 * world)  # flat arena, equal 16-byte aligned shards
    shard = total // world
    g = torch.Generator(device="cpu").manual_seed(1234)
    flat = torch.zeros(total, device=dev)
    grads = torch.zeros(total, device=dev)
    off, views, gviews = 0, [], []
    for s, n in zip(shapes, sizes):
        fan = (s[0] * s[1] * s[2], s[0] * s[1] * s[3]) if len(s) == 4 else (s[0], s[1]) if len(s) == 2 else (s[0], s[0])
        lim = math.sqrt(6.0 / (fan[0] + fan[1]))
        flat[off:off + n] = ((torch.rand(n, generator=g) * 2 - 1) * lim).to(dev)
        views.append(flat[off:off + n].view(s).requires_grad_(True))
        gviews.append(grads[off:off + n].view(s))
        off += n
    m = torch.zeros(shard, device=dev)
    v = torch.zeros(shard, device=dev)
    t_dev = torch.zeros((), device=dev)
    my = slice(rank * shard, (rank + 1) * shard)
    nbatch = 64
    xs = torch.rand(nbatch, B, 784, generator=g).to(dev)
    ys = torch.randint(0, 10, (nbatch, B), generator=g).to(dev)
    x_in = torch.zeros(B, 784, device=dev)
    y_in = torch.zeros(B, dtype=torch.long, device=dev)
    loss_out = torch.zeros((), device=dev)

    def compute():
        logits = forward(views, x_in, keep)
        loss = F.cross_entropy(logits, y_in)
        gs = torch.autograd.grad(loss, views)
        for gv, gg in zip(gviews, gs):
            gv.copy_(gg)
        loss_out.copy_(loss.detach())

    def update():
        gsh = grads[my]
        t_dev.add_(1.0)                                                          # TF1 Adam: eps outside the bias correction
        lr_t = lr * torch.sqrt(1.0 - b2 ** t_dev) / (1.0 - b1 ** t_dev)
        m.mul_(b1).add_(gsh, alpha=1 - b1)
        v.mul_(b2).addcmul_(gsh, gsh, value=1 - b2)
        with torch.no_grad():
            flat[my].sub_(lr_t * m / (v.sqrt() + eps))

    def push():                                                                  # push + sum: the owner gets its shard
        dist.reduce_scatter_tensor(grads[my], grads, op=dist.ReduceOp.SUM)

    def pull():                                                                  # pull: every worker gets every shard
        dist.all_gather_into_tensor(flat, flat[my].clone())

    # warm-up on a side stream (cuDNN autotune, NCCL channels), then capture.  One rank: the whole step is ONE graph.
    # Several ranks: forward + backward is one graph, the optimizer another, and the two NCCL collectives are issued
    # between them (4 launches per step; capturing the collectives too hung this image's NCCL 2.28 / torch 2.11 pair).
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for i in range(max(3, a.warmup)):
            x_in.copy_(xs[i % nbatch]); y_in.copy_(ys[i % nbatch])
            compute()
            if world > 1:
                push()
            update()
            if world > 1:
                pull()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g_compute, g_update = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
    if world == 1:
        with torch.cuda.graph(g_compute):
            compute()
            update()
    else:
        with torch.cuda.graph(g_compute):
            compute()
        with torch.cuda.graph(g_update):
            update()

    class _Step:
        @staticmethod
        def replay():
            g_compute.replay()
            if world > 1:
                push()
                g_update.replay()
                pull()
    graph = _Step

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
    first = None
    sync_all()
    for i in range(a.steps):
        flush.fill_(i & 0xFF)
        x_in.copy_(xs[i % nbatch]); y_in.copy_(ys[i % nbatch])
        evs[i][0].record()
        graph.replay()
        evs[i][1].record()
        if i == 0:
            first = float(loss_out)
    sync_all()
    t = torch.tensor([sum(s.elapsed_time(e) for s, e in evs)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    if rank == 0:
        emit({"metric": "samples/sec", "value": a.steps * B * world / (ms / 1e3), "unit": "samples/s", "n_gpus": world,
              "steps": a.steps, "warmup": max(3, a.warmup), "ms_per_step": ms / a.steps, "higher_is_better": True,
              "scaling": "weak", "vs_baseline": None, "impl": "torch_graph",
              "dtype": "tf32 allowed (cuDNN / cuBLAS), fp32 storage", "data": "synthetic MNIST-shaped, random-init weights",
              "config": {"model": "mnist_cnn_14var" + ("" if a.width_mult == 1 else "_x%d" % a.width_mult),
                         "params": sum(sizes), "global_batch": B * world, "batch_per_worker": B, "optimizer": "adam (TF1 formula)",
                         "parallelism": "flat shards, NCCL reduce_scatter + all_gather; " + ("whole step in one CUDA graph" if world == 1 else "two CUDA graphs (fwd+bwd, optimizer) around the two collectives"),
                         "cuda_graph": True, "l2": "256 MiB written between timed steps"},
              "first_loss": first, "final_loss": float(loss_out)})
    if world > 1:
        dist.destroy_process_group()
    return 0

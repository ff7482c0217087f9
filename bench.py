#!/usr/bin/env python
"""Headline benchmark: training samples/s of the reference MNIST CNN under parameter-server training.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched by torch.distributed.run)

Metric / config are BASELINE.json's: samples/sec (device-timed, max over ranks), sync PS with greedy
size-balanced sharding (``mnist_sync_sharding_greedy``): N GPU-resident PS shards co-located with N
workers, batch 100 per worker (weak scaling), TF1-Adam(1e-4), dropout 0.5, synthetic MNIST-shaped
data, random-init weights.  Storage is fp32 everywhere (the reference's precision); GEMM-shaped ops
run on the tcgen05 tensor cores as kind::tf32 with fp32 accumulation.

Numbers printed (one JSON line, rank 0):
  value      - K steps, each timed with CUDA events on the launching stream (max over ranks of the
               summed step times); a 256 MiB buffer is written between timed steps to flush L2.
               A step = dropout-stream/Adam-scalar update, forward, backward, gradient push,
               reduction, optimizer, parameter pull - nothing skipped.
  e2e        - the same metric through the public API (``SPMDTrainer.train_step``): every step copies
               its batch from pinned host memory (H2D) and the host reads the step's loss (D2H).
  --impl reference  - the unmodified reference cannot be installed here (no setup.py; needs
               TensorFlow 1.x + mpi4py + mpiexec, none present) -> {"unavailable": ...}.
  --impl host_ps / nccl / torch - the stand-in baseline rungs of BASELINE.md section 2.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

REFERENCE_UNAVAILABLE = ("reference is not installable: /root/reference has no setup.py/pyproject.toml and its scripts "
                         "need TensorFlow 1.x graph mode, mpi4py and mpiexec (Python <= 3.7); none exist in this image")


class ClockSampler:
    """Samples SM clocks / throttle reasons with nvidia-smi while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")] + [time.time()])

    def stop(self, t_begin: float = 0.0, t_end: float = 1e18):
        """Median SM clock / reasons over the samples that arrived inside [t_begin, t_end] (the timed regions);
        nvidia-smi is started long before, because its own start-up takes longer than a short timed region."""
        if self.proc:
            time.sleep(0.05)
            self.proc.terminate()
        inside = [r for r in self.rows if len(r) >= 9 and t_begin <= r[-1] <= t_end + 0.05]
        self.rows = inside if inside else self.rows[-3:]
        sm = sorted(int(float(r[1])) for r in self.rows if len(r) >= 8 and r[1].replace(".", "").isdigit())
        mx = [int(float(r[2])) for r in self.rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 8:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "host_ps", "nccl", "torch"])
    ap.add_argument("--variant", default="mnist_sync_sharding_greedy")
    ap.add_argument("--shard", default=None)
    ap.add_argument("--width-mult", type=int, default=1)
    ap.add_argument("--batch-size", type=int, default=100)
    ap.add_argument("--optimizer", default="adam")
    ap.add_argument("--no-graphs", action="store_true")
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--nvls", default="auto", choices=["auto", "on", "off"])
    ap.add_argument("--no-overlap", action="store_true", help="disable the backward-overlapped (bucketed) push")
    ap.add_argument("--k4", action="store_true", help="enable the pull fused into the fc1/fc2 GEMMs (K4)")
    ap.add_argument("--no-k4", action="store_true", help=argparse.SUPPRESS)
    return ap.parse_args()


def main() -> int:
    a = parse()
    # stdout carries exactly ONE line (the JSON): libraries that print banners there (NCCL's version line,
    # torchrun children) are redirected to stderr for the duration of the run.
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    def emit(obj) -> None:
        sys.stdout.flush()
        os.write(real_stdout, (json.dumps(obj) + "\n").encode())

    if a.impl == "reference":
        emit({"impl": "reference", "unavailable": REFERENCE_UNAVAILABLE})
        return 0
    if a.impl == "host_ps":
        from baseline.host_ps.bench_host_ps import run as run_host
        return run_host(a, emit)

    import torch
    import torch.distributed as dist
    from ddl_b200.config import Config, VARIANTS
    from ddl_b200.ops import native
    from ddl_b200.parallel.trainer import SPMDTrainer

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    if world != a.gpus:
        if world == 1 and a.gpus > 1:
            # convenience: re-launch ourselves under torch.distributed.run
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(a.gpus),
                   "--master-addr", "127.0.0.1", "--master-port", "29531", os.path.abspath(__file__)] + sys.argv[1:]
            os.dup2(real_stdout, 1)           # the children do their own redirection
            return subprocess.call(cmd)
        raise SystemExit("WORLD_SIZE=%d does not match --gpus %d" % (world, a.gpus))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device")

    mode, shard = VARIANTS[a.variant]
    shard = a.shard or shard
    if shard == "none" and world > 1:
        num_ps = 1
    else:
        num_ps = world
    transport = {"ours": "fused", "nccl": "nccl", "torch": "nccl" if world > 1 else "local"}[a.impl]
    backend = "torch" if a.impl == "torch" else "native"
    cfg = Config(variant=a.variant, mode=mode, shard=shard, num_ps=num_ps, num_workers=world,
                 batch_size=a.batch_size, optimizer=a.optimizer, transport=transport, backend=backend,
                 synthetic=True, synthetic_train=max(6000, a.batch_size * 32), synthetic_test=1000, seed=1234,
                 eval_every=0, width_mult=a.width_mult, graphs=not a.no_graphs, log_time=False,
                 fused_pull_gemm=a.k4 and not a.no_k4, overlap_push=not a.no_overlap, nvls=a.nvls).validate()
    kind = "colocated" if world > 1 else "auto"
    if world > 1 and num_ps == 1:
        kind = "colocated"
    # nvidia-smi sampler (every 20 ms) is started before anything else: it needs up to a second to produce its
    # first row on an 8-GPU box, longer than a short timed region
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    tr = SPMDTrainer(cfg, rank, world, local, kind)
    dev = tr.device
    B = cfg.batch_size
    nb = tr.data.batches_per_epoch(B, world, False)

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---------------- warm-up (includes the CUDA-graph capture at step 2)
    for i in range(max(3, a.warmup)):
        tr.train_step(i % nb)
    sync_all()
    t_load_begin = time.time()              # clock samples from here to the end of the e2e loop are reported
    graph_mode = tr._graph is not None
    launches_per_step = getattr(tr, "launches_per_step", None)

    # ---------------- device-timed steps (CUDA events per step, L2 flushed between steps)
    flush = None if a.no_flush else torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
    sync_all()
    l0 = native.launches()
    for i in range(a.steps):
        if flush is not None:
            flush.fill_(i & 0xFF)
        evs[i][0].record()
        tr.train_step((a.warmup + i) % nb)
        evs[i][1].record()
    sync_all()
    dev_ms = sum(s.elapsed_time(e) for s, e in evs)
    t = torch.tensor([dev_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms = float(t.item())
    launches = (native.launches() - l0) if not graph_mode else (launches_per_step or 0) * a.steps

    # ---------------- end-to-end through the public API: H2D every step, loss read on the host every step
    host_loss = torch.zeros(2, dtype=torch.float32).pin_memory()
    losses = []
    sync_all()
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    done = [None, None]
    for i in range(a.steps):
        loss = tr.train_step((a.warmup + a.steps + i) % nb)         # H2D from pinned memory inside
        slot = i & 1
        if done[slot] is not None:                                    # read the loss of step i-2 (pipelined D2H)
            done[slot].synchronize()
            losses.append(float(host_loss[slot]))
        host_loss[slot:slot + 1].copy_(loss.reshape(1), non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        done[slot] = ev
    for slot in (a.steps & 1, (a.steps + 1) & 1):
        if done[slot] is not None:
            done[slot].synchronize()
            losses.append(float(host_loss[slot]))
    e1.record()
    sync_all()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s, e0.elapsed_time(e1) / 1e3], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t[0].item())
    clocks = sampler.stop(t_load_begin, time.time())
    if hasattr(tr.transport, "finish"):
        tr.transport.finish()                 # async: drain accumulators, final pull
    acc = tr.evaluate()
    stats = tr.transport.stats() if hasattr(tr.transport, "stats") else {}
    tr.close()

    if rank == 0:
        total = a.steps * B * world
        h2d = B * 784 * 4 + B * 4
        out = {
            "metric": "samples/sec", "value": total / (dev_ms / 1e3), "unit": "samples/s", "n_gpus": world,
            "steps": a.steps, "warmup": max(3, a.warmup), "ms_per_step": dev_ms / a.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "tf32 (fp32 storage, tf32 tensor-core multiply, fp32 accumulate)" if backend == "native" else "fp32",
            "data": "synthetic MNIST-shaped [N,784] fp32, 10 classes; random-init (glorot) weights",
            "impl": a.impl,
            "config": {"model": "mnist_cnn_14var" + ("" if a.width_mult == 1 else "_x%d" % a.width_mult),
                       "params": tr.spec.num_params, "variant": a.variant, "mode": mode, "shard": shard,
                       "global_batch": B * world, "batch_per_worker": B, "seq_len": None,
                       "parallelism": "ps%d+w%d colocated, %s transport" % (num_ps, world, tr.transport.name),
                       "optimizer": a.optimizer, "keep_prob": cfg.keep_prob,
                       "l2": "256 MiB written between timed steps" if flush is not None else "not flushed",
                       "cuda_graph": graph_mode, "backend": backend,
                       "k4_fused_pull_gemm": bool(getattr(tr.transport, "k4_src", None))},
            "clocks": clocks,
            "e2e": {"value": total / e2e_s, "unit": "samples/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                    "ms_per_step": e2e_s * 1e3 / a.steps},
            "gpu_launches": int(launches),
            "final_loss": losses[-1] if losses else None, "first_loss": losses[0] if losses else None,
            "test_accuracy_after_bench": acc, "transport_stats": stats,
        }
        emit(out)
    return 0


if __name__ == "__main__":
    sys.exit(main())

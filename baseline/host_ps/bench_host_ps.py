"""``host_ps`` baseline rung (SURVEY 6.2 / BASELINE.md section 2): the reference's communication
structure re-hosted on torch.distributed/gloo - per-variable blocking send/recv through host memory,
PS-side Python loop with a host sum, optimizer on the PS, serial fan-out - with torch ops (cuDNN /
cuBLAS) for compute on the worker GPUs.  This stands in for "the reference's own build", which
cannot be installed in this image (no TensorFlow 1.x / mpi4py / mpiexec).

Run through ``python bench.py --impl host_ps --gpus N``: N PS ranks (CPU) + N workers (one GPU
each), the ``mnist_sync_sharding_greedy`` message pattern, batch 100 per worker."""
from __future__ import annotations

import glob
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def run(a, emit=lambda o: print(json.dumps(o))) -> int:
    if int(os.environ.get("RANK", "0")) != 0:
        return 0                                   # under torchrun only rank 0 drives the MPMD job
    from ddl_b200.parallel.launch import launch
    n = a.gpus
    steps = a.steps + max(3, a.warmup)
    with tempfile.TemporaryDirectory() as td:
        mpath = os.path.join(td, "m.jsonl")
        ev = int(getattr(a, "eval_every", 0) or 0)          # the reference evaluates every 10 batches INSIDE its timed region
        rows = int(getattr(a, "test_rows", 10000)) if ev else 1000     # (mnist_sync/worker.py:71-72): bench.py --eval-every 10
        extra = ["--synthetic", "--synthetic-train", str(max(6000, 100 * n * (steps + 1))), "--synthetic-test", str(rows),
                 "--seed", "1234", "--eval-every", str(ev), "--max-steps", str(steps), "--backend", "torch",
                 "--metrics-path", mpath, "--batch-size", str(a.batch_size), "--width-mult", str(a.width_mult)]
        env_keep = {k: os.environ.pop(k) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_PORT") if k in os.environ}
        try:
            rc = launch(os.path.join(ROOT, a.variant), n, n, extra, transport="host", timeout=1200)
        finally:
            os.environ.update(env_keep)
        if rc != 0:
            emit({"impl": "host_ps", "error": "launcher exited with %d" % rc})
            return rc
        rows = []
        for f in glob.glob(mpath + "*"):
            rows += [json.loads(l) for l in open(f) if "steady_samples_per_s" in l]
    import torch
    wdev = "gpu" if torch.cuda.is_available() else "cpu"
    slow = max(r["steady_s"] for r in rows)
    steady_steps = rows[0]["steady_steps"]
    emit({
        "metric": "samples/sec", "value": steady_steps * a.batch_size * n / slow, "unit": "samples/s", "n_gpus": n,
        "steps": steady_steps, "warmup": 3, "ms_per_step": slow * 1e3 / steady_steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "fp32 (torch ops, TF32 off)", "impl": "host_ps",
        "data": "synthetic MNIST-shaped", "timing": "host wall clock, max over workers (transport is host-side)",
        "config": {"model": "mnist_cnn_14var", "variant": a.variant, "global_batch": a.batch_size * n,
                   "eval_every": int(getattr(a, "eval_every", 0) or 0),
                   "parallelism": "ps%d(cpu)+w%d(%s), host/gloo transport" % (n, n, wdev)}})
    return 0

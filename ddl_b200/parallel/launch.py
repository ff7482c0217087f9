"""MPMD launcher: the ``mpiexec -n P ... parameter_server.py -np P : -n W ... worker.py -np W`` line of
the reference's run.sh (mnist_sync_sharding/run.sh:3) without MPI.

Spawns one OS process per rank with ``RANK`` / ``WORLD_SIZE`` / ``MASTER_ADDR=127.0.0.1`` set, PS
ranks first (``[0, P)``), then workers (``[P, P+W)``) - the reference's rank layout.  For the device
transports the number of processes is the number of GPUs used (see roles.processes_for): a shard
and a worker that must share a GPU share ONE process.

Failure detection (absent in the reference): if any rank exits non-zero the launcher
terminates the rest instead of letting blocking receives hang forever; ``--timeout`` bounds the run.
"""
from __future__ import annotations

import argparse
import os
import signal
import socket
import subprocess
import sys
import time
from typing import List

from .roles import processes_for


def free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def launch(variant_dir: str, num_ps: int, num_workers: int, extra: List[str], transport: str = "auto",
           timeout: float = 0.0, python: str = sys.executable) -> int:
    try:
        import torch
        ngpu = torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        ngpu = 0
    if transport == "auto":
        transport = "fused" if ngpu else "host"
    if "--device" in extra and extra[extra.index("--device") + 1] == "cpu":
        ngpu = 0
        if transport == "fused":
            transport = "host"
    if transport == "host":
        nproc, kind = num_ps + num_workers, "mpmd"
    elif ngpu == 0:                     # collective transport over gloo on CPU: one process per role
        nproc, kind = num_ps + num_workers, "dedicated"
    else:
        nproc, kind = processes_for(num_ps, num_workers, transport, ngpu)
    port = free_port()
    procs: List[subprocess.Popen] = []
    for r in range(nproc):
        if kind in ("mpmd", "dedicated"):
            role = "ps" if r < num_ps else "worker"
        else:
            role = "worker" if r < num_workers else "ps"
        script = "parameter_server.py" if role == "ps" else "worker.py"
        np_flag = str(num_ps if role == "ps" else num_workers)
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(nproc), LOCAL_RANK=str(r),
                   MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), DDL_NUM_PS=str(num_ps),
                   DDL_NUM_WORKERS=str(num_workers), DDL_LAYOUT="auto" if kind == "mpmd" else kind)
        cmd = [python, script, "-np", np_flag, "--transport", transport] + list(extra)
        procs.append(subprocess.Popen(cmd, cwd=variant_dir, env=env))
    deadline = time.time() + timeout if timeout else None
    rc = 0
    try:
        alive = set(range(nproc))
        while alive:
            for i in list(alive):
                code = procs[i].poll()
                if code is not None:
                    alive.discard(i)
                    if code != 0 and rc == 0:
                        rc = code
                        print("[launch] rank %d exited with %d - stopping the job" % (i, code),
                              file=sys.stderr, flush=True)
                        for j in alive:
                            procs[j].send_signal(signal.SIGTERM)
            if deadline and time.time() > deadline:
                print("[launch] timeout - stopping the job", file=sys.stderr, flush=True)
                rc = rc or 124
                for j in alive:
                    procs[j].send_signal(signal.SIGTERM)
                deadline = None
            time.sleep(0.05)
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
    return rc


def launch_with_retry(variant_dir: str, num_ps: int, num_workers: int, extra: List[str], transport: str = "auto",
                      timeout: float = 0.0, attempts: int = 2) -> int:
    """``launch``; a job that dies within its first seconds (rendezvous port taken between ``free_port`` and the
    bind, a rank killed while importing) is started once more on a fresh port.  Later failures are returned as is."""
    rc = 0
    for k in range(max(1, attempts)):
        t0 = time.time()
        rc = launch(variant_dir, num_ps, num_workers, extra, transport, timeout)
        if rc == 0 or rc == 124 or time.time() - t0 > 20.0 or k + 1 == attempts:
            return rc
        print("[launch] job failed after %.1f s (rc %d) - retrying on a new port" % (time.time() - t0, rc),
              file=sys.stderr, flush=True)
    return rc


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description="run.sh <num_ps> <num_workers> [role flags...]")
    ap.add_argument("--variant-dir", default=".")
    ap.add_argument("--transport", default=os.environ.get("DDL_TRANSPORT", "auto"))
    ap.add_argument("--timeout", type=float, default=float(os.environ.get("DDL_LAUNCH_TIMEOUT", "0")))
    ap.add_argument("num_ps", type=int)
    ap.add_argument("num_workers", type=int)
    ns, extra = ap.parse_known_args(argv)
    return launch_with_retry(os.path.abspath(ns.variant_dir), ns.num_ps, ns.num_workers, extra, ns.transport, ns.timeout,
                             attempts=int(os.environ.get("DDL_LAUNCH_ATTEMPTS", "2")))


if __name__ == "__main__":
    sys.exit(main())

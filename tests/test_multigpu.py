"""Distributed tier, GPU: the fused peer-memory transport on >= 2 GPUs of one box.
Sync: workers bitwise-equal after every step and equal to the NCCL rung; async: no lost updates."""
import os
import subprocess
import sys

import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


NPROC = int(os.environ.get("DDL_TEST_NPROC", "2"))     # 2, 4 or 8 (profiles/mgpu_tests_n*.log are runs of this file)


def run_case(case, nproc=None, timeout=300):
    nproc = nproc or NPROC
    port = 29600 + (abs(hash(case)) % 200)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc),
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "mgpu_worker.py"), case]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT)
    ok = r.returncode == 0 and "MGPU_OK " + case in r.stdout
    if not ok:
        # the workers' own tracebacks sit far above torchrun's summary: keep the first one and every assertion line
        err = r.stderr
        first = err.find("Traceback")
        keep = err[first:first + 3000] if first >= 0 else ""
        lines = [ln for ln in err.splitlines() if "Error" in ln or "assert" in ln.lower()][:20]
        assert ok, "\n".join([r.stdout[-1500:], keep, "\n".join(lines), err[-1500:]])


@pytest.mark.parametrize("case", ["sync_colocated", "sync_dedicated", "sync_vs_nccl", "sync_nvls_off", "sync_nvls_on", "async_sgd", "async_adam", "async_parity", "async_dedicated", "async_train", "ckpt_resume"])
def test_fused_transport(case):
    run_case(case)

import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    # In-process oracle math: on the small virtualised CPUs these suites run on, torch's default (one thread per vCPU)
    # is 10x SLOWER than two threads (bench.py --impl oracle_cpu: 27 vs 430 samples/s) - oversubscribed OpenMP spinning.
    try:
        import torch
        torch.set_num_threads(int(os.environ.get("DDL_TEST_THREADS", "2")))
    except Exception:                                            # noqa: BLE001
        pass
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 CUDA devices")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        have = torch.cuda.is_available()
        ngpu = torch.cuda.device_count() if have else 0
    except Exception:
        have, ngpu = False, 0
    skip_gpu = pytest.mark.skip(reason="no CUDA device")
    skip_multi = pytest.mark.skip(reason="needs >= 2 CUDA devices")
    for it in items:
        if "gpu" in it.keywords and not have:
            it.add_marker(skip_gpu)
        if "multigpu" in it.keywords and ngpu < 2:
            it.add_marker(skip_multi)

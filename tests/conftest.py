import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_sessionstart(session):
    """A fresh checkout has no libpsb200.so (it is git-ignored): build it once, before collection, when nvcc is there
    (cross-compiles without a GPU, under a minute).  Without a compiler the tests of the native code fail loudly."""
    import shutil
    import subprocess
    if hasattr(session.config, "workerinput"):          # pytest-xdist worker: the controller has built it
        return
    lib = os.path.join(ROOT, "parameter-server-architecture-on-apache-kafka_b200", "ops", "libpsb200.so")
    if not os.path.exists(lib) and (shutil.which("nvcc") or os.path.exists("/usr/local/cuda/bin/nvcc")):
        subprocess.run(["bash", os.path.join(ROOT, "build.sh")], cwd=ROOT, check=False,
                       stdout=subprocess.DEVNULL, stderr=subprocess.STDOUT)


def pytest_collection_modifyitems(config, items):
    """GPU tests are skipped (not failed) when collected on a machine without a device."""
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)

import os
import sys

import pytest

# rank-threads of the single-GPU multi-rank tests use several streams each; keep them on distinct hardware queues
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 CUDA devices")


def pytest_sessionstart(session):
    """Test-side convenience only (the product itself fails loudly without its library): on a fresh checkout build
    elemental_b200/libhydrogen_b200.so before collection, as __graft_entry__.build() would.  nvcc cross-compiles without a GPU."""
    import subprocess
    lib = os.path.join(ROOT, "elemental_b200", "libhydrogen_b200.so")
    if not os.path.exists(lib):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "elemental_b200", "csrc"), "-j8", "-s"])

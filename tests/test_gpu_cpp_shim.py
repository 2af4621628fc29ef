"""-m gpu: the C++ shim (include/El.hpp) driver in the style of tests/blas_like/BasicGemm.cpp, 4 rank-threads, 2x2 grid."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_basicgemm_cpp_driver():
    exe = os.path.join(ROOT, "examples", "BasicGemm")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "examples")])
    for args in (["1", "1", "257", "128"], ["4", "2", "300", "96"]):
        out = subprocess.run([exe] + args, capture_output=True, text=True, timeout=120)
        assert out.returncode == 0 and "PASSED" in out.stdout, out.stdout + out.stderr

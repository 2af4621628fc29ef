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


def test_basicsyrk_cpp_driver():
    """examples/BasicSyrk.cpp: tests/blas_like/Syrk.cpp restated against the shim (Align, MakeSymmetric, Syrk, both triangles
    and orientations) plus the shim's alignment rules (adoption in B = A, AlignWith)."""
    exe = os.path.join(ROOT, "examples", "BasicSyrk")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "examples")])
    for args in (["1", "1", "257", "128", "64"], ["4", "2", "300", "200", "64"], ["6", "3", "1301", "96", "32"]):
        out = subprocess.run([exe] + args, capture_output=True, text=True, timeout=180)
        assert out.returncode == 0 and "PASSED" in out.stdout, out.stdout + out.stderr


def test_shimops_cpp_driver():
    """examples/ShimOps.cpp: El::LocalGemm (conformity checks + local product, against a host long-double
    product) and MultiSync (a24: A, B, C on three streams with no host synchronisation, bit-identical to the synchronised run)."""
    exe = os.path.join(ROOT, "examples", "ShimOps")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "examples")])
    for args in (["1", "1", "257", "64"], ["4", "2", "300", "64"], ["6", "3", "333", "32"]):
        out = subprocess.run([exe] + args, capture_output=True, text=True, timeout=180)
        assert out.returncode == 0 and "PASSED" in out.stdout, out.stdout + out.stderr


def test_shim_header_instantiates_every_wrapper():
    """tests/shim_check.cpp: every wrapper of include/El.hpp instantiates for double and float against the C ABI (compile-only)."""
    subprocess.check_call(["g++", "-std=c++17", "-fsyntax-only", "-I" + os.path.join(ROOT, "include"), "-I/usr/local/cuda/include",
                           os.path.join(ROOT, "tests", "shim_check.cpp")])


def test_distgemm_one_process_world_from_the_environment():
    """examples/DistGemm.cpp: the client-program form (El::Environment, mpi::NewWorldComm(), Grid(std::move(comm), height)) with
    the world taken from the launcher's environment (hyd_world_create_env).  One process here; tools/gpu_multi.sh launches N."""
    exe = os.path.join(ROOT, "examples", "DistGemm")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "examples")])
    env = dict(os.environ, RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
    out = subprocess.run([exe, "1", "300", "64"], capture_output=True, text=True, timeout=180, env=env)
    assert out.returncode == 0 and "PASSED" in out.stdout, out.stdout + out.stderr

"""elemental_b200 -- B200-native distributed-GEMM hot path of Hydrogen/Elemental.

Python here is test/bench plumbing over the C ABI (include/hydrogen_b200.h); the
product is libhydrogen_b200.so and the El:: C++ shim in include/El.hpp.
"""
import os as _os

# The engine drives up to 5 streams per rank (compute, 2 comm, upload, unpack); with the default 8 hardware queues
# stream memory operations of different streams can alias onto one queue and serialise. Must precede CUDA init.
_os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

from ._lib import lib, check, HydrogenError, LIB_PATH, hyd_dm  # noqa: F401

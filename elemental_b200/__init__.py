"""elemental_b200 -- B200-native distributed-GEMM hot path of Hydrogen/Elemental.

Python here is test/bench plumbing over the C ABI (include/hydrogen_b200.h); the
product is libhydrogen_b200.so and the El:: C++ shim in include/El.hpp.
"""
from ._lib import lib, check, HydrogenError, LIB_PATH, hyd_dm  # noqa: F401

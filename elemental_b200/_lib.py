"""ctypes binding of libhydrogen_b200.so (the C ABI in include/hydrogen_b200.h).

The library is the product; this module only loads it and declares signatures.
There is no CPU fallback: if the shared object is missing, import fails loudly.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libhydrogen_b200.so")


class HydrogenError(RuntimeError):
    pass


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `make -C elemental_b200/csrc` "
            "(or __graft_entry__.build()); there is no CPU fallback.")
    return C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)


lib = _load()

i32, i64, f32, f64, vp = C.c_int, C.c_int64, C.c_float, C.c_double, C.c_void_p

lib.hyd_last_error.restype = C.c_char_p
lib.hyd_last_error.argtypes = []
lib.hyd_version.restype = i32
lib.hyd_lgemm_f64.argtypes = [i32, i32, i64, i64, i64, f64, vp, i64, vp, i64, f64, vp, i64, vp]
lib.hyd_probe_peak.argtypes = [i32, i32, i32, C.POINTER(f64)]


def check(rc):
    if rc != 0:
        raise HydrogenError(lib.hyd_last_error().decode())

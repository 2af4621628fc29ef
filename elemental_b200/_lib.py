"""ctypes binding of libhydrogen_b200.so (the C ABI in include/hydrogen_b200.h).

The library is the product; this module only loads it and declares signatures.
There is no CPU fallback: if the shared object is missing, import fails loudly.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("HYD_LIB") or os.path.join(_HERE, "libhydrogen_b200.so")  # HYD_LIB: developer A/B builds


class HydrogenError(RuntimeError):
    pass


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `make -C elemental_b200/csrc` "
            "(or __graft_entry__.build()); there is no CPU fallback.")
    return C.CDLL(LIB_PATH)


lib = _load()

i32, i64, f32, f64, vp = C.c_int, C.c_int64, C.c_float, C.c_double, C.c_void_p


class hyd_dm(C.Structure):
    """POD mirror of `hyd_dm` in include/hydrogen_b200.h."""
    _fields_ = [("buf", vp), ("ldim", i64), ("height", i64), ("width", i64),
                ("colDist", i32), ("rowDist", i32), ("colAlign", i32), ("rowAlign", i32),
                ("root", i32), ("reserved", i32)]


dmp = C.POINTER(hyd_dm)
pi64 = C.POINTER(i64)


def _sig(name, res, args):
    f = getattr(lib, name)
    f.restype = res
    f.argtypes = args
    return f


_sig("hyd_last_error", C.c_char_p, [])
_sig("hyd_version", i32, [])
_sig("hyd_launch_count", C.c_longlong, [])
_sig("hyd_alloc_count", C.c_longlong, [])
_sig("hyd_sync_count", C.c_longlong, [])
_sig("hyd_multisync_begin", i32, [vp, C.POINTER(vp), i32])
_sig("hyd_multisync_end", i32, [vp, C.POINTER(vp), i32])
_sig("hyd_grid_set_host_blocks", i32, [vp, i32])
_sig("hyd_grid_profile_trace", i32, [vp, i32, i32, C.POINTER(f64), C.POINTER(f64), i32])
_sig("hyd_local_trrk_static_f64", i32, [i32, i32, i32, i64, i64, i64, i64, i64, i64, i64, f64, vp, i64, vp, i64, f64, vp, i64, vp])
_sig("hyd_host_block_plan", i32, [i64, i64, i64, f64, f64, i32, C.POINTER(i64)])
_sig("hyd_lgemm_f64", i32, [i32, i32, i64, i64, i64, f64, vp, i64, vp, i64, f64, vp, i64, vp])
_sig("hyd_lgemm_f32", i32, [i32, i32, i64, i64, i64, f32, vp, i64, vp, i64, f32, vp, i64, vp])
_sig("hyd_lgemm_f16", i32, [i32, i32, i64, i64, i64, f32, vp, i64, vp, i64, f32, vp, i64, vp])
_sig("hyd_lgemm_bf16", i32, [i32, i32, i64, i64, i64, f32, vp, i64, vp, i64, f32, vp, i64, vp])
_sig("hyd_transpose_f64", i32, [i64, i64, vp, i64, vp, i64, vp])
_sig("hyd_transpose_f32", i32, [i64, i64, vp, i64, vp, i64, vp])
_sig("hyd_copy2d_f64", i32, [i64, i64, vp, i64, vp, i64, vp])
_sig("hyd_copy2d_f32", i32, [i64, i64, vp, i64, vp, i64, vp])
_sig("hyd_axpy2d_f64", i32, [i64, i64, f64, vp, i64, vp, i64, vp])
_sig("hyd_axpy2d_f32", i32, [i64, i64, f32, vp, i64, vp, i64, vp])
_sig("hyd_scale2d_f64", i32, [i64, i64, f64, vp, i64, vp])
_sig("hyd_scale2d_f32", i32, [i64, i64, f32, vp, i64, vp])
_sig("hyd_mod", i64, [i64, i64])
_sig("hyd_shift", i64, [i64, i64, i64])
_sig("hyd_length", i64, [i64, i64, i64])
_sig("hyd_max_length", i64, [i64, i64])
_sig("hyd_nccl_unique_id", i32, [vp])
_sig("hyd_world_create_nccl", i32, [vp, i32, i32, C.POINTER(vp)])
_sig("hyd_local_group_create", i32, [i32, C.POINTER(vp)])
_sig("hyd_local_group_destroy", i32, [vp])
_sig("hyd_world_create_local", i32, [vp, i32, C.POINTER(vp)])
_sig("hyd_world_destroy", i32, [vp])
_sig("hyd_world_barrier", i32, [vp])
_sig("hyd_grid_default_height", i32, [i32])
_sig("hyd_world_size", i32, [vp])
_sig("hyd_world_rank", i32, [vp])
_sig("hyd_world_kind", C.c_char_p, [vp])
_sig("hyd_world_create_env", i32, [C.POINTER(vp)])
_sig("hyd_env_world_info", i32, [C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)])
_sig("hyd_env_exchange_id", i32, [i32, vp, i32])
_sig("hyd_env_exchange_done", i32, [i32])
_sig("hyd_grid_create", i32, [vp, i32, i32, C.POINTER(vp)])
_sig("hyd_grid_destroy", i32, [vp])
for _n in ("hyd_grid_height", "hyd_grid_width", "hyd_grid_size", "hyd_grid_rank", "hyd_grid_row", "hyd_grid_col",
           "hyd_grid_vc_rank", "hyd_grid_vr_rank"):
    _sig(_n, i32, [vp])
_sig("hyd_grid_set_overlap", i32, [vp, i32, i32])
_sig("hyd_grid_profile", i32, [vp, i32])
_sig("hyd_grid_profile_read", i32, [vp, i32, C.POINTER(C.c_longlong), C.POINTER(f64)])
_sig("hyd_dist_stride", i32, [vp, i32])
_sig("hyd_dist_rank", i32, [vp, i32])
_sig("hyd_local_shape", i32, [vp, dmp, pi64, pi64, pi64, pi64])
_sig("hyd_local_shape_of", i32, [i32, i32, i32, i32, dmp, pi64, pi64, pi64, pi64])
_sig("hyd_view", i32, [vp, dmp, i64, i64, i64, i64, i32, dmp])
_sig("hyd_redist_plan", i32, [i32, i32, i32, i32, dmp, dmp, i32, i32, pi64, pi64])
_sig("hyd_redist_f64", i32, [vp, dmp, dmp, i32, vp])
_sig("hyd_redist_f32", i32, [vp, dmp, dmp, i32, vp])
_sig("hyd_axpy_contract_f64", i32, [vp, f64, dmp, dmp, i32, vp])
_sig("hyd_axpy_contract_f32", i32, [vp, f32, dmp, dmp, i32, vp])
_sig("hyd_gemm_f64", i32, [vp, i32, i32, i32, f64, dmp, dmp, f64, dmp, i64, vp])
_sig("hyd_gemm_f32", i32, [vp, i32, i32, i32, f32, dmp, dmp, f32, dmp, i64, vp])
for _sfx, _ft in (("f64", f64), ("f32", f32)):
    _sig("hyd_trrk_" + _sfx, i32, [vp, i32, i32, i32, _ft, dmp, dmp, _ft, dmp, i64, vp])
    _sig("hyd_symm_" + _sfx, i32, [vp, i32, i32, _ft, dmp, dmp, _ft, dmp, i64, vp])
    _sig("hyd_syrk_" + _sfx, i32, [vp, i32, i32, _ft, dmp, _ft, dmp, i64, vp])
    _sig("hyd_scale_trapezoid_" + _sfx, i32, [_ft, i32, i64, i64, vp, i64, i64, vp])
    _sig("hyd_axpy_trapezoid_" + _sfx, i32, [i32, _ft, i64, i64, vp, i64, vp, i64, i64, vp])
    _sig("hyd_make_trapezoidal_" + _sfx, i32, [i32, i64, i64, vp, i64, i64, vp])
    _sig("hyd_scale_trapezoid_dist_" + _sfx, i32, [vp, _ft, i32, dmp, i64, vp])
    _sig("hyd_axpy_trapezoid_dist_" + _sfx, i32, [vp, i32, _ft, dmp, dmp, i64, vp])
    _sig("hyd_make_trapezoidal_dist_" + _sfx, i32, [vp, i32, dmp, i64, vp])
    _sig("hyd_make_symmetric_dist_" + _sfx, i32, [vp, i32, dmp, vp])
_sig("hyd_gemm_default_alg", i32, [i64, i64, i64])
_sig("hyd_adopted_align", i32, [i32, i32, i32, i32, i32])
_sig("hyd_trrk_plan", i32, [i32, i32, i32, i32, dmp, i32, i64, i64, i32, C.POINTER(i64)])
_sig("hyd_gemm_f16", i32, [vp, i32, i32, i32, f32, dmp, dmp, f32, dmp, i64, vp])
_sig("hyd_gemm_bf16", i32, [vp, i32, i32, i32, f32, dmp, dmp, f32, dmp, i64, vp])
_sig("hyd_redist_b16", i32, [vp, dmp, dmp, i32, vp])
_sig("hyd_gemm_host_f64", i32, [vp, i32, i32, i32, f64, dmp, dmp, f64, dmp, i64, vp])
_sig("hyd_gemm_host_f32", i32, [vp, i32, i32, i32, f32, dmp, dmp, f32, dmp, i64, vp])
_sig("hyd_probe_peak", i32, [i32, i32, i32, C.POINTER(f64)])
_sig("hyd_memcopy2d_async", i32, [vp, i64, vp, i64, i64, i64, i32, i32, vp])
_sig("hyd_memcopy1d_async", i32, [vp, vp, i64, i32, i32, vp])
# seam S2: copy::util pack / unpack and axpy::util::InterleaveMatrixUpdate with the reference's argument lists
for _sf, _ft in (("f64", f64), ("f32", f32)):
    _sig("hyd_interleave_" + _sf, i32, [i64, i64, vp, i64, i64, vp, i64, i64, vp])
    _sig("hyd_interleave_update_" + _sf, i32, [_ft, i64, i64, vp, i64, i64, vp, i64, i64, vp])
    for _n in ("row", "col"):
        _sig(f"hyd_{_n}_strided_pack_{_sf}", i32, [i64, i64, i64, i64, vp, i64, vp, i64, vp])
        _sig(f"hyd_{_n}_strided_unpack_{_sf}", i32, [i64, i64, i64, i64, vp, i64, vp, i64, vp])
        _sig(f"hyd_partial_{_n}_strided_pack_{_sf}", i32, [i64] * 8 + [vp, i64, vp, i64, vp])
        _sig(f"hyd_partial_{_n}_strided_unpack_{_sf}", i32, [i64] * 8 + [vp, i64, vp, i64, vp])
    _sig("hyd_strided_pack_" + _sf, i32, [i64] * 6 + [vp, i64, vp, i64, vp])
    _sig("hyd_strided_unpack_" + _sf, i32, [i64] * 6 + [vp, i64, vp, i64, vp])
# the rest of the GPU BLAS-1 layer (level1.cu)
for _sf, _ft in (("f64", f64), ("f32", f32)):
    _sig("hyd_fill_" + _sf, i32, [i64, i64, _ft, vp, i64, vp])
    _sig("hyd_hadamard_" + _sf, i32, [i64, i64, vp, i64, i64, vp, i64, i64, vp, i64, i64, vp])
    _sig("hyd_axpy_trans_" + _sf, i32, [i32, i64, i64, _ft, vp, i64, vp, i64, vp])
    _sig("hyd_copy1d_" + _sf, i32, [i64, vp, i64, vp, i64, vp])
    _sig("hyd_axpy1d_" + _sf, i32, [i64, _ft, vp, i64, vp, i64, vp])
    _sig("hyd_scale1d_" + _sf, i32, [i64, _ft, vp, i64, vp])
    _sig("hyd_axpy_dist_" + _sf, i32, [vp, _ft, dmp, dmp, i32, vp])
# seam S2: collectives on the grid's sub-communicators
for _n in ("hyd_comm_size", "hyd_comm_rank"):
    _sig(_n, i32, [vp, i32])
_sig("hyd_comm_world_rank", i32, [vp, i32, i32])
_sig("hyd_allgather", i32, [vp, i32, vp, i64, vp, vp])
_sig("hyd_alltoall", i32, [vp, i32, vp, i64, vp, vp])
_sig("hyd_sendrecv", i32, [vp, i32, vp, i64, i32, vp, i64, i32, vp])
_sig("hyd_bcast", i32, [vp, i32, vp, i64, i32, vp])
_sig("hyd_reducescatter_f64", i32, [vp, i32, vp, vp, i64, vp])
_sig("hyd_reducescatter_f32", i32, [vp, i32, vp, vp, i64, vp])

# every symbol include/hydrogen_b200.h declares (tests/test_abi.py checks the header against this list)
EXPORTS = sorted(n for n in dir(lib) if n.startswith("hyd_"))


def check(rc):
    if rc != 0:
        raise HydrogenError(lib.hyd_last_error().decode())

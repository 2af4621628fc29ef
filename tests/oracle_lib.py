"""ctypes wrapper of oracle/liboracle.so -- the CPU restatement of the reference (checker only).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/--impl reference leg import this.
"""
import ctypes as C
import glob
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB = os.path.join(ORACLE_DIR, "liboracle.so")

MC, MD, MR, VC, VR, STAR, CIRC = range(7)
DIST_NAMES = {MC: "MC", MR: "MR", VC: "VC", VR: "VR", STAR: "STAR"}
# the 11 GPU-valid element distribution pairs (tests/core/DistMatrix.cpp:100-112 skips MD/CIRC on GPU)
PAIRS = [(MC, MR), (MC, STAR), (MR, MC), (MR, STAR), (STAR, MC), (STAR, MR), (STAR, STAR),
         (STAR, VC), (STAR, VR), (VC, STAR), (VR, STAR)]


def build():
    srcs = glob.glob(os.path.join(ORACLE_DIR, "*.c")) + [os.path.join(ORACLE_DIR, "oracle.h")]
    if not os.path.exists(LIB) or any(os.path.getmtime(s) > os.path.getmtime(LIB) for s in srcs):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "-s"])


build()
lib = C.CDLL(LIB)
i32, i64, f64, vp, u64 = C.c_int, C.c_int64, C.c_double, C.c_void_p, C.c_uint64


class Grid(C.Structure):
    _fields_ = [("r", i32), ("c", i32), ("col_major", i32)]


def _sig(name, res, args):
    f = getattr(lib, name)
    f.restype = res
    f.argtypes = args
    return f


for n in ("orc_mod", "orc_shift", "orc_length"):
    _sig(n, i64, [i64, i64] + ([i64] if n != "orc_mod" else []))
_sig("orc_max_length", i64, [i64, i64])
_sig("orc_dist_stride", i32, [Grid, i32])
_sig("orc_dist_rank", i32, [Grid, i32, i32, i32])
_sig("orc_grid_rank_of", i32, [Grid, i32, i32])
_sig("orc_grid_default_height", i32, [i32])
_sig("orc_dist_collect", i32, [i32])
_sig("orc_dist_partial", i32, [i32])
_sig("orc_dist_partial_union_row", i32, [i32, i32])
_sig("orc_dist_partial_union_col", i32, [i32, i32])
_sig("orc_align_dim_with", i32, [i32] * 8 + [C.POINTER(C.c_int)])
_sig("orc_copy_adopts", i32, [i32] * 4 + [C.POINTER(C.c_int)])
_sig("orc_dm_create", vp, [Grid, i64, i64, i32, i32, i32, i32, i32])
_sig("orc_dm_free", None, [vp])
_sig("orc_dm_local", vp, [vp, i32])
for n in ("orc_dm_lh", "orc_dm_lw", "orc_dm_ld"):
    _sig(n, i64, [vp, i32])
_sig("orc_dm_view", vp, [vp, i64, i64, i64, i64])
_sig("orc_dm_fill", None, [vp, u64])
_sig("orc_dm_fill_index", None, [vp])
_sig("orc_dm_fill_const", None, [vp, f64])
_sig("orc_dm_to_global", i64, [vp, vp])
_sig("orc_dm_from_global", None, [vp, vp])
_sig("orc_uniform", f64, [u64, u64])
_sig("orc_redist_closed", None, [vp, vp, i32])
_sig("orc_axpy_contract_closed", None, [f64, vp, vp, i32])
for n in ("orc_row_allgather", "orc_col_allgather", "orc_col_alltoall_demote", "orc_row_alltoall_demote",
          "orc_partial_col_allgather", "orc_partial_row_allgather", "orc_vector_exchange"):
    _sig(n, i32, [vp, vp])
_sig("orc_local_transpose", None, [vp, vp])
_sig("orc_axpy", i32, [f64, vp, vp, i32])
_sig("orc_hadamard", i32, [vp, vp, vp])
_sig("orc_scale", None, [f64, vp])
_sig("orc_set_blas", i32, [C.c_char_p])
_sig("orc_blas_threads", i32, [i32])
_sig("orc_local_gemm", None, [i32, i32, i64, i64, i64, f64, vp, i64, vp, i64, f64, vp, i64, i32])
_sig("orc_gemm", i32, [i32, i32, i32, f64, vp, vp, f64, vp, i64])
_sig("orc_scale_trapezoid", None, [f64, i32, vp, i64])
_sig("orc_axpy_trapezoid", i32, [i32, f64, vp, vp, i64])
_sig("orc_make_trapezoidal", None, [i32, vp, i64])
_sig("orc_make_symmetric", i32, [i32, vp])
_sig("orc_symm", i32, [i32, i32, f64, vp, vp, f64, vp, i64])
_sig("orc_set_local_trrk_blocksize", None, [i32])
_sig("orc_trrk", i32, [i32, i32, i32, f64, vp, vp, f64, vp, i64])
_sig("orc_syrk", i32, [i32, i32, f64, vp, f64, vp, i64])
_sig("orc_dense_gemm_ref", None, [i32, i32, i64, i64, i64, f64, vp, i64, vp, i64, f64, vp, i64, i32])
# pack/unpack
_P = [i64, i64, i64, i64, vp, i64, vp, i64, i32]
for n in ("orc_row_strided_pack", "orc_row_strided_unpack", "orc_col_strided_pack", "orc_col_strided_unpack"):
    _sig(n, None, _P)
_PP = [i64, i64, i64, i64, i64, i64, i64, i64, vp, i64, vp, i64, i32]
for n in ("orc_partial_row_strided_pack", "orc_partial_row_strided_unpack", "orc_partial_col_strided_pack",
          "orc_partial_col_strided_unpack"):
    _sig(n, None, _PP)
_sig("orc_strided_pack", None, [i64, i64, i64, i64, i64, i64, vp, i64, vp, i64, i32])
_sig("orc_strided_unpack", None, [i64, i64, i64, i64, i64, i64, vp, i64, vp, i64, i32])
_sig("orc_interleave", None, [i64, i64, vp, i64, i64, vp, i64, i64, i32])
_sig("orc_interleave_update", None, [f64, i64, i64, vp, i64, i64, vp, i64, i64, i32])


def find_blas():
    try:
        import scipy
        c = glob.glob(os.path.join(os.path.dirname(scipy.__file__), "..", "scipy.libs", "libscipy_openblas*.so"))
        return c[0] if c else None
    except Exception:
        return None


def use_blas(enable=True):
    p = find_blas() if enable else None
    if p is None:
        lib.orc_set_blas(None)
        return False
    return lib.orc_set_blas(p.encode()) == 0


class DM:
    """A distributed matrix on the simulated grid (all ranks' local blocks live here)."""

    def __init__(self, grid, m, n, cd, rd, ca=0, ra=0, dtype=np.float64, handle=None, parent=None):
        self.grid, self.m, self.n, self.cd, self.rd, self.ca, self.ra = grid, m, n, cd, rd, ca, ra
        self.dtype = np.dtype(dtype)
        self.P = grid.r * grid.c
        self.parent = parent
        self.h = handle if handle is not None else lib.orc_dm_create(grid, m, n, cd, rd, ca, ra, self.dtype.itemsize)

    def __del__(self):
        if getattr(self, "h", None):
            lib.orc_dm_free(self.h)
            self.h = None

    def lh(self, p): return lib.orc_dm_lh(self.h, p)
    def lw(self, p): return lib.orc_dm_lw(self.h, p)
    def ld(self, p): return lib.orc_dm_ld(self.h, p)

    def local(self, p):
        """numpy view (ld x lw, Fortran order) of rank p's local buffer; rows [:lh] are the block."""
        ld, lw = self.ld(p), max(self.lw(p), 0)
        n = ld * max(lw, 1)
        buf = (C.c_char * (n * self.dtype.itemsize)).from_address(lib.orc_dm_local(self.h, p))
        return np.frombuffer(buf, dtype=self.dtype, count=ld * lw).reshape((ld, lw), order="F")

    def block(self, p):
        return self.local(p)[: self.lh(p), :]

    def view(self, i0, j0, h, w):
        g = self.grid
        cs, rs = lib.orc_dist_stride(g, self.cd), lib.orc_dist_stride(g, self.rd)
        return DM(g, h, w, self.cd, self.rd, (self.ca + i0) % cs, (self.ra + j0) % rs, self.dtype,
                  handle=lib.orc_dm_view(self.h, i0, j0, h, w), parent=self)

    def fill(self, seed): lib.orc_dm_fill(self.h, seed); return self
    def fill_index(self): lib.orc_dm_fill_index(self.h); return self
    def fill_const(self, v): lib.orc_dm_fill_const(self.h, float(v)); return self

    def to_global(self):
        out = np.zeros((self.m, self.n), dtype=self.dtype, order="F")
        mism = lib.orc_dm_to_global(self.h, out.ctypes.data)
        return out, mism

    def from_global(self, G):
        G = np.asfortranarray(G, dtype=self.dtype)
        assert G.shape == (self.m, self.n)
        lib.orc_dm_from_global(self.h, G.ctypes.data)
        return self


def gemm(oA, oB, alg, alpha, A, B, beta, Cm, nb):
    rc = lib.orc_gemm(oA, oB, alg, alpha, A.h, B.h, beta, Cm.h, nb)
    assert rc == 0, rc

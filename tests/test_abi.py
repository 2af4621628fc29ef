"""CPU-only: the C-ABI library loads, exports every symbol include/hydrogen_b200.h declares, and its
host-only entry points (index arithmetic, local shapes, views' metadata, redistribution planner) agree
with the oracle.  No kernels are launched here."""
import ctypes as C
import itertools
import os
import re

import numpy as np
import pytest

import oracle_lib as O
from elemental_b200 import lib, hyd_dm, LIB_PATH

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "hydrogen_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(hyd_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    names = declared_symbols()
    assert len(names) >= 50
    raw = C.CDLL(LIB_PATH)
    missing = [n for n in names if not hasattr(raw, n)]
    assert not missing, missing
    assert lib.hyd_version() >= 100


def test_no_link_time_dependency_on_the_oracle_or_nccl():
    import subprocess
    out = subprocess.run(["ldd", LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in out and "libnccl" not in out  # NCCL is bound at run time (comm.cu)


def test_index_arithmetic_matches_oracle():
    for n in range(0, 50):
        for stride in range(1, 9):
            assert lib.hyd_max_length(n, stride) == O.lib.orc_max_length(n, stride)
            for shift in range(stride):
                assert lib.hyd_length(n, shift, stride) == O.lib.orc_length(n, shift, stride)
    for a in range(-20, 20):
        for b in range(1, 9):
            assert lib.hyd_mod(a, b) == O.lib.orc_mod(a, b) == a % b
            assert lib.hyd_shift(a % b, (a * 7) % b, b) == O.lib.orc_shift(a % b, (a * 7) % b, b)


@pytest.mark.parametrize("r,c", [(1, 1), (2, 2), (2, 4), (4, 2), (3, 2)])
def test_local_shapes_match_oracle(r, c):
    og = O.Grid(r, c, 1)
    rng = np.random.default_rng(r + 10 * c)
    for (cd, rd) in O.PAIRS:
        for (m, n) in [(0, 0), (1, 1), (37, 23), (129, 131)]:
            ca = int(rng.integers(0, O.lib.orc_dist_stride(og, cd)))
            ra = int(rng.integers(0, O.lib.orc_dist_stride(og, rd)))
            A = O.DM(og, m, n, cd, rd, ca, ra)
            d = hyd_dm(None, 1, m, n, cd, rd, ca, ra, 0, 0)
            for p in range(r * c):
                cs, rs, lh, lw = (C.c_int64() for _ in range(4))
                assert lib.hyd_local_shape_of(r, c, 1, p, C.byref(d), C.byref(cs), C.byref(rs), C.byref(lh), C.byref(lw)) == 0
                assert (lh.value, lw.value) == (A.lh(p), A.lw(p))


def test_errors_are_reported_not_thrown():
    d = hyd_dm(None, 1, 4, 4, O.MC, O.MC, 0, 0, 0, 0)  # [MC,MC] is not a legal pair
    x = C.c_int64()
    assert lib.hyd_local_shape_of(2, 2, 1, 0, C.byref(d), C.byref(x), C.byref(x), C.byref(x), C.byref(x)) != 0
    assert b"unsupported distribution" in lib.hyd_last_error()
    d = hyd_dm(None, 1, 4, 4, O.MC, O.MR, 5, 0, 0, 0)
    assert lib.hyd_local_shape_of(2, 2, 1, 0, C.byref(d), C.byref(x), C.byref(x), C.byref(x), C.byref(x)) != 0
    assert b"alignment" in lib.hyd_last_error()


@pytest.mark.parametrize("r,c", [(1, 2), (2, 2), (2, 4), (3, 2)])
def test_redist_plan_is_a_consistent_permutation(r, c):
    """Planner invariants for every pair: what p sends to q is what q receives from p, and every rank receives
    exactly its local block (bit-exact redistribution is a permutation of elements)."""
    P = r * c
    og = O.Grid(r, c, 1)
    rng = np.random.default_rng(3)
    for tr in (0, 1):
        for (scd, srd), (dcd, drd) in itertools.product(O.PAIRS, O.PAIRS):
            m, n = 61, 47
            sa = [int(rng.integers(0, O.lib.orc_dist_stride(og, x))) for x in (scd, srd, dcd, drd)]
            s = hyd_dm(None, 1, m, n, scd, srd, sa[0], sa[1], 0, 0)
            dm, dn = (n, m) if tr else (m, n)
            d = hyd_dm(None, 1, dm, dn, dcd, drd, sa[2], sa[3], 0, 0)
            send = np.zeros((P, P), dtype=np.int64)
            recv = np.zeros((P, P), dtype=np.int64)
            for p in range(P):
                assert lib.hyd_redist_plan(r, c, 1, p, C.byref(s), C.byref(d), tr, 8, send[p].ctypes.data_as(C.POINTER(C.c_int64)),
                                           recv[p].ctypes.data_as(C.POINTER(C.c_int64))) == 0
            assert np.array_equal(send, recv.T)
            D = O.DM(og, dm, dn, dcd, drd, sa[2], sa[3])
            for p in range(P):
                assert recv[p].sum() == D.lh(p) * D.lw(p) * 8


def test_default_grid_height_matches_the_reference_rule():
    """Grid::DefaultHeight (src/core/Grid.cpp:58-64): floor(sqrt(p)) upwards until it divides p -- NOT the largest divisor
    below sqrt(p): 8 -> 2 (2x4), 4 -> 2, but 10 -> 5 (5x2) and a prime p >= 5 -> p (p x 1).  Product (host function of the C ABI)
    against the oracle's restatement and the known answers."""
    import oracle_lib as O
    from elemental_b200 import lib
    known = {1: 1, 2: 1, 3: 1, 4: 2, 5: 5, 6: 2, 7: 7, 8: 2, 9: 3, 10: 5, 12: 3, 16: 4, 24: 4, 64: 8}
    for p, h in known.items():
        assert lib.hyd_grid_default_height(p) == h == O.lib.orc_grid_default_height(p), p
    for p in range(1, 2000):
        assert lib.hyd_grid_default_height(p) == O.lib.orc_grid_default_height(p), p
        assert p % lib.hyd_grid_default_height(p) == 0


@pytest.mark.parametrize("r,c", [(1, 1), (2, 2), (2, 4), (4, 2), (3, 2), (1, 8)])
def test_trrk_row_cuts_cover_exactly_the_triangle(r, c):
    """The trapezoidal local update of hyd_trrk_* / hyd_syrk_* (csrc/summa.cu local_trrk) splits each column block of the local
    C into rows updated directly, rows updated under a mask, and rows left alone.  For random sizes, alignments and column
    blocks: direct rows are inside the triangle for every column, untouched rows outside for every column, the two ranges
    are disjoint and ordered, and cuts between them fall on 16-byte row multiples."""
    rng = np.random.default_rng(r * 31 + c)
    P = r * c
    for _ in range(60):
        n = int(rng.integers(1, 700))
        ca, ra = int(rng.integers(0, r)), int(rng.integers(0, c))
        d = hyd_dm(None, 1, n, n, O.MC, O.MR, ca, ra, 0, 0)
        for p in range(P):
            cs, rs, lh, lw = (C.c_int64() for _ in range(4))
            assert lib.hyd_local_shape_of(r, c, 1, p, C.byref(d), C.byref(cs), C.byref(rs), C.byref(lh), C.byref(lw)) == 0
            if lw.value == 0:
                continue
            gi = cs.value + np.arange(lh.value) * r
            for es, uplo in ((8, 0), (8, 1), (4, 0), (4, 1)):
                q = 16 // es
                j0 = int(rng.integers(0, lw.value))
                w = int(rng.integers(1, lw.value - j0 + 1))
                cuts = (C.c_int64 * 4)()
                assert lib.hyd_trrk_plan(r, c, 1, p, C.byref(d), uplo, j0, w, es, cuts) == 0
                d0, d1, t0, t1 = cuts
                gj = rs.value + (j0 + np.arange(w)) * c
                inside = (gi[:, None] >= gj[None, :]) if uplo == 0 else (gi[:, None] <= gj[None, :])
                assert 0 <= d0 <= d1 <= lh.value and 0 <= t0 <= t1 <= lh.value
                assert d1 <= t0 or t1 <= d0
                touched = np.zeros(lh.value, bool)
                touched[d0:d1] = True
                assert inside[d0:d1].all()
                touched[t0:t1] = True
                assert not inside[~touched].any()
                boundary = t1 if uplo == 0 else d1          # the cut the two ranges share
                assert boundary % q == 0 or boundary == lh.value
                assert (t0 if uplo == 0 else d0) % q == 0   # first updated row: keeps the operand pointers 16-byte aligned
                # the strip is never wider than the workspace local_trrk reserves
                assert t1 - t0 <= min(lh.value, ((w - 1) * c) // r + 2 + 2 * q)
    cuts = (C.c_int64 * 4)()
    d = hyd_dm(None, 1, 8, 8, O.MC, O.MR, 0, 0, 0, 0)
    assert lib.hyd_trrk_plan(1, 1, 1, 0, C.byref(d), 2, 0, 4, 8, cuts) != 0
    assert lib.hyd_trrk_plan(1, 1, 1, 0, C.byref(d), 0, 6, 4, 8, cuts) != 0


def test_proxy_alignment_adoption_matches_the_oracle_rule():
    """hyd_adopted_align (used by the library's read proxies) against oracle_core.c's orc_copy_adopts, itself exercised by the
    reference's own SUMMA loops in tests/test_ref_pin.py: every distribution pair, several grids and alignments."""
    for r, c in [(1, 1), (2, 2), (2, 3), (4, 2), (1, 8)]:
        og = O.Grid(r, c, 1)
        for src in range(7):
            for dst in range(7):
                for a in range(max(1, O.lib.orc_dist_stride(og, src))):
                    out = C.c_int(0)
                    k = O.lib.orc_copy_adopts(src, a, dst, O.lib.orc_dist_stride(og, dst), C.byref(out))
                    assert lib.hyd_adopted_align(r, c, src, a, dst) == (out.value if k else 0), (r, c, src, dst, a)


def test_abi_header_is_plain_c():
    """The drop-in boundary is a C ABI: include/hydrogen_b200.h must compile as C99 on its own (what a cgo / JNI / ctypes binding
    generator sees), with no C++ or CUDA types in any signature."""
    import subprocess
    hdr = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "hydrogen_b200.h")
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-fsyntax-only", "-x", "c", hdr])


def test_host_block_plan():
    """hyd_host_block_plan (no GPU work): the column blocks of the host-resident Gemm.  Cuts cover [0, tn] in order; every
    block's download hides under the next block's GEMMs (lambda*w_j <= waves(w_{j+1})), the last block is small (its
    download is the only exposed one) and blocking costs at most a wave or two against the unblocked run."""
    import ctypes as C
    from elemental_b200 import lib
    cuts = (C.c_int64 * 9)()
    for (tm, tn, sms, lam) in ((128, 128, 148, 0.1736), (64, 32, 148, 0.087), (128, 128, 148, 0.49), (128, 256, 148, 0.17), (64, 64, 144, 0.087)):
        nb = lib.hyd_host_block_plan(tm, tn, sms, lam, 0.05, 8, cuts)
        got = list(cuts[: nb + 1])
        w = [y - x for x, y in zip(got, got[1:])]
        waves = [-(-tm * x // sms) for x in w]
        assert 2 <= nb <= 8 and got[0] == 0 and got[-1] == tn and all(x > 0 for x in w), got
        assert all(lam * w[j] <= waves[j + 1] + 1e-9 for j in range(nb - 1)), (got, waves)
        assert sum(waves) <= -(-tm * tn // sms) + 2, (got, sum(waves))
        assert w[-1] * 3 <= tn, got
    assert lib.hyd_host_block_plan(64, 1, 148, 0.1, 0.05, 8, cuts) == 1 and list(cuts[:2]) == [0, 1]
    assert lib.hyd_host_block_plan(64, 32, 148, 0.1, 0.05, 1, cuts) == 1 and list(cuts[:2]) == [0, 32]
    assert lib.hyd_host_block_plan(3, 4, 148, 0.1, 0.05, 8, cuts) == 1     # one wave either way: nothing to gain


def test_static_scratch_trrk_rejects_bad_arguments_before_touching_the_gpu():
    """hyd_local_trrk_static_f64 (the allocation-free entry point): argument errors and a diagonal strip beyond the static
    scratch are reported before any CUDA call, so they can be checked without a GPU (dummy non-null pointers are never
    dereferenced).  The computation itself is tested on a B200 in tests/test_zz_no_alloc.py."""
    import ctypes as C
    from elemental_b200 import lib
    dummy = C.c_void_p(4096)

    def call(uplo, lh, lw, k, cs, cst, rs, rst, lda=None, ldb=None, ldc=None):
        return lib.hyd_local_trrk_static_f64(uplo, 0, 0, lh, lw, k, cs, cst, rs, rst, C.c_double(1.0), dummy, lda or max(lh, 1), dummy,
                                             ldb or max(k, 1), C.c_double(1.0), dummy, ldc or max(lh, 1), None)
    assert call(2, 8, 8, 8, 0, 1, 0, 1) != 0 and b"uplo" in lib.hyd_last_error()
    assert call(0, -1, 8, 8, 0, 1, 0, 1) != 0 and b"negative" in lib.hyd_last_error()
    assert call(0, 8, 8, 8, 2, 2, 0, 1) != 0 and b"shift" in lib.hyd_last_error()
    assert call(0, 8, 8, 8, 0, 1, 0, 0) != 0 and b"stride" in lib.hyd_last_error()
    assert call(0, 8, 8, 8, 0, 1, 0, 1, ldc=4) != 0 and b"ldc" in lib.hyd_last_error()
    assert call(0, 8, 8, 8, 0, 1, 0, 1, lda=4) != 0 and b"leading dimension" in lib.hyd_last_error()
    # 512 local columns, 64 global columns apart, unit row stride: the strip is ~32 k rows x 512 columns of doubles
    assert call(0, 40000, 512, 8, 0, 1, 0, 64) != 0
    msg = lib.hyd_last_error().decode()
    assert "beyond the static scratch" in msg and "hyd_trrk_f64" in msg and str(4 * 1024 * 1024) in msg
    assert call(0, 0, 8, 8, 0, 1, 0, 1) == 0 and call(1, 8, 0, 8, 0, 1, 0, 1) == 0       # empty local block: nothing to do


def test_multisync_and_world_env_fail_loudly_on_bad_input():
    import ctypes as C
    from elemental_b200 import lib
    assert lib.hyd_multisync_begin(None, None, 2) != 0 and b"bad arguments" in lib.hyd_last_error()
    assert lib.hyd_multisync_end(None, None, -1) != 0
    assert lib.hyd_multisync_begin(None, None, 0) == 0 and lib.hyd_multisync_end(None, None, 0) == 0   # nothing to fence
    assert lib.hyd_world_create_env(None) != 0
    import torch
    if not torch.cuda.is_available():       # no device: the product must say so, not fall back to anything
        w = C.c_void_p()
        env_keep = {k: os.environ.get(k) for k in ("RANK", "WORLD_SIZE")}
        os.environ["RANK"], os.environ["WORLD_SIZE"] = "3", "2"
        try:
            assert lib.hyd_world_create_env(C.byref(w)) != 0 and b"rank 3 of 2" in lib.hyd_last_error()
        finally:
            for k, v in env_keep.items():
                if v is None:
                    os.environ.pop(k, None)
                else:
                    os.environ[k] = v
        assert lib.hyd_world_create_env(C.byref(w)) != 0 and b"CUDA" in lib.hyd_last_error()

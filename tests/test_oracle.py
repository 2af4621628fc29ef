"""Pins the CPU oracle (oracle/*.c) with the reference's own test PROPERTIES (no golden files exist):

* tests/core/DistMatrix.cpp:28-76 -- for every distribution pair and random alignments, B = A, gather both
  to [*,*], require bitwise equality.  Here: every restated primitive chain must reproduce the
  closed-form ownership definition bit for bit, on several grids and ragged sizes.
* tests/blas_like/BasicGemm.cpp:93-94,108 -- alpha=2, beta=3, |x-y| <= 1e-10 (fp64) / 1e-3 (fp32).
"""
import itertools

import numpy as np
import pytest

import oracle_lib as O
from oracle_lib import MC, MR, VC, VR, STAR, DM, Grid, lib

GRIDS = [(1, 1), (1, 2), (2, 1), (2, 2), (2, 4), (4, 2), (3, 2)]
SIZES = [(1, 1), (7, 5), (50, 50), (129, 131)]


def rand_align(rng, grid, d):
    s = lib.orc_dist_stride(grid, d)
    return int(rng.integers(0, s))


def test_indexing_known_answers():
    # include/El/core/indexing/impl.hpp:33-36,65-66,229-245
    assert lib.orc_mod(-1, 4) == 3 and lib.orc_mod(5, 4) == 1 and lib.orc_mod(-8, 4) == 0
    assert lib.orc_shift(0, 1, 4) == 3 and lib.orc_shift(2, 2, 4) == 0
    assert lib.orc_length(10, 0, 4) == 3 and lib.orc_length(10, 1, 4) == 3 and lib.orc_length(10, 2, 4) == 2
    assert lib.orc_length(2, 3, 4) == 0 and lib.orc_length(0, 0, 1) == 0
    assert lib.orc_max_length(10, 4) == 3 and lib.orc_max_length(0, 4) == 0
    for n in range(0, 40):
        for s in range(1, 7):
            assert sum(lib.orc_length(n, sh, s) for sh in range(s)) == n


def test_grid_rank_maps():
    # src/core/Grid.cpp:147-148: vcRank = mc + r*mr, vrRank = mr + c*mc; column-major owning rank == vcRank
    g = Grid(2, 4, 1)
    for mc in range(2):
        for mr in range(4):
            assert lib.orc_grid_rank_of(g, mc, mr) == mc + 2 * mr
            assert lib.orc_dist_rank(g, VC, mc, mr) == mc + 2 * mr
            assert lib.orc_dist_rank(g, VR, mc, mr) == mr + 4 * mc
    assert [lib.orc_dist_stride(g, d) for d in (MC, MR, VC, VR, STAR)] == [2, 4, 8, 8, 1]


@pytest.mark.parametrize("r,c", GRIDS)
def test_fill_and_gather_roundtrip(r, c):
    g = Grid(r, c, 1)
    rng = np.random.default_rng(r * 10 + c)
    for (cd, rd) in O.PAIRS:
        A = DM(g, 37, 23, cd, rd, rand_align(rng, g, cd), rand_align(rng, g, rd)).fill(7)
        G, mism = A.to_global()
        assert mism == 0
        exp = np.array([[lib.orc_uniform(7, i + j * 37) for j in range(23)] for i in range(37)])
        assert np.array_equal(G, exp)
        assert sum(A.lh(p) * A.lw(p) for p in range(A.P)) >= 37 * 23


CHAINS = [
    # (primitive, src pair, dst pair)   -- the pairs SUMMA drives
    ("orc_row_allgather", (MC, MR), (MC, STAR)),
    ("orc_row_allgather", (MR, MC), (MR, STAR)),
    ("orc_row_allgather", (STAR, VR), (STAR, STAR)),
    ("orc_col_allgather", (MC, MR), (STAR, MR)),
    ("orc_col_allgather", (MR, MC), (STAR, MC)),
    ("orc_col_allgather", (VC, STAR), (STAR, STAR)),
    ("orc_col_alltoall_demote", (MC, MR), (VC, STAR)),
    ("orc_col_alltoall_demote", (MR, MC), (VR, STAR)),
    ("orc_row_alltoall_demote", (MC, MR), (STAR, VR)),
    ("orc_row_alltoall_demote", (MR, MC), (STAR, VC)),
    ("orc_partial_col_allgather", (VC, STAR), (MC, STAR)),
    ("orc_partial_col_allgather", (VR, STAR), (MR, STAR)),
    ("orc_partial_row_allgather", (STAR, VC), (STAR, MC)),
    ("orc_partial_row_allgather", (STAR, VR), (STAR, MR)),
    ("orc_vector_exchange", (VC, STAR), (VR, STAR)),
    ("orc_vector_exchange", (VR, STAR), (VC, STAR)),
    ("orc_vector_exchange", (STAR, VC), (STAR, VR)),
    ("orc_vector_exchange", (STAR, VR), (STAR, VC)),
]


@pytest.mark.parametrize("r,c", GRIDS)
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_primitive_chains_match_definition(r, c, dtype):
    """The DistMatrix.cpp property: after B = A (restated pack -> collective -> unpack) the gathered
    matrices are bitwise equal, and each rank's block equals the closed-form ownership definition."""
    g = Grid(r, c, 1)
    rng = np.random.default_rng(1000 * r + c)
    checked = 0
    for (prim, (scd, srd), (dcd, drd)) in CHAINS:
        for (m, n) in SIZES:
            for trial in range(3):
                sca, sra = rand_align(rng, g, scd), rand_align(rng, g, srd)
                # aligned branch: destination adopts the source's alignments (AlignColsAndResize, force=false);
                # RowAllGather / ColAllGather also get genuinely unaligned destinations.
                dca = sca % lib.orc_dist_stride(g, dcd) if dcd == scd or prim.startswith("orc_partial_col") else 0
                dra = sra % lib.orc_dist_stride(g, drd) if drd == srd or prim.startswith("orc_partial_row") else 0
                if prim == "orc_col_alltoall_demote":
                    dca = sca + lib.orc_dist_stride(g, scd) * int(rng.integers(0, lib.orc_dist_stride(g, srd)))
                if prim == "orc_row_alltoall_demote":
                    dra = sra + lib.orc_dist_stride(g, srd) * int(rng.integers(0, lib.orc_dist_stride(g, scd)))
                if prim == "orc_vector_exchange":
                    dca, dra = rand_align(rng, g, dcd), rand_align(rng, g, drd)
                if prim == "orc_row_allgather" and trial == 2:
                    dca = rand_align(rng, g, dcd)
                if prim == "orc_col_allgather" and trial == 2:
                    dra = rand_align(rng, g, drd)
                A = DM(g, m, n, scd, srd, sca, sra, dtype).fill(3 + trial)
                B = DM(g, m, n, dcd, drd, dca, dra, dtype).fill_const(-77)
                Bdef = DM(g, m, n, dcd, drd, dca, dra, dtype).fill_const(-77)
                rc = getattr(lib, prim)(A.h, B.h)
                assert rc == 0, (prim, rc, scd, srd, dcd, drd, sca, sra, dca, dra)
                lib.orc_redist_closed(A.h, Bdef.h, 0)
                GA, _ = A.to_global()
                GB, mism = B.to_global()
                assert mism == 0 and np.array_equal(GA, GB), (prim, r, c, m, n)
                for p in range(B.P):
                    assert np.array_equal(B.block(p), Bdef.block(p))
                checked += 1
    assert checked == len(CHAINS) * len(SIZES) * 3


def test_pack_unpack_are_inverse_permutations():
    """copy::util pack followed by the matching unpack is the identity (util.hpp:148-230,363-533,671-722)."""
    rng = np.random.default_rng(5)
    es = 8
    for (h, w, stride) in [(5, 7, 1), (8, 9, 2), (13, 11, 3), (16, 16, 4), (3, 2, 5)]:
        align = int(rng.integers(0, stride))
        A = np.asfortranarray(rng.standard_normal((h + 2, w)))
        portion = max(1, h * -(-w // stride))
        buf = np.full(stride * portion, np.nan)
        out = np.asfortranarray(np.full((h + 3, w), -5.0))
        lib.orc_row_strided_pack(h, w, align, stride, A.ctypes.data, h + 2, buf.ctypes.data, portion, es)
        lib.orc_row_strided_unpack(h, w, align, stride, buf.ctypes.data, portion, out.ctypes.data, h + 3, es)
        assert np.array_equal(out[:h], A[:h]) and np.all(out[h:] == -5.0)
        portion = max(1, -(-h // stride) * w)
        buf = np.full(stride * portion, np.nan)
        out[:] = -5.0
        lib.orc_col_strided_pack(h, w, align, stride, A.ctypes.data, h + 2, buf.ctypes.data, portion, es)
        lib.orc_col_strided_unpack(h, w, align, stride, buf.ctypes.data, portion, out.ctypes.data, h + 3, es)
        assert np.array_equal(out[:h], A[:h]) and np.all(out[h:] == -5.0)
        cs, rs = stride, max(1, stride - 1)
        ca, ra = align, int(rng.integers(0, rs))
        portion = max(1, -(-h // cs) * -(-w // rs))
        buf = np.full(cs * rs * portion, np.nan)
        out[:] = -5.0
        lib.orc_strided_pack(h, w, ca, cs, ra, rs, A.ctypes.data, h + 2, buf.ctypes.data, portion, es)
        lib.orc_strided_unpack(h, w, ca, cs, ra, rs, buf.ctypes.data, portion, out.ctypes.data, h + 3, es)
        assert np.array_equal(out[:h], A[:h]) and np.all(out[h:] == -5.0)


@pytest.mark.parametrize("r,c", [(1, 1), (2, 2), (2, 3)])
def test_axpy_contract_definition(r, c):
    """AxpyContract (AxpyContract.hpp:483-544): B += alpha * (sum of A over the replicated communicator)."""
    g = Grid(r, c, 1)
    rng = np.random.default_rng(11)
    m, n = 19, 14
    for (scd, srd, tr) in [(MC, STAR, 0), (STAR, MR, 0), (STAR, STAR, 0), (MR, STAR, 1), (STAR, MC, 1), (MR, STAR, 0)]:
        sm, sn = (n, m) if tr else (m, n)
        A = DM(g, sm, sn, scd, srd, rand_align(rng, g, scd), rand_align(rng, g, srd))
        # distinct partial per rank
        for p in range(A.P):
            A.block(p)[...] = rng.standard_normal(A.block(p).shape)
        B = DM(g, m, n, MC, MR, rand_align(rng, g, MC), rand_align(rng, g, MR)).fill(9)
        G0, _ = B.to_global()
        # expected: sum over ranks of each rank's contribution scattered to global
        S = np.zeros((sm, sn))
        cs, rs = lib.orc_dist_stride(g, scd), lib.orc_dist_stride(g, srd)
        for v in range(A.P):
            p = lib.orc_grid_rank_of(g, v % r, v // r)
            blk = A.block(p)
            mc, mr = p % r, p // r
            csh = lib.orc_shift(lib.orc_dist_rank(g, scd, mc, mr), A.ca, cs)
            rsh = lib.orc_shift(lib.orc_dist_rank(g, srd, mc, mr), A.ra, rs)
            S[csh::cs, rsh::rs] += blk
        lib.orc_axpy_contract_closed(2.0, A.h, B.h, tr)
        G1, mism = B.to_global()
        assert mism == 0
        np.testing.assert_allclose(G1, G0 + 2.0 * (S.T if tr else S), rtol=1e-14, atol=1e-14)


@pytest.mark.parametrize("use_blas", [False, True])
@pytest.mark.parametrize("r,c", [(1, 1), (2, 2), (2, 4), (3, 2)])
def test_gemm_all_variants_vs_dense(r, c, use_blas):
    """BasicGemm.cpp:96-236 restated: alpha=2, beta=3, every orientation x algorithm, compare to a dense
    long-double product; bound |x - y| <= 1e-10 (fp64) as BasicGemm.cpp:93-94."""
    if use_blas and not O.use_blas(True):
        pytest.skip("no vendor BLAS found")
    if not use_blas:
        O.use_blas(False)
    g = Grid(r, c, 1)
    rng = np.random.default_rng(r * 7 + c)
    m, n, k = 45, 38, 52
    for oA, oB in itertools.product((0, 1), (0, 1)):
        for alg in (1, 2, 3, 4, 0):
            am, an = (k, m) if oA else (m, k)
            bm, bn = (n, k) if oB else (k, n)
            A = DM(g, am, an, MC, MR, rand_align(rng, g, MC), rand_align(rng, g, MR)).fill(1)
            B = DM(g, bm, bn, MC, MR, rand_align(rng, g, MC), rand_align(rng, g, MR)).fill(2)
            Cm = DM(g, m, n, MC, MR, rand_align(rng, g, MC), rand_align(rng, g, MR)).fill(3)
            GA, _ = A.to_global(); GB, _ = B.to_global(); GC, _ = Cm.to_global()
            exp = np.asfortranarray(GC.copy())
            lib.orc_dense_gemm_ref(oA, oB, m, n, k, 2.0, GA.ctypes.data, am, GB.ctypes.data, bm, 3.0, exp.ctypes.data, m, 8)
            O.gemm(oA, oB, alg, 2.0, A, B, 3.0, Cm, 16)
            got, mism = Cm.to_global()
            assert mism == 0
            assert np.max(np.abs(got - exp)) <= 1e-10, (oA, oB, alg, np.max(np.abs(got - exp)))
    O.use_blas(False)


def test_gemm_fp32_and_dot_operands():
    """fp32 bound 1e-3 (BasicGemm.cpp:93) and the LBANN-style [*,VC] operands of config 5 (NT-Dot)."""
    g = Grid(2, 2, 1)
    m, n, k = 40, 24, 300
    A = DM(g, m, k, STAR, VC, 0, 1, np.float32).fill(1)
    B = DM(g, n, k, STAR, VC, 0, 1, np.float32).fill(2)
    Cm = DM(g, m, n, MC, MR, 1, 0, np.float32).fill(3)
    GA, _ = A.to_global(); GB, _ = B.to_global(); GC, _ = Cm.to_global()
    exp = np.asfortranarray(GC.copy())
    lib.orc_dense_gemm_ref(0, 1, m, n, k, 2.0, GA.ctypes.data, m, GB.ctypes.data, n, 3.0, exp.ctypes.data, m, 4)
    O.gemm(0, 1, 0, 2.0, A, B, 3.0, Cm, 16)   # GEMM_DEFAULT -> Dot since 10*m <= k and 10*n <= k is false here? use explicit
    got, _ = Cm.to_global()
    assert np.max(np.abs(got - exp)) <= 1e-3
    Cm2 = DM(g, m, n, MC, MR, 1, 0, np.float32).fill(3)
    O.gemm(0, 1, 4, 2.0, A, B, 3.0, Cm2, 16)
    got2, _ = Cm2.to_global()
    assert np.max(np.abs(got2 - exp)) <= 1e-3


# ---------------------------------------------------------------------------------------------------------------
# distributed BLAS-1 around Gemm (oracle/oracle_level1.c): Axpy / TransposeAxpy / Hadamard / Scale
# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("r,c", [(1, 1), (2, 2), (2, 4), (3, 2)])
def test_level1_axpy_matches_dense_definition(r, c):
    """El::Axpy / TransposeAxpy between any two distributions (Axpy.hpp:161-187, TransposeAxpy.hpp:181-210) is
    Y_global += alpha X_global(^T): checked after gathering, replicas included, for integer data (exact arithmetic)."""
    og = O.Grid(r, c, 1)
    rng = np.random.default_rng(17 * r + c)
    for (xcd, xrd), (ycd, yrd) in itertools.product(O.PAIRS, [(MC, MR), (STAR, VC), (MR, STAR), (STAR, STAR)]):
        for transpose in (0, 1):
            m, n = (37, 22)
            al = lambda d: int(rng.integers(0, O.lib.orc_dist_stride(og, d)))
            X = O.DM(og, n if transpose else m, m if transpose else n, xcd, xrd, al(xcd), al(xrd)).fill_index()
            Y = O.DM(og, m, n, ycd, yrd, al(ycd), al(yrd)).fill_index()
            Y0, _ = Y.to_global()
            Y0 = Y0.copy()
            assert O.lib.orc_axpy(2.0, X.h, Y.h, transpose) == 0
            Xg, _ = X.to_global()
            Yg, mism = Y.to_global()
            assert mism == 0
            assert np.array_equal(Yg, Y0 + 2.0 * (Xg.T if transpose else Xg))
    A = O.DM(og, 5, 4, MC, MR)
    assert O.lib.orc_axpy(1.0, A.h, O.DM(og, 4, 5, MC, MR).h, 0) != 0   # nonconformal


def test_level1_hadamard_and_scale():
    og = O.Grid(2, 2, 1)
    A = O.DM(og, 33, 21, MC, MR, 1, 0).fill(1)
    B = O.DM(og, 33, 21, MC, MR, 1, 0).fill(2)
    Cm = O.DM(og, 33, 21, MC, MR, 1, 0)
    assert O.lib.orc_hadamard(A.h, B.h, Cm.h) == 0
    assert np.array_equal(Cm.to_global()[0], A.to_global()[0] * B.to_global()[0])
    assert O.lib.orc_hadamard(A.h, O.DM(og, 33, 21, MC, MR, 0, 0).h, Cm.h) == 3      # "A and B must be aligned"
    assert O.lib.orc_hadamard(A.h, O.DM(og, 33, 21, MC, STAR, 1, 0).h, Cm.h) == 2    # "must share the same distribution"
    assert O.lib.orc_hadamard(A.h, O.DM(og, 30, 21, MC, MR, 1, 0).h, Cm.h) == 1      # "requires equal dimensions"
    A0 = A.to_global()[0].copy()
    O.lib.orc_scale(0.3, A.h)
    assert np.array_equal(A.to_global()[0], 0.3 * A0)
    O.lib.orc_scale(0.0, A.h)
    assert not A.to_global()[0].any()


# ---------------------------------------------------------------------------------------------------------------
# Trrk / Syrk and the trapezoid routines (oracle/oracle_level3.c)
# ---------------------------------------------------------------------------------------------------------------
def _tri_mask(n, uplo, offset=0, m=None):
    m = n if m is None else m
    i, j = np.meshgrid(np.arange(m), np.arange(n), indexing="ij")
    return (j - i <= offset) if uplo == 0 else (j - i >= offset)


@pytest.mark.parametrize("r,c", [(1, 1), (2, 2), (2, 4), (3, 2)])
def test_trapezoid_routines_match_dense_definition(r, c):
    """ScaleTrapezoid / AxpyTrapezoid / MakeTrapezoidal with offsets, on rectangular matrices in several distributions
    and random alignments: LOWER holds j - i <= offset, UPPER j - i >= offset (ScaleTrapezoid.hpp:22-42)."""
    g = Grid(r, c, 1)
    rng = np.random.default_rng(100 + r * 7 + c)
    for (m, n), (cd, rd), uplo, offset in itertools.product([(37, 37), (23, 41), (50, 19)], [(MC, MR), (VC, STAR), (STAR, VR), (MR, MC)],
                                                            (0, 1), (-3, 0, 2)):
        ca, ra = rand_align(rng, g, cd), rand_align(rng, g, rd)
        mask = _tri_mask(n, uplo, offset, m)
        A = DM(g, m, n, cd, rd, ca, ra).fill(5)
        G0, _ = A.to_global()
        lib.orc_scale_trapezoid(-1.5, uplo, A.h, offset)
        G1, mism = A.to_global()
        assert mism == 0 and np.array_equal(G1, np.where(mask, G0 * -1.5, G0))
        X = DM(g, m, n, cd, rd, ca, ra).fill(6)
        GX, _ = X.to_global()
        assert lib.orc_axpy_trapezoid(uplo, 0.75, X.h, A.h, offset) == 0
        G2, mism = A.to_global()
        assert mism == 0 and np.array_equal(G2, np.where(mask, G1 + 0.75 * GX, G1))
        lib.orc_make_trapezoidal(uplo, A.h, offset)
        G3, mism = A.to_global()
        assert mism == 0 and np.array_equal(G3, np.where(mask, G2, 0.0))
    Y = DM(g, 9, 9, MC, MR, 0, 0)
    assert lib.orc_axpy_trapezoid(0, 1.0, DM(g, 9, 8, MC, MR, 0, 0).h, Y.h, 0) == -1


@pytest.mark.parametrize("use_blas", [False, True])
@pytest.mark.parametrize("r,c", [(1, 1), (2, 2), (2, 4), (3, 2)])
def test_trrk_and_syrk_vs_dense(r, c, use_blas):
    """tests/blas_like/Syrk.cpp restated (alpha, beta, alignments, LocalTrrkBlocksize) with a stronger check than its printed
    associativity residual: the `uplo` triangle equals the dense long-double product to 1e-10, and every entry on the other
    side of the diagonal keeps its bits."""
    if use_blas and not O.use_blas(True):
        pytest.skip("no vendor BLAS found")
    if not use_blas:
        O.use_blas(False)
    g = Grid(r, c, 1)
    rng = np.random.default_rng(200 + r * 7 + c)
    n, k = 150, 45
    for nb_local in (8, 64):
        lib.orc_set_local_trrk_blocksize(nb_local)
        for uplo, oA, oB in itertools.product((0, 1), (0, 1), (0, 1)):
            am, an = (k, n) if oA else (n, k)
            bm, bn = (n, k) if oB else (k, n)
            A = DM(g, am, an, MC, MR, rand_align(rng, g, MC), rand_align(rng, g, MR)).fill(1)
            B = DM(g, bm, bn, MC, MR, rand_align(rng, g, MC), rand_align(rng, g, MR)).fill(2)
            Cm = DM(g, n, n, MC, MR, rand_align(rng, g, MC), rand_align(rng, g, MR)).fill(3)
            GA, _ = A.to_global(); GB, _ = B.to_global(); GC, _ = Cm.to_global()
            exp = np.asfortranarray(GC.copy())
            lib.orc_dense_gemm_ref(oA, oB, n, n, k, 2.0, GA.ctypes.data, am, GB.ctypes.data, bm, 3.0, exp.ctypes.data, n, 8)
            assert lib.orc_trrk(uplo, oA, oB, 2.0, A.h, B.h, 3.0, Cm.h, 16) == 0
            got, mism = Cm.to_global()
            mask = _tri_mask(n, uplo)
            assert mism == 0
            assert np.max(np.abs(got - exp)[mask]) <= 1e-10, (uplo, oA, oB)
            assert np.array_equal(got[~mask], GC[~mask]), (uplo, oA, oB)
        for uplo, orient in itertools.product((0, 1), (0, 1)):
            am, an = (k, n) if orient else (n, k)
            A = DM(g, am, an, MC, MR, rand_align(rng, g, MC), rand_align(rng, g, MR)).fill(1)
            Cm = DM(g, n, n, MC, MR, rand_align(rng, g, MC), rand_align(rng, g, MR)).fill(3)
            GA, _ = A.to_global(); GC, _ = Cm.to_global()
            exp = np.asfortranarray(GC.copy())
            lib.orc_dense_gemm_ref(orient, 1 - orient, n, n, k, 2.0, GA.ctypes.data, am, GA.ctypes.data, am, 3.0, exp.ctypes.data, n, 8)
            assert lib.orc_syrk(uplo, orient, 2.0, A.h, 3.0, Cm.h, 16) == 0
            got, mism = Cm.to_global()
            mask = _tri_mask(n, uplo)
            assert mism == 0
            assert np.max(np.abs(got - exp)[mask]) <= 1e-10, (uplo, orient)
            assert np.array_equal(got[~mask], GC[~mask]), (uplo, orient)
    lib.orc_set_local_trrk_blocksize(64)
    O.use_blas(False)


def test_trrk_operand_proxies_and_fp32():
    """Operands and C in other distributions go through the [MC,MR] proxies (Trrk/NN.hpp:36-41); fp32 bound 1e-3."""
    g = Grid(2, 2, 1)
    n, k = 64, 30
    A = DM(g, n, k, VC, STAR, 1, 0, np.float32).fill(1)
    B = DM(g, k, n, STAR, VR, 0, 2, np.float32).fill(2)
    Cm = DM(g, n, n, MR, MC, 1, 1, np.float32).fill(3)
    GA, _ = A.to_global(); GB, _ = B.to_global(); GC, _ = Cm.to_global()
    exp = np.asfortranarray(GC.copy())
    lib.orc_dense_gemm_ref(0, 0, n, n, k, 2.0, GA.ctypes.data, n, GB.ctypes.data, k, 3.0, exp.ctypes.data, n, 4)
    assert lib.orc_trrk(1, 0, 0, 2.0, A.h, B.h, 3.0, Cm.h, 8) == 0
    got, mism = Cm.to_global()
    mask = _tri_mask(n, 1)
    assert mism == 0 and np.max(np.abs(got - exp)[mask]) <= 1e-3 and np.array_equal(got[~mask], GC[~mask])
    assert lib.orc_trrk(0, 0, 0, 1.0, A.h, B.h, 1.0, DM(g, n, n + 1, MC, MR, 0, 0, np.float32).h, 8) == -1


@pytest.mark.parametrize("r,c", [(1, 1), (2, 2), (2, 4), (3, 2)])
def test_make_symmetric_matches_the_local_definition(r, c):
    """The DistMatrix form (MakeTrapezoidal + Transpose + AxpyTrapezoid, MakeSymmetric.hpp:52-68) against the Matrix form
    (:14-50: copy the triangle across the diagonal), bit for bit."""
    g = Grid(r, c, 1)
    rng = np.random.default_rng(300 + r * 7 + c)
    for n, (cd, rd), uplo in itertools.product((1, 37, 130), [(MC, MR), (VC, STAR), (STAR, VR), (MR, MC), (STAR, STAR)], (0, 1)):
        A = DM(g, n, n, cd, rd, rand_align(rng, g, cd), rand_align(rng, g, rd)).fill(7)
        G0, _ = A.to_global()
        assert lib.orc_make_symmetric(uplo, A.h) == 0
        G1, mism = A.to_global()
        T = np.tril(G0) if uplo == 0 else np.triu(G0)
        exp = T + T.T - np.diag(np.diag(G0))
        assert mism == 0 and np.array_equal(G1, exp)
    assert lib.orc_make_symmetric(0, DM(g, 4, 5, MC, MR, 0, 0).h) == -1


@pytest.mark.parametrize("r,c", [(1, 1), (2, 2), (2, 4), (3, 2)])
def test_symm_vs_dense(r, c):
    """Symm (stationary-C variants, Symm/{LL,LU,RL,RU}.hpp) against the dense product with the symmetric completion of the
    `uplo` triangle of A: the other triangle of A is garbage and must not influence C.  Bound 1e-10 as for Gemm."""
    g = Grid(r, c, 1)
    rng = np.random.default_rng(400 + r * 7 + c)
    m, n = 57, 43
    for side, uplo in itertools.product((0, 1), (0, 1)):
        na = m if side == 0 else n
        A = DM(g, na, na, MC, MR, rand_align(rng, g, MC), rand_align(rng, g, MR)).fill(1)
        B = DM(g, m, n, MC, MR, rand_align(rng, g, MC), rand_align(rng, g, MR)).fill(2)
        Cm = DM(g, m, n, MC, MR, rand_align(rng, g, MC), rand_align(rng, g, MR)).fill(3)
        GA, _ = A.to_global(); GB, _ = B.to_global(); GC, _ = Cm.to_global()
        T = np.tril(GA) if uplo == 0 else np.triu(GA)
        S = np.asfortranarray(T + T.T - np.diag(np.diag(GA)))
        exp = np.asfortranarray(GC.copy())
        if side == 0:
            lib.orc_dense_gemm_ref(0, 0, m, n, m, 2.0, S.ctypes.data, m, GB.ctypes.data, m, 3.0, exp.ctypes.data, m, 8)
        else:
            lib.orc_dense_gemm_ref(0, 0, m, n, n, 2.0, GB.ctypes.data, m, S.ctypes.data, n, 3.0, exp.ctypes.data, m, 8)
        assert lib.orc_symm(side, uplo, 2.0, A.h, B.h, 3.0, Cm.h, 16) == 0
        got, mism = Cm.to_global()
        assert mism == 0 and np.max(np.abs(got - exp)) <= 1e-10, (side, uplo, np.max(np.abs(got - exp)))
        GA2, _ = A.to_global()
        assert np.array_equal(GA, GA2)
    assert lib.orc_symm(0, 0, 1.0, DM(g, 5, 5, MC, MR, 0, 0).h, DM(g, 6, 4, MC, MR, 0, 0).h, 0.0, DM(g, 6, 4, MC, MR, 0, 0).h, 8) == -1

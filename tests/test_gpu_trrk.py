"""-m gpu: Trrk / Syrk / Herk and the trapezoid level-1 routines through the C ABI (hyd_trrk_*, hyd_syrk_*,
hyd_{scale,axpy}_trapezoid_dist_*, hyd_make_trapezoidal_dist_*) against the oracle's restatement of the reference loops
(oracle/oracle_level3.c: Trrk.cpp, Trrk/{Local,NN,NT,TN,TT}.hpp, Syrk/{LN,LT,UN,UT}.hpp, ScaleTrapezoid.hpp, ...).

Modelled on tests/blas_like/Syrk.cpp (alpha, beta, random alignments, non-contiguous C) with a stronger check than its
printed associativity residual: inside the triangle the Gemm tolerances of test_gpu_gemm.py (relative Frobenius error
<= 1e-13*sqrt(k) fp64 / 1e-5*sqrt(k) fp32; |x-y| <= 1e-10 / 1e-3), outside it every entry keeps its bits.  The trapezoid
routines are bit-exact (one multiply / one add per entry; alpha chosen so that no FMA contraction can change the result).

SURVEY 8(f) N2.  Green on a B200 since round 2's first GPU call (profiles/r02_first_call_summary.json)."""
import itertools
import os

import numpy as np
import pytest

import oracle_lib as O
from oracle_lib import MC, MR, VC, VR, STAR
from elemental_b200 import host as H
import gpu_util as U

pytestmark = pytest.mark.gpu


def _mask_dm(og, n, d, ca, ra, uplo, dtype, offset=0, m=None):
    m = n if m is None else m
    i, j = np.meshgrid(np.arange(m), np.arange(n), indexing="ij")
    G = ((j - i <= offset) if uplo == 0 else (j - i >= offset)).astype(dtype)
    return O.DM(og, m, n, d[0], d[1], ca, ra, dtype).from_global(G)


def _case(r, c, dtype, kind, uplo, oA, oB, n, k, nb, rng, dists=None, alpha=2.0, beta=3.0):
    og = O.Grid(r, c, 1)
    dA, dB, dC = dists or ((MC, MR), (MC, MR), (MC, MR))
    mk = lambda h, w, d, seed: O.DM(og, h, w, d[0], d[1], U.rand_align(rng, og, d[0]), U.rand_align(rng, og, d[1]), dtype).fill(seed)
    am, an = (k, n) if oA else (n, k)
    A = mk(am, an, dA, 1)
    B = None
    if kind == "trrk":
        bm, bn = (n, k) if oB else (k, n)
        B = mk(bm, bn, dB, 2)
    Cm = mk(n, n, dC, 3)
    Cin = O.DM(og, n, n, dC[0], dC[1], Cm.ca, Cm.ra, dtype).fill(3)
    M = _mask_dm(og, n, dC, Cm.ca, Cm.ra, uplo, dtype)
    if kind == "trrk":
        assert O.lib.orc_trrk(uplo, oA, oB, alpha, A.h, B.h, beta, Cm.h, nb) == 0
    else:
        assert O.lib.orc_syrk(uplo, oA, alpha, A.h, beta, Cm.h, nb) == 0
    return (kind, uplo, oA, oB, alpha, beta, nb, k), A, B, Cin, Cm, M


def _run(r, c, cases, dtype):
    def body(grid):
        res = []
        for (kind, uplo, oA, oB, alpha, beta, nb, k), A, B, Cin, Cexp, M in cases:
            a, cm = U.to_rank(grid, A), U.to_rank(grid, Cin, ldim_extra=2)
            if kind == "trrk":
                b = U.to_rank(grid, B)
                H.Trrk(uplo, oA, oB, alpha, a, b, beta, cm, blocksize=nb)
            elif kind == "syrk":
                H.Syrk(uplo, oA, alpha, a, beta, cm, blocksize=nb)
            else:
                H.Herk(uplo, oA, alpha, a, beta, cm, blocksize=nb)
            res.append((U.block_of(cm), U.padding_untouched(cm), U.block_of(a)))
        return res

    out = H.run_ranks(r * c, body, height=r)
    tol_rel = 1e-13 if dtype == np.float64 else 1e-5
    tol_abs = 1e-10 if dtype == np.float64 else 1e-3
    for i, (key, A, B, Cin, Cexp, M) in enumerate(cases):
        k = key[-1]
        num = den = 0.0
        for p in range(r * c):
            got, pad_ok, a_after = out[p][i]
            exp, cin, inside = Cexp.block(p), Cin.block(p), M.block(p) != 0
            assert pad_ok and np.array_equal(a_after, A.block(p))
            assert np.array_equal(got[~inside], cin[~inside]), (key, p)       # the other triangle keeps its bits
            d = (got.astype(np.float64) - exp.astype(np.float64))[inside]
            num += float((d * d).sum())
            den += float((exp.astype(np.float64)[inside] ** 2).sum())
            if d.size:
                assert np.abs(d).max() <= tol_abs, (key, p, np.abs(d).max())
        assert (num / max(den, 1e-300)) ** 0.5 <= tol_rel * max(1.0, k) ** 0.5, key


@pytest.mark.parametrize("r,c", [(1, 1), (1, 2), (2, 2), (2, 4), (3, 2)])
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_trrk_all_orientations(r, c, dtype):
    rng = np.random.default_rng(31 * r + c)
    O.use_blas(True)
    try:
        cases = [_case(r, c, dtype, "trrk", uplo, oA, oB, 157, 75, 32, rng)
                 for uplo, oA, oB in itertools.product((H.LOWER, H.UPPER), (0, 1), (0, 1))]
    finally:
        O.use_blas(False)
    _run(r, c, cases, dtype)


@pytest.mark.parametrize("r,c", [(1, 1), (2, 2), (2, 4), (1, 4)])
def test_syrk_herk_several_column_blocks(r, c):
    """Local widths beyond one 512-column block of the trapezoidal update, both triangles and orientations, fp64 and fp32."""
    rng = np.random.default_rng(7 * r + c)
    n = 1203 if r * c == 1 else 2207
    O.use_blas(True)
    try:
        c64 = [_case(r, c, np.float64, kind, uplo, orient, 0, n, 96, 64, rng)
               for kind, (uplo, orient) in zip(("syrk", "herk", "syrk", "herk"), itertools.product((0, 1), (0, 1)))]
        c32 = [_case(r, c, np.float32, "syrk", uplo, orient, 0, n, 96, 64, rng) for uplo, orient in ((0, 1), (1, 0))]
    finally:
        O.use_blas(False)
    _run(r, c, c64, np.float64)
    _run(r, c, c32, np.float32)


def test_trrk_proxies_degenerate_and_errors():
    r, c = 2, 2
    rng = np.random.default_rng(3)
    cases = [
        _case(r, c, np.float32, "trrk", 1, 0, 0, 64, 30, 8, rng, dists=((VC, STAR), (STAR, VR), (MR, MC))),
        _case(r, c, np.float32, "syrk", 0, 1, 0, 40, 33, 16, rng, dists=((STAR, VC), None, (VR, STAR))),
        _case(r, c, np.float32, "trrk", 0, 0, 1, 1, 5, 4, rng),
        _case(r, c, np.float32, "trrk", 1, 1, 1, 9, 1, 4, rng, alpha=1.0, beta=0.0),
        # k > 10 n: routed through the Dot variant of the Gemm engine (Syrk/LN.hpp:152-159)
        _case(r, c, np.float32, "syrk", 0, 0, 0, 24, 400, 64, rng),
        _case(r, c, np.float32, "syrk", 1, 1, 0, 31, 333, 64, rng, dists=((VC, STAR), None, (MC, MR))),
        _case(r, c, np.float32, "trrk", 1, 1, 0, 24, 400, 64, rng),
    ]
    _run(r, c, cases, np.float32)

    def body(grid):
        a, cm = H.DistMatrix(grid, 10, 12), H.DistMatrix(grid, 10, 9)
        with pytest.raises(H.HydrogenError, match="nonconformal"):
            H.Syrk(H.LOWER, H.NORMAL, 1.0, a, 0.0, cm)
        with pytest.raises(H.HydrogenError, match="uplo"):
            H.Syrk(2, H.NORMAL, 1.0, a, 0.0, H.DistMatrix(grid, 10, 10))
        return True
    assert H.run_ranks(1, body) == [True]


@pytest.mark.parametrize("r,c", [(1, 1), (2, 2), (2, 4), (3, 2)])
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_trapezoid_routines_bit_exact(r, c, dtype):
    og = O.Grid(r, c, 1)
    rng = np.random.default_rng(11 * r + c)
    cases = []
    for (m, n), d, uplo, offset in itertools.product([(300, 300), (123, 541), (777, 65)], [(MC, MR), (VC, STAR), (STAR, VR), (MR, MC)],
                                                     (0, 1), (-3, 0, 70)):
        ca, ra = U.rand_align(rng, og, d[0]), U.rand_align(rng, og, d[1])
        Y0 = O.DM(og, m, n, d[0], d[1], ca, ra, dtype).fill(5)
        X = O.DM(og, m, n, d[0], d[1], ca, ra, dtype).fill(6)
        Y1 = O.DM(og, m, n, d[0], d[1], ca, ra, dtype).fill(5)
        O.lib.orc_scale_trapezoid(-1.5, uplo, Y1.h, offset)
        Y2 = O.DM(og, m, n, d[0], d[1], ca, ra, dtype).from_global(Y1.to_global()[0])
        assert O.lib.orc_axpy_trapezoid(uplo, 0.5, X.h, Y2.h, offset) == 0      # 0.5*x is exact: FMA or not, one rounding
        Y3 = O.DM(og, m, n, d[0], d[1], ca, ra, dtype).from_global(Y2.to_global()[0])
        O.lib.orc_make_trapezoidal(uplo, Y3.h, offset)
        cases.append((uplo, offset, Y0, X, Y1, Y2, Y3))

    def body(grid):
        res = []
        for uplo, offset, Y0, X, Y1, Y2, Y3 in cases:
            y, x = U.to_rank(grid, Y0, ldim_extra=3), U.to_rank(grid, X)
            H.ScaleTrapezoid(-1.5, uplo, y, offset)
            s1 = U.block_of(y)
            H.AxpyTrapezoid(uplo, 0.5, x, y, offset)
            s2 = U.block_of(y)
            H.MakeTrapezoidal(uplo, y, offset)
            res.append((s1, s2, U.block_of(y), U.padding_untouched(y)))
        return res

    out = H.run_ranks(r * c, body, height=r)
    for i, (uplo, offset, Y0, X, Y1, Y2, Y3) in enumerate(cases):
        for p in range(r * c):
            s1, s2, s3, pad_ok = out[p][i]
            assert pad_ok
            assert np.array_equal(s1, Y1.block(p)) and np.array_equal(s2, Y2.block(p)) and np.array_equal(s3, Y3.block(p)), (i, p)


@pytest.mark.parametrize("r,c", [(1, 1), (2, 2), (2, 4), (3, 2)])
def test_make_symmetric_bit_exact(r, c):
    og = O.Grid(r, c, 1)
    rng = np.random.default_rng(13 * r + c)
    cases = []
    for n, d, uplo, dtype in itertools.product((1, 300, 1111), [(MC, MR), (VC, STAR), (STAR, VR), (MR, MC)], (0, 1), (np.float64, np.float32)):
        ca, ra = U.rand_align(rng, og, d[0]), U.rand_align(rng, og, d[1])
        A0 = O.DM(og, n, n, d[0], d[1], ca, ra, dtype).fill(7)
        A1 = O.DM(og, n, n, d[0], d[1], ca, ra, dtype).fill(7)
        assert O.lib.orc_make_symmetric(uplo, A1.h) == 0
        cases.append((uplo, A0, A1))

    def body(grid):
        res = []
        for uplo, A0, A1 in cases:
            a = U.to_rank(grid, A0, ldim_extra=1)
            H.MakeSymmetric(uplo, a)
            res.append((U.block_of(a), U.padding_untouched(a)))
        return res

    out = H.run_ranks(r * c, body, height=r)
    for i, (uplo, A0, A1) in enumerate(cases):
        for p in range(r * c):
            got, pad_ok = out[p][i]
            assert pad_ok and np.array_equal(got, A1.block(p)), (i, p)


@pytest.mark.parametrize("r,c", [(1, 1), (2, 2), (2, 4), (3, 2)])
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_symm_hemm_all_sides(r, c, dtype):
    """Symm / Hemm against the oracle's restated stationary-C loops (Symm/{LL,LU,RL,RU}.hpp): Gemm tolerances; A (whose other
    triangle holds unrelated values) is left untouched."""
    og = O.Grid(r, c, 1)
    rng = np.random.default_rng(19 * r + c)
    m, n, nb = 157, 123, 32
    cases = []
    O.use_blas(True)
    try:
        for side, uplo in itertools.product((H.LEFT, H.RIGHT), (H.LOWER, H.UPPER)):
            na = m if side == H.LEFT else n
            dA = (MC, MR) if (side + uplo) % 2 == 0 else (VC, STAR)          # half of the cases through the [MC,MR] copy of a proxy
            mk = lambda h, w, d, seed: O.DM(og, h, w, d[0], d[1], U.rand_align(rng, og, d[0]), U.rand_align(rng, og, d[1]), dtype).fill(seed)
            A, B, Cm = mk(na, na, dA, 1), mk(m, n, (MC, MR), 2), mk(m, n, (MC, MR), 3)
            Cin = O.DM(og, m, n, MC, MR, Cm.ca, Cm.ra, dtype).fill(3)
            assert O.lib.orc_symm(side, uplo, 2.0, A.h, B.h, 3.0, Cm.h, nb) == 0
            cases.append((side, uplo, A, B, Cin, Cm))
    finally:
        O.use_blas(False)

    def body(grid):
        res = []
        for i, (side, uplo, A, B, Cin, Cexp) in enumerate(cases):
            a, b, cm = U.to_rank(grid, A), U.to_rank(grid, B), U.to_rank(grid, Cin, ldim_extra=2)
            (H.Symm if i % 2 == 0 else H.Hemm)(side, uplo, 2.0, a, b, 3.0, cm, blocksize=nb)
            res.append((U.block_of(cm), U.padding_untouched(cm), U.block_of(a)))
        return res

    out = H.run_ranks(r * c, body, height=r)
    tol_rel = 1e-13 if dtype == np.float64 else 1e-5
    tol_abs = 1e-10 if dtype == np.float64 else 1e-3
    for i, (side, uplo, A, B, Cin, Cexp) in enumerate(cases):
        k = m if side == H.LEFT else n
        num = den = 0.0
        for p in range(r * c):
            got, pad_ok, a_after = out[p][i]
            exp = Cexp.block(p)
            assert pad_ok and np.array_equal(a_after, A.block(p))
            d = got.astype(np.float64) - exp.astype(np.float64)
            num += float((d * d).sum())
            den += float((exp.astype(np.float64) ** 2).sum())
            if d.size:
                assert np.abs(d).max() <= tol_abs, (side, uplo, p, np.abs(d).max())
        assert (num / max(den, 1e-300)) ** 0.5 <= tol_rel * max(1.0, k) ** 0.5, (side, uplo)


def test_trapezoid_kernels_reproduce_reference_fixtures():
    """hyd_{scale,axpy}_trapezoid_* / hyd_make_trapezoidal_* (the Matrix<T,GPU> forms) against tests/golden/trapezoid.json: digests of
    what the REFERENCE's own ScaleTrapezoid / AxpyTrapezoid / MakeTrapezoidal loops produce (compiled in place,
    oracle/ref/ref_trapezoid.cpp).  The 1x1-grid cases of tests/trapezoid_cases.py; alpha = +-2^k, so bit-exact."""
    import ctypes as C
    import torch
    import trapezoid_cases as T
    from elemental_b200 import lib, check

    class Cuda:
        def run(self, case):
            Y, X = T.inputs(case)
            m, n = Y.shape
            ld = max(m, 1) + case["pad"]
            tdt = torch.float64 if case["dt"] == "f64" else torch.float32
            y = torch.full((max(n, 1), ld), 7.0, dtype=tdt, device="cuda")      # (n, ld) row-major == column-major ld x n
            x = torch.zeros((max(n, 1), max(m, 1)), dtype=tdt, device="cuda")
            if m and n:
                y[:n, :m].copy_(torch.from_numpy(np.ascontiguousarray(Y.T)))
                x[:n, :m].copy_(torch.from_numpy(np.ascontiguousarray(X.T)))
            sc = (C.c_double if case["dt"] == "f64" else C.c_float)(case["alpha"])
            yp, xp = C.c_void_p(y.data_ptr()), C.c_void_p(x.data_ptr())
            if case["op"] == "scale":
                check(getattr(lib, "hyd_scale_trapezoid_" + case["dt"])(sc, case["uplo"], m, n, yp, ld, case["offset"], None))
            elif case["op"] == "axpy":
                check(getattr(lib, "hyd_axpy_trapezoid_" + case["dt"])(case["uplo"], sc, m, n, xp, max(m, 1), yp, ld, case["offset"], None))
            else:
                check(getattr(lib, "hyd_make_trapezoidal_" + case["dt"])(case["uplo"], m, n, yp, ld, case["offset"], None))
            torch.cuda.synchronize()
            assert bool((y[:, m:] == 7.0).all().item())
            return [y[:n, :m].t().cpu().numpy()]

    n = T.check_against_golden(Cuda(), select=lambda c: c["r"] * c["c"] == 1)
    assert n == 60

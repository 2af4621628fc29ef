"""-m gpu: the rest of the reference's GPU BLAS-1 layer through the C ABI (csrc/level1.cu) -- Fill, Hadamard, transposing
Axpy, the 1-D Copy / Axpy / Scale forms (src/hydrogen/blas/gpu/{Fill,Hadamard,Axpy}.cu, GPU_BLAS_impl.hpp:510-597) and the
distributed El::Axpy / TransposeAxpy / Hadamard / Fill / Scale (include/El/blas_like/level1/*.hpp) against the oracle
(oracle/oracle_level1.c).

Copies, fills, products and scalings are single correctly rounded operations: compared bit for bit.  alpha*x + y may or
may not be contracted into an FMA (in the reference it is cuBLAS geam / the compiler's choice): exact when the products
are exact (integer data), else within one rounding of the larger term."""
import ctypes as C
import itertools

import numpy as np
import pytest
import torch

import oracle_lib as O
from elemental_b200 import host as H
from elemental_b200 import lib, check
import gpu_util as U

pytestmark = pytest.mark.gpu
SENT = -777.0


def _sfx(dt):
    return "f64" if np.dtype(dt) == np.float64 else "f32"


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _p(t):
    return C.c_void_p(t.data_ptr())


def _s():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _sc(dt, v):
    return C.c_double(float(v)) if np.dtype(dt) == np.float64 else C.c_float(float(v))


def _axpy_close(got, y, alpha, x, dt):
    """|got - (y + alpha x)| within one rounding of the product plus one of the sum (FMA or not)."""
    eps = np.finfo(dt).eps
    ref = y.astype(np.float64) + float(alpha) * x.astype(np.float64)
    bound = eps * (np.abs(float(alpha) * x.astype(np.float64)) + np.abs(ref)) + np.finfo(dt).tiny
    return bool(np.all(np.abs(got.astype(np.float64) - ref) <= bound))


@pytest.mark.parametrize("dt", [np.float64, np.float32])
def test_fill(dt):
    for (m, n, pad) in [(1, 1, 0), (64, 9, 0), (64, 9, 4), (33, 17, 3), (1000, 3, 0), (2, 500, 2)]:
        ld = m + pad
        for alpha in (2.5, 0.0):
            a = np.full(ld * n, np.nan if alpha == 0.0 else SENT, dt)
            a.reshape(n, ld)[:, m:] = SENT
            d = _dev(a)
            check(getattr(lib, "hyd_fill_" + _sfx(dt))(m, n, _sc(dt, alpha), _p(d), ld, _s()))
            got = d.cpu().numpy().reshape(n, ld)
            assert np.all(got[:, :m] == dt(alpha)), (m, n, pad)
            assert np.all(got[:, m:] == SENT), "ldim padding touched"
    z = torch.zeros(8, dtype=torch.float64, device="cuda")
    assert lib.hyd_fill_f64(4, 2, 1.0, _p(z), 3, _s()) != 0 and b"leading dimension" in lib.hyd_last_error()
    assert lib.hyd_fill_f64(0, 2, 1.0, _p(z), 1, _s()) == 0


@pytest.mark.parametrize("dt", [np.float64, np.float32])
def test_hadamard(dt):
    rng = np.random.default_rng(3)
    # (m, n, (csX, rsX), (csY, rsY), (csZ, rsZ)): contiguous vector path, padded ldims, general strides
    shapes = [(64, 8, (1, 64), (1, 64), (1, 64)), (64, 8, (1, 68), (1, 64), (1, 72)), (33, 17, (1, 33), (1, 40), (1, 35)),
              (50, 7, (2, 101), (1, 50), (3, 151)), (1, 1, (1, 1), (1, 1), (1, 1)), (5, 300, (1, 5), (1, 8), (1, 5))]
    for (m, n, sx, sy, sz) in shapes:
        nx, ny, nz = [(m - 1) * s[0] + (n - 1) * s[1] + 1 for s in (sx, sy, sz)]
        X, Y = rng.uniform(-1, 1, nx).astype(dt), rng.uniform(-1, 1, ny).astype(dt)
        ii, jj = np.meshgrid(np.arange(m), np.arange(n), indexing="ij")
        ix, iy, iz = [(ii * s[0] + jj * s[1]).ravel() for s in (sx, sy, sz)]
        exp = np.full(nz, SENT, dt)
        exp[iz] = X[ix] * Y[iy]
        dX, dY, dZ = _dev(X), _dev(Y), torch.full((nz,), SENT, dtype=_dev(X).dtype, device="cuda")
        check(getattr(lib, "hyd_hadamard_" + _sfx(dt))(m, n, _p(dX), sx[0], sx[1], _p(dY), sy[0], sy[1], _p(dZ), sz[0], sz[1], _s()))
        assert np.array_equal(dZ.cpu().numpy(), exp), (m, n, sx, sy, sz)
    # in place: Z is X, then Z is Y (the reference's MultAssign branches, Hadamard.cu:78-93)
    m, n = 128, 6
    X, Y = rng.uniform(-1, 1, m * n).astype(dt), rng.uniform(-1, 1, m * n).astype(dt)
    dX, dY = _dev(X), _dev(Y)
    check(getattr(lib, "hyd_hadamard_" + _sfx(dt))(m, n, _p(dX), 1, m, _p(dY), 1, m, _p(dX), 1, m, _s()))
    assert np.array_equal(dX.cpu().numpy(), X * Y)
    dX = _dev(X)
    check(getattr(lib, "hyd_hadamard_" + _sfx(dt))(m, n, _p(dX), 1, m, _p(dY), 1, m, _p(dY), 1, m, _s()))
    assert np.array_equal(dY.cpu().numpy(), X * Y)


@pytest.mark.parametrize("dt", [np.float64, np.float32])
def test_axpy_transpose(dt):
    """gpu_blas::Axpy(TRANSPOSE, m, n, alpha, A, lda, B, ldb): B (m x n) += alpha A^T, A is n x m."""
    rng = np.random.default_rng(4)
    for (m, n, pa, pb) in [(1, 1, 0, 0), (64, 64, 0, 0), (70, 33, 3, 5), (5, 200, 1, 0), (300, 7, 0, 2), (129, 130, 0, 0)]:
        lda, ldb = n + pa, m + pb
        A = np.full((m, lda), SENT, dt)           # column-major n x m with ld lda == row-major (m, lda)
        B = np.full((n, ldb), SENT, dt)
        Ai = rng.integers(-50, 50, (n, m)).astype(dt)   # logical A (n x m): integer data -> exact products and sums
        Bi = rng.integers(-50, 50, (m, n)).astype(dt)
        A[:, :n] = Ai.T
        B[:, :m] = Bi.T
        dA, dB = _dev(A), _dev(B)
        check(getattr(lib, "hyd_axpy_trans_" + _sfx(dt))(1, m, n, _sc(dt, 3.0), _p(dA), lda, _p(dB), ldb, _s()))
        got = dB.cpu().numpy()
        assert np.array_equal(got[:, :m].T, Bi + dt(3.0) * Ai.T), (m, n)
        assert np.all(got[:, m:] == SENT)
        # random data: one-ulp bound; transA = 0 is the plain 2-D axpy
        Ar, Br = rng.uniform(-1, 1, (n, m)).astype(dt), rng.uniform(-1, 1, (m, n)).astype(dt)
        A[:, :n], B[:, :m] = Ar.T, Br.T
        dA, dB = _dev(A), _dev(B)
        check(getattr(lib, "hyd_axpy_trans_" + _sfx(dt))(1, m, n, _sc(dt, 0.7), _p(dA), lda, _p(dB), ldb, _s()))
        assert _axpy_close(dB.cpu().numpy()[:, :m].T, Br, dt(0.7), Ar.T, dt), (m, n)
    m, n = 40, 9
    Xi, Yi = rng.integers(-9, 9, (n, m)).astype(dt), rng.integers(-9, 9, (n, m)).astype(dt)
    dX, dY = _dev(Xi), _dev(Yi)
    check(getattr(lib, "hyd_axpy_trans_" + _sfx(dt))(0, m, n, _sc(dt, 2.0), _p(dX), m, _p(dY), m, _s()))
    assert np.array_equal(dY.cpu().numpy(), Yi + dt(2.0) * Xi)
    z = torch.zeros(16, dtype=torch.float64, device="cuda")
    assert lib.hyd_axpy_trans_f64(1, 4, 3, 1.0, _p(z), 2, _p(z), 4, _s()) != 0   # lda < n


@pytest.mark.parametrize("dt", [np.float64, np.float32])
def test_vector_forms(dt):
    """gpu_blas::Copy / Axpy / Scale (size, X, incx, Y, incy)."""
    rng = np.random.default_rng(5)
    for (n, incx, incy) in [(1, 1, 1), (1000, 1, 1), (1003, 1, 1), (777, 3, 1), (500, 1, 4), (4100, 2, 5)]:
        X = rng.integers(-100, 100, (n - 1) * incx + 1).astype(dt)
        Y = rng.integers(-100, 100, (n - 1) * incy + 1).astype(dt)
        exp = Y.copy(); exp[::incy] = X[::incx]
        dX, dY = _dev(X), _dev(Y)
        check(getattr(lib, "hyd_copy1d_" + _sfx(dt))(n, _p(dX), incx, _p(dY), incy, _s()))
        assert np.array_equal(dY.cpu().numpy(), exp), ("copy", n, incx, incy)
        exp = Y.copy(); exp[::incy] = Y[::incy] + dt(3.0) * X[::incx]
        dY = _dev(Y)
        check(getattr(lib, "hyd_axpy1d_" + _sfx(dt))(n, _sc(dt, 3.0), _p(dX), incx, _p(dY), incy, _s()))
        assert np.array_equal(dY.cpu().numpy(), exp), ("axpy", n, incx, incy)
        Xr = rng.uniform(-1, 1, (n - 1) * incx + 1).astype(dt)
        exp = Xr.copy(); exp[::incx] = dt(0.3) * Xr[::incx]
        dX = _dev(Xr)
        check(getattr(lib, "hyd_scale1d_" + _sfx(dt))(n, _sc(dt, 0.3), _p(dX), incx, _s()))
        assert np.array_equal(dX.cpu().numpy(), exp), ("scale", n, incx)
    bad = torch.zeros(4, dtype=torch.float64, device="cuda")
    assert lib.hyd_copy1d_f64(4, _p(bad), 0, _p(bad), 1, _s()) != 0 and b"increments" in lib.hyd_last_error()


DIST_PAIRS = [((O.MC, O.MR), (O.MC, O.MR)), ((O.MC, O.MR), (O.MR, O.MC)), ((O.MC, O.STAR), (O.MC, O.MR)),
              ((O.STAR, O.VR), (O.MC, O.MR)), ((O.MC, O.MR), (O.VC, O.STAR)), ((O.STAR, O.STAR), (O.MR, O.STAR)),
              ((O.VR, O.STAR), (O.STAR, O.MC)), ((O.MR, O.MC), (O.STAR, O.VC))]


@pytest.mark.parametrize("r,c", [(1, 1), (2, 2), (2, 4), (3, 2)])
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_dist_axpy_and_transpose_axpy(r, c, dtype):
    """El::Axpy / El::TransposeAxpy between distributions: equal DistData -> local, else through Y's distribution.
    Integer-valued operands (fill_index) and alpha = 2: every product and sum is exact, so the comparison is bitwise."""
    og = O.Grid(r, c, 1)
    rng = np.random.default_rng(31 * r + c)
    cases = []
    for k, ((xcd, xrd), (ycd, yrd)) in enumerate(DIST_PAIRS):
        for transpose in (0, 1):
            m, n = [(50, 37), (7, 5), (129, 67)][k % 3]
            same_align = (k == 0)
            if transpose and k == 0:
                xcd, xrd = yrd, ycd   # the transposed distribution of Y: local path of TransposeAxpy when aligned
            yca, yra = U.rand_align(rng, og, ycd), U.rand_align(rng, og, yrd)
            if same_align:
                xca, xra = (yra, yca) if transpose else (yca, yra)
            else:
                xca, xra = U.rand_align(rng, og, xcd), U.rand_align(rng, og, xrd)
            X = O.DM(og, n if transpose else m, m if transpose else n, xcd, xrd, xca, xra, dtype).fill_index()
            Y0 = O.DM(og, m, n, ycd, yrd, yca, yra, dtype).fill_index()
            Ye = O.DM(og, m, n, ycd, yrd, yca, yra, dtype).fill_index()
            assert O.lib.orc_axpy(2.0, X.h, Ye.h, transpose) == 0
            cases.append((X, Y0, Ye, transpose))

    def body(grid):
        res = []
        for (X, Y0, Ye, transpose) in cases:
            x, y = U.to_rank(grid, X, ldim_extra=2), U.to_rank(grid, Y0, ldim_extra=4)
            (H.TransposeAxpy if transpose else H.Axpy)(2.0, x, y)
            res.append((U.block_of(y), U.padding_untouched(y), U.block_of(x)))
        return res

    out = H.run_ranks(r * c, body, height=r)
    for p in range(r * c):
        for i, (X, Y0, Ye, transpose) in enumerate(cases):
            got, pad_ok, x_after = out[p][i]
            assert np.array_equal(got, Ye.block(p)), (r, c, p, i, transpose)
            assert pad_ok
            assert np.array_equal(x_after, X.block(p))


@pytest.mark.parametrize("r,c", [(1, 1), (2, 4)])
def test_dist_hadamard_fill_scale(r, c):
    og = O.Grid(r, c, 1)
    m, n = 61, 45
    A = O.DM(og, m, n, O.MC, O.MR, 1 % r, 2 % c).fill(1)
    B = O.DM(og, m, n, O.MC, O.MR, 1 % r, 2 % c).fill(2)
    Ce = O.DM(og, m, n, O.MC, O.MR, 1 % r, 2 % c).fill_const(SENT)
    assert O.lib.orc_hadamard(A.h, B.h, Ce.h) == 0
    Se = O.DM(og, m, n, O.MC, O.MR, 1 % r, 2 % c).fill(1)
    O.lib.orc_scale(0.3, Se.h)
    Bad = O.DM(og, m, n, O.MC, O.MR, 0, 0)
    assert r * c == 1 or O.lib.orc_hadamard(A.h, Bad.h, Ce.h) == 3   # "A and B must be aligned"

    def body(grid):
        a, b, cm = U.to_rank(grid, A), U.to_rank(grid, B), U.to_rank(grid, Ce, ldim_extra=2)
        H.Hadamard(a, b, cm)
        had = (U.block_of(cm), U.padding_untouched(cm))
        if r * c > 1:
            with pytest.raises(H.HydrogenError):
                H.Hadamard(a, U.to_rank(grid, Bad), cm)
        H.Scale(0.3, a)
        sc = U.block_of(a)
        cm.store.fill_(float("nan"))
        H.Zero(cm)
        zero_ok = bool((cm.store[: cm.lw, : cm.lh] == 0).all().item())
        H.Fill(cm, 4.25)
        fill_ok = bool((cm.store[: cm.lw, : cm.lh] == 4.25).all().item())
        return had, sc, zero_ok, fill_ok

    out = H.run_ranks(r * c, body, height=r)
    for p in range(r * c):
        (had, pad_ok), sc, zero_ok, fill_ok = out[p]
        assert np.array_equal(had, Ce.block(p)) and pad_ok
        assert np.array_equal(sc, Se.block(p))
        assert zero_ok and fill_ok


def test_dist_axpy_errors():
    def body(grid):
        x = H.DistMatrix(grid, 10, 8)
        y = H.DistMatrix(grid, 8, 10)
        with pytest.raises(H.HydrogenError, match="nonconformal"):
            H.Axpy(1.0, x, y)
        H.TransposeAxpy(1.0, x, y)   # conformal once transposed
        with pytest.raises(H.HydrogenError, match="nonconformal"):
            H.TransposeAxpy(1.0, x, H.DistMatrix(grid, 10, 8))
        return True

    assert all(H.run_ranks(2, body, height=1))

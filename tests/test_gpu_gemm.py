"""-m gpu: distributed Gemm parity through hyd_gemm_* (El::Gemm on DistMatrix), every orientation x
SUMMA variant, against the oracle's restated SUMMA loops (oracle/oracle_gemm.c) on the same inputs.

Tolerances (written here, as BASELINE.json asks): relative Frobenius error <= 1e-13*sqrt(k) (fp64),
1e-5*sqrt(k) (fp32); and the reference's own bound |x-y| <= 1e-10 / 1e-3 at alpha=2, beta=3
(tests/blas_like/BasicGemm.cpp:93-94,108)."""
import itertools
import os

import numpy as np
import pytest
import torch

import oracle_lib as O
from oracle_lib import MC, MR, VC, VR, STAR
from elemental_b200 import host as H
import gpu_util as U

pytestmark = pytest.mark.gpu


def _gemm_case(r, c, dtype, oA, oB, alg, m, n, k, nb, rng, dists=None, alpha=2.0, beta=3.0):
    og = O.Grid(r, c, 1)
    am, an = (k, m) if oA else (m, k)
    bm, bn = (n, k) if oB else (k, n)
    dA, dB, dC = dists or ((MC, MR), (MC, MR), (MC, MR))
    mk = lambda h, w, d, seed: O.DM(og, h, w, d[0], d[1], U.rand_align(rng, og, d[0]), U.rand_align(rng, og, d[1]), dtype).fill(seed)
    A, B, Cm = mk(am, an, dA, 1), mk(bm, bn, dB, 2), mk(m, n, dC, 3)
    Cin = O.DM(og, m, n, dC[0], dC[1], Cm.ca, Cm.ra, dtype).fill(3)
    O.gemm(oA, oB, alg, alpha, A, B, beta, Cm, nb)
    return (oA, oB, alg, alpha, beta, nb, k), A, B, Cin, Cm


def _run_cases(r, c, cases, dtype):
    def body(grid):
        res = []
        for (oA, oB, alg, alpha, beta, nb, k), A, B, Cin, Cexp in cases:
            a, b, cm = U.to_rank(grid, A), U.to_rank(grid, B), U.to_rank(grid, Cin, ldim_extra=2)
            H.Gemm(oA, oB, alpha, a, b, beta, cm, alg, blocksize=nb)
            res.append((U.block_of(cm), U.padding_untouched(cm), U.block_of(a), U.block_of(b)))
        return res

    out = H.run_ranks(r * c, body, height=r)
    tol_rel = 1e-13 if dtype == np.float64 else 1e-5
    tol_abs = 1e-10 if dtype == np.float64 else 1e-3
    for i, ((oA, oB, alg, alpha, beta, nb, k), A, B, Cin, Cexp) in enumerate(cases):
        num = den = 0.0
        for p in range(r * c):
            got, pad_ok, a_after, b_after = out[p][i]
            exp = Cexp.block(p)
            assert pad_ok
            assert np.array_equal(a_after, A.block(p)) and np.array_equal(b_after, B.block(p))  # inputs untouched
            d = got.astype(np.float64) - exp.astype(np.float64)
            num += float((d * d).sum())
            den += float((exp.astype(np.float64) ** 2).sum())
            if got.size:
                assert np.abs(d).max() <= tol_abs, (oA, oB, alg, p, np.abs(d).max())
        rel = (num / max(den, 1e-300)) ** 0.5
        assert rel <= tol_rel * max(1.0, k) ** 0.5, (oA, oB, alg, rel)


@pytest.mark.parametrize("r,c", [(1, 1), (1, 2), (2, 2), (2, 4), (3, 2)])
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_gemm_all_variants(r, c, dtype):
    rng = np.random.default_rng(17 * r + c)
    cases = []
    for oA, oB in itertools.product((0, 1), (0, 1)):
        for alg in (H.GEMM_SUMMA_A, H.GEMM_SUMMA_B, H.GEMM_SUMMA_C, H.GEMM_SUMMA_DOT, H.GEMM_DEFAULT):
            cases.append(_gemm_case(r, c, dtype, oA, oB, alg, 145, 138, 152, 32, rng))
    _run_cases(r, c, cases, dtype)


@pytest.mark.parametrize("r,c", [(1, 1), (2, 2), (2, 4)])
def test_gemm_basicgemm_config(r, c):
    """tests/blas_like/BasicGemm.cpp defaults: m=n=k=1000, blocksize 128, alpha=2, beta=3, [MC,MR]."""
    rng = np.random.default_rng(5)
    O.use_blas(True)
    try:
        cases = [_gemm_case(r, c, np.float64, 0, 0, H.GEMM_SUMMA_C, 1000, 1000, 1000, 128, rng),
                 _gemm_case(r, c, np.float32, 0, 0, H.GEMM_SUMMA_C, 1000, 1000, 1000, 128, rng)]
    finally:
        O.use_blas(False)
    _run_cases(r, c, cases[:1], np.float64)
    _run_cases(r, c, cases[1:], np.float32)


def test_gemm_default_heuristic_and_dot_operands(monkeypatch):
    """GEMM_DEFAULT picks Dot/B/A/C by shape (NN.hpp:910-925); [*,VC] operands pass the proxies untouched (config 5)."""
    monkeypatch.setenv("HYD_DOT_TILE_BYTES", "1")  # reference tile (2000 rows): the m=2100 case covers two tiles
    r, c = 2, 2
    rng = np.random.default_rng(9)
    cases = [
        _gemm_case(r, c, np.float32, 0, 1, H.GEMM_DEFAULT, 40, 24, 600, 16, rng, dists=((STAR, VC), (STAR, VC), (MC, MR))),
        _gemm_case(r, c, np.float32, 0, 1, H.GEMM_SUMMA_DOT, 2100, 24, 300, 16, rng, dists=((STAR, VC), (STAR, VC), (MC, MR))),
        _gemm_case(r, c, np.float32, 0, 0, H.GEMM_DEFAULT, 30, 200, 100, 16, rng),   # m <= n, 2m <= k -> B
        _gemm_case(r, c, np.float32, 0, 0, H.GEMM_DEFAULT, 200, 30, 100, 16, rng),   # n <= m, 2n <= k -> A
        _gemm_case(r, c, np.float32, 1, 0, H.GEMM_SUMMA_C, 64, 64, 64, 16, rng, dists=((VC, STAR), (STAR, MR), (MR, MC))),
    ]
    _run_cases(r, c, cases, np.float32)


def test_gemm_degenerate_and_errors():
    rng = np.random.default_rng(2)
    cases = [_gemm_case(2, 2, np.float64, 0, 0, H.GEMM_SUMMA_C, 5, 3, 1, 128, rng),
             _gemm_case(2, 2, np.float64, 0, 0, H.GEMM_SUMMA_A, 1, 1, 7, 2, rng),
             _gemm_case(2, 2, np.float64, 1, 1, H.GEMM_SUMMA_B, 9, 2, 3, 4, rng),
             _gemm_case(2, 2, np.float64, 0, 0, H.GEMM_SUMMA_C, 16, 16, 40, 128, rng, alpha=1.0, beta=0.0)]
    _run_cases(2, 2, cases, np.float64)

    def body(grid):
        a = H.DistMatrix(grid, 10, 12)
        b = H.DistMatrix(grid, 13, 9)
        cm = H.DistMatrix(grid, 10, 9)
        with pytest.raises(H.HydrogenError, match="nonconformal"):
            H.Gemm(0, 0, 1.0, a, b, 0.0, cm)
        with pytest.raises(H.HydrogenError, match="CANNON"):
            H.Gemm(0, 0, 1.0, a, H.DistMatrix(grid, 12, 9), 0.0, cm, H.GEMM_CANNON)
        return True
    assert H.run_ranks(1, body) == [True]


def test_gemm_host_buffers_roundtrip():
    """hyd_gemm_host_*: host-resident operands, copies inside the call."""
    rng = np.random.default_rng(4)
    case = _gemm_case(1, 1, np.float64, 0, 1, H.GEMM_SUMMA_C, 200, 150, 90, 32, rng)
    (oA, oB, alg, alpha, beta, nb, k), A, B, Cin, Cexp = case

    def body(grid):
        mk = lambda X: H.DistMatrix(grid, X.m, X.n, X.cd, X.rd, X.dtype, X.ca, X.ra, device="cpu", pinned=True)
        a, b, cm = mk(A), mk(B), mk(Cin)
        for d, X in ((a, A), (b, B), (cm, Cin)):
            d.store[: d.lw, : d.lh].copy_(torch.from_numpy(np.ascontiguousarray(X.block(0).T)))
        H.Gemm(oA, oB, alpha, a, b, beta, cm, alg, blocksize=nb)
        return cm.local_numpy()
    got = H.run_ranks(1, body)[0]
    assert np.abs(got - Cexp.block(0)).max() <= 1e-10


def _host_dm(grid, X, null_if_empty=False):
    d = H.DistMatrix(grid, X.m, X.n, X.cd, X.rd, X.dtype, X.ca, X.ra, device="cpu", pinned=True)
    p = grid.Rank()
    if d.lh and d.lw:
        d.store[: d.lw, : d.lh].copy_(torch.from_numpy(np.ascontiguousarray(X.block(p).T)))
    elif null_if_empty:
        d.desc = lambda d=d: H.hyd_dm(None, 1, d.m, d.n, d.cd, d.rd, d.ca, d.ra, 0, 0)   # hydrogen_b200.h: buf may be NULL
    return d


@pytest.mark.parametrize("r,c", [(1, 1), (1, 2), (2, 2), (2, 4)])
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_gemm_host_column_blocked_path(r, c, dtype):
    """hyd_gemm_host_* on every grid: the host-resident stationary-C Gemm cuts the local C into column blocks (finished
    blocks download under the next block's GEMMs, B arrives one whole column block ahead).  Every orientation, beta = 0 and
    beta != 0, a ragged last block; against the oracle, and bit-identical between 1 block and several (the k-sum of every
    element is grouped the same way whatever the column cuts)."""
    rng = np.random.default_rng(23 + r + 3 * c)
    n = 384 * c + 70
    cases = [_gemm_case(r, c, dtype, oA, oB, H.GEMM_SUMMA_C, 300, n, 260, 64, rng, alpha=2.0, beta=beta)
             for (oA, oB, beta) in ((0, 0, 0.0), (0, 1, 3.0), (1, 0, 0.0), (1, 1, -0.5))]

    def body(grid):
        res = []
        for nblk in (1, 3, 8):
            grid.SetHostBlocks(nblk)
            for (oA, oB, alg, alpha, beta, nb, k), A, B, Cin, Cexp in cases:
                a, b, cm = _host_dm(grid, A), _host_dm(grid, B), _host_dm(grid, Cin)
                H.Gemm(oA, oB, alpha, a, b, beta, cm, alg, blocksize=nb)
                res.append(cm.local_numpy().copy())
        return res

    out = H.run_ranks(r * c, body, height=r)
    tol = (1e-13 if dtype == np.float64 else 1e-5)
    nc = len(cases)
    for i, ((oA, oB, alg, alpha, beta, nb, k), A, B, Cin, Cexp) in enumerate(cases):
        num = den = 0.0
        for p in range(r * c):
            exp = Cexp.block(p).astype(np.float64)
            d = out[p][i].astype(np.float64) - exp
            num += float((d * d).sum())
            den += float((exp ** 2).sum())
            for alt in (out[p][nc + i], out[p][2 * nc + i]):
                if dtype == np.float64:   # one fp64 kernel, one tile shape: the k-sum of an element does not depend on the cuts
                    assert np.array_equal(out[p][i], alt), (oA, oB, p)
                else:                     # the fp32 dispatcher picks its kernel / tile by the block's width: tolerance only
                    assert np.abs(out[p][i] - alt).max() <= 1e-3, (oA, oB, p)
        assert (num / max(den, 1e-300)) ** 0.5 <= tol * k ** 0.5, (oA, oB, num, den)


def test_gemm_host_null_blocks_on_short_matrices():
    """m smaller than the grid height: some ranks own no row of A and of C and pass NULL host pointers
    (include/hydrogen_b200.h: buf may be NULL for an empty local block).  Every rank must still take the same path -- same
    panel schedule and exchange rounds -- whatever its own pointers are (stream_ops was decided per rank)."""
    r, c = 2, 2
    rng = np.random.default_rng(77)
    cases = [_gemm_case(r, c, np.float64, 0, 0, H.GEMM_SUMMA_C, 1, 300, 700, 128, rng),
             _gemm_case(r, c, np.float64, 0, 1, H.GEMM_SUMMA_C, 1, 1, 520, 128, rng),
             _gemm_case(r, c, np.float64, 1, 0, H.GEMM_SUMMA_A, 1, 200, 300, 64, rng)]

    def body(grid):
        res = []
        for (oA, oB, alg, alpha, beta, nb, k), A, B, Cin, Cexp in cases:
            a, b, cm = _host_dm(grid, A, True), _host_dm(grid, B, True), _host_dm(grid, Cin, True)
            H.Gemm(oA, oB, alpha, a, b, beta, cm, alg, blocksize=nb)
            res.append(cm.local_numpy().copy())
        return res

    out = H.run_ranks(r * c, body, height=r)
    for i, (_, A, B, Cin, Cexp) in enumerate(cases):
        for p in range(r * c):
            exp = Cexp.block(p)
            assert out[p][i].shape == exp.shape
            if exp.size:
                assert np.abs(out[p][i] - exp).max() <= 1e-10


def test_gemm_over_staged_transport(monkeypatch):
    """Every orientation x algorithm on a 2x4 grid with the staged (pack -> exchange -> unpack) transport; the default
    for rank-threads is the push transport with panels written directly into the readers' buffers."""
    monkeypatch.setenv("HYD_TRANSPORT", "staged")
    test_gemm_all_variants(2, 4, np.float64)


def _associativity(grid, dtype, oA, oB, n, nb, alpha, beta, seed):
    """The reference's own Gemm test (tests/blas_like/Gemm.cpp:27-61, TestAssociativity):
    (alpha op(A) op(B) + beta C) X  ==  alpha op(A) (op(B) X) + beta C X   for 100 random right-hand sides.
    Returns this rank's (||E||_F^2, ||Y||_F^2)."""
    tdt = torch.float64 if dtype == np.float64 else torch.float32
    gen = torch.Generator(device="cuda").manual_seed(seed + 31 * grid.Rank())
    mk = lambda h, w: H.DistMatrix(grid, h, w, H.MC, H.MR, dtype=dtype)
    A, B, C0, Cf = mk(n, n), mk(n, n), mk(n, n), mk(n, n)
    X, Y, Z = mk(n, 100), mk(n, 100), mk(n, 100)
    for d in (A, B, C0, X):
        d.store.copy_(torch.rand(d.store.shape, dtype=tdt, device="cuda", generator=gen) * 2 - 1)
    Cf.store.copy_(C0.store)
    H.Gemm(oA, oB, alpha, A, B, beta, Cf, H.GEMM_SUMMA_C, blocksize=nb)        # the Gemm under test
    H.Gemm(oB, H.NORMAL, 1.0, B, X, 0.0, Z, blocksize=nb)
    H.Gemm(oA, H.NORMAL, alpha, A, Z, 0.0, Y, blocksize=nb)
    H.Gemm(H.NORMAL, H.NORMAL, beta, C0, X, 1.0, Y, blocksize=nb)
    torch.cuda.current_stream().synchronize()
    y2 = float((Y.store[: Y.lw, : Y.lh].double() ** 2).sum())
    H.Gemm(H.NORMAL, H.NORMAL, -1.0, Cf, X, 1.0, Y, blocksize=nb)
    torch.cuda.current_stream().synchronize()
    e2 = float((Y.store[: Y.lw, : Y.lh].double() ** 2).sum())
    return e2, y2


@pytest.mark.parametrize("r,c,n,dtype,oA,oB", [(1, 1, 16384, np.float64, 0, 0),     # BASELINE.json configs[1] size, one GPU
                                               (2, 4, 8192, np.float64, 0, 0),      # the 2x4 grid of configs[1] (rank-threads)
                                               (2, 2, 8192, np.float32, 1, 0),      # configs[2] orientations
                                               (2, 2, 8192, np.float32, 0, 1)])
def test_gemm_associativity_at_scale(r, c, n, dtype, oA, oB):
    """Size-independent parity property at (or near) the benchmark sizes, where the oracle is too slow:
    ||E||_F / ||Y||_F <= tol * sqrt(k), tol = 1e-13 (fp64) / 1e-5 (fp32) as in BASELINE.json."""
    out = H.run_ranks(r * c, lambda grid: _associativity(grid, dtype, oA, oB, n, 1024, 2.0, 3.0, 7), height=r)
    e2 = sum(o[0] for o in out)
    y2 = sum(o[1] for o in out)
    rel = (e2 / y2) ** 0.5
    tol = (1e-13 if dtype == np.float64 else 1e-5) * n ** 0.5
    assert rel <= tol, (rel, tol)


def test_multisync_three_streams_bit_identical():
    """hyd_multisync_begin / _end = MakeMultiSync (include/hydrogen/SyncInfo.hpp:136-186, Gemm.cpp:117-119): A produced on
    stream 1, B on stream 2, C on stream 3, no host synchronisation; the fenced Gemm equals the single-stream run bit for bit.
    Both input streams are kept busy first, so an unfenced Gemm would read the zeros A and B still hold."""
    r, c, n, nb = 2, 2, 512, 128

    def body(grid):
        gen = torch.Generator(device="cuda").manual_seed(5 + grid.Rank())
        mk = lambda: H.DistMatrix(grid, n, n, H.MC, H.MR)
        A0, B0, A, B, C1, C2 = mk(), mk(), mk(), mk(), mk(), mk()
        for d in (A0, B0, C1):
            d.store.copy_(torch.rand(d.store.shape, dtype=torch.float64, device="cuda", generator=gen) * 2 - 1)
        C2.store.copy_(C1.store)
        cur = torch.cuda.current_stream()
        H.Gemm(0, 0, 2.0, A0, B0, 3.0, C1, H.GEMM_SUMMA_C, blocksize=nb)
        cur.synchronize()
        s1, s2, s3 = torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.Stream()
        busy = torch.empty(64 << 20, dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        grid.world.barrier()
        with torch.cuda.stream(s1):
            for _ in range(20):
                busy[: 32 << 20].fill_(1.0)
            A.store.copy_(A0.store, non_blocking=True)
        with torch.cuda.stream(s2):
            for _ in range(20):
                busy[32 << 20:].fill_(2.0)
            B.store.copy_(B0.store, non_blocking=True)
        with H.MultiSync(s3, s1, s2):
            H.Gemm(0, 0, 2.0, A, B, 3.0, C2, H.GEMM_SUMMA_C, stream=s3, blocksize=nb)
        with torch.cuda.stream(s1):      # exit half of the fence: A's stream may overwrite A at once
            A.store.fill_(-7.0)
        s3.synchronize()
        s1.synchronize()
        s2.synchronize()
        return bool(torch.equal(C1.store, C2.store))

    assert all(H.run_ranks(r * c, body, height=r))


def test_steady_state_calls_allocate_nothing_and_never_block_the_host():
    """a24: the reference takes its temporaries from a pool (simple_buffer on the CUB allocator) and stays asynchronous.  After a
    warm-up call, Gemm with proxied operands (every operand in a foreign distribution), Syrk, Symm and a redistribution add
    nothing to hyd_alloc_count / hyd_sync_count."""
    from elemental_b200 import lib
    r, c = 2, 2

    def body(grid):
        mk = lambda h, w, cd, rd: H.DistMatrix(grid, h, w, cd, rd)
        A, B, Cm = mk(300, 280, VC, STAR), mk(280, 260, STAR, VR), mk(300, 260, MR, MC)
        S, Sq = mk(300, 300, MC, MR), mk(300, 300, MC, MR)
        out_, A2 = mk(300, 260, MC, MR), mk(300, 280, MR, MC)
        for d in (A, B, Cm, S, Sq):
            d.store.uniform_(-1, 1)

        def work():
            for alg in (H.GEMM_SUMMA_A, H.GEMM_SUMMA_B, H.GEMM_SUMMA_C, H.GEMM_SUMMA_DOT):
                H.Gemm(0, 0, 1.0, A, B, 0.5, Cm, alg, blocksize=64)
            H.Syrk(H.LOWER, H.NORMAL, 1.0, A, 0.0, Sq, blocksize=64)
            H.Symm(H.LEFT, H.UPPER, 1.0, S, Cm, 0.0, out_, blocksize=64)
            H.Copy(A, A2)

        work()
        work()
        torch.cuda.current_stream().synchronize()
        grid.world.barrier()
        a0, s0 = lib.hyd_alloc_count(), lib.hyd_sync_count()
        grid.world.barrier()
        for _ in range(3):
            work()
        torch.cuda.current_stream().synchronize()
        grid.world.barrier()
        return lib.hyd_alloc_count() - a0, lib.hyd_sync_count() - s0

    out = H.run_ranks(r * c, body, height=r)
    assert all(o == (0, 0) for o in out), out

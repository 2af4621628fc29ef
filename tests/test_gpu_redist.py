"""-m gpu: redistribution parity, bit-exact, through hyd_redist_* / hyd_axpy_contract_*.

The reference's own spec for this path is tests/core/DistMatrix.cpp:28-76: for every (src,dst)
distribution pair with random alignments, B = A must reproduce A bit for bit after gathering.
Here every rank's local block is additionally compared with the oracle's restated primitive
chain / closed-form ownership definition (oracle/oracle_core.c, oracle_util.c)."""
import itertools

import numpy as np
import pytest

import oracle_lib as O
from oracle_lib import MC, MR, VC, VR, STAR
from elemental_b200 import host as H
import gpu_util as U

pytestmark = pytest.mark.gpu

GRIDS = [(1, 1), (1, 2), (2, 2), (2, 4), (4, 2), (3, 2)]


def _check_redist(r, c, cases, dtype, transpose=False):
    og = O.Grid(r, c, 1)
    P = r * c
    srcs, exps = [], []
    for (m, n, (scd, srd, sca, sra), (dcd, drd, dca, dra)) in cases:
        A = O.DM(og, m, n, scd, srd, sca, sra, dtype).fill_index()
        dm, dn = (n, m) if transpose else (m, n)
        B = O.DM(og, dm, dn, dcd, drd, dca, dra, dtype).fill_const(-77)
        O.lib.orc_redist_closed(A.h, B.h, int(transpose))
        srcs.append(A)
        exps.append(B)

    def body(grid):
        res = []
        for A, B in zip(srcs, exps):
            a = U.to_rank(grid, A, ldim_extra=2)
            b = U.to_rank(grid, O.DM(og, B.m, B.n, B.cd, B.rd, B.ca, B.ra, dtype).fill_const(-77), ldim_extra=4)
            (H.Transpose if transpose else H.Copy)(a, b)
            res.append((U.block_of(b), U.padding_untouched(b), U.block_of(a)))
        return res

    out = H.run_ranks(P, body, height=r)
    for p in range(P):
        for i, (A, B) in enumerate(zip(srcs, exps)):
            got, pad_ok, src_after = out[p][i]
            assert np.array_equal(got, B.block(p)), (r, c, p, cases[i])
            assert pad_ok
            assert np.array_equal(src_after, A.block(p))


@pytest.mark.parametrize("r,c", GRIDS)
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_all_pairs_random_alignments(r, c, dtype):
    """DistMatrix.cpp:28-76 restated for the 11 GPU-valid distributions (121 pairs), ragged sizes."""
    og = O.Grid(r, c, 1)
    rng = np.random.default_rng(100 * r + c)
    cases = []
    for (scd, srd), (dcd, drd) in itertools.product(O.PAIRS, O.PAIRS):
        m, n = [(50, 50), (7, 5), (129, 67)][len(cases) % 3]
        cases.append((m, n, (scd, srd, U.rand_align(rng, og, scd), U.rand_align(rng, og, srd)),
                      (dcd, drd, U.rand_align(rng, og, dcd), U.rand_align(rng, og, drd))))
    _check_redist(r, c, cases, dtype)


@pytest.mark.parametrize("r,c", [(1, 1), (2, 2), (2, 4), (3, 2)])
def test_transposing_redistribution(r, c):
    """El::Transpose(A,B) for the pairs SUMMA drives ([MC,MR]->[MR,*], [VR,*]->[*,MR], [MC,MR]->[MR,MC], ...)."""
    og = O.Grid(r, c, 1)
    rng = np.random.default_rng(7 * r + c)
    cases = []
    for (scd, srd), (dcd, drd) in itertools.product(O.PAIRS, O.PAIRS):
        m, n = [(33, 50), (1, 9), (130, 61)][len(cases) % 3]
        cases.append((m, n, (scd, srd, U.rand_align(rng, og, scd), U.rand_align(rng, og, srd)),
                      (dcd, drd, U.rand_align(rng, og, dcd), U.rand_align(rng, og, drd))))
    _check_redist(r, c, cases, np.float64, transpose=True)


@pytest.mark.parametrize("r,c", [(2, 2), (2, 4)])
def test_primitive_chains_match_reference_restatement(r, c):
    """The restated reference chains (pack -> collective -> unpack; Copy/*.hpp) and the one-hop engine give the
    same local blocks."""
    og = O.Grid(r, c, 1)
    chains = [("orc_row_allgather", (MC, MR), (MC, STAR)), ("orc_col_allgather", (MC, MR), (STAR, MR)),
              ("orc_col_alltoall_demote", (MC, MR), (VC, STAR)), ("orc_row_alltoall_demote", (MC, MR), (STAR, VR)),
              ("orc_partial_col_allgather", (VC, STAR), (MC, STAR)), ("orc_partial_row_allgather", (STAR, VR), (STAR, MR)),
              ("orc_vector_exchange", (VC, STAR), (VR, STAR)), ("orc_vector_exchange", (STAR, VR), (STAR, VC))]
    srcs, exps = [], []
    for prim, (scd, srd), (dcd, drd) in chains:
        A = O.DM(og, 1000, 131, scd, srd, 0, 0).fill(5)
        B = O.DM(og, 1000, 131, dcd, drd, 0, 0).fill_const(-77)
        assert getattr(O.lib, prim)(A.h, B.h) == 0
        srcs.append(A)
        exps.append(B)

    def body(grid):
        res = []
        for A, B in zip(srcs, exps):
            a = U.to_rank(grid, A)
            b = H.DistMatrix(grid, B.m, B.n, B.cd, B.rd, B.dtype, B.ca, B.ra)
            H.Copy(a, b)
            res.append(U.block_of(b))
        return res

    out = H.run_ranks(r * c, body, height=r)
    for p in range(r * c):
        for i, B in enumerate(exps):
            assert np.array_equal(out[p][i], B.block(p)), (chains[i], p)


def test_empty_and_degenerate():
    og = O.Grid(2, 2, 1)
    cases = [(0, 5, (MC, MR, 0, 0), (STAR, STAR, 0, 0)), (5, 0, (MC, MR, 1, 1), (VC, STAR, 2, 0)),
             (1, 1, (MC, MR, 1, 1), (STAR, VR, 0, 3)), (1, 40, (VC, STAR, 3, 0), (MR, MC, 1, 1)),
             (3, 1, (STAR, STAR, 0, 0), (MC, MR, 1, 0))]
    _check_redist(2, 2, cases, np.float64)


def test_redist_errors():
    def body(grid):
        a = H.DistMatrix(grid, 10, 10, MC, MR)
        b = H.DistMatrix(grid, 10, 11, MC, STAR)
        with pytest.raises(H.HydrogenError, match="nonconformal"):
            H.Copy(a, b)
        c = H.DistMatrix(grid, 10, 10, MC, MR)
        c.cd = H.MD
        with pytest.raises(H.HydrogenError, match="unsupported distribution"):
            H.Copy(a, c)
        return True
    assert H.run_ranks(1, body) == [True]


@pytest.mark.parametrize("r,c", [(1, 1), (2, 2), (2, 4), (3, 2)])
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_axpy_contract(r, c, dtype):
    """AxpyContract / TransposeAxpyContract (AxpyContract.hpp:483-544, TransposeAxpyContract.hpp:14-48): the
    oracle sums partials in ascending VC rank, and so does the GPU reduce kernel -> bit-exact here."""
    og = O.Grid(r, c, 1)
    rng = np.random.default_rng(31 * r + c)
    m, n = 67, 45
    cases = [(MC, STAR, 0), (STAR, MR, 0), (STAR, STAR, 0), (MR, STAR, 1), (STAR, MC, 1), (MR, STAR, 0), (VC, STAR, 0),
             (STAR, VR, 0), (MC, MR, 0)]
    srcs, dsts, exps = [], [], []
    for (scd, srd, tr) in cases:
        sm, sn = (n, m) if tr else (m, n)
        A = O.DM(og, sm, sn, scd, srd, U.rand_align(rng, og, scd), U.rand_align(rng, og, srd), dtype)
        for p in range(A.P):
            A.block(p)[...] = rng.standard_normal(A.block(p).shape).astype(dtype)
        ca, ra = U.rand_align(rng, og, MC), U.rand_align(rng, og, MR)
        B0 = O.DM(og, m, n, MC, MR, ca, ra, dtype).fill(9)
        B1 = O.DM(og, m, n, MC, MR, ca, ra, dtype).fill(9)
        O.lib.orc_axpy_contract_closed(1.5, A.h, B1.h, tr)
        srcs.append(A); dsts.append(B0); exps.append(B1)

    def body(grid):
        res = []
        for (scd, srd, tr), A, B0 in zip(cases, srcs, dsts):
            a, b = U.to_rank(grid, A), U.to_rank(grid, B0, ldim_extra=2)
            (H.TransposeAxpyContract if tr else H.AxpyContract)(1.5, a, b)
            res.append((U.block_of(b), U.padding_untouched(b)))
        return res

    out = H.run_ranks(r * c, body, height=r)
    for p in range(r * c):
        for i, B1 in enumerate(exps):
            got, pad_ok = out[p][i]
            assert pad_ok
            assert np.array_equal(got, B1.block(p)), (cases[i], p, np.abs(got - B1.block(p)).max())


@pytest.fixture
def staged_transport(monkeypatch):
    """Rank-threads default to the push transport (peer stores into receive regions, the layout one-process-per-GPU
    runs use over NVLink); this fixture routes exchanges through the staged pack -> exchange -> unpack path
    (the NCCL send/recv fallback's data layout) instead."""
    monkeypatch.setenv("HYD_TRANSPORT", "staged")
    yield


@pytest.mark.parametrize("r,c", [(1, 2), (2, 2), (2, 4)])
def test_staged_transport_all_pairs(r, c, staged_transport):
    og = O.Grid(r, c, 1)
    rng = np.random.default_rng(41 * r + c)
    for tr in (False, True):
        cases = []
        for (scd, srd), (dcd, drd) in itertools.product(O.PAIRS, O.PAIRS):
            m, n = [(50, 50), (7, 5), (129, 67)][len(cases) % 3]
            cases.append((m, n, (scd, srd, U.rand_align(rng, og, scd), U.rand_align(rng, og, srd)),
                          (dcd, drd, U.rand_align(rng, og, dcd), U.rand_align(rng, og, drd))))
        _check_redist(r, c, cases, np.float64, transpose=tr)


def test_staged_transport_axpy_contract(staged_transport):
    test_axpy_contract(2, 2, np.float64)

"""-m gpu: seam S2 of SURVEY 8(b) through the C ABI -- the reference's pack / unpack utilities with their own
argument lists (copy::util::*, include/El/blas_like/level1/Copy/util.hpp) and the typed collectives on the grid's
sub-communicators (mpi::AllGather / AllToAll / SendRecv / Broadcast / ReduceScatter).

Pack / unpack: bit-exact against the oracle's restatement of the same loops (oracle/oracle_util.c), including bytes
the loops must not touch.  Collectives: bit-exact against their definition on seeded inputs; ReduceScatter adds in
ascending communicator rank."""
import ctypes as C

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
    return "f64" if dt == np.float64 else "f32"


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _p(t):
    return C.c_void_p(t.data_ptr())


def _hp(a):
    return C.c_void_p(a.ctypes.data)


def _s():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _index_matrix(ld, w, dt):
    """Column-major (ld x w) buffer, every element distinct and exact in fp32."""
    return (np.arange(ld * w, dtype=np.int64) % (1 << 24)).astype(dt)


CASES = [(1, 1), (7, 5), (50, 33), (129, 67), (260, 9), (3, 300)]


@pytest.mark.parametrize("dt", [np.float64, np.float32])
@pytest.mark.parametrize("kind", ["row", "col"])
def test_strided_pack_unpack_match_oracle(kind, dt):
    """{Row,Col}Strided{Pack,Unpack} (util.hpp:148-184, 363-421): portions and the round trip, bit for bit."""
    rng = np.random.default_rng(5 if kind == "row" else 6)
    es = np.dtype(dt).itemsize
    for (h, w) in CASES:
        for stride in (1, 2, 3, 4, 8):
            align = int(rng.integers(0, stride))
            ld = h + int(rng.integers(0, 4))
            maxl = -(-w // stride) if kind == "row" else -(-h // stride)
            portion = max(1, (h if kind == "row" else w) * maxl) + int(rng.integers(0, 3))
            A = _index_matrix(ld, w, dt)
            exp = np.full(stride * portion, SENT, dt)
            getattr(O.lib, f"orc_{kind}_strided_pack")(h, w, align, stride, _hp(A), ld, _hp(exp), portion, es)
            dA, dP = _dev(A), torch.full((stride * portion,), SENT, dtype=torch.from_numpy(exp).dtype, device="cuda")
            check(getattr(lib, f"hyd_{kind}_strided_pack_{_sfx(dt)}")(h, w, align, stride, _p(dA), ld, _p(dP), portion, _s()))
            assert np.array_equal(dP.cpu().numpy(), exp), (kind, h, w, stride, align)
            # unpack into a sentinel-filled matrix: the block comes back, the ldim padding stays untouched
            expB = np.full(ld * w, SENT, dt)
            getattr(O.lib, f"orc_{kind}_strided_unpack")(h, w, align, stride, _hp(exp), portion, _hp(expB), ld, es)
            dB = torch.full((ld * w,), SENT, dtype=dP.dtype, device="cuda")
            check(getattr(lib, f"hyd_{kind}_strided_unpack_{_sfx(dt)}")(h, w, align, stride, _p(dP), portion, _p(dB), ld, _s()))
            got = dB.cpu().numpy()
            assert np.array_equal(got, expB), (kind, h, w, stride, align)
            assert np.array_equal(got.reshape(w, ld)[:, :h], A.reshape(w, ld)[:, :h])


@pytest.mark.parametrize("dt", [np.float64, np.float32])
@pytest.mark.parametrize("kind", ["row", "col"])
def test_partial_strided_pack_unpack_match_oracle(kind, dt):
    """Partial{Row,Col}Strided{Pack,Unpack} (util.hpp:186-230, 464-533): the matrix holds the indices of one rank of the
    PARTIAL distribution (stride `part`); portion k collects those of union member k of the full distribution."""
    rng = np.random.default_rng(15 if kind == "row" else 16)
    es = np.dtype(dt).itemsize
    for (h, w) in CASES:
        for (uni, part) in [(1, 1), (2, 2), (4, 2), (2, 4), (3, 2)]:
            stride = uni * part
            align = int(rng.integers(0, stride))
            rank_part = int(rng.integers(0, part))
            shift_m = int(O.lib.orc_shift(rank_part, align % part, part))
            # local extent of the matrix along the partially distributed dimension
            n_dim = w if kind == "row" else h
            loc = int(O.lib.orc_length(n_dim, shift_m, part))
            lh, lw = (h, loc) if kind == "row" else (loc, w)
            ld = max(1, lh) + int(rng.integers(0, 4))
            maxl = -(-n_dim // stride)
            portion = max(1, (h if kind == "row" else w) * maxl) + int(rng.integers(0, 3))
            A = _index_matrix(ld, max(lw, 1), dt)
            exp = np.full(uni * portion, SENT, dt)
            getattr(O.lib, f"orc_partial_{kind}_strided_pack")(h, w, align, stride, uni, part, rank_part, shift_m, _hp(A), ld,
                                                                _hp(exp), portion, es)
            dA = _dev(A)
            dP = torch.full((uni * portion,), SENT, dtype=torch.from_numpy(exp).dtype, device="cuda")
            check(getattr(lib, f"hyd_partial_{kind}_strided_pack_{_sfx(dt)}")(h, w, align, stride, uni, part, rank_part, shift_m,
                                                                            _p(dA), ld, _p(dP), portion, _s()))
            assert np.array_equal(dP.cpu().numpy(), exp), (kind, h, w, uni, part, align, rank_part)
            expB = np.full(ld * max(lw, 1), SENT, dt)
            getattr(O.lib, f"orc_partial_{kind}_strided_unpack")(h, w, align, stride, uni, part, rank_part, shift_m, _hp(exp),
                                                                  portion, _hp(expB), ld, es)
            dB = torch.full((ld * max(lw, 1),), SENT, dtype=dP.dtype, device="cuda")
            check(getattr(lib, f"hyd_partial_{kind}_strided_unpack_{_sfx(dt)}")(h, w, align, stride, uni, part, rank_part,
                                                                              shift_m, _p(dP), portion, _p(dB), ld, _s()))
            got = dB.cpu().numpy()
            assert np.array_equal(got, expB), (kind, h, w, uni, part, align, rank_part)
            if lw and lh:
                assert np.array_equal(got.reshape(max(lw, 1), ld)[:lw, :lh], A.reshape(max(lw, 1), ld)[:lw, :lh])


@pytest.mark.parametrize("dt", [np.float64, np.float32])
def test_2d_strided_pack_unpack_and_interleave_match_oracle(dt):
    """StridedPack/Unpack (util.hpp:671-722; up to 8 x 4 = 32 portions = two kernel batches) and InterleaveMatrix
    (:122-146) with general strides on both sides."""
    rng = np.random.default_rng(25)
    es = np.dtype(dt).itemsize
    tdt = torch.float64 if dt == np.float64 else torch.float32
    for (h, w) in CASES:
        for (cs, rs) in [(1, 1), (2, 2), (2, 4), (8, 4), (3, 2)]:
            ca, ra = int(rng.integers(0, cs)), int(rng.integers(0, rs))
            ld = h + int(rng.integers(0, 4))
            portion = max(1, (-(-h // cs)) * (-(-w // rs))) + int(rng.integers(0, 3))
            A = _index_matrix(ld, w, dt)
            exp = np.full(cs * rs * portion, SENT, dt)
            O.lib.orc_strided_pack(h, w, ca, cs, ra, rs, _hp(A), ld, _hp(exp), portion, es)
            dA, dP = _dev(A), torch.full((cs * rs * portion,), SENT, dtype=tdt, device="cuda")
            check(getattr(lib, "hyd_strided_pack_" + _sfx(dt))(h, w, ca, cs, ra, rs, _p(dA), ld, _p(dP), portion, _s()))
            assert np.array_equal(dP.cpu().numpy(), exp), (h, w, cs, rs, ca, ra)
            expB = np.full(ld * w, SENT, dt)
            O.lib.orc_strided_unpack(h, w, ca, cs, ra, rs, _hp(exp), portion, _hp(expB), ld, es)
            dB = torch.full((ld * w,), SENT, dtype=tdt, device="cuda")
            check(getattr(lib, "hyd_strided_unpack_" + _sfx(dt))(h, w, ca, cs, ra, rs, _p(dP), portion, _p(dB), ld, _s()))
            assert np.array_equal(dB.cpu().numpy(), expB), (h, w, cs, rs, ca, ra)
    # InterleaveMatrix: every combination of unit / non-unit strides (all five kernel paths), incl. a transposing one
    for (h, w) in [(1, 1), (33, 17), (300, 70), (64, 257)]:
        for (csA, rsA, csB, rsB) in [(1, h + 1, 1, h), (2, 2 * h + 3, 1, h), (1, h, 3, 3 * h + 1), (3, 3 * h, 2, 2 * h + 2),
                                     (w + 2, 1, 1, h + 2), (w, 1, 2, 2 * h)]:
            nA = (h - 1) * csA + (w - 1) * rsA + 1
            nB = (h - 1) * csB + (w - 1) * rsB + 1
            A = _index_matrix(nA, 1, dt)
            exp = np.full(nB, SENT, dt)
            O.lib.orc_interleave(h, w, _hp(A), csA, rsA, _hp(exp), csB, rsB, es)
            dA, dB = _dev(A), torch.full((nB,), SENT, dtype=tdt, device="cuda")
            check(getattr(lib, "hyd_interleave_" + _sfx(dt))(h, w, _p(dA), csA, rsA, _p(dB), csB, rsB, _s()))
            assert np.array_equal(dB.cpu().numpy(), exp), (h, w, csA, rsA, csB, rsB)


@pytest.mark.parametrize("dt", [np.float64, np.float32])
def test_interleave_update_is_rounded_axpy(dt):
    """axpy::util::InterleaveMatrixUpdate (Axpy/util.hpp:23-51): B(i,j) += alpha*A(i,j), product and sum each rounded once."""
    rng = np.random.default_rng(35)
    tdt = torch.float64 if dt == np.float64 else torch.float32
    alpha = dt(0.7)
    for (h, w, csA, rsA, csB, rsB) in [(1, 1, 1, 1, 1, 1), (50, 33, 1, 50, 2, 101), (257, 9, 1, 260, 4, 4 * 257), (31, 300, 2, 70, 1, 31)]:
        nA = (h - 1) * csA + (w - 1) * rsA + 1
        nB = (h - 1) * csB + (w - 1) * rsB + 1
        A = rng.uniform(-1, 1, nA).astype(dt)
        B = rng.uniform(-1, 1, nB).astype(dt)
        exp = B.copy()
        ii, jj = np.meshgrid(np.arange(h), np.arange(w), indexing="ij")
        ia, ib = (ii * csA + jj * rsA).ravel(), (ii * csB + jj * rsB).ravel()
        exp[ib] = (B[ib] + (alpha * A[ia]).astype(dt)).astype(dt)
        dA, dB = _dev(A), _dev(B)
        sc = C.c_double(float(alpha)) if dt == np.float64 else C.c_float(float(alpha))
        check(getattr(lib, "hyd_interleave_update_" + _sfx(dt))(sc, h, w, _p(dA), csA, rsA, _p(dB), csB, rsB, _s()))
        assert np.array_equal(dB.cpu().numpy(), exp), (h, w)


def test_pack_argument_errors():
    z = torch.zeros(64, dtype=torch.float64, device="cuda")
    assert lib.hyd_row_strided_pack_f64(4, 4, 0, 2, _p(z), 3, _p(z), 8, _s()) != 0       # ldim < height
    assert b"leading dimension" in lib.hyd_last_error()
    assert lib.hyd_row_strided_pack_f64(4, 4, 2, 2, _p(z), 4, _p(z), 8, _s()) != 0       # align >= stride
    assert lib.hyd_col_strided_pack_f64(4, 4, 0, 2, _p(z), 4, _p(z), 7, _s()) != 0       # portionSize < 2*4
    assert b"portionSize" in lib.hyd_last_error()
    assert lib.hyd_partial_col_strided_pack_f64(4, 4, 0, 4, 2, 3, 0, 0, _p(z), 4, _p(z), 8, _s()) != 0  # union*part != stride
    assert lib.hyd_interleave_f64(-1, 4, _p(z), 1, 4, _p(z), 1, 4, _s()) != 0
    assert lib.hyd_strided_pack_f64(0, 4, 0, 2, 0, 2, _p(z), 1, _p(z), 1, _s()) == 0     # empty: nothing to do


# ---------------------------------------------------------------------------------------------------------------
# collectives on the sub-communicators
# ---------------------------------------------------------------------------------------------------------------
def _val(world_rank, i, salt=0):
    """Deterministic payload: a function of (sender's owning-comm rank, element index)."""
    return np.sin(1.0 + world_rank * 131.0 + i * 0.37 + salt)


@pytest.mark.parametrize("r,c", [(1, 1), (1, 2), (2, 2), (2, 4), (3, 2)])
def test_collectives_on_sub_communicators(r, c):
    n = 1000  # elements per member block
    idx = np.arange(n)

    def body(grid):
        res = {}
        me_w = grid.Rank()
        for comm in (H.COMM_MC, H.COMM_MR, H.COMM_VC, H.COMM_VR):
            size, me = grid.CommSize(comm), grid.CommRank(comm)
            members = [grid.CommWorldRank(comm, k) for k in range(size)]
            assert members[me] == me_w
            # all-gather
            sb = _dev(_val(me_w, idx))
            rb = torch.full((size * n,), SENT, dtype=torch.float64, device="cuda")
            grid.AllGather(comm, sb, rb)
            ag = rb.cpu().numpy()
            # all-to-all: block k of my send buffer is addressed to member k
            sb2 = _dev(np.concatenate([_val(me_w, idx, salt=10.0 * (k + 1)) for k in range(size)]))
            rb2 = torch.full((size * n,), SENT, dtype=torch.float64, device="cuda")
            grid.AllToAll(comm, sb2, rb2)
            a2a = rb2.cpu().numpy()
            # ring shift: send to me+1, receive from me-1
            rb3 = torch.full((n,), SENT, dtype=torch.float64, device="cuda")
            grid.SendRecv(comm, sb, (me + 1) % size, rb3, (me - 1) % size)
            sr = rb3.cpu().numpy()
            # broadcast from the last member
            bb = _dev(_val(me_w, idx, salt=5.0)) if me == size - 1 else torch.full((n,), SENT, dtype=torch.float64, device="cuda")
            grid.Broadcast(comm, bb, size - 1)
            bc = bb.cpu().numpy()
            # reduce-scatter, fp64 and fp32
            rs = {}
            for dt, tdt in ((np.float64, torch.float64), (np.float32, torch.float32)):
                sb4 = _dev(np.concatenate([_val(me_w, idx, salt=3.0 * (k + 1)) for k in range(size)]).astype(dt))
                rb4 = torch.full((n,), SENT, dtype=tdt, device="cuda")
                grid.ReduceScatter(comm, sb4, rb4)
                grid.ReduceScatter(comm, sb4, rb4)  # back to back on one stream: the workspace is reused in order
                rs[dt] = rb4.cpu().numpy()
            res[comm] = (size, me, members, ag, a2a, sr, bc, rs)
        return res

    out = H.run_ranks(r * c, body, height=r)
    for p in range(r * c):
        mc, mr = p % r, p // r
        for comm, (size, me, members, ag, a2a, sr, bc, rs) in out[p].items():
            exp_size = {H.COMM_MC: r, H.COMM_MR: c, H.COMM_VC: r * c, H.COMM_VR: r * c}[comm]
            exp_me = {H.COMM_MC: mc, H.COMM_MR: mr, H.COMM_VC: mc + r * mr, H.COMM_VR: mr + c * mc}[comm]
            assert (size, me) == (exp_size, exp_me)
            if comm == H.COMM_MC:
                assert members == [k + r * mr for k in range(r)]
            if comm == H.COMM_MR:
                assert members == [mc + r * k for k in range(c)]
            assert np.array_equal(ag, np.concatenate([_val(members[k], idx) for k in range(size)]))
            assert np.array_equal(a2a, np.concatenate([_val(members[k], idx, salt=10.0 * (me + 1)) for k in range(size)]))
            assert np.array_equal(sr, _val(members[(me - 1) % size], idx))
            assert np.array_equal(bc, _val(members[size - 1], idx, salt=5.0))
            for dt in (np.float64, np.float32):
                acc = _val(members[0], idx, salt=3.0 * (me + 1)).astype(dt)
                for k in range(1, size):
                    acc = (acc + _val(members[k], idx, salt=3.0 * (me + 1)).astype(dt)).astype(dt)
                assert np.array_equal(rs[dt], acc), (comm, dt)


def test_row_allgather_rebuilt_from_s2_pieces():
    """copy::RowAllGather (Copy/RowAllGather.hpp:18-145) written the reference's way -- pack, mpi::AllGather over the row
    communicator, RowStridedUnpack -- out of the S2 entry points, equals the engine's one-hop redistribution and the oracle."""
    r, c, m, n = 2, 4, 37, 53
    og = O.Grid(r, c, 1)
    A = O.DM(og, m, n, O.MC, O.MR, 1, 2).fill_index()
    B = O.DM(og, m, n, O.MC, O.STAR, 1, 0).fill_const(SENT)
    assert O.lib.orc_row_allgather(A.h, B.h) == 0

    def body(grid):
        a = U.to_rank(grid, A)
        b = H.DistMatrix(grid, m, n, H.MC, H.STAR, colAlign=1)
        lh, maxlw = a.lh, -(-n // c)
        portion = max(1, lh * maxlw)
        send = torch.zeros(portion, dtype=torch.float64, device="cuda")
        recv = torch.zeros(c * portion, dtype=torch.float64, device="cuda")
        check(lib.hyd_interleave_f64(lh, a.lw, _p(a.store), 1, a.ld, _p(send), 1, max(lh, 1), _s()))
        grid.AllGather(H.COMM_MR, send, recv)
        check(lib.hyd_row_strided_unpack_f64(lh, n, a.ra, c, _p(recv), portion, _p(b.store), b.ld, _s()))
        e = H.DistMatrix(grid, m, n, H.MC, H.STAR, colAlign=1)
        H.Copy(a, e)
        return b.local_numpy(), e.local_numpy()

    out = H.run_ranks(r * c, body, height=r)
    for p in range(r * c):
        assert np.array_equal(out[p][0], B.block(p))
        assert np.array_equal(out[p][1], B.block(p))

"""-m "not gpu": the oracle pinned by the REFERENCE's own code.

tests/golden/copy_util.npz holds outputs of the reference's copy::util pack/unpack loops, axpy::util::InterleaveMatrixUpdate
and indexing functions, produced by compiling those headers from /root/reference in place (oracle/Makefile target `ref`,
shim oracle/ref/ref_util.cpp; generator tests/golden/make_golden.py).  Here:
  * the C restatement (oracle/oracle_util.c, oracle_core.c) must reproduce every fixture bit for bit;
  * where oracle/_ref/libref_util.so exists (this container, or shipped prebuilt to the GPU box), the fixtures must still
    be what the reference computes, and oracle == reference on fresh random arguments beyond the fixture list.
"""
import itertools
import os

import numpy as np
import pytest

import golden_cases as G
import oracle_lib as O

HAVE_REF = os.path.exists(G.REF_LIB)
needs_ref = pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref/libref_util.so not built (needs /root/reference)")


def test_fixture_is_the_declared_case_list():
    cases, sha, outs, idx_in, idx_out = G.load_golden()
    assert cases == G.build_cases()
    assert sha.shape == (len(cases), 32) and len(outs) == len(cases)
    assert np.array_equal(idx_in, G.indexing_inputs())
    # every function, both dtypes, empty and ragged shapes are present
    assert {c[0] for c in cases} == set(G.FNS) and {c[1] for c in cases} == {"f64", "f32"}
    assert any(c[2][0] == 0 for c in cases) and any(c[2][1] == 0 for c in cases)


def test_oracle_restatement_reproduces_reference_pack_unpack_fixtures():
    n = G.check_against_golden(G.OracleBackend())
    assert n == len(G.build_cases())


def test_oracle_indexing_reproduces_reference_fixtures():
    _, _, _, idx_in, idx_out = G.load_golden()
    for (n, rank, align, stride), exp in zip(idx_in.tolist(), idx_out.tolist()):
        shift = O.lib.orc_shift(rank, align, stride)
        got = (shift, O.lib.orc_length(n, shift, stride), O.lib.orc_length(n, O.lib.orc_shift(rank, align, stride), stride),
               O.lib.orc_max_length(n, stride), O.lib.orc_mod(rank - align - stride, stride))
        assert got == tuple(exp[:5]), (n, rank, align, stride)


def test_product_host_indexing_reproduces_reference_fixtures():
    """hyd_mod/shift/length/max_length are host functions of the C ABI (no GPU needed)."""
    from elemental_b200 import lib
    _, _, _, idx_in, idx_out = G.load_golden()
    for (n, rank, align, stride), exp in zip(idx_in.tolist(), idx_out.tolist()):
        shift = lib.hyd_shift(rank, align, stride)
        got = (shift, lib.hyd_length(n, shift, stride), lib.hyd_max_length(n, stride), lib.hyd_mod(rank - align - stride, stride))
        assert got == (exp[0], exp[1], exp[3], exp[4]), (n, rank, align, stride)


@needs_ref
def test_fixtures_are_what_the_reference_computes():
    ref = G.RefBackend()
    assert ref.lib.ref_int_bytes() == 4
    assert G.check_against_golden(ref) == len(G.build_cases())


@needs_ref
def test_oracle_equals_reference_on_random_arguments():
    """Beyond the fixture list: 400 random argument sets per family, reference loops vs restatement, bitwise."""
    ref, orc = G.RefBackend(), G.OracleBackend()
    rng = np.random.default_rng(99)
    n = 0
    for _ in range(400):
        h, w = int(rng.integers(0, 90)), int(rng.integers(0, 90))
        dn = "f64" if rng.integers(0, 2) else "f32"
        kind = "row" if rng.integers(0, 2) else "col"
        stride = int(rng.integers(1, 10))
        align = int(rng.integers(0, stride))
        ld = max(1, h) + int(rng.integers(0, 5))
        maxl = -(-w // stride) if kind == "row" else -(-h // stride)
        portion = max(1, (h if kind == "row" else w) * maxl) + int(rng.integers(0, 4))
        cases = [(kind + "_pack", dn, (h, w, align, stride, ld, portion)), (kind + "_unpack", dn, (h, w, align, stride, ld, portion))]
        uni, part = int(rng.integers(1, 5)), int(rng.integers(1, 5))
        st2 = uni * part
        al2, rp = int(rng.integers(0, st2)), int(rng.integers(0, part))
        sm = (rp - al2 % part) % part
        nd = w if kind == "row" else h
        loc = (nd - sm - 1) // part + 1 if nd > sm else 0
        lh, lw = (h, loc) if kind == "row" else (loc, w)
        ld2 = max(1, lh) + int(rng.integers(0, 5))
        por2 = max(1, (h if kind == "row" else w) * (-(-nd // st2))) + int(rng.integers(0, 4))
        pp = (h, w, al2, st2, uni, part, rp, sm, ld2, por2, lh, lw)
        cases += [("p" + kind + "_pack", dn, pp), ("p" + kind + "_unpack", dn, pp)]
        cs, rs = int(rng.integers(1, 6)), int(rng.integers(1, 6))
        p2 = (h, w, int(rng.integers(0, cs)), cs, int(rng.integers(0, rs)), rs, ld, max(1, (-(-h // cs)) * (-(-w // rs))) + 1)
        cases += [("pack2d", dn, p2), ("unpack2d", dn, p2)]
        if h and w:
            csA, csB = int(rng.integers(1, 4)), int(rng.integers(1, 4))
            pi = (h, w, csA, csA * h + int(rng.integers(0, 3)), csB, csB * h + int(rng.integers(0, 3)))
            cases += [("interleave", dn, pi), ("update", dn, pi + (int(rng.integers(0, len(G.ALPHAS))),))]
        for (fn, d, p) in cases:
            assert np.array_equal(ref.run(fn, d, p), orc.run(fn, d, p)), (fn, d, p)
            n += 1
    assert n > 3000


@needs_ref
def test_oracle_indexing_equals_reference_exhaustively_on_small_arguments():
    ref = G.RefBackend().lib
    for stride in range(1, 13):
        for align in range(stride):
            for rank in range(stride):
                s = ref.ref_shift(rank, align, stride)
                assert s == O.lib.orc_shift(rank, align, stride)
                for n in list(range(0, 40)) + [1000, 32768]:
                    assert ref.ref_length(n, s, stride) == O.lib.orc_length(n, s, stride)
                    assert ref.ref_max_length(n, stride) == O.lib.orc_max_length(n, stride)
        for a in range(-30, 30):
            assert ref.ref_mod(a, stride) == O.lib.orc_mod(a, stride)


# ---- trapezoid level-1 routines (ScaleTrapezoid / AxpyTrapezoid / MakeTrapezoidal): tests/trapezoid_cases.py ----
def test_oracle_trapezoid_routines_reproduce_reference_fixtures():
    """oracle/oracle_level3.c against digests of the reference's own loops (tests/golden/trapezoid.json)."""
    import trapezoid_cases as T
    assert T.check_against_golden(T.OracleBackend()) == len(T.build_cases())


@needs_ref
def test_trapezoid_fixtures_are_what_the_reference_computes():
    import trapezoid_cases as T
    assert T.check_against_golden(T.RefBackend()) == len(T.build_cases())


# ---- SUMMA / Trrk loops: tests/summa_cases.py ----
def test_oracle_summa_and_trrk_reproduce_the_reference_loop_code():
    """oracle_gemm.c (SUMMA NN/NT/TN/TT x A/B/C/Dot/default, operands in four distribution sets, random alignments, four grids) and
    oracle_level3.c (Trrk, every orientation and triangle, both LocalTrrk branches) against digests of what the reference's own
    loop headers compute over stand-in matrix classes (tests/golden/summa.json): bit for bit."""
    import summa_cases as S
    assert S.check_against_golden(S.OracleBackend()) == len(S.build_cases())


@pytest.mark.skipif(not os.path.exists(os.path.join(os.path.dirname(G.REF_LIB), "libref_summa.so")),
                    reason="oracle/_ref/libref_summa.so not built (needs /root/reference)")
def test_summa_fixtures_are_what_the_reference_loops_compute():
    import summa_cases as S
    assert S.check_against_golden(S.RefBackend(), select=lambda i, c: i % 3 == 0) > 200


@pytest.mark.skipif(not os.path.exists(os.path.join(os.path.dirname(G.REF_LIB), "libref_summa.so")),
                    reason="oracle/_ref/libref_summa.so not built (needs /root/reference)")
def test_product_default_algorithm_is_the_one_the_reference_dispatcher_picks():
    """GEMM_DEFAULT: run the reference's own dispatchers (gemm::SUMMA_NN/NT/TN/TT, compiled in place) with alg = GEMM_DEFAULT and with
    each explicit variant; the variants differ in the last bits, so the digest tells which one the dispatcher took.  The
    product's host-side choice (hyd_gemm_default_alg, used by hyd_gemm_*) must be that one."""
    import summa_cases as S
    from elemental_b200 import lib
    ref = S.RefBackend()
    shapes = [(8, 8, 90), (8, 40, 90), (40, 8, 90), (12, 30, 30), (30, 12, 30), (40, 40, 40), (9, 9, 89), (9, 9, 91), (20, 40, 39),
              (20, 40, 40), (40, 20, 40), (30, 30, 59), (30, 30, 60)]
    checked = 0
    for (m, n, k), (oA, oB) in itertools.product(shapes, [(0, 0), (0, 1), (1, 0), (1, 1)]):
        base = dict(op="gemm", r=2, c=2, oA=oA, oB=oB, ds=S.DISTS[0], m=m, n=n, k=k, nb=8, nbl=64, dt="f64", al=[1, 0, 0, 1, 1, 1], uplo=0)
        by_alg = {alg: ref.run(dict(base, alg=alg)) for alg in (1, 2, 3, 4)}
        picked = [alg for alg, dg in by_alg.items() if dg == ref.run(dict(base, alg=0))]
        assert picked, (m, n, k)
        assert lib.hyd_gemm_default_alg(m, n, k) in picked, (m, n, k, oA, oB, picked)
        checked += len(picked) == 1               # the digest singles out one variant: the pick is unambiguous
    assert checked >= 16      # Dot and C are singled out; A and B produce the same bits (one local product per element, same reduce order)


# ---- the copy:: redistribution primitives, run SPMD through the reference's own headers (oracle/ref/ref_copy.cpp) ----
_COPY_LIB = os.path.join(os.path.dirname(G.REF_LIB), "libref_copy.so")
_PRIMS = {
    0: ("RowAllGather", [((O.MC, O.MR), (O.MC, O.STAR)), ((O.MR, O.MC), (O.MR, O.STAR)), ((O.STAR, O.VC), (O.STAR, O.STAR)),
                         ((O.STAR, O.MR), (O.STAR, O.STAR)), ((O.VC, O.STAR), (O.VC, O.STAR))]),
    1: ("ColAllGather", [((O.MC, O.MR), (O.STAR, O.MR)), ((O.MR, O.MC), (O.STAR, O.MC)), ((O.VC, O.STAR), (O.STAR, O.STAR)),
                         ((O.MR, O.STAR), (O.STAR, O.STAR))]),
    2: ("PartialColAllGather", [((O.VC, O.STAR), (O.MC, O.STAR)), ((O.VR, O.STAR), (O.MR, O.STAR))]),
    3: ("PartialRowAllGather", [((O.STAR, O.VC), (O.STAR, O.MC)), ((O.STAR, O.VR), (O.STAR, O.MR))]),
    4: ("ColAllToAllDemote", [((O.MC, O.MR), (O.VC, O.STAR)), ((O.MR, O.MC), (O.VR, O.STAR))]),
    5: ("RowAllToAllDemote", [((O.MC, O.MR), (O.STAR, O.VR)), ((O.MR, O.MC), (O.STAR, O.VC))]),
    6: ("ColAllToAllPromote", [((O.VC, O.STAR), (O.MC, O.MR)), ((O.VR, O.STAR), (O.MR, O.MC))]),
    7: ("RowAllToAllPromote", [((O.STAR, O.VR), (O.MC, O.MR)), ((O.STAR, O.VC), (O.MR, O.MC))]),
    8: ("ColwiseVectorExchange", [((O.VC, O.STAR), (O.VR, O.STAR)), ((O.VR, O.STAR), (O.VC, O.STAR))]),
    9: ("RowwiseVectorExchange", [((O.STAR, O.VC), (O.STAR, O.VR)), ((O.STAR, O.VR), (O.STAR, O.VC))]),
    10: ("ColFilter", [((O.STAR, O.MR), (O.MC, O.MR)), ((O.STAR, O.MC), (O.MR, O.MC)), ((O.STAR, O.STAR), (O.VC, O.STAR))]),
    11: ("RowFilter", [((O.MC, O.STAR), (O.MC, O.MR)), ((O.MR, O.STAR), (O.MR, O.MC)), ((O.STAR, O.STAR), (O.STAR, O.VR))]),
    12: ("PartialColFilter", [((O.MC, O.STAR), (O.VC, O.STAR)), ((O.MR, O.STAR), (O.VR, O.STAR))]),
    13: ("PartialRowFilter", [((O.STAR, O.MC), (O.STAR, O.VC)), ((O.STAR, O.MR), (O.STAR, O.VR))]),
    14: ("Translate", [((u, v), (u, v)) for (u, v) in O.PAIRS]),
}


@pytest.mark.skipif(not os.path.exists(_COPY_LIB), reason="oracle/_ref/libref_copy.so not built (needs /root/reference)")
@pytest.mark.parametrize("prim", sorted(_PRIMS))
def test_reference_copy_primitives_realise_the_closed_form_definition(prim):
    """The oracle (and through it every parity test) defines B = A by the ownership formulas (orc_redist_closed).  Here the
    reference's own copy:: primitive -- pack loops, collectives, the unaligned SendRecv branches, the width/height-1 broadcast
    branches -- runs SPMD, one thread per simulated rank, and must produce exactly that: every rank's block of B holds the
    entries of A the formulas give it (gather to [*,*] bitwise equal, replicas consistent), for constrained (possibly
    unaligned) and unconstrained destinations.  With an unconstrained destination the alignments it ends with must be the ones
    orc_copy_adopts predicts -- the rule the shim, the library's read proxies and the oracle's proxies share."""
    import ctypes as C
    ref = C.CDLL(_COPY_LIB)
    ref.ref_copy_primitive.restype = C.c_int
    ref.ref_copy_primitive.argtypes = [C.c_int, C.c_void_p, C.POINTER(C.c_void_p), C.c_int, C.c_int]
    create = O.lib.orc_dm_create
    create.restype = C.c_void_p

    class Head(C.Structure):
        _fields_ = [("g", O.Grid), ("colDist", C.c_int), ("rowDist", C.c_int), ("colAlign", C.c_int), ("rowAlign", C.c_int)]

    rng = np.random.default_rng(500 + prim)
    name, pairs = _PRIMS[prim]
    n = 0
    for (r, c) in [(1, 1), (2, 2), (2, 3), (4, 2)]:
        g = O.Grid(r, c, 1)
        al = lambda d: int(rng.integers(0, O.lib.orc_dist_stride(g, d)))
        for (s, d), (m, w), con in itertools.product(pairs, [(1, 1), (7, 5), (50, 3), (33, 1), (1, 29), (40, 41)], (0, 1)):
            A = O.DM(g, m, w, s[0], s[1], al(s[0]), al(s[1])).fill(1)
            b0 = (al(d[0]), al(d[1]))
            slot = C.c_void_p(create(g, 3, 2, d[0], d[1], b0[0], b0[1], 8))
            assert ref.ref_copy_primitive(prim, A.h, C.byref(slot), con, con) == 0, name
            head = Head.from_address(slot.value)
            got_align = (head.colAlign, head.rowAlign)
            B = O.DM(g, m, w, d[0], d[1], got_align[0], got_align[1], handle=slot.value)
            GB, mism = B.to_global()
            GA, _ = A.to_global()
            assert mism == 0 and np.array_equal(GB, GA), (name, (r, c), s, d, (m, w), con)
            if con:
                assert got_align == b0
            else:
                exp = []
                for sd, sa, dd, init in ((s[0], A.ca, d[0], b0[0]), (s[1], A.ra, d[1], b0[1])):
                    a = C.c_int(0)
                    exp.append(a.value if O.lib.orc_copy_adopts(sd, sa, dd, O.lib.orc_dist_stride(g, dd), C.byref(a)) else init)
                assert got_align == tuple(exp), (name, (r, c), s, d, got_align, exp)
            n += 1
    assert n >= 96


@pytest.mark.skipif(not os.path.exists(_COPY_LIB), reason="oracle/_ref/libref_copy.so not built (needs /root/reference)")
def test_reference_axpy_contract_realises_the_closed_form_definition():
    """B += alpha * (sum of A over the ranks it is replicated / partially summed along): the reference's own scatter routines
    (AxpyContract.hpp: PartialColScatter, PartialRowScatter, ColScatter, RowScatter, Scatter), run SPMD over the stand-ins, against
    orc_axpy_contract_closed -- the definition behind SUMMA-A/B/Dot in the oracle.  Every rank's contribution is a different
    small-integer block, so all sums are exact and the comparison is bitwise whatever order a reduction takes.  Only cases the
    reference implements: its partial scatters exist for aligned operands only ("Unaligned PartialColScatter not implemented"),
    and the unaligned branch of its ColScatter packs B's local width out of A (AxpyContract.hpp:232-240), which overruns the
    buffers when the two differ; the library takes any alignment (tests/test_gpu_redist.py::test_axpy_contract)."""
    import ctypes as C
    ref = C.CDLL(_COPY_LIB)
    ref.ref_axpy_contract.restype = C.c_int
    ref.ref_axpy_contract.argtypes = [C.c_double, C.c_void_p, C.POINTER(C.c_void_p)]
    create = O.lib.orc_dm_create
    create.restype = C.c_void_p
    MC, MR, VC, VR, STAR = O.MC, O.MR, O.VC, O.VR, O.STAR
    pairs = [((MC, MR), (MC, MR)), ((MC, STAR), (VC, STAR)), ((MR, STAR), (VR, STAR)), ((STAR, MC), (STAR, VC)), ((STAR, MR), (STAR, VR)),
             ((STAR, MR), (MC, MR)), ((STAR, MC), (MR, MC)), ((STAR, STAR), (VC, STAR)), ((MC, STAR), (MC, MR)), ((MR, STAR), (MR, MC)),
             ((STAR, STAR), (MC, MR)), ((STAR, STAR), (MR, MC)), ((STAR, STAR), (STAR, VR))]
    part = {VC: MC, VR: MR}
    rng = np.random.default_rng(600)
    n = 0
    for (r, c) in [(1, 1), (2, 2), (2, 3), (4, 2)]:
        g = O.Grid(r, c, 1)
        al = lambda d: int(rng.integers(0, O.lib.orc_dist_stride(g, d)))
        for (s, d), (m, w), trial in itertools.product(pairs, [(1, 1), (7, 5), (50, 3), (33, 1), (1, 29), (40, 41)], (0, 1)):
            sa, da = [al(s[0]), al(s[1])], [al(d[0]), al(d[1])]
            for i in range(2):
                if s[i] == d[i] and (trial == 0 or s == d or (i == 1 and s[0] == STAR and d[0] != STAR)):
                    sa[i] = da[i]
                if part.get(d[i]) == s[i] and s[i] != d[i]:
                    sa[i] = da[i] % O.lib.orc_dist_stride(g, s[i])

            def operand():
                A = O.DM(g, m, w, s[0], s[1], sa[0], sa[1])
                for p in range(r * c):
                    blk = A.block(p)
                    if blk.size:
                        blk[...] = np.random.default_rng(1000 + p).integers(-8, 9, size=blk.shape)
                return A

            B0 = np.random.default_rng(7).integers(-8, 9, size=(m, w)).astype(np.float64)
            A, B1 = operand(), O.DM(g, m, w, d[0], d[1], da[0], da[1]).from_global(B0)
            O.lib.orc_axpy_contract_closed(2.0, A.h, B1.h, 0)
            G1, m1 = B1.to_global()
            G1 = G1.copy()
            A = operand()
            slot = C.c_void_p(create(g, m, w, d[0], d[1], da[0], da[1], 8))
            B2 = O.DM(g, m, w, d[0], d[1], da[0], da[1], handle=slot.value).from_global(B0)
            assert ref.ref_axpy_contract(2.0, A.h, C.byref(slot)) == 0, (s, d, sa, da)
            G2, m2 = B2.to_global()
            assert m1 == 0 and m2 == 0 and np.array_equal(G1, G2), ((r, c), s, d, (m, w), sa, da)
            n += 1
    assert n == 624


@pytest.mark.skipif(not os.path.exists(os.path.join(os.path.dirname(G.REF_LIB), "libref_summa.so")),
                    reason="oracle/_ref/libref_summa.so not built (needs /root/reference)")
def test_oracle_equals_reference_loops_on_random_cases():
    """Beyond the fixture list: random operation (Gemm with any algorithm, Trrk, Syrk, Symm), grid, sizes 0..69 (empty operands,
    blocksize above and below k), any of the eleven distributions for each operand, random alignments, fp64 / fp32 -- the
    restatement and the reference's loop code must still agree bit for bit.  (2 900 such cases were run once with other seeds.)"""
    import summa_cases as S
    ref, orc = S.RefBackend(), S.OracleBackend()
    rng = np.random.default_rng(700)
    pairs = [(O.MC, O.MR), (O.MR, O.MC), (O.MC, O.STAR), (O.STAR, O.MR), (O.MR, O.STAR), (O.STAR, O.MC), (O.VC, O.STAR), (O.STAR, O.VC),
             (O.VR, O.STAR), (O.STAR, O.VR), (O.STAR, O.STAR)]
    for _ in range(200):
        r, c = [(1, 1), (2, 2), (2, 3), (4, 2), (3, 1), (1, 4)][int(rng.integers(0, 6))]
        g = O.Grid(r, c, 1)
        op = ["gemm", "gemm", "trrk", "syrk", "symm"][int(rng.integers(0, 5))]
        ds = tuple(pairs[int(rng.integers(0, len(pairs)))] for _ in range(3))
        m, n, k = (int(x) for x in rng.integers(0, 70, size=3))
        oA, oB = int(rng.integers(0, 2)), int(rng.integers(0, 2))
        case = dict(op=op, r=r, c=c, oA=oA, oB=oB, alg=int(rng.integers(0, 5)), ds=ds, m=m, n=n, k=k, nb=int(rng.integers(1, 40)),
                    nbl=[4, 8, 64][int(rng.integers(0, 3))], dt=["f64", "f32"][int(rng.integers(0, 2))], uplo=int(rng.integers(0, 2)))
        if op in ("trrk", "syrk"):
            case["m"] = n
        if op == "syrk":
            case["oB"] = 1 - oA
        if op == "symm":
            case["oB"], case["k"] = 0, (m if oA == 0 else n)
        case["al"] = [int(rng.integers(0, O.lib.orc_dist_stride(g, d))) for pair in ds for d in pair]
        assert ref.run(case) == orc.run(case), case

"""CPU: the C++ shim's distribution algebra and alignment rules (include/El.hpp) against the reference.

* Collect / Partial / PartialUnionRow / PartialUnionCol and the enumerator values are compared with tables produced by
  the reference's own include/El/core/types.hpp (tests/golden/dist_algebra.json, written by tests/golden/make_golden.py
  from oracle/_ref/libref_util.so; compared live as well where that library exists).
* AlignWith (ElementMatrix.cpp:243-300) is checked at every call site of the reference's SUMMA (Gemm/{NN,NT,TN,TT}.hpp
  `X.AlignWith(Y)`), by the ownership property it exists for, and for the "Nonsensical alignment" cases.
* The alignment an unconstrained B adopts in B = A is checked against the AlignColsAndResize / AlignRowsAndResize argument
  of each copy:: primitive (include/El/blas_like/level1/Copy/*.hpp).
"""
import ctypes
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MC, MD, MR, VC, VR, STAR, CIRC = range(7)
R, C = 2, 3          # the grid align_check.cpp uses
CA, RA = 5, 4        # the DistData alignments align_check.cpp uses


@pytest.fixture(scope="module")
def dump(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("align") / "align_check")
    subprocess.check_call(["g++", "-std=c++17", "-O0", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"),
                           "-I/usr/local/cuda/include", os.path.join(ROOT, "tests", "align_check.cpp"), "-o", exe])
    return json.loads(subprocess.check_output([exe]))


def _golden():
    with open(os.path.join(ROOT, "tests", "golden", "dist_algebra.json")) as f:
        return json.load(f)


def test_dist_algebra_matches_reference_tables(dump):
    g = _golden()
    for key in ("enum", "collect", "partial", "partial_union_row", "partial_union_col"):
        assert dump[key] == g[key], key


def test_dist_algebra_matches_reference_live(dump):
    path = os.path.join(ROOT, "oracle", "_ref", "libref_util.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/libref_util.so not built (no /root/reference here)")
    ref = ctypes.CDLL(path)
    assert dump["collect"] == [ref.ref_dist_collect(d) for d in range(7)]
    assert dump["partial"] == [ref.ref_dist_partial(d) for d in range(7)]
    assert dump["partial_union_row"] == [ref.ref_dist_partial_union_row(u, v) for u in range(7) for v in range(7)]
    assert dump["partial_union_col"] == [ref.ref_dist_partial_union_col(u, v) for u in range(7) for v in range(7)]
    names = {"MC": 0, "MD": 1, "MR": 2, "VC": 3, "VR": 4, "STAR": 5, "CIRC": 6, "NORMAL": 10, "TRANSPOSE": 11, "ADJOINT": 12,
             "ROW_MAJOR": 20, "COLUMN_MAJOR": 21, "ELEMENT": 30, "BLOCK": 31}
    assert dump["enum"] == {k: ref.ref_enum_value(v) for k, v in names.items()}


def _aw(dump, u, v, dc, dr):
    return dump["align_with"][((u * 7 + v) * 7 + dc) * 7 + dr]


# every X.AlignWith(Y) in src/blas_like/level3/Gemm/{NN,NT,TN,TT}.hpp: X's distribution, Y's, and the alignments X must end
# with when Y is aligned (CA, RA) -- None where the dimension is not distributed
SUMMA_SITES = [
    ((MC, STAR), (MC, MR), CA, None),     # A1_MC_STAR.AlignWith(C)          NN.hpp, NT.hpp
    ((MR, STAR), (MC, MR), RA, None),     # B1Trans_MR_STAR.AlignWith(C)     NN.hpp; D1Trans_MR_STAR.AlignWith(B)
    ((STAR, MR), (MC, MR), None, RA),     # B1Trans_STAR_MR.AlignWith(A)     NN.hpp, NT.hpp, TT.hpp
    ((STAR, MC), (MC, MR), None, CA),     # A1_STAR_MC.AlignWith(B)          NN.hpp, TN.hpp, TT.hpp
    ((MC, MR), (MC, MR), CA, RA),         # B1_MCMR.AlignWith(C)             NN.hpp
    ((VC, STAR), (MC, MR), CA, None),     # B1_VC_STAR.AlignWith(A)          NN.hpp (dot): VC rows land in A's MC row
    ((VR, STAR), (MC, MR), RA, None),     # B1_VR_STAR.AlignWith(A)          NN.hpp, NT.hpp, TT.hpp
]


@pytest.mark.parametrize("mine,other,col,row", SUMMA_SITES)
def test_alignwith_at_the_summa_call_sites(dump, mine, other, col, row):
    k0, a0, k1, a1 = _aw(dump, mine[0], mine[1], other[0], other[1])
    if col is None:       # an undistributed dimension: untouched, or "aligned" to 0 through the PartialUnion branch
        assert k0 == 0 or (k0, a0) == (1, 0)
    else:
        assert (k0, a0) == (1, col)
    if row is None:
        assert k1 == 0 or (k1, a1) == (1, 0)
    else:
        assert (k1, a1) == (1, row)


def _coords(dist, align, i):
    """(grid row, grid col) owning index i of a dimension distributed `dist` with alignment `align`; None = not fixed."""
    p = R * C
    if dist == MC:
        return ((i + align) % R, None)
    if dist == MR:
        return (None, (i + align) % C)
    if dist == VC:
        vc = (i + align) % p
        return (vc % R, vc // R)
    if dist == VR:
        vr = (i + align) % p
        return (vr // C, vr % C)
    return (None, None)


def test_alignwith_makes_related_dimensions_share_owners(dump):
    """The point of AlignWith: once X is aligned with Y, the grid coordinate both dimensions depend on agrees for every
    index (so the SUMMA panels need no realignment).  Checked for every related pair the first two branches accept."""
    partial = {MC: MC, MR: MR, VC: MC, VR: MR}
    checked = 0
    for mine in (MC, MR, VC, VR):
        for src in {mine, partial[mine]}:
            # column dimension of [mine,STAR] aligned with DistData {src, CA % stride, STAR, 0}: use the dump's CA directly
            k0, a0, _, _ = _aw(dump, mine, STAR, src, STAR)
            assert k0 == 1
            for i in range(40):
                me, you = _coords(mine, a0, i), _coords(src, CA, i)
                for x, y in zip(me, you):
                    if x is not None and y is not None:
                        assert x == y
                        checked += 1
    assert checked > 0


def test_alignwith_nonsensical_and_untouched_cases(dump):
    # [MC,MR] aligned with a [VC,VR]-like DistData: neither dimension relates, and none is STAR -> LogicError in the reference
    k0, _, k1, _ = _aw(dump, MC, MR, VC, VR)
    assert (k0, k1) == (-1, -1)
    # [MC,MR] aligned with [VC,STAR]: no relation, but the STAR dimension makes it legal: alignment untouched
    k0, _, k1, _ = _aw(dump, MC, MR, VC, STAR)
    assert (k0, k1) == (0, 0)
    # a STAR dimension never aligns and never complains
    for dc in range(7):
        for dr in range(7):
            k0, _, _, _ = _aw(dump, STAR, MR, dc, dr)
            assert k0 in (0, 1) and (k0 == 0 or STAR in (dc, dr))


# the alignment argument of the copy:: primitive that implements dst <- src for one dimension, with the source aligned 5:
# (src, dst) -> expected alignment, or None when the primitive leaves the destination's alignment alone
COPY_ADOPTS = {
    (MC, MC): 5, (MR, MR): 5, (VC, VC): 5, (VR, VR): 5,            # Translate / RowAllGather / ColFilter ...: A.ColAlign()
    (VC, MC): 5 % R, (VR, MR): 5 % C,                              # Partial*AllGather, *AllToAllPromote: Mod(align, stride)
    (MC, VC): 5, (MR, VR): 5,                                      # *AllToAllDemote, Partial*Filter: A.ColAlign()
    (MC, MR): None, (MR, MC): None, (VC, VR): None, (VR, VC): None, (MC, VR): None, (MR, VC): None, (VC, MR): None, (VR, MC): None,
    (STAR, MC): None, (MC, STAR): None, (STAR, STAR): None, (VC, STAR): None, (STAR, VR): None, (CIRC, MC): None, (MC, CIRC): None,
}


def test_copy_adopts_the_alignment_each_primitive_passes(dump):
    for (s, d), want in COPY_ADOPTS.items():
        k, a = dump["copy_adopts"][s * 7 + d]
        if want is None:
            assert k == 0, (s, d)
        else:
            assert (k, a) == (1, want), (s, d)


def test_alignment_rules_match_the_oracle_restatement(dump):
    """All 7^4 (matrix distribution, DistData) combinations of AlignWith and all 49 copy pairs: the shim (El.hpp) against
    the oracle's table-driven restatement (oracle_core.c orc_align_dim_with / orc_copy_adopts)."""
    import ctypes
    from oracle_lib import lib as orc
    for d in range(7):
        assert dump["collect"][d] == orc.orc_dist_collect(d) and dump["partial"][d] == orc.orc_dist_partial(d)
    stride = {MC: R, MR: C, VC: R * C, VR: R * C, MD: 6, STAR: 1, CIRC: 1}
    a = ctypes.c_int(-1)
    for u in range(7):
        for v in range(7):
            assert dump["partial_union_row"][u * 7 + v] == orc.orc_dist_partial_union_row(u, v)
            assert dump["partial_union_col"][u * 7 + v] == orc.orc_dist_partial_union_col(u, v)
            for dc in range(7):
                for dr in range(7):
                    got = _aw(dump, u, v, dc, dr)
                    for which, st in ((0, stride[u]), (1, stride[v])):
                        a.value = -1
                        k = orc.orc_align_dim_with(u, v, which, st, dc, CA, dr, RA, ctypes.byref(a))
                        assert (k, a.value if k == 1 else -1) == (got[2 * which], got[2 * which + 1] if k == 1 else -1), (u, v, dc, dr, which)
    for s in range(7):
        for d in range(7):
            a.value = -1
            k = orc.orc_copy_adopts(s, 5, d, stride[d], ctypes.byref(a))
            got = dump["copy_adopts"][s * 7 + d]
            assert (k, a.value if k else -1) == (got[0], got[1] if k else -1), (s, d)

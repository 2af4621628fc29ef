"""-m gpu: the CUDA pack/unpack path (seam S2 through the C ABI) against the committed golden vectors.

The fixtures (tests/golden/copy_util.npz) are outputs of the REFERENCE's own copy::util loops and of
axpy::util::InterleaveMatrixUpdate, compiled from /root/reference in place (tests/golden/make_golden.py), so this is
parity with reference code rather than with the restatement: every portion, every untouched sentinel, bit for bit.
InterleaveMatrixUpdate: the reference's CPU form rounds alpha*a and the sum separately (?axpy); the kernel does the same
(no FMA contraction), so it too is compared bitwise."""
import pytest

import golden_cases as G

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("family", ["row", "col", "prow", "pcol", "2d", "interleave", "update"])
def test_cuda_pack_unpack_reproduces_reference_fixtures(family):
    be = G.CudaBackend()
    fns = {"row": ("row_pack", "row_unpack"), "col": ("col_pack", "col_unpack"), "prow": ("prow_pack", "prow_unpack"),
           "pcol": ("pcol_pack", "pcol_unpack"), "2d": ("pack2d", "unpack2d"), "interleave": ("interleave",),
           "update": ("update",)}[family]
    n = G.check_against_golden(be, select=lambda fn, dn, p: fn in fns)
    assert n > 0

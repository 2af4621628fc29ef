"""-m gpu: hyd_local_trrk_static_f64, the allocation-free variant of the triangular local update (scratch in a file-scope
static __device__ array; include/hydrogen_b200.h).  No counterpart in the reference: added
for callers that cannot call cudaMalloc.

Checked through the C ABI against the EXISTING entry point hyd_trrk_f64 (a 1x1 grid, one k-panel: the same local update with
the growable workspace) on two calls in a row -- different shapes and triangles, so the second call finds the first one's
contents in the shared scratch.  Tolerance: the Gemm bound of test_gpu_gemm.py / test_gpu_trrk.py, relative Frobenius error
<= 1e-13 * sqrt(k) inside the triangle; outside it every entry keeps its bits.  Also: the library's own allocation counter does
not move, a strip beyond the static scratch is rejected with a clear error, and the strided (local block of an r x c grid) form
agrees with a dense torch product.

Collected last on purpose; passes alone in a fresh process and after the other tests."""
import ctypes as C

import pytest
import torch

from elemental_b200 import host as H
from elemental_b200 import lib

pytestmark = pytest.mark.gpu


def _colmajor(rows, cols, gen):
    """torch (cols, rows) row-major == column-major rows x cols with ld = rows"""
    return torch.rand((cols, rows), dtype=torch.float64, device="cuda", generator=gen) * 2 - 1


def _call_static(uplo, tA, tB, lh, lw, k, cs, cst, rs, rst, alpha, A, lda, B, ldb, beta, Cm, ldc):
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    return lib.hyd_local_trrk_static_f64(uplo, tA, tB, lh, lw, k, cs, cst, rs, rst, C.c_double(alpha), C.c_void_p(A.data_ptr()), lda,
                                         C.c_void_p(B.data_ptr()), ldb, C.c_double(beta), C.c_void_p(Cm.data_ptr()), ldc, st)


def test_static_scratch_trrk_matches_the_existing_entry_point_twice_in_a_row():
    torch.cuda.set_device(0)
    grid = H.Grid(H.World.single(), 1)
    gen = torch.Generator(device="cuda").manual_seed(42)
    try:
        # two calls in a row: (uplo, transA, transB, n, k); the second is smaller and the other triangle, so stale scratch
        # contents of the first call lie under and around its strips
        for (uplo, tA, tB, n, k) in ((H.LOWER, 0, 1, 1501, 700), (H.UPPER, 1, 0, 1100, 320)):
            A = H.DistMatrix(grid, k if tA else n, n if tA else k, H.MC, H.MR)
            B = H.DistMatrix(grid, n if tB else k, k if tB else n, H.MC, H.MR)
            C1 = H.DistMatrix(grid, n, n, H.MC, H.MR)
            C2 = H.DistMatrix(grid, n, n, H.MC, H.MR)
            for d in (A, B, C1):
                d.store.copy_(torch.rand(d.store.shape, dtype=torch.float64, device="cuda", generator=gen) * 2 - 1)
            C2.store.copy_(C1.store)
            C0 = C1.store.clone()
            # existing entry point: blocksize >= k, so the whole update is one panel
            H.Trrk(uplo, tA, tB, 2.0, A, B, 3.0, C1, blocksize=4096)
            torch.cuda.synchronize()
            a0 = lib.hyd_alloc_count()
            rc = _call_static(uplo, tA, tB, n, n, k, 0, 1, 0, 1, 2.0, A.store, A.ld, B.store, B.ld, 3.0, C2.store, C2.ld)
            assert rc == 0, lib.hyd_last_error().decode()
            torch.cuda.synchronize()
            assert lib.hyd_alloc_count() == a0          # the library allocated nothing
            got, want, old = C2.store[:n, :n].t(), C1.store[:n, :n].t(), C0[:n, :n].t()     # (row, col) views
            i, j = torch.meshgrid(torch.arange(n, device="cuda"), torch.arange(n, device="cuda"), indexing="ij")
            inside = (i >= j) if uplo == H.LOWER else (i <= j)
            rel = float((got - want)[inside].norm() / want[inside].norm())
            assert rel <= 1e-13 * k ** 0.5, (uplo, tA, tB, rel)
            assert torch.equal(got[~inside], old[~inside])       # outside the triangle: untouched, bit for bit
            assert torch.equal(C2.store[:, n:], C0[:, n:])          # ldim padding untouched
    finally:
        grid.destroy()


def test_static_scratch_trrk_local_block_of_a_grid_and_size_limit():
    torch.cuda.set_device(0)
    gen = torch.Generator(device="cuda").manual_seed(7)
    # the local block of rank (1, 2) of a 2 x 3 grid: global (1 + 2 i, 2 + 3 j); LOWER triangle
    lh, lw, k, cs, cst, rs, rst = 700, 640, 96, 1, 2, 2, 3
    A, B, Cm = _colmajor(lh, k, gen), _colmajor(k, lw, gen), _colmajor(lh, lw, gen)
    C0 = Cm.clone()
    assert _call_static(0, 0, 0, lh, lw, k, cs, cst, rs, rst, 1.5, A, lh, B, k, 0.5, Cm, lh) == 0, lib.hyd_last_error().decode()
    torch.cuda.synchronize()
    gi = (cs + cst * torch.arange(lh, device="cuda"))[:, None]
    gj = (rs + rst * torch.arange(lw, device="cuda"))[None, :]
    inside = gi >= gj
    want = torch.where(inside, 1.5 * (A.t() @ B.t()) + 0.5 * C0.t(), C0.t())
    assert float((Cm.t() - want).norm() / want.norm()) <= 1e-13 * k ** 0.5
    assert torch.equal(Cm.t()[~inside], C0.t()[~inside])
    # a strip beyond the static scratch: 512 local columns 64 global columns apart on unit row stride -> ~32 k x 512 doubles
    lh2, lw2 = 40000, 512
    A2, B2, C2 = _colmajor(lh2, 8, gen), _colmajor(8, lw2, gen), _colmajor(lh2, lw2, gen)
    keep = C2.clone()
    assert _call_static(0, 0, 0, lh2, lw2, 8, 0, 1, 0, 64, 1.0, A2, lh2, B2, 8, 2.0, C2, lh2) != 0
    msg = lib.hyd_last_error().decode()
    assert "beyond the static scratch" in msg and "hyd_trrk_f64" in msg, msg
    torch.cuda.synchronize()
    assert torch.equal(C2, keep)                                   # rejected before anything was launched
    assert _call_static(2, 0, 0, 8, 8, 8, 0, 1, 0, 1, 1.0, A2, 8, B2, 8, 1.0, C2, 8) != 0   # bad uplo

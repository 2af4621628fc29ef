"""-m gpu parity of the local (per-rank) kernels behind seam S2, through the C ABI.

GEMM: against the oracle's long-double dense product (small) and a plain torch product in the
same precision class (large), tolerance = BASELINE.json's relative Frobenius bound
(1e-13 fp64 / 1e-5 fp32, sqrt(k)-scaled) and the reference's own |x-y| bound
(tests/blas_like/BasicGemm.cpp:93-94: 1e-10 fp64, 1e-3 fp32 at alpha=2, beta=3).
Transpose / copy: bit-exact.  Axpy / scale: exact up to FMA contraction (1 ulp of the product)."""
import ctypes as C
import itertools

import numpy as np
import pytest
import torch

import oracle_lib as O
from elemental_b200 import lib, check

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _colmajor(rows, cols, ld, dtype, gen):
    t = torch.rand((max(cols, 1), ld), dtype=dtype, device=DEV, generator=gen) * 2 - 1
    return t


def _run_gemm(dtype, tA, tB, m, n, k, alpha, beta, pad, check_oracle):
    tdt = torch.float64 if dtype == np.float64 else torch.float32
    gen = torch.Generator(device=DEV).manual_seed(1000 * m + 10 * n + k)
    ar, ac = (k, m) if tA else (m, k)
    br, bc = (n, k) if tB else (k, n)
    q = 16 // np.dtype(dtype).itemsize
    rnd = lambda x: max(q, -(-x // q) * q) + pad
    lda, ldb, ldc = rnd(ar), rnd(br), rnd(m)
    A, B, Cm = _colmajor(ar, ac, lda, tdt, gen), _colmajor(br, bc, ldb, tdt, gen), _colmajor(m, n, ldc, tdt, gen)
    out = Cm.clone()
    fn = lib.hyd_lgemm_f64 if dtype == np.float64 else lib.hyd_lgemm_f32
    sc = C.c_double if dtype == np.float64 else C.c_float
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    check(fn(tA, tB, m, n, k, sc(alpha), C.c_void_p(A.data_ptr()), lda, C.c_void_p(B.data_ptr()), ldb, sc(beta),
             C.c_void_p(out.data_ptr()), ldc, s))
    torch.cuda.synchronize()
    got = out[:n, :m].t().double()
    # padding rows of C (between m and ldc) must not be written
    assert torch.equal(out[:, m:], Cm[:, m:])
    Am, Bm = A[:ac, :ar].t().double(), B[:bc, :br].t().double()
    opA, opB = (Am.t() if tA else Am), (Bm.t() if tB else Bm)
    if check_oracle:
        # oracle's dense long-double reference (oracle/oracle_gemm.c: orc_dense_gemm_ref)
        a = np.asfortranarray(A[:ac, :ar].t().cpu().numpy())
        b = np.asfortranarray(B[:bc, :br].t().cpu().numpy())
        exp = np.asfortranarray(Cm[:n, :m].t().cpu().numpy())
        O.lib.orc_dense_gemm_ref(tA, tB, m, n, k, alpha, a.ctypes.data, ar, b.ctypes.data, br, beta, exp.ctypes.data, m,
                                 np.dtype(dtype).itemsize)
        ref = torch.from_numpy(exp.astype(np.float64)).to(DEV)
    else:
        ref = alpha * (opA @ opB) + beta * Cm[:n, :m].t().double()
    rel = ((got - ref).norm() / ref.norm().clamp_min(1e-300)).item()
    tol = (1e-13 if dtype == np.float64 else 1e-5) * max(1.0, k) ** 0.5
    absmax = (got - ref).abs().max().item()
    assert rel <= tol, (tA, tB, m, n, k, rel, tol)
    if k <= 1000:
        assert absmax <= (1e-10 if dtype == np.float64 else 1e-3), absmax


SHAPES_SMALL = [(1, 1, 1), (128, 128, 16), (7, 5, 3), (129, 131, 17), (256, 384, 64), (300, 200, 100), (64, 1000, 33)]


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("tA,tB", list(itertools.product((0, 1), (0, 1))))
def test_lgemm_small_vs_oracle(dtype, tA, tB):
    for (m, n, k) in SHAPES_SMALL:
        for pad in (0, 1):  # pad=1 -> odd leading dimensions (TMA-incompatible -> generic path)
            _run_gemm(dtype, tA, tB, m, n, k, 2.0, 3.0, pad * 3, True)


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("tA,tB", list(itertools.product((0, 1), (0, 1))))
def test_lgemm_large_vs_torch(dtype, tA, tB):
    _run_gemm(dtype, tA, tB, 1000, 1000, 1000, 2.0, 3.0, 0, False)  # BasicGemm.cpp:245-247 default size
    _run_gemm(dtype, tA, tB, 2048 + 40, 1024 + 24, 512 + 8, 1.0, 0.0, 0, False)
    _run_gemm(dtype, tA, tB, 4096, 2048, 128, 1.0, 1.0, 0, False)  # SUMMA-C rank-nb update shape


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_lgemm_repeatable_under_copy_traffic(dtype):
    """Regression for the stage-release race of the TMA ring (a queued ld.shared overtaken by the consumer's
    mbarrier.arrive read the slot's NEXT fill: one wrong 8x32 block per ~1e13 flops, more often with copy
    traffic on the device).  The kernel is deterministic, so repeated launches must agree bit for bit."""
    tdt = torch.float64 if dtype == np.float64 else torch.float32
    fn = lib.hyd_lgemm_f64 if dtype == np.float64 else lib.hyd_lgemm_f32
    sc = C.c_double if dtype == np.float64 else C.c_float
    m = n = 8192
    k = 4096
    gen = torch.Generator(device=DEV).manual_seed(77)
    A = torch.rand((k, m), dtype=tdt, device=DEV, generator=gen) * 2 - 1
    B = torch.rand((n, k), dtype=tdt, device=DEV, generator=gen) * 2 - 1
    outs = [torch.empty((n, m), dtype=tdt, device=DEV) for _ in range(2)]
    src = torch.empty(64 << 20, dtype=torch.uint8, pin_memory=True)
    dst = torch.empty(64 << 20, dtype=torch.uint8, device=DEV)
    side = torch.cuda.Stream()
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    first = None
    for it in range(6):
        with torch.cuda.stream(side):
            for _ in range(4):
                dst.copy_(src, non_blocking=True)
        o = outs[it & 1]
        check(fn(0, 0, m, n, k, sc(1.0), C.c_void_p(A.data_ptr()), m, C.c_void_p(B.data_ptr()), k, sc(0.0),
                 C.c_void_p(o.data_ptr()), m, s))
        torch.cuda.synchronize()
        if first is None:
            first = o.clone()
        else:
            assert torch.equal(o, first), f"launch {it}: {(o != first).sum().item()} elements differ from launch 0"


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_lgemm_edge_cases(dtype):
    tdt = torch.float64 if dtype == np.float64 else torch.float32
    fn = lib.hyd_lgemm_f64 if dtype == np.float64 else lib.hyd_lgemm_f32
    sc = C.c_double if dtype == np.float64 else C.c_float
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    Cm = torch.rand((8, 8), dtype=tdt, device=DEV)
    A = torch.rand((8, 8), dtype=tdt, device=DEV)
    ref = Cm.clone()
    # k == 0 -> C := beta C (Gemm.cpp:212-219)
    check(fn(0, 0, 8, 8, 0, sc(2.0), C.c_void_p(A.data_ptr()), 8, C.c_void_p(A.data_ptr()), 8, sc(3.0), C.c_void_p(Cm.data_ptr()), 8, s))
    torch.cuda.synchronize()
    assert torch.equal(Cm, ref * 3.0)
    # m == 0 / n == 0 -> no-op
    check(fn(0, 0, 0, 8, 8, sc(2.0), C.c_void_p(A.data_ptr()), 8, C.c_void_p(A.data_ptr()), 8, sc(3.0), C.c_void_p(Cm.data_ptr()), 8, s))
    # beta == 0 must overwrite NaNs in C
    Cm.fill_(float("nan"))
    check(fn(0, 0, 8, 8, 8, sc(1.0), C.c_void_p(A.data_ptr()), 8, C.c_void_p(A.data_ptr()), 8, sc(0.0), C.c_void_p(Cm.data_ptr()), 8, s))
    torch.cuda.synchronize()
    assert torch.isfinite(Cm).all()
    # errors: bad leading dimension (reference: LogicError from the dimension checks, Gemm.cpp:176-209)
    assert fn(0, 0, 8, 8, 8, sc(1.0), C.c_void_p(A.data_ptr()), 4, C.c_void_p(A.data_ptr()), 8, sc(0.0), C.c_void_p(Cm.data_ptr()), 8, s) != 0
    assert b"leading dimension" in lib.hyd_last_error()


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_transpose_copy_axpy_scale(dtype):
    tdt = torch.float64 if dtype == np.float64 else torch.float32
    sfx = "f64" if dtype == np.float64 else "f32"
    sc = C.c_double if dtype == np.float64 else C.c_float
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    gen = torch.Generator(device=DEV).manual_seed(3)
    for (m, n, pa, pb) in [(1, 1, 0, 0), (64, 64, 0, 0), (65, 63, 1, 2), (1000, 37, 0, 3), (37, 1000, 5, 0), (512, 768, 0, 0),
                           (2050, 130, 2, 2)]:
        lda, ldb_t, ldb = m + pa, n + pb, m + pb
        A = torch.rand((n, lda), dtype=tdt, device=DEV, generator=gen)
        # transpose: B (n x m, ld ldb_t) = A^T, bit-exact, padding untouched
        B = torch.full((m, ldb_t), -7.0, dtype=tdt, device=DEV)
        check(getattr(lib, "hyd_transpose_" + sfx)(m, n, C.c_void_p(A.data_ptr()), lda, C.c_void_p(B.data_ptr()), ldb_t, s))
        torch.cuda.synchronize()
        assert torch.equal(B[:, :n], A[:, :m].t()) and bool((B[:, n:] == -7.0).all())
        # copy
        B2 = torch.full((n, ldb), -7.0, dtype=tdt, device=DEV)
        check(getattr(lib, "hyd_copy2d_" + sfx)(m, n, C.c_void_p(A.data_ptr()), lda, C.c_void_p(B2.data_ptr()), ldb, s))
        torch.cuda.synchronize()
        assert torch.equal(B2[:, :m], A[:, :m]) and bool((B2[:, m:] == -7.0).all())
        # axpy: B += alpha*A   (Axpy.cu:90; FMA contraction allowed -> tolerance of one rounding)
        B3 = torch.rand((n, ldb), dtype=tdt, device=DEV, generator=gen)
        ref = B3[:, :m].double() + 1.7 * A[:, :m].double()
        keep = B3[:, m:].clone()
        check(getattr(lib, "hyd_axpy2d_" + sfx)(m, n, sc(1.7), C.c_void_p(A.data_ptr()), lda, C.c_void_p(B3.data_ptr()), ldb, s))
        torch.cuda.synchronize()
        eps = np.finfo(dtype).eps
        assert (B3[:, :m].double() - ref).abs().max().item() <= 4 * eps * 3
        assert torch.equal(B3[:, m:], keep)
        # scale
        B4 = B3.clone()
        check(getattr(lib, "hyd_scale2d_" + sfx)(m, n, sc(3.0), C.c_void_p(B4.data_ptr()), ldb, s))
        torch.cuda.synchronize()
        assert torch.equal(B4[:, :m], B3[:, :m] * 3.0) and torch.equal(B4[:, m:], keep)
        B4.fill_(float("nan"))
        check(getattr(lib, "hyd_scale2d_" + sfx)(m, n, sc(0.0), C.c_void_p(B4.data_ptr()), ldb, s))
        torch.cuda.synchronize()
        assert bool((B4[:, :m] == 0).all())


def test_launch_counter_and_probe():
    n0 = lib.hyd_launch_count()
    t = C.c_double()
    check(lib.hyd_probe_peak(2, 2000, 2, C.byref(t)))
    assert t.value > 1.0  # TFLOP/s on the FP64 tensor pipe
    Cm = torch.zeros((8, 8), dtype=torch.float64, device=DEV)
    check(lib.hyd_scale2d_f64(8, 8, 2.0, C.c_void_p(Cm.data_ptr()), 8, None))
    assert lib.hyd_launch_count() > n0


def test_interdevice_copies_roundtrip():
    """hyd_memcopy{2d,1d}_async = InterDeviceCopy<..>::MemCopy{2D,1D}Async (include/hydrogen/device/gpu/CUDA.hpp:229-420;
    Copy(Matrix<T,CPU>, Matrix<T,GPU>) and back, Copy.hpp:140-207): block with leading dimensions up, down, device to device;
    padding of the destination untouched; argument errors."""
    import ctypes as C
    from elemental_b200 import lib, check
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for dt, es in ((torch.float64, 8), (torch.float32, 4)):
        m, n, lda, ldb = 37, 11, 40, 45
        hA = (torch.arange(lda * n, dtype=dt) % 1000).pin_memory()
        dB = torch.full((ldb * n,), -5.0, dtype=dt, device="cuda")
        dC = torch.full((lda * n,), -6.0, dtype=dt, device="cuda")
        hD = torch.full((ldb * n,), -7.0, dtype=dt).pin_memory()
        p = lambda t: C.c_void_p(t.data_ptr())
        check(lib.hyd_memcopy2d_async(p(dB), ldb, p(hA), lda, m, n, es, 0, st))      # host -> device
        check(lib.hyd_memcopy2d_async(p(dC), lda, p(dB), ldb, m, n, es, 2, st))      # device -> device
        check(lib.hyd_memcopy2d_async(p(hD), ldb, p(dC), lda, m, n, es, 1, st))      # device -> host
        torch.cuda.current_stream().synchronize()
        A2, D2 = hA.view(n, lda), hD.view(n, ldb)
        assert torch.equal(D2[:, :m], A2[:, :m])
        assert bool((D2[:, m:] == -7.0).all()) and bool((dB.view(n, ldb)[:, m:] == -5.0).all())
        h1 = torch.zeros(100, dtype=dt).pin_memory()
        d1 = torch.arange(100, dtype=dt, device="cuda")
        check(lib.hyd_memcopy1d_async(p(h1), p(d1), 100, es, 1, st))
        torch.cuda.current_stream().synchronize()
        assert torch.equal(h1, d1.cpu())
        assert lib.hyd_memcopy2d_async(p(dB), m - 1, p(hA), lda, m, n, es, 0, st) != 0       # ldim < height
        assert lib.hyd_memcopy2d_async(p(dB), ldb, p(hA), lda, m, n, es, 7, st) != 0         # bad kind
        assert lib.hyd_memcopy2d_async(p(dB), ldb, p(hA), lda, 0, n, es, 0, st) == 0         # empty


@pytest.mark.parametrize("tdt,name", [(torch.bfloat16, "bf16"), (torch.float16, "f16")])
def test_lgemm_half(tdt, name):
    """fp16 / bf16 operands and result, fp32 accumulation: against torch's fp32 product of the same half inputs, rounded once
    to the half type.  Tolerance: one half-precision ulp of the result plus fp32 accumulation-order noise
    (|d| <= 2^-8 |c| + 1e-3*sqrt(k)*eps32-scale for bf16, 2^-11 for fp16)."""
    import ctypes as C
    fn = getattr(lib, "hyd_lgemm_" + name)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    gen = torch.Generator(device=DEV).manual_seed(77)
    ulp = 2.0 ** -8 if tdt == torch.bfloat16 else 2.0 ** -11
    for (m, n, k) in [(128, 128, 64), (256, 384, 192), (1000, 520, 328), (2048, 2048, 1024), (100, 60, 40)]:
        for tA, tB in itertools.product((0, 1), (0, 1)):
            for beta in (0.0, 0.5):
                ar, ac = (k, m) if tA else (m, k)
                br, bc = (n, k) if tB else (k, n)
                lda, ldb, ldc = -(-ar // 8) * 8 + 8, -(-br // 8) * 8, m + 2
                A = ((torch.rand((ac, lda), device=DEV, generator=gen) * 2 - 1)).to(tdt)
                B = ((torch.rand((bc, ldb), device=DEV, generator=gen) * 2 - 1)).to(tdt)
                Cm = ((torch.rand((n, ldc), device=DEV, generator=gen) * 2 - 1)).to(tdt)
                C0 = Cm.clone()
                check(fn(tA, tB, m, n, k, C.c_float(0.75), C.c_void_p(A.data_ptr()), lda, C.c_void_p(B.data_ptr()), ldb,
                         C.c_float(beta), C.c_void_p(Cm.data_ptr()), ldc, st))
                torch.cuda.synchronize()
                a = A[:, :ar].float().t()    # ar x ac (column-major storage: A[j, i] = element (i, j))
                b = B[:, :br].float().t()
                opa = a.t() if tA else a     # m x k
                opb = b.t() if tB else b     # k x n
                ref = 0.75 * (opa @ opb) + beta * C0[:, :m].float().t()
                got = Cm[:, :m].float().t()
                err = (got - ref).abs()
                tol = ulp * ref.abs() + 4e-6 * k ** 0.5 + 1e-6
                assert bool((err <= tol).all()), (name, m, n, k, tA, tB, beta, float(err.max()))
                assert torch.equal(Cm[:, m:], C0[:, m:])     # ldc padding untouched

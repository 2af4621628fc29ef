"""Quick GPU check of hyd_lgemm_f64 against torch (cuBLAS) + timing + pipe-peak probes."""
import ctypes as C
import sys
import torch
sys.path.insert(0, ".")
from elemental_b200 import lib, check

dev = torch.device("cuda:0")


def run(tA, tB, m, n, k, alpha=2.0, beta=3.0, lda_pad=0, time_it=False):
    g = torch.Generator(device=dev).manual_seed(1)
    # column-major storage: a torch (cols, ld) row-major tensor == col-major (ld, cols)
    ar, ac = (k, m) if tA else (m, k)
    br, bc = (n, k) if tB else (k, n)
    lda, ldb, ldc = ar + lda_pad, br + lda_pad, m + lda_pad
    A = torch.rand(ac, lda, dtype=torch.float64, device=dev, generator=g) * 2 - 1
    B = torch.rand(bc, ldb, dtype=torch.float64, device=dev, generator=g) * 2 - 1
    Cm = torch.rand(n, ldc, dtype=torch.float64, device=dev, generator=g) * 2 - 1
    Am = A[:, :ar].t()  # (ar x ac) logical
    Bm = B[:, :br].t()
    opA = Am.t() if tA else Am
    opB = Bm.t() if tB else Bm
    ref = alpha * (opA @ opB) + beta * Cm[:, :m].t()
    out = Cm.clone()
    s = torch.cuda.current_stream().cuda_stream
    check(lib.hyd_lgemm_f64(tA, tB, m, n, k, alpha, A.data_ptr(), lda, B.data_ptr(), ldb, beta,
                            out.data_ptr(), ldc, s))
    torch.cuda.synchronize()
    got = out[:, :m].t()
    err = ((got - ref).norm() / ref.norm()).item()
    pad_ok = True
    if lda_pad:
        pad_ok = torch.equal(out[:, m:], Cm[:, m:])
    msg = f"tA={tA} tB={tB} m={m} n={n} k={k} pad={lda_pad} relF={err:.3e} pad_untouched={pad_ok}"
    if time_it:
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        for _ in range(2):
            lib.hyd_lgemm_f64(tA, tB, m, n, k, alpha, A.data_ptr(), lda, B.data_ptr(), ldb, 1.0, out.data_ptr(), ldc, s)
        e0.record()
        reps = 5
        for _ in range(reps):
            lib.hyd_lgemm_f64(tA, tB, m, n, k, alpha, A.data_ptr(), lda, B.data_ptr(), ldb, 1.0, out.data_ptr(), ldc, s)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        msg += f"  hyd {2*m*n*k/ms/1e9:.2f} TF/s ({ms:.3f} ms)"
        e0.record()
        for _ in range(reps):
            torch.addmm(Cm[:, :m].t(), opA, opB)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        msg += f"  cublas {2*m*n*k/ms/1e9:.2f} TF/s"
    print(msg, flush=True)
    assert err < 1e-13 * max(1, k) ** 0.5 and pad_ok, msg


if __name__ == "__main__":
    for kind, name in [(0, "DFMA"), (1, "DMMA884"), (2, "DMMA16816"), (3, "FFMA")]:
        for cps in (2, 4):
            t = C.c_double()
            check(lib.hyd_probe_peak(kind, 20000, cps, C.byref(t)))
            print(f"probe {name} ctas/sm={cps}: {t.value:.2f} TFLOP/s", flush=True)
    for tA in (0, 1):
        for tB in (0, 1):
            run(tA, tB, 128, 128, 16)
            run(tA, tB, 256, 384, 64)
            run(tA, tB, 300, 200, 100, lda_pad=2)
            run(tA, tB, 1000, 1000, 1000)
            run(tA, tB, 131, 77, 50, lda_pad=1)   # odd ld -> generic path
    for tA in (0, 1):
        for tB in (0, 1):
            run(tA, tB, 8192, 8192, 512, time_it=True)
    run(0, 1, 16384, 16384, 128, time_it=True)
    run(0, 1, 16384, 16384, 1024, time_it=True)
    run(0, 0, 8192, 8192, 8192, time_it=True)
    print("ALL OK")

"""Time hyd_lgemm_f32 (all four transposes) against cuBLAS sgemm (torch, TF32 off) and check the error."""
import ctypes as C
import os
import sys
import torch
sys.path.insert(0, ".")
from elemental_b200 import lib, check

torch.backends.cuda.matmul.allow_tf32 = False
dev = torch.device("cuda:0")
s = torch.cuda.current_stream().cuda_stream
shapes = [(8192, 8192, 8192), (16384, 8192, 1024), (16384, 512, 2048)]
if os.environ.get("SGEMM_SHAPES"):
    shapes = [tuple(int(v) for v in x.split("x")) for x in os.environ["SGEMM_SHAPES"].split(",")]
for (m, n, k) in shapes:
    for (tA, tB) in ((0, 0), (0, 1), (1, 0), (1, 1)):
        ar, ac = (k, m) if tA else (m, k)
        br, bc = (n, k) if tB else (k, n)
        A = torch.rand(ac, ar, dtype=torch.float32, device=dev) * 2 - 1
        B = torch.rand(bc, br, dtype=torch.float32, device=dev) * 2 - 1
        Cm = torch.zeros(n, m, dtype=torch.float32, device=dev)
        f = lambda: check(lib.hyd_lgemm_f32(tA, tB, m, n, k, 1.0, A.data_ptr(), ar, B.data_ptr(), br, 0.0, Cm.data_ptr(), m, s))
        f(); f()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(3):
            f()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        opA = A.t() if not tA else A          # (m x k) logical
        opB = B.t() if not tB else B          # (k x n) logical ... A stored (ac, ar) row-major == col-major (ar x ac)
        opA = A.t() if tA == 0 else A         # col-major (ar x ac): logical matrix = A.t(); op = itself or transposed
        Am = A.t()                            # logical ar x ac
        Bm = B.t()
        opA = Am.t() if tA else Am
        opB = Bm.t() if tB else Bm
        ref = (opA.double() @ opB.double())
        got = Cm.t().double()
        rel = ((got - ref).norm() / ref.norm()).item()
        e0.record()
        for _ in range(3):
            torch.matmul(opA, opB)
        e1.record(); torch.cuda.synchronize()
        ms2 = e0.elapsed_time(e1) / 3
        print(f"m={m} n={n} k={k} tA={tA} tB={tB}: hyd {2*m*n*k/ms/1e9:.1f} TF/s ({ms:.2f} ms) relF={rel:.2e} | cublas sgemm {2*m*n*k/ms2/1e9:.1f} TF/s", flush=True)

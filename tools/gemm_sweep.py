"""Time hyd_lgemm_f64 on the SUMMA-C update shape (m x n x k, 'N','T' and 'N','N') for several k."""
import ctypes as C
import os
import sys
import torch
sys.path.insert(0, ".")
from elemental_b200 import lib, check

dev = torch.device("cuda:0")
m = n = int(os.environ.get("SWEEP_N", "16384"))
s = torch.cuda.current_stream().cuda_stream
Cm = torch.rand(n, m, dtype=torch.float64, device=dev)
for (tA, tB) in ((0, 1), (0, 0)):
    for k in (128, 256, 512, 1024, 2048):
        A = torch.rand(k, m, dtype=torch.float64, device=dev)           # m x k col-major
        B = torch.rand(k, n, dtype=torch.float64, device=dev) if tB else torch.rand(n, k, dtype=torch.float64, device=dev)
        ldb = n if tB else k
        for beta in (1.0, 0.0):
            f = lambda: check(lib.hyd_lgemm_f64(tA, tB, m, n, k, 1.0, A.data_ptr(), m, B.data_ptr(), ldb, beta, Cm.data_ptr(), m, s))
            for _ in range(2):
                f()
            e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
            reps = 4
            e0.record()
            for _ in range(reps):
                f()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            print(f"WN={os.environ.get('HYD_DGEMM_WN','def')} tA={tA} tB={tB} k={k} beta={beta}: {2*m*n*k/ms/1e9:.2f} TF/s ({ms:.3f} ms)", flush=True)

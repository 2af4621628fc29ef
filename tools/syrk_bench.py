"""Syrk / Trrk / Symm timing next to Gemm on the same shapes, one GPU (1x1 grid), device-resident operands.

SURVEY 8(f) N2 measurement for the first GPU call of round 2 (the CUDA path was written without a GPU):
    python tools/syrk_bench.py                    # n = k = 8192 fp64
    SYRK_N=16384 SYRK_K=16384 python tools/syrk_bench.py
Prints one JSON line: ms and TFLOP/s per operation.  Syrk / Trrk update n(n+1)/2 entries: their flops are counted as
n*(n+1)*k (useful work), so "tflops" is directly comparable with Gemm's 2*n*n*k.
"""
import json
import os
import sys

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)


def main():
    import numpy as np
    import torch
    from elemental_b200 import host as H
    n = int(os.environ.get("SYRK_N", "8192"))
    k = int(os.environ.get("SYRK_K", str(n)))
    nb = int(os.environ.get("SYRK_NB", "1024"))
    dtype = np.float64 if os.environ.get("SYRK_DTYPE", "f64") == "f64" else np.float32
    torch.cuda.set_device(0)
    grid = H.Grid(H.World.single(), 1)
    tdt = torch.float64 if dtype == np.float64 else torch.float32
    mk = lambda h, w: H.DistMatrix(grid, h, w, H.MC, H.MR, dtype=dtype)
    A, B, Cm, S = mk(n, k), mk(k, n), mk(n, n), mk(n, n)
    for d in (A, B, S):
        d.store.copy_(torch.rand(d.store.shape, dtype=tdt, device=d.store.device) * 2 - 1)
    ops = {
        "gemm_nn": (2.0 * n * n * k, lambda: H.Gemm(0, 0, 1.0, A, B, 0.0, Cm, H.GEMM_SUMMA_C, blocksize=nb)),
        "gemm_nt": (2.0 * n * n * k, lambda: H.Gemm(0, 1, 1.0, A, A, 0.0, Cm, H.GEMM_SUMMA_C, blocksize=nb)),
        "syrk_ln": (1.0 * n * (n + 1) * k, lambda: H.Syrk(H.LOWER, H.NORMAL, 1.0, A, 0.0, Cm, blocksize=nb)),
        "syrk_ut": (1.0 * n * (n + 1) * k, lambda: H.Syrk(H.UPPER, H.TRANSPOSE, 1.0, B, 0.0, Cm, blocksize=nb)),
        "trrk_l_nn": (1.0 * n * (n + 1) * k, lambda: H.Trrk(H.LOWER, 0, 0, 1.0, A, B, 0.0, Cm, blocksize=nb)),
        "symm_ll": (2.0 * n * n * n, lambda: H.Symm(H.LEFT, H.LOWER, 1.0, S, S, 0.0, Cm, blocksize=nb)),
    }
    out = {"tool": "syrk_bench", "n": n, "k": k, "blocksize": nb, "dtype": str(np.dtype(dtype))}
    for name, (flops, fn) in ops.items():
        fn(); fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(3):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        out[name] = {"ms": round(ms, 3), "tflops": round(flops / (ms * 1e-3) / 1e12, 3)}
    print(json.dumps(out))


if __name__ == "__main__":
    main()

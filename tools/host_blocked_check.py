"""Column blocks of the host-resident Gemm hyd_gemm_host_* (csrc/summa.cu summa_c; cuts from hyd_host_block_plan) against
one block (HYD_HOST_CBLOCKS=1: one download after the last panel): same operands, every orientation, beta = 0 and beta != 0;
fp64 results must be bit-identical (the k-sum of every element is grouped the same way whatever the column cuts), then both
are timed at the bench size.  The knob is read when the grid is created, so each variant runs in its own subprocess:
    python tools/host_blocked_check.py            # driver: runs the variants, compares, prints one JSON line
    HBC_N=16384 HBC_TIME=1 python tools/host_blocked_check.py
Round 2, one B200: 30.1 (1 block) -> 33.2 (4 blocks, first planner) -> 34.4 TFLOP/s (wave-aware planner, default 8 blocks).
(The in-process form of the same check, on every grid: tests/test_gpu_gemm.py::test_gemm_host_column_blocked_path.)
"""
import hashlib
import json
import os
import subprocess
import sys

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)


def worker():
    import torch
    from elemental_b200 import host as H
    n = int(os.environ.get("HBC_N", "4096"))
    nb = int(os.environ.get("HBC_NB", "512"))
    timed = os.environ.get("HBC_TIME") == "1"
    torch.cuda.set_device(0)
    grid = H.Grid(H.World.single(), 1)
    out = {}
    gen = torch.Generator(device="cpu").manual_seed(11)
    mk = lambda h, w: H.DistMatrix(grid, h, w, H.MC, H.MR, device="cpu", pinned=True)
    for (oA, oB, beta) in ((0, 0, 0.0), (0, 1, 3.0), (1, 0, 0.0), (1, 1, -0.5)):
        k = n // 2 + 64
        a = mk(k, n) if oA else mk(n, k)
        b = mk(n + 128, k) if oB else mk(k, n + 128)   # C is n x (n+128): ragged last column block
        c = mk(n, n + 128)
        for d in (a, b, c):
            d.store.copy_(torch.rand(d.store.shape, dtype=torch.float64, generator=gen) * 2 - 1)
        H.Gemm(oA, oB, 2.0, a, b, beta, c, H.GEMM_SUMMA_C, blocksize=nb)
        out[f"{oA}{oB}b{beta}"] = hashlib.sha256(c.store[: c.lw, : c.lh].contiguous().numpy().tobytes()).hexdigest()
    if timed:
        a, b, c = mk(n, n), mk(n, n), mk(n, n)
        for d in (a, b):
            d.store.copy_(torch.rand(d.store.shape, dtype=torch.float64, generator=gen) * 2 - 1)
        step = lambda: H.Gemm(0, 0, 1.0, a, b, 0.0, c, H.GEMM_SUMMA_C, blocksize=nb)
        step(); step()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(3):
            step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        out["ms_per_gemm"] = ms
        out["tflops"] = 2.0 * n ** 3 / (ms * 1e-3) / 1e12
    print("HBC " + json.dumps(out), flush=True)


def main():
    if os.environ.get("HBC_WORKER") == "1":
        return worker()
    res = {}
    for cb in ("1", "4", "3"):
        env = dict(os.environ, HBC_WORKER="1", HYD_HOST_CBLOCKS=cb)
        p = subprocess.run([sys.executable, os.path.abspath(__file__)], env=env, capture_output=True, text=True, timeout=600)
        line = [l for l in p.stdout.splitlines() if l.startswith("HBC ")]
        if p.returncode != 0 or not line:
            print(json.dumps({"tool": "host_blocked_check", "error": f"cblocks={cb} rc={p.returncode}", "stderr": p.stderr[-1500:]}))
            sys.exit(1)
        res[cb] = json.loads(line[0][4:])
    keys = [k for k in res["1"] if k not in ("ms_per_gemm", "tflops")]
    same = {cb: all(res[cb][k] == res["1"][k] for k in keys) for cb in ("4", "3")}
    print(json.dumps({"tool": "host_blocked_check", "bit_identical_to_unblocked": same,
                      "timing": {cb: {k: res[cb].get(k) for k in ("ms_per_gemm", "tflops")} for cb in res}}))
    sys.exit(0 if all(same.values()) else 1)


if __name__ == "__main__":
    main()

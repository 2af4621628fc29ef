"""BASELINE.json configs[2] and [4] (the fp32 workloads), one process per GPU:
  C2: SUMMA Gemm TN / NT float m=n=k=32768 on a 1x1 / 1x2 / 2x2 / 2x4 grid, blocksize sweep
  C4: tall-skinny float Gemm m=16384 n=512 k=16384 with [*,VC] / [VC,*] operands (SUMMA_Dot path)
TFLOP/s = 2*m*n*k / device time (CUDA events, max over ranks); operands resident in HBM; data synthetic.
A size-independent parity property is checked once per case: the reference's associativity residual
(tests/blas_like/Gemm.cpp:27-61) within the fp32 tolerance 1e-5*sqrt(k) relative Frobenius.  bench.py's `extra` block
carries the same measurements with clocks; this tool adds the blocksize sweep.

torchrun --nproc-per-node P tools/config_bench.py      (env: CFG_N for C2's size, CFG_NBS="512,1024,2048")
"""
import json
import os
import sys

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from elemental_b200 import host as H  # noqa: E402

GRIDS = {1: 1, 2: 1, 4: 2, 8: 2}


def main():
    rank, size, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(lr)
    if size > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
        world = H.World.from_torch_distributed()
    else:
        world = H.World.single()
    grid = H.Grid(world, GRIDS[size])
    if os.environ.get("CFG_OVERLAP"):
        grid.SetOverlap(int(os.environ["CFG_OVERLAP"]), int(os.environ.get("CFG_SM_RESERVE", "4")))
    st = torch.cuda.Stream()
    out = []
    f32 = np.float32
    import ctypes as C
    from elemental_b200 import lib, check
    _p = C.c_double()
    check(lib.hyd_probe_peak(4, 4000, 1, C.byref(_p)))
    tf32_peak = _p.value

    def timeit(f, reps=3, warm=2):
        for _ in range(warm):
            f()
        if size > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(st)
        for _ in range(reps):
            f()
        e1.record(st)
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / reps], device="cuda")
        if size > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def report(row):
        if rank == 0:
            print(json.dumps(row), flush=True)
            out.append(row)

    def associativity(oa, ob, A, B, Cm, k):
        """The reference's own Gemm test (tests/blas_like/Gemm.cpp:27-61, TestAssociativity), as in bench.py and
        tests/test_gpu_gemm.py: || (op(A) op(B)) X - op(A) (op(B) X) ||_F / || op(A) (op(B) X) ||_F for 16 random right-hand sides,
        Cm = op(A) op(B) as left by the timed run (alpha 1, beta 0).  (Round 1 checked C(2 alpha) = 2 C(alpha) here, which
        holds exactly for ANY kernel because doubling is exact in binary floating point.)"""
        mk = lambda h: H.DistMatrix(grid, h, 16, H.MC, H.MR, dtype=f32)
        X, Z = mk(Cm.n), mk(k)
        Y, W = mk(Cm.m), mk(Cm.m)
        X.store.uniform_(-1, 1)
        H.Gemm(ob, H.NORMAL, 1.0, B, X, 0.0, Z, stream=st)
        H.Gemm(oa, H.NORMAL, 1.0, A, Z, 0.0, Y, stream=st)
        H.Gemm(H.NORMAL, H.NORMAL, 1.0, Cm, X, 0.0, W, stream=st)
        st.synchronize()
        d = W.store[: W.lw, : W.lh].double() - Y.store[: Y.lw, : Y.lh].double()
        t = torch.tensor([float((d ** 2).sum()), float((Y.store[: Y.lw, : Y.lh].double() ** 2).sum())], dtype=torch.float64, device="cuda")
        if size > 1:
            dist.all_reduce(t)
        rel = float((t[0] / t[1].clamp_min(1e-300)) ** 0.5)
        assert rel <= 1e-5 * k ** 0.5, rel
        return rel

    with torch.cuda.stream(st):
        # ---- C2
        n = int(os.environ.get("CFG_N", "32768"))
        nbs = [int(x) for x in os.environ.get("CFG_NBS", "512,1024,2048").split(",")]
        A = H.DistMatrix(grid, n, n, H.MC, H.MR, dtype=f32)
        B = H.DistMatrix(grid, n, n, H.MC, H.MR, dtype=f32)
        Cm = H.DistMatrix(grid, n, n, H.MC, H.MR, dtype=f32)
        for d in (A, B):
            d.store.uniform_(-1, 1)
        for (oa, ob, name) in ((H.TRANSPOSE, H.NORMAL, "TN"), (H.NORMAL, H.TRANSPOSE, "NT")):
            for nb in nbs:
                run = lambda alpha=1.0: H.Gemm(oa, ob, alpha, A, B, 0.0, Cm, H.GEMM_SUMMA_C, stream=st, blocksize=nb)
                ms = timeit(run)
                rel = None
                if nb == nbs[0]:
                    run(1.0)
                    rel = associativity(oa, ob, A, B, Cm, n)
                tf = 2.0 * n * n * n / ms / 1e9
                report({"config": "C2", "op": f"SUMMA-C Gemm {name} fp32 m=n=k={n}", "gpus": size, "blocksize": nb, "ms": round(ms, 3),
                        "tflops": round(tf, 2), "tflops_per_gpu": round(tf / size, 2),
                        "frac_of_tf32_bound_per_gpu": round(tf / size / (tf32_peak / 3), 3),
                        "tf32_bound_note": f"dense TF32 tcgen05 rate measured live ({tf32_peak:.0f} TFLOP/s, hyd_probe_peak kind 4) / 3 MMAs per fp32 product",
                        "associativity_rel": rel})
        del A, B, Cm
        torch.cuda.empty_cache()
        # ---- C4
        m, nn, k = 16384, 512, 16384
        A = H.DistMatrix(grid, m, k, H.STAR, H.VC, dtype=f32)
        B = H.DistMatrix(grid, k, nn, H.VC, H.STAR, dtype=f32)
        Cm = H.DistMatrix(grid, m, nn, H.MC, H.MR, dtype=f32)
        for d in (A, B):
            d.store.uniform_(-1, 1)
        run = lambda alpha=1.0: H.Gemm(H.NORMAL, H.NORMAL, alpha, A, B, 0.0, Cm, H.GEMM_SUMMA_DOT, stream=st)
        ms = timeit(run, reps=10)
        run(1.0)
        rel = associativity(H.NORMAL, H.NORMAL, A, B, Cm, k)
        tf = 2.0 * m * nn * k / ms / 1e9
        report({"config": "C4", "op": f"SUMMA-Dot Gemm NN fp32 m={m} n={nn} k={k}, A[*,VC] B[VC,*] C[MC,MR]", "gpus": size, "ms": round(ms, 3),
                "tflops": round(tf, 2), "tflops_per_gpu": round(tf / size, 2), "associativity_rel": rel})
    if rank == 0:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        json.dump(out, open(os.path.join(ROOT, "gpurun_out", f"config_bench_{size}gpu.json"), "w"), indent=1)
    grid.destroy()
    if size > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

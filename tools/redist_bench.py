"""Redistribution / local-copy bandwidth (BASELINE.json configs[3]): 32768 x 32768 fp64.
1 GPU: local transpose, 2-D copy, axpy, and the engine's [MC,MR]->[*,VR] / Transpose on a 1x1 grid (pure
pack kernels).  Under torchrun: the same redistributions across GPUs + SUMMA panel gathers.
GB/s = algorithmic HBM bytes per GPU (2 * local bytes) / device time (CUDA events, max over ranks)."""
import ctypes as C
import json
import os
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from elemental_b200 import host as H, lib, check  # noqa: E402

GRIDS = {1: 1, 2: 1, 4: 2, 8: 2}
HBM = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0


def timeit(f, st, size, reps=5, warm=2):
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


def main():
    rank, size, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(lr)
    if size > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
        world = H.World.from_torch_distributed()
    else:
        world = H.World.single()
    n = int(os.environ.get("REDIST_N", "32768"))
    grid = H.Grid(world, GRIDS[size])
    st = torch.cuda.Stream()
    out = []

    def breakdown(f):
        """one profiled call: (whole redistribution, pack+push kernels, unpack kernels) in ms on this rank"""
        grid.Profile(True)
        f()
        st.synchronize()
        r = [grid.ProfileRead(k)[1] for k in (1, 3, 4)]
        grid.Profile(False)
        return [round(x, 3) for x in r]

    def report(name, ms, hbm_bytes, nvl_bytes=0, f=None):
        bd = breakdown(f) if f is not None else None
        if rank == 0:
            gbs = hbm_bytes / ms / 1e6
            row = {"op": name, "n": n, "gpus": size, "ms": round(ms, 3), "hbm_GBps_per_gpu": round(gbs, 1), "frac_of_measured_hbm": round(gbs / HBM, 3)}
            if nvl_bytes:
                row["nvlink_GBps_out_per_gpu"] = round(nvl_bytes / ms / 1e6, 1)
                row["frac_of_770_nvlink"] = round(nvl_bytes / ms / 1e6 / 770.0, 3)
                bound = max(hbm_bytes / HBM, nvl_bytes / 770.0) / 1e6
                row["frac_of_roofline"] = round(bound / ms, 3)
            if bd is not None:
                row["rank0_ms_total_pack_unpack"] = bd
            print(json.dumps(row), flush=True)
            out.append(row)

    with torch.cuda.stream(st):
        s = C.c_void_p(st.cuda_stream)
        if size == 1:
            A = torch.rand((n, n), dtype=torch.float64, device="cuda")
            B = torch.empty((n, n), dtype=torch.float64, device="cuda")
            by = 2 * n * n * 8
            report("local transpose (hyd_transpose_f64)", timeit(lambda: check(lib.hyd_transpose_f64(n, n, A.data_ptr(), n, B.data_ptr(), n, s)), st, 1), by)
            report("local copy2d (hyd_copy2d_f64)", timeit(lambda: check(lib.hyd_copy2d_f64(n, n, A.data_ptr(), n, B.data_ptr(), n, s)), st, 1), by)
            report("local axpy2d (hyd_axpy2d_f64)", timeit(lambda: check(lib.hyd_axpy2d_f64(n, n, 2.0, A.data_ptr(), n, B.data_ptr(), n, s)), st, 1), by * 3 // 2)
            report("local scale2d (hyd_scale2d_f64)", timeit(lambda: check(lib.hyd_scale2d_f64(n, n, 1.0000001, B.data_ptr(), n, s)), st, 1), by)
            report("torch copy_ (reference point)", timeit(lambda: B.copy_(A), st, 1), by)
            del A, B
        r, c = grid.Height(), grid.Width()
        P = size
        loc = n * n * 8 / P
        X = H.DistMatrix(grid, n, n, H.MC, H.MR)
        X.store.uniform_(-1, 1)
        cases = [("[MC,MR] -> [*,VR]", (H.STAR, H.VR), 0, loc * (1 - 1.0 / r)),
                 ("[MC,MR] -> [VC,*]", (H.VC, H.STAR), 0, loc * (1 - 1.0 / c)),
                 ("Transpose [MC,MR] -> [MR,MC] (local transpose only)", (H.MR, H.MC), 1, 0),
                 ("Copy [MC,MR] -> [MR,MC] (TransposeDist)", (H.MR, H.MC), 0, loc * (1 - (1.0 / P if r != c else 1.0 / P))),
                 ("Transpose [MC,MR] -> [MC,MR]", (H.MC, H.MR), 1, loc * (1 - 1.0 / P))]
        if size > 1:
            # pure shift: every local block moves whole to the next grid column (contiguous 16-byte path on both sides):
            # the raw peer-store rate of the push transport
            Y = H.DistMatrix(grid, n, n, H.MC, H.MR, rowAlign=1)
            report("[MC,MR] -> [MC,MR] rowAlign+1 (whole-block shift)", timeit(lambda: H.Copy(X, Y, stream=st), st, size), 2 * loc, loc, lambda: H.Copy(X, Y, stream=st))
            del Y
        for name, (cd, rd), tr, nvl in cases:
            Y = H.DistMatrix(grid, n, n, cd, rd)
            f = (lambda: H.Transpose(X, Y, stream=st)) if tr else (lambda: H.Copy(X, Y, stream=st))
            report(name, timeit(f, st, size), 2 * loc, nvl if size > 1 else 0, f)
            # size-independent parity property: the way back restores X bit for bit (chunk-pipelined path included)
            Z = H.DistMatrix(grid, n, n, H.MC, H.MR)
            Z.store.zero_()
            (H.Transpose if tr else H.Copy)(Y, Z, stream=st)
            st.synchronize()
            ok = bool(torch.equal(Z.store[: Z.lw, : Z.lh], X.store[: X.lw, : X.lh]))
            if not ok:
                print(f"[rank {rank}] ROUND TRIP MISMATCH for {name}", flush=True)
            assert ok, name
            del Y, Z
        # SUMMA panel gathers at panel width nb (algorithmic HBM bytes: read local slice + write gathered panel)
        for nb in (128, 512, 2048):
            Ap = H.DistMatrix(grid, n, nb, H.MC, H.STAR)
            Bp = H.DistMatrix(grid, n, nb, H.MR, H.STAR)
            A1 = H.DistMatrix(grid, n, nb, H.MC, H.MR)
            B1 = H.DistMatrix(grid, nb, n, H.MC, H.MR)
            A1.store.uniform_(-1, 1); B1.store.uniform_(-1, 1)
            ga = (n / r) * nb * 8
            gb = (n / c) * nb * 8
            report(f"panel [MC,MR] -> [MC,*] nb={nb}", timeit(lambda: H.Copy(A1, Ap, stream=st), st, size, reps=20), ga / c + ga, ga * (1 - 1.0 / c) if size > 1 else 0)
            report(f"panel ^T [MC,MR] -> [MR,*] nb={nb}", timeit(lambda: H.Transpose(B1, Bp, stream=st), st, size, reps=20), gb / r + gb, gb * (1 - 1.0 / r) if size > 1 else 0)
    if rank == 0:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        json.dump(out, open(os.path.join(ROOT, "gpurun_out", f"redist_bench_{size}gpu.json"), "w"), indent=1)
    grid.destroy()
    if size > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

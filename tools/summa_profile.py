"""Phase breakdown of the SUMMA engine under torchrun: Gemm time with overlap on/off, and the
event-timed share of local GEMM launches vs redistributions (accurate with overlap off: one stream)."""
import os
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
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
    n = int(os.environ.get("PROF_N", "16384"))
    alg = int(os.environ.get("PROF_ALG", "3"))
    oA, oB = int(os.environ.get("PROF_OA", "0")), int(os.environ.get("PROF_OB", "0"))
    import numpy as np
    dtype = np.float32 if os.environ.get("PROF_DTYPE", "f64") == "f32" else np.float64
    grid = H.Grid(world, GRIDS[size])
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        A, B, Cm = (H.DistMatrix(grid, n, n, dtype=dtype) for _ in range(3))
        for d in (A, B, Cm):
            d.store.uniform_(-1, 1)
        for nb in [int(x) for x in os.environ.get("PROF_NB", "256,512,1024,2048").split(",")]:
            modes = [tuple(int(v) for v in x.split(':')) for x in os.environ.get('PROF_MODES', '1:8,1:4,1:16,0:0').split(',')]
            for overlap, res in [(bool(a), b) for a, b in modes]:
                grid.SetOverlap(overlap, res)
                f = lambda: H.Gemm(oA, oB, 1.0, A, B, 0.0, Cm, alg, blocksize=nb)
                f(); f()
                if os.environ.get("PROF_CHECK"):
                    torch.cuda.synchronize()
                    got = Cm.store.clone()
                    grid.SetOverlap(0, 0)
                    f()
                    torch.cuda.synchronize()
                    rel = float((got - Cm.store).norm() / Cm.store.norm())
                    bad = int(((got - Cm.store).abs() > 1e-6 * Cm.store.abs().max()).sum())
                    print(f"[rank {rank}] nb={nb} overlap={overlap}: rel diff vs non-overlapped direct path {rel:.3e}, bad elements {bad}", flush=True)
                    grid.SetOverlap(overlap, res)
                if size > 1:
                    dist.barrier()
                torch.cuda.synchronize()
                grid.Profile(True)
                e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
                e0.record(st)
                reps = 3
                for _ in range(reps):
                    f()
                e1.record(st)
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / reps
                ng, gms = grid.ProfileRead(0)
                nr, rms = grid.ProfileRead(1)
                nd, dms = grid.ProfileRead(2)
                grid.Profile(False)
                t = torch.tensor([ms], device="cuda")
                if size > 1:
                    dist.all_reduce(t, op=dist.ReduceOp.MAX)
                if rank == 0:
                    tf = 2.0 * n ** 3 / (t.item() * 1e-3) / 1e12
                    print(f"P={size} n={n} alg={alg} o={oA}{oB} nb={nb} overlap={overlap} reserve={res}: {t.item():.2f} ms/step {tf:.1f} TF/s | "
                          f"gemm {gms/reps:.2f} ms ({ng//reps} launches) redist {rms/reps:.2f} ms ({nr//reps}) reduce {dms/reps:.2f} ms ({nd//reps})", flush=True)
    grid.destroy()
    if size > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

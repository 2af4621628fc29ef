"""Multi-GPU parity check, one process per GPU over NCCL (run under torch.distributed.run).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/dist_check.py

Every rank builds the oracle's simulated grid (all ranks' blocks) on its host, uploads ITS block, runs the
product path (hyd_redist_* / hyd_axpy_contract_* / hyd_gemm_* over NCCL send/recv), and compares its block
with the oracle's: bit-exact for redistributions, tolerance for Gemm."""
import itertools
import os
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as O  # noqa: E402
from elemental_b200 import host as H  # noqa: E402
import gpu_util as U  # noqa: E402


def main():
    rank, size, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    world = H.World.from_torch_distributed()
    heights = [h for h in (1, 2, 4) if size % h == 0 and h <= size]
    fails = 0
    for r in heights:
        c = size // r
        grid = H.Grid(world, r)
        og = O.Grid(r, c, 1)
        rng = np.random.default_rng(5 + r)  # same stream on every rank
        # --- redistributions: all 121 pairs, random alignments, copy and transpose
        n_re = 0
        for tr in (0, 1):
            for (scd, srd), (dcd, drd) in itertools.product(O.PAIRS, O.PAIRS):
                m, n = [(150, 131), (7, 5), (1029, 67)][n_re % 3]
                A = O.DM(og, m, n, scd, srd, U.rand_align(rng, og, scd), U.rand_align(rng, og, srd)).fill_index()
                dm, dn = (n, m) if tr else (m, n)
                B = O.DM(og, dm, dn, dcd, drd, U.rand_align(rng, og, dcd), U.rand_align(rng, og, drd)).fill_const(-77)
                O.lib.orc_redist_closed(A.h, B.h, tr)
                a = U.to_rank(grid, A)
                b = U.to_rank(grid, O.DM(og, dm, dn, dcd, drd, B.ca, B.ra).fill_const(-77), ldim_extra=2)
                (H.Transpose if tr else H.Copy)(a, b)
                torch.cuda.synchronize()
                if not np.array_equal(U.block_of(b), B.block(rank)) or not U.padding_untouched(b):
                    fails += 1
                    print(f"[rank {rank}] REDIST MISMATCH grid {r}x{c} tr={tr} {(scd, srd)}->{(dcd, drd)} {m}x{n}", flush=True)
                n_re += 1
        # --- reduce side
        for (scd, srd, tr) in [(O.MC, O.STAR, 0), (O.STAR, O.MR, 0), (O.STAR, O.STAR, 0), (O.MR, O.STAR, 1), (O.STAR, O.MC, 1)]:
            m, n = 167, 145
            sm, sn = (n, m) if tr else (m, n)
            A = O.DM(og, sm, sn, scd, srd, U.rand_align(rng, og, scd), U.rand_align(rng, og, srd))
            for p in range(A.P):
                A.block(p)[...] = rng.standard_normal(A.block(p).shape)
            ca, ra = U.rand_align(rng, og, O.MC), U.rand_align(rng, og, O.MR)
            B0 = O.DM(og, m, n, O.MC, O.MR, ca, ra).fill(9)
            B1 = O.DM(og, m, n, O.MC, O.MR, ca, ra).fill(9)
            O.lib.orc_axpy_contract_closed(1.5, A.h, B1.h, tr)
            a, b = U.to_rank(grid, A), U.to_rank(grid, B0)
            (H.TransposeAxpyContract if tr else H.AxpyContract)(1.5, a, b)
            torch.cuda.synchronize()
            if not np.array_equal(U.block_of(b), B1.block(rank)):
                fails += 1
                print(f"[rank {rank}] CONTRACT MISMATCH grid {r}x{c} {(scd, srd, tr)}", flush=True)
        # --- Gemm: every orientation x algorithm, fp64 and fp32, overlap on and off
        for dtype, tol in ((np.float64, 1e-10), (np.float32, 1e-3)):
            for overlap in (True, False):
                grid.SetOverlap(overlap)
                for oA, oB in itertools.product((0, 1), (0, 1)):
                    for alg in (H.GEMM_SUMMA_A, H.GEMM_SUMMA_B, H.GEMM_SUMMA_C, H.GEMM_SUMMA_DOT):
                        m, n, k, nb = 345, 238, 452, 64
                        am, an = (k, m) if oA else (m, k)
                        bm, bn = (n, k) if oB else (k, n)
                        mk = lambda h, w, seed: O.DM(og, h, w, O.MC, O.MR, U.rand_align(rng, og, O.MC), U.rand_align(rng, og, O.MR), dtype).fill(seed)
                        A, B, Cx = mk(am, an, 1), mk(bm, bn, 2), mk(m, n, 3)
                        Cin = O.DM(og, m, n, O.MC, O.MR, Cx.ca, Cx.ra, dtype).fill(3)
                        O.gemm(oA, oB, alg, 2.0, A, B, 3.0, Cx, nb)
                        a, b, cm = U.to_rank(grid, A), U.to_rank(grid, B), U.to_rank(grid, Cin)
                        H.Gemm(oA, oB, 2.0, a, b, 3.0, cm, alg, blocksize=nb)
                        torch.cuda.synchronize()
                        got = U.block_of(cm)
                        err = float(np.abs(got.astype(np.float64) - Cx.block(rank).astype(np.float64)).max()) if got.size else 0.0
                        if err > tol:
                            fails += 1
                            print(f"[rank {rank}] GEMM MISMATCH grid {r}x{c} {dtype.__name__} oA={oA} oB={oB} alg={alg} overlap={overlap} err={err:.3e}", flush=True)
        grid.destroy()
        if rank == 0:
            print(f"grid {r}x{c}: {n_re} redistributions + contracts + 64 gemms checked", flush=True)
    t = torch.tensor([fails], device="cuda")
    dist.all_reduce(t)
    if rank == 0:
        print("DIST_CHECK", "OK" if int(t.item()) == 0 else f"FAILED ({int(t.item())} mismatches)", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if int(t.item()) == 0 else 1)


if __name__ == "__main__":
    main()

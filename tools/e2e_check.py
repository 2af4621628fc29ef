"""Host-operand Gemm (hyd_gemm_host_f64) vs device-resident Gemm, repeated: locates mismatching elements.

torchrun --nproc-per-node P tools/e2e_check.py   (env: E2E_N, E2E_NB, E2E_ITERS)
"""
import os
import sys

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch  # noqa: E402
from elemental_b200 import host as H  # noqa: E402

GRIDS = {1: (1, 1), 2: (1, 2), 4: (2, 2), 8: (2, 4)}


def forensic(A, B, dd, bad):
    """1x1 grid: explain a bad block as (a_wrong - a)[i] * B[kk, j] and look for where a_wrong came from."""
    c0, r0 = int(bad[0, 0]), int(bad[0, 1])
    bad = bad[(bad[:, 0] < c0 + 32) & ((bad[:, 1] - r0).abs() < 8)]   # first block only
    cols = bad[:, 0].unique()
    rows = bad[:, 1].unique()
    D = dd[cols][:, rows].t().contiguous()          # rows x cols
    sv = torch.linalg.svdvals(D)
    print(f"   block {rows.numel()}x{cols.numel()} rows {rows.tolist()} cols {int(cols[0])}.. singular values {sv[:4].tolist()}")
    Bs = B.store[cols]                              # [j, kk] = B[kk, j]
    d0 = D[D.abs().sum(1).argmax()]
    corr = (Bs * d0[:, None]).sum(0) / (Bs.norm(dim=0) * d0.norm() + 1e-300)
    top = corr.abs().topk(4)
    print(f"   best k candidates {top.indices.tolist()} |corr| {top.values.tolist()}")
    kk = int(top.indices[0])
    b = Bs[:, kk]
    alpha = (D @ b) / (b @ b)
    resid = float((D - alpha[:, None] * b[None, :]).norm() / D.norm())
    a_right = A.store[kk, rows]
    a_wrong = a_right + alpha
    print(f"   k={kk} residual {resid:.2e} alpha {alpha.tolist()}")
    print(f"   a_right {a_right.tolist()}\n   a_wrong {a_wrong.tolist()}")
    for i in range(min(rows.numel(), 8)):
        hit = ((A.store - a_wrong[i]).abs() < 1e-11).nonzero()
        hb = ((B.store - a_wrong[i]).abs() < 1e-11).nonzero()
        print(f"   a_wrong[{i}] (row {int(rows[i])}) found in A at [k,row] {hit[:4].tolist()}  in B at [col,k] {hb[:4].tolist()}")


def main():
    ws = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    lr = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(lr)
    if ws > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
        world = H.World.from_torch_distributed()
    else:
        world = H.World.single()
    r, _ = GRIDS[ws]
    grid = H.Grid(world, r)
    n = int(os.environ.get("E2E_N", "16384"))
    nb = int(os.environ.get("E2E_NB", "1024"))
    iters = int(os.environ.get("E2E_ITERS", "4"))
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        A = H.DistMatrix(grid, n, n, H.MC, H.MR)
        B = H.DistMatrix(grid, n, n, H.MC, H.MR)
        Cm = H.DistMatrix(grid, n, n, H.MC, H.MR)
        C2 = H.DistMatrix(grid, n, n, H.MC, H.MR)
        gen = torch.Generator(device="cuda").manual_seed(1234 + rank)
        for d in (A, B):
            d.store.uniform_(-1.0, 1.0, generator=gen)
        H.Gemm(H.NORMAL, H.NORMAL, 1.0, A, B, 0.0, Cm, H.GEMM_SUMMA_C, blocksize=nb)
        hA = H.DistMatrix(grid, n, n, H.MC, H.MR, device="cpu", pinned=True)
        hB = H.DistMatrix(grid, n, n, H.MC, H.MR, device="cpu", pinned=True)
        hC = H.DistMatrix(grid, n, n, H.MC, H.MR, device="cpu", pinned=True)
        hA.store.copy_(A.store)
        hB.store.copy_(B.store)
        torch.cuda.synchronize()
        ref = Cm.store[: Cm.lw, : Cm.lh]
        nrm = float(ref.norm())
        ref0, A0, B0 = ref.clone(), A.store.clone(), B.store.clone()
        for it in range(iters):
            H.Gemm(H.NORMAL, H.NORMAL, 1.0, A, B, 0.0, C2, H.GEMM_SUMMA_C, blocksize=nb)
            torch.cuda.synchronize()
            d2 = C2.store[: Cm.lw, : Cm.lh] - ref
            hC.store.zero_()
            H.Gemm(H.NORMAL, H.NORMAL, 1.0, hA, hB, 0.0, hC, H.GEMM_SUMMA_C, blocksize=nb)
            torch.cuda.synchronize()
            dd = hC.store[: hC.lw, : hC.lh].cuda() - ref
            bad = (dd.abs() > 1e-9).nonzero()
            msg = f"[rank {rank}] it={it} device-rerun rel={float(d2.norm()) / nrm:.2e}  host rel={float(dd.norm()) / nrm:.2e} bad={bad.shape[0]}"
            if bad.shape[0]:
                cols = bad[:, 0].unique()
                rows = bad[:, 1]
                msg += f" cols={cols[:6].tolist()}(n={cols.numel()}) rows=[{int(rows.min())},{int(rows.max())}] std={float(dd[dd.abs() > 1e-9].std()):.3f}"
            print(msg, flush=True)
            for nm, x, x0 in (("Cm", ref, ref0), ("A", A.store, A0), ("B", B.store, B0)):
                ch = (x != x0).nonzero()
                if ch.shape[0]:
                    print(f"   !! resident matrix {nm} changed: {ch.shape[0]} elements, first {ch[:3].tolist()} last {ch[-1].tolist()}", flush=True)
            hch = int((hA.store != A0.cpu()).sum()) if it == iters - 1 else 0
            if hch:
                print(f"   !! pinned hA changed: {hch}")
            if bad.shape[0] and ws == 1:
                forensic(A, B, dd, bad)
            del dd, d2
    if ws > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

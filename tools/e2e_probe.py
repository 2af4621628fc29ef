"""Where the end-to-end (host-resident) Gemm's time goes on N GPUs, one process per GPU (torchrun):
 (1) raw pinned-host <-> device bandwidth of every GPU at once: H2D alone, D2H alone, both directions together, contiguous and
     as the 4 KiB-row 2-D copies a row-sliced operand needs;
 (2) hyd_gemm_host_f64 at the bench size under the engine's knobs (set by the caller through the environment: HYD_PANEL_CE,
     HYD_HOST_CBLOCKS), next to the device-resident time.
Prints one JSON line on rank 0."""
import json
import os
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (NUMA placement helper)
from elemental_b200 import host as H  # noqa: E402

GRIDS = {1: 1, 2: 1, 4: 2, 8: 2}


def main():
    rank, size, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(lr)
    numa = bench.bind_to_gpu_numa_node(lr) if size > 1 else None
    if size > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
        world = H.World.from_torch_distributed()
    else:
        world = H.World.single()
    out = {"gpus": size, "numa": numa, "env": {k: os.environ[k] for k in ("HYD_PANEL_CE", "HYD_HOST_CBLOCKS", "HYD_BENCH_NUMA", "HYD_HOST_B0", "HYD_HOST_FIRST_TILES", "HYD_HOST_GATHER", "E2E_SM_RESERVE") if k in os.environ}}

    def sync():
        if size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def worst(x):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        if size > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    if os.environ.get("E2E_RAW", "1") == "1":
        nbytes = 256 << 20
        hbuf = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
        hbuf2 = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
        dbuf = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
        dbuf2 = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
        s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

        def run(up, down, reps=4):
            sync()
            e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
            e0.record()
            s1.wait_stream(torch.cuda.current_stream())
            s2.wait_stream(torch.cuda.current_stream())
            for _ in range(reps):
                if up:
                    with torch.cuda.stream(s1):
                        dbuf.copy_(hbuf, non_blocking=True)
                if down:
                    with torch.cuda.stream(s2):
                        hbuf2.copy_(dbuf2, non_blocking=True)
            torch.cuda.current_stream().wait_stream(s1)
            torch.cuda.current_stream().wait_stream(s2)
            e1.record()
            torch.cuda.synchronize()
            ms = worst(e0.elapsed_time(e1)) / reps
            return round(nbytes / ms / 1e6, 1)

        run(True, True)
        out["raw_GBps_per_gpu_all_gpus_at_once"] = {"h2d_alone": run(True, False), "d2h_alone": run(False, True), "both_each_direction": run(True, True)}
        # 2-D, as the engine issues it (cudaMemcpy2DAsync through hyd_memcopy2d_async): the k-panels of a row-sliced operand --
        # 16 panels of 512 rows (4 KiB) x 4096 columns out of an 8192-row (64 KiB pitch) column-major block -- and 2 KiB / 16 KiB rows
        import ctypes as C
        from elemental_b200 import lib, check
        for rows in (256, 512, 2048):
            sync()
            e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
            e0.record()
            npan = 8192 // rows
            for pnl in range(npan):
                check(lib.hyd_memcopy2d_async(C.c_void_p(dbuf.data_ptr() + pnl * rows * 8), 8192, C.c_void_p(hbuf.data_ptr() + pnl * rows * 8), 8192,
                                              rows, 4096, 8, 0, C.c_void_p(torch.cuda.current_stream().cuda_stream)))
            e1.record()
            torch.cuda.synchronize()
            out["raw_GBps_per_gpu_all_gpus_at_once"][f"h2d_2d_{rows * 8 // 1024}KiB_rows"] = round(nbytes / worst(e0.elapsed_time(e1)) / 1e6, 1)
        del hbuf, hbuf2, dbuf, dbuf2

    n, nb = int(os.environ.get("E2E_N", "16384")), 1024
    grid = H.Grid(world, GRIDS[size])
    if os.environ.get("E2E_SM_RESERVE"):
        grid.SetOverlap(-1, int(os.environ["E2E_SM_RESERVE"]))
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        mk = lambda dev: H.DistMatrix(grid, n, n, H.MC, H.MR, **({"device": "cpu", "pinned": True} if dev == "cpu" else {}))
        A, B, Cm = mk("cuda"), mk("cuda"), mk("cuda")
        for d in (A, B):
            d.store.uniform_(-1, 1)
        hA, hB, hC = mk("cpu"), mk("cpu"), mk("cpu")
        hA.store.copy_(A.store)
        hB.store.copy_(B.store)

        def timeit(f, reps=4):
            f(); f()
            sync()
            e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
            e0.record(st)
            for _ in range(reps):
                f()
            e1.record(st)
            sync()
            return worst(e0.elapsed_time(e1)) / reps

        dev_ms = timeit(lambda: H.Gemm(0, 0, 1.0, A, B, 0.0, Cm, H.GEMM_SUMMA_C, blocksize=nb))
        host_ms = timeit(lambda: H.Gemm(0, 0, 1.0, hA, hB, 0.0, hC, H.GEMM_SUMMA_C, blocksize=nb))
        # one profiled call: where the device time goes (sum of GEMM launches, upload stream first-to-last copy, downloads)
        grid.Profile(True)
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        sync()
        e0.record(st)
        H.Gemm(0, 0, 1.0, hA, hB, 0.0, hC, H.GEMM_SUMMA_C, blocksize=nb)
        e1.record(st)
        torch.cuda.synchronize()
        prof = {"call_ms": round(e0.elapsed_time(e1), 2)}
        for kind, name in ((0, "gemm"), (5, "upload_stream"), (6, "downloads")):
            cnt, ms_ = grid.ProfileRead(kind)
            prof[name] = {"n": cnt, "ms": round(ms_, 2)}
        import ctypes as C
        from elemental_b200 import lib
        for kind, name in ((0, "gemm"), (5, "upload_stream"), (6, "downloads")):
            a, b = (C.c_double * 64)(), (C.c_double * 64)()
            cnt = lib.hyd_grid_profile_trace(grid.h, kind, 5, a, b, 64)
            prof[name + "_intervals_ms"] = [[round(a[i], 2), round(b[i], 2)] for i in range(max(cnt, 0))]
        grid.Profile(False)
        out["rank0_profile_of_one_call"] = prof
        ok = bool(torch.allclose(hC.store.cuda(), Cm.store, rtol=1e-9, atol=1e-9))
        out.update({"n": n, "device_resident_ms": round(dev_ms, 2), "host_resident_ms": round(host_ms, 2), "exposed_ms": round(host_ms - dev_ms, 2),
                    "e2e_tflops": round(2.0 * n ** 3 / host_ms / 1e9, 1), "matches_device_run": ok,
                    "bytes_per_gpu": {"h2d": int((hA.lh * hA.lw + hB.lh * hB.lw) * 8), "d2h": int(hC.lh * hC.lw * 8)}})
    if rank == 0:
        print("E2E " + json.dumps(out), flush=True)
    grid.destroy()
    if size > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Headline benchmark: distributed fp64 Gemm (El::Gemm NN, SUMMA-C, DistMatrix[MC,MR]) at
m=n=k=16384 -- BASELINE.json configs[1] -- on a 1x1 / 1x2 / 2x2 / 2x4 grid of 1/2/4/8 B200s.

A step = one Gemm over the whole matrices (alpha=1, beta=0 as tests/blas_like/Gemm.cpp:373).
`value` : TFLOP/s (2*m*n*k / device time, max over ranks), operands resident in HBM.
`e2e`   : same metric through hyd_gemm_host_f64 with the local blocks in pinned HOST memory
          (H2D of A,B and D2H of C inside the timed region).
`roofline`: the local DMMA GEMM kernel (dgemm_tma_kernel), algorithmic 2*m_loc*n_loc*k_panel flops per
          launch / mean launch duration from CUDA-event pairs recorded around every launch inside the
          timed region, against the FP64 tensor-pipe peak measured live by hyd_probe_peak (MEASURED_PEAKS.json
          carries no fp64 figure; its bf16 number does not bound an fp64 kernel).
`cpu_baseline` / --impl reference: the CPU restatement of the reference's SUMMA loops (oracle/) with
          OpenBLAS dgemm on all host cores, on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")  # before CUDA init: one hardware queue per engine stream
ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

M = N = K = 16384
GRIDS = {1: (1, 1), 2: (1, 2), 4: (2, 2), 8: (2, 4)}


def workload_name(n, grid):
    return f"SUMMA-C Gemm NN fp64 m=n=k={n} DistMatrix[MC,MR] on {grid[0]}x{grid[1]} grid (BASELINE.json configs[1])"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--nb", type=int, default=int(os.environ.get("HYD_BENCH_NB", "1024")), help="El::Blocksize() (SUMMA panel width)")
    ap.add_argument("--size", type=int, default=M)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------
# CPU arm: the oracle's restated SUMMA loops + OpenBLAS (the reference's CPU path: Gemm.cpp:83-101)
# ---------------------------------------------------------------------------------------------
_REF_SUMMA = None


def ref_summa_lib():
    """oracle/_ref/libref_summa.so: the reference's own SUMMA loop headers (src/blas_like/level3/Gemm/NN.hpp ...) compiled in
    place over the oracle's simulated grid (oracle/ref/ref_summa.cpp).  None when it was not built (no /root/reference at build
    time): the CPU arm then runs the oracle's restatement of the same loops."""
    global _REF_SUMMA
    if _REF_SUMMA is None:
        _REF_SUMMA = False
        path = os.path.join(ROOT, "oracle", "_ref", "libref_summa.so")
        if os.path.exists(path):
            try:
                import ctypes as C
                lib = C.CDLL(path)
                lib.ref_summa_gemm.restype = C.c_int
                lib.ref_summa_gemm.argtypes = [C.c_int] * 3 + [C.c_double, C.c_void_p, C.c_void_p, C.c_double, C.c_void_p]
                _REF_SUMMA = lib
            except OSError:
                pass
    return _REF_SUMMA or None


def cpu_gemm_tflops(n, grid, nb, steps=1, warmup=0):
    import oracle_lib as O
    blas = O.use_blas(True)
    if blas:  # torchrun exports OMP_NUM_THREADS=1: the CPU arm uses (and reports) every host core
        O.lib.orc_blas_threads(os.cpu_count() or 1)
    g = O.Grid(grid[0], grid[1], 1)
    A = O.DM(g, n, n, O.MC, O.MR).fill(1)
    B = O.DM(g, n, n, O.MC, O.MR).fill(2)
    Cm = O.DM(g, n, n, O.MC, O.MR).fill(3)
    ref = ref_summa_lib()
    if ref is not None:
        ref.ref_summa_set_blocksize(nb)

    def step():
        if ref is not None:   # El::Gemm(NORMAL, NORMAL, 1, A, B, 0, C, GEMM_SUMMA_C) through the reference's SUMMA_NNC loop
            assert ref.ref_summa_gemm(0, 0, 3, 1.0, A.h, B.h, 0.0, Cm.h) == 0
        else:
            O.gemm(0, 0, 3, 1.0, A, B, 0.0, Cm, nb)

    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    return 2.0 * n ** 3 / dt / 1e12, dt, blas


def cpu_kind():
    """("reference" | "port", what ran): the reference's loop code when oracle/_ref/libref_summa.so is there, else the restatement."""
    if ref_summa_lib() is not None:
        return "reference", "the reference's SUMMA_NNC loop (src/blas_like/level3/Gemm/NN.hpp:324-367, compiled in place over the oracle's simulated grid)"
    return "port", "oracle SUMMA-C (restated loop)"


def cpu_threads(blas):
    """Threads the CPU arm really ran on: OpenBLAS' own count (1 for the naive loop)."""
    import oracle_lib as O
    return (O.lib.orc_blas_threads(0) or (os.cpu_count() or 1)) if blas else 1


def pick_cpu_sample(budget_s, grid, nb):
    """Largest n in {2048,4096,8192} whose estimated time fits the budget (cubic extrapolation from 2048)."""
    tf, dt, _ = cpu_gemm_tflops(2048, grid, nb)
    n = 2048
    while n < 8192 and dt * 8 <= budget_s:
        n *= 2
        dt *= 8
    return n


def cpu_baseline(grid, nb):
    n = pick_cpu_sample(20.0, grid, 128)
    tf, dt, blas = cpu_gemm_tflops(n, grid, 128)
    cores = cpu_threads(blas)
    kind, what = cpu_kind()
    return {"value": round(tf, 4), "unit": "TFLOP/s", "cores": cores, "kind": kind,
            "sample": f"{what} NN fp64 m=n=k={n} (workload shape scaled from {M}), blocksize 128 (reference default), "
                      f"{grid[0]}x{grid[1]} simulated grid, {'OpenBLAS dgemm all cores' if blas else 'naive triple loop'}, {dt:.2f} s"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    grid = GRIDS[args.gpus]
    total = max(1, args.steps + args.warmup)
    n = pick_cpu_sample(120.0 / total, grid, 128)
    tf, dt, blas = cpu_gemm_tflops(n, grid, 128, steps=args.steps, warmup=args.warmup)
    cores = cpu_threads(blas)
    kind, what = cpu_kind()
    sample = (f"{what} NN fp64 m=n=k={n} per step (bounded sample of the {M}^3 workload), blocksize 128, "
              f"{grid[0]}x{grid[1]} simulated grid in one process, {'OpenBLAS dgemm all cores' if blas else 'naive loop'}")
    line = {"impl": "reference", "metric": "distributed Gemm TFLOP/s (fp64)", "value": round(tf, 4), "unit": "TFLOP/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt * 1e3, 3),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args.size, grid), "sample_n": n},
            "cpu_baseline": {"value": round(tf, 4), "unit": "TFLOP/s", "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": round(tf, 4), "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# ---------------------------------------------------------------------------------------------
class Clocks:
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,power.draw"

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.rows.append([x.strip() for x in ln.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.th.join(timeout=2)
        sm = sorted(int(float(r[0])) for r in self.rows if r and r[0].replace(".", "").isdigit())
        mx = [int(float(r[1])) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [nm for i, nm in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        pw = [float(r[6]) for r in self.rows if len(r) > 6 and r[6].replace(".", "").isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm), "power_w_max": max(pw) if pw else None}


def bind_to_gpu_numa_node(local_rank):
    """Multi-GPU runs: place this rank next to its GPU, so that the pinned host blocks of the end-to-end leg live on the
    GPU's own NUMA node and H2D/D2H do not cross the socket link (launchers such as torchrun do not bind ranks; at 4 and 8
    GPUs the round-1 e2e leg moved only ~55 GB/s host<->device in aggregate).  Two independent steps, each best effort and
    never fatal: (1) CPU affinity = the cores NVML reports as local to the GPU (intersected with the cores this process may
    use); (2) memory policy MPOL_PREFERRED = the NUMA node sysfs reports for the GPU's PCI function.  HYD_BENCH_NUMA=0
    disables both.  Returns what was done, for the JSON line.  (One rank keeps every core at N=1: the CPU baseline leg runs
    there.)"""
    done = {"cpu_bound": False, "mem_node": None}
    if os.environ.get("HYD_BENCH_NUMA", "1") == "0":
        return done
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
    except Exception:  # noqa: BLE001
        return done
    try:
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * w + b for w, word in enumerate(words) for b in range(64) if (int(word) >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            done["cpu_bound"] = True
    except Exception:  # noqa: BLE001
        pass
    try:
        import ctypes
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        bus = bus.lower()
        if len(bus.split(":")[0]) == 8:      # NVML prints an 8-digit PCI domain, sysfs a 4-digit one
            bus = bus[4:]
        with open(f"/sys/bus/pci/devices/{bus}/numa_node") as f:
            node = int(f.read().strip())
        if 0 <= node < 64:
            mask = ctypes.c_ulong(1 << node)
            libc = ctypes.CDLL(None, use_errno=True)
            MPOL_PREFERRED, SYS_set_mempolicy = 1, 238   # x86_64
            if libc.syscall(SYS_set_mempolicy, MPOL_PREFERRED, ctypes.byref(mask), 65) == 0:
                done["mem_node"] = node
    except Exception:  # noqa: BLE001
        pass
    return done


# ---------------------------------------------------------------------------------------------
def run_ours(args):
    import ctypes as C
    import numpy as np
    import torch
    from elemental_b200 import host as H
    from elemental_b200 import lib, check

    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert world_size == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world_size} (launch with torch.distributed.run)"
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path")
    torch.cuda.set_device(local_rank)
    numa_bound = bind_to_gpu_numa_node(local_rank) if world_size > 1 else None
    dist = None
    if world_size > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        world = H.World.from_torch_distributed()
    else:
        world = H.World.single()
    r, c = GRIDS[args.gpus]
    grid = H.Grid(world, r)
    n = args.size
    nb = args.nb
    stream = torch.cuda.Stream()
    flops = 2.0 * n * n * n

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    with torch.cuda.stream(stream):
        # synthetic operands: uniform [-1,1), generated on the device per rank (data: synthetic)
        A = H.DistMatrix(grid, n, n, H.MC, H.MR)
        B = H.DistMatrix(grid, n, n, H.MC, H.MR)
        Cm = H.DistMatrix(grid, n, n, H.MC, H.MR)
        gen = torch.Generator(device="cuda").manual_seed(1234 + rank)
        for d in (A, B, Cm):
            d.store.uniform_(-1.0, 1.0, generator=gen)
        step = lambda: H.Gemm(H.NORMAL, H.NORMAL, 1.0, A, B, 0.0, Cm, H.GEMM_SUMMA_C, blocksize=nb)
        for _ in range(args.warmup):
            step()
        barrier()
        grid.Profile(True)
        l0 = lib.hyd_launch_count()
        clk = Clocks(local_rank)
        if rank == 0:
            clk.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(stream)
        for _ in range(args.steps):
            step()
        e1.record(stream)
        barrier()
        clocks = clk.stop() if rank == 0 else None
        ms = max_over_ranks(e0.elapsed_time(e1)) / args.steps
        launches = (lib.hyd_launch_count() - l0)
        n_gemm, gemm_ms = grid.ProfileRead(0)
        n_red, red_ms = grid.ProfileRead(1)
        grid.Profile(False)
        value = flops / (ms * 1e-3) / 1e12

        # roofline of the dominant kernel (local DMMA GEMM)
        peak = C.c_double()
        check(lib.hyd_probe_peak(2, 20000, 2, C.byref(peak)))
        flops_per_launch = 2.0 * Cm.lh * Cm.lw * min(nb, n)
        achieved = flops_per_launch / (gemm_ms / max(n_gemm, 1) * 1e-3) / 1e12 if n_gemm else None
        roofline = {"bound": "tensor", "kernel": "dgemm_tma_kernel (FP64 DMMA m16n8k8, TMA-fed)",
                    "achieved": round(achieved, 3) if achieved else None, "peak": round(peak.value, 3), "unit": "TFLOP/s",
                    "frac": round(achieved / peak.value, 4) if achieved else None,
                    # DRAM bytes per launch from the ncu --set full capture of this launch shape (profiles/r01_ncu_dgemm_tma_kernel.txt:
                    # 3.37 GB read + 2.10 GB written; algorithmic 8*(2*m*n + (m+n)*k) = 4.56 GB); other shapes were not captured
                    "traffic": 5.48e9 if (args.gpus == 1 and nb == 1024 and n == 16384) else None, "traffic_unit": "bytes/launch",
                    "peak_source": "FP64 tensor pipe measured live (hyd_probe_peak: register-resident DMMA loop); "
                                   "MEASURED_PEAKS.json has no fp64 figure",
                    "flops_per_launch": flops_per_launch, "launches": n_gemm, "mean_launch_ms": round(gemm_ms / max(n_gemm, 1), 4),
                    "kernel_share_of_step": round(gemm_ms / (ms * args.steps), 4),
                    "redist_ms_per_step": round(red_ms / args.steps, 4)}

        # end to end: local blocks in pinned host memory, through hyd_gemm_host_f64
        def measure_e2e():
            hA = H.DistMatrix(grid, n, n, H.MC, H.MR, device="cpu", pinned=True)
            hB = H.DistMatrix(grid, n, n, H.MC, H.MR, device="cpu", pinned=True)
            hC = H.DistMatrix(grid, n, n, H.MC, H.MR, device="cpu", pinned=True)
            hA.store.copy_(A.store)
            hB.store.copy_(B.store)
            estep = lambda: H.Gemm(H.NORMAL, H.NORMAL, 1.0, hA, hB, 0.0, hC, H.GEMM_SUMMA_C, blocksize=nb)
            for _ in range(min(args.warmup, 2)):
                estep()
            esteps = max(1, min(args.steps, 3))
            barrier()
            e0.record(stream)
            for _ in range(esteps):
                estep()
            e1.record(stream)
            barrier()
            ems = max_over_ranks(e0.elapsed_time(e1)) / esteps
            es = 8
            e2e = {"value": round(flops / (ems * 1e-3) / 1e12, 3), "unit": "TFLOP/s", "ms_per_step": round(ems, 3),
                   "h2d_bytes_per_step": int((hA.lh * hA.lw + hB.lh * hB.lw) * es), "d2h_bytes_per_step": int(hC.lh * hC.lw * es),
                   "api": "hyd_gemm_host_f64 (El::Gemm on host-resident DistMatrix blocks)", "steps": esteps}
            # the host path streams a narrower first panel, so the k-sum is grouped differently: compare with the
            # device-resident run in the relative Frobenius norm (BASELINE.json tolerance 1e-13*sqrt(k) for fp64)
            dd = hC.store[: hC.lw, : hC.lh].cuda() - Cm.store[: Cm.lw, : Cm.lh]
            rel = float(dd.norm() / Cm.store[: Cm.lw, : Cm.lh].norm())
            e2e["rel_frobenius_vs_device_run"] = rel
            if rel > 1e-13 * n ** 0.5:
                bad = (dd.abs() > 1e-9).nonzero()
                print(f"[rank {rank}] e2e mismatch: rel={rel:.3e} bad={bad.shape[0]} first={bad[:8].tolist()} max={float(dd.abs().max()):.3e}",
                      file=sys.stderr, flush=True)
            assert rel <= 1e-13 * n ** 0.5, rel
            return e2e

        e2e = None
        if not args.no_e2e:
            # a failure of the end-to-end leg must not take the device-resident measurement down with it: the line is
            # printed with the error in place of the number
            # (one rank only: with several ranks a rank that skipped the leg's barriers would leave the others hanging, so
            # there the exception propagates and the launcher tears the job down at once)
            if world_size > 1:
                e2e = measure_e2e()
            else:
                try:
                    e2e = measure_e2e()
                except Exception as ex:  # noqa: BLE001
                    e2e = {"value": None, "unit": "TFLOP/s", "error": f"{type(ex).__name__}: {ex}"[:300]}
                    print(f"[rank {rank}] e2e leg failed: {ex}", file=sys.stderr, flush=True)

    cpu = None
    if rank == 0 and args.gpus == 1 and not args.no_cpu:
        cpu = cpu_baseline((r, c), nb)
    if rank == 0:
        line = {"metric": "distributed Gemm TFLOP/s (fp64)", "value": round(value, 3), "unit": "TFLOP/s", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms, 3), "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": workload_name(n, (r, c)),
                           "blocksize": nb, "alpha": 1.0, "beta": 0.0, "l2": "operands (3 x %.1f GiB per GPU) exceed the 126 MB L2" %
                           (A.lh * A.lw * 8 / 2 ** 30), "comm": world.kind, "rank0_numa_placement": numa_bound},
                "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
                "frac_of_fp64_peak_all_gpus": round(value / (peak.value * args.gpus), 4)}
        print(json.dumps(line), flush=True)
    barrier()
    grid.destroy()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)

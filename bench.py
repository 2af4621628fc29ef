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
          carries no fp64 figure; its bf16 number does not bound an fp64 kernel); the clock-derived bound
          SMs x 128 flop/clk x sm_mhz is printed beside it.
`parity`: before anything is timed, THIS world (same processes, same transport as the timed run) is checked
          against the CPU oracle: Gemm (the copy-engine gather path, the direct-push path with random alignments,
          the host-resident path) in the relative Frobenius norm, [MC,MR]->[*,VR], distributed Transpose and
          AxpyContract bit for bit.  The run aborts if any of it is off.
`extra` : the rest of BASELINE.json's metric with the same clocks record: fp64 at the reference's default
          Blocksize() = 128, fp32 TN / NT 32768^3 (configs[2]) against the TF32 tensor bound, the 32768^2 fp64
          redistributions (configs[3]) in GB/s against HBM / NVLink, the tall-skinny SUMMA-Dot (configs[4]),
          SUMMA-A/B at configs[1], and -- on 8 GPUs -- the fp64 n=32768 target run.
`cpu_baseline` / --impl reference: the reference's own SUMMA_NNC loop (oracle/_ref) or its restatement (oracle/)
          with OpenBLAS dgemm on all host cores, on a bounded sample of the same workload.
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
NVLINK_GBS = 770.0   # per-direction peer-copy rate of one B200 (B200_PROFILING.md); the redistribution roofline on > 1 GPU


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
    ap.add_argument("--transport", default=os.environ.get("HYD_TRANSPORT", "auto"), choices=["auto", "nccl"],
                    help="auto: NVLink peer-memory push when every rank can map every other (else NCCL); nccl: staged "
                         "pack -> grouped ncclSend/ncclRecv -> unpack")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extra", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
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
            "sample": f"{what} NN fp64 m=n=k={n} (a bounded sample: the workload's shape scaled down from {M}; rate over rate, "
                      f"not the same problem size), blocksize 128 (reference default), "
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
    sample = (f"{what} NN fp64 m=n=k={n} per step (bounded sample of the {M}^3 workload: same shape, smaller size), blocksize 128, "
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
                                          "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
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


def dram_traffic_per_launch(mloc, nloc, kb, dtype):
    """DRAM bytes per launch of the local GEMM from the committed `ncu --set full` captures (profiles/dram_traffic.json:
    dram__bytes_read.sum + dram__bytes_write.sum per launch, keyed by launch shape); None for a shape that was not captured."""
    try:
        table = json.load(open(os.path.join(ROOT, "profiles", "dram_traffic.json")))
        return table.get(f"{dtype}:{mloc}x{nloc}x{kb}", {}).get("bytes")
    except Exception:  # noqa: BLE001
        return None


# ---------------------------------------------------------------------------------------------
def run_ours(args):
    import ctypes as C
    import numpy as np
    import torch
    if args.transport == "nccl":
        os.environ["HYD_TRANSPORT"] = "nccl"     # read when the world is created (csrc/comm.cu init_push)
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
    P = r * c
    grid = H.Grid(world, r)
    n = args.size
    nb = args.nb
    stream = torch.cuda.Stream()
    flops = 2.0 * n * n * n

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce_over_ranks(x, op="max"):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else (dist.ReduceOp.MIN if op == "min" else dist.ReduceOp.SUM))
        return float(t.item())

    max_over_ranks = reduce_over_ranks

    def timed(fn, steps, warmup):
        """ms per step of fn(), CUDA events on `stream`, barrier + synchronize on both sides, max over ranks."""
        for _ in range(warmup):
            fn()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        t0.record(stream)
        for _ in range(steps):
            fn()
        t1.record(stream)
        barrier()
        return max_over_ranks(t0.elapsed_time(t1)) / steps

    # -----------------------------------------------------------------------------------------
    # parity of THIS world against the CPU oracle (small cases, same processes / transport as the timed run)
    # -----------------------------------------------------------------------------------------
    def parity_block():
        import oracle_lib as O
        import gpu_util as U
        O.use_blas(True)
        og = O.Grid(r, c, 1)
        rng = np.random.default_rng(20261105)     # the same stream on every rank
        out = {"transport": world.kind, "gemm": [], "redist_bit_exact": True, "contract_bit_exact": True}
        # --- Gemm fp64 NN SUMMA-C, alpha = 2, beta = 3 (BasicGemm.cpp:108): aligned (copy-engine gather path on the push
        # transport = the path the timed run takes), random alignments (direct push / staged path), host-resident blocks
        pn, pnb = 1024, 256
        for name, random_align, host in (("aligned", False, False), ("random alignments", True, False), ("host-resident", False, True)):
            al = (lambda d: U.rand_align(rng, og, d)) if random_align else (lambda d: 0)
            mk = lambda seed: O.DM(og, pn, pn, O.MC, O.MR, al(O.MC), al(O.MR)).fill(seed)
            A_, B_, Cx = mk(1), mk(2), mk(3)
            Cin = O.DM(og, pn, pn, O.MC, O.MR, Cx.ca, Cx.ra).fill(3)
            O.gemm(0, 0, 3, 2.0, A_, B_, 3.0, Cx, pnb)
            if host:
                def hd(X):
                    d = H.DistMatrix(grid, X.m, X.n, X.cd, X.rd, X.dtype, X.ca, X.ra, device="cpu", pinned=True)
                    if d.lh and d.lw:
                        d.store[: d.lw, : d.lh].copy_(torch.from_numpy(np.ascontiguousarray(X.block(rank).T)))
                    return d
                a, b, cm = hd(A_), hd(B_), hd(Cin)
            else:
                a, b, cm = U.to_rank(grid, A_), U.to_rank(grid, B_), U.to_rank(grid, Cin)
            H.Gemm(0, 0, 2.0, a, b, 3.0, cm, H.GEMM_SUMMA_C, blocksize=pnb)
            torch.cuda.synchronize()
            got, exp = cm.local_numpy(), Cx.block(rank)
            num = reduce_over_ranks(float(((got - exp) ** 2).sum()), "sum")
            den = reduce_over_ranks(float((exp ** 2).sum()), "sum")
            out["gemm"].append({"case": name, "rel_frobenius": (num / den) ** 0.5})
        out["gemm_rel"] = max(g["rel_frobenius"] for g in out["gemm"])
        out["gemm_tol"] = 1e-13 * pn ** 0.5
        # --- redistributions, bit for bit: [MC,MR] -> [*,VR] (RowAllToAllDemote), distributed Transpose, Copy -> [MR,MC], -> [VC,*]
        for (dcd, drd, tr, m_, n_) in ((O.STAR, O.VR, 0, 1029, 515), (O.MC, O.MR, 1, 1029, 515), (O.MR, O.MC, 0, 777, 1030),
                                       (O.VC, O.STAR, 0, 2050, 129)):
            A_ = O.DM(og, m_, n_, O.MC, O.MR, U.rand_align(rng, og, O.MC), U.rand_align(rng, og, O.MR)).fill_index()
            dm_, dn_ = (n_, m_) if tr else (m_, n_)
            B_ = O.DM(og, dm_, dn_, dcd, drd, U.rand_align(rng, og, dcd), U.rand_align(rng, og, drd)).fill_const(-77)
            O.lib.orc_redist_closed(A_.h, B_.h, tr)
            a = U.to_rank(grid, A_)
            b = U.to_rank(grid, O.DM(og, dm_, dn_, dcd, drd, B_.ca, B_.ra).fill_const(-77), ldim_extra=2)
            (H.Transpose if tr else H.Copy)(a, b)
            torch.cuda.synchronize()
            ok = bool(np.array_equal(U.block_of(b), B_.block(rank))) and U.padding_untouched(b)
            out["redist_bit_exact"] = out["redist_bit_exact"] and ok
        # --- AxpyContract [MC,*] -> [MC,MR] and [*,*] -> [MC,MR] (the reduce side of SUMMA-A / Dot), ascending-rank sum
        for (scd, srd) in ((O.MC, O.STAR), (O.STAR, O.STAR)):
            A_ = O.DM(og, 333, 290, scd, srd, U.rand_align(rng, og, scd), U.rand_align(rng, og, srd))
            for p in range(A_.P):
                A_.block(p)[...] = rng.standard_normal(A_.block(p).shape)
            ca, ra = U.rand_align(rng, og, O.MC), U.rand_align(rng, og, O.MR)
            B0 = O.DM(og, 333, 290, O.MC, O.MR, ca, ra).fill(9)
            B1 = O.DM(og, 333, 290, O.MC, O.MR, ca, ra).fill(9)
            O.lib.orc_axpy_contract_closed(1.5, A_.h, B1.h, 0)
            a, b = U.to_rank(grid, A_), U.to_rank(grid, B0)
            H.AxpyContract(1.5, a, b)
            torch.cuda.synchronize()
            out["contract_bit_exact"] = out["contract_bit_exact"] and bool(np.array_equal(U.block_of(b), B1.block(rank)))
        out["redist_bit_exact"] = reduce_over_ranks(float(out["redist_bit_exact"]), "min") == 1.0
        out["contract_bit_exact"] = reduce_over_ranks(float(out["contract_bit_exact"]), "min") == 1.0
        out["checked_against"] = "oracle/liboracle.so (CPU restatement pinned to the reference's loop code), every rank its own block"
        return out

    with torch.cuda.stream(stream):
        parity = None
        if not args.no_parity:
            parity = parity_block()
            assert parity["gemm_rel"] <= parity["gemm_tol"], parity
            assert parity["redist_bit_exact"] and parity["contract_bit_exact"], parity

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
        grid.Profile(False)
        value = flops / (ms * 1e-3) / 1e12

        # roofline of the dominant kernel (local DMMA GEMM)
        peak = C.c_double()
        check(lib.hyd_probe_peak(2, 20000, 2, C.byref(peak)))
        kb_full = min(nb, n)
        # the first panel of a multi-rank run is narrower (nb/4), so launches are not all 2*m_loc*n_loc*nb flops: use the exact
        # flops per step over the launches per step
        flops_per_launch = 2.0 * Cm.lh * Cm.lw * n / max(n_gemm / args.steps, 1)
        achieved = flops_per_launch / (gemm_ms / max(n_gemm, 1) * 1e-3) / 1e12 if n_gemm else None
        gemm_ms_per_step = gemm_ms / args.steps
        sms = torch.cuda.get_device_properties(local_rank).multi_processor_count
        # bytes this GPU receives per full-width panel: (c-1)/c of the A panel over the row comm + (r-1)/r of the B panel
        panel_bytes = int(((c - 1) / c * Cm.lh + (r - 1) / r * Cm.lw) * kb_full * 8)
        roofline = {"bound": "tensor", "kernel": "dgemm_tma_kernel (FP64 DMMA m16n8k8, TMA-fed)",
                    "achieved": round(achieved, 3) if achieved else None, "peak": round(peak.value, 3), "unit": "TFLOP/s",
                    "frac": round(achieved / peak.value, 4) if achieved else None,
                    "traffic": dram_traffic_per_launch(Cm.lh, Cm.lw, kb_full, "f64"), "traffic_unit": "bytes/launch",
                    "traffic_source": "profiles/dram_traffic.json (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum per launch; null = shape not captured)",
                    "algorithmic_bytes_per_launch": 8 * (2 * Cm.lh * Cm.lw + (Cm.lh + Cm.lw) * kb_full),
                    "peak_source": "FP64 tensor pipe measured live (hyd_probe_peak: register-resident DMMA loop); "
                                   "MEASURED_PEAKS.json has no fp64 figure",
                    "flops_per_launch": flops_per_launch, "launches": n_gemm, "mean_launch_ms": round(gemm_ms / max(n_gemm, 1), 4),
                    "kernel_share_of_step": round(gemm_ms / (ms * args.steps), 4),
                    # what the step spends outside the local GEMMs on the compute stream: panel gathers that did not hide,
                    # flag waits, launch gaps (NOT the comm stream's elapsed time, which includes waiting for the GEMMs)
                    "exposed_non_gemm_ms_per_step": round(max(ms - gemm_ms_per_step, 0.0), 4),
                    "panel_gather_bytes_per_gpu": panel_bytes}
        if rank == 0 and clocks and clocks.get("sm_mhz"):
            roofline["clock_bound_tflops"] = round(sms * 128 * clocks["sm_mhz"] * 1e6 / 1e12, 3)
            roofline["clock_bound_note"] = "SMs x 128 fp64 flop/clk/SM (DMMA) x median SM MHz under load"

        # end to end: local blocks in pinned host memory, through hyd_gemm_host_f64
        def measure_e2e():
            hA = H.DistMatrix(grid, n, n, H.MC, H.MR, device="cpu", pinned=True)
            hB = H.DistMatrix(grid, n, n, H.MC, H.MR, device="cpu", pinned=True)
            hC = H.DistMatrix(grid, n, n, H.MC, H.MR, device="cpu", pinned=True)
            hA.store.copy_(A.store)
            hB.store.copy_(B.store)
            estep = lambda: H.Gemm(H.NORMAL, H.NORMAL, 1.0, hA, hB, 0.0, hC, H.GEMM_SUMMA_C, blocksize=nb)
            esteps = max(1, min(args.steps, 5))
            ems = timed(estep, esteps, min(args.warmup, 2))
            es = 8
            h2d_mine, d2h_mine = float((hA.lh * hA.lw + hB.lh * hB.lw) * es), float(hC.lh * hC.lw * es)
            # whole-job bytes (summed over the ranks, like `value`), and this rank's share
            e2e = {"value": round(flops / (ems * 1e-3) / 1e12, 3), "unit": "TFLOP/s", "ms_per_step": round(ems, 3),
                   "h2d_bytes_per_step": int(reduce_over_ranks(h2d_mine, "sum")), "d2h_bytes_per_step": int(reduce_over_ranks(d2h_mine, "sum")),
                   "h2d_bytes_per_step_rank0": int(h2d_mine), "d2h_bytes_per_step_rank0": int(d2h_mine),
                   "api": "hyd_gemm_host_f64 (El::Gemm on host-resident DistMatrix blocks): A by k-panels and B by column blocks "
                          "upload under the GEMMs, C is computed in column blocks and each finished block downloads under the next; "
                          "on several GPUs the panels are gathered by SM kernels because the copy engines carry the host transfers",
                   "steps": esteps}
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

        # -------------------------------------------------------------------------------------
        # extras: the rest of BASELINE.json's metric, each with its own clocks record
        # -------------------------------------------------------------------------------------
        extra = []

        def with_clocks(fn):
            ck = Clocks(local_rank)
            if rank == 0:
                ck.start()
            res = fn()
            rec = ck.stop() if rank == 0 else None
            for x in (res if isinstance(res, list) else [res]):
                x["clocks"] = rec
            return res

        def associativity(oA, oB, Aa, Bb, Cc, dtype, nn, nbb):
            """The reference's own Gemm test (tests/blas_like/Gemm.cpp:27-61): || (op(A) op(B)) X - op(A) (op(B) X) ||_F /
            || op(A) (op(B) X) ||_F for 16 random right-hand sides, Cc = op(A) op(B) as left by the timed run (alpha 1, beta 0)."""
            mk = lambda: H.DistMatrix(grid, nn, 16, H.MC, H.MR, dtype=dtype)
            X, Y, Z, W = mk(), mk(), mk(), mk()
            X.store.uniform_(-1.0, 1.0, generator=torch.Generator(device="cuda").manual_seed(77 + rank))
            H.Gemm(oB, H.NORMAL, 1.0, Bb, X, 0.0, Z, blocksize=nbb)
            H.Gemm(oA, H.NORMAL, 1.0, Aa, Z, 0.0, Y, blocksize=nbb)
            H.Gemm(H.NORMAL, H.NORMAL, 1.0, Cc, X, 0.0, W, blocksize=nbb)
            torch.cuda.synchronize()
            d = (W.store[: W.lw, : W.lh].double() - Y.store[: Y.lw, : Y.lh].double())
            e2 = reduce_over_ranks(float((d ** 2).sum()), "sum")
            y2 = reduce_over_ranks(float((Y.store[: Y.lw, : Y.lh].double() ** 2).sum()), "sum")
            return (e2 / max(y2, 1e-300)) ** 0.5

        def extra_fp64_nb128():
            ms_ = timed(lambda: H.Gemm(H.NORMAL, H.NORMAL, 1.0, A, B, 0.0, Cm, H.GEMM_SUMMA_C, blocksize=128), 2, 1)
            tf = flops / ms_ / 1e9
            return {"name": f"fp64 NN m=n=k={n} SUMMA-C at the reference's default Blocksize() = 128 (src/core/environment.cpp:305-306)",
                    "ms": round(ms_, 3), "tflops": round(tf, 3), "frac_of_fp64_peak_all_gpus": round(tf / (peak.value * P), 4)}

        def extra_variants():
            res = []
            for alg, nm in ((H.GEMM_SUMMA_A, "A"), (H.GEMM_SUMMA_B, "B")):
                grid.Profile(True)
                ms_ = timed(lambda: H.Gemm(H.NORMAL, H.NORMAL, 1.0, A, B, 0.0, Cm, alg, blocksize=nb), 2, 1)
                ng, gms = grid.ProfileRead(0)
                grid.Profile(False)
                tf = flops / ms_ / 1e9
                res.append({"name": f"fp64 NN m=n=k={n} SUMMA-{nm} (stationary {nm}), nb={nb}, gather / GEMM / reduce-scatter pipelined on three streams",
                            "ms": round(ms_, 3), "tflops": round(tf, 3), "frac_of_fp64_peak_all_gpus": round(tf / (peak.value * P), 4),
                            "exposed_non_gemm_ms": round(max(ms_ - gms / 3, 0.0), 3)})
            return res

        def extra_fp32():
            f32 = np.float32
            n32 = int(os.environ.get("HYD_BENCH_N32", "32768"))
            tfpeak = C.c_double()
            check(lib.hyd_probe_peak(4, 4000, 1, C.byref(tfpeak)))
            A32 = H.DistMatrix(grid, n32, n32, H.MC, H.MR, dtype=f32)
            B32 = H.DistMatrix(grid, n32, n32, H.MC, H.MR, dtype=f32)
            C32 = H.DistMatrix(grid, n32, n32, H.MC, H.MR, dtype=f32)
            for d in (A32, B32):
                d.store.uniform_(-1.0, 1.0, generator=gen)
            res = []
            for (oa, ob, nm) in ((H.TRANSPOSE, H.NORMAL, "TN"), (H.NORMAL, H.TRANSPOSE, "NT")):
                ms_ = timed(lambda: H.Gemm(oa, ob, 1.0, A32, B32, 0.0, C32, H.GEMM_SUMMA_C, blocksize=nb), 2, 1)
                tf = 2.0 * n32 ** 3 / ms_ / 1e9
                rel = associativity(oa, ob, A32, B32, C32, f32, n32, nb)
                assert rel <= 1e-5 * n32 ** 0.5, (nm, rel)
                res.append({"name": f"fp32 {nm} m=n=k={n32} SUMMA-C, nb={nb} (BASELINE.json configs[2]); split-TF32 (3 tcgen05 kind::tf32 MMAs per product)",
                            "ms": round(ms_, 3), "tflops": round(tf, 2), "tflops_per_gpu": round(tf / P, 2),
                            "roofline": {"bound": "tensor", "peak": round(tfpeak.value / 3, 2), "unit": "TFLOP/s (fp32-equivalent)",
                                         "frac": round(tf / P / (tfpeak.value / 3), 4),
                                         "peak_source": f"dense TF32 tcgen05 rate measured live (hyd_probe_peak kind 4: {tfpeak.value:.1f} TFLOP/s) / 3 MMAs per fp32 product"},
                            "associativity_rel": rel, "associativity_tol": 1e-5 * n32 ** 0.5})
            return res

        def extra_dot():
            f32 = np.float32
            m_, n_, k_ = 16384, 512, 16384
            Ad = H.DistMatrix(grid, m_, k_, H.STAR, H.VC, dtype=f32)
            Bd = H.DistMatrix(grid, n_, k_, H.STAR, H.VC, dtype=f32)
            Cd = H.DistMatrix(grid, m_, n_, H.MC, H.MR, dtype=f32)
            for d in (Ad, Bd):
                d.store.uniform_(-1.0, 1.0, generator=gen)
            ms_ = timed(lambda: H.Gemm(H.NORMAL, H.TRANSPOSE, 1.0, Ad, Bd, 0.0, Cd, H.GEMM_SUMMA_DOT), 10, 2)
            tf = 2.0 * m_ * n_ * k_ / ms_ / 1e9
            return {"name": f"fp32 NT m={m_} n={n_} k={k_}, A[*,VC] B[*,VC] C[MC,MR], SUMMA-Dot (BASELINE.json configs[4])",
                    "ms": round(ms_, 4), "tflops": round(tf, 2), "tflops_per_gpu": round(tf / P, 2)}

        def extra_redist():
            nr = int(os.environ.get("HYD_BENCH_NR", "32768"))
            hbm = None
            try:
                hbm = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs")
            except Exception:  # noqa: BLE001
                pass
            hbm = hbm or 6500.0
            S = H.DistMatrix(grid, nr, nr, H.MC, H.MR)
            S.store.uniform_(-1.0, 1.0, generator=gen)
            res = []
            for (cd, rd, tr, nm) in ((H.STAR, H.VR, 0, "[MC,MR] -> [*,VR] (RowAllToAllDemote)"),
                                     (H.MC, H.MR, 1, "Transpose [MC,MR] -> [MC,MR]"),
                                     (H.MR, H.MC, 0, "Copy [MC,MR] -> [MR,MC] (TransposeDist)")):
                D = H.DistMatrix(grid, nr, nr, cd, rd)
                Bk = H.DistMatrix(grid, nr, nr, H.MC, H.MR)
                fwd = (lambda: H.Transpose(S, D)) if tr else (lambda: H.Copy(S, D))
                back = (lambda: H.Transpose(D, Bk)) if tr else (lambda: H.Copy(D, Bk))
                ms_ = timed(fwd, 5, 2)
                back()
                torch.cuda.synchronize()
                exact = reduce_over_ranks(float(torch.equal(Bk.store[: Bk.lw, : Bk.lh], S.store[: S.lw, : S.lh])), "min") == 1.0
                assert exact, nm
                # algorithmic bytes per GPU: HBM = read the local block once + write it once; NVLink (one direction)
                # = the part of the local block that leaves the GPU (from the host-side planner)
                local_bytes = S.lh * S.lw * 8
                sb = (C.c_int64 * P)()
                rb = (C.c_int64 * P)()
                sd, dd_ = S.desc(), D.desc()
                check(lib.hyd_redist_plan(r, c, 1, rank, C.byref(sd), C.byref(dd_), tr, 8, sb, rb))
                out_bytes = max_over_ranks(float(sum(sb[p] for p in range(P) if p != rank)))
                hbm_gbs = 2.0 * local_bytes / ms_ / 1e6
                row = {"name": f"redistribution {nm}, {nr}x{nr} fp64 (BASELINE.json configs[3])", "ms": round(ms_, 4),
                       "round_trip_bit_exact": exact, "hbm_gbs_per_gpu": round(hbm_gbs, 1)}
                nv = out_bytes / ms_ / 1e6
                if P == 1 or out_bytes == 0 or hbm_gbs / hbm >= nv / NVLINK_GBS:
                    # nothing (or little) leaves the GPU: the local read + write is what bounds it
                    row["roofline"] = {"bound": "hbm", "achieved": round(hbm_gbs, 1), "peak": hbm, "unit": "GB/s",
                                       "frac": round(hbm_gbs / hbm, 4), "peak_source": "MEASURED_PEAKS.json hbm_gbs (copy rate)"}
                else:
                    row["nvlink_out_bytes_per_gpu"] = int(out_bytes)
                    row["roofline"] = {"bound": "nvlink", "achieved": round(nv, 1), "peak": NVLINK_GBS, "unit": "GB/s out per GPU",
                                       "frac": round(nv / NVLINK_GBS, 4),
                                       "peak_source": "770 GB/s per-direction peer-copy rate, every GPU pushing at once (B200_PROFILING.md)",
                                       "hbm_frac": round(hbm_gbs / hbm, 4)}
                res.append(row)
                del D, Bk
            return res

        def extra_target_32768():
            nt = 32768
            At = H.DistMatrix(grid, nt, nt, H.MC, H.MR)
            Bt = H.DistMatrix(grid, nt, nt, H.MC, H.MR)
            Ct = H.DistMatrix(grid, nt, nt, H.MC, H.MR)
            for d in (At, Bt):
                d.store.uniform_(-1.0, 1.0, generator=gen)
            ms_ = timed(lambda: H.Gemm(H.NORMAL, H.NORMAL, 1.0, At, Bt, 0.0, Ct, H.GEMM_SUMMA_C, blocksize=nb), 3, 1)
            tf = 2.0 * nt ** 3 / ms_ / 1e9
            rel = associativity(H.NORMAL, H.NORMAL, At, Bt, Ct, np.float64, nt, nb)
            assert rel <= 1e-13 * nt ** 0.5, rel
            return {"name": f"TARGET: fp64 NN m=n=k={nt} SUMMA-C DistMatrix[MC,MR] on the {r}x{c} grid, nb={nb} (north_star: >= 70 % of aggregate FP64 tensor peak)",
                    "ms": round(ms_, 3), "tflops": round(tf, 2), "frac_of_fp64_peak_all_gpus": round(tf / (peak.value * P), 4),
                    "associativity_rel": rel, "associativity_tol": 1e-13 * nt ** 0.5}

        if not args.no_extra and os.environ.get("HYD_BENCH_EXTRA", "1") != "0":
            todo = [extra_fp64_nb128]
            if P > 1:
                todo.append(extra_variants)
            if P == 8 or os.environ.get("HYD_BENCH_TARGET") == "1":
                todo.append(extra_target_32768)
            for fn in todo + [extra_dot, extra_redist, extra_fp32]:
                try:
                    res = with_clocks(fn)
                    extra.extend(res if isinstance(res, list) else [res])
                except AssertionError as ex:
                    # every check inside an extra is decided on values reduced over the ranks, so all ranks arrive here together:
                    # the failure is reported in the line (and fails the line's reader), the headline numbers are kept
                    extra.append({"name": fn.__name__, "error": f"CHECK FAILED: {ex}"[:300]})
                except Exception as ex:  # noqa: BLE001
                    if world_size > 1:
                        raise
                    extra.append({"name": fn.__name__, "error": f"{type(ex).__name__}: {ex}"[:300]})
                torch.cuda.empty_cache()

    cpu = None
    if rank == 0 and args.gpus == 1 and not args.no_cpu:
        cpu = cpu_baseline((r, c), nb)
    if rank == 0:
        line = {"metric": "distributed Gemm TFLOP/s (fp64)", "value": round(value, 3), "unit": "TFLOP/s", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms, 3), "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": workload_name(n, (r, c)),
                           "blocksize": nb, "alpha": 1.0, "beta": 0.0, "l2": "operands (3 x %.1f GiB per GPU) exceed the 126 MB L2" %
                           (A.lh * A.lw * 8 / 2 ** 30), "comm": world.kind,
                           "comm_note": "data plane as reported by the library (hyd_world_kind): nvlink-push(ipc)+nccl = panels and "
                                        "redistributions move as peer-memory stores / copy-engine pushes over NVLink, NCCL only bootstraps "
                                        "and synchronises; nccl = staged pack -> grouped ncclSend/ncclRecv -> unpack (--transport nccl, "
                                        "or the fallback when CUDA IPC mapping fails)",
                           "rank0_numa_placement": numa_bound},
                "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
                "frac_of_fp64_peak_all_gpus": round(value / (peak.value * args.gpus), 4), "parity": parity, "extra": extra}
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

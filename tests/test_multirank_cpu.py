"""CPU-only, world_size-2 gloo: the host-side logic of the N>1 path -- NCCL-id bootstrap over torch.distributed,
per-rank local shapes tiling the global matrix, and the redistribution planner's send/recv symmetry ACROSS
processes (each rank plans only for itself; the counts are exchanged with all_gather)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, size, port, ret):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import ctypes as C
    from elemental_b200 import lib, hyd_dm
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=size)
    try:
        # 1. bootstrap payload: rank 0's 128-byte id reaches every rank unchanged
        obj = [bytes(range(128)) if rank == 0 else None]
        dist.broadcast_object_list(obj, src=0)
        assert obj[0] == bytes(range(128))
        # 2. local shapes of a [MC,MR] and a [STAR,VR] matrix tile the global matrix
        r, c = 1, size
        for (cd, rd) in ((0, 2), (5, 4)):
            d = hyd_dm(None, 1, 101, 67, cd, rd, 0, 1 % max(1, size), 0, 0)
            cs, rs, lh, lw = (C.c_int64() for _ in range(4))
            assert lib.hyd_local_shape_of(r, c, 1, rank, C.byref(d), C.byref(cs), C.byref(rs), C.byref(lh), C.byref(lw)) == 0
            t = torch.tensor([lh.value * lw.value])
            dist.all_reduce(t)
            assert int(t.item()) == 101 * 67
        # 3. planner symmetry across processes
        s = hyd_dm(None, 1, 101, 67, 0, 2, 0, 1, 0, 0)        # [MC,MR]
        d = hyd_dm(None, 1, 67, 101, 2, 5, 1, 0, 0, 0)        # ^T -> [MR,*]
        send = np.zeros(size, dtype=np.int64)
        recv = np.zeros(size, dtype=np.int64)
        assert lib.hyd_redist_plan(r, c, 1, rank, C.byref(s), C.byref(d), 1, 8, send.ctypes.data_as(C.POINTER(C.c_int64)),
                                   recv.ctypes.data_as(C.POINTER(C.c_int64))) == 0
        allsend = [torch.zeros(size, dtype=torch.int64) for _ in range(size)]
        dist.all_gather(allsend, torch.from_numpy(send))
        for p in range(size):
            assert int(allsend[p][rank]) == int(recv[p]), (p, rank)
        # 4. max-over-ranks reduction used by bench.py
        t = torch.tensor([float(rank + 1)], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        assert t.item() == float(size)
        # 5. the launcher-environment bootstrap of hyd_world_create_env (mpi::NewWorldComm() of the shim), its host-only part:
        # rank / size from the environment, rank 0's 128 bytes reach the other rank through the rendezvous file (both
        # processes share their parent, as the ranks of one launch do), and rank 0 removes the file afterwards
        os.environ.update(RANK=str(rank), WORLD_SIZE=str(size), LOCAL_RANK=str(rank))
        r_, n_, l_ = C.c_int(-1), C.c_int(-1), C.c_int(-1)
        assert lib.hyd_env_world_info(C.byref(r_), C.byref(n_), C.byref(l_)) == 0 and (r_.value, n_.value, l_.value) == (rank, size, rank)
        ident = (C.c_ubyte * 128)(*([(7 * i + 3) % 256 for i in range(128)] if rank == 0 else [0] * 128))
        assert lib.hyd_env_exchange_id(rank, ident, 30) == 0, lib.hyd_last_error()
        assert list(ident) == [(7 * i + 3) % 256 for i in range(128)]
        dist.barrier()
        assert lib.hyd_env_exchange_done(rank) == 0
        dist.barrier()
        import glob
        assert not glob.glob(f"/tmp/hyd_rdv_{os.getuid()}_{port}_*"), "rendezvous file left behind"
        # 6. the column-block plan of the host-resident Gemm is a pure function of global facts: identical on every rank
        cuts = (C.c_int64 * 9)()
        nb = lib.hyd_host_block_plan(64, 32, 144, 0.087, 0.05, 8, cuts)
        mine = torch.tensor([nb] + list(cuts[: nb + 1]) + [0] * (9 - nb), dtype=torch.int64)
        both = [torch.zeros_like(mine) for _ in range(size)]
        dist.all_gather(both, mine)
        assert all(torch.equal(b, mine) for b in both)
        ret[rank] = 1
    finally:
        dist.destroy_process_group()


def test_two_rank_host_logic_gloo():
    size = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(size, _free_port(), ret), nprocs=size, join=True)
    assert dict(ret) == {0: 1, 1: 1}


_RANK_SNIPPET = r"""
import ctypes as C, os, sys
sys.path.insert(0, sys.argv[1])
from elemental_b200 import lib
rank = int(os.environ["RANK"])
r_, n_, l_ = C.c_int(-1), C.c_int(-1), C.c_int(-1)
assert lib.hyd_env_world_info(C.byref(r_), C.byref(n_), C.byref(l_)) == 0 and (r_.value, n_.value) == (rank, 2)
ident = (C.c_ubyte * 128)(*([(11 * i + 5) % 256 for i in range(128)] if rank == 0 else [0] * 128))
assert lib.hyd_env_exchange_id(rank, ident, 30) == 0, lib.hyd_last_error()
assert list(ident) == [(11 * i + 5) % 256 for i in range(128)]
print("RANK", rank, "GOT THE ID", flush=True)
"""


def test_env_rendezvous_survives_wrapper_processes(tmp_path):
    """The ranks of one launch need not share their parent: under `timeout` (or `env`, `numactl`, ...) every rank has a wrapper
    process of its own.  The first N > 1 run of examples/DistGemm on real GPUs hung exactly there (the rendezvous file was keyed
    by the parent pid; it is keyed by the launcher's job id or the session now).  Two rank processes, each under its own
    `timeout`, in this session: rank 1 must receive rank 0's 128 bytes; also a stale file (older than a minute) is ignored."""
    import subprocess
    import time
    port = str(_free_port())
    env = {k: v for k, v in os.environ.items() if k not in ("TORCHELASTIC_RUN_ID", "SLURM_JOB_ID", "PMIX_NAMESPACE", "HYD_RENDEZVOUS_FILE")}
    env.update(WORLD_SIZE="2", MASTER_PORT=port, HYD_RENDEZVOUS_DIR=str(tmp_path))
    procs = []
    for rank in (1, 0):        # rank 1 first: it has to wait for the file
        procs.append(subprocess.Popen(["timeout", "60", sys.executable, "-c", _RANK_SNIPPET, ROOT], env=dict(env, RANK=str(rank), LOCAL_RANK=str(rank)),
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
        time.sleep(0.3)
    outs = [p.communicate(timeout=90)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert "RANK 1 GOT THE ID" in outs[0] and "RANK 0 GOT THE ID" in outs[1], outs
    files = list(tmp_path.iterdir())
    assert len(files) == 1 and f"_{port}_sid" in files[0].name, files       # keyed by the session, not by a parent pid
    # the same name, left behind by a dead launch long ago: a new rank 1 must not take it for its own
    old = time.time() - 600
    os.utime(files[0], (old, old))
    p = subprocess.run([sys.executable, "-c", _RANK_SNIPPET.replace(", 30)", ", 2)"), ROOT], env=dict(env, RANK="1", LOCAL_RANK="1"),
                       capture_output=True, text=True, timeout=60)
    assert p.returncode != 0 and "never published" in (p.stdout + p.stderr), p.stdout + p.stderr

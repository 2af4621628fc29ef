"""Python mirror of the reference's operator interface for the distributed-GEMM path.

Names, argument meaning and error behaviour follow the reference (C++):
  El::Grid          include/El/core/Grid.hpp:15-146, src/core/Grid.cpp:66-216
  El::DistMatrix    include/El/core/DistMatrix/**  (ELEMENT wrap, Device::GPU)
  El::Gemm          include/El/blas_like/level3.hpp:66-84, src/blas_like/level3/Gemm.cpp:278-307
  El::Copy / operator=, El::Transpose, El::AxpyContract
  El::SetBlocksize / El::Blocksize   src/core/environment.cpp:305-306

This file is plumbing for tests and bench.py (the C++ shim is include/El.hpp); every
operation goes through the C ABI of libhydrogen_b200.so.  torch supplies device
memory and streams only.  There is no CPU fallback.
"""
import ctypes as C
import threading

import numpy as np
import torch

from ._lib import lib, check, hyd_dm, HydrogenError

MC, MD, MR, VC, VR, STAR, CIRC = range(7)
NORMAL, TRANSPOSE, ADJOINT = 0, 1, 2
GEMM_DEFAULT, GEMM_SUMMA_A, GEMM_SUMMA_B, GEMM_SUMMA_C, GEMM_SUMMA_DOT, GEMM_CANNON = range(6)
ROW_MAJOR, COLUMN_MAJOR = 0, 1
LOWER, UPPER = 0, 1   # El::UpperOrLower, include/El/core/types.hpp:489-493
LEFT, RIGHT = 0, 1    # El::LeftOrRight, types.hpp:417-424

_blocksize_stack = [128]  # environment.cpp:305-306


def Blocksize():
    return _blocksize_stack[-1]


def SetBlocksize(nb):
    _blocksize_stack[-1] = int(nb)


def PushBlocksizeStack(nb):
    _blocksize_stack.append(int(nb))


def PopBlocksizeStack():
    _blocksize_stack.pop()


COMM_MC, COMM_MR, COMM_VC, COMM_VR = range(4)  # Grid::MCComm() (my grid column), MRComm() (my grid row), VCComm(), VRComm()


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _stream_ptr(stream=None):
    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)


class World:
    """Communication world: one rank per GPU over NCCL, or P rank-threads on one GPU (tests)."""

    def __init__(self, handle, rank, size):
        self.h, self.rank, self.size = handle, rank, size

    @property
    def kind(self):
        """The data plane the library actually set up (hyd_world_kind): "nvlink-push(ipc)+nccl", "nccl" (also what a failed
        IPC mapping falls back to), "local-push" or "local"."""
        return lib.hyd_world_kind(self.h).decode()

    @staticmethod
    def nccl_unique_id():
        buf = (C.c_char * 128)()
        check(lib.hyd_nccl_unique_id(buf))
        return bytes(buf)

    @classmethod
    def nccl(cls, uid, rank, size):
        h = C.c_void_p()
        check(lib.hyd_world_create_nccl(C.c_char_p(uid), rank, size, C.byref(h)))
        return cls(h, rank, size)

    @classmethod
    def from_torch_distributed(cls):
        """Bootstrap the NCCL world from an initialised torch.distributed process group (any backend)."""
        import torch.distributed as dist
        rank, size = dist.get_rank(), dist.get_world_size()
        obj = [cls.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(obj, src=0)
        return cls.nccl(obj[0], rank, size)

    @classmethod
    def local(cls, group, rank):
        h = C.c_void_p()
        check(lib.hyd_world_create_local(group.h, rank, C.byref(h)))
        return cls(h, rank, group.size)

    @classmethod
    def single(cls):
        grp = LocalGroup(1)
        w = cls.local(grp, 0)
        w._grp = grp
        return w

    def barrier(self):
        check(lib.hyd_world_barrier(self.h))

    def destroy(self):
        if self.h:
            lib.hyd_world_destroy(self.h)
            self.h = None


class LocalGroup:
    def __init__(self, size):
        self.h = C.c_void_p()
        self.size = size
        check(lib.hyd_local_group_create(size, C.byref(self.h)))

    def destroy(self):
        if self.h:
            lib.hyd_local_group_destroy(self.h)
            self.h = None


class Grid:
    """El::Grid(comm, height, order)."""

    def __init__(self, world, height=None, order=COLUMN_MAJOR):
        if height is None:  # Grid::DefaultHeight, Grid.cpp:58-64: from floor(sqrt(p)) upwards until it divides p
            height = lib.hyd_grid_default_height(world.size)
        self.world = world
        self.h = C.c_void_p()
        check(lib.hyd_grid_create(world.h, height, order, C.byref(self.h)))

    def Height(self): return lib.hyd_grid_height(self.h)
    def Width(self): return lib.hyd_grid_width(self.h)
    def Size(self): return lib.hyd_grid_size(self.h)
    def Rank(self): return lib.hyd_grid_rank(self.h)
    def Row(self): return lib.hyd_grid_row(self.h)
    def Col(self): return lib.hyd_grid_col(self.h)
    MCRank, MRRank = Row, Col
    def VCRank(self): return lib.hyd_grid_vc_rank(self.h)
    def VRRank(self): return lib.hyd_grid_vr_rank(self.h)
    def Stride(self, dist): return lib.hyd_dist_stride(self.h, dist)
    def DistRank(self, dist): return lib.hyd_dist_rank(self.h, dist)

    def SetOverlap(self, overlap=True, sm_reserve=-1):
        check(lib.hyd_grid_set_overlap(self.h, int(overlap), sm_reserve))

    def SetHostBlocks(self, max_blocks):
        check(lib.hyd_grid_set_host_blocks(self.h, int(max_blocks)))

    def Profile(self, enable=True):
        check(lib.hyd_grid_profile(self.h, int(enable)))

    def ProfileRead(self, kind=0):
        """(launches, total device ms) of kind 0 = local GEMM, 1 = redistribution, 2 = reduce-scatter,
        3 = pack (+ peer push) kernels, 4 = unpack kernels of redistributions."""
        n, ms = C.c_longlong(), C.c_double()
        check(lib.hyd_grid_profile_read(self.h, kind, C.byref(n), C.byref(ms)))
        return n.value, ms.value

    # -- sub-communicators and the typed collectives on them (mpi::AllGather ... on Grid::MCComm() etc.; seam S2)
    def CommSize(self, comm): return lib.hyd_comm_size(self.h, comm)
    def CommRank(self, comm): return lib.hyd_comm_rank(self.h, comm)
    def CommWorldRank(self, comm, k): return lib.hyd_comm_world_rank(self.h, comm, k)

    def AllGather(self, comm, sbuf, rbuf, stream=None):
        """mpi::AllGather(sbuf, sc, rbuf, rc, comm, syncInfo): member k's sbuf lands at rbuf[k*sc:(k+1)*sc]."""
        check(lib.hyd_allgather(self.h, comm, _ptr(sbuf), sbuf.numel() * sbuf.element_size(), _ptr(rbuf), _stream_ptr(stream)))

    def AllToAll(self, comm, sbuf, rbuf, stream=None):
        n = sbuf.numel() * sbuf.element_size() // max(1, self.CommSize(comm))
        check(lib.hyd_alltoall(self.h, comm, _ptr(sbuf), n, _ptr(rbuf), _stream_ptr(stream)))

    def SendRecv(self, comm, sbuf, to, rbuf, frm, stream=None):
        check(lib.hyd_sendrecv(self.h, comm, _ptr(sbuf), sbuf.numel() * sbuf.element_size(), to, _ptr(rbuf),
                               rbuf.numel() * rbuf.element_size(), frm, _stream_ptr(stream)))

    def Broadcast(self, comm, buf, root, stream=None):
        check(lib.hyd_bcast(self.h, comm, _ptr(buf), buf.numel() * buf.element_size(), root, _stream_ptr(stream)))

    def ReduceScatter(self, comm, sbuf, rbuf, stream=None):
        """mpi::ReduceScatter(sbuf, rbuf, rc, comm, syncInfo): rbuf = sum over members of their block [me*rc:(me+1)*rc]."""
        fn = lib.hyd_reducescatter_f64 if sbuf.dtype == torch.float64 else lib.hyd_reducescatter_f32
        check(fn(self.h, comm, _ptr(sbuf), _ptr(rbuf), rbuf.numel(), _stream_ptr(stream)))

    def destroy(self):
        if self.h:
            lib.hyd_grid_destroy(self.h)
            self.h = None


_TORCH = {np.dtype(np.float64): torch.float64, np.dtype(np.float32): torch.float32, np.dtype(np.float16): torch.float16}
BF16 = "bfloat16"   # dtype tag for bf16 storage (numpy has no such type: host-side arrays of a bf16 matrix are float32)


class DistMatrix:
    """El::DistMatrix<T,U,V,ELEMENT,Device::GPU>: metadata + this rank's local column-major block."""

    def __init__(self, grid, height, width, colDist=MC, rowDist=MR, dtype=np.float64, colAlign=0, rowAlign=0,
                 device="cuda", ldim=None, pinned=False):
        self.grid, self.m, self.n = grid, int(height), int(width)
        self.cd, self.rd, self.ca, self.ra = colDist, rowDist, int(colAlign), int(rowAlign)
        self.bf16 = isinstance(dtype, str) and dtype == BF16
        self.dtype = np.dtype(np.float32) if self.bf16 else np.dtype(dtype)
        self.tdtype = torch.bfloat16 if self.bf16 else _TORCH[self.dtype]
        self.es = 2 if self.bf16 else self.dtype.itemsize
        meta = hyd_dm(None, 1, self.m, self.n, colDist, rowDist, self.ca, self.ra, 0, 0)
        cs, rs, lh, lw = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
        check(lib.hyd_local_shape(grid.h, C.byref(meta), C.byref(cs), C.byref(rs), C.byref(lh), C.byref(lw)))
        self.colShift, self.rowShift, self.lh, self.lw = cs.value, rs.value, lh.value, lw.value
        q = 16 // self.es
        self.ld = int(ldim) if ldim is not None else max(q, -(-self.lh // q) * q)
        if self.ld < max(self.lh, 1):
            raise HydrogenError("DistMatrix: ldim smaller than the local height")
        # torch (lw, ld) row-major == column-major (ld, lw)
        kw = dict(dtype=self.tdtype, device=device)
        if device == "cpu" and pinned:
            kw["pin_memory"] = True
        self.store = torch.zeros((max(self.lw, 1), self.ld), **kw)

    # -- reference accessors
    def Height(self): return self.m
    def Width(self): return self.n
    def LocalHeight(self): return self.lh
    def LocalWidth(self): return self.lw
    def LDim(self): return self.ld
    def ColAlign(self): return self.ca
    def RowAlign(self): return self.ra
    def ColShift(self): return self.colShift
    def RowShift(self): return self.rowShift
    def ColStride(self): return self.grid.Stride(self.cd)
    def RowStride(self): return self.grid.Stride(self.rd)
    def Grid(self): return self.grid

    def desc(self):
        return hyd_dm(C.c_void_p(self.store.data_ptr()), self.ld, self.m, self.n, self.cd, self.rd, self.ca, self.ra, 0, 0)

    def Matrix(self):
        """The local block as an (lh, lw) view (column-major)."""
        return self.store[: self.lw, : self.lh].t()

    def set_from_global(self, G):
        """Fill this rank's block from a full (m, n) host array: local (iLoc, jLoc) = G[colShift + iLoc*colStride, ...]."""
        blk = np.ascontiguousarray(G[self.colShift::self.ColStride(), self.rowShift::self.RowStride()].T, dtype=self.dtype)
        assert blk.shape == (self.lw, self.lh), (blk.shape, self.lw, self.lh)
        if self.lh and self.lw:
            self.store[: self.lw, : self.lh].copy_(torch.from_numpy(blk))
        return self

    def local_numpy(self):
        blk = self.store[: self.lw, : self.lh].t().cpu()
        return (blk.float() if self.bf16 else blk).numpy()

    def scatter_into_global(self, G):
        if self.lh and self.lw:
            G[self.colShift::self.ColStride(), self.rowShift::self.RowStride()] = self.local_numpy()


def _sfx(dtype):
    return "f64" if np.dtype(dtype) == np.float64 else "f32"


def _scalar(dtype, v):
    return C.c_double(v) if np.dtype(dtype) == np.float64 else C.c_float(v)


def _redist_fn(A, B):
    if A.es != B.es or A.bf16 != B.bf16:
        raise HydrogenError("Copy / Transpose: element types differ")
    return lib.hyd_redist_b16 if A.es == 2 else getattr(lib, "hyd_redist_" + _sfx(A.dtype))


def Copy(A, B, stream=None):
    """B = A between any two supported distributions (DistMatrix::operator=)."""
    a, b = A.desc(), B.desc()
    check(_redist_fn(A, B)(A.grid.h, C.byref(a), C.byref(b), 0, _stream_ptr(stream)))


def Transpose(A, B, stream=None):
    """B = A^T (El::Transpose(A, B), Transpose.hpp:318-356)."""
    a, b = A.desc(), B.desc()
    check(_redist_fn(A, B)(A.grid.h, C.byref(a), C.byref(b), 1, _stream_ptr(stream)))


def AxpyContract(alpha, A, B, stream=None):
    a, b = A.desc(), B.desc()
    check(getattr(lib, "hyd_axpy_contract_" + _sfx(A.dtype))(A.grid.h, _scalar(A.dtype, alpha), C.byref(a), C.byref(b), 0,
                                                             _stream_ptr(stream)))


def TransposeAxpyContract(alpha, A, B, stream=None):
    a, b = A.desc(), B.desc()
    check(getattr(lib, "hyd_axpy_contract_" + _sfx(A.dtype))(A.grid.h, _scalar(A.dtype, alpha), C.byref(a), C.byref(b), 1,
                                                             _stream_ptr(stream)))


def Axpy(alpha, X, Y, stream=None):
    """Y += alpha X between any two distributions (El::Axpy on ElementalMatrix, Axpy.hpp:161-187)."""
    x, y = X.desc(), Y.desc()
    check(getattr(lib, "hyd_axpy_dist_" + _sfx(Y.dtype))(Y.grid.h, _scalar(Y.dtype, alpha), C.byref(x), C.byref(y), 0,
                                                         _stream_ptr(stream)))


def TransposeAxpy(alpha, X, Y, stream=None):
    """Y += alpha X^T (El::TransposeAxpy, TransposeAxpy.hpp)."""
    x, y = X.desc(), Y.desc()
    check(getattr(lib, "hyd_axpy_dist_" + _sfx(Y.dtype))(Y.grid.h, _scalar(Y.dtype, alpha), C.byref(x), C.byref(y), 1,
                                                         _stream_ptr(stream)))


class MultiSync:
    """MakeMultiSync(master, others...) (include/hydrogen/SyncInfo.hpp:136-186): on entry the master stream waits for the
    work enqueued so far on the other streams, on exit the others wait for the master.  No host synchronisation."""

    def __init__(self, master, *others):
        self.m = _stream_ptr(master)
        self.n = len(others)
        self.o = (C.c_void_p * max(1, self.n))(*[_stream_ptr(o).value for o in others])

    def __enter__(self):
        check(lib.hyd_multisync_begin(self.m, self.o, self.n))
        return self

    def __exit__(self, *exc):
        check(lib.hyd_multisync_end(self.m, self.o, self.n))
        return False


def _same_dist(A, B):
    return (A.cd, A.rd, A.ca, A.ra) == (B.cd, B.rd, B.ca, B.ra)


def Hadamard(A, B, Cm, stream=None):
    """C(i,j) = A(i,j) B(i,j) on identically distributed and aligned matrices (Hadamard.hpp:109-135)."""
    if (A.m, A.n) != (B.m, B.n):
        raise HydrogenError("Hadamard product requires equal dimensions")
    if (A.cd, A.rd) != (B.cd, B.rd) or (B.cd, B.rd) != (Cm.cd, Cm.rd):
        raise HydrogenError("A, B, and C must share the same distribution")
    if (A.ca, A.ra) != (B.ca, B.ra):
        raise HydrogenError("A and B must be aligned")
    if not _same_dist(A, Cm) or (Cm.m, Cm.n) != (A.m, A.n):
        raise HydrogenError("Hadamard: C must already be aligned with A and have its size (C.AlignWith / Resize)")
    check(getattr(lib, "hyd_hadamard_" + _sfx(Cm.dtype))(A.lh, A.lw, _ptr(A.store), 1, A.ld, _ptr(B.store), 1, B.ld,
                                                        _ptr(Cm.store), 1, Cm.ld, _stream_ptr(stream)))


def Fill(A, alpha, stream=None):
    """Every entry := alpha (El::Fill on the local block, Fill.hpp)."""
    check(getattr(lib, "hyd_fill_" + _sfx(A.dtype))(A.lh, A.lw, _scalar(A.dtype, alpha), _ptr(A.store), A.ld, _stream_ptr(stream)))


def Zero(A, stream=None):
    Fill(A, 0.0, stream)


def Scale(alpha, A, stream=None):
    """A *= alpha (El::Scale on the local block, Scale.hpp)."""
    check(getattr(lib, "hyd_scale2d_" + _sfx(A.dtype))(A.lh, A.lw, _scalar(A.dtype, alpha), _ptr(A.store), A.ld, _stream_ptr(stream)))


def Gemm(orientA, orientB, alpha, A, B, beta, Cm, alg=GEMM_DEFAULT, stream=None, blocksize=None):
    """C := alpha op(A) op(B) + beta C  (El::Gemm on AbstractDistMatrix)."""
    a, b, c = A.desc(), B.desc(), Cm.desc()
    nb = Blocksize() if blocksize is None else int(blocksize)
    if Cm.es == 2:   # fp16 / bf16 storage, fp32 accumulation and scalars
        if (A.es, A.bf16, B.es, B.bf16) != (2, Cm.bf16, 2, Cm.bf16):
            raise HydrogenError("Gemm: A, B and C must share the half-precision storage type")
        fn = lib.hyd_gemm_bf16 if Cm.bf16 else lib.hyd_gemm_f16
        check(fn(Cm.grid.h, orientA, orientB, alg, C.c_float(alpha), C.byref(a), C.byref(b), C.c_float(beta), C.byref(c), nb,
                 _stream_ptr(stream)))
        return
    host = Cm.store.device.type == "cpu"
    fn = getattr(lib, ("hyd_gemm_host_" if host else "hyd_gemm_") + _sfx(Cm.dtype))
    check(fn(Cm.grid.h, orientA, orientB, alg, _scalar(Cm.dtype, alpha), C.byref(a), C.byref(b), _scalar(Cm.dtype, beta),
             C.byref(c), nb, _stream_ptr(stream)))


def Trrk(uplo, orientA, orientB, alpha, A, B, beta, Cm, stream=None, blocksize=None):
    """Triangle `uplo` of C := alpha op(A) op(B) + beta C  (El::Trrk on AbstractDistMatrix, Trrk.cpp:98-115)."""
    a, b, c = A.desc(), B.desc(), Cm.desc()
    nb = Blocksize() if blocksize is None else int(blocksize)
    check(getattr(lib, "hyd_trrk_" + _sfx(Cm.dtype))(Cm.grid.h, uplo, orientA, orientB, _scalar(Cm.dtype, alpha), C.byref(a),
                                                    C.byref(b), _scalar(Cm.dtype, beta), C.byref(c), nb, _stream_ptr(stream)))


def Syrk(uplo, orientation, alpha, A, beta, Cm, conjugate=False, stream=None, blocksize=None):
    """Triangle `uplo` of C := alpha A A^T + beta C (NORMAL) or alpha A^T A + beta C  (El::Syrk, Syrk.cpp:72-90)."""
    a, c = A.desc(), Cm.desc()
    nb = Blocksize() if blocksize is None else int(blocksize)
    check(getattr(lib, "hyd_syrk_" + _sfx(Cm.dtype))(Cm.grid.h, uplo, orientation, _scalar(Cm.dtype, alpha), C.byref(a),
                                                    _scalar(Cm.dtype, beta), C.byref(c), nb, _stream_ptr(stream)))


def Herk(uplo, orientation, alpha, A, beta, Cm, stream=None, blocksize=None):
    """El::Herk (Herk.cpp): Syrk with conjugate = true; the same operation for real T."""
    Syrk(uplo, orientation, alpha, A, beta, Cm, True, stream, blocksize)


def ScaleTrapezoid(alpha, uplo, A, offset=0, stream=None):
    """A *= alpha inside the trapezoid (ScaleTrapezoid.hpp:47-92)."""
    a = A.desc()
    check(getattr(lib, "hyd_scale_trapezoid_dist_" + _sfx(A.dtype))(A.grid.h, _scalar(A.dtype, alpha), uplo, C.byref(a), offset,
                                                                   _stream_ptr(stream)))


def AxpyTrapezoid(uplo, alpha, X, Y, offset=0, stream=None):
    """Y += alpha X inside the trapezoid; X and Y share distribution and alignments (AxpyTrapezoid.hpp:54-100)."""
    x, y = X.desc(), Y.desc()
    check(getattr(lib, "hyd_axpy_trapezoid_dist_" + _sfx(Y.dtype))(Y.grid.h, uplo, _scalar(Y.dtype, alpha), C.byref(x), C.byref(y),
                                                                  offset, _stream_ptr(stream)))


def MakeTrapezoidal(uplo, A, offset=0, stream=None):
    """A := 0 outside the trapezoid (MakeTrapezoidal.hpp:48-90)."""
    a = A.desc()
    check(getattr(lib, "hyd_make_trapezoidal_dist_" + _sfx(A.dtype))(A.grid.h, uplo, C.byref(a), offset, _stream_ptr(stream)))


def Symm(side, uplo, alpha, A, B, beta, Cm, conjugate=False, stream=None, blocksize=None):
    """C := alpha A B + beta C (LEFT) / alpha B A + beta C (RIGHT), A symmetric, only its `uplo` triangle read (Symm.cpp:56-80)."""
    a, b, c = A.desc(), B.desc(), Cm.desc()
    nb = Blocksize() if blocksize is None else int(blocksize)
    check(getattr(lib, "hyd_symm_" + _sfx(Cm.dtype))(Cm.grid.h, side, uplo, _scalar(Cm.dtype, alpha), C.byref(a), C.byref(b),
                                                    _scalar(Cm.dtype, beta), C.byref(c), nb, _stream_ptr(stream)))


def Hemm(side, uplo, alpha, A, B, beta, Cm, stream=None, blocksize=None):
    """El::Hemm (Hemm.cpp): Symm with conjugate = true; the same operation for real T."""
    Symm(side, uplo, alpha, A, B, beta, Cm, True, stream, blocksize)


def MakeSymmetric(uplo, A, conjugate=False, stream=None):
    """A := symmetric completion of its `uplo` triangle (MakeSymmetric.hpp:52-68)."""
    a = A.desc()
    check(getattr(lib, "hyd_make_symmetric_dist_" + _sfx(A.dtype))(A.grid.h, uplo, C.byref(a), _stream_ptr(stream)))


def lgemm(tA, tB, m, n, k, alpha, A, lda, B, ldb, beta, Cm, ldc, dtype=np.float64, stream=None):
    """gpu_blas::Gemm on raw column-major device buffers held by torch tensors."""
    fn = getattr(lib, "hyd_lgemm_" + _sfx(dtype))
    check(fn(tA, tB, m, n, k, _scalar(dtype, alpha), C.c_void_p(A.data_ptr()), lda, C.c_void_p(B.data_ptr()), ldb,
             _scalar(dtype, beta), C.c_void_p(Cm.data_ptr()), ldc, _stream_ptr(stream)))


def run_ranks(P, fn, height=None, order=COLUMN_MAJOR, device=0):
    """Run fn(grid) on P rank-threads sharing ONE GPU (hyd_local_group); returns the per-rank results.

    Each rank-thread works on its own CUDA stream.  This is how the multi-rank engine is
    parity-tested on a single-GPU box; the exchange is stream-ordered device copies."""
    grp = LocalGroup(P)
    out, err = [None] * P, [None] * P

    def body(rank):
        try:
            torch.cuda.set_device(device)
            st = torch.cuda.Stream()
            with torch.cuda.stream(st):
                w = World.local(grp, rank)
                g = Grid(w, height, order)
                try:
                    out[rank] = fn(g)
                    st.synchronize()
                finally:
                    w.barrier()
                    g.destroy()
                    w.destroy()
        except BaseException as e:  # noqa: BLE001
            err[rank] = e

    th = [threading.Thread(target=body, args=(r,)) for r in range(P)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    grp.destroy()
    for e in err:
        if e is not None:
            raise e
    return out

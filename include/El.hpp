// El.hpp -- header-only C++ shim that keeps the reference's names and signatures for the
// distributed-GEMM path and forwards every operation to the C ABI of libhydrogen_b200.so
// (include/hydrogen_b200.h).  Host code written against Hydrogen's
//     El::Grid, El::DistMatrix<T,U,V,ELEMENT,Device::GPU>, El::Gemm, El::Copy / operator=,
//     El::Transpose, El::AxpyContract, El::SetBlocksize / El::Blocksize
// compiles against this header for that path (reference declarations:
// include/El/core/Grid.hpp:15-146, include/El/core/DistMatrix/**, include/El/blas_like/level3.hpp:22-84,
// include/El/blas_like/level1/{Copy,Transpose,AxpyContract}.hpp, include/El/core/environment/decl.hpp).
//
// Errors: non-zero C-ABI status is rethrown as std::runtime_error / std::logic_error, the
// reference's convention (LogicError / RuntimeError, include/El/core/environment/decl.hpp:196-198).
// Memory: DistMatrix owns its local device block (cudaMalloc) unless Attach()ed.
// Streams: every call enqueues on the matrix' SyncInfo stream (default: the null stream),
// as Matrix<T,Device::GPU> does in the reference (include/El/core/Matrix/decl.hpp:470,510-518).
#ifndef EL_B200_SHIM_HPP
#define EL_B200_SHIM_HPP

#include <cuda_runtime.h>
#include <cstdint>
#include <chrono>
#include <initializer_list>
#include <iostream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "hydrogen_b200.h"

namespace El {

typedef long long Int;  // EL_USE_64BIT_INTS flavour: 32768^2 locals overflow 32-bit offsets

enum Dist { MC = HYD_MC, MD = HYD_MD, MR = HYD_MR, VC = HYD_VC, VR = HYD_VR, STAR = HYD_STAR, CIRC = HYD_CIRC };
enum DistWrap { ELEMENT, BLOCK };
enum class Device { CPU, GPU };
enum Orientation { NORMAL = HYD_NORMAL, TRANSPOSE = HYD_TRANSPOSE, ADJOINT = HYD_ADJOINT };
enum GridOrder { ROW_MAJOR = 0, COLUMN_MAJOR = 1 };
enum UpperOrLower { LOWER = 0, UPPER = 1 };   // include/El/core/types.hpp:489-493
enum LeftOrRight { LEFT = 0, RIGHT = 1 };     // types.hpp:417-424
enum GemmAlgorithm {
    GEMM_DEFAULT = HYD_GEMM_DEFAULT, GEMM_SUMMA_A = HYD_GEMM_SUMMA_A, GEMM_SUMMA_B = HYD_GEMM_SUMMA_B,
    GEMM_SUMMA_C = HYD_GEMM_SUMMA_C, GEMM_SUMMA_DOT = HYD_GEMM_SUMMA_DOT, GEMM_CANNON = HYD_GEMM_CANNON
};

inline void LogicError(const std::string& s) { throw std::logic_error(s); }
inline void RuntimeError(const std::string& s) { throw std::runtime_error(s); }
namespace detail {
inline void check(int rc) {
    if (rc == 0) return;
    const std::string msg = hyd_last_error();
    // argument / conformality problems are LogicErrors in the reference, device failures RuntimeErrors
    if (msg.find("CUDA") != std::string::npos || msg.find("NCCL") != std::string::npos) throw std::runtime_error(msg);
    throw std::logic_error(msg);
}
inline void check_cuda(cudaError_t e, const char* what) {
    if (e != cudaSuccess) throw std::runtime_error(std::string(what) + ": " + cudaGetErrorString(e));
}
template <typename T> struct Sfx;
}  // namespace detail

// ---- algorithmic blocksize stack (src/core/environment.cpp:292-330; default 128) ----
namespace detail { inline std::vector<Int>& blocksizes() { static std::vector<Int> s{128}; return s; } }
inline Int Blocksize() { return detail::blocksizes().back(); }
inline void SetBlocksize(Int nb) { detail::blocksizes().back() = nb; }
inline void PushBlocksizeStack(Int nb) { detail::blocksizes().push_back(nb); }
inline void PopBlocksizeStack() { detail::blocksizes().pop_back(); }

// ---- SyncInfo<Device::GPU> (include/hydrogen/SyncInfo.hpp:51-60): the stream work is enqueued on ----
struct SyncInfoGPU { cudaStream_t stream_ = nullptr; };

// ---- MultiSync (include/hydrogen/SyncInfo.hpp:136-186) ----
// MakeMultiSync(master, others...): on construction the master stream -- the OUTPUT matrix's -- waits for everything already
// enqueued on the other (input) streams; on destruction the others wait for the master, so an input may be overwritten on
// its own stream right after the call.  No host synchronisation.  Every distributed operation of this header brackets its
// C-ABI call with one (as the reference does at Gemm.cpp:117-119, Copy/RowAllGather.hpp:25-28, ...), so matrices produced
// on different streams can be combined without any explicit fence by the caller.
class MultiSync {
public:
    MultiSync(cudaStream_t master, std::initializer_list<cudaStream_t> others) : master_(master), n_(0) {
        for (cudaStream_t o : others) if (n_ < 8) others_[n_++] = o;
        if (hyd_multisync_begin(master_, others_, n_)) throw std::runtime_error(hyd_last_error());
    }
    ~MultiSync() { hyd_multisync_end(master_, others_, n_); }
    MultiSync(const MultiSync&) = delete;
    MultiSync& operator=(const MultiSync&) = delete;
    operator SyncInfoGPU() const { SyncInfoGPU si; si.stream_ = master_; return si; }   // usable wherever the master SyncInfo is
private:
    cudaStream_t master_;
    hyd_stream_t others_[8];
    int n_;
};

// ---- communicator: one process per GPU over NCCL (or P rank-threads on one GPU for tests) ----
namespace mpi {
class Comm {
public:
    Comm() = default;
    // every rank passes the same 128-byte id (hyd_nccl_unique_id on rank 0, broadcast by the launcher)
    static Comm NCCL(const void* id128, int rank, int size) { Comm c; detail::check(hyd_world_create_nccl(id128, rank, size, &c.w_)); return c; }
    static Comm Local(hyd_local_group* grp, int rank) { Comm c; detail::check(hyd_world_create_local(grp, rank, &c.w_)); return c; }
    static Comm Self() { Comm c; detail::check(hyd_local_group_create(1, &c.grp_)); detail::check(hyd_world_create_local(c.grp_, 0, &c.w_)); return c; }
    // the world of a one-process-per-GPU launch (mpi::COMM_WORLD to a Hydrogen program): rank / size / device from the
    // launcher's environment, no MPI library involved (hyd_world_create_env)
    static Comm World() { Comm c; detail::check(hyd_world_create_env(&c.w_)); return c; }
    hyd_world* handle() const { return w_; }
    int Size() const { return hyd_world_size(w_); }
    int Rank() const { return hyd_world_rank(w_); }
    void Barrier() const { detail::check(hyd_world_barrier(w_)); }
    void Free() { if (w_) hyd_world_destroy(w_); if (grp_) hyd_local_group_destroy(grp_); w_ = nullptr; grp_ = nullptr; }
private:
    hyd_world* w_ = nullptr;
    hyd_local_group* grp_ = nullptr;
};
// the free functions Hydrogen client code uses on a communicator (include/El/core/imports/mpi.hpp:204-217,263)
inline Comm NewWorldComm() { return Comm::World(); }
inline int Rank(const Comm& comm) { return comm.Rank(); }
inline int Size(const Comm& comm) { return comm.Size(); }
inline void Barrier(const Comm& comm) { comm.Barrier(); }
}  // namespace mpi

// ---- El::Initialize / Finalize / Environment (src/core/environment.cpp:217-300, include/El/core/environment/decl.hpp) ----
// The reference initialises MPI, the GPU manager (device = local rank) and cuBLAS here.  None of that exists in this build:
// the device is chosen when the world communicator is created (mpi::NewWorldComm), there is no global handle.  Kept so that
// client code written against the reference compiles unchanged.
inline void Initialize() {}
inline void Initialize(int& /*argc*/, char**& /*argv*/) {}
inline void Finalize() {}
inline bool Initialized() { return true; }
class Environment {
public:
    Environment() { Initialize(); }
    Environment(int& argc, char**& argv) { Initialize(argc, argv); }
    ~Environment() { Finalize(); }
};
// El::Output(args...) (include/El/core/environment/impl.hpp): one line on stdout
namespace detail {
inline void out_all(std::ostream&) {}
template <typename A, typename... R> void out_all(std::ostream& os, const A& a, const R&... r) { os << a; out_all(os, r...); }
}  // namespace detail
template <typename... Args> void Output(const Args&... args) { std::ostringstream os; detail::out_all(os, args...); os << "\n"; std::cout << os.str() << std::flush; }
// El::Timer (include/El/core/Timer.hpp): Start() / Stop() returns the seconds since Start()
class Timer {
public:
    void Start() { t0_ = std::chrono::steady_clock::now(); }
    double Stop() { return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0_).count(); }
private:
    std::chrono::steady_clock::time_point t0_;
};

// ---- El::Grid (Grid.hpp:15-146; Grid.cpp:66-216) ----
class Grid {
public:
    Grid(const mpi::Comm& comm, int height, GridOrder order = COLUMN_MAJOR) : world_(comm) { detail::check(hyd_grid_create(comm.handle(), height, order, &g_)); }
    // Grid(comm, order): height = Grid::DefaultHeight(size) (Grid.cpp:58-64,70-72)
    explicit Grid(const mpi::Comm& comm, GridOrder order = COLUMN_MAJOR) : Grid(comm, hyd_grid_default_height(comm.Size()), order) {}
    static int DefaultHeight(int gridSize) { return hyd_grid_default_height(gridSize); }
    ~Grid() { if (g_) hyd_grid_destroy(g_); }
    Grid(const Grid&) = delete;
    Grid& operator=(const Grid&) = delete;
    int Height() const { return hyd_grid_height(g_); }
    int Width() const { return hyd_grid_width(g_); }
    int Size() const { return hyd_grid_size(g_); }
    int Rank() const { return hyd_grid_rank(g_); }
    int Row() const { return hyd_grid_row(g_); }
    int Col() const { return hyd_grid_col(g_); }
    int MCRank() const { return Row(); }
    int MRRank() const { return Col(); }
    int VCRank() const { return hyd_grid_vc_rank(g_); }
    int VRRank() const { return hyd_grid_vr_rank(g_); }
    int MCSize() const { return Height(); }
    int MRSize() const { return Width(); }
    int VCSize() const { return Size(); }
    int VRSize() const { return Size(); }
    hyd_grid* handle() const { return g_; }
    // sub-communicators (Grid.hpp:31-54): a (grid, kind) pair -- every member is one NVSwitch hop away, there is no
    // communicator object to create.  ColComm == MCComm (my grid column), RowComm == MRComm (my grid row).
    struct SubComm {
        hyd_grid* g; int kind;
        int Size() const { return hyd_comm_size(g, kind); }
        int Rank() const { return hyd_comm_rank(g, kind); }
    };
    SubComm MCComm() const { return SubComm{g_, HYD_COMM_MC}; }
    SubComm MRComm() const { return SubComm{g_, HYD_COMM_MR}; }
    SubComm VCComm() const { return SubComm{g_, HYD_COMM_VC}; }
    SubComm VRComm() const { return SubComm{g_, HYD_COMM_VR}; }
    SubComm ColComm() const { return MCComm(); }
    SubComm RowComm() const { return MRComm(); }
    // the communicator the grid was built on (Grid.hpp:24-30): Comm() / OwningComm() / ViewingComm() coincide here
    const mpi::Comm& Comm() const { return world_; }
    const mpi::Comm& OwningComm() const { return world_; }
    const mpi::Comm& ViewingComm() const { return world_; }
private:
    hyd_grid* g_ = nullptr;
    mpi::Comm world_;   // a handle copy; the caller keeps ownership (mpi::Comm::Free)
};

// ---- typed collectives on device buffers (src/core/imports/mpi/*.hpp, Device::GPU overloads): counts in elements ----
namespace mpi {
template <typename T>
void AllGather(const T* sbuf, Int sc, T* rbuf, Int /*rc*/, const Grid::SubComm& comm, const SyncInfoGPU& si) {
    detail::check(hyd_allgather(comm.g, comm.kind, sbuf, sc * (Int)sizeof(T), rbuf, si.stream_));
}
template <typename T>
void AllToAll(const T* sbuf, Int sc, T* rbuf, Int /*rc*/, const Grid::SubComm& comm, const SyncInfoGPU& si) {
    detail::check(hyd_alltoall(comm.g, comm.kind, sbuf, sc * (Int)sizeof(T), rbuf, si.stream_));
}
template <typename T>
void SendRecv(const T* sbuf, Int sc, int to, T* rbuf, Int rc, int from, const Grid::SubComm& comm, const SyncInfoGPU& si) {
    detail::check(hyd_sendrecv(comm.g, comm.kind, sbuf, sc * (Int)sizeof(T), to, rbuf, rc * (Int)sizeof(T), from, si.stream_));
}
template <typename T>
void Broadcast(T* buf, Int count, int root, const Grid::SubComm& comm, const SyncInfoGPU& si) {
    detail::check(hyd_bcast(comm.g, comm.kind, buf, count * (Int)sizeof(T), root, si.stream_));
}
inline void ReduceScatter(const double* sbuf, double* rbuf, Int rc, const Grid::SubComm& comm, const SyncInfoGPU& si) {
    detail::check(hyd_reducescatter_f64(comm.g, comm.kind, sbuf, rbuf, rc, si.stream_));
}
inline void ReduceScatter(const float* sbuf, float* rbuf, Int rc, const Grid::SubComm& comm, const SyncInfoGPU& si) {
    detail::check(hyd_reducescatter_f32(comm.g, comm.kind, sbuf, rbuf, rc, si.stream_));
}
}  // namespace mpi

namespace detail {
template <> struct Sfx<double> {
    static int redist(hyd_grid* g, const hyd_dm* a, const hyd_dm* b, int t, void* s) { return hyd_redist_f64(g, a, b, t, s); }
    static int contract(hyd_grid* g, double al, const hyd_dm* a, const hyd_dm* b, int t, void* s) { return hyd_axpy_contract_f64(g, al, a, b, t, s); }
    static int gemm(hyd_grid* g, int oA, int oB, int alg, double al, const hyd_dm* A, const hyd_dm* B, double be, const hyd_dm* C, int64_t nb, void* s) { return hyd_gemm_f64(g, oA, oB, alg, al, A, B, be, C, nb, s); }
    static int gemm_host(hyd_grid* g, int oA, int oB, int alg, double al, const hyd_dm* A, const hyd_dm* B, double be, const hyd_dm* C, int64_t nb, void* s) { return hyd_gemm_host_f64(g, oA, oB, alg, al, A, B, be, C, nb, s); }
    static int lgemm(int tA, int tB, int64_t m, int64_t n, int64_t k, double al, const double* A, int64_t lda, const double* B, int64_t ldb, double be, double* C, int64_t ldc, void* s) { return hyd_lgemm_f64(tA, tB, m, n, k, al, A, lda, B, ldb, be, C, ldc, s); }
    static int transpose(int64_t m, int64_t n, const double* A, int64_t lda, double* B, int64_t ldb, void* s) { return hyd_transpose_f64(m, n, A, lda, B, ldb, s); }
    static int copy2d(int64_t m, int64_t n, const double* A, int64_t lda, double* B, int64_t ldb, void* s) { return hyd_copy2d_f64(m, n, A, lda, B, ldb, s); }
    static int axpy2d(int64_t m, int64_t n, double al, const double* A, int64_t lda, double* B, int64_t ldb, void* s) { return hyd_axpy2d_f64(m, n, al, A, lda, B, ldb, s); }
    static int scale2d(int64_t m, int64_t n, double al, double* A, int64_t lda, void* s) { return hyd_scale2d_f64(m, n, al, A, lda, s); }
    static int fill(int64_t m, int64_t n, double al, double* A, int64_t lda, void* s) { return hyd_fill_f64(m, n, al, A, lda, s); }
    static int hadamard(int64_t m, int64_t n, const double* X, int64_t csX, int64_t rsX, const double* Y, int64_t csY, int64_t rsY, double* Z, int64_t csZ, int64_t rsZ, void* s) { return hyd_hadamard_f64(m, n, X, csX, rsX, Y, csY, rsY, Z, csZ, rsZ, s); }
    static int axpy_trans(int tA, int64_t m, int64_t n, double al, const double* A, int64_t lda, double* B, int64_t ldb, void* s) { return hyd_axpy_trans_f64(tA, m, n, al, A, lda, B, ldb, s); }
    static int copy1d(int64_t n, const double* X, int64_t incx, double* Y, int64_t incy, void* s) { return hyd_copy1d_f64(n, X, incx, Y, incy, s); }
    static int axpy1d(int64_t n, double al, const double* X, int64_t incx, double* Y, int64_t incy, void* s) { return hyd_axpy1d_f64(n, al, X, incx, Y, incy, s); }
    static int scale1d(int64_t n, double al, double* X, int64_t incx, void* s) { return hyd_scale1d_f64(n, al, X, incx, s); }
    static int axpy_dist(hyd_grid* g, double al, const hyd_dm* X, const hyd_dm* Y, int t, void* s) { return hyd_axpy_dist_f64(g, al, X, Y, t, s); }
    static int trrk(hyd_grid* g, int ul, int oA, int oB, double al, const hyd_dm* A, const hyd_dm* B, double be, const hyd_dm* C, int64_t nb, void* s) { return hyd_trrk_f64(g, ul, oA, oB, al, A, B, be, C, nb, s); }
    static int syrk(hyd_grid* g, int ul, int o, double al, const hyd_dm* A, double be, const hyd_dm* C, int64_t nb, void* s) { return hyd_syrk_f64(g, ul, o, al, A, be, C, nb, s); }
    static int scale_trap(hyd_grid* g, double al, int ul, const hyd_dm* A, int64_t off, void* s) { return hyd_scale_trapezoid_dist_f64(g, al, ul, A, off, s); }
    static int axpy_trap(hyd_grid* g, int ul, double al, const hyd_dm* X, const hyd_dm* Y, int64_t off, void* s) { return hyd_axpy_trapezoid_dist_f64(g, ul, al, X, Y, off, s); }
    static int make_trap(hyd_grid* g, int ul, const hyd_dm* A, int64_t off, void* s) { return hyd_make_trapezoidal_dist_f64(g, ul, A, off, s); }
    static int make_symm(hyd_grid* g, int ul, const hyd_dm* A, void* s) { return hyd_make_symmetric_dist_f64(g, ul, A, s); }
    static int symm(hyd_grid* g, int sd, int ul, double al, const hyd_dm* A, const hyd_dm* B, double be, const hyd_dm* C, int64_t nb, void* s) { return hyd_symm_f64(g, sd, ul, al, A, B, be, C, nb, s); }
    static int interleave(int64_t h, int64_t w, const double* A, int64_t csA, int64_t rsA, double* B, int64_t csB, int64_t rsB, void* s) { return hyd_interleave_f64(h, w, A, csA, rsA, B, csB, rsB, s); }
    static int interleave_update(double al, int64_t h, int64_t w, const double* A, int64_t csA, int64_t rsA, double* B, int64_t csB, int64_t rsB, void* s) { return hyd_interleave_update_f64(al, h, w, A, csA, rsA, B, csB, rsB, s); }
    static int row_pack(int64_t h, int64_t w, int64_t al, int64_t st, const double* A, int64_t ld, double* P, int64_t ps, void* s) { return hyd_row_strided_pack_f64(h, w, al, st, A, ld, P, ps, s); }
    static int row_unpack(int64_t h, int64_t w, int64_t al, int64_t st, const double* P, int64_t ps, double* B, int64_t ld, void* s) { return hyd_row_strided_unpack_f64(h, w, al, st, P, ps, B, ld, s); }
    static int col_pack(int64_t h, int64_t w, int64_t al, int64_t st, const double* A, int64_t ld, double* P, int64_t ps, void* s) { return hyd_col_strided_pack_f64(h, w, al, st, A, ld, P, ps, s); }
    static int col_unpack(int64_t h, int64_t w, int64_t al, int64_t st, const double* P, int64_t ps, double* B, int64_t ld, void* s) { return hyd_col_strided_unpack_f64(h, w, al, st, P, ps, B, ld, s); }
    static int prow_pack(int64_t h, int64_t w, int64_t al, int64_t st, int64_t un, int64_t pt, int64_t rp, int64_t sh, const double* A, int64_t ld, double* P, int64_t ps, void* s) { return hyd_partial_row_strided_pack_f64(h, w, al, st, un, pt, rp, sh, A, ld, P, ps, s); }
    static int prow_unpack(int64_t h, int64_t w, int64_t al, int64_t st, int64_t un, int64_t pt, int64_t rp, int64_t sh, const double* P, int64_t ps, double* B, int64_t ld, void* s) { return hyd_partial_row_strided_unpack_f64(h, w, al, st, un, pt, rp, sh, P, ps, B, ld, s); }
    static int pcol_pack(int64_t h, int64_t w, int64_t al, int64_t st, int64_t un, int64_t pt, int64_t rp, int64_t sh, const double* A, int64_t ld, double* P, int64_t ps, void* s) { return hyd_partial_col_strided_pack_f64(h, w, al, st, un, pt, rp, sh, A, ld, P, ps, s); }
    static int pcol_unpack(int64_t h, int64_t w, int64_t al, int64_t st, int64_t un, int64_t pt, int64_t rp, int64_t sh, const double* P, int64_t ps, double* B, int64_t ld, void* s) { return hyd_partial_col_strided_unpack_f64(h, w, al, st, un, pt, rp, sh, P, ps, B, ld, s); }
    static int pack2d(int64_t h, int64_t w, int64_t ca, int64_t cs, int64_t ra, int64_t rs, const double* A, int64_t ld, double* P, int64_t ps, void* s) { return hyd_strided_pack_f64(h, w, ca, cs, ra, rs, A, ld, P, ps, s); }
    static int unpack2d(int64_t h, int64_t w, int64_t ca, int64_t cs, int64_t ra, int64_t rs, const double* P, int64_t ps, double* B, int64_t ld, void* s) { return hyd_strided_unpack_f64(h, w, ca, cs, ra, rs, P, ps, B, ld, s); }
};
template <> struct Sfx<float> {
    static int redist(hyd_grid* g, const hyd_dm* a, const hyd_dm* b, int t, void* s) { return hyd_redist_f32(g, a, b, t, s); }
    static int contract(hyd_grid* g, float al, const hyd_dm* a, const hyd_dm* b, int t, void* s) { return hyd_axpy_contract_f32(g, al, a, b, t, s); }
    static int gemm(hyd_grid* g, int oA, int oB, int alg, float al, const hyd_dm* A, const hyd_dm* B, float be, const hyd_dm* C, int64_t nb, void* s) { return hyd_gemm_f32(g, oA, oB, alg, al, A, B, be, C, nb, s); }
    static int gemm_host(hyd_grid* g, int oA, int oB, int alg, float al, const hyd_dm* A, const hyd_dm* B, float be, const hyd_dm* C, int64_t nb, void* s) { return hyd_gemm_host_f32(g, oA, oB, alg, al, A, B, be, C, nb, s); }
    static int lgemm(int tA, int tB, int64_t m, int64_t n, int64_t k, float al, const float* A, int64_t lda, const float* B, int64_t ldb, float be, float* C, int64_t ldc, void* s) { return hyd_lgemm_f32(tA, tB, m, n, k, al, A, lda, B, ldb, be, C, ldc, s); }
    static int transpose(int64_t m, int64_t n, const float* A, int64_t lda, float* B, int64_t ldb, void* s) { return hyd_transpose_f32(m, n, A, lda, B, ldb, s); }
    static int copy2d(int64_t m, int64_t n, const float* A, int64_t lda, float* B, int64_t ldb, void* s) { return hyd_copy2d_f32(m, n, A, lda, B, ldb, s); }
    static int axpy2d(int64_t m, int64_t n, float al, const float* A, int64_t lda, float* B, int64_t ldb, void* s) { return hyd_axpy2d_f32(m, n, al, A, lda, B, ldb, s); }
    static int scale2d(int64_t m, int64_t n, float al, float* A, int64_t lda, void* s) { return hyd_scale2d_f32(m, n, al, A, lda, s); }
    static int fill(int64_t m, int64_t n, float al, float* A, int64_t lda, void* s) { return hyd_fill_f32(m, n, al, A, lda, s); }
    static int hadamard(int64_t m, int64_t n, const float* X, int64_t csX, int64_t rsX, const float* Y, int64_t csY, int64_t rsY, float* Z, int64_t csZ, int64_t rsZ, void* s) { return hyd_hadamard_f32(m, n, X, csX, rsX, Y, csY, rsY, Z, csZ, rsZ, s); }
    static int axpy_trans(int tA, int64_t m, int64_t n, float al, const float* A, int64_t lda, float* B, int64_t ldb, void* s) { return hyd_axpy_trans_f32(tA, m, n, al, A, lda, B, ldb, s); }
    static int copy1d(int64_t n, const float* X, int64_t incx, float* Y, int64_t incy, void* s) { return hyd_copy1d_f32(n, X, incx, Y, incy, s); }
    static int axpy1d(int64_t n, float al, const float* X, int64_t incx, float* Y, int64_t incy, void* s) { return hyd_axpy1d_f32(n, al, X, incx, Y, incy, s); }
    static int scale1d(int64_t n, float al, float* X, int64_t incx, void* s) { return hyd_scale1d_f32(n, al, X, incx, s); }
    static int axpy_dist(hyd_grid* g, float al, const hyd_dm* X, const hyd_dm* Y, int t, void* s) { return hyd_axpy_dist_f32(g, al, X, Y, t, s); }
    static int trrk(hyd_grid* g, int ul, int oA, int oB, float al, const hyd_dm* A, const hyd_dm* B, float be, const hyd_dm* C, int64_t nb, void* s) { return hyd_trrk_f32(g, ul, oA, oB, al, A, B, be, C, nb, s); }
    static int syrk(hyd_grid* g, int ul, int o, float al, const hyd_dm* A, float be, const hyd_dm* C, int64_t nb, void* s) { return hyd_syrk_f32(g, ul, o, al, A, be, C, nb, s); }
    static int scale_trap(hyd_grid* g, float al, int ul, const hyd_dm* A, int64_t off, void* s) { return hyd_scale_trapezoid_dist_f32(g, al, ul, A, off, s); }
    static int axpy_trap(hyd_grid* g, int ul, float al, const hyd_dm* X, const hyd_dm* Y, int64_t off, void* s) { return hyd_axpy_trapezoid_dist_f32(g, ul, al, X, Y, off, s); }
    static int make_trap(hyd_grid* g, int ul, const hyd_dm* A, int64_t off, void* s) { return hyd_make_trapezoidal_dist_f32(g, ul, A, off, s); }
    static int make_symm(hyd_grid* g, int ul, const hyd_dm* A, void* s) { return hyd_make_symmetric_dist_f32(g, ul, A, s); }
    static int symm(hyd_grid* g, int sd, int ul, float al, const hyd_dm* A, const hyd_dm* B, float be, const hyd_dm* C, int64_t nb, void* s) { return hyd_symm_f32(g, sd, ul, al, A, B, be, C, nb, s); }
    static int interleave(int64_t h, int64_t w, const float* A, int64_t csA, int64_t rsA, float* B, int64_t csB, int64_t rsB, void* s) { return hyd_interleave_f32(h, w, A, csA, rsA, B, csB, rsB, s); }
    static int interleave_update(float al, int64_t h, int64_t w, const float* A, int64_t csA, int64_t rsA, float* B, int64_t csB, int64_t rsB, void* s) { return hyd_interleave_update_f32(al, h, w, A, csA, rsA, B, csB, rsB, s); }
    static int row_pack(int64_t h, int64_t w, int64_t al, int64_t st, const float* A, int64_t ld, float* P, int64_t ps, void* s) { return hyd_row_strided_pack_f32(h, w, al, st, A, ld, P, ps, s); }
    static int row_unpack(int64_t h, int64_t w, int64_t al, int64_t st, const float* P, int64_t ps, float* B, int64_t ld, void* s) { return hyd_row_strided_unpack_f32(h, w, al, st, P, ps, B, ld, s); }
    static int col_pack(int64_t h, int64_t w, int64_t al, int64_t st, const float* A, int64_t ld, float* P, int64_t ps, void* s) { return hyd_col_strided_pack_f32(h, w, al, st, A, ld, P, ps, s); }
    static int col_unpack(int64_t h, int64_t w, int64_t al, int64_t st, const float* P, int64_t ps, float* B, int64_t ld, void* s) { return hyd_col_strided_unpack_f32(h, w, al, st, P, ps, B, ld, s); }
    static int prow_pack(int64_t h, int64_t w, int64_t al, int64_t st, int64_t un, int64_t pt, int64_t rp, int64_t sh, const float* A, int64_t ld, float* P, int64_t ps, void* s) { return hyd_partial_row_strided_pack_f32(h, w, al, st, un, pt, rp, sh, A, ld, P, ps, s); }
    static int prow_unpack(int64_t h, int64_t w, int64_t al, int64_t st, int64_t un, int64_t pt, int64_t rp, int64_t sh, const float* P, int64_t ps, float* B, int64_t ld, void* s) { return hyd_partial_row_strided_unpack_f32(h, w, al, st, un, pt, rp, sh, P, ps, B, ld, s); }
    static int pcol_pack(int64_t h, int64_t w, int64_t al, int64_t st, int64_t un, int64_t pt, int64_t rp, int64_t sh, const float* A, int64_t ld, float* P, int64_t ps, void* s) { return hyd_partial_col_strided_pack_f32(h, w, al, st, un, pt, rp, sh, A, ld, P, ps, s); }
    static int pcol_unpack(int64_t h, int64_t w, int64_t al, int64_t st, int64_t un, int64_t pt, int64_t rp, int64_t sh, const float* P, int64_t ps, float* B, int64_t ld, void* s) { return hyd_partial_col_strided_unpack_f32(h, w, al, st, un, pt, rp, sh, P, ps, B, ld, s); }
    static int pack2d(int64_t h, int64_t w, int64_t ca, int64_t cs, int64_t ra, int64_t rs, const float* A, int64_t ld, float* P, int64_t ps, void* s) { return hyd_strided_pack_f32(h, w, ca, cs, ra, rs, A, ld, P, ps, s); }
    static int unpack2d(int64_t h, int64_t w, int64_t ca, int64_t cs, int64_t ra, int64_t rs, const float* P, int64_t ps, float* B, int64_t ld, void* s) { return hyd_strided_unpack_f32(h, w, ca, cs, ra, rs, P, ps, B, ld, s); }
};
}  // namespace detail

// ---- AbstractDistMatrix<T> / DistMatrix<T,U,V,ELEMENT,Device::GPU> (include/El/core/DistMatrix/**) ----
// Metadata + this rank's local column-major device block.  Shifts follow
// src/core/DistMatrix/AbstractDistMatrix.cpp:399-430; lengths include/El/core/indexing/impl.hpp:33-66.
// Distribution algebra (include/El/core/types.hpp:292-337)
inline Dist Collect(Dist U) { return U == CIRC ? CIRC : STAR; }
inline Dist Partial(Dist U) { return U == VC ? MC : (U == VR ? MR : U); }
inline Dist PartialUnionRow(Dist U, Dist V) { return U == VC ? MR : (U == VR ? MC : V); }
inline Dist PartialUnionCol(Dist U, Dist V) { return PartialUnionRow(V, U); }

namespace detail {
// ElementalMatrix<T>::AlignColsWith / AlignRowsWith (src/core/DistMatrix/ElementMatrix.cpp:243-300) as a pure function:
// `mine` is the distribution of the dimension being aligned, `minePartialUnion` its PartialUnion{Col,Row}Dist, `stride` its
// stride; (dCol, dColAlign, dRow, dRowAlign) the DistData aligned with.  Returns 1 and the new alignment, 0 when the
// reference leaves the alignment alone, -1 where it raises "Nonsensical alignment" (unless allowMismatch).
inline int align_dim_with(Dist mine, Dist minePartialUnion, int stride, Dist dCol, int dColAlign, Dist dRow, int dRowAlign, int* out) {
    if (dCol == mine || dCol == Partial(mine)) { *out = dColAlign; return 1; }
    if (dRow == mine || dRow == Partial(mine)) { *out = dRowAlign; return 1; }
    if (dCol == minePartialUnion) { *out = dColAlign % stride; return 1; }
    if (dRow == minePartialUnion) { *out = dRowAlign % stride; return 1; }
    if (mine != Collect(mine) && dCol != Collect(mine) && dRow != Collect(mine)) return -1;
    return 0;
}
// The alignment an UNCONSTRAINED destination dimension adopts from the same dimension of the source in B = A:
// one line of every copy:: primitive (include/El/blas_like/level1/Copy/*.hpp) --
//   same distribution            Translate.hpp:38-42, RowAllGather.hpp:23, ColAllGather.hpp:36, {Row,Col}Filter.hpp:27-29
//                                -> the source's alignment
//   dst = Partial(src) (VC->MC)  PartialColAllGather.hpp:43-44, PartialRowAllGather.hpp:22-23, {Col,Row}AllToAllPromote.hpp:26-27
//                                -> Mod(srcAlign, dstStride)
//   src = Partial(dst) (MC->VC)  {Col,Row}AllToAllDemote.hpp:25, Partial{Col,Row}Filter.hpp:29-31  -> the source's alignment
// Unrelated distributions (MC vs MR, VC vs VR, anything vs STAR) adopt nothing.
inline bool copy_adopts(Dist src, int srcAlign, Dist dst, int dstStride, int* out) {
    if (dst == STAR || dst == CIRC || src == STAR || src == CIRC) return false;   // stride 1: nothing to align
    if (src == dst) { *out = srcAlign; return true; }
    if (dst == Partial(src)) { *out = srcAlign % dstStride; return true; }
    if (src == Partial(dst)) { *out = srcAlign; return true; }
    return false;
}
}  // namespace detail

template <typename T>
class AbstractDistMatrix {
public:
    virtual ~AbstractDistMatrix() { Empty(); }
    AbstractDistMatrix(const AbstractDistMatrix&) = delete;
    const El::Grid& Grid() const { return *grid_; }
    Int Height() const { return d_.height; }
    Int Width() const { return d_.width; }
    Int LocalHeight() const { return lh_; }
    Int LocalWidth() const { return lw_; }
    Int LDim() const { return d_.ldim; }
    int ColAlign() const { return d_.colAlign; }
    int RowAlign() const { return d_.rowAlign; }
    Int ColShift() const { return colShift_; }
    Int RowShift() const { return rowShift_; }
    int ColStride() const { return hyd_dist_stride(grid_->handle(), d_.colDist); }
    int RowStride() const { return hyd_dist_stride(grid_->handle(), d_.rowDist); }
    Dist ColDist() const { return (Dist)d_.colDist; }
    Dist RowDist() const { return (Dist)d_.rowDist; }
    T* Buffer() { return static_cast<T*>(d_.buf); }
    const T* LockedBuffer() const { return static_cast<const T*>(d_.buf); }
    Int GlobalRow(Int iLoc) const { return colShift_ + iLoc * ColStride(); }
    Int GlobalCol(Int jLoc) const { return rowShift_ + jLoc * RowStride(); }
    SyncInfoGPU& SyncInfo() { return sync_; }
    cudaStream_t Stream() const { return sync_.stream_; }
    void SetStream(cudaStream_t s) { sync_.stream_ = s; }
    const hyd_dm* desc() const { return &d_; }

    // ---- alignment (src/core/DistMatrix/ElementMatrix.cpp:170-370): explicit alignment constrains a dimension; an
    // unconstrained dimension adopts the source's alignment in B = A (AlignColsAndResize(..., force=false, constrain=false))
    bool ColConstrained() const { return colConstrained_; }
    bool RowConstrained() const { return rowConstrained_; }
    void FreeAlignments() { if (!viewing_) { colConstrained_ = false; rowConstrained_ = false; } }   // AbstractDistMatrix.cpp:93-102
    Dist PartialColDist() const { return Partial(ColDist()); }
    Dist PartialRowDist() const { return Partial(RowDist()); }
    Dist PartialUnionColDist() const { return PartialUnionCol(ColDist(), RowDist()); }
    Dist PartialUnionRowDist() const { return PartialUnionRow(ColDist(), RowDist()); }
    void Align(int colAlign, int rowAlign, bool constrain = true) {
        if (viewing_ && (colAlign != d_.colAlign || rowAlign != d_.rowAlign)) LogicError("Tried to realign a view");
        if (constrain) { colConstrained_ = true; rowConstrained_ = true; }
        d_.colAlign = colAlign; d_.rowAlign = rowAlign; Resize(d_.height, d_.width);
    }
    void AlignCols(int colAlign, bool constrain = true) {
        if (viewing_ && colAlign != d_.colAlign) LogicError("Tried to realign a view");
        if (constrain) colConstrained_ = true;
        d_.colAlign = colAlign; Resize(d_.height, d_.width);
    }
    void AlignRows(int rowAlign, bool constrain = true) {
        if (viewing_ && rowAlign != d_.rowAlign) LogicError("Tried to realign a view");
        if (constrain) rowConstrained_ = true;
        d_.rowAlign = rowAlign; Resize(d_.height, d_.width);
    }
    void AlignColsWith(const AbstractDistMatrix<T>& A, bool constrain = true, bool allowMismatch = false) {
        int a;
        const int r = detail::align_dim_with(ColDist(), PartialUnionColDist(), ColStride(), A.ColDist(), A.ColAlign(), A.RowDist(), A.RowAlign(), &a);
        if (r > 0) AlignCols(a, constrain);
        else if (r < 0 && !allowMismatch) LogicError("Nonsensical alignment");
    }
    void AlignRowsWith(const AbstractDistMatrix<T>& A, bool constrain = true, bool allowMismatch = false) {
        int a;
        const int r = detail::align_dim_with(RowDist(), PartialUnionRowDist(), RowStride(), A.ColDist(), A.ColAlign(), A.RowDist(), A.RowAlign(), &a);
        if (r > 0) AlignRows(a, constrain);
        else if (r < 0 && !allowMismatch) LogicError("Nonsensical alignment");
    }
    void AlignWith(const AbstractDistMatrix<T>& A, bool constrain = true, bool allowMismatch = false) {
        AlignColsWith(A, constrain, allowMismatch); AlignRowsWith(A, constrain, allowMismatch);
    }
    void AlignAndResize(int colAlign, int rowAlign, Int height, Int width, bool force = false, bool constrain = false) {
        if (!viewing_) {
            if (force || !colConstrained_) d_.colAlign = colAlign;
            if (force || !rowConstrained_) d_.rowAlign = rowAlign;
        }
        if (constrain) { colConstrained_ = true; rowConstrained_ = true; }
        if (force && (d_.colAlign != colAlign || d_.rowAlign != rowAlign)) LogicError("Could not set alignments");
        Resize(height, width);
    }
    void AlignColsAndResize(int colAlign, Int height, Int width, bool force = false, bool constrain = false) {
        if (!viewing_ && (force || !colConstrained_)) d_.colAlign = colAlign;
        if (constrain) colConstrained_ = true;
        if (force && d_.colAlign != colAlign) LogicError("Could not set col alignment");
        Resize(height, width);
    }
    void AlignRowsAndResize(int rowAlign, Int height, Int width, bool force = false, bool constrain = false) {
        if (!viewing_ && (force || !rowConstrained_)) d_.rowAlign = rowAlign;
        if (constrain) rowConstrained_ = true;
        if (force && d_.rowAlign != rowAlign) LogicError("Could not set row alignment");
        Resize(height, width);
    }
    void Resize(Int height, Int width) {
        d_.height = height; d_.width = width;
        int64_t cs, rs, lh, lw;
        detail::check(hyd_local_shape(grid_->handle(), &d_, &cs, &rs, &lh, &lw));
        colShift_ = cs; rowShift_ = rs; lh_ = lh; lw_ = lw;
        if (viewing_) return;
        const Int q = 16 / (Int)sizeof(T);
        const Int ld = lh > 0 ? (lh + q - 1) / q * q : q;  // 16-byte columns: TMA-friendly
        const size_t need = (size_t)ld * (size_t)(lw > 0 ? lw : 1) * sizeof(T);
        if (need > cap_) {
            release();
            // Device::CPU blocks are pinned host memory (the reference pins through its host memory pool, src/core/MemoryPool):
            // copies to and from the GPU are then asynchronous and hyd_gemm_host_* can stream them under the GEMMs
            if (dev_ == Device::GPU) detail::check_cuda(cudaMalloc(&d_.buf, need), "cudaMalloc");
            else detail::check_cuda(cudaMallocHost(&d_.buf, need), "cudaMallocHost");
            cap_ = need;
        }
        d_.ldim = ld;
    }
    Device GetLocalDevice() const { return dev_; }
    void Attach(Int height, Int width, int colAlign, int rowAlign, T* buffer, Int ldim) {
        Empty(); viewing_ = true; d_.colAlign = colAlign; d_.rowAlign = rowAlign; d_.buf = buffer; d_.ldim = ldim;
        colConstrained_ = rowConstrained_ = true;   // ElementMatrix.cpp Attach: a view's alignments are fixed
        Resize(height, width);
    }
    void Empty() { if (!viewing_) release(); d_.buf = nullptr; cap_ = 0; viewing_ = false; }

    // host <-> device helpers for tests and drivers (the reference's inter-device Copy, MC_STAR.cpp:304-390)
    void SetLocalFromHost(const T* host, Int ldh) {
        if (lh_ && lw_) detail::check_cuda(cudaMemcpy2DAsync(d_.buf, d_.ldim * sizeof(T), host, ldh * sizeof(T), lh_ * sizeof(T), lw_, cudaMemcpyDefault, sync_.stream_), "H2D");
    }
    void GetLocalToHost(T* host, Int ldh) const {
        if (lh_ && lw_) detail::check_cuda(cudaMemcpy2DAsync(host, ldh * sizeof(T), d_.buf, d_.ldim * sizeof(T), lh_ * sizeof(T), lw_, cudaMemcpyDefault, sync_.stream_), "D2H");
        detail::check_cuda(cudaStreamSynchronize(sync_.stream_), "sync");
    }

protected:
    AbstractDistMatrix(const El::Grid& g, int cd, int rd, Device dev = Device::GPU) : grid_(&g), dev_(dev) {
        d_.buf = nullptr; d_.ldim = 1; d_.height = d_.width = 0; d_.colDist = cd; d_.rowDist = rd; d_.colAlign = d_.rowAlign = 0; d_.root = 0; d_.reserved = 0;
    }
    const El::Grid* grid_;
    hyd_dm d_;
    Int colShift_ = 0, rowShift_ = 0, lh_ = 0, lw_ = 0;
    size_t cap_ = 0;
    bool viewing_ = false;
    bool colConstrained_ = false, rowConstrained_ = false;
    Device dev_ = Device::GPU;
    SyncInfoGPU sync_;
    void release() {
        if (!d_.buf) return;
        if (dev_ == Device::GPU) cudaFree(d_.buf); else cudaFreeHost(d_.buf);
        d_.buf = nullptr;
    }
};

namespace detail {
// every operation except assignment and Gemm-on-host-blocks needs its operands on the GPU: there is no CPU path to fall back to
template <typename T>
inline void require_gpu(const AbstractDistMatrix<T>& A, const char* who) {
    if (A.GetLocalDevice() != Device::GPU) LogicError(std::string(who) + ": Device::CPU matrix; this library computes on the GPU only (assign it to a Device::GPU matrix)");
}
}  // namespace detail

// B = A between any two supported distributions (DistMatrix::operator= tables,
// src/core/DistMatrix/ElementMatrix/*.cpp -> copy::* primitives)
// An unconstrained dimension of B first adopts the alignment of the same dimension of A where the two distributions are
// related (detail::copy_adopts: the AlignColsAndResize / AlignRowsAndResize line of the reference's copy:: primitive), so a
// `B = A` into a fresh [MC,*] / [VC,*] / ... matrix lands exactly where the reference puts it, with no realignment traffic.
template <typename T>
void Copy(const AbstractDistMatrix<T>& A, AbstractDistMatrix<T>& B) {
    int ca = B.ColAlign(), ra = B.RowAlign(), a;
    if (!B.ColConstrained() && detail::copy_adopts(A.ColDist(), A.ColAlign(), B.ColDist(), B.ColStride(), &a)) ca = a;
    if (!B.RowConstrained() && detail::copy_adopts(A.RowDist(), A.RowAlign(), B.RowDist(), B.RowStride(), &a)) ra = a;
    B.AlignAndResize(ca, ra, A.Height(), A.Width(), false, false);
    if (A.GetLocalDevice() != Device::GPU || B.GetLocalDevice() != Device::GPU) {
        // Cross-device assignment (the INSTGPU instantiations, src/core/DistMatrix/ElementMatrix/MC_STAR.cpp:304-390 ->
        // Copy(Matrix<T,CPU>, Matrix<T,GPU>), Copy.hpp:140-207): same distribution and alignments, so it is a 2-D copy of the
        // local block.  Redistributions run on the GPU only: there is no CPU path in this library.
        if (A.GetLocalDevice() == Device::CPU && B.GetLocalDevice() == Device::CPU)
            LogicError("B = A between two Device::CPU matrices: this library has no CPU path; assign to a Device::GPU matrix");
        if (A.ColDist() != B.ColDist() || A.RowDist() != B.RowDist() || A.ColAlign() != B.ColAlign() || A.RowAlign() != B.RowAlign())
            LogicError("cross-device B = A needs equal distributions and alignments; redistribute on the GPU side");
        const int kind = A.GetLocalDevice() == Device::CPU ? 0 : 1;
        MultiSync fence(B.Stream(), {A.Stream()});
        detail::check(hyd_memcopy2d_async(B.desc()->buf, B.LDim(), A.desc()->buf, A.LDim(), A.LocalHeight(), A.LocalWidth(),
                                          (int)sizeof(T), kind, B.Stream()));
        return;
    }
    MultiSync fence(B.Stream(), {A.Stream()});
    detail::check(detail::Sfx<T>::redist(B.Grid().handle(), A.desc(), B.desc(), 0, B.Stream()));
}
// B = A^T (Transpose.hpp:318-356); conjugate is a no-op for real T
template <typename T>
void Transpose(const AbstractDistMatrix<T>& A, AbstractDistMatrix<T>& B, bool /*conjugate*/ = false) {
    detail::require_gpu(A, "Transpose");
    detail::require_gpu(B, "Transpose");
    // Transpose.hpp:191-250: B's column dimension relates to A's ROW dimension and vice versa.  With exactly transposed
    // distributions and compatible (or free) alignments the reference calls B.Align(A.RowAlign(), A.ColAlign()), which also
    // constrains B (:202-210); the filter / all-gather branches adopt without constraining.
    int ca = B.ColAlign(), ra = B.RowAlign(), a;
    const bool transposed = A.ColDist() == B.RowDist() && A.RowDist() == B.ColDist();
    if (transposed && (A.ColAlign() == B.RowAlign() || !B.RowConstrained()) && (A.RowAlign() == B.ColAlign() || !B.ColConstrained())) {
        B.Align(A.RowAlign(), A.ColAlign(), true);
    } else {
        if (!B.ColConstrained() && detail::copy_adopts(A.RowDist(), A.RowAlign(), B.ColDist(), B.ColStride(), &a)) ca = a;
        if (!B.RowConstrained() && detail::copy_adopts(A.ColDist(), A.ColAlign(), B.RowDist(), B.RowStride(), &a)) ra = a;
        B.AlignAndResize(ca, ra, A.Width(), A.Height(), false, false);
    }
    B.Resize(A.Width(), A.Height());
    MultiSync fence(B.Stream(), {A.Stream()});
    detail::check(detail::Sfx<T>::redist(B.Grid().handle(), A.desc(), B.desc(), 1, B.Stream()));
}
template <typename T>
void AxpyContract(T alpha, const AbstractDistMatrix<T>& A, AbstractDistMatrix<T>& B) {  // AxpyContract.hpp:483-544
    detail::require_gpu(A, "AxpyContract");
    detail::require_gpu(B, "AxpyContract");
    MultiSync fence(B.Stream(), {A.Stream()});
    detail::check(detail::Sfx<T>::contract(B.Grid().handle(), alpha, A.desc(), B.desc(), 0, B.Stream()));
}
template <typename T>
void TransposeAxpyContract(T alpha, const AbstractDistMatrix<T>& A, AbstractDistMatrix<T>& B, bool = false) {  // TransposeAxpyContract.hpp:14-48
    detail::require_gpu(A, "TransposeAxpyContract");
    detail::require_gpu(B, "TransposeAxpyContract");
    MultiSync fence(B.Stream(), {A.Stream()});
    detail::check(detail::Sfx<T>::contract(B.Grid().handle(), alpha, A.desc(), B.desc(), 1, B.Stream()));
}

template <typename T, Dist U = MC, Dist V = MR, DistWrap W = ELEMENT, Device D = Device::GPU>
class DistMatrix : public AbstractDistMatrix<T> {
    static_assert(W == ELEMENT, "only ELEMENT-wrapped matrices are on the Gemm path");
public:
    // Device::CPU matrices hold their block in pinned host memory: they can be filled, assigned to / from a Device::GPU
    // matrix of the same distribution, and passed to Gemm (all three on the CPU: hyd_gemm_host_*); nothing computes on the CPU
    explicit DistMatrix(const El::Grid& g) : AbstractDistMatrix<T>(g, U, V, D) { this->Resize(0, 0); }
    DistMatrix(Int height, Int width, const El::Grid& g) : AbstractDistMatrix<T>(g, U, V, D) { this->Resize(height, width); }
    template <Dist U2, Dist V2, Device D2>
    DistMatrix& operator=(const DistMatrix<T, U2, V2, W, D2>& A) { Copy<T>(A, *this); return *this; }
    DistMatrix& operator=(const DistMatrix& A) { Copy<T>(A, *this); return *this; }
};

// C := alpha op(A) op(B) + beta C   (level3.hpp:66-84; Gemm.cpp:278-307)
template <typename T>
void Gemm(Orientation orientA, Orientation orientB, T alpha, const AbstractDistMatrix<T>& A, const AbstractDistMatrix<T>& B,
          T beta, AbstractDistMatrix<T>& C, GemmAlgorithm alg = GEMM_DEFAULT) {
    const int ncpu = (A.GetLocalDevice() == Device::CPU) + (B.GetLocalDevice() == Device::CPU) + (C.GetLocalDevice() == Device::CPU);
    MultiSync fence(C.Stream(), {A.Stream(), B.Stream()});   // Gemm.cpp:117-119
    if (ncpu == 3) {
        // all three blocks in (pinned) host memory: the library uploads them panel by panel under the GEMMs and brings C back
        // (what a CPU-resident caller of the reference does with three cross-device assignments around Gemm); returns synchronised
        detail::check(detail::Sfx<T>::gemm_host(C.Grid().handle(), orientA, orientB, alg, alpha, A.desc(), B.desc(), beta, C.desc(),
                                                Blocksize(), C.Stream()));
        return;
    }
    if (ncpu != 0) LogicError("Gemm: A, B and C must live on the same device");
    detail::check(detail::Sfx<T>::gemm(C.Grid().handle(), orientA, orientB, alg, alpha, A.desc(), B.desc(), beta, C.desc(),
                                       Blocksize(), C.Stream()));
}
// C := alpha op(A) op(B), C resized (level3.hpp:76-84)
template <typename T>
void Gemm(Orientation orientA, Orientation orientB, T alpha, const AbstractDistMatrix<T>& A, const AbstractDistMatrix<T>& B,
          AbstractDistMatrix<T>& C, GemmAlgorithm alg = GEMM_DEFAULT) {
    const Int m = orientA == NORMAL ? A.Height() : A.Width(), n = orientB == NORMAL ? B.Width() : B.Height();
    C.Resize(m, n);
    Gemm(orientA, orientB, alpha, A, B, T(0), C, alg);
}

// ---- triangular rank-k updates and the trapezoid routines around them ----
// triangle `uplo` of C := alpha op(A) op(B) + beta C   (src/blas_like/level3/Trrk.cpp:98-115)
template <typename T>
void Trrk(UpperOrLower uplo, Orientation orientA, Orientation orientB, T alpha, const AbstractDistMatrix<T>& A,
          const AbstractDistMatrix<T>& B, T beta, AbstractDistMatrix<T>& C) {
    detail::require_gpu(A, "Trrk");
    detail::require_gpu(B, "Trrk");
    detail::require_gpu(C, "Trrk");
    MultiSync fence(C.Stream(), {A.Stream(), B.Stream()});
    detail::check(detail::Sfx<T>::trrk(C.Grid().handle(), uplo, orientA, orientB, alpha, A.desc(), B.desc(), beta, C.desc(), Blocksize(),
                                       C.Stream()));
}
// triangle `uplo` of C := alpha A A^T + beta C (NORMAL) / alpha A^T A + beta C   (src/blas_like/level3/Syrk.cpp:72-90)
template <typename T>
void Syrk(UpperOrLower uplo, Orientation orientation, T alpha, const AbstractDistMatrix<T>& A, T beta, AbstractDistMatrix<T>& C,
          bool /*conjugate*/ = false) {
    detail::require_gpu(A, "Syrk");
    detail::require_gpu(C, "Syrk");
    MultiSync fence(C.Stream(), {A.Stream()});
    detail::check(detail::Sfx<T>::syrk(C.Grid().handle(), uplo, orientation, alpha, A.desc(), beta, C.desc(), Blocksize(), C.Stream()));
}
// C resized to n x n and zeroed first (Syrk.cpp:92-103)
template <typename T>
void Syrk(UpperOrLower uplo, Orientation orientation, T alpha, const AbstractDistMatrix<T>& A, AbstractDistMatrix<T>& C,
          bool conjugate = false) {
    const Int n = orientation == NORMAL ? A.Height() : A.Width();
    C.Resize(n, n);
    detail::check(detail::Sfx<T>::fill(C.LocalHeight(), C.LocalWidth(), T(0), C.Buffer(), C.LDim(), C.Stream()));
    Syrk(uplo, orientation, alpha, A, T(0), C, conjugate);
}
// El::Herk (src/blas_like/level3/Herk.cpp): Syrk with conjugate = true; identical for real T
template <typename T>
void Herk(UpperOrLower uplo, Orientation orientation, T alpha, const AbstractDistMatrix<T>& A, T beta, AbstractDistMatrix<T>& C) {
    Syrk(uplo, orientation, alpha, A, beta, C, true);
}
template <typename T>
void Herk(UpperOrLower uplo, Orientation orientation, T alpha, const AbstractDistMatrix<T>& A, AbstractDistMatrix<T>& C) {
    Syrk(uplo, orientation, alpha, A, C, true);
}
// include/El/blas_like/level1/{ScaleTrapezoid,AxpyTrapezoid,MakeTrapezoidal}.hpp on DistMatrices
template <typename T>
void ScaleTrapezoid(T alpha, UpperOrLower uplo, AbstractDistMatrix<T>& A, Int offset = 0) {
    detail::require_gpu(A, "ScaleTrapezoid");
    detail::check(detail::Sfx<T>::scale_trap(A.Grid().handle(), alpha, uplo, A.desc(), offset, A.Stream()));
}
template <typename T>
void AxpyTrapezoid(UpperOrLower uplo, T alpha, const AbstractDistMatrix<T>& X, AbstractDistMatrix<T>& Y, Int offset = 0) {
    detail::require_gpu(X, "AxpyTrapezoid");
    detail::require_gpu(Y, "AxpyTrapezoid");
    MultiSync fence(Y.Stream(), {X.Stream()});
    detail::check(detail::Sfx<T>::axpy_trap(Y.Grid().handle(), uplo, alpha, X.desc(), Y.desc(), offset, Y.Stream()));
}
template <typename T>
void MakeTrapezoidal(UpperOrLower uplo, AbstractDistMatrix<T>& A, Int offset = 0) {
    detail::require_gpu(A, "MakeTrapezoidal");
    detail::check(detail::Sfx<T>::make_trap(A.Grid().handle(), uplo, A.desc(), offset, A.Stream()));
}

template <typename T>
void MakeSymmetric(UpperOrLower uplo, AbstractDistMatrix<T>& A, bool /*conjugate*/ = false) {   // MakeSymmetric.hpp:52-68
    detail::require_gpu(A, "MakeSymmetric");
    detail::check(detail::Sfx<T>::make_symm(A.Grid().handle(), uplo, A.desc(), A.Stream()));
}
template <typename T>
void MakeHermitian(UpperOrLower uplo, AbstractDistMatrix<T>& A) { MakeSymmetric(uplo, A, true); }
// C := alpha A B + beta C (LEFT) / alpha B A + beta C (RIGHT), A symmetric, `uplo` triangle read   (Symm.cpp:56-80)
template <typename T>
void Symm(LeftOrRight side, UpperOrLower uplo, T alpha, const AbstractDistMatrix<T>& A, const AbstractDistMatrix<T>& B, T beta,
          AbstractDistMatrix<T>& C, bool /*conjugate*/ = false) {
    detail::require_gpu(A, "Symm");
    detail::require_gpu(B, "Symm");
    detail::require_gpu(C, "Symm");
    MultiSync fence(C.Stream(), {A.Stream(), B.Stream()});
    detail::check(detail::Sfx<T>::symm(C.Grid().handle(), side, uplo, alpha, A.desc(), B.desc(), beta, C.desc(), Blocksize(), C.Stream()));
}
template <typename T>
void Hemm(LeftOrRight side, UpperOrLower uplo, T alpha, const AbstractDistMatrix<T>& A, const AbstractDistMatrix<T>& B, T beta,
          AbstractDistMatrix<T>& C) {   // src/blas_like/level3/Hemm.cpp
    Symm(side, uplo, alpha, A, B, beta, C, true);
}

// local column-major GEMM on raw device buffers (gpu_blas::Gemm, include/hydrogen/blas/GPU_BLAS_impl.hpp:631-646)
namespace gpu_blas {
template <typename T>
void Gemm(Orientation tA, Orientation tB, Int m, Int n, Int k, T alpha, const T* A, Int lda, const T* B, Int ldb, T beta, T* C,
          Int ldc, const SyncInfoGPU& si) {
    detail::check(detail::Sfx<T>::lgemm(tA != NORMAL, tB != NORMAL, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, si.stream_));
}
// 2-D copy / transposing copy / axpy / scale on raw device buffers (include/hydrogen/blas/GPU_BLAS_impl.hpp:84-209,558-567)
template <typename T>
void Copy(Orientation trans, Int m, Int n, const T* A, Int lda, T* B, Int ldb, const SyncInfoGPU& si) {
    // as in the reference, m x n is the shape of A; a transposing copy writes the n x m matrix B
    detail::check(trans == NORMAL ? detail::Sfx<T>::copy2d(m, n, A, lda, B, ldb, si.stream_)
                                  : detail::Sfx<T>::transpose(m, n, A, lda, B, ldb, si.stream_));
}
template <typename T>
void Axpy(Int m, Int n, T alpha, const T* A, Int lda, T* B, Int ldb, const SyncInfoGPU& si) {
    detail::check(detail::Sfx<T>::axpy2d(m, n, alpha, A, lda, B, ldb, si.stream_));
}
template <typename T>
void Scale(Int m, Int n, T alpha, T* A, Int lda, const SyncInfoGPU& si) {
    detail::check(detail::Sfx<T>::scale2d(m, n, alpha, A, lda, si.stream_));
}
// B += alpha op(A), B is m x n (GPU_BLAS_impl.hpp:531-541)
template <typename T>
void Axpy(Orientation transA, Int m, Int n, T alpha, const T* A, Int lda, T* B, Int ldb, const SyncInfoGPU& si) {
    detail::check(detail::Sfx<T>::axpy_trans(transA != NORMAL, m, n, alpha, A, lda, B, ldb, si.stream_));
}
// two-stride 2-D forms (GPU_BLAS_impl.hpp:543-548,570-588): the lattice kernels behind copy::util / axpy::util
template <typename T>
void Copy(Int m, Int n, const T* A, Int rowStrideA, Int lda, T* B, Int rowStrideB, Int ldb, const SyncInfoGPU& si) {
    detail::check(detail::Sfx<T>::interleave(m, n, A, rowStrideA, lda, B, rowStrideB, ldb, si.stream_));
}
template <typename T>
void Axpy(Int m, Int n, T alpha, const T* A, Int rowStrideA, Int lda, T* B, Int rowStrideB, Int ldb, const SyncInfoGPU& si) {
    detail::check(detail::Sfx<T>::interleave_update(alpha, m, n, A, rowStrideA, lda, B, rowStrideB, ldb, si.stream_));
}
// 1-D forms with increments (GPU_BLAS_impl.hpp:510-518,550-556,590-597)
template <typename T>
void Copy(Int size, const T* X, Int incx, T* Y, Int incy, const SyncInfoGPU& si) {
    detail::check(detail::Sfx<T>::copy1d(size, X, incx, Y, incy, si.stream_));
}
template <typename T>
void Axpy(Int size, T alpha, const T* X, Int incx, T* Y, Int incy, const SyncInfoGPU& si) {
    detail::check(detail::Sfx<T>::axpy1d(size, alpha, X, incx, Y, incy, si.stream_));
}
template <typename T>
void Scale(Int size, T alpha, T* X, Int incx, const SyncInfoGPU& si) {
    detail::check(detail::Sfx<T>::scale1d(size, alpha, X, incx, si.stream_));
}
}  // namespace gpu_blas

// hydrogen::Fill_GPU_impl / Hadamard_GPU_impl (include/hydrogen/blas/gpu/{Fill,Hadamard}.hpp)
namespace hydrogen_gpu {
template <typename T>
void Fill_GPU_impl(size_t height, size_t width, T alpha, T* buffer, size_t ldim, cudaStream_t stream) {
    detail::check(detail::Sfx<T>::fill((Int)height, (Int)width, alpha, buffer, (Int)ldim, stream));
}
template <typename T>
void Hadamard_GPU_impl(size_t height, size_t width, const T* X, size_t colStrideX, size_t rowStrideX, const T* Y, size_t colStrideY,
                       size_t rowStrideY, T* Z, size_t colStrideZ, size_t rowStrideZ, cudaStream_t stream) {
    detail::check(detail::Sfx<T>::hadamard((Int)height, (Int)width, X, (Int)colStrideX, (Int)rowStrideX, Y, (Int)colStrideY, (Int)rowStrideY,
                                           Z, (Int)colStrideZ, (Int)rowStrideZ, stream));
}
}  // namespace hydrogen_gpu

// InterDeviceCopy<SrcD,DestD> (include/hydrogen/device/gpu/CUDA.hpp:229-420): the reference's names and argument order
template <Device SrcD, Device DestD> struct InterDeviceCopy {
    static constexpr int kind() { return SrcD == Device::CPU ? (DestD == Device::GPU ? 0 : -1) : (DestD == Device::CPU ? 1 : 2); }
    template <typename T>
    static void MemCopy1DAsync(T* dest, const T* src, size_t size, cudaStream_t stream = nullptr) {
        static_assert(kind() >= 0, "host-to-host copies do not go through the device library");
        detail::check(hyd_memcopy1d_async(dest, src, (Int)size, (int)sizeof(T), kind(), stream));
    }
    template <typename T>
    static void MemCopy2DAsync(T* dest, size_t dest_ldim, const T* src, size_t src_ldim, size_t height, size_t width,
                               cudaStream_t stream = nullptr) {
        static_assert(kind() >= 0, "host-to-host copies do not go through the device library");
        detail::check(hyd_memcopy2d_async(dest, (Int)dest_ldim, src, (Int)src_ldim, (Int)height, (Int)width, (int)sizeof(T), kind(),
                                          stream));
    }
};

// ---- copy::util / axpy::util on device buffers (include/El/blas_like/level1/Copy/util.hpp, Axpy/util.hpp): the
// reference's names and argument order; each call is one kernel launch for all portions ----
namespace copy { namespace util {
template <typename T>
void InterleaveMatrix(Int height, Int width, const T* A, Int colStrideA, Int rowStrideA, T* B, Int colStrideB, Int rowStrideB, const SyncInfoGPU& si) {
    detail::check(detail::Sfx<T>::interleave(height, width, A, colStrideA, rowStrideA, B, colStrideB, rowStrideB, si.stream_));
}
template <typename T>
void RowStridedPack(Int height, Int width, Int rowAlign, Int rowStride, const T* A, Int ALDim, T* BPortions, Int portionSize, const SyncInfoGPU& si) {
    detail::check(detail::Sfx<T>::row_pack(height, width, rowAlign, rowStride, A, ALDim, BPortions, portionSize, si.stream_));
}
template <typename T>
void RowStridedUnpack(Int height, Int width, Int rowAlign, Int rowStride, const T* APortions, Int portionSize, T* B, Int BLDim, const SyncInfoGPU& si) {
    detail::check(detail::Sfx<T>::row_unpack(height, width, rowAlign, rowStride, APortions, portionSize, B, BLDim, si.stream_));
}
template <typename T>
void PartialRowStridedPack(Int height, Int width, Int rowAlign, Int rowStride, Int rowStrideUnion, Int rowStridePart, Int rowRankPart, Int rowShiftA,
                           const T* A, Int ALDim, T* BPortions, Int portionSize, const SyncInfoGPU& si) {
    detail::check(detail::Sfx<T>::prow_pack(height, width, rowAlign, rowStride, rowStrideUnion, rowStridePart, rowRankPart, rowShiftA, A, ALDim, BPortions, portionSize, si.stream_));
}
template <typename T>
void PartialRowStridedUnpack(Int height, Int width, Int rowAlign, Int rowStride, Int rowStrideUnion, Int rowStridePart, Int rowRankPart, Int rowShiftB,
                             const T* APortions, Int portionSize, T* B, Int BLDim, const SyncInfoGPU& si) {
    detail::check(detail::Sfx<T>::prow_unpack(height, width, rowAlign, rowStride, rowStrideUnion, rowStridePart, rowRankPart, rowShiftB, APortions, portionSize, B, BLDim, si.stream_));
}
template <typename T>
void ColStridedPack(Int height, Int width, Int colAlign, Int colStride, const T* A, Int ALDim, T* BPortions, Int portionSize, const SyncInfoGPU& si) {
    detail::check(detail::Sfx<T>::col_pack(height, width, colAlign, colStride, A, ALDim, BPortions, portionSize, si.stream_));
}
template <typename T>
void ColStridedUnpack(Int height, Int width, Int colAlign, Int colStride, const T* APortions, Int portionSize, T* B, Int BLDim, const SyncInfoGPU& si) {
    detail::check(detail::Sfx<T>::col_unpack(height, width, colAlign, colStride, APortions, portionSize, B, BLDim, si.stream_));
}
template <typename T>
void PartialColStridedPack(Int height, Int width, Int colAlign, Int colStride, Int colStrideUnion, Int colStridePart, Int colRankPart, Int colShiftA,
                           const T* A, Int ALDim, T* BPortions, Int portionSize, const SyncInfoGPU& si) {
    detail::check(detail::Sfx<T>::pcol_pack(height, width, colAlign, colStride, colStrideUnion, colStridePart, colRankPart, colShiftA, A, ALDim, BPortions, portionSize, si.stream_));
}
template <typename T>
void PartialColStridedUnpack(Int height, Int width, Int colAlign, Int colStride, Int colStrideUnion, Int colStridePart, Int colRankPart, Int colShiftB,
                             const T* APortions, Int portionSize, T* B, Int BLDim, const SyncInfoGPU& si) {
    detail::check(detail::Sfx<T>::pcol_unpack(height, width, colAlign, colStride, colStrideUnion, colStridePart, colRankPart, colShiftB, APortions, portionSize, B, BLDim, si.stream_));
}
template <typename T>
void StridedPack(Int height, Int width, Int colAlign, Int colStride, Int rowAlign, Int rowStride, const T* A, Int ALDim, T* BPortions, Int portionSize, const SyncInfoGPU& si) {
    detail::check(detail::Sfx<T>::pack2d(height, width, colAlign, colStride, rowAlign, rowStride, A, ALDim, BPortions, portionSize, si.stream_));
}
template <typename T>
void StridedUnpack(Int height, Int width, Int colAlign, Int colStride, Int rowAlign, Int rowStride, const T* APortions, Int portionSize, T* B, Int BLDim, const SyncInfoGPU& si) {
    detail::check(detail::Sfx<T>::unpack2d(height, width, colAlign, colStride, rowAlign, rowStride, APortions, portionSize, B, BLDim, si.stream_));
}
} }  // namespace copy::util
namespace axpy { namespace util {
template <typename T>
void InterleaveMatrixUpdate(T alpha, Int height, Int width, const T* A, Int colStrideA, Int rowStrideA, T* B, Int colStrideB, Int rowStrideB, const SyncInfoGPU& si) {
    detail::check(detail::Sfx<T>::interleave_update(alpha, height, width, A, colStrideA, rowStrideA, B, colStrideB, rowStrideB, si.stream_));
}
} }  // namespace axpy::util

// C := alpha op(A) op(B) + beta C on the LOCAL blocks of conformally distributed matrices (Gemm.cpp:323-431): the
// distribution / alignment / shape checks of the reference's debug build, then the local kernel
template <typename T>
void LocalGemm(Orientation orientA, Orientation orientB, T alpha, const AbstractDistMatrix<T>& A, const AbstractDistMatrix<T>& B,
               T beta, AbstractDistMatrix<T>& C) {
    const bool tA = orientA != NORMAL, tB = orientB != NORMAL;
    const Dist aOut = tA ? A.RowDist() : A.ColDist(), aIn = tA ? A.ColDist() : A.RowDist();
    const Dist bIn = tB ? B.RowDist() : B.ColDist(), bOut = tB ? B.ColDist() : B.RowDist();
    if (aOut != C.ColDist() || aIn != bIn || bOut != C.RowDist()) LogicError("LocalGemm: nonconformal distributions");
    const int aOutAl = tA ? A.RowAlign() : A.ColAlign(), aInAl = tA ? A.ColAlign() : A.RowAlign();
    const int bInAl = tB ? B.RowAlign() : B.ColAlign(), bOutAl = tB ? B.ColAlign() : B.RowAlign();
    if (aOutAl != C.ColAlign() || aInAl != bInAl || bOutAl != C.RowAlign()) LogicError("LocalGemm: operands are not aligned");
    const Int m = tA ? A.Width() : A.Height(), k = tA ? A.Height() : A.Width();
    const Int kB = tB ? B.Width() : B.Height(), n = tB ? B.Height() : B.Width();
    if (m != C.Height() || k != kB || n != C.Width()) LogicError("Nonconformal LocalGemm");
    const Int kLoc = tA ? A.LocalHeight() : A.LocalWidth();
    MultiSync fence(C.Stream(), {A.Stream(), B.Stream()});
    detail::check(detail::Sfx<T>::lgemm(tA, tB, C.LocalHeight(), C.LocalWidth(), kLoc, alpha, A.LockedBuffer(), A.LDim(), B.LockedBuffer(),
                                        B.LDim(), beta, C.Buffer(), C.LDim(), C.Stream()));
}
template <typename T>
void LocalGemm(Orientation orientA, Orientation orientB, T alpha, const AbstractDistMatrix<T>& A, const AbstractDistMatrix<T>& B,
               AbstractDistMatrix<T>& C) {
    C.Resize(orientA == NORMAL ? A.Height() : A.Width(), orientB == NORMAL ? B.Width() : B.Height());
    LocalGemm(orientA, orientB, alpha, A, B, T(0), C);
}

// local scaling of a distributed matrix (El::Scale(beta, C), Gemm.cpp:287) and Zero
template <typename T>
void Scale(T alpha, AbstractDistMatrix<T>& A) {
    detail::check(detail::Sfx<T>::scale2d(A.LocalHeight(), A.LocalWidth(), alpha, A.Buffer(), A.LDim(), A.Stream()));
}
// every local entry := alpha (Fill.hpp); Zero (Zero.hpp) is Fill(0): also clears NaN / Inf, as the reference's memset does
template <typename T>
void Fill(AbstractDistMatrix<T>& A, T alpha) {
    detail::check(detail::Sfx<T>::fill(A.LocalHeight(), A.LocalWidth(), alpha, A.Buffer(), A.LDim(), A.Stream()));
}
template <typename T>
void Zero(AbstractDistMatrix<T>& A) { Fill(A, T(0)); }
// Y += alpha X between any two distributions (Axpy.hpp:161-187): local if the DistData agree, else X is copied into Y's
// distribution and alignment first
template <typename T>
void Axpy(T alpha, const AbstractDistMatrix<T>& X, AbstractDistMatrix<T>& Y) {
    MultiSync fence(Y.Stream(), {X.Stream()});
    detail::check(detail::Sfx<T>::axpy_dist(Y.Grid().handle(), alpha, X.desc(), Y.desc(), 0, Y.Stream()));
}
// Y += alpha X^T (TransposeAxpy.hpp); conjugate is a no-op for real T
template <typename T>
void TransposeAxpy(T alpha, const AbstractDistMatrix<T>& X, AbstractDistMatrix<T>& Y, bool /*conjugate*/ = false) {
    MultiSync fence(Y.Stream(), {X.Stream()});
    detail::check(detail::Sfx<T>::axpy_dist(Y.Grid().handle(), alpha, X.desc(), Y.desc(), 1, Y.Stream()));
}
template <typename T>
void AdjointAxpy(T alpha, const AbstractDistMatrix<T>& X, AbstractDistMatrix<T>& Y) { TransposeAxpy(alpha, X, Y, true); }
// C(i,j) := A(i,j) B(i,j) (Hadamard.hpp:109-135): same distribution, A and B aligned; C is aligned with A and resized
template <typename T>
void Hadamard(const AbstractDistMatrix<T>& A, const AbstractDistMatrix<T>& B, AbstractDistMatrix<T>& C) {
    if (A.Height() != B.Height() || A.Width() != B.Width()) LogicError("Hadamard product requires equal dimensions");
    if (A.ColDist() != B.ColDist() || A.RowDist() != B.RowDist() || B.ColDist() != C.ColDist() || B.RowDist() != C.RowDist())
        LogicError("A, B, and C must share the same distribution");
    if (A.ColAlign() != B.ColAlign() || A.RowAlign() != B.RowAlign()) LogicError("A and B must be aligned");
    if (C.ColAlign() != A.ColAlign() || C.RowAlign() != A.RowAlign()) C.Align(A.ColAlign(), A.RowAlign());
    if (C.Height() != A.Height() || C.Width() != A.Width()) C.Resize(A.Height(), A.Width());
    MultiSync fence(C.Stream(), {A.Stream(), B.Stream()});
    detail::check(detail::Sfx<T>::hadamard(A.LocalHeight(), A.LocalWidth(), A.LockedBuffer(), 1, A.LDim(), B.LockedBuffer(), 1, B.LDim(),
                                           C.Buffer(), 1, C.LDim(), C.Stream()));
}
// B := sum of A over the ranks along which A is a partial sum (Contract.hpp:14-98): align with A, zero, AxpyContract
template <typename T>
void Contract(const AbstractDistMatrix<T>& A, AbstractDistMatrix<T>& B) {
    if (A.ColDist() == B.ColDist() && A.RowDist() == B.RowDist()) { Copy(A, B); return; }
    B.AlignWith(A);
    B.Resize(A.Height(), A.Width());
    Zero(B);
    AxpyContract(T(1), A, B);
}

}  // namespace El
#endif  // EL_B200_SHIM_HPP

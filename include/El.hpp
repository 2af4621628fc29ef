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

// ---- communicator: one process per GPU over NCCL (or P rank-threads on one GPU for tests) ----
namespace mpi {
class Comm {
public:
    Comm() = default;
    // every rank passes the same 128-byte id (hyd_nccl_unique_id on rank 0, broadcast by the launcher)
    static Comm NCCL(const void* id128, int rank, int size) { Comm c; detail::check(hyd_world_create_nccl(id128, rank, size, &c.w_)); return c; }
    static Comm Local(hyd_local_group* grp, int rank) { Comm c; detail::check(hyd_world_create_local(grp, rank, &c.w_)); return c; }
    static Comm Self() { Comm c; detail::check(hyd_local_group_create(1, &c.grp_)); detail::check(hyd_world_create_local(c.grp_, 0, &c.w_)); return c; }
    hyd_world* handle() const { return w_; }
    void Barrier() const { detail::check(hyd_world_barrier(w_)); }
    void Free() { if (w_) hyd_world_destroy(w_); if (grp_) hyd_local_group_destroy(grp_); w_ = nullptr; grp_ = nullptr; }
private:
    hyd_world* w_ = nullptr;
    hyd_local_group* grp_ = nullptr;
};
}  // namespace mpi

// ---- El::Grid (Grid.hpp:15-146; Grid.cpp:66-216) ----
class Grid {
public:
    Grid(const mpi::Comm& comm, int height, GridOrder order = COLUMN_MAJOR) { detail::check(hyd_grid_create(comm.handle(), height, order, &g_)); }
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
private:
    hyd_grid* g_ = nullptr;
};

namespace detail {
template <> struct Sfx<double> {
    static int redist(hyd_grid* g, const hyd_dm* a, const hyd_dm* b, int t, void* s) { return hyd_redist_f64(g, a, b, t, s); }
    static int contract(hyd_grid* g, double al, const hyd_dm* a, const hyd_dm* b, int t, void* s) { return hyd_axpy_contract_f64(g, al, a, b, t, s); }
    static int gemm(hyd_grid* g, int oA, int oB, int alg, double al, const hyd_dm* A, const hyd_dm* B, double be, const hyd_dm* C, int64_t nb, void* s) { return hyd_gemm_f64(g, oA, oB, alg, al, A, B, be, C, nb, s); }
    static int lgemm(int tA, int tB, int64_t m, int64_t n, int64_t k, double al, const double* A, int64_t lda, const double* B, int64_t ldb, double be, double* C, int64_t ldc, void* s) { return hyd_lgemm_f64(tA, tB, m, n, k, al, A, lda, B, ldb, be, C, ldc, s); }
};
template <> struct Sfx<float> {
    static int redist(hyd_grid* g, const hyd_dm* a, const hyd_dm* b, int t, void* s) { return hyd_redist_f32(g, a, b, t, s); }
    static int contract(hyd_grid* g, float al, const hyd_dm* a, const hyd_dm* b, int t, void* s) { return hyd_axpy_contract_f32(g, al, a, b, t, s); }
    static int gemm(hyd_grid* g, int oA, int oB, int alg, float al, const hyd_dm* A, const hyd_dm* B, float be, const hyd_dm* C, int64_t nb, void* s) { return hyd_gemm_f32(g, oA, oB, alg, al, A, B, be, C, nb, s); }
    static int lgemm(int tA, int tB, int64_t m, int64_t n, int64_t k, float al, const float* A, int64_t lda, const float* B, int64_t ldb, float be, float* C, int64_t ldc, void* s) { return hyd_lgemm_f32(tA, tB, m, n, k, al, A, lda, B, ldb, be, C, ldc, s); }
};
}  // namespace detail

// ---- AbstractDistMatrix<T> / DistMatrix<T,U,V,ELEMENT,Device::GPU> (include/El/core/DistMatrix/**) ----
// Metadata + this rank's local column-major device block.  Shifts follow
// src/core/DistMatrix/AbstractDistMatrix.cpp:399-430; lengths include/El/core/indexing/impl.hpp:33-66.
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

    void Align(int colAlign, int rowAlign) { d_.colAlign = colAlign; d_.rowAlign = rowAlign; Resize(d_.height, d_.width); }
    void AlignWith(const AbstractDistMatrix<T>& A) {  // ElementMatrix.cpp AlignWith: adopt alignments of matching dists
        if (A.d_.colDist == d_.colDist) d_.colAlign = A.d_.colAlign; else if (A.d_.rowDist == d_.colDist) d_.colAlign = A.d_.rowAlign;
        if (A.d_.rowDist == d_.rowDist) d_.rowAlign = A.d_.rowAlign; else if (A.d_.colDist == d_.rowDist) d_.rowAlign = A.d_.colAlign;
        Resize(d_.height, d_.width);
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
            if (d_.buf) detail::check_cuda(cudaFree(d_.buf), "cudaFree");
            detail::check_cuda(cudaMalloc(&d_.buf, need), "cudaMalloc");
            cap_ = need;
        }
        d_.ldim = ld;
    }
    void Attach(Int height, Int width, int colAlign, int rowAlign, T* buffer, Int ldim) {
        Empty(); viewing_ = true; d_.colAlign = colAlign; d_.rowAlign = rowAlign; d_.buf = buffer; d_.ldim = ldim; Resize(height, width);
    }
    void Empty() { if (!viewing_ && d_.buf) cudaFree(d_.buf); d_.buf = nullptr; cap_ = 0; viewing_ = false; }

    // host <-> device helpers for tests and drivers (the reference's inter-device Copy, MC_STAR.cpp:304-390)
    void SetLocalFromHost(const T* host, Int ldh) {
        if (lh_ && lw_) detail::check_cuda(cudaMemcpy2DAsync(d_.buf, d_.ldim * sizeof(T), host, ldh * sizeof(T), lh_ * sizeof(T), lw_, cudaMemcpyHostToDevice, sync_.stream_), "H2D");
    }
    void GetLocalToHost(T* host, Int ldh) const {
        if (lh_ && lw_) detail::check_cuda(cudaMemcpy2DAsync(host, ldh * sizeof(T), d_.buf, d_.ldim * sizeof(T), lh_ * sizeof(T), lw_, cudaMemcpyDeviceToHost, sync_.stream_), "D2H");
        detail::check_cuda(cudaStreamSynchronize(sync_.stream_), "sync");
    }

protected:
    AbstractDistMatrix(const El::Grid& g, int cd, int rd) : grid_(&g) {
        d_.buf = nullptr; d_.ldim = 1; d_.height = d_.width = 0; d_.colDist = cd; d_.rowDist = rd; d_.colAlign = d_.rowAlign = 0; d_.root = 0; d_.reserved = 0;
    }
    const El::Grid* grid_;
    hyd_dm d_;
    Int colShift_ = 0, rowShift_ = 0, lh_ = 0, lw_ = 0;
    size_t cap_ = 0;
    bool viewing_ = false;
    SyncInfoGPU sync_;
};

// B = A between any two supported distributions (DistMatrix::operator= tables,
// src/core/DistMatrix/ElementMatrix/*.cpp -> copy::* primitives)
template <typename T>
void Copy(const AbstractDistMatrix<T>& A, AbstractDistMatrix<T>& B) {
    B.Resize(A.Height(), A.Width());
    detail::check(detail::Sfx<T>::redist(B.Grid().handle(), A.desc(), B.desc(), 0, B.Stream()));
}
// B = A^T (Transpose.hpp:318-356); conjugate is a no-op for real T
template <typename T>
void Transpose(const AbstractDistMatrix<T>& A, AbstractDistMatrix<T>& B, bool /*conjugate*/ = false) {
    B.Resize(A.Width(), A.Height());
    detail::check(detail::Sfx<T>::redist(B.Grid().handle(), A.desc(), B.desc(), 1, B.Stream()));
}
template <typename T>
void AxpyContract(T alpha, const AbstractDistMatrix<T>& A, AbstractDistMatrix<T>& B) {  // AxpyContract.hpp:483-544
    detail::check(detail::Sfx<T>::contract(B.Grid().handle(), alpha, A.desc(), B.desc(), 0, B.Stream()));
}
template <typename T>
void TransposeAxpyContract(T alpha, const AbstractDistMatrix<T>& A, AbstractDistMatrix<T>& B, bool = false) {  // TransposeAxpyContract.hpp:14-48
    detail::check(detail::Sfx<T>::contract(B.Grid().handle(), alpha, A.desc(), B.desc(), 1, B.Stream()));
}

template <typename T, Dist U = MC, Dist V = MR, DistWrap W = ELEMENT, Device D = Device::GPU>
class DistMatrix : public AbstractDistMatrix<T> {
    static_assert(W == ELEMENT && D == Device::GPU, "only ELEMENT-wrapped GPU matrices are on the Gemm path");
public:
    explicit DistMatrix(const El::Grid& g) : AbstractDistMatrix<T>(g, U, V) { this->Resize(0, 0); }
    DistMatrix(Int height, Int width, const El::Grid& g) : AbstractDistMatrix<T>(g, U, V) { this->Resize(height, width); }
    template <Dist U2, Dist V2>
    DistMatrix& operator=(const DistMatrix<T, U2, V2, W, D>& A) { Copy<T>(A, *this); return *this; }
    DistMatrix& operator=(const DistMatrix& A) { Copy<T>(A, *this); return *this; }
};

// C := alpha op(A) op(B) + beta C   (level3.hpp:66-84; Gemm.cpp:278-307)
template <typename T>
void Gemm(Orientation orientA, Orientation orientB, T alpha, const AbstractDistMatrix<T>& A, const AbstractDistMatrix<T>& B,
          T beta, AbstractDistMatrix<T>& C, GemmAlgorithm alg = GEMM_DEFAULT) {
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

// local column-major GEMM on raw device buffers (gpu_blas::Gemm, include/hydrogen/blas/GPU_BLAS_impl.hpp:631-646)
namespace gpu_blas {
template <typename T>
void Gemm(Orientation tA, Orientation tB, Int m, Int n, Int k, T alpha, const T* A, Int lda, const T* B, Int ldb, T beta, T* C,
          Int ldc, const SyncInfoGPU& si) {
    detail::check(detail::Sfx<T>::lgemm(tA != NORMAL, tB != NORMAL, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, si.stream_));
}
}  // namespace gpu_blas

}  // namespace El
#endif  // EL_B200_SHIM_HPP

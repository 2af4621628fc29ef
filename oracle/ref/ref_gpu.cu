// ref_gpu.cu -- extern "C" handles onto the REFERENCE's own GPU local layer (seam S2), compiled by nvcc for sm_100a
// from the sources where they lie under /root/reference (never copied); built by oracle/Makefile into
// oracle/_ref/libref_gpu.so together with
//   src/hydrogen/blas/gpu/{Copy,Axpy,Transpose,Fill,Hadamard,Scale}.cu   the reference's CUDA kernels
//   src/core/imports/{cuda,cublas,cub}.cpp                               GPUManager, the cuBLAS wrappers
// and, through this translation unit, the header-only code above them:
//   include/hydrogen/blas/GPU_BLAS_impl.hpp     gpu_blas::{Gemm -> cublas?gemm, Copy -> cublas?geam / copy kernels,
//                                               Axpy, Scale}
//   include/El/blas_like/level1/Copy/util.hpp   the GPU pack/unpack loops (:232-359 cudaMemcpy2DAsync per portion,
//                                               :363-722 InterleaveMatrix -> copy_2d_kernel per portion)
//   include/El/blas_like/level1/Axpy/util.hpp   InterleaveMatrixUpdate (GPU) -> axpy_2d kernel
// TEST / MEASUREMENT INFRASTRUCTURE ONLY: this is "the recompiled reference" the product is compared with on the
// GPU box -- bit-exact parity of every copy (tests/test_gpu_ref_kernels.py, tools/ref_gpu_check.cpp, tools/ref_gpu_check_level1.cpp) and a
// GPU-vs-GPU baseline (the timing half of tools/ref_gpu_check.cpp).  The product never links or loads it.
// One change is forced by the toolkit: cudaPointerAttributes::memoryType (CUDA <= 10) is named ::type since CUDA 11;
// the Makefile passes -DmemoryType=type (include/hydrogen/device/gpu/CUDA.hpp:128).
#include <El/hydrogen_config.h>
#include <hydrogen/Device.hpp>
#include <hydrogen/SyncInfo.hpp>
#include <hydrogen/blas/GPU_BLAS.hpp>
#include <hydrogen/blas/gpu/Fill.hpp>
#include <hydrogen/blas/gpu/Hadamard.hpp>

#define REF_GPU
#include "ref_prelude.hpp"

#include <El/core/indexing/decl.hpp>
#include <El/core/indexing/impl.hpp>
#include <El/blas_like/level1/Copy/util.hpp>
#include <El/blas_like/level1/Axpy/util.hpp>

#include <cstdint>
#include <cstdio>
#include <exception>

namespace El {
void DestroyPinnedHostMemoryPool() {}  // src/core/MemoryPool.cpp:25 (needs El-lite.hpp); only FinalizeCUDA calls it
}

using namespace El;
typedef SyncInfo<Device::GPU> SI;

namespace {
char g_err[512] = "";
cudaEvent_t g_event = nullptr;
int fail(const std::exception& e) {
    std::snprintf(g_err, sizeof(g_err), "%s", e.what());
    return 1;
}
SI si(void* stream) { return SI{(cudaStream_t)stream, g_event}; }
hydrogen::TransposeMode tmode(int t) { return t ? hydrogen::TransposeMode::TRANSPOSE : hydrogen::TransposeMode::NORMAL; }
}  // namespace

#define REF_TRY(...)                 \
    try {                            \
        __VA_ARGS__;                 \
        return 0;                    \
    } catch (const std::exception& e) { \
        return fail(e);              \
    }

extern "C" {

const char* ref_gpu_last_error() { return g_err; }

// GPUManager::Create + cublas::Initialize (src/core/imports/cuda.cpp:66-104, cublas.cpp:21-28)
int ref_gpu_init(int device) {
    REF_TRY(GPUManager::Create(device); cublas::Initialize(); g_event = GPUManager::Event())
}
int ref_gpu_finalize() { REF_TRY(GPUManager::Destroy()) }

#define REF_GPU_DEFINE(sfx, T)                                                                                        \
    /* gpu_blas::Gemm -> cublas?gemm (GPU_BLAS_impl.hpp:631-646, cublas.cpp:130-148): what El::Gemm's local */       \
    /* product is on Device::GPU (Gemm.cpp:152-155) */                                                               \
    int ref_gpu_gemm_##sfx(int tA, int tB, int64_t m, int64_t n, int64_t k, T alpha, const T* A, int64_t lda,       \
                           const T* B, int64_t ldb, T beta, T* C, int64_t ldc, void* s) {                            \
        REF_TRY(gpu_blas::Gemm(tmode(tA), tmode(tB), (Int)m, (Int)n, (Int)k, alpha, A, (Int)lda, B, (Int)ldb, beta, C,    \
                               (Int)ldc, si(s)))                                                                     \
    }                                                                                                                \
    /* gpu_blas::Copy(mode, m, n, A, lda, B, ldb) -> cublas?geam (GPU_BLAS_impl.hpp:558-567): local Transpose */     \
    int ref_gpu_copy_##sfx(int trans, int64_t m, int64_t n, const T* A, int64_t lda, T* B, int64_t ldb, void* s) {  \
        REF_TRY(gpu_blas::Copy(tmode(trans), (Int)m, (Int)n, A, (Int)lda, B, (Int)ldb, si(s)))                         \
    }                                                                                                                \
    int ref_gpu_axpy2d_##sfx(int64_t m, int64_t n, T alpha, const T* A, int64_t lda, T* B, int64_t ldb, void* s) {  \
        REF_TRY(gpu_blas::Axpy((Int)m, (Int)n, alpha, A, (Int)lda, B, (Int)ldb, si(s)))                             \
    }                                                                                                                \
    /* gpu_blas::Axpy(TransposeMode, m, n, ...): B (m x n) += alpha op(A) -> cublas?geam (GPU_BLAS_impl.hpp:110-132,530-540) */ \
    int ref_gpu_axpy_trans_##sfx(int trans, int64_t m, int64_t n, T alpha, const T* A, int64_t lda, T* B, int64_t ldb,  \
                                 void* s) {                                                                          \
        REF_TRY(gpu_blas::Axpy(tmode(trans), (Int)m, (Int)n, alpha, A, (Int)lda, B, (Int)ldb, si(s)))               \
    }                                                                                                                \
    int ref_gpu_scale2d_##sfx(int64_t m, int64_t n, T alpha, T* A, int64_t lda, void* s) {                          \
        REF_TRY(gpu_blas::Scale((Int)m, (Int)n, alpha, A, (Int)lda, si(s)))                                         \
    }                                                                                                                \
    int ref_gpu_fill_##sfx(int64_t m, int64_t n, T alpha, T* A, int64_t lda, void* s) {                             \
        REF_TRY(hydrogen::Fill_GPU_impl((size_t)m, (size_t)n, alpha, A, (size_t)lda, (cudaStream_t)s))              \
    }                                                                                                                \
    int ref_gpu_hadamard_##sfx(int64_t m, int64_t n, const T* A, int64_t rsA, int64_t lda, const T* B, int64_t rsB, \
                               int64_t ldb, T* C, int64_t rsC, int64_t ldc, void* s) {                              \
        REF_TRY(hydrogen::Hadamard_GPU_impl((size_t)m, (size_t)n, A, (size_t)rsA, (size_t)lda, B, (size_t)rsB,      \
                                            (size_t)ldb, C, (size_t)rsC, (size_t)ldc, (cudaStream_t)s))             \
    }                                                                                                                \
    /* copy::util on Device::GPU */                                                                                  \
    int ref_gpu_interleave_##sfx(int64_t h, int64_t w, const T* A, int64_t csA, int64_t rsA, T* B, int64_t csB,     \
                                 int64_t rsB, void* s) {                                                             \
        REF_TRY(copy::util::InterleaveMatrix((Int)h, (Int)w, A, (Int)csA, (Int)rsA, B, (Int)csB, (Int)rsB, si(s)))  \
    }                                                                                                                \
    int ref_gpu_interleave_update_##sfx(T alpha, int64_t h, int64_t w, const T* A, int64_t csA, int64_t rsA, T* B,  \
                                        int64_t csB, int64_t rsB, void* s) {                                         \
        REF_TRY(axpy::util::InterleaveMatrixUpdate(alpha, (Int)h, (Int)w, A, (Int)csA, (Int)rsA, B, (Int)csB,       \
                                                   (Int)rsB, si(s)))                                                 \
    }                                                                                                                \
    int ref_gpu_row_strided_pack_##sfx(int64_t h, int64_t w, int64_t al, int64_t st, const T* A, int64_t ld, T* Bp, \
                                       int64_t portion, void* s) {                                                   \
        REF_TRY(copy::util::RowStridedPack((Int)h, (Int)w, (Int)al, (Int)st, A, (Int)ld, Bp, (Int)portion, si(s)))  \
    }                                                                                                                \
    int ref_gpu_row_strided_unpack_##sfx(int64_t h, int64_t w, int64_t al, int64_t st, const T* Ap,                 \
                                         int64_t portion, T* B, int64_t ld, void* s) {                               \
        REF_TRY(copy::util::RowStridedUnpack((Int)h, (Int)w, (Int)al, (Int)st, Ap, (Int)portion, B, (Int)ld, si(s))) \
    }                                                                                                                \
    int ref_gpu_partial_row_strided_pack_##sfx(int64_t h, int64_t w, int64_t al, int64_t st, int64_t uni,           \
                                               int64_t part, int64_t rankPart, int64_t shiftA, const T* A,           \
                                               int64_t ld, T* Bp, int64_t portion, void* s) {                        \
        REF_TRY(copy::util::PartialRowStridedPack((Int)h, (Int)w, (Int)al, (Int)st, (Int)uni, (Int)part,            \
                                                  (Int)rankPart, (Int)shiftA, A, (Int)ld, Bp, (Int)portion, si(s))) \
    }                                                                                                                \
    int ref_gpu_partial_row_strided_unpack_##sfx(int64_t h, int64_t w, int64_t al, int64_t st, int64_t uni,         \
                                                 int64_t part, int64_t rankPart, int64_t shiftB, const T* Ap,        \
                                                 int64_t portion, T* B, int64_t ld, void* s) {                       \
        REF_TRY(copy::util::PartialRowStridedUnpack((Int)h, (Int)w, (Int)al, (Int)st, (Int)uni, (Int)part,          \
                                                    (Int)rankPart, (Int)shiftB, Ap, (Int)portion, B, (Int)ld,        \
                                                    si(s)))                                                          \
    }                                                                                                                \
    int ref_gpu_col_strided_pack_##sfx(int64_t h, int64_t w, int64_t al, int64_t st, const T* A, int64_t ld, T* Bp, \
                                       int64_t portion, void* s) {                                                   \
        REF_TRY(copy::util::ColStridedPack<T, Device::GPU>((Int)h, (Int)w, (Int)al, (Int)st, A, (Int)ld, Bp,        \
                                                           (Int)portion, si(s)))                                     \
    }                                                                                                                \
    int ref_gpu_col_strided_unpack_##sfx(int64_t h, int64_t w, int64_t al, int64_t st, const T* Ap,                 \
                                         int64_t portion, T* B, int64_t ld, void* s) {                               \
        REF_TRY(copy::util::ColStridedUnpack<T, Device::GPU>((Int)h, (Int)w, (Int)al, (Int)st, Ap, (Int)portion, B, \
                                                             (Int)ld, si(s)))                                        \
    }                                                                                                                \
    int ref_gpu_partial_col_strided_pack_##sfx(int64_t h, int64_t w, int64_t al, int64_t st, int64_t uni,           \
                                               int64_t part, int64_t rankPart, int64_t shiftA, const T* A,           \
                                               int64_t ld, T* Bp, int64_t portion, void* s) {                        \
        REF_TRY(copy::util::PartialColStridedPack<T, Device::GPU>((Int)h, (Int)w, (Int)al, (Int)st, (Int)uni,       \
                                                                  (Int)part, (Int)rankPart, (Int)shiftA, A,          \
                                                                  (Int)ld, Bp, (Int)portion, si(s)))                 \
    }                                                                                                                \
    int ref_gpu_partial_col_strided_unpack_##sfx(int64_t h, int64_t w, int64_t al, int64_t st, int64_t uni,         \
                                                 int64_t part, int64_t rankPart, int64_t shiftB, const T* Ap,        \
                                                 int64_t portion, T* B, int64_t ld, void* s) {                       \
        REF_TRY(copy::util::PartialColStridedUnpack<T, Device::GPU>((Int)h, (Int)w, (Int)al, (Int)st, (Int)uni,     \
                                                                    (Int)part, (Int)rankPart, (Int)shiftB, Ap,       \
                                                                    (Int)portion, B, (Int)ld, si(s)))                \
    }                                                                                                                \
    int ref_gpu_strided_pack_##sfx(int64_t h, int64_t w, int64_t cal, int64_t cst, int64_t ral, int64_t rst,        \
                                   const T* A, int64_t ld, T* Bp, int64_t portion, void* s) {                        \
        REF_TRY(copy::util::StridedPack<T, Device::GPU>((Int)h, (Int)w, (Int)cal, (Int)cst, (Int)ral, (Int)rst, A,  \
                                                        (Int)ld, Bp, (Int)portion, si(s)))                           \
    }                                                                                                                \
    int ref_gpu_strided_unpack_##sfx(int64_t h, int64_t w, int64_t cal, int64_t cst, int64_t ral, int64_t rst,      \
                                     const T* Ap, int64_t portion, T* B, int64_t ld, void* s) {                      \
        REF_TRY(copy::util::StridedUnpack<T, Device::GPU>((Int)h, (Int)w, (Int)cal, (Int)cst, (Int)ral, (Int)rst,   \
                                                          Ap, (Int)portion, B, (Int)ld, si(s)))                      \
    }

REF_GPU_DEFINE(f64, double)
REF_GPU_DEFINE(f32, float)

}  // extern "C"

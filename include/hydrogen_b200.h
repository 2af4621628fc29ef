/* hydrogen_b200.h -- C ABI of libhydrogen_b200.so
 *
 * B200-native (sm_100a) implementation of the distributed-GEMM hot path of
 * Hydrogen/Elemental: El::Gemm on DistMatrix<T,MC,MR> (SUMMA A/B/C/Dot), the
 * element-cyclic redistributions it drives, and the local kernels below them.
 *
 * Every entry point is `extern "C"`, takes plain pointers / integers / a
 * cudaStream_t (passed as void*), returns 0 on success and non-zero on error;
 * hyd_last_error() then holds the message (the C++ shim in include/El.hpp
 * rethrows it as std::runtime_error / std::logic_error, the reference's
 * convention: include/El/core/environment/decl.hpp:196-198).
 *
 * All matrices are column-major with an explicit leading dimension, as in the
 * reference (include/El/core/Matrix/decl.hpp).  Buffers are DEVICE pointers
 * owned by the caller unless a function says "host".
 *
 * "ref:" comments cite the reference interface each function replaces
 * (paths relative to the reference repository root).
 */
#ifndef HYDROGEN_B200_H
#define HYDROGEN_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* hyd_stream_t; /* cudaStream_t */

/* ---- errors / library info ------------------------------------------------ */
const char* hyd_last_error(void);
int hyd_version(void);
/* number of kernels this library has launched in the calling process (bench.py's gpu_launches) */
long long hyd_launch_count(void);

/* ---- seam S2: local (per-rank) kernels ------------------------------------- */

/* C := alpha*op(A)*op(B) + beta*C, column-major.  transX: 0 = NORMAL, 1 = TRANSPOSE (ADJOINT == TRANSPOSE for reals).
 * ref: gpu_blas::Gemm, include/hydrogen/blas/GPU_BLAS_impl.hpp:631-646 (-> cublasDgemm/Sgemm,
 *      src/core/imports/cublas.cpp:130-148); caller El::Gemm(Matrix<T,GPU>) src/blas_like/level3/Gemm.cpp:104-157.
 *      k == 0 scales C by beta (Gemm.cpp:212-219). */
int hyd_lgemm_f64(int transA, int transB, int64_t m, int64_t n, int64_t k, double alpha,
                  const double* A, int64_t lda, const double* B, int64_t ldb, double beta,
                  double* C, int64_t ldc, hyd_stream_t stream);
int hyd_lgemm_f32(int transA, int transB, int64_t m, int64_t n, int64_t k, float alpha,
                  const float* A, int64_t lda, const float* B, int64_t ldb, float beta,
                  float* C, int64_t ldc, hyd_stream_t stream);

/* Pipe-peak probe (register-resident loops): kind 0 = FP64 DFMA, 1 = DMMA m8n8k4, 2 = DMMA m16n8k16,
 * 3 = FP32 FFMA.  Writes achieved TFLOP/s.  Measurement aid for the roofline denominator. */
int hyd_probe_peak(int kind, int iters, int ctas_per_sm, double* tflops);

#ifdef __cplusplus
}
#endif
#endif /* HYDROGEN_B200_H */

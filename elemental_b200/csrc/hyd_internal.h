// Internal declarations shared by the translation units of libhydrogen_b200.so.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <limits.h>

namespace hyd {

// Error convention: C-ABI returns non-zero and hyd_last_error() has the text;
// the C++ shim rethrows (reference: LogicError/RuntimeError, environment/decl.hpp:196-198).
int set_error(const char* fmt, ...);
int check_cuda(cudaError_t e, const char* what);
int num_sms();

typedef CUresult (*tensor_map_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                         const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                         const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                         CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
tensor_map_encode_fn tensor_map_encode();

int lgemm_f64(int tA, int tB, long long m, long long n, long long k, double alpha, const double* A,
              long long lda, const double* B, long long ldb, double beta, double* C, long long ldc,
              cudaStream_t s);
int lgemm_f64_ex(int tA, int tB, long long m, long long n, long long k, double alpha, const double* A,
                 long long lda, const double* B, long long ldb, double beta, double* C, long long ldc,
                 cudaStream_t s, int sm_reserve);
int lgemm_f32_ex(int tA, int tB, long long m, long long n, long long k, float alpha, const float* A,
                 long long lda, const float* B, long long ldb, float beta, float* C, long long ldc,
                 cudaStream_t s, int sm_reserve);
int lgemm_f32_tc(int tA, int tB, long long m, long long n, long long k, float alpha, const float* A, long long lda,
                 const float* B, long long ldb, float beta, float* C, long long ldc, cudaStream_t s, int sm_reserve);
int lgemm_f32(int tA, int tB, long long m, long long n, long long k, float alpha, const float* A,
              long long lda, const float* B, long long ldb, float beta, float* C, long long ldc,
              cudaStream_t s);
// half-precision operands / result, fp32 accumulation (lgemm_f16_tc.cu; pointers to __half / __nv_bfloat16 storage)
int lgemm_f16(int tA, int tB, long long m, long long n, long long k, float alpha, const void* A, long long lda, const void* B,
              long long ldb, float beta, void* C, long long ldc, cudaStream_t s);
int lgemm_bf16(int tA, int tB, long long m, long long n, long long k, float alpha, const void* A, long long lda, const void* B,
               long long ldb, float beta, void* C, long long ldc, cudaStream_t s);

}  // namespace hyd

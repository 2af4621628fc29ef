// Internal declarations shared by the translation units of libhydrogen_b200.so.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <limits.h>
#include <atomic>
#include <nvtx3/nvToolsExt.h>

namespace hyd {

// Error convention: C-ABI returns non-zero and hyd_last_error() has the text;
// the C++ shim rethrows (reference: LogicError/RuntimeError, environment/decl.hpp:196-198).
int set_error(const char* fmt, ...);
int check_cuda(cudaError_t e, const char* what);
int num_sms();

// NVTX ranges carrying the reference's region names (AUTO_PROFILE_REGION, include/El/core/Profiling.hpp:249-263:
// "SUMMA.NNC" src/blas_like/level3/Gemm/NN.hpp:331-334, "SUMMA.NNA" :114, "SUMMA.NNB" :251, "SUMMA.NNDot" :800,
// "Gemm_impl.GPU" src/blas_like/level3/Gemm.cpp:111), so timelines of the two builds line up.  Header-only NVTX v3:
// a no-op unless a profiler is attached.
struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
    NvtxRange(const NvtxRange&) = delete;
    NvtxRange& operator=(const NvtxRange&) = delete;
};

// Process-wide counters behind hyd_launch_count / hyd_alloc_count / hyd_sync_count: kernels this library launched, device
// allocations (cudaMalloc) it made, and host-blocking synchronisations (stream / device) it issued.  Rank-threads of the
// in-process world increment them concurrently, hence atomics.  A steady-state hyd_gemm_* / hyd_redist_* call allocates
// nothing and never blocks the host (tests/test_gpu_gemm.py counts).
extern std::atomic<long long> g_launches, g_allocs, g_syncs;
inline cudaError_t dev_malloc(void** p, size_t bytes) { ++g_allocs; return cudaMalloc(p, bytes); }
inline cudaError_t stream_sync(cudaStream_t s) { ++g_syncs; return cudaStreamSynchronize(s); }
inline cudaError_t device_sync() { ++g_syncs; return cudaDeviceSynchronize(); }

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

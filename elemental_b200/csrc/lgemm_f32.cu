// Local FP32 GEMM for sm_100a:  C := alpha * op(A) * op(B) + beta * C  (column-major).
//
// Replaces gpu_blas::Gemm<float> -> cublasSgemm (include/hydrogen/blas/GPU_BLAS_impl.hpp:277-303,
// src/core/imports/cublas.cpp:130-148).  FFMA path: exact fp32 products and fp32 accumulation,
// i.e. the same arithmetic class as cublasSgemm without tensor-op math, held to the
// 1e-5 (k-scaled) relative Frobenius tolerance of BASELINE.json.
//
// Kernel: 128 x 128 x 16 CTA tile, 256 threads, 8 x 8 register tile per thread, operands staged
// through a 3-deep cp.async (LDGSTS) shared-memory ring in [k][outer] order so that every
// fragment read is a conflict-free 16-byte LDS.  All four transpose combinations are handled by
// the loader (k-contiguous operands are transposed on their way into shared memory).
#include "hyd_internal.h"
#include <algorithm>

namespace hyd {

namespace f32 {
constexpr int BM = 128, BN = 128, BK = 16, STAGES = 3, THREADS = 256;
constexpr int LDS_PITCH = BM + 4;  // floats; +4 keeps 16-byte alignment and skews banks for the transposing store

__device__ __forceinline__ void cp_async4(void* smem, const void* gmem, bool pred) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    const int sz = pred ? 4 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(s), "l"(gmem), "r"(sz));
}
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem, bool pred) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    const int sz = pred ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(gmem), "r"(sz));
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;"); }
template <int N> __device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N)); }

// Load one BK x 128 operand slab into smem[k][o].
//  OUTER_CONTIG: global element (o,k) at o + k*ld   -> 16-byte copies along o when aligned
//  else        : global element (o,k) at k + o*ld   -> 4-byte copies (transposing)
template <bool OUTER_CONTIG>
__device__ __forceinline__ void load_slab(float (*sm)[LDS_PITCH], const float* __restrict__ g, long long ld, int o0,
                                          int k0, int O, int K, bool vec_ok) {
    if (OUTER_CONTIG) {
        if (vec_ok) {
            // 16 k-rows x 32 float4 = 512 copies, 2 per thread
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const int e = threadIdx.x + i * THREADS;
                const int kk = e >> 5, o4 = (e & 31) * 4;
                const bool p = (k0 + kk < K) && (o0 + o4 + 3 < O);
                const float* src = g + (long long)(o0 + o4) + (long long)(k0 + kk) * ld;
                if (p || (k0 + kk >= K) || (o0 + o4 >= O)) {
                    cp_async16(&sm[kk][o4], p ? src : g, p);
                } else {  // ragged edge inside a float4
#pragma unroll
                    for (int q = 0; q < 4; ++q) cp_async4(&sm[kk][o4 + q], (o0 + o4 + q < O) ? src + q : g, o0 + o4 + q < O);
                }
            }
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int e = threadIdx.x + i * THREADS;
                const int kk = e >> 7, o = e & 127;
                const bool p = (k0 + kk < K) && (o0 + o < O);
                cp_async4(&sm[kk][o], p ? g + (long long)(o0 + o) + (long long)(k0 + kk) * ld : g, p);
            }
        }
    } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int e = threadIdx.x + i * THREADS;
            const int kk = e & 15, o = e >> 4;
            const bool p = (k0 + kk < K) && (o0 + o < O);
            cp_async4(&sm[kk][o], p ? g + (long long)(k0 + kk) + (long long)(o0 + o) * ld : g, p);
        }
    }
}

template <bool A_OC, bool B_OC>
__global__ void __launch_bounds__(THREADS, 2)
sgemm_kernel(int M, int N, int K, float alpha, const float* __restrict__ A, long long lda,
             const float* __restrict__ B, long long ldb, float beta, float* __restrict__ C, long long ldc,
             int tiles_m, int vecA, int vecB) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    typedef float Slab[BK][LDS_PITCH];
    Slab* sA = reinterpret_cast<Slab*>(smem_raw);
    Slab* sB = sA + STAGES;
    const int tm = blockIdx.x % tiles_m, tn = blockIdx.x / tiles_m;
    const int m0 = tm * BM, n0 = tn * BN;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;  // 16 x 16 threads, each 8 x 8 (two 4-wide halves)
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
    const int nk = (K + BK - 1) / BK;
#pragma unroll
    for (int st = 0; st < STAGES - 1; ++st) {
        if (st < nk) {
            load_slab<A_OC>(sA[st], A, lda, m0, st * BK, M, K, vecA);
            load_slab<B_OC>(sB[st], B, ldb, n0, st * BK, N, K, vecB);
        }
        cp_commit();
    }
    for (int kb = 0; kb < nk; ++kb) {
        cp_wait<STAGES - 2>();
        __syncthreads();
        const int nxt = kb + STAGES - 1;
        if (nxt < nk) {
            load_slab<A_OC>(sA[nxt % STAGES], A, lda, m0, nxt * BK, M, K, vecA);
            load_slab<B_OC>(sB[nxt % STAGES], B, ldb, n0, nxt * BK, N, K, vecB);
        }
        cp_commit();
        const float (*a)[LDS_PITCH] = sA[kb % STAGES];
        const float (*b)[LDS_PITCH] = sB[kb % STAGES];
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
            const float4 a0 = *reinterpret_cast<const float4*>(&a[kk][tx * 4]);
            const float4 a1 = *reinterpret_cast<const float4*>(&a[kk][64 + tx * 4]);
            const float4 b0 = *reinterpret_cast<const float4*>(&b[kk][ty * 4]);
            const float4 b1 = *reinterpret_cast<const float4*>(&b[kk][64 + ty * 4]);
            const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
    }
    cp_wait<0>();
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int col = n0 + (j < 4 ? ty * 4 + j : 64 + ty * 4 + (j - 4));
        if (col >= N) continue;
        float* cp = C + (long long)col * ldc;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int row = m0 + (i < 4 ? tx * 4 + i : 64 + tx * 4 + (i - 4));
            if (row >= M) continue;
            const float v = alpha * acc[i][j];
            cp[row] = (beta == 0.f) ? v : v + beta * cp[row];
        }
    }
}

__global__ void sscale_kernel(long long M, long long N, float beta, float* __restrict__ C, long long ldc) {
    const long long total = M * N;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
        const long long j = e / M, i = e - j * M;
        float* p = C + i + j * ldc;
        *p = (beta == 0.f) ? 0.f : beta * *p;
    }
}
}  // namespace f32

constexpr long long kSplitTf32MaxK = 200000;   // HYD_LGEMM_F32_SPLIT_TF32_MAX_K in include/hydrogen_b200.h

int lgemm_f32(int tA, int tB, long long m, long long n, long long k, float alpha, const float* A, long long lda,
              const float* B, long long ldb, float beta, float* C, long long ldc, cudaStream_t s) {
    return lgemm_f32_ex(tA, tB, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, s, 0);
}

int lgemm_f32_ex(int tA, int tB, long long m, long long n, long long k, float alpha, const float* A, long long lda,
                 const float* B, long long ldb, float beta, float* C, long long ldc, cudaStream_t s, int sm_reserve) {
    using namespace f32;
    if (m < 0 || n < 0 || k < 0) return set_error("hyd_lgemm_f32: negative dimension");
    if (m == 0 || n == 0) return 0;
    if (m > INT32_MAX || n > INT32_MAX || k > INT32_MAX) return set_error("hyd_lgemm_f32: dimension exceeds 2^31-1");
    const long long a_rows = tA ? k : m, b_rows = tB ? n : k;
    if (lda < std::max<long long>(a_rows, 1) || ldb < std::max<long long>(b_rows, 1) || ldc < m)
        return set_error("hyd_lgemm_f32: leading dimension too small");
    if (k == 0 || alpha == 0.f) {
        if (beta != 1.f) {
            sscale_kernel<<<num_sms() * 8, 256, 0, s>>>(m, n, beta, C, ldc);
            return check_cuda(cudaGetLastError(), "sscale_kernel");
        }
        return 0;
    }
    // The split-TF32 kernel's error grows in proportion to k (the tensor core's fp32 accumulator truncates: measured relative
    // Frobenius error 7.4e-6 at k = 1024, 5.8e-5 at k = 8192, i.e. ~7e-9 * k) while the stated tolerance is 1e-5 * sqrt(k); the two
    // cross near k = 2e6 by that fit, near 2e5 with a margin of three.  Above kSplitTf32MaxK one launch falls back to the exact-
    // product FFMA kernel.  (The SUMMA engines never get there: a local product's k is at most the panel width, or k / P in Dot.)
    if (k <= kSplitTf32MaxK) {   // large, TMA-compatible operands: split-TF32 on the tcgen05 tensor cores (lgemm_f32_tc.cu)
        const int rc = lgemm_f32_tc(tA, tB, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, s, sm_reserve);
        if (rc >= 0) return rc;
    }
    const bool aoc = !tA, boc = tB != 0;  // outer-contiguous?
    const int vecA = aoc && ((uintptr_t)A % 16 == 0) && (lda % 4 == 0);
    const int vecB = boc && ((uintptr_t)B % 16 == 0) && (ldb % 4 == 0);
    const int tiles_m = (int)((m + BM - 1) / BM), tiles_n = (int)((n + BN - 1) / BN);
    const size_t smem = 2 * STAGES * sizeof(float) * BK * LDS_PITCH;
    static bool attr = false;
    if (!attr) {
        cudaFuncSetAttribute(sgemm_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        cudaFuncSetAttribute(sgemm_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        cudaFuncSetAttribute(sgemm_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        cudaFuncSetAttribute(sgemm_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        attr = true;
    }
    const unsigned grid = (unsigned)(tiles_m * (long long)tiles_n);
#define HYD_SGEMM(AO, BO) sgemm_kernel<AO, BO><<<grid, THREADS, smem, s>>>((int)m, (int)n, (int)k, alpha, A, lda, B, ldb, beta, C, ldc, tiles_m, vecA, vecB)
    if (aoc && boc) HYD_SGEMM(true, true);
    else if (aoc) HYD_SGEMM(true, false);
    else if (boc) HYD_SGEMM(false, true);
    else HYD_SGEMM(false, false);
#undef HYD_SGEMM
    return check_cuda(cudaGetLastError(), "sgemm_kernel");
}

}  // namespace hyd

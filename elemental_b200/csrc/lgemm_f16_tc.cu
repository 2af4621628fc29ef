// Local half-precision GEMM on the 5th-generation tensor cores: fp16 or bf16 operands and result,
// fp32 accumulation in TMEM, tcgen05.mma kind::f16.
//
//   C := alpha * op(A) * op(B) + beta * C     (column-major, all four transpose combinations; alpha, beta in fp32)
//
// Reference: Gemm on Matrix<gpu_half_type, Device::GPU> (src/blas_like/level3/Gemm.cpp:460-467, HYDROGEN_GPU_USE_FP16)
// -> gpu_blas::Gemm<__half> -> cublasHgemm (src/core/imports/cublas.cpp:213 ADD_GEMM_IMPL(__half, H)); the reference has
// no bf16 type, bf16 is offered because it is the B200's native training format.
//
// STATUS: parity-tested on a B200 since round 2 (tests/test_gpu_local.py::test_lgemm_half, tests/test_gpu_half.py); the local
// product of hyd_gemm_{f16,bf16} (summa_half.cu).
//
// Structure = lgemm_f32_tc.cu without the splitter stage (one persistent CTA per SM, warp-specialised):
//   warp 0     TMA producer: 128 x 64 tile of A and BN x 64 tile of B per stage (cp.async.bulk.tensor.2d, SWIZZLE_128B,
//              both operand orientations staged in their native layout)
//   warp 1     MMA issuer (one elected lane): four tcgen05.mma.kind::f16 (128 x BN x 16) per stage into one of two TMEM
//              accumulators; tcgen05.commit releases the stage / publishes the accumulator
//   warp 2     TMEM allocator
//   warps 4-7  epilogue: tcgen05.ld -> registers, C = alpha*acc + beta*C, one rounding to the half type, coalesced stores
// Shared-memory descriptors (cute/atom/mma_traits_sm100.hpp: canonical UMMA layouts, in 16-byte units):
//   K-major  (k contiguous):     ((8,n),2):((8,SBO),1)   lines of 64 elements = 128 B along k, 8-line atoms SBO = 1024 B apart;
//                                 one MMA k-step (16 elements) = +32 B inside the swizzled line
//   MN-major (outer contiguous): ((8,n),(8,k)):((1,LBO),(8,SBO))  lines of 64 elements along the outer index, one line per
//                                 k; 8-k-line atoms SBO = 1024 B apart, 64-element chunks of the outer index LBO = BK*128 B
//                                 apart (one TMA box each); one MMA k-step (16 lines) = +2048 B
#include "ptx.cuh"
#include "hyd_internal.h"
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <algorithm>

namespace hyd {
namespace tc16 {

constexpr int BM = 128, BK = 64;
enum { MT = 0, KT = 1 };
enum { F16 = 0, BF16 = 1 };   // UMMA::F16F32Format

template <int BN>
struct Cfg {
    static constexpr int kBytesA = BM * BK * 2, kBytesB = BN * BK * 2;
    static constexpr int kStageBytes = kBytesA + kBytesB;
    static constexpr int kStages = (196608 / kStageBytes) < 6 ? (196608 / kStageBytes) : 6;
    static constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 256;
    static constexpr int kTmemCols = 2 * BN;
    static constexpr int kThreads = 8 * 32;
};

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_f16(uint32_t d_tmem, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(d_tmem),
        "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo, uint32_t layout_type) {
    return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) |
           (1ull << 46) | ((uint64_t)layout_type << 61);
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

template <int E> struct Half;
template <> struct Half<F16> {
    typedef __half type;
    static __device__ __forceinline__ float to_f32(__half x) { return __half2float(x); }
    static __device__ __forceinline__ __half from_f32(float x) { return __float2half_rn(x); }
};
template <> struct Half<BF16> {
    typedef __nv_bfloat16 type;
    static __device__ __forceinline__ float to_f32(__nv_bfloat16 x) { return __bfloat162float(x); }
    static __device__ __forceinline__ __nv_bfloat16 from_f32(float x) { return __float2bfloat16_rn(x); }
};

template <int E, int TA, int TB, int BN>
__global__ void __launch_bounds__(Cfg<BN>::kThreads, 1)
hgemm_tc_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                typename Half<E>::type* __restrict__ C, long long ldc, int M, int N, int K, float alpha, float beta, int tiles_m,
                int tiles_n, int group_m) {
    typedef Cfg<BN> G;
    typedef typename Half<E>::type HT;
    constexpr int kStages = G::kStages, kStageBytes = G::kStageBytes, kBytesA = G::kBytesA;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
    const uint32_t bar_base = smem_base + kStages * kStageBytes;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };                       // TMA landed
    auto empty_bar = [&](int s) { return bar_base + 8u * (kStages + s); };          // MMAs have read the stage
    auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * kStages + a); };      // accumulator complete
    auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * kStages + 2 + a); }; // accumulator drained
    const uint32_t tmem_slot = bar_base + 8u * (2 * kStages + 4);
    volatile uint32_t* tmem_slot_gen = reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(full_bar(s), 1);
            mbar_init(empty_bar(s), 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(tfull_bar(a), 1);
            mbar_init(tempty_bar(a), 4);
        }
        mbar_fence_init();
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"((uint32_t)G::kTmemCols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_gen;

    const int num_kb = (K + BK - 1) / BK;   // the K tail is zero-filled by TMA (out-of-bounds box elements read as 0)
    const int num_tiles = tiles_m * tiles_n;
    const int tiles_per_group = group_m * tiles_n;
    auto tile_coord = [&](int tile, int& tm, int& tn) {
        const int grp = tile / tiles_per_group;
        const int first_m = grp * group_m;
        const int gm = min(group_m, tiles_m - first_m);
        const int in = tile - grp * tiles_per_group;
        tm = first_m + in % gm;
        tn = in / gm;
    };

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            tma_prefetch_desc(&mapA);
            tma_prefetch_desc(&mapB);
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                int tm, tn;
                tile_coord(tile, tm, tn);
                const int m0 = tm * BM, n0 = tn * BN;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(empty_bar(stage), phase ^ 1);
                    const uint32_t sA = smem_base + stage * kStageBytes, sB = sA + kBytesA;
                    mbar_expect_tx(full_bar(stage), kStageBytes);
                    if (TA == MT) {  // boxes of [BK k-lines][64 outer elements]: 128-byte lines along the outer index
#pragma unroll
                        for (int b = 0; b < BM / 64; ++b) tma_load_2d(sA + b * (BK * 128), &mapA, full_bar(stage), m0 + b * 64, kb * BK);
                    } else {
                        tma_load_2d(sA, &mapA, full_bar(stage), kb * BK, m0);
                    }
                    if (TB == MT) {
#pragma unroll
                        for (int b = 0; b < BN / 64; ++b) tma_load_2d(sB + b * (BK * 128), &mapB, full_bar(stage), n0 + b * 64, kb * BK);
                    } else {
                        tma_load_2d(sB, &mapB, full_bar(stage), kb * BK, n0);
                    }
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            // instruction descriptor (UMMA::InstrDescriptor): D = F32 (bits 4-5), A / B format (bits 7-9 / 10-12),
            // A / B major (bits 15 / 16: 1 = MN), N >> 3 (bits 17-22), M >> 4 (bits 24-28)
            const uint32_t idesc = (1u << 4) | ((uint32_t)E << 7) | ((uint32_t)E << 10) | ((TA == MT ? 1u : 0u) << 15) |
                                   ((TB == MT ? 1u : 0u) << 16) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
            const uint32_t lboA = (TA == MT) ? BK * 128 : 16, lboB = (TB == MT) ? BK * 128 : 16;
            const uint32_t stepA = (TA == MT) ? 2048 : 32, stepB = (TB == MT) ? 2048 : 32;
            int stage = 0, acc = 0;
            uint32_t phase = 0, acc_phase = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                mbar_wait(tempty_bar(acc), acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(full_bar(stage), phase);
                    tc_fence_after();
                    const uint32_t sA = smem_base + stage * kStageBytes, sB = sA + kBytesA;
#pragma unroll
                    for (int ks = 0; ks < BK / 16; ++ks) {
                        const uint64_t da = make_desc(sA + ks * stepA, lboA, 1024, 2u /*SWIZZLE_128B*/);
                        const uint64_t db = make_desc(sB + ks * stepB, lboB, 1024, 2u);
                        tc_mma_f16(d_tmem, da, db, idesc, (kb | ks) != 0);
                    }
                    tc_commit(empty_bar(stage));
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
                tc_commit(tfull_bar(acc));
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else if (warp >= 4) {
        // ===================== epilogue =====================
        const int q = warp & 3;  // TMEM lane quarter this warp may read
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            int tm, tn;
            tile_coord(tile, tm, tn);
            mbar_wait(tfull_bar(acc), acc_phase);
            tc_fence_after();
            const int row = tm * BM + q * 32 + lane;
            const bool rd = (beta != 0.f);
#pragma unroll 1
            for (int c0 = 0; c0 < BN; c0 += 32) {
                uint32_t r[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN + c0, r);
                float old[32];
                const int col0 = tn * BN + c0;
                HT* cp = C + (long long)col0 * ldc + row;
                if (rd && row < M) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) old[j] = (col0 + j < N) ? Half<E>::to_f32(cp[(long long)j * ldc]) : 0.f;
                }
                tmem_ld_wait();
                if (c0 + 32 >= BN) {  // accumulator fully read: hand it back before the stores
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(tempty_bar(acc));
                }
                if (row < M) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        if (col0 + j < N) {
                            float v = alpha * __uint_as_float(r[j]);
                            if (rd) v += beta * old[j];
                            cp[(long long)j * ldc] = Half<E>::from_f32(v);
                        }
                    }
                }
            }
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)G::kTmemCols));
    }
}

// Shapes / operands the tensor-core kernel does not take (small, unaligned): one thread per element of C, fp32 accumulation.
template <int E>
__global__ void __launch_bounds__(256)
hgemm_generic_kernel(int tA, int tB, int M, int N, int K, float alpha, const typename Half<E>::type* __restrict__ A, long long lda,
                     const typename Half<E>::type* __restrict__ B, long long ldb, float beta, typename Half<E>::type* __restrict__ C,
                     long long ldc) {
    const long long total = (long long)M * N;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
        const int i = (int)(e % M), j = (int)(e / M);
        float acc = 0.f;
        for (int k = 0; k < K; ++k) {
            const float a = Half<E>::to_f32(tA ? A[k + (long long)i * lda] : A[i + (long long)k * lda]);
            const float b = Half<E>::to_f32(tB ? B[j + (long long)k * ldb] : B[k + (long long)j * ldb]);
            acc = fmaf(a, b, acc);
        }
        typename Half<E>::type* cp = C + i + (long long)j * ldc;
        float v = alpha * acc;
        if (beta != 0.f) v += beta * Half<E>::to_f32(*cp);
        *cp = Half<E>::from_f32(v);
    }
}

static bool make_map_h(CUtensorMap* map, int E, int type, const void* base, long long ld, int outer, int K, int box_outer) {
    cuuint64_t dims[2], strides[1];
    cuuint32_t box[2], estr[2] = {1, 1};
    if (type == MT) { dims[0] = outer; dims[1] = K; box[0] = 64; box[1] = BK; }
    else            { dims[0] = K; dims[1] = outer; box[0] = BK; box[1] = box_outer; }
    strides[0] = (cuuint64_t)ld * 2;
    if (!tensor_map_encode()) return false;
    const CUresult r = tensor_map_encode()(map, E == BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2,
                                           const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS;
}

template <int E, int TA, int TB, int BN>
static cudaError_t launch(const CUtensorMap& ma, const CUtensorMap& mb, void* C, long long ldc, int M, int N, int K, float alpha,
                          float beta, cudaStream_t s) {
    typedef Cfg<BN> G;
    static bool attr_set = false;
    auto kern = hgemm_tc_kernel<E, TA, TB, BN>;
    if (!attr_set) {
        const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, G::kSmemBytes);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    const int tiles_m = (M + BM - 1) / BM, tiles_n = (N + BN - 1) / BN;
    const int grid = max(1, min(tiles_m * tiles_n, num_sms()));
    kern<<<grid, G::kThreads, G::kSmemBytes, s>>>(ma, mb, static_cast<typename Half<E>::type*>(C), ldc, M, N, K, alpha, beta, tiles_m,
                                                  tiles_n, 16);
    return cudaGetLastError();
}

template <int E, int BN>
static cudaError_t dispatch(int typeA, int typeB, const CUtensorMap& ma, const CUtensorMap& mb, void* C, long long ldc, int M, int N,
                            int K, float alpha, float beta, cudaStream_t s) {
    if (typeA == MT && typeB == MT) return launch<E, MT, MT, BN>(ma, mb, C, ldc, M, N, K, alpha, beta, s);
    if (typeA == MT && typeB == KT) return launch<E, MT, KT, BN>(ma, mb, C, ldc, M, N, K, alpha, beta, s);
    if (typeA == KT && typeB == MT) return launch<E, KT, MT, BN>(ma, mb, C, ldc, M, N, K, alpha, beta, s);
    return launch<E, KT, KT, BN>(ma, mb, C, ldc, M, N, K, alpha, beta, s);
}

template <int E>
static int lgemm_half(int tA, int tB, long long m, long long n, long long k, float alpha, const void* A, long long lda, const void* B,
                      long long ldb, float beta, void* C, long long ldc, cudaStream_t s) {
    typedef typename Half<E>::type HT;
    const char* who = E == BF16 ? "hyd_lgemm_bf16" : "hyd_lgemm_f16";
    if (m < 0 || n < 0 || k < 0) return set_error("%s: negative dimension", who);
    if (m == 0 || n == 0) return 0;
    if (m > INT32_MAX || n > INT32_MAX || k > INT32_MAX) return set_error("%s: dimension exceeds 2^31-1", who);
    const long long a_rows = tA ? k : m, b_rows = tB ? n : k;
    if (lda < std::max<long long>(a_rows, 1) || ldb < std::max<long long>(b_rows, 1) || ldc < m)
        return set_error("%s: leading dimension too small", who);
    if (!A || !B || !C) return set_error("%s: null buffer", who);
    const int typeA = tA ? KT : MT, typeB = tB ? MT : KT;
    const bool big = m >= 128 && n >= 128 && k >= 64 && alpha != 0.f;
    const bool tma_ok = ((uintptr_t)A % 16 == 0) && ((uintptr_t)B % 16 == 0) && (lda % 8 == 0) && (ldb % 8 == 0);
    if (big && tma_ok) {
        CUtensorMap ma, mb;
        const int bn = (n >= 256) ? 256 : 128;
        if (make_map_h(&ma, E, typeA, A, lda, (int)m, (int)k, BM) && make_map_h(&mb, E, typeB, B, ldb, (int)n, (int)k, bn)) {
            const cudaError_t e = (bn == 256) ? dispatch<E, 256>(typeA, typeB, ma, mb, C, ldc, (int)m, (int)n, (int)k, alpha, beta, s)
                                              : dispatch<E, 128>(typeA, typeB, ma, mb, C, ldc, (int)m, (int)n, (int)k, alpha, beta, s);
            return check_cuda(e, "hgemm_tc_kernel");
        }
    }
    const long long total = m * n;
    const unsigned grid = (unsigned)std::max<long long>(1, std::min<long long>((total + 255) / 256, (long long)num_sms() * 8));
    hgemm_generic_kernel<E><<<grid, 256, 0, s>>>(tA, tB, (int)m, (int)n, (int)k, alpha, static_cast<const HT*>(A), lda,
                                                 static_cast<const HT*>(B), ldb, beta, static_cast<HT*>(C), ldc);
    return check_cuda(cudaGetLastError(), "hgemm_generic_kernel");
}

}  // namespace tc16

int lgemm_f16(int tA, int tB, long long m, long long n, long long k, float alpha, const void* A, long long lda, const void* B,
              long long ldb, float beta, void* C, long long ldc, cudaStream_t s) {
    return tc16::lgemm_half<tc16::F16>(tA, tB, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, s);
}
int lgemm_bf16(int tA, int tB, long long m, long long n, long long k, float alpha, const void* A, long long lda, const void* B,
               long long ldb, float beta, void* C, long long ldc, cudaStream_t s) {
    return tc16::lgemm_half<tc16::BF16>(tA, tB, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, s);
}

}  // namespace hyd

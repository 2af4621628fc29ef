// Local (per-rank) FP64 GEMM for sm_100a:  C := alpha * op(A) * op(B) + beta * C
// column-major, arbitrary leading dimensions, all four (transA, transB) combos.
//
// Replaces the reference's cuBLAS call for this path:
//   gpu_blas::Gemm  (include/hydrogen/blas/GPU_BLAS_impl.hpp:631-646, :277-303)
//   -> cublasDgemm  (src/core/imports/cublas.cpp:130-148), called from
//   Gemm_impl<GPU>  (src/blas_like/level3/Gemm.cpp:104-157).
//
// Design (B200-first, not a translation):
//   * persistent CTAs (one per SM), 128x128 C tile, BK = 16 per pipeline stage;
//   * one producer warp issues TMA (cp.async.bulk.tensor.2d, SWIZZLE_128B) into a
//     kStages-deep shared-memory ring guarded by full/empty mbarriers;
//   * eight consumer warps (2 x 4, 64x32 warp tile) run DMMA m16n8k16 with
//     accumulators in registers.  FP64 has no tcgen05 kind, so the FP64 tensor pipe
//     is reached through warp-level mma.sync;
//   * both operand layouts ("outer-contiguous" = M-type, "k-contiguous" = K-type) are
//     staged by TMA in their native orientation; the contraction index is permuted
//     (kmap) so that every fragment load is shared-memory bank-conflict free under
//     the 128-byte swizzle.  The permutation is applied to A and B alike, so the
//     product is unchanged.
#include "ptx.cuh"
#include "hyd_internal.h"

namespace hyd {

constexpr int BM = 128, BN = 128, BK = 16;
constexpr int kConsumerWarps = 8;
constexpr int kThreads = (kConsumerWarps + 4) * 32;  // 2 consumer warpgroups + 1 producer warpgroup
constexpr int kStageBytesA = BM * BK * 8;  // 16 KiB
constexpr int kStageBytesB = BN * BK * 8;  // 16 KiB
constexpr int kStageBytes = kStageBytesA + kStageBytesB;
constexpr int kStages = 6;  // 192 KiB ring
constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;

// Operand staging type.
//  MT (outer-contiguous): global element (o,k) at o + k*ld.  smem = BM/16 boxes of
//      [16 k-lines][16 o] (128 B lines, SWIZZLE_128B, XOR key = k & 7).
//  KT (k-contiguous):     global element (o,k) at k + o*ld.  smem = [BM o-lines][16 k]
//      (128 B lines, SWIZZLE_128B, XOR key = o & 7).
enum { MT = 0, KT = 1 };

// kmap: which contraction index (0..15 inside a stage) fragment slot (t, j) carries.
//  KM_M is conflict-free for MT operands, KM_K for KT operands.
template <int KM>
__device__ __forceinline__ int kmap(int t, int j) {
    if (KM == MT) return 2 * t + (j & 1) + 8 * (j >> 1);
    return (t & 1) + 8 * (t >> 1) + 2 * j;
}

// Per-thread byte offset (inside an operand stage) of fragment slot j for the
// 8-row group starting at outer index 0; the caller adds the group's base.
template <int TYPE, int KM>
__device__ __forceinline__ uint32_t frag_off(int g, int t, int j) {
    const int k = kmap<KM>(t, j);
    if (TYPE == MT) return k * 128 + ((g * 8) ^ ((k & 7) << 4));
    return g * 128 + ((k * 8) ^ (g << 4));
}
// Byte offset of the 8-row group with outer base o8 (multiple of 8).
template <int TYPE>
__device__ __forceinline__ uint32_t group_base(int o8) {
    if (TYPE == MT) return (o8 >> 4) * 2048;  // (+ XOR 64 for the odd group, see load)
    return o8 * 128;
}

template <int TA, int TB>
__global__ void __launch_bounds__(kThreads, 1)
dgemm_tma_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                 double* __restrict__ C, long long ldc, int M, int N, int K, double alpha,
                 double beta, int tiles_m, int tiles_n, int group_m) {
    constexpr int KM = (TA == KT && TB == KT) ? KT : MT;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + kStages * kStageBytes;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (kStages + s); };

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(full_bar(s), 1);
            mbar_init(empty_bar(s), kConsumerWarps);
        }
        mbar_fence_init();
    }
    __syncthreads();

    const int num_kb = (K + BK - 1) / BK;
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

    if (warp >= kConsumerWarps) {
        // ===================== TMA producer warpgroup =====================
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        if (warp == kConsumerWarps && lane == 0) {
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
                    const uint32_t sA = smem_base + stage * kStageBytes;
                    const uint32_t sB = sA + kStageBytesA;
                    mbar_expect_tx(full_bar(stage), kStageBytes);
                    if (TA == MT) {
#pragma unroll
                        for (int b = 0; b < BM / 16; ++b)
                            tma_load_2d(sA + b * 2048, &mapA, full_bar(stage), m0 + b * 16, kb * BK);
                    } else {
                        tma_load_2d(sA, &mapA, full_bar(stage), kb * BK, m0);
                    }
                    if (TB == MT) {
#pragma unroll
                        for (int b = 0; b < BN / 16; ++b)
                            tma_load_2d(sB + b * 2048, &mapB, full_bar(stage), n0 + b * 16, kb * BK);
                    } else {
                        tma_load_2d(sB, &mapB, full_bar(stage), kb * BK, n0);
                    }
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
        return;
    }

    // ===================== DMMA consumers =====================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    const int g = lane >> 2, t = lane & 3;
    const int wm = warp & 1, wn = warp >> 1;  // 2 x 4 warps, warp tile 64 x 32
    uint32_t offA[4], offB[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        offA[j] = group_base<TA>(wm * 64) + frag_off<TA, KM>(g, t, j);
        offB[j] = kStageBytesA + group_base<TB>(wn * 32) + frag_off<TB, KM>(g, t, j);
    }
    const uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));

    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int tm, tn;
        tile_coord(tile, tm, tn);
        double acc[4][4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j)
#pragma unroll
                for (int r = 0; r < 4; ++r) acc[i][j][r] = 0.0;

        for (int kb = 0; kb < num_kb; ++kb) {
            mbar_wait(full_bar(stage), phase);
            const uint8_t* st = smem_gen + stage * kStageBytes;
            double bf[4][4];
#pragma unroll
            for (int nt = 0; nt < 4; ++nt) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    uint32_t o;
                    if (TB == MT) o = (offB[j] + (nt >> 1) * 2048) ^ ((nt & 1) * 64);
                    else          o = offB[j] + nt * 1024;
                    bf[nt][j] = *reinterpret_cast<const double*>(st + o);
                }
            }
#pragma unroll
            for (int mt = 0; mt < 4; ++mt) {
                double af[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    uint32_t o;
                    if (TA == MT) o = (offA[i >> 1] + mt * 2048) ^ ((i & 1) * 64);
                    else          o = offA[i >> 1] + mt * 2048 + (i & 1) * 1024;
                    af[i] = *reinterpret_cast<const double*>(st + o);
                }
#pragma unroll
                for (int nt = 0; nt < 4; ++nt) dmma_16816(acc[mt][nt], af, bf[nt]);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(empty_bar(stage));
            if (++stage == kStages) { stage = 0; phase ^= 1; }
        }

        // ---- epilogue: C = alpha*acc + beta*C, straight from registers ----
        const int row0 = tm * BM + wm * 64 + g;
        const int col0 = tn * BN + wn * 32 + 2 * t;
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                const int col = col0 + nt * 8 + c;
                if (col < N) {
                    double* cp = C + (long long)col * ldc;
#pragma unroll
                    for (int mt = 0; mt < 4; ++mt) {
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            const int row = row0 + mt * 16 + h * 8;
                            if (row < M) {
                                const double v = alpha * acc[mt][nt][h * 2 + c];
                                if (beta == 0.0) cp[row] = v;
                                else cp[row] = v + beta * cp[row];
                            }
                        }
                    }
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------
// Generic fallback for operands TMA cannot describe (odd leading dimension or a
// base pointer that is not 16-byte aligned).  Plain tiled DFMA kernel; correctness
// path only -- the SUMMA engine always produces TMA-compatible panels.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
dgemm_generic_kernel(int tA, int tB, int M, int N, int K, double alpha, const double* __restrict__ A,
                     long long lda, const double* __restrict__ B, long long ldb, double beta,
                     double* __restrict__ C, long long ldc) {
    __shared__ double sA[16][65];
    __shared__ double sB[16][65];
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int m0 = blockIdx.x * 64, n0 = blockIdx.y * 64;
    double acc[4][4] = {};
    for (int k0 = 0; k0 < K; k0 += 16) {
        for (int e = threadIdx.x; e < 16 * 64; e += 256) {
            const int kk = e & 15, o = e >> 4;
            const int k = k0 + kk;
            double a = 0.0, b = 0.0;
            if (k < K && m0 + o < M)
                a = tA ? A[k + (long long)(m0 + o) * lda] : A[(m0 + o) + (long long)k * lda];
            if (k < K && n0 + o < N)
                b = tB ? B[(n0 + o) + (long long)k * ldb] : B[k + (long long)(n0 + o) * ldb];
            sA[kk][o] = a;
            sB[kk][o] = b;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < 16; ++kk) {
            double a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) { a[i] = sA[kk][tx + 16 * i]; b[i] = sB[kk][ty + 16 * i]; }
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fma(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int col = n0 + ty + 16 * j;
        if (col >= N) continue;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int row = m0 + tx + 16 * i;
            if (row >= M) continue;
            double* cp = C + row + (long long)col * ldc;
            const double v = alpha * acc[i][j];
            *cp = (beta == 0.0) ? v : v + beta * *cp;
        }
    }
}

// C := beta * C   (k == 0 path; reference Gemm.cpp:212-219 -> Scale)
__global__ void dscale_kernel(int M, int N, double beta, double* __restrict__ C, long long ldc) {
    const long long total = (long long)M * N;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
         e += (long long)gridDim.x * blockDim.x) {
        const long long j = e / M, i = e - j * M;
        double* cp = C + i + j * ldc;
        *cp = (beta == 0.0) ? 0.0 : beta * *cp;
    }
}

static bool make_map_f64(CUtensorMap* map, int type, const double* base, long long ld, int outer,
                         int K, int box_outer) {
    // MT: dims {outer, K}, box {16,16};  KT: dims {K, outer}, box {16, box_outer}.
    cuuint64_t dims[2], strides[1];
    cuuint32_t box[2], estr[2] = {1, 1};
    if (type == MT) { dims[0] = outer; dims[1] = K; box[0] = 16; box[1] = 16; }
    else            { dims[0] = K; dims[1] = outer; box[0] = 16; box[1] = box_outer; }
    strides[0] = (cuuint64_t)ld * 8;
    if (!tensor_map_encode()) return false;
    CUresult r = tensor_map_encode()(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, (void*)base, dims,
                                     strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS;
}

template <int TA, int TB>
static cudaError_t launch_tma(const CUtensorMap& ma, const CUtensorMap& mb, double* C, long long ldc,
                              int M, int N, int K, double alpha, double beta, cudaStream_t s, int sm_reserve) {
    static bool attr_set = false;
    auto kern = dgemm_tma_kernel<TA, TB>;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    const int tiles_m = (M + BM - 1) / BM, tiles_n = (N + BN - 1) / BN;
    const int grid = max(1, min(tiles_m * tiles_n, num_sms() - sm_reserve));
    const int group_m = 16;
    kern<<<grid, kThreads, kSmemBytes, s>>>(ma, mb, C, ldc, M, N, K, alpha, beta, tiles_m, tiles_n, group_m);
    return cudaGetLastError();
}

int lgemm_f64(int tA, int tB, long long m, long long n, long long k, double alpha, const double* A,
              long long lda, const double* B, long long ldb, double beta, double* C, long long ldc,
              cudaStream_t s) {
    return lgemm_f64_ex(tA, tB, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, s, 0);
}

// sm_reserve: SMs left free for concurrently running communication kernels (SUMMA overlap)
int lgemm_f64_ex(int tA, int tB, long long m, long long n, long long k, double alpha, const double* A,
                 long long lda, const double* B, long long ldb, double beta, double* C, long long ldc,
                 cudaStream_t s, int sm_reserve) {
    if (m < 0 || n < 0 || k < 0) return set_error("hyd_lgemm_f64: negative dimension");
    if (m == 0 || n == 0) return 0;
    if (m > INT32_MAX || n > INT32_MAX || k > INT32_MAX) return set_error("hyd_lgemm_f64: dimension exceeds 2^31-1");
    const long long a_rows = tA ? k : m, b_rows = tB ? n : k;
    if (lda < (a_rows > 1 ? a_rows : 1) || ldb < (b_rows > 1 ? b_rows : 1) || ldc < m)
        return set_error("hyd_lgemm_f64: leading dimension too small");
    if (k == 0 || alpha == 0.0) {
        if (beta != 1.0) {
            dscale_kernel<<<num_sms() * 8, 256, 0, s>>>((int)m, (int)n, beta, C, ldc);
            return check_cuda(cudaGetLastError(), "dscale_kernel");
        }
        return 0;
    }
    const int typeA = tA ? KT : MT;   // op(A) is m x k: N -> m contiguous, T -> k contiguous
    const int typeB = tB ? MT : KT;   // op(B) is k x n: N -> k contiguous, T -> n contiguous
    const bool tma_ok = ((uintptr_t)A % 16 == 0) && ((uintptr_t)B % 16 == 0) && (lda % 2 == 0) &&
                        (ldb % 2 == 0);
    CUtensorMap ma, mb;
    if (tma_ok && make_map_f64(&ma, typeA, A, lda, (int)m, (int)k, BM) &&
        make_map_f64(&mb, typeB, B, ldb, (int)n, (int)k, BN)) {
        cudaError_t e;
        if (typeA == MT && typeB == MT) e = launch_tma<MT, MT>(ma, mb, C, ldc, m, n, k, alpha, beta, s, sm_reserve);
        else if (typeA == MT && typeB == KT) e = launch_tma<MT, KT>(ma, mb, C, ldc, m, n, k, alpha, beta, s, sm_reserve);
        else if (typeA == KT && typeB == MT) e = launch_tma<KT, MT>(ma, mb, C, ldc, m, n, k, alpha, beta, s, sm_reserve);
        else e = launch_tma<KT, KT>(ma, mb, C, ldc, m, n, k, alpha, beta, s, sm_reserve);
        return check_cuda(e, "dgemm_tma_kernel");
    }
    dim3 grid((unsigned)((m + 63) / 64), (unsigned)((n + 63) / 64));
    dgemm_generic_kernel<<<grid, 256, 0, s>>>(tA, tB, (int)m, (int)n, (int)k, alpha, A, lda, B, ldb, beta, C, ldc);
    return check_cuda(cudaGetLastError(), "dgemm_generic_kernel");
}

}  // namespace hyd

// Local FP32 GEMM on the 5th-generation tensor cores: split-TF32 (3 x TF32) with tcgen05.mma and TMEM.
//
//   C := alpha * op(A) * op(B) + beta * C     (column-major, all four transpose combinations)
//
// Replaces gpu_blas::Gemm<float> -> cublasSgemm (include/hydrogen/blas/GPU_BLAS_impl.hpp:277-303,
// src/core/imports/cublas.cpp:130-148) for shapes large enough to fill the machine; small or
// TMA-incompatible operands take the FFMA kernel (lgemm_f32.cu).
//
// Arithmetic: every fp32 operand x is split exactly into  hi = x with the low 13 mantissa bits cleared
// (a TF32 number) and lo = x - hi (rounded to TF32), and the product is accumulated in fp32 as
//     A*B  ~=  A_lo*B_hi + A_hi*B_lo + A_hi*B_hi          (the lo*lo term, ~2^-22 relative, is dropped)
// which keeps the result within a few fp32 ulps of an FFMA GEMM -- inside BASELINE.json's fp32 bound
// (relative Frobenius error <= 1e-5 * sqrt(k)) with two orders of magnitude to spare.
//
// Structure (one persistent CTA per SM, warp-specialised, three pipelines):
//   warp 0        TMA producer: raw fp32 tiles of A and B -> shared memory ring (cp.async.bulk.tensor, swizzled)
//   warps 8-11    splitters: raw tile (= hi: the MMA ignores the low 13 bits) -> lo (second buffer), same swizzled byte positions,
//                 fence.proxy.async, then hand the stage to the MMA warp
//   warp 1        MMA issuer (one elected lane): per k-step three tcgen05.mma.kind::tf32 into a TMEM accumulator,
//                 tcgen05.commit releases the stage / publishes the accumulator
//   warps 4-7     epilogue: tcgen05.ld accumulator -> registers, C = alpha*acc + beta*C, coalesced stores;
//                 two TMEM accumulators so the epilogue of tile i overlaps the MMAs of tile i+1
//   warp 2        TMEM allocator
#include "ptx.cuh"
#include "hyd_internal.h"
#include <algorithm>
#include <stdlib.h>
#include <string.h>

namespace hyd {
namespace tc32 {

constexpr int BM = 128;
enum { MT = 0, KT = 1 };  // operand staging: MN-contiguous (outer index contiguous) or K-contiguous

template <int BN, int BK>
struct Cfg {
    static constexpr int kSwizzleBytes = BK * 4;                 // 128 (BK = 32) or 64 (BK = 16): one smem line
    static constexpr int kBytesA = BM * BK * 4, kBytesB = BN * BK * 4;
    static constexpr int kRaw = kBytesA + kBytesB;               // hi tiles (TMA destination)
    static constexpr int kStageBytes = 2 * kRaw;                 // + lo tiles
    static constexpr int kStages = (196608 / kStageBytes) < 6 ? (196608 / kStageBytes) : 6;
    static constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 256;
    static constexpr int kTmemCols = 2 * BN;                     // two accumulators
    static constexpr int kThreads = 12 * 32;
};

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_tf32(uint32_t d_tmem, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(d_tmem),
        "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
        : "memory");
}
// shared-memory matrix descriptor (SM100 UMMA): start address, leading / stride byte offsets (>>4), version 1, swizzle mode
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
__device__ __forceinline__ float to_tf32_rna(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}

template <int TA, int TB, int BN, int BK>
__global__ void __launch_bounds__(Cfg<BN, BK>::kThreads, 1)
sgemm_tc_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, float* __restrict__ C,
                long long ldc, int M, int N, int K, float alpha, float beta, int tiles_m, int tiles_n, int group_m) {
    typedef Cfg<BN, BK> G;
    constexpr int kStages = G::kStages, kStageBytes = G::kStageBytes, kRaw = G::kRaw, kBytesA = G::kBytesA;
    constexpr int kLine = G::kSwizzleBytes;              // bytes of one shared-memory line (BK floats)
    constexpr int kAtom = 8 * kLine;                     // one swizzle atom: 8 lines
    constexpr uint32_t kLayout = (BK == 32) ? 2u : 4u;   // SWIZZLE_128B : SWIZZLE_64B
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
    const uint32_t bar_base = smem_base + kStages * kStageBytes;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };                       // TMA landed
    auto conv_bar = [&](int s) { return bar_base + 8u * (kStages + s); };           // split done
    auto empty_bar = [&](int s) { return bar_base + 8u * (2 * kStages + s); };      // MMAs have read the stage
    auto tfull_bar = [&](int a) { return bar_base + 8u * (3 * kStages + a); };      // accumulator complete
    auto tempty_bar = [&](int a) { return bar_base + 8u * (3 * kStages + 2 + a); }; // accumulator drained
    const uint32_t tmem_slot = bar_base + 8u * (3 * kStages + 4);
    volatile uint32_t* tmem_slot_gen = reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) {
            mbar_init(full_bar(s), 1);
            mbar_init(conv_bar(s), 4);
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
                    mbar_expect_tx(full_bar(stage), kRaw);
                    if (TA == MT) {  // boxes of [BK k-lines][32 outer floats]: 128-byte lines along the outer index
#pragma unroll
                        for (int b = 0; b < BM / 32; ++b) tma_load_2d(sA + b * (BK * 128), &mapA, full_bar(stage), m0 + b * 32, kb * BK);
                    } else {
                        tma_load_2d(sA, &mapA, full_bar(stage), kb * BK, m0);
                    }
                    if (TB == MT) {
#pragma unroll
                        for (int b = 0; b < BN / 32; ++b) tma_load_2d(sB + b * (BK * 128), &mapB, full_bar(stage), n0 + b * 32, kb * BK);
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
            // instruction descriptor: D = F32, A = B = TF32, majors, N >> 3, M >> 4
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((TA == MT ? 1u : 0u) << 15) | ((TB == MT ? 1u : 0u) << 16) |
                                   ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
            // K-major (KT): lines of BK floats along k, SWIZZLE_128B / 64B; 8-row groups kAtom apart (SBO); one k-step
            //   (8 floats) = +32 B inside the swizzled line.
            // MN-major (MT): tf32 allows only the 128B-swizzle-with-32B-atoms layout: 128-byte lines along the outer index,
            //   4-line atoms 512 B apart (SBO), 32-float chunks of the outer index BK*128 B apart (LBO); one k-step = 8 lines.
            const uint32_t lboA = (TA == MT) ? BK * 128 : 16, lboB = (TB == MT) ? BK * 128 : 16;
            const uint32_t sboA = (TA == MT) ? 512 : kAtom, sboB = (TB == MT) ? 512 : kAtom;
            const uint32_t layA = (TA == MT) ? 1u : kLayout, layB = (TB == MT) ? 1u : kLayout;
            const uint32_t stepA = (TA == MT) ? 1024 : 32, stepB = (TB == MT) ? 1024 : 32;
            int stage = 0, acc = 0;
            uint32_t phase = 0, acc_phase = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                mbar_wait(tempty_bar(acc), acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(conv_bar(stage), phase);
                    tc_fence_after();
                    const uint32_t sA = smem_base + stage * kStageBytes, sB = sA + kBytesA;
#pragma unroll
                    for (int ks = 0; ks < BK / 8; ++ks) {
                        const uint64_t a_hi = make_desc(sA + ks * stepA, lboA, sboA, layA);
                        const uint64_t b_hi = make_desc(sB + ks * stepB, lboB, sboB, layB);
                        const uint64_t a_lo = make_desc(sA + kRaw + ks * stepA, lboA, sboA, layA);
                        const uint64_t b_lo = make_desc(sB + kRaw + ks * stepB, lboB, sboB, layB);
                        tc_mma_tf32(d_tmem, a_lo, b_hi, idesc, (kb | ks) != 0);
                        tc_mma_tf32(d_tmem, a_hi, b_lo, idesc, 1);
                        tc_mma_tf32(d_tmem, a_hi, b_hi, idesc, 1);
                    }
                    tc_commit(empty_bar(stage));
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
                tc_commit(tfull_bar(acc));
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else if (warp >= 8) {
        // ===================== splitters: hi (in place) / lo =====================
        const int t = threadIdx.x - 256;  // 0..127
        int stage = 0;
        uint32_t phase = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            for (int kb = 0; kb < num_kb; ++kb) {
                mbar_wait(full_bar(stage), phase);
                float4* raw = reinterpret_cast<float4*>(smem_gen + stage * kStageBytes);
                float4* lo = reinterpret_cast<float4*>(smem_gen + stage * kStageBytes + kRaw);
#pragma unroll 4
                for (int i = t; i < kRaw / 16; i += 128) {
                    const float4 v = raw[i];
                    float4 h, l;
                    h.x = __uint_as_float(__float_as_uint(v.x) & 0xFFFFE000u);
                    h.y = __uint_as_float(__float_as_uint(v.y) & 0xFFFFE000u);
                    h.z = __uint_as_float(__float_as_uint(v.z) & 0xFFFFE000u);
                    h.w = __uint_as_float(__float_as_uint(v.w) & 0xFFFFE000u);
                    l.x = to_tf32_rna(v.x - h.x);
                    l.y = to_tf32_rna(v.y - h.y);
                    l.z = to_tf32_rna(v.z - h.z);
                    l.w = to_tf32_rna(v.w - h.w);
                    // hi is NOT written back: kind::tf32 reads the upper 19 bits of each 32-bit container and ignores the
                    // low 13 mantissa bits, so the raw fp32 tile already is the hi operand (bit-identical results with
                    // and without the store on B200; a rounding tensor core would fail the parity tests at 2^-11).
                    // Saves a third of the splitters' shared-memory traffic, the kernel's limiter (+5 %).
                    lo[i] = l;
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the tensor core
                __syncwarp();
                if (lane == 0) mbar_arrive(conv_bar(stage));
                if (++stage == kStages) { stage = 0; phase ^= 1; }
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
                float* cp = C + (long long)col0 * ldc + row;
                if (rd && row < M) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) old[j] = (col0 + j < N) ? __ldcs(cp + (long long)j * ldc) : 0.f;
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
                            __stcs(cp + (long long)j * ldc, v);
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

static bool make_map_f32(CUtensorMap* map, int type, const float* base, long long ld, int outer, int K, int box_outer, int BK) {
    cuuint64_t dims[2], strides[1];
    cuuint32_t box[2], estr[2] = {1, 1};
    if (type == MT) { dims[0] = outer; dims[1] = K; box[0] = 32; box[1] = BK; }
    else            { dims[0] = K; dims[1] = outer; box[0] = BK; box[1] = box_outer; }
    const CUtensorMapSwizzle sw = (type == MT) ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B
                                               : (BK == 32 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B);
    strides[0] = (cuuint64_t)ld * 4;
    if (!tensor_map_encode()) return false;
    const CUresult r = tensor_map_encode()(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)base, dims, strides, box, estr,
                                           CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                                           CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS;
}

template <int TA, int TB, int BN, int BK>
static cudaError_t launch(const CUtensorMap& ma, const CUtensorMap& mb, float* C, long long ldc, int M, int N, int K, float alpha,
                          float beta, cudaStream_t s, int sm_reserve) {
    typedef Cfg<BN, BK> G;
    static bool attr_set = false;
    auto kern = sgemm_tc_kernel<TA, TB, BN, BK>;
    if (!attr_set) {
        const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, G::kSmemBytes);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    const int tiles_m = (M + BM - 1) / BM, tiles_n = (N + BN - 1) / BN;
    const int grid = max(1, min(tiles_m * tiles_n, num_sms() - sm_reserve));
    kern<<<grid, G::kThreads, G::kSmemBytes, s>>>(ma, mb, C, ldc, M, N, K, alpha, beta, tiles_m, tiles_n, 16);
    return cudaGetLastError();
}

template <int BN, int BK>
static cudaError_t dispatch(int typeA, int typeB, const CUtensorMap& ma, const CUtensorMap& mb, float* C, long long ldc, int M, int N,
                            int K, float alpha, float beta, cudaStream_t s, int sm_reserve) {
    if (typeA == MT && typeB == MT) return launch<MT, MT, BN, BK>(ma, mb, C, ldc, M, N, K, alpha, beta, s, sm_reserve);
    if (typeA == MT && typeB == KT) return launch<MT, KT, BN, BK>(ma, mb, C, ldc, M, N, K, alpha, beta, s, sm_reserve);
    if (typeA == KT && typeB == MT) return launch<KT, MT, BN, BK>(ma, mb, C, ldc, M, N, K, alpha, beta, s, sm_reserve);
    return launch<KT, KT, BN, BK>(ma, mb, C, ldc, M, N, K, alpha, beta, s, sm_reserve);
}

// ---- TF32 tensor-pipe probe: the roofline denominator of the split-TF32 kernel ------------------------------------------
// One CTA per SM; one thread issues `iters` x 4 tcgen05.mma.kind::tf32 of shape 128 x 256 x 8 (the kernel's own instruction
// shape, both operands K-major SWIZZLE_128B in shared memory, zero-filled, never reloaded) into one TMEM accumulator and
// commits once at the end.  No TMA, no splitter, no epilogue: what is left is the rate at which the tensor core can execute
// SS-form MMAs.  TFLOP/s returned = 2*128*256*8 flops per MMA (dense TF32); the 3xTF32 kernel's bound is a third of it.
__global__ void __launch_bounds__(128, 1) tf32_probe_kernel(int iters) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
    const uint32_t sA = smem_base, sB = smem_base + 128 * 128, bar = smem_base + 128 * 128 + 256 * 128, slot = bar + 8;
    volatile uint32_t* slot_gen = reinterpret_cast<volatile uint32_t*>(smem_gen + (slot - smem_base));
    for (int i = threadIdx.x; i < (128 + 256) * 128 / 16; i += blockDim.x) reinterpret_cast<float4*>(smem_gen)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { mbar_init(bar, 1); mbar_fence_init(); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot), "r"(256u));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *slot_gen;
    if (threadIdx.x == 0) {
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(256 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int ks = 0; ks < 4; ++ks)
                tc_mma_tf32(tmem_base, make_desc(sA + ks * 32, 16, 1024, 2u), make_desc(sB + ks * 32, 16, 1024, 2u), idesc, (it | ks) != 0);
        }
        tc_commit(bar);
        mbar_wait(bar, 0);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(256u));
    }
}

}  // namespace tc32

// returns 0 = launched, 1 = error (hyd_last_error set), -1 = not applicable (caller falls back to the FFMA kernel)
int lgemm_f32_tc(int tA, int tB, long long m, long long n, long long k, float alpha, const float* A, long long lda, const float* B,
                 long long ldb, float beta, float* C, long long ldc, cudaStream_t s, int sm_reserve) {
    using namespace tc32;
    static int mode_env = -2;  // HYD_SGEMM: "ffma" disables this path, "tc128" / "tc256" force a tile; default: by shape
    if (mode_env == -2) {
        const char* e = getenv("HYD_SGEMM");
        mode_env = !e ? -1 : (!strcmp(e, "ffma") ? 0 : (!strcmp(e, "tc256") ? 256 : (!strcmp(e, "tc128") ? 128 : -1)));
    }
    if (mode_env == 0) return -1;
    // tile choice by estimated time: waves of the persistent grid x tile area, the 128x256 tile being ~15 % faster per
    // flop when the grid is full (fewer A re-reads per MMA); tall-skinny shapes quantise badly with it
    auto waves_cost = [&](int bn, double speed) {
        const long long tiles = ((m + BM - 1) / BM) * ((n + bn - 1) / bn);
        const long long ctas = std::max(1, num_sms() - sm_reserve);
        return (double)((tiles + ctas - 1) / ctas) * bn / speed;
    };
    const int mode = mode_env > 0 ? mode_env : ((n >= 192 && waves_cost(256, 1.15) <= waves_cost(128, 1.0)) ? 256 : 128);
    if (m < 128 || n < 64 || k < 32 || m * n < 128LL * 128 * 8) return -1;             // too small to fill the machine
    if (((uintptr_t)A % 16) || ((uintptr_t)B % 16) || (lda % 4) || (ldb % 4)) return -1;  // TMA needs 16-byte strides
    const int typeA = tA ? KT : MT, typeB = tB ? MT : KT;
    CUtensorMap ma, mb;
    cudaError_t e;
    if (mode == 256) {
        if (!make_map_f32(&ma, typeA, A, lda, (int)m, (int)k, BM, 16) || !make_map_f32(&mb, typeB, B, ldb, (int)n, (int)k, 256, 16)) return -1;
        e = dispatch<256, 16>(typeA, typeB, ma, mb, C, ldc, (int)m, (int)n, (int)k, alpha, beta, s, sm_reserve);
    } else {
        if (!make_map_f32(&ma, typeA, A, lda, (int)m, (int)k, BM, 32) || !make_map_f32(&mb, typeB, B, ldb, (int)n, (int)k, 128, 32)) return -1;
        e = dispatch<128, 32>(typeA, typeB, ma, mb, C, ldc, (int)m, (int)n, (int)k, alpha, beta, s, sm_reserve);
    }
    return check_cuda(e, "sgemm_tc_kernel");
}

// dense TF32 tcgen05 rate (TFLOP/s) of this GPU: see tf32_probe_kernel
int probe_tf32_peak(int iters, double* tflops) {
    using namespace tc32;
    const int smem = (128 + 256) * 128 + 1024 + 64;
    if (check_cuda(cudaFuncSetAttribute(tf32_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem), "tf32 probe attr")) return 1;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int grid = num_sms();
    for (int rep = 0; rep < 2; ++rep) {   // rep 0 = warm-up
        cudaEventRecord(e0);
        tf32_probe_kernel<<<grid, 128, smem>>>(iters);
        cudaEventRecord(e1);
        if (check_cuda(cudaEventSynchronize(e1), "tf32 probe kernel")) return 1;
    }
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *tflops = 2.0 * 128 * 256 * 8 * 4.0 * (double)iters * grid / (ms * 1e-3) / 1e12;
    return 0;
}

}  // namespace hyd

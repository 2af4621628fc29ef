// Lattice copy / reduce kernels: the pack and unpack side of every redistribution.
//
// A "lattice" is an nr x nc sub-grid of a column-major local matrix picked by two
// arithmetic progressions (rows r0 + a*rstep, columns c0 + b*cstep).  Every pack /
// unpack loop of the reference (copy::util::InterleaveMatrix, {Row,Col}Strided{Pack,Unpack},
// Partial{Row,Col}Strided*, Strided{Pack,Unpack}: include/El/blas_like/level1/Copy/util.hpp:122-230,
// 363-533,671-722; GPU forms :244-421 = one cudaMemcpy2DAsync / copy_2d_kernel per peer)
// is a set of such lattices, one per peer.  Here ALL peers of a redistribution are
// served by ONE launch: block x serves descriptor x mod n and strides
// over that lattice's tiles.  Index math is 64-bit (32768^2 locals overflow int32).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace hyd {

typedef long long i64;

// logical element (a,b), a < nr, b < nc:  dst[a*d_rs + b*d_cs] = src[a*s_rs + b*s_cs]
// (absolute base pointers, so one batch can mix staging buffers and matrices)
struct CopyDesc {
    const void* src;
    void* dst;
    i64 s_rs, s_cs;
    i64 d_rs, d_cs;
    i64 nr, nc;
};
constexpr int kMaxDesc = 24;
struct CopyBatch {
    CopyDesc d[kMaxDesc];
    int n;
};

// dst lattice += alpha * sum_{q<nq} member[q][a + b*nr]   (fixed ascending member order)
constexpr int kMaxMembers = 16;
struct ReduceDesc {
    void* dst;
    i64 d_rs, d_cs;
    i64 nr, nc;
    const void* member[kMaxMembers];
    int nq;
};
constexpr int kMaxRDesc = 8;
struct ReduceBatch {
    ReduceDesc d[kMaxRDesc];
    int n;
};

template <typename T> struct Vec16;
template <> struct Vec16<double> { typedef double2 type; static constexpr int N = 2; };
template <> struct Vec16<float> { typedef float4 type; static constexpr int N = 4; };
// 2-byte elements moved as opaque bits (fp16 / bf16 storage: SURVEY 8f N3); copies and merges only, no arithmetic
struct h16 { unsigned short v; };
template <> struct Vec16<h16> { typedef uint4 type; static constexpr int N = 8; };
template <typename T> __device__ __forceinline__ T ldg_scalar(const T* p) { return __ldg(p); }
template <> __device__ __forceinline__ h16 ldg_scalar<h16>(const h16* p) { h16 r; r.v = __ldg(&p->v); return r; }

// Paths, by the four strides: (1) both sides contiguous along a: 16-byte vector copy; (2) destination contiguous
// along a, source closer to contiguous along b (a transposing pack): (32 V) x 32 tile through shared memory,
// 16-byte stores; (3) destination contiguous, source strided along a (cyclic pack): V gathered loads, one
// 16-byte store; (4) source contiguous, destination strided (cyclic unpack): one 16-byte load, V stores;
// (5) anything else: scalar, thread index along a.
template <typename T>
__global__ void __launch_bounds__(256) lattice_copy_kernel(const __grid_constant__ CopyBatch batch) {
    // descriptors are interleaved along the grid (block x serves descriptor x mod n) so that the NVLink-bound peer
    // parts and the HBM-bound local part are resident together instead of one after the other
    const CopyDesc& D = batch.d[blockIdx.x % batch.n];
    const i64 bx = blockIdx.x / batch.n, gx = gridDim.x / batch.n;
    const i64 nr = D.nr, nc = D.nc;
    if (nr == 0 || nc == 0) return;
    const T* __restrict__ s = static_cast<const T*>(D.src);
    T* __restrict__ d = static_cast<T*>(D.dst);
    constexpr int V = Vec16<T>::N;
    typedef typename Vec16<T>::type VT;

    if (D.s_rs == 1 && D.d_rs == 1) {
        const bool vec_ok = (nr >= V) && (D.s_cs % V == 0) && (D.d_cs % V == 0) &&
                            (((uintptr_t)s % 16) == 0) && (((uintptr_t)d % 16) == 0);
        if (vec_ok) {
            // contiguous columns on both sides: 256 threads x 16 B along a, 4 columns per pass
            const i64 nrv = nr / V;
            const i64 chunks_a = (nrv + 255) / 256;
            const i64 col_groups = (nc + 3) / 4;
            for (i64 tile = bx; tile < chunks_a * col_groups; tile += gx) {
                const i64 ca = tile % chunks_a, cg = tile / chunks_a;
                const i64 av = ca * 256 + threadIdx.x;
                const i64 b0 = cg * 4;
                if (av < nrv) {
                    VT v[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        if (b0 + q < nc) v[q] = __ldg(reinterpret_cast<const VT*>(s + (b0 + q) * D.s_cs) + av);
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        if (b0 + q < nc) reinterpret_cast<VT*>(d + (b0 + q) * D.d_cs)[av] = v[q];
                }
                if (ca == 0) {  // scalar tail rows
                    const i64 tail = nr - nrv * V;
                    for (i64 e = threadIdx.x; e < tail * 4; e += 256) {
                        const i64 a = nrv * V + e % tail, b = b0 + e / tail;
                        if (b < nc) d[a + b * D.d_cs] = s[a + b * D.s_cs];
                    }
                }
            }
            return;
        }
    }
    if (D.s_rs != 1 && D.s_cs < D.s_rs && (D.d_rs == 1 || D.d_rs < D.d_cs)) {
        // transposing: read along b (where the source is contiguous or has the small cyclic stride; a-steps jump whole
        // columns), write along a (destination-contiguous).  The tile is staged
        // transposed ([b][a], pitch TA + V) so the write phase is one 16-byte vector per thread (peer stores are
        // issued per request: see the gather path below); TA = 32 V-vectors of a, 32 columns of b.
        constexpr int NV = 2;            // 16-byte vectors of a per thread in the write phase
        constexpr int TA = 32 * V * NV;  // 128 (fp64) / 256 (fp32) rows: 1 KiB contiguous destination runs
        __shared__ __align__(16) T tile_s[32][TA + V];
        const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
        const bool vec_st = (D.d_rs == 1) && (D.d_cs % V == 0) && (((uintptr_t)d % 16) == 0);
        const i64 ta = (nr + TA - 1) / TA, tb = (nc + 31) / 32;
        for (i64 tile = bx; tile < ta * tb; tile += gx) {
            const i64 a0 = (tile % ta) * TA, b0 = (tile / ta) * 32;
#pragma unroll
            for (int q = 0; q < TA / 8; ++q) {
                const i64 a = a0 + ty + 8 * q, b = b0 + tx;
                if (a < nr && b < nc) tile_s[tx][ty + 8 * q] = s[a * D.s_rs + b * D.s_cs];
            }
            __syncthreads();
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const i64 b = b0 + ty + 8 * q;
                if (b < nc) {
                    if (vec_st && a0 + TA <= nr) {  // uniform per tile: the two branches map lanes to elements differently
#pragma unroll
                        for (int h = 0; h < NV; ++h)
                            *reinterpret_cast<VT*>(d + a0 + (h * 32 + tx) * V + b * D.d_cs) =
                                *reinterpret_cast<const VT*>(&tile_s[ty + 8 * q][(h * 32 + tx) * V]);
                    } else {  // strided destination rows (the local part of a transposing copy) or a ragged edge
#pragma unroll
                        for (int e = 0; e < V * NV; ++e) {
                            const i64 ae = a0 + e * 32 + tx;
                            if (ae < nr) d[ae * D.d_rs + b * D.d_cs] = tile_s[ty + 8 * q][e * 32 + tx];
                        }
                    }
                }
            }
            __syncthreads();
        }
        return;
    }
    if (D.d_rs == 1 && D.s_rs != 1 && nr >= V && D.d_cs % V == 0 && ((uintptr_t)d % 16) == 0) {
        // strided source rows, contiguous destination (the usual pack into a message / a peer's receive region):
        // gather V source elements, ONE 16-byte store per thread -- peer stores over NVLink are issued per request,
        // so halving (fp64) / quartering (fp32) their number is what the push rate follows
        const i64 nrv = nr / V;
        const i64 chunks_a = (nrv + 255) / 256;
        const i64 col_groups = (nc + 3) / 4;
        for (i64 tile = bx; tile < chunks_a * col_groups; tile += gx) {
            const i64 ca = tile % chunks_a, cg = tile / chunks_a;
            const i64 av = ca * 256 + threadIdx.x;
            const i64 b0 = cg * 4;
            if (av < nrv) {
                T v[4][V];
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (b0 + q < nc) {
#pragma unroll
                        for (int e = 0; e < V; ++e) v[q][e] = ldg_scalar(s + (av * V + e) * D.s_rs + (b0 + q) * D.s_cs);
                    }
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (b0 + q < nc) reinterpret_cast<VT*>(d + (b0 + q) * D.d_cs)[av] = *reinterpret_cast<const VT*>(v[q]);
            }
            if (ca == 0) {  // scalar tail rows
                const i64 tail = nr - nrv * V;
                for (i64 e = threadIdx.x; e < tail * 4; e += 256) {
                    const i64 a = nrv * V + e % tail, b = b0 + e / tail;
                    if (b < nc) d[a + b * D.d_cs] = s[a * D.s_rs + b * D.s_cs];
                }
            }
        }
        return;
    }
    if (D.s_rs == 1 && D.d_rs != 1 && nr >= V && D.s_cs % V == 0 && ((uintptr_t)s % 16) == 0) {
        // contiguous source (a received message), strided destination rows: one 16-byte load, V scalar stores
        const i64 nrv = nr / V;
        const i64 chunks_a = (nrv + 255) / 256;
        const i64 col_groups = (nc + 3) / 4;
        for (i64 tile = bx; tile < chunks_a * col_groups; tile += gx) {
            const i64 ca = tile % chunks_a, cg = tile / chunks_a;
            const i64 av = ca * 256 + threadIdx.x;
            const i64 b0 = cg * 4;
            if (av < nrv) {
                VT v[4];
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (b0 + q < nc) v[q] = __ldg(reinterpret_cast<const VT*>(s + (b0 + q) * D.s_cs) + av);
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (b0 + q < nc) {
                        const T* e4 = reinterpret_cast<const T*>(&v[q]);
#pragma unroll
                        for (int e = 0; e < V; ++e) d[(av * V + e) * D.d_rs + (b0 + q) * D.d_cs] = e4[e];
                    }
            }
            if (ca == 0) {
                const i64 tail = nr - nrv * V;
                for (i64 e = threadIdx.x; e < tail * 4; e += 256) {
                    const i64 a = nrv * V + e % tail, b = b0 + e / tail;
                    if (b < nc) d[a * D.d_rs + b * D.d_cs] = s[a + b * D.s_cs];
                }
            }
        }
        return;
    }
    {
        // general strides: thread index along a, 8 columns per pass
        const i64 chunks_a = (nr + 255) / 256;
        const i64 col_groups = (nc + 7) / 8;
        for (i64 tile = bx; tile < chunks_a * col_groups; tile += gx) {
            const i64 a = (tile % chunks_a) * 256 + threadIdx.x;
            const i64 b0 = (tile / chunks_a) * 8;
            if (a < nr) {
                T v[8];
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    if (b0 + q < nc) v[q] = s[a * D.s_rs + (b0 + q) * D.s_cs];
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    if (b0 + q < nc) d[a * D.d_rs + (b0 + q) * D.d_cs] = v[q];
            }
        }
    }
}

// Merged unpack.  When the S parts that make up a destination block interleave its ROWS (part k holds local rows
// k, k+S, k+2S, ...: the element-cyclic case, e.g. two grid rows feeding a [*,VR] block, or a distributed
// transpose), unpacking them one after the other writes every 32-byte sector S times, a quarter or half at a time,
// and each pass is a read-modify-write in DRAM.  Here one pass writes full 16-byte vectors of consecutive
// destination rows, fetching each element from the part that owns it (part k stored contiguously: element t of
// column b at src[k][t*s_rs[k] + b*s_cs[k]], s_rs = 1 for a packed message); across a warp every part is still read as
// contiguous runs.  A part may also be a lattice of the SOURCE matrix itself (the rank's own share of a non-transposing
// redistribution, s_rs = its row step): it is then never packed at all.
constexpr int kMaxMerge = 16;
struct MergeDesc {
    const void* src[kMaxMerge];
    i64 s_rs[kMaxMerge];
    i64 s_cs[kMaxMerge];
    void* dst;      // element (0, 0) of the destination block (rows contiguous)
    i64 d_cs;
    i64 nr, nc;     // rows of the whole block (all parts together), columns
    int S;
};

template <typename T>
__global__ void __launch_bounds__(256) lattice_merge_kernel(const __grid_constant__ MergeDesc D) {
    constexpr int V = Vec16<T>::N;
    typedef typename Vec16<T>::type VT;
    const i64 nr = D.nr, nc = D.nc;
    const int S = D.S;
    T* __restrict__ d = static_cast<T*>(D.dst);
    const bool vec_ok = (D.d_cs % V == 0) && (((uintptr_t)d % 16) == 0);
    const i64 nrv = (nr + V - 1) / V;
    const i64 chunks_a = (nrv + 255) / 256;
    constexpr int CP = 8;  // columns per pass: 8 x V independent loads per thread keep a quarter-occupancy grid near HBM speed
    const i64 col_groups = (nc + CP - 1) / CP;
    for (i64 tile = blockIdx.x; tile < chunks_a * col_groups; tile += gridDim.x) {
        const i64 av = (tile % chunks_a) * 256 + threadIdx.x;
        const i64 b0 = (tile / chunks_a) * CP;
        if (av >= nrv) continue;
        const i64 a = av * V;
        const T* p[V];
        i64 cs[V];
#pragma unroll
        for (int e = 0; e < V; ++e) {
            const i64 ae = (a + e < nr) ? a + e : a;   // clamp: the value is not stored
            const int k = (int)(ae % S);
            p[e] = static_cast<const T*>(D.src[k]) + (ae / S) * D.s_rs[k];
            cs[e] = D.s_cs[k];
        }
        T v[CP][V];
#pragma unroll
        for (int q = 0; q < CP; ++q)
            if (b0 + q < nc) {
#pragma unroll
                for (int e = 0; e < V; ++e) v[q][e] = ldg_scalar(p[e] + (b0 + q) * cs[e]);
            }
#pragma unroll
        for (int q = 0; q < CP; ++q)
            if (b0 + q < nc) {
                T* o = d + a + (b0 + q) * D.d_cs;
                if (vec_ok && a + V <= nr) {
                    *reinterpret_cast<VT*>(o) = *reinterpret_cast<const VT*>(v[q]);
                } else {
#pragma unroll
                    for (int e = 0; e < V; ++e)
                        if (a + e < nr) o[e] = v[q][e];
                }
            }
    }
}

// Reduce-side unpack (reference: ReduceScatter + axpy::util::InterleaveMatrixUpdate,
// include/El/blas_like/level1/AxpyContract.hpp:60-86,  Axpy/util.hpp:23-51):
// the member partial sums are added in ascending member order, then B += alpha * sum
// (one rounding for alpha*sum, one for the add: no FMA contraction, as `B += alpha*A`
// compiled without -fmad would give; see __dmul_rn/__dadd_rn below).
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }

template <typename T>
__global__ void __launch_bounds__(256) lattice_reduce_kernel(T alpha, const __grid_constant__ ReduceBatch batch) {
    const ReduceDesc& D = batch.d[blockIdx.y];
    const i64 nr = D.nr, nc = D.nc;
    if (nr == 0 || nc == 0) return;
    T* d = static_cast<T*>(D.dst);
    const i64 chunks_a = (nr + 255) / 256;
    const i64 col_groups = (nc + 3) / 4;
    for (i64 tile = blockIdx.x; tile < chunks_a * col_groups; tile += gridDim.x) {
        const i64 a = (tile % chunks_a) * 256 + threadIdx.x;
        const i64 b0 = (tile / chunks_a) * 4;
        if (a >= nr) continue;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const i64 b = b0 + q;
            if (b >= nc) break;
            T acc = static_cast<const T*>(D.member[0])[a + b * nr];
            for (int m = 1; m < D.nq; ++m) acc = add_rn(acc, static_cast<const T*>(D.member[m])[a + b * nr]);
            T* p = d + a * D.d_rs + b * D.d_cs;
            *p = add_rn(*p, mul_rn(alpha, acc));
        }
    }
}

}  // namespace hyd

// The rest of the reference's GPU BLAS-1 layer (the level-1 operations LBANN issues around Gemm), on
// raw device buffers (seam S2) and on distributed matrices (seam S1):
//   hydrogen::Fill_GPU_impl(height, width, alpha, buffer, ldim, stream)          src/hydrogen/blas/gpu/Fill.cu:14-113
//   hydrogen::Hadamard_GPU_impl(h, w, X, csX, rsX, Y, csY, rsY, Z, csZ, rsZ, s)  src/hydrogen/blas/gpu/Hadamard.cu:14-127
//   gpu_blas::Axpy(TransposeMode, m, n, alpha, A, lda, B, ldb, si)  -> geam(op(A), B) / axpy_2d_transpose_tiled_kernel
//                                         include/hydrogen/blas/GPU_BLAS_impl.hpp:108-132,531-541; src/hydrogen/blas/gpu/Axpy.cu:14-68
//   gpu_blas::Axpy / Copy / Scale (size, X, incx, ...) 1-D forms                 GPU_BLAS_impl.hpp:66-84,134-150,174-188,510-597
//   El::Axpy(alpha, ElementalMatrix X, ElementalMatrix Y)                        include/El/blas_like/level1/Axpy.hpp:161-187
//   El::TransposeAxpy(alpha, X, Y)                                               include/El/blas_like/level1/TransposeAxpy.hpp
// All of them are HBM-bound; grids are persistent (SMs x 8) with grid-stride loops, 16-byte accesses whenever the
// operands allow it.
#include "../../include/hydrogen_b200.h"
#include "grid.h"
#include "hyd_internal.h"
#include <algorithm>

struct hyd_grid { hyd::Grid* g; };

namespace hyd {
namespace {

template <typename T> struct W16;
template <> struct W16<double> { typedef double2 type; static constexpr int N = 2; };
template <> struct W16<float> { typedef float4 type; static constexpr int N = 4; };

inline unsigned persistent_grid(i64 work_items) {
    return (unsigned)std::max<i64>(1, std::min<i64>(work_items, (i64)num_sms() * 8));
}
inline bool al16(const void* p) { return ((uintptr_t)p % 16) == 0; }

// ---------------------------------------------------------------------------------------------------------------
// Fill: A(i,j) = alpha
// ---------------------------------------------------------------------------------------------------------------
template <typename T, bool VEC>
__global__ void __launch_bounds__(256) fill2d_kernel(i64 m, i64 n, T alpha, T* __restrict__ A, i64 lda) {
    typedef typename W16<T>::type VT;
    constexpr int V = VEC ? W16<T>::N : 1;
    const i64 mv = m / V;
    const i64 row_chunks = (mv + 255) / 256, col_groups = (n + 3) / 4;
    VT v;
    T* vp = reinterpret_cast<T*>(&v);
#pragma unroll
    for (int e = 0; e < W16<T>::N; ++e) vp[e] = alpha;
    for (i64 t = blockIdx.x; t < row_chunks * col_groups; t += gridDim.x) {
        const i64 r = (t % row_chunks) * 256 + threadIdx.x, j0 = (t / row_chunks) * 4;
        if (r >= mv) continue;
#pragma unroll
        for (int q = 0; q < 4; ++q)
            if (j0 + q < n) {
                if (VEC) reinterpret_cast<VT*>(A + (j0 + q) * lda)[r] = v;
                else A[r + (j0 + q) * lda] = alpha;
            }
    }
}

template <typename T>
int fill2d(i64 m, i64 n, T alpha, T* A, i64 lda, cudaStream_t s) {
    const char* who = "hyd_fill";
    if (m < 0 || n < 0) return set_error("%s: negative dimension", who);
    if (m == 0 || n == 0) return 0;
    if (!A) return set_error("%s: null buffer", who);
    if (lda < m) return set_error("%s: leading dimension too small", who);
    constexpr int V = W16<T>::N;
    const bool vec = (m % V == 0) && (lda % V == 0) && al16(A);
    const i64 mv = vec ? m / V : m;
    const unsigned grid = persistent_grid(((mv + 255) / 256) * ((n + 3) / 4));
    if (vec) fill2d_kernel<T, true><<<grid, 256, 0, s>>>(m, n, alpha, A, lda);
    else     fill2d_kernel<T, false><<<grid, 256, 0, s>>>(m, n, alpha, A, lda);
    ++g_launches;
    return check_cuda(cudaGetLastError(), who);
}

// ---------------------------------------------------------------------------------------------------------------
// Hadamard: Z(i,j) = X(i,j) * Y(i,j), each operand with its own (colStride, rowStride).  Z may alias X or Y
// (the reference's MultAssign branches): every thread reads its operands before it writes the same position.
// ---------------------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) hadamard_vec_kernel(i64 mv, i64 n, const T* X, i64 ldx, const T* Y, i64 ldy, T* Z, i64 ldz) {
    typedef typename W16<T>::type VT;
    constexpr int V = W16<T>::N;
    const i64 row_chunks = (mv + 255) / 256, col_groups = (n + 3) / 4;
    for (i64 t = blockIdx.x; t < row_chunks * col_groups; t += gridDim.x) {
        const i64 r = (t % row_chunks) * 256 + threadIdx.x, j0 = (t / row_chunks) * 4;
        if (r >= mv) continue;
        VT x[4], y[4];
#pragma unroll
        for (int q = 0; q < 4; ++q)
            if (j0 + q < n) {
                x[q] = reinterpret_cast<const VT*>(X + (j0 + q) * ldx)[r];
                y[q] = reinterpret_cast<const VT*>(Y + (j0 + q) * ldy)[r];
            }
#pragma unroll
        for (int q = 0; q < 4; ++q)
            if (j0 + q < n) {
                VT o;
                T* op = reinterpret_cast<T*>(&o);
                const T* xp = reinterpret_cast<const T*>(&x[q]);
                const T* yp = reinterpret_cast<const T*>(&y[q]);
#pragma unroll
                for (int e = 0; e < V; ++e) op[e] = xp[e] * yp[e];
                reinterpret_cast<VT*>(Z + (j0 + q) * ldz)[r] = o;
            }
    }
}

template <typename T>
__global__ void __launch_bounds__(256) hadamard_gen_kernel(i64 m, i64 n, const T* X, i64 csx, i64 rsx, const T* Y, i64 csy, i64 rsy,
                                                           T* Z, i64 csz, i64 rsz) {
    const i64 row_chunks = (m + 255) / 256, col_groups = (n + 3) / 4;
    for (i64 t = blockIdx.x; t < row_chunks * col_groups; t += gridDim.x) {
        const i64 i = (t % row_chunks) * 256 + threadIdx.x, j0 = (t / row_chunks) * 4;
        if (i >= m) continue;
        T x[4], y[4];
#pragma unroll
        for (int q = 0; q < 4; ++q)
            if (j0 + q < n) {
                x[q] = X[i * csx + (j0 + q) * rsx];
                y[q] = Y[i * csy + (j0 + q) * rsy];
            }
#pragma unroll
        for (int q = 0; q < 4; ++q)
            if (j0 + q < n) Z[i * csz + (j0 + q) * rsz] = x[q] * y[q];
    }
}

template <typename T>
int hadamard(i64 m, i64 n, const T* X, i64 csx, i64 rsx, const T* Y, i64 csy, i64 rsy, T* Z, i64 csz, i64 rsz, cudaStream_t s) {
    const char* who = "hyd_hadamard";
    if (m < 0 || n < 0) return set_error("%s: negative dimension", who);
    if (m == 0 || n == 0) return 0;
    if (!X || !Y || !Z) return set_error("%s: null buffer", who);
    if (csx <= 0 || rsx <= 0 || csy <= 0 || rsy <= 0 || csz <= 0 || rsz <= 0) return set_error("%s: strides must be positive", who);
    constexpr int V = W16<T>::N;
    const bool vec = csx == 1 && csy == 1 && csz == 1 && (m % V == 0) && (rsx % V == 0) && (rsy % V == 0) && (rsz % V == 0) &&
                     al16(X) && al16(Y) && al16(Z);
    if (vec) {
        const i64 mv = m / V;
        hadamard_vec_kernel<T><<<persistent_grid(((mv + 255) / 256) * ((n + 3) / 4)), 256, 0, s>>>(mv, n, X, rsx, Y, rsy, Z, rsz);
    } else {
        hadamard_gen_kernel<T><<<persistent_grid(((m + 255) / 256) * ((n + 3) / 4)), 256, 0, s>>>(m, n, X, csx, rsx, Y, csy, rsy,
                                                                                                 Z, csz, rsz);
    }
    ++g_launches;
    return check_cuda(cudaGetLastError(), who);
}

// ---------------------------------------------------------------------------------------------------------------
// Transposing axpy: B(i,j) += alpha * A(j,i); B is m x n (ldb), A is n x m (lda).  64 x 64 tile staged through
// shared memory: A is read along its columns (index j), B is read and written along its columns (index i).
// ---------------------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) axpy_trans_kernel(i64 m, i64 n, T alpha, const T* __restrict__ A, i64 lda,
                                                         T* __restrict__ B, i64 ldb) {
    constexpr int TS = 64;
    __shared__ T tile[TS][TS + 1];
    const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;  // 64 x 4
    const i64 ti = (m + TS - 1) / TS, tj = (n + TS - 1) / TS;
    for (i64 t = blockIdx.x; t < ti * tj; t += gridDim.x) {
        const i64 i0 = (t % ti) * TS, j0 = (t / ti) * TS;
#pragma unroll
        for (int q = 0; q < TS / 4; ++q) {  // A(j0 + tx, i0 + ty + 4q): consecutive tx = consecutive addresses
            const i64 j = j0 + tx, i = i0 + ty + 4 * q;
            if (j < n && i < m) tile[ty + 4 * q][tx] = __ldg(A + j + i * lda);
        }
        __syncthreads();
#pragma unroll
        for (int q = 0; q < TS / 4; ++q) {  // B(i0 + tx, j0 + ty + 4q)
            const i64 i = i0 + tx, j = j0 + ty + 4 * q;
            if (i < m && j < n) {
                T* p = B + i + j * ldb;
                *p = *p + alpha * tile[tx][ty + 4 * q];
            }
        }
        __syncthreads();
    }
}

template <typename T>
int axpy_trans(i64 m, i64 n, T alpha, const T* A, i64 lda, T* B, i64 ldb, cudaStream_t s) {
    const char* who = "hyd_axpy_trans";
    if (m < 0 || n < 0) return set_error("%s: negative dimension", who);
    if (m == 0 || n == 0) return 0;
    if (!A || !B) return set_error("%s: null buffer", who);
    if (lda < n || ldb < m) return set_error("%s: leading dimension too small", who);
    axpy_trans_kernel<T><<<persistent_grid(((m + 63) / 64) * ((n + 63) / 64)), 256, 0, s>>>(m, n, alpha, A, lda, B, ldb);
    ++g_launches;
    return check_cuda(cudaGetLastError(), who);
}

// ---------------------------------------------------------------------------------------------------------------
// 1-D forms with increments: OP 0 copy (Y = X), 1 axpy (Y += alpha X), 2 scale (Y *= alpha; exact zeros for alpha == 0)
// ---------------------------------------------------------------------------------------------------------------
template <typename T, int OP>
__global__ void __launch_bounds__(256) vec1d_kernel(i64 n, T alpha, const T* X, i64 incx, T* Y, i64 incy) {
    for (i64 i0 = ((i64)blockIdx.x * 256 + threadIdx.x) * 4; i0 < n; i0 += (i64)gridDim.x * 1024) {
        T x[4], y[4];
#pragma unroll
        for (int q = 0; q < 4; ++q)
            if (i0 + q < n) {
                if (OP != 2) x[q] = X[(i0 + q) * incx];
                if (OP != 0) y[q] = Y[(i0 + q) * incy];
            }
#pragma unroll
        for (int q = 0; q < 4; ++q)
            if (i0 + q < n) {
                T r;
                if (OP == 0) r = x[q];
                else if (OP == 1) r = y[q] + alpha * x[q];
                else r = (alpha == T(0)) ? T(0) : alpha * y[q];
                Y[(i0 + q) * incy] = r;
            }
    }
}

template <typename T, int OP>
int vec1d(i64 n, T alpha, const T* X, i64 incx, T* Y, i64 incy, cudaStream_t s, const char* who) {
    if (n < 0) return set_error("%s: negative length", who);
    if (n == 0) return 0;
    if ((OP != 2 && !X) || !Y) return set_error("%s: null buffer", who);
    if ((OP != 2 && incx <= 0) || incy <= 0) return set_error("%s: increments must be positive", who);
    if ((OP == 2 || incx == 1) && incy == 1) {  // contiguous: the vectorised 2-D kernels on an n x 1 matrix
        if (OP == 0) return copy2d_local<T>(n, 1, X, n, Y, n, s);
        if (OP == 1) return axpy2d_local<T>(n, 1, alpha, X, n, Y, n, s);
        return scale_local<T>(n, 1, alpha, Y, n, s);
    }
    vec1d_kernel<T, OP><<<persistent_grid((n + 1023) / 1024), 256, 0, s>>>(n, alpha, X, incx, Y, incy);
    ++g_launches;
    return check_cuda(cudaGetLastError(), who);
}

// ---------------------------------------------------------------------------------------------------------------
// El::Axpy(alpha, X, Y) / El::TransposeAxpy(alpha, X, Y) on distributed matrices (Axpy.hpp:161-187):
// equal DistData -> local axpy on the blocks; otherwise X is first copied into Y's distribution AND alignment
// (XCopy->AlignWith(YDistData); Copy(X, *XCopy)), then added locally.  The copy is one one-hop redistribution of the
// engine (redist.cu) into a per-grid workspace.
// ---------------------------------------------------------------------------------------------------------------
template <typename T>
int axpy_dist(Grid& g, T alpha, const DM& X, const DM& Y, bool transpose, cudaStream_t s) {
    const char* who = transpose ? "hyd_axpy_dist (transpose)" : "hyd_axpy_dist";
    const i64 xm = transpose ? X.width : X.height, xn = transpose ? X.height : X.width;
    if (xm != Y.height || xn != Y.width) return set_error("%s: nonconformal Axpy: X %lldx%lld, Y %lldx%lld", who, xm, xn, Y.height, Y.width);
    const LocalShape Ly = local_shape(g.shape, Y, g.mc, g.mr);
    if (Y.height == 0 || Y.width == 0) return 0;
    if (Ly.lh > 0 && Ly.lw > 0 && (!Y.buf || Y.ldim < Ly.lh)) return set_error("%s: bad buffer / ldim of Y", who);
    const bool same = transpose ? (X.colDist == Y.rowDist && X.rowDist == Y.colDist && X.colAlign == Y.rowAlign && X.rowAlign == Y.colAlign)
                                : (X.colDist == Y.colDist && X.rowDist == Y.rowDist && X.colAlign == Y.colAlign && X.rowAlign == Y.rowAlign);
    if (same) {
        if (Ly.lh == 0 || Ly.lw == 0) return 0;
        const i64 need = transpose ? Ly.lw : Ly.lh;
        if (!X.buf || X.ldim < need) return set_error("%s: bad buffer / ldim of X", who);
        return transpose ? axpy_trans<T>(Ly.lh, Ly.lw, alpha, static_cast<const T*>(X.buf), X.ldim, static_cast<T*>(Y.buf), Y.ldim, s)
                         : axpy2d_local<T>(Ly.lh, Ly.lw, alpha, static_cast<const T*>(X.buf), X.ldim, static_cast<T*>(Y.buf), Y.ldim, s);
    }
    DM XC = Y;  // Y's distribution and alignments
    const i64 q = 16 / (i64)sizeof(T);
    XC.ldim = std::max<i64>(q, (Ly.lh + q - 1) / q * q);
    if (ws_reserve(g.level1_ws, (size_t)XC.ldim * (size_t)std::max<i64>(Ly.lw, 1) * sizeof(T))) return 1;
    XC.buf = g.level1_ws.buf;
    if (redistribute<T>(g, X, XC, transpose, 3, s)) return 1;  // collective: every rank takes part, also with an empty block
    if (Ly.lh == 0 || Ly.lw == 0) return 0;
    return axpy2d_local<T>(Ly.lh, Ly.lw, alpha, static_cast<const T*>(XC.buf), XC.ldim, static_cast<T*>(Y.buf), Y.ldim, s);
}

DM to_dm(const hyd_dm* a) {
    DM d;
    d.buf = a->buf; d.ldim = a->ldim; d.height = a->height; d.width = a->width;
    d.colDist = a->colDist; d.rowDist = a->rowDist; d.colAlign = a->colAlign; d.rowAlign = a->rowAlign; d.root = a->root;
    return d;
}
int check_dm(const GridShape& sh, const hyd_dm* a, const char* who) {
    if (!a) return set_error("%s: null matrix", who);
    if (a->height < 0 || a->width < 0) return set_error("%s: negative dimension", who);
    if (!pair_supported(a->colDist, a->rowDist)) return set_error("%s: unsupported distribution pair [%d,%d]", who, a->colDist, a->rowDist);
    if (a->colAlign < 0 || a->colAlign >= sh.stride(a->colDist) || a->rowAlign < 0 || a->rowAlign >= sh.stride(a->rowDist))
        return set_error("%s: alignment out of range", who);
    return 0;
}

}  // namespace
}  // namespace hyd

#define S(x) ((cudaStream_t)(x))

extern "C" {

#define HYD_LEVEL1(T, sfx)                                                                                                    \
    int hyd_fill_##sfx(int64_t m, int64_t n, T alpha, T* A, int64_t lda, hyd_stream_t s) {                                    \
        return hyd::fill2d<T>(m, n, alpha, A, lda, S(s));                                                                     \
    }                                                                                                                         \
    int hyd_hadamard_##sfx(int64_t m, int64_t n, const T* X, int64_t csX, int64_t rsX, const T* Y, int64_t csY, int64_t rsY,  \
                           T* Z, int64_t csZ, int64_t rsZ, hyd_stream_t s) {                                                  \
        return hyd::hadamard<T>(m, n, X, csX, rsX, Y, csY, rsY, Z, csZ, rsZ, S(s));                                           \
    }                                                                                                                         \
    int hyd_axpy_trans_##sfx(int transA, int64_t m, int64_t n, T alpha, const T* A, int64_t lda, T* B, int64_t ldb,           \
                             hyd_stream_t s) {                                                                                \
        return transA ? hyd::axpy_trans<T>(m, n, alpha, A, lda, B, ldb, S(s))                                                 \
                      : hyd::axpy2d_local<T>(m, n, alpha, A, lda, B, ldb, S(s));                                              \
    }                                                                                                                         \
    int hyd_copy1d_##sfx(int64_t n, const T* X, int64_t incx, T* Y, int64_t incy, hyd_stream_t s) {                           \
        return hyd::vec1d<T, 0>(n, T(0), X, incx, Y, incy, S(s), "hyd_copy1d");                                               \
    }                                                                                                                         \
    int hyd_axpy1d_##sfx(int64_t n, T alpha, const T* X, int64_t incx, T* Y, int64_t incy, hyd_stream_t s) {                  \
        return hyd::vec1d<T, 1>(n, alpha, X, incx, Y, incy, S(s), "hyd_axpy1d");                                              \
    }                                                                                                                         \
    int hyd_scale1d_##sfx(int64_t n, T alpha, T* X, int64_t incx, hyd_stream_t s) {                                           \
        return hyd::vec1d<T, 2>(n, alpha, nullptr, 1, X, incx, S(s), "hyd_scale1d");                                          \
    }                                                                                                                         \
    int hyd_axpy_dist_##sfx(hyd_grid* g, T alpha, const hyd_dm* X, const hyd_dm* Y, int transpose, hyd_stream_t s) {          \
        if (!g) return hyd::set_error("hyd_axpy_dist: null grid");                                                            \
        if (hyd::check_dm(g->g->shape, X, "hyd_axpy_dist(X)") || hyd::check_dm(g->g->shape, Y, "hyd_axpy_dist(Y)")) return 1; \
        return hyd::axpy_dist<T>(*g->g, alpha, hyd::to_dm(X), hyd::to_dm(Y), transpose != 0, S(s));                           \
    }

HYD_LEVEL1(double, f64)
HYD_LEVEL1(float, f32)
#undef HYD_LEVEL1

}  // extern "C"

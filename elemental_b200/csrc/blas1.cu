// Local BLAS-1-like kernels on column-major matrices with leading dimensions:
// out-of-place transpose, 2-D copy, 2-D axpy, 2-D scale.
//
// Reference interfaces replaced (seam S2):
//   gpu_blas::Copy(TRANSPOSE, m, n, A, lda, B, ldb, si) -> cublas?geam(alpha=1, beta=0)
//       include/hydrogen/blas/GPU_BLAS_impl.hpp:152-172,558-567; caller El::Transpose(Matrix<T,GPU>)
//       include/El/blas_like/level1/Transpose.hpp:161-181  (half: transpose_kernel, src/hydrogen/blas/gpu/Transpose.cu:15-81)
//   gpu_blas::Copy(NORMAL, ...) -> geam / copy_2d_kernel (src/hydrogen/blas/gpu/Copy.cu:24-69)
//   gpu_blas::Axpy 2-D -> geam / axpy_2d_tiled_kernel (src/hydrogen/blas/gpu/Axpy.cu:70-98); B += alpha*A
//   gpu_blas::Scale 2-D -> geam / scale_2d_kernel_naive (src/hydrogen/blas/gpu/Scale.cu:21-32)
//
// All four are HBM-bound (2 x bytes of the matrix, 3 x for axpy).  Grids are persistent
// (SMs x resident CTAs) with grid-stride tile loops; every thread keeps >= 64 B of
// independent loads in flight before its first store.
#include "grid.h"
#include "hyd_internal.h"
#include <algorithm>

#include "lattice.cuh"

namespace hyd {

// ---------------------------------------------------------------------------
// transpose: B(j,i) = A(i,j), A is m x n (lda), B is n x m (ldb).  Bit-exact copy.
// It is the transposing lattice with unit steps, so it runs on lattice_copy_kernel's transposing path
// (lattice.cuh: 128 x 32 / 256 x 32 tile staged transposed in shared memory, 16-byte stores): 95 % of the
// measured HBM copy rate at 32768^2 fp64, against 81 % for the 64 x 64 scalar-store tile it replaces.
// ---------------------------------------------------------------------------

// ---------------------------------------------------------------------------
// elementwise 2-D kernels: OP 0 copy (B = A), 1 axpy (B += alpha*A), 2 scale (B *= alpha; A unused)
// Each CTA pass covers a strip of 2048 rows x 1 column... generalised: a "chunk" is 256 threads x VEC
// elements along the column; 4 chunks (columns or row-blocks) are in flight per thread.
// ---------------------------------------------------------------------------
template <typename T> struct V16;
template <> struct V16<double> { typedef double2 type; static constexpr int N = 2; };
template <> struct V16<float> { typedef float4 type; static constexpr int N = 4; };

template <typename T, int OP>
__device__ __forceinline__ T ew(T alpha, T a, T b) {
    if (OP == 0) return a;
    if (OP == 1) return b + alpha * a;  // one rounding of alpha*a is NOT guaranteed (FMA contraction), as in the reference (Axpy.cu:90)
    return (alpha == T(0)) ? T(0) : alpha * b;
}

template <typename T, int OP, bool VEC>
__global__ void __launch_bounds__(256) ew2d_kernel(i64 m, i64 n, T alpha, const T* __restrict__ A, i64 lda,
                                                   T* __restrict__ B, i64 ldb) {
    typedef typename V16<T>::type VT;
    constexpr int V = VEC ? V16<T>::N : 1;
    constexpr int U = 4;                       // independent accesses per thread
    const i64 mv = VEC ? m / V : m;            // vector rows (VEC requires m % V == 0)
    const i64 row_chunks = (mv + 255) / 256;
    const i64 col_groups = (n + U - 1) / U;
    const i64 total = row_chunks * col_groups;
    for (i64 t = blockIdx.x; t < total; t += gridDim.x) {
        const i64 rc = t % row_chunks, cg = t / row_chunks;
        const i64 r = rc * 256 + threadIdx.x;
        if (r >= mv) continue;
        const i64 j0 = cg * U;
        if (VEC) {
            VT a[U], b[U];
#pragma unroll
            for (int q = 0; q < U; ++q) {
                if (j0 + q < n) {
                    if (OP != 2) a[q] = __ldg(reinterpret_cast<const VT*>(A + (j0 + q) * lda) + r);
                    if (OP != 0) b[q] = reinterpret_cast<const VT*>(B + (j0 + q) * ldb)[r];
                }
            }
#pragma unroll
            for (int q = 0; q < U; ++q) {
                if (j0 + q < n) {
                    VT o;
                    T* op = reinterpret_cast<T*>(&o);
                    const T* ap = reinterpret_cast<const T*>(&a[q]);
                    const T* bp = reinterpret_cast<const T*>(&b[q]);
#pragma unroll
                    for (int e = 0; e < V; ++e) op[e] = ew<T, OP>(alpha, OP != 2 ? ap[e] : T(0), OP != 0 ? bp[e] : T(0));
                    reinterpret_cast<VT*>(B + (j0 + q) * ldb)[r] = o;
                }
            }
        } else {
            T a[U], b[U];
#pragma unroll
            for (int q = 0; q < U; ++q) {
                if (j0 + q < n) {
                    if (OP != 2) a[q] = __ldg(A + r + (j0 + q) * lda);
                    if (OP != 0) b[q] = B[r + (j0 + q) * ldb];
                }
            }
#pragma unroll
            for (int q = 0; q < U; ++q)
                if (j0 + q < n) B[r + (j0 + q) * ldb] = ew<T, OP>(alpha, OP != 2 ? a[q] : T(0), OP != 0 ? b[q] : T(0));
        }
    }
}

template <typename T, int OP>
static int launch_ew(i64 m, i64 n, T alpha, const T* A, i64 lda, T* B, i64 ldb, cudaStream_t s, const char* who) {
    if (m < 0 || n < 0) return set_error("%s: negative dimension", who);
    if (m == 0 || n == 0) return 0;
    if ((OP != 2 && lda < m) || ldb < m) return set_error("%s: leading dimension too small", who);
    constexpr int V = V16<T>::N;
    const bool vec = (m % V == 0) && (ldb % V == 0) && ((uintptr_t)B % 16 == 0) &&
                     (OP == 2 || ((lda % V == 0) && ((uintptr_t)A % 16 == 0)));
    const i64 mv = vec ? m / V : m;
    const i64 total = ((mv + 255) / 256) * ((n + 3) / 4);
    const unsigned grid = (unsigned)std::max<i64>(1, std::min<i64>(total, (i64)num_sms() * 8));
    if (vec) ew2d_kernel<T, OP, true><<<grid, 256, 0, s>>>(m, n, alpha, A, lda, B, ldb);
    else     ew2d_kernel<T, OP, false><<<grid, 256, 0, s>>>(m, n, alpha, A, lda, B, ldb);
    ++g_launches;
    return check_cuda(cudaGetLastError(), who);
}

template <typename T>
int transpose_local(i64 m, i64 n, const T* A, i64 lda, T* B, i64 ldb, cudaStream_t s) {
    if (m < 0 || n < 0) return set_error("hyd_transpose: negative dimension");
    if (m == 0 || n == 0) return 0;
    if (lda < m || ldb < n) return set_error("hyd_transpose: leading dimension too small");
    CopyBatch b;
    b.n = 1;
    CopyDesc& d = b.d[0];
    d.src = A;
    d.dst = B;
    d.s_rs = lda;  // logical row a = column j of A = row j of B
    d.s_cs = 1;    // logical col b = row i of A = column i of B
    d.d_rs = 1;
    d.d_cs = ldb;
    d.nr = n;
    d.nc = m;
    const i64 ta = 64 * (16 / (i64)sizeof(T));
    const i64 tiles = ((n + ta - 1) / ta) * ((m + 31) / 32);
    const unsigned grid = (unsigned)std::max<i64>(1, std::min<i64>(tiles, (i64)num_sms() * 8));
    lattice_copy_kernel<T><<<grid, 256, 0, s>>>(b);
    ++g_launches;
    return check_cuda(cudaGetLastError(), "hyd_transpose (lattice_copy_kernel)");
}
template <typename T>
int copy2d_local(i64 m, i64 n, const T* A, i64 lda, T* B, i64 ldb, cudaStream_t s) {
    return launch_ew<T, 0>(m, n, T(0), A, lda, B, ldb, s, "hyd_copy2d");
}
template <typename T>
int axpy2d_local(i64 m, i64 n, T alpha, const T* A, i64 lda, T* B, i64 ldb, cudaStream_t s) {
    return launch_ew<T, 1>(m, n, alpha, A, lda, B, ldb, s, "hyd_axpy2d");
}
template <typename T>
int scale_local(i64 m, i64 n, T beta, T* C, i64 ldc, cudaStream_t s) {
    if (beta == T(1)) return 0;
    return launch_ew<T, 2>(m, n, beta, nullptr, 0, C, ldc, s, "hyd_scale2d");
}

// ---------------------------------------------------------------------------
// Trapezoid kernels: ScaleTrapezoid / AxpyTrapezoid / MakeTrapezoidal on a local block whose element (i,j) is global
// (gi0 + i*rs, gj0 + j*cs) -- a plain Matrix is gi0 = gj0 = 0, rs = cs = 1; a DistMatrix block passes its shifts and strides
// (include/El/blas_like/level1/{ScaleTrapezoid,AxpyTrapezoid,MakeTrapezoidal}.hpp).  LOWER holds the entries with
// gj - gi <= offset, UPPER those with gj - gi >= offset.  OP 1: Y += alpha X inside; 2: Y *= alpha inside (exactly the
// reference's `*=`, so alpha = 0 keeps NaNs); 3: Y = 0 outside.  Tiles of 256 rows x 4 columns; a tile wholly on the
// untouched side is skipped before any memory access, so the traffic is that of the trapezoid, not of the block.
// ---------------------------------------------------------------------------
template <typename T, int OP>
__global__ void __launch_bounds__(256) trapezoid_kernel(int upper, i64 m, i64 n, i64 gi0, i64 rs, i64 gj0, i64 cs, i64 offset,
                                                        T alpha, const T* __restrict__ X, i64 ldx, T* __restrict__ Y, i64 ldy) {
    constexpr int U = 4;
    const i64 row_chunks = (m + 255) / 256, col_groups = (n + U - 1) / U;
    const i64 total = row_chunks * col_groups;
    for (i64 t = blockIdx.x; t < total; t += gridDim.x) {
        const i64 rc = t % row_chunks, cg = t / row_chunks;
        const i64 i0 = rc * 256, i1 = min(m, i0 + 256) - 1, j0 = cg * U, j1 = min(n, j0 + U) - 1;
        // d = gj - gi over the tile ranges from dmin (last row, first column) to dmax (first row, last column)
        const i64 dmin = (gj0 + j0 * cs) - (gi0 + i1 * rs), dmax = (gj0 + j1 * cs) - (gi0 + i0 * rs);
        const bool none_inside = upper ? dmax < offset : dmin > offset;
        const bool all_inside = upper ? dmin >= offset : dmax <= offset;
        if (OP == 3 ? all_inside : none_inside) continue;
        const i64 i = i0 + threadIdx.x;
        if (i > i1) continue;
        const i64 gi = gi0 + i * rs;
#pragma unroll
        for (int q = 0; q < U; ++q) {
            const i64 j = j0 + q;
            if (j > j1) break;
            const i64 d = (gj0 + j * cs) - gi;
            const bool inside = upper ? d >= offset : d <= offset;
            T* y = Y + i + j * ldy;
            if (OP == 1) { if (inside) *y = *y + alpha * __ldg(X + i + j * ldx); }
            else if (OP == 2) { if (inside) *y = *y * alpha; }
            else { if (!inside) *y = T(0); }
        }
    }
}

template <typename T>
int trapezoid_local(int op, int upper, i64 m, i64 n, i64 gi0, i64 rs, i64 gj0, i64 cs, i64 offset, T alpha, const T* X, i64 ldx,
                    T* Y, i64 ldy, cudaStream_t s) {
    if (m < 0 || n < 0) return set_error("trapezoid: negative dimension");
    if (m == 0 || n == 0) return 0;
    if (ldy < m || (op == 1 && ldx < m)) return set_error("trapezoid: leading dimension too small");
    const i64 total = ((m + 255) / 256) * ((n + 3) / 4);
    const unsigned grid = (unsigned)std::max<i64>(1, std::min<i64>(total, (i64)num_sms() * 8));
    if (op == 1) trapezoid_kernel<T, 1><<<grid, 256, 0, s>>>(upper, m, n, gi0, rs, gj0, cs, offset, alpha, X, ldx, Y, ldy);
    else if (op == 2) trapezoid_kernel<T, 2><<<grid, 256, 0, s>>>(upper, m, n, gi0, rs, gj0, cs, offset, alpha, X, ldx, Y, ldy);
    else trapezoid_kernel<T, 3><<<grid, 256, 0, s>>>(upper, m, n, gi0, rs, gj0, cs, offset, alpha, X, ldx, Y, ldy);
    ++g_launches;
    return check_cuda(cudaGetLastError(), "trapezoid_kernel");
}

#define HYD_INST(T)                                                                             \
    template int trapezoid_local<T>(int, int, i64, i64, i64, i64, i64, i64, i64, T, const T*, i64, T*, i64, cudaStream_t); \
    template int transpose_local<T>(i64, i64, const T*, i64, T*, i64, cudaStream_t);           \
    template int copy2d_local<T>(i64, i64, const T*, i64, T*, i64, cudaStream_t);              \
    template int axpy2d_local<T>(i64, i64, T, const T*, i64, T*, i64, cudaStream_t);           \
    template int scale_local<T>(i64, i64, T, T*, i64, cudaStream_t);
HYD_INST(double)
HYD_INST(float)
#undef HYD_INST

}  // namespace hyd

// Distributed GEMM on fp16 / bf16 storage with fp32 accumulation: C := alpha op(A) op(B) + beta C   (SURVEY 8f N3).
//
// Reference: El::Gemm on DistMatrix<gpu_half_type,...,Device::GPU> -- the same SUMMA loops as every other type
// (src/blas_like/level3/Gemm.cpp:278-307, Gemm/{NN,NT,TN,TT}.hpp) with gpu_blas::Gemm -> cublasHgemm / cublasGemmEx as the
// local product (Gemm.cpp:460-467, src/core/imports/cublas.cpp:213; HYDROGEN_GPU_USE_FP16).  C is stored in half precision,
// so it is rounded once per k-panel there and here alike.
//
// First version, deliberately plain: stationary-C only (the variant LBANN's shapes select), panels gathered with the generic
// one-hop redistribution on 2-byte elements (redistribute<h16>: bits are moved, never interpreted) and multiplied by the
// tcgen05 kind::f16 kernel (lgemm_f16_tc.cu).  No panel double-buffering yet: at ~1.5 PFLOP/s the local product is so short
// that the gathers dominate on more than one GPU; overlapping them is the next step once this has run on a GPU.
// Parity-tested on a B200 since round 2 (tests/test_gpu_half.py: four grids, every orientation, proxied operands).
#include "grid.h"
#include "hyd_internal.h"
#include "lattice.cuh"
#include <algorithm>
#include <vector>

namespace hyd {

namespace {

inline i64 pad8(i64 lh) { return std::max<i64>(8, (lh + 7) / 8 * 8); }   // 16-byte columns of 2-byte elements

// DistMatrixReadProxy<T,T,MC,MR> (include/El/core/Proxy.hpp:174-260): `a` itself when it is already [MC,MR], else a copy
int proxy_mcmr(Grid& g, TempScope& pool, const DM& a, DM& out, cudaStream_t s) {
    if (a.colDist == MC && a.rowDist == MR) { out = a; return 0; }
    out = DM();
    out.height = a.height; out.width = a.width; out.colDist = MC; out.rowDist = MR;
    out.colAlign = adopted_align(g.shape, a.colDist, a.colAlign, MC);   // as the reference's proxies do through B = A
    out.rowAlign = adopted_align(g.shape, a.rowDist, a.rowAlign, MR);
    const LocalShape L = local_shape(g.shape, out, g.mc, g.mr);
    out.ldim = pad8(L.lh);
    out.buf = pool.take((size_t)(out.ldim * std::max<i64>(L.lw, 1)) * 2);
    if (!out.buf) return 1;
    return redistribute<h16>(g, a, out, false, 3, s);
}

int lgemm_h(int bf16, int tA, int tB, i64 m, i64 n, i64 k, float alpha, const void* A, i64 lda, const void* B, i64 ldb, float beta,
            void* C, i64 ldc, cudaStream_t s) {
    ++g_launches;
    return bf16 ? lgemm_bf16(tA, tB, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, s)
                : lgemm_f16(tA, tB, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, s);
}

}  // namespace

int dist_gemm_half(Grid& g, int bf16, int oA, int oB, int alg, float alpha, const DM& A0, const DM& B0, float beta, const DM& C0,
                   i64 blocksize, cudaStream_t s) {
    oA = oA ? 1 : 0;
    oB = oB ? 1 : 0;
    const i64 m = C0.height, n = C0.width;
    const i64 k = oA ? A0.height : A0.width;
    const i64 am = oA ? A0.width : A0.height, bk = oB ? B0.width : B0.height, bn = oB ? B0.height : B0.width;
    if (am != m || bn != n || bk != k)
        return set_error("hyd_gemm(half): nonconformal matrices: op(A) %lldx%lld, op(B) %lldx%lld, C %lldx%lld", am, k, bk, bn, m, n);
    if (blocksize <= 0) return set_error("hyd_gemm(half): blocksize must be positive");
    if (alg != 0 && alg != 3) return set_error("hyd_gemm(half): only GEMM_DEFAULT / GEMM_SUMMA_C are provided in half precision");
    if (m == 0 || n == 0) return 0;
    const GridShape& sh = g.shape;
    TempScope pool(g, s);
    DM A, B, C;
    if (proxy_mcmr(g, pool, C0, C, s) || proxy_mcmr(g, pool, A0, A, s) || proxy_mcmr(g, pool, B0, B, s)) return 1;
    const LocalShape Lc = local_shape(sh, C, g.mc, g.mr);
    char* cbuf = static_cast<char*>(C.buf);
    if (k == 0) {   // Scale(beta, C) only (Gemm.cpp:287): the generic kernel with an empty sum
        if (Lc.lh > 0 && Lc.lw > 0 && lgemm_h(bf16, 0, 0, Lc.lh, Lc.lw, 0, 0.f, cbuf, C.ldim, cbuf, C.ldim, beta, cbuf, C.ldim, s)) return 1;
    }
    for (i64 k0 = 0; k0 < k; k0 += blocksize) {
        const i64 kb = std::min(blocksize, k - k0);
        const float b = k0 == 0 ? beta : 1.f;   // Scale(beta, C) folded into the first panel's update
        if (sh.size() == 1) {
            // 1x1 grid: the panels are views of the operands (RowAllGather.hpp:35-38 degenerates to local copies)
            const char* a = static_cast<const char*>(A.buf) + 2 * (oA ? k0 : k0 * A.ldim);
            const char* bp = static_cast<const char*>(B.buf) + 2 * (oB ? k0 * B.ldim : k0);
            if (lgemm_h(bf16, oA, oB, m, n, kb, alpha, a, A.ldim, bp, B.ldim, b, cbuf, C.ldim, s)) return 1;
            continue;
        }
        // op(A)(:,K1) -> [MC,*] and op(B)(K1,:)^T -> [MR,*], both aligned with C: one redistribution each, transposes folded in
        const DM A1 = oA ? view(sh, A, g.mc, g.mr, k0, 0, kb, m, 2) : view(sh, A, g.mc, g.mr, 0, k0, m, kb, 2);
        const DM B1 = oB ? view(sh, B, g.mc, g.mr, 0, k0, n, kb, 2) : view(sh, B, g.mc, g.mr, k0, 0, kb, n, 2);
        DM Ap, Bp;
        Ap.height = m; Ap.width = kb; Ap.colDist = MC; Ap.rowDist = STAR; Ap.colAlign = C.colAlign; Ap.rowAlign = 0;
        Bp.height = n; Bp.width = kb; Bp.colDist = MR; Bp.rowDist = STAR; Bp.colAlign = C.rowAlign; Bp.rowAlign = 0;
        Ap.ldim = pad8(Lc.lh);
        Bp.ldim = pad8(Lc.lw);
        if (ws_reserve(g.panel[0], (size_t)(Ap.ldim * blocksize) * 2) || ws_reserve(g.panel[2], (size_t)(Bp.ldim * blocksize) * 2)) return 1;
        Ap.buf = g.panel[0].buf;
        Bp.buf = g.panel[2].buf;
        if (redistribute<h16>(g, A1, Ap, oA != 0, 0, s)) return 1;
        if (redistribute<h16>(g, B1, Bp, oB == 0, 1, s)) return 1;
        if (Lc.lh > 0 && Lc.lw > 0 && lgemm_h(bf16, 0, 1, Lc.lh, Lc.lw, kb, alpha, Ap.buf, Ap.ldim, Bp.buf, Bp.ldim, b, cbuf, C.ldim, s))
            return 1;
    }
    if (C.buf != C0.buf && redistribute<h16>(g, C, C0, false, 3, s)) return 1;   // write-back half of the ReadWrite proxy
    return 0;
}

}  // namespace hyd

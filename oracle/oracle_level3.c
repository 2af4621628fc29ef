/* oracle_level3.c -- El::Trrk / El::Syrk and the trapezoid level-1 routines on the simulated grid.
 * TEST INFRASTRUCTURE ONLY (see oracle.h).
 *
 * Follows, loop for loop:
 *   ScaleTrapezoid   include/El/blas_like/level1/ScaleTrapezoid.hpp:47-92   (DistMatrix form: LocalRowOffset per column)
 *   AxpyTrapezoid    include/El/blas_like/level1/AxpyTrapezoid.hpp:54-100
 *   MakeTrapezoidal  include/El/blas_like/level1/MakeTrapezoidal.hpp:48-90
 *   LocalTrrk        src/blas_like/level3/Trrk/Local.hpp:640-830  (recursion on halves of C down to
 *                    g.Width()*LocalTrrkBlocksize, then LocalTrrkKernel :167-215: one LocalGemm on the off-diagonal
 *                    quadrant, the two diagonal quadrants through a temporary + AxpyTrapezoid)
 *   Trrk             src/blas_like/level3/Trrk.cpp:98-115 and Trrk/{NN,NT,TN,TT}.hpp (SUMMA-C panels)
 *   Syrk             src/blas_like/level3/Syrk.cpp:72-90 and Syrk/{LN,LT,UN,UT}.hpp
 * Temporaries are redistributed with the closed-form definition (orc_redist_closed), as in oracle_gemm.c. */
#include "oracle.h"
#include <stdlib.h>
#include <string.h>

static int64_t imin(int64_t a, int64_t b) { return a < b ? a : b; }
static int64_t imax(int64_t a, int64_t b) { return a > b ? a : b; }
static int g_local_trrk_blocksize = 64;   /* src/blas_like/blocksizes.cpp:26 */
void orc_set_local_trrk_blocksize(int b) { g_local_trrk_blocksize = b > 0 ? b : 64; }

static void scal(void* p, int es, double alpha) {
    if (es == 8) *(double*)p *= alpha; else *(float*)p *= (float)alpha;
}
static void axpy1(void* y, const void* x, int es, double alpha) {
    if (es == 8) { volatile double t = alpha * *(const double*)x; *(double*)y += t; }
    else { volatile float t = (float)alpha * *(const float*)x; *(float*)y += t; }
}

/* ScaleTrapezoid.hpp:47-92.  UPPER: rows 0 .. LocalRowOffset(min(j-offset+1, height)) of each local column;
 * LOWER: rows from LocalRowOffset(max(j-offset, 0)) down. */
void orc_scale_trapezoid(double alpha, int uplo, orc_dm* A, int64_t offset) {
    if (alpha == 1.0) return;
    for (int p = 0; p < A->P; ++p) {
        const int64_t cs = orc_dist_stride(A->g, A->colDist), rs = orc_dist_stride(A->g, A->rowDist);
        for (int64_t jl = 0; jl < A->lw[p]; ++jl) {
            const int64_t j = A->rowShift[p] + jl * rs;
            char* col = (char*)A->loc[p] + jl * A->ld[p] * A->es;
            if (uplo == 1) {
                const int64_t boundary = imin(j - offset + 1, A->m);
                const int64_t numRows = orc_length(boundary, A->colShift[p], cs);
                for (int64_t il = 0; il < numRows; ++il) scal(col + il * A->es, A->es, alpha);
            } else {
                const int64_t firstRow = imax(j - offset, 0);
                const int64_t numZeroRows = orc_length(firstRow, A->colShift[p], cs);
                for (int64_t il = numZeroRows; il < A->lh[p]; ++il) scal(col + il * A->es, A->es, alpha);
            }
        }
    }
}

/* AxpyTrapezoid.hpp:54-100: X and Y share distribution and alignments */
int orc_axpy_trapezoid(int uplo, double alpha, const orc_dm* X, orc_dm* Y, int64_t offset) {
    if (X->m != Y->m || X->n != Y->n) return -1;
    if (X->colDist != Y->colDist || X->rowDist != Y->rowDist || X->colAlign != Y->colAlign || X->rowAlign != Y->rowAlign) return -2;
    const int es = Y->es;
    for (int p = 0; p < Y->P; ++p) {
        const int64_t cs = orc_dist_stride(Y->g, Y->colDist), rs = orc_dist_stride(Y->g, Y->rowDist);
        for (int64_t jl = 0; jl < Y->lw[p]; ++jl) {
            const int64_t j = Y->rowShift[p] + jl * rs;
            const char* xc = (const char*)X->loc[p] + jl * X->ld[p] * es;
            char* yc = (char*)Y->loc[p] + jl * Y->ld[p] * es;
            int64_t lo, hi;
            if (uplo == 1) { lo = 0; hi = orc_length(imax(imin(j + 1 - offset, Y->m), 0), Y->colShift[p], cs); }
            else { lo = orc_length(imax(imin(j - offset, Y->m), 0), Y->colShift[p], cs); hi = Y->lh[p]; }
            for (int64_t il = lo; il < hi; ++il) axpy1(yc + il * es, xc + il * es, es, alpha);
        }
    }
    return 0;
}

/* MakeTrapezoidal.hpp:48-90: zero everything outside the trapezoid */
void orc_make_trapezoidal(int uplo, orc_dm* A, int64_t offset) {
    for (int p = 0; p < A->P; ++p) {
        const int64_t cs = orc_dist_stride(A->g, A->colDist), rs = orc_dist_stride(A->g, A->rowDist);
        for (int64_t jl = 0; jl < A->lw[p]; ++jl) {
            const int64_t j = A->rowShift[p] + jl * rs;
            char* col = (char*)A->loc[p] + jl * A->ld[p] * A->es;
            if (uplo == 0) {
                const int64_t boundary = imin(imax(j - offset, 0), A->m);       /* rows < j - offset are zeroed */
                const int64_t numZeroRows = orc_length(boundary, A->colShift[p], cs);
                memset(col, 0, (size_t)(numZeroRows * A->es));
            } else {
                const int64_t firstZeroRow = imax(j - offset + 1, 0);
                const int64_t n0 = orc_length(imin(firstZeroRow, A->m), A->colShift[p], cs);
                if (n0 < A->lh[p]) memset(col + n0 * A->es, 0, (size_t)((A->lh[p] - n0) * A->es));
            }
        }
    }
}

/* LocalGemm on every rank, dimensions taken from the local blocks (Gemm.cpp:323-431) */
static void local_gemm_all(int tA, int tB, double alpha, const orc_dm* A, const orc_dm* B, double beta, orc_dm* C) {
    for (int p = 0; p < C->P; ++p) {
        const int64_t k = tA ? A->lh[p] : A->lw[p];
        orc_local_gemm(tA, tB, C->lh[p], C->lw[p], k, alpha, A->loc[p], A->ld[p], B->loc[p], B->ld[p], beta, C->loc[p],
                       C->ld[p], C->es);
    }
}

/* rows [i0,i1) of op(A) and columns [j0,j1) of op(B) as views of the gathered panels:
 * tA = 0: A is [MC,*] (n x k), else [*,MC] (k x n); tB = 0: B is [*,MR] (k x n), else [MR,*] (n x k) */
static orc_dm* rows_of_opA(orc_dm* A, int tA, int64_t i0, int64_t i1) {
    return tA ? orc_dm_view(A, 0, i0, A->m, i1 - i0) : orc_dm_view(A, i0, 0, i1 - i0, A->n);
}
static orc_dm* cols_of_opB(orc_dm* B, int tB, int64_t j0, int64_t j1) {
    return tB ? orc_dm_view(B, j0, 0, j1 - j0, B->n) : orc_dm_view(B, 0, j0, B->m, j1 - j0);
}

/* the diagonal quadrant through a temporary aligned with it, then AxpyTrapezoid (Local.hpp:200-210) */
static void diag_block(int uplo, int tA, int tB, double alpha, orc_dm* A, orc_dm* B, orc_dm* C) {
    orc_dm* D = orc_dm_create(C->g, C->m, C->n, ORC_MC, ORC_MR, C->colAlign, C->rowAlign, C->es);
    local_gemm_all(tA, tB, alpha, A, B, 0.0, D);
    orc_axpy_trapezoid(uplo, 1.0, D, C, 0);
    orc_dm_free(D);
}

/* LocalTrrk (Local.hpp:640-830): C is a diagonal block of an [MC,MR] matrix, A / B the matching rows / columns */
static void local_trrk(int uplo, int tA, int tB, double alpha, orc_dm* A, orc_dm* B, orc_dm* C) {
    const int64_t n = C->m, half = n / 2;
    orc_dm *AT = rows_of_opA(A, tA, 0, half), *AB = rows_of_opA(A, tA, half, n);
    orc_dm *BL = cols_of_opB(B, tB, 0, half), *BR = cols_of_opB(B, tB, half, n);
    orc_dm *CTL = orc_dm_view(C, 0, 0, half, half), *CBR = orc_dm_view(C, half, half, n - half, n - half);
    if (uplo == 0) {
        orc_dm* CBL = orc_dm_view(C, half, 0, n - half, half);
        local_gemm_all(tA, tB, alpha, AB, BL, 1.0, CBL);
        orc_dm_free(CBL);
    } else {
        orc_dm* CTR = orc_dm_view(C, 0, half, half, n - half);
        local_gemm_all(tA, tB, alpha, AT, BR, 1.0, CTR);
        orc_dm_free(CTR);
    }
    if (n < (int64_t)C->g.c * g_local_trrk_blocksize) {   /* LocalTrrkKernel */
        diag_block(uplo, tA, tB, alpha, AT, BL, CTL);
        diag_block(uplo, tA, tB, alpha, AB, BR, CBR);
    } else {                                              /* recurse */
        local_trrk(uplo, tA, tB, alpha, AT, BL, CTL);
        local_trrk(uplo, tA, tB, alpha, AB, BR, CBR);
    }
    orc_dm_free(AT); orc_dm_free(AB); orc_dm_free(BL); orc_dm_free(BR); orc_dm_free(CTL); orc_dm_free(CBR);
}

#define TMP(m, n, cd, rd, ca, ra) orc_dm_create(g, m, n, cd, rd, ca, ra, es)
static orc_dm* mcmr_proxy(orc_dm* X) { return orc_read_proxy(X, ORC_MC, ORC_MR, 0, 0, 0, 0); }   /* DistMatrixReadProxy<T,T,MC,MR> */

/* El::Trrk on DistMatrices, Trrk.cpp:98-115 */
int orc_trrk(int uplo, int oA, int oB, double alpha, orc_dm* A0, orc_dm* B0, double beta, orc_dm* C0, int64_t bsize) {
    oA = oA ? 1 : 0; oB = oB ? 1 : 0;
    const orc_grid g = C0->g; const int es = C0->es;
    const int64_t n = C0->m, k = oA ? A0->m : A0->n;
    if (C0->n != n || (oA ? A0->n : A0->m) != n || (oB ? B0->m : B0->n) != n || (oB ? B0->n : B0->m) != k) return -1;
    orc_scale_trapezoid(beta, uplo, C0, 0);
    orc_dm *A = mcmr_proxy(A0), *B = mcmr_proxy(B0), *C = mcmr_proxy(C0);
    for (int64_t k0 = 0; k0 < k; k0 += bsize) {
        const int64_t nb = imin(bsize, k - k0);
        orc_dm *A1 = oA ? orc_dm_view(A, k0, 0, nb, n) : orc_dm_view(A, 0, k0, n, nb);
        orc_dm *B1 = oB ? orc_dm_view(B, 0, k0, n, nb) : orc_dm_view(B, k0, 0, nb, n);
        /* A1[MC,*] (N) or A1[*,MC] (T); B1^T[MR,*] (N: NN.hpp:49-52, TN.hpp:45-49) or B1^T[*,MR] via [VR,*] (T: NT.hpp:47-50) */
        orc_dm* Ap = oA ? TMP(nb, n, ORC_STAR, ORC_MC, 0, C->colAlign) : TMP(n, nb, ORC_MC, ORC_STAR, C->colAlign, 0);
        orc_dm* Bp = oB ? TMP(nb, n, ORC_STAR, ORC_MR, 0, C->rowAlign) : TMP(n, nb, ORC_MR, ORC_STAR, C->rowAlign, 0);
        orc_redist_closed(A1, Ap, 0);
        orc_redist_closed(B1, Bp, 1);
        local_trrk(uplo, oA, oB ? 0 : 1, alpha, Ap, Bp, C);
        orc_dm_free(Ap); orc_dm_free(Bp); orc_dm_free(A1); orc_dm_free(B1);
    }
    if (C != C0) { orc_redist_closed(C, C0, 0); orc_dm_free(C); }
    if (A != A0) orc_dm_free(A);
    if (B != B0) orc_dm_free(B);
    return 0;
}

/* El::Syrk / El::Herk (real) on DistMatrices, Syrk.cpp:72-90; Syrk/LN.hpp:14-50 (UN identical with UPPER), LT.hpp:14-52 (UT) */
int orc_syrk(int uplo, int orient, double alpha, orc_dm* A0, double beta, orc_dm* C0, int64_t bsize) {
    orient = orient ? 1 : 0;
    const orc_grid g = C0->g; const int es = C0->es;
    const int64_t n = C0->m, k = orient ? A0->m : A0->n;
    if (C0->n != n || (orient ? A0->n : A0->m) != n) return -1;
    orc_scale_trapezoid(beta, uplo, C0, 0);
    orc_dm *A = mcmr_proxy(A0), *C = mcmr_proxy(C0);
    for (int64_t k0 = 0; k0 < k; k0 += bsize) {
        const int64_t nb = imin(bsize, k - k0);
        if (!orient) {   /* A1[MC,*]; A1[VR,*] -> A1^T[*,MR]; LocalTrrk(uplo, A1[MC,*], A1^T[*,MR]) */
            orc_dm* A1 = orc_dm_view(A, 0, k0, n, nb);
            orc_dm *Ap = TMP(n, nb, ORC_MC, ORC_STAR, C->colAlign, 0), *Bp = TMP(nb, n, ORC_STAR, ORC_MR, 0, C->rowAlign);
            orc_redist_closed(A1, Ap, 0);
            orc_redist_closed(A1, Bp, 1);
            local_trrk(uplo, 0, 0, alpha, Ap, Bp, C);
            orc_dm_free(Ap); orc_dm_free(Bp); orc_dm_free(A1);
        } else {         /* A1^T[MR,*]; A1[*,VR] -> A1[*,MC]; LocalTrrk(uplo, T, T, A1[*,MC], A1^T[MR,*]) */
            orc_dm* A1 = orc_dm_view(A, k0, 0, nb, n);
            orc_dm *Ap = TMP(nb, n, ORC_STAR, ORC_MC, 0, C->colAlign), *Bp = TMP(n, nb, ORC_MR, ORC_STAR, C->rowAlign, 0);
            orc_redist_closed(A1, Ap, 0);
            orc_redist_closed(A1, Bp, 1);
            local_trrk(uplo, 1, 1, alpha, Ap, Bp, C);
            orc_dm_free(Ap); orc_dm_free(Bp); orc_dm_free(A1);
        }
    }
    if (C != C0) { orc_redist_closed(C, C0, 0); orc_dm_free(C); }
    if (A != A0) orc_dm_free(A);
    return 0;
}

/* MakeSymmetric on an ElementalMatrix, MakeSymmetric.hpp:52-68: MakeTrapezoidal(uplo, A); ATrans = A^T (same distribution);
 * AxpyTrapezoid(UPPER, 1, ATrans, A, 1) for LOWER / AxpyTrapezoid(LOWER, 1, ATrans, A, -1) for UPPER */
int orc_make_symmetric(int uplo, orc_dm* A) {
    if (A->m != A->n) return -1;
    orc_make_trapezoidal(uplo, A, 0);
    orc_dm* At = orc_dm_create(A->g, A->m, A->n, A->colDist, A->rowDist, A->colAlign, A->rowAlign, A->es);
    orc_redist_closed(A, At, 1);
    const int rc = uplo == 0 ? orc_axpy_trapezoid(1, 1.0, At, A, 1) : orc_axpy_trapezoid(0, 1.0, At, A, -1);
    orc_dm_free(At);
    return rc;
}

/* El::Symm on DistMatrices, src/blas_like/level3/Symm.cpp:56-80: C *= beta, then symm::{LL,LU,RL,RU}.  side: 0 LEFT (C += alpha
 * A B, A m x m), 1 RIGHT (C += alpha B A, A n x n); only the `uplo` triangle of A is read.  Restated: the stationary-C
 * variants Symm/LL.hpp:153-214 (LLC), LU.hpp (LUC), RL.hpp (RLC), RU.hpp (RUC).  The reference routes tall problems to the
 * stationary-A variants (LL.hpp:226-229) and a one-column B to Symv (Symm.cpp:62-66); those differ from the C variants only
 * in the grouping of the sums, and GEMM-class results are compared by tolerance. */
int orc_symm(int side, int uplo, double alpha, orc_dm* A0, orc_dm* B0, double beta, orc_dm* C0, int64_t bsize) {
    const orc_grid g = C0->g; const int es = C0->es;
    const int64_t m = C0->m, n = C0->n, na = side == 0 ? m : n;
    if (A0->m != na || A0->n != na || B0->m != m || B0->n != n) return -1;
    orc_scale(beta, C0);
    orc_dm *A = mcmr_proxy(A0), *B = mcmr_proxy(B0), *C = mcmr_proxy(C0);
    for (int64_t k = 0; k < na; k += bsize) {
        const int64_t nb = imin(bsize, na - k);
        if (side == 0) {
            orc_dm* B1 = orc_dm_view(B, k, 0, nb, n);
            orc_dm* Bt = TMP(n, nb, ORC_MR, ORC_STAR, C->rowAlign, 0);            /* B1^T[MR,*] aligned with C */
            orc_redist_closed(B1, Bt, 1);
            orc_dm *CT = orc_dm_view(C, 0, 0, k + nb, n), *CB = orc_dm_view(C, k, 0, m - k, n);
            if (uplo == 0) {      /* LLC: AB1[MC,*] lower trapezoid -> CB;  A1L[*,MC] strictly-left part -> CT */
                orc_dm *A1L = orc_dm_view(A, k, 0, nb, k + nb), *AB1 = orc_dm_view(A, k, k, m - k, nb);
                orc_dm *P = TMP(m - k, nb, ORC_MC, ORC_STAR, CB->colAlign, 0), *Q = TMP(nb, k + nb, ORC_STAR, ORC_MC, 0, CT->colAlign);
                orc_redist_closed(AB1, P, 0); orc_redist_closed(A1L, Q, 0);
                orc_make_trapezoidal(0, P, 0); orc_make_trapezoidal(0, Q, k - 1);
                local_gemm_all(0, 1, alpha, P, Bt, 1.0, CB);
                local_gemm_all(1, 1, alpha, Q, Bt, 1.0, CT);
                orc_dm_free(P); orc_dm_free(Q); orc_dm_free(A1L); orc_dm_free(AB1);
            } else {              /* LUC: AT1[MC,*] upper trapezoid (offset -k) -> CT;  A1R[*,MC] strictly-right part -> CB */
                orc_dm *A1R = orc_dm_view(A, k, k, nb, m - k), *AT1 = orc_dm_view(A, 0, k, k + nb, nb);
                orc_dm *P = TMP(k + nb, nb, ORC_MC, ORC_STAR, CT->colAlign, 0), *Q = TMP(nb, m - k, ORC_STAR, ORC_MC, 0, CB->colAlign);
                orc_redist_closed(AT1, P, 0); orc_redist_closed(A1R, Q, 0);
                orc_make_trapezoidal(1, P, -k); orc_make_trapezoidal(1, Q, 1);
                local_gemm_all(0, 1, alpha, P, Bt, 1.0, CT);
                local_gemm_all(1, 1, alpha, Q, Bt, 1.0, CB);
                orc_dm_free(P); orc_dm_free(Q); orc_dm_free(A1R); orc_dm_free(AT1);
            }
            orc_dm_free(CT); orc_dm_free(CB); orc_dm_free(Bt); orc_dm_free(B1);
        } else {
            orc_dm* B1 = orc_dm_view(B, 0, k, m, nb);
            orc_dm* Bm = TMP(m, nb, ORC_MC, ORC_STAR, C->colAlign, 0);            /* B1[MC,*] aligned with C */
            orc_redist_closed(B1, Bm, 0);
            orc_dm *CL = orc_dm_view(C, 0, 0, m, k + nb), *CR = orc_dm_view(C, 0, k, m, n - k);
            if (uplo == 0) {      /* RLC: A1L^T[MR,*] upper (offset -k) -> CL;  AB1^T[*,MR] strictly-upper -> CR */
                orc_dm *A1L = orc_dm_view(A, k, 0, nb, k + nb), *AB1 = orc_dm_view(A, k, k, n - k, nb);
                orc_dm *P = TMP(k + nb, nb, ORC_MR, ORC_STAR, CL->rowAlign, 0), *Q = TMP(nb, n - k, ORC_STAR, ORC_MR, 0, CR->rowAlign);
                orc_redist_closed(A1L, P, 1); orc_redist_closed(AB1, Q, 1);
                orc_make_trapezoidal(1, P, -k); orc_make_trapezoidal(1, Q, 1);
                local_gemm_all(0, 1, alpha, Bm, P, 1.0, CL);
                local_gemm_all(0, 0, alpha, Bm, Q, 1.0, CR);
                orc_dm_free(P); orc_dm_free(Q); orc_dm_free(A1L); orc_dm_free(AB1);
            } else {              /* RUC: A1R^T[MR,*] lower -> CR;  AT1^T[*,MR] strictly-lower (offset k-1) -> CL */
                orc_dm *A1R = orc_dm_view(A, k, k, nb, n - k), *AT1 = orc_dm_view(A, 0, k, k + nb, nb);
                orc_dm *P = TMP(n - k, nb, ORC_MR, ORC_STAR, CR->rowAlign, 0), *Q = TMP(nb, k + nb, ORC_STAR, ORC_MR, 0, CL->rowAlign);
                orc_redist_closed(A1R, P, 1); orc_redist_closed(AT1, Q, 1);
                orc_make_trapezoidal(0, P, 0); orc_make_trapezoidal(0, Q, k - 1);
                local_gemm_all(0, 1, alpha, Bm, P, 1.0, CR);
                local_gemm_all(0, 0, alpha, Bm, Q, 1.0, CL);
                orc_dm_free(P); orc_dm_free(Q); orc_dm_free(A1R); orc_dm_free(AT1);
            }
            orc_dm_free(CL); orc_dm_free(CR); orc_dm_free(Bm); orc_dm_free(B1);
        }
    }
    if (C != C0) { orc_redist_closed(C, C0, 0); orc_dm_free(C); }
    if (A != A0) orc_dm_free(A);
    if (B != B0) orc_dm_free(B);
    return 0;
}

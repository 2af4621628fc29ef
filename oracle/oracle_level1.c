/* oracle_level1.c -- CPU restatement of the distributed BLAS-1 operations around Gemm (TEST INFRASTRUCTURE ONLY).
 *   El::Axpy(alpha, ElementalMatrix X, ElementalMatrix Y)   include/El/blas_like/level1/Axpy.hpp:161-187
 *   El::TransposeAxpy(alpha, A, B)                          include/El/blas_like/level1/TransposeAxpy.hpp:181-210
 *   El::Hadamard(A, B, C) on AbstractDistMatrix              include/El/blas_like/level1/Hadamard.hpp:109-135
 *   El::Fill / El::Scale on the local blocks                 include/El/blas_like/level1/{Fill,Scale}.hpp
 * Pinning: as oracle.h -- the copy inside Axpy is the pinned bit-exact redistribution; the arithmetic is one
 * multiply and one add per element (Axpy.hpp:78-80 `YBuf[i+j*ldY] += alpha*XBuf[i+j*ldX]`), compiled here without
 * contraction.  Whether the reference fuses them (FMA) is compiler / cuBLAS-geam dependent: tests use operands whose
 * products are exact, or a one-ulp tolerance. */
#include "oracle.h"
#include <stdlib.h>
#include <string.h>

static void local_axpy(double alpha, int64_t h, int64_t w, const void* X, int64_t ldx, int transposeX, void* Y, int64_t ldy, int es) {
    for (int64_t j = 0; j < w; ++j)
        for (int64_t i = 0; i < h; ++i) {
            const int64_t xo = transposeX ? j + i * ldx : i + j * ldx;
            if (es == 8) { volatile double t = alpha * ((const double*)X)[xo]; double* q = (double*)Y + i + j * ldy; *q = *q + t; }
            else { volatile float t = (float)alpha * ((const float*)X)[xo]; float* q = (float*)Y + i + j * ldy; *q = *q + t; }
        }
}

/* Y += alpha X (transpose = 0) or Y += alpha X^T (transpose = 1).  Returns non-zero for nonconformal operands. */
int orc_axpy(double alpha, const orc_dm* X, orc_dm* Y, int transpose) {
    const int64_t xm = transpose ? X->n : X->m, xn = transpose ? X->m : X->n;
    if (xm != Y->m || xn != Y->n || X->es != Y->es) return 1;
    const int same = transpose ? (X->colDist == Y->rowDist && X->rowDist == Y->colDist && X->colAlign == Y->rowAlign && X->rowAlign == Y->colAlign)
                               : (X->colDist == Y->colDist && X->rowDist == Y->rowDist && X->colAlign == Y->colAlign && X->rowAlign == Y->rowAlign);
    if (same) {  /* Axpy.hpp:171-174 / TransposeAxpy.hpp:193-199: straight on the local blocks */
        for (int p = 0; p < Y->P; ++p) local_axpy(alpha, Y->lh[p], Y->lw[p], X->loc[p], X->ld[p], transpose, Y->loc[p], Y->ld[p], Y->es);
        return 0;
    }
    /* Axpy.hpp:177-185: XCopy has Y's distribution and alignments, Copy(X, XCopy), local axpy.  TransposeAxpy.hpp:201-208
     * copies into the TRANSPOSED distribution and adds the local transpose: element for element the same sums. */
    orc_dm* XC = orc_dm_create(Y->g, Y->m, Y->n, Y->colDist, Y->rowDist, Y->colAlign, Y->rowAlign, Y->es);
    orc_redist_closed(X, XC, transpose);
    for (int p = 0; p < Y->P; ++p) local_axpy(alpha, Y->lh[p], Y->lw[p], XC->loc[p], XC->ld[p], 0, Y->loc[p], Y->ld[p], Y->es);
    orc_dm_free(XC);
    return 0;
}

/* C = A o B: same distribution for all three, A and B aligned (LogicError otherwise -> non-zero); C adopts A's alignment. */
int orc_hadamard(const orc_dm* A, const orc_dm* B, orc_dm* C) {
    if (A->m != B->m || A->n != B->n) return 1;
    if (A->colDist != B->colDist || A->rowDist != B->rowDist || B->colDist != C->colDist || B->rowDist != C->rowDist) return 2;
    if (A->colAlign != B->colAlign || A->rowAlign != B->rowAlign) return 3;
    if (C->colAlign != A->colAlign || C->rowAlign != A->rowAlign || C->m != A->m || C->n != A->n) return 4;  /* caller aligns / resizes */
    for (int p = 0; p < C->P; ++p)
        for (int64_t j = 0; j < C->lw[p]; ++j)
            for (int64_t i = 0; i < C->lh[p]; ++i) {
                if (C->es == 8) ((double*)C->loc[p])[i + j * C->ld[p]] = ((const double*)A->loc[p])[i + j * A->ld[p]] * ((const double*)B->loc[p])[i + j * B->ld[p]];
                else ((float*)C->loc[p])[i + j * C->ld[p]] = ((const float*)A->loc[p])[i + j * A->ld[p]] * ((const float*)B->loc[p])[i + j * B->ld[p]];
            }
    return 0;
}

/* A *= alpha on every local block (Scale.hpp; alpha == 0 stores zeros, as El::Scale -> Zero does) */
void orc_scale(double alpha, orc_dm* A) {
    for (int p = 0; p < A->P; ++p)
        for (int64_t j = 0; j < A->lw[p]; ++j)
            for (int64_t i = 0; i < A->lh[p]; ++i) {
                if (A->es == 8) { double* q = (double*)A->loc[p] + i + j * A->ld[p]; *q = alpha == 0 ? 0.0 : alpha * *q; }
                else { float* q = (float*)A->loc[p] + i + j * A->ld[p]; *q = alpha == 0 ? 0.0f : (float)alpha * *q; }
            }
}

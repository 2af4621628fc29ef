/* oracle_gemm.c -- El::Gemm on the simulated grid, following the reference's SUMMA loops.
 * TEST INFRASTRUCTURE ONLY (see oracle.h).
 *
 * Follows src/blas_like/level3/Gemm.cpp:278-307 (Scale(beta,C), orientation switch) and
 * Gemm/{NN,NT,TN,TT}.hpp: same temporaries, same AlignWith targets, same local Gemm forms,
 * same panel order.  Redistributions of temporaries use the closed-form definition
 * (orc_redist_closed); tests/test_oracle.py shows the restated primitive chains produce the same
 * buffers.  Reduce steps sum partials in ascending VC rank.  Local products call the vendor BLAS
 * when one was registered with orc_set_blas (reference: blas::Gemm -> dgemm_/sgemm_,
 * src/core/imports/blas/Gemm.hpp:407,450), else a plain triple loop. */
#include "oracle.h"
#include <dlfcn.h>
#include <stdlib.h>
#include <string.h>

typedef void (*dgemm_fn)(const char*, const char*, const int*, const int*, const int*, const double*, const double*,
                         const int*, const double*, const int*, const double*, double*, const int*);
typedef void (*sgemm_fn)(const char*, const char*, const int*, const int*, const int*, const float*, const float*,
                         const int*, const float*, const int*, const float*, float*, const int*);
static dgemm_fn g_dgemm = 0;
static sgemm_fn g_sgemm = 0;
typedef void (*set_threads_fn)(int);
typedef int (*get_threads_fn)(void);
static set_threads_fn g_set_threads = 0;
static get_threads_fn g_get_threads = 0;

int orc_set_blas(const char* path) {
    if (!path) { g_dgemm = 0; g_sgemm = 0; return 0; }
    void* h = dlopen(path, RTLD_NOW | RTLD_LOCAL);
    if (!h) return -1;
    g_dgemm = (dgemm_fn)dlsym(h, "scipy_dgemm_");
    if (!g_dgemm) g_dgemm = (dgemm_fn)dlsym(h, "dgemm_");
    g_sgemm = (sgemm_fn)dlsym(h, "scipy_sgemm_");
    if (!g_sgemm) g_sgemm = (sgemm_fn)dlsym(h, "sgemm_");
    g_set_threads = (set_threads_fn)dlsym(h, "scipy_openblas_set_num_threads");
    if (!g_set_threads) g_set_threads = (set_threads_fn)dlsym(h, "openblas_set_num_threads");
    g_get_threads = (get_threads_fn)dlsym(h, "scipy_openblas_get_num_threads");
    if (!g_get_threads) g_get_threads = (get_threads_fn)dlsym(h, "openblas_get_num_threads");
    return (g_dgemm && g_sgemm) ? 0 : -2;
}

/* Threads the registered BLAS uses (launchers such as torchrun export OMP_NUM_THREADS=1, which OpenBLAS honours at
 * load time; the timed CPU arm sets the count it reports).  n <= 0 only queries.  Returns the count, 0 if unknown. */
int orc_blas_threads(int n) {
    if (n > 0 && g_set_threads) g_set_threads(n);
    return g_get_threads ? g_get_threads() : 0;
}

#define NAIVE_GEMM(T, NAME)                                                                                  \
    static void NAME(int tA, int tB, int64_t m, int64_t n, int64_t k, T alpha, const T* A, int64_t lda,     \
                     const T* B, int64_t ldb, T beta, T* C, int64_t ldc) {                                  \
        for (int64_t j = 0; j < n; ++j)                                                                     \
            for (int64_t i = 0; i < m; ++i) {                                                               \
                T s = 0;                                                                                    \
                for (int64_t p = 0; p < k; ++p) {                                                           \
                    const T a = tA ? A[p + i * lda] : A[i + p * lda];                                       \
                    const T b = tB ? B[j + p * ldb] : B[p + j * ldb];                                       \
                    s += a * b;                                                                             \
                }                                                                                           \
                T* c = C + i + j * ldc;                                                                     \
                *c = (beta == (T)0) ? alpha * s : alpha * s + beta * *c;                                    \
            }                                                                                               \
    }
NAIVE_GEMM(double, naive_dgemm)
NAIVE_GEMM(float, naive_sgemm)

void orc_local_gemm(int tA, int tB, int64_t m, int64_t n, int64_t k, double alpha, const void* A, int64_t lda,
                    const void* B, int64_t ldb, double beta, void* C, int64_t ldc, int es) {
    if (m == 0 || n == 0) return;
    if (es == 8) {
        if (g_dgemm && k > 0) {
            const int im = (int)m, in = (int)n, ik = (int)k, ia = (int)lda, ib = (int)ldb, ic = (int)ldc;
            g_dgemm(tA ? "T" : "N", tB ? "T" : "N", &im, &in, &ik, &alpha, (const double*)A, &ia, (const double*)B, &ib,
                    &beta, (double*)C, &ic);
        } else naive_dgemm(tA, tB, m, n, k, alpha, (const double*)A, lda, (const double*)B, ldb, beta, (double*)C, ldc);
    } else {
        const float fa = (float)alpha, fb = (float)beta;
        if (g_sgemm && k > 0) {
            const int im = (int)m, in = (int)n, ik = (int)k, ia = (int)lda, ib = (int)ldb, ic = (int)ldc;
            g_sgemm(tA ? "T" : "N", tB ? "T" : "N", &im, &in, &ik, &fa, (const float*)A, &ia, (const float*)B, &ib, &fb,
                    (float*)C, &ic);
        } else naive_sgemm(tA, tB, m, n, k, fa, (const float*)A, lda, (const float*)B, ldb, fb, (float*)C, ldc);
    }
}

void orc_dense_gemm_ref(int tA, int tB, int64_t m, int64_t n, int64_t k, double alpha, const void* A, int64_t lda,
                        const void* B, int64_t ldb, double beta, void* C, int64_t ldc, int es) {
    for (int64_t j = 0; j < n; ++j)
        for (int64_t i = 0; i < m; ++i) {
            long double s = 0;
            for (int64_t p = 0; p < k; ++p) {
                const int64_t ia = tA ? p + i * lda : i + p * lda, ib = tB ? j + p * ldb : p + j * ldb;
                const long double a = es == 8 ? ((const double*)A)[ia] : ((const float*)A)[ia];
                const long double b = es == 8 ? ((const double*)B)[ib] : ((const float*)B)[ib];
                s += a * b;
            }
            if (es == 8) { double* c = (double*)C + i + j * ldc; *c = (double)(alpha * s + (long double)beta * *c); }
            else { float* c = (float*)C + i + j * ldc; *c = (float)(alpha * s + (long double)beta * *c); }
        }
}

/* LocalGemm on every rank (Gemm.cpp:323-431): dims from the local blocks */
static void local_gemm_all(int tA, int tB, double alpha, const orc_dm* A, const orc_dm* B, double beta, orc_dm* C) {
    for (int p = 0; p < C->P; ++p) {
        const int64_t k = tA ? A->lh[p] : A->lw[p];
        orc_local_gemm(tA, tB, C->lh[p], C->lw[p], k, alpha, A->loc[p], A->ld[p], B->loc[p], B->ld[p], beta, C->loc[p],
                       C->ld[p], C->es);
    }
}
static void scale_all(double beta, orc_dm* C) {
    if (beta == 1.0) return;
    for (int p = 0; p < C->P; ++p)
        for (int64_t j = 0; j < C->lw[p]; ++j)
            for (int64_t i = 0; i < C->lh[p]; ++i) {
                if (C->es == 8) { double* c = (double*)C->loc[p] + i + j * C->ld[p]; *c = beta == 0 ? 0 : beta * *c; }
                else { float* c = (float*)C->loc[p] + i + j * C->ld[p]; *c = beta == 0 ? 0 : (float)beta * *c; }
            }
}
static int64_t imin(int64_t a, int64_t b) { return a < b ? a : b; }
#define TMP(m, n, cd, rd, ca, ra) orc_dm_create(g, m, n, cd, rd, ca, ra, es)
/* DistMatrixReadProxy<T,T,U,V> of X (Proxy.hpp:174-260): X itself when it is already [U,V] (and has the constrained
 * alignments), else a fresh [U,V] matrix assigned from X -- whose unconstrained dimensions adopt the alignment the copy::
 * primitive passes (orc_copy_adopts: e.g. an [MC,MR] operand read as [VC,*] keeps its MC alignment, ColAllToAllDemote.hpp:25).
 * Found by compiling the reference's own SUMMA_*Dot loops (oracle/ref/ref_summa.cpp): with alignment 0 instead, the k-slices
 * land on other ranks and the partial sums group differently (last-bit differences). */
orc_dm* orc_read_proxy(orc_dm* X, int U, int V, int colCon, int ca, int rowCon, int ra) {
    if (X->colDist == U && X->rowDist == V && (!colCon || X->colAlign == ca) && (!rowCon || X->rowAlign == ra)) return X;
    const orc_grid g = X->g;
    int a = 0;
    if (!colCon) ca = orc_copy_adopts(X->colDist, X->colAlign, U, orc_dist_stride(g, U), &a) ? a : 0;
    if (!rowCon) ra = orc_copy_adopts(X->rowDist, X->rowAlign, V, orc_dist_stride(g, V), &a) ? a : 0;
    orc_dm* P = orc_dm_create(g, X->m, X->n, U, V, ca, ra, X->es);
    orc_redist_closed(X, P, 0);
    return P;
}

/* ---- stationary C: NN.hpp:325-367, NT.hpp:178-221, TN.hpp:178-217, TT.hpp:195-240 ---- */
static void summa_c(int oA, int oB, double alpha, orc_dm* A, orc_dm* B, orc_dm* C, int64_t bsize) {
    const orc_grid g = C->g; const int es = C->es;
    const int64_t m = C->m, n = C->n, sumDim = oA ? A->m : A->n;
    for (int64_t k = 0; k < sumDim; k += bsize) {
        const int64_t nb = imin(bsize, sumDim - k);
        orc_dm *A1 = oA ? orc_dm_view(A, k, 0, nb, m) : orc_dm_view(A, 0, k, m, nb);
        orc_dm *B1 = oB ? orc_dm_view(B, 0, k, n, nb) : orc_dm_view(B, k, 0, nb, n);
        if (!oA && !oB) {          /* NN: A1[MC,*], B1^T[MR,*]; C += a A1 (B1^T)^T */
            orc_dm *Ap = TMP(m, nb, ORC_MC, ORC_STAR, C->colAlign, 0), *Bp = TMP(n, nb, ORC_MR, ORC_STAR, C->rowAlign, 0);
            orc_redist_closed(A1, Ap, 0); orc_redist_closed(B1, Bp, 1);
            local_gemm_all(0, 1, alpha, Ap, Bp, 1.0, C);
            orc_dm_free(Ap); orc_dm_free(Bp);
        } else if (!oA && oB) {    /* NT: A1[MC,*], (B1[VR,*])^T = B1^T[*,MR]; C += a A1 B1^T */
            orc_dm *Ap = TMP(m, nb, ORC_MC, ORC_STAR, C->colAlign, 0), *Bp = TMP(nb, n, ORC_STAR, ORC_MR, 0, C->rowAlign);
            orc_redist_closed(A1, Ap, 0); orc_redist_closed(B1, Bp, 1);
            local_gemm_all(0, 0, alpha, Ap, Bp, 1.0, C);
            orc_dm_free(Ap); orc_dm_free(Bp);
        } else if (oA && !oB) {    /* TN: A1[*,MC], B1^T[MR,*]; C += a A1^T (B1^T)^T */
            orc_dm *Ap = TMP(nb, m, ORC_STAR, ORC_MC, 0, C->colAlign), *Bp = TMP(n, nb, ORC_MR, ORC_STAR, C->rowAlign, 0);
            orc_redist_closed(A1, Ap, 0); orc_redist_closed(B1, Bp, 1);
            local_gemm_all(1, 1, alpha, Ap, Bp, 1.0, C);
            orc_dm_free(Ap); orc_dm_free(Bp);
        } else {                   /* TT: A1[*,MC], B1^T[*,MR]; C += a A1^T B1^T */
            orc_dm *Ap = TMP(nb, m, ORC_STAR, ORC_MC, 0, C->colAlign), *Bp = TMP(nb, n, ORC_STAR, ORC_MR, 0, C->rowAlign);
            orc_redist_closed(A1, Ap, 0); orc_redist_closed(B1, Bp, 1);
            local_gemm_all(1, 0, alpha, Ap, Bp, 1.0, C);
            orc_dm_free(Ap); orc_dm_free(Bp);
        }
        orc_dm_free(A1); orc_dm_free(B1);
    }
}

/* ---- stationary A: NN.hpp:107-206, NT.hpp:15-55, TN.hpp:15-57, TT.hpp:17-61 ---- */
static void summa_a(int oA, int oB, double alpha, orc_dm* A, orc_dm* B, orc_dm* C, int64_t bsize) {
    const orc_grid g = C->g; const int es = C->es;
    const int64_t m = C->m, n = C->n, sumDim = oA ? A->m : A->n;
    for (int64_t k = 0; k < n; k += bsize) {
        const int64_t nb = imin(bsize, n - k);
        orc_dm *B1 = oB ? orc_dm_view(B, k, 0, nb, sumDim) : orc_dm_view(B, 0, k, sumDim, nb);
        orc_dm *C1 = orc_dm_view(C, 0, k, m, nb);
        if (!oA && !oB) {        /* B1^T[*,MR] aligned with A; D1[MC,*] = a A (B1^T)^T; AxpyContract */
            orc_dm *Bt = TMP(nb, sumDim, ORC_STAR, ORC_MR, 0, A->rowAlign), *D1 = TMP(m, nb, ORC_MC, ORC_STAR, A->colAlign, 0);
            orc_redist_closed(B1, Bt, 1); local_gemm_all(0, 1, alpha, A, Bt, 0.0, D1);
            orc_axpy_contract_closed(1.0, D1, C1, 0); orc_dm_free(Bt); orc_dm_free(D1);
        } else if (!oA && oB) {  /* B1^T[MR,*]; D1[MC,*] = a A B1^T */
            orc_dm *Bt = TMP(sumDim, nb, ORC_MR, ORC_STAR, A->rowAlign, 0), *D1 = TMP(m, nb, ORC_MC, ORC_STAR, A->colAlign, 0);
            orc_redist_closed(B1, Bt, 1); local_gemm_all(0, 0, alpha, A, Bt, 0.0, D1);
            orc_axpy_contract_closed(1.0, D1, C1, 0); orc_dm_free(Bt); orc_dm_free(D1);
        } else if (oA && !oB) {  /* B1[MC,*]; D1[MR,*] = a A^T B1; Contract -> [MR,MC]; Axpy into C1 */
            orc_dm *Bm = TMP(sumDim, nb, ORC_MC, ORC_STAR, A->colAlign, 0), *D1 = TMP(m, nb, ORC_MR, ORC_STAR, A->rowAlign, 0);
            orc_redist_closed(B1, Bm, 0); local_gemm_all(1, 0, alpha, A, Bm, 0.0, D1);
            orc_axpy_contract_closed(1.0, D1, C1, 0); orc_dm_free(Bm); orc_dm_free(D1);
        } else {                 /* B1[*,MC]; D1[MR,*] = a A^T B1^T */
            orc_dm *Bm = TMP(nb, sumDim, ORC_STAR, ORC_MC, 0, A->colAlign), *D1 = TMP(m, nb, ORC_MR, ORC_STAR, A->rowAlign, 0);
            orc_redist_closed(B1, Bm, 0); local_gemm_all(1, 1, alpha, A, Bm, 0.0, D1);
            orc_axpy_contract_closed(1.0, D1, C1, 0); orc_dm_free(Bm); orc_dm_free(D1);
        }
        orc_dm_free(B1); orc_dm_free(C1);
    }
}

/* ---- stationary B: NN.hpp:244-287, NT.hpp:96-138, TN.hpp:98-137, TT.hpp:105-152 ---- */
static void summa_b(int oA, int oB, double alpha, orc_dm* A, orc_dm* B, orc_dm* C, int64_t bsize) {
    const orc_grid g = C->g; const int es = C->es;
    const int64_t m = C->m, n = C->n, sumDim = oA ? A->m : A->n;
    for (int64_t k = 0; k < m; k += bsize) {
        const int64_t nb = imin(bsize, m - k);
        orc_dm *A1 = oA ? orc_dm_view(A, 0, k, sumDim, nb) : orc_dm_view(A, k, 0, nb, sumDim);
        orc_dm *C1 = orc_dm_view(C, k, 0, nb, n);
        if (!oA && !oB) {        /* A1[*,MC]; D1^T[MR,*] = a B^T A1^T; TransposeAxpyContract(1, D1^T, C1) */
            orc_dm *Ap = TMP(nb, sumDim, ORC_STAR, ORC_MC, 0, B->colAlign), *Dt = TMP(n, nb, ORC_MR, ORC_STAR, B->rowAlign, 0);
            orc_redist_closed(A1, Ap, 0); local_gemm_all(1, 1, alpha, B, Ap, 0.0, Dt);
            orc_axpy_contract_closed(1.0, Dt, C1, 1); orc_dm_free(Ap); orc_dm_free(Dt);
        } else if (!oA && oB) {  /* A1^T[MR,*]; D1[*,MC] = a (A1^T)^T B^T; Contract -> [MR,MC]; Axpy */
            orc_dm *At = TMP(sumDim, nb, ORC_MR, ORC_STAR, B->rowAlign, 0), *D1 = TMP(nb, n, ORC_STAR, ORC_MC, 0, B->colAlign);
            orc_redist_closed(A1, At, 1); local_gemm_all(1, 1, alpha, At, B, 0.0, D1);
            orc_axpy_contract_closed(1.0, D1, C1, 0); orc_dm_free(At); orc_dm_free(D1);
        } else if (oA && !oB) {  /* A1[MC,*]; D1^T[MR,*] = B^T A1 (alpha = 1); TransposeAxpyContract(alpha, ...) */
            orc_dm *Ap = TMP(sumDim, nb, ORC_MC, ORC_STAR, B->colAlign, 0), *Dt = TMP(n, nb, ORC_MR, ORC_STAR, B->rowAlign, 0);
            orc_redist_closed(A1, Ap, 0); local_gemm_all(1, 0, 1.0, B, Ap, 0.0, Dt);
            orc_axpy_contract_closed(alpha, Dt, C1, 1); orc_dm_free(Ap); orc_dm_free(Dt);
        } else {                 /* A1^T[*,MR]; D1[*,MC] = a A1^T B^T; Contract; Axpy */
            orc_dm *At = TMP(nb, sumDim, ORC_STAR, ORC_MR, 0, B->rowAlign), *D1 = TMP(nb, n, ORC_STAR, ORC_MC, 0, B->colAlign);
            orc_redist_closed(A1, At, 1); local_gemm_all(0, 1, alpha, At, B, 0.0, D1);
            orc_axpy_contract_closed(1.0, D1, C1, 0); orc_dm_free(At); orc_dm_free(D1);
        }
        orc_dm_free(A1); orc_dm_free(C1);
    }
}

/* ---- Dot: NN.hpp:792-841, NT.hpp:265-310, TN.hpp:262-307, TT.hpp:287-333 ---- */
static void summa_dot(int oA, int oB, double alpha, orc_dm* A0, orc_dm* B0, orc_dm* C, int64_t blockSize) {
    const orc_grid g = C->g; const int es = C->es;
    const int64_t m = C->m, n = C->n;
    /* read proxies: A -> [*,VC] (N) / [VC,*] (T); B -> [VC,*] (N) / [*,VC] (T) aligned to A */
    orc_dm* A = oA ? orc_read_proxy(A0, ORC_VC, ORC_STAR, 0, 0, 0, 0) : orc_read_proxy(A0, ORC_STAR, ORC_VC, 0, 0, 0, 0);
    const int kal = oA ? A->colAlign : A->rowAlign;     /* BCtrl.{row,col}Constrain = true, align = A's (TT.hpp:304-307 ...) */
    orc_dm* B = oB ? orc_read_proxy(B0, ORC_STAR, ORC_VC, 0, 0, 1, kal) : orc_read_proxy(B0, ORC_VC, ORC_STAR, 1, kal, 0, 0);
    for (int64_t kO = 0; kO < m; kO += blockSize) {
        const int64_t nbO = imin(blockSize, m - kO);
        orc_dm* A1 = oA ? orc_dm_view(A, 0, kO, A->m, nbO) : orc_dm_view(A, kO, 0, nbO, A->n);
        for (int64_t kI = 0; kI < n; kI += blockSize) {
            const int64_t nbI = imin(blockSize, n - kI);
            orc_dm* B1 = oB ? orc_dm_view(B, kI, 0, nbI, B->n) : orc_dm_view(B, 0, kI, B->m, nbI);
            orc_dm* C11 = orc_dm_view(C, kO, kI, nbO, nbI);
            orc_dm* Z = TMP(nbO, nbI, ORC_STAR, ORC_STAR, 0, 0);
            local_gemm_all(oA, oB, alpha, A1, B1, 0.0, Z);
            orc_axpy_contract_closed(1.0, Z, C11, 0);
            orc_dm_free(Z); orc_dm_free(B1); orc_dm_free(C11);
        }
        orc_dm_free(A1);
    }
    if (A != A0) orc_dm_free(A);
    if (B != B0) orc_dm_free(B);
}

int orc_gemm(int oA, int oB, int alg, double alpha, orc_dm* A0, orc_dm* B0, double beta, orc_dm* C0, int64_t bsize) {
    oA = oA ? 1 : 0; oB = oB ? 1 : 0;
    const orc_grid g = C0->g; const int es = C0->es;
    const int64_t m = C0->m, n = C0->n, k = oA ? A0->m : A0->n;
    if ((oA ? A0->n : A0->m) != m || (oB ? B0->m : B0->n) != n || (oB ? B0->n : B0->m) != k) return -1;
    scale_all(beta, C0);                                     /* Gemm.cpp:287 */
    if (alg == 0) {                                          /* NN.hpp:910-925 */
        if (10.0 * m <= k && 10.0 * n <= k) alg = 4;
        else if (m <= n && 2.0 * m <= k) alg = 2;
        else if (n <= m && 2.0 * n <= k) alg = 1;
        else alg = 3;
    }
    /* ReadWrite proxy of C to [MC,MR] (Proxy.hpp:694-795) */
    orc_dm* C = C0;
    (void)es; (void)g;
    C = orc_read_proxy(C0, ORC_MC, ORC_MR, 0, 0, 0, 0);
    if (alg == 4) summa_dot(oA, oB, alpha, A0, B0, C, 2000);
    else {
        orc_dm *A = orc_read_proxy(A0, ORC_MC, ORC_MR, 0, 0, 0, 0), *B = orc_read_proxy(B0, ORC_MC, ORC_MR, 0, 0, 0, 0);
        if (alg == 1) summa_a(oA, oB, alpha, A, B, C, bsize);
        else if (alg == 2) summa_b(oA, oB, alpha, A, B, C, bsize);
        else if (alg == 3) summa_c(oA, oB, alpha, A, B, C, bsize);
        else return -2;
        if (A != A0) orc_dm_free(A);
        if (B != B0) orc_dm_free(B);
    }
    if (C != C0) { orc_redist_closed(C, C0, 0); orc_dm_free(C); }
    return 0;
}

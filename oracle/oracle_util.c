/* oracle_util.c -- restatement of the reference's pack/unpack loops (copy::util) and of the
 * redistribution primitives built from them, with collectives simulated in-process.
 * TEST INFRASTRUCTURE ONLY (see oracle.h).  Element moves are raw memcpy: bit-exact. */
#include "oracle.h"
#include <stdlib.h>
#include <string.h>

#define EL(p, idx, es) ((char*)(p) + (int64_t)(idx) * (es))
#define CEL(p, idx, es) ((const char*)(p) + (int64_t)(idx) * (es))
static int64_t pad1(int64_t n) { return n > 1 ? n : 1; } /* mpi::Pad, include/El/core/imports/mpi.hpp:183-185 */

/* util.hpp:122-146: B[i*csB + j*rsB] = A[i*csA + j*rsA] */
void orc_interleave(int64_t h, int64_t w, const void* A, int64_t csA, int64_t rsA, void* B, int64_t csB, int64_t rsB, int es) {
    for (int64_t j = 0; j < w; ++j)
        for (int64_t i = 0; i < h; ++i) memcpy(EL(B, i * csB + j * rsB, es), CEL(A, i * csA + j * rsA, es), es);
}
/* util.hpp:148-165 */
void orc_row_strided_pack(int64_t h, int64_t w, int64_t rowAlign, int64_t rowStride, const void* A, int64_t ldA, void* Bp, int64_t portion, int es) {
    for (int64_t k = 0; k < rowStride; ++k) {
        const int64_t rowShift = orc_shift(k, rowAlign, rowStride), lw = orc_length(w, rowShift, rowStride);
        orc_interleave(h, lw, CEL(A, rowShift * ldA, es), 1, rowStride * ldA, EL(Bp, k * portion, es), 1, h, es);
    }
}
/* util.hpp:167-184 */
void orc_row_strided_unpack(int64_t h, int64_t w, int64_t rowAlign, int64_t rowStride, const void* Ap, int64_t portion, void* B, int64_t ldB, int es) {
    for (int64_t k = 0; k < rowStride; ++k) {
        const int64_t rowShift = orc_shift(k, rowAlign, rowStride), lw = orc_length(w, rowShift, rowStride);
        orc_interleave(h, lw, CEL(Ap, k * portion, es), 1, h, EL(B, rowShift * ldB, es), 1, rowStride * ldB, es);
    }
}
/* util.hpp:186-207 */
void orc_partial_row_strided_pack(int64_t h, int64_t w, int64_t rowAlign, int64_t rowStride, int64_t uni, int64_t part, int64_t rankPart, int64_t shiftA, const void* A, int64_t ldA, void* Bp, int64_t portion, int es) {
    for (int64_t k = 0; k < uni; ++k) {
        const int64_t rowShift = orc_shift(rankPart + k * part, rowAlign, rowStride);
        const int64_t rowOffset = (rowShift - shiftA) / part, lw = orc_length(w, rowShift, rowStride);
        orc_interleave(h, lw, CEL(A, rowOffset * ldA, es), 1, uni * ldA, EL(Bp, k * portion, es), 1, h, es);
    }
}
/* util.hpp:209-230 */
void orc_partial_row_strided_unpack(int64_t h, int64_t w, int64_t rowAlign, int64_t rowStride, int64_t uni, int64_t part, int64_t rankPart, int64_t shiftB, const void* Ap, int64_t portion, void* B, int64_t ldB, int es) {
    for (int64_t k = 0; k < uni; ++k) {
        const int64_t rowShift = orc_shift(rankPart + k * part, rowAlign, rowStride);
        const int64_t rowOffset = (rowShift - shiftB) / part, lw = orc_length(w, rowShift, rowStride);
        orc_interleave(h, lw, CEL(Ap, k * portion, es), 1, h, EL(B, rowOffset * ldB, es), 1, uni * ldB, es);
    }
}
/* util.hpp:363-381 */
void orc_col_strided_pack(int64_t h, int64_t w, int64_t colAlign, int64_t colStride, const void* A, int64_t ldA, void* Bp, int64_t portion, int es) {
    for (int64_t k = 0; k < colStride; ++k) {
        const int64_t colShift = orc_shift(k, colAlign, colStride), lh = orc_length(h, colShift, colStride);
        orc_interleave(lh, w, CEL(A, colShift, es), colStride, ldA, EL(Bp, k * portion, es), 1, lh, es);
    }
}
/* util.hpp:403-421 */
void orc_col_strided_unpack(int64_t h, int64_t w, int64_t colAlign, int64_t colStride, const void* Ap, int64_t portion, void* B, int64_t ldB, int es) {
    for (int64_t k = 0; k < colStride; ++k) {
        const int64_t colShift = orc_shift(k, colAlign, colStride), lh = orc_length(h, colShift, colStride);
        orc_interleave(lh, w, CEL(Ap, k * portion, es), 1, lh, EL(B, colShift, es), colStride, ldB, es);
    }
}
/* util.hpp:464-486 */
void orc_partial_col_strided_pack(int64_t h, int64_t w, int64_t colAlign, int64_t colStride, int64_t uni, int64_t part, int64_t rankPart, int64_t shiftA, const void* A, int64_t ldA, void* Bp, int64_t portion, int es) {
    for (int64_t k = 0; k < uni; ++k) {
        const int64_t colShift = orc_shift(rankPart + k * part, colAlign, colStride);
        const int64_t colOffset = (colShift - shiftA) / part, lh = orc_length(h, colShift, colStride);
        orc_interleave(lh, w, CEL(A, colOffset, es), uni, ldA, EL(Bp, k * portion, es), 1, lh, es);
    }
}
/* util.hpp:510-533 */
void orc_partial_col_strided_unpack(int64_t h, int64_t w, int64_t colAlign, int64_t colStride, int64_t uni, int64_t part, int64_t rankPart, int64_t shiftB, const void* Ap, int64_t portion, void* B, int64_t ldB, int es) {
    for (int64_t k = 0; k < uni; ++k) {
        const int64_t colShift = orc_shift(rankPart + k * part, colAlign, colStride);
        const int64_t colOffset = (colShift - shiftB) / part, lh = orc_length(h, colShift, colStride);
        orc_interleave(lh, w, CEL(Ap, k * portion, es), 1, lh, EL(B, colOffset, es), uni, ldB, es);
    }
}
/* util.hpp:671-695 */
void orc_strided_pack(int64_t h, int64_t w, int64_t colAlign, int64_t colStride, int64_t rowAlign, int64_t rowStride, const void* A, int64_t ldA, void* Bp, int64_t portion, int es) {
    for (int64_t l = 0; l < rowStride; ++l) {
        const int64_t rowShift = orc_shift(l, rowAlign, rowStride), lw = orc_length(w, rowShift, rowStride);
        for (int64_t k = 0; k < colStride; ++k) {
            const int64_t colShift = orc_shift(k, colAlign, colStride), lh = orc_length(h, colShift, colStride);
            orc_interleave(lh, lw, CEL(A, colShift + rowShift * ldA, es), colStride, rowStride * ldA,
                           EL(Bp, (k + l * colStride) * portion, es), 1, lh, es);
        }
    }
}
/* util.hpp:697-722 */
void orc_strided_unpack(int64_t h, int64_t w, int64_t colAlign, int64_t colStride, int64_t rowAlign, int64_t rowStride, const void* Ap, int64_t portion, void* B, int64_t ldB, int es) {
    for (int64_t l = 0; l < rowStride; ++l) {
        const int64_t rowShift = orc_shift(l, rowAlign, rowStride), lw = orc_length(w, rowShift, rowStride);
        for (int64_t k = 0; k < colStride; ++k) {
            const int64_t colShift = orc_shift(k, colAlign, colStride), lh = orc_length(h, colShift, colStride);
            orc_interleave(lh, lw, CEL(Ap, (k + l * colStride) * portion, es), 1, lh,
                           EL(B, colShift + rowShift * ldB, es), colStride, rowStride * ldB, es);
        }
    }
}
/* Axpy/util.hpp:23-37: column j of B += alpha * column j of A (blas::Axpy per column = ?axpy: the product and the
 * sum are each rounded once; this file is built with -ffp-contract=off so no FMA is formed) */
void orc_interleave_update(double alpha, int64_t h, int64_t w, const void* A, int64_t csA, int64_t rsA, void* B, int64_t csB, int64_t rsB, int es) {
    for (int64_t j = 0; j < w; ++j)
        for (int64_t i = 0; i < h; ++i) {
            if (es == 8) {
                const double p = alpha * ((const double*)A)[i * csA + j * rsA];
                ((double*)B)[i * csB + j * rsB] += p;
            } else {
                const float p = (float)alpha * ((const float*)A)[i * csA + j * rsA];
                ((float*)B)[i * csB + j * rsB] += p;
            }
        }
}

/* ---------------- simulated communicators ---------------- */
typedef struct { int mc, mr; } coord;
static coord co(orc_grid g, int p) { coord c; orc_grid_coords(g, p, &c.mc, &c.mr); return c; }
/* the communicator a 1-D distribution lives on: rank within it, size, and world rank of member k of p's comm */
static int comm_member(orc_grid g, int dist, int p, int k) {
    coord c = co(g, p);
    switch (dist) {
        case ORC_MC: return orc_grid_rank_of(g, k, c.mr);           /* MC comm: the ranks of my grid column */
        case ORC_MR: return orc_grid_rank_of(g, c.mc, k);           /* MR comm: the ranks of my grid row */
        case ORC_VC: return orc_grid_rank_of(g, k % g.r, k / g.r);
        case ORC_VR: return orc_grid_rank_of(g, k / g.c, k % g.c);
        default: return p;
    }
}
static int partial_of(int d) { return d == ORC_VC ? ORC_MC : d == ORC_VR ? ORC_MR : d; }       /* Partial<U>() */
static int partial_union_of(int d) { return d == ORC_VC ? ORC_MR : d == ORC_VR ? ORC_MC : ORC_STAR; } /* PartialUnionCol */
static int collect(int d) { (void)d; return ORC_STAR; }

static char** alloc_bufs(int P, int64_t elems, int es) {
    char** b = (char**)malloc(sizeof(char*) * P);
    for (int p = 0; p < P; ++p) b[p] = (char*)calloc((size_t)pad1(elems), es);
    return b;
}
static void free_bufs(char** b, int P) { for (int p = 0; p < P; ++p) free(b[p]); free(b); }

/* Copy/RowAllGather.hpp:18-145 : (U,V) -> (U,*) */
int orc_row_allgather(const orc_dm* A, orc_dm* B) {
    if (B->colDist != A->colDist || B->rowDist != collect(A->rowDist) || A->rowDist == ORC_STAR) return -1;
    const orc_grid g = A->g;
    const int P = A->P, es = A->es;
    const int64_t height = A->m, width = A->n;
    const int rowStride = orc_dist_stride(g, A->rowDist), colStride = orc_dist_stride(g, A->colDist);
    const int64_t colDiff = B->colAlign - A->colAlign;
    const int64_t maxLH = orc_max_length(height, colStride), maxLW = orc_max_length(width, rowStride);
    const int64_t portion = pad1(maxLH * maxLW);
    char** first = alloc_bufs(P, portion, es);
    char** second = alloc_bufs(P, (int64_t)rowStride * portion, es);
    /* pack (:59-63 / :107-110) */
    for (int p = 0; p < P; ++p) orc_interleave(A->lh[p], A->lw[p], A->loc[p], 1, A->ld[p], first[p], 1, A->lh[p], es);
    if (colDiff != 0) { /* realign: SendRecv over the column comm (:84-85,113-116) */
        char** tmp = alloc_bufs(P, portion, es);
        for (int p = 0; p < P; ++p) {
            coord c = co(g, p);
            const int colRank = orc_dist_rank(g, A->colDist, c.mc, c.mr);
            const int sendColRank = (int)orc_mod(colRank + colDiff, colStride);
            const int q = comm_member(g, A->colDist, p, sendColRank);
            memcpy(tmp[q], first[p], (size_t)portion * es);
        }
        free_bufs(first, P);
        first = tmp;
    }
    /* AllGather over RowComm: portions concatenated in communicator-rank order (:66-68) */
    for (int p = 0; p < P; ++p)
        for (int k = 0; k < rowStride; ++k)
            memcpy(second[p] + (int64_t)k * portion * es, first[comm_member(g, A->rowDist, p, k)], (size_t)portion * es);
    /* unpack (:71-75 / :141-144) */
    for (int p = 0; p < P; ++p)
        orc_row_strided_unpack(B->lh[p], width, A->rowAlign, rowStride, second[p], portion, B->loc[p], B->ld[p], es);
    free_bufs(first, P); free_bufs(second, P);
    return 0;
}

/* Copy/ColAllGather.hpp:16-181 : (U,V) -> (*,V) */
int orc_col_allgather(const orc_dm* A, orc_dm* B) {
    if (B->rowDist != A->rowDist || B->colDist != ORC_STAR || A->colDist == ORC_STAR) return -1;
    const orc_grid g = A->g;
    const int P = A->P, es = A->es;
    const int64_t height = A->m, width = A->n;
    const int rowStride = orc_dist_stride(g, A->rowDist), colStride = orc_dist_stride(g, A->colDist);
    const int64_t rowDiff = B->rowAlign - A->rowAlign;
    const int64_t maxLH = orc_max_length(height, colStride), maxLW = orc_max_length(width, rowStride);
    const int64_t portion = pad1(maxLH * maxLW);
    char** first = alloc_bufs(P, portion, es);
    char** second = alloc_bufs(P, (int64_t)colStride * portion, es);
    for (int p = 0; p < P; ++p) orc_interleave(A->lh[p], A->lw[p], A->loc[p], 1, A->ld[p], first[p], 1, A->lh[p], es);
    if (rowDiff != 0) { /* SendRecv over the row comm (:96-97,146-155) */
        char** tmp = alloc_bufs(P, portion, es);
        for (int p = 0; p < P; ++p) {
            coord c = co(g, p);
            const int rowRank = orc_dist_rank(g, A->rowDist, c.mc, c.mr);
            const int q = comm_member(g, A->rowDist, p, (int)orc_mod(rowRank + rowDiff, rowStride));
            memcpy(tmp[q], first[p], (size_t)portion * es);
        }
        free_bufs(first, P);
        first = tmp;
    }
    for (int p = 0; p < P; ++p)
        for (int k = 0; k < colStride; ++k)
            memcpy(second[p] + (int64_t)k * portion * es, first[comm_member(g, A->colDist, p, k)], (size_t)portion * es);
    for (int p = 0; p < P; ++p)
        orc_col_strided_unpack(height, B->lw[p], A->colAlign, colStride, second[p], portion, B->loc[p], B->ld[p], es);
    free_bufs(first, P); free_bufs(second, P);
    return 0;
}

/* Copy/ColAllToAllDemote.hpp:15-126 (aligned branch :47-81): (Partial(U), PartialUnionRow(U,V)) -> (U,V),
 * i.e. [MC,MR] -> [VC,*] and [MR,MC] -> [VR,*]. */
int orc_col_alltoall_demote(const orc_dm* A, orc_dm* B) {
    if (!(B->colDist == ORC_VC || B->colDist == ORC_VR) || B->rowDist != ORC_STAR) return -1;
    if (A->colDist != partial_of(B->colDist) || A->rowDist != partial_union_of(B->colDist)) return -1;
    const orc_grid g = A->g;
    const int P = A->P, es = A->es;
    const int64_t height = A->m, width = A->n;
    const int colStride = orc_dist_stride(g, B->colDist);
    const int part = orc_dist_stride(g, partial_of(B->colDist)), uni = orc_dist_stride(g, partial_union_of(B->colDist));
    if (orc_mod(B->colAlign, part) - A->colAlign != 0) return -2; /* unaligned branch not restated */
    const int64_t maxLH = orc_max_length(height, colStride), maxLW = orc_max_length(width, uni);
    const int64_t portion = pad1(maxLH * maxLW);
    char** first = alloc_bufs(P, (int64_t)uni * portion, es);
    char** second = alloc_bufs(P, (int64_t)uni * portion, es);
    for (int p = 0; p < P; ++p) {
        coord c = co(g, p);
        const int colRankPart = orc_dist_rank(g, partial_of(B->colDist), c.mc, c.mr);
        orc_partial_col_strided_pack(height, A->lw[p], B->colAlign, colStride, uni, part, colRankPart, A->colShift[p],
                                     A->loc[p], A->ld[p], first[p], portion, es);
    }
    /* AllToAll over B.PartialUnionColComm(): portion k goes to comm rank k, lands in slot (my comm rank) (:69-72) */
    const int ud = partial_union_of(B->colDist);
    for (int p = 0; p < P; ++p) {
        coord c = co(g, p);
        const int me = orc_dist_rank(g, ud, c.mc, c.mr);
        for (int k = 0; k < uni; ++k) {
            const int q = comm_member(g, ud, p, k);
            memcpy(second[q] + (int64_t)me * portion * es, first[p] + (int64_t)k * portion * es, (size_t)portion * es);
        }
    }
    for (int p = 0; p < P; ++p)
        orc_row_strided_unpack(B->lh[p], width, A->rowAlign, uni, second[p], portion, B->loc[p], B->ld[p], es);
    free_bufs(first, P); free_bufs(second, P);
    return 0;
}

/* Copy/RowAllToAllDemote.hpp:15-124 (aligned branch): (PartialUnionCol(V,U), Partial(V)) -> (U,V),
 * i.e. [MC,MR] -> [*,VR] and [MR,MC] -> [*,VC]. */
int orc_row_alltoall_demote(const orc_dm* A, orc_dm* B) {
    if (!(B->rowDist == ORC_VC || B->rowDist == ORC_VR) || B->colDist != ORC_STAR) return -1;
    if (A->rowDist != partial_of(B->rowDist) || A->colDist != partial_union_of(B->rowDist)) return -1;
    const orc_grid g = A->g;
    const int P = A->P, es = A->es;
    const int64_t height = A->m, width = A->n;
    const int rowStride = orc_dist_stride(g, B->rowDist);
    const int part = orc_dist_stride(g, partial_of(B->rowDist)), uni = orc_dist_stride(g, partial_union_of(B->rowDist));
    if (orc_mod(B->rowAlign, part) - A->rowAlign != 0) return -2;
    const int64_t maxLH = orc_max_length(height, uni), maxLW = orc_max_length(width, rowStride);
    const int64_t portion = pad1(maxLH * maxLW);
    char** first = alloc_bufs(P, (int64_t)uni * portion, es);
    char** second = alloc_bufs(P, (int64_t)uni * portion, es);
    for (int p = 0; p < P; ++p) {
        coord c = co(g, p);
        const int rowRankPart = orc_dist_rank(g, partial_of(B->rowDist), c.mc, c.mr);
        orc_partial_row_strided_pack(A->lh[p], width, B->rowAlign, rowStride, uni, part, rowRankPart, A->rowShift[p],
                                     A->loc[p], A->ld[p], first[p], portion, es);
    }
    const int ud = partial_union_of(B->rowDist);
    for (int p = 0; p < P; ++p) {
        coord c = co(g, p);
        const int me = orc_dist_rank(g, ud, c.mc, c.mr);
        for (int k = 0; k < uni; ++k) {
            const int q = comm_member(g, ud, p, k);
            memcpy(second[q] + (int64_t)me * portion * es, first[p] + (int64_t)k * portion * es, (size_t)portion * es);
        }
    }
    for (int p = 0; p < P; ++p)
        orc_col_strided_unpack(height, B->lw[p], A->colAlign, uni, second[p], portion, B->loc[p], B->ld[p], es);
    free_bufs(first, P); free_bufs(second, P);
    return 0;
}

/* Copy/PartialColAllGather.hpp (aligned branch): (U,V) -> (Partial(U),V): [VC,*] -> [MC,*], [VR,*] -> [MR,*] */
int orc_partial_col_allgather(const orc_dm* A, orc_dm* B) {
    if (!(A->colDist == ORC_VC || A->colDist == ORC_VR) || B->colDist != partial_of(A->colDist) || A->rowDist != B->rowDist)
        return -1;
    if (A->rowDist != ORC_STAR) return -1; /* "This routine assumes rows are not distributed" */
    const orc_grid g = A->g;
    const int P = A->P, es = A->es;
    const int64_t height = A->m, width = A->n;
    const int colStride = orc_dist_stride(g, A->colDist);
    const int part = orc_dist_stride(g, partial_of(A->colDist)), uni = orc_dist_stride(g, partial_union_of(A->colDist));
    if (B->colAlign - orc_mod(A->colAlign, part) != 0) return -2;
    const int64_t portion = pad1(orc_max_length(height, colStride) * width);
    char** first = alloc_bufs(P, portion, es);
    char** second = alloc_bufs(P, (int64_t)uni * portion, es);
    for (int p = 0; p < P; ++p) orc_interleave(A->lh[p], width, A->loc[p], 1, A->ld[p], first[p], 1, A->lh[p], es);
    const int ud = partial_union_of(A->colDist);
    for (int p = 0; p < P; ++p)
        for (int k = 0; k < uni; ++k)
            memcpy(second[p] + (int64_t)k * portion * es, first[comm_member(g, ud, p, k)], (size_t)portion * es);
    for (int p = 0; p < P; ++p) {
        coord c = co(g, p);
        const int rankPart = orc_dist_rank(g, partial_of(A->colDist), c.mc, c.mr);
        orc_partial_col_strided_unpack(height, width, A->colAlign, colStride, uni, part, rankPart, B->colShift[p],
                                       second[p], portion, B->loc[p], B->ld[p], es);
    }
    free_bufs(first, P); free_bufs(second, P);
    return 0;
}

/* Copy/PartialRowAllGather.hpp:16-111 (aligned branch :44-71): (U,V) -> (U,Partial(V)): [*,VC] -> [*,MC], [*,VR] -> [*,MR] */
int orc_partial_row_allgather(const orc_dm* A, orc_dm* B) {
    if (!(A->rowDist == ORC_VC || A->rowDist == ORC_VR) || B->rowDist != partial_of(A->rowDist) || A->colDist != B->colDist)
        return -1;
    if (A->colDist != ORC_STAR) return -1;
    const orc_grid g = A->g;
    const int P = A->P, es = A->es;
    const int64_t height = A->m, width = A->n;
    const int rowStride = orc_dist_stride(g, A->rowDist);
    const int part = orc_dist_stride(g, partial_of(A->rowDist)), uni = orc_dist_stride(g, partial_union_of(A->rowDist));
    if (B->rowAlign - orc_mod(A->rowAlign, part) != 0) return -2;
    const int64_t portion = pad1(height * orc_max_length(width, rowStride));
    char** first = alloc_bufs(P, portion, es);
    char** second = alloc_bufs(P, (int64_t)uni * portion, es);
    for (int p = 0; p < P; ++p) orc_interleave(height, A->lw[p], A->loc[p], 1, A->ld[p], first[p], 1, height, es);
    const int ud = partial_union_of(A->rowDist);
    for (int p = 0; p < P; ++p)
        for (int k = 0; k < uni; ++k)
            memcpy(second[p] + (int64_t)k * portion * es, first[comm_member(g, ud, p, k)], (size_t)portion * es);
    for (int p = 0; p < P; ++p) {
        coord c = co(g, p);
        const int rankPart = orc_dist_rank(g, partial_of(A->rowDist), c.mc, c.mr);
        orc_partial_row_strided_unpack(height, width, A->rowAlign, rowStride, uni, part, rankPart, B->rowShift[p],
                                       second[p], portion, B->loc[p], B->ld[p], es);
    }
    free_bufs(first, P); free_bufs(second, P);
    return 0;
}

/* Copy/Exchange.hpp:130-192: Colwise / RowwiseVectorExchange ([VC,*] <-> [VR,*], [*,VC] <-> [*,VR]):
 * every rank ships its whole local block to the rank whose shift in B equals its shift in A. */
int orc_vector_exchange(const orc_dm* A, orc_dm* B) {
    const orc_grid g = A->g;
    const int P = A->P, es = A->es;
    int colwise;
    if ((A->colDist == ORC_VC || A->colDist == ORC_VR) && (B->colDist == ORC_VC || B->colDist == ORC_VR) &&
        A->colDist != B->colDist && A->rowDist == ORC_STAR && B->rowDist == ORC_STAR) colwise = 1;
    else if ((A->rowDist == ORC_VC || A->rowDist == ORC_VR) && (B->rowDist == ORC_VC || B->rowDist == ORC_VR) &&
             A->rowDist != B->rowDist && A->colDist == ORC_STAR && B->colDist == ORC_STAR) colwise = 0;
    else return -1;
    const int bd = colwise ? B->colDist : B->rowDist;
    for (int p = 0; p < P; ++p) {
        coord c = co(g, p);
        const int64_t diff = colwise ? A->colShift[p] - B->colShift[p] : A->rowShift[p] - B->rowShift[p];
        const int sendRankB = (int)orc_mod(orc_dist_rank(g, bd, c.mc, c.mr) + diff, P);  /* :163 / :184 */
        const int q = comm_member(g, bd, p, sendRankB);
        orc_interleave(A->lh[p], A->lw[p], A->loc[p], 1, A->ld[p], B->loc[q], 1, B->ld[q], es);
    }
    return 0;
}

/* local transpose on every rank: B(j,i) = A(i,j); B's distribution must be A's transposed (Transpose.hpp:202-210) */
void orc_local_transpose(const orc_dm* A, orc_dm* B) {
    const int es = A->es;
    for (int p = 0; p < A->P; ++p)
        for (int64_t j = 0; j < A->lw[p]; ++j)
            for (int64_t i = 0; i < A->lh[p]; ++i)
                memcpy(EL(B->loc[p], j + i * B->ld[p], es), CEL(A->loc[p], i + j * A->ld[p], es), es);
}

/* oracle_core.c -- indexing, grid, simulated distributed matrices, closed-form definitions.
 * TEST INFRASTRUCTURE ONLY (see oracle.h). */
#include "oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* include/El/core/indexing/impl.hpp:229-233 */
int64_t orc_mod(int64_t a, int64_t b) { int64_t r = a % b; return r >= 0 ? r : r + b; }
/* :244-245 */
int64_t orc_shift(int64_t rank, int64_t align, int64_t stride) { return orc_mod(rank - align, stride); }
/* :33-36 */
int64_t orc_length(int64_t n, int64_t shift, int64_t stride) { return n > shift ? (n - shift - 1) / stride + 1 : 0; }
/* :65-66 */
int64_t orc_max_length(int64_t n, int64_t stride) { return n > 0 ? (n - 1) / stride + 1 : 0; }

/* src/core/Grid.cpp:121-148: column-major grids number ranks down the columns (rank == VC rank),
 * row-major grids along the rows (rank == VR rank). */
/* Grid::DefaultHeight, src/core/Grid.cpp:58-64: int(sqrt(double(gridSize))), then ++ until it divides gridSize */
int orc_grid_default_height(int size) {
    int h = (int)sqrt((double)size);
    while (size % h != 0) ++h;
    return h;
}
/* Distribution algebra, include/El/core/types.hpp:292-337, written as lookup tables over the Dist enumerators
 * (MC MD MR VC VR STAR CIRC = 0..6): Collect, Partial, and the partial-union of a [U,V] matrix's row dimension. */
static const int kCollect[7] = {ORC_STAR, ORC_STAR, ORC_STAR, ORC_STAR, ORC_STAR, ORC_STAR, ORC_CIRC};
static const int kPartial[7] = {ORC_MC, ORC_MD, ORC_MR, ORC_MC, ORC_MR, ORC_STAR, ORC_CIRC};
int orc_dist_collect(int u) { return kCollect[u]; }
int orc_dist_partial(int u) { return kPartial[u]; }
int orc_dist_partial_union_row(int u, int v) { return u == ORC_VC ? ORC_MR : (u == ORC_VR ? ORC_MC : v); }
int orc_dist_partial_union_col(int u, int v) { return orc_dist_partial_union_row(v, u); }
/* ElementalMatrix::AlignColsWith / AlignRowsWith, src/core/DistMatrix/ElementMatrix.cpp:243-300, for one dimension of a
 * [U,V] matrix: which = 0 aligns the column dimension (distribution U), 1 the row dimension (V).  `stride` is that
 * dimension's stride.  Returns 1 and *align when the reference calls AlignCols/AlignRows, 0 when it does nothing, -1 where
 * it raises "Nonsensical alignment". */
int orc_align_dim_with(int U, int V, int which, int stride, int dataColDist, int dataColAlign, int dataRowDist, int dataRowAlign,
                       int* align) {
    const int mine = which == 0 ? U : V;
    const int part = kPartial[mine];
    const int uni = which == 0 ? orc_dist_partial_union_col(U, V) : orc_dist_partial_union_row(U, V);
    const int coll = kCollect[mine];
    if (dataColDist == mine || dataColDist == part) *align = dataColAlign;
    else if (dataRowDist == mine || dataRowDist == part) *align = dataRowAlign;
    else if (dataColDist == uni) *align = dataColAlign % stride;
    else if (dataRowDist == uni) *align = dataRowAlign % stride;
    else return (mine != coll && dataColDist != coll && dataRowDist != coll) ? -1 : 0;
    return 1;
}
/* The alignment argument of the copy:: primitive serving one dimension of B = A when that dimension goes src -> dst
 * (Translate.hpp:38-42, RowAllGather.hpp:23, ColAllGather.hpp:36, {Row,Col}Filter.hpp:27-29: the source's alignment;
 * Partial{Col,Row}AllGather.hpp:43-44/22-23, {Col,Row}AllToAllPromote.hpp:26-27: Mod(srcAlign, dstStride);
 * {Col,Row}AllToAllDemote.hpp:25, Partial{Col,Row}Filter.hpp:29-31: the source's alignment).  An unconstrained destination
 * takes it; returns 0 where no primitive passes one (unrelated distributions, STAR, CIRC). */
int orc_copy_adopts(int src, int srcAlign, int dst, int dstStride, int* align) {
    if (dst == ORC_STAR || dst == ORC_CIRC || src == ORC_STAR || src == ORC_CIRC) return 0;
    if (src == dst || kPartial[dst] == src) { *align = srcAlign; return 1; }
    if (kPartial[src] == dst) { *align = srcAlign % dstStride; return 1; }
    return 0;
}
void orc_grid_coords(orc_grid g, int rank, int* mc, int* mr) {
    if (g.col_major) { *mc = rank % g.r; *mr = rank / g.r; }
    else             { *mc = rank / g.c; *mr = rank % g.c; }
}
int orc_grid_rank_of(orc_grid g, int mc, int mr) { return g.col_major ? mc + g.r * mr : mr + g.c * mc; }
/* ColStride()/RowStride(): bottom of src/core/DistMatrix/ElementMatrix/<dist>.cpp, e.g. VC_STAR.cpp:201-205 */
int orc_dist_stride(orc_grid g, int d) {
    switch (d) { case ORC_MC: return g.r; case ORC_MR: return g.c; case ORC_VC: case ORC_VR: return g.r * g.c; default: return 1; }
}
/* ColRank()/RowRank(): e.g. VC_STAR.cpp:232-236; vcRank = mc + r*mr, vrRank = mr + c*mc (Grid.cpp:147-148) */
int orc_dist_rank(orc_grid g, int d, int mc, int mr) {
    switch (d) { case ORC_MC: return mc; case ORC_MR: return mr; case ORC_VC: return mc + g.r * mr; case ORC_VR: return mr + g.c * mc; default: return 0; }
}

static void dm_shapes(orc_dm* a) {
    const int cs = orc_dist_stride(a->g, a->colDist), rs = orc_dist_stride(a->g, a->rowDist);
    for (int p = 0; p < a->P; ++p) {
        int mc, mr;
        orc_grid_coords(a->g, p, &mc, &mr);
        /* AbstractDistMatrix.cpp:399-430: shift = Shift(rank, align, stride) */
        a->colShift[p] = orc_shift(orc_dist_rank(a->g, a->colDist, mc, mr), a->colAlign, cs);
        a->rowShift[p] = orc_shift(orc_dist_rank(a->g, a->rowDist, mc, mr), a->rowAlign, rs);
        a->lh[p] = orc_length(a->m, a->colShift[p], cs);
        a->lw[p] = orc_length(a->n, a->rowShift[p], rs);
    }
}

static orc_dm* dm_alloc_meta(orc_grid g, int64_t m, int64_t n, int cd, int rd, int ca, int ra, int es) {
    orc_dm* a = (orc_dm*)calloc(1, sizeof(orc_dm));
    a->g = g; a->m = m; a->n = n; a->colDist = cd; a->rowDist = rd; a->colAlign = ca; a->rowAlign = ra; a->es = es;
    a->P = g.r * g.c;
    a->loc = (void**)calloc(a->P, sizeof(void*));
    a->ld = (int64_t*)calloc(a->P, sizeof(int64_t));
    a->lh = (int64_t*)calloc(a->P, sizeof(int64_t));
    a->lw = (int64_t*)calloc(a->P, sizeof(int64_t));
    a->colShift = (int64_t*)calloc(a->P, sizeof(int64_t));
    a->rowShift = (int64_t*)calloc(a->P, sizeof(int64_t));
    dm_shapes(a);
    return a;
}

orc_dm* orc_dm_create(orc_grid g, int64_t m, int64_t n, int cd, int rd, int ca, int ra, int es) {
    orc_dm* a = dm_alloc_meta(g, m, n, cd, rd, ca, ra, es);
    a->owner = 1;
    for (int p = 0; p < a->P; ++p) {
        a->ld[p] = a->lh[p] > 0 ? a->lh[p] : 1;
        const int64_t cnt = a->ld[p] * (a->lw[p] > 0 ? a->lw[p] : 1);
        a->loc[p] = calloc((size_t)cnt, (size_t)es);
    }
    return a;
}
void orc_dm_free(orc_dm* a) {
    if (!a) return;
    if (a->owner) for (int p = 0; p < a->P; ++p) free(a->loc[p]);
    free(a->loc); free(a->ld); free(a->lh); free(a->lw); free(a->colShift); free(a->rowShift); free(a);
}
void* orc_dm_local(orc_dm* a, int rank) { return a->loc[rank]; }
int64_t orc_dm_lh(const orc_dm* a, int rank) { return a->lh[rank]; }
int64_t orc_dm_lw(const orc_dm* a, int rank) { return a->lw[rank]; }
int64_t orc_dm_ld(const orc_dm* a, int rank) { return a->ld[rank]; }

/* El::View: alignment (align + i) mod stride; local offset = number of local indices below i */
orc_dm* orc_dm_view(orc_dm* a, int64_t i0, int64_t j0, int64_t h, int64_t w) {
    const int cs = orc_dist_stride(a->g, a->colDist), rs = orc_dist_stride(a->g, a->rowDist);
    orc_dm* v = dm_alloc_meta(a->g, h, w, a->colDist, a->rowDist, (int)orc_mod(a->colAlign + i0, cs),
                              (int)orc_mod(a->rowAlign + j0, rs), a->es);
    v->owner = 0;
    for (int p = 0; p < a->P; ++p) {
        const int64_t iLoc = orc_length(i0, a->colShift[p], cs), jLoc = orc_length(j0, a->rowShift[p], rs);
        v->ld[p] = a->ld[p];
        v->loc[p] = (char*)a->loc[p] + (iLoc + jLoc * a->ld[p]) * a->es;
    }
    return v;
}

/* splitmix64 of (seed, index) -> uniform [-1, 1) */
double orc_uniform(uint64_t seed, uint64_t idx) {
    uint64_t z = seed * 0x9E3779B97F4A7C15ull + idx * 0xBF58476D1CE4E5B9ull + 0x94D049BB133111EBull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z = z ^ (z >> 31);
    return (double)(z >> 11) * (2.0 / 9007199254740992.0) - 1.0;
}

static void put(const orc_dm* a, int p, int64_t iLoc, int64_t jLoc, double v) {
    if (a->es == 8) ((double*)a->loc[p])[iLoc + jLoc * a->ld[p]] = v;
    else ((float*)a->loc[p])[iLoc + jLoc * a->ld[p]] = (float)v;
}

typedef double (*gen_fn)(const orc_dm*, uint64_t, int64_t, int64_t);
static double gen_uniform(const orc_dm* a, uint64_t seed, int64_t i, int64_t j) { return orc_uniform(seed, (uint64_t)(i + j * a->m)); }
static double gen_index(const orc_dm* a, uint64_t seed, int64_t i, int64_t j) { (void)a; (void)seed; return (double)((i * 131 + j) % 16777216); }
static double gen_const(const orc_dm* a, uint64_t seed, int64_t i, int64_t j) { (void)a; (void)i; (void)j; double v; memcpy(&v, &seed, 8); return v; }

static void fill_with(orc_dm* a, gen_fn f, uint64_t seed) {
    const int cs = orc_dist_stride(a->g, a->colDist), rs = orc_dist_stride(a->g, a->rowDist);
    for (int p = 0; p < a->P; ++p)
        for (int64_t jl = 0; jl < a->lw[p]; ++jl)
            for (int64_t il = 0; il < a->lh[p]; ++il)
                put(a, p, il, jl, f(a, seed, a->colShift[p] + il * cs, a->rowShift[p] + jl * rs));
}
void orc_dm_fill(orc_dm* a, uint64_t seed) { fill_with(a, gen_uniform, seed); }
void orc_dm_fill_index(orc_dm* a) { fill_with(a, gen_index, 0); }
void orc_dm_fill_const(orc_dm* a, double v) { uint64_t s; memcpy(&s, &v, 8); fill_with(a, gen_const, s); }

/* gather to a global column-major m x n array (ld = m); counts bitwise disagreements between replicas */
int64_t orc_dm_to_global(const orc_dm* a, void* out) {
    const int cs = orc_dist_stride(a->g, a->colDist), rs = orc_dist_stride(a->g, a->rowDist);
    const int es = a->es;
    int64_t mism = 0;
    char* seen = (char*)calloc((size_t)(a->m * a->n > 0 ? a->m * a->n : 1), 1);
    for (int p = 0; p < a->P; ++p)
        for (int64_t jl = 0; jl < a->lw[p]; ++jl)
            for (int64_t il = 0; il < a->lh[p]; ++il) {
                const int64_t i = a->colShift[p] + il * cs, j = a->rowShift[p] + jl * rs;
                const char* s = (const char*)a->loc[p] + (il + jl * a->ld[p]) * es;
                char* d = (char*)out + (i + j * a->m) * es;
                if (seen[i + j * a->m]) { if (memcmp(d, s, es) != 0) ++mism; }
                else { memcpy(d, s, es); seen[i + j * a->m] = 1; }
            }
    free(seen);
    return mism;
}
void orc_dm_from_global(orc_dm* a, const void* in) {
    const int cs = orc_dist_stride(a->g, a->colDist), rs = orc_dist_stride(a->g, a->rowDist);
    const int es = a->es;
    for (int p = 0; p < a->P; ++p)
        for (int64_t jl = 0; jl < a->lw[p]; ++jl)
            for (int64_t il = 0; il < a->lh[p]; ++il) {
                const int64_t i = a->colShift[p] + il * cs, j = a->rowShift[p] + jl * rs;
                memcpy((char*)a->loc[p] + (il + jl * a->ld[p]) * es, (const char*)in + (i + j * a->m) * es, es);
            }
}

/* B = A (or A^T) by definition: every owner of global (i,j) in dst holds the value src holds there. */
void orc_redist_closed(const orc_dm* src, orc_dm* dst, int transpose) {
    const int es = src->es;
    const int64_t cnt = src->m * src->n > 0 ? src->m * src->n : 1;
    char* G = (char*)malloc((size_t)cnt * es);
    orc_dm_to_global(src, G);
    if (!transpose) { orc_dm_from_global(dst, G); free(G); return; }
    char* Gt = (char*)malloc((size_t)cnt * es);
    for (int64_t j = 0; j < src->n; ++j)
        for (int64_t i = 0; i < src->m; ++i) memcpy(Gt + (j + i * src->n) * es, G + (i + j * src->m) * es, es);
    orc_dm_from_global(dst, Gt);
    free(G); free(Gt);
}

/* dst(i,j) += alpha * sum_{ranks owning src(i,j) [or src(j,i)], ascending VC rank} value
 * (AxpyContract.hpp:483-544: ReduceScatter(SUM) then B += alpha*recv; sum first, one multiply, one add) */
void orc_axpy_contract_closed(double alpha, const orc_dm* src, orc_dm* dst, int transpose) {
    const orc_grid g = src->g;
    const int scs = orc_dist_stride(g, src->colDist), srs = orc_dist_stride(g, src->rowDist);
    const int dcs = orc_dist_stride(g, dst->colDist), drs = orc_dist_stride(g, dst->rowDist);
    const int P = src->P;
    for (int d = 0; d < P; ++d)
        for (int64_t jl = 0; jl < dst->lw[d]; ++jl)
            for (int64_t il = 0; il < dst->lh[d]; ++il) {
                const int64_t i = dst->colShift[d] + il * dcs, j = dst->rowShift[d] + jl * drs;
                const int64_t si = transpose ? j : i, sj = transpose ? i : j;
                int first = 1;
                double accd = 0; float accf = 0;
                for (int v = 0; v < P; ++v) {
                    const int p = orc_grid_rank_of(g, v % g.r, v / g.r);
                    if (orc_mod(si - src->colShift[p], scs) != 0 || orc_mod(sj - src->rowShift[p], srs) != 0) continue;
                    const int64_t off = (si - src->colShift[p]) / scs + (sj - src->rowShift[p]) / srs * src->ld[p];
                    if (src->es == 8) { const double x = ((const double*)src->loc[p])[off]; accd = first ? x : accd + x; }
                    else { const float x = ((const float*)src->loc[p])[off]; accf = first ? x : accf + x; }
                    first = 0;
                }
                if (first) continue;
                if (dst->es == 8) { volatile double t = alpha * accd; double* q = (double*)dst->loc[d] + il + jl * dst->ld[d]; *q = *q + t; }
                else { volatile float t = (float)alpha * accf; float* q = (float*)dst->loc[d] + il + jl * dst->ld[d]; *q = *q + t; }
            }
}

/* oracle.h -- CPU restatement of the reference's distributed-GEMM path (TEST INFRASTRUCTURE ONLY).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may use
 * this library; the product (libhydrogen_b200.so) never links, loads or calls it.
 *
 * All P ranks of an r x c grid are simulated inside one process: a distributed matrix is an
 * array of P local column-major buffers.  Collectives are memcpy / rank-ascending sums.
 *
 * Pinning, strongest first:
 *   - REFERENCE CODE: indexing (include/El/core/indexing/impl.hpp), every copy::util CPU pack/unpack loop
 *     (include/El/blas_like/level1/Copy/util.hpp) and axpy::util::InterleaveMatrixUpdate (Axpy/util.hpp) compile
 *     from where they lie into oracle/_ref/libref_util.so (oracle/Makefile target `ref`, shim oracle/ref/); the
 *     golden vectors tests/golden/copy_util.npz come from it (tests/golden/make_golden.py) and tests/test_ref_pin.py
 *     holds oracle_core.c / oracle_util.c to them bit for bit (plus 3000+ random argument sets where _ref exists).
 *   - REFERENCE CODE, one level up: the SUMMA loop headers themselves (src/blas_like/level3/Gemm/{NN,NT,TN,TT}.hpp, all 16
 *     variants and the GEMM_DEFAULT heuristic; Trrk/{Local,NN,NT,TN,TT}.hpp; the stationary-C loops of Syrk/ and Symm/) and the
 *     trapezoid level-1 headers compile from where they lie over stand-in matrix classes on this library's simulated grid
 *     (oracle/ref/ref_summa.cpp, ref_trapezoid.cpp -> oracle/_ref/libref_summa.so, libref_util.so); oracle_gemm.c and
 *     oracle_level3.c reproduce their results bit for bit on the cases of tests/summa_cases.py / trapezoid_cases.py
 *     (digests in tests/golden/summa.json, trapezoid.json).  That pins the order of operations, temporaries, alignment
 *     targets and panel ranges of the restatement -- not the rounding of the third-party local product.
 *   - REFERENCE CODE under the definition itself: the copy:: redistribution primitives (Copy/RowAllGather.hpp ... Exchange.hpp,
 *     *Filter.hpp: fourteen entry points) run SPMD, one thread per simulated rank, over stand-ins (oracle/ref/ref_copy.cpp ->
 *     oracle/_ref/libref_copy.so) and realise exactly orc_redist_closed's ownership definition, aligned or not; their alignment
 *     adoption is orc_copy_adopts (tests/test_ref_pin.py).
 *   - The reference as a whole cannot be built here (needs MPI, LAPACK, Aluminum, NVSHMEM) and ships no
 *     golden vectors, so what only exists above MPI is pinned by the reference's own test PROPERTIES:
 *       tests/core/DistMatrix.cpp:28-76  : B = A for every distribution pair with random alignments,
 *         gather both to [*,*], require bitwise equality  -> tests/test_oracle.py checks every restated
 *         primitive chain against the closed-form ownership definition this way;
 *       tests/blas_like/BasicGemm.cpp:93-94,108 : |GPU - CPU| <= 1e-10 (fp64) / 1e-3 (fp32), alpha=2, beta=3.
 * The bit pattern of GEMM results is "parity unpinned" (vendor BLAS / cuBLAS / NCCL sum orders
 * are third-party) -- hence the Frobenius tolerance of BASELINE.json.
 */
#ifndef HYD_ORACLE_H
#define HYD_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_MC = 0, ORC_MD = 1, ORC_MR = 2, ORC_VC = 3, ORC_VR = 4, ORC_STAR = 5, ORC_CIRC = 6 };

/* indexing -- include/El/core/indexing/impl.hpp:33-66,229-245 */
int64_t orc_mod(int64_t a, int64_t b);
int64_t orc_shift(int64_t rank, int64_t align, int64_t stride);
int64_t orc_length(int64_t n, int64_t shift, int64_t stride);
int64_t orc_max_length(int64_t n, int64_t stride);

typedef struct { int r, c, col_major; } orc_grid;  /* Grid.cpp:95-150 */
int orc_grid_default_height(int size);             /* Grid::DefaultHeight, Grid.cpp:58-64 */
/* distribution algebra (types.hpp:292-337) and the alignment rules (ElementMatrix.cpp:243-300; the copy:: primitives) */
int orc_dist_collect(int u);
int orc_dist_partial(int u);
int orc_dist_partial_union_row(int u, int v);
int orc_dist_partial_union_col(int u, int v);
int orc_align_dim_with(int U, int V, int which, int stride, int dataColDist, int dataColAlign, int dataRowDist, int dataRowAlign,
                       int* align);
int orc_copy_adopts(int src, int srcAlign, int dst, int dstStride, int* align);
void orc_grid_coords(orc_grid g, int rank, int* mc, int* mr);
int orc_grid_rank_of(orc_grid g, int mc, int mr);
int orc_dist_stride(orc_grid g, int dist);
int orc_dist_rank(orc_grid g, int dist, int mc, int mr);

typedef struct orc_dm {
    orc_grid g;
    int colDist, rowDist, colAlign, rowAlign;
    int64_t m, n;
    int es;         /* 8 = double, 4 = float */
    int P;
    int owner;      /* 1 if loc[] buffers belong to this object */
    void** loc;     /* [P] local buffers, indexed by owning-comm rank */
    int64_t *ld, *lh, *lw, *colShift, *rowShift;
} orc_dm;

orc_dm* orc_dm_create(orc_grid g, int64_t m, int64_t n, int colDist, int rowDist, int colAlign, int rowAlign, int es);
void orc_dm_free(orc_dm* a);
void* orc_dm_local(orc_dm* a, int rank);
int64_t orc_dm_lh(const orc_dm* a, int rank);
int64_t orc_dm_lw(const orc_dm* a, int rank);
int64_t orc_dm_ld(const orc_dm* a, int rank);
/* A(i0:i0+h, j0:j0+w) sharing storage with the parent (El::View) */
orc_dm* orc_dm_view(orc_dm* a, int64_t i0, int64_t j0, int64_t h, int64_t w);
/* synthetic inputs: global (i,j) = u(seed, i + j*m) in [-1,1), independent of the grid */
void orc_dm_fill(orc_dm* a, uint64_t seed);
void orc_dm_fill_index(orc_dm* a);               /* (i*131 + j) mod 2^24, exact in fp32 */
void orc_dm_fill_const(orc_dm* a, double v);
int64_t orc_dm_to_global(const orc_dm* a, void* out); /* returns #replica mismatches (bitwise) */
void orc_dm_from_global(orc_dm* a, const void* in);
double orc_uniform(uint64_t seed, uint64_t idx);

/* ---- closed-form definitions (what "B = A" means, from the ownership formulas) ---- */
void orc_redist_closed(const orc_dm* src, orc_dm* dst, int transpose);
/* dst += alpha * sum over all ranks owning (i,j) in src, ascending VC rank; optionally of src^T */
void orc_axpy_contract_closed(double alpha, const orc_dm* src, orc_dm* dst, int transpose);

/* ---- pack / unpack restatements: include/El/blas_like/level1/Copy/util.hpp ---- */
void orc_interleave(int64_t h, int64_t w, const void* A, int64_t csA, int64_t rsA, void* B, int64_t csB, int64_t rsB, int es); /* :122-146 */
void orc_row_strided_pack(int64_t h, int64_t w, int64_t rowAlign, int64_t rowStride, const void* A, int64_t ldA, void* Bp, int64_t portion, int es); /* :148-165 */
void orc_row_strided_unpack(int64_t h, int64_t w, int64_t rowAlign, int64_t rowStride, const void* Ap, int64_t portion, void* B, int64_t ldB, int es); /* :167-184 */
void orc_partial_row_strided_pack(int64_t h, int64_t w, int64_t rowAlign, int64_t rowStride, int64_t uni, int64_t part, int64_t rankPart, int64_t shiftA, const void* A, int64_t ldA, void* Bp, int64_t portion, int es); /* :186-207 */
void orc_partial_row_strided_unpack(int64_t h, int64_t w, int64_t rowAlign, int64_t rowStride, int64_t uni, int64_t part, int64_t rankPart, int64_t shiftB, const void* Ap, int64_t portion, void* B, int64_t ldB, int es); /* :209-230 */
void orc_col_strided_pack(int64_t h, int64_t w, int64_t colAlign, int64_t colStride, const void* A, int64_t ldA, void* Bp, int64_t portion, int es); /* :363-381 */
void orc_col_strided_unpack(int64_t h, int64_t w, int64_t colAlign, int64_t colStride, const void* Ap, int64_t portion, void* B, int64_t ldB, int es); /* :403-421 */
void orc_partial_col_strided_pack(int64_t h, int64_t w, int64_t colAlign, int64_t colStride, int64_t uni, int64_t part, int64_t rankPart, int64_t shiftA, const void* A, int64_t ldA, void* Bp, int64_t portion, int es); /* :464-486 */
void orc_partial_col_strided_unpack(int64_t h, int64_t w, int64_t colAlign, int64_t colStride, int64_t uni, int64_t part, int64_t rankPart, int64_t shiftB, const void* Ap, int64_t portion, void* B, int64_t ldB, int es); /* :510-533 */
void orc_strided_pack(int64_t h, int64_t w, int64_t colAlign, int64_t colStride, int64_t rowAlign, int64_t rowStride, const void* A, int64_t ldA, void* Bp, int64_t portion, int es); /* :671-695 */
void orc_strided_unpack(int64_t h, int64_t w, int64_t colAlign, int64_t colStride, int64_t rowAlign, int64_t rowStride, const void* Ap, int64_t portion, void* B, int64_t ldB, int es); /* :697-722 */
void orc_interleave_update(double alpha, int64_t h, int64_t w, const void* A, int64_t csA, int64_t rsA, void* B, int64_t csB, int64_t rsB, int es); /* Axpy/util.hpp:23-37 */

/* ---- primitive chains (pack -> simulated collective -> unpack), aligned and unaligned branches.
 * Each returns 0, or -1 if (src,dst) is not the pair the primitive serves. dst alignments are inputs. */
int orc_row_allgather(const orc_dm* A, orc_dm* B);          /* Copy/RowAllGather.hpp:18-145        (U,V) -> (U,*)   */
int orc_col_allgather(const orc_dm* A, orc_dm* B);          /* Copy/ColAllGather.hpp:16-181        (U,V) -> (*,V)   */
int orc_col_alltoall_demote(const orc_dm* A, orc_dm* B);    /* Copy/ColAllToAllDemote.hpp:15-126   [MC,MR]->[VC,*] / [MR,MC]->[VR,*] */
int orc_row_alltoall_demote(const orc_dm* A, orc_dm* B);    /* Copy/RowAllToAllDemote.hpp:15-124   [MC,MR]->[*,VR] / [MR,MC]->[*,VC] */
int orc_partial_col_allgather(const orc_dm* A, orc_dm* B);  /* Copy/PartialColAllGather.hpp        [VC,*]->[MC,*] / [VR,*]->[MR,*]   */
int orc_partial_row_allgather(const orc_dm* A, orc_dm* B);  /* Copy/PartialRowAllGather.hpp:16-111 [*,VC]->[*,MC] / [*,VR]->[*,MR]   */
int orc_vector_exchange(const orc_dm* A, orc_dm* B);        /* Copy/Exchange.hpp:130-192           [VC,*]<->[VR,*], [*,VC]<->[*,VR]  */
void orc_local_transpose(const orc_dm* A, orc_dm* B);       /* Transpose.hpp:99-158 on every rank (B's dist = A's transposed) */

/* ---- GEMM ---- */
/* optional vendor BLAS (dlopen'ed): path to a library exporting scipy_dgemm_/scipy_sgemm_ or dgemm_/sgemm_ */
int orc_set_blas(const char* path);
int orc_blas_threads(int n);   /* set (n > 0) / query the registered BLAS' thread count */
void orc_local_gemm(int tA, int tB, int64_t m, int64_t n, int64_t k, double alpha, const void* A, int64_t lda,
                    const void* B, int64_t ldb, double beta, void* C, int64_t ldc, int es);
/* El::Gemm on the simulated grid, following the reference loops (alg: 0 default, 1 A, 2 B, 3 C, 4 Dot) */
int orc_gemm(int orientA, int orientB, int alg, double alpha, orc_dm* A, orc_dm* B, double beta, orc_dm* C, int64_t blocksize);
/* DistMatrixReadProxy<T,T,U,V> (Proxy.hpp:174-260): X itself if it already is [U,V] with the constrained alignments, else a new
 * matrix assigned from X whose unconstrained dimensions adopt alignments as in B = A (orc_copy_adopts); caller frees if != X */
orc_dm* orc_read_proxy(orc_dm* X, int U, int V, int colCon, int colAlign, int rowCon, int rowAlign);
/* dense check: C := alpha op(A) op(B) + beta C on global column-major arrays (long-double accumulation) */
void orc_dense_gemm_ref(int tA, int tB, int64_t m, int64_t n, int64_t k, double alpha, const void* A, int64_t lda,
                        const void* B, int64_t ldb, double beta, void* C, int64_t ldc, int es);

/* ---- distributed BLAS-1 around Gemm (oracle_level1.c) ---- */
int orc_axpy(double alpha, const orc_dm* X, orc_dm* Y, int transpose);   /* Axpy.hpp:161-187, TransposeAxpy.hpp:181-210 */
int orc_hadamard(const orc_dm* A, const orc_dm* B, orc_dm* C);          /* Hadamard.hpp:109-135 */
void orc_scale(double alpha, orc_dm* A);                                 /* Scale.hpp */

/* ---- Trrk / Syrk and the trapezoid level-1 routines (oracle_level3.c).  uplo: 0 LOWER, 1 UPPER (types.hpp:489-493) ---- */
void orc_scale_trapezoid(double alpha, int uplo, orc_dm* A, int64_t offset);                      /* ScaleTrapezoid.hpp:47-92 */
int orc_axpy_trapezoid(int uplo, double alpha, const orc_dm* X, orc_dm* Y, int64_t offset);       /* AxpyTrapezoid.hpp:54-100 */
void orc_make_trapezoidal(int uplo, orc_dm* A, int64_t offset);                                   /* MakeTrapezoidal.hpp:48-90 */
int orc_make_symmetric(int uplo, orc_dm* A);                                                      /* MakeSymmetric.hpp:52-68 */
void orc_set_local_trrk_blocksize(int b);                                                         /* blocksizes.cpp:82-83 */
int orc_trrk(int uplo, int orientA, int orientB, double alpha, orc_dm* A, orc_dm* B, double beta, orc_dm* C, int64_t blocksize); /* Trrk.cpp:98-115 */
int orc_syrk(int uplo, int orient, double alpha, orc_dm* A, double beta, orc_dm* C, int64_t blocksize);                           /* Syrk.cpp:72-90 */

int orc_symm(int side, int uplo, double alpha, orc_dm* A, orc_dm* B, double beta, orc_dm* C, int64_t blocksize);   /* Symm.cpp:56-80; side 0 LEFT, 1 RIGHT */

#ifdef __cplusplus
}
#endif
#endif

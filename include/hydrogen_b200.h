/* hydrogen_b200.h -- C ABI of libhydrogen_b200.so
 *
 * B200-native (sm_100a) implementation of the distributed-GEMM hot path of
 * Hydrogen/Elemental: El::Gemm on DistMatrix<T,MC,MR> (SUMMA A/B/C/Dot), the
 * element-cyclic redistributions it drives, and the local kernels below them.
 *
 * Every entry point is `extern "C"`, takes plain pointers / integers / a
 * cudaStream_t (passed as void*), returns 0 on success and non-zero on error;
 * hyd_last_error() then holds the message (the C++ shim in include/El.hpp
 * rethrows it as std::runtime_error / std::logic_error, the reference's
 * convention: include/El/core/environment/decl.hpp:196-198).
 *
 * All matrices are column-major with an explicit leading dimension, as in the
 * reference (include/El/core/Matrix/decl.hpp).  Buffers are DEVICE pointers
 * owned by the caller unless a function says "host".
 *
 * "ref:" comments cite the reference interface each function replaces
 * (paths relative to the reference repository root).
 */
#ifndef HYDROGEN_B200_H
#define HYDROGEN_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* hyd_stream_t; /* cudaStream_t */

/* ---- errors / library info ------------------------------------------------ */
const char* hyd_last_error(void);
int hyd_version(void);
/* number of kernels this library has launched in the calling process (bench.py's gpu_launches) */
long long hyd_launch_count(void);
/* device allocations (cudaMalloc) made and host-blocking synchronisations (cudaStream/DeviceSynchronize) issued by this
 * library in the calling process: a steady-state hyd_gemm_* / hyd_redist_* / hyd_trrk_* call adds nothing to either
 * (ref: the pooled simple_buffer temporaries, include/hydrogen/utils/SimpleBuffer.hpp + src/core/imports/cub.cpp, and the
 * asynchronous SyncInfo contract, include/hydrogen/SyncInfo.hpp).  hyd_gemm_host_* ends with one synchronisation by contract. */
long long hyd_alloc_count(void);
long long hyd_sync_count(void);

/* Stream fences at the API edge.  ref: MakeMultiSync(master, others...) include/hydrogen/SyncInfo.hpp:136-186, used at
 * src/blas_like/level3/Gemm.cpp:117-119 and include/El/blas_like/level1/Copy/RowAllGather.hpp:25-28: at construction the
 * master (the OUTPUT matrix's stream) waits for the work already enqueued on every other (INPUT) stream, at destruction
 * every other stream waits for the master.  begin/end bracket any hyd_* call that runs on `master` and reads matrices
 * produced on `others`; streams equal to master (or repeated) are skipped.  No host synchronisation.  (The reference's
 * AddSynchronizationPoint has the stream comparison inverted, SyncInfo.hpp:92-107; this is the documented semantic.) */
int hyd_multisync_begin(hyd_stream_t master, const hyd_stream_t* others, int n);
int hyd_multisync_end(hyd_stream_t master, const hyd_stream_t* others, int n);

/* ---- seam S2: local (per-rank) kernels ------------------------------------- */

/* C := alpha*op(A)*op(B) + beta*C, column-major.  transX: 0 = NORMAL, 1 = TRANSPOSE (ADJOINT == TRANSPOSE for reals).
 * ref: gpu_blas::Gemm, include/hydrogen/blas/GPU_BLAS_impl.hpp:631-646 (-> cublasDgemm/Sgemm,
 *      src/core/imports/cublas.cpp:130-148); caller El::Gemm(Matrix<T,GPU>) src/blas_like/level3/Gemm.cpp:104-157.
 *      k == 0 scales C by beta (Gemm.cpp:212-219). */
int hyd_lgemm_f64(int transA, int transB, int64_t m, int64_t n, int64_t k, double alpha,
                  const double* A, int64_t lda, const double* B, int64_t ldb, double beta,
                  double* C, int64_t ldc, hyd_stream_t stream);
int hyd_lgemm_f32(int transA, int transB, int64_t m, int64_t n, int64_t k, float alpha,
                  const float* A, int64_t lda, const float* B, int64_t ldb, float beta,
                  float* C, int64_t ldc, hyd_stream_t stream);
/* fp32 accuracy: large TMA-compatible shapes run as split TF32 (three tcgen05 kind::tf32 MMAs per product: exact products,
 * the tensor core's truncating fp32 accumulation).  Relative Frobenius error ~ 7e-9 * k (7.4e-6 at k = 1024, 5.8e-5 at 8192),
 * inside the stated tolerance 1e-5 * sqrt(k) for every k up to this bound; a launch with a larger k runs on the exact-product
 * FFMA kernel instead (as does everything under HYD_SGEMM=ffma). */
#define HYD_LGEMM_F32_SPLIT_TF32_MAX_K 200000

/* Half-precision operands and result (IEEE fp16 / bfloat16 storage, 2 bytes per element), fp32 accumulation, alpha and beta
 * in fp32, one rounding to the half type at the store.  SURVEY 8f N3.
 * ref: Gemm on Matrix<gpu_half_type,Device::GPU> (src/blas_like/level3/Gemm.cpp:460-467) -> cublasHgemm
 *      (src/core/imports/cublas.cpp:213); bf16 has no counterpart in the reference.
 * Parity-tested on a B200 since round 2 (tests/test_gpu_local.py::test_lgemm_half); the local product of hyd_gemm_{f16,bf16}. */
int hyd_lgemm_f16(int transA, int transB, int64_t m, int64_t n, int64_t k, float alpha, const void* A, int64_t lda,
                  const void* B, int64_t ldb, float beta, void* C, int64_t ldc, hyd_stream_t stream);
int hyd_lgemm_bf16(int transA, int transB, int64_t m, int64_t n, int64_t k, float alpha, const void* A, int64_t lda,
                   const void* B, int64_t ldb, float beta, void* C, int64_t ldc, hyd_stream_t stream);

/* B := A^T (A is m x n, B is n x m), bit-exact.
 * ref: gpu_blas::Copy(TRANSPOSE,...) -> cublas?geam(1,A,0,-) include/hydrogen/blas/GPU_BLAS_impl.hpp:152-172,558-567;
 *      caller El::Transpose(Matrix<T,GPU>) include/El/blas_like/level1/Transpose.hpp:161-181. */
int hyd_transpose_f64(int64_t m, int64_t n, const double* A, int64_t lda, double* B, int64_t ldb, hyd_stream_t stream);
int hyd_transpose_f32(int64_t m, int64_t n, const float* A, int64_t lda, float* B, int64_t ldb, hyd_stream_t stream);
/* B := A (m x n, both with leading dimensions).  ref: gpu_blas::Copy(NORMAL,...) GPU_BLAS_impl.hpp:84-106;
 *      copy_2d_kernel src/hydrogen/blas/gpu/Copy.cu:24-69. */
int hyd_copy2d_f64(int64_t m, int64_t n, const double* A, int64_t lda, double* B, int64_t ldb, hyd_stream_t stream);
int hyd_copy2d_f32(int64_t m, int64_t n, const float* A, int64_t lda, float* B, int64_t ldb, hyd_stream_t stream);
/* B += alpha*A.  ref: gpu_blas::Axpy 2-D GPU_BLAS_impl.hpp:108-132; axpy_2d_tiled_kernel src/hydrogen/blas/gpu/Axpy.cu:70-98
 *      (the tail of axpy::util::InterleaveMatrixUpdate, include/El/blas_like/level1/Axpy/util.hpp:39-51). */
int hyd_axpy2d_f64(int64_t m, int64_t n, double alpha, const double* A, int64_t lda, double* B, int64_t ldb, hyd_stream_t stream);
int hyd_axpy2d_f32(int64_t m, int64_t n, float alpha, const float* A, int64_t lda, float* B, int64_t ldb, hyd_stream_t stream);
/* A *= alpha (alpha == 0 stores exact zeros).  ref: gpu_blas::Scale 2-D GPU_BLAS_impl.hpp:190-209; El::Scale(beta,C) in
 *      src/blas_like/level3/Gemm.cpp:287. */
int hyd_scale2d_f64(int64_t m, int64_t n, double alpha, double* A, int64_t lda, hyd_stream_t stream);
int hyd_scale2d_f32(int64_t m, int64_t n, float alpha, float* A, int64_t lda, hyd_stream_t stream);

/* A(i,j) := alpha over an m x n block.  ref: hydrogen::Fill_GPU_impl(height, width, alpha, buffer, ldim, stream)
 *      include/hydrogen/blas/gpu/Fill.hpp:19-21, src/hydrogen/blas/gpu/Fill.cu; callers El::Fill / El::Zero
 *      (include/El/blas_like/level1/{Fill,Zero}.hpp). */
int hyd_fill_f64(int64_t m, int64_t n, double alpha, double* A, int64_t lda, hyd_stream_t stream);
int hyd_fill_f32(int64_t m, int64_t n, float alpha, float* A, int64_t lda, hyd_stream_t stream);
/* Z(i,j) := X(i,j) * Y(i,j); every operand has its own (colStride, rowStride) as in the reference; Z may be X or Y.
 * ref: hydrogen::Hadamard_GPU_impl src/hydrogen/blas/gpu/Hadamard.cu:60-112; caller El::Hadamard
 *      include/El/blas_like/level1/Hadamard.hpp:22-107. */
int hyd_hadamard_f64(int64_t m, int64_t n, const double* X, int64_t colStrideX, int64_t rowStrideX, const double* Y,
                     int64_t colStrideY, int64_t rowStrideY, double* Z, int64_t colStrideZ, int64_t rowStrideZ, hyd_stream_t stream);
int hyd_hadamard_f32(int64_t m, int64_t n, const float* X, int64_t colStrideX, int64_t rowStrideX, const float* Y,
                     int64_t colStrideY, int64_t rowStrideY, float* Z, int64_t colStrideZ, int64_t rowStrideZ, hyd_stream_t stream);
/* B += alpha*op(A), B is m x n; transA != 0: A is n x m.  ref: gpu_blas::Axpy(TransposeMode, ...) -> geam(op(A), B)
 *      include/hydrogen/blas/GPU_BLAS_impl.hpp:108-132,531-541; axpy_2d_transpose_tiled_kernel src/hydrogen/blas/gpu/Axpy.cu:14-68. */
int hyd_axpy_trans_f64(int transA, int64_t m, int64_t n, double alpha, const double* A, int64_t lda, double* B, int64_t ldb, hyd_stream_t stream);
int hyd_axpy_trans_f32(int transA, int64_t m, int64_t n, float alpha, const float* A, int64_t lda, float* B, int64_t ldb, hyd_stream_t stream);
/* 1-D forms with increments.  ref: gpu_blas::Copy / Axpy / Scale (size, X, incx, ...) GPU_BLAS_impl.hpp:510-518,550-556,590-597
 *      (-> cublas?copy / ?axpy / ?scal). */
int hyd_copy1d_f64(int64_t n, const double* X, int64_t incx, double* Y, int64_t incy, hyd_stream_t stream);
int hyd_copy1d_f32(int64_t n, const float* X, int64_t incx, float* Y, int64_t incy, hyd_stream_t stream);
int hyd_axpy1d_f64(int64_t n, double alpha, const double* X, int64_t incx, double* Y, int64_t incy, hyd_stream_t stream);
int hyd_axpy1d_f32(int64_t n, float alpha, const float* X, int64_t incx, float* Y, int64_t incy, hyd_stream_t stream);
int hyd_scale1d_f64(int64_t n, double alpha, double* X, int64_t incx, hyd_stream_t stream);
int hyd_scale1d_f32(int64_t n, float alpha, float* X, int64_t incx, hyd_stream_t stream);

/* ---- seam S2: the reference's pack / unpack utilities, same argument lists (device buffers) --------------------
 * ref: copy::util::* in include/El/blas_like/level1/Copy/util.hpp -- InterleaveMatrix :122-146 (GPU :244-267),
 *      RowStridedPack/Unpack :148-184 (:269-309), PartialRowStridedPack/Unpack :186-230 (:311-359),
 *      ColStridedPack/Unpack :363-421, PartialColStridedPack/Unpack :464-533, StridedPack/Unpack :671-722;
 *      axpy::util::InterleaveMatrixUpdate include/El/blas_like/level1/Axpy/util.hpp:23-51.
 * As in the reference, "colStride" is the element distance between consecutive rows of a column (the stride of the
 * column distribution) and "rowStride" the distance between consecutive columns; portion k starts at
 * Portions + k*portionSize.  The reference issues one 2-D copy per portion; here all portions are one kernel launch.
 * Every function exists as _f64 (double) and _f32 (float). */
int hyd_interleave_f64(int64_t height, int64_t width, const double* A, int64_t colStrideA, int64_t rowStrideA,
                       double* B, int64_t colStrideB, int64_t rowStrideB, hyd_stream_t stream);
int hyd_interleave_update_f64(double alpha, int64_t height, int64_t width, const double* A, int64_t colStrideA, int64_t rowStrideA,
                              double* B, int64_t colStrideB, int64_t rowStrideB, hyd_stream_t stream);
int hyd_row_strided_pack_f64(int64_t height, int64_t width, int64_t rowAlign, int64_t rowStride,
                             const double* A, int64_t ALDim, double* BPortions, int64_t portionSize, hyd_stream_t stream);
int hyd_row_strided_unpack_f64(int64_t height, int64_t width, int64_t rowAlign, int64_t rowStride,
                               const double* APortions, int64_t portionSize, double* B, int64_t BLDim, hyd_stream_t stream);
int hyd_partial_row_strided_pack_f64(int64_t height, int64_t width, int64_t rowAlign, int64_t rowStride,
                                     int64_t rowStrideUnion, int64_t rowStridePart, int64_t rowRankPart, int64_t rowShiftA,
                                     const double* A, int64_t ALDim, double* BPortions, int64_t portionSize, hyd_stream_t stream);
int hyd_partial_row_strided_unpack_f64(int64_t height, int64_t width, int64_t rowAlign, int64_t rowStride,
                                       int64_t rowStrideUnion, int64_t rowStridePart, int64_t rowRankPart, int64_t rowShiftB,
                                       const double* APortions, int64_t portionSize, double* B, int64_t BLDim, hyd_stream_t stream);
int hyd_col_strided_pack_f64(int64_t height, int64_t width, int64_t colAlign, int64_t colStride,
                             const double* A, int64_t ALDim, double* BPortions, int64_t portionSize, hyd_stream_t stream);
int hyd_col_strided_unpack_f64(int64_t height, int64_t width, int64_t colAlign, int64_t colStride,
                               const double* APortions, int64_t portionSize, double* B, int64_t BLDim, hyd_stream_t stream);
int hyd_partial_col_strided_pack_f64(int64_t height, int64_t width, int64_t colAlign, int64_t colStride,
                                     int64_t colStrideUnion, int64_t colStridePart, int64_t colRankPart, int64_t colShiftA,
                                     const double* A, int64_t ALDim, double* BPortions, int64_t portionSize, hyd_stream_t stream);
int hyd_partial_col_strided_unpack_f64(int64_t height, int64_t width, int64_t colAlign, int64_t colStride,
                                       int64_t colStrideUnion, int64_t colStridePart, int64_t colRankPart, int64_t colShiftB,
                                       const double* APortions, int64_t portionSize, double* B, int64_t BLDim, hyd_stream_t stream);
int hyd_strided_pack_f64(int64_t height, int64_t width, int64_t colAlign, int64_t colStride, int64_t rowAlign, int64_t rowStride,
                         const double* A, int64_t ALDim, double* BPortions, int64_t portionSize, hyd_stream_t stream);
int hyd_strided_unpack_f64(int64_t height, int64_t width, int64_t colAlign, int64_t colStride, int64_t rowAlign, int64_t rowStride,
                           const double* APortions, int64_t portionSize, double* B, int64_t BLDim, hyd_stream_t stream);

int hyd_interleave_f32(int64_t height, int64_t width, const float* A, int64_t colStrideA, int64_t rowStrideA,
                       float* B, int64_t colStrideB, int64_t rowStrideB, hyd_stream_t stream);
int hyd_interleave_update_f32(float alpha, int64_t height, int64_t width, const float* A, int64_t colStrideA, int64_t rowStrideA,
                              float* B, int64_t colStrideB, int64_t rowStrideB, hyd_stream_t stream);
int hyd_row_strided_pack_f32(int64_t height, int64_t width, int64_t rowAlign, int64_t rowStride,
                             const float* A, int64_t ALDim, float* BPortions, int64_t portionSize, hyd_stream_t stream);
int hyd_row_strided_unpack_f32(int64_t height, int64_t width, int64_t rowAlign, int64_t rowStride,
                               const float* APortions, int64_t portionSize, float* B, int64_t BLDim, hyd_stream_t stream);
int hyd_partial_row_strided_pack_f32(int64_t height, int64_t width, int64_t rowAlign, int64_t rowStride,
                                     int64_t rowStrideUnion, int64_t rowStridePart, int64_t rowRankPart, int64_t rowShiftA,
                                     const float* A, int64_t ALDim, float* BPortions, int64_t portionSize, hyd_stream_t stream);
int hyd_partial_row_strided_unpack_f32(int64_t height, int64_t width, int64_t rowAlign, int64_t rowStride,
                                       int64_t rowStrideUnion, int64_t rowStridePart, int64_t rowRankPart, int64_t rowShiftB,
                                       const float* APortions, int64_t portionSize, float* B, int64_t BLDim, hyd_stream_t stream);
int hyd_col_strided_pack_f32(int64_t height, int64_t width, int64_t colAlign, int64_t colStride,
                             const float* A, int64_t ALDim, float* BPortions, int64_t portionSize, hyd_stream_t stream);
int hyd_col_strided_unpack_f32(int64_t height, int64_t width, int64_t colAlign, int64_t colStride,
                               const float* APortions, int64_t portionSize, float* B, int64_t BLDim, hyd_stream_t stream);
int hyd_partial_col_strided_pack_f32(int64_t height, int64_t width, int64_t colAlign, int64_t colStride,
                                     int64_t colStrideUnion, int64_t colStridePart, int64_t colRankPart, int64_t colShiftA,
                                     const float* A, int64_t ALDim, float* BPortions, int64_t portionSize, hyd_stream_t stream);
int hyd_partial_col_strided_unpack_f32(int64_t height, int64_t width, int64_t colAlign, int64_t colStride,
                                       int64_t colStrideUnion, int64_t colStridePart, int64_t colRankPart, int64_t colShiftB,
                                       const float* APortions, int64_t portionSize, float* B, int64_t BLDim, hyd_stream_t stream);
int hyd_strided_pack_f32(int64_t height, int64_t width, int64_t colAlign, int64_t colStride, int64_t rowAlign, int64_t rowStride,
                         const float* A, int64_t ALDim, float* BPortions, int64_t portionSize, hyd_stream_t stream);
int hyd_strided_unpack_f32(int64_t height, int64_t width, int64_t colAlign, int64_t colStride, int64_t rowAlign, int64_t rowStride,
                           const float* APortions, int64_t portionSize, float* B, int64_t BLDim, hyd_stream_t stream);

/* ---- index arithmetic (host, no GPU needed) ------------------------------------
 * ref: include/El/core/indexing/impl.hpp:33-66 (Length_, MaxLength_), :229-245 (Mod_, Shift_). */
int64_t hyd_mod(int64_t a, int64_t b);
int64_t hyd_shift(int64_t rank, int64_t align, int64_t stride);
int64_t hyd_length(int64_t n, int64_t shift, int64_t stride);
int64_t hyd_max_length(int64_t n, int64_t stride);

/* ---- distributions / matrices ------------------------------------------------------
 * Dist values follow El::Dist (include/El/core/types.hpp:208-216). MD and CIRC are not on the Gemm path. */
enum { HYD_MC = 0, HYD_MD = 1, HYD_MR = 2, HYD_VC = 3, HYD_VR = 4, HYD_STAR = 5, HYD_CIRC = 6 };
/* Orientation (types.hpp:453-458) and GemmAlgorithm (include/El/blas_like/level3.hpp:23-30) */
enum { HYD_NORMAL = 0, HYD_TRANSPOSE = 1, HYD_ADJOINT = 2 };
enum { HYD_GEMM_DEFAULT = 0, HYD_GEMM_SUMMA_A = 1, HYD_GEMM_SUMMA_B = 2, HYD_GEMM_SUMMA_C = 3,
       HYD_GEMM_SUMMA_DOT = 4, HYD_GEMM_CANNON = 5 };

/* One rank's handle on a distributed matrix: the metadata of El::DistMatrix<T,U,V,ELEMENT,GPU>
 * (include/El/core/DistMatrix/Abstract.hpp; shifts derive from alignments as in
 * src/core/DistMatrix/AbstractDistMatrix.cpp:399-430) plus this rank's local device buffer. */
typedef struct hyd_dm {
    void* buf;       /* device pointer to the local column-major block (may be NULL if the block is empty) */
    int64_t ldim;    /* leading dimension of the local block, >= local height */
    int64_t height;  /* GLOBAL height */
    int64_t width;   /* GLOBAL width */
    int32_t colDist, rowDist;   /* HYD_MC ... HYD_STAR */
    int32_t colAlign, rowAlign; /* 0 <= align < stride of the distribution */
    int32_t root;               /* unused for these distributions; kept for layout parity */
    int32_t reserved;
} hyd_dm;

/* ---- communication world + process grid -------------------------------------------
 * ref: El::Grid(mpi::Comm, int height, GridOrder), include/El/core/Grid.hpp:19-56, src/core/Grid.cpp:95-186.
 * One process per GPU: every rank calls hyd_world_create_nccl with the same 128-byte id
 * (hyd_nccl_unique_id on rank 0, broadcast by the launcher: torchrun/MPI/file). */
typedef struct hyd_world hyd_world;
typedef struct hyd_grid hyd_grid;
typedef struct hyd_local_group hyd_local_group;

int hyd_nccl_unique_id(void* out128);
int hyd_world_create_nccl(const void* id128, int rank, int nranks, hyd_world** out);
/* One process per GPU with no launcher library (what mpi::COMM_WORLD / mpi::NewWorldComm() is to a Hydrogen program,
 * include/El/core/imports/mpi.hpp:86,217; src/core/imports/mpi.cpp): rank, size and local rank come from the environment the
 * usual launchers export (RANK / WORLD_SIZE / LOCAL_RANK of torchrun; OMPI_COMM_WORLD_*, PMI_*, SLURM_* otherwise; one rank
 * when none is set), the device is set to local rank mod device count (as hydrogen::gpu::Initialize picks it), and the NCCL id
 * travels through a rendezvous file (HYD_RENDEZVOUS_FILE, else $HYD_RENDEZVOUS_DIR or /tmp /hyd_rdv_<uid>_<MASTER_PORT>_<key>,
 * key = the launcher's job id -- TORCHELASTIC_RUN_ID, SLURM_JOB_ID, PMIX_NAMESPACE -- else the session id, which unlike the
 * parent pid survives wrappers such as `timeout` or `numactl`; two launches at the same time from one shell session without a
 * job id need different MASTER_PORT values or HYD_RENDEZVOUS_FILE).  Rank 0 removes the file once every rank has joined; a
 * file older than a minute is ignored as the leftover of a dead launch.  One node. */
int hyd_world_create_env(hyd_world** out);
/* Its host-only pieces (no GPU needed; exported so that the two-process CPU test can run them): rank / size / local rank
 * from the environment; rank 0 publishes 128 bytes through the rendezvous file and every other rank reads them (polling up to
 * timeout_s seconds); rank 0 removes the file once everybody has joined. */
int hyd_env_world_info(int* rank, int* size, int* local_rank);
int hyd_env_exchange_id(int rank, void* id128, int timeout_s);
int hyd_env_exchange_done(int rank);
/* P rank-threads of one process on one GPU (single-GPU multi-rank runs / tests); each thread takes one rank. */
int hyd_local_group_create(int nranks, hyd_local_group** out);
int hyd_local_group_destroy(hyd_local_group* grp);
int hyd_world_create_local(hyd_local_group* grp, int rank, hyd_world** out);
int hyd_world_destroy(hyd_world* w);
int hyd_world_barrier(hyd_world* w);
/* mpi::Size / mpi::Rank of the owning communicator (include/El/core/imports/mpi.hpp) */
/* the data plane this world moves redistribution traffic on: "nvlink-push(ipc)+nccl" (peer-mapped NVLink stores, NCCL only
 * for bootstrap and barriers), "nccl" (staged pack -> grouped ncclSend/ncclRecv -> unpack; also what a failed IPC mapping
 * falls back to -- the string is how a caller sees which one it got), "local-push" / "local" (rank-threads on one GPU) */
const char* hyd_world_kind(const hyd_world* w);
int hyd_world_size(const hyd_world* w);
int hyd_world_rank(const hyd_world* w);

/* Grid::DefaultHeight (src/core/Grid.cpp:58-64): the grid height El::Grid(comm) picks when none is given --
 * start at floor(sqrt(size)) and INCREASE until it divides size (8 -> 2x4, 4 -> 2x2, a prime p >= 5 -> p x 1). Host function. */
int hyd_grid_default_height(int size);
/* order: 0 = ROW_MAJOR, 1 = COLUMN_MAJOR (types.hpp:409-413); height must divide the world size; height <= 0 picks
 * hyd_grid_default_height(size) as the reference does (Grid.cpp:87). */
int hyd_grid_create(hyd_world* w, int height, int order, hyd_grid** out);
int hyd_grid_destroy(hyd_grid* g);
int hyd_grid_height(const hyd_grid* g);
int hyd_grid_width(const hyd_grid* g);
int hyd_grid_size(const hyd_grid* g);
int hyd_grid_rank(const hyd_grid* g);     /* rank in the owning communicator */
int hyd_grid_row(const hyd_grid* g);      /* MCRank */
int hyd_grid_col(const hyd_grid* g);      /* MRRank */
int hyd_grid_vc_rank(const hyd_grid* g);
int hyd_grid_vr_rank(const hyd_grid* g);
/* engine knobs: overlap panel gathers with the local GEMM (-1 = auto (default): only for the staged NCCL transport,
 * 0 = off, 1 = on); SMs the GEMM leaves to communication kernels while overlapping (default 4, < 0 keeps the value) */
int hyd_grid_set_overlap(hyd_grid* g, int overlap, int sm_reserve);

/* hyd_gemm_host_*: the most column blocks the local C may be cut into so that finished blocks download under the next
 * block's GEMMs (default 8, or HYD_HOST_CBLOCKS at grid creation; 1 = one download after the last panel).  Same value on
 * every rank.  The cuts themselves come from hyd_host_block_plan. */
int hyd_grid_set_host_blocks(hyd_grid* g, int max_blocks);

/* per-kernel device timing: enable (also resets), then read launches and summed CUDA-event time of
 * kind 0 = local GEMM launches, 1 = redistributions, 2 = reduce-scatters issued through this grid,
 * 3 = pack (+ peer push) kernel groups, 4 = unpack kernel groups inside the redistributions,
 * 5 = the upload stream of hyd_gemm_host_* (first to last host-to-device copy of a call), 6 = its downloads of finished blocks. */
int hyd_grid_profile(hyd_grid* g, int enable);
int hyd_grid_profile_read(hyd_grid* g, int kind, long long* launches, double* total_ms);
/* the recorded intervals of `kind` as (start, end) in ms after the first recorded start of `ref_kind` (e.g. kind 0, ref 5: when
 * every local GEMM of a hyd_gemm_host_* call ran, counted from the call's first upload); returns the entries written, -1 on error */
int hyd_grid_profile_trace(hyd_grid* g, int kind, int ref_kind, double* start_ms, double* end_ms, int max_entries);

/* stride / rank / local shape of a distribution on this grid (host; valid for any rank via mc,mr) */
int hyd_dist_stride(const hyd_grid* g, int dist);
int hyd_dist_rank(const hyd_grid* g, int dist);
int hyd_local_shape(const hyd_grid* g, const hyd_dm* a, int64_t* colShift, int64_t* rowShift,
                    int64_t* localHeight, int64_t* localWidth);
/* pure host form for an arbitrary rank of an r x c grid (no handle, no GPU) */
int hyd_local_shape_of(int gridHeight, int gridWidth, int order, int rank, const hyd_dm* a,
                       int64_t* colShift, int64_t* rowShift, int64_t* localHeight, int64_t* localWidth);
/* view of rows [i,i+h) x cols [j,j+w): alignment and buffer offset as El::View / operator()(IR,IR) */
int hyd_view(const hyd_grid* g, const hyd_dm* a, int64_t i, int64_t j, int64_t h, int64_t w, int elemSize,
             hyd_dm* out);

/* Host-only planner of hyd_gemm_host_*: the column blocks (in 128-column tiles of a tm x tn-tile local C) the host-resident
 * stationary-C Gemm computes one after the other so that block j downloads under block j+1's GEMMs.  Minimises
 * sum_j ceil(tm*w_j/sms) + lambda*w_last + mu*(blocks-1) over at most max_blocks (<= 8) blocks; writes the nblk+1 cuts
 * 0 = t_0 < ... < t_nblk = tn and returns nblk.  (No counterpart in the reference: its Gemm has no host-resident form on
 * Device::GPU; the caller copies whole matrices, include/El/blas_like/level1/Copy.hpp:140-207.) */
int hyd_host_block_plan(int64_t tm, int64_t tn, int64_t sms, double lambda, double mu, int max_blocks, int64_t* cuts);

/* Host-only planner: bytes `rank` of an r x c grid sends to / receives from every rank for dst := src (transpose: src^T)
 * in the one-hop exchange that replaces the reference's pack -> collective -> unpack chains
 * (the headers under include/El/blas_like/level1/Copy/).  Arrays have gridHeight*gridWidth entries; entry [rank] is the local part. */
int hyd_redist_plan(int gridHeight, int gridWidth, int order, int rank, const hyd_dm* src, const hyd_dm* dst, int transpose,
                    int elemSize, int64_t* sendBytes, int64_t* recvBytes);

/* ---- seam S2: collectives on the grid's sub-communicators (device buffers, async on `stream`) ---------------------
 * ref: mpi::AllGather / AllToAll / ReduceScatter / SendRecv / Broadcast <T, Device::GPU> with a SyncInfo<GPU>
 *      (src/core/imports/mpi/{AllGather,AllToAll,ReduceScatter,SendRecv,Broadcast}.hpp, include/El/core/imports/mpi.hpp)
 *      on Grid::MCComm() / MRComm() / VCComm() / VRComm() (include/El/core/Grid.hpp:31-54, src/core/Grid.cpp:133-150).
 * HYD_COMM_MC = the ranks of my grid COLUMN indexed by grid row (size = Height), HYD_COMM_MR = my grid ROW indexed by
 * grid column (size = Width), HYD_COMM_VC / VR = all ranks in column-major / row-major order.
 * Counts are BYTES per member (the typed reference calls multiply by sizeof(T)), except hyd_reducescatter_* which is
 * typed because it adds: member k ends with the sum of everybody's k-th block of `count` elements, added in ascending
 * communicator rank (bit-reproducible).  Collective over the GRID: every rank calls, each within its own
 * sub-communicator, in the same order (as every copy:: primitive of the reference does). */
enum { HYD_COMM_MC = 0, HYD_COMM_MR = 1, HYD_COMM_VC = 2, HYD_COMM_VR = 3 };
int hyd_comm_size(const hyd_grid* g, int comm);
int hyd_comm_rank(const hyd_grid* g, int comm);
int hyd_comm_world_rank(const hyd_grid* g, int comm, int commRank); /* owning-communicator rank of member commRank */
int hyd_allgather(hyd_grid* g, int comm, const void* sbuf, int64_t bytesPerMember, void* rbuf, hyd_stream_t stream);
int hyd_alltoall(hyd_grid* g, int comm, const void* sbuf, int64_t bytesPerMember, void* rbuf, hyd_stream_t stream);
int hyd_sendrecv(hyd_grid* g, int comm, const void* sbuf, int64_t sendBytes, int to, void* rbuf, int64_t recvBytes, int from,
                 hyd_stream_t stream);
int hyd_bcast(hyd_grid* g, int comm, void* buf, int64_t bytes, int root, hyd_stream_t stream);
int hyd_reducescatter_f64(hyd_grid* g, int comm, const double* sbuf, double* rbuf, int64_t count, hyd_stream_t stream);
int hyd_reducescatter_f32(hyd_grid* g, int comm, const float* sbuf, float* rbuf, int64_t count, hyd_stream_t stream);

/* ---- seam S1: distributed operations (collective over the grid; async on `stream`) --------- */

/* dst := src (transpose = 0) or dst := src^T (transpose = 1) between ANY two supported distributions/alignments.
 * ref: DistMatrix::operator= tables src/core/DistMatrix/ElementMatrix/<dist>.cpp -> copy:: primitives
 *      (include/El/blas_like/level1/Copy/<primitive>.hpp); El::Transpose(A,B) include/El/blas_like/level1/Transpose.hpp:191-250,318-356. */
int hyd_redist_f64(hyd_grid* g, const hyd_dm* src, const hyd_dm* dst, int transpose, hyd_stream_t stream);
int hyd_redist_f32(hyd_grid* g, const hyd_dm* src, const hyd_dm* dst, int transpose, hyd_stream_t stream);

/* dst += alpha * (sum of src over the ranks along which src's distribution is replicated), optionally of src^T.
 * ref: El::AxpyContract include/El/blas_like/level1/AxpyContract.hpp:483-544 (RowScatter, ColScatter, PartialRowScatter, PartialColScatter, Scatter :17-479),
 *      El::Contract Contract.hpp:14-98, El::TransposeAxpyContract TransposeAxpyContract.hpp:14-48. */
int hyd_axpy_contract_f64(hyd_grid* g, double alpha, const hyd_dm* src, const hyd_dm* dst, int transpose,
                          hyd_stream_t stream);
int hyd_axpy_contract_f32(hyd_grid* g, float alpha, const hyd_dm* src, const hyd_dm* dst, int transpose,
                          hyd_stream_t stream);

/* Y += alpha*X (transpose = 0) or Y += alpha*X^T (transpose = 1) on distributed matrices of ANY two supported distributions:
 * equal distribution and alignments -> local axpy; otherwise X is first redistributed into Y's distribution and alignment.
 * ref: El::Axpy(alpha, ElementalMatrix X, ElementalMatrix Y) include/El/blas_like/level1/Axpy.hpp:161-187;
 *      El::TransposeAxpy include/El/blas_like/level1/TransposeAxpy.hpp. */
int hyd_axpy_dist_f64(hyd_grid* g, double alpha, const hyd_dm* X, const hyd_dm* Y, int transpose, hyd_stream_t stream);
int hyd_axpy_dist_f32(hyd_grid* g, float alpha, const hyd_dm* X, const hyd_dm* Y, int transpose, hyd_stream_t stream);

/* C := alpha op(A) op(B) + beta C.  ref: El::Gemm(orientA,orientB,alpha,A,B,beta,C,alg),
 * include/El/blas_like/level3.hpp:66-84, src/blas_like/level3/Gemm.cpp:278-307; blocksize = El::Blocksize()
 * (src/core/environment.cpp:305-306, default 128). Any supported distribution is accepted for A, B, C
 * (read / read-write proxies, include/El/core/Proxy.hpp:174-260,694-795). */
int hyd_gemm_f64(hyd_grid* g, int orientA, int orientB, int alg, double alpha, const hyd_dm* A, const hyd_dm* B,
                 double beta, const hyd_dm* C, int64_t blocksize, hyd_stream_t stream);
int hyd_gemm_f32(hyd_grid* g, int orientA, int orientB, int alg, float alpha, const hyd_dm* A, const hyd_dm* B,
                 float beta, const hyd_dm* C, int64_t blocksize, hyd_stream_t stream);

/* ---- triangular rank-k updates on the same SUMMA-C engine.  uplo: 0 = LOWER, 1 = UPPER
 * (El::UpperOrLower, include/El/core/types.hpp:489-493).  Only the `uplo` triangle of the square C is read or written.
 * ref: El::Trrk(uplo,orientA,orientB,alpha,A,B,beta,C), src/blas_like/level3/Trrk.cpp:98-115 (ScaleTrapezoid, then
 * trrk::Trrk{NN,NT,TN,TT}: SUMMA-C panels with LocalTrrk, Trrk/Local.hpp:640-740); El::Syrk / El::Herk (real T),
 * src/blas_like/level3/Syrk.cpp:72-90 -> syrk::{LN,LT,UN,UT} (Syrk/LN.hpp:14-50): orient NORMAL is C := alpha A A^T + beta C,
 * TRANSPOSE is C := alpha A^T A + beta C.  Any supported distribution for A, B, C (proxies as in hyd_gemm_*). */
int hyd_trrk_f64(hyd_grid* g, int uplo, int orientA, int orientB, double alpha, const hyd_dm* A, const hyd_dm* B, double beta,
                 const hyd_dm* C, int64_t blocksize, hyd_stream_t stream);
int hyd_trrk_f32(hyd_grid* g, int uplo, int orientA, int orientB, float alpha, const hyd_dm* A, const hyd_dm* B, float beta,
                 const hyd_dm* C, int64_t blocksize, hyd_stream_t stream);
int hyd_syrk_f64(hyd_grid* g, int uplo, int orient, double alpha, const hyd_dm* A, double beta, const hyd_dm* C,
                 int64_t blocksize, hyd_stream_t stream);
int hyd_syrk_f32(hyd_grid* g, int uplo, int orient, float alpha, const hyd_dm* A, float beta, const hyd_dm* C,
                 int64_t blocksize, hyd_stream_t stream);
/* C := alpha A B + beta C (side 0 = LEFT, A m x m) or alpha B A + beta C (side 1 = RIGHT, A n x n), A symmetric with only its
 * `uplo` triangle read (El::LeftOrRight, types.hpp:417-424).  ref: El::Symm / El::Hemm (real T) on AbstractDistMatrix,
 * src/blas_like/level3/Symm.cpp:56-80 -> symm::{LL,LU,RL,RU} (Symm/LL.hpp:153-229 ...).  Synchronises `stream`. */
int hyd_symm_f64(hyd_grid* g, int side, int uplo, double alpha, const hyd_dm* A, const hyd_dm* B, double beta, const hyd_dm* C,
                 int64_t blocksize, hyd_stream_t stream);
int hyd_symm_f32(hyd_grid* g, int side, int uplo, float alpha, const hyd_dm* A, const hyd_dm* B, float beta, const hyd_dm* C,
                 int64_t blocksize, hyd_stream_t stream);
/* Host-only: the SUMMA variant GEMM_DEFAULT selects for an m x n result with inner dimension k (1 A, 2 B, 3 C, 4 Dot).
 * ref: the switch in gemm::SUMMA_NN / NT / TN / TT, src/blas_like/level3/Gemm/NN.hpp:903-925. */
int hyd_gemm_default_alg(int64_t m, int64_t n, int64_t k);
/* Host-only: the alignment an unconstrained dimension distributed `dstDist` adopts in B = A from the same dimension of A
 * (distributed `srcDist`, aligned `srcAlign`) on a gh x gw grid -- the alignment argument of the reference's copy:: primitive for
 * the pair (the headers under include/El/blas_like/level1/Copy/); 0 where none applies.  The library's read proxies use it. */
int hyd_adopted_align(int gh, int gw, int srcDist, int srcAlign, int dstDist);
/* Host-only introspection of the trapezoidal local update (no device call; used by the CPU tests): for local columns
 * [j0, j0+w) of rank `rank`'s block of C on a gh x gw grid, cuts = {d0, d1, t0, t1}: local rows [d0,d1) are updated by a GEMM
 * straight into C (inside the triangle for every column), rows [t0,t1) through the masked temporary, all others not at all. */
int hyd_trrk_plan(int gh, int gw, int order, int rank, const hyd_dm* C, int uplo, int64_t j0, int64_t w, int elem_size,
                  int64_t* cuts);
/* Trapezoid level-1 operations.  LOWER holds the entries with j - i <= offset, UPPER those with j - i >= offset.
 * ref: ScaleTrapezoid(alpha,uplo,A,offset) include/El/blas_like/level1/ScaleTrapezoid.hpp:14-92 (A *= alpha inside);
 *      AxpyTrapezoid(uplo,alpha,X,Y,offset) AxpyTrapezoid.hpp:14-100 (Y += alpha X inside; X, Y same distribution and
 *      alignments); MakeTrapezoidal(uplo,A,offset) MakeTrapezoidal.hpp (A = 0 outside).
 * Local column-major blocks (Matrix<T,GPU>) and DistMatrix forms (any supported distribution). */
int hyd_scale_trapezoid_f64(double alpha, int uplo, int64_t m, int64_t n, double* A, int64_t lda, int64_t offset, hyd_stream_t stream);
int hyd_scale_trapezoid_f32(float alpha, int uplo, int64_t m, int64_t n, float* A, int64_t lda, int64_t offset, hyd_stream_t stream);
int hyd_axpy_trapezoid_f64(int uplo, double alpha, int64_t m, int64_t n, const double* X, int64_t ldx, double* Y, int64_t ldy,
                           int64_t offset, hyd_stream_t stream);
int hyd_axpy_trapezoid_f32(int uplo, float alpha, int64_t m, int64_t n, const float* X, int64_t ldx, float* Y, int64_t ldy,
                           int64_t offset, hyd_stream_t stream);
int hyd_make_trapezoidal_f64(int uplo, int64_t m, int64_t n, double* A, int64_t lda, int64_t offset, hyd_stream_t stream);
int hyd_make_trapezoidal_f32(int uplo, int64_t m, int64_t n, float* A, int64_t lda, int64_t offset, hyd_stream_t stream);
int hyd_scale_trapezoid_dist_f64(hyd_grid* g, double alpha, int uplo, const hyd_dm* A, int64_t offset, hyd_stream_t stream);
int hyd_scale_trapezoid_dist_f32(hyd_grid* g, float alpha, int uplo, const hyd_dm* A, int64_t offset, hyd_stream_t stream);
int hyd_axpy_trapezoid_dist_f64(hyd_grid* g, int uplo, double alpha, const hyd_dm* X, const hyd_dm* Y, int64_t offset,
                                hyd_stream_t stream);
int hyd_axpy_trapezoid_dist_f32(hyd_grid* g, int uplo, float alpha, const hyd_dm* X, const hyd_dm* Y, int64_t offset,
                                hyd_stream_t stream);
int hyd_make_trapezoidal_dist_f64(hyd_grid* g, int uplo, const hyd_dm* A, int64_t offset, hyd_stream_t stream);
int hyd_make_trapezoidal_dist_f32(hyd_grid* g, int uplo, const hyd_dm* A, int64_t offset, hyd_stream_t stream);
/* A := the symmetric matrix whose `uplo` triangle is A's.  ref: MakeSymmetric(uplo, A) / MakeHermitian (real T) on
 * ElementalMatrix, include/El/blas_like/level1/MakeSymmetric.hpp:52-68 (MakeTrapezoidal, Transpose, AxpyTrapezoid of the
 * strictly other triangle).  Synchronises `stream` (a temporary holds A^T). */
int hyd_make_symmetric_dist_f64(hyd_grid* g, int uplo, const hyd_dm* A, hyd_stream_t stream);
int hyd_make_symmetric_dist_f32(hyd_grid* g, int uplo, const hyd_dm* A, hyd_stream_t stream);

/* ---- fp16 / bf16 storage, fp32 accumulation (parity-tested on a B200: tests/test_gpu_half.py). ----
 * El::Gemm on DistMatrix<gpu_half_type,...,Device::GPU> (Gemm.cpp:278-307 with cublasHgemm as the local product, Gemm.cpp:460-467);
 * bf16 has no counterpart in the reference.  Stationary-C only (alg = GEMM_DEFAULT or GEMM_SUMMA_C); C is rounded to its
 * storage type once per k-panel, as in the reference's loop.  Any supported distribution for A, B, C. */
int hyd_gemm_f16(hyd_grid* g, int orientA, int orientB, int alg, float alpha, const hyd_dm* A, const hyd_dm* B, float beta,
                 const hyd_dm* C, int64_t blocksize, hyd_stream_t stream);
int hyd_gemm_bf16(hyd_grid* g, int orientA, int orientB, int alg, float alpha, const hyd_dm* A, const hyd_dm* B, float beta,
                  const hyd_dm* C, int64_t blocksize, hyd_stream_t stream);
/* dst := src (or src^T) for 2-byte elements of any kind (fp16, bf16: bits are moved, never interpreted), bit-exact.
 * ref: DistMatrix<gpu_half_type,U,V>::operator= tables as for hyd_redist_f64. */
int hyd_redist_b16(hyd_grid* g, const hyd_dm* src, const hyd_dm* dst, int transpose, hyd_stream_t stream);

/* Same operation with the local blocks of A, B, C in HOST memory (buf = host pointer; pinned memory makes the copies
 * asynchronous): uploads the blocks, runs hyd_gemm_*, downloads C's block and synchronises `stream`.  This is the call a
 * CPU-resident caller of El::Gemm makes (the reference's inter-device DistMatrix copies: MC_STAR.cpp:304-390, then Gemm). */
int hyd_gemm_host_f64(hyd_grid* g, int orientA, int orientB, int alg, double alpha, const hyd_dm* A, const hyd_dm* B,
                      double beta, const hyd_dm* C, int64_t blocksize, hyd_stream_t stream);
int hyd_gemm_host_f32(hyd_grid* g, int orientA, int orientB, int alg, float alpha, const hyd_dm* A, const hyd_dm* B,
                      float beta, const hyd_dm* C, int64_t blocksize, hyd_stream_t stream);

/* Inter-device copies of local blocks: `height` x `width` elements of `elem_size` bytes (4 or 8), column-major
 * with leading dimensions in elements; kind: 0 = host -> device, 1 = device -> host, 2 = device -> device.  Asynchronous with
 * respect to the host when the host side is pinned; ordered on `stream`.
 * ref: InterDeviceCopy<SrcD,DestD>::MemCopy2DAsync / MemCopy1DAsync (include/hydrogen/device/gpu/CUDA.hpp:229-420), callers
 *      Copy / CopyAsync(Matrix<T,CPU>, Matrix<T,GPU>) and back (include/El/blas_like/level1/Copy.hpp:140-207) and the
 *      cross-device DistMatrix assignments (src/core/DistMatrix/ElementMatrix/MC_STAR.cpp:304-390). */
int hyd_memcopy2d_async(void* dst, int64_t dst_ldim, const void* src, int64_t src_ldim, int64_t height, int64_t width,
                        int elem_size, int kind, hyd_stream_t stream);
int hyd_memcopy1d_async(void* dst, const void* src, int64_t count, int elem_size, int kind, hyd_stream_t stream);

/* ---- allocation-free variant of the triangular local update -----------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).
 * Triangle `uplo` (0 LOWER, 1 UPPER, of the GLOBAL matrix) of   C := alpha op(A) op(B) + beta C   on ONE rank's local block:
 * local entry (i, j) of the localHeight x localWidth block C is global (colShift + i*colStride, rowShift + j*rowStride);
 * op(A) is localHeight x k, op(B) is k x localWidth, all column-major device buffers of the caller.  It is what hyd_trrk_f64 /
 * hyd_syrk_f64 run per k-panel on every rank (shifts 0 and strides 1: a plain local Trrk), with the one temporary of that
 * computation -- the strip of each 512-column block that the diagonal crosses -- held in a file-scope static __device__
 * array of HYD_LOCAL_TRRK_STATIC_SCRATCH_BYTES instead of a workspace that grows with cudaMalloc.
 *   - calls no cudaMalloc / cudaFree / allocating routine; asynchronous on `stream`;
 *   - a strip larger than the array (about 512 * rowStride / colStride rows x 512 columns) is rejected with an error before
 *     anything is launched -- use hyd_trrk_f64 then;
 *   - every call in the process shares the array, which keeps the last call's contents (nothing in it needs to start at zero:
 *     each strip is written with beta = 0 before it is read); calls that overlap on the GPU race on it -- serialise them;
 *   - every GPU that loads the library holds the array until the process exits. */
#define HYD_LOCAL_TRRK_STATIC_SCRATCH_BYTES (4 * 1024 * 1024)
int hyd_local_trrk_static_f64(int uplo, int transA, int transB, int64_t localHeight, int64_t localWidth, int64_t k,
                              int64_t colShift, int64_t colStride, int64_t rowShift, int64_t rowStride, double alpha,
                              const double* A, int64_t lda, const double* B, int64_t ldb, double beta, double* C, int64_t ldc,
                              hyd_stream_t stream);

/* Pipe-peak probe (register-resident loops): kind 0 = FP64 DFMA, 1 = DMMA m8n8k4, 2 = DMMA m16n8k16,
 * 3 = FP32 FFMA, 4 = dense TF32 on the 5th-generation tensor cores (tcgen05.mma.kind::tf32, 128 x 256 x 8 per instruction,
 * operands resident in shared memory, accumulator in TMEM; ctas_per_sm is ignored: one CTA per SM).  Writes achieved
 * TFLOP/s.  Measurement aid for the roofline denominators: the fp64 kernel is bounded by kind 2, the split-TF32 fp32
 * kernel (three TF32 MMAs per fp32 product) by a third of kind 4. */
int hyd_probe_peak(int kind, int iters, int ctas_per_sm, double* tflops);

#ifdef __cplusplus
}
#endif
#endif /* HYDROGEN_B200_H */

// hyd_local_trrk_static_f64: the triangular local update of Trrk / Syrk with its scratch in a file-scope device array.
//
// NO COUNTERPART IN THE REFERENCE -- added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  It is a variant of the one kernel
// sequence of this library whose temporary is not an operand: local_trrk (summa.cu), the reference's LocalTrrk
// (src/blas_like/level3/Trrk/Local.hpp:640-740), which computes the strip of C that the global diagonal crosses into a
// temporary and adds it under a mask.  There the temporary is a per-grid workspace grown with cudaMalloc on first use; here
// it is g_trrk_scratch below: 4 MiB that every GPU which loads this library holds until the process exits.
//
//   * nothing is allocated or freed: no cudaMalloc / cudaFree / cudaMallocAsync, no allocating library routine;
//   * the scratch is SHARED by every call in the process and keeps the last call's contents.  Nothing in it has to start at
//     zero -- every strip is written by a GEMM with beta = 0 (which never reads its output) before the masked add reads it --
//     so it is never cleared;
//   * for the same reason calls that OVERLAP on the GPU (two streams, two threads) race on it: serialise them;
//   * a strip that does not fit (strongly rectangular grids: the strip of a 512-column block is about
//     512 * rowStride / colStride rows) is rejected with an error before anything is launched.
// The existing code paths (hyd_trrk_*, hyd_syrk_*) are untouched and keep their growable workspace.
#include "grid.h"
#include "hyd_internal.h"
#include "../../include/hydrogen_b200.h"
#include <algorithm>
#include <mutex>

namespace hyd {
namespace {

constexpr i64 kScratchDoubles = HYD_LOCAL_TRRK_STATIC_SCRATCH_BYTES / 8;
constexpr i64 kBlock = 512;   // local columns per block, as local_trrk (summa.cu kTrrkBlock)

// In host code this name is a host-side stand-in, not a device address: a kernel given &g_trrk_scratch[0] directly would
// fault (or, where the GPU can reach host memory, silently use the wrong memory).  scratch_ptr() asks the runtime for the
// device address, once per device.
static __device__ __align__(256) double g_trrk_scratch[kScratchDoubles];

double* scratch_ptr() {
    static std::mutex mu;
    static double* ptr[64] = {};
    int dev = -1;
    if (check_cuda(cudaGetDevice(&dev), "hyd_local_trrk_static: cudaGetDevice")) return nullptr;
    if (dev < 0 || dev >= 64) { set_error("hyd_local_trrk_static: device ordinal %d out of range", dev); return nullptr; }
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr[dev]) {   // with device `dev` current: the symbol has one instance per device
        void* p = nullptr;
        if (check_cuda(cudaGetSymbolAddress(&p, g_trrk_scratch), "hyd_local_trrk_static: cudaGetSymbolAddress")) return nullptr;
        ptr[dev] = static_cast<double*>(p);
    }
    return ptr[dev];
}

inline i64 pad2(i64 lh) { return std::max<i64>(2, (lh + 1) / 2 * 2); }   // 16-byte columns of doubles (TMA)

}  // namespace
}  // namespace hyd

extern "C" int hyd_local_trrk_static_f64(int uplo, int transA, int transB, int64_t localHeight, int64_t localWidth, int64_t k,
                                         int64_t colShift, int64_t colStride, int64_t rowShift, int64_t rowStride, double alpha,
                                         const double* A, int64_t lda, const double* B, int64_t ldb, double beta, double* C,
                                         int64_t ldc, hyd_stream_t stream) {
    using namespace hyd;
    const char* who = "hyd_local_trrk_static_f64";
    if (uplo != 0 && uplo != 1) return set_error("%s: uplo must be LOWER (0) or UPPER (1)", who);
    if (localHeight < 0 || localWidth < 0 || k < 0) return set_error("%s: negative dimension", who);
    if (colStride < 1 || rowStride < 1 || colShift < 0 || colShift >= colStride || rowShift < 0 || rowShift >= rowStride)
        return set_error("%s: need stride >= 1 and 0 <= shift < stride", who);
    if (localHeight == 0 || localWidth == 0) return 0;
    const int tA = transA ? 1 : 0, tB = transB ? 1 : 0;
    const i64 lh = localHeight, lw = localWidth;
    if (!C || ldc < lh) return set_error("%s: bad C / ldc", who);
    if (k > 0 && (!A || !B || lda < std::max<i64>(1, tA ? k : lh) || ldb < std::max<i64>(1, tB ? lw : k)))
        return set_error("%s: bad A / B / leading dimension", who);
    LocalShape L;
    L.colShift = colShift; L.rowShift = rowShift; L.lh = lh; L.lw = lw; L.colStride = (int)colStride; L.rowStride = (int)rowStride;
    const i64 q = 2;   // 16 bytes of doubles
    // every strip must fit the static scratch: checked for all column blocks BEFORE anything is launched
    for (i64 j0 = 0; j0 < lw && k > 0; j0 += kBlock) {
        const i64 w = std::min(kBlock, lw - j0);
        const TrrkCuts cut = trrk_cuts(uplo, L, j0, w, q);
        const i64 th = cut.t1 - cut.t0;
        if (th > 0 && pad2(th) * w > kScratchDoubles)
            return set_error("%s: the diagonal strip of local columns [%lld, %lld) is %lld x %lld doubles, beyond the static scratch of "
                             "%lld bytes (HYD_LOCAL_TRRK_STATIC_SCRATCH_BYTES); use hyd_trrk_f64, whose workspace grows",
                             who, (long long)j0, (long long)(j0 + w), (long long)th, (long long)w,
                             (long long)HYD_LOCAL_TRRK_STATIC_SCRATCH_BYTES);
    }
    cudaStream_t s = (cudaStream_t)stream;
    // ScaleTrapezoid(beta, uplo, C) first, as El::Trrk does (Trrk.cpp:98-115); then the update accumulates with beta = 1
    if (beta != 1.0 && trapezoid_local<double>(2, uplo, lh, lw, colShift, colStride, rowShift, rowStride, 0, beta, nullptr, 0, C, ldc, s))
        return 1;
    if (k == 0 || alpha == 0.0) return 0;
    double* D = scratch_ptr();
    if (!D) return 1;
    const i64 r = colStride, c = rowStride;
    for (i64 j0 = 0; j0 < lw; j0 += kBlock) {
        const i64 w = std::min(kBlock, lw - j0);
        const TrrkCuts cut = trrk_cuts(uplo, L, j0, w, q);
        const double* Bj = tB ? B + j0 : B + j0 * ldb;
        double* Cj = C + j0 * ldc;
        if (cut.d1 > cut.d0) {   // rows inside the triangle for every column of the block: straight into C
            const double* Ai = tA ? A + cut.d0 * lda : A + cut.d0;
            ++g_launches;
            if (lgemm_f64_ex(tA, tB, cut.d1 - cut.d0, w, k, alpha, Ai, lda, Bj, ldb, 1.0, Cj + cut.d0, ldc, s, 0)) return 1;
        }
        if (cut.t1 > cut.t0) {   // the strip the diagonal crosses: into the scratch (beta = 0: never read), then a masked add
            const i64 th = cut.t1 - cut.t0, ldt = pad2(th);
            const double* Ai = tA ? A + cut.t0 * lda : A + cut.t0;
            ++g_launches;
            if (lgemm_f64_ex(tA, tB, th, w, k, alpha, Ai, lda, Bj, ldb, 0.0, D, ldt, s, 0)) return 1;
            if (trapezoid_local<double>(1, uplo, th, w, colShift + cut.t0 * r, r, rowShift + j0 * c, c, 0, 1.0, D, ldt, Cj + cut.t0, ldc, s))
                return 1;
        }
    }
    return 0;
}

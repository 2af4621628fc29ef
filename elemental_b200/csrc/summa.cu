// Distributed GEMM: C := alpha op(A) op(B) + beta C on an r x c grid -- SUMMA A / B / C / Dot.
//
// Reference entry point and algorithms this replaces:
//   El::Gemm(orientA,orientB,alpha,A,B,beta,C,alg)   src/blas_like/level3/Gemm.cpp:278-307
//     -> Scale(beta,C) then gemm::SUMMA_{NN,NT,TN,TT}  Gemm/NN.hpp:881-932, NT.hpp:353-403,
//        TN.hpp:350-400, TT.hpp:379-437 (GEMM_DEFAULT heuristic, blockSizeDot = 2000)
//     -> SUMMA_xxA (NN.hpp:107-206, NT.hpp:15-55, TN.hpp:15-57, TT.hpp:17-61)
//        SUMMA_xxB (NN.hpp:244-287, NT.hpp:96-138, TN.hpp:98-137, TT.hpp:105-152)
//        SUMMA_xxC (NN.hpp:325-367, NT.hpp:178-221, TN.hpp:178-217, TT.hpp:195-240)
//        SUMMA_xxDot (NN.hpp:792-841, NT.hpp:265-310, TN.hpp:262-307, TT.hpp:287-333)
//
// Not a port: one engine serves all sixteen variants.
//   C   : per k-panel gather op(A)(:,K1) as [MC,*] and op(B)(K1,:)^T as [MR,*] (each ONE
//         lattice exchange, transposes folded into the pack), then a local rank-nb update
//         C_loc += alpha Ap Bp^T with the sm_100a DMMA kernel.  Panels are double-buffered and
//         gathered on a separate stream so NVLink traffic hides under the previous update.
//   A/B : the moving operand's panel is gathered in the layout the stationary operand's local
//         block needs, multiplied locally, and the partial result is reduce-scattered straight
//         into the [MC,MR] panel of C (one lattice reduce; the reference's Contract -> [MR,MC]
//         -> Axpy detour is a single hop here).
//   Dot : operands in [*,VC]/[VC,*], local products of <= 2000 x 2000 tiles, P-wide lattice reduce.
// The k-summation order differs from the reference only in rounding (tolerance parity);
// the data movement is the same permutation (bit-exact).
#include "grid.h"
#include "hyd_internal.h"
#include <algorithm>
#include <stdlib.h>
#include <string.h>
#include <type_traits>
#include <vector>

namespace hyd {

namespace {

template <typename T>
int lgemm(Grid& g, int tA, int tB, i64 m, i64 n, i64 k, T alpha, const T* A, i64 lda, const T* B, i64 ldb, T beta, T* C,
          i64 ldc, cudaStream_t s, int sm_reserve) {
    ++g_launches;
    NvtxRange nvtx("Gemm_impl.GPU");
    ProfScope prof(g, Grid::PROF_LGEMM, s);
    if constexpr (std::is_same<T, double>::value)
        return lgemm_f64_ex(tA, tB, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, s, sm_reserve);
    else
        return lgemm_f32_ex(tA, tB, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, s, sm_reserve);
}

inline i64 pad_ld(i64 lh, int es) {  // TMA wants 16-byte multiples; keep at least 1
    const i64 q = 16 / es;
    return std::max<i64>(q, (lh + q - 1) / q * q);
}

// temporary distributed matrix living in workspace `w`
template <typename T>
int make_tmp(Grid& g, Workspace& w, i64 h, i64 wd, int cd, int rd, int ca, int ra, DM& out) {
    out = DM();
    out.height = h; out.width = wd; out.colDist = cd; out.rowDist = rd; out.colAlign = ca; out.rowAlign = ra;
    const LocalShape L = local_shape(g.shape, out, g.mc, g.mr);
    out.ldim = pad_ld(L.lh, sizeof(T));
    if (ws_reserve(w, (size_t)(out.ldim * std::max<i64>(L.lw, 1)) * sizeof(T))) return 1;
    out.buf = w.buf;
    return 0;
}

int ensure_engine(Grid& g) {
    if (g.comm_stream[0]) return 0;
    int lo, hi;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    for (int i = 0; i < 2; ++i) {
        if (check_cuda(cudaStreamCreateWithPriority(&g.comm_stream[i], cudaStreamNonBlocking, hi), "stream create")) return 1;
        cudaEventCreateWithFlags(&g.ev_panel_ready[i], cudaEventDisableTiming);
        cudaEventCreateWithFlags(&g.ev_panel_free[i], cudaEventDisableTiming);
    }
    cudaEventCreateWithFlags(&g.ev_fork, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&g.ev_join, cudaEventDisableTiming);
    return 0;
}

// Trapezoidal local update, C_loc(tri) += alpha op(A_loc) op(B_loc): the reference's LocalTrrk
// (src/blas_like/level3/Trrk/Local.hpp:640-740) without its recursion.  `tri` 0 = LOWER, 1 = UPPER, of the GLOBAL matrix:
// local (i,j) is global (colShift + i*colStride, rowShift + j*rowStride).  The local block is cut into column blocks of
// kTrrkBlock local columns.  In each, the rows split three ways by the global diagonal:
//   * wholly inside the triangle for every column of the block -> one GEMM straight into C (beta = 1);
//   * wholly outside -> nothing (the flops the triangle saves);
//   * the strip the diagonal crosses (about kTrrkBlock*rowStride/colStride local rows) -> GEMM into a temporary, then a
//     masked add (trapezoid_kernel OP 1, alpha = 1: one rounding, like the reference's AxpyTrapezoid of its DTL/DBR temporaries).
// Row cuts are rounded to 16 bytes outwards so every sub-GEMM keeps TMA-aligned operands.
constexpr i64 kTrrkBlock = 512;
template <typename T>
int local_trrk(Grid& g, int tri, int tA, int tB, const LocalShape& Lc, i64 kb, T alpha, const T* A, i64 lda, const T* B, i64 ldb,
               T* C, i64 ldc, cudaStream_t s, int sm_reserve) {
    const i64 q = 16 / (i64)sizeof(T);
    const i64 r = Lc.colStride, c = Lc.rowStride, lh = Lc.lh, lw = Lc.lw;
    {   // the widest strip any column block can have, reserved once (growing a workspace synchronises the device)
        const i64 wmax = std::min(kTrrkBlock, lw), thmax = std::min(lh, ((wmax - 1) * c) / r + 2 + 2 * q);
        if (lw > 0 && ws_reserve(g.trrk_ws, (size_t)(pad_ld(thmax, sizeof(T)) * wmax) * sizeof(T))) return 1;
    }
    for (i64 j0 = 0; j0 < lw; j0 += kTrrkBlock) {
        const i64 w = std::min(kTrrkBlock, lw - j0);
        const TrrkCuts cut = trrk_cuts(tri, Lc, j0, w, q);
        const i64 d0 = cut.d0, d1 = cut.d1, t0 = cut.t0, t1 = cut.t1;
        const i64 gj_first = Lc.rowShift + j0 * c;
        const T* Bj = tB ? B + j0 : B + j0 * ldb;
        T* Cj = C + j0 * ldc;
        if (d1 > d0) {
            const T* Ai = tA ? A + d0 * lda : A + d0;
            if (lgemm<T>(g, tA, tB, d1 - d0, w, kb, alpha, Ai, lda, Bj, ldb, T(1), Cj + d0, ldc, s, sm_reserve)) return 1;
        }
        if (t1 > t0) {
            const i64 th = t1 - t0, ldt = pad_ld(th, sizeof(T));
            if ((size_t)(ldt * w) * sizeof(T) > g.trrk_ws.cap) return set_error("local_trrk: strip %lld x %lld exceeds the reserved workspace", th, w);
            T* D = reinterpret_cast<T*>(g.trrk_ws.buf);
            const T* Ai = tA ? A + t0 * lda : A + t0;
            if (lgemm<T>(g, tA, tB, th, w, kb, alpha, Ai, lda, Bj, ldb, T(0), D, ldt, s, sm_reserve)) return 1;
            if (trapezoid_local<T>(1, tri, th, w, Lc.colShift + t0 * r, r, gj_first, c, 0, T(1), D, ldt, Cj + t0, ldc, s)) return 1;
        }
    }
    return 0;
}
// the local update of SUMMA-C: the whole block (Gemm), or only its part of the global triangle (Trrk / Syrk)
template <typename T>
int local_update(Grid& g, int tri, int tA, int tB, const LocalShape& Lc, i64 kb, T alpha, const T* A, i64 lda, const T* B, i64 ldb,
                 T beta, T* C, i64 ldc, cudaStream_t s, int sm_reserve) {
    if (tri < 0) return lgemm<T>(g, tA, tB, Lc.lh, Lc.lw, kb, alpha, A, lda, B, ldb, beta, C, ldc, s, sm_reserve);
    return local_trrk<T>(g, tri, tA, tB, Lc, kb, alpha, A, lda, B, ldb, C, ldc, s, sm_reserve);   // beta applied by the caller
}

// Host images of A's and B's local blocks (hyd_gemm_host_*): SUMMA-C streams them to the device on an upload stream, so
// the H2D copies hide under the GEMMs instead of preceding the whole Gemm.  With a host image of C as well the local C is
// computed column block by column block (host_col_cuts) and each finished block goes back to the host on a download stream
// under the next block's GEMMs: stationary-C keeps every element of C open until the last k-panel, so a whole-C SUMMA could
// only start the download when all the flops are done.
// `download`: the engine also writes C back (decided from global facts, identical on every rank); cblocks: most column blocks.
struct HostOperands { const void* A; i64 lda; const void* B; i64 ldb; void* C; i64 ldc; int cblocks; bool download; };

// size of one host <-> device copy of the host-resident Gemm (HYD_HOST_CHUNK_BYTES, default 4 MiB; 0 = whole pieces)
inline i64 host_chunk_bytes() {
    static const i64 v = getenv("HYD_HOST_CHUNK_BYTES") ? atoll(getenv("HYD_HOST_CHUNK_BYTES")) : ((i64)4 << 20);
    return v > 0 ? v : ((i64)1 << 60);
}

// upload this rank's part of the global sub-block rows [i0, i0+h) x columns [j0, j0+w) of X from its host image
template <typename T>
int upload_rect(const GridShape& sh, const DM& X, int mc, int mr, i64 i0, i64 j0, i64 h, i64 w, const void* host, i64 ldh,
                cudaStream_t us) {
    const LocalShape L = local_shape(sh, X, mc, mr);
    const i64 a0 = length_(i0, L.colShift, L.colStride), a1 = length_(i0 + h, L.colShift, L.colStride);
    const i64 b0 = length_(j0, L.rowShift, L.rowStride), b1 = length_(j0 + w, L.rowShift, L.rowStride);
    if (a1 <= a0 || b1 <= b0) return 0;
    if (!host) return set_error("hyd_gemm_host: NULL host buffer for a non-empty local block");
    // In pieces of host_chunk_bytes(): the copy engines serve the streams' copies one whole copy at a time, and the SUMMA
    // panel gathers are copy-engine peer copies too -- behind a 16 MiB upload piece (or a 184 MiB download issued as one copy)
    // every gather waited for the whole transfer, and a step took 2.6 - 5 ms instead of 1.5 while host traffic was in flight
    // (profiles/r02_e2e_probe_8gpu_trace.log).  Small pieces let the peer copies in after at most one of them.
    const i64 cols_per = std::max<i64>(1, host_chunk_bytes() / std::max<i64>(1, (a1 - a0) * (i64)sizeof(T)));
    for (i64 c0 = b0; c0 < b1; c0 += cols_per) {
        const i64 c1 = std::min(b1, c0 + cols_per);
        T* d = static_cast<T*>(X.buf) + a0 + c0 * X.ldim;
        const T* hsrc = static_cast<const T*>(host) + a0 + c0 * ldh;
        if (a1 - a0 == X.ldim && X.ldim == ldh) {   // whole columns, no padding on either side: a contiguous transfer
            if (check_cuda(cudaMemcpyAsync(d, hsrc, (size_t)(X.ldim * (c1 - c0)) * sizeof(T), cudaMemcpyHostToDevice, us), "hyd_gemm_host: H2D"))
                return 1;
        } else if (check_cuda(cudaMemcpy2DAsync(d, (size_t)X.ldim * sizeof(T), hsrc, (size_t)ldh * sizeof(T), (size_t)(a1 - a0) * sizeof(T),
                                                (size_t)(c1 - c0), cudaMemcpyHostToDevice, us),
                              "hyd_gemm_host: H2D (2-D)"))
            return 1;
    }
    return 0;
}

// ---- copy-engine panel gather for SUMMA-C, NN orientation --------------------------------------------------------
// With peer-mapped panel buffers the two panel gathers need no kernel at all: the A panel is a set of whole local
// columns, and -- if the k index inside the panel is stored owner-by-owner instead of interleaved (any permutation of k
// applied to both panels leaves A_p * B_p unchanged) -- the B panel is a set of contiguous local row blocks kept
// k-contiguous ([*,MR], fed to the GEMM as a non-transposed operand).  Every piece is one strided 2-D copy that the copy
// engines push over NVLink straight into the readers' panels, flags go out as 4-byte copies, waits are stream memory
// operations: zero SMs, so the persistent GEMM keeps all of them while the next panel arrives.
// Position of panel-local k index t: owners q = grid row holding B's row k0+t, in ascending q, each owner's rows in
// ascending order:  pos(t) = off[q(t)] + (t - first[q(t)]) / r.
template <typename T>
int gather_panels_ce(Grid& g, const DM& A, const DM& B, const DM& C, i64 k0, i64 kb, i64 nb, int buf, cudaStream_t cs,
                     DM& Ap, DM& Bp) {
    const GridShape& sh = g.shape;
    const int r = sh.r, c = sh.c, P = sh.size(), me = g.rank;
    const i64 es = sizeof(T), q16 = 16 / es;
    const int slotA = 4 + buf, slotB = 6 + buf;
    ProfScope prof(g, Grid::PROF_REDIST, cs);
    const LocalShape La = local_shape(sh, A, g.mc, g.mr), Lb = local_shape(sh, B, g.mc, g.mr), Lc = local_shape(sh, C, g.mc, g.mr);
    // geometry of the panels: identical on every rank of my grid row (A) / grid column (B)
    i64 maxlh = 0, maxlw = 0;
    for (int p = 0; p < P; ++p) {
        int pmc, pmr;
        sh.coords(p, pmc, pmr);
        const LocalShape L = local_shape(sh, C, pmc, pmr);
        maxlh = std::max(maxlh, L.lh);
        maxlw = std::max(maxlw, L.lw);
    }
    const i64 ldA = std::max<i64>(q16, (Lc.lh + q16 - 1) / q16 * q16);
    const i64 ldAmax = std::max<i64>(q16, (maxlh + q16 - 1) / q16 * q16);
    const i64 ldB = std::max<i64>(q16, (nb + q16 - 1) / q16 * q16);  // k-contiguous columns of the B panel
    if (g.world->shared_reserve(slotA, (size_t)(ldAmax * nb * es))) return 1;
    if (g.world->shared_reserve(slotB, (size_t)(ldB * std::max<i64>(maxlw, 1) * es))) return 1;
    Ap = DM(); Bp = DM();
    Ap.buf = g.world->shared_local(slotA); Ap.ldim = ldA; Ap.height = C.height; Ap.width = kb;
    Bp.buf = g.world->shared_local(slotB); Bp.ldim = ldB; Bp.height = kb; Bp.width = C.width;
    // owner blocks of the panel's k range: B's row k0+t lives on grid row (k0 + t + B.colAlign) mod r
    std::vector<i64> first(r), cnt(r), off(r);
    i64 run = 0;
    for (int q = 0; q < r; ++q) {
        first[q] = mod_((i64)q - B.colAlign - k0, r);       // smallest t >= 0 owned by grid row q
        cnt[q] = length_(kb, first[q], r);
        off[q] = run;
        run += cnt[q];
    }
    std::vector<unsigned char> toA(P, 0), fromA(P, 0), toB(P, 0), fromB(P, 0);
    // ---- A: my panel columns t = tA0 + j*c, j < nA
    const i64 tA0 = mod_((i64)g.mr - A.rowAlign - k0, c);
    const i64 nA = length_(kb, tA0, c);
    const i64 jLoc0 = length_(k0, La.rowShift, c);            // local column index of the first one
    for (int pmr = 0; pmr < c; ++pmr) {
        const int p = sh.rank_of(g.mc, pmr);
        if (p != me) {
            if (nA > 0 && La.lh > 0) toA[p] = 1;
            if (length_(kb, mod_((i64)pmr - A.rowAlign - k0, c), c) > 0 && Lc.lh > 0) fromA[p] = 1;
        }
    }
    // ---- B: my panel rows are the contiguous local rows [iLoc0, iLoc0 + nB)
    const i64 iLoc0 = length_(k0, Lb.colShift, r);
    const i64 nB = cnt[g.mc];
    for (int pmc = 0; pmc < r; ++pmc) {
        const int p = sh.rank_of(pmc, g.mr);
        if (p != me) {
            if (nB > 0 && Lb.lw > 0) toB[p] = 1;
            if (cnt[pmc] > 0 && Lc.lw > 0) fromB[p] = 1;
        }
    }
    if (g.world->push_open(slotA, toA.data(), cs) || g.world->push_open(slotB, toB.data(), cs)) return 1;
    if (nA > 0 && La.lh > 0) {
        const i64 gg = gcd_(r, c), R = r / gg;                // columns of one class (j = u mod R) share their B owner
        for (i64 u = 0; u < R && u < nA; ++u) {
            const i64 t = tA0 + u * c;
            const int q = (int)mod_(k0 + t + B.colAlign, r);
            const i64 pos0 = off[q] + (t - first[q]) / r, nv = (nA - u + R - 1) / R;
            const T* src = static_cast<const T*>(A.buf) + (jLoc0 + u) * A.ldim;
            for (int pmr = 0; pmr < c; ++pmr) {
                const int p = sh.rank_of(g.mc, pmr);
                T* dst = reinterpret_cast<T*>(g.world->shared_peer(slotA, p)) + pos0 * ldA;
                if (check_cuda(cudaMemcpy2DAsync(dst, (size_t)((c / gg) * ldA * es), src, (size_t)(R * A.ldim * es), (size_t)(La.lh * es),
                                                 (size_t)nv, cudaMemcpyDeviceToDevice, cs),
                               "panel push (A)"))
                    return 1;
            }
        }
    }
    if (nB > 0 && Lb.lw > 0) {
        const T* src = static_cast<const T*>(B.buf) + iLoc0;
        for (int pmc = 0; pmc < r; ++pmc) {
            const int p = sh.rank_of(pmc, g.mr);
            T* dst = reinterpret_cast<T*>(g.world->shared_peer(slotB, p)) + off[g.mc];
            if (check_cuda(cudaMemcpy2DAsync(dst, (size_t)(ldB * es), src, (size_t)(B.ldim * es), (size_t)(nB * es), (size_t)Lb.lw,
                                             cudaMemcpyDeviceToDevice, cs),
                           "panel push (B)"))
                return 1;
        }
    }
    if (g.world->push_post_ce(slotA, toA.data(), cs) || g.world->push_post_ce(slotB, toB.data(), cs)) return 1;
    return g.world->push_wait(slotA, fromA.data(), cs) || g.world->push_wait(slotB, fromB.data(), cs);
}

constexpr int kHostGatherSmDefault = 1;   // see host_gather_sm in summa_c (8 GPUs: e2e 140 -> 156 TFLOP/s, profiles/r02_e2e_probe_8gpu_sm_gather.log)

// ---- SUMMA-C (stationary C) -------------------------------------------------
template <typename T>
int summa_c(Grid& g, int oA, int oB, i64 m, i64 n, i64 k, T alpha, const DM& A, const DM& B, T beta, const DM& C, i64 nb,
            cudaStream_t s, const HostOperands* ho, int tri = -1) {
    // tri >= 0 (Trrk / Syrk): only the LOWER (0) / UPPER (1) triangle of C is updated; the caller has already applied beta
    // to it (ScaleTrapezoid) and passes beta = 1.
    // `beta` is applied by the first panel's update (C = alpha A1 B1^T + beta C) instead of a separate
    // Scale(beta, C) pass (reference Gemm.cpp:287): one full read+write of C saved, and beta == 0 never reads C.
    const GridShape& sh = g.shape;
    const int es = sizeof(T);
    if (ensure_engine(g)) return 1;
    // NN with matching alignments on the push transport: panels are gathered by the copy engines alone (gather_panels_ce),
    // so the gather always runs under the previous GEMM and the GEMM keeps every SM.
    static const bool ce_enabled = !(getenv("HYD_PANEL_CE") && atoi(getenv("HYD_PANEL_CE")) == 0);
    // Host-resident operands (ho): HYD_HOST_GATHER=sm gathers the panels with SM kernels (peer stores) under the GEMM instead of
    // with the copy engines, which the host <-> device transfers of the same call keep busy (a copy-engine gather step took
    // 2.3 - 5 ms instead of 1.5 while host DMA was in flight: profiles/r02_e2e_probe_8gpu_*.log)
    static const int host_gather_sm = (getenv("HYD_HOST_GATHER") && !strcmp(getenv("HYD_HOST_GATHER"), "sm")) ? 1 :
                                      (getenv("HYD_HOST_GATHER") && !strcmp(getenv("HYD_HOST_GATHER"), "ce")) ? 0 : kHostGatherSmDefault;
    const bool host_sm = ho && host_gather_sm && sh.size() > 1 && g.world->can_push();
    const bool ce = ce_enabled && sh.size() > 1 && g.world->can_push() && oA == 0 && oB == 0 && A.colAlign == C.colAlign &&
                    B.rowAlign == C.rowAlign && g.overlap != 0 && !host_sm;
    // Otherwise: with the push transport a panel gather costs a few % of its GEMM, less than the wave the persistent GEMM
    // loses when it leaves SMs to communication kernels: auto mode overlaps only the staged (NCCL) transport
    const bool overlap = sh.size() > 1 && (ce || host_sm || (g.overlap < 0 ? !g.world->can_push() : g.overlap != 0));
    // panel schedule.  The first panel's gather / upload cannot hide under a GEMM, so it is narrower
    // (nb/4, >= 128): the exposed prologue shrinks 4x; only the grouping of the k-sum changes.
    std::vector<i64> ks;
    {
        const i64 nb0 = (overlap || ho) ? std::min(nb, std::max<i64>(128, nb / 4)) : nb;
        for (i64 k0 = 0; k0 < k; k0 += (k0 == 0 ? nb0 : nb)) ks.push_back(k0);
        ks.push_back(k);
    }
    const size_t npanels = ks.size() - 1;
    // column blocks of C (GLOBAL column cuts, multiples of 128 * c so that every rank's local cut is the same multiple of
    // 128 local columns): one block unless the caller's C lives on the host (ho->C), then host_col_cuts (dist.h)
    std::vector<i64> js(1, 0);
    if (ho && ho->download && ho->cblocks > 1 && tri < 0) {
        const i64 c = sh.c, r = sh.r;
        const i64 tm = (max_length_(m, r) + 127) / 128, tn = (max_length_(n, c) + 127) / 128;
        // download time of one 128-column tile strip of the local C, in waves of the local GEMM (one wave = one 128 x 128 x k
        // tile per SM).  Rates are round numbers: ~45 GB/s D2H, 250 GFLOP/s (fp64 DMMA) / 1.4 TFLOP/s (3xTF32) per SM.
        const double per_sm = es == 8 ? 250e9 : 1.4e12;
        const double lambda = (double)max_length_(m, r) * es * per_sm / (45e9 * 256.0 * (double)std::max<i64>(k, 1));
        i64 cuts[9];
        const int sms = std::max(1, num_sms() - ((overlap && !ce) ? g.sm_reserve : 0));
        const int nblk = host_col_cuts(tm, tn, sms, lambda, 0.05, std::min(ho->cblocks, 8), cuts);
        // experiment knob: HYD_HOST_FIRST_TILES = t cuts the first block at t tile columns (a short first block makes the
        // whole-column upload of its part of B short: see b0_whole below)
        static const i64 first_tiles = getenv("HYD_HOST_FIRST_TILES") ? atoll(getenv("HYD_HOST_FIRST_TILES")) : 0;
        if (first_tiles > 0 && first_tiles < (nblk > 1 ? cuts[1] : tn)) js.push_back(std::min<i64>(n, first_tiles * 128 * c));
        for (int b = 1; b < nblk; ++b) js.push_back(std::min<i64>(n, cuts[b] * 128 * c));
    }
    js.push_back(n);
    const size_t nblk = js.size() - 1;
    const size_t nsteps = nblk * npanels;
    cudaStream_t us = g.comm_stream[1], ds = nullptr;
    // HYD_HOST_B0 = 1: the part of op(B) under the FIRST column block is uploaded whole (full-height columns) before the first
    // panel instead of panel by panel.  The per-panel slices of an operand whose k runs along its rows (B for NN) are 2-D
    // copies with short rows (k_panel / r elements: 4 KiB on the 2 x 4 grid), which the host link moves far below its
    // contiguous rate when every GPU of the box transfers at once.
    static const int b0_env = getenv("HYD_HOST_B0") ? atoi(getenv("HYD_HOST_B0")) : 0;
    const bool b0_whole = ho && b0_env == 1 && oB == 0 && nblk > 1;
    if (ho) {
        // every upload is enqueued now, on the upload stream, in the order the steps consume them: for block 0 panel by
        // panel (A's panel p and the part of B's panel p under block 0), then the rest of B one whole column block at a time
        // (one big transfer each -- for NN a contiguous one -- a whole block ahead of its first use); consumers wait per event
        while (g.upload_ev.size() < npanels + nblk) {
            cudaEvent_t e;
            cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
            g.upload_ev.push_back(e);
        }
        if (check_cuda(cudaEventRecord(g.ev_join, s), "upload fork")) return 1;
        cudaStreamWaitEvent(us, g.ev_join, 0);
        ProfScope prof_up(g, Grid::PROF_UPLOAD, us);   // first to last H2D of this Gemm on the upload stream
        if (b0_whole) {   // block 0's columns of op(B) as one transfer, before the first panel of A
            if (oB ? upload_rect<T>(sh, B, g.mc, g.mr, js[0], 0, js[1] - js[0], k, ho->B, ho->ldb, us)
                   : upload_rect<T>(sh, B, g.mc, g.mr, 0, js[0], k, js[1] - js[0], ho->B, ho->ldb, us))
                return 1;
            cudaEventRecord(g.upload_ev[npanels], us);
        }
        for (size_t p = 0; p < npanels; ++p) {
            const i64 k0 = ks[p], kb = ks[p + 1] - ks[p];
            if (!b0_whole && (oB ? upload_rect<T>(sh, B, g.mc, g.mr, js[0], k0, js[1] - js[0], kb, ho->B, ho->ldb, us)
                                 : upload_rect<T>(sh, B, g.mc, g.mr, k0, js[0], kb, js[1] - js[0], ho->B, ho->ldb, us)))
                return 1;
            if (oA ? upload_rect<T>(sh, A, g.mc, g.mr, k0, 0, kb, m, ho->A, ho->lda, us)
                   : upload_rect<T>(sh, A, g.mc, g.mr, 0, k0, m, kb, ho->A, ho->lda, us))
                return 1;
            cudaEventRecord(g.upload_ev[p], us);
        }
        for (size_t j = 1; j < nblk; ++j) {
            if (oB ? upload_rect<T>(sh, B, g.mc, g.mr, js[j], 0, js[j + 1] - js[j], k, ho->B, ho->ldb, us)
                   : upload_rect<T>(sh, B, g.mc, g.mr, 0, js[j], k, js[j + 1] - js[j], ho->B, ho->ldb, us))
                return 1;
            cudaEventRecord(g.upload_ev[npanels + j], us);
        }
        if (ho->download) {
            if (!g.download_stream) {
                if (check_cuda(cudaStreamCreateWithFlags(&g.download_stream, cudaStreamNonBlocking), "download stream")) return 1;
                cudaEventCreateWithFlags(&g.ev_block_done, cudaEventDisableTiming);
                cudaEventCreateWithFlags(&g.ev_download_done, cudaEventDisableTiming);
            }
            ds = g.download_stream;
            cudaStreamWaitEvent(ds, g.ev_join, 0);
        }
    }
    // stream `w` (compute or comm) must not touch step (j, p)'s operands before their upload has landed
    auto wait_upload = [&](cudaStream_t w, size_t j, size_t p) {
        if (!ho) return;
        if (j == 0) {
            cudaStreamWaitEvent(w, g.upload_ev[p], 0);
            if (b0_whole && p == 0) cudaStreamWaitEvent(w, g.upload_ev[npanels], 0);
        } else if (p == 0) cudaStreamWaitEvent(w, g.upload_ev[npanels + j], 0);
    };
    // a finished column block of C goes back to the host under the next block's GEMMs
    auto download_block = [&](size_t j, const DM& Cj, const LocalShape& Lj) -> int {
        if (!ds) return 0;
        if (check_cuda(cudaEventRecord(g.ev_block_done, s), "block done")) return 1;
        cudaStreamWaitEvent(ds, g.ev_block_done, 0);
        ProfScope prof_down(g, Grid::PROF_DOWNLOAD, ds);
        if (Lj.lh > 0 && Lj.lw > 0) {
            const LocalShape Lfull = local_shape(sh, C, g.mc, g.mr);
            const i64 jl = length_(js[j], Lfull.rowShift, Lfull.rowStride);
            if (!ho->C) return set_error("hyd_gemm_host: NULL host buffer for a non-empty local block of C");
            const i64 cols_per = std::max<i64>(1, host_chunk_bytes() / std::max<i64>(1, Lj.lh * (i64)sizeof(T)));
            for (i64 c0 = 0; c0 < Lj.lw; c0 += cols_per) {   // in pieces: see upload_rect
                const i64 nc = std::min(cols_per, Lj.lw - c0);
                if (check_cuda(cudaMemcpy2DAsync(static_cast<T*>(ho->C) + (jl + c0) * ho->ldc, (size_t)ho->ldc * sizeof(T),
                                                 static_cast<const T*>(Cj.buf) + c0 * Cj.ldim, (size_t)Cj.ldim * sizeof(T),
                                                 (size_t)Lj.lh * sizeof(T), (size_t)nc, cudaMemcpyDeviceToHost, ds),
                               "hyd_gemm_host: D2H block of C"))
                    return 1;
            }
        }
        if (j + 1 == nblk) {   // the caller's stream is complete only when the last download is
            if (check_cuda(cudaEventRecord(g.ev_download_done, ds), "download join")) return 1;
            cudaStreamWaitEvent(s, g.ev_download_done, 0);
        }
        return 0;
    };
    if (sh.size() == 1) {
        // 1x1 grid: [MC,*] and [MR,*] panels coincide with the operands' own storage
        // (the reference degenerates to local copies, RowAllGather.hpp:35-38); feed views directly.
        for (size_t j = 0; j < nblk; ++j) {
            const i64 j0 = js[j], w = js[j + 1] - js[j];
            const DM Cj = nblk == 1 ? C : view(sh, C, 0, 0, 0, j0, m, w, es);
            const LocalShape Lj = local_shape(sh, Cj, 0, 0);
            for (size_t p = 0; p < npanels; ++p) {
                const i64 k0 = ks[p], kb = ks[p + 1] - ks[p];
                wait_upload(s, j, p);
                const T* a = static_cast<const T*>(A.buf) + (oA ? k0 : k0 * A.ldim);
                const T* b = static_cast<const T*>(B.buf) + (oB ? j0 + k0 * B.ldim : k0 + j0 * B.ldim);
                if (local_update<T>(g, tri, oA, oB, Lj, kb, alpha, a, A.ldim, b, B.ldim, k0 == 0 ? beta : T(1),
                                    static_cast<T*>(Cj.buf), Cj.ldim, s, 0))
                    return 1;
            }
            if (download_block(j, Cj, Lj)) return 1;
        }
        return 0;
    }
    cudaStream_t cs = overlap ? g.comm_stream[0] : s;
    const bool direct = g.world->can_push();
    if (overlap) {
        if (check_cuda(cudaEventRecord(g.ev_fork, s), "fork")) return 1;
        cudaStreamWaitEvent(cs, g.ev_fork, 0);
    }
    for (size_t step = 0; step < nsteps; ++step) {
        const size_t j = step / npanels, p = step % npanels;
        const i64 k0 = ks[p], kb = ks[p + 1] - ks[p];
        const i64 j0 = js[j], w = js[j + 1] - js[j];
        const int buf = (int)(step & 1);
        wait_upload(cs, j, p);
        // this block's columns of C and of op(B): views, so every gather below sees an ordinary (narrower) Gemm
        const DM Cj = nblk == 1 ? C : view(sh, C, g.mc, g.mr, 0, j0, m, w, es);
        const DM Bj = nblk == 1 ? B : (oB ? view(sh, B, g.mc, g.mr, j0, 0, w, k, es) : view(sh, B, g.mc, g.mr, 0, j0, k, w, es));
        const LocalShape Lj = local_shape(sh, Cj, g.mc, g.mr);
        DM Ap, Bp;
        const DM A1 = oA ? view(sh, A, g.mc, g.mr, k0, 0, kb, m, es) : view(sh, A, g.mc, g.mr, 0, k0, m, kb, es);
        const DM B1 = oB ? view(sh, Bj, g.mc, g.mr, 0, k0, w, kb, es) : view(sh, Bj, g.mc, g.mr, k0, 0, kb, w, es);
        if (overlap && step >= 2) cudaStreamWaitEvent(cs, g.ev_panel_free[buf], 0);
        if (ce) {
            if (gather_panels_ce<T>(g, A, Bj, Cj, k0, kb, nb, buf, cs, Ap, Bp)) return 1;
        } else if (direct) {
            // panels live in peer-mapped regions: every owner stores its columns straight into every reader's panel
            Ap.height = m; Ap.width = kb; Ap.colDist = MC; Ap.rowDist = STAR; Ap.colAlign = Cj.colAlign;
            Bp.height = w; Bp.width = kb; Bp.colDist = MR; Bp.rowDist = STAR; Bp.colAlign = Cj.rowAlign;
            if (redistribute_direct<T>(g, A1, Ap, oA != 0, 4 + buf, nb, cs)) return 1;
            if (redistribute_direct<T>(g, B1, Bp, oB == 0, 6 + buf, nb, cs)) return 1;
        } else {
            if (make_tmp<T>(g, g.panel[buf], m, nb, MC, STAR, Cj.colAlign, 0, Ap)) return 1;
            if (make_tmp<T>(g, g.panel[2 + buf], w, nb, MR, STAR, Cj.rowAlign, 0, Bp)) return 1;
            Ap.width = kb;
            Bp.width = kb;
            if (redistribute<T>(g, A1, Ap, oA != 0, 0, cs)) return 1;
            if (redistribute<T>(g, B1, Bp, oB == 0, 1, cs)) return 1;
        }
        if (overlap) {
            cudaEventRecord(g.ev_panel_ready[buf], cs);
            cudaStreamWaitEvent(s, g.ev_panel_ready[buf], 0);
        }
        if (local_update<T>(g, tri, 0, ce ? 0 : 1, Lj, kb, alpha, static_cast<const T*>(Ap.buf), Ap.ldim,
                            static_cast<const T*>(Bp.buf), Bp.ldim, k0 == 0 ? beta : T(1), static_cast<T*>(Cj.buf), Cj.ldim, s,
                            (overlap && !ce) ? g.sm_reserve : 0))
            return 1;
        if (overlap) cudaEventRecord(g.ev_panel_free[buf], s);
        // tell the panel's writers it has been consumed (ordered after the GEMM that read it)
        if (direct && (g.world->push_close(4 + buf, s) || g.world->push_close(6 + buf, s))) return 1;
        if (p + 1 == npanels && download_block(j, Cj, Lj)) return 1;
    }
    return 0;
}

// ---- SUMMA-A / SUMMA-B / SUMMA-Dot: software pipeline ------------------------------------------------------------------
// Each panel is  gather (moving operand's panel)  ->  local product  ->  reduce-scatter into C's panel.  The reference runs
// the three one after the other on one stream (NN.hpp:107-206, 244-287, 792-841).  Here, on more than one rank, they are
// three stages on three streams with double-buffered temporaries: while panel i is multiplied, panel i+1 is being gathered
// (comm stream 0, exchange slots 0/1) and panel i-1's partial result is being reduced into C (comm stream 1, slots 2/3).
// The local GEMM leaves sm_reserve SMs to the pack / reduce kernels.  hyd_grid_set_overlap(g, 0, .) turns it off.
struct Pipe {
    Grid& g; cudaStream_t s, cs, rs; bool on;
    Pipe(Grid& g_, cudaStream_t s_, bool want) : g(g_), s(s_), cs(s_), rs(s_), on(want) {
        if (!on) return;
        if (ensure_engine(g)) { on = false; return; }
        for (cudaEvent_t& e : g.ev_pipe) if (!e) cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
        cs = g.comm_stream[0];
        rs = g.comm_stream[1];
        cudaEventRecord(g.ev_fork, s);
        cudaStreamWaitEvent(cs, g.ev_fork, 0);
        cudaStreamWaitEvent(rs, g.ev_fork, 0);
    }
    enum { IN_READY = 0, IN_FREE = 2, OUT_READY = 4, OUT_FREE = 6 };
    void signal(int what, int buf, cudaStream_t on_stream) { if (on) cudaEventRecord(g.ev_pipe[what + buf], on_stream); }
    void await(int what, int buf, cudaStream_t on_stream) { if (on) cudaStreamWaitEvent(on_stream, g.ev_pipe[what + buf], 0); }
    int sm_reserve() const { return on ? g.sm_reserve : 0; }
    // the caller's stream is complete when the last reduces are (every gather has been consumed by a GEMM on s already)
    void join() {
        if (!on) return;
        cudaEventRecord(g.ev_join, rs);
        cudaStreamWaitEvent(s, g.ev_join, 0);
    }
};

// ---- SUMMA-A (stationary A): panels over n, reduce over the comm A's k-dimension lives on ----
template <typename T>
int summa_a(Grid& g, int oA, int oB, i64 m, i64 n, i64 k, T alpha, const DM& A, const DM& B, const DM& C, i64 nb,
            cudaStream_t s) {
    const GridShape& sh = g.shape;
    const LocalShape La = local_shape(sh, A, g.mc, g.mr);
    const int es = sizeof(T);
    Pipe pipe(g, s, sh.size() > 1 && g.overlap != 0 && n > nb);
    i64 step = 0;
    for (i64 n0 = 0; n0 < n; n0 += nb, ++step) {
        const i64 w = std::min(nb, n - n0);
        const int buf = pipe.on ? (int)(step & 1) : 0;
        // op(B)(:, n0:n0+w) is k x w; gather it with k distributed like A's k-dimension
        DM Bp, D1;
        const int kd = oA ? MC : MR;                     // where A's k index lives
        const int ka = oA ? A.colAlign : A.rowAlign;
        if (step >= 2) pipe.await(Pipe::IN_FREE, buf, pipe.cs);
        if (make_tmp<T>(g, g.panel[buf], k, w, kd, STAR, ka, 0, Bp)) return 1;
        const DM B1 = oB ? view(sh, B, g.mc, g.mr, n0, 0, w, k, es) : view(sh, B, g.mc, g.mr, 0, n0, k, w, es);
        if (redistribute<T>(g, B1, Bp, oB != 0, buf, pipe.cs)) return 1;
        pipe.signal(Pipe::IN_READY, buf, pipe.cs);
        pipe.await(Pipe::IN_READY, buf, s);
        if (step >= 2) pipe.await(Pipe::OUT_FREE, buf, s);
        // D1 = alpha op(A_loc) Bp_loc : m_loc x w, [MC,*] (oA = N) or [MR,*] (oA = T), partial over the other comm
        const int md = oA ? MR : MC;
        const int ma = oA ? A.rowAlign : A.colAlign;
        if (make_tmp<T>(g, g.panel[2 + buf], m, w, md, STAR, ma, 0, D1)) return 1;
        const i64 mloc = oA ? La.lw : La.lh, kloc = oA ? La.lh : La.lw;
        if (lgemm<T>(g, oA, 0, mloc, w, kloc, alpha, static_cast<const T*>(A.buf), A.ldim, static_cast<const T*>(Bp.buf),
                     Bp.ldim, T(0), static_cast<T*>(D1.buf), D1.ldim, s, pipe.sm_reserve()))
            return 1;
        if (kloc == 0 && scale_local<T>(mloc, w, T(0), static_cast<T*>(D1.buf), D1.ldim, s)) return 1;
        pipe.signal(Pipe::IN_FREE, buf, s);
        pipe.signal(Pipe::OUT_READY, buf, s);
        pipe.await(Pipe::OUT_READY, buf, pipe.rs);
        const DM C1 = view(sh, C, g.mc, g.mr, 0, n0, m, w, es);
        if (reduce_into<T>(g, T(1), D1, C1, false, 2 + buf, pipe.rs)) return 1;
        pipe.signal(Pipe::OUT_FREE, buf, pipe.rs);
    }
    pipe.join();
    return 0;
}

// ---- SUMMA-B (stationary B): panels over m ----
template <typename T>
int summa_b(Grid& g, int oA, int oB, i64 m, i64 n, i64 k, T alpha, const DM& A, const DM& B, const DM& C, i64 nb,
            cudaStream_t s) {
    const GridShape& sh = g.shape;
    const LocalShape Lb = local_shape(sh, B, g.mc, g.mr);
    const int es = sizeof(T);
    Pipe pipe(g, s, sh.size() > 1 && g.overlap != 0 && m > nb);
    i64 step = 0;
    for (i64 m0 = 0; m0 < m; m0 += nb, ++step) {
        const i64 h = std::min(nb, m - m0);
        const int buf = pipe.on ? (int)(step & 1) : 0;
        // (op(A)(m0:m0+h, :))^T is k x h; gather it with k distributed like B's k-dimension
        DM Ap, D1t;
        const int kd = oB ? MR : MC;
        const int ka = oB ? B.rowAlign : B.colAlign;
        if (step >= 2) pipe.await(Pipe::IN_FREE, buf, pipe.cs);
        if (make_tmp<T>(g, g.panel[buf], k, h, kd, STAR, ka, 0, Ap)) return 1;
        const DM A1 = oA ? view(sh, A, g.mc, g.mr, 0, m0, k, h, es) : view(sh, A, g.mc, g.mr, m0, 0, h, k, es);
        if (redistribute<T>(g, A1, Ap, oA == 0, buf, pipe.cs)) return 1;
        pipe.signal(Pipe::IN_READY, buf, pipe.cs);
        pipe.await(Pipe::IN_READY, buf, s);
        if (step >= 2) pipe.await(Pipe::OUT_FREE, buf, s);
        // D1t = alpha op(B_loc)^T Ap_loc : n_loc x h, [MR,*] (oB = N) or [MC,*] (oB = T)
        const int nd = oB ? MC : MR;
        const int na = oB ? B.colAlign : B.rowAlign;
        if (make_tmp<T>(g, g.panel[2 + buf], n, h, nd, STAR, na, 0, D1t)) return 1;
        const i64 nloc = oB ? Lb.lh : Lb.lw, kloc = oB ? Lb.lw : Lb.lh;
        if (lgemm<T>(g, oB ? 0 : 1, 0, nloc, h, kloc, alpha, static_cast<const T*>(B.buf), B.ldim,
                     static_cast<const T*>(Ap.buf), Ap.ldim, T(0), static_cast<T*>(D1t.buf), D1t.ldim, s, pipe.sm_reserve()))
            return 1;
        if (kloc == 0 && scale_local<T>(nloc, h, T(0), static_cast<T*>(D1t.buf), D1t.ldim, s)) return 1;
        pipe.signal(Pipe::IN_FREE, buf, s);
        pipe.signal(Pipe::OUT_READY, buf, s);
        pipe.await(Pipe::OUT_READY, buf, pipe.rs);
        const DM C1 = view(sh, C, g.mc, g.mr, m0, 0, h, n, es);
        if (reduce_into<T>(g, T(1), D1t, C1, true, 2 + buf, pipe.rs)) return 1;
        pipe.signal(Pipe::OUT_FREE, buf, pipe.rs);
    }
    pipe.join();
    return 0;
}

// ---- SUMMA-Dot: k distributed over all P ranks, P-wide reduce per C tile ----
template <typename T>
int summa_dot(Grid& g, int oA, int oB, i64 m, i64 n, i64 k, T alpha, const DM& A, const DM& B, const DM& C,
              i64 blockM, i64 blockN, cudaStream_t s) {
    const GridShape& sh = g.shape;
    const int es = sizeof(T);
    // operands must be  A: [*,VC] (N) / [VC,*] (T),  B: [VC,*] (N) / [*,VC] (T), B aligned to A (proxy, Proxy.hpp:160-260)
    const LocalShape La = local_shape(sh, A, g.mc, g.mr);
    const i64 kloc = oA ? La.lh : La.lw;
    // two stages: tile i+1 is multiplied while tile i's [*,*] partial result is reduced into C
    Pipe pipe(g, s, sh.size() > 1 && g.overlap != 0 && (m > blockM || n > blockN));
    i64 step = 0;
    for (i64 i0 = 0; i0 < m; i0 += blockM) {
        const i64 mb = std::min(blockM, m - i0);
        for (i64 j0 = 0; j0 < n; j0 += blockN, ++step) {
            const i64 nbk = std::min(blockN, n - j0);
            const int buf = pipe.on ? (int)(step & 1) : 0;
            DM C11;
            if (step >= 2) pipe.await(Pipe::OUT_FREE, buf, s);
            if (make_tmp<T>(g, g.panel[buf], mb, nbk, STAR, STAR, 0, 0, C11)) return 1;
            const T* a = static_cast<const T*>(A.buf) + (oA ? i0 * A.ldim : i0);
            const T* b = static_cast<const T*>(B.buf) + (oB ? j0 : j0 * B.ldim);
            if (lgemm<T>(g, oA, oB, mb, nbk, kloc, alpha, a, A.ldim, b, B.ldim, T(0), static_cast<T*>(C11.buf), C11.ldim,
                         s, pipe.sm_reserve()))
                return 1;
            if (kloc == 0 && scale_local<T>(mb, nbk, T(0), static_cast<T*>(C11.buf), C11.ldim, s)) return 1;
            pipe.signal(Pipe::OUT_READY, buf, s);
            pipe.await(Pipe::OUT_READY, buf, pipe.rs);
            const DM Cv = view(sh, C, g.mc, g.mr, i0, j0, mb, nbk, es);
            if (reduce_into<T>(g, T(1), C11, Cv, false, 2 + buf, pipe.rs)) return 1;
            pipe.signal(Pipe::OUT_FREE, buf, pipe.rs);
        }
    }
    pipe.join();
    return 0;
}

// read-proxy: return `a` itself if it already has the wanted distribution (and alignment when constrained),
// else a redistributed temporary from the grid's pool (reference: DistMatrixReadProxy, include/El/core/Proxy.hpp:174-260)
template <typename T>
int read_proxy(Grid& g, TempScope& pool, const DM& a, int cd, int rd, bool colCon, int ca, bool rowCon, int ra, DM& out,
               cudaStream_t s) {
    const bool ok = a.colDist == cd && a.rowDist == rd && (!colCon || a.colAlign == ca) && (!rowCon || a.rowAlign == ra);
    if (ok) { out = a; return 0; }
    out = DM();
    out.height = a.height; out.width = a.width; out.colDist = cd; out.rowDist = rd;
    // an unconstrained dimension adopts the source's alignment where the distributions are related, as the reference's proxy
    // does through B = A (an [MC,MR] operand read as [VC,*] keeps its MC alignment): fewer peers in the redistribution
    out.colAlign = colCon ? ca : adopted_align(g.shape, a.colDist, a.colAlign, cd);
    out.rowAlign = rowCon ? ra : adopted_align(g.shape, a.rowDist, a.rowAlign, rd);
    const LocalShape L = local_shape(g.shape, out, g.mc, g.mr);
    out.ldim = pad_ld(L.lh, sizeof(T));
    out.buf = pool.take((size_t)(out.ldim * std::max<i64>(L.lw, 1)) * sizeof(T));   // pooled: no allocation in steady state
    if (!out.buf) return 1;
    return redistribute<T>(g, a, out, false, 3, s);
}

}  // namespace

template <typename T>
int dist_gemm(Grid& g, int oA, int oB, int alg, T alpha, const DM& A0, const DM& B0, T beta, const DM& C0,
              i64 blocksize, cudaStream_t s, HostImages* host) {
    oA = oA ? 1 : 0;
    oB = oB ? 1 : 0;
    const i64 m = C0.height, n = C0.width;
    const i64 k = oA ? A0.height : A0.width;
    const i64 am = oA ? A0.width : A0.height, bk = oB ? B0.width : B0.height, bn = oB ? B0.height : B0.width;
    if (am != m || bn != n || bk != k)
        return set_error("hyd_gemm: nonconformal matrices: op(A) %lldx%lld, op(B) %lldx%lld, C %lldx%lld", am, k, bk, bn, m, n);
    if (blocksize <= 0) return set_error("hyd_gemm: blocksize must be positive");
    if (alg == 5) return set_error("hyd_gemm: GEMM_CANNON is CPU-only in the reference (NN.hpp:20-103) and not provided");
    if (alg < 0 || alg > 5) return set_error("hyd_gemm: unsupported Gemm option");
    if (alg == 0) alg = gemm_default_alg(m, n, k);   // GEMM_DEFAULT heuristic, NN.hpp:910-925 (identical in NT/TN/TT)
    TempScope pool(g, s);
    const LocalShape Lc0 = local_shape(g.shape, C0, g.mc, g.mr);
    if (m == 0 || n == 0) return 0;
    // host-resident A/B (hyd_gemm_host_*): SUMMA-C on [MC,MR] operands streams them panel by panel under the
    // GEMMs; every other path uploads both blocks first.
    HostOperands ho{host ? host->A : nullptr, host ? host->lda : 0, host ? host->B : nullptr, host ? host->ldb : 0,
                    host ? host->C : nullptr, host ? host->ldc : 0, g.host_cblocks, false};
    if (host) host->c_downloaded = 0;
    // Decided from GLOBAL facts only (the call is a host call, the algorithm, k, the distributions): a rank whose local
    // block of A or B is empty may pass a NULL host pointer (hydrogen_b200.h: buf may be NULL for an empty block) and must
    // still take the same path -- same panel schedule, same push rounds -- as every other rank.
    const bool stream_ops = host && alg == 3 && k > 0 && A0.colDist == MC && A0.rowDist == MR &&
                            B0.colDist == MC && B0.rowDist == MR;
    if (host && !stream_ops) {   // each operand on its own; an empty local block uploads nothing
        if (upload_rect<T>(g.shape, A0, g.mc, g.mr, 0, 0, A0.height, A0.width, host->A, host->lda, s)) return 1;
        if (upload_rect<T>(g.shape, B0, g.mc, g.mr, 0, 0, B0.height, B0.width, host->B, host->ldb, s)) return 1;
    }
    // Scale(beta, C) (reference Gemm.cpp:287).  SUMMA-C on a C that needs no proxy folds it into the first panel's
    // epilogue; every other path (and k == 0) scales first, as the reference does.
    const bool fuse_beta = (alg == 3) && k > 0 && C0.colDist == MC && C0.rowDist == MR;
    if (!fuse_beta && scale_local<T>(Lc0.lh, Lc0.lw, beta, static_cast<T*>(C0.buf), C0.ldim, s)) return 1;
    DM A, B, C;
    static const char* const kRegion[4][4] = {{"SUMMA.NNA", "SUMMA.NNB", "SUMMA.NNC", "SUMMA.NNDot"},
                                              {"SUMMA.NTA", "SUMMA.NTB", "SUMMA.NTC", "SUMMA.NTDot"},
                                              {"SUMMA.TNA", "SUMMA.TNB", "SUMMA.TNC", "SUMMA.TNDot"},
                                              {"SUMMA.TTA", "SUMMA.TTB", "SUMMA.TTC", "SUMMA.TTDot"}};
    NvtxRange nvtx(kRegion[2 * oA + oB][alg - 1]);
    if (read_proxy<T>(g, pool, C0, MC, MR, false, 0, false, 0, C, s)) return 1;
    int rc;
    if (alg == 4) {
        if (oA) { if (read_proxy<T>(g, pool, A0, VC, STAR, false, 0, false, 0, A, s)) return 1; }
        else    { if (read_proxy<T>(g, pool, A0, STAR, VC, false, 0, false, 0, A, s)) return 1; }
        const int kalign = oA ? A.colAlign : A.rowAlign;
        if (oB) { if (read_proxy<T>(g, pool, B0, STAR, VC, false, 0, true, kalign, B, s)) return 1; }
        else    { if (read_proxy<T>(g, pool, B0, VC, STAR, true, kalign, false, 0, B, s)) return 1; }
        // tile of C reduced at a time.  The reference's default is 2000 x 2000 (NN.hpp SUMMA_NNDot blockSize) to bound the
        // [*,*] temporary; a 2000-row tile of a tall-skinny C leaves most SMs idle, so the tile grows until the
        // temporary reaches 256 MiB (180 GB of HBM): m=16384, n=512 is one tile
        const i64 bn = std::min<i64>(n, 8192);
        const char* tb = getenv("HYD_DOT_TILE_BYTES");   // bound of the [*,*] temporary (tests shrink it to cover the tile loops)
        const i64 tile_bytes = tb ? std::max<i64>(1, atoll(tb)) : ((i64)256 << 20);
        const i64 bm = std::max<i64>(2000, tile_bytes / (bn * (i64)sizeof(T)));
        rc = summa_dot<T>(g, oA, oB, m, n, k, alpha, A, B, C, bm, bn, s);
    } else {
        if (read_proxy<T>(g, pool, A0, MC, MR, false, 0, false, 0, A, s)) return 1;
        if (read_proxy<T>(g, pool, B0, MC, MR, false, 0, false, 0, B, s)) return 1;
        if (alg == 1) rc = summa_a<T>(g, oA, oB, m, n, k, alpha, A, B, C, blocksize, s);
        else if (alg == 2) rc = summa_b<T>(g, oA, oB, m, n, k, alpha, A, B, C, blocksize, s);
        else {
            // the host path writes C's column blocks back itself when nothing sits between the caller's blocks and the
            // engine: A, B, C all [MC,MR] (stream_ops, fuse_beta), hence no proxies -- global facts, the same on every rank
            ho.download = stream_ops && fuse_beta;
            rc = summa_c<T>(g, oA, oB, m, n, k, alpha, A, B, fuse_beta ? beta : T(1), C, blocksize, s, stream_ops ? &ho : nullptr);
            if (!rc && ho.download) host->c_downloaded = 1;
        }
    }
    if (rc) return rc;
    if (C.buf != C0.buf) {  // write-back half of the ReadWrite proxy
        if (redistribute<T>(g, C, C0, false, 3, s)) return 1;
    }
    return 0;   // proxies go back to the pool with an event on s (TempScope): no host synchronisation
}

// El::Trrk(uplo, orientA, orientB, alpha, A, B, beta, C) -- src/blas_like/level3/Trrk.cpp:98-115 -> ScaleTrapezoid(beta, uplo, C),
// then trrk::Trrk{NN,NT,TN,TT} (Trrk/NN.hpp:17-57 ...): the SUMMA-C panel loop with LocalTrrk as the local update.  Here: the same
// SUMMA-C engine as Gemm (one-hop panel gathers, double buffering, copy-engine / push transports) with local_trrk.
template <typename T>
int dist_trrk(Grid& g, int uplo, int oA, int oB, T alpha, const DM& A0, const DM& B0, T beta, const DM& C0, i64 blocksize,
              cudaStream_t s) {
    oA = oA ? 1 : 0;
    oB = oB ? 1 : 0;
    if (uplo != 0 && uplo != 1) return set_error("hyd_trrk: uplo must be LOWER (0) or UPPER (1)");
    const i64 n = C0.height;
    const i64 k = oA ? A0.height : A0.width;
    const i64 am = oA ? A0.width : A0.height, bk = oB ? B0.width : B0.height, bn = oB ? B0.height : B0.width;
    if (C0.width != n || am != n || bn != n || bk != k)
        return set_error("hyd_trrk: nonconformal matrices: op(A) %lldx%lld, op(B) %lldx%lld, C %lldx%lld", am, k, bk, bn, n, C0.width);
    if (blocksize <= 0) return set_error("hyd_trrk: blocksize must be positive");
    if (n == 0) return 0;
    const LocalShape Lc0 = local_shape(g.shape, C0, g.mc, g.mr);
    if (beta != T(1) && trapezoid_local<T>(2, uplo, Lc0.lh, Lc0.lw, Lc0.colShift, Lc0.colStride, Lc0.rowShift, Lc0.rowStride, 0, beta,
                                            nullptr, 0, static_cast<T*>(C0.buf), C0.ldim, s))
        return 1;
    if (k == 0) return 0;
    TempScope pool(g, s);
    DM A, B, C;
    if (read_proxy<T>(g, pool, C0, MC, MR, false, 0, false, 0, C, s)) return 1;
    if ((double)k > 10.0 * (double)n) {
        // A long inner dimension against a small C: panel-by-panel SUMMA-C would gather k/nb thin panels for little work each.
        // The reference's Syrk switches to its Dot variant here (Syrk/LN.hpp:152-159, weightAwayFromDot = 10: operands as
        // [*,VC] / [VC,*], local products, one reduction); so does this: the Gemm engine's Dot variant computes the full n x n
        // product into a temporary aligned with C (2x the triangle's flops on a small matrix), and only the triangle is added.
        const LocalShape Lc = local_shape(g.shape, C, g.mc, g.mr);
        DM Z = C;
        Z.ldim = pad_ld(Lc.lh, sizeof(T));
        Z.buf = pool.take((size_t)(Z.ldim * std::max<i64>(Lc.lw, 1)) * sizeof(T));
        if (!Z.buf) return 1;
        if (dist_gemm<T>(g, oA, oB, 4, alpha, A0, B0, T(0), Z, blocksize, s, nullptr)) return 1;
        if (trapezoid_local<T>(1, uplo, Lc.lh, Lc.lw, Lc.colShift, Lc.colStride, Lc.rowShift, Lc.rowStride, 0, T(1),
                               static_cast<const T*>(Z.buf), Z.ldim, static_cast<T*>(C.buf), C.ldim, s))
            return 1;
        if (C.buf != C0.buf && redistribute<T>(g, C, C0, false, 3, s)) return 1;
        return 0;
    }
    if (read_proxy<T>(g, pool, A0, MC, MR, false, 0, false, 0, A, s)) return 1;
    if (read_proxy<T>(g, pool, B0, MC, MR, false, 0, false, 0, B, s)) return 1;
    if (summa_c<T>(g, oA, oB, n, n, k, alpha, A, B, T(1), C, blocksize, s, nullptr, uplo)) return 1;
    if (C.buf != C0.buf && redistribute<T>(g, C, C0, false, 3, s)) return 1;
    return 0;
}

// ScaleTrapezoid / AxpyTrapezoid / MakeTrapezoidal on a DistMatrix in any distribution: the local block with its global
// index map (ScaleTrapezoid.hpp:47-92, AxpyTrapezoid.hpp, MakeTrapezoidal.hpp).  op as in trapezoid_local.
template <typename T>
int dist_trapezoid(Grid& g, int op, int uplo, T alpha, const DM* X, const DM& Y, i64 offset, cudaStream_t s) {
    if (uplo != 0 && uplo != 1) return set_error("trapezoid: uplo must be LOWER (0) or UPPER (1)");
    const LocalShape L = local_shape(g.shape, Y, g.mc, g.mr);
    TempScope pool(g, s);
    DM Xt;
    if (X) {
        if (X->height != Y.height || X->width != Y.width) return set_error("AxpyTrapezoid: nonconformal matrices");
        if (X->colDist != Y.colDist || X->rowDist != Y.rowDist || X->colAlign != Y.colAlign || X->rowAlign != Y.rowAlign) {
            // the reference's else-branch (AxpyTrapezoid.hpp:87-98): XCopy(Y.Grid()) aligned with Y, XCopy = X, then the
            // local update.  Collective over the grid (every rank sees the same metadata).
            Xt = Y;
            Xt.ldim = pad_ld(L.lh, sizeof(T));
            Xt.buf = pool.take((size_t)(Xt.ldim * std::max<i64>(L.lw, 1)) * sizeof(T));
            if (!Xt.buf) return 1;
            if (redistribute<T>(g, *X, Xt, false, 3, s)) return 1;
            X = &Xt;
        }
    }
    return trapezoid_local<T>(op, uplo, L.lh, L.lw, L.colShift, L.colStride, L.rowShift, L.rowStride, offset, alpha,
                              X ? static_cast<const T*>(X->buf) : nullptr, X ? X->ldim : 0, static_cast<T*>(Y.buf), Y.ldim, s);
}

// MakeSymmetric(uplo, A) on a DistMatrix (include/El/blas_like/level1/MakeSymmetric.hpp:52-68): MakeTrapezoidal(uplo, A);
// ATrans = A^T; AxpyTrapezoid(the other triangle, 1, ATrans, A, +-1).  ATrans takes A's distribution and alignments, so the
// transpose is one redistribution and the add is local.
template <typename T>
int dist_make_symmetric(Grid& g, int uplo, const DM& A, cudaStream_t s) {
    if (uplo != 0 && uplo != 1) return set_error("hyd_make_symmetric: uplo must be LOWER (0) or UPPER (1)");
    if (A.height != A.width) return set_error("hyd_make_symmetric: cannot make non-square matrix symmetric");
    if (A.height == 0) return 0;
    const LocalShape L = local_shape(g.shape, A, g.mc, g.mr);
    if (trapezoid_local<T>(3, uplo, L.lh, L.lw, L.colShift, L.colStride, L.rowShift, L.rowStride, 0, T(0), nullptr, 0,
                           static_cast<T*>(A.buf), A.ldim, s))
        return 1;
    TempScope pool(g, s);
    DM At = A;
    At.ldim = pad_ld(L.lh, sizeof(T));
    At.buf = pool.take((size_t)(At.ldim * std::max<i64>(L.lw, 1)) * sizeof(T));
    if (!At.buf) return 1;
    if (redistribute<T>(g, A, At, true, 3, s)) return 1;
    return trapezoid_local<T>(1, 1 - uplo, L.lh, L.lw, L.colShift, L.colStride, L.rowShift, L.rowStride, uplo == 0 ? 1 : -1, T(1),
                              static_cast<const T*>(At.buf), At.ldim, static_cast<T*>(A.buf), A.ldim, s);
}
// El::Symm(side, uplo, alpha, A, B, beta, C) -- src/blas_like/level3/Symm.cpp:56-80 (C *= beta; symm::{LL,LU,RL,RU}): C := alpha A B +
// beta C (LEFT) or alpha B A + beta C (RIGHT) with A symmetric and only its `uplo` triangle read.  The reference reads the
// triangle twice per panel (once as is, once transposed, each under MakeTrapezoidal: Symm/LL.hpp:153-214); here the
// symmetric completion of A is built once (one transposing redistribution + two trapezoid passes over m^2 entries, against
// 2 m^2 n flops) and the product is the Gemm engine's -- same flops as the reference, every SUMMA variant and transport
// available.  A itself is never written.
template <typename T>
int dist_symm(Grid& g, int side, int uplo, T alpha, const DM& A0, const DM& B0, T beta, const DM& C0, i64 blocksize, cudaStream_t s) {
    if (side != 0 && side != 1) return set_error("hyd_symm: side must be LEFT (0) or RIGHT (1)");
    if (uplo != 0 && uplo != 1) return set_error("hyd_symm: uplo must be LOWER (0) or UPPER (1)");
    const i64 m = C0.height, n = C0.width, na = side == 0 ? m : n;
    if (A0.height != na || A0.width != na || B0.height != m || B0.width != n)
        return set_error("hyd_symm: nonconformal matrices: A %lldx%lld, B %lldx%lld, C %lldx%lld", A0.height, A0.width, B0.height,
                         B0.width, m, n);
    TempScope pool(g, s);
    DM Ah;
    Ah.height = na; Ah.width = na; Ah.colDist = MC; Ah.rowDist = MR;
    const bool mcmr = A0.colDist == MC && A0.rowDist == MR;
    Ah.colAlign = mcmr ? A0.colAlign : 0;
    Ah.rowAlign = mcmr ? A0.rowAlign : 0;
    const LocalShape L = local_shape(g.shape, Ah, g.mc, g.mr);
    Ah.ldim = pad_ld(L.lh, sizeof(T));
    Ah.buf = pool.take((size_t)(Ah.ldim * std::max<i64>(L.lw, 1)) * sizeof(T));
    if (!Ah.buf) return 1;
    if (na > 0) {
        if (redistribute<T>(g, A0, Ah, false, 3, s)) return 1;
        if (dist_make_symmetric<T>(g, uplo, Ah, s)) return 1;
    }
    return side == 0 ? dist_gemm<T>(g, 0, 0, 0, alpha, Ah, B0, beta, C0, blocksize, s, nullptr)
                     : dist_gemm<T>(g, 0, 0, 0, alpha, B0, Ah, beta, C0, blocksize, s, nullptr);
}
template int dist_symm<double>(Grid&, int, int, double, const DM&, const DM&, double, const DM&, i64, cudaStream_t);
template int dist_symm<float>(Grid&, int, int, float, const DM&, const DM&, float, const DM&, i64, cudaStream_t);

template int dist_make_symmetric<double>(Grid&, int, const DM&, cudaStream_t);
template int dist_make_symmetric<float>(Grid&, int, const DM&, cudaStream_t);

template int dist_trrk<double>(Grid&, int, int, int, double, const DM&, const DM&, double, const DM&, i64, cudaStream_t);
template int dist_trrk<float>(Grid&, int, int, int, float, const DM&, const DM&, float, const DM&, i64, cudaStream_t);
template int dist_trapezoid<double>(Grid&, int, int, double, const DM*, const DM&, i64, cudaStream_t);
template int dist_trapezoid<float>(Grid&, int, int, float, const DM*, const DM&, i64, cudaStream_t);
template int dist_gemm<double>(Grid&, int, int, int, double, const DM&, const DM&, double, const DM&, i64, cudaStream_t, HostImages*);
template int dist_gemm<float>(Grid&, int, int, int, float, const DM&, const DM&, float, const DM&, i64, cudaStream_t, HostImages*);

}  // namespace hyd

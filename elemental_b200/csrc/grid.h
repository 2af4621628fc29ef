// Process grid object behind hyd_grid (reference: El::Grid, include/El/core/Grid.hpp:15-146,
// src/core/Grid.cpp:66-216).  Owns the communication world handle, this rank's grid
// coordinates and the persistent staging workspaces that replace the reference's per-call
// simple_buffer allocations (include/hydrogen/utils/SimpleBuffer.hpp; CUB pool, src/core/imports/cub.cpp).
#pragma once
#include "comm.h"
#include "dist.h"
#include "hyd_internal.h"
#include <cuda_runtime.h>
#include <vector>

namespace hyd {

constexpr int kSlots = 4;  // independent in-flight redistributions (SUMMA double-buffers two panels)

struct Workspace {
    char* buf = nullptr;
    size_t cap = 0;
};

// One pooled temporary (the reference's simple_buffer<T,D> on the CUB pool: include/hydrogen/utils/SimpleBuffer.hpp,
// src/core/imports/cub.cpp): persistent device memory plus the event that marks its last use, so that a later borrower
// on another stream is ordered behind it without any host synchronisation.
struct PoolBuf {
    Workspace w;
    cudaEvent_t released = nullptr;
    cudaStream_t last = nullptr;
};

struct Grid {
    World* world = nullptr;
    bool owns_world = false;
    GridShape shape;
    int rank = 0, mc = 0, mr = 0;
    Workspace ws[kSlots];
    // SUMMA engine resources (created lazily)
    cudaStream_t comm_stream[2] = {nullptr, nullptr};
    cudaEvent_t ev_panel_ready[2] = {nullptr, nullptr}, ev_panel_free[2] = {nullptr, nullptr};
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    cudaEvent_t ev_pipe[8] = {};                    // SUMMA-A/B/Dot pipeline: {in ready, in free, out ready, out free} x 2 buffers
    cudaStream_t unpack_stream = nullptr;           // chunk-pipelined redistributions: unpack of chunk q under the push of q+1
    cudaEvent_t ev_unpack_fork = nullptr, ev_unpack_join = nullptr;
    cudaEvent_t ev_chunk[8] = {};                   // chunk q's own part has been packed
    std::vector<cudaEvent_t> upload_ev;  // per-panel H2D completion (hyd_gemm_host_*)
    cudaStream_t download_stream = nullptr;          // hyd_gemm_host_* with HYD_HOST_CBLOCKS > 1: D2H of finished column blocks of C
    cudaEvent_t ev_block_done = nullptr, ev_download_done = nullptr;
    Workspace panel[4];  // gathered panels (A/B x double buffer) or reduce temporaries
    Workspace host_stage[3];  // device images of host operands (hyd_gemm_host_*)
    Workspace coll_ws;        // receive area of hyd_reducescatter_* (coll.cu)
    Workspace level1_ws;      // X in Y's distribution for hyd_axpy_dist_* (level1.cu)
    Workspace trrk_ws;        // diagonal-strip products of the trapezoidal local update (summa.cu local_trrk)
    std::vector<PoolBuf> pool;  // stack of pooled temporaries (proxies, Trrk / Symm intermediates): see TempScope
    size_t pool_top = 0;
    int overlap = -1;    // -1 auto: gather panels under the previous GEMM only when the transport is SM-heavy (staged NCCL); 0 off; 1 on
    int sm_reserve = 4;  // SMs the local GEMM leaves to NCCL / pack kernels while overlapping
    int host_cblocks = 8; // hyd_gemm_host_*: most column blocks of C (1 = one download after the last panel); hyd_grid_set_host_blocks
    long long launches = 0;
    // optional per-kernel timing (bench.py's roofline leg): CUDA event pairs around each launch of a kind
    enum { PROF_LGEMM = 0, PROF_REDIST = 1, PROF_REDUCE = 2, PROF_PACK = 3, PROF_UNPACK = 4, PROF_UPLOAD = 5, PROF_DOWNLOAD = 6, PROF_KINDS = 7 };
    bool profile = false;
    std::vector<cudaEvent_t> prof_ev[PROF_KINDS];   // start/stop pairs, in launch order
    size_t prof_used[PROF_KINDS] = {};

    int vc() const { return shape.vc(mc, mr); }
    int vr() const { return shape.vr(mc, mr); }
};

int ws_reserve(Workspace& w, size_t bytes);
// RAII event pair around a launch sequence on stream s (no-op unless g.profile)
struct ProfScope {
    Grid& g; int kind; cudaStream_t s; bool on;
    ProfScope(Grid& g_, int kind_, cudaStream_t s_) : g(g_), kind(kind_), s(s_), on(g_.profile) {
        if (!on) return;
        auto& v = g.prof_ev[kind];
        while (v.size() < g.prof_used[kind] + 2) { cudaEvent_t e; cudaEventCreate(&e); v.push_back(e); }
        cudaEventRecord(v[g.prof_used[kind]], s);
    }
    ~ProfScope() {
        if (!on) return;
        cudaEventRecord(g.prof_ev[kind][g.prof_used[kind] + 1], s);
        g.prof_used[kind] += 2;
    }
};
void grid_destroy(Grid* g);

// hyd_gemm_host_*: this rank's HOST images of the local blocks (pinned or pageable).  A NULL pointer is legal for an empty
// local block.  c_downloaded is set when the engine has already written C's blocks back to the host itself.
struct HostImages { const void* A; i64 lda; const void* B; i64 ldb; void* C; i64 ldc; int c_downloaded; };

// Borrow temporaries from the grid's pool for the duration of one operation (proxies of non-[MC,MR] operands, the symmetric
// completion in Symm, ...).  Nested operations stack: an inner scope takes the buffers above the outer one's.  Nothing is
// allocated once the pool has seen the largest shape, and nothing blocks the host: a buffer handed back records an event on
// the borrowing stream, and the next borrower waits for it on ITS stream only if that is a different one.
struct TempScope {
    Grid& g; cudaStream_t s; size_t base;
    TempScope(Grid& g_, cudaStream_t s_) : g(g_), s(s_), base(g_.pool_top) {}
    void* take(size_t bytes) {
        if (g.pool_top == g.pool.size()) g.pool.emplace_back();
        PoolBuf& b = g.pool[g.pool_top];
        if (b.released && b.last != s) cudaStreamWaitEvent(s, b.released, 0);
        if (ws_reserve(b.w, bytes ? bytes : 16)) return nullptr;
        ++g.pool_top;
        return b.w.buf;
    }
    ~TempScope() {
        for (size_t i = base; i < g.pool_top; ++i) {
            PoolBuf& b = g.pool[i];
            if (!b.released) cudaEventCreateWithFlags(&b.released, cudaEventDisableTiming);
            cudaEventRecord(b.released, s);
            b.last = s;
        }
        g.pool_top = base;
    }
};

// dst := src (or src^T), any pair of supported element distributions / alignments.
template <typename T>
int redistribute(Grid& g, const DM& src, const DM& dst, bool transpose, int slot, cudaStream_t s);
// Direct form for library-owned destinations (SUMMA panels): dst lives at the base of push region `slot` on every rank
// (ldim = 16-byte-padded local height, room for alloc_width columns).  Every sender stores its part straight into the
// final positions of every receiver's panel through the peer mapping -- one kernel, no staging, no unpack.  Leaves the
// round open: the consumer calls world->push_close(slot, stream) once it has read the panel.  Fills dst.buf / dst.ldim.
template <typename T>
int redistribute_direct(Grid& g, const DM& src, DM& dst, bool transpose, int slot, i64 alloc_width, cudaStream_t s);
// dst += alpha * (sum over the ranks along which src is a partial sum)  (optionally of src^T)
template <typename T>
int reduce_into(Grid& g, T alpha, const DM& src, const DM& dst, bool transpose, int slot, cudaStream_t s);

int redist_plan(const GridShape& sh, int rank, const DM& src, const DM& dst, bool transpose, int es, i64* sendBytes,
                i64* recvBytes);
// view of rows [i0,i0+h) x cols [j0,j0+w) of a distributed matrix, as seen by rank (mc,mr)
DM view(const GridShape& g, const DM& a, int mc, int mr, i64 i0, i64 j0, i64 h, i64 w, int elem_size);

// local BLAS-1-like kernels (blas1.cu)
template <typename T> int transpose_local(i64 m, i64 n, const T* A, i64 lda, T* B, i64 ldb, cudaStream_t s);
template <typename T> int copy2d_local(i64 m, i64 n, const T* A, i64 lda, T* B, i64 ldb, cudaStream_t s);
template <typename T> int axpy2d_local(i64 m, i64 n, T alpha, const T* A, i64 lda, T* B, i64 ldb, cudaStream_t s);
template <typename T> int scale_local(i64 m, i64 n, T beta, T* C, i64 ldc, cudaStream_t s);
// op 1: Y += alpha X, 2: Y *= alpha (inside the trapezoid), 3: Y = 0 outside it; local (i,j) is global (gi0+i*rs, gj0+j*cs)
template <typename T>
int trapezoid_local(int op, int upper, i64 m, i64 n, i64 gi0, i64 rs, i64 gj0, i64 cs, i64 offset, T alpha, const T* X, i64 ldx,
                    T* Y, i64 ldy, cudaStream_t s);


}  // namespace hyd

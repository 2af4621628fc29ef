// NCCL and in-process communication backends (see comm.h).
#include "comm.h"
#include "hyd_internal.h"
#include <nccl.h>
#include <dlfcn.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

namespace hyd {

// NCCL is bound at run time, not link time: a process that already carries an NCCL (e.g. the
// one bundled with PyTorch, loaded before this library) must keep using that single copy, and a
// plain C++ host program gets the system libnccl.so.2.
namespace nccl {
#define HYD_NCCL_FNS(X) \
    X(ncclGetUniqueId) X(ncclCommInitRank) X(ncclCommInitRankConfig) X(ncclCommDestroy) X(ncclGroupStart) X(ncclGroupEnd) X(ncclSend) \
    X(ncclRecv) X(ncclAllReduce) X(ncclGetErrorString) X(ncclAllGather) X(ncclReduceScatter) X(ncclBroadcast)
#define X(f) static decltype(&::f) f = nullptr;
HYD_NCCL_FNS(X)
#undef X
static int load() {
    static int state = 0;  // 0 = not tried, 1 = ok, -1 = failed
    static pthread_mutex_t mu = PTHREAD_MUTEX_INITIALIZER;
    pthread_mutex_lock(&mu);
    if (state == 0) {
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        state = h ? 1 : -1;
#define X(f) if (h) { f = reinterpret_cast<decltype(f)>(dlsym(h, #f)); if (!f) state = -1; }
        HYD_NCCL_FNS(X)
#undef X
    }
    pthread_mutex_unlock(&mu);
    return state == 1 ? 0 : set_error("NCCL (libnccl.so.2) could not be loaded: %s", dlerror() ? dlerror() : "missing symbol");
}
}  // namespace nccl

// ---------------------------------------------------------------------------
// NCCL: one process per GPU
// ---------------------------------------------------------------------------
static int check_nccl(ncclResult_t r, const char* what) {
    if (r == ncclSuccess) return 0;
    return set_error("%s: NCCL error %d (%s)", what, (int)r, nccl::ncclGetErrorString(r));
}

struct NcclWorld : World {
    ncclComm_t comm = nullptr;
    ~NcclWorld() override { if (comm) nccl::ncclCommDestroy(comm); }
    const char* kind() const override { return "nccl"; }
    int exchange(const void* send, const i64* scount, const i64* soff, void* recv, const i64* rcount,
                 const i64* roff, cudaStream_t s) override {
        bool any = false;
        for (int p = 0; p < size; ++p)
            if (p != rank && (scount[p] > 0 || rcount[p] > 0)) any = true;
        if (!any) return 0;
        if (check_nccl(nccl::ncclGroupStart(), "ncclGroupStart")) return 1;
        for (int p = 0; p < size; ++p) {
            if (p == rank) continue;
            if (scount[p] > 0 &&
                check_nccl(nccl::ncclSend((const char*)send + soff[p], (size_t)scount[p], ncclChar, p, comm, s), "ncclSend"))
                return 1;
            if (rcount[p] > 0 &&
                check_nccl(nccl::ncclRecv((char*)recv + roff[p], (size_t)rcount[p], ncclChar, p, comm, s), "ncclRecv"))
                return 1;
        }
        return check_nccl(nccl::ncclGroupEnd(), "ncclGroupEnd");
    }
    int barrier() override {
        // a zero-payload all-reduce would need a buffer; the engine only needs stream order,
        // so the host barrier is an all-reduce on a 4-byte scratch followed by a sync.
        static thread_local int* scratch = nullptr;
        if (!scratch && check_cuda(cudaMalloc(&scratch, 4), "barrier scratch")) return 1;
        if (check_nccl(nccl::ncclAllReduce(scratch, scratch, 1, ncclInt, ncclSum, comm, 0), "ncclAllReduce")) return 1;
        return check_cuda(cudaStreamSynchronize(0), "barrier sync");
    }
};

int nccl_unique_id(void* out) {
    ncclUniqueId id;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    if (nccl::load()) return 1;
    if (check_nccl(nccl::ncclGetUniqueId(&id), "ncclGetUniqueId")) return 1;
    memcpy(out, &id, 128);
    return 0;
}

World* make_nccl_world(const void* uid, int rank, int size) {
    if (nccl::load()) return nullptr;
    ncclUniqueId id;
    memcpy(&id, uid, 128);
    NcclWorld* w = new NcclWorld;
    w->rank = rank;
    w->size = size;
    // Communication kernels must fit into the few SMs the persistent local GEMM leaves free while panels
    // are gathered under it: cap NCCL's CTA count and turn off CGA clusters (a 4-CTA cluster needs 4 free SMs
    // of one GPC).  HYD_NCCL_MAX_CTAS overrides (0 = NCCL defaults).
    ncclConfig_t cfg = NCCL_CONFIG_INITIALIZER;
    int max_ctas = 8;
    if (const char* e = getenv("HYD_NCCL_MAX_CTAS")) max_ctas = atoi(e);
    if (max_ctas > 0) { cfg.minCTAs = 1; cfg.maxCTAs = max_ctas; cfg.cgaClusterSize = 0; }
    if (check_nccl(nccl::ncclCommInitRankConfig(&w->comm, size, id, rank, &cfg), "ncclCommInitRankConfig")) {
        w->comm = nullptr;
        delete w;
        return nullptr;
    }
    return w;
}

// ---------------------------------------------------------------------------
// Local: P rank-threads, one GPU, stream-ordered copies
// ---------------------------------------------------------------------------
struct LocalSlot {
    const char* send = nullptr;
    std::vector<i64> scount, soff;
    cudaEvent_t ready = nullptr, done = nullptr;
};
struct LocalShared {
    int size = 0;
    pthread_barrier_t bar;
    std::vector<LocalSlot> slots;
    int error = 0;
};

LocalShared* local_shared_create(int size) {
    LocalShared* sh = new LocalShared;
    sh->size = size;
    pthread_barrier_init(&sh->bar, nullptr, size);
    sh->slots.resize(size);
    for (auto& s : sh->slots) {
        s.scount.assign(size, 0);
        s.soff.assign(size, 0);
    }
    return sh;
}
void local_shared_destroy(LocalShared* sh) {
    if (!sh) return;
    for (auto& s : sh->slots) {
        if (s.ready) cudaEventDestroy(s.ready);
        if (s.done) cudaEventDestroy(s.done);
    }
    pthread_barrier_destroy(&sh->bar);
    delete sh;
}

struct LocalWorld : World {
    LocalShared* sh = nullptr;
    const char* kind() const override { return "local"; }
    int barrier() override { pthread_barrier_wait(&sh->bar); return 0; }
    int exchange(const void* send, const i64* scount, const i64* soff, void* recv, const i64* rcount,
                 const i64* roff, cudaStream_t s) override {
        LocalSlot& me = sh->slots[rank];
        if (!me.ready) {
            cudaEventCreateWithFlags(&me.ready, cudaEventDisableTiming);
            cudaEventCreateWithFlags(&me.done, cudaEventDisableTiming);
        }
        me.send = (const char*)send;
        for (int p = 0; p < size; ++p) { me.scount[p] = scount[p]; me.soff[p] = soff[p]; }
        int rc = check_cuda(cudaEventRecord(me.ready, s), "local exchange: record ready");
        pthread_barrier_wait(&sh->bar);  // everyone has published
        for (int p = 0; p < size && !rc; ++p) {
            if (p == rank || rcount[p] <= 0) continue;
            const LocalSlot& q = sh->slots[p];
            if (q.scount[rank] != rcount[p]) {
                rc = set_error("local exchange: rank %d expects %lld B from %d, which sends %lld B", rank,
                               rcount[p], p, q.scount[rank]);
                break;
            }
            rc = check_cuda(cudaStreamWaitEvent(s, q.ready, 0), "local exchange: wait ready");
            if (!rc)
                rc = check_cuda(cudaMemcpyAsync((char*)recv + roff[p], q.send + q.soff[rank], (size_t)rcount[p],
                                                cudaMemcpyDeviceToDevice, s),
                                "local exchange: copy");
        }
        if (!rc) rc = check_cuda(cudaEventRecord(me.done, s), "local exchange: record done");
        pthread_barrier_wait(&sh->bar);  // everyone has enqueued its reads
        for (int p = 0; p < size && !rc; ++p) {
            if (p == rank || scount[p] <= 0) continue;
            rc = check_cuda(cudaStreamWaitEvent(s, sh->slots[p].done, 0), "local exchange: wait done");
        }
        pthread_barrier_wait(&sh->bar);  // slots/events may now be reused
        return rc;
    }
};

World* make_local_world(LocalShared* sh, int rank) {
    LocalWorld* w = new LocalWorld;
    w->sh = sh;
    w->rank = rank;
    w->size = sh->size;
    return w;
}

}  // namespace hyd

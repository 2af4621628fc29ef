// NCCL and in-process communication backends (see comm.h).
#include "comm.h"
#include "hyd_internal.h"
#include "push.h"
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
    // push transport over CUDA-IPC peer mappings (NVLink): bootstrapped through this NCCL communicator
    PushCore pc;
    bool push_ok = false;
    char* hs_dev = nullptr;  // device scratch for the handle all-gather
    ~NcclWorld() override {
        cudaDeviceSynchronize();
        for (int sl = 0; sl < kPushSlots; ++sl) release_region(sl);
        for (int p = 0; p < size; ++p)
            if (p != rank && pc.ctrl_peer[p]) cudaIpcCloseMemHandle(pc.ctrl_peer[p]);
        if (pc.ctrl_local) cudaFree(pc.ctrl_local);
        if (pc.seq_scratch) cudaFree(pc.seq_scratch);
        if (hs_dev) cudaFree(hs_dev);
        if (comm) nccl::ncclCommDestroy(comm);
    }
    const char* kind() const override { return push_ok ? "nvlink-push(ipc)+nccl" : "nccl"; }
    bool can_push() const override { return push_ok; }
    char* shared_local(int slot) override { return pc.region_local[slot]; }
    char* shared_peer(int slot, int p) override { return pc.region_peer[slot][p]; }
    int push_open(int slot, const unsigned char* to, cudaStream_t s) override { return pc.open(slot, to, s); }
    int push_post(int slot, const unsigned char* to, cudaStream_t s) override { return pc.post(slot, to, s); }
    int push_post_ce(int slot, const unsigned char* to, cudaStream_t s) override { return pc.post_ce(slot, to, s); }
    int push_wait(int slot, const unsigned char* from, cudaStream_t s) override { return pc.wait(slot, from, s); }
    int push_close(int slot, cudaStream_t s) override { return pc.close(slot, s); }
    bool push_chunks() const override { return true; }
    void push_next(int slot) override { pc.next(slot); }

    struct HandleMsg { cudaIpcMemHandle_t h; int ok; int pad[15]; };  // 128 bytes
    // all-gather one handle (+ ok flag) per rank through NCCL; returns 1 on a local failure
    int gather_handles(const HandleMsg& mine, std::vector<HandleMsg>& all) {
        static_assert(sizeof(HandleMsg) == 128, "HandleMsg is 128 bytes");
        if (!hs_dev && check_cuda(cudaMalloc(&hs_dev, sizeof(HandleMsg) * (size + 1)), "ipc scratch")) return 1;
        all.resize(size);
        if (check_cuda(cudaMemcpy(hs_dev, &mine, sizeof(HandleMsg), cudaMemcpyHostToDevice), "ipc h2d")) return 1;
        if (check_nccl(nccl::ncclAllGather(hs_dev, hs_dev + sizeof(HandleMsg), sizeof(HandleMsg), ncclChar, comm, 0), "ncclAllGather(handles)")) return 1;
        if (check_cuda(cudaStreamSynchronize(0), "ipc sync")) return 1;
        return check_cuda(cudaMemcpy(all.data(), hs_dev + sizeof(HandleMsg), sizeof(HandleMsg) * size, cudaMemcpyDeviceToHost), "ipc d2h");
    }
    void release_region(int sl) {
        for (int p = 0; p < size; ++p)
            if (p != rank && pc.region_peer[sl][p]) { cudaIpcCloseMemHandle(pc.region_peer[sl][p]); pc.region_peer[sl][p] = nullptr; }
        if (pc.region_local[sl]) cudaFree(pc.region_local[sl]);
        pc.region_local[sl] = nullptr;
        pc.region_cap[sl] = 0;
    }
    // allocate `bytes`, exchange IPC handles, map every peer's allocation; collective
    int map_all(size_t bytes, char** local, char** peers /*[size]*/) {
        HandleMsg mine;
        memset(&mine, 0, sizeof(mine));
        void* p = nullptr;
        mine.ok = (dev_malloc(&p, bytes) == cudaSuccess) && (cudaMemset(p, 0, bytes) == cudaSuccess) &&
                  (cudaIpcGetMemHandle(&mine.h, p) == cudaSuccess) && (device_sync() == cudaSuccess);
        std::vector<HandleMsg> all;
        if (gather_handles(mine, all)) return 1;
        bool ok = true;
        for (int q = 0; q < size; ++q) ok = ok && all[q].ok;
        if (ok) {
            for (int q = 0; q < size && ok; ++q) {
                if (q == rank) { peers[q] = (char*)p; continue; }
                void* m = nullptr;
                if (cudaIpcOpenMemHandle(&m, all[q].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = false; break; }
                peers[q] = (char*)m;
            }
        }
        // agree on the outcome (a rank that failed to map must not leave the others pushing)
        HandleMsg st;
        memset(&st, 0, sizeof(st));
        st.ok = ok ? 1 : 0;
        if (gather_handles(st, all)) return 1;
        bool all_ok = true;
        for (int q = 0; q < size; ++q) all_ok = all_ok && all[q].ok;
        if (!all_ok) {
            cudaGetLastError();
            for (int q = 0; q < size; ++q) {
                if (q != rank && peers[q]) cudaIpcCloseMemHandle(peers[q]);
                peers[q] = nullptr;
            }
            if (p) cudaFree(p);
            *local = nullptr;
            return set_error("CUDA IPC peer mapping failed on at least one rank");
        }
        *local = (char*)p;
        return 0;
    }
    int init_push() {
        const char* e = getenv("HYD_TRANSPORT");
        if (e && strcmp(e, "nccl") == 0) return 0;
        if (size < 2 || size > kPushMaxRanks) return 0;
        pc.rank = rank;
        pc.size = size;
        char* local = nullptr;
        char* peers[kPushMaxRanks] = {};
        if (map_all(PushCore::ctrl_bytes(), &local, peers)) return 0;  // fall back to staged NCCL send/recv
        pc.ctrl_local = (uint32_t*)local;
        for (int p = 0; p < size; ++p) pc.ctrl_peer[p] = (uint32_t*)peers[p];
        void* sq = nullptr;
        if (cudaMalloc(&sq, sizeof(uint32_t) * kPushSlots) == cudaSuccess) pc.seq_scratch = (uint32_t*)sq;
        push_ok = true;
        return 0;
    }
    int shared_reserve(int slot, size_t bytes) override {
        if (bytes <= pc.region_cap[slot]) return 0;
        // every rank takes this branch together (same `bytes`, same history): drain all work that may still
        // touch the old mappings, everywhere, before they go away
        if (check_cuda(device_sync(), "shared_reserve sync")) return 1;
        if (barrier()) return 1;
        release_region(slot);
        const size_t cap = push_grow(bytes);
        char* local = nullptr;
        char* peers[kPushMaxRanks] = {};
        if (map_all(cap, &local, peers)) return 1;
        pc.region_local[slot] = local;
        pc.region_cap[slot] = cap;
        for (int p = 0; p < size; ++p) pc.region_peer[slot][p] = peers[p];
        return barrier();
    }
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
        return check_cuda(stream_sync(0), "barrier sync");
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
    w->init_push();
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
    // push transport: every rank-thread's control block and receive regions (plain device pointers: one process)
    std::vector<uint32_t*> ctrl;
    std::vector<char*> region[kPushSlots];
    // in-process rounds are ordered with CUDA events + host barriers instead of flag waits (see LocalWorld)
    std::vector<cudaEvent_t> posted[kPushSlots], consumed[kPushSlots];
};

LocalShared* local_shared_create(int size) {
    LocalShared* sh = new LocalShared;
    sh->size = size;
    pthread_barrier_init(&sh->bar, nullptr, size);
    sh->slots.resize(size);
    sh->ctrl.assign(size, nullptr);
    for (auto& r : sh->region) r.assign(size, nullptr);
    for (int sl = 0; sl < kPushSlots; ++sl) { sh->posted[sl].assign(size, nullptr); sh->consumed[sl].assign(size, nullptr); }
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
    for (int sl = 0; sl < kPushSlots; ++sl) {
        for (cudaEvent_t e : sh->posted[sl]) if (e) cudaEventDestroy(e);
        for (cudaEvent_t e : sh->consumed[sl]) if (e) cudaEventDestroy(e);
    }
    pthread_barrier_destroy(&sh->bar);
    delete sh;
}

struct LocalWorld : World {
    LocalShared* sh = nullptr;
    PushCore pc;
    bool push_ok = false, ctrl_linked = false;
    ~LocalWorld() override {
        cudaDeviceSynchronize();
        for (int sl = 0; sl < kPushSlots; ++sl) if (pc.region_local[sl]) cudaFree(pc.region_local[sl]);
        if (pc.ctrl_local) cudaFree(pc.ctrl_local);
    }
    const char* kind() const override { return push_ok ? "local-push" : "local"; }
    int barrier() override { pthread_barrier_wait(&sh->bar); return 0; }
    bool can_push() const override { return push_ok; }
    char* shared_local(int slot) override { return pc.region_local[slot]; }
    char* shared_peer(int slot, int p) override { return pc.region_peer[slot][p]; }
    // Rank-threads of ONE process share the GPU's hardware queues: a stream parked in a flag wait could hold up another
    // rank's stream aliased onto the same queue.  So in-process rounds use what one process can use instead -- CUDA
    // events recorded/waited across the ranks' streams, with a host barrier making sure the record was issued before
    // the wait.  Same peer stores, same region layout, same call sequence as the NVLink flag protocol (push.cu).
    cudaEvent_t ev(std::vector<cudaEvent_t>& v, int r) {
        if (!v[r]) cudaEventCreateWithFlags(&v[r], cudaEventDisableTiming);
        return v[r];
    }
    int push_open(int slot, const unsigned char* to, cudaStream_t s) override {
        pthread_barrier_wait(&sh->bar);  // every rank has issued the previous round's close
        for (int p = 0; p < size; ++p)
            if (p != rank && to[p] && sh->consumed[slot][p] &&
                check_cuda(cudaStreamWaitEvent(s, sh->consumed[slot][p], 0), "push_open(local)")) return 1;
        return 0;
    }
    int push_post(int slot, const unsigned char*, cudaStream_t s) override {
        return check_cuda(cudaEventRecord(ev(sh->posted[slot], rank), s), "push_post(local)");
    }
    int push_wait(int slot, const unsigned char* from, cudaStream_t s) override {
        pthread_barrier_wait(&sh->bar);  // every rank has issued its post
        for (int q = 0; q < size; ++q)
            if (q != rank && from[q] && check_cuda(cudaStreamWaitEvent(s, sh->posted[slot][q], 0), "push_wait(local)")) return 1;
        return 0;
    }
    int push_close(int slot, cudaStream_t s) override {
        return check_cuda(cudaEventRecord(ev(sh->consumed[slot], rank), s), "push_close(local)");
    }
    int init_push() {
        const char* e = getenv("HYD_TRANSPORT");
        if ((e && strcmp(e, "staged") == 0) || size < 2 || size > kPushMaxRanks) return 0;
        pc.rank = rank;
        pc.size = size;
        push_ok = true;
        return 0;
    }
    int shared_reserve(int slot, size_t bytes) override {
        if (ctrl_linked && bytes <= pc.region_cap[slot]) return 0;
        // same decision on every rank-thread (same `bytes`, same history)
        device_sync();
        pthread_barrier_wait(&sh->bar);
        int rc = 0;
        if (bytes > pc.region_cap[slot]) {
            if (pc.region_local[slot]) cudaFree(pc.region_local[slot]);
            const size_t cap = push_grow(bytes);
            void* p = nullptr;
            rc = check_cuda(dev_malloc(&p, cap), "push region");
            pc.region_local[slot] = (char*)p;
            pc.region_cap[slot] = rc ? 0 : cap;
            sh->region[slot][rank] = (char*)p;
        }
        pthread_barrier_wait(&sh->bar);
        for (int q = 0; q < size; ++q) {
            pc.region_peer[slot][q] = sh->region[slot][q];
            pc.ctrl_peer[q] = sh->ctrl[q];
        }
        ctrl_linked = true;
        pthread_barrier_wait(&sh->bar);
        return rc;
    }
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
    w->init_push();
    return w;
}

}  // namespace hyd

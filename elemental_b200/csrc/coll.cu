// Seam S2: the typed collectives the reference's copy:: / axpy_contract:: primitives call on the grid's
// sub-communicators, on DEVICE buffers with stream semantics:
//   mpi::AllGather / AllToAll / ReduceScatter / SendRecv / Broadcast <T, Device::GPU>
//     (src/core/imports/mpi/{AllGather,AllToAll,ReduceScatter,SendRecv,Broadcast}.hpp; include/El/core/imports/mpi.hpp)
//   over Grid::MCComm() (the ranks of my grid COLUMN, indexed by grid row), MRComm() (my grid ROW), VCComm(), VRComm()
//     (include/El/core/Grid.hpp:31-54, src/core/Grid.cpp:133-150).
//
// The reference maps AllGather / ReduceScatter / Broadcast to Aluminum -> NCCL and AllToAll / SendRecv to MPI on device
// pointers (include/El/core/imports/aluminum.hpp:99-116).  Here every one of them is the world's single primitive -- a
// variable-size exchange among grid ranks (comm.h), i.e. one NCCL group of ncclSend/ncclRecv over NVSwitch, every
// member one hop away -- plus a device-to-device copy of the caller's own part.  ReduceScatter receives the partial
// sums into a per-grid workspace and adds them in ASCENDING communicator rank (the order the oracle and the engine's
// reduce_into use), so its result is reproducible bit for bit.
//
// Like every collective of the reference these are collective over the GRID: every rank of the grid calls, each inside
// its own sub-communicator, in the same order.  The distributed engine (redist.cu) does not use them -- it plans one
// exchange per redistribution -- they serve callers that keep the reference's pack -> collective -> unpack structure.
#include "../../include/hydrogen_b200.h"
#include "grid.h"
#include "hyd_internal.h"
#include <algorithm>
#include <vector>

struct hyd_grid { hyd::Grid* g; };

namespace hyd {
namespace {

struct SubComm {
    int size = 0, me = 0;
    int member[kPushMaxRanks];  // world rank of communicator rank k
};

int sub_comm(const Grid& g, int comm, SubComm& c, const char* who) {
    const GridShape& sh = g.shape;
    switch (comm) {
        case HYD_COMM_MC: c.size = sh.r; c.me = g.mc; break;
        case HYD_COMM_MR: c.size = sh.c; c.me = g.mr; break;
        case HYD_COMM_VC: c.size = sh.size(); c.me = g.vc(); break;
        case HYD_COMM_VR: c.size = sh.size(); c.me = g.vr(); break;
        default: return set_error("%s: unknown communicator %d", who, comm);
    }
    if (c.size > kPushMaxRanks) return set_error("%s: communicator larger than %d ranks", who, kPushMaxRanks);
    for (int k = 0; k < c.size; ++k) {
        switch (comm) {
            case HYD_COMM_MC: c.member[k] = sh.rank_of(k, g.mr); break;
            case HYD_COMM_MR: c.member[k] = sh.rank_of(g.mc, k); break;
            case HYD_COMM_VC: c.member[k] = sh.rank_of(k % sh.r, k / sh.r); break;   // VC rank = row + r*col
            default:          c.member[k] = sh.rank_of(k / sh.c, k % sh.c); break;   // VR rank = col + c*row
        }
    }
    return 0;
}

struct Counts {
    std::vector<i64> sc, so, rc, ro;
    explicit Counts(int P) : sc(P, 0), so(P, 0), rc(P, 0), ro(P, 0) {}
};

int self_copy(void* dst, const void* src, i64 bytes, cudaStream_t s, const char* who) {
    if (bytes <= 0 || dst == src) return 0;
    return check_cuda(cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDeviceToDevice, s), who);
}

// out[i] = ((p_0[i] + p_1[i]) + p_2[i]) + ...   ascending communicator rank
constexpr int kMaxSum = kPushMaxRanks;
struct SumArgs {
    const void* part[kMaxSum];
    int n;
};
template <typename T>
__global__ void __launch_bounds__(256) sum_parts_kernel(T* __restrict__ out, i64 count, const __grid_constant__ SumArgs a) {
    for (i64 i = (i64)blockIdx.x * 256 + threadIdx.x; i < count; i += (i64)gridDim.x * 256) {
        T acc = static_cast<const T*>(a.part[0])[i];
        for (int q = 1; q < a.n; ++q) acc = acc + static_cast<const T*>(a.part[q])[i];
        out[i] = acc;
    }
}

template <typename T>
int reduce_scatter(Grid& g, int comm, const T* sbuf, T* rbuf, i64 count, cudaStream_t s) {
    const char* who = "hyd_reducescatter";
    SubComm c;
    if (sub_comm(g, comm, c, who)) return 1;
    if (count < 0) return set_error("%s: negative count", who);
    if (count == 0) return 0;
    if (!sbuf || !rbuf) return set_error("%s: null buffer", who);
    const i64 bytes = count * (i64)sizeof(T);
    if (c.size == 1) return self_copy(rbuf, sbuf, bytes, s, who);
    if (ws_reserve(g.coll_ws, (size_t)bytes * c.size)) return 1;
    Counts n(g.world->size);
    for (int k = 0; k < c.size; ++k) {
        if (k == c.me) continue;
        const int p = c.member[k];
        n.sc[p] = bytes; n.so[p] = k * bytes;   // portion k of my partial sums belongs to member k
        n.rc[p] = bytes; n.ro[p] = k * bytes;   // member k's partial of MY portion
    }
    if (g.world->exchange(sbuf, n.sc.data(), n.so.data(), g.coll_ws.buf, n.rc.data(), n.ro.data(), s)) return 1;
    SumArgs a;
    a.n = c.size;
    for (int k = 0; k < c.size; ++k)
        a.part[k] = (k == c.me) ? (const void*)(sbuf + (i64)c.me * count) : (const void*)(g.coll_ws.buf + k * bytes);
    const unsigned grid = (unsigned)std::max<i64>(1, std::min<i64>((count + 255) / 256, (i64)num_sms() * 8));
    sum_parts_kernel<T><<<grid, 256, 0, s>>>(rbuf, count, a);
    ++g_launches;
    return check_cuda(cudaGetLastError(), who);
}

}  // namespace
}  // namespace hyd

using namespace hyd;

extern "C" {

int hyd_comm_size(const hyd_grid* g, int comm) {
    SubComm c;
    return sub_comm(*g->g, comm, c, "hyd_comm_size") ? -1 : c.size;
}
int hyd_comm_rank(const hyd_grid* g, int comm) {
    SubComm c;
    return sub_comm(*g->g, comm, c, "hyd_comm_rank") ? -1 : c.me;
}
int hyd_comm_world_rank(const hyd_grid* g, int comm, int k) {
    SubComm c;
    if (sub_comm(*g->g, comm, c, "hyd_comm_world_rank")) return -1;
    if (k < 0 || k >= c.size) { set_error("hyd_comm_world_rank: rank out of range"); return -1; }
    return c.member[k];
}

int hyd_allgather(hyd_grid* g, int comm, const void* sbuf, int64_t bytes, void* rbuf, hyd_stream_t stream) {
    const char* who = "hyd_allgather";
    Grid& G = *g->g;
    SubComm c;
    if (sub_comm(G, comm, c, who)) return 1;
    if (bytes < 0) return set_error("%s: negative count", who);
    if (bytes == 0) return 0;
    if (!sbuf || !rbuf) return set_error("%s: null buffer", who);
    cudaStream_t s = (cudaStream_t)stream;
    Counts n(G.world->size);
    for (int k = 0; k < c.size; ++k) {
        if (k == c.me) continue;
        const int p = c.member[k];
        n.sc[p] = bytes; n.so[p] = 0;
        n.rc[p] = bytes; n.ro[p] = k * bytes;
    }
    if (self_copy((char*)rbuf + c.me * bytes, sbuf, bytes, s, who)) return 1;
    return G.world->exchange(sbuf, n.sc.data(), n.so.data(), rbuf, n.rc.data(), n.ro.data(), s);
}

int hyd_alltoall(hyd_grid* g, int comm, const void* sbuf, int64_t bytes, void* rbuf, hyd_stream_t stream) {
    const char* who = "hyd_alltoall";
    Grid& G = *g->g;
    SubComm c;
    if (sub_comm(G, comm, c, who)) return 1;
    if (bytes < 0) return set_error("%s: negative count", who);
    if (bytes == 0) return 0;
    if (!sbuf || !rbuf) return set_error("%s: null buffer", who);
    if (sbuf == rbuf && c.size > 1) return set_error("%s: in-place exchange is not supported", who);
    cudaStream_t s = (cudaStream_t)stream;
    Counts n(G.world->size);
    for (int k = 0; k < c.size; ++k) {
        if (k == c.me) continue;
        const int p = c.member[k];
        n.sc[p] = bytes; n.so[p] = k * bytes;
        n.rc[p] = bytes; n.ro[p] = k * bytes;
    }
    if (self_copy((char*)rbuf + c.me * bytes, (const char*)sbuf + c.me * bytes, bytes, s, who)) return 1;
    return G.world->exchange(sbuf, n.sc.data(), n.so.data(), rbuf, n.rc.data(), n.ro.data(), s);
}

int hyd_sendrecv(hyd_grid* g, int comm, const void* sbuf, int64_t sbytes, int to, void* rbuf, int64_t rbytes, int from,
                 hyd_stream_t stream) {
    const char* who = "hyd_sendrecv";
    Grid& G = *g->g;
    SubComm c;
    if (sub_comm(G, comm, c, who)) return 1;
    if (sbytes < 0 || rbytes < 0) return set_error("%s: negative count", who);
    if (to < 0 || to >= c.size || from < 0 || from >= c.size) return set_error("%s: rank out of range", who);
    if ((to == c.me) != (from == c.me)) return set_error("%s: a rank that sends to itself must receive from itself", who);
    if ((sbytes > 0 && !sbuf) || (rbytes > 0 && !rbuf)) return set_error("%s: null buffer", who);
    cudaStream_t s = (cudaStream_t)stream;
    Counts n(G.world->size);
    if (to == c.me) {
        if (sbytes != rbytes) return set_error("%s: self exchange with different counts", who);
        if (self_copy(rbuf, sbuf, sbytes, s, who)) return 1;
    } else {
        n.sc[c.member[to]] = sbytes;
        n.rc[c.member[from]] = rbytes;
    }
    // every rank enters the exchange, also with nothing to move (collective over the grid)
    return G.world->exchange(sbuf, n.sc.data(), n.so.data(), rbuf, n.rc.data(), n.ro.data(), s);
}

int hyd_bcast(hyd_grid* g, int comm, void* buf, int64_t bytes, int root, hyd_stream_t stream) {
    const char* who = "hyd_bcast";
    Grid& G = *g->g;
    SubComm c;
    if (sub_comm(G, comm, c, who)) return 1;
    if (bytes < 0) return set_error("%s: negative count", who);
    if (root < 0 || root >= c.size) return set_error("%s: root out of range", who);
    if (bytes == 0) return 0;
    if (!buf) return set_error("%s: null buffer", who);
    Counts n(G.world->size);
    if (c.me == root) {
        for (int k = 0; k < c.size; ++k)
            if (k != root) n.sc[c.member[k]] = bytes;
    } else {
        n.rc[c.member[root]] = bytes;
    }
    return G.world->exchange(buf, n.sc.data(), n.so.data(), buf, n.rc.data(), n.ro.data(), (cudaStream_t)stream);
}

int hyd_reducescatter_f64(hyd_grid* g, int comm, const double* sbuf, double* rbuf, int64_t count, hyd_stream_t stream) {
    return reduce_scatter<double>(*g->g, comm, sbuf, rbuf, count, (cudaStream_t)stream);
}
int hyd_reducescatter_f32(hyd_grid* g, int comm, const float* sbuf, float* rbuf, int64_t count, hyd_stream_t stream) {
    return reduce_scatter<float>(*g->g, comm, sbuf, rbuf, count, (cudaStream_t)stream);
}

}  // extern "C"

// Communication backends behind the redistribution engine.
//
// The reference routes every redistribution through El::mpi::{AllGather,AllToAll,
// ReduceScatter,SendRecv,Broadcast} on the grid's MC/MR/VC/VR communicators
// (include/El/core/imports/mpi.hpp; src/core/imports/mpi/*.hpp; Aluminum -> NCCL for some,
// MPI for AllToAll/SendRecv, include/El/core/imports/aluminum.hpp:99-116).
// On NVSwitch every peer is one hop at full bandwidth, so this build needs exactly one
// primitive: a variable-size all-to-all ("exchange") among the ranks of the grid, issued
// as one NCCL group of ncclSend/ncclRecv.  Sub-communicators are not needed.
//
// Two implementations:
//   * NcclWorld  : one process per GPU, NCCL over NVLink 5 (the product path);
//   * LocalWorld : P rank-threads of ONE process sharing ONE GPU, exchange done with
//                  stream-ordered device copies (no kernel ever waits on another kernel).
//                  It exists so the complete multi-rank engine can be parity-tested on a
//                  single-GPU box; it is still all-GPU, not a CPU fallback.
#pragma once
#include <cuda_runtime.h>
#include "dist.h"

namespace hyd {

struct World {
    int rank = 0, size = 1;
    virtual ~World() {}
    // Counts/offsets are in BYTES, indexed by world rank.  Entries for `rank` itself are ignored
    // (the engine moves its own part without the network).  Asynchronous on `s`.
    virtual int exchange(const void* send, const i64* scount, const i64* soff, void* recv,
                         const i64* rcount, const i64* roff, cudaStream_t s) = 0;
    virtual int barrier() = 0;  // host-side
    virtual const char* kind() const = 0;

    // ---- push transport over peer-mapped memory (NVLink P2P stores; see push.cu) -------------------
    // When can_push() is true the redistribution engine does not stage-and-send: the pack kernel of rank q
    // stores each message straight into rank p's receive region through a peer-mapped pointer (pack + NVLink
    // transfer fused in one kernel), and rounds are ordered by 32-bit flags in peer-mapped control blocks,
    // waited on with stream memory operations (no kernel ever spins).  A "slot" is one receive region +
    // flag set; all ranks use a slot in the same order (every exchange is collective over the world).
    virtual bool can_push() const { return false; }
    // collective: every rank's region `slot` holds >= bytes afterwards (same `bytes` on every rank)
    virtual int shared_reserve(int /*slot*/, size_t /*bytes*/) { return 1; }
    virtual char* shared_local(int /*slot*/) { return nullptr; }
    virtual char* shared_peer(int /*slot*/, int /*p*/) { return nullptr; }
    // to[p] != 0: I store into p's region this round; from[q] != 0: q stores into mine
    virtual int push_open(int /*slot*/, const unsigned char* /*to*/, cudaStream_t) { return 1; }   // wait: receivers consumed previous round
    virtual int push_post(int /*slot*/, const unsigned char* /*to*/, cudaStream_t) { return 1; }   // signal: my stores have landed
    // same signal issued without any kernel (stream memory op + 4-byte copy-engine writes): for rounds whose data moved
    // with copy engines while a persistent GEMM holds every SM
    virtual int push_post_ce(int slot, const unsigned char* to, cudaStream_t s) { return push_post(slot, to, s); }
    virtual int push_wait(int /*slot*/, const unsigned char* /*from*/, cudaStream_t) { return 1; } // wait for my senders
    virtual int push_close(int /*slot*/, cudaStream_t) { return 1; }                                // broadcast: round consumed
    // chunked rounds (see push.h): true if push_next() is supported; push_next starts the next chunk of the open round
    virtual bool push_chunks() const { return false; }
    virtual void push_next(int /*slot*/) {}
};

constexpr int kPushSlots = 8;   // 0-3: exchange workspaces, 4-7: SUMMA panels (A/B x double buffer)
constexpr int kPushMaxRanks = 64;

World* make_nccl_world(const void* unique_id_128B, int rank, int size);
int nccl_unique_id(void* out_128B);

struct LocalShared;
LocalShared* local_shared_create(int size);
void local_shared_destroy(LocalShared*);
World* make_local_world(LocalShared* sh, int rank);

}  // namespace hyd

// Push transport core: per-rank control block (round flags / acks), receive regions, and the stream-ordered
// round protocol shared by the NCCL-bootstrapped IPC world (one process per GPU, NVLink peer stores) and the
// in-process world (P rank-threads on one GPU).  See comm.h for the interface contract.
//
// Round r on a slot, on every rank, in stream order:
//   open : wait  ack[p] >= r-1  for every p I will store into      (p has consumed the previous round)
//   pack : kernel stores my messages into the peers' regions through mapped pointers
//   post : one tiny kernel: st.release.sys  flag[me] := r  on every p I stored into
//   wait : wait  flag[q] >= r  for every q that stores into me     (cuStreamWaitValue32: no SM is held)
//   unpack kernel reads my region
//   close: one tiny kernel: ack[me] := r on every peer
// A round may be split into chunks (disjoint parts of the region): pack, post, next, pack, post, ... with the
// receiver waiting and unpacking chunk by chunk; every post carries a fresh flag value, only open/close touch acks.
#pragma once
#include "comm.h"
#include <stdint.h>

namespace hyd {

struct PushCore {
    int rank = 0, size = 1;
    uint32_t* ctrl_local = nullptr;                       // [2][kPushSlots][kPushMaxRanks] (flags, acks), zero-initialised
    uint32_t* ctrl_peer[kPushMaxRanks] = {};              // peers' control blocks as mapped here (self = ctrl_local)
    char* region_local[kPushSlots] = {};
    size_t region_cap[kPushSlots] = {};
    char* region_peer[kPushSlots][kPushMaxRanks] = {};    // peers' regions as mapped here (self = region_local)
    uint32_t round[kPushSlots] = {};
    uint32_t closed[kPushSlots] = {};                     // round number of the last close (what the next open waits for)

    static size_t ctrl_bytes() { return sizeof(uint32_t) * 2 * kPushSlots * kPushMaxRanks; }
    static size_t flag_index(int slot, int r) { return (size_t)slot * kPushMaxRanks + r; }
    static size_t ack_index(int slot, int r) { return (size_t)(kPushSlots + slot) * kPushMaxRanks + r; }

    int open(int slot, const unsigned char* to, cudaStream_t s);
    void next(int slot) { ++round[slot]; }             // next chunk of an open round: new flag value, same region, no ack wait
    int post(int slot, const unsigned char* to, cudaStream_t s);
    int post_ce(int slot, const unsigned char* to, cudaStream_t s);
    uint32_t* seq_scratch = nullptr;  // [kPushSlots] device words: round numbers staged for the 4-byte flag copies
    int wait(int slot, const unsigned char* from, cudaStream_t s);
    int close(int slot, cudaStream_t s);
};

size_t push_grow(size_t bytes);  // capacity policy (identical on every rank)

}  // namespace hyd

// Push transport core (see push.h / comm.h).
#include "push.h"
#include "hyd_internal.h"
#include <algorithm>

namespace hyd {

namespace {

typedef CUresult (*wait32_fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);

wait32_fn stream_write32() {
    static wait32_fn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuStreamWriteValue32", &p, cudaEnableDefault, &q) == cudaSuccess) fn = reinterpret_cast<wait32_fn>(p);
    }
    return fn;
}

wait32_fn stream_wait32() {
    static wait32_fn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &p, cudaEnableDefault, &q) == cudaSuccess) fn = reinterpret_cast<wait32_fn>(p);
    }
    return fn;
}

struct SignalArgs {
    uint32_t* addr[kPushMaxRanks];
    int n;
    uint32_t value;
};

// thread i publishes `value` at addr[i] with system-scope release semantics: every store of the preceding
// kernels of this stream (the pack kernel's peer stores, or the unpack kernel's reads) is ordered before it.
__global__ void push_signal_kernel(const __grid_constant__ SignalArgs a) {
    const int i = threadIdx.x;
    if (i < a.n) {
        __threadfence_system();
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(a.addr[i]), "r"(a.value) : "memory");
    }
}

int wait_geq(uint32_t* addr, uint32_t value, cudaStream_t s, const char* what) {
    wait32_fn fn = stream_wait32();
    if (!fn) return set_error("push transport: cuStreamWaitValue32 is not available");
    const CUresult r = fn((CUstream)s, (CUdeviceptr)(uintptr_t)addr, value, CU_STREAM_WAIT_VALUE_GEQ);
    if (r != CUDA_SUCCESS) return set_error("%s: cuStreamWaitValue32 failed (%d)", what, (int)r);
    return 0;
}

}  // namespace

size_t push_grow(size_t bytes) {
    size_t cap = std::max(bytes + bytes / 4, (size_t)1 << 20);
    return (cap + 4095) & ~(size_t)4095;
}

int PushCore::open(int slot, const unsigned char* to, cudaStream_t s) {
    ++round[slot];
    const uint32_t prev = closed[slot];
    if (prev == 0) return 0;
    for (int p = 0; p < size; ++p)
        if (p != rank && to[p] && wait_geq(ctrl_local + ack_index(slot, p), prev, s, "push_open")) return 1;
    return 0;
}

int PushCore::post(int slot, const unsigned char* to, cudaStream_t s) {
    SignalArgs a;
    a.n = 0;
    a.value = round[slot];
    for (int p = 0; p < size; ++p)
        if (p != rank && to[p]) a.addr[a.n++] = ctrl_peer[p] + flag_index(slot, rank);
    if (a.n == 0) return 0;
    push_signal_kernel<<<1, kPushMaxRanks, 0, s>>>(a);
    return check_cuda(cudaGetLastError(), "push_signal_kernel(post)");
}

int PushCore::post_ce(int slot, const unsigned char* to, cudaStream_t s) {
    wait32_fn wr = stream_write32();
    if (!wr || !seq_scratch) return post(slot, to, s);
    bool any = false;
    for (int p = 0; p < size; ++p) any = any || (p != rank && to[p]);
    if (!any) return 0;
    // stage the round number in local memory (stream memory op), then copy it onto every receiver's flag word with the
    // copy engine: both are ordered after the preceding copies of this stream and neither needs an SM
    if (wr((CUstream)s, (CUdeviceptr)(uintptr_t)(seq_scratch + slot), round[slot], 0) != CUDA_SUCCESS)
        return set_error("push_post_ce: cuStreamWriteValue32 failed");
    for (int p = 0; p < size; ++p)
        if (p != rank && to[p] &&
            check_cuda(cudaMemcpyAsync(ctrl_peer[p] + flag_index(slot, rank), seq_scratch + slot, 4, cudaMemcpyDeviceToDevice, s),
                       "push_post_ce: flag copy"))
            return 1;
    return 0;
}

int PushCore::wait(int slot, const unsigned char* from, cudaStream_t s) {
    for (int q = 0; q < size; ++q)
        if (q != rank && from[q] && wait_geq(ctrl_local + flag_index(slot, q), round[slot], s, "push_wait")) return 1;
    return 0;
}

int PushCore::close(int slot, cudaStream_t s) {
    SignalArgs a;
    a.n = 0;
    a.value = closed[slot] = round[slot];
    for (int p = 0; p < size; ++p)
        if (p != rank) a.addr[a.n++] = ctrl_peer[p] + ack_index(slot, rank);
    if (a.n == 0) return 0;
    push_signal_kernel<<<1, kPushMaxRanks, 0, s>>>(a);
    return check_cuda(cudaGetLastError(), "push_signal_kernel(close)");
}

}  // namespace hyd

// Error state, device queries and the driver entry point used for TMA descriptors.
#include "hyd_internal.h"
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

namespace hyd {

static thread_local char g_err[1024] = "";

int set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return 1;
}
const char* last_error() { return g_err; }

int check_cuda(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return 0;
    return set_error("%s: CUDA error %d (%s)", what, (int)e, cudaGetErrorString(e));
}

int num_sms() {
    static int cached[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (cached[dev] == 0) {
        int n = 0;
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        cached[dev] = n > 0 ? n : 148;
    }
    return cached[dev];
}

tensor_map_encode_fn tensor_map_encode() {
    static tensor_map_encode_fn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
        fn = reinterpret_cast<tensor_map_encode_fn>(p);
    }
    return fn;
}

}  // namespace hyd

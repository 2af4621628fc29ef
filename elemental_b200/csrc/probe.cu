// Pipe-peak probes: register-resident FP64 DMMA / DFMA and FP32 FFMA loops.
// Used by bench.py to measure the roofline denominator live on the box
// (MEASURED_PEAKS.json carries only HBM and bf16 numbers).
#include "ptx.cuh"
#include "hyd_internal.h"

namespace hyd {

template <int KIND>
__global__ void __launch_bounds__(256) probe_kernel(int iters, double* sink) {
    const int lane = threadIdx.x & 31;
    if (KIND == 0) {  // DFMA: 16 independent chains per thread
        double a = 1.0 + lane * 1e-9, b = 1.0 - lane * 1e-9, c[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) c[i] = i;
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 16; ++i) c[i] = fma(a, b, c[i]);
        }
        double s = 0;
#pragma unroll
        for (int i = 0; i < 16; ++i) s += c[i];
        if (s == 12345.678) sink[0] = s;
    } else if (KIND == 1) {  // DMMA m8n8k4: 8 independent accumulators
        double a = 1.0 + lane * 1e-9, b = 1.0 - lane * 1e-9, d[8][2];
#pragma unroll
        for (int i = 0; i < 8; ++i) d[i][0] = d[i][1] = i;
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 8; ++i) dmma_884(d[i], a, b);
        }
        double s = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) s += d[i][0] + d[i][1];
        if (s == 12345.678) sink[0] = s;
    } else if (KIND == 2) {  // DMMA m16n8k16: 8 independent accumulators
        double a[8], b[4], d[8][4];
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = 1.0 + (lane + i) * 1e-9;
#pragma unroll
        for (int i = 0; i < 4; ++i) b[i] = 1.0 - (lane + i) * 1e-9;
#pragma unroll
        for (int i = 0; i < 8; ++i) d[i][0] = d[i][1] = d[i][2] = d[i][3] = i;
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 8; ++i) dmma_16816(d[i], a, b);
        }
        double s = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) s += d[i][0] + d[i][1] + d[i][2] + d[i][3];
        if (s == 12345.678) sink[0] = s;
    } else {  // FFMA: 32 independent chains
        float a = 1.0f + lane * 1e-6f, b = 1.0f - lane * 1e-6f, c[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) c[i] = i;
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 32; ++i) c[i] = fmaf(a, b, c[i]);
        }
        float s = 0;
#pragma unroll
        for (int i = 0; i < 32; ++i) s += c[i];
        if (s == 12345.678f) sink[0] = s;
    }
}

int probe_tf32_peak(int iters, double* tflops);   // lgemm_f32_tc.cu

// Returns achieved TFLOP/s for the probe `kind` (0 DFMA, 1 DMMA m8n8k4, 2 DMMA m16n8k16, 3 FFMA, 4 tcgen05 dense TF32).
int probe_peak(int kind, int iters, int ctas_per_sm, double* tflops) {
    if (kind == 4) return probe_tf32_peak(iters, tflops);
    double* sink = nullptr;
    if (check_cuda(cudaMalloc(&sink, 64), "probe malloc")) return 1;
    const int grid = num_sms() * ctas_per_sm, block = 256;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double flops_per_thread_iter = 0;
    for (int rep = 0; rep < 2; ++rep) {  // rep 0 = warm-up
        cudaEventRecord(e0);
        switch (kind) {
            case 0: probe_kernel<0><<<grid, block>>>(iters, sink); flops_per_thread_iter = 16 * 2; break;
            case 1: probe_kernel<1><<<grid, block>>>(iters, sink); flops_per_thread_iter = 8 * 512.0 / 32; break;
            case 2: probe_kernel<2><<<grid, block>>>(iters, sink); flops_per_thread_iter = 8 * 4096.0 / 32; break;
            default: probe_kernel<3><<<grid, block>>>(iters, sink); flops_per_thread_iter = 32 * 2; break;
        }
        cudaEventRecord(e1);
        if (check_cuda(cudaEventSynchronize(e1), "probe kernel")) return 1;
    }
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    *tflops = flops_per_thread_iter * (double)iters * grid * block / (ms * 1e-3) / 1e12;
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(sink);
    return 0;
}

}  // namespace hyd

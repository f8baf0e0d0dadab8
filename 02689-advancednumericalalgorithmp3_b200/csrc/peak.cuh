// Register-resident FP64 throughput probes: the roofline denominators for the DMMA kernels.
// MEASURED_PEAKS.json carries no FP64 figure, so bench.py measures it live through ldc_fp64_peak().
#pragma once
#include "common.cuh"

namespace ldc {

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

constexpr int PEAK_ITERS = 4096;
constexpr int PEAK_ILP = 8;

// 8 independent accumulator tiles per warp, PEAK_ITERS rounds: 2*8*8*4 flop per DMMA.
__global__ void __launch_bounds__(256) k_peak_dmma(double* out, double a0, double b0) {
    double c[PEAK_ILP][2];
#pragma unroll
    for (int q = 0; q < PEAK_ILP; ++q) c[q][0] = c[q][1] = 0.0;
    const double a = a0 + threadIdx.x * 1e-9, b = b0;
    for (int it = 0; it < PEAK_ITERS; ++it) {
#pragma unroll
        for (int q = 0; q < PEAK_ILP; ++q) dmma884(c[q][0], c[q][1], a, b);
    }
    double s = 0;
#pragma unroll
    for (int q = 0; q < PEAK_ILP; ++q) s += c[q][0] + c[q][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// 16 independent DFMA chains per thread.
__global__ void __launch_bounds__(256) k_peak_dfma(double* out, double a0, double b0) {
    double c[16];
#pragma unroll
    for (int q = 0; q < 16; ++q) c[q] = q;
    const double a = a0 + threadIdx.x * 1e-9, b = b0;
    for (int it = 0; it < PEAK_ITERS; ++it) {
#pragma unroll
        for (int q = 0; q < 16; ++q) c[q] = fma(c[q], a, b);
    }
    double s = 0;
#pragma unroll
    for (int q = 0; q < 16; ++q) s += c[q];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// Both pipes in one instruction stream: 8 DMMA + 16 DFMA per round (do the pipes overlap?).
__global__ void __launch_bounds__(256) k_peak_mixed(double* out, double a0, double b0) {
    double c[PEAK_ILP][2], f[16];
#pragma unroll
    for (int q = 0; q < PEAK_ILP; ++q) c[q][0] = c[q][1] = 0.0;
#pragma unroll
    for (int q = 0; q < 16; ++q) f[q] = q;
    const double a = a0 + threadIdx.x * 1e-9, b = b0;
    for (int it = 0; it < PEAK_ITERS; ++it) {
#pragma unroll
        for (int q = 0; q < PEAK_ILP; ++q) {
            dmma884(c[q][0], c[q][1], a, b);
            f[2 * q] = fma(f[2 * q], a, b);
            f[2 * q + 1] = fma(f[2 * q + 1], a, b);
        }
    }
    double s = 0;
#pragma unroll
    for (int q = 0; q < PEAK_ILP; ++q) s += c[q][0] + c[q][1];
#pragma unroll
    for (int q = 0; q < 16; ++q) s += f[q];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

inline int measure_fp64_peak(int device, int which, double* tflops) {
    if (!tflops || which < 0 || which > 2) return fail(LDC_ERR_INVALID, "ldc_fp64_peak: bad argument");
    LDC_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    LDC_CUDA(cudaGetDeviceProperties(&prop, device));
    const int threads = 256, blocks = prop.multiProcessorCount * 8;
    double* out = nullptr;
    LDC_CUDA(cudaMalloc(&out, sizeof(double) * threads * blocks));
    cudaEvent_t e0, e1;
    LDC_CUDA(cudaEventCreate(&e0));
    LDC_CUDA(cudaEventCreate(&e1));
    double best_ms = 1e30;
    for (int rep = 0; rep < 6; ++rep) {
        LDC_CUDA(cudaEventRecord(e0));
        if (which == 0) k_peak_dmma<<<blocks, threads>>>(out, 1.0000001, 0.5);
        if (which == 1) k_peak_dfma<<<blocks, threads>>>(out, 0.9999999, 1e-3);
        if (which == 2) k_peak_mixed<<<blocks, threads>>>(out, 0.9999999, 1e-3);
        LDC_CUDA(cudaEventRecord(e1));
        LDC_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        LDC_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0 && ms < best_ms) best_ms = ms;
    }
    LDC_CUDA(cudaGetLastError());
    const double warps = (double)blocks * threads / 32.0;
    double flop = 0;
    if (which == 0 || which == 2) flop += warps * PEAK_ITERS * PEAK_ILP * 2.0 * 8 * 8 * 4;
    if (which == 1 || which == 2) flop += (double)blocks * threads * PEAK_ITERS * 16 * 2.0;
    *tflops = flop / (best_ms * 1e-3) / 1e12;
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    return LDC_OK;
}

}  // namespace ldc

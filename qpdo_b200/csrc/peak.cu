// peak.cu -- measured FP64 peaks of the device the library runs on (diagnostics; bench.py uses them as the roofline
// denominators of the numeric factorization, SURVEY.md 8(d): "measure a DMMA/DGEMM peak on the box").
//
// Two register-resident kernels over all SMs, timed with CUDA events:
//   * DMMA: mma.sync.aligned.m8n8k4.row.col.f64 (the FP64 tensor instruction the frontal rank-32 update issues),
//     8 independent accumulator tiles per warp, 512 flops per instruction;
//   * DFMA: 8 independent fma chains per thread, 2 flops per lane and instruction.
#include "dev_internal.h"

#define PK_ITERS 2048
#define PK_TILES 8

__global__ void __launch_bounds__(256) k_peak_dmma(double *out, double seed)
{
    double a = seed + 1e-9 * threadIdx.x, b = 1.0 - 1e-9 * threadIdx.x;
    double c0[PK_TILES], c1[PK_TILES];
#pragma unroll
    for (int t = 0; t < PK_TILES; t++) { c0[t] = 0.0; c1[t] = 0.0; }
    for (int i = 0; i < PK_ITERS; i++) {
#pragma unroll
        for (int t = 0; t < PK_TILES; t++)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
                         : "+d"(c0[t]), "+d"(c1[t]) : "d"(a), "d"(b));
    }
    double s = 0.0;
#pragma unroll
    for (int t = 0; t < PK_TILES; t++) s += c0[t] + c1[t];
    if (s == 12345.678) out[0] = s;            // keeps the accumulators live
}

__global__ void __launch_bounds__(256) k_peak_dfma(double *out, double seed)
{
    double x[PK_TILES];
    const double a = 1.0 - 1e-9 * threadIdx.x, b = seed;
#pragma unroll
    for (int t = 0; t < PK_TILES; t++) x[t] = seed * t;
    for (int i = 0; i < PK_ITERS; i++) {
#pragma unroll
        for (int t = 0; t < PK_TILES; t++) x[t] = fma(x[t], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int t = 0; t < PK_TILES; t++) s += x[t];
    if (s == 12345.678) out[0] = s;
}

extern "C" int qpdo_dev_measure_fp64(int device, double *dmma_tflops, double *dfma_tflops)
{
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { qpdo_set_error("no CUDA device available"); return -2; }
    if (device < 0) { if (cudaGetDevice(&device) != cudaSuccess) device = 0; }
    CUDA_OK(cudaSetDevice(device));
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    double *out = nullptr;
    CUDA_OK(cudaMalloc(&out, 64));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int grid = sms * 8, block = 256;
    double best[2] = {0.0, 0.0};
    for (int which = 0; which < 2; which++) {
        for (int rep = 0; rep < 5; rep++) {
            cudaEventRecord(e0);
            if (which == 0) k_peak_dmma<<<grid, block>>>(out, 1.0); else k_peak_dfma<<<grid, block>>>(out, 1e-3);
            cudaEventRecord(e1);
            if (cudaEventSynchronize(e1) != cudaSuccess) { qpdo_set_error("peak kernel failed: %s", cudaGetErrorString(cudaGetLastError())); cudaFree(out); return -1; }
            float ms = 0.f;
            cudaEventElapsedTime(&ms, e0, e1);
            const double warps = (double)grid * block / 32.0;
            const double flops = which == 0 ? warps * PK_ITERS * PK_TILES * 512.0 : warps * 32.0 * PK_ITERS * PK_TILES * 2.0;
            const double tf = flops / (ms * 1e-3) / 1e12;
            if (rep > 0 && tf > best[which]) best[which] = tf;       // first launch is the warm-up
        }
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(out);
    if (dmma_tflops) *dmma_tflops = best[0];
    if (dfma_tflops) *dfma_tflops = best[1];
    return 0;
}

/*
 * peak.cu -- FP32 FFMA throughput microbenchmark: the measured denominator of the tracer's FP32
 * roofline fraction (MEASURED_PEAKS.json carries HBM and bf16 tensor peaks only).  8 independent
 * FMA chains per thread, 1024 threads per CTA, 4 CTAs per SM: issue-bound on the FMA pipe.
 */
#include <cuda_runtime.h>

#include "rt_device.h"

#define PK_ITERS 4096
#define PK_CHAINS 8

__global__ void __launch_bounds__(1024) rt_ffma_peak_kernel(float* out, float a, float b) {
    float v[PK_CHAINS];
#pragma unroll
    for (int k = 0; k < PK_CHAINS; k++) v[k] = (float)(threadIdx.x + k);
    for (int i = 0; i < PK_ITERS; i++) {
#pragma unroll
        for (int k = 0; k < PK_CHAINS; k++) v[k] = __fmaf_rn(v[k], a, b);
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < PK_CHAINS; k++) s += v[k];
    if (s == 12345.678f) out[0] = s;   /* never true: keeps the chains alive */
}

extern "C" cudaError_t rt_launch_ffma_peak(int n_sm, cudaStream_t stream, float* ms, double* flops) {
    float* d = nullptr;
    cudaError_t e = cudaMalloc((void**)&d, 4);
    if (e != cudaSuccess) return e;
    const int grid = n_sm * 8, block = 1024;
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    float best = 1e30f;
    for (int rep = 0; rep < 5; rep++) {
        cudaEventRecord(a, stream);
        rt_ffma_peak_kernel<<<grid, block, 0, stream>>>(d, 0.999f, 0.001f);
        RT_LAUNCHED(1);
        cudaEventRecord(b, stream);
        cudaEventSynchronize(b);
        float t = 0.f;
        cudaEventElapsedTime(&t, a, b);
        if (rep > 0 && t < best) best = t;
    }
    e = cudaGetLastError();
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    cudaFree(d);
    *ms = best;
    *flops = 2.0 * PK_ITERS * PK_CHAINS * (double)grid * block;
    return e;
}

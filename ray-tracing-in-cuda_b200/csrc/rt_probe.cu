// rt_probe.cu — FP32 FMA-chain microbenchmark: the measured denominator of the extend/shade roofline
// (BASELINE.md §4 asks for it; MEASURED_PEAKS.json only carries HBM and bf16 tensor numbers).
#include "rt_ctx.h"

__global__ void __launch_bounds__(256) k_fma_chain(float *out, int iters, float a, float b) {
    float x0 = threadIdx.x * 1e-3f, x1 = x0 + 1.0f, x2 = x0 + 2.0f, x3 = x0 + 3.0f, x4 = x0 + 4.0f, x5 = x0 + 5.0f, x6 = x0 + 6.0f, x7 = x0 + 7.0f;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            x0 = __fmaf_rn(x0, a, b); x1 = __fmaf_rn(x1, a, b); x2 = __fmaf_rn(x2, a, b); x3 = __fmaf_rn(x3, a, b);
            x4 = __fmaf_rn(x4, a, b); x5 = __fmaf_rn(x5, a, b); x6 = __fmaf_rn(x6, a, b); x7 = __fmaf_rn(x7, a, b);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
}

extern "C" int rt_probe_fp32_peak(rt_ctx *ctx, double *tflops) {
    if (!ctx || !tflops) return RT_ERR_INVALID;
    cudaSetDevice(ctx->device);
    const int blocks = ctx->num_sms * 8, iters = 4096;
    float *out = nullptr;
    if (cudaMalloc(&out, sizeof(float) * blocks * 256) != cudaSuccess) return rt_fail(ctx, RT_ERR_CUDA, "rt_probe_fp32_peak: cudaMalloc");
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    double best = 0;
    for (int rep = 0; rep < 6; ++rep) {
        cudaEventRecord(e0, ctx->stream);
        k_fma_chain<<<blocks, 256, 0, ctx->stream>>>(out, iters, 0.999f, 1e-3f);
        cudaEventRecord(e1, ctx->stream);
        cudaEventSynchronize(e1);
        float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
        const double fl = 2.0 * 64.0 * iters * (double)blocks * 256.0;
        if (rep >= 2 && ms > 0) best = best > fl / (ms * 1e-3) / 1e12 ? best : fl / (ms * 1e-3) / 1e12;
        ++ctx->launches;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(out);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return rt_fail(ctx, RT_ERR_CUDA, cudaGetErrorString(e));
    *tflops = best;
    return RT_OK;
}

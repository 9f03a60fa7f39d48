// FP32 FMA-chain microbenchmark: the measured denominator of the FP32 roofline
// (MEASURED_PEAKS.json has no FP32 figure).  8 independent chains per thread,
// every SM filled; 2 flop per FFMA.
#include <cuda_runtime.h>
#include "../../include/rt_b200.h"

namespace {
__global__ void __launch_bounds__(256) fma_chain(float* out, int iters, float a, float b) {
    float x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
            x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
}
}  // namespace

extern "C" int rt_fp32_peak_tflops(int device, int iters, double* tflops, float* ms_out) {
    if (!tflops || iters <= 0) return RT_ERR_BAD_ARG;
    if (device >= 0 && cudaSetDevice(device) != cudaSuccess) return RT_ERR_NO_DEVICE;
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return RT_ERR_NO_DEVICE;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return RT_ERR_CUDA;
    const int blocks = sms * 8, threads = 256;
    float* out = nullptr;
    if (cudaMalloc(&out, (size_t)blocks * threads * sizeof(float)) != cudaSuccess) return RT_ERR_CUDA;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    fma_chain<<<blocks, threads>>>(out, iters / 4 + 1, 0.999f, 0.001f);   // warm-up
    cudaEventRecord(e0);
    fma_chain<<<blocks, threads>>>(out, iters, 0.999f, 0.001f);
    cudaEventRecord(e1);
    cudaError_t e = cudaEventSynchronize(e1);
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(out);
    if (e != cudaSuccess) return RT_ERR_CUDA;
    const double flops = 2.0 * 64.0 * (double)iters * (double)blocks * threads;
    *tflops = flops / ((double)ms * 1e-3) / 1e12;
    if (ms_out) *ms_out = ms;
    return RT_OK;
}

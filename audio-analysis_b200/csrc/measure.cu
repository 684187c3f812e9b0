// Measurement hook (no reference counterpart): the FP32 FFMA throughput this device sustains, so that bench.py can
// state the fused kernel's arithmetic rate as a fraction of a peak MEASURED IN THE SAME RUN instead of a constant.
#include "api_util.h"
#include "cac_common.cuh"

namespace cac {
__global__ void __launch_bounds__(256) ffma_peak_kernel(float* out, int iters, float c) {
    float a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 0.001f + i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
#pragma unroll
            for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], 0.999f, c);      // 16 independent chains per thread
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
}  // namespace cac

// *tflops = 2 flop x FFMA lane-operations per second / 1e12, best of three launches of ~5 ms each.
// Allocates a small buffer and synchronises (a measurement, not part of any data path).
extern "C" int cac_measure_fp32_tflops(int device, float* tflops) {
    if (!tflops) return cac::fail(CAC_ERR_INVALID, "null argument");
    CUDA_TRY(cudaSetDevice(device));
    int sms = 148;
    CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    const int blocks = sms * 8, threads = 256, iters = 4096;
    float* d = nullptr;
    CUDA_TRY(cudaMalloc((void**)&d, (size_t)blocks * threads * sizeof(float)));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cac::ffma_peak_kernel<<<blocks, threads>>>(d, 64, 0.5f);
    float best = 0.f;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0);
        cac::ffma_peak_kernel<<<blocks, threads>>>(d, iters, 0.5f);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        const double flops = 2.0 * (double)blocks * threads * iters * 4 * 16;
        const float t = (float)(flops / (ms * 1e-3) / 1e12);
        if (t > best) best = t;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaError_t e = cudaGetLastError();
    cudaFree(d);
    if (e != cudaSuccess) return cac::fail(CAC_ERR_CUDA, cudaGetErrorString(e));
    *tflops = best;
    return CAC_OK;
}

// Microbenchmark: issue rate of scalar FFMA / FADD vs packed FFMA2 / FADD2 on sm_100a (one-off measurement
// that decided whether the FFT kernels should use packed FP32).  nvcc -gencode arch=compute_100a,code=sm_100a
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(float* out, int iters, float c) {
    float a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 0.001f + i;
    float2* p = reinterpret_cast<float2*>(a);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            if (MODE == 0) {
#pragma unroll
                for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], 0.999f, c);
            } else if (MODE == 1) {
#pragma unroll
                for (int i = 0; i < 8; ++i) p[i] = __ffma2_rn(p[i], make_float2(0.999f, 0.999f), make_float2(c, c));
            } else if (MODE == 2) {
#pragma unroll
                for (int i = 0; i < 16; ++i) a[i] = a[i] + c;
            } else if (MODE == 3) {
#pragma unroll
                for (int i = 0; i < 8; ++i) p[i] = __fadd2_rn(p[i], make_float2(c, c));
            } else if (MODE == 4) {   // 3-register scalar FFMA
#pragma unroll
                for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], a[(i + 1) & 15], c);
            } else {                  // 3-register packed
#pragma unroll
                for (int i = 0; i < 8; ++i) p[i] = __ffma2_rn(p[i], p[(i + 1) & 7], make_float2(c, c));
            }
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE>
void run(const char* name, float* d) {
    const int iters = 4096, blocks = 148 * 8, threads = 256;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<blocks, threads>>>(d, 16, 0.5f);
    cudaEventRecord(e0);
    k<MODE><<<blocks, threads>>>(d, iters, 0.5f);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    double lane_ops = (double)blocks * threads * iters * 4 * 16;      // scalar-equivalent operations
    printf("%-28s %8.3f ms  %7.2f T lane-ops/s  (%.1f per clk per SM at 1.965 GHz)\n", name, ms, lane_ops / ms / 1e9,
           lane_ops / (ms * 1e-3) / 148 / 1.965e9);
}
int main() {
    float* d;
    cudaMalloc(&d, 148 * 8 * 256 * 4);
    run<0>("FFMA imm (scalar)", d);
    run<1>("FFMA2 imm (packed)", d);
    run<2>("FADD (scalar)", d);
    run<3>("FADD2 (packed)", d);
    run<4>("FFMA 3-reg (scalar)", d);
    run<5>("FFMA2 3-reg (packed)", d);
    return 0;
}

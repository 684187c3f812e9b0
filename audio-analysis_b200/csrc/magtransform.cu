// MagTransform.call (src/magtransform.py:17-19, src/magtransformv2.py:19-21, duplicate at
// src/identify_tracks.py:37-39): y = x ** sigmoid(a), elementwise float32.
#include "api_util.h"
#include "cac_common.cuh"

namespace cac {

__global__ void __launch_bounds__(256) magtransform_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                           long long n, float a) {
    // tf.math.sigmoid(a) in float32, then tf.math.pow(x, p)
    const float p = __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-a)));
    const long long stride = (long long)gridDim.x * 256 * 4;
    for (long long i = ((long long)blockIdx.x * 256 + threadIdx.x) * 4; i < n; i += stride) {
        if (i + 3 < n && ((reinterpret_cast<uintptr_t>(in + i) | reinterpret_cast<uintptr_t>(out + i)) & 15) == 0) {
            float4 v = *reinterpret_cast<const float4*>(in + i);
            v.x = powf(v.x, p); v.y = powf(v.y, p); v.z = powf(v.z, p); v.w = powf(v.w, p);
            *reinterpret_cast<float4*>(out + i) = v;
        } else {
            for (int e = 0; e < 4 && i + e < n; ++e) out[i + e] = powf(in[i + e], p);
        }
    }
}

}  // namespace cac

extern "C" int cac_magtransform_run(const float* d_in, float* d_out, int64_t n, float a, void* stream) {
    if (!d_in || !d_out || n < 0) return cac::fail(CAC_ERR_INVALID, "bad argument");
    if (n == 0) return CAC_OK;
    long long blocks = (n + 1023) / 1024;
    if (blocks > 148LL * 16) blocks = 148LL * 16;
    cac::magtransform_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(d_in, d_out, n, a);
    CUDA_TRY(cudaGetLastError());
    return CAC_OK;
}

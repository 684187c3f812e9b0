// get_spect's tail on the device (src/identify_tracks.py:267, 282-287): the classifier hand-off.
// in  [n_rows][n_frames]            (n_rows = segments x mel rows, dB or power values)
// out [n_rows][n_frames][channels]  = (in - mean over frames, if mean_sub) repeated over the channel axis
// One warp per row: the row is read once (coalesced), the mean is accumulated in float64 and rounded once
// (tf.reduce_mean / np.mean in float32 differ from it by summation order only), then written back.
#include "api_util.h"
#include "cac_common.cuh"

namespace cac {

__global__ void __launch_bounds__(256) spect_tail_kernel(const float* in, float* out,
                                                         long long n_rows, int n_frames, int mean_sub, int channels) {
    const int lane = threadIdx.x & 31;
    const long long warps = (long long)gridDim.x * 8;
    for (long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); row < n_rows; row += warps) {
        const float* src = in + row * n_frames;
        float mean = 0.f;
        if (mean_sub) {
            double acc = 0.0;
            for (int f = lane; f < n_frames; f += 32) acc += (double)src[f];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
            mean = (float)(acc / (double)n_frames);
        }
        float* dst = out + row * (long long)n_frames * channels;
        if (channels == 1) {
            for (int f = lane; f < n_frames; f += 32) dst[f] = __fsub_rn(src[f], mean);
        } else {
            const int n = n_frames * channels;                 // consecutive lanes write consecutive floats
            for (int i = lane; i < n; i += 32) dst[i] = __fsub_rn(src[i / channels], mean);
        }
    }
}

}  // namespace cac

extern "C" int cac_spect_tail_run(const float* d_in, int64_t n_rows, int32_t n_frames, int32_t mean_sub,
                                  int32_t channels, float* d_out, void* stream) {
    if (!d_in || !d_out || n_rows < 0 || n_frames < 1 || channels < 1) return cac::fail(CAC_ERR_INVALID, "bad argument");
    if (channels > 1 && d_in == d_out) return cac::fail(CAC_ERR_INVALID, "in place needs channels == 1");
    if (n_rows == 0) return CAC_OK;
    long long blocks = (n_rows + 7) / 8;
    if (blocks > 148LL * 32) blocks = 148LL * 32;
    cac::spect_tail_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(d_in, d_out, n_rows, n_frames, mean_sub,
                                                                             channels);
    CUDA_TRY(cudaGetLastError());
    return CAC_OK;
}

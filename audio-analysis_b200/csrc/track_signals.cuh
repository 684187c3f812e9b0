// Launchers of track_signals.cu (signal_noise after its STFT), used by track_api.cu.
#pragma once
#include <cuda_runtime.h>

namespace cac {
cudaError_t launch_transpose(const float* in, float* out, int rows, long long cols, cudaStream_t st);
cudaError_t launch_row_median(const float* in, float* out, long long rows, long long cols, cudaStream_t st);
cudaError_t launch_signal_mask(const float* spec, const float* rowmed, const float* colmed, unsigned char* mask, int rows,
                               long long cols, cudaStream_t st);
// out = erode (is_max = false) / dilate (true) of `in` with a kh x kw rectangle; tmp is scratch of the same size
cudaError_t launch_morph_rect(const unsigned char* in, unsigned char* tmp, unsigned char* out, int rows, long long cols,
                              int kh, int kw, bool is_max, cudaStream_t st);
cudaError_t launch_ccl(const unsigned char* img, int* label, int* rootid, int rows, int cols, int* counter, int* stats,
                       int max_components, cudaStream_t st);
}  // namespace cac

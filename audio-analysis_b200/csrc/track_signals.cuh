// Launchers of track_signals.cu (generic transpose + row medians: the column medians of cac_signal_components_run,
// whose spectrogram comes from the caller instead of sn_stft_kernel), used by sn_api.cu.
#pragma once
#include <cuda_runtime.h>

namespace cac {
cudaError_t launch_transpose(const float* in, float* out, int rows, long long cols, cudaStream_t st);
cudaError_t launch_row_median(const float* in, float* out, long long rows, long long cols, cudaStream_t st);
}  // namespace cac

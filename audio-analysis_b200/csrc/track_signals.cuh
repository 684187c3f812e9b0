// Launchers of track_signals.cu / track_bits.cu (signal_noise after its STFT), used by track_api.cu.
#pragma once
#include <cuda_runtime.h>

namespace cac {
cudaError_t launch_transpose(const float* in, float* out, int rows, long long cols, cudaStream_t st);
cudaError_t launch_row_median(const float* in, float* out, long long rows, long long cols, cudaStream_t st);
// bit-packed mask: (cols + 31) / 32 words per row, bit i of word w = pixel 32 w + i, padding bits 0
cudaError_t launch_signal_mask_bits(const float* spec, const float* rowmed, const float* colmed, unsigned* bits, int rows,
                                    long long cols, cudaStream_t st);
// out = erode (is_max = false) / dilate (true) of `in` with a kh x kw rectangle; tmp is scratch of the same size
cudaError_t launch_morph_rect_bits(const unsigned* in, unsigned* tmp, unsigned* out, int rows, long long cols, int kh,
                                   int kw, bool is_max, cudaStream_t st);
cudaError_t launch_unpack_bits(const unsigned* bits, unsigned char* out, int rows, long long cols, cudaStream_t st);
// label / rootid: rows * cols ints each (only run-start entries are touched)
cudaError_t launch_ccl_runs(const unsigned* bits, int* label, int* rootid, int rows, int cols, int* counter, int* stats,
                            int max_components, cudaStream_t st);
}  // namespace cac

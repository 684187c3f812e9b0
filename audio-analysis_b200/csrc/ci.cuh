// Declarations shared by ci.cu (kernels) and ci_api.cu (C-ABI) of the Cacophony Index path.
#pragma once
#include <cuda_runtime.h>

namespace cac {
constexpr int kCiLtabRows = 14 + 64;   // 14 FFT twiddle rows + (u, v) untangle twiddles for K2 = 0..31

struct CiArgs {
    const float* pcm;
    const long long* item_off;    // [n_items] first sample of each recording inside pcm
    const long long* win_base;    // [n_items + 1] prefix sum of windows per recording
    int n_items;
    long long n_windows;
    int first;                    // window w of a recording starts at item_off + first + hop * w
    int hop;
    const double* window;         // [2048] numpy.hanning(2048)
    const double2* ltab;          // [kCiLtabRows][32] per-lane twiddles (built in cac_ci_plan_create)
    double* bins;                 // [n_windows][10]
};

cudaError_t launch_ci_bins(const CiArgs& a, cudaStream_t st);
cudaError_t launch_ci_points(const double* bins, const long long* win_base, int n_items, long long n_windows,
                             unsigned char* points, cudaStream_t st);
cudaError_t launch_ci_entries(const unsigned char* points, const long long* entry_start, const int* entry_len,
                              long long n_entries, int* out, cudaStream_t st);

}  // namespace cac

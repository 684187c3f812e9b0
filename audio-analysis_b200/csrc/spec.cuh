// Declarations shared by the framed-spectrum kernels (stft4800.cu, ...) and their C-ABI (spec_api.cu).
#pragma once
#include <cuda_runtime.h>

namespace cac {

// One launch transforms every frame of n_items recordings (librosa.stft framing: centre = True, zero
// padding, frames = 1 + n_samples / hop).  Frames are numbered globally, recording after recording.
struct SpecArgs {
    const float* pcm;
    const long long* item_off;     // [n_items] first sample of each recording
    const long long* item_len;     // [n_items] samples per recording
    const long long* frame_base;   // [n_items + 1] prefix sum of frames
    int n_items;
    long long n_frames_total;      // < 0: read frame_base[n_items] on the device (tables built by a prep kernel)
    int n_fft;                     // see spec4800_supported()
    int hop;
    int power;                     // 1: |X|, 2: |X|^2
    // output selection
    int kmin, nbins;               // FFT bins [kmin, kmin + nbins) are produced
    int n_mels;                    // 0: raw band magnitudes, else banded mel projection of them
    int bin_major;                 // raw output layout: 0 = [frame][bin] (global frames), 1 = per item [bin][frame]
    float* out;
    float* gmax;                   // optional: atomicMax target (float bits as unsigned), may be NULL
    // get_end's chunk scan fused into the mel epilogue (src/identify_tracks.py:404-412): when chunk_mm is set the mel
    // values are not stored; chunk_mm[item * chunk_stride + frame / chunk_frames] = {min, max} (float bits, values >= 0)
    uint2* chunk_mm;
    int chunk_frames, chunk_stride;
    // get_end in two phases (exact: a chunk whose first frame already holds two different values is not constant).
    // chunk_probe = 1: only frame c * chunk_frames of every chunk slot (item, c) is transformed, global index = item *
    // chunk_stride + c.  chunk_list: only the frames of the listed (item, chunk) pairs, *chunk_list_n of them.
    int chunk_probe;
    const int2* chunk_list;
    const int* chunk_list_n;
    // per-plan tables
    const float2* ltab;            // per-lane twiddle / window constants
    const float2* rtab;            // per-round untangle twiddles
    // mel tables (see mel_bands.h), used when n_mels > 0
    const float* wpad;
    const int* mstart;
    int n_taps, n_groups, pbuf_len;
    int gmax_taps[16], goff[16];
};

bool spec4800_supported(int n_fft);   // frame lengths the kernel is instantiated for (64 R points)
size_t spec4800_smem_bytes(const SpecArgs& a);
cudaError_t launch_spec4800(const SpecArgs& a, cudaStream_t st);
// host-built tables for the 4800-point kernel: ltab [21][32] float2, rtab [76] float2
void spec4800_host_tables(int n_fft, int kmin, float2* ltab, float2* rtab);
constexpr int kSpec4800LtabRows = 21;
constexpr int kSpec4800Rounds = 76;

cudaError_t launch_chunk_minmax(const float* in, int n_rows, long long n_cols, int chunk, long long n_chunks,
                                float2* out, cudaStream_t st);
cudaError_t launch_morepork_windows(const float* spec, const long long* win_frame0, const int* win_width,
                                    long long n_windows, int n_bins, float* out, cudaStream_t st);
cudaError_t launch_scale_by_max(float* data, long long n, const float* gmax, cudaStream_t st);
// get_end for a batch (device item tables): tab = [item_off | item_len | frame_base] (3 n + 1), mm = per-chunk {min, max}
cudaError_t launch_get_end_prep(const long long* d_item_off, const long long* d_item_len, int n_items, int n_fft, int hop,
                                long long max_item_samples, long long* tab, uint2* mm, int chunk_stride, int* out,
                                cudaStream_t st);
// after the probe: the examined chunks whose probe frame was constant go on the list (their full evaluation follows)
cudaError_t launch_get_end_select(const long long* tab, int n_items, const uint2* mm, int chunk_stride, int chunk_frames,
                                  int2* list, int* list_n, cudaStream_t st);
cudaError_t launch_get_end_scan(const long long* tab, int n_items, const uint2* mm, int chunk_stride, int chunk_frames,
                                int* out, cudaStream_t st);

}  // namespace cac

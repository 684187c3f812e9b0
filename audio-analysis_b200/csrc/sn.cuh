// Batched signal_noise chain (src/identify_tracks.py:650-691): declarations shared by sn_stft.cu (spectrum + column
// medians), sn_rows.cu (row medians + threshold mask), track_bits.cu (morphology, connected components) and
// sn_api.cu (C-ABI).
//
// A call works on a WAVE of recordings at a time.  Recording j of the wave owns slot j of every scratch array
// (slot stride = the size the longest permitted recording needs), so the kernels only need the item table below and
// the slot strides; recordings go on grid.y / grid.z, no kernel is launched per recording.
#pragma once
#include <cuda_runtime.h>

namespace cac {

constexpr int kSnRows = 2049;                  // 1 + n_fft / 2 with signal_noise's fixed n_fft = 4096 (:652)

struct SnItem {                                // one recording of the wave, written by sn_prep_kernel
    long long pcm_off;                         // first sample in the PCM buffer
    int n_samples;
    int cols;                                  // frames = 1 + n_samples / hop (librosa.stft, centre = True)
    int wpr;                                   // mask words per row = (cols + 31) / 32
    int stage_base;                            // first stage (frames_per_stage consecutive frames) within the wave
    int n_stages;
    int pitch;                                 // floats per row of the recording's matrix (cols rounded up to 32; cols
                                               // itself when the caller supplies the matrix)
};

struct SnWave {                                // by-value kernel argument
    const SnItem* items;
    int n_items;
    int rows;                                  // 2049 for the fused chain; any for cac_signal_components_run
    float* spec;        long long spec_stride;     // [rows][pitch] float32 magnitudes (unscaled), floats per slot
    float* colmed;      long long col_stride;      // [cols] column medians (unscaled)
    float* colthr;                                 // [cols] fl(3 * fl(colmed / a_max)), same stride
    unsigned* amax;                                // [n_items] max magnitude (float bits)
    unsigned* bits;     long long bits_stride;     // three mask planes per slot, words per slot
    long long plane_stride;                        // words per plane (plane p of slot j at bits + j bits_stride + p plane_stride)
    int* label;         long long label_stride;    // union-find forest over run starts, ints per slot
    int* result;        long long result_stride;   // [4 + 6 * max_components] per recording
    int max_components;
};

// sn_stft.cu
size_t sn_stft_smem_bytes(int frames_per_stage, int hop);
int sn_stft_frames_per_stage(int hop);                       // largest stage that fits the shared memory (0: none)
// builds the wave's item table on the device; lengths are clamped to max_item_samples (the slot size)
cudaError_t launch_sn_prep(const long long* d_item_off, const long long* d_item_len, int n_items, int hop,
                           int frames_per_stage, long long max_item_samples, SnItem* items, unsigned* amax, cudaStream_t st);
cudaError_t launch_sn_stft(const float* pcm, const SnWave& w, int hop, int frames_per_stage, cudaStream_t st);
// sn_rows.cu
cudaError_t launch_sn_col_thresholds(const SnWave& w, int max_cols, cudaStream_t st);
cudaError_t launch_sn_rowmed_mask(const SnWave& w, int max_cols, bool aligned_rows, cudaStream_t st);
constexpr int kSnMaxColsFast = 512 * 48;                     // longest row the register-resident select covers
// track_bits.cu (batched over the wave)
cudaError_t launch_morph_rect_bits_batch(const SnWave& w, int src_plane, int tmp_plane, int dst_plane, int max_wpr, int kh,
                                         int kw, bool is_max, cudaStream_t st);
cudaError_t launch_unpack_bits_batch(const SnWave& w, int plane, unsigned char* out, long long out_stride, int max_cols,
                                     cudaStream_t st);
cudaError_t launch_ccl_runs_batch(const SnWave& w, int plane, int max_wpr, cudaStream_t st);

}  // namespace cac

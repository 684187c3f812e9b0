/* cacophony_b200 - C-ABI of the B200-native feature hot path of
 * TheCacophonyProject/audio-analysis.
 *
 * The reference is pure Python and has no FFI/plugin boundary of its own:
 * the seam is a set of Python callables.  Each entry point below names the reference
 * callable (paths relative to /root/reference) whose arithmetic it replaces; the Python
 * mirror of those callables lives in audio-analysis_b200/ and binds this header with
 * ctypes (INTEGRATION.md shows the stub a reference maintainer would add).
 *
 * Conventions: every function returns 0 on success or a negative cac_status; the message
 * for the calling thread's last failure is cac_last_error().  Pointers named d_* are device
 * pointers owned by the caller (e.g. torch tensors' data_ptr()), h_* are host pointers.
 * `stream` is a cudaStream_t passed as void* (NULL = default stream).  *_run functions do
 * not allocate or synchronise (cac_spec_run, cac_ci_run, cac_ci_entries_run, cac_morepork_windows_run and
 * cac_resample_run take small HOST offset tables and stage them with one cudaMemcpyAsync into a plan-owned
 * buffer that grows on demand; the cac_sn_* and cac_get_end_* entry points take device tables and never
 * touch the host); *_host functions copy in, run, copy out and synchronise.  cac_features_run cannot
 * validate device-side segment descriptors: out-of-range descriptors are the caller's responsibility
 * (cac_features_host checks them).
 * There is no CPU fallback anywhere behind this interface.
 */
#ifndef CACOPHONY_B200_H
#define CACOPHONY_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum cac_status {
    CAC_OK = 0,
    CAC_ERR_INVALID = -1,     /* bad argument / unsupported parameter combination */
    CAC_ERR_CUDA = -2,        /* a CUDA runtime call failed (message has the detail) */
    CAC_ERR_NO_DEVICE = -3,   /* no CUDA device: this library never computes on the CPU */
    CAC_ERR_ALLOC = -4
} cac_status;

/* DSP parameters of one classifier model: the keys of the model's metadata.txt read in
 * identify_tracks.classify (src/identify_tracks.py:466-497) and passed to load_samples
 * (src/identify_tracks.py:65-86). */
typedef struct cac_params {
    int32_t sample_rate;     /* sr, 48000 after load_recording (src/identify_tracks.py:49-57) */
    int32_t n_fft;           /* 4096 (src/identify_tracks.py:85) unless the model's metadata sets it (:493): the fused
                              * kernel serves 4096 with bands below bin 1024; 4096 with higher bands, 2048, 1024, 512,
                              * 4800, 3200 and 1600 run unfused (segment rows -> framed spectrum + mel -> dB); any
                              * other length makes cac_features_run fail with CAC_ERR_INVALID */
    int32_t hop_length;      /* 281 default of load_samples (:71); 640 classify fallback (:474) */
    int32_t n_mels;          /* 160 */
    int32_t segment_samples; /* int(sr * segment_length) (:109) */
    int32_t power;           /* 1 or 2: exponent on |STFT| (src/custommel.py:61) */
    int32_t db_scale;        /* librosa.power_to_db(ref=np.max) (src/identify_tracks.py:265-266) */
    int32_t normalize;       /* normalize_data (src/identify_tracks.py:202-209) */
} cac_params;

/* One classifier segment = one get_spect call of the reference's segment loop
 * (src/identify_tracks.py:163-196).  The segment's `segment_samples` samples are
 * pad_left zeros, then n_valid samples pcm[pcm_offset ...], then zeros (np.pad at :165-168). */
typedef struct cac_segment {
    int64_t pcm_offset;
    int32_t n_valid;
    int32_t pad_left;
} cac_segment;

typedef struct cac_plan cac_plan;

const char* cac_version(void);
const char* cac_last_error(void);
/* Number of visible CUDA devices (0 or negative status if none). */
int cac_device_count(void);

/* Plan for the STFT -> custom mel -> dB chain of get_spect (src/identify_tracks.py:212-288,
 * htk=True branch) = librosa.stft (:243) + custommel.mel_spec (src/custommel.py:59-63) +
 * librosa.power_to_db (:266).  `mel_weights` is the dense float32 [n_mels, 1+n_fft/2]
 * filterbank exactly as custommel.mel_f returns it (src/custommel.py:19-56); rows must be
 * contiguous bands (they are, for triangular filters).  The plan is bound to `device`. */
int cac_plan_create(const cac_params* params, const float* h_mel_weights, int device, cac_plan** out);
int cac_plan_destroy(cac_plan* plan);
/* Frames per segment = 1 + segment_samples / hop_length (librosa.stft, center=True). */
int cac_plan_frames(const cac_plan* plan);
/* Measurement hook (no reference counterpart): when enabled, cac_features_run brackets its dominant
 * kernel (the fused STFT->mel kernel) with CUDA events on the launch stream; cac_plan_last_kernel_ms
 * waits for that kernel and returns its duration. */
int cac_plan_set_profiling(cac_plan* plan, int enable);
int cac_plan_last_kernel_ms(cac_plan* plan, float* ms);
/* MagTransform.call (src/magtransform.py:17-19, src/magtransformv2.py:19-21: y = x ** sigmoid(a), the first layer
 * of models whose metadata sets db_scale False, selected by `magv2` at src/identify_tracks.py:313-316) as an epilogue
 * of the fused kernel: with it enabled cac_features_run / cac_features_host write mel ** sigmoid(a) directly - one
 * powf per value in the kernel's flush instead of another read and write of the whole output (cac_magtransform_run).
 * Only for plans with db_scale = 0 that the fused kernel serves (n_fft 4096, bands below bin 1024). */
int cac_plan_set_magtransform(cac_plan* plan, int enable, float a);
/* Measurement hook (no reference counterpart): FP32 FFMA throughput the device sustains (2 flop per lane-operation,
 * TFLOP/s, best of three ~5 ms launches) - the denominator of bench.py's FP32 fraction, measured in the same run. */
int cac_measure_fp32_tflops(int device, float* tflops);
/* Bytes of device scratch cac_features_run needs for n_segments (the unfused path adds the normalised segment rows). */
size_t cac_features_work_bytes(const cac_plan* plan, int64_t n_segments);

/* load_samples' inner loop (src/identify_tracks.py:163-190) for n_segments segments at once:
 * d_out[s][m][f] float32 [n_segments, n_mels, frames] = get_spect(normalize_data(segment s)). */
int cac_features_run(cac_plan* plan, const float* d_pcm, const cac_segment* d_segments,
                     int64_t n_segments, float* d_out, void* d_work, void* stream);
/* Same through host buffers: the PCM upload, the kernels and the download of finished segment chunks
 * run on three streams and overlap when h_pcm / h_out are page-locked (cac_host_register, or any
 * cudaHostAlloc'ed / torch pin_memory buffer); returns after the last byte of h_out has landed. */
int cac_features_host(cac_plan* plan, const float* h_pcm, int64_t n_pcm, const cac_segment* h_segments,
                      int64_t n_segments, float* h_out);

/* Same call for 16-bit PCM as it leaves the decoder: samples are widened on the device to sample / 32768.0f,
 * the value librosa.load / soundfile hand to identify_tracks.load_recording (src/identify_tracks.py:51-58) for
 * 16-bit sources, so results equal cac_features_host on the widened array bit for bit; the upload is half
 * the bytes, which matters because the host path runs at PCIe speed. */
int cac_features_host_pcm16(cac_plan* plan, const int16_t* h_pcm, int64_t n_pcm, const cac_segment* h_segments,
                            int64_t n_segments, float* h_out);
/* Page-lock / release a caller-owned host buffer (cudaHostRegister) so that *_host calls can overlap copies. */
int cac_host_register(void* h_ptr, size_t bytes);
int cac_host_unregister(void* h_ptr);

/* normalize_data (src/identify_tracks.py:202-209) on n_rows rows of row_len samples (bit-exact op order). */
int cac_normalize_run(const float* d_in, float* d_out, int64_t n_rows, int32_t row_len, void* stream);

/* custommel.mel_spec (src/custommel.py:59-63): d_out[n_mels, n_frames] = W . |stft|**power, with
 * d_mag the magnitude matrix [1+n_fft/2, n_frames] row-major (what np.abs(librosa.stft) returns). */
int cac_mel_project_run(cac_plan* plan, const float* d_mag, int32_t n_frames, float* d_out, void* stream);

/* librosa.power_to_db(S, ref=np.max) with top_db=80 on n_items independent blocks of item_len values
 * (in place allowed). */
int cac_power_to_db_run(const float* d_in, float* d_out, int64_t n_items, int64_t item_len, void* d_work,
                        void* stream);

/* cac_power_to_db_run without d_work and without any allocation (no cudaMalloc / cudaFree, no allocating library
 * call): the per-item maxima go to a static device array inside the library.  It has NO counterpart in the
 * reference; it was added for callers that cannot allocate (CUDA graph
 * capture, tight serving loops).  Works on the CURRENT device, like cac_power_to_db_run.  n_items above
 * CAC_POWER_TO_DB_STATIC_MAX_ITEMS is rejected with CAC_ERR_INVALID.  All calls share the one array: calls that
 * overlap on the GPU race; serialise them (one stream) or use cac_power_to_db_run.  Results are bit-identical to
 * cac_power_to_db_run's. */
#define CAC_POWER_TO_DB_STATIC_MAX_ITEMS 65536
int cac_power_to_db_static(const float* d_in, float* d_out, int64_t n_items, int64_t item_len, void* stream);

/* ---- old Cacophony Index (src/cacophony_index.py) ------------------------------------------------
 * float64 throughout, like the reference (numpy.hanning * window -> scipy.fftpack.dct -> band sums). */
typedef struct cac_ci_plan cac_ci_plan;

/* h_edges: the 11 band edges get_ci_bins derives from (window_size, sample_rate)
 * (src/cacophony_index.py:58-63); h_window: numpy.hanning(window_size) (src/common.py:39-40).
 * Only the layout calculate() uses (window 2048 at 16 kHz: edges 25..2048) is built; others fail. */
int cac_ci_plan_create(int32_t window_size, const int32_t* h_edges, const double* h_window, int device,
                       cac_ci_plan** out);
int cac_ci_plan_destroy(cac_ci_plan* plan);
/* len(range(1024, n_samples - 3072, 1024)): windows calculate()'s loop visits (src/cacophony_index.py:89-91). */
int64_t cac_ci_window_count(int64_t n_samples);

/* get_ci_bins (src/cacophony_index.py:53-66) on n_windows windows d_pcm[hop*w .. hop*w + 2048):
 * d_bins[w][10] float64 band energies. */
int cac_ci_bins_run(cac_ci_plan* plan, const float* d_pcm, int64_t n_windows, int32_t hop, double* d_bins,
                    void* stream);

/* calculate()'s hot loop (src/cacophony_index.py:86-97) for n_items recordings stored in d_pcm at
 * h_item_offset[i] (h_item_samples[i] samples each, 16 kHz).  Windows are numbered globally,
 * recording after recording (cac_ci_window_count each).  d_bins[window][10] = get_ci_bins;
 * d_points[window] (may be NULL) = the reference's `points` entry for that window against the
 * previous one, 255 for the first window of each recording (it only seeds previous_bins). */
int cac_ci_run(cac_ci_plan* plan, const float* d_pcm, const int64_t* h_item_offset, const int64_t* h_item_samples,
               int64_t n_items, double* d_bins, uint8_t* d_points, void* stream);
/* Same through host buffers. */
int cac_ci_host(cac_ci_plan* plan, const float* h_pcm, int64_t n_pcm, const int64_t* h_item_offset,
                const int64_t* h_item_samples, int64_t n_items, double* h_bins, uint8_t* h_points);

/* score_from_points' order statistics (src/cacophony_index.py:69-73) for n_entries slices
 * d_points[h_entry_start[e] .. + h_entry_len[e]): d_out[e] = {sum(sorted[k0:k1]), k1 - k0} with
 * k0 = int(len*0.75), k1 = int(len*0.95).  The float tail (mean, correction curve, rounding)
 * stays with the caller. */
int cac_ci_entries_run(cac_ci_plan* plan, const uint8_t* d_points, const int64_t* h_entry_start,
                       const int32_t* h_entry_len, int64_t n_entries, int32_t* d_out, void* stream);

/* ---- framed spectra: identify_morepork, get_end, signal_noise's STFT --------------------------------
 * librosa.stft(frames, n_fft, hop_length) with its defaults (periodic Hann, centre = True, zero
 * padding, 1 + len // hop frames) followed by |.|**power, restricted to the FFT bins the caller
 * consumes, optionally projected through a custommel.mel_f filterbank. */
typedef struct cac_spec_params {
    int32_t n_fft;        /* 4800 = int(sr / 10) at 48 kHz (src/identify_morepork.py:23, src/identify_tracks.py:389).
                           * Mixed-radix kernel: 4800, 3200, 1600, 4096, 2048, 1024, 512.  Any other length up to 19,000,
                           * even or odd (4410 at 44.1 kHz, 2205 at 22.05 kHz, ...): raw bands of at most 512 bins by a
                           * direct band DFT (band_dft.cu) - what _load_sample needs at the file's native rate (:20-25) */
    int32_t hop_length;   /* int(n_fft / 2) for morepork (:24); 281 for get_end (:388) */
    int32_t bin_lo;       /* raw output (n_mels == 0): FFT bins [bin_lo, bin_hi); morepork 60:120 (:25) */
    int32_t bin_hi;
    int32_t n_mels;       /* > 0: mel projection (src/custommel.py:59-63); get_end uses 120 (:390-399) */
    int32_t power;        /* 1 or 2 */
    int32_t bin_major;    /* raw output layout: 0 = [frame][bin] over all frames, 1 = per recording [bin][frame] */
} cac_spec_params;
typedef struct cac_spec_plan cac_spec_plan;

int cac_spec_plan_create(const cac_spec_params* params, const float* h_mel_weights, int device,
                         cac_spec_plan** out);
int cac_spec_plan_destroy(cac_spec_plan* plan);
/* librosa's frame count: 1 + (n_samples + 2 (n_fft // 2) - n_fft) // hop_length (= 1 + n_samples // hop_length for even n_fft) */
int64_t cac_spec_frame_count(const cac_spec_plan* plan, int64_t n_samples);
/* rows per frame of the output: n_mels, or bin_hi - bin_lo */
int32_t cac_spec_rows(const cac_spec_plan* plan);
/* Spectra of n_items recordings stored in d_pcm at h_item_offset[i] (h_item_samples[i] samples).
 * Frames are numbered globally, recording after recording.  Output: mel -> per recording
 * [n_mels][frames_i] (recordings back to back); raw -> see bin_major.  d_max (may be NULL)
 * receives the largest |X|**power produced (signal_noise's a_max, src/identify_tracks.py:656).
 * Raw-band plans with hop_length = n_fft / 2, hop_length a multiple of 16 and at most 62 bins (identify_morepork's
 * spectrogram, src/identify_morepork.py:19-27) run as a tensor-core GEMM (tcgen05, csrc/band_gemm.cu) when every recording
 * starts on a 16-byte boundary - through one TMA tensor map when the recordings are equally long, a whole number of hops
 * and back to back; max error 1-3e-6 of the spectrum maximum instead of 1e-7 (BASELINE's bar for the path: 1e-2 dB).
 * The environment variable CAC_BAND_GEMM=0 keeps the FFT kernels for such plans.  The plan grows a scratch buffer of
 * 512 bytes per frame for it on the first call (cudaMalloc; like the item table this entry point stages). */
int cac_spec_run(cac_spec_plan* plan, const float* d_pcm, const int64_t* h_item_offset, const int64_t* h_item_samples,
                 int64_t n_items, float* d_out, float* d_max, void* stream);
/* d_data[i] /= *d_max  (spectogram / a_max, src/identify_tracks.py:657) */
int cac_scale_by_max_run(float* d_data, int64_t n, const float* d_max, void* stream);
/* get_end's chunk scan (src/identify_tracks.py:404-412): d_out[c] = {min, max} of
 * d_in[:, c*chunk : (c+1)*chunk] for the n_cols // chunk complete chunks of a row-major [n_rows][n_cols] matrix. */
int cac_chunk_minmax_run(const float* d_in, int32_t n_rows, int64_t n_cols, int32_t chunk, float* d_out, void* stream);
/* identify_tracks.get_end (src/identify_tracks.py:387-413) for n_items recordings with one launch chain and no host
 * round trip: `plan` is the n_fft = sr // 10, hop 281, 120-mel, power 1 plan get_end uses (:389-399); recording i =
 * d_pcm[d_item_offset[i] .. + d_item_samples[i]) (DEVICE arrays; lengths above max_item_samples are clamped).  The
 * mel values never reach HBM: the kernel's epilogue keeps each chunk's extrema (chunk_frames = sr // 281 frames,
 * :404) and a scan kernel applies the reference's loop (:406-412).  d_out[i] = index of the first chunk whose maximum
 * equals its minimum (the caller returns chunk * chunk_frames * 281 // sr seconds), or -1 (return len / sr).
 * d_work: cac_get_end_work_bytes(...) bytes, 256-byte aligned.  Neither allocates nor synchronises. */
size_t cac_get_end_work_bytes(const cac_spec_plan* plan, int64_t n_items, int64_t max_item_samples, int32_t chunk_frames);
int cac_get_end_run(cac_spec_plan* plan, const float* d_pcm, const int64_t* d_item_offset, const int64_t* d_item_samples,
                    int64_t n_items, int64_t max_item_samples, int32_t chunk_frames, int32_t* d_out, void* d_work,
                    size_t work_bytes, void* stream);
/* identify_morepork's window loop body (src/identify_morepork.py:150-166): window w covers
 * h_win_width[w] (<= 60) frames of the frame-major raw spectrum d_spec starting at global frame
 * h_win_frame0[w]; librosa.amplitude_to_db(ref=np.max), rescale by |min|, zero-pad to 60 columns;
 * d_out[w][bin][60]. */
int cac_morepork_windows_run(cac_spec_plan* plan, const float* d_spec, const int64_t* h_win_frame0,
                             const int32_t* h_win_width, int64_t n_windows, float* d_out, void* stream);

/* ---- signal_noise (src/identify_tracks.py:650-691) for batches of recordings --------------------------------------
 * |librosa.stft(frames, n_fft=4096, hop_length)| (:652-654), a_max (:656), np.median along both axes of spectogram / a_max
 * (:657-659), signal = (spec > 3 col_median) & (spec > 3 row_median) (:667), cv2.morphologyEx(MORPH_OPEN, ones((4, 4)))
 * (:669-671), cv2.dilate(ones((dilate_h, dilate_w))), cv2.erode(ones((erode_h, dilate_w))) (:683-684; erode_h == 0 is
 * cv2's empty kernel = its 3 x 3 default) and cv2.connectedComponentsWithStats (8-connected, :686) - everything up to
 * the component table.  dilate_h = `height`, dilate_w = `width`, erode_h = height // 10 as derived at :673-684
 * (audio-analysis_b200/identify_tracks.py computes them from sr and hop).  The matrix is written to HBM once and read
 * once: column medians are selected by the warp that transformed the frame, row medians and the threshold mask by the
 * CTA that holds the row in registers, both on the unscaled magnitudes (division by a_max is monotone, the selected
 * values are scaled afterwards - bit-identical to the reference's order of operations).  Recordings are processed in
 * waves of as many as fit the work buffer; no launch, allocation, copy or synchronisation per recording. */
typedef struct cac_sn_plan cac_sn_plan;
int cac_sn_plan_create(int32_t hop_length, int32_t dilate_h, int32_t dilate_w, int32_t erode_h, int device,
                       cac_sn_plan** out);
int cac_sn_plan_destroy(cac_sn_plan* plan);
/* columns of the spectrogram: 1 + n_samples // hop_length (librosa.stft, centre = True) */
int64_t cac_sn_frame_count(const cac_sn_plan* plan, int64_t n_samples);
/* recommended size of d_work for n_items recordings of at most max_item_samples samples (24 wave slots of
 * 4 * 2049 * frames + ~8 MB each; any size that holds one slot works, cac_sn_run just takes more waves) */
size_t cac_sn_work_bytes(const cac_sn_plan* plan, int64_t n_items, int64_t max_item_samples);
/* Recording i = d_pcm[d_item_offset[i] .. + d_item_samples[i]) - DEVICE arrays (a length above max_item_samples is
 * clamped to it).  d_result[i] = 4 + 6 * max_components int32: [0] = number of components (the recording's table is
 * invalid if it exceeds max_components), then from [4] per component {min x, min y, max x, max y, area, order}
 * (cv2's stats: x, y, width = max x - min x + 1, height likewise, area); `order` sorts the components like cv2's
 * labels (first 2 x 2 block in raster order).  d_work: 256-byte aligned.  Neither allocates nor synchronises. */
int cac_sn_run(cac_sn_plan* plan, const float* d_pcm, const int64_t* d_item_offset, const int64_t* d_item_samples,
               int64_t n_items, int64_t max_item_samples, int32_t max_components, int32_t* d_result, void* d_work,
               size_t work_bytes, void* stream);
/* The first stage alone for one recording (:652-656): d_spec [2049][pitch] = |stft| (not yet divided by its maximum;
 * pitch = cols rounded up to a multiple of 32 floats, the kernel writes 16-byte aligned runs; columns >= cols are
 * not written), d_colmed [cols] = np.median(|stft|, axis=0), *d_amax = np.amax(|stft|).  d_work: 512 bytes, 256-byte aligned. */
int cac_sn_spectrum_run(cac_sn_plan* plan, const float* d_pcm, int64_t n_samples, float* d_spec, float* d_colmed,
                        float* d_amax, void* d_work, size_t work_bytes, void* stream);

/* HOST function, no device work: the merge sweeps of track finding - merge_signals (src/identify_tracks.py:725-792)
 * repeated until a sweep merges nothing, as get_tracks_from_signals does before its length filter (:800-802).
 * boxes[n][6] = start, end, freq_start, freq_end, mel_freq_start, mel_freq_end of the candidate signals (float64,
 * the values the reference's Signal objects hold).  A merged box is the bounding box of what it absorbed, so the
 * result names, per surviving signal and in the reference's final order, which INPUT signals supply its fields:
 * out[j][5] = {the survivor's own index, index whose start it has, ... whose end, ... freq_start, ... freq_end}
 * (n rows are written: the *n_out survivors first, then the absorbed signals with the fields they had when they were
 * absorbed - the reference mutates those objects too).  The caller copies the original objects (ints stay ints).  Same IEEE double
 * operations in the same order as the reference's Python (tracks.py:_should_merge is the readable restatement and
 * the test's checker). */
int cac_tracks_merge(const double* boxes, int32_t n, int32_t* out, int32_t* n_out);

/* The stages after the STFT (:658-686) on a spectrogram the caller supplies: d_spec [rows][cols] = spectogram / a_max.
 * Same kernels as cac_sn_run behind the column medians (taken through a transpose here).  d_result as above for one
 * recording; d_stage_masks (may be NULL) receives the three uint8 [rows][cols] masks (threshold, opened, closed).
 * d_work: cac_signal_components_work_bytes(rows, cols) bytes, 256-byte aligned.  Neither allocates nor synchronises.
 * rows <= 65,535, rows * cols < 2^31. */
size_t cac_signal_components_work_bytes(int32_t rows, int64_t cols);
int cac_signal_components_run(const float* d_spec, int32_t rows, int64_t cols, int32_t dilate_h, int32_t dilate_w,
                              int32_t erode_h, int32_t max_components, int32_t* d_result, uint8_t* d_stage_masks,
                              void* d_work, size_t work_bytes, void* stream);

/* Rational-rate resampling front-end: replaces `ffmpeg -ar 16000` in front of the Cacophony
 * Index (src/common.py:76-79) and librosa.resample to 48 kHz (src/identify_tracks.py:55-57).  The filter is the
 * caller's: `h_taps` is the zero-padded polyphase prototype and `n_pre_remove` the number of leading outputs to
 * drop, exactly the quantities scipy.signal.resample_poly derives (audio-analysis_b200/resample.py builds them);
 * y[m] = sum_k h[k] xu[(m + n_pre_remove) down - k], xu = x zero-stuffed by `up`, zeros outside the item.
 * Every item i reads h_in_samples[i] samples at d_in + h_in_offset[i] and writes cac_resample_out_len() samples
 * at d_out + h_out_offset[i].  No allocation besides the item table, no synchronisation. */
typedef struct cac_resample_plan cac_resample_plan;
int cac_resample_plan_create(int32_t up, int32_t down, const float* h_taps, int32_t n_taps, int32_t n_pre_remove,
                             int device, cac_resample_plan** out);
int cac_resample_plan_destroy(cac_resample_plan* plan);
int64_t cac_resample_out_len(const cac_resample_plan* plan, int64_t n_in);
int cac_resample_run(cac_resample_plan* plan, const float* d_in, const int64_t* h_in_offset,
                     const int64_t* h_in_samples, int64_t n_items, float* d_out, const int64_t* h_out_offset,
                     void* stream);

/* get_spect's tail (src/identify_tracks.py:267, 282-287) for the classifier hand-off on the device:
 * d_in [n_rows][n_frames] (n_rows = segments x mel rows) -> d_out [n_rows][n_frames][channels] =
 * (value - mean over the row's frames, when mean_sub) repeated `channels` times (tf.expand_dims /
 * tf.reduce_mean(axis=1) / tf.repeat(axis=2); classify's efficientnet x3 repeat, :539-540, is channels = 3).
 * In place (d_out == d_in) is allowed when channels == 1. */
int cac_spect_tail_run(const float* d_in, int64_t n_rows, int32_t n_frames, int32_t mean_sub, int32_t channels,
                       float* d_out, void* stream);

/* MagTransform.call (src/magtransform.py:17-19, src/magtransformv2.py:19-21): d_out = d_in ** sigmoid(a),
 * elementwise float32 (in place allowed); `a` is the layer's learned scalar (v1 init 0.0, v2 init -1.0). */
int cac_magtransform_run(const float* d_in, float* d_out, int64_t n, float a, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CACOPHONY_B200_H */

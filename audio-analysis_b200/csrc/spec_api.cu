// C-ABI of the framed-spectrum paths (include/cacophony_b200.h, cac_spec_* and helpers):
// identify_morepork._load_sample + window loop, identify_tracks.get_end, signal_noise's STFT.
#include <cstdlib>
#include <cstring>
#include <vector>

#include "api_util.h"
#include "cac_common.cuh"
#include "mel_bands.h"
#include "spec.cuh"

using namespace cac;

struct cac_spec_plan {
    cac_spec_params p;
    int device;
    int kmin, nbins;
    MelBands mb;
    float2* d_ltab = nullptr;
    float2* d_rtab = nullptr;
    float* d_wpad = nullptr;
    int* d_mstart = nullptr;
    long long* d_tab = nullptr;   size_t tab_cap = 0;    // item tables staged per call
    long long* d_wtab = nullptr;  size_t wtab_cap = 0;   // morepork window tables
    bool dft = false;             // any other frame length: direct band DFT (band_dft.cu), raw bands only
    float2* d_tw = nullptr;
    // hop = n_fft / 2 and a raw band of at most 62 bins (identify_morepork): the band DFT as a tensor-core GEMM over
    // non-overlapping chunks (band_gemm.cu); d_P = the chunks' partial transforms, grown on demand by cac_spec_run
    bool gemm = false;
    float* d_bsw = nullptr;
    float* d_P = nullptr;         size_t P_cap = 0;
};

namespace cac {
void band_dft_host_table(int n_fft, float2* tw);
size_t band_dft_smem_bytes(int n_fft);
cudaError_t launch_band_dft(const SpecArgs& a, const float2* tw, cudaStream_t st);
bool band_gemm_supported(int n_fft, int hop, int kmin, int nbins);
size_t band_gemm_table_floats(int hop);
size_t band_gemm_scratch_bytes(long long total_rows);
void band_gemm_host_table(int n_fft, int hop, int kmin, int nbins, float* out);
cudaError_t launch_band_gemm(const SpecArgs& a, const float* base, int uniform_chunks, const float* d_bsw, float* d_P,
                             cudaStream_t st);
}  // namespace cac

// the GEMM tables of a raw-band plan with hop = n_fft / 2 (no error when the shape does not qualify: the other kernels serve it)
static int setup_band_gemm(cac_spec_plan* plan) {
    const cac_spec_params& p = plan->p;
    if (p.n_mels != 0 || p.bin_major || !band_gemm_supported(p.n_fft, p.hop_length, plan->kmin, plan->nbins)) return CAC_OK;
    std::vector<float> tab(band_gemm_table_floats(p.hop_length));
    band_gemm_host_table(p.n_fft, p.hop_length, plan->kmin, plan->nbins, tab.data());
    cudaError_t e = cudaMalloc((void**)&plan->d_bsw, tab.size() * sizeof(float));
    if (e == cudaSuccess) e = cudaMemcpy(plan->d_bsw, tab.data(), tab.size() * sizeof(float), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) return fail(CAC_ERR_ALLOC, std::string("could not upload the band GEMM table: ") + cudaGetErrorString(e));
    plan->gemm = true;
    return CAC_OK;
}

// librosa.stft(center=True): 1 + (n + 2 (n_fft // 2) - n_fft) // hop frames (= 1 + n // hop for even n_fft)
static long long frames_of(const cac_spec_plan* plan, long long n) {
    const long long padded = n + 2LL * (plan->p.n_fft / 2);
    return padded >= plan->p.n_fft ? 1 + (padded - plan->p.n_fft) / plan->p.hop_length : 0;
}

extern "C" int cac_spec_plan_create(const cac_spec_params* params, const float* h_mel_weights, int device,
                                    cac_spec_plan** out) {
    if (!params || !out) return fail(CAC_ERR_INVALID, "null argument");
    const cac_spec_params& p = *params;
    const bool dft = !spec4800_supported(p.n_fft);
    if (dft && (p.n_mels != 0 || p.n_fft < 2 || band_dft_smem_bytes(p.n_fft) > 227 * 1024 || p.bin_hi - p.bin_lo > 512))
        return fail(CAC_ERR_INVALID,
                    "framed-spectrum kernels are built for n_fft 4800 / 3200 / 1600 (sr // 10 at 48 / 32 / 16 kHz), "
                    "4096, 2048, 1024 and 512; any other n_fft up to 19,000 is served for raw bands of at most 512 "
                    "bins only (direct band DFT)");
    if (p.hop_length < 1 || p.hop_length > p.n_fft) return fail(CAC_ERR_INVALID, "hop_length out of range");
    if (p.power != 1 && p.power != 2) return fail(CAC_ERR_INVALID, "power must be 1 or 2");
    if (p.n_mels < 0 || p.n_mels > kMaxMels) return fail(CAC_ERR_INVALID, "n_mels out of range");
    if (p.n_mels > 0 && !h_mel_weights) return fail(CAC_ERR_INVALID, "mel weights missing");
    const int bins = 1 + p.n_fft / 2;
    if (p.n_mels == 0 && (p.bin_lo < 0 || p.bin_hi > bins || p.bin_lo >= p.bin_hi))
        return fail(CAC_ERR_INVALID, "bin range out of bounds");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1)
        return fail(CAC_ERR_NO_DEVICE, "no CUDA device (this library has no CPU path)");
    if (device < 0 || device >= ndev) return fail(CAC_ERR_INVALID, "device index out of range");
    CUDA_TRY(cudaSetDevice(device));

    cac_spec_plan* plan = new cac_spec_plan();
    plan->p = p;
    plan->device = device;
    if (p.n_mels > 0) {
        const std::string berr = plan->mb.build(h_mel_weights, p.n_mels, bins);
        if (!berr.empty()) { delete plan; return fail(CAC_ERR_INVALID, berr); }
        if (plan->mb.n_groups > 16) { delete plan; return fail(CAC_ERR_INVALID, "n_mels > 512 is not supported"); }
        plan->kmin = plan->mb.kmin;
        plan->nbins = plan->mb.kmax - plan->mb.kmin + 1;
    } else {
        plan->kmin = p.bin_lo;
        plan->nbins = p.bin_hi - p.bin_lo;
    }
    if (dft) {
        plan->dft = true;
        std::vector<float2> tw(p.n_fft);
        band_dft_host_table(p.n_fft, tw.data());
        cudaError_t e = cudaMalloc((void**)&plan->d_tw, tw.size() * sizeof(float2));
        if (e == cudaSuccess) e = cudaMemcpy(plan->d_tw, tw.data(), tw.size() * sizeof(float2), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) {
            cudaFree(plan->d_tw);
            delete plan;
            return fail(CAC_ERR_ALLOC, std::string("could not upload the DFT table: ") + cudaGetErrorString(e));
        }
        if (int rc = setup_band_gemm(plan)) { cac_spec_plan_destroy(plan); return rc; }
        *out = plan;
        return CAC_OK;
    }
    SpecArgs probe{};
    probe.n_fft = p.n_fft;
    probe.n_mels = p.n_mels;
    probe.n_taps = (int)plan->mb.wpad.size();
    probe.pbuf_len = plan->mb.pbuf_len;
    if (spec4800_smem_bytes(probe) > 227 * 1024) {
        delete plan;
        return fail(CAC_ERR_INVALID, "filterbank exceeds the shared-memory budget of the 4800-point kernel");
    }
    std::vector<float2> lt(kSpec4800LtabRows * 32), rt(kSpec4800Rounds);
    spec4800_host_tables(p.n_fft, plan->kmin, lt.data(), rt.data());
    cudaError_t e = cudaMalloc((void**)&plan->d_ltab, lt.size() * sizeof(float2));
    if (e == cudaSuccess) e = cudaMalloc((void**)&plan->d_rtab, rt.size() * sizeof(float2));
    if (e == cudaSuccess) e = cudaMemcpy(plan->d_ltab, lt.data(), lt.size() * sizeof(float2), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(plan->d_rtab, rt.data(), rt.size() * sizeof(float2), cudaMemcpyHostToDevice);
    if (e == cudaSuccess && p.n_mels > 0) {
        e = cudaMalloc((void**)&plan->d_wpad, plan->mb.wpad.size() * sizeof(float) + 16);
        if (e == cudaSuccess) e = cudaMalloc((void**)&plan->d_mstart, plan->mb.mstart.size() * sizeof(int) + 16);
        if (e == cudaSuccess)
            e = cudaMemcpy(plan->d_wpad, plan->mb.wpad.data(), plan->mb.wpad.size() * sizeof(float), cudaMemcpyHostToDevice);
        if (e == cudaSuccess)
            e = cudaMemcpy(plan->d_mstart, plan->mb.mstart.data(), plan->mb.mstart.size() * sizeof(int), cudaMemcpyHostToDevice);
    }
    if (e != cudaSuccess) {
        cudaFree(plan->d_ltab); cudaFree(plan->d_rtab); cudaFree(plan->d_wpad); cudaFree(plan->d_mstart);
        delete plan;
        return fail(CAC_ERR_ALLOC, std::string("could not upload the spectrum tables: ") + cudaGetErrorString(e));
    }
    if (int rc = setup_band_gemm(plan)) { cac_spec_plan_destroy(plan); return rc; }
    *out = plan;
    return CAC_OK;
}

extern "C" int cac_spec_plan_destroy(cac_spec_plan* plan) {
    if (!plan) return CAC_OK;
    cudaSetDevice(plan->device);
    cudaFree(plan->d_ltab); cudaFree(plan->d_rtab); cudaFree(plan->d_wpad); cudaFree(plan->d_mstart);
    cudaFree(plan->d_tab); cudaFree(plan->d_wtab); cudaFree(plan->d_tw); cudaFree(plan->d_bsw); cudaFree(plan->d_P);
    delete plan;
    return CAC_OK;
}

extern "C" int64_t cac_spec_frame_count(const cac_spec_plan* plan, int64_t n_samples) {
    if (!plan || n_samples < 0) return CAC_ERR_INVALID;
    return frames_of(plan, n_samples);
}

extern "C" int32_t cac_spec_rows(const cac_spec_plan* plan) {
    if (!plan) return CAC_ERR_INVALID;
    return plan->p.n_mels > 0 ? plan->p.n_mels : plan->nbins;
}

extern "C" int cac_spec_run(cac_spec_plan* plan, const float* d_pcm, const int64_t* h_item_offset,
                            const int64_t* h_item_samples, int64_t n_items, float* d_out, float* d_max, void* stream) {
    if (!plan || !d_pcm || !h_item_offset || !h_item_samples || !d_out) return fail(CAC_ERR_INVALID, "null argument");
    if (n_items < 0 || n_items > 0x7fffffffLL) return fail(CAC_ERR_INVALID, "n_items out of range");
    if (n_items == 0) return CAC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    CUDA_TRY(cudaSetDevice(plan->device));
    std::vector<long long> tab(3 * (size_t)n_items + 1);          // [item_off | item_len | frame_base]
    long long total = 0;
    // the GEMM path copies 16 bytes at a time: every recording must start on a 16-byte boundary
    // (CAC_BAND_GEMM=0 keeps the mixed-radix / direct kernels: A/B runs and the parity test between the two paths)
    const char* gemm_env = getenv("CAC_BAND_GEMM");
    bool gemm = plan->gemm && !(gemm_env && gemm_env[0] == '0') && ((uintptr_t)d_pcm & 15) == 0;
    // uniform batch (equal lengths, a whole number of hops, back to back): the GEMM reads it through one TMA tensor map
    const long long len0 = h_item_samples[0], hop_l = plan->p.hop_length;
    bool uniform = len0 >= hop_l && len0 % hop_l == 0 && len0 / hop_l < (1LL << 30);
    for (int64_t i = 0; i < n_items; ++i) {
        if (h_item_offset[i] < 0 || h_item_samples[i] < 0) return fail(CAC_ERR_INVALID, "negative item offset/length");
        if (h_item_offset[i] & 3) gemm = false;
        if (h_item_samples[i] != len0 || h_item_offset[i] != h_item_offset[0] + i * len0) uniform = false;
        tab[i] = h_item_offset[i];
        tab[n_items + i] = h_item_samples[i];
        tab[2 * n_items + i] = total;
        total += frames_of(plan, h_item_samples[i]);
    }
    tab[3 * n_items] = total;
    int rc = ensure_cap(&plan->d_tab, &plan->tab_cap, tab.size() * sizeof(long long));
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(plan->d_tab, tab.data(), tab.size() * sizeof(long long), cudaMemcpyHostToDevice, st));
    if (d_max) CUDA_TRY(cudaMemsetAsync(d_max, 0, sizeof(float), st));
    SpecArgs a{};
    a.pcm = d_pcm;
    a.item_off = plan->d_tab;
    a.item_len = plan->d_tab + n_items;
    a.frame_base = plan->d_tab + 2 * n_items;
    a.n_items = (int)n_items;
    a.n_frames_total = total;
    a.n_fft = plan->p.n_fft;
    a.hop = plan->p.hop_length;
    a.power = plan->p.power;
    a.kmin = plan->kmin;
    a.nbins = plan->nbins;
    a.n_mels = plan->p.n_mels;
    a.bin_major = plan->p.bin_major;
    a.out = d_out;
    a.gmax = d_max;
    a.ltab = plan->d_ltab;
    a.rtab = plan->d_rtab;
    a.wpad = plan->d_wpad;
    a.mstart = plan->d_mstart;
    a.n_taps = (int)plan->mb.wpad.size();
    a.n_groups = plan->mb.n_groups;
    a.pbuf_len = plan->mb.pbuf_len;
    for (int g = 0; g < 16; ++g) { a.gmax_taps[g] = plan->mb.gmax[g]; a.goff[g] = plan->mb.goff[g]; }
    if (gemm) {
        rc = ensure_cap(&plan->d_P, &plan->P_cap, band_gemm_scratch_bytes(total));
        if (rc) return rc;
        CUDA_TRY(launch_band_gemm(a, d_pcm + h_item_offset[0], uniform ? (int)(len0 / hop_l) : 0, plan->d_bsw, plan->d_P, st));
    } else if (plan->dft) {
        CUDA_TRY(launch_band_dft(a, plan->d_tw, st));
    } else {
        CUDA_TRY(launch_spec4800(a, st));
    }
    return CAC_OK;
}

extern "C" int cac_chunk_minmax_run(const float* d_in, int32_t n_rows, int64_t n_cols, int32_t chunk, float* d_out,
                                    void* stream) {
    if (!d_in || !d_out || n_rows < 1 || n_cols < 0 || chunk < 1) return fail(CAC_ERR_INVALID, "bad argument");
    const long long n_chunks = n_cols / chunk;
    if (n_chunks == 0) return CAC_OK;
    CUDA_TRY(launch_chunk_minmax(d_in, n_rows, n_cols, chunk, n_chunks, reinterpret_cast<float2*>(d_out),
                                 (cudaStream_t)stream));
    return CAC_OK;
}

extern "C" int cac_scale_by_max_run(float* d_data, int64_t n, const float* d_max, void* stream) {
    if (!d_data || !d_max || n < 0) return fail(CAC_ERR_INVALID, "bad argument");
    if (n == 0) return CAC_OK;
    CUDA_TRY(launch_scale_by_max(d_data, n, d_max, (cudaStream_t)stream));
    return CAC_OK;
}

extern "C" int cac_morepork_windows_run(cac_spec_plan* plan, const float* d_spec, const int64_t* h_win_frame0,
                                        const int32_t* h_win_width, int64_t n_windows, float* d_out, void* stream) {
    if (!plan || !d_spec || !h_win_frame0 || !h_win_width || !d_out) return fail(CAC_ERR_INVALID, "null argument");
    if (plan->p.n_mels != 0 || plan->p.bin_major) return fail(CAC_ERR_INVALID, "needs a frame-major raw-band plan");
    if (n_windows < 0) return fail(CAC_ERR_INVALID, "n_windows out of range");
    if (n_windows == 0) return CAC_OK;
    for (int64_t i = 0; i < n_windows; ++i)
        if (h_win_width[i] < 0 || h_win_width[i] > 60 || h_win_frame0[i] < 0)
            return fail(CAC_ERR_INVALID, "window out of range (width must be 0..60)");
    cudaStream_t st = (cudaStream_t)stream;
    CUDA_TRY(cudaSetDevice(plan->device));
    int rc = ensure_cap(&plan->d_wtab, &plan->wtab_cap, (size_t)n_windows * (sizeof(long long) + sizeof(int)));
    if (rc) return rc;
    int* d_width = reinterpret_cast<int*>(plan->d_wtab + n_windows);
    CUDA_TRY(cudaMemcpyAsync(plan->d_wtab, h_win_frame0, (size_t)n_windows * sizeof(long long), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(d_width, h_win_width, (size_t)n_windows * sizeof(int), cudaMemcpyHostToDevice, st));
    CUDA_TRY(launch_morepork_windows(d_spec, plan->d_wtab, d_width, n_windows, plan->nbins, d_out, st));
    return CAC_OK;
}

// ---- identify_tracks.get_end for a batch (device item tables, chunk scan on the device) -----------------------------
namespace {
struct GetEndLayout { size_t tab, mm, list, list_n, total; int chunk_stride; };
GetEndLayout get_end_layout(const cac_spec_plan* plan, int64_t n_items, int64_t max_item_samples, int chunk_frames) {
    GetEndLayout L;
    const long long max_frames = frames_of(plan, max_item_samples);
    L.chunk_stride = (int)(max_frames / chunk_frames + 1);
    size_t o = 0;
    L.tab = o; o += ((3 * (size_t)n_items + 1) * sizeof(long long) + 255) / 256 * 256;
    L.mm = o; o += ((size_t)n_items * L.chunk_stride * sizeof(uint2) + 255) / 256 * 256;
    L.list = o; o += ((size_t)n_items * L.chunk_stride * sizeof(int2) + 255) / 256 * 256;    // candidate (item, chunk) pairs
    L.list_n = o; o += 256;
    L.total = o;
    return L;
}
}  // namespace

extern "C" size_t cac_get_end_work_bytes(const cac_spec_plan* plan, int64_t n_items, int64_t max_item_samples,
                                         int32_t chunk_frames) {
    if (!plan || n_items < 1 || max_item_samples < 0 || chunk_frames < 1) return 0;
    return get_end_layout(plan, n_items, max_item_samples, chunk_frames).total;
}

extern "C" int cac_get_end_run(cac_spec_plan* plan, const float* d_pcm, const int64_t* d_item_offset,
                               const int64_t* d_item_samples, int64_t n_items, int64_t max_item_samples, int32_t chunk_frames,
                               int32_t* d_out, void* d_work, size_t work_bytes, void* stream) {
    if (!plan || !d_pcm || !d_item_offset || !d_item_samples || !d_out || !d_work) return fail(CAC_ERR_INVALID, "null argument");
    if (plan->p.n_mels < 1 || plan->dft) return fail(CAC_ERR_INVALID, "get_end needs a mel plan of the mixed-radix kernel");
    if (n_items < 0 || n_items > 0x7fffffffLL || max_item_samples < 0 || chunk_frames < 1)
        return fail(CAC_ERR_INVALID, "n_items / max_item_samples / chunk_frames out of range");
    if (n_items == 0) return CAC_OK;
    const GetEndLayout L = get_end_layout(plan, n_items, max_item_samples, chunk_frames);
    if (work_bytes < L.total || ((uintptr_t)d_work & 255) != 0)
        return fail(CAC_ERR_INVALID, "work buffer too small or not 256-byte aligned (see cac_get_end_work_bytes)");
    cudaStream_t st = (cudaStream_t)stream;
    CUDA_TRY(cudaSetDevice(plan->device));
    long long* tab = reinterpret_cast<long long*>((char*)d_work + L.tab);
    uint2* mm = reinterpret_cast<uint2*>((char*)d_work + L.mm);
    CUDA_TRY(launch_get_end_prep(reinterpret_cast<const long long*>(d_item_offset),
                                 reinterpret_cast<const long long*>(d_item_samples), (int)n_items, plan->p.n_fft,
                                 plan->p.hop_length, max_item_samples, tab, mm, L.chunk_stride, d_out, st));
    SpecArgs a{};
    a.pcm = d_pcm;
    a.item_off = tab;
    a.item_len = tab + n_items;
    a.frame_base = tab + 2 * n_items;
    a.n_items = (int)n_items;
    a.n_frames_total = -1;                    // on the device: tab[3 n]
    a.n_fft = plan->p.n_fft;
    a.hop = plan->p.hop_length;
    a.power = plan->p.power;
    a.kmin = plan->kmin;
    a.nbins = plan->nbins;
    a.n_mels = plan->p.n_mels;
    a.bin_major = 0;
    a.out = nullptr;                          // the mel values are consumed in the epilogue
    a.gmax = nullptr;
    a.chunk_mm = mm;
    a.chunk_frames = chunk_frames;
    a.chunk_stride = L.chunk_stride;
    // Phase 1: ONE frame per chunk.  A chunk whose first frame already holds two different mel values cannot be the
    // constant chunk the reference looks for (np.amax(data) == np.amin(data), :408); only the chunks whose first frame is
    // constant - digital silence, in practice - are transformed in full (phase 2).  Same values, same answer, 1 / 170 of
    // the frames for a recording without a silent tail.
    int2* list = reinterpret_cast<int2*>((char*)d_work + L.list);
    int* list_n = reinterpret_cast<int*>((char*)d_work + L.list_n);
    CUDA_TRY(cudaMemsetAsync(list_n, 0, sizeof(int), st));
    a.chunk_probe = 1;
    a.n_frames_total = (long long)n_items * L.chunk_stride;
    a.ltab = plan->d_ltab;
    a.rtab = plan->d_rtab;
    a.wpad = plan->d_wpad;
    a.mstart = plan->d_mstart;
    a.n_taps = (int)plan->mb.wpad.size();
    a.n_groups = plan->mb.n_groups;
    a.pbuf_len = plan->mb.pbuf_len;
    for (int g = 0; g < 16; ++g) { a.gmax_taps[g] = plan->mb.gmax[g]; a.goff[g] = plan->mb.goff[g]; }
    CUDA_TRY(launch_spec4800(a, st));
    CUDA_TRY(launch_get_end_select(tab, (int)n_items, mm, L.chunk_stride, chunk_frames, list, list_n, st));
    a.chunk_probe = 0;
    a.chunk_list = list;
    a.chunk_list_n = list_n;
    a.n_frames_total = -1;                    // on the device: *list_n chunks
    CUDA_TRY(launch_spec4800(a, st));
    CUDA_TRY(launch_get_end_scan(tab, (int)n_items, mm, L.chunk_stride, chunk_frames, d_out, st));
    return CAC_OK;
}

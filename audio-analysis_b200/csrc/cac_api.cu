// C-ABI of libcacophony_b200.so (see include/cacophony_b200.h for the contract and the
// reference callables each entry point replaces).  No CPU fallback: without a CUDA device
// every compute entry point fails with CAC_ERR_NO_DEVICE / CAC_ERR_CUDA.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "api_util.h"
#include "cac_common.cuh"
#include "mel_bands.h"
#include "spec.cuh"

namespace cac {
cudaError_t launch_segment_rows(const float* pcm, const cac_segment* segs, long long n, float* rows, int seg_len,
                                int n_frames, int normalize, long long* tab, cudaStream_t st);
size_t feat_smem_bytes(const FeatArgs& a);
cudaError_t launch_features(const FeatArgs& a, int power, int db_scale, cudaStream_t st, cudaEvent_t ev0,
                            cudaEvent_t ev1);
cudaError_t launch_normalize(const float* in, float* out, long long rows, int row_len, cudaStream_t st);
cudaError_t launch_mel_project(const float* mag, float* out, int n_frames, int n_mels, int power,
                               const BandTable* d_bands, const float* d_weights, cudaStream_t st);
cudaError_t launch_power_to_db(const float* in, float* out, long long n_items, long long item_len, unsigned* work,
                               cudaStream_t st);
}  // namespace cac

using namespace cac;

static thread_local std::string g_err;

int cac::fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}

struct cac_plan {
    cac_params p;
    int device;
    int n_frames;
    MelBands mb;                     // banded filterbank (mel_bands.h), host copy
    BandTable* d_bt = nullptr;       // the plan's own device copy of the band table and its taps (cac_mel_project_run)
    float* d_bw = nullptr;
    int frames_per_stage, stage_half;
    float* d_wpad = nullptr;
    int* d_mstart = nullptr;
    std::string fused_unsupported;   // why the fused kernel cannot serve these parameters ("" = it can)
    // general path (fused_unsupported set, n_fft one of the framed-spectrum kernel's lengths): segment rows ->
    // spec4800_kernel with mel epilogue -> power_to_db
    bool general = false;
    MelBands gmb;                    // band table relative to the lowest bin (stft4800.cu convention)
    float2* d_ltab = nullptr;
    float2* d_rtab = nullptr;
    float* d_gwpad = nullptr;
    int* d_gmstart = nullptr;
    // buffers owned by the *_host entry points
    float* d_pcm = nullptr;      size_t pcm_cap = 0;
    short* d_pcm16 = nullptr;    size_t pcm16_cap = 0;   // staging of cac_features_host_pcm16
    cac_segment* d_segs = nullptr; size_t segs_cap = 0;
    float* d_out = nullptr;      size_t out_cap = 0;
    void* d_work = nullptr;      size_t work_cap = 0;
    // MagTransform epilogue (cac_plan_set_magtransform)
    int mag_on = 0;
    float mag_a = 0.f;
    // optional CUDA events bracketing the dominant (stft_mel) kernel of the last cac_features_run
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // cac_features_host pipeline: upload / compute / download streams and per-chunk events
    cudaStream_t s_h2d = nullptr, s_comp = nullptr, s_d2h = nullptr;
    std::vector<cudaEvent_t> ev_up, ev_done;
};

extern "C" const char* cac_version(void) { return "cacophony_b200 0.1.0 (sm_100a)"; }
extern "C" const char* cac_last_error(void) { return g_err.c_str(); }

extern "C" int cac_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) return fail(CAC_ERR_NO_DEVICE, std::string("cudaGetDeviceCount: ") + cudaGetErrorString(e));
    return n;
}

extern "C" int cac_plan_destroy(cac_plan* plan);

extern "C" int cac_plan_create(const cac_params* params, const float* w, int device, cac_plan** out) {
    if (!params || !w || !out) return fail(CAC_ERR_INVALID, "null argument");
    const cac_params& p = *params;
    if (p.n_fft < 8 || (p.n_fft & 1)) return fail(CAC_ERR_INVALID, "n_fft must be even and >= 8");
    if (p.hop_length < 1 || p.hop_length > p.n_fft) return fail(CAC_ERR_INVALID, "hop_length out of range");
    if (p.n_mels < 1 || p.n_mels > kMaxMels) return fail(CAC_ERR_INVALID, "n_mels out of range");
    if (p.power != 1 && p.power != 2) return fail(CAC_ERR_INVALID, "power must be 1 or 2");
    if (p.segment_samples < 1) return fail(CAC_ERR_INVALID, "segment_samples must be positive");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1)
        return fail(CAC_ERR_NO_DEVICE, "no CUDA device (this library has no CPU path)");
    if (device < 0 || device >= ndev) return fail(CAC_ERR_INVALID, "device index out of range");

    cac_plan* plan = new cac_plan();
    plan->p = p;
    plan->device = device;
    plan->n_frames = 1 + p.segment_samples / p.hop_length;
    const int bins = 1 + p.n_fft / 2;
    const std::string berr = plan->mb.build(w, p.n_mels, bins, /*absolute=*/true);
    if (!berr.empty()) { delete plan; return fail(CAC_ERR_INVALID, berr); }
    const int kmin = plan->mb.kmin, kmax = plan->mb.kmax;
    if (p.n_fft != 4096) plan->fused_unsupported = "fused STFT->mel kernel supports n_fft == 4096 only";
    else if (kmax >= 1024) plan->fused_unsupported = "fused STFT->mel kernel covers FFT bins < 1024 (fmax < sr/4)";
    if (plan->mb.n_groups > 16) plan->fused_unsupported = "fused kernel supports n_mels <= 512";
    if (plan->mb.pbuf_len > 2 * 32 * 33 && plan->fused_unsupported.empty())
        plan->fused_unsupported = "band range exceeds the per-warp power-spectrum buffer of the fused kernel";
    // frames staged per round: a multiple of the 12 warps per CTA when it fits the 227 KB of shared
    // memory (hop 281 -> 12), else the largest smaller count (some warps then idle during the FFTs)
    plan->frames_per_stage = 0;
    static const int kFsChoices[] = {24, 12, 10, 8, 6, 4, 3, 2, 1};
    for (int fs : kFsChoices) {
        int half = stage_half_floats(fs, p.hop_length);
        FeatArgs a{};
        a.stage_half = half;
        a.nbins = kmax - kmin + 1;
        a.n_taps = (int)plan->mb.wpad.size();
        a.pbuf_len = plan->mb.pbuf_len;
        a.n_mels = p.n_mels;
        a.frames_per_stage = fs;
        if (feat_smem_bytes(a) <= 227 * 1024) { plan->frames_per_stage = fs; plan->stage_half = half; break; }
    }
    if (!plan->frames_per_stage && plan->fused_unsupported.empty())
        plan->fused_unsupported = "parameters exceed the shared-memory budget of the fused kernel";
    if (!plan->fused_unsupported.empty() && spec4800_supported(p.n_fft)) {
        const std::string gerr = plan->gmb.build(w, p.n_mels, bins);
        SpecArgs probe{};
        probe.n_fft = p.n_fft;
        probe.n_mels = p.n_mels;
        probe.n_taps = (int)plan->gmb.wpad.size();
        probe.pbuf_len = plan->gmb.pbuf_len;
        if (gerr.empty() && plan->gmb.n_groups <= 16 && spec4800_smem_bytes(probe) <= 227 * 1024) {
            cudaSetDevice(device);
            std::vector<float2> lt(kSpec4800LtabRows * 32), rt(kSpec4800Rounds);
            spec4800_host_tables(p.n_fft, plan->gmb.kmin, lt.data(), rt.data());
            cudaError_t e = cudaMalloc((void**)&plan->d_ltab, lt.size() * sizeof(float2));
            if (e == cudaSuccess) e = cudaMalloc((void**)&plan->d_rtab, rt.size() * sizeof(float2));
            if (e == cudaSuccess) e = cudaMalloc((void**)&plan->d_gwpad, plan->gmb.wpad.size() * sizeof(float) + 16);
            if (e == cudaSuccess) e = cudaMalloc((void**)&plan->d_gmstart, plan->gmb.mstart.size() * sizeof(int) + 16);
            if (e == cudaSuccess) e = cudaMemcpy(plan->d_ltab, lt.data(), lt.size() * sizeof(float2), cudaMemcpyHostToDevice);
            if (e == cudaSuccess) e = cudaMemcpy(plan->d_rtab, rt.data(), rt.size() * sizeof(float2), cudaMemcpyHostToDevice);
            if (e == cudaSuccess)
                e = cudaMemcpy(plan->d_gwpad, plan->gmb.wpad.data(), plan->gmb.wpad.size() * sizeof(float), cudaMemcpyHostToDevice);
            if (e == cudaSuccess)
                e = cudaMemcpy(plan->d_gmstart, plan->gmb.mstart.data(), plan->gmb.mstart.size() * sizeof(int), cudaMemcpyHostToDevice);
            if (e != cudaSuccess) {
                cudaFree(plan->d_ltab); cudaFree(plan->d_rtab); cudaFree(plan->d_gwpad); cudaFree(plan->d_gmstart);
                delete plan;
                return fail(CAC_ERR_ALLOC, "could not upload the spectrum tables");
            }
            plan->general = true;
        }
    }
    if (plan->fused_unsupported.empty()) {
        cudaSetDevice(device);
        cudaError_t e1 = cudaMalloc((void**)&plan->d_wpad, plan->mb.wpad.size() * sizeof(float) + 16);
        // slot table of the fused kernel, one int4 per (group, lane): x = first bin of the band (multiple of 4),
        // y = offset of the mel row in the output tile (-1: dead slot), z = offset of the lane's first weight quad,
        // w = quads per row of this group - everything a lane needs for one group in a single 128-bit load
        std::vector<int> slots(plan->mb.mstart.size() * 4);
        for (size_t i = 0; i < plan->mb.mstart.size(); ++i) {
            const int g = (int)(i / 32), lane = (int)(i % 32);
            const unsigned packed = (unsigned)plan->mb.mstart[i];
            const int row = (int)(packed >> 16);
            slots[4 * i + 0] = (int)(packed & 0xFFFFu);
            slots[4 * i + 1] = row == 0xFFFF ? -1 : row * (plan->frames_per_stage + 1);
            slots[4 * i + 2] = (g < 16 ? plan->mb.goff[g] : 0) + lane * 4;
            slots[4 * i + 3] = (g < 16 ? plan->mb.gmax[g] : 0) >> 2;
        }
        cudaError_t e2 = cudaMalloc((void**)&plan->d_mstart, slots.size() * sizeof(int) + 16);
        if (e1 == cudaSuccess && e2 == cudaSuccess) {
            // the fused kernel leaves out the untangle's factor 1/2 (power 2: 1/4): exact powers of two, folded in here
            std::vector<float> wscaled(plan->mb.wpad);
            const float fold = p.power == 2 ? 0.25f : 0.5f;
            for (float& v : wscaled) v *= fold;
            e1 = cudaMemcpy(plan->d_wpad, wscaled.data(), wscaled.size() * sizeof(float), cudaMemcpyHostToDevice);
            e2 = cudaMemcpy(plan->d_mstart, slots.data(), slots.size() * sizeof(int), cudaMemcpyHostToDevice);
        }
        if (e1 != cudaSuccess || e2 != cudaSuccess) {
            cudaFree(plan->d_wpad); cudaFree(plan->d_mstart);
            delete plan;
            return fail(CAC_ERR_ALLOC, "could not upload the filterbank tables");
        }
    }
    {   // band table of cac_mel_project_run: owned by the plan (no process-wide __constant__ state)
        cudaSetDevice(device);
        const size_t wbytes = std::max<size_t>(plan->mb.weights.size(), 1) * sizeof(float);
        cudaError_t e = cudaMalloc((void**)&plan->d_bt, sizeof(BandTable));
        if (e == cudaSuccess) e = cudaMalloc((void**)&plan->d_bw, wbytes);
        if (e == cudaSuccess) e = cudaMemcpy(plan->d_bt, &plan->mb.bt, sizeof(BandTable), cudaMemcpyHostToDevice);
        if (e == cudaSuccess && !plan->mb.weights.empty())
            e = cudaMemcpy(plan->d_bw, plan->mb.weights.data(), plan->mb.weights.size() * sizeof(float), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) {
            cac_plan_destroy(plan);
            return fail(CAC_ERR_ALLOC, "could not upload the band table");
        }
    }
    *out = plan;
    return CAC_OK;
}

extern "C" int cac_plan_destroy(cac_plan* plan) {
    if (!plan) return CAC_OK;
    cudaSetDevice(plan->device);
    if (plan->ev0) { cudaEventDestroy(plan->ev0); cudaEventDestroy(plan->ev1); }
    for (cudaEvent_t e : plan->ev_up) cudaEventDestroy(e);
    for (cudaEvent_t e : plan->ev_done) cudaEventDestroy(e);
    if (plan->s_h2d) { cudaStreamDestroy(plan->s_h2d); cudaStreamDestroy(plan->s_comp); cudaStreamDestroy(plan->s_d2h); }
    cudaFree(plan->d_wpad); cudaFree(plan->d_mstart); cudaFree(plan->d_bt); cudaFree(plan->d_bw);
    cudaFree(plan->d_ltab); cudaFree(plan->d_rtab); cudaFree(plan->d_gwpad); cudaFree(plan->d_gmstart);
    cudaFree(plan->d_pcm); cudaFree(plan->d_pcm16); cudaFree(plan->d_segs); cudaFree(plan->d_out); cudaFree(plan->d_work);
    delete plan;
    return CAC_OK;
}

extern "C" int cac_plan_set_profiling(cac_plan* plan, int enable) {
    if (!plan) return fail(CAC_ERR_INVALID, "null plan");
    CUDA_TRY(cudaSetDevice(plan->device));
    if (enable && !plan->ev0) {
        CUDA_TRY(cudaEventCreate(&plan->ev0));
        CUDA_TRY(cudaEventCreate(&plan->ev1));
    } else if (!enable && plan->ev0) {
        cudaEventDestroy(plan->ev0); cudaEventDestroy(plan->ev1);
        plan->ev0 = plan->ev1 = nullptr;
    }
    return CAC_OK;
}

extern "C" int cac_plan_last_kernel_ms(cac_plan* plan, float* ms) {
    if (!plan || !ms) return fail(CAC_ERR_INVALID, "null argument");
    if (!plan->ev0) return fail(CAC_ERR_INVALID, "profiling is not enabled on this plan");
    CUDA_TRY(cudaEventSynchronize(plan->ev1));
    CUDA_TRY(cudaEventElapsedTime(ms, plan->ev0, plan->ev1));
    return CAC_OK;
}

extern "C" int cac_plan_set_magtransform(cac_plan* plan, int enable, float a) {
    if (!plan) return fail(CAC_ERR_INVALID, "null plan");
    if (enable && plan->p.db_scale)
        return fail(CAC_ERR_INVALID, "MagTransform replaces the dB scale: the plan must be created with db_scale = 0");
    if (enable && (!plan->fused_unsupported.empty()))
        return fail(CAC_ERR_INVALID, "the MagTransform epilogue exists in the fused n_fft 4096 kernel only; use cac_magtransform_run");
    plan->mag_on = enable ? 1 : 0;
    plan->mag_a = a;
    return CAC_OK;
}

extern "C" int cac_plan_frames(const cac_plan* plan) { return plan ? plan->n_frames : CAC_ERR_INVALID; }

static size_t align128(size_t v) { return (v + 127) / 128 * 128; }

extern "C" size_t cac_features_work_bytes(const cac_plan* plan, int64_t n) {
    const size_t nn = (size_t)(n > 0 ? n : 0);
    size_t bytes = nn * (sizeof(float2) + sizeof(unsigned)) + 256;
    if (plan && plan->general)   // item tables + the zero-padded, normalised segments as rows
        bytes = align128(bytes) + align128((3 * nn + 1) * sizeof(long long)) +
                nn * (size_t)plan->p.segment_samples * sizeof(float) + 128;
    return bytes;
}

// general path of cac_features_run (see cac_plan::general)
static int features_run_general(cac_plan* plan, const float* d_pcm, const cac_segment* d_segs, int64_t n, float* d_out,
                                void* d_work, cudaStream_t st) {
    const size_t nn = (size_t)n;
    unsigned* segmax = (unsigned*)((char*)d_work + ((nn * sizeof(float2) + 127) / 128) * 128);   // same slot as the fused path
    char* base = (char*)d_work + align128(nn * (sizeof(float2) + sizeof(unsigned)) + 256);
    long long* tab = (long long*)base;
    float* rows = (float*)(base + align128((3 * nn + 1) * sizeof(long long)));
    CUDA_TRY(launch_segment_rows(d_pcm, d_segs, n, rows, plan->p.segment_samples, plan->n_frames, plan->p.normalize, tab, st));
    SpecArgs a{};
    a.pcm = rows;
    a.item_off = tab;
    a.item_len = tab + n;
    a.frame_base = tab + 2 * n;
    a.n_items = (int)n;
    a.n_frames_total = (long long)n * plan->n_frames;
    a.n_fft = plan->p.n_fft;
    a.hop = plan->p.hop_length;
    a.power = plan->p.power;
    a.kmin = plan->gmb.kmin;
    a.nbins = plan->gmb.kmax - plan->gmb.kmin + 1;
    a.n_mels = plan->p.n_mels;
    a.bin_major = 0;
    a.out = d_out;
    a.gmax = nullptr;
    a.ltab = plan->d_ltab;
    a.rtab = plan->d_rtab;
    a.wpad = plan->d_gwpad;
    a.mstart = plan->d_gmstart;
    a.n_taps = (int)plan->gmb.wpad.size();
    a.n_groups = plan->gmb.n_groups;
    a.pbuf_len = plan->gmb.pbuf_len;
    for (int g = 0; g < 16; ++g) { a.gmax_taps[g] = plan->gmb.gmax[g]; a.goff[g] = plan->gmb.goff[g]; }
    CUDA_TRY(launch_spec4800(a, st));
    if (plan->p.db_scale)
        CUDA_TRY(launch_power_to_db(d_out, d_out, n, (long long)plan->p.n_mels * plan->n_frames, segmax, st));
    return CAC_OK;
}

extern "C" int cac_features_run(cac_plan* plan, const float* d_pcm, const cac_segment* d_segs, int64_t n,
                                float* d_out, void* d_work, void* stream) {
    if (!plan || !d_pcm || !d_segs || !d_out || !d_work) return fail(CAC_ERR_INVALID, "null argument");
    if (!plan->fused_unsupported.empty() && !plan->general)
        return fail(CAC_ERR_INVALID, plan->fused_unsupported + "; the general path needs n_fft 4800 / 3200 / 1600 / "
                                                               "4096 / 2048 / 1024 / 512");
    if (n == 0) return CAC_OK;
    if (((uintptr_t)d_pcm & 15) != 0) return fail(CAC_ERR_INVALID, "d_pcm must be 16-byte aligned");
    if (n < 0 || n > 0x7fffffffLL) return fail(CAC_ERR_INVALID, "n_segments out of range");
    cudaStream_t st = (cudaStream_t)stream;
    CUDA_TRY(cudaSetDevice(plan->device));
    if (plan->general) return features_run_general(plan, d_pcm, d_segs, n, d_out, d_work, st);
    FeatArgs a{};
    a.pcm = d_pcm; a.segs = d_segs; a.out = d_out;
    a.segstat = (float2*)d_work;
    a.segmax = (unsigned*)((char*)d_work + (((size_t)n * sizeof(float2) + 127) / 128) * 128);
    a.n_segs = n;
    a.seg_len = plan->p.segment_samples;
    a.hop = plan->p.hop_length;
    a.n_frames = plan->n_frames;
    a.n_mels = plan->p.n_mels;
    a.kmin = plan->mb.kmin;
    a.nbins = plan->mb.kmax - plan->mb.kmin + 1;
    a.frames_per_stage = plan->frames_per_stage;
    a.stages_per_seg = (plan->n_frames + plan->frames_per_stage - 1) / plan->frames_per_stage;
    a.n_taps = (int)plan->mb.wpad.size();
    a.wpad = plan->d_wpad;
    a.mstart = plan->d_mstart;
    a.n_groups = plan->mb.n_groups;
    a.pbuf_len = plan->mb.pbuf_len;
    for (int g = 0; g < 16; ++g) { a.gmax[g] = plan->mb.gmax[g]; a.goff[g] = plan->mb.goff[g]; }
    a.stage_half = plan->stage_half;
    a.normalize = plan->p.normalize;
    a.mag_on = plan->mag_on;
    a.mag_p = 1.0f / (1.0f + expf(-plan->mag_a));            // tf.math.sigmoid in float32
    CUDA_TRY(launch_features(a, plan->p.power, plan->p.db_scale, st, plan->ev0, plan->ev1));
    return CAC_OK;
}

namespace cac {
// 16-bit PCM -> float32 exactly as the decoders the reference uses do (sample / 32768)
__global__ void __launch_bounds__(256) pcm16_to_float_kernel(const short* __restrict__ in, float* __restrict__ out,
                                                             long long n) {
    const long long stride = (long long)gridDim.x * 256;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += stride)
        out[i] = (float)in[i] * (1.0f / 32768.0f);
}
}  // namespace cac

static int features_host_impl(cac_plan* plan, const void* h_pcm_any, bool pcm16, int64_t n_pcm, const cac_segment* h_segs,
                              int64_t n, float* h_out) {
    const float* h_pcm = static_cast<const float*>(h_pcm_any);
    const short* h_pcm16 = static_cast<const short*>(h_pcm_any);
    if (!plan || !h_pcm_any || !h_segs || !h_out) return fail(CAC_ERR_INVALID, "null argument");
    if (n <= 0) return CAC_OK;
    CUDA_TRY(cudaSetDevice(plan->device));
    for (int64_t i = 0; i < n; ++i) {
        const cac_segment& s = h_segs[i];
        if (s.n_valid < 0 || s.pad_left < 0 || s.pad_left + (int64_t)s.n_valid > plan->p.segment_samples ||
            s.pcm_offset < 0 || s.pcm_offset + s.n_valid > n_pcm)
            return fail(CAC_ERR_INVALID, "segment descriptor out of bounds");
    }
    const size_t item = (size_t)plan->p.n_mels * plan->n_frames;
    const size_t out_bytes = (size_t)n * item * sizeof(float);
    int rc;
    if ((rc = ensure_cap(&plan->d_pcm, &plan->pcm_cap, (size_t)n_pcm * sizeof(float) + 16))) return rc;
    if (pcm16 && (rc = ensure_cap(&plan->d_pcm16, &plan->pcm16_cap, (size_t)n_pcm * sizeof(short) + 16))) return rc;
    if ((rc = ensure_cap(&plan->d_segs, &plan->segs_cap, (size_t)n * sizeof(cac_segment)))) return rc;
    if ((rc = ensure_cap(&plan->d_out, &plan->out_cap, out_bytes))) return rc;
    // Pipeline: the PCM goes up in pieces on one stream, segment chunks are transformed on a second
    // stream as soon as the samples they reference have arrived, and every finished chunk goes back to
    // the host on a third stream - upload, kernels and download overlap (full-duplex PCIe) when the
    // host buffers are page-locked; with pageable buffers the copies serialise but stay correct.
    // chunks of ~156 segments: one wave of the persistent kernel (148 CTAs x ~23 stages) and ~50 MB per copy -
    // small enough that the first download starts ~1.5 ms into the call, large enough for full PCIe rate
    const int n_chunks = (int)std::min<int64_t>(256, std::max<int64_t>(1, (n + 77) / 156));
    const int64_t per_chunk = (n + n_chunks - 1) / n_chunks;
    const size_t work_chunk = (cac_features_work_bytes(plan, per_chunk) + 255) / 256 * 256;
    // the general path's scratch holds whole segment rows: its chunks run one after the other on the compute stream,
    // so they share one slice instead of n_chunks of them
    const size_t work_slices = plan->general ? 1 : (size_t)n_chunks;
    if ((rc = ensure_cap(&plan->d_work, &plan->work_cap, work_chunk * work_slices))) return rc;
    if (!plan->s_h2d) {
        CUDA_TRY(cudaStreamCreateWithFlags(&plan->s_h2d, cudaStreamNonBlocking));
        CUDA_TRY(cudaStreamCreateWithFlags(&plan->s_comp, cudaStreamNonBlocking));
        CUDA_TRY(cudaStreamCreateWithFlags(&plan->s_d2h, cudaStreamNonBlocking));
    }
    while ((int)plan->ev_up.size() < n_chunks) {
        cudaEvent_t a, b;
        CUDA_TRY(cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&b, cudaEventDisableTiming));
        plan->ev_up.push_back(a);
        plan->ev_done.push_back(b);
    }
    CUDA_TRY(cudaMemcpyAsync(plan->d_segs, h_segs, (size_t)n * sizeof(cac_segment), cudaMemcpyHostToDevice, plan->s_h2d));
    // upload piece c = samples up to the furthest one any segment of chunks 0..c touches
    int64_t sent = 0;
    for (int c = 0; c < n_chunks; ++c) {
        const int64_t s0 = c * per_chunk, s1 = std::min<int64_t>(n, s0 + per_chunk);
        int64_t need = sent;
        for (int64_t i = s0; i < s1; ++i) need = std::max<int64_t>(need, h_segs[i].pcm_offset + h_segs[i].n_valid);
        need = std::min<int64_t>(n_pcm, (need + 3) / 4 * 4);
        if (need > sent) {
            if (pcm16) {       // half the bytes over PCIe, widened on the device right behind the copy
                CUDA_TRY(cudaMemcpyAsync(plan->d_pcm16 + sent, h_pcm16 + sent, (size_t)(need - sent) * sizeof(short),
                                         cudaMemcpyHostToDevice, plan->s_h2d));
                const long long cnt = need - sent;
                const unsigned blocks = (unsigned)std::min<long long>((cnt + 255) / 256, 148 * 8);
                pcm16_to_float_kernel<<<blocks, 256, 0, plan->s_h2d>>>(plan->d_pcm16 + sent, plan->d_pcm + sent, cnt);
            } else {
                CUDA_TRY(cudaMemcpyAsync(plan->d_pcm + sent, h_pcm + sent, (size_t)(need - sent) * sizeof(float),
                                         cudaMemcpyHostToDevice, plan->s_h2d));
            }
            sent = need;
        }
        CUDA_TRY(cudaEventRecord(plan->ev_up[c], plan->s_h2d));
    }
    for (int c = 0; c < n_chunks; ++c) {
        const int64_t s0 = c * per_chunk, s1 = std::min<int64_t>(n, s0 + per_chunk);
        if (s1 <= s0) break;
        CUDA_TRY(cudaStreamWaitEvent(plan->s_comp, plan->ev_up[c], 0));
        if ((rc = cac_features_run(plan, plan->d_pcm, plan->d_segs + s0, s1 - s0, plan->d_out + s0 * item,
                                   (char*)plan->d_work + work_chunk * (plan->general ? 0 : c), plan->s_comp)))
            return rc;
        CUDA_TRY(cudaEventRecord(plan->ev_done[c], plan->s_comp));
        CUDA_TRY(cudaStreamWaitEvent(plan->s_d2h, plan->ev_done[c], 0));
        CUDA_TRY(cudaMemcpyAsync(h_out + s0 * item, plan->d_out + s0 * item, (size_t)(s1 - s0) * item * sizeof(float),
                                 cudaMemcpyDeviceToHost, plan->s_d2h));
    }
    CUDA_TRY(cudaStreamSynchronize(plan->s_d2h));
    CUDA_TRY(cudaStreamSynchronize(plan->s_comp));
    return CAC_OK;
}

extern "C" int cac_features_host(cac_plan* plan, const float* h_pcm, int64_t n_pcm, const cac_segment* h_segs,
                                 int64_t n, float* h_out) {
    return features_host_impl(plan, h_pcm, false, n_pcm, h_segs, n, h_out);
}

extern "C" int cac_features_host_pcm16(cac_plan* plan, const int16_t* h_pcm, int64_t n_pcm, const cac_segment* h_segs,
                                       int64_t n, float* h_out) {
    return features_host_impl(plan, h_pcm, true, n_pcm, h_segs, n, h_out);
}

extern "C" int cac_host_register(void* h_ptr, size_t bytes) {
    if (!h_ptr || bytes == 0) return fail(CAC_ERR_INVALID, "bad argument");
    CUDA_TRY(cudaHostRegister(h_ptr, bytes, cudaHostRegisterDefault));
    return CAC_OK;
}

extern "C" int cac_host_unregister(void* h_ptr) {
    if (!h_ptr) return fail(CAC_ERR_INVALID, "bad argument");
    CUDA_TRY(cudaHostUnregister(h_ptr));
    return CAC_OK;
}

extern "C" int cac_normalize_run(const float* d_in, float* d_out, int64_t n_rows, int32_t row_len, void* stream) {
    if (!d_in || !d_out || n_rows < 0 || row_len < 1) return fail(CAC_ERR_INVALID, "bad argument");
    if (n_rows == 0) return CAC_OK;
    CUDA_TRY(launch_normalize(d_in, d_out, n_rows, row_len, (cudaStream_t)stream));
    return CAC_OK;
}

extern "C" int cac_mel_project_run(cac_plan* plan, const float* d_mag, int32_t n_frames, float* d_out, void* stream) {
    if (!plan || !d_mag || !d_out || n_frames < 0) return fail(CAC_ERR_INVALID, "bad argument");
    if (n_frames == 0) return CAC_OK;
    CUDA_TRY(cudaSetDevice(plan->device));
    CUDA_TRY(launch_mel_project(d_mag, d_out, n_frames, plan->p.n_mels, plan->p.power, plan->d_bt, plan->d_bw,
                                (cudaStream_t)stream));
    return CAC_OK;
}

extern "C" int cac_power_to_db_run(const float* d_in, float* d_out, int64_t n_items, int64_t item_len, void* d_work,
                                   void* stream) {
    if (!d_in || !d_out || !d_work || n_items < 0 || item_len < 1) return fail(CAC_ERR_INVALID, "bad argument");
    if (n_items == 0) return CAC_OK;
    CUDA_TRY(launch_power_to_db(d_in, d_out, n_items, item_len, (unsigned*)d_work, (cudaStream_t)stream));
    return CAC_OK;
}

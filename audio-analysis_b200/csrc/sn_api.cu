// C-ABI of the signal_noise chain (include/cacophony_b200.h: cac_sn_*, cac_signal_components_run).
// Nothing here allocates, synchronises or copies from the host: item tables are device arrays, scratch is the
// caller's work buffer, results stay on the device.
#include <algorithm>

#include "api_util.h"
#include "cac_common.cuh"
#include "sn.cuh"
#include "track_signals.cuh"

using namespace cac;

namespace cac {
__global__ void sn_single_item_kernel(SnItem* items, unsigned* amax, long long pcm_off, int n_samples, int cols, int pitch,
                                      int fps, unsigned amax_bits) {
    SnItem im;
    im.pcm_off = pcm_off;
    im.n_samples = n_samples;
    im.cols = cols;
    im.wpr = (cols + 31) / 32;
    im.stage_base = 0;
    im.n_stages = (cols + fps - 1) / fps;
    im.pitch = pitch;
    items[0] = im;
    amax[0] = amax_bits;
}
}  // namespace cac

struct cac_sn_plan {
    int device;
    int hop;
    int dilate_h, dilate_w, erode_h;
    int fps;                       // frames per stage of sn_stft_kernel
};

namespace {
constexpr int kMaxWave = 256;      // recordings per wave (the item table of a wave is built by one thread)
constexpr size_t kAlign = 256;
size_t up(size_t v) { return (v + kAlign - 1) / kAlign * kAlign; }

struct Layout {                    // carve-up of the work buffer for a wave of `slots` recordings
    size_t items, amax, colmed, colthr, bits, spec, total;
    size_t col_stride, plane_stride, bits_stride, spec_stride;   // in elements
    long long max_cols;
    int max_wpr;
};
Layout layout(int rows, long long max_cols, int slots) {
    Layout L;
    L.max_cols = max_cols;
    L.max_wpr = (int)((max_cols + 31) / 32);
    L.col_stride = (size_t)(max_cols + 31) / 32 * 32;
    L.plane_stride = (size_t)rows * L.max_wpr;
    L.bits_stride = 3 * L.plane_stride;
    L.spec_stride = (size_t)rows * (size_t)((max_cols + 31) / 32 * 32);   // rows of the matrix are padded to 32 floats
    size_t o = 0;
    L.items = o; o += up((size_t)slots * sizeof(SnItem));
    L.amax = o; o += up((size_t)slots * sizeof(unsigned));
    L.colmed = o; o += up((size_t)slots * L.col_stride * sizeof(float));
    L.colthr = o; o += up((size_t)slots * L.col_stride * sizeof(float));
    L.bits = o; o += up((size_t)slots * L.bits_stride * sizeof(unsigned));
    L.spec = o; o += up((size_t)slots * L.spec_stride * sizeof(float));
    L.total = o;
    return L;
}
SnWave wave_of(const Layout& L, char* base, int rows, int n_items, int* result, long long result_stride, int max_components) {
    SnWave w{};
    w.items = reinterpret_cast<const SnItem*>(base + L.items);
    w.n_items = n_items;
    w.rows = rows;
    w.spec = reinterpret_cast<float*>(base + L.spec);
    w.spec_stride = (long long)L.spec_stride;
    w.colmed = reinterpret_cast<float*>(base + L.colmed);
    w.col_stride = (long long)L.col_stride;
    w.colthr = reinterpret_cast<float*>(base + L.colthr);
    w.amax = reinterpret_cast<unsigned*>(base + L.amax);
    w.bits = reinterpret_cast<unsigned*>(base + L.bits);
    w.bits_stride = (long long)L.bits_stride;
    w.plane_stride = (long long)L.plane_stride;
    w.label = reinterpret_cast<int*>(base + L.spec);     // the matrix is dead once the mask exists: it becomes the label forest
    w.label_stride = (long long)L.spec_stride;
    w.result = result;
    w.result_stride = result_stride;
    w.max_components = max_components;
    return w;
}

// cv2.morphologyEx(MORPH_OPEN, ones((4, 4))) -> cv2.dilate(ones((h, w))) -> cv2.erode(ones((h // 10, w))) ->
// cv2.connectedComponentsWithStats on the wave's masks (src/identify_tracks.py:669-686); plane 0 holds the threshold mask
int mask_chain(const SnWave& w, const Layout& L, int dilate_h, int dilate_w, int erode_h, unsigned char* d_stage_masks,
               long long mask_stride, cudaStream_t st) {
    const size_t cells = (size_t)w.rows * (size_t)L.max_cols;
    if (d_stage_masks) CUDA_TRY(launch_unpack_bits_batch(w, 0, d_stage_masks, mask_stride, (int)L.max_cols, st));
    CUDA_TRY(launch_morph_rect_bits_batch(w, 0, 1, 2, L.max_wpr, 4, 4, false, st));
    CUDA_TRY(launch_morph_rect_bits_batch(w, 2, 1, 0, L.max_wpr, 4, 4, true, st));
    if (d_stage_masks) CUDA_TRY(launch_unpack_bits_batch(w, 0, d_stage_masks + cells, mask_stride, (int)L.max_cols, st));
    CUDA_TRY(launch_morph_rect_bits_batch(w, 0, 1, 2, L.max_wpr, dilate_h, dilate_w, true, st));
    // erode_h == 0 (height // 10 for sr > ~51 kHz): np.ones((0, w)) is an EMPTY kernel and cv2 then uses its 3 x 3 default
    CUDA_TRY(launch_morph_rect_bits_batch(w, 2, 1, 0, L.max_wpr, erode_h ? erode_h : 3, erode_h ? dilate_w : 3, false, st));
    if (d_stage_masks) CUDA_TRY(launch_unpack_bits_batch(w, 0, d_stage_masks + 2 * cells, mask_stride, (int)L.max_cols, st));
    CUDA_TRY(launch_ccl_runs_batch(w, 0, L.max_wpr, st));
    return CAC_OK;
}

int check_morph(int dilate_h, int dilate_w, int erode_h) {
    if (dilate_h < 1 || dilate_w < 1 || erode_h < 0 || dilate_h > 255 || erode_h > 255 || dilate_w > 4096)
        return fail(CAC_ERR_INVALID, "structuring element out of range (height 1..255, width 1..4096)");
    return CAC_OK;
}
}  // namespace

extern "C" int cac_sn_plan_create(int32_t hop_length, int32_t dilate_h, int32_t dilate_w, int32_t erode_h, int device,
                                  cac_sn_plan** out) {
    if (!out) return fail(CAC_ERR_INVALID, "null argument");
    if (hop_length < 1 || hop_length > 4096) return fail(CAC_ERR_INVALID, "hop_length out of range");
    int rc = check_morph(dilate_h, dilate_w, erode_h);
    if (rc) return rc;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1)
        return fail(CAC_ERR_NO_DEVICE, "no CUDA device (this library has no CPU path)");
    if (device < 0 || device >= ndev) return fail(CAC_ERR_INVALID, "device index out of range");
    const int fps = sn_stft_frames_per_stage(hop_length);
    if (fps < 1) return fail(CAC_ERR_INVALID, "hop_length exceeds the shared-memory budget of the spectrum kernel");
    cac_sn_plan* p = new cac_sn_plan();
    p->device = device;
    p->hop = hop_length;
    p->dilate_h = dilate_h;
    p->dilate_w = dilate_w;
    p->erode_h = erode_h;
    p->fps = fps;
    *out = p;
    return CAC_OK;
}

extern "C" int cac_sn_plan_destroy(cac_sn_plan* plan) {
    delete plan;
    return CAC_OK;
}

extern "C" int64_t cac_sn_frame_count(const cac_sn_plan* plan, int64_t n_samples) {
    if (!plan || n_samples < 0) return CAC_ERR_INVALID;
    return 1 + n_samples / plan->hop;
}

extern "C" size_t cac_sn_work_bytes(const cac_sn_plan* plan, int64_t n_items, int64_t max_item_samples) {
    if (!plan || n_items < 1 || max_item_samples < 0) return 0;
    const int slots = (int)std::min<int64_t>(n_items, 24);     // ~92 MB per one-minute slot; more slots do not make it faster
    return layout(kSnRows, 1 + max_item_samples / plan->hop, slots).total;
}

extern "C" int cac_sn_run(cac_sn_plan* plan, const float* d_pcm, const int64_t* d_item_offset, const int64_t* d_item_samples,
                          int64_t n_items, int64_t max_item_samples, int32_t max_components, int32_t* d_result, void* d_work,
                          size_t work_bytes, void* stream) {
    if (!plan || !d_pcm || !d_item_offset || !d_item_samples || !d_result || !d_work)
        return fail(CAC_ERR_INVALID, "null argument");
    if (n_items < 0 || max_components < 1) return fail(CAC_ERR_INVALID, "n_items / max_components out of range");
    if (n_items == 0) return CAC_OK;
    const long long max_cols = 1 + max_item_samples / plan->hop;
    if (max_item_samples < 0 || (long long)kSnRows * max_cols > 0x7fffffffLL)
        return fail(CAC_ERR_INVALID, "recording too long: the [2049 x frames] matrix must stay below 2^31 cells");
    if (((uintptr_t)d_work & 255) != 0) return fail(CAC_ERR_INVALID, "d_work must be 256-byte aligned");
    // how many recordings fit the work buffer at once
    int slots = 0;
    for (int s = (int)std::min<int64_t>(n_items, kMaxWave); s >= 1; s = (s > 32 ? s - 8 : s - 1))
        if (layout(kSnRows, max_cols, s).total <= work_bytes) { slots = s; break; }
    if (slots < 1) return fail(CAC_ERR_INVALID, "work buffer too small for one recording (see cac_sn_work_bytes)");
    cudaStream_t st = (cudaStream_t)stream;
    CUDA_TRY(cudaSetDevice(plan->device));
    const Layout L = layout(kSnRows, max_cols, slots);
    const long long rstride = 4 + 6LL * max_components;
    for (int64_t i0 = 0; i0 < n_items; i0 += slots) {
        const int n = (int)std::min<int64_t>(slots, n_items - i0);
        SnWave w = wave_of(L, (char*)d_work, kSnRows, n, d_result + i0 * rstride, rstride, max_components);
        CUDA_TRY(launch_sn_prep(reinterpret_cast<const long long*>(d_item_offset + i0),
                                reinterpret_cast<const long long*>(d_item_samples + i0), n, plan->hop, plan->fps,
                                (long long)max_item_samples, const_cast<SnItem*>(w.items), w.amax, st));
        CUDA_TRY(launch_sn_stft(d_pcm, w, plan->hop, plan->fps, st));
        CUDA_TRY(launch_sn_col_thresholds(w, (int)max_cols, st));
        CUDA_TRY(launch_sn_rowmed_mask(w, (int)max_cols, true, st));
        int rc = mask_chain(w, L, plan->dilate_h, plan->dilate_w, plan->erode_h, nullptr, 0, st);
        if (rc) return rc;
    }
    return CAC_OK;
}

extern "C" int cac_sn_spectrum_run(cac_sn_plan* plan, const float* d_pcm, int64_t n_samples, float* d_spec, float* d_colmed,
                                   float* d_amax, void* d_work, size_t work_bytes, void* stream) {
    if (!plan || !d_pcm || !d_spec || !d_colmed || !d_amax || !d_work) return fail(CAC_ERR_INVALID, "null argument");
    const long long cols = 1 + n_samples / plan->hop;
    if (n_samples < 0 || (long long)kSnRows * cols > 0x7fffffffLL) return fail(CAC_ERR_INVALID, "recording length out of range");
    if (work_bytes < 512 || ((uintptr_t)d_work & 255) != 0) return fail(CAC_ERR_INVALID, "d_work: 512 bytes, 256-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    CUDA_TRY(cudaSetDevice(plan->device));
    SnWave w{};
    SnItem* items = reinterpret_cast<SnItem*>(d_work);
    w.items = items;
    w.n_items = 1;
    w.rows = kSnRows;
    w.spec = d_spec;
    w.colmed = d_colmed;
    w.amax = reinterpret_cast<unsigned*>(d_amax);
    sn_single_item_kernel<<<1, 1, 0, st>>>(items, w.amax, 0, (int)n_samples, (int)cols, (int)((cols + 31) / 32 * 32), plan->fps, 0u);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(launch_sn_stft(d_pcm, w, plan->hop, plan->fps, st));
    return CAC_OK;
}

extern "C" size_t cac_signal_components_work_bytes(int32_t rows, int64_t cols) {
    if (rows < 1 || cols < 1) return 0;
    return layout(rows, cols, 1).total;
}

// The post-STFT chain on a spectrogram the caller supplies (already divided by its maximum): column medians through a
// transpose (the fused chain gets them from its spectrum kernel instead), then the same row-median / mask, morphology
// and component kernels as cac_sn_run with a_max = 1 (x / 1 = x exactly).
extern "C" int cac_signal_components_run(const float* d_spec, int32_t rows, int64_t cols, int32_t dilate_h, int32_t dilate_w,
                                         int32_t erode_h, int32_t max_components, int32_t* d_result, uint8_t* d_stage_masks,
                                         void* d_work, size_t work_bytes, void* stream) {
    if (!d_spec || !d_result || !d_work) return fail(CAC_ERR_INVALID, "null argument");
    if (rows < 1 || cols < 1 || (int64_t)rows * cols > 0x7fffffffLL) return fail(CAC_ERR_INVALID, "matrix size out of range");
    if (rows > 65535) return fail(CAC_ERR_INVALID, "more than 65,535 rows (rows ride on a grid dimension)");
    if (max_components < 1) return fail(CAC_ERR_INVALID, "max_components out of range");
    int rc = check_morph(dilate_h, dilate_w, erode_h);
    if (rc) return rc;
    const Layout L = layout(rows, cols, 1);
    if (work_bytes < L.total || ((uintptr_t)d_work & 255) != 0)
        return fail(CAC_ERR_INVALID, "work buffer too small or not 256-byte aligned (see cac_signal_components_work_bytes)");
    cudaStream_t st = (cudaStream_t)stream;
    SnWave w = wave_of(L, (char*)d_work, rows, 1, d_result, 4 + 6LL * max_components, max_components);
    float* scratch = w.spec;                              // transposed copy now, label forest later
    w.spec = const_cast<float*>(d_spec);                  // read-only for every kernel of this path
    sn_single_item_kernel<<<1, 1, 0, st>>>(const_cast<SnItem*>(w.items), w.amax, 0, 0, (int)cols, (int)cols, 1, 0x3f800000u /* 1.0f */);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(launch_transpose(d_spec, scratch, rows, cols, st));
    CUDA_TRY(launch_row_median(scratch, w.colmed, cols, rows, st));
    CUDA_TRY(launch_sn_col_thresholds(w, (int)cols, st));
    const bool aligned_rows = cols % 4 == 0 && ((uintptr_t)d_spec & 15) == 0;     // pitch = cols here
    CUDA_TRY(launch_sn_rowmed_mask(w, (int)cols, aligned_rows, st));
    return mask_chain(w, L, dilate_h, dilate_w, erode_h, d_stage_masks, 0, st);
}

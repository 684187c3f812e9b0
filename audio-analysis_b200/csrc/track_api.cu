// C-ABI of signal_noise's post-STFT stage (include/cacophony_b200.h, cac_signal_components).
#include <vector>

#include "api_util.h"
#include "cac_common.cuh"
#include "track_signals.cuh"

using namespace cac;

namespace {
struct Scratch {   // stream-ordered allocations (cached by the device's default memory pool), freed on every exit path
    cudaStream_t st;
    void* p[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    explicit Scratch(cudaStream_t s) : st(s) {}
    ~Scratch() { for (void* q : p) if (q) cudaFreeAsync(q, st); }
};
// keep freed scratch in the pool instead of returning it to the driver at every synchronisation
void keep_pool_warm() {
    static thread_local int done_for = -1;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev == done_for) return;
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
        unsigned long long thr = ~0ull;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
    }
    done_for = dev;
}
}  // namespace

// Launches the whole chain on `st`; the result block d_small = [count, -, -, -, 6 ints per component ...] stays on the
// device (d_small == nullptr: allocated here and returned through *d_small_out, owned by the Scratch).
static int components_launch(Scratch& s, const float* d_spec, int32_t rows, int64_t cols, int32_t dilate_h, int32_t dilate_w,
                             int32_t erode_h, int32_t max_components, uint8_t* d_stage_masks, int* d_small,
                             int** d_small_out, cudaStream_t st) {
    if (rows < 1 || cols < 1 || (int64_t)rows * cols > 0x7fffffffLL) return fail(CAC_ERR_INVALID, "matrix size out of range");
    if (dilate_h < 1 || dilate_w < 1 || erode_h < 1 || dilate_h > 255 || dilate_w > 4096 || max_components < 1)
        return fail(CAC_ERR_INVALID, "kernel size / max_components out of range");
    const size_t n = (size_t)rows * cols;
    float *d_t, *d_med;
    const size_t wpr = (size_t)((cols + 31) / 32), nw = (size_t)rows * wpr;
    // d_t doubles as the label array later (same size: 4 bytes per cell)
    CUDA_TRY(cudaMallocAsync(&s.p[0], n * sizeof(float), st));
    CUDA_TRY(cudaMallocAsync(&s.p[1], n * sizeof(int), st));
    CUDA_TRY(cudaMallocAsync(&s.p[2], 3 * nw * sizeof(unsigned), st));
    CUDA_TRY(cudaMallocAsync(&s.p[3], ((size_t)rows + cols) * sizeof(float), st));
    if (!d_small) CUDA_TRY(cudaMallocAsync(&s.p[4], ((size_t)max_components * 6 + 4) * sizeof(int), st));
    d_t = (float*)s.p[0];
    int* d_rootid = (int*)s.p[1];
    unsigned* m0 = (unsigned*)s.p[2];              // bit-packed masks (track_bits.cu): 2.6 MB each for a one-minute recording
    unsigned* m1 = m0 + nw;
    unsigned* m2 = m1 + nw;
    d_med = (float*)s.p[3];
    if (!d_small) d_small = (int*)s.p[4];
    if (d_small_out) *d_small_out = d_small;
    float* rowmed = d_med;
    float* colmed = d_med + rows;

    // medians: rows directly, columns as rows of the transposed matrix (:658-659)
    CUDA_TRY(launch_row_median(d_spec, rowmed, rows, cols, st));
    CUDA_TRY(launch_transpose(d_spec, d_t, rows, cols, st));
    CUDA_TRY(launch_row_median(d_t, colmed, cols, rows, st));
    CUDA_TRY(launch_signal_mask_bits(d_spec, rowmed, colmed, m0, rows, cols, st));
    if (d_stage_masks) CUDA_TRY(launch_unpack_bits(m0, d_stage_masks, rows, cols, st));
    // cv2.morphologyEx(MORPH_OPEN, ones((4, 4))) = erode then dilate (:669-671)
    CUDA_TRY(launch_morph_rect_bits(m0, m1, m2, rows, cols, 4, 4, false, st));
    CUDA_TRY(launch_morph_rect_bits(m2, m1, m0, rows, cols, 4, 4, true, st));
    if (d_stage_masks) CUDA_TRY(launch_unpack_bits(m0, d_stage_masks + n, rows, cols, st));
    // cv2.dilate(ones((height, width))) then cv2.erode(ones((height // 10, width))) (:683-684)
    CUDA_TRY(launch_morph_rect_bits(m0, m1, m2, rows, cols, dilate_h, dilate_w, true, st));
    CUDA_TRY(launch_morph_rect_bits(m2, m1, m0, rows, cols, erode_h, dilate_w, false, st));
    if (d_stage_masks) CUDA_TRY(launch_unpack_bits(m0, d_stage_masks + 2 * n, rows, cols, st));
    // cv2.connectedComponentsWithStats (8-connectivity) (:686), on horizontal runs
    int* d_stats = d_small + 4;
    CUDA_TRY(launch_ccl_runs(m0, (int*)d_t, d_rootid, rows, (int)cols, d_small, d_stats, max_components, st));
    return CAC_OK;
}

extern "C" int cac_signal_components(const float* d_spec, int32_t rows, int64_t cols, int32_t dilate_h, int32_t dilate_w,
                                     int32_t erode_h, int32_t max_components, int32_t* h_stats, int32_t* h_count,
                                     uint8_t* d_stage_masks, void* stream) {
    if (!d_spec || !h_stats || !h_count) return fail(CAC_ERR_INVALID, "null argument");
    cudaStream_t st = (cudaStream_t)stream;
    keep_pool_warm();
    Scratch s(st);
    int* d_small = nullptr;
    int rc = components_launch(s, d_spec, rows, cols, dilate_h, dilate_w, erode_h, max_components, d_stage_masks, nullptr,
                               &d_small, st);
    if (rc) return rc;
    int* d_stats = d_small + 4;
    int count = 0;
    CUDA_TRY(cudaMemcpyAsync(&count, d_small, sizeof(int), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    if (count > max_components) return fail(CAC_ERR_INVALID, "more connected components than max_components");
    std::vector<int> raw((size_t)count * 6);
    if (count) CUDA_TRY(cudaMemcpy(raw.data(), d_stats, raw.size() * sizeof(int), cudaMemcpyDeviceToHost));
    for (int i = 0; i < count; ++i) {       // cv2 stats layout: x, y, width, height, area (+ our label-order key)
        const int* r = raw.data() + 6 * i;
        int32_t* o = h_stats + 6 * i;
        o[0] = r[0]; o[1] = r[1]; o[2] = r[2] - r[0] + 1; o[3] = r[3] - r[1] + 1; o[4] = r[4]; o[5] = r[5];
    }
    *h_count = count;
    return CAC_OK;
}

// The same chain without the read-back: everything is queued on `stream` and the result block stays on the device
// (d_result: 4 + 6 * max_components ints - [0] = number of components, then per component
// {min x, min y, max x, max y, area, order}).  For batches: queue one call per recording, synchronise once.
extern "C" int cac_signal_components_async(const float* d_spec, int32_t rows, int64_t cols, int32_t dilate_h,
                                           int32_t dilate_w, int32_t erode_h, int32_t max_components, int32_t* d_result,
                                           void* stream) {
    if (!d_spec || !d_result) return fail(CAC_ERR_INVALID, "null argument");
    cudaStream_t st = (cudaStream_t)stream;
    keep_pool_warm();
    Scratch s(st);          // freed in stream order when the call returns: safe, later work on `st` comes after it
    return components_launch(s, d_spec, rows, cols, dilate_h, dilate_w, erode_h, max_components, nullptr, d_result, nullptr, st);
}

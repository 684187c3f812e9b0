// C-ABI of the old Cacophony Index path (include/cacophony_b200.h, cac_ci_*).
#include <cmath>
#include <cstring>
#include <vector>

#include "api_util.h"
#include "cac_common.cuh"
#include "ci.cuh"

using namespace cac;

struct cac_ci_plan {
    int device;
    double* d_window = nullptr;     // [2048]
    double2* d_ltab = nullptr;      // [kCiLtabRows][32]
    // item / entry tables staged per call (grown on demand)
    long long* d_tab = nullptr;     size_t tab_cap = 0;
    long long* d_etab = nullptr;    size_t etab_cap = 0;
    // buffers of the *_host entry point
    float* d_pcm = nullptr;         size_t pcm_cap = 0;
    double* d_bins = nullptr;       size_t bins_cap = 0;
    unsigned char* d_points = nullptr; size_t points_cap = 0;
};

extern "C" int64_t cac_ci_window_count(int64_t n_samples) {
    // len(range(1024, n_samples - 3072, 1024))   (src/cacophony_index.py:89-91)
    const int64_t stop = n_samples - 3072;
    if (stop <= 1024) return 0;
    return (stop - 1024 + 1023) / 1024;
}

extern "C" int cac_ci_plan_create(int32_t window_size, const int32_t* h_edges, const double* h_window, int device,
                                  cac_ci_plan** out) {
    if (!h_edges || !h_window || !out) return fail(CAC_ERR_INVALID, "null argument");
    static const int32_t kEdges[11] = {25, 38, 60, 93, 145, 226, 351, 546, 848, 1318, 2048};
    if (window_size != 2048 || std::memcmp(h_edges, kEdges, sizeof(kEdges)) != 0)
        return fail(CAC_ERR_INVALID,
                    "the Cacophony Index kernel is built for window 2048 at 16 kHz (band edges 25..2048) only");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1)
        return fail(CAC_ERR_NO_DEVICE, "no CUDA device (this library has no CPU path)");
    if (device < 0 || device >= ndev) return fail(CAC_ERR_INVALID, "device index out of range");
    CUDA_TRY(cudaSetDevice(device));
    // per-lane twiddles: rows 0-3 W_1024^(8 a lane), rows 4-11 W_1024^(b lane), row 12 W_8192^lane,
    // row 13 -i W_8192^(5 lane)
    // rows 14 + 2 K2 / 15 + 2 K2: W_8192^k and -i W_8192^(5k) for k = lane + 32 K2 (the untangle's twiddles)
    std::vector<double2> lt(kCiLtabRows * 32);
    const long double tau = 6.283185307179586476925286766559005768L;
    for (int lane = 0; lane < 32; ++lane) {
        for (int a = 0; a < 4; ++a) {
            long double ang = -tau * (long double)((8 * a * lane) % 1024) / 1024.0L;
            lt[a * 32 + lane] = make_double2((double)cosl(ang), (double)sinl(ang));
        }
        for (int b = 0; b < 8; ++b) {
            long double ang = -tau * (long double)(b * lane) / 1024.0L;
            lt[(4 + b) * 32 + lane] = make_double2((double)cosl(ang), (double)sinl(ang));
        }
        long double a1 = -tau * (long double)lane / 8192.0L, a5 = -tau * (long double)(5 * lane) / 8192.0L;
        lt[12 * 32 + lane] = make_double2((double)cosl(a1), (double)sinl(a1));
        lt[13 * 32 + lane] = make_double2((double)sinl(a5), (double)-cosl(a5));   // -i (c + i s) = s - i c
        for (int k2 = 0; k2 < 32; ++k2) {
            const int k = lane + 32 * k2;
            const long double b1 = -tau * (long double)k / 8192.0L, b5 = -tau * (long double)((5 * k) % 8192) / 8192.0L;
            lt[(14 + 2 * k2) * 32 + lane] = make_double2((double)cosl(b1), (double)sinl(b1));
            lt[(15 + 2 * k2) * 32 + lane] = make_double2((double)sinl(b5), (double)-cosl(b5));
        }
    }
    cac_ci_plan* plan = new cac_ci_plan();
    plan->device = device;
    cudaError_t e1 = cudaMalloc((void**)&plan->d_window, 2048 * sizeof(double));
    cudaError_t e2 = cudaMalloc((void**)&plan->d_ltab, lt.size() * sizeof(double2));
    if (e1 == cudaSuccess && e2 == cudaSuccess) {
        e1 = cudaMemcpy(plan->d_window, h_window, 2048 * sizeof(double), cudaMemcpyHostToDevice);
        e2 = cudaMemcpy(plan->d_ltab, lt.data(), lt.size() * sizeof(double2), cudaMemcpyHostToDevice);
    }
    if (e1 != cudaSuccess || e2 != cudaSuccess) {
        cudaFree(plan->d_window); cudaFree(plan->d_ltab);
        delete plan;
        return fail(CAC_ERR_ALLOC, "could not upload the Cacophony Index tables");
    }
    *out = plan;
    return CAC_OK;
}

extern "C" int cac_ci_plan_destroy(cac_ci_plan* plan) {
    if (!plan) return CAC_OK;
    cudaSetDevice(plan->device);
    cudaFree(plan->d_window); cudaFree(plan->d_ltab); cudaFree(plan->d_tab); cudaFree(plan->d_etab);
    cudaFree(plan->d_pcm); cudaFree(plan->d_bins); cudaFree(plan->d_points);
    delete plan;
    return CAC_OK;
}

extern "C" int cac_ci_run(cac_ci_plan* plan, const float* d_pcm, const int64_t* h_item_offset,
                          const int64_t* h_item_samples, int64_t n_items, double* d_bins, uint8_t* d_points,
                          void* stream) {
    if (!plan || !d_pcm || !h_item_offset || !h_item_samples || !d_bins)
        return fail(CAC_ERR_INVALID, "null argument");
    if (n_items < 0 || n_items > 0x7fffffffLL) return fail(CAC_ERR_INVALID, "n_items out of range");
    if (n_items == 0) return CAC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    CUDA_TRY(cudaSetDevice(plan->device));
    std::vector<long long> tab(2 * (size_t)n_items + 1);           // [item_off | win_base]
    long long total = 0;
    for (int64_t i = 0; i < n_items; ++i) {
        if (h_item_offset[i] < 0 || h_item_samples[i] < 0) return fail(CAC_ERR_INVALID, "negative item offset/length");
        tab[i] = h_item_offset[i];
        tab[n_items + i] = total;
        total += cac_ci_window_count(h_item_samples[i]);
    }
    tab[2 * n_items] = total;
    if (total == 0) return CAC_OK;
    int rc = ensure_cap(&plan->d_tab, &plan->tab_cap, tab.size() * sizeof(long long));
    if (rc) return rc;
    // pageable source: the copy is staged before the call returns, so `tab` may go out of scope
    CUDA_TRY(cudaMemcpyAsync(plan->d_tab, tab.data(), tab.size() * sizeof(long long), cudaMemcpyHostToDevice, st));
    CiArgs a{};
    a.pcm = d_pcm;
    a.item_off = plan->d_tab;
    a.win_base = plan->d_tab + n_items;
    a.n_items = (int)n_items;
    a.n_windows = total;
    a.first = 1024;
    a.hop = 1024;
    a.window = plan->d_window;
    a.ltab = plan->d_ltab;
    a.bins = d_bins;
    CUDA_TRY(launch_ci_bins(a, st));
    if (d_points) CUDA_TRY(launch_ci_points(d_bins, a.win_base, a.n_items, total, d_points, st));
    return CAC_OK;
}

extern "C" int cac_ci_bins_run(cac_ci_plan* plan, const float* d_pcm, int64_t n_windows, int32_t hop, double* d_bins,
                               void* stream) {
    if (!plan || !d_pcm || !d_bins) return fail(CAC_ERR_INVALID, "null argument");
    if (n_windows < 0 || hop < 0) return fail(CAC_ERR_INVALID, "n_windows / hop out of range");
    if (n_windows == 0) return CAC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    CUDA_TRY(cudaSetDevice(plan->device));
    const long long tab[3] = {0, 0, n_windows};                    // one item at offset 0 holding every window
    int rc = ensure_cap(&plan->d_tab, &plan->tab_cap, sizeof(tab));
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(plan->d_tab, tab, sizeof(tab), cudaMemcpyHostToDevice, st));
    CiArgs a{};
    a.pcm = d_pcm;
    a.item_off = plan->d_tab;
    a.win_base = plan->d_tab + 1;
    a.n_items = 1;
    a.n_windows = n_windows;
    a.first = 0;
    a.hop = hop;
    a.window = plan->d_window;
    a.ltab = plan->d_ltab;
    a.bins = d_bins;
    CUDA_TRY(launch_ci_bins(a, st));
    return CAC_OK;
}

extern "C" int cac_ci_entries_run(cac_ci_plan* plan, const uint8_t* d_points, const int64_t* h_entry_start,
                                  const int32_t* h_entry_len, int64_t n_entries, int32_t* d_out, void* stream) {
    if (!plan || !d_points || !h_entry_start || !h_entry_len || !d_out) return fail(CAC_ERR_INVALID, "null argument");
    if (n_entries < 0) return fail(CAC_ERR_INVALID, "n_entries out of range");
    if (n_entries == 0) return CAC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    CUDA_TRY(cudaSetDevice(plan->device));
    const size_t bytes = (size_t)n_entries * (sizeof(long long) + sizeof(int));
    int rc = ensure_cap(&plan->d_etab, &plan->etab_cap, bytes);
    if (rc) return rc;
    int* d_len = reinterpret_cast<int*>(plan->d_etab + n_entries);
    CUDA_TRY(cudaMemcpyAsync(plan->d_etab, h_entry_start, (size_t)n_entries * sizeof(long long), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(d_len, h_entry_len, (size_t)n_entries * sizeof(int), cudaMemcpyHostToDevice, st));
    CUDA_TRY(launch_ci_entries(d_points, plan->d_etab, d_len, n_entries, d_out, st));
    return CAC_OK;
}

extern "C" int cac_ci_host(cac_ci_plan* plan, const float* h_pcm, int64_t n_pcm, const int64_t* h_item_offset,
                           const int64_t* h_item_samples, int64_t n_items, double* h_bins, uint8_t* h_points) {
    if (!plan || !h_pcm || !h_item_offset || !h_item_samples || !h_bins) return fail(CAC_ERR_INVALID, "null argument");
    if (n_items <= 0) return CAC_OK;
    CUDA_TRY(cudaSetDevice(plan->device));
    long long total = 0;
    for (int64_t i = 0; i < n_items; ++i) {
        if (h_item_offset[i] < 0 || h_item_samples[i] < 0 || h_item_offset[i] + h_item_samples[i] > n_pcm)
            return fail(CAC_ERR_INVALID, "item out of bounds");
        total += cac_ci_window_count(h_item_samples[i]);
    }
    if (total == 0) return CAC_OK;
    int rc;
    if ((rc = ensure_cap(&plan->d_pcm, &plan->pcm_cap, (size_t)n_pcm * sizeof(float) + 16))) return rc;
    if ((rc = ensure_cap(&plan->d_bins, &plan->bins_cap, (size_t)total * 10 * sizeof(double)))) return rc;
    if ((rc = ensure_cap(&plan->d_points, &plan->points_cap, (size_t)total))) return rc;
    CUDA_TRY(cudaMemcpyAsync(plan->d_pcm, h_pcm, (size_t)n_pcm * sizeof(float), cudaMemcpyHostToDevice, 0));
    if ((rc = cac_ci_run(plan, plan->d_pcm, h_item_offset, h_item_samples, n_items, plan->d_bins,
                         h_points ? plan->d_points : nullptr, nullptr)))
        return rc;
    CUDA_TRY(cudaMemcpyAsync(h_bins, plan->d_bins, (size_t)total * 10 * sizeof(double), cudaMemcpyDeviceToHost, 0));
    if (h_points) CUDA_TRY(cudaMemcpyAsync(h_points, plan->d_points, (size_t)total, cudaMemcpyDeviceToHost, 0));
    CUDA_TRY(cudaStreamSynchronize(0));
    return CAC_OK;
}

// Rational-rate resampling front-end: 48 kHz -> 16 kHz for the Cacophony Index
// (the reference shells out to `ffmpeg -ar 16000`, src/common.py:76-79) and native rate -> 48 kHz for the
// classifier path (librosa.resample, src/identify_tracks.py:55-57).  Neither ffmpeg's nor soxr's filter
// can be pinned here, so the filter is our own choice and stated as such: scipy.signal.resample_poly's
// design (Kaiser beta 5 windowed sinc, 20 max(up, down) + 1 taps, zero padding at the ends), built on the
// host by the caller and passed in already padded, so results are pinned against scipy itself.
//   y[m] = sum_k h[k] xu[(m + n_pre_remove) down - k],   xu[j up] = x[j], zero elsewhere.
// One thread per output sample; the taps live in shared memory, the input span of a CTA is contiguous
// and read through L1 (consecutive threads start `down / up` samples apart).
#include <vector>

#include "api_util.h"
#include "cac_common.cuh"

namespace cac {

__global__ void __launch_bounds__(256) resample_poly_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                            const long long* __restrict__ tab, int n_items,
                                                            const float* __restrict__ taps, int n_taps, int up, int down,
                                                            int n_pre_remove) {
    extern __shared__ float h[];
    for (int i = threadIdx.x; i < n_taps; i += 256) h[i] = taps[i];
    __syncthreads();
    const int item = blockIdx.y;
    const long long in_off = tab[item], n_in = tab[n_items + item], out_off = tab[2 * n_items + item];
    const long long n_out = (n_in * up + down - 1) / down;
    const float* x = in + in_off;
    for (long long m = (long long)blockIdx.x * 256 + threadIdx.x; m < n_out; m += (long long)gridDim.x * 256) {
        const long long t = (m + n_pre_remove) * down;       // position on the up-sampled grid
        const int k0 = (int)(t % up);                         // first tap that meets a real sample
        long long j_hi = t / up;                              // ... which is x[j_hi]; tap k0 + up i meets x[j_hi - i]
        int k = k0;
        if (j_hi >= n_in) {                                   // taps that fall past the end meet zeros
            const long long skip = j_hi - (n_in - 1);
            k += (int)(skip * up);
            j_hi -= skip;
        }
        float acc = 0.f;
        if (k < n_taps && j_hi >= 0) {
            // oldest sample first, separate multiply and add: scipy's upfirdn loop order and float32 arithmetic
            long long i_max = (n_taps - 1 - k) / up;
            if (i_max > j_hi) i_max = j_hi;
            int kk = k + (int)i_max * up;
            for (long long j = j_hi - i_max; j <= j_hi; ++j, kk -= up) acc = __fadd_rn(acc, __fmul_rn(__ldg(x + j), h[kk]));
        }
        out[out_off + m] = acc;
    }
}

}  // namespace cac

struct cac_resample_plan {
    int device, up, down, n_taps, n_pre_remove;
    float* d_taps = nullptr;
    long long* d_tab = nullptr;  size_t tab_cap = 0;
};

using namespace cac;

extern "C" int cac_resample_plan_create(int32_t up, int32_t down, const float* h_taps, int32_t n_taps,
                                        int32_t n_pre_remove, int device, cac_resample_plan** out) {
    if (!h_taps || !out) return fail(CAC_ERR_INVALID, "null argument");
    if (up < 1 || down < 1 || n_taps < 1 || n_taps > 12000 || n_pre_remove < 0)
        return fail(CAC_ERR_INVALID, "up / down / taps out of range (at most 12,000 taps)");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1)
        return fail(CAC_ERR_NO_DEVICE, "no CUDA device (this library has no CPU path)");
    if (device < 0 || device >= ndev) return fail(CAC_ERR_INVALID, "device index out of range");
    CUDA_TRY(cudaSetDevice(device));
    cac_resample_plan* plan = new cac_resample_plan();
    plan->device = device; plan->up = up; plan->down = down; plan->n_taps = n_taps; plan->n_pre_remove = n_pre_remove;
    cudaError_t e = cudaMalloc((void**)&plan->d_taps, (size_t)n_taps * sizeof(float));
    if (e == cudaSuccess) e = cudaMemcpy(plan->d_taps, h_taps, (size_t)n_taps * sizeof(float), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        cudaFree(plan->d_taps);
        delete plan;
        return fail(CAC_ERR_ALLOC, "could not upload the resampling filter");
    }
    *out = plan;
    return CAC_OK;
}

extern "C" int cac_resample_plan_destroy(cac_resample_plan* plan) {
    if (!plan) return CAC_OK;
    cudaSetDevice(plan->device);
    cudaFree(plan->d_taps); cudaFree(plan->d_tab);
    delete plan;
    return CAC_OK;
}

extern "C" int64_t cac_resample_out_len(const cac_resample_plan* plan, int64_t n_in) {
    if (!plan || n_in < 0) return CAC_ERR_INVALID;
    return (n_in * plan->up + plan->down - 1) / plan->down;
}

extern "C" int cac_resample_run(cac_resample_plan* plan, const float* d_in, const int64_t* h_in_offset,
                                const int64_t* h_in_samples, int64_t n_items, float* d_out, const int64_t* h_out_offset,
                                void* stream) {
    if (!plan || !d_in || !d_out || !h_in_offset || !h_in_samples || !h_out_offset)
        return fail(CAC_ERR_INVALID, "null argument");
    if (n_items < 0 || n_items > 65535) return fail(CAC_ERR_INVALID, "n_items out of range (at most 65,535 per call)");
    if (n_items == 0) return CAC_OK;
    cudaStream_t st = (cudaStream_t)stream;
    CUDA_TRY(cudaSetDevice(plan->device));
    std::vector<long long> tab(3 * (size_t)n_items);
    long long longest = 0;
    for (int64_t i = 0; i < n_items; ++i) {
        if (h_in_offset[i] < 0 || h_in_samples[i] < 0 || h_out_offset[i] < 0)
            return fail(CAC_ERR_INVALID, "negative item offset / length");
        tab[i] = h_in_offset[i];
        tab[n_items + i] = h_in_samples[i];
        tab[2 * n_items + i] = h_out_offset[i];
        const long long n_out = cac_resample_out_len(plan, h_in_samples[i]);
        longest = n_out > longest ? n_out : longest;
    }
    if (longest == 0) return CAC_OK;
    int rc = ensure_cap(&plan->d_tab, &plan->tab_cap, tab.size() * sizeof(long long));
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(plan->d_tab, tab.data(), tab.size() * sizeof(long long), cudaMemcpyHostToDevice, st));
    long long bx = (longest + 255) / 256;
    if (bx > 148LL * 64) bx = 148LL * 64;
    dim3 grid((unsigned)bx, (unsigned)n_items);
    resample_poly_kernel<<<grid, 256, (size_t)plan->n_taps * sizeof(float), st>>>(
        d_in, d_out, plan->d_tab, (int)n_items, plan->d_taps, plan->n_taps, plan->up, plan->down, plan->n_pre_remove);
    CUDA_TRY(cudaGetLastError());
    return CAC_OK;
}

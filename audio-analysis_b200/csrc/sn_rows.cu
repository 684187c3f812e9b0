// signal_noise's row medians and threshold mask for a wave of recordings (src/identify_tracks.py:657-667):
//   spectogram = spectogram / a_max
//   row_medians = np.median(spectogram, axis=1)
//   signal = (spectogram > 3 * column_medians) & (spectogram > 3 * row_medians)
// One CTA per (row, recording).  The row is read from HBM ONCE into registers; the median is selected there by a
// radix select on the unscaled values (sn_select.cuh: below the common prefix of the row's smallest and largest key,
// 12 bits per pass - two passes for a typical row; division by a_max is monotone, so the order statistic of the
// quotients is the quotient of the order statistic), the one or two selected values are scaled, and the mask of the
// row is formed on the spot from fl(x / a_max) - bit-identical to the reference's arithmetic - and written bit-packed
// (32 cells per ballot).  The scaled matrix, the transposed matrix and the separate mask pass of the first version
// (5 reads + 2 writes of the 84 MB matrix per minute of audio) are gone.
#include <cstdlib>

#include "cac_common.cuh"
#include "sn.cuh"
#include "sn_select.cuh"

namespace cac {

// The reference compares fl(x / a_max) with a threshold t.  fl(x / a) is monotone in x, so that comparison equals
// x > x_t with x_t = the largest float whose quotient still rounds to <= t: one or two divisions per THRESHOLD (the
// product t a is within an ulp or two of x_t; it is walked to the exact boundary) instead of one per CELL.
__device__ __forceinline__ float unscaled_threshold(float t, float a) {
    if (!(a > 0.f)) return INFINITY;                  // silent recording: 0 / 0 = NaN, nothing exceeds a threshold
    unsigned y = __float_as_uint(__fmul_rn(t, a));
    while (y > 0u && __fdiv_rn(__uint_as_float(y), a) > t) --y;
    while (y < 0x7f7fffffu && __fdiv_rn(__uint_as_float(y + 1u), a) <= t) ++y;
    return __uint_as_float(y);
}

// colthr[c] = x_t of t = fl(3 * fl(colmed[c] / a_max)): the column threshold in the unscaled domain, once per column
__global__ void __launch_bounds__(256) sn_col_thresholds_kernel(const SnWave w) {
    const int item = blockIdx.y;
    const int cols = w.items[item].cols;
    const int c = blockIdx.x * 256 + threadIdx.x;
    if (c >= cols) return;
    const float a = __uint_as_float(w.amax[item]);
    const size_t o = (size_t)item * w.col_stride + c;
    w.colthr[o] = unscaled_threshold(__fmul_rn(3.0f, __fdiv_rn(w.colmed[o], a)), a);
}

struct RowSel {                       // shared scratch of one CTA's select
    unsigned hist[kSelCtaBins];
    unsigned tot[32];
    unsigned out[2];
    unsigned mn, mx, cnt_le, min_gt;
};

// CTA-wide minimum / maximum of per-thread values (all threads call; ends with a barrier)
template <int NT>
__device__ __forceinline__ void cta_minmax(RowSel& s, unsigned mn, unsigned mx) {
    if (threadIdx.x == 0) { s.mn = 0xffffffffu; s.mx = 0u; }
    __syncthreads();
    mn = __reduce_min_sync(0xffffffffu, mn);
    mx = __reduce_max_sync(0xffffffffu, mx);
    if ((threadIdx.x & 31) == 0) { atomicMin(&s.mn, mn); atomicMax(&s.mx, mx); }
    __syncthreads();
}

// median of the SCALED row from the selected order statistic(s): v1 = rank (cols - 1) / 2 of the unscaled keys,
// cnt_le / min_gt = how many keys are <= v1 and the smallest key above it (only read for an even count)
__device__ __forceinline__ float scaled_median(unsigned v1, unsigned cnt_le, unsigned min_gt, int cols, float a) {
    float med = __fdiv_rn(__uint_as_float(v1), a);
    if ((cols & 1) == 0) {
        // even count: the next order statistic is v1 again when enough elements are <= v1, else the smallest one above;
        // np.median -> np.mean of the two middle float32 values
        const unsigned k0 = (unsigned)((cols - 1) / 2);
        const unsigned v2 = (cnt_le >= k0 + 2u) ? v1 : min_gt;
        med = __fdiv_rn(__fadd_rn(med, __fdiv_rn(__uint_as_float(v2), a)), 2.0f);
    }
    return med;
}

// NT threads hold PER keys each, element i of thread t = column t + NT i (a warp's 32 lanes = one mask word).
template <int NT, int PER, int MINB = 1>
__global__ void __launch_bounds__(NT, MINB) sn_rowmed_mask_kernel(const SnWave w) {
    __shared__ __align__(16) RowSel s;
    const int item = blockIdx.y, row = blockIdx.x;
    const SnItem im = w.items[item];
    const int cols = im.cols;
    if (cols > NT * PER) return;                            // served by a wider instantiation (launch_sn_rowmed_mask)
    const int tid = threadIdx.x, lane = tid & 31;
    const unsigned* rowp = reinterpret_cast<const unsigned*>(w.spec + (size_t)item * w.spec_stride) + (size_t)row * im.pitch;
    unsigned key[PER];
    unsigned mn = 0xffffffffu, mx = 0u;
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        const int c = tid + NT * i;
        key[i] = c < cols ? __ldg(rowp + c) : 0u;
    }
#pragma unroll
    for (int i = 0; i < PER; ++i)
        if (tid + NT * i < cols) { mn = min(mn, key[i]); mx = max(mx, key[i]); }
    cta_minmax<NT>(s, mn, mx);
    mn = s.mn;
    mx = s.mx;
    const unsigned k0 = (unsigned)((cols - 1) / 2);
    unsigned v1 = mx;
    if (mn != mx) {                                         // uniform for the CTA
        int remaining = 32 - __clz(mn ^ mx);                // bits [0, remaining) vary
        unsigned prefix = remaining >= 32 ? 0u : (mx >> remaining) << remaining;
        unsigned k = k0;
#pragma unroll 1
        while (remaining > 0) {
            const int wd = min(kSelCtaBits, remaining), shift = remaining - wd;
            const unsigned dmask = (1u << wd) - 1u;
            const unsigned above = remaining >= 32 ? 0u : 0xffffffffu << remaining;   // bits fixed so far
            uint4* h4 = reinterpret_cast<uint4*>(s.hist);
            for (int i = tid; i < kSelCtaBins / 4; i += NT) h4[i] = make_uint4(0u, 0u, 0u, 0u);
            __syncthreads();
#pragma unroll
            for (int i = 0; i < PER; ++i)
                if (tid + NT * i < cols && ((key[i] ^ prefix) & above) == 0u) atomicAdd(s.hist + ((key[i] >> shift) & dmask), 1u);
            __syncthreads();
            sel_scan_cta<NT>(s.hist, k, s.tot, s.out);
            prefix |= s.out[0] << shift;
            k = s.out[1];
            remaining = shift;
            __syncthreads();                                // s.out is read before the next pass rewrites it
        }
        v1 = prefix;
    }
    if ((cols & 1) == 0) {
        if (tid == 0) { s.cnt_le = 0u; s.min_gt = 0xffffffffu; }
        __syncthreads();
        unsigned cnt = 0u, gt = 0xffffffffu;
#pragma unroll
        for (int i = 0; i < PER; ++i)
            if (tid + NT * i < cols) {
                if (key[i] <= v1) ++cnt; else gt = min(gt, key[i]);
            }
        cnt = __reduce_add_sync(0xffffffffu, cnt);
        gt = __reduce_min_sync(0xffffffffu, gt);
        if (lane == 0) { atomicAdd(&s.cnt_le, cnt); atomicMin(&s.min_gt, gt); }
        __syncthreads();
    }
    const float a = __uint_as_float(w.amax[item]);
    const float tr = unscaled_threshold(__fmul_rn(3.0f, scaled_median(v1, s.cnt_le, s.min_gt, cols, a)), a);
    const float* thr = w.colthr + (size_t)item * w.col_stride;
    unsigned* bits = w.bits + (size_t)item * w.bits_stride + (size_t)row * im.wpr;
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        const int c = tid + NT * i;
        bool on = false;
        if (c < cols) {
            const float x = __uint_as_float(key[i]);
            on = x > __ldg(thr + c) && x > tr;           // == fl(x / a) > 3 column median && fl(x / a) > 3 row median
        }
        const unsigned word = __ballot_sync(0xffffffffu, on);
        if (lane == 0 && (c >> 5) < im.wpr) bits[c >> 5] = word;
    }
}

// Rows longer than the register-resident variants cover (recordings of more than ~4 minutes at hop 281):
// the same select with the row re-read from global memory (L2) in every pass.
__global__ void __launch_bounds__(256) sn_rowmed_mask_long_kernel(const SnWave w) {
    constexpr int NT = 256;
    __shared__ __align__(16) RowSel s;
    const int item = blockIdx.y, row = blockIdx.x;
    const SnItem im = w.items[item];
    const int cols = im.cols;
    if (cols <= kSnMaxColsFast) return;
    const int tid = threadIdx.x, lane = tid & 31;
    const unsigned* rowp = reinterpret_cast<const unsigned*>(w.spec + (size_t)item * w.spec_stride) + (size_t)row * im.pitch;
    unsigned mn = 0xffffffffu, mx = 0u;
    for (int c = tid; c < cols; c += NT) { const unsigned key = rowp[c]; mn = min(mn, key); mx = max(mx, key); }
    cta_minmax<NT>(s, mn, mx);
    mn = s.mn;
    mx = s.mx;
    const unsigned k0 = (unsigned)((cols - 1) / 2);
    unsigned v1 = mx;
    if (mn != mx) {
        int remaining = 32 - __clz(mn ^ mx);
        unsigned prefix = remaining >= 32 ? 0u : (mx >> remaining) << remaining;
        unsigned k = k0;
        while (remaining > 0) {
            const int wd = min(kSelCtaBits, remaining), shift = remaining - wd;
            const unsigned dmask = (1u << wd) - 1u;
            const unsigned above = remaining >= 32 ? 0u : 0xffffffffu << remaining;
            uint4* h4 = reinterpret_cast<uint4*>(s.hist);
            for (int i = tid; i < kSelCtaBins / 4; i += NT) h4[i] = make_uint4(0u, 0u, 0u, 0u);
            __syncthreads();
            for (int c = tid; c < cols; c += NT) {
                const unsigned key = rowp[c];
                if (((key ^ prefix) & above) == 0u) atomicAdd(s.hist + ((key >> shift) & dmask), 1u);
            }
            __syncthreads();
            sel_scan_cta<NT>(s.hist, k, s.tot, s.out);
            prefix |= s.out[0] << shift;
            k = s.out[1];
            remaining = shift;
            __syncthreads();
        }
        v1 = prefix;
    }
    if ((cols & 1) == 0) {
        if (tid == 0) { s.cnt_le = 0u; s.min_gt = 0xffffffffu; }
        __syncthreads();
        unsigned cnt = 0u, gt = 0xffffffffu;
        for (int c = tid; c < cols; c += NT) {
            const unsigned key = rowp[c];
            if (key <= v1) ++cnt; else gt = min(gt, key);
        }
        cnt = __reduce_add_sync(0xffffffffu, cnt);
        gt = __reduce_min_sync(0xffffffffu, gt);
        if (lane == 0) { atomicAdd(&s.cnt_le, cnt); atomicMin(&s.min_gt, gt); }
        __syncthreads();
    }
    const float a = __uint_as_float(w.amax[item]);
    const float tr = unscaled_threshold(__fmul_rn(3.0f, scaled_median(v1, s.cnt_le, s.min_gt, cols, a)), a);
    const float* thr = w.colthr + (size_t)item * w.col_stride;
    unsigned* bits = w.bits + (size_t)item * w.bits_stride + (size_t)row * im.wpr;
    for (int c0 = 0; c0 < cols; c0 += NT) {                 // uniform trip count: every lane reaches the ballot
        const int c = c0 + tid;
        bool on = false;
        if (c < cols) {
            const float x = __uint_as_float(rowp[c]);
            on = x > thr[c] && x > tr;
        }
        const unsigned word = __ballot_sync(0xffffffffu, on);
        if (lane == 0 && (c >> 5) < im.wpr) bits[c >> 5] = word;
    }
}

cudaError_t launch_sn_col_thresholds(const SnWave& w, int max_cols, cudaStream_t st) {
    dim3 grid((unsigned)((max_cols + 255) / 256), (unsigned)w.n_items);
    sn_col_thresholds_kernel<<<grid, 256, 0, st>>>(w);
    return cudaGetLastError();
}

// max_cols = upper bound of the wave's row length: picks the narrowest register-resident variant that covers it
// (recordings of a wave that are much shorter just predicate the tail off)
cudaError_t launch_sn_rowmed_mask(const SnWave& w, int max_cols, cudaStream_t st) {
    dim3 grid((unsigned)w.rows, (unsigned)w.n_items);
    if (max_cols <= 256 * 8) sn_rowmed_mask_kernel<256, 8><<<grid, 256, 0, st>>>(w);
    else if (max_cols <= 256 * 16) sn_rowmed_mask_kernel<256, 16><<<grid, 256, 0, st>>>(w);
    else if (max_cols <= 256 * 32) sn_rowmed_mask_kernel<256, 32><<<grid, 256, 0, st>>>(w);
    else if (max_cols <= 256 * 44) {                      // one-minute recordings: 10,250 columns
        // 512 threads x 22 keys at two CTAs per SM: measured fastest of {256 x 44, 512 x 22, 1024 x 11} (0.38 / 0.28 /
        // 0.30-0.34 ms per recording for the whole chain, profiles/r3_experiments.md); CAC_SN_ROWS_VARIANT picks another
        static const int variant = getenv("CAC_SN_ROWS_VARIANT") ? atoi(getenv("CAC_SN_ROWS_VARIANT")) : 1;
        if (variant == 0) sn_rowmed_mask_kernel<256, 44><<<grid, 256, 0, st>>>(w);
        else if (variant == 2) sn_rowmed_mask_kernel<1024, 11, 1><<<grid, 1024, 0, st>>>(w);
        else if (variant == 3) sn_rowmed_mask_kernel<1024, 11, 2><<<grid, 1024, 0, st>>>(w);
        else sn_rowmed_mask_kernel<512, 22, 2><<<grid, 512, 0, st>>>(w);
    }
    else {
        sn_rowmed_mask_kernel<512, 48><<<grid, 512, 0, st>>>(w);         // rows up to 24,576 columns
        if (max_cols > kSnMaxColsFast) sn_rowmed_mask_long_kernel<<<grid, 256, 0, st>>>(w);   // the longer ones
    }
    return cudaGetLastError();
}

}  // namespace cac

// signal_noise's row medians and threshold mask for a wave of recordings (src/identify_tracks.py:657-667):
//   spectogram = spectogram / a_max
//   row_medians = np.median(spectogram, axis=1)
//   signal = (spectogram > 3 * column_medians) & (spectogram > 3 * row_medians)
// One CTA per (row, recording).  The row is read from HBM ONCE into registers; the median is selected there by a
// four-pass 8-bit radix select on the unscaled values (division by a_max is monotone: the order statistic of the
// quotients is the quotient of the order statistic), the one or two selected values are scaled, and the mask of the
// row is formed on the spot from fl(x / a_max) - bit-identical to the reference's arithmetic - and written bit-packed
// (32 cells per ballot).  The scaled matrix, the transposed matrix and the separate mask pass of the first version
// (5 reads + 2 writes of the 84 MB matrix per minute of audio) are gone.
#include "cac_common.cuh"
#include "sn.cuh"
#include "sn_select.cuh"

namespace cac {

// colthr[c] = fl(3 * fl(colmed[c] / a_max)): the column threshold, once per column instead of once per cell
__global__ void __launch_bounds__(256) sn_col_thresholds_kernel(const SnWave w) {
    const int item = blockIdx.y;
    const int cols = w.items[item].cols;
    const int c = blockIdx.x * 256 + threadIdx.x;
    if (c >= cols) return;
    const float a = __uint_as_float(w.amax[item]);
    const size_t o = (size_t)item * w.col_stride + c;
    w.colthr[o] = __fmul_rn(3.0f, __fdiv_rn(w.colmed[o], a));
}

// shared tail: median of the scaled row from the selected order statistics, then the row's mask words.
// NT threads hold PER keys each, element i of thread t = column t + NT i (a warp's 32 lanes = one mask word).
template <int NT, int PER>
__global__ void __launch_bounds__(NT) sn_rowmed_mask_kernel(const SnWave w) {
    __shared__ __align__(16) unsigned hist[kSelHistWords];
    __shared__ unsigned s_prefix, s_k, s_cnt_le, s_min_gt;
    const int item = blockIdx.y, row = blockIdx.x;
    const SnItem im = w.items[item];
    const int cols = im.cols;
    if (cols > NT * PER) return;                            // served by a wider instantiation (launch_sn_rowmed_mask)
    const int tid = threadIdx.x, lane = tid & 31;
    const unsigned* rowp = reinterpret_cast<const unsigned*>(w.spec + (size_t)item * w.spec_stride) + (size_t)row * cols;
    unsigned key[PER];
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        const int c = tid + NT * i;
        key[i] = c < cols ? __ldg(rowp + c) : 0xffffffffu;
    }
    const unsigned k0 = (unsigned)((cols - 1) / 2);
    unsigned prefix = 0u, k = k0;
    unsigned* hl = hist + (lane & 3);
#pragma unroll 1
    for (int pass = 3; pass >= 0; --pass) {
        const int shift = 8 * pass;
        const unsigned himask = pass == 3 ? 0u : (0xffffffffu << (shift + 8));
        for (int i = tid; i < kSelHistWords; i += NT) hist[i] = 0u;
        __syncthreads();
#pragma unroll
        for (int i = 0; i < PER; ++i)
            if (tid + NT * i < cols && (key[i] & himask) == prefix) atomicAdd(hl + ((key[i] >> shift) & 255u) * 4, 1u);
        __syncthreads();
        if (tid < 32) {
            unsigned kk = k;
            const unsigned digit = sel_scan_digit(hist, lane, kk);
            if (lane == 0) { s_prefix = prefix | (digit << shift); s_k = kk; }
        }
        __syncthreads();
        prefix = s_prefix;
        k = s_k;
    }
    const unsigned v1 = prefix;
    const float a = __uint_as_float(w.amax[item]);
    float med = __fdiv_rn(__uint_as_float(v1), a);
    if ((cols & 1) == 0) {
        // even count: the next order statistic is v1 again when enough elements are <= v1, else the smallest one above
        if (tid == 0) { s_cnt_le = 0u; s_min_gt = 0xffffffffu; }
        __syncthreads();
        unsigned cnt = 0u, mn = 0xffffffffu;
#pragma unroll
        for (int i = 0; i < PER; ++i)
            if (tid + NT * i < cols) {
                if (key[i] <= v1) ++cnt; else mn = min(mn, key[i]);
            }
        cnt = __reduce_add_sync(0xffffffffu, cnt);
        mn = __reduce_min_sync(0xffffffffu, mn);
        if (lane == 0) { atomicAdd(&s_cnt_le, cnt); atomicMin(&s_min_gt, mn); }
        __syncthreads();
        const unsigned v2 = (s_cnt_le >= k0 + 2u) ? v1 : s_min_gt;
        // np.median -> np.mean of the two middle float32 values (of the SCALED row)
        med = __fdiv_rn(__fadd_rn(med, __fdiv_rn(__uint_as_float(v2), a)), 2.0f);
    }
    const float tr = __fmul_rn(3.0f, med);
    const float* thr = w.colthr + (size_t)item * w.col_stride;
    unsigned* bits = w.bits + (size_t)item * w.bits_stride + (size_t)row * im.wpr;
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        const int c = tid + NT * i;
        bool on = false;
        if (c < cols) {
            const float xn = __fdiv_rn(__uint_as_float(key[i]), a);
            on = xn > __ldg(thr + c) && xn > tr;
        }
        const unsigned word = __ballot_sync(0xffffffffu, on);
        if (lane == 0 && (c >> 5) < im.wpr) bits[c >> 5] = word;
    }
}

// Rows longer than the register-resident variants cover (recordings of more than ~2.4 minutes at hop 281):
// the same select with the row re-read from global memory (L2) in every pass.
__global__ void __launch_bounds__(256) sn_rowmed_mask_long_kernel(const SnWave w) {
    __shared__ __align__(16) unsigned hist[kSelHistWords];
    __shared__ unsigned s_prefix, s_k, s_cnt_le, s_min_gt;
    const int item = blockIdx.y, row = blockIdx.x;
    const SnItem im = w.items[item];
    const int cols = im.cols;
    if (cols <= kSnMaxColsFast) return;
    const int tid = threadIdx.x, lane = tid & 31;
    const unsigned* rowp = reinterpret_cast<const unsigned*>(w.spec + (size_t)item * w.spec_stride) + (size_t)row * cols;
    const unsigned k0 = (unsigned)((cols - 1) / 2);
    unsigned prefix = 0u, k = k0;
    unsigned* hl = hist + (lane & 3);
    for (int pass = 3; pass >= 0; --pass) {
        const int shift = 8 * pass;
        const unsigned himask = pass == 3 ? 0u : (0xffffffffu << (shift + 8));
        for (int i = tid; i < kSelHistWords; i += 256) hist[i] = 0u;
        __syncthreads();
        for (int c = tid; c < cols; c += 256) {
            const unsigned key = rowp[c];
            if ((key & himask) == prefix) atomicAdd(hl + ((key >> shift) & 255u) * 4, 1u);
        }
        __syncthreads();
        if (tid < 32) {
            unsigned kk = k;
            const unsigned digit = sel_scan_digit(hist, lane, kk);
            if (lane == 0) { s_prefix = prefix | (digit << shift); s_k = kk; }
        }
        __syncthreads();
        prefix = s_prefix;
        k = s_k;
    }
    const unsigned v1 = prefix;
    const float a = __uint_as_float(w.amax[item]);
    float med = __fdiv_rn(__uint_as_float(v1), a);
    if ((cols & 1) == 0) {
        if (tid == 0) { s_cnt_le = 0u; s_min_gt = 0xffffffffu; }
        __syncthreads();
        unsigned cnt = 0u, mn = 0xffffffffu;
        for (int c = tid; c < cols; c += 256) {
            const unsigned key = rowp[c];
            if (key <= v1) ++cnt; else mn = min(mn, key);
        }
        cnt = __reduce_add_sync(0xffffffffu, cnt);
        mn = __reduce_min_sync(0xffffffffu, mn);
        if (lane == 0) { atomicAdd(&s_cnt_le, cnt); atomicMin(&s_min_gt, mn); }
        __syncthreads();
        const unsigned v2 = (s_cnt_le >= k0 + 2u) ? v1 : s_min_gt;
        med = __fdiv_rn(__fadd_rn(med, __fdiv_rn(__uint_as_float(v2), a)), 2.0f);
    }
    const float tr = __fmul_rn(3.0f, med);
    const float* thr = w.colthr + (size_t)item * w.col_stride;
    unsigned* bits = w.bits + (size_t)item * w.bits_stride + (size_t)row * im.wpr;
    for (int c0 = 0; c0 < cols; c0 += 256) {                // uniform trip count: every lane reaches the ballot
        const int c = c0 + tid;
        bool on = false;
        if (c < cols) {
            const float xn = __fdiv_rn(__uint_as_float(rowp[c]), a);
            on = xn > thr[c] && xn > tr;
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
    else if (max_cols <= 256 * 48) sn_rowmed_mask_kernel<256, 48><<<grid, 256, 0, st>>>(w);
    else {
        sn_rowmed_mask_kernel<512, 48><<<grid, 512, 0, st>>>(w);         // rows up to 24,576 columns
        if (max_cols > kSnMaxColsFast) sn_rowmed_mask_long_kernel<<<grid, 256, 0, st>>>(w);   // the longer ones
    }
    return cudaGetLastError();
}

}  // namespace cac

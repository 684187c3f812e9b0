// Generic pieces of signal_noise's median stage (src/identify_tracks.py:658-659) for a spectrogram the CALLER
// supplies (cac_signal_components_run): a tiled transpose and an exact row median, i.e. column medians as row
// medians of the transposed matrix.  The batched chain does not use them: its column medians come out of the
// spectrum kernel (sn_stft.cu) and its row medians are fused with the mask (sn_rows.cu).
#include "cac_common.cuh"
#include "track_signals.cuh"

namespace cac {

// ---- out[c][r] = in[r][c] ------------------------------------------------------------------------
__global__ void __launch_bounds__(256) transpose_kernel(const float* __restrict__ in, float* __restrict__ out, int rows,
                                                        long long cols) {
    __shared__ float tile[32][33];
    const long long c0 = (long long)blockIdx.x * 32;
    const int r0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int j = ty; j < 32; j += 8)
        if (r0 + j < rows && c0 + tx < cols) tile[j][tx] = in[(size_t)(r0 + j) * cols + c0 + tx];
    __syncthreads();
    for (int j = ty; j < 32; j += 8)
        if (c0 + j < cols && r0 + tx < rows) out[(size_t)(c0 + j) * rows + r0 + tx] = tile[tx][j];
}

// ---- np.median of every row (values >= 0: the float bit pattern orders like the value) -----------
// One CTA per row; 8-bit radix select, four passes over the row (L2-resident), then one pass for the
// upper middle element when the count is even.
__global__ void __launch_bounds__(256) row_median_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                         long long cols) {
    __shared__ unsigned hist[256];
    __shared__ unsigned s_prefix, s_k;
    __shared__ unsigned s_cnt_le, s_min_gt;
    const unsigned* row = reinterpret_cast<const unsigned*>(in) + (size_t)blockIdx.x * cols;
    const int tid = threadIdx.x;
    if (tid == 0) { s_prefix = 0u; s_k = (unsigned)((cols - 1) / 2); }
    for (int pass = 3; pass >= 0; --pass) {
        hist[tid] = 0u;
        __syncthreads();
        const unsigned prefix = s_prefix;
        const unsigned himask = (pass == 3) ? 0u : (0xffffffffu << (8 * (pass + 1)));
        for (long long i = tid; i < cols; i += 256) {
            const unsigned key = row[i];
            if ((key & himask) == prefix) atomicAdd(&hist[(key >> (8 * pass)) & 255u], 1u);
        }
        __syncthreads();
        if (tid == 0) {
            unsigned k = s_k, b = 0;
            while (b < 255u && hist[b] <= k) { k -= hist[b]; ++b; }
            s_k = k;
            s_prefix = prefix | (b << (8 * pass));
        }
        __syncthreads();
    }
    const unsigned v1 = s_prefix;
    if ((cols & 1) == 1) {
        if (tid == 0) out[blockIdx.x] = __uint_as_float(v1);
        return;
    }
    // even count: the next order statistic is v1 again when enough elements are <= v1, else the smallest one above
    if (tid == 0) { s_cnt_le = 0u; s_min_gt = 0xffffffffu; }
    __syncthreads();
    unsigned cnt = 0, mn = 0xffffffffu;
    for (long long i = tid; i < cols; i += 256) {
        const unsigned key = row[i];
        if (key <= v1) ++cnt; else mn = min(mn, key);
    }
    atomicAdd(&s_cnt_le, cnt);
    atomicMin(&s_min_gt, mn);
    __syncthreads();
    if (tid == 0) {
        const unsigned klo = (unsigned)((cols - 1) / 2);
        const unsigned v2 = (s_cnt_le >= klo + 2u) ? v1 : s_min_gt;
        // np.median -> np.mean of the two middle float32 values
        out[blockIdx.x] = __fdiv_rn(__fadd_rn(__uint_as_float(v1), __uint_as_float(v2)), 2.0f);
    }
}

// ---- launchers --------------------------------------------------------------------------------------
cudaError_t launch_transpose(const float* in, float* out, int rows, long long cols, cudaStream_t st) {
    dim3 grid((unsigned)((cols + 31) / 32), (unsigned)((rows + 31) / 32));
    transpose_kernel<<<grid, 256, 0, st>>>(in, out, rows, cols);
    return cudaGetLastError();
}

cudaError_t launch_row_median(const float* in, float* out, long long rows, long long cols, cudaStream_t st) {
    row_median_kernel<<<(unsigned)rows, 256, 0, st>>>(in, out, cols);
    return cudaGetLastError();
}

}  // namespace cac

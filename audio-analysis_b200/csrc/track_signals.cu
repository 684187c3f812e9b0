// signal_noise after its STFT (src/identify_tracks.py:658-691): exact row / column medians of the
// normalised magnitude spectrogram, the 3x-median mask, cv2-compatible rectangular morphology
// (4x4 open, dilate (height, width), erode (height // 10, width)) and 8-connected components with
// bounding boxes.  Everything is small integer / byte work on a [2049 x frames] matrix that is
// already resident in HBM (84 MB per minute, mostly L2-resident): plain coalesced kernels.
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

// ---- signal = (spec > 3 * column_medians) & (spec > 3 * row_medians)  (:667) ---------------------
__global__ void __launch_bounds__(256) signal_mask_kernel(const float* __restrict__ spec, const float* __restrict__ rowmed,
                                                          const float* __restrict__ colmed, unsigned char* __restrict__ mask,
                                                          int rows, long long cols) {
    const long long c = (long long)blockIdx.x * 256 + threadIdx.x;
    const int r = blockIdx.y;
    if (c >= cols) return;
    const float v = spec[(size_t)r * cols + c];
    const float tc = __fmul_rn(3.0f, colmed[c]), tr = __fmul_rn(3.0f, rowmed[r]);
    mask[(size_t)r * cols + c] = (v > tc && v > tr) ? 1 : 0;
}

// ---- cv2.erode / cv2.dilate with a rectangular all-ones kernel, separable passes -----------------
// out(y, x) = min|max over d in [-anchor, k - 1 - anchor] of in(.. + d), pixels outside the image ignored
// (cv2's default border value for morphology); anchor = k / 2 (cv2's default centre).
template <bool IS_MAX>
__global__ void __launch_bounds__(256) morph_h_kernel(const unsigned char* __restrict__ in, unsigned char* __restrict__ out,
                                                      int rows, long long cols, int k) {
    extern __shared__ unsigned char seg[];          // 256 + k pixels of this row
    const int anchor = k / 2;
    const long long c0 = (long long)blockIdx.x * 256;
    const int r = blockIdx.y;
    const unsigned char* row = in + (size_t)r * cols;
    const unsigned char neutral = IS_MAX ? 0 : 255;
    for (int i = threadIdx.x; i < 256 + k; i += 256) {
        const long long c = c0 - anchor + i;
        seg[i] = (c >= 0 && c < cols) ? row[c] : neutral;
    }
    __syncthreads();
    const long long c = c0 + threadIdx.x;
    if (c >= cols) return;
    unsigned char v = neutral;
    for (int d = 0; d < k; ++d) {
        const unsigned char s = seg[threadIdx.x + d];
        v = IS_MAX ? max(v, s) : min(v, s);
    }
    out[(size_t)r * cols + c] = v;
}

template <bool IS_MAX>
__global__ void __launch_bounds__(256) morph_v_kernel(const unsigned char* __restrict__ in, unsigned char* __restrict__ out,
                                                      int rows, long long cols, int k) {
    const int anchor = k / 2;
    const long long c = (long long)blockIdx.x * 256 + threadIdx.x;
    const int r = blockIdx.y;
    if (c >= cols) return;
    unsigned char v = IS_MAX ? 0 : 255;
    for (int d = 0; d < k; ++d) {
        const int rr = r - anchor + d;
        if (rr >= 0 && rr < rows) {
            const unsigned char s = in[(size_t)rr * cols + c];
            v = IS_MAX ? max(v, s) : min(v, s);
        }
    }
    out[(size_t)r * cols + c] = v;
}

// ---- 8-connected components: union-find on pixel indices ------------------------------------------
__device__ __forceinline__ int ccl_find(const int* label, int a) {
    int p = label[a];
    while (p != a) { a = p; p = label[a]; }
    return a;
}
__device__ __forceinline__ void ccl_union(int* label, int a, int b) {
    while (true) {
        a = ccl_find(label, a);
        b = ccl_find(label, b);
        if (a == b) return;
        if (a < b) { const int t = a; a = b; b = t; }     // a > b: hang the larger root under the smaller
        const int old = atomicMin(&label[a], b);
        if (old == a) return;
        a = old;
    }
}

__global__ void __launch_bounds__(256) ccl_init_kernel(const unsigned char* __restrict__ img, int* __restrict__ label,
                                                       long long n) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i < n) label[i] = img[i] ? (int)i : -1;
}

__global__ void __launch_bounds__(256) ccl_merge_kernel(const unsigned char* __restrict__ img, int* __restrict__ label,
                                                        int rows, int cols) {
    const int x = blockIdx.x * 256 + threadIdx.x;
    const int y = blockIdx.y;
    if (x >= cols) return;
    const int p = y * cols + x;
    if (!img[p]) return;
    if (x > 0 && img[p - 1]) ccl_union(label, p, p - 1);
    if (y > 0) {
        const int q = p - cols;
        if (img[q]) ccl_union(label, p, q);
        if (x > 0 && img[q - 1]) ccl_union(label, p, q - 1);
        if (x + 1 < cols && img[q + 1]) ccl_union(label, p, q + 1);
    }
}

// roots get consecutive ids (arbitrary order); rootid[p] is valid for root pixels only
__global__ void __launch_bounds__(256) ccl_roots_kernel(const int* __restrict__ label, int* __restrict__ rootid,
                                                        long long n, int* __restrict__ counter, int max_components) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i >= n || label[i] != (int)i) return;
    const int id = atomicAdd(counter, 1);
    rootid[i] = id < max_components ? id : -1;
}

// stats[id] = {min x, min y, max x, max y, area, first 2x2 block in raster order}
__global__ void __launch_bounds__(256) ccl_stats_kernel(const int* __restrict__ label, const int* __restrict__ rootid,
                                                        int rows, int cols, int* __restrict__ stats) {
    const int x = blockIdx.x * 256 + threadIdx.x;
    const int y = blockIdx.y;
    if (x >= cols) return;
    const int p = y * cols + x;
    if (label[p] < 0) return;
    const int id = rootid[ccl_find(label, p)];
    if (id < 0) return;
    int* s = stats + 6 * id;
    atomicMin(s + 0, x);
    atomicMin(s + 1, y);
    atomicMax(s + 2, x);
    atomicMax(s + 3, y);
    atomicAdd(s + 4, 1);
    atomicMin(s + 5, (y >> 1) * ((cols + 1) >> 1) + (x >> 1));
}

__global__ void __launch_bounds__(256) ccl_stats_init_kernel(int* stats, int max_components) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= max_components) return;
    int* s = stats + 6 * i;
    s[0] = 0x7fffffff; s[1] = 0x7fffffff; s[2] = -1; s[3] = -1; s[4] = 0; s[5] = 0x7fffffff;
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

cudaError_t launch_signal_mask(const float* spec, const float* rowmed, const float* colmed, unsigned char* mask, int rows,
                               long long cols, cudaStream_t st) {
    dim3 grid((unsigned)((cols + 255) / 256), (unsigned)rows);
    signal_mask_kernel<<<grid, 256, 0, st>>>(spec, rowmed, colmed, mask, rows, cols);
    return cudaGetLastError();
}

cudaError_t launch_morph_rect(const unsigned char* in, unsigned char* tmp, unsigned char* out, int rows, long long cols,
                              int kh, int kw, bool is_max, cudaStream_t st) {
    dim3 grid((unsigned)((cols + 255) / 256), (unsigned)rows);
    if (is_max) {
        morph_h_kernel<true><<<grid, 256, 256 + kw, st>>>(in, tmp, rows, cols, kw);
        morph_v_kernel<true><<<grid, 256, 0, st>>>(tmp, out, rows, cols, kh);
    } else {
        morph_h_kernel<false><<<grid, 256, 256 + kw, st>>>(in, tmp, rows, cols, kw);
        morph_v_kernel<false><<<grid, 256, 0, st>>>(tmp, out, rows, cols, kh);
    }
    return cudaGetLastError();
}

cudaError_t launch_ccl(const unsigned char* img, int* label, int* rootid, int rows, int cols, int* counter, int* stats,
                       int max_components, cudaStream_t st) {
    const long long n = (long long)rows * cols;
    dim3 grid((unsigned)((cols + 255) / 256), (unsigned)rows);
    cudaError_t e = cudaMemsetAsync(counter, 0, sizeof(int), st);
    if (e != cudaSuccess) return e;
    ccl_stats_init_kernel<<<(max_components + 255) / 256, 256, 0, st>>>(stats, max_components);
    ccl_init_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(img, label, n);
    ccl_merge_kernel<<<grid, 256, 0, st>>>(img, label, rows, cols);
    ccl_roots_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(label, rootid, n, counter, max_components);
    ccl_stats_kernel<<<grid, 256, 0, st>>>(label, rootid, rows, cols, stats);
    return cudaGetLastError();
}

}  // namespace cac

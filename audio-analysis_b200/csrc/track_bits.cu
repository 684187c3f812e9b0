// signal_noise's mask stages on a BIT-PACKED mask (src/identify_tracks.py:667-686): the 3x-median threshold,
// cv2-compatible rectangular morphology and 8-connected components with stats.  The [2049 x 10250] mask of a
// one-minute recording is 21 M cells; as bytes every stage moves 21 MB per pass and the union-find works on
// 21 M labels.  Packed 32 pixels per word it is 2.6 MB (L2-resident), the horizontal passes become funnel
// shifts, the vertical passes word-wise OR / AND, and the components are built from horizontal RUNS (a label
// per run start, unions only between runs of adjacent rows), a few thousand objects instead of millions.
// Bit i of word w of a row = pixel 32 w + i; the padding bits of the last word of a row are always 0.
#include "cac_common.cuh"
#include "track_signals.cuh"

namespace cac {

__device__ __forceinline__ unsigned valid_bits(int w, long long cols) {   // pixels of word w that exist
    const long long left = cols - 32LL * w;
    return left >= 32 ? 0xffffffffu : (left <= 0 ? 0u : ((1u << (int)left) - 1u));
}

// ---- signal = (spec > 3 * column_medians) & (spec > 3 * row_medians)  (:667), 32 pixels per ballot ----
__global__ void __launch_bounds__(256) signal_mask_bits_kernel(const float* __restrict__ spec,
                                                               const float* __restrict__ rowmed,
                                                               const float* __restrict__ colmed,
                                                               unsigned* __restrict__ bits, int rows, long long cols,
                                                               int wpr) {
    const long long c = (long long)blockIdx.x * 256 + threadIdx.x;
    const int r = blockIdx.y;
    bool on = false;
    if (c < cols) {
        const float v = spec[(size_t)r * cols + c];
        const float tc = __fmul_rn(3.0f, colmed[c]), tr = __fmul_rn(3.0f, rowmed[r]);
        on = v > tc && v > tr;
    }
    const unsigned word = __ballot_sync(0xffffffffu, on);
    const long long w = c >> 5;
    if ((threadIdx.x & 31) == 0 && w < wpr) bits[(size_t)r * wpr + w] = word;
}

// ---- cv2.erode / cv2.dilate with a 1 x k rectangle: out(x) = OR | AND over d in [0, k) of in(x - k/2 + d) ----
// pixels outside the image are ignored (cv2's morphology border).  Erosion runs as the complement of the
// dilation of the complement (outside = 0 in the complemented domain is exactly "ignored").
template <bool IS_MAX>
__global__ void __launch_bounds__(128) morph_h_bits_kernel(const unsigned* __restrict__ in, unsigned* __restrict__ out,
                                                           long long cols, int wpr, int k) {
    extern __shared__ unsigned seg[];                 // words w0 - halo .. w0 + 127 + halo of this row
    const int halo = (k + 31) / 32 + 1;
    const int w0 = blockIdx.x * 128;
    const int r = blockIdx.y;
    const unsigned* row = in + (size_t)r * wpr;
    for (int i = threadIdx.x; i < 128 + 2 * halo; i += 128) {
        const int w = w0 - halo + i;
        unsigned v = 0u;
        if (w >= 0 && w < wpr) {
            v = row[w];
            if (!IS_MAX) v = ~v & valid_bits(w, cols);
        }
        seg[i] = v;
    }
    __syncthreads();
    const int w = w0 + threadIdx.x;
    if (w >= wpr) return;
    const int anchor = k / 2;
    unsigned acc = 0u;
    // window of 32 pixels starting at pixel 32 w - anchor + d, as a funnel shift over two staged words
    int bit = 32 * (threadIdx.x + halo) - anchor;     // position inside seg, in pixels
    for (int d = 0; d < k; ++d, ++bit) {
        const int i = bit >> 5;
        acc |= __funnelshift_r(seg[i], seg[i + 1], bit & 31);
    }
    if (!IS_MAX) acc = ~acc;
    out[(size_t)r * wpr + w] = acc & valid_bits(w, cols);
}

// ---- k x 1 rectangle: OR | AND over the rows r - k/2 .. r - k/2 + k - 1 that exist --------------------
template <bool IS_MAX>
__global__ void __launch_bounds__(128) morph_v_bits_kernel(const unsigned* __restrict__ in, unsigned* __restrict__ out,
                                                           int rows, int wpr, int k) {
    const int w = blockIdx.x * 128 + threadIdx.x;
    const int r = blockIdx.y;
    if (w >= wpr) return;
    const int r0 = max(r - k / 2, 0), r1 = min(r - k / 2 + k - 1, rows - 1);
    unsigned acc = IS_MAX ? 0u : 0xffffffffu;        // AND of real rows keeps the (zero) padding bits zero
    for (int rr = r0; rr <= r1; ++rr) {
        const unsigned v = in[(size_t)rr * wpr + w];
        acc = IS_MAX ? (acc | v) : (acc & v);
    }
    out[(size_t)r * wpr + w] = acc;
}

__global__ void __launch_bounds__(256) unpack_bits_kernel(const unsigned* __restrict__ bits, unsigned char* __restrict__ out,
                                                          long long cols, int wpr) {
    const long long c = (long long)blockIdx.x * 256 + threadIdx.x;
    const int r = blockIdx.y;
    if (c >= cols) return;
    out[(size_t)r * cols + c] = (bits[(size_t)r * wpr + (c >> 5)] >> (c & 31)) & 1u;
}

// ---- 8-connected components on horizontal runs ---------------------------------------------------------
// A run is a maximal horizontal sequence of set pixels; it is named by the pixel index (y * cols + x) of its
// first pixel, and label[] holds a union-find forest over those names (only run-start entries are ever touched).
__device__ __forceinline__ int run_end(const unsigned* row, int wpr, int x) {          // last pixel of the run holding x
    int w = x >> 5;
    unsigned z = ~row[w] & (0xffffffffu << (x & 31));                                   // zero pixels at or after x
    while (z == 0u) {
        if (++w >= wpr) return 32 * wpr - 1;     // cannot happen: padding bits are 0
        z = ~row[w];
    }
    return 32 * w + __ffs(z) - 2;
}
__device__ __forceinline__ int run_start(const unsigned* row, int x) {                  // first pixel of the run holding x
    int w = x >> 5;
    const int b = x & 31;
    unsigned z = ~row[w] & (b == 31 ? 0xffffffffu : ((2u << b) - 1u));                  // zero pixels at or before x
    while (z == 0u) {
        if (--w < 0) return 0;
        z = ~row[w];
    }
    return 32 * w + (32 - __clz(z));
}
__device__ __forceinline__ int uf_find(const int* label, int a) {
    int p = label[a];
    while (p != a) { a = p; p = label[a]; }
    return a;
}
__device__ __forceinline__ void uf_union(int* label, int a, int b) {
    while (true) {
        a = uf_find(label, a);
        b = uf_find(label, b);
        if (a == b) return;
        if (a < b) { const int t = a; a = b; b = t; }     // a > b: hang the larger root under the smaller
        const int old = atomicMin(&label[a], b);
        if (old == a) return;
        a = old;
    }
}
// run starts inside word w of a row: set pixels whose left neighbour is clear
__device__ __forceinline__ unsigned start_bits(const unsigned* row, int w) {
    const unsigned v = row[w];
    const unsigned prev = w > 0 ? row[w - 1] >> 31 : 0u;
    return v & ~((v << 1) | prev);
}

__global__ void __launch_bounds__(128) runs_init_kernel(const unsigned* __restrict__ bits, int* __restrict__ label, int rows,
                                                        int cols, int wpr) {
    const int w = blockIdx.x * 128 + threadIdx.x;
    const int y = blockIdx.y;
    if (w >= wpr) return;
    unsigned s = start_bits(bits + (size_t)y * wpr, w);
    while (s) {
        const int x = 32 * w + __ffs(s) - 1;
        s &= s - 1;
        label[y * cols + x] = y * cols + x;
    }
}

// every run is joined with the runs of the row above that touch [xs - 1, xe + 1] (8-connectivity)
__global__ void __launch_bounds__(128) runs_merge_kernel(const unsigned* __restrict__ bits, int* __restrict__ label, int rows,
                                                         int cols, int wpr) {
    const int w = blockIdx.x * 128 + threadIdx.x;
    const int y = blockIdx.y + 1;
    if (w >= wpr || y >= rows) return;
    const unsigned* row = bits + (size_t)y * wpr;
    const unsigned* up = row - wpr;
    unsigned s = start_bits(row, w);
    while (s) {
        const int xs = 32 * w + __ffs(s) - 1;
        s &= s - 1;
        const int xe = run_end(row, wpr, xs);
        const int lo = max(xs - 1, 0), hi = min(xe + 1, cols - 1);
        int x = lo;
        while (x <= hi) {
            // next set pixel of the row above at or after x
            int uw = x >> 5;
            unsigned m = up[uw] & (0xffffffffu << (x & 31));
            while (m == 0u && 32 * (uw + 1) <= hi) m = up[++uw];
            if (m == 0u) break;
            const int ux = 32 * uw + __ffs(m) - 1;
            if (ux > hi) break;
            const int us = run_start(up, ux);
            uf_union(label, y * cols + xs, (y - 1) * cols + us);
            x = run_end(up, wpr, ux) + 2;            // the pixel after a run is clear
        }
    }
}

// roots get consecutive ids (arbitrary order); rootid[] is valid at root positions only
__global__ void __launch_bounds__(128) runs_roots_kernel(const unsigned* __restrict__ bits, const int* __restrict__ label,
                                                         int* __restrict__ rootid, int rows, int cols, int wpr,
                                                         int* __restrict__ counter, int max_components) {
    const int w = blockIdx.x * 128 + threadIdx.x;
    const int y = blockIdx.y;
    if (w >= wpr) return;
    unsigned s = start_bits(bits + (size_t)y * wpr, w);
    while (s) {
        const int p = y * cols + 32 * w + __ffs(s) - 1;
        s &= s - 1;
        if (label[p] == p) {
            const int id = atomicAdd(counter, 1);
            rootid[p] = id < max_components ? id : -1;
        }
    }
}

// stats[id] = {min x, min y, max x, max y, area, first 2x2 block in raster order}, one update per run
__global__ void __launch_bounds__(128) runs_stats_kernel(const unsigned* __restrict__ bits, const int* __restrict__ label,
                                                         const int* __restrict__ rootid, int rows, int cols, int wpr,
                                                         int* __restrict__ stats) {
    const int w = blockIdx.x * 128 + threadIdx.x;
    const int y = blockIdx.y;
    if (w >= wpr) return;
    const unsigned* row = bits + (size_t)y * wpr;
    unsigned s = start_bits(row, w);
    while (s) {
        const int xs = 32 * w + __ffs(s) - 1;
        s &= s - 1;
        const int xe = run_end(row, wpr, xs);
        const int id = rootid[uf_find(label, y * cols + xs)];
        if (id < 0) continue;
        int* st = stats + 6 * id;
        atomicMin(st + 0, xs);
        atomicMin(st + 1, y);
        atomicMax(st + 2, xe);
        atomicMax(st + 3, y);
        atomicAdd(st + 4, xe - xs + 1);
        atomicMin(st + 5, (y >> 1) * ((cols + 1) >> 1) + (xs >> 1));
    }
}

__global__ void __launch_bounds__(256) stats_init_kernel(int* stats, int max_components) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= max_components) return;
    int* s = stats + 6 * i;
    s[0] = 0x7fffffff; s[1] = 0x7fffffff; s[2] = -1; s[3] = -1; s[4] = 0; s[5] = 0x7fffffff;
}

// ---- launchers --------------------------------------------------------------------------------------
cudaError_t launch_signal_mask_bits(const float* spec, const float* rowmed, const float* colmed, unsigned* bits, int rows,
                                    long long cols, cudaStream_t st) {
    const int wpr = (int)((cols + 31) / 32);
    dim3 grid((unsigned)((cols + 255) / 256), (unsigned)rows);
    signal_mask_bits_kernel<<<grid, 256, 0, st>>>(spec, rowmed, colmed, bits, rows, cols, wpr);
    return cudaGetLastError();
}

cudaError_t launch_morph_rect_bits(const unsigned* in, unsigned* tmp, unsigned* out, int rows, long long cols, int kh,
                                   int kw, bool is_max, cudaStream_t st) {
    const int wpr = (int)((cols + 31) / 32);
    dim3 grid((unsigned)((wpr + 127) / 128), (unsigned)rows);
    const size_t smem = (128 + 2 * ((kw + 31) / 32 + 1)) * sizeof(unsigned);
    if (is_max) {
        morph_h_bits_kernel<true><<<grid, 128, smem, st>>>(in, tmp, cols, wpr, kw);
        morph_v_bits_kernel<true><<<grid, 128, 0, st>>>(tmp, out, rows, wpr, kh);
    } else {
        morph_h_bits_kernel<false><<<grid, 128, smem, st>>>(in, tmp, cols, wpr, kw);
        morph_v_bits_kernel<false><<<grid, 128, 0, st>>>(tmp, out, rows, wpr, kh);
    }
    return cudaGetLastError();
}

cudaError_t launch_unpack_bits(const unsigned* bits, unsigned char* out, int rows, long long cols, cudaStream_t st) {
    const int wpr = (int)((cols + 31) / 32);
    dim3 grid((unsigned)((cols + 255) / 256), (unsigned)rows);
    unpack_bits_kernel<<<grid, 256, 0, st>>>(bits, out, cols, wpr);
    return cudaGetLastError();
}

cudaError_t launch_ccl_runs(const unsigned* bits, int* label, int* rootid, int rows, int cols, int* counter, int* stats,
                            int max_components, cudaStream_t st) {
    const int wpr = (cols + 31) / 32;
    dim3 grid((unsigned)((wpr + 127) / 128), (unsigned)rows);
    cudaError_t e = cudaMemsetAsync(counter, 0, sizeof(int), st);
    if (e != cudaSuccess) return e;
    stats_init_kernel<<<(max_components + 255) / 256, 256, 0, st>>>(stats, max_components);
    runs_init_kernel<<<grid, 128, 0, st>>>(bits, label, rows, cols, wpr);
    if (rows > 1) {
        dim3 gridm((unsigned)((wpr + 127) / 128), (unsigned)(rows - 1));
        runs_merge_kernel<<<gridm, 128, 0, st>>>(bits, label, rows, cols, wpr);
    }
    runs_roots_kernel<<<grid, 128, 0, st>>>(bits, label, rootid, rows, cols, wpr, counter, max_components);
    runs_stats_kernel<<<grid, 128, 0, st>>>(bits, label, rootid, rows, cols, wpr, stats);
    return cudaGetLastError();
}

}  // namespace cac

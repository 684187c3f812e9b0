// signal_noise's mask stages on a BIT-PACKED mask (src/identify_tracks.py:667-686): the 3x-median threshold,
// cv2-compatible rectangular morphology and 8-connected components with stats.  The [2049 x 10250] mask of a
// one-minute recording is 21 M cells; as bytes every stage moves 21 MB per pass and the union-find works on
// 21 M labels.  Packed 32 pixels per word it is 2.6 MB (L2-resident), the horizontal passes become funnel
// shifts, the vertical passes word-wise OR / AND, and the components are built from horizontal RUNS (a label
// per run start, unions only between runs of adjacent rows), a few thousand objects instead of millions.
// Bit i of word w of a row = pixel 32 w + i; the padding bits of the last word of a row are always 0.
// Every kernel works on a WAVE of recordings (sn.cuh): recording = blockIdx.z, its geometry comes from the item table,
// its masks / labels / result block from slot blockIdx.z of the wave's scratch arrays.
#include "cac_common.cuh"
#include "sn.cuh"

namespace cac {

__device__ __forceinline__ unsigned valid_bits(int w, int cols) {          // pixels of word w that exist
    const int left = cols - 32 * w;
    return left >= 32 ? 0xffffffffu : (left <= 0 ? 0u : ((1u << left) - 1u));
}

// ---- cv2.erode / cv2.dilate with a 1 x k rectangle: out(x) = OR | AND over d in [0, k) of in(x - k/2 + d) ----
// pixels outside the image are ignored (cv2's morphology border).  Erosion runs as the complement of the
// dilation of the complement (outside = 0 in the complemented domain is exactly "ignored").
// A thread owns ONE word column of a strip of kMorphStrip rows (grid: word chunks x strips x recordings): no index
// division, the per-thread set-up is paid once per strip, and the two to four input words an output needs (and, in the
// vertical pass, the k rows) come from L1 / L2 - the whole mask plane of a recording is 2.6 MB.  (The first version
// ran one thread per output word with a 64-bit division to find its row: ~150 instructions of overhead for ~20 of
// work, 70 us for the 1 x 1 "copy" pass of a wave.)
constexpr int kMorphStrip = 8, kMorphThreads = 128;

template <bool IS_MAX>
__device__ __forceinline__ unsigned morph_word(const unsigned* row, int w, int wpr, int cols) {
    if (w < 0 || w >= wpr) return 0u;
    const unsigned v = __ldg(row + w);
    return IS_MAX ? v : (~v & valid_bits(w, cols));
}
template <bool IS_MAX>
__global__ void __launch_bounds__(kMorphThreads) morph_h_bits_kernel(const SnWave wv, int src_plane, int dst_plane, int k) {
    const SnItem im = wv.items[blockIdx.z];
    const int wpr = im.wpr, cols = im.cols;
    const int w = blockIdx.x * kMorphThreads + threadIdx.x;
    if (w >= wpr) return;
    const unsigned* in = wv.bits + (size_t)blockIdx.z * wv.bits_stride + (size_t)src_plane * wv.plane_stride;
    unsigned* out = wv.bits + (size_t)blockIdx.z * wv.bits_stride + (size_t)dst_plane * wv.plane_stride;
    const unsigned vb = valid_bits(w, cols);
    // out = OR over d in [0, k) of the 32-pixel window starting at pixel 32 w - anchor + d
    const int bit0 = 32 * w - k / 2;                       // may be negative: arithmetic shift = floor division
    const int i0 = bit0 >> 5;
    const int r_begin = blockIdx.y * kMorphStrip, r_end = min(r_begin + kMorphStrip, wv.rows);
    if (k <= 64) {
        // the 32 + k - 1 pixels an output word depends on lie inside words i0 .. i0 + 2 (bit0 mod 32 + k <= 64 for every
        // k <= 64 with the anchor k / 2): OR of shifts 0 .. m - 1 by doubling (1, 2, 4, ... ), then
        // [0, k) = [0, m) u [k - m, k) for the largest power of two m <= k: ~40 operations instead of 6 per pixel of k.
        // Which of the three words exist, and which of their pixels, depends on the column only: once per thread.
        bool ok[3];
        unsigned keep[3];                                  // erosion: the pixels of the complement that exist
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            ok[j] = i0 + j >= 0 && i0 + j < wpr;
            keep[j] = ok[j] ? valid_bits(i0 + j, cols) : 0u;
        }
        int m_top = 1;
        while (2 * m_top <= k) m_top <<= 1;
        const int rest = k - m_top, sh = bit0 & 31;        // rest < m_top <= 32
        const unsigned* row = in + (size_t)r_begin * wpr + i0;
        unsigned* orow = out + (size_t)r_begin * wpr + w;
        for (int r = r_begin; r < r_end; ++r, row += wpr, orow += wpr) {
            unsigned A[4];
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                const unsigned v = ok[j] ? __ldg(row + j) : 0u;
                A[j] = IS_MAX ? v : (~v & keep[j]);
            }
            A[3] = 0u;
            for (int m = 1; m < m_top; m <<= 1) {
#pragma unroll
                for (int j = 0; j < 3; ++j) A[j] |= __funnelshift_rc(A[j], A[j + 1], m);
            }
            if (rest) {
#pragma unroll
                for (int j = 0; j < 3; ++j) A[j] |= __funnelshift_rc(A[j], A[j + 1], rest);
            }
            unsigned acc = __funnelshift_rc(A[0], A[1], sh);
            if (!IS_MAX) acc = ~acc;
            *orow = acc & vb;
        }
        return;
    }
    for (int r = r_begin; r < r_end; ++r) {
        const unsigned* row = in + (size_t)r * wpr;
        unsigned acc = 0u;
        int bit = bit0, i = i0;
        unsigned lo = morph_word<IS_MAX>(row, i, wpr, cols), hi = morph_word<IS_MAX>(row, i + 1, wpr, cols);
        for (int d = 0; d < k; ++d, ++bit) {
            if ((bit >> 5) != i) {
                i = bit >> 5;
                lo = hi;
                hi = morph_word<IS_MAX>(row, i + 1, wpr, cols);
            }
            acc |= __funnelshift_r(lo, hi, bit & 31);
        }
        if (!IS_MAX) acc = ~acc;
        out[(size_t)r * wpr + w] = acc & vb;
    }
}

// ---- k x 1 rectangle: OR | AND over the rows r - k/2 .. r - k/2 + k - 1 that exist --------------------
// KH > 0: the height is a compile-time constant - the strip's 7 + KH input rows are loaded ONCE into registers and
// every output is a fixed combination of them (the heights of the chain at its native rates: 2, 3, 4, 6, 10; a height of 1 skips this pass);
// KH = 0: any height, one loop over the rows per output.
template <bool IS_MAX, int KH>
__global__ void __launch_bounds__(kMorphThreads) morph_v_bits_kernel(const SnWave wv, int src_plane, int dst_plane, int k) {
    const int wpr = wv.items[blockIdx.z].wpr;
    const int rows = wv.rows;
    const int w = blockIdx.x * kMorphThreads + threadIdx.x;
    if (w >= wpr) return;
    const int r_begin = blockIdx.y * kMorphStrip, r_end = min(r_begin + kMorphStrip, rows);
    const unsigned* in = wv.bits + (size_t)blockIdx.z * wv.bits_stride + (size_t)src_plane * wv.plane_stride + w;
    unsigned* out = wv.bits + (size_t)blockIdx.z * wv.bits_stride + (size_t)dst_plane * wv.plane_stride + w;
    constexpr unsigned kIdentity = IS_MAX ? 0u : 0xffffffffu;   // AND of real rows keeps the (zero) padding bits zero
    if constexpr (KH > 0) {
        constexpr int NIN = kMorphStrip + KH - 1;
        unsigned L[NIN];
        const int first = r_begin - KH / 2;
#pragma unroll
        for (int t = 0; t < NIN; ++t) {
            const int rr = first + t;
            L[t] = (rr >= 0 && rr < rows) ? __ldg(in + (unsigned)(rr * wpr)) : kIdentity;
        }
#pragma unroll
        for (int j = 0; j < kMorphStrip; ++j) {
            unsigned acc = L[j];
#pragma unroll
            for (int d = 1; d < KH; ++d) acc = IS_MAX ? (acc | L[j + d]) : (acc & L[j + d]);
            if (r_begin + j < r_end) out[(unsigned)((r_begin + j) * wpr)] = acc;
        }
    } else {
        for (int r = r_begin; r < r_end; ++r) {
            const int r0 = max(r - k / 2, 0), r1 = min(r - k / 2 + k - 1, rows - 1);
            unsigned acc = kIdentity;
            for (int rr = r0; rr <= r1; ++rr) {
                const unsigned v = __ldg(in + (size_t)rr * wpr);
                acc = IS_MAX ? (acc | v) : (acc & v);
            }
            out[(size_t)r * wpr] = acc;
        }
    }
}

__global__ void __launch_bounds__(256) unpack_bits_kernel(const SnWave wv, int plane, unsigned char* __restrict__ out_all,
                                                          long long out_stride) {
    const SnItem im = wv.items[blockIdx.z];
    const long long cols = im.cols;
    const int wpr = im.wpr;
    const unsigned* bits = wv.bits + (size_t)blockIdx.z * wv.bits_stride + (size_t)plane * wv.plane_stride;
    unsigned char* out = out_all + (size_t)blockIdx.z * out_stride;
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
// label[p] >= 0: parent of run start p (p itself for a root); label[p] < 0: p is a numbered root, id = -1 - label[p]
__device__ __forceinline__ int uf_find(const int* label, int a) {
    int p = label[a];
    while (p != a) { a = p; p = label[a]; }
    return a;
}
__device__ __forceinline__ int uf_find_id(const int* label, int a) {     // after runs_roots_kernel
    int p = label[a];
    while (p >= 0) { a = p; p = label[a]; }
    return -1 - p;
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

struct CclView {                                  // recording blockIdx.z of the wave
    const unsigned* bits;
    int* label;
    int* result;
    int cols, wpr;
};
__device__ __forceinline__ CclView ccl_view(const SnWave& wv, int plane) {
    const SnItem im = wv.items[blockIdx.z];
    CclView v;
    v.bits = wv.bits + (size_t)blockIdx.z * wv.bits_stride + (size_t)plane * wv.plane_stride;
    v.label = wv.label + (size_t)blockIdx.z * wv.label_stride;
    v.result = wv.result + (size_t)blockIdx.z * wv.result_stride;
    v.cols = im.cols;
    v.wpr = im.wpr;
    return v;
}

__global__ void __launch_bounds__(128) runs_init_kernel(const SnWave wv, int plane) {
    const CclView v = ccl_view(wv, plane);
    const int w = blockIdx.x * 128 + threadIdx.x;
    if (w >= v.wpr) return;
    const int y_end = min((int)(blockIdx.y + 1) * kMorphStrip, wv.rows);
    for (int y = blockIdx.y * kMorphStrip; y < y_end; ++y) {
        unsigned s = start_bits(v.bits + (size_t)y * v.wpr, w);
        while (s) {
            const int x = 32 * w + __ffs(s) - 1;
            s &= s - 1;
            v.label[y * v.cols + x] = y * v.cols + x;
        }
    }
}

// every run is joined with the runs of the row above that touch [xs - 1, xe + 1] (8-connectivity)
__global__ void __launch_bounds__(128) runs_merge_kernel(const SnWave wv, int plane) {
    const CclView v = ccl_view(wv, plane);
    const int wpr = v.wpr, cols = v.cols;
    int* label = v.label;
    const int w = blockIdx.x * 128 + threadIdx.x;
    if (w >= wpr) return;
    const int y_end = min((int)(blockIdx.y + 1) * kMorphStrip, wv.rows);
    for (int y = max((int)blockIdx.y * kMorphStrip, 1); y < y_end; ++y) {
        const unsigned* row = v.bits + (size_t)y * wpr;
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
                x = run_end(up, wpr, ux) + 2;        // the pixel after a run is clear
            }
        }
    }
}

// roots get consecutive ids (arbitrary order), written in place as -1 - id; result[0] counts them (a count above
// max_components marks the recording's table as overflowed - its stats rows beyond the limit are never written)
__global__ void __launch_bounds__(128) runs_roots_kernel(const SnWave wv, int plane) {
    const CclView v = ccl_view(wv, plane);
    const int w = blockIdx.x * 128 + threadIdx.x;
    if (w >= v.wpr) return;
    const int y_end = min((int)(blockIdx.y + 1) * kMorphStrip, wv.rows);
    for (int y = blockIdx.y * kMorphStrip; y < y_end; ++y) {
        unsigned s = start_bits(v.bits + (size_t)y * v.wpr, w);
        while (s) {
            const int p = y * v.cols + 32 * w + __ffs(s) - 1;
            s &= s - 1;
            if (v.label[p] == p) v.label[p] = -1 - atomicAdd(v.result, 1);
        }
    }
}

// stats[id] = {min x, min y, max x, max y, area, first 2x2 block in raster order}, one update per run
__global__ void __launch_bounds__(128) runs_stats_kernel(const SnWave wv, int plane) {
    const CclView v = ccl_view(wv, plane);
    const int wpr = v.wpr, cols = v.cols;
    const int w = blockIdx.x * 128 + threadIdx.x;
    if (w >= wpr) return;
    const int y_end = min((int)(blockIdx.y + 1) * kMorphStrip, wv.rows);
    for (int y = blockIdx.y * kMorphStrip; y < y_end; ++y) {
        const unsigned* row = v.bits + (size_t)y * wpr;
        unsigned s = start_bits(row, w);
        while (s) {
            const int xs = 32 * w + __ffs(s) - 1;
            s &= s - 1;
            const int xe = run_end(row, wpr, xs);
            const int id = uf_find_id(v.label, y * cols + xs);
            if (id >= wv.max_components) continue;
            int* st = v.result + 4 + 6 * id;
            atomicMin(st + 0, xs);
            atomicMin(st + 1, y);
            atomicMax(st + 2, xe);
            atomicMax(st + 3, y);
            atomicAdd(st + 4, xe - xs + 1);
            atomicMin(st + 5, (y >> 1) * ((cols + 1) >> 1) + (xs >> 1));
        }
    }
}

__global__ void __launch_bounds__(256) stats_init_kernel(const SnWave wv) {
    int* result = wv.result + (size_t)blockIdx.y * wv.result_stride;
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i < 4) result[i] = 0;                         // [0] = component counter
    if (i >= wv.max_components) return;
    int* s = result + 4 + 6 * i;
    s[0] = 0x7fffffff; s[1] = 0x7fffffff; s[2] = -1; s[3] = -1; s[4] = 0; s[5] = 0x7fffffff;
}

// ---- launchers (whole wave per launch) ---------------------------------------------------------------------
// out = erode (is_max = false) / dilate (true) of plane `src` with a kh x kw rectangle; `tmp` is scratch
cudaError_t launch_morph_rect_bits_batch(const SnWave& w, int src_plane, int tmp_plane, int dst_plane, int max_wpr, int kh,
                                         int kw, bool is_max, cudaStream_t st) {
    // (rows ride on grid.y here and in the component kernels: matrices of up to 65,535 rows)
    dim3 grid((unsigned)((max_wpr + kMorphThreads - 1) / kMorphThreads), (unsigned)((w.rows + kMorphStrip - 1) / kMorphStrip),
              (unsigned)w.n_items);
    const int h_dst = kh == 1 ? dst_plane : tmp_plane;     // a height of 1: the horizontal pass is the whole operation
    if (is_max) morph_h_bits_kernel<true><<<grid, kMorphThreads, 0, st>>>(w, src_plane, h_dst, kw);
    else morph_h_bits_kernel<false><<<grid, kMorphThreads, 0, st>>>(w, src_plane, h_dst, kw);
    if (kh == 1) return cudaGetLastError();
#define CAC_MORPH_V(KH)                                                                                         \
    do {                                                                                                        \
        if (is_max) morph_v_bits_kernel<true, KH><<<grid, kMorphThreads, 0, st>>>(w, tmp_plane, dst_plane, kh); \
        else morph_v_bits_kernel<false, KH><<<grid, kMorphThreads, 0, st>>>(w, tmp_plane, dst_plane, kh);       \
    } while (0)
    switch (kh) {
        case 2: CAC_MORPH_V(2); break;
        case 3: CAC_MORPH_V(3); break;
        case 4: CAC_MORPH_V(4); break;
        case 6: CAC_MORPH_V(6); break;
        case 10: CAC_MORPH_V(10); break;
        default: CAC_MORPH_V(0); break;
    }
#undef CAC_MORPH_V
    return cudaGetLastError();
}

cudaError_t launch_unpack_bits_batch(const SnWave& w, int plane, unsigned char* out, long long out_stride, int max_cols,
                                     cudaStream_t st) {
    dim3 grid((unsigned)((max_cols + 255) / 256), (unsigned)w.rows, (unsigned)w.n_items);
    unpack_bits_kernel<<<grid, 256, 0, st>>>(w, plane, out, out_stride);
    return cudaGetLastError();
}

cudaError_t launch_ccl_runs_batch(const SnWave& w, int plane, int max_wpr, cudaStream_t st) {
    // a thread owns one word column of a strip of kMorphStrip rows, as in the morphology kernels
    dim3 grid((unsigned)((max_wpr + 127) / 128), (unsigned)((w.rows + kMorphStrip - 1) / kMorphStrip), (unsigned)w.n_items);
    const int init_n = w.max_components > 4 ? w.max_components : 4;
    stats_init_kernel<<<dim3((unsigned)((init_n + 255) / 256), (unsigned)w.n_items), 256, 0, st>>>(w);
    runs_init_kernel<<<grid, 128, 0, st>>>(w, plane);
    if (w.rows > 1) runs_merge_kernel<<<grid, 128, 0, st>>>(w, plane);
    runs_roots_kernel<<<grid, 128, 0, st>>>(w, plane);
    runs_stats_kernel<<<grid, 128, 0, st>>>(w, plane);
    return cudaGetLastError();
}

}  // namespace cac

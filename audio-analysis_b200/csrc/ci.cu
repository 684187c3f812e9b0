// Old Cacophony Index hot loop (src/cacophony_index.py:53-66 get_ci_bins, :86-97 calculate's loop,
// :69-73/:99-114 the 312-point score windows) in float64, as the reference computes it.
//
// get_ci_bins = numpy.hanning(2048) * window -> scipy.fftpack.dct (type II, unnormalised) -> 10
// log-spaced band energies.  One warp transforms one window: the 2048-point DCT-II is a 2048-point
// real FFT of the even/odd-reordered samples (Makhoul), computed as ONE 1024-point complex FFT
// (radix 32 x 32 in registers, one shared-memory transpose) followed by a fused untangle + quarter-
// wave twiddle that yields X[k] and X[2048-k] together; squares are accumulated into the 10 bands in
// registers (band membership is resolved at compile time for the 16 kHz / 2048 layout).
// Everything is FP64: the 2x comparisons of calculate() then agree with the reference except for
// band-energy ratios within ~1e-13 of the threshold (BASELINE.json allows 1e-6).
#include "cac_common.cuh"
#include "ci.cuh"
#include "fft_regs.cuh"

namespace cac {

constexpr int kCiWarps = 8;
constexpr int kCiExStride = 33;                       // double2 per exchange row (odd: conflict-free transpose)
constexpr int kCiExElems = 32 * kCiExStride;          // double2 per warp
// logspace(log10(25), log10(2048), 11, dtype=int) for sample_rate 16000, window 2048 (cacophony_index.py:58-63)
__host__ __device__ constexpr int ci_edge(int b) {
    constexpr int e[11] = {25, 38, 60, 93, 145, 226, 351, 546, 848, 1318, 2048};
    return e[b];
}

// acc[b] += v*v for the band that holds DCT bin k; k lies in [LO, LO+31] (compile-time range)
template <int LO, int B>
__device__ __forceinline__ void band_acc(double v, int k, double (&acc)[10]) {
    if constexpr (B < 10) {
        constexpr int e0 = ci_edge(B), e1 = ci_edge(B + 1);
        if constexpr (e0 <= LO + 31 && e1 > LO) {
            if constexpr (e0 <= LO && e1 > LO + 31) {
                acc[B] = fma(v, v, acc[B]);
            } else {
                const double m = (k >= e0 && k < e1) ? v : 0.0;
                acc[B] = fma(m, m, acc[B]);
            }
        }
        band_acc<LO, B + 1>(v, k, acc);
    }
}

__device__ __forceinline__ double shfl_d(double v, int src) {
    return __shfl_sync(0xffffffffu, v, src);
}

// pass-1 twiddle W_1024^(lane * K1) = A[K1 >> 3] * B[K1 & 7] applied to the value at out_pos(32, K1)
template <int K1>
struct CiTw1 {
    static __device__ __forceinline__ void run(double (&zr)[32], double (&zi)[32], const double2* lt) {
        if constexpr (K1 > 0) {
            constexpr int A = K1 >> 3, B = K1 & 7, P = out_pos(32, K1);
            double tr, ti;
            if constexpr (A == 0) { const double2 b = lt[(4 + B) * 32]; tr = b.x; ti = b.y; }
            else if constexpr (B == 0) { const double2 a = lt[A * 32]; tr = a.x; ti = a.y; }
            else {
                const double2 a = lt[A * 32], b = lt[(4 + B) * 32];
                tr = fma(-a.y, b.y, a.x * b.x);
                ti = fma(a.x, b.y, a.y * b.x);
            }
            const double xr = zr[P], xi = zi[P];
            zr[P] = fma(-xi, ti, xr * tr);
            zi[P] = fma(xr, ti, xi * tr);
        }
        if constexpr (K1 + 1 < 32) CiTw1<K1 + 1>::run(zr, zi, lt);
    }
};

template <int K1>
struct CiExWrite {
    static __device__ __forceinline__ void run(const double (&zr)[32], const double (&zi)[32], double2* ex_lane) {
        ex_lane[K1 * kCiExStride] = make_double2(zr[out_pos(32, K1)], zi[out_pos(32, K1)]);
        if constexpr (K1 + 1 < 32) CiExWrite<K1 + 1>::run(zr, zi, ex_lane);
    }
};
template <int N2>
struct CiExRead {
    static __device__ __forceinline__ void run(double (&zr)[32], double (&zi)[32], const double2* ex_row) {
        const double2 t = ex_row[N2];
        zr[N2] = t.x;
        zi[N2] = t.y;
        if constexpr (N2 + 1 < 32) CiExRead<N2 + 1>::run(zr, zi, ex_row);
    }
};

// Untangle + DCT twiddle for FFT bin k = lane + 32*K2 (Z at out_pos(32, K2)):
//   G = W_8192^k (Z[k] + conj Z[1024-k]) - i W_8192^(5k) (Z[k] - conj Z[1024-k])
//   X[k] = Re G,  X[2048-k] = -Im G   (X = scipy.fftpack.dct type II of the windowed samples)
template <int K2>
struct CiUntangle {
    static __device__ __forceinline__ void run(const double (&zr)[32], const double (&zi)[32], int lane, int partner,
                                               const double2* lt, double (&acc)[10]) {
        constexpr int pown = out_pos(32, K2);
        constexpr int psend = out_pos(32, 31 - K2);
        constexpr int psend0 = out_pos(32, (32 - K2) & 31);
        const double ar = zr[pown], ai = zi[pown];
        const double sr = (lane == 0) ? zr[psend0] : zr[psend];
        const double si = (lane == 0) ? zi[psend0] : zi[psend];
        const double br = shfl_d(sr, partner);
        const double bi = -shfl_d(si, partner);
        const double sumr = ar + br, sumi = ai + bi;
        const double dr = ar - br, di = ai - bi;
        // u = W_8192^k, v = -i W_8192^(5k) for k = lane + 32 K2, from the host-built table (rows 14 + 2 K2, 15 + 2 K2)
        const double2 u = lt[(14 + 2 * K2) * 32], v = lt[(15 + 2 * K2) * 32];
        const double ur = u.x, ui = u.y, vr = v.x, vi = v.y;
        const double gr = fma(vr, dr, fma(-vi, di, fma(ur, sumr, -ui * sumi)));
        const double gi = fma(vr, di, fma(vi, dr, fma(ur, sumi, ui * sumr)));
        const int k = lane + 32 * K2;
        band_acc<32 * K2, 0>(gr, k, acc);
        if constexpr (K2 > 0) {
            band_acc<2017 - 32 * K2, 0>(gi, 2048 - k, acc);    // (-gi)^2 == gi^2
        } else {
            // lane 0 holds k = 0: its mirror slot carries X[1024] = sqrt(2) (Re Z0 - Im Z0) (band 8);
            // the other lanes hold X[2048 - lane] (band 9)
            const double x1024 = 1.41421356237309504880168872420969808 * (ar - ai);
            const double m8 = (lane == 0) ? x1024 : 0.0;
            const double m9 = (lane == 0) ? 0.0 : gi;
            acc[8] = fma(m8, m8, acc[8]);
            acc[9] = fma(m9, m9, acc[9]);
        }
        if constexpr (K2 + 1 < 32) CiUntangle<K2 + 1>::run(zr, zi, lane, partner, lt, acc);
    }
};

// load z[lane + 32*I]: I < 16 from the float4 at m = lane + 32 I as (x0*w, x2*w); I >= 16 from the float4
// at m' = (31 - lane) + 32 (31 - I) as (x3*w, x1*w)   (Makhoul reorder folded into the complex packing)
// ALIGNED is decided once per window, outside the unrolled sequence
template <int I, bool ALIGNED>
struct CiLoad {
    static __device__ __forceinline__ void run(double (&zr)[32], double (&zi)[32], const float* p,
                                               const double2* wz, const double2* wz2, int lane) {
        const int m = (I < 16) ? lane + 32 * I : (31 - lane) + 32 * (31 - I);
        float4 f;
        if constexpr (ALIGNED) {
            f = __ldg(reinterpret_cast<const float4*>(p) + m);
        } else {
            f.x = __ldg(p + 4 * m); f.y = __ldg(p + 4 * m + 1); f.z = __ldg(p + 4 * m + 2); f.w = __ldg(p + 4 * m + 3);
        }
        if constexpr (I < 16) {
            const double2 w = wz[m];
            zr[I] = (double)f.x * w.x;
            zi[I] = (double)f.z * w.y;
        } else {
            const double2 w = wz2[m];
            zr[I] = (double)f.w * w.x;
            zi[I] = (double)f.y * w.y;
        }
        if constexpr (I + 1 < 32) CiLoad<I + 1, ALIGNED>::run(zr, zi, p, wz, wz2, lane);
    }
};

__global__ void __launch_bounds__(kCiWarps * 32, 1) ci_bins_kernel(const CiArgs a) {
    extern __shared__ __align__(16) unsigned char ci_smem_raw[];
    double2* wz = reinterpret_cast<double2*>(ci_smem_raw);          // [512] (w[4m], w[4m+2])
    double2* wz2 = wz + 512;                                         // [512] (w[4m+3], w[4m+1])
    double2* ltab = wz2 + 512;                                       // [kCiLtabRows][32]
    double2* exall = ltab + kCiLtabRows * 32;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int m = tid; m < 512; m += kCiWarps * 32) {
        wz[m] = make_double2(a.window[4 * m], a.window[4 * m + 2]);
        wz2[m] = make_double2(a.window[4 * m + 3], a.window[4 * m + 1]);
    }
    for (int i = tid; i < kCiLtabRows * 32; i += kCiWarps * 32) ltab[i] = a.ltab[i];
    __syncthreads();
    double2* ex = exall + warp * kCiExElems;
    const double2* lt = ltab + lane;
    const int partner = (32 - lane) & 31;
    const long long stride = (long long)gridDim.x * kCiWarps;

#pragma unroll 1
    for (long long g = (long long)blockIdx.x * kCiWarps + warp; g < a.n_windows; g += stride) {
        // recording that owns global window g (binary search over the prefix table)
        int lo = 0, hi = a.n_items - 1;
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (a.win_base[mid] <= g) lo = mid; else hi = mid - 1;
        }
        const long long w = g - a.win_base[lo];
        // calculate()'s loop: first = hop = 1024 -> offsets 1024, 2048, ... (cacophony_index.py:89-91)
        const float* p = a.pcm + a.item_off[lo] + a.first + (long long)a.hop * w;
        const bool aligned = ((reinterpret_cast<uintptr_t>(p) & 15) == 0);

        double zr[32], zi[32];
        if (aligned) CiLoad<0, true>::run(zr, zi, p, wz, wz2, lane);
        else CiLoad<0, false>::run(zr, zi, p, wz, wz2, lane);
        fft_dif<32>(zr, zi);
        CiTw1<0>::run(zr, zi, lt);
        __syncwarp();
        CiExWrite<0>::run(zr, zi, ex + lane);
        __syncwarp();
        CiExRead<0>::run(zr, zi, ex + lane * kCiExStride);
        fft_dif<32>(zr, zi);

        double acc[10];
#pragma unroll
        for (int b = 0; b < 10; ++b) acc[b] = 0.0;
        CiUntangle<0>::run(zr, zi, lane, partner, lt, acc);
        double mine = 0.0;
#pragma unroll
        for (int b = 0; b < 10; ++b) {
            double v = acc[b];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == b) mine = v;
        }
        if (lane < 10) a.bins[g * 10 + lane] = mine;
    }
}

// points.append(sum(bins*2 < prev) + sum(bins > prev*2))  (src/cacophony_index.py:94-96); the first
// window of every recording only seeds `prev` and gets 255.
__global__ void __launch_bounds__(256) ci_points_kernel(const double* __restrict__ bins,
                                                        const long long* __restrict__ win_base, int n_items,
                                                        long long n_windows, unsigned char* __restrict__ points) {
    const long long g = (long long)blockIdx.x * 256 + threadIdx.x;
    if (g >= n_windows) return;
    int lo = 0, hi = n_items - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (win_base[mid] <= g) lo = mid; else hi = mid - 1;
    }
    if (g == win_base[lo]) { points[g] = 255; return; }
    int c = 0;
#pragma unroll
    for (int b = 0; b < 10; ++b) {
        const double cur = bins[g * 10 + b], prev = bins[(g - 1) * 10 + b];
        c += (cur * 2.0 < prev) + (cur > prev * 2.0);
    }
    points[g] = (unsigned char)c;
}

// score_from_points' sorted slice (src/cacophony_index.py:69-73) per 312-point entry: with points in
// 0..10 the sum of sorted[k0:k1] follows exactly from an 11-bin histogram.  One warp per entry.
__global__ void __launch_bounds__(256) ci_entries_kernel(const unsigned char* __restrict__ points,
                                                         const long long* __restrict__ entry_start,
                                                         const int* __restrict__ entry_len, long long n_entries,
                                                         int* __restrict__ out /* [n_entries][2] = sum, count */) {
    const long long e = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (e >= n_entries) return;
    const unsigned char* p = points + entry_start[e];
    const int n = entry_len[e];
    int hist[11];
#pragma unroll
    for (int v = 0; v < 11; ++v) hist[v] = 0;
    for (int i = lane; i < n; i += 32) {
        const int pv = p[i];
#pragma unroll
        for (int v = 0; v < 11; ++v) hist[v] += (pv == v);
    }
#pragma unroll
    for (int v = 0; v < 11; ++v)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) hist[v] += __shfl_xor_sync(0xffffffffu, hist[v], o);
    if (lane == 0) {
        const int k0 = (int)(n * 0.75), k1 = (int)(n * 0.95);      // Python int(len * 0.75), float64 product
        int pos = 0, sum = 0;
        for (int v = 0; v < 11; ++v) {
            const int a0 = max(pos, k0), a1 = min(pos + hist[v], k1);
            if (a1 > a0) sum += v * (a1 - a0);
            pos += hist[v];
        }
        out[2 * e] = sum;
        out[2 * e + 1] = k1 - k0;
    }
}

size_t ci_smem_bytes() { return sizeof(double2) * (512 + 512 + kCiLtabRows * 32 + (size_t)kCiWarps * kCiExElems); }

cudaError_t launch_ci_bins(const CiArgs& a, cudaStream_t st) {
    static bool attr_set[64] = {false};
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    const size_t smem = ci_smem_bytes();
    if (dev < 64 && !attr_set[dev]) {
        cudaError_t e = cudaFuncSetAttribute(ci_bins_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        attr_set[dev] = true;
    }
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    long long grid = (a.n_windows + kCiWarps - 1) / kCiWarps;
    if (grid > sms) grid = sms;                                 // persistent: one CTA per SM
    ci_bins_kernel<<<(unsigned)grid, kCiWarps * 32, smem, st>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_ci_points(const double* bins, const long long* win_base, int n_items, long long n_windows,
                             unsigned char* points, cudaStream_t st) {
    ci_points_kernel<<<(unsigned)((n_windows + 255) / 256), 256, 0, st>>>(bins, win_base, n_items, n_windows, points);
    return cudaGetLastError();
}

cudaError_t launch_ci_entries(const unsigned char* points, const long long* entry_start, const int* entry_len,
                              long long n_entries, int* out, cudaStream_t st) {
    ci_entries_kernel<<<(unsigned)((n_entries + 7) / 8), 256, 0, st>>>(points, entry_start, entry_len, n_entries, out);
    return cudaGetLastError();
}

}  // namespace cac

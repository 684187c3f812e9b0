// Framed 64*R-point real FFT, R = 75 (n_fft 4800 = sr // 10 at 48 kHz) or R = 25 (n_fft 1600 at 16 kHz),
// with selectable epilogue:
//   * identify_morepork._load_sample (src/identify_morepork.py:19-27): |stft(n_fft=4800, hop=2400)|[60:120]
//   * identify_tracks.get_end (src/identify_tracks.py:387-403): |stft(n_fft=4800, hop=281)| -> 120-mel (power 1)
// librosa.stft semantics: periodic Hann, centre = True with zero padding, frames = 1 + len // hop.
// One warp per frame: 4800 real points = 2400-point complex FFT = radix-75 (3 x 5 x 5, registers,
// lane = n2) -> twiddle -> transpose through shared memory -> radix-32 per row (lane = k1, three
// passes over the 75 rows, results written back in place) -> real-FFT untangle for the requested bins
// only, read straight from the Z matrix in shared memory.  No staging buffer: samples come from
// global memory (the overlap between neighbouring frames is served by L1/L2).
#include <algorithm>

#include "cac_common.cuh"
#include "fft_regs.cuh"
#include "spec.cuh"

namespace cac {

constexpr int kWarps48 = 8;
constexpr int kExStride48 = 33;                         // float2 per row of the Z matrix (odd: conflict-free)

// z[I] = (x[2n] w[2n], x[2n+1] w[2n+1]), n = lane + 32 I; w[s] = 0.5 - 0.5 cos(2 pi s / (64 R)), s = 64 I + (2 lane + e)
// MODE 1: the whole frame lies inside the recording and is 8-byte aligned (64-bit loads at immediate offsets from
// one base pointer); MODE 2: inside but not aligned (odd hop: every other frame) - two 32-bit loads, no checks;
// MODE 0: every sample is bounds-checked (the centre-padded frames at both ends of a recording).  The case is
// decided once per frame, outside the unrolled sequence (a per-element branch costs more than the load it guards).
template <int R, int I, int MODE>
struct Load48 {
    static __device__ __forceinline__ void run(float (&re)[R], float (&im)[R], const float* pl, long long tl,
                                               long long n, float hcb0, float hsb0, float hcb1, float hsb1) {
        // pl = p + 2 lane, tl = t0 + 2 lane: sample index of this lane's first element
        constexpr float ca = (float)cx_cos2pi(I, R);
        constexpr float sa = (float)cx_sin2pi(I, R);
        const float w0 = fmaf(-ca, hcb0, fmaf(sa, hsb0, 0.5f));
        const float w1 = fmaf(-ca, hcb1, fmaf(sa, hsb1, 0.5f));
        float x0, x1;
        if constexpr (MODE == 1) {
            const float2 v = __ldg(reinterpret_cast<const float2*>(pl + 64 * I));
            x0 = v.x; x1 = v.y;
        } else if constexpr (MODE == 2) {
            x0 = __ldg(pl + 64 * I);
            x1 = __ldg(pl + 64 * I + 1);
        } else {
            const long long a0 = tl + 64 * I;
            x0 = (a0 >= 0 && a0 < n) ? __ldg(pl + 64 * I) : 0.f;
            x1 = (a0 + 1 >= 0 && a0 + 1 < n) ? __ldg(pl + 64 * I + 1) : 0.f;
        }
        re[I] = x0 * w0;
        im[I] = x1 * w1;
        if constexpr (I + 1 < R) Load48<R, I + 1, MODE>::run(re, im, pl, tl, n, hcb0, hsb0, hcb1, hsb1);
    }
};

// pass-1 twiddle W_(32R)^(lane * K1) = A[K1 >> 3] * B[K1 & 7] on the value at out_pos(75, K1), then the
// value goes to row K1, column lane of the Z matrix
template <int R, int K1>
struct Tw48 {
    static __device__ __forceinline__ void run(const float (&re)[R], const float (&im)[R], const float2* lt,
                                               float2* ex_lane) {
        constexpr int A = K1 >> 3, B = K1 & 7, P = out_pos(R, K1);
        float xr = re[P], xi = im[P];
        if constexpr (K1 > 0) {
            float tr, ti;
            if constexpr (A == 0) { const float2 b = lt[(10 + B) * 32]; tr = b.x; ti = b.y; }
            else if constexpr (B == 0) { const float2 a = lt[A * 32]; tr = a.x; ti = a.y; }
            else {
                const float2 a = lt[A * 32], b = lt[(10 + B) * 32];
                tr = fmaf(-a.y, b.y, a.x * b.x);
                ti = fmaf(a.x, b.y, a.y * b.x);
            }
            const float yr = fmaf(-xi, ti, xr * tr);
            xi = fmaf(xr, ti, xi * tr);
            xr = yr;
        }
        ex_lane[K1 * kExStride48] = make_float2(xr, xi);
        if constexpr (K1 + 1 < R) Tw48<R, K1 + 1>::run(re, im, lt, ex_lane);
    }
};

template <int N2>
struct RowRead48 {
    static __device__ __forceinline__ void run(float (&zr)[32], float (&zi)[32], const float2* row) {
        const float2 t = row[N2];
        zr[N2] = t.x;
        zi[N2] = t.y;
        if constexpr (N2 + 1 < 32) RowRead48<N2 + 1>::run(zr, zi, row);
    }
};
template <int K2>
struct RowWrite48 {
    static __device__ __forceinline__ void run(const float (&zr)[32], const float (&zi)[32], float2* row) {
        row[K2] = make_float2(zr[out_pos(32, K2)], zi[out_pos(32, K2)]);
        if constexpr (K2 + 1 < 32) RowWrite48<K2 + 1>::run(zr, zi, row);
    }
};

struct Smem48 {
    int lt, rt, wts, mstart, ex, pbuf, total;   // byte offsets
};
__host__ __device__ inline Smem48 smem48_layout(int rows, int n_mels, int n_taps, int pbuf_len) {
    Smem48 L;
    int o = 0;
    L.lt = o; o += kSpec4800LtabRows * 32 * 8;
    L.rt = o; o += ((kSpec4800Rounds * 8 + 15) / 16) * 16;
    L.wts = o; o += ((n_mels ? n_taps : 0) * 4 + 15) / 16 * 16;
    L.mstart = o; o += (n_mels + 31) / 32 * 32 * 4;
    L.ex = o; o += kWarps48 * rows * kExStride48 * 8;
    L.pbuf = o; o += kWarps48 * (n_mels ? pbuf_len : 0) * 4;
    L.total = o;
    return L;
}

template <int R>
__global__ void __launch_bounds__(kWarps48 * 32, 1) spec4800_kernel(const SpecArgs a) {
    constexpr int kHalf = 32 * R;                      // complex FFT length = n_fft / 2
    constexpr int kExElems48 = R * kExStride48;
    extern __shared__ __align__(16) unsigned char smem48[];
    const Smem48 L = smem48_layout(R, a.n_mels, a.n_taps, a.pbuf_len);
    float2* ltab = reinterpret_cast<float2*>(smem48 + L.lt);
    float2* rtab = reinterpret_cast<float2*>(smem48 + L.rt);
    float* wts = reinterpret_cast<float*>(smem48 + L.wts);
    int* mstart = reinterpret_cast<int*>(smem48 + L.mstart);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float2* ex = reinterpret_cast<float2*>(smem48 + L.ex) + warp * kExElems48;
    float* pbuf = reinterpret_cast<float*>(smem48 + L.pbuf) + warp * (a.n_mels ? a.pbuf_len : 0);

    for (int i = tid; i < kSpec4800LtabRows * 32; i += kWarps48 * 32) ltab[i] = a.ltab[i];
    for (int i = tid; i < kSpec4800Rounds; i += kWarps48 * 32) rtab[i] = a.rtab[i];
    if (a.n_mels) {
        for (int i = tid; i < a.n_taps; i += kWarps48 * 32) wts[i] = a.wpad[i];
        for (int i = tid; i < a.n_groups * 32; i += kWarps48 * 32) mstart[i] = a.mstart[i];
    }
    __syncthreads();
    const float2* lt = ltab + lane;
    const long long stride = (long long)gridDim.x * kWarps48;
    float vmax = 0.f;

    // recording that owns the warp's first frame: binary search once, then walked forward (a warp's frames ascend).
    // Measured: the walk saves 9 % on morepork and 4 % on get_end (R = 75), but the R = 64 instantiation
    // (signal_noise) comes out 2.3x slower with it, so the other lengths keep the per-frame binary search.
    constexpr bool kWalk = (R == 75);
    int lo = 0;
    if (kWalk) {
        const long long g = (long long)blockIdx.x * kWarps48 + warp;
        int hi = a.n_items - 1;
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (a.frame_base[mid] <= g) lo = mid; else hi = mid - 1;
        }
    }
    const long long n_total = a.chunk_list ? (long long)(*a.chunk_list_n) * a.chunk_frames
                            : a.chunk_probe ? (long long)a.n_items * a.chunk_stride
                            : a.n_frames_total >= 0 ? a.n_frames_total : a.frame_base[a.n_items];
#pragma unroll 1
    for (long long g = (long long)blockIdx.x * kWarps48 + warp; g < n_total; g += stride) {
        int fsel = -1;                                                    // frame within the item when a chunk mode picks it
        if (a.chunk_list) {
            const int e = (int)(g / a.chunk_frames);
            const int2 ic = a.chunk_list[e];
            lo = ic.x;
            fsel = ic.y * a.chunk_frames + (int)(g - (long long)e * a.chunk_frames);
        } else if (a.chunk_probe) {
            lo = (int)(g / a.chunk_stride);
            fsel = (int)(g - (long long)lo * a.chunk_stride) * a.chunk_frames;
        } else if (kWalk) {
            while (a.frame_base[lo + 1] <= g) ++lo;           // g < n_frames_total = frame_base[n_items]: stops in range
        } else {
            lo = 0;
            int hi = a.n_items - 1;
            while (lo < hi) {
                const int mid = (lo + hi + 1) >> 1;
                if (a.frame_base[mid] <= g) lo = mid; else hi = mid - 1;
            }
        }
        const long long fbase = a.frame_base[lo];
        const int f = fsel >= 0 ? fsel : (int)(g - fbase);
        const int nfi = (int)(a.frame_base[lo + 1] - fbase);
        if (a.chunk_probe && (long long)f + a.chunk_frames >= nfi) continue;   // the reference's loop never looks at that chunk
        const long long n = a.item_len[lo];
        const long long t0 = (long long)f * a.hop - kHalf;                // centre = True
        const float* p = a.pcm + a.item_off[lo] + t0;
        const bool inside = t0 >= 0 && t0 + 2 * kHalf <= n;
        const bool aligned = (reinterpret_cast<uintptr_t>(p) & 7) == 0;
        {
            float re[R], im[R];
            const float2 h0 = lt[18 * 32], h1 = lt[19 * 32];
            if (inside && aligned) Load48<R, 0, 1>::run(re, im, p + 2 * lane, t0 + 2 * lane, n, h0.x, h0.y, h1.x, h1.y);
            else if (inside) Load48<R, 0, 2>::run(re, im, p + 2 * lane, t0 + 2 * lane, n, h0.x, h0.y, h1.x, h1.y);
            else Load48<R, 0, 0>::run(re, im, p + 2 * lane, t0 + 2 * lane, n, h0.x, h0.y, h1.x, h1.y);
            fft_dif<R>(re, im);
            __syncwarp();                                                 // previous frame's readers are done
            Tw48<R, 0>::run(re, im, lt, ex + lane);
        }
        __syncwarp();
#pragma unroll 1
        for (int h = 0; h * 32 < R; ++h) {
            const int k1 = lane + 32 * h;
            if (k1 < R) {
                float zr[32], zi[32];
                float2* row = ex + k1 * kExStride48;
                RowRead48<0>::run(zr, zi, row);
                fft_dif<32>(zr, zi);
                RowWrite48<0>::run(zr, zi, row);                          // Z[k1 + R k2] at row k1, column k2
            }
        }
        __syncwarp();
        // real-FFT untangle of the requested bins, H = n_fft / 2:
        //   X[k] = (Z[k] + conj Z[H-k])/2 - (i/2) W_(2H)^k (Z[k] - conj Z[H-k])
        const float2 lc = lt[20 * 32];                                    // cos / sin of 2 pi (kmin + lane) / n_fft
        // bin k = kmin + lane + 32 r sits at row k % R, column k / R of the Z matrix, its partner H - k likewise: both
        // (quotient, remainder) pairs are carried from round to round instead of being divided out again, and the
        // output position walks by a fixed step (bin-major: 32 rows of the recording's [bins][frames] matrix)
        float* op;
        size_t ostep = 32;
        if (a.n_mels) op = pbuf + lane;
        else if (a.bin_major) { op = a.out + (size_t)fbase * a.nbins + f + (size_t)lane * nfi; ostep = (size_t)32 * nfi; }
        else op = a.out + (size_t)g * a.nbins + lane;
        int k = a.kmin + lane;
        int qa = k / R, ra = k - qa * R;
        int qb = (kHalf - k) / R, rb = (kHalf - k) - qb * R;
#pragma unroll 1
        for (int r = 0; r * 32 < a.nbins; ++r, k += 32, op += ostep) {
            if (r * 32 + lane < a.nbins) {
                const int ia = (k == kHalf) ? 0 : ra * kExStride48 + qa;
                const int ib = (k == 0 || k == kHalf) ? 0 : rb * kExStride48 + qb;
                const float2 za = ex[ia];
                const float2 zb = ex[ib];
                const float2 rc = rtab[r];
                const float c = fmaf(-rc.y, lc.y, rc.x * lc.x);
                const float s = fmaf(rc.x, lc.y, rc.y * lc.x);
                const float br = zb.x, bi = -zb.y;
                const float sumr = za.x + br, sumi = za.y + bi;
                const float dr = za.x - br, di = za.y - bi;
                const float tr = fmaf(di, s, dr * c);
                const float ti = fmaf(-dr, s, di * c);
                const float xr = 0.5f * (sumr + ti);
                const float xi = 0.5f * (sumi - tr);
                float pw = fmaf(xr, xr, xi * xi);
                if (a.power == 1) pw = sqrtf(pw);
                vmax = fmaxf(vmax, pw);
                *op = pw;
            }
            ra += 32 % R; qa += 32 / R;
            if (ra >= R) { ra -= R; ++qa; }
            rb -= 32 % R; qb -= 32 / R;
            if (rb < 0) { rb += R; --qb; }
        }
        if (a.n_mels) {
            for (int i = a.nbins + lane; i < a.pbuf_len; i += 32) pbuf[i] = 0.f;
            __syncwarp();
            float* oitem = a.out + (size_t)fbase * a.n_mels;
            float fmn = INFINITY, fmx = 0.f;
#pragma unroll 1
            for (int gidx = 0; gidx < a.n_groups; ++gidx) {
                const int slot = mstart[gidx * 32 + lane];   // (band start) | row << 16, mel_bands.h
                const int m = (int)((unsigned)slot >> 16);
                const bool live = m != 0xFFFF;
                const float4* pp = reinterpret_cast<const float4*>(pbuf + (slot & 0xFFFF));
                const float4* wg = reinterpret_cast<const float4*>(wts + a.goff[gidx]) + lane;
                const int gm4 = a.gmax_taps[gidx] >> 2;
                float acc = 0.f;
                for (int i = 0; i < gm4; ++i) {
                    const float4 w = wg[i * 32];
                    const float4 q = pp[i];
                    acc = fmaf(w.x, q.x, acc);
                    acc = fmaf(w.y, q.y, acc);
                    acc = fmaf(w.z, q.z, acc);
                    acc = fmaf(w.w, q.w, acc);
                }
                if (a.chunk_mm) { if (live) { fmn = fminf(fmn, acc); fmx = fmaxf(fmx, acc); } }
                else if (live) oitem[(size_t)m * nfi + f] = acc;
            }
            if (a.chunk_mm) {                                 // get_end: only the chunk's extrema are kept
                fmn = warp_min(fmn);
                fmx = warp_max(fmx);
                if (lane == 0) {
                    uint2* mm = a.chunk_mm + (size_t)lo * a.chunk_stride + f / a.chunk_frames;
                    atomicMin(&mm->x, __float_as_uint(fmn));
                    atomicMax(&mm->y, __float_as_uint(fmx));
                }
            }
            __syncwarp();
        }
    }
    if (a.gmax) {
        vmax = warp_max(vmax);
        if (lane == 0) atomicMax(reinterpret_cast<unsigned*>(a.gmax), __float_as_uint(vmax));
    }
}

// min / max of every block in[:, c*chunk : (c+1)*chunk] of a row-major [n_rows][n_cols] matrix
// (get_end's 1-second chunks, src/identify_tracks.py:404-412)
__global__ void __launch_bounds__(256) chunk_minmax_kernel(const float* __restrict__ in, int n_rows, long long n_cols,
                                                           int chunk, float2* __restrict__ out) {
    __shared__ float smn[8], smx[8];
    const long long c = blockIdx.x;
    float mn = INFINITY, mx = -INFINITY;
    for (int i = threadIdx.x; i < n_rows * chunk; i += 256) {
        const int rr = i / chunk, cc = i - rr * chunk;
        const float v = in[(size_t)rr * n_cols + c * chunk + cc];
        mn = fminf(mn, v);
        mx = fmaxf(mx, v);
    }
    mn = warp_min(mn); mx = warp_max(mx);
    if ((threadIdx.x & 31) == 0) { smn[threadIdx.x >> 5] = mn; smx[threadIdx.x >> 5] = mx; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) { mn = fminf(mn, smn[w]); mx = fmaxf(mx, smx[w]); }
        out[c] = make_float2(mn, mx);
    }
}

// identify_morepork's window loop body (src/identify_morepork.py:150-166) for one window per CTA:
// spec is frame-major [frame][n_bins]; the window covers `width` frames from win_frame0;
// amplitude_to_db(ref=np.max) (amin 1e-5, top_db 80), rescale by |min|, zero-pad to 60 columns,
// output [n_bins][60].
// 10 log10(max(amin, m^2)) - refdb, floored at -80 (librosa.power_to_db with ref = max, top_db = 80, on float32 as the
// reference calls it).  10 log10(x) = 3.0103 log2(x) on the hardware log2 (absolute error below 3e-5 dB over the whole
// range, 3e-7 after the division by 80; the library log10f costs ~20 instructions per value in kernels that are bound by
// instruction issue).  refdb is computed by the same expression, so the reference cell is exactly 0.
__device__ __forceinline__ float morepork_db_raw(float m) {
    const float amin = 1e-10f;                 // (1e-5)**2 rounded to float32
    return __fmul_rn(3.0102999566398120f, __log2f(fmaxf(amin, __fmul_rn(m, m))));
}
__device__ __forceinline__ float morepork_db(float m, float refdb) {
    return fmaxf(__fsub_rn(morepork_db_raw(m), refdb), -80.0f);   // log_spec.max() is exactly 0 at the reference cell
}
// v / den + 1 with one reciprocal per thread: q = v r corrected by one residual step (correctly rounded but for rare ties)
__device__ __forceinline__ float morepork_unit(float v, float den, float rden) {
    const float q = __fmul_rn(v, rden);
    return __fadd_rn(fmaf(fmaf(-q, den, v), rden, q), 1.0f);
}

__global__ void __launch_bounds__(256) morepork_windows_kernel(const float* __restrict__ spec,
                                                               const long long* __restrict__ win_frame0,
                                                               const int* __restrict__ win_width, int n_bins,
                                                               float* __restrict__ out) {
    extern __shared__ float tile[];           // [width][n_bins]
    __shared__ float sred[8];
    __shared__ float s_bcast;
    const long long w = blockIdx.x;
    const int width = win_width[w];
    const float* src = spec + (size_t)win_frame0[w] * n_bins;
    const int n = width * n_bins;
    const int tid = threadIdx.x;
    float mx = 0.f;
    for (int i = tid; i < n; i += 256) {
        const float v = fabsf(src[i]);
        tile[i] = v;
        mx = fmaxf(mx, v);
    }
    mx = warp_max(mx);
    if ((tid & 31) == 0) sred[tid >> 5] = mx;
    __syncthreads();
    if (tid == 0) {
        for (int k = 1; k < 8; ++k) mx = fmaxf(mx, sred[k]);
        s_bcast = mx;
    }
    __syncthreads();
    const float ref = s_bcast;
    const float refdb = morepork_db_raw(ref);
    float mn = INFINITY;
    for (int i = tid; i < n; i += 256) {
        const float v = morepork_db(tile[i], refdb);
        tile[i] = v;
        mn = fminf(mn, v);
    }
    mn = warp_min(mn);
    __syncthreads();
    if ((tid & 31) == 0) sred[tid >> 5] = mn;
    __syncthreads();
    if (tid == 0) {
        for (int k = 1; k < 8; ++k) mn = fminf(mn, sred[k]);
        s_bcast = mn;
    }
    __syncthreads();
    const float vmin = s_bcast;
    const float den = fabsf(vmin), rden = __frcp_rn(den);
    float* o = out + (size_t)w * n_bins * 60;
    for (int i = tid; i < n_bins * 60; i += 256) {
        const int b = i / 60, c = i - b * 60;
        float v = 0.f;                         // np.pad columns
        if (c < width) {
            v = tile[c * n_bins + b];
            if (vmin != 0.f) v = morepork_unit(v, den, rden);
        }
        o[i] = v;
    }
}

__global__ void __launch_bounds__(256) scale_by_max_kernel(float* __restrict__ data, long long n,
                                                           const float* __restrict__ gmax) {
    const float d = *gmax;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256)
        data[i] = __fdiv_rn(data[i], d);
}

// ---- get_end for a batch: device-built item tables and the first-silent-chunk scan ------------------------------
__global__ void get_end_prep_kernel(const long long* __restrict__ off, const long long* __restrict__ len, int n, int n_fft,
                                    int hop, long long max_len, long long* __restrict__ tab, uint2* __restrict__ mm,
                                    int chunk_stride, int* __restrict__ out) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        long long total = 0;
        for (int i = 0; i < n; ++i) {
            const long long ns = max(0LL, min(len[i], max_len));
            tab[i] = off[i];
            tab[n + i] = ns;
            tab[2 * n + i] = total;
            const long long padded = ns + 2LL * (n_fft / 2);
            total += padded >= n_fft ? 1 + (padded - n_fft) / hop : 0;
        }
        tab[3 * n] = total;
    }
    const long long cells = (long long)n * chunk_stride;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < cells; i += (long long)gridDim.x * blockDim.x)
        mm[i] = make_uint2(0x7f800000u, 0u);                  // {+inf, 0}
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) out[i] = -1;
}

__global__ void get_end_select_kernel(const long long* __restrict__ tab, int n, const uint2* __restrict__ mm, int chunk_stride,
                                      int chunk_frames, int2* __restrict__ list, int* __restrict__ list_n) {
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)n * chunk_stride) return;
    const int i = (int)(idx / chunk_stride), c = (int)(idx - (long long)i * chunk_stride);
    const long long nf = tab[2 * n + i + 1] - tab[2 * n + i];
    if ((long long)(c + 1) * chunk_frames >= nf) return;
    const uint2 v = mm[idx];
    if (v.x == v.y) list[atomicAdd(list_n, 1)] = make_int2(i, c);
}

// the reference's loop (src/identify_tracks.py:406-412): chunks [c L, (c + 1) L) while (c + 1) L < n_frames
__global__ void get_end_scan_kernel(const long long* __restrict__ tab, int n, const uint2* __restrict__ mm, int chunk_stride,
                                    int chunk_frames, int* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const long long nf = tab[2 * n + i + 1] - tab[2 * n + i];
    for (int c = 0; (long long)(c + 1) * chunk_frames < nf; ++c) {
        const uint2 v = mm[(size_t)i * chunk_stride + c];
        if (v.x == v.y) { out[i] = c; return; }
    }
}

// ---- host side -------------------------------------------------------------------------------------
void spec4800_host_tables(int n_fft, int kmin, float2* ltab, float2* rtab) {
    const long double tau = 6.283185307179586476925286766559005768L;
    const int half = n_fft / 2;
    for (int lane = 0; lane < 32; ++lane) {
        for (int a = 0; a < 10; ++a) {
            const long double ang = tau * (long double)((8 * a * lane) % half) / (long double)half;
            ltab[a * 32 + lane] = make_float2((float)cosl(ang), (float)-sinl(ang));
        }
        for (int b = 0; b < 8; ++b) {
            const long double ang = tau * (long double)(b * lane) / (long double)half;
            ltab[(10 + b) * 32 + lane] = make_float2((float)cosl(ang), (float)-sinl(ang));
        }
        for (int e = 0; e < 2; ++e) {
            const long double ang = tau * (long double)(2 * lane + e) / (long double)n_fft;
            ltab[(18 + e) * 32 + lane] = make_float2((float)(0.5L * cosl(ang)), (float)(0.5L * sinl(ang)));
        }
        const long double ang = tau * (long double)((kmin + lane) % n_fft) / (long double)n_fft;
        ltab[20 * 32 + lane] = make_float2((float)cosl(ang), (float)sinl(ang));
    }
    for (int r = 0; r < kSpec4800Rounds; ++r) {
        const long double ang = tau * (long double)((32 * r) % n_fft) / (long double)n_fft;
        rtab[r] = make_float2((float)cosl(ang), (float)sinl(ang));
    }
}

bool spec4800_supported(int n_fft) {
    return n_fft == 4800 || n_fft == 1600 || n_fft == 3200 || n_fft == 4096 || n_fft == 2048 || n_fft == 1024 ||
           n_fft == 512;
}

size_t spec4800_smem_bytes(const SpecArgs& a) {
    return (size_t)smem48_layout(a.n_fft / 64, a.n_mels, a.n_taps, a.pbuf_len).total;
}

cudaError_t launch_spec4800(const SpecArgs& a, cudaStream_t st) {
    const size_t smem = spec4800_smem_bytes(a);
    void (*kern)(const SpecArgs) = nullptr;
    switch (a.n_fft) {
        case 4800: kern = spec4800_kernel<75>; break;   // sr // 10 at 48 kHz (get_end, morepork)
        case 1600: kern = spec4800_kernel<25>; break;   // sr // 10 at 16 kHz
        case 3200: kern = spec4800_kernel<50>; break;   // sr // 10 at 32 kHz
        case 4096: kern = spec4800_kernel<64>; break;   // signal_noise's full spectrum (src/identify_tracks.py:652-654)
        case 2048: kern = spec4800_kernel<32>; break;   // load_samples with a model whose metadata sets n_fft
        case 1024: kern = spec4800_kernel<16>; break;
        case 512:  kern = spec4800_kernel<8>; break;
        default: return cudaErrorInvalidValue;
    }
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    long long grid = a.n_frames_total >= 0 ? (a.n_frames_total + kWarps48 - 1) / kWarps48 : sms;
    if (grid > sms) grid = sms;
    kern<<<(unsigned)grid, kWarps48 * 32, smem, st>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_get_end_prep(const long long* d_item_off, const long long* d_item_len, int n_items, int n_fft, int hop,
                                long long max_item_samples, long long* tab, uint2* mm, int chunk_stride, int* out,
                                cudaStream_t st) {
    const long long cells = (long long)n_items * chunk_stride;
    const unsigned blocks = (unsigned)std::min<long long>(std::max<long long>((cells + 255) / 256, 1), 1024);
    get_end_prep_kernel<<<blocks, 256, 0, st>>>(d_item_off, d_item_len, n_items, n_fft, hop, max_item_samples, tab, mm,
                                                chunk_stride, out);
    return cudaGetLastError();
}

cudaError_t launch_get_end_select(const long long* tab, int n_items, const uint2* mm, int chunk_stride, int chunk_frames,
                                  int2* list, int* list_n, cudaStream_t st) {
    const long long cells = (long long)n_items * chunk_stride;
    get_end_select_kernel<<<(unsigned)((cells + 255) / 256), 256, 0, st>>>(tab, n_items, mm, chunk_stride, chunk_frames, list,
                                                                           list_n);
    return cudaGetLastError();
}

cudaError_t launch_get_end_scan(const long long* tab, int n_items, const uint2* mm, int chunk_stride, int chunk_frames,
                                int* out, cudaStream_t st) {
    get_end_scan_kernel<<<(unsigned)((n_items + 127) / 128), 128, 0, st>>>(tab, n_items, mm, chunk_stride, chunk_frames, out);
    return cudaGetLastError();
}

cudaError_t launch_chunk_minmax(const float* in, int n_rows, long long n_cols, int chunk, long long n_chunks,
                                float2* out, cudaStream_t st) {
    chunk_minmax_kernel<<<(unsigned)n_chunks, 256, 0, st>>>(in, n_rows, n_cols, chunk, out);
    return cudaGetLastError();
}

// The same for the native layout (60 bins per frame, 16-byte aligned spectrogram) in well under half the instructions
// (the kernel above runs ~100 per output value at 87 % of the issue slots): 16-byte loads, the smallest dB value taken
// from the smallest magnitude (every step of the dB formula is monotone, so min over the values = the value of the
// min; no second pass over the tile, no second reduction), rows of the tile padded to 61 floats so that the transposed
// read is conflict-free, four outputs per thread written as one 16-byte store.
__global__ void __launch_bounds__(256) morepork_windows60_kernel(const float* __restrict__ spec,
                                                                 const long long* __restrict__ win_frame0,
                                                                 const int* __restrict__ win_width, float* __restrict__ out) {
    constexpr int NB = 60, PITCH = 61, Q = NB / 4;          // 15 quads per frame
    __shared__ float tile[60 * PITCH];                      // [width][61]: magnitudes (signs kept, only squares are used)
    __shared__ float smx[8], smn[8];
    const long long w = blockIdx.x;
    const int width = win_width[w];
    const float4* src = reinterpret_cast<const float4*>(spec + (size_t)win_frame0[w] * NB);
    const int tid = threadIdx.x;
    float mx = 0.f, mn = INFINITY;
    for (int i = tid; i < width * Q; i += 256) {
        const float4 v = __ldg(src + i);
        const int c = i / Q, b0 = 4 * (i - c * Q);
        float* t = tile + c * PITCH + b0;
        t[0] = v.x; t[1] = v.y; t[2] = v.z; t[3] = v.w;
        mx = fmaxf(fmaxf(mx, fabsf(v.x)), fmaxf(fabsf(v.y), fmaxf(fabsf(v.z), fabsf(v.w))));
        mn = fminf(fminf(mn, fabsf(v.x)), fminf(fabsf(v.y), fminf(fabsf(v.z), fabsf(v.w))));
    }
    mx = warp_max(mx);
    mn = warp_min(mn);
    if ((tid & 31) == 0) { smx[tid >> 5] = mx; smn[tid >> 5] = mn; }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 8; ++k) { mx = fmaxf(mx, smx[k]); mn = fminf(mn, smn[k]); }
    const float refdb = morepork_db_raw(mx);
    const float vmin = morepork_db(mn, refdb);
    const float den = fabsf(vmin), rden = __frcp_rn(den);
    float4* o = reinterpret_cast<float4*>(out + (size_t)w * NB * 60);
    for (int i = tid; i < NB * Q; i += 256) {               // output quad i = bin i / 15, columns 4 (i % 15) ..
        const int b = i / Q, c0 = 4 * (i - b * Q);
        float r[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float v = 0.f;                                  // np.pad columns
            if (c0 + j < width) {
                v = morepork_db(tile[(c0 + j) * PITCH + b], refdb);
                if (vmin != 0.f) v = morepork_unit(v, den, rden);
            }
            r[j] = v;
        }
        o[i] = make_float4(r[0], r[1], r[2], r[3]);
    }
}

cudaError_t launch_morepork_windows(const float* spec, const long long* win_frame0, const int* win_width,
                                    long long n_windows, int n_bins, float* out, cudaStream_t st) {
    if (n_bins == 60 && ((uintptr_t)spec & 15) == 0 && ((uintptr_t)out & 15) == 0)
        morepork_windows60_kernel<<<(unsigned)n_windows, 256, 0, st>>>(spec, win_frame0, win_width, out);
    else
        morepork_windows_kernel<<<(unsigned)n_windows, 256, 60 * n_bins * sizeof(float), st>>>(spec, win_frame0, win_width,
                                                                                              n_bins, out);
    return cudaGetLastError();
}

cudaError_t launch_scale_by_max(float* data, long long n, const float* gmax, cudaStream_t st) {
    long long blocks = (n + 256 * 8 - 1) / (256 * 8);
    if (blocks > 148LL * 32) blocks = 148LL * 32;
    if (blocks < 1) blocks = 1;
    scale_by_max_kernel<<<(unsigned)blocks, 256, 0, st>>>(data, n, gmax);
    return cudaGetLastError();
}

}  // namespace cac

// Exact order statistics by radix select on float bit patterns (values >= 0: the bit pattern orders like the value)
// - np.median of signal_noise (src/identify_tracks.py:658-659).
// The keys of a spectrogram row / column share their leading bits (sign, most of the exponent), so the select
// starts below the common prefix of the smallest and the largest key and takes wide digits (11 bits per pass for a
// warp's private histogram, 12 for a CTA's): two or three passes instead of four 8-bit ones, and the first pass
// spreads over hundreds of bins instead of the two or three exponent bytes (no same-address serialisation).
// Histogram updates are unconditional atomics whose ADDRESS is selected (elements outside the current prefix go to a
// throw-away word): no branch per element.
#pragma once
#include <cuda_runtime.h>

namespace cac {

constexpr int kSelWarpBits = 11, kSelWarpBins = 1 << kSelWarpBits;   // 8 KB per warp
constexpr int kSelCtaBits = 12, kSelCtaBins = 1 << kSelCtaBits;      // 16 KB per CTA

// hist[bin] += 1 when p, else a throw-away word: ONE unconditional shared-memory atomic.  (`if (p) atomicAdd` and a
// predicated `red.shared` both compile to a branch with a reconvergence barrier around ATOMS - ncu showed
// branch_resolving at 20-25 % of the select's stall samples; a select of the address costs nothing of the kind.)
__device__ __forceinline__ void smem_inc_sel(unsigned* hist, unsigned bin, bool p, unsigned* trash) {
    atomicAdd(p ? hist + bin : trash, 1u);
}

// One warp finds the bin of its 2048-bin histogram that holds rank k (0-based, k < total): returns the bin, k becomes
// the rank inside the bin.  Lane owns the quads (4 consecutive bins) lane + 32 j.  All 32 lanes must call it.
__device__ __forceinline__ unsigned sel_scan_warp(const unsigned* hist, int lane, unsigned& k) {
    const uint4* h4 = reinterpret_cast<const uint4*>(hist);
    constexpr int J = kSelWarpBins / 4 / 32;            // 16 quads per lane
    unsigned c[J];
    int js = 0;
    unsigned kk = k;
    bool found = false;
#pragma unroll
    for (int j = 0; j < J; ++j) {
        const uint4 q = h4[lane + 32 * j];
        c[j] = q.x + q.y + q.z + q.w;
        const unsigned t = __reduce_add_sync(0xffffffffu, c[j]);     // bins [128 j, 128 j + 128)
        if (!found) {
            if (kk >= t && j + 1 < J) { kk -= t; js = j + 1; } else found = true;
        }
    }
    unsigned cj = c[0];
#pragma unroll
    for (int j = 1; j < J; ++j)
        if (js == j) cj = c[j];
    unsigned inc = cj;                                   // inclusive prefix over the lanes of block js
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned up = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += up;
    }
    const unsigned ball = __ballot_sync(0xffffffffu, inc > kk);
    const int ls = ball ? __ffs(ball) - 1 : 31;
    kk -= __shfl_sync(0xffffffffu, inc - cj, ls);
    const uint4 q = h4[ls + 32 * js];                    // the quad that holds the rank (same address for all lanes)
    unsigned b = 0;
    if (kk >= q.x) { kk -= q.x; b = 1; if (kk >= q.y) { kk -= q.y; b = 2; if (kk >= q.z) { kk -= q.z; b = 3; } } }
    k = kk;
    return 4u * (unsigned)(ls + 32 * js) + b;
}

// The whole CTA (NT threads, all must call) finds the bin of its 4096-bin histogram that holds rank k: the bin goes to
// s_out[0], the rank inside it to s_out[1].  s_tot: NT / 32 words of scratch.  Ends with a barrier.
template <int NT>
__device__ __forceinline__ void sel_scan_cta(const unsigned* hist, unsigned k, unsigned* s_tot, unsigned* s_out) {
    constexpr int BPT = kSelCtaBins / NT;                // consecutive bins per thread
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint4* h4 = reinterpret_cast<const uint4*>(hist) + tid * (BPT / 4);
    unsigned bins[BPT];
    unsigned s = 0;
#pragma unroll
    for (int i = 0; i < BPT / 4; ++i) {
        const uint4 q = h4[i];
        bins[4 * i] = q.x; bins[4 * i + 1] = q.y; bins[4 * i + 2] = q.z; bins[4 * i + 3] = q.w;
        s += q.x + q.y + q.z + q.w;
    }
    unsigned inc = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned up = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += up;
    }
    if (lane == 31) s_tot[warp] = inc;
    __syncthreads();
    unsigned base = 0;
#pragma unroll
    for (int w = 0; w < NT / 32; ++w)
        if (w < warp) base += s_tot[w];
    const unsigned exc = base + inc - s;
    if (k >= exc && k < exc + s) {                       // exactly one thread
        unsigned kk = k - exc;
        unsigned b = 0;
#pragma unroll
        for (int i = 0; i < BPT - 1; ++i)
            if (b == (unsigned)i && kk >= bins[i]) { kk -= bins[i]; b = i + 1; }
        s_out[0] = (unsigned)(tid * BPT) + b;
        s_out[1] = kk;
    }
    __syncthreads();
}

}  // namespace cac

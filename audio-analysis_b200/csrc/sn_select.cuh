// Exact order statistics by 8-bit radix select on float bit patterns (values >= 0: the bit pattern orders like the
// value) - np.median of signal_noise (src/identify_tracks.py:658-659).  The histogram of one pass is 256 digits x 4
// replicas in shared memory, word index = digit * 4 + (lane & 3): the lanes of a warp that meet in one digit spread
// over four counters in four different banks, so one ATOMS instruction serialises ~2 deep instead of ~8 deep when
// most of a row / column shares its exponent byte.
#pragma once
#include <cuda_runtime.h>

namespace cac {

constexpr int kSelHistWords = 256 * 4;

// One warp reads the 1024-word histogram and finds the digit that holds rank k (0-based, k < total count):
// returns the digit, k becomes the rank inside that digit's bucket.  All 32 lanes must call it.
__device__ __forceinline__ unsigned sel_scan_digit(const unsigned* hist, int lane, unsigned& k) {
    const uint4* h4 = reinterpret_cast<const uint4*>(hist);
    unsigned c[8], t[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {                       // lane owns digits lane + 32 j: conflict-free 128-bit reads
        const uint4 q = h4[lane + 32 * j];
        c[j] = q.x + q.y + q.z + q.w;
        t[j] = __reduce_add_sync(0xffffffffu, c[j]);    // count of digits [32 j, 32 j + 32)
    }
    int js = 0;
    unsigned kk = k;
#pragma unroll
    for (int j = 0; j < 7; ++j)
        if (js == j && kk >= t[j]) { kk -= t[j]; js = j + 1; }
    unsigned cj = c[0];
#pragma unroll
    for (int j = 1; j < 8; ++j)
        if (js == j) cj = c[j];
    unsigned inc = cj;                                  // inclusive prefix over the lanes of block js
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned up = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += up;
    }
    const unsigned ball = __ballot_sync(0xffffffffu, inc > kk);
    const int ls = ball ? __ffs(ball) - 1 : 31;         // first lane whose cumulative count exceeds kk
    const unsigned exc = __shfl_sync(0xffffffffu, inc - cj, ls);
    k = kk - exc;
    return (unsigned)(32 * js + ls);
}

}  // namespace cac

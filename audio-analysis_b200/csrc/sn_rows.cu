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

// ---- the same row step for matrices with 16-byte aligned rows (every matrix the fused chain produces) ------------------
// ncu's source view of the kernel above (profiles/r3j_sn_rowmed_lines.txt): 2,100 instructions per thread and row, the
// issue slots 64 % busy - the kernel is bound by its INSTRUCTION COUNT, not by the histogram atomics or by HBM.  This
// version does the same select and mask in well under half the instructions:
//   * thread t holds the columns 4 t + 4 NT v + {0..3}: one 16-byte load per four keys, one threshold load per four
//     columns, the mask word of 32 columns assembled from eight lanes' nibbles by three shuffles;
//   * a BRACKET [L, L + span] for the median from 32 keys spread over the row (sorted by every warp with shuffles; the
//     bracket is the sample's ranks 6 and 25, which misses the median of one row in about a thousand);
//   * ONE pass over the keys: keys below L are counted, keys inside the bracket go to bin (key - L) >> shift of a
//     4,096-bin histogram that spans exactly the bracket - so the bin that holds the median has a few keys, where the
//     top 12 varying bits of the whole row (the kernel above) leave hundreds of keys for two more passes;
//   * those few keys are compacted into shared memory and the order statistic is found among them by counting;
//   * the second middle value of an even row comes from the same candidates, or - when the first one is the largest of
//     its bin - from one min-pass over the keys above the bin.
// The bracket only steers: the counts are exact, and when the rank turns out to lie outside it (or the row is too
// short to sample) the pass is redone with the row's minimum and maximum as the bracket; a bin with more keys than
// threads (heavy ties) becomes the next, narrower bracket.
constexpr unsigned kNoKey = 0x80000000u;      // -0.0f: above every magnitude as unsigned, below as signed, never > t

struct RowSelV {
    unsigned hist[kSelCtaBins + 32];          // + one overflow word per lane: keys outside the bracket
    unsigned tot[32];
    unsigned out[4];                          // bin, rank inside the bin, keys in the bin (0: rank not in the histogram)
    unsigned mn;
    int mx;
    unsigned below, ncand, v1, v2, min_gt;
    float tr;
    unsigned sample[32];
    unsigned cand[1024];
};

// bin of `hist` that holds rank k, the rank inside it and the bin's count -> out[0..2] (untouched when k is not below
// the histogram's total).  All NT threads call; ends with a barrier.
template <int NT>
__device__ __forceinline__ void sel_scan_cta_v(const unsigned* hist, unsigned k, unsigned* s_tot, unsigned* out) {
    constexpr int BPT = kSelCtaBins / NT, NW = NT / 32;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint4* h4 = reinterpret_cast<const uint4*>(hist) + tid * (BPT / 4);
    unsigned bins[BPT];
    unsigned sum = 0;
#pragma unroll
    for (int i = 0; i < BPT / 4; ++i) {
        const uint4 q = h4[i];
        bins[4 * i] = q.x; bins[4 * i + 1] = q.y; bins[4 * i + 2] = q.z; bins[4 * i + 3] = q.w;
        sum += q.x + q.y + q.z + q.w;
    }
    unsigned inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned up = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += up;
    }
    if (lane == 31) s_tot[warp] = inc;
    __syncthreads();
    const unsigned wt = lane < NW ? s_tot[lane] : 0u;       // exclusive scan of the warp totals, redone by every warp
    unsigned winc = wt;
#pragma unroll
    for (int o = 1; o < NW; o <<= 1) {
        const unsigned up = __shfl_up_sync(0xffffffffu, winc, o);
        if (lane >= o) winc += up;
    }
    const unsigned base = __shfl_sync(0xffffffffu, winc - wt, warp);
    const unsigned exc = base + inc - sum;
    if (k >= exc && k < exc + sum) {                        // at most one thread
        unsigned kk = k - exc, b = 0, cnt = bins[0];
#pragma unroll
        for (int i = 0; i < BPT - 1; ++i)
            if (b == (unsigned)i && kk >= bins[i]) { kk -= bins[i]; b = i + 1; cnt = bins[i + 1]; }
        out[0] = (unsigned)(tid * BPT) + b;
        out[1] = kk;
        out[2] = cnt;
    }
    __syncthreads();
}

// ascending bitonic sort of one value per lane
__device__ __forceinline__ unsigned warp_sort32(unsigned x, int lane) {
#pragma unroll
    for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            const unsigned y = __shfl_xor_sync(0xffffffffu, x, j);
            const bool keep_min = ((lane & j) == 0) == ((lane & k) == 0);
            x = keep_min ? min(x, y) : max(x, y);
        }
    }
    return x;
}

template <int NT, int V, int MINB>
__global__ void __launch_bounds__(NT, MINB) sn_rowmed_mask_vec_kernel(const SnWave w) {
    static_assert(NT <= 1024 && NT % 32 == 0 && kSelCtaBins % (4 * NT) == 0, "threads per CTA");
    __shared__ __align__(16) RowSelV s;
    const int item = blockIdx.y, row = blockIdx.x;
    const SnItem im = w.items[item];
    const int cols = im.cols;
    if (cols < 1 || cols > 4 * NT * V) return;              // served by a wider instantiation (launch_sn_rowmed_mask)
    const int tid = threadIdx.x, lane = tid & 31;
    const unsigned* rowp = reinterpret_cast<const unsigned*>(w.spec + (size_t)item * w.spec_stride) + (size_t)row * im.pitch;
    const int c0 = 4 * tid;
    unsigned key[4 * V];
#pragma unroll
    for (int v = 0; v < V; ++v) {
        const int c = c0 + 4 * NT * v;
        uint4 q = make_uint4(kNoKey, kNoKey, kNoKey, kNoKey);
        if (c < cols) q = __ldg(reinterpret_cast<const uint4*>(rowp + c));      // the pitch is a multiple of 32 floats
        key[4 * v] = q.x; key[4 * v + 1] = q.y; key[4 * v + 2] = q.z; key[4 * v + 3] = q.w;
    }
    if (tid == 0) {
        s.mn = 0xffffffffu; s.mx = 0; s.below = 0u; s.ncand = 0u; s.v1 = 0u; s.v2 = 0u; s.min_gt = 0xffffffffu; s.out[2] = 0u;
    }
    uint4* h4 = reinterpret_cast<uint4*>(s.hist);
#pragma unroll
    for (int i = 0; i < kSelCtaBins / 4 / NT; ++i) h4[tid + NT * i] = make_uint4(0u, 0u, 0u, 0u);
    {                                                       // the vector that straddles the end of the row
        const int cb = cols & ~3, vb = cb / (4 * NT);
        if ((cols & 3) && c0 + 4 * NT * vb == cb) {
#pragma unroll
            for (int v = 0; v < V; ++v)
                if (v == vb) {
#pragma unroll
                    for (int j = 1; j < 4; ++j)
                        if (j >= (cols & 3)) key[4 * v + j] = kNoKey;
                }
        }
    }
    const int nv = cols / (4 * NT);                         // vector slots that are inside the row for every thread
    bool have = nv >= 1;                                    // uniform
    if (have && tid % (NT / 32) == 0) {                     // sample i: column 4 tid + 4 NT (i nv / 32)
        const int i = tid / (NT / 32), vs = (i * nv) >> 5;   // 0 .. nv - 1, spread over the sample
        unsigned sv = key[0];
#pragma unroll
        for (int v = 1; v < V; ++v)
            if (v == vs) sv = key[4 * v];
        s.sample[i] = sv;
    }
    __syncthreads();
    unsigned L = 0u, span = 0u;
    if (have) {
        const unsigned x = warp_sort32(s.sample[lane], lane);
        L = __shfl_sync(0xffffffffu, x, 6);
        span = __shfl_sync(0xffffffffu, x, 25) - L;
    }
    const unsigned k0 = (unsigned)((cols - 1) / 2);
    const bool even = (cols & 1) == 0;
    unsigned v1 = 0u, v2 = 0u;
#pragma unroll 1
    for (int attempt = 0;; ++attempt) {                     // one trip for all but a handful of rows
        if (attempt > 0) {
            __syncthreads();
            if (tid == 0) { s.below = 0u; s.ncand = 0u; s.v1 = 0u; s.v2 = 0u; s.out[2] = 0u; }
#pragma unroll
            for (int i = 0; i < kSelCtaBins / 4 / NT; ++i) h4[tid + NT * i] = make_uint4(0u, 0u, 0u, 0u);
            __syncthreads();
        }
        if (!have) {                                        // bracket = the row's extremes
            unsigned mn = 0xffffffffu;
            int mx = 0;
#pragma unroll
            for (int i = 0; i < 4 * V; ++i) { mn = min(mn, key[i]); mx = max(mx, (int)key[i]); }
            mn = __reduce_min_sync(0xffffffffu, mn);
            mx = __reduce_max_sync(0xffffffffu, mx);
            if (lane == 0) { atomicMin(&s.mn, mn); atomicMax(&s.mx, mx); }
            __syncthreads();
            L = s.mn;
            span = (unsigned)s.mx - L;
            have = true;
        }
        const int shift = max(0, 32 - __clz(span) - kSelCtaBits);       // span >> shift < 4,096
        span |= (1u << shift) - 1u;                         // the last bin in full: the bins tile the bracket exactly
        unsigned below = 0u;
#pragma unroll
        for (int i = 0; i < 4 * V; ++i) {                   // no branch: a key outside the bracket (below: d wraps) is
            const unsigned d = key[i] - L;                  // clamped to the lane's overflow word
            below += d >> 31;                               // key < L (both below 2^31; a pad counts only when L = 0)
            atomicAdd(s.hist + min(d >> shift, (unsigned)(kSelCtaBins + lane)), 1u);
        }
        if (L == 0u) below = 0u;
        below = __reduce_add_sync(0xffffffffu, below);
        if (lane == 0 && below) atomicAdd(&s.below, below);
        __syncthreads();
        sel_scan_cta_v<NT>(s.hist, k0 - s.below, s.tot, s.out);          // a rank below the bracket wraps: not found
        const unsigned b = s.out[0], kin = s.out[1], hb = s.out[2];
        if (hb == 0u) {                                     // the median is outside the sampled bracket
            if (attempt > 8) break;                         // (cannot happen with the extremes as the bracket)
            have = false;
            continue;
        }
        const unsigned base = L + (b << shift);             // the bin = keys [base, base + 2^shift)
        bool second_above = even && kin + 1u >= hb;         // the next order statistic lies above this bin
        if (shift == 0) {
            v1 = v2 = base;
        } else if (hb <= (unsigned)NT) {
            // the bin's keys -> shared memory; rank k of them = the largest key with at most k keys below it
            const unsigned width = 1u << shift;
#pragma unroll
            for (int v = 0; v < V; ++v) {                   // one branch per four keys: a hit is rare
                const unsigned e0 = key[4 * v] - base, e1 = key[4 * v + 1] - base, e2 = key[4 * v + 2] - base, e3 = key[4 * v + 3] - base;
                if (min(min(e0, e1), min(e2, e3)) < width) {
                    if (e0 < width) s.cand[atomicAdd(&s.ncand, 1u)] = key[4 * v];
                    if (e1 < width) s.cand[atomicAdd(&s.ncand, 1u)] = key[4 * v + 1];
                    if (e2 < width) s.cand[atomicAdd(&s.ncand, 1u)] = key[4 * v + 2];
                    if (e3 < width) s.cand[atomicAdd(&s.ncand, 1u)] = key[4 * v + 3];
                }
            }
            __syncthreads();
            if ((unsigned)(tid & ~31) < hb) {               // warps that own a candidate
                const bool mine = (unsigned)tid < hb;
                const unsigned my = mine ? s.cand[tid] : 0u;
                unsigned lt = 0u;
#pragma unroll 4
                for (unsigned i = 0; i < hb; ++i) lt += (s.cand[i] - my) >> 31;       // keys < 2^31: x < my <=> borrow
                const unsigned a1 = __reduce_max_sync(0xffffffffu, mine && lt <= kin ? my : 0u);
                const unsigned a2 = __reduce_max_sync(0xffffffffu, mine && lt <= kin + 1u ? my : 0u);
                if (lane == 0) { atomicMax(&s.v1, a1); atomicMax(&s.v2, a2); }
            }
            __syncthreads();
            v1 = s.v1;
            v2 = s.v2;
        } else {                                            // heavy ties: the bin is the next bracket
            L = base;
            span = (1u << shift) - 1u;
            continue;
        }
        if (second_above) {                                 // smallest key above the bin (the pad value is above all)
            const unsigned lim = base + (1u << shift);
            unsigned gt = 0xffffffffu;                      // min of key - lim over the keys >= lim (the others wrap high)
#pragma unroll
            for (int i = 0; i < 4 * V; ++i) gt = min(gt, key[i] - lim);
            gt = __reduce_min_sync(0xffffffffu, gt);
            if (lane == 0) atomicMin(&s.min_gt, gt);
            __syncthreads();
            v2 = s.min_gt + lim;
        }
        break;
    }
    if (tid < 32) {                                         // one warp turns the order statistics into the row's threshold
        const float a = __uint_as_float(w.amax[item]);      // (divisions and a walk to the rounding boundary: ~100
        float med = __fdiv_rn(__uint_as_float(v1), a);      // instructions that 16 warps would each repeat)
        if (even) med = __fdiv_rn(__fadd_rn(med, __fdiv_rn(__uint_as_float(v2), a)), 2.0f);   // np.mean of the two middle values
        const float t = unscaled_threshold(__fmul_rn(3.0f, med), a);
        if (lane == 0) s.tr = t;
    }
    __syncthreads();
    const float tr = s.tr;
    const float* thr = w.colthr + (size_t)item * w.col_stride;
    unsigned* bits = w.bits + (size_t)item * w.bits_stride + (size_t)row * im.wpr;
    const int sh = 4 * (lane & 7);
#pragma unroll
    for (int v = 0; v < V; ++v) {
        const int c = c0 + 4 * NT * v;
        if (4 * NT * v + 128 * (tid >> 5) >= 32 * im.wpr) break;          // warp-uniform: none of its words exists
        unsigned nib = 0u;
        if (c < cols) {                                     // the stride of the thresholds is a multiple of 32 floats
            const float4 t4 = __ldg(reinterpret_cast<const float4*>(thr + c));
            const float x0 = __uint_as_float(key[4 * v]), x1 = __uint_as_float(key[4 * v + 1]);
            const float x2 = __uint_as_float(key[4 * v + 2]), x3 = __uint_as_float(key[4 * v + 3]);
            nib = (x0 > t4.x && x0 > tr ? 1u : 0u) | (x1 > t4.y && x1 > tr ? 2u : 0u) | (x2 > t4.z && x2 > tr ? 4u : 0u) |
                  (x3 > t4.w && x3 > tr ? 8u : 0u);
        }
        unsigned word = nib << sh;
        word |= __shfl_xor_sync(0xffffffffu, word, 1);
        word |= __shfl_xor_sync(0xffffffffu, word, 2);
        word |= __shfl_xor_sync(0xffffffffu, word, 4);
        if ((lane & 7) == 0 && (c >> 5) < im.wpr) bits[c >> 5] = word;
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
// (recordings of a wave that are much shorter just predicate the tail off).  aligned_rows: every row of every matrix
// starts on a 16-byte boundary and is padded to whole 16-byte groups (the fused chain's own matrices: pitch = a
// multiple of 32 floats) -> the vector kernel; a caller's matrix of any pitch -> the scalar one.
// CAC_SN_ROWS_VARIANT=0 forces the scalar kernel (A/B runs).
cudaError_t launch_sn_rowmed_mask(const SnWave& w, int max_cols, bool aligned_rows, cudaStream_t st) {
    dim3 grid((unsigned)w.rows, (unsigned)w.n_items);
    static const int variant = getenv("CAC_SN_ROWS_VARIANT") ? atoi(getenv("CAC_SN_ROWS_VARIANT")) : 1;
    if (aligned_rows && variant != 0) {
        if (max_cols <= 256 * 8) sn_rowmed_mask_vec_kernel<256, 2, 4><<<grid, 256, 0, st>>>(w);
        else if (max_cols <= 256 * 16) sn_rowmed_mask_vec_kernel<256, 4, 4><<<grid, 256, 0, st>>>(w);
        else if (max_cols <= 256 * 32) sn_rowmed_mask_vec_kernel<256, 8, 4><<<grid, 256, 0, st>>>(w);
        else if (max_cols <= 512 * 24) sn_rowmed_mask_vec_kernel<512, 6, 2><<<grid, 512, 0, st>>>(w);   // one minute: 10,250
        else {
            sn_rowmed_mask_vec_kernel<512, 12, 1><<<grid, 512, 0, st>>>(w);      // rows up to 24,576 columns
            if (max_cols > kSnMaxColsFast) sn_rowmed_mask_long_kernel<<<grid, 256, 0, st>>>(w);   // the longer ones
        }
        return cudaGetLastError();
    }
    if (max_cols <= 256 * 8) sn_rowmed_mask_kernel<256, 8><<<grid, 256, 0, st>>>(w);
    else if (max_cols <= 256 * 16) sn_rowmed_mask_kernel<256, 16><<<grid, 256, 0, st>>>(w);
    else if (max_cols <= 256 * 32) sn_rowmed_mask_kernel<256, 32><<<grid, 256, 0, st>>>(w);
    else if (max_cols <= 256 * 44) sn_rowmed_mask_kernel<512, 22, 2><<<grid, 512, 0, st>>>(w);
    else {
        sn_rowmed_mask_kernel<512, 48><<<grid, 512, 0, st>>>(w);         // rows up to 24,576 columns
        if (max_cols > kSnMaxColsFast) sn_rowmed_mask_long_kernel<<<grid, 256, 0, st>>>(w);   // the longer ones
    }
    return cudaGetLastError();
}

}  // namespace cac

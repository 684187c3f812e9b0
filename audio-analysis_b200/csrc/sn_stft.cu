// signal_noise's spectrum for a wave of recordings (src/identify_tracks.py:652-659):
//   spectogram = np.abs(librosa.stft(frames, n_fft=4096, hop_length=hop))      -> spec [2049][cols], unscaled
//   a_max = np.amax(spectogram)                                                -> amax (float bits, atomicMax)
//   column_medians = np.median(spectogram / a_max, axis=0)                     -> colmed [cols], unscaled
// Division by a_max is monotone, so the order statistic is selected on the unscaled magnitudes here and scaled
// where it is consumed (sn_rows.cu) - bit-identical to dividing first, and a_max is not known before this kernel ends.
//
// Same structure as stft_mel_kernel: persistent, one CTA of 12 warps per SM, work unit = a stage of
// frames_per_stage consecutive frames of one recording, samples of stage q + 1 arrive by cp.async while stage q is
// transformed, one warp per frame (fft4096.cuh).  The 12 warps of the CTA work on 12 consecutive frames; their
// buffers together are a [12][2049] tile that the CTA writes bin-major as contiguous 48-byte runs (rows of the
// matrix are padded to a pitch of 32 floats, so the runs are 16-byte aligned: full sectors, no partial-sector
// read-modify-write in L2 / HBM - per-lane 4-byte stores had cost 1.75x the bytes and 40 % of the kernel).  The warp
// that transformed a frame still holds its 2,049 magnitudes in registers and selects their median by a radix
// select whose histogram reuses the warp's buffer - no transpose, no second matrix, no extra pass over HBM.
#include "cac_common.cuh"
#include "fft4096.cuh"
#if CAC_FFT_PACKED
#include "fft4096p.cuh"
#endif
#include "sn.cuh"
#include "sn_select.cuh"

namespace cac {

struct SnStftSmem { int raw, stage, ex, ltab, total; };   // offsets in floats
__host__ __device__ inline SnStftSmem sn_stft_layout(int stage_half) {
    SnStftSmem L;
    int o = 0;
    L.raw = o; o += 2 * stage_half;
    L.stage = o; o += 2 * stage_half;
    L.ex = o; o += kWarpsPerCta * kExFloats;
    L.ltab = o; o += 32 * 32;
    L.total = o;
    return L;
}

// median (rank 1024 of the 2,049 magnitudes of one frame): lane holds bins lane + 32 i in v[], lane 0 also bin 2048;
// mn / mx = smallest / largest of the 2,049 keys (warp-uniform).  Radix select below their common prefix, 11 bits per pass.
__device__ __forceinline__ unsigned frame_median_bits(const float (&v)[64], float nyq, unsigned mn, unsigned mx,
                                                      unsigned* hist, int lane) {
    if (mn == mx) return mx;
    const int top = 32 - __clz(mn ^ mx);                  // bits [0, top) vary; keys are < 2^31, so top <= 31
    unsigned prefix = (mx >> top) << top;
    unsigned k = 1024u;
    uint4* h4 = reinterpret_cast<uint4*>(hist);
    unsigned* trash = hist + kSelWarpBins + lane;         // the warp's buffer has 2,112 words: 32 spare ones per lane
    const unsigned ekey = __float_as_uint(nyq);
    int remaining = top;
#pragma unroll 1
    while (remaining > 0) {
        const int w = min(kSelWarpBits, remaining), shift = remaining - w;
        const unsigned dmask = (1u << w) - 1u;
#pragma unroll
        for (int j = 0; j < kSelWarpBins / 4 / 32; ++j) h4[lane + 32 * j] = make_uint4(0u, 0u, 0u, 0u);
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 64; ++i) {
            const unsigned key = __float_as_uint(v[i]);
            smem_inc_sel(hist, (key >> shift) & dmask, ((key ^ prefix) >> remaining) == 0u, trash);
        }
        if (lane == 0) smem_inc_sel(hist, (ekey >> shift) & dmask, ((ekey ^ prefix) >> remaining) == 0u, trash);
        __syncwarp();
        const unsigned digit = sel_scan_warp(hist, lane, k);
        prefix |= digit << shift;
        remaining = shift;
        __syncwarp();                                   // every lane has read the histogram before it is cleared again
    }
    return prefix;
}

#if !CAC_SN_COLSEL_RADIX
// The same median through a sampled bracket (the idea of sn_rowmed_mask_vec_kernel at warp level): `samp` = 32 of the
// frame's magnitudes spread over the spectrum (one per lane), sorted with shuffles; ranks 6 and 25 bracket the median
// of all but one frame in about a thousand.  ONE pass then counts the keys below the bracket and puts the keys inside
// it into a 2,048-bin histogram that spans exactly the bracket ((key - L) >> shift), the bin that holds rank 1024 has
// one to three keys, they go to shared memory and are ranked by counting.  The counts are exact whatever the bracket;
// when the rank is outside it the pass is redone with the frame's extremes, a bin of more than 32 keys (ties) becomes
// the next bracket.  About half the instructions of the three radix passes above.
__device__ __forceinline__ unsigned sorted_lane(unsigned x, int lane) {
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

// bin of the warp's 2,048-bin histogram that holds rank k -> returned, k becomes the rank inside it, cnt the bin's
// count; 0xffffffff when k is not below the histogram's total.  All 32 lanes call.
__device__ __forceinline__ unsigned sel_scan_warp_total(const unsigned* hist, int lane, unsigned& k, unsigned& cnt) {
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
            if (kk >= t) { kk -= t; js = j + 1; } else found = true;
        }
    }
    if (!found) return 0xffffffffu;
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
    const int ls = __ffs(ball) - 1;                      // kk < the block's total: some lane's prefix exceeds it
    kk -= __shfl_sync(0xffffffffu, inc - cj, ls);
    const uint4 q = h4[ls + 32 * js];                    // the quad that holds the rank (same address for all lanes)
    unsigned b = 0, n = q.x;
    if (kk >= q.x) { kk -= q.x; b = 1; n = q.y; if (kk >= q.y) { kk -= q.y; b = 2; n = q.z; if (kk >= q.z) { kk -= q.z; b = 3; n = q.w; } } }
    k = kk;
    cnt = n;
    return 4u * (unsigned)(ls + 32 * js) + b;
}

__device__ __forceinline__ unsigned frame_median_bracketed(const float (&v)[64], float nyq, float samp, unsigned mxkey,
                                                           unsigned* hist, int lane) {
    uint4* h4 = reinterpret_cast<uint4*>(hist);           // the warp's buffer has 2,112 words: 32 spare ones per lane
    const unsigned ekey = __float_as_uint(nyq);
    unsigned L, span;
    {
        const unsigned x = sorted_lane(__float_as_uint(samp), lane);
        L = __shfl_sync(0xffffffffu, x, 6);
        span = __shfl_sync(0xffffffffu, x, 25) - L;
    }
#pragma unroll 1
    for (int attempt = 0; attempt < 12; ++attempt) {
#pragma unroll
        for (int j = 0; j < kSelWarpBins / 4 / 32; ++j) h4[lane + 32 * j] = make_uint4(0u, 0u, 0u, 0u);
        __syncwarp();
        const int shift = max(0, 32 - __clz(span) - kSelWarpBits);     // span >> shift < 2,048
        span |= (1u << shift) - 1u;                         // the last bin in full: the bins tile the bracket exactly
        unsigned below = 0u;
#pragma unroll
        for (int i = 0; i < 64; ++i) {
            const unsigned d = __float_as_uint(v[i]) - L;   // outside the bracket (below: d wraps): the lane's spare word
            below += d >> 31;                               // key < L: both are below 2^31
            atomicAdd(hist + min(d >> shift, (unsigned)(kSelWarpBins + lane)), 1u);
        }
        if (lane == 0) {
            const unsigned d = ekey - L;
            below += d >> 31;
            atomicAdd(hist + min(d >> shift, (unsigned)kSelWarpBins), 1u);
        }
        __syncwarp();
        below = __reduce_add_sync(0xffffffffu, below);
        unsigned k = 1024u - below, hb = 0u;                // a rank below the bracket wraps: not found
        const unsigned digit = sel_scan_warp_total(hist, lane, k, hb);
        __syncwarp();                                       // every lane has read the histogram before it is reused
        if (digit == 0xffffffffu) {                         // the median is outside the sampled bracket: the extremes
            float mn = nyq;
#pragma unroll
            for (int i = 0; i < 64; ++i) mn = fminf(mn, v[i]);
            L = __float_as_uint(warp_min(mn));
            span = mxkey - L;
            continue;
        }
        const unsigned base = L + (digit << shift);
        if (shift == 0) return base;
        const unsigned width = 1u << shift;
        if (hb > 32u) {                                     // ties: the bin is the next bracket
            L = base;
            span = width - 1u;
            continue;
        }
        // the bin's keys -> hist[0..31] (counter in hist[32]); rank k = the largest with at most k keys below it
        if (lane == 0) hist[32] = 0u;
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 64; i += 4) {                   // one branch per four keys: a hit is rare
            const unsigned e0 = __float_as_uint(v[i]) - base, e1 = __float_as_uint(v[i + 1]) - base;
            const unsigned e2 = __float_as_uint(v[i + 2]) - base, e3 = __float_as_uint(v[i + 3]) - base;
            if (min(min(e0, e1), min(e2, e3)) < width) {
                if (e0 < width) hist[atomicAdd(hist + 32, 1u)] = e0 + base;
                if (e1 < width) hist[atomicAdd(hist + 32, 1u)] = e1 + base;
                if (e2 < width) hist[atomicAdd(hist + 32, 1u)] = e2 + base;
                if (e3 < width) hist[atomicAdd(hist + 32, 1u)] = e3 + base;
            }
        }
        if (lane == 0 && ekey - base < width) hist[atomicAdd(hist + 32, 1u)] = ekey;
        __syncwarp();
        const bool mine = (unsigned)lane < hb;
        const unsigned my = mine ? hist[lane] : 0u;
        unsigned lt = 0u;
        for (unsigned i = 0; i < hb; ++i) lt += (__shfl_sync(0xffffffffu, my, (int)i) - my) >> 31;
        const unsigned r = __reduce_max_sync(0xffffffffu, mine && lt <= k ? my : 0u);
        __syncwarp();
        return r;
    }
    return 0u;                                              // not reached: the extremes always hold the rank
}
#endif

__global__ void __launch_bounds__(kWarpsPerCta * 32, 1) sn_stft_kernel(const float* __restrict__ pcm, const SnWave w,
                                                                       int hop, int fps, int stage_half) {
    extern __shared__ __align__(16) float smem[];
    const SnStftSmem L = sn_stft_layout(stage_half);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float* ex = smem + L.ex + warp * kExFloats;
    float* pbuf = ex;                                    // magnitudes reuse the exchange buffer (fft_frame), then the histogram does
    unsigned* hist = reinterpret_cast<unsigned*>(ex);
    float* ltab = smem + L.ltab;
    if (warp == 0) init_lane_table(ltab, lane);
    const float* lt = ltab + lane;

    const SnItem last = w.items[w.n_items - 1];
    const long long n_stages = last.stage_base + last.n_stages;
    const long long q_begin = n_stages * blockIdx.x / gridDim.x, q_end = n_stages * (blockIdx.x + 1) / gridDim.x;
    const int n_my = (int)(q_end - q_begin);
    if (n_my <= 0) return;
    int item = 0;
    while (w.items[item].stage_base + w.items[item].n_stages <= q_begin) ++item;
    int sidx = (int)(q_begin - w.items[item].stage_base);

    auto plane_buf = [&](int parity) { return smem + (parity ? L.stage : L.raw); };
    // staged element u of stage (item, sidx) = sample f0 * hop - 2048 + u of the recording, zero outside it
    // (librosa's centre padding); thread t owns u = t, t + 384, ...: fixed plane (384 is even)
    auto issue_stage = [&](int it_idx, int s, int parity) {
        const SnItem im = w.items[it_idx];
        const int f0 = s * fps;
        const int nfs = min(fps, im.cols - f0);
        const int len = (nfs - 1) * hop + kNfft;
        const float* src = pcm + im.pcm_off;
        const int r0 = f0 * hop - kNfft / 2;
        constexpr int kT = kWarpsPerCta * 32;
        float* dst = plane_buf(parity) + (tid & 1) * stage_half + (tid >> 1);
        const int n_steps = (len - tid + kT - 1) / kT;
        const int first = r0 + tid;
        int k_lo = first >= 0 ? 0 : (-first + kT - 1) / kT;
        int k_hi = im.n_samples > first ? (im.n_samples - first + kT - 1) / kT : 0;
        k_lo = min(k_lo, n_steps);
        k_hi = max(k_lo, min(k_hi, n_steps));
        for (int k = 0; k < k_lo; ++k) dst[k * (kT / 2)] = 0.f;
        const float* g = src + first + (long long)k_lo * kT;
        float* d = dst + k_lo * (kT / 2);
#pragma unroll 4
        for (int k = k_lo; k < k_hi; ++k, g += kT, d += kT / 2) cp_async4(d, g, true);
        for (int k = k_hi; k < n_steps; ++k) dst[k * (kT / 2)] = 0.f;
        cp_async_commit();
    };

    issue_stage(item, sidx, 0);
    for (int it = 0; it < n_my; ++it) {
        const SnItem im = w.items[item];
        const int f0 = sidx * fps;
        const int nfs = min(fps, im.cols - f0);
        float* se = plane_buf(it & 1);
        float* so = se + stage_half;
        const bool last_of_item = sidx + 1 == im.n_stages;

        cp_async_wait_all();
        __syncthreads();                       // stage `it` has landed; every warp is done with stage it - 1
        if (it + 1 < n_my) issue_stage(last_of_item ? item + 1 : item, last_of_item ? 0 : sidx + 1, (it + 1) & 1);

        float* spec = w.spec + (size_t)item * w.spec_stride;
        float* colmed = w.colmed + (size_t)item * w.col_stride;
        float vmax = 0.f;
        // rounds of 12 consecutive frames, one per warp: transform, then the CTA writes each bin's 12 values as ONE
        // contiguous 48-byte run (the 12 warps' buffers together are a [12][2049] tile), then every warp selects its
        // frame's median with its buffer as the histogram
#pragma unroll 1
        for (int r0 = 0; r0 < nfs; r0 += kWarpsPerCta) {
            const int fi = r0 + warp;
            const bool active = fi < nfs;
            float v[64];
            float nyq = 0.f, samp = 0.f;
            if (active) {
#if CAC_FFT_PACKED
                fft_frame_p<1, 32, true>(se, so, fi * hop, ex, pbuf, lane, lt);
#else
                fft_frame<1, 32, true>(se, so, fi * hop, ex, pbuf, lane, lt);
#endif
                __syncwarp();                  // lane 0's bin 2048 is visible
#pragma unroll
                for (int i = 0; i < 64; ++i) v[i] = pbuf[lane + 32 * i];     // the lane's own values, 2 |X[k]|
                nyq = pbuf[2048];
#if !CAC_SN_COLSEL_RADIX
                samp = pbuf[64 * lane + 21];                                  // 32 magnitudes spread over the spectrum
#endif
            }
            __syncthreads();                   // the round's magnitudes are complete in the 12 buffers
            {
                const int nr = min(kWarpsPerCta, nfs - r0);
                const float* tile = smem + L.ex;                          // frame j of the round: tile[j * kExFloats + bin]
                float* orow = spec + (size_t)tid * im.pitch + (f0 + r0);
                const size_t ostep = (size_t)(kWarpsPerCta * 32) * im.pitch;
                if (nr == kWarpsPerCta) {      // f0 + r0 is a multiple of 4 and the pitch of 32: 16-byte aligned runs
                    // three consecutive lanes write the three 16-byte quarters of one bin's 48-byte run: a warp's store
                    // covers 10 contiguous runs (20 sectors) instead of 32 scattered 16-byte pieces (32 sectors)
                    constexpr int kRowsPerIt = kWarpsPerCta * 32 / 3;
                    const int sub = tid / 3, q = tid - 3 * sub;
                    const float* t = tile + (4 * q) * kExFloats;
                    float* o = spec + (size_t)sub * im.pitch + (f0 + r0) + 4 * q;
                    for (int bin = sub; bin < kSnRows; bin += kRowsPerIt, o += (size_t)kRowsPerIt * im.pitch)
                        *reinterpret_cast<float4*>(o) = make_float4(0.5f * t[bin], 0.5f * t[kExFloats + bin],
                                                                    0.5f * t[2 * kExFloats + bin], 0.5f * t[3 * kExFloats + bin]);
                } else {
                    for (int bin = tid; bin < kSnRows; bin += kWarpsPerCta * 32, orow += ostep)
                        for (int j = 0; j < nr; ++j) orow[j] = 0.5f * tile[j * kExFloats + bin];
                }
            }
            __syncthreads();                   // the buffers are free: they become the histograms
            if (active) {
#if CAC_SN_COLSEL_RADIX
                float mx = nyq, mn = nyq;
#pragma unroll
                for (int i = 0; i < 64; ++i) { mx = fmaxf(mx, v[i]); mn = fminf(mn, v[i]); }
                mx = warp_max(mx);
                mn = warp_min(mn);
                vmax = fmaxf(vmax, mx);
                const unsigned med = frame_median_bits(v, nyq, __float_as_uint(mn), __float_as_uint(mx), hist, lane);
#else
                float mx = nyq;
#pragma unroll
                for (int i = 0; i < 64; ++i) mx = fmaxf(mx, v[i]);
                mx = warp_max(mx);
                vmax = fmaxf(vmax, mx);
                const unsigned med = frame_median_bracketed(v, nyq, samp, __float_as_uint(mx), hist, lane);
#endif
                if (lane == 0) colmed[f0 + fi] = 0.5f * __uint_as_float(med);
            }
        }
        if (lane == 0) atomicMax(w.amax + item, __float_as_uint(0.5f * vmax));
        if (last_of_item) { ++item; sidx = 0; } else ++sidx;
    }
}

// item table of a wave (one thread: a wave holds at most a few dozen recordings)
__global__ void sn_prep_kernel(const long long* __restrict__ off, const long long* __restrict__ len, int n, int hop, int fps,
                               long long max_len, SnItem* __restrict__ items, unsigned* __restrict__ amax) {
    if (threadIdx.x == 0) {
        int base = 0;
        for (int i = 0; i < n; ++i) {
            SnItem im;
            im.pcm_off = off[i];
            im.n_samples = (int)max(0LL, min(len[i], max_len));   // never beyond the slot the caller sized
            im.cols = 1 + im.n_samples / hop;
            im.wpr = (im.cols + 31) / 32;
            im.stage_base = base;
            im.n_stages = (im.cols + fps - 1) / fps;
            im.pitch = (im.cols + 31) / 32 * 32;
            base += im.n_stages;
            items[i] = im;
        }
    }
    for (int i = threadIdx.x; i < n; i += blockDim.x) amax[i] = 0u;
}

// ---- host side ------------------------------------------------------------------------------------------------------
size_t sn_stft_smem_bytes(int frames_per_stage, int hop) {
    return sizeof(float) * (size_t)sn_stft_layout(stage_half_floats(frames_per_stage, hop)).total;
}

int sn_stft_frames_per_stage(int hop) {
    static const int kChoices[] = {24, 12, 10, 8, 6, 4, 3, 2, 1};
    for (int fs : kChoices)
        if (sn_stft_smem_bytes(fs, hop) <= 227 * 1024) return fs;
    return 0;
}

cudaError_t launch_sn_prep(const long long* d_item_off, const long long* d_item_len, int n_items, int hop,
                           int frames_per_stage, long long max_item_samples, SnItem* items, unsigned* amax, cudaStream_t st) {
    sn_prep_kernel<<<1, 64, 0, st>>>(d_item_off, d_item_len, n_items, hop, frames_per_stage, max_item_samples, items, amax);
    return cudaGetLastError();
}

cudaError_t launch_sn_stft(const float* pcm, const SnWave& w, int hop, int frames_per_stage, cudaStream_t st) {
    const size_t smem = sn_stft_smem_bytes(frames_per_stage, hop);
    cudaError_t e = cudaFuncSetAttribute(sn_stft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // persistent: one CTA per SM; the stage count lives on the device, CTAs without stages leave at once
    sn_stft_kernel<<<(unsigned)sms, kWarpsPerCta * 32, smem, st>>>(pcm, w, hop, frames_per_stage,
                                                                   stage_half_floats(frames_per_stage, hop));
    return cudaGetLastError();
}

}  // namespace cac

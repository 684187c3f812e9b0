// Fused framed real-FFT -> power -> banded mel projection for classifier segments.
//
// Replaces, for every segment of load_samples' loop (src/identify_tracks.py:163-190):
//   normalize_data (:202-209) -> librosa.stft(n_fft=4096, hop) (:243) -> |.|**power
//   (src/custommel.py:61) -> mel_f(...).dot(magnitude) (src/custommel.py:62-63)
// The [2049 x frames] spectrum never leaves the SM: one warp transforms one frame (fft4096.cuh:
// 4096 real points = 2048-point complex FFT as radix-64 x radix-32 in registers with a single
// shared-memory transpose) and parks the frame's power spectrum in its own shared-memory buffer;
// the sparse triangular filterbank is applied by the same warp as banded dot products with
// lanes = mel rows (128-bit loads of zero-padded tap-major weights kept in shared memory), and the
// CTA writes each stage's mel values as contiguous row segments.
#include "cac_common.cuh"
#include "fft4096.cuh"
#if CAC_FFT_PACKED
#include "fft4096p.cuh"
#endif

namespace cac {


// Shared-memory carve-up of stft_mel_kernel (floats unless noted); the launcher uses the same function.
struct FeatSmem {
    int raw;           // staging buffer of the odd stages: 2 planes (even / odd samples) of stage_half floats
    int stage;         // staging buffer of the even stages, same shape
    int ex;            // 8 warps x kExFloats
    int wts;           // padded tap-major band weights (FeatArgs::wpad)
    int bands;         // int4 slot table [n_groups * 32] (cac_api.cu)
    int ltab;          // 32 x 32
    int otile;         // n_mels x (frames_per_stage + 1): this stage's mel values, flushed as row segments
    int total;
};
__host__ __device__ inline FeatSmem feat_smem_layout(int stage_half, int n_taps, int n_mels, int frames_per_stage) {
    FeatSmem L;
    int o = 0;
    L.raw = o; o += 2 * stage_half;
    L.stage = o; o += 2 * stage_half;
    L.ex = o; o += kWarpsPerCta * kExFloats;
    L.wts = o; o += (n_taps + 3) / 4 * 4;
    L.bands = o; o += (n_mels + 31) / 32 * 32 * 4;
    o = (o + 31) / 32 * 32;
    L.ltab = o; o += 32 * 32;
    L.otile = o; o += n_mels * (frames_per_stage + 1);
    L.total = o;
    return L;
}

// Persistent kernel, one CTA (12 warps) per SM.  Work unit = a "stage": frames_per_stage consecutive
// frames of one segment.  The samples of stage q+1 are fetched with cp.async (zero-filled outside the
// real samples) straight into the other pair of even / odd sample planes while the warps transform
// stage q; each warp owns whole frames: FFT -> power spectrum (own buffer) -> banded mel gather -> tile.
template <int POWER, int K2N>
__global__ void __launch_bounds__(kWarpsPerCta * 32, 1) stft_mel_kernel(const FeatArgs a) {
    extern __shared__ __align__(16) float smem[];
    const FeatSmem L = feat_smem_layout(a.stage_half, a.n_taps, a.n_mels, a.frames_per_stage);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float* ex = smem + L.ex + warp * kExFloats;
    float* pbuf = ex;   // the power spectrum reuses the warp's exchange buffer once pass 2 has read it
    float* wts = smem + L.wts;
    int4* slots = reinterpret_cast<int4*>(smem + L.bands);
    float* ltab = smem + L.ltab;
    float* otile = smem + L.otile;
    const int tstride = a.frames_per_stage + 1;       // odd row stride: conflict-free both ways

    // ---- one-time tables ------------------------------------------------------------------------
    for (int i = tid; i < a.n_taps; i += kWarpsPerCta * 32) wts[i] = a.wpad[i];
    for (int m = tid; m < a.n_groups * 32; m += kWarpsPerCta * 32) slots[m] = reinterpret_cast<const int4*>(a.mstart)[m];
    if (warp == 0) init_lane_table(ltab, lane);
    const float* lt = ltab + lane;

    const long long n_stages = a.n_segs * a.stages_per_seg;
    // two staging buffers, each a pair of planes (even / odd staged samples): stage q lives in buffer q & 1
    auto plane_buf = [&](int parity) { return smem + (parity ? L.stage : L.raw); };
    // Stage q = frames [f0, f0+nfs) of one segment = staged elements u in [0, len), element u being sample
    // s = s0 + u of the zero-padded segment: the RAW sample (no normalisation, see below), zero in the np.pad
    // region and outside the segment (librosa's centre padding).  4-byte cp.async straight into the planes:
    // no alignment cases, no staging copy, no de-interleave pass.
    // stages are addressed as (segment, stage within the segment, buffer parity): no divisions in the loop
    auto issue_stage = [&](int seg, int sidx, int parity) {
        const int f0 = sidx * a.frames_per_stage;
        const int nfs = min(a.frames_per_stage, a.n_frames - f0);
        const int len = (nfs - 1) * a.hop + kNfft;
        const cac_segment sd = a.segs[seg];
        const float* src = a.pcm + sd.pcm_offset;
        const int r0 = f0 * a.hop - kNfft / 2 - sd.pad_left;      // index into the segment's real samples of u = 0
        // thread t owns u = t, t + 384, ...: fixed plane (384 is even), plane index + 192 and source + 384 per step;
        // [k_lo, k_hi) are its steps that meet a real sample, the others are zero-filled with plain stores
        constexpr int kT = kWarpsPerCta * 32;
        float* dst = plane_buf(parity) + (tid & 1) * a.stage_half + (tid >> 1);
        const int n_steps = (len - tid + kT - 1) / kT;                   // u = tid + kT k < len
        const int first = r0 + tid;                                      // r of step 0
        int k_lo = first >= 0 ? 0 : (-first + kT - 1) / kT;
        int k_hi = sd.n_valid > first ? (sd.n_valid - first + kT - 1) / kT : 0;
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

    // The warps park a stage's mel values in otile ([mel][frame]); the whole CTA then writes every mel row's
    // frames_per_stage consecutive frames as one contiguous run (full 32-byte sectors instead of 4-byte
    // stores scattered over n_mels rows: no read-modify-write fills in L2 / HBM).
    auto flush_tile = [&](int seg, int sidx) {
        const int f0 = sidx * a.frames_per_stage;
        const int nfs = min(a.frames_per_stage, a.n_frames - f0);
        float* outseg = a.out + (size_t)seg * a.n_mels * a.n_frames + f0;
        constexpr int kThreads = kWarpsPerCta * 32;
        if (kThreads % nfs == 0) {               // full stages (24 frames): a fixed (row, frame) walk, no divisions
            const int rows_per_iter = kThreads / nfs;
            int m = tid / nfs;
            const int f = tid - m * nfs;
            float* gp = outseg + (size_t)m * a.n_frames + f;
            const float* sp = otile + m * tstride + f;
            const size_t gstep = (size_t)rows_per_iter * a.n_frames;
            const int sstep = rows_per_iter * tstride;
            if (a.mag_on) { for (; m < a.n_mels; m += rows_per_iter, gp += gstep, sp += sstep) *gp = powf(*sp, a.mag_p); }
            else { for (; m < a.n_mels; m += rows_per_iter, gp += gstep, sp += sstep) *gp = *sp; }
        } else {
            for (int t = tid; t < a.n_mels * nfs; t += kThreads) {
                const int m = t / nfs, f = t - m * nfs;
                const float v = otile[m * tstride + f];
                outseg[(size_t)m * a.n_frames + f] = a.mag_on ? powf(v, a.mag_p) : v;
            }
        }
    };

    // normalize_data is affine, x_n = alpha x + beta inside the segment (alpha = 2 / range,
    // beta = 2 (1e-6 - 0.5 - min / range)), and the FFT is linear: for a frame that lies inside the segment
    // FFT(w x_n) = alpha FFT(w x) + beta FFT(w), and the periodic Hann window's spectrum FFT(w) is zero
    // from bin 2 on.  So the frames are transformed un-normalised and the mel values are scaled by alpha^2
    // (power 2) or |alpha| (power 1).  Only stages that reach over a segment end (centre padding stays
    // zero, so x_n = beta there would be wrong) - and every stage when a band touches bins 0 / 1 - get
    // gamma = beta / alpha added to their in-segment samples first.
    // blocked distribution: every CTA owns a contiguous run of stages, so the samples two consecutive
    // stages (and the two 50 %-overlapping segments) share are re-read by the SAME CTA a moment later
    const long long q_begin = n_stages * blockIdx.x / gridDim.x, q_end = n_stages * (blockIdx.x + 1) / gridDim.x;
    const int n_my = (int)(q_end - q_begin);
    int seg = (int)(q_begin / a.stages_per_seg);
    int sidx = (int)(q_begin - (long long)seg * a.stages_per_seg);
    if (n_my > 0) issue_stage(seg, sidx, 0);
    for (int it = 0; it < n_my; ++it) {
        const int f0 = sidx * a.frames_per_stage;
        const int nfs = min(a.frames_per_stage, a.n_frames - f0);
        const int len = (nfs - 1) * a.hop + kNfft;
        float* se = plane_buf(it & 1);
        float* so = se + a.stage_half;
        const bool last_of_seg = sidx + 1 == a.stages_per_seg;

        cp_async_wait_all();
        __syncthreads();                       // stage q has landed; every warp is done with stage q - 1
        if (it + 1 < n_my) issue_stage(last_of_seg ? seg + 1 : seg, last_of_seg ? 0 : sidx + 1, (it + 1) & 1);   // into the buffer stage q - 1 used
        if (it > 0) flush_tile(sidx > 0 ? seg : seg - 1, sidx > 0 ? sidx - 1 : a.stages_per_seg - 1);
        float sc = 1.f;                        // output scale: alpha^2 / |alpha|
        bool seg_nan = false;                  // constant segment: normalize_data divides 0 by 0 -> every value is NaN
        if (a.normalize) {
            const float2 st = a.segstat[seg];
            const float range = __fsub_rn(st.y, st.x);
            seg_nan = !(range > 0.f);
            const float alpha = __fdividef(2.0f, range);
            // ranges far from 1 (not audio in [-1, 1]): normalise the staged samples explicitly, scale stays 1
            const bool explicit_norm = !(range > 1e-9f && range < 1e9f);
            const int s0 = f0 * a.hop - kNfft / 2;                                  // segment index of u = 0
            const bool edge = s0 < 0 || s0 + len > a.seg_len || a.kmin < 2;
            if (!explicit_norm) sc = (POWER == 2) ? alpha * alpha : alpha;
            if (edge || explicit_norm) {
                const double gamma_d = (1e-6 - 0.5) * (double)range - (double)st.x;     // beta / alpha
                const float mul = explicit_norm ? (float)(2.0 / (double)range) : 1.f;
                const float add = explicit_norm ? (float)(2.0 / (double)range * gamma_d) : (float)gamma_d;
                for (int u = tid; u < len; u += kWarpsPerCta * 32) {
                    const int sidx = s0 + u;
                    if (sidx >= 0 && sidx < a.seg_len) {
                        float* e = se + (u & 1) * a.stage_half + (u >> 1);
                        *e = fmaf(*e, mul, add);
                    }
                }
            }
            if (seg_nan) sc = __uint_as_float(0x7fc00000u);
        }
        __syncthreads();                       // the tile is flushed, the planes are final

        float vmax = 0.f;
#pragma unroll 1
        for (int fi = warp; fi < nfs; fi += kWarpsPerCta) {
#if CAC_FFT_PACKED
            fft_frame_p<POWER, K2N>(se, so, fi * a.hop, ex, pbuf, lane, lt);
#else
            fft_frame<POWER, K2N>(se, so, fi * a.hop, ex, pbuf, lane, lt);
#endif
#pragma unroll 1
            for (int i = 64 * K2N + lane; i < a.pbuf_len; i += 32) pbuf[i] = 0.f;  // zero-weight taps past the computed bins read zeros
            __syncwarp();
            // banded mel projection of this frame: lane owns mel rows lane, lane+32, ...
#pragma unroll 1
            for (int g = 0; g < a.n_groups; ++g) {
                const int4 slot = slots[g * 32 + lane];      // band start | tile offset of the row | weight offset | quads
                // zero weights in front of the band: both loads are 128-bit
                const float4* pp = reinterpret_cast<const float4*>(pbuf + slot.x);
                const float4* wg = reinterpret_cast<const float4*>(wts + slot.z);
                const int gm4 = slot.w;
                // four independent partial sums (one per tap residue mod 4): fixed order, so deterministic,
                // and the FMA chains of one group overlap instead of serialising on one accumulator
                const float4* const pend = pp + gm4;
#if CAC_FFT_PACKED
                float2 acc01 = make_float2(0.f, 0.f), acc23 = make_float2(0.f, 0.f);   // the same four partial sums, packed
#pragma unroll 2
                for (; pp != pend; ++pp, wg += 32) {
                    const float4 w = *wg;
                    const float4 p = *pp;
                    acc01 = __ffma2_rn(make_float2(w.x, w.y), make_float2(p.x, p.y), acc01);
                    acc23 = __ffma2_rn(make_float2(w.z, w.w), make_float2(p.z, p.w), acc23);
                }
                const float acc = ((acc01.x + acc01.y) + (acc23.x + acc23.y)) * sc;
#else
                float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, acc3 = 0.f;
#pragma unroll 2
                for (; pp != pend; ++pp, wg += 32) {        // the padding taps add +0
                    const float4 w = *wg;
                    const float4 p = *pp;
                    acc0 = fmaf(w.x, p.x, acc0);
                    acc1 = fmaf(w.y, p.y, acc1);
                    acc2 = fmaf(w.z, p.z, acc2);
                    acc3 = fmaf(w.w, p.w, acc3);
                }
                const float acc = ((acc0 + acc1) + (acc2 + acc3)) * sc;
#endif
                if (slot.y >= 0) {
                    otile[slot.y + fi] = acc;
                    vmax = fmaxf(vmax, acc);
                }
            }
            __syncwarp();
        }
        vmax = warp_max(vmax);
        if (seg_nan) vmax = __uint_as_float(0x7fc00000u);     // fmaxf drops NaNs; np.max keeps them (bits order above every finite value)
        if (lane == 0) atomicMax(a.segmax + seg, __float_as_uint(vmax));
        if (last_of_seg) { ++seg; sidx = 0; } else ++sidx;
    }
    __syncthreads();
    if (n_my > 0) flush_tile(sidx > 0 ? seg : seg - 1, sidx > 0 ? sidx - 1 : a.stages_per_seg - 1);
}

// ---- per-segment min / max of the zero-padded segment (normalize_data's reductions) ---------
__global__ void __launch_bounds__(256) seg_minmax_kernel(const float* __restrict__ pcm,
                                                         const cac_segment* __restrict__ segs,
                                                         float2* __restrict__ segstat, unsigned* segmax,
                                                         int seg_len) {
    __shared__ float smn[8], smx[8];
    const long long seg = blockIdx.x;
    const cac_segment sd = segs[seg];
    const float* p = pcm + sd.pcm_offset;
    float mn = INFINITY, mx = -INFINITY;
    // scalar head up to the first 16-byte boundary, 128-bit body (four loads in flight per thread), scalar tail
    const int head = min(sd.n_valid, (int)((4 - ((reinterpret_cast<uintptr_t>(p) >> 2) & 3)) & 3));
    const int n4 = (sd.n_valid - head) >> 2;
    if ((int)threadIdx.x < head) { const float v = __ldg(p + threadIdx.x); mn = fminf(mn, v); mx = fmaxf(mx, v); }
    const float4* p4 = reinterpret_cast<const float4*>(p + head);
#pragma unroll 4
    for (int i = threadIdx.x; i < n4; i += 256) {
        const float4 v = __ldg(p4 + i);
        mn = fminf(fminf(mn, v.x), fminf(v.y, fminf(v.z, v.w)));
        mx = fmaxf(fmaxf(mx, v.x), fmaxf(v.y, fmaxf(v.z, v.w)));
    }
    for (int i = head + 4 * n4 + threadIdx.x; i < sd.n_valid; i += 256) {
        const float v = __ldg(p + i);
        mn = fminf(mn, v);
        mx = fmaxf(mx, v);
    }
    mn = warp_min(mn);
    mx = warp_max(mx);
    if ((threadIdx.x & 31) == 0) { smn[threadIdx.x >> 5] = mn; smx[threadIdx.x >> 5] = mx; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) { mn = fminf(mn, smn[w]); mx = fmaxf(mx, smx[w]); }
        if (sd.n_valid < seg_len) { mn = fminf(mn, 0.f); mx = fmaxf(mx, 0.f); }   // the np.pad zeros
        segstat[seg] = make_float2(mn, mx);
        segmax[seg] = 0u;
    }
}

// ---- librosa.power_to_db(S, ref=np.max), top_db = 80, per item ------------------------------
// explicit _rn intrinsics: no FMA contraction, so the block maximum maps to exactly 0 dB as in numpy;
// log_spec.max() is refdb - refdb = 0 when ref is the block maximum, so the floor is -top_db.
// Comparisons instead of fmaxf so that NaN goes through as in np.maximum.
__device__ __forceinline__ float db_ref(float vmax) {
    return __fmul_rn(10.0f, log10f(vmax < 1e-10f ? 1e-10f : vmax));
}
__device__ __forceinline__ float db_value(float v, float refdb) {
    const float d = __fsub_rn(db_ref(v), refdb);
    return d < -80.0f ? -80.0f : d;
}
// grid = (chunks per item, items): no per-element index arithmetic, 128-bit accesses when aligned
__global__ void __launch_bounds__(256) power_to_db_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                          const unsigned* __restrict__ itemmax, long long item_len) {
    const long long item = blockIdx.y;
    const float refdb = db_ref(__uint_as_float(itemmax[item]));
    const float* src = in + item * item_len;
    float* dst = out + item * item_len;
    const long long stride = (long long)gridDim.x * 256;
    const long long t = (long long)blockIdx.x * 256 + threadIdx.x;
    if ((item_len & 3) == 0 && ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0) {
        const float4* s4 = reinterpret_cast<const float4*>(src);
        float4* d4 = reinterpret_cast<float4*>(dst);
        for (long long i = t; i < (item_len >> 2); i += stride) {
            float4 v = s4[i];
            v.x = db_value(v.x, refdb); v.y = db_value(v.y, refdb);
            v.z = db_value(v.z, refdb); v.w = db_value(v.w, refdb);
            d4[i] = v;
        }
    } else {
        for (long long i = t; i < item_len; i += stride) dst[i] = db_value(src[i], refdb);
    }
}

__global__ void __launch_bounds__(256) item_max_kernel(const float* __restrict__ in, unsigned* itemmax,
                                                       long long item_len) {
    __shared__ float smx[8];
    const float* p = in + (long long)blockIdx.x * item_len;
    float mx = -INFINITY;
    for (long long i = threadIdx.x; i < item_len; i += 256) mx = fmaxf(mx, p[i]);
    mx = warp_max(mx);
    if ((threadIdx.x & 31) == 0) smx[threadIdx.x >> 5] = mx;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) mx = fmaxf(mx, smx[w]);
        itemmax[blockIdx.x] = __float_as_uint(mx);
    }
}

// ---- normalize_data on independent rows ------------------------------------------------------
__global__ void __launch_bounds__(256) normalize_rows_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                             int row_len) {
    __shared__ float smn[8], smx[8];
    __shared__ float s_mn, s_range;
    const float* p = in + (long long)blockIdx.x * row_len;
    float* q = out + (long long)blockIdx.x * row_len;
    float mn = INFINITY, mx = -INFINITY;
    for (int i = threadIdx.x; i < row_len; i += 256) { float v = p[i]; mn = fminf(mn, v); mx = fmaxf(mx, v); }
    mn = warp_min(mn); mx = warp_max(mx);
    if ((threadIdx.x & 31) == 0) { smn[threadIdx.x >> 5] = mn; smx[threadIdx.x >> 5] = mx; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) { mn = fminf(mn, smn[w]); mx = fmaxf(mx, smx[w]); }
        s_mn = mn; s_range = __fsub_rn(mx, mn);
    }
    __syncthreads();
    mn = s_mn; const float range = s_range;
    for (int i = threadIdx.x; i < row_len; i += 256) q[i] = normalize_sample(p[i], mn, range);
}

// ---- general-parameter path: the segments as dense rows -------------------------------------------
// For parameter sets the fused kernel does not cover (n_fft other than 4096, bands above bin 1023) every
// segment is written out zero-padded (np.pad, src/identify_tracks.py:165-168) and normalised in
// normalize_data's exact float32 op order (:202-209); the rows then go through the framed-spectrum kernel
// of stft4800.cu with its banded mel epilogue and through power_to_db_kernel (cac_api.cu).
__global__ void __launch_bounds__(256) segment_rows_kernel(const float* __restrict__ pcm,
                                                           const cac_segment* __restrict__ segs,
                                                           float* __restrict__ rows, int seg_len, int normalize) {
    __shared__ float smn[8], smx[8];
    __shared__ float s_mn, s_range;
    const long long seg = blockIdx.x;
    const cac_segment sd = segs[seg];
    const float* p = pcm + sd.pcm_offset;
    float* q = rows + seg * seg_len;
    float mn = 0.f, range = 1.f;
    if (normalize) {
        float lo = INFINITY, hi = -INFINITY;
        for (int i = threadIdx.x; i < sd.n_valid; i += 256) { const float v = p[i]; lo = fminf(lo, v); hi = fmaxf(hi, v); }
        lo = warp_min(lo); hi = warp_max(hi);
        if ((threadIdx.x & 31) == 0) { smn[threadIdx.x >> 5] = lo; smx[threadIdx.x >> 5] = hi; }
        __syncthreads();
        if (threadIdx.x == 0) {
            for (int w = 1; w < 8; ++w) { lo = fminf(lo, smn[w]); hi = fmaxf(hi, smx[w]); }
            if (sd.n_valid < seg_len) { lo = fminf(lo, 0.f); hi = fmaxf(hi, 0.f); }   // the np.pad zeros
            s_mn = lo; s_range = __fsub_rn(hi, lo);
        }
        __syncthreads();
        mn = s_mn; range = s_range;
    }
    for (int j = threadIdx.x; j < seg_len; j += 256) {
        const int r = j - sd.pad_left;
        const float v = (r >= 0 && r < sd.n_valid) ? p[r] : 0.f;
        q[j] = normalize ? normalize_sample(v, mn, range) : v;
    }
}

// item tables of the framed-spectrum kernel for n equal rows: [offset | length | frame prefix sum]
__global__ void item_table_kernel(long long* tab, long long n, int seg_len, int n_frames) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i < n) { tab[i] = i * seg_len; tab[n + i] = seg_len; tab[2 * n + i] = i * n_frames; }
    if (i == n) tab[3 * n] = n * n_frames;
}

// ---- custommel.mel_spec on a caller-supplied magnitude matrix [bins, frames] ----------------
__global__ void __launch_bounds__(256) mel_project_kernel(const float* __restrict__ mag, float* __restrict__ out,
                                                          int n_frames, int n_mels, int power,
                                                          const BandTable* __restrict__ bands,
                                                          const float* __restrict__ weights) {
    const int f = blockIdx.x * 256 + threadIdx.x;
    const int m = blockIdx.y;
    if (f >= n_frames) return;
    const int k0 = bands->start[m], len = bands->len[m];
    const float* w = weights + bands->off[m];
    float acc = 0.f;
    for (int t = 0; t < len; ++t) {
        float v = fabsf(mag[(size_t)(k0 + t) * n_frames + f]);
        if (power == 2) v = v * v;
        else if (power != 1) v = powf(v, (float)power);
        acc = fmaf(w[t], v, acc);
    }
    out[(size_t)m * n_frames + f] = acc;
}

}  // namespace cac

// host-side launchers (called from cac_api.cu)
namespace cac {

// items go on grid.y in slices of 65535; each item is covered by ~item_len / (256 * 4 * 4) CTAs
static void launch_db(const float* in, float* out, const unsigned* itemmax, long long n_items, long long item_len,
                      cudaStream_t st) {
    long long bx = (item_len + 256 * 16 - 1) / (256 * 16);
    if (bx < 1) bx = 1;
    if (bx > 1024) bx = 1024;
    for (long long i0 = 0; i0 < n_items; i0 += 65535) {
        const long long ny = n_items - i0 < 65535 ? n_items - i0 : 65535;
        power_to_db_kernel<<<dim3((unsigned)bx, (unsigned)ny), 256, 0, st>>>(in + i0 * item_len, out + i0 * item_len,
                                                                             itemmax + i0, item_len);
    }
}

size_t feat_smem_bytes(const FeatArgs& a) {
    return sizeof(float) * (size_t)feat_smem_layout(a.stage_half, a.n_taps, a.n_mels, a.frames_per_stage).total;
}

cudaError_t launch_features(const FeatArgs& a, int power, int db_scale, cudaStream_t st, cudaEvent_t ev0,
                            cudaEvent_t ev1) {
    seg_minmax_kernel<<<(unsigned)a.n_segs, 256, 0, st>>>(a.pcm, a.segs, a.segstat, a.segmax, a.seg_len);
    const size_t smem = feat_smem_bytes(a);
    // K2N = number of 64-bin blocks the untangle stage must produce (bins < 64*K2N)
    const int k2n = (a.kmin + a.nbins - 1) / 64 + 1;
    void (*kern)(const FeatArgs);
    if (k2n <= 15) kern = (power == 1) ? stft_mel_kernel<1, 15> : stft_mel_kernel<2, 15>;
    else           kern = (power == 1) ? stft_mel_kernel<1, 16> : stft_mel_kernel<2, 16>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    long long grid = a.n_segs * a.stages_per_seg;
    if (grid > sms) grid = sms;                               // persistent: one CTA per SM
    if (ev0) cudaEventRecord(ev0, st);
    kern<<<(unsigned)grid, kWarpsPerCta * 32, smem, st>>>(a);
    if (ev1) cudaEventRecord(ev1, st);
    if (db_scale) {
        const long long item = (long long)a.n_mels * a.n_frames;
        launch_db(a.out, a.out, a.segmax, a.n_segs, item, st);
    }
    return cudaGetLastError();
}

cudaError_t launch_segment_rows(const float* pcm, const cac_segment* segs, long long n, float* rows, int seg_len,
                                int n_frames, int normalize, long long* tab, cudaStream_t st) {
    segment_rows_kernel<<<(unsigned)n, 256, 0, st>>>(pcm, segs, rows, seg_len, normalize);
    item_table_kernel<<<(unsigned)((n + 1 + 255) / 256), 256, 0, st>>>(tab, n, seg_len, n_frames);
    return cudaGetLastError();
}

cudaError_t launch_normalize(const float* in, float* out, long long rows, int row_len, cudaStream_t st) {
    normalize_rows_kernel<<<(unsigned)rows, 256, 0, st>>>(in, out, row_len);
    return cudaGetLastError();
}

cudaError_t launch_mel_project(const float* mag, float* out, int n_frames, int n_mels, int power,
                               const BandTable* d_bands, const float* d_weights, cudaStream_t st) {
    dim3 grid((n_frames + 255) / 256, n_mels);
    mel_project_kernel<<<grid, 256, 0, st>>>(mag, out, n_frames, n_mels, power, d_bands, d_weights);
    return cudaGetLastError();
}

cudaError_t launch_power_to_db(const float* in, float* out, long long n_items, long long item_len, unsigned* work,
                               cudaStream_t st) {
    item_max_kernel<<<(unsigned)n_items, 256, 0, st>>>(in, work, item_len);
    launch_db(in, out, work, n_items, item_len, st);
    return cudaGetLastError();
}

}  // namespace cac

// ---- cac_power_to_db_static: power_to_db without any allocation ------------------------------------------------------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The per-item maxima of
// cac_power_to_db_run's caller-owned work buffer live in a file-scope device array instead: 256 KB that every GPU
// which loads the library holds until exit.  ALL calls share the array: calls that overlap on the GPU (two streams,
// two threads) race on it.  item_max_kernel overwrites slots [0, n_items) before power_to_db_kernel reads them, so
// nothing has to start at zero and the previous call's contents do not matter.
#include <mutex>

#include "api_util.h"

namespace cac {
constexpr int kDbStaticDevices = 64;
static __device__ __align__(256) unsigned g_db_static_max[CAC_POWER_TO_DB_STATIC_MAX_ITEMS];
static unsigned* g_db_static_ptr[kDbStaticDevices] = {nullptr};   // device address of the array, per device
static std::mutex g_db_static_mu;

// In host code `g_db_static_max` is only the host-side stand-in of the symbol; its device address is looked up once
// per device, with that device current.
static int db_static_scratch(unsigned** out) {
    int dev = -1;
    CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kDbStaticDevices) return fail(CAC_ERR_INVALID, "cac_power_to_db_static: device index out of range");
    std::lock_guard<std::mutex> lock(g_db_static_mu);
    if (!g_db_static_ptr[dev]) {
        unsigned* p = nullptr;
        CUDA_TRY(cudaGetSymbolAddress((void**)&p, g_db_static_max));
        g_db_static_ptr[dev] = p;
    }
    *out = g_db_static_ptr[dev];
    return CAC_OK;
}
}  // namespace cac

extern "C" int cac_power_to_db_static(const float* d_in, float* d_out, int64_t n_items, int64_t item_len, void* stream) {
    using namespace cac;
    if (!d_in || !d_out || n_items < 0 || item_len < 1) return fail(CAC_ERR_INVALID, "bad argument");
    if (n_items > CAC_POWER_TO_DB_STATIC_MAX_ITEMS)
        return fail(CAC_ERR_INVALID, "cac_power_to_db_static: n_items = " + std::to_string((long long)n_items) +
                                         " exceeds the static scratch of " +
                                         std::to_string(CAC_POWER_TO_DB_STATIC_MAX_ITEMS) +
                                         " items; use cac_power_to_db_run with a caller-owned work buffer");
    if (n_items == 0) return CAC_OK;
    unsigned* scratch = nullptr;
    const int rc = db_static_scratch(&scratch);
    if (rc) return rc;
    CUDA_TRY(launch_power_to_db(d_in, d_out, n_items, item_len, scratch, (cudaStream_t)stream));
    return CAC_OK;
}

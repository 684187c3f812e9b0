// Framed band DFT as a GEMM on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM) for the
// identify_morepork spectrogram (src/identify_morepork.py:19-27): n_fft = sr // 10, hop = n_fft / 2, only the 60 bins
// of 600-1200 Hz are kept.
//
// hop = n_fft / 2 means frame f is chunk f-1 followed by chunk f of the recording (chunks of hop samples, librosa's
// centre padding = an all-zero chunk -1), so with the partial transform of a chunk
//     P_c[k] = sum_{m < hop} x[c hop + m] exp(-2 pi i k m / n_fft)
// the un-windowed DFT of frame f is  X_f[k] = P_{f-1}[k] + (-1)^k P_f[k],  and the periodic Hann window is the
// three-tap filter  X^w[k] = X[k] / 2 - (X[k-1] + X[k+1]) / 4  in the frequency domain.  Every sample is therefore
// multiplied with the transform matrix ONCE (not twice, as in the overlapping frames), and P for ALL chunks is one
// GEMM  P[chunks x 128] = A[chunks x hop] . B[hop x 128]  whose A operand is the PCM buffer itself (row stride hop)
// and whose B holds cos / -sin of the 62 bins kmin-1 .. kmin+60 (124 columns + 4 of padding).
//
// Precision: kind::tf32 multiplies 11-bit mantissas, so both operands are split, x = hi + lo with hi = x with the low
// 13 mantissa bits cleared (exactly representable in TF32, whatever the rounding of the hardware conversion) and
// lo = x - hi (exact in FP32); three products hi.hi + lo.hi + hi.lo accumulate in FP32 in TMEM, the dropped lo.lo term
// and the TF32 rounding of lo are ~2^-21 relative.  The tensor core adds into its FP32 accumulator with truncation:
// over the 900 accumulations of a whole chunk (hop 2400) that is a systematic -1.7e-5 relative shrink (measured,
// profiles/r3_experiments.md).  So the accumulator is restarted every kBgSegKb k-blocks (18 accumulations) and the
// segments are summed in FP32 registers with round-to-nearest by the epilogue warps, which own a whole 128-column row
// each (168 registers through setmaxnreg, taken from the producer / MMA warpgroups).
//
// One persistent CTA per SM (640 threads), warp-specialised around four 48 KB stages [A hi | A lo | B hi | B lo] of one
// 16-sample k-block each; every hand-over is an mbarrier.  The kernel is templated on how A arrives:
//   TMA form (equally long recordings of a whole number of hops, back to back: the buffer IS a [chunks x hop] matrix)
//     warp  8     one lane: per k-block ONE cp.async.bulk.tensor.2d box (256 chunk rows x 16 samples, 64-byte swizzle, rows
//                 past the end zero-filled) into A hi, and the bulk copy (cp.async.bulk) of the pre-swizzled B k-block
//     warps 0-7   converters, 4 chunks of every k-block per thread: A lo from the landed samples (the raw sample stays
//                 as the hi operand: the MMA ignores the low 13 mantissa bits), fence.proxy.async, arrive
//   cp.async form (any batch whose recordings start on 16-byte boundaries, ragged lengths included)
//     warps 0-7   producers: 16-byte cp.async of the raw samples through a per-tile row table, three k-blocks ahead,
//                 then hi / lo split in place (own chunks only, no cross-thread hand-over)
//     warp  8     one lane: bulk copy of the B k-block
//   both forms
//     warp  9     one lane: 6 tcgen05.mma per k-block (M = 128 columns of P, N = 256 chunk rows, K = 8: three products x
//                 two k-steps), tcgen05.commit to the stage's empty barrier / the accumulator's full barrier
//     warps 12-19 epilogue: tcgen05.ld of every finished segment accumulator (128 x 256, double-buffered: all 512 TMEM
//                 columns), running sums in registers, P rows to global memory after the tile's last segment
// band_frames_kernel / band_frames_uniform_kernel then form the frames (sum of two P rows, Hann taps, magnitude).
#include <cuda.h>
#include <cstdint>
#include <cstring>
#include <vector>

#include "cac_common.cuh"
#include "spec.cuh"

#ifndef CAC_BG_TRACE
#define CAC_BG_TRACE 0
#endif
#ifndef CAC_BG_SKIP_HI
#define CAC_BG_SKIP_HI 0      // 1: leave the raw sample as the hi operand (correct only if the MMA truncates FP32 -> TF32)
#endif

namespace cac {

#if CAC_BG_TRACE      // experiments only: timestamps of the first k-blocks of CTA 0 (tools/dbg_band_gemm.py trace)
__device__ long long g_bg_trace[8][256];
#define BG_TRACE(row, idx) do { if (blockIdx.x == 0 && (idx) < 256) g_bg_trace[row][idx] = clock64(); } while (0)
#else
#define BG_TRACE(row, idx) do { } while (0)
#endif

constexpr int kBgM = 256;                 // chunk rows per tile
constexpr int kBgN = 128;                 // columns of P: (re, im) of 62 bins + 4 unused
constexpr int kBgK = 16;                  // floats per k-block = one 64-byte swizzle row
constexpr int kBgStages = 4;
constexpr int kBgProducerWarps = 8;
constexpr int kBgEpiWarp0 = 12, kBgEpiWarps = 8;                      // warps 8-11: B loader, MMA issuer, two spare (warpgroup-aligned roles)
constexpr int kBgThreads = (kBgEpiWarp0 + kBgEpiWarps) * 32;          // 640
constexpr int kBgSegKb = 3;               // k-blocks per accumulation segment (18 MMA accumulations; see "Precision")
// setmaxnreg: the kernel starts with 96 registers per thread (640 threads); the producer warpgroups and the MMA warpgroup
// release what the two epilogue warpgroups take (only RELEASED registers can be taken): 256 (96 - 56) + 128 (96 - 32) =
// 18,432 = 256 (168 - 96)
constexpr int kBgRegsProd = 56, kBgRegsLow = 32, kBgRegsEpi = 168;
static_assert(256 * (96 - kBgRegsProd) + 128 * (96 - kBgRegsLow) >= 256 * (kBgRegsEpi - 96), "register hand-over does not add up");
constexpr int kBgABytes = kBgM * kBgK * 4;                            // 16 KB per split
constexpr int kBgBBytes = kBgN * kBgK * 4;                            // 8 KB per split
constexpr int kBgStageBytes = 2 * kBgABytes + 2 * kBgBBytes;          // A_hi | A_lo | B_hi | B_lo = 48 KB
constexpr int kBgTmemCols = 512;                                      // 2 accumulator stages x 2 row halves x 128

struct BgSmem {                                                       // after the stages
    const float* row_ptr[2][kBgM];
    int row_len[2][kBgM];
    unsigned long long landed[kBgStages], full[kBgStages], empty[kBgStages], acc_full[2], acc_empty[2];
    unsigned tmem_base;
};
constexpr size_t kBgSmemBytes = (size_t)kBgStages * kBgStageBytes + sizeof(BgSmem);

// ---- PTX wrappers ----------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(void* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(void* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(void* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(void* bar, unsigned parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// the whole warp waits, one lane polls (a poll is a shared-memory access per lane: 32 polling lanes cost 32 wavefronts)
__device__ __forceinline__ void mbar_wait_warp(void* bar, unsigned parity, int lane) {
    if (lane == 0) mbar_wait(bar, parity);
    __syncwarp();
}
// for the roles that wait long (epilogue, B loader): polls spaced out, the shared-memory pipe is the kernel's bottleneck
__device__ __forceinline__ bool mbar_try(void* bar, unsigned parity) {
    unsigned ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait_relaxed(void* bar, unsigned parity, unsigned ns) {
    while (!mbar_try(bar, parity)) __nanosleep(ns);
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, void* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// one box of a 2-D tensor map (cuTensorMapEncodeTiled) -> shared memory, completion counted in bytes on the mbarrier
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, void* bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                     smem_u32(dst)),
                 "l"(reinterpret_cast<unsigned long long>(map)), "r"(c0), "r"(c1), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void cp_async16_zfill(void* dst, const void* src, int src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(src_bytes) : "memory");
}
template <int N>
__device__ __forceinline__ void cp_async_wait_group() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(void* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[smem] . B[smem]^T, TF32 inputs, FP32 accumulate
__device__ __forceinline__ void tc_mma_tf32(unsigned d_tmem, unsigned long long a_desc, unsigned long long b_desc,
                                            unsigned idesc, unsigned accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// K-major operand tile with 64-byte rows, 64-byte swizzle (cute::UMMA::SmemDescriptor: start >> 4 | LBO 1 | SBO = 8 rows
// = 512 B | version 1 | layout SWIZZLE_64B = 4)
__device__ __forceinline__ unsigned long long umma_desc_sw64(unsigned smem_addr) {
    return (unsigned long long)((smem_addr & 0x3FFFFu) >> 4) | (1ull << 16) | ((unsigned long long)(512 >> 4) << 32) |
           (1ull << 46) | (4ull << 61);
}
// cute::UMMA::InstrDescriptor: D = F32 (1 << 4), A = B = TF32 (2 << 7, 2 << 10), both K-major, N >> 3 at 17, M >> 4 at 24.
// The MMA computes the TRANSPOSE of the tile, D^T[128 columns of P x 256 chunk rows] = B-tile . A-tile^T (M = 128, N = 256):
// one instruction reads 4 KB + 8 KB of operands per 128 cycles instead of two (M = 128 halves of the rows) reading
// 2 x (4 + 4) KB - the operand reads of TF32 MMAs are what fills the shared-memory pipe here.
constexpr unsigned kBgIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((unsigned)(kBgM >> 3) << 17) | ((unsigned)(kBgN >> 4) << 24);

// byte offset of the 16-byte chunk c (0..3) of row r in a 64-byte-swizzled tile (Swizzle<2,4,3>: chunk ^= (r >> 1) & 3)
__device__ __host__ __forceinline__ int sw64_offset(int r, int c) { return r * 64 + ((c ^ ((r >> 1) & 3)) << 4); }

// kTma = false: any batch whose recordings start on 16-byte boundaries; rows = chunks 0 .. frames - 1 of every recording
//                (the last one partial or empty), fetched by the producer warps with cp.async through a per-tile row table.
// kTma = true:  equally long recordings of a whole number of hops, back to back: the buffer IS a [chunks x hop] matrix and
//                ONE lane fetches a whole 256-row k-block per TMA instruction (a_map); the producer warps only convert.
template <bool kTma>
__global__ void __launch_bounds__(kBgThreads, 1)
band_gemm_kernel(const __grid_constant__ CUtensorMap a_map, const float* __restrict__ pcm, const long long* __restrict__ item_off, const long long* __restrict__ item_len,
                 const long long* __restrict__ frame_base, int n_items, long long total_rows, int hop,
                 const float* __restrict__ bsw, float* __restrict__ P) {
    // no static shared memory in this kernel: the dynamic window starts at its declared alignment.  Plain indexing of the
    // array keeps every access in the shared state space (LDS / STS); pointer arithmetic through an integer made them
    // generic LD / ST, whose latency (4 dependent ones per k-block) was the critical path of the first version.
    extern __shared__ __align__(1024) unsigned char stages[];
    BgSmem& S = *reinterpret_cast<BgSmem*>(stages + (size_t)kBgStages * kBgStageBytes);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if ((smem_u32(stages) & 1023u) != 0u) __trap();      // the swizzle is a function of the address bits
    const int n_kb = hop / kBgK;
    const long long n_tiles = (total_rows + kBgM - 1) / kBgM;

    if (tid == 0) {
        for (int s = 0; s < kBgStages; ++s) {
            mbar_init(&S.landed[s], 1);
            mbar_init(&S.full[s], kBgProducerWarps + 1);
            mbar_init(&S.empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) { mbar_init(&S.acc_full[a], 1); mbar_init(&S.acc_empty[a], kBgEpiWarps); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 9) {                          // the MMA warp owns the tensor memory
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&S.tmem_base)), "n"(kBgTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const unsigned tmem = S.tmem_base;
    const int n_seg = (n_kb + kBgSegKb - 1) / kBgSegKb;
    if (kTma && warp < kBgProducerWarps) {
        // ===== converters (TMA mode): the lo operand of the landed samples, lo = x - (x with the low 13 mantissa bits cleared).
        // The hi operand is the raw sample as it landed: the MMA reads the top 19 bits of an FP32 operand and ignores the
        // rest (measured: rewriting the truncated value changes no bit of the result).  All eight warps work on every
        // k-block (two teams alternating k-blocks, 8 chunks per thread, had the longer landed -> arrived chain and were
        // 5 % slower). =====
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kBgRegsProd));
        const int c = tid & 3, r0 = tid >> 2;                              // all eight warps on every k-block: rows r0 + 64 i
        const int soff0 = sw64_offset(r0, c);                              // + 4096 i: (row >> 1) & 3 does not depend on i
        long long n = 0;
        for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x)
            for (int kb = 0; kb < n_kb; ++kb, ++n) {
                const int st = (int)(n & (kBgStages - 1));
                mbar_wait_warp(&S.landed[st], (unsigned)((n / kBgStages) & 1), lane);
                if (tid == 0) BG_TRACE(0, (int)n);
                const unsigned char* a_hi = stages + (size_t)st * kBgStageBytes + soff0;
                unsigned char* a_lo = stages + (size_t)st * kBgStageBytes + kBgABytes + soff0;
                float4 v[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) v[i] = *reinterpret_cast<const float4*>(a_hi + 4096 * i);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    float4 l;
                    l.x = v[i].x - __uint_as_float(__float_as_uint(v[i].x) & 0xffffe000u);
                    l.y = v[i].y - __uint_as_float(__float_as_uint(v[i].y) & 0xffffe000u);
                    l.z = v[i].z - __uint_as_float(__float_as_uint(v[i].z) & 0xffffe000u);
                    l.w = v[i].w - __uint_as_float(__float_as_uint(v[i].w) & 0xffffe000u);
                    *reinterpret_cast<float4*>(a_lo + 4096 * i) = l;
                }
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) mbar_arrive(&S.full[st]);
                if (tid == 0) BG_TRACE(1, (int)n);
            }
    } else if (warp < kBgProducerWarps) {
        // ===== A producers (cp.async mode): thread owns chunk c of rows r0 + 64 i =====
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kBgRegsProd));
        const int c = tid & 3, r0 = tid >> 2;
        int soff[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) soff[i] = sw64_offset(r0 + 64 * i, c);
        int st_i = 0, ph_i = 1, st_d = 0;     // ring positions: next stage to fill (and the parity its empty barrier must have passed), next to split
        int tpar = 0;
        for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, tpar ^= 1) {
            {   // row table of the tile: source pointer and valid length of every chunk row
                const long long g = tile * kBgM + tid;
                const float* ptr = pcm;
                int len = 0;
                if (g < total_rows) {
                    int lo = 0, hi = n_items - 1;
                    while (lo < hi) {
                        const int mid = (lo + hi + 1) >> 1;
                        if (frame_base[mid] <= g) lo = mid; else hi = mid - 1;
                    }
                    const long long off = (g - frame_base[lo]) * hop;
                    const long long rem = item_len[lo] - off;
                    len = (int)(rem < 0 ? 0 : (rem > hop ? hop : rem));
                    ptr = pcm + item_off[lo] + off;
                }
                S.row_ptr[tpar][tid] = ptr;
                S.row_len[tpar][tid] = len;
            }
            asm volatile("bar.sync 1, 256;" ::: "memory");
            const float* rp[4];
            int rl[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) { rp[i] = S.row_ptr[tpar][r0 + 64 * i] + 4 * c; rl[i] = S.row_len[tpar][r0 + 64 * i] - 4 * c; }
            auto issue = [&](int kb) {
                mbar_wait_warp(&S.empty[st_i], (unsigned)ph_i, lane);
                unsigned char* a_hi = stages + (size_t)st_i * kBgStageBytes;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int rem = rl[i] - kb * kBgK;                     // valid floats from this chunk's first one on
                    const int nb = rem >= 4 ? 16 : (rem > 0 ? 4 * rem : 0);
                    cp_async16_zfill(a_hi + soff[i], nb > 0 ? rp[i] + kb * kBgK : pcm, nb);
                }
                asm volatile("cp.async.commit_group;" ::: "memory");
                if (++st_i == kBgStages) { st_i = 0; ph_i ^= 1; }
            };
            // the copies run kBgStages - 1 k-blocks ahead; k-block kb is split BEFORE the copy of kb + 3 is issued (that
            // one waits for the MMAs of kb - 1 to release the stage), so the split of kb overlaps the MMAs of kb - 1
            constexpr int kPre = kBgStages - 1;
            for (int kb = 0; kb < kPre; ++kb) issue(kb);                   // n_kb >= kBgStages (checked by the launcher)
            for (int kb = 0; kb < n_kb; ++kb) {
                if (kb + kPre <= n_kb) cp_async_wait_group<kPre - 1>(); else cp_async_wait_group<0>();
                if (tid == 0) BG_TRACE(0, kb);
                const int st = st_d;
                unsigned char* a_hi = stages + (size_t)st * kBgStageBytes;
                unsigned char* a_lo = a_hi + kBgABytes;
                float4 v[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) v[i] = *reinterpret_cast<const float4*>(a_hi + soff[i]);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    float4 h;
                    h.x = __uint_as_float(__float_as_uint(v[i].x) & 0xffffe000u);
                    h.y = __uint_as_float(__float_as_uint(v[i].y) & 0xffffe000u);
                    h.z = __uint_as_float(__float_as_uint(v[i].z) & 0xffffe000u);
                    h.w = __uint_as_float(__float_as_uint(v[i].w) & 0xffffe000u);
#if !CAC_BG_SKIP_HI
                    *reinterpret_cast<float4*>(a_hi + soff[i]) = h;
#endif
                    *reinterpret_cast<float4*>(a_lo + soff[i]) = make_float4(v[i].x - h.x, v[i].y - h.y, v[i].z - h.z, v[i].w - h.w);
                }
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) mbar_arrive(&S.full[st]);
                if (tid == 0) BG_TRACE(1, kb);
                if (++st_d == kBgStages) st_d = 0;
                if (kb + kPre < n_kb) issue(kb + kPre);
                if (tid == 0) BG_TRACE(2, kb);
            }
        }
    } else if (warp < kBgEpiWarp0) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kBgRegsLow));
        if (warp == 8 && lane == 0) {
            // ===== loader: one bulk copy of the k-block's pre-swizzled B tile [hi | lo] per stage (and the A box in TMA mode) =====
            {
            long long issued = 0;
            for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x)
                for (int kb = 0; kb < n_kb; ++kb, ++issued) {
                    const int st = (int)(issued % kBgStages);
                    mbar_wait_relaxed(&S.empty[st], (unsigned)(((issued / kBgStages) & 1) ^ 1), 32);
                    BG_TRACE(5, (int)issued);
                    if (kTma) {                // the whole A k-block of the tile: 256 chunk rows x 16 samples, 64-byte swizzle,
                        mbar_arrive_expect_tx(&S.landed[st], kBgABytes);     // rows past the end of the buffer arrive as zeros
                        tma_load_2d(stages + (size_t)st * kBgStageBytes, &a_map, kb * kBgK, (int)(tile * kBgM), &S.landed[st]);
                    }
                    mbar_arrive_expect_tx(&S.full[st], 2 * kBgBBytes);
                    bulk_g2s(stages + (size_t)st * kBgStageBytes + 2 * kBgABytes, bsw + (size_t)kb * (2 * kBgBBytes / 4),
                             2 * kBgBBytes, &S.full[st]);
                }
        }
        } else if (warp == 9 && lane == 0) {
            // ===== MMA issuer =====
            {
            long long used = 0, u = 0;              // k-blocks / accumulation segments so far
            for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                for (int kb0 = 0; kb0 < n_kb; kb0 += kBgSegKb, ++u) {
                    const int acc = (int)(u & 1);
                    mbar_wait(&S.acc_empty[acc], (unsigned)(((u >> 1) & 1) ^ 1));
                    tc_fence_after();
                    const int kb1 = kb0 + kBgSegKb < n_kb ? kb0 + kBgSegKb : n_kb;
                    for (int kb = kb0; kb < kb1; ++kb, ++used) {
                        const int st = (int)(used % kBgStages);
                        mbar_wait(&S.full[st], (unsigned)((used / kBgStages) & 1));
                        BG_TRACE(3, (int)used);
                        tc_fence_after();
                        const unsigned base = smem_u32(stages + (size_t)st * kBgStageBytes);
                        const unsigned long long a_hi = umma_desc_sw64(base), a_lo = umma_desc_sw64(base + kBgABytes);
                        const unsigned long long b_hi = umma_desc_sw64(base + 2 * kBgABytes);
                        const unsigned long long b_lo = umma_desc_sw64(base + 2 * kBgABytes + kBgBBytes);
                        const unsigned d = tmem + (unsigned)(acc * 256);
#pragma unroll
                        for (int ks = 0; ks < 2; ++ks) {                   // 8 floats = 32 bytes per k-step
                            const unsigned long long o = (unsigned long long)(ks * 2);
                            tc_mma_tf32(d, b_hi + o, a_hi + o, kBgIdesc, (unsigned)((kb != kb0) | ks));
                            tc_mma_tf32(d, b_hi + o, a_lo + o, kBgIdesc, 1u);
                            tc_mma_tf32(d, b_lo + o, a_hi + o, kBgIdesc, 1u);
                        }
                        tc_commit(&S.empty[st]);                           // the stage is free once these MMAs have read it
                        BG_TRACE(4, (int)used);
                    }
                    tc_commit(&S.acc_full[acc]);
                }
            }
        }
        }
    } else {
        // ===== epilogue: warp (q, h) owns TMEM lanes [32 q, 32 q + 32) = columns 32 q + lane of P, for the chunk rows
        // [128 h, 128 h + 128) of the tile (TMEM columns) =====
        asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kBgRegsEpi));
        const int q = warp & 3, h = (warp - kBgEpiWarp0) >> 2;
        long long u = 0;
        for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            float sum[128];
#pragma unroll
            for (int j = 0; j < 128; ++j) sum[j] = 0.f;
#pragma unroll 1
            for (int sgm = 0; sgm < n_seg; ++sgm, ++u) {
                const int acc = (int)(u & 1);
                if (lane == 0) mbar_wait_relaxed(&S.acc_full[acc], (unsigned)((u >> 1) & 1), 200);
                __syncwarp();
                tc_fence_after();
#pragma unroll
                for (int cc = 0; cc < 128 / 16; ++cc) {
                    unsigned v[16];
                    const unsigned taddr = tmem + ((unsigned)(q * 32) << 16) + (unsigned)(acc * 256 + h * 128 + cc * 16);
                    asm volatile(
                        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
                        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                        : "r"(taddr));
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                    for (int j = 0; j < 16; ++j) sum[cc * 16 + j] += __uint_as_float(v[j]);
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&S.acc_empty[acc]);
            }
            const long long g0 = tile * kBgM + h * 128;                    // a warp writes 128 consecutive bytes of one P row
            float* out = P + (size_t)g0 * kBgN + q * 32 + lane;
#pragma unroll
            for (int j = 0; j < 128; ++j)
                if (g0 + j < total_rows) out[(size_t)j * kBgN] = sum[j];
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 9) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(kBgTmemCols) : "memory");
    }
}

// Frame g of an item = P rows g - 1 (none for the item's first frame) and g: X[k] = P_prev[k] + (-1)^k P_cur[k] for the 62
// bins kmin - 1 .. kmin + 60, Hann in the frequency domain, magnitude (power 1) or its square (power 2).
// Block = 64 bins x 4 frames, thread (b, y) = bin b of frame 4 blockIdx.x + y: no division to find the frame (the first
// version numbered the cells flat and divided a 64-bit index by nbins in every thread - more instructions than the work);
// out[g * nbins + b] like spec4800_kernel's frame-major raw output.
constexpr int kBfFrames = 4;

// prev / cur: the two P rows of the frame (nullptr = no such chunk), already offset to column 2 b
__device__ __forceinline__ float band_frame_cell(const float* prev, const float* cur, int bin0, int power) {
    float xr[3], xi[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        const float s = ((bin0 + j) & 1) ? -1.f : 1.f;
        float2 p = make_float2(0.f, 0.f), c = make_float2(0.f, 0.f);
        if (prev) p = __ldg(reinterpret_cast<const float2*>(prev) + j);
        if (cur) c = __ldg(reinterpret_cast<const float2*>(cur) + j);
        xr[j] = fmaf(s, c.x, p.x);
        xi[j] = fmaf(s, c.y, p.y);
    }
    const float wr = 0.5f * xr[1] - 0.25f * (xr[0] + xr[2]);
    const float wi = 0.5f * xi[1] - 0.25f * (xi[0] + xi[2]);
    float mag = fmaf(wr, wr, wi * wi);
    if (power == 1) mag = sqrtf(mag);
    return mag;
}

__device__ __forceinline__ void band_frames_max(float mag, float* smax, float* gmax) {   // all threads of the block call
    const int t = threadIdx.y * 64 + threadIdx.x;
    float m = warp_max(mag);
    if ((t & 31) == 0) smax[t >> 5] = m;
    __syncthreads();
    if (t == 0) {
        for (int w = 1; w < 2 * kBfFrames; ++w) m = fmaxf(m, smax[w]);
        atomicMax(reinterpret_cast<unsigned*>(gmax), __float_as_uint(m));
    }
}

__global__ void __launch_bounds__(64 * kBfFrames) band_frames_kernel(const float* __restrict__ P, const long long* __restrict__ frame_base,
                                                                     int n_items, long long total_rows, int kmin, int nbins, int power,
                                                                     float* __restrict__ out, float* gmax) {
    __shared__ float smax[2 * kBfFrames];
    __shared__ int s_lo;
    const long long g = (long long)blockIdx.x * kBfFrames + threadIdx.y;
    const int b = threadIdx.x;
    if (threadIdx.x == 0 && threadIdx.y == 0) {   // the item of the block's first frame: one search per block, the threads walk on
        const long long g0 = (long long)blockIdx.x * kBfFrames;
        int lo = 0, hi = n_items - 1;
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (frame_base[mid] <= g0) lo = mid; else hi = mid - 1;
        }
        s_lo = lo;
    }
    __syncthreads();
    float mag = 0.f;
    if (g < total_rows && b < nbins) {
        int lo = s_lo;
        while (frame_base[lo + 1] <= g) ++lo;                          // g < total_rows = frame_base[n_items]
        const float* cur = P + (size_t)g * kBgN + 2 * b;               // columns 2 j, 2 j + 1 = bin kmin - 1 + j; j = b, b + 1, b + 2
        mag = band_frame_cell(frame_base[lo] == g ? nullptr : cur - kBgN, cur, kmin - 1 + b, power);
        out[g * nbins + b] = mag;
    }
    if (gmax) band_frames_max(mag, smax, gmax);
}

// The same for the TMA form's row numbering: item i owns the chunk rows [i C, i C + C) of P and the frames [i (C + 1),
// (i + 1) (C + 1)); frame f = chunk f - 1 (none for f = 0) followed by chunk f (all zeros for f = C: centre padding).
__global__ void __launch_bounds__(64 * kBfFrames) band_frames_uniform_kernel(const float* __restrict__ P, int chunks_per_item,
                                                                             long long total_frames, int kmin, int nbins, int power,
                                                                             float* __restrict__ out, float* gmax) {
    __shared__ float smax[2 * kBfFrames];
    const long long g = (long long)blockIdx.x * kBfFrames + threadIdx.y;
    const int b = threadIdx.x;
    float mag = 0.f;
    if (g < total_frames && b < nbins) {
        long long item;
        int f;
        if (total_frames <= 0x7fffffffLL) {                            // the usual case: 32-bit division
            const unsigned gi = (unsigned)g, per = (unsigned)(chunks_per_item + 1);
            item = gi / per;
            f = (int)(gi - (unsigned)item * per);
        } else {
            item = g / (chunks_per_item + 1);
            f = (int)(g - item * (chunks_per_item + 1));
        }
        const float* cur = P + (size_t)(item * chunks_per_item + f) * kBgN + 2 * b;
        mag = band_frame_cell(f > 0 ? cur - kBgN : nullptr, f < chunks_per_item ? cur : nullptr, kmin - 1 + b, power);
        out[g * nbins + b] = mag;
    }
    if (gmax) band_frames_max(mag, smax, gmax);
}

// ---- host side -------------------------------------------------------------------------------------------------------
bool band_gemm_supported(int n_fft, int hop, int kmin, int nbins) {
    return n_fft == 2 * hop && hop % kBgK == 0 && hop / kBgK >= kBgStages && kmin >= 1 && nbins + 2 <= kBgN / 2 &&
           kmin + nbins <= n_fft / 2;
}
size_t band_gemm_table_floats(int hop) { return (size_t)(hop / kBgK) * 2 * (kBgBBytes / 4); }
size_t band_gemm_scratch_bytes(long long total_rows) {
    const long long tiles = (total_rows + kBgM - 1) / kBgM;
    return (size_t)(tiles * kBgM + 1) * kBgN * sizeof(float);        // + 1: the uniform frames kernel forms (never reads) an address one row on
}

// B, k-block by k-block, [hi | lo], each a [128 columns of P][16 samples] K-major tile in the 64-byte swizzle the MMA
// reads: column 2 j = cos(2 pi k m / n_fft), 2 j + 1 = -sin(.) for k = kmin - 1 + j, sample m = 16 kb + kk of the chunk
void band_gemm_host_table(int n_fft, int hop, int kmin, int nbins, float* out) {
    const long double tau = 6.283185307179586476925286766559005768L;
    const int n_kb = hop / kBgK, per = kBgBBytes / 4;
    for (int kb = 0; kb < n_kb; ++kb) {
        float* hi = out + (size_t)kb * 2 * per;
        float* lo = hi + per;
        for (int n = 0; n < kBgN; ++n)
            for (int kk = 0; kk < kBgK; ++kk) {
                float v = 0.f;
                if (n < 2 * (nbins + 2)) {
                    const long long k = kmin - 1 + (n >> 1), m = (long long)kb * kBgK + kk;
                    const long double ang = tau * (long double)((k * m) % n_fft) / (long double)n_fft;
                    v = (n & 1) ? (float)-sinl(ang) : (float)cosl(ang);
                }
                uint32_t bits;
                memcpy(&bits, &v, 4);
                bits &= 0xffffe000u;
                float h;
                memcpy(&h, &bits, 4);
                const int o = (sw64_offset(n, kk >> 2) >> 2) + (kk & 3);
                hi[o] = h;
                lo[o] = v - h;
            }
    }
}

// The PCM buffer as a [chunks x hop] row-major FP32 matrix, boxes of 256 rows x 16 samples, 64-byte swizzle.
// cuTensorMapEncodeTiled comes from the driver through the runtime's entry-point query (no link against libcuda).
static cudaError_t encode_a_map(const float* base, long long total_chunks, int hop, CUtensorMap* map) {
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeFn encode = nullptr;
    if (!encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
        if (e != cudaSuccess) return e;
        if (qres != cudaDriverEntryPointSuccess || !fn) return cudaErrorNotSupported;
        encode = reinterpret_cast<EncodeFn>(fn);
    }
    const cuuint64_t dims[2] = {(cuuint64_t)hop, (cuuint64_t)total_chunks};
    const cuuint64_t strides[1] = {(cuuint64_t)hop * sizeof(float)};
    const cuuint32_t box[2] = {(cuuint32_t)kBgK, (cuuint32_t)kBgM};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue;
}

template <bool kTma>
static cudaError_t set_smem_attr() {
    static bool done[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 64 && done[dev]) return cudaSuccess;
    cudaError_t e = cudaFuncSetAttribute(band_gemm_kernel<kTma>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBgSmemBytes);
    if (e == cudaSuccess && dev < 64) done[dev] = true;
    return e;
}

// uniform_chunks > 0: the recordings are uniform_chunks * hop samples each and lie back to back from a.pcm + item 0's offset
// (checked by the caller, which passes that address as `base`): the TMA form.  Otherwise the cp.async form.
cudaError_t launch_band_gemm(const SpecArgs& a, const float* base, int uniform_chunks, const float* d_bsw, float* d_P,
                             cudaStream_t st) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    alignas(64) CUtensorMap a_map;
    memset(&a_map, 0, sizeof(a_map));
    if (uniform_chunks > 0) {
        const long long rows = (long long)a.n_items * uniform_chunks, frames = (long long)a.n_items * (uniform_chunks + 1);
        cudaError_t e = set_smem_attr<true>();
        if (e == cudaSuccess) e = encode_a_map(base, rows, a.hop, &a_map);
        if (e != cudaSuccess) return e;
        const long long tiles = (rows + kBgM - 1) / kBgM;
        band_gemm_kernel<true><<<(unsigned)(tiles < sms ? tiles : sms), kBgThreads, kBgSmemBytes, st>>>(
            a_map, nullptr, nullptr, nullptr, nullptr, 0, rows, a.hop, d_bsw, d_P);
        band_frames_uniform_kernel<<<(unsigned)((frames + kBfFrames - 1) / kBfFrames), dim3(64, kBfFrames), 0, st>>>(
            d_P, uniform_chunks, frames, a.kmin, a.nbins, a.power, a.out, a.gmax);
        return cudaGetLastError();
    }
    const long long rows = a.n_frames_total;
    if (rows < 1) return cudaSuccess;
    cudaError_t e = set_smem_attr<false>();
    if (e != cudaSuccess) return e;
    const long long tiles = (rows + kBgM - 1) / kBgM;
    band_gemm_kernel<false><<<(unsigned)(tiles < sms ? tiles : sms), kBgThreads, kBgSmemBytes, st>>>(
        a_map, a.pcm, a.item_off, a.item_len, a.frame_base, a.n_items, rows, a.hop, d_bsw, d_P);
    band_frames_kernel<<<(unsigned)((rows + kBfFrames - 1) / kBfFrames), dim3(64, kBfFrames), 0, st>>>(
        d_P, a.frame_base, a.n_items, rows, a.kmin, a.nbins, a.power, a.out, a.gmax);
    return cudaGetLastError();
}

}  // namespace cac

#if CAC_BG_TRACE
extern "C" int cac_debug_band_gemm_trace(long long* out) {
    return (int)cudaMemcpyFromSymbol(out, cac::g_bg_trace, sizeof(long long) * 8 * 256);
}
#endif

// Shared declarations for the cacophony_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/cacophony_b200.h"

namespace cac {

constexpr int kWarp = 32;
constexpr int kMaxMels = 512;
constexpr int kMaxTaps = 12288;   // packed band weights of one plan (device copy owned by the plan)

// Banded filterbank (one contiguous band per mel row), uploaded per plan.
struct BandTable {
    int16_t start[kMaxMels];   // first FFT bin of the band
    int16_t len[kMaxMels];     // taps
    int32_t off[kMaxMels];     // offset into weights[]
};

struct FeatArgs {
    const float* pcm;
    const cac_segment* segs;
    float* out;            // [n_segs][n_mels][n_frames]
    float2* segstat;       // [n_segs] (min, max) of the padded segment
    unsigned* segmax;      // [n_segs] max mel value (float bits, values >= 0)
    long long n_segs;
    int seg_len;
    int hop;
    int n_frames;
    int n_mels;
    int kmin;              // lowest FFT bin any band touches
    int nbins;             // kmax - kmin + 1
    int stages_per_seg;    // ceil(n_frames / frames_per_stage)
    int frames_per_stage;  // multiple of 8: frames staged (and transformed) per round
    int n_taps;            // floats in wpad
    // Filterbank for the fused kernel: mel rows in groups of 32 (lane = row within the group), every
    // band starts at a multiple of 4 bins (zero weights in front) and is zero-padded to the group's
    // longest band (multiple of 4); stored as float4 per (tap quad, lane):
    // wpad[goff[g] + ((t/4)*32 + lane)*4 + t%4] -> conflict-free 128-bit reads without predicates.
    const float* wpad;
    const int* mstart;     // [n_groups * 32] int4 slot table (cac_api.cu): band start, tile offset of the row (-1 dead), weight offset, quads
    int n_groups;
    int pbuf_len;          // floats per warp power-spectrum buffer (covers the padded reads)
    int gmax[16];          // padded band length of group g
    int goff[16];          // offset of group g inside wpad
    int stage_half;        // floats reserved for each of the even/odd sample planes
    int normalize;
    // MagTransform fused into the flush (src/magtransform.py:17-19, src/magtransformv2.py:19-21): when mag_on the
    // value leaving the kernel is mel ** mag_p, mag_p = sigmoid(a) (only without db_scale; cac_plan_set_magtransform)
    int mag_on;
    float mag_p;
};

// floats reserved for each of the even / odd sample planes of a stage of `frames` consecutive 4096-sample frames
// (stft_mel.cu, sn_stft.cu)
__host__ __device__ inline int stage_half_floats(int frames, int hop) {
    const int stage_len = (frames - 1) * hop + 4096;
    return (((stage_len + 8 + 1) / 2 + 1 + 31) / 32) * 32 + 16;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// normalize_data (src/identify_tracks.py:202-209), same float32 op order:
// x - min; / max(x - min); + 1e-6; - 0.5; * 2
__device__ __forceinline__ float normalize_sample(float v, float mn, float range) {
    float x = __fsub_rn(v, mn);
    x = __fdiv_rn(x, range);
    x = __fadd_rn(x, 0.000001f);
    x = __fsub_rn(x, 0.5f);
    return __fmul_rn(x, 2.0f);
}

}  // namespace cac

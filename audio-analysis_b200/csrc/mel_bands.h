// Host-side preparation of a custommel.mel_f filterbank (src/custommel.py:19-56) for the kernels:
// every mel row is one contiguous band; rows are grouped by 32 (lane = row within the group), bands
// start at a multiple of 4 bins (zero weights in front) and are zero-padded to the group's longest
// band, stored tap-major as float4 per (tap quad, lane) -> conflict-free, predicate-free 128-bit reads.
#pragma once
#include <string>
#include <vector>
#include "cac_common.cuh"

namespace cac {

struct MelBands {
    BandTable bt;
    std::vector<float> weights;   // packed band weights (for the standalone projection kernel)
    int kmin = 0, kmax = 0;       // lowest / highest FFT bin any band touches
    std::vector<float> wpad;      // wpad[goff[g] + ((t/4)*32 + lane)*4 + t%4]
    std::vector<int> mstart;      // [n_mels] (first bin - kmin) rounded down to a multiple of 4
    int n_groups = 0, pbuf_len = 0, gmax[16] = {0}, goff[16] = {0};

    // returns "" or the reason the table cannot be built
    std::string build(const float* w, int n_mels, int bins) {
        if (n_mels < 1 || n_mels > kMaxMels) return "n_mels out of range";
        int lo = bins, hi = -1;
        for (int m = 0; m < n_mels; ++m) {
            int first = -1, last = -1;
            for (int k = 0; k < bins; ++k)
                if (w[(size_t)m * bins + k] != 0.0f) { if (first < 0) first = k; last = k; }
            if (first < 0) { first = 0; last = -1; }          // empty filter: contributes zero
            bt.start[m] = (int16_t)first;
            bt.len[m] = (int16_t)(last - first + 1);
            bt.off[m] = (int32_t)weights.size();
            for (int k = first; k <= last; ++k) weights.push_back(w[(size_t)m * bins + k]);
            if (last >= first) { lo = first < lo ? first : lo; hi = last > hi ? last : hi; }
        }
        if (hi < 0) { lo = 0; hi = 0; }
        for (int m = 0; m < n_mels; ++m)
            if (bt.len[m] == 0) bt.start[m] = (int16_t)lo;
        if ((int)weights.size() > kMaxTaps) return "filterbank has too many taps";
        kmin = lo;
        kmax = hi;
        n_groups = (n_mels + 31) / 32;
        mstart.assign(n_mels, 0);
        int need = kmax - kmin + 1;
        for (int g = 0; g < n_groups && g < 16; ++g) {
            int gm = 0;
            for (int m = g * 32; m < n_mels && m < g * 32 + 32; ++m) {
                int front = (bt.start[m] - kmin) & 3;
                gm = bt.len[m] + front > gm ? bt.len[m] + front : gm;
            }
            gm = (gm + 3) / 4 * 4;
            gmax[g] = gm;
            goff[g] = (int)wpad.size();
            wpad.resize(wpad.size() + (size_t)gm * 32, 0.0f);
            for (int m = g * 32; m < n_mels && m < g * 32 + 32; ++m) {
                const int s0 = bt.start[m] - kmin, front = s0 & 3, lane = m - g * 32;
                mstart[m] = s0 - front;
                for (int t = 0; t < bt.len[m]; ++t) {
                    const int tt = t + front;
                    wpad[goff[g] + ((size_t)(tt >> 2) * 32 + lane) * 4 + (tt & 3)] = weights[bt.off[m] + t];
                }
                if (mstart[m] + gm > need) need = mstart[m] + gm;
            }
        }
        pbuf_len = (need + 31) / 32 * 32;
        return "";
    }
};

}  // namespace cac

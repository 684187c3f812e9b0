// Host-side preparation of a custommel.mel_f filterbank (src/custommel.py:19-56) for the kernels:
// every mel row is one contiguous band; rows are dealt to (group, lane) slots, 32 per group, bands
// start at a multiple of 4 bins (zero weights in front) and are zero-padded to the group's longest
// band, stored tap-major as float4 per (tap quad, lane) -> conflict-free, predicate-free 128-bit reads
// of the weights, and (by the choice of slots) nearly conflict-free 128-bit reads of the spectrum.
#pragma once
#include <algorithm>
#include <string>
#include <vector>
#include "cac_common.cuh"

namespace cac {

struct MelBands {
    BandTable bt;
    std::vector<float> weights;   // packed band weights (for the standalone projection kernel)
    int kmin = 0, kmax = 0;       // lowest / highest FFT bin any band touches
    std::vector<float> wpad;      // wpad[goff[g] + ((t/4)*32 + lane)*4 + t%4]
    std::vector<int> mstart;      // [n_groups * 32] slot table: (first bin - base bin, rounded down to 4) | mel row << 16; row 0xFFFF = dead slot
    int n_groups = 0, pbuf_len = 0, gmax[16] = {0}, goff[16] = {0};

    // returns "" or the reason the table cannot be built
    // absolute = false: band starts (mstart) and pbuf_len are relative to bin kmin; true: relative to bin 0
    std::string build(const float* w, int n_mels, int bins, bool absolute = false) {
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
        const int kbase = absolute ? 0 : kmin;
        kmax = hi;
        n_groups = (n_mels + 31) / 32;
        // Row -> (group, lane) slots.  A 128-bit shared load serves 8 lanes per wavefront, so the 8 rows of a
        // quarter-warp should start in 8 different 16-byte columns (start / 4 mod 8): rows are taken longest
        // first (keeps the padding per group small) and dealt to quarter-warps by that residue, preferring
        // rows whose index mod 32 is still free in the group (conflict-free tile stores).
        const int n_slots = n_groups * 32;
        std::vector<int> s0(n_mels), plen(n_mels), order(n_mels), slot_row(n_slots, -1);
        for (int m = 0; m < n_mels; ++m) {
            s0[m] = bt.start[m] - kbase;
            plen[m] = bt.len[m] + (s0[m] & 3);
            order[m] = m;
        }
        std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return plen[x] > plen[y]; });
        std::vector<char> taken(n_mels, 0);
        for (int g = 0; g < n_groups; ++g) {
            bool used32[32] = {false};
            for (int q = 0; q < 4; ++q) {
                bool usedres[8] = {false};
                int filled = 0;
                for (int crit = 0; crit < 3 && filled < 8; ++crit)
                    for (int oi = 0; oi < n_mels && filled < 8; ++oi) {
                        const int m = order[oi];
                        if (taken[m]) continue;
                        const int res = ((s0[m] - (s0[m] & 3)) >> 2) & 7;
                        if ((crit == 0 && (usedres[res] || used32[m & 31])) || (crit == 1 && usedres[res])) continue;
                        taken[m] = 1;
                        usedres[res] = true;
                        used32[m & 31] = true;
                        slot_row[g * 32 + q * 8 + filled++] = m;
                    }
            }
        }
        mstart.assign(n_slots, (int)0xFFFF0000u);            // dead slot: row 0xFFFF, start 0
        int need = kmax - kbase + 1;
        for (int g = 0; g < n_groups && g < 16; ++g) {
            int gm = 0;
            for (int l = 0; l < 32; ++l) {
                const int m = slot_row[g * 32 + l];
                if (m >= 0 && plen[m] > gm) gm = plen[m];
            }
            gm = (gm + 3) / 4 * 4;
            gmax[g] = gm;
            goff[g] = (int)wpad.size();
            wpad.resize(wpad.size() + (size_t)gm * 32, 0.0f);
            for (int l = 0; l < 32; ++l) {
                const int m = slot_row[g * 32 + l];
                if (m < 0) continue;
                const int front = s0[m] & 3, ms = s0[m] - front;
                mstart[g * 32 + l] = ms | (m << 16);
                for (int t = 0; t < bt.len[m]; ++t) {
                    const int tt = t + front;
                    wpad[goff[g] + ((size_t)(tt >> 2) * 32 + l) * 4 + (tt & 3)] = weights[bt.off[m] + t];
                }
                if (ms + gm > need) need = ms + gm;
            }
        }
        pbuf_len = (need + 31) / 32 * 32;
        return "";
    }
};

}  // namespace cac

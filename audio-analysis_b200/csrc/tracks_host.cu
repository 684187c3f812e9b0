// Host-side merge sweeps of track finding (src/identify_tracks.py:725-792 merge_signals, repeated until nothing merges
// as get_tracks_from_signals does, :800-802).  No device code: the reference runs this scalar O(n^2) logic in Python on
// ~10-100 boxes per recording; with the GPU chain at 0.19 ms per recording the Python loop (0.2 ms per recording) had
// become the slowest stage of classify_prep_batch.  Same IEEE double operations in the same order as tracks.py
// (_should_merge), which stays as the readable mirror of the reference and as the checker of this function
// (tests/test_tracks_host.py).  A merged box only ever takes minima / maxima of the boxes it absorbed, so the result
// is returned as INDICES of the input signals that supply each field: the caller copies the original Python objects
// (ints stay ints, floats stay floats), and the mel values follow their frequencies (mel is monotone).
#include <algorithm>
#include <cstdint>
#include <vector>

#include "../../include/cacophony_b200.h"
#include "api_util.h"

namespace {

struct Sig {
    double s, e, ms, me;          // start, end, mel_freq_start, mel_freq_end
    double fs, fe;                // freq_start, freq_end (only min / max are taken)
    int self, i_s, i_e, i_fs, i_fe;
};

// tracks.py:_should_merge(s, u) - src/identify_tracks.py:745-785
inline bool should_merge(const Sig& s, const Sig& u) {
    const double sme = s.me, ume = u.me;
    if (!((ume < 1500 && sme < 1500) || (ume > 1500 && sme > 1500))) return false;
    const double ss = s.s, se = s.e, us = u.s, ue = u.e, sms = s.ms, ums = u.ms;
    const double t_overlap = (se - ss) + (ue - us) - (std::max(se, ue) - std::min(ss, us));
    const double need = (sms > 1000 && ums > 1000) ? 0.5 : 0.75;
    const double m_overlap = (sme - sms) + (ume - ums) - (std::max(sme, ume) - std::min(sms, ums));
    const double u_range = ume - ums;
    if (t_overlap > (ue - us) * 0.75 && m_overlap > -20) return true;
    if (t_overlap > 0 && m_overlap > u_range * need) return true;
    const double gap = ss > ue ? ss - ue : us - se;
    if (m_overlap > u_range * need && gap <= 2) {
        const double s_range = sme - sms;
        const double similar = ume > s_range ? s_range / u_range : u_range / s_range;
        return similar >= 0.75;
    }
    return false;
}

}  // namespace

extern "C" int cac_tracks_merge(const double* boxes, int32_t n, int32_t* out, int32_t* n_out) {
    if (n < 0 || (n > 0 && (!boxes || !out)) || !n_out) return cac::fail(CAC_ERR_INVALID, "null argument");
    std::vector<Sig> v((size_t)n);
    for (int i = 0; i < n; ++i) {
        const double* b = boxes + 6 * (size_t)i;
        v[i] = Sig{b[0], b[1], b[4], b[5], b[2], b[3], i, i, i, i, i};
    }
    std::vector<char> dead;
    std::vector<Sig> gone;
    bool merged = n > 0;
    while (merged) {
        // sorted(sorted(signals, key=mel_freq_end, reverse=True), key=start): both stable, ties keep the previous order
        std::stable_sort(v.begin(), v.end(), [](const Sig& a, const Sig& b) { return a.me > b.me; });
        std::stable_sort(v.begin(), v.end(), [](const Sig& a, const Sig& b) { return a.s < b.s; });
        dead.assign(v.size(), 0);
        merged = false;
        for (size_t i = 0; i < v.size(); ++i) {
            if (dead[i]) continue;
            Sig& s = v[i];
            for (size_t j = 0; j < v.size(); ++j) {
                if (j == i || dead[j]) continue;
                const Sig& u = v[j];
                if (!should_merge(s, u)) continue;
                // Signal.merge: min / max return their FIRST argument on ties
                if (u.s < s.s) { s.s = u.s; s.i_s = u.i_s; }
                if (u.e > s.e) { s.e = u.e; s.i_e = u.i_e; }
                if (u.fs < s.fs) { s.fs = u.fs; s.i_fs = u.i_fs; }
                if (u.fe > s.fe) { s.fe = u.fe; s.i_fe = u.i_fe; }
                if (u.ms < s.ms) s.ms = u.ms;
                if (u.me > s.me) s.me = u.me;
                dead[j] = 1;
                merged = true;
                break;
            }
        }
        if (merged) {
            size_t k = 0;
            for (size_t i = 0; i < v.size(); ++i) {
                if (dead[i]) gone.push_back(v[i]); else v[k++] = v[i];
            }
            v.resize(k);
        }
    }
    *n_out = (int32_t)v.size();
    v.insert(v.end(), gone.begin(), gone.end());            // the absorbed ones, as they were when they were absorbed
    for (size_t i = 0; i < v.size(); ++i) {
        int32_t* o = out + 5 * i;
        o[0] = v[i].self; o[1] = v[i].i_s; o[2] = v[i].i_e; o[3] = v[i].i_fs; o[4] = v[i].i_fe;
    }
    return CAC_OK;
}

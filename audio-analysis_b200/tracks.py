"""Host half of track finding: the `Signal` record and the interval-merge heuristics that turn signal_noise's
component boxes into tracks (src/identify_tracks.py:709-842, 915-1033).  O(100) rectangles per recording and pure
scalar logic, so it stays on the host; it is here so that the GPU stages on either side of it
(signal_noise before, load_samples after) form one drop-in chain.  Results are pinned against the reference's own
functions on seeded rectangles (tests/golden/tracks.npz, tests/test_tracks_host.py).
"""
import functools

import numpy as np


def segment_overlap(first, second):
    """Overlap of two (lo, hi) intervals, negative = size of the gap (src/identify_tracks.py:709-714)."""
    return (first[1] - first[0]) + (second[1] - second[0]) - (max(first[1], second[1]) - min(first[0], second[0]))


def mel_freq(f):
    """HTK mel (src/identify_tracks.py:717-718)."""
    return 2595.0 * np.log10(1.0 + f / 700.0)


@functools.lru_cache(maxsize=1 << 16)
def _mel_cached(f):
    """mel_freq as a Python float, memoised: the frequencies are FFT bin centres or small integers, a few thousand
    distinct values, and numpy's scalar log10 is the most expensive thing the merge heuristics do."""
    return float(mel_freq(f))


class Signal:
    """src/identify_tracks.py:915-1033: a time / frequency box plus the classifier's verdict on it.  Compared and
    removed by identity, like the reference's class (no __eq__)."""

    def __init__(self, start, end, freq_start, freq_end):
        self.start = start
        self.end = end
        self.freq_start = freq_start
        self.freq_end = freq_end
        self._refresh_mel()
        self.results = []
        self.master_tag = None
        self.master_model = None
        self.master_below_thresh = True
        self.track_id = None

    def _refresh_mel(self):
        # the same float64 values as the reference's numpy scalars, held as Python floats: the merge heuristics compare
        # them O(n^2) times per sweep and numpy scalar arithmetic is several times slower than float arithmetic
        self.mel_freq_start = _mel_cached(float(self.freq_start))
        self.mel_freq_end = _mel_cached(float(self.freq_end))

    # -- geometry ------------------------------------------------------------------------------------------
    @property
    def length(self):
        return self.end - self.start

    @property
    def freq_range(self):
        return self.freq_end - self.freq_start

    @property
    def mel_freq_range(self):
        return self.mel_freq_end - self.mel_freq_start

    def time_overlap(self, other):
        return segment_overlap((self.start, self.end), (other.start, other.end))

    def freq_overlap(self, other):
        return segment_overlap((self.freq_start, self.freq_end), (other.freq_start, other.freq_end))

    def mel_freq_overlap(self, other):
        return segment_overlap((self.mel_freq_start, self.mel_freq_end), (other.mel_freq_start, other.mel_freq_end))

    def copy(self):
        return Signal(self.start, self.end, self.freq_start, self.freq_end)

    def to_array(self, decimals=1):
        box = [self.start, self.end, self.freq_start, self.freq_end]
        return box if decimals is None else list(np.round(np.array(box), decimals))

    def merge(self, other):
        """Bounding box of both (:1004-1010)."""
        self.start, self.end = min(self.start, other.start), max(self.end, other.end)
        self.freq_start, self.freq_end = min(self.freq_start, other.freq_start), max(self.freq_end, other.freq_end)
        self._refresh_mel()

    def enlarge(self, scale, min_track_length):
        """Grow around the centre in time (at least min_track_length, clipped at 0) and frequency, the latter
        truncated to ints (:985-1002)."""
        grow_t = (max(self.length * scale, min_track_length) - self.length) / 2
        self.start = max(self.start - grow_t, 0)
        self.end = self.end + grow_t
        grow_f = (self.freq_range * scale - self.freq_range) / 2
        lo = self.freq_start - grow_f
        self.freq_end = int(self.freq_end + grow_f)
        self.freq_start = int(max(lo, 0))
        self._refresh_mel()

    # -- metadata ------------------------------------------------------------------------------------------
    def get_meta(self):
        """The per-track entry of analysis_result["species_identify"] (:1015-1033)."""
        meta = {"begin_s": self.start, "end_s": self.end, "freq_start": self.freq_start, "freq_end": self.freq_end}
        if self.master_tag is not None:
            meta["master_tag"] = {"below_thresh": self.master_below_thresh, "prediction": self.master_tag.get_meta(),
                                  "model": self.master_model}
        meta["model_results"] = [r.get_meta() for r in self.results]
        if self.track_id is not None:
            meta["track_id"] = self.track_id
        return meta

    def __str__(self):
        return f"Signal: {self.start}-{self.end} f: {self.freq_start}-{self.freq_end}"


def _should_merge(s, u):
    """The three merge rules of one (s, u) pair (src/identify_tracks.py:745-785).  segment_overlap is written out
    (same operations in the same order) because this runs O(n^2) times per sweep."""
    sme, ume = s.mel_freq_end, u.mel_freq_end
    if not ((ume < 1500 and sme < 1500) or (ume > 1500 and sme > 1500)):
        return False
    ss, se, us, ue = s.start, s.end, u.start, u.end
    sms, ums = s.mel_freq_start, u.mel_freq_start
    t_overlap = (se - ss) + (ue - us) - (max(se, ue) - min(ss, us))
    need = 0.5 if (sms > 1000 and ums > 1000) else 0.75
    m_overlap = (sme - sms) + (ume - ums) - (max(sme, ume) - min(sms, ums))
    u_range = ume - ums
    if t_overlap > (ue - us) * 0.75 and m_overlap > -20:
        return True
    if t_overlap > 0 and m_overlap > u_range * need:
        return True
    gap = ss - ue if ss > ue else us - se
    if m_overlap > u_range * need and gap <= 2:
        s_range = sme - sms
        similar = s_range / u_range if ume > s_range else u_range / s_range
        return similar >= 0.75
    return False


def merge_signals(signals):
    """One merging sweep (src/identify_tracks.py:725-792): signals ordered by start (ties: highest mel_freq_end first);
    each surviving signal absorbs the FIRST other survivor it may merge with.  Returns (signals, anything merged)."""
    ordered = sorted(sorted(signals, key=lambda s: s.mel_freq_end, reverse=True), key=lambda s: s.start)
    dead = set()                                   # ids: identity, like the reference's `in to_delete`
    for s in ordered:
        if id(s) in dead:
            continue
        for u in ordered:
            if u is s or id(u) in dead:
                continue
            if _should_merge(s, u):
                s.merge(u)
                dead.add(id(u))
                break
    survivors = [s for s in ordered if id(s) not in dead]
    return survivors, bool(dead)


def merge_until_stable(signals):
    """`while merged: signals, merged = merge_signals(signals)` (src/identify_tracks.py:800-802) through the library's
    host function cac_tracks_merge: the sweeps run in C on the boxes' float64 values (the same operations in the same
    order as _should_merge), and every surviving Signal then takes its fields from the ORIGINAL objects the result
    names - the same objects, types and values the Python loop above leaves behind (merge() only ever takes minima
    and maxima).  tests/test_tracks_host.py runs both on the golden rectangles and on random ones."""
    import ctypes
    from . import engine
    n = len(signals)
    if n == 0:
        return []
    orig = [(s.start, s.end, s.freq_start, s.freq_end, s.mel_freq_start, s.mel_freq_end) for s in signals]
    boxes = np.array(orig, dtype=np.float64)
    out = np.empty((n, 5), dtype=np.int32)
    n_out = ctypes.c_int32()
    engine.check(engine.lib.cac_tracks_merge(boxes.ctypes.data, n, out.ctypes.data, ctypes.byref(n_out)))
    survivors = []
    for j, (me, i_s, i_e, i_fs, i_fe) in enumerate(out.tolist()):
        s = signals[me]                                # rows beyond n_out: absorbed signals, as the loop above leaves them
        s.start, s.end = orig[i_s][0], orig[i_e][1]
        s.freq_start, s.mel_freq_start = orig[i_fs][2], orig[i_fs][4]
        s.freq_end, s.mel_freq_end = orig[i_fe][3], orig[i_fe][5]
        if j < n_out.value:
            survivors.append(s)
    return survivors


def get_tracks_from_signals(signals, end):
    """src/identify_tracks.py:795-842: merge until nothing merges, drop short signals, enlarge the rest by 1.4x (at
    least 0.7 s, clipped to [0, end]) absorbing signals that overlap more than 70 % of the shorter one, drop tracks
    narrower than 50 mel.  Quirk kept: the minimum length starts at 0.35 s but is overwritten by every overlap test
    with the shorter length of the pair just compared (:803, :826)."""
    signals = merge_until_stable(signals)
    dead = set()
    min_length = 0.35
    for s in signals:
        if id(s) in dead:
            continue
        if s.length < min_length:
            dead.add(id(s))
            continue
        s.enlarge(1.4, min_track_length=0.7)
        s.end = min(end, s.end)
        for other in signals:
            if other is s or id(other) in dead:
                continue
            min_length = min(s.length, other.length)
            if s.time_overlap(other) > 0.7 * min_length:
                s.merge(other)
                dead.add(id(other))
    return [s for s in signals if id(s) not in dead and not s.mel_freq_range < 50]

"""Generate tests/golden/tracks.json and tests/golden/analyse.json by executing the REFERENCE's own host code unchanged
(src/identify_tracks.py: Signal, merge_signals, get_tracks_from_signals, get_max_chirps; src/analyse.py:
calc_cacophony_index, get_chirps, filter_tracks and the result dictionary of species_identify) under the stand-in
modules of oracle/stubs, on seeded random rectangles.  Run in the build container only:
    python tests/golden/make_golden_host.py
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import reference_loader as rl  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def boxes(rng, n, length):
    """Rectangles shaped like signal_noise's output: 0.2-3 s long, multiples of the 11.7 Hz bin width in frequency."""
    out = []
    for _ in range(n):
        t0 = float(rng.uniform(0, length - 0.3))
        dur = float(rng.choice([rng.uniform(0.15, 0.5), rng.uniform(0.5, 3.0)]))
        f0 = float(int(rng.uniform(8, 600)) * 11.71875)
        f1 = f0 + float(int(rng.uniform(8, 300)) * 11.71875)
        out.append([t0, min(t0 + dur, length), f0, f1])
    out.sort(key=lambda b: b[0])
    return out


class Tag:                                   # stands for identify_tracks.Prediction: the code under test reads .what / get_meta()
    def __init__(self, what):
        self.what = what

    def get_meta(self):
        return {"label": self.what}


def main():
    it = rl.load("identify_tracks")
    an = rl.load("analyse")
    rng = np.random.default_rng(20250101)
    tracks_cases, analyse_cases = [], []
    labels = ["bird", "kiwi", "morepork", "human", "noise", "insect", None]
    for case in range(24):
        length = float(rng.choice([10.0, 20.5, 59.9, 60.0, 61.0, 61.9, 125.3]))
        n = int(rng.integers(0, 40))
        bx = boxes(rng, n, length)
        sigs = [it.Signal(*b) for b in bx]
        tracks = it.get_tracks_from_signals([s.copy() for s in sigs], length)
        tracks_cases.append({"length": length, "signals": bx,
                             "tracks": [[t.start, t.end, t.freq_start, t.freq_end] for t in tracks]})
        # verdicts for the result assembly
        tags = [labels[int(rng.integers(0, len(labels)))] for _ in tracks]
        for t, w in zip(tracks, tags):
            t.master_tag = Tag(w) if w is not None else None
        bird_labels = ["bird", "kiwi", "morepork"]
        ci, version = an.calc_cacophony_index(an.filter_tracks(tracks), length)
        chirp_signals = [s.copy() for s in sigs]
        chirps = an.get_chirps(tracks, bird_labels, chirp_signals)
        analyse_cases.append({"length": length, "signals": bx,
                              "tracks": [[t.start, t.end, t.freq_start, t.freq_end] for t in tracks], "tags": tags,
                              "bird_labels": bird_labels, "cacophony_index": ci, "version": version, "chirps": chirps,
                              "signals_left": len(chirp_signals), "max_chirps": it.get_max_chirps(length),
                              "rec_signals": [[float(v) for v in s.to_array()] for s in sigs]})
    # a few hand-made coverage cases (spans crossing several periods, out-of-order tracks, int lengths)
    for length, spans in ((60, [(1.0, 5.0), (4.0, 45.5), (44.0, 59.0)]), (61, [(0.0, 61.0)]), (40.0, [(30.0, 35.0), (5.0, 8.0)]),
                          (19.0, [(2.0, 3.0)]), (100.5, [(10.0, 95.0), (96.0, 100.5)])):
        tr = [it.Signal(a, b, 1000, 3000) for a, b in spans]
        ci, version = an.calc_cacophony_index(tr, length)
        analyse_cases.append({"length": length, "signals": [], "tracks": [[a, b, 1000, 3000] for a, b in spans],
                              "tags": ["bird"] * len(spans), "bird_labels": ["bird"], "cacophony_index": ci,
                              "version": version, "chirps": 0, "signals_left": 0, "max_chirps": it.get_max_chirps(length),
                              "rec_signals": []})
    json.dump(tracks_cases, open(os.path.join(OUT, "tracks.json"), "w"))
    json.dump(analyse_cases, open(os.path.join(OUT, "analyse.json"), "w"))
    print(len(tracks_cases), "track cases;", len(analyse_cases), "analyse cases;",
          sum(len(c["tracks"]) for c in tracks_cases), "tracks from", sum(len(c["signals"]) for c in tracks_cases), "signals")


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Timings of the general-parameter paths beside the CPU oracle (one 60 s recording, host arrays in / out):
load_samples with n_fft 2048 / 1024 (unfused: segment rows -> framed spectrum + mel -> dB) against the fused
n_fft 4096 call, and identify_morepork's spectrogram + windows at 48 kHz (mixed-radix kernel), 32 kHz and
44.1 / 22.05 kHz (direct band DFT).  Prints one JSON line."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from audio_analysis_b200 import synth, identify_tracks as T, identify_morepork as M  # noqa: E402
from oracle import oracle_np as O  # noqa: E402


def timed(fn, reps):
    fn()
    t0 = time.perf_counter()
    for _ in range(reps):
        out = fn()
    return (time.perf_counter() - t0) / reps * 1e3, out


def main():
    res = {"config": "one 60 s recording per call, host arrays in / host results out, single process"}
    rec = synth.recording(4242, 60.0, 48000)
    for n_fft in (4096, 2048, 1024):
        kw = dict(hop_length=281, mel_break=1750, n_mels=160, fmin=50, fmax=11000, power=2, db_scale=True,
                  normalize=True, n_fft=n_fft)
        ms, got = timed(lambda: T.load_samples(rec, 48000, [T.Signal(0, 60.0, 0, 24000)], 3.0, 1.5, **kw), 10)
        cms, ref = timed(lambda: O.load_samples(rec, 48000, [O.Signal(0, 60.0, 0, 24000)], 3.0, 1.5, **kw), 1)
        worst = max(float(np.abs(a - np.asarray(b)).max()) for a, b in zip(got[0], ref[0]))
        res["load_samples_n_fft_%d" % n_fft] = {"ours_ms": round(ms, 3), "cpu_oracle_ms": round(cms, 1),
                                                 "max_db_diff": worst}
    for sr in (48000, 32000, 44100, 22050):
        r = synth.recording(77, 60.0, sr)
        ms, (got, _) = timed(lambda: M.samples_from_recording((r, sr)), 10)
        cms, (ref, _) = timed(lambda: O.morepork_windows(O.morepork_spectrogram(r, sr)), 1)
        res["morepork_%d" % sr] = {"ours_ms": round(ms, 3), "cpu_oracle_ms": round(cms, 1), "n_fft": int(sr / 10),
                                   "max_window_diff": float(np.abs(got - ref).max())}
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()

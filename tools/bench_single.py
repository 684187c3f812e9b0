#!/usr/bin/env python
"""BASELINE.json configs[0]: ONE 60 s 48 kHz recording through the drop-in calls a reference user makes
(get_end, signal_noise, load_samples S1, cacophony_index on the 16 kHz twin), host arrays in, host results out,
next to the CPU oracle port on the same box.  Prints one JSON line."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from audio_analysis_b200 import synth, identify_tracks as T, cacophony_index as C, resample as R  # noqa: E402
from oracle import oracle_np as O  # noqa: E402

S1 = dict(hop_length=281, mel_break=1750, n_mels=160, fmin=50, fmax=11000, power=2, db_scale=True, normalize=True,
          n_fft=4096)


def timed(fn, reps):
    fn()
    t0 = time.perf_counter()
    for _ in range(reps):
        out = fn()
    return (time.perf_counter() - t0) / reps * 1e3, out


def main():
    import torch
    rec = synth.recording(4242, 60.0, 48000)
    rec16 = R.to_16k(rec)
    ours, ref = {}, {}
    ours["get_end"], _ = timed(lambda: T.get_end(rec, 48000), 10)
    ours["signal_noise"], sig = timed(lambda: T.signal_noise(rec, 48000), 10)
    ours["load_samples"], seg = timed(lambda: T.load_samples(rec, 48000, [T.Signal(0, 60.0, 0, 24000)], 3.0, 1.5, **S1), 10)
    ours["to_16k"], _ = timed(lambda: R.to_16k(rec), 10)
    ours["cacophony_index"], ci = timed(lambda: C.calculate(rec16), 10)
    torch.cuda.synchronize()

    def once():     # the same three identify_tracks calls on ONE upload of the recording (T.to_device)
        dev = T.to_device(rec)
        return (T.get_end(dev, 48000), T.signal_noise(dev, 48000),
                T.load_samples(dev, 48000, [T.Signal(0, 60.0, 0, 24000)], 3.0, 1.5, **S1))
    ours_once, _ = timed(once, 10)
    ref["get_end"], _ = timed(lambda: O.get_end(rec, 48000), 1)
    ref["signal_noise"], rsig = timed(lambda: O.signal_noise(rec, 48000), 1)
    ref["load_samples"], rseg = timed(lambda: O.load_samples(rec, 48000, [O.Signal(0, 60.0, 0, 24000)], 3.0, 1.5, **S1), 1)
    ref["to_16k"], _ = timed(lambda: O.resample_poly(rec, 1, 3), 1)
    ref["cacophony_index"], rci = timed(lambda: O.calculate_from_array(rec16), 1)
    assert ci == rci and len(sig) == len(rsig) and len(seg[0]) == len(rseg[0]) == 39
    worst = max(float(np.abs(a - np.asarray(b)).max()) for a, b in zip(seg[0], rseg[0]))
    print(json.dumps({"config": "configs[0]: one 60 s 48 kHz recording, host arrays in / host results out, single process",
                      "ours_ms": {k: round(v, 3) for k, v in ours.items()}, "ours_total_ms": round(sum(ours.values()), 3),
                      "ours_identify_tracks_one_upload_ms": round(ours_once, 3),
                      "cpu_oracle_ms": {k: round(v, 1) for k, v in ref.items()},
                      "cpu_oracle_total_ms": round(sum(ref.values()), 1),
                      "speedup_total": round(sum(ref.values()) / sum(ours.values()), 1),
                      "max_db_diff_load_samples": worst, "n_signals": len(sig),
                      "ci_entries": len(ci["cacophony_index_old"])}), flush=True)


if __name__ == "__main__":
    main()

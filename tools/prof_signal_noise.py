#!/usr/bin/env python
"""signal_noise on one 60 s recording: wall time per call (for an ncu launch list: run under
`ncu --metrics gpu__time_duration.sum -c 400 --csv`)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from audio_analysis_b200 import synth, identify_tracks as T  # noqa: E402

rec = synth.recording(4242, 60.0, 48000)
T.signal_noise(rec, 48000)
t0 = time.perf_counter()
n = int(os.environ.get("REPS", "5"))
for _ in range(n):
    s = T.signal_noise(rec, 48000)
print("signal_noise: %.3f ms per call, %d signals" % ((time.perf_counter() - t0) / n * 1e3, len(s)))

#!/usr/bin/env python
"""Device-resident run of the batched classify prep for profiling (ncu launch lists): `n` one-minute recordings,
`reps` repetitions of get_end_batch + signal_noise_batch's GPU part.  usage: python tools/prof_signal_noise.py [n] [reps]"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from audio_analysis_b200 import synth, identify_tracks as T  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 48
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
pool = np.stack([synth.recording(700 + i, 60.0, 48000) for i in range(8)])
pcm = torch.from_numpy(pool).cuda().repeat((n + 7) // 8, 1)[:n].contiguous()
m = pcm.shape[1]
tabs = torch.from_numpy(np.stack([np.arange(n, dtype=np.int64) * m, np.full(n, m, dtype=np.int64)])).cuda()
sn = T._signal_noise_plan(48000, 281)
ge = T._get_end_plan(48000)
flat = pcm.view(-1)
for _ in range(2):
    ge.get_end_chunks(flat, tabs[0], tabs[1], 170, max_item_samples=m)
    sn.run(flat, tabs[0], tabs[1], max_components=4096, max_item_samples=m)
torch.cuda.synchronize()
for name, fn in (("get_end_batch", lambda: ge.get_end_chunks(flat, tabs[0], tabs[1], 170, max_item_samples=m)),
                 ("signal_noise_batch", lambda: sn.run(flat, tabs[0], tabs[1], max_components=4096, max_item_samples=m))):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print("%s: %.3f ms per recording (%d recordings, %d reps)" % (name, e0.elapsed_time(e1) / reps / n, n, reps), flush=True)

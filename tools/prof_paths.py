#!/usr/bin/env python
"""Small fixed workloads of the secondary paths for ncu captures: python tools/prof_paths.py ci|morepork|getend"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from audio_analysis_b200 import synth  # noqa: E402

which = sys.argv[1]
if which == "ci":
    from audio_analysis_b200 import cacophony_index as C
    pool = np.stack([synth.recording(900 + i, 60.0, 16000) for i in range(8)])
    pcm = torch.from_numpy(pool).cuda().repeat(15, 1).contiguous().reshape(-1)      # 2 hours
    plan = C._plan(2048, 16000)
    for _ in range(3):
        plan.run(pcm, [0], [pcm.numel()])
elif which == "morepork":
    from audio_analysis_b200 import identify_morepork as M
    pool = np.stack([synth.recording(300 + i, 60.0, 48000) for i in range(8)])
    pcm = torch.from_numpy(pool).cuda().repeat(32, 1).contiguous()                  # 256 recordings
    plan = M.BatchPlan(48000, pcm.shape[1])
    for _ in range(3):
        plan.run(pcm)
elif which == "getend":
    from audio_analysis_b200 import identify_tracks as T
    pool = np.stack([synth.recording(300 + i, 60.0, 48000) for i in range(8)])
    pcm = torch.from_numpy(pool).cuda().repeat(2, 1).contiguous()
    plan = T._get_end_plan(48000)
    n = pcm.shape[1]
    for _ in range(3):
        plan.run(pcm.view(-1), np.arange(16) * n, np.full(16, n))
torch.cuda.synchronize()
print("done", which)

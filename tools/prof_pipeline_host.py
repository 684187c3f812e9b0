import cProfile, pstats, sys, os, io
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tools"))
import numpy as np, torch
from audio_analysis_b200 import identify_tracks as T, synth
pool = [synth.recording(700 + i, seconds=60.0) for i in range(16)]
pcm = torch.from_numpy(np.stack(pool)).cuda().repeat(16, 1).contiguous()
recs = [pcm[i] for i in range(256)]
S1 = dict(hop_length=281, mel_break=1750, n_mels=160, fmin=50, fmax=11000, power=2, db_scale=True, normalize=True, n_fft=4096)
np.random.seed(1); T.classify_prep_batch(recs, 48000, 3.0, 1.5, **S1); torch.cuda.synchronize()
import time
t0 = time.perf_counter(); np.random.seed(1); T.classify_prep_batch(recs, 48000, 3.0, 1.5, **S1); torch.cuda.synchronize(); print("wall ms", (time.perf_counter() - t0) * 1e3)
for ch in (64, 32, 24, 16, 12, 8):
    best = 1e9
    for rep in range(3):
        t0 = time.perf_counter(); np.random.seed(1); T.classify_prep_batch(recs, 48000, 3.0, 1.5, chunk=ch, **S1); torch.cuda.synchronize()
        best = min(best, (time.perf_counter() - t0) * 1e3)
    print("chunk %d: %.1f ms for 256 recordings" % (ch, best))
pr = cProfile.Profile()
np.random.seed(1)
pr.enable(); T.classify_prep_batch(recs, 48000, 3.0, 1.5, **S1); torch.cuda.synchronize(); pr.disable()
st = io.StringIO(); pstats.Stats(pr, stream=st).sort_stats("cumulative").print_stats(28); print(st.getvalue()[:6000])

"""Generate tests/golden/signal_noise.npz by executing the REFERENCE's own signal_noise
(src/identify_tracks.py:650-706, unchanged, cv2 4.x as in requirements.txt) under the stand-in
librosa.  Run in the build container only:  python tests/golden/make_golden_signal_noise.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import reference_loader as rl  # noqa: E402
from audio_analysis_b200 import synth  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def main():
    it = rl.load("identify_tracks")
    out = {}
    for name, seed, seconds in (("a", 1234, 12.0), ("b", 77, 20.0), ("c", 4242, 6.5)):
        rec = synth.recording(seed, seconds=seconds)
        sigs = it.signal_noise(rec, 48000)
        out[name + "_seed"] = np.int64(seed)
        out[name + "_seconds"] = np.float64(seconds)
        out[name + "_signals"] = np.array([[s.start, s.end, s.freq_start, s.freq_end] for s in sigs],
                                          dtype=np.float64).reshape(-1, 4)
        print(name, len(sigs), "signals")
    np.savez_compressed(os.path.join(OUT, "signal_noise.npz"), **out)


if __name__ == "__main__":
    main()


def slaney():
    """tests/golden/get_spect_slaney.npz: the reference's get_spect(htk=False) under the librosa stand-in."""
    from oracle import oracle_np as O
    it = rl.load("identify_tracks")
    rec = synth.recording(1234, seconds=12.0)
    x = O.normalize_data(rec[48000:48000 + 144000])
    ref = np.asarray(it.get_spect(x, 48000, 281, False, False, 1750, False, 160, 50, 11000, 4096, 2, True))
    np.savez_compressed(os.path.join(OUT, "get_spect_slaney.npz"), seed=1234, seg_start=48000,
                        s1_db=ref[:, :, 0].astype(np.float32))


if __name__ == "__main__" and "slaney" in sys.argv:
    slaney()

"""GPU polyphase resampler vs scipy.signal.resample_poly (the front-end's stated filter; the reference's
own ffmpeg / soxr filters are not available here, so parity for this row is against scipy, as DESIGN.md says)."""
import numpy as np
import pytest

from audio_analysis_b200 import synth
from oracle import oracle_np as O

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("orig,target,n", [(48000, 16000, 480001), (48000, 16000, 7), (44100, 48000, 44100 * 2 + 13),
                                           (16000, 48000, 16000), (22050, 48000, 30011), (48000, 48000, 100)])
def test_resample_matches_scipy(engine, orig, target, n):
    from audio_analysis_b200 import resample as R
    import math
    x = synth.recording(900 + n % 97, seconds=max(n / orig, 0.01) + 0.01, sr=orig)[:n]
    g = math.gcd(orig, target)
    ref = O.resample_poly(x, target // g, orig // g)
    got = R.resample(x, orig, target)
    assert got.shape == ref.shape and got.dtype == np.float32
    d = np.abs(got - ref).max() if len(ref) else 0.0
    print("resample %d -> %d: max abs diff %.3e, bitwise equal %.4f" % (orig, target, d, float(np.mean(got == ref))))
    assert np.array_equal(got, ref)          # same taps, same float64 accumulation order: bit for bit


def test_batch_and_cacophony_index_feed(engine):
    """One 48 kHz upload feeds the Cacophony Index: to_16k on the device, then calculate on the result."""
    import torch
    from audio_analysis_b200 import resample as R, cacophony_index as CI
    recs = synth.batch(950, 3, seconds=30.0)
    dev = torch.from_numpy(recs).cuda()
    low = R.to_16k(dev)
    assert low.is_cuda and tuple(low.shape) == (3, 480000)
    for i in range(3):
        ref = O.resample_poly(recs[i], 1, 3)
        assert np.array_equal(low[i].cpu().numpy(), ref)
    res = CI.calculate_batch(low.reshape(-1), np.arange(3) * 480000, [480000] * 3)
    for i in range(3):
        assert res[i] == O.calculate_from_array(O.resample_poly(recs[i], 1, 3), fast=True)

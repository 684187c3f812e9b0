"""GPU parity of the chained classify prep (src/identify_tracks.py:416-440, 502-524) and of the MagTransform epilogue
(src/magtransform.py:17-19) through the C-ABI."""
import numpy as np
import pytest

from audio_analysis_b200 import synth
from oracle import oracle_np as O

pytestmark = pytest.mark.gpu
S1 = dict(hop_length=281, mel_break=1750, n_mels=160, fmin=50, fmax=11000, power=2, db_scale=True, normalize=True,
          n_fft=4096)


def _box(s):
    return (s.start, s.end, s.freq_start, s.freq_end)


def test_classify_prep_batch_equals_the_separate_calls(engine):
    """get_end -> signal_noise(frames[: int(sr * length)]) -> get_tracks_from_signals -> load_samples, chained on the
    device with two synchronisations, equals the reference's sequence of calls (oracle) recording by recording."""
    from audio_analysis_b200 import identify_tracks as T, tracks as TR
    recs = [synth.recording(41, seconds=14.0), synth.recording(42, seconds=9.0, zero_tail_s=3.0),
            synth.recording(43, seconds=11.3), np.zeros(48000 * 3, np.float32)]
    np.random.seed(5)
    got = T.classify_prep_batch(recs, 48000, 3.0, 1.5, **S1)
    assert len(got) == len(recs)
    np.random.seed(5)
    for rec, g in zip(recs, got):
        length = O.get_end(rec, 48000)
        assert g["length"] == length and g["raw_length"] == len(rec) / 48000
        ref_sig = O.signal_noise(rec[: int(48000 * length)], 48000)
        assert [_box(s) for s in g["signals"]] == [tuple(r) for r in ref_sig]
        ref_tracks = TR.get_tracks_from_signals([TR.Signal(*r) for r in ref_sig], length)
        assert [_box(t) for t in g["tracks"]] == [_box(t) for t in ref_tracks]
        if not ref_tracks:
            assert g["samples"] is None
            continue
        ref = O.load_samples(rec, 48000, [O.Signal(*_box(t)) for t in ref_tracks], 3.0, 1.5, **S1)
        assert len(g["samples"]) == len(ref)
        for a, b in zip(g["samples"], ref):
            a = a.cpu().numpy()
            assert a.shape[0] == len(b)
            for x, y in zip(a, b):
                assert x.shape == y.shape and np.abs(x - y).max() <= 1e-2
    # one recording through the single-recording entry = the same record of the batch
    np.random.seed(5)
    one = T.classify_prep(recs[0], 48000, **dict(S1, segment_length=3.0, segment_stride=1.5))
    assert [_box(t) for t in one["tracks"]] == [_box(t) for t in got[0]["tracks"]] and one["length"] == got[0]["length"]
    assert all(bool((a == b).all()) for a, b in zip(one["samples"], got[0]["samples"]))


def test_magtransform_epilogue_equals_separate_pass_and_reference(engine, golden):
    """cac_plan_set_magtransform: mel ** sigmoid(a) written by the fused kernel's flush == cac_magtransform_run on the
    plain mel output (same powf, bit for bit) and the reference layer's formula on the oracle's mel power."""
    import torch
    from audio_analysis_b200 import custommel, identify_tracks as T
    rec = synth.recording(9, seconds=7.5)
    w = custommel.mel_f(48000, 160, 50, 11000, 4096, 1750)
    segs, _ = T.segment_descriptors(len(rec), 48000, [T.Signal(0.0, 7.5, 0, 24000)], 3.0, 1.5)
    dev = torch.from_numpy(rec).cuda()
    for hop in (281, 640):
        plan = engine.FeaturePlan(w, 48000, 4096, hop, 160, 144000, power=2, db_scale=False, normalize=True)
        plain = plan.run(dev, segs).clone()
        for a in (0.0, -1.0, 0.7):
            plan.set_magtransform(a)
            fused = plan.run(dev, segs).clone()
            assert torch.equal(fused, engine.magtransform(plain, a))
            ref = O.magtransform(plain.cpu().numpy(), a)
            assert np.allclose(fused.cpu().numpy(), ref, rtol=2e-6, atol=0)
        plan.set_magtransform(None)
        assert torch.equal(plan.run(dev, segs), plain)
    db_plan = engine.FeaturePlan(w, 48000, 4096, 281, 160, 144000, power=2, db_scale=True)
    with pytest.raises(engine.CacError):
        db_plan.set_magtransform(0.0)

"""GPU parity of the framed-spectrum paths through the C-ABI: identify_morepork spectrogram +
windows and identify_tracks.get_end, against the reference-generated fixtures and the oracle.
Tolerances: magnitudes <= 1e-4 relative on >= 90 % of cells and <= 2e-6 of the spectrum maximum
everywhere (float32 FFT vs the reference's float64); morepork window values (dB / 80, in [0, 1])
<= 1.25e-4 = 1e-2 dB; get_end bit-exact."""
import numpy as np
import pytest

from audio_analysis_b200 import synth
from oracle import oracle_np as O

pytestmark = pytest.mark.gpu


def _check_mag(got, ref):
    assert got.shape == ref.shape
    scale = float(ref.max())
    assert np.abs(got - ref).max() <= 2e-6 * scale
    frac = float(np.mean(np.abs(got - ref) <= 1e-4 * np.abs(ref)))
    assert frac >= 0.9, frac
    return frac


@pytest.mark.parametrize("sr,key,seed,seconds", [(48000, "48", 1234, 12.0), (16000, "16", 55, 8.0)])
def test_morepork_spectrogram_and_windows_vs_reference_fixture(engine, golden, sr, key, seed, seconds):
    from audio_analysis_b200 import identify_morepork as M
    g = golden("morepork")
    rec = synth.recording(seed, seconds=seconds, sr=sr)
    got_sr, spec = M._load_sample((rec, sr))
    assert got_sr == sr
    frac = _check_mag(spec, g["spec" + key])
    samples, starts = M.samples_from_recording((rec, sr))
    ref = g["win" + key]
    assert samples.shape == ref.shape + (1,) and samples.dtype == np.float32
    assert np.array_equal(np.array(starts), g["starts" + key])
    d = np.abs(samples[..., 0] - ref)
    print("morepork %s: spec match@1e-4 %.5f, window max diff %.3e" % (key, frac, d.max()))
    assert d.max() <= 1.25e-4
    assert samples.max() <= 1.0 and samples.min() >= 0.0


@pytest.mark.parametrize("n", [2400 * 10, 2400 * 40 + 17, 2400 * 61, 2400 * 75 - 1, 100])
def test_morepork_short_recordings_and_wraparound_slices(engine, n):
    """fewer than 60 frames -> the reference's negative-start slice (src/identify_morepork.py:146-149)"""
    from audio_analysis_b200 import identify_morepork as M
    rec = synth.recording(9, seconds=max(n / 48000.0, 0.05) + 0.01)[:n]
    samples, starts = M.samples_from_recording((rec, 48000))
    ref, ref_starts = O.morepork_windows(O.morepork_spectrogram(rec, 48000))
    assert samples.shape == ref.shape and list(starts) == list(ref_starts)
    assert np.abs(samples - ref).max() <= 1.25e-4


@pytest.mark.parametrize("sr,seconds", [(32000, 7.0), (44100, 6.5), (22050, 9.0), (8000, 12.0)])
def test_morepork_other_native_rates_vs_oracle(engine, sr, seconds):
    """_load_sample keeps the file's own rate (librosa.load(sr=None), src/identify_morepork.py:20): 32 kHz runs on the
    mixed-radix kernel (n_fft 3200), the rates whose n_fft = int(sr / 10) has other prime factors or is odd
    (4410, 2205 with hop 1102, 800) on the direct band DFT (band_dft.cu)."""
    from audio_analysis_b200 import identify_morepork as M
    rec = synth.recording(4000 + sr // 100, seconds=seconds, sr=sr)
    ref_spec = O.morepork_spectrogram(rec, sr)
    got_sr, spec = M._load_sample((rec, sr))
    assert got_sr == sr
    frac = _check_mag(spec, ref_spec)
    samples, starts = M.samples_from_recording((rec, sr))
    ref, ref_starts = O.morepork_windows(ref_spec)
    assert samples.shape == ref.shape and list(starts) == list(ref_starts)
    d = np.abs(samples - ref)
    print("morepork %d Hz: spec match@1e-4 %.5f, window max diff %.3e" % (sr, frac, d.max()))
    assert d.max() <= 1.25e-4


def test_morepork_batch_equals_single(engine):
    import torch
    from audio_analysis_b200 import identify_morepork as M
    recs = synth.batch(40, 3, seconds=10.0)
    bp = M.BatchPlan(48000, recs.shape[1])
    out, starts = bp.run(torch.from_numpy(recs).cuda())
    assert out.shape == (3, bp.n_windows, 60, 60)
    for i in range(3):
        single, s1 = M.samples_from_recording((recs[i], 48000))
        assert np.array_equal(out[i].cpu().numpy(), single[..., 0]) and s1 == starts


def test_get_end_vs_reference_fixture(engine, golden):
    from audio_analysis_b200 import identify_tracks as T
    g = golden("get_end")
    rec_tail = synth.recording(77, seconds=6.0, zero_tail_s=3.0)
    assert T.get_end(rec_tail, 48000) == float(g["end_tail"]) == O.get_end(rec_tail, 48000)
    rec = synth.recording(1234, seconds=12.0)
    assert T.get_end(rec, 48000) == float(g["end_full"])
    for tail in (0.5, 1.2, 2.0):
        r = synth.recording(31, seconds=4.0, zero_tail_s=tail)
        assert T.get_end(r, 48000) == O.get_end(r, 48000)


def test_get_end_mel_matches_oracle(engine):
    import torch
    from audio_analysis_b200 import identify_tracks as T
    rec = synth.recording(8, seconds=3.0)
    plan = T._get_end_plan(48000)
    mel, base = plan.run(torch.from_numpy(rec).cuda(), [0], [rec.shape[0]])
    n = int(base[-1])
    got = mel.view(120, n).cpu().numpy()
    ref = O.mel_spec(np.abs(O.stft(rec, 4800, 281)), 48000, 4800, 281, 120, 50, 11000, 1750, power=1)
    assert got.shape == ref.shape == (120, 1 + rec.shape[0] // 281)
    assert np.abs(got - ref).max() <= 1e-5 * ref.max()
    assert float(np.mean(np.abs(got - ref) <= 1e-4 * np.abs(ref))) >= 0.9


def test_spec_plan_raw_layouts_and_max(engine):
    import torch
    rec = synth.recording(12, seconds=2.0)
    ref = np.abs(O.stft(rec, 4800, 1000))
    pcm = torch.from_numpy(np.concatenate([rec, rec[:50000]])).cuda()
    for bin_major in (False, True):
        plan = engine.SpecPlan(4800, 1000, bin_lo=0, bin_hi=2401, power=1, bin_major=bin_major)
        out, base, mx = plan.run(pcm, [0, rec.shape[0]], [rec.shape[0], 50000], want_max=True)
        n0 = int(base[1])
        a = out[:n0 * 2401]
        got = (a.view(2401, n0) if bin_major else a.view(n0, 2401).t()).cpu().numpy()
        _check_mag(got, ref)
        ref2 = np.abs(O.stft(rec[:50000], 4800, 1000))
        b = out[n0 * 2401:]
        n1 = int(base[2] - base[1])
        got2 = (b.view(2401, n1) if bin_major else b.view(n1, 2401).t()).cpu().numpy()
        _check_mag(got2, ref2)
        assert abs(float(mx.item()) - max(ref.max(), ref2.max())) <= 2e-6 * ref.max()
    p2 = engine.SpecPlan(4800, 2400, bin_lo=0, bin_hi=2401, power=2)
    out, base = p2.run(pcm, [0], [rec.shape[0]])
    got = out.view(int(base[1]), 2401).t().cpu().numpy()
    _check_mag(got, np.abs(O.stft(rec, 4800, 2400)) ** 2)


def test_spec_plan_rejects_unsupported(engine):
    with pytest.raises(engine.CacError):      # the direct band DFT serves raw bands only, not a mel epilogue
        engine.SpecPlan(4410, 281, n_mels=120, power=1, mel_weights=np.ones((120, 2206), np.float32))
    with pytest.raises(engine.CacError):      # ... and not more than 512 bins
        engine.SpecPlan(4410, 2205, bin_lo=0, bin_hi=2206)
    with pytest.raises(engine.CacError):
        engine.SpecPlan(4800, 2400, bin_lo=100, bin_hi=3000)


def test_signal_noise_spectrogram_vs_reference_fixture(engine, golden):
    from audio_analysis_b200 import identify_tracks as T
    g = golden("get_end")
    rec = synth.recording(1234, seconds=12.0)
    got = T.signal_noise_spectrogram(rec[:96000], 48000).cpu().numpy()
    assert got.shape == tuple(g["signal_noise_shape"]) == (2049, 1 + 96000 // 281)
    ref_sub = g["signal_noise_sub"]                     # the reference's own output, every 8th row / column
    assert np.abs(got[::8, ::8] - ref_sub).max() <= 2e-6
    ref = O.signal_noise_spectrogram(rec[:96000], 48000)
    assert got.max() == 1.0
    _check_mag(got, ref)


def test_magtransform_vs_reference_fixture(engine, golden):
    from audio_analysis_b200 import magtransform as M1, magtransformv2 as M2
    g = golden("magtransform")
    x = g["x"]
    y1, y2 = M1.MagTransform()(x), M2.MagTransform()(x)
    assert y1.dtype == np.float32 and y1.shape == x.shape
    assert np.allclose(y1, g["v1"], rtol=2e-6, atol=0) and np.allclose(y2, g["v2"], rtol=2e-6, atol=0)
    big = np.abs(np.random.default_rng(0).standard_normal(100003)).astype(np.float32) * 50
    assert np.allclose(M1.MagTransform(a=0.7)(big), O.magtransform(big, 0.7), rtol=2e-6, atol=0)
    assert float(M2.MagTransform().a) == -1.0 and float(M1.MagTransform().a) == 0.0

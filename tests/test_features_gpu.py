"""GPU parity: STFT -> custom mel -> dB through the C-ABI vs the oracle (same seeded inputs).
Tolerances are BASELINE.json's: mel power <= 1e-4 relative, power_to_db <= 1e-2 dB max-abs,
>= 90 % of outputs matching; slicing bit-exact."""
import numpy as np
import pytest

from audio_analysis_b200 import synth
from oracle import oracle_np as O

pytestmark = pytest.mark.gpu

S1 = dict(hop_length=281, mel_break=1750, n_mels=160, fmin=50, fmax=11000, power=2, n_fft=4096)


def _match_fraction(got, ref, rtol):
    return float(np.mean(np.abs(got - ref) <= rtol * np.abs(ref)))


def test_library_loaded(engine):
    assert "sm_100a" in engine.version()
    assert engine.lib.cac_device_count() >= 1


def test_normalize_bit_exact(engine):
    from audio_analysis_b200 import identify_tracks as T
    x = synth.recording(3, seconds=3.0)
    assert np.array_equal(T.normalize_data(x), O.normalize_data(x))
    xs = np.stack([x[:48000], x[48000:96000] * 0.3 + 0.1])
    assert np.array_equal(T.normalize_data(xs), O.normalize_data(xs))


def test_get_spect_power_vs_golden_and_oracle(engine, golden):
    from audio_analysis_b200 import identify_tracks as T
    g = golden("get_spect")
    rec = synth.recording(int(g["seed"]), seconds=12.0)
    s0 = int(g["seg_start"])
    x = O.normalize_data(rec[s0:s0 + 144000])
    got = T.get_spect(x, 48000, 640, False, False, 1750, True, 160, 50, 11000, 4096, 2, False)
    assert got.shape == (160, 226, 1) and got.dtype == np.float32
    ref = g["hop640_power"]
    frac = _match_fraction(got[:, :, 0], ref, 1e-4)
    worst = float((np.abs(got[:, :, 0] - ref) / np.abs(ref)).max())
    print("mel power hop640: match@1e-4 = %.5f worst rel = %.3e" % (frac, worst))
    assert frac >= 0.9
    assert np.abs(got[:, :, 0] - ref).max() <= 1e-5 * ref.max()


def test_get_spect_db_vs_golden(engine, golden):
    from audio_analysis_b200 import identify_tracks as T
    g = golden("get_spect")
    rec = synth.recording(int(g["seed"]), seconds=12.0)
    s0 = int(g["seg_start"])
    x = O.normalize_data(rec[s0:s0 + 144000])
    got = T.get_spect(x, 48000, 281, False, False, 1750, True, 160, 50, 11000, 4096, 2, True)
    assert got.shape == (160, 513, 1)
    d = np.abs(got[:, :, 0] - g["s1_db"])
    print("dB: max abs diff %.3e, frac<=1e-2: %.5f" % (d.max(), (d <= 1e-2).mean()))
    assert (d <= 1e-2).mean() >= 0.9
    assert d.max() <= 1e-2
    assert got.max() == 0.0 and got.min() >= -80.0


def test_power1_meansub_channels(engine, golden):
    from audio_analysis_b200 import identify_tracks as T
    g = golden("get_spect")
    rec = synth.recording(int(g["seed"]), seconds=12.0)
    s0 = int(g["seg_start"])
    x = O.normalize_data(rec[s0:s0 + 144000])
    got = T.get_spect(x, 48000, 281, True, False, 1750, True, 160, 50, 11000, 4096, 1, False, 3)
    assert got.shape == (160, 513, 3)
    ref = g["power1_meansub_ch3"]
    scale = np.abs(ref).max()
    assert np.abs(got[::4, ::4, :] - ref).max() <= 2e-5 * scale


def test_load_samples_matches_oracle_with_ragged_tracks(engine):
    from audio_analysis_b200 import identify_tracks as T
    rec = synth.recording(1234, seconds=12.0)
    spec = [(0.0, 12.0, 0, 24000), (1.3, 2.1, 800, 3000), (0.05, 0.9, 500, 900), (11.2, 11.95, 2000, 9000),
            (3.0, 9.7, 100, 12000), (2.0, 5.0, 12000, 15000), (6.1, 9.1000001, 50, 11000)]
    kw = dict(db_scale=True, normalize=True, **S1)
    np.random.seed(5)
    got = T.load_samples(rec, 48000, [T.Signal(*s) for s in spec], 3.0, 1.5, **kw)
    np.random.seed(5)
    ref = O.load_samples(rec, 48000, [O.Signal(*s) for s in spec], 3.0, 1.5, **kw)
    assert [len(t) for t in got] == [len(t) for t in ref]
    worst, total, ok = 0.0, 0, 0
    for tg, tr in zip(got, ref):
        for a, b in zip(tg, tr):
            assert a.shape == b.shape
            d = np.abs(a - b)
            worst = max(worst, float(d.max()))
            total += d.size
            ok += int((d <= 1e-2).sum())
    print("load_samples: %d cells, max |dB diff| %.3e, match %.6f" % (total, worst, ok / total))
    assert ok / total >= 0.9 and worst <= 1e-2


def test_load_samples_golden_reference_segments(engine, golden):
    """Directly against outputs of the reference's own load_samples (tests/golden)."""
    from audio_analysis_b200 import identify_tracks as T
    g = golden("load_samples")
    rec = synth.recording(int(g["rec_seed"]), seconds=12.0)
    np.random.seed(int(g["seed"]))
    got = T.load_samples(rec, 48000, [T.Signal(0.0, 12.0, 0, 24000), T.Signal(5.2, 6.0, 1000, 4000)], 3.0, 1.5,
                         db_scale=True, normalize=True, **S1)
    assert [len(t) for t in got] == list(g["counts"])
    for a, b in ((got[0][0], g["t0_first"]), (got[0][-1], g["t0_last"]), (got[1][0], g["t1_only"])):
        assert np.abs(a[:, :, 0] - b).max() <= 1e-2


def test_mel_spec_projection(engine):
    from audio_analysis_b200 import custommel
    x = synth.recording(8, seconds=2.0)
    mag = np.abs(O.stft(x, 4096, 281))
    for power in (1, 2):
        got = custommel.mel_spec(mag, 48000, 4096, 281, 160, 50, 11000, 1750, power=power)
        ref = O.mel_spec(mag, 48000, 4096, 281, 160, 50, 11000, 1750, power=power)
        assert got.shape == ref.shape
        assert np.abs(got - ref).max() <= 1e-5 * np.abs(ref).max()
        assert _match_fraction(got, ref, 1e-4) >= 0.999
    mag2 = np.abs(O.stft(x, 4800, 281))
    got = custommel.mel_spec(mag2, 48000, 4800, 281, 120, 50, 11000, 1750, power=1)
    ref = O.mel_spec(mag2, 48000, 4800, 281, 120, 50, 11000, 1750, power=1)
    assert _match_fraction(got, ref, 1e-4) >= 0.999


def test_device_batch_equals_host_path_and_is_deterministic(engine):
    import torch
    from audio_analysis_b200 import identify_tracks as T
    recs = synth.batch(100, 3, seconds=6.0)
    segs, counts = T.segment_descriptors(recs.shape[1], 48000, [T.Signal(0, 6.0, 0, 24000)], 3.0, 1.5)
    allsegs = np.concatenate([segs.copy() for _ in range(3)])
    for i in range(3):
        allsegs["pcm_offset"][i * len(segs):(i + 1) * len(segs)] += i * recs.shape[1]
    plan = T._plan(48000, 4096, 281, 160, 50, 11000, 1750, 144000, 2, True, True)
    pcm = torch.from_numpy(recs.reshape(-1)).cuda()
    a = plan.run(pcm, allsegs).cpu().numpy()
    b = plan.run(pcm, allsegs).cpu().numpy()
    c = plan.run_host(recs.reshape(-1), allsegs)
    assert np.array_equal(a, b) and np.array_equal(a, c)
    ref = O.load_samples(recs[1], 48000, [O.Signal(0, 6.0, 0, 24000)], 3.0, 1.5, db_scale=True, **S1)
    assert np.abs(a[len(segs)] - ref[0][0][:, :, 0]).max() <= 1e-2


def test_error_paths(engine):
    import ctypes
    w = np.zeros((160, 2049), np.float32)
    p = engine.CacParams(48000, 4096, 281, 160, 144000, 3, 1, 1)
    h = ctypes.c_void_p()
    rc = engine.lib.cac_plan_create(ctypes.byref(p), w.ctypes.data_as(ctypes.c_void_p), 0, ctypes.byref(h))
    assert rc == -1 and b"power" in engine.lib.cac_last_error()
    from audio_analysis_b200 import custommel
    odd = custommel.mel_f(48000, 160, 50, 11000, 3000, 1750)    # a frame length no kernel is built for
    plan = engine.FeaturePlan(odd, 48000, 3000, 281, 160, 144000)
    with pytest.raises(engine.CacError):
        plan.run_host(np.zeros(144000, np.float32), np.array([(0, 144000, 0)], dtype=engine.SEGMENT_DTYPE))
    good = engine.FeaturePlan(custommel.mel_f(48000, 160, 50, 11000, 4096, 1750), 48000, 4096, 281, 160, 144000)
    with pytest.raises(engine.CacError):   # descriptor past the end of the pcm buffer
        good.run_host(np.zeros(1000, np.float32), np.array([(0, 144000, 0)], dtype=engine.SEGMENT_DTYPE))


def test_htk_false_slaney_branch(engine, golden):
    """get_spect(htk=False) = librosa.feature.melspectrogram (Slaney filterbank, fmin 50, fmax 11000, power 2)."""
    from audio_analysis_b200 import identify_tracks as T
    g = golden("get_spect_slaney")
    rec = synth.recording(int(g["seed"]), seconds=12.0)
    s0 = int(g["seg_start"])
    x = O.normalize_data(rec[s0:s0 + 144000])
    got = T.get_spect(x, 48000, 281, False, False, 1750, False, 160, 50, 11000, 4096, 2, True)
    assert got.shape == (160, 513, 1)
    assert np.abs(got[:, :, 0] - g["s1_db"]).max() <= 1e-2
    tracks = [(0.0, 6.0, 0, 24000)]
    a = T.load_samples(rec[:288000], 48000, [T.Signal(*t) for t in tracks], 3.0, 1.5, htk=False, db_scale=True, **S1)
    b = O.load_samples(rec[:288000], 48000, [O.Signal(*t) for t in tracks], 3.0, 1.5, htk=False, db_scale=True, **S1)
    assert len(a[0]) == len(b[0]) == 3
    for u, v in zip(a[0], b[0]):
        assert np.abs(u - v).max() <= 1e-2


def test_load_samples_device_tail_matches_host_path(engine):
    """Classifier hand-off on the device: [n_seg, n_mels, frames, channels] CUDA tensors per track equal the
    host path (load_samples) - bitwise without mean_sub, within float32 rounding of the row mean with it."""
    import torch
    from audio_analysis_b200 import identify_tracks as T
    rec = synth.recording(41, seconds=9.0)
    tracks = [T.Signal(0.0, 9.0, 0, 24000), T.Signal(1.0, 2.0, 13000, 15000), T.Signal(2.5, 7.1, 500, 9000)]
    kw = dict(segment_length=3.0, stride=1.5, db_scale=True, **S1)
    for mean_sub, channels in ((False, 1), (True, 1), (False, 3), (True, 3)):
        np.random.seed(5)
        host = T.load_samples(rec, 48000, tracks, mean_sub=mean_sub, channels=channels, **kw)
        np.random.seed(5)
        dev = T.load_samples_device(torch.from_numpy(rec).cuda(), 48000, tracks, mean_sub=mean_sub, channels=channels, **kw)
        assert len(dev) == len(host) == 3 and dev[1].shape[0] == 0 and len(host[1]) == 0
        for d, h in zip(dev, host):
            assert d.is_cuda and d.shape[0] == len(h)
            if len(h) == 0:
                continue
            hh = np.stack(h)
            assert tuple(d.shape) == hh.shape and d.shape[3] == channels
            if mean_sub:
                assert np.abs(d.cpu().numpy() - hh).max() <= 1e-5
                assert float(d.double().mean(dim=2).abs().max()) <= 2e-5
            else:
                assert np.array_equal(d.cpu().numpy(), hh)
    ref = O.load_samples(rec, 48000, [O.Signal(0.0, 9.0, 0, 24000)], 3.0, 1.5, db_scale=True, mean_sub=True, channels=3, **S1)
    np.random.seed(5)
    got = T.load_samples_device(rec, 48000, [T.Signal(0.0, 9.0, 0, 24000)], mean_sub=True, channels=3, **kw)[0]
    assert np.abs(got.cpu().numpy() - np.stack(ref[0])).max() <= 1e-2


def test_constant_segment_is_nan_like_numpy_and_raises_like_librosa(engine):
    """Digital silence inside a track: normalize_data divides 0 by 0 (src/identify_tracks.py:202-209), numpy
    carries NaN through stft / mel / power_to_db, librosa.stft raises ParameterError.  The C-ABI returns the NaNs
    (only for the constant segments), the drop-in raises."""
    from audio_analysis_b200 import identify_tracks as T
    rec = synth.recording(77, seconds=9.0)
    rec[3 * 48000:6 * 48000 + 10] = 0.25                      # segment k = 2 ([3 s, 6 s)) is constant
    segs, counts = T.segment_descriptors(len(rec), 48000, [T.Signal(0.0, 9.0, 0, 24000)], 3.0, 1.5)
    plan = T._plan(48000, 4096, 281, 160, 50, 11000, 1750, 144000, 2, True, True)
    out = plan.run_host(rec, segs)
    nan_seg = np.isnan(out).reshape(len(segs), -1)
    assert nan_seg[2].all() and not nan_seg[[0, 1, 3, 4]].any()
    power = T._plan(48000, 4096, 281, 160, 50, 11000, 1750, 144000, 2, False, True).run_host(rec, segs)
    assert np.isnan(power[2]).all() and np.isfinite(power[[0, 1, 3, 4]]).all()
    with pytest.raises(T.ParameterError):
        T.load_samples(rec, 48000, [T.Signal(0.0, 9.0, 0, 24000)], 3.0, 1.5, db_scale=True, **S1)
    with pytest.raises(O.ParameterError):
        O.load_samples(rec, 48000, [O.Signal(0.0, 9.0, 0, 24000)], 3.0, 1.5, db_scale=True, **S1)
    # without normalisation silence is legal: all-zero power -> 0 dB everywhere (max(1e-10, 0) on both sides)
    quiet = np.zeros(144000, np.float32)
    got = T.get_spect(quiet, 48000, 281, False, False, 1750, True, 160, 50, 11000, 4096, 2, True)
    ref = O.get_spect(quiet, 48000, 281, False, False, 1750, True, 160, 50, 11000, 4096, 2, True)
    assert np.array_equal(got, np.asarray(ref)) and float(np.abs(got).max()) == 0.0


@pytest.mark.parametrize("hop,n_mels,brk,fmin,fmax,power,db,norm,seg_len,stride", [
    (640, 160, 1000, 50, 11000, 2, True, True, 3.0, 1.5),      # classify's fallback hop / load_samples' default break
    (281, 120, 1750, 50, 11000, 1, False, True, 3.0, 1.0),     # power 1, no dB, 4 row groups with dead slots
    (281, 80, 1750, 100, 8000, 2, True, False, 2.5, 2.5),      # 3 row groups, no normalisation, other segment length
    (512, 33, 1000, 500, 3000, 2, True, True, 1.0, 0.5),       # two groups with 31 dead slots, narrow band, short segments
    (281, 200, 1750, 50, 11000, 2, False, True, 3.0, 1.5),     # 7 row groups (general gather path)
])
def test_parameter_sweep_vs_oracle(engine, hop, n_mels, brk, fmin, fmax, power, db, norm, seg_len, stride):
    """Every parameter of load_samples is runtime data: other hops, filterbank sizes (exercising the
    slot tables of mel_bands.h), band limits, power, dB and normalisation switches against the oracle."""
    from audio_analysis_b200 import identify_tracks as T
    rec = synth.recording(300 + hop + n_mels, seconds=7.0)
    tracks = [(0.0, 7.0, 0, 24000), (0.4, 1.1, 700, 2500), (5.5, 7.0, 1000, 4000)]
    kw = dict(hop_length=hop, mel_break=brk, n_mels=n_mels, fmin=fmin, fmax=fmax, power=power, db_scale=db,
              normalize=norm, n_fft=4096)
    np.random.seed(11)
    got = T.load_samples(rec, 48000, [T.Signal(*t) for t in tracks], seg_len, stride, **kw)
    np.random.seed(11)
    ref = O.load_samples(rec, 48000, [O.Signal(*t) for t in tracks], seg_len, stride, **kw)
    assert [len(g) for g in got] == [len(r) for r in ref]
    n_cells = n_ok = 0
    for g_track, r_track in zip(got, ref):
        for g, r in zip(g_track, r_track):
            r = np.asarray(r)
            assert g.shape == r.shape
            if db:
                assert np.abs(g - r).max() <= 1e-2
            else:
                n_cells += r.size
                n_ok += int(np.sum(np.abs(g - r) <= 1e-4 * np.abs(r) + 1e-30))
    if not db:
        assert n_ok >= 0.999 * n_cells


@pytest.mark.parametrize("n_fft,hop,n_mels,fmax,power,db,norm", [
    (2048, 281, 160, 11000, 2, True, True),       # a model whose metadata sets n_fft (src/identify_tracks.py:493)
    (1024, 256, 80, 11000, 2, True, False),
    (512, 128, 40, 8000, 1, False, True),
    (4096, 281, 160, 16000, 2, True, True),       # bands above bin 1023: outside the fused kernel's range
])
def test_general_path_other_frame_lengths_vs_oracle(engine, n_fft, hop, n_mels, fmax, power, db, norm):
    """Parameter sets the fused kernel does not cover go through segment rows -> framed-spectrum kernel with
    mel epilogue -> power_to_db (cac_plan general path) behind the same load_samples call."""
    from audio_analysis_b200 import identify_tracks as T
    rec = synth.recording(900 + n_fft // 64, seconds=6.0)
    tracks = [(0.0, 6.0, 0, 24000), (0.5, 1.4, 700, 2500), (4.6, 6.0, 1000, 4000)]
    kw = dict(hop_length=hop, mel_break=1750, n_mels=n_mels, fmin=50, fmax=fmax, power=power, db_scale=db,
              normalize=norm, n_fft=n_fft)
    np.random.seed(5)
    got = T.load_samples(rec, 48000, [T.Signal(*t) for t in tracks], 3.0, 1.5, **kw)
    np.random.seed(5)
    ref = O.load_samples(rec, 48000, [O.Signal(*t) for t in tracks], 3.0, 1.5, **kw)
    assert [len(g) for g in got] == [len(r) for r in ref]
    n_cells = n_ok = 0
    for g_track, r_track in zip(got, ref):
        for g, r in zip(g_track, r_track):
            r = np.asarray(r)
            assert g.shape == r.shape == (n_mels, 1 + 144000 // hop, 1)
            if db:
                assert np.abs(g - r).max() <= 1e-2
            else:
                n_cells += r.size
                n_ok += int(np.sum(np.abs(g - r) <= 1e-4 * np.abs(r) + 1e-30))
    if not db:
        assert n_ok >= 0.999 * n_cells


def test_unsupported_frame_length_fails_loudly(engine):
    from audio_analysis_b200 import identify_tracks as T
    rec = synth.recording(1, seconds=3.0)
    with pytest.raises(engine.CacError):
        T.load_samples(rec, 48000, [T.Signal(0.0, 3.0, 0, 24000)], 3.0, 1.5, n_fft=3000, **{k: v for k, v in S1.items() if k != "n_fft"})


def test_pcm16_host_path_equals_float_path(engine):
    """16-bit PCM as decoded goes up as int16 (half the PCIe bytes) and is widened on the device to sample / 32768:
    bit for bit the result of handing over the widened float32 array."""
    from audio_analysis_b200 import identify_tracks as T
    rec16 = np.round(synth.recording(5, seconds=7.5) * 32767.0).astype(np.int16)
    recf = rec16.astype(np.float32) / np.float32(32768.0)
    segs, _ = T.segment_descriptors(len(rec16), 48000, [T.Signal(0.0, 7.5, 0, 24000)], 3.0, 1.5)
    plan = T._plan(48000, 4096, 281, 160, 50, 11000, 1750, 144000, 2, True, True)
    a = plan.run_host(rec16, segs)
    b = plan.run_host(recf, segs)
    assert a.dtype == np.float32 and np.array_equal(a, b)
    kw = dict(segment_length=3.0, stride=1.5, db_scale=True, **S1)
    np.random.seed(3)
    x = T.load_samples(rec16, 48000, [T.Signal(0.0, 7.5, 0, 24000), T.Signal(1.0, 2.0, 500, 5000)], **kw)
    np.random.seed(3)
    y = T.load_samples(recf, 48000, [T.Signal(0.0, 7.5, 0, 24000), T.Signal(1.0, 2.0, 500, 5000)], **kw)
    assert [len(t) for t in x] == [len(t) for t in y]
    assert all(np.array_equal(p, q) for s, t in zip(x, y) for p, q in zip(s, t))

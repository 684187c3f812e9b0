"""Pins oracle/oracle_np.py (the numpy restatement that travels to the GPU box) against
fixtures produced by executing the reference's own files (tests/golden/make_golden.py)."""
import numpy as np
import pytest

from audio_analysis_b200 import synth
from oracle import oracle_np as O


def test_filterbanks_bit_exact(golden):
    g = golden("filterbanks")
    for name in ("s1", "get_end", "break1000", "sr16k"):
        a = g[name + "_args"]
        w = O.mel_f(int(a[0]), int(a[1]), a[2], a[3], int(a[4]), a[5])
        assert tuple(g[name + "_shape"]) == w.shape
        r, c = np.nonzero(w)
        assert np.array_equal(r, g[name + "_rows"]) and np.array_equal(c, g[name + "_cols"])
        assert np.array_equal(w[r, c], g[name + "_vals"])
    # structural known answers
    w = O.mel_f(48000, 160, 50, 11000, 4096, 1750)
    assert (w != 0).sum() == 1853
    cols = np.nonzero(w.any(axis=0))[0]
    assert cols.min() == 5 and cols.max() == 938


def test_get_spect_bit_exact(golden):
    g = golden("get_spect")
    rec = synth.recording(int(g["seed"]), seconds=12.0)
    s0 = int(g["seg_start"])
    x = O.normalize_data(rec[s0:s0 + 144000])
    assert x.dtype == np.float32
    assert np.array_equal(x[:4096], g["normalized_head"])
    a = O.get_spect(x, 48000, 281, False, False, 1750, True, 160, 50, 11000, 4096, 2, True)
    assert a.shape == (160, 513, 1) and a.dtype == np.float32
    assert np.array_equal(a[:, :, 0], g["s1_db"])
    b = O.get_spect(x, 48000, 640, False, False, 1750, True, 160, 50, 11000, 4096, 2, False)
    assert b.shape == (160, 226, 1)
    assert np.array_equal(b[:, :, 0], g["hop640_power"])
    c = O.get_spect(x, 48000, 281, True, False, 1750, True, 160, 50, 11000, 4096, 1, False, 3)
    assert c.shape == (160, 513, 3)
    assert np.array_equal(c[::4, ::4, :], g["power1_meansub_ch3"])


@pytest.mark.parametrize("case,kw", [("default", {}), ("pad_short", {"pad_short_tracks": True}),
                                     ("seg2_stride1", {"segment_length": 2.0, "stride": 1.0})])
def test_slicing_bit_exact(golden, case, kw):
    g = golden("slicing")
    tracks = [O.Signal(*row) for row in g["tracks"]]
    kw = dict(kw)
    seg_len = kw.pop("segment_length", 3.0)
    stride = kw.pop("stride", 1.5)
    np.random.seed(99)
    table = O.segment_table(12 * 48000, 48000, tracks, seg_len, stride, **kw)
    assert [len(t) for t in table] == list(g[case + "_counts"])
    rows = [(ti, a, n, p) for ti, segs in enumerate(table) for (a, n, p) in segs]
    ref = g[case]
    assert len(rows) == len(ref)
    for (ti, a, n, p), r in zip(rows, ref):
        assert ti == r[0] and n == r[2]
        if n:
            assert a == r[1] and p == r[3]
    assert (ref[:, 4] == int(48000 * seg_len)).all()


def test_load_samples_bit_exact(golden):
    g = golden("load_samples")
    rec = synth.recording(int(g["rec_seed"]), seconds=12.0)
    np.random.seed(int(g["seed"]))
    tr = [O.Signal(0.0, 12.0, 0, 24000), O.Signal(5.2, 6.0, 1000, 4000)]
    res = O.load_samples(rec, 48000, tr, 3.0, 1.5, hop_length=281, mel_break=1750, n_mels=160,
                         fmin=50, fmax=11000, power=2, db_scale=True, normalize=True, n_fft=4096)
    assert [len(t) for t in res] == list(g["counts"])
    assert np.array_equal(res[0][0][:, :, 0], g["t0_first"])
    assert np.array_equal(res[0][-1][:, :, 0], g["t0_last"])
    assert np.array_equal(res[1][0][:, :, 0], g["t1_only"])


def test_whole_minute_track_has_39_segments():
    table = O.segment_table(60 * 48000, 48000, [O.Signal(0, 60, 0, 24000)], 3, 1.5)
    assert len(table[0]) == 39
    assert all(n == 144000 and p == 0 for (_, n, p) in table[0])
    assert [a for (a, _, _) in table[0]][:3] == [0, 72000, 144000]


def test_get_end_and_signal_noise(golden):
    g = golden("get_end")
    rec = synth.recording(1234, seconds=12.0)
    tail = synth.recording(77, seconds=6.0, zero_tail_s=3.0)
    assert float(O.get_end(tail, 48000)) == float(g["end_tail"])
    assert float(O.get_end(rec, 48000)) == float(g["end_full"])
    sn = O.signal_noise_spectrogram(rec[:96000], 48000)
    assert tuple(g["signal_noise_shape"]) == sn.shape
    assert np.array_equal(sn[::8, ::8].astype(np.float32), g["signal_noise_sub"])


def test_morepork_bit_exact(golden):
    g = golden("morepork")
    rec = synth.recording(1234, seconds=12.0)
    s = O.morepork_spectrogram(rec, 48000)
    assert np.array_equal(s, g["spec48"])
    w, t = O.morepork_windows(s)
    assert np.array_equal(w[:, :, :, 0], g["win48"]) and np.array_equal(np.array(t), g["starts48"])
    rec16 = synth.recording(55, seconds=8.0, sr=16000)
    s = O.morepork_spectrogram(rec16, 16000)
    assert np.array_equal(s, g["spec16"])
    w, t = O.morepork_windows(s)
    assert np.array_equal(w[:, :, :, 0], g["win16"]) and np.array_equal(np.array(t), g["starts16"])


def test_morepork_minute_has_61_windows():
    s = np.abs(np.random.default_rng(0).standard_normal((60, 1201))).astype(np.float32)
    w, t = O.morepork_windows(s)
    assert w.shape == (61, 60, 60, 1) and t[0] == 0.0 and t[-1] == (1201 - 60) / 20


def test_cacophony_index_bit_exact(golden):
    g = golden("cacophony_index")
    x = synth.recording(int(g["seed"]), seconds=60.0, sr=16000)
    assert list(O.ci_edges()) == [25, 38, 60, 93, 145, 226, 351, 546, 848, 1318, 2048]
    for i in range(8):
        off = 1024 * (i + 1)
        assert np.array_equal(O.get_ci_bins(x[off:off + 2048], 16000), g["bins_first"][i])
    pts = O.ci_points(x)
    assert len(pts) == 933 and np.array_equal(np.array(pts), g["points"])
    assert np.array_equal(np.array(O.ci_points_fast(x)), g["points"])
    res = O.ci_table(pts, x.shape[0])
    tab = np.array([[e["begin_s"], e["end_s"], e["index_percent"]] for e in res["cacophony_index_old"]])
    assert np.array_equal(tab, g["table"]) and len(tab) == 3
    assert res["cacophony_index_old_version"] == str(g["version"])
    assert O.score_from_points(pts[:312]) == g["scores"][0]
    short = O.calculate_from_array(synth.recording(5, seconds=10.0, sr=16000))
    assert short["cacophony_index_old"] == [] and short["ci_warning"] == str(g["short_warning"])


def test_magtransform(golden):
    g = golden("magtransform")
    assert np.allclose(O.magtransform(g["x"], 0.0), g["v1"], rtol=1e-6)
    assert np.allclose(O.magtransform(g["x"], -1.0), g["v2"], rtol=1e-6)


def test_stft_standin_matches_torch():
    """Cross-check of the librosa.stft restatement against torch.stft (float32 FFT)."""
    import torch
    x = synth.recording(9, seconds=3.0)
    ref = np.abs(O.stft(x, 4096, 281))
    t = torch.stft(torch.from_numpy(x), 4096, hop_length=281, window=torch.hann_window(4096, periodic=True),
                   center=True, pad_mode="constant", return_complex=True).abs().numpy()
    assert ref.shape == t.shape == (2049, 513)
    assert np.abs(ref - t).max() < 2e-3 * max(1.0, ref.max() / 100)


def test_slaney_branch_matches_reference_under_stub(golden):
    g = golden("get_spect_slaney")
    rec = synth.recording(int(g["seed"]), seconds=12.0)
    s0 = int(g["seg_start"])
    x = O.normalize_data(rec[s0:s0 + 144000])
    mine = O.get_spect(x, 48000, 281, False, False, 1750, False, 160, 50, 11000, 4096, 2, True)
    assert np.abs(mine[:, :, 0] - g["s1_db"]).max() <= 1e-4
    from audio_analysis_b200 import custommel
    assert np.array_equal(custommel.slaney_mel_f(48000, 4096, 160, 50, 11000), O.slaney_mel_filterbank(48000, 4096, 160, 50, 11000))


def test_signal_noise_oracle_equals_reference_fixture(golden):
    g = golden("signal_noise")
    for name in "abc":
        rec = synth.recording(int(g[name + "_seed"]), seconds=float(g[name + "_seconds"]))
        sig = O.signal_noise(rec, 48000)
        assert np.array_equal(np.array(sig, dtype=np.float64).reshape(-1, 4), g[name + "_signals"])


def test_librosa_stand_ins_against_torchaudio():
    """Independent cross-check of the restated librosa layer (librosa itself is not installable here): torchaudio ships its
    own Slaney mel filterbank, dB conversion and STFT.  Agreement is to float32 rounding, not bit-exact (different
    operation order), which is what pins the SEMANTICS: periodic Hann, centre zero padding, 1 + len // hop frames;
    Slaney scale and area normalisation; amin / ref=max / top_db clamp."""
    import os
    import sys
    import torch
    ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    pytest.importorskip("torchaudio")
    import torchaudio.functional as F
    stubs = os.path.join(ROOT, "oracle", "stubs")
    sys.path.insert(0, stubs)
    saved = {k: sys.modules.pop(k) for k in list(sys.modules) if k == "librosa" or k.startswith("librosa.")}
    try:
        import librosa
        from librosa import feature
        # filters.mel (htk=False branch of get_spect, src/identify_tracks.py:229-238)
        w = feature.filters_mel(48000, 4096, n_mels=160, fmin=50, fmax=11000)
        fb = F.melscale_fbanks(n_freqs=2049, f_min=50.0, f_max=11000.0, n_mels=160, sample_rate=48000, norm="slaney",
                               mel_scale="slaney").numpy().T
        # torchaudio builds the triangles in float32, librosa in float64: 2e-7 absolute on weights of up to 0.049
        assert w.shape == fb.shape and np.abs(w - fb).max() <= 1e-6
        assert np.array_equal(w.argmax(axis=1), fb.argmax(axis=1))
        # power_to_db(ref=np.max) (src/identify_tracks.py:265-266)
        x = (np.abs(np.random.default_rng(0).standard_normal((64, 100))).astype(np.float32) ** 2) * 1e-3
        x[3, 5] = 0.0
        db = F.amplitude_to_DB(torch.from_numpy(x), multiplier=10.0, amin=1e-10,
                               db_multiplier=float(np.log10(max(1e-10, x.max()))), top_db=80.0).numpy()
        assert np.abs(db - librosa.power_to_db(x, ref=np.max)).max() <= 1e-4
        # amplitude_to_db(ref=np.max) (src/identify_morepork.py:152): 20 log10 with amin 1e-5
        a = np.sqrt(x)
        dba = F.amplitude_to_DB(torch.from_numpy(a), multiplier=20.0, amin=1e-5,
                                db_multiplier=float(np.log10(max(1e-5, a.max()))), top_db=80.0).numpy()
        assert np.abs(dba - librosa.amplitude_to_db(a, ref=np.max)).max() <= 1e-4
        # stft framing (periodic Hann, centre, zero padding) against torchaudio's Spectrogram
        rec = synth.recording(3, seconds=1.0)
        for n_fft, hop in ((4096, 281), (4800, 2400)):
            ref = np.abs(librosa.stft(rec, n_fft=n_fft, hop_length=hop))
            got = F.spectrogram(torch.from_numpy(rec).double(), pad=0, window=torch.hann_window(n_fft, periodic=True, dtype=torch.float64),
                                n_fft=n_fft, hop_length=hop, win_length=n_fft, power=1.0, normalized=False, center=True,
                                pad_mode="constant").numpy()
            assert got.shape == ref.shape == (1 + n_fft // 2, 1 + len(rec) // hop)
            assert np.abs(got - ref).max() <= 1e-5 * ref.max()
    finally:
        sys.path.remove(stubs)
        for k in [k for k in sys.modules if k == "librosa" or k.startswith("librosa.")]:
            sys.modules.pop(k)
        sys.modules.update(saved)

"""GPU parity of signal_noise (src/identify_tracks.py:650-706) through the C-ABI.
* the post-STFT chain (medians, 3x threshold, cv2 morphology, connected components) is bit-exact when it
  is fed the oracle's spectrogram;
* end to end (our float32 STFT) the threshold mask may differ only in cells within 1e-6 (of the
  normalised scale) of a threshold, and the returned signals equal the reference's own output."""
import numpy as np
import pytest

from audio_analysis_b200 import synth
from oracle import oracle_np as O

pytestmark = pytest.mark.gpu
CASES = [("a", 1234, 12.0), ("b", 77, 20.0), ("c", 4242, 6.5)]


def _stages_from_oracle_spec(engine, rec):
    import torch
    spec = O.signal_noise_spectrogram(rec, 48000).astype(np.float32)
    stats, masks = engine.signal_components(torch.from_numpy(np.ascontiguousarray(spec)).cuda(), 10, 42, 1, want_masks=True)
    return spec, stats, masks.cpu().numpy()


@pytest.mark.parametrize("name,seed,seconds", CASES)
def test_post_stft_chain_bit_exact_on_oracle_spectrogram(engine, name, seed, seconds):
    import cv2
    rec = synth.recording(seed, seconds=seconds)
    sig, mask, opened, closed, ref_stats = O.signal_noise_stages(rec, 48000)
    spec, stats, masks = _stages_from_oracle_spec(engine, rec)
    assert np.array_equal(masks[0], mask)
    assert np.array_equal(masks[1], opened)
    assert np.array_equal(masks[2], closed)
    _, _, cv_stats, _ = cv2.connectedComponentsWithStats(closed)
    order = np.argsort(stats[:, 5], kind="stable")
    assert np.array_equal(stats[order][:, :5], cv_stats[1:])          # same components in cv2's label order


def test_row_medians_exact(engine):
    import torch
    rng = np.random.default_rng(0)
    for shape in ((7, 10250), (5, 2049), (3, 2), (4, 1), (6, 1000)):
        x = np.abs(rng.standard_normal(shape)).astype(np.float32)
        x[0, : shape[1] // 2] = 0.25                                   # heavy ties
        t = torch.from_numpy(x).cuda()
        stats, masks = engine.signal_components(t, 1, 1, 1, want_masks=True)
        ref = (x > 3 * np.median(x, axis=0)[None, :]) & (x > 3 * np.median(x, axis=1)[:, None])
        assert np.array_equal(masks[0].cpu().numpy(), ref.astype(np.uint8)), shape


@pytest.mark.parametrize("shape", [(9, 10252), (3, 4096), (3, 8192), (2, 20000), (2, 24576), (5, 1000), (4, 4), (3, 12288)])
def test_row_medians_exact_on_16_byte_aligned_rows(engine, shape):
    """Matrices whose rows are whole 16-byte groups take sn_rowmed_mask_vec_kernel (one histogram pass, candidates ranked
    in shared memory) - the kernel the fused chain uses.  Rows built to reach each of its branches: plain noise, a wide
    dynamic range, ties by the thousand around the median (more candidates than threads: histogram passes), a constant
    row, a row of zeros, keys that differ only in their low 12 bits (no second digit), two-valued rows whose middle
    values sit in different bins, and a row with a handful of duplicates of the median."""
    import torch
    rng = np.random.default_rng(shape[1])
    rows, cols = shape
    x = np.abs(rng.standard_normal(shape)).astype(np.float32)
    x[0, : cols // 2] = 0.25
    if rows > 1:
        x[1] = np.exp(5.0 * rng.standard_normal(cols)).astype(np.float32)
    if rows > 2:
        x[2] = 0.5
    if rows > 3:
        x[3] = 0.0
    if rows > 4:
        x[4] = (np.uint32(0x3f000000) + rng.integers(0, 4096, cols).astype(np.uint32)).view(np.float32)
    if rows > 5:
        x[5] = np.where(np.arange(cols) % 2 == 0, np.float32(0.001), np.float32(7.0))
    if rows > 6:
        x[6, :5] = np.sort(x[6])[cols // 2]
    if rows > 7:
        x[7] = np.round(x[7] * 8) / 8                                  # ~30 distinct values
    t = torch.from_numpy(x).cuda()
    assert t.data_ptr() % 16 == 0
    stats, masks = engine.signal_components(t, 1, 1, 1, want_masks=True)
    ref = (x > 3 * np.median(x, axis=0)[None, :]) & (x > 3 * np.median(x, axis=1)[:, None])
    got = masks[0].cpu().numpy()
    bad = np.argwhere(got != ref.astype(np.uint8))
    assert len(bad) == 0, (shape, bad[:5])


@pytest.mark.parametrize("shape,dh,dw,eh", [((60001, 8), 10, 3, 1), ((37, 700), 7, 70, 2), ((300, 33), 1, 1, 0),
                                               ((50, 300), 3, 64, 1), ((50, 300), 2, 63, 2), ((40, 200), 6, 33, 3)])
def test_morphology_on_other_shapes_follows_cv2(engine, shape, dh, dw, eh):
    """The strip kernels on matrices unlike the fused chain's: 60,001 rows (a last strip of one row), a structuring
    element wider than 64 pixels, rows shorter than two words, erode height 0 (3 x 3)."""
    import cv2
    import torch
    rng = np.random.default_rng(shape[0])
    x = np.abs(rng.standard_normal(shape)).astype(np.float32)
    blocks = rng.random(((shape[0] + 5) // 6, (shape[1] + 4) // 5)) < 0.3      # 6 x 5 blobs survive the 4 x 4 opening
    x[np.kron(blocks, np.ones((6, 5), bool))[: shape[0], : shape[1]]] += 40.0
    _, masks = engine.signal_components(torch.from_numpy(x).cuda(), dh, dw, eh, max_components=1 << 20, want_masks=True)
    masks = masks.cpu().numpy()
    opened = cv2.morphologyEx(masks[0], cv2.MORPH_OPEN, np.ones((4, 4), np.uint8))
    closed = cv2.dilate(opened, np.ones((dh, dw), np.uint8))
    closed = cv2.erode(closed, np.ones((eh, dw) if eh else (3, 3), np.uint8))
    assert masks[0].any() and opened.any()
    assert np.array_equal(masks[1], opened)
    assert np.array_equal(masks[2], closed)


@pytest.mark.parametrize("name,seed,seconds", CASES)
def test_signal_noise_equals_reference_fixture(engine, golden, name, seed, seconds):
    from audio_analysis_b200 import identify_tracks as T
    g = golden("signal_noise")
    rec = synth.recording(seed, seconds=seconds)
    got = T.signal_noise(rec, 48000)
    ref = g[name + "_signals"]
    assert len(got) == len(ref)
    for s, r in zip(got, ref):
        assert (s.start, s.end, s.freq_start, s.freq_end) == tuple(r)


def test_threshold_mask_differs_only_near_thresholds(engine):
    from audio_analysis_b200 import identify_tracks as T
    rec = synth.recording(1234, seconds=12.0)
    spec = T.signal_noise_spectrogram(rec, 48000).contiguous()
    _, masks = engine.signal_components(spec, 10, 42, 1, want_masks=True)
    ref_spec = O.signal_noise_spectrogram(rec, 48000)
    _, mask, _, _, _ = O.signal_noise_stages(rec, 48000)
    got = masks[0].cpu().numpy()
    bad = got != mask
    tc = 3 * np.median(ref_spec, axis=0)[None, :]
    tr = 3 * np.median(ref_spec, axis=1)[:, None]
    near = (np.abs(ref_spec - tc) <= 1e-6) | (np.abs(ref_spec - tr) <= 1e-6)
    print("mask cells differing: %d of %d, all near a threshold: %s" % (bad.sum(), bad.size, bool((~bad | near).all())))
    assert (~bad | near).all()
    assert bad.mean() <= 1e-4


def test_silent_recording_has_no_signals(engine):
    from audio_analysis_b200 import identify_tracks as T
    assert T.signal_noise(np.zeros(48000 * 2, np.float32), 48000) == []


def test_frames_uploaded_once_give_the_same_results(engine):
    """to_device(frames): get_end / signal_noise / load_samples on the CUDA tensor = the calls on the numpy array."""
    from audio_analysis_b200 import identify_tracks as T
    rec = synth.recording(31, seconds=9.0, zero_tail_s=2.0)
    dev = T.to_device(rec)
    assert T.get_end(dev, 48000) == T.get_end(rec, 48000)
    a, b = T.signal_noise(dev, 48000), T.signal_noise(rec, 48000)
    assert [(s.start, s.end, s.freq_start, s.freq_end) for s in a] == [(s.start, s.end, s.freq_start, s.freq_end) for s in b]
    kw = dict(hop_length=281, mel_break=1750, n_mels=160, fmin=50, fmax=11000, power=2, db_scale=True, n_fft=4096)
    tracks = [T.Signal(0.0, 9.0, 0, 24000), T.Signal(1.0, 2.2, 500, 5000)]
    np.random.seed(2)
    x = T.load_samples(dev, 48000, tracks, 3.0, 1.5, **kw)
    np.random.seed(2)
    y = T.load_samples(rec, 48000, tracks, 3.0, 1.5, **kw)
    assert [len(t) for t in x] == [len(t) for t in y]
    assert all(np.array_equal(p, q) for s, t in zip(x, y) for p, q in zip(s, t))


def test_signal_noise_batch_equals_per_recording_calls(engine):
    """signal_noise_batch: one cac_sn_run call for all recordings (ragged lengths), one synchronisation."""
    import time
    import torch
    from audio_analysis_b200 import identify_tracks as T
    recs = [synth.recording(70 + i, seconds=s) for i, s in enumerate((8.0, 11.5, 6.0))]
    key = lambda sigs: [(s.start, s.end, s.freq_start, s.freq_end) for s in sigs]
    single = [key(T.signal_noise(r, 48000)) for r in recs]
    batch = T.signal_noise_batch(recs, 48000)
    assert [key(b) for b in batch] == single
    assert T.signal_noise_batch([], 48000) == []
    # throughput note (printed with -s): 32 one-minute recordings already on the device (two waves of 24 slots)
    pool = torch.from_numpy(synth.batch(500, 32, seconds=60.0)).cuda()
    T.signal_noise_batch(pool[:2].contiguous(), 48000)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = T.signal_noise_batch(pool, 48000)
    dt = time.perf_counter() - t0
    print("signal_noise_batch: %.3f ms per one-minute recording (32 per call)" % (dt / 32 * 1e3))
    assert len(out) == 32
    for i in (3, 23, 24, 31):                                      # both waves, first and last slots
        assert key(out[i]) == key(T.signal_noise(pool[i], 48000))


def test_spectrum_stage_medians_and_maximum_are_exact(engine):
    """cac_sn_spectrum_run: the column medians selected inside the spectrum kernel equal np.median of the matrix the
    same kernel wrote, a_max equals its maximum, and the matrix matches the oracle's float64 STFT."""
    from audio_analysis_b200 import identify_tracks as T
    for seed, seconds in ((1234, 12.0), (5, 1.3), (6, 0.05)):
        rec = synth.recording(seed, seconds=seconds)
        plan = T._signal_noise_plan(48000, 281)
        spec, colmed, amax = plan.spectrum(T.to_device(rec))
        spec, colmed, amax = spec.cpu().numpy(), colmed.cpu().numpy(), float(amax.cpu().numpy()[0])
        assert spec.shape == (2049, 1 + len(rec) // 281)
        assert np.array_equal(colmed, np.median(spec, axis=0))
        assert amax == spec.max()
        ref = O.signal_noise_spectrogram(rec, 48000)               # normalised by its own maximum
        assert np.abs(spec / np.float32(amax) - ref).max() <= 2e-6


def test_batch_chain_equals_standalone_chain_on_its_own_spectrum(engine):
    """cac_sn_run (medians from inside the kernels, unscaled matrix) == cac_signal_components_run fed
    spectrum / a_max (the reference's order of operations), component for component."""
    import torch
    from audio_analysis_b200 import identify_tracks as T
    rec = synth.recording(99, seconds=14.0)
    plan = T._signal_noise_plan(48000, 281)
    dev = T.to_device(rec)
    got = engine.collect_components(plan.run(dev, [0], [len(rec)], max_components=4096).cpu().numpy()[0])
    spec = T.signal_noise_spectrogram(dev, 48000)
    ref = engine.signal_components(spec, plan.height, plan.width, plan.height // 10)
    srt = lambda a: a[np.argsort(a[:, 5], kind="stable")]
    assert len(got) == len(ref) and np.array_equal(srt(got), srt(ref))


def test_component_overflow_falls_back_per_recording(engine):
    """A recording with more components than max_components is redone alone; the others of the batch are kept."""
    from audio_analysis_b200 import identify_tracks as T
    recs = [synth.recording(70, seconds=8.0), synth.recording(71, seconds=11.5)]
    key = lambda sigs: [(s.start, s.end, s.freq_start, s.freq_end) for s in sigs]
    full = T.signal_noise_batch(recs, 48000)
    tiny = T.signal_noise_batch(recs, 48000, max_components=1)
    assert [key(a) for a in tiny] == [key(b) for b in full]


def test_other_sample_rates_follow_cv2(engine):
    """sr = 16 kHz: height 27 -> erode (2, 13); sr = 96 kHz: height 6 -> erode height 0 = cv2's empty kernel (3 x 3)."""
    from audio_analysis_b200 import identify_tracks as T
    for sr in (16000, 96000):
        rec = synth.recording(12, seconds=6.0, sr=sr)
        ref = O.signal_noise(rec, sr)                             # [(start, end, freq_start, freq_end)]
        got = T.signal_noise(rec, sr)
        assert [(s.start, s.end, s.freq_start, s.freq_end) for s in got] == [tuple(r) for r in ref]


def test_get_end_batch_equals_get_end(engine):
    from audio_analysis_b200 import identify_tracks as T
    recs = [synth.recording(31, seconds=9.0, zero_tail_s=2.0), synth.recording(32, seconds=7.5),
            synth.recording(33, seconds=12.0, zero_tail_s=5.5), np.zeros(48000 * 3, np.float32)]
    want = [O.get_end(r, 48000) for r in recs]
    assert T.get_end_batch(recs, 48000) == want
    assert [T.get_end(r, 48000) for r in recs] == want
    assert T.get_end_batch([], 48000) == []


def test_get_end_probe_and_full_evaluation_agree_with_the_reference_loop(engine):
    """cac_get_end_run transforms ONE frame per chunk first and only the chunks whose first frame is constant in full.
    Cases that separate the two phases: a silent gap that covers a whole chunk in the MIDDLE of a recording (the reference
    returns there), a gap that covers a chunk's first frame but not the whole chunk (candidate after the probe, rejected
    by the full evaluation), silence that starts inside a chunk, a recording of silence only, one shorter than a chunk."""
    from audio_analysis_b200 import identify_tracks as T
    sr, hop, L = 48000, 281, 48000 // 281                 # 170 frames per chunk = 47,770 samples
    base = [synth.recording(40 + i, seconds=9.0) for i in range(4)]
    a = base[0].copy(); a[3 * L * hop - 3000: 5 * L * hop + 3000] = 0.0           # chunks 3 and 4 fully silent (plus margins)
    b = base[1].copy(); b[2 * L * hop - 3000: 2 * L * hop + 9000] = 0.0           # only the first frames of chunk 2
    c = base[2].copy(); c[4 * L * hop + 20000:] = 0.0                             # silence starts inside chunk 4
    d = base[3].copy(); d[: 2 * L * hop + 4000] = 0.0                             # silent from the start for two chunks
    recs = [a, b, c, d, np.zeros(sr * 4, np.float32), synth.recording(44, seconds=0.9), base[0]]
    want = [O.get_end(r, sr) for r in recs]
    assert want[0] == 3 * L * hop // sr and want[1] == len(b) / sr and want[3] == 0
    assert T.get_end_batch(recs, sr) == want
    assert [T.get_end(r, sr) for r in recs] == want

"""Out-of-bounds WRITE check of every device buffer the host layer hands to the C-ABI (outputs, work buffers, item
tables), without compute-sanitizer (closed on this GPU pool: profiles/r3n_sanitizer.md).

torch.empty / torch.zeros / torch.full are patched for the duration of a test: every CUDA allocation gets a 4 KB band
of a known byte pattern on either side.  The public calls of each path are run on small ragged inputs, their results
are compared with an unguarded run (different base addresses, same values), and every band must still hold the pattern
afterwards.  Reads out of bounds are not seen by this; the parity tests with ragged tails are what covers those."""
import numpy as np
import pytest

from audio_analysis_b200 import synth

pytestmark = pytest.mark.gpu
GUARD = 4096
PATTERN = 0xA5
S1 = dict(hop_length=281, mel_break=1750, n_mels=160, fmin=50, fmax=11000, power=2, db_scale=True, normalize=True,
          n_fft=4096)


class Guards:
    def __init__(self, monkeypatch):
        import torch
        self.torch = torch
        self.held = []                                  # (base uint8 tensor, payload bytes)
        self._empty, self._zeros, self._full = torch.empty, torch.zeros, torch.full
        monkeypatch.setattr(torch, "empty", self._wrap(self._empty, None))
        monkeypatch.setattr(torch, "zeros", self._wrap(self._zeros, 0))
        monkeypatch.setattr(torch, "full", self._full_wrapped)

    def _is_cuda(self, device):
        return device is not None and self.torch.device(device).type == "cuda"

    def _alloc(self, size, dtype, device, fill):
        torch = self.torch
        dtype = dtype or torch.float32
        size = tuple(size[0]) if len(size) == 1 and isinstance(size[0], (tuple, list, torch.Size)) else tuple(size)
        numel = int(np.prod(size)) if size else 1
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        padded = (nbytes + 255) // 256 * 256
        base = self._empty(padded + 2 * GUARD, dtype=torch.uint8, device=device)
        base[:GUARD] = PATTERN
        base[GUARD + nbytes:] = PATTERN
        self.held.append((base, nbytes))
        t = base[GUARD:GUARD + nbytes].view(dtype).view(size)
        if fill is not None:
            t.fill_(fill)
        return t

    def _wrap(self, orig, fill):
        def f(*size, dtype=None, device=None, **kw):
            if "size" in kw:
                size = (kw.pop("size"),)
            if not self._is_cuda(device) or kw.get("pin_memory"):
                return orig(*size, dtype=dtype, device=device, **kw)
            return self._alloc(size, dtype, device, fill)
        return f

    def _full_wrapped(self, size, fill_value, dtype=None, device=None, **kw):
        if not self._is_cuda(device):
            return self._full(size, fill_value, dtype=dtype, device=device, **kw)
        if dtype is None:
            dtype = self.torch.float32 if isinstance(fill_value, float) else self.torch.int64
        return self._alloc((size,) if isinstance(size, int) else (tuple(size),), dtype, device, fill_value)

    def check(self):
        self.torch.cuda.synchronize()
        assert self.held, "no CUDA allocation went through the guarded allocator"
        for i, (base, nbytes) in enumerate(self.held):
            lo = base[:GUARD]
            hi = base[GUARD + nbytes:]
            assert bool((lo == PATTERN).all()), "buffer %d (%d bytes): write BELOW the buffer" % (i, nbytes)
            assert bool((hi == PATTERN).all()), "buffer %d (%d bytes): write BEYOND the buffer" % (i, nbytes)
        return len(self.held)


def _clear_plan_caches():
    """Cached plans hold work buffers allocated earlier in the process: drop them so they are re-made under guard."""
    from audio_analysis_b200 import cacophony_index as CI, identify_morepork as M, identify_tracks as T, resample as R, custommel
    for f in (CI._plan, M._plan, T._plan, T._get_end_plan, T._signal_noise_plan, R._plan, custommel._projection_plan):
        f.cache_clear()


def _recs():
    return [synth.recording(61, seconds=12.7), synth.recording(62, seconds=8.1, zero_tail_s=2.0),
            synth.recording(63, seconds=5.03), synth.recording(64, seconds=3.0)]


def _box(s):
    return (s.start, s.end, s.freq_start, s.freq_end)


@pytest.fixture
def guards(engine, monkeypatch):
    _clear_plan_caches()
    g = Guards(monkeypatch)
    yield g
    monkeypatch.undo()
    _clear_plan_caches()


def test_guard_bands_classify_prep_chain(engine, monkeypatch):
    """get_end_batch, signal_noise_batch (incl. the overflow re-run), classify_prep_batch, load_samples."""
    from audio_analysis_b200 import identify_tracks as T
    recs = _recs()
    np.random.seed(3)
    plain = T.classify_prep_batch(recs, 48000, 3.0, 1.5, **S1)
    plain_sig = T.signal_noise_batch(recs, 48000, max_components=8)         # forces the overflow fallback
    plain_end = T.get_end_batch(recs, 48000)
    _clear_plan_caches()
    g = Guards(monkeypatch)
    np.random.seed(3)
    got = T.classify_prep_batch(recs, 48000, 3.0, 1.5, **S1)
    got_sig = T.signal_noise_batch(recs, 48000, max_components=8)
    got_end = T.get_end_batch(recs, 48000)
    n = g.check()
    monkeypatch.undo()
    _clear_plan_caches()
    assert n >= 6
    assert list(got_end) == list(plain_end)
    for a, b in zip(got_sig, plain_sig):
        assert [_box(s) for s in a] == [_box(s) for s in b]
    for a, b in zip(got, plain):
        assert a["length"] == b["length"] and [_box(t) for t in a["tracks"]] == [_box(t) for t in b["tracks"]]
        if b["samples"] is None:
            assert a["samples"] is None
        else:
            assert all(bool((x == y).all()) for x, y in zip(a["samples"], b["samples"]))


def test_guard_bands_features_general_path_and_standalone_ops(engine, guards):
    """load_samples on the fused kernel and on the general (n_fft 2048) path, MagTransform epilogue, the standalone ops."""
    import torch
    from audio_analysis_b200 import custommel, identify_tracks as T
    rec = synth.recording(65, seconds=9.4)
    tracks = [T.Signal(0.4, 8.9, 200, 9000), T.Signal(2.0, 3.1, 50, 11000)]
    T.load_samples(rec, 48000, tracks, 3.0, 1.5, **S1)
    T.load_samples(rec, 48000, tracks, 3.0, 1.5, **dict(S1, n_fft=2048, hop_length=140, n_mels=96))
    T.load_samples(rec, 48000, tracks, 2.0, 1.0, **dict(S1, power=1, db_scale=False))
    x = torch.from_numpy(np.abs(np.random.default_rng(1).normal(size=(3, 77, 41))).astype(np.float32)).cuda()
    engine.power_to_db(x)
    T.normalize_data(rec[:12345])
    mag = np.abs(np.random.default_rng(2).normal(size=(2049, 37))).astype(np.float32)
    custommel.mel_spec(mag, 48000, 4096, 281, 160, 50, 11000)
    assert guards.check() >= 1


def test_guard_bands_morepork_ci_resample(engine, guards):
    """morepork spectrogram + windows, Cacophony Index (batch and streamed), polyphase resampler."""
    import torch
    from audio_analysis_b200 import cacophony_index as CI, identify_morepork as M, resample as R
    recs = [synth.recording(71, seconds=11.0), synth.recording(72, seconds=6.3)]
    for r in recs:
        M.samples_from_recording((r, 48000))
    low = [R.resample(r, 48000, 16000) for r in recs]
    R.resample(recs[1][:30011], 22050, 48000)
    pcm = np.concatenate(low)
    offs = np.array([0, len(low[0])], np.int64)
    lens = [len(low[0]), len(low[1])]
    ref = CI.calculate_batch(torch.from_numpy(pcm).cuda(), offs, lens)
    s = CI.Stream()
    for lo in range(0, len(low[0]), 50000):
        s.feed(low[0][lo:lo + 50000])
    assert s.result() == ref[0]
    assert guards.check() >= 1

"""The tensor-core band DFT (csrc/band_gemm.cu: tcgen05.mma kind::tf32 with a hi / lo split, behind cac_spec_run for
raw-band plans with hop = n_fft / 2 - identify_morepork's spectrogram, src/identify_morepork.py:19-27) against a float64
band DFT and against the mixed-radix FFT kernel it replaces on that path.

Tolerances: <= 1e-4 relative on >= 90 % of the cells as in tests/test_spec_gpu.py (measured: 99.99-100 %), and a maximum
absolute error of 2e-6 of the spectrum maximum on the multi-recording fixtures, 4e-6 where stated: the three-product TF32
split (hi.hi + hi.lo + lo.hi, lo itself rounded to 11 bits) has a floor of 1-3e-6 of the maximum depending on how much of
the frame's energy lies outside the band (measured 0.8e-6 .. 2.7e-6); BASELINE's bar for this path is 1e-2 dB = 1.15e-3."""
import os

import numpy as np
import pytest

from audio_analysis_b200 import synth

pytestmark = pytest.mark.gpu
SR = 48000


def _reference(rec, n_fft=4800, lo=60, hi=120):
    """librosa.stft(n_fft, hop = n_fft / 2, centre = True, periodic Hann) restricted to bins [lo, hi), in float64"""
    hop = n_fft // 2
    w = 0.5 - 0.5 * np.cos(2 * np.pi * np.arange(n_fft) / n_fft)
    E = np.exp(-2j * np.pi * np.outer(np.arange(n_fft), np.arange(lo, hi)) / n_fft)
    x = np.concatenate([np.zeros(n_fft // 2), rec.astype(np.float64), np.zeros(n_fft // 2)])
    fr = np.stack([x[f * hop:f * hop + n_fft] for f in range(1 + len(rec) // hop)]) * w
    return np.abs(fr @ E)


def _pack(recs, align):
    offs, pos = [], 0
    for r in recs:
        offs.append(pos)
        pos += (len(r) + align - 1) // align * align
    flat = np.zeros(pos + 1, np.float32)
    for o, r in zip(offs, recs):
        flat[o:o + len(r)] = r
    return flat, np.array(offs, np.int64), np.array([len(r) for r in recs], np.int64)


def _run(plan, flat, offs, lens, gemm, shift=0):
    import torch
    old = os.environ.get("CAC_BAND_GEMM")
    os.environ["CAC_BAND_GEMM"] = "1" if gemm else "0"
    try:
        dev = torch.from_numpy(np.concatenate([np.zeros(shift, np.float32), flat])).cuda()
        spec, base, mx = plan.run(dev[shift:] if shift else dev, offs, lens, want_max=True)
        torch.cuda.synchronize()
        return spec.cpu().numpy(), base, float(mx.item())
    finally:
        if old is None:
            del os.environ["CAC_BAND_GEMM"]
        else:
            os.environ["CAC_BAND_GEMM"] = old


def _uniform(seed, n_rec, chunks, hop, sr=SR):
    n = chunks * hop
    recs = [synth.recording(seed + i, seconds=n / sr + 0.01, sr=sr)[:n] for i in range(n_rec)]
    return recs, np.concatenate(recs), np.arange(n_rec, dtype=np.int64) * n, np.full(n_rec, n, np.int64)


@pytest.mark.parametrize("sr,n_rec,chunks", [(48000, 5, 137), (48000, 1, 300), (32000, 3, 53), (16000, 3, 53), (8000, 2, 41)])
def test_band_gemm_uniform_batch_tma_form(engine, sr, n_rec, chunks):
    """Equally long recordings of a whole number of hops, back to back (identify_morepork.BatchPlan's layout): the A operand
    comes through one TMA tensor map over the buffer as a [chunks x hop] matrix.  5 x 137 chunks = 685 rows = 2 full
    256-row tiles + a partial one, every item boundary inside a tile.  Checked against float64, the FFT kernel and the
    cp.async form of the same GEMM (the same recordings with a 4-sample gap between them)."""
    from audio_analysis_b200 import identify_morepork as M
    hop = sr // 20
    recs, flat, offs, ns = _uniform(300, n_rec, chunks, hop, sr)
    plan = M._plan(sr)
    got, base, mx = _run(plan, flat, offs, ns, gemm=True)
    other, base2, mx2 = _run(plan, flat, offs, ns, gemm=False)
    assert list(base) == list(base2) and got.shape == other.shape == (n_rec * (chunks + 1) * 60,)
    assert not np.array_equal(got, other)                            # two different kernels ran
    ref = np.concatenate([_reference(r, sr // 10) for r in recs]).reshape(-1)
    tol = 4e-6                               # the 3-product TF32 split's floor is 1-3e-6 of the maximum (module docstring)
    for name, x in (("tensor-core GEMM, TMA form", got), ("other kernel", other)):
        d = np.abs(x - ref)
        print("%d Hz %s: max abs error %.2e of the maximum, <= 1e-4 relative on %.4f of the cells" %
              (sr, name, d.max() / ref.max(), float(np.mean(d <= 1e-4 * np.abs(ref)))))
        assert d.max() <= tol * ref.max() and float(np.mean(d <= 1e-4 * np.abs(ref))) >= 0.9
    assert mx == got.max()
    flat2 = np.zeros(len(flat) + 4 * n_rec + 4, np.float32)
    offs2 = offs + 4 * np.arange(n_rec)
    for o, r in zip(offs2, recs):
        flat2[o:o + len(r)] = r
    cpa, _, _ = _run(plan, flat2, offs2, ns, gemm=True)              # not back to back: the cp.async form
    assert np.abs(cpa - got).max() <= tol * ref.max() and np.abs(cpa - ref).max() <= tol * ref.max()


def test_band_gemm_ragged_batch_vs_float64_and_fft_kernel(engine):
    """Ragged recordings - shorter than one chunk, an exact multiple of the hop, a partial last chunk, a long one that
    spans several 256-row tiles - through both paths."""
    from audio_analysis_b200 import identify_morepork as M
    lens = [int(SR * 3.3) + 7, SR * 60, 1200, 2400 * 5, 4801, SR * 11 + 1234, 1]
    recs = [synth.recording(300 + i, seconds=n / SR + 0.01)[:n] for i, n in enumerate(lens)]
    flat, offs, ns = _pack(recs, 4)
    plan = M._plan(SR)
    got, base, mx = _run(plan, flat, offs, ns, gemm=True)
    fft, base2, mx2 = _run(plan, flat, offs, ns, gemm=False)
    assert list(base) == list(base2) and got.shape == fft.shape
    ref = np.concatenate([_reference(r) for r in recs]).reshape(-1)
    scale = ref.max()
    for name, x in (("tensor-core GEMM", got), ("mixed-radix FFT", fft)):
        d = np.abs(x - ref)
        print("%s: max abs error %.2e of the maximum, <= 1e-4 relative on %.4f of the cells" %
              (name, d.max() / scale, float(np.mean(d <= 1e-4 * np.abs(ref)))))
        assert d.max() <= 2e-6 * scale
        assert float(np.mean(d <= 1e-4 * np.abs(ref))) >= 0.9
    assert abs(mx - got.max()) == 0.0 and abs(mx - mx2) <= 4e-6 * scale


def test_band_gemm_falls_back_when_a_recording_is_not_16_byte_aligned(engine):
    """The GEMM path copies 16 bytes at a time; item offsets that are not multiples of 4 samples (or a PCM base that is
    not 16-byte aligned) take the FFT kernel - same call, same result layout."""
    from audio_analysis_b200 import identify_morepork as M
    recs = [synth.recording(310 + i, seconds=2.1 + 0.37 * i)[: 96001 + 17762 * i] for i in range(3)]
    plan = M._plan(SR)
    flat, offs, ns = _pack(recs, 1)                                 # offsets 0, 96001, 209764: not multiples of 4
    assert any(o % 4 for o in offs)
    a, _, _ = _run(plan, flat, offs, ns, gemm=True)
    b, _, _ = _run(plan, flat, offs, ns, gemm=False)
    assert np.array_equal(a, b)                                     # both ran the FFT kernel
    flat4, offs4, _ = _pack(recs, 4)
    c, _, _ = _run(plan, flat4, offs4, ns, gemm=True, shift=1)      # aligned offsets, misaligned base pointer
    d, _, _ = _run(plan, flat4, offs4, ns, gemm=False)
    assert np.array_equal(c, d)
    e, _, _ = _run(plan, flat4, offs4, ns, gemm=True)               # and the GEMM path proper agrees to its tolerance
    assert np.abs(e - d).max() <= 4e-6 * d.max()


@pytest.mark.parametrize("sr", [32000, 16000, 8000])
def test_band_gemm_other_rates(engine, sr):
    """n_fft = sr // 10 at other native rates (hop 1600 / 800 / 400 = 100 / 50 / 25 k-blocks; 8 kHz has no mixed-radix
    kernel: the direct band DFT is the other path)."""
    from audio_analysis_b200 import identify_morepork as M
    rec = synth.recording(320, seconds=7.3, sr=sr)
    plan = M._plan(sr)
    flat, offs, ns = _pack([rec, rec[: sr * 2 + 3]], 4)
    got, _, _ = _run(plan, flat, offs, ns, gemm=True)
    other, _, _ = _run(plan, flat, offs, ns, gemm=False)
    n_fft = sr // 10
    ref = np.concatenate([_reference(r, n_fft) for r in (rec, rec[: sr * 2 + 3])]).reshape(-1)
    # at 8 kHz the 600-1200 Hz band holds a small share of the (pink) spectrum, and the rounding errors of either path scale
    # with the energy of the whole frame: 4e-6 of the band maximum there
    tol = 4e-6 if sr == 8000 else 2e-6
    for name, x in (("gemm", got), ("other", other)):
        err = np.abs(x - ref).max() / ref.max()
        print("%d Hz %s: %.2e of the maximum" % (sr, name, err))
        assert err <= tol

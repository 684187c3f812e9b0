"""Seeded synthetic recordings: pink noise + Hann-tapered chirps.

float32 mono in [-1, 1]; `np.random.default_rng(seed0 + rec_idx)`; pink noise (white
Gaussian shaped by 1/sqrt(f) in the rfft domain, RMS 0.05) plus 8-16 Hann-tapered
linear/log chirps per minute (0.3-2 s, 500-9000 Hz, amplitude 0.1-0.5), clipped.
Used identically by the oracle and the GPU path; generation is I/O, outside any timed region.
"""
import numpy as np


def pink_noise(rng, n, rms=0.05):
    spec = np.fft.rfft(rng.standard_normal(n))
    f = np.arange(spec.shape[0], dtype=np.float64)
    f[0] = 1.0
    spec /= np.sqrt(f)
    x = np.fft.irfft(spec, n)
    x *= rms / np.sqrt(np.mean(x * x))
    return x


def recording(seed, seconds=60.0, sr=48000, zero_tail_s=0.0, return_chirps=False):
    """One synthetic recording.  `zero_tail_s` appends digital silence (for get_end)."""
    rng = np.random.default_rng(seed)
    n = int(round(seconds * sr))
    x = pink_noise(rng, n)
    n_chirps = int(rng.integers(8, 17) * max(seconds, 1.0) / 60.0 + 0.5) or 1
    chirps = []
    nyq_cap = min(9000.0, 0.45 * sr)
    f_lo = min(500.0, 0.05 * sr)
    for _ in range(n_chirps):
        dur = float(rng.uniform(0.3, 2.0))
        dur = min(dur, seconds * 0.5)
        t0 = float(rng.uniform(0.0, max(seconds - dur, 1e-3)))
        f0, f1 = (float(v) for v in rng.uniform(f_lo, nyq_cap, size=2))
        amp = float(rng.uniform(0.1, 0.5))
        log_sweep = bool(rng.integers(0, 2))
        m = int(dur * sr)
        t = np.arange(m) / sr
        if log_sweep and f0 != f1:
            k = (f1 / f0) ** (1.0 / dur)
            phase = 2 * np.pi * f0 * (k ** t - 1.0) / np.log(k)
        else:
            phase = 2 * np.pi * (f0 * t + 0.5 * (f1 - f0) / dur * t * t)
        i0 = int(t0 * sr)
        m = min(m, n - i0)
        x[i0:i0 + m] += (amp * np.hanning(m) * np.sin(phase[:m]))
        chirps.append((t0, t0 + dur, min(f0, f1), max(f0, f1)))
    np.clip(x, -1.0, 1.0, out=x)
    x = x.astype(np.float32)
    if zero_tail_s > 0:
        x = np.concatenate([x, np.zeros(int(zero_tail_s * sr), np.float32)])
    if return_chirps:
        return x, chirps
    return x


def batch(seed0, n_recordings, seconds=60.0, sr=48000):
    return np.stack([recording(seed0 + i, seconds, sr) for i in range(n_recordings)])

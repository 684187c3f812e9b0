"""Drop-in for the reference's custommel module (src/custommel.py).

`mel_f` builds the break-frequency mel filterbank on the host, once per parameter set (a
constant table, like FFT twiddles; the reference rebuilds it on every mel_spec call,
src/custommel.py:62).  `mel_spec` applies it on the GPU as banded dot products
(cac_mel_project_run) instead of the reference's dense sgemm (src/custommel.py:63).
"""
import functools

import numpy as np


def hz_to_mel(frequencies, break_freq):
    """src/custommel.py:6-8."""
    return 2595.0 * np.log10(1.0 + np.asarray(frequencies, dtype=np.float64) / break_freq)


def mel_to_hz(mels, break_freq):
    return break_freq * (10.0 ** (np.asarray(mels, dtype=np.float64) / 2595.0) - 1.0)


def mel_frequencies(n_mels, fmin, fmax, break_freq):
    """src/custommel.py:11-16: n_mels points uniformly spaced on the warped axis."""
    lo, hi = hz_to_mel(fmin, break_freq), hz_to_mel(fmax, break_freq)
    return mel_to_hz(np.linspace(lo, hi, n_mels), break_freq)


@functools.lru_cache(maxsize=32)
def _mel_f_cached(sr, n_mels, fmin, fmax, n_fft, break_freq):
    n_mels = int(n_mels)
    bin_hz = np.fft.rfftfreq(n=n_fft, d=1.0 / sr)                  # librosa.fft_frequencies
    pts = mel_frequencies(n_mels + 2, fmin, fmax, break_freq)
    width = np.diff(pts)
    dist = pts[:, None] - bin_hz[None, :]                          # float64, like subtract.outer
    rising = -dist[:n_mels] / width[:n_mels, None]
    falling = dist[2:n_mels + 2] / width[1:n_mels + 1, None]
    tri = np.maximum(0, np.minimum(rising, falling)).astype(np.float32)   # stored float32 first ...
    area = 2.0 / (pts[2:n_mels + 2] - pts[:n_mels])
    tri = (tri.astype(np.float64) * area[:, None]).astype(np.float32)     # ... then Slaney-scaled in place
    if not np.all((pts[:-2] == 0) | (tri.max(axis=1) > 0)):
        print("Empty filters detected in mel frequency basis. Some channels will produce empty "
              "responses. Try increasing your sampling rate (and fmax) or reducing n_mels.")
    tri.setflags(write=False)
    return tri


def mel_f(sr, n_mels, fmin, fmax, n_fft, break_freq):
    """src/custommel.py:19-56 -> float32 [n_mels, 1 + n_fft//2] (bit-identical values)."""
    return np.array(_mel_f_cached(sr, int(n_mels), fmin, fmax, int(n_fft), break_freq))


def mel_spec(stft, sr, n_fft, hop_length, n_mels, fmin, fmax, break_freq=1750, power=2):
    """src/custommel.py:59-63: mel_f(...) . |stft|**power, computed on the GPU."""
    from . import engine
    mag = np.abs(np.asarray(stft)).astype(np.float32, copy=False)
    plan = _projection_plan(sr, int(n_fft), int(hop_length), int(n_mels), fmin, fmax, break_freq, int(power))
    return plan.mel_project(mag)


@functools.lru_cache(maxsize=16)
def _projection_plan(sr, n_fft, hop_length, n_mels, fmin, fmax, break_freq, power):
    from . import engine
    w = _mel_f_cached(sr, n_mels, fmin, fmax, n_fft, break_freq)
    return engine.FeaturePlan(w, sr, n_fft, hop_length, n_mels, segment_samples=n_fft, power=power,
                              db_scale=False, normalize=False)


# ---- librosa's own (Slaney) filterbank: the htk=False branch of get_spect --------------------------
def _slaney_hz_to_mel(f):
    """librosa.hz_to_mel(htk=False): linear below 1 kHz (200/3 Hz per mel), logarithmic above."""
    f = np.asarray(f, dtype=np.float64)
    f_sp = 200.0 / 3
    mels = f / f_sp
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(f >= min_log_hz, min_log_mel + np.log(np.maximum(f, 1e-300) / min_log_hz) / logstep, mels)


def _slaney_mel_to_hz(m):
    m = np.asarray(m, dtype=np.float64)
    f_sp = 200.0 / 3
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), f_sp * m)


@functools.lru_cache(maxsize=16)
def _slaney_mel_f_cached(sr, n_fft, n_mels, fmin, fmax):
    """librosa.filters.mel(sr, n_fft, n_mels, fmin, fmax, htk=False, norm="slaney") as reached through
    librosa.feature.melspectrogram at src/identify_tracks.py:230-238 (librosa ~= 0.11; restated from its
    documented algorithm - librosa itself is not installable here, so this boundary is unpinned)."""
    bin_hz = np.fft.rfftfreq(n=n_fft, d=1.0 / sr)
    pts = _slaney_mel_to_hz(np.linspace(_slaney_hz_to_mel(fmin), _slaney_hz_to_mel(fmax), n_mels + 2))
    width = np.diff(pts)
    dist = pts[:, None] - bin_hz[None, :]
    rising = -dist[:n_mels] / width[:n_mels, None]
    falling = dist[2:n_mels + 2] / width[1:n_mels + 1, None]
    tri = np.maximum(0, np.minimum(rising, falling)).astype(np.float32)
    area = 2.0 / (pts[2:n_mels + 2] - pts[:n_mels])
    tri = (tri.astype(np.float64) * area[:, None]).astype(np.float32)
    tri.setflags(write=False)
    return tri


def slaney_mel_f(sr, n_fft, n_mels, fmin, fmax):
    return np.array(_slaney_mel_f_cached(sr, int(n_fft), int(n_mels), fmin, fmax))

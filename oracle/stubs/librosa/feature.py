"""librosa.feature stand-in (TEST INFRASTRUCTURE ONLY).  melspectrogram restates librosa 0.11's
documented algorithm for the htk=False branch of get_spect (src/identify_tracks.py:229-238):
power spectrogram |stft|**2 (hann, center, zero padding) projected through librosa.filters.mel
(Slaney scale, slaney norm, float32).  librosa itself is un-vendored and not installable here:
parity at this boundary is unpinned.  The use_mfcc branch (:269-280) is not restated - it cannot
run in the reference either (tf.image.resize_with_pad on a 2-D mfcc with *mel.shape)."""
import numpy as np

_stft = None   # set by the package __init__ (librosa.stft stand-in)


def _hz_to_mel(frequencies):
    frequencies = np.asanyarray(frequencies, dtype=float)
    f_min, f_sp = 0.0, 200.0 / 3
    mels = (frequencies - f_min) / f_sp
    min_log_hz = 1000.0
    min_log_mel = (min_log_hz - f_min) / f_sp
    logstep = np.log(6.4) / 27.0
    if frequencies.ndim:
        log_t = frequencies >= min_log_hz
        mels[log_t] = min_log_mel + np.log(frequencies[log_t] / min_log_hz) / logstep
    elif frequencies >= min_log_hz:
        mels = min_log_mel + np.log(frequencies / min_log_hz) / logstep
    return mels


def _mel_to_hz(mels):
    mels = np.asanyarray(mels, dtype=float)
    f_min, f_sp = 0.0, 200.0 / 3
    freqs = f_min + f_sp * mels
    min_log_hz = 1000.0
    min_log_mel = (min_log_hz - f_min) / f_sp
    logstep = np.log(6.4) / 27.0
    log_t = mels >= min_log_mel
    freqs[log_t] = min_log_hz * np.exp(logstep * (mels[log_t] - min_log_mel))
    return freqs


def filters_mel(sr, n_fft, n_mels=128, fmin=0.0, fmax=None):
    if fmax is None:
        fmax = float(sr) / 2
    weights = np.zeros((n_mels, int(1 + n_fft // 2)), dtype=np.float32)
    fftfreqs = np.fft.rfftfreq(n=n_fft, d=1.0 / sr)
    mel_f = _mel_to_hz(np.linspace(_hz_to_mel(fmin), _hz_to_mel(fmax), n_mels + 2))
    fdiff = np.diff(mel_f)
    ramps = np.subtract.outer(mel_f, fftfreqs)
    for i in range(n_mels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        weights[i] = np.maximum(0, np.minimum(lower, upper))
    enorm = 2.0 / (mel_f[2:n_mels + 2] - mel_f[:n_mels])
    weights *= enorm[:, np.newaxis]
    return weights


def melspectrogram(y=None, sr=22050, n_fft=2048, hop_length=512, n_mels=128, fmin=0.0, fmax=None, power=2.0, **kw):
    S = np.abs(_stft(y, n_fft=n_fft, hop_length=hop_length)) ** power
    return np.einsum("...ft,mf->...mt", S, filters_mel(sr, n_fft, n_mels, fmin, fmax), optimize=True)


def mfcc(*a, **k):
    raise NotImplementedError("librosa.feature.mfcc not restated (the reference's use_mfcc branch cannot run)")

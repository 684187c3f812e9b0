"""Stand-in for librosa ~=0.11.0 (TEST INFRASTRUCTURE ONLY - never imported by the product path).

librosa is not installed in the build container and its source is not under
/root/reference, so the few functions the reference's hot path calls are restated
here from librosa 0.11's documented semantics.  Call sites in the reference:
  stft            src/identify_tracks.py:243,389,654  src/identify_morepork.py:24
  power_to_db     src/identify_tracks.py:266
  amplitude_to_db src/identify_morepork.py:152
  fft_frequencies src/custommel.py:25  src/identify_tracks.py:677
It lets the reference's own files execute UNCHANGED (oracle/reference_loader.py) so
that golden vectors can be generated from the reference's slicing/quirk code.
Parity vs real librosa is unpinned (no librosa here); the stft stand-in is
cross-checked against torch.stft and scipy.signal.ShortTimeFFT in tests/.
"""
import numpy as np
import scipy.fft
import scipy.signal


__version__ = "0.11.0-standin"


def fft_frequencies(*, sr=22050, n_fft=2048):
    return np.fft.rfftfreq(n=n_fft, d=1.0 / sr)


class ParameterError(Exception):
    """librosa.util.exceptions.ParameterError"""


def stft(y, *, n_fft=2048, hop_length=None, win_length=None, window="hann",
         center=True, dtype=None, pad_mode="constant", out=None):
    """librosa.stft defaults: periodic Hann of win_length=n_fft, center=True with zero
    padding of n_fft//2 on both sides, frames = 1 + len(y)//hop, float32 input gives a
    complex64 result (the FFT itself runs in double inside scipy.fft/pocketfft because
    the float64 window promotes the product)."""
    if hop_length is None:
        hop_length = n_fft // 4
    if win_length is None:
        win_length = n_fft
    assert win_length == n_fft and window == "hann" and center and pad_mode == "constant"
    y = np.asarray(y)
    if not np.isfinite(y).all():                              # librosa.util.valid_audio
        raise ParameterError("Audio buffer is not finite everywhere")
    if dtype is None:
        dtype = np.complex64 if y.dtype == np.float32 else np.complex128
    fft_window = scipy.signal.get_window("hann", win_length, fftbins=True)
    ypad = np.pad(y, (n_fft // 2, n_fft // 2), mode="constant")
    n_frames = 1 + (len(ypad) - n_fft) // hop_length
    out = np.empty((1 + n_fft // 2, n_frames), dtype=dtype)
    # block the work so the float64 frame matrix stays small
    block = max(1, (1 << 22) // n_fft)
    idx = np.arange(n_fft)[None, :]
    for b0 in range(0, n_frames, block):
        b1 = min(n_frames, b0 + block)
        starts = (np.arange(b0, b1) * hop_length)[:, None]
        frames = ypad[starts + idx]                       # [nb, n_fft] in y.dtype
        spec = scipy.fft.rfft(fft_window[None, :] * frames, axis=1)
        out[:, b0:b1] = spec.T
    return out


def power_to_db(S, *, ref=1.0, amin=1e-10, top_db=80.0):
    S = np.asarray(S)
    magnitude = np.abs(S) if np.iscomplexobj(S) else S
    ref_value = ref(magnitude) if callable(ref) else np.abs(ref)
    log_spec = 10.0 * np.log10(np.maximum(amin, magnitude))
    log_spec -= 10.0 * np.log10(np.maximum(amin, ref_value))
    if top_db is not None:
        log_spec = np.maximum(log_spec, log_spec.max() - top_db)
    return log_spec


def amplitude_to_db(S, *, ref=1.0, amin=1e-5, top_db=80.0):
    S = np.asarray(S)
    magnitude = np.abs(S)
    ref_value = ref(magnitude) if callable(ref) else np.abs(ref)
    power = np.square(magnitude)
    return power_to_db(power, ref=ref_value ** 2, amin=amin ** 2, top_db=top_db)


def load(*a, **k):
    raise NotImplementedError("librosa.load is decode I/O (out of scope); stand-in only")


def resample(*a, **k):
    raise NotImplementedError("librosa.resample (soxr) is out of scope; stand-in only")


from . import feature  # noqa: E402,F401  (after stft is defined: feature.melspectrogram calls it)
feature._stft = stft

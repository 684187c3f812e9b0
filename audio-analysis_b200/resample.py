"""Resampling front-end on the GPU.

The reference resamples in two places, both through third-party code that cannot be pinned here:
`ffmpeg -ar 16000` in front of the Cacophony Index (src/common.py:70-96) and
`librosa.resample(frames, orig_sr=sr, target_sr=48000)` (soxr_hq) in load_recording
(src/identify_tracks.py:51-58).  This module does NOT claim their filters: it implements
scipy.signal.resample_poly's documented design (Kaiser beta = 5 windowed sinc with 20 max(up, down) + 1
taps, zero padding at both ends) so that the result is pinned against scipy itself (float32 arithmetic in
scipy's own accumulation order), and one 48 kHz upload can feed the classifier, morepork and Cacophony
Index paths."""
import functools
import math

import numpy as np


def poly_filter(up, down, window=("kaiser", 5.0)):
    """(up, down, padded taps float32, n_pre_remove) as scipy.signal.resample_poly derives them for a
    float32 input (scipy/signal/_signaltools.py: resample_poly; _upfirdn.py: _output_len)."""
    from scipy.signal import firwin
    g = math.gcd(int(up), int(down))
    up, down = int(up) // g, int(down) // g
    max_rate = max(up, down)
    half_len = 10 * max_rate
    h = firwin(2 * half_len + 1, 1.0 / max_rate, window=window).astype(np.float32)
    h *= up
    n_pre_pad = down - half_len % down
    n_pre_remove = (half_len + n_pre_pad) // down
    return up, down, h, n_pre_pad, n_pre_remove


def _output_len(len_h, in_len, up, down):
    return (((in_len - 1) * up + len_h) - 1) // down + 1


@functools.lru_cache(maxsize=8)
def _plan(up, down, device=0):
    from . import engine
    up, down, h, n_pre_pad, n_pre_remove = poly_filter(up, down)
    taps = np.concatenate([np.zeros(n_pre_pad, np.float32), h])     # trailing zero padding adds nothing
    return engine.ResamplePlan(up, down, taps, n_pre_remove, device)


def resample_poly(x, up, down, device=0):
    """scipy.signal.resample_poly(x, up, down) (default window, padtype='constant') for a 1-D or
    [n, samples] float32 numpy array or CUDA tensor; returns the same kind of object."""
    import torch
    g = math.gcd(int(up), int(down))
    if up // g == down // g == 1:
        return x.copy() if isinstance(x, np.ndarray) else x.clone()
    plan = _plan(int(up), int(down), device)
    was_np = isinstance(x, np.ndarray)
    t = torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)).cuda(device) if was_np else x.contiguous()
    rows = 1 if t.dim() == 1 else t.shape[0]
    n_in = t.shape[-1]
    n_out = plan.out_len(n_in)
    out = torch.empty((rows, n_out), dtype=torch.float32, device=t.device)
    plan.run(t.reshape(-1), np.arange(rows, dtype=np.int64) * n_in, np.full(rows, n_in, np.int64), out=out.view(-1),
             out_offset=np.arange(rows, dtype=np.int64) * n_out)
    out = out[0] if t.dim() == 1 else out
    return out.cpu().numpy() if was_np else out


def resample(frames, orig_sr, target_sr, device=0):
    """Drop-in shape of librosa.resample(frames, orig_sr=..., target_sr=...) with the filter described above."""
    g = math.gcd(int(orig_sr), int(target_sr))
    return resample_poly(frames, int(target_sr) // g, int(orig_sr) // g, device)


def to_16k(frames, sr=48000, device=0):
    """What `ffmpeg -ar 16000` does in front of cacophony_index.calculate (src/common.py:76-79), on the GPU."""
    return resample(frames, sr, 16000, device)

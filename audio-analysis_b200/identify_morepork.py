"""Drop-in for the spectrogram half of the reference's identify_morepork module
(src/identify_morepork.py:6-27 and the window loop :142-167): native-rate STFT with
n_fft = int(sr / 10), hop = n_fft / 2, rows 60:120 (600-1200 Hz), 60-frame windows sliding by
20 frames, librosa.amplitude_to_db(ref=np.max), rescale to [0, 1], zero-pad.  All sample- and
cell-touching work runs in CUDA kernels (engine.SpecPlan); the SavedModel ensemble and the
span heuristics (:30-134, :169-200) are outside this package (models are not in the repo).
"""
import functools

import numpy as np

frequency_min = 600
frequency_max = 1200
num_bands = int((frequency_max - frequency_min) / 10)
slices_per_second = 20
seconds_per_sample = 3.0
slices_per_sample = int(slices_per_second * seconds_per_sample)
sample_slide_seconds = 1.0
sample_slide_slices = int(sample_slide_seconds * slices_per_second)
activation_threshold = 1.0


@functools.lru_cache(maxsize=8)
def _plan(sr, device=0):
    from . import engine
    nfft = int(sr / 10)
    return engine.SpecPlan(nfft, int(nfft / 2), bin_lo=int(frequency_min / 10), bin_hi=int(frequency_max / 10),
                           n_mels=0, power=1, bin_major=False, device=device)


def window_table(n_frames):
    """Column range of every window of the loop at src/identify_morepork.py:144-149, including the
    Python negative-index slice when the recording has fewer than 60 frames:
    returns (first_column, width, start_time) lists."""
    first, width, start_times = [], [], []
    for base in range(0, n_frames, sample_slide_slices):
        limit = min(base + slices_per_sample, n_frames)
        start = limit - slices_per_sample
        start_times.append(start / slices_per_second)
        lo = start if start >= 0 else max(n_frames + start, 0)     # npspec[:, start:limit] with start < 0
        first.append(lo)
        width.append(limit - lo)
    return first, width, start_times


def _decode(recording, loader=None):
    if isinstance(recording, tuple):
        frames, sr = recording
    else:
        if loader is None:
            import librosa  # the reference's decoder (src/identify_morepork.py:20); I/O outside this package
            loader = lambda p: librosa.load(p, sr=None)  # noqa: E731
        frames, sr = loader(recording)
    return np.ascontiguousarray(frames, dtype=np.float32), int(sr)


def _load_sample(path, loader=None):
    """src/identify_morepork.py:19-27 -> (sr, npspec[60, n_frames]).  `path` may also be an already
    decoded (frames, sr) tuple."""
    import torch
    frames, sr = _decode(path, loader)
    plan = _plan(sr)
    spec, base = plan.run(torch.from_numpy(frames).cuda(plan.device), [0], [frames.shape[0]])
    n = int(base[-1])
    return sr, spec.view(n, plan.rows).t().contiguous().cpu().numpy()


def samples_from_recording(recording, loader=None):
    """identify_morepork's preparation (src/identify_morepork.py:139-167): returns
    (samples float32 [n_windows, 60, 60, 1], start_times) for one recording, spectrogram and
    windows both computed on the GPU."""
    import torch
    frames, sr = _decode(recording, loader)
    plan = _plan(sr)
    out, starts = BatchPlan(sr, frames.shape[0], plan.device).run(torch.from_numpy(frames).cuda(plan.device)[None, :])
    return out[0].cpu().numpy()[..., None], starts


class BatchPlan:
    """Spectrogram + windows for a batch of equally long recordings held in one CUDA tensor
    [n_recordings, n_samples] (BASELINE.json configs[2])."""

    def __init__(self, sr, n_samples, device=0):
        self.plan = _plan(int(sr), device)
        self.n_samples = int(n_samples)
        self.n_frames = self.plan.frame_count(n_samples)
        self.first, self.width, self.start_times = window_table(self.n_frames)
        self.n_windows = len(self.first)
        self._spec = None

    def run(self, pcm, out=None):
        torch = __import__("torch")
        assert pcm.dim() == 2 and pcm.shape[1] == self.n_samples and pcm.is_contiguous()
        n_rec = pcm.shape[0]
        offs = np.arange(n_rec, dtype=np.int64) * self.n_samples
        need = n_rec * self.n_frames * self.plan.rows
        if self._spec is None or self._spec.numel() < need:
            self._spec = torch.empty((need,), dtype=torch.float32, device=pcm.device)
        self.plan.run(pcm.view(-1), offs, np.full(n_rec, self.n_samples, dtype=np.int64), out=self._spec)
        f0 = (np.arange(n_rec, dtype=np.int64)[:, None] * self.n_frames + np.array(self.first, dtype=np.int64)[None, :])
        wd = np.tile(np.array(self.width, dtype=np.int32), n_rec)
        if out is None:
            out = torch.empty((n_rec, self.n_windows, self.plan.rows, 60), dtype=torch.float32, device=pcm.device)
        self.plan.morepork_windows(self._spec, f0.reshape(-1), wd, out=out)
        return out, list(self.start_times)


def identify_morepork(recording, models):
    """The classifier half (src/identify_morepork.py:169-200) needs the SavedModel ensemble, which is
    not part of this package; use samples_from_recording() for its inputs."""
    raise NotImplementedError("morepork SavedModel inference is outside cacophony-b200; "
                              "samples_from_recording() produces the model inputs")

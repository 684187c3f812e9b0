"""Drop-in for the feature-extraction half of the reference's identify_tracks module:
load_samples / normalize_data / get_spect (src/identify_tracks.py:65-288) with the same
signatures.  The slicing arithmetic (including the legacy global-RNG draws) runs on the
host and produces a segment descriptor table; every sample-touching step (normalise, STFT,
mel, dB) runs in the CUDA kernels behind engine.FeaturePlan.
"""
import functools
import logging

import numpy as np

from . import custommel


from .tracks import Signal, get_tracks_from_signals, merge_signals  # noqa: F401  (src/identify_tracks.py:725-842, 915-1033)


@functools.lru_cache(maxsize=16)
def _plan(sr, n_fft, hop_length, n_mels, fmin, fmax, mel_break, segment_samples, power, db_scale, normalize):
    """mel_break None selects librosa's own Slaney filterbank (get_spect's htk=False branch)."""
    from . import engine
    if mel_break is None:
        w = custommel._slaney_mel_f_cached(sr, n_fft, n_mels, fmin, fmax)
    else:
        w = custommel._mel_f_cached(sr, n_mels, fmin, fmax, n_fft, mel_break)
    return engine.FeaturePlan(w, sr, n_fft, hop_length, n_mels, segment_samples, power=power,
                              db_scale=db_scale, normalize=normalize)


def segment_descriptors(n_samples, sr, tracks, segment_length, stride, fmin=50, fmax=11000,
                        pad_short_tracks=False, randint=None):
    """Host half of load_samples (src/identify_tracks.py:109-196): returns
    (descriptors[engine.SEGMENT_DTYPE], counts_per_track).  Consumes np.random.randint in the
    same order as the reference (:132 once per short track, :167 once per short segment)."""
    from .engine import SEGMENT_DTYPE
    draw = randint or np.random.randint
    seg_n = int(sr * segment_length)
    rows, counts = [], []
    for t in tracks:
        if t.freq_start > fmax or t.freq_end < fmin:
            counts.append(0)                      # placeholder track, nothing to classify
            continue
        lo, hi = int(sr * t.start), int(t.end * sr)
        if not pad_short_tracks:
            short_by = seg_n - (hi - lo)
            if short_by > 0:                      # widen to exactly one segment around the track
                shift = draw(0, short_by)
                lo -= shift
                if lo <= 0:
                    lo, hi = 0, min(seg_n, n_samples)
                else:
                    hi = hi + short_by - shift
                    if hi > n_samples:
                        hi = n_samples
                        lo = max(hi - seg_n, 0)
                assert hi - lo == seg_n
        base = min(max(lo, 0), n_samples)         # numpy slice clipping of frames[lo:hi]
        avail = min(max(hi, base), n_samples) - base
        a, b = 0, min(hi, seg_n)                  # first window ends at the ABSOLUTE index hi (:150-151)
        t_start = 0
        n_here = 0
        while True:
            a_c = min(a, avail)
            n_valid = min(max(b, a_c), avail) - a_c
            pad_left = int(draw(0, seg_n - n_valid)) if n_valid != seg_n else 0
            rows.append((base + a_c, n_valid, pad_left))
            n_here += 1
            t_start = t_start + stride
            t_end = t_start + segment_length
            a = int(t_start * sr)
            b = min(int(t_end * sr), a + seg_n)
            if t_end > t.length:
                break
        counts.append(n_here)
    return np.array(rows, dtype=SEGMENT_DTYPE).reshape(-1), counts


class ParameterError(Exception):
    """Stands in for librosa.util.exceptions.ParameterError (librosa itself is not a dependency of this package)."""


def _check_finite(first_values):
    """librosa.stft refuses non-finite audio (util.valid_audio).  The one way load_samples gets there is a
    constant segment: normalize_data (src/identify_tracks.py:202-209) divides 0 by 0 and the whole segment is
    NaN.  The kernels carry the NaNs through to the output like numpy would; the drop-in raises like librosa."""
    if np.isnan(first_values).any():
        raise ParameterError("Audio buffer is not finite everywhere")


def _is_cuda(frames):
    return hasattr(frames, "is_cuda") and frames.is_cuda


def to_device(frames, device=0):
    """Upload a decoded recording ONCE; get_end / signal_noise / load_samples accept the returned CUDA float32
    tensor in place of the numpy array (classify calls all three on the same frames, src/identify_tracks.py:419-420,
    502), so the 11.5 MB per minute cross PCIe a single time."""
    import torch
    if _is_cuda(frames):
        return frames
    return torch.from_numpy(np.ascontiguousarray(frames, dtype=np.float32)).cuda(device)


def _finish(block, mean_sub, channels):
    """get_spect tail (src/identify_tracks.py:267,282-287) on [n, n_mels, frames]."""
    block = block[..., None]
    if mean_sub:
        block = block - np.mean(block, axis=2, keepdims=True, dtype=block.dtype)
    if channels > 1:
        block = np.repeat(block, channels, axis=3)
    return block


def load_samples(frames, sr, tracks, segment_length, stride, hop_length=281, mean_sub=False, use_mfcc=False,
                 mel_break=1000, htk=True, n_mels=160, fmin=50, fmax=11000, channels=1, power=2,
                 db_scale=False, filter_freqs=False, filter_below=None, normalize=True, n_fft=4096,
                 pad_short_tracks=False):
    """Same signature and return structure as src/identify_tracks.py:65-199: one list per
    track (empty for out-of-band tracks), each holding [n_mels, frames, channels] arrays."""
    logging.debug("Loading samples with length %s stride %s hop length %s", segment_length, stride, hop_length)
    if use_mfcc:
        raise NotImplementedError("use_mfcc: the reference branch (src/identify_tracks.py:269-280) cannot run "
                                  "(tf.image.resize_with_pad on a 2-D mfcc with *mel.shape); not built")
    if filter_freqs or filter_below:
        raise NotImplementedError("Butterworth pre-filter is a host pre-step outside this path")
    on_device = _is_cuda(frames)      # to_device(frames): already uploaded, only the results come back
    if not on_device:
        # int16 frames (PCM as decoded) are uploaded as they are and widened on the device to sample / 32768
        frames = np.ascontiguousarray(frames, dtype=np.int16 if getattr(frames, "dtype", None) == np.int16 else np.float32)
    segs, counts = segment_descriptors(len(frames), sr, tracks, segment_length, stride, fmin, fmax,
                                       pad_short_tracks)
    if len(segs) == 0:
        return [[] for _ in counts]
    if not htk:     # librosa.feature.melspectrogram(fmin=50, fmax=11000): Slaney mel, power 2 (:229-238)
        fmin, fmax, power, mel_break = 50, 11000, 2, None
    plan = _plan(sr, int(n_fft), int(hop_length), int(n_mels), fmin, fmax, mel_break,
                 int(sr * segment_length), int(power), bool(db_scale), bool(normalize))
    raw = plan.run(frames, segs).cpu().numpy() if on_device else plan.run_host(frames, segs)
    _check_finite(raw[:, 0, 0])
    block = _finish(raw, mean_sub, channels)
    out, k = [], 0
    for c in counts:
        out.append([block[k + i] for i in range(c)])
        k += c
    return out


def load_samples_device(frames, sr, tracks, segment_length, stride, hop_length=281, mean_sub=False,
                        mel_break=1000, htk=True, n_mels=160, fmin=50, fmax=11000, channels=1, power=2,
                        db_scale=False, normalize=True, n_fft=4096, pad_short_tracks=False, device=0):
    """load_samples for a classifier that lives on the GPU: `frames` is a numpy array or a
    CUDA float32 tensor; returns one CUDA tensor [n_segments, n_mels, frames, channels] per track (views of
    one block; a [0, ...] tensor for out-of-band tracks), i.e. what `np.array(d)` builds at
    src/identify_tracks.py:544 - nothing is copied back to the host.  mean_sub / channels (and classify's
    efficientnet x3 repeat, :539-540 = channels 3) run in cac_spect_tail_run."""
    import torch
    from . import engine
    dev = torch.device("cuda", device)
    if isinstance(frames, np.ndarray):
        frames = torch.from_numpy(np.ascontiguousarray(frames, dtype=np.float32)).to(dev)
    assert frames.is_cuda and frames.dtype == torch.float32 and frames.dim() == 1 and frames.is_contiguous()
    segs, counts = segment_descriptors(frames.numel(), sr, tracks, segment_length, stride, fmin, fmax, pad_short_tracks)
    if not htk:
        fmin, fmax, power, mel_break = 50, 11000, 2, None
    plan = _plan(sr, int(n_fft), int(hop_length), int(n_mels), fmin, fmax, mel_break,
                 int(sr * segment_length), int(power), bool(db_scale), bool(normalize))
    if len(segs) == 0:
        empty = torch.empty((0, int(n_mels), plan.n_frames, int(channels)), dtype=torch.float32, device=dev)
        return [empty for _ in counts]
    raw = plan.run(frames, segs)
    _check_finite(raw[:, 0, 0].cpu().numpy())
    block = engine.spect_tail(raw, mean_sub, channels)
    out, k = [], 0
    for c in counts:
        out.append(block[k:k + c])
        k += c
    return out


def normalize_data(x):
    """src/identify_tracks.py:202-209 on the GPU (bit-exact float32 op order)."""
    from . import engine
    return engine.normalize_rows(np.asarray(x, dtype=np.float32))


def get_spect(data, sr, hop_length, mean_sub, use_mfcc, mel_break, htk, n_mels, fmin, fmax, n_fft, power,
              db_scale, channels=1, pass_freqs=None):
    """src/identify_tracks.py:212-288 for one already-normalised segment."""
    from .engine import SEGMENT_DTYPE
    if use_mfcc:
        raise NotImplementedError("use_mfcc: the reference branch cannot run as written; not built")
    data = np.ascontiguousarray(data, dtype=np.float32)
    if not htk:     # librosa.feature.melspectrogram(fmin=50, fmax=11000): Slaney mel, power 2 (:229-238)
        plan = _plan(sr, int(n_fft), int(hop_length), int(n_mels), 50, 11000, None, len(data), 2, bool(db_scale), False)
    else:
        plan = _plan(sr, int(n_fft), int(hop_length), int(n_mels), 50 if fmin is None else fmin,
                     11000 if fmin is None else fmax, mel_break, len(data), int(power), bool(db_scale), False)
    seg = np.array([(0, len(data), 0)], dtype=SEGMENT_DTYPE)
    return _finish(plan.run_host(data, seg), mean_sub, channels)[0]


@functools.lru_cache(maxsize=4)
def _get_end_plan(sr, device=0):
    from . import engine
    n_fft = sr // 10
    w = custommel._mel_f_cached(sr, 120, 50, 11000, n_fft, 1750)
    return engine.SpecPlan(n_fft, 281, n_mels=120, power=1, mel_weights=w, device=device)


def _as_batch(recordings, device):
    """list of 1-D arrays / CUDA tensors, or a 2-D array / tensor [n, samples] -> (flat CUDA float32 tensor,
    offsets, lengths) without copying what is already one contiguous CUDA block."""
    import torch
    dev = torch.device("cuda", device)
    if hasattr(recordings, "dim") and recordings.dim() == 2 and recordings.is_cuda and recordings.is_contiguous() \
            and recordings.dtype == torch.float32:
        n, m = recordings.shape
        return recordings.reshape(-1), np.arange(n, dtype=np.int64) * m, np.full(n, m, dtype=np.int64)
    items = [recordings[i] for i in range(len(recordings))]
    lens = np.array([int(r.shape[0]) for r in items], dtype=np.int64)
    if len(items) == 1 and _is_cuda(items[0]) and items[0].dtype == torch.float32 and items[0].is_contiguous():
        return items[0].reshape(-1), np.zeros(1, np.int64), lens              # one uploaded recording: used in place
    padded = (lens + 3) // 4 * 4                                   # keep every recording 16-byte aligned
    offs = np.concatenate([[0], np.cumsum(padded)[:-1]]).astype(np.int64) if len(items) else np.zeros(0, np.int64)
    flat = torch.zeros((int(padded.sum()),), dtype=torch.float32, device=dev)
    for r, o, n in zip(items, offs, lens):
        src = r if _is_cuda(r) else torch.from_numpy(np.ascontiguousarray(r, dtype=np.float32))
        flat[o:o + n].copy_(src, non_blocking=True)
    return flat, offs, lens


def get_end_batch(recordings, sr):
    """get_end (src/identify_tracks.py:387-413) for many recordings with one launch chain and one synchronisation:
    `recordings` as for signal_noise_batch.  STFT (n_fft = sr // 10, hop 281), the 120-mel projection and the
    1-second chunk scan all run on the device (cac_get_end_run); returns one value per recording, equal to get_end()."""
    hop_length = 281
    plan = _get_end_plan(int(sr))
    flat, offs, lens = _as_batch(recordings, plan.device)
    if len(lens) == 0:
        return []
    chunk_length = sr // hop_length
    first = plan.get_end_chunks(flat, offs, lens, chunk_length).cpu().numpy()          # the one synchronisation
    return [int(c) * chunk_length * hop_length // sr if c >= 0 else int(n) / sr for c, n in zip(first.tolist(), lens.tolist())]


def get_end(frames, sr):
    """src/identify_tracks.py:387-413: first 1-second chunk of the 120-mel magnitude spectrogram
    (n_fft = sr // 10, hop 281) whose values are all equal (digital silence) -> its start in whole
    seconds; else the file length.  STFT, mel and the chunk scan run on the GPU (one recording of get_end_batch).
    `frames`: numpy array or the CUDA tensor to_device() returned."""
    plan = _get_end_plan(int(sr))
    return get_end_batch([to_device(frames, plan.device)], sr)[0]


@functools.lru_cache(maxsize=8)
def _signal_noise_plan(sr, hop_length, device=0):
    """Structuring elements as signal_noise derives them (src/identify_tracks.py:673-684): width = int(0.25 sr / hop),
    height = index of the first FFT bin above 100 Hz, plus one."""
    from . import engine
    width = int(SIGNAL_WIDTH * sr / hop_length)
    freqs = np.fft.rfftfreq(n=4096, d=1.0 / sr)                  # librosa.fft_frequencies
    above = np.nonzero(freqs > 100)[0]
    height = int(above[0]) + 1 if len(above) else 0
    plan = engine.SignalNoisePlan(hop_length, height, width, height // 10, device=device)
    plan.width, plan.height, plan.freqs = width, height, freqs
    return plan


def signal_noise_spectrogram(frames, sr, hop_length=281):
    """First stage of signal_noise (src/identify_tracks.py:650-657): |STFT(n_fft=4096, hop)| of the
    whole recording divided by its maximum, as a CUDA float32 tensor [2049, n_frames]."""
    plan = _signal_noise_plan(int(sr), int(hop_length))
    spec, _, a_max = plan.spectrum(to_device(frames, plan.device))
    return (spec / a_max).contiguous()                            # IEEE float32 division, as numpy's


SIGNAL_WIDTH = 0.25   # src/identify_tracks.py:21


def _signals_from_stats(stats, sr, width, height, freqs, signal_cls):
    """Reference filter (:689-691) and time / frequency conversion (:697-704, with its hard-coded 281) of the
    component table [n, 6] = (x, y, w, h, area, order)."""
    stats = stats[np.argsort(stats[:, 5], kind="stable")]         # cv2's label order ...
    stats = stats[np.argsort(stats[:, 0], kind="stable")]         # ... then sorted(stats, key=x) (stable)
    min_width = 0.65 * width
    min_height = height - height // 10
    signals = []
    for x, y, w, h, _, _ in stats.tolist():
        if not (w > min_width and h > min_height):
            continue
        max_freq = min(len(freqs) - 1, y + h)
        signals.append((signal_cls or Signal)(x * 281 / sr, (x + w) * 281 / sr, freqs[y], freqs[max_freq]))
    return signals


def signal_noise_batch(recordings, sr, hop_length=281, signal_cls=None, max_components=4096):
    """signal_noise (src/identify_tracks.py:650-706) for MANY recordings: `recordings` is a list of 1-D numpy arrays /
    CUDA tensors or a 2-D array / CUDA tensor [n, samples].  One cac_sn_run call queues the whole chain for all of them
    (spectrum with column medians, row medians + threshold mask, morphology, connected components; recordings go on
    the grid, nothing is launched per recording), the small component tables come back with one copy.  A recording
    with more than max_components components (pathological noise) is redone alone with room for 65,536.  Returns
    one list of Signal per recording, equal to signal_noise() called on each."""
    plan = _signal_noise_plan(int(sr), int(hop_length))
    flat, offs, lens = _as_batch(recordings, plan.device)
    if len(lens) == 0:
        return []
    from . import engine
    host = plan.run(flat, offs, lens, max_components=max_components).cpu().numpy()     # the one synchronisation
    out = []
    for i in range(len(lens)):
        if int(host[i, 0]) > max_components:
            row = plan.run(flat, offs[i:i + 1], lens[i:i + 1], max_components=65536).cpu().numpy()[0]
        else:
            row = host[i]
        out.append(_signals_from_stats(engine.collect_components(row), sr, plan.width, plan.height, plan.freqs, signal_cls))
    return out


def signal_noise(frames, sr, hop_length=281, signal_cls=None):
    """src/identify_tracks.py:650-706: candidate track signals of a recording (`signal_cls`: the class to
    build, e.g. tracks.Signal with the reference's merge helpers; default: this module's Signal).  STFT magnitude,
    medians, threshold, morphology and connected components run on the GPU (one recording of signal_noise_batch);
    the handful of component boxes comes back to the host, where the reference's filter (:689-691) and time /
    frequency conversion (:697-704, including its hard-coded 281) are applied.  Returns a list of Signal.
    Supported range: any sr whose structuring element height (first FFT bin above 100 Hz, + 1) is at most 255
    (sr >= 1.7 kHz; erode height 0, sr > ~51 kHz, behaves like cv2's empty kernel) and recordings of up to
    2^31 / 2049 frames (~100 minutes at hop 281 / 48 kHz)."""
    plan = _signal_noise_plan(int(sr), int(hop_length))
    return signal_noise_batch([to_device(frames, plan.device)], sr, hop_length, signal_cls, max_components=65536)[0]


def classify_prep_batch(recordings, sr, segment_length=3, segment_stride=1.5, hop_length=281, mean_sub=False,
                        mel_break=1000, htk=True, n_mels=160, fmin=50, fmax=11000, channels=1, power=2, db_scale=False,
                        normalize=True, n_fft=4096, pad_short_tracks=False, max_components=1024, want_samples=True,
                        chunk=24):
    """Everything classify does to a recording before its model runs (src/identify_tracks.py:416-440, 502-524) for a
    batch of decoded recordings.  The batch goes through in chunks of `chunk` recordings, and per chunk:
      1. get_end -> signal_noise(frames[: int(sr * length)]) chained on the device (the trimmed lengths are computed on
         the GPU from get_end's chunk index and handed to cac_sn_run as its device item table); one asynchronous copy
         brings the end chunks and the component tables of the chunk back into page-locked memory;
      2. the host turns components into signals and tracks (get_tracks_from_signals, O(100) boxes per recording) and
         ONE cac_features_run per chunk produces the classifier segments of every track on the device.
    The GPU work of chunk i + 1 is queued before the host waits for chunk i, so the host's scalar merge logic overlaps
    the kernels; a single recording costs two synchronisations.
    Returns one dict per recording: tracks, length, signals (copies taken before the merge, as classify keeps them for
    the chirp count), raw_length, and samples = one CUDA tensor [n_segments, n_mels, frames, channels] per track
    (what load_samples_device returns; None when want_samples is False).  Consumes np.random like load_samples."""
    import torch
    from . import engine
    sn = _signal_noise_plan(int(sr), 281)                         # classify passes 281 (:420)
    ge = _get_end_plan(int(sr))
    flat, offs, lens = _as_batch(recordings, sn.device)
    n = len(lens)
    if n == 0:
        return []
    dev = flat.device
    chunk_length = sr // 281
    max_len = int(lens.max())
    rstride = 4 + 6 * int(max_components)
    tables = torch.from_numpy(np.stack([offs, lens, np.array([int(sr * (int(m) / sr)) for m in lens], dtype=np.int64)])).to(dev)
    first_h = torch.empty((n,), dtype=torch.int32).pin_memory()
    comp_h = torch.empty((n, rstride), dtype=torch.int32).pin_memory()
    feat_plan = None
    if want_samples:
        f_lo, f_hi, pw, brk = (50, 11000, 2, None) if not htk else (fmin, fmax, power, mel_break)
        feat_plan = _plan(sr, int(n_fft), int(hop_length), int(n_mels), f_lo, f_hi, brk, int(sr * segment_length),
                          int(pw), bool(db_scale), bool(normalize))

    def enqueue(lo, hi):
        """phase 1 of recordings [lo, hi): kernels + the copy of their small results, nothing waits"""
        first = ge.get_end_chunks(flat, tables[0, lo:hi], tables[1, lo:hi], chunk_length, max_item_samples=max_len)
        # frames[: int(sr * length)] with length = chunk * chunk_length * 281 // sr whole seconds, or len / sr (:387-420)
        first64 = first.to(torch.int64)
        secs = torch.div(first64 * (chunk_length * 281), sr, rounding_mode="floor")
        trimmed = torch.minimum(torch.where(first64 >= 0, secs * sr, tables[2, lo:hi]), tables[1, lo:hi])
        comp = sn.run(flat, tables[0, lo:hi], trimmed, max_components=max_components, max_item_samples=max_len)
        first_h[lo:hi].copy_(first, non_blocking=True)
        comp_h[lo:hi].copy_(comp, non_blocking=True)
        done = torch.cuda.Event()
        done.record()
        return done

    out = [None] * n
    feat_kw = dict(fmin=fmin, fmax=fmax, pad_short_tracks=pad_short_tracks)
    first_np, comp_np = first_h.numpy(), comp_h.numpy()

    def finish(lo, hi):
        """host half of recordings [lo, hi) + their classifier segments (queued, not waited for)"""
        all_segs, per_rec = [], []
        for i in range(lo, hi):
            c = int(first_np[i])
            length = c * chunk_length * 281 // sr if c >= 0 else int(lens[i]) / sr
            row = comp_np[i]
            if int(row[0]) > max_components:                       # pathological recording: redo it alone
                n_trim = min(int(sr * length), int(lens[i]))
                row = sn.run(flat, offs[i:i + 1], np.array([n_trim]), max_components=65536).cpu().numpy()[0]
            signals = _signals_from_stats(engine.collect_components(row), sr, sn.width, sn.height, sn.freqs, Signal)
            tracks = get_tracks_from_signals([s.copy() for s in signals], length)
            out[i] = {"tracks": tracks, "length": length, "signals": signals, "raw_length": int(lens[i]) / sr,
                      "samples": None}
            if want_samples and tracks:
                segs, counts = segment_descriptors(int(lens[i]), sr, tracks, segment_length, segment_stride, **feat_kw)
                segs["pcm_offset"] += int(offs[i])
                all_segs.append(segs)
                per_rec.append((i, counts))
        if all_segs:
            block = engine.spect_tail(feat_plan.run(flat, np.concatenate(all_segs)), mean_sub, channels)
            k = 0
            for i, counts in per_rec:
                samples = []
                for cnt in counts:
                    samples.append(block[k:k + cnt])
                    k += cnt
                out[i]["samples"] = samples

    bounds = [(lo, min(lo + int(chunk), n)) for lo in range(0, n, int(chunk))]
    pending = enqueue(*bounds[0])
    for j, (lo, hi) in enumerate(bounds):
        nxt = enqueue(*bounds[j + 1]) if j + 1 < len(bounds) else None     # the GPU stays busy while the host merges
        pending.synchronize()
        finish(lo, hi)
        pending = nxt
    return out


def classify_prep(frames, sr, **kw):
    """classify_prep_batch for one recording: one upload, two synchronisations (see there)."""
    return classify_prep_batch([to_device(frames)], sr, **kw)[0]

"""CPU restatement (numpy/scipy, float64 FFTs like the reference) of the reference's
feature hot path.  TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs, never by the product package.

Every function cites the reference lines it restates (paths relative to
/root/reference).  The restatement is pinned by tests/test_oracle.py against fixtures
under tests/golden/ that were produced by executing the reference's own files
(tests/golden/make_golden.py).  librosa itself (un-vendored, ~=0.11.0,
requirements.txt:1) is restated from its documented semantics: that boundary is
"parity unpinned" - no librosa is installable here and the reference holds no tests.
"""
import math

import numpy as np
import scipy.fft
import scipy.fftpack
import scipy.signal


# ----------------------------------------------------------------------------- librosa
def fft_frequencies(sr, n_fft):
    """librosa.fft_frequencies == np.fft.rfftfreq (src/custommel.py:25)."""
    return np.fft.rfftfreq(n=n_fft, d=1.0 / sr)


class ParameterError(Exception):
    """librosa.util.exceptions.ParameterError"""


def stft(y, n_fft, hop_length):
    """librosa.stft defaults as called at src/identify_tracks.py:243,389,654 and
    src/identify_morepork.py:24: periodic Hann, win_length=n_fft, center=True with
    zero padding, 1 + len(y)//hop frames, float64 rfft rounded to complex64.  Like librosa it refuses
    non-finite audio (util.valid_audio), which is what a constant segment becomes after normalize_data."""
    y = np.asarray(y, dtype=np.float32)
    if not np.isfinite(y).all():
        raise ParameterError("Audio buffer is not finite everywhere")
    win = scipy.signal.get_window("hann", n_fft, fftbins=True)
    ypad = np.pad(y, (n_fft // 2, n_fft // 2))
    n_frames = 1 + (len(ypad) - n_fft) // hop_length
    out = np.empty((1 + n_fft // 2, n_frames), dtype=np.complex64)
    block = max(1, (1 << 22) // n_fft)
    idx = np.arange(n_fft)[None, :]
    for b0 in range(0, n_frames, block):
        b1 = min(n_frames, b0 + block)
        starts = (np.arange(b0, b1) * hop_length)[:, None]
        out[:, b0:b1] = scipy.fft.rfft(win[None, :] * ypad[starts + idx], axis=1).T
    return out


def power_to_db(S, ref=1.0, amin=1e-10, top_db=80.0):
    """librosa.power_to_db (src/identify_tracks.py:266 with ref=np.max)."""
    S = np.asarray(S)
    ref_value = ref(S) if callable(ref) else np.abs(ref)
    log_spec = 10.0 * np.log10(np.maximum(amin, S))
    log_spec -= 10.0 * np.log10(np.maximum(amin, ref_value))
    if top_db is not None:
        log_spec = np.maximum(log_spec, log_spec.max() - top_db)
    return log_spec


def amplitude_to_db(S, ref=1.0, amin=1e-5, top_db=80.0):
    """librosa.amplitude_to_db (src/identify_morepork.py:152 with ref=np.max)."""
    magnitude = np.abs(np.asarray(S))
    ref_value = ref(magnitude) if callable(ref) else np.abs(ref)
    return power_to_db(np.square(magnitude), ref=ref_value ** 2, amin=amin ** 2, top_db=top_db)


# --------------------------------------------------------------------------- custommel
def hz_to_mel(frequencies, break_freq):
    """src/custommel.py:6-8."""
    return 2595.0 * np.log10(1.0 + np.array(frequencies) / break_freq)


def mel_frequencies(n_mels, fmin, fmax, break_freq):
    """src/custommel.py:11-16."""
    mels = np.linspace(hz_to_mel(fmin, break_freq), hz_to_mel(fmax, break_freq), n_mels)
    return break_freq * (10.0 ** (mels / 2595.0) - 1.0)


def mel_f(sr, n_mels, fmin, fmax, n_fft, break_freq):
    """src/custommel.py:19-56: triangular filters in float64 stored to float32, then the
    Slaney area normalisation applied in place to the float32 array."""
    n_mels = int(n_mels)
    weights = np.zeros((n_mels, int(1 + n_fft // 2)), dtype=np.float32)
    fftfreqs = fft_frequencies(sr, n_fft)
    pts = mel_frequencies(n_mels + 2, fmin, fmax, break_freq)
    fdiff = np.diff(pts)
    ramps = np.subtract.outer(pts, fftfreqs)
    for i in range(n_mels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        weights[i] = np.maximum(0, np.minimum(lower, upper))
    enorm = 2.0 / (pts[2:n_mels + 2] - pts[:n_mels])
    weights *= enorm[:, np.newaxis]
    return weights


def slaney_mel_filterbank(sr, n_fft, n_mels, fmin, fmax):
    """librosa.filters.mel(htk=False, norm="slaney") as used by librosa.feature.melspectrogram
    (src/identify_tracks.py:230-238); Slaney scale: 200/3 Hz per mel below 1 kHz, log(6.4)/27 per mel above."""
    f_sp, min_log_hz, logstep = 200.0 / 3, 1000.0, np.log(6.4) / 27.0

    def h2m(f):
        return f / f_sp if f < min_log_hz else min_log_hz / f_sp + np.log(f / min_log_hz) / logstep

    mels = np.linspace(h2m(float(fmin)), h2m(float(fmax)), n_mels + 2)
    pts = np.where(mels >= min_log_hz / f_sp, min_log_hz * np.exp(logstep * (mels - min_log_hz / f_sp)), f_sp * mels)
    weights = np.zeros((n_mels, 1 + n_fft // 2), dtype=np.float32)
    fftfreqs = fft_frequencies(sr, n_fft)
    fdiff = np.diff(pts)
    ramps = np.subtract.outer(pts, fftfreqs)
    for i in range(n_mels):
        weights[i] = np.maximum(0, np.minimum(-ramps[i] / fdiff[i], ramps[i + 2] / fdiff[i + 1]))
    weights *= (2.0 / (pts[2:n_mels + 2] - pts[:n_mels]))[:, np.newaxis]
    return weights


def mel_spec(stft_mag, sr, n_fft, hop_length, n_mels, fmin, fmax, break_freq=1750, power=2):
    """src/custommel.py:59-63."""
    magnitude = np.abs(stft_mag) ** power
    return mel_f(sr, n_mels, fmin, fmax, n_fft, break_freq).dot(magnitude)


# --------------------------------------------------------------------- identify_tracks
class Signal:
    """The attributes of identify_tracks.Signal that load_samples reads
    (src/identify_tracks.py:915-986)."""

    def __init__(self, start, end, freq_start, freq_end):
        self.start, self.end, self.freq_start, self.freq_end = start, end, freq_start, freq_end

    @property
    def length(self):
        return self.end - self.start


def normalize_data(x):
    """src/identify_tracks.py:202-209 (note: +1e-6 is added after the divide)."""
    min_v = np.min(x, -1, keepdims=True)
    x = x - min_v
    max_v = np.max(x, -1, keepdims=True)
    x = x / max_v + 0.000001
    x = x - 0.5
    x = x * 2
    return x


def get_spect(data, sr, hop_length, mean_sub, use_mfcc, mel_break, htk, n_mels, fmin, fmax,
              n_fft, power, db_scale, channels=1):
    """src/identify_tracks.py:212-288, use_mfcc=False.  htk=False = librosa.feature.melspectrogram with
    librosa's own Slaney filterbank (restated from librosa 0.11's documented algorithm: unpinned)."""
    assert not use_mfcc, "the use_mfcc branch cannot run in the reference either"
    spectogram = np.abs(stft(data, n_fft, hop_length))
    if not htk:
        mel = slaney_mel_filterbank(sr, n_fft, n_mels, 50, 11000).dot(spectogram ** 2.0)
    else:
        mel = mel_spec(spectogram, sr, n_fft, hop_length, n_mels,
                       50 if fmin is None else fmin, 11000 if fmin is None else fmax,
                       mel_break, power=power)
    if db_scale:
        mel = power_to_db(mel, ref=np.max)
    mel = np.expand_dims(mel, axis=2)
    if mean_sub:
        mel = mel - np.expand_dims(np.mean(mel, axis=1, dtype=mel.dtype), axis=1)
    if channels > 1:
        mel = np.repeat(mel, channels, axis=2)
    return mel


def segment_table(n_frames_total, sr, tracks, segment_length, stride, fmin=50, fmax=11000,
                  pad_short_tracks=False, randint=None):
    """The slicing arithmetic of load_samples (src/identify_tracks.py:109-196) without the
    spectrogram work: returns, per track, a list of (abs_start, n_valid, pad_left) with
    abs_start the absolute sample index of the first real sample, n_valid the number of
    real samples and pad_left the zeros in front (np.pad offset).  `randint` defaults to
    the legacy global np.random.randint the reference consumes (:132,:167)."""
    randint = randint or np.random.randint
    sample_size = int(sr * segment_length)
    out = []
    for t in tracks:
        segs = []
        if t.freq_start > fmax or t.freq_end < fmin:
            out.append(segs)
            continue
        start = 0
        end = start + segment_length
        sr_end = int(t.end * sr)
        sr_start = int(sr * t.start)
        if pad_short_tracks:
            end = min(end, t.length)
        else:
            missing = sample_size - (sr_end - sr_start)
            if missing > 0:
                offset = randint(0, missing)
                sr_start = sr_start - offset
                if sr_start <= 0:
                    sr_start = 0
                    sr_end = min(sr_start + sample_size, n_frames_total)
                else:
                    end_offset = sr_end + missing - offset
                    if end_offset > n_frames_total:
                        end_offset = n_frames_total
                        sr_start = max(end_offset - sample_size, 0)
                    sr_end = end_offset
                assert sr_end - sr_start == sample_size
        # frames[sr_start:sr_end] with numpy clipping semantics
        t0 = min(max(sr_start, 0), n_frames_total)
        t1 = min(max(sr_end, t0), n_frames_total)
        track_len = t1 - t0
        s0 = 0
        s1 = min(sr_end, sample_size)          # quirk: absolute end index (:150-151)
        while True:
            a = min(s0, track_len)
            b = min(max(s1, a), track_len)
            n_valid = b - a
            pad_left = 0
            if n_valid != sample_size:
                extra = sample_size - n_valid
                pad_left = int(randint(0, extra))
            segs.append((t0 + a, n_valid, pad_left))
            start = start + stride
            end = start + segment_length
            s0 = int(start * sr)
            s1 = min(int(end * sr), s0 + sample_size)
            if end > t.length:
                break
        out.append(segs)
    return out


def load_samples(frames, sr, tracks, segment_length, stride, hop_length=281, mean_sub=False,
                 use_mfcc=False, mel_break=1000, htk=True, n_mels=160, fmin=50, fmax=11000,
                 channels=1, power=2, db_scale=False, filter_freqs=False, filter_below=None,
                 normalize=True, n_fft=4096, pad_short_tracks=False):
    """src/identify_tracks.py:65-199 (Butterworth options off, as by default :466-467)."""
    assert not filter_freqs and not filter_below
    frames = np.asarray(frames)
    sample_size = int(sr * segment_length)
    table = segment_table(len(frames), sr, tracks, segment_length, stride, fmin, fmax,
                          pad_short_tracks)
    result = []
    for segs in table:
        track_data = []
        for (a, n_valid, pad_left) in segs:
            data = frames[a:a + n_valid]
            if n_valid != sample_size:
                data = np.pad(data, (pad_left, sample_size - n_valid - pad_left))
            if normalize:
                data = normalize_data(data)
            track_data.append(get_spect(data, sr, hop_length, mean_sub, use_mfcc, mel_break, htk,
                                        n_mels, fmin, fmax, n_fft, power, db_scale, channels))
        result.append(track_data)
    return result


def get_end(frames, sr):
    """src/identify_tracks.py:387-413."""
    hop_length = 281
    spectogram = np.abs(stft(frames, sr // 10, hop_length))
    mel = mel_spec(spectogram, sr, sr // 10, hop_length, 120, 50, 11000, 1750, power=1)
    start = 0
    chunk_length = sr // hop_length
    end = start + chunk_length
    file_length = len(frames) / sr
    while end < mel.shape[1]:
        data = mel[:, start:end]
        if np.amax(data) == np.amin(data):
            return start * hop_length // sr
        start = end
        end = start + chunk_length
    return file_length


def signal_noise_spectrogram(frames, sr, hop_length=281):
    """First stage of signal_noise (src/identify_tracks.py:650-657): |STFT(4096)| / max."""
    spectogram = np.abs(stft(frames, 4096, hop_length))
    return spectogram / np.amax(spectogram)


SIGNAL_WIDTH = 0.25


def signal_noise_stages(frames, sr, hop_length=281):
    """signal_noise (src/identify_tracks.py:650-706) with its intermediates: returns
    (signals, mask, opened, closed, stats) where signals = [(start, end, freq_start, freq_end)],
    mask = the 3x-median threshold (:667), opened = after the 4x4 MORPH_OPEN (:671),
    closed = after dilate (height, width) / erode (height // 10, width) (:683-684), stats = the
    filtered, x-sorted component stats (:686-691).  cv2 is the reference's own dependency."""
    import cv2
    n_fft = 4096
    spectogram = np.abs(stft(frames, n_fft, hop_length))
    a_max = np.amax(spectogram)
    spectogram = spectogram / a_max
    row_medians = np.median(spectogram, axis=1)
    column_medians = np.median(spectogram, axis=0)
    signal = (spectogram > 3 * column_medians[np.newaxis, :]) & (spectogram > 3 * row_medians[:, np.newaxis])
    mask = signal.astype(np.uint8)
    opened = cv2.morphologyEx(mask, cv2.MORPH_OPEN, np.ones((4, 4), np.uint8))
    width = int(SIGNAL_WIDTH * sr / hop_length)
    freqs = fft_frequencies(sr, n_fft)
    height = 0
    for i, f in enumerate(freqs):
        if f > 100:
            height = i + 1
            break
    closed = cv2.dilate(opened, np.ones((height, width), np.uint8))
    closed = cv2.erode(closed, np.ones((height // 10, width), np.uint8))
    _, _, stats, _ = cv2.connectedComponentsWithStats(closed)
    stats = sorted(stats[1:], key=lambda stat: stat[0])
    min_width = 0.65 * width
    min_height = height - height // 10
    stats = [st for st in stats if st[2] > min_width and st[3] > min_height]
    signals = []
    for st in stats:
        max_freq = min(len(freqs) - 1, st[1] + st[3])
        signals.append((st[0] * 281 / sr, (st[0] + st[2]) * 281 / sr, freqs[st[1]], freqs[max_freq]))
    return signals, mask, opened, closed, np.array(stats, dtype=np.int64).reshape(-1, 5)


def signal_noise(frames, sr, hop_length=281):
    return signal_noise_stages(frames, sr, hop_length)[0]


# ------------------------------------------------------------------- identify_morepork
MOREPORK_FMIN, MOREPORK_FMAX = 600, 1200
SLICES_PER_SECOND = 20
SLICES_PER_SAMPLE = 60
SAMPLE_SLIDE_SLICES = 20


def morepork_spectrogram(frames, sr):
    """_load_sample after decode (src/identify_morepork.py:22-27)."""
    nfft = int(sr / 10)
    s = stft(frames, nfft, int(nfft / 2))
    return np.abs(s)[int(MOREPORK_FMIN / 10):int(MOREPORK_FMAX / 10)]


def morepork_windows(npspec):
    """The window loop of identify_morepork (src/identify_morepork.py:142-167)."""
    samples, start_times = [], []
    for base in range(0, npspec.shape[1], SAMPLE_SLIDE_SLICES):
        limit = min(base + SLICES_PER_SAMPLE, npspec.shape[1])
        start = limit - SLICES_PER_SAMPLE
        start_times.append(start / SLICES_PER_SECOND)
        sample = npspec[:, start:limit]
        sample = amplitude_to_db(sample, ref=np.max)
        if sample.min() != 0:
            sample = sample / abs(sample.min()) + 1.0
        if sample.shape[1] < 60:
            sample = np.pad(sample, ((0, 0), (0, 60 - sample.shape[1])))
        samples.append(sample.reshape(sample.shape + (1,)))
    return np.array(samples), start_times


# --------------------------------------------------------------------- cacophony_index
CI_SAMPLE_RATE = 16000
CI_WINDOW = 2048
CI_BIN_20 = 312


def ci_edges(window_size=CI_WINDOW, sample_rate=CI_SAMPLE_RATE):
    """src/cacophony_index.py:58-63."""
    bass_cut_off_band = 100 * 2 * window_size // sample_rate
    return np.logspace(math.log10(bass_cut_off_band), math.log10(window_size), num=11, dtype=int)


def get_ci_bins(source_trim, sample_rate):
    """src/cacophony_index.py:53-66 (numpy.hanning = symmetric Hann, src/common.py:39-40;
    scipy.fftpack.dct type II unnormalised; Python sequential sum of squares)."""
    window_size = source_trim.shape[0]
    signal = np.hanning(window_size) * source_trim
    dct = scipy.fftpack.dct(signal)
    edges = ci_edges(window_size, sample_rate)
    return np.array([sum(x * x) for x in np.split(dct, edges)[1:-1]])


def ci_points(source_data, sample_rate=CI_SAMPLE_RATE, window_size=CI_WINDOW):
    """The hot loop of calculate (src/cacophony_index.py:86-97)."""
    half = window_size // 2
    previous_bins = None
    points = []
    for offset in range(half, source_data.shape[0] - half * 3, half):
        bins = get_ci_bins(source_data[offset:offset + window_size], sample_rate)
        if previous_bins is not None:
            points.append(int(sum(bins * 2 < previous_bins)) + int(sum(bins > previous_bins * 2)))
        previous_bins = bins
    return points


def ci_points_fast(source_data, sample_rate=CI_SAMPLE_RATE, window_size=CI_WINDOW, block=4096):
    """Vectorised equivalent of ci_points (same float64 DCT; band sums via np.add.reduceat,
    whose pairwise order can differ from Python's sequential sum in the last ulp) - used
    only for long recordings where the per-window loop would take minutes."""
    half = window_size // 2
    offsets = np.arange(half, source_data.shape[0] - half * 3, half)
    if offsets.size == 0:
        return []
    win = np.hanning(window_size)
    edges = ci_edges(window_size, sample_rate)
    bins_all = np.empty((offsets.size, 10))
    idx = np.arange(window_size)[None, :]
    for b0 in range(0, offsets.size, block):
        o = offsets[b0:b0 + block, None]
        d = scipy.fftpack.dct(win[None, :] * source_data[o + idx], axis=1)
        sq = d * d
        for j in range(10):
            bins_all[b0:b0 + block, j] = sq[:, edges[j]:edges[j + 1]].sum(axis=1)
    prev, cur = bins_all[:-1], bins_all[1:]
    return list(((cur * 2 < prev).sum(axis=1) + (cur > prev * 2).sum(axis=1)).astype(int))


def score_from_points(points):
    """src/cacophony_index.py:69-73."""
    points_sorted = sorted(points)
    k0 = int(len(points) * 0.75)
    k1 = int(len(points) * 0.95)
    return 10 * np.mean(points_sorted[k0:k1])


def apply_correction_curve_202001C(raw_score):
    """src/cacophony_index.py:76-78."""
    s = raw_score - 10
    return max(100 * s / (s + 18), 0)


def ci_table(points, n_samples, sample_rate=CI_SAMPLE_RATE, window_size=CI_WINDOW):
    """src/cacophony_index.py:99-124."""
    half = window_size // 2
    table = []
    entry_count = (len(points) + 31) // CI_BIN_20
    for e in range(entry_count):
        q = 0
        if e:
            q = e * (len(points) - CI_BIN_20) // (entry_count - 1)
        score = apply_correction_curve_202001C(score_from_points(points[q:q + CI_BIN_20]))
        table.append({"begin_s": round(q * half / sample_rate),
                      "end_s": round((q + CI_BIN_20) * half / sample_rate),
                      "index_percent": round(score, 1)})
    result = {"cacophony_index_old": table, "cacophony_index_old_version": "2020-01-20_A"}
    if table == []:
        result["ci_warning"] = (
            "Cacophony Index requires at least 20 seconds of audio, but only %d seconds of audio were provided."
            % (n_samples / sample_rate))
    return result


def calculate_from_array(source_data, fast=False):
    """cacophony_index.calculate (src/cacophony_index.py:81-125) after the decode step."""
    pts = ci_points_fast(source_data) if fast else ci_points(source_data)
    return ci_table(pts, source_data.shape[0])


# ------------------------------------------------------------------------ magtransform
def magtransform(x, a):
    """MagTransform.call: x ** sigmoid(a) (src/magtransform.py:17-19,
    src/magtransformv2.py:19-21); a=0.0 (v1 init) -> 0.5, a=-1.0 (v2 init) -> 0.2689."""
    a = np.float32(a)
    p = np.float32(1.0) / (np.float32(1.0) + np.exp(-a, dtype=np.float32))
    return np.power(np.asarray(x, dtype=np.float32), p).astype(np.float32)


# --------------------------------------------------------------------- resampling front-end
def resample_poly(x, up, down):
    """The resampling front-end's stated filter (the reference itself uses ffmpeg
    -ar 16000, src/common.py:76-79, and librosa.resample, src/identify_tracks.py:55-57, neither of
    which exists here): scipy.signal.resample_poly with its default Kaiser(5.0) design and zero padding."""
    import scipy.signal
    return scipy.signal.resample_poly(np.asarray(x, dtype=np.float32), up, down)


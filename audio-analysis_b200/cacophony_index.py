"""Drop-in for the reference's cacophony_index module (src/cacophony_index.py): same callables
(get_ci_bins, score_from_points, apply_correction_curve_202001C, calculate) with the hot loop -
Hann window, 2048-point DCT-II, 10 band energies, the 2x comparisons and the per-entry order
statistics - in float64 CUDA kernels (engine.CiPlan).  Only the scalar tail the reference runs
once per ~20 s entry (mean, correction curve, Python round) stays on the host, in the same
numpy/Python types so that the JSON comes out identical.
"""
import functools
import math

import numpy

WINDOW_SIZE = 2048
SAMPLE_RATE = 16000
BIN_20_WIDTH = 312  # ~20 seconds (src/cacophony_index.py:99)
VERSION = "2020-01-20_A"


def _edges(window_size, sample_rate):
    """src/cacophony_index.py:58-63."""
    bass_cut_off_band = 100 * 2 * window_size // sample_rate
    return numpy.logspace(math.log10(bass_cut_off_band), math.log10(window_size), num=11, dtype=int)


@functools.lru_cache(maxsize=4)
def _plan(window_size, sample_rate, device=0):
    from . import engine
    # numpy.hanning: common.get_window_const(window_size, "hanning") (src/common.py:39-40)
    return engine.CiPlan(_edges(window_size, sample_rate), numpy.hanning(window_size), device=device)


def get_ci_bins(source_trim, sample_rate):
    """src/cacophony_index.py:53-66: float64[10] band energies of one window."""
    import torch
    source_trim = numpy.ascontiguousarray(source_trim, dtype=numpy.float32)
    plan = _plan(int(source_trim.shape[0]), int(sample_rate))
    pcm = torch.from_numpy(source_trim).cuda(plan.device)
    return plan.bins(pcm, 1, 0).cpu().numpy()[0]


def score_from_points(points):
    """src/cacophony_index.py:69-73 (host scalar version, for callers that hold a points list)."""
    points_sorted = sorted(points)
    k0 = int(len(points) * 0.75)
    k1 = int(len(points) * 0.95)
    return 10 * numpy.mean(points_sorted[k0:k1])


def apply_correction_curve_202001C(raw_score):
    """src/cacophony_index.py:76-78."""
    s = raw_score - 10
    return max(100 * s / (s + 18), 0)


def entry_layout(n_points):
    """Start index and length of every 312-point entry (src/cacophony_index.py:101-107)."""
    entry_count = (n_points + 31) // BIN_20_WIDTH
    starts, lens = [], []
    for e in range(entry_count):
        q = 0
        if e:
            q = e * (n_points - BIN_20_WIDTH) // (entry_count - 1)
        starts.append(q)
        lens.append(max(0, min(q + BIN_20_WIDTH, n_points) - q))
    return starts, lens


def _table_from_stats(starts, stats, sample_rate=SAMPLE_RATE, window_size=WINDOW_SIZE):
    half_window_size = window_size // 2
    table = []
    for q, (total, count) in zip(starts, stats):
        # numpy.mean of the sorted int slice == float64 sum / count (exact: small integers)
        raw_score = 10 * (numpy.float64(int(total)) / int(count)) if count else 10 * numpy.float64("nan")
        score = apply_correction_curve_202001C(raw_score)
        table.append({"begin_s": round(q * half_window_size / sample_rate),
                      "end_s": round((q + BIN_20_WIDTH) * half_window_size / sample_rate),
                      "index_percent": round(score, 1)})
    return table


def _result(table, n_samples, sample_rate=SAMPLE_RATE):
    result = {"cacophony_index_old": table, "cacophony_index_old_version": VERSION}
    if table == []:
        p = n_samples / sample_rate
        result["ci_warning"] = (
            "Cacophony Index requires at least 20 seconds of audio, but only %d seconds of audio were provided." % p)
    return result


def calculate_batch(pcm, item_offset, item_samples, device=0):
    """calculate() for many 16 kHz recordings held in one flat CUDA float32 tensor: returns one
    result dict per recording.  All sample-touching work and the order statistics run on the GPU."""
    plan = _plan(WINDOW_SIZE, SAMPLE_RATE, device)
    _, points, base = plan.run(pcm, item_offset, item_samples)
    starts_all, lens_all, owner = [], [], []
    for i in range(len(item_offset)):
        n_points = max(0, int(base[i + 1] - base[i]) - 1)
        starts, lens = entry_layout(n_points)
        owner.append((len(starts_all), starts))
        starts_all += [int(base[i]) + 1 + q for q in starts]       # +1: the first window only seeds previous_bins
        lens_all += lens
    stats = plan.entries(points, starts_all, lens_all) if starts_all else numpy.zeros((0, 2), numpy.int32)
    out = []
    for i, (k, starts) in enumerate(owner):
        table = _table_from_stats(starts, stats[k:k + len(starts)])
        out.append(_result(table, int(item_samples[i])))
    return out


def calculate(source_file_name, loader=None):
    """src/cacophony_index.py:81-125.  `source_file_name` may also be an already decoded 16 kHz
    float32 array; decoding a path uses `loader(path, 16000)` (default: the reference's
    common.load_audio_file_as_numpy_array, i.e. ffmpeg - I/O outside this package)."""
    import torch
    if isinstance(source_file_name, (str, bytes)) or hasattr(source_file_name, "__fspath__"):
        if loader is None:
            import common  # the reference's helper module (src/common.py:70-96)
            loader = common.load_audio_file_as_numpy_array
        source_data = loader(source_file_name, SAMPLE_RATE)
    else:
        source_data = source_file_name
    source_data = numpy.ascontiguousarray(source_data, dtype=numpy.float32)
    plan = _plan(WINDOW_SIZE, SAMPLE_RATE)
    pcm = torch.from_numpy(source_data).cuda(plan.device)
    return calculate_batch(pcm, [0], [source_data.shape[0]], plan.device)[0]


class Stream:
    """Long-form recordings fed in chunks (BASELINE.json configs[3]): the carried state is the tail of samples from
    the last finished window on (that window is recomputed to seed previous_bins) plus the points so far.  feed()
    accepts chunks of any size - numpy float32 arrays, page-locked torch CPU tensors (uploaded as they are) or CUDA
    tensors; result() returns exactly what calculate() returns on the concatenation.
    The carry lives on the DEVICE and nothing synchronises before result(): a chunk goes up on a copy stream (numpy
    chunks through two alternating page-locked staging buffers) while the kernels of the previous chunk run."""

    def __init__(self, device=0):
        import torch
        self._torch = torch
        self.plan = _plan(WINDOW_SIZE, SAMPLE_RATE, device)
        self.device = device
        self._carry = None             # device tensor: samples from _carry_start on
        self._carry_start = 0          # absolute sample index of _carry[0]
        self._next_window = 0          # index of the next window to compute (offset 1024 * (w + 1))
        self._total = 0
        self._points = []              # CUDA uint8 tensors
        self._copy_stream = None       # created with the first upload
        self._stage = [None, None]     # page-locked staging buffers for pageable chunks
        self._stage_done = [None, None]
        self._k = 0

    def _upload(self, chunk):
        """chunk -> CUDA float32 tensor, the copy queued on the copy stream; the compute stream waits for it."""
        torch = self._torch
        device = torch.device("cuda", self.device)
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=device)
        if hasattr(chunk, "is_cuda") and chunk.is_cuda:
            return chunk.to(device, torch.float32).reshape(-1)
        if isinstance(chunk, numpy.ndarray):
            chunk = numpy.ascontiguousarray(chunk, dtype=numpy.float32).reshape(-1)
            k = self._k
            self._k ^= 1
            if self._stage_done[k] is not None:
                self._stage_done[k].synchronize()              # the copy that last used this staging buffer is done
            if self._stage[k] is None or self._stage[k].numel() < chunk.shape[0]:
                self._stage[k] = torch.empty(chunk.shape[0], dtype=torch.float32).pin_memory()
            src = self._stage[k][:chunk.shape[0]]
            src.numpy()[:] = chunk
        else:
            src, k = chunk.reshape(-1), None                   # torch CPU tensor (page-locked: truly asynchronous)
        cur = torch.cuda.current_stream(device)
        with torch.cuda.stream(self._copy_stream):
            dev = src.to(device, non_blocking=True)
            done = torch.cuda.Event()
            done.record(self._copy_stream)
        if k is not None:
            self._stage_done[k] = done
        cur.wait_event(done)
        dev.record_stream(cur)
        return dev

    def feed(self, chunk):
        torch = self._torch
        dev_chunk = self._upload(chunk)
        buf = torch.cat([self._carry, dev_chunk]) if self._carry is not None and self._carry.numel() else dev_chunk
        self._total += int(dev_chunk.numel())
        n_final = self.plan.window_count(self._total)          # windows that exist whatever follows
        if n_final <= self._next_window:
            self._carry = buf
            return
        # recompute the previous window (if any) so that the first new window has its previous_bins
        w0 = max(self._next_window - 1, 0)
        # a pseudo-recording starting 1024 samples before window w0: its windows are w0, w0 + 1, ...
        start = 1024 * w0 - self._carry_start
        n_new = n_final - w0
        n_samples = 1024 * (n_new + 2) + 2048                  # window_count(n_samples) == n_new
        assert self.plan.window_count(n_samples) == n_new
        if start + n_samples > buf.numel():                    # only padding beyond the last window's end
            buf = torch.cat([buf, torch.zeros(start + n_samples - buf.numel(), dtype=torch.float32, device=buf.device)])
        _, points, _ = self.plan.run(buf, [start], [n_samples])
        self._points.append(points[1:])                        # points[0] = 255: window w0 only seeds previous_bins
        self._next_window = n_final
        keep_from = 1024 * (n_final - 1) - self._carry_start   # next pseudo-recording starts at window n_final - 1
        self._carry = buf[keep_from:self._total - self._carry_start].clone()
        self._carry_start += keep_from

    def result(self):
        torch = self._torch
        if not self._points:
            return _result([], self._total)
        points = torch.cat(self._points)
        starts, lens = entry_layout(int(points.numel()))
        stats = self.plan.entries(points, starts, lens) if starts else numpy.zeros((0, 2), numpy.int32)
        return _result(_table_from_stats(starts, stats), self._total)

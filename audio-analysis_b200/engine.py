"""ctypes binding of libcacophony_b200.so (include/cacophony_b200.h).

PyTorch is used only for device memory and streams (tensors' data_ptr() are handed to the
C-ABI).  There is no CPU fallback: a missing library raises ImportError here, and every
compute call raises CacError when no CUDA device is present.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libcacophony_b200.so")

if not os.path.isfile(LIB_PATH):
    raise ImportError(
        "libcacophony_b200.so is not built (run `python -c 'import __graft_entry__ as g; g.build()'` "
        "or audio-analysis_b200/csrc/build.sh); there is no CPU fallback")

lib = ctypes.CDLL(LIB_PATH)


class CacParams(ctypes.Structure):
    _fields_ = [("sample_rate", ctypes.c_int32), ("n_fft", ctypes.c_int32), ("hop_length", ctypes.c_int32),
                ("n_mels", ctypes.c_int32), ("segment_samples", ctypes.c_int32), ("power", ctypes.c_int32),
                ("db_scale", ctypes.c_int32), ("normalize", ctypes.c_int32)]


class CacSegment(ctypes.Structure):
    _fields_ = [("pcm_offset", ctypes.c_int64), ("n_valid", ctypes.c_int32), ("pad_left", ctypes.c_int32)]


SEGMENT_DTYPE = np.dtype([("pcm_offset", "<i8"), ("n_valid", "<i4"), ("pad_left", "<i4")])
assert SEGMENT_DTYPE.itemsize == ctypes.sizeof(CacSegment) == 16

_vp, _i64, _i32 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32
lib.cac_version.restype = ctypes.c_char_p
lib.cac_last_error.restype = ctypes.c_char_p
lib.cac_device_count.restype = ctypes.c_int
lib.cac_plan_create.argtypes = [ctypes.POINTER(CacParams), _vp, ctypes.c_int, ctypes.POINTER(_vp)]
lib.cac_plan_destroy.argtypes = [_vp]
lib.cac_plan_frames.argtypes = [_vp]
lib.cac_plan_set_profiling.argtypes = [_vp, ctypes.c_int]
lib.cac_plan_last_kernel_ms.argtypes = [_vp, ctypes.POINTER(ctypes.c_float)]
lib.cac_measure_fp32_tflops.argtypes = [ctypes.c_int, ctypes.POINTER(ctypes.c_float)]
lib.cac_plan_set_magtransform.argtypes = [_vp, ctypes.c_int, ctypes.c_float]
lib.cac_features_work_bytes.argtypes = [_vp, _i64]
lib.cac_features_work_bytes.restype = ctypes.c_size_t
lib.cac_features_run.argtypes = [_vp, _vp, _vp, _i64, _vp, _vp, _vp]
lib.cac_features_host.argtypes = [_vp, _vp, _i64, _vp, _i64, _vp]
lib.cac_features_host_pcm16.argtypes = [_vp, _vp, _i64, _vp, _i64, _vp]
lib.cac_normalize_run.argtypes = [_vp, _vp, _i64, _i32, _vp]
lib.cac_mel_project_run.argtypes = [_vp, _vp, _i32, _vp, _vp]
lib.cac_power_to_db_run.argtypes = [_vp, _vp, _i64, _i64, _vp, _vp]
lib.cac_power_to_db_static.argtypes = [_vp, _vp, _i64, _i64, _vp]
POWER_TO_DB_STATIC_MAX_ITEMS = 65536      # CAC_POWER_TO_DB_STATIC_MAX_ITEMS

lib.cac_ci_plan_create.argtypes = [_i32, _vp, _vp, ctypes.c_int, ctypes.POINTER(_vp)]
lib.cac_ci_plan_destroy.argtypes = [_vp]
lib.cac_ci_window_count.argtypes = [_i64]
lib.cac_ci_window_count.restype = _i64
lib.cac_ci_bins_run.argtypes = [_vp, _vp, _i64, _i32, _vp, _vp]
lib.cac_ci_run.argtypes = [_vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp]
lib.cac_ci_host.argtypes = [_vp, _vp, _i64, _vp, _vp, _i64, _vp, _vp]
lib.cac_ci_entries_run.argtypes = [_vp, _vp, _vp, _vp, _i64, _vp, _vp]



class CacSpecParams(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int32) for n in ("n_fft", "hop_length", "bin_lo", "bin_hi", "n_mels", "power",
                                              "bin_major")]


lib.cac_spec_plan_create.argtypes = [ctypes.POINTER(CacSpecParams), _vp, ctypes.c_int, ctypes.POINTER(_vp)]
lib.cac_spec_plan_destroy.argtypes = [_vp]
lib.cac_spec_frame_count.argtypes = [_vp, _i64]
lib.cac_spec_frame_count.restype = _i64
lib.cac_spec_rows.argtypes = [_vp]
lib.cac_spec_run.argtypes = [_vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp]
lib.cac_scale_by_max_run.argtypes = [_vp, _i64, _vp, _vp]
lib.cac_chunk_minmax_run.argtypes = [_vp, _i32, _i64, _i32, _vp, _vp]
lib.cac_morepork_windows_run.argtypes = [_vp, _vp, _vp, _vp, _i64, _vp, _vp]

lib.cac_sn_plan_create.argtypes = [_i32, _i32, _i32, _i32, ctypes.c_int, ctypes.POINTER(_vp)]
lib.cac_sn_plan_destroy.argtypes = [_vp]
lib.cac_sn_frame_count.argtypes = [_vp, _i64]
lib.cac_sn_frame_count.restype = _i64
lib.cac_sn_work_bytes.argtypes = [_vp, _i64, _i64]
lib.cac_sn_work_bytes.restype = ctypes.c_size_t
lib.cac_sn_run.argtypes = [_vp, _vp, _vp, _vp, _i64, _i64, _i32, _vp, _vp, ctypes.c_size_t, _vp]
lib.cac_sn_spectrum_run.argtypes = [_vp, _vp, _i64, _vp, _vp, _vp, _vp, ctypes.c_size_t, _vp]
lib.cac_signal_components_work_bytes.argtypes = [_i32, _i64]
lib.cac_signal_components_work_bytes.restype = ctypes.c_size_t
lib.cac_signal_components_run.argtypes = [_vp, _i32, _i64, _i32, _i32, _i32, _i32, _vp, _vp, _vp, ctypes.c_size_t, _vp]
lib.cac_get_end_work_bytes.argtypes = [_vp, _i64, _i64, _i32]
lib.cac_get_end_work_bytes.restype = ctypes.c_size_t
lib.cac_get_end_run.argtypes = [_vp, _vp, _vp, _vp, _i64, _i64, _i32, _vp, _vp, ctypes.c_size_t, _vp]
lib.cac_tracks_merge.argtypes = [_vp, _i32, _vp, ctypes.POINTER(ctypes.c_int32)]
lib.cac_host_register.argtypes = [_vp, ctypes.c_size_t]
lib.cac_host_unregister.argtypes = [_vp]
lib.cac_magtransform_run.argtypes = [_vp, _vp, _i64, ctypes.c_float, _vp]
lib.cac_spect_tail_run.argtypes = [_vp, _i64, _i32, _i32, _i32, _vp, _vp]
lib.cac_resample_plan_create.argtypes = [_i32, _i32, _vp, _i32, _i32, ctypes.c_int, ctypes.POINTER(_vp)]
lib.cac_resample_plan_destroy.argtypes = [_vp]
lib.cac_resample_out_len.argtypes = [_vp, _i64]
lib.cac_resample_out_len.restype = _i64
lib.cac_resample_run.argtypes = [_vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp]

EXPORTS = ["cac_resample_plan_create", "cac_resample_plan_destroy", "cac_resample_out_len", "cac_resample_run", "cac_spect_tail_run", "cac_sn_plan_create", "cac_sn_plan_destroy", "cac_sn_frame_count", "cac_sn_work_bytes", "cac_sn_run", "cac_sn_spectrum_run", "cac_signal_components_work_bytes", "cac_signal_components_run", "cac_get_end_work_bytes", "cac_get_end_run", "cac_host_register", "cac_host_unregister", "cac_magtransform_run", "cac_spec_plan_create", "cac_spec_plan_destroy", "cac_spec_frame_count", "cac_spec_rows", "cac_spec_run",
           "cac_scale_by_max_run", "cac_chunk_minmax_run", "cac_morepork_windows_run", "cac_ci_plan_create", "cac_ci_plan_destroy", "cac_ci_window_count", "cac_ci_bins_run", "cac_ci_run",
           "cac_ci_host", "cac_ci_entries_run", "cac_version", "cac_last_error", "cac_device_count", "cac_plan_create", "cac_plan_destroy",
           "cac_plan_frames", "cac_measure_fp32_tflops", "cac_plan_set_magtransform", "cac_plan_set_profiling", "cac_plan_last_kernel_ms", "cac_features_work_bytes", "cac_features_run", "cac_features_host", "cac_features_host_pcm16",
           "cac_normalize_run", "cac_mel_project_run", "cac_power_to_db_run", "cac_power_to_db_static", "cac_tracks_merge"]


class CacError(RuntimeError):
    pass


def check(rc):
    if rc != 0:
        raise CacError("cacophony_b200 error %d: %s" % (rc, lib.cac_last_error().decode()))


def version():
    return lib.cac_version().decode()


def measure_fp32_tflops(device=0):
    """FP32 FFMA throughput of the device measured now (cac_measure_fp32_tflops): TFLOP/s."""
    _torch()
    v = ctypes.c_float()
    check(lib.cac_measure_fp32_tflops(int(device), ctypes.byref(v)))
    return float(v.value)


def _torch():
    import torch
    if not torch.cuda.is_available():
        raise CacError("no CUDA device: cacophony_b200 has no CPU path")
    return torch


def _stream_ptr(torch):
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


class FeaturePlan:
    """STFT -> custom mel -> dB plan (cac_plan_*) for one DSP parameter set."""

    def __init__(self, mel_weights, sr, n_fft, hop_length, n_mels, segment_samples, power=2, db_scale=True,
                 normalize=True, device=0):
        w = np.ascontiguousarray(mel_weights, dtype=np.float32)
        assert w.shape == (n_mels, 1 + n_fft // 2), w.shape
        self.params = CacParams(int(sr), int(n_fft), int(hop_length), int(n_mels), int(segment_samples),
                                int(power), int(bool(db_scale)), int(bool(normalize)))
        self.device = device
        self._h = ctypes.c_void_p()
        check(lib.cac_plan_create(ctypes.byref(self.params), w.ctypes.data_as(_vp), device, ctypes.byref(self._h)))
        self.n_frames = lib.cac_plan_frames(self._h)
        self.n_mels = int(n_mels)
        self._work = None

    def close(self):
        if self._h:
            lib.cac_plan_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_profiling(self, enable=True):
        check(lib.cac_plan_set_profiling(self._h, int(enable)))

    def set_magtransform(self, a=None):
        """Fuse MagTransform (y = x ** sigmoid(a)) into the kernel's output stage; a=None switches it off."""
        check(lib.cac_plan_set_magtransform(self._h, int(a is not None), float(a or 0.0)))

    def last_kernel_ms(self):
        ms = ctypes.c_float()
        check(lib.cac_plan_last_kernel_ms(self._h, ctypes.byref(ms)))
        return ms.value

    # ---- device-resident path -----------------------------------------------------------
    def run(self, pcm, segments, out=None):
        """pcm: float32 CUDA tensor (flat); segments: CUDA uint8/int tensor holding cac_segment
        records (or a numpy SEGMENT_DTYPE array, uploaded here); returns [n, n_mels, frames]."""
        torch = _torch()
        dev = torch.device("cuda", self.device)
        if isinstance(segments, np.ndarray):
            assert segments.dtype == SEGMENT_DTYPE
            n = segments.shape[0]
            segments = self._stage_segments(torch, dev, segments)
        else:
            n = segments.numel() * segments.element_size() // 16
        assert pcm.is_cuda and pcm.dtype == torch.float32 and pcm.is_contiguous()
        if out is None:
            out = torch.empty((n, self.n_mels, self.n_frames), dtype=torch.float32, device=dev)
        need = lib.cac_features_work_bytes(self._h, n)
        if self._work is None or self._work.numel() < need:
            self._work = torch.empty(need, dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            check(lib.cac_features_run(self._h, pcm.data_ptr(), segments.data_ptr(), n, out.data_ptr(),
                                       self._work.data_ptr(), _stream_ptr(torch)))
        return out

    def _stage_segments(self, torch, dev, segments):
        """Descriptor table -> device through a page-locked staging buffer and an ASYNCHRONOUS copy: a pageable upload
        would make the host wait for everything already queued on the stream (classify_prep_batch queues the next chunk's
        kernels before it gets here)."""
        raw = np.ascontiguousarray(segments).view(np.uint8).reshape(-1)
        if getattr(self, "_seg_done", None) is not None:
            self._seg_done.synchronize()                   # the previous copy out of the staging buffer has finished
        if getattr(self, "_seg_pin", None) is None or self._seg_pin.numel() < raw.size:
            self._seg_pin = torch.empty(max(raw.size, 1 << 16), dtype=torch.uint8).pin_memory()
        self._seg_pin[:raw.size].numpy()[:] = raw
        out = self._seg_pin[:raw.size].to(dev, non_blocking=True)
        self._seg_done = torch.cuda.Event()
        self._seg_done.record()
        return out

    def last_segment_max(self, n):
        """Per-segment maximum mel power of the last run() over n segments (the power_to_db reference
        value, kept in the work buffer): CUDA float32 view [n] - a small per-segment result to gather."""
        torch = _torch()
        off = ((n * 8 + 127) // 128) * 128
        return self._work[off:off + 4 * n].view(torch.float32)

    # ---- host-buffer path (H2D + run + D2H inside the call) ----------------------------------
    def run_host(self, pcm, segments, out=None):
        """Host arrays in, host array out.  int16 PCM (as decoded) is uploaded as is and widened on the device
        to sample / 32768 (cac_features_host_pcm16); anything else is taken as float32."""
        pcm16 = isinstance(pcm, np.ndarray) and pcm.dtype == np.int16
        pcm = np.ascontiguousarray(pcm, dtype=np.int16 if pcm16 else np.float32).reshape(-1)
        segments = np.ascontiguousarray(segments, dtype=SEGMENT_DTYPE)
        n = segments.shape[0]
        if out is None:
            out = np.empty((n, self.n_mels, self.n_frames), dtype=np.float32)
        fn = lib.cac_features_host_pcm16 if pcm16 else lib.cac_features_host
        check(fn(self._h, pcm.ctypes.data_as(_vp), pcm.size, segments.ctypes.data_as(_vp), n, out.ctypes.data_as(_vp)))
        return out

    def mel_project(self, mag):
        """custommel.mel_spec on a magnitude matrix [bins, frames] (numpy or CUDA tensor)."""
        torch = _torch()
        dev = torch.device("cuda", self.device)
        was_np = isinstance(mag, np.ndarray)
        t = torch.as_tensor(np.ascontiguousarray(mag, dtype=np.float32) if was_np else mag).to(dev).contiguous()
        out = torch.empty((self.n_mels, t.shape[1]), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            check(lib.cac_mel_project_run(self._h, t.data_ptr(), t.shape[1], out.data_ptr(), _stream_ptr(torch)))
        return out.cpu().numpy() if was_np else out


def normalize_rows(x, device=0):
    """normalize_data on the last axis (bit-exact float32 op order)."""
    torch = _torch()
    dev = torch.device("cuda", device)
    was_np = isinstance(x, np.ndarray)
    t = torch.as_tensor(np.ascontiguousarray(x, dtype=np.float32) if was_np else x).to(dev).contiguous()
    out = torch.empty_like(t)
    rows = t.numel() // t.shape[-1]
    with torch.cuda.device(dev):
        check(lib.cac_normalize_run(t.data_ptr(), out.data_ptr(), rows, t.shape[-1], _stream_ptr(torch)))
    return out.cpu().numpy() if was_np else out


def power_to_db(x, device=0):
    """librosa.power_to_db(ref=np.max, top_db=80) per leading-axis item."""
    torch = _torch()
    dev = torch.device("cuda", device)
    was_np = isinstance(x, np.ndarray)
    t = torch.as_tensor(np.ascontiguousarray(x, dtype=np.float32) if was_np else x).to(dev).contiguous()
    n_items = t.shape[0] if t.dim() > 2 else 1
    out = torch.empty_like(t)
    work = torch.empty(n_items * 4 + 16, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(lib.cac_power_to_db_run(t.data_ptr(), out.data_ptr(), n_items, t.numel() // n_items, work.data_ptr(),
                                      _stream_ptr(torch)))
    return out.cpu().numpy() if was_np else out


class CiPlan:
    """Old Cacophony Index plan (cac_ci_*): float64 DCT band energies, points and score order statistics."""

    def __init__(self, edges, window, device=0):
        edges = np.ascontiguousarray(edges, dtype=np.int32)
        window = np.ascontiguousarray(window, dtype=np.float64)
        assert edges.shape == (11,)
        self.device = device
        self._h = ctypes.c_void_p()
        check(lib.cac_ci_plan_create(int(window.shape[0]), edges.ctypes.data_as(_vp), window.ctypes.data_as(_vp),
                                     device, ctypes.byref(self._h)))

    def close(self):
        if self._h:
            lib.cac_ci_plan_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @staticmethod
    def window_count(n_samples):
        return int(lib.cac_ci_window_count(int(n_samples)))

    def bins(self, pcm, n_windows, hop):
        """get_ci_bins on windows pcm[hop*w : hop*w + 2048] of a CUDA float32 tensor -> CUDA float64 [n, 10]."""
        torch = _torch()
        dev = torch.device("cuda", self.device)
        assert pcm.is_cuda and pcm.dtype == torch.float32 and pcm.is_contiguous()
        assert n_windows == 0 or hop * (n_windows - 1) + 2048 <= pcm.numel()
        out = torch.empty((n_windows, 10), dtype=torch.float64, device=dev)
        with torch.cuda.device(dev):
            check(lib.cac_ci_bins_run(self._h, pcm.data_ptr(), n_windows, hop, out.data_ptr(), _stream_ptr(torch)))
        return out

    def run(self, pcm, item_offset, item_samples, want_points=True):
        """calculate()'s loop on recordings stored in the flat CUDA tensor `pcm`; returns
        (bins [W, 10] float64, points [W] uint8 or None, win_base [n_items + 1] int64 numpy)."""
        torch = _torch()
        dev = torch.device("cuda", self.device)
        off = np.ascontiguousarray(item_offset, dtype=np.int64)
        ns = np.ascontiguousarray(item_samples, dtype=np.int64)
        assert pcm.is_cuda and pcm.dtype == torch.float32 and pcm.is_contiguous()
        assert off.shape == ns.shape and (off + ns <= pcm.numel()).all()
        counts = np.array([self.window_count(v) for v in ns], dtype=np.int64)
        base = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
        total = int(base[-1])
        bins = torch.empty((total, 10), dtype=torch.float64, device=dev)
        points = torch.empty((total,), dtype=torch.uint8, device=dev) if want_points else None
        with torch.cuda.device(dev):
            check(lib.cac_ci_run(self._h, pcm.data_ptr(), off.ctypes.data_as(_vp), ns.ctypes.data_as(_vp), len(off),
                                 bins.data_ptr(), points.data_ptr() if want_points else None, _stream_ptr(torch)))
        return bins, points, base

    def run_host(self, pcm, item_offset, item_samples):
        pcm = np.ascontiguousarray(pcm, dtype=np.float32).reshape(-1)
        off = np.ascontiguousarray(item_offset, dtype=np.int64)
        ns = np.ascontiguousarray(item_samples, dtype=np.int64)
        counts = np.array([self.window_count(v) for v in ns], dtype=np.int64)
        base = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
        total = int(base[-1])
        bins = np.empty((total, 10), dtype=np.float64)
        points = np.empty((total,), dtype=np.uint8)
        check(lib.cac_ci_host(self._h, pcm.ctypes.data_as(_vp), pcm.size, off.ctypes.data_as(_vp),
                              ns.ctypes.data_as(_vp), len(off), bins.ctypes.data_as(_vp), points.ctypes.data_as(_vp)))
        return bins, points, base

    def entries(self, points, entry_start, entry_len):
        """Order statistics of score_from_points for slices of the CUDA uint8 tensor `points`:
        returns int32 numpy [n_entries, 2] = (sum of sorted[k0:k1], k1 - k0)."""
        torch = _torch()
        dev = torch.device("cuda", self.device)
        es = np.ascontiguousarray(entry_start, dtype=np.int64)
        el = np.ascontiguousarray(entry_len, dtype=np.int32)
        out = torch.empty((len(es), 2), dtype=torch.int32, device=dev)
        if len(es):
            with torch.cuda.device(dev):
                check(lib.cac_ci_entries_run(self._h, points.data_ptr(), es.ctypes.data_as(_vp), el.ctypes.data_as(_vp),
                                             len(es), out.data_ptr(), _stream_ptr(torch)))
        return out.cpu().numpy()


def stft_frame_counts(n_samples, n_fft, hop_length):
    """Frames librosa.stft(center=True) produces for recordings of n_samples samples (array or scalar):
    1 + (n + 2 (n_fft // 2) - n_fft) // hop, i.e. 1 + n // hop for even n_fft and 1 + (n - 1) // hop for odd n_fft
    (0 for an empty recording).  Same rule as cac_spec_frame_count."""
    ns = np.asarray(n_samples, dtype=np.int64)
    return np.maximum(0, 1 + (ns + 2 * (int(n_fft) // 2) - int(n_fft)) // int(hop_length))


class SpecPlan:
    """Framed-spectrum plan (cac_spec_*): librosa.stft framing -> |X|**power of selected bins,
    optionally through a custommel filterbank."""

    def __init__(self, n_fft, hop_length, bin_lo=0, bin_hi=0, n_mels=0, power=1, bin_major=False, mel_weights=None,
                 device=0):
        self.params = CacSpecParams(int(n_fft), int(hop_length), int(bin_lo), int(bin_hi), int(n_mels), int(power),
                                    int(bool(bin_major)))
        w = None
        if n_mels:
            w = np.ascontiguousarray(mel_weights, dtype=np.float32)
            assert w.shape == (n_mels, 1 + n_fft // 2), w.shape
        self.device = device
        self.hop_length = int(hop_length)
        self._h = ctypes.c_void_p()
        check(lib.cac_spec_plan_create(ctypes.byref(self.params), w.ctypes.data_as(_vp) if w is not None else None,
                                       device, ctypes.byref(self._h)))
        self.rows = int(lib.cac_spec_rows(self._h))
        self._work = None

    def close(self):
        if self._h:
            lib.cac_spec_plan_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def frame_count(self, n_samples):
        return int(lib.cac_spec_frame_count(self._h, int(n_samples)))   # librosa's count, also for odd n_fft

    def run(self, pcm, item_offset, item_samples, out=None, want_max=False):
        """pcm: flat CUDA float32 tensor.  Returns (out, frame_base[, max tensor]); `out` is a flat
        CUDA float32 tensor of rows * total_frames values laid out as the plan's layout says."""
        torch = _torch()
        dev = torch.device("cuda", self.device)
        off = np.ascontiguousarray(item_offset, dtype=np.int64)
        ns = np.ascontiguousarray(item_samples, dtype=np.int64)
        assert pcm.is_cuda and pcm.dtype == torch.float32 and pcm.is_contiguous()
        assert off.shape == ns.shape and (off + ns <= pcm.numel()).all()
        base = np.concatenate([[0], np.cumsum(stft_frame_counts(ns, self.params.n_fft, self.hop_length))]).astype(np.int64)
        total = int(base[-1])
        if out is None:
            out = torch.empty((total * self.rows,), dtype=torch.float32, device=dev)
        assert out.numel() >= total * self.rows
        mx = torch.zeros((1,), dtype=torch.float32, device=dev) if want_max else None
        with torch.cuda.device(dev):
            check(lib.cac_spec_run(self._h, pcm.data_ptr(), off.ctypes.data_as(_vp), ns.ctypes.data_as(_vp), len(off),
                                   out.data_ptr(), mx.data_ptr() if want_max else None, _stream_ptr(torch)))
        return (out, base, mx) if want_max else (out, base)

    def get_end_chunks(self, pcm, item_offset, item_samples, chunk_frames, max_item_samples=None):
        """cac_get_end_run: for every recording of the flat CUDA float32 tensor `pcm` the index of the first
        chunk of `chunk_frames` frames whose mel values are all equal, -1 if there is none; CUDA int32 [n], queued
        on the current stream (tables: numpy / list, or CUDA int64 tensors together with max_item_samples)."""
        torch = _torch()
        dev = torch.device("cuda", self.device)
        assert pcm.is_cuda and pcm.dtype == torch.float32 and pcm.is_contiguous()
        off, ns, mx = _item_tables(pcm, item_offset, item_samples)
        if max_item_samples is None:
            assert mx is not None, "device item tables need max_item_samples"
            max_item_samples = mx
        n = int(off.numel())
        out = torch.empty((n,), dtype=torch.int32, device=dev)
        if n == 0:
            return out
        need = lib.cac_get_end_work_bytes(self._h, n, int(max_item_samples), int(chunk_frames))
        work = _work_buffer(self, need, dev)
        with torch.cuda.device(dev):
            check(lib.cac_get_end_run(self._h, pcm.data_ptr(), off.data_ptr(), ns.data_ptr(), n, int(max_item_samples),
                                      int(chunk_frames), out.data_ptr(), work.data_ptr(), work.numel(), _stream_ptr(torch)))
        return out

    def morepork_windows(self, spec, win_frame0, win_width, out=None):
        torch = _torch()
        dev = torch.device("cuda", self.device)
        f0 = np.ascontiguousarray(win_frame0, dtype=np.int64)
        wd = np.ascontiguousarray(win_width, dtype=np.int32)
        if out is None:
            out = torch.empty((len(f0), self.rows, 60), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            check(lib.cac_morepork_windows_run(self._h, spec.data_ptr(), f0.ctypes.data_as(_vp), wd.ctypes.data_as(_vp),
                                               len(f0), out.data_ptr(), _stream_ptr(torch)))
        return out


def chunk_minmax(x, chunk):
    """(min, max) of x[:, c*chunk:(c+1)*chunk] for every complete chunk of a CUDA float32 [rows, cols] tensor."""
    torch = _torch()
    assert x.is_cuda and x.dtype == torch.float32 and x.is_contiguous() and x.dim() == 2
    n = x.shape[1] // chunk
    out = torch.empty((n, 2), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        check(lib.cac_chunk_minmax_run(x.data_ptr(), x.shape[0], x.shape[1], chunk, out.data_ptr(), _stream_ptr(torch)))
    return out


def scale_by_max(x, mx):
    torch = _torch()
    with torch.cuda.device(x.device):
        check(lib.cac_scale_by_max_run(x.data_ptr(), x.numel(), mx.data_ptr(), _stream_ptr(torch)))
    return x


def magtransform(x, a, device=0):
    """x ** sigmoid(a) (MagTransform.call) on a numpy array or CUDA float32 tensor."""
    torch = _torch()
    dev = torch.device("cuda", device)
    was_np = isinstance(x, np.ndarray)
    t = torch.as_tensor(np.ascontiguousarray(x, dtype=np.float32) if was_np else x).to(dev).contiguous()
    out = torch.empty_like(t)
    with torch.cuda.device(dev):
        check(lib.cac_magtransform_run(t.data_ptr(), out.data_ptr(), t.numel(), float(a), _stream_ptr(torch)))
    return out.cpu().numpy() if was_np else out


class ResamplePlan:
    """cac_resample_plan: polyphase FIR resampling by up / down with the caller's (padded) filter."""

    def __init__(self, up, down, taps, n_pre_remove, device=0):
        taps = np.ascontiguousarray(taps, dtype=np.float32)
        self.up, self.down, self.device = int(up), int(down), int(device)
        self._h = _vp()
        check(lib.cac_resample_plan_create(self.up, self.down, taps.ctypes.data_as(_vp), len(taps), int(n_pre_remove),
                                           self.device, ctypes.byref(self._h)))

    def __del__(self):
        if getattr(self, "_h", None):
            lib.cac_resample_plan_destroy(self._h)
            self._h = None

    def out_len(self, n_in):
        return int(lib.cac_resample_out_len(self._h, int(n_in)))

    def run(self, pcm, in_offset, in_samples, out=None, out_offset=None):
        """Resample items of the flat CUDA float32 tensor `pcm`; returns (out, out_offset [n + 1] int64):
        item i occupies out[out_offset[i] : out_offset[i + 1]] unless explicit out_offset are given."""
        torch = _torch()
        off = np.ascontiguousarray(in_offset, dtype=np.int64)
        ns = np.ascontiguousarray(in_samples, dtype=np.int64)
        assert pcm.is_cuda and pcm.dtype == torch.float32 and pcm.is_contiguous()
        assert off.shape == ns.shape and (off + ns <= pcm.numel()).all()
        lens = np.array([self.out_len(v) for v in ns], dtype=np.int64)
        if out_offset is None:
            out_offset = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
        oo = np.ascontiguousarray(out_offset, dtype=np.int64)
        need = int((oo[:len(lens)] + lens).max()) if len(lens) else 0
        if out is None:
            out = torch.empty((need,), dtype=torch.float32, device=pcm.device)
        assert out.is_cuda and out.dtype == torch.float32 and out.is_contiguous() and out.numel() >= need
        with torch.cuda.device(pcm.device):
            check(lib.cac_resample_run(self._h, pcm.data_ptr(), off.ctypes.data_as(_vp), ns.ctypes.data_as(_vp), len(off),
                                       out.data_ptr(), oo.ctypes.data_as(_vp), _stream_ptr(torch)))
        return out, oo


def spect_tail(block, mean_sub=False, channels=1):
    """get_spect's tail on a CUDA float32 [n, n_mels, frames] block: returns [n, n_mels, frames, channels]
    (mean over frames subtracted per mel row when mean_sub), still on the device."""
    torch = _torch()
    assert block.is_cuda and block.dtype == torch.float32 and block.is_contiguous() and block.dim() == 3
    n, n_mels, frames = block.shape
    channels = int(channels)
    if not mean_sub and channels == 1:
        return block.unsqueeze(-1)
    out = block if channels == 1 else torch.empty((n, n_mels, frames, channels), dtype=torch.float32, device=block.device)
    with torch.cuda.device(block.device):
        check(lib.cac_spect_tail_run(block.data_ptr(), n * n_mels, frames, int(bool(mean_sub)), channels,
                                     out.data_ptr(), _stream_ptr(torch)))
    return out.view(n, n_mels, frames, channels)


def bind_to_device_numa(device=0):
    """Pin the calling process to the CPUs next to CUDA device `device` (NVML's ideal affinity) so that the
    page-locked buffers it allocates afterwards live on that GPU's NUMA node.  One process per GPU that feeds
    its GPU through *_host calls should call this first: without it two ranks can end up streaming through
    the same socket's memory and the inter-socket link.  Returns the CPU list, or None when NVML cannot tell."""
    torch = _torch()
    try:
        import pynvml
        pynvml.nvmlInit()
        uuid = str(torch.cuda.get_device_properties(device).uuid)
        if not uuid.startswith("GPU-"):
            uuid = "GPU-" + uuid
        try:
            h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByUUID(uuid)
        n_words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, n_words)
        cpus = [64 * w + b for w, word in enumerate(mask) for b in range(64) if (int(word) >> b) & 1]
        allowed = sorted(set(cpus) & set(os.sched_getaffinity(0)))
        if not allowed:
            return None
        os.sched_setaffinity(0, allowed)
        return allowed
    except Exception:
        return None


class pinned:
    """Context manager page-locking a numpy array for the duration of *_host calls (cac_host_register)."""

    def __init__(self, arr):
        self.arr = arr

    def __enter__(self):
        check(lib.cac_host_register(self.arr.ctypes.data_as(_vp), self.arr.nbytes))
        return self.arr

    def __exit__(self, *exc):
        lib.cac_host_unregister(self.arr.ctypes.data_as(_vp))
        return False


def _work_buffer(holder, need, dev):
    """A 256-byte aligned CUDA scratch tensor of at least `need` bytes cached on `holder` (grown, never shrunk)."""
    torch = _torch()
    buf = getattr(holder, "_work", None)
    if buf is None or buf.numel() < need or buf.device != dev:
        buf = torch.empty(int(need), dtype=torch.uint8, device=dev)      # torch allocations are 512-byte aligned
        holder._work = buf
    return buf


def _item_tables(pcm, item_offset, item_samples):
    """(offsets, lengths) as CUDA int64 tensors + the longest length; accepts numpy / lists (uploaded) or CUDA tensors."""
    torch = _torch()
    if hasattr(item_offset, "is_cuda") and item_offset.is_cuda:
        assert item_samples.is_cuda and item_offset.dtype == item_samples.dtype == torch.int64
        return item_offset, item_samples, None
    off = np.ascontiguousarray(item_offset, dtype=np.int64)
    ns = np.ascontiguousarray(item_samples, dtype=np.int64)
    assert off.shape == ns.shape and off.ndim == 1
    assert len(off) == 0 or ((off >= 0).all() and (ns >= 0).all() and int((off + ns).max()) <= pcm.numel())
    both = torch.from_numpy(np.stack([off, ns])).to(pcm.device, non_blocking=True)
    return both[0], both[1], (int(ns.max()) if len(ns) else 0)


def collect_components(result_host):
    """[count, -, -, -, (min x, min y, max x, max y, area, order) ...] block (numpy int32, downloaded) ->
    [n, 6] = (x, y, width, height, area, order), cv2's stats layout + the label-order key."""
    count = int(result_host[0])
    max_components = (result_host.size - 4) // 6
    if count > max_components:
        raise CacError("more connected components (%d) than max_components (%d)" % (count, max_components))
    raw = result_host[4:4 + 6 * count].reshape(count, 6)
    out = raw.copy()
    out[:, 2] = raw[:, 2] - raw[:, 0] + 1
    out[:, 3] = raw[:, 3] - raw[:, 1] + 1
    return out


class SignalNoisePlan:
    """cac_sn_*: signal_noise (src/identify_tracks.py:650-691) up to the connected-component table, for batches of
    recordings on the device.  dilate_h / dilate_w / erode_h are the reference's height / width / height // 10."""

    def __init__(self, hop_length, dilate_h, dilate_w, erode_h, device=0):
        self.hop_length, self.device = int(hop_length), int(device)
        self._h = _vp()
        self._work = None
        check(lib.cac_sn_plan_create(self.hop_length, int(dilate_h), int(dilate_w), int(erode_h), self.device,
                                     ctypes.byref(self._h)))

    def __del__(self):
        if getattr(self, "_h", None):
            lib.cac_sn_plan_destroy(self._h)
            self._h = None

    def frame_count(self, n_samples):
        return int(lib.cac_sn_frame_count(self._h, int(n_samples)))

    def run(self, pcm, item_offset, item_samples, max_components=4096, max_item_samples=None, result=None):
        """Queue the whole chain for every recording pcm[off_i : off_i + n_i] of the flat CUDA float32 tensor `pcm`
        (tables: numpy / list, or CUDA int64 tensors together with max_item_samples).  Returns the CUDA int32 tensor
        [n, 4 + 6 * max_components]; nothing is synchronised."""
        torch = _torch()
        dev = torch.device("cuda", self.device)
        assert pcm.is_cuda and pcm.dtype == torch.float32 and pcm.is_contiguous()
        off, ns, mx = _item_tables(pcm, item_offset, item_samples)
        if max_item_samples is None:
            assert mx is not None, "device item tables need max_item_samples"
            max_item_samples = mx
        n = int(off.numel())
        if result is None:
            result = torch.empty((n, 4 + 6 * int(max_components)), dtype=torch.int32, device=dev)
        assert result.is_cuda and result.dtype == torch.int32 and result.is_contiguous() and result.numel() >= n * (4 + 6 * int(max_components))
        if n == 0:
            return result
        need = lib.cac_sn_work_bytes(self._h, n, int(max_item_samples))
        work = _work_buffer(self, need, dev)
        with torch.cuda.device(dev):
            check(lib.cac_sn_run(self._h, pcm.data_ptr(), off.data_ptr(), ns.data_ptr(), n, int(max_item_samples),
                                 int(max_components), result.data_ptr(), work.data_ptr(), work.numel(), _stream_ptr(torch)))
        return result

    def spectrum(self, pcm):
        """First stage for one recording (1-D CUDA float32 tensor): (|stft| [2049, cols] unscaled, column medians [cols],
        a_max [1]) as CUDA tensors."""
        torch = _torch()
        dev = torch.device("cuda", self.device)
        assert pcm.is_cuda and pcm.dtype == torch.float32 and pcm.is_contiguous() and pcm.dim() == 1
        cols = self.frame_count(pcm.numel())
        pitch = (cols + 31) // 32 * 32                            # rows padded to 32 floats (aligned 48-byte runs)
        spec = torch.empty((2049, pitch), dtype=torch.float32, device=dev)
        colmed = torch.empty((cols,), dtype=torch.float32, device=dev)
        amax = torch.zeros((1,), dtype=torch.float32, device=dev)
        work = _work_buffer(self, 4096, dev)
        with torch.cuda.device(dev):
            check(lib.cac_sn_spectrum_run(self._h, pcm.data_ptr(), pcm.numel(), spec.data_ptr(), colmed.data_ptr(),
                                          amax.data_ptr(), work.data_ptr(), work.numel(), _stream_ptr(torch)))
        return spec[:, :cols], colmed, amax


class _ComponentsScratch:
    _work = None


def signal_components(spec, dilate_h, dilate_w, erode_h, max_components=65536, want_masks=False):
    """cac_signal_components_run on a CUDA float32 [rows, cols] normalised spectrogram: returns the int32
    numpy array [n, 6] = (x, y, width, height, area, order) (and the three stage masks if asked)."""
    torch = _torch()
    assert spec.is_cuda and spec.dtype == torch.float32 and spec.is_contiguous() and spec.dim() == 2
    rows, cols = spec.shape
    result = torch.empty((4 + 6 * int(max_components),), dtype=torch.int32, device=spec.device)
    masks = torch.empty((3, rows, cols), dtype=torch.uint8, device=spec.device) if want_masks else None
    need = lib.cac_signal_components_work_bytes(rows, cols)
    work = _work_buffer(_ComponentsScratch, need, spec.device)
    with torch.cuda.device(spec.device):
        check(lib.cac_signal_components_run(spec.data_ptr(), rows, cols, int(dilate_h), int(dilate_w), int(erode_h),
                                            int(max_components), result.data_ptr(),
                                            masks.data_ptr() if want_masks else None, work.data_ptr(), work.numel(),
                                            _stream_ptr(torch)))
    stats = collect_components(result.cpu().numpy())
    return (stats, masks) if want_masks else stats

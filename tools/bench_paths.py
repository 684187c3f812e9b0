#!/usr/bin/env python
"""Timings of the paths next to the headline (BASELINE.json configs[0], [2], [3] and the per-recording classify-prep
pipeline).  bench.py imports these functions and puts their dictionaries into the `paths` object of its JSON line;
run directly it prints one JSON line per path:  python tools/bench_paths.py [single] [morepork] [ci] [pipeline]
Every function returns {ms, recording_hours_per_s, algorithmic_bytes, hbm_frac, dominant_kernel, cpu_baseline, ...};
device work is timed with CUDA events on the current stream after warm-up, host-inclusive legs with the wall clock
between synchronisations.  CPU baselines are the oracle port on a bounded sample (seconds), one process."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from audio_analysis_b200 import synth  # noqa: E402

S1 = dict(hop_length=281, mel_break=1750, n_mels=160, fmin=50, fmax=11000, power=2, db_scale=True, normalize=True,
          n_fft=4096)


def hbm_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0          # B200_PROFILING.md fallback


def timed(fn, warmup=3, steps=10):
    """mean ms per call, CUDA events on the current stream"""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def wall(fn, warmup=1, steps=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        out = fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / steps * 1e3, out


def _pool(seed0, n, sr=48000):
    return np.stack([synth.recording(seed0 + i, 60.0, sr) for i in range(n)])


def path_ci(hours=8.0, cpu=True):
    """configs[3]: one 8-hour 16 kHz night recording; device-resident kernels, the whole calculate_batch call, and the
    same recording STREAMED from page-locked host memory in 8-minute chunks (upload inside the timed region)."""
    from audio_analysis_b200 import cacophony_index as C
    pool = _pool(900, 16, 16000)
    n_min = int(hours * 60)
    host = torch.from_numpy(pool).repeat((n_min + 15) // 16, 1)[:n_min].contiguous().reshape(-1).pin_memory()
    pcm = host.cuda()
    n = pcm.numel()
    plan = C._plan(2048, 16000)
    ms = timed(lambda: plan.run(pcm, [0], [n]))
    call_ms, res = wall(lambda: C.calculate_batch(pcm, [0], [n])[0])
    n_win = plan.window_count(n)
    chunk = 8 * 60 * 16000

    def streamed():
        st = C.Stream()
        for p in range(0, n, chunk):
            st.feed(host[p:p + chunk])
        return st.result()
    stream_ms, res2 = wall(streamed)
    assert res2 == res
    algo = n * 4 + n_win * (80 + 1)
    out = {"config": "configs[3]: %.0f h at 16 kHz (%d samples), cacophony_index per ~20 s entry" % (hours, n),
           "windows": n_win, "entries": len(res["cacophony_index_old"]), "ms": ms,
           "recording_hours_per_s": hours / (ms * 1e-3), "algorithmic_bytes": algo,
           "hbm_frac": algo / ms / 1e6 / hbm_peak(), "dominant_kernel": "ci_bins_kernel", "dtype": "f64",
           "calculate_batch_call_ms": call_ms,
           "streamed": {"ms": stream_ms, "recording_hours_per_s": hours / (stream_ms * 1e-3), "chunks": (n + chunk - 1) // chunk,
                        "h2d_bytes": n * 4, "h2d_gbs": n * 4 / stream_ms / 1e6,
                        "api": "cacophony_index.Stream.feed on page-locked 8-minute chunks + result(); equals the one-shot table"}}
    if cpu:
        from oracle import oracle_np as O
        x = np.tile(pool[0], 10)
        t0 = time.perf_counter()
        O.calculate_from_array(x)
        w = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": 600.0 / 3600.0 / w, "unit": "recording-hours/s", "cores": 1, "kind": "port",
                               "sample": "cacophony_index loop on a 10-minute 16 kHz slice, %.1f s" % w}
    del pcm, host
    return out


def path_morepork(n_rec=4096, cpu=True):
    """configs[2]: identify_morepork's spectrogram (n_fft 4800, hop 2400, bins 60:120) + 60 x 60 windows."""
    from audio_analysis_b200 import identify_morepork as M
    pool = _pool(300, 16)
    n_rec = int(os.environ.get("CAC_MOREPORK_RECORDINGS", n_rec))
    pcm = torch.from_numpy(pool).cuda().repeat((n_rec + 15) // 16, 1)[:n_rec].contiguous()
    plan = M.BatchPlan(48000, pcm.shape[1])
    out = None

    def step():
        nonlocal out
        out = plan.run(pcm, out=out)[0]
    ms = timed(step, steps=5)
    algo = pcm.numel() * 4 + out.numel() * 4
    res = {"config": "configs[2]: %d x 60 s 48 kHz recordings, morepork spectrogram + windows" % n_rec, "ms": ms,
           "recording_hours_per_s": n_rec / 60.0 / (ms * 1e-3), "algorithmic_bytes": algo,
           "hbm_frac": algo / ms / 1e6 / hbm_peak(),
           "dominant_kernel": "spec4800_kernel<75>" if os.environ.get("CAC_BAND_GEMM", "1")[:1] == "0" else
           "band_gemm_kernel (tcgen05.mma kind::tf32, 3-product split, FP32 accumulate in TMEM)", "dtype": "f32"}
    if cpu:
        from oracle import oracle_np as O
        t0 = time.perf_counter()
        for i in range(8):
            O.morepork_windows(O.morepork_spectrogram(pool[i], 48000))
        w = time.perf_counter() - t0
        res["cpu_baseline"] = {"value": 8 * 60.0 / 3600.0 / w, "unit": "recording-hours/s", "cores": 1, "kind": "port",
                               "sample": "_load_sample STFT + window loop on 8 one-minute recordings, %.1f s" % w}
    del pcm, out
    plan._spec = None
    return res


def path_pipeline(n_rec=256, cpu=True):
    """The per-recording classify prep (src/identify_tracks.py:416-440, 502-524): get_end -> signal_noise on the trimmed
    recording -> get_tracks_from_signals (host) -> load_samples S1 on the detected tracks, for n_rec device-resident
    one-minute recordings (identify_tracks.classify_prep_batch), with the phases timed separately."""
    from audio_analysis_b200 import identify_tracks as T
    pool = _pool(700, 16)
    n_rec = int(os.environ.get("CAC_PIPELINE_RECORDINGS", n_rec))
    pcm = torch.from_numpy(pool).cuda().repeat((n_rec + 15) // 16, 1)[:n_rec].contiguous()
    n = pcm.shape[1]
    offs = np.arange(n_rec, dtype=np.int64) * n
    lens = np.full(n_rec, n, dtype=np.int64)
    tabs = torch.from_numpy(np.stack([offs, lens])).cuda()
    sn = T._signal_noise_plan(48000, 281)
    ge = T._get_end_plan(48000)
    flat = pcm.view(-1)
    ge_ms = timed(lambda: ge.get_end_chunks(flat, tabs[0], tabs[1], 48000 // 281, max_item_samples=n), steps=5)
    sn_ms = timed(lambda: sn.run(flat, tabs[0], tabs[1], max_components=4096, max_item_samples=n), steps=5)
    import gc
    gc.collect()                                           # start the wall-clock region without another path's garbage
    np.random.seed(0)
    total_ms, res = wall(lambda: T.classify_prep_batch(pcm, 48000, 3.0, 1.5, **S1), warmup=2, steps=4)
    n_tracks = sum(len(r["tracks"]) for r in res)
    n_segs = sum(int(s.shape[0]) for r in res if r["samples"] for s in r["samples"])
    out_bytes = n_segs * 160 * 513 * 4
    algo = pcm.numel() * 4 + out_bytes
    out = {"config": "pipeline: %d x 60 s 48 kHz recordings, get_end -> signal_noise -> get_tracks_from_signals (host) -> "
                     "load_samples S1 on the detected tracks" % n_rec,
           "ms": total_ms, "ms_per_recording": total_ms / n_rec, "recording_hours_per_s": n_rec / 60.0 / (total_ms * 1e-3),
           "tracks": n_tracks, "segments": n_segs, "algorithmic_bytes": algo, "hbm_frac": algo / total_ms / 1e6 / hbm_peak(),
           "phases_ms_per_recording": {"get_end_batch": ge_ms / n_rec, "signal_noise_batch": sn_ms / n_rec,
                                       "host_merge_and_load_samples": max(0.0, total_ms - ge_ms - sn_ms) / n_rec},
           "dominant_kernel": "sn_stft_kernel" if sn_ms >= ge_ms else "spec4800_kernel<75>", "dtype": "f32",
           "synchronisations": "one per 24-recording chunk; the host merge of chunk i overlaps the kernels of chunk i + 1"}
    del res
    if cpu:
        from oracle import oracle_np as O
        from audio_analysis_b200 import tracks as TR
        t0 = time.perf_counter()
        for i in range(2):
            x = pool[i]
            length = O.get_end(x, 48000)
            sig = O.signal_noise(x[: int(48000 * length)], 48000)
            tr = TR.get_tracks_from_signals([TR.Signal(*s) for s in sig], length)
            if tr:
                O.load_samples(x, 48000, [O.Signal(t.start, t.end, t.freq_start, t.freq_end) for t in tr], 3.0, 1.5, **S1)
        w = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": 2 * 60.0 / 3600.0 / w, "unit": "recording-hours/s", "cores": 1, "kind": "port",
                               "sample": "the same chain through the oracle on 2 one-minute recordings, %.1f s" % w}
    del pcm
    return out


def path_single(cpu=True):
    """configs[0]: ONE 60 s 48 kHz recording, host array in: classify prep with one upload and two synchronisations
    (identify_tracks.classify_prep), the 48 -> 16 kHz resampling and cacophony_index.calculate; next to the same work
    through the separate drop-in calls (each uploads and synchronises on its own)."""
    from audio_analysis_b200 import identify_tracks as T, cacophony_index as C, resample as R
    rec = synth.recording(4242, 60.0, 48000)
    whole = [T.Signal(0, 60.0, 0, 24000)]
    kw = dict(S1, segment_length=3.0, segment_stride=1.5)

    def chained():
        dev = T.to_device(rec)
        prep = T.classify_prep(dev, 48000, **kw)
        ci = C.calculate_batch(R.to_16k(dev), [0], [rec.shape[0] // 3])[0]
        return prep, ci
    np.random.seed(0)
    chained_ms, (prep, ci) = wall(chained, warmup=2, steps=10)
    sep = {}
    sep["get_end"], _ = wall(lambda: T.get_end(rec, 48000), 1, 10)
    sep["signal_noise"], _ = wall(lambda: T.signal_noise(rec, 48000), 1, 10)
    sep["load_samples_39_segments"], _ = wall(lambda: T.load_samples(rec, 48000, whole, 3.0, 1.5, **S1), 1, 10)
    sep["to_16k"], rec16 = wall(lambda: R.to_16k(rec), 1, 10)
    sep["cacophony_index"], _ = wall(lambda: C.calculate(rec16), 1, 10)
    n_segs = sum(int(s.shape[0]) for s in (prep["samples"] or []))
    out = {"config": "configs[0]: one 60 s 48 kHz recording, host array in, single process",
           "ms": chained_ms, "recording_hours_per_s": 60.0 / 3600.0 / (chained_ms * 1e-3),
           "api": "to_device + classify_prep (get_end -> signal_noise -> tracks -> load_samples on the device) + to_16k + "
                  "cacophony_index.calculate_batch: one upload, 3 synchronisations",
           "tracks": len(prep["tracks"]), "segments": n_segs, "ci_entries": len(ci["cacophony_index_old"]),
           "algorithmic_bytes": rec.nbytes + n_segs * 160 * 513 * 4,
           "hbm_frac": (rec.nbytes + n_segs * 160 * 513 * 4) / chained_ms / 1e6 / hbm_peak(),
           "dominant_kernel": "latency-bound (launch + synchronisation); largest kernel sn_stft_kernel",
           "separate_calls_ms": {k: round(v, 3) for k, v in sep.items()}, "separate_calls_total_ms": round(sum(sep.values()), 3),
           "dtype": "f32"}
    if cpu:
        from oracle import oracle_np as O
        from audio_analysis_b200 import tracks as TR
        t0 = time.perf_counter()
        length = O.get_end(rec, 48000)
        sig = O.signal_noise(rec[: int(48000 * length)], 48000)
        tr = TR.get_tracks_from_signals([TR.Signal(*s) for s in sig], length)
        if tr:
            O.load_samples(rec, 48000, [O.Signal(t.start, t.end, t.freq_start, t.freq_end) for t in tr], 3.0, 1.5, **S1)
        O.calculate_from_array(O.resample_poly(rec, 1, 3))
        w = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": 60.0 / 3600.0 / w, "unit": "recording-hours/s", "cores": 1, "kind": "port",
                               "sample": "the same recording through the oracle, %.1f s" % w}
    return out


PATHS = {"single": path_single, "morepork": path_morepork, "ci": path_ci, "pipeline": path_pipeline}

if __name__ == "__main__":
    for name in sys.argv[1:] or ["ci"]:
        print(json.dumps({name: PATHS[name](cpu=os.environ.get("CAC_CPU_BASELINE", "1") != "0")}), flush=True)

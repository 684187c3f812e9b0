#!/usr/bin/env python
"""Device-resident timings of the secondary paths (BASELINE.json configs[2] and [3]); bench.py keeps the
headline configs[1].  Prints one JSON line per path.  usage: python tools/bench_paths.py [ci] [morepork]"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from audio_analysis_b200 import synth  # noqa: E402

PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
    os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0


def timed(fn, warmup=3, steps=10):
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


def bench_ci(hours=8.0):
    from audio_analysis_b200 import cacophony_index as C
    pool = np.stack([synth.recording(900 + i, 60.0, 16000) for i in range(16)])
    n_min = int(hours * 60)
    reps = (n_min + len(pool) - 1) // len(pool)
    pcm = torch.from_numpy(pool).cuda().repeat(reps, 1)[:n_min].contiguous().reshape(-1)
    n = pcm.numel()
    plan = C._plan(2048, 16000)
    ms = timed(lambda: plan.run(pcm, [0], [n]))
    t0 = time.perf_counter()
    res = C.calculate_batch(pcm, [0], [n])[0]
    wall = time.perf_counter() - t0
    n_win = plan.window_count(n)
    cpu = None
    if os.environ.get("CAC_CPU_BASELINE", "1") != "0":       # the oracle port on a 10-minute slice, one process
        from oracle import oracle_np as O
        x = np.tile(pool[0], 10)
        t0 = time.perf_counter()
        O.calculate_from_array(x)
        w = time.perf_counter() - t0
        cpu = {"value": 600.0 / 3600.0 / w, "unit": "recording-hours/s", "cores": 1, "kind": "port",
               "sample": "cacophony_index loop on a 10-minute 16 kHz slice, %.1f s wall" % w}
    algo = n * 4 + n_win * (80 + 1)
    print(json.dumps({"path": "cacophony_index (configs[3])", "recording_hours": hours, "samples": n, "windows": n_win,
                      "entries": len(res["cacophony_index_old"]), "ms_bins_plus_points": ms,
                      "recording_hours_per_s": hours / (ms * 1e-3), "algorithmic_GBps": algo / ms / 1e6,
                      "hbm_frac_of_measured": algo / ms / 1e6 / PEAK, "calculate_batch_wall_ms": wall * 1e3,
                      "dtype": "f64", "cpu_baseline": cpu}), flush=True)


def bench_morepork(n_rec=4096):
    from audio_analysis_b200 import identify_morepork as M
    pool = np.stack([synth.recording(300 + i, 60.0, 48000) for i in range(16)])
    n_rec = int(os.environ.get("CAC_MOREPORK_RECORDINGS", n_rec))
    reps = (n_rec + len(pool) - 1) // len(pool)
    pcm = torch.from_numpy(pool).cuda().repeat(reps, 1)[:n_rec].contiguous()
    plan = M.BatchPlan(48000, pcm.shape[1])
    out = None

    def step():
        nonlocal out
        out = plan.run(pcm, out=out)[0]
    ms = timed(step, steps=5)
    algo = pcm.numel() * 4 + out.numel() * 4
    cpu = None
    if os.environ.get("CAC_CPU_BASELINE", "1") != "0":       # the oracle port on 8 recordings, one process
        from oracle import oracle_np as O
        t0 = time.perf_counter()
        for i in range(8):
            O.morepork_windows(O.morepork_spectrogram(pool[i], 48000))
        w = time.perf_counter() - t0
        cpu = {"value": 8 * 60.0 / 3600.0 / w, "unit": "recording-hours/s", "cores": 1, "kind": "port",
               "sample": "_load_sample STFT + window loop on 8 one-minute recordings, %.1f s wall" % w}
    print(json.dumps({"path": "identify_morepork (configs[2])", "recordings": n_rec, "ms": ms, "cpu_baseline": cpu,
                      "recording_hours_per_s": n_rec / 60.0 / (ms * 1e-3), "algorithmic_GBps": algo / ms / 1e6,
                      "hbm_frac_of_measured": algo / ms / 1e6 / PEAK, "dtype": "f32"}), flush=True)


def bench_hop281(n_rec=64):
    """The two whole-recording spectra of identify_tracks.classify (src/identify_tracks.py:419-420): get_end's
    120-mel magnitude spectrogram (n_fft 4800, hop 281) and signal_noise's full magnitude spectrum (n_fft 4096,
    hop 281, [2049, frames] per recording), device-resident, n_rec one-minute recordings per launch."""
    from audio_analysis_b200 import engine, custommel
    pool = np.stack([synth.recording(500 + i, 60.0, 48000) for i in range(16)])
    n_rec = int(os.environ.get("CAC_HOP281_RECORDINGS", n_rec))
    reps = (n_rec + len(pool) - 1) // len(pool)
    pcm = torch.from_numpy(pool).cuda().repeat(reps, 1)[:n_rec].contiguous()
    n = pcm.shape[1]
    offs, lens = np.arange(n_rec, dtype=np.int64) * n, np.full(n_rec, n, dtype=np.int64)
    w = custommel._mel_f_cached(48000, 120, 50, 11000, 4800, 1750)
    plans = {"get_end_mel_spectrogram": engine.SpecPlan(4800, 281, n_mels=120, power=1, mel_weights=w),
             "signal_noise_spectrum": engine.SpecPlan(4096, 281, bin_lo=0, bin_hi=2049, n_mels=0, power=1, bin_major=True)}
    for name, plan in plans.items():
        out = plan.run(pcm.view(-1), offs, lens)[0]
        ms = timed(lambda: plan.run(pcm.view(-1), offs, lens, out=out), steps=5)
        algo = pcm.numel() * 4 + out.numel() * 4
        print(json.dumps({"path": name, "recordings": n_rec, "ms": ms, "recording_hours_per_s": n_rec / 60.0 / (ms * 1e-3),
                          "algorithmic_GBps": algo / ms / 1e6, "hbm_frac_of_measured": algo / ms / 1e6 / PEAK,
                          "dtype": "f32"}), flush=True)
        del out


def bench_signal_noise_batch(n_rec=32):
    """Track finding (signal_noise, src/identify_tracks.py:650-706) for a batch of one-minute recordings that are
    already on the device: spectrum, scaling, medians, bit-packed morphology, run-based components, one
    synchronisation for all component tables (identify_tracks.signal_noise_batch); wall clock."""
    from audio_analysis_b200 import identify_tracks as T
    pool = np.stack([synth.recording(700 + i, 60.0, 48000) for i in range(16)])
    pcm = torch.from_numpy(pool).cuda().repeat((n_rec + 15) // 16, 1)[:n_rec].contiguous()
    T.signal_noise_batch(pcm[:2], 48000)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = T.signal_noise_batch(pcm, 48000)
    wall = time.perf_counter() - t0
    t0 = time.perf_counter()
    for i in range(4):
        T.signal_noise(pcm[i], 48000)
    single = (time.perf_counter() - t0) / 4
    print(json.dumps({"path": "signal_noise_batch", "recordings": n_rec, "ms_per_recording": wall / n_rec * 1e3,
                      "ms_per_recording_single_calls": single * 1e3, "recording_hours_per_s": n_rec / 60.0 / wall,
                      "signals_first_recording": len(out[0]), "dtype": "f32"}), flush=True)


if __name__ == "__main__":
    which = sys.argv[1:] or ["ci"]
    if "signal_noise" in which:
        bench_signal_noise_batch()
    if "hop281" in which:
        bench_hop281()
    if "ci" in which:
        bench_ci()
    if "morepork" in which:
        bench_morepork()

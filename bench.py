#!/usr/bin/env python
"""Benchmark of the feature hot path (BASELINE.json metric: recording-hours/sec, STFT->mel->dB).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one pass of load_samples' inner loop (normalise -> STFT 4096/281 -> custom mel
160 -> power_to_db) over the workload of BASELINE.json configs[1]: 1,024 one-minute 48 kHz
synthetic recordings PER GPU, one whole-file track each = 39 segments -> [39936,160,513] f32.
Multi-GPU = whole recordings sharded across ranks (weak scaling, no data-path collective).
Prints ONE JSON line on rank 0.  `--impl reference` times the CPU oracle port of the
reference path (oracle/oracle_np.py = the reference's numpy/scipy arithmetic) on host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SR = 48000
REC_SECONDS = 60.0
N_REC = int(os.environ.get("CAC_BENCH_RECORDINGS", "1024"))   # per GPU
POOL = int(os.environ.get("CAC_BENCH_POOL", "64"))            # unique synthetic recordings per rank
WAVES = int(os.environ.get("CAC_BENCH_WAVES", "1"))           # passes over the per-GPU batch per step (configs[4]:
#                                                              8 GPUs x 8 waves x 1,024 = 65,536 recordings per step)
S1 = dict(hop_length=281, mel_break=1750, n_mels=160, fmin=50, fmax=11000, power=2, db_scale=True,
          normalize=True, n_fft=4096)
SEG_LEN, STRIDE = 3.0, 1.5
ALGO_BYTES_PER_REC = 11_520_000 + 39 * 160 * 513 * 4          # unique PCM in + dB out
FLOP_PER_REC = 2.65e9
FP32_PEAK_FALLBACK_TFLOPS = 73.8    # only if the in-run measurement (engine.measure_fp32_tflops) fails


def _gen(seed):
    from audio_analysis_b200 import synth
    return synth.recording(seed, REC_SECONDS, SR)


def make_pool(rank, n):
    import multiprocessing as mp
    workers = max(1, min(n, (os.cpu_count() or 8) // max(1, int(os.environ.get("WORLD_SIZE", "1")))))
    seeds = [1000 + rank * 100003 + i for i in range(n)]
    if workers == 1:
        return np.stack([_gen(s) for s in seeds])
    with mp.get_context("fork").Pool(workers) as p:
        return np.stack(p.map(_gen, seeds))


def _oracle_one(seed):
    from oracle import oracle_np as O
    x = _gen(seed)
    t0 = time.perf_counter()
    res = O.load_samples(x, SR, [O.Signal(0, REC_SECONDS, 0, 24000)], SEG_LEN, STRIDE, **S1)
    assert len(res[0]) == 39
    return time.perf_counter() - t0


def _one_thread_per_worker():
    """Pool initializer: one BLAS / OpenMP thread per worker process.  Without it every worker's sgemm
    (custommel.mel_spec's dense dot) spawns a thread per core and `cores` workers oversubscribe the host so badly
    that the pool is SLOWER than one process - an unfairly low CPU baseline."""
    try:
        from threadpoolctl import threadpool_limits
        global _TP_LIMIT
        _TP_LIMIT = threadpool_limits(limits=1)
    except Exception:
        for k in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
            os.environ[k] = "1"


def cpu_oracle_rate(n_recordings, workers):
    """rec-hours/s of the CPU oracle on `n_recordings` one-minute recordings with `workers` processes
    (one compute thread each; a single worker keeps the libraries' default threading, which is how the reference runs)."""
    import multiprocessing as mp
    seeds = [5000 + i for i in range(n_recordings)]
    if workers == 1:
        t0 = time.perf_counter()
        for s in seeds:
            _oracle_one(s)
        wall = time.perf_counter() - t0
    else:
        with mp.get_context("fork").Pool(workers, initializer=_one_thread_per_worker) as p:
            p.map(_gen, seeds[:workers])          # warm the workers (imports) outside the timing
            t0 = time.perf_counter()
            p.map(_oracle_one, seeds, chunksize=1)
            wall = time.perf_counter() - t0
    return n_recordings * REC_SECONDS / 3600.0 / wall, wall


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "100"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        rows = [r.split(",") for r in open(self.f.name).read().strip().splitlines() if r.count(",") >= 5]
        os.unlink(self.f.name)
        sm = [float(r[0]) for r in rows if r[0].strip().replace(".", "").isdigit()]
        if sm:
            out["sm_mhz"] = float(np.median(sm))
            out["sm_max_mhz"] = float(rows[0][1])
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            out["reasons"] = [n for j, n in enumerate(names) if any("Active" == r[2 + j].strip() for r in rows)]
        return out


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU path (oracle port) on the host cores, bounded sample."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    per_step = max(4 * cores, 8)          # ~15-20 s of CPU work per step with one thread per core
    for _ in range(min(args.warmup, 1)):      # the CPU arm has no clocks or caches to warm beyond the first import
        cpu_oracle_rate(min(cores, per_step), cores)
    rates, walls = [], []
    for _ in range(args.steps):
        r, w = cpu_oracle_rate(per_step, cores)
        rates.append(r)
        walls.append(w)
    value = per_step * len(walls) * REC_SECONDS / 3600.0 / sum(walls)
    sample = ("%d one-minute recordings per step (39 segments each), %d worker processes with one BLAS / FFT thread each"
              % (per_step, cores))
    line = {
        "impl": "reference", "metric": "recording-hours/sec (STFT->mel->dB)", "value": value,
        "unit": "recording-hours/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1000.0 * sum(walls) / len(walls), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(), "sample": sample,
        "cpu_baseline": {"value": value, "unit": "recording-hours/s", "cores": cores, "kind": "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": "recording-hours/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config():
    c = {"workload": "configs[1]: %d x 60 s 48 kHz recordings per GPU, load_samples S1 (n_fft 4096, hop 281, "
                     "160 mels 50-11000 Hz break 1750, power 2, power_to_db, normalize), 39 segments each" % N_REC,
         "recordings_per_gpu": N_REC, "segments_per_recording": 39, "unique_recordings_per_gpu": POOL,
         "l2": "inputs (11.8 GB PCM per GPU) are larger than L2; no flush needed"}
    if WAVES > 1:
        c["waves_per_step"] = WAVES
        c["workload"] += "; %d waves per step (configs[4]-style sharded run, per-recording results gathered once per step)" % WAVES
    return c


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="profiling runs only: skip the host-buffer leg")
    ap.add_argument("--no-paths", action="store_true", help="skip the `paths` object (other configs, pipeline)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch
    import torch.distributed as dist
    from audio_analysis_b200 import engine, identify_tracks as T, sharding
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    all_cpus = os.sched_getaffinity(0)
    numa_cpus = engine.bind_to_device_numa(local_rank)    # page-locked buffers below land next to this rank's GPU
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL prints its version banner on stdout when the communicator comes up; stdout is for the one JSON line
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.all_reduce(torch.zeros(1, device=dev))
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)

    # ---- inputs: a pool of unique synthetic recordings, replicated into distinct HBM buffers -----
    pool = make_pool(rank, POOL)                                   # [POOL, 2.88M] float32 host
    n_samp = pool.shape[1]
    segs1, counts = T.segment_descriptors(n_samp, SR, [T.Signal(0, REC_SECONDS, 0, 24000)], SEG_LEN, STRIDE)
    nseg1 = len(segs1)
    assert counts == [39]
    plan = T._plan(SR, 4096, 281, 160, 50, 11000, 1750, int(SR * SEG_LEN), 2, True, True) if local_rank == 0 else \
        engine.FeaturePlan(T.custommel._mel_f_cached(SR, 160, 50, 11000, 4096, 1750), SR, 4096, 281, 160,
                           int(SR * SEG_LEN), power=2, db_scale=True, normalize=True, device=local_rank)
    plan.set_profiling(True)

    def descriptors(n_rec):
        s = np.tile(segs1, n_rec)
        s["pcm_offset"] += np.repeat(np.arange(n_rec, dtype=np.int64) * n_samp, nseg1)
        return s

    pool_pinned = torch.from_numpy(pool).pin_memory()
    pool_dev = pool_pinned.to(dev, non_blocking=True)
    pcm = pool_dev.repeat((N_REC + POOL - 1) // POOL, 1)[:N_REC].contiguous().reshape(-1)   # distinct buffers
    segs_dev = torch.from_numpy(descriptors(N_REC).view(np.uint8).reshape(-1)).to(dev)
    out = torch.empty((N_REC * nseg1, 160, plan.n_frames), dtype=torch.float32, device=dev)
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    wave_results = torch.empty((WAVES, N_REC, nseg1), dtype=torch.float32, device=dev)

    def step():
        """WAVES passes of the hot path over this rank's recordings + (N > 1) the per-recording result gather"""
        for w in range(WAVES):
            plan.run(pcm, segs_dev, out)
            if world > 1 or WAVES > 1:
                wave_results[w].copy_(plan.last_segment_max(N_REC * nseg1).view(N_REC, nseg1))
        if world > 1:
            return sharding.gather_per_recording(wave_results.view(WAVES * N_REC, nseg1), world * WAVES * N_REC)
        return None

    # ---- device-resident timing ---------------------------------------------------------------------
    warmup_done = max(args.warmup, 3)                      # the timing rules ask for at least three
    for _ in range(warmup_done):
        step()
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kernel_ms = []
    ev0.record()
    for _ in range(args.steps):
        gathered = step()
    ev1.record()
    barrier()
    total_ms = ev0.elapsed_time(ev1)
    # dominant-kernel duration (events recorded by the library around the fused kernel), one more pass
    for _ in range(3):
        plan.run(pcm, segs_dev, out)
        kernel_ms.append(plan.last_kernel_ms())
    clocks = sampler.stop() if sampler else None
    checksum = float(out[:nseg1].double().sum().item())
    if gathered is not None:
        assert gathered.shape == (world * WAVES * N_REC, nseg1) and bool((gathered > 0).all())

    # ---- end to end through host buffers (pinned H2D of the inputs + D2H of the result per step) -----
    chunk = min(N_REC, int(os.environ.get("CAC_BENCH_E2E_CHUNK", "256")))     # recordings per API call
    n_chunks = (N_REC + chunk - 1) // chunk
    seg_chunk = descriptors(chunk)
    h2d_bytes = int(n_chunks * (chunk * n_samp * 4 + seg_chunk.nbytes))
    d2h_bytes = int(n_chunks * chunk * nseg1 * 160 * plan.n_frames * 4)
    e2e_steps = args.steps
    e2e_s = e2e16_s = e2e_dev_s = float("nan")
    ceiling = None
    if not args.no_e2e:
        out_host_t = torch.empty((chunk * nseg1, 160, plan.n_frames), dtype=torch.float32).pin_memory()
        out_host = out_host_t.numpy()
        # physically distinct host copies of every recording of a call (the pool repeated), page-locked
        host_in_t = pool_pinned.repeat((chunk + POOL - 1) // POOL, 1)[:chunk].contiguous().pin_memory()
        pool_np = host_in_t.numpy().reshape(-1)
        # the same recordings as 16-bit PCM (what a decoder delivers): half the upload
        host_in16_t = torch.from_numpy(np.round(pool_np * 32767.0).astype(np.int16)).pin_memory()
        pool_np16 = host_in16_t.numpy()

        def e2e_step(pcm16=False):
            for _ in range(n_chunks):
                plan.run_host(pool_np16 if pcm16 else pool_np, seg_chunk, out_host)

        def timed_wall(fn):
            fn()
            barrier()
            t0 = time.perf_counter()
            for _ in range(e2e_steps):
                fn()
            barrier()
            return (time.perf_counter() - t0) / e2e_steps

        e2e_s = timed_wall(e2e_step)
        e2e16_s = timed_wall(lambda: e2e_step(pcm16=True))

        # the deployment north_star describes for a classifier on the GPU: PCM up, segments stay on the device
        # (load_samples_device), only the per-segment maxima come back.  Uploads of call k + 1 overlap the kernels of call k.
        up_stream = torch.cuda.Stream(device=dev)
        dev_in = [torch.empty(chunk * n_samp, dtype=torch.float32, device=dev) for _ in range(2)]
        segc_dev = torch.from_numpy(seg_chunk.view(np.uint8).reshape(-1)).to(dev)
        out_c = out[: chunk * nseg1]
        maxima_host = torch.empty((n_chunks, chunk * nseg1), dtype=torch.float32).pin_memory()
        flat_in = host_in_t.reshape(-1)

        def e2e_device_step():
            cur = torch.cuda.current_stream(dev)
            ups, done = [], []
            for c in range(n_chunks):
                with torch.cuda.stream(up_stream):
                    if c >= 2:
                        up_stream.wait_event(done[c - 2])           # the buffer's previous consumer has finished
                    dev_in[c & 1].copy_(flat_in, non_blocking=True)
                    ev = torch.cuda.Event()
                    ev.record(up_stream)
                    ups.append(ev)
                cur.wait_event(ups[c])
                plan.run(dev_in[c & 1], segc_dev, out_c)
                maxima_host[c].copy_(plan.last_segment_max(chunk * nseg1), non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(cur)
                done.append(ev)
            cur.synchronize()

        e2e_dev_s = timed_wall(e2e_device_step)
        assert bool((maxima_host > 0).all())

        # what this box lets `world` ranks copy at the same time: H2D and D2H of 1 GiB each, concurrently, from
        # page-locked memory allocated after the NUMA binding - the ceiling e2e can be read against
        probe_n = 1 << 28
        ph = torch.empty(probe_n, dtype=torch.float32).pin_memory()
        pd = torch.empty(probe_n, dtype=torch.float32, device=dev)
        qh = torch.empty(probe_n, dtype=torch.float32).pin_memory()
        qd = torch.empty(probe_n, dtype=torch.float32, device=dev)
        s1, s2 = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)

        def both():
            with torch.cuda.stream(s1):
                pd.copy_(ph, non_blocking=True)
            with torch.cuda.stream(s2):
                qh.copy_(qd, non_blocking=True)
            s1.synchronize()
            s2.synchronize()

        def up_only():
            with torch.cuda.stream(s1):
                pd.copy_(ph, non_blocking=True)
            s1.synchronize()

        reps = 4
        both()
        barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            both()
        barrier()
        t_both = (time.perf_counter() - t0) / reps
        up_only()
        barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            up_only()
        barrier()
        t_up = (time.perf_counter() - t0) / reps
        ceiling = [probe_n * 4 / t_both / 1e9, probe_n * 4 / t_up / 1e9]
        ph = pd = qh = qd = dev_in = out_c = None

    # ---- reduce over ranks (max time) -------------------------------------------------------------------
    t = torch.tensor([total_ms, e2e_s, float(np.median(kernel_ms)), e2e16_s, e2e_dev_s] + [-c for c in (ceiling or [0.0, 0.0])],
                     dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, e2e_s, k_ms, e2e16_s, e2e_dev_s, neg_both, neg_up = (float(v) for v in t.tolist())
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    try:
        fp32_peak, fp32_src = engine.measure_fp32_tflops(local_rank), "measured in this run (cac_measure_fp32_tflops, FFMA)"
    except Exception as exc:          # keep the line; say what happened
        fp32_peak, fp32_src = FP32_PEAK_FALLBACK_TFLOPS, "fallback constant (%s)" % exc
    rec_hours = world * WAVES * N_REC * REC_SECONDS / 3600.0
    job_hours = world * N_REC * REC_SECONDS / 3600.0
    ms_per_step = total_ms / args.steps
    achieved = ALGO_BYTES_PER_REC * N_REC / (k_ms * 1e-3) / 1e9
    traffic, traffic_src = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        traffic = tj.get("stft_mel_bytes_per_launch")
        traffic_src = "NOT measured in this run: dram__bytes_read.sum + dram__bytes_write.sum from %s (profiles/traffic.json)" % tj.get("source", "?")
    except Exception:
        pass
    line = {
        "metric": "recording-hours/sec (STFT->mel->dB)", "value": rec_hours / (ms_per_step * 1e-3),
        "unit": "recording-hours/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "warmup_done": warmup_done, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(),
        "host_affinity": ("%d CPUs next to the GPU (NVML)" % len(numa_cpus)) if numa_cpus else "unchanged",
        "e2e": {"value": job_hours / e2e_s, "unit": "recording-hours/s",
                "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes, "ms_per_step": e2e_s * 1e3, "steps": e2e_steps,
                "gbs_per_rank": (h2d_bytes + d2h_bytes) / e2e_s / 1e9 if e2e_s == e2e_s else None,
                "copy_ceiling_gbs": None if ceiling is None else {
                    "h2d_while_d2h_per_rank": -neg_both, "h2d_alone_per_rank": -neg_up,
                    "note": "slowest rank; %d ranks copying 1 GiB each way at the same time from page-locked memory" % world},
                "api": "FeaturePlan.run_host -> cac_features_host (3-stream upload/compute/download pipeline, pinned host buffers), %d recordings per call" % chunk},
        "e2e_pcm16": {"value": job_hours / e2e16_s, "unit": "recording-hours/s",
                      "h2d_bytes_per_step": int(n_chunks * (chunk * n_samp * 2 + seg_chunk.nbytes)),
                      "d2h_bytes_per_step": d2h_bytes, "ms_per_step": e2e16_s * 1e3,
                      "note": "same call with the recordings as 16-bit PCM (cac_features_host_pcm16): not the headline, the reference hands over float32"},
        "e2e_device_consumer": {"value": job_hours / e2e_dev_s, "unit": "recording-hours/s", "h2d_bytes_per_step": h2d_bytes,
                                "d2h_bytes_per_step": int(n_chunks * chunk * nseg1 * 4), "ms_per_step": e2e_dev_s * 1e3,
                                "note": "NOT the headline: PCM uploaded from pinned memory every step, the segments stay on the GPU "
                                        "for a classifier there (load_samples_device), only the per-segment maxima come back"},
        "gpu_launches": 3 * WAVES * args.steps,
        "kernels_per_step": ["seg_minmax_kernel", "stft_mel_kernel<2>", "power_to_db_kernel"],
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                     "frac": achieved / hbm_peak, "traffic": traffic, "traffic_source": traffic_src,
                     "kernel": "stft_mel_kernel<2>",
                     "kernel_ms": k_ms, "peak_source": "MEASURED_PEAKS.json (of measured)" if peaks else "fallback",
                     "fp32_tflops": FLOP_PER_REC * N_REC / (k_ms * 1e-3) / 1e12,
                     "fp32_peak_tflops": fp32_peak, "fp32_peak_source": fp32_src,
                     "fp32_frac_of_measured_peak": FLOP_PER_REC * N_REC / (k_ms * 1e-3) / 1e12 / fp32_peak,
                     "note": "FP32-pipe bound (75-110 flop/B, SURVEY.md 8d); HBM fraction reported as the metric asks"},
        "clocks": clocks, "checksum_first_recording": checksum,
    }
    if world == 1 and not args.no_cpu_baseline:
        os.sched_setaffinity(0, all_cpus)                  # the CPU baseline may use every host core again
        cores = os.cpu_count() or 1
        n = max(4 * cores, 8)
        v, wall = cpu_oracle_rate(n, cores)
        v1, wall1 = cpu_oracle_rate(2, 1)
        line["cpu_baseline"] = {"value": v, "unit": "recording-hours/s", "cores": cores, "kind": "port",
                                "sample": "%d one-minute recordings, %d processes x 1 thread, %.1f s wall; single process with default threading: %.5f rec-h/s"
                                          % (n, cores, wall, v1)}
    if world == 1 and not args.no_paths:
        # the other BASELINE.json configs and the per-recording pipeline, each a few seconds (tools/bench_paths.py)
        out = pcm = segs_dev = wave_results = pool_dev = None      # (cells shared with the step closures)
        if not args.no_e2e:
            out_host_t = out_host = host_in_t = pool_np = host_in16_t = pool_np16 = maxima_host = flat_in = None
        torch.cuda.empty_cache()
        if numa_cpus:
            os.sched_setaffinity(0, numa_cpus)             # back next to the GPU: the paths have host halves (track merging)
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import bench_paths
        line["paths"] = {}
        for name in ("single", "pipeline", "morepork", "ci"):
            try:
                line["paths"][name] = bench_paths.PATHS[name](cpu=not args.no_cpu_baseline)
            except Exception as exc:     # a failing side path must not take the headline line with it
                line["paths"][name] = {"error": "%s: %s" % (type(exc).__name__, exc)}
            torch.cuda.empty_cache()
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

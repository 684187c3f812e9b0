"""Generate tests/golden/*.npz by executing the REFERENCE's own files (unchanged, from
/root/reference/src) under the stand-in modules in oracle/stubs.  Run in the build
container only:  python tests/golden/make_golden.py
The reference has no tests or golden vectors of its own; these
fixtures are what pins oracle/oracle_np.py (tests/test_oracle.py) and, through it, the
CUDA path.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import reference_loader as rl  # noqa: E402
from audio_analysis_b200 import synth  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
S1 = dict(hop_length=281, mean_sub=False, use_mfcc=False, mel_break=1750, htk=True, n_mels=160,
          fmin=50, fmax=11000, channels=1, power=2, db_scale=True, normalize=True, n_fft=4096)


def main():
    it = rl.load("identify_tracks")
    cm = rl.load("custommel")
    ci = rl.load("cacophony_index")
    mp = rl.load("identify_morepork")
    common = sys.modules.get("common") or rl.load("common")
    mt1 = rl.load("magtransform")
    mt2 = rl.load("magtransformv2")

    # ---- filterbanks (custommel.mel_f), stored sparse ---------------------------------
    fb = {}
    for name, args in {"s1": (48000, 160, 50, 11000, 4096, 1750),
                       "get_end": (48000, 120, 50, 11000, 4800, 1750),
                       "break1000": (48000, 160, 50, 11000, 4096, 1000),
                       "sr16k": (16000, 64, 50, 7000, 1024, 1750)}.items():
        w = cm.mel_f(*args)
        r, c = np.nonzero(w)
        fb[name + "_args"] = np.array(args, dtype=np.float64)
        fb[name + "_rows"] = r.astype(np.int32)
        fb[name + "_cols"] = c.astype(np.int32)
        fb[name + "_vals"] = w[r, c]
        fb[name + "_shape"] = np.array(w.shape)
    np.savez_compressed(os.path.join(OUT, "filterbanks.npz"), **fb)

    # ---- get_spect on single 3 s segments ----------------------------------------------
    rec = synth.recording(1234, seconds=12.0)
    seg = rec[48000:48000 + 144000]
    gs = {"seed": 1234, "seg_start": 48000}
    x = it.normalize_data(seg)
    gs["normalized_head"] = x[:4096]
    gs["s1_db"] = np.array(it.get_spect(x, 48000, 281, False, False, 1750, True, 160, 50, 11000,
                                        4096, 2, True))[:, :, 0]
    gs["hop640_power"] = np.array(it.get_spect(x, 48000, 640, False, False, 1750, True, 160, 50,
                                               11000, 4096, 2, False))[:, :, 0]
    gs["power1_meansub_ch3"] = np.array(it.get_spect(x, 48000, 281, True, False, 1750, True, 160,
                                                     50, 11000, 4096, 1, False, 3))[::4, ::4, :]
    np.savez_compressed(os.path.join(OUT, "get_spect.npz"), **gs)

    # ---- slicing (load_samples) on a ramp so every index is recoverable ----------------
    ramp = np.arange(1, 12 * 48000 + 1, dtype=np.float32)
    tracks = [it.Signal(0.0, 12.0, 0, 24000),        # whole file
              it.Signal(1.3, 2.1, 800, 3000),        # short, interior -> random widen
              it.Signal(0.05, 0.9, 500, 900),        # short, clamps at file start
              it.Signal(11.2, 11.95, 2000, 9000),    # short, clamps at file end
              it.Signal(3.0, 9.7, 100, 12000),       # ragged tail -> zero pad with offset
              it.Signal(2.0, 5.0, 12000, 15000),     # out of band -> empty placeholder
              it.Signal(4.25, 8.0, 20, 40),          # out of band (below fmin)
              it.Signal(6.1, 9.1000001, 50, 11000)]
    captured = []
    real_get_spect = it.get_spect

    def fake_get_spect(data, *a, **k):
        captured.append(np.array(data))
        return np.zeros((1, 1, 1), np.float32)

    it.get_spect = fake_get_spect
    sl = {}
    for case, kw in {"default": {}, "pad_short": {"pad_short_tracks": True},
                     "seg2_stride1": {"segment_length": 2.0, "stride": 1.0}}.items():
        captured.clear()
        np.random.seed(99)
        seg_len = kw.pop("segment_length", 3.0)
        stride = kw.pop("stride", 1.5)
        res = it.load_samples(ramp, 48000, tracks, seg_len, stride, normalize=False, **kw)
        rows = []
        k = 0
        for ti, tr in enumerate(res):
            for _ in tr:
                d = captured[k]
                k += 1
                nz = np.nonzero(d)[0]
                if len(nz):
                    rows.append((ti, int(d[nz[0]]) - 1, len(nz), int(nz[0]), len(d)))
                else:
                    rows.append((ti, 0, 0, 0, len(d)))
        sl[case] = np.array(rows, dtype=np.int64)
        sl[case + "_counts"] = np.array([len(tr) for tr in res])
    it.get_spect = real_get_spect
    sl["tracks"] = np.array([[t.start, t.end, t.freq_start, t.freq_end] for t in tracks])
    np.savez_compressed(os.path.join(OUT, "slicing.npz"), **sl)

    # ---- load_samples end to end on audio (first/last segments kept) -------------------
    np.random.seed(7)
    tr2 = [it.Signal(0.0, 12.0, 0, 24000), it.Signal(5.2, 6.0, 1000, 4000)]
    res = it.load_samples(rec, 48000, tr2, 3.0, 1.5, **S1)
    np.savez_compressed(os.path.join(OUT, "load_samples.npz"),
                        counts=np.array([len(t) for t in res]),
                        t0_first=np.array(res[0][0])[:, :, 0], t0_last=np.array(res[0][-1])[:, :, 0],
                        t1_only=np.array(res[1][0])[:, :, 0], seed=7, rec_seed=1234)

    # ---- get_end / signal_noise STFT ---------------------------------------------------
    rec_tail = synth.recording(77, seconds=6.0, zero_tail_s=3.0)
    lib = it.librosa  # the stand-in the reference module itself imported
    sn = np.abs(lib.stft(rec[:96000], n_fft=4096, hop_length=281))
    sn = sn / np.amax(sn)
    np.savez_compressed(os.path.join(OUT, "get_end.npz"),
                        end_tail=np.float64(it.get_end(rec_tail, 48000)),
                        end_full=np.float64(it.get_end(rec, 48000)),
                        signal_noise_sub=sn[::8, ::8].astype(np.float32),
                        signal_noise_shape=np.array(sn.shape))

    # ---- morepork ----------------------------------------------------------------------
    def mp_run(frames, sr):
        nfft = int(sr / 10)
        npspec = np.abs(lib.stft(frames, n_fft=nfft, hop_length=int(nfft / 2)))[60:120]
        samples, starts = [], []
        # the loop body of identify_morepork (src/identify_morepork.py:144-166) is inline in a
        # function that also loads TF models, so it is re-driven here line for line
        for base in range(0, npspec.shape[1], mp.sample_slide_slices):
            limit = base + mp.slices_per_sample
            if limit > npspec.shape[1]:
                limit = npspec.shape[1]
            start = limit - mp.slices_per_sample
            starts.append(start / mp.slices_per_second)
            sample = npspec[:, start:limit]
            sample = lib.amplitude_to_db(sample, ref=np.max)
            if sample.min() != 0:
                sample = sample / abs(sample.min()) + 1.0
            if sample.shape[1] < 60:
                sample = np.pad(sample, ((0, 0), (0, 60 - sample.shape[1])))
            samples.append(sample.reshape(sample.shape + (1,)))
        return npspec, np.array(samples), np.array(starts)

    s48, w48, t48 = mp_run(rec, 48000)
    rec16 = synth.recording(55, seconds=8.0, sr=16000)
    s16, w16, t16 = mp_run(rec16, 16000)
    np.savez_compressed(os.path.join(OUT, "morepork.npz"), spec48=s48, win48=w48[:, :, :, 0],
                        starts48=t48, spec16=s16, win16=w16[:, :, :, 0], starts16=t16)

    # ---- cacophony index ---------------------------------------------------------------
    ci16 = synth.recording(4321, seconds=60.0, sr=16000)
    common.load_audio_file_as_numpy_array = lambda f, sr: ci16
    ci.common.load_audio_file_as_numpy_array = lambda f, sr: ci16
    table = ci.calculate("synthetic")
    half = 1024
    prev, pts, bins_first = None, [], []
    for off in range(half, ci16.shape[0] - 3 * half, half):
        b = ci.get_ci_bins(ci16[off:off + 2048], 16000)
        if len(bins_first) < 8:
            bins_first.append(b)
        if prev is not None:
            pts.append(int(sum(b * 2 < prev) + sum(b > prev * 2)))
        prev = b
    short = synth.recording(5, seconds=10.0, sr=16000)
    ci.common.load_audio_file_as_numpy_array = lambda f, sr: short
    table_short = ci.calculate("short")
    np.savez_compressed(
        os.path.join(OUT, "cacophony_index.npz"), seed=4321, bins_first=np.array(bins_first),
        points=np.array(pts, dtype=np.int32),
        table=np.array([[e["begin_s"], e["end_s"], e["index_percent"]]
                        for e in table["cacophony_index_old"]], dtype=np.float64),
        version=table["cacophony_index_old_version"],
        short_warning=table_short.get("ci_warning", ""),
        scores=np.array([ci.score_from_points(pts[:312]),
                         ci.apply_correction_curve_202001C(ci.score_from_points(pts[:312]))]))

    # ---- magtransform ------------------------------------------------------------------
    xin = np.abs(np.random.default_rng(3).standard_normal((4, 33)).astype(np.float32))
    np.savez_compressed(os.path.join(OUT, "magtransform.npz"), x=xin,
                        v1=np.asarray(mt1.MagTransform()(xin)), v2=np.asarray(mt2.MagTransform()(xin)))
    for f in sorted(os.listdir(OUT)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Band-DFT GEMM path (csrc/band_gemm.cu) against the mixed-radix kernel and a float64 band DFT.
usage: python tools/dbg_band_gemm.py            # runs itself twice (CAC_BAND_GEMM=1 / 0) and compares
       python tools/dbg_band_gemm.py run <tag>  # one mode, writes gpurun_out/bg_<tag>.npz"""
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "gpurun_out")
SR = 48000
LENS = [int(SR * 3.3) + 7, SR * 60, 1200, 2400 * 5, 4801, SR * 11 + 1234]   # ragged, with and without a partial chunk
if os.environ.get("CAC_DBG_UNIFORM"):                        # e.g. 600: a uniform batch of that many 40-second recordings (TMA form)
    LENS = [2400 * 800] * int(os.environ["CAC_DBG_UNIFORM"])


def recordings():
    from audio_analysis_b200 import synth
    pool = {}
    recs = []
    for i, n in enumerate(LENS):
        if (i % 8, n) not in pool:
            pool[(i % 8, n)] = synth.recording(300 + i % 8, seconds=n / SR + 0.01)[:n]
        recs.append(pool[(i % 8, n)])
    offs, pos = [], 0
    for r in recs:
        offs.append(pos)
        pos += (len(r) + 3) // 4 * 4          # every recording on a 16-byte boundary
    flat = np.zeros(pos, np.float32)
    for o, r in zip(offs, recs):
        flat[o:o + len(r)] = r
    return recs, flat, np.array(offs, np.int64), np.array([len(r) for r in recs], np.int64)


def run(tag):
    import torch
    from audio_analysis_b200 import identify_morepork as M
    recs, flat, offs, lens = recordings()
    plan = M._plan(SR)
    dev = torch.from_numpy(flat).cuda()
    spec, base, mx = plan.run(dev, offs, lens, want_max=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        plan.run(dev, offs, lens)
    torch.cuda.synchronize()
    print(tag, "5 runs %.3f ms each" % ((time.perf_counter() - t0) * 200), "max", float(mx.item()), flush=True)
    np.savez(os.path.join(OUT, "bg_%s.npz" % tag), spec=spec.cpu().numpy(), base=base, mx=mx.cpu().numpy())


def reference(recs):
    """float64 band DFT of librosa.stft(n_fft=4800, hop=2400, centre=True, periodic Hann), bins 60..119, magnitude"""
    n_fft, hop = 4800, 2400
    w = 0.5 - 0.5 * np.cos(2 * np.pi * np.arange(n_fft) / n_fft)
    k = np.arange(60, 120)
    E = np.exp(-2j * np.pi * np.outer(np.arange(n_fft), k) / n_fft)
    out = []
    for r in recs:
        x = np.concatenate([np.zeros(n_fft // 2), r.astype(np.float64), np.zeros(n_fft // 2)])
        nf = 1 + len(r) // hop
        fr = np.stack([x[f * hop:f * hop + n_fft] for f in range(nf)]) * w
        out.append(np.abs(fr @ E))
    return np.concatenate(out).reshape(-1)


def trace():
    """needs a library built with -DCAC_BG_TRACE=1: per-k-block timestamps of CTA 0's first tile"""
    import ctypes
    import torch
    from audio_analysis_b200 import engine, identify_morepork as M
    recs, flat, offs, lens = recordings()
    plan = M._plan(SR)
    dev = torch.from_numpy(flat).cuda()
    for _ in range(3):
        plan.run(dev, offs, lens)
    torch.cuda.synchronize()
    buf = np.zeros((8, 256), np.int64)
    rc = engine.lib.cac_debug_band_gemm_trace(buf.ctypes.data_as(ctypes.c_void_p))
    assert rc == 0, rc
    t0 = buf[0, 0]
    names = ["landed seen", "conv arrived", "(cp) issued +3", "M full seen", "M committed", "L empty seen"]
    print("kb   " + "  ".join("%-14s" % n for n in names))
    for kb in list(range(0, 24)) + list(range(100, 108)):
        print("%3d  " % kb + "  ".join("%-14d" % (buf[r, kb] - t0) for r in range(6)))
    d = np.diff(buf[4, 20:140])
    print("M commit period over kb 20..140: mean %.0f min %d max %d clk" % (d.mean(), d.min(), d.max()))
    for r, n in enumerate(names):
        print("%-16s mean period %.0f" % (n, np.diff(buf[r, 20:140]).mean()))
    sl = slice(20, 140)
    print("means: landed - empty seen %.0f | arrived - landed %.0f | M full seen - arrived %.0f | commit - full seen %.0f | "
          "next full seen - commit %.0f" %
          ((buf[0, sl] - buf[5, sl]).mean(), (buf[1, sl] - buf[0, sl]).mean(), (buf[3, sl] - buf[1, sl]).mean(),
           (buf[4, sl] - buf[3, sl]).mean(), (buf[3, 21:141] - buf[4, sl]).mean()))


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    if len(sys.argv) > 2 and sys.argv[1] == "run":
        run(sys.argv[2])
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "trace":
        trace()
        sys.exit(0)
    for tag, flag in (("gemm", "1"), ("fft", "0")):
        env = dict(os.environ, CAC_BAND_GEMM=flag)
        rc = subprocess.run(["timeout", "40", sys.executable, __file__, "run", tag], env=env).returncode
        print(tag, "exit", rc, flush=True)
    recs = recordings()[0]
    ref = reference(recs)
    for tag in ("gemm", "fft"):
        p = os.path.join(OUT, "bg_%s.npz" % tag)
        if not os.path.exists(p):
            continue
        got = np.load(p)["spec"]
        d = np.abs(got - ref)
        print("%-4s vs float64: max abs %.3e (scale %.3e), rel<=1e-4 on %.4f of cells, nan %d, max %.6e" %
              (tag, d.max(), ref.max(), float(np.mean(d <= 1e-4 * np.abs(ref))), int(np.isnan(got).sum()),
               float(np.load(p)["mx"][0])))
        worst = int(np.argmax(d))
        print("     worst cell: frame %d bin %d got %.6e ref %.6e" % (worst // 60, worst % 60, got[worst], ref[worst]))



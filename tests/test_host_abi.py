"""CPU-side checks: the C-ABI library builds, loads and exports every symbol the header
declares; host slicing logic is bit-exact vs the reference-generated fixtures; the
filterbank is bit-exact; no compute call is made (no GPU here)."""
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def eng():
    subprocess.check_call(["bash", os.path.join(ROOT, "audio-analysis_b200", "csrc", "build.sh")])
    from audio_analysis_b200 import engine
    return engine


def test_header_symbols_exported(eng):
    hdr = open(os.path.join(ROOT, "include", "cacophony_b200.h")).read()
    declared = set(re.findall(r"\b(cac_[a-z_0-9]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    for name in sorted(declared):
        assert hasattr(eng.lib, name), "missing export " + name
    assert declared == set(eng.EXPORTS)
    assert eng.version().startswith("cacophony_b200")


def test_no_cpu_fallback_without_device(eng):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from audio_analysis_b200 import custommel
    w = custommel.mel_f(48000, 160, 50, 11000, 4096, 1750)
    with pytest.raises(eng.CacError):
        eng.FeaturePlan(w, 48000, 4096, 281, 160, 144000)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "audio-analysis_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f


def test_mel_f_bit_exact_vs_reference_fixture(golden):
    from audio_analysis_b200 import custommel
    g = golden("filterbanks")
    for name in ("s1", "get_end", "break1000", "sr16k"):
        a = g[name + "_args"]
        w = custommel.mel_f(int(a[0]), int(a[1]), a[2], a[3], int(a[4]), a[5])
        r, c = np.nonzero(w)
        assert np.array_equal(r, g[name + "_rows"]) and np.array_equal(c, g[name + "_cols"])
        assert np.array_equal(w[r, c], g[name + "_vals"])


@pytest.mark.parametrize("case,kw", [("default", {}), ("pad_short", {"pad_short_tracks": True}),
                                     ("seg2_stride1", {"segment_length": 2.0, "stride": 1.0})])
def test_segment_descriptors_bit_exact_vs_reference_fixture(eng, golden, case, kw):
    from audio_analysis_b200 import identify_tracks as T
    g = golden("slicing")
    kw = dict(kw)
    seg_len, stride = kw.pop("segment_length", 3.0), kw.pop("stride", 1.5)
    np.random.seed(99)
    segs, counts = T.segment_descriptors(12 * 48000, 48000, [T.Signal(*r) for r in g["tracks"]], seg_len, stride, **kw)
    assert counts == list(g[case + "_counts"])
    ref = g[case]
    assert len(segs) == len(ref)
    for s, r in zip(segs, ref):
        assert s["n_valid"] == r[2]
        if r[2]:
            assert s["pcm_offset"] == r[1] and s["pad_left"] == r[3]


def test_whole_minute_is_39_segments(eng):
    from audio_analysis_b200 import identify_tracks as T
    segs, counts = T.segment_descriptors(60 * 48000, 48000, [T.Signal(0, 60, 0, 24000)], 3, 1.5)
    assert counts == [39] and (segs["n_valid"] == 144000).all() and (segs["pad_left"] == 0).all()
    assert list(segs["pcm_offset"][:3]) == [0, 72000, 144000]


def test_host_fft_emulation():
    """The register-FFT templates (fft_regs.cuh) compiled for the host: warp algorithm emulated."""
    exe = "/tmp/cac_fft_emul"
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-o", exe, os.path.join(ROOT, "tests", "host", "fft_emul.cpp")])
    out = subprocess.check_output([exe]).decode()
    vals = dict(line.split()[:2] for line in out.strip().splitlines() if line.startswith("small"))
    assert all(float(v) < 2e-6 for v in vals.values()), out
    rel = float(re.search(r"rel ([0-9.e+-]+)", out).group(1))
    assert rel < 1e-6, out

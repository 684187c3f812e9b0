"""BASELINE.json's FULL sizes on the GPU, checked through size-independent properties (the oracle cannot
run these sizes in minutes): replicas of one recording must come out bitwise equal wherever the batch
puts them, a big batch must equal the small batch of its unique members, dB blocks must peak at exactly
0 dB and stay above -80 dB, periodic input must give periodic Cacophony Index bins, a chunked Stream must
equal the one-shot run, and the structural known answers of SURVEY.md 8(c) must hold
(39 segments x 513 frames per minute, 61 morepork windows, 449,996 windows / 1,442 entries per 8 hours)."""
import numpy as np
import pytest

from audio_analysis_b200 import synth
from oracle import oracle_np as O

pytestmark = pytest.mark.gpu

S1 = dict(hop_length=281, mel_break=1750, n_mels=160, fmin=50, fmax=11000, power=2, n_fft=4096)


def test_configs1_1024_recordings_replica_invariance_and_db_range(engine):
    """configs[1]: 1,024 one-minute 48 kHz recordings, load_samples S1 -> [39936, 160, 513]."""
    import torch
    from audio_analysis_b200 import identify_tracks as T
    n_unique, n_rec, sr = 4, 1024, 48000
    pool = synth.batch(7000, n_unique, seconds=60.0)                       # [4, 2,880,000]
    n_samples = pool.shape[1]
    segs, counts = T.segment_descriptors(n_samples, sr, [T.Signal(0, 60.0, 0, 24000)], 3.0, 1.5)
    assert len(segs) == 39                                                 # 39 segments per whole-minute track
    plan = T._plan(sr, 4096, 281, 160, 50, 11000, 1750, 144000, 2, True, True)
    assert plan.n_frames == 513

    def descriptors(n):
        d = np.concatenate([segs.copy() for _ in range(n)])
        d["pcm_offset"] += np.repeat(np.arange(n, dtype=np.int64) * n_samples, len(segs))
        return d

    pool_dev = torch.from_numpy(pool).cuda()
    small = plan.run(pool_dev.reshape(-1), descriptors(n_unique)).clone()  # [156, 160, 513]
    pcm = pool_dev.repeat(n_rec // n_unique, 1).reshape(-1)               # physically distinct copies: 11.8 GB
    assert pcm.numel() == n_rec * n_samples
    out = plan.run(pcm, descriptors(n_rec))
    assert tuple(out.shape) == (n_rec * 39, 160, 513)
    rep = out.view(n_rec // n_unique, n_unique * 39, 160, 513)
    # every replica equals the small batch bitwise, whichever CTA / stage run / launch size produced it
    for r in range(rep.shape[0]):
        assert torch.equal(rep[r], small), "replica %d differs from the 4-recording batch" % r
    # power_to_db(ref=max, top_db=80): every segment peaks at exactly 0 dB and never goes below -80
    flat = out.view(n_rec * 39, -1)
    assert torch.equal(flat.max(dim=1).values, torch.zeros(n_rec * 39, device=out.device))
    assert float(flat.min()) >= -80.0
    # one WHOLE recording from the middle of the big batch (all 39 segments: every stage / CTA boundary inside it)
    # and the last recording's edge segments against the oracle
    mid = rep[rep.shape[0] // 2, 2 * 39:3 * 39].cpu().numpy()
    ref = O.load_samples(pool[2], sr, [O.Signal(0, 60.0, 0, 24000)], 3.0, 1.5, db_scale=True, **S1)[0]
    assert len(ref) == 39
    for k in range(39):
        assert np.abs(mid[k] - ref[k][:, :, 0]).max() <= 1e-2             # BASELINE.json: 1e-2 dB max-abs
    last = rep[rep.shape[0] - 1, 3 * 39:4 * 39].cpu().numpy()
    ref3 = O.load_samples(pool[3], sr, [O.Signal(0, 60.0, 0, 24000)], 3.0, 1.5, db_scale=True, **S1)[0]
    for k in (0, 38):
        assert np.abs(last[k] - ref3[k][:, :, 0]).max() <= 1e-2
    # run-to-run determinism at full size: a checksum of per-segment checksums
    chk1 = flat.sum(dim=1, dtype=torch.float64).cpu().numpy().copy()
    out.zero_()
    plan.run(pcm, descriptors(n_rec), out=out)
    chk2 = flat.sum(dim=1, dtype=torch.float64).cpu().numpy()
    assert np.array_equal(chk1, chk2)
    del out, pcm
    torch.cuda.empty_cache()


def test_configs2_4096_morepork_recordings(engine):
    """configs[2]: 4,096 one-minute 48 kHz recordings through the morepork spectrogram + 60 x 60 windows."""
    import torch
    from audio_analysis_b200 import identify_morepork as M
    n_unique, n_rec = 8, 4096
    pool = synth.batch(7100, n_unique, seconds=60.0)
    plan = M.BatchPlan(48000, pool.shape[1])
    assert plan.n_frames == 1201 and plan.n_windows == 61                  # SURVEY.md 8(c) known answers
    pool_dev = torch.from_numpy(pool).cuda()
    small, start_times = plan.run(pool_dev)
    small = small.clone()
    assert start_times[0] == 0.0 and len(start_times) == 61
    pcm = pool_dev.repeat(n_rec // n_unique, 1)                            # 47 GB of physically distinct copies
    out, _ = plan.run(pcm)
    assert tuple(out.shape) == (n_rec, 61, 60, 60)
    rep = out.view(n_rec // n_unique, n_unique, 61, 60, 60)
    assert bool((rep == small.unsqueeze(0)).all())
    # amplitude_to_db(ref=max) / abs(min) + 1 (src/identify_morepork.py:152-154): windows live in [0, 1], peak exactly 1
    assert float(out.min()) >= 0.0 and float(out.max()) == 1.0
    # one recording of the big batch against the oracle
    ref, ref_times = O.morepork_windows(O.morepork_spectrogram(pool[5], 48000))
    got = rep[-1, 5].cpu().numpy()
    assert list(ref_times) == list(start_times)
    assert np.abs(got - np.asarray(ref).reshape(got.shape)).max() <= 1.25e-4
    del out, pcm
    torch.cuda.empty_cache()


def test_configs3_eight_hour_cacophony_index(engine):
    """configs[3]: one 8-hour 16 kHz recording; periodic input (period = 960 hops) so that every property
    can be checked without an 8-hour oracle run."""
    import torch
    from audio_analysis_b200 import cacophony_index as CI
    period = 960 * 1024                                                    # 61.44 s at 16 kHz
    base = synth.recording(7200, seconds=period / 16000.0, sr=16000)[:period]
    n_total = 8 * 3600 * 16000                                             # 460,800,000 samples = 1.84 GB
    pcm_host = np.tile(base, n_total // period + 1)[:n_total]
    plan = CI._plan(CI.WINDOW_SIZE, CI.SAMPLE_RATE)
    pcm = torch.from_numpy(pcm_host).cuda()
    bins, points, wbase = plan.run(pcm, [0], [n_total])
    n_win = int(wbase[-1])
    assert n_win == 449996                                                 # 449,996 windows, 449,995 points
    bins = bins[:n_win]
    points = points[:n_win]
    # window i and window i + 960 see the same samples: bins bitwise equal, points equal (from the 2nd period on)
    k = (n_win // 960 - 1) * 960
    assert torch.equal(bins[:k], bins[960:960 + k])
    assert torch.equal(points[1:k], points[961:960 + k])
    # the first period against the float64 oracle, window by window (threshold cells within 1e-6 excepted)
    ref_bins = np.stack([O.get_ci_bins(base[1024 * (i + 1):1024 * (i + 1) + 2048].astype(np.float32), 16000) for i in range(64)])
    got_bins = bins[:64].cpu().numpy()
    assert np.allclose(got_bins, ref_bins, rtol=1e-9, atol=0)
    # whole result: one-shot == chunked Stream (37 uneven chunks), 1,442 entries
    one = CI.calculate(pcm_host)
    assert len(one["cacophony_index_old"]) == 1442 and "ci_warning" not in one
    st = CI.Stream()
    edges = np.linspace(0, n_total, 38).astype(np.int64)
    edges[1:-1] += np.arange(1, 37) * 37                                   # not multiples of the hop
    for a, b in zip(edges[:-1], edges[1:]):
        st.feed(pcm_host[a:b])
    assert st.result() == one
    # periodic input -> the entries repeat too once their 312-point spans fall on the same phase; at least all
    # index values must be finite percentages
    vals = [e["index_percent"] for e in one["cacophony_index_old"]]
    assert all(0.0 <= v <= 100.0 for v in vals)

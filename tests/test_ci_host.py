"""CPU checks of the Cacophony Index host logic (entry layout, result assembly, chunked
streaming bookkeeping) with the CUDA plan replaced by an oracle-backed fake: no compute call
reaches the library here."""
import numpy as np
import pytest

from audio_analysis_b200 import synth
from oracle import oracle_np as O


class _FakePlan:
    device = 0

    @staticmethod
    def window_count(n):
        return len(range(1024, n - 3072, 1024))

    def run(self, pcm, off, ns, want_points=True):
        import torch
        assert list(off) == [0]
        x = pcm.numpy()[:ns[0]]
        pts = O.ci_points_fast(x)
        n_win = self.window_count(ns[0])
        assert len(pts) == max(n_win - 1, 0)
        return None, torch.from_numpy(np.array([255] + list(pts), dtype=np.uint8)), np.array([0, n_win])

    def entries(self, points, starts, lens):
        p = points.numpy()
        out = []
        for s, n in zip(starts, lens):
            q = sorted(p[s:s + n].tolist())
            k0, k1 = int(n * 0.75), int(n * 0.95)
            out.append((sum(q[k0:k1]), k1 - k0))
        return np.array(out, dtype=np.int32).reshape(-1, 2)


@pytest.fixture()
def fake(monkeypatch):
    import torch
    from audio_analysis_b200 import cacophony_index as C
    plan = _FakePlan()
    monkeypatch.setattr(C, "_plan", lambda *a, **k: plan)
    monkeypatch.setattr(C.Stream, "_upload", lambda self, seg: torch.from_numpy(seg))
    return C


def test_window_count_matches_range():
    import ctypes
    from audio_analysis_b200 import engine
    for n in list(range(0, 9000, 37)) + [4096, 4097, 5120, 5121, 960000, 460800000]:
        assert engine.CiPlan.window_count(n) == len(range(1024, n - 3072, 1024)), n


@pytest.mark.parametrize("n_points", [0, 5, 280, 281, 311, 312, 313, 592, 593, 933, 5000])
def test_entry_layout_and_table_match_oracle(fake, n_points):
    rng = np.random.default_rng(n_points)
    pts = rng.integers(0, 11, size=n_points).tolist()
    starts, lens = fake.entry_layout(n_points)
    import torch
    stats = _FakePlan().entries(torch.from_numpy(np.array(pts, dtype=np.uint8)), starts, lens) if starts else []
    got = fake._result(fake._table_from_stats(starts, stats), 1024 * (n_points + 4))
    ref = O.ci_table(pts, 1024 * (n_points + 4))
    assert got == ref
    for a, b in zip(got["cacophony_index_old"], ref["cacophony_index_old"]):
        assert all(type(a[k]) is type(b[k]) for k in a)      # identical JSON (0 vs 0.0)


def test_stream_bookkeeping_equals_one_shot(fake):
    x = synth.recording(5, seconds=70.0, sr=16000)
    ref = O.calculate_from_array(x, fast=True)
    rng = np.random.default_rng(1)
    for sizes in ([16000 * 20] * 4, list(rng.integers(1, 30000, size=100)), [x.shape[0]], [3000] * 400):
        st, pos = fake.Stream(), 0
        for n in sizes:
            if pos >= x.shape[0]:
                break
            st.feed(x[pos:pos + int(n)])
            pos += int(n)
        if pos < x.shape[0]:
            st.feed(x[pos:])
        assert st.result() == ref
    st = fake.Stream()
    st.feed(x[:5000])
    assert st.result() == O.calculate_from_array(x[:5000])

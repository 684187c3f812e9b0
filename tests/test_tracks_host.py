"""tracks.py (merge heuristics) and analyse.py (result assembly) against fixtures produced by the reference's own
functions (tests/golden/make_golden_host.py executes src/identify_tracks.py:725-842 and src/analyse.py:18-126)."""
import json
import os

import pytest

from audio_analysis_b200 import analyse as A
from audio_analysis_b200 import tracks as TR

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


class Tag:
    def __init__(self, what):
        self.what = what

    def get_meta(self):
        return {"label": self.what}


def _load(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


@pytest.mark.parametrize("case", _load("tracks.json"))
def test_get_tracks_from_signals_equals_reference(case):
    got = TR.get_tracks_from_signals([TR.Signal(*b) for b in case["signals"]], case["length"])
    assert [[t.start, t.end, t.freq_start, t.freq_end] for t in got] == case["tracks"]


def _python_sweeps(signals):
    merged = True
    while merged:
        signals, merged = TR.merge_signals(signals)
    return signals


def _boxes(rng, n, quantised):
    out = []
    for _ in range(n):
        t0 = rng.uniform(0, 30)
        f0 = rng.uniform(0, 9000)
        box = [t0, t0 + rng.exponential(1.5), f0, f0 + rng.uniform(0, 5000)]
        if quantised:                                  # many exact ties in every field; ints where the reference has ints
            box = [round(box[0] * 2) / 2, round(box[1] * 2) / 2 + 0.5, int(box[2] // 500 * 500), int(box[3] // 500 * 500 + 500)]
        out.append(box)
    return out


@pytest.mark.parametrize("seed", range(40))
def test_native_merge_sweeps_equal_the_python_loop(seed):
    """cac_tracks_merge (the path get_tracks_from_signals takes) against merge_signals repeated in Python: the same
    survivors in the same order, every field the same VALUE and the same TYPE, and the absorbed objects left in the
    state they had when they were absorbed."""
    import numpy as np
    rng = np.random.default_rng(seed)
    n = int(rng.integers(0, 60)) if seed else 0
    boxes = _boxes(rng, n, quantised=seed % 2 == 1)
    a, b = [TR.Signal(*x) for x in boxes], [TR.Signal(*x) for x in boxes]
    want, got = _python_sweeps(a), TR.merge_until_stable(b)
    key = lambda s: (s.start, s.end, s.freq_start, s.freq_end, s.mel_freq_start, s.mel_freq_end)
    assert [key(s) for s in got] == [key(s) for s in want]
    assert [[type(v) for v in key(s)] for s in got] == [[type(v) for v in key(s)] for s in want]
    assert [b.index(s) for s in got] == [a.index(s) for s in want]                # the same objects survive
    assert [key(s) for s in b] == [key(s) for s in a]                            # absorbed signals: as they were then


def _tracks(case):
    out = []
    for box, tag in zip(case["tracks"], case["tags"]):
        t = TR.Signal(*box)
        t.master_tag = Tag(tag) if tag is not None else None
        out.append(t)
    return out


@pytest.mark.parametrize("case", _load("analyse.json"))
def test_result_assembly_equals_reference(case):
    tracks = _tracks(case)
    ci, version = A.calc_cacophony_index(A.filter_tracks(tracks), case["length"])
    assert ci == case["cacophony_index"] and version == case["version"]
    signals = [TR.Signal(*b) for b in case["signals"]]
    assert [[float(v) for v in s.to_array()] for s in signals] == case["rec_signals"]
    assert A.get_chirps(tracks, case["bird_labels"], signals) == case["chirps"]
    assert len(signals) == case["signals_left"]                     # counted signals are removed, like the reference
    assert A.get_max_chirps(case["length"]) == case["max_chirps"]


def test_analysis_result_schema_and_batch():
    case = next(c for c in _load("analyse.json") if c["chirps"] > 0)
    tracks = _tracks(case)
    for i, t in enumerate(tracks):
        t.track_id = i
    res = A.analysis_result(tracks, case["length"], [TR.Signal(*b) for b in case["signals"]], case["length"] + 0.25,
                            case["bird_labels"])
    assert set(res) == {"duration", "cacophony_index", "cacophony_index_version", "chirps", "non_bird_tags",
                        "species_identify", "species_identify_version"}
    assert res["cacophony_index"] == case["cacophony_index"] and res["cacophony_index_version"] == "2.0"
    assert res["chirps"] == {"chirps": case["chirps"], "max_chirps": case["max_chirps"],
                             "chirp_index": round(100 * case["chirps"] / case["max_chirps"]),
                             "signals": [list(map(float, s)) for s in case["rec_signals"]]}
    assert res["duration"] == case["length"] + 0.25 and res["non_bird_tags"] == ["human", "noise", "insect"]
    meta = res["species_identify"][0]
    assert set(meta) >= {"begin_s", "end_s", "freq_start", "freq_end", "model_results", "track_id"}
    # analyse_tracks=True: only the track list (src/analyse.py:161-177)
    only = A.analysis_result(_tracks(case), case["length"], [], case["length"], case["bird_labels"], analyse_tracks=True)
    assert set(only) == {"non_bird_tags", "species_identify", "species_identify_version"}
    batch = A.analysis_result_batch([None, (_tracks(case), case["length"], [], case["length"], case["bird_labels"])])
    assert batch[0] == A.empty_result() and batch[1]["cacophony_index"] == case["cacophony_index"]
    text = A.merge_into_metadata({"location": 1}, dict(batch[1]), processing_time_seconds=1.234)
    back = json.loads(text)
    assert back["location"] == 1 and back["analysis_result"]["processing_time_seconds"] == 1.2
    assert text == json.dumps(back, sort_keys=True, indent=4)
    assert A.calc_cacophony_index_batch([A.filter_tracks(_tracks(case))], [case["length"]])[0][0] == case["cacophony_index"]

"""Result assembly: what the reference's CLI builds from classify()'s output before it writes
`analysis_result` into the recording's metadata file - the track-coverage Cacophony Index (src/analyse.py:18-80), the
chirp count (:92-126), get_max_chirps (src/identify_tracks.py:45-46) and the `analysis_result` dictionary itself
(:152-177, 453-468) - for one recording or a batch.  Scalar host code over a handful of tracks per recording; it is
here so that a batch run over many files (GPU feature extraction for all of them, then one assembly pass) produces the
reference's JSON.  Pinned against the reference's own functions (tests/golden/analyse.json, tests/test_analyse_host.py).
"""
import json
import math

from .tracks import segment_overlap

NON_BIRD = ["human", "noise", "insect"]      # src/identify_tracks.py:16
SIGNAL_WIDTH = 0.25                          # src/identify_tracks.py:21
SPECIES_IDENTIFY_VERSION = "2025-12-01"      # src/analyse.py:176


def get_max_chirps(length):
    """src/identify_tracks.py:43-46: at most one chirp per SIGNAL_WIDTH + 10 ms."""
    return int(length / (SIGNAL_WIDTH + 0.01))


def filter_tracks(tracks):
    """Tracks whose master tag is a bird (src/analyse.py:83-89)."""
    return [t for t in tracks if t.master_tag is not None and t.master_tag.what not in NON_BIRD]


def _periods(length, period_length=20):
    """The 20-second bins of calc_cacophony_index (src/analyse.py:22-43): a last bin shorter than 2 s is folded into
    the one before it."""
    bins = math.ceil(length / period_length)
    tail = length - period_length * (bins - 1)
    fold = bins > 1 and tail < 2
    if fold:
        bins -= 1
    out = [{"begin_s": i * period_length, "end_s": min(length, (i + 1) * period_length), "index_percent": 0}
           for i in range(bins)]
    if fold:
        out[-1]["end_s"] = length
    return out


def calc_cacophony_index(tracks, length):
    """src/analyse.py:18-80: percentage of every 20-second period covered by the given (bird) tracks, walked in the
    order given with a running `bird_until` front; returns (periods, "1.0").  The period bookkeeping reproduces the
    reference step for step (including its recomputation of the current period from bird_until / period_length)."""
    percents = _periods(length)
    bins = len(percents)
    covered = 0
    bird_until = -1
    period = 0
    period_length = percents[0]["end_s"] - percents[0]["begin_s"] if percents else 20
    period_end = period_length
    for track in tracks:
        if track.start <= bird_until < track.end:
            lo, hi = bird_until, track.end                      # the bird started inside the span already counted
        elif bird_until < track.end:
            lo, hi = track.start, track.end
        else:
            continue
        while hi > period_end:                                   # the span runs past the current period: close it
            if lo < period_end:
                covered += period_end - lo
                lo = period_end
            percents[period]["index_percent"] = round(100 * covered / period_length, 1)
            covered = 0
            period = min(period + 1, bins - 1)
            period_length = percents[period]["end_s"] - percents[period]["begin_s"]
            period_end += period_length
        covered += hi - lo
        bird_until = hi
        period = min(len(percents) - 1, int(bird_until / period_length), bins - 1)
        period_length = percents[period]["end_s"] - percents[period]["begin_s"]
    if period < len(percents):
        percents[period]["index_percent"] = round(100 * covered / period_length, 1)
    return percents, "1.0"


def get_chirps(tracks, bird_labels, signals):
    """src/analyse.py:92-126: signals that overlap a bird-labelled track in time and (within 200 mel) in frequency;
    every signal counts once and is REMOVED from `signals` when counted, as in the reference."""
    birds = sorted((t for t in tracks if t.master_tag is not None and t.master_tag.what in bird_labels),
                   key=lambda t: t.start)
    chirps = 0
    last_end = 0
    for t in birds:
        start, end = t.start, t.end
        if start < last_end:
            start = last_end
            end = max(start, end)
        i = 0
        while i < len(signals):
            s = signals[i]
            if segment_overlap((start, end), (s.start, s.end)) > 0 and t.mel_freq_overlap(s) > -200:
                chirps += 1
                del signals[i]
            elif s.start > end:
                break
            else:
                i += 1
        last_end = t.end
    return chirps


def analysis_result(tracks, length, signals, raw_length, bird_labels, analyse_tracks=False, region_code=None):
    """The dictionary species_identify builds from classify()'s tuple (src/analyse.py:143-177): tracks must already
    carry their master tags (Signal.master_tag).  `signals` is consumed by get_chirps like in the reference."""
    result = {}
    rec_signals = [s.to_array() for s in signals]
    chirps = get_chirps(tracks, bird_labels, signals)
    cacophony_index, _ = calc_cacophony_index(filter_tracks(tracks), length)
    if not analyse_tracks:
        max_chirps = get_max_chirps(length)
        if region_code is not None:
            result["region_code"] = region_code
        result["duration"] = raw_length
        result["cacophony_index"] = cacophony_index
        result["cacophony_index_version"] = "2.0"
        result["chirps"] = {"chirps": chirps, "max_chirps": max_chirps,
                            "chirp_index": 0 if max_chirps == 0 else round(100 * chirps / max_chirps),
                            "signals": rec_signals}
    result["non_bird_tags"] = NON_BIRD
    result["species_identify"] = [t.get_meta() for t in tracks]
    result["species_identify_version"] = SPECIES_IDENTIFY_VERSION
    return result


def empty_result():
    """species_identify when classify returned nothing (src/analyse.py:174-177)."""
    return {"non_bird_tags": NON_BIRD, "species_identify": [], "species_identify_version": SPECIES_IDENTIFY_VERSION}


def analysis_result_batch(classified, analyse_tracks=False):
    """analysis_result for many recordings: `classified` holds one classify() tuple (tracks, length, signals,
    raw_length, bird_labels) or None per recording."""
    return [empty_result() if c is None else analysis_result(*c, analyse_tracks=analyse_tracks) for c in classified]


def calc_cacophony_index_batch(tracks_per_recording, lengths):
    """calc_cacophony_index for many recordings (bird tracks of each, its length in seconds)."""
    return [calc_cacophony_index(t, n) for t, n in zip(tracks_per_recording, lengths)]


def merge_into_metadata(metadata, summary, processing_time_seconds=None):
    """main()'s last step (src/analyse.py:451-468): the summary goes under "analysis_result" of the recording's
    metadata; returns the JSON text the reference writes (sorted keys, indent 4)."""
    if processing_time_seconds is not None:
        summary["processing_time_seconds"] = round(processing_time_seconds, 1)
    metadata = dict(metadata or {})
    metadata["analysis_result"] = summary
    return json.dumps(metadata, sort_keys=True, indent=4)

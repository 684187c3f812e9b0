"""CPU check of identify_morepork's window table (host logic): column ranges must equal the
reference loop's Python slices, including negative-start wrap-around for short recordings."""
import numpy as np
import pytest


@pytest.mark.parametrize("n", [1, 5, 19, 20, 21, 25, 39, 40, 41, 50, 59, 60, 61, 79, 80, 81, 100, 241, 1201])
def test_window_table_equals_reference_slices(n):
    from audio_analysis_b200 import identify_morepork as M
    spec = np.arange(3 * n, dtype=np.float32).reshape(3, n)
    first, width, starts = M.window_table(n)
    k = 0
    for base in range(0, n, M.sample_slide_slices):        # src/identify_morepork.py:144-149
        limit = base + M.slices_per_sample
        if limit > n:
            limit = n
        start = limit - M.slices_per_sample
        ref = spec[:, start:limit]
        assert starts[k] == start / M.slices_per_second
        assert np.array_equal(spec[:, first[k]:first[k] + width[k]], ref) and ref.shape[1] > 0
        k += 1
    assert k == len(first)


@pytest.mark.parametrize("n_fft,hop", [(4800, 2400), (4410, 2205), (2205, 1102), (800, 400), (4096, 281), (15, 7)])
def test_stft_frame_counts_equal_the_oracle_stft(n_fft, hop):
    """engine.stft_frame_counts (and cac_spec_frame_count) follow librosa.stft(center=True) for even AND odd frame
    lengths: checked against the shape the oracle's stft returns."""
    from audio_analysis_b200 import engine
    from oracle import oracle_np as O
    for n in (1, hop - 1, hop, hop + 1, n_fft - 1, n_fft, 3 * n_fft + 5, 10 * hop):
        ref = O.stft(np.zeros(n, np.float32), n_fft, hop).shape[1]
        assert int(engine.stft_frame_counts(n, n_fft, hop)) == ref, (n, n_fft, hop)
    assert list(engine.stft_frame_counts([0, 1], 4800, 2400)) == [1, 1]
    assert list(engine.stft_frame_counts([0, 1], 2205, 1102)) == [0, 1]

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

"""cac_power_to_db_static: the allocation-free variant of cac_power_to_db_run (scratch in a static device array).

No reference counterpart - added for callers that cannot call cudaMalloc.  Collected LAST on purpose (pytest -x).
The entry point is called through the C-ABI twice in a row with different item counts: the second call sees the
first call's leftovers in the shared scratch and must not care."""
import ctypes

import numpy as np
import pytest

from oracle import oracle_np as O

pytestmark = pytest.mark.gpu


def _both(engine, torch, x):
    """x: CUDA float32 [n_items, item_len] -> (existing entry point's output, static entry point's output)."""
    n_items, item_len = x.shape
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    ref = torch.empty_like(x)
    work = torch.empty(n_items * 4 + 16, dtype=torch.uint8, device=x.device)
    engine.check(engine.lib.cac_power_to_db_run(x.data_ptr(), ref.data_ptr(), n_items, item_len, work.data_ptr(), stream))
    got = torch.full_like(x, float("nan"))
    engine.check(engine.lib.cac_power_to_db_static(x.data_ptr(), got.data_ptr(), n_items, item_len, stream))
    torch.cuda.synchronize()
    return ref.cpu().numpy(), got.cpu().numpy()


def test_static_scratch_matches_work_buffer_entry_point(engine):
    import torch
    rng = np.random.default_rng(20261105)
    # mel-power-like magnitudes over ~9 decades, so that the top_db floor is active; second call: fewer, louder items
    shapes = [(37, 120 * 45), (5, 4099)]
    for call, (n_items, item_len) in enumerate(shapes):
        host = (10.0 ** rng.uniform(-9.0, 0.5 + call, size=(n_items, item_len))).astype(np.float32)
        host[0, :7] = 0.0                                          # below amin
        ref, got = _both(engine, torch, torch.from_numpy(host).cuda())
        assert np.array_equal(got, ref), "call %d differs from cac_power_to_db_run" % call
        want = np.stack([O.power_to_db(h, ref=np.max) for h in host])
        assert np.abs(got - want).max() <= 1e-2                    # BASELINE.json's dB tolerance (test_features_gpu.py)
        assert got.max(axis=1).tolist() == [0.0] * n_items and got.min() >= -80.0


def test_static_scratch_in_place_and_limits(engine):
    import torch
    rng = np.random.default_rng(7)
    host = rng.uniform(1e-6, 3.0, size=(3, 1000)).astype(np.float32)
    x = torch.from_numpy(host).cuda()
    ref, _ = _both(engine, torch, x)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    engine.check(engine.lib.cac_power_to_db_static(x.data_ptr(), x.data_ptr(), 3, 1000, stream))       # in place
    torch.cuda.synchronize()
    assert np.array_equal(x.cpu().numpy(), ref)
    # empty input: nothing to do; too many items for the static scratch: refused before anything is launched
    assert engine.lib.cac_power_to_db_static(x.data_ptr(), x.data_ptr(), 0, 1000, stream) == 0
    rc = engine.lib.cac_power_to_db_static(x.data_ptr(), x.data_ptr(), engine.POWER_TO_DB_STATIC_MAX_ITEMS + 1, 1, stream)
    assert rc != 0 and b"exceeds the static scratch" in engine.lib.cac_last_error()
    torch.cuda.synchronize()
    assert np.array_equal(x.cpu().numpy(), ref)

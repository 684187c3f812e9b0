"""world_size-2 gloo test of the multi-GPU host logic: whole-recording sharding and
the per-recording result gather give the same table as a single rank, for even and ragged splits."""
import os
import socket

import numpy as np
import pytest


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_items, q):
    import torch
    import torch.distributed as dist
    from audio_analysis_b200 import sharding
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = sharding.shard_bounds(n_items, rank, world)
    # "per-recording result" = a deterministic function of the global recording index
    idx = torch.arange(lo, hi, dtype=torch.float64)
    local = torch.stack([idx, idx * idx + 0.5, -idx], dim=1)
    full = sharding.gather_per_recording(local, n_items)
    # timing reduction as in bench.py: max over ranks
    t = torch.tensor([10.0 + rank], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    q.put((rank, full.numpy(), float(t.item())))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n_items", [8, 7, 1])
def test_two_rank_gather_equals_single_rank(n_items):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_items, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    idx = np.arange(n_items, dtype=np.float64)
    ref = np.stack([idx, idx * idx + 0.5, -idx], axis=1)
    for rank, full, tmax in got:
        assert np.array_equal(full, ref), rank
        assert tmax == 11.0


def test_shard_bounds_cover_everything_once():
    from audio_analysis_b200 import sharding
    for n in (0, 1, 5, 1024, 65536, 65537):
        for world in (1, 2, 4, 8):
            b = [sharding.shard_bounds(n, r, world) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            assert max(h - l for l, h in b) - min(h - l for l, h in b) <= 1


def test_shard_descriptors_slices_whole_recordings():
    from audio_analysis_b200 import sharding
    from audio_analysis_b200.engine import SEGMENT_DTYPE
    counts = [3, 0, 5, 2, 4]
    segs = np.zeros(sum(counts), dtype=SEGMENT_DTYPE)
    segs["pcm_offset"] = np.arange(len(segs))
    parts = [sharding.shard_descriptors(segs, counts, r, 2) for r in range(2)]
    assert [list(c) for _, c in parts] == [[3, 0], [5, 2, 4]]
    assert np.array_equal(np.concatenate([p for p, _ in parts]), segs)

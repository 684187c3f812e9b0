"""Recording-level data parallelism: every rank owns whole recordings, the kernels
need no exchange, and only small per-recording results are gathered at the end (torch.distributed:
NCCL on GPUs, gloo in the CPU tests).  Nothing here touches samples."""
import numpy as np


def shard_bounds(n_items, rank, world):
    """Contiguous block [lo, hi) of recordings owned by `rank` (sizes differ by at most one)."""
    assert 0 <= rank < world
    return (n_items * rank) // world, (n_items * (rank + 1)) // world


def shard_sizes(n_items, world):
    return [shard_bounds(n_items, r, world)[1] - shard_bounds(n_items, r, world)[0] for r in range(world)]


def gather_per_recording(local, n_items_total, group=None):
    """All-gather of per-recording result rows: `local` is this rank's [n_local, ...] tensor for its
    shard_bounds block; returns the [n_items_total, ...] tensor in recording order on every rank.
    Ragged shards are padded to the largest shard for the collective and trimmed afterwards."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        assert local.shape[0] == n_items_total
        return local
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    sizes = shard_sizes(n_items_total, world)
    assert local.shape[0] == sizes[rank], (local.shape, sizes, rank)
    biggest = max(sizes)
    if local.shape[0] < biggest:
        pad = torch.zeros((biggest - local.shape[0],) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        local = torch.cat([local, pad])
    out = torch.empty((world * biggest,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, local.contiguous(), group=group)
    if all(s == biggest for s in sizes):
        return out
    return torch.cat([out[r * biggest:r * biggest + sizes[r]] for r in range(world)])


def shard_descriptors(segments, counts_per_recording, rank, world):
    """Slice a segment descriptor table (engine.SEGMENT_DTYPE, recordings back to back with
    counts_per_recording[i] segments each) down to the recordings this rank owns."""
    counts = np.asarray(counts_per_recording, dtype=np.int64)
    lo, hi = shard_bounds(len(counts), rank, world)
    start = np.concatenate([[0], np.cumsum(counts)])
    return segments[start[lo]:start[hi]], counts[lo:hi]

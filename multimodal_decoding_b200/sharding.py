"""Vertex sharding across the GPUs of one box.

Centres are independent, so the only collective on the path is the final gather of the per-centre score block.
The partition is `np.array_split(arange(V), world_size)`: contiguous ranges, the first `V % world_size` one longer —
the chunking the reference's GroupIterator applies to its joblib workers (searchlight.py:130), so concatenating
the shards in rank order reproduces the reference's vertex order.
"""
from __future__ import annotations

import numpy as np


def split_bounds(n: int, parts: int):
    sizes = np.full(parts, n // parts, dtype=np.int64)
    sizes[: n % parts] += 1
    ends = np.cumsum(sizes)
    return [(int(e - s), int(e)) for s, e in zip(sizes, ends)]


def _world():
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank(), dist.get_world_size()
    except ImportError:
        pass
    return 0, 1


def shard_range(n_centres: int, rank: int | None = None, world_size: int | None = None):
    r, w = _world()
    rank = r if rank is None else rank
    world_size = w if world_size is None else world_size
    return split_bounds(n_centres, world_size)[rank]


def gather_scores(local_scores: np.ndarray, n_total: int) -> np.ndarray:
    """All-gather of the (V_shard x 8) float64 score blocks into (V x 8) on every rank (NCCL when the process
    group is NCCL — device tensors over NVLink —, gloo on CPU in the tests)."""
    rank, world = _world()
    if world == 1:
        assert len(local_scores) == n_total
        return local_scores
    import torch
    import torch.distributed as dist
    bounds = split_bounds(n_total, world)
    width = local_scores.shape[1] if local_scores.ndim == 2 else 8
    longest = max(hi - lo for lo, hi in bounds)
    on_gpu = dist.get_backend() == "nccl"
    dev = torch.device("cuda", torch.cuda.current_device()) if on_gpu else torch.device("cpu")
    mine = torch.zeros((longest, width), dtype=torch.float64, device=dev)
    mine[: len(local_scores)] = torch.from_numpy(np.ascontiguousarray(local_scores)).to(dev)
    everyone = torch.empty((world, longest, width), dtype=torch.float64, device=dev)
    dist.all_gather_into_tensor(everyone, mine)
    everyone = everyone.cpu().numpy()
    return np.concatenate([everyone[r, : hi - lo] for r, (lo, hi) in enumerate(bounds)])

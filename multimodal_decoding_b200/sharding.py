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


def spatial_order(coords, bits: int = 10) -> np.ndarray:
    """Permutation that sorts vertices along a Morton (Z-order) curve of their coordinates.

    fsaverage / icosphere vertex ids are in subdivision order, so a contiguous id range is scattered over the
    whole surface.  Sharding the *Morton-ordered* rows instead gives every GPU a compact patch, which keeps the
    neighbourhood unions of its centre clusters small (csrc/api.cu build_plan)."""
    c = np.asarray(coords, dtype=np.float64)
    lo, span = c.min(axis=0), np.ptp(c, axis=0)
    span[span == 0] = 1.0
    q = np.minimum(((c - lo) / span * (1 << bits)).astype(np.uint64), (1 << bits) - 1)
    code = np.zeros(len(c), dtype=np.uint64)
    for b in range(bits):
        for axis in range(3):
            code |= ((q[:, axis] >> np.uint64(b)) & np.uint64(1)) << np.uint64(3 * b + axis)
    return np.argsort(code, kind="stable")


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
    everyone = torch.empty((world * longest, width), dtype=torch.float64, device=dev)   # flat: accepted by NCCL and gloo
    dist.all_gather_into_tensor(everyone, mine)
    everyone = everyone.cpu().numpy().reshape(world, longest, width)
    return np.concatenate([everyone[r, : hi - lo] for r, (lo, hi) in enumerate(bounds)])


def device_copy_via_peers(a: np.ndarray):
    """(n x width) HOST array that every rank holds -> the same array as a CUDA tensor on this rank's GPU, moved by
    all ranks together: rank r copies only its r-th block of rows over PCIe and the blocks are all-gathered over NVLink
    (NCCL).  With N ranks the host-to-device traffic per rank falls to 1/N of the array.  Returns None when there is no NCCL
    process group (the caller then hands the host array to the library, which copies all of it)."""
    rank, world = _world()
    if world == 1:
        return None
    import torch
    import torch.distributed as dist
    if dist.get_backend() != "nccl":
        return None
    a = np.asarray(a)
    n, width = a.shape
    bounds = split_bounds(n, world)
    longest = max(hi - lo for lo, hi in bounds)
    dev = torch.device("cuda", torch.cuda.current_device())
    lo, hi = bounds[rank]
    mine = torch.zeros((longest, width), dtype=torch.from_numpy(a[:0]).dtype, device=dev)
    mine[: hi - lo].copy_(torch.from_numpy(a[lo:hi]), non_blocking=True)
    everyone = torch.empty((world * longest, width), dtype=mine.dtype, device=dev)
    dist.all_gather_into_tensor(everyone, mine)
    if all(b - a_ == longest for a_, b in bounds):
        return everyone[:n]
    return torch.cat([everyone[r * longest: r * longest + (b - a_)] for r, (a_, b) in enumerate(bounds)])

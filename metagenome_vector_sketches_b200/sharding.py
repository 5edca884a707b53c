"""Rank sharding of a CSR batch: samples are independent, so each rank
(one process per GPU) takes one contiguous, size-balanced range of samples and there is no
data-path collective.  The partition itself is computed by the C++ driver code
(mvs_partition_samples) so the CLI (--gpus) and the torch.distributed launchers agree."""
import numpy as np

from ._lib import load_library


def partition_bounds(offsets, world):
    offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
    n = len(offsets) - 1
    bounds = np.zeros(world + 1, dtype=np.int64)
    load_library().mvs_partition_samples(offsets.ctypes.data, n, int(world), bounds.ctypes.data)
    return bounds


def shard_for_rank(hashes, offsets, rank, world):
    """Returns (first_sample, hashes_view, local_offsets) of `rank`'s shard."""
    b = partition_bounds(offsets, world)
    a, e = int(b[rank]), int(b[rank + 1])
    h0, h1 = int(offsets[a]), int(offsets[e])
    return a, hashes[h0:h1], (np.asarray(offsets[a:e + 1], dtype=np.uint64) - np.uint64(h0))


def max_over_ranks(x, group=None):
    """Timing rule of bench.py: the job's step time is the slowest rank's."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return float(x)
    dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    t = torch.tensor([float(x)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())

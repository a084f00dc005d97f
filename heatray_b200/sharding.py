"""Multi-GPU partitioning of the progressive passes. Pure host logic, no device code:
 - sample-index sharding: rank r of G renders sampleIndex r, r+G, ...; one fp32 sum-reduce of the accumulation buffers;
 - screen-band sharding: interleaved bands of rows, rank = band mod G; disjoint pixels, gathered by a sum of zero-padded buffers.
The reduce is a torch.distributed collective (NCCL over NVLink on the GPU box, gloo in the CPU tests)."""
from __future__ import annotations

import numpy as np


def sample_indices(rank: int, world: int, steps: int, table_length: int):
    """Sample indices rendered by `rank` in `steps` steps; wraps into the QMC table like getSequenceValue does."""
    return [(rank + s * world) % table_length for s in range(steps)]


def band_rows(height: int, band_height: int, rank: int, world: int) -> np.ndarray:
    """Boolean mask of the rows owned by `rank` under interleaved band sharding."""
    return (np.arange(height) // max(1, band_height)) % max(1, world) == rank


def reduce_accumulation(tensor, dst: int = 0):
    """Sum the per-rank fp32 RGBA accumulation buffers into rank `dst` (alpha adds up to the total sample count)."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(tensor, dst=dst, op=dist.ReduceOp.SUM)
    return tensor

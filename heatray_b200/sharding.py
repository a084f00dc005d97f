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


_ROWS = {}      # row-index tensors per (height, band height, world, device)


def gather_bands(acc, band_height: int, rank: int, world: int, dst: int = 0):
    """Screen-band exchange step: every rank sends ONLY the rows it rendered (frame / world bytes) and rank `dst` puts them in
    place; the other ranks' buffers are left as they are. `acc` is the (H, W, 4) fp32 accumulation buffer of this rank (pixels
    outside its bands are zero). Bands are disjoint, so the result is bit-identical to a single-rank render
    ("ncclGather of disjoint bands" instead of a sum of zero-padded frames)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_initialized() and world > 1):
        return acc
    height = acc.shape[0]
    key = (height, band_height, world, str(acc.device))
    if key not in _ROWS:
        _ROWS[key] = [torch.from_numpy(np.flatnonzero(band_rows(height, band_height, r, world))).to(acc.device) for r in range(world)]
    rows = _ROWS[key]
    most = max(len(x) for x in rows)
    mine = torch.zeros((most,) + tuple(acc.shape[1:]), dtype=acc.dtype, device=acc.device)      # equal-sized pieces for the collective
    mine[:len(rows[rank])] = acc[rows[rank]]
    pieces = [torch.empty_like(mine) for _ in range(world)] if rank == dst else None
    dist.gather(mine, pieces, dst=dst)
    if rank == dst:
        for r in range(world):
            if r != dst:
                acc[rows[r]] = pieces[r][:len(rows[r])]
    return acc

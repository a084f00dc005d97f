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


class PeerExchange:
    """The exchange step as ONE kernel per rank over NVLink peer memory (csrc/peer_exchange.cu, hrcExchange* in include/heatray_cuda.h)
    instead of an NCCL collective. torch.distributed is used once, to hand the CUDA-IPC handles round (plumbing); every later call
    only enqueues a device copy and a kernel on the given CUDA stream.

        ex = PeerExchange(lib, device, rank, world, height, width)        # collective: every rank constructs it
        ex.sum(acc_ptr, stream)  /  ex.bands(acc_ptr, band_height, stream)  # collective: same number of calls on every rank
        frame = ex.result()                                               # rank 0: (H, W, 4) fp32 torch view of the combined frame
    """

    def __init__(self, lib, device: int, rank: int, world: int, height: int, width: int):
        import ctypes as C
        import torch
        import torch.distributed as dist
        from . import capi
        self.lib, self.rank, self.world, self.height, self.width, self.device = lib, rank, world, height, width, device
        self._C = C
        for name, args in (("hrcExchangeCreate", [C.c_int32, C.c_int32, C.c_int32, C.c_uint64, C.POINTER(C.c_void_p), C.c_void_p]),
                           ("hrcExchangeConnect", [C.c_void_p, C.c_void_p]), ("hrcExchangeSum", [C.c_void_p, C.c_void_p, C.c_void_p]),
                           ("hrcExchangeBands", [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]), ("hrcExchangeResult", [C.c_void_p, C.POINTER(C.c_void_p)]),
                           ("hrcExchangeStatus", [C.c_void_p]), ("hrcExchangeDestroy", [C.c_void_p])):
            getattr(lib, name).argtypes = args
            getattr(lib, name).restype = C.c_int
        lib.hrcExchangeGetLastError.argtypes = [C.c_void_p]
        lib.hrcExchangeGetLastError.restype = C.c_char_p
        self.ex = C.c_void_p()
        handle = (C.c_uint8 * capi.EXCHANGE_HANDLE_BYTES)()
        self._check(lib.hrcExchangeCreate(device, rank, world, height * width, C.byref(self.ex), handle), created=False)
        if world > 1:
            gathered = [None] * world
            dist.all_gather_object(gathered, bytes(handle))
            blob = b"".join(gathered)
            self._check(lib.hrcExchangeConnect(self.ex, C.c_char_p(blob)))
            dist.barrier()                                    # every rank has mapped every peer before the first exchange
        self._result = None
        if rank == 0:
            p = C.c_void_p()
            self._check(lib.hrcExchangeResult(self.ex, C.byref(p)))

            class _Arr:
                __cuda_array_interface__ = {"shape": (height, width, 4), "typestr": "<f4", "data": (p.value, False), "version": 2}
            self._result = torch.as_tensor(_Arr(), device=f"cuda:{device}")

    def _check(self, rc, created=True):
        if rc != 0:
            from . import capi
            raise capi.HrcError(rc, (self.lib.hrcExchangeGetLastError(self.ex if created else None) or b"").decode())

    def sum(self, src_ptr: int, stream: int = 0):
        self._check(self.lib.hrcExchangeSum(self.ex, self._C.c_void_p(src_ptr), self._C.c_void_p(stream)))

    def bands(self, src_ptr: int, band_height: int, stream: int = 0):
        self._check(self.lib.hrcExchangeBands(self.ex, self._C.c_void_p(src_ptr), self.width, band_height, self._C.c_void_p(stream)))

    def result(self):
        return self._result

    def status(self):
        self._check(self.lib.hrcExchangeStatus(self.ex))

    def close(self):
        if self.ex:
            self.lib.hrcExchangeDestroy(self.ex)
            self.ex = None

"""One rank of tests/test_zy_gpu_peer_exchange.py: RANK / WORLD_SIZE / MASTER_* from the environment, gloo for the handle exchange,
ranks spread over the visible GPUs (all on cuda:0 when there is only one: CUDA IPC works between processes on one device too)."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from heatray_b200 import capi, sharding        # noqa: E402


def frame(rank, call, h, w, device):
    g = torch.Generator().manual_seed(1000 * call + rank)
    return (torch.rand((h, w, 4), generator=g) * 3.0).to(device)


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    h, w, band = 270, 480, 8
    dist.init_process_group("gloo", rank=rank, world_size=world)
    dev = rank % torch.cuda.device_count()
    torch.cuda.set_device(dev)
    device = f"cuda:{dev}"
    lib = capi.load_library()
    try:
        ex = sharding.PeerExchange(lib, dev, rank, world, h, w)
    except capi.HrcError as e:
        if "cudaIpc" in str(e):                                    # a box that does not allow CUDA IPC between these processes: nothing to test
            print(f"rank {rank}: SKIP {e}", flush=True)
            os._exit(77)
        raise
    stream = torch.cuda.current_stream().cuda_stream
    ok = True
    for call in range(3):                                          # sums: three calls in a row, new data each time
        mine = frame(rank, call, h, w, device)
        torch.cuda.synchronize()
        ex.sum(mine.data_ptr(), stream)
        mine.zero_()                                               # the exchange works on its own snapshot: the source may change at once (stream order)
        torch.cuda.synchronize()
        if rank == 0:
            want = frame(0, call, h, w, device)
            for p in range(1, world):
                want = want + frame(p, call, h, w, device)         # rank order, like the kernel
            ok &= bool(torch.equal(ex.result(), want))
        dist.barrier()
    rows = torch.arange(h, device=device)
    for call in range(3, 5):                                       # bands: each rank holds only its rows, the rest is zero
        full = frame(99, call, h, w, device)                       # what a single GPU would have rendered
        own = ((rows // band) % world == rank)
        mine = full * own[:, None, None]
        torch.cuda.synchronize()
        ex.bands(mine.data_ptr(), band, stream)
        torch.cuda.synchronize()
        if rank == 0:
            ok &= bool(torch.equal(ex.result(), full))
        dist.barrier()
    ex.status()
    ex.close()
    dist.destroy_process_group()
    print(f"rank {rank}: {'OK' if ok else 'MISMATCH'}", flush=True)
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()

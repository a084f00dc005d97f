"""Exchange step alone, back to back (torchrun, one rank per GPU): the peer-memory kernel (hrcExchangeSum / hrcExchangeBands) against NCCL reduce /
all_reduce on the same fp32 frames. us per exchange, max over ranks, 50 exchanges after 10 warm-up."""
import os, sys
import torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from heatray_b200 import capi, sharding

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
lib = capi.load_library()
for (h, w, label) in ((1080, 1920, "1080p 33 MB"), (2160, 3840, "4K 133 MB")):
    frame = torch.rand((h, w, 4), device="cuda")
    work = frame.clone()
    ex = sharding.PeerExchange(lib, lr, rank, world, h, w)
    st = torch.cuda.current_stream().cuda_stream
    cases = {"peer sum": lambda: ex.sum(frame.data_ptr(), st), "peer bands(8)": lambda: ex.bands(frame.data_ptr(), 8, st),
             "nccl reduce": lambda: dist.reduce(work, dst=0), "nccl all_reduce": lambda: dist.all_reduce(work)}
    for name, fn in cases.items():
        for _ in range(10): fn()
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(50): fn()
        e1.record(); e1.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / 50 * 1e3], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0: print(f"{world} GPUs {label:12s} {name:16s} {t.item():8.1f} us", flush=True)
    ex.status(); ex.close()
dist.destroy_process_group()

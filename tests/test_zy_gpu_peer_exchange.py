"""hrcExchangeSum / hrcExchangeBands (csrc/peer_exchange.cu): the multi-GPU frame exchange as one kernel per rank over peer memory.
Ranks are separate processes (CUDA IPC needs them to be); they use different GPUs when the box has several, else they share cuda:0.
Bar: the sum equals the rank-ordered fp32 sum bit for bit; the band exchange reproduces the single-GPU frame bit for bit; five
exchanges in a row on the same buffers. The file sorts next to last on purpose: this is the only test that needs several
processes to share the GPU."""
import os
import socket
import subprocess
import sys
from pathlib import Path

import pytest

pytestmark = pytest.mark.gpu
WORKER = Path(__file__).parent / "_peer_exchange_worker.py"


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2, 3])
def test_peer_exchange_matches_rank_ordered_sum_and_single_gpu_bands(lib, world):
    port = _free_port()
    procs = []
    for rank in range(world):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, str(WORKER)], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = []
    try:
        for p in procs:
            out, _ = p.communicate(timeout=240)
            outs.append(out)
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
    if any(p.returncode == 77 for p in procs):
        pytest.skip("CUDA IPC between processes is not available on this box:\n" + outs[0][-500:])
    for rank, (p, out) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, f"rank {rank} failed:\n{out[-3000:]}"
        assert f"rank {rank}: OK" in out

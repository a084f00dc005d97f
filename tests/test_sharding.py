"""Multi-rank partitioning without a GPU: world_size-2 gloo processes shard a convergence run by sample index (and by
screen bands), reduce the accumulation buffers, and must reproduce the single-rank image. The renderer
behind each rank is the CPU oracle here; on the GPU box bench.py runs the same logic over NCCL."""
import os
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from heatray_b200 import sharding

ROOT = Path(__file__).resolve().parent.parent


def test_sample_index_partition_is_a_partition():
    for world in (1, 2, 4, 8):
        steps, table = 16, 128
        seen = sorted(i for r in range(world) for i in sharding.sample_indices(r, world, steps, table))
        assert seen == list(range(world * steps))
    assert sharding.sample_indices(1, 2, 3, 4) == [1, 3, 1]       # wraps like getSequenceValue (sequence.rlsl:25)


def test_band_rows_are_disjoint_and_cover():
    for world in (2, 3, 8):
        masks = [sharding.band_rows(1080, 64, r, world) for r in range(world)]
        assert np.array_equal(sum(m.astype(int) for m in masks), np.ones(1080, int))


def _worker(rank, world, port, mode, out_path):
    sys.path.insert(0, str(ROOT))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from heatray_b200 import renderer, scenes
    from oracle import loader
    spec = scenes.cornell_box(width=32, height=32, passes=4)
    b = loader.backend(threads=1)
    r = renderer.Renderer(b)
    r.load_scene(spec)
    if mode == "samples":
        for s in sharding.sample_indices(rank, world, spec.passes // world, spec.passes):
            r.render_pass(sample_index=s)
    else:
        r.set_band(4, rank, world)
        for s in range(spec.passes):
            r.render_pass(sample_index=s)
    acc = torch.from_numpy(r.pixels().copy())
    if mode == "bands_gather":
        sharding.gather_bands(acc, 4, rank, world, dst=0)
    else:
        sharding.reduce_accumulation(acc, dst=0)
    if rank == 0:
        np.save(out_path, acc.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["samples", "bands", "bands_gather"])
def test_two_rank_gloo_equals_single_rank(tmp_path, mode, oracle_mod):
    from heatray_b200 import renderer, scenes
    spec = scenes.cornell_box(width=32, height=32, passes=4)
    b = oracle_mod.backend(threads=1)
    r = renderer.Renderer(b)
    r.load_scene(spec)
    for s in range(spec.passes):
        r.render_pass(sample_index=s)
    single = r.pixels()
    b.close()
    out = tmp_path / "acc.npy"
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, mode, str(out)), nprocs=2, join=True)
    multi = np.load(out)
    assert np.array_equal(multi[..., 3], single[..., 3])
    if mode.startswith("bands"):
        assert np.array_equal(multi, single)                       # disjoint pixels: bit-identical
    else:
        assert np.allclose(multi, single, rtol=1e-6, atol=1e-7)    # same samples, different fp32 summation order

"""Traversal CTAs per SM for pipelined passes (HRC_TUNE_BLOCKS_PER_SM_PIPE) at 8 pass lanes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from heatray_b200 import capi, renderer, scenes
lib = capi.load_library()
specs = {"sphere_grid": scenes.sphere_grid(), "terrain": scenes.terrain(), "material_grid": scenes.material_grid()}
for name, band in (("sphere_grid", None), ("material_grid", None), ("terrain", (8, 0, 8))):
    for lanes, blocks in ((4, 4), (8, 2), (8, 3), (8, 4), (8, 6), (8, 8)):
        os.environ["HRC_TUNE_PASS_LANES"] = str(lanes); os.environ["HRC_TUNE_BLOCKS_PER_SM_PIPE"] = str(blocks)
        b = capi.Backend(lib, "hrc", device=0, collect_stats=False)
        r = renderer.Renderer(b); r.load_scene(specs[name])
        if band: r.set_band(*band)
        n = specs[name].passes
        for i in range(16): r.enqueue_pass(sample_index=i % n)
        b.synchronize(); b.clear(); b.reset_stats()
        b.timer_start()
        for i in range(32): r.enqueue_pass(sample_index=i % n)
        ms = b.timer_stop()
        st = b.get_stats(); rays = st.raysPrimary + st.raysExtension + st.raysOcclusion
        print("%-13s band %-10s lanes %d blocks/SM %d: %.3f ms/pass %.0f Mrays/s" % (name, band, lanes, blocks, ms / 32, rays / ms / 1e3), flush=True)
        b.close()

"""Pass lanes (HRC_TUNE_PASS_LANES) on one GPU: whole frames of configs #2 / #4 and ONE rank's share of an 8-way band split of config #4
(what each GPU of the 8-GPU strong-scaling run sees). Mrays/s over 32 enqueued passes."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from heatray_b200 import capi, renderer, scenes
lib = capi.load_library()
specs = {"sphere_grid": scenes.sphere_grid(), "terrain": scenes.terrain()}
for name, band in (("sphere_grid", None), ("terrain", None), ("terrain", (16, 0, 8)), ("terrain", (8, 0, 8))):
    for lanes in (1, 2, 4, 6, 8):
        os.environ["HRC_TUNE_PASS_LANES"] = str(lanes)
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
        print("%-12s band %-10s lanes %d: %.3f ms/pass %.0f Mrays/s" % (name, band, lanes, ms / 32, rays / ms / 1e3), flush=True)
        b.close()

"""Distribution of per-pixel relative differences between the CUDA path and the oracle (experiment helper, not a test)."""
import sys, time
sys.path.insert(0, ".")
import numpy as np
from heatray_b200 import capi, renderer, scenes
from oracle import loader
lib = capi.load_library()
def run(spec, passes):
    out = []
    for mk in ("cpu", "gpu"):
        b = loader.backend() if mk == "cpu" else capi.Backend(lib, "hrc")
        r = renderer.Renderer(b); r.load_scene(spec)
        b.set_option(capi.OPT_PIXEL_RAY_COUNTS, 1); b.clear()
        for i in range(passes): r.render_pass(sample_index=i)
        px = r.pixels().copy(); rc = b.ray_counts().copy()
        out.append((px, rc)); b.close()
    (pc, rc), (pg, rg) = out
    worst = (np.abs(pg[..., :3] - pc[..., :3]) / np.maximum(np.abs(pc[..., :3]), 1e-2)).max(-1)
    print(spec.name, spec.width, "within 1e-6/5/4/3/2:", " ".join("%.5f" % (worst <= t).mean() for t in (1e-6, 1e-5, 1e-4, 1e-3, 1e-2)), "ray count differs %.5f" % (rc != rg).mean(), "bit-equal %.4f" % (pg == pc).all(-1).mean())
small = len(sys.argv) > 1
run(scenes.sphere_grid(width=480, height=270, grid=4, u_slices=32, v_slices=16, env_size=(512, 256)), 4)
run(scenes.material_grid(width=480, height=270, grid=5, u_slices=32, v_slices=16), 4)
if not small:
    run(scenes.sphere_grid(), 2)
    run(scenes.material_grid(), 2)

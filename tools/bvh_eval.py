"""Development tool: node steps / triangle tests per ray of the product's BVH8 builder on the CPU (tools/bvh_eval.cpp), on rays captured from an
oracle pass over the same scene (primary, extension and occlusion rays of real paths). No GPU needed.
    python tools/bvh_eval.py [sphere_grid|material_grid|terrain] ...      (env HRC_TUNE_* knobs of the builder apply)"""
import os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from heatray_b200 import capi, renderer, scenes
from oracle import loader

EXE = "/tmp/bvh_eval"
CONFIGS = {"sphere_grid": lambda: scenes.sphere_grid(), "material_grid": lambda: scenes.material_grid(), "terrain": lambda: scenes.terrain(n=1024, blobs=200),
           "sphere_grid_1m": lambda: scenes.sphere_grid_1m()}


def dump(name, path):
    spec = CONFIGS[name]()
    tris = []
    for m in spec.meshes:
        M = np.asarray(m.transform, np.float32).reshape(4, 4)          # mathematical 4x4 (row-major numpy), translation in the last column
        p = np.asarray(m.positions, np.float32).reshape(-1, 3)
        w = p @ M[:3, :3].T + M[:3, 3]
        idx = np.asarray(m.indices).reshape(-1)
        if m.draw_mode == capi.DRAW_TRIANGLES:
            tri = idx.reshape(-1, 3)
        else:                                                          # triangle strip (Mesh.cpp draw mode): (i, i+1, i+2), degenerate ones dropped
            tri = np.stack([idx[:-2], idx[1:-1], idx[2:]], 1)
            tri = tri[(tri[:, 0] != tri[:, 1]) & (tri[:, 1] != tri[:, 2]) & (tri[:, 0] != tri[:, 2])]
        tris.append(w[tri].reshape(-1, 9))
    tris = np.concatenate(tris).astype(np.float32)
    b = loader.backend()
    try:
        r = renderer.Renderer(b)
        r.load_scene(spec, width=240, height=135)
        loader.capture_rays(b, 2_000_000)
        for i in range(2):
            r.render_pass(sample_index=i)
        rays, kinds = loader.captured_rays(b)
    finally:
        b.close()
    rec = np.zeros(len(rays), np.dtype([("o", "<f4", 3), ("tmin", "<f4"), ("d", "<f4", 3), ("tmax", "<f4"), ("kind", "<i4")]))
    rec["o"], rec["tmin"], rec["d"], rec["tmax"], rec["kind"] = rays["origin"], rays["tmin"], rays["direction"], rays["tmax"], kinds
    with open(path, "wb") as f:
        f.write(np.uint64(len(tris)).tobytes()); f.write(tris.tobytes()); f.write(np.uint64(len(rec)).tobytes()); f.write(rec.tobytes())
    return len(tris), np.bincount(kinds, minlength=3)


if __name__ == "__main__":
    subprocess.run(["g++", "-O2", "-std=c++17", "-pthread", f"{ROOT}/tools/bvh_eval.cpp", f"{ROOT}/heatray_b200/csrc/bvh_builder.cpp", "-o", EXE], check=True)
    for name in sys.argv[1:] or ["sphere_grid"]:
        path = f"/tmp/bvh_eval_{name}.bin"
        if not os.path.exists(path):
            print(name, dump(name, path))
        print(f"== {name}"); sys.stdout.flush()
        subprocess.run([EXE, path], check=True)

#!/usr/bin/env python
"""Writes tests/golden/shader_ref.npz from oracle/_ref/libref_shaders.so — the reference's shader programs compiled as C++
from the text under /root/reference (oracle/ref_probe/ref_shaders.cpp). Run in the build container after build_ref.sh.
The inputs are regenerated from fixed seeds by oracle/shader_probe.py (material_suite / miss_suite); the file stores the
reference side's answers plus a digest of the inputs, so the test notices if the generator ever drifts."""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle import loader, shader_probe as sp          # noqa: E402

SEED, CASES = 20261105, 96


def main():
    ref = sp.RefSide(loader.ref_lib("shaders"))
    out = {}
    digests = []
    for i, (mtype, flags, params, world, cases, si) in enumerate(sp.material_suite(SEED, CASES)):
        ref.set_world(world)
        r = ref.shade(mtype, flags, params, si, cases)
        out[f"shade_{i}"] = r
        digests.append(sp.input_digest(cases, world.seq, bytes(params)))
    for i, (world, rays) in enumerate(sp.miss_suite(SEED + 1, 512)):
        ref.set_world(world)
        out[f"miss_{i}"] = ref.miss(rays)
        digests.append(sp.input_digest(rays))
    out["digests"] = np.array(digests)
    out["meta"] = np.array([SEED, CASES])
    path = ROOT / "tests" / "golden" / "shader_ref.npz"
    np.savez_compressed(path, **out)
    print(path, path.stat().st_size, "bytes,", len(digests), "groups")


if __name__ == "__main__" and "--vertex" not in sys.argv:
    main()


def vertex_golden():
    """tests/golden/vertex_ref.npz: vertex.rlsl compiled as C++ (oracle/_ref/libref_shaders.so::ref_vertex) on vertex_cases(SEED + 2, 400) per #define set."""
    lib = loader.ref_lib("shaders")
    cases = sp.vertex_cases(SEED + 2, 400)
    out = {"digest": np.array(sp.input_digest(cases))}
    for f in sp.VERTEX_VARIANTS:
        out[f"vertex_{f}"] = sp.ref_vertex(lib, f, cases)
    path = ROOT / "tests" / "golden" / "vertex_ref.npz"
    np.savez_compressed(path, **out)
    print(path, path.stat().st_size, "bytes")


if __name__ == "__main__" and "--vertex" in sys.argv:
    vertex_golden()

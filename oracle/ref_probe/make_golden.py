#!/usr/bin/env python
"""Writes tests/golden/*.json from oracle/_ref/*.so, i.e. from the reference's OWN code compiled in place
(Random.h, MultiScatterUtil.cpp, utility.rlsl + brdfs.rlsl through glm, OrbitCamera.h + glm). Run in the build container
(where /root/reference is mounted) after oracle/ref_probe/build_ref.sh; the JSON files are committed and are what the
tests on the GPU box compare against. Floats are stored as uint32 bit patterns so the comparison is bit-exact."""
import ctypes as C
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
REF = ROOT / "oracle" / "_ref"
OUT = ROOT / "tests" / "golden"
u32, i32, f32 = C.c_uint32, C.c_int32, C.c_float


def bits(a):
    return np.asarray(a, np.float32).view(np.uint32).reshape(-1).tolist()


def qmc():
    lib = C.CDLL(str(REF / "libref_qmc.so"))
    out = {}

    def run(fn, n, *args):
        a = np.zeros((n, 2), np.float32)
        fn(a.ctypes.data_as(C.c_void_p), *args)
        return bits(a)
    for s in (0, 1, 3, 7, 15):
        for n in (8, 64, 257):
            out[f"sobol/{s}/{n}"] = run(lib.ref_sobol, n, u32(n), u32(s))
            out[f"halton/{s}/{n}"] = run(lib.ref_halton, n, u32(n), i32(s))
            out[f"hammersley/{s}/{n}"] = run(lib.ref_hammersley, n, u32(n), i32(s))
            out[f"radial_sobol/{s}/{n}"] = run(lib.ref_radial_sobol, n, u32(n), u32(s))
            out[f"uniform/{s}/{n}"] = run(lib.ref_uniform, n, u32(n), u32(s), f32(0.0), f32(1.0))
            for e in (5, 6, 8):
                out[f"polygonal{e}/{s}/{n}"] = run(lib.ref_random_polygonal, n, u32(e), u32(n), u32(s))
        out[f"blue_noise/{s}/48"] = run(lib.ref_blue_noise, 48, u32(48), i32(s))
    (OUT / "qmc_ref.json").write_text(json.dumps(out))


def lut():
    lib = C.CDLL(str(REF / "libref_lut.so"))
    lib.ref_lut_value.restype = f32
    pts = [(0, 0), (0, 127), (127, 0), (127, 1), (127, 2), (127, 3), (127, 127), (64, 64), (10, 100), (100, 10), (33, 77), (5, 5)]
    out = {f"{r}/{c}": bits([lib.ref_lut_value(r, c, 128, 4096)])[0] for r, c in pts}
    rows = np.zeros((2, 128), np.float32)
    lib.ref_lut_rows(rows.ctypes.data_as(C.c_void_p), 50, 52, 128, 4096)
    out["rows50_51"] = bits(rows)
    (OUT / "lut_ref.json").write_text(json.dumps(out))


def rlsl():
    lib = C.CDLL(str(REF / "libref_rlsl.so"))
    rng = np.random.default_rng(1234)
    cases = []
    # the survey's own probe point first
    first = [0.3, 0.8, -0.2, 0.25, 0.6, 0.5, 0.7, 0.1, 0.25, 0.9, 0.4, 0.7, 1 / 1.57, 0.6, 0.04, 0.3, 10.5 / 512, 20.5 / 512]
    inputs = [first]
    for _ in range(63):
        n = rng.normal(size=3); v = rng.normal(size=3); v[1] = abs(v[1]) + 0.05
        inputs.append([*n, rng.uniform(), rng.uniform(), *v, rng.uniform(0.0001, 1.0), rng.uniform(), rng.uniform(0.01, 1), rng.uniform(0.01, 1),
                       rng.uniform(0.4, 2.2), rng.uniform(), rng.uniform(0, 0.2), rng.uniform(), rng.uniform(), rng.uniform()])
    for inp in inputs:
        a = np.asarray(inp, np.float32)
        o = np.zeros(18, np.float32)
        lib.ref_rlsl_probe(a.ctypes.data_as(C.c_void_p), o.ctypes.data_as(C.c_void_p))
        cases.append({"in": bits(a), "out": bits(o)})
    (OUT / "rlsl_helpers.json").write_text(json.dumps(cases))


def host():
    lib = C.CDLL(str(REF / "libref_host.so"))
    rng = np.random.default_rng(99)
    cams, lights = [], []
    params = [(19.0, 0.0, 0.0, (0.0, 0.0, 0.0)), (19.0, 0.0, 0.6, (0.0, 0.0, 0.0)), (5.4, 0.0, 0.0, (0.0, 1.0, 0.0))]
    params += [(float(rng.uniform(1, 50)), float(rng.uniform(0, 6.28)), float(rng.uniform(-1.5, 1.5)), tuple(map(float, rng.uniform(-3, 3, 3)))) for _ in range(13)]
    for d, phi, th, tgt in params:
        t = np.asarray(tgt, np.float32); o = np.zeros(16, np.float32)
        lib.ref_orbit_view(f32(d), f32(phi), f32(th), t.ctypes.data_as(C.c_void_p), o.ctypes.data_as(C.c_void_p))
        cams.append({"in": bits([d, phi, th, *tgt]), "out": bits(o)})
    angles = [(np.pi / 2, 0.0), (0.0, 0.0), (np.pi / 3, 0.4)] + [(float(rng.uniform(-3, 3)), float(rng.uniform(0, 6.28))) for _ in range(13)]
    for th, phi in angles:
        o = np.zeros(3, np.float32)
        lib.ref_light_forward(f32(th), f32(phi), o.ctypes.data_as(C.c_void_p))
        lights.append({"in": bits([th, phi]), "out": bits(o)})
    # ProcessLight's euler-angle recovery for imported lights (AssimpMeshProvider.cpp:564-586) on random rigid + translated node transforms
    la = []
    for k in range(24):
        q = rng.normal(size=4); q /= np.linalg.norm(q)
        x, y, z, w = q
        R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                      [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                      [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
        M = np.eye(4); M[:3, :3] = R
        M[:3, 3] = rng.uniform(-4, 4, 3) if k % 2 else 0.0
        m = np.ascontiguousarray(M.T, np.float32).reshape(-1)          # column-major
        o = np.zeros(2, np.float32)
        lib.ref_light_angles(m.ctypes.data_as(C.c_void_p), o.ctypes.data_as(C.c_void_p))
        la.append({"in": bits(m), "out": bits(o)})
    # SceneTransform::transform() (HeatrayRenderer.h:343-346) and util::AABB::center/radius/valid under that transform (AABB.h:34-52)
    st, ab = [], []
    for k in range(16):
        y, p, r = (float(v) for v in rng.uniform(-3.2, 3.2, 3)) if k else (0.0, 0.0, 0.0)
        sc = float(rng.uniform(0.01, 20.0)) if k else 1.0
        o = np.zeros(16, np.float32)
        lib.ref_scene_transform(f32(y), f32(p), f32(r), f32(sc), o.ctypes.data_as(C.c_void_p))
        st.append({"in": bits([y, p, r, sc]), "out": bits(o)})
        mn = rng.uniform(-5, 1, 3).astype(np.float32); mx = (mn + rng.uniform(-0.5, 6, 3)).astype(np.float32)
        if k == 3:
            mx = mn.copy()                                               # invalid box
        o5 = np.zeros(6, np.float32)
        lib.ref_aabb_camera(mn.ctypes.data_as(C.c_void_p), mx.ctypes.data_as(C.c_void_p), o.ctypes.data_as(C.c_void_p), o5.ctypes.data_as(C.c_void_p))
        ab.append({"min": bits(mn), "max": bits(mx), "transform": bits(o), "out": bits(o5)})
    (OUT / "host_glm_ref.json").write_text(json.dumps({"orbit_view": cams, "light_forward": lights, "light_angles": la,
                                                       "scene_transform": st, "aabb_camera": ab}))


def display():
    """displayGL.frag (compiled as C++) at random texels / uv / parameter sets, plus the defaults and each stage alone."""
    lib = C.CDLL(str(REF / "libref_display.so"))
    rng = np.random.default_rng(4242)
    #          tonemap exposure bright contrast hue  sat  vib  r    g    b    vigI vigF
    default = [0.0, 1.0, 0.0, 1.0, 1.0, 1.0, 0.0, 1.0, 1.0, 1.0, 0.0, 1.0]      # HeatrayRenderer.h:104-117
    sets = [default, [1.0] + default[1:]]
    for i, val in ((1, 2.0), (2, 0.15), (3, 1.4), (4, 0.7), (5, 1.6), (6, 0.8), (7, 0.5), (8, 1.3), (9, 0.6), (10, 1.2), (11, 0.4)):
        q = list(default); q[i] = val; sets.append(q)
    for _ in range(12):
        sets.append([float(rng.integers(0, 2)), float(2.0 ** rng.uniform(-2, 2)), float(rng.uniform(-0.3, 0.3)), float(rng.uniform(0.5, 1.6)), float(rng.uniform(0.3, 1.5)),
                     float(rng.uniform(0, 2)), float(rng.uniform(0, 1)), float(rng.uniform(0.3, 1.5)), float(rng.uniform(0.3, 1.5)), float(rng.uniform(0.3, 1.5)),
                     float(rng.uniform(0, 2)), float(rng.uniform(0.2, 1.5))])
    cases = []
    for q in sets:
        for k in range(24):
            n = float(rng.integers(1, 65))
            rgb = rng.uniform(0, 1, 3) ** 2 * n * (float(rng.uniform(0, 4)) if k % 3 else 1.0)
            if k == 0: rgb = np.zeros(3)
            if k == 1: rgb = np.full(3, 0.5 * n)                      # grey: d == 0 in the HSV conversion
            texel = np.asarray([*rgb, n], np.float32)
            uv = np.asarray([rng.uniform(), rng.uniform()], np.float32)
            pa = np.asarray(q, np.float32); o = np.zeros(4, np.float32)
            lib.ref_display(texel.ctypes.data_as(C.c_void_p), uv.ctypes.data_as(C.c_void_p), pa.ctypes.data_as(C.c_void_p), o.ctypes.data_as(C.c_void_p))
            cases.append({"texel": bits(texel), "uv": bits(uv), "params": bits(pa), "out": bits(o)})
    (OUT / "display_ref.json").write_text(json.dumps(cases))


if __name__ == "__main__":
    if not (REF / "libref_qmc.so").exists():
        sys.exit("oracle/_ref is missing: run oracle/ref_probe/build_ref.sh where /root/reference is mounted")
    OUT.mkdir(parents=True, exist_ok=True)
    qmc(); lut(); rlsl(); host(); display()
    print("golden vectors written to", OUT)

"""Golden session files written by the reference's own Session class (oracle/_ref/libref_session.so, built from
/root/reference/Source/HeatrayRenderer/Session/Session.cpp + its vendored tinyxml2). Run in the build container only:
    python oracle/ref_probe/make_session_golden.py
Writes tests/golden/session_ref_<k>.xml and tests/golden/session_ref.json (the values that went in and what the reference read back)."""
import ctypes as C
import json
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
ref = C.CDLL(str(ROOT / "oracle/_ref/libref_session.so"))
N = ref.ref_session_count()
assert N == 50
KINDS = "".join("bufs"[ref.ref_session_kind(i)] for i in range(N))     # per variable in table order (Session.h:21-78)


def case(seed):
    rng = np.random.default_rng(seed)
    vals = []
    for k in KINDS:
        if k == "b":
            vals.append(float(rng.integers(0, 2)))
        elif k == "u":
            vals.append(float(rng.integers(0, 5000)))
        elif k == "f":
            mag = 10.0 ** rng.integers(-6, 7)
            vals.append(float(np.float32(rng.normal() * mag)))
        else:
            vals.append(0.0)
    return vals


CASES = [
    (case(1), "Multi-Material", "studio.hdr"),
    (case(2), "/home/user/My \"Scenes\" & <things>/sponza's.gltf", "solid color"),
    (case(3), "C:\\assets\\helmet.glb", "<none>"),
    ([0.0 if k != "f" else float(np.float32(v)) for k, v in zip(KINDS, np.linspace(-1e-7, 3.4e38, N))], "", "noon_grass.exr"),
]

out = []
for i, (vals, scene, env) in enumerate(CASES):
    path = ROOT / f"tests/golden/session_ref_{i}.xml"
    arr = (C.c_double * N)(*vals)
    assert ref.ref_session_write(str(path).encode(), arr, scene.encode(), env.encode()) == 1
    back = (C.c_double * N)()
    s, e = C.create_string_buffer(1024), C.create_string_buffer(1024)
    assert ref.ref_session_read(str(path).encode(), back, s, e) == 1
    out.append({"file": path.name, "written": vals, "scene": scene, "env": env,
                "read_back": list(back), "scene_back": s.value.decode(), "env_back": e.value.decode()})
(ROOT / "tests/golden/session_ref.json").write_text(json.dumps({"kinds": KINDS, "cases": out}, indent=1))
print("wrote", len(out), "session files")

"""The product's acceleration-structure build (heatray_b200/csrc/bvh_builder.cpp: binned-SAH BVH2, 8-wide collapse, 80-byte quantised nodes with
octant-ordered slots) checked WITHOUT a GPU: tools/bvh_eval.cpp links the builder and walks the compressed nodes on the CPU in k_trace's order
(nearest pending child first, conservative quantised slab test, leaf triangles after the node step). Every ray's answer must equal the
brute-force answer over all triangles, bit for bit (same triangle test): a lost child, a wrong slot permutation, a box quantised too tight
or a broken depth guard shows up as a mismatch."""
import subprocess
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
RAY = np.dtype([("o", "<f4", 3), ("tmin", "<f4"), ("d", "<f4", 3), ("tmax", "<f4"), ("kind", "<i4")])


@pytest.fixture(scope="module")
def bvh_eval(tmp_path_factory):
    exe = tmp_path_factory.mktemp("bvh") / "bvh_eval"
    subprocess.run(["g++", "-O2", "-std=c++17", "-pthread", str(ROOT / "tools" / "bvh_eval.cpp"), str(ROOT / "heatray_b200" / "csrc" / "bvh_builder.cpp"), "-o", str(exe)], check=True)
    return exe


def _rays(rng, n, lo, hi):
    r = np.zeros(n, RAY)
    r["o"] = rng.uniform(lo, hi, (n, 3))
    d = rng.normal(size=(n, 3)); r["d"] = d / np.linalg.norm(d, axis=1, keepdims=True)
    axis = rng.random(n) < 0.15                                    # some axis-parallel rays: zero direction components in the slab test
    r["d"][axis] = np.eye(3, dtype=np.float32)[rng.integers(0, 3, axis.sum())] * rng.choice([-1.0, 1.0], (axis.sum(), 1))
    r["tmax"] = np.where(rng.random(n) < 0.3, rng.uniform(0.1, 5.0, n), np.inf)
    r["kind"] = rng.choice([1, 2], n)                              # closest-hit and any-hit
    return r


def _scene(kind, rng):
    if kind == "soup":                                             # random triangles of very different sizes
        c = rng.uniform(-5, 5, (3000, 1, 3)); s = rng.choice([0.01, 0.3, 4.0], (3000, 1, 1))
        return (c + rng.normal(size=(3000, 3, 3)) * s).astype(np.float32), (-7, 7)
    if kind == "grid":                                             # a flat regular grid: flat boxes (zero extent on one axis), many ties
        n = 40
        x, z = np.meshgrid(np.arange(n, dtype=np.float32), np.arange(n, dtype=np.float32))
        p = np.stack([x, np.zeros_like(x), z], -1)
        a, b, c, d = p[:-1, :-1], p[1:, :-1], p[:-1, 1:], p[1:, 1:]
        return np.concatenate([np.stack([a, b, c], 2).reshape(-1, 3, 3), np.stack([b, d, c], 2).reshape(-1, 3, 3)]).astype(np.float32), (-2, 42)
    if kind == "skewed":                                           # sizes spread over six decades along one axis: the SAH peels one triangle per level (depth guard)
        k = np.arange(600, dtype=np.float64)
        c = np.stack([1.5 ** (k / 12.0), np.zeros_like(k), np.zeros_like(k)], 1)[:, None, :]
        return (c + rng.normal(size=(600, 3, 3)) * (0.2 * 1.5 ** (k / 12.0))[:, None, None]).astype(np.float32), (-10, 200)
    if kind == "tiny":
        return rng.normal(size=(2, 3, 3)).astype(np.float32), (-3, 3)
    raise ValueError(kind)


@pytest.mark.parametrize("kind", ["soup", "grid", "skewed", "tiny"])
def test_wide_bvh_walk_equals_brute_force(bvh_eval, tmp_path, kind):
    rng = np.random.default_rng({"soup": 11, "grid": 12, "skewed": 13, "tiny": 14}[kind])
    tris, (lo, hi) = _scene(kind, rng)
    rays = _rays(rng, 4000, lo, hi)
    path = tmp_path / f"{kind}.bin"
    with open(path, "wb") as f:
        f.write(np.uint64(len(tris)).tobytes()); f.write(np.ascontiguousarray(tris.reshape(-1, 9)).tobytes())
        f.write(np.uint64(len(rays)).tobytes()); f.write(rays.tobytes())
    res = subprocess.run([str(bvh_eval), str(path), "--check"], capture_output=True, text=True)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert "0 mismatches" in res.stdout

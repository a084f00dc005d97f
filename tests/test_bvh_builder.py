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


@pytest.mark.gpu
def test_gpu_traversal_counters_match_the_cpu_walk(bvh_eval, tmp_path, lib, oracle_mod):
    """the roofline's algorithmic bytes are 80 B x node steps + 48 B x triangle tests, counted by the kernels themselves
    (HRC_OPT_TRAVERSAL_COUNTERS). The same counts from an independent place: the rays of the same pass captured from the oracle, walked over the
    same BVH8 on the CPU. They agree to a fraction of a percent (not exactly: the two slab tests round differently in the last place, and a path
    whose last-place difference flips a decision contributes different rays)."""
    from heatray_b200 import capi, renderer, scenes
    from oracle import loader
    spec = scenes.sphere_grid(width=240, height=135, passes=4, grid=4, u_slices=32, v_slices=16, env_size=(128, 64), max_depth=6)
    tris = []
    for m in spec.meshes:
        M = np.asarray(m.transform, np.float32).reshape(4, 4)
        w = np.asarray(m.positions, np.float32).reshape(-1, 3) @ M[:3, :3].T + M[:3, 3]
        idx = np.asarray(m.indices).reshape(-1)
        if m.draw_mode == capi.DRAW_TRIANGLES:
            tri = idx.reshape(-1, 3)
        else:
            tri = np.stack([idx[:-2], idx[1:-1], idx[2:]], 1)
            tri = tri[(tri[:, 0] != tri[:, 1]) & (tri[:, 1] != tri[:, 2]) & (tri[:, 0] != tri[:, 2])]
        tris.append(w[tri].reshape(-1, 9))
    tris = np.concatenate(tris).astype(np.float32)
    ob = loader.backend()
    try:
        ro = renderer.Renderer(ob)
        ro.load_scene(spec)
        loader.capture_rays(ob, 4_000_000)
        ro.render_pass(sample_index=0)
        rays, kinds = loader.captured_rays(ob)
    finally:
        ob.close()
    rec = np.zeros(len(rays), RAY)
    rec["o"], rec["tmin"], rec["d"], rec["tmax"], rec["kind"] = rays["origin"], rays["tmin"], rays["direction"], rays["tmax"], kinds
    path = tmp_path / "pass.bin"
    with open(path, "wb") as f:
        f.write(np.uint64(len(tris)).tobytes()); f.write(tris.tobytes()); f.write(np.uint64(len(rec)).tobytes()); f.write(rec.tobytes())
    out = subprocess.run([str(bvh_eval), str(path)], capture_output=True, text=True, check=True).stdout
    tot = out.strip().splitlines()[-1].split()
    cpu = {tot[i]: float(tot[i + 1]) for i in range(1, len(tot), 2)}
    gb = capi.Backend(lib, "hrc", device=0, collect_stats=True)
    try:
        rg = renderer.Renderer(gb)
        rg.load_scene(spec)
        gb.reset_stats()
        st = rg.render_pass(sample_index=0)
    finally:
        gb.close()
    gpu = {"closest_nodes": st.nodesVisited, "closest_tris": st.trianglesTested, "any_nodes": st.nodesVisitedAny, "any_tris": st.trianglesTestedAny}
    assert int(st.raysPrimary + st.raysExtension + st.raysOcclusion) == pytest.approx(len(rays), rel=2e-3)
    print({k: (int(gpu[k]), int(cpu[k])) for k in gpu})
    for k in gpu:
        assert gpu[k] == pytest.approx(cpu[k], rel=1e-3), (k, gpu[k], cpu[k])          # measured on a B200: 730070 vs 730072 node steps, triangle tests and the occlusion counts equal

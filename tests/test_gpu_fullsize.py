"""BASELINE.json configs #2, #3 and #4 at their FULL sizes (262 K / 496 K triangles at 1920x1080; 10 M triangles at 3840x2160; 8 bounces) on the GPU, checked through
the oracle (image parity at the full frame size: two whole passes of #2 and #3 at 1920x1080, #4's full 10 M-triangle scene on a 960x540
frame (two passes) and at its full 3840x2160 (one pass)) and through size-independent properties:
  * closest-hit IDs and t,u,v bit-exact against the oracle's BVH on random rays plus the secondary / occlusion rays of an oracle
    pass over a low-resolution frame of the SAME full scene;
  * progressive contract: alpha == pass count on every pixel, sums finite and bounded by passes * rays-per-pixel * maxChannelValue;
  * linearity of accumulation: pass(s0) + pass(s1) rendered separately == both accumulated (sample-index sharding);
  * screen-band sharding: union of interleaved band renders == full frame, bit for bit;
  * ray budget: rays per pixel per pass <= 2 D + 2 (no alpha pass-throughs in these scenes).
"""
import numpy as np
import pytest

from heatray_b200 import capi, renderer, scenes

pytestmark = pytest.mark.gpu

FULL = {"config2_sphere_grid": scenes.sphere_grid, "config3_material_grid": scenes.material_grid, "config4_terrain": scenes.terrain}
SIZE = {"config2_sphere_grid": (1920, 1080), "config3_material_grid": (1920, 1080), "config4_terrain": (3840, 2160)}


@pytest.fixture(scope="module", params=list(FULL))
def full_scene(request, lib):
    spec = FULL[request.param]()
    assert (spec.width, spec.height) == SIZE[request.param] and spec.max_depth == 8
    if request.param == "config4_terrain":
        assert sum(len(m.indices) // 3 for m in spec.meshes) == 10018818
    b = capi.Backend(lib, "hrc", device=0, collect_stats=True)
    r = renderer.Renderer(b)
    r.load_scene(spec)
    yield spec, r
    b.close()


def test_full_scene_hits_bit_exact_vs_oracle(full_scene, oracle_mod):
    spec, rg = full_scene
    ob = oracle_mod.backend()
    try:
        ro = renderer.Renderer(ob)
        ro.load_scene(spec, width=96, height=54)                 # same full scene, small frame: the oracle's pass supplies real secondary rays
        oracle_mod.capture_rays(ob, 300000)
        ro.render_pass()
        rays, kinds = oracle_mod.captured_rays(ob)
        oracle_mod.capture_rays(ob, 0)
        assert (kinds == 1).sum() > 1000 and (kinds == 2).sum() > 1000
        rng = np.random.default_rng(42)
        n = 200000
        rnd = np.zeros(n, capi.RAY_DTYPE)
        rnd["origin"] = rng.uniform(-14, 14, (n, 3)).astype(np.float32)
        d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
        rnd["direction"] = d.astype(np.float32)
        rnd["tmin"], rnd["tmax"] = 0.0, np.inf
        allrays = np.concatenate([rays, rnd])
        ho, hg = ob.trace_rays(allrays), rg.b.trace_rays(allrays)
        assert np.array_equal(ho["primId"], hg["primId"])
        hit = ho["primId"] >= 0
        assert hit.sum() > 20000
        for k in ("t", "u", "v"):
            assert np.array_equal(ho[k][hit], hg[k][hit]), k
        occ = allrays[np.concatenate([kinds == 2, np.zeros(n, bool)])]
        assert np.array_equal(ob.trace_rays(occ, any_hit=True)["primId"] >= 0, rg.b.trace_rays(occ, any_hit=True)["primId"] >= 0)
    finally:
        ob.close()


def test_full_size_progressive_contract_linearity_and_budget(full_scene):
    spec, rg = full_scene
    b = rg.b
    b.clear()
    st0 = rg.render_pass(sample_index=0)
    a = rg.pixels().copy()
    b.clear()
    st1 = rg.render_pass(sample_index=1)
    bb = rg.pixels().copy()
    b.clear()
    rg.render_pass(sample_index=0)
    rg.render_pass(sample_index=1)
    both = rg.pixels().copy()
    assert (a[..., 3] == 1).all() and (bb[..., 3] == 1).all() and (both[..., 3] == 2).all()
    assert np.isfinite(both).all() and (both[..., :3] >= 0).all()
    per_pixel_cap = (2 * spec.max_depth + 2) * spec.max_channel_value          # every contribution is clamped to maxChannelValue (accumulator.rlsl:24)
    assert both[..., :3].max() <= 2 * per_pixel_cap
    assert not np.array_equal(a, bb)                                           # different sample indices give different samples
    # accumulating pass 1 onto pass 0 adds the same per-contribution values in a different association: fp32 rounding only
    assert np.allclose(both[..., :3], a[..., :3] + bb[..., :3], rtol=2e-6, atol=1e-6)
    for st in (st0, st1):
        rays = st.raysPrimary + st.raysExtension + st.raysOcclusion
        assert st.raysPrimary == spec.width * spec.height
        assert rays <= spec.width * spec.height * (2 * spec.max_depth + 2)
        assert rays > 2 * spec.width * spec.height                             # a real multi-bounce workload (the open terrain frame is half sky: 2.7 rays per pixel)
    # same sample index again: bit-identical (no atomics on pixels, deterministic queues per pixel)
    b.clear()
    rg.render_pass(sample_index=0)
    assert np.array_equal(rg.pixels(), a)


def test_full_size_band_sharding_is_bit_identical(full_scene):
    spec, rg = full_scene
    b = rg.b
    b.clear()
    rg.set_band(1, 0, 1)
    rg.render_pass(sample_index=3)
    full = rg.pixels().copy()
    total = np.zeros_like(full)
    try:
        for rank in range(2):
            b.clear()
            rg.set_band(16, rank, 2)
            rg.render_pass(sample_index=3)
            part = rg.pixels()
            rows = (np.arange(spec.height) // 16) % 2 == rank
            assert (part[~rows] == 0).all()
            total += part
    finally:
        rg.set_band(1, 0, 1)
    assert np.array_equal(total, full)


# ---- image parity against the oracle at BASELINE size ---------------------------------------------------------------------------
def _parity_report(pc, pg, rc, rg):
    """Per-pixel worst relative difference over rgb (denominator floored at 1e-2 of a channel value) and how the pixels distribute over it."""
    worst = (np.abs(pg[..., :3] - pc[..., :3]) / np.maximum(np.abs(pc[..., :3]), 1e-2)).max(-1)
    frac = {t: float((worst <= t).mean()) for t in (1e-6, 1e-5, 1e-4, 1e-3, 1e-2)}
    return worst, frac, float((rc != rg).mean())


@pytest.mark.parametrize("name,size,passes", [("config2_sphere_grid", (1920, 1080), 2), ("config3_material_grid", (1920, 1080), 2), ("config4_terrain", (960, 540), 2), ("config4_terrain", (3840, 2160), 1)])
def test_full_size_image_parity_vs_oracle(name, size, passes, lib, oracle_mod):
    """Whole passes of the full scenes on the CPU oracle (shading pinned bit for bit to the reference's shader text, tests/test_shader_pinning.py)
    and on the GPU. The sample counts are equal on every pixel; the colours differ only where the two libms' sin / cos / exp / atan2 differ in the
    last place and a sharp lobe, an environment texel edge or a Russian-roulette / lobe-selection threshold amplifies that (the kernels are
    compiled without FMA contraction, so every other operation rounds as on the CPU). Bars are set just under what is measured
    (profiles/r02_notes.md): >= 99.9 % of the pixels within 1e-3, >= 99.9 % within 1e-4, >= 99 % within 1e-5, <= 0.005 % beyond 1e-2, and the
    paths of >= 99.99 % of the pixels submit the same number of rays (a different count = a discrete decision flipped somewhere on the path)."""
    spec = FULL[name]()
    w, h = size
    out = []
    for mk in ("cpu", "gpu"):
        b = oracle_mod.backend() if mk == "cpu" else capi.Backend(lib, "hrc", device=0)
        try:
            r = renderer.Renderer(b)
            r.load_scene(spec, width=w, height=h)
            b.set_option(capi.OPT_PIXEL_RAY_COUNTS, 1)
            b.clear()
            for i in range(passes):
                r.render_pass(sample_index=i)
            out.append((r.pixels().copy(), b.ray_counts().copy()))
        finally:
            b.close()
    (pc, rc), (pg, rg) = out
    assert np.array_equal(pc[..., 3], pg[..., 3]) and (pg[..., 3] == passes).all()
    worst, frac, rays_differ = _parity_report(pc, pg, rc, rg)
    print(f"{name} {w}x{h} x{passes}: within 1e-6/1e-5/1e-4/1e-3/1e-2 = " + " ".join(f"{frac[t]:.5f}" for t in sorted(frac)) + f"; ray count differs on {rays_differ:.6f} of the pixels; rays {rc.sum()} vs {rg.sum()}")
    assert frac[1e-3] >= 0.999 and frac[1e-4] >= 0.999 and frac[1e-5] >= 0.99 and frac[1e-2] >= 0.99995
    assert rays_differ <= 1e-4
    assert abs(int(rc.sum()) - int(rg.sum())) <= 1e-5 * rc.sum()
    mse = float(np.mean((pg[..., :3] - pc[..., :3]) ** 2)) / passes ** 2
    assert 10 * np.log10(spec.max_channel_value ** 2 / max(mse, 1e-30)) >= 70.0          # PSNR of the mean image against the clamp value

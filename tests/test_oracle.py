"""Pins the CPU oracle: restated RLSL helpers against values produced by the reference's own shader text compiled as C++
(tests/golden/rlsl_helpers.json), its BVH against its brute-force intersector, and analytic lighting checks."""
import ctypes as C
import json
import math
from pathlib import Path

import numpy as np
import pytest

from heatray_b200 import capi, renderer, scenes

GOLD = Path(__file__).parent / "golden"


def test_rlsl_helpers_vs_reference_shader_text(oracle_mod):
    """orthonormalFrame, cosineWeightedSample, sampleVisibleGGX, D_GGX, G1/G2_Smith_GGX, F_Fresnel, F_Schlick, random():
    utility.rlsl + brdfs.rlsl compiled through glm vs the oracle's restatement. Same libm, same op order => <= 2 ulp."""
    o = oracle_mod.load()
    cases = json.loads((GOLD / "rlsl_helpers.json").read_text())
    worst = 0
    for c in cases:
        a = np.asarray(c["in"], np.uint32).view(np.float32).copy()
        want = np.asarray(c["out"], np.uint32).view(np.float32)
        got = np.zeros(18, np.float32)
        o.orcProbeHelpers(a.ctypes.data_as(C.c_void_p), got.ctypes.data_as(C.c_void_p))
        # random() (index 17) is a hash of sin(): compared exactly below where it decides sequence ids
        ulps = np.abs(got[:17].view(np.int32).astype(np.int64) - want[:17].view(np.int32).astype(np.int64))
        near_zero = np.abs(want[:17]) < 1e-6
        assert (ulps[~near_zero] <= 4).all(), (ulps, got, want)
        assert np.allclose(got[:17][near_zero], want[:17][near_zero], atol=1e-6)
        worst = max(worst, int(ulps[~near_zero].max()))
        assert abs(got[17] - want[17]) < 1e-3 or abs(abs(got[17] - want[17]) - 1) < 1e-3
    # SURVEY.md Appendix C.3, first case
    a = np.asarray(cases[0]["in"], np.uint32).view(np.float32).copy()
    got = np.zeros(18, np.float32)
    o.orcProbeHelpers(a.ctypes.data_as(C.c_void_p), got.ctypes.data_as(C.c_void_p))
    assert np.allclose(got[0:6], [0.93885857, -0.34188175, 0.040760949, 0.040760949, 0.22792116, 0.972826], atol=1e-6)
    assert np.allclose(got[6:9], [-0.40450838, 0.86602539, -0.29389277], atol=1e-6)
    assert np.allclose(got[12:17], [0.34359634, 0.92917675, 0.91453815, 0.07509616, 0.20134717], atol=1e-6)
    assert got[17] == np.float32(0.00390625)


def test_live_rlsl_reference_if_present(oracle_mod):
    ref = oracle_mod.ref_lib("rlsl")
    if ref is None:
        pytest.skip("oracle/_ref not built")
    o = oracle_mod.load()
    rng = np.random.default_rng(5)
    for _ in range(200):
        a = rng.uniform(0.05, 1.0, 18).astype(np.float32)
        g1, g2 = np.zeros(18, np.float32), np.zeros(18, np.float32)
        o.orcProbeHelpers(a.ctypes.data_as(C.c_void_p), g1.ctypes.data_as(C.c_void_p))
        ref.ref_rlsl_probe(a.ctypes.data_as(C.c_void_p), g2.ctypes.data_as(C.c_void_p))
        assert np.allclose(g1[:17], g2[:17], rtol=2e-6, atol=1e-7)


def _random_rays(n, extent, seed=3):
    rng = np.random.default_rng(seed)
    rays = np.zeros(n, capi.RAY_DTYPE)
    rays["origin"] = rng.uniform(-extent, extent, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays["direction"] = d.astype(np.float32)
    rays["tmax"] = np.inf
    return rays


def test_oracle_bvh_equals_brute_force(oracle_mod, oracle_backend):
    spec = scenes.sphere_grid(width=16, height=16, passes=1, grid=2, u_slices=16, v_slices=8, env_size=(8, 4))
    r = renderer.Renderer(oracle_backend)
    r.load_scene(spec)
    rays = _random_rays(20000, 5.0)
    h1 = oracle_backend.trace_rays(rays)
    a1 = oracle_backend.trace_rays(rays, any_hit=True)
    oracle_mod.load().orcSetBruteForce(oracle_backend.ctx, 1)
    h2 = oracle_backend.trace_rays(rays)
    a2 = oracle_backend.trace_rays(rays, any_hit=True)
    assert np.array_equal(h1, h2) and np.array_equal(a1, a2)
    assert (h1["primId"] >= 0).sum() > 2000
    assert np.array_equal(a1["primId"] == 0, h1["primId"] >= 0)


def test_strip_equals_triangles_and_winding(oracle_backend):
    """GL-style strip expansion with the odd-triangle swap (policy A-9) and front face by determinant (Mesh.cpp:86-91)."""
    b = oracle_backend
    b.resize(4, 4)
    mat = b.create_material(capi.MATERIAL_PBR, capi.MAT_DOUBLE_SIDED, capi.PbrParams())
    p, n, i = scenes.plane_mesh(2.0, 2.0)
    b.create_mesh(p, n, i, mat, draw_mode=capi.DRAW_TRIANGLE_STRIP)
    b.commit()
    rays = np.zeros(3, capi.RAY_DTYPE)
    rays["origin"] = [[-0.5, 1, 0.5], [0.5, 1, -0.5], [0.5, -1, -0.5]]
    rays["direction"] = [[0, -1, 0], [0, -1, 0], [0, 1, 0]]
    rays["tmax"] = np.inf
    h = b.trace_rays(rays)
    assert (h["primId"] & 0x3fffffff).tolist() == [0, 1, 1] and np.allclose(h["t"], 1.0)
    assert [(x & capi.HIT_BACKFACE) != 0 for x in h["primId"].tolist()] == [False, False, True]
    b.clear_scene()
    mat = b.create_material(capi.MATERIAL_PBR, capi.MAT_DOUBLE_SIDED, capi.PbrParams())
    b.create_mesh(p, n, np.array([0, 1, 3, 3, 1, 2], np.uint32), mat)
    b.commit()
    h2 = b.trace_rays(rays)
    assert np.array_equal(h, h2)
    # mirrored transform: det < 0 => CW front face; the same side must stay "front" for the shaders
    b.clear_scene()
    mat = b.create_material(capi.MATERIAL_PBR, capi.MAT_DOUBLE_SIDED, capi.PbrParams())
    b.create_mesh(p, n, i, mat, transform=scenes.scale(-1, 1, 1), draw_mode=capi.DRAW_TRIANGLE_STRIP)
    b.commit()
    h3 = b.trace_rays(rays)
    assert [(x & capi.HIT_BACKFACE) != 0 for x in h3["primId"].tolist()] == [True, True, False]   # geometric winding flipped ...


def test_set_mesh_transform_equals_creating_with_that_transform(oracle_backend):
    """Scene::applyTransform (Scene.cpp:38-49): rewriting a mesh's worldFromEntity and re-committing == building it there."""
    b = oracle_backend
    b.resize(4, 4)
    p, n, i = scenes.sphere_mesh(12, 8, 1.0)
    T = scenes.translate(1.5, -0.5, 2.0) @ scenes.scale(2.0, 2.0, 2.0)
    mat = b.create_material(capi.MATERIAL_PBR, 0, capi.PbrParams())
    m0 = b.create_mesh(p, n, i, mat)
    b.commit()
    rays = _random_rays(4000, 4.0)
    before = b.trace_rays(rays)
    b.set_mesh_transform(m0, T)
    b.commit()
    moved = b.trace_rays(rays)
    assert not np.array_equal(before["primId"], moved["primId"])
    b.clear_scene()
    mat = b.create_material(capi.MATERIAL_PBR, 0, capi.PbrParams())
    b.create_mesh(p, n, i, mat, transform=T)
    b.commit()
    assert np.array_equal(b.trace_rays(rays), moved)
    with pytest.raises(capi.HrcError):
        b.set_mesh_transform(77, T)


def _flat_scene(base=(0.8, 0.8, 0.8), **kw):
    p, n, i = scenes.plane_mesh(200.0, 200.0)
    mats = [scenes._pbr(base, rough=1.0, metal=0.0, f0=0.0)]
    meshes = [scenes.MeshSpec(p, n, i, 0, draw_mode=capi.DRAW_TRIANGLE_STRIP)]
    cam = dict(distance=10.0, phi=0.0, theta=math.pi / 2 - 1e-3, target=(0.0, 0.0, 0.0), focal_length=200.0, fstop=None, focus_distance=10.0)
    args = dict(env=dict(enabled=False, solid=None, map=None, ev=0.0, theta=0.0), camera=cam)
    args.update(kw)
    return scenes.SceneSpec("flat", 16, 16, 4, 0, meshes, mats, **args)


def test_lambert_plane_under_directional_light(oracle_backend):
    """Analytic check: depth-0 radiance of a Lambert plane = Cdiff/pi * N.L * colour, per sample."""
    th = math.pi / 3
    spec = _flat_scene(dir_lights=[dict(color=(1.0, 0.5, 0.25), illuminance=683.0 * 2.0, theta=th, phi=0.0)])
    r = renderer.Renderer(oracle_backend)
    r.load_scene(spec)
    for _ in range(spec.passes):
        r.render_pass()
    img = renderer.resolve(r.pixels())
    L = np.array([0.0, math.sin(th), math.cos(th)])       # direction to the light for (theta, phi=0)
    d = capi.DirLight()
    capi.load_library().hrhPackDirectionalLight((C.c_float * 3)(1, 0.5, 0.25), 683.0 * 2.0, th, 0.0, C.byref(d))
    ndl = float(d.direction[1])
    want = 0.8 / math.pi * ndl * np.array([2.0, 1.0, 0.5])
    assert np.allclose(img.reshape(-1, 3), want, rtol=2e-5), (img[0, 0], want)


def test_point_light_inverse_square(oracle_backend):
    """Point light: colour = c * (I/683) * 4pi, radiance = Cdiff/pi * N.L * colour / d^2 (pointLight.rlsl:20-29)."""
    hgt = 3.0
    spec = _flat_scene(point_lights=[dict(color=(1.0, 1.0, 1.0), intensity=683.0 / (4 * math.pi), position=(0.0, hgt, 0.0))])
    r = renderer.Renderer(oracle_backend)
    r.load_scene(spec)
    r.render_pass()
    img = renderer.resolve(r.pixels())
    centre = img[8, 8]
    want = 0.8 / math.pi * 1.0 * 1.0 / (hgt * hgt)
    assert np.allclose(centre, want, rtol=2e-3)


def test_furnace_sphere_array_stays_below_environment(oracle_backend):
    """The reference's own energy check (HeatrayRenderer.cpp:241-286): white spheres in a 0.8 environment must not
    emit more than 0.8 (allowing Monte-Carlo noise on the mean)."""
    sp, sn, si = scenes.sphere_mesh(24, 12, 0.5)
    mats, meshes = [], []
    for k, (rough, metal) in enumerate([(0.0, 0.0), (0.5, 0.0), (0.9, 0.0), (0.0, 1.0), (0.5, 1.0), (0.9, 1.0)]):
        mats.append(scenes._pbr((1.0, 1.0, 1.0), rough=rough, metal=metal, f0=0.0))
        meshes.append(scenes.MeshSpec(sp, sn, si, k, scenes.translate((k % 3 - 1) * 1.2, (k // 3) * 1.5 - 0.75, 0.0)))
    spec = scenes.SceneSpec("furnace", 96, 64, 24, 6, meshes, mats, env=dict(enabled=True, solid=(0.8, 0.8, 0.8), map=None, ev=0.0, theta=0.0),
                            camera=dict(distance=9.0, phi=0.0, theta=0.0, target=(0, 0, 0), focal_length=50.0, fstop=None, focus_distance=9.0))
    r = renderer.Renderer(oracle_backend)
    r.load_scene(spec)
    for _ in range(spec.passes):
        r.render_pass()
    img = renderer.resolve(r.pixels())
    assert np.isfinite(img).all()
    assert img.mean() <= 0.8 * 1.02
    assert np.allclose(img[0, 0], 0.8, rtol=1e-6)          # background pixels see the environment directly


def test_progressive_contract(oracle_backend):
    spec = scenes.cornell_box(width=24, height=24, passes=3)
    r = renderer.Renderer(oracle_backend)
    r.load_scene(spec)
    st = r.render_pass()
    assert st.raysPrimary == 24 * 24 and st.rays <= 24 * 24 * (2 * spec.max_depth + 2)
    p1 = r.pixels()
    r.render_pass()
    p2 = r.pixels()
    assert (p1[..., 3] == 1).all() and (p2[..., 3] == 2).all() and (p2[..., :3] >= p1[..., :3]).all()
    assert p1[..., :3].max() <= 3 * math.pi * (spec.max_depth + 2)      # per-contribution clamp (quirk B-Q14)


def _fence_scene(alpha, masked=True):
    """Ground plane under a directional light with a horizontal textured sheet in between."""
    import math
    from heatray_b200 import scenes as sc
    tex = np.zeros((4, 4, 4), np.float32); tex[..., :3] = 0.5; tex[..., 3] = alpha
    mats = [sc._pbr((0.8, 0.8, 0.8), rough=1.0, f0=0.0), sc._pbr((1, 1, 1), rough=1.0, f0=0.0, alpha_mask=masked, textures=dict(baseColorTexture=tex))]
    p, n, i, uv = sc.plane_mesh(1.0, 1.0, with_uv=True)
    meshes = [sc.MeshSpec(p, n, i, 0, sc.scale(20, 1, 20), capi.DRAW_TRIANGLE_STRIP),
              sc.MeshSpec(p, n, i, 1, sc.translate(0, 5.0, 0) @ sc.scale(40, 1, 40), capi.DRAW_TRIANGLE_STRIP, uvs=uv)]
    return sc.SceneSpec("fence", 16, 16, 1, 0, meshes, mats, env=dict(enabled=False, solid=None, map=None, ev=0.0, theta=0.0),
                        dir_lights=[dict(color=(1, 1, 1), illuminance=683.0, theta=1.45, phi=0.0)],
                        camera=dict(distance=3.0, phi=0.0, theta=0.9, target=(0, 0, 0), focal_length=50.0, fstop=None, focus_distance=3.0))


def test_alpha_masked_surface_lets_occlusion_rays_through(oracle_backend):
    """physicallyBased.rlsl:70-90 + Mesh.cpp:95-100 (policy A-5): alpha < 1 passes shadow rays on (one extra segment each,
    light distance bookkeeping in extraT), alpha = 1 shadows, and a non-masked sheet is an ordinary occluder."""
    from heatray_b200 import renderer
    lit = {}
    for name, spec in (("open", _fence_scene(0.0)), ("closed", _fence_scene(1.0)), ("solid", _fence_scene(0.0, masked=False))):
        r = renderer.Renderer(oracle_backend)
        r.load_scene(spec)
        st = r.render_pass()
        lit[name] = (r.pixels()[..., :3].sum(), st.raysOcclusion, st.raysPrimary)
    n = 16 * 16
    assert lit["closed"][0] == 0 and lit["solid"][0] == 0 and lit["open"][0] > 0
    assert lit["open"][1] == 2 * n and lit["closed"][1] == n and lit["solid"][1] == n     # the pass-through re-emit is one more occlusion segment


def test_textured_base_colour_is_bilinear(oracle_backend):
    """showBaseColor on a textured quad reproduces level-0 bilinear filtering with REPEAT (policy A-12)."""
    import math
    from heatray_b200 import renderer, scenes as sc
    rng = np.random.default_rng(5)
    tex = rng.uniform(0, 1, (4, 4, 4)).astype(np.float32)
    mats = [sc._pbr((1, 1, 1), textures=dict(baseColorTexture=tex))]
    p, n, i, uv = sc.plane_mesh(2.0, 2.0, with_uv=True)
    spec = sc.SceneSpec("quad", 8, 8, 1, 0, [sc.MeshSpec(p, n, i, 0, sc.rotate("x", math.pi / 2), capi.DRAW_TRIANGLE_STRIP, uvs=uv)], mats,
                        env=dict(enabled=False, solid=None, map=None, ev=0.0, theta=0.0),
                        camera=dict(distance=4.0, phi=0.0, theta=0.0, target=(0, 0, 0), focal_length=50.0, fstop=None, focus_distance=4.0))
    r = renderer.Renderer(oracle_backend)
    r.load_scene(spec)
    r.reset(debug="showUVs"); r.render_pass(); uvs = r.pixels()
    r.reset(debug="showBaseColor"); r.render_pass(); col = r.pixels()
    hit = uvs[..., 3] == 2      # frame shader +1, visualiser +1
    assert hit.sum() >= 16
    def bilinear(u, v):
        x, y = u * 4 - 0.5, v * 4 - 0.5
        x0, y0 = int(np.floor(x)), int(np.floor(y)); fx, fy = np.float32(x - x0), np.float32(y - y0)
        t = lambda xx, yy: tex[yy % 4, xx % 4, :3]
        return (t(x0, y0) * (1 - fx) + t(x0 + 1, y0) * fx) * (1 - fy) + (t(x0, y0 + 1) * (1 - fx) + t(x0 + 1, y0 + 1) * fx) * fy
    for y in range(8):
        for x in range(8):
            if not hit[y, x]:
                continue
            assert np.allclose(col[y, x, :3], bilinear(uvs[y, x, 0], uvs[y, x, 1]), atol=2e-6)


def test_interactive_block_mode_covers_each_pixel_once(oracle_backend):
    """perspective.rlsl:42-57 + PassGenerator.cpp:372-381: nine interactive passes sample every pixel exactly once, all at
    sample index 0, so their sum equals one ordinary pass."""
    from heatray_b200 import renderer, scenes
    spec = scenes.cornell_box(width=40, height=31, passes=2)
    r = renderer.Renderer(oracle_backend)
    r.load_scene(spec)
    for k in range(9):
        st = r.render_pass(interactive=True)
        a = r.pixels()[..., 3]
        assert a.max() == 1 and st.raysPrimary == a.sum() - (0 if k == 0 else prev)
        prev = a.sum()
    inter = r.pixels()
    assert (inter[..., 3] == 1).all() and r.sample_index == 1 and r.block_sample == [0, 0]
    r.reset()
    r.render_pass()
    assert np.array_equal(r.pixels(), inter)


def test_deferred_passes_commit_oldest_first(oracle_backend):
    """The oracle's mirror of hrcEnqueuePassDeferred / Commit / Discard (the contract the CUDA library is tested against on the GPU): up to
    three pending, invisible until committed, committed oldest first, a discard drops all pending."""
    spec = scenes.cornell_box(width=32, height=32, passes=4)
    b = oracle_backend
    r = renderer.Renderer(b)
    r.load_scene(spec)
    singles = []
    for i in range(4):
        b.clear()
        r.render_pass(sample_index=i)
        singles.append(r.pixels().copy())
    b.clear()
    r.render_pass(sample_index=0)
    for i in (1, 2, 3):
        r.frame.sampleIndex = i
        b.enqueue_pass_deferred(r.frame)
    with pytest.raises(capi.HrcError):
        b.enqueue_pass_deferred(r.frame)
    assert np.array_equal(r.pixels(), singles[0])
    b.commit_deferred()
    assert np.array_equal(r.pixels(), singles[0] + singles[1])
    b.discard_deferred()
    with pytest.raises(capi.HrcError):
        b.commit_deferred()
    assert np.array_equal(r.pixels(), singles[0] + singles[1])

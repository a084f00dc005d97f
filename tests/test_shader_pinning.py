"""Pins the oracle's shading to the reference's shader TEXT.

The reference's RLSL programs — perspective.rlsl, physicallyBased.rlsl, glass.rlsl (one build per #define set), environmentLight /
directionalLight / pointLight / spotLight.rlsl and everything they include — are compiled as C++ through glm from the files where
they lie under /root/reference (oracle/ref_probe/rlsl_to_cpp.py + rlsl_shim.h + ref_shaders.cpp -> oracle/_ref/libref_shaders.so).
The oracle's restatement (oracle/oracle.cpp: shadePBR, shadeGlass, the sample functions of microfacet.rlsl, computeLightSample,
shadeMiss, generatePrimaryRay, performAccumulate) runs the same randomised shader invocations through orcProbeShade / orcProbeMiss /
orcProbeRaygen. Bar: every discrete decision equal (rays emitted, their kind, miss primitive, depth, sequence attributes) and every
float (direction, weight, maxT, extraT, accumulated value) within MAX_ULP units in the last place.
  * the golden tests run everywhere against tests/golden/shader_ref.npz (written by oracle/ref_probe/make_shader_golden.py);
  * the live tests run where oracle/_ref/libref_shaders.so exists and use >= 1e5 hits per #define set."""
import ctypes as C
from pathlib import Path

import numpy as np
import pytest

from heatray_b200 import capi, renderer, scenes
from oracle import shader_probe as sp

GOLD = Path(__file__).parent / "golden" / "shader_ref.npz"
MAX_ULP = 2


def _oracle_shade(oracle_mod, mtype, flags, params, world, cases, si):
    b = oracle_mod.backend()
    try:
        orc = sp.OracleSide(b)
        orc.set_world(world)
        cases = cases.copy()
        cases["material"] = orc.material(mtype, flags, params)
        return orc.shade(si, cases)
    finally:
        b.close()


def _assert_same(ref, orc, what, compare_origin=False):
    ok, worst = sp.compare(ref, orc, MAX_ULP, compare_origin)
    if not ok.all():
        i = int(np.flatnonzero(~ok)[0])
        raise AssertionError(f"{what}: {int((~ok).sum())} of {len(ok)} invocations differ; first #{i}\n  reference {ref[i]}\n  oracle    {orc[i]}")
    return int(worst.max()) if len(worst) else 0


def test_material_shaders_match_reference_golden(oracle_mod, lib):
    """physicallyBased.rlsl:55-331, glass.rlsl:138-280, microfacet.rlsl:25-220, lightSampling.rlsl:11-161, accumulator.rlsl:12-28."""
    gold = np.load(GOLD)
    seed, n = (int(v) for v in gold["meta"])
    groups = 0
    worst = 0
    for i, (mtype, flags, params, world, cases, si) in enumerate(sp.material_suite(seed, n)):
        assert sp.input_digest(cases, world.seq, bytes(params)) == str(gold["digests"][i]), \
            "the case generator drifted from the one that wrote tests/golden/shader_ref.npz: rerun oracle/ref_probe/make_shader_golden.py"
        orc = _oracle_shade(oracle_mod, mtype, flags, params, world, cases, si)
        worst = max(worst, _assert_same(gold[f"shade_{i}"], orc, f"material type {mtype} flags {flags:#05x} group {i}"))
        groups += 1
    assert groups == (len(sp.PBR_VARIANTS) + len(sp.GLASS_VARIANTS)) * len(sp.LIGHT_SETS)
    print(f"{groups} groups x {n} shader invocations, worst float difference {worst} ulp")


def test_light_shaders_match_reference_golden(oracle_mod, lib):
    """environmentLight.rlsl:19-34, directionalLight.rlsl:20-26, pointLight.rlsl:20-29, spotLight.rlsl:20-36 incl. the NaN / Inf visualiser."""
    gold = np.load(GOLD)
    seed = int(gold["meta"][0]) + 1
    base = (len(sp.PBR_VARIANTS) + len(sp.GLASS_VARIANTS)) * len(sp.LIGHT_SETS)
    for i, (world, rays) in enumerate(sp.miss_suite(seed, 512)):
        assert sp.input_digest(rays) == str(gold["digests"][base + i])
        b = oracle_mod.backend()
        try:
            orc = sp.OracleSide(b); orc.set_world(world)
            got = orc.miss(rays)
        finally:
            b.close()
        want = gold[f"miss_{i}"]
        assert (sp.ulp_diff(want["accumSum"], got["accumSum"]) <= MAX_ULP).all()
        if world.globals_.enableAccumulatorVisualizer:            # accumulator.rlsl:16-21: flagged values come out as 100
            assert (want["accumSum"][::7][want["accumCount"][::7] > 0][:, :3] == 100.0).any()


@pytest.fixture(scope="module")
def ref_side(oracle_mod):
    lib = oracle_mod.ref_lib("shaders")
    if lib is None:
        pytest.skip("oracle/_ref/libref_shaders.so not built (needs /root/reference)")
    return sp.RefSide(lib)


def test_compiled_variants_cover_the_suite(ref_side):
    assert ref_side.variants(sp.PBR) == sp.PBR_VARIANTS and ref_side.variants(sp.GLASS) == sp.GLASS_VARIANTS


def test_material_shaders_live_1e5_per_define_set(oracle_mod, lib, ref_side):
    """>= 1e5 randomised hits per compiled #define set: depths 0..11 (Russian roulette, the mod-16 sequence wrap), front and back faces,
    all four light types alone and together, environment on / off, every debug visualiser, alpha-masked shadow rays."""
    per_world = 14300                                             # x 7 light sets = 100,100 per #define set
    total = 0
    worst = 0
    exact = 0
    for mtype, flags, params, world, cases, si in sp.material_suite(77, per_world):
        ref_side.set_world(world)
        ref = ref_side.shade(mtype, flags, params, si, cases)
        orc = _oracle_shade(oracle_mod, mtype, flags, params, world, cases, si)
        worst = max(worst, _assert_same(ref, orc, f"material type {mtype} flags {flags:#05x}"))
        # what the shader leaves the hit point with is rl_IntersectionPoint; the oracle then applies the ray-epsilon policy A-8
        live = ref["rayCount"] > 0
        assert np.array_equal(ref["rays"]["origin"][live, 0], cases["P"][live])
        exact += int(sp.compare(ref, orc, 0)[0].sum())
        total += len(cases)
    sets = len(sp.PBR_VARIANTS) + len(sp.GLASS_VARIANTS)
    assert total >= 100000 * sets
    print(f"{total} shader invocations over {sets} #define sets: worst {worst} ulp, {exact / total:.6f} bit-equal")


def test_light_shaders_live(oracle_mod, lib, ref_side):
    for world, rays in sp.miss_suite(5, 50000):
        ref_side.set_world(world)
        ref = ref_side.miss(rays)
        b = oracle_mod.backend()
        try:
            orc = sp.OracleSide(b); orc.set_world(world)
            got = orc.miss(rays)
        finally:
            b.close()
        assert (sp.ulp_diff(ref["accumSum"], got["accumSum"]) <= MAX_ULP).all()


@pytest.mark.parametrize("w,h,dof,interactive", [(96, 64, False, False), (64, 64, True, False), (50, 70, True, True), (64, 48, False, True)])
def test_frame_shader_live(oracle_mod, lib, ref_side, w, h, dof, interactive):
    """perspective.rlsl:39-93 on every pixel: which pixels an interactive pass samples, sequenceID (the sin hash, quirk B-Q2 — the host
    table must agree with the shader's own evaluation), the height-strided offset lookup (B-Q1), jitter, thin-lens origin / direction."""
    spec = scenes.cornell_box(width=w, height=h, passes=16)
    if dof:
        spec.camera["fstop"] = 1.4; spec.bokeh = capi.BOKEH_HEXAGON
    b = oracle_mod.backend()
    try:
        r = renderer.Renderer(b); r.load_scene(spec)
        b.set_environment(True, -1, 1.0, 0.0)
        t = r.tables
        L = ref_side.lib
        L.ref_set_globals(C.byref(r.globals))
        env = capi.EnvLight(1, -1, 1.0, 0.0)
        L.ref_set_lights(C.byref(env), 0, None, 0, None, 0, None)
        L.ref_set_sequences(t["xy"].shape[0], t["xy"].shape[1], sp._p(t["xy"]), sp._p(t["aperture"]))
        L.ref_set_pixel_offsets(sp._p(t["offsets"]), t["offsets"].size)
        bc = np.ascontiguousarray(t["block_coords"])
        L.ref_set_interactive_blocks(bc.shape[1], bc.shape[0], sp._p(bc))
        ys, xs = np.mgrid[0:h, 0:w]
        xy = np.stack([xs.ravel(), ys.ravel()], 1).astype(np.int32)
        orc = sp.OracleSide(b)
        for si in (0, 5, 15):
            for bs in ((0, 0), (1, 2), (2, 1)) if interactive else ((0, 0),):
                f = r.frame
                f.sampleIndex = si; f.interactiveMode = int(interactive); f.currentBlockPixelSample = (C.c_int32 * 2)(*bs)
                ref = sp.normalize_emitted(ref_side.raygen(f, w, h, xy))            # OpenRL normalises emitted directions (policy A-11)
                got = orc.raygen(f, xy)
                _assert_same(ref, got, f"frame shader {w}x{h} sample {si} block {bs}", compare_origin=True)
                if interactive:
                    assert 0 < ref["rayCount"].sum() < w * h / 6
    finally:
        b.close()


def test_deviation_d1_second_extension_ray_without_environment(oracle_mod, lib, ref_side):
    """DESIGN.md deviation D-1. microfacet.rlsl:64-65,93-97 (quirk B-Q11): a punctual light picked with a light sample of EXACTLY 0.0
    falls into the 'environment' branch; with no environment light its miss primitive is rl_NullPrimitive, so the shader emits a second
    NON-occlusion ray. A sample value of 0.0 needs a zero 32-bit Owen-scrambled integer (2^-32 per table entry); the wavefront carries one
    extension ray per hit and drops this one. This test pins that the dropped ray is the ONLY difference."""
    rng = np.random.default_rng(3)
    w = sp.random_world(rng, "dpps", 8)
    w.seq[:, :, 1] = 0.0                                            # every light pick is exactly 0
    tex0 = sp.add_material_textures(rng, w)
    params = sp.random_pbr_params(rng, sp.DS, tex0)
    cases = sp.random_cases(rng, 4000, 0, w)
    ref_side.set_world(w)
    ref = ref_side.shade(sp.PBR, sp.DS, params, 3, cases)
    orc = _oracle_shade(oracle_mod, sp.PBR, sp.DS, params, w, cases, 3)
    extra = ref["rayCount"] - orc["rayCount"]
    assert set(np.unique(extra)) <= {0, 1} and (extra == 1).sum() > 100
    first = ref["rays"][:, 0]
    assert (first["occlusionTest"][extra == 1] == 0).all() and (first["missKind"][extra == 1] == 0).all()
    trimmed = ref.copy()
    sel = extra == 1
    trimmed["rays"][sel, 0] = ref["rays"][sel, 1]; trimmed["rays"][sel, 1] = 0; trimmed["rayCount"][sel] -= 1
    _assert_same(trimmed, orc, "all other outputs")


# ---- vertex.rlsl: the geometry flattening ----------------------------------------------------------------------------------------------
VERTEX_GOLD = Path(__file__).parent / "golden" / "vertex_ref.npz"
SEED_VERTEX = 20261105 + 2


def _same_vertices(ref, got, what):
    for f in ref.dtype.names:
        assert np.array_equal(ref[f].view(np.uint32), got[f].view(np.uint32)), f"{what}: {f} differs"


def test_vertex_shader_matches_reference_golden(oracle_mod):
    """vertex.rlsl:25-43 (rl_Position = worldFromEntity * vec4(p, 1); normal / tangent / bitangent = mat3(worldFromEntity) * v, quirk B-Q7: not the
    inverse transpose) against the oracle's xformPoint / xformNormal, which commitScene applies to every vertex. Bar: bit for bit."""
    gold = np.load(VERTEX_GOLD)
    cases = sp.vertex_cases(SEED_VERTEX, 400)
    assert sp.input_digest(cases) == str(gold["digest"]), "vertex_cases drifted: rerun oracle/ref_probe/make_shader_golden.py --vertex"
    o = oracle_mod.load()
    for f in sp.VERTEX_VARIANTS:
        _same_vertices(gold[f"vertex_{f}"], sp.oracle_vertex(o, f, cases), f"#define set {f}")


def test_vertex_shader_live(oracle_mod, ref_side):
    cases = sp.vertex_cases(9, 200000)
    o = oracle_mod.load()
    for f in sp.VERTEX_VARIANTS:
        ref = sp.ref_vertex(ref_side.lib, f, cases)
        _same_vertices(ref, sp.oracle_vertex(o, f, cases), f"#define set {f}")
        assert (ref["position"][:, 3] == 1.0).all()
        if not f & 2:
            assert not ref["tangent"].any() and not ref["bitangent"].any()

"""Session files (reference: Session/Session.h:21-78 variable table, Session.cpp:88-214 parse/write, HeatrayRenderer.cpp:503-690 mapping).
The golden XML files under tests/golden/ were written by the reference's own Session class + tinyxml2 compiled in place
(oracle/ref_probe/ref_session.cpp, make_session_golden.py); session_ref.json holds what went in and what the reference read back."""
import ctypes as C
import filecmp
import json
import math
from pathlib import Path

import numpy as np
import pytest

from heatray_b200 import capi

GOLDEN = Path(__file__).parent / "golden"
META = json.loads((GOLDEN / "session_ref.json").read_text())
KINDS = META["kinds"]
REF_SO = Path(__file__).resolve().parents[1] / "oracle/_ref/libref_session.so"


def numbers_and_strings(session):
    flat = session.flat()
    assert len(flat) == 50 == len(KINDS)
    nums = [0.0 if k == "s" else v for k, v in zip(KINDS, flat)]
    strs = [v for k, v in zip(KINDS, flat) if k == "s"]
    return nums, strs


def fill(session, values, scene, env):
    i = 0
    for name, _ in session._fields_:
        cur = getattr(session, name)
        if isinstance(cur, bytes):
            setattr(session, name, (scene if name == "scene" else env).encode())
            i += 1
        elif hasattr(cur, "__len__"):
            for j in range(len(cur)):
                cur[j] = values[i]
                i += 1
        else:
            setattr(session, name, int(values[i]) if KINDS[i] in "bu" else values[i])
            i += 1
    assert i == 50


def test_defaults_match_the_reference_table():
    s = capi.session_defaults()
    assert (s.interactiveMode, s.maxRenderPasses, s.maxChannelValue, s.maxRayDepth) == (0, 32, 1.0, 10)
    assert s.scene == b"Multi-Material" and s.environmentMap == b"studio.hdr" and s.environmentBuiltIn == 1
    assert (s.cameraAspectRatio, s.cameraFocusDistance, s.cameraFocalLength) == (-1.0, 1.0, 50.0)
    assert s.orbitDistance == 10.0 and s.distanceScale == 1.0 and s.scale == 1.0


@pytest.mark.parametrize("case", META["cases"], ids=lambda c: c["file"])
def test_reads_reference_written_files_like_the_reference(case):
    s, missing = capi.read_session(GOLDEN / case["file"])
    assert missing == 0
    nums, strs = numbers_and_strings(s)
    assert strs == [case["scene_back"], case["env_back"]]
    for k, ours, ref in zip(KINDS, nums, case["read_back"]):
        assert ours == ref or (math.isnan(ours) and math.isnan(ref)), (k, ours, ref)


@pytest.mark.parametrize("case", META["cases"], ids=lambda c: c["file"])
def test_writes_byte_identical_files(case, tmp_path):
    s = capi.session_defaults()
    fill(s, case["written"], case["scene"], case["env"])
    out = tmp_path / "ours.xml"
    capi.write_session(out, s)
    assert out.read_bytes() == (GOLDEN / case["file"]).read_bytes()


def test_missing_variables_keep_their_value_and_are_counted(tmp_path):
    p = tmp_path / "partial.xml"
    p.write_text('<?xml version="1.0"?>\n<!-- hand edited -->\n<HeatraySession>\n  <RenderOptions>\n    <MaxRayDepth value="3"/>\n'
                 "    <Scene value='a&lt;b&#65;&#x42;'/>\n    <Unknown value=\"7\"/>\n  </RenderOptions>\n"
                 '  <PostProcessing><TonemapEnable value="1"/><Exposure value="1.5e0" other="x"/></PostProcessing>\n</HeatraySession>\n')
    s, missing = capi.read_session(p)
    assert missing == 50 - 4
    assert s.maxRayDepth == 3 and s.scene == b"a<bAB" and s.tonemapEnable == 1 and s.exposure == 1.5
    assert s.maxRenderPasses == 32 and s.environmentMap == b"studio.hdr"      # untouched defaults


def test_errors(tmp_path):
    with pytest.raises(capi.HrcError):
        capi.read_session(tmp_path / "nope.xml")
    bad = tmp_path / "bad.xml"
    bad.write_text("<NotASession><RenderOptions/></NotASession>")
    with pytest.raises(capi.HrcError):
        capi.read_session(bad)
    with pytest.raises(capi.HrcError):
        capi.write_session(tmp_path / "no_such_dir" / "x.xml", capi.session_defaults())


def test_apply_maps_onto_render_options_and_post_processing():
    s = capi.session_defaults()
    s.interactiveMode, s.maxRenderPasses, s.maxRayDepth, s.maxChannelValue = 1, 77, 5, 3.5
    s.sampleMode, s.bokehShape = 2, 3
    s.environmentMap = b"solid color"
    s.environmentSolidColor[:] = [0.25, 0.5, 0.75]
    s.environmentExposureCompensation, s.environmentThetaRotation = 1.5, 0.3
    s.cameraAspectRatio, s.cameraFocusDistance, s.cameraFocalLength, s.cameraApertureRadius, s.cameraFStop = 1.5, 4.0, 35.0, 0.01, 2.8
    s.orbitDistance, s.orbitPhi, s.orbitTheta = 7.0, 0.4, -0.2
    s.orbitTarget[:] = [1.0, 2.0, 3.0]
    s.tonemapEnable, s.exposure, s.brightness, s.vignetteIntensity = 1, 2.0, 0.1, 0.4
    o, p = capi.apply_session(s)
    assert (o.enableInteractiveMode, o.enableOfflineMode) == (1, 1)           # HeatrayRenderer.cpp:594-595 feeds both from InteractiveMode
    assert (o.maxRenderPasses, o.maxRayDepth, o.maxChannelValue, o.sampleMode, o.bokehShape) == (77, 5, 3.5, 2, 3)
    assert (o.envEnabled, o.envUseSolidColor) == (1, 1) and list(o.envSolidColor) == [0.25, 0.5, 0.75]
    assert (o.envExposureCompensation, o.envThetaRotation) == (1.5, np.float32(0.3))
    assert (o.aspectRatio, o.focusDistance, o.focalLength, o.apertureRadius, o.fstop) == (1.5, 4.0, 35.0, np.float32(0.01), np.float32(2.8))
    assert o.resetInternalState == 1
    view = (capi.f32 * 16)()
    capi.load_library().hrhOrbitCameraViewMatrix(capi.f32(7.0), capi.f32(0.4), capi.f32(-0.2), (capi.f32 * 3)(1.0, 2.0, 3.0), view)
    assert list(o.viewMatrix) == list(view)
    assert (p.tonemappingEnabled, p.cameraExposure, p.brightness, p.vignetteIntensity) == (1, 4.0, np.float32(0.1), np.float32(0.4))
    s.environmentMap = b"<none>"
    o, _ = capi.apply_session(s)
    assert (o.envEnabled, o.envUseSolidColor) == (0, 0)
    s.environmentMap = b"studio.hdr"
    o, _ = capi.apply_session(s)
    assert (o.envEnabled, o.envUseSolidColor) == (1, 0)


@pytest.mark.skipif(not REF_SO.exists(), reason="reference probe not built (container only)")
def test_reference_parses_our_files(tmp_path):
    ref = C.CDLL(str(REF_SO))
    rng = np.random.default_rng(9)
    for trial in range(20):
        vals = [float(rng.integers(0, 2)) if k == "b" else float(rng.integers(0, 1 << 31)) if k == "u"
                else float(np.float32(rng.normal() * 10.0 ** rng.integers(-8, 9))) if k == "f" else 0.0 for k in KINDS]
        s = capi.session_defaults()
        fill(s, vals, f"dir {trial}/a&b<c>\"d'.glb", "env.hdr")
        ours, theirs = tmp_path / "ours.xml", tmp_path / "theirs.xml"
        capi.write_session(ours, s)
        arr = (C.c_double * 50)(*vals)
        assert ref.ref_session_write(str(theirs).encode(), arr, s.scene, s.environmentMap) == 1
        assert filecmp.cmp(ours, theirs, shallow=False)
        back = (C.c_double * 50)()
        sc, ev = C.create_string_buffer(1024), C.create_string_buffer(1024)
        assert ref.ref_session_read(str(ours).encode(), back, sc, ev) == 1
        again, missing = capi.read_session(ours)
        nums, strs = numbers_and_strings(again)
        assert missing == 0 and strs == [sc.value.decode(), ev.value.decode()] and nums == list(back)


# ---- scene transform + AABB framing used when a session is applied (HeatrayRenderer.h:343-346, AABB.h:34-52) -------------------
def _unbits(b):
    return np.array(b, np.uint32).view(np.float32)


def test_scene_transform_and_aabb_camera_match_the_reference_glm():
    lib = capi.load_library()
    g = json.loads((GOLDEN / "host_glm_ref.json").read_text())
    assert len(g["scene_transform"]) >= 16 and len(g["aabb_camera"]) >= 16
    for c in g["scene_transform"]:
        y, p, r, s = _unbits(c["in"])
        o = (capi.f32 * 16)()
        lib.hrhSceneTransformMatrix(y, p, r, s, o)
        assert np.array_equal(np.array(o, np.float32).view(np.uint32), np.array(c["out"], np.uint32))
    valid = 0
    for c in g["aabb_camera"]:
        mn, mx, t = (_unbits(c[k]) for k in ("min", "max", "transform"))
        cen, rad = (capi.f32 * 3)(), capi.f32()
        v = lib.hrhAABBCamera((capi.f32 * 3)(*mn), (capi.f32 * 3)(*mx), (capi.f32 * 16)(*t), cen, C.byref(rad))
        bottom = lib.hrhAABBBottom((capi.f32 * 3)(*mn), (capi.f32 * 3)(*mx), (capi.f32 * 16)(*t))
        got = np.array([*cen, rad.value, v, bottom], np.float32)
        assert np.array_equal(got.view(np.uint32), np.array(c["out"], np.uint32))
        valid += v
    assert 0 < valid < len(g["aabb_camera"])                                   # the set holds an invalid box too
    cen, rad = (capi.f32 * 3)(), capi.f32()
    assert lib.hrhAABBCamera((capi.f32 * 3)(-1, -2, -3), (capi.f32 * 3)(1, 2, 3), None, cen, C.byref(rad)) == 1     # NULL transform = identity
    assert list(cen) == [0, 0, 0] and rad.value == np.float32(np.sqrt(np.float32(56.0)))


def _headless():
    from heatray_b200.build import APP, build_apps
    if not APP.exists():
        build_apps()
    return APP


def test_headless_rejects_a_bad_session(tmp_path):
    import subprocess
    res = subprocess.run([str(_headless()), "--session", str(tmp_path / "missing.xml")], capture_output=True, text=True)
    assert res.returncode == 1 and "session" in res.stderr


@pytest.mark.gpu
def test_headless_session_round_trip_reproduces_the_image(tmp_path):
    """Render with flags and --save-session, then render again from that session alone: same pixels, bit for bit. Then change the
    session's scene transform and check the geometry moved."""
    import subprocess
    from heatray_b200 import scenes
    obj = scenes.write_obj(scenes.cornell_box(), tmp_path / "cornell.obj")
    a, b, c, sess = tmp_path / "a.png", tmp_path / "b.png", tmp_path / "c.png", tmp_path / "s.xml"
    res = subprocess.run([str(_headless()), str(obj), "--size", "80", "64", "--passes", "6", "--depth", "5", "--env", "0.7", "0.6", "0.5", "--orbit", "0.3", "-0.1",
                          "--tonemap", "--exposure", "0.5", "--out", str(a), "--save-session", str(sess)], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    s, missing = capi.read_session(sess)
    assert missing == 0 and s.scene.decode() == str(obj) and (s.maxRenderPasses, s.maxRayDepth, s.tonemapEnable) == (6, 5, 1)
    assert s.environmentMap == b"solid color" and list(s.environmentSolidColor) == [np.float32(0.7), np.float32(0.6), np.float32(0.5)]
    assert s.exposure == 0.5 and (s.orbitPhi, s.orbitTheta) == (np.float32(0.3), np.float32(-0.1))
    res = subprocess.run([str(_headless()), "--size", "80", "64", "--session", str(sess), "--out", str(b)], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    assert json.loads(res.stdout.strip().splitlines()[-1])["passes"] == 6
    ia, ib = capi.load_image(a), capi.load_image(b)
    # session floats are written with %.8g (Session.cpp / tinyxml2), so the stored AABB and with it the camera can move by an ulp:
    # the same picture up to a few re-decided paths
    same = np.abs(ia.astype(int) - ib.astype(int))
    print("session round trip: mean abs diff", same.mean(), "equal pixels", (same.max(-1) == 0).mean())
    assert ia.std() > 5 and same.mean() < 1.0 and (same.max(-1) <= 2).mean() > 0.9
    s.rotationYaw, s.scale = 0.6, 2.0                                          # scene transform: every mesh gets scale * yawPitchRoll in front of its own
    capi.write_session(sess, s)
    res = subprocess.run([str(_headless()), "--size", "80", "64", "--session", str(sess), "--out", str(c)], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    ic = capi.load_image(c)
    moved = np.abs(ic.astype(int) - ia.astype(int)).mean()
    print("scene transform: mean abs diff", moved)
    assert ic.std() > 5 and moved > 4 * max(same.mean(), 0.25)


@pytest.mark.gpu
def test_headless_renders_the_default_session_scene(tmp_path):
    """The default session names the built-in "Multi-Material" scene (Session.h:25): heatray_headless builds it like changeScene does."""
    import subprocess
    s = capi.session_defaults()
    assert s.scene == b"Multi-Material"
    s.maxRenderPasses, s.maxRayDepth = 4, 4
    s.environmentMap = b"solid color"
    s.environmentSolidColor[:] = [0.6, 0.6, 0.7]
    s.orbitDistance, s.orbitPhi, s.orbitTheta = 19.0, 0.0, 0.0                 # OrbitCamera defaults (the table's own defaults are placeholders)
    s.orbitTarget[:] = [0.0, 0.0, 0.0]
    s.cameraFocusDistance = 19.0
    # post-processing: the table's defaults are all zero (a black picture); the viewer always writes its own PostProcessingParams
    s.contrast = s.hue = s.saturation = s.red = s.green = s.blue = s.vignetteFalloff = 1.0
    sess, out = tmp_path / "default.xml", tmp_path / "mm.png"
    capi.write_session(sess, s)
    res = subprocess.run([str(_headless()), "--session", str(sess), "--size", "96", "64", "--out", str(out)], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    info = json.loads(res.stdout.strip().splitlines()[-1])
    assert info["ok"] and info["meshes"] == 3 and info["materials"] == 3 and info["passes"] == 4
    img = capi.load_image(out)
    assert img.shape == (64, 96, 3) and img.std() > 1 and len(np.unique(img)) > 16      # a bright ground under a uniform sky: low contrast, not flat
    # and by name on the command line
    res = subprocess.run([str(_headless()), "Sphere Array", "--size", "96", "64", "--passes", "2", "--out", str(tmp_path / "sa.png")], capture_output=True, text=True)
    assert res.returncode == 0 and json.loads(res.stdout.strip().splitlines()[-1])["meshes"] == 20

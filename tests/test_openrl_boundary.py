"""The drop-in boundary, compiled: the reference's OWN host code — PassGenerator.cpp, Scene.cpp, Mesh.cpp, Lighting.cpp, the materials, the
lights, ShaderCodeLoader.cpp and the header-only RLWrapper classes, unmodified, built where they lie by oracle/ref_probe/build_ref.sh into
oracle/_ref/libref_scene.so — runs against libheatray_openrl.so, the product's implementation of the OpenRL entry points (rl.h:383-528)
over the backend C-ABI. What comes out of PassGenerator::renderPass's PixelPackBuffer must be the picture the repo's own host path
(hrhBuildPresetScene + renderer.Renderer) renders on the same backend: bit for bit, because every table, matrix, material block and mesh the
reference code builds is what the host layer builds (each pinned separately elsewhere) and the shim only re-addresses them.
CPU tests bind the shim to the oracle's implementation of the backend ABI (the checker); the GPU tests use the CUDA backend."""
import ctypes as C
import math
import os
from pathlib import Path

import numpy as np
import pytest

from heatray_b200 import capi, renderer, scenes

ROOT = Path(__file__).resolve().parent.parent
REF_SCENE = ROOT / "oracle" / "_ref" / "libref_scene.so"
SHIM = capi.PKG / "libheatray_openrl.so"
SHADERS = ("perspective", "vertex", "passthrough", "physicallyBased", "glass", "environmentLight", "directionalLight", "pointLight", "spotLight")


class Options(C.Structure):
    _fields_ = [("maxRenderPasses", C.c_int32), ("maxRayDepth", C.c_int32), ("interactive", C.c_int32), ("bokehShape", C.c_int32),
                ("envSolidColor", C.c_float * 3), ("envExposure", C.c_float), ("envTheta", C.c_float), ("envEnabled", C.c_int32),
                ("cameraDistance", C.c_float), ("cameraPhi", C.c_float), ("cameraTheta", C.c_float), ("cameraTarget", C.c_float * 3),
                ("focalLength", C.c_float), ("fstop", C.c_float), ("focusDistance", C.c_float)]


class Light(C.Structure):
    _fields_ = [("type", C.c_int32), ("color", C.c_float * 3), ("power", C.c_float), ("position", C.c_float * 3), ("theta", C.c_float), ("phi", C.c_float),
                ("inner", C.c_float), ("outer", C.c_float)]


class RefHost:
    """The reference's PassGenerator + Scene + Lighting behind libref_scene.so, talking to `backend_lib` (prefix) through the OpenRL shim."""

    def __init__(self, backend_lib, prefix, tmp_path, w, h):
        if not REF_SCENE.exists():
            pytest.skip("oracle/_ref/libref_scene.so not built (needs /root/reference)")
        self.shim = C.CDLL(str(SHIM), mode=C.RTLD_GLOBAL)
        self.shim.hrlBindBackend.argtypes = [C.c_void_p, C.c_char_p]
        self.shim.hrlBackendCallCount.restype = C.c_uint64
        self.shim.hrlBackendCallCount.argtypes = [C.c_char_p]
        self.shim.hrlGetLastErrorString.restype = C.c_char_p
        assert self.shim.hrlBindBackend(C.c_void_p(backend_lib._handle), prefix.encode()) == 0
        assert self.shim.hrlBindHostLibrary(C.c_void_p(capi.load_host_library()._handle)) == 0
        self.lib = C.CDLL(str(REF_SCENE))
        self.lib.refscene_create.restype = C.c_void_p
        for name in ("refscene_destroy", "refscene_load_preset", "refscene_add_light", "refscene_remove_light", "refscene_render", "refscene_apply_transform", "refscene_resize"):
            getattr(self.lib, name).argtypes = {"refscene_destroy": [C.c_void_p], "refscene_load_preset": [C.c_void_p, C.c_char_p],
                                                "refscene_add_light": [C.c_void_p, C.POINTER(Light)], "refscene_remove_light": [C.c_void_p, C.c_int],
                                                "refscene_render": [C.c_void_p, C.POINTER(Options), C.c_int, C.c_int, C.c_void_p],
                                                "refscene_apply_transform": [C.c_void_p, C.c_void_p], "refscene_resize": [C.c_void_p, C.c_int, C.c_int]}[name]
        lut = np.ascontiguousarray(renderer.multiscatter_lut(), np.float32)
        self.lib.refscene_set_lut(lut.ctypes.data_as(C.c_void_p), lut.shape[0])
        # ShaderCodeLoader.cpp reads "Resources/shaders/<name>.rlsl" relative to the working directory. The backend never compiles RLSL (it
        # recognises a program by its RL_SHADER_NAME and #define prefix), so stand-in files do: the GPU box has no reference checkout.
        self.cwd = os.getcwd()
        real = Path("/root/reference/Resources/shaders")
        if real.is_dir() and os.environ.get("HR_TEST_STANDIN_SHADERS") != "1":
            os.chdir("/root/reference")
        else:
            d = Path(tmp_path) / "Resources" / "shaders"
            d.mkdir(parents=True, exist_ok=True)
            for s in SHADERS:
                (d / f"{s}.rlsl").write_text("// stand-in: the backend does not compile RLSL\n")
            os.chdir(tmp_path)
        self.w, self.h = w, h
        self.p = C.c_void_p(self.lib.refscene_create(w, h))
        assert self.p.value, (self.shim.hrlGetLastErrorString() or b"").decode()

    def close(self):
        if self.p:
            self.lib.refscene_destroy(self.p)
            self.p = None
        os.chdir(self.cwd)

    def calls(self, name):
        return int(self.shim.hrlBackendCallCount(name.encode()))

    def render(self, opt, passes, reset=True):
        out = np.zeros((self.h, self.w, 4), np.float32)
        wh = self.lib.refscene_render(self.p, C.byref(opt), passes, int(reset), out.ctypes.data_as(C.c_void_p))
        assert wh == self.w * 65536 + self.h, (wh, (self.shim.hrlGetLastErrorString() or b"").decode())
        return out


def _options(spec, passes):
    cam, env = spec.camera, spec.env
    o = Options()
    o.maxRenderPasses, o.maxRayDepth, o.interactive, o.bokehShape = passes, spec.max_depth, 0, spec.bokeh
    o.envEnabled = int(env["enabled"])
    o.envSolidColor[:] = env["solid"] or (0.5, 0.5, 0.5)
    o.envExposure, o.envTheta = env["ev"], env["theta"]
    o.cameraDistance, o.cameraPhi, o.cameraTheta = cam["distance"], cam["phi"], cam["theta"]
    o.cameraTarget[:] = cam["target"]
    o.focalLength, o.fstop, o.focusDistance = cam["focal_length"], cam["fstop"] or -1.0, cam["focus_distance"]
    return o


DIR = dict(color=(1.0, 0.9, 0.8), illuminance=683.0 * 2.0, theta=math.pi / 3, phi=0.4)
DIR2 = dict(color=(0.2, 0.4, 1.0), illuminance=683.0 * 1.5, theta=1.2, phi=2.0)
DIR3 = dict(color=(0.9, 0.2, 0.2), illuminance=683.0 * 1.0, theta=0.7, phi=4.0)
POINT = dict(color=(1.0, 1.0, 1.0), intensity=683.0 * 3.0, position=(0.5, 2.5, 1.0))
SPOT = dict(color=(1.0, 1.0, 0.9), intensity=683.0 * 30.0, position=(-1.0, 4.0, 2.0), theta=math.pi / 2.2, phi=0.3, inner=math.radians(15.0), outer=math.radians(35.0))


def _light(kind, d):
    L = Light()
    L.type = {"dir": 0, "point": 1, "spot": 2}[kind]
    L.color[:] = d["color"]
    L.power = d.get("illuminance", d.get("intensity", 0.0))
    L.position[:] = d.get("position", (0, 0, 0))
    L.theta, L.phi, L.inner, L.outer = d.get("theta", 0.0), d.get("phi", 0.0), d.get("inner", 0.0), d.get("outer", 0.0)
    return L


def _own_path(backend, spec, passes):
    r = renderer.Renderer(backend)
    r.load_scene(spec)
    for i in range(passes):
        r.render_pass(sample_index=i)
    return r.pixels().copy()


def _run_boundary(backend_lib, prefix, make_backend, tmp_path, exact=True):
    w, h, passes = 96, 64, 3
    env = dict(enabled=True, solid=(0.6, 0.7, 0.9), map=None, ev=0.5, theta=0.3)
    cam = dict(distance=7.0, phi=0.4, theta=0.35, target=(0.0, 0.0, 0.0), focal_length=35.0, fstop=2.8, focus_distance=7.0)
    host = RefHost(backend_lib, prefix, tmp_path, w, h)
    try:
        # 1. every built-in scene of HeatrayRenderer::changeScene through the reference's own Scene::addMesh / Material::build / Mesh::Mesh
        for name in scenes.PRESET_SCENES:
            spec = scenes.preset_scene(name, width=w, height=h, passes=passes, max_depth=5, env=env, camera=cam)
            assert host.lib.refscene_load_preset(host.p, name.encode()) == 1
            meshes_before = host.calls("hrcCreateMesh")
            got = host.render(_options(spec, passes), passes)
            assert host.calls("hrcCreateMesh") - meshes_before == len(spec.meshes)
            b = make_backend()
            try:
                want = _own_path(b, spec, passes)
            finally:
                b.close()
            assert np.array_equal(got[..., 3], want[..., 3]) and (got[..., 3] == passes).all()
            if exact:
                assert np.array_equal(got, want), f"{name}: {np.abs(got - want).max()}"
            else:
                assert np.allclose(got, want, rtol=1e-5, atol=1e-6), name
        # 2. Lighting container semantics (Lighting.cpp:201-271): add three directional lights, a point and a spot light, remove the FIRST
        #    directional one: the last takes its slot (swap-with-last + updateLightIndex), so the arrays read [DIR3, DIR2]
        name = "Multi-Material"
        assert host.lib.refscene_load_preset(host.p, name.encode()) == 1
        h1 = host.lib.refscene_add_light(host.p, C.byref(_light("dir", DIR)))
        host.lib.refscene_add_light(host.p, C.byref(_light("dir", DIR2)))
        host.lib.refscene_add_light(host.p, C.byref(_light("dir", DIR3)))
        host.lib.refscene_add_light(host.p, C.byref(_light("point", POINT)))
        host.lib.refscene_add_light(host.p, C.byref(_light("spot", SPOT)))
        host.lib.refscene_remove_light(host.p, h1)
        spec = scenes.preset_scene(name, width=w, height=h, passes=passes, max_depth=5, env=env, camera=cam)
        spec.dir_lights, spec.point_lights, spec.spot_lights = [DIR3, DIR2], [POINT], [SPOT]
        got = host.render(_options(spec, passes), passes)
        b = make_backend()
        try:
            want = _own_path(b, spec, passes)
            spec.dir_lights = [DIR, DIR2, DIR3]
            unremoved = _own_path(b, spec, passes)
        finally:
            b.close()
        assert np.array_equal(got, want) if exact else np.allclose(got, want, rtol=1e-5, atol=1e-6)
        assert not np.allclose(got, unremoved, rtol=1e-3)                      # the removal is visible
        # a sixth light of a type is refused (Lighting.cpp:201-217 returns nullptr)
        for _ in range(3):
            assert host.lib.refscene_add_light(host.p, C.byref(_light("dir", DIR))) >= 0
        assert host.lib.refscene_add_light(host.p, C.byref(_light("dir", DIR))) == -1
        # 3. progressive contract through the reference's own loop: two more passes without a reset continue the same accumulation
        passes_before = host.calls("hrcRenderPass")
        more = host.render(_options(spec, passes + 2), 2, reset=True)
        assert (more[..., 3] == 2).all() and host.calls("hrcRenderPass") - passes_before == 2
        again = host.render(_options(spec, passes + 2), 2, reset=False)
        assert (again[..., 3] == 4).all()
        # 4. Scene::applyTransform rewrites the worldFromEntity uniform of every primitive: the backend mesh moves
        t = np.eye(4, dtype=np.float32); t[3, 1] = 0.75                        # column-major translation by +0.75 in y
        host.lib.refscene_apply_transform(host.p, t.ctypes.data_as(C.c_void_p))
        moved = host.render(_options(spec, passes + 2), 2, reset=True)
        assert not np.allclose(moved[..., :3], more[..., :3], rtol=1e-3)
        assert host.calls("hrcSetMeshTransform") >= len(spec.meshes)
        # 5. PassGenerator::resize: framebuffer, pixel pack buffer and the per-pixel tables follow
        host.lib.refscene_resize(host.p, 64, 48)
        host.w, host.h = 64, 48
        small = host.render(_options(spec, passes), 1, reset=True)
        assert small.shape == (48, 64, 4) and (small[..., 3] == 1).all() and small[..., :3].std() > 0.01
    finally:
        host.close()


def test_reference_host_code_over_the_openrl_entry_points_cpu(oracle_mod, lib, tmp_path):
    """The reference's PassGenerator / Scene / Mesh / Lighting / materials / lights, unmodified, against libheatray_openrl.so bound to the oracle's
    implementation of the backend ABI: bit-identical to the repo's own host path on the same backend."""
    o = oracle_mod.load()
    _run_boundary(o, "orc", lambda: oracle_mod.backend(), tmp_path)


@pytest.mark.gpu
def test_reference_host_code_over_the_openrl_entry_points_gpu(lib, tmp_path):
    """The same through the CUDA backend: what the reference's viewer would show on a B200."""
    _run_boundary(lib, "hrc", lambda: capi.Backend(lib, "hrc", device=0), tmp_path)


@pytest.mark.gpu
def test_reference_viewer_loop_with_backend_lookahead_gpu(lib, tmp_path, monkeypatch):
    """HEATRAY_OPENRL_SPECULATIVE_PASSES: the unmodified reference loop (one rlRenderFrame per pass, PassGenerator.cpp:325-401) with the backend running
    ahead. The Globals block PassGenerator rewrites every pass only to bump sampleIndex must not count as a state change (it would end the look-ahead
    every pass), and the frames must equal the plain loop's: same sample counts, sums up to fp32 addition order."""
    w, h, passes = 96, 64, 8
    env = dict(enabled=True, solid=(0.6, 0.7, 0.9), map=None, ev=0.5, theta=0.3)
    cam = dict(distance=7.0, phi=0.4, theta=0.35, target=(0.0, 0.0, 0.0), focal_length=35.0, fstop=2.8, focus_distance=7.0)
    spec = scenes.preset_scene("Multi-Material", width=w, height=h, passes=passes, max_depth=5, env=env, camera=cam)
    frames = {}
    for depth in (0, 3):
        if depth:
            monkeypatch.setenv("HEATRAY_OPENRL_SPECULATIVE_PASSES", str(depth))
        else:
            monkeypatch.delenv("HEATRAY_OPENRL_SPECULATIVE_PASSES", raising=False)
        host = RefHost(lib, "hrc", tmp_path, w, h)
        try:
            assert host.lib.refscene_load_preset(host.p, b"Multi-Material") == 1
            first = host.render(_options(spec, passes), 3)
            globals_before, options = host.calls("hrcSetGlobals"), host.calls("hrcSetOption")
            frames[depth] = (first, host.render(_options(spec, passes), 5, reset=False))
            assert host.calls("hrcSetGlobals") == globals_before, "the per-pass rewrite of Globals.sampleIndex was forwarded as a state change"
            assert options == (1 if depth else 0)                  # hrcSetOption(HRC_OPT_SPECULATIVE_PASSES) once, at context creation
        finally:
            host.close()
    for k in (0, 1):
        plain, ahead = frames[0][k], frames[3][k]
        assert np.array_equal(plain[..., 3], ahead[..., 3]) and (ahead[..., 3] == (3, 8)[k]).all()
        assert np.allclose(plain[..., :3], ahead[..., :3], rtol=2e-5, atol=1e-6)

"""Host layer (materials, lights, camera, mesh providers, LUT generator) against golden values produced by reference code
(tests/golden/*.json from oracle/ref_probe/make_golden.py) and against hand-computed packing rules."""
import ctypes as C
import json
import math
from pathlib import Path

import numpy as np
import pytest

from heatray_b200 import capi, renderer, scenes

GOLD = Path(__file__).parent / "golden"
f32 = C.c_float


def _f3(v):
    return (C.c_float * 3)(*v)


def _bits(a):
    return np.asarray(a, np.float32).view(np.uint32).reshape(-1)


def test_pbr_material_packing(lib):
    src = capi.HrhPbrMaterial(_f3((1.5, 0.5, -0.2)), _f3((2.0, 0.25, 0.0)), 1.7, 0.0, 0.5, 1.0, 0.3, 1, 0)
    out = capi.PbrParams()
    flags = lib.hrhPackPbrMaterial(C.byref(src), 7, C.byref(out))
    assert flags == capi.MAT_DOUBLE_SIDED
    assert list(out.baseColor) == [1.0, 0.5, 0.0] and list(out.emissiveColor) == [1.0, 0.25, 0.0]
    assert out.metallic == 1.0
    assert out.roughness == np.float32(0.01) and out.roughnessAlpha == np.float32(0.01) * np.float32(0.01)       # roughness floor (quirk B-Q16)
    assert out.specularF0 == np.float32(0.5) * np.float32(0.08)
    assert out.clearCoat == np.float32(1.0) * np.float32(0.2)
    assert out.clearCoatRoughness == np.float32(0.3) and out.clearCoatRoughnessAlpha == np.float32(0.3) * np.float32(0.3)
    assert out.multiscatterLUT == 7 and out.baseColorTexture == -1


def test_material_feature_flags(lib):
    """The #define selection of PhysicallyBasedMaterial.cpp:49-113 / GlassMaterial.cpp:40-74 as HRC_MAT_* flags."""
    def pbr(**kw):
        src = capi.HrhPbrMaterial(_f3((1, 1, 1)), _f3((0, 0, 0)), 0.0, 0.5, 0.5, 0.0, 0.0, 0, 0, **kw)
        out = capi.PbrParams()
        return lib.hrhPackPbrMaterial(C.byref(src), -1, C.byref(out)), out
    assert pbr()[0] == 0
    f, out = pbr(baseColorTexture=3, normalmap=5)
    assert f == (1 << 2) | (1 << 5) and out.baseColorTexture == 3 and out.normalmap == 5 and out.emissiveTexture == -1
    assert pbr(forceEnableAllTextures=1)[0] == (1 << 2) | (1 << 3) | (1 << 6) | (1 << 7)      # emissive / normal maps are never forced
    assert pbr(emissiveTexture=0, clearCoatNormalmap=1, enableVertexColors=1)[0] == (1 << 4) | (1 << 8) | (1 << 9)
    src = capi.HrhPbrMaterial(_f3((1, 1, 1)), _f3((0, 0, 0)), 0.0, 0.5, 0.5, 0.0, 0.0, 1, 1)
    assert lib.hrhPackPbrMaterial(C.byref(src), -1, C.byref(capi.PbrParams())) == 1 | 2
    g = capi.HrhGlassMaterial(_f3((1, 1, 1)), 0.1, 1.5, 0.0, baseColorTexture=2, metallicRoughnessTexture=4)
    out = capi.GlassParams()
    assert lib.hrhPackGlassMaterial(C.byref(g), C.byref(out)) == (1 << 2) | (1 << 3) and out.baseColorTexture == 2 and out.normalmap == -1


def test_glass_material_packing(lib):
    src = capi.HrhGlassMaterial(_f3((0.9, 1.6, 0.6)), 0.0, 1.57, 0.5)
    out = capi.GlassParams()
    lib.hrhPackGlassMaterial(C.byref(src), C.byref(out))
    assert list(out.baseColor) == [np.float32(0.9), 1.0, np.float32(0.6)]
    assert out.roughness == np.float32(0.01) and out.ior == np.float32(1.57) and out.density == 0.5
    f0 = abs((np.float32(1) - np.float32(1.57)) / (np.float32(1) + np.float32(1.57)))
    assert out.specularF0 == np.float32(f0 * f0)


def test_light_packing_units(lib):
    d = capi.DirLight()
    lib.hrhPackDirectionalLight(_f3((1, 0.5, 0.25)), 683.0 * math.pi, math.pi / 2, 0.0, C.byref(d))
    assert np.allclose(list(d.color), np.array([1, 0.5, 0.25]) * math.pi, rtol=1e-6)        # colour * E / 683
    assert np.allclose(list(d.direction), [0, 1, 0], atol=1e-6)                            # default light points straight up TO the light
    p = capi.PointLight()
    lib.hrhPackPointLight(_f3((1, 1, 1)), 683.0, _f3((0, 3, 0)), C.byref(p))
    assert np.allclose(list(p.color), [4 * math.pi] * 3, rtol=1e-6) and list(p.position) == [0, 3, 0]
    s = capi.SpotLight()
    lib.hrhPackSpotLight(_f3((1, 1, 1)), 683.0, _f3((0, 3, 0)), math.pi / 2, 0.0, math.radians(50.0), math.radians(40.0), C.byref(s))
    assert np.allclose(list(s.color), [math.pi] * 3, rtol=1e-6)
    assert np.allclose(list(s.direction), [0, -1, 0], atol=1e-6)                           # FROM the light (quirk B-Q18)
    assert np.isclose(s.angles[1], math.cos(math.radians(40.0)), rtol=1e-6)
    assert np.isclose(s.angles[0], math.cos(math.radians(39.0)), rtol=1e-6)                # inner > outer => outer - 1 degree
    lib.hrhEnvironmentExposure.restype = f32
    assert lib.hrhEnvironmentExposure(f32(3.0)) == 8.0


def test_camera_and_light_direction_vs_glm(lib):
    g = json.loads((GOLD / "host_glm_ref.json").read_text())
    for c in g["orbit_view"]:
        d, phi, th, tx, ty, tz = np.asarray(c["in"], np.uint32).view(np.float32)
        out = (C.c_float * 16)()
        lib.hrhOrbitCameraViewMatrix(f32(d), f32(phi), f32(th), _f3((tx, ty, tz)), out)
        assert np.array_equal(_bits(list(out)), np.asarray(c["out"], np.uint32)), "OrbitCamera::createViewMatrix mismatch"
    for c in g["light_forward"]:
        th, phi = np.asarray(c["in"], np.uint32).view(np.float32)
        s = capi.SpotLight()
        lib.hrhPackSpotLight(_f3((1, 1, 1)), 1.0, _f3((0, 0, 0)), f32(th), f32(phi), 0.1, 0.2, C.byref(s))
        assert np.array_equal(_bits(list(s.direction)), np.asarray(c["out"], np.uint32))
        d = capi.DirLight()
        lib.hrhPackDirectionalLight(_f3((1, 1, 1)), 1.0, f32(th), f32(phi), C.byref(d))
        assert np.array_equal(np.asarray(list(d.direction), np.float32), -np.asarray(c["out"], np.uint32).view(np.float32))


def test_camera_helpers(lib):
    assert np.isclose(lib.hrhFovTan(f32(50.0)), 24.0 / 100.0, rtol=1e-6)                   # tan(atan2(24, 2f))
    assert np.isclose(lib.hrhApertureRadius(f32(50.0), f32(2.8)), 0.017857, rtol=1e-4)


def test_multiscatter_lut_vs_the_lut_file_the_reference_ships(lib):
    """Resources/multiscatter_lut.tiff is what the reference actually samples (MultiScatterUtil.cpp:141-150 loads it; :49-139 is the
    generator that wrote it). Its texels (decoded by Pillow into tests/golden/tiff_ref.npz; FreeImage hands them over bottom-up) against
    the table hrhGenerateMultiscatterLUT builds: the file was generated on another machine and libm, so the bar is 3e-7 absolute with
    at least 99 % of the 16384 texels bit-equal."""
    g = np.load(GOLD / "tiff_ref.npz")
    shipped = g["ref/reference_multiscatter_lut"][::-1, :, 0]
    lut = renderer.multiscatter_lut()
    assert shipped.shape == lut.shape == (128, 128)
    assert np.abs(shipped - lut).max() <= 3e-7
    assert (_bits(shipped) == _bits(lut)).mean() >= 0.99


def test_multiscatter_lut_vs_reference_generator(lib):
    g = json.loads((GOLD / "lut_ref.json").read_text())
    lut = renderer.multiscatter_lut()
    assert lut.shape == (128, 128)
    for key, bits in g.items():
        if key == "rows50_51":
            assert np.array_equal(_bits(lut[50:52]), np.asarray(bits, np.uint32))
        else:
            r, c = map(int, key.split("/"))
            assert _bits(lut[r, c])[0] == bits, key
    # SURVEY.md Appendix C.2
    assert np.isclose(lut[127, 127], 2.21144, rtol=1e-6) and np.isclose(lut[64, 64], 0.1716799, rtol=1e-6) and lut[0, 127] == 0.0
    fresh = np.zeros((8, 8), np.float32)
    lib.hrhGenerateMultiscatterLUT(fresh.ctypes.data_as(C.c_void_p), 8, 256, 2)
    assert np.isfinite(fresh).all() and fresh[7, 7] > fresh[0, 7]


def test_mesh_providers():
    p, n, i = scenes.plane_mesh(15.0, 15.0)
    assert p.tolist() == [[-7.5, 0, 7.5], [7.5, 0, 7.5], [7.5, 0, -7.5], [-7.5, 0, -7.5]] and i.tolist() == [0, 1, 3, 2]
    assert (n == [0, 1, 0]).all()
    sp, sn, si = scenes.sphere_mesh(64, 32, 0.8)
    assert sp.shape == ((64 + 1) * (32 + 2), 3) and si.size == 3 * 2 * 64 * 32            # SphereMeshProvider.h:29,81
    assert np.allclose(np.linalg.norm(sp, axis=1), 0.8, atol=1e-5)
    assert np.allclose(sn * 0.8, sp, atol=1e-5)
    assert si.max() < len(sp)
    # poles: first and last vertex of every slice column
    assert np.allclose(sp[0], [0, 0.8, 0], atol=1e-6) and np.allclose(sp[33], [0, -0.8, 0], atol=1e-6)


def test_scene_triangle_counts_match_baseline():
    assert scenes.cornell_box().triangle_count == 36
    assert scenes.sphere_grid(env_size=(8, 4)).triangle_count == 262150
    assert scenes.material_grid().triangle_count == 495618
    # config 4 is 2*(2047^2) + 400*4096 = 10,018,818; checked on a scaled-down instance to keep the CPU suite fast
    t = scenes.terrain(n=65, blobs=3, env_size=(8, 4))
    assert t.triangle_count == 2 * 64 * 64 + 3 * 4096
    assert 2 * 2047 * 2047 + 400 * 4096 == 10018818

"""Scene import (heatray_b200/csrc/host_import.cpp; reference Scene::loadFromDisk -> AssimpMeshProvider.cpp:65-723).
Own OBJ/MTL and glTF 2.0 readers that apply the reference's Assimp->Heatray mapping rules; the glm-dependent part (light
orientation recovery) is pinned against the reference's own glm (tests/golden/host_glm_ref.json)."""
import base64
import ctypes as C
import json
import math
import struct

import numpy as np
import pytest

from heatray_b200 import capi, renderer, scenes

OBJ = """# quad + triangle, two materials, relative indices, a polygon without normals
mtllib test.mtl
o first
v 0 0 0
v 1 0 0
v 1 1 0
v 0 1 0
vt 0 0
vt 1 0
vt 1 1
vt 0 1
vn 0 0 1
usemtl red
f 1/1/1 2/2/1 3/3/1 4/4/1
o second
v 2 0 0
v 3 0 0
v 2.5 1 0.5
v 2.5 0.5 -1
usemtl shiny
f -4 -3 -2
f -3 -4 -1
g third
usemtl unknown_material
f 1 2 3
"""
MTL = """newmtl red
Kd 0.8 0.1 0.1
Ke 0.5 0.25 0
map_Kd -s 1 1 1 tex/albedo file.png
norm bump.png

newmtl shiny
Kd 0.2 0.3 0.4
Ni 1.5
Pr 0.25
Pm 1.0
Pc 0.6
Pcr 0.1
"""


def _import(path, **kw):
    return scenes.import_scene(path, load_textures=False, **kw)


def test_obj_import_follows_the_assimp_mapping(lib, tmp_path):
    (tmp_path / "test.obj").write_text(OBJ)
    (tmp_path / "test.mtl").write_text(MTL)
    spec = _import(tmp_path / "test.obj")
    assert [m.triangle_count for m in spec.meshes] == [2, 2, 1]                 # the quad is fanned (ProcessMesh :242-250)
    red, shiny, unknown = (spec.materials[m.material].params for m in spec.meshes)
    assert np.allclose(red["baseColor"], (0.8, 0.1, 0.1)) and np.allclose(red["emissiveColor"], (0.5, 0.25, 0))
    assert red["specularF0"] == 0.0                                            # Ni defaults to 1.0 and is always present: ((1-1)/(1+1))^2
    assert red["roughness"] == 1.0 and red["metallic"] == 0.0 and red["doubleSided"] is True
    assert math.isclose(shiny["specularF0"], ((1 - 1.5) / (1 + 1.5)) ** 2, rel_tol=1e-6)
    assert (shiny["roughness"], shiny["metallic"]) == (0.25, 1.0) and math.isclose(shiny["clearCoat"], 0.6, rel_tol=1e-6) and math.isclose(shiny["clearCoatRoughness"], 0.1, rel_tol=1e-6)
    assert np.allclose(unknown["baseColor"], 0.6)                               # Assimp's default grey
    q = spec.meshes[0]
    assert q.positions.shape == (4, 3) and q.uvs.shape == (4, 2) and np.allclose(q.normals, (0, 0, 1))
    assert q.tangents is not None and np.allclose(q.tangents, (1, 0, 0), atol=1e-6) and np.allclose(q.bitangents, (0, 1, 0), atol=1e-6)   # cross(n, t)
    t = spec.meshes[1]                                                           # no normals in the file: smooth normals, unit length, shared edge vertices average the two faces
    assert t.positions.shape == (4, 3) and t.uvs is None
    assert np.allclose(np.linalg.norm(t.normals, axis=1), 1.0, atol=1e-6)
    # texture paths resolve relative to the asset, options are skipped, spaces survive
    h = C.c_void_p()
    assert lib.hrhImportScene(str(tmp_path / "test.obj").encode(), 0, C.byref(h)) == 0
    m = capi.HrhSceneMaterial()
    lib.hrhSceneGetMaterial(h, 0, C.byref(m))
    assert m.texturePath[0].decode() == str(tmp_path) + "/tex/albedo file.png" and m.texturePath[3].decode().endswith("/bump.png") and not m.texturePath[1]
    lib.hrhFreeScene(h)
    # AABB quirk + camera framing (HeatrayRenderer.cpp:302-308, AABB.h:34-44)
    assert np.allclose(spec.camera["target"], (1.5, 0.5, -0.25)) and math.isclose(spec.camera["distance"], 3 * math.sqrt(9 + 1 + 2.25), rel_tol=1e-6)


def test_obj_global_scale_and_errors(lib, tmp_path):
    (tmp_path / "a.obj").write_text("v 0 0 0\nv 100 0 0\nv 0 100 0\nf 1 2 3\n")
    spec = _import(tmp_path / "a.obj", convert_to_meters=True)                   # centimetres -> metres (AssimpMeshProvider.cpp:693-695)
    assert np.allclose(spec.meshes[0].positions.max(0), (1, 1, 0))
    assert spec.materials[0].params["baseColor"] == (pytest.approx(0.6),) * 3
    (tmp_path / "bad.obj").write_text("v 0 0 0\nf 1 2 9\n")
    with pytest.raises(capi.HrcError):
        _import(tmp_path / "bad.obj")
    with pytest.raises(capi.HrcError):
        _import(tmp_path / "missing.obj")
    with pytest.raises(capi.HrcError):
        _import(tmp_path / "scene.fbx")


def _gltf(tmp_path, glb=False):
    pos = np.asarray([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0]], np.float32)
    nrm = np.tile(np.asarray([[0, 0, 1]], np.float32), (4, 1))
    uv = np.asarray([[0, 0], [1, 0], [1, 1], [0, 1]], np.float32)
    col = np.asarray([[255, 0, 0, 255], [0, 255, 0, 255], [0, 0, 255, 255], [255, 255, 255, 255]], np.uint8)
    idx = np.asarray([0, 1, 2, 0, 2, 3], np.uint16)
    strip = np.asarray([0, 1, 3, 2], np.uint8)
    blob, views = b"", []
    for a, target in ((pos, 34962), (nrm, 34962), (uv, 34962), (col, 34962), (idx, 34963), (strip, 34963)):
        while len(blob) % 4:
            blob += b"\0"
        views.append({"buffer": 0, "byteOffset": len(blob), "byteLength": a.nbytes, "target": target})
        blob += a.tobytes()
    doc = {
        "asset": {"version": "2.0"}, "scene": 0, "scenes": [{"nodes": [0, 3]}],
        "nodes": [
            {"name": "root", "translation": [10, 0, 0], "children": [1, 2]},
            {"name": "quadNode", "mesh": 0, "rotation": [0, 0, math.sin(math.pi / 4), math.cos(math.pi / 4)], "scale": [2, 2, 2]},
            {"name": "spotNode", "translation": [0, 5, 0], "rotation": [math.sin(-math.pi / 4), 0, 0, math.cos(math.pi / 4)],
             "extensions": {"KHR_lights_punctual": {"light": 0}}},
            {"name": "matrixNode", "mesh": 1, "matrix": [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, -3, 0, 0, 1],
             "extensions": {"KHR_lights_punctual": {"light": 1}}},
        ],
        "meshes": [
            {"name": "quad", "primitives": [{"attributes": {"POSITION": 0, "NORMAL": 1, "TEXCOORD_0": 2, "COLOR_0": 3}, "indices": 4, "material": 0}]},
            {"name": "strip", "primitives": [{"attributes": {"POSITION": 0}, "indices": 5, "mode": 5, "material": 1},
                                             {"attributes": {"POSITION": 0, "NORMAL": 1}, "indices": 4, "material": 2},
                                             {"attributes": {"POSITION": 0, "NORMAL": 1}, "indices": 4, "material": 3},
                                             {"attributes": {"POSITION": 0, "NORMAL": 1}, "indices": 4}]},
        ],
        "materials": [
            {"name": "opaque", "pbrMetallicRoughness": {"baseColorFactor": [0.5, 0.25, 0.125, 1], "metallicFactor": 0.3, "roughnessFactor": 0.7,
                                                        "baseColorTexture": {"index": 0}}, "emissiveFactor": [1, 0.5, 0], "doubleSided": True,
             "normalTexture": {"index": 0},
             "extensions": {"KHR_materials_clearcoat": {"clearcoatFactor": 0.8, "clearcoatRoughnessFactor": 0.2}, "KHR_materials_ior": {"ior": 1.4}}},
            {"name": "masked", "alphaMode": "MASK"},
            {"name": "blend", "alphaMode": "BLEND", "pbrMetallicRoughness": {"baseColorFactor": [0.9, 0.9, 1, 0.5], "roughnessFactor": 0.1}},
            {"name": "transmissive", "extensions": {"KHR_materials_transmission": {"transmissionFactor": 0.9}, "KHR_materials_ior": {"ior": 1.7}}},
        ],
        "textures": [{"source": 0}], "images": [{"uri": "maps/base%20color.png"}],
        "extensions": {"KHR_lights_punctual": {"lights": [
            {"type": "spot", "color": [1, 0.5, 0.25], "intensity": 100, "spot": {"innerConeAngle": 0.2, "outerConeAngle": 0.6}},
            {"type": "point", "color": [0, 1, 0], "intensity": 40}]}},
        "accessors": [
            {"bufferView": 0, "componentType": 5126, "count": 4, "type": "VEC3", "min": [0, 0, 0], "max": [1, 1, 0]},
            {"bufferView": 1, "componentType": 5126, "count": 4, "type": "VEC3"},
            {"bufferView": 2, "componentType": 5126, "count": 4, "type": "VEC2"},
            {"bufferView": 3, "componentType": 5121, "normalized": True, "count": 4, "type": "VEC4"},
            {"bufferView": 4, "componentType": 5123, "count": 6, "type": "SCALAR"},
            {"bufferView": 5, "componentType": 5121, "count": 4, "type": "SCALAR"},
        ],
        "bufferViews": views,
    }
    if glb:
        doc["buffers"] = [{"byteLength": len(blob)}]
        js = json.dumps(doc).encode()
        js += b" " * (-len(js) % 4)
        bin_ = blob + b"\0" * (-len(blob) % 4)
        body = struct.pack("<II", len(js), 0x4E4F534A) + js + struct.pack("<II", len(bin_), 0x004E4942) + bin_
        p = tmp_path / "scene.glb"
        p.write_bytes(b"glTF" + struct.pack("<II", 2, 12 + len(body)) + body)
    else:
        doc["buffers"] = [{"byteLength": len(blob), "uri": "data:application/octet-stream;base64," + base64.b64encode(blob).decode()}]
        p = tmp_path / "scene.gltf"
        p.write_text(json.dumps(doc))
    return p


@pytest.mark.parametrize("glb", [False, True])
def test_gltf_import_follows_the_assimp_mapping(lib, tmp_path, glb):
    spec = _import(_gltf(tmp_path, glb))
    assert [m.triangle_count for m in spec.meshes] == [2, 2, 2, 2, 2]
    quad = spec.meshes[0]
    # node transform: root translation * (rotation 90 deg about z, scale 2)
    want = np.asarray([[0, -2, 0, 10], [2, 0, 0, 0], [0, 0, 2, 0], [0, 0, 0, 1]], np.float32)
    assert np.allclose(quad.transform, want, atol=1e-6)
    assert np.allclose(quad.uvs[:, 1], [1, 1, 0, 0])                              # v flipped like Assimp's importer (textures load flipped too)
    assert np.allclose(quad.colors[0], (1, 0, 0)) and np.allclose(quad.colors[3], (1, 1, 1))   # normalised u8
    assert quad.tangents is not None
    strip = spec.meshes[1]
    assert strip.indices.tolist() == [0, 1, 3, 1, 3, 2] or strip.indices.tolist() == [0, 1, 3, 3, 1, 2]     # strip expanded with the winding kept
    assert np.allclose(strip.transform[:3, 3], (-3, 0, 0))
    assert np.allclose(np.abs(strip.normals), (0, 0, 1), atol=1e-6)              # generated
    opaque, masked, blend, transmissive, default = (spec.materials[m.material] for m in spec.meshes)
    p = opaque.params
    assert opaque.kind == "pbr" and np.allclose(p["baseColor"], (0.5, 0.25, 0.125)) and np.allclose(p["emissiveColor"], (1, 0.5, 0))
    assert math.isclose(p["metallic"], 0.3, rel_tol=1e-6) and math.isclose(p["roughness"], 0.7, rel_tol=1e-6) and p["doubleSided"] and p["vertexColors"]
    assert math.isclose(p["clearCoat"], 0.8, rel_tol=1e-6) and math.isclose(p["clearCoatRoughness"], 0.2, rel_tol=1e-6)
    assert math.isclose(p["specularF0"], ((1 - 1.4) / (1 + 1.4)) ** 2, rel_tol=1e-5)
    assert masked.kind == "pbr" and masked.params["alphaMask"] and not masked.params["doubleSided"] and masked.params["metallic"] == 1.0   # glTF defaults
    assert masked.params["specularF0"] == 0.5
    assert blend.kind == "glass" and math.isclose(blend.params["ior"], 1.33, rel_tol=1e-6) and math.isclose(blend.params["roughness"], 0.1, rel_tol=1e-6)
    assert transmissive.kind == "glass" and math.isclose(transmissive.params["ior"], 1.7, rel_tol=1e-6) and math.isclose(transmissive.params["density"], 0.05, rel_tol=1e-6)
    assert default.kind == "pbr" and np.allclose(default.params["baseColor"], 1.0)
    # lights: intensity folded into the colour by Assimp and split again by vector length (ProcessLight :541-545)
    (spot,), (point,) = spec.spot_lights, spec.point_lights
    c = np.asarray([1, 0.5, 0.25]) * 100
    assert math.isclose(spot["intensity"], float(np.linalg.norm(c)), rel_tol=1e-6) and np.allclose(spot["color"], c / np.linalg.norm(c))
    assert np.allclose(spot["position"], (10, 5, 0)) and math.isclose(spot["inner"], 0.2, rel_tol=1e-6) and math.isclose(spot["outer"], 0.6, rel_tol=1e-6)
    assert np.allclose(point["position"], (-3, 0, 0)) and math.isclose(point["intensity"], 40.0, rel_tol=1e-6)
    h = C.c_void_p()
    lib.hrhImportScene(str(_gltf(tmp_path, glb)).encode(), 0, C.byref(h))
    m = capi.HrhSceneMaterial()
    lib.hrhSceneGetMaterial(h, 0, C.byref(m))
    assert m.texturePath[0].decode() == str(tmp_path) + "/maps/base color.png" and m.texturePath[3] == m.texturePath[0]
    lib.hrhFreeScene(h)


def test_imported_light_orientation_matches_reference_glm(lib):
    from pathlib import Path
    cases = json.loads((Path(__file__).parent / "golden" / "host_glm_ref.json").read_text())["light_angles"]
    assert len(cases) >= 20
    for c in cases:
        m = np.asarray(c["in"], np.uint32).view(np.float32).copy()
        t, p = C.c_float(), C.c_float()
        lib.hrhLightAnglesFromTransform(m.ctypes.data_as(capi.fp), C.byref(t), C.byref(p))
        got = np.asarray([t.value, p.value], np.float32).view(np.uint32)
        assert np.array_equal(got, np.asarray(c["out"], np.uint32))


def _triangle_set(spec):
    out = []
    for ms in spec.meshes:
        M = np.asarray(ms.transform, np.float64)
        P = np.asarray(ms.positions, np.float64).reshape(-1, 3) @ M[:3, :3].T + M[:3, 3]
        idx = np.asarray(ms.indices).reshape(-1)
        tris = idx.reshape(-1, 3) if ms.draw_mode == capi.DRAW_TRIANGLES else np.asarray([(idx[j], idx[j + 1], idx[j + 2]) for j in range(len(idx) - 2)])
        for t in tris:
            out.append(tuple(sorted(tuple(np.round(P[v], 5)) for v in t)))
    return sorted(out)


def test_cornell_box_obj_round_trip(lib, tmp_path):
    """BASELINE config #1 names a Cornell-box OBJ: the procedural scene written as OBJ + MTL and read back is the same 36 triangles,
    the same materials and (through the oracle) the same picture up to the F0 rounding of the Ni round trip."""
    spec = scenes.cornell_box(width=48, height=48, passes=2)
    path = scenes.write_obj(spec, tmp_path / "cornell.obj")
    imp = scenes.import_scene(path, width=48, height=48, passes=2, max_depth=4, env=spec.env, camera=spec.camera)
    assert imp.triangle_count == 36 and _triangle_set(imp) == _triangle_set(spec)
    for a, b in zip(spec.materials, imp.materials):
        assert np.allclose(a.params["baseColor"], b.params["baseColor"]) and np.allclose(a.params["emissiveColor"], b.params["emissiveColor"])
        assert math.isclose(a.params["specularF0"], b.params["specularF0"], rel_tol=1e-5)


def test_imported_scene_renders_like_the_procedural_one(oracle_backend, lib, tmp_path):
    spec = scenes.cornell_box(width=40, height=40, passes=3)
    imp = scenes.import_scene(scenes.write_obj(spec, tmp_path / "cornell.obj"), width=40, height=40, passes=3, max_depth=4, env=spec.env, camera=spec.camera)
    imp.point_lights = spec.point_lights                                       # OBJ has no lights
    r = renderer.Renderer(oracle_backend)
    imgs = []
    for s in (spec, imp):
        r.load_scene(s)
        for _ in range(s.passes):
            r.render_pass()
        imgs.append(r.pixels().copy())
    a, b = imgs
    assert np.array_equal(a[..., 3], b[..., 3])
    assert np.isclose(a[..., :3], b[..., :3], rtol=2e-2, atol=2e-3).mean() > 0.97    # same geometry and lights; primitive order / F0 rounding differ


@pytest.mark.gpu
def test_gpu_renders_imported_gltf_like_the_oracle(oracle_backend, gpu_backend, tmp_path):
    """An imported glTF asset (nodes, strip primitive, vertex colours, glass by BLEND / transmission, punctual lights) end to end on both backends."""
    spec = scenes.import_scene(_gltf(tmp_path, True), width=96, height=64, passes=3, max_depth=5, load_textures=False,
                               camera=dict(distance=22.0, target=(4.0, 1.0, 0.0)))
    ro, rg = renderer.Renderer(oracle_backend), renderer.Renderer(gpu_backend)
    ro.load_scene(spec)
    rg.load_scene(spec)
    for _ in range(spec.passes):
        ro.render_pass()
        rg.render_pass()
    po, pg = ro.pixels(), rg.pixels()
    assert np.array_equal(po[..., 3], pg[..., 3])
    assert np.isclose(pg[..., :3], po[..., :3], rtol=1e-3, atol=1e-4).all(-1).mean() >= 0.99
    assert po[..., :3].std() > 1e-3


def test_builtin_scenes_follow_change_scene(lib, oracle_backend):
    """hrhBuildPresetScene vs HeatrayRenderer::changeScene (HeatrayRenderer.cpp:123-286): meshes, transforms and material parameters."""
    assert scenes.PRESET_SCENES == ("Editable PBR Material", "Editable Glass Material", "Multi-Material", "Sphere Array")
    sp = scenes.preset_scene("Multi-Material")
    assert [len(m.indices) // 3 for m in sp.meshes] == [2, 5000, 5000]                  # PlaneMeshProvider(15, 15), SphereMeshProvider(50, 50, 1)
    plane, s1, s2 = sp.meshes
    assert np.allclose(plane.transform[:3, 3], (0.0, -1.5, 0.0)) and np.abs(plane.positions).max() == 7.5
    assert np.allclose(s1.transform[:3, 3], (-0.9, -0.5, -0.8)) and np.allclose(s2.transform[:3, 3], (1.2, -0.5, 0.8))
    assert np.allclose(np.linalg.norm(s1.positions, axis=1), 1.0, atol=1e-6)
    g, m1, m2 = (sp.materials[m.material] for m in sp.meshes)
    assert g.kind == "pbr" and np.allclose(g.params["baseColor"], 0.9) and g.params["roughness"] == 1.0 and g.params["specularF0"] == 0.0
    assert m1.params["metallic"] == 1.0 and m1.params["roughness"] == np.float32(0.1) and np.allclose(m1.params["baseColor"], 0.4) and m1.params["specularF0"] == np.float32(0.3)
    assert m2.kind == "glass" and np.allclose(m2.params["baseColor"], (0.9, 0.6, 0.6)) and m2.params["ior"] == np.float32(1.57)
    assert m2.params["roughness"] == np.float32(0.1) and m2.params["density"] == 0.5
    arr = scenes.preset_scene("Sphere Array")
    assert len(arr.meshes) == 20 and len(arr.materials) == 20
    rough, x = np.float32(0.0), np.float32(-5.0 * 1.1 + 0.55)
    for i in range(10):
        for row, metal, y in ((0, 0.0, 0.0), (1, 1.0, 1.5)):
            mesh = arr.meshes[row * 10 + i]
            p = arr.materials[mesh.material].params
            assert p["metallic"] == metal and p["roughness"] == rough and np.allclose(p["baseColor"], 1.0) and p["specularF0"] == 0.0
            assert np.allclose(mesh.transform[:3, 3], (x, y, 0.0), atol=1e-6)
            assert np.allclose(np.linalg.norm(mesh.positions, axis=1), 0.5, atol=1e-6)
        rough = np.float32(rough + np.float32(0.1)); x = np.float32(x + np.float32(1.1))      # the reference accumulates both in fp32
    for name, kind in (("Editable PBR Material", "pbr"), ("Editable Glass Material", "glass")):
        e = scenes.preset_scene(name)
        assert len(e.meshes) == 1 and e.materials[0].kind == kind and np.allclose(e.materials[0].params["baseColor"], 0.8)
    with pytest.raises(capi.HrcError):
        scenes.preset_scene("Cornell")
    # and it renders: the default camera (OrbitCamera.h:22-25) looks at the two spheres on the ground plane
    r = renderer.Renderer(oracle_backend)
    r.load_scene(scenes.preset_scene("Multi-Material", width=48, height=32, passes=2, max_depth=3))
    r.render_pass(); r.render_pass()
    px = r.pixels()
    assert np.isfinite(px).all() and (px[..., 3] == 2).all() and px[..., :3].std() > 0.01


def test_ground_plane_like_the_viewers_button(lib, tmp_path):
    """HeatrayRenderer.cpp:1167-1193: side = max(1, size_t(AABB diagonal)) * 5, at the AABB's bottom, PBR 0 / 0.9 / 0.9 / 0.2."""
    obj = scenes.write_obj(scenes.cornell_box(), tmp_path / "cornell.obj")
    plain = scenes.import_scene(obj, load_textures=False)
    sp = scenes.import_scene(obj, load_textures=False, ground_plane=True)
    assert len(sp.meshes) == len(plain.meshes) + 1 and len(sp.materials) == len(plain.materials) + 1
    world = np.concatenate([m.positions @ np.asarray(m.transform)[:3, :3].T + np.asarray(m.transform)[:3, 3] for m in plain.meshes])
    lo, hi = world.min(0), world.max(0)
    side = max(1, int(np.float32(np.linalg.norm((hi - lo).astype(np.float32))))) * 5
    gp = sp.meshes[-1]
    assert len(gp.indices) == 6 and np.abs(gp.positions[:, [0, 2]]).max() == side / 2 and (gp.positions[:, 1] == 0).all()
    assert np.isclose(np.asarray(gp.transform)[1, 3], lo[1], atol=1e-6) and np.allclose(np.asarray(gp.transform)[[0, 2], 3], 0.0)
    p = sp.materials[gp.material].params
    assert (p["metallic"], p["roughness"], p["specularF0"]) == (0.0, np.float32(0.9), np.float32(0.2)) and np.allclose(p["baseColor"], 0.9)


def test_malformed_scene_files_are_rejected_not_crashed_on(lib, tmp_path):
    """Seeded mutations of a .gltf, a .glb, an .obj and its .mtl (bytes, digits, cuts, truncation) and a bufferView with a negative
    offset: the importers must return an error or a scene, never crash or read outside a buffer."""
    from test_images import _mutate
    rng = np.random.default_rng(77)
    glb, gltf = _gltf(tmp_path, True), _gltf(tmp_path, False)
    obj = scenes.write_obj(scenes.cornell_box(), tmp_path / "cornell.obj")
    bases = {p: p.read_bytes() for p in (glb, gltf, obj, tmp_path / "cornell.mtl")}
    imported = rejected = 0
    for t in range(800):
        target = list(bases)[t % 4]
        target.write_bytes(_mutate(rng, bases[target], structural=b"0123456789-e.{}[],:\""))
        opened = obj if target.suffix == ".mtl" else target
        try:
            spec = scenes.import_scene(opened, load_textures=False)
            assert spec is not None
            imported += 1
        except (capi.HrcError, ValueError):
            rejected += 1
        target.write_bytes(bases[target])
    assert imported > 50 and rejected > 50
    doc = json.loads(bases[gltf].decode())
    doc["bufferViews"][0]["byteOffset"] = -4096
    gltf.write_text(json.dumps(doc))
    with pytest.raises(capi.HrcError):
        scenes.import_scene(gltf, load_textures=False)
    doc["bufferViews"][0]["byteOffset"] = 0
    doc["accessors"][0]["count"] = 1 << 30
    gltf.write_text(json.dumps(doc))
    with pytest.raises(capi.HrcError):
        scenes.import_scene(gltf, load_textures=False)


# ---- heatray_headless: load -> passes -> resolve -> screenshot as one native program ---------------------------------------
def _headless():
    from heatray_b200.build import APP, build_apps
    if not APP.exists():
        build_apps()
    return APP


def test_headless_program_fails_loudly_without_a_gpu(lib, tmp_path):
    import subprocess
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    spec = scenes.cornell_box()
    obj = scenes.write_obj(spec, tmp_path / "cornell.obj")
    res = subprocess.run([str(_headless()), str(obj), "--size", "32", "32", "--passes", "1", "--out", str(tmp_path / "o.png")], capture_output=True, text=True)
    assert res.returncode == 1 and "no CPU path" in res.stderr and not (tmp_path / "o.png").exists()
    assert subprocess.run([str(_headless())], capture_output=True).returncode == 2


@pytest.mark.gpu
def test_headless_program_renders_assets_to_png_and_hdr(lib, gold_texture_dir, tmp_path):
    """OBJ with an MTL texture and a glTF with lights through heatray_headless; the PNG is the display pipeline's output of the HDR means."""
    import subprocess
    spec = scenes.cornell_box()
    obj = scenes.write_obj(spec, tmp_path / "cornell.obj")
    with open(tmp_path / "cornell.mtl", "a") as f:
        f.write("\n")
    mtl = (tmp_path / "cornell.mtl").read_text().replace("newmtl m0\n", "newmtl m0\nmap_Kd albedo.png\n", 1)
    (tmp_path / "cornell.mtl").write_text(mtl)
    (tmp_path / "albedo.png").write_bytes(gold_texture_dir["file/png_mixed_RGB"].tobytes())
    out, hdr = tmp_path / "c.png", tmp_path / "c.pfm"
    res = subprocess.run([str(_headless()), str(obj), "--size", "96", "80", "--passes", "8", "--depth", "4", "--distance-scale", "0.45", "--env", "0.8", "0.8", "0.8",
                          "--out", str(out), "--hdr", str(hdr), "--tonemap"], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    info = json.loads(res.stdout.strip().splitlines()[-1])
    assert info["ok"] and info["passes"] == 8 and info["meshes"] == len(spec.meshes) and info["textures"] == 1 and info["textures_missing"] == 0
    png = capi.load_image(out)
    assert png.shape == (80, 96, 3) and png.std() > 10 and png.max() > 120
    raw = hdr.read_bytes()
    head = b"PF\n96 80\n-1.0\n"
    mean = np.frombuffer(raw[len(head):], "<f4").reshape(80, 96, 3)            # bottom-up rows, rgb / alpha
    assert np.isfinite(mean).all() and mean.max() > 0.2
    # the PNG must be the colour pipeline applied to those means: re-run the pipeline through the C-ABI on a buffer holding them
    b = capi.Backend(lib, "hrc", device=0)
    b.resize(96, 80)
    b.add_pixels(np.concatenate([mean, np.ones((80, 96, 1), np.float32)], -1))
    ldr = b.resolve(capi.PostProcess(tonemappingEnabled=1), capi.RESOLVE_DISPLAY_RGB8)
    b.close()
    assert np.abs(np.flipud(ldr).astype(int) - png.astype(int)).max() <= 1
    # glTF with punctual lights
    res = subprocess.run([str(_headless()), str(_gltf(tmp_path, True)), "--size", "64", "48", "--passes", "4", "--out", str(tmp_path / "g.png")], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    info = json.loads(res.stdout.strip().splitlines()[-1])
    assert info["ok"] and info["lights"] == [0, 1, 1] and info["meshes"] == 5 and info["textures_missing"] >= 1     # the glTF names a texture file that does not exist
    assert capi.load_image(tmp_path / "g.png").std() > 5


@pytest.fixture()
def gold_texture_dir():
    from pathlib import Path
    return np.load(Path(__file__).parent / "golden" / "images_ref.npz")

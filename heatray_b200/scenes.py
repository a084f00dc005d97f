"""Procedural bench/test scenes — the concrete synthetic inputs of BASELINE.json's configs.

Geometry comes from the host layer's mesh providers (hrhPlaneMesh / hrhSphereMesh, the restated
PlaneMeshProvider / SphereMeshProvider) plus numpy generators for boxes and terrain. Placement and material
parameters are drawn from SplitMix64(seed = config number). The reference's studio.hdr environment is not
copied into this repository; `studio_environment()` synthesises a lat-long HDR map of the same size.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass, field
from pathlib import Path

import numpy as np

from . import capi


class SplitMix64:
    def __init__(self, seed: int):
        self.s = seed & 0xFFFFFFFFFFFFFFFF

    def next(self) -> int:
        self.s = (self.s + 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
        z = self.s
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
        return z ^ (z >> 31)

    def uniform(self, lo=0.0, hi=1.0) -> float:
        return lo + (hi - lo) * ((self.next() >> 11) * (1.0 / (1 << 53)))


@dataclass
class MeshSpec:
    positions: np.ndarray
    normals: np.ndarray
    indices: np.ndarray
    material: int
    transform: np.ndarray = field(default_factory=lambda: np.eye(4, dtype=np.float32))
    draw_mode: int = capi.DRAW_TRIANGLES
    uvs: np.ndarray = None          # optional per-vertex attributes (MeshProvider.h VertexBufferUsage)
    tangents: np.ndarray = None
    bitangents: np.ndarray = None
    colors: np.ndarray = None

    @property
    def triangle_count(self):
        n = self.indices.size
        return n // 3 if self.draw_mode == capi.DRAW_TRIANGLES else max(0, n - 2)


@dataclass
class MaterialSpec:
    kind: str                      # "pbr" | "glass"
    params: dict
    textures: dict = field(default_factory=dict)   # slot name (HrhPbrMaterial field) -> float32 array (H, W, 4), row 0 = t = 0


@dataclass
class SceneSpec:
    name: str
    width: int
    height: int
    passes: int
    max_depth: int
    meshes: list
    materials: list
    env: dict                      # {"enabled": bool, "solid": (r,g,b) | None, "map": ndarray | None, "ev": float, "theta": float}
    dir_lights: list = field(default_factory=list)     # dicts: color, illuminance, theta, phi
    point_lights: list = field(default_factory=list)   # dicts: color, intensity, position
    spot_lights: list = field(default_factory=list)    # dicts: color, intensity, position, theta, phi, inner, outer
    camera: dict = field(default_factory=dict)         # distance, phi, theta, target, focal_length, fstop, focus_distance
    bokeh: int = capi.BOKEH_CIRCULAR
    max_channel_value: float = math.pi

    @property
    def triangle_count(self):
        return sum(m.triangle_count for m in self.meshes)


# ---- transforms (row-major mathematical matrices; capi transposes into glm layout) ---------------
def translate(x, y, z):
    m = np.eye(4, dtype=np.float32)
    m[:3, 3] = (x, y, z)
    return m


def scale(x, y, z):
    return np.diag([x, y, z, 1.0]).astype(np.float32)


def rotate(axis, angle):
    c, s = math.cos(angle), math.sin(angle)
    m = np.eye(4, dtype=np.float32)
    if axis == "x":
        m[1, 1], m[1, 2], m[2, 1], m[2, 2] = c, -s, s, c
    elif axis == "y":
        m[0, 0], m[0, 2], m[2, 0], m[2, 2] = c, s, -s, c
    else:
        m[0, 0], m[0, 1], m[1, 0], m[1, 1] = c, -s, s, c
    return m


# ---- mesh providers -------------------------------------------------------------------------------
def plane_mesh(width, length, with_uv=False):
    lib = capi.load_host_library()
    pos, nrm, uv, idx = np.zeros((4, 3), np.float32), np.zeros((4, 3), np.float32), np.zeros((4, 2), np.float32), np.zeros(4, np.uint32)
    lib.hrhPlaneMesh(width, length, capi._ptr(pos), capi._ptr(nrm), capi._ptr(uv), capi._ptr(idx))
    return (pos, nrm, idx, uv) if with_uv else (pos, nrm, idx)


_sphere_cache = {}


def sphere_mesh_uv(u_slices, v_slices, radius):
    """Sphere with texture coordinates plus a tangent frame (tangent = d position / d longitude, bitangent = n x t)."""
    pos, nrm, idx = sphere_mesh(u_slices, v_slices, radius)
    uv = _sphere_cache[(u_slices, v_slices, radius, "uv")]
    up = np.array([0.0, 1.0, 0.0], np.float32)
    tan = np.cross(up, nrm)
    ln = np.linalg.norm(tan, axis=1, keepdims=True)
    tan = np.where(ln > 1e-6, tan / np.maximum(ln, 1e-6), np.array([1.0, 0.0, 0.0], np.float32)).astype(np.float32)
    bit = np.cross(nrm, tan).astype(np.float32)
    return pos, nrm, idx, uv, tan, bit


def sphere_mesh(u_slices, v_slices, radius):
    key = (u_slices, v_slices, radius)
    if key not in _sphere_cache:
        lib = capi.load_host_library()
        nv, ni = C.c_uint32(), C.c_uint32()
        lib.hrhSphereMesh(u_slices, v_slices, radius, None, None, None, None, C.byref(nv), C.byref(ni))
        pos, nrm = np.zeros((nv.value, 3), np.float32), np.zeros((nv.value, 3), np.float32)
        uv, idx = np.zeros((nv.value, 2), np.float32), np.zeros(ni.value, np.uint32)
        lib.hrhSphereMesh(u_slices, v_slices, radius, capi._ptr(pos), capi._ptr(nrm), capi._ptr(uv), capi._ptr(idx), None, None)
        _sphere_cache[key] = (pos, nrm, idx)
        _sphere_cache[key + ("uv",)] = uv
    return _sphere_cache[key]


def box_mesh(sx, sy, sz):
    """Axis-aligned box centred at the origin: 6 faces x 2 triangles with per-face normals (CCW outward)."""
    hx, hy, hz = sx / 2, sy / 2, sz / 2
    faces = [((1, 0, 0), (0, 1, 0), (0, 0, 1)), ((-1, 0, 0), (0, 0, 1), (0, 1, 0)),
             ((0, 1, 0), (0, 0, 1), (1, 0, 0)), ((0, -1, 0), (1, 0, 0), (0, 0, 1)),
             ((0, 0, 1), (1, 0, 0), (0, 1, 0)), ((0, 0, -1), (0, 1, 0), (1, 0, 0))]
    h = np.array([hx, hy, hz], np.float32)
    pos, nrm, idx = [], [], []
    for n, u, v in faces:
        n, u, v = np.array(n, np.float32), np.array(u, np.float32), np.array(v, np.float32)
        base = len(pos)
        for a, b in ((-1, -1), (1, -1), (1, 1), (-1, 1)):
            pos.append((n + a * u + b * v) * h)
            nrm.append(n)
        idx += [base, base + 1, base + 2, base, base + 2, base + 3]
    return np.array(pos, np.float32), np.array(nrm, np.float32), np.array(idx, np.uint32)


def _value_noise(x, y, seed):
    """Deterministic lattice value noise on float64 arrays (hash of integer lattice points, smoothstep blend)."""
    xi, yi = np.floor(x).astype(np.int64), np.floor(y).astype(np.int64)
    fx, fy = x - xi, y - yi

    k = np.uint64((seed * 2654435761) & 0xFFFFFFFF)

    def h(a, b):
        v = (a.astype(np.uint64) * np.uint64(374761393) + b.astype(np.uint64) * np.uint64(668265263) + k) & np.uint64(0xFFFFFFFF)
        v = ((v ^ (v >> np.uint64(13))) * np.uint64(1274126177)) & np.uint64(0xFFFFFFFF)
        return ((v ^ (v >> np.uint64(16))) & np.uint64(0xFFFF)).astype(np.float64) / 65535.0

    sx, sy = fx * fx * (3 - 2 * fx), fy * fy * (3 - 2 * fy)
    a, b = h(xi, yi), h(xi + 1, yi)
    c, d = h(xi, yi + 1), h(xi + 1, yi + 1)
    return (a * (1 - sx) + b * sx) * (1 - sy) + (c * (1 - sx) + d * sx) * sy


def terrain_mesh(n, size, height, seed):
    """n x n vertex heightfield over [-size/2, size/2]^2 displaced by 5-octave value noise; 2*(n-1)^2 triangles."""
    g = np.linspace(0.0, 1.0, n)
    u, v = np.meshgrid(g, g, indexing="xy")
    hgt = np.zeros_like(u)
    amp, freq = 1.0, 4.0
    for o in range(5):
        hgt += amp * _value_noise(u * freq, v * freq, seed + o)
        amp *= 0.5
        freq *= 2.0
    hgt = (hgt / 1.9375 - 0.5) * height
    x, z = (u - 0.5) * size, (v - 0.5) * size
    pos = np.stack([x, hgt, z], -1).astype(np.float32)
    dx = np.gradient(hgt, size / (n - 1), axis=1)
    dz = np.gradient(hgt, size / (n - 1), axis=0)
    nrm = np.stack([-dx, np.ones_like(dx), -dz], -1)
    nrm = (nrm / np.linalg.norm(nrm, axis=-1, keepdims=True)).astype(np.float32)
    i = (np.arange(n - 1)[:, None] * n + np.arange(n - 1)[None, :]).reshape(-1).astype(np.uint32)
    idx = np.stack([i, i + n, i + 1, i + 1, i + n, i + n + 1], -1).reshape(-1)
    return pos.reshape(-1, 3), nrm.reshape(-1, 3), idx.astype(np.uint32)


def studio_environment(width=2048, height=1024):
    """Synthetic stand-in for Resources/Environments/studio.hdr (2048x1024): dim grey room, bright ceiling softboxes,
    a warm key light and a cool fill. Row 0 is the bottom row (t = 0)."""
    t = (np.arange(height, dtype=np.float32) + 0.5) / height
    s = (np.arange(width, dtype=np.float32) + 0.5) / width
    S, T = np.meshgrid(s, t, indexing="xy")
    img = np.empty((height, width, 3), np.float32)
    base = 0.08 + 0.25 * T
    img[..., 0], img[..., 1], img[..., 2] = base, base, base * 1.05

    def softbox(cs, ct, ws, wt, col):
        m = np.clip(1 - np.abs(S - cs) / ws, 0, 1) * np.clip(1 - np.abs(T - ct) / wt, 0, 1)
        m = np.minimum(1.0, m * 4.0)
        for k in range(3):
            img[..., k] += m * col[k]

    softbox(0.25, 0.80, 0.06, 0.08, (18.0, 17.0, 15.0))
    softbox(0.62, 0.85, 0.10, 0.06, (9.0, 9.5, 11.0))
    softbox(0.85, 0.55, 0.03, 0.15, (5.0, 4.0, 3.0))
    softbox(0.45, 0.60, 0.02, 0.04, (30.0, 28.0, 24.0))
    return img


# ---- the configs -----------------------------------------------------------------------------------
def _pbr(base, rough=1.0, metal=0.0, f0=0.5, emissive=(0, 0, 0), cc=0.0, ccr=0.0, double_sided=True, alpha_mask=False,
         vertex_colors=False, force_textures=False, textures=None):
    return MaterialSpec("pbr", dict(baseColor=base, emissiveColor=emissive, metallic=metal, roughness=rough, specularF0=f0,
                                    clearCoat=cc, clearCoatRoughness=ccr, doubleSided=double_sided, alphaMask=alpha_mask,
                                    vertexColors=vertex_colors, forceEnableAllTextures=force_textures), textures or {})


def _glass(base, rough, ior, density, textures=None):
    return MaterialSpec("glass", dict(baseColor=base, roughness=rough, ior=ior, density=density), textures or {})


def cornell_box(width=512, height=512, passes=64, max_depth=4):
    """Config #1: 5 walls (10 tris) + tall box (12) + short box (12) + ceiling emitter quad (2) = 36 triangles,
    one point light just below the ceiling, no environment."""
    mats = [_pbr((0.73, 0.73, 0.73)), _pbr((0.65, 0.05, 0.05)), _pbr((0.12, 0.45, 0.15)),
            _pbr((0.0, 0.0, 0.0), emissive=(1.0, 1.0, 1.0))]
    p, n, i = plane_mesh(2.0, 2.0)
    strip = capi.DRAW_TRIANGLE_STRIP
    hp = math.pi / 2
    meshes = [
        MeshSpec(p, n, i, 0, translate(0, 0, 0), strip),                                  # floor
        MeshSpec(p, n, i, 0, translate(0, 2, 0) @ rotate("x", math.pi), strip),           # ceiling
        MeshSpec(p, n, i, 0, translate(0, 1, -1) @ rotate("x", hp), strip),               # back wall
        MeshSpec(p, n, i, 1, translate(-1, 1, 0) @ rotate("z", -hp), strip),              # left (red)
        MeshSpec(p, n, i, 2, translate(1, 1, 0) @ rotate("z", hp), strip),                # right (green)
    ]
    bp, bn, bi = box_mesh(0.6, 1.2, 0.6)
    meshes.append(MeshSpec(bp, bn, bi, 0, translate(-0.35, 0.6, -0.3) @ rotate("y", 0.3)))
    bp, bn, bi = box_mesh(0.6, 0.6, 0.6)
    meshes.append(MeshSpec(bp, bn, bi, 0, translate(0.35, 0.3, 0.35) @ rotate("y", -0.3)))
    ep, en, ei = plane_mesh(0.5, 0.5)
    meshes.append(MeshSpec(ep, en, ei, 3, translate(0, 1.995, 0) @ rotate("x", math.pi), strip))
    return SceneSpec("cornell", width, height, passes, max_depth, meshes, mats,
                     env=dict(enabled=False, solid=None, map=None, ev=0.0, theta=0.0),
                     point_lights=[dict(color=(1.0, 1.0, 1.0), intensity=683.0 * 0.5, position=(0.0, 1.9, 0.0))],
                     camera=dict(distance=5.4, phi=0.0, theta=0.0, target=(0.0, 1.0, 0.0), focal_length=50.0, fstop=None, focus_distance=5.4))


def sphere_grid(width=1920, height=1080, passes=32, max_depth=8, grid=8, u_slices=64, v_slices=32, seed=2, env_size=(2048, 1024), distance=None, name="sphere_grid"):
    """Config #2 ("Sponza-scale"): grid x grid lat-long spheres (2*u*v tris each) + floor + 2 walls (6 tris)."""
    rng = SplitMix64(seed)
    mats = [_pbr((0.6, 0.6, 0.6), rough=0.9, f0=0.5)]
    meshes = []
    p, n, i = plane_mesh(1.0, 1.0)
    ext = grid * 1.9 + 4.0
    strip = capi.DRAW_TRIANGLE_STRIP
    meshes.append(MeshSpec(p, n, i, 0, translate(0, -0.8, 0) @ scale(ext, 1, ext), strip))
    meshes.append(MeshSpec(p, n, i, 0, translate(0, ext / 4 - 0.8, -ext / 2) @ rotate("x", math.pi / 2) @ scale(ext, 1, ext / 2), strip))
    meshes.append(MeshSpec(p, n, i, 0, translate(-ext / 2, ext / 4 - 0.8, 0) @ rotate("z", -math.pi / 2) @ scale(ext / 2, 1, ext), strip))
    sp, sn, si = sphere_mesh(u_slices, v_slices, 0.8)
    for gz in range(grid):
        for gx in range(grid):
            rough = rng.uniform(0.05, 1.0)
            metal = 1.0 if rng.uniform() < 0.5 else 0.0
            base = (rng.uniform(0.2, 0.9), rng.uniform(0.2, 0.9), rng.uniform(0.2, 0.9))
            mats.append(_pbr(base, rough=rough, metal=metal, f0=0.5))
            x = (gx - (grid - 1) / 2) * 1.9
            z = (gz - (grid - 1) / 2) * 1.9
            meshes.append(MeshSpec(sp, sn, si, len(mats) - 1, translate(x, 0.0, z)))
    distance = 19.0 * grid / 8 if distance is None else distance            # OrbitCamera.h:22 default distance for the 8 x 8 grid, scaled with the grid
    return SceneSpec(name, width, height, passes, max_depth, meshes, mats,
                     env=dict(enabled=True, solid=None, map=studio_environment(*env_size), ev=0.0, theta=0.0),
                     camera=dict(distance=distance, phi=0.0, theta=0.6, target=(0.0, 0.0, 0.0), focal_length=50.0, fstop=None, focus_distance=distance))


def sphere_grid_1m(width=1920, height=1080, passes=32, max_depth=8):
    """The workload BASELINE.json's target is quoted on: ~1 M triangles at 1080p, 8 bounces — a 16 x 16 grid of the reference's own
    64 x 32 lat-long spheres (SphereMeshProvider.h:45-131, 4096 triangles each) = 1,048,576 + floor / walls (6) = 1,048,582 triangles."""
    return sphere_grid(width, height, passes, max_depth, grid=16, seed=6, name="sphere_grid_1m")


def material_grid(width=1920, height=1080, passes=32, max_depth=8, grid=11, u_slices=64, v_slices=32, seed=3):
    """Config #3: grid x grid spheres, odd cells glass, even cells clear-coated PBR, ground strip (2 tris);
    spot + directional light, solid grey environment, f/2.8 with hexagonal bokeh."""
    rng = SplitMix64(seed)
    mats = [_pbr((0.5, 0.5, 0.5), rough=0.8, f0=0.5)]
    meshes = []
    p, n, i = plane_mesh(1.0, 1.0)
    ext = grid * 1.4 + 4.0
    meshes.append(MeshSpec(p, n, i, 0, translate(0, -0.6, 0) @ scale(ext, 1, ext), capi.DRAW_TRIANGLE_STRIP))
    sp, sn, si = sphere_mesh(u_slices, v_slices, 0.6)
    for gz in range(grid):
        for gx in range(grid):
            cell = gz * grid + gx
            base = (rng.uniform(0.3, 0.95), rng.uniform(0.3, 0.95), rng.uniform(0.3, 0.95))
            if cell & 1:
                mats.append(_glass(base, rng.uniform(0.0, 0.5), rng.uniform(1.1, 2.0), rng.uniform(0.0, 1.0)))
            else:
                mats.append(_pbr(base, rough=rng.uniform(0.1, 0.9), metal=0.0, f0=0.5, cc=rng.uniform(0.0, 1.0), ccr=rng.uniform(0.0, 0.5)))
            x = (gx - (grid - 1) / 2) * 1.4
            z = (gz - (grid - 1) / 2) * 1.4
            meshes.append(MeshSpec(sp, sn, si, len(mats) - 1, translate(x, 0.0, z)))
    return SceneSpec("material_grid", width, height, passes, max_depth, meshes, mats,
                     env=dict(enabled=True, solid=(0.5, 0.5, 0.5), map=None, ev=0.0, theta=0.0),
                     dir_lights=[dict(color=(1.0, 0.95, 0.9), illuminance=683.0 * 2.0, theta=math.pi / 3, phi=0.4)],
                     spot_lights=[dict(color=(1.0, 1.0, 1.0), intensity=683.0 * 40.0, position=(0.0, 8.0, 0.0), theta=math.pi / 2, phi=0.0,
                                       inner=math.radians(20.0), outer=math.radians(40.0))],
                     camera=dict(distance=16.0, phi=0.0, theta=0.55, target=(0.0, 0.0, 0.0), focal_length=50.0, fstop=2.8, focus_distance=16.0),
                     bokeh=capi.BOKEH_HEXAGON)


def terrain(width=3840, height=2160, passes=16, max_depth=8, n=2048, blobs=400, seed=4, env_size=(2048, 1024)):
    """Config #4: n x n displaced heightfield (2(n-1)^2 tris) + `blobs` flattened 4096-triangle foliage spheres."""
    rng = SplitMix64(seed)
    mats = [_pbr((0.35, 0.3, 0.22), rough=0.95, f0=0.3), _pbr((0.15, 0.4, 0.12), rough=0.7, f0=0.5)]
    tp, tn, ti = terrain_mesh(n, 60.0, 6.0, seed)
    meshes = [MeshSpec(tp, tn, ti, 0)]
    sp, sn, si = sphere_mesh(64, 32, 1.0)
    grid = tp.reshape(n, n, 3)
    for _ in range(blobs):
        gx, gz = int(rng.uniform(0.05, 0.95) * (n - 1)), int(rng.uniform(0.05, 0.95) * (n - 1))
        c = grid[gz, gx]
        r = rng.uniform(0.25, 0.7)
        meshes.append(MeshSpec(sp, sn, si, 1, translate(float(c[0]), float(c[1]) + r * 0.8, float(c[2])) @ scale(r, r * 1.3, r)))
    return SceneSpec("terrain", width, height, passes, max_depth, meshes, mats,
                     env=dict(enabled=True, solid=None, map=studio_environment(*env_size), ev=0.0, theta=0.0),
                     dir_lights=[dict(color=(1.0, 0.96, 0.9), illuminance=683.0 * 3.0, theta=math.pi / 3.5, phi=0.7)],
                     camera=dict(distance=55.0, phi=0.3, theta=0.45, target=(0.0, 0.0, 0.0), focal_length=35.0, fstop=None, focus_distance=55.0))


def _checker(n, cells, a, b):
    """n x n RGBA checker of `cells` squares per side between colours a and b (4 floats each)."""
    g = (np.arange(n) * cells // n)
    m = ((g[:, None] + g[None, :]) & 1).astype(np.float32)[..., None]
    return (np.asarray(a, np.float32) * (1 - m) + np.asarray(b, np.float32) * m).astype(np.float32)


def _bump_normalmap(n, waves):
    """Tangent-space normal map of a sinusoidal egg-crate height field, encoded (n*0.5+0.5), alpha 1."""
    t = (np.arange(n, dtype=np.float32) + 0.5) / n
    x, y = np.meshgrid(t, t, indexing="xy")
    dx = 0.6 * np.cos(2 * np.pi * waves * x) * np.abs(np.sin(2 * np.pi * waves * y))
    dy = 0.6 * np.cos(2 * np.pi * waves * y) * np.abs(np.sin(2 * np.pi * waves * x))
    nrm = np.stack([-dx, -dy, np.ones_like(dx)], -1)
    nrm /= np.linalg.norm(nrm, axis=-1, keepdims=True)
    return np.concatenate([nrm * 0.5 + 0.5, np.ones_like(dx)[..., None]], -1).astype(np.float32)


def _gradient(n, c0, c1):
    t = ((np.arange(n, dtype=np.float32) + 0.5) / n)[None, :, None]
    img = np.asarray(c0, np.float32) * (1 - t) + np.asarray(c1, np.float32) * t
    return np.broadcast_to(img, (n, n, 4)).astype(np.float32).copy()


def textured_showcase(width=192, height=128, passes=4, max_depth=5, u_slices=32, v_slices=16):
    """Every texture slot, vertex colours, an alpha-masked (non-occluder) fence and textured glass in one small scene:
    the parity case for physicallyBased.rlsl:57-156 / glass.rlsl:145-176 (not one of BASELINE.json's bench configs)."""
    tex_base = _checker(64, 8, (0.9, 0.85, 0.8, 1.0), (0.25, 0.3, 0.6, 1.0))
    tex_mr = _gradient(32, (0.0, 0.15, 0.0, 1.0), (0.0, 1.0, 1.0, 1.0))                 # .g roughness, .b metallic
    tex_nm = _bump_normalmap(64, 4)
    tex_fence = _checker(32, 8, (0.8, 0.7, 0.3, 1.0), (0.0, 0.0, 0.0, 0.0))              # alpha 0 holes
    tex_em = _checker(16, 4, (0.0, 0.0, 0.0, 1.0), (0.9, 0.4, 0.1, 1.0))
    tex_cc = _gradient(16, (0.2, 0.0, 0.0, 1.0), (1.0, 0.0, 0.0, 1.0))
    tex_ccr = _gradient(16, (0.1, 0.0, 0.0, 1.0), (0.9, 0.0, 0.0, 1.0))
    tex_ccn = _bump_normalmap(32, 2)
    tex_glass = _checker(32, 6, (0.95, 0.95, 0.95, 1.0), (0.6, 0.9, 0.7, 1.0))
    mats = [
        _pbr((1.0, 1.0, 1.0), rough=0.9, f0=0.5, textures=dict(baseColorTexture=tex_base, metallicRoughnessTexture=tex_mr, normalmap=tex_nm)),      # 0 ground
        _pbr((1.0, 1.0, 1.0), rough=0.8, f0=0.5, alpha_mask=True, textures=dict(baseColorTexture=tex_fence)),                                        # 1 fence
        _pbr((0.7, 0.7, 0.75), rough=0.5, f0=0.5, cc=1.0, ccr=0.4,
             textures=dict(emissiveTexture=tex_em, clearCoatTexture=tex_cc, clearCoatRoughnessTexture=tex_ccr, clearCoatNormalmap=tex_ccn, normalmap=tex_nm)),   # 2
        _pbr((1.0, 1.0, 1.0), rough=0.6, metal=0.0, f0=0.5, vertex_colors=True, force_textures=True),                                                # 3 vertex colours + forced dummy textures
        _glass((1.0, 1.0, 1.0), 0.3, 1.5, 0.5, textures=dict(baseColorTexture=tex_glass, metallicRoughnessTexture=tex_mr, normalmap=tex_nm)),        # 4
        _pbr((0.6, 0.6, 0.6), rough=1.0, f0=0.5, double_sided=False),                                                                                # 5 single-sided back wall
    ]
    meshes = []
    p, n, i, uv = plane_mesh(1.0, 1.0, with_uv=True)
    strip = capi.DRAW_TRIANGLE_STRIP
    tx = np.tile(np.array([1.0, 0.0, 0.0], np.float32), (4, 1))
    bz = np.tile(np.array([0.0, 0.0, -1.0], np.float32), (4, 1))
    meshes.append(MeshSpec(p, n, i, 0, scale(8, 1, 8), strip, uvs=uv * 4.0, tangents=tx, bitangents=bz))
    meshes.append(MeshSpec(p, n, i, 1, translate(0.0, 1.0, 1.2) @ rotate("x", math.pi / 2) @ scale(5, 1, 2), strip, uvs=uv, tangents=tx, bitangents=bz))
    meshes.append(MeshSpec(p, n, i, 5, translate(0.0, 1.5, -3.0) @ rotate("x", math.pi / 2) @ scale(8, 1, 3), strip))
    sp, sn, si, suv, st, sb = sphere_mesh_uv(u_slices, v_slices, 0.7)
    meshes.append(MeshSpec(sp, sn, si, 2, translate(-1.7, 0.7, -0.3), uvs=suv, tangents=st, bitangents=sb))
    col = (sn * 0.5 + 0.5).astype(np.float32)
    meshes.append(MeshSpec(sp, sn, si, 3, translate(0.0, 0.7, -0.6), uvs=suv, colors=col))
    meshes.append(MeshSpec(sp, sn, si, 4, translate(1.7, 0.7, -0.3) @ scale(1.0, 1.1, 1.0), uvs=suv, tangents=st, bitangents=sb))
    return SceneSpec("textured_showcase", width, height, passes, max_depth, meshes, mats,
                     env=dict(enabled=True, solid=(0.35, 0.4, 0.5), map=None, ev=0.0, theta=0.0),
                     dir_lights=[dict(color=(1.0, 0.95, 0.9), illuminance=683.0 * 1.5, theta=math.pi / 3, phi=0.6)],
                     point_lights=[dict(color=(1.0, 0.9, 0.8), intensity=683.0 * 1.2, position=(0.0, 2.2, 3.0))],
                     camera=dict(distance=6.5, phi=0.15, theta=0.35, target=(0.0, 0.6, 0.0), focal_length=35.0, fstop=None, focus_distance=6.5))


def build(name: str, **kw) -> SceneSpec:
    return {"cornell": cornell_box, "sphere_grid": sphere_grid, "material_grid": material_grid, "terrain": terrain,
            "textured_showcase": textured_showcase}[name](**kw)


# ---- assets from disk (Scene::loadFromDisk, Scene.cpp:14-27 -> hrhImportScene) -----------------------------------------
_SLOT_FIELDS = ["baseColorTexture", "metallicRoughnessTexture", "emissiveTexture", "normalmap", "clearCoatTexture", "clearCoatRoughnessTexture", "clearCoatNormalmap"]
_SLOT_LINEAR = [True, False, True, False, False, False, False]     # convertToLinear per slot (AssimpMeshProvider.cpp:448-505)


def import_scene(path, convert_to_meters=False, width=512, height=512, passes=16, max_depth=5, env=None, camera=None, load_textures=True, ground_plane=False) -> SceneSpec:
    """OBJ / glTF asset -> SceneSpec through the C++ importer (host_import.cpp) and texture loader (host_image.cpp).
    Camera framing follows HeatrayRenderer.cpp:302-308: target = AABB centre, distance = 3 x AABB diagonal, focus on the target."""
    import ctypes as C
    lib = capi.load_host_library()
    h = C.c_void_p()
    rc = lib.hrhImportScene(str(path).encode(), int(convert_to_meters), C.byref(h))
    if rc != capi.HRC_OK:
        raise capi.HrcError(rc, (lib.hrhImportLastError() or b"").decode())
    if ground_plane:                                                   # the viewer's "Add Ground Plane" (HeatrayRenderer.cpp:1167-1193)
        info = capi.HrhSceneInfo()
        lib.hrhSceneGetInfo(h, C.byref(info))
        lib.hrhSceneAddGroundPlane(h, info.aabbMin, info.aabbMax, None)
    return _spec_from_scene_handle(lib, h, Path(str(path)).stem, width, height, passes, max_depth, env, camera, load_textures)


PRESET_SCENES = ("Editable PBR Material", "Editable Glass Material", "Multi-Material", "Sphere Array")


def preset_scene(name, width=512, height=512, passes=16, max_depth=5, env=None, camera=None) -> SceneSpec:
    """One of the reference's built-in scenes (HeatrayRenderer::changeScene, HeatrayRenderer.cpp:123-286) through hrhBuildPresetScene.
    The reference does not re-frame the camera for them: the default OrbitCamera applies (distance 19, OrbitCamera.h:22-25), focus on it."""
    import ctypes as C
    lib = capi.load_host_library()
    h = C.c_void_p()
    rc = lib.hrhBuildPresetScene(str(name).encode(), C.byref(h))
    if rc != capi.HRC_OK:
        raise capi.HrcError(rc, (lib.hrhImportLastError() or b"").decode())
    cam = dict(distance=19.0, phi=0.0, theta=0.0, target=(0.0, 0.0, 0.0), focal_length=50.0, fstop=None, focus_distance=19.0)
    cam.update(camera or {})
    return _spec_from_scene_handle(lib, h, str(name), width, height, passes, max_depth, env, cam, False)


def _spec_from_scene_handle(lib, h, name, width, height, passes, max_depth, env, camera, load_textures) -> SceneSpec:
    """HrhScene -> SceneSpec; frees the handle."""
    import ctypes as C
    try:
        info = capi.HrhSceneInfo()
        lib.hrhSceneGetInfo(h, C.byref(info))
        tex_cache = {}

        def texture(pth, linear):
            key = (pth, linear)
            if key not in tex_cache:
                try:
                    px, fmt, _, _ = capi.load_texture(pth, True, linear)
                except capi.HrcError:
                    tex_cache[key] = None                      # the reference logs and carries on without the texture (TextureLoader.cpp:29-33)
                else:
                    if px.dtype == np.uint8:
                        px = px.astype(np.float32) / np.float32(255.0)
                    elif px.shape[2] == 3:
                        px = np.concatenate([px, np.ones(px.shape[:2] + (1,), np.float32)], -1)
                    elif px.shape[2] == 1:
                        px = np.concatenate([px, px, px, np.ones_like(px)], -1)
                    tex_cache[key] = np.ascontiguousarray(px, np.float32)
            return tex_cache[key]
        materials = []
        for i in range(info.materialCount):
            m = capi.HrhSceneMaterial()
            lib.hrhSceneGetMaterial(h, i, C.byref(m))
            textures = {}
            if load_textures:
                for s, fld in enumerate(_SLOT_FIELDS):
                    if m.texturePath[s]:
                        t = texture(m.texturePath[s].decode(), _SLOT_LINEAR[s])
                        if t is not None:
                            textures[fld] = t
            if m.type == capi.MATERIAL_GLASS:
                g = m.glass
                materials.append(_glass(tuple(g.baseColor), g.roughness, g.ior, g.density,
                                        {k: v for k, v in textures.items() if k in ("baseColorTexture", "normalmap", "metallicRoughnessTexture")}))
            else:
                p = m.pbr
                materials.append(_pbr(tuple(p.baseColor), p.roughness, p.metallic, p.specularF0, tuple(p.emissiveColor), p.clearCoat, p.clearCoatRoughness,
                                      bool(p.doubleSided), bool(p.alphaMask), bool(p.enableVertexColors), False, textures))
        meshes = []
        for i in range(info.meshCount):
            sm = capi.HrhSceneMesh()
            lib.hrhSceneGetMesh(h, i, C.byref(sm))
            nv, ni = sm.vertexCount, sm.indexCount

            def arr(ptr, comps):
                return np.ctypeslib.as_array(ptr, shape=(nv, comps)).copy() if ptr else None
            idx = np.ctypeslib.as_array(sm.indices, shape=(ni,)).copy().astype(np.uint32)
            xf = np.asarray(list(sm.localTransform), np.float32).reshape(4, 4).T          # column-major -> mathematical (row-major) matrix
            meshes.append(MeshSpec(arr(sm.positions, 3), arr(sm.normals, 3), idx, sm.materialIndex, xf, capi.DRAW_TRIANGLES,
                                   uvs=arr(sm.texCoords, 2), tangents=arr(sm.tangents, 3), bitangents=arr(sm.bitangents, 3), colors=arr(sm.colors, 3)))
        dl, pl, sl = [], [], []
        for i in range(info.lightCount):
            L = capi.HrhSceneLight()
            lib.hrhSceneGetLight(h, i, C.byref(L))
            if L.type == 0:
                dl.append(dict(color=tuple(L.color), illuminance=L.intensity, theta=L.theta, phi=L.phi))
            elif L.type == 1:
                pl.append(dict(color=tuple(L.color), intensity=L.intensity, position=tuple(L.position)))
            else:
                sl.append(dict(color=tuple(L.color), intensity=L.intensity, position=tuple(L.position), theta=L.theta, phi=L.phi, inner=L.innerAngle, outer=L.outerAngle))
        lo, hi = np.asarray(list(info.aabbMin), np.float32), np.asarray(list(info.aabbMax), np.float32)
        centre, radius = (lo + hi) * 0.5, float(np.linalg.norm(hi - lo))                   # AABB.h:34-44 (radius() is the full diagonal)
        cam = dict(distance=radius * 3.0, phi=0.0, theta=0.0, target=tuple(float(v) for v in centre), focal_length=50.0, fstop=None, focus_distance=radius * 3.0)
        cam.update(camera or {})
        return SceneSpec(name, width, height, passes, max_depth, meshes, materials,
                         env=env or dict(enabled=True, solid=(0.5, 0.5, 0.5), map=None, ev=0.0, theta=0.0),
                         dir_lights=dl[:5], point_lights=pl[:5], spot_lights=sl[:5], camera=cam)
    finally:
        lib.hrhFreeScene(h)


def write_obj(spec: SceneSpec, path) -> Path:
    """Writes the meshes of `spec`, flattened to world space, as Wavefront OBJ + MTL (BASELINE config #1 names a Cornell-box OBJ).
    Strips are expanded GL-style; normals are written transformed by mat3(worldFromEntity) like vertex.rlsl:28. PBR parameters go out
    through the keys Assimp's OBJ importer understands (Kd, Ke, Pr, Pm, Pc, Pcr, Ni with ((1-Ni)/(1+Ni))^2 = specularF0)."""
    path = Path(path)
    mtl = path.with_suffix(".mtl")
    with open(mtl, "w") as f:
        for i, m in enumerate(spec.materials):
            p = m.params
            f.write(f"newmtl m{i}\n")
            f.write("Kd %.9g %.9g %.9g\nKe %.9g %.9g %.9g\n" % (*p["baseColor"], *p.get("emissiveColor", (0, 0, 0))))
            s = math.sqrt(max(p.get("specularF0", 0.5), 0.0))
            f.write("Pr %.9g\nPm %.9g\nPc %.9g\nPcr %.9g\nNi %.9g\n\n" % (p.get("roughness", 1.0), p.get("metallic", 0.0), p.get("clearCoat", 0.0),
                                                                       p.get("clearCoatRoughness", 0.0), (1 + s) / (1 - s) if s < 1 else 1e9))
    with open(path, "w") as f:
        f.write(f"mtllib {mtl.name}\n")
        base = 1
        for k, ms in enumerate(spec.meshes):
            M = np.asarray(ms.transform, np.float64)
            P = np.asarray(ms.positions, np.float64).reshape(-1, 3) @ M[:3, :3].T + M[:3, 3]
            N = np.asarray(ms.normals, np.float64).reshape(-1, 3) @ M[:3, :3].T
            f.write(f"o mesh{k}\nusemtl m{ms.material}\n")
            for v in P.astype(np.float32):
                f.write("v %.9g %.9g %.9g\n" % tuple(v))
            for v in N.astype(np.float32):
                f.write("vn %.9g %.9g %.9g\n" % tuple(v))
            idx = np.asarray(ms.indices).reshape(-1)
            if ms.draw_mode == capi.DRAW_TRIANGLES:
                tris = idx.reshape(-1, 3)
            else:
                tris = np.asarray([(idx[j], idx[j + 1], idx[j + 2]) if j % 2 == 0 else (idx[j + 1], idx[j], idx[j + 2]) for j in range(len(idx) - 2)])
            if np.linalg.det(M[:3, :3]) < 0:
                tris = tris[:, ::-1]
            for a, b, c in tris:
                f.write("f %d//%d %d//%d %d//%d\n" % (a + base, a + base, b + base, b + base, c + base, c + base))
            base += len(P)
    return path

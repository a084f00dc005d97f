"""TEST INFRASTRUCTURE: drives the shader-level probes of oracle/shade_probe.h.

The same randomised shader invocations (frame shader, PBR / glass material shaders per #define set, the four light shaders) run
  (a) through the reference's RLSL text compiled as C++  -> oracle/_ref/libref_shaders.so (ref_probe/ref_shaders.cpp), and
  (b) through the oracle's restatement                    -> orcProbeShade / orcProbeMiss / orcProbeRaygen (oracle.cpp),
and are compared field by field. Used by tests/test_shader_pinning.py and by oracle/ref_probe/make_shader_golden.py, which
stores (a)'s answers for a bounded set of cases under tests/golden/ so the comparison also runs where the reference is absent.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from heatray_b200 import capi

RAY = np.dtype([("origin", np.float32, 3), ("maxT", np.float32), ("direction", np.float32, 3), ("extraT", np.float32),
                ("weight", np.float32, 3), ("depth", np.int32), ("sequenceID", np.int32), ("sequenceIndexOffset", np.int32),
                ("occlusionTest", np.int32), ("missKind", np.int32), ("missIndex", np.int32), ("pad", np.int32)])
CASE = np.dtype([("in", RAY), ("P", np.float32, 3), ("T", np.float32), ("Ng", np.float32, 3), ("frontFacing", np.int32),
                 ("normal", np.float32, 3), ("material", np.int32), ("texCoord", np.float32, 2), ("pad0", np.float32, 2),
                 ("tangent", np.float32, 3), ("pad1", np.float32), ("bitangent", np.float32, 3), ("pad2", np.float32),
                 ("vertexColor", np.float32, 3), ("pad3", np.float32)])
RESULT = np.dtype([("rayCount", np.int32), ("accumCount", np.int32), ("rays", RAY, 2), ("accumSum", np.float32, 4)])
assert RAY.itemsize == 72 and CASE.itemsize == 72 + 16 * 7 and RESULT.itemsize == 8 + 144 + 16

DS, AM, BCT, MRT, EMT, NM, CCT, CCRT, CCNM, VC = (1 << i for i in range(10))
ANY_TEXTURE = BCT | MRT | EMT | NM | CCT | CCRT | CCNM
PBR, GLASS = 0, 1


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def unit(v):
    return v / np.linalg.norm(v, axis=-1, keepdims=True)


@dataclass
class World:
    """Everything a shader invocation reads besides the hit: lights, QMC tables, globals, textures."""
    env: capi.EnvLight
    dirs: list
    points: list
    spots: list
    seq: np.ndarray           # (ns, len, 2)
    aperture: np.ndarray
    globals_: capi.Globals
    textures: dict = field(default_factory=dict)     # handle -> (rgba (h, w, 4) float32, wrapS, wrapT)


def random_world(rng, lights: str, max_depth: int, debug: str | None = None, ns=16, length=64, env_texture=True) -> World:
    def col(lo=0.1, hi=3.0):
        return rng.uniform(lo, hi, 3).astype(np.float32)
    dirs, points, spots = [], [], []
    env_on = "e" in lights
    counts = {"d": lights.count("d"), "p": lights.count("p"), "s": lights.count("s")}
    for _ in range(counts["d"]):
        dirs.append(capi.DirLight((C.c_float * 3)(*unit(rng.normal(size=3))), (C.c_float * 3)(*col())))
    for _ in range(counts["p"]):
        points.append(capi.PointLight((C.c_float * 3)(*rng.uniform(-8, 8, 3)), (C.c_float * 3)(*col(1, 60))))
    for _ in range(counts["s"]):
        inner = rng.uniform(0.05, 0.6); outer = inner + rng.uniform(0.05, 0.6)
        spots.append(capi.SpotLight((C.c_float * 3)(*rng.uniform(-8, 8, 3)), (C.c_float * 3)(*unit(rng.normal(size=3))), (C.c_float * 3)(*col(1, 60)),
                                    (C.c_float * 2)(float(np.cos(np.float32(inner))), float(np.cos(np.float32(outer))))))
    textures = {}
    env_tex = -1
    if env_on and env_texture:
        textures[0] = (rng.uniform(0, 2.5, (8, 16, 4)).astype(np.float32), capi.WRAP_REPEAT, capi.WRAP_CLAMP)
        env_tex = 0
    env = capi.EnvLight(1 if env_on else 0, env_tex, float(2.0 ** rng.uniform(-2, 1)), float(rng.uniform(0, 6.2)))
    seq = rng.uniform(0, 1, (ns, length, 2)).astype(np.float32)
    seq[rng.integers(0, ns, 8), rng.integers(0, length, 8), rng.integers(0, 2, 8)] = 0.0      # exact 0 and 1 do occur (quirk B-Q4, B-Q11)
    seq[rng.integers(0, ns, 8), rng.integers(0, length, 8), rng.integers(0, 2, 8)] = 1.0
    aperture = rng.uniform(0, 1, (ns, length, 2)).astype(np.float32)
    g = capi.Globals()
    g.maxRayDepth = max_depth; g.maxChannelValue = float(np.float32(np.pi))
    if debug in ("showNANs", "showInf"):
        g.enableAccumulatorVisualizer = 1; setattr(g, debug, 1)
    elif debug:
        g.enableDebugVisualizer = 1; setattr(g, debug, 1)
    return World(env, dirs, points, spots, seq, aperture, g, textures)


def add_material_textures(rng, world: World, first_handle=1, lut=None):
    """Eight small random RGBA textures (handles first_handle..+6) and the multiscatter LUT (handle first_handle+7)."""
    for k in range(7):
        t = rng.uniform(0, 1, (6 + k, 5 + 2 * k, 4)).astype(np.float32)
        if k == 0:
            t[..., 3] = np.where(rng.uniform(size=t.shape[:2]) < 0.5, 1.0, t[..., 3])            # base colour: half the texels opaque (alpha mask)
        world.textures[first_handle + k] = (t, capi.WRAP_REPEAT, capi.WRAP_REPEAT)
    if lut is None:
        lut = rng.uniform(0, 1.5, (16, 16)).astype(np.float32)
    rgba = np.zeros(lut.shape + (4,), np.float32); rgba[..., 0] = lut; rgba[..., 3] = 1.0
    world.textures[first_handle + 7] = (rgba, capi.WRAP_CLAMP, capi.WRAP_CLAMP)
    return first_handle


def random_pbr_params(rng, flags: int, tex0: int) -> capi.PbrParams:
    f32 = np.float32
    p = capi.PbrParams()
    handles = [tex0 + k if flags & bit else -1 for k, bit in enumerate((BCT, MRT, EMT, NM, CCT, CCRT, CCNM))]
    (p.baseColorTexture, p.metallicRoughnessTexture, p.emissiveTexture, p.normalmap, p.clearCoatTexture, p.clearCoatRoughnessTexture,
     p.clearCoatNormalmap) = handles
    p.multiscatterLUT = tex0 + 7
    p.baseColor[:] = rng.uniform(0, 1, 3).tolist()
    p.emissiveColor[:] = (rng.uniform(0, 4, 3) if rng.uniform() < 0.3 else np.zeros(3)).tolist()
    p.metallic = float(rng.choice([0.0, 1.0, rng.uniform()]))
    r = f32(max(rng.uniform(0, 1), 0.01)); p.roughness = float(r); p.roughnessAlpha = float(r * r)
    p.specularF0 = float(f32(rng.uniform(0, 1)) * f32(0.08))
    p.clearCoat = float(f32(rng.choice([0.0, rng.uniform()])) * f32(0.2))
    cr = f32(max(rng.uniform(0, 1), 0.01)); p.clearCoatRoughness = float(cr); p.clearCoatRoughnessAlpha = float(cr * cr)
    return p


def random_glass_params(rng, flags: int, tex0: int) -> capi.GlassParams:
    f32 = np.float32
    p = capi.GlassParams()
    p.baseColorTexture = tex0 if flags & BCT else -1
    p.metallicRoughnessTexture = tex0 + 1 if flags & MRT else -1
    p.normalmap = tex0 + 3 if flags & NM else -1
    p.baseColor[:] = rng.uniform(0, 1, 3).tolist()
    r = f32(max(rng.uniform(0, 0.6), 0.01)); p.roughness = float(r); p.roughnessAlpha = float(r * r)
    p.ior = float(rng.uniform(1.05, 2.2)); p.density = float(rng.uniform(0, 2))
    p.specularF0 = float(((1 - p.ior) / (1 + p.ior)) ** 2)
    return p


def random_cases(rng, n: int, material: int, world: World, occlusion_fraction=0.0) -> np.ndarray:
    c = np.zeros(n, CASE)
    N = unit(rng.normal(size=(n, 3)))
    # incoming direction: mostly arriving against the normal, a share grazing or from behind
    d = unit(rng.normal(size=(n, 3)))
    flip = (np.einsum("ij,ij->i", d, N) > 0) & (rng.uniform(size=n) < 0.9)
    d[flip] = -d[flip]
    c["in"]["direction"] = d
    c["in"]["origin"] = rng.uniform(-6, 6, (n, 3))
    c["in"]["maxT"] = np.inf
    c["in"]["weight"] = rng.uniform(0, 1.6, (n, 3)) * rng.choice([1.0, 1.0, 0.2, 0.02], (n, 1))
    c["in"]["depth"] = rng.integers(0, 12, n)
    ns, length = world.seq.shape[:2]
    c["in"]["sequenceID"] = rng.integers(0, ns, n)
    c["in"]["sequenceIndexOffset"] = rng.integers(0, length, n)
    c["in"]["missKind"] = np.where(rng.uniform(size=n) < 0.5, 1 if world.env.enabled else 0, 0)
    occ = rng.uniform(size=n) < occlusion_fraction
    c["in"]["occlusionTest"] = occ
    c["in"]["maxT"][occ] = rng.uniform(0.5, 30, occ.sum())
    c["in"]["extraT"][occ] = rng.uniform(0, 5, occ.sum()) * (rng.uniform(size=occ.sum()) < 0.5)
    c["in"]["missKind"][occ] = 3; c["in"]["missIndex"][occ] = 0
    c["T"] = rng.uniform(0.01, 20, n)
    c["T"][occ] = c["in"]["maxT"][occ] * rng.uniform(0.05, 0.95, occ.sum())
    c["P"] = c["in"]["origin"] + c["in"]["direction"] * c["T"][:, None]
    c["Ng"] = N
    c["frontFacing"] = rng.uniform(size=n) < 0.75
    c["normal"] = unit(N + rng.normal(scale=0.15, size=(n, 3))) * rng.uniform(0.5, 2.0, (n, 1))       # interpolated normals are not unit length
    t = unit(np.cross(N, unit(rng.normal(size=(n, 3)))))
    c["tangent"] = t * rng.uniform(0.5, 1.5, (n, 1))
    c["bitangent"] = np.cross(N, t) * rng.uniform(0.5, 1.5, (n, 1))
    c["texCoord"] = rng.uniform(-1, 2, (n, 2))
    c["vertexColor"] = rng.uniform(0, 1, (n, 3))
    c["material"] = material
    return c


# ---- the two sides -----------------------------------------------------------------------------------------------------
class RefSide:
    def __init__(self, lib: C.CDLL):
        self.lib = lib
        lib.ref_set_pixel_offsets.argtypes = [C.c_void_p, C.c_size_t]
        lib.ref_shade.argtypes = [C.c_int, C.c_uint32, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        lib.ref_variant_flags.restype = C.c_uint32

    def variants(self, mtype):
        return [int(self.lib.ref_variant_flags(mtype, i)) for i in range(self.lib.ref_variant_count(mtype))]

    def set_world(self, w: World):
        L = self.lib
        L.ref_set_globals(C.byref(w.globals_))
        da = (capi.DirLight * 5)(*w.dirs); pa = (capi.PointLight * 5)(*w.points); sa = (capi.SpotLight * 5)(*w.spots)
        L.ref_set_lights(C.byref(w.env), len(w.dirs), da, len(w.points), pa, len(w.spots), sa)
        seq = np.ascontiguousarray(w.seq, np.float32); ap = np.ascontiguousarray(w.aperture, np.float32)
        L.ref_set_sequences(seq.shape[0], seq.shape[1], _p(seq), _p(ap))
        L.ref_clear_textures()
        for h, (rgba, ws, wt) in w.textures.items():
            rgba = np.ascontiguousarray(rgba, np.float32)
            L.ref_set_texture(h, rgba.shape[1], rgba.shape[0], _p(rgba), ws, wt)

    def shade(self, mtype, flags, params, sample_index, cases):
        res = np.zeros(len(cases), RESULT)
        rc = self.lib.ref_shade(mtype, flags, C.byref(params), sample_index, len(cases), _p(cases), _p(res))
        if rc != 0:
            raise KeyError(f"material variant {mtype}/{flags:#x} is not compiled into libref_shaders.so")
        return res

    def miss(self, rays):
        res = np.zeros(len(rays), RESULT)
        self.lib.ref_miss(len(rays), _p(rays), _p(res))
        return res

    def raygen(self, frame: capi.FrameParams, w, h, xy):
        xy = np.ascontiguousarray(xy, np.int32)
        res = np.zeros(len(xy), RESULT)
        self.lib.ref_raygen(C.byref(frame), w, h, len(xy), _p(xy), _p(res))
        return res


class OracleSide:
    """The oracle context is set up through the same public C-ABI calls a renderer makes (textures, materials, lights, tables)."""
    def __init__(self, backend: capi.Backend):
        self.b = backend
        self.tex = {}

    def set_world(self, w: World):
        b = self.b
        b.set_globals(w.globals_)
        self.tex = {}
        for h in sorted(w.textures):
            rgba, ws, wt = w.textures[h]
            self.tex[h] = b.create_texture(np.ascontiguousarray(rgba, np.float32), capi.TEX_RGBA32F, ws, wt)
        env = w.env
        b.set_environment(bool(env.enabled), self.tex.get(env.texture, -1), env.exposureCompensation, env.thetaRotation)
        b.set_directional_lights(w.dirs); b.set_point_lights(w.points); b.set_spot_lights(w.spots)
        b.set_sequences(w.seq, w.aperture)

    def material(self, mtype, flags, params):
        q = type(params)()
        C.memmove(C.byref(q), C.byref(params), C.sizeof(params))
        for name, _ in q._fields_:
            v = getattr(q, name)
            if isinstance(v, int) and (name.endswith("Texture") or name.endswith("map") or name == "multiscatterLUT"):
                setattr(q, name, self.tex.get(v, -1))
        return self.b.create_material(mtype, flags, q)

    def shade(self, sample_index, cases):
        res = np.zeros(len(cases), RESULT)
        rc = self.b.lib.orcProbeShade(self.b.ctx, C.c_int32(sample_index), C.c_int32(len(cases)), _p(cases), _p(res))
        assert rc == 0, rc
        return res

    def miss(self, rays):
        res = np.zeros(len(rays), RESULT)
        rc = self.b.lib.orcProbeMiss(self.b.ctx, C.c_int32(len(rays)), _p(rays), _p(res))
        assert rc == 0, rc
        return res

    def raygen(self, frame: capi.FrameParams, xy):
        xy = np.ascontiguousarray(xy, np.int32)
        res = np.zeros(len(xy), RESULT)
        rc = self.b.lib.orcProbeRaygen(self.b.ctx, C.byref(frame), C.c_int32(len(xy)), _p(xy), _p(res))
        assert rc == 0, rc
        return res


# ---- comparison -----------------------------------------------------------------------------------------------------------
def ulp_diff(a, b):
    """Distance in units in the last place between float32 arrays (0 for equal values incl. +-0 and inf == inf; huge for NaN vs number)."""
    a = np.asarray(a, np.float32); b = np.asarray(b, np.float32)
    ia = a.view(np.int32).astype(np.int64); ib = b.view(np.int32).astype(np.int64)
    ia = np.where(ia < 0, -(ia & 0x7fffffff), ia); ib = np.where(ib < 0, -(ib & 0x7fffffff), ib)
    d = np.abs(ia - ib)
    both_nan = np.isnan(a) & np.isnan(b)
    return np.where(both_nan, 0, d)


def normalize_emitted(res: np.ndarray) -> np.ndarray:
    """"OpenRL auto-normalizes the direction for emitted rays" (perspective.rlsl:86; policy A-11): d * (1 / sqrt(dot(d, d))) in fp32,
    the operation order of glm::normalize. Applied to the reference side's frame-shader rays before they are compared."""
    out = res.copy()
    d = out["rays"]["direction"].astype(np.float32)
    f = np.float32
    dd = (d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]).astype(f) + (d[..., 2] * d[..., 2]).astype(f)
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = (f(1.0) / np.sqrt(dd, dtype=f)).astype(f)
        live = np.arange(2)[None, :] < out["rayCount"][:, None]
        out["rays"]["direction"] = np.where(live[..., None], d * inv[..., None], d)
    return out


def compare(ref: np.ndarray, orc: np.ndarray, max_ulp=2, compare_origin=False):
    """Returns (ok mask per case, worst ulp distance over the float fields of the agreeing cases). Discrete fields must be equal."""
    n = len(ref)
    ok = ref["rayCount"] == orc["rayCount"]
    worst = np.zeros(n, np.int64)
    for k in range(2):
        live = ref["rayCount"] > k
        r, o = ref["rays"][:, k], orc["rays"][:, k]
        same = np.ones(n, bool)
        for f in ("depth", "sequenceID", "sequenceIndexOffset", "occlusionTest", "missKind", "missIndex"):
            same &= r[f] == o[f]
        u = np.maximum.reduce([ulp_diff(r["direction"], o["direction"]).max(-1), ulp_diff(r["weight"], o["weight"]).max(-1),
                               ulp_diff(r["maxT"], o["maxT"]), ulp_diff(r["extraT"], o["extraT"])])
        if compare_origin:
            u = np.maximum(u, ulp_diff(r["origin"], o["origin"]).max(-1))
        ok &= ~live | (same & (u <= max_ulp))
        worst = np.maximum(worst, np.where(live & same, u, 0))
    ua = ulp_diff(ref["accumSum"], orc["accumSum"]).max(-1)
    ok &= ua <= max_ulp
    worst = np.maximum(worst, ua)
    return ok, worst


# ---- the fixed suite behind tests/golden/shader_ref.npz -----------------------------------------------------------------------
DEBUG_MODES = [None, "showGeometricNormals", "showUVs", "showTangents", "showBitangents", "showNormalmap", "showFinalNormals", "showBaseColor",
               "showRoughness", "showMetallic", "showEmissive", "showClearcoat", "showClearcoatRoughness", "showClearcoatNormalmap", "showShader",
               "showNANs", "showInf"]
LIGHT_SETS = ("e", "ed", "dpps", "", "edddddpppppsssss", "ep", "es")
PBR_VARIANTS = [0, DS, AM | BCT, DS | AM | BCT, DS | BCT, DS | MRT, DS | EMT, DS | NM, DS | CCT, DS | CCRT, DS | CCNM, DS | VC, VC,
                DS | BCT | NM, DS | BCT | MRT | CCT | CCRT, DS | BCT | MRT | NM, BCT | MRT | NM | VC,
                DS | BCT | MRT | EMT | NM | CCT | CCRT | CCNM | VC, DS | AM | BCT | MRT | EMT | NM | CCT | CCRT | CCNM | VC]     # = rlsl_to_cpp.py
GLASS_VARIANTS = [0, BCT, MRT, NM, BCT | MRT, BCT | MRT | NM]


def material_suite(seed: int, cases_per_world: int, worlds=LIGHT_SETS):
    """Deterministic list of (mtype, flags, params, world, cases-without-material, sampleIndex): every compiled #define set of both
    material programs under every light set; the second light set of each variant also cycles through the debug / accumulator visualisers.
    Exact zeros in the QMC table are only planted where the environment light exists (DESIGN.md deviation D-1: with no environment
    a light-pick sample of exactly 0.0 makes the shader emit a second, non-occlusion ray the wavefront does not carry)."""
    rng = np.random.default_rng(seed)
    k = 0
    for mtype, variants in ((PBR, PBR_VARIANTS), (GLASS, GLASS_VARIANTS)):
        for flags in variants:
            for li, lights in enumerate(worlds):
                debug = DEBUG_MODES[k % len(DEBUG_MODES)] if li == 1 else None
                k += li == 1
                w = random_world(rng, lights, int(rng.integers(3, 11)), debug=debug)
                if not w.env.enabled:
                    w.seq[w.seq == 0.0] = 0.5
                tex0 = add_material_textures(rng, w)
                params = random_pbr_params(rng, flags, tex0) if mtype == PBR else random_glass_params(rng, flags, tex0)
                cases = random_cases(rng, cases_per_world, 0, w, occlusion_fraction=0.2 if (flags & AM) else 0.0)
                yield mtype, flags, params, w, cases, int(rng.integers(0, w.seq.shape[1]))


def miss_suite(seed: int, n: int):
    rng = np.random.default_rng(seed)
    for trial, debug in enumerate((None, None, "showNANs", "showInf")):
        w = random_world(rng, "edddddpppppsssss", 8, debug=debug)
        rays = np.zeros(n, RAY)
        rays["direction"] = unit(rng.normal(size=(n, 3)))
        rays["weight"] = rng.uniform(0, 2, (n, 3))
        rays["maxT"] = rng.uniform(0.1, 40, n); rays["extraT"] = rng.uniform(0, 5, n) * (rng.uniform(size=n) < 0.3)
        rays["missKind"] = rng.integers(1, 5, n); rays["missIndex"] = np.where(rays["missKind"] == 1, 0, rng.integers(0, 5, n))
        if debug:
            rays["weight"][::7, 0] = np.nan if debug == "showNANs" else np.inf
        yield w, rays


def input_digest(*arrays) -> str:
    import hashlib
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


# ---- vertex.rlsl (the geometry flattening of commitScene) ---------------------------------------------------------------------------
VERTEX_CASE = np.dtype([("worldFromEntity", "<f4", (16,)), ("position", "<f4", (3,)), ("normal", "<f4", (3,)), ("texCoord", "<f4", (2,)),
                        ("tangent", "<f4", (3,)), ("bitangent", "<f4", (3,)), ("color", "<f4", (3,))])
VERTEX_RESULT = np.dtype([("position", "<f4", (4,)), ("normal", "<f4", (3,)), ("texCoord", "<f4", (2,)), ("tangent", "<f4", (3,)),
                          ("bitangent", "<f4", (3,)), ("color", "<f4", (3,))])
VERTEX_VARIANTS = [0, 1, 3, 4, 5, 7]          # bit 1 HAS_TEXTURES, 2 USE_TANGENT_SPACE, 4 VERTEX_COLORS (oracle/ref_probe/rlsl_to_cpp.py)


def vertex_cases(seed: int, n: int) -> np.ndarray:
    """Random vertices under random worldFromEntity matrices: rotations, non-uniform and negative scales, shears, large translations
    (column-major, affine: last row 0 0 0 1 as every transform the reference builds)."""
    rng = np.random.default_rng(seed)
    c = np.zeros(n, VERTEX_CASE)
    A = rng.normal(size=(n, 3, 3)) * rng.choice([0.01, 1.0, 100.0], size=(n, 1, 1))
    q, _ = np.linalg.qr(rng.normal(size=(n, 3, 3)))
    pure = rng.random(n) < 0.3
    A[pure] = q[pure]                                             # some pure rotations / reflections
    M = np.zeros((n, 4, 4))
    M[:, :3, :3] = A
    M[:, :3, 3] = rng.normal(size=(n, 3)) * rng.choice([0.0, 1.0, 1000.0], size=(n, 1))
    M[:, 3, 3] = 1.0
    c["worldFromEntity"] = M.transpose(0, 2, 1).reshape(n, 16).astype(np.float32)      # column-major
    c["position"] = (rng.normal(size=(n, 3)) * rng.choice([1e-3, 1.0, 1e3], size=(n, 1))).astype(np.float32)
    for k in ("normal", "tangent", "bitangent"):
        v = rng.normal(size=(n, 3)); c[k] = (v / np.linalg.norm(v, axis=1, keepdims=True)).astype(np.float32)
    c["texCoord"] = rng.uniform(-2, 3, (n, 2)).astype(np.float32)
    c["color"] = rng.random((n, 3)).astype(np.float32)
    return c


def ref_vertex(lib, flags: int, cases: np.ndarray) -> np.ndarray:
    out = np.zeros(len(cases), VERTEX_RESULT)
    lib.ref_vertex.argtypes = [C.c_uint32, C.c_int, C.c_void_p, C.c_void_p]
    assert lib.ref_vertex(flags, len(cases), _p(cases), _p(out)) == 0, f"vertex.rlsl was not compiled for #define set {flags}"
    return out


def oracle_vertex(lib, flags: int, cases: np.ndarray) -> np.ndarray:
    out = np.zeros(len(cases), VERTEX_RESULT)
    lib.orcProbeVertex.argtypes = [C.c_uint32, C.c_int32, C.c_void_p, C.c_void_p]
    assert lib.orcProbeVertex(flags, len(cases), _p(cases), _p(out)) == 0
    return out

"""ctypes mirror of include/heatray_cuda.h and include/heatray_host.h.

`load_library()` returns the in-tree CUDA backend and FAILS LOUDLY when it is missing — the product has
no CPU path. `Backend` is a thin object wrapper over a library that exports the C-ABI with a given
prefix; the tests build the same wrapper around the CPU oracle (see oracle/loader.py).
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

PKG = Path(__file__).resolve().parent
LIB_PATH = PKG / "libheatray_cuda.so"

HRC_OK = 0
ERR_NAMES = {1: "INVALID_VALUE", 2: "INVALID_OPERATION", 3: "OUT_OF_MEMORY", 4: "CUDA", 5: "UNSUPPORTED", 6: "NO_DEVICE"}
MATERIAL_PBR, MATERIAL_GLASS = 0, 1
MAT_DOUBLE_SIDED = 1
TEX_R32F, TEX_RGB32F, TEX_RGBA32F, TEX_RGBA8 = 0, 1, 2, 3
WRAP_CLAMP, WRAP_REPEAT = 0, 1
DRAW_TRIANGLES, DRAW_TRIANGLE_STRIP = 0, 1
HIT_BACKFACE = 0x40000000
OPT_TRAVERSAL_COUNTERS, OPT_KERNEL_TIMING, OPT_SPECULATIVE_PASSES, OPT_PIXEL_RAY_COUNTS, OPT_SPECULATIVE_STRIDE = 0, 1, 2, 3, 4
SAMPLE_RANDOM, SAMPLE_HALTON, SAMPLE_HAMMERSLEY, SAMPLE_BLUE_NOISE, SAMPLE_SOBOL = range(5)
BOKEH_CIRCULAR, BOKEH_PENTAGON, BOKEH_HEXAGON, BOKEH_OCTAGON = range(4)

f32, i32, u32, u64 = C.c_float, C.c_int32, C.c_uint32, C.c_uint64
fp = C.POINTER(C.c_float)


class HrcError(RuntimeError):
    def __init__(self, code, text):
        super().__init__(f"HRC_ERR_{ERR_NAMES.get(code, code)}: {text}")
        self.code = code


class ContextDesc(C.Structure):
    _fields_ = [("device", i32), ("collectStats", i32)]


class TextureDesc(C.Structure):
    _fields_ = [("width", i32), ("height", i32), ("format", i32), ("wrapS", i32), ("wrapT", i32)]


class PbrParams(C.Structure):
    _fields_ = [("baseColorTexture", i32), ("metallicRoughnessTexture", i32), ("emissiveTexture", i32), ("normalmap", i32),
                ("clearCoatTexture", i32), ("clearCoatRoughnessTexture", i32), ("clearCoatNormalmap", i32), ("multiscatterLUT", i32),
                ("baseColor", f32 * 3), ("emissiveColor", f32 * 3),
                ("metallic", f32), ("roughness", f32), ("specularF0", f32), ("roughnessAlpha", f32),
                ("clearCoat", f32), ("clearCoatRoughness", f32), ("clearCoatRoughnessAlpha", f32)]


class GlassParams(C.Structure):
    _fields_ = [("baseColorTexture", i32), ("normalmap", i32), ("metallicRoughnessTexture", i32),
                ("baseColor", f32 * 3), ("roughness", f32), ("roughnessAlpha", f32), ("ior", f32), ("density", f32), ("specularF0", f32)]


class MeshDesc(C.Structure):
    _fields_ = [("positions", C.c_void_p), ("positionStride", i32), ("normals", C.c_void_p), ("normalStride", i32),
                ("texCoords", C.c_void_p), ("texCoordStride", i32), ("tangents", C.c_void_p), ("tangentStride", i32),
                ("bitangents", C.c_void_p), ("bitangentStride", i32), ("colors", C.c_void_p), ("colorStride", i32),
                ("vertexCount", u32), ("indices", C.c_void_p), ("elementCount", u32), ("drawMode", i32),
                ("worldFromEntity", f32 * 16), ("frontFaceCW", i32), ("isOccluder", i32), ("material", i32)]


class EnvLight(C.Structure):
    _fields_ = [("enabled", i32), ("texture", i32), ("exposureCompensation", f32), ("thetaRotation", f32)]


class DirLight(C.Structure):
    _fields_ = [("direction", f32 * 3), ("color", f32 * 3)]


class PointLight(C.Structure):
    _fields_ = [("position", f32 * 3), ("color", f32 * 3)]


class SpotLight(C.Structure):
    _fields_ = [("position", f32 * 3), ("direction", f32 * 3), ("color", f32 * 3), ("angles", f32 * 2)]


class Globals(C.Structure):
    _fields_ = [("maxRayDepth", i32), ("sampleIndex", i32), ("maxChannelValue", f32), ("enableDebugVisualizer", i32),
                ("showGeometricNormals", i32), ("showUVs", i32), ("showTangents", i32), ("showBitangents", i32), ("showNormalmap", i32),
                ("showFinalNormals", i32), ("showBaseColor", i32), ("showRoughness", i32), ("showMetallic", i32), ("showEmissive", i32),
                ("showClearcoat", i32), ("showClearcoatRoughness", i32), ("showClearcoatNormalmap", i32), ("showShader", i32),
                ("enableAccumulatorVisualizer", i32), ("showNANs", i32), ("showInf", i32)]


class PostProcess(C.Structure):        # HrcPostProcess ↔ HeatrayRenderer::PostProcessingParams (HeatrayRenderer.h:104-117)
    _fields_ = [("tonemappingEnabled", i32), ("cameraExposure", f32), ("brightness", f32), ("contrast", f32), ("hue", f32), ("saturation", f32),
                ("vibrance", f32), ("red", f32), ("green", f32), ("blue", f32), ("vignetteIntensity", f32), ("vignetteFalloff", f32)]

    def __init__(self, *a, **kw):
        defaults = dict(tonemappingEnabled=0, cameraExposure=1.0, brightness=0.0, contrast=1.0, hue=1.0, saturation=1.0, vibrance=0.0,
                        red=1.0, green=1.0, blue=1.0, vignetteIntensity=0.0, vignetteFalloff=1.0)
        defaults.update(kw)
        super().__init__(*a, **defaults)


RESOLVE_DISPLAY_RGBA32F, RESOLVE_DISPLAY_RGB8, RESOLVE_HDR_RGB32F = 0, 1, 2


class HrhImage(C.Structure):
    _fields_ = [("width", i32), ("height", i32), ("channels", i32), ("isFloat", i32), ("data", C.c_void_p)]


class HrhLoadedTexture(C.Structure):
    _fields_ = [("image", HrhImage), ("format", i32), ("wrapS", i32), ("wrapT", i32)]


class FrameParams(C.Structure):
    _fields_ = [("fovTan", f32), ("aspectRatio", f32), ("focusDistance", f32), ("apertureRadius", f32), ("viewMatrix", f32 * 16),
                ("sampleIndex", i32), ("maxSampleIndex", f32), ("interactiveMode", i32), ("blockSize", i32 * 2),
                ("currentBlockPixelSample", i32 * 2), ("bandHeight", i32), ("bandIndex", i32), ("bandCount", i32)]


class PassStats(C.Structure):
    _fields_ = [("gpuMs", f32), ("raysPrimary", u64), ("raysExtension", u64), ("raysOcclusion", u64), ("nodesVisited", u64),
                ("trianglesTested", u64), ("nodesVisitedAny", u64), ("trianglesTestedAny", u64), ("closestLaunches", u64), ("anyLaunches", u64), ("shadeLaunches", u64),
                ("msClosest", f32), ("msAny", f32), ("msShade", f32), ("msRaygen", f32),
                ("msClosestPrimary", f32), ("reserved0", f32), ("nodesVisitedPrimary", u64), ("trianglesTestedPrimary", u64)]

    @property
    def rays(self):
        return self.raysPrimary + self.raysExtension + self.raysOcclusion


class SceneInfo(C.Structure):
    _fields_ = [("triangleCount", u64), ("nodeCount", u64), ("nodeBytes", u64), ("triangleBytes", u64), ("shadingBytes", u64), ("buildSeconds", f32)]


RAY_DTYPE = np.dtype([("origin", np.float32, 3), ("tmin", np.float32), ("direction", np.float32, 3), ("tmax", np.float32)])
HIT_DTYPE = np.dtype([("t", np.float32), ("u", np.float32), ("v", np.float32), ("primId", np.int32)])


class HrhPbrMaterial(C.Structure):
    _fields_ = [("baseColor", f32 * 3), ("emissiveColor", f32 * 3), ("metallic", f32), ("roughness", f32), ("specularF0", f32),
                ("clearCoat", f32), ("clearCoatRoughness", f32), ("doubleSided", i32), ("alphaMask", i32),
                ("baseColorTexture", i32), ("metallicRoughnessTexture", i32), ("emissiveTexture", i32), ("normalmap", i32),
                ("clearCoatTexture", i32), ("clearCoatRoughnessTexture", i32), ("clearCoatNormalmap", i32),
                ("forceEnableAllTextures", i32), ("enableVertexColors", i32)]
    _TEX = ("baseColorTexture", "metallicRoughnessTexture", "emissiveTexture", "normalmap", "clearCoatTexture", "clearCoatRoughnessTexture", "clearCoatNormalmap")

    def __init__(self, *a, **kw):
        for t in self._TEX[max(0, len(a) - 9):]:
            kw.setdefault(t, -1)              # HRC_NULL_TEXTURE
        super().__init__(*a, **kw)


class HrhGlassMaterial(C.Structure):
    _fields_ = [("baseColor", f32 * 3), ("roughness", f32), ("ior", f32), ("density", f32),
                ("baseColorTexture", i32), ("normalmap", i32), ("metallicRoughnessTexture", i32), ("forceEnableAllTextures", i32)]
    _TEX = ("baseColorTexture", "normalmap", "metallicRoughnessTexture")

    def __init__(self, *a, **kw):
        for t in self._TEX[max(0, len(a) - 4):]:
            kw.setdefault(t, -1)
        super().__init__(*a, **kw)


class HrhSceneInfo(C.Structure):
    _fields_ = [("materialCount", i32), ("meshCount", i32), ("lightCount", i32), ("aabbMin", f32 * 3), ("aabbMax", f32 * 3)]


class HrhSceneMaterial(C.Structure):
    _fields_ = [("type", i32), ("name", C.c_char_p), ("pbr", HrhPbrMaterial), ("glass", HrhGlassMaterial), ("texturePath", C.c_char_p * 7)]


class HrhSceneMesh(C.Structure):
    _fields_ = [("name", C.c_char_p), ("vertexCount", u32), ("indexCount", u32), ("positions", fp), ("normals", fp), ("texCoords", fp),
                ("tangents", fp), ("bitangents", fp), ("colors", fp), ("indices", C.POINTER(u32)), ("materialIndex", i32), ("localTransform", f32 * 16)]


class HrhSceneLight(C.Structure):
    _fields_ = [("name", C.c_char_p), ("type", i32), ("color", f32 * 3), ("intensity", f32), ("position", f32 * 3), ("theta", f32), ("phi", f32),
                ("innerAngle", f32), ("outerAngle", f32)]


class RenderOptions(C.Structure):       # HrhRenderOptions ↔ PassGenerator::RenderOptions (PassGenerator.h:49-150)
    _fields_ = [("enableInteractiveMode", i32), ("enableOfflineMode", i32), ("resetInternalState", i32),
                ("maxRenderPasses", u32), ("maxRayDepth", u32), ("maxChannelValue", f32),
                ("envEnabled", i32), ("envUseSolidColor", i32), ("envSolidColor", f32 * 3), ("envTexture", i32),
                ("envExposureCompensation", f32), ("envThetaRotation", f32),
                ("aspectRatio", f32), ("focusDistance", f32), ("focalLength", f32), ("apertureRadius", f32), ("fstop", f32),
                ("viewMatrix", f32 * 16), ("sampleMode", i32), ("bokehShape", i32), ("debugVisMode", i32),
                ("debugPassRendering", i32), ("debugPassIndex", i32)]


SESSION_STRING_SIZE = 512


class HrhSession(C.Structure):          # one field per HEATRAY_SESSION_VARIABLES entry, in table order (Session.h:21-78)
    _fields_ = [("interactiveMode", i32), ("maxRenderPasses", u32), ("maxChannelValue", f32), ("scene", C.c_char * SESSION_STRING_SIZE),
                ("maxRayDepth", u32), ("sampleMode", u32), ("bokehShape", u32),
                ("environmentMap", C.c_char * SESSION_STRING_SIZE), ("environmentBuiltIn", i32),
                ("environmentExposureCompensation", f32), ("environmentThetaRotation", f32), ("environmentSolidColor", f32 * 3),
                ("cameraAspectRatio", f32), ("cameraFocusDistance", f32), ("cameraFocalLength", f32), ("cameraApertureRadius", f32), ("cameraFStop", f32),
                ("orbitDistance", f32), ("orbitPhi", f32), ("orbitTheta", f32), ("orbitTarget", f32 * 3), ("orbitMaxDistance", f32),
                ("units", u32), ("aabbMin", f32 * 3), ("aabbMax", f32 * 3), ("distanceScale", f32),
                ("rotationYaw", f32), ("rotationPitch", f32), ("rotationRoll", f32), ("scale", f32),
                ("tonemapEnable", i32), ("exposure", f32), ("brightness", f32), ("contrast", f32), ("hue", f32), ("saturation", f32), ("vibrance", f32),
                ("red", f32), ("green", f32), ("blue", f32), ("vignetteIntensity", f32), ("vignetteFalloff", f32)]

    def flat(self):
        """The 50 variables in table order: numbers as float, the two strings as str."""
        out = []
        for name, t in self._fields_:
            v = getattr(self, name)
            if isinstance(v, bytes):
                out.append(v.decode())
            elif hasattr(v, "__len__"):
                out.extend(float(x) for x in v)
            else:
                out.append(float(v))
        return out


PASS_COMPLETE_CALLBACK = C.CFUNCTYPE(None, C.c_void_p, i32, fp, i32, i32, f32, u64)
CONTEXT_TASK = C.CFUNCTYPE(None, C.c_void_p, C.c_void_p)


# names of every exported symbol of the backend C-ABI (checked by tests/test_abi.py against the header)
BACKEND_SYMBOLS = [
    "CreateContext", "DestroyContext", "SetOption", "GetLastError", "Resize", "Clear", "ReadPixels", "ReleasePixels", "AddPixels", "GetDevicePixels", "Resolve",
    "CreateTexture2D", "DestroyTexture", "CreateMaterial", "UpdateMaterial", "CreateMesh", "DestroyMesh", "SetMeshTransform", "ClearScene", "CommitScene",
    "SetEnvironment", "SetDirectionalLights", "SetPointLights", "SetSpotLights", "SetSequences", "SetPixelTables", "SetInteractiveBlocks", "SetGlobals",
    "RenderPass", "EnqueuePass", "EnqueuePassDeferred", "CommitDeferred", "DiscardDeferred", "Synchronize", "ResetStats", "TimerStart", "TimerStop", "GetStats", "TraceRays", "GetSceneInfo", "ReadRayCounts", "TraceRaysNoAlloc",
    "ExchangeCreate", "ExchangeConnect", "ExchangeSum", "ExchangeBands", "ExchangeResult", "ExchangeStatus", "ExchangeDestroy", "ExchangeGetLastError",
]

_lib = None


def load_library() -> C.CDLL:
    """The in-tree CUDA backend. Raises if it has not been built: there is no fallback."""
    global _lib
    if _lib is None:
        path = Path(os.environ.get("HEATRAY_CUDA_LIB", LIB_PATH))      # override: experimental kernel variants (tools/build_variant.sh)
        if not path.exists():
            raise RuntimeError(f"{path} is missing: run `python -m heatray_b200.build` (nvcc, sm_100a). There is no CPU fallback.")
        _lib = C.CDLL(str(path))
        _declare_host(_lib)
    return _lib


_host = None


def load_host_library() -> C.CDLL:
    """The host layer alone (libheatray_host.so: QMC tables, packing, mesh providers, image / scene / session readers — no device code).
    What the oracle-side renderer and bench.py's CPU reference arm use, so that those processes never map the CUDA backend."""
    global _host
    if _host is None:
        path = PKG / "libheatray_host.so"
        if not path.exists():
            raise RuntimeError(f"{path} is missing: run `python -m heatray_b200.build`.")
        _host = C.CDLL(str(path))
        _declare_host(_host, backend=False)
    return _host


def _declare_host(lib, backend=True):
    lib.hrhSequenceOffsetCount.restype = C.c_size_t
    lib.hrhSequenceOffsetCount.argtypes = [i32, i32]
    lib.hrhGenerateSequenceOffsets.argtypes = [i32, i32, C.c_void_p, C.c_size_t]
    lib.hrhGenerateSequenceIds.argtypes = [i32, i32, i32, C.c_void_p]
    lib.hrhGenerateRandomSequences.argtypes = [i32, i32, u32, u32, C.c_void_p, C.c_void_p]
    for n in ("hrhSobol", "hrhHalton", "hrhHammersley", "hrhRadialSobol"):
        getattr(lib, n).argtypes = [C.c_void_p, u32, u32]
    lib.hrhBlueNoise.argtypes = [C.c_void_p, u32, i32]
    lib.hrhUniformRandom.argtypes = [C.c_void_p, u32, u32, f32, f32]
    lib.hrhRandomPolygonal.argtypes = [C.c_void_p, u32, u32, u32]
    lib.hrhPackPbrMaterial.restype = u32
    lib.hrhPackPbrMaterial.argtypes = [C.POINTER(HrhPbrMaterial), i32, C.POINTER(PbrParams)]
    lib.hrhPackGlassMaterial.restype = u32
    lib.hrhPackGlassMaterial.argtypes = [C.POINTER(HrhGlassMaterial), C.POINTER(GlassParams)]
    lib.hrhGenerateMultiscatterLUT.argtypes = [C.c_void_p, i32, i32, i32]
    lib.hrhPackDirectionalLight.argtypes = [fp, f32, f32, f32, C.POINTER(DirLight)]
    lib.hrhPackPointLight.argtypes = [fp, f32, fp, C.POINTER(PointLight)]
    lib.hrhPackSpotLight.argtypes = [fp, f32, fp, f32, f32, f32, f32, C.POINTER(SpotLight)]
    lib.hrhEnvironmentExposure.restype = f32
    lib.hrhEnvironmentExposure.argtypes = [f32]
    lib.hrhOrbitCameraViewMatrix.argtypes = [f32, f32, f32, fp, fp]
    lib.hrhFovTan.restype = f32
    lib.hrhFovTan.argtypes = [f32]
    lib.hrhApertureRadius.restype = f32
    lib.hrhApertureRadius.argtypes = [f32, f32]
    if backend:            # hr::PassGenerator drives the backend: it lives in libheatray_cuda.so
        lib.hrhDefaultRenderOptions.argtypes = [C.POINTER(RenderOptions)]
        lib.hrhPassGeneratorCreate.argtypes = [i32, i32, i32, C.POINTER(C.c_void_p)]
        lib.hrhPassGeneratorDestroy.argtypes = [C.c_void_p]
        lib.hrhPassGeneratorResize.argtypes = [C.c_void_p, i32, i32]
        lib.hrhPassGeneratorRenderPass.argtypes = [C.c_void_p, C.POINTER(RenderOptions), PASS_COMPLETE_CALLBACK, C.c_void_p]
        lib.hrhPassGeneratorRunTask.argtypes = [C.c_void_p, CONTEXT_TASK, C.c_void_p]
        lib.hrhPassGeneratorWaitIdle.argtypes = [C.c_void_p]
        lib.hrhPassGeneratorLastError.restype = C.c_char_p
        lib.hrhPassGeneratorLastError.argtypes = [C.c_void_p]
    lib.hrhLoadImage.argtypes = [C.c_char_p, i32, C.POINTER(HrhImage)]
    lib.hrhLoadImageFromMemory.argtypes = [C.c_void_p, C.c_size_t, i32, C.POINTER(HrhImage)]
    lib.hrhFreeImage.argtypes = [C.POINTER(HrhImage)]
    lib.hrhLoadTexture.argtypes = [C.c_char_p, i32, i32, C.POINTER(HrhLoadedTexture)]
    lib.hrhExpandToRGBA8.argtypes = [C.POINTER(HrhImage), C.c_void_p]
    lib.hrhWritePNG.argtypes = [C.c_char_p, C.c_void_p, i32, i32, i32, i32]
    lib.hrhWriteHDR.argtypes = [C.c_char_p, C.c_void_p, i32, i32, i32]
    lib.hrhWritePFM.argtypes = [C.c_char_p, C.c_void_p, i32, i32, i32]
    lib.hrhImageLastError.restype = C.c_char_p
    lib.hrhImportScene.argtypes = [C.c_char_p, i32, C.POINTER(C.c_void_p)]
    lib.hrhBuildPresetScene.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
    lib.hrhFreeScene.argtypes = [C.c_void_p]
    lib.hrhSceneGetInfo.argtypes = [C.c_void_p, C.POINTER(HrhSceneInfo)]
    lib.hrhSceneGetMaterial.argtypes = [C.c_void_p, i32, C.POINTER(HrhSceneMaterial)]
    lib.hrhSceneGetMesh.argtypes = [C.c_void_p, i32, C.POINTER(HrhSceneMesh)]
    lib.hrhSceneGetLight.argtypes = [C.c_void_p, i32, C.POINTER(HrhSceneLight)]
    lib.hrhImportLastError.restype = C.c_char_p
    lib.hrhLightAnglesFromTransform.argtypes = [fp, fp, fp]
    lib.hrhSceneTransformMatrix.argtypes = [f32, f32, f32, f32, fp]
    lib.hrhSceneTransformMatrix.restype = None
    lib.hrhSceneApplyTransform.argtypes = [C.c_void_p, fp]
    lib.hrhSceneApplyTransform.restype = None
    lib.hrhAABBCamera.argtypes = [fp, fp, fp, fp, fp]
    lib.hrhAABBBottom.argtypes = [fp, fp, fp]
    lib.hrhAABBBottom.restype = f32
    lib.hrhSceneAddGroundPlane.argtypes = [C.c_void_p, fp, fp, fp]
    lib.hrhSessionDefaults.argtypes = [C.POINTER(HrhSession)]
    lib.hrhSessionDefaults.restype = None
    lib.hrhSessionRead.argtypes = [C.c_char_p, C.POINTER(HrhSession), C.POINTER(i32)]
    lib.hrhSessionWrite.argtypes = [C.c_char_p, C.POINTER(HrhSession)]
    lib.hrhSessionApply.argtypes = [C.POINTER(HrhSession), C.POINTER(RenderOptions), C.POINTER(PostProcess)]
    lib.hrhSessionApply.restype = None
    lib.hrhSessionLastError.restype = C.c_char_p
    lib.hrhPlaneMesh.argtypes = [f32, f32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.hrhSphereMesh.argtypes = [i32, i32, f32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(u32), C.POINTER(u32)]


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Backend:
    """Object view of one backend context: the calls a reference-side RLWrapper shim would make."""

    def __init__(self, lib: C.CDLL, prefix: str = "hrc", device: int = 0, collect_stats: bool = False, borrowed_ctx=None):
        self.lib, self.prefix = lib, prefix
        self._fn = {}
        self._owns = borrowed_ctx is None
        for name in BACKEND_SYMBOLS:
            try:
                self._fn[name] = getattr(lib, prefix + name)
            except AttributeError:
                if prefix == "hrc":
                    raise
        self._fn["GetLastError"].restype = C.c_char_p
        self._fn["GetLastError"].argtypes = [C.c_void_p]
        self.width = self.height = 0
        if borrowed_ctx is not None:          # a context owned by hr::PassGenerator (inside a task callback)
            self.ctx = C.c_void_p(borrowed_ctx)
            return
        self.ctx = C.c_void_p()
        desc = ContextDesc(device, int(collect_stats))
        rc = self._fn["CreateContext"](C.byref(desc), C.byref(self.ctx))
        if rc != HRC_OK:
            raise HrcError(rc, (self._fn["GetLastError"](None) or b"").decode())
        self._keep = []

    # -- plumbing ---------------------------------------------------------------------------
    def _call(self, name, *args):
        rc = self._fn[name](self.ctx, *args)
        if rc != HRC_OK:
            raise HrcError(rc, (self._fn["GetLastError"](self.ctx) or b"").decode())

    def close(self):
        if self.ctx and self._owns:
            self._fn["DestroyContext"](self.ctx)
            self.ctx = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def set_option(self, option: int, value: int):
        self._call("SetOption", i32(option), i32(value))

    # -- framebuffer ------------------------------------------------------------------------
    def resize(self, w, h):
        self._call("Resize", i32(w), i32(h))
        self.width, self.height = w, h

    def clear(self):
        self._call("Clear")

    def read_pixels(self) -> np.ndarray:
        p = fp()
        self._call("ReadPixels", C.byref(p))
        out = np.ctypeslib.as_array(p, shape=(self.height, self.width, 4)).copy()
        self._call("ReleasePixels")
        return out

    def map_pixels(self):
        p = fp()
        self._call("ReadPixels", C.byref(p))
        return np.ctypeslib.as_array(p, shape=(self.height, self.width, 4))

    def unmap_pixels(self):
        self._call("ReleasePixels")

    def add_pixels(self, rgba: np.ndarray):
        a = np.ascontiguousarray(rgba, np.float32)
        self._call("AddPixels", _ptr(a))

    def resolve(self, post: "PostProcess | None" = None, fmt: int = RESOLVE_DISPLAY_RGB8) -> np.ndarray:
        """hrcResolve: the display colour pipeline (displayGL.frag) or the HDR screenshot divide, rows bottom-up."""
        shape, dt = {RESOLVE_DISPLAY_RGBA32F: ((self.height, self.width, 4), np.float32), RESOLVE_DISPLAY_RGB8: ((self.height, self.width, 3), np.uint8),
                     RESOLVE_HDR_RGB32F: ((self.height, self.width, 3), np.float32)}.get(fmt, ((self.height, self.width, 4), np.float32))
        out = np.empty(shape, dt)
        self._call("Resolve", C.byref(post) if post is not None else None, i32(fmt), _ptr(out))
        return out

    def device_pixels(self) -> int:
        p = C.c_void_p()
        self._call("GetDevicePixels", C.byref(p))
        return p.value

    # -- resources --------------------------------------------------------------------------
    def create_texture(self, pixels: np.ndarray, fmt: int, wrap_s=WRAP_CLAMP, wrap_t=WRAP_CLAMP) -> int:
        a = np.ascontiguousarray(pixels)
        d = TextureDesc(a.shape[1], a.shape[0], fmt, wrap_s, wrap_t)
        t = i32()
        self._call("CreateTexture2D", C.byref(d), _ptr(a), C.byref(t))
        return t.value

    def destroy_texture(self, t):
        self._call("DestroyTexture", i32(t))

    def create_material(self, mtype: int, flags: int, params) -> int:
        m = i32()
        self._call("CreateMaterial", i32(mtype), u32(flags), C.byref(params), C.c_size_t(C.sizeof(params)), C.byref(m))
        return m.value

    def update_material(self, mat: int, params):
        self._call("UpdateMaterial", i32(mat), C.byref(params), C.c_size_t(C.sizeof(params)))

    def create_mesh(self, positions, normals, indices, material, transform=None, draw_mode=DRAW_TRIANGLES,
                    front_face_cw=None, is_occluder=True, uvs=None, tangents=None, bitangents=None, colors=None) -> int:
        pos = np.ascontiguousarray(positions, np.float32).reshape(-1, 3)
        nrm = np.ascontiguousarray(normals, np.float32).reshape(-1, 3)
        idx = np.ascontiguousarray(indices, np.uint32).reshape(-1)
        M = np.eye(4, dtype=np.float32) if transform is None else np.asarray(transform, np.float32).reshape(4, 4)
        d = MeshDesc()
        d.positions, d.positionStride = _ptr(pos), 12
        d.normals, d.normalStride = _ptr(nrm), 12
        d.vertexCount, d.indices, d.elementCount, d.drawMode = pos.shape[0], _ptr(idx), idx.shape[0], draw_mode
        keep = []
        for name, arr, comps in (("texCoords", uvs, 2), ("tangents", tangents, 3), ("bitangents", bitangents, 3), ("colors", colors, 3)):
            if arr is not None:
                a = np.ascontiguousarray(arr, np.float32).reshape(-1, comps)
                assert a.shape[0] == pos.shape[0], f"{name}: one entry per vertex"
                keep.append(a)
                setattr(d, name, _ptr(a))
                setattr(d, name[:-1] + "Stride", comps * 4)
        # `transform` is given as a mathematical 4x4 (row-major numpy); the ABI wants glm column-major
        d.worldFromEntity = (f32 * 16)(*M.T.reshape(-1))
        if front_face_cw is None:
            front_face_cw = bool(np.linalg.det(M.astype(np.float64)) < 0.0)     # Mesh.cpp:86-91
        d.frontFaceCW, d.isOccluder, d.material = int(front_face_cw), int(is_occluder), material
        m = i32()
        self._call("CreateMesh", C.byref(d), C.byref(m))
        return m.value

    def destroy_mesh(self, mesh):
        self._call("DestroyMesh", i32(mesh))

    def set_mesh_transform(self, mesh, transform):
        """Scene::applyTransform for one mesh: `transform` is a mathematical 4x4 (row-major numpy); re-commit afterwards."""
        M = np.asarray(transform, np.float32).reshape(4, 4)
        self._call("SetMeshTransform", i32(mesh), (f32 * 16)(*M.T.reshape(-1)))

    def clear_scene(self):
        self._call("ClearScene")

    def commit(self):
        self._call("CommitScene")

    def scene_info(self) -> SceneInfo:
        s = SceneInfo()
        self._call("GetSceneInfo", C.byref(s))
        return s

    # -- lights / tables / globals ----------------------------------------------------------
    def set_environment(self, enabled, texture=-1, exposure=1.0, theta=0.0):
        e = EnvLight(int(enabled), texture, exposure, theta)
        self._call("SetEnvironment", C.byref(e))

    def set_directional_lights(self, lights):
        arr = (DirLight * max(1, len(lights)))(*lights)
        self._call("SetDirectionalLights", i32(len(lights)), arr)

    def set_point_lights(self, lights):
        arr = (PointLight * max(1, len(lights)))(*lights)
        self._call("SetPointLights", i32(len(lights)), arr)

    def set_spot_lights(self, lights):
        arr = (SpotLight * max(1, len(lights)))(*lights)
        self._call("SetSpotLights", i32(len(lights)), arr)

    def set_sequences(self, xy: np.ndarray, aperture: np.ndarray):
        xy = np.ascontiguousarray(xy, np.float32)
        ap = np.ascontiguousarray(aperture, np.float32)
        self._call("SetSequences", i32(xy.shape[0]), i32(xy.shape[1]), _ptr(xy), _ptr(ap))

    def set_pixel_tables(self, seq_offset_x: np.ndarray, sequence_id: np.ndarray):
        a = np.ascontiguousarray(seq_offset_x, np.float32)
        b = np.ascontiguousarray(sequence_id, np.uint8)
        self._call("SetPixelTables", _ptr(a), C.c_size_t(a.size), _ptr(b))

    def set_interactive_blocks(self, block_coords: np.ndarray, block_rand: np.ndarray):
        """block_coords: (blockH, blockW, 3) texels; block_rand: (blocksY, blocksX, 2)."""
        bc = np.ascontiguousarray(block_coords, np.float32); br = np.ascontiguousarray(block_rand, np.float32)
        self._call("SetInteractiveBlocks", i32(bc.shape[1]), i32(bc.shape[0]), _ptr(bc), i32(br.shape[1]), i32(br.shape[0]), _ptr(br))

    def set_globals(self, g: Globals):
        self._call("SetGlobals", C.byref(g))

    # -- passes -----------------------------------------------------------------------------
    def render_pass(self, frame: FrameParams, want_stats=True):
        st = PassStats()
        self._call("RenderPass", C.byref(frame), C.byref(st) if want_stats else None)
        return st

    def enqueue_pass(self, frame: FrameParams):
        self._call("EnqueuePass", C.byref(frame))

    def enqueue_pass_deferred(self, frame: FrameParams):
        self._call("EnqueuePassDeferred", C.byref(frame))

    def commit_deferred(self):
        self._call("CommitDeferred")

    def discard_deferred(self):
        self._call("DiscardDeferred")

    def synchronize(self):
        self._call("Synchronize")

    def timer_start(self):
        self._call("TimerStart")

    def timer_stop(self) -> float:
        ms = f32()
        self._call("TimerStop", C.byref(ms))
        return ms.value

    def reset_stats(self):
        self._call("ResetStats")

    def get_stats(self) -> PassStats:
        st = PassStats()
        self._call("GetStats", C.byref(st))
        return st

    def ray_counts(self) -> np.ndarray:
        """Rays per pixel since the last clear (OPT_PIXEL_RAY_COUNTS), (h, w) uint32, row 0 = bottom."""
        out = np.zeros((self.height, self.width), np.uint32)
        self._call("ReadRayCounts", _ptr(out))
        return out

    def trace_rays(self, rays: np.ndarray, any_hit=False) -> np.ndarray:
        r = np.ascontiguousarray(rays, RAY_DTYPE)
        hits = np.zeros(r.shape[0], HIT_DTYPE)
        self._call("TraceRays", i32(r.shape[0]), _ptr(r), i32(int(any_hit)), _ptr(hits))
        return hits

    def trace_rays_no_alloc(self, rays: np.ndarray, any_hit=False) -> np.ndarray:
        """hrcTraceRaysNoAlloc: the same, staged in the library's static device array (at most NO_ALLOC_MAX_RAYS rays, no cudaMalloc)."""
        r = np.ascontiguousarray(rays, RAY_DTYPE)
        hits = np.zeros(r.shape[0], HIT_DTYPE)
        self._call("TraceRaysNoAlloc", i32(r.shape[0]), _ptr(r), i32(int(any_hit)), _ptr(hits))
        return hits


NO_ALLOC_MAX_RAYS = 262144
EXCHANGE_HANDLE_BYTES, EXCHANGE_MAX_RANKS = 192, 8


# ---- images (host_image.cpp) --------------------------------------------------------------------
def _image_to_numpy(lib, im: HrhImage) -> np.ndarray:
    n = im.width * im.height * im.channels
    ct = C.c_float if im.isFloat else C.c_uint8
    a = np.ctypeslib.as_array(C.cast(im.data, C.POINTER(ct)), shape=(n,)).copy().reshape(im.height, im.width, im.channels)
    lib.hrhFreeImage(C.byref(im))
    return a


def load_image(path_or_bytes, flip=False) -> np.ndarray:
    """hrhLoadImage / hrhLoadImageFromMemory -> (h, w, channels) uint8 or float32 array."""
    lib = load_host_library()
    im = HrhImage()
    if isinstance(path_or_bytes, (bytes, bytearray)):
        buf = (C.c_uint8 * len(path_or_bytes)).from_buffer_copy(path_or_bytes)
        rc = lib.hrhLoadImageFromMemory(buf, len(path_or_bytes), int(flip), C.byref(im))
    else:
        rc = lib.hrhLoadImage(str(path_or_bytes).encode(), int(flip), C.byref(im))
    if rc != HRC_OK:
        raise HrcError(rc, (lib.hrhImageLastError() or b"").decode())
    return _image_to_numpy(lib, im)


def load_texture(path, generate_mips=True, convert_to_linear=False):
    """util::loadTexture (TextureLoader.cpp:17-137): returns (pixels, HRC_TEX_* format, wrapS, wrapT); uint8 images come back expanded to RGBA8."""
    lib = load_library()
    lt = HrhLoadedTexture()
    rc = lib.hrhLoadTexture(str(path).encode(), int(generate_mips), int(convert_to_linear), C.byref(lt))
    if rc != HRC_OK:
        raise HrcError(rc, (lib.hrhImageLastError() or b"").decode())
    fmt, ws, wt = lt.format, lt.wrapS, lt.wrapT
    if lt.image.isFloat:
        return _image_to_numpy(lib, lt.image), fmt, ws, wt
    out = np.empty((lt.image.height, lt.image.width, 4), np.uint8)
    lib.hrhExpandToRGBA8(C.byref(lt.image), _ptr(out))
    lib.hrhFreeImage(C.byref(lt.image))
    return out, fmt, ws, wt


def session_defaults() -> HrhSession:
    s = HrhSession()
    load_library().hrhSessionDefaults(C.byref(s))
    return s


def read_session(path, into: HrhSession = None):
    """Session::parseSessionFile (Session.cpp:88-150): returns (session, number of variables the file did not hold)."""
    lib = load_library()
    s = into if into is not None else session_defaults()
    missing = i32()
    rc = lib.hrhSessionRead(str(path).encode(), C.byref(s), C.byref(missing))
    if rc != HRC_OK:
        raise HrcError(rc, (lib.hrhSessionLastError() or b"").decode())
    return s, missing.value


def write_session(path, session: HrhSession):
    lib = load_library()
    rc = lib.hrhSessionWrite(str(path).encode(), C.byref(session))
    if rc != HRC_OK:
        raise HrcError(rc, (lib.hrhSessionLastError() or b"").decode())


def apply_session(session: HrhSession):
    """HeatrayRenderer::readSessionFile's mapping (HeatrayRenderer.cpp:503-600): returns (RenderOptions, PostProcess)."""
    o, p = RenderOptions(), PostProcess()
    load_library().hrhSessionApply(C.byref(session), C.byref(o), C.byref(p))
    return o, p


def write_png(path, pixels: np.ndarray, rows_bottom_up=True):
    a = np.ascontiguousarray(pixels, np.uint8)
    h, w = a.shape[:2]
    ch = 1 if a.ndim == 2 else a.shape[2]
    rc = load_library().hrhWritePNG(str(path).encode(), _ptr(a), w, h, ch, int(rows_bottom_up))
    if rc != HRC_OK:
        raise HrcError(rc, (load_library().hrhImageLastError() or b"").decode())


def write_hdr(path, rgb: np.ndarray, rows_bottom_up=True, pfm=False):
    a = np.ascontiguousarray(rgb, np.float32)
    h, w = a.shape[:2]
    lib = load_library()
    rc = (lib.hrhWritePFM if pfm else lib.hrhWriteHDR)(str(path).encode(), _ptr(a), w, h, int(rows_bottom_up))
    if rc != HRC_OK:
        raise HrcError(rc, (lib.hrhImageLastError() or b"").decode())

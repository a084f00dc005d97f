/*
 * heatray_host.h — flat C view of the host layer that sits ABOVE the backend C-ABI
 * (heatray_cuda.h): the parts of the reference that run on the CPU before/around a pass and
 * that the reference keeps in PassGenerator / Materials / Lights / Scene / Utility.
 *
 * In a real integration these stay reference code (they are unchanged by swapping OpenRL for
 * the CUDA backend); they are rebuilt here because the reference host cannot be linked without
 * OpenRL + OpenGL, and they are what the headless bench and the tests drive. The C++ classes
 * behind these functions live in heatray_b200/csrc/host_*.{h,cpp} (namespace hr) and mirror the
 * reference's names: hr::PassGenerator, hr::RenderOptions, hr::PhysicallyBasedMaterial, ...
 */
#ifndef HEATRAY_HOST_H
#define HEATRAY_HOST_H

#include "heatray_cuda.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- QMC (reference Source/Utility/Random.h, BlueNoise.h; PassGenerator.cpp:150-159,603-684) ---- */
enum { HRH_SAMPLE_RANDOM = 0, HRH_SAMPLE_HALTON = 1, HRH_SAMPLE_HAMMERSLEY = 2, HRH_SAMPLE_BLUE_NOISE = 3, HRH_SAMPLE_SOBOL = 4 }; /* RenderOptions::SampleMode, PassGenerator.h:99-105 */
enum { HRH_BOKEH_CIRCULAR = 0, HRH_BOKEH_PENTAGON = 1, HRH_BOKEH_HEXAGON = 2, HRH_BOKEH_OCTAGON = 3 };                           /* RenderOptions::BokehShape, PassGenerator.h:111-116 */
HRC_API void hrhSobol(float* xy, uint32_t count, uint32_t sequenceIndex);           /* util::sobol, Random.h:221-264 */
HRC_API void hrhHalton(float* xy, uint32_t count, uint32_t sequenceIndex);          /* util::halton, Random.h:169-217 */
HRC_API void hrhHammersley(float* xy, uint32_t count, uint32_t sequenceIndex);      /* util::hammersley, Random.h:134-154 */
HRC_API void hrhBlueNoise(float* xy, uint32_t count, int32_t sequenceIndex);        /* util::blueNoise, Random.h:158-165 */
HRC_API void hrhUniformRandom(float* xy, uint32_t count, uint32_t seed, float lo, float hi); /* util::uniformRandomFloats<vec2>, Random.h:113-130 */
HRC_API void hrhRadialSobol(float* xy, uint32_t count, uint32_t sequenceIndex);     /* util::radialSobol, Random.h:268-289 */
HRC_API void hrhRandomPolygonal(float* xy, uint32_t numEdges, uint32_t count, uint32_t seed); /* util::randomPolygonal, Random.h:293-355 */
/* generateRandomSequences (PassGenerator.cpp:603-684): xy and aperture are vec2[numSequences*sampleCount]. */
HRC_API void hrhGenerateRandomSequences(int32_t sampleMode, int32_t bokehShape, uint32_t numSequences, uint32_t sampleCount, float* xy, float* aperture);
/* generateSequenceOffsets (PassGenerator.cpp:150-159), extended to the entries the shader's index formula reaches (quirk B-Q1). */
HRC_API size_t hrhSequenceOffsetCount(int32_t w, int32_t h);
HRC_API void hrhGenerateSequenceOffsets(int32_t w, int32_t h, float* outX, size_t count);
/* sequenceID = floor(random(FrameCoord/FrameSize) * numSequences) per pixel (perspective.rlsl:62, quirk B-Q2). */
HRC_API void hrhGenerateSequenceIds(int32_t w, int32_t h, int32_t numSequences, uint8_t* out);

/* Interactive block mode (PassGenerator.cpp:266-294, perspective.rlsl:42-57): the shuffled (row, col, 0) block-coordinate
 * texture — the reference seeds the shuffle from std::random_device (quirk B-Q6); here the seed is explicit — and the
 * per-block hash pair of the frame shader. */
HRC_API void hrhGenerateInteractiveBlockCoords(int32_t blockW, int32_t blockH, uint32_t seed, float* outRGB /* blockW*blockH*3 */);
HRC_API void hrhInteractiveBlockCount(int32_t w, int32_t h, int32_t blockW, int32_t blockH, int32_t* blocksX, int32_t* blocksY);
HRC_API void hrhGenerateInteractiveBlockRand(int32_t blocksX, int32_t blocksY, float* outXY /* blocksX*blocksY*2 */);

/* ---- materials (PhysicallyBasedMaterial::Parameters → ShaderParams; GlassMaterial likewise) ---- */
typedef struct HrhPbrMaterial {       /* PhysicallyBasedMaterial.h Parameters (user-facing, unclamped) */
    float baseColor[3]; float emissiveColor[3];
    float metallic, roughness, specularF0, clearCoat, clearCoatRoughness;
    int32_t doubleSided, alphaMask;
    /* textures: HRC_NULL_TEXTURE = none (the shader then sees the 1x1 white dummy, Texture.h:178-192) */
    HrcTexture baseColorTexture, metallicRoughnessTexture, emissiveTexture, normalmap;
    HrcTexture clearCoatTexture, clearCoatRoughnessTexture, clearCoatNormalmap;
    int32_t forceEnableAllTextures;   /* PhysicallyBasedMaterial.h Parameters::forceEnableAllTextures */
    int32_t enableVertexColors;       /* Material::m_enableVertexColors (Material.h) */
} HrhPbrMaterial;
typedef struct HrhGlassMaterial {     /* GlassMaterial.h Parameters */
    float baseColor[3]; float roughness, ior, density;
    HrcTexture baseColorTexture, normalmap, metallicRoughnessTexture;
    int32_t forceEnableAllTextures;
} HrhGlassMaterial;
/* modify(): clamps and derived values (PhysicallyBasedMaterial.cpp:127-145, quirk B-Q16); returns the feature flags. */
HRC_API uint32_t hrhPackPbrMaterial(const HrhPbrMaterial* in, HrcTexture multiscatterLUT, HrcPbrParams* out);
HRC_API uint32_t hrhPackGlassMaterial(const HrhGlassMaterial* in, HrcGlassParams* out);   /* GlassMaterial.cpp:88-126 */
/* generateMultiScatterTexture (MultiScatterUtil.cpp:49-139): dim x dim floats, row = roughness bin, col = NdotV bin. */
HRC_API void hrhGenerateMultiscatterLUT(float* out, int32_t dim, int32_t sampleCount, int32_t threads);

/* ---- lights (photometric → radiometric packing, quirks B-Q17/B-Q18) ---- */
HRC_API void hrhPackDirectionalLight(const float color[3], float illuminance, float theta, float phi, HrcDirLight* out);   /* DirectionalLight.cpp:42-78 */
HRC_API void hrhPackPointLight(const float color[3], float luminousIntensity, const float position[3], HrcPointLight* out); /* PointLight.cpp:41-50 */
HRC_API void hrhPackSpotLight(const float color[3], float luminousIntensity, const float position[3], float theta, float phi,
                              float innerAngle, float outerAngle, HrcSpotLight* out);                                        /* SpotLight.cpp:44-91 */
HRC_API float hrhEnvironmentExposure(float exposureCompensationEV);                                                         /* EnvironmentLight.cpp:95 */

/* ---- camera ---- */
HRC_API void hrhOrbitCameraViewMatrix(float distance, float phi, float theta, const float target[3], float out[16]);       /* OrbitCamera.h:32-45 */
HRC_API float hrhFovTan(float focalLengthMm);                          /* PassGenerator.cpp:341-343,357 */
HRC_API float hrhApertureRadius(float focalLengthMm, float fstop);    /* RenderOptions::Camera::setApertureRadius, PassGenerator.h:92-94 */

/* ---- procedural mesh providers (Scene/PlaneMeshProvider.h, Scene/SphereMeshProvider.h) ---- */
/* Plane: 4 vertices, strip indices {0,1,3,2}. Buffers: positions[12], normals[12], uvs[8], indices[4]. */
HRC_API void hrhPlaneMesh(float width, float length, float* positions, float* normals, float* uvs, uint32_t* indices);
/* Sphere: (u+1)*(v+2) vertices, 2*u*v triangles. Pass NULL buffers to query counts only. */
HRC_API void hrhSphereMesh(int32_t uSlices, int32_t vSlices, float radius, float* positions, float* normals, float* uvs,
                           uint32_t* indices, uint32_t* vertexCount, uint32_t* indexCount);

/* ---- the pass interface: hr::PassGenerator (reference PassGenerator.h:35-193) ---- */
typedef struct HrhPassGenerator_t* HrhPassGenerator;
typedef struct HrhRenderOptions {      /* PassGenerator::RenderOptions, PassGenerator.h:49-150 (strings → handled by callers) */
    int32_t enableInteractiveMode, enableOfflineMode, resetInternalState;
    uint32_t maxRenderPasses, maxRayDepth;
    float maxChannelValue;
    /* environment */
    int32_t envEnabled; int32_t envUseSolidColor; float envSolidColor[3]; HrcTexture envTexture;
    float envExposureCompensation, envThetaRotation;
    /* camera */
    float aspectRatio, focusDistance, focalLength, apertureRadius, fstop; float viewMatrix[16];
    int32_t sampleMode, bokehShape, debugVisMode, debugPassRendering, debugPassIndex;
} HrhRenderOptions;
/* PassCompleteCallback (PassGenerator.h:161): invoked on the worker thread. `pixels` stays mapped until the next pass. */
typedef void (*HrhPassCompleteCallback)(void* user, int32_t frameDataAvailable, const float* pixels, int32_t width, int32_t height,
                                        float passTimeSeconds, uint64_t passIndex);
HRC_API void hrhDefaultRenderOptions(HrhRenderOptions* out);
HRC_API int  hrhPassGeneratorCreate(int32_t device, int32_t width, int32_t height, HrhPassGenerator* out);  /* init(), PassGenerator.cpp:30-42,161-299 */
HRC_API int  hrhPassGeneratorDestroy(HrhPassGenerator pg);                                                    /* destroy(), :100-106 */
HRC_API int  hrhPassGeneratorResize(HrhPassGenerator pg, int32_t width, int32_t height);                      /* resize(), :108-113 */
HRC_API int  hrhPassGeneratorRenderPass(HrhPassGenerator pg, const HrhRenderOptions* opts, HrhPassCompleteCallback cb, void* user); /* renderPass(), :115-122 (asynchronous enqueue) */
/* loadScene / changeLighting / modifyScene / runOpenRLTask (PassGenerator.cpp:124-148): the task runs on the worker thread and gets the backend context. */
typedef void (*HrhContextTask)(void* user, HrcContext ctx);
HRC_API int  hrhPassGeneratorRunTask(HrhPassGenerator pg, HrhContextTask task, void* user);
HRC_API int  hrhPassGeneratorWaitIdle(HrhPassGenerator pg);       /* drains the job queue (test/bench helper) */
HRC_API const char* hrhPassGeneratorLastError(HrhPassGenerator pg);

/* ---- images: texture loading (reference Source/Utility/TextureLoader.cpp:17-159, which decodes with stb_image / FreeImage)
 * and screenshots (HeatrayRenderer::saveScreenshot, HeatrayRenderer.cpp:1619-1662, FreeImage_Save). Own decoders/encoders:
 * PNG (all colour types and bit depths, Adam7, tRNS), baseline + progressive JPEG, Radiance .hdr (RGBE, flat and RLE). ---- */
typedef struct HrhImage {
    int32_t width, height, channels;   /* channels as stored in the file: 1 (luminance), 2, 3 (RGB) or 4 (RGBA) */
    int32_t isFloat;                   /* 1: data is float32 (Radiance .hdr; stbi_loadf path); 0: uint8 */
    void*   data;                      /* width*height*channels elements, rows top-down unless flipped at load; free with hrhFreeImage */
} HrhImage;
/* ↔ stbi_load / stbi_loadf with stbi_set_flip_vertically_on_load(flip) (TextureLoader.cpp:90-107); 16-bit PNG samples keep their high byte. */
HRC_API int  hrhLoadImage(const char* path, int32_t flipVertically, HrhImage* out);
HRC_API int  hrhLoadImageFromMemory(const void* bytes, size_t size, int32_t flipVertically, HrhImage* out);
HRC_API void hrhFreeImage(HrhImage* img);
/* ↔ util::loadTextureInternal (TextureLoader.cpp:17-137): "../" path fallback (:24-35), vertical flip, 8-bit sRGB->linear->8-bit
 * requantisation of the colour channels when convertToLinear (alpha untouched, :109-133), channel count -> format (:138-151,
 * 2-channel images are rejected like the reference), sampler = LINEAR + CLAMP_TO_EDGE when !generateMips else the default
 * REPEAT (+ mip filter, which policy A-12 replaces by level-0 bilinear). */
typedef struct HrhLoadedTexture {
    HrhImage image;
    int32_t  format;                   /* HRC_TEX_* to hand to hrcCreateTexture2D after hrhExpandTexture */
    int32_t  wrapS, wrapT;             /* HRC_WRAP_* */
} HrhLoadedTexture;
HRC_API int  hrhLoadTexture(const char* path, int32_t generateMips, int32_t convertToLinear, HrhLoadedTexture* out);
/* Expands 1/3/4-channel uint8 or float texels to what hrcCreateTexture2D accepts (RGBA8 or RGBA32F/RGB32F/R32F; luminance replicates to RGB
 * with alpha 1 like RL_LUMINANCE). `out` must hold width*height*4 bytes (uint8 input) or is the image itself (float input, no copy needed). */
HRC_API int  hrhExpandToRGBA8(const HrhImage* img, uint8_t* out);
/* Screenshot writers (rows given bottom-up like hrcResolve output; written top-down as the formats require). */
HRC_API int  hrhWritePNG(const char* path, const uint8_t* pixels, int32_t width, int32_t height, int32_t channels, int32_t rowsBottomUp);
HRC_API int  hrhWriteHDR(const char* path, const float* rgb, int32_t width, int32_t height, int32_t rowsBottomUp);   /* Radiance RGBE */
HRC_API int  hrhWritePFM(const char* path, const float* rgb, int32_t width, int32_t height, int32_t rowsBottomUp);
HRC_API const char* hrhImageLastError(void);

/* ---- session files (reference Session/Session.h:21-78, Session.cpp:88-214; HeatrayRenderer::read/writeSessionFile, HeatrayRenderer.cpp:503-690) ---- */
#define HRH_SESSION_STRING_SIZE 512    /* Variable::Value::c, Session.cpp:41 */
typedef struct HrhSession {            /* one field per HEATRAY_SESSION_VARIABLES entry, same order */
    int32_t interactiveMode; uint32_t maxRenderPasses; float maxChannelValue; char scene[HRH_SESSION_STRING_SIZE];
    uint32_t maxRayDepth, sampleMode, bokehShape;
    char environmentMap[HRH_SESSION_STRING_SIZE]; int32_t environmentBuiltIn;
    float environmentExposureCompensation, environmentThetaRotation, environmentSolidColor[3];
    float cameraAspectRatio, cameraFocusDistance, cameraFocalLength, cameraApertureRadius, cameraFStop;
    float orbitDistance, orbitPhi, orbitTheta, orbitTarget[3], orbitMaxDistance;
    uint32_t units; float aabbMin[3], aabbMax[3], distanceScale, rotationYaw, rotationPitch, rotationRoll, scale;
    int32_t tonemapEnable; float exposure, brightness, contrast, hue, saturation, vibrance, red, green, blue, vignetteIntensity, vignetteFalloff;
} HrhSession;
HRC_API void hrhSessionDefaults(HrhSession* session);                                 /* the table's defaults (Session.h:21-78) */
HRC_API int  hrhSessionRead(const char* path, HrhSession* session, int32_t* missing); /* parseSessionFile; absent variables keep their value */
HRC_API int  hrhSessionWrite(const char* path, const HrhSession* session);            /* writeSessionFile, tinyxml2 number formatting */
HRC_API void hrhSessionApply(const HrhSession* session, HrhRenderOptions* options, HrcPostProcess* post);   /* readSessionFile's mapping */
HRC_API const char* hrhSessionLastError(void);

/* ---- scene import (reference Scene::loadFromDisk, Scene.cpp:14-27 -> AssimpMeshProvider, AssimpMeshProvider.cpp:65-723).
 * Own readers for Wavefront OBJ/MTL and glTF 2.0 (.gltf/.glb) that apply the reference's Assimp->Heatray mapping rules and hand
 * back what MeshProvider exposes (MeshProvider.h:50-81). Pointers stay valid until hrhFreeScene. ---- */
typedef struct HrhScene_t* HrhScene;
typedef struct HrhSceneInfo { int32_t materialCount, meshCount, lightCount; float aabbMin[3], aabbMax[3]; } HrhSceneInfo;
enum { HRH_TEX_BASE_COLOR = 0, HRH_TEX_METALLIC_ROUGHNESS = 1, HRH_TEX_EMISSIVE = 2, HRH_TEX_NORMAL = 3, HRH_TEX_CLEARCOAT = 4,
       HRH_TEX_CLEARCOAT_ROUGHNESS = 5, HRH_TEX_CLEARCOAT_NORMAL = 6, HRH_TEX_SLOT_COUNT = 7 };
typedef struct HrhSceneMaterial {
    int32_t type;                        /* HRC_MATERIAL_PBR or HRC_MATERIAL_GLASS (alphaMode BLEND or transmission != 0, :365-376) */
    const char* name;
    HrhPbrMaterial pbr;                  /* user-facing parameters (ProcessMaterial :378-425); texture handles are HRC_NULL_TEXTURE: the caller creates them */
    HrhGlassMaterial glass;              /* ProcessGlassMaterial :264-283 (uses slots base colour, normal, metallic-roughness) */
    const char* texturePath[7];          /* HRH_TEX_*: file to load, NULL = none. Base colour and emissive load with convertToLinear = true, the rest false (:448-505) */
} HrhSceneMaterial;
typedef struct HrhSceneMesh {            /* one Submesh of the provider (ProcessMesh :100-262 + ProcessNode :73-98) */
    const char* name;
    uint32_t vertexCount, indexCount;    /* triangles: indexCount / 3 */
    const float *positions, *normals;    /* xyz per vertex, always present (smooth normals are generated when the file has none) */
    const float *texCoords;              /* uv per vertex or NULL */
    const float *tangents, *bitangents;  /* xyz per vertex or NULL (present whenever texCoords are) */
    const float *colors;                 /* rgb per vertex or NULL; the material then has enableVertexColors */
    const uint32_t* indices;
    int32_t materialIndex;
    float localTransform[16];            /* column-major world-from-mesh transform of the owning node */
} HrhSceneMesh;
typedef struct HrhSceneLight {           /* ProcessLight :516-660; feed hrhPackDirectionalLight / PointLight / SpotLight */
    const char* name; int32_t type;      /* 0 directional, 1 point, 2 spot */
    float color[3]; float intensity;     /* |colour * intensity| split as the reference does (:541-545) */
    float position[3]; float theta, phi, innerAngle, outerAngle;
} HrhSceneLight;
HRC_API int  hrhImportScene(const char* path, int32_t convertToMeters, HrhScene* out);
HRC_API void hrhFreeScene(HrhScene scene);
HRC_API int  hrhSceneGetInfo(HrhScene scene, HrhSceneInfo* info);
HRC_API int  hrhSceneGetMaterial(HrhScene scene, int32_t index, HrhSceneMaterial* out);
HRC_API int  hrhSceneGetMesh(HrhScene scene, int32_t index, HrhSceneMesh* out);
HRC_API int  hrhSceneGetLight(HrhScene scene, int32_t index, HrhSceneLight* out);
/* Scene::loadFromDisk's second half (Scene.cpp:14-27, Mesh.cpp:55-155): loads the textures, builds the materials, creates one backend mesh
 * per submesh, commits, replaces the punctual lights. multiscatterLUT: an existing LUT texture or HRC_NULL_TEXTURE to generate one. */
typedef struct HrhSceneUploadResult { int32_t meshCount, materialCount, textureCount, texturesMissing; HrcTexture multiscatterLUT;
                                      int32_t dirLights, pointLights, spotLights; } HrhSceneUploadResult;
HRC_API int  hrhSceneUpload(HrhScene scene, HrcContext ctx, int32_t clearOldScene, HrcTexture multiscatterLUT, HrhSceneUploadResult* result);
/* The reference's built-in scenes (HeatrayRenderer::changeScene, HeatrayRenderer.cpp:123-286): "Editable PBR Material", "Editable Glass Material",
 * "Multi-Material", "Sphere Array". The scene AABB is reported empty: the reference does not re-frame the camera for them. */
HRC_API int  hrhBuildPresetScene(const char* name, HrhScene* out);
/* Session scene transform: scale(scale) * yawPitchRoll (HeatrayRenderer.h:343-346); applied at upload as transform * mesh transform (Scene.cpp:38-49). */
HRC_API void hrhSceneTransformMatrix(float yaw, float pitch, float roll, float scale, float out16[16]);
HRC_API void hrhSceneApplyTransform(HrhScene scene, const float transform16[16]);
/* util::AABB::center / radius / valid under a transform (AABB.h:34-52; NULL = identity): camera framing uses distance = 3 radius (HeatrayRenderer.cpp:300-310). Returns valid(). */
HRC_API int  hrhAABBCamera(const float aabbMin[3], const float aabbMax[3], const float transform16[16], float center[3], float* radius);
HRC_API float hrhAABBBottom(const float aabbMin[3], const float aabbMax[3], const float transform16[16]);   /* util::AABB::bottom(), AABB.h:46-51 */
/* The viewer's "Add Ground Plane" (HeatrayRenderer.cpp:1167-1193): side max(1, size_t(radius)) * 5 at the AABB's bottom. Returns the mesh index or -1. */
HRC_API int  hrhSceneAddGroundPlane(HrhScene scene, const float aabbMin[3], const float aabbMax[3], const float transform16[16]);
/* theta/phi of an imported directional or spot light from its node transform (ProcessLight :564-586, incl. its vec4(axis, 1) quirk). */
HRC_API void hrhLightAnglesFromTransform(const float worldFromLight[16], float* theta, float* phi);
HRC_API const char* hrhImportLastError(void);

#ifdef __cplusplus
}
#endif
#endif /* HEATRAY_HOST_H */

/*
 * heatray_cuda.h — C-ABI of the B200-native path-tracing backend (libheatray_cuda.so).
 *
 * This is the drop-in boundary: the entry points below are what Heatray's
 * RLWrapper layer (reference the headers in Source/RLWrapper/, a header-only C++ veneer over the
 * closed OpenRL C API, 3rdParty/OpenRL/rl.h:383-528) would bind instead of rl*().
 * RLSL compilation has no CUDA meaning, so the shader/program string API collapses
 * into enums + POD parameter blocks; everything else maps 1:1.
 *
 * Conventions (mirroring OpenRL as the reference uses it):
 *  - plain C, plain pointers and sizes, no C++/torch types;
 *  - every call returns HRC_OK (0) or an HRC_ERR_* code; hrcGetLastError() gives text
 *    (reference: RLFunc()/rlGetError(), RLWrapper/Error.h:18-41);
 *  - a context is thread-affine: one owner thread at a time, like an OpenRL context
 *    (PassGenerator.h:229-231);
 *  - host memory passed in is copied at call time; the caller keeps ownership;
 *  - matrices are column-major float[16] (glm layout);
 *  - the frame buffer holds SUMS of clamped radiance in RGB and the sample count in A,
 *    row 0 = bottom row (PassGenerator.cpp:176-188, displayGL.frag:77-78).
 *
 * The same declarations with the prefix `orc` are exported by the CPU oracle
 * (oracle/liboracle.so) — test infrastructure only, never linked by this library.
 */
#ifndef HEATRAY_CUDA_H
#define HEATRAY_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef HRC_API
#define HRC_API __attribute__((visibility("default")))
#endif

/* ---- error codes ---------------------------------------------------------------- */
enum {
    HRC_OK = 0,
    HRC_ERR_INVALID_VALUE = 1,   /* ↔ RL_INVALID_VALUE */
    HRC_ERR_INVALID_OPERATION = 2, /* ↔ RL_INVALID_OPERATION (e.g. double map, render without commit) */
    HRC_ERR_OUT_OF_MEMORY = 3,   /* ↔ RL_OUT_OF_MEMORY */
    HRC_ERR_CUDA = 4,            /* a CUDA runtime call failed; text has the CUDA error string */
    HRC_ERR_UNSUPPORTED = 5,     /* feature of the reference not built yet; never a silent fallback */
    HRC_ERR_NO_DEVICE = 6        /* no CUDA device: the product has no CPU path */
};

typedef struct HrcContext_t* HrcContext;
typedef int32_t HrcMesh;      /* ↔ RLprimitive (one per submesh, Mesh.cpp:69) */
typedef int32_t HrcMaterial;  /* ↔ RLprogram + "Material" uniform block */
typedef int32_t HrcTexture;   /* ↔ RLtexture */
#define HRC_NULL_TEXTURE (-1)

#define HRC_MAX_LIGHTS_PER_TYPE 5   /* ShaderLightingDefines.h:18-20 */
#define HRC_NUM_RANDOM_SEQUENCES 16 /* PassGenerator.h:193 (informative; any count is accepted) */

/* ---- context (↔ OpenRLCreateContext / OpenRLDestroyContext, PassGenerator.cpp:164,432) */
typedef struct HrcContextDesc {
    int32_t device;          /* CUDA ordinal */
    int32_t collectStats;    /* 1: start with HRC_OPT_TRAVERSAL_COUNTERS and HRC_OPT_KERNEL_TIMING on */
} HrcContextDesc;

HRC_API int hrcCreateContext(const HrcContextDesc* desc, HrcContext* out);
/* Runtime switches (no reference equivalent; OpenRL exposes RL_PROFILE-style enums it never uses, rl.h:346-355). */
enum {
    HRC_OPT_TRAVERSAL_COUNTERS = 0,  /* count nodes visited / triangles tested (slightly slower kernels) */
    HRC_OPT_KERNEL_TIMING = 1,       /* CUDA events between kernels: HrcPassStats.ms* are filled by hrcRenderPass */
    HRC_OPT_PIXEL_RAY_COUNTS = 3,    /* parity hook: count the rays (primary + extension + occlusion segments) every pixel's paths submit to traversal,
                                      * readable with hrcReadRayCounts; hrcClear zeroes the counts. A pixel whose value differs from another
                                      * implementation's together with its ray count took a different discrete decision somewhere on its path. */
    HRC_OPT_SPECULATIVE_STRIDE = 4,  /* s = 1..4096 (default 1): the sample-index step HRC_OPT_SPECULATIVE_PASSES predicts with (pass i is followed by i + s).
                                      * A rank of a sample-index-sharded render (rank q of N renders q, q + N, q + 2N, ...) sets s = N. */
    HRC_OPT_SPECULATIVE_PASSES = 2   /* n = 0..3 (default 0): hrcRenderPass(i) also starts passes i+1..i+n of the same progressive render (same frame
                                      * parameters, sample indices i + s, i + 2s, ... below maxSampleIndex; s = HRC_OPT_SPECULATIVE_STRIDE) into side buffers nothing else sees. When the next
                                      * hrcRenderPass asks for exactly the predicted pass and no state changed in between, it is committed instead of
                                      * rendered; any other request discards the predictions. For callers that render and read back one pass at a
                                      * time (the reference viewer's loop, HeatrayRenderer.cpp:366-408): the readback then overlaps the next passes.
                                      * The picture differs from the unspeculated one by fp32 addition order only. With n > 0 HrcPassStats counts the
                                      * work that finished during the call (one pass in steady state), and hrcEnqueuePassDeferred is unavailable. */
};
HRC_API int hrcSetOption(HrcContext ctx, int32_t option, int32_t value);
HRC_API int hrcDestroyContext(HrcContext ctx);
/* ctx may be NULL for errors raised before a context exists. */
HRC_API const char* hrcGetLastError(HrcContext ctx);

/* ---- framebuffer + readback (↔ Framebuffer/Texture RGBA32F + PixelPackBuffer) ------- */
/* ↔ runInitJob FBO part + runResizeJob (PassGenerator.cpp:172-197, 301-323); clears.
 * Limit: width * height <= 2^24 pixels (16.7 M; 3840x2160 is 8.3 M): a ray record carries its pixel index in 24 bits next to its depth and
 * flags. Larger frames return HRC_ERR_UNSUPPORTED; render them as tiles or by screen bands on several contexts (HrcFrameParams.bandHeight / bandIndex / bandCount). */
HRC_API int hrcResize(HrcContext ctx, int32_t width, int32_t height);
/* ↔ rlClear(RL_COLOR_BUFFER_BIT) (PassGenerator.cpp:439) */
HRC_API int hrcClear(HrcContext ctx);
/* ↔ PixelPackBuffer::setPixelData + mapPixelData (PixelPackBuffer.h:39-60): copies the
 * accumulation buffer into pinned host memory and returns the pointer (valid until release).
 * A second read without release is HRC_ERR_INVALID_OPERATION (the reference asserts). */
HRC_API int hrcReadPixels(HrcContext ctx, const float** rgba);
HRC_API int hrcReleasePixels(HrcContext ctx);   /* ↔ unmapPixelData (PixelPackBuffer.h:64-70) */
/* Add `rgba` (w*h*4 floats, host) into the accumulation buffer: the multi-GPU reduce path
 * writes the summed buffer back through this (no reference equivalent). */
HRC_API int hrcAddPixels(HrcContext ctx, const float* rgba);
/* Device pointer of the accumulation buffer (w*h float4), for NCCL reduce without staging. */
HRC_API int hrcGetDevicePixels(HrcContext ctx, void** devPtr);

/* ---- multi-GPU frame exchange over NVLink peer memory (no reference equivalent) ----
 * One process per GPU renders its share (sample indices or screen bands); the frames are combined on rank 0 by ONE kernel launch per rank
 * that reads the peers' snapshots / writes rank 0's result directly through peer mappings (csrc/peer_exchange.cu): no NCCL call, no host
 * synchronisation. Set-up: every rank calls hrcExchangeCreate and hands its HRC_EXCHANGE_HANDLE_BYTES of CUDA-IPC handles to all others
 * through whatever process group the caller has (torch.distributed all_gather in heatray_b200/sharding.py), then hrcExchangeConnect with
 * all of them in rank order. Every rank must then call hrcExchangeSum / hrcExchangeBands the same number of times, like any collective;
 * the calls only enqueue work on `cudaStream` (a cudaStream_t, NULL = the default stream). On rank 0 the combined frame is complete in the
 * buffer hrcExchangeResult returns once that stream has run the call. */
typedef struct HrcExchange_t* HrcExchange;
#define HRC_EXCHANGE_MAX_RANKS 8
#define HRC_EXCHANGE_HANDLE_BYTES 192
HRC_API int hrcExchangeCreate(int32_t device, int32_t rank, int32_t world, uint64_t float4Count, HrcExchange* out, uint8_t* handleOut);
HRC_API int hrcExchangeConnect(HrcExchange ex, const uint8_t* allHandles /* world * HRC_EXCHANGE_HANDLE_BYTES */);
/* Sample-index sharding: result = sum over the ranks, in rank order, of `srcDevice` (float4Count float4 on this rank's GPU, e.g. hrcGetDevicePixels).
 * The call copies `srcDevice` into its own exported snapshot buffer first, in stream order: the caller may go on writing it afterwards. */
HRC_API int hrcExchangeSum(HrcExchange ex, const float* srcDevice, void* cudaStream);
/* Screen-band sharding: row y of the result comes from rank (y / bandHeight) % world (bit-identical to a single-GPU frame). Each rank's kernel reads its
 * rows straight from `srcDevice` (no snapshot: nobody else reads it), which must stay unchanged until `cudaStream` has run the call. */
HRC_API int hrcExchangeBands(HrcExchange ex, const float* srcDevice, int32_t width, int32_t bandHeight, void* cudaStream);
HRC_API int hrcExchangeResult(HrcExchange ex, const float** devicePtr);   /* rank 0 only */
/* Waits for the device and reports a peer that never arrived / finished (the kernel's spins are bounded, ~3 s). */
HRC_API int hrcExchangeStatus(HrcExchange ex);
HRC_API int hrcExchangeDestroy(HrcExchange ex);
HRC_API const char* hrcExchangeGetLastError(HrcExchange ex);

/* ---- resolve + colour pipeline (↔ displayGL.frag:28-151 drawn by HeatrayRenderer.h:228-262, and the
 * screenshot paths of HeatrayRenderer::saveScreenshot, HeatrayRenderer.cpp:1619-1662). The reference runs this as a
 * GL fragment shader over the window; here it is one CUDA kernel over the accumulation buffer (no graphics API). */
typedef struct HrcPostProcess {       /* ↔ HeatrayRenderer::PostProcessingParams (HeatrayRenderer.h:104-117) */
    int32_t tonemappingEnabled;       /* ACES fit, displayGL.frag:82-90 */
    float   cameraExposure;           /* already 2^exposure (HeatrayRenderer.h:238) */
    float   brightness, contrast;     /* :95-97 */
    float   hue, saturation, vibrance;/* :100-129 */
    float   red, green, blue;         /* :132-136 */
    float   vignetteIntensity, vignetteFalloff; /* :139-143 */
} HrcPostProcess;
enum {
    HRC_RESOLVE_DISPLAY_RGBA32F = 0,  /* fragColor of displayGL.frag as floats (sRGB-encoded, unclamped), w*h*4 floats */
    HRC_RESOLVE_DISPLAY_RGB8 = 1,     /* the same quantised like an 8-bit unorm colour buffer: round(clamp(c,0,1)*255), w*h*3 bytes (the LDR screenshot) */
    HRC_RESOLVE_HDR_RGB32F = 2        /* rgb / alpha only, no colour pipeline (the HDR screenshot path, HeatrayRenderer.cpp:1629-1643); `post` may be NULL */
};
/* Rows are written bottom-up like the accumulation buffer (row 0 = bottom row; glReadPixels / FreeImage order).
 * `hostOut` is caller-owned host memory of the size the format implies. Pixels with alpha == 0 divide by zero exactly as
 * the shader does (inf/NaN in float formats, 0 or 255 after the 8-bit clamp). */
HRC_API int hrcResolve(HrcContext ctx, const HrcPostProcess* post, int32_t format, void* hostOut);

/* ---- textures (↔ openrl::Texture::create, Texture.h:59-91) ------------------------- */
enum { HRC_TEX_R32F = 0, HRC_TEX_RGB32F = 1, HRC_TEX_RGBA32F = 2, HRC_TEX_RGBA8 = 3 };
enum { HRC_WRAP_CLAMP = 0, HRC_WRAP_REPEAT = 1 };
typedef struct HrcTextureDesc {
    int32_t width, height;
    int32_t format;          /* HRC_TEX_* */
    int32_t wrapS, wrapT;    /* HRC_WRAP_* */
} HrcTextureDesc;
/* Row 0 of `pixels` is the row addressed by t = 0 (bottom row, GL convention).
 * Filtering is level-0 bilinear in fp32 (policy A-12 in SURVEY.md). */
HRC_API int hrcCreateTexture2D(HrcContext ctx, const HrcTextureDesc* desc, const void* pixels, HrcTexture* out);
HRC_API int hrcDestroyTexture(HrcContext ctx, HrcTexture tex);

/* ---- materials (↔ Material::build/modify + the "Material" uniform block) ---------- */
enum { HRC_MATERIAL_PBR = 0, HRC_MATERIAL_GLASS = 1 };
/* feature flags ↔ the #defines injected by PhysicallyBasedMaterial.cpp:49-113 / GlassMaterial.cpp:40-74 */
enum {
    HRC_MAT_DOUBLE_SIDED                = 1 << 0,
    HRC_MAT_ALPHA_MASK                  = 1 << 1,
    HRC_MAT_HAS_BASE_COLOR_TEXTURE      = 1 << 2,
    HRC_MAT_HAS_METALLIC_ROUGHNESS_TEXTURE = 1 << 3,
    HRC_MAT_HAS_EMISSIVE_TEXTURE        = 1 << 4,
    HRC_MAT_HAS_NORMALMAP               = 1 << 5,
    HRC_MAT_HAS_CLEARCOAT_TEXTURE       = 1 << 6,
    HRC_MAT_HAS_CLEARCOAT_ROUGHNESS_TEXTURE = 1 << 7,
    HRC_MAT_HAS_CLEARCOAT_NORMALMAP     = 1 << 8,
    HRC_MAT_VERTEX_COLORS               = 1 << 9
};
/* ↔ ShaderParams, PhysicallyBasedMaterial.cpp:16-36 / uniformblock Material, physicallyBased.rlsl:29-48.
 * Values are the ALREADY CLAMPED shader-side values (modify(), PhysicallyBasedMaterial.cpp:127-145). */
typedef struct HrcPbrParams {
    HrcTexture baseColorTexture, metallicRoughnessTexture, emissiveTexture, normalmap;
    HrcTexture clearCoatTexture, clearCoatRoughnessTexture, clearCoatNormalmap, multiscatterLUT;
    float baseColor[3];
    float emissiveColor[3];
    float metallic, roughness, specularF0, roughnessAlpha;
    float clearCoat, clearCoatRoughness, clearCoatRoughnessAlpha;
} HrcPbrParams;
/* ↔ ShaderParams, GlassMaterial.cpp:15-28 / uniformblock Material, glass.rlsl:30-40 */
typedef struct HrcGlassParams {
    HrcTexture baseColorTexture, normalmap, metallicRoughnessTexture;
    float baseColor[3];
    float roughness, roughnessAlpha, ior, density, specularF0;
} HrcGlassParams;
HRC_API int hrcCreateMaterial(HrcContext ctx, int32_t type, uint32_t featureFlags,
                              const void* params, size_t paramBytes, HrcMaterial* out);
/* ↔ Material::modify → Buffer::modify (PhysicallyBasedMaterial.cpp:190) */
HRC_API int hrcUpdateMaterial(HrcContext ctx, HrcMaterial mat, const void* params, size_t paramBytes);

/* ---- geometry (↔ Buffer + Primitive + rlDrawElements, Mesh.cpp:28-155) -------------- */
enum { HRC_DRAW_TRIANGLES = 0, HRC_DRAW_TRIANGLE_STRIP = 1 };   /* MeshProvider.h DrawMode */
typedef struct HrcMeshDesc {
    const float* positions;  int32_t positionStride;   /* bytes; 3 floats per vertex; required */
    const float* normals;    int32_t normalStride;     /* required */
    const float* texCoords;  int32_t texCoordStride;   /* optional (NULL) */
    const float* tangents;   int32_t tangentStride;    /* optional */
    const float* bitangents; int32_t bitangentStride;  /* optional */
    const float* colors;     int32_t colorStride;      /* optional */
    uint32_t vertexCount;
    const uint32_t* indices; /* already offset to the submesh's first element */
    uint32_t elementCount;   /* number of indices (Submesh::elementCount) */
    int32_t drawMode;        /* HRC_DRAW_* */
    float worldFromEntity[16]; /* submesh.localTransform * transform (Mesh.cpp:85) */
    int32_t frontFaceCW;     /* det(worldFromEntity) < 0 (Mesh.cpp:86-91) */
    int32_t isOccluder;      /* 0 for PBR + alphaMask (Mesh.cpp:95-100) */
    HrcMaterial material;
} HrcMeshDesc;
HRC_API int hrcCreateMesh(HrcContext ctx, const HrcMeshDesc* desc, HrcMesh* out);
HRC_API int hrcDestroyMesh(HrcContext ctx, HrcMesh mesh);
/* Replaces a mesh's worldFromEntity ↔ Scene::applyTransform (Scene.cpp:38-49: the "worldFromEntity" uniform is rewritten with
 * transform * submesh.transform; frontFaceCW stays what it was at creation). Takes effect at the next hrcCommitScene. */
HRC_API int hrcSetMeshTransform(HrcContext ctx, HrcMesh mesh, const float worldFromEntity[16]);
HRC_API int hrcClearScene(HrcContext ctx);  /* ↔ Scene::clearMeshesAndMaterials (PassGenerator.cpp:408) */
/* Flattens all meshes to world space (vertex.rlsl:25-43), builds the SAH BVH on the host,
 * collapses it to 8-wide compressed nodes and uploads. Explicit where OpenRL is implicit. */
HRC_API int hrcCommitScene(HrcContext ctx);

/* ---- lights (↔ the four light uniform blocks, ShaderLightingDefines.h:33-64) ------- */
typedef struct HrcEnvLight {
    int32_t enabled;             /* ↔ EnvironmentLight.lightPrimitive != rl_NullPrimitive */
    HrcTexture texture;          /* lat-long map or 1x1 solid colour */
    float exposureCompensation;  /* already 2^EV (EnvironmentLight.cpp:95) */
    float thetaRotation;         /* radians */
} HrcEnvLight;
typedef struct HrcDirLight   { float direction[3]; float color[3]; } HrcDirLight;   /* direction TO the light */
typedef struct HrcPointLight { float position[3];  float color[3]; } HrcPointLight;
typedef struct HrcSpotLight  { float position[3];  float direction[3]; float color[3]; float angles[2]; } HrcSpotLight; /* direction FROM the light; angles = cos(inner), cos(outer) */
HRC_API int hrcSetEnvironment(HrcContext ctx, const HrcEnvLight* env);
HRC_API int hrcSetDirectionalLights(HrcContext ctx, int32_t n, const HrcDirLight* lights);
HRC_API int hrcSetPointLights(HrcContext ctx, int32_t n, const HrcPointLight* lights);
HRC_API int hrcSetSpotLights(HrcContext ctx, int32_t n, const HrcSpotLight* lights);

/* ---- QMC tables + globals ---------------------------------------------------------- */
/* ↔ RandomSequences / RandomSequenceMetadata / ApertureSamples blocks (PassGenerator.cpp:603-684):
 * xy and apertureXY are vec2[numSequences*sequenceLength], sequence-major. */
HRC_API int hrcSetSequences(HrcContext ctx, int32_t numSequences, int32_t sequenceLength,
                            const float* xy, const float* apertureXY);
/* ↔ SequenceOffsets block (PassGenerator.cpp:150-159; only .x is read, perspective.rlsl:64) and the
 * per-pixel sequenceID = floor(random(FrameCoord/FrameSize)*numSequences) (perspective.rlsl:62),
 * which the host evaluates once (quirk B-Q2). offsetCount must cover the reference's index formula
 * int(FrameCoord.y*FrameSize.y + FrameCoord.x) for every pixel (quirk B-Q1). */
HRC_API int hrcSetPixelTables(HrcContext ctx, const float* seqOffsetX, size_t offsetCount,
                              const uint8_t* sequenceId /* w*h */);
/* ↔ the interactive-mode inputs of the frame shader (perspective.rlsl:42-57): the blockW x blockH RGB32F texture of shuffled
 * (row, col, 0) block coordinates (PassGenerator.cpp:266-294; NEAREST + REPEAT) and, per block of pixels, the two hash values
 * random(vec2(blockIndex)), random(vec2(blockIndex.yx)) the shader derives from `sin` — evaluated once on the host like the
 * per-pixel sequenceID (quirk B-Q2). blockRandXY holds blocksX*blocksY vec2, row-major by block row. */
HRC_API int hrcSetInteractiveBlocks(HrcContext ctx, int32_t blockW, int32_t blockH, const float* blockCoordsRGB,
                                    int32_t blocksX, int32_t blocksY, const float* blockRandXY);
/* ↔ PassGenerator::GlobalData (PassGenerator.h:269-294) / uniformblock Globals (globalData.rlsl:9-34): 84 bytes */
typedef struct HrcGlobals {
    int32_t maxRayDepth;
    int32_t sampleIndex;       /* overwritten per pass by HrcFrameParams.sampleIndex */
    float   maxChannelValue;
    int32_t enableDebugVisualizer;
    int32_t showGeometricNormals, showUVs, showTangents, showBitangents, showNormalmap;
    int32_t showFinalNormals, showBaseColor, showRoughness, showMetallic, showEmissive;
    int32_t showClearcoat, showClearcoatRoughness, showClearcoatNormalmap, showShader;
    int32_t enableAccumulatorVisualizer, showNANs, showInf;
} HrcGlobals;
HRC_API int hrcSetGlobals(HrcContext ctx, const HrcGlobals* globals);

/* ---- the pass (↔ frame uniforms + rlRenderFrame, PassGenerator.cpp:349-386) -------- */
typedef struct HrcFrameParams {
    float fovTan, aspectRatio, focusDistance, apertureRadius;   /* perspective.rlsl:15-18 */
    float viewMatrix[16];        /* camera-to-world (quirk B-Q19) */
    int32_t sampleIndex;         /* Globals.sampleIndex for this pass */
    float maxSampleIndex;        /* float(maxRenderPasses) (PassGenerator.cpp:368) */
    int32_t interactiveMode;     /* 1: render one pixel of every blockSize block per pass (needs hrcSetInteractiveBlocks) */
    int32_t blockSize[2], currentBlockPixelSample[2];
    /* screen-band sharding: this context renders rows y with (y / bandHeight) % bandCount == bandIndex.
     * bandCount <= 1 renders the whole frame. */
    int32_t bandHeight, bandIndex, bandCount;
} HrcFrameParams;
typedef struct HrcPassStats {
    float    gpuMs;                 /* CUDA-event time of the pass on the context's stream */
    uint64_t raysPrimary, raysExtension, raysOcclusion;   /* segments submitted to traversal */
    uint64_t nodesVisited, trianglesTested;               /* closest-hit kernel; only with HRC_OPT_TRAVERSAL_COUNTERS */
    uint64_t nodesVisitedAny, trianglesTestedAny;         /* occlusion kernel; only with HRC_OPT_TRAVERSAL_COUNTERS */
    uint64_t closestLaunches, anyLaunches, shadeLaunches; /* kernels launched by this pass */
    float    msClosest, msAny, msShade, msRaygen;         /* only with HRC_OPT_KERNEL_TIMING (per-kernel events, summed over the pass) */
    /* the primary-ray share of the closest-hit figures above (wavefront iteration 0), so that callers can report secondary
     * (extension) rays on their own: secondary = total - primary */
    float    msClosestPrimary; float reserved0;
    uint64_t nodesVisitedPrimary, trianglesTestedPrimary;
} HrcPassStats;
/* Enqueues and completes one pass; stats may be NULL. */
HRC_API int hrcRenderPass(HrcContext ctx, const HrcFrameParams* frame, HrcPassStats* stats);
/* Asynchronous variant: enqueue only (no host sync, no stats read-back). hrcSynchronize() waits.
 * hrcGetStats() returns the totals accumulated since the last hrcResetStats().
 * Consecutive enqueued passes are pipelined: they rotate over up to 8 internal sets of queues and accumulation buffers so that the
 * latency-bound last bounces of one pass overlap the start of the next (passes are independent given sampleIndex and the result is
 * a sum, accumulator.rlsl:27). Every other entry point folds those buffers into the one the caller sees first, so this is not
 * observable except through fp32 addition order; the result is deterministic for a given sequence of calls. */
HRC_API int hrcEnqueuePass(HrcContext ctx, const HrcFrameParams* frame);
/* A deferred pass renders like hrcEnqueuePass but into a side buffer that no other call sees until hrcCommitDeferred (then the next
 * read / resolve / pass folds it in). Up to HRC_MAX_DEFERRED_PASSES may be pending; hrcCommitDeferred commits the oldest one,
 * hrcDiscardDeferred drops all that are pending. It lets a caller that already knows its next passes start them before reading the
 * current result back: the host copy of pass i then overlaps the rendering of passes i+1 .. i+3. */
#define HRC_MAX_DEFERRED_PASSES 3
HRC_API int hrcEnqueuePassDeferred(HrcContext ctx, const HrcFrameParams* frame);
HRC_API int hrcCommitDeferred(HrcContext ctx);
HRC_API int hrcDiscardDeferred(HrcContext ctx);
HRC_API int hrcSynchronize(HrcContext ctx);
HRC_API int hrcResetStats(HrcContext ctx);
/* CUDA-event stopwatch on the context's stream: Start records, Stop records + waits and returns the device time between them. */
HRC_API int hrcTimerStart(HrcContext ctx);
HRC_API int hrcTimerStop(HrcContext ctx, float* milliseconds);
HRC_API int hrcGetStats(HrcContext ctx, HrcPassStats* stats);

/* ---- parity hooks (no reference equivalent; the fixed-ray-set test) ---- */
typedef struct HrcRay { float origin[3]; float tmin; float direction[3]; float tmax; } HrcRay; /* 32 B */
typedef struct HrcHit { float t, u, v; int32_t primId; /* -1 = miss; bit 30 set = back face */ } HrcHit; /* 16 B */
#define HRC_HIT_BACKFACE 0x40000000
/* Closest hit (anyHit = 0) or occlusion test (anyHit = 1: primId = 0 if occluded else -1) for n host rays. */
HRC_API int hrcTraceRays(HrcContext ctx, int32_t n, const HrcRay* rays, int32_t anyHit, HrcHit* hits);
/* hrcTraceRays for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * allocation). No counterpart in the reference. Same inputs, same outputs and the same
 * kernel as hrcTraceRays, but the device-side staging of rays and hits lives in a static device array inside the library (12 MiB on
 * every GPU that loads it, held until the process exits) instead of per-call allocations: the call allocates and frees nothing.
 * Limit: n <= HRC_NO_ALLOC_MAX_RAYS, more is HRC_ERR_INVALID_VALUE. All calls on one GPU share the array, which keeps the last call's
 * contents (nothing in it has to start at zero: every hit a call returns was written by that call). Host threads are serialised
 * inside the call, and it returns after its work has finished; anything that makes two of these calls overlap on the GPU races. */
#define HRC_NO_ALLOC_MAX_RAYS 262144
HRC_API int hrcTraceRaysNoAlloc(HrcContext ctx, int32_t n, const HrcRay* rays, int32_t anyHit, HrcHit* hits);
/* Rays per pixel since the last hrcClear (needs HRC_OPT_PIXEL_RAY_COUNTS): w*h counts, row 0 = bottom row like the accumulation buffer. */
HRC_API int hrcReadRayCounts(HrcContext ctx, uint32_t* hostOut);
/* Scene facts after commit. */
typedef struct HrcSceneInfo {
    uint64_t triangleCount, nodeCount;
    uint64_t nodeBytes, triangleBytes, shadingBytes;
    float    buildSeconds;
} HrcSceneInfo;
HRC_API int hrcGetSceneInfo(HrcContext ctx, HrcSceneInfo* info);

#ifdef __cplusplus
}
#endif
#endif /* HEATRAY_CUDA_H */

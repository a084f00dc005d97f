// device_types.h — data layout shared by the host side of the backend (hrc_api.cu) and the kernels.
// Everything here lives in HBM unless noted; see DESIGN.md "Data layout".
#pragma once
#include "../../include/heatray_cuda.h"
#include <cuda_runtime.h>
#include <cstdint>

namespace hr {

constexpr int kMaxIterations = 64;        // wavefront iterations per pass: maxRayDepth + 1 (+16 pass-through cap, policy A-14)
constexpr int kMaxLanes = 8;              // hrcEnqueuePass rotates consecutive passes over up to this many sets of queues / accumulation buffers
constexpr int kDeferredLanes = 4;         // side lanes of hrcEnqueuePassDeferred: up to kDeferredLanes - 1 passes pending at a time
constexpr int kTraceStack = 32;           // per-thread traversal stack entries (uint2)

// miss-shader selector carried by a ray (policy A-6)
enum : uint32_t { kMissNone = 0, kMissEnv = 1, kMissDir = 2, kMissPoint = 3, kMissSpot = 4 };

struct alignas(16) DMaterial {            // 96 B, one per HrcMaterial
    float baseColor[3]; uint32_t typeFlags;       // bits 0..3 type, bits 4.. feature flags (HRC_MAT_*)
    float emissive[3];  float metallic;
    float roughness, specularF0, roughnessAlpha, clearCoat;
    float clearCoatRoughness, clearCoatRoughnessAlpha, ior, density;
    int32_t lutTex, baseColorTex, metallicRoughnessTex, emissiveTex;      // texture handles, -1 = the 1x1 white dummy (Texture.h:178-192)
    int32_t normalTex, clearCoatTex, clearCoatRoughnessTex, clearCoatNormalTex;
};
constexpr uint32_t kMatAnyTexture = HRC_MAT_HAS_BASE_COLOR_TEXTURE | HRC_MAT_HAS_METALLIC_ROUGHNESS_TEXTURE | HRC_MAT_HAS_EMISSIVE_TEXTURE | HRC_MAT_HAS_NORMALMAP |
                                    HRC_MAT_HAS_CLEARCOAT_TEXTURE | HRC_MAT_HAS_CLEARCOAT_ROUGHNESS_TEXTURE | HRC_MAT_HAS_CLEARCOAT_NORMALMAP;   // HAS_TEXTURES
constexpr uint32_t kMatTangentSpace = HRC_MAT_HAS_NORMALMAP | HRC_MAT_HAS_CLEARCOAT_NORMALMAP;                                                   // USE_TANGENT_SPACE
constexpr int kAttrStride = 11;           // float4 per triangle in DScene::attr
constexpr int kMaxMaskedSurfaces = 16;    // alpha-mask pass-throughs per occlusion ray (policy A-14)
constexpr uint32_t kTriNonOccluder = 4u;  // bit 2 of the triangle record's e1.w (bits 0..1: material type)

struct DTexture {                          // all textures are expanded to RGBA32F
    const float4* texels; int32_t w, h, wrapS, wrapT;
};

struct DLights {
    int32_t nDir, nPoint, nSpot, envEnabled;
    float dirDirection[HRC_MAX_LIGHTS_PER_TYPE][3], dirColor[HRC_MAX_LIGHTS_PER_TYPE][3];
    float pointPosition[HRC_MAX_LIGHTS_PER_TYPE][3], pointColor[HRC_MAX_LIGHTS_PER_TYPE][3];
    float spotPosition[HRC_MAX_LIGHTS_PER_TYPE][3], spotDirection[HRC_MAX_LIGHTS_PER_TYPE][3], spotColor[HRC_MAX_LIGHTS_PER_TYPE][3];
    float spotAngles[HRC_MAX_LIGHTS_PER_TYPE][2];
    int32_t envTex; float envExposure, envTheta; int32_t pad;
};

// Per-pass constants: uploaded into a constant-memory slot per lane (kernels.cu g_frames), kernels are passed the slot number.
struct DFrame {
    HrcGlobals g;
    float fovTan, aspectRatio, focusDistance, apertureRadius;
    float view[16];
    float maxSampleIndex;
    int32_t w, h;
    int32_t bandHeight, bandIndex, bandCount, localRows;   // screen-band sharding
    int32_t numSequences, sequenceLength;
    uint32_t offsetCount;
    int32_t interactiveMode, blockW, blockH, blockSampleX, blockSampleY, blocksX;   // perspective.rlsl:20-22,42-57
    int32_t rayTileW, rayTileH;        // k_raygen: a warp's 32 primary rays cover a rayTileW x rayTileH pixel tile (0: 32 consecutive pixels of a row)
    DLights lights;
};

// Wavefront queues (structure of arrays; one slot per in-flight ray).
// Extension rays are two arrays of 32-byte records (one L2 sector each), because the shade kernels gather them through
// the per-class index lists while the traversal kernel streams only the first one:
struct RayQueue {
    float4* od;    // [2i]   origin.xyz, tmax            [2i+1] direction.xyz, extraT
    float4* wm;    // [2i]   weight.xyz, pixel (bits)    [2i+1] x: depth | missKind<<8 | sequenceID<<16 (bits), y: sequenceIndexOffset (bits)
};
struct ShadowQueue {
    float4* o;     // origin.xyz, tmax
    float4* d;     // direction.xyz, extraT
    float4* w;     // weight.xyz, pixel (24 bits) | missKind<<24 | missIndex<<27
};

// Per-iteration control words, zeroed at the start of every pass.
struct IterCtrl { uint32_t extCount, shadowCount, extHead, shadowHead; uint32_t classCount[3]; uint32_t maskedContinuations; uint32_t litCount; uint32_t pad[3]; };   // class 0: miss, 1: PBR, 2: glass; litCount: occlusion rays that reached their light
// One more entry behind the per-iteration ones describes the pass's tail (k_tail_decide / k_path): extCount = rays the path kernel
// takes over (0: the wavefront runs to its end), extHead = the iteration whose extension queue they sit in.
constexpr int kCtrlEntries = kMaxIterations + 1;
static_assert(kCtrlEntries <= 128, "k_clear_ctrl zeroes the control words with one 128-thread block");
constexpr uint32_t kQueueDrained = 0x80000000u;          // extHead of an iteration the path kernel took: k_trace finds nothing to pull

// Coherence sort of the extension rays between wavefront iterations (HRC_TUNE_SORT_BITS > 0): k_shade files every emitted ray under
// key = octant << 3b | morton(cell of its origin in a 2^b-per-axis grid over the scene box), k_sort_scan turns the histogram into bin
// offsets, k_sort_scatter writes the ray indices bin by bin into `order`, and k_trace<closest> pulls rays through `order`, so the 32 rays
// a warp refills with start in the same part of the scene heading the same way. Ray records stay where k_shade wrote them.
constexpr int kSortMaxBits = 5;                          // cells per axis = 2^bits; bins = 8 octants * 2^(3 bits)
constexpr uint32_t kSortMinRays = 32768;                 // below this the iteration is latency-bound: not sorted
struct SortCtrl { uint32_t sorted; uint32_t pad[3]; };   // per iteration: 1 when `order` is valid for its extension rays

struct DeviceStats {                       // accumulated by k_finish_pass
    unsigned long long raysPrimary, raysExtension, raysOcclusion, nodesVisited, trianglesTested, nodesVisitedAny, trianglesTestedAny, stackOverflows;
    unsigned long long nodesVisitedPrimary, trianglesTestedPrimary;      // iteration 0 of the closest-hit kernel (included in nodesVisited / trianglesTested)
};

struct DScene {
    const float4* nodes;      // 5 x float4 per wide node
    const float4* tris;       // 3 x float4 per triangle, leaf order: (v0, primId) (e1, material type) (e2, 0)
    const float4* shade;      // 4 x float4 per triangle, primId order: (n0, material) (n1, frontFaceCW) (n2, 0) (Ng, 0)
    const float4* attr;       // optional varyings, kAttrStride x float4 per triangle, primId order (nullptr when no mesh has any):
                              // (uv0, uv1) (uv2, -, -) (tangent0..2) (bitangent0..2) (colour0..2)
    const DMaterial* materials;
    const DTexture* textures;
    const float2* seq;        // RandomSequences
    const float2* aperture;   // ApertureSamples
    const float* seqOffsetX;  // SequenceOffsets.x
    const uint8_t* sequenceId;
    const float* blockCoords; // interactive mode: blockW*blockH RGB texels of shuffled (row, col, 0)
    const float2* blockRand;  // interactive mode: per block (random(blockIndex), random(blockIndex.yx))
    float sortMin[3], sortScale[3];   // coherence sort grid: cell = (origin - sortMin) * sortScale, clamped to [0, 2^bits)
    uint32_t nodeCount, pad0;         // wide BVH nodes
};

// Launch wrappers implemented in kernels.cu
struct PassBuffers {
    RayQueue ext[2];
    ShadowQueue shadow;
    float4* hits;             // t, u, v, primId | material class << 28 | HRC_HIT_BACKFACE  (-1: miss)
    uint32_t* classIdx[3];    // per-class lists of ray indices of the current iteration, filled by k_trace<closest>
    uint32_t* litIdx;         // occlusion rays of the current iteration that reached their light, filled by k_trace<any> for k_shade_lights (nullptr: the light shaders run inside k_trace<any>)
    IterCtrl* ctrl;           // kCtrlEntries entries
    DeviceStats* stats;
    float4* accum;
    uint32_t capacity;        // rays per queue
    // coherence sort (nullptr when off)
    uint32_t* sortKeys;       // key of extension ray j of the NEXT iteration
    uint32_t* sortOrder;      // sorted position -> ray index
    uint32_t* sortHist;       // bins counters (zeroed by k_sort_scan after use)
    uint32_t* sortOffs;       // bins running offsets
    SortCtrl* sortCtrl;       // kMaxIterations entries
    uint32_t* rayCounts;      // HRC_OPT_PIXEL_RAY_COUNTS: rays per pixel (shared by all lanes, hence atomics), else nullptr
};

struct LaunchConfig { int smCount; int traceBlocksPerSM; int traceThreads; bool collectStats; bool masked; bool shadeMiss, shadePbr, shadeGlass; int refillIdle, refillIdleAny; int sortBits; int tailRays; };

constexpr int kFrameSlots = 1 + (kMaxLanes - 1) + kDeferredLanes;     // one constant-memory DFrame per lane: main, extra lanes, deferred lanes
void uploadFrame(const DFrame* pinnedFrame, int slot, cudaStream_t s);  // stream-ordered copy into the lane's slot; the source must stay valid until it ran
void launchClearCtrl(const PassBuffers& pb, cudaStream_t s);
void launchRaygen(const DScene& sc, int slot, uint32_t total, const PassBuffers& pb, cudaStream_t s);
void launchTraceClosest(const DScene& sc, int slot, const PassBuffers& pb, int iter, const LaunchConfig& lc, cudaStream_t s);
void launchShade(const DScene& sc, int slot, const PassBuffers& pb, int iter, const LaunchConfig& lc, cudaStream_t s);
void launchTraceAny(const DScene& sc, int slot, const PassBuffers& pb, int iter, const LaunchConfig& lc, cudaStream_t s);
void launchShadeLights(const DScene& sc, int slot, const PassBuffers& pb, int iter, const LaunchConfig& lc, cudaStream_t s);   // after launchTraceAny, same stream (no-op without pb.litIdx)
void launchSortRays(const PassBuffers& pb, int iter, const LaunchConfig& lc, cudaStream_t s);   // orders the extension rays of iteration `iter` (emitted by shade(iter - 1))
// Tail of a pass (lc.tailRays > 0): once an iteration's extension queue holds no more than lc.tailRays rays, ONE launch of k_path follows every
// remaining path to its end (trace, shade, occlusion ray, next bounce, in registers) and the wavefront kernels behind it find their queues empty.
void launchTailDecide(const PassBuffers& pb, int iter, const LaunchConfig& lc, cudaStream_t s);          // after shade(iter - 1) (or raygen), same stream
void launchPathTail(const DScene& sc, int slot, const PassBuffers& pb, int iter, int iterations, const LaunchConfig& lc, cudaStream_t s);   // after any(iter - 1)
void launchFinishPass(const PassBuffers& pb, int iterations, bool rearm, cudaStream_t s);   // rearm: also zero the control words for the next pass
void launchTraceRays(const DScene& sc, const HrcRay* rays, HrcHit* hits, int n, bool anyHit, const LaunchConfig& lc, cudaStream_t s);
void launchAddPixels(float4* accum, const float4* src, size_t n, cudaStream_t s);
// hrcTraceRaysNoAlloc: staging buffers for up to kNoAllocMaxRays rays / hits inside a static __device__ array (kernels.cu); `device` must be current.
constexpr int kNoAllocMaxRays = HRC_NO_ALLOC_MAX_RAYS;
cudaError_t traceScratchBuffers(int device, HrcRay** rays, HrcHit** hits);
void launchResolve(const float4* accum, const HrcPostProcess& post, int format, int w, int h, void* out, cudaStream_t s);   // resolve.cu

} // namespace hr

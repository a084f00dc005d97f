// hrc_api.cu — implementation of the backend C-ABI (include/heatray_cuda.h) over CUDA device memory.
// Reference-side counterparts are cited in the header next to each entry point.
#include <mutex>
#include "device_types.h"
#include "bvh_builder.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

using namespace hr;

namespace {

// One captured pass: every kernel argument of a pass is fixed by these values (the per-pass constants travel through a constant-memory
// slot), so a pass whose key matches replays the instantiated graph: one upload + one graph launch instead of ~40 launches + ~36 event calls.
struct GraphKey { DScene sc; PassBuffers pb; LaunchConfig lc; int32_t iterations; uint32_t total; int32_t slot; int32_t overlap; };
struct GraphEntry { GraphKey key; cudaGraphExec_t exec = nullptr; uint64_t lastUse = 0; };
constexpr int kFrameRing = 32;            // pinned staging entries for the per-pass constants: the host runs at most this many passes ahead of the device

struct TextureRec { float4* dev = nullptr; int w = 0, h = 0, wrapS = 0, wrapT = 0; bool alive = false; };
struct MaterialRec { DMaterial d; int type; uint32_t flags; };
struct MeshRec {
    bool alive = false;
    std::vector<float> pos, nrm;       // entity space, tightly packed xyz
    std::vector<float> uv, tan, bit, col;   // optional attributes (empty when absent): 2 / 3 / 3 / 3 floats per vertex
    std::vector<uint32_t> idx;
    int drawMode = 0; float M[16]; int frontFaceCW = 0, isOccluder = 1, material = 0;
};

thread_local std::string g_err;

} // namespace

struct HrcContext_t {
    int device = 0;
    bool counters = false, kernelTiming = false;
    cudaStream_t stream = nullptr, stream2 = nullptr;   // stream2: occlusion traversal of iteration i overlaps closest-hit traversal of i+1
    std::vector<cudaEvent_t> evSync;                     // 2 per iteration, timing disabled
    bool overlap = true;
    cudaEvent_t evStart = nullptr, evStop = nullptr, evTimer0 = nullptr, evTimer1 = nullptr;
    std::vector<cudaEvent_t> evPool;
    std::string err;
    int smCount = 148;
    int tuneRefillIdle = 10, tuneRefillIdleAny = 10, tuneBlocksPerSM = 8;
    // Extra lanes for hrcEnqueuePass: consecutive enqueued passes rotate over passLanes sets of queues / streams / accumulation buffers, so
    // the latency-bound tail of one pass (a few thousand rays per kernel) runs next to the head of the following one. Every other entry
    // point first joins the lanes (accum += the extra lanes' buffers), so callers see one buffer; the result is deterministic for a given call sequence.
    struct Lane { PassBuffers pb{}; cudaStream_t s = nullptr, s2 = nullptr; std::vector<cudaEvent_t> evSync; float4* accum = nullptr;
                  cudaEvent_t evDone = nullptr; bool active = false; std::vector<GraphEntry> graphs; } extra[kMaxLanes - 1];      // lanes 1 .. passLanes-1; lane 0 is the main set above
    // Speculative look-ahead (HRC_OPT_SPECULATIVE_PASSES = n > 0): hrcRenderPass(i) also starts passes i+1 .. i+n of the same progressive
    // render in the deferred lanes. If the caller's next hrcRenderPass asks for exactly what was predicted (same frame parameters, next
    // sample index, no state change in between) the pass is already rendered or under way and is simply committed; anything else discards
    // the predictions. A caller that renders and reads back one pass at a time (the reference viewer's loop) then gets its readback
    // overlapped with the next passes without queueing anything itself.
    uint32_t* dRayCounts = nullptr;        // HRC_OPT_PIXEL_RAY_COUNTS
    int specDepth = 0, specStride = 1; uint64_t stateVersion = 0;
    HrcFrameParams deferredParams[kDeferredLanes] = {}; uint64_t deferredVersion[kDeferredLanes] = {};
    std::vector<GraphEntry> graphs;        // captured passes of the main lane
    bool useGraphs = true; uint64_t graphClock = 0;
    DFrame* frameRing = nullptr; cudaEvent_t frameEv[kFrameRing] = {}; bool frameUsed[kFrameRing] = {}; int frameNext = 0;
    int passLanes = 8, nextLane = 0;
    // Deferred passes (hrcEnqueuePassDeferred): rendered into side lanes that joins ignore until hrcCommitDeferred, so a caller can
    // already render passes i+1 .. i+3 while pass i is being read back (hr::PassGenerator does that when its next job is queued).
    // A ring: the deferredCount lanes before deferredNext are pending, oldest first; one lane more than may be pending, so the lane
    // a new pass takes is either free or holds a committed pass that the next join folds in.
    Lane deferred[kDeferredLanes]; int deferredNext = 0, deferredCount = 0;
    int oldestDeferred() const { return (deferredNext - deferredCount + kDeferredLanes) % kDeferredLanes; }
    // Band bookkeeping for the readback: while every pass since the last clear rendered the same set of screen bands, pixels outside
    // them are zero on the device and hrcReadPixels copies only the context's own rows.
    int rbBandHeight = 0, rbBandIndex = 0, rbBandCount = 0; bool accumOutsideBandsZero = true;
    bool devicePixelsExported = false;     // hrcGetDevicePixels handed the buffer out: the caller may write any row at any time (in-place NCCL reduce), so reads copy whole frames until hrcResize
    long long pinnedZeroKey = -1;          // which rows of the pinned frame may be non-zero: -1 any, 0 none (fresh), else the band set's key
    bool tuneDeferLights = true;           // HRC_TUNE_DEFER_LIGHTS: occlusion rays that reach their light are listed by k_trace<any> and shaded by k_shade_lights (0: inline in k_trace<any>)
    int tuneRayTileW = 4;                  // HRC_TUNE_RAY_TILE_W: width of the pixel tile of a warp's 32 primary rays (8 -> 8 x 4, 4 -> 4 x 8, 0 -> a 32 x 1 strip)
    int tuneTailRays = 0;                  // HRC_TUNE_TAIL_RAYS: an iteration with at most this many extension rays hands the rest of the pass to k_path. Measured: no gain at any threshold (profiles/r02_notes.md), so 0 = off
    int tuneSortBits = 0;                  // HRC_TUNE_SORT_BITS: cells per axis (2^bits) of the coherence sort grid, 0 = off
    float sceneMin[3] = {0, 0, 0}, sceneMax[3] = {1, 1, 1};
    int tunePipelinedBlocksPerSM = 4;      // traversal CTAs per SM for pipelined passes: kernels of several lanes are resident together, so each takes fewer slots
    cudaEvent_t evMerged = nullptr;   // traversal scheduling knobs (HRC_TUNE_* environment overrides, for experiments)

    int w = 0, h = 0;
    float4* accum = nullptr;
    float* pinned = nullptr; size_t pinnedBytes = 0; bool mapped = false;
    float4* staging = nullptr;        // device staging for hrcAddPixels
    void* resolveDev = nullptr; size_t resolveBytes = 0;   // device output of hrcResolve

    std::vector<TextureRec> textures;
    std::vector<MaterialRec> materials;
    std::vector<MeshRec> meshes;
    bool committed = false, materialsDirty = true, texturesDirty = true;
    bool anySingleSided = false, anyPbr = false, anyGlass = false;

    float4 *dNodes = nullptr, *dTris = nullptr, *dShade = nullptr, *dAttr = nullptr;
    bool anyMasked = false;
    DMaterial* dMaterials = nullptr; size_t dMaterialsCap = 0;
    DTexture* dTextures = nullptr; size_t dTexturesCap = 0;
    float2 *dSeq = nullptr, *dAperture = nullptr; float* dSeqOffsetX = nullptr; uint8_t* dSequenceId = nullptr;
    int numSequences = 0, sequenceLength = 0; size_t offsetCount = 0; bool pixelTables = false;
    float* dBlockCoords = nullptr; float2* dBlockRand = nullptr; int blockW = 0, blockH = 0, blocksX = 0, blocksY = 0;

    PassBuffers pb{};
    DLights lights{};
    HrcGlobals globals{};
    HrcSceneInfo info{};
    DeviceStats lastStats{};
};

namespace {

int fail(HrcContext c, int code, const std::string& msg) { if (c) c->err = msg; else g_err = msg; return code; }
int failCuda(HrcContext c, cudaError_t e, const char* what) {
    return fail(c, e == cudaErrorMemoryAllocation ? HRC_ERR_OUT_OF_MEMORY : HRC_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}
#define CU(c, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return failCuda(c, e_, #call); } while (0)

template <class T> void freeDev(T*& p) { if (p) { cudaFree(p); p = nullptr; } }

void destroyGraphList(std::vector<GraphEntry>& v) { for (auto& g : v) if (g.exec) cudaGraphExecDestroy(g.exec); v.clear(); }
void destroyGraphs(HrcContext c) {
    destroyGraphList(c->graphs);
    for (auto& L : c->extra) destroyGraphList(L.graphs);
    for (auto& L : c->deferred) destroyGraphList(L.graphs);
}

void freeQueues(PassBuffers& pb) {
    for (int k = 0; k < 2; ++k) { freeDev(pb.ext[k].od); freeDev(pb.ext[k].wm); }
    for (int k = 0; k < 3; ++k) freeDev(pb.classIdx[k]);
    freeDev(pb.litIdx);
    freeDev(pb.shadow.o); freeDev(pb.shadow.d); freeDev(pb.shadow.w); freeDev(pb.hits);
    freeDev(pb.sortKeys); freeDev(pb.sortOrder); freeDev(pb.sortHist); freeDev(pb.sortOffs); freeDev(pb.sortCtrl);
    pb.capacity = 0;
}
int allocQueues(HrcContext c, PassBuffers& pb, size_t n) {
    for (int k = 0; k < 2; ++k) {
        CU(c, cudaMalloc(&pb.ext[k].od, 2 * n * sizeof(float4))); CU(c, cudaMalloc(&pb.ext[k].wm, 2 * n * sizeof(float4)));
    }
    for (int k = 0; k < 3; ++k) CU(c, cudaMalloc(&pb.classIdx[k], n * sizeof(uint32_t)));
    if (c->tuneDeferLights) CU(c, cudaMalloc(&pb.litIdx, n * sizeof(uint32_t)));
    CU(c, cudaMalloc(&pb.shadow.o, n * sizeof(float4))); CU(c, cudaMalloc(&pb.shadow.d, n * sizeof(float4)));
    CU(c, cudaMalloc(&pb.shadow.w, n * sizeof(float4))); CU(c, cudaMalloc(&pb.hits, n * sizeof(float4)));
    if (c->tuneSortBits > 0) {                       // coherence sort of the extension rays (device_types.h)
        const size_t bins = std::max<size_t>(1024, (size_t)8 << (3 * c->tuneSortBits));
        CU(c, cudaMalloc(&pb.sortKeys, n * sizeof(uint32_t))); CU(c, cudaMalloc(&pb.sortOrder, n * sizeof(uint32_t)));
        CU(c, cudaMalloc(&pb.sortHist, bins * sizeof(uint32_t))); CU(c, cudaMalloc(&pb.sortOffs, bins * sizeof(uint32_t)));
        CU(c, cudaMalloc(&pb.sortCtrl, kMaxIterations * sizeof(SortCtrl)));
        CU(c, cudaMemset(pb.sortHist, 0, bins * sizeof(uint32_t))); CU(c, cudaMemset(pb.sortCtrl, 0, kMaxIterations * sizeof(SortCtrl)));
    }
    pb.capacity = (uint32_t)n;
    return HRC_OK;
}
// Folds the extra lanes into the main one, in stream order on the main stream: after this, work on c->stream sees (and a synchronize
// on it covers) everything that was enqueued on any lane.
void joinLanes(HrcContext c) {
    c->nextLane = 0;                       // a join between two enqueued passes keeps both on the main lane
    const size_t n = (size_t)c->w * c->h;
    bool any = false;
    auto fold = [&](HrcContext_t::Lane& L) {
        if (!L.active) return;
        cudaEventRecord(L.evDone, L.s);
        cudaStreamWaitEvent(c->stream, L.evDone, 0);
        launchAddPixels(c->accum, L.accum, n, c->stream);
        cudaMemsetAsync(L.accum, 0, n * sizeof(float4), c->stream);
        any = true;
    };
    for (auto& L : c->extra) fold(L);
    for (int k = 0; k < kDeferredLanes; ++k) fold(c->deferred[(c->deferredNext + k) % kDeferredLanes]);   // committed deferred passes only, in commit order (ring order from the next free lane); pending ones keep running, invisible
    if (!any) return;
    cudaEventRecord(c->evMerged, c->stream);
    auto release = [&](HrcContext_t::Lane& L) { if (L.active) { cudaStreamWaitEvent(L.s, c->evMerged, 0); L.active = false; } };
    for (auto& L : c->extra) release(L);
    for (auto& L : c->deferred) release(L);
}
// For calls that free or rewrite what a pending deferred pass still reads (resize, scene / table uploads): wait for it.
void waitPendingDeferred(HrcContext c) {
    for (int k = 0, d = c->oldestDeferred(); k < c->deferredCount; ++k, d = (d + 1) % kDeferredLanes)
        if (c->deferred[d].s) cudaStreamSynchronize(c->deferred[d].s);
}
void freeExtraLanes(HrcContext c) {         // buffers that depend on the frame size; streams and counters stay
    for (auto& L : c->extra) { freeQueues(L.pb); freeDev(L.accum); L.active = false; destroyGraphList(L.graphs); }
    for (auto& L : c->deferred) { freeQueues(L.pb); freeDev(L.accum); L.active = false; destroyGraphList(L.graphs); }
    c->nextLane = 0; c->deferredCount = 0;
}
int ensureLane(HrcContext c, HrcContext_t::Lane& L) {
    if (L.accum) return HRC_OK;
    const size_t n = (size_t)c->w * c->h;
    if (!L.s) {
        CU(c, cudaStreamCreateWithFlags(&L.s, cudaStreamNonBlocking)); CU(c, cudaStreamCreateWithFlags(&L.s2, cudaStreamNonBlocking));
        for (int i = 0; i < 2 * kMaxIterations; ++i) { cudaEvent_t ev; CU(c, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming)); L.evSync.push_back(ev); }
        CU(c, cudaEventCreateWithFlags(&L.evDone, cudaEventDisableTiming));
        if (!c->evMerged) CU(c, cudaEventCreateWithFlags(&c->evMerged, cudaEventDisableTiming));
        CU(c, cudaMalloc(&L.pb.ctrl, sizeof(IterCtrl) * kCtrlEntries)); CU(c, cudaMalloc(&L.pb.stats, sizeof(DeviceStats)));
        CU(c, cudaMemset(L.pb.stats, 0, sizeof(DeviceStats))); CU(c, cudaMemset(L.pb.ctrl, 0, sizeof(IterCtrl) * kCtrlEntries));
    }
    auto allocate = [&]() -> int {
        CU(c, cudaMalloc(&L.accum, n * sizeof(float4)));
        CU(c, cudaMemset(L.accum, 0, n * sizeof(float4)));
        return allocQueues(c, L.pb, n);
    };
    if (int r = allocate()) { freeQueues(L.pb); freeDev(L.accum); return r; }      // a retry starts from scratch instead of finding a lane without queues
    L.pb.accum = L.accum; L.pb.rayCounts = c->dRayCounts;
    return HRC_OK;
}

DScene sceneOf(HrcContext c) {
    DScene s{}; s.nodes = c->dNodes; s.tris = c->dTris; s.shade = c->dShade; s.attr = c->dAttr; s.materials = c->dMaterials; s.textures = c->dTextures;
    s.seq = c->dSeq; s.aperture = c->dAperture; s.seqOffsetX = c->dSeqOffsetX; s.sequenceId = c->dSequenceId; s.blockCoords = c->dBlockCoords; s.blockRand = c->dBlockRand;
    s.nodeCount = (uint32_t)c->info.nodeCount;
    for (int k = 0; k < 3; ++k) {
        const float extent = c->sceneMax[k] - c->sceneMin[k];
        s.sortMin[k] = c->sceneMin[k]; s.sortScale[k] = extent > 0.0f ? (float)(1 << c->tuneSortBits) / extent : 0.0f;
    }
    return s;
}
LaunchConfig configOf(HrcContext c) {
    LaunchConfig lc{}; lc.smCount = c->smCount; lc.traceBlocksPerSM = 8; lc.traceThreads = 128; lc.collectStats = c->counters;
    lc.refillIdle = c->tuneRefillIdle; lc.refillIdleAny = c->tuneRefillIdleAny; lc.traceBlocksPerSM = c->tuneBlocksPerSM;
    lc.masked = c->anyMasked; lc.sortBits = c->tuneSortBits;
    lc.tailRays = c->counters ? 0 : c->tuneTailRays;          // with traversal counters on every ray goes through the instrumented wavefront kernels
    lc.shadeMiss = c->lights.envEnabled != 0; lc.shadePbr = c->anyPbr; lc.shadeGlass = c->anyGlass;
    return lc;
}

int uploadMaterials(HrcContext c) {
    if (!c->materialsDirty) return HRC_OK;
    size_t n = std::max<size_t>(1, c->materials.size());
    if (n > c->dMaterialsCap) { freeDev(c->dMaterials); CU(c, cudaMalloc(&c->dMaterials, n * sizeof(DMaterial))); c->dMaterialsCap = n; }
    std::vector<DMaterial> host(n);
    c->anySingleSided = c->anyPbr = c->anyGlass = false;
    for (size_t i = 0; i < c->materials.size(); ++i) {
        host[i] = c->materials[i].d;
        if (c->materials[i].type == HRC_MATERIAL_PBR) c->anyPbr = true; else c->anyGlass = true;
        if (c->materials[i].type == HRC_MATERIAL_PBR && !(c->materials[i].flags & HRC_MAT_DOUBLE_SIDED)) c->anySingleSided = true;
    }
    CU(c, cudaMemcpyAsync(c->dMaterials, host.data(), n * sizeof(DMaterial), cudaMemcpyHostToDevice, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    c->materialsDirty = false;
    return HRC_OK;
}
int uploadTextures(HrcContext c) {
    if (!c->texturesDirty) return HRC_OK;
    size_t n = std::max<size_t>(1, c->textures.size());
    if (n > c->dTexturesCap) { freeDev(c->dTextures); CU(c, cudaMalloc(&c->dTextures, n * sizeof(DTexture))); c->dTexturesCap = n; }
    std::vector<DTexture> host(n);
    for (size_t i = 0; i < c->textures.size(); ++i) { const TextureRec& t = c->textures[i]; host[i] = DTexture{t.dev, t.w, t.h, t.wrapS, t.wrapT}; }
    CU(c, cudaMemcpyAsync(c->dTextures, host.data(), n * sizeof(DTexture), cudaMemcpyHostToDevice, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    c->texturesDirty = false;
    return HRC_OK;
}

int packMaterial(HrcContext c, MaterialRec& m, int type, const void* params, size_t bytes) {
    DMaterial d; std::memset(&d, 0, sizeof(d));
    if (type == HRC_MATERIAL_PBR) {
        if (bytes != sizeof(HrcPbrParams)) return fail(c, HRC_ERR_INVALID_VALUE, "PBR parameter block has the wrong size");
        const HrcPbrParams* p = (const HrcPbrParams*)params;
        for (int k = 0; k < 3; ++k) { d.baseColor[k] = p->baseColor[k]; d.emissive[k] = p->emissiveColor[k]; }
        d.metallic = p->metallic; d.roughness = p->roughness; d.specularF0 = p->specularF0; d.roughnessAlpha = p->roughnessAlpha;
        d.clearCoat = p->clearCoat; d.clearCoatRoughness = p->clearCoatRoughness; d.clearCoatRoughnessAlpha = p->clearCoatRoughnessAlpha;
        d.lutTex = p->multiscatterLUT;
        d.baseColorTex = p->baseColorTexture; d.metallicRoughnessTex = p->metallicRoughnessTexture; d.emissiveTex = p->emissiveTexture;
        d.normalTex = p->normalmap; d.clearCoatTex = p->clearCoatTexture; d.clearCoatRoughnessTex = p->clearCoatRoughnessTexture;
        d.clearCoatNormalTex = p->clearCoatNormalmap;
    } else if (type == HRC_MATERIAL_GLASS) {
        if (bytes != sizeof(HrcGlassParams)) return fail(c, HRC_ERR_INVALID_VALUE, "glass parameter block has the wrong size");
        const HrcGlassParams* p = (const HrcGlassParams*)params;
        for (int k = 0; k < 3; ++k) d.baseColor[k] = p->baseColor[k];
        d.roughness = p->roughness; d.roughnessAlpha = p->roughnessAlpha; d.ior = p->ior; d.density = p->density; d.specularF0 = p->specularF0;
        d.lutTex = -1;
        d.baseColorTex = p->baseColorTexture; d.metallicRoughnessTex = p->metallicRoughnessTexture; d.normalTex = p->normalmap;
        d.emissiveTex = d.clearCoatTex = d.clearCoatRoughnessTex = d.clearCoatNormalTex = -1;
    } else return fail(c, HRC_ERR_INVALID_VALUE, "unknown material type");
    for (int32_t* h : {&d.lutTex, &d.baseColorTex, &d.metallicRoughnessTex, &d.emissiveTex, &d.normalTex, &d.clearCoatTex, &d.clearCoatRoughnessTex, &d.clearCoatNormalTex}) {
        if (*h < 0) { *h = -1; continue; }
        if (*h >= (int)c->textures.size() || !c->textures[*h].alive) return fail(c, HRC_ERR_INVALID_VALUE, "bad material texture handle");
    }
    if (type == HRC_MATERIAL_GLASS && (m.flags & ~(uint32_t)(HRC_MAT_HAS_BASE_COLOR_TEXTURE | HRC_MAT_HAS_METALLIC_ROUGHNESS_TEXTURE | HRC_MAT_HAS_NORMALMAP)))
        return fail(c, HRC_ERR_INVALID_VALUE, "glass materials take only the base-colour, metallic-roughness and normal-map flags (GlassMaterial.cpp:40-74)");
    d.typeFlags = (uint32_t)type | (m.flags << 4);
    m.d = d; m.type = type;
    return HRC_OK;
}

// vertex.rlsl:25-43 on the host; glm evaluates mat4*vec4 as (m0*x + m1*y) + (m2*z + m3*w)
inline void xformPoint(const float* M, const float* p, float* o) {
    o[0] = (M[0] * p[0] + M[4] * p[1]) + (M[8] * p[2] + M[12] * 1.0f);
    o[1] = (M[1] * p[0] + M[5] * p[1]) + (M[9] * p[2] + M[13] * 1.0f);
    o[2] = (M[2] * p[0] + M[6] * p[1]) + (M[10] * p[2] + M[14] * 1.0f);
}
inline void xformNormal(const float* M, const float* n, float* o) {   // mat3(worldFromEntity) * n (quirk B-Q7)
    o[0] = (M[0] * n[0] + M[4] * n[1]) + M[8] * n[2];
    o[1] = (M[1] * n[0] + M[5] * n[1]) + M[9] * n[2];
    o[2] = (M[2] * n[0] + M[6] * n[1]) + M[10] * n[2];
}

} // namespace

extern "C" {

HRC_API int hrcCreateContext(const HrcContextDesc* desc, HrcContext* out) {
    if (!out) return fail(nullptr, HRC_ERR_INVALID_VALUE, "null out pointer");
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(nullptr, HRC_ERR_NO_DEVICE, std::string("no CUDA device available (this backend has no CPU path): ") + cudaGetErrorString(e));
    int dev = desc ? desc->device : 0;
    if (dev < 0 || dev >= count) return fail(nullptr, HRC_ERR_INVALID_VALUE, "bad CUDA device ordinal");
    HrcContext c = new HrcContext_t();
    c->device = dev; c->counters = c->kernelTiming = desc && desc->collectStats;
    if ((e = cudaSetDevice(dev)) != cudaSuccess) { int r = failCuda(nullptr, e, "cudaSetDevice"); delete c; return r; }
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, dev)) != cudaSuccess) { int r = failCuda(nullptr, e, "cudaGetDeviceProperties"); delete c; return r; }
    c->smCount = prop.multiProcessorCount;
    if ((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess) { int r = failCuda(nullptr, e, "cudaStreamCreate"); delete c; return r; }
    cudaStreamCreateWithFlags(&c->stream2, cudaStreamNonBlocking);
    for (int i = 0; i < 2 * kMaxIterations; ++i) { cudaEvent_t ev; cudaEventCreateWithFlags(&ev, cudaEventDisableTiming); c->evSync.push_back(ev); }
    if (const char* e2 = std::getenv("HRC_TUNE_OVERLAP")) c->overlap = std::atoi(e2) != 0;
    if (const char* e2 = std::getenv("HRC_TUNE_GRAPHS")) c->useGraphs = std::atoi(e2) != 0;      // 0: launch every kernel of a pass individually (experiments)
    if (cudaMallocHost(&c->frameRing, sizeof(DFrame) * kFrameRing) != cudaSuccess) { int r = fail(nullptr, HRC_ERR_OUT_OF_MEMORY, "pinned frame ring"); delete c; return r; }
    for (auto& ev : c->frameEv) cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
    if (const char* e2 = std::getenv("HRC_TUNE_DEFER_LIGHTS")) c->tuneDeferLights = std::atoi(e2) != 0;
    if (const char* e2 = std::getenv("HRC_TUNE_RAY_TILE_W")) { const int v = std::atoi(e2); c->tuneRayTileW = (v == 2 || v == 4 || v == 8 || v == 16) ? v : 0; }
    if (const char* e2 = std::getenv("HRC_TUNE_TAIL_RAYS")) c->tuneTailRays = std::max(0, std::min(1 << 22, std::atoi(e2)));
    if (const char* e2 = std::getenv("HRC_TUNE_SORT_BITS")) c->tuneSortBits = std::max(0, std::min(kSortMaxBits, std::atoi(e2)));
    if (const char* e2 = std::getenv("HRC_TUNE_BLOCKS_PER_SM_PIPE")) c->tunePipelinedBlocksPerSM = std::max(1, std::min(16, std::atoi(e2)));
    if (const char* e2 = std::getenv("HRC_TUNE_PASS_LANES")) c->passLanes = std::max(1, std::min(kMaxLanes, std::atoi(e2)));   // 1: no overlap between passes
    cudaEventCreate(&c->evStart); cudaEventCreate(&c->evStop); cudaEventCreate(&c->evTimer0); cudaEventCreate(&c->evTimer1);
    cudaMalloc(&c->pb.ctrl, sizeof(IterCtrl) * kCtrlEntries);
    cudaMalloc(&c->pb.stats, sizeof(DeviceStats));
    cudaMemsetAsync(c->pb.stats, 0, sizeof(DeviceStats), c->stream);
    cudaMemsetAsync(c->pb.ctrl, 0, sizeof(IterCtrl) * kCtrlEntries, c->stream);
    if ((e = cudaStreamSynchronize(c->stream)) != cudaSuccess) { int r = failCuda(nullptr, e, "context setup"); delete c; return r; }
    std::memset(&c->globals, 0, sizeof(c->globals));
    c->globals.maxRayDepth = 5; c->globals.maxChannelValue = 3.14159265358979f;   // GlobalData defaults, PassGenerator.h:270-272
    c->lights.envTex = -1;
    if (const char* e = std::getenv("HRC_TUNE_REFILL_IDLE")) c->tuneRefillIdle = c->tuneRefillIdleAny = std::max(1, std::min(32, std::atoi(e)));
    if (const char* e = std::getenv("HRC_TUNE_REFILL_IDLE_ANY")) c->tuneRefillIdleAny = std::max(1, std::min(32, std::atoi(e)));      // occlusion traversal only
    if (const char* e = std::getenv("HRC_TUNE_BLOCKS_PER_SM")) c->tuneBlocksPerSM = std::max(1, std::min(16, std::atoi(e)));
    *out = c;
    return HRC_OK;
}

HRC_API int hrcDestroyContext(HrcContext c) {
    if (!c) return HRC_OK;
    cudaSetDevice(c->device);
    joinLanes(c); waitPendingDeferred(c);
    cudaStreamSynchronize(c->stream);
    destroyGraphs(c);
    for (auto ev : c->frameEv) if (ev) cudaEventDestroy(ev);
    if (c->frameRing) cudaFreeHost(c->frameRing);
    freeQueues(c->pb); freeExtraLanes(c);
    std::vector<HrcContext_t::Lane*> lanes;
    for (auto& L : c->extra) lanes.push_back(&L);
    for (auto& L : c->deferred) lanes.push_back(&L);
    for (HrcContext_t::Lane* Lp : lanes) {
        HrcContext_t::Lane& L = *Lp;
        if (!L.s) continue;
        cudaStreamSynchronize(L.s); cudaStreamSynchronize(L.s2);
        for (auto ev : L.evSync) cudaEventDestroy(ev);
        cudaEventDestroy(L.evDone);
        freeDev(L.pb.ctrl); freeDev(L.pb.stats);
        cudaStreamDestroy(L.s2); cudaStreamDestroy(L.s);
    }
    if (c->evMerged) cudaEventDestroy(c->evMerged);
    freeDev(c->accum); freeDev(c->staging); if (c->resolveDev) { cudaFree(c->resolveDev); c->resolveDev = nullptr; } if (c->pinned) cudaFreeHost(c->pinned);
    for (auto& t : c->textures) freeDev(t.dev);
    freeDev(c->dNodes); freeDev(c->dTris); freeDev(c->dShade); freeDev(c->dAttr); freeDev(c->dMaterials); freeDev(c->dTextures);
    freeDev(c->dSeq); freeDev(c->dAperture); freeDev(c->dSeqOffsetX); freeDev(c->dSequenceId); freeDev(c->dBlockCoords); freeDev(c->dBlockRand);
    freeDev(c->pb.ctrl); freeDev(c->pb.stats);
    for (auto ev : c->evPool) cudaEventDestroy(ev);
    for (auto ev : c->evSync) cudaEventDestroy(ev);
    if (c->stream2) { cudaStreamSynchronize(c->stream2); cudaStreamDestroy(c->stream2); }
    cudaEventDestroy(c->evStart); cudaEventDestroy(c->evStop); cudaEventDestroy(c->evTimer0); cudaEventDestroy(c->evTimer1);
    cudaStreamDestroy(c->stream);
    delete c;
    return HRC_OK;
}

HRC_API int hrcSetOption(HrcContext c, int32_t option, int32_t value) {
    if (!c) return fail(nullptr, HRC_ERR_INVALID_VALUE, "null context");
    if (option == HRC_OPT_TRAVERSAL_COUNTERS) c->counters = value != 0;
    else if (option == HRC_OPT_KERNEL_TIMING) c->kernelTiming = value != 0;
    else if (option == HRC_OPT_PIXEL_RAY_COUNTS) {
        if (!c->accum) return fail(c, HRC_ERR_INVALID_OPERATION, "ray counts need a framebuffer size first");
        cudaSetDevice(c->device); joinLanes(c); waitPendingDeferred(c);
        CU(c, cudaStreamSynchronize(c->stream));
        freeDev(c->dRayCounts);
        if (value) { CU(c, cudaMalloc(&c->dRayCounts, (size_t)c->w * c->h * sizeof(uint32_t))); CU(c, cudaMemset(c->dRayCounts, 0, (size_t)c->w * c->h * sizeof(uint32_t))); }
        c->pb.rayCounts = c->dRayCounts;
        for (auto& L : c->extra) L.pb.rayCounts = c->dRayCounts;
        for (auto& L : c->deferred) L.pb.rayCounts = c->dRayCounts;
    }
    else if (option == HRC_OPT_SPECULATIVE_PASSES) {
        if (value < 0 || value > HRC_MAX_DEFERRED_PASSES) return fail(c, HRC_ERR_INVALID_VALUE, "0..3 speculative passes");
        if (c->deferredCount > 0) { int r = hrcDiscardDeferred(c); if (r) return r; }
        c->specDepth = value;
    }
    else if (option == HRC_OPT_SPECULATIVE_STRIDE) {
        if (value < 1 || value > 4096) return fail(c, HRC_ERR_INVALID_VALUE, "speculation stride 1..4096");
        if (c->deferredCount > 0) { int r = hrcDiscardDeferred(c); if (r) return r; }
        c->specStride = value;
    }
    else return fail(c, HRC_ERR_INVALID_VALUE, "unknown option");
    return HRC_OK;
}

HRC_API const char* hrcGetLastError(HrcContext c) { return c ? c->err.c_str() : g_err.c_str(); }

HRC_API int hrcResize(HrcContext c, int32_t w, int32_t h) {
    if (c) c->stateVersion++;
    if (!c) return fail(nullptr, HRC_ERR_INVALID_VALUE, "null context");
    if (w <= 0 || h <= 0) return fail(c, HRC_ERR_INVALID_VALUE, "non-positive framebuffer size");
    if ((uint64_t)w * (uint64_t)h > (1u << 24)) return fail(c, HRC_ERR_UNSUPPORTED, "framebuffers above 2^24 pixels are not supported (pixel index is packed in 24 bits)");
    if (c->mapped) return fail(c, HRC_ERR_INVALID_OPERATION, "resize while pixels are mapped");
    cudaSetDevice(c->device);
    joinLanes(c); waitPendingDeferred(c);
    CU(c, cudaStreamSynchronize(c->stream));
    auto dropFrame = [&]() {      // the context has no framebuffer from here until every allocation below has succeeded
        freeQueues(c->pb); freeExtraLanes(c); freeDev(c->accum); freeDev(c->staging); freeDev(c->dSeqOffsetX); freeDev(c->dSequenceId);
        if (c->pinned) { cudaFreeHost(c->pinned); c->pinned = nullptr; }
        c->pinnedBytes = 0; c->pb.accum = nullptr; c->w = c->h = 0;
        freeDev(c->dRayCounts); c->pb.rayCounts = nullptr;
        for (auto& L : c->extra) L.pb.rayCounts = nullptr;
        for (auto& L : c->deferred) L.pb.rayCounts = nullptr;
        c->pixelTables = false; c->offsetCount = 0;
        freeDev(c->dBlockCoords); freeDev(c->dBlockRand); c->blocksX = c->blocksY = 0;   // block tables depend on the frame size
        destroyGraphs(c);
    };
    dropFrame();
    const size_t n = (size_t)w * h;
    auto allocate = [&]() -> int {
        CU(c, cudaMalloc(&c->accum, n * sizeof(float4)));
        CU(c, cudaMemsetAsync(c->accum, 0, n * sizeof(float4), c->stream));
        CU(c, cudaMallocHost(&c->pinned, n * sizeof(float4))); c->pinnedBytes = n * sizeof(float4);
        std::memset(c->pinned, 0, n * sizeof(float4));
        { int r = allocQueues(c, c->pb, n); if (r) return r; }
        CU(c, cudaStreamSynchronize(c->stream));
        return HRC_OK;
    };
    if (int r = allocate()) { const std::string why = c->err; dropFrame(); c->err = why; return r; }
    c->pb.accum = c->accum;
    c->w = w; c->h = h; c->accumOutsideBandsZero = true; c->devicePixelsExported = false; c->rbBandCount = 0; c->pinnedZeroKey = 0;
    return HRC_OK;
}

HRC_API int hrcClear(HrcContext c) {
    if (c) c->stateVersion++;
    if (!c || !c->accum) return fail(c, HRC_ERR_INVALID_OPERATION, "no framebuffer");
    cudaSetDevice(c->device); joinLanes(c);
    CU(c, cudaMemsetAsync(c->accum, 0, (size_t)c->w * c->h * sizeof(float4), c->stream));
    if (c->dRayCounts) CU(c, cudaMemsetAsync(c->dRayCounts, 0, (size_t)c->w * c->h * sizeof(uint32_t), c->stream));
    c->accumOutsideBandsZero = !c->devicePixelsExported; c->rbBandCount = 0;
    return HRC_OK;
}

HRC_API int hrcReadPixels(HrcContext c, const float** rgba) {
    if (!c || !rgba) return fail(c, HRC_ERR_INVALID_VALUE, "null pointer");
    if (!c->accum) return fail(c, HRC_ERR_INVALID_OPERATION, "no framebuffer");
    if (c->mapped) return fail(c, HRC_ERR_INVALID_OPERATION, "pixels already mapped");
    cudaSetDevice(c->device); joinLanes(c);
    if (c->accumOutsideBandsZero && c->rbBandCount > 1) {
        // Screen-band sharding: this context only ever wrote its own interleaved bands since the last clear, so only those rows
        // cross PCIe; the rest of the pinned frame is zero, as it is on the device.
        const long long key = ((long long)c->rbBandHeight << 40) | ((long long)c->rbBandIndex << 20) | c->rbBandCount;
        if (c->pinnedZeroKey != key && c->pinnedZeroKey != 0) std::memset(c->pinned, 0, (size_t)c->w * c->h * sizeof(float4));
        c->pinnedZeroKey = key;
        const size_t rowBytes = (size_t)c->w * sizeof(float4);
        for (int y0 = c->rbBandIndex * c->rbBandHeight; y0 < c->h; y0 += c->rbBandHeight * c->rbBandCount) {
            const int rows = std::min(c->rbBandHeight, c->h - y0);
            CU(c, cudaMemcpyAsync((char*)c->pinned + (size_t)y0 * rowBytes, (const char*)c->accum + (size_t)y0 * rowBytes, (size_t)rows * rowBytes, cudaMemcpyDeviceToHost, c->stream));
        }
    } else {
        CU(c, cudaMemcpyAsync(c->pinned, c->accum, (size_t)c->w * c->h * sizeof(float4), cudaMemcpyDeviceToHost, c->stream));
        c->pinnedZeroKey = -1;
    }
    CU(c, cudaStreamSynchronize(c->stream));
    c->mapped = true; *rgba = c->pinned;
    return HRC_OK;
}
HRC_API int hrcReleasePixels(HrcContext c) {
    if (!c) return fail(nullptr, HRC_ERR_INVALID_VALUE, "null context");
    if (!c->mapped) return fail(c, HRC_ERR_INVALID_OPERATION, "pixels are not mapped");
    c->mapped = false; return HRC_OK;
}
HRC_API int hrcAddPixels(HrcContext c, const float* rgba) {
    if (c) c->stateVersion++;
    if (!c || !rgba) return fail(c, HRC_ERR_INVALID_VALUE, "null pointer");
    if (!c->accum) return fail(c, HRC_ERR_INVALID_OPERATION, "no framebuffer");
    cudaSetDevice(c->device); joinLanes(c);
    const size_t n = (size_t)c->w * c->h;
    if (!c->staging) CU(c, cudaMalloc(&c->staging, n * sizeof(float4)));
    CU(c, cudaMemcpyAsync(c->staging, rgba, n * sizeof(float4), cudaMemcpyHostToDevice, c->stream));
    launchAddPixels(c->accum, c->staging, n, c->stream);
    c->accumOutsideBandsZero = false; c->rbBandCount = 0;
    CU(c, cudaStreamSynchronize(c->stream));
    return HRC_OK;
}
HRC_API int hrcGetDevicePixels(HrcContext c, void** devPtr) {
    if (c) c->stateVersion++;
    if (!c || !devPtr) return fail(c, HRC_ERR_INVALID_VALUE, "null pointer");
    if (!c->accum) return fail(c, HRC_ERR_INVALID_OPERATION, "no framebuffer");
    cudaSetDevice(c->device); joinLanes(c);
    CU(c, cudaStreamSynchronize(c->stream));
    c->accumOutsideBandsZero = false; c->rbBandCount = 0; c->devicePixelsExported = true;      // the caller may write through the pointer (in-place NCCL reduce), now and after any later hrcClear
    *devPtr = c->accum; return HRC_OK;
}

HRC_API int hrcResolve(HrcContext c, const HrcPostProcess* post, int32_t format, void* hostOut) {
    if (!c || !hostOut) return fail(c, HRC_ERR_INVALID_VALUE, "null pointer");
    if (!c->accum) return fail(c, HRC_ERR_INVALID_OPERATION, "no framebuffer");
    if (format < HRC_RESOLVE_DISPLAY_RGBA32F || format > HRC_RESOLVE_HDR_RGB32F) return fail(c, HRC_ERR_INVALID_VALUE, "unknown resolve format");
    if (!post && format != HRC_RESOLVE_HDR_RGB32F) return fail(c, HRC_ERR_INVALID_VALUE, "post-processing parameters are required for the display formats");
    cudaSetDevice(c->device); joinLanes(c);
    const size_t n = (size_t)c->w * c->h;
    const size_t bytes = n * (format == HRC_RESOLVE_DISPLAY_RGBA32F ? 16 : format == HRC_RESOLVE_DISPLAY_RGB8 ? 3 : 12);
    const size_t padded = (bytes + 15) & ~(size_t)15;
    if (c->resolveBytes < padded) {
        CU(c, cudaStreamSynchronize(c->stream));
        if (c->resolveDev) { cudaFree(c->resolveDev); c->resolveDev = nullptr; } c->resolveBytes = 0;
        CU(c, cudaMalloc(&c->resolveDev, padded)); c->resolveBytes = padded;
    }
    HrcPostProcess pp{}; if (post) pp = *post;
    launchResolve(c->accum, pp, format, c->w, c->h, c->resolveDev, c->stream);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return failCuda(c, e, "resolve launch");
    CU(c, cudaMemcpyAsync(hostOut, c->resolveDev, bytes, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    return HRC_OK;
}

HRC_API int hrcCreateTexture2D(HrcContext c, const HrcTextureDesc* d, const void* pixels, HrcTexture* out) {
    if (c) c->stateVersion++;
    if (!c || !d || !pixels || !out || d->width <= 0 || d->height <= 0) return fail(c, HRC_ERR_INVALID_VALUE, "bad texture descriptor");
    cudaSetDevice(c->device); joinLanes(c);
    const size_t n = (size_t)d->width * d->height;
    std::vector<float4> host(n);
    const float* f = (const float*)pixels; const uint8_t* b = (const uint8_t*)pixels;
    for (size_t i = 0; i < n; ++i) {
        switch (d->format) {
        case HRC_TEX_R32F: host[i] = make_float4(f[i], 0.f, 0.f, 1.f); break;
        case HRC_TEX_RGB32F: host[i] = make_float4(f[3 * i], f[3 * i + 1], f[3 * i + 2], 1.f); break;
        case HRC_TEX_RGBA32F: host[i] = make_float4(f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]); break;
        case HRC_TEX_RGBA8: host[i] = make_float4((float)b[4 * i] / 255.0f, (float)b[4 * i + 1] / 255.0f, (float)b[4 * i + 2] / 255.0f, (float)b[4 * i + 3] / 255.0f); break;
        default: return fail(c, HRC_ERR_INVALID_VALUE, "unknown texture format");
        }
    }
    TextureRec t; t.w = d->width; t.h = d->height; t.wrapS = d->wrapS; t.wrapT = d->wrapT; t.alive = true;
    CU(c, cudaMalloc(&t.dev, n * sizeof(float4)));
    CU(c, cudaMemcpyAsync(t.dev, host.data(), n * sizeof(float4), cudaMemcpyHostToDevice, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    c->textures.push_back(t); c->texturesDirty = true;
    *out = (HrcTexture)c->textures.size() - 1;
    return HRC_OK;
}
HRC_API int hrcDestroyTexture(HrcContext c, HrcTexture tex) {
    if (c) c->stateVersion++;
    if (!c || tex < 0 || tex >= (int)c->textures.size() || !c->textures[tex].alive) return fail(c, HRC_ERR_INVALID_VALUE, "bad texture handle");
    cudaSetDevice(c->device); joinLanes(c); waitPendingDeferred(c);
    CU(c, cudaStreamSynchronize(c->stream));
    freeDev(c->textures[tex].dev); c->textures[tex].alive = false; c->texturesDirty = true;
    return HRC_OK;
}

HRC_API int hrcCreateMaterial(HrcContext c, int32_t type, uint32_t flags, const void* params, size_t bytes, HrcMaterial* out) {
    if (c) c->stateVersion++;
    if (!c || !params || !out) return fail(c, HRC_ERR_INVALID_VALUE, "null pointer");
    if (flags >> 10) return fail(c, HRC_ERR_INVALID_VALUE, "unknown material feature flags");
    MaterialRec m; m.flags = flags;
    int r = packMaterial(c, m, type, params, bytes); if (r) return r;
    c->materials.push_back(m); c->materialsDirty = true;
    *out = (HrcMaterial)c->materials.size() - 1;
    return HRC_OK;
}
HRC_API int hrcUpdateMaterial(HrcContext c, HrcMaterial mat, const void* params, size_t bytes) {
    if (c) c->stateVersion++;
    if (!c || mat < 0 || mat >= (int)c->materials.size() || !params) return fail(c, HRC_ERR_INVALID_VALUE, "bad material handle");
    int r = packMaterial(c, c->materials[mat], c->materials[mat].type, params, bytes); if (r) return r;
    c->materialsDirty = true; return HRC_OK;
}

HRC_API int hrcCreateMesh(HrcContext c, const HrcMeshDesc* d, HrcMesh* out) {
    if (c) c->stateVersion++;
    if (!c || !d || !out || !d->positions || !d->normals || !d->indices) return fail(c, HRC_ERR_INVALID_VALUE, "mesh needs positions, normals and indices");
    if (d->material < 0 || d->material >= (int)c->materials.size()) return fail(c, HRC_ERR_INVALID_VALUE, "bad material handle");
    if (d->drawMode != HRC_DRAW_TRIANGLES && d->drawMode != HRC_DRAW_TRIANGLE_STRIP) return fail(c, HRC_ERR_INVALID_VALUE, "bad draw mode");
    MeshRec m; m.alive = true;
    m.pos.resize((size_t)3 * d->vertexCount); m.nrm.resize((size_t)3 * d->vertexCount);
    const int ps = d->positionStride ? d->positionStride : 12, ns = d->normalStride ? d->normalStride : 12;
    for (uint32_t i = 0; i < d->vertexCount; ++i) {
        std::memcpy(&m.pos[3 * (size_t)i], (const char*)d->positions + (size_t)i * ps, 12);
        std::memcpy(&m.nrm[3 * (size_t)i], (const char*)d->normals + (size_t)i * ns, 12);
    }
    auto copyAttr = [&](const float* src, int stride, int comps, std::vector<float>& dst) {
        if (!src) return;
        dst.resize((size_t)comps * d->vertexCount);
        const int st = stride ? stride : comps * 4;
        for (uint32_t i = 0; i < d->vertexCount; ++i) std::memcpy(&dst[(size_t)comps * i], (const char*)src + (size_t)i * st, (size_t)comps * 4);
    };
    copyAttr(d->texCoords, d->texCoordStride, 2, m.uv); copyAttr(d->tangents, d->tangentStride, 3, m.tan);
    copyAttr(d->bitangents, d->bitangentStride, 3, m.bit); copyAttr(d->colors, d->colorStride, 3, m.col);
    m.idx.assign(d->indices, d->indices + d->elementCount);
    for (uint32_t i : m.idx) if (i >= d->vertexCount) return fail(c, HRC_ERR_INVALID_VALUE, "index out of range");
    m.drawMode = d->drawMode; std::memcpy(m.M, d->worldFromEntity, sizeof(m.M));
    m.frontFaceCW = d->frontFaceCW; m.isOccluder = d->isOccluder; m.material = d->material;
    c->meshes.push_back(std::move(m)); c->committed = false;
    *out = (HrcMesh)c->meshes.size() - 1;
    return HRC_OK;
}
HRC_API int hrcDestroyMesh(HrcContext c, HrcMesh mesh) {
    if (c) c->stateVersion++;
    if (!c || mesh < 0 || mesh >= (int)c->meshes.size() || !c->meshes[mesh].alive) return fail(c, HRC_ERR_INVALID_VALUE, "bad mesh handle");
    c->meshes[mesh] = MeshRec(); c->committed = false; return HRC_OK;
}
HRC_API int hrcSetMeshTransform(HrcContext c, HrcMesh mesh, const float* worldFromEntity) {
    if (c) c->stateVersion++;
    if (!c || mesh < 0 || mesh >= (int)c->meshes.size() || !c->meshes[mesh].alive || !worldFromEntity) return fail(c, HRC_ERR_INVALID_VALUE, "bad mesh handle");
    std::memcpy(c->meshes[mesh].M, worldFromEntity, sizeof(c->meshes[mesh].M)); c->committed = false; return HRC_OK;
}
HRC_API int hrcClearScene(HrcContext c) {
    if (c) c->stateVersion++;
    if (!c) return fail(nullptr, HRC_ERR_INVALID_VALUE, "null context");
    c->meshes.clear(); c->materials.clear(); c->committed = false; c->materialsDirty = true;
    return HRC_OK;
}

HRC_API int hrcCommitScene(HrcContext c) {
    if (c) c->stateVersion++;
    if (!c) return fail(nullptr, HRC_ERR_INVALID_VALUE, "null context");
    cudaSetDevice(c->device); joinLanes(c); waitPendingDeferred(c);
    CU(c, cudaStreamSynchronize(c->stream));
    // 1. flatten to world space; primitive id = running triangle index in (mesh, element) order (policy A-9)
    std::vector<BuildTriangle> bt;
    std::vector<float4> shade;      // 4 per triangle
    std::vector<int32_t> primType;  // material type per triangle: selects the shade kernel (ray class) at hit time
    size_t triTotal = 0;
    for (const MeshRec& m : c->meshes) if (m.alive) triTotal += m.drawMode == HRC_DRAW_TRIANGLES ? m.idx.size() / 3 : (m.idx.size() >= 3 ? m.idx.size() - 2 : 0);
    if (triTotal >= (1u << 28)) return fail(c, HRC_ERR_UNSUPPORTED, "more than 2^28 triangles");
    bt.reserve(triTotal); shade.reserve(triTotal * 4);
    std::vector<float4> attr;       // kAttrStride per triangle, only when some mesh carries uv / tangent / bitangent / colour
    bool anyAttr = false;
    for (const MeshRec& m : c->meshes) if (m.alive && (!m.uv.empty() || !m.tan.empty() || !m.bit.empty() || !m.col.empty())) anyAttr = true;
    if (anyAttr) attr.reserve(triTotal * kAttrStride);
    c->anyMasked = false;
    for (const MeshRec& m : c->meshes) {
        if (!m.alive) continue;
        const size_t nv = m.pos.size() / 3;
        std::vector<float> wp(3 * nv), wn(3 * nv), wt(m.tan.size()), wb(m.bit.size());
        for (size_t i = 0; i < nv; ++i) { xformPoint(m.M, &m.pos[3 * i], &wp[3 * i]); xformNormal(m.M, &m.nrm[3 * i], &wn[3 * i]); }
        for (size_t i = 0; i < m.tan.size() / 3; ++i) xformNormal(m.M, &m.tan[3 * i], &wt[3 * i]);     // vertex.rlsl:36-37
        for (size_t i = 0; i < m.bit.size() / 3; ++i) xformNormal(m.M, &m.bit[3 * i], &wb[3 * i]);
        if (!m.isOccluder) c->anyMasked = true;
        auto emit = [&](uint32_t a, uint32_t b, uint32_t cc) {
            BuildTriangle t;
            for (int k = 0; k < 3; ++k) { t.v0[k] = wp[3 * a + k]; t.v1[k] = wp[3 * b + k]; t.v2[k] = wp[3 * cc + k]; }
            bt.push_back(t);
            primType.push_back(c->materials[m.material].type | (m.isOccluder ? 0 : (int32_t)kTriNonOccluder));
            if (anyAttr) {
                const uint32_t vi[3] = {a, b, cc};
                float4 rec[kAttrStride];
                for (auto& r : rec) r = make_float4(0.f, 0.f, 0.f, 0.f);
                if (!m.uv.empty()) { rec[0] = make_float4(m.uv[2 * a], m.uv[2 * a + 1], m.uv[2 * b], m.uv[2 * b + 1]); rec[1] = make_float4(m.uv[2 * cc], m.uv[2 * cc + 1], 0.f, 0.f); }
                for (int k = 0; k < 3; ++k) {
                    if (!wt.empty()) rec[2 + k] = make_float4(wt[3 * vi[k]], wt[3 * vi[k] + 1], wt[3 * vi[k] + 2], 0.f);
                    if (!wb.empty()) rec[5 + k] = make_float4(wb[3 * vi[k]], wb[3 * vi[k] + 1], wb[3 * vi[k] + 2], 0.f);
                    if (!m.col.empty()) rec[8 + k] = make_float4(m.col[3 * vi[k]], m.col[3 * vi[k] + 1], m.col[3 * vi[k] + 2], 0.f);
                }
                attr.insert(attr.end(), rec, rec + kAttrStride);
            }
            // geometric normal = normalize(cross(e1, e2)) with the same fp32 operation order as the shaders' restatement
            float e1[3], e2[3];
            for (int k = 0; k < 3; ++k) { e1[k] = t.v1[k] - t.v0[k]; e2[k] = t.v2[k] - t.v0[k]; }
            float nx = e1[1] * e2[2] - e1[2] * e2[1], ny = e1[2] * e2[0] - e1[0] * e2[2], nz = e1[0] * e2[1] - e1[1] * e2[0];
            float inv = 1.0f / std::sqrt((nx * nx + ny * ny) + nz * nz);
            shade.push_back(make_float4(wn[3 * a], wn[3 * a + 1], wn[3 * a + 2], __builtin_bit_cast(float, (int32_t)m.material)));
            shade.push_back(make_float4(wn[3 * b], wn[3 * b + 1], wn[3 * b + 2], __builtin_bit_cast(float, (int32_t)m.frontFaceCW)));
            shade.push_back(make_float4(wn[3 * cc], wn[3 * cc + 1], wn[3 * cc + 2], 0.0f));
            shade.push_back(make_float4(nx * inv, ny * inv, nz * inv, 0.0f));
        };
        if (m.drawMode == HRC_DRAW_TRIANGLES) {
            for (size_t k = 0; k + 2 < m.idx.size(); k += 3) emit(m.idx[k], m.idx[k + 1], m.idx[k + 2]);
        } else {
            for (size_t k = 0; k + 2 < m.idx.size(); ++k) {
                if (k & 1) emit(m.idx[k + 1], m.idx[k], m.idx[k + 2]);
                else       emit(m.idx[k], m.idx[k + 1], m.idx[k + 2]);
            }
        }
    }
    // 2. build
    WideBVH bvh;
    buildWideBVH(bt.data(), bt.size(), 0, bvh);
    if (bvh.maxDepth >= (uint32_t)kTraceStack) return fail(c, HRC_ERR_UNSUPPORTED, "BVH deeper than the traversal stack");
    // 3. leaf-ordered triangle records (v0 | primId, e1, e2)
    std::vector<float4> tris(std::max<size_t>(1, bvh.triOrder.size()) * 3);
    for (size_t s = 0; s < bvh.triOrder.size(); ++s) {
        const uint32_t prim = bvh.triOrder[s];
        const BuildTriangle& t = bt[prim];
        tris[3 * s + 0] = make_float4(t.v0[0], t.v0[1], t.v0[2], __builtin_bit_cast(float, (int32_t)prim));
        tris[3 * s + 1] = make_float4(t.v1[0] - t.v0[0], t.v1[1] - t.v0[1], t.v1[2] - t.v0[2], __builtin_bit_cast(float, primType[prim]));
        tris[3 * s + 2] = make_float4(t.v2[0] - t.v0[0], t.v2[1] - t.v0[1], t.v2[2] - t.v0[2], 0.0f);
    }
    if (shade.empty()) shade.resize(4, make_float4(0, 0, 0, 0));
    // 4. upload
    freeDev(c->dNodes); freeDev(c->dTris); freeDev(c->dShade); freeDev(c->dAttr);
    if (!attr.empty()) {
        CU(c, cudaMalloc(&c->dAttr, attr.size() * sizeof(float4)));
        CU(c, cudaMemcpyAsync(c->dAttr, attr.data(), attr.size() * sizeof(float4), cudaMemcpyHostToDevice, c->stream));
    }
    CU(c, cudaMalloc(&c->dNodes, bvh.nodes.size() * sizeof(WideNode)));
    CU(c, cudaMalloc(&c->dTris, tris.size() * sizeof(float4)));
    CU(c, cudaMalloc(&c->dShade, shade.size() * sizeof(float4)));
    CU(c, cudaMemcpyAsync(c->dNodes, bvh.nodes.data(), bvh.nodes.size() * sizeof(WideNode), cudaMemcpyHostToDevice, c->stream));
    CU(c, cudaMemcpyAsync(c->dTris, tris.data(), tris.size() * sizeof(float4), cudaMemcpyHostToDevice, c->stream));
    CU(c, cudaMemcpyAsync(c->dShade, shade.data(), shade.size() * sizeof(float4), cudaMemcpyHostToDevice, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    for (int k = 0; k < 3; ++k) { c->sceneMin[k] = bvh.sceneMin[k]; c->sceneMax[k] = bvh.sceneMax[k]; }
    c->info.triangleCount = bt.size(); c->info.nodeCount = bvh.nodes.size();
    c->info.nodeBytes = bvh.nodes.size() * sizeof(WideNode); c->info.triangleBytes = bt.size() * 48; c->info.shadingBytes = bt.size() * 64 + attr.size() * sizeof(float4);
    c->info.buildSeconds = (float)bvh.buildSeconds;
    c->committed = true;
    return HRC_OK;
}

HRC_API int hrcSetEnvironment(HrcContext c, const HrcEnvLight* env) {
    if (c) c->stateVersion++;
    if (!c || !env) return fail(c, HRC_ERR_INVALID_VALUE, "null pointer");
    if (env->enabled && env->texture >= 0 && (env->texture >= (int)c->textures.size() || !c->textures[env->texture].alive))
        return fail(c, HRC_ERR_INVALID_VALUE, "bad environment texture handle");
    c->lights.envEnabled = env->enabled ? 1 : 0; c->lights.envTex = env->texture; c->lights.envExposure = env->exposureCompensation; c->lights.envTheta = env->thetaRotation;
    return HRC_OK;
}
HRC_API int hrcSetDirectionalLights(HrcContext c, int32_t n, const HrcDirLight* l) {
    if (c) c->stateVersion++;
    if (!c || n < 0 || n > HRC_MAX_LIGHTS_PER_TYPE || (n && !l)) return fail(c, HRC_ERR_INVALID_VALUE, "0..5 directional lights");
    c->lights.nDir = n;
    for (int i = 0; i < n; ++i) for (int k = 0; k < 3; ++k) { c->lights.dirDirection[i][k] = l[i].direction[k]; c->lights.dirColor[i][k] = l[i].color[k]; }
    return HRC_OK;
}
HRC_API int hrcSetPointLights(HrcContext c, int32_t n, const HrcPointLight* l) {
    if (c) c->stateVersion++;
    if (!c || n < 0 || n > HRC_MAX_LIGHTS_PER_TYPE || (n && !l)) return fail(c, HRC_ERR_INVALID_VALUE, "0..5 point lights");
    c->lights.nPoint = n;
    for (int i = 0; i < n; ++i) for (int k = 0; k < 3; ++k) { c->lights.pointPosition[i][k] = l[i].position[k]; c->lights.pointColor[i][k] = l[i].color[k]; }
    return HRC_OK;
}
HRC_API int hrcSetSpotLights(HrcContext c, int32_t n, const HrcSpotLight* l) {
    if (c) c->stateVersion++;
    if (!c || n < 0 || n > HRC_MAX_LIGHTS_PER_TYPE || (n && !l)) return fail(c, HRC_ERR_INVALID_VALUE, "0..5 spot lights");
    c->lights.nSpot = n;
    for (int i = 0; i < n; ++i) {
        for (int k = 0; k < 3; ++k) { c->lights.spotPosition[i][k] = l[i].position[k]; c->lights.spotDirection[i][k] = l[i].direction[k]; c->lights.spotColor[i][k] = l[i].color[k]; }
        c->lights.spotAngles[i][0] = l[i].angles[0]; c->lights.spotAngles[i][1] = l[i].angles[1];
    }
    return HRC_OK;
}

HRC_API int hrcSetSequences(HrcContext c, int32_t ns, int32_t len, const float* xy, const float* ap) {
    if (c) c->stateVersion++;
    if (!c || ns <= 0 || len <= 0 || !xy || !ap) return fail(c, HRC_ERR_INVALID_VALUE, "bad sequence tables");
    cudaSetDevice(c->device); joinLanes(c); waitPendingDeferred(c);
    CU(c, cudaStreamSynchronize(c->stream));
    freeDev(c->dSeq); freeDev(c->dAperture);
    const size_t bytes = (size_t)ns * len * sizeof(float2);
    CU(c, cudaMalloc(&c->dSeq, bytes)); CU(c, cudaMalloc(&c->dAperture, bytes));
    CU(c, cudaMemcpy(c->dSeq, xy, bytes, cudaMemcpyHostToDevice));
    CU(c, cudaMemcpy(c->dAperture, ap, bytes, cudaMemcpyHostToDevice));
    c->numSequences = ns; c->sequenceLength = len;
    return HRC_OK;
}
HRC_API int hrcSetPixelTables(HrcContext c, const float* off, size_t count, const uint8_t* sid) {
    if (c) c->stateVersion++;
    if (!c || !off || !sid || count == 0) return fail(c, HRC_ERR_INVALID_VALUE, "bad pixel tables");
    if (c->w <= 0) return fail(c, HRC_ERR_INVALID_OPERATION, "pixel tables need a framebuffer size first");
    cudaSetDevice(c->device); joinLanes(c); waitPendingDeferred(c);
    CU(c, cudaStreamSynchronize(c->stream));
    freeDev(c->dSeqOffsetX); freeDev(c->dSequenceId);
    CU(c, cudaMalloc(&c->dSeqOffsetX, count * sizeof(float)));
    CU(c, cudaMalloc(&c->dSequenceId, (size_t)c->w * c->h));
    CU(c, cudaMemcpy(c->dSeqOffsetX, off, count * sizeof(float), cudaMemcpyHostToDevice));
    CU(c, cudaMemcpy(c->dSequenceId, sid, (size_t)c->w * c->h, cudaMemcpyHostToDevice));
    c->offsetCount = count; c->pixelTables = true;
    return HRC_OK;
}
HRC_API int hrcSetInteractiveBlocks(HrcContext c, int32_t bw, int32_t bh, const float* coords, int32_t bx, int32_t by, const float* rnd) {
    if (c) c->stateVersion++;
    if (!c || !coords || !rnd || bw <= 0 || bh <= 0 || bx <= 0 || by <= 0) return fail(c, HRC_ERR_INVALID_VALUE, "bad interactive block tables");
    if (c->w <= 0) return fail(c, HRC_ERR_INVALID_OPERATION, "interactive block tables need a framebuffer size first");
    if ((int64_t)bx * bw < c->w || (int64_t)by * bh < c->h) return fail(c, HRC_ERR_INVALID_VALUE, "block tables do not cover the frame");
    cudaSetDevice(c->device); joinLanes(c); waitPendingDeferred(c);
    CU(c, cudaStreamSynchronize(c->stream));
    freeDev(c->dBlockCoords); freeDev(c->dBlockRand);
    CU(c, cudaMalloc(&c->dBlockCoords, (size_t)bw * bh * 3 * sizeof(float)));
    CU(c, cudaMalloc(&c->dBlockRand, (size_t)bx * by * sizeof(float2)));
    CU(c, cudaMemcpy(c->dBlockCoords, coords, (size_t)bw * bh * 3 * sizeof(float), cudaMemcpyHostToDevice));
    CU(c, cudaMemcpy(c->dBlockRand, rnd, (size_t)bx * by * sizeof(float2), cudaMemcpyHostToDevice));
    c->blockW = bw; c->blockH = bh; c->blocksX = bx; c->blocksY = by;
    return HRC_OK;
}
HRC_API int hrcSetGlobals(HrcContext c, const HrcGlobals* g) {
    if (c) c->stateVersion++;
    if (!c || !g) return fail(c, HRC_ERR_INVALID_VALUE, "null pointer");
    c->globals = *g; return HRC_OK;
}

// The kernels of one pass, in launch order, on the lane's two streams. Called either directly or under stream capture.
static void issuePass(const DScene& sc, int slot, uint32_t total, const PassBuffers& pb, const LaunchConfig& lc, int iterations, bool overlap,
                      cudaStream_t s, cudaStream_t s2, const std::vector<cudaEvent_t>& evSync) {
    launchRaygen(sc, slot, total, pb, s);
    const bool tail = lc.tailRays > 0;
    if (tail && total <= (uint32_t)lc.tailRays) {         // a frame this small is all tail: the primary rays go straight to the path kernel
        launchTailDecide(pb, 0, lc, s);
        launchPathTail(sc, slot, pb, 0, iterations, lc, s);
    }
    if (!overlap || iterations < 2) {
        for (int it = 0; it < iterations; ++it) {
            launchTraceClosest(sc, slot, pb, it, lc, s);
            launchShade(sc, slot, pb, it, lc, s);
            if (tail && it + 1 < iterations) launchTailDecide(pb, it + 1, lc, s);
            if (it + 1 < iterations) launchSortRays(pb, it + 1, lc, s);
            launchTraceAny(sc, slot, pb, it, lc, s);
            launchShadeLights(sc, slot, pb, it, lc, s);
            if (tail && it + 1 < iterations) launchPathTail(sc, slot, pb, it + 1, iterations, lc, s);
        }
    } else {
        // Occlusion rays of iteration i and extension rays of iteration i+1 are independent (both were emitted by shade(i)):
        // k_trace<any>(i) runs on the second stream next to k_trace<closest>(i+1) and fills the SMs its tail leaves idle.
        // shade(i+1) waits for both: it reuses the shadow queue and accumulates into the same pixels as the light shaders.
        for (int it = 0; it < iterations; ++it) {
            launchTraceClosest(sc, slot, pb, it, lc, s);
            if (it > 0) cudaStreamWaitEvent(s, evSync[2 * (it - 1) + 1], 0);
            launchShade(sc, slot, pb, it, lc, s);
            if (tail && it + 1 < iterations) launchTailDecide(pb, it + 1, lc, s);
            cudaEventRecord(evSync[2 * it], s);
            if (it + 1 < iterations) launchSortRays(pb, it + 1, lc, s);
            cudaStreamWaitEvent(s2, evSync[2 * it], 0);
            launchTraceAny(sc, slot, pb, it, lc, s2);
            launchShadeLights(sc, slot, pb, it, lc, s2);
            // the path kernel accumulates into the pixels any(it) writes: behind it on the second stream, and shade(it + 1) (a no-op once the
            // tail is taken) waits for both
            if (tail && it + 1 < iterations) launchPathTail(sc, slot, pb, it + 1, iterations, lc, s2);
            cudaEventRecord(evSync[2 * it + 1], s2);
        }
        cudaStreamWaitEvent(s, evSync[2 * (iterations - 1) + 1], 0);
    }
    launchFinishPass(pb, iterations, true, s);          // also re-arms the control words: a lane's IterCtrl is all zero whenever no pass is in flight on it
}

static int enqueuePass(HrcContext c, const HrcFrameParams* fp, bool timeKernels, float* kernelMs /*[5]: raygen, closest, shade, any, closest of iteration 0*/, int* launches /*[3]*/, int lane = 0, bool pipelined = false) {
    if (!c || !fp) return fail(c, HRC_ERR_INVALID_VALUE, "null pointer");
    if (!c->committed) return fail(c, HRC_ERR_INVALID_OPERATION, "render pass without a committed scene");
    if (!c->accum) return fail(c, HRC_ERR_INVALID_OPERATION, "render pass without a framebuffer");
    if (!c->dSeq || !c->pixelTables) return fail(c, HRC_ERR_INVALID_OPERATION, "render pass without QMC / pixel tables");
    if (fp->interactiveMode) {
        if (!c->dBlockCoords) return fail(c, HRC_ERR_INVALID_OPERATION, "interactive mode without block tables (hrcSetInteractiveBlocks)");
        if (fp->blockSize[0] != c->blockW || fp->blockSize[1] != c->blockH) return fail(c, HRC_ERR_INVALID_VALUE, "blockSize differs from the uploaded block tables");
    }
    if (c->mapped) return fail(c, HRC_ERR_INVALID_OPERATION, "render pass while pixels are mapped");
    if (fp->sampleIndex < 0) return fail(c, HRC_ERR_INVALID_VALUE, "negative sampleIndex");
    const size_t maxIdx = (size_t)(int)(((float)(c->h - 1) + 0.5f) * (float)c->h + ((float)(c->w - 1) + 0.5f));
    if (maxIdx >= c->offsetCount) return fail(c, HRC_ERR_INVALID_VALUE, "sequence-offset table too short for the reference index formula");
    cudaSetDevice(c->device);
    if (c->materialsDirty || c->texturesDirty) { joinLanes(c); waitPendingDeferred(c); }   // the uploads rewrite tables that passes in flight on other lanes read
    int r = uploadMaterials(c); if (r) return r;
    r = uploadTextures(c); if (r) return r;
    if (c->globals.maxRayDepth < 0 || c->globals.maxRayDepth + 1 + (c->anySingleSided ? 16 : 0) > kMaxIterations)
        return fail(c, HRC_ERR_INVALID_VALUE, "maxRayDepth out of range (a pass runs at most 64 wavefront iterations: maxRayDepth + 1, + 16 pass-through segments with single-sided materials)");

    DFrame fr; std::memset(&fr, 0, sizeof(fr));
    fr.g = c->globals; fr.g.sampleIndex = fp->sampleIndex;
    fr.fovTan = fp->fovTan; fr.aspectRatio = fp->aspectRatio; fr.focusDistance = fp->focusDistance; fr.apertureRadius = fp->apertureRadius;
    std::memcpy(fr.view, fp->viewMatrix, sizeof(fr.view));
    fr.maxSampleIndex = fp->maxSampleIndex; fr.w = c->w; fr.h = c->h;
    fr.bandCount = fp->bandCount > 1 ? fp->bandCount : 1; fr.bandHeight = fp->bandHeight > 0 ? fp->bandHeight : 1; fr.bandIndex = fp->bandIndex;
    if (fr.bandCount > 1) {
        if (fr.bandIndex < 0 || fr.bandIndex >= fr.bandCount) return fail(c, HRC_ERR_INVALID_VALUE, "band index out of range");
        int rows = 0; for (int y = 0; y < c->h; ++y) if ((y / fr.bandHeight) % fr.bandCount == fr.bandIndex) rows++;
        fr.localRows = rows;
    } else fr.localRows = c->h;
    if (fr.bandCount > 1 && c->accumOutsideBandsZero && (c->rbBandCount == 0 || (c->rbBandCount == fr.bandCount && c->rbBandIndex == fr.bandIndex && c->rbBandHeight == fr.bandHeight))) {
        c->rbBandCount = fr.bandCount; c->rbBandIndex = fr.bandIndex; c->rbBandHeight = fr.bandHeight;
    } else { c->accumOutsideBandsZero = false; c->rbBandCount = 0; }
    fr.numSequences = c->numSequences; fr.sequenceLength = c->sequenceLength; fr.offsetCount = (uint32_t)c->offsetCount;
    fr.interactiveMode = fp->interactiveMode ? 1 : 0; fr.blockW = c->blockW; fr.blockH = c->blockH;
    fr.blockSampleX = fp->currentBlockPixelSample[0]; fr.blockSampleY = fp->currentBlockPixelSample[1]; fr.blocksX = c->blocksX;
    fr.lights = c->lights;
    // tiles need whole tiles: frame width a multiple of the tile width, the context's rows a multiple of the tile height; interactive passes compact their rays anyway
    const int tileW = c->tuneRayTileW, tileH = tileW > 0 ? 32 / tileW : 0;
    if (tileW > 0 && !fr.interactiveMode && fr.w % tileW == 0 && fr.localRows % tileH == 0) { fr.rayTileW = tileW; fr.rayTileH = tileH; }
    if (fr.lights.envEnabled && fr.lights.envTex >= 0 && !c->textures[fr.lights.envTex].alive) return fail(c, HRC_ERR_INVALID_OPERATION, "environment texture was destroyed");

    const int iterations = fr.g.maxRayDepth + 1 + (c->anySingleSided ? 16 : 0);
    const DScene sc = sceneOf(c);
    LaunchConfig lc = configOf(c);
    if (pipelined) lc.traceBlocksPerSM = c->tunePipelinedBlocksPerSM;
    HrcContext_t::Lane* L = lane > 0 ? &c->extra[lane - 1] : lane < 0 ? &c->deferred[-lane - 1] : nullptr;      // lane < 0: deferred lane -lane-1
    const int slot = lane > 0 ? lane : lane < 0 ? kMaxLanes + (-lane - 1) : 0;
    PassBuffers pb = L ? L->pb : c->pb;
    // Deferred light shading pays where the light shader is long (the environment lookup: atan2, texture fetch) and loses a little on scenes with punctual
    // lights only, whose light shaders are a few multiplies (config #1: -4 %, profiles/r02_notes.md): there they stay inside k_trace<any>.
    if (!fr.lights.envEnabled) pb.litIdx = nullptr;
    const std::vector<cudaEvent_t>& evSync = L ? L->evSync : c->evSync;
    cudaStream_t s = L ? L->s : c->stream, s2 = L ? L->s2 : c->stream2;
    const uint32_t total = (uint32_t)fr.w * (uint32_t)fr.localRows;

    // per-pass constants: pinned ring entry -> the lane's constant-memory slot, in stream order
    const int k = c->frameNext; c->frameNext = (k + 1) % kFrameRing;
    if (c->frameUsed[k]) CU(c, cudaEventSynchronize(c->frameEv[k]));
    c->frameRing[k] = fr; c->frameUsed[k] = true;
    uploadFrame(&c->frameRing[k], slot, s);
    CU(c, cudaEventRecord(c->frameEv[k], s));

    if (launches) { launches[0] = iterations; launches[1] = iterations * ((int)lc.shadeMiss + (int)lc.shadePbr + (int)lc.shadeGlass + (pb.litIdx ? 1 : 0)); launches[2] = iterations; }
    if (timeKernels) {
        // instrumented pass: CUDA events between the kernels, nothing overlapped
        const size_t evNeeded = (size_t)(3 * iterations + 2);
        while (c->evPool.size() < evNeeded) { cudaEvent_t ev; CU(c, cudaEventCreate(&ev)); c->evPool.push_back(ev); }
        size_t ev = 0;
        auto mark = [&]() { cudaEventRecord(c->evPool[ev++], s); };
        mark();
        launchRaygen(sc, slot, total, pb, s);
        mark();
        for (int it = 0; it < iterations; ++it) {
            launchTraceClosest(sc, slot, pb, it, lc, s); mark();
            launchShade(sc, slot, pb, it, lc, s);    if (it + 1 < iterations) launchSortRays(pb, it + 1, lc, s); mark();
            launchTraceAny(sc, slot, pb, it, lc, s); launchShadeLights(sc, slot, pb, it, lc, s); mark();
        }
        launchFinishPass(pb, iterations, false, s);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return failCuda(c, e, "kernel launch");
        CU(c, cudaStreamSynchronize(s));
        if (kernelMs) {
            float ms = 0; for (int q = 0; q < 5; ++q) kernelMs[q] = 0;
            cudaEventElapsedTime(&ms, c->evPool[0], c->evPool[1]); kernelMs[0] = ms;
            for (int it = 0; it < iterations; ++it) {
                cudaEventElapsedTime(&ms, c->evPool[1 + 3 * it], c->evPool[2 + 3 * it]); kernelMs[1] += ms; if (it == 0) kernelMs[4] = ms;
                cudaEventElapsedTime(&ms, c->evPool[2 + 3 * it], c->evPool[3 + 3 * it]); kernelMs[2] += ms;
                cudaEventElapsedTime(&ms, c->evPool[3 + 3 * it], c->evPool[4 + 3 * it]); kernelMs[3] += ms;
            }
        }
        if (std::getenv("HRC_DEBUG_ITERATIONS")) {          // per-iteration ray counts and kernel times of this pass, to stderr
            std::vector<IterCtrl> hc(iterations);
            cudaMemcpy(hc.data(), pb.ctrl, sizeof(IterCtrl) * iterations, cudaMemcpyDeviceToHost);
            for (int it = 0; it < iterations; ++it) {
                float a = 0, b = 0, d = 0;
                cudaEventElapsedTime(&a, c->evPool[1 + 3 * it], c->evPool[2 + 3 * it]);
                cudaEventElapsedTime(&b, c->evPool[2 + 3 * it], c->evPool[3 + 3 * it]);
                cudaEventElapsedTime(&d, c->evPool[3 + 3 * it], c->evPool[4 + 3 * it]);
                std::fprintf(stderr, "[hrc] iter %2d ext %8u (miss %8u pbr %8u glass %8u) closest %.3f ms | shade %.3f ms | occ %8u any %.3f ms\n", it, hc[it].extCount,
                             hc[it].classCount[0], hc[it].classCount[1], hc[it].classCount[2], a, b, hc[it].shadowCount, d);
            }
        }
        launchClearCtrl(pb, s);                             // the invariant every other pass relies on: control words are zero between passes
        return HRC_OK;
    }

    if (c->useGraphs) {
        GraphKey key; std::memset(&key, 0, sizeof(key));
        key.sc = sc; key.pb = pb; key.lc = lc; key.iterations = iterations; key.total = total; key.slot = slot; key.overlap = c->overlap ? 1 : 0;
        std::vector<GraphEntry>& cache = L ? L->graphs : c->graphs;
        GraphEntry* g = nullptr;
        for (auto& e : cache) if (std::memcmp(&e.key, &key, sizeof(key)) == 0) { g = &e; break; }
        if (!g) {
            cudaGraph_t graph = nullptr; cudaGraphExec_t exec = nullptr;
            cudaError_t e = cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
            if (e == cudaSuccess) {
                issuePass(sc, slot, total, pb, lc, iterations, c->overlap, s, s2, evSync);
                e = cudaStreamEndCapture(s, &graph);
            }
            if (e == cudaSuccess) e = cudaGraphInstantiate(&exec, graph, 0);
            if (graph) cudaGraphDestroy(graph);
            if (e != cudaSuccess) { cudaGetLastError(); return failCuda(c, e, "capturing the pass as a CUDA graph (HRC_TUNE_GRAPHS=0 launches the kernels one by one)"); }
            if (cache.size() >= 4) {                        // a lane sees few distinct shapes (band set, depth, counters on / off): drop the least recently used
                size_t lru = 0; for (size_t i = 1; i < cache.size(); ++i) if (cache[i].lastUse < cache[lru].lastUse) lru = i;
                cudaGraphExecDestroy(cache[lru].exec); cache.erase(cache.begin() + lru);
            }
            GraphEntry ne; ne.key = key; ne.exec = exec; cache.push_back(ne); g = &cache.back();
            if (std::getenv("HRC_DEBUG_GRAPHS")) std::fprintf(stderr, "[hrc] captured a pass graph: lane slot %d, %d iterations, %u pixels (%zu cached on this lane)\n", slot, iterations, total, cache.size());
        }
        g->lastUse = ++c->graphClock;
        CU(c, cudaGraphLaunch(g->exec, s));
        return HRC_OK;
    }
    issuePass(sc, slot, total, pb, lc, iterations, c->overlap, s, s2, evSync);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return failCuda(c, e, "kernel launch");
    return HRC_OK;
}

static int readStats(HrcContext c, DeviceStats* out) {     // callers have joined the lanes: the main stream is ordered after lane 1
    DeviceStats x[kMaxLanes - 1 + kDeferredLanes] = {};
    CU(c, cudaMemcpyAsync(out, c->pb.stats, sizeof(DeviceStats), cudaMemcpyDeviceToHost, c->stream));
    for (int k = 0; k < kMaxLanes - 1; ++k)
        if (c->extra[k].pb.stats) CU(c, cudaMemcpyAsync(&x[k], c->extra[k].pb.stats, sizeof(DeviceStats), cudaMemcpyDeviceToHost, c->stream));
    for (int k = 0; k < kDeferredLanes; ++k)            // deferred lanes: counters of passes that finished (a pending pass adds its own when it ends)
        if (c->deferred[k].pb.stats) CU(c, cudaMemcpyAsync(&x[kMaxLanes - 1 + k], c->deferred[k].pb.stats, sizeof(DeviceStats), cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    for (const DeviceStats& b : x) {
    out->raysPrimary += b.raysPrimary; out->raysExtension += b.raysExtension; out->raysOcclusion += b.raysOcclusion;
    out->nodesVisited += b.nodesVisited; out->trianglesTested += b.trianglesTested; out->nodesVisitedAny += b.nodesVisitedAny;
    out->trianglesTestedAny += b.trianglesTestedAny; out->stackOverflows += b.stackOverflows;
    out->nodesVisitedPrimary += b.nodesVisitedPrimary; out->trianglesTestedPrimary += b.trianglesTestedPrimary;
    }
    return HRC_OK;
}

// hrcRenderPass under HRC_OPT_SPECULATIVE_PASSES: commit the predicted pass if the prediction holds, else render it now; then keep
// specDepth passes of the same progressive render running ahead in the deferred lanes.
static int renderOrCommitPredicted(HrcContext c, const HrcFrameParams* fp, int* launches) {
    if (!fp) return fail(c, HRC_ERR_INVALID_VALUE, "null pointer");
    int r = HRC_OK;
    const int oldest = c->oldestDeferred();
    const bool hit = c->deferredCount > 0 && c->deferredVersion[oldest] == c->stateVersion && std::memcmp(&c->deferredParams[oldest], fp, sizeof(*fp)) == 0;
    if (hit) { c->deferred[oldest].active = true; c->deferredCount--; }            // the next join folds it in (and waits for it, in stream order)
    else {
        if (c->deferredCount > 0) { r = hrcDiscardDeferred(c); if (r) return r; }
        r = enqueuePass(c, fp, false, nullptr, launches); if (r) return r;
    }
    if (fp->interactiveMode) return HRC_OK;                                         // block passes change their parameters every pass: not predicted
    HrcFrameParams nx = *fp;
    for (int ahead = c->deferredCount + 1; ahead <= c->specDepth; ++ahead) {
        nx.sampleIndex = fp->sampleIndex + ahead * c->specStride;
        if ((float)nx.sampleIndex >= fp->maxSampleIndex) break;                      // the progressive render ends at maxRenderPasses (PassGenerator.cpp:368)
        const int d = c->deferredNext;
        if (c->deferred[d].active) joinLanes(c);                                    // its previous pass was committed but not folded in yet
        r = ensureLane(c, c->deferred[d]); if (r) return r;
        r = enqueuePass(c, &nx, false, nullptr, nullptr, -(d + 1), true); if (r) return r;
        c->deferredParams[d] = nx; c->deferredVersion[d] = c->stateVersion;
        c->deferredNext = (d + 1) % kDeferredLanes; c->deferredCount++;
    }
    return HRC_OK;
}

HRC_API int hrcRenderPass(HrcContext c, const HrcFrameParams* fp, HrcPassStats* stats) {
    if (!c) return fail(nullptr, HRC_ERR_INVALID_VALUE, "null context");
    cudaSetDevice(c->device); joinLanes(c);
    float kms[5] = {0, 0, 0, 0, 0}; int launches[3] = {0, 0, 0};
    DeviceStats before; int r = readStats(c, &before); if (r) return r;
    CU(c, cudaEventRecord(c->evStart, c->stream));
    if (c->specDepth > 0 && !(c->kernelTiming && stats)) { r = renderOrCommitPredicted(c, fp, launches); if (r) return r; joinLanes(c); }
    else { if (c->deferredCount > 0 && c->specDepth > 0) { r = hrcDiscardDeferred(c); if (r) return r; }
    r = enqueuePass(c, fp, c->kernelTiming && stats, kms, launches); if (r) return r; }
    CU(c, cudaEventRecord(c->evStop, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    cudaError_t e = cudaGetLastError(); if (e != cudaSuccess) return failCuda(c, e, "render pass");
    if (stats) {
        DeviceStats after; r = readStats(c, &after); if (r) return r;
        std::memset(stats, 0, sizeof(*stats));
        cudaEventElapsedTime(&stats->gpuMs, c->evStart, c->evStop);
        stats->raysPrimary = after.raysPrimary - before.raysPrimary; stats->raysExtension = after.raysExtension - before.raysExtension;
        stats->raysOcclusion = after.raysOcclusion - before.raysOcclusion;
        stats->nodesVisited = after.nodesVisited - before.nodesVisited; stats->trianglesTested = after.trianglesTested - before.trianglesTested;
        stats->nodesVisitedAny = after.nodesVisitedAny - before.nodesVisitedAny; stats->trianglesTestedAny = after.trianglesTestedAny - before.trianglesTestedAny;
        stats->closestLaunches = launches[0]; stats->shadeLaunches = launches[1]; stats->anyLaunches = launches[2];
        stats->msRaygen = kms[0]; stats->msClosest = kms[1]; stats->msShade = kms[2]; stats->msAny = kms[3]; stats->msClosestPrimary = kms[4];
        stats->nodesVisitedPrimary = after.nodesVisitedPrimary - before.nodesVisitedPrimary; stats->trianglesTestedPrimary = after.trianglesTestedPrimary - before.trianglesTestedPrimary;
        if (after.stackOverflows != before.stackOverflows) return fail(c, HRC_ERR_INVALID_OPERATION, "traversal stack overflow");
    }
    return HRC_OK;
}
// Asynchronous pass. Consecutive calls rotate over the lanes (see HrcContext_t::extra); any other entry point joins them.
HRC_API int hrcEnqueuePass(HrcContext c, const HrcFrameParams* fp) {
    if (!c) return fail(nullptr, HRC_ERR_INVALID_VALUE, "null context");
    int lane = 0;
    const bool pipelined = c->passLanes > 1 && c->accum && !c->kernelTiming;
    if (pipelined) {
        lane = c->nextLane;
        if (lane > 0) { cudaSetDevice(c->device); int r = ensureLane(c, c->extra[lane - 1]); if (r) return r; }
        c->nextLane = (c->nextLane + 1) % c->passLanes;
    }
    const int r = enqueuePass(c, fp, false, nullptr, nullptr, lane, pipelined);
    if (r == HRC_OK && lane > 0) c->extra[lane - 1].active = true;
    return r;
}
// Deferred pass: rendered like hrcEnqueuePass, but into a side buffer that no other call sees until hrcCommitDeferred. Up to
// kDeferredLanes - 1 may be pending; they are committed oldest first. Lets a caller that already knows its next passes start them
// before it reads the current result back (hr::PassGenerator look-ahead).
HRC_API int hrcEnqueuePassDeferred(HrcContext c, const HrcFrameParams* fp) {
    if (!c) return fail(nullptr, HRC_ERR_INVALID_VALUE, "null context");
    if (c->specDepth > 0) return fail(c, HRC_ERR_INVALID_OPERATION, "the deferred lanes are in use by HRC_OPT_SPECULATIVE_PASSES");
    if (c->deferredCount >= kDeferredLanes - 1) return fail(c, HRC_ERR_INVALID_OPERATION, "too many deferred passes are pending (commit or discard first)");
    if (!c->accum) return fail(c, HRC_ERR_INVALID_OPERATION, "render pass without a framebuffer");
    cudaSetDevice(c->device);
    const int d = c->deferredNext;
    if (c->deferred[d].active) joinLanes(c);                      // its previous pass was committed but not folded in yet
    int r = ensureLane(c, c->deferred[d]); if (r) return r;
    r = enqueuePass(c, fp, false, nullptr, nullptr, -(d + 1), c->deferredCount >= 1);     // several in flight: the pipelined launch shape (fewer traversal CTAs per SM each)
    if (r == HRC_OK) { c->deferredNext = (d + 1) % kDeferredLanes; c->deferredCount++; }
    return r;
}
HRC_API int hrcCommitDeferred(HrcContext c) {
    if (!c) return fail(nullptr, HRC_ERR_INVALID_VALUE, "null context");
    if (c->deferredCount <= 0) return fail(c, HRC_ERR_INVALID_OPERATION, "no deferred pass is pending");
    c->deferred[c->oldestDeferred()].active = true;               // the next join folds it in (and waits for it, in stream order)
    c->deferredCount--;
    return HRC_OK;
}
HRC_API int hrcDiscardDeferred(HrcContext c) {
    if (!c) return fail(nullptr, HRC_ERR_INVALID_VALUE, "null context");
    if (c->deferredCount <= 0) return HRC_OK;
    cudaSetDevice(c->device);
    for (int d = c->oldestDeferred(); c->deferredCount > 0; d = (d + 1) % kDeferredLanes, c->deferredCount--) {      // all pending passes
        HrcContext_t::Lane& L = c->deferred[d];
        CU(c, cudaMemsetAsync(L.accum, 0, (size_t)c->w * c->h * sizeof(float4), L.s));      // after its kernels, in stream order
        CU(c, cudaStreamSynchronize(L.s));                      // rare (reset, scene change): whatever follows may free what the pass was reading
    }
    return HRC_OK;
}
HRC_API int hrcSynchronize(HrcContext c) {
    if (!c) return fail(nullptr, HRC_ERR_INVALID_VALUE, "null context");
    cudaSetDevice(c->device); joinLanes(c); waitPendingDeferred(c);
    CU(c, cudaStreamSynchronize(c->stream));
    return HRC_OK;
}
HRC_API int hrcResetStats(HrcContext c) {
    if (!c) return fail(nullptr, HRC_ERR_INVALID_VALUE, "null context");
    cudaSetDevice(c->device); joinLanes(c);
    CU(c, cudaMemsetAsync(c->pb.stats, 0, sizeof(DeviceStats), c->stream));
    for (auto& L : c->extra) if (L.pb.stats) CU(c, cudaMemsetAsync(L.pb.stats, 0, sizeof(DeviceStats), L.s));   // the lane's own stream already waits for the join
    for (auto& L : c->deferred) if (L.pb.stats) CU(c, cudaMemsetAsync(L.pb.stats, 0, sizeof(DeviceStats), L.s));
    return HRC_OK;
}
HRC_API int hrcTimerStart(HrcContext c) {
    if (!c) return fail(nullptr, HRC_ERR_INVALID_VALUE, "null context");
    cudaSetDevice(c->device); joinLanes(c);
    CU(c, cudaEventRecord(c->evTimer0, c->stream));
    return HRC_OK;
}
HRC_API int hrcTimerStop(HrcContext c, float* ms) {
    if (!c || !ms) return fail(c, HRC_ERR_INVALID_VALUE, "null pointer");
    cudaSetDevice(c->device); joinLanes(c);
    CU(c, cudaEventRecord(c->evTimer1, c->stream));
    CU(c, cudaEventSynchronize(c->evTimer1));
    CU(c, cudaEventElapsedTime(ms, c->evTimer0, c->evTimer1));
    return HRC_OK;
}
HRC_API int hrcGetStats(HrcContext c, HrcPassStats* stats) {
    if (!c || !stats) return fail(c, HRC_ERR_INVALID_VALUE, "null pointer");
    cudaSetDevice(c->device); joinLanes(c);
    DeviceStats d; int r = readStats(c, &d); if (r) return r;
    std::memset(stats, 0, sizeof(*stats));
    stats->raysPrimary = d.raysPrimary; stats->raysExtension = d.raysExtension; stats->raysOcclusion = d.raysOcclusion;
    stats->nodesVisited = d.nodesVisited; stats->trianglesTested = d.trianglesTested;
    stats->nodesVisitedAny = d.nodesVisitedAny; stats->trianglesTestedAny = d.trianglesTestedAny;
    stats->nodesVisitedPrimary = d.nodesVisitedPrimary; stats->trianglesTestedPrimary = d.trianglesTestedPrimary;
    return HRC_OK;
}

HRC_API int hrcTraceRays(HrcContext c, int32_t n, const HrcRay* rays, int32_t anyHit, HrcHit* hits) {
    if (!c) return fail(nullptr, HRC_ERR_INVALID_VALUE, "null context");
    if (!c->committed) return fail(c, HRC_ERR_INVALID_OPERATION, "trace without a committed scene");
    if (n < 0 || (n && (!rays || !hits))) return fail(c, HRC_ERR_INVALID_VALUE, "bad ray buffers");
    if (n == 0) return HRC_OK;
    cudaSetDevice(c->device);
    HrcRay* dr = nullptr; HrcHit* dh = nullptr;
    CU(c, cudaMalloc(&dr, (size_t)n * sizeof(HrcRay)));
    cudaError_t e = cudaMalloc(&dh, (size_t)n * sizeof(HrcHit));
    if (e != cudaSuccess) { cudaFree(dr); return failCuda(c, e, "cudaMalloc"); }
    cudaMemcpyAsync(dr, rays, (size_t)n * sizeof(HrcRay), cudaMemcpyHostToDevice, c->stream);
    launchTraceRays(sceneOf(c), dr, dh, n, anyHit != 0, configOf(c), c->stream);
    cudaMemcpyAsync(hits, dh, (size_t)n * sizeof(HrcHit), cudaMemcpyDeviceToHost, c->stream);
    e = cudaStreamSynchronize(c->stream);
    cudaFree(dr); cudaFree(dh);
    if (e != cudaSuccess) return failCuda(c, e, "hrcTraceRays");
    return HRC_OK;
}

// hrcTraceRays without allocations (include/heatray_cuda.h): rays and hits are staged in the library's static device array.
static std::mutex g_noAllocCallMutex;      // one call at a time per process: the array is shared by every context on a device
HRC_API int hrcTraceRaysNoAlloc(HrcContext c, int32_t n, const HrcRay* rays, int32_t anyHit, HrcHit* hits) {
    if (!c) return fail(nullptr, HRC_ERR_INVALID_VALUE, "null context");
    if (!c->committed) return fail(c, HRC_ERR_INVALID_OPERATION, "trace without a committed scene");
    if (n < 0 || (n && (!rays || !hits))) return fail(c, HRC_ERR_INVALID_VALUE, "bad ray buffers");
    if (n > HRC_NO_ALLOC_MAX_RAYS) return fail(c, HRC_ERR_INVALID_VALUE, "hrcTraceRaysNoAlloc takes at most HRC_NO_ALLOC_MAX_RAYS (262144) rays per call: split the set or use hrcTraceRays");
    if (n == 0) return HRC_OK;
    cudaSetDevice(c->device);
    std::lock_guard<std::mutex> lock(g_noAllocCallMutex);
    HrcRay* dr = nullptr; HrcHit* dh = nullptr;
    cudaError_t e = traceScratchBuffers(c->device, &dr, &dh);
    if (e != cudaSuccess) return failCuda(c, e, "cudaGetSymbolAddress of the static trace scratch");
    CU(c, cudaMemcpyAsync(dr, rays, (size_t)n * sizeof(HrcRay), cudaMemcpyHostToDevice, c->stream));
    launchTraceRays(sceneOf(c), dr, dh, n, anyHit != 0, configOf(c), c->stream);
    CU(c, cudaMemcpyAsync(hits, dh, (size_t)n * sizeof(HrcHit), cudaMemcpyDeviceToHost, c->stream));
    e = cudaStreamSynchronize(c->stream);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) return failCuda(c, e, "hrcTraceRaysNoAlloc");
    return HRC_OK;
}

HRC_API int hrcReadRayCounts(HrcContext c, uint32_t* hostOut) {
    if (!c || !hostOut) return fail(c, HRC_ERR_INVALID_VALUE, "null pointer");
    if (!c->dRayCounts) return fail(c, HRC_ERR_INVALID_OPERATION, "HRC_OPT_PIXEL_RAY_COUNTS is off");
    cudaSetDevice(c->device); joinLanes(c);
    CU(c, cudaMemcpyAsync(hostOut, c->dRayCounts, (size_t)c->w * c->h * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    return HRC_OK;
}

HRC_API int hrcGetSceneInfo(HrcContext c, HrcSceneInfo* info) {
    if (!c || !info) return fail(c, HRC_ERR_INVALID_VALUE, "null pointer");
    *info = c->info; return HRC_OK;
}

} // extern "C"

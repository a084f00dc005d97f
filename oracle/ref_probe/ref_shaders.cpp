// ref_shaders.cpp — TEST INFRASTRUCTURE (build container only) -> oracle/_ref/libref_shaders.so.
// The reference's shader programs themselves — perspective.rlsl, physicallyBased.rlsl, glass.rlsl, environmentLight.rlsl,
// directionalLight.rlsl, pointLight.rlsl, spotLight.rlsl with everything they include (sequence, utility, brdfs, lightDefines,
// lightSampling, microfacet, accumulator, globalData, rayAttributes) — compiled as C++ through glm from the text where it lies
// under /root/reference (rewritten by rlsl_to_cpp.py: only RLSL-only declarations change, no arithmetic), one namespace per
// program and per #define set. rlsl_shim.h supplies the RLSL built-ins. The exports run a shader main() on the probe records of
// oracle/shade_probe.h so tests can pin the oracle's restatement (oracle/oracle.cpp) to the shader text.
// Build: oracle/ref_probe/build_ref.sh (g++ -std=gnu++20 -D_MSC_EXTENSIONS -O2 -ffp-contract=off -fsingle-precision-constant).
#include "rlsl_shim.h"
#include <glm/glm/gtc/type_ptr.hpp>

#include "../../include/heatray_cuda.h"
#include "../shade_probe.h"

#include <map>

#define MAX_NUM_DIRECTIONAL_LIGHTS 5      // ShaderLightingDefines::appendLightingShaderDefines (ShaderLightingDefines.h:26-31)
#define MAX_NUM_POINT_LIGHTS 5
#define MAX_NUM_SPOT_LIGHTS 5

namespace {

struct ProbeTexture { std::vector<float> rgba; int w = 0, h = 0, wrapS = 0, wrapT = 0; };

struct ProbeState {
    HrcGlobals globals{};
    HrcEnvLight env{};
    int nDir = 0, nPoint = 0, nSpot = 0;
    HrcDirLight dir[5]{}; HrcPointLight point[5]{}; HrcSpotLight spot[5]{};
    int numSequences = 0, sequenceLength = 0;
    std::vector<float> seq, aperture, offsets;      // vec2 tables
    std::map<int, ProbeTexture> textures;
    std::vector<float> blockCoords; int blockW = 0, blockH = 0;   // RGBA expansion of the interactive block table
    sampler2D sampler(int handle) const {
        sampler2D s;
        auto it = textures.find(handle);
        if (handle < 0 || it == textures.end()) return s;          // dummy 1x1 white
        s.rgba = it->second.rgba.data(); s.w = it->second.w; s.h = it->second.h; s.wrapS = it->second.wrapS; s.wrapT = it->second.wrapT;
        return s;
    }
    sampler2D blockTexture() const {
        sampler2D s;
        if (blockCoords.empty()) return s;
        s.rgba = blockCoords.data(); s.w = blockW; s.h = blockH; s.wrapS = s.wrapT = 1; s.nearest = 1;   // NEAREST + REPEAT (PassGenerator.cpp:266-294)
        return s;
    }
};
ProbeState g_state;

RLRay toRLRay(const OrcProbeRay& r) {
    RLRay o;
    o.origin = vec3(r.origin[0], r.origin[1], r.origin[2]); o.direction = vec3(r.direction[0], r.direction[1], r.direction[2]);
    o.depth = r.depth; o.maxT = r.maxT; o.occlusionTest = r.occlusionTest != 0;
    o.defaultPrimitive = r.missKind ? makePrimitive(r.missKind, r.missIndex) : rl_NullPrimitive;
    o.weight = vec3(r.weight[0], r.weight[1], r.weight[2]); o.sequenceID = r.sequenceID; o.sequenceIndexOffset = r.sequenceIndexOffset; o.extraT = r.extraT;
    return o;
}
OrcProbeRay fromRLRay(const RLRay& r) {
    OrcProbeRay o{};
    for (int k = 0; k < 3; ++k) { o.origin[k] = r.origin[k]; o.direction[k] = r.direction[k]; o.weight[k] = r.weight[k]; }
    o.maxT = r.maxT; o.extraT = r.extraT; o.depth = r.depth; o.sequenceID = r.sequenceID; o.sequenceIndexOffset = r.sequenceIndexOffset;
    o.occlusionTest = r.occlusionTest ? 1 : 0; o.missKind = r.defaultPrimitive / 16; o.missIndex = r.defaultPrimitive % 16;
    return o;
}
void setHit(const OrcShadeCase& c) {
    rl_isFrameShader = false;
    rl_InRay = toRLRay(c.in);
    rl_IntersectionPoint = vec3(c.P[0], c.P[1], c.P[2]);
    rl_IntersectionT = c.T;
    rl_FrontFacing = c.frontFacing != 0;
}
void collect(OrcShadeResult& out) {
    out = OrcShadeResult{};
    out.rayCount = (int)rl_emitted.size(); out.accumCount = (int)rl_accumulated.size();
    for (int i = 0; i < out.rayCount && i < 2; ++i) out.rays[i] = fromRLRay(rl_emitted[i]);
    for (const RLAccum& a : rl_accumulated) for (int k = 0; k < 4; ++k) out.accumSum[k] += a.value[k];
    rl_emitted.clear(); rl_accumulated.clear();
}

typedef void (*MaterialRun)(const void* params, const OrcShadeCase& c, int sampleIndex);
struct MaterialVariant { uint32_t flags; MaterialRun run; };
typedef void (*VertexRun)(const OrcVertexCase& c, OrcVertexResult& o);
struct VertexVariant { uint32_t flags; VertexRun run; };

namespace frame_shader {
#define PROGRAM_KIND 0
#include "perspective.inc"
#include "ref_shaders_program.inc"
#undef PROGRAM_KIND
}
namespace env_light {
#define PROGRAM_KIND 3
#include "environmentLight.inc"
#include "ref_shaders_program.inc"
#undef PROGRAM_KIND
}
namespace dir_light {
#define PROGRAM_KIND 4
#include "directionalLight.inc"
#include "ref_shaders_program.inc"
#undef PROGRAM_KIND
}
namespace point_light {
#define PROGRAM_KIND 5
#include "pointLight.inc"
#include "ref_shaders_program.inc"
#undef PROGRAM_KIND
}
namespace spot_light {
#define PROGRAM_KIND 6
#include "spotLight.inc"
#include "ref_shaders_program.inc"
#undef PROGRAM_KIND
}
#include "variants.inc"

} // namespace

#define REF_API extern "C" __attribute__((visibility("default")))

REF_API void ref_set_globals(const HrcGlobals* g) { g_state.globals = *g; }
REF_API void ref_set_lights(const HrcEnvLight* env, int nDir, const HrcDirLight* d, int nPoint, const HrcPointLight* p, int nSpot, const HrcSpotLight* s) {
    g_state.env = *env; g_state.nDir = nDir; g_state.nPoint = nPoint; g_state.nSpot = nSpot;
    for (int i = 0; i < 5; ++i) {
        g_state.dir[i] = i < nDir ? d[i] : HrcDirLight{}; g_state.point[i] = i < nPoint ? p[i] : HrcPointLight{}; g_state.spot[i] = i < nSpot ? s[i] : HrcSpotLight{};
    }
}
REF_API void ref_set_sequences(int ns, int len, const float* xy, const float* aperture) {
    g_state.numSequences = ns; g_state.sequenceLength = len;
    g_state.seq.assign(xy, xy + 2 * (size_t)ns * len); g_state.aperture.assign(aperture, aperture + 2 * (size_t)ns * len);
}
// SequenceOffsets block: vec2 per entry in the reference (PassGenerator.cpp:150-159); the backend keeps .x only, so .y is filled with 0 here.
REF_API void ref_set_pixel_offsets(const float* offsetX, size_t count) {
    g_state.offsets.assign(2 * count, 0.0f);
    for (size_t i = 0; i < count; ++i) g_state.offsets[2 * i] = offsetX[i];
}
REF_API void ref_set_interactive_blocks(int bw, int bh, const float* rgb) {
    g_state.blockW = bw; g_state.blockH = bh; g_state.blockCoords.assign(4 * (size_t)bw * bh, 0.0f);
    for (int i = 0; i < bw * bh; ++i) for (int k = 0; k < 3; ++k) g_state.blockCoords[4 * i + k] = rgb[3 * i + k];
}
REF_API void ref_set_texture(int handle, int w, int h, const float* rgba, int wrapS, int wrapT) {
    ProbeTexture& t = g_state.textures[handle];
    t.w = w; t.h = h; t.wrapS = wrapS; t.wrapT = wrapT; t.rgba.assign(rgba, rgba + 4 * (size_t)w * h);
}
REF_API void ref_clear_textures() { g_state.textures.clear(); }

// Runs the material program compiled for (type, flags) on n hits. Returns -1 when that #define set was not compiled.
REF_API int ref_shade(int type, uint32_t flags, const void* params, int sampleIndex, int n, const OrcShadeCase* cases, OrcShadeResult* results) {
    const MaterialVariant* table = type == HRC_MATERIAL_PBR ? kPbrVariants : kGlassVariants;
    const size_t count = type == HRC_MATERIAL_PBR ? sizeof(kPbrVariants) / sizeof(kPbrVariants[0]) : sizeof(kGlassVariants) / sizeof(kGlassVariants[0]);
    for (size_t v = 0; v < count; ++v) {
        if (table[v].flags != flags) continue;
        for (int i = 0; i < n; ++i) { table[v].run(params, cases[i], sampleIndex); collect(results[i]); }
        return 0;
    }
    return -1;
}
REF_API int ref_variant_count(int type) { return type == HRC_MATERIAL_PBR ? (int)(sizeof(kPbrVariants) / sizeof(kPbrVariants[0])) : (int)(sizeof(kGlassVariants) / sizeof(kGlassVariants[0])); }
REF_API uint32_t ref_variant_flags(int type, int i) { return type == HRC_MATERIAL_PBR ? kPbrVariants[i].flags : kGlassVariants[i].flags; }

// Light ("miss") shaders: rays[i].missKind / missIndex select the program and its lightIndex uniform; maxT is what the shader sees as rl_IntersectionT.
REF_API void ref_miss(int n, const OrcProbeRay* rays, OrcShadeResult* results) {
    for (int i = 0; i < n; ++i) {
        switch (rays[i].missKind) {
        case 1: env_light::run(rays[i]); break;
        case 2: dir_light::run(rays[i]); break;
        case 3: point_light::run(rays[i]); break;
        case 4: spot_light::run(rays[i]); break;
        default: break;
        }
        collect(results[i]);
    }
}
// Frame shader for pixels xy[2i], xy[2i+1] of a w x h frame.
REF_API void ref_raygen(const HrcFrameParams* fp, int w, int h, int n, const int32_t* xy, OrcShadeResult* results) {
    for (int i = 0; i < n; ++i) { frame_shader::run(*fp, w, h, xy[2 * i], xy[2 * i + 1]); collect(results[i]); }
}
// vertex.rlsl for vertices of one #define set (bit 1 HAS_TEXTURES, 2 USE_TANGENT_SPACE, 4 VERTEX_COLORS); -1 when that set was not compiled.
REF_API int ref_vertex(uint32_t flags, int n, const OrcVertexCase* cases, OrcVertexResult* results) {
    for (const VertexVariant& v : kVertexVariants)
        if (v.flags == flags) { for (int i = 0; i < n; ++i) v.run(cases[i], results[i]); return 0; }
    return -1;
}

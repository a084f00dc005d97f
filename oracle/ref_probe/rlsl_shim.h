// rlsl_shim.h — TEST INFRASTRUCTURE. The RLSL built-ins the reference's shaders use, as plain C++ over glm, so that the shader
// programs rewritten by rlsl_to_cpp.py compile and run on the host. Nothing here is reference code: these are the OpenRL
// semantics the shaders rely on, each with the policy that fixes what OpenRL leaves unpinned.
#pragma once
#define GLM_FORCE_SWIZZLE
#include <glm/glm/glm.hpp>

#include <cmath>
#include <cstring>
#include <limits>
#include <vector>

using namespace glm;

// GLSL's two-argument atan
static inline float atan(float y, float x) { return atan2f(y, x); }

// ---- primitives ------------------------------------------------------------------------------------------------------
typedef int primitive;
static const primitive rl_NullPrimitive = 0;
// The probe encodes a light primitive as kind * 16 + index (kind: 1 environment, 2 directional, 3 point, 4 spot).
static inline primitive makePrimitive(int kind, int index) { return kind * 16 + index; }

// ---- textures: policy A-12 (level-0 bilinear in fp32, texel centres at (i + 0.5) / N; NEAREST for the interactive block table) --
struct sampler2D {
    const float* rgba = nullptr;   // 4 floats per texel, row 0 = t 0; nullptr = the 1x1 white dummy (Texture.h:178-192)
    int w = 1, h = 1, wrapS = 0, wrapT = 0, nearest = 0;   // wrap: 0 clamp, 1 repeat
};
static inline int shimWrap(int i, int n, int mode) {
    if (mode == 1) { i %= n; if (i < 0) i += n; return i; }
    return std::min(std::max(i, 0), n - 1);
}
static inline vec4 texture2D(const sampler2D& s, vec2 uv) {
    if (!s.rgba) return vec4(1.0f);
    if (s.nearest) {
        int x = shimWrap((int)std::floor(uv.x * (float)s.w), s.w, s.wrapS), y = shimWrap((int)std::floor(uv.y * (float)s.h), s.h, s.wrapT);
        const float* p = s.rgba + 4 * ((size_t)y * s.w + x);
        return vec4(p[0], p[1], p[2], p[3]);
    }
    float x = uv.x * (float)s.w - 0.5f, y = uv.y * (float)s.h - 0.5f;
    float fx0 = std::floor(x), fy0 = std::floor(y);
    float fx = x - fx0, fy = y - fy0;
    int x0 = shimWrap((int)fx0, s.w, s.wrapS), x1 = shimWrap((int)fx0 + 1, s.w, s.wrapS);
    int y0 = shimWrap((int)fy0, s.h, s.wrapT), y1 = shimWrap((int)fy0 + 1, s.h, s.wrapT);
    const float* p00 = s.rgba + 4 * ((size_t)y0 * s.w + x0);
    const float* p10 = s.rgba + 4 * ((size_t)y0 * s.w + x1);
    const float* p01 = s.rgba + 4 * ((size_t)y1 * s.w + x0);
    const float* p11 = s.rgba + 4 * ((size_t)y1 * s.w + x1);
    vec4 r;
    for (int k = 0; k < 4; ++k) {
        float a = p00[k] * (1.0f - fx) + p10[k] * fx;
        float b = p01[k] * (1.0f - fx) + p11[k] * fx;
        r[k] = a * (1.0f - fy) + b * fy;
    }
    return r;
}

// ---- rays ------------------------------------------------------------------------------------------------------------
struct RLRay {
    vec3 origin = vec3(0.0f), direction = vec3(0.0f);
    int depth = 0;
    float maxT = std::numeric_limits<float>::infinity();
    bool occlusionTest = false;
    primitive defaultPrimitive = rl_NullPrimitive;
    // rayattributes (rayAttributes.rlsl:8-11)
    vec3 weight = vec3(0.0f);
    int sequenceID = 0;
    int sequenceIndexOffset = 0;
    float extraT = 0.0f;
};

struct RLAccum { vec4 value; bool hasAlpha; };

// Per-invocation state shared by all programs.
static RLRay rl_InRay, rl_OutRay;
static vec3 rl_IntersectionPoint;
static float rl_IntersectionT;
static bool rl_FrontFacing;
static vec4 rl_FrameCoord, rl_FrameSize;
static vec4 rl_Position;                             // vertex shaders
static bool rl_isFrameShader = false;
static std::vector<RLRay> rl_emitted;
static std::vector<RLAccum> rl_accumulated;
struct RLOutputRayCount { int n[1]; int& operator[](int i) { return n[i]; } RLOutputRayCount& operator=(int v) { n[0] = v; return *this; } };
static RLOutputRayCount rl_OutputRayCount;

// createRay(): policy A-3 — a ray shader's new ray starts as a copy of the incoming ray (attributes, maxT, occlusionTest,
// defaultPrimitive) leaving the intersection point, one level deeper; the frame shader's starts from the defaults.
static inline void createRay() {
    if (rl_isFrameShader) { rl_OutRay = RLRay(); return; }
    rl_OutRay = rl_InRay;
    rl_OutRay.origin = rl_IntersectionPoint;
    rl_OutRay.depth = rl_InRay.depth + 1;
}
static inline void emitRay() { rl_emitted.push_back(rl_OutRay); }
// accumulate(): policy A-2 — additive; the vec3 form leaves alpha untouched.
static inline void accumulate(vec3 c) { rl_accumulated.push_back(RLAccum{vec4(c, 0.0f), false}); }
static inline void accumulate(vec4 c) { rl_accumulated.push_back(RLAccum{c, true}); }

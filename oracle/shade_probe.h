/* shade_probe.h — TEST INFRASTRUCTURE. Plain-C records for the shader-level parity probes: the same randomised shader
 * invocations are run (a) through the reference's RLSL shader text compiled as C++ (oracle/_ref/libref_shaders.so,
 * oracle/ref_probe/ref_shaders.cpp) and (b) through the oracle's restatement (orcProbeShade / orcProbeMiss / orcProbeRaygen in
 * oracle/oracle.cpp), and compared field by field by tests/test_shader_pinning.py.
 * A record describes what a ray shader sees (rl_InRay + the intersection built-ins + the interpolated varyings) and what it
 * does (rays passed to emitRay(), values passed to accumulate()). */
#ifndef HEATRAY_SHADE_PROBE_H
#define HEATRAY_SHADE_PROBE_H
#include <stdint.h>

typedef struct OrcProbeRay {
    float origin[3];    float maxT;                 /* rl_*Ray.origin, .maxT */
    float direction[3]; float extraT;               /* .direction, rayattribute extraT */
    float weight[3];    int32_t depth;              /* rayattribute weight, .depth */
    int32_t sequenceID, sequenceIndexOffset;        /* rayattributes (rayAttributes.rlsl:8-11) */
    int32_t occlusionTest;                          /* .occlusionTest */
    int32_t missKind, missIndex;                    /* .defaultPrimitive as (kind, index): 0 rl_NullPrimitive, 1 environment, 2 directional[i], 3 point[i], 4 spot[i] */
    int32_t pad;
} OrcProbeRay;                                      /* 72 bytes */

typedef struct OrcShadeCase {
    OrcProbeRay in;                                 /* rl_InRay */
    float P[3];  float T;                           /* rl_IntersectionPoint, rl_IntersectionT */
    float Ng[3]; int32_t frontFacing;               /* geometric normal (ray-epsilon policy A-8 only; no shader reads it), rl_FrontFacing */
    float normal[3]; int32_t material;              /* varying normal; material handle (oracle side) */
    float texCoord[2]; float pad0[2];               /* varying texCoord */
    float tangent[3]; float pad1;                   /* varying tangent */
    float bitangent[3]; float pad2;                 /* varying bitangent */
    float vertexColor[3]; float pad3;               /* varying vertexColor */
} OrcShadeCase;

typedef struct OrcShadeResult {
    int32_t rayCount;                               /* emitRay() calls, in call order (direct before indirect) */
    int32_t accumCount;                             /* accumulate() calls (reference side; the oracle reports 1 when the sum is non-zero) */
    OrcProbeRay rays[2];
    float accumSum[4];                              /* sum of everything passed to accumulate(): rgb (+ a for the vec4 form) */
} OrcShadeResult;

/* vertex.rlsl:25-43: one vertex through the vertex program of a material (#define set as bits: 1 HAS_TEXTURES, 2 USE_TANGENT_SPACE, 4 VERTEX_COLORS). */
typedef struct OrcVertexCase {
    float worldFromEntity[16];                      /* column-major, like glm */
    float position[3], normal[3], texCoord[2], tangent[3], bitangent[3], color[3];
} OrcVertexCase;
typedef struct OrcVertexResult {
    float position[4];                              /* rl_Position */
    float normal[3], texCoord[2], tangent[3], bitangent[3], color[3];      /* the varyings; zero where the #define set has none */
} OrcVertexResult;

#endif

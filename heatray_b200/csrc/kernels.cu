// kernels.cu — the wavefront path-tracing kernels (sm_100a).
//
// What each kernel replaces in the reference (all of it ran inside the closed OpenRL library):
//   k_raygen        frame shader          Resources/shaders/perspective.rlsl:39-93
//   k_trace<false>  closest-hit traversal OpenRL internal (policy A-9 in SURVEY.md)
//   k_trace<true>   occlusion traversal + the four light ("miss") shaders
//                   environmentLight.rlsl:19-34, directionalLight.rlsl:20-26, pointLight.rlsl:20-29, spotLight.rlsl:20-36
//   k_shade         material shaders      physicallyBased.rlsl:55-331, glass.rlsl:138-280 (+ microfacet.rlsl,
//                   brdfs.rlsl, lightSampling.rlsl, utility.rlsl, sequence.rlsl, accumulator.rlsl)
//
// No tensor cores: this is pointer chasing + fp32 ALU, bound by the memory system.
#include <mutex>
#include "device_types.h"

#include <cuda_fp16.h>

#include <cfloat>
#include <cstdio>

namespace hr {
namespace {

// Per-pass constants (DFrame) live in constant memory, one slot per lane of passes in flight: a pass uploads its DFrame into its lane's
// slot in stream order (uploadFrame) and every kernel of the pass is told the slot number. The kernel arguments of a pass are then the
// same from pass to pass, which is what lets hrc_api.cu capture the whole pass once as a CUDA graph and replay it.
__constant__ DFrame g_frames[kFrameSlots];

// ---------------------------------------------------------------------------------------
// float3 helpers
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
__device__ __forceinline__ float3 f3(float s) { return make_float3(s, s, s); }
__device__ __forceinline__ float3 f3(const float* p) { return make_float3(p[0], p[1], p[2]); }
__device__ __forceinline__ float3 xyz(float4 v) { return make_float3(v.x, v.y, v.z); }
__device__ __forceinline__ float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ float3 operator-(float3 a) { return f3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ float3 operator*(float3 a, float3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ float3 operator*(float3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ float3 operator*(float s, float3 a) { return f3(a.x * s, a.y * s, a.z * s); }
#ifndef HR_EXACT_DIV
#define HR_EXACT_DIV 0
#endif
#if HR_EXACT_DIV
__device__ __forceinline__ float3 operator/(float3 a, float s) { return f3(__fdiv_rn(a.x, s), __fdiv_rn(a.y, s), __fdiv_rn(a.z, s)); }   // component-wise IEEE division, as GLSL / glm define vec3 / float
#else
__device__ __forceinline__ float3 operator/(float3 a, float s) { const float r = 1.0f / s; return f3(a.x * r, a.y * r, a.z * r); }   // one IEEE reciprocal, <= 1 ulp from a true division
#endif
__device__ __forceinline__ float dot(float3 a, float3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
__device__ __forceinline__ float3 cross(float3 a, float3 b) { return f3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
__device__ __forceinline__ float3 normalize(float3 a) { return a * (1.0f / sqrtf(dot(a, a))); }
__device__ __forceinline__ float length(float3 a) { return sqrtf(dot(a, a)); }
struct M3 { float3 c0, c1, c2; };
__device__ __forceinline__ float3 mul(const M3& m, float3 v) { return (m.c0 * v.x + m.c1 * v.y) + m.c2 * v.z; }
__device__ __forceinline__ float3 mulT(const M3& m, float3 v) { return f3(dot(m.c0, v), dot(m.c1, v), dot(m.c2, v)); }

// ---------------------------------------------------------------------------------------
// utility.rlsl / brdfs.rlsl
// ---------------------------------------------------------------------------------------
#define K_PI 3.14159265359f
#define K_TWO_PI (2.0f * K_PI)
#define K_ONE_OVER_PI (1.0f / K_PI)

__device__ __forceinline__ float saturate(float f) { return fminf(fmaxf(f, 0.0f), 1.0f); }     // utility.rlsl:20-23
__device__ __forceinline__ float square(float x) { return x * x; }
__device__ __forceinline__ float pow5(float x) { return x * square(x) * square(x); }            // :146-149
__device__ __forceinline__ float greaterThanZero(float f) { return fmaxf(1e-5f, f); }           // :158-161
__device__ __forceinline__ float luminosity(float3 c) { return dot(c, f3(0.33f, 0.59f, 0.11f)); } // :168-171

__device__ __forceinline__ M3 orthonormalFrame(float3 N) {                                      // :45-61
    float3 l = f3(N.x, N.z, N.y);
    float s = l.z < 0.f ? -1.f : 1.f;
    float a = -1.0f / (s + l.z);
    float b = l.x * l.y * a;
    float3 X = f3(1.0f + s * l.x * l.x * a, s * b, -s * l.x);
    float3 Z = f3(b, s + l.y * l.y * a, -l.y);
    M3 m; m.c0 = f3(X.x, X.z, X.y); m.c1 = N; m.c2 = f3(Z.x, Z.z, Z.y);
    return m;
}
__device__ __forceinline__ float3 cosineWeightedSample(float u1, float u2) {                    // :65-76
    float theta = sqrtf(u1);
    float phi = K_TWO_PI * u2;
    float sp, cp; sincosf(phi, &sp, &cp);
    return normalize(f3(theta * cp, sqrtf(fmaxf(0.0f, 1.0f - u1)), theta * sp));
}
__device__ __forceinline__ float3 sampleVisibleGGX(float3 localV, float u1, float u2, float alpha) {   // :112-144
    float3 zUpV = f3(localV.x, localV.z, localV.y);
    float3 Vh = normalize(f3(zUpV.x * alpha, zUpV.y * alpha, zUpV.z));
    float lengthSquared = (Vh.x * Vh.x) + (Vh.y * Vh.y);
    float3 T1 = (lengthSquared > 0.0f) ? f3(-Vh.y, Vh.x, 0.0f) * (1.0f / sqrtf(lengthSquared)) : f3(1.0f, 0.0f, 0.0f);
    float3 T2 = cross(Vh, T1);
    float r = sqrtf(u1);
    float phi = K_TWO_PI * u2;
    float sp, cp; sincosf(phi, &sp, &cp);
    float t1 = r * cp, t2 = r * sp;
    float s = 0.5f * (1.0f + Vh.z);
    float t1Squared = square(t1);
    t2 = (1.0f - s) * sqrtf(1.0f - t1Squared) + (s * t2);
    float3 Nh = (t1 * T1) + (t2 * T2) + sqrtf(fmaxf(0.0f, 1.0f - t1Squared - square(t2))) * Vh;
    float3 zUp = normalize(f3(alpha * Nh.x, alpha * Nh.y, fmaxf(0.0f, Nh.z)));
    return f3(zUp.x, zUp.z, zUp.y);
}
__device__ __forceinline__ float3 F_Schlick3(float3 Cspec, float cosTheta) { return Cspec + (f3(1.0f) - Cspec) * pow5(1.0f - cosTheta); }   // brdfs.rlsl:47-51
__device__ __forceinline__ float F_Schlick1(float f0, float cosTheta) { return f0 + (1.0f - f0) * pow5(1.0f - cosTheta); }                  // :54-58
__device__ __forceinline__ float F_Fresnel(float eta, float cosThetaI) {                                                                    // :60-72
    float sinThetaT2 = square(eta) * (1.0f - square(cosThetaI));
    if (sinThetaT2 < 1.0f) {
        float cosThetaT = sqrtf(1.0f - sinThetaT2);
        float perpendicular = square((eta * cosThetaI - cosThetaT) / (eta * cosThetaI + cosThetaT));
        float parallel = square((cosThetaI - eta * cosThetaT) / (cosThetaI + eta * cosThetaT));
        return 0.5f * (perpendicular + parallel);
    }
    return 1.0f;
}
__device__ __forceinline__ float D_GGX(float NdotH, float alpha) {                               // :74-79
    float alpha2 = square(alpha);
    float denominator = square(square(NdotH) * (alpha2 - 1.0f) + 1.0f);
    return K_ONE_OVER_PI * (alpha2 / greaterThanZero(denominator));
}
__device__ __forceinline__ float G1_Smith_GGX(float NdotI, float alpha) {                        // :89-94
    float alpha2 = square(alpha);
    float denom = sqrtf(alpha2 + (1.0f - alpha2) * greaterThanZero(square(NdotI))) + NdotI;
    return (2.0f * NdotI) / greaterThanZero(denom);
}
__device__ __forceinline__ float G2_Smith_GGX(float NdotO, float NdotI, float alpha) { return G1_Smith_GGX(NdotO, alpha) * G1_Smith_GGX(NdotI, alpha); }   // :96-99
__device__ __forceinline__ float glslSmoothstep(float e0, float e1, float x) {                   // policy A-16
    float t = saturate((x - e0) / (e1 - e0));
    return t * t * (3.0f - 2.0f * t);
}
__device__ __forceinline__ float3 mix3(float3 a, float3 b, float t) { return a * (1.0f - t) + b * t; }

// ---------------------------------------------------------------------------------------
// Texture fetch: level-0 bilinear in fp32 over RGBA32F texels (policy A-12). Manual filtering keeps
// full fp32 weights (the texture unit would use 9-bit fixed point) so the oracle can be matched.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ int wrapCoord(int i, int n, int mode) {
    if (mode == HRC_WRAP_REPEAT) { i %= n; if (i < 0) i += n; return i; }
    return min(max(i, 0), n - 1);
}
__device__ __forceinline__ float4 texFetch(const DTexture& t, float u, float v) {
    float x = u * (float)t.w - 0.5f, y = v * (float)t.h - 0.5f;
    float fx0 = floorf(x), fy0 = floorf(y);
    float fx = x - fx0, fy = y - fy0;
    int x0 = wrapCoord((int)fx0, t.w, t.wrapS), x1 = wrapCoord((int)fx0 + 1, t.w, t.wrapS);
    int y0 = wrapCoord((int)fy0, t.h, t.wrapT), y1 = wrapCoord((int)fy0 + 1, t.h, t.wrapT);
    float4 p00 = __ldg(&t.texels[(size_t)y0 * t.w + x0]);
    float4 p10 = __ldg(&t.texels[(size_t)y0 * t.w + x1]);
    float4 p01 = __ldg(&t.texels[(size_t)y1 * t.w + x0]);
    float4 p11 = __ldg(&t.texels[(size_t)y1 * t.w + x1]);
    float4 r;
    r.x = (p00.x * (1.0f - fx) + p10.x * fx) * (1.0f - fy) + (p01.x * (1.0f - fx) + p11.x * fx) * fy;
    r.y = (p00.y * (1.0f - fx) + p10.y * fx) * (1.0f - fy) + (p01.y * (1.0f - fx) + p11.y * fx) * fy;
    r.z = (p00.z * (1.0f - fx) + p10.z * fx) * (1.0f - fy) + (p01.z * (1.0f - fx) + p11.z * fx) * fy;
    r.w = (p00.w * (1.0f - fx) + p10.w * fx) * (1.0f - fy) + (p01.w * (1.0f - fx) + p11.w * fx) * fy;
    return r;
}

struct Surface { float3 P, Ng; float T; float2 uv; float3 tangent, bitangent, color; };   // + interpolated varyings (vertex.rlsl:31-42), zero when absent
__device__ __forceinline__ float3 offsetOrigin(const Surface& s, float3 dir) {          // ray epsilon policy A-8
    float m = fmaxf(1.0f, fmaxf(fabsf(s.P.x), fmaxf(fabsf(s.P.y), fabsf(s.P.z))));
    float eps = 1e-4f * m;
    float sgn = dot(s.Ng, dir) >= 0.0f ? 1.0f : -1.0f;
    return s.P + s.Ng * (sgn * eps);
}
// texture2D(sampler, texCoord) of a material texture; handle -1 samples the 1x1 white dummy (Texture.h:178-192)
__device__ __forceinline__ float4 sampleMaterialTexture(const DScene& sc, int handle, float2 uv) {
    if (handle < 0) return make_float4(1.0f, 1.0f, 1.0f, 1.0f);
    return texFetch(sc.textures[handle], uv.x, uv.y);
}
// Barycentric interpolation of the optional varyings of triangle `prim` (same operation order as the normals).
__device__ __forceinline__ void loadVaryings(const DScene& sc, int prim, float u, float v, bool wantTangents, bool wantColor, Surface& s) {
    const float4* ap = sc.attr + (size_t)prim * kAttrStride;
    const float bw = (1.0f - u) - v;
    const float4 a0 = __ldg(ap), a1 = __ldg(ap + 1);
    s.uv = make_float2((a0.x * bw + a0.z * u) + a1.x * v, (a0.y * bw + a0.w * u) + a1.y * v);
    if (wantTangents) {
        s.tangent = (xyz(__ldg(ap + 2)) * bw + xyz(__ldg(ap + 3)) * u) + xyz(__ldg(ap + 4)) * v;
        s.bitangent = (xyz(__ldg(ap + 5)) * bw + xyz(__ldg(ap + 6)) * u) + xyz(__ldg(ap + 7)) * v;
    }
    if (wantColor) s.color = (xyz(__ldg(ap + 8)) * bw + xyz(__ldg(ap + 9)) * u) + xyz(__ldg(ap + 10)) * v;
}

// ---------------------------------------------------------------------------------------
// accumulator.rlsl:12-28. One thread owns a pixel inside any one kernel, so plain RMW is exclusive.
// ---------------------------------------------------------------------------------------
// accumulator.rlsl:27 (the built-in accumulate): pixel += value. Within one kernel launch a pixel receives contributions from one thread
// only (one path per pixel per pass), and launches that touch the same buffer are ordered by streams and events, so a plain
// read-modify-write is safe. RED = true issues it as one 128-bit reduction at L2 instead (red.global.add.v4.f32, sm_90+): same sum, but
// the thread does not wait for the old value. Experiment switch HR_ACCUM_RED (light shaders inside k_trace<any> only), off: the occlusion
// kernel alone gains 2-3 %, the whole pass loses (profiles/r01_v9_notes.md); in k_shade the reductions are slower than the read-modify-write.
#ifndef HR_ACCUM_RED
#define HR_ACCUM_RED 0
#endif
#ifndef HR_ACCUM_SKIP_ZERO
#define HR_ACCUM_SKIP_ZERO 1
#endif
template <bool RED>
__device__ __forceinline__ void addToPixel(float4* accum, uint32_t pixel, float x, float y, float z, float w) {
    if (RED) {
        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" :: "l"(accum + pixel), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
    } else {
        float4 p = accum[pixel];
        p.x += x; p.y += y; p.z += z; p.w += w;
        accum[pixel] = p;
    }
}
// GLSL min(x, y): "returns y if y < x, otherwise x" (a NaN in x comes back as NaN, a NaN in y is dropped); fminf would drop both.
__device__ __forceinline__ float glslMin(float x, float y) { return (y < x) ? y : x; }
template <bool RED = false>
__device__ __forceinline__ void performAccumulate(const HrcGlobals& g, float4* accum, uint32_t pixel, float3 color) {
    float3 value = f3(0.0f);
    if (g.enableAccumulatorVisualizer == 1) {
        if (g.showNANs == 1) {
            bool any = isnan(color.x) || isnan(color.y) || isnan(color.z);
            value = any ? f3(100.0f) : f3(glslMin(color.x, 1.0f), glslMin(color.y, 1.0f), glslMin(color.z, 1.0f)) * 0.1f;
        } else if (g.showInf == 1) {
            bool any = isinf(color.x) || isinf(color.y) || isinf(color.z);
            value = any ? f3(100.0f) : f3(glslMin(color.x, 1.0f), glslMin(color.y, 1.0f), glslMin(color.z, 1.0f)) * 0.1f;
        }
    } else {
        value = f3(glslMin(g.maxChannelValue, color.x), glslMin(g.maxChannelValue, color.y), glslMin(g.maxChannelValue, color.z));
    }
#if HR_ACCUM_SKIP_ZERO
    if (value.x == 0.0f && value.y == 0.0f && value.z == 0.0f) return;      // adding +0 to a non-negative sum changes nothing: skip the memory round trip
#endif
    addToPixel<RED>(accum, pixel, value.x, value.y, value.z, 0.0f);
}
__device__ __forceinline__ void accumulate4(float4* accum, uint32_t pixel, float3 rgb, float a) {
    addToPixel<false>(accum, pixel, rgb.x, rgb.y, rgb.z, a);
}

// sequence.rlsl:18-28
__device__ __forceinline__ float2 getSequenceValue(const DScene& sc, const DFrame& fr, int sequenceIndex, int sampleIndex) {
    int ws = sequenceIndex % fr.numSequences;
    int wv = sampleIndex % fr.sequenceLength;
    return __ldg(&sc.seq[ws * fr.sequenceLength + wv]);
}
// The material shaders read sequences (id + depth + k), k = 0..4, at one sample index: the two integer modulos
// (by run-time divisors) are hoisted out of the five lookups.
struct SeqCursor {
    const float2* col; int base, ns, len;
    __device__ __forceinline__ SeqCursor(const DScene& sc, const DFrame& fr, int sequenceIndex, int sampleIndex)
        : col(sc.seq + sampleIndex % fr.sequenceLength), base(sequenceIndex % fr.numSequences), ns(fr.numSequences), len(fr.sequenceLength) {}
    __device__ __forceinline__ float2 at(int k) const {
        int r = base + k;
        while (r >= ns) r -= ns;
        return __ldg(col + r * len);
    }
};

// ---------------------------------------------------------------------------------------
// Light ("miss") shaders
// ---------------------------------------------------------------------------------------
template <bool RED = false>
__device__ __forceinline__ void shadeMiss(const DScene& sc, const DFrame& fr, float4* accum, uint32_t pixel,
                                          uint32_t missKind, uint32_t missIndex, float3 dir, float3 weight, float tmax, float extraT) {
    const DLights& L = fr.lights;
    switch (missKind) {
    case kMissEnv: {                                                    // environmentLight.rlsl:19-34
        if (!L.envEnabled) return;
        float theta = atan2f(dir.x, -dir.z) + L.envTheta;
        if (theta > K_TWO_PI) theta = theta - K_TWO_PI;
        float phi = atan2f(dir.y, sqrtf(dir.x * dir.x + dir.z * dir.z));
        float u = (theta / K_TWO_PI) + 0.5f;
        float v = (-phi * K_ONE_OVER_PI) + 0.5f;
        float3 sample = f3(0.0f);
        if (L.envTex >= 0) sample = xyz(texFetch(sc.textures[L.envTex], u, 1.0f - v));
        sample = sample * L.envExposure;
        performAccumulate<RED>(fr.g, accum, pixel, weight * sample);
        break; }
    case kMissDir:                                                      // directionalLight.rlsl:20-26
        performAccumulate<RED>(fr.g, accum, pixel, weight * f3(L.dirColor[missIndex]));
        break;
    case kMissPoint: {                                                  // pointLight.rlsl:20-29 (rl_IntersectionT = maxT, policy A-7)
        float attenuation = 1.0f / square(tmax + extraT);
        performAccumulate<RED>(fr.g, accum, pixel, weight * f3(L.pointColor[missIndex]) * attenuation);
        break; }
    case kMissSpot: {                                                   // spotLight.rlsl:20-36
        float rayAngle = dot(-dir, f3(L.spotDirection[missIndex]));
        if (rayAngle >= 0.0f) {
            float attenuation = 1.0f / square(tmax + extraT);
            float3 result = weight * f3(L.spotColor[missIndex]) * attenuation;
            result = result * (1.0f - glslSmoothstep(L.spotAngles[missIndex][0], L.spotAngles[missIndex][1], rayAngle));
            performAccumulate<RED>(fr.g, accum, pixel, result);
        }
        break; }
    default: break;
    }
}

// ---------------------------------------------------------------------------------------
// Traversal of the 8-wide compressed BVH
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t byteOf(uint32_t x, int i) { return (x >> (i * 8)) & 0xffu; }

// Möller–Trumbore as ONE fixed sequence of IEEE operations that the CPU oracle repeats (std::fmaf where an FMA stands here), so
// hit/miss decisions and (t,u,v) agree bit for bit (policy A-9):
//   cross(a,b).x = fma(a.y, b.z, -(a.z*b.y));  dot(a,b) = fma(a.z, b.z, fma(a.y, b.y, a.x*b.x));
//   the barycentric tests run on the undivided numerators against |det| (sign of det folded in by XOR on the sign bit) and the one
//   division 1/|det| is only reached by candidates that passed them — most leaf triangles are rejected before it.
__device__ __forceinline__ float dotFma(float ax, float ay, float az, float bx, float by, float bz) {
    return __fmaf_rn(az, bz, __fmaf_rn(ay, by, __fmul_rn(ax, bx)));
}
__device__ __forceinline__ bool triangleTest(float4 a, float4 b, float4 c, float3 o, float3 d, float tmin, float tmax,
                                             float& t, float& u, float& v, bool& back) {
    // pv = cross(d, e2)
    const float pvx = __fmaf_rn(d.y, c.z, -__fmul_rn(d.z, c.y));
    const float pvy = __fmaf_rn(d.z, c.x, -__fmul_rn(d.x, c.z));
    const float pvz = __fmaf_rn(d.x, c.y, -__fmul_rn(d.y, c.x));
    const float det = dotFma(b.x, b.y, b.z, pvx, pvy, pvz);
    if (det == 0.0f) return false;
    const uint32_t sgn = __float_as_uint(det) & 0x80000000u;
    const float adet = fabsf(det);
    const float tvx = __fsub_rn(o.x, a.x), tvy = __fsub_rn(o.y, a.y), tvz = __fsub_rn(o.z, a.z);
    const float U = __uint_as_float(__float_as_uint(dotFma(tvx, tvy, tvz, pvx, pvy, pvz)) ^ sgn);
    if (!(U >= 0.0f && U <= adet)) return false;
    // qv = cross(tv, e1)
    const float qvx = __fmaf_rn(tvy, b.z, -__fmul_rn(tvz, b.y));
    const float qvy = __fmaf_rn(tvz, b.x, -__fmul_rn(tvx, b.z));
    const float qvz = __fmaf_rn(tvx, b.y, -__fmul_rn(tvy, b.x));
    const float V = __uint_as_float(__float_as_uint(dotFma(d.x, d.y, d.z, qvx, qvy, qvz)) ^ sgn);
    if (!(V >= 0.0f && __fadd_rn(U, V) <= adet)) return false;
    const float T = __uint_as_float(__float_as_uint(dotFma(c.x, c.y, c.z, qvx, qvy, qvz)) ^ sgn);
    const float inv = __fdiv_rn(1.0f, adet);
    t = __fmul_rn(T, inv);
    if (!(t > tmin && t < tmax)) return false;
    u = __fmul_rn(U, inv); v = __fmul_rn(V, inv);
    back = sgn != 0u;
    return true;
}

struct TraceCounters { uint32_t nodes, tris, overflow; };

// Byte J of q spliced into mantissa bits 8..15 of 1.0f with one PRMT: the float 1 + q * 2^-15. The child-plane distance
// (origin + q*cell - o) / d is then ONE FMA, f * A + B with A = cell/d * 2^15 and B = (origin - o)/d - A: no I2F (quarter
// rate) and no separate subtract. B absorbs a rounding error of at most 2^-9 cell, which the conservative bias on B
// (nodeSetup) covers.
template <int J>
__device__ __forceinline__ float byteToUnitFloat(uint32_t q) {
    return __uint_as_float(__byte_perm(q, 0x3F800000u, 0x7604u + (J << 4)));
}

// Traversal stack of a lane. With HR_SMEM_STACK = N > 0 the N entries nearest the root of the stack live in shared memory
// ([entry][thread], conflict-free: a lane always hits its own bank pair) and only deeper entries go to thread-local memory.
#ifndef HR_SMEM_STACK
#define HR_SMEM_STACK 0
#endif
struct TraceStack {
    uint2 deep[kTraceStack - HR_SMEM_STACK];
#if HR_SMEM_STACK > 0
    uint2* sh;
    __device__ __forceinline__ void bind(uint2* blockShared) { sh = blockShared + threadIdx.x; }
    __device__ __forceinline__ void put(int i, uint2 v) { if (i < HR_SMEM_STACK) sh[i * 128] = v; else deep[i - HR_SMEM_STACK] = v; }
    __device__ __forceinline__ uint2 get(int i) const { return i < HR_SMEM_STACK ? sh[i * 128] : deep[i - HR_SMEM_STACK]; }
#define HR_DECLARE_TRACE_STACK(NAME) __shared__ uint2 NAME##_sh[HR_SMEM_STACK * 128]; TraceStack NAME; NAME.bind(NAME##_sh)
#else
    __device__ __forceinline__ void put(int i, uint2 v) { deep[i] = v; }
    __device__ __forceinline__ uint2 get(int i) const { return deep[i]; }
#define HR_DECLARE_TRACE_STACK(NAME) TraceStack NAME
#endif
};

#ifndef HR_TOP_SMEM
#define HR_TOP_SMEM 0            // nodes of the top BVH levels kept in shared memory by k_trace (1 + 8 = 9, 1 + 8 + 64 = 73); measured neutral, off (profiles/r02_notes.md)
#endif
#ifndef HR_TRI_PREFETCH
#define HR_TRI_PREFETCH 0        // 128-byte lines of a node's leaf triangles prefetched at the start of its node step (experiment, see profiles/r02_notes.md)
#endif
#ifndef HR_TMA_REFILL
#define HR_TMA_REFILL 0          // k_trace refills a warp's idle lanes with ONE bulk copy (cp.async.bulk, TMA 1-D) of the contiguous ray records instead of per-lane loads; measured, off
#endif
// Per-lane traversal state. A lane keeps it across the warp-synchronous phases of k_trace (node step, triangle rounds,
// refill), so a warp tops itself up with fresh rays while slow rays are still in flight.
struct Traversal {
    float3 o, d, idir;
    float tmin, tmaxRay;
    uint32_t octInv;
    float bestT, bestU, bestV; int bestPrim;      // bestPrim carries HRC_HIT_BACKFACE and, in bits 28..29, the material class
    bool found;
    uint2 group;       // x: first child node index; y: bits 24..31 pending internal children (octant order), bits 0..7 imask
    uint2 tri;         // leaf triangles hit by the last node step (x: first triangle, y: bit mask)
    int sp;

    __device__ __forceinline__ void begin(float3 org, float3 dir, float tmin_, float tmax_) {
        o = org; d = dir; tmin = tmin_; tmaxRay = tmax_;
        // guard the slab arithmetic against zero direction components (box tests only)
        const float tiny = 1e-20f;
        const float3 dd = f3(fabsf(d.x) < tiny ? copysignf(tiny, d.x) : d.x,
                             fabsf(d.y) < tiny ? copysignf(tiny, d.y) : d.y,
                             fabsf(d.z) < tiny ? copysignf(tiny, d.z) : d.z);
        idir = f3(1.0f / dd.x, 1.0f / dd.y, 1.0f / dd.z);
        // slot bit set = positive side; a ray going positive must visit negative-side children first
        octInv = (dd.x < 0.0f ? 0u : 4u) | (dd.y < 0.0f ? 0u : 2u) | (dd.z < 0.0f ? 0u : 1u);
        bestT = tmax_; bestPrim = 0x0fffffff; bestU = 0.f; bestV = 0.f;
        found = false;
        sp = 0;
        group = make_uint2(0u, 0x80000000u);     // root: "child 7 of a virtual parent whose childBase is 0"
        tri = make_uint2(0u, 0u);
    }
    __device__ __forceinline__ bool hasNodeWork() const { return (group.y & 0xff000000u) != 0u; }

    // Pops the nearest pending child of `group`, tests its 8 children, leaves the hit internal children in `group` and
    // files the hit leaf triangles as a pending triangle group.
    template <bool STATS, bool ANY>
    __device__ __forceinline__ void nodeStep(const float4* __restrict__ nodes, TraceStack& stack, TraceCounters& tc, const float4* topNodes = nullptr, const float4* trisForPrefetch = nullptr) {
        const uint32_t hits = group.y;
        const uint32_t bit = 31u - __clz(hits);                  // highest set bit = nearest remaining child
        group.y &= ~(1u << bit);
        if (group.y & 0xff000000u) {
            if (sp < kTraceStack) stack.put(sp++, group); else if (STATS) tc.overflow++;
        }
        // Visiting order of a node's children: front to back for closest-hit rays. Occlusion rays only need ANY occluder and find one sooner BACK to front
        // (they start on a surface and head for a light: what lies near the origin rarely blocks; measured on the CPU walk of the same tree,
        // tools/bvh_eval.cpp: -5 % node steps, -4 % triangle tests per occlusion ray), so they use the mirrored octant. Plane selection below keeps octInv.
        const uint32_t ordInv = ANY ? (octInv ^ 7u) : octInv;
        const uint32_t slot = (bit - 24u) ^ (ordInv & 7u);
        const uint32_t rel = __popc(hits & ~(0xffffffffu << slot));   // internal children stored before this slot
        const uint32_t nodeIdx = group.x + rel;

        float4 n0, n1, n2, n3, n4;
#if HR_TOP_SMEM > 0
        if (topNodes && nodeIdx < HR_TOP_SMEM) {          // experiment (north_star: "hot top-level nodes staged in shared memory"): the first levels of the breadth-first node array
            const float4* sp = topNodes + nodeIdx * 5;
            n0 = sp[0]; n1 = sp[1]; n2 = sp[2]; n3 = sp[3]; n4 = sp[4];
        } else
#endif
        {
            const float4* np = nodes + (size_t)nodeIdx * 5;
            n0 = __ldg(np + 0); n1 = __ldg(np + 1); n2 = __ldg(np + 2); n3 = __ldg(np + 3); n4 = __ldg(np + 4);
        }
        if (STATS) tc.nodes++;
#if HR_TRI_PREFETCH > 0
        // experiment: the node's leaf triangles sit contiguously at its triBase; pull their first lines towards L1 while the slab tests run
        // (the dependent triangle fetch is 41 % of the closest-hit kernel's stall samples)
        if (trisForPrefetch) {
            const char* tp = reinterpret_cast<const char*>(trisForPrefetch + (size_t)__float_as_uint(n1.y) * 3);
#pragma unroll
            for (int k = 0; k < HR_TRI_PREFETCH; ++k) asm volatile("prefetch.global.L1 [%0];" :: "l"(tp + k * 128));
        }
#endif

        const uint32_t eimask = __float_as_uint(n0.w);
        // Child plane distance (origin + q*cell - o) / d as ONE FMA on the widened magic number (1024 + q):
        // A = cell / d (cell = 2^(e-127); e == 0 marks a flat axis with cell size 0), B = (origin - o)/d - 1024 A.
        const uint32_t ex = byteOf(eimask, 0), ey = byteOf(eimask, 1), ez = byteOf(eimask, 2);
        const float Ax = __uint_as_float(ex << 23) * idir.x;
        const float Ay = __uint_as_float(ey << 23) * idir.y;
        const float Az = __uint_as_float(ez << 23) * idir.z;
        const float Bx = fmaf(-1024.0f, Ax, (n0.x - o.x) * idir.x), By = fmaf(-1024.0f, Ay, (n0.y - o.y) * idir.y), Bz = fmaf(-1024.0f, Az, (n0.z - o.z) * idir.z);
        // conservative bias: entry planes a little earlier, exit planes a little later (2^-9 cell >> the 2^-14 cell rounding of B)
        const float kBias = 1.0f / 512.0f;
        const float ex_ = fabsf(Ax) * kBias, ey_ = fabsf(Ay) * kBias, ez_ = fabsf(Az) * kBias;
        const float Bnx = Bx - ex_, Bny = By - ey_, Bnz = Bz - ez_;
        const float Bfx = Bx + ex_, Bfy = By + ey_, Bfz = Bz + ez_;
        const bool negX = !(octInv & 4u), negY = !(octInv & 2u), negZ = !(octInv & 1u);
#if HR_TFAR_PRESCALE
        const float tfar = (ANY ? tmaxRay : bestT) * 1.000001f;   // experiment: the tie guard once per node step instead of once per child
#else
        const float tfar = ANY ? tmaxRay : bestT;   // occlusion rays cull against maxT only: bestT then tracks the nearest alpha-masked hit, which must not hide occluders
#endif
        const uint32_t octInv4 = ordInv * 0x01010101u;

        uint32_t hitMask = 0;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            const uint32_t meta4 = __float_as_uint(half == 0 ? n1.z : n1.w);
            const uint32_t isInner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
            const uint32_t innerMask4 = (isInner4 >> 4) * 0xffu;       // 0xff in every byte whose slot holds an internal child
            const uint32_t bitIndex4 = (meta4 ^ (octInv4 & innerMask4)) & 0x1f1f1f1fu;
            const uint32_t childBits4 = (meta4 >> 5) & 0x07070707u;
            const uint32_t qlox = __float_as_uint(half == 0 ? n2.x : n2.y), qloy = __float_as_uint(half == 0 ? n2.z : n2.w);
            const uint32_t qloz = __float_as_uint(half == 0 ? n3.x : n3.y), qhix = __float_as_uint(half == 0 ? n3.z : n3.w);
            const uint32_t qhiy = __float_as_uint(half == 0 ? n4.x : n4.y), qhiz = __float_as_uint(half == 0 ? n4.z : n4.w);
            const uint32_t nx = negX ? qhix : qlox, fx = negX ? qlox : qhix;
            const uint32_t ny = negY ? qhiy : qloy, fy = negY ? qloy : qhiy;
            const uint32_t nz = negZ ? qhiz : qloz, fz = negZ ? qloz : qhiz;
#if HR_TFAR_PRESCALE
#define HR_TIE_GUARD(X) (X)
#else
#define HR_TIE_GUARD(X) ((X) * 1.000001f)
#endif
#define HR_PLANES(Q, SEL, A, B, LO, HI)                                                                                \
            { const uint32_t h_ = __byte_perm(Q, 0x64646464u, SEL);   /* two fp16 magic numbers 1024 + q in one PRMT */  \
              const __half2 hv_ = *reinterpret_cast<const __half2*>(&h_);                                               \
              LO = fmaf(__low2float(hv_), A, B); HI = fmaf(__high2float(hv_), A, B); }   /* widening on the FMA pipe */
#define HR_HIT(J, TNX, TNY, TNZ, TFX, TFY, TFZ)                                                                        \
            {                                                                                                          \
                const float tn = fmaxf(fmaxf(TNX, TNY), fmaxf(TNZ, tmin));                                              \
                const float tf = fminf(fminf(TFX, TFY), fminf(TFZ, tfar));                                              \
                if (tn <= HR_TIE_GUARD(tf)) /* conservative: rounding must never cull an exact tie */                   \
                    hitMask |= byteOf(childBits4, J) << byteOf(bitIndex4, J);                                           \
            }
#define HR_PAIR(J0, J1, SEL)                                                                                           \
            {                                                                                                          \
                float anx, bnx, any_, bny, anz, bnz, afx, bfx, afy, bfy, afz, bfz;                                      \
                HR_PLANES(nx, SEL, Ax, Bnx, anx, bnx) HR_PLANES(ny, SEL, Ay, Bny, any_, bny) HR_PLANES(nz, SEL, Az, Bnz, anz, bnz) \
                HR_PLANES(fx, SEL, Ax, Bfx, afx, bfx) HR_PLANES(fy, SEL, Ay, Bfy, afy, bfy) HR_PLANES(fz, SEL, Az, Bfz, afz, bfz) \
                HR_HIT(J0, anx, any_, anz, afx, afy, afz) HR_HIT(J1, bnx, bny, bnz, bfx, bfy, bfz)                       \
            }
            HR_PAIR(0, 1, 0x4140) HR_PAIR(2, 3, 0x4342)
#undef HR_PAIR
#undef HR_HIT
#undef HR_PLANES
#undef HR_TIE_GUARD
        }
        group.x = __float_as_uint(n1.x);
        group.y = (hitMask & 0xff000000u) | (eimask >> 24);
        const uint32_t triBits = hitMask & 0x00ffffffu;
        tri = make_uint2(__float_as_uint(n1.y), triBits);
    }

    // Takes the next node group from the stack when the current one is used up.
    __device__ __forceinline__ void popIfEmpty(const TraceStack& stack) {
        if (!hasNodeWork() && sp > 0) group = stack.get(--sp);
    }

    // Tests ONE pending triangle. Returns true when an any-hit ray is finished.
    template <bool ANY, bool STATS, bool MASKED>
    __device__ __forceinline__ bool triStep(const float4* __restrict__ tris, TraceCounters& tc) {
        const uint32_t ti = 31u - __clz(tri.y);
        tri.y &= ~(1u << ti);
        const float4* tp = tris + (size_t)(tri.x + ti) * 3;
        const float4 a = __ldg(tp), b = __ldg(tp + 1), c = __ldg(tp + 2);
        if (STATS) tc.tris++;
        float t, u, v; bool back;
        if (triangleTest(a, b, c, o, d, tmin, tmaxRay, t, u, v, back)) {
            if (ANY && !(MASKED && (__float_as_uint(b.w) & kTriNonOccluder))) { found = true; return true; }
            const int prim = __float_as_int(a.w);
            if (t < bestT || (t == bestT && prim < (bestPrim & 0x0fffffff))) {
                bestT = t; bestU = u; bestV = v; found = true;
                bestPrim = prim | (back ? HRC_HIT_BACKFACE : 0) | ((__float_as_int(b.w) & 3) << 28);
            }
        }
        return false;
    }
    __device__ __forceinline__ bool finished() const { return !hasNodeWork() && tri.y == 0u && sp == 0; }
};

// Plain per-thread traversal (the hrcTraceRays parity hook): same steps, no warp scheduling.
// Returns true when an any-hit ray stopped at an occluder.
template <bool ANY, bool MASKED = false>
__device__ __forceinline__ bool traverseSimple(const float4* __restrict__ nodes, const float4* __restrict__ tris, Traversal& tr) {
    HR_DECLARE_TRACE_STACK(stack);
    TraceCounters tc{0u, 0u, 0u};
    for (;;) {
        if (tr.hasNodeWork()) tr.nodeStep<false, ANY>(nodes, stack, tc);
        while (tr.tri.y) if (tr.triStep<ANY, false, MASKED>(tris, tc)) return true;
        tr.popIfEmpty(stack);
        if (tr.finished()) return false;
    }
}

// ---------------------------------------------------------------------------------------
// Kernels
// ---------------------------------------------------------------------------------------
__global__ void k_clear_ctrl(IterCtrl* ctrl) {
    int i = threadIdx.x;
    if (i < kCtrlEntries) ctrl[i] = IterCtrl{};
}

// perspective.rlsl:39-93 — one thread per pixel of this context's screen bands.
__global__ void __launch_bounds__(256) k_raygen(const DScene sc, const int slot, RayQueue q, float4* accum, IterCtrl* ctrl) {
    const DFrame& fr = g_frames[slot];
    const uint32_t total = (uint32_t)fr.w * (uint32_t)fr.localRows;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t == 0 && !fr.interactiveMode) ctrl[0].extCount = total;
    const int W = fr.w, H = fr.h;
    bool sampled = t < total;
    uint32_t pixel = 0; float fx = 0.f, fy = 0.f;
    if (sampled) {
        int lr = t / W, x = t - lr * W;
        if (fr.rayTileW > 0) {
            // ray t of the pass = pixel (t % 32) of tile (t / 32): the 32 rays a traversal warp pulls together start through one compact pixel tile
            // instead of a 32 x 1 strip (narrower frustum: the warp's lanes walk the same nodes for longer). Which queue slot a pixel's ray sits in
            // has no influence on the pixel's value (one path per pixel per pass).
            const uint32_t tile = t >> 5, in = t & 31u, tilesX = (uint32_t)W / (uint32_t)fr.rayTileW;
            const uint32_t ty = tile / tilesX, tx = tile - ty * tilesX;
            x = (int)(tx * fr.rayTileW + in % (uint32_t)fr.rayTileW);
            lr = (int)(ty * fr.rayTileH + in / (uint32_t)fr.rayTileW);
        }
        int y = lr;
        if (fr.bandCount > 1) { const int lb = lr / fr.bandHeight; y = (lb * fr.bandCount + fr.bandIndex) * fr.bandHeight + (lr - lb * fr.bandHeight); }
        pixel = (uint32_t)y * W + x;
        fx = (float)x + 0.5f; fy = (float)y + 0.5f;                       // rl_FrameCoord (policy A-1)
        if (fr.interactiveMode) {                                          // :42-57: one pixel of every block per pass
            const int bx = x / fr.blockW, by = y / fr.blockH;
            const float2 rnd = __ldg(&sc.blockRand[(size_t)by * fr.blocksX + bx]);   // random(blockIndex), random(blockIndex.yx) (host, quirk B-Q2)
            const float su = __fadd_rn(__fmul_rn(1.0f / (float)fr.blockW, (float)fr.blockSampleX), rnd.x);
            const float sv = __fadd_rn(__fmul_rn(1.0f / (float)fr.blockH, (float)fr.blockSampleY), rnd.y);
            int ts = (int)floorf(__fmul_rn(su, (float)fr.blockW)) % fr.blockW, tt = (int)floorf(__fmul_rn(sv, (float)fr.blockH)) % fr.blockH;   // NEAREST + REPEAT
            if (ts < 0) ts += fr.blockW;
            if (tt < 0) tt += fr.blockH;
            const float* texel = sc.blockCoords + 3 * (tt * fr.blockW + ts);
            sampled = (x % fr.blockW) == (int)__ldg(texel) && (y % fr.blockH) == (int)__ldg(texel + 1);
        }
    }
    uint32_t i = t;
    if (fr.interactiveMode) {                                              // compaction: one atomic per warp
        const uint32_t m = __ballot_sync(0xffffffffu, sampled);
        const uint32_t lane = threadIdx.x & 31u;
        uint32_t base = 0;
        if (m && lane == (uint32_t)(__ffs(m) - 1)) base = atomicAdd(&ctrl[0].extCount, (uint32_t)__popc(m));
        base = __shfl_sync(0xffffffffu, base, m ? __ffs(m) - 1 : 0);
        i = base + __popc(m & ((1u << lane) - 1u));
    }
    if (!sampled) return;

    float4 px = accum[pixel]; px.w += 1.0f; accum[pixel] = px;            // accumulate(vec4(0,0,0,1)), :60

    const int sequenceID = sc.sequenceId[pixel];                           // :62 via the host table (quirk B-Q2)
    const uint32_t oi = (uint32_t)(int)__fadd_rn(__fmul_rn(fy, (float)H), fx);   // :64 (quirk B-Q1)
    const float rnd = __ldg(&sc.seqOffsetX[min(oi, fr.offsetCount - 1)]);
    const int sequenceIndex = (int)floorf(rnd * fr.maxSampleIndex);       // :65
    const float2 so = getSequenceValue(sc, fr, sequenceID, fr.g.sampleIndex + sequenceIndex);   // :67
    const float spx = (fx - 0.5f) + so.x, spy = (fy - 0.5f) + so.y;       // :68
    const float u = spx / (float)W, v = spy / (float)H;                   // :69
    const float cx = (2.0f * u - 1.0f) * fr.aspectRatio * fr.fovTan;      // :72
    const float cy = (1.0f - 2.0f * v) * fr.fovTan * -1.0f;               // :73
    const float3 dirCam = normalize(f3(cx, cy, -1.0f));                   // :74
    const float3 focal = fr.focusDistance * dirCam;                       // :77
    const uint32_t tableSize = (uint32_t)fr.numSequences * (uint32_t)fr.sequenceLength;
    const uint32_t ai = ((uint32_t)sequenceID * (uint32_t)fr.sequenceLength + (uint32_t)fr.g.sampleIndex) % tableSize;   // :78 (quirk B-Q3)
    float2 ap = __ldg(&sc.aperture[ai]);
    ap.x = ((ap.x * 2.0f) - 1.0f) * fr.apertureRadius;                    // :80
    ap.y = ((ap.y * 2.0f) - 1.0f) * fr.apertureRadius;
    const float3 org = f3(ap.x, ap.y, 0.0f);
    const float3 dir = focal - org;
    const float* M = fr.view;
    const float3 wo = f3((M[0] * org.x + M[4] * org.y) + (M[8] * org.z + M[12]),
                         (M[1] * org.x + M[5] * org.y) + (M[9] * org.z + M[13]),
                         (M[2] * org.x + M[6] * org.y) + (M[10] * org.z + M[14]));   // :85
    const float3 wd = normalize(f3((M[0] * dir.x + M[4] * dir.y) + M[8] * dir.z,
                                   (M[1] * dir.x + M[5] * dir.y) + M[9] * dir.z,
                                   (M[2] * dir.x + M[6] * dir.y) + M[10] * dir.z)); // :86 (normalised on emit, A-11)
    q.od[2 * (size_t)i] = make_float4(wo.x, wo.y, wo.z, __int_as_float(0x7f800000));   // tmax = +inf
    q.od[2 * (size_t)i + 1] = make_float4(wd.x, wd.y, wd.z, 0.0f);                     // extraT = 0
    q.wm[2 * (size_t)i] = make_float4(1.0f, 1.0f, 1.0f, __uint_as_float(pixel));
    const uint32_t missKind = fr.lights.envEnabled ? kMissEnv : kMissNone;   // defaultPrimitive = env light, :87
    q.wm[2 * (size_t)i + 1] = make_float4(__uint_as_float(0u | (missKind << 8) | ((uint32_t)sequenceID << 16)), __int_as_float(sequenceIndex), 0.0f, 0.0f);
}

// Persistent traversal kernel. Every lane owns one ray and runs the classic while-while loop (one 8-wide node step, then
// that node's leaf triangles at once). When kRefillIdle lanes of a warp have finished, the warp pauses at a
// warp-synchronous point: finished rays are appended to the per-material-class index lists (one atomic per warp per
// class: the compaction that lets each shade kernel run one material) and the idle lanes pull fresh rays from the queue
// (one atomic per warp) while the other lanes keep their traversal state.
// An occlusion ray found no occluder on its way, but an alpha-masked (non-occluder) surface, tr.bestPrim: physicallyBased.rlsl:70-90
// (policy A-5). Returns true when the surface lets the ray through (tr then holds the continuation ray behind it, extraT the
// distance travelled so far, :80-84); false when the ray is shadowed (:88-90).
__device__ __forceinline__ bool maskedSurfaceStep(const DScene& sc, Traversal& tr, float& extraT, int& maskedPasses) {
    const int prim = tr.bestPrim & 0x0fffffff;
    const float4* sp = sc.shade + (size_t)prim * 4;
    const int mat = __float_as_int(__ldg(sp).w);
    const uint32_t tf = sc.materials[mat].typeFlags;
    float alpha = 1.0f;
    if ((tf & 0xfu) == HRC_MATERIAL_PBR && ((tf >> 4) & HRC_MAT_HAS_BASE_COLOR_TEXTURE) && sc.attr) {
        Surface sv; loadVaryings(sc, prim, tr.bestU, tr.bestV, false, false, sv);
        alpha = sampleMaterialTexture(sc, sc.materials[mat].baseColorTex, sv.uv).w;
    }
    if ((tf & 0xfu) == HRC_MATERIAL_PBR && ((tf >> 4) & HRC_MAT_ALPHA_MASK) && alpha < 1.0f && maskedPasses < kMaxMaskedSurfaces) {
        Surface sv; sv.P = tr.o + tr.d * tr.bestT; sv.Ng = xyz(__ldg(sp + 3));
        const float t = tr.bestT;
        extraT += t; ++maskedPasses;                                     // :84
        tr.begin(offsetOrigin(sv, tr.d), tr.d, 0.0f, tr.tmaxRay - t);    // :80
        return true;
    }
    return false;
}

// Tried and measured slower on B200 (profiles/r01_v3_notes.md): parking leaf triangles in a per-lane ring and testing
// them in warp-wide rounds (triangle-test lane utilisation doubled, node-step utilisation and scheduling overhead ate the
// gain), and "one node step or one triangle per trip".
constexpr uint32_t kFull = 0xffffffffu;
__device__ __forceinline__ void prefetchL2(const void* p) { asm volatile("prefetch.global.L2 [%0];" :: "l"(p)); }
#ifndef HR_RAY_PREFETCH_AHEAD
#define HR_RAY_PREFETCH_AHEAD 0
#endif

#ifndef HR_TRACE_MIN_BLOCKS
#define HR_TRACE_MIN_BLOCKS 8
#endif
template <bool ANY, bool STATS, bool MASKED>
__global__ void __launch_bounds__(128, HR_TRACE_MIN_BLOCKS) k_trace(const DScene sc, const int slot, const PassBuffers pb, int iter, int kRefillIdle) {
    const DFrame& fr = g_frames[slot];
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t below = (1u << lane) - 1u;
    IterCtrl* ctrl = pb.ctrl + iter;
    const uint32_t count = ANY ? ctrl->shadowCount : ctrl->extCount;
    uint32_t* head = ANY ? &ctrl->shadowHead : &ctrl->extHead;
    const float4* __restrict__ qod = (iter & 1) ? pb.ext[1].od : pb.ext[0].od;   // no dynamic indexing of kernel parameters (it forces a local copy)
    const float4* __restrict__ nodes = sc.nodes;
    const float4* __restrict__ tris = sc.tris;
    // coherence sort: sorted position -> ray index (extension rays of iterations >= 1 only; see device_types.h)
    const uint32_t* __restrict__ order = (!ANY && pb.sortCtrl && pb.sortCtrl[iter].sorted) ? pb.sortOrder : nullptr;
    TraceCounters tc{0u, 0u, 0u};
    Traversal tr;
    HR_DECLARE_TRACE_STACK(stack);
#if HR_TOP_SMEM > 0
    __shared__ float4 topNodes[HR_TOP_SMEM * 5];
    for (int k = threadIdx.x; k < HR_TOP_SMEM * 5; k += blockDim.x) topNodes[k] = (k / 5 < (int)sc.nodeCount) ? __ldg(nodes + k) : make_float4(0.f, 0.f, 0.f, 0.f);
    __syncthreads();
#else
    const float4* topNodes = nullptr;
#endif
#if HR_TMA_REFILL
    __shared__ __align__(128) float4 refillBuf[4][64];        // per warp: 32 ray records of 32 B
    __shared__ __align__(8) unsigned long long refillBar[4];
    const uint32_t warpInBlock = threadIdx.x >> 5;
    uint32_t refillPhase = 0;
    if (lane == 0) {
        const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&refillBar[warpInBlock]);
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(bar));
    }
    __syncwarp();
#endif
    bool active = false, exhausted = false;
    uint32_t rayIdx = 0, doneIdx = 0;
    int doneClass = -1;                      // closest-hit only: class of a finished ray not yet appended to its list
    float extraT = 0.0f; int maskedPasses = 0;   // MASKED only: distance already travelled through alpha-masked surfaces

    for (;;) {
        const uint32_t idle = __ballot_sync(kFull, !active);
        if (idle == kFull || (!exhausted && __popc(idle) >= kRefillIdle)) {
            if (!ANY) {                                                   // retire: append finished rays to the class lists
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const uint32_t m = __ballot_sync(kFull, doneClass == c);
                    if (m) {
                        const int leader = __ffs(m) - 1;
                        uint32_t base = 0;
                        if ((int)lane == leader) base = atomicAdd(&ctrl->classCount[c], (uint32_t)__popc(m));
                        base = __shfl_sync(kFull, base, leader);
                        if (doneClass == c) pb.classIdx[c][base + __popc(m & below)] = doneIdx;
                    }
                }
                doneClass = -1;
            }
            if (ANY && pb.litIdx) {                                       // retire: rays that reached their light go to k_shade_lights (the light shaders at 32 lanes instead of ~5)
                const uint32_t m = __ballot_sync(kFull, doneClass == 0);
                if (m) {
                    const int leader = __ffs(m) - 1;
                    uint32_t base = 0;
                    if ((int)lane == leader) base = atomicAdd(&ctrl->litCount, (uint32_t)__popc(m));
                    base = __shfl_sync(kFull, base, leader);
                    if (doneClass == 0) pb.litIdx[base + __popc(m & below)] = doneIdx;
                }
                doneClass = -1;
            }
            if (!exhausted) {                                             // refill
                const uint32_t n = __popc(idle);
                const int leader = __ffs(idle) - 1;
                uint32_t base = 0;
                if ((int)lane == leader) base = atomicAdd(head, n);
                base = __shfl_sync(kFull, base, leader);
#if HR_TMA_REFILL
                if (!ANY && !order) {       // one bulk copy of the warp's contiguous records [base, base + n) into shared memory, completion on the warp's mbarrier
                    const uint32_t take = base < count ? min(n, count - base) : 0u;
                    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&refillBar[warpInBlock]);
                    if (take) {
                        if ((int)lane == leader) {
                            const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&refillBuf[warpInBlock][0]);
                            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(take * 32u) : "memory");
                            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                         :: "r"(dst), "l"(qod + 2 * (size_t)base), "r"(take * 32u), "r"(bar) : "memory");
                        }
                        uint32_t done = 0;
                        while (!done) asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(bar), "r"(refillPhase) : "memory");
                        refillPhase ^= 1u;
                    }
                }
#endif
                if (!active) {
                    const uint32_t slotInBatch = __popc(idle & below);
                    uint32_t i = base + slotInBatch;
                    if (i < count) {
                        if (!ANY && order) i = __ldcs(order + i);
#if HR_TMA_REFILL
                        const bool bulk = !ANY && !order;
                        const float4 ro = bulk ? refillBuf[warpInBlock][2 * slotInBatch] : __ldcs(ANY ? pb.shadow.o + i : qod + 2 * (size_t)i);
                        const float4 rd = bulk ? refillBuf[warpInBlock][2 * slotInBatch + 1] : __ldcs(ANY ? pb.shadow.d + i : qod + 2 * (size_t)i + 1);
#else
                        const float4 ro = __ldcs(ANY ? pb.shadow.o + i : qod + 2 * (size_t)i), rd = __ldcs(ANY ? pb.shadow.d + i : qod + 2 * (size_t)i + 1);   // streamed: keep the BVH in L2
#endif
#if HR_RAY_PREFETCH_AHEAD > 0
                        // the queue is consumed in index order by all warps together: pull the ray HR_RAY_PREFETCH_AHEAD places further on into L2
                        if (i + HR_RAY_PREFETCH_AHEAD < count) {
                            if (ANY) { prefetchL2(pb.shadow.o + i + HR_RAY_PREFETCH_AHEAD); prefetchL2(pb.shadow.d + i + HR_RAY_PREFETCH_AHEAD); }
                            else prefetchL2(qod + 2 * (size_t)(i + HR_RAY_PREFETCH_AHEAD));
                        }
#endif
                        rayIdx = i;
                        tr.begin(xyz(ro), xyz(rd), 0.0f, ro.w);
                        if (MASKED) { extraT = rd.w; maskedPasses = 0; }
                        active = true;
                    }
                }
                if (base + n >= count) exhausted = true;
            }
            if (!__any_sync(kFull, active)) break;
        }

        while (active) {
            if (tr.hasNodeWork()) tr.nodeStep<STATS, ANY>(nodes, stack, tc, topNodes, tris);
            bool occluded = false;
            while (tr.tri.y) if (tr.triStep<ANY, STATS, MASKED>(tris, tc)) { occluded = true; break; }
            tr.popIfEmpty(stack);
            if (occluded) { active = false; break; }                       // any-hit: nothing to shade
            if (tr.finished()) {
                if (ANY && MASKED && (tr.bestPrim & 0x0fffffff) != 0x0fffffff) {
                    if (maskedSurfaceStep(sc, tr, extraT, maskedPasses)) {               // passes through: the ray goes on behind the surface
                        atomicAdd(&ctrl->maskedContinuations, 1u);
                        if (pb.rayCounts) atomicAdd(pb.rayCounts + (__float_as_uint(pb.shadow.w[rayIdx].w) & 0x00ffffffu), 1u);
                        continue;
                    }
                    active = false;                                                      // :88-90 shadowed
                    break;
                }
                active = false;
                if (ANY && pb.litIdx) {          // the ray reached its light: file it for k_shade_lights
                    if (MASKED && maskedPasses > 0) {                       // what the light shader needs of the continuation ray goes back into the (now dead) queue record
                        pb.shadow.o[rayIdx].w = tr.tmaxRay; pb.shadow.d[rayIdx].w = extraT;
                    }
                    doneIdx = rayIdx; doneClass = 0;
                } else if (ANY) {      // the ray reached its light: run the light shader here (policy A-5/A-6)
                    const float4 w = __ldcs(pb.shadow.w + rayIdx);
                    const uint32_t bits = __float_as_uint(w.w);
                    shadeMiss<HR_ACCUM_RED != 0>(sc, fr, pb.accum, bits & 0x00ffffffu, (bits >> 24) & 7u, bits >> 27, tr.d, xyz(w), tr.tmaxRay, MASKED ? extraT : pb.shadow.d[rayIdx].w);
                } else {
                    __stcs(pb.hits + rayIdx, tr.found ? make_float4(tr.bestT, tr.bestU, tr.bestV, __int_as_float(tr.bestPrim))
                                                      : make_float4(0.f, 0.f, 0.f, __int_as_float(-1)));
                    doneIdx = rayIdx;
                    doneClass = tr.found ? 1 + ((tr.bestPrim >> 28) & 3) : 0;
                    if (pb.rayCounts && !tr.found && !fr.lights.envEnabled)     // parity hook: without an environment light no k_shade<0> runs to count the missed ray
                        atomicAdd(pb.rayCounts + __float_as_uint(((iter & 1) ? pb.ext[1].wm : pb.ext[0].wm)[2 * (size_t)rayIdx].w), 1u);
                }
                break;
            }
            if (!exhausted && 32 - __popc(__activemask()) >= kRefillIdle) break;
        }
    }
    if (STATS) {
        uint32_t n = tc.nodes, t = tc.tris, ov = tc.overflow;
        for (int s = 16; s > 0; s >>= 1) { n += __shfl_xor_sync(kFull, n, s); t += __shfl_xor_sync(kFull, t, s); ov += __shfl_xor_sync(kFull, ov, s); }
        if (lane == 0) {
            atomicAdd(ANY ? &pb.stats->nodesVisitedAny : &pb.stats->nodesVisited, (unsigned long long)n);
            atomicAdd(ANY ? &pb.stats->trianglesTestedAny : &pb.stats->trianglesTested, (unsigned long long)t);
            if (ov) atomicAdd(&pb.stats->stackOverflows, (unsigned long long)ov);
            if (!ANY && iter == 0) { atomicAdd(&pb.stats->nodesVisitedPrimary, (unsigned long long)n); atomicAdd(&pb.stats->trianglesTestedPrimary, (unsigned long long)t); }
        }
    }
}

// ---- material shading ------------------------------------------------------------------
struct InRay { float3 o, d, weight; float tmax; uint32_t depth, sequenceID, pixel; int sequenceIndexOffset; };
struct Emit {
    bool hasOcc, hasExt;
    float3 occO, occD, occW; float occTmax; uint32_t occMissKind, occMissIndex;
    float3 extO, extD, extW; uint32_t extMissKind;
};

__device__ __forceinline__ void emitOcc(Emit& e, const Surface& s, float3 dir, float3 w, uint32_t kind, uint32_t idx, float tmax) {
    e.hasOcc = true; e.occO = offsetOrigin(s, dir); e.occD = dir; e.occW = w; e.occTmax = tmax; e.occMissKind = kind; e.occMissIndex = idx;
}
__device__ __forceinline__ void emitExt(Emit& e, const Surface& s, float3 dir, float3 w, uint32_t kind) {
    e.hasExt = true; e.extO = offsetOrigin(s, dir); e.extD = dir; e.extW = w; e.extMissKind = kind;
}

struct LightSample { float3 dir; uint32_t missKind, missIndex; float probability, maxDistance; int lightType; };
#define LIGHT_TYPE_DIRECTIONAL 1
#define LIGHT_TYPE_POINT 2
#define LIGHT_TYPE_SPOT 3
#define LIGHT_TYPE_ENVIRONMENT 4

// lightSampling.rlsl:11-161. The shader fills three 5-entry weight arrays, normalises them and walks the CDF; here the
// weights are evaluated twice (sum, then walk) by one small function instead, so nothing lives in local memory and the
// loops stay rolled: identical fp32 values, a fraction of the code.
__device__ __forceinline__ float lightWeight(const DLights& L, int type, int i, float3 P, float3 N, float3& dir) {
    if (type == 0) {
        dir = f3(L.dirDirection[i]);
        return saturate(dot(N, dir)) * luminosity(f3(L.dirColor[i]));
    }
    if (type == 1) {
        dir = normalize(f3(L.pointPosition[i]) - P);
        return saturate(dot(N, dir)) * luminosity(f3(L.pointColor[i]));
    }
    dir = normalize(f3(L.spotPosition[i]) - P);
    const float rayAngle = dot(f3(L.spotDirection[i]), -dir);
    return saturate(dot(N, dir)) * luminosity(f3(L.spotColor[i])) *
           ((rayAngle > 0.0f) ? 1.0f : 0.0f) *
           ((rayAngle < L.spotAngles[i][1]) ? 0.0f : 1.0f) *
           (1.0f - glslSmoothstep(L.spotAngles[i][0], L.spotAngles[i][1], rayAngle));
}
__device__ __noinline__ void computeLightSample(const DLights& L, float3 P, float3 N, float lightProbability, LightSample& ls) {
    float probabilitySum = 0.0f;
    float3 dir;
#pragma unroll 1
    for (int type = 0; type < 3; ++type) {
        const int n = type == 0 ? L.nDir : (type == 1 ? L.nPoint : L.nSpot);
#pragma unroll 1
        for (int i = 0; i < n; ++i) probabilitySum += lightWeight(L, type, i, P, N, dir);
    }
    float environment = 0.0f;
    if (L.envEnabled) { environment = 50.0f * L.envExposure; probabilitySum += environment; }   // :77
    const float norm = 1.0f / greaterThanZero(probabilitySum);
    environment *= norm;
    float current = 0.0f;
#pragma unroll 1
    for (int type = 0; type < 3; ++type) {
        const int n = type == 0 ? L.nDir : (type == 1 ? L.nPoint : L.nSpot);
#pragma unroll 1
        for (int i = 0; i < n; ++i) {
            const float w = lightWeight(L, type, i, P, N, dir) * norm;
            current += w;
            if (w > 0.0f && lightProbability <= current) {
                ls.dir = dir; ls.missIndex = i; ls.probability = w;
                if (type == 0) { ls.missKind = kMissDir; ls.lightType = LIGHT_TYPE_DIRECTIONAL; }
                else if (type == 1) { ls.missKind = kMissPoint; ls.lightType = LIGHT_TYPE_POINT; ls.maxDistance = length(f3(L.pointPosition[i]) - P); }
                else { ls.missKind = kMissSpot; ls.lightType = LIGHT_TYPE_SPOT; ls.maxDistance = length(f3(L.spotPosition[i]) - P); }
                return;
            }
        }
    }
    ls.lightType = LIGHT_TYPE_ENVIRONMENT;
    ls.probability = environment;
}

__device__ __forceinline__ float3 computeMultiscattering(const DScene& sc, int lut, float3 Cspec, float NdotI, float roughness) {   // microfacet.rlsl:17-23
    float ms = 0.0f;
    if (lut >= 0) ms = texFetch(sc.textures[lut], NdotI, roughness).x;
    return f3(1.0f) + Cspec * ms;
}

// One BSDF lobe of the PBR material as the sample functions of microfacet.rlsl see it.
struct Lobe { bool diffuse; float3 N; float NdotI; float3 C; float alpha, roughness, probability; };
__device__ __forceinline__ float kInf() { return __int_as_float(0x7f800000); }

// physicallyBased.rlsl:55-331 with microfacet.rlsl:25-220 folded in. The direct (:236-274) and indirect (:277-329) halves
// run as two trips of ONE loop so that every sample function exists once in the instruction stream: the first version
// inlined each of them at 2-3 call sites, the kernel grew to ~200 KB of SASS and stalled on instruction fetch.
__device__ void shadePBR(const DScene& sc, const DFrame& fr, float4* accum, const InRay& in, const DMaterial& m, const Surface& s,
                         float3 interpNormal, bool frontFacing, Emit& e) {
    const HrcGlobals& g = fr.g;
    const uint32_t flags = m.typeFlags >> 4;
    float3 baseColor = f3(m.baseColor);
    float alpha = 1.0f;
    if (flags & HRC_MAT_HAS_BASE_COLOR_TEXTURE) {                             // :59-65
        const float4 t = sampleMaterialTexture(sc, m.baseColorTex, s.uv);
        baseColor = baseColor * xyz(t); alpha = t.w;
    }
    if (flags & HRC_MAT_VERTEX_COLORS) baseColor = baseColor * s.color;       // :66-68
    if ((flags & HRC_MAT_ALPHA_MASK) && alpha < 1.0f) {                       // :70-87 (extension rays; occlusion rays: k_trace<any>)
        e.hasExt = true; e.extO = offsetOrigin(s, in.d); e.extD = in.d; e.extW = in.weight; e.extMissKind = 0xffu;
        return;
    }
    float3 N = normalize(interpNormal);
    if (flags & HRC_MAT_DOUBLE_SIDED) { if (!frontFacing) N = -N; }          // :96-100
    else if (!frontFacing) {                                                  // :102-107 pass-through re-emit
        e.hasExt = true; e.extO = offsetOrigin(s, in.d); e.extD = in.d; e.extW = in.weight; e.extMissKind = 0xffu;   // 0xff: keep the incoming miss shader
        return;
    }
    float3 clearCoatN = N;
    float3 nmSample = f3(0.0f), ccnmSample = f3(0.0f);
    float metallic = m.metallic, roughness = m.roughness, roughnessAlpha = m.roughnessAlpha;
    float clearCoat = m.clearCoat, clearCoatRoughness = m.clearCoatRoughness, clearCoatRoughnessAlpha = m.clearCoatRoughnessAlpha;
    float3 emissive = f3(m.emissive);
    if (flags & kMatAnyTexture) {
        if (flags & kMatTangentSpace) {
            const float3 T = normalize(s.tangent), B = normalize(s.bitangent);
            if (flags & HRC_MAT_HAS_NORMALMAP) {                              // :112-118
                nmSample = xyz(sampleMaterialTexture(sc, m.normalTex, s.uv));
                const float3 ts = nmSample * 2.0f - f3(1.0f);
                N = normalize((T * ts.x + B * ts.y) + N * ts.z);
            }
            if (flags & HRC_MAT_HAS_CLEARCOAT_NORMALMAP) {                    // :120-126
                ccnmSample = xyz(sampleMaterialTexture(sc, m.clearCoatNormalTex, s.uv));
                const float3 ts = ccnmSample * 2.0f - f3(1.0f);
                clearCoatN = normalize((T * ts.x + B * ts.y) + clearCoatN * ts.z);
            }
        }
        if (flags & HRC_MAT_HAS_METALLIC_ROUGHNESS_TEXTURE) {                 // :135-140 (.bg)
            const float4 t = sampleMaterialTexture(sc, m.metallicRoughnessTex, s.uv);
            metallic *= t.z; roughness *= t.y; roughnessAlpha = roughness * roughness;
        }
        if (flags & HRC_MAT_HAS_CLEARCOAT_TEXTURE) clearCoat *= sampleMaterialTexture(sc, m.clearCoatTex, s.uv).x;            // :145-147
        if (flags & HRC_MAT_HAS_CLEARCOAT_ROUGHNESS_TEXTURE) {                // :148-151
            clearCoatRoughness *= sampleMaterialTexture(sc, m.clearCoatRoughnessTex, s.uv).x;
            clearCoatRoughnessAlpha = clearCoatRoughness * clearCoatRoughness;
        }
        if (flags & HRC_MAT_HAS_EMISSIVE_TEXTURE) emissive = xyz(sampleMaterialTexture(sc, m.emissiveTex, s.uv));             // :153-156 (replaces)
    }
    const M3 frame = orthonormalFrame(N);
    const float3 V = -in.d;
    const float NdotV = saturate(dot(N, V));

    if (g.enableDebugVisualizer == 1) {                                       // :159-203
        const bool ht = (flags & kMatAnyTexture) != 0, uts = (flags & kMatTangentSpace) != 0;
        if (g.showGeometricNormals == 1) accumulate4(accum, in.pixel, (interpNormal + f3(1.0f)) * 0.5f, 1.0f);
        else if (ht && g.showUVs == 1) accumulate4(accum, in.pixel, f3(s.uv.x, s.uv.y, 0.0f), 1.0f);
        else if (uts && g.showTangents == 1) accumulate4(accum, in.pixel, (s.tangent + f3(1.0f)) * 0.5f, 1.0f);
        else if (uts && g.showBitangents == 1) accumulate4(accum, in.pixel, (s.bitangent + f3(1.0f)) * 0.5f, 1.0f);
        else if ((flags & HRC_MAT_HAS_NORMALMAP) && g.showNormalmap == 1) accumulate4(accum, in.pixel, nmSample, 1.0f);
        else if (g.showFinalNormals == 1) accumulate4(accum, in.pixel, (N + f3(1.0f)) * 0.5f, 1.0f);
        else if (g.showBaseColor == 1) accumulate4(accum, in.pixel, baseColor, 1.0f);
        else if (g.showEmissive == 1) accumulate4(accum, in.pixel, emissive, 1.0f);
        else if (g.showRoughness == 1) accumulate4(accum, in.pixel, f3(roughness), 1.0f);
        else if (g.showMetallic == 1) accumulate4(accum, in.pixel, f3(metallic), 1.0f);
        else if (g.showClearcoat == 1) accumulate4(accum, in.pixel, f3(clearCoat), 1.0f);
        else if (g.showClearcoatRoughness == 1) accumulate4(accum, in.pixel, f3(clearCoatRoughness), 1.0f);
        else if (g.showShader == 1) accumulate4(accum, in.pixel, f3(1.0f, 0.0f, 0.0f), 1.0f);
        else if ((flags & HRC_MAT_HAS_CLEARCOAT_NORMALMAP) && g.showClearcoatNormalmap == 1) accumulate4(accum, in.pixel, ccnmSample, 1.0f);
        return;
    }

    float3 weight = in.weight;
    performAccumulate(g, accum, in.pixel, weight * emissive);                // :205
    const float clearCoatNdotV = saturate(dot(clearCoatN, V));
    const float clearCoatScale = F_Schlick1(0.04f, clearCoatNdotV) * clearCoat;
    const float bottom = 1.0f - clearCoatScale;
    const float3 Cdiff = (baseColor * (1.0f - metallic)) * bottom;         // :214
    const float3 Cspec = mix3(f3(m.specularF0), baseColor, metallic) * bottom;   // :220
    const float diffuseLuminance = luminosity(Cdiff), specularLuminance = luminosity(Cspec);
    const float pn = 1.0f / greaterThanZero(diffuseLuminance + specularLuminance + clearCoatScale);
    const float diffuseProbability = diffuseLuminance * pn, specularProbability = specularLuminance * pn, clearCoatProbability = clearCoatScale * pn;
    const SeqCursor seq(sc, fr, (int)in.sequenceID + (int)in.depth, g.sampleIndex + in.sequenceIndexOffset);

#pragma unroll 1
    for (int phase = 0; phase < 2; ++phase) {                                 // 0: direct lighting :236-274, 1: indirect :277-329
        if (phase == 1) {
            if (!((int)in.depth < g.maxRayDepth)) return;
            if (in.depth > 3) {                                               // Russian roulette, :279-288 (quirk B-Q12)
                const float2 rr = seq.at(2);
                const float probability = fmaxf(weight.x, fmaxf(weight.y, weight.z));
                if (rr.x >= probability) return;
                weight = weight / probability;
            }
        }
        const float2 rnd = seq.at(phase ? 3 : 0);
        const float2 pr = seq.at(phase ? 4 : 1);
        Lobe lb;
        if (pr.x <= diffuseProbability) {
            lb.diffuse = true; lb.N = N; lb.NdotI = NdotV; lb.C = Cdiff; lb.alpha = 0.f; lb.roughness = 0.f; lb.probability = diffuseProbability;
        } else if (pr.x <= (diffuseProbability + clearCoatProbability)) {
            lb.diffuse = false; lb.N = clearCoatN; lb.NdotI = clearCoatNdotV; lb.C = f3(clearCoatScale); lb.alpha = clearCoatRoughnessAlpha;
            lb.roughness = clearCoatRoughness; lb.probability = clearCoatProbability;
        } else if (pr.x <= (diffuseProbability + clearCoatProbability + specularProbability)) {
            lb.diffuse = false; lb.N = N; lb.NdotI = NdotV; lb.C = Cspec; lb.alpha = roughnessAlpha; lb.roughness = roughness; lb.probability = specularProbability;
        } else continue;

        float optLightProb = 1.0f;
        bool missEnv = false;
        if (phase == 0) {                                                     // directDiffuseSample :53-98 / directSpecularSample :155-220
            LightSample ls; ls.maxDistance = 0.f; ls.missKind = 0; ls.missIndex = 0; ls.dir = f3(0.f);
            computeLightSample(fr.lights, s.P, lb.N, pr.y, ls);
            if (ls.lightType != LIGHT_TYPE_ENVIRONMENT && pr.y > 0.0f) {      // quirk B-Q11
                float NdotO = dot(lb.N, ls.dir);
                if (NdotO > 0.0f) {
                    NdotO = saturate(NdotO);
                    float3 reflectance;
                    if (lb.diffuse) {
                        reflectance = ((lb.C / K_PI) * NdotO) * weight;
                    } else {
                        const float3 H = normalize(V + ls.dir);
                        const float NdotH = saturate(dot(lb.N, H));
                        const float IdotH = saturate(dot(V, H));
                        const float D = D_GGX(NdotH, lb.alpha);
                        const float3 F = F_Schlick3(lb.C, IdotH);
                        const float G = G2_Smith_GGX(NdotO, lb.NdotI, lb.alpha);
                        float3 specular = (D * F * G) / greaterThanZero(4.0f * lb.NdotI);
                        specular = specular * computeMultiscattering(sc, m.lutTex, lb.C, lb.NdotI, lb.roughness);
                        reflectance = specular * weight;
                    }
                    reflectance = reflectance / lb.probability;
                    reflectance = reflectance / ls.probability;
                    if (dot(reflectance, reflectance) > 1e-5f) {
                        const float maxT = (ls.lightType == LIGHT_TYPE_POINT || ls.lightType == LIGHT_TYPE_SPOT) ? ls.maxDistance : kInf();
                        emitOcc(e, s, ls.dir, reflectance, ls.missKind, ls.missIndex, maxT);
                    }
                }
                continue;
            }
            if (!(ls.probability > 0.0f)) continue;
            optLightProb = ls.probability;                                    // environment: BSDF-sampled occlusion ray (quirk B-Q8)
            missEnv = true;
        }
        // indirectDiffuseSample :25-51 / indirectSpecularSample :100-153
        float3 O, reflectance;
        if (lb.diffuse) {
            O = mul(frame, cosineWeightedSample(rnd.x, rnd.y));
            if (!(dot(lb.N, O) > 0.0f)) continue;
            reflectance = lb.C * weight;
        } else {
            const float3 H = mul(frame, sampleVisibleGGX(mulT(frame, V), rnd.x, rnd.y, lb.alpha));
            const float IdotH = saturate(dot(V, H));
            O = normalize(2.0f * IdotH * H - V);
            float NdotO = dot(lb.N, O);
            if (!(NdotO > 0.0f)) continue;
            NdotO = saturate(NdotO);
            const float3 F = F_Schlick3(lb.C, IdotH);
            const float G2 = G2_Smith_GGX(lb.NdotI, NdotO, lb.alpha);
            const float G1 = G1_Smith_GGX(lb.NdotI, lb.alpha);
            float3 specular = (F * G2) / greaterThanZero(G1);
            specular = specular * computeMultiscattering(sc, m.lutTex, lb.C, lb.NdotI, lb.roughness);
            reflectance = specular * weight;
        }
        reflectance = reflectance / lb.probability;
        reflectance = reflectance / optLightProb;
        if (dot(reflectance, reflectance) > 1e-5f) {
            if (missEnv) { if (fr.lights.envEnabled) emitOcc(e, s, O, reflectance, kMissEnv, 0, kInf()); }
            else emitExt(e, s, O, reflectance, kMissNone);
        }
    }
}

// glass.rlsl:138-280 with the sample functions (:47-129) folded in; the reflection branch runs its direct and indirect
// halves as two trips of one loop (see shadePBR).
__device__ void shadeGlass(const DScene& sc, const DFrame& fr, float4* accum, const InRay& in, const DMaterial& m, const Surface& s,
                           float3 interpNormal, bool frontFacing, Emit& e) {
    const HrcGlobals& g = fr.g;
    const uint32_t flags = m.typeFlags >> 4;
    float3 N = normalize(interpNormal);
    float nIn = 1.0f, nOut = m.ior;
    float3 weight = in.weight;
    float3 nmSample = f3(0.0f);
    if (flags & HRC_MAT_HAS_NORMALMAP) {                                      // :145-151
        const float3 T = normalize(s.tangent), B = normalize(s.bitangent);
        nmSample = xyz(sampleMaterialTexture(sc, m.normalTex, s.uv));
        const float3 ts = nmSample * 2.0f - f3(1.0f);
        N = normalize((T * ts.x + B * ts.y) + N * ts.z);
    }
    float3 baseColor = f3(m.baseColor);
    if (flags & HRC_MAT_HAS_BASE_COLOR_TEXTURE) baseColor = baseColor * xyz(sampleMaterialTexture(sc, m.baseColorTex, s.uv));   // :154-156
    if (!frontFacing) {                                                       // :161-167
        N = -N; nIn = m.ior; nOut = 1.0f;
        const float3 absorption = f3(1.0f) - baseColor;                       // beersLaw :131-136
        const float3 k = (absorption * m.density) * -s.T;                    // exp(absorption * Material.density * -rayLength): left to right, as the shader text associates it
        weight = in.weight * f3(expf(k.x), expf(k.y), expf(k.z));
    }
    float roughness = m.roughness, roughnessAlpha = m.roughnessAlpha;
    if (flags & HRC_MAT_HAS_METALLIC_ROUGHNESS_TEXTURE) {                     // :171-176 (.g; metallic ignored)
        roughness *= sampleMaterialTexture(sc, m.metallicRoughnessTex, s.uv).y;
        roughnessAlpha = roughness * roughness;
    }
    if (g.enableDebugVisualizer == 1) {                                       // :179-211
        const bool ht = (flags & kMatAnyTexture) != 0, uts = (flags & kMatTangentSpace) != 0;
        if (g.showGeometricNormals == 1) accumulate4(accum, in.pixel, (interpNormal + f3(1.0f)) * 0.5f, 1.0f);
        else if (g.showFinalNormals == 1) accumulate4(accum, in.pixel, (N + f3(1.0f)) * 0.5f, 1.0f);
        else if (g.showBaseColor == 1) accumulate4(accum, in.pixel, baseColor, 1.0f);
        else if (g.showRoughness == 1) accumulate4(accum, in.pixel, f3(roughness), 1.0f);
        else if (g.showShader == 1) accumulate4(accum, in.pixel, f3(0.0f, 1.0f, 0.0f), 1.0f);
        else if (ht && g.showUVs == 1) accumulate4(accum, in.pixel, f3(s.uv.x, s.uv.y, 0.0f), 1.0f);
        else if (uts && g.showTangents == 1) accumulate4(accum, in.pixel, (s.tangent + f3(1.0f)) * 0.5f, 1.0f);
        else if (uts && g.showBitangents == 1) accumulate4(accum, in.pixel, (s.bitangent + f3(1.0f)) * 0.5f, 1.0f);
        else if ((flags & HRC_MAT_HAS_NORMALMAP) && g.showNormalmap == 1) accumulate4(accum, in.pixel, nmSample, 1.0f);
        return;
    }
    const M3 frame = orthonormalFrame(N);
    const float3 I = -in.d;
    const float3 localI = mulT(frame, I);
    const float eta = nIn / nOut;
    const SeqCursor seq(sc, fr, (int)in.sequenceID + (int)in.depth, g.sampleIndex + in.sequenceIndexOffset);
    float2 rnd = seq.at(0);
    const float3 H0 = mul(frame, sampleVisibleGGX(localI, rnd.x, rnd.y, roughnessAlpha));
    const float HdotI = saturate(dot(H0, I));
    float2 rp = seq.at(1);
    const float F = F_Fresnel(eta, HdotI);
    const float NdotI = saturate(dot(N, I));
    if (!frontFacing) rp.y = 0.0f;                                            // :227-231
    if (rp.y < (1.0f - F)) {                                                  // :234-256 refraction
        const float3 mI = -I;
        const float dNI = dot(H0, mI);
        const float k = 1.0f - eta * eta * (1.0f - dNI * dNI);
        if (k < 0.0f) return;                                                 // refract() == 0 (policy A-16)
        const float3 R = eta * mI - (eta * dNI + sqrtf(k)) * H0;
        if (dot(R, R) == 0.0f) return;
        const float3 O = normalize(R);
        const float NdotO = fabsf(dot(N, O));
        const float G2 = G2_Smith_GGX(NdotI, NdotO, roughnessAlpha);
        const float G1 = G1_Smith_GGX(NdotI, roughnessAlpha);
        float3 transmission = baseColor * G2 / greaterThanZero(G1);
        transmission = transmission * weight;
        if (dot(transmission, transmission) > 1e-5f && (int)in.depth < g.maxRayDepth)
            emitExt(e, s, O, transmission, fr.lights.envEnabled ? kMissEnv : kMissNone);
        return;
    }
    // :257-279 reflection
#pragma unroll 1
    for (int phase = 0; phase < 2; ++phase) {
        float optLightProb = 1.0f;
        bool missEnv = false;
        if (phase == 0) {                                                     // directSpecularGlassSample :80-129
            rnd = seq.at(2);
            LightSample ls; ls.maxDistance = 0.f; ls.missKind = 0; ls.missIndex = 0; ls.dir = f3(0.f);
            computeLightSample(fr.lights, s.P, N, rp.x, ls);
            if (ls.lightType != LIGHT_TYPE_ENVIRONMENT) {
                float NdotO = dot(N, ls.dir);
                if (NdotO > 0.0f) {
                    NdotO = saturate(NdotO);
                    const float3 H = normalize(I + ls.dir);
                    const float NdotH = saturate(dot(N, H));
                    const float D = D_GGX(NdotH, roughnessAlpha);
                    const float G = G2_Smith_GGX(NdotO, NdotI, roughnessAlpha);
                    const float specular = (D * G) / greaterThanZero(4.0f * NdotI);
                    float3 reflectance = specular * baseColor;
                    reflectance = reflectance * weight;
                    reflectance = reflectance / ls.probability;
                    if (dot(reflectance, reflectance) > 1e-5f) {
                        const float maxT = (ls.lightType == LIGHT_TYPE_POINT || ls.lightType == LIGHT_TYPE_SPOT) ? ls.maxDistance : kInf();
                        emitOcc(e, s, ls.dir, reflectance, ls.missKind, ls.missIndex, maxT);
                    }
                }
                continue;
            }
            if (!(ls.probability > 0.0f)) continue;
            optLightProb = ls.probability; missEnv = true;
        } else {
            if (!((int)in.depth < g.maxRayDepth)) return;
            if (in.depth > 3) {                                               // :265-274
                const float2 rr = seq.at(3);
                const float probability = fmaxf(weight.x, fmaxf(weight.y, weight.z));
                if (rr.x >= probability) return;
                weight = weight / probability;
            }
            rnd = seq.at(4);
        }
        // indirectSpecularGlassSample :47-78
        const float3 H = mul(frame, sampleVisibleGGX(localI, rnd.x, rnd.y, roughnessAlpha));
        const float IdotH = saturate(dot(I, H));
        const float3 O = normalize(2.0f * IdotH * H - I);
        float NdotO = dot(N, O);
        if (NdotO > 0.0f) {
            NdotO = saturate(NdotO);
            const float NdotH = saturate(dot(N, H));
            const float G = G2_Smith_GGX(NdotO, NdotI, m.roughnessAlpha);     // Material.roughnessAlpha (quirk B-Q20)
            float3 reflectance = baseColor * ((G * IdotH) / (NdotH * NdotI));
            reflectance = reflectance * weight;
            reflectance = reflectance / optLightProb;
            if (dot(reflectance, reflectance) > 1e-5f) {
                if (missEnv) { if (fr.lights.envEnabled) emitOcc(e, s, O, reflectance, kMissEnv, 0, kInf()); }
                else emitExt(e, s, O, reflectance, kMissNone);
            }
        }
    }
}

// Material / miss shading, one kernel per ray class (CLS 0: miss, 1: PBR, 2: glass) over the index list k_trace<closest>
// compacted for that class: a warp runs ONE material program, and each kernel's code stays small enough for the
// instruction cache (the single all-materials kernel of the first version stalled on instruction fetch).
// The (<= 1) occlusion ray and (<= 1) extension ray of every hit are appended to the next queues with warp-aggregated
// atomics (ballot + popc prefix) — the compaction of live rays.
#ifndef HR_SHADE_MIN_BLOCKS
#define HR_SHADE_MIN_BLOCKS 4
#endif
#ifndef HR_SHADE_PREFETCH
#define HR_SHADE_PREFETCH 0
#endif
// Sort key of an extension ray: octant of its direction above the Morton code of its origin's cell (device_types.h).
__device__ __forceinline__ uint32_t spreadBits3(uint32_t v) {          // 5 low bits -> every third bit
    v = (v | (v << 8)) & 0x0000100fu;
    v = (v | (v << 4)) & 0x000010c3u;
    v = (v | (v << 2)) & 0x00001249u;
    return v;
}
__device__ __forceinline__ uint32_t raySortKey(const DScene& sc, float3 o, float3 d, int bits) {
    const int top = (1 << bits) - 1;
    const uint32_t ix = (uint32_t)min(max((int)((o.x - sc.sortMin[0]) * sc.sortScale[0]), 0), top);
    const uint32_t iy = (uint32_t)min(max((int)((o.y - sc.sortMin[1]) * sc.sortScale[1]), 0), top);
    const uint32_t iz = (uint32_t)min(max((int)((o.z - sc.sortMin[2]) * sc.sortScale[2]), 0), top);
    const uint32_t oct = (d.x < 0.0f ? 1u : 0u) | (d.y < 0.0f ? 2u : 0u) | (d.z < 0.0f ? 4u : 0u);
    return (oct << (3 * bits)) | spreadBits3(ix) | (spreadBits3(iy) << 1) | (spreadBits3(iz) << 2);
}

// The material shader of one hit: `hr` is the hit record k_trace<closest> writes (t, u, v, primitive | class << 28 | HRC_HIT_BACKFACE).
// CLS 1: physicallyBased.rlsl, 2: glass.rlsl. One body for the wavefront's k_shade<CLS> and the path kernel k_path.
template <int CLS>
__device__ __forceinline__ void shadeHit(const DScene& sc, const DFrame& fr, float4* accum, const InRay& in, float4 hr, Emit& e) {
    const int primBits = __float_as_int(hr.w);
    const int prim = primBits & 0x0fffffff;
    const bool back = (primBits & HRC_HIT_BACKFACE) != 0;
    const float4* sp = sc.shade + (size_t)prim * 4;
    const float4 s0 = __ldg(sp), s1 = __ldg(sp + 1), s2 = __ldg(sp + 2), s3 = __ldg(sp + 3);
    Surface s;
    s.P = in.o + in.d * hr.x;
    s.Ng = xyz(s3);
    s.T = hr.x;
    const bool frontFacing = __float_as_int(s1.w) ? back : !back;   // policy A-10
    const float bw = (1.0f - hr.y) - hr.z;
    const float3 n = (xyz(s0) * bw + xyz(s1) * hr.y) + xyz(s2) * hr.z;
    const DMaterial m = sc.materials[__float_as_int(s0.w)];
    s.uv = make_float2(0.f, 0.f); s.tangent = f3(0.f); s.bitangent = f3(0.f); s.color = f3(0.f);
    const uint32_t mflags = m.typeFlags >> 4;
    if (sc.attr && (mflags & (kMatAnyTexture | HRC_MAT_VERTEX_COLORS)))
        loadVaryings(sc, prim, hr.y, hr.z, (mflags & kMatTangentSpace) != 0, (mflags & HRC_MAT_VERTEX_COLORS) != 0, s);
    if (CLS == 1) shadePBR(sc, fr, accum, in, m, s, n, frontFacing, e);
    else          shadeGlass(sc, fr, accum, in, m, s, n, frontFacing, e);
}

template <int CLS>
__global__ void __launch_bounds__(128, HR_SHADE_MIN_BLOCKS) k_shade(const DScene sc, const int slot, PassBuffers pb, int iter, int sortBits) {
    const DFrame& fr = g_frames[slot];
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warpsTotal = (gridDim.x * blockDim.x) >> 5;
    const uint32_t warpId = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    IterCtrl* ctrl = pb.ctrl + iter;
    const uint32_t count = ctrl->classCount[CLS];
    const uint32_t* __restrict__ list = pb.classIdx[CLS];
    const RayQueue rq = (iter & 1) ? pb.ext[1] : pb.ext[0];
    const RayQueue nq = (iter & 1) ? pb.ext[0] : pb.ext[1];
    const int hardCap = fr.g.maxRayDepth + 16;                                // policy A-14
    const bool lastIter = (iter + 1 >= kMaxIterations);

#if HR_SHADE_PREFETCH
    // The list entries of the next two trips are loaded ahead; the ray and hit records of the next trip are pulled into L2
    // while this trip shades (they are streamed from DRAM otherwise: three dependent misses before the first FMA).
    const uint32_t stride = warpsTotal * 32u;
    uint32_t iNext = 0xffffffffu, iNext2 = 0xffffffffu;
    if (CLS != 0) {
        const uint32_t k0 = warpId * 32u + lane;
        if (k0 < count) iNext = list[k0];
        if (k0 + stride < count) iNext2 = list[k0 + stride];
    }
#endif
    for (uint32_t base = warpId * 32u; base < count; base += warpsTotal * 32u) {
        const uint32_t k = base + lane;
        Emit e; e.hasOcc = false; e.hasExt = false;
        InRay in; in.depth = 0; in.pixel = 0; in.sequenceID = 0; in.sequenceIndexOffset = 0;
        uint32_t inMissKind = 0;
#if HR_SHADE_PREFETCH
        const uint32_t iCur = iNext;
        if (CLS != 0) {
            iNext = iNext2;
            iNext2 = (k + 2u * stride < count) ? list[k + 2u * stride] : 0xffffffffu;
            if (iNext != 0xffffffffu) {
                prefetchL2(rq.od + 2 * (size_t)iNext); prefetchL2(rq.wm + 2 * (size_t)iNext); prefetchL2(pb.hits + iNext);
            }
        }
#endif
        if (k < count) {
#if HR_SHADE_PREFETCH
            const size_t i = CLS != 0 ? iCur : list[k];
#else
            const size_t i = list[k];
#endif
            const float4 rd = __ldcs(rq.od + 2 * i + 1), rw = __ldcs(rq.wm + 2 * i), rm = __ldcs(rq.wm + 2 * i + 1);
            const uint32_t mx = __float_as_uint(rm.x);
            in.d = xyz(rd); in.weight = xyz(rw);
            in.pixel = __float_as_uint(rw.w);
            in.depth = mx & 0xffu; inMissKind = (mx >> 8) & 0xfu; in.sequenceID = mx >> 16; in.sequenceIndexOffset = __float_as_int(rm.y);
            if (CLS == 0) {
                shadeMiss(sc, fr, pb.accum, in.pixel, inMissKind, 0u, in.d, in.weight, 0.0f, 0.0f);   // extension rays only miss into the environment
            } else {
                const float4 ro = __ldcs(rq.od + 2 * i);
                const float4 hr = __ldcs(pb.hits + i);
                in.o = xyz(ro); in.tmax = ro.w;
                shadeHit<CLS>(sc, fr, pb.accum, in, hr, e);
                if (e.hasExt && ((int)in.depth + 1 > hardCap || lastIter)) e.hasExt = false;
            }
        }
        if (pb.rayCounts && k < count) atomicAdd(pb.rayCounts + in.pixel, 1u + (e.hasOcc ? 1u : 0u));   // parity hook: this ray and the occlusion ray it emits
        if (CLS == 0) continue;
        // ---- compaction: one atomic per warp per queue ------------------------------------
        const uint32_t occMask = __ballot_sync(0xffffffffu, e.hasOcc);
        const uint32_t extMask = __ballot_sync(0xffffffffu, e.hasExt);
        uint32_t occBase = 0, extBase = 0;
        if (lane == 0) {
            if (occMask) occBase = atomicAdd(&ctrl->shadowCount, (uint32_t)__popc(occMask));
            if (extMask) extBase = atomicAdd(&ctrl[1].extCount, (uint32_t)__popc(extMask));
        }
        occBase = __shfl_sync(0xffffffffu, occBase, 0);
        extBase = __shfl_sync(0xffffffffu, extBase, 0);
        const uint32_t below = (1u << lane) - 1u;
        if (e.hasOcc) {
            const uint32_t j = occBase + __popc(occMask & below);
            __stcs(pb.shadow.o + j, make_float4(e.occO.x, e.occO.y, e.occO.z, e.occTmax));
            __stcs(pb.shadow.d + j, make_float4(e.occD.x, e.occD.y, e.occD.z, 0.0f));
            __stcs(pb.shadow.w + j, make_float4(e.occW.x, e.occW.y, e.occW.z, __uint_as_float(in.pixel | (e.occMissKind << 24) | (e.occMissIndex << 27))));
        }
        if (e.hasExt) {
            const size_t j = extBase + __popc(extMask & below);
            const uint32_t mk = e.extMissKind == 0xffu ? inMissKind : e.extMissKind;
            __stcs(nq.od + 2 * j, make_float4(e.extO.x, e.extO.y, e.extO.z, __int_as_float(0x7f800000)));
            __stcs(nq.od + 2 * j + 1, make_float4(e.extD.x, e.extD.y, e.extD.z, 0.0f));
            __stcs(nq.wm + 2 * j, make_float4(e.extW.x, e.extW.y, e.extW.z, __uint_as_float(in.pixel)));
            __stcs(nq.wm + 2 * j + 1, make_float4(__uint_as_float((in.depth + 1u) | (mk << 8) | (in.sequenceID << 16)), __int_as_float(in.sequenceIndexOffset), 0.0f, 0.0f));
            if (sortBits > 0) {
                const uint32_t key = raySortKey(sc, e.extO, e.extD, sortBits);
                __stcs(pb.sortKeys + j, key);
                atomicAdd(pb.sortHist + key, 1u);
            }
        }
    }
}

// ---- tail of a pass: the remaining paths in one launch ---------------------------------
// Late wavefront iterations hold a few thousand rays each and cost a fixed ~60-90 us of kernel boundaries (trace, three shade classes,
// occlusion trace). k_tail_decide runs after shade(iter - 1): when the extension queue of iteration `iter` has come down to `threshold`
// rays it hands them to k_path and marks the queue drained, so the wavefront kernels still to be launched (the launch sequence is fixed
// when the pass graph is captured) pull nothing and return at once.
__global__ void k_tail_decide(PassBuffers pb, int iter, uint32_t threshold) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    IterCtrl* tail = pb.ctrl + kMaxIterations;
    if (tail->extCount) return;                                   // the tail of this pass has been handed over already
    const uint32_t n = pb.ctrl[iter].extCount;
    if (n == 0u || n > threshold) return;
    tail->extCount = n; tail->extHead = (uint32_t)iter;
    pb.ctrl[iter].extHead = kQueueDrained;                        // extCount stays: k_finish_pass counts these rays as before
}

// One thread per remaining path: what k_trace<closest>, k_shade<0..2> and k_trace<any> of iterations iter, iter + 1, ... would do for it, in
// the same order (shade(i) accumulates, then the light shader of its occlusion ray, then shade(i + 1)), with the same device functions:
// the pixel sums are bit-identical to the wavefront's. Rays stay in registers; only the counters and the pixels are written.
template <bool MASKED>
__global__ void __launch_bounds__(128, 3) k_path(const DScene sc, const int slot, const PassBuffers pb, int iter, int iterations) {
    const IterCtrl* tail = pb.ctrl + kMaxIterations;
    const uint32_t count = tail->extCount;
    if (count == 0u || tail->extHead != (uint32_t)iter) return;
    const DFrame& fr = g_frames[slot];
    const RayQueue rq = (iter & 1) ? pb.ext[1] : pb.ext[0];
    const float4* __restrict__ nodes = sc.nodes;
    const float4* __restrict__ tris = sc.tris;
    const int hardCap = fr.g.maxRayDepth + 16;                                // policy A-14
    uint32_t extRays = 0, occRays = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        const float4 ro = __ldcs(rq.od + 2 * (size_t)i), rd = __ldcs(rq.od + 2 * (size_t)i + 1), rw = __ldcs(rq.wm + 2 * (size_t)i), rm = __ldcs(rq.wm + 2 * (size_t)i + 1);
        const uint32_t mx = __float_as_uint(rm.x);
        InRay in;
        in.o = xyz(ro); in.tmax = ro.w; in.d = xyz(rd); in.weight = xyz(rw); in.pixel = __float_as_uint(rw.w);
        in.depth = mx & 0xffu; in.sequenceID = mx >> 16; in.sequenceIndexOffset = __float_as_int(rm.y);
        uint32_t missKind = (mx >> 8) & 0xfu;
        for (int it = iter;; ++it) {
            Traversal tr;
            tr.begin(in.o, in.d, 0.0f, in.tmax);
            traverseSimple<false>(nodes, tris, tr);
            Emit e; e.hasOcc = false; e.hasExt = false;
            if (!tr.found) {
                shadeMiss(sc, fr, pb.accum, in.pixel, missKind, 0u, in.d, in.weight, 0.0f, 0.0f);   // extension rays only miss into the environment
            } else {
                const float4 hr = make_float4(tr.bestT, tr.bestU, tr.bestV, __int_as_float(tr.bestPrim));
                if (((tr.bestPrim >> 28) & 3) == 0) shadeHit<1>(sc, fr, pb.accum, in, hr, e);
                else                                shadeHit<2>(sc, fr, pb.accum, in, hr, e);
                if (e.hasExt && ((int)in.depth + 1 > hardCap || it + 1 >= kMaxIterations)) e.hasExt = false;
            }
            if (pb.rayCounts) atomicAdd(pb.rayCounts + in.pixel, 1u + (e.hasOcc ? 1u : 0u));
            if (e.hasOcc) {                                        // k_trace<any>(it): the light shader runs when nothing occludes
                ++occRays;
                float extraT = 0.0f; int maskedPasses = 0;
                tr.begin(e.occO, e.occD, 0.0f, e.occTmax);
                bool lit;
                for (;;) {
                    if (traverseSimple<true, MASKED>(nodes, tris, tr)) { lit = false; break; }
                    if (MASKED && (tr.bestPrim & 0x0fffffff) != 0x0fffffff) {
                        if (maskedSurfaceStep(sc, tr, extraT, maskedPasses)) {
                            ++occRays;
                            if (pb.rayCounts) atomicAdd(pb.rayCounts + in.pixel, 1u);
                            continue;
                        }
                        lit = false; break;
                    }
                    lit = true; break;
                }
                if (lit) shadeMiss(sc, fr, pb.accum, in.pixel, e.occMissKind, e.occMissIndex, tr.d, e.occW, tr.tmaxRay, extraT);
            }
            if (!e.hasExt || it + 1 >= iterations) break;          // rays emitted by the last iteration of the pass are not traced
            ++extRays;
            if (e.extMissKind != 0xffu) missKind = e.extMissKind;
            in.o = e.extO; in.d = e.extD; in.weight = e.extW; in.tmax = __int_as_float(0x7f800000); in.depth += 1u;
        }
    }
    for (int sft = 16; sft > 0; sft >>= 1) { extRays += __shfl_xor_sync(0xffffffffu, extRays, sft); occRays += __shfl_xor_sync(0xffffffffu, occRays, sft); }
    if ((threadIdx.x & 31u) == 0u) {
        if (extRays) atomicAdd(&pb.stats->raysExtension, (unsigned long long)extRays);
        if (occRays) atomicAdd(&pb.stats->raysOcclusion, (unsigned long long)occRays);
    }
}

// Histogram -> exclusive bin offsets for the extension rays of iteration `iter`, one CTA of 1024 threads (bins is a multiple of 1024);
// re-arms the histogram for the next iteration. Iterations with few rays are latency-bound anyway and stay unsorted.
__global__ void __launch_bounds__(1024) k_sort_scan(PassBuffers pb, int iter, uint32_t bins) {
    __shared__ uint32_t warpSums[32];
    const uint32_t t = threadIdx.x, lane = t & 31u, warp = t >> 5;
    const uint32_t per = bins >> 10;
    const bool sorted = pb.ctrl[iter].extCount >= kSortMinRays;
    if (t == 0) pb.sortCtrl[iter].sorted = sorted ? 1u : 0u;
    uint32_t* hist = pb.sortHist + (size_t)t * per;
    uint32_t* offs = pb.sortOffs + (size_t)t * per;
    uint32_t sum = 0;
    if (sorted) for (uint32_t k = 0; k < per; ++k) sum += hist[k];
    uint32_t incl = sum;
    for (int d = 1; d < 32; d <<= 1) { const uint32_t v = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= (uint32_t)d) incl += v; }
    if (lane == 31) warpSums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = warpSums[lane];
        for (int d = 1; d < 32; d <<= 1) { const uint32_t v = __shfl_up_sync(0xffffffffu, w, d); if (lane >= (uint32_t)d) w += v; }
        warpSums[lane] = w;
    }
    __syncthreads();
    uint32_t run = incl - sum + (warp ? warpSums[warp - 1] : 0u);
    for (uint32_t k = 0; k < per; ++k) {
        const uint32_t h = hist[k];
        if (sorted) offs[k] = run;
        run += h;
        hist[k] = 0u;
    }
}
__global__ void __launch_bounds__(256) k_sort_scatter(PassBuffers pb, int iter) {
    if (!pb.sortCtrl[iter].sorted) return;
    const uint32_t count = pb.ctrl[iter].extCount;
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < count; j += gridDim.x * blockDim.x) {
        const uint32_t key = __ldcs(pb.sortKeys + j);
        pb.sortOrder[atomicAdd(pb.sortOffs + key, 1u)] = j;
    }
}

// The light ("miss") shaders of the occlusion rays of iteration `iter` that reached their light (policy A-5 / A-6), one ray per thread over the list
// k_trace<any> filed them in. Same shadeMiss, same inputs as the inline call it replaces, one occlusion ray per pixel and iteration: pixel sums are bit-identical.
__global__ void __launch_bounds__(128) k_shade_lights(const DScene sc, const int slot, const PassBuffers pb, int iter) {
    const DFrame& fr = g_frames[slot];
    const uint32_t count = pb.ctrl[iter].litCount;
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < count; k += gridDim.x * blockDim.x) {
        const uint32_t i = pb.litIdx[k];
        const float4 w = __ldcs(pb.shadow.w + i), d = __ldcs(pb.shadow.d + i);
        const float tmax = __ldcs(pb.shadow.o + i).w;
        const uint32_t bits = __float_as_uint(w.w);
        shadeMiss(sc, fr, pb.accum, bits & 0x00ffffffu, (bits >> 24) & 7u, bits >> 27, xyz(d), xyz(w), tmax, d.w);
    }
}

__global__ void k_finish_pass(PassBuffers pb, int iterations, int rearm) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    unsigned long long ext = 0, occ = 0;
    for (int i = 1; i < iterations; ++i) ext += pb.ctrl[i].extCount;
    for (int i = 0; i < iterations; ++i) occ += pb.ctrl[i].shadowCount + pb.ctrl[i].maskedContinuations;
    pb.stats->raysPrimary += pb.ctrl[0].extCount;
    pb.stats->raysExtension += ext;
    pb.stats->raysOcclusion += occ;
    if (rearm) for (int i = 0; i < kCtrlEntries; ++i) pb.ctrl[i] = IterCtrl{};   // the next pass on these buffers starts without a clear kernel
}

// hrcTraceRays: the fixed-ray-set parity hook. Same node / triangle steps as the pass kernels.
template <bool ANY>
__global__ void __launch_bounds__(128) k_trace_rays(const DScene sc, const HrcRay* __restrict__ rays, HrcHit* __restrict__ hits, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const HrcRay r = rays[i];
    Traversal tr;
    tr.begin(f3(r.origin[0], r.origin[1], r.origin[2]), f3(r.direction[0], r.direction[1], r.direction[2]), r.tmin, r.tmax);
    traverseSimple<ANY>(sc.nodes, sc.tris, tr);
    HrcHit out;
    if (ANY) { out.t = 0.f; out.u = 0.f; out.v = 0.f; out.primId = tr.found ? 0 : -1; }
    else if (tr.found) { out.t = tr.bestT; out.u = tr.bestU; out.v = tr.bestV; out.primId = tr.bestPrim & (0x0fffffff | HRC_HIT_BACKFACE); }
    else { out.t = 0.f; out.u = 0.f; out.v = 0.f; out.primId = -1; }
    hits[i] = out;
}

__global__ void k_add_pixels(float4* accum, const float4* src, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { float4 a = accum[i], b = src[i]; a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; accum[i] = a; }
}

} // namespace

// ---------------------------------------------------------------------------------------
// Launch wrappers
// ---------------------------------------------------------------------------------------
void launchClearCtrl(const PassBuffers& pb, cudaStream_t s) { k_clear_ctrl<<<1, 128, 0, s>>>(pb.ctrl); }

void uploadFrame(const DFrame* pinnedFrame, int slot, cudaStream_t s) {
    cudaMemcpyToSymbolAsync(g_frames, pinnedFrame, sizeof(DFrame), (size_t)slot * sizeof(DFrame), cudaMemcpyHostToDevice, s);
}

void launchRaygen(const DScene& sc, int slot, uint32_t total, const PassBuffers& pb, cudaStream_t s) {
    if (total == 0) return;
    k_raygen<<<(total + 255) / 256, 256, 0, s>>>(sc, slot, pb.ext[0], pb.accum, pb.ctrl);
}

void launchTraceClosest(const DScene& sc, int slot, const PassBuffers& pb, int iter, const LaunchConfig& lc, cudaStream_t s) {
    const int grid = lc.smCount * lc.traceBlocksPerSM;
    if (lc.collectStats) k_trace<false, true, false><<<grid, 128, 0, s>>>(sc, slot, pb, iter, lc.refillIdle);
    else                 k_trace<false, false, false><<<grid, 128, 0, s>>>(sc, slot, pb, iter, lc.refillIdle);
}

void launchTraceAny(const DScene& sc, int slot, const PassBuffers& pb, int iter, const LaunchConfig& lc, cudaStream_t s) {
    const int grid = lc.smCount * lc.traceBlocksPerSM;
    if (lc.masked) {
        if (lc.collectStats) k_trace<true, true, true><<<grid, 128, 0, s>>>(sc, slot, pb, iter, lc.refillIdleAny);
        else                 k_trace<true, false, true><<<grid, 128, 0, s>>>(sc, slot, pb, iter, lc.refillIdleAny);
    } else {
        if (lc.collectStats) k_trace<true, true, false><<<grid, 128, 0, s>>>(sc, slot, pb, iter, lc.refillIdleAny);
        else                 k_trace<true, false, false><<<grid, 128, 0, s>>>(sc, slot, pb, iter, lc.refillIdleAny);
    }
}

void launchShadeLights(const DScene& sc, int slot, const PassBuffers& pb, int iter, const LaunchConfig& lc, cudaStream_t s) {
    if (pb.litIdx) k_shade_lights<<<lc.smCount * 8, 128, 0, s>>>(sc, slot, pb, iter);
}

void launchShade(const DScene& sc, int slot, const PassBuffers& pb, int iter, const LaunchConfig& lc, cudaStream_t s) {
    const int grid = lc.smCount * 8;
    const int sortBits = pb.sortKeys ? lc.sortBits : 0;
    if (lc.shadeMiss)  k_shade<0><<<grid, 128, 0, s>>>(sc, slot, pb, iter, 0);
    if (lc.shadePbr)   k_shade<1><<<grid, 128, 0, s>>>(sc, slot, pb, iter, sortBits);
    if (lc.shadeGlass) k_shade<2><<<grid, 128, 0, s>>>(sc, slot, pb, iter, sortBits);
}

void launchSortRays(const PassBuffers& pb, int iter, const LaunchConfig& lc, cudaStream_t s) {
    if (!pb.sortKeys || lc.sortBits <= 0) return;
    const uint32_t bins = 8u << (3 * lc.sortBits);
    k_sort_scan<<<1, 1024, 0, s>>>(pb, iter, bins < 1024u ? 1024u : bins);
    k_sort_scatter<<<lc.smCount * 4, 256, 0, s>>>(pb, iter);
}

void launchTailDecide(const PassBuffers& pb, int iter, const LaunchConfig& lc, cudaStream_t s) {
    if (lc.tailRays > 0) k_tail_decide<<<1, 32, 0, s>>>(pb, iter, (uint32_t)lc.tailRays);
}
void launchPathTail(const DScene& sc, int slot, const PassBuffers& pb, int iter, int iterations, const LaunchConfig& lc, cudaStream_t s) {
    if (lc.tailRays <= 0) return;
    const int grid = (lc.tailRays + 127) / 128;                    // one thread per ray of the largest tail that is taken
    if (lc.masked) k_path<true><<<grid, 128, 0, s>>>(sc, slot, pb, iter, iterations);
    else           k_path<false><<<grid, 128, 0, s>>>(sc, slot, pb, iter, iterations);
}

void launchFinishPass(const PassBuffers& pb, int iterations, bool rearm, cudaStream_t s) { k_finish_pass<<<1, 32, 0, s>>>(pb, iterations, rearm ? 1 : 0); }

void launchTraceRays(const DScene& sc, const HrcRay* rays, HrcHit* hits, int n, bool anyHit, const LaunchConfig&, cudaStream_t s) {
    if (n <= 0) return;
    if (anyHit) k_trace_rays<true><<<(n + 127) / 128, 128, 0, s>>>(sc, rays, hits, n);
    else        k_trace_rays<false><<<(n + 127) / 128, 128, 0, s>>>(sc, rays, hits, n);
}

// ---- hrcTraceRaysNoAlloc: the fixed-ray-set hook with its staging buffers in a static device array --------------------------------
// hrcTraceRays stages the host rays / hits in device buffers it allocates per call; callers that must not allocate (see the header)
// get the same kernel over this array instead. One array per GPU that loads the library, held until the process exits:
// kNoAllocMaxRays * (32 B ray + 16 B hit) = 12 MiB. Every call on a device shares it.
namespace {
static __device__ __align__(256) unsigned char g_traceScratch[(size_t)kNoAllocMaxRays * (sizeof(HrcRay) + sizeof(HrcHit))];
constexpr int kScratchDevices = 64;
std::mutex g_scratchMutex;
unsigned char* g_scratchAddr[kScratchDevices] = {};      // device address of g_traceScratch on each device, resolved on first use
}

// In host code `g_traceScratch` is only the host-side handle of the symbol, never a device address: the address the kernel needs comes
// from cudaGetSymbolAddress, per device, with that device current (the caller has made `device` current).
cudaError_t traceScratchBuffers(int device, HrcRay** rays, HrcHit** hits) {
    if (device < 0 || device >= kScratchDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_scratchMutex);
    if (!g_scratchAddr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_traceScratch);
        if (e != cudaSuccess) return e;
        g_scratchAddr[device] = static_cast<unsigned char*>(p);
    }
    *rays = reinterpret_cast<HrcRay*>(g_scratchAddr[device]);
    *hits = reinterpret_cast<HrcHit*>(g_scratchAddr[device] + (size_t)kNoAllocMaxRays * sizeof(HrcRay));
    return cudaSuccess;
}

void launchAddPixels(float4* accum, const float4* src, size_t n, cudaStream_t s) {
    if (n) k_add_pixels<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(accum, src, n);
}

} // namespace hr

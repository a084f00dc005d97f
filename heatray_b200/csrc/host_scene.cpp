// host_scene.cpp — host-side parameter packing of the pass interface: materials, lights, camera,
// procedural mesh providers and the multiscatter LUT generator. Each function restates one piece of
// reference host code (cited inline); the arithmetic is fp32 with glm's evaluation order.
#include "host_api.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

namespace hr {
namespace {

inline float saturate(float v) { return std::min(std::max(v, 0.0f), 1.0f); }
constexpr float kPi = 3.14159265358979323846f;
constexpr float kLumensToWatts = 1.0f / 683.0f;     // Lights/Light.h LUMENS_TO_WATTS

struct Quat { float w, x, y, z; };
inline Quat angleAxis(float angle, float ax, float ay, float az) {          // glm::angleAxis
    const float s = std::sin(angle * 0.5f);
    return Quat{std::cos(angle * 0.5f), ax * s, ay * s, az * s};
}
inline Quat mul(const Quat& p, const Quat& q) {                             // glm quat operator*
    return Quat{p.w * q.w - p.x * q.x - p.y * q.y - p.z * q.z,
                p.w * q.x + p.x * q.w + p.y * q.z - p.z * q.y,
                p.w * q.y + p.y * q.w + p.z * q.x - p.x * q.z,
                p.w * q.z + p.z * q.w + p.x * q.y - p.y * q.x};
}
inline Quat inverse(const Quat& q) {                                        // conjugate(q) / dot(q, q)
    const float d = (q.w * q.w + q.x * q.x) + (q.y * q.y + q.z * q.z);   // glm's quaternion dot order
    return Quat{q.w / d, -q.x / d, -q.y / d, -q.z / d};
}
inline void mat3Cast(const Quat& q, float m[9]) {                           // glm::mat3_cast, column-major
    const float qxx = q.x * q.x, qyy = q.y * q.y, qzz = q.z * q.z, qxz = q.x * q.z, qxy = q.x * q.y, qyz = q.y * q.z;
    const float qwx = q.w * q.x, qwy = q.w * q.y, qwz = q.w * q.z;
    m[0] = 1.0f - 2.0f * (qyy + qzz); m[1] = 2.0f * (qxy + qwz);        m[2] = 2.0f * (qxz - qwy);
    m[3] = 2.0f * (qxy - qwz);        m[4] = 1.0f - 2.0f * (qxx + qzz); m[5] = 2.0f * (qyz + qwx);
    m[6] = 2.0f * (qxz + qwy);        m[7] = 2.0f * (qyz - qwx);        m[8] = 1.0f - 2.0f * (qxx + qyy);
}
// -viewMatrix[2].xyz of mat4_cast(inverse(angleAxis(theta, X) * angleAxis(phi, Y))), normalised
inline void orientationForward(float theta, float phi, float out[3]) {
    const Quat o = mul(angleAxis(theta, 1.0f, 0.0f, 0.0f), angleAxis(phi, 0.0f, 1.0f, 0.0f));
    float m[9]; mat3Cast(inverse(o), m);
    float d[3] = {m[6] * -1.0f, m[7] * -1.0f, m[8] * -1.0f};
    const float inv = 1.0f / std::sqrt((d[0] * d[0] + d[1] * d[1]) + d[2] * d[2]);
    out[0] = d[0] * inv; out[1] = d[1] * inv; out[2] = d[2] * inv;
}

// --- multiscatter LUT (MultiScatterUtil.cpp:21-89) ---
inline float lutG1(float NdotI, float alpha) {
    const float a2 = alpha * alpha;
    const float denom = sqrtf(a2 + (1.0f - a2) * (NdotI * NdotI)) + NdotI;
    return (2.0f * NdotI) / std::max(denom, 1e-5f);
}
float lutValue(float NdotV, float alpha, const std::vector<float>& seq) {
    const size_t n = seq.size() / 2;
    const float Vx = sqrtf(1.0f - (NdotV * NdotV)), Vz = NdotV;       // V = (Vx, 0, Vz), N = +Z
    float result = 0.0f;
    for (size_t i = 0; i < n; ++i) {
        const float rx = seq[2 * i], ry = seq[2 * i + 1];
        const float a2 = alpha * alpha;
        const float cosTheta = sqrtf(std::max(0.0f, (1.0f - rx) / ((a2 - 1.0f) * rx + 1.0f)));
        const float sinTheta = sqrtf(std::max(0.0f, 1.0f - cosTheta * cosTheta));
        const float phi = 6.28318530717958647692f * ry;
        float hx = sinTheta * cosf(phi), hy = sinTheta * sinf(phi), hz = cosTheta;
        const float hl = 1.0f / std::sqrt((hx * hx + hy * hy) + hz * hz);           // glm::normalize = v * inversesqrt(dot)
        hx *= hl; hy *= hl; hz *= hl;
        const float VdotHraw = (Vx * hx + 0.0f * hy) + Vz * hz;
        const float k = 2.0f * VdotHraw;
        const float Lz = k * hz - Vz;                                  // only N.L = L.z is needed
        const float NdotL = saturate(Lz);
        if (NdotL > 0.0f) {
            const float VdotH = saturate(VdotHraw);
            const float NdotH = saturate(hz);
            result += (lutG1(NdotL, alpha) * lutG1(NdotV, alpha) * VdotH) / (NdotV * NdotH);
        }
    }
    return result / float(n);
}

} // namespace

uint32_t packPbrMaterial(const HrhPbrMaterial& in, HrcTexture lut, HrcPbrParams& out) {   // PhysicallyBasedMaterial.cpp:127-191
    const float kMinRoughness = 0.01f, kMaxSpecularF0 = 0.08f, kMaxClearcoat = 0.2f;
    std::memset(&out, 0, sizeof(out));
    out.baseColorTexture = out.metallicRoughnessTexture = out.emissiveTexture = out.normalmap = HRC_NULL_TEXTURE;
    out.clearCoatTexture = out.clearCoatRoughnessTexture = out.clearCoatNormalmap = HRC_NULL_TEXTURE;
    out.multiscatterLUT = lut;
    for (int k = 0; k < 3; ++k) { out.baseColor[k] = saturate(in.baseColor[k]); out.emissiveColor[k] = saturate(in.emissiveColor[k]); }
    out.metallic = saturate(in.metallic);
    out.roughness = std::max(saturate(in.roughness), kMinRoughness);
    out.specularF0 = in.specularF0 * kMaxSpecularF0;
    out.roughnessAlpha = out.roughness * out.roughness;
    out.clearCoat = in.clearCoat * kMaxClearcoat;
    out.clearCoatRoughness = std::max(saturate(in.clearCoatRoughness), kMinRoughness);
    out.clearCoatRoughnessAlpha = out.clearCoatRoughness * out.clearCoatRoughness;
    out.baseColorTexture = in.baseColorTexture; out.metallicRoughnessTexture = in.metallicRoughnessTexture;
    out.emissiveTexture = in.emissiveTexture; out.normalmap = in.normalmap; out.clearCoatTexture = in.clearCoatTexture;
    out.clearCoatRoughnessTexture = in.clearCoatRoughnessTexture; out.clearCoatNormalmap = in.clearCoatNormalmap;
    uint32_t flags = 0;                                               // feature #defines, PhysicallyBasedMaterial.cpp:49-113
    const bool force = in.forceEnableAllTextures != 0;
    if (in.baseColorTexture >= 0 || force) flags |= HRC_MAT_HAS_BASE_COLOR_TEXTURE;
    if (in.metallicRoughnessTexture >= 0 || force) flags |= HRC_MAT_HAS_METALLIC_ROUGHNESS_TEXTURE;
    if (in.emissiveTexture >= 0) flags |= HRC_MAT_HAS_EMISSIVE_TEXTURE;
    if (in.normalmap >= 0) flags |= HRC_MAT_HAS_NORMALMAP;
    if (in.clearCoatTexture >= 0 || force) flags |= HRC_MAT_HAS_CLEARCOAT_TEXTURE;
    if (in.clearCoatRoughnessTexture >= 0 || force) flags |= HRC_MAT_HAS_CLEARCOAT_ROUGHNESS_TEXTURE;
    if (in.clearCoatNormalmap >= 0) flags |= HRC_MAT_HAS_CLEARCOAT_NORMALMAP;
    if (in.doubleSided) flags |= HRC_MAT_DOUBLE_SIDED;
    if (in.alphaMask) flags |= HRC_MAT_ALPHA_MASK;
    if (in.enableVertexColors) flags |= HRC_MAT_VERTEX_COLORS;
    return flags;
}

uint32_t packGlassMaterial(const HrhGlassMaterial& in, HrcGlassParams& out) {              // GlassMaterial.cpp:88-126
    std::memset(&out, 0, sizeof(out));
    out.baseColorTexture = out.normalmap = out.metallicRoughnessTexture = HRC_NULL_TEXTURE;
    for (int k = 0; k < 3; ++k) out.baseColor[k] = saturate(in.baseColor[k]);
    out.roughness = std::max(saturate(in.roughness), 0.01f);
    out.roughnessAlpha = out.roughness * out.roughness;
    out.density = in.density;
    out.ior = std::max(0.0f, in.ior);
    const float f0 = std::fabs((1.0f - out.ior) / (1.0f + out.ior));
    out.specularF0 = f0 * f0;
    out.baseColorTexture = in.baseColorTexture; out.normalmap = in.normalmap; out.metallicRoughnessTexture = in.metallicRoughnessTexture;
    uint32_t flags = 0;                                               // GlassMaterial.cpp:40-74
    const bool force = in.forceEnableAllTextures != 0;
    if (in.baseColorTexture >= 0 || force) flags |= HRC_MAT_HAS_BASE_COLOR_TEXTURE;
    if (in.metallicRoughnessTexture >= 0 || force) flags |= HRC_MAT_HAS_METALLIC_ROUGHNESS_TEXTURE;
    if (in.normalmap >= 0) flags |= HRC_MAT_HAS_NORMALMAP;
    return flags;
}

void generateMultiscatterLUT(float* out, int dim, int sampleCount, int threads) {           // MultiScatterUtil.cpp:91-139
    std::vector<float> seq(2 * (size_t)sampleCount);
    sobol(seq.data(), (uint32_t)sampleCount, 0);
    if (threads <= 0) threads = (int)std::thread::hardware_concurrency();
    threads = std::max(1, std::min(threads, dim));
    auto rows = [&](int r0, int r1) {
        for (int row = r0; row < r1; ++row) {
            const float roughness = saturate(((float)row + 0.5f) / float(dim));
            const float alpha = roughness * roughness;
            for (int col = 0; col < dim; ++col) {
                const float NdotV = saturate(((float)col + 0.5f) / float(dim));
                const float v = lutValue(NdotV, alpha, seq);
                out[(size_t)row * dim + col] = (1.0f - v) / v;
            }
        }
    };
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t) pool.emplace_back(rows, dim * t / threads, dim * (t + 1) / threads);
    for (auto& t : pool) t.join();
}

} // namespace hr

extern "C" {

HRC_API uint32_t hrhPackPbrMaterial(const HrhPbrMaterial* in, HrcTexture lut, HrcPbrParams* out) { return hr::packPbrMaterial(*in, lut, *out); }
HRC_API uint32_t hrhPackGlassMaterial(const HrhGlassMaterial* in, HrcGlassParams* out) { return hr::packGlassMaterial(*in, *out); }
HRC_API void hrhGenerateMultiscatterLUT(float* out, int32_t dim, int32_t sampleCount, int32_t threads) { hr::generateMultiscatterLUT(out, dim, sampleCount, threads); }

HRC_API void hrhPackDirectionalLight(const float color[3], float illuminance, float theta, float phi, HrcDirLight* out) {
    float f[3]; hr::orientationForward(theta, phi, f);                 // DirectionalLight.cpp:64-78: direction TO the light
    for (int k = 0; k < 3; ++k) { out->direction[k] = f[k] * -1.0f; out->color[k] = color[k] * (illuminance * hr::kLumensToWatts); }
}
HRC_API void hrhPackPointLight(const float color[3], float luminousIntensity, const float position[3], HrcPointLight* out) {
    const float watts = (luminousIntensity * hr::kLumensToWatts) * (4.0f * hr::kPi);   // PointLight.cpp:41-50
    for (int k = 0; k < 3; ++k) { out->position[k] = position[k]; out->color[k] = color[k] * watts; }
}
HRC_API void hrhPackSpotLight(const float color[3], float luminousIntensity, const float position[3], float theta, float phi,
                              float innerAngle, float outerAngle, HrcSpotLight* out) {
    // SpotLight::setParams range fix-up (SpotLight.cpp:58-69)
    const float oneDegree = 0.01745329251994329576923690768489f;
    if (innerAngle > outerAngle) innerAngle = std::max(0.0f, outerAngle - oneDegree);
    if (innerAngle > 0.0f && innerAngle == outerAngle) innerAngle -= oneDegree;
    float f[3]; hr::orientationForward(theta, phi, f);                 // SpotLight.cpp:76-91: direction FROM the light
    const float watts = (luminousIntensity * hr::kLumensToWatts) * hr::kPi;              // SpotLight.cpp:44-56
    for (int k = 0; k < 3; ++k) { out->position[k] = position[k]; out->direction[k] = f[k]; out->color[k] = color[k] * watts; }
    out->angles[0] = cosf(innerAngle); out->angles[1] = cosf(outerAngle);
}
HRC_API float hrhEnvironmentExposure(float ev) { return powf(2.0f, ev); }               // EnvironmentLight.cpp:95

HRC_API void hrhOrbitCameraViewMatrix(float distance, float phi, float theta, const float target[3], float out[16]) {   // OrbitCamera.h:32-45
    const hr::Quat o = hr::mul(hr::angleAxis(theta, 1.0f, 0.0f, 0.0f), hr::angleAxis(phi, 0.0f, 1.0f, 0.0f));
    float r[9]; hr::mat3Cast(hr::inverse(o), r);
    const float t[3] = {target[0] + 0.0f, target[1] + 0.0f, target[2] + distance};
    // mat4_cast(q) * translate(I, t): rotation columns, then R * t as the translation column (glm mat4*mat4 order)
    for (int c = 0; c < 3; ++c) { out[4 * c + 0] = r[3 * c + 0]; out[4 * c + 1] = r[3 * c + 1]; out[4 * c + 2] = r[3 * c + 2]; out[4 * c + 3] = 0.0f; }
    for (int k = 0; k < 3; ++k) out[12 + k] = ((r[k] * t[0] + r[3 + k] * t[1]) + r[6 + k] * t[2]) + 0.0f;
    out[15] = 1.0f;
}
HRC_API float hrhFovTan(float focalLengthMm) {                                           // PassGenerator.cpp:341-343,357
    const float fovY = 2.0f * std::atan2(24.0f, 2.0f * focalLengthMm);
    return tanf(fovY * 0.5f);
}
HRC_API float hrhApertureRadius(float focalLengthMm, float fstop) { return (focalLengthMm / fstop) / 1000.0f; }   // PassGenerator.h:92-94

HRC_API void hrhPlaneMesh(float width, float length, float* positions, float* normals, float* uvs, uint32_t* indices) {   // PlaneMeshProvider.h:40-95
    const float sx = width * 0.5f, sz = length * 0.5f;
    const float corners[4][3] = {{-1.0f, 0.0f, 1.0f}, {1.0f, 0.0f, 1.0f}, {1.0f, 0.0f, -1.0f}, {-1.0f, 0.0f, -1.0f}};
    const float uv[4][2] = {{-1.0f, -1.0f}, {1.0f, -1.0f}, {1.0f, 1.0f}, {-1.0f, 1.0f}};
    const uint32_t strip[4] = {0, 1, 3, 2};
    for (int v = 0; v < 4; ++v) {
        if (positions) { positions[3 * v] = corners[v][0] * sx; positions[3 * v + 1] = corners[v][1] * 0.0f; positions[3 * v + 2] = corners[v][2] * sz; }
        if (normals) { normals[3 * v] = 0.0f; normals[3 * v + 1] = 1.0f; normals[3 * v + 2] = 0.0f; }
        if (uvs) { uvs[2 * v] = uv[v][0]; uvs[2 * v + 1] = uv[v][1]; }
        if (indices) indices[v] = strip[v];
    }
}

HRC_API void hrhSphereMesh(int32_t uSlices, int32_t vSlices, float radius, float* positions, float* normals, float* uvs,
                           uint32_t* indices, uint32_t* vertexCount, uint32_t* indexCount) {                              // SphereMeshProvider.h:17-131
    const int vSteps = vSlices + 2;
    if (vertexCount) *vertexCount = (uint32_t)((uSlices + 1) * vSteps);
    if (indexCount) *indexCount = (uint32_t)(3 * 2 * uSlices * vSlices);
    const float twoPi = 6.28318530717958647692f;
    size_t vtx = 0;
    if (positions || normals || uvs) {
        for (int ii = 0; ii < uSlices + 1; ++ii) {
            const float u = (float)ii / (float)uSlices;
            for (int jj = 0; jj < vSteps; ++jj, ++vtx) {
                const float v = (float)jj / (float)(vSlices + 1);
                const float az = float(u * twoPi), pol = float(v * hr::kPi);
                // CartesianFromSpherical (SphereMeshProvider.h:17-22); std::cos/sin float overloads
                const float px = radius * std::cos(az) * std::sin(pol);
                const float py = radius * std::cos(pol);
                const float pz = radius * std::sin(az) * std::sin(-pol);
                if (positions) { positions[3 * vtx] = px; positions[3 * vtx + 1] = py; positions[3 * vtx + 2] = pz; }
                if (normals) {
                    const float inv = 1.0f / std::sqrt((px * px + py * py) + pz * pz);
                    normals[3 * vtx] = px * inv; normals[3 * vtx + 1] = py * inv; normals[3 * vtx + 2] = pz * inv;
                }
                if (uvs) { uvs[2 * vtx] = u; uvs[2 * vtx + 1] = 1.f - v; }
            }
        }
    }
    if (indices) {
        uint32_t* p = indices;
        for (int ii = 0; ii < uSlices; ++ii) {
            for (int jj = 0; jj < vSteps - 1; ++jj) {
                const uint32_t a = (uint32_t)(ii * vSteps + jj), b = (uint32_t)((ii + 1) * vSteps + jj);
                if (jj == 0) { *p++ = a; *p++ = a + 1; *p++ = b + 1; }                       // first pole: one triangle
                else if (jj == vSteps - 2) { *p++ = b; *p++ = a; *p++ = a + 1; }              // second pole
                else { *p++ = a; *p++ = a + 1; *p++ = b + 1; *p++ = b + 1; *p++ = b; *p++ = a; }
            }
        }
    }
}

} // extern "C"

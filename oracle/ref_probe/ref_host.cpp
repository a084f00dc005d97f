#include <glm/glm/gtx/euler_angles.hpp>
#include <algorithm>
// Host-side math of the reference evaluated with the reference's own glm: OrbitCamera::createViewMatrix from the
// UNMODIFIED header (Source/HeatrayRenderer/OrbitCamera.h) and the light-direction expressions of
// DirectionalLight.cpp:64-78 / SpotLight.cpp:76-91 written with the same glm calls (the light classes themselves
// cannot link without OpenRL). Output: oracle/_ref/libref_host.so. Test infrastructure.
#include <glm/glm/glm.hpp>
#include <glm/glm/gtc/quaternion.hpp>
#include <glm/glm/gtc/matrix_transform.hpp>
#include <glm/glm/gtx/quaternion.hpp>
#include <HeatrayRenderer/OrbitCamera.h>
#include <Utility/AABB.h>
extern "C" {
// HeatrayRenderer::SceneTransform::transform() (HeatrayRenderer.h:343-346), same glm calls.
void ref_scene_transform(float yaw, float pitch, float roll, float scale, float* out16) {
    glm::mat4 m = glm::scale(glm::mat4(1.0f), glm::vec3(scale)) * glm::yawPitchRoll(yaw, pitch, roll);
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) out16[4 * i + j] = m[i][j];
}
// util::AABB (unmodified header): center(), radius(), valid(), bottom() with a transform set -> out = cx, cy, cz, radius, valid, bottom.
void ref_aabb_camera(const float* mn, const float* mx, const float* in16, float* out5) {
    util::AABB b; b.min = glm::vec3(mn[0], mn[1], mn[2]); b.max = glm::vec3(mx[0], mx[1], mx[2]);
    for (int c = 0; c < 4; ++c) for (int r = 0; r < 4; ++r) b.transform[c][r] = in16[4 * c + r];
    glm::vec3 c = b.center(); out5[0] = c.x; out5[1] = c.y; out5[2] = c.z; out5[3] = b.radius(); out5[4] = b.valid() ? 1.0f : 0.0f; out5[5] = b.bottom();
}
void ref_orbit_view(float distance, float phi, float theta, const float* target, float* out16) {
    OrbitCamera c; c.distance = distance; c.phi = phi; c.theta = theta; c.target = glm::vec3(target[0], target[1], target[2]);
    glm::mat4 m = c.createViewMatrix();
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) out16[4 * i + j] = m[i][j];
}
void ref_light_forward(float theta, float phi, float* out3) {   // normalize(viewMatrix[2].xyz * -1)
    constexpr glm::vec3 right = glm::vec3(1.0f, 0.0f, 0.0f);
    constexpr glm::vec3 up = glm::vec3(0.0f, 1.0f, 0.0f);
    glm::quat rotTheta = glm::angleAxis(theta, right);
    glm::quat rotPhi = glm::angleAxis(phi, up);
    glm::quat orientation = rotTheta * rotPhi;
    const glm::mat4 viewMatrix = glm::mat4_cast(glm::inverse(orientation));
    glm::vec3 direction = glm::vec3(viewMatrix[2]) * -1.0f;
    glm::vec3 n = glm::normalize(direction);
    out3[0] = n.x; out3[1] = n.y; out3[2] = n.z;
}
// ProcessLight's orientation recovery (AssimpMeshProvider.cpp:564-586): local frame with vec4(axis, 1) columns, lightTransform * local,
// glm::extractEulerAngleXYZ, clamps. in16 = column-major node transform; up/dir as Assimp's glTF importer sets them. out = (theta, phi).
void ref_light_angles(const float* in16, float* out2) {
    glm::mat4 lightTransform; for (int c = 0; c < 4; ++c) for (int r = 0; r < 4; ++r) lightTransform[c][r] = in16[4 * c + r];
    glm::vec4 yAxis = glm::vec4(0.0f, 1.0f, 0.0f, 1.0f);
    glm::vec4 zAxis = glm::vec4(0.0f, 0.0f, -1.0f, 1.0f);
    glm::vec4 xAxis = glm::vec4(glm::cross(glm::vec3(yAxis), glm::vec3(zAxis)), 1.0f);
    glm::mat4x4 localTransform = glm::mat4x4(xAxis, yAxis, zAxis, glm::vec4(0.0f, 0.0f, 0.0f, 1.0f));
    glm::mat4x4 finalTransform = lightTransform * localTransform;
    float roll = 0.0f, pitch = 0.0f, yaw = 0.0f;
    glm::extractEulerAngleXYZ(finalTransform, pitch, yaw, roll);
    yaw = std::clamp(yaw, 0.0f, glm::two_pi<float>());
    pitch = std::clamp(pitch, -glm::half_pi<float>(), glm::half_pi<float>());
    out2[0] = pitch; out2[1] = yaw;
}
}

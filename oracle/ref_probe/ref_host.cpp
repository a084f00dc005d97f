// Host-side math of the reference evaluated with the reference's own glm: OrbitCamera::createViewMatrix from the
// UNMODIFIED header (Source/HeatrayRenderer/OrbitCamera.h) and the light-direction expressions of
// DirectionalLight.cpp:64-78 / SpotLight.cpp:76-91 written with the same glm calls (the light classes themselves
// cannot link without OpenRL). Output: oracle/_ref/libref_host.so. Test infrastructure.
#include <glm/glm/glm.hpp>
#include <glm/glm/gtc/quaternion.hpp>
#include <glm/glm/gtc/matrix_transform.hpp>
#include <glm/glm/gtx/quaternion.hpp>
#include <HeatrayRenderer/OrbitCamera.h>
extern "C" {
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
}

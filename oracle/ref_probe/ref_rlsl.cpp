// utility.rlsl + brdfs.rlsl of the reference compiled as C++ through glm.
// gen_rlsl_includes.sh writes the two shader files into /tmp with (a) their #include lines stripped exactly
// as ShaderCodeLoader.cpp:34-43 resolves them and (b) `dot(st.xy,` -> `dot(st,` at utility.rlsl:17 (glm swizzle
// proxies do not deduce in templates). Compiled with -fsingle-precision-constant. Output: oracle/_ref/libref_rlsl.so.
#define GLM_FORCE_SWIZZLE
#include <glm/glm/glm.hpp>
using namespace glm;
static float atan(float y, float x) { return atan2f(y, x); }
#include "/tmp/ref_rlsl_utility.inc"
#include "/tmp/ref_rlsl_brdfs.inc"
extern "C" void ref_rlsl_probe(const float* in, float* out) {
    // same packing as orcProbeHelpers (oracle/oracle.cpp)
    vec3 n = normalize(vec3(in[0], in[1], in[2]));
    mat3 f = orthonormalFrame(n);
    out[0] = f[0].x; out[1] = f[0].y; out[2] = f[0].z; out[3] = f[2].x; out[4] = f[2].y; out[5] = f[2].z;
    vec3 cw = cosineWeightedSample(in[3], in[4]); out[6] = cw.x; out[7] = cw.y; out[8] = cw.z;
    vec3 vg = sampleVisibleGGX(normalize(vec3(in[5], in[6], in[7])), in[3], in[4], in[8]); out[9] = vg.x; out[10] = vg.y; out[11] = vg.z;
    out[12] = D_GGX(in[9], in[8]); out[13] = G1_Smith_GGX(in[10], in[8]); out[14] = G2_Smith_GGX(in[10], in[11], in[8]);
    out[15] = F_Fresnel(in[12], in[13]); out[16] = F_Schlick(in[14], in[15]); out[17] = random(vec2(in[16], in[17]));
}

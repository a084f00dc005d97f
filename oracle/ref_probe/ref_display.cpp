// displayGL.frag of the reference compiled as C++ through glm (the resolve + colour pipeline).
// build_ref.sh writes the shader into /tmp with `#version` dropped, `uniform`/`in`/`out` globals turned into statics and
// `main` renamed; nothing else is touched. texture() returns the one texel under test. -fsingle-precision-constant.
// Output: oracle/_ref/libref_display.so.
#define GLM_FORCE_SWIZZLE
#include <glm/glm/glm.hpp>
using namespace glm;
static vec4 g_texel;
struct sampler2D;
static vec4 texture(const sampler2D&, vec2) { return g_texel; }
#include "/tmp/ref_display.inc"
// params: tonemappingEnabled, cameraExposure, brightness, contrast, hue, saturation, vibrance, red, green, blue, vignetteIntensity, vignetteFalloff
extern "C" void ref_display(const float* texel, const float* uv, const float* p, float* out) {
    g_texel = vec4(texel[0], texel[1], texel[2], texel[3]);
    textureCoords = vec2(uv[0], uv[1]);
    tonemappingEnabled = (int)p[0]; cameraExposure = p[1]; brightness = p[2]; contrast = p[3]; hue = p[4]; saturation = p[5]; vibrance = p[6];
    red = p[7]; green = p[8]; blue = p[9]; vignetteIntensity = p[10]; vignetteFalloff = p[11];
    display_main();
    out[0] = fragColor.x; out[1] = fragColor.y; out[2] = fragColor.z; out[3] = fragColor.w;
}

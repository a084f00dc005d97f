// resolve.cu — the resolve + colour pipeline of the viewer as one CUDA kernel over the accumulation buffer.
//
// Reference: Resources/shaders/displayGL.frag:72-151 (main), :28-70 (ACES fit, sRGB helpers), drawn as a window-sized quad
// by HeatrayRenderer.h:228-262 with texture coordinates 0..1 across the render area (displayGL.vert:5-24), and the HDR
// screenshot loop of HeatrayRenderer::saveScreenshot (HeatrayRenderer.cpp:1629-1643). No graphics API is involved here:
// the "texture" is the accumulation buffer itself (NEAREST, 1:1), the "framebuffer" a plain device array.
//
// HBM-bound streaming kernel: 16 B read per pixel, 3 / 12 / 16 B written; each thread owns four consecutive pixels so
// the 8-bit output is written as three 32-bit words.
#include "device_types.h"

namespace hr {
namespace {

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }

__device__ __forceinline__ float linearToSRGB1(float c) {                       // displayGL.frag:48-59
    return (c <= 0.0031308f) ? 12.92f * c : 1.055f * powf(c, 1.0f / 2.4f) - 0.055f;
}
__device__ __forceinline__ float srgbToLinear1(float c) {                       // :61-72
    return (c <= 0.04045f) ? c / 12.92f : powf((c + 0.055f) / (1.0f + 0.055f), 2.4f);
}
__device__ __forceinline__ float rrtAndOdtFit1(float v) {                       // :41-46
    const float a = v * (v + 0.0245786f) - 0.000090537f;
    const float b = v * (0.983729f * v + 0.4329510f) + 0.238081f;
    return a / b;
}
__device__ __forceinline__ float clamp01(float x) { return fminf(fmaxf(x, 0.0f), 1.0f); }
__device__ __forceinline__ float mix1(float x, float y, float a) { return x * (1.0f - a) + y * a; }
__device__ __forceinline__ float fract1(float x) { return x - floorf(x); }

// One pixel of displayGL.frag main(): `t` is the accumulation texel, (u, v) the texture coordinate of the fragment.
__device__ __forceinline__ V3 displayPixel(const HrcPostProcess& p, float4 t, float u, float v) {
    V3 c = v3(t.x / t.w, t.y / t.w, t.z / t.w);                                 // :77-78
    if (p.tonemappingEnabled == 1) {                                            // :82-90
        c = v3(linearToSRGB1(c.x), linearToSRGB1(c.y), linearToSRGB1(c.z));
        V3 a = v3(0.59719f * c.x + 0.35458f * c.y + 0.04823f * c.z,            // ACESInputMat (column-major constructor)
                  0.07600f * c.x + 0.90834f * c.y + 0.01566f * c.z,
                  0.02840f * c.x + 0.13383f * c.y + 0.83777f * c.z);
        a = v3(rrtAndOdtFit1(a.x), rrtAndOdtFit1(a.y), rrtAndOdtFit1(a.z));
        c = v3(1.60475f * a.x + -0.53108f * a.y + -0.07367f * a.z,             // ACESOutputMat
               -0.10208f * a.x + 1.10813f * a.y + -0.00605f * a.z,
               -0.00327f * a.x + -0.07276f * a.y + 1.07602f * a.z);
        c = v3(clamp01(c.x), clamp01(c.y), clamp01(c.z));
        c = v3(srgbToLinear1(c.x), srgbToLinear1(c.y), srgbToLinear1(c.z));
    }
    c = v3((c.x - 0.5f) * p.contrast + 0.5f + p.brightness,                     // :95-97
           (c.y - 0.5f) * p.contrast + 0.5f + p.brightness,
           (c.z - 0.5f) * p.contrast + 0.5f + p.brightness);
    float h, s, val;
    {                                                                           // RGB -> HSV, :103-112
        const float kx = 0.0f, ky = -1.0f / 3.0f, kz = 2.0f / 3.0f, kw = -1.0f;
        const float st1 = (c.y < c.z) ? 0.0f : 1.0f;                            // step(b, g)
        const float px = mix1(c.z, c.y, st1), py = mix1(c.y, c.z, st1), pz = mix1(kw, kx, st1), pw = mix1(kz, ky, st1);
        const float st2 = (c.x < px) ? 0.0f : 1.0f;                             // step(p.x, r)
        const float qx = mix1(px, c.x, st2), qy = mix1(py, py, st2), qz = mix1(pw, pz, st2), qw = mix1(c.x, px, st2);
        const float d = qx - fminf(qw, qy);
        const float e = 1.0e-10f;
        h = fabsf(qz + (qw - qy) / (6.0f * d + e)); s = d / (qx + e); val = qx;
    }
    h *= p.hue; s *= p.saturation;                                              // :114-115
    s *= 1.0f + sqrtf(s) * p.vibrance;                                          // :117-121
    {                                                                           // HSV -> RGB, :124-128
        const float px = fabsf(fract1(h + 1.0f) * 6.0f - 3.0f), py = fabsf(fract1(h + 2.0f / 3.0f) * 6.0f - 3.0f), pz = fabsf(fract1(h + 1.0f / 3.0f) * 6.0f - 3.0f);
        c = v3(val * mix1(1.0f, clamp01(px - 1.0f), s), val * mix1(1.0f, clamp01(py - 1.0f), s), val * mix1(1.0f, clamp01(pz - 1.0f), s));
    }
    c = v3(c.x * p.red, c.y * p.green, c.z * p.blue);                           // :132-136
    {                                                                           // vignette, :139-143 (`+ blue` is the shader's own)
        const float dx = 0.5f - u, dy = 0.5f - v;
        const float dist = sqrtf(dx * dx + dy * dy);
        const float e0 = 0.8f, e1 = p.vignetteFalloff * 0.799f;
        const float x = dist * (p.vignetteIntensity + p.blue);
        const float tt = clamp01((x - e0) / (e1 - e0));
        const float sm = tt * tt * (3.0f - 2.0f * tt);
        c = v3(c.x * sm, c.y * sm, c.z * sm);
    }
    c = v3(c.x * p.cameraExposure, c.y * p.cameraExposure, c.z * p.cameraExposure);   // :146
    return v3(linearToSRGB1(c.x), linearToSRGB1(c.y), linearToSRGB1(c.z));      // :149
}

__device__ __forceinline__ uint32_t unorm8(float c) {                           // float -> 8-bit unorm colour buffer; NaN -> 0
    return (uint32_t)__float2int_rn(clamp01(c) * 255.0f);
}

template <int FORMAT>
__global__ void __launch_bounds__(256) k_resolve(const float4* __restrict__ accum, const __grid_constant__ HrcPostProcess p, int w, int h, void* __restrict__ out) {
    const size_t n = (size_t)w * h;
    const size_t quad = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t first = quad * 4;
    if (first >= n) return;
    V3 c[4];
    const int cnt = (int)min((size_t)4, n - first);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (k < cnt) {
            const size_t i = first + k;
            const float4 t = __ldcs(accum + i);
            if (FORMAT == HRC_RESOLVE_HDR_RGB32F) {
                const float divisor = 1.0f / t.w;                               // HeatrayRenderer.cpp:1634-1638
                c[k] = v3(t.x * divisor, t.y * divisor, t.z * divisor);
            } else {
                const int x = (int)(i % (size_t)w), y = (int)(i / (size_t)w);
                c[k] = displayPixel(p, t, ((float)x + 0.5f) / (float)w, ((float)y + 0.5f) / (float)h);
            }
        } else c[k] = v3(0.f, 0.f, 0.f);
    }
    if (FORMAT == HRC_RESOLVE_DISPLAY_RGBA32F) {
        float4* o = (float4*)out + first;
#pragma unroll
        for (int k = 0; k < 4; ++k) if (k < cnt) __stcs(o + k, make_float4(c[k].x, c[k].y, c[k].z, 1.0f));
    } else if (FORMAT == HRC_RESOLVE_HDR_RGB32F) {
        float* o = (float*)out + first * 3;
        if (cnt == 4) {
            float4* o4 = (float4*)o;                                            // 48 B per thread, 16-byte aligned (first % 4 == 0)
            __stcs(o4 + 0, make_float4(c[0].x, c[0].y, c[0].z, c[1].x));
            __stcs(o4 + 1, make_float4(c[1].y, c[1].z, c[2].x, c[2].y));
            __stcs(o4 + 2, make_float4(c[2].z, c[3].x, c[3].y, c[3].z));
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) if (k < cnt) { o[3 * k] = c[k].x; o[3 * k + 1] = c[k].y; o[3 * k + 2] = c[k].z; }
        }
    } else {
        uint8_t* o = (uint8_t*)out + first * 3;
        uint32_t b[12];
#pragma unroll
        for (int k = 0; k < 4; ++k) { b[3 * k] = unorm8(c[k].x); b[3 * k + 1] = unorm8(c[k].y); b[3 * k + 2] = unorm8(c[k].z); }
        if (cnt == 4) {
            uint32_t* o32 = (uint32_t*)o;                                       // 12 B per thread, 4-byte aligned
            o32[0] = b[0] | (b[1] << 8) | (b[2] << 16) | (b[3] << 24);
            o32[1] = b[4] | (b[5] << 8) | (b[6] << 16) | (b[7] << 24);
            o32[2] = b[8] | (b[9] << 8) | (b[10] << 16) | (b[11] << 24);
        } else {
#pragma unroll
            for (int k = 0; k < 12; ++k) if (k < 3 * cnt) o[k] = (uint8_t)b[k];
        }
    }
}

} // namespace

void launchResolve(const float4* accum, const HrcPostProcess& post, int format, int w, int h, void* out, cudaStream_t s) {
    const size_t quads = ((size_t)w * h + 3) / 4;
    const unsigned grid = (unsigned)((quads + 255) / 256);
    if (!grid) return;
    if (format == HRC_RESOLVE_DISPLAY_RGBA32F)   k_resolve<HRC_RESOLVE_DISPLAY_RGBA32F><<<grid, 256, 0, s>>>(accum, post, w, h, out);
    else if (format == HRC_RESOLVE_DISPLAY_RGB8) k_resolve<HRC_RESOLVE_DISPLAY_RGB8><<<grid, 256, 0, s>>>(accum, post, w, h, out);
    else                                         k_resolve<HRC_RESOLVE_HDR_RGB32F><<<grid, 256, 0, s>>>(accum, post, w, h, out);
}

} // namespace hr

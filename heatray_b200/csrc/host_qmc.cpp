// host_qmc.cpp — host-side sample-table generators of the pass interface.
//
// Restates what the reference's PassGenerator builds on the CPU before a render starts:
//   util::sobol / halton / hammersley / blueNoise / uniformRandomFloats / radialSobol /
//   randomPolygonal            (reference Source/Utility/Random.h:26-355, BlueNoise.h:52-101, Hash.h:17-36)
//   generateRandomSequences    (Source/HeatrayRenderer/PassGenerator.cpp:603-684)
//   generateSequenceOffsets    (PassGenerator.cpp:150-159)
//   per-pixel sequence id      (Resources/shaders/perspective.rlsl:62 + utility.rlsl:15-18; quirk B-Q2)
// The tables are tiny and built once per reset, so they stay on the host; the backend receives
// them through hrcSetSequences()/hrcSetPixelTables().
//
// Bit-exactness against the reference generators is checked by tests/test_qmc.py against
// tests/golden/qmc_ref.json (tables written by the reference's own Random.h compiled in place).

#include "host_api.h"

#include <algorithm>
#include <array>
#include <cmath>
#include <random>
#include <vector>
#include <cstdint>
#include <cstring>
#include <limits>
#include <random>
#include <vector>

namespace hr {
namespace {

// ---- integer hashing / scrambling (Random.h:36-78) ------------------------------------
inline uint32_t hashBurley(uint32_t v) {
    v ^= v >> 16; v *= 0x85ebca6bu;
    v ^= v >> 13; v *= 0xc2b2ae35u;
    v ^= v >> 16;
    return v;
}
inline uint32_t hashCombineBurley(uint32_t seed, uint32_t v) { return seed ^ (v + (seed << 6) + (seed >> 2)); }
inline uint32_t permuteLaineKarras(uint32_t v, uint32_t seed) {
    v += seed;
    v ^= v * 0x6c50b47cu; v ^= v * 0xb82f1e52u; v ^= v * 0xc7afe638u; v ^= v * 0x8d22f6e6u;
    return v;
}
inline uint32_t bitReverse(uint32_t v) {
    v = (v << 16) | (v >> 16);
    v = ((v & 0x55555555u) << 1) | ((v & 0xAAAAAAAAu) >> 1);
    v = ((v & 0x33333333u) << 2) | ((v & 0xCCCCCCCCu) >> 2);
    v = ((v & 0x0F0F0F0Fu) << 4) | ((v & 0xF0F0F0F0u) >> 4);
    v = ((v & 0x00FF00FFu) << 8) | ((v & 0xFF00FF00u) >> 8);
    return v;
}
inline uint32_t scrambleNested(uint32_t v, uint32_t seed) { return bitReverse(permuteLaineKarras(bitReverse(v), seed)); }

// Random.h:26-34. float(UINT32_MAX) rounds to 2^32, so 1.0f maps past the range; the reference's
// conversion is UB there and x86-64 gcc yields the low 32 bits of the 64-bit truncation (quirk B-Q4).
inline uint32_t unitToBits(float f) { return (uint32_t)(uint64_t)(int64_t)(f * 4294967296.0f); }
inline float bitsToUnit(uint32_t u) { return float(u) * (1.0f / 4294967296.0f); }

// owenScrambleSequence (Random.h:85-108): shuffle the index, then Owen-scramble each dimension.
template <class Gen>
void scrambledSequence(float* xy, uint32_t count, uint32_t sequenceIndex, Gen gen) {
    const uint32_t seed = hashBurley(sequenceIndex + 1);
    const uint32_t seedX = hashCombineBurley(seed, 0), seedY = hashCombineBurley(seed, 1);
    for (uint32_t i = 0; i < count; ++i) {
        float sx, sy;
        gen(scrambleNested(i, seed), i, sx, sy);
        xy[2 * i + 0] = bitsToUnit(scrambleNested(unitToBits(sx), seedX));
        xy[2 * i + 1] = bitsToUnit(scrambleNested(unitToBits(sy), seedY));
    }
}

// Sobol' dimension 0 is the bit-reversed index; dimension 1 uses the classic direction numbers,
// generated here by the recurrence v[k] = v[k-1] ^ (v[k-1] >> 1) (equals the table at Random.h:226-244).
struct SobolDirections {
    uint32_t d1[32];
    SobolDirections() { uint32_t v = 0x80000000u; for (int k = 0; k < 32; ++k) { d1[k] = v; v ^= v >> 1; } }
};
const SobolDirections kSobol;

inline void sobolPoint(uint32_t index, float& x, float& y) {
    uint32_t a = 0, b = 0;
    for (uint32_t bit = 0; bit < 32; ++bit) {
        if ((index >> bit) & 1u) { a ^= 0x80000000u >> bit; b ^= kSobol.d1[bit]; }
    }
    x = bitsToUnit(a); y = bitsToUnit(b);
}

inline float radicalInverse(uint32_t index, int base) {   // Random.h:193-205
    float result = 0.0f, f = 1.0f, denom = float(base);
    uint32_t n = index;
    while (n > 0) { f = f / denom; result += f * float(n % (uint32_t)base); n = n / (uint32_t)base; }
    return result;
}

// FNV-1a over the object bytes read as (signed) char, as Hash.h:17-30 does.
inline uint64_t fnv1a(const void* p, size_t n) {
    uint64_t h = 0xcbf29ce484222325ull;
    const signed char* b = (const signed char*)p;
    for (size_t i = 0; i < n; ++i) { h ^= (uint64_t)(int64_t)b[i]; h *= 0x100000001b3ull; }
    return h;
}
inline float fnvUnit(uint32_t seed) {                      // BlueNoise.h:97-100
    uint64_t a = fnv1a(&seed, sizeof(seed));
    uint64_t b = fnv1a(&a, sizeof(a));
    return float(b) / float(std::numeric_limits<uint64_t>::max());
}

} // namespace

void sobol(float* xy, uint32_t count, uint32_t sequenceIndex) {
    scrambledSequence(xy, count, sequenceIndex, [](uint32_t idx, uint32_t, float& x, float& y) { sobolPoint(idx, x, y); });
}

void halton(float* xy, uint32_t count, uint32_t sequenceIndex) {
    static const int bases[16][2] = {{2, 3}, {2, 5}, {2, 7}, {3, 7}, {4, 5}, {5, 7}, {5, 9}, {5, 11},
                                     {6, 11}, {5, 11}, {8, 11}, {3, 5}, {11, 15}, {2, 15}, {3, 19}, {7, 10}};   // Random.h:172-189
    const int bx = bases[sequenceIndex & 15][0], by = bases[sequenceIndex & 15][1];
    scrambledSequence(xy, count, sequenceIndex, [bx, by](uint32_t idx, uint32_t, float& x, float& y) {
        x = radicalInverse(idx, bx); y = radicalInverse(idx, by);
    });
}

void hammersley(float* xy, uint32_t count, uint32_t sequenceIndex) {   // Random.h:134-154
    const float divisor = 1.0f / float(count);
    scrambledSequence(xy, count, sequenceIndex, [divisor](uint32_t idx, uint32_t arrayIndex, float& x, float& y) {
        x = float(arrayIndex) * divisor;
        y = float(bitReverse(idx)) * 2.3283064365386963e-10f;
    });
}

void blueNoise(float* xy, uint32_t count, int sequenceIndex) {          // BlueNoise.h:52-101 (best candidate of 30)
    if (count == 0) return;
    int seed = int(fnv1a(&sequenceIndex, sizeof(sequenceIndex)));
    auto next = [&seed]() { return fnvUnit(uint32_t(seed++)); };
    std::vector<float> pts; pts.reserve(2 * (size_t)count);
    { float a = next(); float b = next(); pts.push_back(a); pts.push_back(b); }
    const float diag = std::sqrt(1.0f * 1.0f + 1.0f * 1.0f);
    for (uint32_t i = 1; i < count; ++i) {
        float bestDist = 0.0f, bx = 0.0f, by = 0.0f;
        for (int k = 0; k < 30; ++k) {
            float cx = next(); float cy = next();
            float nearest = diag;
            for (size_t p = 0; p + 1 < pts.size(); p += 2) {
                float dx = cx - pts[p], dy = cy - pts[p + 1];
                float d = std::sqrt(dx * dx + dy * dy);
                if (d < nearest) nearest = d;
            }
            if (nearest > bestDist) { bestDist = nearest; bx = cx; by = cy; }
        }
        pts.push_back(bx); pts.push_back(by);
    }
    std::memcpy(xy, pts.data(), sizeof(float) * 2 * (size_t)count);
}

void uniformRandom(float* xy, uint32_t count, uint32_t seed, float lo, float hi) {   // Random.h:113-130 (libstdc++ streams, quirk B-Q5)
    std::mt19937 gen; gen.seed(seed);
    std::uniform_real_distribution<float> dist(lo, hi);
    for (uint32_t i = 0; i < 2 * count; ++i) xy[i] = dist(gen);
}

void radialSobol(float* xy, uint32_t count, uint32_t sequenceIndex) {   // Random.h:268-289
    sobol(xy, count, sequenceIndex);
    const float twoPi = 6.28318530717958647692f;
    for (uint32_t i = 0; i < count; ++i) {
        float s = xy[2 * i], t = xy[2 * i + 1];
        float r = sqrtf(t), a = twoPi * s;
        xy[2 * i + 0] = (r * cosf(a) + 1.0f) * 0.5f;
        xy[2 * i + 1] = (r * sinf(a) + 1.0f) * 0.5f;
    }
}

void randomPolygonal(float* xy, uint32_t numEdges, uint32_t count, uint32_t seed) {   // Random.h:293-355
    const float twoPi = 6.28318530717958647692f;
    std::vector<float> vx(numEdges + 1), vy(numEdges + 1);
    const float step = twoPi / numEdges;
    for (uint32_t e = 0; e < numEdges; ++e) { float th = step * float(e); vx[e] = cosf(th); vy[e] = sinf(th); }
    vx[numEdges] = 0.0f; vy[numEdges] = 0.0f;
    std::mt19937 gen; gen.seed(seed);
    std::uniform_real_distribution<float> unit(0.0f, 1.0f);
    std::uniform_int_distribution<int> pick(0, int(numEdges) - 1);
    for (uint32_t i = 0; i < count; ++i) {
        int tri = pick(gen);
        float alpha, beta;
        do { alpha = unit(gen); beta = unit(gen); } while (alpha + beta > 1.0f);
        float gamma = 1.0f - (alpha + beta);
        uint32_t b = uint32_t(tri), c = (uint32_t(tri) + 1) % numEdges;
        // centre * alpha + v[b] * beta + v[c] * gamma, evaluated left to right like glm
        float px = (vx[numEdges] * alpha + vx[b] * beta) + vx[c] * gamma;
        float py = (vy[numEdges] * alpha + vy[b] * beta) + vy[c] * gamma;
        xy[2 * i + 0] = (px + 1.0f) * 0.5f;
        xy[2 * i + 1] = (py + 1.0f) * 0.5f;
    }
}

void generateRandomSequences(int sampleMode, int bokehShape, uint32_t numSequences, uint32_t sampleCount, float* xy, float* aperture) {
    for (uint32_t s = 0; s < numSequences; ++s) {
        float* dst = xy + 2 * (size_t)s * sampleCount;
        switch (sampleMode) {
            case HRH_SAMPLE_RANDOM:     uniformRandom(dst, sampleCount, s, 0.0f, 1.0f); break;
            case HRH_SAMPLE_HALTON:     halton(dst, sampleCount, s); break;
            case HRH_SAMPLE_HAMMERSLEY: hammersley(dst, sampleCount, s); break;
            case HRH_SAMPLE_BLUE_NOISE: blueNoise(dst, sampleCount, int(s)); break;
            default:                    sobol(dst, sampleCount, s); break;
        }
        float* ap = aperture + 2 * (size_t)s * sampleCount;
        switch (bokehShape) {
            case HRH_BOKEH_PENTAGON: randomPolygonal(ap, 5, sampleCount, s); break;
            case HRH_BOKEH_HEXAGON:  randomPolygonal(ap, 6, sampleCount, s); break;
            case HRH_BOKEH_OCTAGON:  randomPolygonal(ap, 8, sampleCount, s); break;
            default:                 radialSobol(ap, sampleCount, s); break;
        }
    }
}

// Number of SequenceOffsets entries the reference's index formula can touch (quirk B-Q1):
// index = int((y+0.5)*H + (x+0.5)); the largest is reached at the last pixel.
size_t sequenceOffsetCount(int w, int h) {
    size_t maxIdx = (size_t)(int)(((float)(h - 1) + 0.5f) * (float)h + ((float)(w - 1) + 0.5f));
    size_t n = (size_t)w * (size_t)h;
    return maxIdx + 1 > n ? maxIdx + 1 : n;
}

void generateSequenceOffsets(int w, int h, float* outX, size_t count) {
    std::vector<float> xy(2 * count);
    sobol(xy.data(), (uint32_t)count, 0);     // prefix-stable: the first w*h entries equal the reference's table
    for (size_t i = 0; i < count; ++i) outX[i] = xy[2 * i];
}

void generateSequenceIds(int w, int h, int numSequences, uint8_t* out) {
    for (int y = 0; y < h; ++y) {
        for (int x = 0; x < w; ++x) {
            float sx = ((float)x + 0.5f) / (float)w, sy = ((float)y + 0.5f) / (float)h;
            float d = sx * 12.9898f + sy * 78.233f;
            float v = sinf(d) * 43758.5453123f;
            float r = v - floorf(v);
            out[(size_t)y * w + x] = (uint8_t)(int)floorf(r * (float)numSequences);
        }
    }
}

// utility.rlsl:15-18 random(st), evaluated with the host's libm (the ground truth for every `sin`-hash decision, quirk B-Q2)
static inline float rlslRandom(float sx, float sy) {
    float d = sx * 12.9898f + sy * 78.233f;
    float v = sinf(d) * 43758.5453123f;
    return v - floorf(v);
}

void generateInteractiveBlockCoords(int blockW, int blockH, uint32_t seed, float* outRGB) {      // PassGenerator.cpp:266-294
    std::vector<std::array<float, 3>> coords;
    for (int row = 0; row < blockW; ++row)
        for (int col = 0; col < blockH; ++col) coords.push_back({(float)row, (float)col, 0.0f});
    std::mt19937 generator(seed);                                                                // reference: std::random_device (quirk B-Q6)
    std::shuffle(coords.begin(), coords.end(), generator);
    for (size_t i = 0; i < coords.size(); ++i) for (int k = 0; k < 3; ++k) outRGB[3 * i + k] = coords[i][k];
}

void generateInteractiveBlockRand(int blocksX, int blocksY, float* outXY) {                      // perspective.rlsl:44-46
    for (int by = 0; by < blocksY; ++by)
        for (int bx = 0; bx < blocksX; ++bx) {
            outXY[2 * ((size_t)by * blocksX + bx)] = rlslRandom((float)bx, (float)by);
            outXY[2 * ((size_t)by * blocksX + bx) + 1] = rlslRandom((float)by, (float)bx);
        }
}

} // namespace hr

// ---- flat C exports ---------------------------------------------------------------------
extern "C" {
HRC_API void hrhSobol(float* xy, uint32_t count, uint32_t seq) { hr::sobol(xy, count, seq); }
HRC_API void hrhHalton(float* xy, uint32_t count, uint32_t seq) { hr::halton(xy, count, seq); }
HRC_API void hrhHammersley(float* xy, uint32_t count, uint32_t seq) { hr::hammersley(xy, count, seq); }
HRC_API void hrhBlueNoise(float* xy, uint32_t count, int32_t seq) { hr::blueNoise(xy, count, seq); }
HRC_API void hrhUniformRandom(float* xy, uint32_t count, uint32_t seed, float lo, float hi) { hr::uniformRandom(xy, count, seed, lo, hi); }
HRC_API void hrhRadialSobol(float* xy, uint32_t count, uint32_t seq) { hr::radialSobol(xy, count, seq); }
HRC_API void hrhRandomPolygonal(float* xy, uint32_t edges, uint32_t count, uint32_t seed) { hr::randomPolygonal(xy, edges, count, seed); }
HRC_API void hrhGenerateRandomSequences(int32_t mode, int32_t bokeh, uint32_t numSeq, uint32_t count, float* xy, float* ap) {
    hr::generateRandomSequences(mode, bokeh, numSeq, count, xy, ap);
}
HRC_API size_t hrhSequenceOffsetCount(int32_t w, int32_t h) { return hr::sequenceOffsetCount(w, h); }
HRC_API void hrhGenerateSequenceOffsets(int32_t w, int32_t h, float* outX, size_t count) { hr::generateSequenceOffsets(w, h, outX, count); }
HRC_API void hrhGenerateInteractiveBlockCoords(int32_t bw, int32_t bh, uint32_t seed, float* out) { hr::generateInteractiveBlockCoords(bw, bh, seed, out); }
HRC_API void hrhInteractiveBlockCount(int32_t w, int32_t h, int32_t bw, int32_t bh, int32_t* bx, int32_t* by) { *bx = (w + bw - 1) / bw; *by = (h + bh - 1) / bh; }
HRC_API void hrhGenerateInteractiveBlockRand(int32_t bx, int32_t by, float* out) { hr::generateInteractiveBlockRand(bx, by, out); }
HRC_API void hrhGenerateSequenceIds(int32_t w, int32_t h, int32_t numSeq, uint8_t* out) { hr::generateSequenceIds(w, h, numSeq, out); }
}

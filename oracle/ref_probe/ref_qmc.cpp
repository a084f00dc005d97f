// Thin C exports around the UNMODIFIED reference generators (Source/Utility/Random.h), compiled in
// place from /root/reference. Output: oracle/_ref/libref_qmc.so (git-ignored). Test infrastructure.
#include <Utility/Random.h>
extern "C" {
void ref_sobol(float* xy, uint32_t n, uint32_t s) { util::sobol((glm::vec2*)xy, n, s); }
void ref_halton(float* xy, uint32_t n, int s) { util::halton((glm::vec2*)xy, n, s); }
void ref_hammersley(float* xy, uint32_t n, int s) { util::hammersley((glm::vec2*)xy, n, s); }
void ref_blue_noise(float* xy, uint32_t n, int s) { util::blueNoise((glm::vec2*)xy, n, s); }
void ref_uniform(float* xy, uint32_t n, uint32_t seed, float lo, float hi) { util::uniformRandomFloats<glm::vec2>((glm::vec2*)xy, n, seed, lo, hi); }
void ref_radial_sobol(float* xy, uint32_t n, uint32_t s) { util::radialSobol((glm::vec2*)xy, n, s); }
void ref_random_polygonal(float* xy, uint32_t edges, uint32_t n, uint32_t seed) { util::randomPolygonal((glm::vec2*)xy, edges, n, seed); }
}

// The UNMODIFIED reference LUT generator (Source/HeatrayRenderer/Materials/MultiScatterUtil.cpp) included
// as a translation unit, with link stubs for the OpenRL / FreeImage / logging symbols it references.
// Output: oracle/_ref/libref_lut.so (git-ignored). Test infrastructure.
#include <HeatrayRenderer/Materials/MultiScatterUtil.cpp>
extern "C" {
// value stored in the LUT for (row, col) of a dim x dim table with `samples` Sobol(0) samples
float ref_lut_value(int row, int col, int dim, int samples) {
    std::vector<glm::vec2> seq(samples);
    util::sobol(seq.data(), samples, 0);
    float roughness = glm::saturate<float, glm::defaultp>((float(row) + 0.5f) / float(dim));
    float alpha = roughness * roughness;
    float NdotV = glm::saturate<float, glm::defaultp>((float(col) + 0.5f) / float(dim));
    float value = generateValue(NdotV, alpha, samples, seq);
    return (1.0f - value) / value;
}
void ref_lut_rows(float* out, int row0, int row1, int dim, int samples) {
    std::vector<glm::vec2> seq(samples);
    util::sobol(seq.data(), samples, 0);
    for (int row = row0; row < row1; ++row) {
        float roughness = glm::saturate<float, glm::defaultp>((float(row) + 0.5f) / float(dim));
        float alpha = roughness * roughness;
        for (int col = 0; col < dim; ++col) {
            float NdotV = glm::saturate<float, glm::defaultp>((float(col) + 0.5f) / float(dim));
            float value = generateValue(NdotV, alpha, samples, seq);
            out[(size_t)(row - row0) * dim + col] = (1.0f - value) / value;
        }
    }
}
}

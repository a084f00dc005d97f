// host_api.h — C++ declarations of the host layer (namespace hr); the flat C view is include/heatray_host.h.
#pragma once
#include "../../include/heatray_host.h"
#include <cstddef>
#include <cstdint>

namespace hr {
// host_qmc.cpp
void sobol(float* xy, uint32_t count, uint32_t sequenceIndex);
void halton(float* xy, uint32_t count, uint32_t sequenceIndex);
void hammersley(float* xy, uint32_t count, uint32_t sequenceIndex);
void blueNoise(float* xy, uint32_t count, int sequenceIndex);
void uniformRandom(float* xy, uint32_t count, uint32_t seed, float lo, float hi);
void radialSobol(float* xy, uint32_t count, uint32_t sequenceIndex);
void randomPolygonal(float* xy, uint32_t numEdges, uint32_t count, uint32_t seed);
void generateRandomSequences(int sampleMode, int bokehShape, uint32_t numSequences, uint32_t sampleCount, float* xy, float* aperture);
size_t sequenceOffsetCount(int w, int h);
void generateSequenceOffsets(int w, int h, float* outX, size_t count);
void generateSequenceIds(int w, int h, int numSequences, uint8_t* out);
void generateInteractiveBlockCoords(int blockW, int blockH, uint32_t seed, float* outRGB);
void generateInteractiveBlockRand(int blocksX, int blocksY, float* outXY);
} // namespace hr

namespace hr {
// host_scene.cpp
uint32_t packPbrMaterial(const HrhPbrMaterial& in, HrcTexture lut, HrcPbrParams& out);
uint32_t packGlassMaterial(const HrhGlassMaterial& in, HrcGlassParams& out);
void generateMultiscatterLUT(float* out, int dim, int sampleCount, int threads);
} // namespace hr

// bvh_builder.h — host-side acceleration-structure build for the CUDA backend.
//
// Replaces OpenRL's opaque acceleration structure (triggered by rlDrawElements, reference
// Source/HeatrayRenderer/Scene/Mesh.cpp:152). Binned-SAH BVH2 -> greedy collapse to an 8-wide BVH ->
// 80-byte compressed nodes (quantised child boxes, octant-ordered child slots) laid out breadth-first.
#pragma once
#include <cstdint>
#include <vector>

namespace hr {

struct BuildTriangle { float v0[3], v1[3], v2[3]; };

// 80-byte compressed wide node = 5 x 16-byte loads on the device.
struct alignas(16) WideNode {
    float    origin[3];        // quantisation-grid origin (node AABB min)
    uint8_t  exp[3];           // per-axis biased exponent: cell size = 2^(exp-127)
    uint8_t  imask;            // bit s set: slot s holds an internal child
    uint32_t childBase;        // index of the first internal child (children are contiguous, slot order)
    uint32_t triBase;          // index of the node's first triangle in the leaf-ordered triangle array
    uint8_t  meta[8];          // per slot: 0 empty | internal: 0x20 | (24 + slot) | leaf: unary count << 5 | first triangle offset
    uint8_t  qlo[3][8];        // quantised child AABB min, per axis, per slot
    uint8_t  qhi[3][8];        // quantised child AABB max
};
static_assert(sizeof(WideNode) == 80, "compressed wide node must be 80 bytes");

struct WideBVH {
    std::vector<WideNode> nodes;       // nodes[0] is the root
    std::vector<uint32_t> triOrder;    // leaf order -> input triangle index
    float sceneMin[3], sceneMax[3];
    uint32_t maxDepth = 0;
    double buildSeconds = 0.0;
};

// threads <= 0: use all hardware threads.
void buildWideBVH(const BuildTriangle* tris, size_t count, int threads, WideBVH& out);

} // namespace hr

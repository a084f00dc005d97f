// bvh_eval.cpp — DEVELOPMENT TOOL (not product, not oracle): builds the product's 8-wide compressed BVH (csrc/bvh_builder.cpp) for a triangle
// soup and walks it on the CPU the way k_trace does (octant-ordered slots, nearest pending child first, conservative quantised slab test,
// leaf triangles right after the node step), counting node steps and triangle tests per ray. Used to compare builder variants without a GPU:
// the counts are what the GPU's HRC_OPT_TRAVERSAL_COUNTERS report (same structure, same order), up to last-place rounding in the slab test.
//
//   g++ -O2 -std=c++17 -pthread tools/bvh_eval.cpp heatray_b200/csrc/bvh_builder.cpp -o /tmp/bvh_eval
//   /tmp/bvh_eval scene.bin          (written by tools/bvh_eval.py: triangles + rays captured from an oracle pass)
#include "../heatray_b200/csrc/bvh_builder.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

using namespace hr;

struct Ray { float o[3], tmin, d[3], tmax; int kind; };      // kind 0 primary, 1 extension (closest hit), 2 occlusion (any hit)

static bool triTest(const BuildTriangle& t, const Ray& r, float& tt) {
    const float e1[3] = {t.v1[0] - t.v0[0], t.v1[1] - t.v0[1], t.v1[2] - t.v0[2]}, e2[3] = {t.v2[0] - t.v0[0], t.v2[1] - t.v0[1], t.v2[2] - t.v0[2]};
    const float pv[3] = {r.d[1] * e2[2] - r.d[2] * e2[1], r.d[2] * e2[0] - r.d[0] * e2[2], r.d[0] * e2[1] - r.d[1] * e2[0]};
    const float det = e1[0] * pv[0] + e1[1] * pv[1] + e1[2] * pv[2];
    if (det == 0.0f) return false;
    const float inv = 1.0f / det;
    const float tv[3] = {r.o[0] - t.v0[0], r.o[1] - t.v0[1], r.o[2] - t.v0[2]};
    const float u = (tv[0] * pv[0] + tv[1] * pv[1] + tv[2] * pv[2]) * inv;
    if (u < 0.0f || u > 1.0f) return false;
    const float qv[3] = {tv[1] * e1[2] - tv[2] * e1[1], tv[2] * e1[0] - tv[0] * e1[2], tv[0] * e1[1] - tv[1] * e1[0]};
    const float v = (r.d[0] * qv[0] + r.d[1] * qv[1] + r.d[2] * qv[2]) * inv;
    if (v < 0.0f || u + v > 1.0f) return false;
    tt = (e2[0] * qv[0] + e2[1] * qv[1] + e2[2] * qv[2]) * inv;
    return tt > r.tmin && tt < r.tmax;
}

struct Counts { double nodes = 0, tris = 0, rays = 0; };

// Returns the nearest hit distance (closest) or 0 / -1 for occluded / not occluded (any); r.tmax when nothing is hit.
static float trace(const WideBVH& bvh, const std::vector<BuildTriangle>& tris, const Ray& r, bool any, Counts& c) {
    float idir[3]; uint32_t octInv = 0;
    for (int a = 0; a < 3; ++a) {
        float d = r.d[a]; if (std::fabs(d) < 1e-20f) d = std::copysign(1e-20f, d);
        idir[a] = 1.0f / d;
        if (!(d < 0.0f)) octInv |= (4u >> a);                 // x -> bit 2, y -> bit 1, z -> bit 0: slot bit set = positive side
    }
    if (any) octInv ^= 7u;                                  // occlusion rays visit children back to front (kernels.cu: Traversal::nodeStep); the slab test below does not use octInv
    float bestT = r.tmax;
    struct Group { uint32_t base, hits, imask; float tn[8]; };      // tn: entry distance of each pending internal child, by bit position (BVH_EVAL_GROUP_CULL only)
    std::vector<Group> stack;
    Group g{0u, 0x80000000u, 0u, {0, 0, 0, 0, 0, 0, 0, 0}};                            // root = "child 7 of a virtual parent whose childBase is 0"
    bool rootPending = true;
    c.rays += 1;
    for (;;) {
        uint32_t triBase = 0, triBits = 0;
        if (g.hits & 0xff000000u) {
            const uint32_t bit = 31u - __builtin_clz(g.hits);
            g.hits &= ~(1u << bit);
            uint32_t nodeIdx;
            if (rootPending) { nodeIdx = 0; rootPending = false; }
            else {
                const uint32_t slot = (bit - 24u) ^ (octInv & 7u);
                nodeIdx = g.base + __builtin_popcount(g.imask & ~(0xffffffffu << slot));
            }
            if (g.hits & 0xff000000u) stack.push_back(g);
            const WideNode& n = bvh.nodes[nodeIdx];
            c.nodes += 1;
            uint32_t hitMask = 0; float newTn[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            const float tfar = any ? r.tmax : bestT;
            float cell[3]; for (int a = 0; a < 3; ++a) cell[a] = n.exp[a] ? std::ldexp(1.0f, (int)n.exp[a] - 127) : 0.0f;
            for (int s = 0; s < 8; ++s) {
                const uint8_t m = n.meta[s];
                if (!m) continue;
                float tn = r.tmin, tf = tfar;
                for (int a = 0; a < 3; ++a) {
                    const float lo = n.origin[a] + n.qlo[a][s] * cell[a], hi = n.origin[a] + n.qhi[a][s] * cell[a];
                    const float t0 = (lo - r.o[a]) * idir[a], t1 = (hi - r.o[a]) * idir[a];
                    const float bias = std::fabs(cell[a] * idir[a]) * (1.0f / 512.0f);
                    tn = std::fmax(tn, std::fmin(t0, t1) - bias); tf = std::fmin(tf, std::fmax(t0, t1) + bias);
                }
                if (!(tn <= tf * 1.000001f)) continue;
                const bool inner = (m & 0x20) && ((m & 0x1f) >= 24);
                if (inner) { hitMask |= 1u << (24u + ((uint32_t)s ^ (octInv & 7u))); newTn[(uint32_t)s ^ (octInv & 7u)] = tn; }
                else hitMask |= (uint32_t)(m >> 5) << (m & 0x1f);
            }
            g.base = n.childBase; g.imask = n.imask; g.hits = hitMask & 0xff000000u; for (int q = 0; q < 8; ++q) g.tn[q] = newTn[q];
            triBase = n.triBase; triBits = hitMask & 0x00ffffffu;
        }
        while (triBits) {
            const uint32_t ti = 31u - __builtin_clz(triBits);
            triBits &= ~(1u << ti);
            c.tris += 1;
            float t;
            Ray rr = r; rr.tmax = r.tmax;
            if (triTest(tris[bvh.triOrder[triBase + ti]], rr, t)) {
                if (any) return 0.0f;
                if (t < bestT) bestT = t;
            }
        }
        if (!(g.hits & 0xff000000u)) {
            if (stack.empty()) return any ? -1.0f : bestT;
            g = stack.back(); stack.pop_back();
            // What-if (not in the kernel): a group coming off the stack is dropped when the nearest of its remaining children already lies beyond the
            // best hit. The key models 16 spare bits of the stack word: the entry distance truncated to its upper 16 bits (rounds down: conservative).
            static const bool groupCull = std::getenv("BVH_EVAL_GROUP_CULL") != nullptr;
            if (groupCull && !any) {
                float mn = 3.4e38f; for (int q = 0; q < 8; ++q) if (g.hits & (1u << (24 + q))) mn = std::fmin(mn, g.tn[q]);
                uint32_t bits; std::memcpy(&bits, &mn, 4); bits &= 0xffff0000u; std::memcpy(&mn, &bits, 4);
                if (mn > bestT) { g.hits = 0; continue; }
            }
        }
    }
}

int main(int argc, char** argv) {
    if (argc < 2) { std::fprintf(stderr, "usage: bvh_eval scene.bin\n"); return 2; }
    FILE* f = std::fopen(argv[1], "rb");
    if (!f) { std::perror(argv[1]); return 1; }
    uint64_t nt = 0, nr = 0;
    if (std::fread(&nt, 8, 1, f) != 1) return 1;
    std::vector<BuildTriangle> tris(nt);
    if (std::fread(tris.data(), sizeof(BuildTriangle), nt, f) != nt) return 1;
    if (std::fread(&nr, 8, 1, f) != 1) return 1;
    std::vector<Ray> rays(nr);
    if (std::fread(rays.data(), sizeof(Ray), nr, f) != nr) return 1;
    std::fclose(f);
    WideBVH bvh;
    buildWideBVH(tris.data(), tris.size(), 0, bvh);
    Counts c[3];
    if (argc > 2 && std::strcmp(argv[2], "--check") == 0) {
        // every ray against every triangle: the walk over the compressed wide BVH must find exactly the brute-force answer (same triangle test)
        size_t bad = 0;
        for (const Ray& r : rays) {
            const bool any = r.kind == 2;
            const float got = trace(bvh, tris, r, any, c[r.kind]);
            float best = r.tmax; bool hit = false;
            for (const BuildTriangle& t : tris) { float tt; if (triTest(t, r, tt)) { hit = true; if (tt < best) best = tt; } }
            const float want = any ? (hit ? 0.0f : -1.0f) : best;
            if (got != want) { if (bad < 5) std::printf("MISMATCH kind %d: bvh %.9g brute force %.9g\n", r.kind, got, want); ++bad; }
        }
        std::printf("check: %zu rays, %zu mismatches, %zu nodes, depth %u\n", rays.size(), bad, bvh.nodes.size(), bvh.maxDepth);
        return bad ? 1 : 0;
    }
    for (const Ray& r : rays) trace(bvh, tris, r, r.kind == 2, c[r.kind]);
    const char* names[3] = {"primary", "extension", "occlusion"};
    double cost = 0, total = 0;
    std::printf("%zu triangles, %zu nodes, depth %u, build %.2f s\n", tris.size(), bvh.nodes.size(), bvh.maxDepth, bvh.buildSeconds);
    for (int k = 0; k < 3; ++k) {
        if (c[k].rays == 0) continue;
        std::printf("  %-10s %9.0f rays  nodes/ray %6.3f  tris/ray %6.3f\n", names[k], c[k].rays, c[k].nodes / c[k].rays, c[k].tris / c[k].rays);
        // lane-adjusted issue cost from the ncu source page (profiles/r02_v5_ncu_trace_regions.txt): a node step ~322 instructions at ~25 lanes,
        // a triangle test ~50 instructions at ~3.5 (closest) / 4.5 (any) lanes
        cost += c[k].nodes * (322.0 / 25.0) + c[k].tris * (50.0 / (k == 2 ? 4.5 : 3.5));
        total += c[k].rays;
    }
    std::printf("  lane-adjusted cost per ray %.2f\n", cost / total);
    std::printf("totals closest_nodes %.0f closest_tris %.0f any_nodes %.0f any_tris %.0f\n", c[0].nodes + c[1].nodes, c[0].tris + c[1].tris, c[2].nodes, c[2].tris);
    return 0;
}

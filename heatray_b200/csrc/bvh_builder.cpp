// bvh_builder.cpp — binned-SAH BVH2 build, greedy 8-wide collapse, 80-byte node compression.
// See bvh_builder.h. The layout follows the compressed-wide-BVH idea (8 children, 8-bit quantised
// boxes on a per-node power-of-two grid, octant-ordered slots); the code is this project's own.
#include "bvh_builder.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <thread>

namespace hr {
namespace {

struct Box {
    float lo[3], hi[3];
    void reset() { for (int a = 0; a < 3; ++a) { lo[a] = std::numeric_limits<float>::infinity(); hi[a] = -std::numeric_limits<float>::infinity(); } }
    void grow(const Box& b) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], b.lo[a]); hi[a] = std::max(hi[a], b.hi[a]); } }
    void grow(const float* p) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], p[a]); hi[a] = std::max(hi[a], p[a]); } }
    float halfArea() const {
        float ex = hi[0] - lo[0], ey = hi[1] - lo[1], ez = hi[2] - lo[2];
        return ex * ey + ey * ez + ez * ex;
    }
};

struct Prim { Box box; float c[3]; uint32_t id; };

struct Node2 {
    Box box;
    uint32_t left = 0, right = 0;   // children (internal)
    uint32_t first = 0, count = 0;  // prim range (leaf when count > 0)
};

constexpr int kMaxBins = 64;
int g_bins = 32;                            // SAH bins per axis; 32 measured best on configs #2 / #3 (8: +1.7 / +2.4 %, 16: 0, 32: +1.2 / +4.0 %, 64: +1.0 / +3.4 %
                                            // in whole-pass Mrays/s; exact sweep SAH below 64..4096 triangles changed nothing). HRC_TUNE_BINS overrides.
#define kBins g_bins
constexpr int kMaxBinaryDepth = 30;        // < kTraceStack (device_types.h); hrcCommitScene keeps its depth check as an assertion
constexpr uint32_t kMaxLeafLimit = 3;      // a child slot encodes at most 3 triangles (unary count in 3 bits)
// Leaf policy: at most g_maxLeaf triangles, made a leaf when n * area <= g_travCost * area + SAH(best split). Defaults measured on B200
// (profiles/); HRC_TUNE_MAX_LEAF / HRC_TUNE_TRAV_COST override them for experiments.
uint32_t g_maxLeaf = 3; float g_travCost = 0.5f;
bool g_splitLeaves = true;                // HRC_TUNE_SPLIT_LEAVES=0: keep multi-triangle leaves together even when the wide node has free slots
#define kMaxLeaf g_maxLeaf

struct Builder2 {
    std::vector<Prim>& prims;
    std::vector<Node2>& nodes;
    std::atomic<uint32_t> nextNode{1};
    int spawnDepth;

    Builder2(std::vector<Prim>& p, std::vector<Node2>& n, int sd) : prims(p), nodes(n), spawnDepth(sd) {}

    void build(uint32_t nodeIdx, uint32_t begin, uint32_t end, int depth) {
        Box bounds, cbounds; bounds.reset(); cbounds.reset();
        for (uint32_t i = begin; i < end; ++i) { bounds.grow(prims[i].box); cbounds.grow(prims[i].c); }
        Node2& node = nodes[nodeIdx];
        node.box = bounds;
        const uint32_t n = end - begin;
        if (n == 1) { node.first = begin; node.count = 1; return; }

        // binned SAH over the three axes
        float bestCost = std::numeric_limits<float>::infinity(); int bestAxis = -1, bestSplit = -1;
        for (int a = 0; a < 3; ++a) {
            const float cmin = cbounds.lo[a], cmax = cbounds.hi[a];
            if (!(cmax > cmin)) continue;
            Box bb[kMaxBins]; uint32_t bc[kMaxBins];
            for (int b = 0; b < kBins; ++b) { bb[b].reset(); bc[b] = 0; }
            const float scale = kBins / (cmax - cmin);
            for (uint32_t i = begin; i < end; ++i) {
                int b = std::min(kBins - 1, std::max(0, (int)((prims[i].c[a] - cmin) * scale)));
                bb[b].grow(prims[i].box); bc[b]++;
            }
            float ra[kMaxBins]; uint32_t rc[kMaxBins];
            Box acc; acc.reset(); uint32_t cnt = 0;
            for (int b = kBins - 1; b > 0; --b) { acc.grow(bb[b]); cnt += bc[b]; ra[b] = cnt ? acc.halfArea() : 0.0f; rc[b] = cnt; }
            acc.reset(); cnt = 0;
            for (int b = 0; b < kBins - 1; ++b) {
                acc.grow(bb[b]); cnt += bc[b];
                if (cnt == 0 || rc[b + 1] == 0) continue;
                float cost = acc.halfArea() * cnt + ra[b + 1] * rc[b + 1];
                if (cost < bestCost) { bestCost = cost; bestAxis = a; bestSplit = b; }
            }
        }

        // Depth guard: the traversal stack holds kTraceStack entries (one per wide level, and a wide level spans at least one binary
        // level). A branch that is unusually deep for the primitives it still holds (SAH peeling one triangle per level off a skewed
        // distribution) finishes with median splits, which need ceil(log2 n) more levels, so the tree can never outgrow the stack.
        int log2n = 0; while ((1u << log2n) < n) ++log2n;
        if (depth + log2n >= kMaxBinaryDepth) bestAxis = -1;

        uint32_t mid;
        if (bestAxis < 0) {                       // all centroids coincide (or the depth guard fired): median split
            if (n <= kMaxLeaf) { node.first = begin; node.count = n; return; }
            mid = begin + n / 2;
            int wide = 0;
            for (int a = 1; a < 3; ++a) if (cbounds.hi[a] - cbounds.lo[a] > cbounds.hi[wide] - cbounds.lo[wide]) wide = a;
            if (cbounds.hi[wide] > cbounds.lo[wide])
                std::nth_element(prims.begin() + begin, prims.begin() + mid, prims.begin() + end, [wide](const Prim& x, const Prim& y) { return x.c[wide] < y.c[wide]; });
        } else {
            const float area = bounds.halfArea();
            if (n <= kMaxLeaf && (float)n * area <= g_travCost * area + bestCost) { node.first = begin; node.count = n; return; }
            const float cmin = cbounds.lo[bestAxis], cmax = cbounds.hi[bestAxis];
            const float scale = kBins / (cmax - cmin);
            auto it = std::partition(prims.begin() + begin, prims.begin() + end, [&](const Prim& p) {
                int b = std::min(kBins - 1, std::max(0, (int)((p.c[bestAxis] - cmin) * scale)));
                return b <= bestSplit;
            });
            mid = (uint32_t)(it - prims.begin());
            if (mid == begin || mid == end) mid = begin + n / 2;
        }
        const uint32_t l = nextNode.fetch_add(2);
        node.left = l; node.right = l + 1;
        if (depth < spawnDepth && n > 32768) {
            std::thread t([this, l, begin, mid, depth]() { build(l, begin, mid, depth + 1); });
            build(l + 1, mid, end, depth + 1);
            t.join();
        } else {
            build(l, begin, mid, depth + 1);
            build(l + 1, mid, end, depth + 1);
        }
    }
};

// direction sign vector of slot s: bit 2 -> x, bit 1 -> y, bit 0 -> z; set bit = positive side
inline float slotAffinity(const float d[3], int s) {
    return ((s & 4) ? d[0] : -d[0]) + ((s & 2) ? d[1] : -d[1]) + ((s & 1) ? d[2] : -d[2]);
}

inline uint8_t gridExponent(float extent) {
    if (!(extent > 0.0f)) return 0;               // flat axis: cell size 0
    int e; std::frexp(extent / 255.0f, &e);       // extent/255 = m * 2^e, m in [0.5, 1)  => 2^e >= extent/255
    while (std::ldexp(255.0, e) < (double)extent) ++e;
    while (e > -126 && std::ldexp(255.0, e - 1) >= (double)extent) --e;
    e = std::max(-126, std::min(127, e));
    return (uint8_t)(e + 127);
}

} // namespace

void buildWideBVH(const BuildTriangle* tris, size_t count, int threads, WideBVH& out) {
    auto t0 = std::chrono::steady_clock::now();
    out.nodes.clear(); out.triOrder.clear(); out.maxDepth = 0;
    for (int a = 0; a < 3; ++a) { out.sceneMin[a] = 0; out.sceneMax[a] = 0; }
    if (count == 0) {
        WideNode root; std::memset(&root, 0, sizeof(root));
        out.nodes.push_back(root);
        return;
    }
    if (const char* e = std::getenv("HRC_TUNE_MAX_LEAF")) g_maxLeaf = (uint32_t)std::max(1, std::min((int)kMaxLeafLimit, std::atoi(e)));
    if (const char* e = std::getenv("HRC_TUNE_TRAV_COST")) g_travCost = (float)std::atof(e);
    if (const char* e = std::getenv("HRC_TUNE_SPLIT_LEAVES")) g_splitLeaves = std::atoi(e) != 0;
    if (const char* e = std::getenv("HRC_TUNE_BINS")) g_bins = std::max(2, std::min(kMaxBins, std::atoi(e)));
    if (threads <= 0) threads = (int)std::thread::hardware_concurrency();
    if (threads < 1) threads = 1;
    int spawnDepth = 0; while ((1 << spawnDepth) < threads) ++spawnDepth;
    spawnDepth += 1;

    std::vector<Prim> prims(count);
    for (size_t i = 0; i < count; ++i) {
        Prim& p = prims[i]; p.box.reset();
        p.box.grow(tris[i].v0); p.box.grow(tris[i].v1); p.box.grow(tris[i].v2);
        for (int a = 0; a < 3; ++a) p.c[a] = 0.5f * (p.box.lo[a] + p.box.hi[a]);
        p.id = (uint32_t)i;
    }
    std::vector<Node2> nodes2(2 * count);
    {
        Builder2 b(prims, nodes2, spawnDepth);
        b.build(0, 0, (uint32_t)count, 0);
        nodes2.resize(b.nextNode.load());
    }
    for (int a = 0; a < 3; ++a) { out.sceneMin[a] = nodes2[0].box.lo[a]; out.sceneMax[a] = nodes2[0].box.hi[a]; }

    // ---- collapse to 8-wide, breadth first ------------------------------------------------
    struct Pending { uint32_t n2; uint32_t wide; uint32_t depth; };
    std::vector<Pending> queue; queue.reserve(count / 2 + 16);
    out.nodes.reserve(count / 2 + 16);
    out.triOrder.reserve(count);
    out.nodes.push_back(WideNode{});
    queue.push_back(Pending{0, 0, 1});
    for (size_t qi = 0; qi < queue.size(); ++qi) {
        const Pending cur = queue[qi];
        out.maxDepth = std::max(out.maxDepth, cur.depth);
        const Node2& src = nodes2[cur.n2];
        uint32_t kids[8]; int nk = 0;
        if (src.count > 0) kids[nk++] = cur.n2;                // degenerate: the whole scene is one leaf
        else { kids[nk++] = src.left; kids[nk++] = src.right; }
        while (nk < 8) {
            int best = -1; float bestArea = -1.0f;
            for (int k = 0; k < nk; ++k) {
                const Node2& c = nodes2[kids[k]];
                if (c.count == 0) { float ar = c.box.halfArea(); if (ar > bestArea) { bestArea = ar; best = k; } }
            }
            if (best < 0) break;
            const Node2& c = nodes2[kids[best]];
            kids[best] = c.left; kids[nk++] = c.right;
        }
        // Child records: inner kids keep their binary node; leaf kids carry (first, count, box) so that a leaf can be split further here.
        struct Kid { uint32_t n2, first, count; Box box; };
        Kid kid[8];
        for (int k = 0; k < nk; ++k) { const Node2& c = nodes2[kids[k]]; kid[k] = Kid{kids[k], c.first, c.count, c.box}; }
        // Free slots cost nothing in the node step (all 8 are evaluated): while some are left, peel triangles off multi-triangle leaves
        // into children of their own, largest expected saving in triangle tests first (area(leaf) * n - area(tri) - area(rest) * (n - 1)).
        if (g_splitLeaves) {
            auto boxOf = [&](uint32_t first, uint32_t n) { Box b; b.reset(); for (uint32_t i = 0; i < n; ++i) { b.grow(prims[first + i].box.lo); b.grow(prims[first + i].box.hi); } return b; };
            while (nk < 8) {
                int best = -1; float bestGain = 0.0f; Box bestHead, bestRest;
                for (int k = 0; k < nk; ++k) {
                    if (kid[k].count < 2) continue;
                    const Box head = boxOf(kid[k].first, 1), rest = boxOf(kid[k].first + 1, kid[k].count - 1);
                    const float gain = kid[k].box.halfArea() * (float)kid[k].count - head.halfArea() - rest.halfArea() * (float)(kid[k].count - 1);
                    if (gain > bestGain) { bestGain = gain; best = k; bestHead = head; bestRest = rest; }
                }
                if (best < 0) break;
                kid[nk++] = Kid{0u, kid[best].first + 1, kid[best].count - 1, bestRest};
                kid[best].count = 1; kid[best].box = bestHead;
            }
        }
        // octant-ordered slot assignment: greedily give each slot the child lying furthest along its sign vector
        int slotOf[8]; bool slotUsed[8] = {false}; bool kidDone[8] = {false};
        float pc[3]; for (int a = 0; a < 3; ++a) pc[a] = 0.5f * (src.box.lo[a] + src.box.hi[a]);
        float aff[8][8];
        for (int k = 0; k < nk; ++k) {
            const Box& cb = kid[k].box;
            float d[3]; for (int a = 0; a < 3; ++a) d[a] = 0.5f * (cb.lo[a] + cb.hi[a]) - pc[a];
            for (int s = 0; s < 8; ++s) aff[k][s] = slotAffinity(d, s);
        }
        for (int round = 0; round < nk; ++round) {
            int bk = -1, bs = -1; float bv = -std::numeric_limits<float>::infinity();
            for (int k = 0; k < nk; ++k) if (!kidDone[k])
                for (int s = 0; s < 8; ++s) if (!slotUsed[s] && aff[k][s] > bv) { bv = aff[k][s]; bk = k; bs = s; }
            if (bk < 0) {   // NaN affinities (degenerate boxes): fall back to first free pair
                for (int k = 0; k < nk && bk < 0; ++k) if (!kidDone[k]) bk = k;
                for (int s = 0; s < 8 && bs < 0; ++s) if (!slotUsed[s]) bs = s;
            }
            slotOf[bk] = bs; slotUsed[bs] = true; kidDone[bk] = true;
        }
        int kidAt[8]; for (int s = 0; s < 8; ++s) kidAt[s] = -1;
        for (int k = 0; k < nk; ++k) kidAt[slotOf[k]] = k;

        WideNode w; std::memset(&w, 0, sizeof(w));
        for (int a = 0; a < 3; ++a) { w.origin[a] = src.box.lo[a]; w.exp[a] = gridExponent(src.box.hi[a] - src.box.lo[a]); }
        w.childBase = (uint32_t)out.nodes.size();
        w.triBase = (uint32_t)out.triOrder.size();
        uint32_t triOffset = 0, innerCount = 0;
        for (int s = 0; s < 8; ++s) {
            if (kidAt[s] < 0) continue;
            const Kid& c = kid[kidAt[s]];
            for (int a = 0; a < 3; ++a) {
                double cell = w.exp[a] ? std::ldexp(1.0, (int)w.exp[a] - 127) : 0.0;
                int lo = 0, hi = 0;
                if (cell > 0.0) {
                    lo = (int)std::floor(((double)c.box.lo[a] - (double)w.origin[a]) / cell);
                    hi = (int)std::ceil(((double)c.box.hi[a] - (double)w.origin[a]) / cell);
                }
                w.qlo[a][s] = (uint8_t)std::max(0, std::min(255, lo));
                w.qhi[a][s] = (uint8_t)std::max(0, std::min(255, hi));
            }
            if (c.count == 0) {
                w.imask |= (uint8_t)(1u << s);
                w.meta[s] = (uint8_t)(0x20 | (24 + s));
                queue.push_back(Pending{c.n2, w.childBase + innerCount, cur.depth + 1});
                innerCount++;
            } else {
                const uint32_t unary = (1u << c.count) - 1u;   // 1 -> 001, 2 -> 011, 3 -> 111
                w.meta[s] = (uint8_t)((unary << 5) | triOffset);
                for (uint32_t i = 0; i < c.count; ++i) out.triOrder.push_back(prims[c.first + i].id);
                triOffset += c.count;
            }
        }
        out.nodes.resize(out.nodes.size() + innerCount);
        out.nodes[cur.wide] = w;
    }
    out.buildSeconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

} // namespace hr

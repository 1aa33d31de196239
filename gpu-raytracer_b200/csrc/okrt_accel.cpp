/*
 * okrt_accel.cpp — builds the device acceleration layout described in okrt_accel.h from the
 * reference-format arrays handed to okrt_set_* (t0 Triangle, t2 BvhNode, t5 OctTreeNode, t7 Mesh;
 * resources/shaders/GPU-Structs.hlsli:95-98,113-125,159-179).  Host C++, runs once per scene change.
 */
#include "okrt_accel.h"
#include "okrt_internal.h"

#include <algorithm>
#include <atomic>
#include <cfloat>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <future>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

namespace okrt {
namespace {

inline float bitsToFloat(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }

struct Builder {
    const okrt_triangle* tris; uint32_t numTris;
    const okrt_bvh_node* nodes; uint32_t numNodes;
    FastLayout& out;
    uint32_t deepestPair = 0;     /* pair-node levels below a mesh root (for the fast stack bound) */
    uint32_t deepestLeafLevel = 0;/* reference node levels (for the exact stack bound) */
    bool failed = false;

    uint32_t newPair()
    {
        const uint32_t idx = (uint32_t)(out.nodes.size() / 4);
        out.nodes.resize(out.nodes.size() + 4, Float4{0, 0, 0, 0});
        return idx;
    }
    void setChild(uint32_t pair, int which, const okrt_aabb& box, uint32_t ref)
    {
        Float4* q = &out.nodes[(size_t)pair * 4];
        if (ref == kRefNone) {                 /* empty child: a NaN box fails every comparison of the slab test */
            const float nan = std::nanf("");
            q[which * 2 + 0].x = q[which * 2 + 0].y = q[which * 2 + 0].z = nan;
            q[which * 2 + 1].x = q[which * 2 + 1].y = q[which * 2 + 1].z = nan;
            if (which == 0) q[0].w = bitsToFloat(ref); else q[1].w = bitsToFloat(ref);
            return;
        }
        q[which * 2 + 0].x = box.min[0]; q[which * 2 + 0].y = box.min[1]; q[which * 2 + 0].z = box.min[2];
        q[which * 2 + 1].x = box.max[0]; q[which * 2 + 1].y = box.max[1]; q[which * 2 + 1].z = box.max[2];
        if (which == 0) q[0].w = bitsToFloat(ref); else q[1].w = bitsToFloat(ref);
    }
    bool fail(const std::string& why) { if (!failed) out.error = why; failed = true; return false; }

    /* child reference for triangles [start, start+count) bounded by `box` */
    uint32_t leafRef(uint32_t start, uint32_t count, const okrt_aabb& box, uint32_t level, uint32_t pairLevel)
    {
        if (count == 0) return kRefNone;
        if ((uint64_t)start + count > numTris || start >= kMaxTriIndex) { fail("BVH leaf references triangles out of range"); return kRefNone; }
        if (count <= kMaxLeafTris) return kRefTriLeaf | ((count - 1u) << 27) | start;
        /* oversized leaf (depth-limited or maxLeafTriangles > 8 trees): a balanced tree of pair nodes that repeat the
         * leaf's box over 8-triangle records — log2(count / 8) stack levels, not count / 8 */
        const uint32_t p = newPair();
        deepestPair = std::max(deepestPair, pairLevel + 1);
        const uint32_t records = (count + kMaxLeafTris - 1u) / kMaxLeafTris;
        const uint32_t left = ((records + 1u) / 2u) * kMaxLeafTris;
        const uint32_t r0 = leafRef(start, left, box, level, pairLevel + 1);
        const uint32_t r1 = leafRef(start + left, count - left, box, level, pairLevel + 1);
        setChild(p, 0, box, r0);
        setChild(p, 1, box, r1);
        return p;
    }

    /* a reference BVH that passed flattenable(): records its depth (the reference-order stack bound) and says whether
     * any leaf holds a triangle */
    bool hasTriangles(uint32_t root)
    {
        bool any = false;
        std::vector<std::pair<uint32_t, uint32_t>> st{{root, 0u}};
        while (!st.empty()) {
            const auto [idx, level] = st.back(); st.pop_back();
            const okrt_bvh_node& n = nodes[idx];
            if (n.firstChildIdx != OKRT_INVALID_UINT) { st.push_back({n.firstChildIdx, level + 1}); st.push_back({n.firstChildIdx + 1, level + 1}); continue; }
            deepestLeafLevel = std::max(deepestLeafLevel, level);
            any = any || (n.triStart != OKRT_INVALID_UINT && n.triEnd > n.triStart);
        }
        return any;
    }

    uint32_t convert(uint32_t nodeIdx, uint32_t level, uint32_t pairLevel)
    {
        if (failed) return kRefNone;
        if (nodeIdx >= numNodes) { fail("BVH child index out of range"); return kRefNone; }
        if (level > 4096) { fail("BVH deeper than 4096 levels (cycle?)"); return kRefNone; }
        const okrt_bvh_node& n = nodes[nodeIdx];
        if (n.firstChildIdx == OKRT_INVALID_UINT) {
            deepestLeafLevel = std::max(deepestLeafLevel, level);
            const uint32_t count = (n.triEnd >= n.triStart && n.triStart != OKRT_INVALID_UINT) ? n.triEnd - n.triStart : 0u;
            return leafRef(n.triStart, count, n.boundingBox, level, pairLevel);
        }
        if ((uint64_t)n.firstChildIdx + 1 >= numNodes) { fail("BVH child index out of range"); return kRefNone; }
        const uint32_t p = newPair();
        deepestPair = std::max(deepestPair, pairLevel + 1);
        const uint32_t r0 = convert(n.firstChildIdx, level + 1, pairLevel + 1);
        const uint32_t r1 = convert(n.firstChildIdx + 1, level + 1, pairLevel + 1);
        setChild(p, 0, nodes[n.firstChildIdx].boundingBox, r0);
        setChild(p, 1, nodes[n.firstChildIdx + 1].boundingBox, r1);
        return p;
    }
};

/* explicit node of the scene-wide tree while it is built and optimised; a leaf has left < 0 and carries `ref` */
/* wide-layout payload of a leaf: triangles [wstart, wstart + wcount) of a reference leaf and its slack, or (wcount == 0) the
 * instance record wstart */
struct TopNode { double lo[3], hi[3]; int32_t left = -1, right = -1, parent = -1; uint32_t ref = kRefNone; uint32_t wstart = 0, wcount = 0; float slack = 0.f; };
struct TopLeaf { double lo[3], hi[3]; uint32_t ref; uint32_t weight; uint32_t wstart = 0, wcount = 0; float slack = 0.f; };   /* weight: triangles below (SAH cost) */

static inline double topArea(const TopNode& n)
{
    const double dx = n.hi[0] - n.lo[0], dy = n.hi[1] - n.lo[1], dz = n.hi[2] - n.lo[2];
    return 2.0 * (dx * dy + dy * dz + dz * dx);
}
static inline double topUnionArea(const TopNode& a, const TopNode& c)
{
    const double dx = std::max(a.hi[0], c.hi[0]) - std::min(a.lo[0], c.lo[0]);
    const double dy = std::max(a.hi[1], c.hi[1]) - std::min(a.lo[1], c.lo[1]);
    const double dz = std::max(a.hi[2], c.hi[2]) - std::min(a.lo[2], c.lo[2]);
    return 2.0 * (dx * dy + dy * dz + dz * dx);
}
static uint32_t topTreeDepth(const std::vector<TopNode>& tn, int32_t root)
{
    uint32_t deepest = 0;
    std::vector<std::pair<int32_t, uint32_t>> st{{root, 0u}};
    while (!st.empty()) {
        const auto [n, level] = st.back(); st.pop_back();
        if (tn[n].left < 0) continue;
        deepest = std::max(deepest, level + 1);
        st.push_back({tn[n].left, level + 1}); st.push_back({tn[n].right, level + 1});
    }
    return deepest;
}

/* Insertion-based optimisation of a binary BVH (after Bittner, Hapala, Havran 2013, restated): take a subtree out,
 * let its sibling take the parent's place, and put it back where the total surface area of the tree grows least
 * (branch-and-bound search from the root).  Interior boxes stay exact unions of their children, so everything the
 * order-independence argument needs from the tree (conservative boxes) is untouched.  Deterministic: nodes are visited
 * in order of decreasing area, no random numbers.  Returns the (possibly new) root. */
static int32_t reinsertOptimize(std::vector<TopNode>& tn, int32_t root, int passes)
{
    auto refitUp = [&](int32_t n) {
        while (n >= 0) {
            const TopNode &l = tn[tn[n].left], &r = tn[tn[n].right];
            bool changed = false;
            for (int k = 0; k < 3; k++) {
                const double lo = std::min(l.lo[k], r.lo[k]), hi = std::max(l.hi[k], r.hi[k]);
                if (lo != tn[n].lo[k] || hi != tn[n].hi[k]) { tn[n].lo[k] = lo; tn[n].hi[k] = hi; changed = true; }
            }
            if (!changed) break;
            n = tn[n].parent;
        }
    };
    struct Entry { double induced; int32_t node; bool operator<(const Entry& o) const { return induced > o.induced; } };
    std::vector<Entry> heap;
    std::vector<int32_t> order;
    for (int pass = 0; pass < passes; pass++) {
        order.clear();
        for (int32_t i = 0; i < (int32_t)tn.size(); i++) if (i != root && tn[i].parent >= 0 && tn[i].parent != root) order.push_back(i);
        std::vector<double> areaAt(tn.size());
        for (size_t i = 0; i < tn.size(); i++) areaAt[i] = topArea(tn[i]);
        std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t c) { return areaAt[a] > areaAt[c]; });
        size_t moved = 0;
        for (int32_t x : order) {
            const int32_t parent = tn[x].parent;
            if (parent < 0 || parent == root) continue;                     /* became a child of the root meanwhile */
            const int32_t grand = tn[parent].parent;
            const int32_t sibling = tn[parent].left == x ? tn[parent].right : tn[parent].left;
            /* detach x: the sibling takes the parent's place */
            if (tn[grand].left == parent) tn[grand].left = sibling; else tn[grand].right = sibling;
            tn[sibling].parent = grand;
            refitUp(grand);
            /* best place for x */
            const double areaX = topArea(tn[x]);
            double best = DBL_MAX; int32_t bestNode = sibling;
            heap.clear();
            heap.push_back({0.0, root});
            while (!heap.empty()) {
                std::pop_heap(heap.begin(), heap.end());
                const Entry e = heap.back(); heap.pop_back();
                if (e.induced + areaX >= best) break;
                const double direct = topUnionArea(tn[e.node], tn[x]);
                const double total = e.induced + direct;
                if (total < best) { best = total; bestNode = e.node; }
                const double below = total - topArea(tn[e.node]);          /* what every ancestor-to-be already grows by */
                if (tn[e.node].left >= 0 && below + areaX < best) {
                    heap.push_back({below, tn[e.node].left}); std::push_heap(heap.begin(), heap.end());
                    heap.push_back({below, tn[e.node].right}); std::push_heap(heap.begin(), heap.end());
                }
            }
            /* attach: `parent` becomes the new interior node above bestNode and x */
            const int32_t above = tn[bestNode].parent;
            tn[parent].left = bestNode; tn[parent].right = x; tn[parent].parent = above;
            tn[bestNode].parent = parent; tn[x].parent = parent;
            if (above < 0) root = parent;
            else if (tn[above].left == bestNode) tn[above].left = parent; else tn[above].right = parent;
            for (int k = 0; k < 3; k++) { tn[parent].lo[k] = DBL_MAX; tn[parent].hi[k] = -DBL_MAX; }   /* force the refit */
            refitUp(parent);
            if (bestNode != sibling) moved++;
        }
        if (moved == 0) break;
    }
    return root;
}

/* true if every interior node's box contains both child boxes (component-wise, exactly) and nothing is NaN */
bool flattenable(const okrt_bvh_node* nodes, uint32_t numNodes, uint32_t root)
{
    std::vector<uint32_t> st{root};
    size_t visited = 0;
    while (!st.empty()) {
        const uint32_t idx = st.back(); st.pop_back();
        if (idx >= numNodes || ++visited > numNodes) return false;
        const okrt_bvh_node& n = nodes[idx];
        for (int k = 0; k < 3; k++) if (!(n.boundingBox.min[k] <= n.boundingBox.max[k])) { if (n.firstChildIdx != OKRT_INVALID_UINT || (n.triStart != OKRT_INVALID_UINT && n.triEnd > n.triStart)) return false; }
        if (n.firstChildIdx == OKRT_INVALID_UINT) continue;
        if ((uint64_t)n.firstChildIdx + 1 >= numNodes) return false;
        for (int c = 0; c < 2; c++) {
            const okrt_bvh_node& ch = nodes[n.firstChildIdx + c];
            const bool empty = ch.firstChildIdx == OKRT_INVALID_UINT && (ch.triStart == OKRT_INVALID_UINT || ch.triEnd <= ch.triStart);
            if (empty) continue;
            for (int k = 0; k < 3; k++)
                if (!(n.boundingBox.min[k] <= ch.boundingBox.min[k] && ch.boundingBox.max[k] <= n.boundingBox.max[k])) return false;
            st.push_back(n.firstChildIdx + c);
        }
    }
    return true;
}

/* How far a triangle test may report a hit BEFORE the ray enters the triangle's (exact) leaf box, in distance along the ray,
 * beyond the relative window: float32 Moeller-Trumbore has an absolute error of ~10 u |o - p2| / (|d.n| sin(phi)) in t (u = 2^-24,
 * phi = angle between the edges), and |o - p2| <= t + longest edge.  The t-proportional part is what OKRT_DELTA covers; the
 * edge-proportional part is this slack: 2^-13 x longest edge x |E1||E2| / |E1 x E2| (capped at 64), the largest over the
 * leaf's triangles.  2^-13 ~ 1.2e-4 covers incidence down to ~0.3 degrees; every ACCEPTED hit is checked against it at run time. */
float leafSlack(const okrt_triangle* tris, uint32_t start, uint32_t count)
{
    double worst = 0.0;
    for (uint32_t t = start; t < start + count; t++) {
        const float (*p)[3] = tris[t].position;
        double e1[3], e2[3], e3[3];
        for (int k = 0; k < 3; k++) { e1[k] = (double)p[0][k] - p[2][k]; e2[k] = (double)p[1][k] - p[2][k]; e3[k] = (double)p[0][k] - p[1][k]; }
        const double l1 = std::sqrt(e1[0] * e1[0] + e1[1] * e1[1] + e1[2] * e1[2]), l2 = std::sqrt(e2[0] * e2[0] + e2[1] * e2[1] + e2[2] * e2[2]);
        const double l3 = std::sqrt(e3[0] * e3[0] + e3[1] * e3[1] + e3[2] * e3[2]);
        const double cx = e1[1] * e2[2] - e1[2] * e2[1], cy = e1[2] * e2[0] - e1[0] * e2[2], cz = e1[0] * e2[1] - e1[1] * e2[0];
        const double area2 = std::sqrt(cx * cx + cy * cy + cz * cz);
        const double shape = area2 > 0.0 ? std::min(64.0, std::max(1.0, l1 * l2 / area2)) : 64.0;
        const double longest = std::max(l1, std::max(l2, l3));
        if (std::isfinite(longest)) worst = std::max(worst, longest * shape);
    }
    const double s = worst * (1.0 / 8192.0);
    return s < 3.0e38 ? (float)s : 3.0e38f;
}


/* ------------------------------------------------------------------ wide layout (okrt_accel.h)
 * Collapses binary trees (the scene-wide SAH tree; the reference's own tree of an instance that is walked with a ray switch)
 * into 8-wide nodes with quantised child boxes. */
struct WideEmitter {
    std::vector<TopNode>& tn;
    const okrt_triangle* tris;
    FastLayout& out;
    std::vector<double> cull;          /* 6 per binary node: what a wide child box has to contain (leaf: exact box grown by its slack) */
    uint32_t deepest = 0;
    std::string error;

    static double area(const double* b)
    {
        const double dx = b[3] - b[0], dy = b[4] - b[1], dz = b[5] - b[2];
        return (dx < 0 || dy < 0 || dz < 0) ? 0.0 : 2.0 * (dx * dy + dy * dz + dz * dx);
    }
    /* leaves with more triangles than a leaf record can name are cut into a small subtree of records that repeat the box */
    void splitOversized(int32_t n)
    {
        if (tn[n].wcount <= kWideMaxLeafTris) return;
        const uint32_t half = tn[n].wcount / 2u;
        TopNode a = tn[n], c = tn[n];
        a.wcount = half; c.wstart = tn[n].wstart + half; c.wcount = tn[n].wcount - half;
        a.parent = c.parent = n;
        a.slack = leafSlack(tris, a.wstart, a.wcount); c.slack = leafSlack(tris, c.wstart, c.wcount);
        const int32_t ia = (int32_t)tn.size(); tn.push_back(a);
        const int32_t ic = (int32_t)tn.size(); tn.push_back(c);
        tn[n].left = ia; tn[n].right = ic; tn[n].wcount = 0;
        splitOversized(ia); splitOversized(ic);
    }
    void prepare(int32_t root)
    {
        /* post-order over the binary tree: culling boxes bottom-up */
        std::vector<int32_t> order, st{root};
        while (!st.empty()) {
            const int32_t n = st.back(); st.pop_back();
            if (tn[n].left < 0 && tn[n].wcount > kWideMaxLeafTris) splitOversized(n);
            order.push_back(n);
            if (tn[n].left >= 0) { st.push_back(tn[n].left); st.push_back(tn[n].right); }
        }
        cull.resize(tn.size() * 6);
        for (size_t i = order.size(); i-- > 0;) {
            const int32_t n = order[i];
            double* c = &cull[(size_t)n * 6];
            if (tn[n].left < 0) {
                const double s = tn[n].wcount ? (double)tn[n].slack : 0.0;
                for (int k = 0; k < 3; k++) { c[k] = tn[n].lo[k] - s; c[3 + k] = tn[n].hi[k] + s; }
            } else {
                const double *a = &cull[(size_t)tn[n].left * 6], *d = &cull[(size_t)tn[n].right * 6];
                for (int k = 0; k < 3; k++) { c[k] = std::min(a[k], d[k]); c[3 + k] = std::max(a[3 + k], d[3 + k]); }
            }
        }
    }
    uint32_t newNodes(uint32_t k)
    {
        const uint32_t first = (uint32_t)(out.wideNodes.size() / kWideNodeWords);
        out.wideNodes.resize(out.wideNodes.size() + (size_t)k * kWideNodeWords, 0u);
        out.wideLeaves.resize(out.wideLeaves.size() + (size_t)k * 8u * kWideLeafWords, 0u);
        return first;
    }
    bool fill(uint32_t node, int32_t bin, uint32_t level)
    {
        deepest = std::max(deepest, level + 1);
        /* children: open the interior child of the largest area until there are eight */
        int32_t child[8]; int n = 0;
        if (tn[bin].left < 0) child[n++] = bin;
        else { child[n++] = tn[bin].left; child[n++] = tn[bin].right; }
        while (n < 8) {
            int best = -1; double bestArea = -1.0;
            for (int i = 0; i < n; i++)
                if (tn[child[i]].left >= 0) { const double a = area(&cull[(size_t)child[i] * 6]); if (a > bestArea) { bestArea = a; best = i; } }
            if (best < 0) break;
            const int32_t open = child[best];
            child[best] = tn[open].left; child[n++] = tn[open].right;
        }
        double box[6] = {DBL_MAX, DBL_MAX, DBL_MAX, -DBL_MAX, -DBL_MAX, -DBL_MAX};
        for (int i = 0; i < n; i++) { const double* c = &cull[(size_t)child[i] * 6]; for (int k = 0; k < 3; k++) { box[k] = std::min(box[k], c[k]); box[3 + k] = std::max(box[3 + k], c[3 + k]); } }
        /* slots: slot s is visited first by rays of direction octant s (bit 2: -x, bit 1: -y, bit 0: -z), so it should hold the
         * child that lies farthest AGAINST that direction; greedy assignment of the cheapest (child, slot) pair */
        int slotOf[8]; bool slotUsed[8] = {false, false, false, false, false, false, false, false};
        {
            double cost[8][8];
            for (int i = 0; i < n; i++) {
                const double* c = &cull[(size_t)child[i] * 6];
                for (int s = 0; s < 8; s++) {
                    double v = 0.0;
                    for (int k = 0; k < 3; k++) { const double rel = 0.5 * (c[k] + c[3 + k]) - 0.5 * (box[k] + box[3 + k]); v += ((s >> (2 - k)) & 1) ? -rel : rel; }
                    cost[i][s] = v;
                }
            }
            bool done[8] = {false, false, false, false, false, false, false, false};
            for (int round = 0; round < n; round++) {
                int bi = -1, bs = -1; double bc = DBL_MAX;
                for (int i = 0; i < n; i++) if (!done[i]) for (int s = 0; s < 8; s++) if (!slotUsed[s] && cost[i][s] < bc) { bc = cost[i][s]; bi = i; bs = s; }
                if (bi < 0) { bi = 0; while (done[bi]) bi++; bs = 0; while (slotUsed[bs]) bs++; }       /* NaN boxes: any free slot */
                done[bi] = true; slotUsed[bs] = true; slotOf[bi] = bs;
            }
        }
        /* quantisation grid: cell 2^e per axis, origin a multiple of the cell, 255 cells cover the box, every plane exact in float32 */
        float p[3]; uint32_t eb[3]; double cell[3];
        for (int k = 0; k < 3; k++) {
            const double lo = box[k], hi = box[3 + k];
            if (!(lo <= hi) || !std::isfinite(lo) || !std::isfinite(hi)) { error = "wide layout: box is not finite"; return false; }
            int e = -100;
            if (hi > lo) e = std::max(e, (int)std::ceil(std::log2((hi - lo) / 255.0)));
            for (;; e++) {
                if (e > 100) { error = "wide layout: box too large to quantise"; return false; }
                const double c = std::ldexp(1.0, e), origin = std::floor(lo / c) * c;
                if ((hi - origin) / c > 255.0) continue;
                if (std::fabs(origin) / c + 256.0 >= 16777216.0) continue;      /* origin + q * cell must be a float32 */
                p[k] = (float)origin; cell[k] = c; eb[k] = (uint32_t)(e + 127);
                if ((double)p[k] != origin) continue;
                break;
            }
        }
        uint32_t* w = &out.wideNodes[(size_t)node * kWideNodeWords];
        uint8_t q[6][8];
        for (int s = 0; s < 8; s++) { for (int a = 0; a < 3; a++) { q[a][s] = 255; q[3 + a][s] = 0; } }
        uint32_t imask = 0u;
        for (int i = 0; i < n; i++) {
            const int s = slotOf[i];
            const double* c = &cull[(size_t)child[i] * 6];
            for (int k = 0; k < 3; k++) {
                double ql = std::floor((c[k] - (double)p[k]) / cell[k]), qh = std::ceil((c[3 + k] - (double)p[k]) / cell[k]);
                ql = std::min(255.0, std::max(0.0, ql)); qh = std::min(255.0, std::max(0.0, qh));
                if ((double)p[k] + ql * cell[k] > c[k] || (double)p[k] + qh * cell[k] < c[3 + k]) { error = "wide layout: quantised box does not contain its child"; return false; }
                q[k][s] = (uint8_t)ql; q[3 + k][s] = (uint8_t)qh;
            }
            if (tn[child[i]].left >= 0) imask |= 1u << s;
        }
        for (int k = 0; k < 3; k++) std::memcpy(&w[k], &p[k], 4);
        w[3] = eb[0] | (eb[1] << 8) | (eb[2] << 16) | (imask << 24);
        const uint32_t numInterior = (uint32_t)__builtin_popcount(imask);
        const uint32_t childBase = numInterior ? newNodes(numInterior) : 0u;
        w = &out.wideNodes[(size_t)node * kWideNodeWords];                     /* newNodes may have moved the array */
        w[4] = childBase; w[5] = 0u;
        for (int a = 0; a < 6; a++) {
            w[6 + 2 * a] = (uint32_t)q[a][0] | ((uint32_t)q[a][1] << 8) | ((uint32_t)q[a][2] << 16) | ((uint32_t)q[a][3] << 24);
            w[7 + 2 * a] = (uint32_t)q[a][4] | ((uint32_t)q[a][5] << 8) | ((uint32_t)q[a][6] << 16) | ((uint32_t)q[a][7] << 24);
        }
        for (int i = 0; i < n; i++) {
            const int s = slotOf[i];
            const TopNode& c = tn[child[i]];
            if (c.left >= 0) continue;
            uint32_t* lw = &out.wideLeaves[((size_t)node * 8u + (size_t)s) * kWideLeafWords];
            for (int k = 0; k < 3; k++) {                                      /* a triangle leaf's box is float32 already; an instance's world box rounds outward */
                float lo = (float)c.lo[k], hi = (float)c.hi[k];
                if ((double)lo > c.lo[k]) lo = std::nextafterf(lo, -INFINITY);
                if ((double)hi < c.hi[k]) hi = std::nextafterf(hi, INFINITY);
                std::memcpy(&lw[k], &lo, 4); std::memcpy(&lw[3 + k], &hi, 4);
            }
            if (c.wcount) {
                if (c.wstart >= (1u << 26)) { error = "wide layout: more than 2^26 triangles"; return false; }
                lw[6] = (c.wcount << 26) | c.wstart;
                std::memcpy(&lw[7], &c.slack, 4);
                out.maxSlack = std::max(out.maxSlack, c.slack);
            } else {
                lw[6] = (kWideLeafInstance << 26) | c.wstart;
                lw[7] = 0u;
            }
        }
        for (int i = 0; i < n; i++) {
            if (tn[child[i]].left < 0) continue;
            const uint32_t rank = (uint32_t)__builtin_popcount(imask & ((1u << slotOf[i]) - 1u));
            if (!fill(childBase + rank, child[i], level + 1)) return false;
        }
        return true;
    }
    /* returns the wide root of the binary tree under `root` and its depth in node levels */
    bool emitTree(int32_t root, uint32_t& wideRoot, uint32_t& depth)
    {
        prepare(root);
        deepest = 0;
        wideRoot = newNodes(1);
        if (!fill(wideRoot, root, 0)) return false;
        depth = deepest;
        return true;
    }
};

/* the reference's own tree of one mesh as a binary tree of its LEAVES (interior boxes are rebuilt from the leaves: the
 * reference's are supersets, flattenable() has checked that) */
int32_t appendReferenceTree(std::vector<TopNode>& tn, const okrt_triangle* tris, const okrt_bvh_node* nodes, uint32_t idx)
{
    const okrt_bvh_node& n = nodes[idx];
    if (n.firstChildIdx == OKRT_INVALID_UINT) {
        if (n.triStart == OKRT_INVALID_UINT || n.triEnd <= n.triStart) return -1;
        TopNode leaf;
        for (int k = 0; k < 3; k++) { leaf.lo[k] = n.boundingBox.min[k]; leaf.hi[k] = n.boundingBox.max[k]; }
        leaf.wstart = n.triStart; leaf.wcount = n.triEnd - n.triStart; leaf.slack = leafSlack(tris, n.triStart, n.triEnd - n.triStart);
        tn.push_back(leaf);
        return (int32_t)tn.size() - 1;
    }
    const int32_t l = appendReferenceTree(tn, tris, nodes, n.firstChildIdx), r = appendReferenceTree(tn, tris, nodes, n.firstChildIdx + 1);
    if (l < 0) return r;
    if (r < 0) return l;
    TopNode in;
    for (int k = 0; k < 3; k++) { in.lo[k] = std::min(tn[l].lo[k], tn[r].lo[k]); in.hi[k] = std::max(tn[l].hi[k], tn[r].hi[k]); }
    in.left = l; in.right = r;
    tn.push_back(in);
    const int32_t self = (int32_t)tn.size() - 1;
    tn[l].parent = self; tn[r].parent = self;
    return self;
}

inline float roundDown(double v) { float f = (float)v; return (double)f > v ? std::nextafterf(f, -INFINITY) : f; }
inline float roundUp(double v) { float f = (float)v; return (double)f < v ? std::nextafterf(f, INFINITY) : f; }

} // namespace


/* Validates the oct-tree arrays and derives what both layout builders need: parent per node, whether every box on the way up
 * from a node contains the one below it, which (reachable) node lists each instance, the reference traversal's stack need. */
bool octTreeChecks(const okrt_octree_node* octNodes, uint32_t numOctNodes, uint32_t numInstances, std::vector<uint32_t>& octParent,
                   std::vector<uint8_t>& octNested, std::vector<uint32_t>& instOct, uint32_t& exactOctDepth, std::string& error)
{
    octParent.assign(numOctNodes, 0xFFFFFFFFu);
    std::vector<uint8_t> reachable(numOctNodes, 0);
    exactOctDepth = 0;
    if (numOctNodes) {
        std::vector<uint32_t> stack{0u};
        size_t maxOcc = 1, visited = 0;
        while (!stack.empty()) {
            const uint32_t n = stack.back(); stack.pop_back();
            if (reachable[n]) { error = "oct-tree node reachable twice"; return false; }
            reachable[n] = 1;
            if (++visited > numOctNodes) { error = "oct-tree cycle"; return false; }
            const okrt_octree_node& on = octNodes[n];
            for (uint32_t c = 0; c < on.numChildren; c++) {
                const uint64_t ci = (uint64_t)on.firstChildIdx + c;
                if (ci >= numOctNodes) { error = "oct-tree child index out of range"; return false; }
                octParent[ci] = n;
                stack.push_back((uint32_t)ci);
            }
            maxOcc = std::max(maxOcc, stack.size());
        }
        exactOctDepth = (uint32_t)maxOcc;
    }
    /* octNested[n]: every box on the way from n up to the root contains the box below it (component-wise, exactly) */
    octNested.assign(numOctNodes, 0);
    for (uint32_t n = 0; n < numOctNodes; n++) {
        if (!reachable[n]) continue;
        bool nested = true;
        for (uint32_t c = n, p = octParent[n]; p != 0xFFFFFFFFu && nested; c = p, p = octParent[p])
            for (int k = 0; k < 3; k++)
                nested = nested && octNodes[p].boundingBox.min[k] <= octNodes[c].boundingBox.min[k] && octNodes[c].boundingBox.max[k] <= octNodes[p].boundingBox.max[k];
        octNested[n] = nested ? 1 : 0;
    }
    instOct.assign(numInstances, 0xFFFFFFFFu);
    for (uint32_t n = 0; n < numOctNodes; n++) {
        if (!reachable[n]) continue;
        const okrt_octree_node& on = octNodes[n];
        if (on.meshesStartIdx == OKRT_INVALID_UINT || on.meshesEndIdx == OKRT_INVALID_UINT) continue;
        if (on.meshesEndIdx > numInstances && on.meshesStartIdx < on.meshesEndIdx) { error = "oct-tree node lists instances out of range"; return false; }
        for (uint32_t i = on.meshesStartIdx; i < on.meshesEndIdx; i++) {
            if (instOct[i] != 0xFFFFFFFFu) { error = "instance listed by two oct-tree nodes"; return false; }
            instOct[i] = n;
        }
    }
    return true;
}

bool buildFastLayout(const okrt_triangle* tris, uint32_t numTris, const okrt_bvh_node* nodes, uint32_t numNodes,
                     const okrt_mesh_instance* instances, uint32_t numInstances,
                     const okrt_octree_node* octNodes, uint32_t numOctNodes, FastLayout& out, bool wantWide)
{
    out = FastLayout();
    const bool verbose = std::getenv("OKRT_LAYOUT_VERBOSE") != nullptr;
    auto phaseStart = std::chrono::steady_clock::now();
    auto phase = [&](const char* name) {
        const auto now = std::chrono::steady_clock::now();
        if (verbose) std::fprintf(stderr, "[okrt] layout: %-28s %8.1f ms\n", name, std::chrono::duration<double, std::milli>(now - phaseStart).count());
        phaseStart = now;
    };

    /* triangles: 3 x float4 */
    /* triangles: p2 | E1 = p0 - p2 | E2 = p1 - p2 (GPU-Utilities.hlsli:124: the same float32 subtractions, done once) */
    out.tris.assign((size_t)numTris * kTriQuads, Float4{0.f, 0.f, 0.f, 0.f});
    for (uint32_t t = 0; t < numTris; t++) {
        const float (*p)[3] = tris[t].position;
        volatile float e[6];
        for (int k = 0; k < 3; k++) { e[k] = p[0][k] - p[2][k]; e[3 + k] = p[1][k] - p[2][k]; }
        out.tris[(size_t)t * kTriQuads + 0] = Float4{p[2][0], p[2][1], p[2][2], e[0]};
        out.tris[(size_t)t * kTriQuads + 1] = Float4{e[1], e[2], e[3], e[4]};
        out.tris[(size_t)t * kTriQuads + 2] = Float4{e[5], 0.f, 0.f, 0.f};
    }

    phase("triangle records");
    /* oct-tree: parents, reachability, reference stack bound, which node lists each instance */
    std::vector<uint8_t> octNested;
    std::vector<uint32_t> instOct;
    if (!octTreeChecks(octNodes, numOctNodes, numInstances, out.octParent, octNested, instOct, out.exactOctDepth, out.error)) return false;
    phase("oct-tree checks");
    /* bottom level: one conversion per distinct BVH root */
    Builder b{tris, numTris, nodes, numNodes, out};
    std::unordered_map<uint32_t, uint32_t> rootRefOf;
    std::unordered_map<uint32_t, uint32_t> rootUsers;            /* reachable instances per BVH root */
    for (uint32_t i = 0; i < numInstances; i++)
        if (instOct[i] != 0xFFFFFFFFu) rootUsers[instances[i].bvhNodeStartIdx]++;
    out.triOwner.assign(numTris, 0xFFFFFFFFu);
    std::vector<TopLeaf> leaves;
    std::vector<std::pair<uint32_t, uint32_t>> wideInstanceRoots;   /* (instance record, reference BVH root) walked with a ray switch */
    uint32_t chainDepth = 0;                                     /* pair levels of oversized-leaf chains below a flattened leaf */
    for (uint32_t i = 0; i < numInstances; i++) {
        if (instOct[i] == 0xFFFFFFFFu) continue;                 /* never visited by the reference either */
        const okrt_mesh_instance& inst = instances[i];
        if (inst.bvhNodeStartIdx >= numNodes) { out.error = "instance bvhNodeStartIdx out of range"; return false; }
        /* pair-node copy of the instance's reference BVH: only instances that are walked as instances need it, so it
         * is made on demand (a merged, flattened instance contributes its leaves to the scene-wide tree instead) */
        auto convertedRoot = [&]() -> uint32_t {
            auto it = rootRefOf.find(inst.bvhNodeStartIdx);
            if (it != rootRefOf.end()) return it->second;
            const uint32_t before = b.deepestPair;
            b.deepestPair = 0;
            const uint32_t ref = b.convert(inst.bvhNodeStartIdx, 0, 0);
            b.deepestPair = std::max(before, b.deepestPair);
            rootRefOf.emplace(inst.bvhNodeStartIdx, ref);
            return ref;
        };
        const okrt_aabb& rb = nodes[inst.bvhNodeStartIdx].boundingBox;
        const float* T = inst.transformMatrix;
        const float* Ti = inst.inverseTransformMatrix;
        static const float I16[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
        bool identity = true;
        for (int k = 0; k < 16; k++) identity = identity && T[k] == I16[k] && Ti[k] == I16[k];
        const bool flatCandidate = identity && rootUsers[inst.bvhNodeStartIdx] == 1u && flattenable(nodes, numNodes, inst.bvhNodeStartIdx);
        uint32_t rootRef = kRefNone;
        if (flatCandidate) {
            if (!b.hasTriangles(inst.bvhNodeStartIdx)) continue; /* empty mesh */
        } else {
            rootRef = convertedRoot();
            if (b.failed) return false;
            if (rootRef == kRefNone) continue;                   /* empty mesh */
        }
        bool finite = true;
        for (int k = 0; k < 3; k++) finite = finite && std::isfinite(rb.min[k]) && std::isfinite(rb.max[k]) && rb.min[k] <= rb.max[k];
        for (int k = 0; k < 16; k++) finite = finite && std::isfinite(T[k]) && std::isfinite(Ti[k]);
        if (!finite) continue;                                   /* the reference's box test can never pass */

        /* rigidity: (T3x3)^T (T3x3) == I within 1e-4 — otherwise local and world distances differ and
         * the reference's local-vs-world culling (RaytracerCS.hlsl:285) is order dependent */
        for (int a = 0; a < 3; a++)
            for (int c = 0; c < 3; c++) {
                double d = 0;
                for (int r = 0; r < 3; r++) d += (double)T[r * 4 + a] * (double)T[r * 4 + c];
                if (std::fabs(d - (a == c ? 1.0 : 0.0)) > 1e-4) out.allRigid = false;
            }
        /* world box of the root box (padded unless the transform is exactly the identity): a primitive of the scene-wide tree,
         * and what the oct-tree boxes above the instance have to contain */
        double worldLo[3], worldHi[3];
        if (identity) { for (int k = 0; k < 3; k++) { worldLo[k] = rb.min[k]; worldHi[k] = rb.max[k]; } }
        else {
            for (int k = 0; k < 3; k++) { worldLo[k] = DBL_MAX; worldHi[k] = -DBL_MAX; }
            for (int corner = 0; corner < 8; corner++) {
                const double v[3] = {(corner & 1) ? rb.max[0] : rb.min[0], (corner & 2) ? rb.max[1] : rb.min[1], (corner & 4) ? rb.max[2] : rb.min[2]};
                for (int row = 0; row < 3; row++) {
                    const double p = (double)T[row * 4 + 0] * v[0] + (double)T[row * 4 + 1] * v[1] + (double)T[row * 4 + 2] * v[2] + (double)T[row * 4 + 3];
                    worldLo[row] = std::min(worldLo[row], p); worldHi[row] = std::max(worldHi[row], p);
                }
            }
            double diag = 0, mag = 0;
            for (int k = 0; k < 3; k++) { diag += (worldHi[k] - worldLo[k]) * (worldHi[k] - worldLo[k]); mag = std::max(mag, std::max(std::fabs(worldLo[k]), std::fabs(worldHi[k]))); }
            const double pad = 1e-4 * std::sqrt(diag) + 1e-5 * mag + 1e-6;
            for (int k = 0; k < 3; k++) { worldLo[k] -= pad; worldHi[k] += pad; }
        }
        /* The reference culls an oct-tree node AND everything below it when the ray enters the node's box at or beyond the
         * closest hit so far (RaytracerCS.hlsl:262-264), and createOctTree keeps untransformed occupant boxes (RayTracer.cpp:562)
         * that are rounded through centre / extents: a node's box need not contain what it lists.  The traversal does not rely
         * on containment: the distNear of every oct-tree box above the winning hit's instance joins the winner's gate
         * (okrt_wide.cuh, wideResult / wideEnterInstance), a missed box sends the ray to the reference order.
         * A BVH whose interior boxes do not contain their children (possible for uploaded arrays, never for BvhBuilder's) makes
         * the reference's own culls order dependent inside the mesh: such an instance is walked by the reference order only. */
        const bool orderDependent = !flattenable(nodes, numNodes, inst.bvhNodeStartIdx);
        /* An identity instance is MERGED into the scene-wide tree only if the oct-tree boxes above it contain its root box up to
         * the rounding of createOctTree's centre / extents arithmetic (a few ulp): then a ray enters those boxes no later than
         * (1 + ~1e-6) x the hit, which the traversal's window covers.  Otherwise it is walked as an instance, whose oct-tree
         * distNear is known before anything inside it is tested. */
        bool octContains = true;
        for (uint32_t on = instOct[i]; on != 0xFFFFFFFFu && octContains; on = out.octParent[on])
            for (int k = 0; k < 3; k++) {
                const double tol = 1e-5 * std::max(1.0, std::max(std::fabs(worldLo[k]), std::fabs(worldHi[k])));
                octContains = octContains && (double)octNodes[on].boundingBox.min[k] <= worldLo[k] + tol && worldHi[k] - tol <= (double)octNodes[on].boundingBox.max[k];
            }
        const uint32_t rec = (uint32_t)(out.instances.size() / kInstQuads);
        out.instances.resize(out.instances.size() + kInstQuads);
        Float4* r = &out.instances[(size_t)rec * kInstQuads];
        for (int row = 0; row < 3; row++) r[row] = Float4{Ti[row * 4 + 0], Ti[row * 4 + 1], Ti[row * 4 + 2], Ti[row * 4 + 3]};
        for (int row = 0; row < 3; row++) r[3 + row] = Float4{T[row * 4 + 0], T[row * 4 + 1], T[row * 4 + 2], rb.min[row]};
        r[6] = Float4{rb.max[0], rb.max[1], rb.max[2], bitsToFloat((identity ? 1u : 0u) | (orderDependent ? 2u : 0u))};
        r[7] = Float4{bitsToFloat(rootRef), bitsToFloat(instOct[i]), bitsToFloat(i), bitsToFloat(octNested[instOct[i]] ? 1u : 0u)};
        const okrt_aabb& ob = octNodes[instOct[i]].boundingBox;
        r[8] = Float4{ob.min[0], ob.min[1], ob.min[2], 0.f};
        r[9] = Float4{ob.max[0], ob.max[1], ob.max[2], 0.f};

        /* An identity-transform instance that is the only user of its mesh is MERGED into the top level: its
         * local ray is the same for all such instances (normalize(d*I), o*I), so the top-level tree can link
         * straight to its BVH root with the root's exact box — no ray switch, no exit marker.  Its triangles
         * remember their owner so that a hit can name the instance (and check its oct-tree chain). */
        bool merged = identity && rootUsers[inst.bvhNodeStartIdx] == 1u && octContains;
        std::vector<std::pair<uint32_t, uint32_t>> ranges;
        if (merged) {
            std::vector<uint32_t> st{inst.bvhNodeStartIdx};
            while (!st.empty() && merged) {
                const okrt_bvh_node& n = nodes[st.back()]; st.pop_back();
                if (n.firstChildIdx != OKRT_INVALID_UINT) { st.push_back(n.firstChildIdx); st.push_back(n.firstChildIdx + 1); continue; }
                if (n.triStart == OKRT_INVALID_UINT || n.triEnd <= n.triStart) continue;
                for (uint32_t j = n.triStart; j < n.triEnd; j++) if (out.triOwner[j] != 0xFFFFFFFFu) { merged = false; break; }
                ranges.emplace_back(n.triStart, n.triEnd);
            }
        }
        if (!merged && rootRef == kRefNone) {                    /* a flat candidate whose triangles are shared after all */
            rootRef = convertedRoot();
            if (b.failed) return false;
            r[7].x = bitsToFloat(rootRef);
        }
        if (merged) {
            for (auto& rg : ranges) for (uint32_t j = rg.first; j < rg.second; j++) out.triOwner[j] = rec;
            out.numMerged++;
            if (flatCandidate) {
                /* every REFERENCE LEAF becomes a primitive of one scene-wide SAH tree.  Its box is the reference's own
                 * (bit copy) and is tested with the reference's arithmetic; a reference parent box contains its
                 * children's boxes exactly (BvhBuilder.cpp:209-222 takes min/max over the same points), and the slab
                 * test is monotone under rounding, so "leaf box hit" implies that every ancestor the reference
                 * tests on its way down is hit too — the per-mesh interior nodes need not be visited. */
                std::vector<uint32_t> st{inst.bvhNodeStartIdx};
                while (!st.empty()) {
                    const okrt_bvh_node& n = nodes[st.back()]; st.pop_back();
                    if (n.firstChildIdx != OKRT_INVALID_UINT) { st.push_back(n.firstChildIdx); st.push_back(n.firstChildIdx + 1); continue; }
                    if (n.triStart == OKRT_INVALID_UINT || n.triEnd <= n.triStart) continue;
                    const uint32_t before = b.deepestPair;
                    b.deepestPair = 0;
                    TopLeaf leaf; leaf.ref = b.leafRef(n.triStart, n.triEnd - n.triStart, n.boundingBox, 0, 0);
                    chainDepth = std::max(chainDepth, b.deepestPair);
                    b.deepestPair = before;
                    if (b.failed) return false;
                    leaf.weight = n.triEnd - n.triStart;
                    leaf.wstart = n.triStart; leaf.wcount = n.triEnd - n.triStart; leaf.slack = wantWide ? leafSlack(tris, n.triStart, n.triEnd - n.triStart) : 0.f;
                    for (int k = 0; k < 3; k++) { leaf.lo[k] = n.boundingBox.min[k]; leaf.hi[k] = n.boundingBox.max[k]; }
                    leaves.push_back(leaf);
                    out.numFlattenedLeaves++;
                }
            } else {
                TopLeaf leaf; leaf.ref = rootRef; leaf.weight = std::max<uint32_t>(1u, inst.triEnd > inst.triStart ? inst.triEnd - inst.triStart : 1u);
                for (int k = 0; k < 3; k++) { leaf.lo[k] = rb.min[k]; leaf.hi[k] = rb.max[k]; }   /* exact: this IS the reference's root-box test */
                leaf.wstart = rec; leaf.wcount = 0;              /* wide layout: an instance record (its tree is not nested: flagged order dependent) */
                leaves.push_back(leaf);
            }
            continue;
        }

        /* the (padded) world box: the top level only has to be conservative */
        TopLeaf leaf; leaf.ref = kRefInstance | rec; leaf.weight = std::max<uint32_t>(1u, inst.triEnd > inst.triStart ? inst.triEnd - inst.triStart : 1u);
        for (int k = 0; k < 3; k++) { leaf.lo[k] = worldLo[k]; leaf.hi[k] = worldHi[k]; }
        leaf.wstart = rec; leaf.wcount = 0;
        if (!orderDependent) wideInstanceRoots.emplace_back(rec, inst.bvhNodeStartIdx);
        leaves.push_back(leaf);
    }
    out.exactBvhDepth = b.deepestLeafLevel + 1;
    phase("instances and leaves");

    /* top level: binned SAH (32 bins per axis, cost = area x triangles below) over reference leaves and instances;
     * median split when the SAH finds nothing or the tree gets too deep for the traversal stack */
    uint32_t topDepth = 0;
    constexpr int kBins = 32;
    constexpr uint32_t kSahDepthLimit = 44;
    const char* sweepEnv = std::getenv("OKRT_TOP_SWEEP");
    const size_t sweepBelow = sweepEnv ? (size_t)std::atoll(sweepEnv) : (size_t)-1;
    const char* sweepMinEnv = std::getenv("OKRT_TOP_SWEEP_MIN");
    const size_t sweepAbove = sweepMinEnv ? (size_t)std::atoll(sweepMinEnv) : 8192;     /* smaller ranges use the 32-bin SAH: the sweep pays at the top levels only (measured) */
    std::vector<TopNode> tn;
    tn.resize(leaves.size() * 2 + 2);                   /* slots are claimed with an atomic counter: subtrees are built in parallel */
    std::atomic<int32_t> nextTopNode{0};
    std::function<int32_t(size_t, size_t, uint32_t)> buildTop = [&](size_t lo, size_t hi, uint32_t level) -> int32_t {
        double bl[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, bh[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
        double cl[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, ch[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
        for (size_t i = lo; i < hi; i++)
            for (int k = 0; k < 3; k++) {
                bl[k] = std::min(bl[k], leaves[i].lo[k]); bh[k] = std::max(bh[k], leaves[i].hi[k]);
                const double c = 0.5 * (leaves[i].lo[k] + leaves[i].hi[k]);
                cl[k] = std::min(cl[k], c); ch[k] = std::max(ch[k], c);
            }
        const int32_t self = nextTopNode.fetch_add(1);
        for (int k = 0; k < 3; k++) { tn[self].lo[k] = bl[k]; tn[self].hi[k] = bh[k]; }
        if (hi - lo == 1) { tn[self].ref = leaves[lo].ref; tn[self].wstart = leaves[lo].wstart; tn[self].wcount = leaves[lo].wcount; tn[self].slack = leaves[lo].slack; return self; }

        size_t mid = lo;
        if (hi - lo > 2 && hi - lo <= sweepBelow && hi - lo >= sweepAbove && level < kSahDepthLimit) {
            /* small ranges: every split position of every axis (full sweep), not just the 31 bin borders */
            auto area = [](const double* l, const double* h) { const double dx = h[0] - l[0], dy = h[1] - l[1], dz = h[2] - l[2]; return 2.0 * (dx * dy + dy * dz + dz * dx); };
            const size_t n = hi - lo;
            double bestCost = DBL_MAX; int bestAxis = -1; size_t bestCut = 0;
            std::vector<double> rightArea(n); std::vector<uint64_t> rightW(n);
            std::vector<std::pair<double, uint32_t>> keyed(n), bestOrder;          /* (centroid x 2, position in the range) */
            for (int axis = 0; axis < 3; axis++) {
                for (size_t i = 0; i < n; i++) keyed[i] = {leaves[lo + i].lo[axis] + leaves[lo + i].hi[axis], (uint32_t)i};
                std::sort(keyed.begin(), keyed.end());                              /* ties by position: deterministic */
                double rl[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, rh[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX}; uint64_t rw = 0;
                for (size_t i = n - 1; i > 0; i--) {
                    const TopLeaf& q = leaves[lo + keyed[i].second];
                    for (int k = 0; k < 3; k++) { rl[k] = std::min(rl[k], q.lo[k]); rh[k] = std::max(rh[k], q.hi[k]); }
                    rw += q.weight; rightArea[i] = area(rl, rh); rightW[i] = rw;
                }
                double ll[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, lh[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX}; uint64_t lw = 0;
                bool improved = false;
                for (size_t i = 0; i + 1 < n; i++) {
                    const TopLeaf& q = leaves[lo + keyed[i].second];
                    for (int k = 0; k < 3; k++) { ll[k] = std::min(ll[k], q.lo[k]); lh[k] = std::max(lh[k], q.hi[k]); }
                    lw += q.weight;
                    const double cost = area(ll, lh) * (double)lw + rightArea[i + 1] * (double)rightW[i + 1];
                    if (cost < bestCost) { bestCost = cost; bestAxis = axis; bestCut = i + 1; improved = true; }
                }
                if (improved) bestOrder = keyed;
            }
            if (bestAxis >= 0) {
                std::vector<TopLeaf> moved(n);
                for (size_t i = 0; i < n; i++) moved[i] = leaves[lo + bestOrder[i].second];
                std::copy(moved.begin(), moved.end(), leaves.begin() + (ptrdiff_t)lo);
                mid = lo + bestCut;
            }
        } else if (hi - lo > 2 && level < kSahDepthLimit) {
            double bestCost = DBL_MAX; int bestAxis = -1, bestBin = -1;
            for (int axis = 0; axis < 3; axis++) {
                const double ext = ch[axis] - cl[axis];
                if (!(ext > 0)) continue;
                struct Bin { double lo[3], hi[3]; uint64_t w; } bins[kBins];
                for (auto& bn : bins) { for (int k = 0; k < 3; k++) { bn.lo[k] = DBL_MAX; bn.hi[k] = -DBL_MAX; } bn.w = 0; }
                const double scale = kBins / ext;
                for (size_t i = lo; i < hi; i++) {
                    const double c = 0.5 * (leaves[i].lo[axis] + leaves[i].hi[axis]);
                    const int bi = std::min(kBins - 1, std::max(0, (int)((c - cl[axis]) * scale)));
                    Bin& bn = bins[bi];
                    for (int k = 0; k < 3; k++) { bn.lo[k] = std::min(bn.lo[k], leaves[i].lo[k]); bn.hi[k] = std::max(bn.hi[k], leaves[i].hi[k]); }
                    bn.w += leaves[i].weight;
                }
                double rightArea[kBins]; uint64_t rightW[kBins];
                double rl[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, rh[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX}; uint64_t rw = 0;
                auto area = [](const double* l, const double* h) { const double dx = h[0] - l[0], dy = h[1] - l[1], dz = h[2] - l[2]; return (dx < 0 || dy < 0 || dz < 0) ? 0.0 : 2.0 * (dx * dy + dy * dz + dz * dx); };
                for (int bi = kBins - 1; bi > 0; bi--) {
                    for (int k = 0; k < 3; k++) { rl[k] = std::min(rl[k], bins[bi].lo[k]); rh[k] = std::max(rh[k], bins[bi].hi[k]); }
                    rw += bins[bi].w; rightArea[bi] = area(rl, rh); rightW[bi] = rw;
                }
                double ll[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, lh[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX}; uint64_t lw = 0;
                for (int bi = 0; bi < kBins - 1; bi++) {
                    for (int k = 0; k < 3; k++) { ll[k] = std::min(ll[k], bins[bi].lo[k]); lh[k] = std::max(lh[k], bins[bi].hi[k]); }
                    lw += bins[bi].w;
                    if (lw == 0 || rightW[bi + 1] == 0) continue;
                    const double cost = area(ll, lh) * (double)lw + rightArea[bi + 1] * (double)rightW[bi + 1];
                    if (cost < bestCost) { bestCost = cost; bestAxis = axis; bestBin = bi; }
                }
            }
            if (bestAxis >= 0) {
                const double scale = kBins / (ch[bestAxis] - cl[bestAxis]), c0 = cl[bestAxis];
                const int axis = bestAxis, split = bestBin;
                auto it = std::partition(leaves.begin() + (ptrdiff_t)lo, leaves.begin() + (ptrdiff_t)hi, [=](const TopLeaf& a) {
                    const double c = 0.5 * (a.lo[axis] + a.hi[axis]);
                    return std::min(kBins - 1, std::max(0, (int)((c - c0) * scale))) <= split;
                });
                mid = (size_t)(it - leaves.begin());
            }
        }
        if (mid == lo || mid == hi) {
            int axis = 0;
            if (ch[1] - cl[1] > ch[axis] - cl[axis]) axis = 1;
            if (ch[2] - cl[2] > ch[axis] - cl[axis]) axis = 2;
            mid = lo + (hi - lo) / 2;
            std::nth_element(leaves.begin() + (ptrdiff_t)lo, leaves.begin() + (ptrdiff_t)mid, leaves.begin() + (ptrdiff_t)hi,
                             [axis](const TopLeaf& a, const TopLeaf& c) { return a.lo[axis] + a.hi[axis] < c.lo[axis] + c.hi[axis]; });
        }
        int32_t l, r;
        if (hi - lo > 8192 && level < 4) {                      /* disjoint ranges of `leaves`, disjoint slots of `tn` */
            auto left = std::async(std::launch::async, [&]() { return buildTop(lo, mid, level + 1); });
            r = buildTop(mid, hi, level + 1);
            l = left.get();
        } else { l = buildTop(lo, mid, level + 1); r = buildTop(mid, hi, level + 1); }
        tn[self].left = l; tn[self].right = r;
        tn[l].parent = self; tn[r].parent = self;
        return self;
    };
    /* pair nodes in depth-first order; a leaf's box stays the reference's own floats, interior boxes (exact unions) go one
     * ulp outward, which keeps "child hit => parent hit" safe in every rounding case */
    std::function<uint32_t(int32_t, uint32_t, okrt_aabb&)> emitTop = [&](int32_t n, uint32_t level, okrt_aabb& box) -> uint32_t {
        for (int k = 0; k < 3; k++) { box.min[k] = roundDown(tn[n].lo[k]); box.max[k] = roundUp(tn[n].hi[k]); }
        if (tn[n].left < 0) return tn[n].ref;
        for (int k = 0; k < 3; k++) { box.min[k] = std::nextafterf(box.min[k], -INFINITY); box.max[k] = std::nextafterf(box.max[k], INFINITY); }
        const uint32_t p = b.newPair();
        topDepth = std::max(topDepth, level + 1);
        okrt_aabb b0, b1;
        const uint32_t r0 = emitTop(tn[n].left, level + 1, b0);
        const uint32_t r1 = emitTop(tn[n].right, level + 1, b1);
        b.setChild(p, 0, b0, r0);
        b.setChild(p, 1, b1, r1);
        return p;
    };
    int32_t wideBinaryRoot = -1;
    if (!leaves.empty()) {
        okrt_aabb rootBox;
        int32_t root = buildTop(0, leaves.size(), 0);
        tn.resize((size_t)nextTopNode.load());
        const char* passesEnv = std::getenv("OKRT_TOP_REINSERT");
        const int passes = passesEnv ? std::atoi(passesEnv) : 0;
        if (passes > 0 && leaves.size() > 2) {
            const std::vector<TopNode> before = tn;
            const int32_t rootBefore = root;
            root = reinsertOptimize(tn, root, passes);
            if (topTreeDepth(tn, root) > kSahDepthLimit + 4) { tn = before; root = rootBefore; }      /* stack budget */
        }
        phase("scene-wide tree (SAH)");
        wideBinaryRoot = root;
        out.topRef = emitTop(root, 0, rootBox);
        phase("pair-node emission");
        if (leaves.size() == 1) {
            /* single instance: wrap it in a pair node so the kernel always starts at an interior node (and tests its box) */
            const uint32_t p = b.newPair();
            okrt_aabb empty; for (int k = 0; k < 3; k++) { empty.min[k] = FLT_MAX; empty.max[k] = -FLT_MAX; }
            b.setChild(p, 0, rootBox, out.topRef);
            b.setChild(p, 1, empty, kRefNone);
            out.topRef = p;
            topDepth = 1;
        }
    }
    /* wide layout: the same scene-wide tree collapsed into 8-wide quantised nodes; instances that are walked with a ray switch
     * get their own wide tree (one per distinct mesh), its root in r8.w of the record */
    if (!leaves.empty() && wantWide) {
        WideEmitter we{tn, tris, out};
        uint32_t topWideDepth = 0, instWideDepth = 0;
        if (!we.emitTree(wideBinaryRoot, out.wideRoot, topWideDepth)) { out.error = we.error; return false; }
        std::unordered_map<uint32_t, uint32_t> wideRootOf;
        for (const auto& ir : wideInstanceRoots) {
            auto it = wideRootOf.find(ir.second);
            if (it == wideRootOf.end()) {
                const int32_t bin = appendReferenceTree(tn, tris, nodes, ir.second);
                uint32_t wr = kRefNone, depth = 0;
                if (bin >= 0 && !we.emitTree(bin, wr, depth)) { out.error = we.error; return false; }
                instWideDepth = std::max(instWideDepth, depth);
                it = wideRootOf.emplace(ir.second, wr).first;
            }
            out.instances[(size_t)ir.first * kInstQuads + 8].w = bitsToFloat(it->second);
        }
        out.wideDepth = topWideDepth;
        /* sentinel + one deferred group per level + (at an instance: the two groups of the level, the exit marker) */
        out.wideStackDepth = 1 + topWideDepth + 3 + instWideDepth;
        phase("wide nodes");
    }
    if ((out.nodes.size() / 4) >= (1u << 30) || (out.instances.size() / kInstQuads) >= (1u << 30)) { out.error = "scene too large for 30-bit references"; return false; }

    /* fast stack: one deferred sibling per level in each tree, plus the instance-exit marker */
    out.maxStackDepth = topDepth + std::max(b.deepestPair, chainDepth) + 2;
    if (b.failed) return false;
    for (uint32_t t = 0; t < numTris; t++) out.tris[(size_t)t * kTriQuads + 2].y = bitsToFloat(out.triOwner[t]);
    return true;
}

} // namespace okrt

/* ------------------------------------------------------------------ okrt_layout_report (include/okrt.h) */
extern "C" OKRT_API int okrt_layout_report(const okrt_triangle* tris, uint32_t numTris, const okrt_bvh_node* nodes, uint32_t numNodes,
                                           const okrt_mesh_instance* instances, uint32_t numInstances,
                                           const okrt_octree_node* octNodes, uint32_t numOctNodes, okrt_layout_info* out)
{
    using namespace okrt;
    if (!out) { setError("okrt_layout_report: null output"); return OKRT_ERR_INVALID_ARGUMENT; }
    std::memset(out, 0, sizeof(*out));
    FastLayout fl;
    const auto t0 = std::chrono::steady_clock::now();
    if (!buildFastLayout(tris, numTris, nodes, numNodes, instances, numInstances, octNodes, numOctNodes, fl)) {
        setError("scene rejected: " + fl.error);
        return OKRT_ERR_INVALID_ARGUMENT;
    }
    out->buildMs = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    out->numPairNodes = (uint32_t)(fl.nodes.size() / 4);
    out->numFlattenedLeaves = fl.numFlattenedLeaves; out->numMergedInstances = fl.numMerged;
    out->numInstanceRecords = (uint32_t)(fl.instances.size() / kInstQuads);
    out->fastStackDepth = fl.maxStackDepth; out->exactBvhDepth = fl.exactBvhDepth; out->exactOctDepth = fl.exactOctDepth;
    /* expected work of a random ray through the tree reachable from topRef (an instance counts as one triangle) */
    auto boxArea = [](const Float4& lo, const Float4& hi) {
        const double dx = (double)hi.x - lo.x, dy = (double)hi.y - lo.y, dz = (double)hi.z - lo.z;
        return (dx >= 0 && dy >= 0 && dz >= 0) ? 2.0 * (dx * dy + dy * dz + dz * dx) : 0.0;
    };
    if (fl.topRef != kRefNone && (fl.topRef & kRefTagMask) == kRefInterior) {
        std::vector<uint32_t> stack{fl.topRef};
        const Float4* rq = &fl.nodes[(size_t)fl.topRef * 4];
        const Float4 rlo{std::fmin(rq[0].x, rq[2].x), std::fmin(rq[0].y, rq[2].y), std::fmin(rq[0].z, rq[2].z), 0.f};
        const Float4 rhi{std::fmax(rq[1].x, rq[3].x), std::fmax(rq[1].y, rq[3].y), std::fmax(rq[1].z, rq[3].z), 0.f};
        const double rootArea = boxArea(rlo, rhi);
        double sum = 0.0, interior = 0.0;
        while (!stack.empty() && rootArea > 0.0) {
            const uint32_t p = stack.back(); stack.pop_back();
            const Float4* q = &fl.nodes[(size_t)p * 4];
            for (int c = 0; c < 2; c++) {
                uint32_t ref; std::memcpy(&ref, c ? &q[1].w : &q[0].w, 4);
                if (ref == kRefNone) continue;
                const double a = boxArea(q[2 * c], q[2 * c + 1]) / rootArea;
                const uint32_t tag = ref & kRefTagMask;
                if (tag == kRefInterior) { sum += a; interior += a; stack.push_back(ref); }
                else if (tag == kRefTriLeaf) sum += a * (double)(((ref >> 27) & 7u) + 1u);
                else sum += a;
            }
        }
        out->sahCost = sum;
        if (std::getenv("OKRT_LAYOUT_VERBOSE")) std::fprintf(stderr, "[okrt] layout: interior term %.4f, leaf term %.4f\n", interior, sum - interior);
    }
    return OKRT_OK;
}

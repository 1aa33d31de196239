/*
 * okrt_host.cpp — host-side builders of the path-tracing core (rows a10-a13 of SURVEY.md §8).
 *
 * These produce, bit for bit, the arrays the reference's CPU code uploads to the GPU:
 *   okrt_geometry_build  <->  RayTracer::loadMeshAndBvhData (source/Graphics/RayTracer.cpp:120-217)
 *                             + BvhBuilder (source/Graphics/BvhBuilder.cpp:12-230)
 *   okrt_octree_build    <->  RayTracer::createOctTree / loadOctTree (RayTracer.cpp:523-788)
 *   okrt_calculate_projection_data <-> RayTracer::calculateProjectionData (RayTracer.cpp:268-308)
 *   okrt_scale_texture / okrt_environment_cross_to_faces <-> RayTracer.cpp:310-432, DX11.cpp:232-270
 *
 * The BVH builder is NOT the reference's algorithm re-typed: the reference evaluates each of the
 * 3 x 99 SAH candidate planes with a full pass over the node's triangles (O(297 N) per node).  Here
 * every triangle is binned once per axis by the exact index of the first candidate plane that puts
 * it on the left, and the 99 (count, box) pairs are recovered by prefix/suffix unions.  min/max and
 * integer counts are order-independent, and the cost expression is evaluated with the same float32
 * operations in the same order, so the chosen plane, the partition and therefore the whole node
 * array are identical to the reference's — at O(N log 99 + 99) per axis.  Large nodes are binned by
 * several threads.  tests/test_bvh_builder.py checks bit-equality against the reference's own
 * BvhBuilder.cpp (oracle/_ref) and the committed golden fixtures.
 *
 * Compiled with -ffp-contract=off: float32 results must not depend on FMA contraction.
 */
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "okrt.h"
#include "okrt_internal.h"

namespace okrt {

thread_local std::string g_lastError;
void setError(const std::string& msg) { g_lastError = msg; }

/* ---------------------------------------------------------------- boxes */
struct Box {
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX};      /* Okay::AABB defaults, Utilities.h:69-70 */
    float hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    inline void grow(const float* p)
    {
        for (int k = 0; k < 3; k++) { lo[k] = p[k] < lo[k] ? p[k] : lo[k]; hi[k] = hi[k] < p[k] ? p[k] : hi[k]; }
    }
    inline void merge(const Box& b)
    {
        for (int k = 0; k < 3; k++) { lo[k] = b.lo[k] < lo[k] ? b.lo[k] : lo[k]; hi[k] = hi[k] < b.hi[k] ? b.hi[k] : hi[k]; }
    }
    /* Okay::AABB::getArea, Utilities.h:78-82 — an empty box yields +inf, which the cost rule relies on */
    inline float area() const
    {
        const float ex = hi[0] - lo[0], ey = hi[1] - lo[1], ez = hi[2] - lo[2];
        return ex * ey + ey * ez + ez * ex;
    }
};

static inline void growTriangle(Box& b, const okrt_triangle& t)
{
    b.grow(t.position[0]); b.grow(t.position[1]); b.grow(t.position[2]);
}

/* ----------------------------------------------------------- BVH builder */
namespace {

constexpr uint32_t kPlanes = 99;        /* NUM_TESTS - 1, BvhBuilder.cpp:74,84 */
constexpr uint32_t kBins = kPlanes + 1;

struct Bins {
    Box box[kBins];
    uint32_t count[kBins];
    void clear() { for (uint32_t i = 0; i < kBins; i++) { box[i] = Box(); count[i] = 0; } }
};

struct BuildNode {
    Box bounds;
    uint32_t begin = 0, end = 0;        /* range in the working index array (stable order) */
    uint32_t firstChild = OKRT_INVALID_UINT;
};

struct SplitChoice { float cost = FLT_MAX; uint32_t axis = 0; float pos = 0.f; };

class SahBuilder {
public:
    SahBuilder(const okrt_triangle* tris, uint32_t n, uint32_t maxDepth, uint32_t maxLeaf)
        : tris_(tris), n_(n), maxDepth_(maxDepth), maxLeaf_(maxLeaf)
    {
        centroid_.resize((size_t)n * 3);
        /* OkayMath::getMiddle, SMath.h:9-12 */
        for (uint32_t i = 0; i < n; i++)
            for (int k = 0; k < 3; k++)
                centroid_[(size_t)i * 3 + k] =
                    ((tris[i].position[0][k] + tris[i].position[1][k]) + tris[i].position[2][k]) * (1.f / 3.f);
        order_.resize(n);
        scratch_.resize(n);
        for (uint32_t i = 0; i < n; i++) order_[i] = i;
        unsigned hw = std::thread::hardware_concurrency();
        workers_ = hw ? std::min(hw, 16u) : 1u;
    }

    void run(const okrt_aabb& rootBounds)
    {
        nodes_.clear();
        BuildNode root;
        for (int k = 0; k < 3; k++) { root.bounds.lo[k] = rootBounds.min[k]; root.bounds.hi[k] = rootBounds.max[k]; }
        root.begin = 0; root.end = n_;
        nodes_.push_back(root);

        struct Pending { uint32_t node, depth; };
        std::vector<Pending> todo;
        todo.push_back({0u, 0u});
        while (!todo.empty()) {
            const Pending cur = todo.back();
            todo.pop_back();
            const uint32_t count = nodes_[cur.node].end - nodes_[cur.node].begin;
            /* BvhBuilder.cpp:143: uint32 arithmetic, maxDepth 0 wraps like the reference */
            if (cur.depth >= maxDepth_ - 1u || count <= maxLeaf_)
                continue;

            const SplitChoice best = chooseSplit(nodes_[cur.node]);
            const float parentCost = count * nodes_[cur.node].bounds.area();
            if (best.cost >= parentCost)
                continue;

            const uint32_t mid = partition(nodes_[cur.node], best);
            if (mid == nodes_[cur.node].begin || mid == nodes_[cur.node].end)
                continue;

            BuildNode left, right;
            left.begin = nodes_[cur.node].begin; left.end = mid;
            right.begin = mid; right.end = nodes_[cur.node].end;
            fit(left); fit(right);
            const uint32_t first = (uint32_t)nodes_.size();
            nodes_[cur.node].firstChild = first;
            nodes_.push_back(left);
            nodes_.push_back(right);
            todo.push_back({first, cur.depth + 1u});
            todo.push_back({first + 1u, cur.depth + 1u});
        }
    }

    /* RayTracer.cpp:172-199: leaves claim [triStart,triEnd) in node order */
    uint32_t emit(okrt_bvh_node* outNodes, uint32_t* outTriOrder) const
    {
        uint32_t cursor = 0;
        for (size_t k = 0; k < nodes_.size(); k++) {
            const BuildNode& bn = nodes_[k];
            okrt_bvh_node& g = outNodes[k];
            for (int a = 0; a < 3; a++) { g.boundingBox.min[a] = bn.bounds.lo[a]; g.boundingBox.max[a] = bn.bounds.hi[a]; }
            g.triStart = g.triEnd = OKRT_INVALID_UINT;
            g.firstChildIdx = bn.firstChild;
            if (bn.firstChild != OKRT_INVALID_UINT) continue;
            g.triStart = cursor;
            g.triEnd = cursor + (bn.end - bn.begin);
            for (uint32_t i = bn.begin; i < bn.end; i++) outTriOrder[cursor++] = order_[i];
        }
        return (uint32_t)nodes_.size();
    }

    size_t numNodes() const { return nodes_.size(); }

private:
    const okrt_triangle* tris_;
    uint32_t n_, maxDepth_, maxLeaf_;
    std::vector<float> centroid_;
    std::vector<uint32_t> order_, scratch_;
    std::vector<BuildNode> nodes_;
    unsigned workers_;

    void fit(BuildNode& nd) const
    {
        nd.bounds = Box();
        for (uint32_t i = nd.begin; i < nd.end; i++) growTriangle(nd.bounds, tris_[order_[i]]);
    }

    /* number of candidate planes p[1..99] with p[i] <= c, p non-decreasing */
    static inline uint32_t planesNotAbove(const float* planes, float c)
    {
        uint32_t lo = 0, hi = kPlanes;            /* answer in [0, 99] */
        while (lo < hi) {
            const uint32_t m = (lo + hi) >> 1;    /* planes[m] is candidate m+1 */
            if (planes[m] <= c) lo = m + 1; else hi = m;
        }
        return lo;
    }

    void binRange(uint32_t begin, uint32_t end, uint32_t axis, const float* planes, Bins& bins) const
    {
        for (uint32_t i = begin; i < end; i++) {
            const uint32_t t = order_[i];
            const float c = centroid_[(size_t)t * 3 + axis];
            /* NaN centroids compare false with `<` on every plane => always right => bin 99 */
            const uint32_t b = (c == c) ? planesNotAbove(planes, c) : kPlanes;
            bins.count[b]++;
            growTriangle(bins.box[b], tris_[t]);
        }
    }

    /* the reference's evaluateSAH for one plane (BvhBuilder.cpp:38-70), used only if the candidate
     * positions of an axis are not monotone (never observed; kept for exactness) */
    float bruteCost(const BuildNode& nd, uint32_t axis, float pos) const
    {
        Box l, r; uint32_t lc = 0, rc = 0;
        for (uint32_t i = nd.begin; i < nd.end; i++) {
            const uint32_t t = order_[i];
            if (centroid_[(size_t)t * 3 + axis] < pos) { lc++; growTriangle(l, tris_[t]); }
            else { rc++; growTriangle(r, tris_[t]); }
        }
        const float cost = lc * l.area() + rc * r.area();
        return cost > 0.f ? cost : FLT_MAX;
    }

    SplitChoice chooseSplit(const BuildNode& nd) const
    {
        SplitChoice best;
        const uint32_t count = nd.end - nd.begin;
        for (uint32_t axis = 0; axis < 3; axis++) {
            const float bmin = nd.bounds.lo[axis], bmax = nd.bounds.hi[axis];
            if (bmin == bmax) continue;

            float planes[kPlanes];
            bool monotone = true;
            for (uint32_t i = 1; i <= kPlanes; i++) {
                const float a = i / (float)100;
                planes[i - 1] = bmin * (1.0f - a) + bmax * a;      /* glm::mix */
                if (i > 1 && !(planes[i - 2] <= planes[i - 1])) monotone = false;
            }
            if (!monotone) {
                for (uint32_t i = 1; i <= kPlanes; i++) {
                    const float cost = bruteCost(nd, axis, planes[i - 1]);
                    if (cost < best.cost) { best.cost = cost; best.axis = axis; best.pos = planes[i - 1]; }
                }
                continue;
            }

            Bins bins; bins.clear();
            const unsigned nthreads = count >= 65536u ? workers_ : 1u;
            if (nthreads <= 1) {
                binRange(nd.begin, nd.end, axis, planes, bins);
            } else {
                std::vector<Bins> part(nthreads);
                std::vector<std::thread> pool;
                const uint32_t chunk = (count + nthreads - 1) / nthreads;
                for (unsigned w = 0; w < nthreads; w++) {
                    part[w].clear();
                    const uint32_t b = nd.begin + std::min(count, w * chunk), e = nd.begin + std::min(count, (w + 1) * chunk);
                    pool.emplace_back([this, b, e, axis, &planes, &part, w]() { binRange(b, e, axis, planes, part[w]); });
                }
                for (auto& th : pool) th.join();
                for (unsigned w = 0; w < nthreads; w++)
                    for (uint32_t k = 0; k < kBins; k++) { bins.count[k] += part[w].count[k]; bins.box[k].merge(part[w].box[k]); }
            }

            /* suffix unions: right side of plane i = bins i..99 */
            Box rightBox[kBins + 1]; uint32_t rightCount[kBins + 1];
            rightBox[kBins] = Box(); rightCount[kBins] = 0;
            for (int k = (int)kBins - 1; k >= 0; k--) {
                rightBox[k] = rightBox[k + 1]; rightBox[k].merge(bins.box[k]);
                rightCount[k] = rightCount[k + 1] + bins.count[k];
            }
            Box leftBox; uint32_t leftCount = 0;
            for (uint32_t i = 1; i <= kPlanes; i++) {
                leftBox.merge(bins.box[i - 1]); leftCount += bins.count[i - 1];   /* bins 0..i-1 */
                float cost = leftCount * leftBox.area() + rightCount[i] * rightBox[i].area();
                cost = cost > 0.f ? cost : FLT_MAX;
                if (cost < best.cost) { best.cost = cost; best.axis = axis; best.pos = planes[i - 1]; }
            }
        }
        return best;
    }

    /* stable partition by `centroid < pos` (BvhBuilder.cpp:157-173) */
    uint32_t partition(const BuildNode& nd, const SplitChoice& sc)
    {
        uint32_t l = nd.begin, r = 0;
        for (uint32_t i = nd.begin; i < nd.end; i++) {
            const uint32_t t = order_[i];
            if (centroid_[(size_t)t * 3 + sc.axis] < sc.pos) order_[l++] = t;
            else scratch_[r++] = t;
        }
        std::memcpy(order_.data() + l, scratch_.data(), (size_t)r * sizeof(uint32_t));
        return l;
    }
};

} // namespace

int buildBvh(const okrt_triangle* tris, uint32_t numTriangles, const okrt_aabb& rootBounds, uint32_t maxDepth,
             uint32_t maxLeaf, std::vector<okrt_bvh_node>& nodes, std::vector<uint32_t>& triOrder)
{
    SahBuilder b(tris, numTriangles, maxDepth, maxLeaf);
    b.run(rootBounds);
    nodes.resize(b.numNodes());
    triOrder.resize(numTriangles);
    b.emit(nodes.data(), triOrder.data());
    return OKRT_OK;
}

/* --------------------------------------------------------- Transform math
 * Column-major 4x4 like glm (m[c][r]); every expression keeps glm 0.9.9.9's operation order. */
struct Mat4 { float m[4][4]; };

static Mat4 identity()
{
    Mat4 r; std::memset(&r, 0, sizeof(r));
    r.m[0][0] = r.m[1][1] = r.m[2][2] = r.m[3][3] = 1.f;
    return r;
}

static Mat4 mul(const Mat4& a, const Mat4& b)
{
    Mat4 r;
    for (int c = 0; c < 4; c++)
        for (int row = 0; row < 4; row++)
            r.m[c][row] = ((a.m[0][row] * b.m[c][0] + a.m[1][row] * b.m[c][1]) + a.m[2][row] * b.m[c][2]) + a.m[3][row] * b.m[c][3];
    return r;
}

static inline float radiansOf(float deg) { return deg * 0.01745329251994329576923690768489f; }

static void rotationColumns(const float* rotationDeg, float R[3][3])
{
    const float ex = radiansOf(rotationDeg[0]), ey = radiansOf(rotationDeg[1]), ez = radiansOf(rotationDeg[2]);
    const float cx = std::cos(ex * 0.5f), cy = std::cos(ey * 0.5f), cz = std::cos(ez * 0.5f);
    const float sx = std::sin(ex * 0.5f), sy = std::sin(ey * 0.5f), sz = std::sin(ez * 0.5f);
    const float w = cx * cy * cz + sx * sy * sz;
    const float x = sx * cy * cz - cx * sy * sz;
    const float y = cx * sy * cz + sx * cy * sz;
    const float z = cx * cy * sz - sx * sy * cz;
    const float xx = x * x, yy = y * y, zz = z * z, xz = x * z, xy = x * y, yz = y * z, wx = w * x, wy = w * y, wz = w * z;
    R[0][0] = 1.f - 2.f * (yy + zz); R[0][1] = 2.f * (xy + wz);       R[0][2] = 2.f * (xz - wy);
    R[1][0] = 2.f * (xy - wz);       R[1][1] = 1.f - 2.f * (xx + zz); R[1][2] = 2.f * (yz + wx);
    R[2][0] = 2.f * (xz + wy);       R[2][1] = 2.f * (yz - wx);       R[2][2] = 1.f - 2.f * (xx + yy);
}

/* Transform::calculateMatrix, Components.h:13-18: translate(I,p) * toMat4(quat(radians(r))) * scale(I,s) */
static Mat4 composeTransform(const okrt_transform& t)
{
    Mat4 T = identity();
    for (int row = 0; row < 4; row++) {
        const Mat4 I = identity();
        T.m[3][row] = ((I.m[0][row] * t.position[0] + I.m[1][row] * t.position[1]) + I.m[2][row] * t.position[2]) + I.m[3][row];
    }
    float R3[3][3];
    rotationColumns(t.rotation, R3);
    Mat4 R = identity();
    for (int c = 0; c < 3; c++) for (int row = 0; row < 3; row++) R.m[c][row] = R3[c][row];
    Mat4 S = identity();                       /* glm::scale: column c of I times scale[c] */
    for (int c = 0; c < 3; c++) for (int row = 0; row < 4; row++) S.m[c][row] = S.m[c][row] * t.scale[c];
    return mul(mul(T, R), S);
}

/* glm::inverse(mat4): cofactor expansion in glm's association order */
static Mat4 inverse(const Mat4& M)
{
    const float (*m)[4] = M.m;
    const float c00 = m[2][2] * m[3][3] - m[3][2] * m[2][3], c02 = m[1][2] * m[3][3] - m[3][2] * m[1][3], c03 = m[1][2] * m[2][3] - m[2][2] * m[1][3];
    const float c04 = m[2][1] * m[3][3] - m[3][1] * m[2][3], c06 = m[1][1] * m[3][3] - m[3][1] * m[1][3], c07 = m[1][1] * m[2][3] - m[2][1] * m[1][3];
    const float c08 = m[2][1] * m[3][2] - m[3][1] * m[2][2], c10 = m[1][1] * m[3][2] - m[3][1] * m[1][2], c11 = m[1][1] * m[2][2] - m[2][1] * m[1][2];
    const float c12 = m[2][0] * m[3][3] - m[3][0] * m[2][3], c14 = m[1][0] * m[3][3] - m[3][0] * m[1][3], c15 = m[1][0] * m[2][3] - m[2][0] * m[1][3];
    const float c16 = m[2][0] * m[3][2] - m[3][0] * m[2][2], c18 = m[1][0] * m[3][2] - m[3][0] * m[1][2], c19 = m[1][0] * m[2][2] - m[2][0] * m[1][2];
    const float c20 = m[2][0] * m[3][1] - m[3][0] * m[2][1], c22 = m[1][0] * m[3][1] - m[3][0] * m[1][1], c23 = m[1][0] * m[2][1] - m[2][0] * m[1][1];
    const float f0[4] = {c00, c00, c02, c03}, f1[4] = {c04, c04, c06, c07}, f2[4] = {c08, c08, c10, c11};
    const float f3[4] = {c12, c12, c14, c15}, f4[4] = {c16, c16, c18, c19}, f5[4] = {c20, c20, c22, c23};
    const float v0[4] = {m[1][0], m[0][0], m[0][0], m[0][0]}, v1[4] = {m[1][1], m[0][1], m[0][1], m[0][1]};
    const float v2[4] = {m[1][2], m[0][2], m[0][2], m[0][2]}, v3[4] = {m[1][3], m[0][3], m[0][3], m[0][3]};
    static const float sa[4] = {+1, -1, +1, -1}, sb[4] = {-1, +1, -1, +1};
    Mat4 inv;
    for (int i = 0; i < 4; i++) {
        inv.m[0][i] = ((v1[i] * f0[i] - v2[i] * f1[i]) + v3[i] * f2[i]) * sa[i];
        inv.m[1][i] = ((v0[i] * f0[i] - v2[i] * f3[i]) + v3[i] * f4[i]) * sb[i];
        inv.m[2][i] = ((v0[i] * f1[i] - v1[i] * f3[i]) + v3[i] * f5[i]) * sa[i];
        inv.m[3][i] = ((v0[i] * f2[i] - v1[i] * f4[i]) + v2[i] * f5[i]) * sb[i];
    }
    const float d0 = m[0][0] * inv.m[0][0], d1 = m[0][1] * inv.m[1][0], d2 = m[0][2] * inv.m[2][0], d3 = m[0][3] * inv.m[3][0];
    const float oneOverDet = 1.f / ((d0 + d1) + (d2 + d3));
    for (int c = 0; c < 4; c++) for (int row = 0; row < 4; row++) inv.m[c][row] *= oneOverDet;
    return inv;
}

static Mat4 transpose(const Mat4& a)
{
    Mat4 r;
    for (int c = 0; c < 4; c++) for (int row = 0; row < 4; row++) r.m[c][row] = a.m[row][c];
    return r;
}

static void normalize3(float* v)
{
    const float inv = 1.f / std::sqrt((v[0] * v[0] + v[1] * v[1]) + v[2] * v[2]);
    v[0] *= inv; v[1] *= inv; v[2] *= inv;
}

} // namespace okrt

using namespace okrt;

/* =============================================================== C ABI */
/* struct okrt_geometry: okrt_internal.h (shared with the device builder) */

struct okrt_octree {
    std::vector<okrt_octree_node> nodes;
    std::vector<okrt_mesh_instance> instances;
};

extern "C" {

const char* okrt_version(void) { return "okrt 0.1 (sm_100a)"; }
const char* okrt_last_error(void) { return g_lastError.c_str(); }

void okrt_material_default(okrt_material* m)
{
    if (!m) return;
    std::memset(m, 0, sizeof(*m));
    m->albedoTextureIdx = m->roughnessTextureIdx = m->metallicTextureIdx = m->specularTextureIdx = OKRT_INVALID_UINT;
    m->normalMapIdx = OKRT_INVALID_UINT;
    m->roughness = 1.f; m->specular = 1.f;
    m->emissionColour[0] = m->emissionColour[1] = m->emissionColour[2] = 1.f;
    m->indexOfRefraction = 1.f;
}

void okrt_options_default(okrt_options* o)
{
    if (!o) return;
    std::memset(o, 0, sizeof(*o));
    o->maxBounces = 1;                 /* NUM_BOUNCES at HEAD */
    o->traversal = OKRT_TRAVERSAL_FAST;
    o->shardCount = 1;
    o->frameStride = 1;
}

void okrt_render_data_default(okrt_render_data* rd)
{
    if (!rd) return;
    std::memset(rd, 0, sizeof(*rd));
    rd->accumulationEnabled = 1u;
    rd->debugMaxCount = 500u;
    static const float I[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    std::memcpy(rd->cameraInverseProjectionMatrix, I, 64);
    std::memcpy(rd->cameraInverseViewMatrix, I, 64);
}

int okrt_build_bvh(const okrt_triangle* tris, uint32_t numTriangles, const okrt_aabb* rootBounds, uint32_t maxDepth,
                   uint32_t maxLeafTriangles, okrt_bvh_node* outNodes, uint32_t nodeCapacity, uint32_t* outNumNodes,
                   uint32_t* outTriOrder)
{
    if ((!tris && numTriangles) || !rootBounds || !outNodes || !outNumNodes || !outTriOrder) {
        setError("okrt_build_bvh: null argument");
        return OKRT_ERR_INVALID_ARGUMENT;
    }
    std::vector<okrt_bvh_node> nodes; std::vector<uint32_t> order;
    buildBvh(tris, numTriangles, *rootBounds, maxDepth, maxLeafTriangles, nodes, order);
    if (nodes.size() > nodeCapacity) { setError("okrt_build_bvh: node capacity too small"); return OKRT_ERR_CAPACITY; }
    std::memcpy(outNodes, nodes.data(), nodes.size() * sizeof(okrt_bvh_node));
    if (numTriangles) std::memcpy(outTriOrder, order.data(), (size_t)numTriangles * sizeof(uint32_t));
    *outNumNodes = (uint32_t)nodes.size();
    return OKRT_OK;
}

int okrt_geometry_build(const okrt_mesh_view* meshes, uint32_t numMeshes, uint32_t maxDepth, uint32_t maxLeafTriangles,
                        okrt_geometry** out)
{
    if (!out || (!meshes && numMeshes)) { setError("okrt_geometry_build: null argument"); return OKRT_ERR_INVALID_ARGUMENT; }
    okrt_geometry* g = new okrt_geometry();
    size_t total = 0;
    for (uint32_t i = 0; i < numMeshes; i++) total += meshes[i].numTriangles;
    g->triPos.reserve(total); g->triInfo.reserve(total);
    g->descs.resize(numMeshes); g->bounds.resize(numMeshes);

    uint32_t triBase = 0;
    std::vector<okrt_bvh_node> nodes; std::vector<uint32_t> order;
    for (uint32_t i = 0; i < numMeshes; i++) {
        const okrt_mesh_view& mv = meshes[i];
        if (mv.numTriangles && (!mv.trianglesPos || !mv.trianglesInfo)) {
            delete g; setError("okrt_geometry_build: mesh with null triangle arrays"); return OKRT_ERR_INVALID_ARGUMENT;
        }
        buildBvh(mv.trianglesPos, mv.numTriangles, mv.boundingBox, maxDepth, maxLeafTriangles, nodes, order);
        const uint32_t nodeBase = (uint32_t)g->nodes.size();
        for (const okrt_bvh_node& n : nodes) {
            okrt_bvh_node gn = n;
            if (gn.firstChildIdx != OKRT_INVALID_UINT) gn.firstChildIdx += nodeBase;        /* tryOffsetIdx */
            else { gn.triStart += triBase; gn.triEnd += triBase; }
            g->nodes.push_back(gn);
        }
        for (uint32_t k = 0; k < mv.numTriangles; k++) {
            g->triPos.push_back(mv.trianglesPos[order[k]]);
            g->triInfo.push_back(mv.trianglesInfo[order[k]]);
        }
        g->descs[i] = okrt_mesh_desc{triBase, triBase + mv.numTriangles, nodeBase, (uint32_t)nodes.size()};
        g->bounds[i] = mv.boundingBox;
        triBase += mv.numTriangles;
    }
    *out = g;
    return OKRT_OK;
}

void okrt_geometry_free(okrt_geometry* g) { delete g; }
uint32_t okrt_geometry_num_triangles(const okrt_geometry* g) { return g ? (uint32_t)g->triPos.size() : 0; }
uint32_t okrt_geometry_num_nodes(const okrt_geometry* g) { return g ? (uint32_t)g->nodes.size() : 0; }
uint32_t okrt_geometry_num_meshes(const okrt_geometry* g) { return g ? (uint32_t)g->descs.size() : 0; }
const okrt_triangle* okrt_geometry_triangles_pos(const okrt_geometry* g) { return g ? g->triPos.data() : nullptr; }
const okrt_triangle_info* okrt_geometry_triangles_info(const okrt_geometry* g) { return g ? g->triInfo.data() : nullptr; }
const okrt_bvh_node* okrt_geometry_bvh_nodes(const okrt_geometry* g) { return g ? g->nodes.data() : nullptr; }
const okrt_mesh_desc* okrt_geometry_mesh_descs(const okrt_geometry* g) { return g ? g->descs.data() : nullptr; }
const okrt_aabb* okrt_geometry_mesh_bounds(const okrt_geometry* g) { return g ? g->bounds.data() : nullptr; }

void okrt_transform_matrix(const okrt_transform* t, float outRowMajor[16])
{
    const Mat4 M = composeTransform(*t);
    for (int r = 0; r < 4; r++) for (int c = 0; c < 4; c++) outRowMajor[r * 4 + c] = M.m[c][r];
}

/* ------------------------------------------------------------ oct-tree */
namespace {

struct Occupant { uint32_t id; Box box; };      /* id < numMeshEntities: mesh entity; else sphere entity */

struct Cell {
    Box box;
    std::vector<Occupant> occupants;
    uint32_t child[8];
    Cell() { for (uint32_t& c : child) c = OKRT_INVALID_UINT; }
};

/* Okay::AABB::intersects(a, b), Utilities.h:87-93: strict `<` on every axis */
bool overlaps(const Box& a, const Box& b)
{
    for (int k = 0; k < 3; k++) {
        const float centreGap = std::fabs((b.hi[k] + b.lo[k]) * 0.5f - (a.hi[k] + a.lo[k]) * 0.5f);
        const float reach = (a.hi[k] - a.lo[k]) * 0.5f + (b.hi[k] - b.lo[k]) * 0.5f;
        if (!(centreGap < reach)) return false;
    }
    return true;
}

} // namespace

int okrt_octree_build(const okrt_geometry* g, const okrt_mesh_entity* meshEntities, uint32_t numMeshEntities,
                      const okrt_sphere_entity* sphereEntities, uint32_t numSphereEntities, uint32_t maxDepth,
                      uint32_t maxLeafObjects, okrt_octree** out)
{
    if (!g || !out || (!meshEntities && numMeshEntities) || (!sphereEntities && numSphereEntities)) {
        setError("okrt_octree_build: null argument");
        return OKRT_ERR_INVALID_ARGUMENT;
    }
    for (uint32_t e = 0; e < numMeshEntities; e++)
        if (meshEntities[e].meshID >= g->descs.size()) { setError("okrt_octree_build: meshID out of range"); return OKRT_ERR_INVALID_ARGUMENT; }

    std::vector<Cell> cells(1);
    cells.reserve(128);

    /* root: union of transformed mesh-box corners and sphere boxes (RayTracer.cpp:539-577); every
     * occupant keeps its UNtransformed mesh-local box (:562) */
    for (uint32_t e = 0; e < numMeshEntities; e++) {
        const okrt_aabb& mb = g->bounds[meshEntities[e].meshID];
        const Mat4 M = composeTransform(meshEntities[e].transform);
        float centre[3], half[3];
        for (int k = 0; k < 3; k++) { centre[k] = (mb.min[k] + mb.max[k]) * 0.5f; half[k] = (mb.min[k] - mb.max[k]) * 0.5f; }
        for (int corner = 0; corner < 8; corner++) {
            /* DirectX::BoundingBox::GetCorners: half * (+-1) + centre */
            const float sx = (corner & 1) ? 1.f : -1.f, sy = (corner & 2) ? 1.f : -1.f, sz = (corner & 4) ? 1.f : -1.f;
            const float v[3] = {half[0] * sx + centre[0], half[1] * sy + centre[1], half[2] * sz + centre[2]};
            float p[3];
            for (int r = 0; r < 3; r++) p[r] = (M.m[0][r] * v[0] + M.m[1][r] * v[1]) + (M.m[2][r] * v[2] + M.m[3][r] * 1.f);
            cells[0].box.grow(p);
        }
        Occupant o; o.id = e;
        for (int k = 0; k < 3; k++) { o.box.lo[k] = mb.min[k]; o.box.hi[k] = mb.max[k]; }
        cells[0].occupants.push_back(o);
    }
    for (uint32_t e = 0; e < numSphereEntities; e++) {
        Occupant o; o.id = numMeshEntities + e;
        for (int k = 0; k < 3; k++) {
            o.box.lo[k] = -sphereEntities[e].radius + sphereEntities[e].transform.position[k];
            o.box.hi[k] = sphereEntities[e].radius + sphereEntities[e].transform.position[k];
        }
        cells[0].box.grow(o.box.lo); cells[0].box.grow(o.box.hi);
        cells[0].occupants.push_back(o);
    }

    static const float kOctant[8][3] = {{1, -1, 1}, {-1, -1, 1}, {1, 1, 1}, {-1, 1, 1}, {1, -1, -1}, {-1, -1, -1}, {1, 1, -1}, {-1, 1, -1}};
    struct Visit { uint32_t cell, depth; };
    std::vector<Visit> todo{{0u, 0u}};
    std::vector<Occupant> movable, taken;

    while (!todo.empty()) {
        const Visit v = todo.back();
        todo.pop_back();
        if (v.depth >= maxDepth || cells[v.cell].occupants.size() <= maxLeafObjects)
            continue;

        /* half-size box about the centre (RayTracer.cpp:621-632) */
        Box half = cells[v.cell].box;
        float quarter[3];
        for (int k = 0; k < 3; k++) {
            const float c = (half.hi[k] + half.lo[k]) * 0.5f;
            half.hi[k] -= c; half.lo[k] -= c;
            half.lo[k] *= 0.5f; half.hi[k] *= 0.5f;
            half.hi[k] += c; half.lo[k] += c;
            quarter[k] = (half.hi[k] - half.lo[k]) * 0.5f;
        }

        /* only occupants small relative to the cell may move down; scanned back to front (:634-643) */
        movable.clear();
        const float cellArea = cells[v.cell].box.area();
        const size_t population = cells[v.cell].occupants.size();
        for (size_t k = population; k-- > 0;)
            if (cells[v.cell].occupants[k].box.area() / cellArea < 0.75f)
                movable.push_back(cells[v.cell].occupants[k]);

        for (uint32_t oct = 0; oct < 8u; oct++) {
            Box octBox = half;
            for (int k = 0; k < 3; k++) { octBox.hi[k] += kOctant[oct][k] * quarter[k]; octBox.lo[k] += kOctant[oct][k] * quarter[k]; }

            /* first fit, back to front, removing from the pool (:655-670) */
            taken.clear();
            for (size_t k = movable.size(); k-- > 0;) {
                if (!overlaps(octBox, movable[k].box)) continue;
                taken.push_back(movable[k]);
                movable.erase(movable.begin() + (ptrdiff_t)k);
            }
            if (taken.empty() || taken.size() == population)
                continue;

            std::vector<Occupant>& stay = cells[v.cell].occupants;
            for (const Occupant& t : taken) {
                auto it = std::find_if(stay.begin(), stay.end(), [&](const Occupant& o) { return o.id == t.id; });
                if (it != stay.end()) stay.erase(it);
            }

            Cell child;
            child.occupants = taken;
            for (const Occupant& o : child.occupants) { child.box.grow(o.box.lo); child.box.grow(o.box.hi); }   /* refit :503-519 */
            cells.push_back(std::move(child));
            cells[v.cell].child[oct] = (uint32_t)cells.size() - 1u;
            todo.push_back({(uint32_t)cells.size() - 1u, v.depth + 1u});
        }
    }

    /* loadOctTree (RayTracer.cpp:706-788) */
    okrt_octree* t = new okrt_octree();
    t->nodes.resize(cells.size());
    for (size_t i = 0; i < cells.size(); i++) {
        okrt_octree_node& n = t->nodes[i];
        for (int k = 0; k < 3; k++) { n.boundingBox.min[k] = cells[i].box.lo[k]; n.boundingBox.max[k] = cells[i].box.hi[k]; }
        n.spheresStartIdx = n.spheresEndIdx = OKRT_INVALID_UINT;
        n.firstChildIdx = OKRT_INVALID_UINT;
        n.numChildren = 0;
        for (uint32_t c : cells[i].child) {
            if (c == OKRT_INVALID_UINT) continue;
            if (n.firstChildIdx == OKRT_INVALID_UINT) n.firstChildIdx = c;
            n.numChildren++;
        }
        n.meshesStartIdx = (uint32_t)t->instances.size();
        for (const Occupant& o : cells[i].occupants) {
            if (o.id >= numMeshEntities) continue;
            const okrt_mesh_entity& me = meshEntities[o.id];
            okrt_mesh_instance inst;
            const Mat4 uploaded = transpose(composeTransform(me.transform));       /* glm::transpose(T), :747 */
            const Mat4 uploadedInv = inverse(uploaded);                            /* glm::inverse(that), :749 */
            std::memcpy(inst.transformMatrix, uploaded.m, 64);
            std::memcpy(inst.inverseTransformMatrix, uploadedInv.m, 64);
            inst.triStart = g->descs[me.meshID].startIdx;
            inst.triEnd = g->descs[me.meshID].endIdx;
            inst.boundingBox = g->bounds[me.meshID];
            inst.material = me.material;
            inst.bvhNodeStartIdx = g->descs[me.meshID].bvhTreeStartIdx;
            t->instances.push_back(inst);
        }
        n.meshesEndIdx = (uint32_t)t->instances.size();
    }
    *out = t;
    return OKRT_OK;
}

void okrt_octree_free(okrt_octree* t) { delete t; }
uint32_t okrt_octree_num_nodes(const okrt_octree* t) { return t ? (uint32_t)t->nodes.size() : 0; }
uint32_t okrt_octree_num_instances(const okrt_octree* t) { return t ? (uint32_t)t->instances.size() : 0; }
const okrt_octree_node* okrt_octree_nodes(const okrt_octree* t) { return t ? t->nodes.data() : nullptr; }
const okrt_mesh_instance* okrt_octree_instances(const okrt_octree* t) { return t ? t->instances.data() : nullptr; }

/* -------------------------------------------------------------- camera */
int okrt_calculate_projection_data(const okrt_camera* cam, uint32_t width, uint32_t height, okrt_render_data* rd)
{
    if (!cam || !rd || !width || !height) { setError("okrt_calculate_projection_data: bad argument"); return OKRT_ERR_INVALID_ARGUMENT; }
    const float w = (float)width, h = (float)height;
    const float aspect = w / h;
    const float halfPlane = std::tan(radiansOf(cam->fov * 0.5f)) * cam->nearZ;
    rd->viewPlaneDims[0] = halfPlane * 2.f;
    rd->viewPlaneDims[1] = rd->viewPlaneDims[0] / aspect;
    std::memcpy(rd->cameraPosition, cam->position, 12);
    rd->cameraNearZ = cam->nearZ;

    /* glm::perspectiveFovLH (x64 config: LH, depth -1..1) */
    const float rad = radiansOf(cam->fov);
    const float ph = std::cos(0.5f * rad) / std::sin(0.5f * rad);
    const float pw = ph * h / w;
    Mat4 P; std::memset(&P, 0, sizeof(P));
    P.m[0][0] = pw; P.m[1][1] = ph;
    P.m[2][2] = (cam->farZ + cam->nearZ) / (cam->farZ - cam->nearZ);
    P.m[2][3] = 1.f;
    P.m[3][2] = -(2.f * cam->farZ * cam->nearZ) / (cam->farZ - cam->nearZ);
    const Mat4 Pinv = inverse(P);
    for (int r = 0; r < 4; r++) for (int c = 0; c < 4; c++) rd->cameraInverseProjectionMatrix[r * 4 + c] = Pinv.m[c][r];

    float R3[3][3];
    rotationColumns(cam->rotation, R3);
    float fwd[3] = {R3[2][0], R3[2][1], R3[2][2]}, right[3] = {R3[0][0], R3[0][1], R3[0][2]}, up[3] = {R3[1][0], R3[1][1], R3[1][2]};
    normalize3(fwd); normalize3(right); normalize3(up);
    std::memcpy(rd->cameraRightDir, right, 12);
    std::memcpy(rd->cameraUpDir, up, 12);

    /* glm::lookAtLH(eye, eye + fwd, +Y) */
    const float* eye = cam->position;
    float f[3] = {(eye[0] + fwd[0]) - eye[0], (eye[1] + fwd[1]) - eye[1], (eye[2] + fwd[2]) - eye[2]};
    normalize3(f);
    float s[3] = {1.f * f[2] - f[1] * 0.f, 0.f * f[0] - f[2] * 0.f, 0.f * f[1] - f[0] * 1.f};   /* cross((0,1,0), f) */
    normalize3(s);
    const float u[3] = {f[1] * s[2] - s[1] * f[2], f[2] * s[0] - s[2] * f[0], f[0] * s[1] - s[0] * f[1]};
    Mat4 V = identity();
    V.m[0][0] = s[0]; V.m[1][0] = s[1]; V.m[2][0] = s[2];
    V.m[0][1] = u[0]; V.m[1][1] = u[1]; V.m[2][1] = u[2];
    V.m[0][2] = f[0]; V.m[1][2] = f[1]; V.m[2][2] = f[2];
    V.m[3][0] = -((s[0] * eye[0] + s[1] * eye[1]) + s[2] * eye[2]);
    V.m[3][1] = -((u[0] * eye[0] + u[1] * eye[1]) + u[2] * eye[2]);
    V.m[3][2] = -((f[0] * eye[0] + f[1] * eye[1]) + f[2] * eye[2]);
    const Mat4 Vinv = inverse(V);
    for (int r = 0; r < 4; r++) for (int c = 0; c < 4; c++) rd->cameraInverseViewMatrix[r * 4 + c] = Vinv.m[c][r];
    return OKRT_OK;
}

/* ------------------------------------------------------------ sharding */
/* longest-processing-time-first: bands by decreasing cost (ties: lower band first), each to the least-loaded rank (ties: lower rank) */
int okrt_shard_balance(const double* bandCost, uint32_t numBands, uint32_t shardCount, uint32_t* ownerOfBand)
{
    if (!bandCost || !ownerOfBand || shardCount == 0u) { okrt::setError("okrt_shard_balance: bad argument"); return OKRT_ERR_INVALID_ARGUMENT; }
    std::vector<uint32_t> order(numBands);
    for (uint32_t b = 0; b < numBands; b++) order[b] = b;
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t c) { return bandCost[a] > bandCost[c]; });
    std::vector<double> load(shardCount, 0.0);
    for (uint32_t b : order) {
        uint32_t best = 0;
        for (uint32_t r = 1; r < shardCount; r++) if (load[r] < load[best]) best = r;
        ownerOfBand[b] = best;
        load[best] += bandCost[b] > 0.0 ? bandCost[b] : 0.0;
    }
    return OKRT_OK;
}

uint32_t okrt_shard_owner_of_row(uint32_t y, uint32_t bandRows, uint32_t shardCount)
{
    if (bandRows == 0u) bandRows = 8u;
    if (shardCount == 0u) shardCount = 1u;
    return (y / bandRows) % shardCount;
}

uint32_t okrt_shard_seed_frame(const okrt_options* o, uint32_t localFrame, uint32_t accumulationEnabled)
{
    const uint32_t stride = o && o->frameStride ? o->frameStride : 1u;
    const uint32_t offset = o ? o->frameOffset : 0u;
    return accumulationEnabled ? offset + (localFrame - 1u) * stride + 1u : offset;
}

/* ------------------------------------------------------------ textures */
uint32_t okrt_texture_array_size(const uint32_t* widths, const uint32_t* heights, uint32_t count)
{
    if (!count || !widths || !heights) return 0;
    uint32_t area = 0;
    for (uint32_t i = 0; i < count; i++) area += widths[i] * heights[i];           /* uint32 wrap as the reference */
    return (uint32_t)std::sqrt((double)(area / count));
}

static inline uint8_t quantise(float x)
{
    x = std::fmin(std::fmax(x, 0.f), 1.f);
    return (uint8_t)(x * 255.0f + 0.5f);
}

int okrt_scale_texture(const uint8_t* src, uint32_t sw, uint32_t sh, uint8_t* dst, uint32_t dw, uint32_t dh)
{
    if (!src || !dst || !sw || !sh || !dw || !dh) { setError("okrt_scale_texture: bad argument"); return OKRT_ERR_INVALID_ARGUMENT; }
    if (sw == dw && sh == dh) { std::memcpy(dst, src, (size_t)sw * sh * 4); return OKRT_OK; }
    std::vector<int> xa(dw), xb(dw); std::vector<float> xw(dw);
    for (uint32_t x = 0; x < dw; x++) {
        const float fx = ((float)x + 0.5f) / (float)dw * (float)sw - 0.5f;
        const float fl = std::floor(fx);
        xw[x] = fx - fl;
        xa[x] = std::min(std::max((int)fl, 0), (int)sw - 1);
        xb[x] = std::min(std::max((int)fl + 1, 0), (int)sw - 1);
    }
    for (uint32_t y = 0; y < dh; y++) {
        const float fy = ((float)y + 0.5f) / (float)dh * (float)sh - 0.5f;
        const float fl = std::floor(fy);
        const float wy = fy - fl;
        const int ya = std::min(std::max((int)fl, 0), (int)sh - 1), yb = std::min(std::max((int)fl + 1, 0), (int)sh - 1);
        const uint8_t* rowA = src + (size_t)ya * sw * 4;
        const uint8_t* rowB = src + (size_t)yb * sw * 4;
        uint8_t* o = dst + (size_t)y * dw * 4;
        for (uint32_t x = 0; x < dw; x++) {
            const float wx = xw[x];
            for (int ch = 0; ch < 4; ch++) {
                const float c00 = rowA[xa[x] * 4 + ch] / 255.0f, c10 = rowA[xb[x] * 4 + ch] / 255.0f;
                const float c01 = rowB[xa[x] * 4 + ch] / 255.0f, c11 = rowB[xb[x] * 4 + ch] / 255.0f;
                const float top = c00 * (1.0f - wx) + c10 * wx, bot = c01 * (1.0f - wx) + c11 * wx;
                o[x * 4 + ch] = quantise(top * (1.0f - wy) + bot * wy);
            }
        }
    }
    return OKRT_OK;
}

int okrt_environment_cross_to_faces(const uint8_t* cross, uint32_t imgW, uint32_t imgH, uint8_t* outFaces6)
{
    if (!cross || !outFaces6 || imgW < 4 || imgH < 3) { setError("okrt_environment_cross_to_faces: bad argument"); return OKRT_ERR_INVALID_ARGUMENT; }
    const uint32_t fw = imgW / 4u, fh = imgH / 3u;
    /* (column, row) of each face in the 4x3 cross: +X, -X, +Y, -Y, +Z, -Z (RayTracer.cpp:377-399) */
    static const uint32_t cell[6][2] = {{2, 1}, {0, 1}, {1, 0}, {1, 2}, {1, 1}, {3, 1}};
    for (int f = 0; f < 6; f++)
        for (uint32_t y = 0; y < fh; y++)
            std::memcpy(outFaces6 + (((size_t)f * fh + y) * fw) * 4,
                        cross + (((size_t)(cell[f][1] * fh + y)) * imgW + cell[f][0] * fw) * 4, (size_t)fw * 4);
    return OKRT_OK;
}

/* ------------------------------------------------------------ debug overlay (row f4)
 * DebugRenderer::renderBvhNodeBBs / renderOctTreeNodeBBs (DebugRenderer.cpp:264-303, 335-364) draw the 12 edges of
 * node boxes as a line list through DebugBBVS.hlsl:13-33 (unit cube scaled to the box, object matrix, camera
 * view-projection of updateCameraData :366-380), in the order executeDrawMode (:447-506) visits the nodes, with the
 * colours of drawNodeBoundingBox / drawOctTreeNodeBoundingBox (:407-445), no depth test.  Here the host produces the
 * clipped screen-space segments (double arithmetic, then float32) and the device rasterises them (okrt_draw_overlay). */
int okrt_overlay_build(uint32_t kind, const void* nodeArray, uint32_t numNodes, uint32_t nodeIdx, uint32_t drawMode,
                       const okrt_transform* objectTransform, const okrt_camera* cam, uint32_t width, uint32_t height, okrt_overlay** out)
{
    if (!out || !cam || !width || !height || kind > OKRT_OVERLAY_OCTREE || drawMode > OKRT_DRAW_WITH_DESCENDANTS || (!nodeArray && numNodes)) {
        setError("okrt_overlay_build: bad argument"); return OKRT_ERR_INVALID_ARGUMENT;
    }
    okrt_overlay* ov = new okrt_overlay();
    *out = ov;
    if (drawMode == OKRT_DRAW_NONE || nodeIdx >= numNodes) return OKRT_OK;
    const okrt_bvh_node* bvh = kind == OKRT_OVERLAY_BVH ? (const okrt_bvh_node*)nodeArray : nullptr;
    const okrt_octree_node* oct = kind == OKRT_OVERLAY_OCTREE ? (const okrt_octree_node*)nodeArray : nullptr;
    auto firstChild = [&](uint32_t i) { return bvh ? bvh[i].firstChildIdx : oct[i].firstChildIdx; };
    auto childCount = [&](uint32_t i) { return bvh ? 2u : oct[i].numChildren; };
    /* executeDrawMode: the node order of the draws */
    std::vector<uint32_t> order;
    if (drawMode == OKRT_DRAW_SINGLE) order.push_back(nodeIdx);
    else if (drawMode == OKRT_DRAW_WITH_CHILDREN) {
        order.push_back(nodeIdx);
        /* the caller passes numChildren of the selected node (2 for a BVH node) :299,:359-360 */
        if (firstChild(nodeIdx) != OKRT_INVALID_UINT) for (uint32_t c = 0; c < childCount(nodeIdx); c++) order.push_back(firstChild(nodeIdx) + c);
    } else {
        std::vector<uint32_t> st{nodeIdx};
        while (!st.empty()) {
            const uint32_t cur = st.back(); st.pop_back();
            if (cur >= numNodes) { delete ov; *out = nullptr; setError("okrt_overlay_build: child index out of range"); return OKRT_ERR_INVALID_ARGUMENT; }
            if (firstChild(cur) != OKRT_INVALID_UINT) for (uint32_t c = 0; c < childCount(cur); c++) st.push_back(firstChild(cur) + c);
            order.push_back(cur);
        }
    }
    /* camera: the float32 basis of updateCameraData, then double */
    float R3[3][3];
    rotationColumns(cam->rotation, R3);
    float f[3] = {(cam->position[0] + R3[2][0]) - cam->position[0], (cam->position[1] + R3[2][1]) - cam->position[1], (cam->position[2] + R3[2][2]) - cam->position[2]};
    normalize3(f);
    float sv[3] = {1.f * f[2] - f[1] * 0.f, 0.f * f[0] - f[2] * 0.f, 0.f * f[1] - f[0] * 1.f};
    normalize3(sv);
    const float uv[3] = {f[1] * sv[2] - sv[1] * f[2], f[2] * sv[0] - sv[2] * f[0], f[0] * sv[1] - sv[0] * f[1]};
    const float rad = radiansOf(cam->fov);
    const float ph = std::cos(0.5f * rad) / std::sin(0.5f * rad);
    const float pw = ph * (float)height / (float)width;
    Mat4 M = identity();
    if (objectTransform) M = composeTransform(*objectTransform);
    const double zn = cam->nearZ, zf = cam->farZ, W = width, H = height;
    static const int corner[12][2] = {{0, 1}, {2, 3}, {4, 5}, {6, 7}, {0, 2}, {1, 3}, {4, 6}, {5, 7}, {0, 4}, {1, 5}, {2, 6}, {3, 7}};
    for (size_t k = 0; k < order.size(); k++) {
        const uint32_t n = order[k];
        if (n >= numNodes) { delete ov; *out = nullptr; setError("okrt_overlay_build: child index out of range"); return OKRT_ERR_INVALID_ARGUMENT; }
        const okrt_aabb& box = bvh ? bvh[n].boundingBox : oct[n].boundingBox;
        const float strength = 1.f - (float)n / (float)numNodes;
        const float rgb[3] = {bvh ? 0.9f * strength : 0.4f, bvh ? 0.7f * strength : 0.7f, bvh ? 0.5f * strength : 0.4f};
        uint32_t packed = 0xFF000000u;
        for (int c = 0; c < 3; c++) packed |= (uint32_t)quantise(rgb[c]) << (8 * c);
        ov->colours.push_back(packed);
        double view[8][3];
        for (int c = 0; c < 8; c++) {
            /* DebugBBVS.hlsl:27-30: unit-cube vertex * half extent + centre, in float32 */
            float local[3];
            for (int a = 0; a < 3; a++) {
                const float unit = ((c >> a) & 1) ? 1.f : -1.f;
                local[a] = unit * ((box.max[a] - box.min[a]) * 0.5f) + (box.max[a] + box.min[a]) * 0.5f;
            }
            double world[3];
            for (int r = 0; r < 3; r++) world[r] = (double)M.m[0][r] * local[0] + (double)M.m[1][r] * local[1] + (double)M.m[2][r] * local[2] + (double)M.m[3][r];
            const double rel[3] = {world[0] - cam->position[0], world[1] - cam->position[1], world[2] - cam->position[2]};
            view[c][0] = sv[0] * rel[0] + sv[1] * rel[1] + sv[2] * rel[2];
            view[c][1] = uv[0] * rel[0] + uv[1] * rel[1] + uv[2] * rel[2];
            view[c][2] = f[0] * rel[0] + f[1] * rel[1] + f[2] * rel[2];
        }
        for (int e = 0; e < 12; e++) {
            double a[3] = {view[corner[e][0]][0], view[corner[e][0]][1], view[corner[e][0]][2]};
            double c2[3] = {view[corner[e][1]][0], view[corner[e][1]][1], view[corner[e][1]][2]};
            /* depth clip: nearZ <= view z <= farZ */
            double t0 = 0.0, t1 = 1.0;
            const double dz = c2[2] - a[2];
            bool visible = true;
            for (int side = 0; side < 2 && visible; side++) {
                const double plane = side ? zf : zn, sign = side ? -1.0 : 1.0;
                const double da = sign * (a[2] - plane), dc = sign * (c2[2] - plane);
                if (da < 0 && dc < 0) visible = false;
                else if (da < 0) t0 = std::max(t0, (plane - a[2]) / dz);
                else if (dc < 0) t1 = std::min(t1, (plane - a[2]) / dz);
            }
            if (!visible || t0 > t1) continue;
            float seg[4];
            for (int end = 0; end < 2; end++) {
                const double t = end ? t1 : t0;
                const double x = a[0] + (c2[0] - a[0]) * t, y = a[1] + (c2[1] - a[1]) * t, z = a[2] + dz * t;
                seg[2 * end + 0] = (float)((((double)pw * x / z) * 0.5 + 0.5) * W);
                seg[2 * end + 1] = (float)((1.0 - (((double)ph * y / z) * 0.5 + 0.5)) * H);
            }
            ov->segments.insert(ov->segments.end(), seg, seg + 4);
            ov->keys.push_back((uint32_t)k + 1u);
        }
    }
    return OKRT_OK;
}
void okrt_overlay_free(okrt_overlay* ov) { delete ov; }
uint32_t okrt_overlay_num_segments(const okrt_overlay* ov) { return ov ? (uint32_t)ov->keys.size() : 0; }
uint32_t okrt_overlay_num_draws(const okrt_overlay* ov) { return ov ? (uint32_t)ov->colours.size() : 0; }
const float* okrt_overlay_segments(const okrt_overlay* ov) { return ov ? ov->segments.data() : nullptr; }
const uint32_t* okrt_overlay_keys(const okrt_overlay* ov) { return ov ? ov->keys.data() : nullptr; }
const uint32_t* okrt_overlay_colours(const okrt_overlay* ov) { return ov ? ov->colours.data() : nullptr; }

} // extern "C"

/*
 * okrt_layout_device.cu — the traversal layout (okrt_accel.h: child-pair nodes, triangle records) built ON THE DEVICE from the
 * reference-format arrays that okrt_set_* has already uploaded.  Finishes row f2 of SURVEY.md §8: the reference rebuilds its
 * acceleration data on every scene edit (Application.cpp:312-317,405-411,580-585,613-618 -> RayTracer::createOctTree /
 * loadMeshAndBvhData), so the first okrt_render after okrt_set_octree / okrt_set_mesh_instances / okrt_set_bvh_nodes must not
 * cost tens of frames.  The host builder (okrt_accel.cpp: full-sweep SAH on host threads, 150-600 ms) stays as the complete
 * path; this one covers the case every BASELINE config is — all instances identity, each the only user of its mesh — and hands
 * anything else back to the host builder (OKRT_ERR_STATE + reason).
 *
 *   1 thread per reference BVH node : parent links, "parent box contains child boxes" (what lets the traversal skip the
 *                                     reference's interior nodes), leaf flags; bad input -> host builder
 *   1 thread per reference leaf     : walk up to the root -> owning instance record, triangle owners, reference stack depth
 *   cub::DeviceSelect               : the leaves in node order (deterministic)
 *   1 thread per triangle           : p2 | E1 | E2 | owner records (the same float32 subtractions as GPU-Utilities.hlsli:124)
 *   PLOC over the LEAVES            : parallel locally-ordered clustering (after Meister & Bittner 2018, restated): Morton order
 *                                     of the leaf boxes, then rounds of "every cluster finds the neighbour (+-16 in the order) whose
 *                                     union has the smallest area; mutual pairs merge"; node numbers come from prefix sums, so
 *                                     the tree is the same on every run.  Leaf boxes stay the reference's floats (bit copies);
 *                                     interior boxes are exact float unions, emitted one ulp outward.
 *   1 thread per interior node      : the 64-byte child-pair node
 * The tree differs from the host builder's; results do not (every box and triangle test is the reference's arithmetic, the
 * visiting order is free: DESIGN.md §4) — tests/test_gpu_layout.py renders both and compares the images bit for bit.
 */
#include <cuda_runtime.h>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include <thrust/iterator/counting_iterator.h>

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>

#include "okrt.h"
#include "okrt_accel.h"
#include "okrt_internal.h"
#include "okrt_layout_device.h"

namespace okrt {
namespace {

#define OKRT_LD_CUDA(call)                                                                   \
    do {                                                                                     \
        cudaError_t e__ = (call);                                                            \
        if (e__ != cudaSuccess) {                                                            \
            setError(std::string(#call) + ": " + cudaGetErrorString(e__));                   \
            return OKRT_ERR_CUDA;                                                            \
        }                                                                                    \
    } while (0)

constexpr int kPlocRadius = 16;
constexpr uint32_t kNone = 0xFFFFFFFFu;

/* status words on the device */
enum { kStBad = 0, kStRefDepth = 1, kStTopDepth = 2, kStMerged = 3, kStKept = 4, kStWords = 8 };
enum { kBadChildRange = 1, kBadNotNested = 2, kBadLeafRange = 4, kBadOversized = 8, kBadDeep = 16, kBadSharedTriangle = 32, kBadBox = 64 };

__global__ void scatterRootsKernel(const uint2* __restrict__ rootRec, uint32_t n, int32_t* __restrict__ recOfRoot)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) recOfRoot[rootRec[i].x] = (int32_t)rootRec[i].y;
}

__global__ void scanNodesKernel(const okrt_bvh_node* __restrict__ nodes, uint32_t numNodes, uint32_t numTris,
                                uint32_t* __restrict__ parent, uint8_t* __restrict__ isLeaf, uint32_t* __restrict__ status)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= numNodes) return;
    const okrt_bvh_node n = nodes[i];
    uint32_t bad = 0u;
    uint8_t leaf = 0;
    if (n.firstChildIdx != OKRT_INVALID_UINT) {
        if ((uint64_t)n.firstChildIdx + 1u >= numNodes) bad |= kBadChildRange;
        else {
            for (uint32_t c = 0; c < 2u; c++) {
                const okrt_bvh_node ch = nodes[n.firstChildIdx + c];
                parent[n.firstChildIdx + c] = i;
                const bool empty = ch.firstChildIdx == OKRT_INVALID_UINT && (ch.triStart == OKRT_INVALID_UINT || ch.triEnd <= ch.triStart);
                if (empty) continue;
                for (int k = 0; k < 3; k++)
                    if (!(n.boundingBox.min[k] <= ch.boundingBox.min[k] && ch.boundingBox.max[k] <= n.boundingBox.max[k])) bad |= kBadNotNested;
            }
            for (int k = 0; k < 3; k++) if (!(n.boundingBox.min[k] <= n.boundingBox.max[k])) bad |= kBadBox;
        }
    } else if (n.triStart != OKRT_INVALID_UINT && n.triEnd > n.triStart) {
        const uint32_t count = n.triEnd - n.triStart;
        if ((uint64_t)n.triStart + count > numTris || n.triStart >= kMaxTriIndex) bad |= kBadLeafRange;
        else if (count > kMaxLeafTris) bad |= kBadOversized;
        else leaf = 1;
        for (int k = 0; k < 3; k++) if (!(n.boundingBox.min[k] <= n.boundingBox.max[k])) bad |= kBadBox;
    }
    isLeaf[i] = leaf;
    if (bad) atomicOr(status + kStBad, bad);
}

/* every reference leaf finds its root and with it the instance record that owns its triangles; leaves of meshes no
 * instance uses are dropped (the reference never reaches them either) */
__global__ void leafOwnerKernel(const okrt_bvh_node* __restrict__ nodes, uint32_t numNodes, const uint32_t* __restrict__ parent,
                                const int32_t* __restrict__ recOfRoot, uint8_t* __restrict__ isLeaf, int32_t* __restrict__ triOwner,
                                uint32_t* __restrict__ status)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= numNodes || !isLeaf[i]) return;
    uint32_t j = i, steps = 0u;
    while (parent[j] != kNone && steps < 4096u) { j = parent[j]; steps++; }
    if (steps >= 4096u) { atomicOr(status + kStBad, (uint32_t)kBadDeep); isLeaf[i] = 0; return; }
    const int32_t rec = recOfRoot[j];
    if (rec < 0) { isLeaf[i] = 0; return; }
    atomicMax(status + kStRefDepth, steps);
    const okrt_bvh_node n = nodes[i];
    for (uint32_t t = n.triStart; t < n.triEnd; t++) {
        const int32_t old = atomicCAS(triOwner + t, -1, rec);
        if (old != -1 && old != rec) atomicOr(status + kStBad, (uint32_t)kBadSharedTriangle);
    }
}

__global__ void triangleRecordsKernel(const float* __restrict__ triPos, const int32_t* __restrict__ triOwner, uint32_t numTris, float4* __restrict__ out)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= numTris) return;
    const float* p = triPos + (size_t)t * 9;
    /* E1 = p0 - p2, E2 = p1 - p2: individually rounded float32 subtractions (this file is compiled with -fmad=false) */
    const float e1x = p[0] - p[6], e1y = p[1] - p[7], e1z = p[2] - p[8], e2x = p[3] - p[6], e2y = p[4] - p[7], e2z = p[5] - p[8];
    out[(size_t)t * kTriQuads + 0] = make_float4(p[6], p[7], p[8], e1x);
    out[(size_t)t * kTriQuads + 1] = make_float4(e1y, e1z, e2x, e2y);
    out[(size_t)t * kTriQuads + 2] = make_float4(e2z, __int_as_float(triOwner[t]), 0.f, 0.f);
}

/* order-preserving map float -> uint for atomicMin / atomicMax */
__device__ __forceinline__ uint32_t orderedBits(float f) { const uint32_t u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__device__ __forceinline__ float orderedFloat(uint32_t u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u); }

/* primitives of the scene-wide tree: one per selected leaf (its box: bit copy; its reference word), scene bounds */
__global__ void leafPrimsKernel(const okrt_bvh_node* __restrict__ nodes, const uint32_t* __restrict__ leafNode, const uint32_t* __restrict__ numLeavesPtr,
                                float* __restrict__ primBox, uint32_t* __restrict__ primRef, uint32_t* __restrict__ bounds /* 6 ordered words */)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= *numLeavesPtr) return;
    const okrt_bvh_node n = nodes[leafNode[p]];
    for (int k = 0; k < 3; k++) {
        primBox[(size_t)p * 6 + k] = n.boundingBox.min[k]; primBox[(size_t)p * 6 + 3 + k] = n.boundingBox.max[k];
        atomicMin(bounds + k, orderedBits(n.boundingBox.min[k])); atomicMax(bounds + 3 + k, orderedBits(n.boundingBox.max[k]));
    }
    primRef[p] = kRefTriLeaf | ((n.triEnd - n.triStart - 1u) << 27) | n.triStart;
}

__device__ __forceinline__ uint32_t expandBits10b(uint32_t v)
{
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
__global__ void primMortonKernel(const float* __restrict__ primBox, uint32_t n, const uint32_t* __restrict__ bounds, uint32_t* __restrict__ codes, uint32_t* __restrict__ ids)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    uint32_t q[3];
    for (int k = 0; k < 3; k++) {
        const float lo = orderedFloat(bounds[k]), hi = orderedFloat(bounds[3 + k]);
        const float c = (primBox[(size_t)p * 6 + k] + primBox[(size_t)p * 6 + 3 + k]) * 0.5f;
        const float ext = hi - lo;
        const float u = ext > 0.f ? (c - lo) / ext : 0.f;
        q[k] = (uint32_t)fminf(fmaxf(u * 1024.0f, 0.0f), 1023.0f);
    }
    codes[p] = (expandBits10b(q[0]) << 2) | (expandBits10b(q[1]) << 1) | expandBits10b(q[2]);
    ids[p] = p;
}

/* tree nodes 0 .. n-1: the leaves in Morton order; n .. 2n-2: interior nodes in creation order */
__global__ void treeLeavesKernel(const float* __restrict__ primBox, const uint32_t* __restrict__ sortedPrim, uint32_t n,
                                 float* __restrict__ box, uint32_t* __restrict__ cluster)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t p = sortedPrim[i];
    for (int k = 0; k < 6; k++) box[(size_t)i * 6 + k] = primBox[(size_t)p * 6 + k];
    cluster[i] = i;
}

__device__ __forceinline__ float unionArea(const float* a, const float* b)
{
    const float dx = fmaxf(a[3], b[3]) - fminf(a[0], b[0]), dy = fmaxf(a[4], b[4]) - fminf(a[1], b[1]), dz = fmaxf(a[5], b[5]) - fminf(a[2], b[2]);
    return dx * dy + dy * dz + dz * dx;
}

/* The clustering rounds keep their state on the device — state[0] = clusters left, state[1] = next free tree node, double
 * buffered by round parity — so that the host can queue many rounds without reading anything back (a round with one cluster
 * left changes nothing). */
__global__ void plocNearestKernel(const float* __restrict__ box, const uint32_t* __restrict__ cluster, const uint32_t* __restrict__ state, uint32_t* __restrict__ nearest)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t count = state[0];
    if (i >= count) return;
    float mine[6];
    const float* b = box + (size_t)cluster[i] * 6;
    for (int k = 0; k < 6; k++) mine[k] = b[k];
    const int lo = max(0, (int)i - kPlocRadius), hi = min((int)count - 1, (int)i + kPlocRadius);
    float best = FLT_MAX; uint32_t bestJ = kNone;
    for (int j = lo; j <= hi; j++) {
        if (j == (int)i) continue;
        const float a = unionArea(mine, box + (size_t)cluster[j] * 6);
        if (a < best || bestJ == kNone) { best = a; bestJ = (uint32_t)j; }      /* ties: the smaller position */
    }
    nearest[i] = bestJ;
}

/* flags[i] = merge | keep << 32 for the clusters left, 0 up to `bound` (what the scan covers) */
__global__ void plocFlagsKernel(const uint32_t* __restrict__ nearest, const uint32_t* __restrict__ state, uint32_t bound, unsigned long long* __restrict__ flags)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= bound) return;
    if (i >= state[0]) { flags[i] = 0ull; return; }
    const uint32_t j = nearest[i];
    const bool mutual = j != kNone && nearest[j] == i;
    const unsigned long long merge = (mutual && i < j) ? 1ull : 0ull, keep = (mutual && i > j) ? 0ull : 1ull;
    flags[i] = merge | (keep << 32);
}

__global__ void plocApplyKernel(const uint32_t* __restrict__ nearest, const unsigned long long* __restrict__ flags, const unsigned long long* __restrict__ idx,
                                const uint32_t* __restrict__ stateIn, uint32_t* __restrict__ stateOut,
                                const uint32_t* __restrict__ cluster, uint32_t* __restrict__ clusterOut, float* __restrict__ box,
                                uint32_t* __restrict__ child /* 2 per node */, uint32_t* __restrict__ treeParent)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t count = stateIn[0], nextNode = stateIn[1];
    if (i >= count) return;
    const uint32_t mergeFlag = (uint32_t)(flags[i] & 0xFFFFFFFFull), keepFlag = (uint32_t)(flags[i] >> 32);
    const uint32_t mergeIdx = (uint32_t)(idx[i] & 0xFFFFFFFFull), keepIdx = (uint32_t)(idx[i] >> 32);
    if (mergeFlag) {
        const uint32_t id = nextNode + mergeIdx, a = cluster[i], b = cluster[nearest[i]];
        const float *ba = box + (size_t)a * 6, *bb = box + (size_t)b * 6;
        float* bo = box + (size_t)id * 6;
        for (int k = 0; k < 3; k++) { bo[k] = fminf(ba[k], bb[k]); bo[3 + k] = fmaxf(ba[3 + k], bb[3 + k]); }
        child[2 * (size_t)id] = a; child[2 * (size_t)id + 1] = b;
        treeParent[a] = id; treeParent[b] = id;
        clusterOut[keepIdx] = id;
    } else if (keepFlag) clusterOut[keepIdx] = cluster[i];
    if (i == count - 1u) { stateOut[0] = keepIdx + keepFlag; stateOut[1] = nextNode + mergeIdx + mergeFlag; }
}

__global__ void emitPairsKernel(const float* __restrict__ box, const uint32_t* __restrict__ child, const uint32_t* __restrict__ sortedPrim,
                                const uint32_t* __restrict__ primRef, uint32_t n, float4* __restrict__ pairs)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;          /* interior node n + k -> pair node k */
    if (k + 1u >= n) return;
    float4 q[4];
    uint32_t refs[2];
    for (int c = 0; c < 2; c++) {
        const uint32_t id = child[2 * (size_t)(n + k) + c];
        const float* b = box + (size_t)id * 6;
        float lo[3] = {b[0], b[1], b[2]}, hi[3] = {b[3], b[4], b[5]};
        if (id < n) refs[c] = primRef[sortedPrim[id]];                 /* leaf: the reference's own box, bit for bit */
        else {
            refs[c] = id - n;                                          /* interior: exact union, one ulp outward */
            for (int a = 0; a < 3; a++) { lo[a] = nextafterf(lo[a], -INFINITY); hi[a] = nextafterf(hi[a], INFINITY); }
        }
        q[2 * c] = make_float4(lo[0], lo[1], lo[2], 0.f);
        q[2 * c + 1] = make_float4(hi[0], hi[1], hi[2], 0.f);
    }
    q[0].w = __uint_as_float(refs[0]); q[1].w = __uint_as_float(refs[1]);
    for (int c = 0; c < 4; c++) pairs[(size_t)k * 4 + c] = q[c];
}

__global__ void treeDepthKernel(const uint32_t* __restrict__ treeParent, uint32_t n, uint32_t* __restrict__ status)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t j = i, steps = 0u;
    while (treeParent[j] != kNone && steps < 4096u) { j = treeParent[j]; steps++; }
    atomicMax(status + kStTopDepth, steps);
}

struct Arena {
    uint8_t* base = nullptr; size_t size = 0, used = 0;
    template <typename T> T* take(size_t count)
    {
        used = (used + 255) & ~(size_t)255;
        T* p = reinterpret_cast<T*>(base + used);
        used += count * sizeof(T);
        return used <= size ? p : nullptr;
    }
};

} // namespace

int buildFastLayoutDevice(const okrt_triangle* dTris, uint32_t numTris, const okrt_bvh_node* dNodes, const okrt_bvh_node* hNodes, uint32_t numNodes,
                          const okrt_mesh_instance* hInstances, uint32_t numInstances, const okrt_octree_node* hOct, uint32_t numOctNodes,
                          cudaStream_t stream, DeviceLayout& out, std::string& why)
{
    out = DeviceLayout();
    why.clear();
    /* ---- host part: O(instances + oct-tree nodes) */
    std::vector<uint8_t> octNested;
    std::vector<uint32_t> instOct;
    std::string error;
    if (!octTreeChecks(hOct, numOctNodes, numInstances, out.octParent, octNested, instOct, out.exactOctDepth, error)) { why = error; return OKRT_ERR_STATE; }
    std::unordered_map<uint32_t, uint32_t> rootUsers;
    for (uint32_t i = 0; i < numInstances; i++)
        if (instOct[i] != kNone) { if (hInstances[i].bvhNodeStartIdx >= numNodes) { why = "instance bvhNodeStartIdx out of range"; return OKRT_ERR_STATE; } rootUsers[hInstances[i].bvhNodeStartIdx]++; }
    std::vector<uint2> rootRec;
    static const float I16[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    for (uint32_t i = 0; i < numInstances; i++) {
        if (instOct[i] == kNone) continue;
        const okrt_mesh_instance& inst = hInstances[i];
        const okrt_aabb& rb = hNodes[inst.bvhNodeStartIdx].boundingBox;
        const float* T = inst.transformMatrix;
        const float* Ti = inst.inverseTransformMatrix;
        bool identity = true;
        for (int k = 0; k < 16; k++) identity = identity && T[k] == I16[k] && Ti[k] == I16[k];
        if (!identity) { why = "an instance carries a transform (walked with its own ray: host builder)"; return OKRT_ERR_STATE; }
        if (rootUsers[inst.bvhNodeStartIdx] != 1u) { why = "a mesh is used by several instances (host builder)"; return OKRT_ERR_STATE; }
        bool finite = true;
        for (int k = 0; k < 3; k++) finite = finite && std::isfinite(rb.min[k]) && std::isfinite(rb.max[k]) && rb.min[k] <= rb.max[k];
        if (!finite) { why = "an instance has an empty or non-finite root box (host builder)"; return OKRT_ERR_STATE; }
        for (uint32_t on = instOct[i]; on != kNone; on = out.octParent[on])
            for (int k = 0; k < 3; k++) {
                const double tol = 1e-5 * std::max(1.0, std::max(std::fabs((double)rb.min[k]), std::fabs((double)rb.max[k])));
                if (!((double)hOct[on].boundingBox.min[k] <= rb.min[k] + tol && rb.max[k] - tol <= (double)hOct[on].boundingBox.max[k])) {
                    why = "an oct-tree box does not contain an instance it lists (host builder)"; return OKRT_ERR_STATE;
                }
            }
        const uint32_t rec = (uint32_t)(out.instances.size() / kInstQuads);
        out.instances.resize(out.instances.size() + kInstQuads);
        Float4* r = &out.instances[(size_t)rec * kInstQuads];
        auto bits = [](uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; };
        for (int row = 0; row < 3; row++) r[row] = Float4{Ti[row * 4 + 0], Ti[row * 4 + 1], Ti[row * 4 + 2], Ti[row * 4 + 3]};
        for (int row = 0; row < 3; row++) r[3 + row] = Float4{T[row * 4 + 0], T[row * 4 + 1], T[row * 4 + 2], rb.min[row]};
        r[6] = Float4{rb.max[0], rb.max[1], rb.max[2], bits(1u)};
        r[7] = Float4{bits(kRefNone), bits(instOct[i]), bits(i), bits(octNested[instOct[i]] ? 1u : 0u)};
        const okrt_aabb& ob = hOct[instOct[i]].boundingBox;
        r[8] = Float4{ob.min[0], ob.min[1], ob.min[2], bits(kRefNone)};
        r[9] = Float4{ob.max[0], ob.max[1], ob.max[2], 0.f};
        rootRec.push_back(make_uint2(inst.bvhNodeStartIdx, rec));
    }
    out.numMerged = (uint32_t)rootRec.size();
    out.allRigid = true;
    if (rootRec.empty() || numNodes == 0 || numTris == 0) { why = "empty scene (host builder)"; return OKRT_ERR_STATE; }

    /* ---- device part */
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    OKRT_LD_CUDA(cudaEventCreate(&ev0)); OKRT_LD_CUDA(cudaEventCreate(&ev1));
    struct Cleanup { cudaEvent_t a, b; void* scratch; ~Cleanup() { if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); if (scratch) cudaFree(scratch); } } cleanup{ev0, ev1, nullptr};
    const size_t maxLeaves = numNodes, treeNodes = 2 * maxLeaves;
    size_t selectBytes = 0, sortBytes = 0, scanBytes = 0;
    cub::DeviceSelect::Flagged(nullptr, selectBytes, thrust::counting_iterator<uint32_t>(0), (const uint8_t*)nullptr, (uint32_t*)nullptr, (uint32_t*)nullptr, (int)numNodes, stream);
    cub::DeviceRadixSort::SortPairs(nullptr, sortBytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)maxLeaves, 0, 30, stream);
    cub::DeviceScan::ExclusiveSum(nullptr, scanBytes, (const unsigned long long*)nullptr, (unsigned long long*)nullptr, (int)maxLeaves, stream);
    const size_t cubBytes = std::max(selectBytes, std::max(sortBytes, scanBytes));
    Arena ar;
    ar.size = cubBytes + 4096 + (size_t)numNodes * (4 + 1 + 4) + (size_t)numTris * 4 + rootRec.size() * 8 +
              maxLeaves * (4 + 24 + 4 + 4 + 4 + 4 + 4) + treeNodes * (24 + 8 + 4) + maxLeaves * 4 * 7 + 64 * 256;
    OKRT_LD_CUDA(cudaMalloc(&cleanup.scratch, ar.size));
    OKRT_LD_CUDA(cudaMalloc(&out.tris, (size_t)numTris * kTriQuads * sizeof(float4)));
    OKRT_LD_CUDA(cudaEventRecord(ev0, stream));                        /* deviceMs: the build, not the two allocations above */
    ar.base = (uint8_t*)cleanup.scratch;
    uint32_t* status = ar.take<uint32_t>(kStWords + 8);                  /* + 6 ordered bound words, + number of leaves */
    uint32_t* bounds = status + kStWords;
    uint32_t* numLeavesDev = status + kStWords + 6;
    void* cubTemp = ar.take<uint8_t>(cubBytes);
    uint32_t* parent = ar.take<uint32_t>(numNodes);
    uint8_t* isLeaf = ar.take<uint8_t>(numNodes);
    int32_t* recOfRoot = ar.take<int32_t>(numNodes);
    int32_t* triOwner = ar.take<int32_t>(numTris);
    uint2* dRootRec = ar.take<uint2>(rootRec.size());
    uint32_t* leafNode = ar.take<uint32_t>(maxLeaves);
    float* primBox = ar.take<float>(maxLeaves * 6);
    uint32_t* primRef = ar.take<uint32_t>(maxLeaves);
    uint32_t *codes = ar.take<uint32_t>(maxLeaves), *codesSorted = ar.take<uint32_t>(maxLeaves), *ids = ar.take<uint32_t>(maxLeaves), *sortedPrim = ar.take<uint32_t>(maxLeaves);
    float* box = ar.take<float>(treeNodes * 6);
    uint32_t* child = ar.take<uint32_t>(treeNodes * 2);
    uint32_t* treeParent = ar.take<uint32_t>(treeNodes);
    uint32_t *clusterA = ar.take<uint32_t>(maxLeaves), *clusterB = ar.take<uint32_t>(maxLeaves), *nearest = ar.take<uint32_t>(maxLeaves);
    unsigned long long *flags64 = ar.take<unsigned long long>(maxLeaves), *idx64 = ar.take<unsigned long long>(maxLeaves);
    uint32_t* plocState = ar.take<uint32_t>(4);
    if (!plocState) { cudaFree(out.tris); out.tris = nullptr; setError("device layout: scratch arena too small"); return OKRT_ERR_CAPACITY; }

    const unsigned B = 256;
    auto grid = [&](size_t n) { return (unsigned)((n + B - 1) / B); };
    OKRT_LD_CUDA(cudaMemsetAsync(status, 0, (kStWords + 8) * 4, stream));
    {
        uint32_t init[6];
        for (int k = 0; k < 3; k++) { init[k] = 0xFFFFFFFFu; init[3 + k] = 0u; }
        OKRT_LD_CUDA(cudaMemcpyAsync(bounds, init, sizeof(init), cudaMemcpyHostToDevice, stream));
    }
    OKRT_LD_CUDA(cudaMemsetAsync(parent, 0xFF, (size_t)numNodes * 4, stream));
    OKRT_LD_CUDA(cudaMemsetAsync(recOfRoot, 0xFF, (size_t)numNodes * 4, stream));
    OKRT_LD_CUDA(cudaMemsetAsync(triOwner, 0xFF, (size_t)numTris * 4, stream));
    OKRT_LD_CUDA(cudaMemcpyAsync(dRootRec, rootRec.data(), rootRec.size() * sizeof(uint2), cudaMemcpyHostToDevice, stream));
    scatterRootsKernel<<<grid(rootRec.size()), B, 0, stream>>>(dRootRec, (uint32_t)rootRec.size(), recOfRoot);
    scanNodesKernel<<<grid(numNodes), B, 0, stream>>>(dNodes, numNodes, numTris, parent, isLeaf, status);
    leafOwnerKernel<<<grid(numNodes), B, 0, stream>>>(dNodes, numNodes, parent, recOfRoot, isLeaf, triOwner, status);
    {
        size_t bytes = cubBytes;
        OKRT_LD_CUDA(cub::DeviceSelect::Flagged(cubTemp, bytes, thrust::counting_iterator<uint32_t>(0), isLeaf, leafNode, numLeavesDev, (int)numNodes, stream));
    }
    /* the triangle records and the leaf primitives do not need the host to know the leaf count */
    triangleRecordsKernel<<<grid(numTris), B, 0, stream>>>((const float*)dTris, triOwner, numTris, out.tris);
    leafPrimsKernel<<<grid(maxLeaves), B, 0, stream>>>(dNodes, leafNode, numLeavesDev, primBox, primRef, bounds);
    uint32_t hStatus[kStWords + 8];
    OKRT_LD_CUDA(cudaMemcpyAsync(hStatus, status, sizeof(hStatus), cudaMemcpyDeviceToHost, stream));
    OKRT_LD_CUDA(cudaStreamSynchronize(stream));
    auto giveUp = [&](const std::string& reason) { if (out.tris) cudaFree(out.tris); if (out.nodes) cudaFree(out.nodes); out.tris = nullptr; out.nodes = nullptr; why = reason; return OKRT_ERR_STATE; };
    if (hStatus[kStBad]) {
        char msg[128];
        std::snprintf(msg, sizeof(msg), "reference BVH needs the host builder's checks (flags 0x%x)", hStatus[kStBad]);
        return giveUp(msg);
    }
    const uint32_t n = hStatus[kStWords + 6];
    if (n == 0) return giveUp("no triangles reachable (host builder)");
    out.exactBvhDepth = hStatus[kStRefDepth] + 1u;
    out.numFlattenedLeaves = n;

    /* ---- PLOC over the n leaves */
    primMortonKernel<<<grid(n), B, 0, stream>>>(primBox, n, bounds, codes, ids);
    {
        size_t bytes = cubBytes;
        OKRT_LD_CUDA(cub::DeviceRadixSort::SortPairs(cubTemp, bytes, codes, codesSorted, ids, sortedPrim, (int)n, 0, 30, stream));
    }
    OKRT_LD_CUDA(cudaMemsetAsync(treeParent, 0xFF, (size_t)2 * n * 4, stream));
    treeLeavesKernel<<<grid(n), B, 0, stream>>>(primBox, sortedPrim, n, box, clusterA);
    uint32_t count = n, nextNode = n, rounds = 0, parity = 0;
    uint32_t *cur = clusterA, *nxt = clusterB;
    {
        const uint32_t init[2] = {n, n};
        OKRT_LD_CUDA(cudaMemcpyAsync(plocState, init, sizeof(init), cudaMemcpyHostToDevice, stream));
    }
    /* rounds are queued kPlocChunk at a time; the host reads the state once per chunk (a read per round made the build wait
     * for the host 50 times, which a busy host stretches to tens of milliseconds) */
    constexpr uint32_t kPlocChunk = 16;
    while (count > 1u) {
        const uint32_t bound = count;
        for (uint32_t k = 0; k < kPlocChunk; k++) {
            const uint32_t* in = plocState + 2u * parity;
            uint32_t* outState = plocState + 2u * (parity ^ 1u);
            plocNearestKernel<<<grid(bound), B, 0, stream>>>(box, cur, in, nearest);
            plocFlagsKernel<<<grid(bound), B, 0, stream>>>(nearest, in, bound, flags64);
            size_t bytes = cubBytes;
            OKRT_LD_CUDA(cub::DeviceScan::ExclusiveSum(cubTemp, bytes, flags64, idx64, (int)bound, stream));
            plocApplyKernel<<<grid(bound), B, 0, stream>>>(nearest, flags64, idx64, in, outState, cur, nxt, box, child, treeParent);
            std::swap(cur, nxt);
            parity ^= 1u;
            rounds++;
        }
        uint32_t two[2];
        OKRT_LD_CUDA(cudaMemcpyAsync(two, plocState + 2u * parity, 8, cudaMemcpyDeviceToHost, stream));
        OKRT_LD_CUDA(cudaStreamSynchronize(stream));
        if (two[0] == 0u || two[0] >= count) return giveUp("clustering made no progress (host builder)");
        count = two[0]; nextNode = two[1];
        if (rounds > 4096u) return giveUp("clustering did not finish (host builder)");
    }
    if (n > 1u && nextNode != 2u * n - 1u) return giveUp("clustering lost nodes (host builder)");

    /* ---- pair nodes */
    const uint32_t numPairs = n > 1u ? n - 1u : 1u;
    OKRT_LD_CUDA(cudaMalloc(&out.nodes, (size_t)numPairs * 4 * sizeof(float4)));
    if (n > 1u) {
        emitPairsKernel<<<grid(n - 1u), B, 0, stream>>>(box, child, sortedPrim, primRef, n, out.nodes);
        treeDepthKernel<<<grid(n), B, 0, stream>>>(treeParent, n, status);
        /* the last interior node created is the root */
        out.topRef = 2u * n - 2u - n;
    } else {
        /* a single leaf: one pair node whose second child is empty (NaN box), as the host builder does */
        float hBox[6]; uint32_t hRef;
        OKRT_LD_CUDA(cudaMemcpyAsync(hBox, primBox, sizeof(hBox), cudaMemcpyDeviceToHost, stream));
        OKRT_LD_CUDA(cudaMemcpyAsync(&hRef, primRef, 4, cudaMemcpyDeviceToHost, stream));
        OKRT_LD_CUDA(cudaStreamSynchronize(stream));
        const float nan = std::nanf("");
        float refBits, noneBits; std::memcpy(&refBits, &hRef, 4); const uint32_t none = kRefNone; std::memcpy(&noneBits, &none, 4);
        const float4 q[4] = {make_float4(hBox[0], hBox[1], hBox[2], refBits), make_float4(hBox[3], hBox[4], hBox[5], noneBits),
                             make_float4(nan, nan, nan, 0.f), make_float4(nan, nan, nan, 0.f)};
        OKRT_LD_CUDA(cudaMemcpyAsync(out.nodes, q, sizeof(q), cudaMemcpyHostToDevice, stream));
        out.topRef = 0u;
    }
    OKRT_LD_CUDA(cudaEventRecord(ev1, stream));
    OKRT_LD_CUDA(cudaMemcpyAsync(hStatus, status, sizeof(hStatus), cudaMemcpyDeviceToHost, stream));
    OKRT_LD_CUDA(cudaStreamSynchronize(stream));
    out.numPairNodes = numPairs;
    out.maxStackDepth = (n > 1u ? hStatus[kStTopDepth] : 1u) + 2u;
    out.plocRounds = rounds;
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, ev0, ev1) == cudaSuccess) out.deviceMs = ms;
    if (out.maxStackDepth > kFastStackCapacity) return giveUp("clustered tree deeper than the traversal stack (host builder)");
    OKRT_LD_CUDA(cudaGetLastError());
    return OKRT_OK;
}

} // namespace okrt

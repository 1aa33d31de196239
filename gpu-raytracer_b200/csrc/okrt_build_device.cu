/*
 * okrt_build_device.cu — row f2 of SURVEY.md §8: a BVH builder that runs on the GPU and replaces
 * BvhBuilder::buildTree (source/Graphics/BvhBuilder.cpp:38-207, minutes for the 871k-triangle dragon on one CPU
 * thread) where a *different but valid* tree is acceptable: the UI's "Rebuild BVH" (Application.cpp:405-411), scene
 * edits, first display.  Output is the reference's own format — BvhNode 36 B with adjacent children
 * (firstChildIdx, firstChildIdx + 1), leaf-ordered Triangle / TriangleInfo arrays, MeshDesc — so everything
 * downstream (okrt_set_bvh_nodes, the scene-wide tree of okrt_accel.cpp, the reference-order traversal, the CPU
 * oracle) consumes it unchanged.  Parity statement: the oracle and the kernels given THIS node array agree bit for
 * bit; against the reference's SAH tree the image is the same up to tie cases (tests/test_gpu_builder.py: PSNR).
 *
 * Algorithm (LBVH): 30-bit Morton code of each triangle's centroid in the mesh box; stable radix sort of
 * (code, triangle) pairs (cub::DeviceRadixSort — NVIDIA's library, plumbing, not the hot path); Karras' parallel
 * radix-tree construction (one thread per internal node, ties broken by sorted position); every subtree of at most
 * maxLeafTriangles triangles becomes one leaf; boxes are fitted bottom-up (second arrival at a node merges), as exact
 * float min/max, so a parent box contains its children's boxes exactly (what okrt_accel.cpp:flattenable() requires).
 * Internal node i puts its two children at slots 2i+1 and 2i+2 of the output array (root at 0): children are adjacent
 * as the format demands, no compaction pass; slots of pruned subtrees stay as unreachable empty leaves.
 */
#include <cuda_runtime.h>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <algorithm>
#include <cfloat>
#include <cstring>
#include <string>
#include <vector>

#include "okrt.h"
#include "okrt_internal.h"

namespace {

#define OKRT_BUILD_CUDA(call)                                                                \
    do {                                                                                     \
        cudaError_t e__ = (call);                                                            \
        if (e__ != cudaSuccess) {                                                            \
            okrt::setError(std::string(#call) + ": " + cudaGetErrorString(e__));             \
            return OKRT_ERR_CUDA;                                                            \
        }                                                                                    \
    } while (0)

__device__ __forceinline__ uint32_t expandBits10(uint32_t v)
{
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

/* triangle boxes (exact min/max of the corners) and Morton codes of the box centres */
__global__ void mortonKernel(const float* __restrict__ triPos, uint32_t n, float3 lo, float3 invExtent,
                             float* __restrict__ primBox, uint32_t* __restrict__ codes, uint32_t* __restrict__ ids)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* p = triPos + (size_t)i * 9;
    float bl[3], bh[3];
    for (int k = 0; k < 3; k++) { bl[k] = fminf(fminf(p[k], p[3 + k]), p[6 + k]); bh[k] = fmaxf(fmaxf(p[k], p[3 + k]), p[6 + k]); }
    for (int k = 0; k < 3; k++) { primBox[(size_t)i * 6 + k] = bl[k]; primBox[(size_t)i * 6 + 3 + k] = bh[k]; }
    const float cx = ((bl[0] + bh[0]) * 0.5f - lo.x) * invExtent.x, cy = ((bl[1] + bh[1]) * 0.5f - lo.y) * invExtent.y, cz = ((bl[2] + bh[2]) * 0.5f - lo.z) * invExtent.z;
    const uint32_t x = (uint32_t)fminf(fmaxf(cx * 1024.0f, 0.0f), 1023.0f), y = (uint32_t)fminf(fmaxf(cy * 1024.0f, 0.0f), 1023.0f), z = (uint32_t)fminf(fmaxf(cz * 1024.0f, 0.0f), 1023.0f);
    codes[i] = (expandBits10(x) << 2) | (expandBits10(y) << 1) | expandBits10(z);
    ids[i] = i;
}

constexpr uint32_t kLeafBit = 0x80000000u;      /* child reference: leaf k of the sorted order, else internal node */
constexpr uint32_t kNoParent = 0xFFFFFFFFu;

/* common-prefix length of sorted keys i and j; equal codes fall back to the positions, so all keys are distinct */
__device__ __forceinline__ int prefixLen(const uint32_t* __restrict__ codes, int n, int i, int j)
{
    if (j < 0 || j >= n) return -1;
    const uint32_t a = codes[i], b = codes[j];
    return a != b ? __clz(a ^ b) : 32 + __clz((uint32_t)i ^ (uint32_t)j);
}

/* one thread per internal node of the binary radix tree over n sorted keys (n - 1 internal nodes, node 0 = root) */
__global__ void radixTreeKernel(const uint32_t* __restrict__ codes, int n, uint32_t* __restrict__ child /* 2 per node */,
                                uint32_t* __restrict__ rangeFirst, uint32_t* __restrict__ rangeLast,
                                uint32_t* __restrict__ nodeParent, uint32_t* __restrict__ leafParent)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = prefixLen(codes, n, i, i + 1) > prefixLen(codes, n, i, i - 1) ? 1 : -1;
    const int floorLen = prefixLen(codes, n, i, i - d);
    int reach = 2;
    while (prefixLen(codes, n, i, i + reach * d) > floorLen) reach <<= 1;
    int len = 0;
    for (int t = reach >> 1; t >= 1; t >>= 1)
        if (prefixLen(codes, n, i, i + (len + t) * d) > floorLen) len += t;
    const int j = i + len * d;
    const int nodeLen = prefixLen(codes, n, i, j);
    int s = 0, t = len;
    do {
        t = (t + 1) >> 1;
        if (prefixLen(codes, n, i, i + (s + t) * d) > nodeLen) s += t;
    } while (t > 1);
    const int split = i + s * d + min(d, 0);
    const int first = min(i, j), last = max(i, j);
    const uint32_t left = first == split ? (kLeafBit | (uint32_t)split) : (uint32_t)split;
    const uint32_t right = last == split + 1 ? (kLeafBit | (uint32_t)(split + 1)) : (uint32_t)(split + 1);
    child[2 * i] = left; child[2 * i + 1] = right;
    rangeFirst[i] = (uint32_t)first; rangeLast[i] = (uint32_t)last;
    if (left & kLeafBit) leafParent[split] = (uint32_t)i; else nodeParent[split] = (uint32_t)i;
    if (right & kLeafBit) leafParent[split + 1] = (uint32_t)i; else nodeParent[split + 1] = (uint32_t)i;
    if (i == 0) nodeParent[0] = kNoParent;
}

__device__ __forceinline__ void loadBox(const float* __restrict__ boxes, uint32_t idx, float* b)
{
    for (int k = 0; k < 6; k++) b[k] = __ldcg(boxes + (size_t)idx * 6 + k);
}

/* bottom-up fit: the second thread to reach a node owns it and merges its two children (exact min / max) */
__global__ void fitKernel(int n, const uint32_t* __restrict__ sortedIds, const float* __restrict__ primBox,
                          const uint32_t* __restrict__ child, const uint32_t* __restrict__ nodeParent,
                          const uint32_t* __restrict__ leafParent, float* nodeBox, uint32_t* arrivals)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    uint32_t node = leafParent[k];
    while (node != kNoParent) {
        __threadfence();
        if (atomicAdd(arrivals + node, 1u) == 0u) return;
        __threadfence();
        float a[6], b[6];
        const uint32_t cl = child[2 * node], cr = child[2 * node + 1];
        if (cl & kLeafBit) loadBox(primBox, sortedIds[cl & ~kLeafBit], a); else loadBox(nodeBox, cl, a);
        if (cr & kLeafBit) loadBox(primBox, sortedIds[cr & ~kLeafBit], b); else loadBox(nodeBox, cr, b);
        for (int c = 0; c < 3; c++) {
            __stcg(nodeBox + (size_t)node * 6 + c, fminf(a[c], b[c]));
            __stcg(nodeBox + (size_t)node * 6 + 3 + c, fmaxf(a[3 + c], b[3 + c]));
        }
        node = nodeParent[node];
    }
}

/* kept[i] = 1 when internal node i stays an interior node of the output (its range exceeds the leaf budget) */
__global__ void keptKernel(int numInternal, const uint32_t* __restrict__ rangeFirst, const uint32_t* __restrict__ rangeLast,
                           uint32_t maxLeaf, uint32_t* __restrict__ kept)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < numInternal) kept[i] = (rangeLast[i] - rangeFirst[i] + 1u > maxLeaf) ? 1u : 0u;
}

struct OutNode { float box[6]; uint32_t triStart, triEnd, firstChild; };     /* == okrt_bvh_node, 36 B */

__device__ __forceinline__ void writeNode(OutNode* out, uint32_t slot, const float* box, uint32_t t0, uint32_t t1, uint32_t fc)
{
    OutNode nd;
    for (int k = 0; k < 6; k++) nd.box[k] = box[k];
    nd.triStart = t0; nd.triEnd = t1; nd.firstChild = fc;
    out[slot] = nd;
}

/* every kept node of compact rank c writes its two children to slots 2c + 1 and 2c + 2; node 0 also writes the root */
__global__ void emitKernel(int numInternal, const uint32_t* __restrict__ kept, const uint32_t* __restrict__ rank,
                           const uint32_t* __restrict__ child, const uint32_t* __restrict__ rangeFirst,
                           const uint32_t* __restrict__ rangeLast, const uint32_t* __restrict__ sortedIds,
                           const float* __restrict__ primBox, const float* __restrict__ nodeBox, okrt_aabb rootBounds,
                           OutNode* __restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= numInternal || !kept[i]) return;
    const uint32_t base = 2u * rank[i] + 1u;
    if (i == 0) {
        const float rb[6] = {rootBounds.min[0], rootBounds.min[1], rootBounds.min[2], rootBounds.max[0], rootBounds.max[1], rootBounds.max[2]};
        writeNode(out, 0, rb, OKRT_INVALID_UINT, OKRT_INVALID_UINT, 1u);
    }
    for (int side = 0; side < 2; side++) {
        const uint32_t c = child[2 * i + side];
        float box[6];
        if (c & kLeafBit) {
            const uint32_t k = c & ~kLeafBit;
            for (int q = 0; q < 6; q++) box[q] = primBox[(size_t)sortedIds[k] * 6 + q];
            writeNode(out, base + side, box, k, k + 1u, OKRT_INVALID_UINT);
        } else {
            for (int q = 0; q < 6; q++) box[q] = nodeBox[(size_t)c * 6 + q];
            if (kept[c]) writeNode(out, base + side, box, OKRT_INVALID_UINT, OKRT_INVALID_UINT, 2u * rank[c] + 1u);
            else writeNode(out, base + side, box, rangeFirst[c], rangeLast[c] + 1u, OKRT_INVALID_UINT);
        }
    }
}

/* leaf-ordered copies of the two triangle arrays (36 B and 132 B records, moved as 4-byte words) */
__global__ void gatherKernel(uint32_t n, const uint32_t* __restrict__ sortedIds, const uint32_t* __restrict__ src,
                             uint32_t wordsPerRecord, uint32_t* __restrict__ dst)
{
    const size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= (size_t)n * wordsPerRecord) return;
    const uint32_t k = (uint32_t)(w / wordsPerRecord), q = (uint32_t)(w % wordsPerRecord);
    dst[w] = src[(size_t)sortedIds[k] * wordsPerRecord + q];
}

/* device scratch of one build call; grown to the largest mesh, released at scope exit */
struct Scratch {
    std::vector<void*> blocks;
    ~Scratch() { for (void* p : blocks) cudaFree(p); }
    template <typename T> cudaError_t get(T** out, size_t count)
    {
        void* p = nullptr;
        cudaError_t e = cudaMalloc(&p, std::max<size_t>(count, 1) * sizeof(T));
        if (e == cudaSuccess) blocks.push_back(p);
        *out = (T*)p;
        return e;
    }
};

inline unsigned gridFor(size_t items, unsigned block) { return (unsigned)((items + block - 1) / block); }

} // namespace

/* f2: GPU replacement of the per-mesh loop of RayTracer::loadMeshAndBvhData (RayTracer.cpp:143-201) with an LBVH in
 * place of BvhBuilder::buildTree.  Same outputs, same index conventions as okrt_geometry_build. */
int okrt_geometry_build_device(const okrt_mesh_view* meshes, uint32_t numMeshes, uint32_t maxLeafTriangles, int device,
                               okrt_geometry** out, float* outBuildMs)
{
    using okrt::setError;
    if (!out || (!meshes && numMeshes)) { setError("okrt_geometry_build_device: null argument"); return OKRT_ERR_INVALID_ARGUMENT; }
    *out = nullptr;
    size_t total = 0; uint32_t largest = 0;
    for (uint32_t i = 0; i < numMeshes; i++) {
        if (meshes[i].numTriangles && (!meshes[i].trianglesPos || !meshes[i].trianglesInfo)) {
            setError("okrt_geometry_build_device: mesh with null triangle arrays"); return OKRT_ERR_INVALID_ARGUMENT;
        }
        if (meshes[i].numTriangles > 0x3FFFFFFFu) { setError("okrt_geometry_build_device: mesh too large"); return OKRT_ERR_CAPACITY; }
        total += meshes[i].numTriangles; largest = std::max(largest, meshes[i].numTriangles);
    }
    int deviceCount = 0;
    if (cudaGetDeviceCount(&deviceCount) != cudaSuccess || device < 0 || device >= deviceCount) {
        setError("okrt_geometry_build_device: no such CUDA device (this entry point has no CPU fallback; use okrt_geometry_build)");
        return OKRT_ERR_NO_DEVICE;
    }
    OKRT_BUILD_CUDA(cudaSetDevice(device));

    constexpr uint32_t kPosWords = sizeof(okrt_triangle) / 4, kInfoWords = sizeof(okrt_triangle_info) / 4;
    static_assert(sizeof(okrt_bvh_node) == sizeof(OutNode), "node record");
    Scratch sc;
    cudaStream_t stream = nullptr; cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    OKRT_BUILD_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    struct StreamGuard { cudaStream_t& s; cudaEvent_t& a; cudaEvent_t& b;
        ~StreamGuard() { if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); if (s) cudaStreamDestroy(s); } } guard{stream, ev0, ev1};
    OKRT_BUILD_CUDA(cudaEventCreate(&ev0)); OKRT_BUILD_CUDA(cudaEventCreate(&ev1));

    uint32_t *dPos = nullptr, *dInfo = nullptr, *dPosSorted = nullptr, *dInfoSorted = nullptr;
    uint32_t *codes = nullptr, *codesSorted = nullptr, *ids = nullptr, *idsSorted = nullptr;
    uint32_t *child = nullptr, *rangeFirst = nullptr, *rangeLast = nullptr, *nodeParent = nullptr, *leafParent = nullptr;
    uint32_t *arrivals = nullptr, *kept = nullptr, *rank = nullptr;
    float *primBox = nullptr, *nodeBox = nullptr; OutNode* outNodes = nullptr; void* cubTemp = nullptr; size_t cubBytes = 0;
    if (largest > 1) {
        const size_t n = largest;
        OKRT_BUILD_CUDA(sc.get(&dPos, n * kPosWords)); OKRT_BUILD_CUDA(sc.get(&dInfo, n * kInfoWords));
        OKRT_BUILD_CUDA(sc.get(&dPosSorted, n * kPosWords)); OKRT_BUILD_CUDA(sc.get(&dInfoSorted, n * kInfoWords));
        OKRT_BUILD_CUDA(sc.get(&codes, n)); OKRT_BUILD_CUDA(sc.get(&codesSorted, n));
        OKRT_BUILD_CUDA(sc.get(&ids, n)); OKRT_BUILD_CUDA(sc.get(&idsSorted, n));
        OKRT_BUILD_CUDA(sc.get(&child, 2 * n)); OKRT_BUILD_CUDA(sc.get(&rangeFirst, n)); OKRT_BUILD_CUDA(sc.get(&rangeLast, n));
        OKRT_BUILD_CUDA(sc.get(&nodeParent, n)); OKRT_BUILD_CUDA(sc.get(&leafParent, n));
        OKRT_BUILD_CUDA(sc.get(&arrivals, n)); OKRT_BUILD_CUDA(sc.get(&kept, n)); OKRT_BUILD_CUDA(sc.get(&rank, n));
        OKRT_BUILD_CUDA(sc.get(&primBox, 6 * n)); OKRT_BUILD_CUDA(sc.get(&nodeBox, 6 * n));
        OKRT_BUILD_CUDA(sc.get(&outNodes, 2 * n + 1));
        size_t sortBytes = 0, scanBytes = 0;
        OKRT_BUILD_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, sortBytes, codes, codesSorted, ids, idsSorted, (int)n, 0, 30, stream));
        OKRT_BUILD_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, scanBytes, kept, rank, (int)n, stream));
        cubBytes = std::max(sortBytes, scanBytes);
        uint8_t* tmp = nullptr; OKRT_BUILD_CUDA(sc.get(&tmp, cubBytes)); cubTemp = tmp;
    }

    okrt_geometry* g = new okrt_geometry();
    struct GeometryGuard { okrt_geometry* p; ~GeometryGuard() { delete p; } } owned{g};
    g->triPos.resize(total); g->triInfo.resize(total);
    g->descs.resize(numMeshes); g->bounds.resize(numMeshes);
    OKRT_BUILD_CUDA(cudaEventRecord(ev0, stream));

    constexpr unsigned kBlock = 256;
    uint32_t triBase = 0;
    std::vector<okrt_bvh_node> nodes;
    for (uint32_t m = 0; m < numMeshes; m++) {
        const okrt_mesh_view& mv = meshes[m];
        const uint32_t n = mv.numTriangles, nodeBase = (uint32_t)g->nodes.size();
        if (n <= std::max(maxLeafTriangles, 1u)) {
            /* the root is the only node (BvhBuilder.cpp:52-57: a node within the leaf budget is not split) */
            okrt_bvh_node root; root.boundingBox = mv.boundingBox; root.triStart = 0; root.triEnd = n; root.firstChildIdx = OKRT_INVALID_UINT;
            nodes.assign(1, root);
            if (n) {
                std::memcpy(g->triPos.data() + triBase, mv.trianglesPos, (size_t)n * sizeof(okrt_triangle));
                std::memcpy(g->triInfo.data() + triBase, mv.trianglesInfo, (size_t)n * sizeof(okrt_triangle_info));
            }
        } else {
            const int numInternal = (int)n - 1;
            OKRT_BUILD_CUDA(cudaMemcpyAsync(dPos, mv.trianglesPos, (size_t)n * sizeof(okrt_triangle), cudaMemcpyHostToDevice, stream));
            OKRT_BUILD_CUDA(cudaMemcpyAsync(dInfo, mv.trianglesInfo, (size_t)n * sizeof(okrt_triangle_info), cudaMemcpyHostToDevice, stream));
            const float3 lo = make_float3(mv.boundingBox.min[0], mv.boundingBox.min[1], mv.boundingBox.min[2]);
            float inv[3];
            for (int k = 0; k < 3; k++) { const float e = mv.boundingBox.max[k] - mv.boundingBox.min[k]; inv[k] = e > 0.f ? 1.0f / e : 0.f; }
            mortonKernel<<<gridFor(n, kBlock), kBlock, 0, stream>>>((const float*)dPos, n, lo, make_float3(inv[0], inv[1], inv[2]), primBox, codes, ids);
            size_t tempBytes = cubBytes;
            OKRT_BUILD_CUDA(cub::DeviceRadixSort::SortPairs(cubTemp, tempBytes, codes, codesSorted, ids, idsSorted, (int)n, 0, 30, stream));
            radixTreeKernel<<<gridFor(numInternal, kBlock), kBlock, 0, stream>>>(codesSorted, (int)n, child, rangeFirst, rangeLast, nodeParent, leafParent);
            OKRT_BUILD_CUDA(cudaMemsetAsync(arrivals, 0, (size_t)numInternal * sizeof(uint32_t), stream));
            fitKernel<<<gridFor(n, kBlock), kBlock, 0, stream>>>((int)n, idsSorted, primBox, child, nodeParent, leafParent, nodeBox, arrivals);
            keptKernel<<<gridFor(numInternal, kBlock), kBlock, 0, stream>>>(numInternal, rangeFirst, rangeLast, std::max(maxLeafTriangles, 1u), kept);
            tempBytes = cubBytes;
            OKRT_BUILD_CUDA(cub::DeviceScan::ExclusiveSum(cubTemp, tempBytes, kept, rank, numInternal, stream));
            emitKernel<<<gridFor(numInternal, kBlock), kBlock, 0, stream>>>(numInternal, kept, rank, child, rangeFirst, rangeLast, idsSorted,
                                                                            primBox, nodeBox, mv.boundingBox, outNodes);
            gatherKernel<<<gridFor((size_t)n * kPosWords, kBlock), kBlock, 0, stream>>>(n, idsSorted, dPos, kPosWords, dPosSorted);
            gatherKernel<<<gridFor((size_t)n * kInfoWords, kBlock), kBlock, 0, stream>>>(n, idsSorted, dInfo, kInfoWords, dInfoSorted);
            OKRT_BUILD_CUDA(cudaGetLastError());
            uint32_t tail[2] = {0, 0};
            OKRT_BUILD_CUDA(cudaMemcpyAsync(&tail[0], rank + numInternal - 1, sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
            OKRT_BUILD_CUDA(cudaMemcpyAsync(&tail[1], kept + numInternal - 1, sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
            OKRT_BUILD_CUDA(cudaMemcpyAsync(g->triPos.data() + triBase, dPosSorted, (size_t)n * sizeof(okrt_triangle), cudaMemcpyDeviceToHost, stream));
            OKRT_BUILD_CUDA(cudaMemcpyAsync(g->triInfo.data() + triBase, dInfoSorted, (size_t)n * sizeof(okrt_triangle_info), cudaMemcpyDeviceToHost, stream));
            OKRT_BUILD_CUDA(cudaStreamSynchronize(stream));
            const uint32_t numKept = tail[0] + tail[1];
            nodes.resize(2 * (size_t)numKept + 1);
            OKRT_BUILD_CUDA(cudaMemcpyAsync(nodes.data(), outNodes, nodes.size() * sizeof(okrt_bvh_node), cudaMemcpyDeviceToHost, stream));
            OKRT_BUILD_CUDA(cudaStreamSynchronize(stream));
        }
        for (const okrt_bvh_node& nd : nodes) {
            okrt_bvh_node gn = nd;
            if (gn.firstChildIdx != OKRT_INVALID_UINT) gn.firstChildIdx += nodeBase;
            else { gn.triStart += triBase; gn.triEnd += triBase; }
            g->nodes.push_back(gn);
        }
        g->descs[m] = okrt_mesh_desc{triBase, triBase + n, nodeBase, (uint32_t)nodes.size()};
        g->bounds[m] = mv.boundingBox;
        triBase += n;
    }
    OKRT_BUILD_CUDA(cudaEventRecord(ev1, stream));
    OKRT_BUILD_CUDA(cudaEventSynchronize(ev1));
    if (outBuildMs) OKRT_BUILD_CUDA(cudaEventElapsedTime(outBuildMs, ev0, ev1));
    owned.p = nullptr;
    *out = g;
    return OKRT_OK;
}

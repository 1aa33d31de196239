/*
 * okrt_accel.h — device-side acceleration layout ("fast layout") built at upload time from the
 * reference's canonical arrays (BvhNode 36 B, Triangle 36 B, Mesh 232 B, OctTreeNode 48 B).
 *
 * Child-pair node, 64 B, 16 B aligned, four LDG.128 per visit (BASELINE.json north_star: "node layout
 * re-packed into ... 16B-aligned AABB pairs for vectorised LDG.128 loads"):
 *     q0 = { c0.min.x, c0.min.y, c0.min.z, bits(ref0) }
 *     q1 = { c0.max.x, c0.max.y, c0.max.z, bits(ref1) }
 *     q2 = { c1.min.x, c1.min.y, c1.min.z, 0 }
 *     q3 = { c1.max.x, c1.max.y, c1.max.z, 0 }
 * The box floats are bit copies of the reference node boxes (BLAS) so that the slab test gives the
 * reference's distNear/distFar exactly.  A child reference is one 32-bit word:
 *     00 nnnnn...   interior: index of a child-pair node
 *     10 ccc sssss  triangle leaf: count-1 (3 bits, <= 8 triangles), first triangle (27 bits)
 *     11 iiiii...   instance leaf (top level only): index into the instance-record array
 * Leaves with more than 8 triangles (possible when the SAH refuses to split or maxDepth is hit) are
 * chained through extra pair nodes that repeat the leaf's box.
 *
 * Triangle: 48 B = three float4 { p2.xyz, E1.x | E1.yz, E2.xy | E2.z, bits(owner), 0, 0 } with E1 = p0 - p2 and
 * E2 = p1 - p2 (GPU-Utilities.hlsli:124; the same float32 subtraction, done once at upload); owner = instance record
 * of the merged instance the triangle belongs to (what triOwner holds), so a hit names its instance without a lookup.
 *
 * Instance record: 160 B = ten float4
 *     r0..r2  rows 0..2 of inverseTransformMatrix           (world -> local)
 *     r3..r5  rows 0..2 of transformMatrix, .w = root box min (local -> world for directions)
 *     r6      root box max.xyz, .w = bits(flags)  (bit0: transform is exactly identity)
 *     r7      bits{ rootRef, octNodeIdx, instanceIdx (reference order), flags } (flags bit0: the oct-tree boxes above the
 *             instance are nested, i.e. a point inside r8/r9 is inside every ancestor box)
 *     r8,r9   min.xyz / max.xyz of the oct-tree node that lists the instance
 * The top-level tree over instance world boxes replaces the reference's oct-tree walk for CULLING
 * only; every oct-tree box the reference would test on the way to an instance is still tested (with
 * its arithmetic) before the instance is entered, so hit/miss decisions are unchanged.
 */
#pragma once
#include <cstdint>
#include <string>
#include <vector>
#include "okrt.h"

namespace okrt {

constexpr uint32_t kRefTagMask = 0xC0000000u;
constexpr uint32_t kRefInterior = 0x00000000u;
constexpr uint32_t kRefTriLeaf = 0x80000000u;
constexpr uint32_t kRefInstance = 0xC0000000u;
constexpr uint32_t kRefNone = 0xFFFFFFFFu;        /* empty scene / empty child */
constexpr uint32_t kMaxLeafTris = 8;
constexpr uint32_t kInstQuads = 10;               /* float4 per instance record */
constexpr uint32_t kMaxTriIndex = (1u << 27);
constexpr uint32_t kTriQuads = 3;       /* float4 per triangle record (48 B: p2 | E1 | E2 | owner); 64-byte records measured slower */
constexpr uint32_t kFastStackCapacity = 64;       /* entries; validated against tree depth at upload */
constexpr uint32_t kExactBvhStackCapacity = 64;   /* reference: 50, RaytracerCS.hlsl:249 */
constexpr uint32_t kExactOctStackCapacity = 32;   /* reference: 20, RaytracerCS.hlsl:244 */


/* ---- wide layout (8-wide nodes with quantised child boxes; leaf boxes stay exact) ----------------------------------
 * Node, 96 B = 24 words, 32 B aligned (three 256-bit loads, three sectors):
 *     w0..w2   p.xyz      origin of the node's quantisation grid (a multiple of the cell size: p + q*cell is exact in float32)
 *     w3       e.x | e.y << 8 | e.z << 16 | imask << 24      cell size of axis k = 2^(e.k - 127); imask bit s: slot s is interior
 *     w4       childBase  interior children are stored contiguously, in slot order: child of slot s = childBase + popc(imask & ((1<<s)-1))
 *     w5       0
 *     w6..w17  qlo.x[8] qlo.y[8] qlo.z[8] qhi.x[8] qhi.y[8] qhi.z[8]   one byte per slot; an empty slot has qlo = 255, qhi = 0
 *     w18..23  0
 * A child's quantised box CONTAINS the culling box of everything below it: the exact box of every leaf below, grown by that
 * leaf's slack (below).  Slots are assigned so that visiting slots in the order (7 - octant) XOR priority is roughly near to far
 * for a ray of that direction octant (after Ylitie, Karras, Laine 2017, restated): no sorting in the kernel.
 * Leaf record, 32 B = 8 words, at index node * 8 + slot (one 256-bit load):
 *     w0..w5   min.xyz, max.xyz: bit copy of the reference leaf's box — tested with the reference's own arithmetic
 *     w6       count << 26 | first triangle (count 1..62), or 63 << 26 | instance record
 *     w7       slack: how far (in distance along a ray) a triangle test of this leaf may report a hit before the ray enters the
 *              leaf's box, beyond the relative window delta: 2^-13 x longest edge x edge-shape factor of its triangles */
constexpr uint32_t kWideNodeWords = 24;
constexpr uint32_t kWideLeafWords = 8;
constexpr uint32_t kWideMaxLeafTris = 62;
constexpr uint32_t kWideLeafInstance = 63;
constexpr uint32_t kWideStackCapacity = 32;       /* 8-byte entries; validated against the tree depth at upload */

struct Float4 { float x, y, z, w; };

struct FastLayout {
    std::vector<Float4> nodes;          /* 4 per pair node */
    std::vector<Float4> tris;           /* kTriQuads per triangle */
    std::vector<Float4> instances;      /* 8 per instance record, top-level order */
    std::vector<uint32_t> octParent;    /* parent index per oct-tree node (root: 0xFFFFFFFF) */
    std::vector<uint32_t> triOwner;     /* per triangle: instance record of the merged instance that owns it, else ~0u */
    uint32_t numMerged = 0;             /* instances linked straight into the top level */
    uint32_t numFlattenedLeaves = 0;    /* reference leaves that became primitives of the scene-wide SAH tree */
    uint32_t topRef = kRefNone;         /* reference to the top-level root (interior or instance leaf) */
    uint32_t maxStackDepth = 0;         /* worst-case fast-stack occupancy */
    uint32_t exactBvhDepth = 0;         /* worst-case reference BVH stack occupancy */
    uint32_t exactOctDepth = 0;         /* worst-case reference oct-tree stack occupancy */
    bool allRigid = true;               /* every instance transform has unit scale */
    std::vector<uint32_t> wideNodes;    /* kWideNodeWords per node */
    std::vector<uint32_t> wideLeaves;   /* kWideLeafWords per (node, slot) */
    uint32_t wideRoot = kRefNone;       /* root node of the scene-wide tree */
    uint32_t wideStackDepth = 0;        /* worst-case occupancy of the wide traversal's stack */
    uint32_t wideDepth = 0;             /* node levels of the scene-wide tree */
    float maxSlack = 0.f;               /* largest leaf slack of the scene */
    std::string error;
};

/* Returns false (and sets out.error) if the scene cannot be laid out. */
bool buildFastLayout(const okrt_triangle* tris, uint32_t numTris,
                     const okrt_bvh_node* nodes, uint32_t numNodes,
                     const okrt_mesh_instance* instances, uint32_t numInstances,
                     const okrt_octree_node* octNodes, uint32_t numOctNodes,
                     FastLayout& out, bool wantWide = false);

bool octTreeChecks(const okrt_octree_node* octNodes, uint32_t numOctNodes, uint32_t numInstances, std::vector<uint32_t>& octParent,
                   std::vector<uint8_t>& octNested, std::vector<uint32_t>& instOct, uint32_t& exactOctDepth, std::string& error);

} // namespace okrt

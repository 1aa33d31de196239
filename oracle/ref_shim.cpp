/*
 * ref_shim.cpp — TEST INFRASTRUCTURE.  Thin C entry points around the REFERENCE's own code.
 *
 * This file is ours; it contains no reference source.  It is compiled together with
 *   /root/reference/source/Graphics/BvhBuilder.cpp      (unmodified, from where it lies)
 * against the reference's own headers (source/, deps/include/glm) into
 * oracle/_ref/libokrt_ref.so by oracle/Makefile, only where /root/reference exists.
 * It is used to pin the oracle's restatement (tests/test_oracle_vs_ref.py), to generate the golden
 * fixtures under tests/golden/ (tests/golden/make_golden.py), and — as the reference's own CPU code
 * for row a10 — it is what the parity tests of the product's BVH builder compare against.
 *
 * The shader (HLSL) and RayTracer.cpp (D3D11) cannot be compiled here; only the parts below can.
 */
#include <cstdint>
#include <cstring>
#include <vector>

#include "Graphics/BvhBuilder.h"      /* reference: source/Graphics/BvhBuilder.h (+ Mesh.h, Utilities.h) */
#include "Scene/Components.h"         /* reference: source/Scene/Components.h (Transform, Material, ...) */
#include "glm/gtc/matrix_transform.hpp"

#define REF_API extern "C" __attribute__((visibility("default")))

struct ShimNode { float mn[3]; float mx[3]; uint32_t triStart, triEnd, firstChildIdx; };

/* BvhBuilder::buildTree on one mesh, then the flattening loop of RayTracer::loadMeshAndBvhData
 * (RayTracer.cpp:172-199) with mesh-local indices. */
REF_API uint32_t ref_build_bvh(const float* triPos /* n*9 */, uint32_t numTris, const float* bbMin, const float* bbMax,
                               uint32_t maxDepth, uint32_t maxLeafTriangles, ShimNode* outNodes, uint32_t nodeCapacity,
                               uint32_t* outTriOrder)
{
    MeshData md;
    md.positions.resize((size_t)numTris * 3);
    md.normals.assign((size_t)numTris * 3, glm::vec3(0.f));
    md.uvs.assign((size_t)numTris * 3, glm::vec2(0.f));
    md.tangents.assign((size_t)numTris * 3, glm::vec3(0.f));
    md.bitangents.assign((size_t)numTris * 3, glm::vec3(0.f));
    for (size_t i = 0; i < (size_t)numTris * 3; i++)
        md.positions[i] = glm::vec3(triPos[i * 3 + 0], triPos[i * 3 + 1], triPos[i * 3 + 2]);
    md.boundingBox.min = glm::vec3(bbMin[0], bbMin[1], bbMin[2]);
    md.boundingBox.max = glm::vec3(bbMax[0], bbMax[1], bbMax[2]);
    Mesh mesh(md, "shim");

    BvhBuilder builder(maxLeafTriangles, maxDepth);
    builder.buildTree(mesh);
    const std::vector<BvhNode>& nodes = builder.getTree();
    if (nodes.size() > nodeCapacity) return ~0u;

    uint32_t localTriStart = 0u;
    for (uint32_t k = 0; k < (uint32_t)nodes.size(); k++) {
        ShimNode& g = outNodes[k];
        const BvhNode& n = nodes[k];
        for (int a = 0; a < 3; a++) { g.mn[a] = n.boundingBox.min[a]; g.mx[a] = n.boundingBox.max[a]; }
        g.triStart = g.triEnd = Okay::INVALID_UINT;
        g.firstChildIdx = n.firstChildIdx;
        if (!n.isLeaf()) continue;
        g.triStart = localTriStart;
        g.triEnd = localTriStart + (uint32_t)n.triIndicies.size();
        for (uint32_t ti : n.triIndicies) outTriOrder[localTriStart++] = ti;
    }
    return (uint32_t)nodes.size();
}

/* glm::transpose(Transform::calculateMatrix()) as 16 floats (what RayTracer.cpp:747-748 uploads), and its glm::inverse */
REF_API void ref_transform_matrices(const float* pos, const float* rotDeg, const float* scale, float* outT16, float* outInv16)
{
    Transform t;
    t.position = glm::vec3(pos[0], pos[1], pos[2]);
    t.rotation = glm::vec3(rotDeg[0], rotDeg[1], rotDeg[2]);
    t.scale = glm::vec3(scale[0], scale[1], scale[2]);
    glm::mat4 m = glm::transpose(t.calculateMatrix());
    glm::mat4 inv = glm::inverse(m);
    std::memcpy(outT16, &m[0][0], 64);
    std::memcpy(outInv16, &inv[0][0], 64);
}

/* The glm calls of RayTracer::calculateProjectionData (RayTracer.cpp:273-307), in its order.
 * out: viewPlaneDims[2], invProj[16], invView[16], right[3], up[3] */
REF_API void ref_projection_data(const float* camPos, const float* camRotDeg, float fov, float nearZ, float farZ,
                                 uint32_t width, uint32_t height, float* viewPlaneDims, float* invProj16, float* invView16,
                                 float* right3, float* up3)
{
    Transform camTra;
    camTra.position = glm::vec3(camPos[0], camPos[1], camPos[2]);
    camTra.rotation = glm::vec3(camRotDeg[0], camRotDeg[1], camRotDeg[2]);
    Camera camData(fov, nearZ);
    camData.farZ = farZ;

    const glm::vec2 windowDimsVec((float)width, (float)height);
    const float aspectRatio = windowDimsVec.x / windowDimsVec.y;
    const float sideB = camData.nearZ;
    const float vAngleRadians = glm::radians(camData.fov * 0.5f);
    const float sideA = glm::tan(vAngleRadians) * sideB;
    viewPlaneDims[0] = sideA * 2.f;
    viewPlaneDims[1] = viewPlaneDims[0] / aspectRatio;

    glm::mat4 ip = glm::transpose(glm::inverse(
        glm::perspectiveFovLH(glm::radians(camData.fov), windowDimsVec.x, windowDimsVec.y, camData.nearZ, camData.farZ)));
    std::memcpy(invProj16, &ip[0][0], 64);

    glm::vec3 camForward = camTra.getForwardVec();
    glm::vec3 camRight = camTra.getRightVec();
    glm::vec3 camUp = camTra.getUpVec();
    for (int k = 0; k < 3; k++) { right3[k] = camRight[k]; up3[k] = camUp[k]; }

    glm::mat4 iv = glm::transpose(glm::inverse(
        glm::lookAtLH(camTra.position, camTra.position + camForward, glm::vec3(0.f, 1.f, 0.f))));
    std::memcpy(invView16, &iv[0][0], 64);
}

/* Okay::AABB helpers used by the oct-tree build (Utilities.h:72-93) */
REF_API float ref_aabb_area(const float* mn, const float* mx)
{
    return Okay::AABB(glm::vec3(mn[0], mn[1], mn[2]), glm::vec3(mx[0], mx[1], mx[2])).getArea();
}
REF_API int ref_aabb_intersects(const float* amn, const float* amx, const float* bmn, const float* bmx)
{
    return Okay::AABB::intersects(Okay::AABB(glm::vec3(amn[0], amn[1], amn[2]), glm::vec3(amx[0], amx[1], amx[2])),
                                  Okay::AABB(glm::vec3(bmn[0], bmn[1], bmn[2]), glm::vec3(bmx[0], bmx[1], bmx[2]))) ? 1 : 0;
}

/* sizeof checks of the reference's CPU mirrors */
REF_API void ref_struct_sizes(uint32_t* out)
{
    out[0] = sizeof(Okay::Triangle); out[1] = sizeof(Okay::VertexInfo); out[2] = sizeof(Okay::TriangleInfo);
    out[3] = sizeof(Okay::AABB); out[4] = sizeof(Material); out[5] = sizeof(GPU_MeshComponent);
    out[6] = sizeof(glm::vec3) + sizeof(Sphere); out[7] = sizeof(GPU_DirectionalLight); out[8] = sizeof(GPU_PointLight);
    out[9] = sizeof(GPU_SpotLight);
}

/* Material{} defaults, Components.h:42-56 */
REF_API void ref_material_default(void* out68)
{
    Material m;
    std::memcpy(out68, &m, sizeof(Material));
}

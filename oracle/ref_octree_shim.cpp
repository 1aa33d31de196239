/*
 * ref_octree_shim.cpp — TEST INFRASTRUCTURE.  C entry point around the REFERENCE's own oct-tree builder.
 *
 * This file is ours; it contains no reference source.  oracle/Makefile extracts the text of
 *   RayTracer::refitOctTreeNode / createOctTree / loadOctTree   (/root/reference/source/Graphics/RayTracer.cpp:503-788)
 * at build time into oracle/_ref/octree_extract.inc (git-ignored, deleted after the compile) and compiles it here, unmodified,
 * as members of the reference's own `class RayTracer` (its header is included as it lies), together with the reference's
 * Scene.cpp, EnTT (vendored header), glm, Components.h, Mesh.h, ResourceManager.h.  What is NOT the reference's: <d3d11.h> and
 * <DirectXCollision.h> (oracle/ref_stubs: opaque types, BoundingBox::GetCorners) and the handful of D3D-side definitions below
 * that the three functions link against but whose effects are irrelevant here (GPUStorage::updateRaw, Okay::updateBuffer).
 * The rest of RayTracer.cpp (D3D11 device code) is not compiled.
 * Result: row a11 of SURVEY.md §8 is pinned to the reference itself (tests/test_cpu_builders.py, tests/golden/octree_*.npz).
 */
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <cstring>
#include <string_view>
#include <stack>
#include <vector>

#define private public          /* the shim fills ResourceManager::m_meshes / RayTracer::m_meshDescs and reads the results */
#include "Graphics/ResourceManager.h"
#include "Graphics/RayTracer.h"
#undef private
#include "Scene/Scene.h"
#include "Scene/Components.h"
#include "glm/gtc/matrix_transform.hpp"
#include "glm/gtx/quaternion.hpp"
#include <DirectXCollision.h>

/* ---- D3D-side symbols the class needs to link; none of them does anything */
namespace Okay {
ID3D11DeviceContext* getDeviceContext() { static ID3D11DeviceContext ctx; return &ctx; }
void updateBuffer(ID3D11Buffer*, const void*, size_t) {}
}
GPUStorage::GPUStorage() : m_pBuffer(nullptr), m_pSRV(nullptr), m_capacity(0), m_elementByteWidth(0) {}
GPUStorage::~GPUStorage() {}
void GPUStorage::updateRaw(uint32_t, const void*) {}
/* initiate keeps the initial data a caller hands over (RayTracer::loadMeshAndBvhData's triangle arrays); no buffer is created */
static std::map<const GPUStorage*, std::vector<uint8_t>> g_storageData;
void GPUStorage::initiate(uint32_t elementByteWidth, uint32_t capacity, const void* pData)
{
    std::vector<uint8_t>& v = g_storageData[this];
    v.clear();
    if (pData) v.assign((const uint8_t*)pData, (const uint8_t*)pData + (size_t)elementByteWidth * capacity);
}
RenderTexture::RenderTexture() {}
RenderTexture::~RenderTexture() {}
RayTracer::RayTracer() : m_pScene(nullptr), m_pResourceManager(nullptr), m_pTargetTexture(nullptr), m_pRenderDataBuffer(nullptr),
                         m_pMainRaytracingCS(nullptr), m_pTextures(nullptr), m_pEnvironmentMapSRV(nullptr) {}
RayTracer::~RayTracer() {}

/* ---- the reference's own text: RayTracer.cpp:503-788 */
static int okrt_silent_printf(const char*, ...) { return 0; }
#define printf okrt_silent_printf
#include "octree_extract.inc"
#undef printf

/* ---- the reference's own text: RayTracer::loadEnvironmentMap, RayTracer.cpp:343-432 (row a13, the cube part).  stb_image and
 * the D3D11 device are replaced: stbi_load hands out the caller's cross image, ID3D11Device::CreateTexture2D (ref_stubs/d3d11.h)
 * passes the six faces the function has cut out of it to the sink below. */
OkrtTextureSink g_okrtTextureSink = {nullptr};
namespace Okay { ID3D11Device* getDevice() { static ID3D11Device dev; return &dev; } }
enum { STBI_rgb_alpha = 4 };
static const uint8_t* g_crossImage = nullptr;
static int g_crossW = 0, g_crossH = 0;
static unsigned char* stbi_load(const char*, int* w, int* h, int*, int channels)
{
    if (!g_crossImage) return nullptr;
    const size_t bytes = (size_t)g_crossW * g_crossH * channels;
    unsigned char* copy = (unsigned char*)std::malloc(bytes);
    std::memcpy(copy, g_crossImage, bytes);
    *w = g_crossW; *h = g_crossH;
    return copy;
}
static void stbi_image_free(void* p) { std::free(p); }
#include "envmap_extract.inc"

/* ---- the reference's own text: RayTracer::calculateProjectionData, RayTracer.cpp:268-308 (row a12) */
#include "projection_extract.inc"

/* ---- the reference's own text: RayTracer::loadMeshAndBvhData, RayTracer.cpp:120-217 (row a10: the flattening of the per-mesh
 * trees into t0 / t1 / t2 and the mesh descriptors); BvhBuilder.cpp is compiled into this library unmodified */
#include "Graphics/BvhBuilder.h"
#define printf okrt_silent_printf
#include "meshload_extract.inc"
#undef printf

#define REF_API extern "C" __attribute__((visibility("default")))

/* The reference's whole constant buffer (RayTracer::RenderData, 240 bytes: its own default member initialisers plus what
 * calculateProjectionData writes) for a camera entity with this Transform / Camera and a target of width x height. */
static_assert(sizeof(RayTracer::RenderData) == 240, "cbuffer");
REF_API void ref_render_data_after_projection(const float* camPos, const float* camRotDeg, float fov, float nearZ, float farZ,
                                              uint32_t width, uint32_t height, void* out240)
{
    Scene scene;
    Entity cam = scene.createEntity();
    Transform& t = cam.getComponent<Transform>();
    t.position = glm::vec3(camPos[0], camPos[1], camPos[2]);
    t.rotation = glm::vec3(camRotDeg[0], camRotDeg[1], camRotDeg[2]);
    Camera& c = cam.addComponent<Camera>(fov, nearZ);
    c.farZ = farZ;
    RayTracer rt;
    rt.setScene(scene);
    rt.m_renderData.textureDims = glm::uvec2(width, height);
    rt.calculateProjectionData();
    std::memcpy(out240, &rt.m_renderData, 240);
}

struct ShimMeshView { const float* trianglesPos; const float* trianglesInfo; uint32_t numTriangles; float bounds[6]; };
static_assert(sizeof(Okay::Triangle) == 36 && sizeof(Okay::TriangleInfo) == 132 && sizeof(GPUNode) == 36 && sizeof(MeshDesc) == 16, "wire structs");

/* views: as okrt_mesh_view (include/okrt.h).  Outputs: t0 (36 B per triangle), t1 (132 B), t2 (36 B per node), 16-byte mesh
 * descriptors.  Returns the node count (~0u if a capacity is too small); *outNumTris = triangles written. */
REF_API uint32_t ref_load_mesh_and_bvh(const ShimMeshView* views, uint32_t numMeshes, uint32_t maxDepth, uint32_t maxLeafTriangles,
                                       void* outTriPos, void* outTriInfo, uint32_t triCapacity, void* outNodes, uint32_t nodeCapacity,
                                       void* outDescs, uint32_t* outNumTris)
{
    ResourceManager rm;
    for (uint32_t m = 0; m < numMeshes; m++) {
        const uint32_t nv = views[m].numTriangles * 3u;
        MeshData md;
        md.positions.resize(nv); md.normals.resize(nv); md.uvs.resize(nv); md.tangents.resize(nv); md.bitangents.resize(nv);
        for (uint32_t v = 0; v < nv; v++) {
            const float* p = views[m].trianglesPos + (size_t)v * 3;
            const float* q = views[m].trianglesInfo + (size_t)v * 11;
            md.positions[v] = glm::vec3(p[0], p[1], p[2]);
            md.normals[v] = glm::vec3(q[0], q[1], q[2]);
            md.uvs[v] = glm::vec2(q[3], q[4]);
            md.tangents[v] = glm::vec3(q[5], q[6], q[7]);
            md.bitangents[v] = glm::vec3(q[8], q[9], q[10]);
        }
        md.boundingBox.min = glm::vec3(views[m].bounds[0], views[m].bounds[1], views[m].bounds[2]);
        md.boundingBox.max = glm::vec3(views[m].bounds[3], views[m].bounds[4], views[m].bounds[5]);
        rm.m_meshes.emplace_back(md, "shim");
    }
    RayTracer rt;
    rt.m_pResourceManager = &rm;
    rt.loadMeshAndBvhData(maxDepth, maxLeafTriangles);
    const std::vector<uint8_t>& pos = g_storageData[&rt.m_trianglePositions];
    const std::vector<uint8_t>& info = g_storageData[&rt.m_triangleInfo];
    const uint32_t numTris = (uint32_t)(pos.size() / sizeof(Okay::Triangle));
    uint32_t result = (uint32_t)rt.m_bvhTreeNodes.size();
    if (numTris > triCapacity || rt.m_bvhTreeNodes.size() > nodeCapacity) result = ~0u;
    else {
        std::memcpy(outTriPos, pos.data(), pos.size());
        std::memcpy(outTriInfo, info.data(), info.size());
        std::memcpy(outNodes, rt.m_bvhTreeNodes.data(), rt.m_bvhTreeNodes.size() * sizeof(GPUNode));
        std::memcpy(outDescs, rt.m_meshDescs.data(), rt.m_meshDescs.size() * sizeof(MeshDesc));
        *outNumTris = numTris;
    }
    g_storageData.erase(&rt.m_trianglePositions); g_storageData.erase(&rt.m_triangleInfo); g_storageData.erase(&rt.m_bvhTree);
    return result;
}

static uint8_t* g_facesOut = nullptr;
static uint32_t g_faceW = 0, g_faceH = 0;
static void takeCubeFaces(const D3D11_TEXTURE2D_DESC* desc, const D3D11_SUBRESOURCE_DATA* data)
{
    g_faceW = desc->Width; g_faceH = desc->Height;
    if (!g_facesOut || desc->ArraySize != 6u) return;
    for (uint32_t f = 0; f < 6u; f++)
        for (uint32_t y = 0; y < desc->Height; y++)
            std::memcpy(g_facesOut + ((size_t)f * desc->Height + y) * desc->Width * 4, (const uint8_t*)data[f].pSysMem + (size_t)y * data[f].SysMemPitch,
                        (size_t)desc->Width * 4);
}

/* crossRgba8: imgW x imgH RGBA8 (what stbi_load returns for the cross image); outFaces: 6 x (imgH/3) x (imgW/4) RGBA8 in the
 * order the reference fills the cube's array slices (+X -X +Y -Y +Z -Z).  Returns faceW << 16 | faceH (0 on failure). */
REF_API uint32_t ref_environment_cross_to_faces(const uint8_t* crossRgba8, uint32_t imgW, uint32_t imgH, uint8_t* outFaces)
{
    g_crossImage = crossRgba8; g_crossW = (int)imgW; g_crossH = (int)imgH;
    g_facesOut = outFaces; g_faceW = g_faceH = 0;
    g_okrtTextureSink.take = takeCubeFaces;
    RayTracer rt;
    rt.loadEnvironmentMap("cross");
    g_okrtTextureSink.take = nullptr; g_crossImage = nullptr; g_facesOut = nullptr;
    return (g_faceW << 16) | g_faceH;
}

struct ShimMeshEntity { float position[3], rotation[3], scale[3]; Material material; uint32_t meshID; };
struct ShimSphereEntity { float position[3], rotation[3], scale[3]; Material material; float radius; };
struct ShimMeshDesc { uint32_t startIdx, endIdx, bvhTreeStartIdx, numBvhNodes; };
static_assert(sizeof(GPU_OctTreeNode) == 48 && sizeof(GPU_MeshComponent) == 232, "wire structs");

/* Same arguments as oracle_build_octree (oracle/okrt_oracle.cpp) and okrt_octree_build's inputs.  Returns the node count
 * (~0u if a capacity is too small). */
REF_API uint32_t ref_build_octree(const ShimMeshDesc* descs, const float* meshBounds /* 6 per mesh */, uint32_t numMeshes,
                                  const ShimMeshEntity* meshEnts, uint32_t numMeshEnts, const ShimSphereEntity* sphereEnts, uint32_t numSphereEnts,
                                  uint32_t maxDepth, uint32_t maxLeafObjects, void* outNodes, uint32_t nodeCapacity,
                                  void* outInstances, uint32_t instanceCapacity, uint32_t* outNumInstances)
{
    ResourceManager rm;
    for (uint32_t m = 0; m < numMeshes; m++) {
        MeshData md;
        md.boundingBox.min = glm::vec3(meshBounds[m * 6 + 0], meshBounds[m * 6 + 1], meshBounds[m * 6 + 2]);
        md.boundingBox.max = glm::vec3(meshBounds[m * 6 + 3], meshBounds[m * 6 + 4], meshBounds[m * 6 + 5]);
        rm.m_meshes.emplace_back(md, "shim");
    }
    Scene scene;
    /* The arrays are given in the order the registry views ENUMERATE the entities (the contract of okrt_octree_build, include/okrt.h).
     * EnTT walks a pool from its newest element to its oldest, so the entities are created last-to-first. */
    for (uint32_t e = numMeshEnts; e-- > 0u;) {
        Entity ent = scene.createEntity();
        Transform& t = ent.getComponent<Transform>();
        t.position = glm::vec3(meshEnts[e].position[0], meshEnts[e].position[1], meshEnts[e].position[2]);
        t.rotation = glm::vec3(meshEnts[e].rotation[0], meshEnts[e].rotation[1], meshEnts[e].rotation[2]);
        t.scale = glm::vec3(meshEnts[e].scale[0], meshEnts[e].scale[1], meshEnts[e].scale[2]);
        MeshComponent& mc = ent.addComponent<MeshComponent>();
        mc.meshID = meshEnts[e].meshID;
        mc.material = meshEnts[e].material;
    }
    for (uint32_t e = numSphereEnts; e-- > 0u;) {
        Entity ent = scene.createEntity();
        Transform& t = ent.getComponent<Transform>();
        t.position = glm::vec3(sphereEnts[e].position[0], sphereEnts[e].position[1], sphereEnts[e].position[2]);
        Sphere& sp = ent.addComponent<Sphere>();
        sp.material = sphereEnts[e].material;
        sp.radius = sphereEnts[e].radius;
    }
    RayTracer rt;
    rt.m_pResourceManager = &rm;
    rt.setScene(scene);
    rt.m_meshDescs.resize(numMeshes);
    for (uint32_t m = 0; m < numMeshes; m++) {
        rt.m_meshDescs[m].startIdx = descs[m].startIdx; rt.m_meshDescs[m].endIdx = descs[m].endIdx;
        rt.m_meshDescs[m].bvhTreeStartIdx = descs[m].bvhTreeStartIdx; rt.m_meshDescs[m].numBvhNodes = descs[m].numBvhNodes;
    }
    rt.createOctTree(scene, maxDepth, maxLeafObjects);
    if (rt.m_octTreeNodes.size() > nodeCapacity || rt.m_gpuMeshes.size() > instanceCapacity) return ~0u;
    std::memcpy(outNodes, rt.m_octTreeNodes.data(), rt.m_octTreeNodes.size() * sizeof(GPU_OctTreeNode));
    std::memcpy(outInstances, rt.m_gpuMeshes.data(), rt.m_gpuMeshes.size() * sizeof(GPU_MeshComponent));
    *outNumInstances = (uint32_t)rt.m_gpuMeshes.size();
    return (uint32_t)rt.m_octTreeNodes.size();
}

/*
 * RayTracer.hpp — C++ host-side mirror of the reference's `RayTracer` class (source/Graphics/RayTracer.h:71-193)
 * on top of the C ABI (include/okrt.h).  Header only.  Same method names, argument meaning and call order:
 *
 *     okrt::ResourceManager rm;  rm.addMesh(...); rm.addTexture(...);
 *     okrt::Scene scene;         scene.meshEntities.push_back(...); scene.camera = ...;
 *     okrt::RayTracer rt;        rt.initiate(1920, 1080, rm, envFaces, faceW, faceH);   // RayTracer.cpp:60
 *     rt.setScene(scene);
 *     rt.createOctTree(scene, 2, 50);                                                  // Application.cpp:104
 *     rt.render();                                                                     // RayTracer.cpp:219
 *
 * Errors: the reference asserts (OKAY_ASSERT); here every failing ABI call throws okrt::Error with the
 * library's message.  There is no CPU fallback anywhere in this header.
 */
#pragma once
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "okrt.h"

namespace okrt {

struct Error : std::runtime_error {
    int code;
    Error(int c, const std::string& what) : std::runtime_error(what), code(c) {}
};
inline void check(int rc) { if (rc != OKRT_OK) throw Error(rc, okrt_last_error()); }

/* source/Graphics/Mesh.h:17-58 */
struct Mesh {
    std::vector<okrt_triangle> trianglesPos;
    std::vector<okrt_triangle_info> trianglesInfo;
    okrt_aabb boundingBox;
};

/* source/Graphics/ResourceManager.h: asset id == load order */
struct Texture { std::vector<uint8_t> rgba8; uint32_t width = 0, height = 0; };
struct ResourceManager {
    std::vector<Mesh> meshes;
    std::vector<Texture> textures;
    uint32_t addMesh(Mesh m) { meshes.push_back(std::move(m)); return (uint32_t)meshes.size() - 1; }
    uint32_t addTexture(Texture t) { textures.push_back(std::move(t)); return (uint32_t)textures.size() - 1; }
};

/* the EnTT registry as plain arrays, in registry-view order (source/Scene) */
struct Scene {
    std::vector<okrt_mesh_entity> meshEntities;
    std::vector<okrt_sphere_entity> sphereEntities;
    std::vector<okrt_directional_light> directionalLights;
    std::vector<okrt_point_light> pointLights;
    std::vector<okrt_spot_light> spotLights;
    okrt_camera camera{{0, 0, 0}, {0, 0, 0}, 90.f, 0.1f, 1000.f};     /* Application.cpp:67-70, Components.h:130-139 */
};

class RayTracer {
public:
    enum DebugDisplayMode : uint32_t { None = 0, BBCheckCount = 1, TriCheckCount = 2 };   /* RayTracer.h:74-79 */

    RayTracer() = default;
    RayTracer(const RayTracer&) = delete;
    RayTracer& operator=(const RayTracer&) = delete;
    ~RayTracer() { shutdown(); }

    void shutdown()                                                   /* RayTracer.cpp:34-58 */
    {
        if (m_geometry) okrt_geometry_free(m_geometry);
        if (m_handle) okrt_destroy(m_handle);
        m_geometry = nullptr; m_handle = nullptr;
    }

    /* RayTracer::initiate (RayTracer.cpp:60-118); environmentFaces: 6 x faceH x faceW RGBA8 or nullptr */
    void initiate(uint32_t width, uint32_t height, const ResourceManager& resourceManager,
                  const uint8_t* environmentFaces = nullptr, uint32_t faceW = 0, uint32_t faceH = 0, int device = 0)
    {
        shutdown();
        m_pResourceManager = &resourceManager;
        check(okrt_create(device, width, height, &m_handle));
        okrt_render_data_default(&m_renderData);
        m_renderData.textureDims[0] = width; m_renderData.textureDims[1] = height;
        okrt_options_default(&m_options);
        loadTextureData();
        if (environmentFaces) check(okrt_set_environment_map(m_handle, environmentFaces, faceW, faceH));
        loadMeshAndBvhData(30, 5);                                    /* RayTracer.cpp:97 */
    }

    /* RayTracer::loadMeshAndBvhData (RayTracer.cpp:120-217) */
    void loadMeshAndBvhData(uint32_t maxDepth, uint32_t maxLeafTriangles)
    {
        const auto& meshes = m_pResourceManager->meshes;
        if (meshes.empty()) return;
        std::vector<okrt_mesh_view> views(meshes.size());
        for (size_t i = 0; i < meshes.size(); i++)
            views[i] = okrt_mesh_view{meshes[i].trianglesPos.data(), meshes[i].trianglesInfo.data(),
                                      (uint32_t)meshes[i].trianglesPos.size(), meshes[i].boundingBox};
        if (m_geometry) okrt_geometry_free(m_geometry);
        m_geometry = nullptr;
        check(okrt_geometry_build(views.data(), (uint32_t)views.size(), maxDepth, maxLeafTriangles, &m_geometry));
        check(okrt_set_triangles(m_handle, okrt_geometry_triangles_pos(m_geometry), okrt_geometry_triangles_info(m_geometry),
                                 okrt_geometry_num_triangles(m_geometry)));
        check(okrt_set_bvh_nodes(m_handle, okrt_geometry_bvh_nodes(m_geometry), okrt_geometry_num_nodes(m_geometry)));
    }

    /* RayTracer::createOctTree + loadOctTree (RayTracer.cpp:523-788) */
    void createOctTree(const Scene& scene, uint32_t maxDepth, uint32_t maxLeafObjects)
    {
        m_pScene = &scene;
        okrt_octree* tree = nullptr;
        check(okrt_octree_build(m_geometry, scene.meshEntities.data(), (uint32_t)scene.meshEntities.size(),
                                scene.sphereEntities.data(), (uint32_t)scene.sphereEntities.size(), maxDepth, maxLeafObjects, &tree));
        int rc = okrt_set_mesh_instances(m_handle, okrt_octree_instances(tree), okrt_octree_num_instances(tree));
        if (rc == OKRT_OK) rc = okrt_set_octree(m_handle, okrt_octree_nodes(tree), okrt_octree_num_nodes(tree));
        okrt_octree_free(tree);
        check(rc);
        m_renderData.numMeshes = 0u;                                  /* RayTracer.cpp:780 */
        m_renderData.numSpheres = 0u;                                 /* :781 — spheres are not fed at HEAD */
    }

    void setScene(const Scene& scene) { m_pScene = &scene; }          /* RayTracer.h:195 */
    const std::vector<okrt_mesh_desc> getMeshDescriptors() const      /* RayTracer.h:211 */
    {
        const okrt_mesh_desc* d = okrt_geometry_mesh_descs(m_geometry);
        return std::vector<okrt_mesh_desc>(d, d + okrt_geometry_num_meshes(m_geometry));
    }

    /* RayTracer::render (RayTracer.cpp:219-260) */
    void render()
    {
        calculateProjectionData();
        updateBuffers();
        check(okrt_render(m_handle));
    }

    void toggleAccumulation(bool enable) { check(okrt_toggle_accumulation(m_handle, enable ? 1 : 0)); }   /* RayTracer.h:197 */
    void resetAccumulation() { check(okrt_reset_accumulation(m_handle)); }                                 /* :203 */
    uint32_t getNumAccumulationFrames() const { return okrt_get_num_accumulation_frames(m_handle); }       /* :211 */
    float& getDOFStrength() { return m_renderData.dofStrength; }
    float& getDOFDistance() { return m_renderData.dofDistance; }
    void setDebugMode(DebugDisplayMode mode) { m_renderData.debugMode = mode; }
    uint32_t& getDebugMaxCount() { return m_renderData.debugMaxCount; }

    void onResize(uint32_t width, uint32_t height)                    /* RayTracer.cpp:494-501 */
    {
        check(okrt_resize(m_handle, width, height));
        m_renderData.textureDims[0] = width; m_renderData.textureDims[1] = height;
    }

    /* not in the reference: NUM_BOUNCES is a #define there (RaytracerCS.hlsl:6) */
    okrt_options& options() { return m_options; }
    void applyOptions() { check(okrt_set_options(m_handle, &m_options)); }
    void readResult(std::vector<uint8_t>& rgba8)
    {
        rgba8.resize((size_t)m_renderData.textureDims[0] * m_renderData.textureDims[1] * 4);
        check(okrt_read_result(m_handle, rgba8.data()));
    }
    okrt_renderer* handle() { return m_handle; }
    /* new: `frames` frames in one call (okrt_render_frames), the non-blocking read-back, and the single exchange step of the
     * sample-sharded mode (include/okrt.h: okrt_read_result_async, okrt_reduce_accumulation) */
    void renderFrames(uint32_t frames)
    {
        calculateProjectionData();
        updateBuffers();
        check(okrt_render_frames(m_handle, frames));
    }
    void readResultAsync(uint8_t* pinnedRgba8, uint32_t slot) { check(okrt_read_result_async(m_handle, pinnedRgba8, nullptr, slot)); }
    void waitRead(uint32_t slot) { check(okrt_wait_read(m_handle, slot)); }
    void reduceAccumulation(okrt_allreduce_fn allreduce, void* user, uint32_t totalFrames) { check(okrt_reduce_accumulation(m_handle, allreduce, user, totalFrames)); }
    void sync() { check(okrt_sync(m_handle)); }

    /* RayTracer::getGlobalNodeIdx (RayTracer.h:226-230) */
    uint32_t getGlobalNodeIdx(uint32_t meshID, uint32_t localNodeIdx) const
    {
        return okrt_geometry_mesh_descs(m_geometry)[meshID].bvhTreeStartIdx + localNodeIdx;
    }

    /* the "Rebuild BVH" path with the tree built on the GPU (row f2): reference node format, not the reference's tree */
    float loadMeshAndBvhDataOnDevice(uint32_t maxLeafTriangles, int device = 0)
    {
        const auto& meshes = m_pResourceManager->meshes;
        if (meshes.empty()) return 0.f;
        std::vector<okrt_mesh_view> views(meshes.size());
        for (size_t i = 0; i < meshes.size(); i++)
            views[i] = okrt_mesh_view{meshes[i].trianglesPos.data(), meshes[i].trianglesInfo.data(),
                                      (uint32_t)meshes[i].trianglesPos.size(), meshes[i].boundingBox};
        if (m_geometry) okrt_geometry_free(m_geometry);
        m_geometry = nullptr;
        float ms = 0.f;
        check(okrt_geometry_build_device(views.data(), (uint32_t)views.size(), maxLeafTriangles, device, &m_geometry, &ms));
        check(okrt_set_triangles(m_handle, okrt_geometry_triangles_pos(m_geometry), okrt_geometry_triangles_info(m_geometry),
                                 okrt_geometry_num_triangles(m_geometry)));
        check(okrt_set_bvh_nodes(m_handle, okrt_geometry_bvh_nodes(m_geometry), okrt_geometry_num_nodes(m_geometry)));
        return ms;
    }

    /* Application::saveScreenshot (Application.cpp:661-674) */
    void saveScreenshot(const char* path = "render.png")
    {
        std::vector<uint8_t> rgba8;
        readResult(rgba8);
        check(okrt_write_png(path, rgba8.data(), m_renderData.textureDims[0], m_renderData.textureDims[1]));
    }

    /* DebugRenderer::renderBvhNodeBBs (DebugRenderer.cpp:264-303); drawMode = DrawMode enum value (0-3) */
    void renderBvhNodeBBs(const okrt_mesh_entity& entity, uint32_t localNodeIdx, uint32_t drawMode)
    {
        okrt_overlay* ov = nullptr;
        check(okrt_overlay_build(OKRT_OVERLAY_BVH, okrt_geometry_bvh_nodes(m_geometry), okrt_geometry_num_nodes(m_geometry),
                                 getGlobalNodeIdx(entity.meshID, localNodeIdx), drawMode, &entity.transform, &m_pScene->camera,
                                 m_renderData.textureDims[0], m_renderData.textureDims[1], &ov));
        const int rc = okrt_draw_overlay(m_handle, ov);
        okrt_overlay_free(ov);
        check(rc);
    }

private:
    void loadTextureData()                                            /* RayTracer.cpp:310-342 */
    {
        const auto& tex = m_pResourceManager->textures;
        if (tex.empty()) return;
        std::vector<uint32_t> w(tex.size()), h(tex.size());
        for (size_t i = 0; i < tex.size(); i++) { w[i] = tex[i].width; h[i] = tex[i].height; }
        const uint32_t n = okrt_texture_array_size(w.data(), h.data(), (uint32_t)tex.size());
        std::vector<uint8_t> layers((size_t)n * n * 4 * tex.size());
        for (size_t i = 0; i < tex.size(); i++)
            check(okrt_scale_texture(tex[i].rgba8.data(), tex[i].width, tex[i].height, layers.data() + i * (size_t)n * n * 4, n, n));
        check(okrt_set_textures(m_handle, layers.data(), n, (uint32_t)tex.size()));
    }

    void calculateProjectionData()                                    /* RayTracer.cpp:268-308 */
    {
        check(okrt_calculate_projection_data(&m_pScene->camera, m_renderData.textureDims[0], m_renderData.textureDims[1], &m_renderData));
    }

    void updateBuffers()                                              /* RayTracer.cpp:434-492 */
    {
        const Scene& s = *m_pScene;
        check(okrt_set_directional_lights(m_handle, s.directionalLights.data(), (uint32_t)s.directionalLights.size()));
        check(okrt_set_point_lights(m_handle, s.pointLights.data(), (uint32_t)s.pointLights.size()));
        check(okrt_set_spot_lights(m_handle, s.spotLights.data(), (uint32_t)s.spotLights.size()));
        m_renderData.numDirLights = (uint32_t)s.directionalLights.size();
        m_renderData.numPointLights = (uint32_t)s.pointLights.size();
        m_renderData.numSpotLights = (uint32_t)s.spotLights.size();
        /* the frame counter lives in the library (it advances inside okrt_render, RayTracer.cpp:489) */
        okrt_render_data cur;
        check(okrt_get_render_data(m_handle, &cur));
        m_renderData.numAccumulationFrames = cur.numAccumulationFrames;
        m_renderData.accumulationEnabled = cur.accumulationEnabled;
        check(okrt_set_render_data(m_handle, &m_renderData));
    }

    const Scene* m_pScene = nullptr;
    const ResourceManager* m_pResourceManager = nullptr;
    okrt_renderer* m_handle = nullptr;
    okrt_geometry* m_geometry = nullptr;
    okrt_render_data m_renderData{};
    okrt_options m_options{};
};

} // namespace okrt

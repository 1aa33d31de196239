/* okrt_internal.h — declarations shared between the host builders and the device side. */
#ifndef OKRT_INTERNAL_H
#define OKRT_INTERNAL_H
#include <cstdint>
#include <string>
#include <vector>
#include "okrt.h"

static_assert(sizeof(okrt_triangle) == 36, "Triangle");
static_assert(sizeof(okrt_vertex_info) == 44, "VertexInfo");
static_assert(sizeof(okrt_triangle_info) == 132, "TriangleInfo");
static_assert(sizeof(okrt_aabb) == 24, "AABB");
static_assert(sizeof(okrt_bvh_node) == 36, "BvhNode");
static_assert(sizeof(okrt_octree_node) == 48, "OctTreeNode");
static_assert(sizeof(okrt_material) == 68, "Material");
static_assert(sizeof(okrt_mesh_instance) == 232, "Mesh");
static_assert(sizeof(okrt_sphere) == 84, "Sphere");
static_assert(sizeof(okrt_directional_light) == 32, "DirectionalLight");
static_assert(sizeof(okrt_point_light) == 32, "PointLight");
static_assert(sizeof(okrt_spot_light) == 48, "SpotLight");
static_assert(sizeof(okrt_render_data) == 240, "RenderData");
static_assert(sizeof(okrt_aov) == 56, "okrt_aov");

/* result of loadMeshAndBvhData: the three arrays the reference uploads + per-mesh descriptors (RayTracer.h:15-21) */
struct okrt_geometry {
    std::vector<okrt_triangle> triPos;
    std::vector<okrt_triangle_info> triInfo;
    std::vector<okrt_bvh_node> nodes;
    std::vector<okrt_mesh_desc> descs;
    std::vector<okrt_aabb> bounds;
};

/* screen-space line list of a debug overlay (row f4): 4 floats per segment, the draw it belongs to (1-based; later draws
 * win a pixel, like consecutive Draw calls), one RGBA8 colour per draw */
struct okrt_overlay {
    std::vector<float> segments;
    std::vector<uint32_t> keys;
    std::vector<uint32_t> colours;
};

namespace okrt {
void setError(const std::string& msg);
int buildBvh(const okrt_triangle* tris, uint32_t numTriangles, const okrt_aabb& rootBounds, uint32_t maxDepth,
             uint32_t maxLeaf, std::vector<okrt_bvh_node>& nodes, std::vector<uint32_t>& triOrder);
}
#endif

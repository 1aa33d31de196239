/*
 * okrt.h — C ABI of the B200-native path-tracing core.
 *
 * This is the drop-in boundary for ONE hot path of Olowerr/GPU-Raytracer: the
 * compute-shader path tracer (resources/shaders/RaytracerCS.hlsl) and the host
 * code that feeds it (source/Graphics/RayTracer.cpp, BvhBuilder.cpp).  The
 * reference has no FFI layer; its boundary is the C++ class `RayTracer`
 * (source/Graphics/RayTracer.h:71-193) on top and the D3D11 binding table
 * (resources/shaders/ShaderResourceRegisters.h:9-29, bound at
 * source/Graphics/RayTracer.cpp:230-253) underneath.  Every entry point below
 * names the reference interface it replaces.
 *
 * Plain C: PODs, pointers and sizes only.  All wire structs are 4-byte aligned
 * with no padding and are byte-identical to the reference's GPU structs
 * (resources/shaders/GPU-Structs.hlsli) and their CPU mirrors.
 *
 * Ownership: the caller owns every host array; `okrt_set_*` copies to the
 * device before returning (reference: CreateBuffer with initial data /
 * Map(WRITE_DISCARD)+memcpy, source/DirectX/DX11.cpp:104-156).
 * Errors: every call returns an int status (0 = OKRT_OK); the library never
 * aborts (reference: OKAY_ASSERT + bool, source/Utilities.h:12).
 * Threading: one host thread per renderer handle; one CUDA stream per handle
 * (internal helper streams fork from and join it inside a call);
 * `okrt_render*` is asynchronous, `okrt_sync`/`okrt_read_*` synchronise.
 */
#ifndef OKRT_H
#define OKRT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define OKRT_API __attribute__((visibility("default")))
#else
#define OKRT_API
#endif

#define OKRT_INVALID_UINT 0xFFFFFFFFu /* Okay::INVALID_UINT, source/Utilities.h:23 */

/* ------------------------------------------------------------------ status */
enum {
    OKRT_OK = 0,
    OKRT_ERR_INVALID_ARGUMENT = 1,
    OKRT_ERR_CUDA = 2,          /* a CUDA runtime call failed; see okrt_last_error() */
    OKRT_ERR_STATE = 3,         /* call order violated (e.g. render before geometry) */
    OKRT_ERR_CAPACITY = 4,      /* a fixed traversal stack would overflow for this scene */
    OKRT_ERR_NO_DEVICE = 5      /* no CUDA device: the product path has no CPU fallback */
};

/* ------------------------------------------------------------- wire structs
 * Byte layouts: SURVEY.md §8.2.  Sizes are static_assert'ed in the library. */

typedef struct okrt_triangle {          /* 36 B  GPU-Structs.hlsli:95-98, Utilities.h:103-106 */
    float position[3][3];
} okrt_triangle;

typedef struct okrt_vertex_info {       /* 44 B  GPU-Structs.hlsli:100-106, Utilities.h:108-114 */
    float normal[3];
    float uv[2];
    float tangent[3];
    float bitangent[3];
} okrt_vertex_info;

typedef struct okrt_triangle_info {     /* 132 B GPU-Structs.hlsli:108-111, Utilities.h:116-119 */
    okrt_vertex_info vertexInfo[3];
} okrt_triangle_info;

typedef struct okrt_aabb {              /* 24 B  GPU-Structs.hlsli:73-78, Utilities.h:63-70 */
    float min[3];
    float max[3];
} okrt_aabb;

typedef struct okrt_bvh_node {          /* 36 B  GPU-Structs.hlsli:159-165, RayTracer.h:23-29 */
    okrt_aabb boundingBox;
    uint32_t triStart;                  /* leaf: [triStart, triEnd) in the leaf-ordered arrays */
    uint32_t triEnd;
    uint32_t firstChildIdx;             /* OKRT_INVALID_UINT <=> leaf; children at idx, idx+1 */
} okrt_bvh_node;

typedef struct okrt_octree_node {       /* 48 B  GPU-Structs.hlsli:167-179, RayTracer.h:57-69 */
    okrt_aabb boundingBox;
    uint32_t meshesStartIdx;
    uint32_t meshesEndIdx;
    uint32_t spheresStartIdx;           /* unused by the shader */
    uint32_t spheresEndIdx;             /* unused by the shader */
    uint32_t firstChildIdx;
    uint32_t numChildren;
} okrt_octree_node;

typedef struct okrt_material {          /* 68 B  GPU-Structs.hlsli:36-71, Components.h:42-56 */
    float    albedo[3];
    uint32_t albedoTextureIdx;
    float    roughness;
    uint32_t roughnessTextureIdx;
    float    metallic;
    uint32_t metallicTextureIdx;
    float    specular;
    uint32_t specularTextureIdx;
    uint32_t normalMapIdx;
    float    emissionColour[3];
    float    emissionPower;
    float    transparency;
    float    indexOfRefraction;
} okrt_material;

typedef struct okrt_mesh_instance {     /* 232 B GPU-Structs.hlsli:113-125, Components.h:117-128 */
    float    transformMatrix[16];       /* 16 floats in memory = row-major T (host uploads glm::transpose(T),
                                           RayTracer.cpp:747-749); world = T * local */
    float    inverseTransformMatrix[16];/* row-major inverse(T) */
    uint32_t triStart;
    uint32_t triEnd;
    okrt_aabb boundingBox;              /* mesh-local */
    okrt_material material;
    uint32_t bvhNodeStartIdx;
} okrt_mesh_instance;

typedef struct okrt_sphere {            /* 84 B  GPU-Structs.hlsli:86-93, RayTracer.cpp:88 */
    float position[3];
    okrt_material material;
    float radius;
} okrt_sphere;

typedef struct okrt_directional_light { /* 32 B  GPU-Structs.hlsli:127-135 */
    float colour[3];
    float intensity;
    float effectiveAngle;               /* cos(angle) — RayTracer.cpp:448 */
    float direction[3];
} okrt_directional_light;

typedef struct okrt_point_light {       /* 32 B  GPU-Structs.hlsli:137-145 */
    float colour[3];
    float intensity;
    float radius;
    float position[3];
} okrt_point_light;

typedef struct okrt_spot_light {        /* 48 B  GPU-Structs.hlsli:147-157 */
    float colour[3];
    float intensity;
    float radius;
    float maxAngle;                     /* cos(angle) — RayTracer.cpp:477 */
    float position[3];
    float direction[3];
} okrt_spot_light;

typedef struct okrt_render_data {       /* 240 B RaytracerCS.hlsl:19-47, RayTracer.h:129-157 */
    uint32_t accumulationEnabled;
    uint32_t numAccumulationFrames;
    uint32_t numSpheres;
    uint32_t numMeshes;
    uint32_t numDirLights;
    uint32_t numPointLights;
    uint32_t numSpotLights;
    uint32_t debugMode;                 /* 0 none, 1 bbCheckCount, 2 triCheckCount (RayTracer.h:74-79) */
    uint32_t textureDims[2];
    float    viewPlaneDims[2];
    float    cameraInverseProjectionMatrix[16]; /* row-major inverse(P) (transposed upload, RayTracer.cpp:295) */
    float    cameraInverseViewMatrix[16];       /* row-major inverse(V) (RayTracer.cpp:306) */
    float    cameraPosition[3];
    float    cameraNearZ;
    float    cameraUpDir[3];
    float    dofStrength;
    float    cameraRightDir[3];
    float    dofDistance;
    uint32_t debugMaxCount;
    float    pad0[3];
} okrt_render_data;

/* ------------------------------------------------ host-side scene inputs
 * These replace the EnTT registry (source/Scene) with plain arrays. */

typedef struct okrt_mesh_desc {         /* RayTracer.h:15-21 */
    uint32_t startIdx;
    uint32_t endIdx;
    uint32_t bvhTreeStartIdx;
    uint32_t numBvhNodes;
} okrt_mesh_desc;

typedef struct okrt_mesh_view {         /* one Mesh asset: Mesh.h:44-46 */
    const okrt_triangle*      trianglesPos;
    const okrt_triangle_info* trianglesInfo;
    uint32_t                  numTriangles;
    okrt_aabb                 boundingBox;   /* Importer.cpp:52-53 */
} okrt_mesh_view;

typedef struct okrt_transform {         /* Components.h:6-24 (rotation = Euler degrees) */
    float position[3];
    float rotation[3];
    float scale[3];
} okrt_transform;

typedef struct okrt_mesh_entity {       /* Transform + MeshComponent, Components.h:108-112 */
    okrt_transform transform;
    okrt_material  material;
    uint32_t       meshID;
} okrt_mesh_entity;

typedef struct okrt_sphere_entity {     /* Transform + Sphere, Components.h:102-106 */
    okrt_transform transform;
    okrt_material  material;
    float          radius;
} okrt_sphere_entity;

typedef struct okrt_camera {            /* Transform + Camera, Components.h:130-139 */
    float position[3];
    float rotation[3];                  /* Euler degrees */
    float fov;                          /* degrees */
    float nearZ;
    float farZ;
} okrt_camera;

/* ------------------------------------------------------------- new fields
 * Parameters the reference fixes at compile time or does not have. */

enum { OKRT_TRAVERSAL_FAST = 0,   /* near-first traversal of the re-packed child-pair BVH with
                                     exact-arithmetic commit; falls back to reference order on ties */
       OKRT_TRAVERSAL_EXACT = 1,  /* reference visit order, RaytracerCS.hlsl:244-334 (debug counters exact) */
       OKRT_TRAVERSAL_WIDE = 2 }; /* near-first traversal of 8-wide nodes with quantised child boxes (wavefront pipeline only);
                                    * same results; measured slower than OKRT_TRAVERSAL_FAST on B200, kept as the
                                    * memory-lean alternative: 3 sectors per 8 children instead of 2 per 2 */

enum { OKRT_PIPELINE_WAVEFRONT = 0,  /* generate / trace / shade kernels with ray queues in HBM (default) */
       OKRT_PIPELINE_MEGAKERNEL = 1 }; /* one persistent kernel with per-lane path regeneration */

enum { OKRT_SHARD_NONE = 0,
       OKRT_SHARD_TILE = 1 };     /* this renderer traces only its interleaved row bands */

typedef struct okrt_options {
    uint32_t maxBounces;          /* NUM_BOUNCES, RaytracerCS.hlsl:6 (loop is i <= maxBounces) */
    uint32_t traversal;           /* OKRT_TRAVERSAL_* */
    uint32_t coverAllPixels;      /* 0: reference coverage Dispatch(W/16,H/9) (RayTracer.cpp:256), uncovered
                                     pixels keep the clear colour; 1: trace every pixel */
    uint32_t writeAov;            /* 1: record okrt_aov per pixel after the first findClosestHit */
    uint32_t shardMode;           /* OKRT_SHARD_* */
    uint32_t shardRank;
    uint32_t shardCount;
    uint32_t shardBandRows;       /* rows per interleaved band (tile mode); 0 = default */
    uint32_t frameOffset;         /* sample sharding: seed frame of local frame k (1-based) is */
    uint32_t frameStride;         /*   frameOffset + (k-1)*frameStride + 1; defaults 0 / 1 */
    uint32_t pipeline;            /* OKRT_PIPELINE_*: same results, different kernel organisation */
    uint32_t batchesInFlight;     /* wavefront pipeline: how many batches of frames may be on the device at once, each on its
                                     own stream with its own path state (1..4; 0 = default 3).  Frames still reach the
                                     accumulation image in frame order: same images for every value */
    uint32_t reserved[4];
} okrt_options;

typedef struct okrt_aov {         /* parity tap, one per pixel (new) */
    uint32_t hitKind;             /* 0 miss, 1 sphere, 2 mesh triangle */
    uint32_t hitIdx;              /* sphere index or mesh-instance index */
    uint32_t triIdx;              /* index into the leaf-ordered triangle arrays */
    float    baryU;
    float    baryV;
    float    distance;            /* payload.distance (world space) */
    uint32_t bbCheckCount;        /* totals over the whole path, exact in OKRT_TRAVERSAL_EXACT */
    uint32_t triCheckCount;
    uint32_t numSegments;         /* findClosestHit calls on this path */
    uint32_t flags;               /* bit0: first segment resolved by the reference-order fallback */
    float    radiance[3];         /* this frame's `light` before accumulation */
    uint32_t reserved;
} okrt_aov;

typedef struct okrt_stats {
    uint64_t raySegments;         /* findClosestHit calls since the last reset_stats */
    uint64_t fallbackSegments;    /* of those, resolved by the reference-order fallback */
    uint64_t pixelsTraced;
    uint64_t kernelLaunches;      /* CUDA kernels launched by this handle since reset_stats */
    float    lastRenderMs;        /* device time of the last okrt_render/_frames call (CUDA events) */
    float    traceKernelMs;       /* of that call: summed device time of the traversal kernel launches (CUDA events around each) */
    float    traceKernelLaunches; /* how many launches traceKernelMs covers */
    float    reserved[1];
} okrt_stats;

typedef struct okrt_renderer okrt_renderer;   /* RayTracer instance */
typedef struct okrt_geometry okrt_geometry;   /* result of loadMeshAndBvhData */
typedef struct okrt_octree   okrt_octree;     /* result of createOctTree/loadOctTree */

/* -------------------------------------------------------------- utilities */
OKRT_API const char* okrt_version(void);
OKRT_API const char* okrt_last_error(void);      /* thread-local message of the last failing call */
OKRT_API int  okrt_device_count(void);           /* 0 when no CUDA device is visible */
OKRT_API void okrt_material_default(okrt_material* m);   /* Components.h:42-56 defaults */
OKRT_API void okrt_options_default(okrt_options* o);     /* maxBounces 1 (HEAD), FAST, reference coverage */
OKRT_API void okrt_render_data_default(okrt_render_data* rd); /* RayTracer.h:131-156 defaults */

/* ------------------------------------------------ host builders (CPU, C++)
 * a10: RayTracer::loadMeshAndBvhData (RayTracer.cpp:120-217) + BvhBuilder
 * (BvhBuilder.cpp:12-230).  Produces the identical node array, node order and
 * leaf-ordered triangle arrays as the reference builder (bit-exact), using an
 * exact binned evaluation of the same 3 x 99 SAH candidates. */
OKRT_API int okrt_geometry_build(const okrt_mesh_view* meshes, uint32_t numMeshes,
                                 uint32_t maxDepth, uint32_t maxLeafTriangles,
                                 okrt_geometry** out);
OKRT_API void okrt_geometry_free(okrt_geometry* g);
OKRT_API uint32_t okrt_geometry_num_triangles(const okrt_geometry* g);
OKRT_API uint32_t okrt_geometry_num_nodes(const okrt_geometry* g);
OKRT_API uint32_t okrt_geometry_num_meshes(const okrt_geometry* g);
OKRT_API const okrt_triangle*      okrt_geometry_triangles_pos(const okrt_geometry* g);   /* t0 */
OKRT_API const okrt_triangle_info* okrt_geometry_triangles_info(const okrt_geometry* g);  /* t1 */
OKRT_API const okrt_bvh_node*      okrt_geometry_bvh_nodes(const okrt_geometry* g);       /* t2 */
OKRT_API const okrt_mesh_desc*     okrt_geometry_mesh_descs(const okrt_geometry* g);
OKRT_API const okrt_aabb*          okrt_geometry_mesh_bounds(const okrt_geometry* g);
/* f2: the same outputs from a builder that runs on the GPU (LBVH: Morton sort, parallel radix tree, bottom-up box
 * fit, subtrees of <= maxLeafTriangles collapsed to leaves).  Replaces BvhBuilder::buildTree (BvhBuilder.cpp:38-207)
 * behind RayTracer::loadMeshAndBvhData (RayTracer.cpp:143-201) where a different-but-valid tree is acceptable, e.g.
 * the UI's rebuild button (Application.cpp:405-411).  Node format, child adjacency and leaf-ordered triangle arrays
 * are the reference's; the tree is NOT the reference's SAH tree, so images agree with the reference's only up to its
 * traversal-order tie cases.  No depth limit: okrt_render reports OKRT_ERR_CAPACITY if a tree outgrows the stacks.
 * outBuildMs (may be NULL): device-timed duration including the host<->device copies.  No CPU fallback. */
OKRT_API int okrt_geometry_build_device(const okrt_mesh_view* meshes, uint32_t numMeshes, uint32_t maxLeafTriangles,
                                        int device, okrt_geometry** out, float* outBuildMs);
/* Single-mesh form of BvhBuilder::buildTree (BvhBuilder.cpp:12-36): nodes use
 * mesh-local indices, triOrder[k] = source triangle stored at leaf slot k.
 * outNodes needs capacity 2*numTriangles+1, outTriOrder numTriangles. */
OKRT_API int okrt_build_bvh(const okrt_triangle* tris, uint32_t numTriangles, const okrt_aabb* rootBounds,
                            uint32_t maxDepth, uint32_t maxLeafTriangles,
                            okrt_bvh_node* outNodes, uint32_t nodeCapacity, uint32_t* outNumNodes,
                            uint32_t* outTriOrder);

/* a11: RayTracer::createOctTree + loadOctTree (RayTracer.cpp:523-788).
 * `meshEntities` / `sphereEntities` are given in the order the reference's
 * registry views would enumerate them. */
OKRT_API int okrt_octree_build(const okrt_geometry* g,
                               const okrt_mesh_entity* meshEntities, uint32_t numMeshEntities,
                               const okrt_sphere_entity* sphereEntities, uint32_t numSphereEntities,
                               uint32_t maxDepth, uint32_t maxLeafObjects,
                               okrt_octree** out);
OKRT_API void okrt_octree_free(okrt_octree* t);
OKRT_API uint32_t okrt_octree_num_nodes(const okrt_octree* t);
OKRT_API uint32_t okrt_octree_num_instances(const okrt_octree* t);
OKRT_API const okrt_octree_node*   okrt_octree_nodes(const okrt_octree* t);       /* t5 */
OKRT_API const okrt_mesh_instance* okrt_octree_instances(const okrt_octree* t);   /* t7 */

/* a12: Transform::calculateMatrix (Components.h:13-18) and
 * RayTracer::calculateProjectionData (RayTracer.cpp:268-308). */
OKRT_API void okrt_transform_matrix(const okrt_transform* t, float outRowMajor[16]);
OKRT_API int  okrt_calculate_projection_data(const okrt_camera* cam, uint32_t width, uint32_t height,
                                             okrt_render_data* inout);

/* a13: RayTracer::loadTextureData (RayTracer.cpp:310-342) + Okay::scaleTexture
 * (DX11.cpp:232-270) and RayTracer::loadEnvironmentMap (RayTracer.cpp:344-432). */
OKRT_API uint32_t okrt_texture_array_size(const uint32_t* widths, const uint32_t* heights, uint32_t count);
OKRT_API int okrt_scale_texture(const uint8_t* srcRgba8, uint32_t srcW, uint32_t srcH,
                                uint8_t* dstRgba8, uint32_t dstW, uint32_t dstH);
OKRT_API int okrt_environment_cross_to_faces(const uint8_t* crossRgba8, uint32_t imgW, uint32_t imgH,
                                             uint8_t* outFaces6 /* 6*(imgW/4)*(imgH/3)*4 */);

/* f1: Importer::loadObjects (Importer.cpp:77-120) + aiMeshToMeshData (:15-56) for Wavefront OBJ / MTL: one mesh
 * per (object or group, material) as assimp's OBJ importer forms them, run through the reference's post-process
 * flags Triangulate | ConvertToLeftHanded | CalcTangentSpace, de-indexed, position * scale.  assimp is not
 * available to this build (prebuilt Windows library in the reference): its behaviour is restated, parity at the
 * import boundary is unpinned (csrc/okrt_obj.cpp). */
typedef struct okrt_obj okrt_obj;
enum { OKRT_OBJ_TEX_ALBEDO = 0, OKRT_OBJ_TEX_ROUGHNESS = 1, OKRT_OBJ_TEX_METALLIC = 2, OKRT_OBJ_TEX_SPECULAR = 3,
       OKRT_OBJ_TEX_NORMAL = 4 };                /* ObjectDecriptionStr texture paths, Importer.h */
OKRT_API int okrt_obj_load(const char* path, float scale, okrt_obj** out);       /* reads <path> and the .mtl files it names */
OKRT_API int okrt_obj_load_memory(const char* objText, size_t objLen, const char* mtlText, size_t mtlLen,
                                  float scale, okrt_obj** out);
OKRT_API void okrt_obj_free(okrt_obj* a);
OKRT_API uint32_t okrt_obj_num_meshes(const okrt_obj* a);
/* view->trianglesPos / trianglesInfo stay owned by the asset; *name (optional) is the aiMesh name */
OKRT_API int okrt_obj_mesh(const okrt_obj* a, uint32_t i, okrt_mesh_view* view, const char** name);
OKRT_API const char* okrt_obj_texture_path(const okrt_obj* a, uint32_t mesh, uint32_t slot);   /* "" when the material binds none */

/* f3: Application::saveScreenshot (Application.cpp:661-674) = stbi_write_png(path, W, H, 4, data, W * 4) on the RGBA8
 * target.  Own encoder (8-bit RGBA, filter 0, fixed-Huffman deflate); the decoded image equals the input bit for bit. */
OKRT_API int okrt_write_png(const char* path, const uint8_t* rgba8, uint32_t width, uint32_t height);
/* Progressive state on disk (checkpoint / resume, offline merge of sample shards): the RGBA32F accumulation (u1) and the
 * frame count.  okrt_read_accumulation_file with rgba32f == NULL reads only the header. */
OKRT_API int okrt_write_accumulation_file(const char* path, const float* rgba32f, uint32_t width, uint32_t height,
                                          uint32_t numAccumulationFrames);
OKRT_API int okrt_read_accumulation_file(const char* path, float* rgba32f, uint32_t* width, uint32_t* height,
                                         uint32_t* numAccumulationFrames);

/* f4: the box overlay of DebugRenderer::renderBvhNodeBBs / renderOctTreeNodeBBs (DebugRenderer.cpp:264-303, 335-364,
 * DebugBBVS.hlsl:13-33, DebugBBPS.hlsl): the 12 edges of node boxes as lines over the result image, nodes in the order
 * of executeDrawMode (:447-506), colours of :407-445, no depth test.  okrt_overlay_build (host) makes the clipped
 * screen-space line list; okrt_draw_overlay (device) rasterises it into the RGBA8 result.  Rasterisation rule of this
 * library (D3D11's own rasteriser is unavailable on this box): a line owns, in its major axis, every pixel whose
 * centre lies in [start, end) and, in the minor axis, the pixel containing the interpolated coordinate. */
typedef struct okrt_overlay okrt_overlay;
enum { OKRT_OVERLAY_BVH = 0, OKRT_OVERLAY_OCTREE = 1 };                                   /* DBGRenderData.mode */
enum { OKRT_DRAW_NONE = 0, OKRT_DRAW_SINGLE = 1, OKRT_DRAW_WITH_CHILDREN = 2, OKRT_DRAW_WITH_DESCENDANTS = 3 };  /* DrawMode */
OKRT_API int okrt_overlay_build(uint32_t kind, const void* nodeArray, uint32_t numNodes, uint32_t nodeIdx, uint32_t drawMode,
                                const okrt_transform* objectTransform /* NULL: identity */, const okrt_camera* cam,
                                uint32_t width, uint32_t height, okrt_overlay** out);
OKRT_API void okrt_overlay_free(okrt_overlay* ov);
OKRT_API uint32_t okrt_overlay_num_segments(const okrt_overlay* ov);
OKRT_API uint32_t okrt_overlay_num_draws(const okrt_overlay* ov);
OKRT_API const float*    okrt_overlay_segments(const okrt_overlay* ov);   /* x0 y0 x1 y1 per segment, pixels */
OKRT_API const uint32_t* okrt_overlay_keys(const okrt_overlay* ov);       /* draw (1-based) of each segment */
OKRT_API const uint32_t* okrt_overlay_colours(const okrt_overlay* ov);    /* RGBA8 per draw */

/* Host-only report on the traversal layout okrt_render builds from the binding-table arrays (no GPU needed): lets a
 * test or a tool check stack needs and compare tree quality.  sahCost = sum over child boxes of the scene-wide tree of
 * area(box)/area(root) x (1 for an interior child, triangle count for a leaf child): the expected work of a random ray. */
typedef struct okrt_layout_info {
    uint32_t numPairNodes;          /* 64-byte child-pair nodes */
    uint32_t numFlattenedLeaves;    /* reference BVH leaves that became primitives of the scene-wide tree */
    uint32_t numMergedInstances;
    uint32_t numInstanceRecords;
    uint32_t fastStackDepth;        /* worst-case occupancy; capacity 64 */
    uint32_t exactBvhDepth;
    uint32_t exactOctDepth;
    uint32_t reserved;
    double   sahCost;
    double   buildMs;
} okrt_layout_info;
OKRT_API int okrt_layout_report(const okrt_triangle* tris, uint32_t numTris, const okrt_bvh_node* nodes, uint32_t numNodes,
                                const okrt_mesh_instance* instances, uint32_t numInstances,
                                const okrt_octree_node* octNodes, uint32_t numOctNodes, okrt_layout_info* out);

/* How the traversal layout the kernels walk right now was built.
 *   builder 1: on the device (okrt_layout_device.cu), milliseconds after a scene edit: every instance identity and the only user
 *              of its mesh — all BASELINE configs.  Its clustered tree costs rays/s against the host builder's SAH tree, so the
 *              host builder starts in the background at once (OKRT_LAYOUT_REFINE=0: not) ...
 *   builder 2: ... and its layout has replaced the clustered one (at an okrt_render* call after it finished, or in
 *              okrt_wait_layout, which waits for it).  Images are the same with either tree.
 *   builder 0: on host threads, synchronously (the complete path: instances with transforms, shared meshes,
 *              OKRT_TRAVERSAL_WIDE, OKRT_LAYOUT_HOST=1); `why` says what sent it there. */
typedef struct okrt_layout_build {
    float    wallMs;             /* host wall clock of the synchronous part of the rebuild, uploads included */
    float    deviceMs;           /* CUDA events around the device builder's kernels (0 for the host builder) */
    float    refineMs;           /* builder 2: duration of the background host build */
    uint32_t builder;
    uint32_t numPairNodes;
    uint32_t fastStackDepth;
    uint32_t rounds;             /* clustering rounds of the device builder */
    char     why[100];
} okrt_layout_build;
OKRT_API int okrt_get_layout_build(okrt_renderer* r, okrt_layout_build* out);
OKRT_API int okrt_wait_layout(okrt_renderer* r);      /* builds the layout if an edit is pending and waits for its refinement */

/* Host-only check of the near-first traversal (no GPU): runs the traversal kernels' lane functions, compiled for the host,
 * over arbitrary rays (6 floats per ray: origin, direction) on the layout okrt_render would build from the same arrays.
 * out: 12 words per ray — distance, u, v (float bits), triangle, kind (0 miss / 2 mesh), instance index, 1 if the reference
 * order has to decide, the bound left for that traversal (float bits), node steps, leaf entries, triangle tests, 0.
 * Lets tests/ compare the traversal with the oracle's findClosestHit (RaytracerCS.hlsl:218-334) where no GPU exists. */
OKRT_API int okrt_trace_rays_host(const okrt_triangle* tris, uint32_t numTris, const okrt_bvh_node* nodes, uint32_t numNodes,
                                  const okrt_mesh_instance* instances, uint32_t numInstances,
                                  const okrt_octree_node* octNodes, uint32_t numOctNodes,
                                  uint32_t traversal /* OKRT_TRAVERSAL_FAST or _WIDE */, const float* rays, uint32_t numRays, uint32_t* out);

/* ------------------------------------------------------- renderer (device) */
OKRT_API int  okrt_create(int device, uint32_t width, uint32_t height, okrt_renderer** out); /* RayTracer::initiate :60 */
OKRT_API void okrt_destroy(okrt_renderer* r);                                                 /* RayTracer::shutdown :34 */
OKRT_API int  okrt_resize(okrt_renderer* r, uint32_t width, uint32_t height);                 /* RayTracer::onResize :494 */

/* binding table (ShaderResourceRegisters.h:9-29) */
OKRT_API int okrt_set_triangles(okrt_renderer* r, const okrt_triangle* pos, const okrt_triangle_info* info, uint32_t n); /* t0,t1 RayTracer.cpp:209-210 */
OKRT_API int okrt_set_bvh_nodes(okrt_renderer* r, const okrt_bvh_node* nodes, uint32_t n);          /* t2 :211 */
OKRT_API int okrt_set_textures(okrt_renderer* r, const uint8_t* rgba8, uint32_t size, uint32_t layers); /* t3 :335 */
OKRT_API int okrt_set_environment_map(okrt_renderer* r, const uint8_t* faces6, uint32_t faceW, uint32_t faceH); /* t4 :420-429 */
OKRT_API int okrt_set_octree(okrt_renderer* r, const okrt_octree_node* nodes, uint32_t n);           /* t5 :778 */
OKRT_API int okrt_set_spheres(okrt_renderer* r, const okrt_sphere* spheres, uint32_t n);             /* t6 :88 */
OKRT_API int okrt_set_mesh_instances(okrt_renderer* r, const okrt_mesh_instance* inst, uint32_t n);  /* t7 :777 */
OKRT_API int okrt_set_directional_lights(okrt_renderer* r, const okrt_directional_light* l, uint32_t n); /* t8 :438-451 */
OKRT_API int okrt_set_point_lights(okrt_renderer* r, const okrt_point_light* l, uint32_t n);         /* t9 :453-465 */
OKRT_API int okrt_set_spot_lights(okrt_renderer* r, const okrt_spot_light* l, uint32_t n);           /* t10 :467-481 */
OKRT_API int okrt_set_render_data(okrt_renderer* r, const okrt_render_data* rd);                     /* b0 :491 */
OKRT_API int okrt_get_render_data(const okrt_renderer* r, okrt_render_data* out);
OKRT_API int okrt_set_options(okrt_renderer* r, const okrt_options* o);

/* RayTracer::render (RayTracer.cpp:219-260): numAccumulationFrames += accumulationEnabled (:489), clear the
 * result to (0.2,0.4,0.6,1) (:227-228), dispatch.  okrt_render_frames repeats it `frames` times. */
OKRT_API int okrt_render(okrt_renderer* r);
OKRT_API int okrt_render_frames(okrt_renderer* r, uint32_t frames);
OKRT_API int okrt_sync(okrt_renderer* r);

OKRT_API int okrt_toggle_accumulation(okrt_renderer* r, int enable);   /* RayTracer.h:197-201 */
OKRT_API int okrt_reset_accumulation(okrt_renderer* r);                /* RayTracer.h:203-209 */
OKRT_API uint32_t okrt_get_num_accumulation_frames(const okrt_renderer* r); /* RayTracer.h:211 */

/* u0 / u1 read-back (reference: staging copy, DX11.cpp:196-230) */
OKRT_API int okrt_read_result(okrt_renderer* r, uint8_t* outRgba8);            /* W*H*4 bytes */
/* Tile-sharded renderer: copies only the row bands this rank owns into their place in a full-size W*H*4 image (other
 * rows untouched), so N ranks fill one host image without moving N copies of it; *outBytes (may be NULL) = bytes moved.
 * Unsharded renderers behave like okrt_read_result. */
OKRT_API int okrt_read_result_owned(okrt_renderer* r, uint8_t* outRgba8, uint64_t* outBytes);
/* Non-blocking form: snapshots the result image as of the frames submitted so far (device-to-device, on the renderer's
 * stream) and starts the device-to-host copy of the rows this handle owns on a second stream; returns at once, so the caller
 * can submit the next okrt_render_frames while the copy runs (the reference's render() is fire-and-forget on the device queue,
 * RayTracer.cpp:219-260).  `slot` (0 .. OKRT_READ_SLOTS-1) names one of four snapshots; okrt_wait_read(slot) blocks until that
 * copy has landed in outRgba8, which must stay valid (and should be page-locked, okrt_host_register) until then.  A loop that
 * waits for step k-2's image after submitting step k (three slots in turn) keeps three batches on the device
 * (okrt_options::batchesInFlight). */
#define OKRT_READ_SLOTS 4
OKRT_API int okrt_read_result_async(okrt_renderer* r, uint8_t* outRgba8, uint64_t* outBytes, uint32_t slot);
OKRT_API int okrt_wait_read(okrt_renderer* r, uint32_t slot);
OKRT_API int okrt_read_accumulation(okrt_renderer* r, float* outRgba32f);
/* Tile sharding with an explicit band table: `bands` = the global band indices (ascending) this handle traces, instead of every
 * shardCount-th band; n = 0 restores the interleave.  okrt_shard_balance (host-only) assigns bands to ranks from measured band
 * costs (e.g. ray segments per band of one calibration frame) by longest-processing-time-first, deterministically, so that
 * every rank computes the same table without talking to the others. */
OKRT_API int okrt_set_shard_bands(okrt_renderer* r, const uint32_t* bands, uint32_t n);
OKRT_API int okrt_shard_balance(const double* bandCost, uint32_t numBands, uint32_t shardCount, uint32_t* ownerOfBand);
/* f4: rasterises the overlay's lines into the RGBA8 result (u0); the accumulation (u1) is untouched, the next
 * okrt_render_frames rewrites u0. */
OKRT_API int okrt_draw_overlay(okrt_renderer* r, const okrt_overlay* ov);
/* okrt_draw_overlay for callers that cannot allocate device memory (CUDA graph capture around them, serving loops, harnesses
 * that forbid cudaMalloc).  It has no counterpart in the reference.
 * Same kernels and the same pixels as okrt_draw_overlay, but the scratch (pixel keys, line list, draw colours) is one
 * file-scope device array of fixed size — 9.3 MiB held by every GPU that loads the library, until exit — so the call makes
 * no cudaMalloc / cudaFree.  Image, line list or draw count beyond the limits below: OKRT_ERR_INVALID_ARGUMENT, nothing drawn
 * (use okrt_draw_overlay).  All callers on a device share the array, which keeps the last call's contents (the call clears
 * what must start at zero): calls that overlap on one GPU — two renderers on different streams, two host threads — race;
 * serialise them.  A call with an empty overlay only does the per-device set-up (the address look-up), as a warm-up. */
#define OKRT_OVERLAY_STATIC_MAX_PIXELS   2097152u   /* 2^21 >= 1920 x 1080 */
#define OKRT_OVERLAY_STATIC_MAX_SEGMENTS 65536u
#define OKRT_OVERLAY_STATIC_MAX_DRAWS    16384u
OKRT_API int okrt_draw_overlay_static(okrt_renderer* r, const okrt_overlay* ov);
      /* W*H*4 floats */
OKRT_API int okrt_read_aov(okrt_renderer* r, okrt_aov* out);                   /* W*H entries; needs writeAov */
/* progressive-state restore / merge (new): replaces the accumulation buffer and frame counter */
OKRT_API int okrt_write_accumulation(okrt_renderer* r, const float* rgba32f, uint32_t numAccumulationFrames);
/* recompute u0 = saturate(u1 / frames) (RaytracerCS.hlsl:529) after an external reduction of u1 */
OKRT_API int okrt_resolve(okrt_renderer* r, uint32_t totalFrames);

/* Sample-sharded mode: every rank has rendered its own frames into u1; `allreduce` sums `count` floats in place
 * across the ranks on `cudaStream` (ncclAllReduce(ptr, ptr, count, ncclFloat, ncclSum, comm, stream) in a C++ host —
 * host/sample_sharded_nccl.cpp —, torch.distributed.all_reduce in bench.py); the library enqueues it on the renderer's stream
 * after the frames submitted so far, then u0 = saturate(u1 / totalFrames) (RaytracerCS.hlsl:529).  Does not block.  The
 * callback returns 0 on success.  NCCL itself stays out of the library's link line: the caller owns the communicator. */
typedef int (*okrt_allreduce_fn)(void* deviceFloats, size_t count, void* cudaStream, void* user);
OKRT_API int okrt_reduce_accumulation(okrt_renderer* r, okrt_allreduce_fn allreduce, void* user, uint32_t totalFrames);

/* Unit rates of this device, measured by two small kernels (about 50 ms): what the traversal kernel's roofline is quoted
 * against (bench.py).  sectorsPerSecond: 32-byte sectors per second that divergent 256-bit loads (every lane its own sector, an
 * L2-resident 16 MB array, the traversal kernel's access shape) deliver over the whole device; warpInstructionsPerSecond: issue
 * rate of a dependent-free FP32 instruction stream with 8 warps per scheduler. */
typedef struct okrt_unit_peaks {
    double sectorsPerSecond;       /* 16 MB window: every load misses the L1 and hits the L2 */
    double sectorsPerSecondL1;     /* 16 KB window per block: every load hits the L1 — the LSU data stage's own rate */
    double warpInstructionsPerSecond;
    double smClockMHz;             /* (cycles of the issue kernel / its duration): the SM clock during the measurement */
    uint32_t numSms;
    uint32_t reserved;
} okrt_unit_peaks;
OKRT_API int okrt_measure_unit_peaks(int device, okrt_unit_peaks* out);

/* device pointers for peer/NCCL exchange (sample-sharded mode reduces u1 in place) */
/* Page-locks a caller-owned host buffer so that okrt_read_* / okrt_write_accumulation copy at full PCIe rate and
 * without a staging copy (the reference maps a staging texture, DX11.cpp:196-230). */
OKRT_API int okrt_host_register(void* ptr, size_t bytes);
OKRT_API int okrt_host_unregister(void* ptr);
OKRT_API void* okrt_accumulation_device_ptr(okrt_renderer* r);
OKRT_API void* okrt_result_device_ptr(okrt_renderer* r);
OKRT_API void* okrt_cuda_stream(okrt_renderer* r);

/* sharding arithmetic (host-only, no device needed): which rank traces image row y in tile mode, and which
 * global frame's seed the k-th (1-based) local frame of a sample-sharded renderer draws */
OKRT_API uint32_t okrt_shard_owner_of_row(uint32_t y, uint32_t bandRows, uint32_t shardCount);
OKRT_API uint32_t okrt_shard_seed_frame(const okrt_options* o, uint32_t localFrame, uint32_t accumulationEnabled);

OKRT_API int okrt_get_stats(okrt_renderer* r, okrt_stats* out);   /* synchronises */
/* step counters of the traversal kernel since okrt_reset_stats (8 x uint64 per bounce: rays, node steps, leaf entries, triangle
 * tests, node-phase iterations, triangle-phase iterations, lanes busy in either); OKRT_ERR_STATE unless built with -DOKRT_TRACE_STATS */
OKRT_API int okrt_get_trace_stats(okrt_renderer* r, uint64_t* out, uint32_t bounces);
OKRT_API int okrt_reset_stats(okrt_renderer* r);

#ifdef __cplusplus
}
#endif
#endif /* OKRT_H */

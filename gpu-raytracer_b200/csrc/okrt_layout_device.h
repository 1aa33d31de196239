/* okrt_layout_device.h — device-side builder of the traversal layout (okrt_layout_device.cu). */
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <string>
#include <vector>
#include "okrt.h"
#include "okrt_accel.h"

namespace okrt {

struct DeviceLayout {
    float4* nodes = nullptr;            /* device, 4 float4 per child-pair node; owned by the caller after OKRT_OK */
    float4* tris = nullptr;             /* device, kTriQuads float4 per triangle */
    uint32_t numPairNodes = 0;
    std::vector<Float4> instances;      /* kInstQuads per instance record (host; a few hundred records) */
    std::vector<uint32_t> octParent;
    uint32_t topRef = kRefNone, numMerged = 0, numFlattenedLeaves = 0;
    uint32_t maxStackDepth = 0, exactBvhDepth = 0, exactOctDepth = 0;
    uint32_t plocRounds = 0;
    bool allRigid = true;
    float deviceMs = 0.f;               /* CUDA events around the device part */
};

/* OKRT_OK: `out` holds the layout.  OKRT_ERR_STATE: the scene needs the host builder (`why` says what was met).
 * dTris / dNodes: the reference-format arrays on the device (t0, t2); hNodes etc.: the host copies okrt_set_* keeps. */
int buildFastLayoutDevice(const okrt_triangle* dTris, uint32_t numTris, const okrt_bvh_node* dNodes, const okrt_bvh_node* hNodes, uint32_t numNodes,
                          const okrt_mesh_instance* hInstances, uint32_t numInstances, const okrt_octree_node* hOct, uint32_t numOctNodes,
                          cudaStream_t stream, DeviceLayout& out, std::string& why);

} // namespace okrt

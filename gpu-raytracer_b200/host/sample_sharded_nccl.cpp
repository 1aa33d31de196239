/* Sample-sharded progressive rendering from a C++ host with NCCL directly (BASELINE configs[4]): one process per
 * GPU; rank r renders global frames r+1, r+1+N, ... into its own accumulation image; ONE ncclAllReduce sums the images in
 * place over NVLink (okrt_reduce_accumulation hands the device pointer and the renderer's stream to the callback), then every
 * rank holds saturate(sum / totalFrames).  The library itself does not link NCCL — the host owns the communicator.
 *
 *   g++ -std=c++17 sample_sharded_nccl.cpp -I../../include -I. -I/usr/local/cuda/include -L.. -lokrt -lnccl -L/usr/local/cuda/lib64 -lcudart
 *   RANK=r WORLD_SIZE=N ./a.out /tmp/okrt_nccl_id [frames-per-rank]       (rank 0 writes the unique id to that file)
 */
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <thread>

#include <cuda_runtime.h>
#include <nccl.h>

#include "RayTracer.hpp"

struct Reduce { ncclComm_t comm; };
static int allreduce(void* deviceFloats, size_t count, void* cudaStream, void* user)
{
    const Reduce* r = static_cast<const Reduce*>(user);
    return ncclAllReduce(deviceFloats, deviceFloats, count, ncclFloat, ncclSum, r->comm, static_cast<cudaStream_t>(cudaStream)) == ncclSuccess ? 0 : 1;
}

int main(int argc, char** argv)
{
    const int rank = std::getenv("RANK") ? std::atoi(std::getenv("RANK")) : 0;
    const int world = std::getenv("WORLD_SIZE") ? std::atoi(std::getenv("WORLD_SIZE")) : 1;
    const char* idPath = argc > 1 ? argv[1] : "/tmp/okrt_nccl_id";
    const uint32_t frames = argc > 2 ? (uint32_t)std::atoi(argv[2]) : 8u;
    if (okrt_device_count() <= rank) { std::printf("no CUDA device for rank %d (there is no CPU path)\n", rank); return 3; }
    cudaSetDevice(rank);
    ncclUniqueId id;
    if (rank == 0) {
        ncclGetUniqueId(&id);
        std::FILE* f = std::fopen(idPath, "wb"); std::fwrite(&id, sizeof(id), 1, f); std::fclose(f);
    } else {
        for (int tries = 0;; tries++) {
            std::FILE* f = std::fopen(idPath, "rb");
            if (f) { const size_t n = std::fread(&id, sizeof(id), 1, f); std::fclose(f); if (n == 1) break; }
            if (tries > 600) { std::printf("rank %d: no unique id at %s\n", rank, idPath); return 1; }
            std::this_thread::sleep_for(std::chrono::milliseconds(100));
        }
    }
    Reduce red;
    if (ncclCommInitRank(&red.comm, world, id, rank) != ncclSuccess) { std::printf("ncclCommInitRank failed\n"); return 1; }

    /* a lit quad in front of the camera */
    okrt::ResourceManager rm;
    okrt::Mesh quad;
    const float p[2][3][3] = {{{-5, 0, 5}, {5, 0, 5}, {5, 10, 5}}, {{-5, 0, 5}, {5, 10, 5}, {-5, 10, 5}}};
    for (int t = 0; t < 2; t++) {
        okrt_triangle tri; std::memcpy(tri.position, p[t], sizeof(tri.position));
        okrt_triangle_info info; std::memset(&info, 0, sizeof(info));
        for (int v = 0; v < 3; v++) info.vertexInfo[v].normal[2] = -1.f;
        quad.trianglesPos.push_back(tri); quad.trianglesInfo.push_back(info);
    }
    quad.boundingBox = okrt_aabb{{-5, 0, 5}, {5, 10, 5}};
    rm.addMesh(quad);
    okrt::Scene scene;
    okrt_mesh_entity e; std::memset(&e, 0, sizeof(e));
    e.transform.scale[0] = e.transform.scale[1] = e.transform.scale[2] = 1.f;
    okrt_material_default(&e.material);
    e.material.albedo[0] = 0.8f; e.material.emissionPower = 1.f;
    scene.meshEntities.push_back(e);
    scene.camera.position[1] = 5.f;
    try {
        okrt::RayTracer rt;
        rt.initiate(640, 360, rm, nullptr, 0, 0, rank);
        rt.options().maxBounces = 4;
        rt.options().frameOffset = (uint32_t)rank;         /* local frame k draws the seed of global frame rank + (k-1) N + 1 */
        rt.options().frameStride = (uint32_t)world;
        rt.applyOptions();
        rt.setScene(scene);
        rt.createOctTree(scene, 2, 50);
        rt.renderFrames(frames);
        rt.reduceAccumulation(allreduce, &red, frames * (uint32_t)world);
        rt.sync();
        std::vector<uint8_t> img;
        rt.readResult(img);
        std::printf("rank %d of %d: %u frames in the reduced image, centre pixel %u %u %u\n", rank, world, rt.getNumAccumulationFrames(),
                    img[(180 * 640 + 320) * 4], img[(180 * 640 + 320) * 4 + 1], img[(180 * 640 + 320) * 4 + 2]);
    } catch (const okrt::Error& err) {
        std::printf("okrt error %d: %s\n", err.code, err.what());
        return 1;
    }
    ncclCommDestroy(red.comm);
    return 0;
}

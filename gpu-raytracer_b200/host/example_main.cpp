/* Headless example / link test of the C++ mirror: one quad, one frame.  Exits 0 on a GPU box, prints the
 * library's error and exits 3 where no CUDA device exists (there is no CPU path). */
#include <cstdio>
#include "RayTracer.hpp"

int main()
{
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
        rt.initiate(160, 90, rm);
        rt.setScene(scene);
        rt.createOctTree(scene, 2, 50);
        rt.render();
        std::vector<uint8_t> img;
        rt.readResult(img);
        std::printf("frames %u, centre pixel %u %u %u\n", rt.getNumAccumulationFrames(), img[(45 * 160 + 80) * 4], img[(45 * 160 + 80) * 4 + 1], img[(45 * 160 + 80) * 4 + 2]);
    } catch (const okrt::Error& err) {
        std::printf("okrt error %d: %s\n", err.code, err.what());
        return err.code == OKRT_ERR_NO_DEVICE ? 3 : 1;
    }
    return 0;
}

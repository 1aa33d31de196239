/*
 * ref_shader_driver.cpp — TEST INFRASTRUCTURE.  Runs the reference's OWN compute shader text as C++ on the host.
 *
 * oracle/Makefile (`ref`) concatenates resources/shaders/GPU-Structs.hlsli, GPU-Utilities.hlsli and RaytracerCS.hlsl from
 * /root/reference where they lie into oracle/_ref/shader_extract.inc (deleted after the compile; nothing is copied into the
 * repository), with these mechanical edits only (sed): #include lines dropped, `: register(...)`, `: SV_DispatchThreadID` and
 * the `[numthreads(16, 9, 1)]` line dropped, `cbuffer X` -> `inline namespace`, `inout T x` -> `T& x`, and the compile-time
 * constant NUM_BOUNCES (1 at HEAD) bound to a variable so that the BASELINE configs' bounce counts can be run.  Two places
 * where HLSL and C++ give the same text different meanings are edited as well: the switch in findClosestHit jumps over initialised
 * locals, which HLSL allows (each case body gets braces), and HLSL's `||` evaluates both operands where C++ stops at the first
 * true one — the only operand with a side effect in the whole text is the random draw in findTransparencyBounce
 * (`cannot_refract || reflectance(...) > randomFloat(seed)`, RaytracerCS.hlsl:127), rewritten with `|` [Q7].  Everything else
 * the text needs is oracle/ref_stubs/hlsl_shim.h, whose header lists the conventions it fixes.
 *
 * ref_shader_render runs main() for every thread of Dispatch(W/16, H/9, 1) x numthreads(16, 9, 1) (RayTracer.cpp:256) over the
 * caller's accumulation image — exactly one RayTracer::render() — and is compared word for word with oracle_render
 * (tests/test_cpu_ref_shader.py) and with the CUDA kernels (tests/test_gpu_ref_shader.py).
 */
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#include "oracle_math.h"

#ifdef UINT_MAX
#undef UINT_MAX
#endif
#ifdef FLT_MAX
#undef FLT_MAX
#endif
#ifdef PI
#undef PI
#endif

namespace hlslref {
#include "hlsl_shim.h"
static uint hlslref_num_bounces = 1u;
#include "shader_extract.inc"
static_assert(sizeof(Material) == 68 && sizeof(Sphere) == 84 && sizeof(Triangle) == 36 && sizeof(TriangleInfo) == 132, "wire structs");
static_assert(sizeof(Mesh) == 232 && sizeof(BvhNode) == 36 && sizeof(OctTreeNode) == 48 && sizeof(RenderData) == 240, "wire structs");
static_assert(sizeof(DirectionalLight) == 32 && sizeof(PointLight) == 32 && sizeof(SpotLight) == 48, "wire structs");
} // namespace hlslref

/* the binding table as oracle/binding.py builds it (struct OracleScene of okrt_oracle.cpp) */
struct RefShaderScene {
    const void *triPos, *triInfo, *bvhNodes;
    const uint8_t *textures, *envFaces;
    const void *octNodes, *spheres, *meshes, *dirLights, *pointLights, *spotLights;
    uint32_t numTris, numBvhNodes, texSize, texLayers, envW, envH, numOctNodes;
    uint32_t numSpheres, numMeshes, numDirLights, numPointLights, numSpotLights;
};

extern "C" __attribute__((visibility("default")))
const char* ref_shader_description(void)
{
    return "the reference's RaytracerCS.hlsl + GPU-Utilities.hlsli + GPU-Structs.hlsli compiled as C++ (oracle/ref_stubs/hlsl_shim.h)";
}

/* One RayTracer::render(): rd is the 240-byte cbuffer as uploaded (numAccumulationFrames already advanced, RayTracer.cpp:489).
 * accum: W*H float4 (u1, read and written); resultF: W*H float4 as main() stores it; result8 (may be NULL): the same as UNORM8.
 * coverAll = 0 runs the reference's thread grid (pixels beyond 16*(W/16) x 9*(H/9) are not touched). */
extern "C" __attribute__((visibility("default")))
uint32_t ref_shader_render_rows(const RefShaderScene* s, const void* rd, uint32_t maxBounces, uint32_t coverAll,
                                float* accum, float* resultF, uint8_t* result8, const uint32_t* rowList, uint32_t numRows, int threads);

extern "C" __attribute__((visibility("default")))
uint32_t ref_shader_render(const RefShaderScene* s, const void* rd, uint32_t maxBounces, uint32_t coverAll,
                           float* accum, float* resultF, uint8_t* result8, int threads)
{
    return ref_shader_render_rows(s, rd, maxBounces, coverAll, accum, resultF, result8, nullptr, 0u, threads);
}

/* the same for the thread rows in rowList only (NULL: every row): a bounded sample of a frame for timing */
extern "C" __attribute__((visibility("default")))
uint32_t ref_shader_render_rows(const RefShaderScene* s, const void* rd, uint32_t maxBounces, uint32_t coverAll,
                                float* accum, float* resultF, uint8_t* result8, const uint32_t* rowList, uint32_t numRows, int threads)
{
    using namespace hlslref;
    std::memcpy(&renderData, rd, sizeof(RenderData));
    hlslref_num_bounces = maxBounces;
    trianglePosData.data = (const Triangle*)s->triPos; trianglePosData.count = s->numTris;
    triangleInfoData.data = (const TriangleInfo*)s->triInfo; triangleInfoData.count = s->numTris;
    bvhNodes.data = (const BvhNode*)s->bvhNodes; bvhNodes.count = s->numBvhNodes;
    textures.data = s->textures; textures.size = s->texSize; textures.layers = s->texLayers;
    environmentMap.faces = s->envFaces; environmentMap.w = s->envW; environmentMap.h = s->envH;
    octTreeNodes.data = (const OctTreeNode*)s->octNodes; octTreeNodes.count = s->numOctNodes;
    sphereData.data = (const Sphere*)s->spheres; sphereData.count = s->numSpheres;
    meshData.data = (const Mesh*)s->meshes; meshData.count = s->numMeshes;
    directionalLights.data = (const DirectionalLight*)s->dirLights; directionalLights.count = s->numDirLights;
    pointLights.data = (const PointLight*)s->pointLights; pointLights.count = s->numPointLights;
    spotLights.data = (const SpotLight*)s->spotLights; spotLights.count = s->numSpotLights;
    const uint32_t W = renderData.textureDims.x, H = renderData.textureDims.y;
    accumulationBuffer.data = (float4*)accum; accumulationBuffer.width = W;
    resultBuffer.data = (float4*)resultF; resultBuffer.width = W;
    const uint32_t coveredW = coverAll ? W : 16u * (W / 16u), coveredH = coverAll ? H : 9u * (H / 9u);
    const unsigned n = threads > 0 ? (unsigned)threads : std::max(1u, std::thread::hardware_concurrency());
    std::atomic<uint32_t> nextRow(0);
    std::vector<std::thread> pool;
    for (unsigned t = 0; t < n; t++)
        pool.emplace_back([&]() {
            for (;;) {
                const uint32_t k = nextRow.fetch_add(1);
                if (k >= (rowList ? numRows : coveredH)) break;
                const uint32_t y = rowList ? rowList[k] : k;
                if (y >= coveredH) continue;
                for (uint32_t x = 0; x < coveredW; x++) {
                    uint3 id; id.x = x; id.y = y; id.z = 0u;
                    hlslref::main(id);
                }
            }
        });
    for (auto& th : pool) th.join();
    if (result8)
        for (uint32_t k = 0; k < (rowList ? numRows : coveredH); k++)
            for (uint32_t x = 0, y = rowList ? rowList[k] : k; x < coveredW && y < coveredH; x++)
                for (int k = 0; k < 4; k++) {
                    const float v = resultF[((size_t)y * W + x) * 4 + k];
                    /* D3D11 float -> UNORM8 [Q14]: saturate, scale by 255, add 0.5, truncate; NaN -> 0 */
                    result8[((size_t)y * W + x) * 4 + k] = (uint8_t)(std::fmin(std::fmax(v, 0.0f), 1.0f) * 255.0f + 0.5f);
                }
    return coveredW * (rowList ? numRows : coveredH);
}

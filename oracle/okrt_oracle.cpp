/*
 * okrt_oracle.cpp — TEST INFRASTRUCTURE.  CPU oracle for the path-tracing hot path.
 *
 * A scalar, single-source-of-truth restatement of the reference's device code
 *   resources/shaders/RaytracerCS.hlsl:77-537
 *   resources/shaders/GPU-Utilities.hlsli:1-198
 *   resources/shaders/GPU-Structs.hlsli:1-194
 * and of the D3D-free parts of its host feed
 *   source/Graphics/RayTracer.cpp:120-217 (flatten), :268-308 (camera), :503-788 (oct-tree)
 *   source/Graphics/BvhBuilder.cpp:12-230 (brute-force SAH builder)
 *   source/Scene/Components.h:13-23 (Transform math; glm 0.9.9.9 formulas restated)
 *   source/DirectX/DX11.cpp:232-270 + Scale shaders (bilinear texture rescale)
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library, and only as the checker or the timed CPU
 * baseline.  The product (gpu-raytracer_b200/) never links, imports or executes it.
 *
 * PARITY PIN STATUS
 *   - BVH builder, Transform/camera matrices: PINNED against the reference's own
 *     code (BvhBuilder.cpp compiled unmodified + its vendored glm) through
 *     oracle/_ref/libokrt_ref.so — see tests/test_oracle_vs_ref.py and the
 *     committed fixtures under tests/golden/.
 *   - Oct-tree (createOctTree / loadOctTree / refitOctTreeNode): PINNED against the
 *     reference's own function text compiled into oracle/_ref/libokrt_ref_octree.so.
 *   - Device path (HLSL): PINNED against the reference's own shader text compiled as
 *     C++ (oracle/ref_shader_driver.cpp + oracle/ref_stubs/hlsl_shim.h ->
 *     oracle/_ref/libokrt_ref_shader.so): oracle_render equals it word for word,
 *     tests/test_cpu_ref_shader.py, fixtures tests/golden/ref_shader_*.npz.  What
 *     nothing outside this repository pins: the conventions hlsl_shim.h lists
 *     (intrinsic math bits, texture filtering, mul's summation order) — D3D11 leaves
 *     them to the driver or hardware, and cs_5_0 cannot run without D3D11
 *     (unavailable on this box).  The restatement follows the shader line by line;
 *     each function cites the lines it restates.
 *
 * Semantics that differ between HLSL and C++ are restated deliberately; see the
 * Q-list in SURVEY.md §8.0 (Q1..Q15), referenced inline as [Qn].
 *
 * Build: g++ -O2 -ffp-contract=off -pthread -shared -fPIC (see oracle/Makefile).
 */
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>
#include <atomic>
#include <thread>

#include "oracle_math.h"

#define ORACLE_API extern "C" __attribute__((visibility("default")))

namespace {

/* ------------------------------------------------------------------ shim */
struct float2 { float x, y; };
struct float3 { float x, y, z; };
struct float4 { float x, y, z, w; };

inline float3 f3(float x, float y, float z) { return float3{x, y, z}; }
inline float3 f3(const float* p) { return float3{p[0], p[1], p[2]}; }
inline float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline float3 operator*(float3 a, float3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline float3 operator*(float3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
inline float3 operator*(float s, float3 a) { return f3(s * a.x, s * a.y, s * a.z); }
inline float3 operator/(float3 a, float s) { return f3(a.x / s, a.y / s, a.z / s); }
inline float3 operator-(float3 a) { return f3(-a.x, -a.y, -a.z); }
inline float2 operator+(float2 a, float2 b) { return float2{a.x + b.x, a.y + b.y}; }
inline float2 operator*(float2 a, float s) { return float2{a.x * s, a.y * s}; }

/* dp3: products summed left to right */
inline float dot(float3 a, float3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline float3 cross(float3 a, float3 b)
{
    return f3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
inline float length(float3 a) { return std::sqrt(dot(a, a)); }
/* normalize(v) = v * rsqrt(dot(v,v)), with rsqrt taken as the IEEE 1/sqrt */
inline float3 normalize(float3 a) { float inv = 1.0f / std::sqrt(dot(a, a)); return a * inv; }
/* lerp(a,b,s) = a + s*(b-a) */
inline float lerp(float a, float b, float s) { return a + s * (b - a); }
inline float3 lerp(float3 a, float3 b, float s) { return f3(lerp(a.x, b.x, s), lerp(a.y, b.y, s), lerp(a.z, b.z, s)); }
/* reflect(i,n) = i - 2*dot(i,n)*n */
inline float3 reflect(float3 i, float3 n) { float d2 = 2.0f * dot(i, n); return i - n * d2; }
/* HLSL min/max return the non-NaN operand */
inline float hmin(float a, float b) { return std::fmin(a, b); }
inline float hmax(float a, float b) { return std::fmax(a, b); }
inline float saturate(float x) { return hmin(hmax(x, 0.0f), 1.0f); }

/* ---------------------------------------------------- wire structs (HLSL) */
const uint32_t UINT_MAX_ = ~0u;
const float FLT_MAX_ = 3.402823466e+38F;
const float PI_ = 3.14159265f;           /* GPU-Structs.hlsli:3 */
const float AIR_REFRACTION_INDEX = 1.0f;

struct Material {                         /* GPU-Structs.hlsli:36-71 */
    float albedo[3]; uint32_t albedoTex;
    float roughness; uint32_t roughnessTex;
    float metallic; uint32_t metallicTex;
    float specular; uint32_t specularTex;
    uint32_t normalMapIdx;
    float emissionColour[3];
    float emissionPower;
    float transparency;
    float indexOfRefraction;
};
struct AABB { float min[3]; float max[3]; };
struct Ray { float3 origin; float3 direction; };
struct Sphere { float position[3]; Material material; float radius; };
struct Triangle { float position[3][3]; };
struct VertexInfo { float normal[3]; float uv[2]; float tangent[3]; float bitangent[3]; };
struct TriangleInfo { VertexInfo vertexInfo[3]; };
struct Mesh {
    float transformMatrix[16];
    float inverseTransformMatrix[16];
    uint32_t triangleStartIdx, triangleEndIdx;
    AABB boundingBox;
    Material material;
    uint32_t bvhNodeStartIdx;
};
struct DirectionalLight { float colour[3]; float intensity; float effectiveAngle; float direction[3]; };
struct PointLight { float colour[3]; float intensity; float radius; float position[3]; };
struct SpotLight { float colour[3]; float intensity; float radius; float maxAngle; float position[3]; float direction[3]; };
struct BvhNode { AABB boundingBox; uint32_t triStart, triEnd, firstChildIdx; };
struct OctTreeNode { AABB boundingBox; uint32_t meshesStartIdx, meshesEndIdx, spheresStartIdx, spheresEndIdx, firstChildIdx, numChildren; };
struct RenderData {                       /* RaytracerCS.hlsl:19-47 */
    uint32_t accumulationEnabled, numAccumulationFrames, numSpheres, numMeshes;
    uint32_t numDirLights, numPointLights, numSpotLights, debugMode;
    uint32_t textureDims[2]; float viewPlaneDims[2];
    float cameraInverseProjectionMatrix[16];
    float cameraInverseViewMatrix[16];
    float cameraPosition[3]; float cameraNearZ;
    float cameraUpDir[3]; float dofStrength;
    float cameraRightDir[3]; float dofDistance;
    uint32_t debugMaxCount; float pad0[3];
};
static_assert(sizeof(Material) == 68, "Material");
static_assert(sizeof(Sphere) == 84, "Sphere");
static_assert(sizeof(Triangle) == 36, "Triangle");
static_assert(sizeof(TriangleInfo) == 132, "TriangleInfo");
static_assert(sizeof(Mesh) == 232, "Mesh");
static_assert(sizeof(BvhNode) == 36, "BvhNode");
static_assert(sizeof(OctTreeNode) == 48, "OctTreeNode");
static_assert(sizeof(RenderData) == 240, "RenderData");
static_assert(sizeof(DirectionalLight) == 32 && sizeof(PointLight) == 32 && sizeof(SpotLight) == 48, "lights");

struct OracleScene {                      /* the binding table, t0..t10 */
    const Triangle* triPos;               /* t0 */
    const TriangleInfo* triInfo;          /* t1 */
    const BvhNode* bvhNodes;              /* t2 */
    const uint8_t* textures;              /* t3: layers x size x size RGBA8 */
    const uint8_t* envFaces;              /* t4: 6 x envH x envW RGBA8 (+X,-X,+Y,-Y,+Z,-Z) */
    const OctTreeNode* octNodes;          /* t5 */
    const Sphere* spheres;                /* t6 */
    const Mesh* meshes;                   /* t7 */
    const DirectionalLight* dirLights;    /* t8 */
    const PointLight* pointLights;        /* t9 */
    const SpotLight* spotLights;          /* t10 */
    uint32_t numTris, numBvhNodes, texSize, texLayers, envW, envH, numOctNodes;
    uint32_t numSpheres, numMeshes, numDirLights, numPointLights, numSpotLights; /* buffer element counts */
};

struct Aov {                              /* mirrors okrt_aov */
    uint32_t hitKind, hitIdx, triIdx;
    float baryU, baryV, distance;
    uint32_t bbCheckCount, triCheckCount, numSegments, flags;
    float radiance[3];
    uint32_t reserved;
};
static_assert(sizeof(Aov) == 56, "Aov");

enum { ORACLE_MATH_CANONICAL = 0, ORACLE_MATH_LIBM = 1 };

struct Ctx {
    const OracleScene* s;
    const RenderData* rd;
    int mathMode;
    uint32_t stackOverflow;               /* times the reference's fixed stacks would have overflowed */
    /* visit totals for the algorithmic-bytes roofline: oct-tree nodes, instances, BVH nodes,
     * triangle tests, segments that hit something, texture taps */
    uint64_t nOct = 0, nInst = 0, nBvh = 0, nTri = 0, nHit = 0, nTap = 0;
};

inline float m_log(const Ctx& c, float x) { return c.mathMode == ORACLE_MATH_LIBM ? std::log(x) : omath::c_log(x); }
inline float m_cos(const Ctx& c, float x) { return c.mathMode == ORACLE_MATH_LIBM ? std::cos(x) : omath::c_cos(x); }
inline float m_sin(const Ctx& c, float x) { return c.mathMode == ORACLE_MATH_LIBM ? std::sin(x) : omath::c_sin(x); }
inline float m_atan2(const Ctx& c, float y, float x) { return c.mathMode == ORACLE_MATH_LIBM ? std::atan2(y, x) : omath::c_atan2(y, x); }
inline float m_pow(const Ctx& c, float x, float y)
{
    if (c.mathMode == ORACLE_MATH_LIBM) {
        if (x < 0.0f) return std::numeric_limits<float>::quiet_NaN();  /* exp2(y*log2(x)) */
        return std::pow(x, y);
    }
    return omath::c_pow(x, y);
}

/* [Q1] mul(float4(v,w), M): the 16 floats in memory are row-major M; result_j = dot(row j, v). */
inline float3 mulPoint(float3 v, const float* m)
{
    return f3(((v.x * m[0] + v.y * m[1]) + v.z * m[2]) + m[3],
              ((v.x * m[4] + v.y * m[5]) + v.z * m[6]) + m[7],
              ((v.x * m[8] + v.y * m[9]) + v.z * m[10]) + m[11]);
}
inline float3 mulDir(float3 v, const float* m)
{
    return f3((v.x * m[0] + v.y * m[1]) + v.z * m[2],
              (v.x * m[4] + v.y * m[5]) + v.z * m[6],
              (v.x * m[8] + v.y * m[9]) + v.z * m[10]);
}
inline float4 mulPoint4(float3 v, const float* m)
{
    float4 r;
    r.x = ((v.x * m[0] + v.y * m[1]) + v.z * m[2]) + m[3];
    r.y = ((v.x * m[4] + v.y * m[5]) + v.z * m[6]) + m[7];
    r.z = ((v.x * m[8] + v.y * m[9]) + v.z * m[10]) + m[11];
    r.w = ((v.x * m[12] + v.y * m[13]) + v.z * m[14]) + m[15];
    return r;
}

inline bool isValidIdx(uint32_t idx) { return idx != UINT_MAX_; }  /* GPU-Utilities.hlsli:13-16 */

/* ------------------------------------------------------------------- RNG */
/* GPU-Utilities.hlsli:18-25 [F7]: `seed *= 747796405u + 2891336453u` multiplies by the SUM. */
inline uint32_t pcg_hash(uint32_t& seed)
{
    seed *= 747796405u + 2891336453u;
    seed = ((seed >> ((seed >> 28u) + 4u)) ^ seed) * 277803737u;
    seed = (seed >> 22u) ^ seed;
    return seed;
}
/* GPU-Utilities.hlsli:51-54: (float)UINT_MAX == 4294967296.0f */
inline float randomFloat(uint32_t& seed) { return (float)pcg_hash(seed) / (float)UINT_MAX_; }

/* GPU-Utilities.hlsli:56-62 */
inline float randomFloatNormalDistribution(const Ctx& c, uint32_t& seed)
{
    float theta = (2 * PI_) * randomFloat(seed);
    float rho = std::sqrt(-2 * m_log(c, randomFloat(seed)));
    return rho * m_cos(c, theta);
}
/* GPU-Utilities.hlsli:64-70 */
inline float3 getRandomVector(const Ctx& c, uint32_t& seed)
{
    float x = randomFloatNormalDistribution(c, seed);
    float y = randomFloatNormalDistribution(c, seed);
    float z = randomFloatNormalDistribution(c, seed);
    return normalize(f3(x, y, z));
}
/* GPU-Utilities.hlsli:72-78 */
inline float2 randomPointInCircle(const Ctx& c, uint32_t& seed)
{
    float angle = randomFloat(seed) * 2.f * PI_;
    float2 pointOnCircle = float2{m_cos(c, angle), m_sin(c, angle)};
    return pointOnCircle * std::sqrt(randomFloat(seed));
}
inline float clampedDot(float3 v, float3 u) { return hmax(dot(v, u), 0.f); }   /* :80-83 */

/* GPU-Utilities.hlsli:85-93 — shadows the HLSL intrinsic */
inline float3 refract_(float3 direction, float3 normal, float refractionRatio)
{
    float cosTheta = hmin(dot(-direction, normal), 1.f);
    float3 rayOutPerpendicular = refractionRatio * (direction + cosTheta * normal);
    float ropLengthSqrd = dot(rayOutPerpendicular, rayOutPerpendicular);
    float3 rayOutParallel = -std::sqrt(std::fabs(1.f - ropLengthSqrd)) * normal;
    return rayOutPerpendicular + rayOutParallel;
}

/* ---------------------------------------------------------- Collision:: */
/* GPU-Utilities.hlsli:98-114 [Q15] */
inline float RayAndSphere(const Ray& ray, float3 spherePos, float sphereRadius)
{
    float3 rayToSphere = spherePos - ray.origin;
    float distToClosestPoint = dot(rayToSphere, ray.direction);
    float rayToSphereMagSqrd = dot(rayToSphere, rayToSphere);
    float sphereRadiusSqrd = sphereRadius * sphereRadius;

    if (distToClosestPoint < 0.f && rayToSphereMagSqrd > sphereRadiusSqrd)
        return -1.f;

    float sideA = rayToSphereMagSqrd - distToClosestPoint * distToClosestPoint;
    if (sideA > sphereRadiusSqrd)
        return -1.f;

    float sideB = std::sqrt(sphereRadiusSqrd - sideA);
    return rayToSphereMagSqrd > sphereRadiusSqrd ? distToClosestPoint - sideB : distToClosestPoint + sideB;
}

/* GPU-Utilities.hlsli:117-155 [Q6] */
inline float RayAndTriangle(const Ray& ray, float3 p0, float3 p1, float3 p2, float2& baryUVCoord)
{
    const float EPSILON = 0.000001f;
    float3 E1 = p0 - p2, E2 = p1 - p2;
    float3 cross1 = cross(ray.direction, E2);
    float determinant = dot(E1, cross1);

    if (determinant < EPSILON && determinant > -EPSILON)
        return -1.f;

    float inverseDet = 1.f / determinant;

    float3 rayMoved = ray.origin - p2;
    baryUVCoord.x = dot(rayMoved, cross1) * inverseDet;
    if (baryUVCoord.x < -EPSILON)
        return -1.f;

    float3 cross2 = cross(rayMoved, E1);
    baryUVCoord.y = dot(ray.direction, cross2) * inverseDet;
    if (baryUVCoord.y < -EPSILON || baryUVCoord.x + baryUVCoord.y > 1.0f)
        return -1.f;

    float t1 = dot(E2, cross2) * inverseDet;
    if (t1 < -EPSILON)
        return -1.f;

    return t1;
}

/* GPU-Utilities.hlsli:183-197 */
inline float RayAndAABBDist(const Ray& ray, const AABB& aabb)
{
    float3 inverseRayDir = f3(1.f / ray.direction.x, 1.f / ray.direction.y, 1.f / ray.direction.z);
    float3 tMin = (f3(aabb.min) - ray.origin) * inverseRayDir;
    float3 tMax = (f3(aabb.max) - ray.origin) * inverseRayDir;
    float3 t1 = f3(hmin(tMin.x, tMax.x), hmin(tMin.y, tMax.y), hmin(tMin.z, tMax.z));
    float3 t2 = f3(hmax(tMin.x, tMax.x), hmax(tMin.y, tMax.y), hmax(tMin.z, tMax.z));

    float distFar = hmin(hmin(t2.x, t2.y), t2.z);
    float distNear = hmax(hmax(t1.x, t1.y), t1.z);

    bool didHit = distFar >= distNear && distFar > 0.f;
    return didHit ? distNear : FLT_MAX_;
}

/* ------------------------------------------------------------- textures */
/* D3D11 UNORM8 -> float is exactly c/255 */
inline float3 texel(const uint8_t* p) { return f3((float)p[0] / 255.0f, (float)p[1] / 255.0f, (float)p[2] / 255.0f); }

inline bool finitef(float x) { return x - x == 0.0f; }

/* bilinear weights in plain float: ((c00*(1-fx) + c10*fx)*(1-fy)) + ((c01*(1-fx) + c11*fx)*fy) */
inline float3 bilerp(float3 c00, float3 c10, float3 c01, float3 c11, float fx, float fy)
{
    float wx0 = 1.0f - fx, wy0 = 1.0f - fy;
    float3 top = c00 * wx0 + c10 * fx;
    float3 bot = c01 * wx0 + c11 * fx;
    return top * wy0 + bot * fy;
}

/* RaytracerCS.hlsl:133-136: textures.SampleLevel(simp, float3(uv, idx), 0) — bilinear, WRAP, lod 0 [Q12].
 * D3D's fixed-point filter weights are hardware behaviour; restated as exact float bilinear. */
inline float3 sampleTexture(const Ctx& c, uint32_t textureIdx, float2 uv)
{
    const OracleScene& s = *c.s;
    const_cast<Ctx&>(c).nTap++;
    if (!s.textures || s.texLayers == 0 || s.texSize == 0) return f3(0, 0, 0);   /* null SRV reads 0 */
    uint32_t layer = textureIdx < s.texLayers ? textureIdx : s.texLayers - 1;     /* slice index clamps */
    const int N = (int)s.texSize;
    float u = finitef(uv.x) ? uv.x : 0.0f, v = finitef(uv.y) ? uv.y : 0.0f;
    float x = u * (float)N - 0.5f, y = v * (float)N - 0.5f;
    if (!(std::fabs(x) < 1.0e9f)) x = 0.0f;
    if (!(std::fabs(y) < 1.0e9f)) y = 0.0f;
    float xf = std::floor(x), yf = std::floor(y);
    float fx = x - xf, fy = y - yf;
    int x0 = (int)xf % N; if (x0 < 0) x0 += N;
    int y0 = (int)yf % N; if (y0 < 0) y0 += N;
    int x1 = x0 + 1 == N ? 0 : x0 + 1;
    int y1 = y0 + 1 == N ? 0 : y0 + 1;
    const uint8_t* base = s.textures + (size_t)layer * N * N * 4;
    return bilerp(texel(base + ((size_t)y0 * N + x0) * 4), texel(base + ((size_t)y0 * N + x1) * 4),
                  texel(base + ((size_t)y1 * N + x0) * 4), texel(base + ((size_t)y1 * N + x1) * 4), fx, fy);
}

/* RaytracerCS.hlsl:86-88: environmentMap.SampleLevel(simp, direction, 0) [Q13].
 * D3D cube addressing: major axis picks the face, (sc,tc)/|ma| -> [0,1]^2, bilinear inside the face
 * (edge texels clamp; cross-face filtering at seams is hardware behaviour and not restated). */
inline float3 getEnvironmentLight(const Ctx& c, float3 d)
{
    const OracleScene& s = *c.s;
    if (!s.envFaces || s.envW == 0 || s.envH == 0) return f3(0, 0, 0);
    float ax = std::fabs(d.x), ay = std::fabs(d.y), az = std::fabs(d.z);
    int face; float sc, tc, ma;
    if (ax >= ay && ax >= az) { ma = ax; if (d.x >= 0) { face = 0; sc = -d.z; tc = -d.y; } else { face = 1; sc = d.z; tc = -d.y; } }
    else if (ay >= az)        { ma = ay; if (d.y >= 0) { face = 2; sc = d.x; tc = d.z; } else { face = 3; sc = d.x; tc = -d.z; } }
    else                      { ma = az; if (d.z >= 0) { face = 4; sc = d.x; tc = -d.y; } else { face = 5; sc = -d.x; tc = -d.y; } }
    if (!(ma > 0.0f) || !finitef(ma) || !finitef(sc) || !finitef(tc)) return f3(0, 0, 0);
    float u = (sc / ma + 1.0f) * 0.5f, v = (tc / ma + 1.0f) * 0.5f;
    const int W = (int)s.envW, H = (int)s.envH;
    float x = u * (float)W - 0.5f, y = v * (float)H - 0.5f;
    float xf = std::floor(x), yf = std::floor(y);
    float fx = x - xf, fy = y - yf;
    int x0 = (int)xf, y0 = (int)yf, x1 = x0 + 1, y1 = y0 + 1;
    x0 = std::min(std::max(x0, 0), W - 1); x1 = std::min(std::max(x1, 0), W - 1);
    y0 = std::min(std::max(y0, 0), H - 1); y1 = std::min(std::max(y1, 0), H - 1);
    const uint8_t* base = s.envFaces + (size_t)face * W * H * 4;
    return bilerp(texel(base + ((size_t)y0 * W + x0) * 4), texel(base + ((size_t)y0 * W + x1) * 4),
                  texel(base + ((size_t)y1 * W + x0) * 4), texel(base + ((size_t)y1 * W + x1) * 4), fx, fy);
}

/* RaytracerCS.hlsl:77-84 */
inline float2 normalToUV(const Ctx& c, float3 normal)
{
    float2 uv;
    uv.x = m_atan2(c, normal.x, normal.z) / (2 * PI_) + 0.5f;
    uv.y = normal.y * 0.5f + 0.5f;
    return uv;
}

/* RaytracerCS.hlsl:100-105 [Q7]: (1 + r0), sic */
inline float reflectance(const Ctx& c, float cosine, float reflectionIdx)
{
    float r0 = (1.f - reflectionIdx) / (1.f + reflectionIdx);
    r0 *= r0;
    return r0 + (1.f + r0) * m_pow(c, 1.f - cosine, 5.f);
}

/* RaytracerCS.hlsl:107-112 */
inline float3 findReflectDirection(const Ctx& c, float3 direction, float3 normal, float roughness, uint32_t& seed)
{
    float3 diffuseReflection = normalize(normal + getRandomVector(c, seed));
    float3 specularReflection = reflect(direction, normal);
    return normalize(lerp(specularReflection, diffuseReflection, roughness));
}

/* RaytracerCS.hlsl:114-131 [Q7]: no short-circuit — randomFloat is always consumed */
inline float3 findTransparencyBounce(const Ctx& c, float3 direction, float3 normal, float refractionIdx, uint32_t& seed)
{
    bool hitFrontFace = dot(direction, normal) < 0.f;
    float refractionRatio = hitFrontFace ? AIR_REFRACTION_INDEX / refractionIdx : refractionIdx;
    if (!hitFrontFace)
        normal = normal * -1.f;

    float cos_theta = dot(-direction, normal);
    float sin_theta = std::sqrt(1.f - cos_theta * cos_theta);
    bool cannot_refract = refractionRatio * sin_theta > 1.f;

    float refl = reflectance(c, cos_theta, refractionRatio);
    float rnd = randomFloat(seed);
    if (cannot_refract || refl > rnd)
        direction = reflect(direction, normal);

    return refract_(direction, normal, refractionRatio);
}

/* RaytracerCS.hlsl:138-165 */
inline void findMaterialTextureColours(const Ctx& c, Material& material, float2 meshUVs)
{
    if (isValidIdx(material.albedoTex)) {
        float3 t = sampleTexture(c, material.albedoTex, meshUVs);
        material.albedo[0] = t.x; material.albedo[1] = t.y; material.albedo[2] = t.z;
    }
    /* clamp(x, lo, 1) = min(max(x, lo), 1) */
    if (isValidIdx(material.roughnessTex)) {
        float r = sampleTexture(c, material.roughnessTex, meshUVs).x;
        material.roughness = hmin(hmax(r, material.roughness), 1.f);
    }
    if (isValidIdx(material.metallicTex)) {
        float m = sampleTexture(c, material.metallicTex, meshUVs).x;
        material.metallic = hmin(hmax(m, material.metallic), 1.f);
    }
    if (isValidIdx(material.specularTex)) {
        float sp = sampleTexture(c, material.specularTex, meshUVs).x;
        material.specular = hmin(hmax(sp, material.specular), 1.f);
    }
}

/* RaytracerCS.hlsl:167-175: mul(s, float3x3(T,B,N)) = s.x*T + s.y*B + s.z*N */
inline float3 sampleNormalMap(const Ctx& c, uint32_t textureIdx, float2 meshUVs, float3 normal, float3 tangent, float3 bitangent)
{
    float3 sn = sampleTexture(c, textureIdx, meshUVs);
    sn = sn * 2.f - f3(1.f, 1.f, 1.f);
    float3 r = f3((sn.x * tangent.x + sn.y * bitangent.x) + sn.z * normal.x,
                  (sn.x * tangent.y + sn.y * bitangent.y) + sn.z * normal.y,
                  (sn.x * tangent.z + sn.y * bitangent.z) + sn.z * normal.z);
    return normalize(r);
}

/* RaytracerCS.hlsl:177-206 [Q2] */
inline Ray createRay(const Ctx& c, uint32_t px, uint32_t py, uint32_t& seed)
{
    const RenderData& rd = *c.rd;
    float3 pos = f3((float)px, (float)(rd.textureDims[1] - py), rd.cameraNearZ);
    pos.x += randomFloat(seed) - 0.5f;
    pos.y += randomFloat(seed) - 0.5f;

    pos.x /= (float)rd.textureDims[0]; pos.y /= (float)rd.textureDims[1];
    pos.x *= 2.f; pos.y *= 2.f;
    pos.x -= 1.f; pos.y -= 1.f;

    float4 target = mulPoint4(pos, rd.cameraInverseProjectionMatrix);
    Ray ray;
    ray.origin = f3(rd.cameraPosition);
    float3 t = f3(target.x / target.z, target.y / target.z, target.z / target.z);
    ray.direction = mulDir(normalize(t), rd.cameraInverseViewMatrix);
    return ray;
}

/* RaytracerCS.hlsl:208-216 */
inline void applyDOF(const Ctx& c, Ray& ray, uint32_t& seed)
{
    const RenderData& rd = *c.rd;
    float2 rayJitter = randomPointInCircle(c, seed) * rd.dofStrength;
    float3 rayOffset = f3(rd.cameraRightDir) * rayJitter.x + f3(rd.cameraUpDir) * rayJitter.y;
    float3 focusPoint = ray.origin + ray.direction * (rd.dofDistance + rd.cameraNearZ);
    ray.origin = ray.origin + rayOffset;
    ray.direction = normalize(focusPoint - ray.origin);
}

struct Payload {                          /* RaytracerCS.hlsl:10-17 */
    bool hit;
    float distance;
    Material material;
    float3 worldPosition;
    float3 worldNormal;
};

inline Material materialCreate()          /* GPU-Structs.hlsli:52-70 */
{
    Material m;
    std::memset(&m, 0, sizeof(m));
    m.albedoTex = m.roughnessTex = m.metallicTex = m.specularTex = UINT_MAX_;
    m.normalMapIdx = UINT_MAX_;
    return m;
}

struct HitTap { uint32_t hitKind, hitIdx, triIdx; float u, v; };

/* RaytracerCS.hlsl:218-395 [Q4,Q5,Q8] */
Payload findClosestHit(Ctx& c, const Ray& ray, uint32_t& bbCheckCount, uint32_t& triCheckCount, HitTap* tap)
{
    const OracleScene& s = *c.s;
    const RenderData& rd = *c.rd;
    Payload payload;
    payload.hit = false;
    payload.distance = FLT_MAX_;
    payload.material = materialCreate();
    payload.worldPosition = f3(0, 0, 0);
    payload.worldNormal = f3(0, 0, 0);

    uint32_t hitIdx = UINT_MAX_;
    uint32_t hitType = 0;

    for (uint32_t i = 0; i < rd.numSpheres; i++) {
        if (i >= s.numSpheres) break;     /* OOB structured reads return 0: radius 0 at the origin never wins */
        float distanceToHit = RayAndSphere(ray, f3(s.spheres[i].position), s.spheres[i].radius);
        if (distanceToHit > 0.f && distanceToHit < payload.distance) {
            payload.distance = distanceToHit;
            hitIdx = i;
            hitType = 0;
        }
    }

    /* [Q5] `half` stacks are float32 under cs_5_0: exact for indices < 2^24.  The reference has no
     * overflow check on its 20/50-entry stacks; the oracle keeps larger ones and counts overflows. */
    const uint32_t OCT_MAX_STACK_SIZE = 20, BVH_MAX_STACK_SIZE = 50;
    uint32_t octStack[256];
    octStack[0] = 0;
    uint32_t octStackSize = s.numOctNodes ? 1 : 0;
    uint32_t bvhStack[256];
    uint32_t bvhStackSize = 0;

    uint32_t triHitIdx = UINT_MAX_;
    float3 hitBary = f3(0, 0, 0);

    while (octStackSize > 0) {
        uint32_t octCurrentNodeIdx = octStack[--octStackSize];
        const OctTreeNode& octNode = s.octNodes[octCurrentNodeIdx];

        bbCheckCount += 1;
        c.nOct++;
        if (RayAndAABBDist(ray, octNode.boundingBox) >= payload.distance)
            continue;

        for (uint32_t i = octNode.meshesStartIdx; i < octNode.meshesEndIdx; i++) {
            bbCheckCount += 1;
            c.nInst++;
            const Mesh& mesh = s.meshes[i];
            bvhStack[0] = mesh.bvhNodeStartIdx;
            bvhStackSize = 1;

            Ray localRay;
            localRay.origin = mulPoint(ray.origin, mesh.inverseTransformMatrix);
            localRay.direction = normalize(mulDir(ray.direction, mesh.inverseTransformMatrix));

            while (bvhStackSize > 0) {
                uint32_t currentNodeIdx = bvhStack[--bvhStackSize];
                const BvhNode& bvhNode = s.bvhNodes[currentNodeIdx];

                bbCheckCount += 1;
                c.nBvh++;
                if (RayAndAABBDist(localRay, bvhNode.boundingBox) >= payload.distance)
                    continue;

                if (isValidIdx(bvhNode.firstChildIdx)) {
                    if (bvhStackSize + 2 > BVH_MAX_STACK_SIZE) c.stackOverflow++;
                    if (bvhStackSize + 2 > 256) continue;
                    bvhStack[bvhStackSize++] = bvhNode.firstChildIdx;
                    bvhStack[bvhStackSize++] = bvhNode.firstChildIdx + 1;
                    continue;
                }

                for (uint32_t j = bvhNode.triStart; j < bvhNode.triEnd; j++) {
                    const Triangle& tri = s.triPos[j];
                    float2 baryUVCoords = float2{0.f, 0.f};
                    triCheckCount += 1;
                    c.nTri++;
                    float distanceToHit = RayAndTriangle(localRay, f3(tri.position[0]), f3(tri.position[1]), f3(tri.position[2]), baryUVCoords);
                    if (distanceToHit <= 0.f)
                        continue;

                    float3 localRayToHit = localRay.direction * distanceToHit;
                    distanceToHit = length(mulDir(localRayToHit, mesh.transformMatrix));

                    if (distanceToHit < payload.distance) {
                        payload.distance = distanceToHit;
                        hitIdx = i;
                        hitType = 1;
                        triHitIdx = j;
                        hitBary.x = baryUVCoords.x; hitBary.y = baryUVCoords.y;
                        hitBary.z = 1.f - (hitBary.x + hitBary.y);
                    }
                }
            }
        }

        for (uint32_t i = 0; i < octNode.numChildren; i++) {
            if (octStackSize + 1 > OCT_MAX_STACK_SIZE) c.stackOverflow++;
            if (octStackSize + 1 > 256) break;
            octStack[octStackSize++] = octNode.firstChildIdx + i;
        }
    }

    payload.hit = hitIdx != UINT_MAX_;
    payload.worldPosition = ray.origin + ray.direction * payload.distance;

    if (tap) {
        tap->hitKind = !payload.hit ? 0u : (hitType == 0 ? 1u : 2u);
        tap->hitIdx = hitIdx;
        tap->triIdx = (payload.hit && hitType == 1) ? triHitIdx : UINT_MAX_;
        tap->u = hitType == 1 ? hitBary.x : 0.f;
        tap->v = hitType == 1 ? hitBary.y : 0.f;
    }

    /* [Q8] on a miss the shader runs the sphere branch on zeros and discards the result. */
    if (!payload.hit)
        return payload;
    c.nHit++;

    if (hitType == 0) {                   /* RaytracerCS.hlsl:342-362 */
        const Sphere& sp = s.spheres[hitIdx];
        payload.material = sp.material;
        float3 hitNormal = normalize(payload.worldPosition - f3(sp.position));
        float2 sphereUVs = normalToUV(c, hitNormal);
        if (isValidIdx(payload.material.normalMapIdx)) {
            float3 normal = hitNormal;
            float3 tangent = normalize(cross(f3(0.f, 1.f, 0.f), hitNormal));
            float3 bitangent = normalize(cross(tangent, hitNormal));
            payload.worldNormal = sampleNormalMap(c, payload.material.normalMapIdx, sphereUVs, normal, tangent, bitangent);
        } else {
            payload.worldNormal = hitNormal;
        }
        findMaterialTextureColours(c, payload.material, sphereUVs);
    } else {                              /* RaytracerCS.hlsl:364-391 */
        const Mesh& mesh = s.meshes[hitIdx];
        payload.material = mesh.material;
        const TriangleInfo& tri = s.triInfo[triHitIdx];
        const VertexInfo& p0 = tri.vertexInfo[0];
        const VertexInfo& p1 = tri.vertexInfo[1];
        const VertexInfo& p2 = tri.vertexInfo[2];
        /* barycentricInterpolation: value0*u + value1*v + value2*w (GPU-Utilities.hlsli:3-11) */
        float2 lerpedUV = float2{(p0.uv[0] * hitBary.x + p1.uv[0] * hitBary.y) + p2.uv[0] * hitBary.z,
                                 (p0.uv[1] * hitBary.x + p1.uv[1] * hitBary.y) + p2.uv[1] * hitBary.z};
        auto bary3 = [&](const float* a, const float* b, const float* cc) {
            return (f3(a) * hitBary.x + f3(b) * hitBary.y) + f3(cc) * hitBary.z;
        };
        float3 normal = mulDir(bary3(p0.normal, p1.normal, p2.normal), mesh.transformMatrix);
        float3 tangent = mulDir(bary3(p0.tangent, p1.tangent, p2.tangent), mesh.transformMatrix);
        float3 bitangent = mulDir(bary3(p0.bitangent, p1.bitangent, p2.bitangent), mesh.transformMatrix);
        normal = normalize(normal);
        tangent = normalize(tangent);
        bitangent = normalize(bitangent);
        if (isValidIdx(payload.material.normalMapIdx))
            payload.worldNormal = sampleNormalMap(c, payload.material.normalMapIdx, lerpedUV, normal, tangent, bitangent);
        else
            payload.worldNormal = normal;
        findMaterialTextureColours(c, payload.material, lerpedUV);
    }
    return payload;
}

/* RaytracerCS.hlsl:397-418 [Q11] */
inline float3 getLighting(const Ray& ray, const Payload& hitPayload, float3 lightPos, float lightRadius, float3 lightColour, float lightIntensity)
{
    float dist = RayAndSphere(ray, lightPos, lightRadius);
    if (dist < 0.f)
        return f3(0.f, 0.f, 0.f);

    float lightStrengthModifier = 1.f;
    if (hitPayload.hit && hitPayload.distance > dist)
        lightStrengthModifier *= clampedDot(ray.direction, -hitPayload.worldNormal);
    else if (hitPayload.hit && hitPayload.distance < dist)
        lightStrengthModifier *= hitPayload.material.transparency * clampedDot(ray.direction, -hitPayload.worldNormal);

    return (lightColour * lightIntensity) * lightStrengthModifier;
}

struct PixelOut { float3 light; uint32_t bb, tri, segments; HitTap tap; float tapDistance; };

/* RaytracerCS.hlsl:421-520 (everything before the accumulate/store) */
PixelOut tracePixel(Ctx& c, uint32_t px, uint32_t py, uint32_t seedFrame, uint32_t maxBounces)
{
    const OracleScene& s = *c.s;
    const RenderData& rd = *c.rd;
    PixelOut out;
    std::memset(&out, 0, sizeof(out));

    /* [Q3] */
    uint32_t seed = px + (py + 74813u) * rd.textureDims[0] * (seedFrame + 1u);

    Ray ray = createRay(c, px, py, seed);
    applyDOF(c, ray, seed);

    float3 light = f3(0.f, 0.f, 0.f);
    float3 contribution = f3(1.f, 1.f, 1.f);
    uint32_t bbCheckCount = 0, triCheckCount = 0;

    for (uint32_t i = 0; i <= maxBounces; i++) {
        HitTap tap;
        Payload hitData = findClosestHit(c, ray, bbCheckCount, triCheckCount, &tap);
        out.segments++;
        if (i == 0) { out.tap = tap; out.tapDistance = hitData.distance; }

        for (uint32_t p = 0u; p < rd.numPointLights && p < s.numPointLights; p++) {
            const PointLight& pl = s.pointLights[p];
            float3 lightning = getLighting(ray, hitData, f3(pl.position), pl.radius, f3(pl.colour), pl.intensity);
            light = light + lightning * contribution;
        }
        for (uint32_t sp = 0u; sp < rd.numSpotLights && sp < s.numSpotLights; sp++) {
            const SpotLight& sl = s.spotLights[sp];
            float3 rayToLight = normalize((f3(sl.position) + getRandomVector(c, seed) * 0.01f) - ray.origin);
            float cosTheta = dot(rayToLight, -f3(sl.direction));
            if (cosTheta < sl.maxAngle)
                continue;
            float3 lightning = getLighting(ray, hitData, f3(sl.position), sl.radius, f3(sl.colour), sl.intensity);
            light = light + lightning * contribution;
        }
        for (uint32_t d = 0u; d < rd.numDirLights && d < s.numDirLights; d++) {
            const DirectionalLight& dl = s.dirLights[d];
            if (clampedDot(ray.direction, -f3(dl.direction)) < dl.effectiveAngle)
                continue;
            float lightStrengthModifier = 1.f;
            if (hitData.hit && i)
                lightStrengthModifier *= hitData.material.transparency * clampedDot(ray.direction, -hitData.worldNormal);
            else if (!i)
                lightStrengthModifier = 0.f;
            light = light + ((f3(dl.colour) * dl.intensity) * contribution) * lightStrengthModifier;
        }

        if (!hitData.hit) {
            light = light + getEnvironmentLight(c, ray.direction) * contribution;
            break;
        }

        const Material& material = hitData.material;

        /* [Q10] */
        float metallicFactor = (material.metallic * (1.f - material.roughness)) >= randomFloat(seed) ? 1.f : 0.f;
        float specularFactor = (material.specular * (1.f - material.roughness)) >= randomFloat(seed) ? 1.f : 0.f;
        float transparencyFactor = material.transparency >= randomFloat(seed) ? 1.f : 0.f;

        float3 reflectDir = findReflectDirection(c, ray.direction, hitData.worldNormal, material.roughness * (1.f - specularFactor), seed);
        float3 refractDir = findTransparencyBounce(c, ray.direction, hitData.worldNormal, material.indexOfRefraction, seed);

        float3 bounceDir = normalize(lerp(reflectDir, refractDir, transparencyFactor));
        float3 hitPoint = hitData.worldPosition + bounceDir * 0.001f;

        contribution = contribution * lerp(f3(material.albedo), f3(1.f, 1.f, 1.f), metallicFactor);
        light = light + ((f3(material.emissionColour) * material.emissionPower) * contribution) * (1.f - transparencyFactor);

        ray.origin = hitPoint;
        ray.direction = bounceDir;
    }

    /* RaytracerCS.hlsl:516-520 */
    uint32_t debugModeMaxCount = rd.debugMaxCount;
    if (rd.debugMode == 1)
        light = bbCheckCount > debugModeMaxCount ? f3(1.f, 0.f, 0.f) : f3(1.f, 1.f, 1.f) * ((float)bbCheckCount / (float)debugModeMaxCount);
    else if (rd.debugMode == 2)
        light = triCheckCount > debugModeMaxCount ? f3(1.f, 0.f, 0.f) : f3(1.f, 1.f, 1.f) * ((float)triCheckCount / (float)debugModeMaxCount);

    out.light = light;
    out.bb = bbCheckCount;
    out.tri = triCheckCount;
    return out;
}

/* D3D11 float -> UNORM8: saturate, scale by 255, add 0.5, truncate; NaN -> 0 */
inline uint8_t toUnorm8(float x) { return (uint8_t)(saturate(x) * 255.0f + 0.5f); }

} // namespace

/* =================================================================== C API */

ORACLE_API const char* oracle_description(void)
{
    return "okrt CPU oracle: scalar restatement of RaytracerCS.hlsl / GPU-Utilities.hlsli (test infrastructure)";
}

/* visit totals of the last oracle_render call: segments, oct-tree nodes, instances, BVH nodes, triangle tests,
 * hit segments, texture taps */
static uint64_t g_lastVisits[7];
ORACLE_API void oracle_last_visit_totals(uint64_t* out7) { for (int k = 0; k < 7; k++) out7[k] = g_lastVisits[k]; }

/* Full frame (or row band [y0,y1)) of RaytracerCS::main, RaytracerCS.hlsl:421-537, preceded by the host's
 * clear of the result target (RayTracer.cpp:227-228) for those rows.
 *   seedFrame   value of renderData.numAccumulationFrames used in the seed [Q3]
 *   coverAll    0: only pixels the reference dispatch reaches, x < 16*(W/16), y < 9*(H/9) [Q9]
 *   accum       RGBA32F accumulation image, read-modify-write (u1)
 *   result      RGBA8 result image (u0)
 * Returns the number of stack overflows the reference's fixed stacks would have had. */
static uint32_t renderRows(const OracleScene* scene, const RenderData* rd, uint32_t maxBounces, uint32_t seedFrame,
                           uint32_t coverAll, float* accum, uint8_t* result, Aov* aov,
                           const uint32_t* rowList, uint32_t y0, uint32_t y1, int threads, int mathMode, uint64_t* outSegments)
{
    const uint32_t W = rd->textureDims[0], H = rd->textureDims[1];
    if (y1 > H) y1 = H;
    const uint32_t coveredW = coverAll ? W : 16u * (W / 16u);
    const uint32_t coveredH = coverAll ? H : 9u * (H / 9u);
    uint64_t segments = 0;
    uint32_t overflow = 0;
    /* rows are handed out dynamically to `threads` std::threads (<= 0: all hardware threads) */
    unsigned nthreads = threads > 0 ? (unsigned)threads : std::max(1u, std::thread::hardware_concurrency());
    std::atomic<uint32_t> nextRow(y0);
    std::atomic<uint64_t> segTotal(0);
    std::atomic<uint32_t> ovfTotal(0);
    std::atomic<uint64_t> visit[6];
    for (auto& v : visit) v = 0;
    auto worker = [&]() {
        Ctx c{scene, rd, mathMode, 0};
        uint64_t localSegments = 0;
        for (;;) {
            const uint32_t slotIdx = nextRow.fetch_add(1);
            if (slotIdx >= y1) break;
            const uint32_t y = rowList ? rowList[slotIdx] : slotIdx;
            if (y >= H) continue;
            for (uint32_t x = 0; x < W; x++) {
                const size_t pi = (size_t)y * W + x;
                if (result) { result[pi * 4 + 0] = 51; result[pi * 4 + 1] = 102; result[pi * 4 + 2] = 153; result[pi * 4 + 3] = 255; }
                if (aov) std::memset(&aov[pi], 0, sizeof(Aov));
                if (x >= coveredW || y >= coveredH) continue;
                PixelOut po = tracePixel(c, x, y, seedFrame, maxBounces);
                localSegments += po.segments;
                if (aov) {
                    Aov& a = aov[pi];
                    a.hitKind = po.tap.hitKind; a.hitIdx = po.tap.hitIdx; a.triIdx = po.tap.triIdx;
                    a.baryU = po.tap.u; a.baryV = po.tap.v; a.distance = po.tapDistance;
                    a.bbCheckCount = po.bb; a.triCheckCount = po.tri; a.numSegments = po.segments; a.flags = 0;
                    a.radiance[0] = po.light.x; a.radiance[1] = po.light.y; a.radiance[2] = po.light.z;
                }
                /* RaytracerCS.hlsl:525-536 [Q14] */
                float r4[4];
                if (rd->accumulationEnabled == 1) {
                    float* acc = accum + pi * 4;
                    acc[0] += po.light.x; acc[1] += po.light.y; acc[2] += po.light.z; acc[3] += 1.f;
                    const float n = (float)rd->numAccumulationFrames;
                    for (int k = 0; k < 4; k++) r4[k] = saturate(acc[k] / n);
                } else {
                    r4[0] = saturate(po.light.x); r4[1] = saturate(po.light.y); r4[2] = saturate(po.light.z); r4[3] = 1.f;
                }
                if (result) for (int k = 0; k < 4; k++) result[pi * 4 + k] = toUnorm8(r4[k]);
            }
        }
        segTotal += localSegments;
        ovfTotal += c.stackOverflow;
        visit[0] += c.nOct; visit[1] += c.nInst; visit[2] += c.nBvh; visit[3] += c.nTri; visit[4] += c.nHit; visit[5] += c.nTap;
    };
    if (nthreads <= 1) worker();
    else {
        std::vector<std::thread> pool;
        for (unsigned t = 0; t < nthreads; t++) pool.emplace_back(worker);
        for (auto& t : pool) t.join();
    }
    segments = segTotal.load();
    overflow = ovfTotal.load();
    g_lastVisits[0] = segments;
    for (int k = 0; k < 6; k++) g_lastVisits[k + 1] = visit[k].load();
    if (outSegments) *outSegments = segments;
    return overflow;
}

ORACLE_API uint32_t oracle_render(const OracleScene* scene, const RenderData* rd, uint32_t maxBounces, uint32_t seedFrame,
                                  uint32_t coverAll, float* accum, uint8_t* result, Aov* aov,
                                  uint32_t y0, uint32_t y1, int threads, int mathMode, uint64_t* outSegments)
{
    return renderRows(scene, rd, maxBounces, seedFrame, coverAll, accum, result, aov, nullptr, y0, y1, threads, mathMode, outSegments);
}

/* findClosestHit (RaytracerCS.hlsl:218-334) for arbitrary rays: 6 floats per ray in (origin, direction), 6 words per ray
 * out (distance, u, v as float bits, triangle, kind 0/1/2, instance-or-sphere index).  Lets the tests compare a traversal of
 * the product with the reference order ray by ray, without rendering. */
ORACLE_API void oracle_trace_rays(const OracleScene* scene, const RenderData* rd, const float* rays, uint32_t numRays, uint32_t* out, int threads)
{
    unsigned nthreads = threads > 0 ? (unsigned)threads : std::max(1u, std::thread::hardware_concurrency());
    std::atomic<uint32_t> nextRay(0);
    auto worker = [&]() {
        Ctx c;
        c.s = scene; c.rd = rd; c.mathMode = ORACLE_MATH_CANONICAL; c.stackOverflow = 0;
        for (;;) {
            const uint32_t begin = nextRay.fetch_add(1024u);
            if (begin >= numRays) break;
            for (uint32_t i = begin; i < std::min(numRays, begin + 1024u); i++) {
                const float* r = rays + (size_t)i * 6;
                Ray ray; ray.origin = float3{r[0], r[1], r[2]}; ray.direction = float3{r[3], r[4], r[5]};
                uint32_t bb = 0, tc = 0;
                HitTap tap;
                const Payload p = findClosestHit(c, ray, bb, tc, &tap);
                uint32_t* o = out + (size_t)i * 6;
                std::memcpy(o + 0, &p.distance, 4); std::memcpy(o + 1, &tap.u, 4); std::memcpy(o + 2, &tap.v, 4);
                o[3] = tap.triIdx; o[4] = tap.hitKind; o[5] = tap.hitIdx;
            }
        }
    };
    std::vector<std::thread> pool;
    for (unsigned t = 0; t < nthreads; t++) pool.emplace_back(worker);
    for (auto& th : pool) th.join();
}

/* same, for an explicit list of image rows (bounded samples of a frame, tile-sharded ranks) */
ORACLE_API uint32_t oracle_render_rows(const OracleScene* scene, const RenderData* rd, uint32_t maxBounces, uint32_t seedFrame,
                                       uint32_t coverAll, float* accum, uint8_t* result, Aov* aov,
                                       const uint32_t* rows, uint32_t numRows, int threads, int mathMode, uint64_t* outSegments)
{
    return renderRows(scene, rd, maxBounces, seedFrame, coverAll, accum, result, aov, rows, 0u, numRows, threads, mathMode, outSegments);
}

/* ---------------------------------------------------------------- KATs */
ORACLE_API uint32_t oracle_pcg_hash(uint32_t* seed) { return pcg_hash(*seed); }
ORACLE_API float oracle_random_float(uint32_t* seed) { return randomFloat(*seed); }
ORACLE_API float oracle_ray_triangle(const float* o, const float* d, const float* p0, const float* p1, const float* p2, float* uv)
{
    Ray r{f3(o), f3(d)}; float2 b{0, 0};
    float t = RayAndTriangle(r, f3(p0), f3(p1), f3(p2), b);
    uv[0] = b.x; uv[1] = b.y;
    return t;
}
ORACLE_API float oracle_ray_aabb_dist(const float* o, const float* d, const float* bmin, const float* bmax)
{
    Ray r{f3(o), f3(d)}; AABB a;
    for (int k = 0; k < 3; k++) { a.min[k] = bmin[k]; a.max[k] = bmax[k]; }
    return RayAndAABBDist(r, a);
}
ORACLE_API float oracle_ray_sphere(const float* o, const float* d, const float* c, float radius)
{
    Ray r{f3(o), f3(d)};
    return RayAndSphere(r, f3(c), radius);
}
/* which: 0 log, 1 cos, 2 sin, 3 exp, 4 atan2(x,y=arg2), 5 pow(x, arg2) — canonical implementations */
ORACLE_API void oracle_math_eval(int which, const float* x, const float* arg2, float* out, uint32_t n)
{
    for (uint32_t i = 0; i < n; i++) {
        switch (which) {
        case 0: out[i] = omath::c_log(x[i]); break;
        case 1: out[i] = omath::c_cos(x[i]); break;
        case 2: out[i] = omath::c_sin(x[i]); break;
        case 3: out[i] = omath::c_exp(x[i]); break;
        case 4: out[i] = omath::c_atan2(x[i], arg2[i]); break;
        default: out[i] = omath::c_pow(x[i], arg2[i]); break;
        }
    }
}
ORACLE_API void oracle_sample_texture(const OracleScene* scene, uint32_t idx, float u, float v, float* rgb)
{
    RenderData rd{}; Ctx c{scene, &rd, 0, 0};
    float3 t = sampleTexture(c, idx, float2{u, v});
    rgb[0] = t.x; rgb[1] = t.y; rgb[2] = t.z;
}
ORACLE_API void oracle_sample_environment(const OracleScene* scene, const float* d, float* rgb)
{
    RenderData rd{}; Ctx c{scene, &rd, 0, 0};
    float3 t = getEnvironmentLight(c, f3(d));
    rgb[0] = t.x; rgb[1] = t.y; rgb[2] = t.z;
}

/* ================================================================ host side
 * Restatement of the reference's CPU feed.  AABB helpers: Utilities.h:63-101. */
namespace {

struct HAABB {
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    float mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    void growTo(const float* p) { for (int k = 0; k < 3; k++) { mn[k] = std::min(p[k], mn[k]); mx[k] = std::max(p[k], mx[k]); } }
    float getArea() const
    {
        float ex = mx[0] - mn[0], ey = mx[1] - mn[1], ez = mx[2] - mn[2];
        return ex * ey + ey * ez + ez * ex;
    }
};
/* glm::min(a,b) = (b < a) ? b : a ; glm::max(a,b) = (a < b) ? b : a — std::min/max agree for non-NaN */

/* Utilities.h:87-93 */
bool aabbIntersects(const HAABB& a, const HAABB& b)
{
    for (int k = 0; k < 3; k++) {
        float aToB = std::fabs((b.mx[k] + b.mn[k]) * 0.5f - (a.mx[k] + a.mn[k]) * 0.5f);
        float ext = (a.mx[k] - a.mn[k]) * 0.5f + (b.mx[k] - b.mn[k]) * 0.5f;
        if (!(aToB < ext)) return false;
    }
    return true;
}

struct BNode { HAABB bb; std::vector<uint32_t> tris; uint32_t firstChild = UINT_MAX_; };

/* BvhBuilder.cpp:38-70 */
float evaluateSAH(const Triangle* tris, const std::vector<float>& mids, const BNode& node, uint32_t axis, float pos)
{
    HAABB leftBox, rightBox;
    uint32_t leftCount = 0, rightCount = 0;
    for (uint32_t triIndex : node.tris) {
        const Triangle& t = tris[triIndex];
        if (mids[triIndex * 3 + axis] < pos) {
            leftCount++;
            leftBox.growTo(t.position[0]); leftBox.growTo(t.position[1]); leftBox.growTo(t.position[2]);
        } else {
            rightCount++;
            rightBox.growTo(t.position[0]); rightBox.growTo(t.position[1]); rightBox.growTo(t.position[2]);
        }
    }
    float cost = leftCount * leftBox.getArea() + rightCount * rightBox.getArea();
    return cost > 0.f ? cost : FLT_MAX;
}

/* glm::mix(x, y, a) for floats = x * (1 - a) + y * a */
inline float glmMix(float x, float y, float a) { return x * (1.0f - a) + y * a; }

} // namespace

/* BvhBuilder::buildTree / buildTreeInternal, BvhBuilder.cpp:12-36,99-207.
 * outNodes capacity >= 2*numTris+1.  Node order = allocation order; children adjacent.
 * Leaves get [triStart,triEnd) in leaf order as RayTracer.cpp:172-199 assigns them (mesh-local),
 * outTriOrder[k] = source triangle index stored at leaf slot k. */
ORACLE_API uint32_t oracle_build_bvh(const Triangle* tris, uint32_t numTris, const float* rootMin, const float* rootMax,
                                     uint32_t maxDepth, uint32_t maxLeafTriangles,
                                     BvhNode* outNodes, uint32_t* outTriOrder)
{
    std::vector<BNode> nodes;
    nodes.emplace_back();
    for (int k = 0; k < 3; k++) { nodes[0].bb.mn[k] = rootMin[k]; nodes[0].bb.mx[k] = rootMax[k]; }
    nodes[0].tris.resize(numTris);
    for (uint32_t i = 0; i < numTris; i++) nodes[0].tris[i] = i;

    /* OkayMath::getMiddle, SMath.h:9-12: (p0 + p1 + p2) * (1/3) */
    std::vector<float> mids((size_t)numTris * 3);
    for (uint32_t i = 0; i < numTris; i++)
        for (int k = 0; k < 3; k++)
            mids[(size_t)i * 3 + k] = ((tris[i].position[0][k] + tris[i].position[1][k]) + tris[i].position[2][k]) * (1.f / 3.f);

    struct NodeStack { uint32_t nodeIndex, depth; };
    std::vector<NodeStack> stack;
    stack.push_back({0u, 0u});

    while (!stack.empty()) {
        NodeStack nd = stack.back();
        stack.pop_back();
        uint32_t nodeNumTris = (uint32_t)nodes[nd.nodeIndex].tris.size();
        if (nd.depth >= maxDepth - 1u || nodeNumTris <= maxLeafTriangles)
            continue;

        /* findBestSplitPlane, BvhBuilder.cpp:72-97 */
        uint32_t axis = 0; float splitPos = 0.f;
        float bestCost = FLT_MAX;
        {
            const BNode& node = nodes[nd.nodeIndex];
            for (uint32_t a = 0; a < 3; a++) {
                float boundsMin = node.bb.mn[a], boundsMax = node.bb.mx[a];
                if (boundsMin == boundsMax) continue;
                for (uint32_t i = 1; i < 100; i++) {
                    float candidatePos = glmMix(boundsMin, boundsMax, i / (float)100);
                    float cost = evaluateSAH(tris, mids, node, a, candidatePos);
                    if (cost < bestCost) { splitPos = candidatePos; axis = a; bestCost = cost; }
                }
            }
        }
        float parentCost = nodeNumTris * nodes[nd.nodeIndex].bb.getArea();
        if (bestCost >= parentCost)
            continue;

        std::vector<uint32_t> left, right;
        for (uint32_t triIndex : nodes[nd.nodeIndex].tris) {
            if (mids[(size_t)triIndex * 3 + axis] < splitPos) left.push_back(triIndex);
            else right.push_back(triIndex);
        }
        if (left.empty() || right.empty())
            continue;

        nodes.emplace_back();
        nodes.emplace_back();
        uint32_t first = (uint32_t)nodes.size() - 2u;
        nodes[nd.nodeIndex].firstChild = first;
        nodes[first].tris = left;
        nodes[first + 1].tris = right;
        /* BvhBuilder::findAABB, BvhBuilder.cpp:209-223 */
        for (uint32_t ch = first; ch < first + 2; ch++)
            for (uint32_t ti : nodes[ch].tris)
                for (int k = 0; k < 3; k++) nodes[ch].bb.growTo(tris[ti].position[k]);

        stack.push_back({first, nd.depth + 1u});
        stack.push_back({first + 1, nd.depth + 1u});
    }

    /* flatten: RayTracer.cpp:172-199 (mesh-local indices) */
    uint32_t localTriStart = 0;
    for (uint32_t k = 0; k < (uint32_t)nodes.size(); k++) {
        BvhNode& g = outNodes[k];
        for (int a = 0; a < 3; a++) { g.boundingBox.min[a] = nodes[k].bb.mn[a]; g.boundingBox.max[a] = nodes[k].bb.mx[a]; }
        g.triStart = g.triEnd = UINT_MAX_;
        g.firstChildIdx = nodes[k].firstChild;
        if (nodes[k].firstChild != UINT_MAX_) continue;
        g.triStart = localTriStart;
        g.triEnd = localTriStart + (uint32_t)nodes[k].tris.size();
        for (uint32_t ti : nodes[k].tris) outTriOrder[localTriStart++] = ti;
    }
    return (uint32_t)nodes.size();
}

/* ------------------------------------------------- glm 0.9.9.9 restated
 * Column-major 4x4: m[c][r].  Only what Components.h / RayTracer.cpp call. */
namespace {
struct M4 { float m[4][4]; };

M4 m4Identity() { M4 r; std::memset(&r, 0, sizeof(r)); for (int i = 0; i < 4; i++) r.m[i][i] = 1.f; return r; }

/* glm operator*(mat4, mat4): Result[c] = A[0]*B[c][0] + A[1]*B[c][1] + A[2]*B[c][2] + A[3]*B[c][3] */
M4 m4Mul(const M4& A, const M4& B)
{
    M4 R;
    for (int c = 0; c < 4; c++)
        for (int r = 0; r < 4; r++)
            R.m[c][r] = ((A.m[0][r] * B.m[c][0] + A.m[1][r] * B.m[c][1]) + A.m[2][r] * B.m[c][2]) + A.m[3][r] * B.m[c][3];
    return R;
}

/* glm::translate(mat4(1), v): Result[3] = m[0]*v0 + m[1]*v1 + m[2]*v2 + m[3] */
M4 m4Translate(const float* v)
{
    M4 m = m4Identity(), R = m;
    for (int r = 0; r < 4; r++)
        R.m[3][r] = ((m.m[0][r] * v[0] + m.m[1][r] * v[1]) + m.m[2][r] * v[2]) + m.m[3][r];
    return R;
}
M4 m4Scale(const float* v)
{
    M4 m = m4Identity(), R;
    for (int r = 0; r < 4; r++) { R.m[0][r] = m.m[0][r] * v[0]; R.m[1][r] = m.m[1][r] * v[1]; R.m[2][r] = m.m[2][r] * v[2]; R.m[3][r] = m.m[3][r]; }
    return R;
}

struct Quat { float w, x, y, z; };
/* glm::qua(vec3 eulerAngle) (glm/detail/type_quat.inl): c = cos(e*0.5), s = sin(e*0.5) */
Quat quatFromEuler(const float* e)
{
    float cx = std::cos(e[0] * 0.5f), cy = std::cos(e[1] * 0.5f), cz = std::cos(e[2] * 0.5f);
    float sx = std::sin(e[0] * 0.5f), sy = std::sin(e[1] * 0.5f), sz = std::sin(e[2] * 0.5f);
    Quat q;
    q.w = cx * cy * cz + sx * sy * sz;
    q.x = sx * cy * cz - cx * sy * sz;
    q.y = cx * sy * cz + sx * cy * sz;
    q.z = cx * cy * sz - sx * sy * cz;
    return q;
}
/* glm::mat3_cast (glm/gtc/quaternion.inl) */
void quatToMat3(const Quat& q, float R[3][3])
{
    float qxx = q.x * q.x, qyy = q.y * q.y, qzz = q.z * q.z;
    float qxz = q.x * q.z, qxy = q.x * q.y, qyz = q.y * q.z;
    float qwx = q.w * q.x, qwy = q.w * q.y, qwz = q.w * q.z;
    R[0][0] = 1.f - 2.f * (qyy + qzz); R[0][1] = 2.f * (qxy + qwz);       R[0][2] = 2.f * (qxz - qwy);
    R[1][0] = 2.f * (qxy - qwz);       R[1][1] = 1.f - 2.f * (qxx + qzz); R[1][2] = 2.f * (qyz + qwx);
    R[2][0] = 2.f * (qxz + qwy);       R[2][1] = 2.f * (qyz - qwx);       R[2][2] = 1.f - 2.f * (qxx + qyy);
}
inline float radiansf(float deg) { return deg * 0.01745329251994329576923690768489f; }

/* Transform::calculateMatrix, Components.h:13-18 */
M4 transformMatrix(const float* position, const float* rotationDeg, const float* scale)
{
    float e[3] = {radiansf(rotationDeg[0]), radiansf(rotationDeg[1]), radiansf(rotationDeg[2])};
    float R3[3][3];
    quatToMat3(quatFromEuler(e), R3);
    M4 R = m4Identity();
    for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) R.m[c][r] = R3[c][r];
    return m4Mul(m4Mul(m4Translate(position), R), m4Scale(scale));
}

/* glm::inverse(mat4) — compute_inverse<4,4,float> in glm/detail/func_matrix.inl */
M4 m4Inverse(const M4& M)
{
    const float (*m)[4] = M.m;
    float Coef00 = m[2][2] * m[3][3] - m[3][2] * m[2][3];
    float Coef02 = m[1][2] * m[3][3] - m[3][2] * m[1][3];
    float Coef03 = m[1][2] * m[2][3] - m[2][2] * m[1][3];
    float Coef04 = m[2][1] * m[3][3] - m[3][1] * m[2][3];
    float Coef06 = m[1][1] * m[3][3] - m[3][1] * m[1][3];
    float Coef07 = m[1][1] * m[2][3] - m[2][1] * m[1][3];
    float Coef08 = m[2][1] * m[3][2] - m[3][1] * m[2][2];
    float Coef10 = m[1][1] * m[3][2] - m[3][1] * m[1][2];
    float Coef11 = m[1][1] * m[2][2] - m[2][1] * m[1][2];
    float Coef12 = m[2][0] * m[3][3] - m[3][0] * m[2][3];
    float Coef14 = m[1][0] * m[3][3] - m[3][0] * m[1][3];
    float Coef15 = m[1][0] * m[2][3] - m[2][0] * m[1][3];
    float Coef16 = m[2][0] * m[3][2] - m[3][0] * m[2][2];
    float Coef18 = m[1][0] * m[3][2] - m[3][0] * m[1][2];
    float Coef19 = m[1][0] * m[2][2] - m[2][0] * m[1][2];
    float Coef20 = m[2][0] * m[3][1] - m[3][0] * m[2][1];
    float Coef22 = m[1][0] * m[3][1] - m[3][0] * m[1][1];
    float Coef23 = m[1][0] * m[2][1] - m[2][0] * m[1][1];

    float Fac0[4] = {Coef00, Coef00, Coef02, Coef03};
    float Fac1[4] = {Coef04, Coef04, Coef06, Coef07};
    float Fac2[4] = {Coef08, Coef08, Coef10, Coef11};
    float Fac3[4] = {Coef12, Coef12, Coef14, Coef15};
    float Fac4[4] = {Coef16, Coef16, Coef18, Coef19};
    float Fac5[4] = {Coef20, Coef20, Coef22, Coef23};

    float Vec0[4] = {m[1][0], m[0][0], m[0][0], m[0][0]};
    float Vec1[4] = {m[1][1], m[0][1], m[0][1], m[0][1]};
    float Vec2[4] = {m[1][2], m[0][2], m[0][2], m[0][2]};
    float Vec3[4] = {m[1][3], m[0][3], m[0][3], m[0][3]};

    float Inv0[4], Inv1[4], Inv2[4], Inv3[4];
    for (int i = 0; i < 4; i++) {
        Inv0[i] = (Vec1[i] * Fac0[i] - Vec2[i] * Fac1[i]) + Vec3[i] * Fac2[i];
        Inv1[i] = (Vec0[i] * Fac0[i] - Vec2[i] * Fac3[i]) + Vec3[i] * Fac4[i];
        Inv2[i] = (Vec0[i] * Fac1[i] - Vec1[i] * Fac3[i]) + Vec3[i] * Fac5[i];
        Inv3[i] = (Vec0[i] * Fac2[i] - Vec1[i] * Fac4[i]) + Vec2[i] * Fac5[i];
    }
    const float SignA[4] = {+1, -1, +1, -1};
    const float SignB[4] = {-1, +1, -1, +1};
    M4 Inverse;
    for (int i = 0; i < 4; i++) {
        Inverse.m[0][i] = Inv0[i] * SignA[i];
        Inverse.m[1][i] = Inv1[i] * SignB[i];
        Inverse.m[2][i] = Inv2[i] * SignA[i];
        Inverse.m[3][i] = Inv3[i] * SignB[i];
    }
    float Row0[4] = {Inverse.m[0][0], Inverse.m[1][0], Inverse.m[2][0], Inverse.m[3][0]};
    float Dot0[4] = {m[0][0] * Row0[0], m[0][1] * Row0[1], m[0][2] * Row0[2], m[0][3] * Row0[3]};
    float Dot1 = (Dot0[0] + Dot0[1]) + (Dot0[2] + Dot0[3]);
    float OneOverDeterminant = 1.f / Dot1;
    for (int c = 0; c < 4; c++) for (int r = 0; r < 4; r++) Inverse.m[c][r] *= OneOverDeterminant;
    return Inverse;
}

void storeTransposed(const M4& M, float* out16)
{
    /* memory of glm::transpose(M) (column-major) == row-major M */
    for (int r = 0; r < 4; r++) for (int c = 0; c < 4; c++) out16[r * 4 + c] = M.m[c][r];
}

void v3normalize(float* v)
{
    /* glm::normalize = v * inversesqrt(dot(v,v)), inversesqrt(x) = 1/sqrt(x) */
    float d = (v[0] * v[0] + v[1] * v[1]) + v[2] * v[2];
    float inv = 1.f / std::sqrt(d);
    v[0] *= inv; v[1] *= inv; v[2] *= inv;
}
void v3cross(const float* x, const float* y, float* o)
{
    o[0] = x[1] * y[2] - y[1] * x[2];
    o[1] = x[2] * y[0] - y[2] * x[0];
    o[2] = x[0] * y[1] - y[0] * x[1];
}
float v3dot(const float* a, const float* b) { return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]; }
} // namespace

ORACLE_API void oracle_transform_matrix(const float* position, const float* rotationDeg, const float* scale, float* outRowMajor16)
{
    storeTransposed(transformMatrix(position, rotationDeg, scale), outRowMajor16);
}

/* RayTracer::calculateProjectionData, RayTracer.cpp:268-308 (x64 glm config: LH, depth -1..1).
 * Writes viewPlaneDims, cameraPosition, cameraNearZ, both matrices, right/up dirs into rd. */
ORACLE_API void oracle_calculate_projection_data(const float* camPos, const float* camRotDeg, float fovDeg, float nearZ, float farZ,
                                                 uint32_t width, uint32_t height, RenderData* rd)
{
    const float w = (float)width, h = (float)height;
    const float aspectRatio = w / h;
    const float sideB = nearZ;
    const float vAngleRadians = radiansf(fovDeg * 0.5f);
    const float sideA = std::tan(vAngleRadians) * sideB;
    rd->viewPlaneDims[0] = sideA * 2.f;
    rd->viewPlaneDims[1] = rd->viewPlaneDims[0] / aspectRatio;
    for (int k = 0; k < 3; k++) rd->cameraPosition[k] = camPos[k];
    rd->cameraNearZ = nearZ;

    /* glm::perspectiveFovLH_NO(fov, width, height, zNear, zFar) */
    const float rad = radiansf(fovDeg);
    const float hh = std::cos(0.5f * rad) / std::sin(0.5f * rad);
    const float ww = hh * h / w;
    M4 P; std::memset(&P, 0, sizeof(P));
    P.m[0][0] = ww;
    P.m[1][1] = hh;
    P.m[2][2] = (farZ + nearZ) / (farZ - nearZ);
    P.m[2][3] = 1.f;
    P.m[3][2] = -(2.f * farZ * nearZ) / (farZ - nearZ);
    storeTransposed(m4Inverse(P), rd->cameraInverseProjectionMatrix);

    /* Transform::getForwardVec/RightVec/UpVec, Components.h:20-23 */
    float e[3] = {radiansf(camRotDeg[0]), radiansf(camRotDeg[1]), radiansf(camRotDeg[2])};
    float R3[3][3];
    quatToMat3(quatFromEuler(e), R3);
    float fwd[3] = {R3[2][0], R3[2][1], R3[2][2]};
    float right[3] = {R3[0][0], R3[0][1], R3[0][2]};
    float up[3] = {R3[1][0], R3[1][1], R3[1][2]};
    v3normalize(fwd); v3normalize(right); v3normalize(up);
    for (int k = 0; k < 3; k++) { rd->cameraRightDir[k] = right[k]; rd->cameraUpDir[k] = up[k]; }

    /* glm::lookAtLH(eye, eye + fwd, (0,1,0)) */
    float center[3] = {camPos[0] + fwd[0], camPos[1] + fwd[1], camPos[2] + fwd[2]};
    float f[3] = {center[0] - camPos[0], center[1] - camPos[1], center[2] - camPos[2]};
    v3normalize(f);
    const float upv[3] = {0.f, 1.f, 0.f};
    float s[3]; v3cross(upv, f, s); v3normalize(s);
    float u[3]; v3cross(f, s, u);
    M4 V = m4Identity();
    V.m[0][0] = s[0]; V.m[1][0] = s[1]; V.m[2][0] = s[2];
    V.m[0][1] = u[0]; V.m[1][1] = u[1]; V.m[2][1] = u[2];
    V.m[0][2] = f[0]; V.m[1][2] = f[1]; V.m[2][2] = f[2];
    V.m[3][0] = -v3dot(s, camPos);
    V.m[3][1] = -v3dot(u, camPos);
    V.m[3][2] = -v3dot(f, camPos);
    storeTransposed(m4Inverse(V), rd->cameraInverseViewMatrix);
}

/* ------------------------------------------------------------- oct-tree
 * RayTracer::createOctTree + loadOctTree + refitOctTreeNode, RayTracer.cpp:503-788.
 * Entities arrive in registry-view order: mesh entities first, then sphere entities (:539-577). */
struct OracleMeshEntity { float position[3], rotation[3], scale[3]; Material material; uint32_t meshID; };
struct OracleSphereEntity { float position[3], rotation[3], scale[3]; Material material; float radius; };
struct OracleMeshDesc { uint32_t startIdx, endIdx, bvhTreeStartIdx, numBvhNodes; };

ORACLE_API uint32_t oracle_build_octree(const OracleMeshDesc* descs, const AABB* meshBounds, uint32_t numMeshes,
                                        const OracleMeshEntity* meshEnts, uint32_t numMeshEnts,
                                        const OracleSphereEntity* sphereEnts, uint32_t numSphereEnts,
                                        uint32_t maxDepth, uint32_t maxLeafObjects,
                                        OctTreeNode* outNodes, uint32_t nodeCapacity, Mesh* outInstances, uint32_t* outNumInstances)
{
    (void)numMeshes;
    struct Ent { uint32_t id; HAABB aabb; };           /* id < numMeshEnts: mesh entity, else sphere */
    struct ONode { HAABB bb; std::vector<Ent> entities; uint32_t children[8]; };
    std::vector<ONode> nodes;
    nodes.emplace_back();
    for (int k = 0; k < 8; k++) nodes[0].children[k] = UINT_MAX_;

    for (uint32_t e = 0; e < numMeshEnts; e++) {
        const AABB& mb = meshBounds[meshEnts[e].meshID];
        /* DirectX::BoundingBox::GetCorners: center + extents * (+-1,+-1,+-1); extents are negative
         * here (min - max, RayTracer.cpp:546) — the corner SET is unchanged, the ORDER is not, and
         * growTo is order-independent. */
        float center[3], extents[3];
        for (int k = 0; k < 3; k++) { center[k] = (mb.min[k] + mb.max[k]) * 0.5f; extents[k] = (mb.min[k] - mb.max[k]) * 0.5f; }
        M4 matrix = transformMatrix(meshEnts[e].position, meshEnts[e].rotation, meshEnts[e].scale);
        static const float BOX_OFFSET[8][3] = {{-1, -1, 1}, {1, -1, 1}, {1, 1, 1}, {-1, 1, 1}, {-1, -1, -1}, {1, -1, -1}, {1, 1, -1}, {-1, 1, -1}};
        for (int cI = 0; cI < 8; cI++) {
            /* XMVectorMultiplyAdd(extents, offset, center) — evaluated as extents*offset + center (no FMA in the oracle) */
            float v[3];
            for (int k = 0; k < 3; k++) v[k] = extents[k] * BOX_OFFSET[cI][k] + center[k];
            /* matrix * vec4(v, 1): glm mat4*vec4 = (m[0]*v0 + m[1]*v1) + (m[2]*v2 + m[3]*v3) */
            float p[3];
            for (int r = 0; r < 3; r++)
                p[r] = (matrix.m[0][r] * v[0] + matrix.m[1][r] * v[1]) + (matrix.m[2][r] * v[2] + matrix.m[3][r] * 1.f);
            nodes[0].bb.growTo(p);
        }
        Ent en; en.id = e;
        for (int k = 0; k < 3; k++) { en.aabb.mn[k] = mb.min[k]; en.aabb.mx[k] = mb.max[k]; }   /* untransformed, :562 */
        nodes[0].entities.push_back(en);
    }
    for (uint32_t e = 0; e < numSphereEnts; e++) {
        Ent en; en.id = numMeshEnts + e;
        for (int k = 0; k < 3; k++) {
            en.aabb.mn[k] = -sphereEnts[e].radius + sphereEnts[e].position[k];
            en.aabb.mx[k] = sphereEnts[e].radius + sphereEnts[e].position[k];
        }
        nodes[0].bb.growTo(en.aabb.mn);
        nodes[0].bb.growTo(en.aabb.mx);
        nodes[0].entities.push_back(en);
    }

    static const float CHILD_BBS_OFFSETS[8][3] = {{1, -1, 1}, {-1, -1, 1}, {1, 1, 1}, {-1, 1, 1}, {1, -1, -1}, {-1, -1, -1}, {1, 1, -1}, {-1, 1, -1}};
    struct St { uint32_t nodeIndex, depth; };
    std::vector<St> stack;
    stack.push_back({0u, 0u});
    std::vector<Ent> childEntities, pool;

    while (!stack.empty()) {
        St sd = stack.back();
        stack.pop_back();
        if (sd.depth >= maxDepth || nodes[sd.nodeIndex].entities.size() <= maxLeafObjects)
            continue;

        HAABB def = nodes[sd.nodeIndex].bb;
        float bbCenter[3];
        for (int k = 0; k < 3; k++) bbCenter[k] = (def.mx[k] + def.mn[k]) * 0.5f;
        for (int k = 0; k < 3; k++) {
            def.mx[k] -= bbCenter[k]; def.mn[k] -= bbCenter[k];
            def.mn[k] *= 0.5f; def.mx[k] *= 0.5f;
            def.mx[k] += bbCenter[k]; def.mn[k] += bbCenter[k];
        }

        pool.clear();
        float nodeBBArea = nodes[sd.nodeIndex].bb.getArea();
        uint32_t totalChildren = (uint32_t)nodes[sd.nodeIndex].entities.size();
        for (int k = (int)nodes[sd.nodeIndex].entities.size() - 1; k >= 0; k--)
            if (nodes[sd.nodeIndex].entities[k].aabb.getArea() / nodeBBArea < 0.75f)
                pool.push_back(nodes[sd.nodeIndex].entities[k]);

        float ext[3];
        for (int k = 0; k < 3; k++) ext[k] = (def.mx[k] - def.mn[k]) * 0.5f;
        for (uint32_t i = 0; i < 8u; i++) {
            HAABB childBB = def;
            for (int k = 0; k < 3; k++) {
                childBB.mx[k] += CHILD_BBS_OFFSETS[i][k] * ext[k];
                childBB.mn[k] += CHILD_BBS_OFFSETS[i][k] * ext[k];
            }
            childEntities.clear();
            for (int k = (int)pool.size() - 1; k >= 0; k--) {
                if (!aabbIntersects(childBB, pool[k].aabb))
                    continue;
                childEntities.push_back(pool[k]);
                pool.erase(pool.begin() + k);
            }
            if (childEntities.empty())
                continue;
            if (childEntities.size() == totalChildren)
                continue;

            std::vector<Ent>& pe = nodes[sd.nodeIndex].entities;
            for (const Ent& ce : childEntities)
                for (size_t j = 0; j < pe.size(); j++)
                    if (pe[j].id == ce.id) { pe.erase(pe.begin() + j); break; }

            nodes.emplace_back();
            ONode& child = nodes.back();
            for (int k = 0; k < 8; k++) child.children[k] = UINT_MAX_;
            nodes[sd.nodeIndex].children[i] = (uint32_t)nodes.size() - 1;
            child.entities = childEntities;
            /* refitOctTreeNode, RayTracer.cpp:503-519 */
            child.bb = HAABB();
            for (const Ent& ce : child.entities) { child.bb.growTo(ce.aabb.mn); child.bb.growTo(ce.aabb.mx); }
            stack.push_back({(uint32_t)nodes.size() - 1, sd.depth + 1u});
        }
    }

    /* loadOctTree, RayTracer.cpp:706-788 */
    if ((uint32_t)nodes.size() > nodeCapacity) return UINT_MAX_;
    uint32_t numInst = 0;
    for (uint32_t i = 0; i < (uint32_t)nodes.size(); i++) {
        const ONode& node = nodes[i];
        OctTreeNode& g = outNodes[i];
        for (int k = 0; k < 3; k++) { g.boundingBox.min[k] = node.bb.mn[k]; g.boundingBox.max[k] = node.bb.mx[k]; }
        g.meshesStartIdx = g.meshesEndIdx = g.spheresStartIdx = g.spheresEndIdx = UINT_MAX_;
        g.firstChildIdx = UINT_MAX_;
        g.numChildren = 0;
        for (int k = 0; k < 8; k++) {
            if (node.children[k] == UINT_MAX_) continue;
            if (g.firstChildIdx == UINT_MAX_) g.firstChildIdx = node.children[k];
            g.numChildren += 1;
        }
        uint32_t numMeshesHere = 0;
        for (const Ent& en : node.entities) {
            if (en.id >= numMeshEnts) continue;
            const OracleMeshEntity& me = meshEnts[en.id];
            Mesh& gm = outInstances[numInst++];
            gm.triangleStartIdx = descs[me.meshID].startIdx;
            gm.triangleEndIdx = descs[me.meshID].endIdx;
            gm.boundingBox = meshBounds[me.meshID];
            M4 T = transformMatrix(me.position, me.rotation, me.scale);
            /* transformMatrix = glm::transpose(T); inverseTransformMatrix = glm::inverse(transpose(T)) */
            M4 Tt; for (int c = 0; c < 4; c++) for (int r = 0; r < 4; r++) Tt.m[c][r] = T.m[r][c];
            M4 TtInv = m4Inverse(Tt);
            std::memcpy(gm.transformMatrix, Tt.m, 64);
            std::memcpy(gm.inverseTransformMatrix, TtInv.m, 64);
            gm.material = me.material;
            gm.bvhNodeStartIdx = descs[me.meshID].bvhTreeStartIdx;
            numMeshesHere++;
        }
        g.meshesStartIdx = numInst - numMeshesHere;
        g.meshesEndIdx = g.meshesStartIdx + numMeshesHere;
    }
    *outNumInstances = numInst;
    return (uint32_t)nodes.size();
}

/* ------------------------------------------------------------- textures
 * RayTracer::loadTextureData, RayTracer.cpp:318-324: N = uint32(sqrt(totArea / count)) */
ORACLE_API uint32_t oracle_texture_array_size(const uint32_t* widths, const uint32_t* heights, uint32_t count)
{
    if (!count) return 0;
    uint32_t totArea = 0;
    for (uint32_t i = 0; i < count; i++) totArea += widths[i] * heights[i];
    /* totArea / textures.size() is an integer (size_t) division; glm::sqrt on the integer promotes to double */
    return (uint32_t)std::sqrt((double)(totArea / count));
}

/* Okay::scaleTexture, DX11.cpp:232-270 + ScaleVS/PS: full-target quad, uv = (pixel centre)/(dst size),
 * bilinear CLAMP sample of the source (the default sampler after ClearState) [Q12], UNORM8 re-quantise. */
ORACLE_API void oracle_scale_texture(const uint8_t* src, uint32_t sw, uint32_t sh, uint8_t* dst, uint32_t dw, uint32_t dh)
{
    if (sw == dw && sh == dh) { std::memcpy(dst, src, (size_t)sw * sh * 4); return; }
    for (uint32_t y = 0; y < dh; y++)
        for (uint32_t x = 0; x < dw; x++) {
            float u = ((float)x + 0.5f) / (float)dw, v = ((float)y + 0.5f) / (float)dh;
            float fx = u * (float)sw - 0.5f, fy = v * (float)sh - 0.5f;
            float x0f = std::floor(fx), y0f = std::floor(fy);
            float ax = fx - x0f, ay = fy - y0f;
            int x0 = (int)x0f, y0 = (int)y0f, x1 = x0 + 1, y1 = y0 + 1;
            x0 = std::min(std::max(x0, 0), (int)sw - 1); x1 = std::min(std::max(x1, 0), (int)sw - 1);
            y0 = std::min(std::max(y0, 0), (int)sh - 1); y1 = std::min(std::max(y1, 0), (int)sh - 1);
            for (int ch = 0; ch < 4; ch++) {
                float c00 = src[((size_t)y0 * sw + x0) * 4 + ch] / 255.0f, c10 = src[((size_t)y0 * sw + x1) * 4 + ch] / 255.0f;
                float c01 = src[((size_t)y1 * sw + x0) * 4 + ch] / 255.0f, c11 = src[((size_t)y1 * sw + x1) * 4 + ch] / 255.0f;
                float top = c00 * (1.0f - ax) + c10 * ax, bot = c01 * (1.0f - ax) + c11 * ax;
                dst[((size_t)y * dw + x) * 4 + ch] = toUnorm8(top * (1.0f - ay) + bot * ay);
            }
        }
}

/* RayTracer::loadEnvironmentMap, RayTracer.cpp:352-399: 4x3 horizontal cross -> +X,-X,+Y,-Y,+Z,-Z [Q13] */
ORACLE_API void oracle_environment_cross_to_faces(const uint8_t* img, uint32_t imgW, uint32_t imgH, uint8_t* faces)
{
    const uint32_t width = imgW / 4u, height = imgH / 3u;
    const uint32_t* px = (const uint32_t*)img;
    const size_t starts[6] = {
        (size_t)imgW * height + width * 2u,   /* +X */
        (size_t)imgW * height,                /* -X */
        (size_t)width,                        /* +Y */
        (size_t)imgW * height * 2u + width,   /* -Y */
        (size_t)imgW * height + width,        /* +Z */
        (size_t)imgW * height + width * 3u};  /* -Z */
    uint32_t* out = (uint32_t*)faces;
    for (int f = 0; f < 6; f++)
        for (uint32_t i = 0; i < height; i++)
            std::memcpy(out + ((size_t)f * height + i) * width, px + starts[f] + (size_t)i * imgW, (size_t)width * 4);
}

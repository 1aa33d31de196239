/*
 * okrt_kernels.cuh — sm_100a device code of the path tracer.
 *
 * Replaces resources/shaders/RaytracerCS.hlsl + GPU-Utilities.hlsli.  Arithmetic contract: every
 * float32 operation that decides a result (slab test, Moeller-Trumbore, world-distance conversion,
 * hit resolve, BSDF sampling, accumulation) is evaluated with the reference's expression order and
 * individually rounded (-fmad=false), so that a hit is (triangle, u, v, distance) bit-for-bit what
 * the scalar restatement of the HLSL computes.  What is free to change — and does — is the ORDER in
 * which boxes and triangles are visited and the hierarchy used for culling; see traverseFast().
 */
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#ifdef OKRT_DEBUG_CHECKS
#include <cassert>
#define OKRT_CHECK(c) assert(c)
#else
#define OKRT_CHECK(c) ((void)0)
#endif

#include "okrt.h"
#include "okrt_accel.h"
#include "okrt_math.cuh"

/* functions the CPU check of the wide traversal (okrt_trace_rays_host) shares with the kernels */
#define OKRT_HD __host__ __device__ __forceinline__

namespace okrt {

/* ------------------------------------------------------------------ shim */
struct V3 { float x, y, z; };
struct V2 { float x, y; };

OKRT_HD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
OKRT_HD V3 v3(const float* p) { return v3(p[0], p[1], p[2]); }
OKRT_HD V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
OKRT_HD V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
OKRT_HD V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
OKRT_HD V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
OKRT_HD V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
OKRT_HD V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
OKRT_HD float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
OKRT_HD V3 cross(V3 a, V3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
OKRT_HD float length(V3 a) { return sqrtf(dot(a, a)); }
OKRT_HD V3 normalize(V3 a) { const float inv = 1.0f / sqrtf(dot(a, a)); return a * inv; }
__device__ __forceinline__ float lerp1(float a, float b, float s) { return a + s * (b - a); }
__device__ __forceinline__ V3 lerp3(V3 a, V3 b, float s) { return v3(lerp1(a.x, b.x, s), lerp1(a.y, b.y, s), lerp1(a.z, b.z, s)); }
__device__ __forceinline__ V3 reflect(V3 i, V3 n) { const float d2 = 2.0f * dot(i, n); return i - n * d2; }
__device__ __forceinline__ float saturate(float x) { return fminf(fmaxf(x, 0.0f), 1.0f); }

#define OKRT_FLT_MAX 3.402823466e+38F
#define OKRT_PI 3.14159265f

/* [Q1] the 16 floats are row-major M; mul(float4(v,w), M_hlsl) == M * v */
__device__ __forceinline__ V3 mulPoint(V3 v, const float* m)
{
    return v3(((v.x * m[0] + v.y * m[1]) + v.z * m[2]) + m[3],
              ((v.x * m[4] + v.y * m[5]) + v.z * m[6]) + m[7],
              ((v.x * m[8] + v.y * m[9]) + v.z * m[10]) + m[11]);
}
__device__ __forceinline__ V3 mulDir(V3 v, const float* m)
{
    return v3((v.x * m[0] + v.y * m[1]) + v.z * m[2],
              (v.x * m[4] + v.y * m[5]) + v.z * m[6],
              (v.x * m[8] + v.y * m[9]) + v.z * m[10]);
}
OKRT_HD V3 mulPointRows(V3 v, float4 r0, float4 r1, float4 r2)
{
    return v3(((v.x * r0.x + v.y * r0.y) + v.z * r0.z) + r0.w,
              ((v.x * r1.x + v.y * r1.y) + v.z * r1.z) + r1.w,
              ((v.x * r2.x + v.y * r2.y) + v.z * r2.z) + r2.w);
}
OKRT_HD V3 mulDirRows(V3 v, float4 r0, float4 r1, float4 r2)
{
    return v3((v.x * r0.x + v.y * r0.y) + v.z * r0.z,
              (v.x * r1.x + v.y * r1.y) + v.z * r1.z,
              (v.x * r2.x + v.y * r2.y) + v.z * r2.z);
}

/* ------------------------------------------------------------ scene view */
struct DeviceScene {
    /* reference-format buffers (binding table t0..t10) */
    const float* triPos;                      /* t0: 9 floats per triangle */
    const float* triInfo;                     /* t1: 33 floats per triangle */
    const okrt_bvh_node* bvhNodes;            /* t2 */
    cudaTextureObject_t textures;             /* t3: layered 2D, uchar4, point fetch */
    cudaTextureObject_t environment;          /* t4: 6-layer 2D, uchar4, point fetch */
    const okrt_octree_node* octNodes;         /* t5 */
    const okrt_sphere* spheres;               /* t6 */
    const okrt_mesh_instance* instances;      /* t7 */
    const okrt_directional_light* dirLights;  /* t8 */
    const okrt_point_light* pointLights;      /* t9 */
    const okrt_spot_light* spotLights;        /* t10 */
    uint32_t numTris, numBvhNodes, numOctNodes, numSpheres, numInstances, numDirLights, numPointLights, numSpotLights;
    uint32_t texSize, texLayers, envW, envH;
    float texInvSize;                          /* 1 / texSize */
    /* fast layout */
    const float4* fastNodes;
    const float4* fastTris;
    const float4* fastInstances;
    const uint32_t* octParent;
    uint32_t topRef;
    uint32_t numFastNodes, numFastInstances;
    uint32_t pad0;
    cudaTextureObject_t fastTrisTex;          /* fastTris seen through the texture pipe (float4 texels) */
    /* wide layout (okrt_accel.h): 6 uint4 per node, 2 uint4 per leaf record */
    const uint4* wideNodes;
    const uint4* wideLeaves;
    uint32_t wideRoot, numWideNodes;
    float maxSlack;
    uint32_t pad1;
    /* shading record, 64 B per triangle (4 float4), packed on the device from t1 when the triangles are set: what every mesh hit
     * reads (3 vertex normals, 3 uvs) in two sectors instead of six; tangents / bitangents stay in t1 for normal-mapped hits.
     *   {n0.xyz, uv0.x} {n1.xyz, uv0.y} {n2.xyz, uv1.x} {uv1.y, uv2.x, uv2.y, 0} */
    const float4* shadeInfo;
};

struct FrameParams {
    okrt_render_data rd;                      /* b0 */
    uint32_t maxBounces;
    uint32_t seedFrame;                       /* numAccumulationFrames as used by the seed [Q3] (first frame of a batch) */
    uint32_t seedStride, batchFrames;         /* wavefront: frames traced together; frame f uses seedFrame + f*seedStride */
    uint32_t traversal;
    uint32_t coveredW, coveredH;              /* [Q9] */
    uint32_t bandRows, shardRank, shardCount; /* tile sharding: row band b belongs to rank b % shardCount */
    const uint32_t* ownedBands;               /* tile sharding with an explicit table (okrt_set_shard_bands): global band of this rank's k-th band */
    uint32_t tilesX, numTiles;                /* 16x8 tiles this launch owns */
    uint32_t allCovered;                      /* every pixel slot of those tiles is a covered pixel [Q9]: bounce 0 needs no queue */
    uint32_t* workCounter;                    /* next tile to hand out (zeroed before the launch) */
    float4* accumulation;                     /* u1 */
    uchar4* result;                           /* u0 */
    okrt_aov* aov;                            /* optional */
    unsigned long long* counters;             /* [0] segments, [1] fallback segments, [2] pixels */
};

/* ------------------------------------------------------------------ portable intrinsics (host + device) */
OKRT_HD uint32_t wBits(float f)
{
#ifdef __CUDA_ARCH__
    return __float_as_uint(f);
#else
    uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}
OKRT_HD float wFloat(uint32_t u)
{
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}
OKRT_HD float wFma(float a, float b, float c)
{
#ifdef __CUDA_ARCH__
    return __fmaf_rn(a, b, c);
#else
    return fmaf(a, b, c);
#endif
}
OKRT_HD float wMax3(float a, float b, float c)
{
#ifdef __CUDA_ARCH__
    float d; asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d;      /* FMNMX3 */
#else
    return fmaxf(fmaxf(a, b), c);
#endif
}
OKRT_HD float wMin3(float a, float b, float c)
{
#ifdef __CUDA_ARCH__
    float d; asm("min.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d;
#else
    return fminf(fminf(a, b), c);
#endif
}
OKRT_HD uint32_t wPopc(uint32_t x)
{
#ifdef __CUDA_ARCH__
    return (uint32_t)__popc(x);
#else
    return (uint32_t)__builtin_popcount(x);
#endif
}
OKRT_HD uint32_t wHighBit(uint32_t x)       /* x != 0 */
{
#ifdef __CUDA_ARCH__
    return 31u - (uint32_t)__clz((int)x);
#else
    return 31u - (uint32_t)__builtin_clz(x);
#endif
}
OKRT_HD uint32_t wLowBit(uint32_t x)        /* x != 0 */
{
#ifdef __CUDA_ARCH__
    return (uint32_t)__ffs((int)x) - 1u;
#else
    return (uint32_t)__builtin_ctz(x);
#endif
}
/* 32768 + byte `k` of `w` as a float: one PRMT (the byte lands in bits 8..15 of 0x47000000 = 32768.0f, whose ulp is 2^-8) */
template <int k>
OKRT_HD float wPlane(uint32_t w)
{
#ifdef __CUDA_ARCH__
    return __uint_as_float(__byte_perm(w, 0x47000000u, 0x7404 | (k << 4)));
#else
    return wFloat(0x47000000u | (((w >> (8 * k)) & 0xFFu) << 8));
#endif
}
/* one 256-bit read-only load; p is 32-byte aligned */
OKRT_HD void wLoad256(const uint4* p, uint4& a, uint4& b)
{
#ifdef __CUDA_ARCH__
    asm("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w), "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w) : "l"(p));
#else
    a = p[0]; b = p[1];
#endif
}
OKRT_HD float4 wLdg(const float4* p)
{
#ifdef __CUDA_ARCH__
    return __ldg(p);
#else
    return *p;
#endif
}
OKRT_HD void wTriangle(const DeviceScene& s, uint32_t j, float4& a, float4& b, float4& c)
{
#ifdef __CUDA_ARCH__
    const int tj = (int)(j * kTriQuads);             /* through the texture pipe: the LSU data stage is the busier unit */
    a = tex1Dfetch<float4>(s.fastTrisTex, tj); b = tex1Dfetch<float4>(s.fastTrisTex, tj + 1); c = tex1Dfetch<float4>(s.fastTrisTex, tj + 2);
#else
    const float4* tp = s.fastTris + (size_t)j * kTriQuads;
    a = tp[0]; b = tp[1]; c = tp[2];
#endif
}

/* tile sharding: image tile row of this rank's local tile row (its k-th band is global band k*shardCount + shardRank, or the
 * k-th entry of the explicit band table) */
__device__ __forceinline__ uint32_t ownedTileRow(const FrameParams& fp, uint32_t tileRowLocal)
{
    const uint32_t tilesPerBand = fp.bandRows / 8u;
    const uint32_t k = tileRowLocal / tilesPerBand;
    const uint32_t band = fp.ownedBands ? fp.ownedBands[k] : k * fp.shardCount + fp.shardRank;
    return band * tilesPerBand + tileRowLocal % tilesPerBand;
}

/* ------------------------------------------------------------------- RNG */
__device__ __forceinline__ uint32_t pcgHash(uint32_t& seed)           /* GPU-Utilities.hlsli:18-25 [F7] */
{
    seed *= 747796405u + 2891336453u;
    seed = ((seed >> ((seed >> 28u) + 4u)) ^ seed) * 277803737u;
    seed = (seed >> 22u) ^ seed;
    return seed;
}
__device__ __forceinline__ float randomFloat(uint32_t& seed)          /* :51-54; (float)UINT_MAX == 2^32 */
{
    return __uint2float_rn(pcgHash(seed)) / 4294967296.0f;
}
__device__ __forceinline__ float randomNormal(uint32_t& seed)         /* :56-62 */
{
    const float theta = (2 * OKRT_PI) * randomFloat(seed);
    const float rho = sqrtf(-2 * okm::log_(randomFloat(seed)));
    return rho * okm::cos_(theta);
}
__device__ __forceinline__ V3 getRandomVector(uint32_t& seed)         /* :64-70 */
{
    const float x = randomNormal(seed);
    const float y = randomNormal(seed);
    const float z = randomNormal(seed);
    return normalize(v3(x, y, z));
}
__device__ __forceinline__ V2 randomPointInCircle(uint32_t& seed)     /* :72-78 */
{
    const float angle = randomFloat(seed) * 2.f * OKRT_PI;
    const float c = okm::cos_(angle), s = okm::sin_(angle);
    const float r = sqrtf(randomFloat(seed));
    V2 o; o.x = c * r; o.y = s * r;
    return o;
}
__device__ __forceinline__ float clampedDot(V3 v, V3 u) { return fmaxf(dot(v, u), 0.f); }

__device__ __forceinline__ V3 refractCustom(V3 direction, V3 normal, float refractionRatio)   /* :85-93 */
{
    const float cosTheta = fminf(dot(-direction, normal), 1.f);
    const V3 perp = refractionRatio * (direction + cosTheta * normal);
    const float lenSq = dot(perp, perp);
    const V3 parallel = -sqrtf(fabsf(1.f - lenSq)) * normal;
    return perp + parallel;
}

/* ----------------------------------------------------------- intersection */
OKRT_HD float rayAndSphere(V3 o, V3 d, V3 c, float radius)   /* :98-114 [Q15] */
{
    const V3 rayToSphere = c - o;
    const float distToClosestPoint = dot(rayToSphere, d);
    const float magSq = dot(rayToSphere, rayToSphere);
    const float rSq = radius * radius;
    if (distToClosestPoint < 0.f && magSq > rSq) return -1.f;
    const float sideA = magSq - distToClosestPoint * distToClosestPoint;
    if (sideA > rSq) return -1.f;
    const float sideB = sqrtf(rSq - sideA);
    return magSq > rSq ? distToClosestPoint - sideB : distToClosestPoint + sideB;
}

/* :117-155 [Q6] */
__device__ __forceinline__ float rayAndTriangle(V3 o, V3 d, V3 p0, V3 p1, V3 p2, float& u, float& v)
{
    const float EPSILON = 0.000001f;
    const V3 E1 = p0 - p2, E2 = p1 - p2;
    const V3 cross1 = cross(d, E2);
    const float determinant = dot(E1, cross1);
    if (determinant < EPSILON && determinant > -EPSILON) return -1.f;
    const float inverseDet = 1.f / determinant;
    const V3 rayMoved = o - p2;
    u = dot(rayMoved, cross1) * inverseDet;
    if (u < -EPSILON) return -1.f;
    const V3 cross2 = cross(rayMoved, E1);
    v = dot(d, cross2) * inverseDet;
    if (v < -EPSILON || u + v > 1.0f) return -1.f;
    const float t1 = dot(E2, cross2) * inverseDet;
    if (t1 < -EPSILON) return -1.f;
    return t1;
}

/* :183-197 with 1/direction hoisted (same value the shader recomputes per box) */
OKRT_HD float rayAndAABBDist(V3 o, V3 idir, V3 bmin, V3 bmax)
{
    const V3 tMin = (bmin - o) * idir;
    const V3 tMax = (bmax - o) * idir;
    const float distFar = fminf(fminf(fmaxf(tMin.x, tMax.x), fmaxf(tMin.y, tMax.y)), fmaxf(tMin.z, tMax.z));
    const float distNear = fmaxf(fmaxf(fminf(tMin.x, tMax.x), fminf(tMin.y, tMax.y)), fminf(tMin.z, tMax.z));
    const bool didHit = distFar >= distNear && distFar > 0.f;
    return didHit ? distNear : OKRT_FLT_MAX;
}
OKRT_HD V3 rcp3(V3 d) { return v3(1.f / d.x, 1.f / d.y, 1.f / d.z); }

/* --------------------------------------------------------------- textures */
/* (float)k / 255.0f for k = 0..255, correctly rounded, without the IEEE division sequence: one Newton step on
 * q0 = k * RN(1/255) with fused multiply-adds.  Equal to the division for all 256 inputs (checked exhaustively in
 * exact rational arithmetic, tests/test_cpu_oracle.py::test_unorm8_division_by_refinement). */
__device__ __forceinline__ float unorm8ToFloat(unsigned char k)
{
    const float x = (float)k, r = __uint_as_float(0x3B808081u);
    const float q0 = __fmul_rn(x, r);
    return __fmaf_rn(__fmaf_rn(-255.0f, q0, x), r, q0);
}
__device__ __forceinline__ V3 texelRgb(uchar4 t) { return v3(unorm8ToFloat(t.x), unorm8ToFloat(t.y), unorm8ToFloat(t.z)); }
/* floor-mod of an integer-valued float by the texture size (WRAP addressing): the exact remainder, through a
 * float quotient estimate that is off by at most one below 2^22 */
__device__ __forceinline__ int wrapIndex(float xf, int N, float invN)
{
    const int xi = (int)xf;
    if (fabsf(xf) < 4194304.0f) {
        int r = xi - __float2int_rd(xf * invN) * N;
        if (r < 0) r += N;
        if (r >= N) r -= N;
        return r;
    }
    int r = xi % N;
    return r < 0 ? r + N : r;
}
__device__ __forceinline__ bool finitef_(float x) { return x - x == 0.0f; }
__device__ __forceinline__ V3 bilerp(V3 c00, V3 c10, V3 c01, V3 c11, float fx, float fy)
{
    const float wx0 = 1.0f - fx, wy0 = 1.0f - fy;
    const V3 top = c00 * wx0 + c10 * fx;
    const V3 bot = c01 * wx0 + c11 * fx;
    return top * wy0 + bot * fy;
}

/* RaytracerCS.hlsl:133-136 [Q12]: bilinear, WRAP, lod 0.  Four point fetches through the texture
 * unit (uchar4, no hardware filtering) and a float32 lerp: D3D's 8-bit fixed-point filter weights
 * are not reproducible on a CPU, the float lerp is. */
/* the four taps of a uv: every layer of the array has the same size, so all maps of a hit (albedo, roughness, metallic, specular,
 * normal) share them — computed once per hit, not once per map */
struct TexTaps { float cx0, cx1, cy0, cy1, fx, fy; };
__device__ __forceinline__ TexTaps textureTaps(const DeviceScene& s, V2 uv)
{
    TexTaps t;
    const int N = (int)s.texSize;
    const float u = finitef_(uv.x) ? uv.x : 0.0f, v = finitef_(uv.y) ? uv.y : 0.0f;
    float x = u * (float)N - 0.5f, y = v * (float)N - 0.5f;
    if (!(fabsf(x) < 1.0e9f)) x = 0.0f;
    if (!(fabsf(y) < 1.0e9f)) y = 0.0f;
    const float xf = floorf(x), yf = floorf(y);
    t.fx = x - xf; t.fy = y - yf;
    const int x0 = wrapIndex(xf, N, s.texInvSize), y0 = wrapIndex(yf, N, s.texInvSize);
    const int x1 = x0 + 1 == N ? 0 : x0 + 1;
    const int y1 = y0 + 1 == N ? 0 : y0 + 1;
    t.cx0 = (float)x0 + 0.5f; t.cx1 = (float)x1 + 0.5f; t.cy0 = (float)y0 + 0.5f; t.cy1 = (float)y1 + 0.5f;
    return t;
}
struct TexQuad { uchar4 t00, t10, t01, t11; };
__device__ __forceinline__ TexQuad fetchQuad(const DeviceScene& s, uint32_t textureIdx, const TexTaps& t)
{
    const int layer = (int)(textureIdx < s.texLayers ? textureIdx : s.texLayers - 1u);
    TexQuad q;
    q.t00 = tex2DLayered<uchar4>(s.textures, t.cx0, t.cy0, layer);
    q.t10 = tex2DLayered<uchar4>(s.textures, t.cx1, t.cy0, layer);
    q.t01 = tex2DLayered<uchar4>(s.textures, t.cx0, t.cy1, layer);
    q.t11 = tex2DLayered<uchar4>(s.textures, t.cx1, t.cy1, layer);
    return q;
}
__device__ __forceinline__ V3 filterQuad(const TexQuad& q, const TexTaps& t)
{
    return bilerp(texelRgb(q.t00), texelRgb(q.t10), texelRgb(q.t01), texelRgb(q.t11), t.fx, t.fy);
}
/* RaytracerCS.hlsl:133-136 [Q12]: bilinear, WRAP, lod 0.  Four point fetches through the texture
 * unit (uchar4, no hardware filtering) and a float32 lerp: D3D's 8-bit fixed-point filter weights
 * are not reproducible on a CPU, the float lerp is. */
__device__ __forceinline__ V3 sampleTexture(const DeviceScene& s, uint32_t textureIdx, const TexTaps& t)
{
    if (s.texLayers == 0u) return v3(0.f, 0.f, 0.f);
    return filterQuad(fetchQuad(s, textureIdx, t), t);
}

/* RaytracerCS.hlsl:86-88 [Q13] */
__device__ __forceinline__ V3 environmentLight(const DeviceScene& s, V3 d)
{
    if (s.envW == 0u || s.envH == 0u) return v3(0.f, 0.f, 0.f);
    const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    int face; float sc, tc, ma;
    if (ax >= ay && ax >= az) { ma = ax; if (d.x >= 0) { face = 0; sc = -d.z; tc = -d.y; } else { face = 1; sc = d.z; tc = -d.y; } }
    else if (ay >= az)        { ma = ay; if (d.y >= 0) { face = 2; sc = d.x; tc = d.z; } else { face = 3; sc = d.x; tc = -d.z; } }
    else                      { ma = az; if (d.z >= 0) { face = 4; sc = d.x; tc = -d.y; } else { face = 5; sc = -d.x; tc = -d.y; } }
    if (!(ma > 0.0f) || !finitef_(ma) || !finitef_(sc) || !finitef_(tc)) return v3(0.f, 0.f, 0.f);
    const float u = (sc / ma + 1.0f) * 0.5f, v = (tc / ma + 1.0f) * 0.5f;
    const int W = (int)s.envW, H = (int)s.envH;
    const float x = u * (float)W - 0.5f, y = v * (float)H - 0.5f;
    const float xf = floorf(x), yf = floorf(y);
    const float fx = x - xf, fy = y - yf;
    int x0 = (int)xf, y0 = (int)yf, x1 = x0 + 1, y1 = y0 + 1;
    x0 = min(max(x0, 0), W - 1); x1 = min(max(x1, 0), W - 1);
    y0 = min(max(y0, 0), H - 1); y1 = min(max(y1, 0), H - 1);
    const uchar4 t00 = tex2DLayered<uchar4>(s.environment, (float)x0 + 0.5f, (float)y0 + 0.5f, face);
    const uchar4 t10 = tex2DLayered<uchar4>(s.environment, (float)x1 + 0.5f, (float)y0 + 0.5f, face);
    const uchar4 t01 = tex2DLayered<uchar4>(s.environment, (float)x0 + 0.5f, (float)y1 + 0.5f, face);
    const uchar4 t11 = tex2DLayered<uchar4>(s.environment, (float)x1 + 0.5f, (float)y1 + 0.5f, face);
    return bilerp(texelRgb(t00), texelRgb(t10), texelRgb(t01), texelRgb(t11), fx, fy);
}

/* ------------------------------------------------------------ hit record */
struct Hit {
    float distance;        /* payload.distance */
    uint32_t kind;         /* 0 miss, 1 sphere, 2 mesh */
    uint32_t idx;          /* sphere or instance index (reference order) */
    uint32_t tri;
    float u, v;
};

struct Surface {           /* what shading needs from Payload (RaytracerCS.hlsl:10-17) */
    okrt_material material;
    V3 worldPosition;
    V3 worldNormal;
};

/* ------------------------------------------- reference-order traversal
 * RaytracerCS.hlsl:228-334 [Q4,Q5], visit order and counters exactly as the shader. */
/* `limit` = FLT_MAX gives the reference's traversal verbatim.  The fallback of the near-first traversal passes
 * best*(1+delta) of the distance it found: boxes that start at or beyond it hold no candidate inside the window
 * and cannot influence which window candidate the reference returns (order independence), so they
 * are skipped — the traversal then costs a few dozen boxes instead of a few hundred. */
__device__ __noinline__ void traverseExact(const DeviceScene& s, uint32_t numSpheres, V3 o, V3 d, Hit& hit,
                                           uint32_t& bbCheckCount, uint32_t& triCheckCount, float limit = OKRT_FLT_MAX)
{
    hit.distance = OKRT_FLT_MAX; hit.kind = 0u; hit.idx = 0xFFFFFFFFu; hit.tri = 0xFFFFFFFFu; hit.u = 0.f; hit.v = 0.f;

    for (uint32_t i = 0; i < numSpheres; i++) {
        const okrt_sphere& sp = s.spheres[i];
        const float t = rayAndSphere(o, d, v3(sp.position), sp.radius);
        if (t > 0.f && t < hit.distance) { hit.distance = t; hit.idx = i; hit.kind = 1u; }
    }

    uint32_t octStack[kExactOctStackCapacity];
    uint32_t bvhStack[kExactBvhStackCapacity];
    uint32_t octStackSize = s.numOctNodes ? 1u : 0u;
    octStack[0] = 0u;
    const V3 idirWorld = rcp3(d);

    while (octStackSize > 0u) {
        const uint32_t octIdx = octStack[--octStackSize];
        const okrt_octree_node& on = s.octNodes[octIdx];
        bbCheckCount += 1u;
        const float octNear = rayAndAABBDist(o, idirWorld, v3(on.boundingBox.min), v3(on.boundingBox.max));
        if (octNear >= hit.distance || octNear >= limit)
            continue;

        const uint32_t mStart = on.meshesStartIdx, mEnd = on.meshesEndIdx;
        for (uint32_t i = mStart; i < mEnd; i++) {
            bbCheckCount += 1u;
            const okrt_mesh_instance& inst = s.instances[i];
            uint32_t bvhStackSize = 1u;
            bvhStack[0] = inst.bvhNodeStartIdx;
            const V3 lo = mulPoint(o, inst.inverseTransformMatrix);
            const V3 ld = normalize(mulDir(d, inst.inverseTransformMatrix));
            const V3 lidir = rcp3(ld);

            while (bvhStackSize > 0u) {
                const uint32_t nodeIdx = bvhStack[--bvhStackSize];
                const okrt_bvh_node& bn = s.bvhNodes[nodeIdx];
                bbCheckCount += 1u;
                const float boxNear = rayAndAABBDist(lo, lidir, v3(bn.boundingBox.min), v3(bn.boundingBox.max));
                if (boxNear >= hit.distance || boxNear >= limit)
                    continue;
                const uint32_t first = bn.firstChildIdx;
                if (first != 0xFFFFFFFFu) {
                    bvhStack[bvhStackSize++] = first;
                    bvhStack[bvhStackSize++] = first + 1u;
                    continue;
                }
                const uint32_t tEnd = bn.triEnd;
                for (uint32_t j = bn.triStart; j < tEnd; j++) {
                    const float* tp = s.triPos + (size_t)j * 9;
                    float bu = 0.f, bv = 0.f;
                    triCheckCount += 1u;
                    float t = rayAndTriangle(lo, ld, v3(tp), v3(tp + 3), v3(tp + 6), bu, bv);
                    if (t <= 0.f) continue;
                    t = length(mulDir(ld * t, inst.transformMatrix));     /* RaytracerCS.hlsl:312-314 */
                    if (t < hit.distance) { hit.distance = t; hit.idx = i; hit.kind = 2u; hit.tri = j; hit.u = bu; hit.v = bv; }
                }
            }
        }
        for (uint32_t c = 0; c < on.numChildren; c++)
            octStack[octStackSize++] = on.firstChildIdx + c;
    }
}

/* The bounded reference-order traversal for ONE ray, run by a whole warp (traceFallbackKernel).  All lanes hold the
 * same ray and the same state; control flow is uniform except for the instance lists, where lane l prepares instance
 * base + l (local ray, distance to the root box of its BVH — the value the reference's first pop of that instance
 * compares with payload.distance, RaytracerCS.hlsl:285).  The warp then walks the instances whose root box starts
 * inside the window in the reference's order (ascending index), applying the reference's own cull against the
 * current payload.distance in that order.  An oct-tree node lists ~70 instances on the bench scene and a few of
 * them survive: the 70 dependent (instance record -> root node) fetches of the scalar form become 3 parallel rounds. */
__device__ __noinline__ void traverseExactWarp(const DeviceScene& s, uint32_t numSpheres, V3 o, V3 d, Hit& hit, float limit)
{
    const uint32_t lane = threadIdx.x & 31u;
    hit.distance = OKRT_FLT_MAX; hit.kind = 0u; hit.idx = 0xFFFFFFFFu; hit.tri = 0xFFFFFFFFu; hit.u = 0.f; hit.v = 0.f;
    for (uint32_t i = 0; i < numSpheres; i++) {
        const okrt_sphere& sp = s.spheres[i];
        const float t = rayAndSphere(o, d, v3(sp.position), sp.radius);
        if (t > 0.f && t < hit.distance) { hit.distance = t; hit.idx = i; hit.kind = 1u; }
    }
    uint32_t octStack[kExactOctStackCapacity];
    uint32_t bvhStack[kExactBvhStackCapacity];
    uint32_t octStackSize = s.numOctNodes ? 1u : 0u;
    octStack[0] = 0u;
    const V3 idirWorld = rcp3(d);
    while (octStackSize > 0u) {
        const uint32_t octIdx = octStack[--octStackSize];
        const okrt_octree_node& on = s.octNodes[octIdx];
        const float octNear = rayAndAABBDist(o, idirWorld, v3(on.boundingBox.min), v3(on.boundingBox.max));
        if (octNear >= hit.distance || octNear >= limit) continue;
        const uint32_t mStart = on.meshesStartIdx, mEnd = on.meshesEndIdx;
        for (uint32_t base = mStart; base < mEnd; base += 32u) {
            const uint32_t mine = base + lane;
            float rootNear = OKRT_FLT_MAX;
            if (mine < mEnd) {
                const okrt_mesh_instance& inst = s.instances[mine];
                const V3 lo = mulPoint(o, inst.inverseTransformMatrix);
                const V3 lidir = rcp3(normalize(mulDir(d, inst.inverseTransformMatrix)));
                const okrt_bvh_node& root = s.bvhNodes[inst.bvhNodeStartIdx];
                rootNear = rayAndAABBDist(lo, lidir, v3(root.boundingBox.min), v3(root.boundingBox.max));
            }
            uint32_t pending = __ballot_sync(0xFFFFFFFFu, rootNear < limit);
            while (pending) {
                const uint32_t l = (uint32_t)__ffs(pending) - 1u;
                pending &= pending - 1u;
                if (__shfl_sync(0xFFFFFFFFu, rootNear, l) >= hit.distance) continue;      /* the reference's cull at the root pop */
                const uint32_t i = base + l;
                const okrt_mesh_instance& inst = s.instances[i];
                const V3 lo = mulPoint(o, inst.inverseTransformMatrix);
                const V3 ld = normalize(mulDir(d, inst.inverseTransformMatrix));
                const V3 lidir = rcp3(ld);
                uint32_t bvhStackSize = 1u;
                bvhStack[0] = inst.bvhNodeStartIdx;
                while (bvhStackSize > 0u) {
                    const uint32_t nodeIdx = bvhStack[--bvhStackSize];
                    const okrt_bvh_node& bn = s.bvhNodes[nodeIdx];
                    const float boxNear = rayAndAABBDist(lo, lidir, v3(bn.boundingBox.min), v3(bn.boundingBox.max));
                    if (boxNear >= hit.distance || boxNear >= limit) continue;
                    const uint32_t first = bn.firstChildIdx;
                    if (first != 0xFFFFFFFFu) { bvhStack[bvhStackSize++] = first; bvhStack[bvhStackSize++] = first + 1u; continue; }
                    const uint32_t tEnd = bn.triEnd;
                    for (uint32_t j = bn.triStart; j < tEnd; j++) {
                        const float* tp = s.triPos + (size_t)j * 9;
                        float bu = 0.f, bv = 0.f;
                        float t = rayAndTriangle(lo, ld, v3(tp), v3(tp + 3), v3(tp + 6), bu, bv);
                        if (t <= 0.f) continue;
                        t = length(mulDir(ld * t, inst.transformMatrix));     /* RaytracerCS.hlsl:312-314 */
                        if (t < hit.distance) { hit.distance = t; hit.idx = i; hit.kind = 2u; hit.tri = j; hit.u = bu; hit.v = bv; }
                    }
                }
            }
        }
        for (uint32_t c = 0; c < on.numChildren; c++) octStack[octStackSize++] = on.firstChildIdx + c;
    }
}

/* ------------------------------------------------ near-first traversal
 * Same boxes, same triangles, same arithmetic — different visiting order and a different culling
 * hierarchy at the top.  Correctness argument (DESIGN.md "Order independence"): with relaxed culling
 * (a box is skipped only if distNear >= best*(1+delta)) every triangle whose distance is within
 * delta of the final minimum is tested; if exactly one candidate lies in that window the reference's
 * order-dependent culls cannot change the winner, so the result equals the reference's bit for bit.
 * Otherwise (ties, coplanar duplicates, hits on a shared edge) `ambiguous` is set and the caller
 * re-traces the ray with traverseExact. */
#define OKRT_DELTA 1.0e-4f
/* a missed box reports FLT_MAX and must never pass, also while nothing has been hit yet */
OKRT_HD float relaxedLimit(float best) { return fminf(best * (1.0f + OKRT_DELTA), OKRT_FLT_MAX); }

constexpr uint32_t kTraceDone = 0xFFFFFFFDu;
constexpr uint32_t kTraceExitInstance = 0xFFFFFFFEu;
constexpr uint32_t kTraceTop = 0xFFFFFFFFu;

/* ------------------------------------------- near-first traversal, step form
 * The same traversal as traceRound(), cut into STEPS of similar cost so that a warp can choose which kind of step
 * to run next (traceFastKernel, okrt_wavefront.cuh).  ncu on the while-while form (profiles/r1_d_*): 9.2 of 32
 * lanes active per instruction — a round lasts as long as its slowest lane's run of interior nodes, and the
 * per-candidate oct-tree chain test ran with 2.3 lanes.  Here
 *   - one node step = one child-pair node, one triangle step = one triangle of the leaf the lane stands on;
 *   - the stack holds a sentinel at the bottom, so a pop is one unconditional load;
 *   - the oct-tree chain of a merged instance is tested once, for the winning triangle, when the ray finishes
 *     (a failed chain or two candidates inside the window send the ray to the reference-order kernel).
 * Arithmetic per box and per triangle is unchanged, so are the results. */
#ifndef OKRT_STACK_TOPREG
#define OKRT_STACK_TOPREG 0
#endif
#ifndef OKRT_COLD_IN_LOCAL
#define OKRT_COLD_IN_LOCAL 1
#endif
/* OKRT_SMEM_STACK = K > 0 (device code only): stack entries 0..K-1 of a lane live in shared memory, word-interleaved over the
 * block's 128 threads (entry e of thread t at word e*128 + t: every lane owns its bank, a push or pop is one wavefront whatever
 * the lanes' depths are), entries K.. stay in local memory.  In local memory lanes at different depths touch different 128-byte
 * rows: 2-3 wavefronts per push or pop (profiles/r2_j: stack traffic is 28 % of the traversal kernel's L1 wavefronts). */
#ifndef OKRT_SMEM_STACK
#define OKRT_SMEM_STACK 0
#endif
#if OKRT_SMEM_STACK && defined(__CUDA_ARCH__)
#define OKRT_SMEM_STACK_ON 1
#else
#define OKRT_SMEM_STACK_ON 0
#endif
/* OKRT_COLD_IN_LOCAL: the lane's hit record (written once or twice per ray, read when the ray finishes) lives in the slots
 * behind the traversal stack instead of registers, which buys resident warps */
constexpr uint32_t kColdSlots = 8;
constexpr uint32_t kColdHitRec = kFastStackCapacity + 0, kColdHitTri = kFastStackCapacity + 1, kColdHitU = kFastStackCapacity + 2,
                   kColdHitV = kFastStackCapacity + 3, kColdSecond = kFastStackCapacity + 4;
struct StepLane {
    V3 ro, rd, ridir;        /* current ray: identity-local at the top level, instance-local inside an instance */
    float best;
#if !OKRT_COLD_IN_LOCAL
    float second;
    uint32_t hitRec;         /* 1<<30 | sphere, or 2<<30 | instance record */
    uint32_t hitTri;
    float hitU, hitV;
#endif
    uint32_t cur;            /* reference the lane stands on: interior / triangle leaf / instance / exit marker / done */
    uint32_t triNext;        /* next triangle of the leaf in `cur` */
    int sp;                  /* next free stack slot; slot 0 holds kTraceDone */
#if OKRT_STACK_TOPREG
    uint32_t top;            /* newest stack entry, kept in a register: a pop hands it out and re-loads it from memory, so the
                              * load's latency is hidden behind the step that follows instead of stalling it */
#endif
    uint32_t inst;           /* instance record inside a generic instance; kTraceTop at the top level */
    uint32_t path;
#if OKRT_SMEM_STACK_ON
    uint32_t smBase;         /* shared-window address of this thread's entry 0 */
#endif
};
#if OKRT_SMEM_STACK_ON
__device__ __forceinline__ void stackSharedStore(uint32_t addr, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" :: "r"(addr), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t stackSharedLoad(uint32_t addr) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory"); return v; }
#define OKRT_SMEM_STACK_DECLARE(L) __shared__ uint32_t okrtSmemStack[OKRT_SMEM_STACK * 128]; \
    (L).smBase = (uint32_t)__cvta_generic_to_shared(okrtSmemStack + threadIdx.x);
#else
#define OKRT_SMEM_STACK_DECLARE(L)
#endif

OKRT_HD void stepCandidate(StepLane& L, uint32_t* stack, float t, uint32_t rec, uint32_t tri, float u, float v)
{
#if OKRT_COLD_IN_LOCAL
    if (t < L.best) {
        stack[kColdSecond] = wBits(L.best); L.best = t;
        stack[kColdHitRec] = rec; stack[kColdHitTri] = tri; stack[kColdHitU] = wBits(u); stack[kColdHitV] = wBits(v);
    } else if (t < wFloat(stack[kColdSecond])) stack[kColdSecond] = wBits(t);
#else
    if (t < L.best) { L.second = L.best; L.best = t; L.hitRec = rec; L.hitTri = tri; L.hitU = u; L.hitV = v; }
    else if (t < L.second) L.second = t;
#endif
}

OKRT_HD void stepTopRay(StepLane& L, V3 o, V3 d)
{
    const float4 i0 = make_float4(1, 0, 0, 0), i1 = make_float4(0, 1, 0, 0), i2 = make_float4(0, 0, 1, 0);
    L.ro = mulPointRows(o, i0, i1, i2);
    L.rd = normalize(mulDirRows(d, i0, i1, i2));
    L.ridir = rcp3(L.rd);
    L.inst = kTraceTop;
}

/* the lane moves to `ref`; a triangle leaf starts at its first triangle */
OKRT_HD void stepEnter(StepLane& L, uint32_t ref)
{
    L.cur = ref;
    if ((ref & kRefTagMask) == kRefTriLeaf) L.triNext = ref & (kMaxTriIndex - 1u);
}
OKRT_HD void stackPush(StepLane& L, uint32_t* stack, uint32_t v)
{
    OKRT_CHECK(L.sp >= 1 && L.sp < (int)kFastStackCapacity);
#if OKRT_SMEM_STACK_ON
    if (L.sp < OKRT_SMEM_STACK) stackSharedStore(L.smBase + (uint32_t)L.sp * 512u, v); else stack[L.sp] = v;
    L.sp++;
#elif OKRT_STACK_TOPREG
    stack[L.sp++] = L.top; L.top = v;
#else
    stack[L.sp++] = v;
#endif
}
OKRT_HD uint32_t stackPop(StepLane& L, const uint32_t* stack)
{
    OKRT_CHECK(L.sp >= 1 && L.sp <= (int)kFastStackCapacity);
#if OKRT_SMEM_STACK_ON
    --L.sp;
    return L.sp < OKRT_SMEM_STACK ? stackSharedLoad(L.smBase + (uint32_t)L.sp * 512u) : stack[L.sp];
#elif OKRT_STACK_TOPREG
    const uint32_t v = L.top; L.top = stack[--L.sp]; return v;
#else
    return stack[--L.sp];
#endif
}
OKRT_HD void stepPop(StepLane& L, const uint32_t* stack) { stepEnter(L, stackPop(L, stack)); }

OKRT_HD void stepInit(const DeviceScene& s, StepLane& L, uint32_t* stack, uint32_t numSpheres, V3 o, V3 d)
{
    L.best = OKRT_FLT_MAX;
#if OKRT_COLD_IN_LOCAL
    stack[kColdSecond] = wBits(OKRT_FLT_MAX);
    stack[kColdHitRec] = 0x3FFFFFFFu; stack[kColdHitTri] = 0xFFFFFFFFu; stack[kColdHitU] = 0u; stack[kColdHitV] = 0u;
#else
    L.second = OKRT_FLT_MAX;
    L.hitRec = 0x3FFFFFFFu; L.hitTri = 0xFFFFFFFFu; L.hitU = 0.f; L.hitV = 0.f;
#endif
    stack[0] = kTraceDone; L.sp = 1;
#if OKRT_SMEM_STACK_ON
    stackSharedStore(L.smBase, kTraceDone);
#endif
#if OKRT_STACK_TOPREG
    L.top = kTraceDone;
#endif
    stepTopRay(L, o, d);
    for (uint32_t i = 0; i < numSpheres; i++) {                /* RaytracerCS.hlsl:232-242 */
        const okrt_sphere& sp = s.spheres[i];
        const float t = rayAndSphere(o, d, v3(sp.position), sp.radius);
        if (t > 0.f) stepCandidate(L, stack, t, (1u << 30) | i, 0xFFFFFFFFu, 0.f, 0.f);
    }
    stepEnter(L, s.topRef == kRefNone ? kTraceDone : s.topRef);
}

/* GPU-Utilities.hlsli:183-197 for one box, as the three conditions the traversal needs: hit, in front, before `limit` */
OKRT_HD bool boxPasses(V3 o, V3 idir, float4 lo, float4 hi, float limit, float& distNear)
{
    const V3 tMin = (v3(lo.x, lo.y, lo.z) - o) * idir;
    const V3 tMax = (v3(hi.x, hi.y, hi.z) - o) * idir;
    const float distFar = fminf(fminf(fmaxf(tMin.x, tMax.x), fmaxf(tMin.y, tMax.y)), fmaxf(tMin.z, tMax.z));
    distNear = fmaxf(fmaxf(fminf(tMin.x, tMax.x), fminf(tMin.y, tMax.y)), fminf(tMin.z, tMax.z));
    return distFar >= distNear && distFar > 0.f && distNear < limit;
}

#ifndef OKRT_LD256
#define OKRT_LD256 1
#endif
#ifndef OKRT_TRI_TEX
#define OKRT_TRI_TEX 1          /* triangle records through the texture pipe (tex1Dfetch) instead of LDG */
#endif
/* one 256-bit read-only load (LDG.E.ENL2.256.CONSTANT, sm_100+): two quads for one L1 tag lookup instead of two.
 * p must be 32-byte aligned. */
OKRT_HD void ldg256(const float4* p, float4& a, float4& b)
{
#ifndef OKRT_NODE_LD_HINT
#define OKRT_NODE_LD_HINT ""
#endif
#ifdef __CUDA_ARCH__
    asm("ld.global.nc" OKRT_NODE_LD_HINT ".v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
                 : "l"(p));
#else
    a = p[0]; b = p[1];
#endif
}

/* one child-pair node.  Empty children carry NaN boxes (okrt_accel.cpp), which never pass. */
OKRT_HD void stepNode(const DeviceScene& s, StepLane& L, uint32_t* stack)
{
    OKRT_CHECK(L.cur < s.numFastNodes);
    const float4* q = s.fastNodes + (size_t)L.cur * 4;
#if OKRT_LD256
    float4 q0, q1, q2, q3;
    ldg256(q, q0, q1); ldg256(q + 2, q2, q3);
#else
    const float4 q0 = wLdg(q), q1 = wLdg(q + 1), q2 = wLdg(q + 2), q3 = wLdg(q + 3);
#endif
    const float limit = relaxedLimit(L.best);
    float n0, n1;
    const uint32_t r0 = wBits(q0.w), r1 = wBits(q1.w);
    const bool h0 = boxPasses(L.ro, L.ridir, q0, q1, limit, n0);
    const bool h1 = boxPasses(L.ro, L.ridir, q2, q3, limit, n1);
    const bool firstIs1 = h1 && (!h0 || n1 < n0);
    if (h0 && h1) stackPush(L, stack, firstIs1 ? r0 : r1);
    uint32_t nxt = firstIs1 ? r1 : r0;
    if (!(h0 || h1)) nxt = stackPop(L, stack);
    stepEnter(L, nxt);
}

/* GPU-Utilities.hlsli:117-155 [Q6] with E1 = p0 - p2 and E2 = p1 - p2 formed at upload (the same float32 subtraction) */
OKRT_HD float rayAndTriangleEdges(V3 o, V3 d, V3 p2, V3 E1, V3 E2, float& u, float& v)
{
    const float EPSILON = 0.000001f;
    const V3 cross1 = cross(d, E2);
    const float determinant = dot(E1, cross1);
    if (determinant < EPSILON && determinant > -EPSILON) return -1.f;
    const float inverseDet = 1.f / determinant;
    const V3 rayMoved = o - p2;
    u = dot(rayMoved, cross1) * inverseDet;
    if (u < -EPSILON) return -1.f;
    const V3 cross2 = cross(rayMoved, E1);
    v = dot(d, cross2) * inverseDet;
    if (v < -EPSILON || u + v > 1.0f) return -1.f;
    const float t1 = dot(E2, cross2) * inverseDet;
    if (t1 < -EPSILON) return -1.f;
    return t1;
}

/* one triangle of the leaf the lane stands on */
OKRT_HD void stepTriangle(const DeviceScene& s, StepLane& L, uint32_t* stack)
{
    const uint32_t j = L.triNext++;
    OKRT_CHECK(j < s.numTris);
    /* through the texture pipe: the LSU data stage (one 32-byte sector per cycle for divergent lanes) is the busiest unit
     * of this kernel (80 %), the texture pipe is otherwise idle */
    float4 a, b, c;                                              /* p2 | E1 | E2 | owner */
    wTriangle(s, j, a, b, c);
    float bu = 0.f, bv = 0.f;
    float t = rayAndTriangleEdges(L.ro, L.rd, v3(a.x, a.y, a.z), v3(a.w, b.x, b.y), v3(b.z, b.w, c.x), bu, bv);
    if (t > 0.f) {
        V3 w = L.rd * t;                                        /* RaytracerCS.hlsl:312-314; T == I for merged instances */
        uint32_t rec = wBits(c.y);                    /* owner of a merged instance's triangle */
        if (L.inst != kTraceTop) {
            const float4* r = s.fastInstances + (size_t)L.inst * kInstQuads;
            w = mulDirRows(w, wLdg(r + 3), wLdg(r + 4), wLdg(r + 5));
            rec = L.inst;
        }
        t = length(w);
        stepCandidate(L, stack, t, (2u << 30) | rec, j, bu, bv);
    }
    const uint32_t end = (L.cur & (kMaxTriIndex - 1u)) + ((L.cur >> 27) & 7u) + 1u;
    if (L.triNext == end) stepPop(L, stack);
}

/* Oct-tree boxes on the way to an instance, world ray (RaytracerCS.hlsl:262-264).  The reference skips a node, and everything
 * below it, when the box is missed OR entered at or beyond the closest hit so far — and createOctTree's boxes need not contain
 * what they list (untransformed occupant boxes, RayTracer.cpp:562).  All boxes must be HIT; their largest distNear is returned
 * and joins the GATE of whatever is found in the instance: nothing the reference tests earlier may be as close as the gate. */
OKRT_HD bool octChainNear(const DeviceScene& s, V3 o, V3 d, uint32_t octNode, float& maxNear)
{
    const V3 idirWorld = rcp3(d);
    maxNear = 0.f;
    for (uint32_t on = octNode; on != 0xFFFFFFFFu; on = s.octParent[on]) {
        const okrt_octree_node& node = s.octNodes[on];
        const float n = rayAndAABBDist(o, idirWorld, v3(node.boundingBox.min), v3(node.boundingBox.max));
        if (!(n < OKRT_FLT_MAX)) return false;
        maxNear = fmaxf(maxNear, n);
    }
    return true;
}

/* the rare references: entering a generic (non-merged) instance, leaving one */
OKRT_HD void stepMisc(const DeviceScene& s, StepLane& L, uint32_t* stack, V3 o, V3 d)
{
    if (L.cur == kTraceExitInstance) {
        stepTopRay(L, o, d);
    } else {                                                     /* generic instance leaf */
        const uint32_t recIdx = L.cur & ~kRefTagMask;
        OKRT_CHECK(recIdx < s.numFastInstances);
        const float4* rec = s.fastInstances + (size_t)recIdx * kInstQuads;
        const float4 r6 = wLdg(rec + 6), r7 = wLdg(rec + 7);
        float octNear;
        if (octChainNear(s, o, d, wBits(r7.y), octNear)) {
            const float4 i0 = wLdg(rec), i1 = wLdg(rec + 1), i2 = wLdg(rec + 2);
            const V3 lo = mulPointRows(o, i0, i1, i2);
            const V3 ld = normalize(mulDirRows(d, i0, i1, i2));
            const V3 lidir = rcp3(ld);
            const float4 a0 = wLdg(rec + 3), a1 = wLdg(rec + 4), a2 = wLdg(rec + 5);
            /* the mesh's root box (RaytracerCS.hlsl:285 on the first popped node) */
            if (rayAndAABBDist(lo, lidir, v3(a0.w, a1.w, a2.w), v3(r6.x, r6.y, r6.z)) < relaxedLimit(L.best)) {
                stackPush(L, stack, kTraceExitInstance);
                L.ro = lo; L.rd = ld; L.ridir = lidir;
                L.inst = recIdx;
                stepEnter(L, wBits(r7.x));
                return;
            }
        }
    }
    stepPop(L, stack);
}

/* the finished lane's hit; returns true when the reference order has to decide.  `L` stands at the top level
 * again (the exit marker of an instance restores the top ray), so L.ro is the world origin. */
OKRT_HD bool stepResult(const DeviceScene& s, const StepLane& L, const uint32_t* stack, const float4* rayD, Hit& hit, float& seed)
{
#if OKRT_COLD_IN_LOCAL
    const uint32_t hitRec = stack[kColdHitRec];
    const float second = wFloat(stack[kColdSecond]);
    hit.tri = stack[kColdHitTri]; hit.u = wFloat(stack[kColdHitU]); hit.v = wFloat(stack[kColdHitV]);
#else
    const uint32_t hitRec = L.hitRec;
    const float second = L.second;
    hit.tri = L.hitTri; hit.u = L.hitU; hit.v = L.hitV;
#endif
    hit.distance = L.best;
    hit.kind = L.best < OKRT_FLT_MAX ? (hitRec >> 30) : 0u;
    hit.idx = hit.kind ? (hitRec & 0x3FFFFFFFu) : 0xFFFFFFFFu;
    /* GATE of the closest hit W: the reference tests W unless something it tested earlier is at least as close as the distNear
     * of a box on its way to W (it culls on `>= payload.distance`, RaytracerCS.hlsl:264,285).  The leaf box and the BVH boxes
     * above it start before W (checked by the differential stress test, tests/test_gpu_stress.py); the oct-tree boxes need not —
     * createOctTree sorts occupants by their UNtransformed boxes (RayTracer.cpp:562), so a node need not even contain what it
     * lists.  A chain that is entered before W's window ends cannot be culled by anything beyond the window (and what lies in
     * the window is known: `second`).  A chain that is MISSED, or entered BEHIND the window (moved props), is different: the reference never reaches W, or culls
     * its node against something farther that this traversal has already culled against W — only the unbounded reference
     * order can decide. */
    bool chainMissed = false;
    if (hit.kind == 2u) {
        const uint32_t rec = hit.idx;
        OKRT_CHECK(rec < s.numFastInstances);
        const float4* r = s.fastInstances + (size_t)rec * kInstQuads;
        const float4 r6 = wLdg(r + 6), r7 = wLdg(r + 7), lo = wLdg(r + 8), hi = wLdg(r + 9);
        hit.idx = wBits(r7.z);
        /* An origin strictly inside the box of an identity instance's node hits it with distNear <= 0 — both slabs of every
         * axis have opposite signs, |1/d| >= 1/2 for the unit directions the shader produces — and the boxes above are nested. */
        const float m = 1.0e-20f;
        const bool inside = (wBits(r6.w) & 1u) != 0u && (wBits(r7.w) & 1u) != 0u &&
                            L.ro.x - lo.x > m && hi.x - L.ro.x > m && L.ro.y - lo.y > m && hi.y - L.ro.y > m && L.ro.z - lo.z > m && hi.z - L.ro.z > m;
        if (!inside) {
            const float4 d = *rayD;
            float octNear;
            if (!octChainNear(s, L.ro, v3(d.x, d.y, d.z), wBits(r7.y), octNear) || octNear >= relaxedLimit(L.best)) chainMissed = true;
        }
    }
    /* two candidates inside the window above the closest hit, or a box the reference never gets through: the reference order
     * decides, searching up to the window (everything the reference could return lies inside it) or without a bound */
    seed = chainMissed ? OKRT_FLT_MAX : L.best;
    return chainMissed || (hit.kind != 0u && second < relaxedLimit(L.best));
}

/* whole-ray form of the near-first traversal over the child-pair layout, used by the megakernel and by the CPU check
 * (okrt_trace_rays_host): the same steps, one ray per thread */
OKRT_HD bool traversePairs(const DeviceScene& s, uint32_t numSpheres, V3 o, V3 d, Hit& hit, float& seed)
{
    uint32_t stack[kFastStackCapacity + kColdSlots];
    StepLane L;
    L.path = 0u;
    OKRT_SMEM_STACK_DECLARE(L)
    stepInit(s, L, stack, numSpheres, o, d);
    while (L.cur != kTraceDone) {
        const uint32_t tag = L.cur & kRefTagMask;
        if (tag == kRefInterior) stepNode(s, L, stack);
        else if (tag == kRefTriLeaf) stepTriangle(s, L, stack);
        else stepMisc(s, L, stack, o, d);
    }
    const float4 dd = make_float4(d.x, d.y, d.z, 0.f);
    return stepResult(s, L, stack, &dd, hit, seed);
}

/* returns true when the reference order has to decide; `limit` bounds that traversal */
OKRT_HD bool traverseFast(const DeviceScene& s, uint32_t numSpheres, V3 o, V3 d, Hit& hit, float& limit)
{
    float seed;
    const bool ambiguous = traversePairs(s, numSpheres, o, d, hit, seed);
    limit = relaxedLimit(seed);
    return ambiguous;
}

} // namespace okrt
#include "okrt_wide.cuh"          /* the same traversal over 8-wide quantised nodes (OKRT_TRAVERSAL_WIDE) */
namespace okrt {

/* ------------------------------------------------------------ hit resolve
 * RaytracerCS.hlsl:336-392 [Q8] */
__device__ __forceinline__ void findMaterialTextureColours(const DeviceScene& s, okrt_material& m, const TexTaps& taps)   /* :138-165 */
{
    if (s.texLayers == 0u) {                                   /* no array bound: a map reads as (0, 0, 0) */
        if (m.albedoTextureIdx != 0xFFFFFFFFu) { m.albedo[0] = 0.f; m.albedo[1] = 0.f; m.albedo[2] = 0.f; }
        if (m.roughnessTextureIdx != 0xFFFFFFFFu) m.roughness = fminf(fmaxf(0.f, m.roughness), 1.f);
        if (m.metallicTextureIdx != 0xFFFFFFFFu) m.metallic = fminf(fmaxf(0.f, m.metallic), 1.f);
        if (m.specularTextureIdx != 0xFFFFFFFFu) m.specular = fminf(fmaxf(0.f, m.specular), 1.f);
        return;
    }
    if (m.albedoTextureIdx != 0xFFFFFFFFu) {
        const V3 t = sampleTexture(s, m.albedoTextureIdx, taps);
        m.albedo[0] = t.x; m.albedo[1] = t.y; m.albedo[2] = t.z;
    }
    if (m.roughnessTextureIdx != 0xFFFFFFFFu)
        m.roughness = fminf(fmaxf(sampleTexture(s, m.roughnessTextureIdx, taps).x, m.roughness), 1.f);
    if (m.metallicTextureIdx != 0xFFFFFFFFu)
        m.metallic = fminf(fmaxf(sampleTexture(s, m.metallicTextureIdx, taps).x, m.metallic), 1.f);
    if (m.specularTextureIdx != 0xFFFFFFFFu)
        m.specular = fminf(fmaxf(sampleTexture(s, m.specularTextureIdx, taps).x, m.specular), 1.f);
}

__device__ __forceinline__ V3 sampleNormalMap(const DeviceScene& s, uint32_t textureIdx, const TexTaps& taps, V3 normal, V3 tangent, V3 bitangent)   /* :167-175 */
{
    V3 sn = sampleTexture(s, textureIdx, taps);
    sn = sn * 2.f - v3(1.f, 1.f, 1.f);
    const V3 r = v3((sn.x * tangent.x + sn.y * bitangent.x) + sn.z * normal.x,
                    (sn.x * tangent.y + sn.y * bitangent.y) + sn.z * normal.y,
                    (sn.x * tangent.z + sn.y * bitangent.z) + sn.z * normal.z);
    return normalize(r);
}

__device__ __forceinline__ void loadMaterial(okrt_material& m, const okrt_material* src)
{
    const uint32_t* p = reinterpret_cast<const uint32_t*>(src);
    uint32_t* q = reinterpret_cast<uint32_t*>(&m);
#pragma unroll
    for (int k = 0; k < 17; k++) q[k] = __ldg(p + k);
}

__device__ __forceinline__ void resolveHit(const DeviceScene& s, V3 o, V3 d, const Hit& hit, Surface& sf)
{
    sf.worldPosition = o + d * hit.distance;
    if (hit.kind == 1u) {                                     /* sphere, :342-362 */
        const okrt_sphere& sp = s.spheres[hit.idx];
        loadMaterial(sf.material, &sp.material);
        const V3 hitNormal = normalize(sf.worldPosition - v3(sp.position));
        V2 uv;                                                /* normalToUV :77-84 */
        uv.x = okm::atan2_(hitNormal.x, hitNormal.z) / (2 * OKRT_PI) + 0.5f;
        uv.y = hitNormal.y * 0.5f + 0.5f;
        const TexTaps taps = textureTaps(s, uv);
        if (sf.material.normalMapIdx != 0xFFFFFFFFu) {
            const V3 tangent = normalize(cross(v3(0.f, 1.f, 0.f), hitNormal));
            const V3 bitangent = normalize(cross(tangent, hitNormal));
            sf.worldNormal = sampleNormalMap(s, sf.material.normalMapIdx, taps, hitNormal, tangent, bitangent);
        } else sf.worldNormal = hitNormal;
        findMaterialTextureColours(s, sf.material, taps);
    } else {                                                  /* mesh, :364-391 */
        const okrt_mesh_instance& inst = s.instances[hit.idx];
        loadMaterial(sf.material, &inst.material);
        const float* ti = s.triInfo + (size_t)hit.tri * 33;   /* VertexInfo: normal 0-2, uv 3-4, tangent 5-7, bitangent 8-10 */
        const float4* si = s.shadeInfo + (size_t)hit.tri * 4;  /* bit copies of the same floats (DeviceScene::shadeInfo) */
        const float4 s0 = __ldg(si), s1 = __ldg(si + 1), s2 = __ldg(si + 2), s3 = __ldg(si + 3);
        const float bu = hit.u, bv = hit.v, bw = 1.f - (hit.u + hit.v);
        V2 uv;
        uv.x = (s0.w * bu + s2.w * bv) + s3.y * bw;
        uv.y = (s1.w * bu + s3.x * bv) + s3.z * bw;
        const V3 n0 = v3(s0.x, s0.y, s0.z), n1 = v3(s1.x, s1.y, s1.z), n2 = v3(s2.x, s2.y, s2.z);
        V3 normal = normalize(mulDir((n0 * bu + n1 * bv) + n2 * bw, inst.transformMatrix));
        const TexTaps taps = textureTaps(s, uv);
        if (sf.material.normalMapIdx != 0xFFFFFFFFu) {
            const V3 g0 = v3(__ldg(ti + 5), __ldg(ti + 6), __ldg(ti + 7)), g1 = v3(__ldg(ti + 16), __ldg(ti + 17), __ldg(ti + 18)), g2 = v3(__ldg(ti + 27), __ldg(ti + 28), __ldg(ti + 29));
            const V3 b0 = v3(__ldg(ti + 8), __ldg(ti + 9), __ldg(ti + 10)), b1 = v3(__ldg(ti + 19), __ldg(ti + 20), __ldg(ti + 21)), b2 = v3(__ldg(ti + 30), __ldg(ti + 31), __ldg(ti + 32));
            const V3 tangent = normalize(mulDir((g0 * bu + g1 * bv) + g2 * bw, inst.transformMatrix));
            const V3 bitangent = normalize(mulDir((b0 * bu + b1 * bv) + b2 * bw, inst.transformMatrix));
            sf.worldNormal = sampleNormalMap(s, sf.material.normalMapIdx, taps, normal, tangent, bitangent);
        } else sf.worldNormal = normal;
        findMaterialTextureColours(s, sf.material, taps);
    }
}

/* ------------------------------------------------------------- shading */
__device__ __forceinline__ float reflectance(float cosine, float reflectionIdx)     /* :100-105 [Q7] */
{
    float r0 = (1.f - reflectionIdx) / (1.f + reflectionIdx);
    r0 *= r0;
    return r0 + (1.f + r0) * okm::pow_(1.f - cosine, 5.f);
}

__device__ __forceinline__ V3 findReflectDirection(V3 direction, V3 normal, float roughness, uint32_t& seed)   /* :107-112 */
{
    const V3 diffuse = normalize(normal + getRandomVector(seed));
    const V3 specular = reflect(direction, normal);
    return normalize(lerp3(specular, diffuse, roughness));
}

__device__ __forceinline__ V3 findTransparencyBounce(V3 direction, V3 normal, float refractionIdx, uint32_t& seed)   /* :114-131 [Q7] */
{
    const bool hitFrontFace = dot(direction, normal) < 0.f;
    const float refractionRatio = hitFrontFace ? 1.f / refractionIdx : refractionIdx;
    if (!hitFrontFace) normal = normal * -1.f;
    const float cos_theta = dot(-direction, normal);
    const float sin_theta = sqrtf(1.f - cos_theta * cos_theta);
    const bool cannot_refract = refractionRatio * sin_theta > 1.f;
    const float refl = reflectance(cos_theta, refractionRatio);
    const float rnd = randomFloat(seed);                      /* always consumed: no short-circuit in HLSL */
    if (cannot_refract || refl > rnd) direction = reflect(direction, normal);
    return refractCustom(direction, normal, refractionRatio);
}

/* :397-418 [Q11] */
__device__ __forceinline__ V3 getLighting(V3 o, V3 d, bool hitAny, float hitDistance, float transparency, V3 worldNormal,
                                          V3 lightPos, float lightRadius, V3 lightColour, float lightIntensity)
{
    const float dist = rayAndSphere(o, d, lightPos, lightRadius);
    if (dist < 0.f) return v3(0.f, 0.f, 0.f);
    float modifier = 1.f;
    if (hitAny && hitDistance > dist) modifier *= clampedDot(d, -worldNormal);
    else if (hitAny && hitDistance < dist) modifier *= transparency * clampedDot(d, -worldNormal);
    return (lightColour * lightIntensity) * modifier;
}

__device__ __forceinline__ uchar4 toUnorm8(float r, float g, float b, float a)
{
    return make_uchar4((unsigned char)(saturate(r) * 255.0f + 0.5f), (unsigned char)(saturate(g) * 255.0f + 0.5f),
                       (unsigned char)(saturate(b) * 255.0f + 0.5f), (unsigned char)(saturate(a) * 255.0f + 0.5f));
}

/* ------------------------------------------------------ one path segment
 * Everything RaytracerCS::main does between two findClosestHit calls (RaytracerCS.hlsl:448-513): lights,
 * miss -> environment, lobe pick, bounce.  Shared by the wavefront shade kernel and the megakernel. */
struct PathState {
    V3 o, d, light, contribution;
    uint32_t seed, pixel, bounce;       /* pixel = y * width + x */
};

__device__ __forceinline__ bool shadeSegment(const DeviceScene& s, const okrt_render_data& rd, uint32_t maxBounces, const Hit& hit, PathState& ps)
{
    const V3 o = ps.o, d = ps.d;
    const uint32_t i = ps.bounce;
    const bool hitAny = hit.kind != 0u;
    Surface sf;
    if (hitAny) resolveHit(s, o, d, hit, sf);
    else { sf.worldNormal = v3(0.f, 0.f, 0.f); sf.worldPosition = v3(0.f, 0.f, 0.f); sf.material.transparency = 0.f; }

    V3 light = ps.light, contribution = ps.contribution;
    /* lights :448-487 [Q11] */
    const uint32_t nPoint = rd.numPointLights < s.numPointLights ? rd.numPointLights : s.numPointLights;
    for (uint32_t p = 0u; p < nPoint; p++) {
        const okrt_point_light& pl = s.pointLights[p];
        const V3 l = getLighting(o, d, hitAny, hit.distance, sf.material.transparency, sf.worldNormal, v3(pl.position), pl.radius, v3(pl.colour), pl.intensity);
        light = light + l * contribution;
    }
    const uint32_t nSpot = rd.numSpotLights < s.numSpotLights ? rd.numSpotLights : s.numSpotLights;
    for (uint32_t q = 0u; q < nSpot; q++) {
        const okrt_spot_light& sl = s.spotLights[q];
        const V3 rayToLight = normalize((v3(sl.position) + getRandomVector(ps.seed) * 0.01f) - o);
        const float cosTheta = dot(rayToLight, -v3(sl.direction));
        if (cosTheta < sl.maxAngle) continue;
        const V3 l = getLighting(o, d, hitAny, hit.distance, sf.material.transparency, sf.worldNormal, v3(sl.position), sl.radius, v3(sl.colour), sl.intensity);
        light = light + l * contribution;
    }
    const uint32_t nDir = rd.numDirLights < s.numDirLights ? rd.numDirLights : s.numDirLights;
    for (uint32_t q = 0u; q < nDir; q++) {
        const okrt_directional_light& dl = s.dirLights[q];
        if (clampedDot(d, -v3(dl.direction)) < dl.effectiveAngle) continue;
        float modifier = 1.f;
        if (hitAny && i) modifier *= sf.material.transparency * clampedDot(d, -sf.worldNormal);
        else if (!i) modifier = 0.f;
        light = light + ((v3(dl.colour) * dl.intensity) * contribution) * modifier;
    }

    bool finished;
    if (!hitAny) {                                     /* :489-493 */
        light = light + environmentLight(s, d) * contribution;
        finished = true;
    } else {
        const okrt_material& m = sf.material;          /* :495-513 [Q10] */
        const float metallicFactor = (m.metallic * (1.f - m.roughness)) >= randomFloat(ps.seed) ? 1.f : 0.f;
        const float specularFactor = (m.specular * (1.f - m.roughness)) >= randomFloat(ps.seed) ? 1.f : 0.f;
        const float transparencyFactor = m.transparency >= randomFloat(ps.seed) ? 1.f : 0.f;
        const V3 reflectDir = findReflectDirection(d, sf.worldNormal, m.roughness * (1.f - specularFactor), ps.seed);
        const V3 refractDir = findTransparencyBounce(d, sf.worldNormal, m.indexOfRefraction, ps.seed);
        const V3 bounceDir = normalize(lerp3(reflectDir, refractDir, transparencyFactor));
        const V3 hitPoint = sf.worldPosition + bounceDir * 0.001f;
        contribution = contribution * lerp3(v3(m.albedo), v3(1.f, 1.f, 1.f), metallicFactor);
        light = light + ((v3(m.emissionColour) * m.emissionPower) * contribution) * (1.f - transparencyFactor);
        ps.o = hitPoint; ps.d = bounceDir;
        ps.bounce = i + 1u;
        finished = ps.bounce > maxBounces;
    }
    ps.light = light; ps.contribution = contribution;
    return finished;
}

/* --------------------------------------------------------------- kernel
 * RaytracerCS::main (RaytracerCS.hlsl:421-537) as a persistent kernel.  The shader maps one thread to one
 * pixel and lets it loop over its bounces; paths end after 1..maxBounces+1 segments, so most lanes of a
 * warp idle while the longest path finishes.  Here a lane owns a PATH for as long as it lives and picks
 * the next pixel as soon as it ends ("path regeneration"): warps stay full until the image runs out.
 * Pixels are handed out in 16x8 tiles (one atomic per 128 pixels), lanes of a warp are refilled from the
 * warp's current tile by ballot + prefix popcount.  Per-pixel results do not depend on which lane traced
 * them: the seed is a function of (x, y, frame) only [Q3]. */
constexpr int kTileW = 16, kTileH = 8;
constexpr int kTilePixels = kTileW * kTileH;
constexpr int kBlockThreads = 128;

template <bool kInstrumented>
__global__ void __launch_bounds__(kBlockThreads)
pathTraceKernel(const __grid_constant__ DeviceScene s, const __grid_constant__ FrameParams fp)
{
    const okrt_render_data& rd = fp.rd;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t laneMaskLt = (1u << lane) - 1u;
    const uint32_t width = rd.textureDims[0];
    const uint32_t numSpheres = rd.numSpheres < s.numSpheres ? rd.numSpheres : s.numSpheres;
    const bool exactMode = fp.traversal == OKRT_TRAVERSAL_EXACT;

    uint32_t slotNext = 0u, slotEnd = 0u;     /* warp-uniform: unassigned pixel slots of the warp's current tile */
    bool exhausted = false;                   /* warp-uniform: the tile queue is empty */
    bool alive = false;
    PathState ps;
    ps.o = ps.d = ps.light = ps.contribution = v3(0.f, 0.f, 0.f); ps.seed = 0u; ps.pixel = 0u; ps.bounce = 0u;
    uint32_t segments = 0u, fallbacks = 0u, pixels = 0u;
    /* instrumentation (AOV tap, debug heat maps): only in the kInstrumented build of the kernel */
    uint32_t bbCheckCount = 0u, triCheckCount = 0u, pathSegments = 0u, firstFlags = 0u;
    Hit first; first.kind = 0u; first.idx = 0xFFFFFFFFu; first.tri = 0xFFFFFFFFu; first.u = first.v = 0.f; first.distance = OKRT_FLT_MAX;

    for (;;) {
        /* ---- refill dead lanes */
        uint32_t dead = __ballot_sync(0xFFFFFFFFu, !alive);
        while (dead != 0u && !exhausted) {
            if (slotNext == slotEnd) {
                uint32_t tile = 0u;
                if (lane == 0u) tile = atomicAdd(fp.workCounter, 1u);
                tile = __shfl_sync(0xFFFFFFFFu, tile, 0);
                if (tile >= fp.numTiles) { exhausted = true; break; }
                slotNext = tile * kTilePixels; slotEnd = slotNext + kTilePixels;
            }
            const uint32_t want = __popc(dead), avail = slotEnd - slotNext;
            const uint32_t take = want < avail ? want : avail;
            const uint32_t rank = __popc(dead & laneMaskLt);
            if (!alive && rank < take) {
                const uint32_t slot = slotNext + rank;
                const uint32_t tile = slot / kTilePixels, j = slot % kTilePixels;
                const uint32_t tileX = tile % fp.tilesX, tileRowLocal = tile / fp.tilesX;
                const uint32_t tileRow = ownedTileRow(fp, tileRowLocal);
                const uint32_t px = tileX * kTileW + j % kTileW, py = tileRow * kTileH + j / kTileW;
                if (px < fp.coveredW && py < fp.coveredH) {
                    /* [Q3] seed, createRay :177-206 [Q2], applyDOF :208-216 */
                    uint32_t seed = px + (py + 74813u) * width * (fp.seedFrame + 1u);
                    V3 pos = v3((float)px, (float)(rd.textureDims[1] - py), rd.cameraNearZ);
                    pos.x += randomFloat(seed) - 0.5f;
                    pos.y += randomFloat(seed) - 0.5f;
                    pos.x /= (float)rd.textureDims[0]; pos.y /= (float)rd.textureDims[1];
                    pos.x *= 2.f; pos.y *= 2.f;
                    pos.x -= 1.f; pos.y -= 1.f;
                    const V3 target = mulPoint(pos, rd.cameraInverseProjectionMatrix);
                    V3 o = v3(rd.cameraPosition);
                    V3 d = mulDir(normalize(v3(target.x / target.z, target.y / target.z, target.z / target.z)), rd.cameraInverseViewMatrix);
                    const V2 jit = randomPointInCircle(seed);
                    const float jx = jit.x * rd.dofStrength, jy = jit.y * rd.dofStrength;
                    const V3 rayOffset = v3(rd.cameraRightDir) * jx + v3(rd.cameraUpDir) * jy;
                    const V3 focusPoint = o + d * (rd.dofDistance + rd.cameraNearZ);
                    o = o + rayOffset;
                    d = normalize(focusPoint - o);
                    ps.o = o; ps.d = d; ps.seed = seed; ps.pixel = py * width + px; ps.bounce = 0u;
                    ps.light = v3(0.f, 0.f, 0.f); ps.contribution = v3(1.f, 1.f, 1.f);
                    alive = true; pixels++;
                    if (kInstrumented) { bbCheckCount = 0u; triCheckCount = 0u; pathSegments = 0u; firstFlags = 0u; }
                }
            }
            slotNext += take;
            dead = __ballot_sync(0xFFFFFFFFu, !alive);
        }
        if (__ballot_sync(0xFFFFFFFFu, alive) == 0u) break;

        if (alive) {
            /* ---- one path segment: findClosestHit + lights + bounce, RaytracerCS.hlsl:444-514 */
            const V3 o = ps.o, d = ps.d;
            const uint32_t i = ps.bounce;
            Hit hit;
            bool usedFallback = false;
            float fallbackLimit = OKRT_FLT_MAX;
            if (exactMode) traverseExact(s, numSpheres, o, d, hit, bbCheckCount, triCheckCount);
            else if (traverseFast(s, numSpheres, o, d, hit, fallbackLimit)) {
                uint32_t bb = 0u, tc = 0u;
                traverseExact(s, numSpheres, o, d, hit, bb, tc, fallbackLimit);
                usedFallback = true; fallbacks++;
            }
            segments++;
            if (kInstrumented) { pathSegments++; if (i == 0u) { first = hit; firstFlags = usedFallback ? 1u : 0u; } }

            const bool finished = shadeSegment(s, rd, fp.maxBounces, hit, ps);
            V3 light = ps.light;

            if (finished) {
                if (kInstrumented) {                           /* debug heat maps :516-520 */
                    if (rd.debugMode == 1u)
                        light = bbCheckCount > rd.debugMaxCount ? v3(1.f, 0.f, 0.f) : v3(1.f, 1.f, 1.f) * ((float)bbCheckCount / (float)rd.debugMaxCount);
                    else if (rd.debugMode == 2u)
                        light = triCheckCount > rd.debugMaxCount ? v3(1.f, 0.f, 0.f) : v3(1.f, 1.f, 1.f) * ((float)triCheckCount / (float)rd.debugMaxCount);
                }
                /* accumulate + store :525-536 [Q14] */
                const size_t pi = ps.pixel;
                float r, g, b, a;
                if (rd.accumulationEnabled == 1u) {
                    float4 acc = fp.accumulation[pi];
                    acc.x += light.x; acc.y += light.y; acc.z += light.z; acc.w += 1.f;
                    fp.accumulation[pi] = acc;
                    const float n = (float)rd.numAccumulationFrames;
                    r = acc.x / n; g = acc.y / n; b = acc.z / n; a = acc.w / n;
                } else { r = light.x; g = light.y; b = light.z; a = 1.f; }
                fp.result[pi] = toUnorm8(r, g, b, a);
                if (kInstrumented && fp.aov) {
                    okrt_aov av;
                    av.hitKind = first.kind; av.hitIdx = first.idx; av.triIdx = first.kind == 2u ? first.tri : 0xFFFFFFFFu;
                    av.baryU = first.kind == 2u ? first.u : 0.f; av.baryV = first.kind == 2u ? first.v : 0.f;
                    av.distance = first.distance;
                    av.bbCheckCount = bbCheckCount; av.triCheckCount = triCheckCount;
                    av.numSegments = pathSegments; av.flags = firstFlags;
                    av.radiance[0] = light.x; av.radiance[1] = light.y; av.radiance[2] = light.z; av.reserved = 0u;
                    fp.aov[pi] = av;
                }
                alive = false;
            }
        }
    }

    /* ray counters: one atomic per warp */
    const uint32_t segSum = __reduce_add_sync(0xFFFFFFFFu, segments);
    const uint32_t fbSum = __reduce_add_sync(0xFFFFFFFFu, fallbacks);
    const uint32_t pxSum = __reduce_add_sync(0xFFFFFFFFu, pixels);
    if (lane == 0u && pxSum) {
        atomicAdd(fp.counters + 0, (unsigned long long)segSum);
        if (fbSum) atomicAdd(fp.counters + 1, (unsigned long long)fbSum);
        atomicAdd(fp.counters + 2, (unsigned long long)pxSum);
    }
}

/* host clear of the result target, RayTracer.cpp:227-228: (0.2, 0.4, 0.6, 1) as UNORM8 */
__global__ void clearResultKernel(uchar4* result, size_t n)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) result[i] = make_uchar4(51, 102, 153, 255);
}

/* result = saturate(accumulation / frames), RaytracerCS.hlsl:529 — used after an external reduction */
__global__ void resolveKernel(const float4* accumulation, uchar4* result, size_t n, float frames, uint32_t width, uint32_t coveredW, uint32_t coveredH)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t x = (uint32_t)(i % width), y = (uint32_t)(i / width);
    if (x >= coveredW || y >= coveredH) { result[i] = make_uchar4(51, 102, 153, 255); return; }
    const float4 a = accumulation[i];
    result[i] = toUnorm8(a.x / frames, a.y / frames, a.z / frames, a.w / frames);
}

} // namespace okrt

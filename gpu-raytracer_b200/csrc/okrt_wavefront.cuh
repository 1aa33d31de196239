/*
 * okrt_wavefront.cuh — the wavefront form of RaytracerCS::main (RaytracerCS.hlsl:421-537).
 *
 * The shader is a megakernel: one thread per pixel loops over bounces, and inside each bounce over BVH
 * nodes.  On a 32-wide SIMT machine both loops diverge: ncu on the persistent megakernel of this repo
 * (profiles/r1_b_*) shows ~10 of 32 lanes active in the node loop, because a warp waits for its slowest
 * ray before any lane can shade and start its next segment.  The wavefront form removes that coupling:
 *
 *   generate  : one thread per pixel -> primary ray + path state in HBM (64 B per path)
 *   per bounce: trace  — persistent warps pull rays from a queue; a lane whose ray is finished writes its
 *                        hit record (20 B) and immediately pulls the next ray (ballot/popc refill), so the
 *                        traversal loop always runs with (nearly) full warps;
 *               shade  — one thread per queued path: hit resolve, lights, environment, lobe pick, bounce;
 *                        surviving paths are appended to the next queue by warp-aggregated atomics,
 *                        finished paths accumulate into u1 and store u0.
 *
 * Per-path arithmetic is the same code as the megakernel (traverseExact / the fast traversal's box and
 * triangle tests / shadeSegment), so results are bit-identical between the two pipelines and the oracle.
 * Path state traffic: (64 B read + 64 B write + 2 x 20 B hit) per segment ~ 1.5 GB per 1080p frame,
 * i.e. ~0.25 ms of HBM time against several ms of traversal.
 */
#pragma once
#include "okrt_kernels.cuh"

namespace okrt {

struct WavefrontBuffers {
    float4* rayO;          /* o.xyz, bits(seed) */
    float4* rayD;          /* d.xyz, bits(pixel index) */
    float4* radiance;      /* light.xyz, bits(bounce) */
    float4* throughput;    /* contribution.xyz, unused */
    float4* hitA;          /* distance, u, v, bits(triangle) */
    uint32_t* hitB;        /* kind << 30 | instance-or-sphere index */
    uint32_t* queue[2];    /* ping-pong queues of path slots */
    uint32_t* counts;      /* [2*bounce] number of queued paths, [2*bounce+1] trace work counter; zeroed per frame */
    unsigned long long* traceStats;   /* OKRT_TRACE_STATS builds: 8 counters per bounce (diagnostic) */
    uint32_t* fallbackQueue;   /* paths whose hit the reference-order kernel has to decide (this bounce) */
    uint32_t* fallbackCounts;  /* [bounce]; zeroed per frame (lives behind counts) */
};
constexpr uint32_t kMaxWavefrontBounces = 63;
/* hitB of a path whose hit the reference-order kernel is still deciding: the main shade launch skips it, the late one (after
 * traceFallbackKernel, which may run beside the main shade launch on a second stream) shades it */
constexpr uint32_t kHitPending = 0xFFFFFFFFu;
constexpr uint32_t kMaxQueueCounts = 2 * (kMaxWavefrontBounces + 2);
#ifndef OKRT_TRACE_BLOCK_THREADS
#define OKRT_TRACE_BLOCK_THREADS 128
#endif
constexpr int kTraceBlockThreads = OKRT_TRACE_BLOCK_THREADS;
#ifndef OKRT_TRACE_MIN_BLOCKS
#define OKRT_TRACE_MIN_BLOCKS 12         /* 40 registers, 48 warps per SM, the lane's hit record behind its stack (OKRT_COLD_IN_LOCAL):
                                         * 8 blocks / 63 registers 2540, 9 / 56 2655, 10 / 48 2677, 12 / 40 2704, 16 / 32 2434 Mrays/s */
#endif
#ifndef OKRT_KEEP_RUNNING
#define OKRT_KEEP_RUNNING 16
#endif
#ifndef OKRT_TRACE_CHUNK
#define OKRT_TRACE_CHUNK 64
#endif
constexpr uint32_t kTraceChunk = OKRT_TRACE_CHUNK;         /* rays a warp claims per atomic */


/* ---------------------------------------------------------------- generate
 * [Q3] seed, createRay :177-206 [Q2], applyDOF :208-216.  Slot order = 16x8 tile order of this rank's tiles. */
__global__ void __launch_bounds__(128)
generateKernel(const __grid_constant__ FrameParams fp, const WavefrontBuffers wb)
{
    const okrt_render_data& rd = fp.rd;
    const uint32_t gslot = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t numSlots = fp.numTiles * kTilePixels;            /* a multiple of 128: warps never straddle frames */
    const uint32_t frame = gslot / numSlots, slot = gslot % numSlots;
    if (frame >= fp.batchFrames) return;
    const uint32_t tile = slot / kTilePixels, j = slot % kTilePixels;
    const uint32_t tileX = tile % fp.tilesX, tileRowLocal = tile / fp.tilesX;
    const uint32_t tileRow = ownedTileRow(fp, tileRowLocal);
    const uint32_t px = tileX * kTileW + j % kTileW, py = tileRow * kTileH + j / kTileW;
    const bool covered = px < fp.coveredW && py < fp.coveredH;

    /* path slot = frame * numSlots + pixel slot (so that the final gather can add frames in order); covered
     * pixels are compacted into queue 0 (keeps tile order inside a warp).  When every slot is a covered pixel (1920x1080,
     * 3840x2160, ...: width a multiple of 16, height of 72) the queue of bounce 0 is the identity: nothing is written, the
     * kernels of bounce 0 index the paths directly, and the count is stored once (a million same-address atomics less). */
    const uint32_t path = gslot;
    if (fp.allCovered) {
        if (gslot == 0u) wb.counts[0] = numSlots * fp.batchFrames;
    } else {
        const uint32_t mask = __ballot_sync(0xFFFFFFFFu, covered);
        uint32_t base = 0u;
        if ((threadIdx.x & 31u) == 0u && mask) base = atomicAdd(wb.counts + 0, (uint32_t)__popc(mask));
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (!covered) return;
        wb.queue[0][base + __popc(mask & ((1u << (threadIdx.x & 31u)) - 1u))] = path;
    }

    const uint32_t width = rd.textureDims[0];
    uint32_t seed = px + (py + 74813u) * width * ((fp.seedFrame + frame * fp.seedStride) + 1u);
    V3 pos = v3((float)px, (float)(rd.textureDims[1] - py), rd.cameraNearZ);
    pos.x += randomFloat(seed) - 0.5f;
    pos.y += randomFloat(seed) - 0.5f;
    pos.x /= (float)rd.textureDims[0]; pos.y /= (float)rd.textureDims[1];
    pos.x *= 2.f; pos.y *= 2.f;
    pos.x -= 1.f; pos.y -= 1.f;
    const V3 target = mulPoint(pos, rd.cameraInverseProjectionMatrix);
    V3 o = v3(rd.cameraPosition);
    V3 d = mulDir(normalize(v3(target.x / target.z, target.y / target.z, target.z / target.z)), rd.cameraInverseViewMatrix);
    /* applyDOF with dofStrength == 0: the disc sample is multiplied by zero, so the offset is (+-0, +-0, +-0) and
     * o + offset == o bit for bit unless a camera coordinate is -0.0f (-0 + +0 = +0) — then, and for every non-zero
     * strength, the full sample is evaluated.  The two RNG draws happen either way (the seed must advance). */
    V2 jit;
    const bool plainZeroOffset = rd.dofStrength == 0.f && __float_as_uint(rd.cameraPosition[0]) != 0x80000000u &&
                                 __float_as_uint(rd.cameraPosition[1]) != 0x80000000u && __float_as_uint(rd.cameraPosition[2]) != 0x80000000u;
    if (plainZeroOffset) { (void)randomFloat(seed); (void)randomFloat(seed); jit.x = 0.f; jit.y = 0.f; }
    else jit = randomPointInCircle(seed);
    const float jx = jit.x * rd.dofStrength, jy = jit.y * rd.dofStrength;
    const V3 rayOffset = v3(rd.cameraRightDir) * jx + v3(rd.cameraUpDir) * jy;
    const V3 focusPoint = o + d * (rd.dofDistance + rd.cameraNearZ);
    o = o + rayOffset;
    d = normalize(focusPoint - o);

    wb.rayO[path] = make_float4(o.x, o.y, o.z, __uint_as_float(seed));
    wb.rayD[path] = make_float4(d.x, d.y, d.z, __uint_as_float(py * width + px));
    /* radiance (0, 0, 0 | bounce 0) and throughput (1, 1, 1) of a new path are not stored: shadeKernel knows them at bounce 0 */
    if (fp.aov) {
        okrt_aov av;
        av.hitKind = 0u; av.hitIdx = 0xFFFFFFFFu; av.triIdx = 0xFFFFFFFFu; av.baryU = 0.f; av.baryV = 0.f; av.distance = OKRT_FLT_MAX;
        av.bbCheckCount = 0u; av.triCheckCount = 0u; av.numSegments = 0u; av.flags = 0u;
        av.radiance[0] = av.radiance[1] = av.radiance[2] = 0.f; av.reserved = 0u;
        fp.aov[py * width + px] = av;
    }
}

/* ------------------------------------------------------------------- trace
 * Reference-order traversal (debug counters, scaled instances, forced exact mode): one thread per ray. */
__global__ void __launch_bounds__(128)
traceExactKernel(const __grid_constant__ DeviceScene s, const __grid_constant__ FrameParams fp, const WavefrontBuffers wb,
                 const uint32_t* __restrict__ queueIn, uint32_t bounce)
{
    const uint32_t count = wb.counts[2u * bounce];
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= count) return;
    const uint32_t path = queueIn ? queueIn[k] : k;
    const float4 ro = wb.rayO[path], rdv = wb.rayD[path];
    const uint32_t numSpheres = fp.rd.numSpheres < s.numSpheres ? fp.rd.numSpheres : s.numSpheres;
    Hit hit;
    uint32_t bb = 0u, tc = 0u;
    traverseExact(s, numSpheres, v3(ro.x, ro.y, ro.z), v3(rdv.x, rdv.y, rdv.z), hit, bb, tc);
    wb.hitA[path] = make_float4(hit.distance, hit.u, hit.v, __uint_as_float(hit.tri));
    wb.hitB[path] = (hit.kind << 30) | (hit.idx & 0x3FFFFFFFu);
    if (fp.aov) {
        okrt_aov* av = fp.aov + __float_as_uint(rdv.w);
        av->bbCheckCount += bb; av->triCheckCount += tc;
    }
}

/* The rays traceFastKernel could not decide (two candidates inside the window, or an oct-tree chain that the
 * winner's instance fails): reference order, full warps, one launch per bounce. */
__global__ void __launch_bounds__(128)
traceFallbackKernel(const __grid_constant__ DeviceScene s, const __grid_constant__ FrameParams fp, const WavefrontBuffers wb, uint32_t bounce)
{
    const uint32_t count = wb.fallbackCounts[bounce];
    const uint32_t numSpheres = fp.rd.numSpheres < s.numSpheres ? fp.rd.numSpheres : s.numSpheres;
    /* one ray per WARP: 32 unrelated reference-order traversals in one warp diverge completely and serialise
     * (measured: ~1 ms per launch for ~2000 rays); a warp per ray lasts one traversal, and the warp shares the
     * instance tests of that traversal among its lanes (traverseExactWarp) */
    const uint32_t warpsPerGrid = gridDim.x * (blockDim.x >> 5);
    for (uint32_t k = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); k < count; k += warpsPerGrid) {
        const uint32_t path = wb.fallbackQueue[k];
        const float4 ro = wb.rayO[path], rdv = wb.rayD[path];
        const float limit = wb.hitA[path].x;                        /* the bound the near-first traversal left for this ray */
        Hit hit;
        traverseExactWarp(s, numSpheres, v3(ro.x, ro.y, ro.z), v3(rdv.x, rdv.y, rdv.z), hit, limit);
        __syncwarp();
        if ((threadIdx.x & 31u) == 0u) {
            /* hitB keeps its pending marker — the main shade launch may be reading it at this moment; kind | index travel in the
             * spare word of the path's throughput */
            wb.hitA[path] = make_float4(hit.distance, hit.u, hit.v, __uint_as_float(hit.tri));
            wb.throughput[path].w = __uint_as_float((hit.kind << 30) | (hit.idx & 0x3FFFFFFFu));
            if (fp.aov && bounce == 0u) fp.aov[__float_as_uint(rdv.w)].flags = 1u;
        }
    }
    if (blockIdx.x == 0u && threadIdx.x == 0u && count) atomicAdd(fp.counters + 1, (unsigned long long)count);
}

/* Near-first traversal in step form (StepLane, okrt_kernels.cuh): persistent warps with lane refill.  A warp
 * alternates between a NODE phase (every lane that stands on an interior node tests one child pair per
 * iteration) and a TRIANGLE phase (every lane that stands on a leaf tests one triangle per iteration); a phase
 * ends when fewer than a threshold of its lanes are left, so that no phase drags on for a few stragglers while
 * the other lanes wait (the while-while form waits for the LAST lane: 9.2 of 32 lanes active, profiles/r1_d_*). */
#ifndef OKRT_REFILL_IDLE
#define OKRT_REFILL_IDLE 16         /* refill as soon as this many lanes have no ray */
#endif
#ifndef OKRT_NODE_KEEP
#define OKRT_NODE_KEEP 6            /* stay in the node phase while at least this many lanes want a node step */
#endif
#ifndef OKRT_KEEP_CONST
#define OKRT_KEEP_CONST 0           /* 1: compile-time keep thresholds (measured slower: the register allocation changes) */
#endif
#ifndef OKRT_TRI_FIRST
#define OKRT_TRI_FIRST 0
#endif
#ifndef OKRT_TRI_KEEP
#define OKRT_TRI_KEEP 6             /* stay in the triangle phase while at least this many lanes want a triangle step */
#endif
__global__ void __launch_bounds__(kTraceBlockThreads, OKRT_TRACE_MIN_BLOCKS)
traceFastKernel(const __grid_constant__ DeviceScene s, const __grid_constant__ FrameParams fp, const WavefrontBuffers wb,
                const uint32_t* __restrict__ queueIn, uint32_t bounce)
{
    const uint32_t count = wb.counts[2u * bounce];
    uint32_t* workCounter = wb.counts + 2u * bounce + 1u;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t laneMaskLt = (1u << lane) - 1u;
    const uint32_t numSpheres = fp.rd.numSpheres < s.numSpheres ? fp.rd.numSpheres : s.numSpheres;

    uint32_t stack[kFastStackCapacity + kColdSlots];
    StepLane L;
    L.cur = kTraceDone; L.sp = 1; L.path = 0u; L.inst = kTraceTop; L.triNext = 0u;
    stack[0] = kTraceDone;
    OKRT_SMEM_STACK_DECLARE(L)
#if OKRT_SMEM_STACK_ON
    stackSharedStore(L.smBase, kTraceDone);
#endif
#if OKRT_STACK_TOPREG
    L.top = kTraceDone;
#endif
    bool have = false;                        /* this lane owns a ray whose result is not written yet */
    uint32_t next = 0u, end = 0u;             /* warp-uniform: unclaimed part of the warp's current chunk */
    bool exhausted = false;                   /* warp-uniform */
#ifdef OKRT_TRACE_STATS
    uint32_t stNodeIters = 0u, stNodeSteps = 0u, stTriIters = 0u, stTriSteps = 0u, stRefills = 0u, stRays = 0u;
    const long long stStart = clock64();
#endif

    /* lanes without a ray stand on kTraceDone, which belongs to no phase */
    for (;;) {
        const uint32_t tag = L.cur & kRefTagMask;
        const uint32_t mNode = __ballot_sync(0xFFFFFFFFu, tag == kRefInterior);
        const uint32_t mTri = __ballot_sync(0xFFFFFFFFu, tag == kRefTriLeaf);
        const uint32_t mMisc = __ballot_sync(0xFFFFFFFFu, tag == kRefInstance && L.cur != kTraceDone);
        const uint32_t nNode = (uint32_t)__popc(mNode), nTri = (uint32_t)__popc(mTri);

        if (mMisc) {                                                    /* instance entry / exit: rare, handled at once */
            if (tag == kRefInstance && L.cur != kTraceDone) {
                const float4 ro = wb.rayO[L.path], rdv = wb.rayD[L.path];
                stepMisc(s, L, stack, v3(ro.x, ro.y, ro.z), v3(rdv.x, rdv.y, rdv.z));
            }
            continue;
        }

        const uint32_t active = nNode + nTri;
        if (active == 0u || (!exhausted && active <= 32u - (uint32_t)OKRT_REFILL_IDLE)) {
            /* ---- finished lanes write their hit record (or queue the ray for the reference order) */
            if (have && L.cur == kTraceDone) {
                Hit hit;
                float seed;
                if (stepResult(s, L, stack, wb.rayD + L.path, hit, seed)) {
                    wb.fallbackQueue[atomicAdd(wb.fallbackCounts + bounce, 1u)] = L.path;
                    wb.hitA[L.path].x = relaxedLimit(seed);              /* bounds the reference-order traversal's search */
                    wb.hitB[L.path] = kHitPending;
                } else {
                    wb.hitA[L.path] = make_float4(hit.distance, hit.u, hit.v, __uint_as_float(hit.tri));
                    wb.hitB[L.path] = (hit.kind << 30) | (hit.idx & 0x3FFFFFFFu);
                }
                have = false;
            }
            /* ---- refill */
#ifdef OKRT_TRACE_STATS
            stRefills++;
#endif
            uint32_t need = __ballot_sync(0xFFFFFFFFu, !have);
            while (need != 0u && !exhausted) {
                if (next == end) {
                    uint32_t base = 0u;
                    if (lane == 0u) base = atomicAdd(workCounter, kTraceChunk);
                    base = __shfl_sync(0xFFFFFFFFu, base, 0);
                    if (base >= count) { exhausted = true; break; }
                    next = base; end = min(base + kTraceChunk, count);
                }
                const uint32_t want = __popc(need), avail = end - next;
                const uint32_t take = want < avail ? want : avail;
                const uint32_t rank = __popc(need & laneMaskLt);
                if (!have && rank < take) {
                    const uint32_t k = next + rank;
                    L.path = queueIn ? __ldg(queueIn + k) : k;
                    const float4 ro = wb.rayO[L.path], rdv = wb.rayD[L.path];
                    stepInit(s, L, stack, numSpheres, v3(ro.x, ro.y, ro.z), v3(rdv.x, rdv.y, rdv.z));
                    have = true;
#ifdef OKRT_TRACE_STATS
                    stRays++;
#endif
                }
                next += take;
                need = __ballot_sync(0xFFFFFFFFu, !have);
            }
            if (__ballot_sync(0xFFFFFFFFu, have) == 0u) break;
            continue;
        }

#if OKRT_TRI_FIRST
        if (nNode > nTri) {
#else
        if (nNode >= nTri) {
#endif
            /* ---- node phase */
#if OKRT_KEEP_CONST
            constexpr uint32_t keep = (uint32_t)OKRT_NODE_KEEP;
#else
            const uint32_t keep = min((uint32_t)OKRT_NODE_KEEP, nNode);
#endif
            do {
#ifdef OKRT_TRACE_STATS
                stNodeIters++; if ((L.cur & kRefTagMask) == kRefInterior) stNodeSteps++;
#endif
                if ((L.cur & kRefTagMask) == kRefInterior) stepNode(s, L, stack);
            } while ((uint32_t)__popc(__ballot_sync(0xFFFFFFFFu, (L.cur & kRefTagMask) == kRefInterior)) >= keep);
        } else {
            /* ---- triangle phase */
#if OKRT_KEEP_CONST
            constexpr uint32_t keep = (uint32_t)OKRT_TRI_KEEP;
#else
            const uint32_t keep = min((uint32_t)OKRT_TRI_KEEP, nTri);
#endif
            do {
#ifdef OKRT_TRACE_STATS
                stTriIters++; if ((L.cur & kRefTagMask) == kRefTriLeaf) stTriSteps++;
#endif
                if ((L.cur & kRefTagMask) == kRefTriLeaf) stepTriangle(s, L, stack);
            } while ((uint32_t)__popc(__ballot_sync(0xFFFFFFFFu, (L.cur & kRefTagMask) == kRefTriLeaf)) >= keep);
        }
    }
#ifdef OKRT_TRACE_STATS
    {
        /* per bounce, the layout okrt_get_trace_stats documents: [0] rays [1] node steps [2] leaf entries (wide layout only)
         * [3] triangle tests [4] node-phase iterations (warps) [5] triangle-phase iterations [6] lanes busy in node iterations
         * [7] lanes busy in triangle iterations */
        unsigned long long* st = wb.traceStats + 8u * bounce;
        const uint32_t a0 = __reduce_add_sync(0xFFFFFFFFu, stRays), a1 = __reduce_add_sync(0xFFFFFFFFu, stNodeSteps), a2 = __reduce_add_sync(0xFFFFFFFFu, stTriSteps);
        if (lane == 0u) {
            atomicAdd(st + 0, (unsigned long long)a0); atomicAdd(st + 1, (unsigned long long)a1); atomicAdd(st + 3, (unsigned long long)a2);
            atomicAdd(st + 4, (unsigned long long)stNodeIters); atomicAdd(st + 5, (unsigned long long)stTriIters);
            atomicAdd(st + 6, (unsigned long long)a1); atomicAdd(st + 7, (unsigned long long)a2);
        }
        (void)stStart; (void)stRefills;
    }
#endif
}

/* ------------------------------------------------------------- trace, wide layout
 * The same persistent-warp scheme as traceFastKernel over the wide layout (okrt_wide.cuh): a NODE phase (every lane with a
 * pending node group visits one 8-wide node: three 256-bit loads, eight quantised slab tests) and a TRIANGLE phase (every lane
 * with hit leaves enters one leaf — exact box test — and / or tests one triangle).  A lane's leaves are tested before it
 * descends further (they belong to the node it just visited: the nearest geometry it knows). */
#ifndef OKRT_WIDE_MIN_BLOCKS
#define OKRT_WIDE_MIN_BLOCKS 8
#endif
#ifndef OKRT_WIDE_NODE_KEEP
#define OKRT_WIDE_NODE_KEEP 8
#endif
#ifndef OKRT_WIDE_TRI_KEEP
#define OKRT_WIDE_TRI_KEEP 8
#endif
#ifndef OKRT_WIDE_REFILL_IDLE
#define OKRT_WIDE_REFILL_IDLE 12
#endif
__global__ void __launch_bounds__(kTraceBlockThreads, OKRT_WIDE_MIN_BLOCKS)
traceWideKernel(const __grid_constant__ DeviceScene s, const __grid_constant__ FrameParams fp, const WavefrontBuffers wb,
                const uint32_t* __restrict__ queueIn, uint32_t bounce)
{
    const uint32_t count = wb.counts[2u * bounce];
    uint32_t* workCounter = wb.counts + 2u * bounce + 1u;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t laneMaskLt = (1u << lane) - 1u;
    const uint32_t numSpheres = fp.rd.numSpheres < s.numSpheres ? fp.rd.numSpheres : s.numSpheres;

    uint32_t stack[kWideStackWords + kWideColdWords];
    WideLane L;
    L.sp = 0; L.path = 0u; L.inst = kTraceTop; L.lg = 0u; L.triNext = L.triEnd = 0u; L.ngHits = 0u; L.ngBase = 0u; L.flags = 0u;
    L.best = L.lam = OKRT_FLT_MAX; L.leafNear = L.leafSlack = 0.f; L.octInv = 0u;
    L.ro = L.rd = L.ridir = v3(0.f, 0.f, 0.f);
#ifdef OKRT_TRACE_STATS
    L.stNodes = L.stLeaves = L.stTris = 0u;
    uint32_t stNodeIters = 0u, stTriIters = 0u, stRefills = 0u, stRays = 0u, stNodeSum = 0u, stLeafSum = 0u, stTriSum = 0u, stNodeLanes = 0u, stTriLanes = 0u;
    const long long stStart = clock64();
#endif
    bool have = false;                        /* this lane owns a ray whose result is not written yet */
    uint32_t next = 0u, end = 0u;             /* warp-uniform: unclaimed part of the warp's current chunk */
    bool exhausted = false;                   /* warp-uniform */

    for (;;) {
        const bool wantsTri = have && wideWantsTri(L), wantsNode = have && wideWantsNode(L);
        const uint32_t nNode = (uint32_t)__popc(__ballot_sync(0xFFFFFFFFu, wantsNode));
        const uint32_t nTri = (uint32_t)__popc(__ballot_sync(0xFFFFFFFFu, wantsTri));
        const uint32_t active = nNode + nTri;
        if (active == 0u || (!exhausted && active <= 32u - (uint32_t)OKRT_WIDE_REFILL_IDLE)) {
            /* ---- finished lanes write their hit record (or queue the ray for the reference order) */
            if (have && wideDone(L)) {
                Hit hit;
                float limit;
                if (wideResult(s, L, stack, wb.rayD + L.path, hit, limit)) {
                    wb.fallbackQueue[atomicAdd(wb.fallbackCounts + bounce, 1u)] = L.path;
                    wb.hitA[L.path].x = limit;                           /* bounds the reference-order traversal's search */
                    wb.hitB[L.path] = kHitPending;
                } else {
                    wb.hitA[L.path] = make_float4(hit.distance, hit.u, hit.v, __uint_as_float(hit.tri));
                    wb.hitB[L.path] = (hit.kind << 30) | (hit.idx & 0x3FFFFFFFu);
                }
#ifdef OKRT_TRACE_STATS
                stNodeSum += L.stNodes; stLeafSum += L.stLeaves; stTriSum += L.stTris;
#endif
                have = false;
            }
            /* ---- refill */
#ifdef OKRT_TRACE_STATS
            stRefills++;
#endif
            uint32_t need = __ballot_sync(0xFFFFFFFFu, !have);
            while (need != 0u && !exhausted) {
                if (next == end) {
                    uint32_t base = 0u;
                    if (lane == 0u) base = atomicAdd(workCounter, kTraceChunk);
                    base = __shfl_sync(0xFFFFFFFFu, base, 0);
                    if (base >= count) { exhausted = true; break; }
                    next = base; end = min(base + kTraceChunk, count);
                }
                const uint32_t want = __popc(need), avail = end - next;
                const uint32_t take = want < avail ? want : avail;
                const uint32_t rank = __popc(need & laneMaskLt);
                if (!have && rank < take) {
                    const uint32_t k = next + rank;
                    L.path = queueIn ? __ldg(queueIn + k) : k;
                    const float4 ro = wb.rayO[L.path], rdv = wb.rayD[L.path];
                    wideInit(s, L, stack, numSpheres, v3(ro.x, ro.y, ro.z), v3(rdv.x, rdv.y, rdv.z));
                    have = true;
#ifdef OKRT_TRACE_STATS
                    stRays++;
#endif
                }
                next += take;
                need = __ballot_sync(0xFFFFFFFFu, !have);
            }
            if (__ballot_sync(0xFFFFFFFFu, have) == 0u) break;
            continue;
        }
        if (nNode >= nTri) {
            const uint32_t keep = min((uint32_t)OKRT_WIDE_NODE_KEEP, nNode);
            uint32_t left;
            do {
                const bool mine = have && wideWantsNode(L);
#ifdef OKRT_TRACE_STATS
                stNodeIters++; if (mine) stNodeLanes++;
#endif
                if (mine) wideNodeStep(s, L, stack, wb.rayO, wb.rayD);
                left = (uint32_t)__popc(__ballot_sync(0xFFFFFFFFu, have && wideWantsNode(L)));
            } while (left >= keep);
        } else {
            const uint32_t keep = min((uint32_t)OKRT_WIDE_TRI_KEEP, nTri);
            uint32_t left;
            do {
                const bool mine = have && wideWantsTri(L);
#ifdef OKRT_TRACE_STATS
                stTriIters++; if (mine) stTriLanes++;
#endif
                if (mine) wideTriStep(s, L, stack, wb.rayO, wb.rayD);
                left = (uint32_t)__popc(__ballot_sync(0xFFFFFFFFu, have && wideWantsTri(L)));
            } while (left >= keep);
        }
    }
#ifdef OKRT_TRACE_STATS
    {
        /* per bounce: [0] rays [1] node steps [2] leaf entries [3] triangle tests [4] node-phase iterations (warps) [5] triangle-phase
         * iterations [6] lanes busy in node iterations [7] lanes busy in triangle iterations */
        unsigned long long* st = wb.traceStats + 8u * bounce;
        const uint32_t a0 = __reduce_add_sync(0xFFFFFFFFu, stRays), a1 = __reduce_add_sync(0xFFFFFFFFu, stNodeSum), a2 = __reduce_add_sync(0xFFFFFFFFu, stLeafSum);
        const uint32_t a3 = __reduce_add_sync(0xFFFFFFFFu, stTriSum), a6 = __reduce_add_sync(0xFFFFFFFFu, stNodeLanes), a7 = __reduce_add_sync(0xFFFFFFFFu, stTriLanes);
        if (lane == 0u) {
            atomicAdd(st + 0, (unsigned long long)a0); atomicAdd(st + 1, (unsigned long long)a1); atomicAdd(st + 2, (unsigned long long)a2);
            atomicAdd(st + 3, (unsigned long long)a3); atomicAdd(st + 4, (unsigned long long)stNodeIters); atomicAdd(st + 5, (unsigned long long)stTriIters);
            atomicAdd(st + 6, (unsigned long long)a6); atomicAdd(st + 7, (unsigned long long)a7);
        }
        (void)stStart; (void)stRefills;
    }
#endif
}

/* ------------------------------------------------------------------- shade
 * RaytracerCS.hlsl:336-392 (hit resolve) + :448-536 for one queued path per thread. */
#ifndef OKRT_SHADE_MIN_BLOCKS
#define OKRT_SHADE_MIN_BLOCKS 8
#endif
/* paths a thread shades per trip before the block appends its survivors to the next queue (one atomic and two barriers per trip) */
#ifndef OKRT_SHADE_PATHS_PER_THREAD
#define OKRT_SHADE_PATHS_PER_THREAD 1
#endif
constexpr uint32_t kShadePathsPerThread = OKRT_SHADE_PATHS_PER_THREAD;
static_assert(kShadePathsPerThread == 1u || kShadePathsPerThread == 2u, "one or two");
template <bool kInstrumented>
__global__ void __launch_bounds__(128, OKRT_SHADE_MIN_BLOCKS)
shadeKernel(const __grid_constant__ DeviceScene s, const __grid_constant__ FrameParams fp, const WavefrontBuffers wb,
            const uint32_t* __restrict__ queueIn, uint32_t* __restrict__ queueOut, uint32_t bounce, uint32_t late)
{
    /* late == 0: every path of the bounce's queue except those still with the reference-order kernel (kHitPending);
     * late == 1: exactly those, from the fallback queue, once that kernel is done */
    const okrt_render_data& rd = fp.rd;
    const uint32_t count = late ? wb.fallbackCounts[bounce] : wb.counts[2u * bounce];
    __shared__ uint32_t warpSurvivors[2][4];     /* two copies, used by alternate trips: a trip's readers and the next trip's */
    __shared__ uint32_t blockBase[2];            /* writers are a barrier apart without a third barrier per trip */
    uint32_t trip = 0u;
    if (blockIdx.x == 0u && threadIdx.x == 0u && !late) {
        atomicAdd(fp.counters + 0, (unsigned long long)count);
        if (bounce == 0u) atomicAdd(fp.counters + 2, (unsigned long long)count);
    }
    /* the grid is sized for the machine, not for the queue (whose length only the device knows): a block walks the
     * queue in grid strides, the trip count is uniform within a block */
    constexpr uint32_t kSpt = kShadePathsPerThread;
    for (uint32_t first = blockIdx.x * blockDim.x * kSpt; first < count; first += gridDim.x * blockDim.x * kSpt) {
    bool survives = false, keep0 = false;
    uint32_t path = 0u, keepPath0 = 0u;
#pragma unroll 1
    for (uint32_t h = 0u; h < kSpt; h++) {
    if (kSpt == 2u && h == 1u) { keep0 = survives; keepPath0 = path; }
    const uint32_t k = first + h * blockDim.x + threadIdx.x;
    const bool active = k < count;
    survives = false;
    path = 0u;
    uint32_t hbEarly = 0u;
    float4 ro, rdv, li, th, ha;
    if (active) {
        /* the whole path record is requested at once, before hitB says whether the path is this launch's to shade (one memory
         * latency instead of two in a row; a path still with the reference-order kernel is one in ten thousand) */
        path = late ? wb.fallbackQueue[k] : (queueIn ? __ldg(queueIn + k) : k);
        hbEarly = wb.hitB[path];
        ro = wb.rayO[path]; rdv = wb.rayD[path]; ha = wb.hitA[path];
        if (bounce == 0u) {                      /* a new path: light 0, bounce 0, contribution 1 (generateKernel stores neither) */
            li = make_float4(0.f, 0.f, 0.f, __uint_as_float(0u));
            th = make_float4(1.f, 1.f, 1.f, late ? wb.throughput[path].w : 0.f);
        } else { li = wb.radiance[path]; th = wb.throughput[path]; }
    }
    if (active && (late || hbEarly != kHitPending)) {
        const uint32_t hb = late ? __float_as_uint(th.w) : hbEarly;
        PathState ps;
        ps.o = v3(ro.x, ro.y, ro.z); ps.seed = __float_as_uint(ro.w);
        ps.d = v3(rdv.x, rdv.y, rdv.z); ps.pixel = __float_as_uint(rdv.w);
        ps.light = v3(li.x, li.y, li.z); ps.bounce = __float_as_uint(li.w);
        ps.contribution = v3(th.x, th.y, th.z);
        Hit hit;
        hit.distance = ha.x; hit.u = ha.y; hit.v = ha.z; hit.tri = __float_as_uint(ha.w);
        hit.kind = hb >> 30; hit.idx = hit.kind ? (hb & 0x3FFFFFFFu) : 0xFFFFFFFFu;

        const bool finished = shadeSegment(s, rd, fp.maxBounces, hit, ps);
        V3 light = ps.light;
        const size_t pi = ps.pixel;

        if (kInstrumented && fp.aov) {
            okrt_aov* av = fp.aov + pi;
            av->numSegments = bounce + 1u;
            if (bounce == 0u) {
                av->hitKind = hit.kind; av->hitIdx = hit.idx; av->triIdx = hit.kind == 2u ? hit.tri : 0xFFFFFFFFu;
                av->baryU = hit.kind == 2u ? hit.u : 0.f; av->baryV = hit.kind == 2u ? hit.v : 0.f; av->distance = hit.distance;
            }
        }
        if (finished) {
            if (kInstrumented && fp.aov) {                      /* debug heat maps :516-520 */
                const uint32_t bb = fp.aov[pi].bbCheckCount, tc = fp.aov[pi].triCheckCount;
                if (rd.debugMode == 1u)
                    light = bb > rd.debugMaxCount ? v3(1.f, 0.f, 0.f) : v3(1.f, 1.f, 1.f) * ((float)bb / (float)rd.debugMaxCount);
                else if (rd.debugMode == 2u)
                    light = tc > rd.debugMaxCount ? v3(1.f, 0.f, 0.f) : v3(1.f, 1.f, 1.f) * ((float)tc / (float)rd.debugMaxCount);
                fp.aov[pi].radiance[0] = light.x; fp.aov[pi].radiance[1] = light.y; fp.aov[pi].radiance[2] = light.z;
            }
            /* the path's radiance waits in its slot; accumulateKernel adds the batch's frames in frame order */
            wb.radiance[path] = make_float4(light.x, light.y, light.z, __uint_as_float(ps.bounce));
        } else {
            wb.rayO[path] = make_float4(ps.o.x, ps.o.y, ps.o.z, __uint_as_float(ps.seed));
            wb.rayD[path] = make_float4(ps.d.x, ps.d.y, ps.d.z, rdv.w);
            wb.radiance[path] = make_float4(ps.light.x, ps.light.y, ps.light.z, __uint_as_float(ps.bounce));
            wb.throughput[path] = make_float4(ps.contribution.x, ps.contribution.y, ps.contribution.z, 0.f);
            survives = true;
        }
    }
    }
    /* surviving paths -> next queue, one atomic per block (every warp of every block adding to the same counter is
     * half a million same-address atomics per launch) */
    const uint32_t mask = __ballot_sync(0xFFFFFFFFu, survives);
    const uint32_t mask0 = kSpt == 2u ? __ballot_sync(0xFFFFFFFFu, keep0) : 0u;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    uint32_t* const ws = warpSurvivors[trip & 1u];
    if (lane == 0u) ws[warp] = (uint32_t)(__popc(mask) + __popc(mask0));
    __syncthreads();
    if (threadIdx.x == 0u) {
        const uint32_t total = ws[0] + ws[1] + ws[2] + ws[3];
        blockBase[trip & 1u] = total ? atomicAdd(wb.counts + 2u * (bounce + 1u), total) : 0u;
    }
    __syncthreads();
    if (survives || keep0) {
        uint32_t base = blockBase[trip & 1u];
        for (uint32_t w = 0; w < warp; w++) base += ws[w];
        if (keep0) queueOut[base + __popc(mask0 & ((1u << lane) - 1u))] = keepPath0;
        if (survives) queueOut[base + __popc(mask0) + __popc(mask & ((1u << lane) - 1u))] = path;
    }
    trip++;                          /* the next trip writes the other copy; the one after it is two barriers away */
    }
}

/* ------------------------------------------------------------ accumulate
 * RaytracerCS.hlsl:525-536 [Q14] for a batch: u1 += light of frame 0, 1, ... in that order (the order the
 * reference's frame loop adds them, so u1 stays bit-identical), then u0 = saturate(u1 / numAccumulationFrames). */
__global__ void __launch_bounds__(128)
accumulateKernel(const __grid_constant__ FrameParams fp, const WavefrontBuffers wb)
{
    const okrt_render_data& rd = fp.rd;
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t numSlots = fp.numTiles * kTilePixels;
    if (slot >= numSlots) return;
    const uint32_t tile = slot / kTilePixels, j = slot % kTilePixels;
    const uint32_t tileX = tile % fp.tilesX, tileRowLocal = tile / fp.tilesX;
    const uint32_t tileRow = ownedTileRow(fp, tileRowLocal);
    const uint32_t px = tileX * kTileW + j % kTileW, py = tileRow * kTileH + j / kTileW;
    if (px >= fp.coveredW || py >= fp.coveredH) return;
    const size_t pi = (size_t)py * rd.textureDims[0] + px;
    float r, g, b, a;
    if (rd.accumulationEnabled == 1u) {
        float4 acc = fp.accumulation[pi];
        for (uint32_t f = 0; f < fp.batchFrames; f++) {
            const float4 l = wb.radiance[(size_t)f * numSlots + slot];
            acc.x += l.x; acc.y += l.y; acc.z += l.z; acc.w += 1.f;
        }
        fp.accumulation[pi] = acc;
        const float n = (float)rd.numAccumulationFrames;
        r = acc.x / n; g = acc.y / n; b = acc.z / n; a = acc.w / n;
    } else {
        const float4 l = wb.radiance[(size_t)(fp.batchFrames - 1u) * numSlots + slot];
        r = l.x; g = l.y; b = l.z; a = 1.f;
    }
    fp.result[pi] = toUnorm8(r, g, b, a);
}

} // namespace okrt

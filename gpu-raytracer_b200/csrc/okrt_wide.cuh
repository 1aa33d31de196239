/*
 * okrt_wide.cuh — near-first traversal over the WIDE layout (okrt_accel.h): 8-wide nodes whose child boxes are quantised to
 * 8 bits per plane against an exact per-node grid, exact reference leaf boxes next to the triangles.
 *
 * Replaces the per-ray work of findClosestHit (resources/shaders/RaytracerCS.hlsl:256-334) for rays the near-first path can
 * decide; everything that DECIDES a result is still the reference's arithmetic on the reference's floats:
 *   - a triangle is tested only if the ray HITS its reference leaf box (GPU-Utilities.hlsli:183-197, bit copy of the box);
 *   - the triangle test is GPU-Utilities.hlsli:117-155 [Q6], the distance RaytracerCS.hlsl:312-314.
 * The quantised boxes only CULL.  A child box contains the exact boxes of every leaf below it (grown by the leaf's slack), and
 * its slab test is widened by a bound on its own rounding error, so a leaf the reference could reach with a candidate inside
 * the window is never culled.  What the window is, and when a ray is handed to the reference-order kernel instead, is written
 * down at WideLane / wideCandidate / wideResult.
 *
 * The lane functions are __host__ __device__: okrt_trace_rays_host (okrt_device.cu) runs the same code on the CPU so that the
 * layout and the certification logic can be compared with the oracle without a GPU (tests/test_cpu_traversal.py).
 */
#pragma once
#include "okrt_kernels.cuh"

namespace okrt {

#ifndef OKRT_DELTA
#define OKRT_DELTA 1.0e-4f
#endif

/* ------------------------------------------------------------------ lane state
 * WINDOW AND CERTIFICATION.  n(l) is the reference's distNear of leaf l's box for this ray, s(l) the leaf's
 * slack.  Assumption A1 (checked for every accepted hit, assumed for triangles that are never tested): a triangle test does not
 * report a hit closer than (n - s)(1 - delta/2) — hits lie in or behind their leaf box up to rounding.
 *   lam      boxes whose (slack-grown) entry distance is >= lam are skipped.  lam only shrinks: after a new closest hit with
 *            distance t in a leaf with entry n it is min(lam, max(t, n)(1 + 2 delta)).  By A1 every candidate in a skipped box
 *            lies beyond max(t, n)(1 + delta) of every winner found so far.
 *   gate     max(t, n) of the closest hit W: the reference tests W unless something it tested EARLIER has a distance <= n
 *            (RaytracerCS.hlsl:285: it skips a box when distNear >= payload.distance), and returns W unless something ties.
 *   verdict  W is returned without the reference order iff no other tested candidate has a distance <= gate (1 + delta), the
 *            boxes skipped before W was found started beyond its gate, A1 held for every accepted hit, and the oct-tree boxes
 *            above W's instance are hit (their distNear joins the gate).  Everything else goes to traceFallbackKernel, which
 *            walks the reference order itself, bounded by gate (1 + 16 delta) + 4 maxSlack (or unbounded after an A1 violation,
 *            a failed oct-tree chain or an order-dependent instance). */
constexpr uint32_t kWideFlagAmbiguous = 1u, kWideFlagUnbounded = 2u;
constexpr uint32_t kWideColdWords = 8;
constexpr uint32_t kWideStackWords = 2u * kWideStackCapacity;
constexpr uint32_t kWcSecond = kWideStackWords + 0, kWcHitRec = kWideStackWords + 1, kWcHitTri = kWideStackWords + 2, kWcHitU = kWideStackWords + 3,
                   kWcHitV = kWideStackWords + 4, kWcGate = kWideStackWords + 5, kWcInstGate = kWideStackWords + 6;

struct WideLane {
    V3 ro, rd, ridir;          /* current ray: identity-local at the top level, instance-local inside a ray-switched instance */
    float best, lam;
    uint32_t ngBase, ngHits;   /* node group: childBase of a visited node, (its interior hits in visiting order) << 24 | imask */
    uint32_t lg;               /* leaf group: node << 8 | hit leaf slots */
    uint32_t triNext, triEnd;  /* triangles of the leaf being tested */
    float leafNear, leafSlack; /* its box entry distance and slack */
    int sp;                    /* next free stack entry; entry 0 is the sentinel; 0 = ray finished */
    uint32_t inst;             /* instance record inside a ray-switched instance; kTraceTop at the top level */
    uint32_t path;
    uint32_t octInv;           /* 7 - direction octant (bit 2: -x, bit 1: -y, bit 0: -z) of the current ray */
    uint32_t flags;
#ifdef OKRT_TRACE_STATS
    uint32_t stNodes, stLeaves, stTris;
#endif
};

OKRT_HD bool wideWantsTri(const WideLane& L) { return (L.lg & 0xFFu) != 0u || L.triNext < L.triEnd; }
OKRT_HD bool wideWantsNode(const WideLane& L) { return L.sp > 0 && !wideWantsTri(L); }
OKRT_HD bool wideDone(const WideLane& L) { return L.sp == 0; }

OKRT_HD uint32_t wideOctInv(V3 ridir)
{
    /* sign bits: 1/(-0) = -inf keeps the sign */
    const uint32_t oct = ((wBits(ridir.x) >> 31) << 2) | ((wBits(ridir.y) >> 31) << 1) | (wBits(ridir.z) >> 31);
    return 7u - oct;
}
OKRT_HD void wideTopRay(WideLane& L, V3 o, V3 d)
{
    const float4 i0 = make_float4(1, 0, 0, 0), i1 = make_float4(0, 1, 0, 0), i2 = make_float4(0, 0, 1, 0);
    L.ro = mulPointRows(o, i0, i1, i2);                        /* the reference's local ray of an identity instance */
    L.rd = normalize(mulDirRows(d, i0, i1, i2));
    L.ridir = rcp3(L.rd);
    L.inst = kTraceTop;
    L.octInv = wideOctInv(L.ridir);
}
OKRT_HD void widePush(WideLane& L, uint32_t* stack, uint32_t x, uint32_t y)
{
    OKRT_CHECK(L.sp >= 1 && L.sp < (int)kWideStackCapacity);
    stack[2 * L.sp] = x; stack[2 * L.sp + 1] = y; L.sp++;
}
/* "enter node `node`": a group whose only pending child is slot 0 of an imask-less parent */
OKRT_HD void wideEnterRoot(WideLane& L, uint32_t node)
{
    L.ngBase = node; L.ngHits = 1u << (24u + L.octInv);
}

/* a hit at world distance t in the leaf the lane stands on (or a sphere: leafNear = t, leafSlack = 0) */
OKRT_HD void wideCandidate(WideLane& L, uint32_t* stack, float t, uint32_t rec, uint32_t tri, float u, float v)
{
    const float n = L.leafNear;
    if (t < (n - L.leafSlack) * (1.0f - 0.5f * OKRT_DELTA)) L.flags |= kWideFlagUnbounded;          /* A1 violated */
    const float gate = fmaxf(t, n);
    if (t < L.best) {
        if (gate >= L.lam * (1.0f - 0.5f * OKRT_DELTA)) L.flags |= kWideFlagAmbiguous;              /* boxes were skipped that may hold something <= gate */
        stack[kWcSecond] = wBits(L.best); L.best = t;
        stack[kWcHitRec] = rec; stack[kWcHitTri] = tri; stack[kWcHitU] = wBits(u); stack[kWcHitV] = wBits(v); stack[kWcGate] = wBits(gate);
        L.lam = fminf(L.lam, gate * (1.0f + 2.0f * OKRT_DELTA));
    } else if (t < wFloat(stack[kWcSecond])) stack[kWcSecond] = wBits(t);
}

OKRT_HD void wideInit(const DeviceScene& s, WideLane& L, uint32_t* stack, uint32_t numSpheres, V3 o, V3 d)
{
    L.best = OKRT_FLT_MAX; L.lam = OKRT_FLT_MAX; L.flags = 0u;
    stack[kWcSecond] = wBits(OKRT_FLT_MAX); stack[kWcHitRec] = 0x3FFFFFFFu; stack[kWcHitTri] = 0xFFFFFFFFu;
    stack[kWcHitU] = 0u; stack[kWcHitV] = 0u; stack[kWcGate] = 0u; stack[kWcInstGate] = 0u;
    stack[0] = kTraceDone; stack[1] = 0u; L.sp = 1;
    L.lg = 0u; L.triNext = 0u; L.triEnd = 0u; L.leafNear = 0.f; L.leafSlack = 0.f;
    wideTopRay(L, o, d);
    for (uint32_t i = 0; i < numSpheres; i++) {                /* RaytracerCS.hlsl:232-242 */
        const okrt_sphere& sp = s.spheres[i];
        const float t = rayAndSphere(o, d, v3(sp.position), sp.radius);
        if (t > 0.f) { L.leafNear = t; L.leafSlack = 0.f; wideCandidate(L, stack, t, (1u << 30) | i, 0xFFFFFFFFu, 0.f, 0.f); }
    }
    L.ngBase = 0u; L.ngHits = 0u;
    if (s.wideRoot != kRefNone) wideEnterRoot(L, s.wideRoot);
#ifdef OKRT_TRACE_STATS
    L.stNodes = L.stLeaves = L.stTris = 0u;
#endif
}

/* ------------------------------------------------------------------ one wide node
 * Quantised slab tests of the eight children; returns the hit slots.  Per axis k the child planes are p + q * cell (exact in
 * float32); with A = cell / d and C = (p - o) / d a plane's ray distance is q A + C.  The kernel evaluates
 * fma(32768 + q, A, C - 32768 A): the byte goes straight into the mantissa of 32768.0f (one PRMT, no conversion).  `err` bounds
 * the rounding of C (2 ulp), of C - 32768 A (2^-9 |A|) and of the fma (1 ulp of |C| + 255 |A|), and widens the slab by that
 * much on both sides, so the quantised test never rejects a ray whose exact leaf test would pass.  1/d is clamped to +-2^100:
 * an axis-parallel ray gives huge finite distances of the right sign instead of inf - inf. */
OKRT_HD uint32_t wideTestNode(const uint4& n0, const uint4& n1, const uint4& n2, const uint4& n3, const uint4& n4, V3 ro, V3 ridir, float lam)
{
    const float big = 1.2676506e30f;                           /* 2^100 */
    const V3 ic = v3(fminf(fmaxf(ridir.x, -big), big), fminf(fmaxf(ridir.y, -big), big), fminf(fmaxf(ridir.z, -big), big));
    const float ax = ic.x * wFloat((n0.w & 0xFFu) << 23), ay = ic.y * wFloat(((n0.w >> 8) & 0xFFu) << 23), az = ic.z * wFloat(((n0.w >> 16) & 0xFFu) << 23);
    const float cx = (wFloat(n0.x) - ro.x) * ic.x, cy = (wFloat(n0.y) - ro.y) * ic.y, cz = (wFloat(n0.z) - ro.z) * ic.z;
    const float err = wMax3(wFma(fabsf(ax), 5500.f, fabsf(cx)), wFma(fabsf(ay), 5500.f, fabsf(cy)), wFma(fabsf(az), 5500.f, fabsf(cz))) * 4.0e-7f;
    const float mx = wFma(-32768.f, ax, cx), my = wFma(-32768.f, ay, cy), mz = wFma(-32768.f, az, cz);
    const float lx = mx - err, ly = my - err, lz = mz - err, hx = mx + err, hy = my + err, hz = mz + err;
    /* near / far plane bytes by direction sign: words n1.zw = qlo.x, n2.xy = qlo.y, n2.zw = qlo.z, n3.xy = qhi.x, n3.zw = qhi.y, n4.xy = qhi.z */
    const bool nx = ic.x < 0.f, ny = ic.y < 0.f, nz = ic.z < 0.f;
    const uint32_t nearX0 = nx ? n3.x : n1.z, nearX1 = nx ? n3.y : n1.w, farX0 = nx ? n1.z : n3.x, farX1 = nx ? n1.w : n3.y;
    const uint32_t nearY0 = ny ? n3.z : n2.x, nearY1 = ny ? n3.w : n2.y, farY0 = ny ? n2.x : n3.z, farY1 = ny ? n2.y : n3.w;
    const uint32_t nearZ0 = nz ? n4.x : n2.z, nearZ1 = nz ? n4.y : n2.w, farZ0 = nz ? n2.z : n4.x, farZ1 = nz ? n2.w : n4.y;
    const float lamBelow = wFloat(wBits(lam) - 1u);            /* near < lam  <=>  near <= the float below lam (lam > 0) */
    const float tiny = wFloat(1u);                              /* far > 0     <=>  far >= the smallest denormal */
    uint32_t hits = 0u;
#define OKRT_WIDE_CHILD(k, NX, NY, NZ, FX, FY, FZ, bit)                                                                  \
    {                                                                                                                    \
        const float tn = wMax3(wFma(wPlane<k>(NX), ax, lx), wFma(wPlane<k>(NY), ay, ly), wFma(wPlane<k>(NZ), az, lz));   \
        const float tf = wMin3(wFma(wPlane<k>(FX), ax, hx), wFma(wPlane<k>(FY), ay, hy), wFma(wPlane<k>(FZ), az, hz));   \
        if (fmaxf(tn, tiny) <= fminf(tf, lamBelow)) hits |= (bit);                                                       \
    }
    OKRT_WIDE_CHILD(0, nearX0, nearY0, nearZ0, farX0, farY0, farZ0, 1u)
    OKRT_WIDE_CHILD(1, nearX0, nearY0, nearZ0, farX0, farY0, farZ0, 2u)
    OKRT_WIDE_CHILD(2, nearX0, nearY0, nearZ0, farX0, farY0, farZ0, 4u)
    OKRT_WIDE_CHILD(3, nearX0, nearY0, nearZ0, farX0, farY0, farZ0, 8u)
    OKRT_WIDE_CHILD(0, nearX1, nearY1, nearZ1, farX1, farY1, farZ1, 16u)
    OKRT_WIDE_CHILD(1, nearX1, nearY1, nearZ1, farX1, farY1, farZ1, 32u)
    OKRT_WIDE_CHILD(2, nearX1, nearY1, nearZ1, farX1, farY1, farZ1, 64u)
    OKRT_WIDE_CHILD(3, nearX1, nearY1, nearZ1, farX1, farY1, farZ1, 128u)
#undef OKRT_WIDE_CHILD
    return hits;
}

/* bit j of the result = bit (j ^ x) of m, for 8-bit m: the visiting priority of a slot is slot ^ octInv */
OKRT_HD uint32_t widePermute(uint32_t m, uint32_t x)
{
    const uint32_t a = ((m & 0x55u) << 1) | ((m >> 1) & 0x55u); m = (x & 1u) ? a : m;
    const uint32_t b = ((m & 0x33u) << 2) | ((m >> 2) & 0x33u); m = (x & 2u) ? b : m;
    const uint32_t c = ((m & 0x0Fu) << 4) | ((m >> 4) & 0x0Fu); m = (x & 4u) ? c : m;
    return m;
}

/* One node step: take the nearest pending child of the lane's node group (popping a group first if it has none), test its
 * eight children.  rayO / rayD: the path's world ray, needed when an instance's exit marker comes off the stack. */
OKRT_HD void wideNodeStep(const DeviceScene& s, WideLane& L, uint32_t* stack, const float4* rayO, const float4* rayD)
{
    if ((L.ngHits >> 24) == 0u) {
        OKRT_CHECK(L.sp >= 1);
        L.sp--;
        const uint32_t x = stack[2 * L.sp], y = stack[2 * L.sp + 1];
        if ((y >> 24) != 0u) { L.ngBase = x; L.ngHits = y; }
        else {
            if (x == kTraceDone) { L.sp = 0; }
            else if (x == kTraceExitInstance) {
                const float4 o = rayO[L.path], d = rayD[L.path];
                wideTopRay(L, v3(o.x, o.y, o.z), v3(d.x, d.y, d.z));
                stack[kWcInstGate] = 0u;
            } else L.lg = x;                                    /* the leaves that waited while the instance was walked */
            return;
        }
    }
    const uint32_t bit = wHighBit(L.ngHits);
    L.ngHits &= ~(1u << bit);
    const uint32_t slot = (bit - 24u) ^ L.octInv;
    const uint32_t node = L.ngBase + wPopc(L.ngHits & 0xFFu & ((1u << slot) - 1u));
    if ((L.ngHits >> 24) != 0u) widePush(L, stack, L.ngBase, L.ngHits);
    OKRT_CHECK(node < s.numWideNodes);
    const uint4* q = s.wideNodes + (size_t)node * 6;
    uint4 n0, n1, n2, n3, n4, n5;
    wLoad256(q, n0, n1); wLoad256(q + 2, n2, n3); wLoad256(q + 4, n4, n5);
    const uint32_t hits = wideTestNode(n0, n1, n2, n3, n4, L.ro, L.ridir, L.lam);
    const uint32_t imask = n0.w >> 24;
    L.ngBase = n1.x;
    L.ngHits = (widePermute(hits & imask, L.octInv) << 24) | imask;
    L.lg = (node << 8) | (hits & ~imask);
#ifdef OKRT_TRACE_STATS
    L.stNodes++;
#endif
}

/* oct-tree boxes on the way to an instance, world ray (RaytracerCS.hlsl:262): all must be HIT; their largest distNear joins
 * the gate of what is found inside */
OKRT_HD bool wideOctChain(const DeviceScene& s, V3 o, V3 d, uint32_t octNode, float& maxNear)
{
    const V3 idirWorld = rcp3(d);
    maxNear = -OKRT_FLT_MAX;
    for (uint32_t on = octNode; on != 0xFFFFFFFFu; on = s.octParent[on]) {
        const okrt_octree_node& node = s.octNodes[on];
        const float n = rayAndAABBDist(o, idirWorld, v3(node.boundingBox.min), v3(node.boundingBox.max));
        if (!(n < OKRT_FLT_MAX)) return false;
        maxNear = fmaxf(maxNear, n);
    }
    return true;
}

/* the lane reached the leaf record of an instance that is walked with its own (local) ray */
OKRT_HD void wideEnterInstance(const DeviceScene& s, WideLane& L, uint32_t* stack, uint32_t recIdx, const float4* rayO, const float4* rayD)
{
    OKRT_CHECK(recIdx < s.numFastInstances);
    const float4* rec = s.fastInstances + (size_t)recIdx * kInstQuads;
    const float4 r6 = wLdg(rec + 6), r7 = wLdg(rec + 7), r8 = wLdg(rec + 8);
    if ((wBits(r6.w) & 2u) != 0u) {
        /* order-dependent instance (its oct-tree boxes do not contain it, or its tree is not nested): the reference order decides */
        L.flags |= kWideFlagUnbounded;
        L.ngHits = 0u; L.lg = 0u; L.triNext = L.triEnd = 0u; L.sp = 0;
        return;
    }
    const float4 o4 = rayO[L.path], d4 = rayD[L.path];
    const V3 o = v3(o4.x, o4.y, o4.z), d = v3(d4.x, d4.y, d4.z);
    float octNear;
    if (!wideOctChain(s, o, d, wBits(r7.y), octNear)) return;
    const float4 i0 = wLdg(rec), i1 = wLdg(rec + 1), i2 = wLdg(rec + 2);
    const V3 lo = mulPointRows(o, i0, i1, i2);
    const V3 ld = normalize(mulDirRows(d, i0, i1, i2));
    const V3 lidir = rcp3(ld);
    const float4 a0 = wLdg(rec + 3), a1 = wLdg(rec + 4), a2 = wLdg(rec + 5);
    /* the mesh's root box (RaytracerCS.hlsl:285 on the first popped node) */
    const float rootNear = rayAndAABBDist(lo, lidir, v3(a0.w, a1.w, a2.w), v3(r6.x, r6.y, r6.z));
    const uint32_t root = wBits(r8.w);
    if (!(rootNear < L.lam) || root == kRefNone) return;
    if ((L.ngHits >> 24) != 0u) widePush(L, stack, L.ngBase, L.ngHits);
    if ((L.lg & 0xFFu) != 0u) widePush(L, stack, L.lg, 0u);
    widePush(L, stack, kTraceExitInstance, 0u);
    L.ro = lo; L.rd = ld; L.ridir = lidir; L.inst = recIdx; L.octInv = wideOctInv(lidir);
    stack[kWcInstGate] = wBits(fmaxf(octNear, 0.f));
    L.lg = 0u;
    wideEnterRoot(L, root);
}

/* One triangle step: enter the next hit leaf of the lane's leaf group if no triangle is pending (exact box test), then test
 * one triangle. */
OKRT_HD void wideTriStep(const DeviceScene& s, WideLane& L, uint32_t* stack, const float4* rayO, const float4* rayD)
{
    if (L.triNext == L.triEnd) {
        const uint32_t slot = wLowBit(L.lg & 0xFFu);
        L.lg &= ~(1u << slot);
        uint4 b0, b1;
        wLoad256(s.wideLeaves + ((size_t)(L.lg >> 8) * 8u + slot) * 2u, b0, b1);
        /* GPU-Utilities.hlsli:183-197 on the reference's own box */
        const float n = rayAndAABBDist(L.ro, L.ridir, v3(wFloat(b0.x), wFloat(b0.y), wFloat(b0.z)), v3(wFloat(b0.w), wFloat(b1.x), wFloat(b1.y)));
        const float slack = wFloat(b1.w);
#ifdef OKRT_TRACE_STATS
        L.stLeaves++;
#endif
        if (n < OKRT_FLT_MAX && n - slack < L.lam) {
            const uint32_t count = b1.z >> 26, first = b1.z & 0x03FFFFFFu;
            if (count == kWideLeafInstance) { wideEnterInstance(s, L, stack, first, rayO, rayD); return; }
            L.triNext = first; L.triEnd = first + count;
            L.leafNear = L.inst == kTraceTop ? n : fmaxf(n, wFloat(stack[kWcInstGate]));
            L.leafSlack = slack;
        }
    }
    if (L.triNext < L.triEnd) {
        const uint32_t j = L.triNext++;
        OKRT_CHECK(j < s.numTris);
        float4 a, b, c;
        wTriangle(s, j, a, b, c);                                /* p2 | E1 | E2 | owner */
        float bu = 0.f, bv = 0.f;
        float t = rayAndTriangleEdges(L.ro, L.rd, v3(a.x, a.y, a.z), v3(a.w, b.x, b.y), v3(b.z, b.w, c.x), bu, bv);
#ifdef OKRT_TRACE_STATS
        L.stTris++;
#endif
        if (t > 0.f) {
            V3 w = L.rd * t;                                     /* RaytracerCS.hlsl:312-314; T == I for merged instances */
            uint32_t rec = wBits(c.y);                           /* owner of a merged instance's triangle */
            if (L.inst != kTraceTop) {
                const float4* r = s.fastInstances + (size_t)L.inst * kInstQuads;
                w = mulDirRows(w, wLdg(r + 3), wLdg(r + 4), wLdg(r + 5));
                rec = L.inst;
            }
            t = length(w);
            wideCandidate(L, stack, t, (2u << 30) | rec, j, bu, bv);
        }
    }
}

/* The finished lane's verdict (see WideLane).  Returns true when the reference order has to decide; `limit` then bounds
 * that traversal.  L stands at the top level again, so L.ro is the world origin. */
OKRT_HD bool wideResult(const DeviceScene& s, const WideLane& L, const uint32_t* stack, const float4* rayD, Hit& hit, float& limit)
{
    const uint32_t hitRec = stack[kWcHitRec];
    const float second = wFloat(stack[kWcSecond]);
    hit.tri = stack[kWcHitTri]; hit.u = wFloat(stack[kWcHitU]); hit.v = wFloat(stack[kWcHitV]);
    hit.distance = L.best;
    hit.kind = L.best < OKRT_FLT_MAX ? (hitRec >> 30) : 0u;
    hit.idx = hit.kind ? (hitRec & 0x3FFFFFFFu) : 0xFFFFFFFFu;
    float gate = fmaxf(L.best, wFloat(stack[kWcGate]));
    bool unbounded = (L.flags & kWideFlagUnbounded) != 0u;
    if (hit.kind == 2u) {
        const uint32_t rec = hit.idx;
        OKRT_CHECK(rec < s.numFastInstances);
        const float4* r = s.fastInstances + (size_t)rec * kInstQuads;
        const float4 r6 = wLdg(r + 6), r7 = wLdg(r + 7), lo = wLdg(r + 8), hi = wLdg(r + 9);
        hit.idx = wBits(r7.z);
        if ((wBits(r6.w) & 1u) != 0u && L.inst == kTraceTop) {
            /* merged identity instance: the reference reaches it only through the oct-tree boxes above it (RaytracerCS.hlsl:262).
             * An origin strictly inside a box hits it with distNear <= 0 — both slabs of every axis have opposite signs — and the
             * boxes above are nested. */
            const float m = 1.0e-20f;
            const bool inside = (wBits(r7.w) & 1u) != 0u &&
                                L.ro.x - lo.x > m && hi.x - L.ro.x > m && L.ro.y - lo.y > m && hi.y - L.ro.y > m && L.ro.z - lo.z > m && hi.z - L.ro.z > m;
            if (!inside) {
                const float4 d = *rayD;
                float octNear;
                if (!wideOctChain(s, L.ro, v3(d.x, d.y, d.z), wBits(r7.y), octNear)) unbounded = true;
                else gate = fmaxf(gate, octNear);
            }
        }
    }
    const bool ambiguous = unbounded || (hit.kind != 0u && ((L.flags & kWideFlagAmbiguous) != 0u || second <= gate * (1.0f + OKRT_DELTA)));
    limit = unbounded ? OKRT_FLT_MAX : fminf(gate * (1.0f + 16.0f * OKRT_DELTA) + 4.0f * s.maxSlack, OKRT_FLT_MAX);
    return ambiguous;
}

/* whole-ray form: the megakernel and the CPU check */
OKRT_HD bool wideTraverse(const DeviceScene& s, uint32_t numSpheres, V3 o, V3 d, Hit& hit, float& limit, uint32_t* counts = nullptr)
{
    uint32_t stack[kWideStackWords + kWideColdWords];
    WideLane L;
    L.path = 0u;
    const float4 o4 = make_float4(o.x, o.y, o.z, 0.f), d4 = make_float4(d.x, d.y, d.z, 0.f);
    wideInit(s, L, stack, numSpheres, o, d);
    while (!wideDone(L)) {
        if (wideWantsTri(L)) wideTriStep(s, L, stack, &o4, &d4);
        else wideNodeStep(s, L, stack, &o4, &d4);
    }
#ifdef OKRT_TRACE_STATS
    if (counts) { counts[0] = L.stNodes; counts[1] = L.stLeaves; counts[2] = L.stTris; }
#endif
    if (counts) counts[3] = L.flags;
    return wideResult(s, L, stack, &d4, hit, limit);
}

} // namespace okrt

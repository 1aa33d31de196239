/*
 * okrt_device.cu — the renderer object behind the C ABI (include/okrt.h): device buffers for the
 * reference's binding table, the re-packed acceleration layout, launch of the path-trace kernel.
 * Replaces RayTracer::{initiate, render, updateBuffers, onResize, resetAccumulation}
 * (source/Graphics/RayTracer.cpp:60-118,219-260,434-501; RayTracer.h:197-211) and GPUStorage
 * (source/Graphics/GPUStorage.cpp:27-50) with cudaMalloc-backed buffers on one stream per handle.
 *
 * There is no CPU fallback: without a CUDA device okrt_create fails with OKRT_ERR_NO_DEVICE.
 */
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <memory>
#include <thread>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <mutex>
#include <sys/resource.h>
#include <sys/syscall.h>
#include <unistd.h>
#include <vector>

#include "okrt.h"
#include "okrt_accel.h"
#include "okrt_internal.h"
#include "okrt_layout_device.h"
#include "okrt_kernels.cuh"
#include "okrt_wavefront.cuh"

using namespace okrt;

namespace {

struct CudaError { cudaError_t code; const char* what; };

#define OKRT_CUDA(call)                                                                      \
    do {                                                                                     \
        cudaError_t e__ = (call);                                                            \
        if (e__ != cudaSuccess) {                                                            \
            setError(std::string(#call) + ": " + cudaGetErrorString(e__));                   \
            return OKRT_ERR_CUDA;                                                            \
        }                                                                                    \
    } while (0)

/* GPUStorage equivalent: initiate / updateRaw with re-allocation when the element count changes */
struct DeviceBuffer {
    void* ptr = nullptr;
    size_t bytes = 0;
    uint32_t count = 0;

    int upload(const void* host, uint32_t n, size_t elemBytes, cudaStream_t stream)
    {
        const size_t need = (size_t)n * elemBytes;
        if (need != bytes) {
            if (ptr) { cudaFree(ptr); ptr = nullptr; bytes = 0; }
            if (need) OKRT_CUDA(cudaMalloc(&ptr, need));
            bytes = need;
        }
        count = n;
        if (need) {
            OKRT_CUDA(cudaMemcpyAsync(ptr, host, need, cudaMemcpyHostToDevice, stream));
            OKRT_CUDA(cudaStreamSynchronize(stream));         /* caller's array may be freed on return */
        }
        return OKRT_OK;
    }
    void release() { if (ptr) cudaFree(ptr); ptr = nullptr; bytes = 0; count = 0; }
};

struct TextureArray {
    cudaArray_t array = nullptr;
    cudaTextureObject_t object = 0;
    uint32_t width = 0, height = 0, layers = 0;

    void release()
    {
        if (object) cudaDestroyTextureObject(object);
        if (array) cudaFreeArray(array);
        object = 0; array = nullptr; width = height = layers = 0;
    }
    int upload(const uint8_t* rgba8, uint32_t w, uint32_t h, uint32_t n, cudaStream_t stream)
    {
        release();
        if (!w || !h || !n) return OKRT_OK;
        cudaChannelFormatDesc desc = cudaCreateChannelDesc<uchar4>();
        OKRT_CUDA(cudaMalloc3DArray(&array, &desc, make_cudaExtent(w, h, n), cudaArrayLayered));
        cudaMemcpy3DParms p;
        std::memset(&p, 0, sizeof(p));
        p.srcPtr = make_cudaPitchedPtr(const_cast<uint8_t*>(rgba8), (size_t)w * 4, w, h);
        p.dstArray = array;
        p.extent = make_cudaExtent(w, h, n);
        p.kind = cudaMemcpyHostToDevice;
        OKRT_CUDA(cudaMemcpy3DAsync(&p, stream));
        OKRT_CUDA(cudaStreamSynchronize(stream));
        cudaResourceDesc res;
        std::memset(&res, 0, sizeof(res));
        res.resType = cudaResourceTypeArray;
        res.res.array.array = array;
        cudaTextureDesc td;
        std::memset(&td, 0, sizeof(td));
        td.addressMode[0] = td.addressMode[1] = td.addressMode[2] = cudaAddressModeClamp;
        td.filterMode = cudaFilterModePoint;                  /* filtering is done in float32 in the kernel */
        td.readMode = cudaReadModeElementType;
        td.normalizedCoords = 0;
        OKRT_CUDA(cudaCreateTextureObject(&object, &res, &td, nullptr));
        width = w; height = h; layers = n;
        return OKRT_OK;
    }
};

} // namespace

/* The device builder's clustered tree is ready in milliseconds but costs ~18 % of the rays/s of the host builder's full-sweep SAH
 * tree (measured).  After a device build the host builder therefore runs in the background on the same arrays;
 * its layout replaces the clustered one at the next okrt_render* call after it finishes (okrt_wait_layout waits for it).
 * Images do not change: both trees hold the reference's leaves, the visiting order is free. */
struct RefineJob {
    std::shared_ptr<const std::vector<okrt_triangle>> tris;
    std::shared_ptr<const std::vector<okrt_bvh_node>> nodes;
    std::shared_ptr<const std::vector<okrt_mesh_instance>> instances;
    std::shared_ptr<const std::vector<okrt_octree_node>> oct;
    FastLayout layout;
    bool ok = false;
    float ms = 0.f;
    std::atomic<bool> done{false};
};

/* BATCH LANES.  A batch (generate -> (trace, fallback, shade) x bounces -> accumulate over its own path state) runs on one of
 * up to kMaxLanes streams; consecutive batches — of one okrt_render_frames call or of consecutive calls — take the lanes in turn,
 * so that the thin late bounces of batch k run beside the full early bounces of batch k+1 (every kernel pulls its work from a
 * device counter and makes progress with whatever share of the machine it gets).  What orders them:
 *   - the accumulate stage of a batch waits for an event recorded on the renderer's main stream when the call began; the main
 *     stream has by then been made to wait for every earlier batch (end of okrt_render_frames) and carries whatever else the
 *     caller queued (result snapshots, resets, overlays): frames reach u1 in frame order, u0 is never overwritten under a reader;
 *   - a lane's chain START waits for the main stream only after a call that may have changed what the chain reads
 *     (`mainTouched`: every entry point except render / set_render_data / read_result_async / wait_read): the steady
 *     render-and-read loop never serialises its batches, an edit always does.
 * One GPU and N tile-sharded GPUs run the same code; what changes with N is only how small a 16-frame batch is. */
constexpr int kMaxLanes = 4;
struct BatchLane {
    cudaStream_t stream = nullptr, streamBeside = nullptr;   /* the chain; the reference-order kernel beside the main shade launch */
    cudaEvent_t evFork = nullptr, evJoin = nullptr;          /* chain <-> beside */
    cudaEvent_t evDone = nullptr;                            /* after the lane's latest accumulate */
    WavefrontBuffers wb{};                                   /* path state of the lane's batch */
    size_t wbPaths = 0;
    uint64_t forkSeen = 0;                                   /* generation of the main-stream fork the lane has waited for */
};

struct okrt_renderer {
    int device = 0;
    cudaStream_t stream = nullptr;             /* main stream: uploads, reads, resets; always downstream of every batch issued */
    BatchLane lanes[kMaxLanes];
    uint32_t nextLane = 0;
    int lastLane = -1;                         /* lane of the most recent batch */
    bool mainTouched = true;                   /* the next batch's chain must wait for the main stream */
    uint64_t forkGen = 0;
    cudaEvent_t evMainFork = nullptr, evMainNow = nullptr;
    unsigned long long* traceStats = nullptr;  /* OKRT_TRACE_STATS builds: 8 counters per bounce, all lanes add to it */
    cudaStream_t streamCopy = nullptr;         /* okrt_read_result_async: device-to-host copies overlap the next frames */
    uchar4* snapshot[OKRT_READ_SLOTS] = {};    /* result image as of the call, one per slot */
    cudaEvent_t evSnapshot[OKRT_READ_SLOTS] = {}, evCopied[OKRT_READ_SLOTS] = {};
    bool copyPending[OKRT_READ_SLOTS] = {};
    cudaEvent_t evStart = nullptr, evStop = nullptr;
    bool timingPending = false;                /* evStart goes onto the lane of the call's first batch */
    std::vector<cudaEvent_t> traceEvents;      /* pairs around every traversal-kernel launch of the last render call */
    size_t traceEventsUsed = 0;
    uint32_t width = 0, height = 0;

    DeviceBuffer triPos, triInfo, shadeInfo, bvhNodes, octNodes, spheres, instances, dirLights, pointLights, spotLights;
    TextureArray textures, environment;
    DeviceBuffer fastNodes, fastTris, fastInstances, octParent, wideNodes, wideLeaves;
    uint32_t wideRoot = kRefNone;
    float maxSlack = 0.f;
    cudaTextureObject_t fastTrisTex = 0;

    /* host copies needed to (re)build the fast layout */
    /* shared, immutable once set: a background refinement (RefineJob) may still read the arrays of an older edit */
    std::shared_ptr<const std::vector<okrt_triangle>> hTris = std::make_shared<std::vector<okrt_triangle>>();
    std::shared_ptr<const std::vector<okrt_bvh_node>> hNodes = std::make_shared<std::vector<okrt_bvh_node>>();
    std::shared_ptr<const std::vector<okrt_mesh_instance>> hInstances = std::make_shared<std::vector<okrt_mesh_instance>>();
    std::shared_ptr<const std::vector<okrt_octree_node>> hOct = std::make_shared<std::vector<okrt_octree_node>>();
    std::shared_ptr<struct RefineJob> refine;   /* host SAH build of the layout the device builder has just put in place */
    bool layoutDirty = true;
    uint32_t topRef = kRefNone;
    bool allRigid = true;
    uint32_t fastStackNeed = 0, exactBvhNeed = 0, exactOctNeed = 0;

    float4* accumulation = nullptr;
    uchar4* result = nullptr;
    okrt_aov* aov = nullptr;
    unsigned long long* counters = nullptr;     /* 4 x u64 on device */

    okrt_render_data rd;
    okrt_options opt;
    uint64_t launches = 0;
    float lastRenderMs = 0.f;
    int numSms = 0, blocksPerSmLean = 1, blocksPerSmInstrumented = 1;   /* persistent grid = SMs x resident blocks */
    int blocksPerSmTrace = 1, blocksPerSmWide = 1;
    bool layoutHasWide = false;                 /* the wide layout is built only for OKRT_TRAVERSAL_WIDE */
    std::vector<uint32_t> ownedBands;           /* okrt_set_shard_bands: this rank's bands (ascending); empty: interleaved by arithmetic */
    DeviceBuffer dOwnedBands;
    DeviceBuffer ovSegments, ovKeys, ovColours, ovPixelKey;   /* okrt_draw_overlay scratch */
    okrt_layout_build layoutBuild{};            /* how the last rebuild went (okrt_get_layout_build) */
};

namespace {

int allocateImages(okrt_renderer* r)
{
    const size_t n = (size_t)r->width * r->height;
    if (r->accumulation) cudaFree(r->accumulation);
    if (r->result) cudaFree(r->result);
    if (r->aov) cudaFree(r->aov);
    r->accumulation = nullptr; r->result = nullptr; r->aov = nullptr;
    OKRT_CUDA(cudaMalloc(&r->accumulation, n * sizeof(float4)));
    OKRT_CUDA(cudaMalloc(&r->result, n * sizeof(uchar4)));
    OKRT_CUDA(cudaMemsetAsync(r->accumulation, 0, n * sizeof(float4), r->stream));
    clearResultKernel<<<(unsigned)((n + 255) / 256), 256, 0, r->stream>>>(r->result, n);
    r->launches++;
    OKRT_CUDA(cudaGetLastError());
    return OKRT_OK;
}

int ensureAov(okrt_renderer* r)
{
    if (r->aov) return OKRT_OK;
    const size_t n = (size_t)r->width * r->height;
    OKRT_CUDA(cudaMalloc(&r->aov, n * sizeof(okrt_aov)));
    OKRT_CUDA(cudaMemsetAsync(r->aov, 0, n * sizeof(okrt_aov), r->stream));
    return OKRT_OK;
}

int finishLayout(okrt_renderer* r);
int installHostLayout(okrt_renderer* r, const FastLayout& fl, bool wantWide);
void startRefine(okrt_renderer* r);

/* the layout built on the device where the scene allows it (okrt_layout_device.cu) */
int rebuildLayoutDevice(okrt_renderer* r, std::string& why)
{
    DeviceLayout dl;
    const int rc = buildFastLayoutDevice((const okrt_triangle*)r->triPos.ptr, (uint32_t)r->hTris->size(), (const okrt_bvh_node*)r->bvhNodes.ptr, r->hNodes->data(),
                                         (uint32_t)r->hNodes->size(), r->hInstances->data(), (uint32_t)r->hInstances->size(), r->hOct->data(), (uint32_t)r->hOct->size(),
                                         r->stream, dl, why);
    if (rc) return rc;
    if (dl.exactBvhDepth > kExactBvhStackCapacity || dl.exactOctDepth > kExactOctStackCapacity) {
        cudaFree(dl.nodes); cudaFree(dl.tris);
        why = "reference-order stacks would overflow (the host builder reports it)";
        return OKRT_ERR_STATE;
    }
    r->fastNodes.release(); r->fastTris.release();
    r->fastNodes.ptr = dl.nodes; r->fastNodes.count = dl.numPairNodes * 4u; r->fastNodes.bytes = (size_t)r->fastNodes.count * sizeof(Float4);
    r->fastTris.ptr = dl.tris; r->fastTris.count = (uint32_t)r->hTris->size() * kTriQuads; r->fastTris.bytes = (size_t)r->fastTris.count * sizeof(Float4);
    int rc2;
    if ((rc2 = r->fastInstances.upload(dl.instances.data(), (uint32_t)dl.instances.size(), sizeof(Float4), r->stream))) return rc2;
    if ((rc2 = r->octParent.upload(dl.octParent.data(), (uint32_t)dl.octParent.size(), sizeof(uint32_t), r->stream))) return rc2;
    r->wideNodes.release(); r->wideLeaves.release(); r->wideRoot = kRefNone; r->maxSlack = 0.f; r->layoutHasWide = false;
    r->topRef = dl.topRef; r->allRigid = dl.allRigid;
    r->fastStackNeed = dl.maxStackDepth; r->exactBvhNeed = dl.exactBvhDepth; r->exactOctNeed = dl.exactOctDepth;
    r->layoutBuild.builder = 1u; r->layoutBuild.deviceMs = dl.deviceMs; r->layoutBuild.numPairNodes = dl.numPairNodes;
    r->layoutBuild.fastStackDepth = dl.maxStackDepth; r->layoutBuild.rounds = dl.plocRounds;
    return finishLayout(r);
}

int rebuildLayout(okrt_renderer* r)
{
    const auto t0 = std::chrono::steady_clock::now();
    std::memset(&r->layoutBuild, 0, sizeof(r->layoutBuild));
    r->refine.reset();                                                  /* a refinement of older arrays is of no use any more */
    const char* hostEnv = std::getenv("OKRT_LAYOUT_HOST");            /* read at every rebuild: tests switch builders in one process */
    const bool forceHost = hostEnv && std::atoi(hostEnv) != 0;
    std::string why = forceHost ? "OKRT_LAYOUT_HOST=1" : (r->opt.traversal == OKRT_TRAVERSAL_WIDE ? "OKRT_TRAVERSAL_WIDE is laid out by the host builder" : "");
    if (why.empty()) {
        const int rc = rebuildLayoutDevice(r, why);
        if (rc == OKRT_OK) {
            r->layoutBuild.wallMs = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
            startRefine(r);
            return OKRT_OK;
        }
        if (rc != OKRT_ERR_STATE) return rc;
    }
    std::snprintf(r->layoutBuild.why, sizeof(r->layoutBuild.why), "%s", why.c_str());
    FastLayout fl;
    const bool wantWide = r->opt.traversal == OKRT_TRAVERSAL_WIDE;
    if (!buildFastLayout(r->hTris->data(), (uint32_t)r->hTris->size(), r->hNodes->data(), (uint32_t)r->hNodes->size(),
                         r->hInstances->data(), (uint32_t)r->hInstances->size(), r->hOct->data(), (uint32_t)r->hOct->size(), fl, wantWide)) {
        setError("scene rejected: " + fl.error);
        return OKRT_ERR_INVALID_ARGUMENT;
    }
    int rc = installHostLayout(r, fl, wantWide);
    if (rc) return rc;
    r->layoutBuild.builder = 0u;
    r->layoutBuild.wallMs = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return OKRT_OK;
}

/* uploads a layout the host builder made and makes it the one the kernels walk */
int installHostLayout(okrt_renderer* r, const FastLayout& fl, bool wantWide)
{
    if (fl.maxStackDepth > kFastStackCapacity || fl.wideStackDepth > kWideStackCapacity || fl.exactBvhDepth > kExactBvhStackCapacity || fl.exactOctDepth > kExactOctStackCapacity) {
        char msg[256];
        std::snprintf(msg, sizeof(msg), "traversal stack would overflow: pairs %u/%u, wide %u/%u, bvh %u/%u, oct-tree %u/%u",
                      fl.maxStackDepth, kFastStackCapacity, fl.wideStackDepth, kWideStackCapacity, fl.exactBvhDepth, kExactBvhStackCapacity, fl.exactOctDepth, kExactOctStackCapacity);
        setError(msg);
        return OKRT_ERR_CAPACITY;
    }
    int rc;
    if ((rc = r->fastNodes.upload(fl.nodes.data(), (uint32_t)fl.nodes.size(), sizeof(Float4), r->stream))) return rc;
    if ((rc = r->fastTris.upload(fl.tris.data(), (uint32_t)fl.tris.size(), sizeof(Float4), r->stream))) return rc;
    if ((rc = r->fastInstances.upload(fl.instances.data(), (uint32_t)fl.instances.size(), sizeof(Float4), r->stream))) return rc;
    if ((rc = r->octParent.upload(fl.octParent.data(), (uint32_t)fl.octParent.size(), sizeof(uint32_t), r->stream))) return rc;
    if (fl.wideNodes.size() / kWideNodeWords >= (1u << 24) - 1u) { setError("scene too large: the wide layout names nodes with 24 bits"); return OKRT_ERR_CAPACITY; }
    if ((rc = r->wideNodes.upload(fl.wideNodes.data(), (uint32_t)fl.wideNodes.size(), sizeof(uint32_t), r->stream))) return rc;
    if ((rc = r->wideLeaves.upload(fl.wideLeaves.data(), (uint32_t)fl.wideLeaves.size(), sizeof(uint32_t), r->stream))) return rc;
    r->wideRoot = fl.wideRoot; r->maxSlack = fl.maxSlack; r->layoutHasWide = wantWide;
    r->topRef = fl.topRef;
    r->allRigid = fl.allRigid;
    r->fastStackNeed = fl.maxStackDepth; r->exactBvhNeed = fl.exactBvhDepth; r->exactOctNeed = fl.exactOctDepth;
    r->layoutBuild.numPairNodes = (uint32_t)(fl.nodes.size() / 4); r->layoutBuild.fastStackDepth = fl.maxStackDepth;
    return finishLayout(r);
}

/* starts the host builder in the background on the arrays of this edit (see RefineJob) */
void startRefine(okrt_renderer* r)
{
    const char* e = std::getenv("OKRT_LAYOUT_REFINE");
    if (e && std::atoi(e) == 0) { r->refine.reset(); return; }
    auto job = std::make_shared<RefineJob>();
    job->tris = r->hTris; job->nodes = r->hNodes; job->instances = r->hInstances; job->oct = r->hOct;
    r->refine = job;
    std::thread([job]() {
        /* lowest scheduling priority for this thread and the builder threads it starts (they inherit it): the refinement must
         * not take the cores from the thread that is feeding the GPU */
        setpriority(PRIO_PROCESS, (id_t)syscall(SYS_gettid), 19);
        const auto t0 = std::chrono::steady_clock::now();
        job->ok = buildFastLayout(job->tris->data(), (uint32_t)job->tris->size(), job->nodes->data(), (uint32_t)job->nodes->size(),
                                  job->instances->data(), (uint32_t)job->instances->size(), job->oct->data(), (uint32_t)job->oct->size(), job->layout, false);
        job->ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
        job->done.store(true, std::memory_order_release);
    }).detach();
}

/* installs the refined layout if it is ready (or waits for it); the job of an older edit was dropped when the edit arrived */
int adoptRefined(okrt_renderer* r, bool wait)
{
    if (!r->refine) return OKRT_OK;
    if (wait) while (!r->refine->done.load(std::memory_order_acquire)) std::this_thread::sleep_for(std::chrono::microseconds(200));
    if (!r->refine->done.load(std::memory_order_acquire)) return OKRT_OK;
    std::shared_ptr<RefineJob> job = r->refine;
    r->refine.reset();
    if (!job->ok || job->layout.maxStackDepth > kFastStackCapacity || !job->layout.allRigid) return OKRT_OK;       /* keep the clustered tree */
    OKRT_CUDA(cudaStreamSynchronize(r->stream));             /* downstream of every batch: the old layout is no longer read */
    const int rc = installHostLayout(r, job->layout, false);
    if (rc) return rc;
    r->mainTouched = true;
    r->layoutBuild.builder = 2u;
    r->layoutBuild.refineMs = job->ms;
    return OKRT_OK;
}

/* the triangle records as float4 texels: stepTriangle fetches them through the texture pipe */
int finishLayout(okrt_renderer* r)
{
    if (r->fastTrisTex) { cudaDestroyTextureObject(r->fastTrisTex); r->fastTrisTex = 0; }
    if (r->fastTris.bytes) {
        int maxTexels = 0;
        OKRT_CUDA(cudaDeviceGetAttribute(&maxTexels, cudaDevAttrMaxTexture1DLinearWidth, r->device));
        if (r->fastTris.bytes / sizeof(Float4) > (size_t)maxTexels) {
            char msg[160];
            std::snprintf(msg, sizeof(msg), "scene too large: %zu triangles, the triangle texture holds %d", r->fastTris.bytes / sizeof(Float4) / kTriQuads, maxTexels / (int)kTriQuads);
            setError(msg);
            return OKRT_ERR_CAPACITY;
        }
        cudaResourceDesc res; std::memset(&res, 0, sizeof(res));
        res.resType = cudaResourceTypeLinear; res.res.linear.devPtr = r->fastTris.ptr; res.res.linear.sizeInBytes = r->fastTris.bytes;
        res.res.linear.desc = cudaCreateChannelDesc<float4>();
        cudaTextureDesc td; std::memset(&td, 0, sizeof(td));
        td.readMode = cudaReadModeElementType; td.filterMode = cudaFilterModePoint; td.addressMode[0] = cudaAddressModeClamp;
        OKRT_CUDA(cudaCreateTextureObject(&r->fastTrisTex, &res, &td, nullptr));
    }
    r->layoutDirty = false;
    return OKRT_OK;
}

DeviceScene sceneView(const okrt_renderer* r)
{
    DeviceScene s;
    std::memset(&s, 0, sizeof(s));
    s.triPos = (const float*)r->triPos.ptr;
    s.triInfo = (const float*)r->triInfo.ptr;
    s.shadeInfo = (const float4*)r->shadeInfo.ptr;
    s.bvhNodes = (const okrt_bvh_node*)r->bvhNodes.ptr;
    s.textures = r->textures.object;
    s.environment = r->environment.object;
    s.octNodes = (const okrt_octree_node*)r->octNodes.ptr;
    s.spheres = (const okrt_sphere*)r->spheres.ptr;
    s.instances = (const okrt_mesh_instance*)r->instances.ptr;
    s.dirLights = (const okrt_directional_light*)r->dirLights.ptr;
    s.pointLights = (const okrt_point_light*)r->pointLights.ptr;
    s.spotLights = (const okrt_spot_light*)r->spotLights.ptr;
    s.numTris = r->triPos.count; s.numBvhNodes = r->bvhNodes.count; s.numOctNodes = r->octNodes.count;
    s.numSpheres = r->spheres.count; s.numInstances = r->instances.count;
    s.numDirLights = r->dirLights.count; s.numPointLights = r->pointLights.count; s.numSpotLights = r->spotLights.count;
    s.texSize = r->textures.width; s.texLayers = r->textures.layers;
    s.texInvSize = r->textures.width ? 1.0f / (float)r->textures.width : 0.f;
    s.envW = r->environment.width; s.envH = r->environment.height;
    s.fastNodes = (const float4*)r->fastNodes.ptr;
    s.fastTris = (const float4*)r->fastTris.ptr;
    s.fastTrisTex = r->fastTrisTex;
    s.fastInstances = (const float4*)r->fastInstances.ptr;
    s.octParent = (const uint32_t*)r->octParent.ptr;
    s.topRef = r->topRef;
    s.numFastNodes = r->fastNodes.count / 4u; s.numFastInstances = r->fastInstances.count / kInstQuads;
    s.wideNodes = (const uint4*)r->wideNodes.ptr; s.wideLeaves = (const uint4*)r->wideLeaves.ptr;
    s.wideRoot = r->wideRoot; s.numWideNodes = r->wideNodes.count / kWideNodeWords; s.maxSlack = r->maxSlack;
    return s;
}

void coveredExtent(const okrt_renderer* r, uint32_t& cw, uint32_t& ch)
{
    /* [Q9] Dispatch(W/16, H/9) x numthreads(16,9) — RayTracer.cpp:256 */
    cw = r->opt.coverAllPixels ? r->width : 16u * (r->width / 16u);
    ch = r->opt.coverAllPixels ? r->height : 9u * (r->height / 9u);
}

void freeLaneBuffers(BatchLane& l)
{
    void* ptrs[] = {l.wb.rayO, l.wb.rayD, l.wb.radiance, l.wb.throughput, l.wb.hitA, l.wb.hitB, l.wb.queue[0], l.wb.queue[1], l.wb.counts, l.wb.fallbackQueue};
    for (void* p : ptrs) if (p) cudaFree(p);
    std::memset(&l.wb, 0, sizeof(l.wb));
    l.wbPaths = 0;
}
void freeWavefront(okrt_renderer* r)
{
    for (BatchLane& l : r->lanes) freeLaneBuffers(l);
    if (r->traceStats) cudaFree(r->traceStats);
    r->traceStats = nullptr;
}

constexpr uint32_t kCountWords = kMaxQueueCounts + kMaxWavefrontBounces + 2;

/* streams and events of a lane on first use; path state grown to `paths` (freeing waits for the device: rare, sizes repeat) */
int ensureLane(okrt_renderer* r, BatchLane& l, size_t paths)
{
    if (!l.stream) {
        OKRT_CUDA(cudaStreamCreateWithFlags(&l.stream, cudaStreamNonBlocking));
        OKRT_CUDA(cudaStreamCreateWithFlags(&l.streamBeside, cudaStreamNonBlocking));
        OKRT_CUDA(cudaEventCreateWithFlags(&l.evFork, cudaEventDisableTiming));
        OKRT_CUDA(cudaEventCreateWithFlags(&l.evJoin, cudaEventDisableTiming));
        OKRT_CUDA(cudaEventCreateWithFlags(&l.evDone, cudaEventDisableTiming));
    }
    if (!r->traceStats) {
        OKRT_CUDA(cudaMalloc(&r->traceStats, 8 * (kMaxWavefrontBounces + 2) * sizeof(unsigned long long)));
        OKRT_CUDA(cudaMemset(r->traceStats, 0, 8 * (kMaxWavefrontBounces + 2) * sizeof(unsigned long long)));
    }
    if (l.wbPaths >= paths && l.wb.counts) { l.wb.traceStats = r->traceStats; return OKRT_OK; }
    freeLaneBuffers(l);
    OKRT_CUDA(cudaMalloc(&l.wb.rayO, paths * sizeof(float4)));
    OKRT_CUDA(cudaMalloc(&l.wb.rayD, paths * sizeof(float4)));
    OKRT_CUDA(cudaMalloc(&l.wb.radiance, paths * sizeof(float4)));
    OKRT_CUDA(cudaMalloc(&l.wb.throughput, paths * sizeof(float4)));
    OKRT_CUDA(cudaMalloc(&l.wb.hitA, paths * sizeof(float4)));
    OKRT_CUDA(cudaMalloc(&l.wb.hitB, paths * sizeof(uint32_t)));
    OKRT_CUDA(cudaMalloc(&l.wb.queue[0], paths * sizeof(uint32_t)));
    OKRT_CUDA(cudaMalloc(&l.wb.queue[1], paths * sizeof(uint32_t)));
    OKRT_CUDA(cudaMalloc(&l.wb.fallbackQueue, paths * sizeof(uint32_t)));
    OKRT_CUDA(cudaMalloc(&l.wb.counts, kCountWords * sizeof(uint32_t)));
    l.wb.fallbackCounts = l.wb.counts + kMaxQueueCounts;
    l.wb.traceStats = r->traceStats;
    l.wbPaths = paths;
    return OKRT_OK;
}

/* how many batches may be in flight: okrt_options::batchesInFlight (0 = default), OKRT_LANES overrides (tuning) */
int laneCount(const okrt_renderer* r)
{
    static const int env = [] { const char* e = std::getenv("OKRT_LANES"); return e ? std::atoi(e) : 0; }();
    int n = env > 0 ? env : (r->opt.batchesInFlight ? (int)r->opt.batchesInFlight : 3);
    return std::max(1, std::min(kMaxLanes, n));
}

/* Takes the next lane for a batch and orders its chain after the main stream where that is needed (see BatchLane).  `serial`:
 * the batch may not overlap with anything (instrumented runs share the AOV image, the megakernel adds to u1 itself). */
int beginBatch(okrt_renderer* r, size_t paths, bool serial, BatchLane*& out)
{
    const int count = serial ? 1 : laneCount(r);
    if (r->nextLane >= (uint32_t)count) r->nextLane = 0;
    BatchLane& l = r->lanes[serial ? 0 : r->nextLane];
    if (!serial) r->nextLane = (r->nextLane + 1u) % (uint32_t)count;
    int rc = ensureLane(r, l, paths);
    if (rc) return rc;
    if (r->mainTouched) {
        r->forkGen++;
        OKRT_CUDA(cudaEventRecord(r->evMainFork, r->stream));
        r->mainTouched = false;
    }
    if (serial) {                      /* after everything issued so far (the main stream is downstream of every batch) */
        OKRT_CUDA(cudaEventRecord(r->evMainFork, r->stream));
        OKRT_CUDA(cudaStreamWaitEvent(l.stream, r->evMainFork, 0));
        l.forkSeen = r->forkGen;
    } else if (l.forkSeen != r->forkGen) {
        OKRT_CUDA(cudaStreamWaitEvent(l.stream, r->evMainFork, 0));
        l.forkSeen = r->forkGen;
    }
    out = &l;
    return OKRT_OK;
}

/* the batch's writes to u1 / u0 come after everything on the main stream as of now */
int gateAccumulate(okrt_renderer* r, BatchLane& l)
{
    OKRT_CUDA(cudaEventRecord(r->evMainNow, r->stream));
    OKRT_CUDA(cudaStreamWaitEvent(l.stream, r->evMainNow, 0));
    return OKRT_OK;
}

/* the main stream waits for the batch: whatever the caller queues next sees its images */
int endBatch(okrt_renderer* r, BatchLane& l)
{
    OKRT_CUDA(cudaEventRecord(l.evDone, l.stream));
    OKRT_CUDA(cudaStreamWaitEvent(r->stream, l.evDone, 0));
    r->lastLane = (int)(&l - r->lanes);
    return OKRT_OK;
}

/* generate -> (trace, fallback, shade) x (maxBounces + 1) -> accumulate, for `fp.batchFrames` frames at once, on the next batch
 * lane; queue sizes stay on the device, nothing syncs with the host */
int launchWavefront(okrt_renderer* r, const FrameParams& fp, bool instrumented)
{
    const size_t slots = (size_t)fp.numTiles * kTilePixels;
    const size_t paths = slots * fp.batchFrames;
    BatchLane* lanePtr = nullptr;
    int rc = beginBatch(r, paths, instrumented, lanePtr);
    if (rc) return rc;
    BatchLane& lane = *lanePtr;
    const WavefrontBuffers& wb = lane.wb;
    cudaStream_t st = lane.stream;
    const DeviceScene scene = sceneView(r);
    const char* shadeWaves = std::getenv("OKRT_SHADE_WAVES");     /* tuning knob: resident-block multiples per SM */
    static const int traceBlocksEnv = [] { const char* e = std::getenv("OKRT_TRACE_BLOCKS"); return e ? std::max(1, std::atoi(e)) : 1 << 20; }();
    static const int shadeBlocksEnv = [] { const char* e = std::getenv("OKRT_SHADE_BLOCKS"); return e ? std::max(1, std::atoi(e)) : 0; }();
    const unsigned flatGrid = (unsigned)((paths + 127) / 128);
    /* persistent traversal grid: the whole machine (12 blocks per SM) when batches run one after the other; with several batches
     * in flight a few block slots per SM are left to the other lanes' kernels, which would otherwise wait for this kernel's last
     * block (measured: 10 per SM +0.1-0.4 % for full-image batches, 8 per SM +1.0-1.1 % for batches of a quarter or an eighth;
     * alone, a kernel on 8 or 10 per SM loses 2-6 %) */
    bool othersInFlight = false;                /* is the previous batch still on the device?  (a caller that synchronises after
                                                 * every call never has two batches in flight, whatever the lane count) */
    if (laneCount(r) > 1 && !instrumented && r->lastLane >= 0 && r->lanes[r->lastLane].evDone) {
        othersInFlight = cudaEventQuery(r->lanes[r->lastLane].evDone) == cudaErrorNotReady;
        cudaGetLastError();
    }
    const int traceBlocksShared = othersInFlight ? (paths <= ((size_t)1 << 23) ? 8 : 10) : r->blocksPerSmTrace;
    const unsigned traceGrid = std::min<unsigned>((unsigned)(r->numSms * std::min(std::min(r->blocksPerSmTrace, traceBlocksShared), traceBlocksEnv)), std::max(1u, flatGrid));
    const unsigned wideGrid = std::min<unsigned>((unsigned)(r->numSms * std::min(r->blocksPerSmWide, traceBlocksEnv)), std::max(1u, flatGrid));
    unsigned shadeGrid = std::min<unsigned>((unsigned)(r->numSms * OKRT_SHADE_MIN_BLOCKS * (shadeWaves ? std::max(1, std::atoi(shadeWaves)) : 16)), std::max(1u, flatGrid));
    if (shadeBlocksEnv) shadeGrid = std::min<unsigned>((unsigned)(r->numSms * shadeBlocksEnv), std::max(1u, flatGrid));
    shadeGrid = std::min<unsigned>(shadeGrid, std::max(1u, (flatGrid + kShadePathsPerThread - 1u) / kShadePathsPerThread));
    const unsigned fallbackGrid = std::min<unsigned>((unsigned)(r->numSms * 8), std::max(1u, flatGrid));   /* 4 rays (warps) per block */
    if (r->timingPending) { OKRT_CUDA(cudaEventRecord(r->evStart, st)); r->timingPending = false; }
    OKRT_CUDA(cudaMemsetAsync(wb.counts, 0, kCountWords * sizeof(uint32_t), st));
    generateKernel<<<flatGrid, 128, 0, st>>>(fp, wb);
    r->launches++;
    for (uint32_t b = 0; b <= fp.maxBounces; b++) {
        const uint32_t* qin = (b == 0u && fp.allCovered) ? nullptr : wb.queue[b & 1u];
        uint32_t* qout = wb.queue[(b + 1u) & 1u];
        if (fp.traversal == OKRT_TRAVERSAL_EXACT) {
            traceExactKernel<<<flatGrid, 128, 0, st>>>(scene, fp, wb, qin, b);
            if (instrumented) shadeKernel<true><<<shadeGrid, 128, 0, st>>>(scene, fp, wb, qin, qout, b, 0u);
            else shadeKernel<false><<<shadeGrid, 128, 0, st>>>(scene, fp, wb, qin, qout, b, 0u);
            r->launches += 2;
            continue;
        }
        if (r->traceEventsUsed + 2 > r->traceEvents.size()) {
            r->traceEvents.resize(r->traceEventsUsed + 2, nullptr);
            OKRT_CUDA(cudaEventCreate(&r->traceEvents[r->traceEventsUsed]));
            OKRT_CUDA(cudaEventCreate(&r->traceEvents[r->traceEventsUsed + 1]));
        }
        cudaEvent_t e0 = r->traceEvents[r->traceEventsUsed], e1 = r->traceEvents[r->traceEventsUsed + 1];
        r->traceEventsUsed += 2;
        OKRT_CUDA(cudaEventRecord(e0, st));
        if (fp.traversal == OKRT_TRAVERSAL_WIDE) traceWideKernel<<<wideGrid, kTraceBlockThreads, 0, st>>>(scene, fp, wb, qin, b);
        else traceFastKernel<<<traceGrid, kTraceBlockThreads, 0, st>>>(scene, fp, wb, qin, b);
        OKRT_CUDA(cudaEventRecord(e1, st));
        /* The reference-order kernel decides a few hundred rays per launch and lasts one bounded traversal (~100 us of
         * dependent loads): on the lane's second stream it runs BESIDE the main shade launch, which skips its paths (kHitPending);
         * a late shade launch over the fallback queue follows. */
        static const bool beside = [] { const char* e = std::getenv("OKRT_FALLBACK_BESIDE"); return e ? std::atoi(e) != 0 : true; }();
        if (beside) {
            OKRT_CUDA(cudaEventRecord(lane.evFork, st));
            OKRT_CUDA(cudaStreamWaitEvent(lane.streamBeside, lane.evFork, 0));
            traceFallbackKernel<<<fallbackGrid, 128, 0, lane.streamBeside>>>(scene, fp, wb, b);
            OKRT_CUDA(cudaEventRecord(lane.evJoin, lane.streamBeside));
        } else traceFallbackKernel<<<fallbackGrid, 128, 0, st>>>(scene, fp, wb, b);
        if (instrumented) shadeKernel<true><<<shadeGrid, 128, 0, st>>>(scene, fp, wb, qin, qout, b, 0u);
        else shadeKernel<false><<<shadeGrid, 128, 0, st>>>(scene, fp, wb, qin, qout, b, 0u);
        if (beside) OKRT_CUDA(cudaStreamWaitEvent(st, lane.evJoin, 0));
        if (instrumented) shadeKernel<true><<<fallbackGrid, 128, 0, st>>>(scene, fp, wb, qin, qout, b, 1u);
        else shadeKernel<false><<<fallbackGrid, 128, 0, st>>>(scene, fp, wb, qin, qout, b, 1u);
        r->launches += 4;                                  /* traversal, fallback, main shade, late shade */
    }
    if ((rc = gateAccumulate(r, lane))) return rc;
    accumulateKernel<<<(unsigned)((slots + 127) / 128), 128, 0, st>>>(fp, wb);
    r->launches++;
    OKRT_CUDA(cudaGetLastError());
    if ((rc = endBatch(r, lane))) return rc;
    if (std::getenv("OKRT_DEBUG_COUNTS")) {          /* diagnostic: rays per bounce of this batch */
        std::vector<uint32_t> h(kCountWords);
        cudaMemcpyAsync(h.data(), wb.counts, kCountWords * sizeof(uint32_t), cudaMemcpyDeviceToHost, st);
        cudaStreamSynchronize(st);
        std::fprintf(stderr, "okrt: batch of %u frames, rays per bounce:", fp.batchFrames);
        for (uint32_t b = 0; b <= fp.maxBounces; b++) std::fprintf(stderr, " %u(fb %u)", h[2 * b], h[kMaxQueueCounts + b]);
        std::fprintf(stderr, "\n");
#ifdef OKRT_TRACE_STATS
        std::vector<unsigned long long> ts(8 * (fp.maxBounces + 1));
        cudaMemcpy(ts.data(), r->traceStats, ts.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
        for (uint32_t b = 0; b <= fp.maxBounces; b++) {
            const unsigned long long* t = &ts[8 * b];
            const double rays = (double)std::max(1ull, t[0]);
            std::fprintf(stderr, "  bounce %u (since reset): rays %llu  node steps/ray %.2f  leaf entries/ray %.2f  triangle tests/ray %.2f  lanes per node iteration %.1f  per triangle iteration %.1f\n",
                         b, t[0], t[1] / rays, t[2] / rays, t[3] / rays, (double)t[6] / std::max(1ull, t[4]), (double)t[7] / std::max(1ull, t[5]));
        }
#endif
    }
    return OKRT_OK;
}

/* Renders up to `want` frames in one go and reports how many (`done`).  The wavefront pipeline traces several
 * progressive frames together (more rays per launch, fewer launches, better strong scaling); u1 still receives
 * them in frame order.  Instrumented runs, the megakernel and accumulation-off render one frame per call. */
constexpr uint32_t kMaxBatchFrames = 128;
constexpr size_t kMaxBatchPaths = (size_t)1 << 25;

/* how many of `want` frames the next batch takes, and the pixel slots of one frame (this handle's share of the 8-row bands
 * under tile sharding) */
uint32_t planBatch(const okrt_renderer* r, uint32_t want, size_t& slotsPerFrame)
{
    const bool instrumentedRun = r->opt.writeAov != 0u || r->rd.debugMode != 0u;
    const bool wavefront = r->opt.pipeline == OKRT_PIPELINE_WAVEFRONT && r->opt.maxBounces <= kMaxWavefrontBounces;
    uint32_t ownedRows = r->height;
    if (r->opt.shardMode == OKRT_SHARD_TILE && r->opt.shardCount > 1u) {
        const uint32_t band = r->opt.shardBandRows ? r->opt.shardBandRows : 8u;
        const uint32_t bands = (r->height + band - 1u) / band;
        ownedRows = (r->ownedBands.empty() ? bands / r->opt.shardCount + ((bands % r->opt.shardCount) > r->opt.shardRank ? 1u : 0u) : (uint32_t)r->ownedBands.size()) * band;
    }
    slotsPerFrame = (size_t)((r->width + kTileW - 1) / kTileW) * ((ownedRows + kTileH - 1) / kTileH) * kTilePixels;
    if (!(wavefront && !instrumentedRun && r->rd.accumulationEnabled == 1u)) return 1u;
    const size_t most = std::max<size_t>(1, std::min<size_t>(kMaxBatchFrames, kMaxBatchPaths / std::max<size_t>(slotsPerFrame, 1)));
    /* split `want` into equal batches rather than full ones and a small remainder */
    const size_t batches = (want + most - 1) / most;
    return (uint32_t)((want + batches - 1) / batches);
}

/* path state of the lanes this call will use, before its first kernel: growing a lane frees memory, which waits for the device */
int presizeLanes(okrt_renderer* r, uint32_t frames)
{
    if (!(r->opt.pipeline == OKRT_PIPELINE_WAVEFRONT && r->opt.maxBounces <= kMaxWavefrontBounces) || !frames) return OKRT_OK;
    if (r->opt.writeAov != 0u || r->rd.debugMode != 0u) return OKRT_OK;
    size_t slotsPerFrame = 0;
    const uint32_t done = planBatch(r, frames, slotsPerFrame);
    const uint32_t batches = (frames + done - 1u) / done;
    const int count = laneCount(r);
    const uint32_t first = r->nextLane >= (uint32_t)count ? 0u : r->nextLane;
    for (uint32_t i = 0; i < std::min<uint32_t>(batches, (uint32_t)count); i++) {
        const int rc = ensureLane(r, r->lanes[(first + i) % (uint32_t)count], slotsPerFrame * done);
        if (rc) return rc;
    }
    return OKRT_OK;
}

int launchFrames(okrt_renderer* r, uint32_t want, uint32_t& done)
{
    size_t slotsPerFrame = 0;
    done = planBatch(r, want, slotsPerFrame);
    /* updateBuffers: the counter advances before the constants are uploaded (RayTracer.cpp:489) */
    const uint32_t firstLocalFrame = r->rd.numAccumulationFrames + r->rd.accumulationEnabled;
    r->rd.numAccumulationFrames += r->rd.accumulationEnabled * done;
    r->rd.textureDims[0] = r->width; r->rd.textureDims[1] = r->height;

    FrameParams fp;
    std::memset(&fp, 0, sizeof(fp));
    fp.rd = r->rd;
    fp.maxBounces = r->opt.maxBounces;
    /* sample sharding: local frame k (1-based) draws the seed of global frame offset + (k-1)*stride + 1 */
    fp.seedFrame = okrt_shard_seed_frame(&r->opt, firstLocalFrame, r->rd.accumulationEnabled);
    fp.seedStride = r->rd.accumulationEnabled ? r->opt.frameStride : 0u;
    fp.batchFrames = done;
    /* scaled instances make the reference's local-vs-world culling order dependent; debug heat maps
     * display the reference's own visit counters: both need the reference visiting order */
    fp.traversal = (r->opt.traversal == OKRT_TRAVERSAL_EXACT || !r->allRigid || r->rd.debugMode != 0u) ? OKRT_TRAVERSAL_EXACT : r->opt.traversal;
    coveredExtent(r, fp.coveredW, fp.coveredH);
    fp.accumulation = r->accumulation;
    fp.result = r->result;
    /* AOV taps and the debug heat maps (which display the reference's visit counters) are instrumented builds */
    const bool instrumented = r->opt.writeAov != 0u || r->rd.debugMode != 0u;
    if (instrumented) { const int rcA = ensureAov(r); if (rcA) return rcA; }
    fp.aov = instrumented ? r->aov : nullptr;
    fp.counters = r->counters;

    uint32_t rows = r->height;
    fp.shardCount = 1; fp.shardRank = 0; fp.bandRows = kTileH;
    if (r->opt.shardMode == OKRT_SHARD_TILE && r->opt.shardCount > 1u) {
        fp.shardCount = r->opt.shardCount; fp.shardRank = r->opt.shardRank;
        fp.bandRows = r->opt.shardBandRows ? r->opt.shardBandRows : 8u;
        const uint32_t bands = (r->height + fp.bandRows - 1u) / fp.bandRows;
        const uint32_t mine = r->ownedBands.empty() ? bands / fp.shardCount + ((bands % fp.shardCount) > fp.shardRank ? 1u : 0u) : (uint32_t)r->ownedBands.size();
        rows = mine * fp.bandRows;
        if (!r->ownedBands.empty()) {
            if (r->ownedBands.back() >= bands) { setError("okrt_set_shard_bands: a band lies outside the image"); return OKRT_ERR_INVALID_ARGUMENT; }
            fp.ownedBands = (const uint32_t*)r->dOwnedBands.ptr;
        }
    }
    /* the reference clears the whole target to (0.2,0.4,0.6,1) every frame (RayTracer.cpp:227-228) and then
     * overwrites every covered pixel; uncovered pixels [Q9] are cleared once, when the images are created
     * or the coverage option changes, and never written afterwards */
    if (rows == 0) return OKRT_OK;
    fp.tilesX = (r->width + kTileW - 1) / kTileW;
    fp.numTiles = fp.tilesX * ((rows + kTileH - 1) / kTileH);
    /* every slot of the owned tiles is a covered pixel: full tiles in x, and in y the covered extent is the image and ends on a band */
    fp.allCovered = (fp.coveredW == fp.tilesX * (uint32_t)kTileW && fp.coveredH == r->height && r->height % fp.bandRows == 0u) ? 1u : 0u;
    fp.workCounter = reinterpret_cast<uint32_t*>(r->counters + 3);
    if (r->opt.pipeline == OKRT_PIPELINE_WAVEFRONT && fp.maxBounces <= kMaxWavefrontBounces) return launchWavefront(r, fp, instrumented);
    /* the megakernel adds to u1 itself: a serial batch on lane 0, after everything issued so far */
    BatchLane* lane = nullptr;
    int rc = beginBatch(r, 1, true, lane);
    if (rc) return rc;
    if (r->timingPending) { OKRT_CUDA(cudaEventRecord(r->evStart, lane->stream)); r->timingPending = false; }
    OKRT_CUDA(cudaMemsetAsync(fp.workCounter, 0, sizeof(unsigned long long), lane->stream));
    const int blocksPerSm = instrumented ? r->blocksPerSmInstrumented : r->blocksPerSmLean;
    const unsigned wantBlocks = (unsigned)((fp.numTiles * (uint32_t)kTilePixels + kBlockThreads - 1) / kBlockThreads);
    const unsigned grid = std::min<unsigned>((unsigned)(r->numSms * blocksPerSm), std::max(1u, wantBlocks));
    if (instrumented) pathTraceKernel<true><<<grid, kBlockThreads, 0, lane->stream>>>(sceneView(r), fp);
    else pathTraceKernel<false><<<grid, kBlockThreads, 0, lane->stream>>>(sceneView(r), fp);
    r->launches++;
    OKRT_CUDA(cudaGetLastError());
    return endBatch(r, *lane);
}

/* the steady loop's calls (render, set_render_data, read_result_async, wait_read): queue nothing a batch's chain reads */
int checkHandleQuiet(const okrt_renderer* r, const char* fn)
{
    if (!r) { setError(std::string(fn) + ": null renderer"); return OKRT_ERR_INVALID_ARGUMENT; }
    cudaError_t e = cudaSetDevice(r->device);
    if (e != cudaSuccess) { setError(std::string(fn) + ": cudaSetDevice: " + cudaGetErrorString(e)); return OKRT_ERR_CUDA; }
    return OKRT_OK;
}
/* every other entry point may queue work on the main stream that the next batch has to see from its first kernel on */
int checkHandle(const okrt_renderer* r, const char* fn)
{
    const int rc = checkHandleQuiet(r, fn);
    if (!rc) const_cast<okrt_renderer*>(r)->mainTouched = true;
    return rc;
}

/* rows of the result this handle owns, as contiguous byte runs of the image (a strided family + extra runs) */
struct OwnedRuns {
    size_t first = 0, pitch = 0, runBytes = 0; uint32_t regular = 0;      /* `regular` runs of runBytes at first + k * pitch */
    std::vector<std::pair<size_t, size_t>> extra;                         /* (offset, bytes) */
};
OwnedRuns ownedRuns(const okrt_renderer* r)
{
    OwnedRuns o;
    const size_t rowBytes = (size_t)r->width * 4;
    if (r->opt.shardMode != OKRT_SHARD_TILE || r->opt.shardCount <= 1) { o.pitch = o.runBytes = rowBytes * r->height; o.regular = 1; return o; }
    const uint32_t band = r->opt.shardBandRows ? r->opt.shardBandRows : 8u, count = r->opt.shardCount, rank = r->opt.shardRank;
    const uint32_t bands = (r->height + band - 1u) / band;
    if (!r->ownedBands.empty()) {                                         /* explicit table: neighbouring bands merge into one run */
        for (size_t i = 0; i < r->ownedBands.size();) {
            size_t j = i + 1;
            while (j < r->ownedBands.size() && r->ownedBands[j] == r->ownedBands[j - 1] + 1u) j++;
            const size_t rowStart = (size_t)r->ownedBands[i] * band, rowEnd = std::min<size_t>((size_t)(r->ownedBands[j - 1] + 1u) * band, r->height);
            if (rowEnd > rowStart) o.extra.emplace_back(rowStart * rowBytes, (rowEnd - rowStart) * rowBytes);
            i = j;
        }
        return o;
    }
    const uint32_t mine = bands / count + ((bands % count) > rank ? 1u : 0u);
    o.runBytes = (size_t)band * rowBytes; o.first = (size_t)rank * band * rowBytes; o.pitch = (size_t)count * band * rowBytes;
    if (!mine) return o;
    const uint32_t lastBand = rank + (mine - 1u) * count;
    const uint32_t lastRows = std::min(band, r->height - lastBand * band);
    o.regular = lastRows == band ? mine : mine - 1u;
    if (o.regular != mine) o.extra.emplace_back((size_t)lastBand * band * rowBytes, (size_t)lastRows * rowBytes);
    return o;
}
/* copies the owned runs from `src` to `dst` (same layout: full-size images) */
cudaError_t copyOwned(const OwnedRuns& o, uint8_t* dst, const uint8_t* src, cudaMemcpyKind kind, cudaStream_t stream, uint64_t* moved)
{
    cudaError_t e = cudaSuccess;
    if (o.regular) {
        e = cudaMemcpy2DAsync(dst + o.first, o.pitch, src + o.first, o.pitch, o.runBytes, o.regular, kind, stream);
        if (moved) *moved += (uint64_t)o.regular * o.runBytes;
    }
    for (const auto& run : o.extra) {
        if (e != cudaSuccess) break;
        e = cudaMemcpyAsync(dst + run.first, src + run.first, run.second, kind, stream);
        if (moved) *moved += run.second;
    }
    return e;
}


} // namespace

extern "C" {

int okrt_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int okrt_create(int device, uint32_t width, uint32_t height, okrt_renderer** out)
{
    if (!out || !width || !height) { setError("okrt_create: bad argument"); return OKRT_ERR_INVALID_ARGUMENT; }
    *out = nullptr;
    const int n = okrt_device_count();
    if (n <= 0) { setError("okrt_create: no CUDA device visible; this library has no CPU path"); return OKRT_ERR_NO_DEVICE; }
    if (device < 0 || device >= n) { setError("okrt_create: device index out of range"); return OKRT_ERR_INVALID_ARGUMENT; }
    OKRT_CUDA(cudaSetDevice(device));
    okrt_renderer* r = new okrt_renderer();
    r->device = device; r->width = width; r->height = height;
    okrt_render_data_default(&r->rd);
    okrt_options_default(&r->opt);
    r->rd.textureDims[0] = width; r->rd.textureDims[1] = height;
    cudaError_t e;
    if ((e = cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&r->evMainFork, cudaEventDisableTiming)) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&r->evMainNow, cudaEventDisableTiming)) != cudaSuccess ||
        (e = cudaEventCreate(&r->evStart)) != cudaSuccess || (e = cudaEventCreate(&r->evStop)) != cudaSuccess ||
        (e = cudaMalloc(&r->counters, 4 * sizeof(unsigned long long))) != cudaSuccess ||
        (e = cudaMemsetAsync(r->counters, 0, 4 * sizeof(unsigned long long), r->stream)) != cudaSuccess) {
        setError(std::string("okrt_create: ") + cudaGetErrorString(e));
        okrt_destroy(r);
        return OKRT_ERR_CUDA;
    }
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess ||
        (e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&r->blocksPerSmLean, pathTraceKernel<false>, kBlockThreads, 0)) != cudaSuccess ||
        (e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&r->blocksPerSmInstrumented, pathTraceKernel<true>, kBlockThreads, 0)) != cudaSuccess ||
        (e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&r->blocksPerSmTrace, traceFastKernel, kTraceBlockThreads, 0)) != cudaSuccess ||
        (e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&r->blocksPerSmWide, traceWideKernel, kTraceBlockThreads, 0)) != cudaSuccess) {
        setError(std::string("okrt_create: ") + cudaGetErrorString(e));
        okrt_destroy(r);
        return OKRT_ERR_CUDA;
    }
    if (const char* carve = std::getenv("OKRT_L1_CARVEOUT")) {     /* tuning knob: shared-memory share of the L1 (percent) */
        const int pct = std::atoi(carve);
        cudaFuncSetAttribute(traceFastKernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
        cudaFuncSetAttribute(traceWideKernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
        cudaFuncSetAttribute(shadeKernel<false>, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
        cudaFuncSetAttribute(traceFallbackKernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
        cudaFuncSetAttribute(generateKernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
        cudaFuncSetAttribute(accumulateKernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
    }
    r->numSms = prop.multiProcessorCount;
    r->blocksPerSmLean = std::max(1, r->blocksPerSmLean);
    r->blocksPerSmInstrumented = std::max(1, r->blocksPerSmInstrumented);
    r->blocksPerSmTrace = std::max(1, r->blocksPerSmTrace);
    r->blocksPerSmWide = std::max(1, r->blocksPerSmWide);
    const int rc = allocateImages(r);
    if (rc) { okrt_destroy(r); return rc; }
    *out = r;
    return OKRT_OK;
}

void okrt_destroy(okrt_renderer* r)
{
    if (!r) return;
    cudaSetDevice(r->device);
    for (BatchLane& l : r->lanes) {
        if (l.stream) cudaStreamSynchronize(l.stream);
        if (l.streamBeside) cudaStreamSynchronize(l.streamBeside);
    }
    if (r->stream) cudaStreamSynchronize(r->stream);
    r->triPos.release(); r->triInfo.release(); r->shadeInfo.release(); r->bvhNodes.release(); r->octNodes.release(); r->spheres.release();
    r->instances.release(); r->dirLights.release(); r->pointLights.release(); r->spotLights.release();
    if (r->fastTrisTex) { cudaDestroyTextureObject(r->fastTrisTex); r->fastTrisTex = 0; }
    r->fastNodes.release(); r->fastTris.release(); r->fastInstances.release(); r->octParent.release(); r->wideNodes.release(); r->wideLeaves.release();
    r->dOwnedBands.release();
    r->ovSegments.release(); r->ovKeys.release(); r->ovColours.release(); r->ovPixelKey.release();
    r->textures.release(); r->environment.release();
    if (r->accumulation) cudaFree(r->accumulation);
    if (r->result) cudaFree(r->result);
    if (r->aov) cudaFree(r->aov);
    if (r->counters) cudaFree(r->counters);
    freeWavefront(r);
    for (cudaEvent_t e : r->traceEvents) if (e) cudaEventDestroy(e);
    if (r->evStart) cudaEventDestroy(r->evStart);
    if (r->evStop) cudaEventDestroy(r->evStop);
    if (r->evMainFork) cudaEventDestroy(r->evMainFork);
    if (r->evMainNow) cudaEventDestroy(r->evMainNow);
    for (BatchLane& l : r->lanes) {
        if (l.evFork) cudaEventDestroy(l.evFork);
        if (l.evJoin) cudaEventDestroy(l.evJoin);
        if (l.evDone) cudaEventDestroy(l.evDone);
        if (l.streamBeside) cudaStreamDestroy(l.streamBeside);
        if (l.stream) cudaStreamDestroy(l.stream);
    }
    for (int k = 0; k < OKRT_READ_SLOTS; k++) {
        if (r->snapshot[k]) cudaFree(r->snapshot[k]);
        if (r->evSnapshot[k]) cudaEventDestroy(r->evSnapshot[k]);
        if (r->evCopied[k]) cudaEventDestroy(r->evCopied[k]);
    }
    if (r->streamCopy) { cudaStreamSynchronize(r->streamCopy); cudaStreamDestroy(r->streamCopy); }
    if (r->stream) cudaStreamDestroy(r->stream);
    delete r;
}

int okrt_resize(okrt_renderer* r, uint32_t width, uint32_t height)
{
    int rc = checkHandle(r, "okrt_resize"); if (rc) return rc;
    if (!width || !height) { setError("okrt_resize: zero extent"); return OKRT_ERR_INVALID_ARGUMENT; }
    OKRT_CUDA(cudaStreamSynchronize(r->stream));
    if (r->streamCopy) OKRT_CUDA(cudaStreamSynchronize(r->streamCopy));
    for (int k = 0; k < OKRT_READ_SLOTS; k++) { if (r->snapshot[k]) cudaFree(r->snapshot[k]); r->snapshot[k] = nullptr; r->copyPending[k] = false; }
    r->width = width; r->height = height;
    r->rd.textureDims[0] = width; r->rd.textureDims[1] = height;
    r->rd.numAccumulationFrames = 0u;                        /* onResize -> resetAccumulation, RayTracer.cpp:494-501 */
    return allocateImages(r);
}

/* DeviceScene::shadeInfo from t1 (TriangleInfo = 3 x VertexInfo of 11 floats: normal 0-2, uv 3-4), bit copies */
__global__ void packShadeInfoKernel(const float* __restrict__ triInfo, float4* __restrict__ out, uint32_t n)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const float* ti = triInfo + (size_t)t * 33;
    out[4 * (size_t)t + 0] = make_float4(ti[0], ti[1], ti[2], ti[3]);
    out[4 * (size_t)t + 1] = make_float4(ti[11], ti[12], ti[13], ti[4]);
    out[4 * (size_t)t + 2] = make_float4(ti[22], ti[23], ti[24], ti[14]);
    out[4 * (size_t)t + 3] = make_float4(ti[15], ti[25], ti[26], 0.f);
}

int okrt_set_triangles(okrt_renderer* r, const okrt_triangle* pos, const okrt_triangle_info* info, uint32_t n)
{
    int rc = checkHandle(r, "okrt_set_triangles"); if (rc) return rc;
    if (n && (!pos || !info)) { setError("okrt_set_triangles: null array"); return OKRT_ERR_INVALID_ARGUMENT; }
    if ((rc = r->triPos.upload(pos, n, sizeof(okrt_triangle), r->stream))) return rc;
    if ((rc = r->triInfo.upload(info, n, sizeof(okrt_triangle_info), r->stream))) return rc;
    {
        const size_t need = (size_t)n * 4 * sizeof(float4);
        if (need != r->shadeInfo.bytes) {
            r->shadeInfo.release();
            if (need) OKRT_CUDA(cudaMalloc(&r->shadeInfo.ptr, need));
            r->shadeInfo.bytes = need;
        }
        r->shadeInfo.count = n;
        if (n) {
            packShadeInfoKernel<<<(n + 255u) / 256u, 256, 0, r->stream>>>((const float*)r->triInfo.ptr, (float4*)r->shadeInfo.ptr, n);
            OKRT_CUDA(cudaGetLastError());
        }
    }
    r->hTris = std::make_shared<std::vector<okrt_triangle>>(pos, pos + n);
    r->layoutDirty = true;
    return OKRT_OK;
}

int okrt_set_bvh_nodes(okrt_renderer* r, const okrt_bvh_node* nodes, uint32_t n)
{
    int rc = checkHandle(r, "okrt_set_bvh_nodes"); if (rc) return rc;
    if (n && !nodes) { setError("okrt_set_bvh_nodes: null array"); return OKRT_ERR_INVALID_ARGUMENT; }
    if ((rc = r->bvhNodes.upload(nodes, n, sizeof(okrt_bvh_node), r->stream))) return rc;
    r->hNodes = std::make_shared<std::vector<okrt_bvh_node>>(nodes, nodes + n);
    r->layoutDirty = true;
    return OKRT_OK;
}

int okrt_set_textures(okrt_renderer* r, const uint8_t* rgba8, uint32_t size, uint32_t layers)
{
    int rc = checkHandle(r, "okrt_set_textures"); if (rc) return rc;
    if (size && layers && !rgba8) { setError("okrt_set_textures: null array"); return OKRT_ERR_INVALID_ARGUMENT; }
    return r->textures.upload(rgba8, size, size, layers, r->stream);
}

int okrt_set_environment_map(okrt_renderer* r, const uint8_t* faces6, uint32_t faceW, uint32_t faceH)
{
    int rc = checkHandle(r, "okrt_set_environment_map"); if (rc) return rc;
    if (faceW && faceH && !faces6) { setError("okrt_set_environment_map: null array"); return OKRT_ERR_INVALID_ARGUMENT; }
    return r->environment.upload(faces6, faceW, faceH, (faceW && faceH) ? 6u : 0u, r->stream);
}

int okrt_set_octree(okrt_renderer* r, const okrt_octree_node* nodes, uint32_t n)
{
    int rc = checkHandle(r, "okrt_set_octree"); if (rc) return rc;
    if (n && !nodes) { setError("okrt_set_octree: null array"); return OKRT_ERR_INVALID_ARGUMENT; }
    if ((rc = r->octNodes.upload(nodes, n, sizeof(okrt_octree_node), r->stream))) return rc;
    r->hOct = std::make_shared<std::vector<okrt_octree_node>>(nodes, nodes + n);
    r->layoutDirty = true;
    return OKRT_OK;
}

int okrt_set_spheres(okrt_renderer* r, const okrt_sphere* spheres, uint32_t n)
{
    int rc = checkHandle(r, "okrt_set_spheres"); if (rc) return rc;
    if (n && !spheres) { setError("okrt_set_spheres: null array"); return OKRT_ERR_INVALID_ARGUMENT; }
    return r->spheres.upload(spheres, n, sizeof(okrt_sphere), r->stream);
}

int okrt_set_mesh_instances(okrt_renderer* r, const okrt_mesh_instance* inst, uint32_t n)
{
    int rc = checkHandle(r, "okrt_set_mesh_instances"); if (rc) return rc;
    if (n && !inst) { setError("okrt_set_mesh_instances: null array"); return OKRT_ERR_INVALID_ARGUMENT; }
    if ((rc = r->instances.upload(inst, n, sizeof(okrt_mesh_instance), r->stream))) return rc;
    r->hInstances = std::make_shared<std::vector<okrt_mesh_instance>>(inst, inst + n);
    r->layoutDirty = true;
    return OKRT_OK;
}

int okrt_set_directional_lights(okrt_renderer* r, const okrt_directional_light* l, uint32_t n)
{
    int rc = checkHandle(r, "okrt_set_directional_lights"); if (rc) return rc;
    if (n && !l) { setError("okrt_set_directional_lights: null array"); return OKRT_ERR_INVALID_ARGUMENT; }
    return r->dirLights.upload(l, n, sizeof(okrt_directional_light), r->stream);
}

int okrt_set_point_lights(okrt_renderer* r, const okrt_point_light* l, uint32_t n)
{
    int rc = checkHandle(r, "okrt_set_point_lights"); if (rc) return rc;
    if (n && !l) { setError("okrt_set_point_lights: null array"); return OKRT_ERR_INVALID_ARGUMENT; }
    return r->pointLights.upload(l, n, sizeof(okrt_point_light), r->stream);
}

int okrt_set_spot_lights(okrt_renderer* r, const okrt_spot_light* l, uint32_t n)
{
    int rc = checkHandle(r, "okrt_set_spot_lights"); if (rc) return rc;
    if (n && !l) { setError("okrt_set_spot_lights: null array"); return OKRT_ERR_INVALID_ARGUMENT; }
    return r->spotLights.upload(l, n, sizeof(okrt_spot_light), r->stream);
}

int okrt_set_render_data(okrt_renderer* r, const okrt_render_data* rd)
{
    if (!r || !rd) { setError("okrt_set_render_data: null argument"); return OKRT_ERR_INVALID_ARGUMENT; }
    r->rd = *rd;
    r->rd.textureDims[0] = r->width; r->rd.textureDims[1] = r->height;
    return OKRT_OK;
}

int okrt_get_render_data(const okrt_renderer* r, okrt_render_data* out)
{
    if (!r || !out) { setError("okrt_get_render_data: null argument"); return OKRT_ERR_INVALID_ARGUMENT; }
    *out = r->rd;
    return OKRT_OK;
}

int okrt_set_options(okrt_renderer* r, const okrt_options* o)
{
    if (!r || !o) { setError("okrt_set_options: null argument"); return OKRT_ERR_INVALID_ARGUMENT; }
    if (o->traversal > OKRT_TRAVERSAL_WIDE || o->shardMode > OKRT_SHARD_TILE || o->pipeline > OKRT_PIPELINE_MEGAKERNEL) { setError("okrt_set_options: unknown mode"); return OKRT_ERR_INVALID_ARGUMENT; }
    if (o->traversal == OKRT_TRAVERSAL_WIDE && o->pipeline != OKRT_PIPELINE_WAVEFRONT) { setError("okrt_set_options: OKRT_TRAVERSAL_WIDE needs the wavefront pipeline"); return OKRT_ERR_INVALID_ARGUMENT; }
    if (o->traversal == OKRT_TRAVERSAL_WIDE && !r->layoutHasWide) r->layoutDirty = true;
    if (o->shardMode == OKRT_SHARD_TILE && (o->shardCount == 0u || o->shardRank >= o->shardCount)) { setError("okrt_set_options: shardRank/shardCount"); return OKRT_ERR_INVALID_ARGUMENT; }
    if (o->shardBandRows % kTileH) { setError("okrt_set_options: shardBandRows must be a multiple of 8"); return OKRT_ERR_INVALID_ARGUMENT; }
    const bool coverageChanged = (r->opt.coverAllPixels != 0u) != (o->coverAllPixels != 0u);
    r->opt = *o;
    if (coverageChanged) {
        const size_t n = (size_t)r->width * r->height;
        if (cudaSetDevice(r->device) == cudaSuccess) {
            clearResultKernel<<<(unsigned)((n + 255) / 256), 256, 0, r->stream>>>(r->result, n);
            r->launches++;
        }
    }
    if (r->opt.frameStride == 0u) r->opt.frameStride = 1u;
    if (r->opt.shardCount == 0u) r->opt.shardCount = 1u;
    return OKRT_OK;
}

int okrt_render_frames(okrt_renderer* r, uint32_t frames)
{
    int rc = checkHandleQuiet(r, "okrt_render"); if (rc) return rc;
    if (r->layoutDirty) { r->mainTouched = true; if ((rc = rebuildLayout(r))) return rc; }
    if (r->refine && (rc = adoptRefined(r, false))) return rc;
    if (r->opt.writeAov && !r->aov) { r->mainTouched = true; if ((rc = ensureAov(r))) return rc; }
    if ((rc = presizeLanes(r, frames))) return rc;
    r->traceEventsUsed = 0;
    r->timingPending = true;                                  /* evStart: where the first batch's chain begins */
    for (uint32_t f = 0; f < frames;) {
        uint32_t done = 1u;
        if ((rc = launchFrames(r, frames - f, done))) return rc;
        f += done;
    }
    if (r->timingPending) { OKRT_CUDA(cudaEventRecord(r->evStart, r->stream)); r->timingPending = false; }   /* nothing was launched */
    OKRT_CUDA(cudaEventRecord(r->evStop, r->stream));         /* the main stream waits for every batch (endBatch) */
    return OKRT_OK;
}

int okrt_render(okrt_renderer* r) { return okrt_render_frames(r, 1u); }

int okrt_sync(okrt_renderer* r)
{
    int rc = checkHandle(r, "okrt_sync"); if (rc) return rc;
    OKRT_CUDA(cudaStreamSynchronize(r->stream));
    return OKRT_OK;
}

int okrt_toggle_accumulation(okrt_renderer* r, int enable)
{
    int rc = checkHandle(r, "okrt_toggle_accumulation"); if (rc) return rc;
    r->rd.accumulationEnabled = enable ? 1u : 0u;
    return okrt_reset_accumulation(r);
}

int okrt_reset_accumulation(okrt_renderer* r)
{
    int rc = checkHandle(r, "okrt_reset_accumulation"); if (rc) return rc;
    r->rd.numAccumulationFrames = 0u;
    OKRT_CUDA(cudaMemsetAsync(r->accumulation, 0, (size_t)r->width * r->height * sizeof(float4), r->stream));
    return OKRT_OK;
}

uint32_t okrt_get_num_accumulation_frames(const okrt_renderer* r) { return r ? r->rd.numAccumulationFrames : 0u; }

int okrt_read_result(okrt_renderer* r, uint8_t* outRgba8)
{
    int rc = checkHandle(r, "okrt_read_result"); if (rc) return rc;
    if (!outRgba8) { setError("okrt_read_result: null output"); return OKRT_ERR_INVALID_ARGUMENT; }
    OKRT_CUDA(cudaMemcpyAsync(outRgba8, r->result, (size_t)r->width * r->height * 4, cudaMemcpyDeviceToHost, r->stream));
    OKRT_CUDA(cudaStreamSynchronize(r->stream));
    return OKRT_OK;
}

/* Tile-sharded ranks own interleaved row bands; only those rows of the result carry this rank's pixels.  One strided
 * DMA moves them into place in a full-size host image (the caller's compositor sees every rank's rows side by side). */
int okrt_read_result_owned(okrt_renderer* r, uint8_t* outRgba8, uint64_t* outBytes)
{
    int rc = checkHandle(r, "okrt_read_result_owned"); if (rc) return rc;
    if (!outRgba8) { setError("okrt_read_result_owned: null output"); return OKRT_ERR_INVALID_ARGUMENT; }
    if (r->opt.shardMode != OKRT_SHARD_TILE || r->opt.shardCount <= 1) {
        if (outBytes) *outBytes = (uint64_t)r->width * r->height * 4;
        return okrt_read_result(r, outRgba8);
    }
    uint64_t moved = 0;
    OKRT_CUDA(copyOwned(ownedRuns(r), outRgba8, (const uint8_t*)r->result, cudaMemcpyDeviceToHost, r->stream, &moved));
    OKRT_CUDA(cudaStreamSynchronize(r->stream));
    if (outBytes) *outBytes = moved;
    return OKRT_OK;
}

/* Tile sharding with an explicit band table: `bands` = the global band indices (ascending, no duplicates) this handle traces,
 * instead of every shardCount-th band.  Lets the host balance ranks by measured band cost (okrt_shard_balance); n = 0 restores
 * the interleave.  Takes effect with the next okrt_render* call; shardBandRows must be set first. */
int okrt_set_shard_bands(okrt_renderer* r, const uint32_t* bands, uint32_t n)
{
    int rc = checkHandle(r, "okrt_set_shard_bands"); if (rc) return rc;
    if (n && !bands) { setError("okrt_set_shard_bands: null array"); return OKRT_ERR_INVALID_ARGUMENT; }
    for (uint32_t i = 1; i < n; i++) if (bands[i] <= bands[i - 1]) { setError("okrt_set_shard_bands: bands must ascend"); return OKRT_ERR_INVALID_ARGUMENT; }
    OKRT_CUDA(cudaStreamSynchronize(r->stream));
    r->ownedBands.assign(bands, bands + n);
    return r->dOwnedBands.upload(bands, n, sizeof(uint32_t), r->stream);
}

int okrt_read_result_async(okrt_renderer* r, uint8_t* outRgba8, uint64_t* outBytes, uint32_t slot)
{
    int rc = checkHandleQuiet(r, "okrt_read_result_async"); if (rc) return rc;
    if (!outRgba8 || slot >= OKRT_READ_SLOTS) { setError("okrt_read_result_async: bad argument"); return OKRT_ERR_INVALID_ARGUMENT; }
    const size_t imageBytes = (size_t)r->width * r->height * 4;
    if (!r->streamCopy) {
        OKRT_CUDA(cudaStreamCreateWithFlags(&r->streamCopy, cudaStreamNonBlocking));
        for (int k = 0; k < OKRT_READ_SLOTS; k++) {
            OKRT_CUDA(cudaEventCreateWithFlags(&r->evSnapshot[k], cudaEventDisableTiming));
            OKRT_CUDA(cudaEventCreateWithFlags(&r->evCopied[k], cudaEventDisableTiming));
        }
    }
    if (!r->snapshot[slot]) OKRT_CUDA(cudaMalloc(&r->snapshot[slot], imageBytes));
    /* the snapshot of this slot may still be on its way to the host from two calls ago */
    if (r->copyPending[slot]) OKRT_CUDA(cudaStreamWaitEvent(r->stream, r->evCopied[slot], 0));
    const OwnedRuns runs = ownedRuns(r);
    uint64_t moved = 0;
    const uint8_t* src = (const uint8_t*)r->result;
    uint8_t* snap = (uint8_t*)r->snapshot[slot];
    OKRT_CUDA(copyOwned(runs, snap, src, cudaMemcpyDeviceToDevice, r->stream, nullptr));
    OKRT_CUDA(cudaEventRecord(r->evSnapshot[slot], r->stream));
    OKRT_CUDA(cudaStreamWaitEvent(r->streamCopy, r->evSnapshot[slot], 0));
    OKRT_CUDA(copyOwned(runs, outRgba8, snap, cudaMemcpyDeviceToHost, r->streamCopy, &moved));
    OKRT_CUDA(cudaEventRecord(r->evCopied[slot], r->streamCopy));
    r->copyPending[slot] = true;
    if (outBytes) *outBytes = moved;
    return OKRT_OK;
}

int okrt_wait_read(okrt_renderer* r, uint32_t slot)
{
    int rc = checkHandleQuiet(r, "okrt_wait_read"); if (rc) return rc;
    if (slot >= OKRT_READ_SLOTS) { setError("okrt_wait_read: bad slot"); return OKRT_ERR_INVALID_ARGUMENT; }
    if (!r->copyPending[slot]) return OKRT_OK;
    OKRT_CUDA(cudaEventSynchronize(r->evCopied[slot]));
    r->copyPending[slot] = false;
    return OKRT_OK;
}

int okrt_reduce_accumulation(okrt_renderer* r, okrt_allreduce_fn allreduce, void* user, uint32_t totalFrames)
{
    int rc = checkHandle(r, "okrt_reduce_accumulation"); if (rc) return rc;
    if (!allreduce || !totalFrames) { setError("okrt_reduce_accumulation: bad argument"); return OKRT_ERR_INVALID_ARGUMENT; }
    if (allreduce((void*)r->accumulation, (size_t)r->width * r->height * 4, (void*)r->stream, user) != 0) {
        setError("okrt_reduce_accumulation: the all-reduce callback failed");
        return OKRT_ERR_STATE;
    }
    r->rd.numAccumulationFrames = totalFrames;
    return okrt_resolve(r, totalFrames);
}

/* ------------------------------------------------------------ unit rates (bench.py's roofline denominators) */
__global__ void __launch_bounds__(128) sectorRateKernel(const uint4* __restrict__ data, uint32_t mask, uint32_t blockWindow, uint32_t iters, uint32_t* sink)
{
    /* blockWindow != 0: each block stays inside its own window of that many sectors (L1-resident) */
    if (blockWindow) { data += (size_t)(blockIdx.x % 64u) * blockWindow * 2u; mask = blockWindow - 1u; }
    /* every lane walks its own pseudo-random chain of 32-byte sectors: the traversal kernel's access shape (two 256-bit loads per
     * node, no two lanes in one sector) with enough independent loads in flight per lane to measure throughput, not latency */
    uint32_t x = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    uint32_t acc = 0u;
    for (uint32_t i = 0; i < iters; i++) {
        uint4 a[4], b[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            x = x * 1664525u + 1013904223u;
            const uint4* p = data + (size_t)((x >> 8) & mask) * 2u;
            asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                         : "=r"(a[k].x), "=r"(a[k].y), "=r"(a[k].z), "=r"(a[k].w), "=r"(b[k].x), "=r"(b[k].y), "=r"(b[k].z), "=r"(b[k].w) : "l"(p));
        }
#pragma unroll
        for (int k = 0; k < 4; k++) acc += a[k].x ^ b[k].w;
    }
    if (acc == 0x12345678u) *sink = acc;
}

__global__ void __launch_bounds__(128) issueRateKernel(uint32_t iters, float seed, float* sink, unsigned long long* cycles)
{
    float a0 = seed + threadIdx.x, a1 = a0 + 1.f, a2 = a0 + 2.f, a3 = a0 + 3.f, a4 = a0 + 4.f, a5 = a0 + 5.f, a6 = a0 + 6.f, a7 = a0 + 7.f;
    const float m = 1.0000001f, c = 0.5f;
    unsigned long long g0, g1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g0));
    const long long t0 = clock64();
    for (uint32_t i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 8; k++) {           /* 64 independent FFMA per trip: 8 chains x 8 */
            a0 = __fmaf_rn(a0, m, c); a1 = __fmaf_rn(a1, m, c); a2 = __fmaf_rn(a2, m, c); a3 = __fmaf_rn(a3, m, c);
            a4 = __fmaf_rn(a4, m, c); a5 = __fmaf_rn(a5, m, c); a6 = __fmaf_rn(a6, m, c); a7 = __fmaf_rn(a7, m, c);
        }
    }
    const long long t1 = clock64();
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g1));
    const float sum = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
    if (sum == 123.456f) *sink = sum;
    /* SM clock under this load: block 0's own cycles over its own nanoseconds (the grid runs several waves of blocks) */
    if (blockIdx.x == 0 && threadIdx.x == 0) { cycles[0] = (unsigned long long)(t1 - t0); cycles[1] = g1 - g0; }
}

int okrt_measure_unit_peaks(int device, okrt_unit_peaks* out)
{
    if (!out) { setError("okrt_measure_unit_peaks: null output"); return OKRT_ERR_INVALID_ARGUMENT; }
    std::memset(out, 0, sizeof(*out));
    if (okrt_device_count() <= device || device < 0) { setError("okrt_measure_unit_peaks: no such CUDA device"); return OKRT_ERR_NO_DEVICE; }
    OKRT_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    OKRT_CUDA(cudaGetDeviceProperties(&prop, device));
    const uint32_t sectors = 1u << 19;                          /* 2^19 x 32 B = 16 MB: inside the 126 MB L2, far outside the L1 */
    uint4* data = nullptr; uint32_t* sink = nullptr; float* fsink = nullptr; unsigned long long* cycles = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    cudaError_t e = cudaMalloc(&data, (size_t)sectors * 32);
    if (e == cudaSuccess) e = cudaMemset(data, 1, (size_t)sectors * 32);
    if (e == cudaSuccess) e = cudaMalloc(&sink, 4);
    if (e == cudaSuccess) e = cudaMalloc(&fsink, 4);
    if (e == cudaSuccess) e = cudaMalloc(&cycles, 16);
    if (e == cudaSuccess) e = cudaEventCreate(&e0);
    if (e == cudaSuccess) e = cudaEventCreate(&e1);
    const unsigned grid = (unsigned)prop.multiProcessorCount * 8u;
    float msSector = 0.f, msSectorL1 = 0.f, msIssue = 0.f;
    const uint32_t sIters = 256u, iIters = 4096u;
    unsigned long long hCycles[2] = {0, 0};
    if (e == cudaSuccess) {
        for (int rep = 0; rep < 3 && e == cudaSuccess; rep++) {     /* the last repetition is the warm one */
            cudaEventRecord(e0);
            sectorRateKernel<<<grid, 128>>>(data, sectors - 1u, 0u, sIters, sink);
            cudaEventRecord(e1);
            e = cudaEventSynchronize(e1);
            if (e == cudaSuccess) e = cudaEventElapsedTime(&msSector, e0, e1);
        }
        for (int rep = 0; rep < 3 && e == cudaSuccess; rep++) {
            cudaEventRecord(e0);
            sectorRateKernel<<<grid, 128>>>(data, sectors - 1u, 512u, sIters * 4u, sink);       /* 512 sectors = 16 KB per block, 8 blocks per SM */
            cudaEventRecord(e1);
            e = cudaEventSynchronize(e1);
            if (e == cudaSuccess) e = cudaEventElapsedTime(&msSectorL1, e0, e1);
        }
        for (int rep = 0; rep < 3 && e == cudaSuccess; rep++) {
            cudaEventRecord(e0);
            issueRateKernel<<<grid, 128>>>(iIters, 1.0f, fsink, cycles);
            cudaEventRecord(e1);
            e = cudaEventSynchronize(e1);
            if (e == cudaSuccess) e = cudaEventElapsedTime(&msIssue, e0, e1);
        }
        if (e == cudaSuccess) e = cudaMemcpy(hCycles, cycles, 16, cudaMemcpyDeviceToHost);
    }
    if (data) cudaFree(data); if (sink) cudaFree(sink); if (fsink) cudaFree(fsink); if (cycles) cudaFree(cycles);
    if (e0) cudaEventDestroy(e0); if (e1) cudaEventDestroy(e1);
    if (e != cudaSuccess) { setError(std::string("okrt_measure_unit_peaks: ") + cudaGetErrorString(e)); return OKRT_ERR_CUDA; }
    const double threads = (double)grid * 128.0;
    out->sectorsPerSecond = threads * sIters * 4.0 * 2.0 / (msSector * 1e-3);         /* 4 loads x 2 sectors per trip */
    out->sectorsPerSecondL1 = threads * (sIters * 4.0) * 4.0 * 2.0 / (msSectorL1 * 1e-3);
    out->warpInstructionsPerSecond = (threads / 32.0) * iIters * 64.0 / (msIssue * 1e-3);
    out->smClockMHz = hCycles[1] ? (double)hCycles[0] / (double)hCycles[1] * 1e3 : 0.0;
    out->numSms = (uint32_t)prop.multiProcessorCount;
    return OKRT_OK;
}

/* ------------------------------------------------------------ debug overlay (row f4)
 * One warp per line segment: lanes stride over the pixels of the major axis.  A pixel keeps the highest draw index that
 * touched it (atomicMax), so overlapping boxes resolve like consecutive Draw calls; the resolve pass writes colours. */
__global__ void overlayRasterKernel(const float4* __restrict__ segments, const uint32_t* __restrict__ keys, uint32_t numSegments,
                                    uint32_t width, uint32_t height, uint32_t* __restrict__ pixelKey)
{
    const uint32_t seg = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (seg >= numSegments) return;
    const float4 s = segments[seg];
    const float dx = s.z - s.x, dy = s.w - s.y;
    const bool xMajor = fabsf(dx) >= fabsf(dy);
    const float a0 = xMajor ? s.x : s.y, a1 = xMajor ? s.z : s.w;          /* major axis */
    const float b0 = xMajor ? s.y : s.x, db = xMajor ? dy : dx, da = xMajor ? dx : dy;
    if (!(da != 0.f) || !isfinite(a0) || !isfinite(a1) || !isfinite(b0) || !isfinite(db)) return;
    const float lo = fminf(a0, a1), hi = fmaxf(a0, a1);
    const float majorLimit = (float)(xMajor ? width : height), minorLimit = (float)(xMajor ? height : width);
    /* pixel centres c + 0.5 in [lo, hi), clamped to the image */
    const float first = fmaxf(ceilf(lo - 0.5f), 0.f), last = fminf(ceilf(hi - 0.5f), majorLimit);   /* [first, last) */
    const uint32_t key = keys[seg];
    for (float c = first + (float)lane; c < last; c += 32.f) {
        const float t = ((c + 0.5f) - a0) / da;
        const float b = floorf(b0 + t * db);
        if (!(b >= 0.f) || !(b < minorLimit)) continue;
        const uint32_t x = xMajor ? (uint32_t)c : (uint32_t)b, y = xMajor ? (uint32_t)b : (uint32_t)c;
        atomicMax(pixelKey + (size_t)y * width + x, key);
    }
}

__global__ void overlayResolveKernel(const uint32_t* __restrict__ pixelKey, const uint32_t* __restrict__ colours, size_t pixels, uchar4* result)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= pixels) return;
    const uint32_t k = pixelKey[i];
    if (!k) return;
    const uint32_t c = colours[k - 1u];
    result[i] = make_uchar4((unsigned char)(c & 0xFFu), (unsigned char)((c >> 8) & 0xFFu), (unsigned char)((c >> 16) & 0xFFu), (unsigned char)(c >> 24));
}

int okrt_draw_overlay(okrt_renderer* r, const okrt_overlay* ov)
{
    int rc = checkHandle(r, "okrt_draw_overlay"); if (rc) return rc;
    if (!ov) { setError("okrt_draw_overlay: null overlay"); return OKRT_ERR_INVALID_ARGUMENT; }
    const uint32_t n = (uint32_t)ov->keys.size();
    if (!n) return OKRT_OK;
    const size_t pixels = (size_t)r->width * r->height;
    /* scratch kept with the renderer (grown on demand): the reference redraws its overlays every frame (DebugRenderer.cpp:264-364) */
    auto grow = [](DeviceBuffer& b, size_t bytes) -> cudaError_t {
        if (b.bytes >= bytes) return cudaSuccess;
        if (b.ptr) cudaFree(b.ptr);
        b.ptr = nullptr; b.bytes = 0;
        const cudaError_t err = cudaMalloc(&b.ptr, bytes);
        if (err == cudaSuccess) b.bytes = bytes;
        return err;
    };
    cudaError_t e = grow(r->ovSegments, (size_t)n * sizeof(float4));
    if (e == cudaSuccess) e = grow(r->ovKeys, (size_t)n * 4);
    if (e == cudaSuccess) e = grow(r->ovColours, ov->colours.size() * 4);
    if (e == cudaSuccess) e = grow(r->ovPixelKey, pixels * 4);
    float4* dSeg = (float4*)r->ovSegments.ptr; uint32_t *dKeys = (uint32_t*)r->ovKeys.ptr, *dColours = (uint32_t*)r->ovColours.ptr, *dPixelKey = (uint32_t*)r->ovPixelKey.ptr;
    if (e == cudaSuccess) e = cudaMemcpyAsync(dSeg, ov->segments.data(), (size_t)n * sizeof(float4), cudaMemcpyHostToDevice, r->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dKeys, ov->keys.data(), (size_t)n * 4, cudaMemcpyHostToDevice, r->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dColours, ov->colours.data(), ov->colours.size() * 4, cudaMemcpyHostToDevice, r->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(dPixelKey, 0, pixels * 4, r->stream);
    if (e == cudaSuccess) {
        overlayRasterKernel<<<(unsigned)(((size_t)n * 32 + 127) / 128), 128, 0, r->stream>>>(dSeg, dKeys, n, r->width, r->height, dPixelKey);
        overlayResolveKernel<<<(unsigned)((pixels + 255) / 256), 256, 0, r->stream>>>(dPixelKey, dColours, pixels, (uchar4*)r->result);
        r->launches += 2;
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(r->stream);          /* the overlay's host arrays may go away on return */
    if (e != cudaSuccess) { setError(std::string("okrt_draw_overlay: ") + cudaGetErrorString(e)); return OKRT_ERR_CUDA; }
    return OKRT_OK;
}

/* ------------------------------------------------------------ okrt_draw_overlay without allocation
 * No counterpart in the reference; added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, serving loops, harnesses that forbid it).  Same two kernels as okrt_draw_overlay; their scratch
 * (pixel keys, the line list, the draw colours) lives in one file-scope device array of a fixed maximum size instead of
 * buffers grown with cudaMalloc.  Every device that loads the library holds the array (9.3 MiB) until exit.
 *   layout: [pixelKey: MAX_PIXELS u32][segments: MAX_SEGMENTS float4][keys: MAX_SEGMENTS u32][colours: MAX_DRAWS u32]
 * In host code `gOverlayStaticScratch` is a host-side stand-in, not a device address: the kernels get the pointer that
 * cudaGetSymbolAddress returns for the current device (looked up once per device, below). */
constexpr size_t kOvStaticPixelKeyBytes = (size_t)OKRT_OVERLAY_STATIC_MAX_PIXELS * 4;
constexpr size_t kOvStaticSegmentsBytes = (size_t)OKRT_OVERLAY_STATIC_MAX_SEGMENTS * sizeof(float4);
constexpr size_t kOvStaticKeysBytes = (size_t)OKRT_OVERLAY_STATIC_MAX_SEGMENTS * 4;
constexpr size_t kOvStaticColoursBytes = (size_t)OKRT_OVERLAY_STATIC_MAX_DRAWS * 4;
constexpr size_t kOvStaticBytes = kOvStaticPixelKeyBytes + kOvStaticSegmentsBytes + kOvStaticKeysBytes + kOvStaticColoursBytes;
static_assert(kOvStaticPixelKeyBytes % 256 == 0 && kOvStaticSegmentsBytes % 256 == 0 && kOvStaticKeysBytes % 256 == 0, "sub-arrays stay 256-byte aligned");
static __device__ __align__(256) unsigned char gOverlayStaticScratch[kOvStaticBytes];

namespace {
constexpr int kOvStaticMaxDevices = 64;
std::mutex gOvStaticMutex;
unsigned char* gOvStaticPtr[kOvStaticMaxDevices] = {};       /* device address of gOverlayStaticScratch, per device ordinal */

/* set-up, once per device, with that device current (checkHandle has made it so) */
int overlayStaticAddress(int device, unsigned char** out)
{
    if (device < 0 || device >= kOvStaticMaxDevices) { setError("okrt_draw_overlay_static: device ordinal beyond 63"); return OKRT_ERR_INVALID_ARGUMENT; }
    std::lock_guard<std::mutex> lock(gOvStaticMutex);
    if (!gOvStaticPtr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, gOverlayStaticScratch);
        if (e != cudaSuccess || !p) {
            setError(std::string("okrt_draw_overlay_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
            return OKRT_ERR_CUDA;
        }
        gOvStaticPtr[device] = (unsigned char*)p;
    }
    *out = gOvStaticPtr[device];
    return OKRT_OK;
}
} // namespace

int okrt_draw_overlay_static(okrt_renderer* r, const okrt_overlay* ov)
{
    int rc = checkHandle(r, "okrt_draw_overlay_static"); if (rc) return rc;
    if (!ov) { setError("okrt_draw_overlay_static: null overlay"); return OKRT_ERR_INVALID_ARGUMENT; }
    const size_t n = ov->keys.size(), draws = ov->colours.size(), pixels = (size_t)r->width * r->height;
    if (pixels > OKRT_OVERLAY_STATIC_MAX_PIXELS || n > OKRT_OVERLAY_STATIC_MAX_SEGMENTS || draws > OKRT_OVERLAY_STATIC_MAX_DRAWS) {
        setError("okrt_draw_overlay_static: " + std::to_string(pixels) + " pixels / " + std::to_string(n) + " segments / " + std::to_string(draws) +
                 " draws exceed the static scratch (" + std::to_string(OKRT_OVERLAY_STATIC_MAX_PIXELS) + " / " + std::to_string(OKRT_OVERLAY_STATIC_MAX_SEGMENTS) +
                 " / " + std::to_string(OKRT_OVERLAY_STATIC_MAX_DRAWS) + "): use okrt_draw_overlay");
        return OKRT_ERR_INVALID_ARGUMENT;
    }
    unsigned char* base = nullptr;
    if ((rc = overlayStaticAddress(r->device, &base))) return rc;       /* an empty overlay still does the set-up: a warm-up call */
    if (!n) return OKRT_OK;
    uint32_t* dPixelKey = (uint32_t*)base;
    float4* dSeg = (float4*)(base + kOvStaticPixelKeyBytes);
    uint32_t* dKeys = (uint32_t*)(base + kOvStaticPixelKeyBytes + kOvStaticSegmentsBytes);
    uint32_t* dColours = (uint32_t*)(base + kOvStaticPixelKeyBytes + kOvStaticSegmentsBytes + kOvStaticKeysBytes);
    /* the array keeps the last call's contents: the pixel keys must start at zero, the rest is overwritten up to what is read */
    cudaError_t e = cudaMemsetAsync(dPixelKey, 0, pixels * 4, r->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dSeg, ov->segments.data(), n * sizeof(float4), cudaMemcpyHostToDevice, r->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dKeys, ov->keys.data(), n * 4, cudaMemcpyHostToDevice, r->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dColours, ov->colours.data(), draws * 4, cudaMemcpyHostToDevice, r->stream);
    if (e == cudaSuccess) {
        overlayRasterKernel<<<(unsigned)((n * 32 + 127) / 128), 128, 0, r->stream>>>(dSeg, dKeys, (uint32_t)n, r->width, r->height, dPixelKey);
        overlayResolveKernel<<<(unsigned)((pixels + 255) / 256), 256, 0, r->stream>>>(dPixelKey, dColours, pixels, (uchar4*)r->result);
        r->launches += 2;
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(r->stream);          /* the overlay's host arrays may go away on return */
    if (e != cudaSuccess) { setError(std::string("okrt_draw_overlay_static: ") + cudaGetErrorString(e)); return OKRT_ERR_CUDA; }
    return OKRT_OK;
}

int okrt_read_accumulation(okrt_renderer* r, float* outRgba32f)
{
    int rc = checkHandle(r, "okrt_read_accumulation"); if (rc) return rc;
    if (!outRgba32f) { setError("okrt_read_accumulation: null output"); return OKRT_ERR_INVALID_ARGUMENT; }
    OKRT_CUDA(cudaMemcpyAsync(outRgba32f, r->accumulation, (size_t)r->width * r->height * 16, cudaMemcpyDeviceToHost, r->stream));
    OKRT_CUDA(cudaStreamSynchronize(r->stream));
    return OKRT_OK;
}

int okrt_read_aov(okrt_renderer* r, okrt_aov* out)
{
    int rc = checkHandle(r, "okrt_read_aov"); if (rc) return rc;
    if (!out) { setError("okrt_read_aov: null output"); return OKRT_ERR_INVALID_ARGUMENT; }
    if (!r->aov) { setError("okrt_read_aov: render with options.writeAov = 1 first"); return OKRT_ERR_STATE; }
    OKRT_CUDA(cudaMemcpyAsync(out, r->aov, (size_t)r->width * r->height * sizeof(okrt_aov), cudaMemcpyDeviceToHost, r->stream));
    OKRT_CUDA(cudaStreamSynchronize(r->stream));
    return OKRT_OK;
}

int okrt_write_accumulation(okrt_renderer* r, const float* rgba32f, uint32_t numAccumulationFrames)
{
    int rc = checkHandle(r, "okrt_write_accumulation"); if (rc) return rc;
    if (!rgba32f) { setError("okrt_write_accumulation: null input"); return OKRT_ERR_INVALID_ARGUMENT; }
    OKRT_CUDA(cudaMemcpyAsync(r->accumulation, rgba32f, (size_t)r->width * r->height * 16, cudaMemcpyHostToDevice, r->stream));
    OKRT_CUDA(cudaStreamSynchronize(r->stream));
    r->rd.numAccumulationFrames = numAccumulationFrames;
    return OKRT_OK;
}

int okrt_resolve(okrt_renderer* r, uint32_t totalFrames)
{
    int rc = checkHandle(r, "okrt_resolve"); if (rc) return rc;
    const size_t n = (size_t)r->width * r->height;
    uint32_t cw, ch;
    coveredExtent(r, cw, ch);
    resolveKernel<<<(unsigned)((n + 255) / 256), 256, 0, r->stream>>>(r->accumulation, r->result, n, (float)totalFrames, r->width, cw, ch);
    r->launches++;
    OKRT_CUDA(cudaGetLastError());
    return OKRT_OK;
}

int okrt_host_register(void* ptr, size_t bytes)
{
    if (!ptr || !bytes) { setError("okrt_host_register: null argument"); return OKRT_ERR_INVALID_ARGUMENT; }
    OKRT_CUDA(cudaHostRegister(ptr, bytes, cudaHostRegisterDefault));
    return OKRT_OK;
}

int okrt_host_unregister(void* ptr)
{
    if (!ptr) { setError("okrt_host_unregister: null argument"); return OKRT_ERR_INVALID_ARGUMENT; }
    OKRT_CUDA(cudaHostUnregister(ptr));
    return OKRT_OK;
}

void* okrt_accumulation_device_ptr(okrt_renderer* r) { return r ? (void*)r->accumulation : nullptr; }
void* okrt_result_device_ptr(okrt_renderer* r) { return r ? (void*)r->result : nullptr; }
void* okrt_cuda_stream(okrt_renderer* r) { return r ? (void*)r->stream : nullptr; }

int okrt_get_stats(okrt_renderer* r, okrt_stats* out)
{
    int rc = checkHandle(r, "okrt_get_stats"); if (rc) return rc;
    if (!out) { setError("okrt_get_stats: null output"); return OKRT_ERR_INVALID_ARGUMENT; }
    unsigned long long h[4] = {0, 0, 0, 0};
    OKRT_CUDA(cudaMemcpyAsync(h, r->counters, sizeof(h), cudaMemcpyDeviceToHost, r->stream));
    OKRT_CUDA(cudaStreamSynchronize(r->stream));
    std::memset(out, 0, sizeof(*out));
    out->raySegments = h[0]; out->fallbackSegments = h[1]; out->pixelsTraced = h[2];
    out->kernelLaunches = r->launches;
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, r->evStart, r->evStop) == cudaSuccess) r->lastRenderMs = ms; else cudaGetLastError();
    out->lastRenderMs = r->lastRenderMs;
    float traceMs = 0.f;
    for (size_t i = 0; i + 1 < r->traceEventsUsed; i += 2) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, r->traceEvents[i], r->traceEvents[i + 1]) == cudaSuccess) traceMs += t; else cudaGetLastError();
    }
    out->traceKernelMs = traceMs;
    out->traceKernelLaunches = (float)(r->traceEventsUsed / 2);
    return OKRT_OK;
}

/* Step counters of the traversal kernel since okrt_reset_stats, 8 per bounce (okrt_wavefront.cuh, traceWideKernel).  Only a
 * library built with -DOKRT_TRACE_STATS counts (tools/: libokrt_stats.so); the product build returns OKRT_ERR_STATE. */
int okrt_get_trace_stats(okrt_renderer* r, uint64_t* out, uint32_t bounces)
{
    int rc = checkHandle(r, "okrt_get_trace_stats"); if (rc) return rc;
    if (!out || bounces > kMaxWavefrontBounces + 1) { setError("okrt_get_trace_stats: bad argument"); return OKRT_ERR_INVALID_ARGUMENT; }
#ifdef OKRT_TRACE_STATS
    if (!r->traceStats) { std::memset(out, 0, (size_t)bounces * 8 * sizeof(uint64_t)); return OKRT_OK; }
    OKRT_CUDA(cudaMemcpyAsync(out, r->traceStats, (size_t)bounces * 8 * sizeof(uint64_t), cudaMemcpyDeviceToHost, r->stream));
    OKRT_CUDA(cudaStreamSynchronize(r->stream));
    return OKRT_OK;
#else
    setError("okrt_get_trace_stats: this library was built without -DOKRT_TRACE_STATS");
    return OKRT_ERR_STATE;
#endif
}

/* Host-only check of the wide traversal (no GPU): builds the traversal layout from the binding-table arrays and runs the
 * kernels' own lane functions (okrt_wide.cuh, compiled for the host) over `numRays` rays (6 floats each: origin, direction).
 * out: 12 words per ray — distance, u, v (float bits), triangle, kind, instance-or-sphere index, needs-reference-order flag,
 * the bound left for the reference-order traversal (float bits), node steps, leaf entries, triangle tests, 0.
 * The step counts need a -DOKRT_TRACE_STATS build (0 otherwise). */
int okrt_trace_rays_host(const okrt_triangle* tris, uint32_t numTris, const okrt_bvh_node* nodes, uint32_t numNodes,
                         const okrt_mesh_instance* instances, uint32_t numInstances, const okrt_octree_node* octNodes, uint32_t numOctNodes,
                         uint32_t traversal, const float* rays, uint32_t numRays, uint32_t* out)
{
    if ((numRays && (!rays || !out)) || (traversal != OKRT_TRAVERSAL_FAST && traversal != OKRT_TRAVERSAL_WIDE)) { setError("okrt_trace_rays_host: bad argument"); return OKRT_ERR_INVALID_ARGUMENT; }
    const bool wide = traversal == OKRT_TRAVERSAL_WIDE;
    FastLayout fl;
    if (!buildFastLayout(tris, numTris, nodes, numNodes, instances, numInstances, octNodes, numOctNodes, fl, wide)) {
        setError("scene rejected: " + fl.error);
        return OKRT_ERR_INVALID_ARGUMENT;
    }
    if (fl.wideStackDepth > kWideStackCapacity || fl.maxStackDepth > kFastStackCapacity) { setError("traversal stack would overflow"); return OKRT_ERR_CAPACITY; }
    if (!fl.allRigid) { setError("okrt_trace_rays_host: scaled instances use the reference order"); return OKRT_ERR_STATE; }
    DeviceScene s;
    std::memset(&s, 0, sizeof(s));
    s.octNodes = octNodes; s.instances = instances; s.numOctNodes = numOctNodes; s.numInstances = numInstances; s.numTris = numTris;
    s.fastTris = (const float4*)fl.tris.data(); s.fastInstances = (const float4*)fl.instances.data(); s.octParent = fl.octParent.data();
    s.numFastInstances = (uint32_t)(fl.instances.size() / kInstQuads);
    s.topRef = fl.topRef; s.numFastNodes = (uint32_t)(fl.nodes.size() / 4);
    /* the node and leaf arrays are read with 32-byte loads */
    std::vector<uint4> pn((fl.nodes.size() * sizeof(Float4) + 31) / 32 * 2);
    if (!fl.nodes.empty()) std::memcpy(pn.data(), fl.nodes.data(), fl.nodes.size() * sizeof(Float4));
    s.fastNodes = (const float4*)pn.data();
    std::vector<uint4> wn((fl.wideNodes.size() + 3) / 4), wl((fl.wideLeaves.size() + 3) / 4);
    if (!fl.wideNodes.empty()) std::memcpy(wn.data(), fl.wideNodes.data(), fl.wideNodes.size() * 4);
    if (!fl.wideLeaves.empty()) std::memcpy(wl.data(), fl.wideLeaves.data(), fl.wideLeaves.size() * 4);
    s.wideNodes = wn.data(); s.wideLeaves = wl.data(); s.wideRoot = fl.wideRoot; s.numWideNodes = (uint32_t)(fl.wideNodes.size() / kWideNodeWords);
    s.maxSlack = fl.maxSlack;
    const unsigned nthreads = std::max(1u, std::thread::hardware_concurrency());
    std::vector<std::thread> pool;
    std::atomic<uint32_t> nextRay(0);
    for (unsigned t = 0; t < nthreads; t++)
        pool.emplace_back([&]() {
            for (;;) {
                const uint32_t begin = nextRay.fetch_add(1024u);
                if (begin >= numRays) break;
                for (uint32_t i = begin; i < std::min(numRays, begin + 1024u); i++) {
                    const float* r6 = rays + (size_t)i * 6;
                    Hit hit; float limit = 0.f; uint32_t counts[4] = {0u, 0u, 0u, 0u};
                    const V3 ro = v3(r6[0], r6[1], r6[2]), rd = v3(r6[3], r6[4], r6[5]);
                    const bool ambiguous = wide ? wideTraverse(s, 0u, ro, rd, hit, limit, counts) : traverseFast(s, 0u, ro, rd, hit, limit);
                    uint32_t* o = out + (size_t)i * 12;
                    std::memcpy(o + 0, &hit.distance, 4); std::memcpy(o + 1, &hit.u, 4); std::memcpy(o + 2, &hit.v, 4);
                    o[3] = hit.tri; o[4] = hit.kind; o[5] = hit.idx; o[6] = ambiguous ? 1u : 0u; std::memcpy(o + 7, &limit, 4);
                    o[8] = counts[0]; o[9] = counts[1]; o[10] = counts[2]; o[11] = counts[3];
                }
            }
        });
    for (auto& th : pool) th.join();
    return OKRT_OK;
}

int okrt_wait_layout(okrt_renderer* r)
{
    int rc = checkHandle(r, "okrt_wait_layout"); if (rc) return rc;
    if (r->layoutDirty && (rc = rebuildLayout(r))) return rc;
    return adoptRefined(r, true);
}

int okrt_get_layout_build(okrt_renderer* r, okrt_layout_build* out)
{
    if (!r || !out) { setError("okrt_get_layout_build: null argument"); return OKRT_ERR_INVALID_ARGUMENT; }
    *out = r->layoutBuild;
    return OKRT_OK;
}

int okrt_reset_stats(okrt_renderer* r)
{
    int rc = checkHandle(r, "okrt_reset_stats"); if (rc) return rc;
    OKRT_CUDA(cudaMemsetAsync(r->counters, 0, 4 * sizeof(unsigned long long), r->stream));
    if (r->traceStats) OKRT_CUDA(cudaMemsetAsync(r->traceStats, 0, 8 * (kMaxWavefrontBounces + 2) * sizeof(unsigned long long), r->stream));
    r->launches = 0;
    return OKRT_OK;
}

} // extern "C"

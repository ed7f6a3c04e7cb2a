// rt_api.cu — host side of librt_b200.so: the C ABI declared in include/rt_b200.h.
//
// Replaces the reference's PyOpenGL call sites (see the table at the top of rt_b200.h).  A context owns
// one CUDA device's state: the re-laid-out scene, the uniform block, two RGBA32F accumulation targets
// (ping-pong, screen.py:37-47) and a stream.  All device work is stream-ordered; read-backs synchronise.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <limits>
#include <string>
#include <type_traits>
#include <vector>

#include "../../include/rt_b200.h"
#include "rt_types.h"

// launchers (rt_kernels.cu)
size_t rtPixelKernelSmem(int stackDepth);
cudaError_t rtLaunchPixel(const PassParams& P, bool count, bool ids, cudaStream_t stream);
cudaError_t rtLaunchPersistent(const PassParams& P, bool count, bool ids, int sms, cudaStream_t stream);
cudaError_t rtLaunchPool(const PassParams& P, bool count, bool ids, int sms, cudaStream_t stream);
cudaError_t rtLaunchWavefront(const PassParams& P, bool count, bool ids, int sms, cudaStream_t stream);
cudaError_t rtLaunchWavefrontStatic(const PassParams& P, int device, int sms, cudaStream_t stream);
cudaError_t rtLaunchSamples(const PassParams& P, bool ids, int sms, uint32_t* todoA, uint32_t* todoB, cudaStream_t stream);
cudaError_t rtLaunchChain(const PassParams& P, bool ids, int sms, cudaStream_t stream);
size_t rtChainRingBytes(int sms);
cudaError_t rtMeasurePeaks(int sms, cudaStream_t stream, double* out);
int rtSpecRounds();
int rtSpecLaunches();
int rtSpecWorkUints();
cudaError_t rtSpecInit();
int rtSpecTableK(int raysPerPixel, int nBounces);
cudaError_t rtLaunchToRgba8(const float4* src, uchar4* dst, size_t n, cudaStream_t stream);
cudaError_t rtLaunchBandCopy(float4* target, float4* dense, int width, int rows, int tileSizeY, int numTilesY,
                             int tile, int toDense, cudaStream_t stream);
cudaError_t rtLaunchFinishSum(float4* target, size_t n, float frames, cudaStream_t stream);
cudaError_t rtLaunchScaleSum(float4* target, size_t n, float frames, cudaStream_t stream);
cudaError_t rtLaunchBuildRecords(const float4* raw, const uint32_t* triId, uint32_t nSlots, TriRec* tris, ShadeRec* shade,
                                 cudaStream_t stream);

struct rt_ctx {
    int device = 0;
    int smCount = 148;
    cudaStream_t ownStream = nullptr, stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    std::string err;
    // scene
    // Two sets of scene buffers: an upload fills the set that is NOT in use, on its own stream, while passes already
    // launched keep reading the other one (they hold its pointers by value) — a re-upload overlaps the previous render.
    struct SceneSet {
        NodeRec* dNodes = nullptr;
        TriRec* dTris = nullptr;
        ShadeRec* dShade = nullptr;
        uint32_t* dTriId = nullptr;
        uint2* dBig = nullptr;
        float4* dRawTris = nullptr;       // the caller's tris[] as uploaded (144 B each); records are built from it on the device
        size_t capNodes = 0, capSlots = 0, capSlotRecs = 0, capBig = 0, capRaw = 0;   // element capacities: re-uploads reuse the buffers
        void* hStage = nullptr;           // pinned staging for nodes + slot ids + big leaves
        size_t hStageBytes = 0;
        cudaEvent_t lastUse = nullptr;    // recorded after every pass that reads this set
    } sets[2];
    int curSet = 0;
    cudaStream_t upStream = nullptr, downStream = nullptr;   // H2D of an upload / D2H of an asynchronous read-back
    cudaEvent_t evRendered = nullptr;                        // main stream -> downStream
    cudaEvent_t evSwitch = nullptr;                          // old main stream -> new main stream (rt_set_stream)
    cudaEvent_t evDown[2] = {nullptr, nullptr};              // read-back of target[i] finished (awaited before target[i] is written)
    bool downPending[2] = {false, false};
    DeviceScene scene{};
    bool haveScene = false;
    // uniforms
    rt_params params{};
    bool haveParams = false;
    int variant = RT_VARIANT_AUTO;
    // targets
    int width = 0, height = 0;
    float4* target[2] = {nullptr, nullptr};
    float4* peerTarget[2] = {nullptr, nullptr};              // the assembling rank's targets, mapped through CUDA IPC (rt_peer_open)
    int accumIndex = 0;
    int32_t* dIds = nullptr;
    uchar4* dRgba8 = nullptr;
    unsigned long long* dCounters = nullptr;
    unsigned int* dWork = nullptr;
    // wavefront variant: queue of pixels whose primary ray hit (allocated on first use, sized to the target)
    float4* dQueueA = nullptr;
    float4* dQueueB = nullptr;
    uint32_t* dQueuePix = nullptr;
    uint4* dQueueWork = nullptr;
    unsigned int* dQueueCount = nullptr;
    size_t queueCapacity = 0;
    // samples variant: three more record planes and the table of F(k)
    float4* dQueueC = nullptr;
    float4* dQueueD = nullptr;
    float4* dQueueE = nullptr;
    float4* dSpecBuf = nullptr;
    size_t specRecords = 0, specBufEntries = 0, recPlanes = 0;
    unsigned int* dWorkArr = nullptr;     // work counters of the rounds + todo counts
    uint32_t* dTodo[2] = {nullptr, nullptr};
    uint32_t* dChainClass = nullptr;
    int lastLaunches = 0;                 // kernel launches of the last pass (rt_last_pass_launches)
    int lastVariant = 0;                  // what the last pass ran (rt_last_pass_variant)
    bool lastGated = false;
    long long lastOwned = 0;
    uint32_t lastCapacity = 0;
    uint32_t* dChainK = nullptr;
    float4* dChainColor = nullptr;
    float4* dChainRing = nullptr;         // chain variant: rings of the helped pixel slots (allocated on first use)
};

namespace {

int fail(rt_ctx* c, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (c) c->err = buf;
    return code;
}

#define RT_CUDA(c, call)                                                                                  \
    do {                                                                                                  \
        cudaError_t e_ = (call);                                                                          \
        if (e_ != cudaSuccess) return fail((c), RT_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_));   \
    } while (0)

struct Bind {   // make the context's device current for the duration of a call
    int prev = -1;
    explicit Bind(rt_ctx* c) {
        cudaGetDevice(&prev);
        if (prev != c->device) cudaSetDevice(c->device);
        else prev = -1;
    }
    ~Bind() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

// std430 records as the reference uploads them (scene.py:45-70).
struct HostTri {
    float v0p[3], p0, v0n[3], p1, v1p[3], p2, v1n[3], p3, v2p[3], p4, v2n[3], p5;
    float color[3], p6, emc[3], p7, surface[2], p8[2];
};
struct HostBox {
    uint32_t numTriangles, triangleOffset;
    int32_t childA, childB;
    float posMin[3], p0, posMax[3], p1;
};
static_assert(sizeof(HostTri) == 144 && sizeof(HostBox) == 48, "std430 layout");

void freeScene(rt_ctx* c) {
    for (auto& S : c->sets) {
        cudaFree(S.dNodes); cudaFree(S.dTris); cudaFree(S.dShade); cudaFree(S.dTriId); cudaFree(S.dBig);
        cudaFree(S.dRawTris); cudaFreeHost(S.hStage);
        S.dNodes = nullptr; S.dTris = nullptr; S.dShade = nullptr; S.dTriId = nullptr; S.dBig = nullptr;
        S.dRawTris = nullptr; S.hStage = nullptr;
        S.capNodes = S.capSlots = S.capSlotRecs = S.capBig = S.capRaw = S.hStageBytes = 0;
    }
    c->haveScene = false;
    c->scene = DeviceScene{};
}

// Main-stream work that writes an accumulation target first waits for read-backs still copying out of it.
void waitDownloads(rt_ctx* c, int which = -1) {      // which: target 0 / 1, or -1 for both
    for (int i = 0; i < 2; i++)
        if ((which < 0 || which == i) && c->downPending[i]) {
            cudaStreamWaitEvent(c->stream, c->evDown[i], 0);
            c->downPending[i] = false;
        }
}

void freeQueue(rt_ctx* c) {
    cudaFree(c->dQueueA); cudaFree(c->dQueueB); cudaFree(c->dQueuePix); cudaFree(c->dQueueWork);
    cudaFree(c->dQueueC); cudaFree(c->dQueueD); cudaFree(c->dQueueE); cudaFree(c->dSpecBuf);
    cudaFree(c->dTodo[0]); cudaFree(c->dTodo[1]); cudaFree(c->dChainK); cudaFree(c->dChainColor); cudaFree(c->dChainClass);
    c->dTodo[0] = c->dTodo[1] = nullptr; c->dChainK = nullptr; c->dChainColor = nullptr; c->dChainClass = nullptr;
    c->dQueueA = c->dQueueB = nullptr; c->dQueuePix = nullptr; c->dQueueWork = nullptr;
    c->dQueueC = c->dQueueD = c->dQueueE = c->dSpecBuf = nullptr;
    c->queueCapacity = 0; c->specRecords = 0; c->specBufEntries = 0; c->recPlanes = 0;
}

cudaError_t ensureQueue(rt_ctx* c, bool withWork);

// Samples and chain variants: the five record planes of the queue (80 bytes per pixel of the target).
cudaError_t ensureRecordPlanes(rt_ctx* c) {
    cudaError_t e = ensureQueue(c, false);
    if (e != cudaSuccess) return e;
    if (c->recPlanes < c->queueCapacity) {
        cudaFree(c->dQueueC); cudaFree(c->dQueueD); cudaFree(c->dQueueE);
        c->dQueueC = c->dQueueD = c->dQueueE = nullptr; c->recPlanes = 0;
        if ((e = cudaMalloc(&c->dQueueC, c->queueCapacity * sizeof(float4))) != cudaSuccess) return e;
        if ((e = cudaMalloc(&c->dQueueD, c->queueCapacity * sizeof(float4))) != cudaSuccess) return e;
        if ((e = cudaMalloc(&c->dQueueE, c->queueCapacity * sizeof(float4))) != cudaSuccess) return e;
        c->recPlanes = c->queueCapacity;
    }
    return cudaSuccess;
}

// Samples variant: record planes for every pixel of the target, and a table of K entries for up to half of the
// pixels this pass owns (the speculative mode is only entered when at most half of them were queued).
cudaError_t ensureSpec(rt_ctx* c, size_t ownedPixels, int K, uint32_t* capacityOut) {
    cudaError_t e = ensureRecordPlanes(c);
    if (e != cudaSuccess) return e;
    if (c->specRecords < c->queueCapacity) {
        cudaFree(c->dTodo[0]); cudaFree(c->dTodo[1]); cudaFree(c->dChainK); cudaFree(c->dChainColor); cudaFree(c->dChainClass);
        c->specRecords = 0;
        c->dTodo[0] = c->dTodo[1] = nullptr; c->dChainK = nullptr; c->dChainColor = nullptr; c->dChainClass = nullptr;
        if ((e = cudaMalloc(&c->dTodo[0], c->queueCapacity * sizeof(uint32_t))) != cudaSuccess) return e;
        if ((e = cudaMalloc(&c->dTodo[1], c->queueCapacity * sizeof(uint32_t))) != cudaSuccess) return e;
        if ((e = cudaMalloc(&c->dChainK, c->queueCapacity * sizeof(uint32_t))) != cudaSuccess) return e;
        if ((e = cudaMalloc(&c->dChainClass, c->queueCapacity * sizeof(uint32_t))) != cudaSuccess) return e;
        if ((e = cudaMalloc(&c->dChainColor, c->queueCapacity * sizeof(float4))) != cudaSuccess) return e;
        c->specRecords = c->queueCapacity;
    }
    // (test tooling: RT_VOTE_TUNE=..,..,1 forces the speculative mode, which then needs room for every pixel)
    int t0 = 0, t1 = 0, t2 = 0;
    if (const char* v = getenv("RT_VOTE_TUNE")) sscanf(v, "%d,%d,%d", &t0, &t1, &t2);
    const size_t cap = t2 == 1 ? ownedPixels : (ownedPixels + 1) / 2;
    const size_t need = std::max<size_t>(cap * size_t(std::max(K, 0)), 1);
    if (c->specBufEntries < need) {
        cudaFree(c->dSpecBuf); c->dSpecBuf = nullptr; c->specBufEntries = 0;
        if ((e = cudaMalloc(&c->dSpecBuf, need * sizeof(float4))) != cudaSuccess) {
            cudaGetLastError();
            *capacityOut = 0;          // no room for the table: stage 3 evaluates every sample on demand
            return cudaSuccess;
        }
        c->specBufEntries = need;
    }
    *capacityOut = uint32_t(std::min<size_t>(K > 0 ? c->specBufEntries / size_t(K) : 0, 0xffffffffu));
    return cudaSuccess;
}

// The queue can hold every pixel of the target (all primary rays may hit); the work array only exists once a
// counting pass asked for it.
cudaError_t ensureQueue(rt_ctx* c, bool withWork) {
    const size_t n = size_t(c->width) * c->height;
    cudaError_t e = cudaSuccess;
    if (c->queueCapacity < n) {
        freeQueue(c);
        if ((e = cudaMalloc(&c->dQueueA, n * sizeof(float4))) != cudaSuccess) return e;
        if ((e = cudaMalloc(&c->dQueueB, n * sizeof(float4))) != cudaSuccess) return e;
        if ((e = cudaMalloc(&c->dQueuePix, n * sizeof(uint32_t))) != cudaSuccess) return e;
        c->queueCapacity = n;
    }
    if (withWork && !c->dQueueWork) e = cudaMalloc(&c->dQueueWork, c->queueCapacity * sizeof(uint4));
    return e;
}

void freeTargets(rt_ctx* c) {
    freeQueue(c);
    cudaFree(c->target[0]); cudaFree(c->target[1]); cudaFree(c->dIds); cudaFree(c->dRgba8);
    for (int i = 0; i < 2; i++) {           // peer mappings are of the old size: a new target means a new rt_peer_open
        if (c->peerTarget[i]) cudaIpcCloseMemHandle(c->peerTarget[i]);
        c->peerTarget[i] = nullptr;
    }
    c->target[0] = c->target[1] = nullptr; c->dIds = nullptr; c->dRgba8 = nullptr;
    c->width = c->height = 0;
    c->accumIndex = 0;
}

// How many coordinates c in [0, extent) satisfy (c / tileSize) % numTiles == tile.
// Whether triangleTest defers the fetch of a triangle's edges (rt_device.cuh): yes when the pass is large enough to keep
// the L1 data pipe busy (8 Mi pixel-samples: a 4K frame at 1 spp, 1080p at 4), no for small, latency-bound passes.
// Measured on B200 (profiles/r02_variants.md section 10); RT_DEFER_EDGES=0/1 overrides.
int deferEdgesFor(long long ownedPixels, int raysPerPixel) {
    if (const char* e = getenv("RT_DEFER_EDGES")) return atoi(e) != 0;
    return ownedPixels * std::max(raysPerPixel, 1) >= (8ll << 20) ? 1 : 0;
}

int ownedCount(int extent, int tileSize, int numTiles, int tile) {
    if (tileSize <= 0 || numTiles <= 0 || tile < 0 || tile >= numTiles) return 0;
    const long long period = (long long)tileSize * numTiles;
    const long long full = extent / period;
    long long rest = extent - full * period - (long long)tile * tileSize;
    rest = std::max(0LL, std::min<long long>(rest, tileSize));
    return int(full * tileSize + rest);
}

}  // namespace

extern "C" {

const char* rt_version(void) { return "librt_b200 0.1 (sm_100a)"; }

const char* rt_last_error(rt_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

int rt_create(rt_ctx** out, int device) {
    if (!out) return RT_ERR_INVALID;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) return RT_ERR_CUDA;
    if (device < 0 || device >= n) return RT_ERR_INVALID;
    rt_ctx* c = new rt_ctx();
    c->device = device;
    Bind bind(c);
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { delete c; return RT_ERR_CUDA; }
    c->smCount = prop.multiProcessorCount;
    if (cudaStreamCreateWithFlags(&c->ownStream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->upStream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->downStream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->evRendered, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->evSwitch, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->evDown[0], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->evDown[1], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->sets[0].lastUse, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->sets[1].lastUse, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreate(&c->ev0) != cudaSuccess || cudaEventCreate(&c->ev1) != cudaSuccess ||
        cudaMalloc(&c->dCounters, RT_CNT_COUNT * sizeof(unsigned long long)) != cudaSuccess ||
        cudaMalloc(&c->dWork, sizeof(unsigned int)) != cudaSuccess ||
        cudaMalloc(&c->dQueueCount, sizeof(unsigned int)) != cudaSuccess ||
        cudaMalloc(&c->dWorkArr, size_t(rtSpecWorkUints()) * sizeof(unsigned int)) != cudaSuccess || rtSpecInit() != cudaSuccess) {
        cudaGetLastError();
        rt_destroy(c);          // releases whatever was created (every member starts as null)
        return RT_ERR_CUDA;
    }
    c->stream = c->ownStream;
    cudaMemset(c->dCounters, 0, RT_CNT_COUNT * sizeof(unsigned long long));
    *out = c;
    return RT_OK;
}

void rt_destroy(rt_ctx* c) {
    if (!c) return;
    Bind bind(c);
    cudaStreamSynchronize(c->stream);
    if (c->upStream) cudaStreamSynchronize(c->upStream);
    if (c->downStream) cudaStreamSynchronize(c->downStream);
    freeScene(c);
    freeTargets(c);
    cudaFree(c->dCounters); cudaFree(c->dWork); cudaFree(c->dQueueCount); cudaFree(c->dWorkArr); cudaFree(c->dChainRing);
    for (cudaEvent_t e : {c->ev0, c->ev1, c->evRendered, c->evSwitch, c->evDown[0], c->evDown[1], c->sets[0].lastUse, c->sets[1].lastUse})
        if (e) cudaEventDestroy(e);
    for (cudaStream_t st : {c->ownStream, c->upStream, c->downStream})
        if (st) cudaStreamDestroy(st);
    delete c;
}

// The new stream is ordered behind everything already queued on the old one (an event recorded there, awaited here), so
// a caller may move the context between its own stream and, say, torch's current stream at any point without a host
// synchronisation and without losing the order of the work (distributed.py does that around its collectives).
int rt_set_stream(rt_ctx* c, void* s) {
    if (!c) return RT_ERR_INVALID;
    cudaStream_t next = s ? static_cast<cudaStream_t>(s) : c->ownStream;
    if (next == c->stream) return RT_OK;
    Bind bind(c);
    RT_CUDA(c, cudaEventRecord(c->evSwitch, c->stream));
    RT_CUDA(c, cudaStreamWaitEvent(next, c->evSwitch, 0));
    c->stream = next;
    return RT_OK;
}

void* rt_get_stream(rt_ctx* c) { return c ? static_cast<void*>(c->stream) : nullptr; }

int rt_sync(rt_ctx* c) {
    if (!c) return RT_ERR_INVALID;
    Bind bind(c);
    RT_CUDA(c, cudaStreamSynchronize(c->stream));
    RT_CUDA(c, cudaStreamSynchronize(c->downStream));     // read-backs started with rt_read_accum_f32_async
    return RT_OK;
}

int rt_scene_free(rt_ctx* c) {
    if (!c) return RT_ERR_INVALID;
    Bind bind(c);
    cudaStreamSynchronize(c->stream);
    cudaStreamSynchronize(c->upStream);
    freeScene(c);
    return RT_OK;
}

// scene.py:223-236 — validates the three SSBO payloads, then builds the device layout (rt_types.h):
// internal nodes in depth-first order from the root, triangles in leaf order.  Unreachable nodes
// (the reference's builder leaves dead single-child chains in the array) are dropped.
int rt_scene_upload(rt_ctx* c, const void* trisV, uint32_t nTris, const void* boxesV, uint32_t nBoxes,
                    const uint32_t* indices, uint32_t nIndices) {
    if (!c) return RT_ERR_INVALID;
    if ((nTris && !trisV) || (nBoxes && !boxesV) || (nIndices && !indices))
        return fail(c, RT_ERR_INVALID, "rt_scene_upload: null buffer with non-zero count");
    Bind bind(c);
    const HostTri* tris = static_cast<const HostTri*>(trisV);
    const HostBox* boxes = static_cast<const HostBox*>(boxesV);

    std::vector<NodeRec> nodes;
    std::vector<uint32_t> triIds;                  // slot -> index into the caller's tris[]
    std::vector<uint2> bigLeaves;
    DeviceScene sc{};
    sc.rootRef = RT_REF_NONE;
    sc.stackDepth = 1;

    if (nBoxes > 0) {
        // 1. reachability, validation, cycle check and height (iterative DFS)
        std::vector<int8_t> state(nBoxes, 0);          // 0 new, 1 open, 2 done
        std::vector<int32_t> height(nBoxes, 0);        // internal nodes on the longest path below (incl. self)
        std::vector<uint32_t> work{0};
        while (!work.empty()) {
            const uint32_t i = work.back();
            const HostBox& b = boxes[i];
            if (state[i] == 2) { work.pop_back(); continue; }
            if (b.childA == -1) {                       // leaf to the shader (fragment.glsl:266)
                if (uint64_t(b.triangleOffset) + b.numTriangles > nIndices)
                    return fail(c, RT_ERR_INVALID, "box %u: triangle range [%u,+%u) exceeds %u indices", i,
                                b.triangleOffset, b.numTriangles, nIndices);
                for (uint32_t k = 0; k < b.numTriangles; k++)
                    if (indices[b.triangleOffset + k] >= nTris)
                        return fail(c, RT_ERR_INVALID, "box %u: triangle index %u >= %u", i,
                                    indices[b.triangleOffset + k], nTris);
                state[i] = 2; height[i] = 0; work.pop_back();
                continue;
            }
            if (b.childA < 0 || uint32_t(b.childA) >= nBoxes || b.childB < -1 ||
                (b.childB >= 0 && uint32_t(b.childB) >= nBoxes))
                return fail(c, RT_ERR_INVALID, "box %u: child index out of range (%d, %d)", i, b.childA, b.childB);
            const int32_t kids[2] = {b.childA, b.childB};
            if (state[i] == 0) {
                state[i] = 1;
                for (int32_t k : kids) {
                    if (k < 0) continue;
                    if (state[k] == 1) return fail(c, RT_ERR_INVALID, "box %u: child %d closes a cycle", i, k);
                    if (state[k] == 0) work.push_back(uint32_t(k));
                }
            } else {
                int32_t h = 0;
                for (int32_t k : kids)
                    if (k >= 0) {
                        if (state[k] != 2) return fail(c, RT_ERR_INVALID, "box %u: child %d closes a cycle", i, k);
                        h = std::max(h, height[k]);
                    }
                height[i] = h + 1; state[i] = 2; work.pop_back();
            }
        }
        if (height[0] > RT_MAX_STACK)
            return fail(c, RT_ERR_INVALID, "tree height %d exceeds the traversal stack (%d)", height[0], RT_MAX_STACK);
        sc.stackDepth = std::max(1, height[0]);

        // 2. numbering: internal nodes depth-first (below), leaves in order of first reference; the loop further down
        //    numbers whatever is left breadth-first (everything, with RT_NODE_ORDER=bfs)
        std::vector<int32_t> ref(nBoxes, RT_REF_NONE);
        std::vector<uint32_t> order;
        uint32_t nSlots = 0;
        auto leafRef = [&](uint32_t i) -> int32_t {
            const HostBox& b = boxes[i];
            const uint32_t first = nSlots;
            triIds.insert(triIds.end(), indices + b.triangleOffset, indices + b.triangleOffset + b.numTriangles);
            nSlots += b.numTriangles;
            uint32_t code;
            if (b.numTriangles <= RT_LEAF_MAX) {
                code = (first << RT_LEAF_BITS) | b.numTriangles;
            } else {
                code = (uint32_t(bigLeaves.size()) << RT_LEAF_BITS) | RT_LEAF_BIG;
                bigLeaves.push_back(make_uint2(first, b.numTriangles));
            }
            return int32_t(~code);
        };
        auto assign = [&](int32_t child) -> int32_t {
            if (child < 0) return RT_REF_NONE;
            if (ref[child] != RT_REF_NONE) return ref[child];
            if (boxes[child].childA == -1) {
                ref[child] = leafRef(uint32_t(child));
            } else {
                ref[child] = int32_t(order.size());
                order.push_back(uint32_t(child));
            }
            return ref[child];
        };
        sc.rootRef = assign(0);
        // Depth-first (pre-order, child A's subtree before child B's): a subtree's nodes, and the triangles of its leaves,
        // are contiguous in memory, so the rays an SM works on (neighbouring pixels: neighbouring parts of the scene) keep
        // hitting the same lines.  Measured against breadth-first numbering (RT_NODE_ORDER=bfs keeps it for comparison):
        // dragon 4K 64 spp 26.5 -> 26.2 ms, closed room 8 spp 155.3 -> 154.2, cfg3_open 10.83 -> 10.54
        // (profiles/r02_variants.md).  Only the numbering changes: child order inside a node, triangle order inside a leaf
        // and therefore the traversal are the reference's.
        const char* orderMode = getenv("RT_NODE_ORDER");
        if (!(orderMode && strncmp(orderMode, "bfs", 3) == 0) && sc.rootRef >= 0) {
            std::function<void(uint32_t)> visit = [&](uint32_t i) {        // depth <= tree height <= RT_MAX_STACK
                const int32_t kids[2] = {boxes[i].childA, boxes[i].childB};
                for (int32_t child : kids) {
                    if (child < 0 || ref[child] != RT_REF_NONE) continue;
                    if (boxes[child].childA == -1) {
                        ref[child] = leafRef(uint32_t(child));
                    } else {
                        ref[child] = int32_t(order.size());
                        order.push_back(uint32_t(child));
                        visit(uint32_t(child));
                    }
                }
            };
            visit(order[0]);
        }
        for (int k = 0; k < 3; k++) { sc.rootMin[k] = boxes[0].posMin[k]; sc.rootMax[k] = boxes[0].posMax[k]; }
        nodes.reserve(nBoxes);
        triIds.reserve(nIndices);
        for (size_t at = 0; at < order.size(); at++) {
            const HostBox& b = boxes[order[at]];
            NodeRec n{};
            n.refA = assign(b.childA);
            n.refB = assign(b.childB);
            const HostBox& a = boxes[b.childA];
            n.aMinX = a.posMin[0]; n.aMinY = a.posMin[1]; n.aMinZ = a.posMin[2];
            n.aMaxX = a.posMax[0]; n.aMaxY = a.posMax[1]; n.aMaxZ = a.posMax[2];
            if (b.childB >= 0) {
                const HostBox& bb = boxes[b.childB];
                n.bMinX = bb.posMin[0]; n.bMinY = bb.posMin[1]; n.bMinZ = bb.posMin[2];
                n.bMaxX = bb.posMax[0]; n.bMaxY = bb.posMax[1]; n.bMaxZ = bb.posMax[2];
            } else {
                // absent child (fragment.glsl:287 leaves tB = -1): NaN bounds can never pass the slab test
                const float nan = std::numeric_limits<float>::quiet_NaN();
                n.bMinX = n.bMinY = n.bMinZ = n.bMaxX = n.bMaxY = n.bMaxZ = nan;
            }
            nodes.push_back(n);
            if (triIds.size() >= RT_MAX_SLOTS)
                return fail(c, RT_ERR_INVALID, "scene needs more than %u triangle slots", RT_MAX_SLOTS);
        }
    }

    // device side: the set of buffers that is not in use is filled on the upload stream, so passes still in flight on the
    // main stream (they read the other set) are not waited for.  Buffers are reused while they are large enough (a
    // re-upload then costs no allocation); the three small tables go through one pinned staging block, the caller's
    // tris[] is copied as it is and the 48 B + 80 B records of every slot are built from it by a kernel (same operations,
    // same order, no contraction: fragment.glsl:112-116).
    rt_ctx::SceneSet& S = c->sets[c->curSet ^ 1];
    RT_CUDA(c, cudaEventSynchronize(S.lastUse));      // passes that read this set two uploads ago
    auto grow = [&](auto** ptr, size_t& cap, size_t need) -> cudaError_t {
        using T = typename std::remove_pointer<typename std::remove_pointer<decltype(ptr)>::type>::type;
        if (need <= cap && *ptr) return cudaSuccess;
        cudaFree(*ptr); *ptr = nullptr; cap = 0;
        const size_t n = std::max<size_t>(need + need / 8, 1);
        cudaError_t e = cudaMalloc(reinterpret_cast<void**>(ptr), n * sizeof(T));
        if (e == cudaSuccess) cap = n;
        return e;
    };
    const size_t nSlotsAll = triIds.size();
    RT_CUDA(c, grow(&S.dNodes, S.capNodes, nodes.size()));
    RT_CUDA(c, grow(&S.dTriId, S.capSlots, nSlotsAll));
    if (!S.dTris || !S.dShade || S.capSlots != S.capSlotRecs) {
        cudaFree(S.dTris); cudaFree(S.dShade); S.dTris = nullptr; S.dShade = nullptr; S.capSlotRecs = 0;
        RT_CUDA(c, cudaMalloc(reinterpret_cast<void**>(&S.dTris), S.capSlots * sizeof(TriRec)));
        RT_CUDA(c, cudaMalloc(reinterpret_cast<void**>(&S.dShade), S.capSlots * sizeof(ShadeRec)));
        S.capSlotRecs = S.capSlots;
    }
    RT_CUDA(c, grow(&S.dBig, S.capBig, bigLeaves.size()));
    RT_CUDA(c, grow(&S.dRawTris, S.capRaw, size_t(nTris) * (sizeof(HostTri) / sizeof(float4))));
    const size_t bytesNodes = nodes.size() * sizeof(NodeRec), bytesIds = nSlotsAll * sizeof(uint32_t),
                 bytesBig = bigLeaves.size() * sizeof(uint2);
    const size_t offIds = bytesNodes, offBig = (offIds + bytesIds + 15) & ~size_t(15), stageBytes = offBig + bytesBig;
    if (stageBytes > S.hStageBytes) {
        cudaFreeHost(S.hStage); S.hStage = nullptr; S.hStageBytes = 0;
        RT_CUDA(c, cudaMallocHost(&S.hStage, stageBytes + stageBytes / 8));
        S.hStageBytes = stageBytes + stageBytes / 8;
    }
    char* stage = static_cast<char*>(S.hStage);
    if (bytesNodes) std::memcpy(stage, nodes.data(), bytesNodes);
    if (bytesIds) std::memcpy(stage + offIds, triIds.data(), bytesIds);
    if (bytesBig) std::memcpy(stage + offBig, bigLeaves.data(), bytesBig);
    cudaStream_t up = c->upStream;
    if (nTris) RT_CUDA(c, cudaMemcpyAsync(S.dRawTris, tris, size_t(nTris) * sizeof(HostTri), cudaMemcpyHostToDevice, up));
    if (bytesNodes) RT_CUDA(c, cudaMemcpyAsync(S.dNodes, stage, bytesNodes, cudaMemcpyHostToDevice, up));
    if (bytesIds) RT_CUDA(c, cudaMemcpyAsync(S.dTriId, stage + offIds, bytesIds, cudaMemcpyHostToDevice, up));
    if (bytesBig) RT_CUDA(c, cudaMemcpyAsync(S.dBig, stage + offBig, bytesBig, cudaMemcpyHostToDevice, up));
    if (nSlotsAll)
        RT_CUDA(c, rtLaunchBuildRecords(S.dRawTris, S.dTriId, uint32_t(nSlotsAll), S.dTris, S.dShade, up));
    // glBufferData semantics: the caller's arrays are free at return; and every later launch on the main stream is
    // issued after this point, so it sees the finished set
    RT_CUDA(c, cudaStreamSynchronize(up));
    c->curSet ^= 1;
    sc.nodes = S.dNodes; sc.tris = S.dTris; sc.shade = S.dShade; sc.triId = S.dTriId; sc.bigLeaves = S.dBig;
    sc.nNodes = uint32_t(nodes.size());
    sc.nSlots = uint32_t(nSlotsAll);
    c->scene = sc;
    c->haveScene = true;
    return RT_OK;
}

int rt_set_params(rt_ctx* c, const rt_params* p) {
    if (!c || !p) return RT_ERR_INVALID;
    if (p->tileSizeX <= 0 || p->tileSizeY <= 0 || p->numTilesX <= 0 || p->numTilesY <= 0)
        return fail(c, RT_ERR_INVALID, "tile sizes and counts must be positive (the shader divides by them)");
    if (p->nBounces < 0 || p->rays_per_pixel < 0) return fail(c, RT_ERR_INVALID, "negative nBounces / rays_per_pixel");
    c->params = *p;
    c->haveParams = true;
    return RT_OK;
}

int rt_set_variant(rt_ctx* c, int variant) {
    if (!c) return RT_ERR_INVALID;
    if (variant < RT_VARIANT_AUTO || variant > RT_VARIANT_CHAIN) return fail(c, RT_ERR_INVALID, "unknown variant %d", variant);
    c->variant = variant;
    return RT_OK;
}

int rt_target_free(rt_ctx* c) {
    if (!c) return RT_ERR_INVALID;
    Bind bind(c);
    cudaStreamSynchronize(c->stream);
    cudaStreamSynchronize(c->downStream);
    c->downPending[0] = c->downPending[1] = false;
    freeTargets(c);
    return RT_OK;
}

int rt_target_clear(rt_ctx* c) {
    if (!c) return RT_ERR_INVALID;
    if (!c->target[0]) return fail(c, RT_ERR_STATE, "no target");
    Bind bind(c);
    waitDownloads(c);
    const size_t n = size_t(c->width) * c->height;
    RT_CUDA(c, cudaMemsetAsync(c->target[0], 0, n * sizeof(float4), c->stream));
    RT_CUDA(c, cudaMemsetAsync(c->target[1], 0, n * sizeof(float4), c->stream));
    RT_CUDA(c, cudaMemsetAsync(c->dIds, 0xff, n * sizeof(int32_t), c->stream));
    c->accumIndex = 0;   // main.py:262
    return RT_OK;
}

int rt_target_create(rt_ctx* c, int width, int height) {
    if (!c) return RT_ERR_INVALID;
    if (width <= 0 || height <= 0) return fail(c, RT_ERR_INVALID, "target size %d x %d", width, height);
    Bind bind(c);
    cudaStreamSynchronize(c->stream);
    cudaStreamSynchronize(c->downStream);
    c->downPending[0] = c->downPending[1] = false;
    freeTargets(c);
    const size_t n = size_t(width) * height;
    if (cudaMalloc(&c->target[0], n * sizeof(float4)) != cudaSuccess || cudaMalloc(&c->target[1], n * sizeof(float4)) != cudaSuccess ||
        cudaMalloc(&c->dIds, n * sizeof(int32_t)) != cudaSuccess || cudaMalloc(&c->dRgba8, n * sizeof(uchar4)) != cudaSuccess) {
        freeTargets(c);
        cudaGetLastError();
        return fail(c, RT_ERR_NOMEM, "cannot allocate %d x %d targets", width, height);
    }
    c->width = width; c->height = height;
    return rt_target_clear(c);
}

}  // extern "C"

namespace {

// One draw (main.py:376-394, 401): read target[accumIndex], write the other, flip.
// sumMode: 0 = the shader's running mean; 1 = next = curr; 2 = next = prev + curr.
int renderPass(rt_ctx* c, int frameNumber, int tileX, int tileY, int flags, int sumMode) {
    if (!c->haveScene) return fail(c, RT_ERR_STATE, "rt_render_pass before rt_scene_upload");
    if (!c->haveParams) return fail(c, RT_ERR_STATE, "rt_render_pass before rt_set_params");
    if (!c->target[0]) return fail(c, RT_ERR_STATE, "rt_render_pass before rt_target_create");
    waitDownloads(c, 1 - c->accumIndex);                 // the target this pass writes
    const rt_params& u = c->params;
    PassParams P{};
    P.scene = c->scene;
    P.tileX = tileX; P.tileY = tileY; P.tileSizeX = u.tileSizeX; P.tileSizeY = u.tileSizeY;
    P.numTilesX = u.numTilesX; P.numTilesY = u.numTilesY;
    P.lambertian = u.lambertian; P.nBounces = u.nBounces; P.raysPerPixel = u.rays_per_pixel; P.frameNumber = frameNumber;
    P.skyBrightness = u.skyBrightness; P.xStep = u.xStep; P.yStep = u.yStep; P.dirStartX = u.dirStartX;
    P.dirStartY = u.dirStartY; P.jitterAmount = u.jitterAmount;
    for (int k = 0; k < 3; k++) {
        P.camPos[k] = u.camPos[k]; P.camRight[k] = u.camRight[k]; P.camUp[k] = u.camUp[k]; P.camForward[k] = u.camForward[k];
    }
    P.width = c->width; P.height = c->height;
    const int prev = c->accumIndex, next = 1 - prev;
    P.prev = c->target[prev]; P.next = c->target[next];
    if (flags & RT_PASS_TO_PEER) {
        if (!c->peerTarget[0]) return fail(c, RT_ERR_STATE, "RT_PASS_TO_PEER without rt_peer_open");
        if (sumMode != 0) return fail(c, RT_ERR_INVALID, "RT_PASS_TO_PEER is for tile mode, not for frame sums");
        P.peerNext = c->peerTarget[next];
    }
    P.primaryIds = (flags & RT_PASS_PRIMARY_IDS) ? c->dIds : nullptr;
    P.counters = (flags & RT_PASS_COUNTERS) ? c->dCounters : nullptr;
    P.ownedW = ownedCount(c->width, u.tileSizeX, u.numTilesX, tileX);
    P.ownedH = ownedCount(c->height, u.tileSizeY, u.numTilesY, tileY);
    P.sumMode = sumMode;
    P.workCounter = c->dWork;
    P.scene.deferEdges = deferEdgesFor((long long)P.ownedW * P.ownedH, u.rays_per_pixel);
    const bool covers = P.ownedW == c->width && P.ownedH == c->height;
    if (!covers && !(flags & RT_PASS_SKIP_MASKED)) {
        // pixels outside the tile keep the previous target's texel, all four channels (fragment.glsl:382-386)
        RT_CUDA(c, cudaMemcpyAsync(c->target[next], c->target[prev], size_t(c->width) * c->height * sizeof(float4),
                                   cudaMemcpyDeviceToDevice, c->stream));
    }
    cudaError_t e;
    int variant = c->variant;
    // AUTO, by measurement on B200 (profiles/r01_variants.md, r02_variants.md).  Fewer than 8 samples per pixel: the
    // one-thread-per-pixel kernel.  More make the per-pixel work uneven (sky pixels finish at once, surface pixels run
    // rays_per_pixel paths): the persistent warp-vote kernel.  From 16 samples per pixel the serial sample chain of a pixel
    // (the RNG state is carried from sample to sample) starts to bound the pass — its tail on a whole frame, all of it on a
    // tile-sharded frame's 1/N share — and where the sky is visible the two variants that evaluate samples ahead of the
    // chain win: CHAIN (a pixel per lane, idle lanes help their neighbours' chains; no table, little waste) for a large
    // share, SAMPLES (every offset of every hit pixel through a table, in rounds) for a small one (dragon 4K 64 spp, share
    // 1 / 1/2 / 1/4 / 1/8: persistent 27.2 / 16.5 / 14.0 / 12.3 ms, chain 26.2 / 16.2 / 13.0 / 10.6, samples 30.3 / 16.4 /
    // 8.9 / 4.6).  In a closed room, or where most pixels hit, both lose to the persistent kernel (box 4K 8 spp: chain 162
    // vs 155 ms).  Which case it is only the primary rays tell, so the pass is launched with BOTH branches and stage 1's
    // count of hit pixels decides on the device (specGate).
    bool gated = false;
    const bool canSpeculate = P.counters == nullptr && c->width <= 65535 && c->height <= 65535;   // no counting instantiation; px | py << 16
    if (variant == RT_VARIANT_AUTO) {
        const long long owned = (long long)P.ownedW * P.ownedH, manyPixels = (long long)c->smCount * 2048 * 16;
        variant = u.rays_per_pixel < 8 ? RT_VARIANT_PIXEL : RT_VARIANT_PERSISTENT;
        if (u.rays_per_pixel >= 16 && canSpeculate && sumMode == 0) {
            variant = owned <= manyPixels ? RT_VARIANT_SAMPLES : RT_VARIANT_CHAIN;
            gated = true;
        }
    }
    // the speculative variants evaluate paths the reference would not: they have no counting instantiation
    if ((variant == RT_VARIANT_SAMPLES || variant == RT_VARIANT_CHAIN) && !canSpeculate) variant = RT_VARIANT_PERSISTENT;
    if (variant == RT_VARIANT_WAVEFRONT && (c->width > 65535 || c->height > 65535)) variant = RT_VARIANT_PIXEL;
    if (variant == RT_VARIANT_CHAIN && (u.nBounces > 63 || (long long)u.rays_per_pixel * std::max(u.nBounces, 1) >= (1ll << 25)))
        variant = RT_VARIANT_PERSISTENT;       // the chain variant's ring tags hold hits < 64 and offsets < 2^26
    if (variant == RT_VARIANT_CHAIN) {
        e = ensureRecordPlanes(c);
        if (e != cudaSuccess) return fail(c, RT_ERR_NOMEM, "chain queue: %s", cudaGetErrorString(e));
        P.queueA = c->dQueueA; P.queueB = c->dQueueB; P.queueC = c->dQueueC; P.queueD = c->dQueueD; P.queueE = c->dQueueE;
        P.queueCount = c->dQueueCount; P.workCounters = c->dWorkArr;
        if (!c->dChainRing && cudaMalloc(&c->dChainRing, rtChainRingBytes(c->smCount)) != cudaSuccess) {
            cudaGetLastError();
            return fail(c, RT_ERR_NOMEM, "chain rings: %zu bytes", rtChainRingBytes(c->smCount));
        }
        P.chainRing = c->dChainRing;
        P.specCapacity = 0xffffffffu;          // no table to fit: the gate is the hit fraction alone
        P.specGate = gated ? 1 : 0;            // AUTO: the persistent kernel is the other branch; forced: the chain kernel alone
        e = rtLaunchChain(P, P.primaryIds != nullptr, c->smCount, c->stream);
        if (e == cudaSuccess && gated) e = rtLaunchPersistent(P, false, P.primaryIds != nullptr, c->smCount, c->stream);
    } else if (variant == RT_VARIANT_SAMPLES) {
        const int K = rtSpecTableK(u.rays_per_pixel, u.nBounces);
        uint32_t capacity = 0;
        e = ensureSpec(c, size_t(P.ownedW) * size_t(P.ownedH), K, &capacity);
        if (e != cudaSuccess) return fail(c, RT_ERR_NOMEM, "samples queue: %s", cudaGetErrorString(e));
        P.queueA = c->dQueueA; P.queueB = c->dQueueB; P.queueC = c->dQueueC; P.queueD = c->dQueueD; P.queueE = c->dQueueE;
        P.queueCount = c->dQueueCount; P.specBuf = c->dSpecBuf; P.specK = K; P.specCapacity = capacity;
        P.specBufEntries = uint32_t(std::min<size_t>(c->specBufEntries, 0xffffffffu));
        P.workCounters = c->dWorkArr; P.todoCount = c->dWorkArr + 32; P.classCount = c->dWorkArr + 64;
        P.chainK = c->dChainK; P.chainColor = c->dChainColor; P.chainClass = c->dChainClass;
        P.specGate = gated ? 1 : 0;
        e = rtLaunchSamples(P, P.primaryIds != nullptr, c->smCount, c->dTodo[0], c->dTodo[1], c->stream);
        if (e == cudaSuccess && gated) e = rtLaunchPersistent(P, false, P.primaryIds != nullptr, c->smCount, c->stream);
    } else if (variant == RT_VARIANT_WAVEFRONT) {
        e = ensureQueue(c, P.counters != nullptr);
        if (e != cudaSuccess) return fail(c, RT_ERR_NOMEM, "wavefront queue: %s", cudaGetErrorString(e));
        P.queueA = c->dQueueA; P.queueB = c->dQueueB; P.queuePix = c->dQueuePix; P.queueWork = c->dQueueWork;
        P.queueCount = c->dQueueCount;
        e = rtLaunchWavefront(P, P.counters != nullptr, P.primaryIds != nullptr, c->smCount, c->stream);
    } else if (variant == RT_VARIANT_POOL)
        e = rtLaunchPool(P, P.counters != nullptr, P.primaryIds != nullptr, c->smCount, c->stream);
    else if (variant == RT_VARIANT_PERSISTENT)
        e = rtLaunchPersistent(P, P.counters != nullptr, P.primaryIds != nullptr, c->smCount, c->stream);
    else
        e = rtLaunchPixel(P, P.counters != nullptr, P.primaryIds != nullptr, c->stream);
    if (e != cudaSuccess) return fail(c, RT_ERR_CUDA, "kernel launch: %s", cudaGetErrorString(e));
    c->lastLaunches = variant == RT_VARIANT_SAMPLES ? rtSpecLaunches() + (gated ? 1 : 0) : variant == RT_VARIANT_CHAIN ? (gated ? 3 : 2)
                      : variant == RT_VARIANT_WAVEFRONT ? 2 : 1;
    c->lastVariant = variant; c->lastGated = gated;
    c->lastOwned = (long long)P.ownedW * P.ownedH; c->lastCapacity = P.specCapacity;
    cudaEventRecord(c->sets[c->curSet].lastUse, c->stream);     // the next upload into this set waits for it
    c->accumIndex = next;
    return RT_OK;
}

}  // namespace

extern "C" {

int rt_render_pass_ex(rt_ctx* c, int frameNumber, int tileX, int tileY, int flags) {
    if (!c) return RT_ERR_INVALID;
    Bind bind(c);
    return renderPass(c, frameNumber, tileX, tileY, flags, 0);
}

int rt_render_pass(rt_ctx* c, int frameNumber, int tileX, int tileY) {
    return rt_render_pass_ex(c, frameNumber, tileX, tileY, RT_PASS_DEFAULT);
}

// main.py:409-430: tileX fastest, then tileY, then the frame counter.
int rt_render_frames(rt_ctx* c, int firstFrame, int nFrames) {
    if (!c) return RT_ERR_INVALID;
    if (!c->haveParams) return fail(c, RT_ERR_STATE, "rt_render_frames before rt_set_params");
    Bind bind(c);
    for (int f = firstFrame; f < firstFrame + nFrames; f++)
        for (int ty = 0; ty < c->params.numTilesY; ty++)
            for (int tx = 0; tx < c->params.numTilesX; tx++) {
                const int rc = renderPass(c, f, tx, ty, RT_PASS_DEFAULT, 0);
                if (rc != RT_OK) return rc;
            }
    return RT_OK;
}

// Frame-sharded progressive mode: this context renders frames firstFrame,
// firstFrame + stride, ... of tile (params.tileX, params.tileY) and keeps their SUM; the owner of the
// image adds the shares (NCCL reduce) and calls rt_finish_sum.  Masked pixels are left untouched.
int rt_render_frames_sum_ex(rt_ctx* c, int firstFrame, int nFrames, int stride, int flags) {
    if (!c) return RT_ERR_INVALID;
    if (!c->haveParams) return fail(c, RT_ERR_STATE, "rt_render_frames_sum before rt_set_params");
    if (!c->target[0]) return fail(c, RT_ERR_STATE, "rt_render_frames_sum before rt_target_create");
    if (stride <= 0) return fail(c, RT_ERR_INVALID, "frame stride must be positive");
    if (flags & ~RT_SUM_CONTINUE) return fail(c, RT_ERR_INVALID, "unknown rt_render_frames_sum_ex flags %d", flags);
    const rt_params& u = c->params;
    if (ownedCount(c->width, u.tileSizeX, u.numTilesX, u.tileX) != c->width ||
        ownedCount(c->height, u.tileSizeY, u.numTilesY, u.tileY) != c->height)
        return fail(c, RT_ERR_INVALID, "sum mode needs a tile that covers the target (numTilesX = numTilesY = 1): masked "
                                       "pixels would carry the other target's stale value into the reduce");
    Bind bind(c);
    for (int k = 0; k < nFrames; k++) {
        const int mode = (k == 0 && !(flags & RT_SUM_CONTINUE)) ? 1 : 2;
        const int rc = renderPass(c, firstFrame + k * stride, u.tileX, u.tileY, RT_PASS_SKIP_MASKED, mode);
        if (rc != RT_OK) return rc;
    }
    return RT_OK;
}

int rt_render_frames_sum(rt_ctx* c, int firstFrame, int nFrames, int stride) {
    return rt_render_frames_sum_ex(c, firstFrame, nFrames, stride, RT_SUM_DEFAULT);
}

int rt_sum_resume(rt_ctx* c, int priorFrames) {
    if (!c) return RT_ERR_INVALID;
    if (!c->target[0]) return fail(c, RT_ERR_STATE, "no target");
    if (priorFrames < 0) return fail(c, RT_ERR_INVALID, "prior_frames must not be negative");
    Bind bind(c);
    waitDownloads(c, c->accumIndex);
    const size_t n = size_t(c->width) * c->height;
    if (priorFrames == 0) RT_CUDA(c, cudaMemsetAsync(c->target[c->accumIndex], 0, n * sizeof(float4), c->stream));
    else RT_CUDA(c, rtLaunchScaleSum(c->target[c->accumIndex], n, float(priorFrames), c->stream));
    return RT_OK;
}

int rt_finish_sum(rt_ctx* c, int totalFrames) {
    if (!c) return RT_ERR_INVALID;
    if (!c->target[0]) return fail(c, RT_ERR_STATE, "no target");
    if (totalFrames <= 0) return fail(c, RT_ERR_INVALID, "total_frames must be positive");
    Bind bind(c);
    waitDownloads(c, c->accumIndex);
    RT_CUDA(c, rtLaunchFinishSum(c->target[c->accumIndex], size_t(c->width) * c->height, float(totalFrames), c->stream));
    return RT_OK;
}

// The read-back runs on its own stream behind everything the main stream has been given so far, so the next passes
// overlap it (the GL counterpart is glReadPixels into a pixel-buffer object); a pass that is about to overwrite the
// target being copied waits for the copy on the device (waitDownloads), never on the host.
int rt_read_accum_f32_async(rt_ctx* c, float* dst) {
    if (!c || !dst) return RT_ERR_INVALID;
    if (!c->target[0]) return fail(c, RT_ERR_STATE, "no target");
    Bind bind(c);
    const int idx = c->accumIndex;
    RT_CUDA(c, cudaEventRecord(c->evRendered, c->stream));
    RT_CUDA(c, cudaStreamWaitEvent(c->downStream, c->evRendered, 0));
    RT_CUDA(c, cudaMemcpyAsync(dst, c->target[idx], size_t(c->width) * c->height * sizeof(float4), cudaMemcpyDeviceToHost,
                               c->downStream));
    RT_CUDA(c, cudaEventRecord(c->evDown[idx], c->downStream));
    c->downPending[idx] = true;
    return RT_OK;
}

int rt_read_wait(rt_ctx* c) {
    if (!c) return RT_ERR_INVALID;
    Bind bind(c);
    RT_CUDA(c, cudaStreamSynchronize(c->downStream));
    return RT_OK;
}

int rt_read_accum_f32(rt_ctx* c, float* dst) {
    const int rc = rt_read_accum_f32_async(c, dst);
    return rc != RT_OK ? rc : rt_read_wait(c);
}

// Upload an accumulation image into the current target (checkpoint restore; the GL counterpart would be
// glTexSubImage2D on accum_tex[accum_index], which the reference never does).
int rt_last_pass_launches(rt_ctx* c) { return c ? c->lastLaunches : 0; }

// Which variant rendered the last pass.  Where AUTO launched both branches this reads stage 1's count back (it waits for
// the pass) and repeats the device's decision (vote::samplesSpeculate).
int rt_last_pass_variant(rt_ctx* c) {
    if (!c) return RT_ERR_INVALID;
    if (!c->lastGated) return c->lastVariant;
    Bind bind(c);
    unsigned int queued = 0;
    RT_CUDA(c, cudaMemcpyAsync(&queued, c->dQueueCount, sizeof queued, cudaMemcpyDeviceToHost, c->stream));
    RT_CUDA(c, cudaStreamSynchronize(c->stream));
    int forced = 0, t0 = 0, t1 = 0;
    if (const char* v = getenv("RT_VOTE_TUNE")) sscanf(v, "%d,%d,%d", &t0, &t1, &forced);
    const bool speculate = queued <= c->lastCapacity && (forced == 1 || (forced != 2 && 4ll * queued <= c->lastOwned));
    return speculate ? c->lastVariant : RT_VARIANT_PERSISTENT;
}

int rt_measure_peaks(rt_ctx* c, double* out3) {
    if (!c || !out3) return RT_ERR_INVALID;
    Bind bind(c);
    out3[0] = out3[1] = out3[2] = 0.0;
    RT_CUDA(c, rtMeasurePeaks(c->smCount, c->stream, out3));
    return RT_OK;
}

int rt_write_accum_f32(rt_ctx* c, const float* src) {
    if (!c || !src) return RT_ERR_INVALID;
    if (!c->target[0]) return fail(c, RT_ERR_STATE, "no target");
    Bind bind(c);
    waitDownloads(c, c->accumIndex);
    RT_CUDA(c, cudaMemcpyAsync(c->target[c->accumIndex], src, size_t(c->width) * c->height * sizeof(float4),
                               cudaMemcpyHostToDevice, c->stream));
    RT_CUDA(c, cudaStreamSynchronize(c->stream));
    return RT_OK;
}

int rt_read_rgba8(rt_ctx* c, uint8_t* dst) {
    if (!c || !dst) return RT_ERR_INVALID;
    if (!c->target[0]) return fail(c, RT_ERR_STATE, "no target");
    Bind bind(c);
    const size_t n = size_t(c->width) * c->height;
    RT_CUDA(c, rtLaunchToRgba8(c->target[c->accumIndex], c->dRgba8, n, c->stream));
    RT_CUDA(c, cudaMemcpyAsync(dst, c->dRgba8, n * sizeof(uchar4), cudaMemcpyDeviceToHost, c->stream));
    RT_CUDA(c, cudaStreamSynchronize(c->stream));
    return RT_OK;
}

int rt_read_primary_ids(rt_ctx* c, int32_t* dst) {
    if (!c || !dst) return RT_ERR_INVALID;
    if (!c->dIds) return fail(c, RT_ERR_STATE, "no target");
    Bind bind(c);
    RT_CUDA(c, cudaMemcpyAsync(dst, c->dIds, size_t(c->width) * c->height * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
    RT_CUDA(c, cudaStreamSynchronize(c->stream));
    return RT_OK;
}

int rt_read_counters(rt_ctx* c, uint64_t* dst, int n) {
    if (!c || !dst || n < 0) return RT_ERR_INVALID;
    Bind bind(c);
    unsigned long long host[RT_CNT_COUNT];
    RT_CUDA(c, cudaMemcpyAsync(host, c->dCounters, sizeof host, cudaMemcpyDeviceToHost, c->stream));
    RT_CUDA(c, cudaStreamSynchronize(c->stream));
    for (int k = 0; k < n; k++) dst[k] = k < RT_CNT_COUNT ? host[k] : 0;
    return RT_OK;
}

int rt_reset_counters(rt_ctx* c) {
    if (!c) return RT_ERR_INVALID;
    Bind bind(c);
    RT_CUDA(c, cudaMemsetAsync(c->dCounters, 0, RT_CNT_COUNT * sizeof(unsigned long long), c->stream));
    return RT_OK;
}

void* rt_accum_device_ptr(rt_ctx* c, int which) {
    if (!c || !c->target[0]) return nullptr;
    if (which < 0) which = c->accumIndex;
    return which <= 1 ? c->target[which] : nullptr;
}

int rt_accum_index(rt_ctx* c) { return c ? c->accumIndex : RT_ERR_INVALID; }

int rt_set_accum_index(rt_ctx* c, int index) {
    if (!c || (index != 0 && index != 1)) return RT_ERR_INVALID;
    c->accumIndex = index;
    return RT_OK;
}

int rt_band_rows(rt_ctx* c, int tileY) {
    if (!c) return RT_ERR_INVALID;
    if (!c->haveParams || !c->target[0]) return fail(c, RT_ERR_STATE, "rt_band_rows needs params and a target");
    return ownedCount(c->height, c->params.tileSizeY, c->params.numTilesY, tileY);
}

static int bandCopy(rt_ctx* c, int tileY, void* dense, int toDense) {
    if (!c || !dense) return RT_ERR_INVALID;
    if (!c->haveParams || !c->target[0]) return fail(c, RT_ERR_STATE, "band copy needs params and a target");
    Bind bind(c);
    if (!toDense) waitDownloads(c, c->accumIndex);
    const int rows = ownedCount(c->height, c->params.tileSizeY, c->params.numTilesY, tileY);
    RT_CUDA(c, rtLaunchBandCopy(c->target[c->accumIndex], static_cast<float4*>(dense), c->width, rows, c->params.tileSizeY,
                                c->params.numTilesY, tileY, toDense, c->stream));
    return RT_OK;
}

int rt_band_gather(rt_ctx* c, int tileY, void* dense) { return bandCopy(c, tileY, dense, 1); }
int rt_band_scatter(rt_ctx* c, int tileY, const void* dense) { return bandCopy(c, tileY, const_cast<void*>(dense), 0); }

// One draw into caller-provided buffers with no allocation anywhere on the path (include/rt_b200.h).
int rt_render_pass_static(rt_ctx* c, int frameNumber, int tileX, int tileY, int width, int height, const void* prevDev,
                          void* nextDev) {
    if (!c) return RT_ERR_INVALID;
    if (!c->haveScene) return fail(c, RT_ERR_STATE, "rt_render_pass_static before rt_scene_upload");
    if (!c->haveParams) return fail(c, RT_ERR_STATE, "rt_render_pass_static before rt_set_params");
    if (width <= 0 || height <= 0 || width > 65535 || height > 65535)
        return fail(c, RT_ERR_INVALID, "rt_render_pass_static: frame %d x %d (1 .. 65535 each way)", width, height);
    if (!nextDev) return fail(c, RT_ERR_INVALID, "rt_render_pass_static: no output buffer");
    const rt_params& u = c->params;
    PassParams P{};
    P.scene = c->scene;
    P.tileX = tileX; P.tileY = tileY; P.tileSizeX = u.tileSizeX; P.tileSizeY = u.tileSizeY;
    P.numTilesX = u.numTilesX; P.numTilesY = u.numTilesY;
    P.lambertian = u.lambertian; P.nBounces = u.nBounces; P.raysPerPixel = u.rays_per_pixel; P.frameNumber = frameNumber;
    P.skyBrightness = u.skyBrightness; P.xStep = u.xStep; P.yStep = u.yStep; P.dirStartX = u.dirStartX;
    P.dirStartY = u.dirStartY; P.jitterAmount = u.jitterAmount;
    for (int k = 0; k < 3; k++) {
        P.camPos[k] = u.camPos[k]; P.camRight[k] = u.camRight[k]; P.camUp[k] = u.camUp[k]; P.camForward[k] = u.camForward[k];
    }
    P.width = width; P.height = height;
    P.prev = static_cast<const float4*>(prevDev); P.next = static_cast<float4*>(nextDev);
    P.ownedW = ownedCount(width, u.tileSizeX, u.numTilesX, tileX);
    P.ownedH = ownedCount(height, u.tileSizeY, u.numTilesY, tileY);
    P.sumMode = 0;
    const long long owned = (long long)P.ownedW * P.ownedH;
    P.scene.deferEdges = deferEdgesFor(owned, u.rays_per_pixel);
    if (owned > (long long)RT_STATIC_MAX_PIXELS)
        return fail(c, RT_ERR_INVALID, "rt_render_pass_static: the tile owns %lld pixels, the static queue holds %d "
                    "(RT_STATIC_MAX_PIXELS); draw smaller tiles or use rt_render_pass", owned, int(RT_STATIC_MAX_PIXELS));
    const bool covers = P.ownedW == width && P.ownedH == height;
    if (!prevDev && (frameNumber > 0 || !covers))
        return fail(c, RT_ERR_INVALID, "rt_render_pass_static: the previous frame is read (frameNumber > 0 or a partial tile) but prev_dev is NULL");
    Bind bind(c);
    if (!covers)    // pixels outside the tile keep the previous target's texel, all four channels (fragment.glsl:382-386)
        RT_CUDA(c, cudaMemcpyAsync(nextDev, prevDev, size_t(width) * height * sizeof(float4), cudaMemcpyDeviceToDevice, c->stream));
    const cudaError_t e = rtLaunchWavefrontStatic(P, c->device, c->smCount, c->stream);
    if (e != cudaSuccess) return fail(c, RT_ERR_CUDA, "rt_render_pass_static: %s", cudaGetErrorString(e));
    cudaEventRecord(c->sets[c->curSet].lastUse, c->stream);     // the next upload into this set waits for it
    return RT_OK;
}

int rt_peer_export(rt_ctx* c, void* handlesOut) {
    if (!c || !handlesOut) return RT_ERR_INVALID;
    if (!c->target[0]) return fail(c, RT_ERR_STATE, "rt_peer_export before rt_target_create");
    Bind bind(c);
    static_assert(sizeof(cudaIpcMemHandle_t) == RT_IPC_HANDLE_BYTES, "handle size");
    for (int i = 0; i < 2; i++) {
        cudaIpcMemHandle_t h;
        RT_CUDA(c, cudaIpcGetMemHandle(&h, c->target[i]));
        memcpy(static_cast<char*>(handlesOut) + i * RT_IPC_HANDLE_BYTES, &h, sizeof h);
    }
    return RT_OK;
}

int rt_peer_close(rt_ctx* c) {
    if (!c) return RT_ERR_INVALID;
    Bind bind(c);
    if (c->peerTarget[0] || c->peerTarget[1]) cudaStreamSynchronize(c->stream);
    for (int i = 0; i < 2; i++) {
        if (c->peerTarget[i]) cudaIpcCloseMemHandle(c->peerTarget[i]);
        c->peerTarget[i] = nullptr;
    }
    return RT_OK;
}

int rt_peer_open(rt_ctx* c, const void* handles) {
    if (!c || !handles) return RT_ERR_INVALID;
    if (!c->target[0]) return fail(c, RT_ERR_STATE, "rt_peer_open before rt_target_create");
    rt_peer_close(c);
    Bind bind(c);
    for (int i = 0; i < 2; i++) {
        cudaIpcMemHandle_t h;
        memcpy(&h, static_cast<const char*>(handles) + i * RT_IPC_HANDLE_BYTES, sizeof h);
        void* p = nullptr;
        const cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            cudaGetLastError();
            rt_peer_close(c);
            return fail(c, RT_ERR_CUDA, "cudaIpcOpenMemHandle: %s", cudaGetErrorString(e));
        }
        c->peerTarget[i] = static_cast<float4*>(p);
    }
    return RT_OK;
}

int rt_peer_is_open(rt_ctx* c) { return c && c->peerTarget[0] ? 1 : 0; }

int rt_target_wait_reads(rt_ctx* c) {
    if (!c) return RT_ERR_INVALID;
    if (!c->target[0]) return fail(c, RT_ERR_STATE, "no target");
    Bind bind(c);
    waitDownloads(c, 1 - c->accumIndex);
    return RT_OK;
}

int rt_timer_start(rt_ctx* c) {
    if (!c) return RT_ERR_INVALID;
    Bind bind(c);
    RT_CUDA(c, cudaEventRecord(c->ev0, c->stream));
    return RT_OK;
}

int rt_timer_stop(rt_ctx* c, float* ms) {
    if (!c || !ms) return RT_ERR_INVALID;
    Bind bind(c);
    RT_CUDA(c, cudaEventRecord(c->ev1, c->stream));
    RT_CUDA(c, cudaEventSynchronize(c->ev1));
    RT_CUDA(c, cudaEventElapsedTime(ms, c->ev0, c->ev1));
    return RT_OK;
}

}  // extern "C"

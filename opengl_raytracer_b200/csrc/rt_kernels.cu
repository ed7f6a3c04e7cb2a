// rt_kernels.cu — sm_100a kernels of librt_b200: the per-pixel loop of the reference's fragment.glsl.
//
// Variant PIXEL: one thread per owned pixel, the whole pixel-frame (rays_per_pixel paths of up to
// nBounces rays, RNG state carried across the samples — fragment.glsl:352-366 makes them sequentially
// dependent) in one kernel.  A warp covers an 8x4 pixel tile so primary rays are coherent.  Traversal
// stacks live in shared memory ((ref, tNear) pairs, entry-major so lanes are conflict-free); scene
// records come through the read-only path as 16-byte loads; the scene (<= 8 MB) is L2-resident.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <mutex>

#include "../../include/rt_b200.h"
#include "rt_shade.cuh"
#include "rt_vote.cuh"
#include "rt_pool.cuh"
#include "rt_wave.cuh"
#include "rt_spec.cuh"
#include "rt_chain.cuh"

namespace {

constexpr int kBlockThreads = 128;   // 4 warps: 16 x 8 pixel block, each warp an 8 x 4 tile

struct PathCounters {
    uint32_t v[RT_CNT_COUNT];
};

// One pixel-frame: fragment.glsl main() from :388 on.  Returns the pixel's colour for this frame.
template <bool COUNT, bool IDS, bool DEFER>
RT_DEV f3 pixelFrame(const PassParams& P, int px, int py, StackRef st, PathCounters& pc, int32_t& primaryId) {
    const DeviceScene& sc = P.scene;
    uint32_t seed;
    const f3 camPos = ld3(P.camPos);
    const f3 primaryDir = primaryDirection(P, px, py, seed);
    const f3 sky = scale3(mk3(0.1f, 0.6f, 0.92f), P.skyBrightness);   // :242-244, :388
    const bool lambertian = P.lambertian != 0;

    // The primary ray is the same for every sample of the pixel-frame (:361-362): trace it once.
    WorkCount primaryWork = {0, 0, 0, 0};
    RayHit primaryHit;
    primaryHit.slot = -1; primaryHit.t = 0.f; primaryHit.u = 0.f; primaryHit.v = 0.f;
    if (P.nBounces > 0 && P.raysPerPixel > 0) {
        primaryHit = traverse<COUNT, DEFER>(sc, camPos, primaryDir, st, primaryWork);
        if (COUNT) pc.v[RT_CNT_RAYS_TRACED]++;
    }
    if (IDS) primaryId = primaryHit.slot >= 0 ? int32_t(__ldg(sc.triId + primaryHit.slot)) : -1;

    f3 color = mk3(0.f, 0.f, 0.f);
    for (int s = 0; s < P.raysPerPixel; s++) {
        // raytrace(), fragment.glsl:309-350
        f3 incoming = mk3(0.f, 0.f, 0.f), rayColor = mk3(1.f, 1.f, 1.f);
        f3 o = camPos, d = primaryDir;
        if (COUNT) pc.v[RT_CNT_PATHS]++;
        for (int i = 0; i < P.nBounces; i++) {
            RayHit h;
            if (i == 0) {
                h = primaryHit;
                if (COUNT) {
                    pc.v[RT_CNT_INTERNAL] += primaryWork.internal; pc.v[RT_CNT_LEAVES] += primaryWork.leaves;
                    pc.v[RT_CNT_TRI_TESTS] += primaryWork.tris;    pc.v[RT_CNT_POPS] += primaryWork.pops;
                }
            } else {
                WorkCount wc = {0, 0, 0, 0};
                h = traverse<COUNT, DEFER>(sc, o, d, st, wc);
                if (COUNT) {
                    pc.v[RT_CNT_RAYS_TRACED]++;
                    pc.v[RT_CNT_INTERNAL] += wc.internal; pc.v[RT_CNT_LEAVES] += wc.leaves;
                    pc.v[RT_CNT_TRI_TESTS] += wc.tris;    pc.v[RT_CNT_POPS] += wc.pops;
                }
            }
            if (COUNT) pc.v[RT_CNT_RAYS]++;
            if (h.slot >= 0) {
                if (COUNT) pc.v[RT_CNT_HITS]++;
                const SurfaceHit hit = resolveHit(sc, h, o, d);
                const f3 dDiff = diffuseDir(hit.normal, seed, lambertian);
                const f3 dSpec = reflect3(d, hit.normal);
                d = lerpDir(dDiff, dSpec, hit.roughness);
                o = add3(hit.point, scale3(hit.normal, 1e-4f));
                f3 emitted = scale3(hit.emissionColor, hit.emission);
                if (lambertian) emitted = scale3(emitted, 2.0f);
                incoming = add3(incoming, mul3(emitted, rayColor));
                rayColor = mul3(rayColor, hit.color);
                if (hit.emission > 0.0f) break;
            } else {
                incoming = add3(incoming, sky);   // un-attenuated, as the shader does (:343)
                break;
            }
        }
        color = add3(color, incoming);
    }
    return div3(color, float(P.raysPerPixel));   // :365
}

template <bool COUNT>
RT_DEV void flushCounters(const PassParams& P, PathCounters& pc) {
    if constexpr (COUNT) {
#pragma unroll
        for (int k = 0; k < RT_CNT_COUNT; k++) {
            unsigned long long v = pc.v[k];
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
            if ((threadIdx.x & 31) == 0 && v) atomicAdd(P.counters + k, v);
        }
    }
}

template <bool COUNT, bool IDS, bool DEFER = false>      // DEFER: see triangleTest (rt_device.cuh)
__global__ void __launch_bounds__(kBlockThreads) rtPixelKernel(const __grid_constant__ PassParams P) {
    extern __shared__ uint2 stackMem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ox = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
    const int oy = blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
    const bool valid = ox < P.ownedW && oy < P.ownedH;
    PathCounters pc;
#pragma unroll
    for (int k = 0; k < RT_CNT_COUNT; k++) pc.v[k] = 0;
    if (valid) {
        const int px = ownedToPixel(ox, P.tileSizeX, P.numTilesX, P.tileX);
        const int py = ownedToPixel(oy, P.tileSizeY, P.numTilesY, P.tileY);
        StackRef st;
        st.base = stackMem; st.stride = kBlockThreads;
        int32_t primaryId = -1;
        const f3 curr = pixelFrame<COUNT, IDS, DEFER>(P, px, py, st, pc, primaryId);
        const size_t at = size_t(py) * size_t(P.width) + size_t(px);
        storePixel(P, at, accumulate(P, curr, at));
        if (IDS) P.primaryIds[at] = primaryId;
    }
    flushCounters<COUNT>(P, pc);
}

// ---------------------------------------------------------------------------------- utility kernels
// glReadPixels(GL_RGBA, GL_UNSIGNED_BYTE) of an RGBA32F target (main.py:434): clamp to [0,1], scale by
// 255, round to nearest.  NaN clamps to 0.
__global__ void rtToRgba8Kernel(const float4* __restrict__ src, uchar4* __restrict__ dst, size_t n) {
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    const size_t step = size_t(gridDim.x) * blockDim.x;
    for (; i < n; i += step) {
        const float4 p = src[i];
        auto conv = [](float x) -> unsigned char {
            x = (x != x) ? 0.0f : fminf(fmaxf(x, 0.0f), 1.0f);
            return (unsigned char)floorf(__fadd_rn(__fmul_rn(x, 255.0f), 0.5f));
        };
        dst[i] = make_uchar4(conv(p.x), conv(p.y), conv(p.z), conv(p.w));
    }
}

// Rows owned by tile row `tile` <-> dense buffer (multi-GPU band sharding).
__global__ void rtBandCopyKernel(float4* __restrict__ target, float4* __restrict__ dense, int width, int rows,
                                 int tileSizeY, int numTilesY, int tile, int toDense) {
    const size_t n = size_t(rows) * size_t(width);
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    const size_t step = size_t(gridDim.x) * blockDim.x;
    for (; i < n; i += step) {
        const int k = int(i / size_t(width)), x = int(i % size_t(width));
        const int y = (numTilesY == 1) ? k : (k / tileSizeY) * (numTilesY * tileSizeY) + tile * tileSizeY + (k % tileSizeY);
        const size_t at = size_t(y) * size_t(width) + size_t(x);
        if (toDense) dense[i] = target[at];
        else target[at] = dense[i];
    }
}

// After the frame-sharded reduce: the target holds the SUM of total frames -> mean (IEEE division, alpha 1).
__global__ void rtFinishSumKernel(float4* __restrict__ target, size_t n, float frames) {
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    const size_t step = size_t(gridDim.x) * blockDim.x;
    for (; i < n; i += step) {
        float4 p = target[i];
        p.x = __fdiv_rn(p.x, frames); p.y = __fdiv_rn(p.y, frames); p.z = __fdiv_rn(p.z, frames); p.w = 1.0f;
        target[i] = p;
    }
}

// rt_sum_resume: running mean of `frames` frames -> their sum (one multiply per colour channel, alpha kept).
__global__ void rtScaleSumKernel(float4* __restrict__ target, size_t n, float frames) {
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    const size_t step = size_t(gridDim.x) * blockDim.x;
    for (; i < n; i += step) {
        float4 p = target[i];
        p.x = __fmul_rn(p.x, frames); p.y = __fmul_rn(p.y, frames); p.z = __fmul_rn(p.z, frames);
        target[i] = p;
    }
}

// rt_scene_upload: one thread per triangle slot builds its TriRec (v0, the two edges and their cross product, exactly
// the operations of fragment.glsl:112-116 in that order) and ShadeRec from the caller's 144-byte triangle (nine
// float4: v0p v0n v1p v1n v2p v2n color emc surface; scene.py:13-30).
__global__ void rtBuildRecordsKernel(const float4* __restrict__ raw, const uint32_t* __restrict__ triId, uint32_t nSlots,
                                     TriRec* __restrict__ tris, ShadeRec* __restrict__ shade) {
    const uint32_t step = gridDim.x * blockDim.x;
    for (uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x; slot < nSlots; slot += step) {
        const float4* q = raw + size_t(triId[slot]) * 9;
        const float4 v0 = __ldg(q), n0 = __ldg(q + 1), v1 = __ldg(q + 2), n1 = __ldg(q + 3), v2 = __ldg(q + 4),
                     n2 = __ldg(q + 5), col = __ldg(q + 6), emc = __ldg(q + 7), surf = __ldg(q + 8);
        const float abx = __fsub_rn(v1.x, v0.x), aby = __fsub_rn(v1.y, v0.y), abz = __fsub_rn(v1.z, v0.z);
        const float acx = __fsub_rn(v2.x, v0.x), acy = __fsub_rn(v2.y, v0.y), acz = __fsub_rn(v2.z, v0.z);
        const float nx = __fsub_rn(__fmul_rn(aby, acz), __fmul_rn(acy, abz));
        const float ny = __fsub_rn(__fmul_rn(abz, acx), __fmul_rn(acz, abx));
        const float nz = __fsub_rn(__fmul_rn(abx, acy), __fmul_rn(acx, aby));
        float4* t = reinterpret_cast<float4*>(tris + slot);
        t[0] = make_float4(v0.x, v0.y, v0.z, nx);
        t[1] = make_float4(ny, nz, 0.0f, 0.0f);
        t[2] = make_float4(abx, aby, abz, acx);
        t[3] = make_float4(acy, acz, 0.0f, 0.0f);
        float4* s = reinterpret_cast<float4*>(shade + slot);
        s[0] = make_float4(n0.x, n0.y, n0.z, n1.x);
        s[1] = make_float4(n1.y, n1.z, n2.x, n2.y);
        s[2] = make_float4(n2.z, col.x, col.y, col.z);
        s[3] = make_float4(emc.x, emc.y, emc.z, surf.x);
        s[4] = make_float4(surf.y, 0.0f, 0.0f, 0.0f);
        s[5] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    }
}

}  // namespace

// ------------------------------------------------------------------------------------- launchers
// (C++ linkage, called from rt_api.cu)
int rtChainMaxBlocksPerSm();
size_t rtPixelKernelSmem(int stackDepth) { return size_t(stackDepth) * kBlockThreads * sizeof(uint2); }

cudaError_t rtLaunchPixel(const PassParams& P, bool count, bool ids, cudaStream_t stream) {
    dim3 grid((P.ownedW + 15) / 16, (P.ownedH + 7) / 8);
    if (grid.x == 0 || grid.y == 0) return cudaSuccess;
    const size_t smem = rtPixelKernelSmem(P.scene.stackDepth);
#define RT_LAUNCH(C, I)                                                                                   \
    do {                                                                                                  \
        if (smem > 48 * 1024)                                                                             \
            cudaFuncSetAttribute(rtPixelKernel<C, I>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)); \
        rtPixelKernel<C, I><<<grid, kBlockThreads, smem, stream>>>(P);                                   \
    } while (0)
    // the deferring instantiation (triangleTest) exists for the plain pass only; counting / id passes are not timed
    if (!count && !ids && P.scene.deferEdges) {
        if (smem > 48 * 1024)
            cudaFuncSetAttribute(rtPixelKernel<false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
        rtPixelKernel<false, false, true><<<grid, kBlockThreads, smem, stream>>>(P);
    } else if (count && ids) RT_LAUNCH(true, true);
    else if (count) RT_LAUNCH(true, false);
    else if (ids) RT_LAUNCH(false, true);
    else RT_LAUNCH(false, false);
#undef RT_LAUNCH
    return cudaGetLastError();
}

cudaError_t rtLaunchPersistent(const PassParams& Pin, bool count, bool ids, int sms, cudaStream_t stream) {
    PassParams P = Pin;
    if (P.ownedW <= 0 || P.ownedH <= 0) return cudaSuccess;
    const size_t base = vote::smemBytes(P.scene.stackDepth);
    if (const char* e = getenv("RT_VOTE_TUNE")) sscanf(e, "%d,%d,%d", &P.tune[0], &P.tune[1], &P.tune[2]);   // [2]: same forced decision as the samples rounds (specGate)
    const size_t smem = base;
    cudaError_t e = cudaMemsetAsync(P.workCounter, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    // no more CTAs than there is work for: one warp per 32 pixel indices
    const long long tiles = (long long)((P.ownedW + 7) / 8) * ((P.ownedH + 3) / 4);
    const int needed = int(std::min<long long>((tiles * 32 + vote::kThreads - 1) / vote::kThreads, 1 << 30));
#define RT_LAUNCH(C, I, D)                                                                                       \
    do {                                                                                                         \
        if (smem > 48 * 1024)                                                                                    \
            cudaFuncSetAttribute(vote::rtVoteKernel<C, I, D>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)); \
        int perSm = 0;                                                                                           \
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, vote::rtVoteKernel<C, I, D>, vote::kThreads, smem); \
        if (e != cudaSuccess) return e;                                                                          \
        if (perSm < 1) perSm = 1;                                                                                \
        if (const char* v = getenv("RT_VOTE_CTAS")) perSm = std::max(1, std::min(perSm, atoi(v)));                \
        vote::rtVoteKernel<C, I, D><<<std::min(sms * perSm, std::max(needed, 1)), vote::kThreads, smem, stream>>>(P); \
    } while (0)
    // the deferring instantiation (triangleTest) exists for the plain pass only; counting / id passes are not timed
    if (!count && !ids && P.scene.deferEdges) RT_LAUNCH(false, false, true);
    else if (count && ids) RT_LAUNCH(true, true, false);
    else if (count) RT_LAUNCH(true, false, false);
    else if (ids) RT_LAUNCH(false, true, false);
    else RT_LAUNCH(false, false, false);
#undef RT_LAUNCH
    return cudaGetLastError();
}

// ---- POOL variant: one persistent CTA per SM; shared memory = ray pool + the top of the tree
namespace {
template <int T, int K, int S>
cudaError_t launchPoolCfg(PassParams P, bool count, bool ids, int sms, int topWanted, cudaStream_t stream) {
    int dev = 0, maxSmem = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&maxSmem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    const size_t slots = pool::Smem<T, K>::slotBytes(P.scene.stackDepth);
    if (slots + 1024 > size_t(maxSmem)) return cudaErrorInvalidConfiguration;
    long long top = (long long)((size_t(maxSmem) - slots - 512) / 56);
    if (topWanted >= 0) top = std::min<long long>(top, topWanted);
    P.topCount = int(std::min<long long>(top, P.scene.nNodes));
    const size_t smem = pool::Smem<T, K>::bytes(P.scene.stackDepth, P.topCount);
    const long long tiles = (long long)((P.ownedW + 7) / 8) * ((P.ownedH + 3) / 4);
    const int needed = int(std::min<long long>((tiles * 32 + T * K - 1) / (T * K), 1 << 30));
    const int grid = std::min(sms, std::max(needed, 1));
    cudaError_t e = cudaSuccess;
#define RT_LAUNCH(C, I)                                                                                        \
    do {                                                                                                       \
        e = cudaFuncSetAttribute(pool::rtPoolKernel<T, K, S, C, I>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)); \
        if (e != cudaSuccess) return e;                                                                        \
        pool::rtPoolKernel<T, K, S, C, I><<<grid, T, smem, stream>>>(P);                                       \
    } while (0)
    if (count && ids) RT_LAUNCH(true, true);
    else if (count) RT_LAUNCH(true, false);
    else if (ids) RT_LAUNCH(false, true);
    else RT_LAUNCH(false, false);
#undef RT_LAUNCH
    return cudaGetLastError();
}
}  // namespace

cudaError_t rtLaunchPool(const PassParams& P, bool count, bool ids, int sms, cudaStream_t stream) {
    if (P.ownedW <= 0 || P.ownedH <= 0) return cudaSuccess;
    if (P.scene.stackDepth > pool::kMaxDepth) return rtLaunchPixel(P, count, ids, stream);   // very deep (degenerate) trees
    cudaError_t e = cudaMemsetAsync(P.workCounter, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    // tuning knobs (measurement only; the defaults are what profiles/ documents)
    int cfg = 0, top = 1024;
    if (const char* v = getenv("RT_POOL_CFG")) cfg = atoi(v);
    if (const char* v = getenv("RT_TOP_NODES")) top = atoi(v);
    switch (cfg) {
        case 1: return launchPoolCfg<256, 2, 4>(P, count, ids, sms, top, stream);
        case 2: return launchPoolCfg<256, 3, 4>(P, count, ids, sms, top, stream);
        case 3: return launchPoolCfg<384, 2, 2>(P, count, ids, sms, top, stream);
        case 4: return launchPoolCfg<384, 2, 8>(P, count, ids, sms, top, stream);
        case 5: return launchPoolCfg<512, 1, 4>(P, count, ids, sms, top, stream);
        default: return launchPoolCfg<384, 2, 4>(P, count, ids, sms, top, stream);
    }
}

// ---- WAVEFRONT variant: primary stage (one thread per owned pixel) -> queue -> path stage (persistent warps)
cudaError_t rtLaunchWavefront(const PassParams& P, bool count, bool ids, int sms, cudaStream_t stream) {
    if (P.ownedW <= 0 || P.ownedH <= 0) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(P.workCounter, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(P.queueCount, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    const size_t smem = wave::smemBytes(P.scene.stackDepth);
    dim3 grid((P.ownedW + 15) / 16, (P.ownedH + 7) / 8);
#define RT_LAUNCH1(C, I)                                                                                       \
    do {                                                                                                       \
        if (smem > 48 * 1024)                                                                                  \
            cudaFuncSetAttribute(wave::rtPrimaryKernel<C, I>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)); \
        wave::rtPrimaryKernel<C, I><<<grid, wave::kThreads, smem, stream>>>(P);                                \
    } while (0)
    if (count && ids) RT_LAUNCH1(true, true);
    else if (count) RT_LAUNCH1(true, false);
    else if (ids) RT_LAUNCH1(false, true);
    else RT_LAUNCH1(false, false);
#undef RT_LAUNCH1
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    if (P.nBounces <= 0 || P.raysPerPixel <= 0) return cudaSuccess;   // nothing was queued
    const long long pixels = (long long)P.ownedW * P.ownedH;
#define RT_LAUNCH2(C)                                                                                          \
    do {                                                                                                       \
        if (smem > 48 * 1024)                                                                                  \
            cudaFuncSetAttribute(wave::rtPathKernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)); \
        int perSm = 0;                                                                                         \
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, wave::rtPathKernel<C>, wave::kThreads, smem); \
        if (e != cudaSuccess) return e;                                                                        \
        if (perSm < 1) perSm = 1;                                                                              \
        const long long want = (pixels + wave::kThreads - 1) / wave::kThreads;                                 \
        wave::rtPathKernel<C><<<int(std::min<long long>((long long)sms * perSm, std::max<long long>(want, 1))), \
                                wave::kThreads, smem, stream>>>(P);                                            \
    } while (0)
    if (count) RT_LAUNCH2(true);
    else RT_LAUNCH2(false);
#undef RT_LAUNCH2
    return cudaGetLastError();
}

// ---- SAMPLES variant: primary stage -> queue -> rounds of [parallel evaluation of F(k) -> chain]
int rtSpecTableK(int raysPerPixel, int nBounces) {
    (void)nBounces;
    if (const char* e = getenv("RT_SPEC_K0")) {           // experiments: round 0's table size as a percentage of rays_per_pixel
        const int pct = atoi(e);
        if (pct >= 100) return std::min(spec::kMaxK / 4, std::max(8, int((long long)raysPerPixel * pct / 100)));
    }
    return spec::tableK(raysPerPixel);
}
int rtSpecRounds() { return spec::kRounds; }
// layout of the samples variant's counter array: [0,32) work counters (2 per round), [32,64) parked pixels per round,
// [64, 64 + (rounds + 1) * 32) size-class counts per round
constexpr int kSpecWorkUints = 64 + (spec::kRounds + 1) * spec::kClassStride;

cudaError_t rtSpecInit() {
    static uint2 lo[spec::kJumpLo], hi[spec::kJumpHi];
    spec::fillJumpTables(lo, hi);
    cudaError_t e = cudaMemcpyToSymbol(spec::cJumpLo, lo, sizeof lo);
    if (e != cudaSuccess) return e;
    return cudaMemcpyToSymbol(spec::cJumpHi, hi, sizeof hi);
}

// ---- WAVEFRONT pass without allocation (rt_render_pass_static): the queue and both counters live in a file-scope
// array instead of cudaMalloc'ed buffers.  Every GPU that loads the library holds the array until exit.  All calls share
// it: a call clears the counters it needs at zero on its stream, and calls that overlap on the GPU race.
namespace {
constexpr size_t kStaticMaxPixels = RT_STATIC_MAX_PIXELS;
constexpr size_t kStaticHeaderBytes = 256;              // [0] queue count, [1] work counter
constexpr size_t kStaticBytes = kStaticHeaderBytes + kStaticMaxPixels * (sizeof(float4) + sizeof(float4) + sizeof(uint32_t));
}  // namespace
static __device__ __align__(256) unsigned char gStaticScratch[kStaticBytes];

namespace {
constexpr int kStaticMaxDevices = 64;
std::mutex gStaticLock;
unsigned char* gStaticAddress[kStaticMaxDevices] = {};
}  // namespace

// The array's device address on `device`, which must be current.  In host code `gStaticScratch` is only a stand-in for
// the symbol, never a device pointer: the address comes from cudaGetSymbolAddress, once per device.
static cudaError_t staticScratchAddress(int device, unsigned char** out) {
    if (device < 0 || device >= kStaticMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> guard(gStaticLock);
    if (gStaticAddress[device] == nullptr) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, gStaticScratch);
        if (e != cudaSuccess) return e;
        gStaticAddress[device] = static_cast<unsigned char*>(p);
    }
    *out = gStaticAddress[device];
    return cudaSuccess;
}

cudaError_t rtLaunchWavefrontStatic(const PassParams& Pin, int device, int sms, cudaStream_t stream) {
    PassParams P = Pin;
    if (P.ownedW <= 0 || P.ownedH <= 0) return cudaSuccess;
    if ((unsigned long long)P.ownedW * (unsigned long long)P.ownedH > kStaticMaxPixels) return cudaErrorInvalidValue;
    unsigned char* base = nullptr;
    cudaError_t e = staticScratchAddress(device, &base);
    if (e != cudaSuccess) return e;
    P.queueCount = reinterpret_cast<unsigned int*>(base);
    P.workCounter = reinterpret_cast<unsigned int*>(base) + 1;
    P.queueA = reinterpret_cast<float4*>(base + kStaticHeaderBytes);
    P.queueB = P.queueA + kStaticMaxPixels;
    P.queuePix = reinterpret_cast<uint32_t*>(P.queueB + kStaticMaxPixels);
    P.queueWork = nullptr;                                  // counting passes only
    P.counters = nullptr; P.primaryIds = nullptr;
    if ((e = cudaMemsetAsync(base, 0, kStaticHeaderBytes, stream)) != cudaSuccess) return e;   // the last call's counts
    const size_t smem = wave::smemBytes(P.scene.stackDepth);
    dim3 grid((P.ownedW + 15) / 16, (P.ownedH + 7) / 8);
    if (smem > 48 * 1024) {
        cudaFuncSetAttribute(wave::rtPrimaryKernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
        cudaFuncSetAttribute(wave::rtPathKernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    }
    wave::rtPrimaryKernel<false, false><<<grid, wave::kThreads, smem, stream>>>(P);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    if (P.nBounces <= 0 || P.raysPerPixel <= 0) return cudaSuccess;   // nothing was queued
    int perSm = 0;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, wave::rtPathKernel<false>, wave::kThreads, smem)) != cudaSuccess) return e;
    if (perSm < 1) perSm = 1;
    const long long pixels = (long long)P.ownedW * P.ownedH;
    const long long want = (pixels + wave::kThreads - 1) / wave::kThreads;
    wave::rtPathKernel<false><<<int(std::min<long long>((long long)sms * perSm, std::max<long long>(want, 1))), wave::kThreads, smem, stream>>>(P);
    return cudaGetLastError();
}

// kernel launches of one samples pass that uses tables: stage 1 + round 0 (evaluate, chain) + 3 per later round
int rtSpecLaunches() { return 3 + 3 * (spec::kRounds - 1); }
int rtSpecWorkUints() { return kSpecWorkUints; }

cudaError_t rtLaunchSamples(const PassParams& Pin, bool ids, int sms, uint32_t* todoA, uint32_t* todoB, cudaStream_t stream) {
    PassParams P = Pin;
    if (P.ownedW <= 0 || P.ownedH <= 0) return cudaSuccess;
    if (const char* e = getenv("RT_VOTE_TUNE")) sscanf(e, "%d,%d,%d", &P.tune[0], &P.tune[1], &P.tune[2]);
    if (const char* e = getenv("RT_SPEC_MARGIN")) { int a = 0, b = 0; if (sscanf(e, "%d,%d", &a, &b) == 2) P.tune[3] = (a << 8) | (b & 255); }
    // workCounters, todoCount and classCount are one array (rt_api.cu): cleared with one memset
    cudaError_t e = cudaMemsetAsync(P.workCounters, 0, size_t(kSpecWorkUints) * sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    if ((e = cudaMemsetAsync(P.queueCount, 0, sizeof(unsigned int), stream)) != cudaSuccess) return e;
    const size_t smem = spec::smemBytes(P.scene.stackDepth);
    const size_t smemEval = vote::smemBytes(P.scene.stackDepth);
    dim3 grid((P.ownedW + 15) / 16, (P.ownedH + 7) / 8);
    if (smem > 48 * 1024) {
        cudaFuncSetAttribute(spec::rtSpecPrimaryKernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
        cudaFuncSetAttribute(spec::rtSpecPrimaryKernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
        cudaFuncSetAttribute(spec::rtSpecChainKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    }
    if (smemEval > 48 * 1024)
        cudaFuncSetAttribute(spec::rtSpecEvalKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smemEval));
    if (ids) spec::rtSpecPrimaryKernel<true><<<grid, spec::kThreads, smem, stream>>>(P);
    else spec::rtSpecPrimaryKernel<false><<<grid, spec::kThreads, smem, stream>>>(P);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    if (P.nBounces <= 1 || P.raysPerPixel <= 0) return cudaSuccess;   // nothing was queued
    const long long pixels = (long long)P.ownedW * P.ownedH;
    int perSmEval = 0, perSmChain = 0;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSmEval, spec::rtSpecEvalKernel, spec::kThreads, smemEval)) != cudaSuccess) return e;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSmChain, spec::rtSpecChainKernel, spec::kThreads, smem)) != cudaSuccess) return e;
    const long long wantEval = (pixels / 2 * std::max(P.specK, 1) + spec::kThreads - 1) / spec::kThreads;
    const long long wantChain = (pixels + spec::kThreads - 1) / spec::kThreads;
    const int gridEval = int(std::min<long long>((long long)sms * std::max(perSmEval, 1), std::max<long long>(wantEval, 1)));
    const int gridChain = int(std::min<long long>((long long)sms * std::max(perSmChain, 1), std::max<long long>(wantChain, 1)));
    // the rounds are launched unconditionally: a round without pixels (or the whole speculation, when stage 1 found
    // that most primary rays hit) is three kernels that read a counter and return
    P.todoFlat = todoA;
    P.todoSorted = todoB;
    const int gridSort = sms * 2;
    for (int round = 0; round < spec::kRounds; round++) {
        P.specRound = round;
        const bool tables = P.specK > 0 && P.specCapacity > 0;
        if (round > 0) spec::rtSpecSortKernel<<<gridSort, spec::kThreads, 0, stream>>>(P);
        if (tables) spec::rtSpecEvalKernel<<<gridEval, spec::kThreads, smemEval, stream>>>(P);
        spec::rtSpecChainKernel<<<gridChain, spec::kThreads, smem, stream>>>(P);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        if (!tables) break;
    }
    return cudaSuccess;
}

// ---- CHAIN variant: primary stage -> queue -> ONE persistent kernel (a pixel per lane + in-warp sample stealing)
int rtChainMaxBlocksPerSm() { return 8; }
size_t rtChainRingBytes(int sms) { return chain::ringBytes(sms * rtChainMaxBlocksPerSm()); }

namespace {
// ring entry e of lane t <- "F(k) = 1e9 with one hit" for k = e, e + kRing, e + 2 kRing or e + 3 kRing (by lane)
__global__ void rtChainPoisonKernel(float4* ring, size_t entries) {
    for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < entries; i += size_t(gridDim.x) * blockDim.x) {
        const uint32_t e = uint32_t(i % chain::kRing), t = uint32_t(i / chain::kRing);
        const uint32_t k = e + chain::kRing * (t & 3u);
        ring[i] = make_float4(1e9f, 1e9f, 1e9f, __uint_as_float(((k + 1u) << chain::kHitBits) | 1u));
    }
}
}  // namespace

cudaError_t rtLaunchChain(const PassParams& Pin, bool ids, int sms, cudaStream_t stream) {
    PassParams P = Pin;
    if (P.ownedW <= 0 || P.ownedH <= 0) return cudaSuccess;
    if (const char* e = getenv("RT_VOTE_TUNE")) sscanf(e, "%d,%d,%d,%d", &P.tune[0], &P.tune[1], &P.tune[2], &P.tune[3]);
    P.chainTune[0] = 16; P.chainTune[1] = 1; P.chainTune[2] = chain::kRing; P.chainTune[3] = 0;   // quota scale (sixteenths), helping, window, scattered queue order (off: see r02_variants.md)
    if (const char* e = getenv("RT_CHAIN_TUNE")) sscanf(e, "%d,%d,%d,%d", &P.chainTune[0], &P.chainTune[1], &P.chainTune[2], &P.chainTune[3]);
    cudaError_t e = cudaMemsetAsync(P.workCounters, 0, 64 * sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    if ((e = cudaMemsetAsync(P.queueCount, 0, sizeof(unsigned int), stream)) != cudaSuccess) return e;
    const size_t smem = spec::smemBytes(P.scene.stackDepth);
    dim3 grid((P.ownedW + 15) / 16, (P.ownedH + 7) / 8);
    if (smem > 48 * 1024) {
        cudaFuncSetAttribute(spec::rtSpecPrimaryKernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
        cudaFuncSetAttribute(spec::rtSpecPrimaryKernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    }
#ifdef RT_CHAIN_STATS
    struct StatsPrinter {
        const PassParams& P; cudaStream_t stream;
        ~StatsPrinter() {
            unsigned long long v[8]; unsigned int q = 0;
            cudaStreamSynchronize(stream);
            cudaMemcpy(v, P.workCounters + 16, sizeof v, cudaMemcpyDeviceToHost);
            cudaMemcpy(&q, P.queueCount, sizeof q, cudaMemcpyDeviceToHost);
            fprintf(stderr, "chain stats: queued %u  started %llu own %llu busy-iters %llu dropped %llu rays %llu idle-iters %llu shade-sections(lanes) %llu used %llu\n",
                    q, v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7]);
            unsigned long long x[12];
            cudaMemcpy(x, P.workCounters + 32, sizeof x, cudaMemcpyDeviceToHost);
            fprintf(stderr, "   warp iterations max %llu sum %llu; warp life max %.3f ms, sum %.1f ms; lives in 2 ms bins:", x[0], x[1], x[2] * 1e-6, x[3] * 1e-6);
            for (int i = 0; i < 8; i++) fprintf(stderr, " %llu", x[4 + i]);
            fprintf(stderr, "\n");
            static unsigned int tl[3][256];
            cudaMemcpyFromSymbol(tl, chain::gTimeline, sizeof tl);
            const char* names[3] = {"started ", "consumed", "finished"};
            for (int w = 0; w < 3; w++) {
                fprintf(stderr, "   timeline %s per 0.25 ms:", names[w]);
                int last = 255;
                while (last > 0 && tl[w][last] == 0) last--;
                for (int i = 0; i <= last; i++) fprintf(stderr, " %u", tl[w][i]);
                fprintf(stderr, "\n");
            }
        }
    } printer{P, stream};
    {
        static unsigned int zeros[3][256];
        unsigned long long z = 0;
        cudaMemcpyToSymbolAsync(chain::gTimeline, zeros, sizeof zeros, 0, cudaMemcpyHostToDevice, stream);
        cudaMemcpyToSymbolAsync(chain::gTimelineStart, &z, sizeof z, 0, cudaMemcpyHostToDevice, stream);
        cudaStreamSynchronize(stream);
    }
#endif
    if (ids) spec::rtSpecPrimaryKernel<true><<<grid, spec::kThreads, smem, stream>>>(P);
    else spec::rtSpecPrimaryKernel<false><<<grid, spec::kThreads, smem, stream>>>(P);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    if (P.nBounces <= 1 || P.raysPerPixel <= 0) return cudaSuccess;   // nothing was queued
    const size_t smemChain = chain::smemBytes(P.scene.stackDepth);
    if (smemChain > 48 * 1024) {
        e = cudaFuncSetAttribute(chain::rtChainKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smemChain));
        if (e != cudaSuccess) return e;
    }
    int perSm = 0;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, chain::rtChainKernel, chain::kThreads, smemChain)) != cudaSuccess) return e;
    if (perSm < 1) perSm = 1;
    if (const char* v = getenv("RT_VOTE_CTAS")) perSm = std::max(1, std::min(perSm, atoi(v)));
    perSm = std::min(perSm, rtChainMaxBlocksPerSm());      // what rt_api.cu sized the rings for
    const long long pixels = (long long)P.ownedW * P.ownedH;
    const long long want = (pixels + chain::kThreads - 1) / chain::kThreads;
    const int gridChain = int(std::min<long long>((long long)sms * perSm, std::max<long long>(want, 1)));
    if (getenv("RT_CHAIN_POISON")) {
        // test aid (tests/test_gpu_parity.py): the rings are never initialised — a slot clears its own at its first help — so
        // fill them with entries that LOOK like results of small offsets; a read before the clear would poison the image
        rtChainPoisonKernel<<<sms * 8, 256, 0, stream>>>(P.chainRing, size_t(gridChain) * chain::kThreads * chain::kRing);
    }
    chain::rtChainKernel<<<gridChain, chain::kThreads, smemChain, stream>>>(P);
    return cudaGetLastError();
}

cudaError_t rtLaunchToRgba8(const float4* src, uchar4* dst, size_t n, cudaStream_t stream) {
    rtToRgba8Kernel<<<148 * 8, 256, 0, stream>>>(src, dst, n);
    return cudaGetLastError();
}

cudaError_t rtLaunchBandCopy(float4* target, float4* dense, int width, int rows, int tileSizeY, int numTilesY,
                             int tile, int toDense, cudaStream_t stream) {
    if (rows <= 0) return cudaSuccess;
    rtBandCopyKernel<<<148 * 8, 256, 0, stream>>>(target, dense, width, rows, tileSizeY, numTilesY, tile, toDense);
    return cudaGetLastError();
}

cudaError_t rtLaunchFinishSum(float4* target, size_t n, float frames, cudaStream_t stream) {
    rtFinishSumKernel<<<148 * 8, 256, 0, stream>>>(target, n, frames);
    return cudaGetLastError();
}

cudaError_t rtLaunchScaleSum(float4* target, size_t n, float frames, cudaStream_t stream) {
    rtScaleSumKernel<<<148 * 8, 256, 0, stream>>>(target, n, frames);
    return cudaGetLastError();
}

cudaError_t rtLaunchBuildRecords(const float4* raw, const uint32_t* triId, uint32_t nSlots, TriRec* tris, ShadeRec* shade,
                                 cudaStream_t stream) {
    const unsigned blocks = std::min<unsigned>(148 * 8, (nSlots + 255) / 256);
    rtBuildRecordsKernel<<<blocks, 256, 0, stream>>>(raw, triId, nSlots, tris, shade);
    return cudaGetLastError();
}

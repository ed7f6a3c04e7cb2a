// rt_wave.cuh — variant WAVEFRONT: the pixel-frame split at its one natural seam.
//
// Every sample of a pixel-frame starts with the SAME primary ray (fragment.glsl:361-362), so the frame is
//   stage 1  rtPrimaryKernel  one thread per owned pixel: seed, jittered primary direction, primary traversal
//                             (coherent rays, no divergence problem).  A pixel whose primary ray misses is
//                             finished here (every sample of it is "0 + sky", :342-346).  A pixel that hits
//                             appends a 36-byte record (direction, RNG state, hit) to a compact queue with one
//                             warp-aggregated atomic.
//   stage 2  rtPathKernel     persistent warps pull 32 records at a time and run the sample/bounce loops.  All
//                             lanes of a warp now own pixels with the same amount of work (rays_per_pixel paths
//                             that all start on a surface), so the shading blocks run with full warps — in the
//                             one-kernel variants a warp that touched the dragon's silhouette ran its 64 samples
//                             with a handful of lanes while the sky lanes idled (ncu: shading at 4-5 of 32 lanes).
//                             What is identical for all samples of the pixel is computed once: the primary
//                             surface point, normal and material, the normalised mirror direction, the first
//                             bounce's origin, emitted light and throughput.
// Bit-exactness: every value is produced by the same operations on the same operands as in pixelFrame()
// (rt_kernels.cu); hoisting a computation out of the sample loop does not change its result.
#pragma once
#include "rt_shade.cuh"

namespace wave {

constexpr int kThreads = 128;

struct WorkAcc {
    uint32_t v[RT_CNT_COUNT];
};

template <bool COUNT>
RT_DEV void flush(const PassParams& P, WorkAcc& pc) {
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

// ------------------------------------------------------------------------------------------ stage 1
template <bool COUNT, bool IDS>
__global__ void __launch_bounds__(kThreads) rtPrimaryKernel(const __grid_constant__ PassParams P) {
    extern __shared__ uint2 stackMem[];
    const DeviceScene& sc = P.scene;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ox = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
    const int oy = blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
    const bool valid = ox < P.ownedW && oy < P.ownedH;
    WorkAcc pc;
#pragma unroll
    for (int k = 0; k < RT_CNT_COUNT; k++) pc.v[k] = 0;

    bool queued = false;
    f3 dir = mk3(0.f, 0.f, 0.f);
    uint32_t seed = 0;
    RayHit hit;
    hit.slot = -1; hit.t = 0.f; hit.u = 0.f; hit.v = 0.f;
    WorkCount work = {0, 0, 0, 0};
    int px = 0, py = 0;
    if (valid) {
        px = ownedToPixel(ox, P.tileSizeX, P.numTilesX, P.tileX);
        py = ownedToPixel(oy, P.tileSizeY, P.numTilesY, P.tileY);
        const size_t at = size_t(py) * size_t(P.width) + size_t(px);
        const bool casts = P.nBounces > 0 && P.raysPerPixel > 0;
        if (casts) {
            StackRef st;
            st.base = stackMem; st.stride = kThreads;
            dir = primaryDirection(P, px, py, seed);
            hit = traverse<COUNT>(sc, ld3(P.camPos), dir, st, work);
            if (COUNT) pc.v[RT_CNT_RAYS_TRACED]++;
        }
        if (IDS) P.primaryIds[at] = hit.slot >= 0 ? int32_t(__ldg(sc.triId + hit.slot)) : -1;
        queued = hit.slot >= 0;
        if (!queued) {
            // no ray, or the primary ray missed: every sample is incoming = 0 + sky (:342-346), summed as trace() does
            f3 color = mk3(0.f, 0.f, 0.f);
            if (casts) {
                const f3 inc = add3(mk3(0.f, 0.f, 0.f), scale3(mk3(0.1f, 0.6f, 0.92f), P.skyBrightness));
                for (int s = 0; s < P.raysPerPixel; s++) color = add3(color, inc);
            }
            storePixel(P, at, accumulate(P, div3(color, float(P.raysPerPixel)), at));   // :365, :409-416
            if (COUNT && P.raysPerPixel > 0) {
                pc.v[RT_CNT_PATHS] += P.raysPerPixel;
                if (casts) {
                    pc.v[RT_CNT_RAYS] += P.raysPerPixel;
                    pc.v[RT_CNT_INTERNAL] += work.internal * P.raysPerPixel; pc.v[RT_CNT_LEAVES] += work.leaves * P.raysPerPixel;
                    pc.v[RT_CNT_TRI_TESTS] += work.tris * P.raysPerPixel;    pc.v[RT_CNT_POPS] += work.pops * P.raysPerPixel;
                }
            }
        }
    }
    // append the pixels that go on to stage 2 (one atomic per warp)
    const uint32_t m = __ballot_sync(0xffffffffu, queued);
    if (m != 0u) {
        unsigned int base = 0;
        const int leader = __ffs(m) - 1;
        if (lane == leader) base = atomicAdd(P.queueCount, (unsigned int)__popc(m));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (queued) {
            const unsigned int j = base + (unsigned int)__popc(m & ((1u << lane) - 1u));
            P.queueA[j] = make_float4(dir.x, dir.y, dir.z, __uint_as_float(seed));
            P.queueB[j] = make_float4(hit.t, hit.u, hit.v, __int_as_float(hit.slot));
            P.queuePix[j] = uint32_t(px) | (uint32_t(py) << 16);
            if (COUNT) P.queueWork[j] = make_uint4(work.internal, work.leaves, work.tris, work.pops);
        }
    }
    flush<COUNT>(P, pc);
}

// ------------------------------------------------------------------------------------------ stage 2
// fragment.glsl:234-240 with the specular half precomputed: s1 = normalize(specular) * t, k0 = 1 - t, t = 1 - roughness.
RT_DEV f3 lerpDirPre(f3 diffuse, float k0, f3 s1) {
    const f3 zero = mk3(0.f, 0.f, 0.f);
    const f3 d0 = length3(diffuse) > 0.0f ? normalize3(diffuse) : zero;
    return normalize3(add3(scale3(d0, k0), s1));
}

template <bool COUNT>
__global__ void __launch_bounds__(kThreads) rtPathKernel(const __grid_constant__ PassParams P) {
    extern __shared__ uint2 stackMem[];
    const DeviceScene& sc = P.scene;
    const uint32_t lane = threadIdx.x & 31u;
    StackRef st;
    st.base = stackMem; st.stride = kThreads;
    const unsigned int total = *P.queueCount;
    const f3 camPos = ld3(P.camPos);
    const f3 zero = mk3(0.f, 0.f, 0.f), one = mk3(1.f, 1.f, 1.f);
    const f3 sky = scale3(mk3(0.1f, 0.6f, 0.92f), P.skyBrightness);   // :242-244, :388
    const bool lambertian = P.lambertian != 0;
    WorkAcc pc;
#pragma unroll
    for (int k = 0; k < RT_CNT_COUNT; k++) pc.v[k] = 0;

    for (;;) {
        // the warp claims the next 32 records
        unsigned int base = 0;
        if (lane == 0) base = atomicAdd(P.workCounter, 32u);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= total) break;
        const unsigned int j = base + lane;
        if (j < total) {
            const float4 ra = P.queueA[j], rb = P.queueB[j];
            const uint32_t pix = P.queuePix[j];
            const f3 primaryDir = mk3(ra.x, ra.y, ra.z);
            uint32_t seed = __float_as_uint(ra.w);
            RayHit ph;
            ph.t = rb.x; ph.u = rb.y; ph.v = rb.z; ph.slot = __float_as_int(rb.w);
            uint4 pw = make_uint4(0u, 0u, 0u, 0u);
            if (COUNT) pw = P.queueWork[j];

            // once per pixel-frame: everything about the first bounce that does not depend on the sample
            const SurfaceHit hit0 = resolveHit(sc, ph, camPos, primaryDir);
            const f3 spec0 = reflect3(primaryDir, hit0.normal);
            const float t0 = sub(1.0f, hit0.roughness);
            const float k0 = sub(1.0f, t0);
            const f3 s1 = scale3(length3(spec0) > 0.0f ? normalize3(spec0) : zero, t0);
            const f3 origin0 = add3(hit0.point, scale3(hit0.normal, 1e-4f));
            f3 emitted0 = scale3(hit0.emissionColor, hit0.emission);
            if (lambertian) emitted0 = scale3(emitted0, 2.0f);
            const f3 incoming0 = add3(zero, mul3(emitted0, one));
            const f3 rayColor0 = mul3(one, hit0.color);
            const bool stop0 = hit0.emission > 0.0f;
            const f3 normal0 = hit0.normal;

            f3 color = zero;
            for (int s = 0; s < P.raysPerPixel; s++) {
                // raytrace(), fragment.glsl:309-350; bounce 0 is the cached primary hit
                if (COUNT) {
                    pc.v[RT_CNT_PATHS]++; pc.v[RT_CNT_RAYS]++; pc.v[RT_CNT_HITS]++;
                    pc.v[RT_CNT_INTERNAL] += pw.x; pc.v[RT_CNT_LEAVES] += pw.y;
                    pc.v[RT_CNT_TRI_TESTS] += pw.z; pc.v[RT_CNT_POPS] += pw.w;
                }
                const f3 dDiff0 = diffuseDir(normal0, seed, lambertian);
                f3 d = lerpDirPre(dDiff0, k0, s1);
                f3 o = origin0;
                f3 incoming = incoming0, rayColor = rayColor0;
                if (!stop0) {
                    for (int i = 1; i < P.nBounces; i++) {
                        WorkCount wc = {0, 0, 0, 0};
                        const RayHit h = traverse<COUNT>(sc, o, d, st, wc);
                        if (COUNT) {
                            pc.v[RT_CNT_RAYS_TRACED]++; pc.v[RT_CNT_RAYS]++;
                            pc.v[RT_CNT_INTERNAL] += wc.internal; pc.v[RT_CNT_LEAVES] += wc.leaves;
                            pc.v[RT_CNT_TRI_TESTS] += wc.tris;    pc.v[RT_CNT_POPS] += wc.pops;
                        }
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
                }
                color = add3(color, incoming);
            }
            const int px = int(pix & 0xffffu), py = int(pix >> 16);
            const size_t at = size_t(py) * size_t(P.width) + size_t(px);
            storePixel(P, at, accumulate(P, div3(color, float(P.raysPerPixel)), at));   // :365, :409-416
        }
    }
    flush<COUNT>(P, pc);
}

inline size_t smemBytes(int stackDepth) { return size_t(stackDepth) * kThreads * sizeof(uint2); }

}  // namespace wave

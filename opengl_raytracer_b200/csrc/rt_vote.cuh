// rt_vote.cuh — variant PERSISTENT: persistent warps, warp-vote scheduled state machine.
//
// Why: in the one-thread-per-pixel kernel every lane walks its own sample/bounce/traversal loops, so a warp
// pays max-over-lanes at every level, and — worse — the rare, expensive blocks (leaf triangle tests: 3 % of the
// node steps; shading) are executed for one or two lanes at a time (ncu capture B: 11.3 of 32 lanes active,
// leaf/pop code with 4-7 lanes).  Here each lane is a small state machine
//
//     INTERNAL (at an internal node)   LEAF (at a leaf)   SHADE (traversal finished / needs a pixel)   DONE
//
// and every iteration of the warp's loop executes exactly ONE section, chosen by ballot: lanes that reach a
// leaf or finish a ray wait until enough lanes are in the same state (or nobody else can make progress), so
// the expensive sections run densely populated and the hot internal-node section is never interleaved with
// them.  All branches that select code are warp-uniform, so lanes stay converged.  Lanes fetch their next
// pixel from a global counter (warp-aggregated atomic), which also removes the sky-versus-mesh imbalance.
//
// Per-ray semantics are exactly those of traverse()/pixelFrame() in rt_device.cuh / rt_kernels.cu — the
// visiting order inside a ray never changes, only which lanes run when.
#pragma once
#include "rt_shade.cuh"

namespace vote {

#ifndef RT_VOTE_THREADS
#define RT_VOTE_THREADS 128
#endif
constexpr int kThreads = RT_VOTE_THREADS;
constexpr int kScratchWords = 7;          // per lane: primary direction (3) + primary hit (slot, t, u, v)
constexpr int MODE_INTERNAL = 0, MODE_LEAF = 1, MODE_SHADE = 2, MODE_DONE = 3;

inline size_t smemBytes(int stackDepth) {
    return size_t(stackDepth) * kThreads * sizeof(uint2) + size_t(kScratchWords) * kThreads * sizeof(uint32_t);
}
// Top of the tree in shared memory (nodes are numbered breadth-first, so "the first topCount nodes" are the top
// levels): four float4 arrays (q0 | q1 | q2 | refs) so that lanes reading record `cur` spread over all bank groups.
inline size_t topBytes(int topCount) { return size_t(topCount) * 64; }

struct Lane {
    // traversal
    f3 o, d, inv;
    RayHit best;
    int32_t cur;
    uint32_t sp;
    int32_t mode;
    // path
    f3 incoming, rayColor, color;
    uint32_t seed;
    int32_t sample, bounce;
    int32_t px, py;
    bool fresh;          // the traversal in flight is the pixel's primary ray
    bool needPixel;
};

// Start the traversal of (o, d): root slab test (fragment.glsl:253-262).
template <bool COUNT>
RT_DEV void beginRay(const DeviceScene& sc, Lane& L, uint32_t* cnt) {
    L.inv = mk3(__frcp_rn(L.d.x), __frcp_rn(L.d.y), __frcp_rn(L.d.z));
    L.best.t = 1e30f; L.best.u = 0.f; L.best.v = 0.f; L.best.slot = -1;
    L.sp = 0;
    L.cur = sc.rootRef;
    bool running = L.cur != RT_REF_NONE;
    if (COUNT) { cnt[RT_CNT_POPS]++; cnt[RT_CNT_RAYS_TRACED]++; }
    if (running) {
        bool hitRoot;
        const float tRoot = slabTest(L.o, L.inv, sc.rootMin[0], sc.rootMin[1], sc.rootMin[2], sc.rootMax[0],
                                     sc.rootMax[1], sc.rootMax[2], hitRoot);
        running = hitRoot && !(tRoot > L.best.t);
    }
    L.mode = running ? (L.cur >= 0 ? MODE_INTERNAL : MODE_LEAF) : MODE_SHADE;
}

// Pop until an entry survives the cull against the current closestT (fragment.glsl:256-262).
RT_DEV void popNext(Lane& L, const uint2* stack, uint32_t stride) {
    bool found = false;
    while (L.sp > 0 && !found) {
        L.sp--;
        const uint2 e = stack[L.sp * stride];
        if (!(__uint_as_float(e.y) > L.best.t)) { L.cur = int32_t(e.x); found = true; }
    }
    L.mode = found ? (L.cur >= 0 ? MODE_INTERNAL : MODE_LEAF) : MODE_SHADE;
}

// One internal-node step (fragment.glsl:281-304), see traverse() for the order-preserving details.
template <bool COUNT>
RT_DEV void stepInternal(const DeviceScene& sc, Lane& L, uint2* stack, uint32_t stride, uint32_t* cnt,
                         const float4* top, int topCount) {
    float4 q0, q1, q2;
    int2 refs;
    if (L.cur < topCount) {
        q0 = top[L.cur]; q1 = top[topCount + L.cur]; q2 = top[2 * topCount + L.cur];
        const float4 r = top[3 * topCount + L.cur];
        refs = make_int2(__float_as_int(r.x), __float_as_int(r.y));
    } else {
        const float4* np = reinterpret_cast<const float4*>(sc.nodes + L.cur);
        q0 = __ldg(np); q1 = __ldg(np + 1); q2 = __ldg(np + 2);
        refs = __ldg(reinterpret_cast<const int2*>(np + 3));
    }
    bool hitA, hitB;
    const float tA = slabTest(L.o, L.inv, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, hitA);
    const float tB = slabTest(L.o, L.inv, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, hitB);
    if (COUNT) { cnt[RT_CNT_INTERNAL]++; cnt[RT_CNT_POPS] += (hitA ? 1u : 0u) + (hitB ? 1u : 0u); }
    const bool aFirst = hitA && (!hitB || tA < tB);
    const int32_t nearRef = aFirst ? refs.x : refs.y, farRef = aFirst ? refs.y : refs.x;
    const float nearT = aFirst ? tA : tB, farT = aFirst ? tB : tA;
    const bool goNear = (hitA || hitB) && !(nearT > L.best.t);
    const bool pushFar = hitA && hitB && !(farT > L.best.t);
    if (pushFar) {
        stack[L.sp * stride] = make_uint2(uint32_t(farRef), __float_as_uint(farT));
        L.sp++;
    }
    if (goNear) {
        L.cur = nearRef;
        L.mode = nearRef >= 0 ? MODE_INTERNAL : MODE_LEAF;
    } else {
        popNext(L, stack, stride);
    }
}

// One leaf (fragment.glsl:266-279): every triangle in slot order, first found wins ties; then pop.
template <bool COUNT>
RT_DEV void stepLeaf(const DeviceScene& sc, Lane& L, const uint2* stack, uint32_t stride, uint32_t* cnt) {
    const uint32_t code = uint32_t(~L.cur);
    uint32_t count = code & ((1u << RT_LEAF_BITS) - 1u);
    uint32_t first = code >> RT_LEAF_BITS;
    if (count == RT_LEAF_BIG) {
        const uint2 big = __ldg(sc.bigLeaves + first);
        first = big.x; count = big.y;
    }
    if (COUNT) { cnt[RT_CNT_LEAVES]++; cnt[RT_CNT_TRI_TESTS] += count; }
    for (uint32_t k = 0; k < count; k++) {
        float t, u, v;
        if (triangleTest(sc.tris + first + k, L.o, L.d, L.best.t, t, u, v)) {
            L.best.t = t; L.best.u = u; L.best.v = v; L.best.slot = int32_t(first + k);
        }
    }
    popNext(L, stack, stride);
}


// Scheduling thresholds: a section runs when at least this many lanes wait for it, or when no lane can do
// anything cheaper.  Tuned on B200 (profiles/r01_ncu_notes.md).
#ifndef RT_VOTE_LEAF_LANES
#define RT_VOTE_LEAF_LANES 8
#endif
#ifndef RT_VOTE_SHADE_LANES
#define RT_VOTE_SHADE_LANES 8
#endif

template <bool COUNT, bool IDS>
__global__ void __launch_bounds__(kThreads) rtVoteKernel(const __grid_constant__ PassParams P) {
    extern __shared__ uint2 smem[];
    const DeviceScene& sc = P.scene;
    const uint32_t lane = threadIdx.x & 31u;
    uint2* const stack = smem + threadIdx.x;
    const uint32_t stride = kThreads;
    uint32_t* const scratch = reinterpret_cast<uint32_t*>(smem + size_t(sc.stackDepth) * kThreads) + threadIdx.x;
    float4* const top = reinterpret_cast<float4*>(reinterpret_cast<uint32_t*>(smem + size_t(sc.stackDepth) * kThreads) +
                                                  size_t(kScratchWords) * kThreads);
    const int topCount = P.topCount;
    for (int i = threadIdx.x; i < topCount * 4; i += kThreads) {
        const int node = i >> 2, part = i & 3;
        top[part * topCount + node] = __ldg(reinterpret_cast<const float4*>(sc.nodes + node) + part);
    }
    __syncthreads();

    const int tilesX = (P.ownedW + 7) >> 3, tilesY = (P.ownedH + 3) >> 2;
    const unsigned int totalIdx = (unsigned int)tilesX * (unsigned int)tilesY * 32u;
    const f3 camPos = ld3(P.camPos);
    const f3 zero = mk3(0.f, 0.f, 0.f);
    const f3 sky = scale3(mk3(0.1f, 0.6f, 0.92f), P.skyBrightness);   // fragment.glsl:242-244, :388
    const bool lambertian = P.lambertian != 0;
    const bool degenerate = P.nBounces <= 0 || P.raysPerPixel <= 0;   // no ray is ever cast

    uint32_t cnt[RT_CNT_COUNT];
    uint32_t primWork[4] = {0, 0, 0, 0}, primBase[4] = {0, 0, 0, 0};
#pragma unroll
    for (int k = 0; k < RT_CNT_COUNT; k++) cnt[k] = 0;

    Lane L;
    L.o = camPos; L.d = zero; L.inv = zero;
    L.best.t = 0.f; L.best.u = 0.f; L.best.v = 0.f; L.best.slot = -1;
    L.cur = 0; L.sp = 0; L.mode = MODE_SHADE;
    L.incoming = zero; L.rayColor = zero; L.color = zero;
    L.seed = 0; L.sample = 0; L.bounce = 0; L.px = 0; L.py = 0;
    L.fresh = false; L.needPixel = true;

    for (;;) {
        const uint32_t mI = __ballot_sync(0xffffffffu, L.mode == MODE_INTERNAL);
        const uint32_t mL = __ballot_sync(0xffffffffu, L.mode == MODE_LEAF);
        const uint32_t mS = __ballot_sync(0xffffffffu, L.mode == MODE_SHADE);
        if ((mI | mL | mS) == 0u) break;

        if (__popc(mS) >= RT_VOTE_SHADE_LANES || (mI | mL) == 0u) {
            // ------------------------------------------------------------------ SHADE section
            if (L.mode == MODE_SHADE && !L.needPixel) {
                bool pathEnds = false, pixelEnds = false;
                if (L.fresh) {
                    // the pixel's primary ray just finished: keep its hit for the other samples (:361-362)
                    L.fresh = false;
                    scratch[3 * kThreads] = uint32_t(L.best.slot);
                    scratch[4 * kThreads] = __float_as_uint(L.best.t);
                    scratch[5 * kThreads] = __float_as_uint(L.best.u);
                    scratch[6 * kThreads] = __float_as_uint(L.best.v);
                    if (COUNT) {
                        primWork[0] = cnt[RT_CNT_INTERNAL] - primBase[0]; primWork[1] = cnt[RT_CNT_LEAVES] - primBase[1];
                        primWork[2] = cnt[RT_CNT_TRI_TESTS] - primBase[2]; primWork[3] = cnt[RT_CNT_POPS] - primBase[3];
                    }
                    if (IDS) {
                        P.primaryIds[size_t(L.py) * size_t(P.width) + size_t(L.px)] =
                            L.best.slot >= 0 ? int32_t(__ldg(sc.triId + L.best.slot)) : -1;
                    }
                    if (L.best.slot < 0) {
                        // primary miss: every sample of the pixel-frame is "incoming = 0 + sky" (:342-346)
                        const f3 inc = add3(zero, sky);
                        for (int s = 0; s < P.raysPerPixel; s++) L.color = add3(L.color, inc);
                        if (COUNT) {
                            cnt[RT_CNT_RAYS] += P.raysPerPixel; cnt[RT_CNT_PATHS] += P.raysPerPixel - 1;
                            cnt[RT_CNT_INTERNAL] += primWork[0] * (P.raysPerPixel - 1);
                            cnt[RT_CNT_LEAVES] += primWork[1] * (P.raysPerPixel - 1);
                            cnt[RT_CNT_TRI_TESTS] += primWork[2] * (P.raysPerPixel - 1);
                            cnt[RT_CNT_POPS] += primWork[3] * (P.raysPerPixel - 1);
                        }
                        pixelEnds = true;
                    }
                }
                if (!pixelEnds) {
                    // one iteration of raytrace()'s loop body (fragment.glsl:313-347) for the ray just traced
                    if (COUNT) cnt[RT_CNT_RAYS]++;
                    if (L.best.slot >= 0) {
                        if (COUNT) cnt[RT_CNT_HITS]++;
                        const SurfaceHit hit = resolveHit(sc, L.best, L.o, L.d);
                        const f3 dDiff = diffuseDir(hit.normal, L.seed, lambertian);
                        const f3 dSpec = reflect3(L.d, hit.normal);
                        L.d = lerpDir(dDiff, dSpec, hit.roughness);
                        L.o = add3(hit.point, scale3(hit.normal, 1e-4f));
                        f3 emitted = scale3(hit.emissionColor, hit.emission);
                        if (lambertian) emitted = scale3(emitted, 2.0f);
                        L.incoming = add3(L.incoming, mul3(emitted, L.rayColor));
                        L.rayColor = mul3(L.rayColor, hit.color);
                        L.bounce++;
                        pathEnds = hit.emission > 0.0f || L.bounce >= P.nBounces;
                        if (!pathEnds) beginRay<COUNT>(sc, L, cnt);
                    } else {
                        L.incoming = add3(L.incoming, sky);   // un-attenuated, as the shader does (:343)
                        pathEnds = true;
                    }
                    if (pathEnds) {
                        L.color = add3(L.color, L.incoming);   // trace(), :359
                        L.sample++;
                        if (L.sample >= P.raysPerPixel) {
                            pixelEnds = true;
                        } else {
                            // next sample: same primary ray, RNG state carried over (:358-363)
                            L.incoming = zero; L.rayColor = mk3(1.f, 1.f, 1.f); L.bounce = 0;
                            L.o = camPos;
                            L.d = mk3(__uint_as_float(scratch[0]), __uint_as_float(scratch[kThreads]),
                                      __uint_as_float(scratch[2 * kThreads]));
                            L.best.slot = int32_t(scratch[3 * kThreads]);
                            L.best.t = __uint_as_float(scratch[4 * kThreads]);
                            L.best.u = __uint_as_float(scratch[5 * kThreads]);
                            L.best.v = __uint_as_float(scratch[6 * kThreads]);
                            L.mode = MODE_SHADE;
                            if (COUNT) {
                                cnt[RT_CNT_PATHS]++;
                                cnt[RT_CNT_INTERNAL] += primWork[0]; cnt[RT_CNT_LEAVES] += primWork[1];
                                cnt[RT_CNT_TRI_TESTS] += primWork[2]; cnt[RT_CNT_POPS] += primWork[3];
                            }
                        }
                    }
                }
                if (pixelEnds) {
                    const size_t at = size_t(L.py) * size_t(P.width) + size_t(L.px);
                    P.next[at] = accumulate(P, div3(L.color, float(P.raysPerPixel)), at);   // :365, :409-416
                    L.needPixel = true;
                    L.mode = MODE_SHADE;
                }
            }
            // ---- work fetch: lanes without a pixel claim the next indices (one atomic per warp)
            for (;;) {
                const uint32_t want = __ballot_sync(0xffffffffu, L.mode == MODE_SHADE && L.needPixel);
                if (want == 0u) break;
                unsigned int base = 0;
                if (lane == uint32_t(__ffs(want) - 1)) base = atomicAdd(P.workCounter, (unsigned int)__popc(want));
                base = __shfl_sync(0xffffffffu, base, __ffs(want) - 1);
                if (L.mode == MODE_SHADE && L.needPixel) {
                    const unsigned int idx = base + (unsigned int)__popc(want & ((1u << lane) - 1u));
                    if (idx >= totalIdx) {
                        L.mode = MODE_DONE; L.needPixel = false;
                    } else {
                        const unsigned int tile = idx >> 5, within = idx & 31u;
                        const int ox = int(tile % (unsigned int)tilesX) * 8 + int(within & 7u);
                        const int oy = int(tile / (unsigned int)tilesX) * 4 + int(within >> 3);
                        if (ox < P.ownedW && oy < P.ownedH) {
                            L.needPixel = false;
                            L.px = ownedToPixel(ox, P.tileSizeX, P.numTilesX, P.tileX);
                            L.py = ownedToPixel(oy, P.tileSizeY, P.numTilesY, P.tileY);
                            L.color = zero; L.incoming = zero; L.rayColor = mk3(1.f, 1.f, 1.f);
                            L.sample = 0; L.bounce = 0;
                            if (degenerate) {
                                const size_t at = size_t(L.py) * size_t(P.width) + size_t(L.px);
                                P.next[at] = accumulate(P, div3(zero, float(P.raysPerPixel)), at);
                                if (IDS) P.primaryIds[at] = -1;
                                if (COUNT && P.raysPerPixel > 0) cnt[RT_CNT_PATHS] += P.raysPerPixel;
                                L.needPixel = true;
                            } else {
                                L.o = camPos;
                                L.d = primaryDirection(P, L.px, L.py, L.seed);
                                scratch[0] = __float_as_uint(L.d.x);
                                scratch[kThreads] = __float_as_uint(L.d.y);
                                scratch[2 * kThreads] = __float_as_uint(L.d.z);
                                L.fresh = true;
                                if (COUNT) {
                                    cnt[RT_CNT_PATHS]++;
                                    primBase[0] = cnt[RT_CNT_INTERNAL]; primBase[1] = cnt[RT_CNT_LEAVES];
                                    primBase[2] = cnt[RT_CNT_TRI_TESTS]; primBase[3] = cnt[RT_CNT_POPS];
                                }
                                beginRay<COUNT>(sc, L, cnt);
                            }
                        }
                    }
                }
            }
        } else if (__popc(mL) >= RT_VOTE_LEAF_LANES || mI == 0u) {
            // ------------------------------------------------------------------- LEAF section
            if (L.mode == MODE_LEAF) stepLeaf<COUNT>(sc, L, stack, stride, cnt);
        } else {
            // --------------------------------------------------------------- INTERNAL section
            if (L.mode == MODE_INTERNAL) stepInternal<COUNT>(sc, L, stack, stride, cnt, top, topCount);
        }
    }

    if constexpr (COUNT) {
#pragma unroll
        for (int k = 0; k < RT_CNT_COUNT; k++) {
            unsigned long long v = cnt[k];
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
            if (lane == 0 && v) atomicAdd(P.counters + k, v);
        }
    }
}

}  // namespace vote

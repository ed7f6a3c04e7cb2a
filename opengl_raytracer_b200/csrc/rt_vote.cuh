// rt_vote.cuh — variant PERSISTENT: persistent warps, warp-vote scheduled state machine.
//
// Why: in the one-thread-per-pixel kernel every lane walks its own sample/bounce/traversal loops, so a warp
// pays max-over-lanes at every level, and the rare, expensive blocks (leaf triangle tests, shading) are executed
// for two to five lanes at a time (profiles/r01_ncu_notes.md).  Here each lane is a small state machine
//
//     INTERNAL (at an internal node)   LEAF (at a leaf)   SHADE (traversal finished / needs a pixel)   DONE
//
// and every iteration of the warp's loop executes exactly ONE section, chosen by ballot: lanes that reach a
// leaf or finish a ray wait until enough lanes are in the same state (or nobody else can make progress), so the
// expensive sections run densely populated and the hot internal-node section is never interleaved with them.
// All branches that select code are warp-uniform, so lanes stay converged.  Lanes fetch their next pixel from a
// global counter (warp-aggregated atomic), which also removes the sky-versus-mesh imbalance.
//
// Second version (ncu capture G): the INTERNAL section runs a burst of kBurst node steps per vote (the ballots
// were 10 % of the instructions); a step pops at most once, branch-free (the pop loop ran at 4 of 32 lanes);
// everything about a pixel-frame's first bounce that does not depend on the sample — surface point, normal,
// material, normalised mirror direction, next origin, emitted light, throughput — is computed once when the
// primary ray lands and parked in shared memory (shading was 25 % of the instructions at 4-5 lanes).
//
// Per-ray semantics are exactly those of traverse()/pixelFrame() in rt_device.cuh / rt_kernels.cu — the
// visiting order inside a ray never changes, only which lanes run when; hoisting a computation out of the
// sample loop does not change its result.
#pragma once
#include "rt_shade.cuh"

namespace vote {

#ifndef RT_VOTE_THREADS
#define RT_VOTE_THREADS 128
#endif
constexpr int kThreads = RT_VOTE_THREADS;
#ifndef RT_VOTE_BURST
#define RT_VOTE_BURST 4
#endif
constexpr int kBurst = RT_VOTE_BURST;
// per lane: normal0 (3), k0, s1 (3), origin0 (3), incoming0 (3), rayColor0 (3), stop0
constexpr int kScratchWords = 17;
constexpr int MODE_INTERNAL = 0, MODE_LEAF = 1, MODE_SHADE = 2, MODE_DONE = 3;

inline size_t smemBytes(int stackDepth) {
    return size_t(stackDepth) * kThreads * sizeof(uint2) + size_t(kScratchWords) * kThreads * sizeof(uint32_t);
}

struct Lane {
    // traversal
    f3 o, d, inv;
    RayHit best;
    int32_t cur;      // RT_REF_NONE while INTERNAL: "the last pop was culled, pop again"
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

// One internal-node step (fragment.glsl:281-304), see traverse() for the order-preserving details.  The step
// pops at most ONE entry: closestT only changes in the LEAF section, so a popped entry fails the cull only if a
// leaf improved closestT after it was pushed; then cur = RT_REF_NONE and the lane pops again on its next step.
template <bool COUNT>
RT_DEV void stepInternal(const DeviceScene& sc, Lane& L, uint2* stack, uint32_t stride, uint32_t* cnt) {
    bool goNear = false;
    if (L.cur != RT_REF_NONE) {
        float4 q0, q1, q2;
        int2 refs;
        // (prefetching both children's records here — prefetch.global.L1 — was measured: dragon 27.2 -> 30.6 ms, box 155.7 ->
        // 402.8 ms; the L1 data pipe is the ceiling, not its latency; profiles/r02_variants.md)
        loadNode(sc.nodes + L.cur, q0, q1, q2, refs);
        bool hitA, hitB;
        const float tA = slabTest(L.o, L.inv, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, hitA);
        const float tB = slabTest(L.o, L.inv, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, hitB);
        if (COUNT) { cnt[RT_CNT_INTERNAL]++; cnt[RT_CNT_POPS] += (hitA ? 1u : 0u) + (hitB ? 1u : 0u); }
        const bool aFirst = hitA && (!hitB || tA < tB);
        const int32_t nearRef = aFirst ? refs.x : refs.y, farRef = aFirst ? refs.y : refs.x;
        const float nearT = aFirst ? tA : tB, farT = aFirst ? tB : tA;
        goNear = (hitA || hitB) && !(nearT > L.best.t);
        const bool pushFar = hitA && hitB && !(farT > L.best.t);
        if (pushFar) {
            stack[L.sp * stride] = make_uint2(uint32_t(farRef), __float_as_uint(farT));
            L.sp++;
        }
        L.cur = nearRef;
    }
    const bool havePop = !goNear && L.sp > 0;
    uint2 e = make_uint2(uint32_t(RT_REF_NONE), 0u);
    if (havePop) { L.sp--; e = stack[L.sp * stride]; }
    if (!goNear) L.cur = (__uint_as_float(e.y) > L.best.t) ? RT_REF_NONE : int32_t(e.x);
    L.mode = (goNear || havePop) ? (L.cur >= 0 ? MODE_INTERNAL : MODE_LEAF) : MODE_SHADE;
}

// One leaf (fragment.glsl:266-279): every triangle in slot order, first found wins ties; then pop.
template <bool COUNT, bool DEFER = false>      // DEFER: see triangleTest (rt_device.cuh)
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
        if (triangleTest<DEFER>(sc.tris + first + k, L.o, L.d, L.best.t, t, u, v)) {
            L.best.t = t; L.best.u = u; L.best.v = v; L.best.slot = int32_t(first + k);
        }
    }
    popNext(L, stack, stride);
}

// fragment.glsl:234-240 with the specular half precomputed: s1 = normalize(specular) * t, k0 = 1 - t, t = 1 - roughness.
RT_DEV f3 lerpDirPre(f3 diffuse, float k0, f3 s1) {
    const f3 zero = mk3(0.f, 0.f, 0.f);
    const f3 d0 = length3(diffuse) > 0.0f ? normalize3(diffuse) : zero;
    return normalize3(add3(scale3(d0, k0), s1));
}

// Scheduling thresholds: a section runs when at least this many lanes wait for it, or when no lane can do
// anything cheaper.  Defaults tuned on B200 (profiles/r01_variants.md); PassParams.tune overrides (bench tooling).
#ifndef RT_VOTE_LEAF_LANES
#define RT_VOTE_LEAF_LANES 8
#endif
#ifndef RT_VOTE_SHADE_LANES
#define RT_VOTE_SHADE_LANES 12
#endif

// The samples variant's device-side decision (rt_spec.cuh), made from what stage 1 queued: speculate when few pixels hit
// the scene (sky is visible: paths are short and most F(k) are used) and the table has room.  tune[2] forces it on / off.
// With PassParams::specGate the same launch sequence carries BOTH branches — the rounds of rt_spec.cuh, which return at
// once when the answer is no, and this file's kernel over the whole pass, which returns at once when it is yes.
RT_DEV bool samplesSpeculate(const PassParams& P, unsigned int queued) {
    if (queued > P.specCapacity) return false;
    if (P.tune[2] == 1) return true;
    if (P.tune[2] == 2) return false;
    return 4ull * queued <= (unsigned long long)P.ownedW * (unsigned long long)P.ownedH;
}

#ifndef RT_VOTE_MIN_CTAS
#define RT_VOTE_MIN_CTAS 6
#endif
template <bool COUNT, bool IDS, bool DEFER = false>
__global__ void __launch_bounds__(kThreads, COUNT ? 1 : RT_VOTE_MIN_CTAS) rtVoteKernel(const __grid_constant__ PassParams P) {
    extern __shared__ uint2 smem[];
    if (P.specGate != 0 && samplesSpeculate(P, *P.queueCount)) return;    // the samples rounds did this pass
    const DeviceScene& sc = P.scene;
    const uint32_t lane = threadIdx.x & 31u;
    uint2* const stack = smem + threadIdx.x;
    const uint32_t stride = kThreads;
    float* const scratch = reinterpret_cast<float*>(smem + size_t(sc.stackDepth) * kThreads) + threadIdx.x;
    const int leafLanes = P.tune[0] > 0 ? P.tune[0] : RT_VOTE_LEAF_LANES;
    const int shadeLanes = P.tune[1] > 0 ? P.tune[1] : RT_VOTE_SHADE_LANES;

    const int tilesX = (P.ownedW + 7) >> 3, tilesY = (P.ownedH + 3) >> 2;
    const unsigned int totalIdx = (unsigned int)tilesX * (unsigned int)tilesY * 32u;
    const f3 camPos = ld3(P.camPos);
    const f3 zero = mk3(0.f, 0.f, 0.f), one = mk3(1.f, 1.f, 1.f);
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

        // thresholds shrink with the number of lanes that still have work: a sparsely populated warp (the tail of a
        // frame, or a rank's thin share of a tile-sharded frame) must not make its few rays wait for each other —
        // that wait is pure latency on the serial sample chain of a pixel (measured: 12 ms floor at 1/8 frame)
        const int active = __popc(mI | mL | mS);
        const int shadeT = min(shadeLanes, max(1, (active * 3) >> 3));
        const int leafT = min(leafLanes, max(1, active >> 2));
        if (__popc(mS) >= shadeT || (mI | mL) == 0u) {
            // ------------------------------------------------------------------ SHADE section
            if (L.mode == MODE_SHADE && !L.needPixel) {
                bool pathEnds = false, pixelEnds = false, startSample = false;
                if (L.fresh) {
                    // the pixel's primary ray just landed (every sample of the pixel-frame shares it, :361-362)
                    L.fresh = false;
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
                    } else {
                        // once per pixel-frame: the part of bounce 0 that does not depend on the sample (:319-334)
                        const SurfaceHit hit0 = resolveHit(sc, L.best, camPos, L.d);
                        const f3 spec0 = reflect3(L.d, hit0.normal);
                        const float t0 = sub(1.0f, hit0.roughness);
                        const f3 s1 = scale3(length3(spec0) > 0.0f ? normalize3(spec0) : zero, t0);
                        const f3 origin0 = add3(hit0.point, scale3(hit0.normal, 1e-4f));
                        f3 emitted0 = scale3(hit0.emissionColor, hit0.emission);
                        if (lambertian) emitted0 = scale3(emitted0, 2.0f);
                        const f3 incoming0 = add3(zero, mul3(emitted0, one));
                        const f3 rayColor0 = mul3(one, hit0.color);
                        scratch[0 * kThreads] = hit0.normal.x; scratch[1 * kThreads] = hit0.normal.y; scratch[2 * kThreads] = hit0.normal.z;
                        scratch[3 * kThreads] = sub(1.0f, t0);
                        scratch[4 * kThreads] = s1.x; scratch[5 * kThreads] = s1.y; scratch[6 * kThreads] = s1.z;
                        scratch[7 * kThreads] = origin0.x; scratch[8 * kThreads] = origin0.y; scratch[9 * kThreads] = origin0.z;
                        scratch[10 * kThreads] = incoming0.x; scratch[11 * kThreads] = incoming0.y; scratch[12 * kThreads] = incoming0.z;
                        scratch[13 * kThreads] = rayColor0.x; scratch[14 * kThreads] = rayColor0.y; scratch[15 * kThreads] = rayColor0.z;
                        scratch[16 * kThreads] = hit0.emission > 0.0f ? 1.0f : 0.0f;
                        startSample = true;
                    }
                } else {
                    // one iteration of raytrace()'s loop body (fragment.glsl:313-347) for the (non-primary) ray just traced
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
                }
                if (pathEnds) {
                    L.color = add3(L.color, L.incoming);   // trace(), :359
                    L.sample++;
                    if (L.sample >= P.raysPerPixel) pixelEnds = true;
                    else startSample = true;   // next sample: same primary ray, RNG state carried over (:358-363)
                }
                if (startSample) {
                    // bounce 0 of a sample: the cached primary hit
                    if (COUNT) {
                        cnt[RT_CNT_RAYS]++; cnt[RT_CNT_HITS]++;
                        if (L.sample > 0) {
                            cnt[RT_CNT_PATHS]++;
                            cnt[RT_CNT_INTERNAL] += primWork[0]; cnt[RT_CNT_LEAVES] += primWork[1];
                            cnt[RT_CNT_TRI_TESTS] += primWork[2]; cnt[RT_CNT_POPS] += primWork[3];
                        }
                    }
                    const f3 incoming0 = mk3(scratch[10 * kThreads], scratch[11 * kThreads], scratch[12 * kThreads]);
                    if (scratch[16 * kThreads] != 0.0f || P.nBounces <= 1) {
                        // the path ends at bounce 0 (emissive surface or no further bounce) for every remaining sample:
                        // each adds 0 + incoming0; the RNG draws of diffuse() never reach the output
                        const int remaining = P.raysPerPixel - L.sample;
                        for (int s = 0; s < remaining; s++) L.color = add3(L.color, incoming0);
                        if (COUNT) {
                            cnt[RT_CNT_RAYS] += remaining - 1; cnt[RT_CNT_HITS] += remaining - 1; cnt[RT_CNT_PATHS] += remaining - 1;
                            cnt[RT_CNT_INTERNAL] += primWork[0] * (remaining - 1); cnt[RT_CNT_LEAVES] += primWork[1] * (remaining - 1);
                            cnt[RT_CNT_TRI_TESTS] += primWork[2] * (remaining - 1); cnt[RT_CNT_POPS] += primWork[3] * (remaining - 1);
                        }
                        pixelEnds = true;
                    } else {
                        const f3 normal0 = mk3(scratch[0 * kThreads], scratch[1 * kThreads], scratch[2 * kThreads]);
                        const f3 s1 = mk3(scratch[4 * kThreads], scratch[5 * kThreads], scratch[6 * kThreads]);
                        const f3 dDiff0 = diffuseDir(normal0, L.seed, lambertian);
                        L.d = lerpDirPre(dDiff0, scratch[3 * kThreads], s1);
                        L.o = mk3(scratch[7 * kThreads], scratch[8 * kThreads], scratch[9 * kThreads]);
                        L.incoming = incoming0;
                        L.rayColor = mk3(scratch[13 * kThreads], scratch[14 * kThreads], scratch[15 * kThreads]);
                        L.bounce = 1;
                        beginRay<COUNT>(sc, L, cnt);
                    }
                }
                if (pixelEnds) {
                    const size_t at = size_t(L.py) * size_t(P.width) + size_t(L.px);
                    storePixel(P, at, accumulate(P, div3(L.color, float(P.raysPerPixel)), at));   // :365, :409-416
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
                            L.color = zero; L.incoming = zero; L.rayColor = one;
                            L.sample = 0; L.bounce = 0;
                            if (degenerate) {
                                const size_t at = size_t(L.py) * size_t(P.width) + size_t(L.px);
                                storePixel(P, at, accumulate(P, div3(zero, float(P.raysPerPixel)), at));
                                if (IDS) P.primaryIds[at] = -1;
                                if (COUNT && P.raysPerPixel > 0) cnt[RT_CNT_PATHS] += P.raysPerPixel;
                                L.needPixel = true;
                            } else {
                                L.o = camPos;
                                L.d = primaryDirection(P, L.px, L.py, L.seed);
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
        } else if (__popc(mL) >= leafT || mI == 0u) {
            // ------------------------------------------------------------------- LEAF section
            if (L.mode == MODE_LEAF) stepLeaf<COUNT, DEFER>(sc, L, stack, stride, cnt);
        } else {
            // ------------------------------------------- INTERNAL section: a burst of node steps per vote
#pragma unroll 1
            for (int s = 0; s < kBurst; s++)
                if (L.mode == MODE_INTERNAL) stepInternal<COUNT>(sc, L, stack, stride, cnt);
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

// rt_pool.cuh — variant POOL: persistent CTAs, an on-chip ray pool per lane, warp-vote scheduled stages.
//
// What the ncu captures of the two earlier variants said (profiles/r01_ncu_notes.md):
//   * one-thread-per-pixel kernel: 11.7 of 32 lanes active per executed instruction — a warp pays the longest
//     traversal of its 32 rays at every bounce, and the rare expensive blocks (leaf triangle tests, shading)
//     run for two or three lanes at a time;  L1 data-pipe wavefronts at 74 % of peak, because every lane's
//     64-byte node record is four separate 16-byte loads to its own cache line (1 wavefront per line);
//   * first warp-vote kernel: 16 of 32 lanes, but 40 % more thread-instructions — a lane that waits for its
//     leaf/shade stage idles, and the state machine was paid on every node step.
//
// This variant keeps K rays ("slots") per lane in SHARED MEMORY (227 KB per SM on B200 is what makes it fit):
// a lane whose ray waits for the leaf or shade stage simply steps its other ray.  Each iteration of the warp
// runs ONE stage, picked by ballot:
//     EXTEND  — a burst of S internal-node steps for the lane's first ray that is at an internal node whose record
//               is staged in shared memory; the default stage: it runs whenever no other stage has reached its
//               threshold (so this hot stage is the densely populated one);
//     DEEP    — the same step for rays at nodes that must come from global memory;
//     LEAF    — triangle tests of one waiting leaf per lane, then the pop;
//     SHADE   — finished rays: terminal events, hit shading, next ray / next sample / next pixel (work fetched
//               from a global counter with one warp-aggregated atomic).
// SHADE, LEAF and DEEP run when RT_POOL_TS / RT_POOL_TL / RT_POOL_TD lanes have a ray waiting for them (checked in
// that order), or when nothing else can run.  Measured outcome: profiles/r01_variants.md §5 — not what AUTO picks.
// The first `topCount` nodes of the tree (under breadth-first numbering, RT_NODE_ORDER=bfs, which this variant was built
// and measured with: its top levels — 78-89 % of all node visits with 512-1024 nodes, tools/node_hist.py) are staged in
// shared memory, where a divergent 16-byte read costs bank conflicts instead of one L1 wavefront per lane.  (Since round 2
// the default numbering is depth-first: the staged prefix is then the root's leftmost subtree — correct, less useful.)
//
// Per-ray semantics are exactly those of traverse()/pixelFrame() (rt_device.cuh / rt_kernels.cu): the visiting
// order inside a ray never changes, only which ray a lane advances when.
#pragma once
#include "rt_shade.cuh"

namespace pool {

// A slot's mode.  INTERNAL: at a node whose record is staged in shared memory; DEEP: at a node that must come
// from global memory (index >= topCount), or about to re-pop after a culled pop (cur == RT_REF_NONE).
constexpr int MODE_INTERNAL = 0, MODE_LEAF = 1, MODE_SHADE = 2, MODE_DONE = 3, MODE_DEEP = 4;
constexpr int kMaxDepth = 31;   // sp is kept in 5 bits
constexpr uint32_t FULL = 0xffffffffu;

// Stage thresholds (lanes of the warp that have a ray waiting for the stage).
#ifndef RT_POOL_TL
#define RT_POOL_TL 16
#endif
#ifndef RT_POOL_TS
#define RT_POOL_TS 12
#endif
#ifndef RT_POOL_TD
#define RT_POOL_TD 12
#endif

// Per-pixel path state of a slot.  Lives in local memory (indexed by the slot number at run time); it is only
// touched by the SHADE stage, once per ray.
struct PathState {
    f3 incoming, rayColor, color;
    f3 primDir;                 // the pixel-frame's primary direction (shared by all its samples, :361-362)
    float primT, primU, primV;  // ... and its closest hit, traced once
    int32_t primSlot;
    uint32_t seed;
    int32_t sample, bounce;
    int32_t px, py;
    int32_t fresh;              // the ray in flight is the primary ray of the pixel
    uint32_t work[4];           // COUNT only: internal / leaves / tri tests / pops of the ray in flight
    uint32_t primWork[4];       // COUNT only: ... of the primary ray (re-added per extra sample)
};

// Shared-memory layout for T threads x K slots, slot index si = k * T + tid (lanes -> consecutive addresses).
template <int T, int K>
struct Smem {
    float4* A;        // o.xyz, bestT
    float4* B;        // inv.xyz, (unused)
    float4* C;        // d.xyz, bestSlot (bits)
    float2* D;        // bestU, bestV
    int32_t* cur;     // node reference the ray is at
    uint2* stack;     // [depth][K*T]  (ref, tNear)
    float4* top;      // [3][topCount] child bounds of the first topCount nodes
    int2* topRefs;    // [topCount]
    static __host__ __device__ size_t slotBytes(int depth) { return size_t(K) * T * (16 * 3 + 8 + 4 + 8 * size_t(depth)); }
    static __host__ __device__ size_t bytes(int depth, int topCount) { return slotBytes(depth) + size_t(topCount) * 56; }
    __device__ Smem(unsigned char* base, int depth, int topCount) {
        A = reinterpret_cast<float4*>(base);
        B = A + K * T;
        C = B + K * T;
        top = C + K * T;
        D = reinterpret_cast<float2*>(top + 3 * size_t(topCount));
        stack = reinterpret_cast<uint2*>(D + K * T);
        topRefs = reinterpret_cast<int2*>(stack + size_t(depth) * K * T);
        cur = reinterpret_cast<int32_t*>(topRefs + topCount);
    }
};

RT_DEV int modeOf(uint32_t modes, int k) { return int((modes >> (8 * k)) & 7u); }
RT_DEV uint32_t spOf(uint32_t modes, int k) { return (modes >> (8 * k + 3)) & 31u; }
RT_DEV uint32_t withSlot(uint32_t modes, int k, int mode, uint32_t sp) {
    return (modes & ~(0xffu << (8 * k))) | ((uint32_t(mode) | (sp << 3)) << (8 * k));
}
RT_DEV int classify(int32_t cur, int topCount) {
    return cur < 0 ? MODE_LEAF : (cur >= topCount ? MODE_DEEP : MODE_INTERNAL);
}
RT_DEV void prefetchL1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
// lowest slot in `mode`, or -1
template <int K>
RT_DEV int firstIn(uint32_t modes, int mode) {
    int found = -1;
#pragma unroll
    for (int k = K - 1; k >= 0; k--)
        if (modeOf(modes, k) == mode) found = k;
    return found;
}

// Pop until an entry survives the cull against the current closestT (fragment.glsl:256-262).
RT_DEV void popNext(const uint2* stack, uint32_t stride, float bestT, uint32_t& sp, int32_t& cur, int& mode, int topCount) {
    bool found = false;
    while (sp > 0 && !found) {
        sp--;
        const uint2 e = stack[sp * stride];
        if (!(__uint_as_float(e.y) > bestT)) { cur = int32_t(e.x); found = true; }
    }
    mode = found ? classify(cur, topCount) : MODE_SHADE;
}

// The arithmetic of one internal-node step (fragment.glsl:281-304) on a fetched record; see traverse() in
// rt_device.cuh for why this preserves the shader's visiting order.
struct StepOut {
    bool goNear, pushFar;
    int32_t nearRef, farRef;
    float farT;
    uint32_t hits;
};
RT_DEV StepOut nodeMath(f3 o, f3 inv, float bestT, float4 q0, float4 q1, float4 q2, int2 refs) {
    bool hitA, hitB;
    const float tA = slabTest(o, inv, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, hitA);
    const float tB = slabTest(o, inv, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, hitB);
    const bool aFirst = hitA && (!hitB || tA < tB);
    StepOut r;
    r.nearRef = aFirst ? refs.x : refs.y;
    r.farRef = aFirst ? refs.y : refs.x;
    const float nearT = aFirst ? tA : tB;
    r.farT = aFirst ? tB : tA;
    r.goNear = (hitA || hitB) && !(nearT > bestT);
    r.pushFar = hitA && hitB && !(r.farT > bestT);
    r.hits = (hitA ? 1u : 0u) + (hitB ? 1u : 0u);
    return r;
}

// Start the traversal of (o, d) in slot si: root slab test (fragment.glsl:253-262).  Returns the slot's mode.
template <bool COUNT, int T, int K>
RT_DEV int beginRay(const DeviceScene& sc, const Smem<T, K>& sm, int si, f3 o, f3 d, uint32_t* cnt, PathState& s, int topCount) {
    const f3 inv = mk3(__frcp_rn(d.x), __frcp_rn(d.y), __frcp_rn(d.z));
    sm.A[si] = make_float4(o.x, o.y, o.z, 1e30f);
    sm.B[si] = make_float4(inv.x, inv.y, inv.z, 0.f);
    sm.C[si] = make_float4(d.x, d.y, d.z, __int_as_float(-1));
    sm.cur[si] = sc.rootRef;
    bool running = sc.rootRef != RT_REF_NONE;   // empty scene: the shader's root index fails its bounds check (:259)
    if (COUNT) {
        cnt[RT_CNT_POPS]++; cnt[RT_CNT_RAYS_TRACED]++;
        s.work[0] = 0; s.work[1] = 0; s.work[2] = 0; s.work[3] = 1;
    }
    if (running) {
        bool hitRoot;
        const float tRoot = slabTest(o, inv, sc.rootMin[0], sc.rootMin[1], sc.rootMin[2], sc.rootMax[0],
                                     sc.rootMax[1], sc.rootMax[2], hitRoot);
        running = hitRoot && !(tRoot > 1e30f);
    }
    return running ? classify(sc.rootRef, topCount) : MODE_SHADE;
}

template <int T, int K, int S, bool COUNT, bool IDS>
__global__ void __launch_bounds__(T, 1) rtPoolKernel(const __grid_constant__ PassParams P) {
    extern __shared__ __align__(16) unsigned char poolSmem[];
    const DeviceScene& sc = P.scene;
    const int tid = threadIdx.x;
    const uint32_t lane = uint32_t(tid) & 31u;
    const int topCount = P.topCount;
    const Smem<T, K> sm(poolSmem, sc.stackDepth, topCount);
    constexpr uint32_t STRIDE = K * T;

    // stage the first nodes of the array (its top levels under RT_NODE_ORDER=bfs, the numbering this variant was built
    // for; with the default depth-first numbering: the root and the leftmost subtree — still correct, less useful)
    for (int i = tid; i < topCount * 3; i += T) {
        const int node = i / 3, part = i - node * 3;
        sm.top[part * topCount + node] = __ldg(reinterpret_cast<const float4*>(sc.nodes + node) + part);
    }
    for (int i = tid; i < topCount; i += T)
        sm.topRefs[i] = __ldg(reinterpret_cast<const int2*>(reinterpret_cast<const float4*>(sc.nodes + i) + 3));
    __syncthreads();

    const int tilesX = (P.ownedW + 7) >> 3, tilesY = (P.ownedH + 3) >> 2;
    const unsigned int totalIdx = (unsigned int)tilesX * (unsigned int)tilesY * 32u;
    const f3 camPos = ld3(P.camPos);
    const f3 zero = mk3(0.f, 0.f, 0.f);
    const f3 sky = scale3(mk3(0.1f, 0.6f, 0.92f), P.skyBrightness);   // fragment.glsl:242-244, :388
    const bool lambertian = P.lambertian != 0;
    const bool degenerate = P.nBounces <= 0 || P.raysPerPixel <= 0;   // no ray is ever cast

    uint32_t cnt[RT_CNT_COUNT];
#pragma unroll
    for (int k = 0; k < RT_CNT_COUNT; k++) cnt[k] = 0;

    PathState ps[K];
    // every slot starts in SHADE mode with "needs a pixel" (sp field = 1 while in SHADE mode)
    uint32_t modes = 0;
#pragma unroll
    for (int k = 0; k < K; k++) modes = withSlot(modes, k, MODE_SHADE, 1u);

    for (;;) {
        const int kI = firstIn<K>(modes, MODE_INTERNAL);
        const int kD = firstIn<K>(modes, MODE_DEEP);
        const int kL = firstIn<K>(modes, MODE_LEAF);
        const int kS = firstIn<K>(modes, MODE_SHADE);
        const uint32_t mI = __ballot_sync(FULL, kI >= 0);
        const uint32_t mD = __ballot_sync(FULL, kD >= 0);
        const uint32_t mL = __ballot_sync(FULL, kL >= 0);
        const uint32_t mS = __ballot_sync(FULL, kS >= 0);
        if ((mI | mD | mL | mS) == 0u) break;

        // stage choice (warp-uniform): the rarer, slower stages run when enough lanes wait for them, the hot
        // EXTEND stage otherwise; when nothing can extend, whichever of the others has most lanes
        const int nI = __popc(mI), nD = __popc(mD), nL = __popc(mL), nS = __popc(mS);
        int stage;
        if (nS >= RT_POOL_TS) stage = MODE_SHADE;
        else if (nL >= RT_POOL_TL) stage = MODE_LEAF;
        else if (nD >= RT_POOL_TD) stage = MODE_DEEP;
        else if (nI > 0) stage = MODE_INTERNAL;
        else stage = (nL >= nS && nL >= nD) ? MODE_LEAF : (nD >= nS ? MODE_DEEP : MODE_SHADE);

        if (stage == MODE_INTERNAL) {
            // --------------------------------------------------- EXTEND stage: node records from shared memory only
            if (kI >= 0) {
                const int si = kI * T + tid;
                const float4 a = sm.A[si], b = sm.B[si];
                const f3 o = mk3(a.x, a.y, a.z), inv = mk3(b.x, b.y, b.z);
                const float bestT = a.w;
                int32_t cur = sm.cur[si];
                uint32_t sp = spOf(modes, kI);
                int mode = MODE_INTERNAL;
                uint2* const stack = sm.stack + si;
                uint32_t wInternal = 0, wPops = 0;
#pragma unroll 1
                for (int s = 0; s < S; s++) {
                    if (mode == MODE_INTERNAL) {
                        const StepOut r = nodeMath(o, inv, bestT, sm.top[cur], sm.top[topCount + cur],
                                                   sm.top[2 * topCount + cur], sm.topRefs[cur]);
                        if (COUNT) { wInternal++; wPops += r.hits; }
                        if (r.pushFar) {
                            stack[sp * STRIDE] = make_uint2(uint32_t(r.farRef), __float_as_uint(r.farT));
                            sp++;
                        }
                        // next node: the near child, else ONE pop (fragment.glsl:256-262).  closestT is constant inside
                        // a burst, so a popped entry is culled only if a leaf improved closestT after it was pushed; the
                        // rare culled pop parks the slot in DEEP mode (cur = RT_REF_NONE), whose stage pops in a loop.
                        const bool havePop = !r.goNear && sp > 0;
                        uint2 e = make_uint2(uint32_t(RT_REF_NONE), 0u);
                        if (havePop) { sp--; e = stack[sp * STRIDE]; }
                        const bool culled = __uint_as_float(e.y) > bestT;
                        cur = r.goNear ? r.nearRef : (culled ? RT_REF_NONE : int32_t(e.x));
                        mode = (r.goNear || havePop) ? classify(cur, topCount) : MODE_SHADE;
                    }
                }
                if (mode == MODE_DEEP && cur != RT_REF_NONE) prefetchL1(sc.nodes + cur);
                sm.cur[si] = cur;
                modes = withSlot(modes, kI, mode, mode == MODE_SHADE ? 0u : sp);
                if (COUNT) {
                    cnt[RT_CNT_INTERNAL] += wInternal; cnt[RT_CNT_POPS] += wPops;
                    ps[kI].work[0] += wInternal; ps[kI].work[3] += wPops;
                }
            }
        } else if (stage == MODE_DEEP) {
            // ------------------------------------------- DEEP stage: node records from global memory (L1/L2), loop pops
            if (kD >= 0) {
                const int si = kD * T + tid;
                const float4 a = sm.A[si], b = sm.B[si];
                const f3 o = mk3(a.x, a.y, a.z), inv = mk3(b.x, b.y, b.z);
                const float bestT = a.w;
                int32_t cur = sm.cur[si];
                uint32_t sp = spOf(modes, kD);
                int mode = MODE_DEEP;
                uint2* const stack = sm.stack + si;
                uint32_t wInternal = 0, wPops = 0;
#pragma unroll 1
                for (int s = 0; s < S; s++) {
                    if (mode == MODE_DEEP) {
                        bool goNear = false;
                        if (cur != RT_REF_NONE) {
                            float4 q0, q1, q2;
                            int2 refs;
                            loadNode(sc.nodes + cur, q0, q1, q2, refs);
                            const StepOut r = nodeMath(o, inv, bestT, q0, q1, q2, refs);
                            if (COUNT) { wInternal++; wPops += r.hits; }
                            if (r.pushFar) {
                                stack[sp * STRIDE] = make_uint2(uint32_t(r.farRef), __float_as_uint(r.farT));
                                sp++;
                            }
                            goNear = r.goNear;
                            cur = r.nearRef;
                        }
                        if (goNear) mode = classify(cur, topCount);
                        else popNext(stack, STRIDE, bestT, sp, cur, mode, topCount);
                    }
                }
                if (mode == MODE_DEEP) prefetchL1(sc.nodes + cur);
                sm.cur[si] = cur;
                modes = withSlot(modes, kD, mode, mode == MODE_SHADE ? 0u : sp);
                if (COUNT) {
                    cnt[RT_CNT_INTERNAL] += wInternal; cnt[RT_CNT_POPS] += wPops;
                    ps[kD].work[0] += wInternal; ps[kD].work[3] += wPops;
                }
            }
        } else if (stage == MODE_LEAF) {
            // -------------------------------------------------------------------------- LEAF stage
            if (kL >= 0) {
                const int si = kL * T + tid;
                const float4 a = sm.A[si], c = sm.C[si];
                const f3 o = mk3(a.x, a.y, a.z), d = mk3(c.x, c.y, c.z);
                float bestT = a.w;
                int32_t cur = sm.cur[si];
                uint32_t sp = spOf(modes, kL);
                // every triangle in slot order, first found wins ties (fragment.glsl:266-279)
                const uint32_t code = uint32_t(~cur);
                uint32_t count = code & ((1u << RT_LEAF_BITS) - 1u);
                uint32_t first = code >> RT_LEAF_BITS;
                if (count == RT_LEAF_BIG) {
                    const uint2 big = __ldg(sc.bigLeaves + first);
                    first = big.x; count = big.y;
                }
                if (COUNT) {
                    cnt[RT_CNT_LEAVES]++; cnt[RT_CNT_TRI_TESTS] += count;
                    ps[kL].work[1]++; ps[kL].work[2] += count;
                }
                int32_t bestSlot = -1;
                float bestU = 0.f, bestV = 0.f;
                for (uint32_t k = 0; k < count; k++) {
                    float t, u, v;
                    if (triangleTest<false>(sc.tris + first + k, o, d, bestT, t, u, v)) {
                        bestT = t; bestU = u; bestV = v; bestSlot = int32_t(first + k);
                    }
                }
                if (bestSlot >= 0) {
                    sm.A[si].w = bestT;
                    sm.C[si].w = __int_as_float(bestSlot);
                    sm.D[si] = make_float2(bestU, bestV);
                }
                int mode;
                popNext(sm.stack + si, STRIDE, bestT, sp, cur, mode, topCount);
                if (mode == MODE_DEEP) prefetchL1(sc.nodes + cur);
                sm.cur[si] = cur;
                modes = withSlot(modes, kL, mode, mode == MODE_SHADE ? 0u : sp);
            }
        } else {
            // ------------------------------------------------------------------------- SHADE stage
            {
                const bool mine = kS >= 0;
                const int kk = mine ? kS : 0;
                const int si = kk * T + tid;
                bool need = mine && spOf(modes, kk) != 0u;      // slot waits for a pixel
                int mode = MODE_SHADE;
                PathState s;
                if (mine) s = ps[kk];
                if (mine && !need) {
                    const float4 a = sm.A[si], c = sm.C[si];
                    const float2 uv = sm.D[si];
                    f3 o = mk3(a.x, a.y, a.z), d = mk3(c.x, c.y, c.z);
                    RayHit best;
                    best.t = a.w; best.slot = __float_as_int(c.w); best.u = uv.x; best.v = uv.y;
                    bool pathEnds = false, pixelEnds = false, shadeHit = false;
                    if (s.fresh) {
                        // the pixel's primary ray just finished: keep its hit for the other samples (:361-362)
                        s.fresh = 0;
                        s.primSlot = best.slot; s.primT = best.t; s.primU = best.u; s.primV = best.v;
                        if (COUNT) {
#pragma unroll
                            for (int j = 0; j < 4; j++) s.primWork[j] = s.work[j];
                        }
                        if (IDS) {
                            P.primaryIds[size_t(s.py) * size_t(P.width) + size_t(s.px)] =
                                best.slot >= 0 ? int32_t(__ldg(sc.triId + best.slot)) : -1;
                        }
                        if (best.slot < 0) {
                            // primary miss: every sample of the pixel-frame is "incoming = 0 + sky" (:342-346)
                            const f3 inc = add3(zero, sky);
                            for (int j = 0; j < P.raysPerPixel; j++) s.color = add3(s.color, inc);
                            if (COUNT) {
                                cnt[RT_CNT_RAYS] += P.raysPerPixel; cnt[RT_CNT_PATHS] += P.raysPerPixel - 1;
                                cnt[RT_CNT_INTERNAL] += s.primWork[0] * (P.raysPerPixel - 1);
                                cnt[RT_CNT_LEAVES] += s.primWork[1] * (P.raysPerPixel - 1);
                                cnt[RT_CNT_TRI_TESTS] += s.primWork[2] * (P.raysPerPixel - 1);
                                cnt[RT_CNT_POPS] += s.primWork[3] * (P.raysPerPixel - 1);
                            }
                            pixelEnds = true;
                        }
                    }
                    if (!pixelEnds) {
                        if (best.slot < 0) {
                            if (COUNT) cnt[RT_CNT_RAYS]++;
                            s.incoming = add3(s.incoming, sky);   // un-attenuated, as the shader does (:343)
                            pathEnds = true;
                        } else {
                            shadeHit = true;
                        }
                    }
                    // up to two rounds: a path that ends on a miss hands over to the next sample, whose first ray is
                    // the cached primary hit — shade it right away
#pragma unroll 1
                    for (int round = 0; round < 2; round++) {
                        if (pathEnds) {
                            pathEnds = false;
                            s.color = add3(s.color, s.incoming);   // trace(), :359
                            s.sample++;
                            if (s.sample >= P.raysPerPixel) {
                                pixelEnds = true;
                            } else {
                                // next sample: same primary ray, RNG state carried over (:358-363)
                                s.incoming = zero; s.rayColor = mk3(1.f, 1.f, 1.f); s.bounce = 0;
                                o = camPos; d = s.primDir;
                                best.slot = s.primSlot; best.t = s.primT; best.u = s.primU; best.v = s.primV;
                                shadeHit = round == 0;
                                if (round != 0) {
                                    // shaded next time: park the cached primary hit in the slot
                                    sm.A[si] = make_float4(o.x, o.y, o.z, best.t);
                                    sm.C[si] = make_float4(d.x, d.y, d.z, __int_as_float(best.slot));
                                    sm.D[si] = make_float2(best.u, best.v);
                                }
                                if (COUNT) {
                                    cnt[RT_CNT_PATHS]++;
                                    cnt[RT_CNT_INTERNAL] += s.primWork[0]; cnt[RT_CNT_LEAVES] += s.primWork[1];
                                    cnt[RT_CNT_TRI_TESTS] += s.primWork[2]; cnt[RT_CNT_POPS] += s.primWork[3];
                                }
                            }
                        }
                        if (shadeHit) {
                            // one iteration of raytrace()'s loop body (fragment.glsl:313-340) for a ray that hit
                            shadeHit = false;
                            if (COUNT) { cnt[RT_CNT_RAYS]++; cnt[RT_CNT_HITS]++; }
                            const SurfaceHit hit = resolveHit(sc, best, o, d);
                            const f3 dDiff = diffuseDir(hit.normal, s.seed, lambertian);
                            const f3 dSpec = reflect3(d, hit.normal);
                            d = lerpDir(dDiff, dSpec, hit.roughness);
                            o = add3(hit.point, scale3(hit.normal, 1e-4f));
                            f3 emitted = scale3(hit.emissionColor, hit.emission);
                            if (lambertian) emitted = scale3(emitted, 2.0f);
                            s.incoming = add3(s.incoming, mul3(emitted, s.rayColor));
                            s.rayColor = mul3(s.rayColor, hit.color);
                            s.bounce++;
                            pathEnds = hit.emission > 0.0f || s.bounce >= P.nBounces;
                            if (!pathEnds) mode = beginRay<COUNT, T, K>(sc, sm, si, o, d, cnt, s, topCount);
                        }
                    }
                    if (pixelEnds) {
                        const size_t at = size_t(s.py) * size_t(P.width) + size_t(s.px);
                        storePixel(P, at, accumulate(P, div3(s.color, float(P.raysPerPixel)), at));   // :365, :409-416
                        need = true;
                    }
                }
                // ---- work fetch: slots without a pixel claim the next indices (one atomic per warp and round)
                for (;;) {
                    const uint32_t want = __ballot_sync(FULL, need);
                    if (want == 0u) break;
                    unsigned int base = 0;
                    const int leader = __ffs(want) - 1;
                    if (int(lane) == leader) base = atomicAdd(P.workCounter, (unsigned int)__popc(want));
                    base = __shfl_sync(FULL, base, leader);
                    if (need) {
                        const unsigned int idx = base + (unsigned int)__popc(want & ((1u << lane) - 1u));
                        if (idx >= totalIdx) {
                            mode = MODE_DONE; need = false;
                        } else {
                            const unsigned int tile = idx >> 5, within = idx & 31u;
                            const int ox = int(tile % (unsigned int)tilesX) * 8 + int(within & 7u);
                            const int oy = int(tile / (unsigned int)tilesX) * 4 + int(within >> 3);
                            if (ox < P.ownedW && oy < P.ownedH) {
                                s.px = ownedToPixel(ox, P.tileSizeX, P.numTilesX, P.tileX);
                                s.py = ownedToPixel(oy, P.tileSizeY, P.numTilesY, P.tileY);
                                s.color = zero; s.incoming = zero; s.rayColor = mk3(1.f, 1.f, 1.f);
                                s.sample = 0; s.bounce = 0;
                                if (degenerate) {
                                    const size_t at = size_t(s.py) * size_t(P.width) + size_t(s.px);
                                    storePixel(P, at, accumulate(P, div3(zero, float(P.raysPerPixel)), at));
                                    if (IDS) P.primaryIds[at] = -1;
                                    if (COUNT && P.raysPerPixel > 0) cnt[RT_CNT_PATHS] += P.raysPerPixel;
                                } else {
                                    need = false;
                                    s.primDir = primaryDirection(P, s.px, s.py, s.seed);
                                    s.fresh = 1;
                                    if (COUNT) cnt[RT_CNT_PATHS]++;
                                    mode = beginRay<COUNT, T, K>(sc, sm, si, camPos, s.primDir, cnt, s, topCount);
                                }
                            }
                        }
                    }
                }
                if (mine) {
                    ps[kk] = s;
                    // SHADE mode with sp = 0: a finished ray (or a parked primary hit) waits; need never survives here
                    modes = withSlot(modes, kk, mode, 0u);
                }
            }
        }
    }

    if constexpr (COUNT) {
#pragma unroll
        for (int k = 0; k < RT_CNT_COUNT; k++) {
            unsigned long long v = cnt[k];
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(FULL, v, off);
            if (lane == 0 && v) atomicAdd(P.counters + k, v);
        }
    }
}

}  // namespace pool

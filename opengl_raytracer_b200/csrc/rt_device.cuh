// rt_device.cuh — device-side arithmetic of the reference's shaders/fragment.glsl, restated for sm_100a.
//
// Numeric contract (identical to oracle/rt_oracle.cpp, which the parity tests
// compare against bit for bit): IEEE binary32, round-to-nearest-even, NO contraction — every product
// and sum below is an explicit __fmul_rn / __fadd_rn so the result does not depend on -fmad; division
// and square root are the correctly rounded __fdiv_rn / __frcp_rn / __fsqrt_rn (1/x correctly rounded is
// unique, so __frcp_rn(x) == 1.0f / x); min/max are fminf/fmaxf (FMNMX: a NaN operand is ignored).
// dot = (x*x' + y*y') + z*z'; cross(a,b) = (a.y*b.z - b.y*a.z, a.z*b.x - b.z*a.x, a.x*b.y - b.x*a.y).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "rt_types.h"

#define RT_DEV __device__ __forceinline__

struct f3 {
    float x, y, z;
};

RT_DEV f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
RT_DEV float mul(float a, float b) { return __fmul_rn(a, b); }
RT_DEV float add(float a, float b) { return __fadd_rn(a, b); }
RT_DEV float sub(float a, float b) { return __fsub_rn(a, b); }
RT_DEV f3 add3(f3 a, f3 b) { return mk3(add(a.x, b.x), add(a.y, b.y), add(a.z, b.z)); }
RT_DEV f3 sub3(f3 a, f3 b) { return mk3(sub(a.x, b.x), sub(a.y, b.y), sub(a.z, b.z)); }
RT_DEV f3 mul3(f3 a, f3 b) { return mk3(mul(a.x, b.x), mul(a.y, b.y), mul(a.z, b.z)); }
RT_DEV f3 scale3(f3 a, float s) { return mk3(mul(a.x, s), mul(a.y, s), mul(a.z, s)); }
RT_DEV f3 neg3(f3 a) { return mk3(-a.x, -a.y, -a.z); }
RT_DEV float dot3(f3 a, f3 b) { return add(add(mul(a.x, b.x), mul(a.y, b.y)), mul(a.z, b.z)); }
RT_DEV f3 cross3(f3 a, f3 b) {
    return mk3(sub(mul(a.y, b.z), mul(b.y, a.z)), sub(mul(a.z, b.x), mul(b.z, a.x)),
               sub(mul(a.x, b.y), mul(b.x, a.y)));
}
RT_DEV float length3(f3 v) { return __fsqrt_rn(dot3(v, v)); }
RT_DEV f3 div3(f3 a, float s) { return mk3(__fdiv_rn(a.x, s), __fdiv_rn(a.y, s), __fdiv_rn(a.z, s)); }
RT_DEV f3 normalize3(f3 v) { return div3(v, length3(v)); }
RT_DEV f3 ld3(const float* p) { return mk3(p[0], p[1], p[2]); }

// 32 bytes through the read-only path in ONE instruction (LDG.E.256.CONSTANT, sm_100+).  Not volatile: scene records
// never change while a pass runs, so the compiler may schedule, hoist or drop the load like any __ldg.
#ifndef RT_LDG256
#define RT_LDG256 1
#endif
RT_DEV void ld256(const void* p, float4& lo, float4& hi) {
#if RT_LDG256
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(lo.x), "=f"(lo.y), "=f"(lo.z), "=f"(lo.w), "=f"(hi.x), "=f"(hi.y), "=f"(hi.z), "=f"(hi.w)
        : "l"(p));
#else
    lo = __ldg(reinterpret_cast<const float4*>(p));
    hi = __ldg(reinterpret_cast<const float4*>(p) + 1);
#endif
}
// One NodeRec: the twelve child bounds in q0..q2 (rt_types.h order) and the two child references.
RT_DEV void loadNode(const NodeRec* node, float4& q0, float4& q1, float4& q2, int2& refs) {
    float4 q3;
    ld256(node, q0, q1);
    ld256(reinterpret_cast<const char*>(node) + 32, q2, q3);
    refs = make_int2(__float_as_int(q3.x), __float_as_int(q3.y));
}
// One TriRec: v0 | eAB | eAC | N in q0..q2 (the padding half-line rides along with q2).
// TriRec in two halves (rt_types.h): the plane (v0, N) first, the edges when they are needed.
RT_DEV void loadTriPlane(const TriRec* rec, f3& v0, f3& N) {
    float4 a, b;
    ld256(rec, a, b);
    v0 = mk3(a.x, a.y, a.z); N = mk3(a.w, b.x, b.y);
}
RT_DEV void loadTriEdges(const TriRec* rec, f3& eAB, f3& eAC) {
    float4 a, b;
    ld256(reinterpret_cast<const char*>(rec) + 32, a, b);
    eAB = mk3(a.x, a.y, a.z); eAC = mk3(a.w, b.x, b.y);
}

// fragment.glsl:206-218.  float(result)/2^32*2-1: the scalings are exact, so one rounding at the -1.
RT_DEV float randomValue(uint32_t& state) {
    state = state * 747796405u + 2891336453u;
    uint32_t t = state >> ((state >> 28u) + 4u);
    uint32_t result = (t ^ state) * 277803737u;
    result = (result >> 22u) ^ result;
    float f = __uint2float_rn(result);
    return sub(mul(mul(f, 2.3283064365386963e-10f), 2.0f), 1.0f);
}

// fragment.glsl:181-204 with invDir hoisted per ray (bit-identical: same operands, same operations).
// Returns the entry distance max(dstNear, 0) and sets `hit` to (dstFar >= dstNear && !(dstFar < 0)), which
// equals dstFar >= max(dstNear, 0): dstNear can only be NaN when every slab value is NaN, and then dstFar
// is NaN too and the comparison is false, as in the shader.  (The sign of a zero distance is not observable:
// the value is only compared.)
RT_DEV float slabTest(f3 o, f3 inv, float minX, float minY, float minZ, float maxX, float maxY, float maxZ, bool& hit) {
    const float tx0 = mul(sub(minX, o.x), inv.x), tx1 = mul(sub(maxX, o.x), inv.x);
    const float ty0 = mul(sub(minY, o.y), inv.y), ty1 = mul(sub(maxY, o.y), inv.y);
    const float tz0 = mul(sub(minZ, o.z), inv.z), tz1 = mul(sub(maxZ, o.z), inv.z);
    const float dstNear = fmaxf(fmaxf(fminf(tx0, tx1), fminf(ty0, ty1)), fminf(tz0, tz1));
    const float dstFar = fminf(fminf(fmaxf(tx0, tx1), fmaxf(ty0, ty1)), fmaxf(tz0, tz1));
    const float entry = fmaxf(dstNear, 0.0f);
    hit = dstFar >= entry;
    return entry;
}

struct RayHit {
    float t, u, v;
    int32_t slot;     // triangle slot of the closest hit, -1 = miss
};

struct WorkCount {
    uint32_t internal, leaves, tris, pops;
};

// Shared-memory traversal stack: entry e of thread t lives at stackBase[e * blockDim.x + t], an 8-byte
// (ref, tNear) pair — lanes touch consecutive 8-byte words, so a push/pop is conflict-free.
struct StackRef {
    uint2* base;
    uint32_t stride;
};

// fragment.glsl:106-147 up to the acceptance test, on a precomputed TriRec.  Returns true when the
// shader would have taken this candidate (h.didHit && h.t < closestT && h.t > 1e-6).  The shader's early
// returns are folded into one predicate; each comparison is the shader's own (NaN-faithful): reject when
// abs(det) < EPS, t <= EPS, u < 0, v < 0 or u + v > 1.  Values computed past a rejected test are unused
// (1/det of a tiny det may be inf/NaN — harmless, the predicate is already false).
// DEFER: fetch the edges only in the lanes whose hit distance is in range — one L1 wavefront less per rejected candidate,
// one more dependent memory access for the accepted ones.  It pays where the data pipe is the ceiling (closed room 4K 8 spp
// 154.2 -> 149.8 ms) and costs where latency is (small frames: cfg2_open 1.46 -> 1.62 ms; the speculating variants), so
// the pass decides on the host (DeviceScene::deferEdges, rt_api.cu) between two instantiations of the pixel and persistent
// kernels; everything else does not defer.  (A run-time flag instead of the template was measured: slower than either.)
template <bool DEFER>
RT_DEV bool triangleTest(const TriRec* __restrict__ rec, f3 o, f3 d, float closestT, float& tOut, float& uOut,
                         float& vOut) {
    const float EPS = 1e-6f;
    f3 v0, N, eAB, eAC;
    loadTriPlane(rec, v0, N);
    if (!DEFER) loadTriEdges(rec, eAB, eAC);
    const float det = dot3(d, N);
    bool ok = !(fabsf(det) < EPS);
    const float invDet = __frcp_rn(det);
    const f3 ao = sub3(o, v0);
    const float t = mul(-dot3(ao, N), invDet);
    ok = ok && !(t <= EPS) && (t < closestT && t > 1e-6f);
    if (ok) {
        if (DEFER) loadTriEdges(rec, eAB, eAC);
        const f3 dao = cross3(ao, d);
        const float u = mul(-dot3(eAC, dao), invDet);
        const float v = mul(dot3(eAB, dao), invDet);
        ok = !(u < 0.0f || v < 0.0f || add(u, v) > 1.0f);
        tOut = t; uOut = u; vOut = v;
    }
    return ok;
}

// fragment.glsl:246-307.  Same visiting order and the same accept/cull comparisons as the shader:
//  * a node's tNear is computed once (when its parent is expanded) and carried on the stack; the pop-time
//    cull `tNear > closestT` uses the closestT current at the pop, as in the shader (:261-262);
//  * the nearer child (tA < tB strictly; on a tie B first, :292-297) is visited next without a round trip
//    through the stack — it would have been popped immediately;
//  * an entry already farther than closestT when pushed can never survive its pop (closestT only
//    shrinks), so it is not pushed; the POPS counter still counts it.
// Control flow is strictly structured (one loop exit, no continue/return inside): with independent thread
// scheduling the lanes of a warp reconverge at the immediate post-dominator of each branch, and an early
// `return` inside the loop moves that point to the function exit — the first version did that and ran with
// 2 of 32 lanes active (profiles/r01_ncu_notes.md).
template <bool COUNT, bool DEFER = false>
RT_DEV RayHit traverse(const DeviceScene& sc, f3 o, f3 d, StackRef st, WorkCount& wc) {
    RayHit best;
    best.t = 1e30f; best.u = 0.f; best.v = 0.f; best.slot = -1;
    const f3 inv = mk3(__frcp_rn(d.x), __frcp_rn(d.y), __frcp_rn(d.z));
    uint32_t sp = 0;
    int32_t cur = sc.rootRef;
    bool running = cur != RT_REF_NONE;   // empty scene: the shader's root index fails its bounds check (:259)
    if (COUNT) wc.pops++;
    if (running) {
        bool hitRoot;
        const float tRoot = slabTest(o, inv, sc.rootMin[0], sc.rootMin[1], sc.rootMin[2], sc.rootMax[0],
                                     sc.rootMax[1], sc.rootMax[2], hitRoot);
        running = hitRoot && !(tRoot > best.t);
    }
    uint2* const stack = st.base + threadIdx.x;
    const uint32_t stride = st.stride;
    // cur == RT_REF_NONE inside the loop means "the last pop was culled: pop again" (no node this iteration).
    while (running) {
        bool goOn = false;      // cur already holds the next node (the near child)
        if (cur >= 0) {
            if (cur != RT_REF_NONE) {
                // internal node: slab-test both children (:281-287); an absent child has NaN bounds -> never hit
                float4 q0, q1, q2;
                int2 refs;
                loadNode(sc.nodes + cur, q0, q1, q2, refs);
                if (COUNT) wc.internal++;
                bool hitA, hitB;
                const float tA = slabTest(o, inv, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, hitA);
                const float tB = slabTest(o, inv, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, hitB);
                if (COUNT) wc.pops += (hitA ? 1u : 0u) + (hitB ? 1u : 0u);
                const bool aFirst = hitA && (!hitB || tA < tB);     // both hit: A first only if strictly nearer
                const int32_t nearRef = aFirst ? refs.x : refs.y, farRef = aFirst ? refs.y : refs.x;
                const float nearT = aFirst ? tA : tB, farT = aFirst ? tB : tA;
                goOn = (hitA || hitB) && !(nearT > best.t);
                const bool pushFar = hitA && hitB && !(farT > best.t);
                if (pushFar) {
                    stack[sp * stride] = make_uint2(uint32_t(farRef), __float_as_uint(farT));
                    sp++;
                }
                cur = nearRef;
            }
        } else {
            // leaf: every triangle in slot order, first found wins ties (:266-279)
            const uint32_t code = uint32_t(~cur);
            uint32_t count = code & ((1u << RT_LEAF_BITS) - 1u);
            uint32_t first = code >> RT_LEAF_BITS;
            if (count == RT_LEAF_BIG) {
                const uint2 big = __ldg(sc.bigLeaves + first);
                first = big.x; count = big.y;
            }
            if (COUNT) { wc.leaves++; wc.tris += count; }
            for (uint32_t k = 0; k < count; k++) {
                float t, u, v;
                if (triangleTest<DEFER>(sc.tris + first + k, o, d, best.t, t, u, v)) {
                    best.t = t; best.u = u; best.v = v; best.slot = int32_t(first + k);
                }
            }
        }
        // ONE pop (:256-262), executed by all lanes together instead of a loop run by the few lanes that need it
        // (ncu: that loop was 19 % of the kernel's instructions at 4 of 32 lanes).  An entry that fails the cull
        // against the current closestT leaves cur = RT_REF_NONE and the lane pops again next iteration.
        const bool havePop = !goOn && sp > 0;
        uint2 e = make_uint2(uint32_t(RT_REF_NONE), 0u);
        if (havePop) { sp--; e = stack[sp * stride]; }
        if (!goOn) cur = (__uint_as_float(e.y) > best.t) ? RT_REF_NONE : int32_t(e.x);
        running = goOn || havePop;
    }
    return best;
}

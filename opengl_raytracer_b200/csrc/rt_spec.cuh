// rt_spec.cuh — variant SAMPLES: the samples of a pixel-frame evaluated in parallel, bit-exactly.
//
// The reference carries the RNG state across the rays_per_pixel samples of a pixel-frame (fragment.glsl:358-363),
// which chains them: a 64-spp pixel on the mesh is ~77 strictly sequential rays.  Measured on B200 (tools/
// chain_probe.py): one 8x4 tile of the 4K/64-spp dragon frame takes 1.3-4.8 ms ALONE on an idle GPU, and 1/8 of the
// frame takes 12.5 ms where the work is worth 3.5 ms — the chain, not the throughput, bounds a tile-sharded frame.
//
// The chain is thinner than it looks.  All samples share the primary ray, and RandomValue() is an LCG step
// (state = state * 747796405 + 2891336453), so the only thing sample s+1 inherits from sample s is HOW MANY draws
// it consumed: 3 per surface hit.  Let F(k) be the path that starts with the RNG advanced by 3k draws (an O(1)
// jump with a table of (a^3k, c_3k)), returning its radiance and its hit count.  Then the pixel is the chain
//        k_0 = 0,  k_{s+1} = k_s + hits(F(k_s)),   colour = (((0 + F(k_0)) + F(k_1)) + ...) / rays_per_pixel
// and every F(k), k < K, can be evaluated independently:
//   stage 1  rtSpecPrimaryKernel  primary ray per pixel (as the wavefront variant); sky pixels and pixels whose path
//                                 ends at bounce 0 are finished here; the others queue an 80-byte record holding the
//                                 sample-independent half of bounce 0.
//   stage 2  rtSpecEvalKernel     items (pixel, k), k < K = spp + margin, evaluated by persistent, vote-scheduled
//                                 warps (the state machine of rt_vote.cuh, fed with items instead of pixels);
//                                 result: radiance + hits, 16 bytes per item.
//   stage 3  rtSpecChainKernel    per pixel: follow the chain through the table, add in sample order, write the
//                                 pixel.  A chain that leaves the table continues by evaluating F on demand.
// Cost: F is evaluated for every k < K but only every hits-th is used, so the work grows by the mean hit count per
// path: 1.21 under open sky (worth it for a small share of a frame, which becomes embarrassingly parallel; a whole frame
// is better served by rt_chain.cuh), 4.7 in a closed room (not worth
// it).  The mode is decided ON THE DEVICE from stage 1's hit fraction (vote::samplesSpeculate): when more than a
// quarter of the pixels hit, stage 2 exits at once and stage 3 evaluates every sample on demand — the plain sequential
// pixel-frame — or, under AUTO (PassParams::specGate), leaves the pass to the persistent kernel launched behind it.
// Exactness: F(k) runs the same operations on the same operands as the sequential loop would at that RNG state, and
// the additions happen in sample order.
#pragma once
#include "rt_shade.cuh"
#include "rt_vote.cuh"

namespace spec {

constexpr int kThreads = 128;
constexpr int kRounds = 4;                   // speculative rounds: 0 = every queued pixel, 1.. = the chains that ran further
constexpr int kJumpLo = 256, kJumpHi = 64;   // offsets k < kJumpLo * kJumpHi can be jumped to
constexpr int kMaxK = kJumpLo * kJumpHi;
__constant__ uint2 cJumpLo[kJumpLo];         // (a, c) of 3k draws, k < 256: state' = a * state + c
__constant__ uint2 cJumpHi[kJumpHi];         // (a, c) of 3 * 256 * h draws
// h blocks of 3 * 256 draws applied to state s by square-and-multiply, `block` = the (a, c) of one block: affine maps
// of one LCG commute, so any order of composition is the same map.  Only for offsets beyond the tables.
__host__ __device__ inline uint32_t jumpBlocks(uint32_t s, unsigned int h, uint2 block) {
    uint32_t ba = block.x, bc = block.y;
    for (; h != 0u; h >>= 1) {
        if (h & 1u) s = s * ba + bc;
        bc = ba * bc + bc;
        ba = ba * ba;
    }
    return s;
}
RT_DEV uint32_t jumpSeed(uint32_t seed, unsigned int k) {
    const uint2 lo = cJumpLo[k & (kJumpLo - 1)];
    const uint32_t s = seed * lo.x + lo.y;
    const unsigned int h = k >> 8;
    if (h < (unsigned int)kJumpHi) {
        const uint2 hi = cJumpHi[h];
        return s * hi.x + hi.y;
    }
    // k >= kMaxK (rays_per_pixel x hits >= 16384: thousands of samples in a closed room): no table entry, compose
    return jumpBlocks(s, h, cJumpHi[1]);
}

// Table sizes.  Round 0 gives every queued pixel rays_per_pixel samples at up to 1.15 hits each (open-sky chains run
// at 1.0-1.3).  A chain that leaves its table knows its own rate (hits so far / samples so far), so the pixel asks for
// what its remaining samples should need, with a margin, rounded up to a size class (8, 10, 12, 14, 16, 20, ...); the
// last round asks for the bound (remaining samples x nBounces hits), which always suffices.  A round is then ONE
// evaluation launch over all classes, and a 64-spp frame is done in three rounds where fixed doubling top-ups took
// six (each of which costs at least one path latency, 0.3 ms, however few pixels are left).
// Size classes, four per octave: 8, 10, 12, 14, 16, 20, 24, 28, 32, 40, ... — a chain is given at most 1.25 x what it asked
// for (two per octave, round 1, gave up to 1.5 x; with the tighter margins below: 1/2, 1/4, 1/8 share of the 4K dragon
// frame 15.6 / 8.47 / 4.60 -> 15.0 / 8.05 / 4.49 ms, profiles/r02_variants.md).
constexpr int kClasses = 44;
constexpr int kClassStride = 64;             // uints per round in PassParams::classCount: [c] pixels of class c, [62] entries reserved
constexpr int kReserved = 62;
constexpr int kClassBits = 6;                // chainClass = rank << kClassBits | class
__host__ __device__ inline unsigned int classK(int c) { return (4u + (unsigned int)(c & 3)) << ((c >> 2) + 1); }
inline int laterK(int raysPerPixel) {      // round 0's allowance over one hit per sample: 15 % (measured best of 0-25 %)
    const int k = (raysPerPixel * 3 + 10) / 20;
    return k > 8 ? (k < 256 ? k : 256) : 8;
}
inline int tableK(int raysPerPixel) {
    const int k = raysPerPixel + laterK(raysPerPixel);
    return k < kMaxK / 4 ? k : kMaxK / 4;
}

struct Record {          // the sample-independent half of bounce 0 (fragment.glsl:319-334), 5 x float4 per queued pixel
    f3 normal0, s1, origin0, incoming0, rayColor0;
    float k0;
    uint32_t seed, pix;
};

RT_DEV void storeRecord(const PassParams& P, unsigned int j, const Record& r) {
    P.queueA[j] = make_float4(r.normal0.x, r.normal0.y, r.normal0.z, r.k0);
    P.queueB[j] = make_float4(r.s1.x, r.s1.y, r.s1.z, __uint_as_float(r.seed));
    P.queueC[j] = make_float4(r.origin0.x, r.origin0.y, r.origin0.z, __uint_as_float(r.pix));
    P.queueD[j] = make_float4(r.incoming0.x, r.incoming0.y, r.incoming0.z, 0.f);
    P.queueE[j] = make_float4(r.rayColor0.x, r.rayColor0.y, r.rayColor0.z, 0.f);
}
RT_DEV Record loadRecord(const PassParams& P, unsigned int j) {
    const float4 a = __ldg(P.queueA + j), b = __ldg(P.queueB + j), c = __ldg(P.queueC + j), d = __ldg(P.queueD + j),
                 e = __ldg(P.queueE + j);
    Record r;
    r.normal0 = mk3(a.x, a.y, a.z); r.k0 = a.w;
    r.s1 = mk3(b.x, b.y, b.z); r.seed = __float_as_uint(b.w);
    r.origin0 = mk3(c.x, c.y, c.z); r.pix = __float_as_uint(c.w);
    r.incoming0 = mk3(d.x, d.y, d.z);
    r.rayColor0 = mk3(e.x, e.y, e.z);
    return r;
}

RT_DEV bool specMode(const PassParams& P, unsigned int queued) { return vote::samplesSpeculate(P, queued); }

// Round 0's table entries per pixel: the launch's size, clamped so that the queued pixels fit the buffer.
RT_DEV unsigned int roundK(const PassParams& P, unsigned int pixelsThisRound) {
    const unsigned int fit = P.specBufEntries / (pixelsThisRound > 0u ? pixelsThisRound : 1u);
    return min((unsigned int)P.specK, fit);
}

// Where a round's pixels and table entries are: class c holds pixels [pixBase[c], pixBase[c+1]) of the round's sorted
// list, each with K[c] entries starting at itemBase[c] (an item's index is its entry in the table).  Round 0 is one
// class holding every queued pixel.
struct ClassTable {
    unsigned int itemBase[kClasses + 1], pixBase[kClasses + 1], K[kClasses];
};
RT_DEV void buildClassTable(const PassParams& P, int round, unsigned int queued, bool speculative, ClassTable& T) {
    if (threadIdx.x == 0) {
        unsigned int items = 0, pix = 0;
        for (int c = 0; c < kClasses; c++) {
            unsigned int n, K;
            if (round == 0) { n = c == 0 ? queued : 0u; K = speculative ? roundK(P, queued) : 0u; }
            else { n = P.classCount[round * kClassStride + c]; K = classK(c); }
            T.itemBase[c] = items; T.pixBase[c] = pix; T.K[c] = K;
            items += n * K; pix += n;
        }
        T.itemBase[kClasses] = items; T.pixBase[kClasses] = pix;
    }
    __syncthreads();
}

// ------------------------------------------------------------------------------------------ stage 1
template <bool IDS>
__global__ void __launch_bounds__(kThreads) rtSpecPrimaryKernel(const __grid_constant__ PassParams P) {
    extern __shared__ uint2 stackMem[];
    const DeviceScene& sc = P.scene;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ox = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
    const int oy = blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
    const bool valid = ox < P.ownedW && oy < P.ownedH;
    const f3 zero = mk3(0.f, 0.f, 0.f), one = mk3(1.f, 1.f, 1.f);
    bool queued = false;
    Record rec;
    rec.normal0 = zero; rec.s1 = zero; rec.origin0 = zero; rec.incoming0 = zero; rec.rayColor0 = zero;
    rec.k0 = 0.f; rec.seed = 0; rec.pix = 0;
    if (valid) {
        const int px = ownedToPixel(ox, P.tileSizeX, P.numTilesX, P.tileX);
        const int py = ownedToPixel(oy, P.tileSizeY, P.numTilesY, P.tileY);
        const size_t at = size_t(py) * size_t(P.width) + size_t(px);
        const bool casts = P.nBounces > 0 && P.raysPerPixel > 0;
        const f3 camPos = ld3(P.camPos);
        RayHit hit;
        hit.slot = -1; hit.t = 0.f; hit.u = 0.f; hit.v = 0.f;
        f3 dir = zero;
        if (casts) {
            StackRef st;
            st.base = stackMem; st.stride = kThreads;
            WorkCount work = {0, 0, 0, 0};
            dir = primaryDirection(P, px, py, rec.seed);
            hit = traverse<false>(sc, camPos, dir, st, work);
        }
        if (IDS) P.primaryIds[at] = hit.slot >= 0 ? int32_t(__ldg(sc.triId + hit.slot)) : -1;
        f3 color = zero;
        bool finished = true;
        if (casts && hit.slot < 0) {
            // primary miss: every sample is incoming = 0 + sky (:342-346), summed as trace() does
            const f3 inc = add3(zero, scale3(mk3(0.1f, 0.6f, 0.92f), P.skyBrightness));
            for (int s = 0; s < P.raysPerPixel; s++) color = add3(color, inc);
        } else if (casts) {
            // the sample-independent half of bounce 0 (:319-334)
            const SurfaceHit hit0 = resolveHit(sc, hit, camPos, dir);
            const f3 spec0 = reflect3(dir, hit0.normal);
            const float t0 = sub(1.0f, hit0.roughness);
            f3 emitted0 = scale3(hit0.emissionColor, hit0.emission);
            if (P.lambertian != 0) emitted0 = scale3(emitted0, 2.0f);
            rec.incoming0 = add3(zero, mul3(emitted0, one));
            if (hit0.emission > 0.0f || P.nBounces <= 1) {
                // every path ends at bounce 0: each sample adds 0 + incoming0; diffuse()'s draws never reach the output
                for (int s = 0; s < P.raysPerPixel; s++) color = add3(color, rec.incoming0);
            } else {
                rec.normal0 = hit0.normal;
                rec.k0 = sub(1.0f, t0);
                rec.s1 = scale3(length3(spec0) > 0.0f ? normalize3(spec0) : zero, t0);
                rec.origin0 = add3(hit0.point, scale3(hit0.normal, 1e-4f));
                rec.rayColor0 = mul3(one, hit0.color);
                rec.pix = uint32_t(px) | (uint32_t(py) << 16);
                finished = false;
            }
        }
        if (finished) storePixel(P, at, accumulate(P, div3(color, float(P.raysPerPixel)), at));   // :365, :409-416
        queued = !finished;
    }
    const uint32_t m = __ballot_sync(0xffffffffu, queued);
    if (m != 0u) {
        unsigned int base = 0;
        const int leader = __ffs(m) - 1;
        if (lane == leader) base = atomicAdd(P.queueCount, (unsigned int)__popc(m));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (queued) storeRecord(P, base + (unsigned int)__popc(m & ((1u << lane) - 1u)), rec);
    }
}

// F(k) evaluated start to finish by one thread (stage 3's on-demand path; the same arithmetic as stage 2's
// state machine).  `seed` is the RNG state at the start of the sample; returns the sample's incoming light,
// advances seed by 3 draws per surface hit.
RT_DEV f3 evalPath(const PassParams& P, const Record& r, uint32_t& seed, StackRef st, f3 sky, bool lambertian) {
    const DeviceScene& sc = P.scene;
    const f3 dDiff0 = diffuseDir(r.normal0, seed, lambertian);
    f3 d = vote::lerpDirPre(dDiff0, r.k0, r.s1);
    f3 o = r.origin0;
    f3 incoming = r.incoming0, rayColor = r.rayColor0;
    for (int i = 1; i < P.nBounces; i++) {
        WorkCount wc = {0, 0, 0, 0};
        const RayHit h = traverse<false>(sc, o, d, st, wc);
        if (h.slot >= 0) {
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
    return incoming;
}

// ------------------------------------------------------------------------------------------ stage 2
// Items (pixel j, offset k) through the warp-vote state machine of rt_vote.cuh: INTERNAL bursts, LEAF and SHADE
// sections picked by ballot.  A lane that finishes an item writes (incoming, hits) and fetches the next item;
// consecutive items are consecutive k of one pixel, so the lanes of a warp start from the same surface point.
__global__ void __launch_bounds__(kThreads) rtSpecEvalKernel(const __grid_constant__ PassParams P) {
    extern __shared__ uint2 smem[];
    using namespace vote;
    const DeviceScene& sc = P.scene;
    const unsigned int queued = *P.queueCount;
    if (!specMode(P, queued)) return;
    const int round = P.specRound;
    __shared__ ClassTable T;
    buildClassTable(P, round, queued, true, T);
    const uint32_t lane = threadIdx.x & 31u;
    uint2* const stack = smem + threadIdx.x;
    const uint32_t stride = kThreads;
    const unsigned int totalItems = T.itemBase[kClasses];
    unsigned int* const workCounter = P.workCounters + 2 * round;
    const f3 zero = mk3(0.f, 0.f, 0.f);
    const f3 sky = scale3(mk3(0.1f, 0.6f, 0.92f), P.skyBrightness);
    const bool lambertian = P.lambertian != 0;
    const int leafLanes = P.tune[0] > 0 ? P.tune[0] : RT_VOTE_LEAF_LANES;
    const int shadeLanes = P.tune[1] > 0 ? P.tune[1] : RT_VOTE_SHADE_LANES;
    uint32_t cnt[RT_CNT_COUNT];   // unused (COUNT = false), keeps the shared step functions' signature

    Lane L;
    L.o = zero; L.d = zero; L.inv = zero;
    L.best.t = 0.f; L.best.u = 0.f; L.best.v = 0.f; L.best.slot = -1;
    L.cur = 0; L.sp = 0; L.mode = MODE_SHADE;
    L.incoming = zero; L.rayColor = zero; L.color = zero;
    L.seed = 0; L.sample = 0; L.bounce = 0; L.px = 0; L.py = 0;
    L.fresh = false; L.needPixel = true;    // needPixel: the lane needs an item; px holds the item index

    for (;;) {
        const uint32_t mI = __ballot_sync(0xffffffffu, L.mode == MODE_INTERNAL);
        const uint32_t mL = __ballot_sync(0xffffffffu, L.mode == MODE_LEAF);
        const uint32_t mS = __ballot_sync(0xffffffffu, L.mode == MODE_SHADE);
        if ((mI | mL | mS) == 0u) break;
        const int active = __popc(mI | mL | mS);
        const int shadeT = min(shadeLanes, max(1, (active * 3) >> 3));
        const int leafT = min(leafLanes, max(1, active >> 2));
        if (__popc(mS) >= shadeT || (mI | mL) == 0u) {
            // ------------------------------------------------------------------ SHADE section
            if (L.mode == MODE_SHADE && !L.needPixel) {
                // one iteration of raytrace()'s loop body (fragment.glsl:313-347) for the ray just traced
                bool pathEnds;
                if (L.best.slot >= 0) {
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
                    if (!pathEnds) beginRay<false>(sc, L, cnt);
                } else {
                    L.incoming = add3(L.incoming, sky);   // un-attenuated, as the shader does (:343)
                    pathEnds = true;
                }
                if (pathEnds) {
                    // F(k) = (incoming, number of surface hits = draws consumed / 3)
                    P.specBuf[L.px] = make_float4(L.incoming.x, L.incoming.y, L.incoming.z, __int_as_float(L.bounce));
                    L.needPixel = true;
                    L.mode = MODE_SHADE;
                }
            }
            // ---- item fetch (one atomic per warp)
            const uint32_t want = __ballot_sync(0xffffffffu, L.mode == MODE_SHADE && L.needPixel);
            if (want != 0u) {
                unsigned int base = 0;
                if (lane == uint32_t(__ffs(want) - 1)) base = atomicAdd(workCounter, (unsigned int)__popc(want));
                base = __shfl_sync(0xffffffffu, base, __ffs(want) - 1);
                if (L.mode == MODE_SHADE && L.needPixel) {
                    const unsigned int item = base + (unsigned int)__popc(want & ((1u << lane) - 1u));
                    if (item >= totalItems) {
                        L.mode = MODE_DONE; L.needPixel = false;
                    } else {
                        // item -> (t-th pixel of this round, kk-th offset past the pixel's chain position)
                        int c = 0;
                        while (item >= T.itemBase[c + 1]) c++;
                        const unsigned int K = T.K[c], rel = item - T.itemBase[c], tIn = rel / K, kk = rel - tIn * K;
                        const unsigned int t = T.pixBase[c] + tIn;
                        const unsigned int j = round == 0 ? t : P.todoSorted[t];
                        const unsigned int k = (round == 0 ? 0u : P.chainK[j]) + kk;
                        const Record r = loadRecord(P, j);
                        L.seed = jumpSeed(r.seed, k);           // RNG state after 3k draws
                        L.px = int(item);
                        L.needPixel = false;
                        // bounce 0 of the sample: the cached primary hit (:319-334)
                        const f3 dDiff0 = diffuseDir(r.normal0, L.seed, lambertian);
                        L.d = lerpDirPre(dDiff0, r.k0, r.s1);
                        L.o = r.origin0;
                        L.incoming = r.incoming0;
                        L.rayColor = r.rayColor0;
                        L.bounce = 1;
                        beginRay<false>(sc, L, cnt);
                    }
                }
            }
        } else if (__popc(mL) >= leafT || mI == 0u) {
            if (L.mode == MODE_LEAF) stepLeaf<false, false>(sc, L, stack, stride, cnt);
        } else {
#pragma unroll 1
            for (int s = 0; s < kBurst; s++)
                if (L.mode == MODE_INTERNAL) stepInternal<false>(sc, L, stack, stride, cnt);
        }
    }
}

// ------------------------------------------------------------------------------------------ stage 3
// One thread per pixel of the round: trace() (fragment.glsl:352-366) as a chain through the table of F.  A chain
// that leaves its table parks (k, sample, colour), reserves a table of the size class its own rate predicts and joins
// the next round; after the last round — or when there is no table at all, or no room for one — the remaining samples
// are evaluated one after the other.
__global__ void __launch_bounds__(kThreads) rtSpecChainKernel(const __grid_constant__ PassParams P) {
    extern __shared__ uint2 stackMem[];
    const uint32_t lane = threadIdx.x & 31u;
    StackRef st;
    st.base = stackMem; st.stride = kThreads;
    const unsigned int queued = *P.queueCount;
    const int round = P.specRound;
    const bool speculative = specMode(P, queued);
    if (!speculative && (round > 0 || P.specGate != 0)) return;   // gated: the persistent kernel renders this pass
    __shared__ ClassTable T;
    buildClassTable(P, round, queued, speculative, T);
    const unsigned int pixelsThisRound = T.pixBase[kClasses];
    const f3 sky = scale3(mk3(0.1f, 0.6f, 0.92f), P.skyBrightness);
    const bool lambertian = P.lambertian != 0;
    unsigned int* const workCounter = P.workCounters + 2 * round + 1;
    unsigned int* const nextClass = P.classCount + (round + 1) * kClassStride;
    for (;;) {
        unsigned int base = 0;
        if (lane == 0) base = atomicAdd(workCounter, 32u);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= pixelsThisRound) break;
        const unsigned int t = base + lane;
        bool again = false;
        unsigned int j = 0;
        if (t < pixelsThisRound) {
            int c = 0;
            while (t >= T.pixBase[c + 1]) c++;
            const unsigned int K = T.K[c];                     // 0: every sample on demand
            j = round == 0 ? t : P.todoSorted[t];
            const Record r = loadRecord(P, j);
            f3 color = mk3(0.f, 0.f, 0.f);
            unsigned int k = 0;
            int s = 0;
            if (round > 0) {
                const float4 saved = P.chainColor[j];
                color = mk3(saved.x, saved.y, saved.z);
                s = __float_as_int(saved.w);
                k = P.chainK[j];
            }
            const unsigned int k0 = k;
            const float4* table = P.specBuf + size_t(T.itemBase[c]) + size_t(t - T.pixBase[c]) * K;
            for (; s < P.raysPerPixel && k - k0 < K; s++) {
                const float4 f = table[k - k0];
                color = add3(color, mk3(f.x, f.y, f.z));      // trace(), :359
                k += (unsigned int)__float_as_int(f.w);
            }
            if (s < P.raysPerPixel && K > 0 && round + 1 < kRounds) {
                // the table this chain asks for: its rate so far x the samples left, with a margin (a sample's hit count
                // varies by a few) — or the bound, in the last round and where there is no rate yet
                const unsigned long long rem = (unsigned long long)(P.raysPerPixel - s);
                const unsigned long long bound = rem * (unsigned long long)max(P.nBounces, 1);
                unsigned long long need = bound;
                if (round + 2 < kRounds && s > 0) {
                    // margins: tune[3] = (per-mille factor << 8) | (tenths of sqrt(rem)); 0 = the defaults
                    const float mFactor = P.tune[3] > 0 ? float(P.tune[3] >> 8) * 1e-3f : 1.03f;
                    const float mSqrt = P.tune[3] > 0 ? float(P.tune[3] & 255) * 0.1f : 1.0f;
                    const float predicted = float(k) / float(s) * float(rem) * mFactor + mSqrt * sqrtf(float(rem)) + 2.0f;
                    need = min(bound, (unsigned long long)predicted);
                }
                int cls = 0;
                while (cls < kClasses && (unsigned long long)classK(cls) < need) cls++;
                if (cls < kClasses && (unsigned long long)k + classK(cls) + 8ull < (unsigned long long)kMaxK) {
                    const unsigned int Kc = classK(cls);
                    const unsigned int before = atomicAdd(nextClass + kReserved, Kc);
                    if ((unsigned long long)before + Kc <= (unsigned long long)P.specBufEntries) {
                        const unsigned int rank = atomicAdd(nextClass + cls, 1u);
                        P.chainColor[j] = make_float4(color.x, color.y, color.z, __int_as_float(s));
                        P.chainK[j] = k;
                        P.chainClass[j] = (rank << kClassBits) | (unsigned int)cls;
                        again = true;
                    } else {
                        atomicSub(nextClass + kReserved, Kc);   // the buffer is full: this pixel finishes below
                    }
                }
            }
            if (!again) {
                if (s < P.raysPerPixel) {
                    uint32_t seed = jumpSeed(r.seed, k);
                    for (; s < P.raysPerPixel; s++) color = add3(color, evalPath(P, r, seed, st, sky, lambertian));
                }
                const int px = int(r.pix & 0xffffu), py = int(r.pix >> 16);
                const size_t at = size_t(py) * size_t(P.width) + size_t(px);
                storePixel(P, at, accumulate(P, div3(color, float(P.raysPerPixel)), at));   // :365, :409-416
            }
        }
        const uint32_t m = __ballot_sync(0xffffffffu, again);
        if (m != 0u) {
            unsigned int at = 0;
            const int leader = __ffs(m) - 1;
            if (int(lane) == leader) at = atomicAdd(P.todoCount + round + 1, (unsigned int)__popc(m));
            at = __shfl_sync(0xffffffffu, at, leader);
            if (again) P.todoFlat[at + (unsigned int)__popc(m & ((1u << lane) - 1u))] = j;
        }
    }
}

// Between two rounds: the parked pixels, listed in the order they parked, are put in class order (a pixel's place is
// its class's first place + the rank it drew when it parked), which is the order the round's table is laid out in.
__global__ void __launch_bounds__(kThreads) rtSpecSortKernel(const __grid_constant__ PassParams P) {
    const unsigned int queued = *P.queueCount;
    if (!specMode(P, queued)) return;
    const int round = P.specRound;
    __shared__ unsigned int pixBase[kClasses];
    if (threadIdx.x == 0) {
        unsigned int pix = 0;
        for (int c = 0; c < kClasses; c++) { pixBase[c] = pix; pix += P.classCount[round * kClassStride + c]; }
    }
    __syncthreads();
    const unsigned int n = P.todoCount[round];
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const unsigned int j = P.todoFlat[i], cr = P.chainClass[j];
        P.todoSorted[pixBase[cr & ((1u << kClassBits) - 1u)] + (cr >> kClassBits)] = j;
    }
}

inline size_t smemBytes(int stackDepth) { return size_t(stackDepth) * kThreads * sizeof(uint2); }

// Host: the jump tables.  (a_n, c_n) with state_n = a_n * state_0 + c_n after n LCG steps (mod 2^32).
inline void fillJumpTables(uint2* lo, uint2* hi) {
    const uint32_t A = 747796405u, C = 2891336453u;
    uint32_t a = 1u, c = 0u;
    for (int k = 0; k < kJumpLo; k++) {
        lo[k] = make_uint2(a, c);
        for (int step = 0; step < 3; step++) { a = a * A; c = c * A + C; }
    }
    const uint32_t aBlock = a, cBlock = c;      // 3 * 256 steps
    a = 1u; c = 0u;
    for (int h = 0; h < kJumpHi; h++) {
        hi[h] = make_uint2(a, c);
        c = c * aBlock + cBlock; a = a * aBlock;
    }
}

}  // namespace spec

// rt_chain.cuh — variant CHAIN: persistent warps over the queue of hit pixels, every lane owning one pixel's sample chain,
// with IN-WARP SAMPLE STEALING: a lane that has no pixel of its own evaluates samples of its neighbours' pixels ahead of
// them, bit-exactly.
//
// Why.  The reference carries the RNG state across the rays_per_pixel samples of a pixel-frame (fragment.glsl:358-363):
// a 64-spp pixel on the mesh is ~77 strictly sequential rays (rt_spec.cuh has the measurements).  The persistent kernel
// (rt_vote.cuh) gives each lane one pixel, so a pass ends with a tail in which a few lanes finish long chains while the
// rest of the GPU idles — and a GPU that owns 1/8 of a tile-sharded frame is ALL tail (12.3 ms for work worth 3.4 ms).
// rt_spec.cuh cuts the chain by evaluating F(k) = "the path whose RNG starts 3k draws in" for every offset k at once
// (all samples share the primary ray and RandomValue() is an LCG, so sample s+1 inherits from sample s only the number
// of draws it consumed, 3 per surface hit: k_{s+1} = k_s + hits(F(k_s))), but every offset the chain jumps over was
// evaluated for nothing: hits per sample - 1 = 21 % more paths on the dragon under open sky, 370 % in a closed room.
//
// Here speculation is only what idle lanes do:
//   stage 1  spec::rtSpecPrimaryKernel  (shared with SAMPLES) primary ray per pixel; sky pixels and pixels whose paths end at
//            bounce 0 are finished, the others queue the sample-independent half of bounce 0 (80 bytes).
//   stage 2  rtChainKernel  persistent warps (6 CTAs x 4 warps per SM), the warp-vote state machine of rt_vote.cuh.  Lane a
//            owns slot a: one queued pixel, its chain position kCur, its colour sum.  A lane's first choice of work is its
//            own pixel's F(kCur) — never wasted, exactly the sequential loop.  A lane without a pixel takes a new one from
//            the queue while the warp's QUOTA allows (quota = pixels left in the queue / warps in the grid, at most 32: as
//            the queue drains, warps own fewer pixels and free lanes).  A lane with neither becomes a HELPER: it claims the
//            next unclaimed offset of the neighbour with the most samples left (within a window of kRing offsets), evaluates
//            F there and drops (radiance, hits) into that slot's ring (global memory, L2-resident).  Every lane walks its own chain
//            through its ring whenever a SHADE section runs: colour += F(kCur), kCur += hits, in sample order.  An item
//            whose offset the chain has jumped over is dropped where it stands.
// So a whole frame on one GPU runs like the persistent kernel (no speculation until the queue holds less than one pixel per
// lane), its tail is shared out inside the warps, and a 1/N share — fewer hit pixels than lanes from the start — runs with
// quota < 32: e.g. 20 owners and 12 helpers per warp.  Wasted evaluations are only those of helpers (and of owners whose own
// offset a helper already holds) that the chain then skips.
//
// Exactness: F(k) runs the same operations on the same operands as the sequential loop would at that RNG state (the
// state is jumped to, spec::jumpSeed), and the additions happen in sample order, as in rt_spec.cuh.
#pragma once
#include "rt_shade.cuh"
#include "rt_vote.cuh"
#include "rt_spec.cuh"

namespace chain {

#ifdef RT_CHAIN_STATS
__device__ unsigned long long gTimelineStart;
__device__ unsigned int gTimeline[3][256];      // per 0.25 ms of the kernel: evaluations started, samples consumed, pixels finished
#endif

constexpr int kThreads = 128;
constexpr int kRing = 32;                  // ring entries per slot: offsets [kCur, kCur + window), window <= kRing, can be claimed
constexpr int kHitBits = 6;                // ring tag = (k + 1) << kHitBits | hits: nBounces < 64 (the host checks)

// The rings live in global memory (L2): kRing x 16 bytes per lane of the grid.  Only helped slots ever touch them — an
// owner's own result stays in registers — so a whole frame on one GPU reads no ring until its tail.
inline size_t smemBytes(int stackDepth) { return size_t(stackDepth) * kThreads * sizeof(uint2); }
inline size_t ringBytes(int gridBlocks) { return size_t(gridBlocks) * kThreads * kRing * sizeof(float4); }


__global__ void __launch_bounds__(kThreads, 6) rtChainKernel(const __grid_constant__ PassParams P) {
    extern __shared__ uint2 smem[];
    using namespace vote;
    const DeviceScene& sc = P.scene;
    const unsigned int queued = *P.queueCount;
    if (P.specGate != 0 && !samplesSpeculate(P, queued)) return;   // gated: the persistent kernel behind this launch renders the pass
    const uint32_t lane = threadIdx.x & 31u;
    uint2* const stack = smem + threadIdx.x;
    const uint32_t stride = kThreads;
    // ring entry e of this warp's lane t: warpRing[t * kRing + e] = (incoming xyz, tag)
    float4* const warpRing = P.chainRing + (size_t(blockIdx.x) * kThreads + (threadIdx.x & ~31u)) * kRing;
    float4* const myRing = warpRing + lane * kRing;
    unsigned int* const pixelCounter = P.workCounters;   // next queued pixel (zeroed per pass)
    const f3 zero = mk3(0.f, 0.f, 0.f);
    const f3 sky = scale3(mk3(0.1f, 0.6f, 0.92f), P.skyBrightness);
    const bool lambertian = P.lambertian != 0;
    const int leafLanes = P.tune[0] > 0 ? P.tune[0] : RT_VOTE_LEAF_LANES;
    const int shadeLanes = P.tune[1] > 0 ? P.tune[1] : RT_VOTE_SHADE_LANES;
    const uint32_t spp = uint32_t(P.raysPerPixel);
    const uint32_t totalWarps = gridDim.x * (kThreads / 32);
    const uint32_t quotaScale = uint32_t(max(P.chainTune[0], 1));      // sixteenths: quota = pixels left * scale / 16 / warps
    const bool helping = P.chainTune[1] != 0;
    const uint32_t window = uint32_t(min(max(P.chainTune[2], 2), kRing));
    // The queue is in image order and a pixel's cost is a matter of where it lies (a fold of the mesh: five bounces per
    // sample; a convex flank: one), so consecutive entries would give a warp 32 dear or 32 cheap pixels and in-warp help
    // nothing to even out.  Queue position i is therefore read at (i * step) mod queued, step ~ 0.618 * queued and coprime
    // to it: a bijection that hands every warp pixels from all over the frame.
    uint32_t step = 1u;
    if (P.chainTune[3] != 0 && queued > 2u) {
        step = uint32_t((unsigned long long)queued * 618ull / 1000ull) | 1u;
        for (;;) {
            uint32_t a = queued, b = step;
            while (b != 0u) { const uint32_t t = a % b; a = b; b = t; }
            if (a == 1u) break;
            step += 2u;
        }
    }
    uint32_t cnt[RT_CNT_COUNT];   // unused (COUNT = false), keeps the shared step functions' signature
#ifdef RT_CHAIN_STATS
    {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        atomicCAS(&gTimelineStart, 0ull, t);
    }
    const unsigned long long stStart = *reinterpret_cast<volatile unsigned long long*>(&gTimelineStart);
#define RT_TIMELINE(which)                                                                   \
    do {                                                                                     \
        unsigned long long t_;                                                               \
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                               \
        const unsigned long long bin_ = (t_ - stStart) / 250000ull;                          \
        atomicAdd(&gTimeline[which][bin_ < 255ull ? bin_ : 255ull], 1u);                     \
    } while (0)
    uint32_t stStarted = 0, stOwn = 0, stRefused = 0, stDropped = 0, stRays = 0, stIdle = 0, stSections = 0, stUsed = 0, stIters = 0;
    unsigned long long stT0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(stT0));
#define RT_STAT(x) (x)
#else
#define RT_STAT(x) ((void)0)
#define RT_TIMELINE(which) ((void)0)
#endif

    Lane L;
    L.o = zero; L.d = zero; L.inv = zero;
    L.best.t = 0.f; L.best.u = 0.f; L.best.v = 0.f; L.best.slot = -1;
    L.cur = 0; L.sp = 0; L.mode = MODE_SHADE;
    L.incoming = zero; L.rayColor = zero; L.color = zero;     // color: this lane's OWN pixel's sum, in sample order
    L.seed = 0; L.sample = 0; L.bounce = 0; L.px = 0; L.py = 0;
    L.fresh = false; L.needPixel = true;    // needPixel: the lane holds no item
    // the item in flight: (slot = owner lane, offset, the slot's pixel serial when it was claimed)
    uint32_t itemLane = lane, itemK = 0, itemSerial = 0;
    // this lane's slot
    bool slotActive = false;
    uint32_t slotPixel = 0;                 // queue index of the pixel
    uint32_t slotSerial = 0;                // bumped whenever the slot changes pixel
    uint32_t kCur = 0, sDone = 0, nextK = 0;   // chain position, samples added, first offset nobody has claimed (>= kCur)
    bool helped = false;                    // some offset of this pixel went to another lane (or to this lane out of order)
    bool exhausted = false;                 // warp-uniform: the queue is empty
    uint32_t left = queued;                 // warp-uniform estimate of the pixels left in the queue (last value seen)

    for (;;) {
        const uint32_t mI = __ballot_sync(0xffffffffu, L.mode == MODE_INTERNAL);
        const uint32_t mL = __ballot_sync(0xffffffffu, L.mode == MODE_LEAF);
        const uint32_t mS = __ballot_sync(0xffffffffu, L.mode == MODE_SHADE);
        if ((mI | mL | mS) == 0u) break;    // every lane idle: no slot is active and the queue is empty (see phase 4)
        RT_STAT(L.mode == MODE_DONE ? stIdle++ : stRefused++);     // (stats build: loop iterations idle / busy)
        RT_STAT(stIters++);
        const int active = __popc(mI | mL | mS);
        const int shadeT = min(shadeLanes, max(1, (active * 3) >> 3));
        const int leafT = min(leafLanes, max(1, active >> 2));
        if (__popc(mS) >= shadeT || (mI | mL) == 0u) {
            // ---------------------------------------------------------------------------------------- SHADE section
            // phase 1: the ray just traced -- one iteration of raytrace()'s loop body (fragment.glsl:313-346).  (Shading the
            // hits in a section of their own that waits for more of them was measured and is slower: 27.3 vs 26.2 ms, the
            // wait lengthens every chain; profiles/r02_variants.md)
            bool done = false;
            if (L.mode == MODE_SHADE && !L.needPixel) {
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
                    if (hit.emission > 0.0f || L.bounce >= P.nBounces) done = true;
                    else { beginRay<false>(sc, L, cnt); RT_STAT(stRays++); }
                } else {
                    L.incoming = add3(L.incoming, sky);   // un-attenuated, as the shader does (:343)
                    done = true;
                }
            }
            // deliver F(k) = (incoming, hits): an owner's own next sample is added right here (trace(), fragment.glsl:359); any
            // other result goes to the slot's ring, unless the chain has moved past k or the slot has moved on
            {
                const uint32_t tSerial = __shfl_sync(0xffffffffu, slotSerial, itemLane);
                const uint32_t tK = __shfl_sync(0xffffffffu, kCur, itemLane);
                if (done) {
                    if (itemLane == lane && itemK == kCur && itemSerial == slotSerial) {
                        L.color = add3(L.color, L.incoming);
                        kCur += uint32_t(L.bounce);
                        sDone++;
                        RT_STAT(stUsed++); RT_TIMELINE(1);
                    } else if (tSerial == itemSerial && itemK >= tK) {
                        __stcg(warpRing + itemLane * kRing + (itemK & (kRing - 1)),
                               make_float4(L.incoming.x, L.incoming.y, L.incoming.z, __uint_as_float(((itemK + 1u) << kHitBits) | uint32_t(L.bounce))));
                    }
                    L.needPixel = true;
                }
            }
            __syncwarp();
            // phase 2: every lane walks its own chain on through its ring (trace(), fragment.glsl:352-366)
            if (slotActive) {
                if (helped) {
                    while (sDone < spp) {
                        const float4 f = __ldcg(myRing + (kCur & (kRing - 1)));
                        const uint32_t tag = __float_as_uint(f.w);
                        if ((tag >> kHitBits) != kCur + 1u) break;
                        L.color = add3(L.color, mk3(f.x, f.y, f.z));      // trace(), :359
                        kCur += tag & ((1u << kHitBits) - 1u);
                        sDone++;
                        RT_STAT(stUsed++); RT_TIMELINE(1);
                    }
                }
                nextK = max(nextK, kCur);
                if (sDone >= spp) {
                    const spec::Record r = spec::loadRecord(P, slotPixel);
                    const size_t at = size_t(r.pix >> 16) * size_t(P.width) + size_t(r.pix & 0xffffu);
                    storePixel(P, at, accumulate(P, div3(L.color, float(P.raysPerPixel)), at));   // :365, :409-416
                    slotActive = false;
                    slotSerial++;
                    RT_TIMELINE(2);
                }
            }
            // phase 3: an item whose offset the chain has passed, or whose pixel is finished, is dropped where it stands
            {
                const uint32_t tSerial = __shfl_sync(0xffffffffu, slotSerial, itemLane);
                const uint32_t tK = __shfl_sync(0xffffffffu, kCur, itemLane);
                if (!L.needPixel && (tSerial != itemSerial || itemK < tK)) { L.needPixel = true; RT_STAT(stDropped++); }
            }
            RT_STAT(stSections++);
            // phase 4: work for the lanes that hold no item
            bool got = !L.needPixel;
            uint32_t takeLane = lane, takeK = 0;
            // (a) the own pixel's next sample, if nobody has claimed it
            if (!got && slotActive && nextK == kCur) { got = true; takeLane = lane; takeK = kCur; nextK = kCur + 1u; RT_STAT(stOwn++); }
            // (b) a new pixel, while the warp owns fewer than its quota
            if (!exhausted) {
                const uint32_t owners = __popc(__ballot_sync(0xffffffffu, slotActive));
                const uint32_t quota = min(32u, max(1u, (left * quotaScale / 16u + totalWarps - 1u) / totalWarps));
                const uint32_t wantNew = __ballot_sync(0xffffffffu, !got && !slotActive);
                const uint32_t allowed = quota > owners ? quota - owners : 0u;
                const uint32_t myRank = __popc(wantNew & ((1u << lane) - 1u));
                const uint32_t n = min(allowed, uint32_t(__popc(wantNew)));
                if (n > 0u) {
                    unsigned int base = 0;
                    if (lane == 0) base = atomicAdd(pixelCounter, n);
                    base = __shfl_sync(0xffffffffu, base, 0);
                    left = base + n < queued ? queued - (base + n) : 0u;
                    exhausted = base + n >= queued;
                    if (!got && !slotActive && myRank < n && base + myRank < queued) {
                        slotActive = true; slotSerial++;
                        slotPixel = uint32_t((unsigned long long)(base + myRank) * step % queued);
                        kCur = 0; sDone = 0; nextK = 1u; helped = false;
                        L.color = zero;
                        got = true; takeLane = lane; takeK = 0; RT_STAT(stOwn++);
                    }
                }
            }
            __syncwarp();
            // (c) help: the slot with the most unclaimed samples that still fit its window hands out its next offsets
            if (helping) {
                for (;;) {
                    const uint32_t helpers = __ballot_sync(0xffffffffu, !got);
                    if (helpers == 0u) break;
                    uint32_t room = 0;
                    if (slotActive) {
                        const uint32_t endNeed = kCur + (spp - sDone), endWin = kCur + window;   // exclusive bounds
                        const uint32_t end = min(endNeed, endWin);
                        room = end > nextK ? end - nextK : 0u;
                    }
                    const uint32_t best = __reduce_max_sync(0xffffffffu, room > 0u ? ((room << 5) | lane) : 0u);
                    if (best == 0u) break;
                    const uint32_t v = best & 31u;
                    const uint32_t n = min(best >> 5, uint32_t(__popc(helpers)));
                    const uint32_t vNext = __shfl_sync(0xffffffffu, nextK, v);
                    const uint32_t myRank = __popc(helpers & ((1u << lane) - 1u));
                    if (!got && myRank < n) { got = true; takeLane = v; takeK = vNext + myRank; }
                    if (lane == v) {
                        // first help for this pixel: entries of earlier pixels (or passes) must not look like its own
                        if (!helped)
                            for (int e = 0; e < kRing; e++) __stcg(&myRing[e].w, 0.f);
                        helped = true;
                        nextK += n;
                    }
                }
            }
            // start the items handed out (the slot's record comes from the queue planes: L1/L2-resident)
            {
                const uint32_t tSerial = __shfl_sync(0xffffffffu, slotSerial, takeLane);
                const uint32_t tPixel = __shfl_sync(0xffffffffu, slotPixel, takeLane);
                if (L.needPixel) {
                    if (got) {
                        itemLane = takeLane; itemK = takeK; itemSerial = tSerial;
                        L.needPixel = false;
                        const spec::Record r = spec::loadRecord(P, tPixel);
                        L.seed = spec::jumpSeed(r.seed, takeK);                          // RNG state after 3k draws
                        // bounce 0 of the sample: the cached primary hit (:319-334)
                        const f3 dDiff0 = diffuseDir(r.normal0, L.seed, lambertian);
                        L.d = lerpDirPre(dDiff0, r.k0, r.s1);
                        L.o = r.origin0;
                        L.incoming = r.incoming0;
                        L.rayColor = r.rayColor0;
                        L.bounce = 1;
                        beginRay<false>(sc, L, cnt);
                        RT_STAT(stStarted++); RT_STAT(stRays++); RT_TIMELINE(0);
                    } else {
                        L.mode = MODE_DONE;     // idle until a SHADE section has something for it
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
#ifdef RT_CHAIN_STATS
    {
        uint32_t v[8] = {stStarted, stOwn, stRefused, stDropped, stRays, stIdle, stSections, stUsed};
#pragma unroll
        for (int i = 0; i < 8; i++) {
            unsigned long long x = v[i];
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) x += __shfl_xor_sync(0xffffffffu, x, off);
            if (lane == 0) atomicAdd(reinterpret_cast<unsigned long long*>(P.workCounters + 16) + i, x);
        }
        unsigned long long t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        if (lane == 0) {
            unsigned long long* extra = reinterpret_cast<unsigned long long*>(P.workCounters + 32);
            atomicMax(extra + 0, (unsigned long long)stIters);          // most loop iterations of any warp
            atomicAdd(extra + 1, (unsigned long long)stIters);
            atomicMax(extra + 2, t1 - stT0);                            // longest-living warp, ns
            atomicAdd(extra + 3, t1 - stT0);
            // histogram of warp life times in 8 bins of 2 ms
            const unsigned long long bin = min((t1 - stT0) / 2000000ull, 7ull);
            atomicAdd(extra + 4 + bin, 1ull);
        }
    }
#endif
}

}  // namespace chain

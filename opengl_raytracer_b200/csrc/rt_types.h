// rt_types.h — device-side data layout of librt_b200 (private to the library).
//
// The reference uploads three SSBOs (fragment.glsl:94-104): tris[] 144 B, boundingBoxes[] 48 B,
// triangleIndices[] u32.  rt_scene_upload re-lays them out, bit-preservingly, so the
// traversal touches only what it needs, with 16-byte vector loads through the read-only path:
//
// Records are 32-byte aligned and fetched with 256-bit loads (ld.global.nc.v8.f32 = LDG.E.256, sm_100+): the L1 data pipe
// charges a divergent load one wavefront per 128-byte line per INSTRUCTION whatever its width, so a 64-byte node costs
// two wavefronts per lane instead of four (profiles/r02_variants.md §1: 200 G vs 120 G records/s L1-resident).
//
//   NodeRec  64 B  one per REACHABLE internal node: the bounds of BOTH children (the shader slab-tests
//                  both children when it expands a node, fragment.glsl:285-287) + two child references.
//                  A popped node's own bounds are never re-fetched: its tNear travels with the stack
//                  entry (the re-test at fragment.glsl:261 recomputes the same value from the same inputs).
//   TriRec   64 B  one per triangle slot, in leaf order (removes the triangleIndices indirection):
//                  v0, eAB = v1-v0, eAC = v2-v0, N = cross(eAB,eAC) (48 B + padding to two 32-byte halves), computed at
//                  upload with the shader's operation order and no contraction (fragment.glsl:112-116).
//   ShadeRec 96 B  one per triangle slot: the three vertex normals + material, fetched once per ray for
//                  the winning triangle only (the shader copies them for every candidate, :151-173).
//   triId    4 B   per slot: the original triangle index (primary-id image only).
//
// Child reference (int32): >= 0 : index of an internal NodeRec
//                          <  0 : leaf, ~ref = (first_slot << RT_LEAF_BITS) | count,  count in [1, RT_LEAF_MAX]
//                                 count == RT_LEAF_BIG means: first_slot indexes bigLeaves[] = {first, count}
//                          RT_REF_NONE : absent child (never hit)
#pragma once
#include <stdint.h>

#define RT_LEAF_BITS 6
#define RT_LEAF_BIG 63
#define RT_LEAF_MAX 62
#define RT_REF_NONE 0x7fffffff
#define RT_MAX_SLOTS (1u << (31 - RT_LEAF_BITS))
#define RT_MAX_STACK 64

struct alignas(32) NodeRec {
    float aMinX, aMinY, aMinZ, aMaxX;
    float aMaxY, aMaxZ, bMinX, bMinY;
    float bMinZ, bMaxX, bMaxY, bMaxZ;
    int32_t refA, refB, pad0, pad1;
};

struct alignas(32) TriRec {
    // first 32 bytes: what the distance test needs (plane: v0 and N); second 32 bytes: the edges, fetched only by the
    // lanes whose hit distance is in range (rt_device.cuh triangleTest) — one L1 wavefront less for the others
    float v0x, v0y, v0z, nx;
    float ny, nz, pad0, pad1;
    float abx, aby, abz, acx;
    float acy, acz, pad2, pad3;
};

struct alignas(32) ShadeRec {
    float n0x, n0y, n0z, n1x;
    float n1y, n1z, n2x, n2y;
    float n2z, colR, colG, colB;
    float emcR, emcG, emcB, emission;
    float roughness, pad[7];
};
static_assert(sizeof(NodeRec) == 64 && sizeof(TriRec) == 64 && sizeof(ShadeRec) == 96, "record sizes");

struct DeviceScene {
    const NodeRec* nodes;
    const TriRec* tris;
    const ShadeRec* shade;
    const uint32_t* triId;
    const uint2* bigLeaves;
    float rootMin[3], rootMax[3];   // the root's own bounds (slab-tested once per ray, fragment.glsl:253-262)
    int32_t rootRef;
    int32_t stackDepth;             // entries a traversal can need (tree depth + 1)
    uint32_t nNodes, nSlots;
    int32_t deferEdges;             // per pass: triangleTest fetches a triangle's edges only where the distance test passed
};

// Everything a pass needs, passed by value as a __grid_constant__ kernel parameter.
struct PassParams {
    DeviceScene scene;
    // uniform block (fragment.glsl:7-38)
    int32_t tileX, tileY, tileSizeX, tileSizeY, numTilesX, numTilesY;
    int32_t lambertian, nBounces, raysPerPixel, frameNumber;
    float skyBrightness, xStep, yStep, dirStartX, dirStartY, jitterAmount;
    float camPos[3], camRight[3], camUp[3], camForward[3];
    // targets
    int32_t width, height;
    const float4* prev;
    float4* next;
    float4* peerNext;               // nullable: the same target on the assembling rank (peer tile mode), written as well
    int32_t* primaryIds;            // nullable
    unsigned long long* counters;   // nullable (RT_CNT_COUNT entries)
    // compact launch domain: the pixels of this tile only
    int32_t ownedW, ownedH;
    int32_t sumMode;                // 1: next = prev + curr (frame-sharded progressive mode), 0: running mean
    unsigned int* workCounter;      // persistent variant
    // wavefront variant: queue of the pixels whose primary ray hit (stage 1 -> stage 2)
    float4* queueA;                 // primary direction xyz, RNG state
    float4* queueB;                 // primary hit t, u, v, slot
    uint32_t* queuePix;             // px | py << 16
    uint4* queueWork;               // counting passes only: work of the primary ray
    unsigned int* queueCount;
    // samples variant: three more record planes, the table of F(k) (specK entries per queued pixel) and a 2nd work counter
    float4* queueC;
    float4* queueD;
    float4* queueE;
    float4* specBuf;
    int32_t specK;
    uint32_t specCapacity;          // queued pixels the table has room for
    uint32_t specBufEntries;        // float4 entries of specBuf
    unsigned int* workCounters;     // 2 per round (evaluation, chain), zeroed per pass
    unsigned int* todoCount;        // [rounds + 1] pixels whose chain left its table
    uint32_t* todoFlat;             // pixels parked for the next round, in parking order
    uint32_t* todoSorted;           // this round's pixels in class order (round > 0)
    unsigned int* classCount;       // [rounds + 1][32]: pixels per size class, [30] = table entries reserved
    uint32_t* chainClass;           // per queued pixel: (rank within its class << 5) | class
    uint32_t* chainK;               // parked chain position per queued pixel
    float4* chainColor;             // parked colour sum (xyz) and sample index (w)
    int32_t specRound;
    int32_t specGate;               // AUTO: samples variant with the persistent kernel as its other branch (see vote::samplesSpeculate)
    int32_t tune[4];                // scheduling thresholds (0 = built-in default); bench tooling only
    int32_t topCount;               // persistent variant: nodes [0, topCount) are staged in shared memory
    int32_t chainTune[4];           // chain variant: quota scale (sixteenths), helping on/off, window, scattered queue order
    float4* chainRing;              // chain variant: kRing entries per lane of the grid (global memory)
};

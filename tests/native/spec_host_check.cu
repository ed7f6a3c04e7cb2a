// Host-side check of the samples variant's arithmetic (no GPU needed): the LCG jump tables against stepping
// RandomValue()'s state update (fragment.glsl:206-208) one draw at a time, and the size-class ladder.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../include/rt_b200.h"
#include "../../opengl_raytracer_b200/csrc/rt_spec.cuh"

static uint32_t stepLcg(uint32_t s) { return s * 747796405u + 2891336453u; }

int main() {
    static uint2 lo[spec::kJumpLo], hi[spec::kJumpHi];
    spec::fillJumpTables(lo, hi);
    const uint32_t seeds[] = {0u, 1u, 0xdeadbeefu, 0xffffffffu, 123456789u};
    long checked = 0;
    for (uint32_t seed : seeds) {
        uint32_t s = seed;
        for (unsigned int k = 0; k < (unsigned int)spec::kMaxK; k++) {
            const uint2 a = lo[k & (spec::kJumpLo - 1)], b = hi[k >> 8];
            const uint32_t jumped = (seed * a.x + a.y) * b.x + b.y;
            if (jumped != s) { std::printf("jump mismatch seed %u k %u: %u vs %u\n", seed, k, jumped, s); return 1; }
            s = stepLcg(stepLcg(stepLcg(s)));        // one surface hit = three draws
            checked++;
        }
    }
    // offsets beyond the tables (k >= kMaxK) go through jumpBlocks: compare with stepping, incl. the table boundary
    for (uint32_t seed : seeds) {
        uint32_t s = seed;
        const unsigned int last = 3u * (unsigned int)spec::kMaxK + 77u;
        for (unsigned int k = 0; k <= last; k++) {
            if (k >= (unsigned int)spec::kMaxK - 2u && (k < (unsigned int)spec::kMaxK + 600u || k % 97u == 0u || k == last)) {
                const uint2 a = lo[k & (spec::kJumpLo - 1)];
                const uint32_t jumped = spec::jumpBlocks(seed * a.x + a.y, k >> 8, hi[1]);
                if (jumped != s) { std::printf("far jump mismatch seed %u k %u: %u vs %u\n", seed, k, jumped, s); return 1; }
                checked++;
            }
            s = stepLcg(stepLcg(stepLcg(s)));
        }
    }
    unsigned int prev = 0;
    for (int c = 0; c < spec::kClasses; c++) {
        const unsigned int K = spec::classK(c);
        if (K <= prev || K % 2u != 0u) { std::printf("class %d: %u after %u\n", c, K, prev); return 1; }
        if (c > 0 && 4u * K > 5u * prev) { std::printf("class %d grows by more than 1.25x\n", c); return 1; }
        prev = K;
    }
    if (spec::classK(0) != 8u || prev < (unsigned int)spec::kMaxK / 2u) { std::printf("class range\n"); return 1; }
    for (int spp : {1, 2, 16, 64, 1024, 100000}) {
        const int K = spec::tableK(spp);
        if (K < 8 || K > spec::kMaxK / 4 || (spp < 1000 && K < spp)) { std::printf("tableK(%d) = %d\n", spp, K); return 1; }
    }
    std::printf("ok %ld jumps\n", checked);
    return 0;
}

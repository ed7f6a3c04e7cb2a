// ldg256_probe.cu — micro-benchmark behind the node/triangle record loads of librt_b200 (profiles/r02_variants.md section 1).
// Every lane of a warp reads one 64-byte record at an unrelated index of a window (what a node step does), either as
// 4 x LDG.128 or as 2 x LDG.256 (ld.global.nc.v8.f32, sm_100+).  Reports records/s for an L1-resident and an
// L2-resident window.  build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ldg256_probe ldg256_probe.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>

struct alignas(32) Rec { float f[16]; };

__device__ __forceinline__ void ld256(const void* p, float (&v)[8]) {
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
                 : "l"(p));
}

template <int MODE>   // 0: 4 x LDG.128, 1: 2 x LDG.256, 2: 1 x LDG.256 (32-byte record half), 3: 2 x LDG.128
__global__ void __launch_bounds__(128) probe(const Rec* __restrict__ recs, uint32_t mask, int iters, float* out) {
    uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    float acc = 0.f;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 4; u++) {
            s = s * 747796405u + 2891336453u;
            const Rec* r = recs + ((s >> 8) & mask);
            if (MODE == 0) {
                const float4* q = reinterpret_cast<const float4*>(r);
                const float4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2), d = __ldg(q + 3);
                acc += ((a.x + a.y) + (a.z + a.w)) + ((b.x + b.y) + (b.z + b.w)) + ((c.x + c.y) + (c.z + c.w)) + ((d.x + d.y) + (d.z + d.w));
            } else if (MODE == 1) {
                float a[8], b[8];
                ld256(r, a); ld256(reinterpret_cast<const char*>(r) + 32, b);
                acc += ((a[0] + a[1]) + (a[2] + a[3])) + ((a[4] + a[5]) + (a[6] + a[7])) + ((b[0] + b[1]) + (b[2] + b[3])) + ((b[4] + b[5]) + (b[6] + b[7]));
            } else if (MODE == 2) {
                float a[8];
                ld256(r, a);
                acc += ((a[0] + a[1]) + (a[2] + a[3])) + ((a[4] + a[5]) + (a[6] + a[7]));
            } else {
                const float4* q = reinterpret_cast<const float4*>(r);
                const float4 a = __ldg(q), b = __ldg(q + 1);
                acc += ((a.x + a.y) + (a.z + a.w)) + ((b.x + b.y) + (b.z + b.w));
            }
        }
    }
    if (acc == 12345.678f) out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int MODE>
float run(const Rec* recs, uint32_t mask, int grid, int iters, float* sink) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    probe<MODE><<<grid, 128>>>(recs, mask, iters, sink);
    float best = 1e30f;
    for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(e0);
        probe<MODE><<<grid, 128>>>(recs, mask, iters, sink);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return best;
}

int main() {
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
    const int sms = prop.multiProcessorCount;
    const size_t maxRecs = size_t(1) << 20;   // 64 MB
    Rec* recs; float* sink;
    cudaMalloc(&recs, maxRecs * sizeof(Rec)); cudaMemset(recs, 0, maxRecs * sizeof(Rec));
    cudaMalloc(&sink, size_t(sms) * 6 * 128 * sizeof(float));
    const int grid = sms * 6, iters = 2048;
    const double n = double(grid) * 128.0 * iters * 4.0;
    printf("device %s, %d SMs; grid %d x 128, records per launch %.3g\n", prop.name, sms, grid, n);
    const char* names[4] = {"4 x LDG.128 (64 B)", "2 x LDG.256 (64 B)", "1 x LDG.256 (32 B)", "2 x LDG.128 (32 B)"};
    for (int lg = 8; lg <= 20; lg += 3) {   // 256 records = 16 KB (L1), 2 K = 128 KB, 16 K = 1 MB, 128 K = 8 MB, 1 M = 64 MB (L2)
        const uint32_t mask = (1u << lg) - 1u;
        float ms[4];
        ms[0] = run<0>(recs, mask, grid, iters, sink);
        ms[1] = run<1>(recs, mask, grid, iters, sink);
        ms[2] = run<2>(recs, mask, grid, iters, sink);
        ms[3] = run<3>(recs, mask, grid, iters, sink);
        for (int m = 0; m < 4; m++)
            printf("window %8.0f KB  %-20s %8.3f ms  %8.2f G records/s  %8.2f GB/s\n", (mask + 1.0) * 64 / 1024, names[m], ms[m],
                   n / (ms[m] * 1e-3) * 1e-9, n * (m < 2 ? 64 : 32) / (ms[m] * 1e-3) * 1e-9);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e == cudaSuccess ? 0 : 1;
}

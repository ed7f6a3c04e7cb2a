// rt_peaks.cu — micro-benchmarks for the two on-SM ceilings the traversal kernels sit under, and for L2 -> SM bandwidth
// (MEASURED_PEAKS.json has the HBM and tensor peaks only).  bench.py runs them on the same GPU, at the
// same clocks, right after the timed region and reports roofline.issue / roofline.l1_wavefronts against them.
//
//   issue          warp instructions per second: 8 warps per scheduler, 8 independent FFMA chains each (inline PTX, so
//                  -fmad=false cannot split them); counted = the FFMAs only (loop overhead is 3 in 131 instructions).
//   l1 wavefronts  every lane of a warp loads 16 bytes from its own 128-byte line of an L1-resident 16 KB window: 32
//                  wavefronts per load instruction, the cost of a node fetch by 32 lanes at 32 different nodes.
//   l2             coalesced 16-byte loads over a 64 MB buffer (L2-resident, far beyond L1), 16 sweeps.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/rt_b200.h"

namespace {

constexpr int kIssueUnroll = 16, kIssueChains = 8;

__global__ void __launch_bounds__(256) peakIssueKernel(float* out, int iters, float seed) {
    float a0 = seed, a1 = seed + 1.f, a2 = seed + 2.f, a3 = seed + 3.f, a4 = seed + 4.f, a5 = seed + 5.f, a6 = seed + 6.f,
          a7 = seed + 7.f;
    const float m = 0.999f, c = 0.001f;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < kIssueUnroll; u++) {
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a0) : "f"(m), "f"(c));
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a1) : "f"(m), "f"(c));
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a2) : "f"(m), "f"(c));
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a3) : "f"(m), "f"(c));
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a4) : "f"(m), "f"(c));
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a5) : "f"(m), "f"(c));
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a6) : "f"(m), "f"(c));
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a7) : "f"(m), "f"(c));
        }
    }
    const float s = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
    if (s == 12345.678f) out[blockIdx.x * blockDim.x + threadIdx.x] = s;   // never true: keeps the chains alive
}

constexpr int kWaveUnroll = 8;

// window: 16 KB = 128 lines of 128 B; lane l reads line (l * 4 + j) & 127 at a 16-byte offset that moves with i
__global__ void __launch_bounds__(256) peakWavefrontKernel(const float4* __restrict__ window, float* out, int iters) {
    const uint32_t lane = threadIdx.x & 31u;
    float acc = 0.f;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < kWaveUnroll; j++) {
            const uint32_t line = (lane * 4u + uint32_t(j) + uint32_t(i)) & 127u;
            const float4 v = __ldg(window + line * 8u + (uint32_t(i + j) & 7u));
            acc += v.x + v.w;
        }
    }
    if (acc == 12345.678f) out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

__global__ void __launch_bounds__(256) peakL2Kernel(const float4* __restrict__ buf, size_t n, int sweeps, float* out) {
    float acc = 0.f;
    const size_t step = size_t(gridDim.x) * blockDim.x;
    for (int s = 0; s < sweeps; s++)
        for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += step) {
            const float4 v = __ldcg(buf + i);     // L2 only: the sweep would not fit L1 anyway
            acc += v.x + v.y + v.z + v.w;
        }
    if (acc == 12345.678f) out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <typename F>
cudaError_t timeMs(cudaStream_t stream, F&& launch, float* ms) {
    cudaEvent_t e0, e1;
    cudaError_t e = cudaEventCreate(&e0);
    if (e != cudaSuccess) return e;
    if ((e = cudaEventCreate(&e1)) != cudaSuccess) { cudaEventDestroy(e0); return e; }
    float best = 1e30f;
    launch();                                       // warm-up
    for (int rep = 0; rep < 3 && e == cudaSuccess; rep++) {
        cudaEventRecord(e0, stream);
        launch();
        cudaEventRecord(e1, stream);
        e = cudaEventSynchronize(e1);
        float t = 0.f;
        if (e == cudaSuccess) e = cudaEventElapsedTime(&t, e0, e1);
        if (t < best) best = t;
    }
    if (e == cudaSuccess) e = cudaGetLastError();
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    *ms = best;
    return e;
}

}  // namespace

// out[0] = G warp-instructions / s, out[1] = G L1 data-pipe wavefronts / s, out[2] = L2 -> SM GB/s
cudaError_t rtMeasurePeaks(int sms, cudaStream_t stream, double* out) {
    const size_t l2Bytes = size_t(64) << 20;
    float4* buf = nullptr;
    float* sink = nullptr;
    cudaError_t e = cudaMalloc(&buf, l2Bytes);
    if (e != cudaSuccess) return e;
    if ((e = cudaMalloc(&sink, size_t(sms) * 8 * 256 * sizeof(float))) != cudaSuccess) { cudaFree(buf); return e; }
    cudaMemsetAsync(buf, 0, l2Bytes, stream);
    float ms = 0.f;
    // issue: 8 CTAs x 256 threads per SM = 16 warps per scheduler
    const int issueIters = 4096, issueGrid = sms * 8;
    e = timeMs(stream, [&] { peakIssueKernel<<<issueGrid, 256, 0, stream>>>(sink, issueIters, 1.0f); }, &ms);
    if (e == cudaSuccess) {
        const double warpInst = double(issueGrid) * 8.0 * double(issueIters) * kIssueUnroll * kIssueChains;
        out[0] = warpInst / (double(ms) * 1e-3) * 1e-9;
        const int waveIters = 4096, waveGrid = sms * 8;
        e = timeMs(stream, [&] { peakWavefrontKernel<<<waveGrid, 256, 0, stream>>>(buf, sink, waveIters); }, &ms);
        if (e == cudaSuccess) {
            const double wavefronts = double(waveGrid) * 8.0 * double(waveIters) * kWaveUnroll * 32.0;
            out[1] = wavefronts / (double(ms) * 1e-3) * 1e-9;
            const int sweeps = 16;
            e = timeMs(stream, [&] { peakL2Kernel<<<sms * 8, 256, 0, stream>>>(buf, l2Bytes / sizeof(float4), sweeps, sink); }, &ms);
            if (e == cudaSuccess) out[2] = double(l2Bytes) * sweeps / (double(ms) * 1e-3) * 1e-9;
        }
    }
    cudaFree(buf); cudaFree(sink);
    return e;
}

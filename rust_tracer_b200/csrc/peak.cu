// peak.cu -- FP32 roofline denominator: an FFMA-saturation micro-kernel.
// MEASURED_PEAKS.json carries no FP32 figure, so bench.py measures one on the box it runs
// on: 16 independent FFMA chains per thread, 3-register form (what the sweep issues).
#include "kernels.h"

namespace rt {

constexpr int PEAK_BLOCK = 256;
constexpr int PEAK_CHAINS = 16;

__global__ void __launch_bounds__(PEAK_BLOCK) fp32_peak_kernel(int iters, float *sink, float seed_a, float seed_b) {
    float acc[PEAK_CHAINS];
    // operands come from kernel arguments so nothing is constant-folded
    float a = seed_a + (float)threadIdx.x * 1e-9f;
    float b = seed_b;
#pragma unroll
    for (int c = 0; c < PEAK_CHAINS; ++c) acc[c] = (float)c;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
#pragma unroll
            for (int c = 0; c < PEAK_CHAINS; ++c) acc[c] = fmaf(acc[c], a, b);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < PEAK_CHAINS; ++c) s += acc[c];
    if (s == 123.456f) sink[blockIdx.x * blockDim.x + threadIdx.x] = s; // never true in practice; keeps the chains live
}

double launch_fp32_peak(int blocks, int iters, float *sink, cudaStream_t stream) {
    fp32_peak_kernel<<<blocks, PEAK_BLOCK, 0, stream>>>(iters, sink, 0.999999f, 1e-7f);
    return 2.0 * (double)blocks * PEAK_BLOCK * (double)iters * 4.0 * PEAK_CHAINS;
}

} // namespace rt

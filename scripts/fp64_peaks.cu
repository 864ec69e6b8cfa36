// scripts/fp64_peaks.cu — measured FP64 pipe peaks of the B200 (roofline denominators of the NN closure).
//
// MEASURED_PEAKS.json carries HBM and bf16 numbers only; the closure's GEMMs run on the FP64 tensor
// cores (mma.sync.m8n8k4.f64 -> DMMA.8x8x4) and its tanh on the FP64 FMA pipe, so this program times
//   dmma      : back-to-back DMMAs, CH independent accumulator chains per warp
//   dfma      : back-to-back DFMAs, CH independent chains per thread
//   mix_warp  : every warp alternates 1 DMMA : R DFMAs  (same warp feeds both pipes)
//   mix_split : half of the warps issue only DMMAs, the other half only DFMAs
// and prints one JSON line with TFLOP/s for each (512 flop per DMMA warp-instruction, 64 per DFMA
// warp-instruction).  mix_* tell whether the two pipes overlap or share hardware.
//
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o scripts/_build/fp64_peaks scripts/fp64_peaks.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { std::printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); std::exit(1); } } while (0)

__device__ __forceinline__ void dmma(double& d0, double& d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

template <int CH>
__global__ void __launch_bounds__(1024) k_dmma(double* out, int iters, double a, double b) {
    double c0[CH], c1[CH];
#pragma unroll
    for (int i = 0; i < CH; ++i) { c0[i] = threadIdx.x * 1e-3 + i; c1[i] = i; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < CH; ++i) dmma(c0[i], c1[i], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < CH; ++i) s += c0[i] + c1[i];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int CH>
__global__ void __launch_bounds__(1024) k_dfma(double* out, int iters, double a, double b) {
    double c[CH];
#pragma unroll
    for (int i = 0; i < CH; ++i) c[i] = threadIdx.x * 1e-3 + i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < CH; ++i) asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(c[i]) : "d"(a), "d"(b));
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < CH; ++i) s += c[i];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// every warp: per iteration 4 DMMAs (4 chains) + 4*R DFMAs (8 chains)
template <int R>
__global__ void __launch_bounds__(1024) k_mix_warp(double* out, int iters, double a, double b) {
    double c0[4], c1[4], f[8];
#pragma unroll
    for (int i = 0; i < 4; ++i) { c0[i] = threadIdx.x * 1e-3 + i; c1[i] = i; }
#pragma unroll
    for (int i = 0; i < 8; ++i) f[i] = threadIdx.x * 1e-3 + i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            dmma(c0[i], c1[i], a, b);
#pragma unroll
            for (int r = 0; r < R; ++r)
                asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(f[(i * R + r) & 7]) : "d"(a), "d"(b));
        }
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 4; ++i) s += c0[i] + c1[i];
#pragma unroll
    for (int i = 0; i < 8; ++i) s += f[i];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// even warps: DMMA only (8 chains); odd warps: DFMA only (8 chains), `R` DFMAs per DMMA of the partner
template <int R>
__global__ void __launch_bounds__(1024) k_mix_split(double* out, int iters, double a, double b) {
    const int warp = threadIdx.x >> 5;
    double c0[8], c1[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { c0[i] = threadIdx.x * 1e-3 + i; c1[i] = i; }
    if (warp & 1) {
        for (int it = 0; it < iters * R; ++it) {
#pragma unroll
            for (int i = 0; i < 8; ++i) asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(c0[i]) : "d"(a), "d"(b));
        }
    } else {
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 8; ++i) dmma(c0[i], c1[i], a, b);
        }
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += c0[i] + c1[i];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
static double time_ms(F&& launch, int reps) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    launch(); launch();
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < reps; ++r) {
        CK(cudaEventRecord(e0));
        launch();
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    return best;
}

int main() {
    int nsm = 0;
    CK(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, 0));
    double* out;
    CK(cudaMalloc(&out, sizeof(double) * 1024 * 4096));
    const int iters = 4096;
    std::printf("{\"sms\": %d", nsm);
    // warps per SM sweep: block sizes 128..1024, one block per SM
    for (int threads : {128, 256, 512, 1024}) {
        const int blocks = nsm;
        const double warps = (double)blocks * threads / 32;
        double ms = time_ms([&] { k_dmma<8><<<blocks, threads>>>(out, iters, 1.0000001, 0.9999999); }, 5);
        std::printf(", \"dmma_ch8_w%d_TF\": %.3f", threads / 32, warps * iters * 8 * 512.0 / (ms * 1e-3) / 1e12);
        ms = time_ms([&] { k_dmma<4><<<blocks, threads>>>(out, iters, 1.0000001, 0.9999999); }, 5);
        std::printf(", \"dmma_ch4_w%d_TF\": %.3f", threads / 32, warps * iters * 4 * 512.0 / (ms * 1e-3) / 1e12);
        ms = time_ms([&] { k_dmma<2><<<blocks, threads>>>(out, iters, 1.0000001, 0.9999999); }, 5);
        std::printf(", \"dmma_ch2_w%d_TF\": %.3f", threads / 32, warps * iters * 2 * 512.0 / (ms * 1e-3) / 1e12);
        ms = time_ms([&] { k_dmma<1><<<blocks, threads>>>(out, iters, 1.0000001, 0.9999999); }, 5);
        std::printf(", \"dmma_ch1_w%d_TF\": %.3f", threads / 32, warps * iters * 1 * 512.0 / (ms * 1e-3) / 1e12);
        ms = time_ms([&] { k_dfma<8><<<blocks, threads>>>(out, iters, 1.0000001, 0.9999999); }, 5);
        std::printf(", \"dfma_ch8_w%d_TF\": %.3f", threads / 32, warps * iters * 8 * 64.0 / (ms * 1e-3) / 1e12);
    }
    {
        const int threads = 512, blocks = nsm;
        const double warps = (double)blocks * threads / 32;
        double ms = time_ms([&] { k_mix_warp<1><<<blocks, threads>>>(out, iters, 1.0000001, 0.9999999); }, 5);
        std::printf(", \"mixwarp_r1\": {\"dmma_TF\": %.3f, \"dfma_TF\": %.3f}", warps * iters * 4 * 512.0 / (ms * 1e-3) / 1e12,
                    warps * iters * 4 * 64.0 / (ms * 1e-3) / 1e12);
        ms = time_ms([&] { k_mix_warp<2><<<blocks, threads>>>(out, iters, 1.0000001, 0.9999999); }, 5);
        std::printf(", \"mixwarp_r2\": {\"dmma_TF\": %.3f, \"dfma_TF\": %.3f}", warps * iters * 4 * 512.0 / (ms * 1e-3) / 1e12,
                    warps * iters * 8 * 64.0 / (ms * 1e-3) / 1e12);
        ms = time_ms([&] { k_mix_warp<4><<<blocks, threads>>>(out, iters, 1.0000001, 0.9999999); }, 5);
        std::printf(", \"mixwarp_r4\": {\"dmma_TF\": %.3f, \"dfma_TF\": %.3f}", warps * iters * 4 * 512.0 / (ms * 1e-3) / 1e12,
                    warps * iters * 16 * 64.0 / (ms * 1e-3) / 1e12);
        ms = time_ms([&] { k_mix_warp<8><<<blocks, threads>>>(out, iters, 1.0000001, 0.9999999); }, 5);
        std::printf(", \"mixwarp_r8\": {\"dmma_TF\": %.3f, \"dfma_TF\": %.3f}", warps * iters * 4 * 512.0 / (ms * 1e-3) / 1e12,
                    warps * iters * 32 * 64.0 / (ms * 1e-3) / 1e12);
        ms = time_ms([&] { k_mix_split<2><<<blocks, threads>>>(out, iters, 1.0000001, 0.9999999); }, 5);
        std::printf(", \"mixsplit_r2\": {\"dmma_TF\": %.3f, \"dfma_TF\": %.3f}", warps / 2 * iters * 8 * 512.0 / (ms * 1e-3) / 1e12,
                    warps / 2 * iters * 2 * 8 * 64.0 / (ms * 1e-3) / 1e12);
        ms = time_ms([&] { k_mix_split<8><<<blocks, threads>>>(out, iters, 1.0000001, 0.9999999); }, 5);
        std::printf(", \"mixsplit_r8\": {\"dmma_TF\": %.3f, \"dfma_TF\": %.3f}", warps / 2 * iters * 8 * 512.0 / (ms * 1e-3) / 1e12,
                    warps / 2 * iters * 8 * 8 * 64.0 / (ms * 1e-3) / 1e12);
    }
    std::printf("}\n");
    cudaFree(out);
    return 0;
}
